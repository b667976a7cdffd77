import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run on the GPU box with -m gpu)")


@pytest.fixture(scope="session")
def mcp():
    import autonomous_microservice_composition_via_llm_agents_in_an_mcp_control_plane_b200 as pkg

    return pkg


@pytest.fixture(scope="session")
def oracle():
    import oracle as o

    o.build()
    return o
