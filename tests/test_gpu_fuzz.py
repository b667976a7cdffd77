"""A slice of tools/fuzz_parity.py in the suite: random shapes / metrics / paths / K2 shapes, every case bit-exact."""
import subprocess
import sys
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu


def test_randomised_shapes_are_bit_exact():
    tool = Path(__file__).resolve().parent.parent / "tools" / "fuzz_parity.py"
    r = subprocess.run([sys.executable, str(tool), "60", "4242"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert "60/60 cases bit-exact" in r.stdout
