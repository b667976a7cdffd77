"""B200-native retrieval + level scheduling behind the reference control plane's Python surface.

Only the hot path lives here: the CUDA library (``csrc/`` ->
``libmcp_b200.so``, C ABI in ``include/mcp_b200.h``) and the host-side mirror of the reference
interface for that path (``control_plane.py``).  No CPU fallback: importing is cheap, but any
compute call needs the library and a B200.
"""
from . import _native  # noqa: F401
from ._native import McpError, build_info, device_count  # noqa: F401
from .index import ServiceIndex, merge_topk_device  # noqa: F401
from . import dag, telemetry  # noqa: F401

__all__ = ["ServiceIndex", "merge_topk_device", "dag", "telemetry", "McpError", "build_info", "device_count"]
