"""ray-tracer-rust_b200 -- B200-native per-pixel ray-casting + shading path behind the reference's entry points.

The directory name is not a Python identifier; import it through the root-level `rtb200` shim.
"""
from . import binding  # noqa: F401
from .binding import *  # noqa: F401,F403
from .binding import lib, LIB_PATH, EXPORTS, Scene, RtbError  # noqa: F401
