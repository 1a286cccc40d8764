"""Import shim: exposes the package in `ray-tracer-rust_b200/` (not a valid identifier) as module `rtb200`."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ray-tracer-rust_b200")
_spec = importlib.util.spec_from_file_location("ray_tracer_rust_b200", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_pkg = importlib.util.module_from_spec(_spec)
sys.modules["ray_tracer_rust_b200"] = _pkg
_spec.loader.exec_module(_pkg)
globals().update({k: v for k, v in vars(_pkg).items() if not k.startswith("__")})
