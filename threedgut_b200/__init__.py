"""``threedgut_b200`` -- the importable name of this repository's package.

The package directory is called ``3dgrut_b200`` (prescribed by the build's layout), which is not a Python identifier:
``import 3dgrut_b200`` is a SyntaxError and every consumer would have to go through ``importlib``.  This alias makes

    from threedgut_b200 import Tracer, abi          # drop-in for `from threedgut_tracer import Tracer`
    import threedgut_b200.native

work, and it is the SAME module objects (``threedgut_b200 is importlib.import_module("3dgrut_b200")``): no code lives here.
"""
import importlib
import sys

_real = importlib.import_module("3dgrut_b200")
for _sub in ("abi", "camera", "dist", "native", "optimizers", "protocols", "scene_io", "synthetic", "tracer"):
    sys.modules[f"{__name__}.{_sub}"] = importlib.import_module(f"3dgrut_b200.{_sub}")
sys.modules[__name__] = _real
