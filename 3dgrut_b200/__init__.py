"""B200-native 3DGUT rasteriser: drop-in for ``threedgut_tracer`` (reference:
threedgut_tracer/__init__.py, threedgut_tracer/tracer.py).

The directory name starts with a digit, so import it with
``importlib.import_module("3dgrut_b200")``; ``Tracer`` and the other heavy
symbols are resolved lazily so that ``3dgrut_b200.abi`` stays importable on a
machine without the CUDA library.
"""
import importlib

_LAZY = {
    "Tracer": ".tracer",
    "Batch": ".protocols",
    "SplatRaster": ".native",
    "load_library": ".native",
    "ShutterType": ".abi",
    "default_config": ".abi",
    "SelectiveAdam": ".optimizers",
    "update_gradient_buffer": ".optimizers",
    "read_ply": ".scene_io",
    "write_ply": ".scene_io",
    "read_colmap_cameras": ".scene_io",
    "read_colmap_images": ".scene_io",
    "colmap_views": ".scene_io",
}


def __getattr__(name):
    if name in _LAZY:
        mod = importlib.import_module(_LAZY[name], __name__)
        return getattr(mod, name)
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")


__all__ = sorted(_LAZY)
