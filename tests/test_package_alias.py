"""The package is importable under a valid identifier (INTEGRATION.md stub B)."""
import importlib


def test_alias_is_the_same_package():
    import threedgut_b200
    from threedgut_b200 import Tracer, abi, camera
    from threedgut_b200.native import SplatRaster
    import threedgut_b200.dist as d

    real = importlib.import_module("3dgrut_b200")
    assert threedgut_b200 is real and abi is real.abi and Tracer is real.Tracer
    assert SplatRaster is importlib.import_module("3dgrut_b200.native").SplatRaster
    assert d is importlib.import_module("3dgrut_b200.dist") and camera is importlib.import_module("3dgrut_b200.camera")
