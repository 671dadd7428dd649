"""CPU-side checks of the C-ABI library: it loads without a GPU, exports every symbol include/gut_b200.h declares,
and its struct layouts match the ctypes mirror.  No compute calls."""
import ctypes as C
import importlib
import os
import re

from tests.helpers import ROOT, abi


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "gut_b200.h")).read()
    return sorted(set(re.findall(r"GUT_API\s+[\w\s\*]+?\b(gut_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    native = importlib.import_module("3dgrut_b200.native")
    assert os.path.exists(native.LIB_PATH), "libgut_b200.so not built (run __graft_entry__.build())"
    lib = C.CDLL(native.LIB_PATH)
    names = _declared_symbols()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), f"{n} declared in gut_b200.h but not exported"


def test_struct_layouts_match():
    native = importlib.import_module("3dgrut_b200.native")
    lib = native.load_library()
    assert lib.gut_abi_version() == abi.GUT_ABI_VERSION
    assert lib.gut_sizeof_config() == C.sizeof(abi.GutConfig)
    assert lib.gut_sizeof_camera() == C.sizeof(abi.GutCamera)
    assert lib.gut_sizeof_stage_outputs() == C.sizeof(abi.GutStageOutputs)
    assert lib.gut_sizeof_frame_stats() == C.sizeof(abi.GutFrameStats)


def test_default_config_matches_python_mirror():
    native = importlib.import_module("3dgrut_b200.native")
    lib = native.load_library()
    c = abi.GutConfig()
    lib.gut_default_config(C.byref(c))
    d = abi.default_config()
    for name, _ in abi.GutConfig._fields_:
        assert getattr(c, name) == getattr(d, name), name


def test_stage_names_follow_reference_tags():
    native = importlib.import_module("3dgrut_b200.native")
    lib = native.load_library()
    names = [lib.gut_stage_name(i).decode() for i in range(lib.gut_num_stages())]
    # DeviceLaunchesLogger tags, threedgut_tracer/src/gutRenderer.cu:256-378,447,475
    for tag in ("render::project", "render::prepare-expand", "render::expand", "render::sort", "render::render",
                "render-backward::render", "render-backward::project"):
        assert tag in names


def test_colour_slot_sizes():
    """gut_colour_slot_floats (host arithmetic, no GPU): dense = [N,3] padded to 16 bytes; compact = 4 header floats + one
    (mask, first row) pair per 32 particles (padded to 16 bytes) + room for every row."""
    native = importlib.import_module("3dgrut_b200.native")
    lib = native.load_library()
    for n in (0, 1, 31, 32, 33, 3001, 3_000_000):
        rows = (3 * n + 3) // 4 * 4
        assert lib.gut_colour_slot_floats(n, abi.GUT_COLOUR_DENSE) == rows
        assert lib.gut_colour_slot_floats(n, abi.GUT_COLOUR_COMPACT) == 4 + (2 * ((n + 31) // 32) + 3) // 4 * 4 + rows
        assert lib.gut_colour_slot_floats(n, abi.GUT_COLOUR_COMPACT) % 4 == 0
