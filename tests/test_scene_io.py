"""Scene and camera file formats, CPU only.

COLMAP: tests/golden/colmap holds small model files and what the REFERENCE's own readers returned for them
(tests/golden/make_scene_io_golden.py imported threedgrut/datasets/utils.py in the build container) -- pinned.
PLY: the reference reads and writes through `plyfile`, which is not installed, so the layout is checked against the
exporter's attribute list (ply_exporter.py:34-47) restated here and a hand-written ASCII file -- unpinned."""
import importlib
import json
import os

import numpy as np
import pytest

from tests.helpers import ROOT

sio = importlib.import_module("3dgrut_b200.scene_io")
GOLD = os.path.join(ROOT, "tests", "golden", "colmap")


@pytest.fixture(scope="module")
def expected():
    return json.load(open(os.path.join(GOLD, "expected_from_reference.json")))


@pytest.mark.parametrize("ext", ["bin", "txt"])
def test_colmap_readers_match_the_reference_readers(expected, ext):
    cams = sio.read_colmap_cameras(os.path.join(GOLD, f"cameras.{ext}"))
    imgs = sio.read_colmap_images(os.path.join(GOLD, f"images.{ext}"))
    exp = expected[ext]
    assert sorted(cams) == sorted(int(k) for k in exp["cameras"])
    for k, e in exp["cameras"].items():
        c = cams[int(k)]
        assert (c["id"], c["model"], c["width"], c["height"]) == (e["id"], e["model"], e["width"], e["height"])
        assert np.array_equal(c["params"], np.array(e["params"]))          # float64, bit-exact
    assert [im["name"] for im in imgs] == [e["name"] for e in exp["images"]]  # sorted by name like the reference
    for im, e in zip(imgs, exp["images"]):
        assert (im["id"], im["camera_id"]) == (e["id"], e["camera_id"])
        assert np.array_equal(im["qvec"], np.array(e["qvec"])) and np.array_equal(im["tvec"], np.array(e["tvec"]))
        assert np.array_equal(sio.qvec_to_so3(im["qvec"]), np.array(e["so3"]))


def test_colmap_views_follow_the_dataset_conventions(expected):
    cams = sio.read_colmap_cameras(os.path.join(GOLD, "cameras.bin"))
    imgs = sio.read_colmap_images(os.path.join(GOLD, "images.bin"))
    views = sio.colmap_views(cams, imgs, downsample=1)
    assert len(views) == len(imgs)
    for v, im in zip(views, imgs):
        # C2W = inv([R | t])  (dataset_colmap.py:261-267): the camera centre is -R^T t
        R = sio.qvec_to_so3(im["qvec"])
        np.testing.assert_allclose(v["T_to_world"][:3, 3], -R.T @ im["tvec"], rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(v["T_to_world"][:3, :3], R.T, rtol=1e-5, atol=1e-6)
        cam = cams[im["camera_id"]]
        if cam["model"] == "OPENCV_FISHEYE":
            intr = v["intrinsics_OpenCVFisheyeCameraModelParameters"]
            m = expected["max_radius"][0]
            assert intr["max_angle"] == pytest.approx(max(2 * m["value"] / 860.0, 2 * m["value"] / 861.5) / 2, rel=1e-6)
            assert intr["radial_coeffs"].shape == (4,)
        else:
            intr = v["intrinsics_OpenCVPinholeCameraModelParameters"]
            # the principal point is forced to the image centre, distortion is zero (dataset_colmap.py:133-147)
            assert np.array_equal(intr["principal_point"], np.array([cam["width"], cam["height"]], np.float32) / 2)
            assert not intr["radial_coeffs"].any() and intr["radial_coeffs"].shape == (6,)
            fy = cam["params"][0] if cam["model"] == "SIMPLE_PINHOLE" else cam["params"][1]
            assert np.array_equal(intr["focal_length"], np.array([cam["params"][0], fy], np.float32))
    half = sio.colmap_intrinsics(cams[1], downsample=2)[1]
    assert np.array_equal(half["focal_length"], np.array([1230.5 / 2, 1228.25 / 2], np.float32))
    with pytest.raises(ValueError):
        sio.colmap_intrinsics({"model": "OPENCV", "width": 10, "height": 10, "params": np.zeros(8)})
    # the views are accepted by the camera conversion of the tracer
    camera = importlib.import_module("3dgrut_b200.camera")
    protocols = importlib.import_module("3dgrut_b200.protocols")
    for v in views:
        kw = {k: v[k] for k in v if k.startswith("intrinsics")}
        rays = np.zeros((1, 2, 2, 3), np.float32)
        cam = camera.camera_from_batch(protocols.Batch(rays_ori=rays, rays_dir=rays, T_to_world=v["T_to_world"][None], **kw))
        assert cam.focal_length[0] > 0 and abs(sum(x * x for x in cam.start_pose[3:7]) - 1.0) < 1e-5


def test_max_radius_matches_reference(expected):
    for m in expected["max_radius"]:
        w, h = m["size"]
        pp = np.array(m["pp"], np.float32)
        corners = np.array([[0.0, 0.0], [w, 0.0], [0.0, h], [w, h]])
        assert float(np.max(np.linalg.norm(corners - pp[None].astype(np.float64), axis=1))) == pytest.approx(m["value"], rel=1e-12)


def _params(n, seed):
    rng = np.random.default_rng(seed)
    return {"positions": rng.normal(size=(n, 3)).astype(np.float32), "rotation": rng.normal(size=(n, 4)).astype(np.float32),
            "scale": rng.normal(size=(n, 3)).astype(np.float32), "density": rng.normal(size=(n, 1)).astype(np.float32),
            "features_albedo": rng.normal(size=(n, 3)).astype(np.float32), "features_specular": rng.normal(size=(n, 45)).astype(np.float32)}


@pytest.mark.parametrize("n", [0, 1, 257])
def test_ply_round_trip_and_layout(tmp_path, n):
    p = _params(n, n)
    path = str(tmp_path / "scene.ply")
    sio.write_ply(path, p)
    header = open(path, "rb").read().split(b"end_header\n")[0].decode().splitlines()
    assert header[0] == "ply" and header[1] == "format binary_little_endian 1.0" and header[2] == f"element vertex {n}"
    props = [ln.split()[-1] for ln in header if ln.startswith("property float")]
    # PLYExporter._construct_list_of_attributes (ply_exporter.py:34-47)
    want = ["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(3)] + [f"f_rest_{i}" for i in range(45)] + ["opacity"] + \
        [f"scale_{i}" for i in range(3)] + [f"rot_{i}" for i in range(4)]
    assert props == want == sio.ply_attribute_names()
    assert os.path.getsize(path) == len("\n".join(header)) + len("\nend_header\n") + n * 62 * 4
    back = sio.read_ply(path)
    for k in sio.PARAMETER_NAMES:
        assert back[k].dtype == np.float32 and np.array_equal(back[k], p[k]), k


def test_ply_specular_is_channel_major_in_the_file(tmp_path):
    """f_rest_j holds channel j // 15, coefficient j % 15 (init_from_ply reshapes [N,3,S] and transposes, model.py:692-693)."""
    p = _params(3, 7)
    path = str(tmp_path / "s.ply")
    sio.write_ply(path, p)
    v = sio._read_ply_vertices(path)
    spec = p["features_specular"].reshape(3, 15, 3)
    for j in (0, 1, 14, 15, 29, 30, 44):
        assert np.array_equal(v[f"f_rest_{j}"], spec[:, j % 15, j // 15])
    assert np.array_equal(v["nz"], np.ones(3, np.float32)) and not v["nx"].any()


def test_ascii_ply_with_reordered_properties(tmp_path):
    names = sio.ply_attribute_names()
    order = list(reversed(names))
    rows = np.arange(2 * len(names), dtype=np.float32).reshape(2, len(names)) / 8
    path = tmp_path / "a.ply"
    with open(path, "w") as fh:
        fh.write("ply\nformat ascii 1.0\ncomment hand written\nelement vertex 2\n" + "".join(f"property float {k}\n" for k in order) +
                 "element face 0\nproperty list uchar int vertex_indices\nend_header\n")
        for r in rows:
            fh.write(" ".join(repr(float(x)) for x in r) + "\n")
    got = sio.read_ply(str(path))
    col = {k: i for i, k in enumerate(order)}
    assert np.array_equal(got["positions"], rows[:, [col["x"], col["y"], col["z"]]])
    assert np.array_equal(got["rotation"], rows[:, [col[f"rot_{i}"] for i in range(4)]])
    assert np.array_equal(got["density"][:, 0], rows[:, col["opacity"]])
    with pytest.raises(ValueError):
        sio.read_ply(str(path), max_n_features=2)


def test_parameters_from_checkpoint():
    torch = pytest.importorskip("torch")
    p = _params(5, 1)
    ckpt = {k: torch.nn.Parameter(torch.from_numpy(v)) for k, v in p.items()}
    ckpt.update({"n_active_features": 2, "max_n_features": 3, "scene_extent": 4.5, "optimizer": {}, "background": {}})
    params, meta = sio.parameters_from_checkpoint(ckpt)
    assert meta == {"n_active_features": 2, "max_n_features": 3, "scene_extent": 4.5}
    for k in sio.PARAMETER_NAMES:
        assert np.array_equal(params[k], p[k])
