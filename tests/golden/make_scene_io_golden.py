"""Generates tests/golden/colmap/* : small COLMAP model files plus what the REFERENCE's own readers return for them.
Run in the build container only (it imports /root/reference/threedgrut/datasets/utils.py by path; that module needs numpy and
torch only).  The GPU box and the tests read the committed outputs, never /root/reference.

    python tests/golden/make_scene_io_golden.py
"""
import importlib.util
import json
import os
import struct
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "colmap")


def load_reference_utils():
    spec = importlib.util.spec_from_file_location("ref_dataset_utils", "/root/reference/threedgrut/datasets/utils.py")
    mod = importlib.util.module_from_spec(spec)
    sys.modules[spec.name] = mod     # dataclasses looks the defining module up while the file executes
    spec.loader.exec_module(mod)
    return mod


def main():
    os.makedirs(OUT, exist_ok=True)
    rng = np.random.default_rng(42)
    cameras = [(1, 1, 1237, 822, [1230.5, 1228.25, 618.5, 411.0]),                    # PINHOLE
               (2, 0, 640, 480, [512.0, 320.0, 240.0]),                              # SIMPLE_PINHOLE
               (7, 5, 1920, 1080, [860.0, 861.5, 955.25, 541.5, -0.02, 0.003, 1e-4, -2e-5])]   # OPENCV_FISHEYE
    images = []
    for i in range(5):
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        images.append((10 + i, q.tolist(), rng.normal(size=3).tolist(), cameras[i % 3][0], f"images/frame_{4 - i:03d}.png",
                       rng.uniform(0, 500, size=(i + 1, 2)).tolist(), rng.integers(-1, 99, size=i + 1).tolist()))
    with open(os.path.join(OUT, "cameras.bin"), "wb") as fh:
        fh.write(struct.pack("<Q", len(cameras)))
        for cid, mid, w, h, prm in cameras:
            fh.write(struct.pack("<iiQQ", cid, mid, w, h))
            fh.write(struct.pack("<" + "d" * len(prm), *prm))
    names = {0: "SIMPLE_PINHOLE", 1: "PINHOLE", 5: "OPENCV_FISHEYE"}
    with open(os.path.join(OUT, "cameras.txt"), "w") as fh:
        fh.write("# Camera list with one line of data per camera:\n#   CAMERA_ID, MODEL, WIDTH, HEIGHT, PARAMS[]\n")
        for cid, mid, w, h, prm in cameras:
            fh.write(f"{cid} {names[mid]} {w} {h} " + " ".join(repr(float(x)) for x in prm) + "\n")
    with open(os.path.join(OUT, "images.bin"), "wb") as fh:
        fh.write(struct.pack("<Q", len(images)))
        for iid, q, t, cid, name, xys, p3d in images:
            fh.write(struct.pack("<idddddddi", iid, *q, *t, cid))
            fh.write(name.encode("utf-8") + b"\x00")
            fh.write(struct.pack("<Q", len(xys)))
            for (x, y), pid in zip(xys, p3d):
                fh.write(struct.pack("<ddq", x, y, pid))
    with open(os.path.join(OUT, "images.txt"), "w") as fh:
        fh.write("# Image list with two lines of data per image:\n")
        for iid, q, t, cid, name, xys, p3d in images:
            fh.write(f"{iid} " + " ".join(repr(float(x)) for x in q + t) + f" {cid} {name}\n")
            fh.write(" ".join(f"{x!r} {y!r} {pid}" for (x, y), pid in zip(xys, p3d)) + "\n")

    ref = load_reference_utils()
    expected = {}
    for ext in ("bin", "txt"):
        cams = (ref.read_colmap_intrinsics_binary if ext == "bin" else ref.read_colmap_intrinsics_text)(os.path.join(OUT, f"cameras.{ext}"))
        imgs = (ref.read_colmap_extrinsics_binary if ext == "bin" else ref.read_colmap_extrinsics_text)(os.path.join(OUT, f"images.{ext}"))
        expected[ext] = {
            "cameras": {str(k): {"id": int(c.id), "model": c.model, "width": int(c.width), "height": int(c.height),
                                 "params": [float(x) for x in c.params]} for k, c in cams.items()},
            "images": [{"id": int(im.id), "qvec": [float(x) for x in im.qvec], "tvec": [float(x) for x in im.tvec],
                        "camera_id": int(im.camera_id), "name": im.name,
                        "so3": ref.qvec_to_so3(im.qvec).tolist()} for im in imgs],
        }
    # compute_max_radius as the fisheye path of ColmapDataset uses it (dataset_colmap.py:166-172)
    expected["max_radius"] = [{"size": [1920.0, 1080.0], "pp": [955.25, 541.5],
                               "value": ref.compute_max_radius(np.array([1920.0, 1080.0]), np.array([955.25, 541.5], dtype=np.float32))},
                              {"size": [640.0, 480.0], "pp": [400.0, 100.0],
                               "value": ref.compute_max_radius(np.array([640.0, 480.0]), np.array([400.0, 100.0], dtype=np.float32))}]
    with open(os.path.join(OUT, "expected_from_reference.json"), "w") as fh:
        json.dump(expected, fh, indent=1)
    print("wrote", OUT)


if __name__ == "__main__":
    main()
