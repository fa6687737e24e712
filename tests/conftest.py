import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
SCENE_DIR = os.path.join(ROOT, "tests", "scenes")          # this repo's own small scenes
REF_INPUTS = os.path.join(ROOT, "oracle", "_ref", "inputs")  # the reference's scene XMLs (placed by oracle/Makefile)
ASSETS = os.path.join(ROOT, "assets")                      # stand-in meshes/textures (tools/gen_assets.py)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden_names():
    if not os.path.isdir(GOLDEN_DIR):
        return []
    return sorted(f[:-4] for f in os.listdir(GOLDEN_DIR) if f.endswith(".npz") and not f.startswith(("img_", "imghi_", "cls_", "feat_", "edge_", "rays_", "photon_", "cliff_")))


CLASS_ID = {"pathtracing": 1, "montecarlogi": 2, "blinn_basic": 3, "phong_basic": 4}  # QG_CLASS_* (include/qaray_b200.h)


def class_golden_cases():
    """material-class fixtures (tools/make_class_golden.py): cls_<class>_<scene>.npz -> [(class, scene)]"""
    out = []
    if os.path.isdir(GOLDEN_DIR):
        for f in sorted(os.listdir(GOLDEN_DIR)):
            for cls in CLASS_ID:
                if f.startswith("cls_%s_" % cls) and f.endswith(".npz"):
                    out.append((cls, f[len("cls_%s_" % cls):-4]))
    return out


def image_golden_names():
    """converged-image fixtures (tools/make_image_golden.py): img_<scene>.npz"""
    if not os.path.isdir(GOLDEN_DIR):
        return []
    return sorted(f[4:-4] for f in os.listdir(GOLDEN_DIR) if f.endswith(".npz") and f.startswith("img_"))


def load_golden(name):
    return np.load(os.path.join(GOLDEN_DIR, name + ".npz"))


def scene_xml(name):
    """path of a reference input scene; skips the test when the inputs/assets are not in place"""
    p = os.path.join(REF_INPUTS, name + ".xml")
    if not os.path.exists(p):
        pytest.skip("reference inputs missing (run `make -C oracle ref` where /root/reference exists)")
    if not os.path.exists(os.path.join(ASSETS, "STANDIN_ASSETS.txt")):
        pytest.skip("stand-in assets missing (run tools/gen_assets.py)")
    return p


def rays18_to_struct(rays18, has_diff):
    from qaray_b200 import capi
    rays18 = np.asarray(rays18, np.float32).reshape(-1, 18)
    r = np.zeros(len(rays18), capi.RAY_DTYPE)
    r["p"], r["dir"] = rays18[:, 0:3], rays18[:, 3:6]
    r["xp"], r["xdir"] = rays18[:, 6:9], rays18[:, 9:12]
    r["yp"], r["ydir"] = rays18[:, 12:15], rays18[:, 15:18]
    r["has_diff"] = np.asarray(has_diff, np.int32)
    return r


def struct_to_rays18(r):
    return np.concatenate([r["p"], r["dir"], r["xp"], r["xdir"], r["yp"], r["ydir"]], 1).astype(np.float32)
