#!/usr/bin/env python3
"""Fixtures for the material classes the reference's `MtlBlinn` alias can select -> tests/golden/cls_<class>_<scene>.npz.
TEST INFRASTRUCTURE.

The reference picks the class at compile time (src/materials/materials.h:57-61); oracle/Makefile builds one parity
library per class (libqaref_parity_<class>.so = the reference with that one line switched). For each (class, scene):

* blinn_basic / phong_basic on scenes without glossy materials: the Whitted-style recursion draws no random number, so
  the reference image is deterministic -> `ref` (linear fp32 means), compared per pixel.
* pathtracing / montecarlogi (and the Basic classes on glossy scenes): two reference renders with different seeds
  -> `ref_a`, `ref_b`, compared statistically like tests/golden/img_*.npz.
* pathtracing additionally: single-sample radiance with the counter-based stream (oracle/Makefile patch P4) -> `rad_*`;
  the class spawns one child per hit, so the GPU path consumes the stream in the reference's order.

    python tools/make_class_golden.py [class:scene ...]
"""
import os
import re
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
INPUTS = os.path.join(ROOT, "oracle", "_ref", "inputs")
ASSETS = os.path.join(ROOT, "assets")
GOLDEN = os.path.join(ROOT, "tests", "golden")
SEED = 0x5EED

# (class, scene) -> (width, height, spp, deterministic)
CASES = {
    ("blinn_basic", "example_project3_sphere"): (160, 120, 4, True),
    ("blinn_basic", "example_project7_object"): (160, 120, 2, True),
    ("blinn_basic", "trc_scene_tower"): (160, 90, 2, True),
    ("blinn_basic", "example_project12_caustics_glossy"): (128, 96, 32, False),
    ("phong_basic", "example_project3_sphere"): (160, 120, 4, True),
    ("phong_basic", "example_project4"): (160, 120, 2, True),
    ("pathtracing", "example_project3_sphere"): (160, 120, 64, False),
    ("pathtracing", "example_project12_caustics_glossy"): (128, 96, 64, False),
    ("pathtracing", "trc_scene_tower"): (128, 72, 64, False),
    ("montecarlogi", "example_project3_sphere"): (128, 96, 16, False),
    ("montecarlogi", "example_project12_caustics_glossy"): (96, 72, 16, False),
}


def worker(cls, name):
    from oracle import refapi
    W, H, spp, det = CASES[(cls, name)]
    xml = open(os.path.join(INPUTS, name + ".xml")).read()
    xml = re.sub(r'<width\s+value="\d+"\s*/>', '<width value="%d"/>' % W, xml)
    xml = re.sub(r'<height\s+value="\d+"\s*/>', '<height value="%d"/>' % H, xml)
    with tempfile.NamedTemporaryFile("w", suffix=".xml", delete=False) as fp:
        fp.write(xml)
        path = fp.name
    ref = refapi.Ref(cls)
    ref.set_params(spp, spp, 5, True, 1)
    ref.load(path, ASSETS)
    os.unlink(path)
    assert (ref.W, ref.H) == (W, H), (ref.W, ref.H)
    out = {"size": np.array([W, H, spp], np.int32)}
    imgs = []
    for seed in (12345, 987654321):
        ref.seed(seed)
        rgb, depth = ref.render_linear_serial(spp)
        imgs.append(rgb)
    out["depth"] = depth
    if det:
        assert np.array_equal(imgs[0], imgs[1]), "class %s on %s is not deterministic" % (cls, name)
        out["ref"] = imgs[0]
    else:
        out["ref_a"], out["ref_b"] = imgs
    if cls == "pathtracing":
        rs = np.random.RandomState(SEED)
        n_rad = 1500
        rp = np.stack([rs.randint(0, W, n_rad), rs.randint(0, H, n_rad), rs.randint(0, 64, n_rad)], 1).astype(np.int32)
        rgb, z, hit = ref.samples(rp, SEED)
        out["rad_pix"], out["rad_rgb"], out["rad_seed"] = rp, rgb, np.array([SEED], np.uint32)
    np.savez_compressed(os.path.join(GOLDEN, "cls_%s_%s.npz" % (cls, name)), **out)
    d = np.clip(imgs[0], 0, 1) - np.clip(imgs[1], 0, 1)
    print("%s:%s %dx%d %d spp  rmse(a,b) %.5f  mean %.5f" % (cls, name, W, H, spp, float(np.sqrt((d * d).mean())), float(imgs[0].mean())))


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--worker":
        worker(*sys.argv[2].split(":"))
        sys.exit(0)
    for key in sys.argv[1:] or ["%s:%s" % k for k in CASES]:
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--worker", key], cwd=ROOT, stdout=subprocess.PIPE,
                           stderr=subprocess.STDOUT, text=True)
        tail = [l for l in r.stdout.splitlines() if l.startswith(key)]
        print(tail[-1] if tail else "%s FAILED (%d)\n%s" % (key, r.returncode, r.stdout[-1500:]))
