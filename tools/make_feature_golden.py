#!/usr/bin/env python3
"""Fixtures for the sampling features outside the five BASELINE configs -> tests/golden/feat_<feature>_<scene>.npz.
TEST INFRASTRUCTURE. All from the reference itself (oracle/_ref/libqaref_parity.so):

* dof       depth of field (renderer.cpp:319-322, scene.cpp:104-111): two fixed-spp renders with different seeds, and
            single samples with the counter-based stream (the lens sample is the stream's first two numbers)
* area      PointLight with size > 0.01 (lights.cpp:50-65): two fixed-spp renders with different seeds
* adaptive  sppMin < sppMax, SuperSamplerHalton's variance stop (scene.cpp:92-97): two renders + samples taken per pixel

    python tools/make_feature_golden.py [feature:scene ...]
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

# (feature, scene) -> (width, height, spp_min, spp_max)
CASES = {
    ("dof", "example_project9"): (160, 120, 32, 32),
    ("area", "example_project10_test"): (128, 96, 16, 16),
    ("adaptive", "example_project3_sphere"): (160, 120, 4, 64),
    ("adaptive", "example_project12_caustics_glossy"): (96, 72, 4, 32),
}


def worker(feat, name):
    from oracle import refapi
    W, H, smin, smax = CASES[(feat, name)]
    xml = open(os.path.join(INPUTS, name + ".xml")).read()
    xml = re.sub(r'<width\s+value="\d+"\s*/>', '<width value="%d"/>' % W, xml)
    xml = re.sub(r'<height\s+value="\d+"\s*/>', '<height value="%d"/>' % H, xml)
    with tempfile.NamedTemporaryFile("w", suffix=".xml", delete=False) as fp:
        fp.write(xml)
        path = fp.name
    ref = refapi.Ref()
    ref.set_params(smin, smax, 5, True, 1)
    ref.load(path, ASSETS)
    os.unlink(path)
    assert (ref.W, ref.H) == (W, H), (ref.W, ref.H)
    out = {"size": np.array([W, H, smin, smax], np.int32)}
    imgs, counts = [], []
    for seed in (12345, 987654321):
        ref.seed(seed)
        rgb, depth, cnt = ref.render_adaptive_serial(smin, smax)
        imgs.append(rgb)
        counts.append(cnt)
    out["ref_a"], out["ref_b"], out["depth"] = imgs[0], imgs[1], depth
    if feat == "adaptive":
        out["count_a"], out["count_b"] = counts
    if feat == "dof":
        rs = np.random.RandomState(SEED)
        n_rad = 1500
        rp = np.stack([rs.randint(0, W, n_rad), rs.randint(0, H, n_rad), rs.randint(0, 64, n_rad)], 1).astype(np.int32)
        rgb, z, hit = ref.samples(rp, SEED)
        out["rad_pix"], out["rad_rgb"], out["rad_z"], out["rad_seed"] = rp, rgb, z, np.array([SEED], np.uint32)
    np.savez_compressed(os.path.join(GOLDEN, "feat_%s_%s.npz" % (feat, name)), **out)
    d = np.clip(imgs[0], 0, 1) - np.clip(imgs[1], 0, 1)
    print("%s:%s %dx%d spp %d..%d  rmse(a,b) %.5f  mean %.5f  mean samples %.2f / %.2f" % (
        feat, name, W, H, smin, smax, float(np.sqrt((d * d).mean())), float(imgs[0].mean()), counts[0].mean(), counts[1].mean()))


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--worker":
        worker(*sys.argv[2].split(":"))
        sys.exit(0)
    for key in sys.argv[1:] or ["%s:%s" % k for k in CASES]:
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--worker", key], cwd=ROOT, stdout=subprocess.PIPE,
                           stderr=subprocess.STDOUT, text=True)
        tail = [l for l in r.stdout.splitlines() if l.startswith(key)]
        print(tail[-1] if tail else "%s FAILED (%d)\n%s" % (key, r.returncode, r.stdout[-1500:]))
