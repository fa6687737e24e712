#!/usr/bin/env python3
"""Converged-image fixtures from the REAL reference -> tests/golden/img_<scene>.npz. TEST INFRASTRUCTURE.

For each BASELINE config: the scene XML with <width>/<height> reduced (written to a temp file; everything else
verbatim), rendered twice by the reference's own Shade recursion and running-mean accumulation
(oracle/ref_driver.cpp qaref_render_linear_serial = PixelRender's loop body, renderer.cpp:310-346, serial) with two
different xorshift seeds, fixed spp, bounce 5. The reference is not reproducible run to run in normal use (time-seeded
RNG), so image parity is statistical: the GPU image must be as close to ref_a as ref_b is.

    python tools/make_image_golden.py [scene ...]
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

# scene -> (width, height, spp)
SCENES = {
    "example_project3_sphere": (160, 120, 64),
    "example_project7_object": (160, 120, 64),
    "example_project12_caustics_glossy": (160, 120, 64),
    "trc_scene_tower": (160, 90, 64),
    "trc_scene_xmas": (128, 96, 32),
}


def worker(name):
    from oracle import refapi
    W, H, spp = SCENES[name]
    xml = open(os.path.join(INPUTS, name + ".xml")).read()
    xml = re.sub(r'<width\s+value="\d+"\s*/>', '<width value="%d"/>' % W, xml)
    xml = re.sub(r'<height\s+value="\d+"\s*/>', '<height value="%d"/>' % H, xml)
    with tempfile.NamedTemporaryFile("w", suffix=".xml", delete=False) as fp:
        fp.write(xml)
        path = fp.name
    ref = refapi.Ref()
    ref.set_params(spp, spp, 5, True, 1)
    ref.load(path, ASSETS)
    os.unlink(path)
    assert (ref.W, ref.H) == (W, H), (ref.W, ref.H)
    imgs = []
    for seed in (12345, 987654321):
        ref.seed(seed)
        rgb, depth = ref.render_linear_serial(spp)
        imgs.append(rgb)
    np.savez_compressed(os.path.join(GOLDEN, "img_%s.npz" % name), ref_a=imgs[0], ref_b=imgs[1], depth=depth,
                        size=np.array([W, H, spp], np.int32))
    d = np.clip(imgs[0], 0, 1) - np.clip(imgs[1], 0, 1)
    print("%-40s %dx%d %d spp  rmse(a,b) clipped-linear %.5f  mean %.5f" % (name, W, H, spp, float(np.sqrt((d * d).mean())), float(imgs[0].mean())))


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--worker":
        worker(sys.argv[2])
        sys.exit(0)
    for nm in sys.argv[1:] or list(SCENES):
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--worker", nm], cwd=ROOT, stdout=subprocess.PIPE,
                           stderr=subprocess.STDOUT, text=True)
        tail = [l for l in r.stdout.splitlines() if l.startswith(nm)]
        print(tail[-1] if tail else "%s FAILED (%d)\n%s" % (nm, r.returncode, r.stdout[-1500:]))
