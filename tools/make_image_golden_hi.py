#!/usr/bin/env python3
"""High-spp converged-image fixtures from the REAL reference -> tests/golden/imghi_<scene>.npz. TEST INFRASTRUCTURE.

north_star: "the converged image is checked by per-pixel RMSE within a stated tolerance at matched high spp";
SURVEY.md 8(c)2: N in {64, 1024}. tools/make_image_golden.py holds the 64-spp thumbnails; this tool renders the five
BASELINE configs at 320x240 (tower 320x180) with 1024 spp (C1, C3, C4) or 256 spp (C2, C5: 32-tap textures / 24 lights
make the serial reference ~10x slower per sample), twice with different xorshift seeds, through the reference's own
Shade recursion and running mean (oracle/ref_driver.cpp qaref_render_linear_rows = PixelRender's loop body,
renderer.cpp:310-346). The frame is split into row bands over worker processes (the reference's scene is a process
global); each band gets its own seed. Images are stored as UNCLIPPED linear fp32 means.

    python tools/make_image_golden_hi.py [-j PROCS] [scene ...]
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
    "example_project3_sphere": (320, 240, 1024),
    "example_project12_caustics_glossy": (320, 240, 1024),
    "trc_scene_tower": (320, 180, 1024),
    "example_project7_object": (320, 240, 256),
    "trc_scene_xmas": (320, 240, 256),
}
SEEDS = (12345, 987654321)
BANDS = 12


def worker(name, which, band, out_path):
    from oracle import refapi
    import ctypes as C
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
    j0, j1 = H * band // BANDS, H * (band + 1) // BANDS
    ref.seed((SEEDS[which] * 2654435761 + band * 40503 + 1) & 0xFFFFFFFF or 1)
    rgb = np.zeros((H, W, 3), np.float32)
    depth = np.zeros((H, W), np.float32)
    ref.lib.qaref_render_linear_rows.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    ref.lib.qaref_render_linear_rows(spp, j0, j1, rgb.ctypes.data_as(C.c_void_p), depth.ctypes.data_as(C.c_void_p))
    np.savez(out_path, rgb=rgb[j0:j1], depth=depth[j0:j1])


def main(argv):
    procs = 6
    if argv[:1] == ["-j"]:
        procs, argv = int(argv[1]), argv[2:]
    names = argv or list(SCENES)
    tmp = tempfile.mkdtemp(prefix="imghi_")
    jobs = [(n, w, b) for n in names for w in range(2) for b in range(BANDS)]
    running = []
    while jobs or running:
        while jobs and len(running) < procs:
            n, w, b = jobs.pop(0)
            out = os.path.join(tmp, "%s_%d_%d.npz" % (n, w, b))
            p = subprocess.Popen([sys.executable, os.path.abspath(__file__), "--worker", n, str(w), str(b), out], cwd=ROOT,
                                 stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            running.append((p, n, w, b))
        p, n, w, b = running.pop(0)
        if p.wait() != 0:
            raise SystemExit("worker %s %d %d failed" % (n, w, b))
    for n in names:
        W, H, spp = SCENES[n]
        imgs, depth = [], None
        for w in range(2):
            parts = [np.load(os.path.join(tmp, "%s_%d_%d.npz" % (n, w, b))) for b in range(BANDS)]
            imgs.append(np.concatenate([q["rgb"] for q in parts], 0))
            depth = np.concatenate([q["depth"] for q in parts], 0)
        # fp16 storage: 2^-11 relative, two orders below the Monte-Carlo noise of either image; halves the fixture
        np.savez_compressed(os.path.join(GOLDEN, "imghi_%s.npz" % n), ref_a=imgs[0].astype(np.float16), ref_b=imgs[1].astype(np.float16), depth=depth,
                            size=np.array([W, H, spp], np.int32))
        d = imgs[0] - imgs[1]
        rel = d / (0.5 * (imgs[0] + imgs[1]) + 0.05)
        print("%-40s %dx%d %d spp  rmse(a,b) linear %.5f  relative %.5f  mean %.5f  max %.2f" % (
            n, W, H, spp, float(np.sqrt((d * d).mean())), float(np.sqrt((rel * rel).mean())), float(imgs[0].mean()), float(imgs[0].max())))


if __name__ == "__main__":
    if len(sys.argv) > 5 and sys.argv[1] == "--worker":
        worker(sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), sys.argv[5])
        sys.exit(0)
    main(sys.argv[1:])
