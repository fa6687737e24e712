#!/usr/bin/env python3
"""Photon / caustics map fixtures from the REAL reference -> tests/golden/photon_<scene>.npz. TEST INFRASTRUCTURE.

Three stages of the reference's -use-photon-map path, each pinned separately:
  emit      the body of the emission loops (src/renderers/renderer.cpp:143-196, :214-266: RandomPhoton, TraceNodeNormal,
            IsPhotonSurface, MtlBlinn_PhotonMap::RandomPhotonBounce) for emissions 0..EMISSIONS-1 of both maps, fed with the
            counter-based stream of (emission, map, seed) (oracle patch P4) -> photons per emission
  estimate  cyPhotonMap itself (src/ext/cyPhotonMap.h): the photons of those emissions through AddPhoton / ScalePhotonPowers /
            PrepareForIrradianceEstimation, EstimateIrradiance<100> (quadratic filter) at QUERIES surface points; plus the
            photons as the map hands them back (direction / power quantisation)
  image     the reference's own render with the maps on (its own emission pass and random streams), twice with different
            seeds: 160x120, 64 spp, linear radiance

    python tools/make_photon_golden.py [scene ...]
"""
import ctypes as C
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
# trc_mtl_glass has a glass object without a diffuse colour: its caustics map can fill. The example "caustics" scenes give every
# material a diffuse colour, so no photon ever reaches a diffuse surface by a specular-only route and the reference's caustics
# loop (renderer.cpp:214-266) never ends on them: they get the emission and estimate stages of the photon map only.
SCENES = ["trc_mtl_glass", "example_project12_caustics_glossy"]
WITH_IMAGE = {"trc_mtl_glass"}
SEED = 0x5EED
EMISSIONS = 4096
BOUNCE = 20
QUERIES = 2048
IMG = (160, 120, 64)
MAPS = (10000, 1000)  # RendererParam's default sizes (renderer.h:52,55)


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


def emit(ref, which, first, n):
    counts = np.zeros(n, np.int32)
    rec = np.zeros((n, BOUNCE, 9), np.float32)
    ref.lib.qaref_emit_photons.argtypes = [C.c_int, C.c_uint32, C.c_int64, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p]
    ref.lib.qaref_emit_photons.restype = C.c_int
    assert ref.lib.qaref_emit_photons(which, first, n, SEED, BOUNCE, vp(counts), vp(rec)) == 1, "scene has no photon source"
    return counts, rec


def flat_photons(counts, rec):
    return np.concatenate([rec[e, :counts[e]] for e in range(len(counts)) if counts[e]] or [np.zeros((0, 9), np.float32)], 0)


def estimate(ref, ph, scale, radius, qpos, qnrm):
    n, nq = len(ph), len(qpos)
    irrad, dout = np.zeros((nq, 3), np.float32), np.zeros((nq, 3), np.float32)
    bdir, bpow = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
    pos, d, pw = (np.ascontiguousarray(ph[:, k:k + 3]) for k in (0, 3, 6))
    ref.lib.qaref_photon_estimate.argtypes = [C.c_int64] + [C.c_void_p] * 3 + [C.c_float, C.c_float, C.c_int64] + [C.c_void_p] * 6
    ref.lib.qaref_photon_estimate(n, vp(pos), vp(d), vp(pw), scale, radius, nq, vp(qpos), vp(qnrm), vp(irrad), vp(dout), vp(bdir), vp(bpow))
    return irrad, dout, bdir, bpow


def image_worker(name, which, out):
    from oracle import refapi
    W, H, spp = IMG
    xml = open(os.path.join(INPUTS, name + ".xml")).read()
    xml = re.sub(r'<width\s+value="\d+"\s*/>', '<width value="%d"/>' % W, xml)
    xml = re.sub(r'<height\s+value="\d+"\s*/>', '<height value="%d"/>' % H, xml)
    with tempfile.NamedTemporaryFile("w", suffix=".xml", delete=False) as fp:
        fp.write(xml)
        path = fp.name
    ref = refapi.Ref()
    ref.set_params(spp, spp, 5, True, 1)
    ref.lib.qaref_set_photon.argtypes = [C.c_int] * 3
    ref.lib.qaref_set_photon(1, MAPS[0], MAPS[1])
    ref.seed(0x1234567 + 977 * which)  # the emission pass of ComputeScene draws from this stream
    cwd = os.getcwd()
    ref.load(path, ASSETS)  # chdir()s into the asset directory; ComputeScene writes photonmap.dat / caustics.dat there
    for f in ("photonmap.dat", "caustics.dat"):
        if os.path.exists(f):
            os.unlink(f)
    os.chdir(cwd)
    os.unlink(path)
    rgb, depth = ref.render_linear_serial(spp)
    np.savez(out, rgb=rgb)


def one(name):
    from oracle import refapi
    ref = refapi.Ref()
    ref.set_params(1, 1, 5, True, 1)
    ref.load(os.path.join(INPUTS, name + ".xml"), ASSETS)
    out = {}
    rng = np.random.default_rng(SEED)
    for which, radius in ((0, 0.2), (1, 1.0)):
        counts, rec = emit(ref, which, 0, EMISSIONS)
        ph = flat_photons(counts, rec)
        if len(ph) == 0:
            print("%-36s map %d: no photon stored in %d emissions — stage skipped" % (name, which, EMISSIONS))
            continue
        out["counts%d" % which] = counts
        out["photons%d" % which] = ph
        emitted = int((counts > 0).sum())
        scale = np.float32(1.0) / np.float32(max(1, emitted))
        # query points: on the surfaces the photons landed on (jittered inside the gather radius), facing the photons
        pick = rng.integers(0, max(1, len(ph)), QUERIES)
        base = ph[pick] if len(ph) else np.zeros((QUERIES, 9), np.float32)
        qpos = (base[:, :3] + rng.normal(scale=0.3 * radius, size=(QUERIES, 3))).astype(np.float32)
        qn = -base[:, 3:6] + rng.normal(scale=0.3, size=(QUERIES, 3))
        qn = (qn / np.maximum(1e-9, np.linalg.norm(qn, axis=1, keepdims=True))).astype(np.float32)
        irrad, dout, bdir, bpow = estimate(ref, ph, scale, radius, qpos, qn)
        out.update({"scale%d" % which: scale, "radius%d" % which: np.float32(radius), "qpos%d" % which: qpos, "qnrm%d" % which: qn,
                    "irrad%d" % which: irrad, "dir%d" % which: dout, "back_dir%d" % which: bdir, "back_power%d" % which: bpow})
        print("%-36s map %d: %d photons from %d emissions (%d stored something), %d of %d estimates non-zero" % (
            name, which, len(ph), EMISSIONS, emitted, int((irrad.sum(1) > 0).sum()), QUERIES))
    if name not in WITH_IMAGE:
        np.savez_compressed(os.path.join(GOLDEN, "photon_%s.npz" % name), seed=np.uint32(SEED), bounce=np.int32(BOUNCE), **out)
        return
    tmp = tempfile.mkdtemp(prefix="photon_")
    procs = []
    for w in range(2):
        o = os.path.join(tmp, "img%d.npz" % w)
        procs.append((subprocess.Popen([sys.executable, os.path.abspath(__file__), "--image", name, str(w), o], cwd=ROOT), o))
    imgs = []
    for p, o in procs:
        assert p.wait(timeout=900) == 0
        imgs.append(np.load(o)["rgb"])
    out["img_a"], out["img_b"] = imgs[0], imgs[1]  # heavy-tailed (photon powers divided by small pick probabilities): fp32
    out["img_size"] = np.array(IMG, np.int32)
    out["map_sizes"] = np.array(MAPS, np.int32)
    d = imgs[0] - imgs[1]
    print("%-36s image %dx%d %d spp with maps %s: rmse(a, b) %.5f  mean %.5f" % (name, IMG[0], IMG[1], IMG[2], MAPS, float(np.sqrt((d * d).mean())),
                                                                           float(imgs[0].mean())))
    np.savez_compressed(os.path.join(GOLDEN, "photon_%s.npz" % name), seed=np.uint32(SEED), bounce=np.int32(BOUNCE), **out)


if __name__ == "__main__":
    if len(sys.argv) > 4 and sys.argv[1] == "--image":
        image_worker(sys.argv[2], int(sys.argv[3]), sys.argv[4])
        sys.exit(0)
    if len(sys.argv) > 2 and sys.argv[1] == "--one":
        one(sys.argv[2])
        sys.exit(0)
    for n in sys.argv[1:] or SCENES:
        subprocess.run([sys.executable, os.path.abspath(__file__), "--one", n], cwd=ROOT, check=True, timeout=1500)
