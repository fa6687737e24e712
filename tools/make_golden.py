#!/usr/bin/env python3
"""Golden vectors from the REAL reference (oracle/_ref/libqaref_parity.so = the reference's own sources compiled by
oracle/Makefile) -> tests/golden/<scene>.npz. TEST INFRASTRUCTURE.

The reference keeps no tests or known-answer vectors for its pixel loop, so the vectors that
pin this path are outputs of the reference itself, run here on the committed stand-in assets (tools/gen_assets.py):

  nodes / camera / screen   scene tree with every Transformation, camera frame of Renderer::ComputeScene
  rays, hits                primary DiffRays of scattered pixels (samples 0..2) and incoherent secondary-style rays,
                            with what Scene::TraceNodeNormal returned (hit, node, face id, z, p, N, uvw, duvw, ...)
  srays, stmax, socc        shadow rays from those hit points toward every light, with Scene::TraceNodeShadow's answer
  pix, rgb, z               single-sample radiance of (i, j, s) from the reference's Shade recursion, its RNG replaced
                            by the counter-based stream (oracle/Makefile patch P4), seed recorded

    python tools/make_golden.py            # all scenes (one fresh process each: the reference scene is a global)
"""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
INPUTS = os.path.join(ROOT, "oracle", "_ref", "inputs")
ASSETS = os.path.join(ROOT, "assets")
GOLDEN = os.path.join(ROOT, "tests", "golden")
SEED = 0x5EED

# name -> (n primary pixels, n radiance samples, with radiance)
SCENES = {
    "example_project3_sphere": (700, 1500, True),          # config 1
    "example_project7_object": (700, 1200, True),          # config 2
    "example_project12_caustics_glossy": (700, 1200, True),  # config 3
    "trc_scene_tower": (700, 1000, True),                  # config 4
    "trc_scene_xmas": (700, 800, True),                    # config 5
    "example_project4": (400, 800, True),
    "example_project7_checkboard": (400, 800, True),
    "example_project12_box": (400, 800, True),
    "trc_mtl_glass": (400, 800, True),
    "trc_mtl_glossy": (400, 800, True),
    "trc_mtl_coffee": (300, 600, True),
    "example_project3_box": (300, 600, True),
    "example_project6": (300, 0, False),                   # traversal only
}


def light_table(xml, assets):
    """lights of the scene from this repo's host loader (needs no GPU) -> list of (type, position, direction)"""
    import ctypes as C
    from qaray_b200 import capi
    lib = capi.load()
    h = C.c_void_p()
    capi.check(lib, lib.qg_host_scene_load_xml(os.fsencode(xml), os.fsencode(assets), C.byref(h)))
    d = lib.qg_host_scene_desc(h).contents
    out = [(d.lights[k].type, np.array(d.lights[k].position[:], np.float32), np.array(d.lights[k].direction[:], np.float32))
           for k in range(d.num_lights)]
    lib.qg_host_scene_destroy(h)
    return out


def worker(name):
    from oracle import refapi
    n_pix, n_rad, with_rad = SCENES[name]
    xml = os.path.join(INPUTS, name + ".xml")
    ref = refapi.Ref()
    ref.set_params(1, 1, 5, True, 1)
    ref.load(xml, ASSETS)
    W, H = ref.W, ref.H
    rs = np.random.RandomState(SEED)
    out = {}
    meta, xf = ref.nodes()
    out["node_meta"], out["node_xf"] = meta, xf
    out["camera"], out["screen"] = ref.camera(), ref.screen()
    # primary rays
    pix = np.stack([rs.randint(0, W, n_pix), rs.randint(0, H, n_pix)], 1).astype(np.int32)
    rays, samp = [], []
    for s in range(3):
        rays.append(ref.primary_rays(pix, s))
        samp.append(np.full(n_pix, s, np.int32))
    rays = np.concatenate(rays)
    out["prim_pix"] = np.concatenate([np.concatenate([pix] * 3), np.concatenate(samp)[:, None]], 1).astype(np.int32)
    has_diff = np.ones(len(rays), np.uint8)
    hits = ref.trace_closest(rays, has_diff)
    # incoherent secondary-style rays from the hit points (and a few from the camera position)
    hp = hits["p"][hits["hit"] == 1]
    if len(hp) == 0:
        hp = np.zeros((1, 3), np.float32)
    n_sec = n_pix
    org = hp[rs.randint(0, len(hp), n_sec)]
    d = rs.normal(size=(n_sec, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    sec = np.zeros((n_sec, 18), np.float32)
    for k in range(3):
        sec[:, 6 * k:6 * k + 3] = org
        sec[:, 6 * k + 3:6 * k + 6] = d
    sec_hits = ref.trace_closest(sec, np.zeros(n_sec, np.uint8))
    out["rays"] = np.concatenate([rays, sec])
    out["has_diff"] = np.concatenate([has_diff, np.zeros(n_sec, np.uint8)])
    out["hits"] = np.concatenate([hits, sec_hits])
    # shadow rays toward every light from the primary hit points
    lights = light_table(xml, ASSETS)
    sr, st = [], []
    pts = hits["p"][hits["hit"] == 1][: max(1, 1200 // max(1, len(lights)))]
    for (lt, lpos, ldir) in lights:
        if lt == 0 or len(pts) == 0:
            continue
        if lt == 1:
            dd = np.tile(-ldir, (len(pts), 1)).astype(np.float32)
            tm = np.full(len(pts), 1e30, np.float32)
        else:
            dv = (lpos[None, :] - pts).astype(np.float32)
            tm = np.linalg.norm(dv, axis=1).astype(np.float32)
            dd = (dv / np.maximum(tm, 1e-20)[:, None]).astype(np.float32)
        sr.append(np.concatenate([pts, dd], 1).astype(np.float32))
        st.append(tm)
    if sr:
        out["srays"], out["stmax"] = np.concatenate(sr), np.concatenate(st)
        out["socc"] = ref.trace_shadow(out["srays"], out["stmax"])
    # single-sample radiance with the counter-based stream
    if with_rad and n_rad:
        rp = np.stack([rs.randint(0, W, n_rad), rs.randint(0, H, n_rad), rs.randint(0, 64, n_rad)], 1).astype(np.int32)
        rgb, z, hit = ref.samples(rp, SEED)
        out["rad_pix"], out["rad_rgb"], out["rad_z"], out["rad_hit"] = rp, rgb, z, hit
        out["rad_seed"] = np.array([SEED], np.uint32)
    os.makedirs(GOLDEN, exist_ok=True)
    np.savez_compressed(os.path.join(GOLDEN, name + ".npz"), **out)
    print("%-40s nodes %3d  rays %5d (hit %5d)  shadow %5d  radiance %5d" % (
        name, len(meta), len(out["rays"]), int(out["hits"]["hit"].sum()), len(out.get("srays", [])), n_rad if with_rad else 0))


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--worker":
        worker(sys.argv[2])
        sys.exit(0)
    names = sys.argv[1:] or list(SCENES)
    for nm in names:
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--worker", nm], cwd=ROOT,
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        tail = [l for l in r.stdout.splitlines() if l.startswith(nm)]
        print(tail[-1] if tail else "%s FAILED (%d)\n%s" % (nm, r.returncode, r.stdout[-2000:]))
