#!/usr/bin/env python3
"""Bit-exactness stress fixtures from the REAL reference -> tests/golden/rays_<scene>.npz. TEST INFRASTRUCTURE.

2^20 closest-hit rays (classes A, B, C of tests/raystress.py) and 2^20 any-hit rays per scene, traced by the reference
(Scene::TraceNodeNormal / TraceNodeShadow, src/scene/scene.cpp:35-74, through oracle/_ref/libqaref_parity.so). Committed per
scene: a SHA-256 digest per 4096 rays of (hit, node, face id, distance bits) / of the occlusion flags — 16 KB — plus every
64th record in full, so a mismatching block can be narrowed down without the oracle.

    python tools/make_ray_stress.py [scene ...]      (one process per scene: the reference's scene is a process global)
"""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
INPUTS = os.path.join(ROOT, "oracle", "_ref", "inputs")
ASSETS = os.path.join(ROOT, "assets")
GOLDEN = os.path.join(ROOT, "tests", "golden")
SCENES = ["trc_scene_tower", "trc_scene_xmas", "example_project7_object", "example_project12_caustics_glossy"]
N = 1 << 20


def to18(r6):
    return np.concatenate([r6, r6, r6], 1).astype(np.float32)  # DiffRay(p, d): x and y are copies (src/core/ray.h:51-56)


def one(name):
    import raystress as rs
    from oracle import refapi
    xml = os.path.join(INPUTS, name + ".xml")
    c, r = rs.scene_sphere(xml, ASSETS)
    ref = refapi.Ref()
    ref.set_params(1, 1, 5, True, 1)
    ref.load(xml, ASSETS)
    ab = rs.rays_ab(c, r, N // 2)
    hab = ref.trace_closest(to18(ab), np.zeros(len(ab), np.uint8))
    cc = rs.rays_c(ab, hab["hit"], hab["node"], hab["z"], N // 2)
    hc = ref.trace_closest(to18(cc), np.zeros(len(cc), np.uint8))
    s6, tmax = rs.rays_s(ab, hab["hit"], hab["z"], c, r, N)
    occ = ref.trace_shadow(s6, tmax)
    rec_ab = rs.pack_closest(hab["hit"], hab["node"], hab["face"], hab["z"])
    rec_c = rs.pack_closest(hc["hit"], hc["node"], hc["face"], hc["z"])
    np.savez_compressed(os.path.join(GOLDEN, "rays_%s.npz" % name), centre=c, radius=np.float32(r), n=np.int64(N),
                        dig_ab=rs.digests(rec_ab), dig_c=rs.digests(rec_c), dig_s=rs.digests(occ.astype(np.uint8)),
                        sub_ab=rec_ab[::64], sub_c=rec_c[::64], sub_s=occ[::64],
                        hits=np.array([hab["hit"].sum(), hc["hit"].sum(), occ.sum()], np.int64))
    print("%-40s R %.1f  hits A/B %d of %d  C %d of %d  occluded %d of %d" % (name, r, hab["hit"].sum(), len(ab), hc["hit"].sum(), len(cc),
                                                                       occ.sum(), len(occ)))


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--one":
        one(sys.argv[2])
        sys.exit(0)
    for n in sys.argv[1:] or SCENES:
        subprocess.run([sys.executable, os.path.abspath(__file__), "--one", n], cwd=ROOT, check=True)
