#!/usr/bin/env python3
"""Edge-case ray fixtures from the REAL reference -> tests/golden/edge_<scene>.npz. TEST INFRASTRUCTURE.

Rays the random fixtures of tools/make_golden.py never contain: directions with one or two components exactly zero or
below the reference's 1e-7 "parallel" threshold (the slab test of src/core/box.cpp:94-128 and TriObj::TraceBVHNode then
ignores that axis — the case the CUDA path's degenerate-axis cull must not change), rays starting exactly on geometry
(t <= bias rejection), and tmax-limited shadow rays that end just before / just after a surface.

    python tools/make_edge_golden.py [scene ...]
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
SCENES = ["trc_scene_tower", "example_project7_object", "example_project3_sphere"]


def worker(name):
    from oracle import refapi
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    ref = refapi.Ref()
    ref.set_params(1, 1, 5, True, 1)
    ref.load(os.path.join(INPUTS, name + ".xml"), ASSETS)
    rs = np.random.RandomState(0xED6E)
    hp = g["hits"]["p"][g["hits"]["hit"] == 1]
    cam = g["camera"][:3]
    n = 3000
    org = np.where(rs.rand(n, 1) < 0.5, hp[rs.randint(0, len(hp), n)], cam[None, :] + rs.normal(size=(n, 3)) * 5).astype(np.float32)
    d = rs.normal(size=(n, 3)).astype(np.float32)
    kind = rs.randint(0, 4, n)
    for k in range(n):
        ax = rs.permutation(3)
        if kind[k] == 0:
            d[k, ax[0]] = 0.0                      # one exactly-zero component
        elif kind[k] == 1:
            d[k, ax[0]] = 0.0
            d[k, ax[1]] = 0.0                      # axis-parallel ray
        elif kind[k] == 2:
            d[k, ax[0]] = np.float32(rs.choice([3e-8, -3e-8, 9.9e-8, 1.01e-7, -1.2e-7]))  # around the 1e-7 threshold
        else:
            d[k, ax[0]] = np.float32(-0.0)
    nz = np.linalg.norm(d, axis=1, keepdims=True)
    d = (d / nz).astype(np.float32)
    d[np.abs(d) < 1e-30] = 0.0
    rays = np.zeros((n, 18), np.float32)
    for k in range(3):
        rays[:, 6 * k:6 * k + 3] = org
        rays[:, 6 * k + 3:6 * k + 6] = d
    hits = ref.trace_closest(rays, np.zeros(n, np.uint8))
    # shadow rays: same rays, tmax just below / above the closest hit distance, and unlimited
    z = hits["z"].copy()
    z[hits["hit"] == 0] = 10.0
    scale = np.float32(1.0) + rs.choice(np.array([-1e-3, -1e-6, 0.0, 1e-6, 1e-3], np.float32), n)
    tmax = (z * scale).astype(np.float32)
    occ = ref.trace_shadow(rays[:, :6], tmax)
    np.savez_compressed(os.path.join(GOLDEN, "edge_%s.npz" % name), rays=rays, hits=hits, tmax=tmax, occ=occ)
    print("%s rays %d hit %d occluded %d" % (name, n, int(hits["hit"].sum()), int(occ.sum())))


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--worker":
        worker(sys.argv[2])
        sys.exit(0)
    for nm in sys.argv[1:] or SCENES:
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--worker", nm], cwd=ROOT, stdout=subprocess.PIPE,
                           stderr=subprocess.STDOUT, text=True)
        tail = [l for l in r.stdout.splitlines() if l.startswith(nm)]
        print(tail[-1] if tail else "%s FAILED (%d)\n%s" % (nm, r.returncode, r.stdout[-1500:]))
