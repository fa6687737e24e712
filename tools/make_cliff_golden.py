#!/usr/bin/env python3
"""Fixture of the any-hit cliff from the REAL reference -> tests/golden/cliff_directional.npz. TEST INFRASTRUCTURE.
tests/scenes/cliff_directional.xml: a directional light shining straight down on an unrotated mesh. The shadow rays a render
would trace — from the camera's first hits toward the light, tmax = 1e30 — traced by the reference's TraceNodeShadow."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import refapi  # noqa: E402

xml = os.path.join(ROOT, "tests", "scenes", "cliff_directional.xml")
ref = refapi.Ref()
ref.set_params(1, 1, 5, True, 1)
ref.load(xml, os.path.join(ROOT, "assets"))
W, H = ref.W, ref.H
pix = np.array([(i, j) for j in range(0, H, 2) for i in range(0, W, 2)], np.int32)
rays = ref.primary_rays(pix, 0)
hits = ref.trace_closest(rays, np.ones(len(rays), np.uint8))
m = hits["hit"] == 1
p = hits["p"][m]
srays = np.concatenate([p, np.tile(np.array([[0, 0, 1]], np.float32), (len(p), 1))], 1).astype(np.float32)
tmax = np.full(len(p), 1e30, np.float32)
occ = ref.trace_shadow(srays, tmax)
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "cliff_directional.npz"), srays=srays, tmax=tmax, occ=occ)
print("cliff fixture: %d shadow rays, %d occluded" % (len(occ), int(occ.sum())))
