"""The any-hit cliff (round-1 verdict, weak point 7): a directional light whose direction has zero components in
a mesh's frame. Every unoccluded shadow ray toward it has degenerate axes, the reference's slab test ignores them
(src/objects/objects.cpp:353-376) and its walk visits every box above the ray's start, accepting rounding-noise hits on the
way — so the exact answer needs that whole walk. tests/scenes/cliff_directional.xml + tests/golden/cliff_directional.npz
(tools/make_cliff_golden.py: the reference's TraceNodeShadow on the shadow rays a render of that scene traces).
Parity: occlusion exact, through the warp-per-ray exact walk (anyexact.cuh) and, with the hand-over queue forced empty,
through the in-lane repeat. Time: the hand-over must not be slower than the in-lane repeat, and a render of the scene stays
within a bound per shadow ray."""
import os
import time

import numpy as np
import pytest

from conftest import ASSETS, GOLDEN_DIR, ROOT
from qaray_b200.renderer import Renderer_B200, RendererParam, Scene

pytestmark = pytest.mark.gpu
XML = os.path.join(ROOT, "tests", "scenes", "cliff_directional.xml")


def trace(monkeypatch, cap):
    if cap is None:
        monkeypatch.delenv("QB_ANYQ_CAP", raising=False)
    else:
        monkeypatch.setenv("QB_ANYQ_CAP", str(cap))
    g = np.load(os.path.join(GOLDEN_DIR, "cliff_directional.npz"))
    sc = Scene.load(XML, ASSETS)
    sc.trace_any(g["srays"][:64], g["tmax"][:64])  # warm up (allocations, module load)
    t0 = time.time()
    occ = sc.trace_any(g["srays"], g["tmax"])
    dt = time.time() - t0
    sc.close()
    return occ, g["occ"], dt


def test_cliff_rays_are_exact_and_the_hand_over_is_not_slower(monkeypatch):
    occ, ref, dt_queue = trace(monkeypatch, None)
    assert np.array_equal(occ, ref), "%d of %d shadow rays differ from the reference" % ((occ != ref).sum(), len(ref))
    occ2, _, dt_lane = trace(monkeypatch, 0)
    assert np.array_equal(occ2, ref), "in-lane repeat: %d rays differ" % (occ2 != ref).sum()
    occ3, _, _ = trace(monkeypatch, 100)  # a queue that overflows: both paths in one launch
    assert np.array_equal(occ3, ref)
    print("cliff: %d rays, hand-over %.1f ms, in-lane %.1f ms" % (len(ref), dt_queue * 1e3, dt_lane * 1e3))
    assert dt_queue <= 1.5 * dt_lane + 0.02
    # the reference needs ~3.5 ms per such ray on one core (68 s for these 19 k rays): a bound three orders below that
    assert dt_queue <= len(ref) * 4e-6 + 0.05


def test_cliff_scene_renders_within_a_time_bound():
    p = RendererParam(spp=4, max_bounce=2)
    r = Renderer_B200(p)
    r.Init(0)
    assert r.LoadScene(XML, ASSETS) == 1
    r.ComputeScene()
    r.Render()
    st = r.Render()
    r.Terminate()
    rays = st.closest_rays + st.shadow_rays
    assert st.shadow_rays > 100000 and np.isfinite(r.linearBuffer).all()
    print("cliff render: %d shadow rays, %.1f ms" % (st.shadow_rays, st.device_ms))
    assert st.device_ms <= st.shadow_rays * 4e-3 * 1e-3 * 1e3 + 50  # 4 us per shadow ray + 50 ms
