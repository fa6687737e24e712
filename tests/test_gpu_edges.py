"""GPU edge cases through the C ABI.

* tests/golden/edge_<scene>.npz (tools/make_edge_golden.py, the reference itself): rays with exactly-zero, negative-zero
  and around-1e-7 direction components (the reference's slab tests ignore such an axis; the CUDA path's degenerate-axis
  cull must not change any result), rays starting on geometry, shadow rays whose tmax ends just before / after a
  surface. Bar: hit/miss, node, face, front flag exact and t bit-equal; occlusion exact.
* empty and ragged inputs: zero rays, a scene without objects or lights, 1x1 and 5x3 frames, max_bounce 0.
"""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import ASSETS, GOLDEN_DIR, load_golden, rays18_to_struct, scene_xml
from qaray_b200 import capi
from qaray_b200.renderer import Renderer_B200, RendererParam, Scene

pytestmark = pytest.mark.gpu


def edge_names():
    if not os.path.isdir(GOLDEN_DIR):
        return []
    return sorted(f[5:-4] for f in os.listdir(GOLDEN_DIR) if f.startswith("edge_") and f.endswith(".npz"))


@pytest.mark.parametrize("name", edge_names())
def test_degenerate_direction_rays_match_reference(name):
    g = load_golden("edge_" + name)
    sc = Scene.load(scene_xml(name), ASSETS)
    out = sc.trace_closest(rays18_to_struct(g["rays"], np.zeros(len(g["rays"]), np.int32)))
    ref = g["hits"]
    m = ref["hit"] == 1
    assert np.array_equal(out["hit"], ref["hit"])
    assert np.array_equal(out["node"][m], ref["node"][m]) and np.array_equal(out["face"][m], ref["face"][m])
    assert np.array_equal(out["front"][m], ref["front"][m])
    assert np.array_equal(out["z"][m].view(np.uint32), ref["z"][m].view(np.uint32)), "hit distance not bit-equal"
    occ = sc.trace_any(g["rays"][:, :6], g["tmax"])
    assert np.array_equal(occ, g["occ"])
    sc.close()


@pytest.mark.parametrize("cap", [0, 1, 7])
def test_cull_check_queue_overflow_stays_exact(cap, monkeypatch):
    """QB_REDO_CAP forces the deferred cull-check queue (deviation 1) to overflow: pairs that do not fit are
    repeated exactly inside their lane, in the first pass and in the redo pass alike (round-1 advice: a ray with one
    queued and one overflowed pair must not lose the overflowed pair's exact result)."""
    monkeypatch.setenv("QB_REDO_CAP", str(cap))
    name = "trc_scene_tower"
    g = load_golden("edge_" + name)
    sc = Scene.load(scene_xml(name), ASSETS)
    out = sc.trace_closest(rays18_to_struct(g["rays"], np.zeros(len(g["rays"]), np.int32)))
    sc.close()
    ref = g["hits"]
    m = ref["hit"] == 1
    assert np.array_equal(out["hit"], ref["hit"])
    assert np.array_equal(out["node"][m], ref["node"][m]) and np.array_equal(out["face"][m], ref["face"][m])
    assert np.array_equal(out["z"][m].view(np.uint32), ref["z"][m].view(np.uint32)), "hit distance not bit-equal"


def test_zero_rays_and_zero_pixels_are_ok():
    sc = Scene.load(scene_xml("example_project3_sphere"), ASSETS)
    assert len(sc.trace_closest(np.zeros(0, capi.RAY_DTYPE))) == 0
    assert len(sc.trace_any(np.zeros((0, 6), np.float32), np.zeros(0, np.float32))) == 0
    rgb, z = sc.sample_radiance(RendererParam(spp=1), np.zeros((0, 3), np.int32))
    assert rgb.shape == (0, 3)
    sc.close()


EMPTY_XML = """<xml><scene><background r="0.25" g="0.5" b="0.75"/></scene>
<camera><position x="0" y="0" z="5"/><target x="0" y="0" z="0"/><up x="0" y="1" z="0"/><fov value="40"/>
<width value="%d"/><height value="%d"/></camera></xml>"""

NOLIGHT_XML = """<xml><scene><background r="0.1" g="0.1" b="0.1"/>
<object type="sphere" name="s" material="m"/>
<material type="blinn" name="m"><diffuse r="1" g="0" b="0"/><emission r="0.5" g="0.25" b="0"/></material></scene>
<camera><position x="0" y="0" z="5"/><target x="0" y="0" z="0"/><up x="0" y="1" z="0"/><fov value="40"/>
<width value="%d"/><height value="%d"/></camera></xml>"""


def render_xml(tmp_path, text, W, H, spp=2, bounce=5, cls=0):
    f = tmp_path / "s.xml"
    f.write_text(text % (W, H))
    p = RendererParam(spp=spp, max_bounce=bounce)
    p.material_class = cls
    r = Renderer_B200(p)
    r.Init(0)
    assert r.LoadScene(str(f), str(tmp_path)) == 1
    r.ComputeScene()
    st = r.Render()
    out = (r.linearBuffer.copy(), r.depthBuffer.copy(), r.maskBuffer.copy(), r.colorBuffer.copy(), st)
    r.Terminate()
    return out


@pytest.mark.parametrize("W,H", [(1, 1), (5, 3), (37, 23)])
def test_scene_without_objects_renders_the_background(tmp_path, W, H):
    lin, depth, mask, rgb8, st = render_xml(tmp_path, EMPTY_XML, W, H)
    assert mask.all() and lin.shape == (H, W, 3)
    assert np.allclose(lin, np.array([0.25, 0.5, 0.75], np.float32), atol=1e-6)
    assert (depth == np.float32(1e30)).all()
    assert st.samples == W * H * 2 and st.shadow_rays == 0


@pytest.mark.parametrize("cls", [0, 1, 2, 3, 4])
def test_scene_without_lights_and_bounce_zero(tmp_path, cls):
    lin, depth, mask, rgb8, st = render_xml(tmp_path, NOLIGHT_XML, 33, 21, spp=1, bounce=0, cls=cls)
    assert mask.all() and np.isfinite(lin).all() and st.shadow_rays == 0
    hit = depth < 1e29
    assert hit.any() and (~hit).any()
    assert np.allclose(lin[~hit], 0.1, atol=1e-6)
    # the Whitted-style Basic classes have no emission term (MtlBlinn_Basic.cpp:36); the others add it
    want = np.array([0.0, 0.0, 0.0] if cls in (3, 4) else [0.5, 0.25, 0.0], np.float32)
    assert np.allclose(lin[hit], want, atol=1e-6)
