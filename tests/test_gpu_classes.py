"""GPU parity of the material classes the reference's `MtlBlinn` alias selects (src/materials/materials.h:57-61):
MtlBlinn_PathTracing, MtlBlinn_MonteCarloGI, MtlBlinn_Basic, MtlPhong_Basic — qgRenderParams.material_class.

Fixtures: tests/golden/cls_<class>_<scene>.npz from tools/make_class_golden.py = the reference itself compiled with
that alias (oracle/_ref/libqaref_parity_<class>.so).

* deterministic fixtures (`ref`; Whitted-style classes on scenes without glossy materials draw no random number):
  per-pixel comparison of the linear means, tolerance 2e-3 absolute + 2e-3 relative; at most 0.5 % of the pixels may
  differ more (a secondary ray grazing a silhouette can flip between fmad and non-fmad arithmetic of the two builds).
* stochastic fixtures (`ref_a`, `ref_b`): RMSE(gpu, a) <= 1.25 * RMSE(a, b) on clamped linear radiance and frame means
  within 4 standard errors, as for the default class (tests/test_gpu_image.py).
* MtlBlinn_PathTracing single samples with the counter-based stream (`rad_*`): 2e-4 relative + 2e-5 absolute, at most
  0.5 % of the samples may branch differently at a selection threshold.
"""
import os

import numpy as np
import pytest

from conftest import ASSETS, CLASS_ID, class_golden_cases as class_cases, load_golden, scene_xml
from qaray_b200.renderer import Renderer_B200, RendererParam, Scene

pytestmark = pytest.mark.gpu

def render(name, cls, spp, width, height, seed=0x5EED):
    p = RendererParam(spp=spp, max_bounce=5, srgb=True, seed=seed)
    p.width, p.height = width, height
    p.material_class = CLASS_ID[cls]
    r = Renderer_B200(p)
    r.Init(0)
    assert r.LoadScene(scene_xml(name), ASSETS) == 1
    r.ComputeScene()
    st = r.Render()
    out = (r.linearBuffer.copy(), r.depthBuffer.copy(), r.maskBuffer.copy(), st)
    r.Terminate()
    return out


@pytest.mark.parametrize("cls,name", class_cases())
def test_class_image_matches_reference(cls, name):
    g = load_golden("cls_%s_%s" % (cls, name))
    W, H, spp = [int(v) for v in g["size"]]
    lin, depth, mask, st = render(name, cls, spp, W, H)
    assert mask.all() and np.isfinite(lin).all()
    assert np.array_equal(depth.view(np.uint32), g["depth"].view(np.uint32))
    if "ref" in g.files:
        ref = g["ref"]
        bad = (np.abs(lin - ref) > 2e-3 + 2e-3 * np.abs(ref)).any(axis=2)
        assert bad.mean() <= 0.005, "%d of %d pixels differ from the deterministic reference image (max %.4f)" % (
            bad.sum(), bad.size, float(np.abs(lin - ref).max()))
        assert abs(float(lin.mean()) - float(ref.mean())) <= 1e-3 * max(1.0, float(ref.mean()))
    else:
        a, b = np.clip(g["ref_a"], 0, 1), np.clip(g["ref_b"], 0, 1)
        m = np.clip(lin, 0, 1)
        floor = float(np.sqrt(((a - b) ** 2).mean()))
        err = float(np.sqrt(((m - a) ** 2).mean()))
        assert err <= 1.25 * floor + 1e-4, "RMSE(gpu, ref_a) %.5f vs reference noise floor %.5f" % (err, floor)
        se = float((a - b).std() / np.sqrt(a.size))
        assert abs(float(m.mean()) - float(a.mean())) <= 4 * se + 1e-4, (m.mean(), a.mean(), se)


@pytest.mark.parametrize("cls,name", [c for c in class_cases() if c[0] == "pathtracing"])
def test_pathtracing_single_samples_match_reference(cls, name):
    g = load_golden("cls_%s_%s" % (cls, name))
    W, H, spp = [int(v) for v in g["size"]]
    sc = Scene.load(scene_xml(name), ASSETS)
    p = RendererParam(spp=64, seed=int(g["rad_seed"][0]))
    p.width, p.height = W, H
    p.material_class = CLASS_ID[cls]
    rgb, z = sc.sample_radiance(p, g["rad_pix"])
    sc.close()
    ref = g["rad_rgb"]
    bad = (np.abs(rgb - ref) > 2e-5 + 2e-4 * np.abs(ref)).any(axis=1).sum()
    assert bad <= max(1, len(rgb) // 200), "%d of %d radiance samples differ" % (bad, len(rgb))


def test_branching_classes_are_shard_and_wave_invariant():
    """the union over tile shards and a small wave size give the same image as one big wave (per-path RNG streams)"""
    name, cls = "example_project3_sphere", "montecarlogi"
    whole = render(name, cls, 2, 96, 72)[0]
    p = RendererParam(spp=2, seed=0x5EED)
    p.width, p.height, p.material_class, p.batch_paths = 96, 72, CLASS_ID[cls], 4096
    r = Renderer_B200(p)
    r.Init(0)
    assert r.LoadScene(scene_xml(name), ASSETS) == 1
    r.ComputeScene()
    r.Render()
    small = r.linearBuffer.copy()
    r.Terminate()
    assert np.allclose(whole, small, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("cls", ["photonmap", "pathtracing", "blinn_basic"])
def test_material_sorted_shading_gives_the_same_image(cls):
    """qgRenderParams.sort_by_material only reorders the shade queue: per-path RNG streams and per-pixel sums (float atomics
    apart) are unchanged, on a scene with several materials and a MultiMtl"""
    name = "example_project7_object"
    outs = []
    for sort in (0, 1):
        p = RendererParam(spp=4, max_bounce=5, seed=0x5EED)
        p.width, p.height, p.material_class, p.sort_by_material = 160, 120, CLASS_ID.get(cls, 0), sort
        r = Renderer_B200(p)
        r.Init(0)
        assert r.LoadScene(scene_xml(name), ASSETS) == 1
        r.ComputeScene()
        st = r.Render()
        outs.append((r.linearBuffer.copy(), r.depthBuffer.copy(), st.closest_rays, st.shadow_rays))
        r.Terminate()
    assert np.array_equal(outs[0][1].view(np.uint32), outs[1][1].view(np.uint32))
    assert outs[0][2:] == outs[1][2:]
    assert np.allclose(outs[0][0], outs[1][0], rtol=1e-4, atol=1e-5)
