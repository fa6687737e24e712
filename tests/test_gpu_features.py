"""GPU parity of the sampling features the five BASELINE configs do not use: depth of field,
area lights, adaptive sampling. Fixtures: tests/golden/feat_<feature>_<scene>.npz (tools/make_feature_golden.py, the
reference itself). All three are stochastic in the reference (time-seeded RNG), so image parity is statistical:
RMSE(gpu, a) <= 1.25 * RMSE(a, b) on clamped linear radiance, frame means within 4 standard errors.

* dof       additionally single samples with the counter-based stream: the lens sample is the sample's first two random
            numbers (scene.cpp:104-111), so hit distance and radiance must match sample for sample
* adaptive  additionally the per-pixel sample counts: mean count within 10 % of the reference's, count planes as close
            to reference run A as reference run B is (x1.5), and the sample-count plane = 255 * s / sppMax
"""
import os

import numpy as np
import pytest

from conftest import ASSETS, GOLDEN_DIR, load_golden, scene_xml
from qaray_b200.renderer import Renderer_B200, RendererParam, Scene

pytestmark = pytest.mark.gpu


def render(name, W, H, spp_min, spp_max, seed=0x5EED, rank=0, ranks=1):
    p = RendererParam(spp=spp_max, max_bounce=5, srgb=True, seed=seed)
    p.sppMin = spp_min
    p.width, p.height = W, H
    r = Renderer_B200(p)
    r.Init(0)
    assert r.LoadScene(scene_xml(name), ASSETS) == 1
    r.mpiRank, r.mpiSize = rank, ranks
    r.ComputeScene()
    st = r.Render()
    out = (r.linearBuffer.copy(), r.depthBuffer.copy(), r.sampleCountBuffer.copy(), r.maskBuffer.copy(), st)
    r.Terminate()
    return out


def check_image(lin, g):
    a, b = np.clip(g["ref_a"], 0, 1), np.clip(g["ref_b"], 0, 1)
    m = np.clip(lin, 0, 1)
    floor = float(np.sqrt(((a - b) ** 2).mean()))
    err = float(np.sqrt(((m - a) ** 2).mean()))
    assert err <= 1.25 * floor + 1e-4, "RMSE(gpu, ref_a) %.5f vs reference noise floor %.5f" % (err, floor)
    se = float((a - b).std() / np.sqrt(a.size))
    assert abs(float(m.mean()) - float(a.mean())) <= 4 * se + 1e-4, (m.mean(), a.mean(), se)


def test_depth_of_field_image():
    g = load_golden("feat_dof_example_project9")
    W, H, smin, smax = [int(v) for v in g["size"]]
    lin, depth, sc, mask, st = render("example_project9", W, H, smin, smax)
    assert mask.all() and np.isfinite(lin).all()
    check_image(lin, g)


def test_depth_of_field_single_samples():
    g = load_golden("feat_dof_example_project9")
    W, H = int(g["size"][0]), int(g["size"][1])
    sc = Scene.load(scene_xml("example_project9"), ASSETS)
    p = RendererParam(spp=64, seed=int(g["rad_seed"][0]))
    p.width, p.height = W, H
    rgb, z = sc.sample_radiance(p, g["rad_pix"])
    sc.close()
    ref, rz = g["rad_rgb"], g["rad_z"]
    # the lens point goes through sinf/cosf/sqrtf, so hit distances agree to rounding, not to the bit
    hit = rz < 1e29
    assert np.array_equal(z < 1e29, hit)
    assert np.abs(z[hit] - rz[hit]).max() <= 1e-4 * np.abs(rz[hit]).max()
    bad = (np.abs(rgb - ref) > 2e-5 + 1e-3 * np.abs(ref)).any(axis=1).sum()
    assert bad <= max(1, len(rgb) // 100), "%d of %d radiance samples differ" % (bad, len(rgb))


def test_area_light_image():
    g = load_golden("feat_area_example_project10_test")
    W, H, smin, smax = [int(v) for v in g["size"]]
    lin, depth, sc, mask, st = render("example_project10_test", W, H, smin, smax)
    assert mask.all() and np.isfinite(lin).all()
    check_image(lin, g)
    assert np.array_equal(depth.view(np.uint32), g["depth"].view(np.uint32))
    # 64 shadow samples per (hit, area light) group on top of the point-light rays
    assert st.shadow_rays >= 64 * W * H


@pytest.mark.parametrize("name", ["example_project3_sphere", "example_project12_caustics_glossy"])
def test_adaptive_sampling(name):
    g = load_golden("feat_adaptive_" + name)
    W, H, smin, smax = [int(v) for v in g["size"]]
    lin, depth, sc, mask, st = render(name, W, H, smin, smax)
    assert mask.all() and np.isfinite(lin).all()
    check_image(lin, g)
    assert np.array_equal(depth.view(np.uint32), g["depth"].view(np.uint32))
    ca, cb = g["count_a"].astype(np.float64), g["count_b"].astype(np.float64)
    # the 8-bit plane is 255 * s / sppMax truncated (renderer.cpp:362-363): recover s to within one count
    s_gpu = sc.astype(np.float64) * smax / 255.0
    assert s_gpu.min() >= smin - 1 and s_gpu.max() <= smax
    assert abs(s_gpu.mean() - ca.mean()) <= 0.10 * ca.mean() + 0.5, (s_gpu.mean(), ca.mean(), cb.mean())
    floor = float(np.sqrt(((ca - cb) ** 2).mean()))
    err = float(np.sqrt(((s_gpu - ca) ** 2).mean()))
    assert err <= 1.5 * floor + 1.0, "sample-count plane: rms(gpu, a) %.3f vs rms(a, b) %.3f" % (err, floor)
    # pixels that stop at sppMin in both reference runs (mostly flat, noise-free regions) mostly stop there on the GPU
    # too — "mostly" because a third independent run of a stochastic pixel need not repeat the first two
    flat = (ca == smin) & (cb == smin)
    if flat.any():
        assert (np.abs(s_gpu[flat] - smin) <= 1).mean() >= 0.90


def test_adaptive_tile_shards_union_equals_whole_frame():
    name = "example_project3_sphere"
    whole = render(name, 96, 72, 4, 32)
    acc = np.zeros_like(whole[0])
    cover = np.zeros(whole[3].shape, np.int32)
    for r in range(2):
        lin, depth, sc, mask, st = render(name, 96, 72, 4, 32, rank=r, ranks=2)
        acc += lin * mask[..., None]
        cover += mask
        assert np.array_equal(sc[mask == 1], whole[2][mask == 1])
    assert (cover == 1).all()
    assert np.allclose(acc, whole[0], rtol=1e-5, atol=1e-6)


FUSED_WORKER = r'''
import os, sys
import numpy as np
sys.path.insert(0, %(root)r)
sys.path.insert(0, os.path.join(%(root)r, "tests"))
from conftest import ASSETS, scene_xml
from qaray_b200.renderer import RendererParam, Scene
name, out = sys.argv[1], sys.argv[2]
sc = Scene.load(scene_xml(name), ASSETS)
info = sc.info()
W, H = info.camera.width, info.camera.height
rng = np.random.default_rng(7)
n = 20000
pix = np.stack([rng.integers(0, W, n), rng.integers(0, H, n), rng.integers(0, 64, n)], 1).astype(np.int32)
rgb, z = sc.sample_radiance(RendererParam(spp=64, max_bounce=5, srgb=True, seed=0x5EED), pix)
sc.close()
np.savez(out, rgb=rgb, z=z)
'''


@pytest.mark.parametrize("name", ["trc_scene_tower", "trc_scene_xmas", "example_project7_object"])
def test_fused_shade_gives_the_same_samples_as_the_separate_attribute_pass(name, tmp_path):
    """k_wf_shade<true> computes the hit attributes itself at depths >= 1; QB_FUSED_FINISH=0 brings the separate
    k_wf_finish pass back for every depth. Single samples (one accumulation slot each, so no atomic order is involved) must come
    out the same: position, normal, texture coordinates and footprints are the same arithmetic either way. Distances bit for bit;
    radiance up to the order in which the shadow kernel's lanes add the lights of one hit into the sample's slot (fp32 atomics:
    an ulp or two when two or more lights are visible), which differs from run to run in either mode."""
    import subprocess
    import sys
    from conftest import ROOT
    res = []
    for fused in ("1", "0"):
        out = str(tmp_path / ("fused%s.npz" % fused))
        env = dict(os.environ, QB_FUSED_FINISH=fused)
        r = subprocess.run([sys.executable, "-c", FUSED_WORKER % {"root": ROOT}, name, out], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                           text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        res.append(np.load(out))
    a, b = res
    assert np.isfinite(a["rgb"]).mean() > 0.99 and a["rgb"].sum() > 0
    assert np.array_equal(a["z"].view(np.uint32), b["z"].view(np.uint32))
    fin = np.isfinite(a["rgb"]).all(axis=1)
    assert np.array_equal(fin, np.isfinite(b["rgb"]).all(axis=1))
    same = (a["rgb"].view(np.uint32) == b["rgb"].view(np.uint32)).all(axis=1)
    assert same.mean() >= 0.9, "only %.1f %% of the samples are bit-equal" % (100 * same.mean())
    close = np.abs(a["rgb"][fin] - b["rgb"][fin]) <= 1e-7 + 4e-6 * np.abs(b["rgb"][fin])
    assert close.all(), "%d samples differ between the fused and the separate pass beyond the order of additions" % (~close.all(axis=1)).sum()
