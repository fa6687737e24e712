"""GPU tests of the multi-GPU product path (include/qaray_b200.h: qg_shard_params, qgComm, qgMulti; the C++ CLI's -gpus N):
the reference's rank model (src/renderers/Renderer_MPI.cpp:33-55,123-207) over NCCL. What one device can prove is proved
on one device — the shards of a G_t x G_s split, rendered through the same entry points the ranks use, sum to the whole
frame; qgMulti over the devices present equals qg_render — and the N >= 2 cases run when the box has the GPUs."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest
from PIL import Image

from conftest import ASSETS, ROOT, scene_xml
from qaray_b200 import capi

pytestmark = pytest.mark.gpu

CLI = os.path.join(ROOT, "qaray_b200", "qaray_b200_cli")


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def device_count():
    import torch
    return torch.cuda.device_count()


def load(lib, name):
    h = C.c_void_p()
    capi.check(lib, lib.qg_scene_load_xml(os.fsencode(scene_xml(name)), os.fsencode(ASSETS), C.byref(h)))
    info = capi.SceneInfo()
    capi.check(lib, lib.qg_scene_info(h, C.byref(info)))
    return h, info.camera.width, info.camera.height


def accumulate(lib, h, p, W, H):
    import torch
    acc = torch.zeros(W * H * 4, dtype=torch.float32, device="cuda")
    dep = torch.zeros(W * H, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    st = capi.RenderStats()
    capi.check(lib, lib.qg_render_accumulate(h, C.byref(p), C.c_void_p(acc.data_ptr()), C.c_void_p(dep.data_ptr()), None, C.byref(st)))
    return acc.cpu().numpy().reshape(H, W, 4), dep.cpu().numpy().reshape(H, W), st


@pytest.mark.parametrize("name,nranks,tile_split", [("example_project12_caustics_glossy", 2, 1), ("example_project7_object", 4, 2),
                                                    ("trc_scene_tower", 8, 2), ("example_project3_sphere", 3, 3)])
def test_rank_shards_sum_to_the_whole_frame(name, nranks, tile_split):
    """bench.py's split (sample ranges) and SURVEY 8(e)'s tile x sample split: the ranks' accumulation buffers added up — what
    the ncclReduce does — equal the single-GPU buffer up to the order of the fp32 additions; weights and depth exactly."""
    lib = capi.load()
    capi.check(lib, lib.qg_init(0))
    h, W, H = load(lib, name)
    if name == "trc_scene_tower":
        W, H = 240, 135
    spp = 12
    whole = capi.RenderParams(spp, 5, 1, 0x5EED, W, H, 0, 1, 0, 0, 0, 0)
    ref_acc, ref_dep, ref_st = accumulate(lib, h, whole, W, H)
    acc = np.zeros_like(ref_acc, dtype=np.float64)
    dep = np.zeros_like(ref_dep)
    rays = 0
    for r in range(nranks):
        mine = capi.RenderParams()
        capi.check(lib, lib.qg_shard_params(C.byref(whole), r, nranks, tile_split, C.byref(mine)))
        a, d, st = accumulate(lib, h, mine, W, H)
        acc += a
        dep += d
        rays += st.closest_rays + st.shadow_rays
    lib.qg_scene_destroy(h)
    assert np.array_equal(acc[..., 3], ref_acc[..., 3]) and (ref_acc[..., 3] == spp).all()
    assert np.array_equal(dep, ref_dep)
    assert rays == ref_st.closest_rays + ref_st.shadow_rays   # same paths, same rays: the RNG is keyed by (pixel, sample)
    assert np.allclose(acc[..., :3], ref_acc[..., :3], rtol=2e-5, atol=1e-5 * spp)


def build_maps(lib, h, sizes=(5000, 300)):
    pp = capi.PhotonParams.defaults(0x5EED)
    pp.photon_map_size, pp.caustics_map_size = sizes
    capi.check(lib, lib.qg_photon_maps_build(h, C.byref(pp), None))


def test_photon_mapped_shards_sum_to_the_whole_frame():
    """-use-photon-map under the rank model (the reference's own "TODO: Photon Map for MPI Mode", renderer.cpp:116): the maps are
    a function of the seed, so every rank builds the same ones; the gather of a path depends on its (pixel, sample) stream only,
    so the shards of a tile x sample split still add up to the single-GPU frame."""
    lib = capi.load()
    capi.check(lib, lib.qg_init(0))
    name, spp, nranks, tile_split = "trc_mtl_glass", 8, 4, 2
    h, W, H = load(lib, name)
    W, H = W // 4, H // 4
    build_maps(lib, h)
    whole = capi.RenderParams(spp, 5, 1, 0x5EED, W, H, 0, 1, 0, 0, 0, 0)
    whole.use_photon_map = 1
    ref_acc, ref_dep, ref_st = accumulate(lib, h, whole, W, H)
    # the ranks' scenes are replicas that built their own maps: a second scene stands in for them
    h2, _, _ = load(lib, name)
    build_maps(lib, h2)
    acc = np.zeros_like(ref_acc, dtype=np.float64)
    rays = 0
    for r in range(nranks):
        mine = capi.RenderParams()
        capi.check(lib, lib.qg_shard_params(C.byref(whole), r, nranks, tile_split, C.byref(mine)))
        assert mine.use_photon_map == 1
        a, d, st = accumulate(lib, h2, mine, W, H)
        acc += a
        rays += st.closest_rays + st.shadow_rays
    lib.qg_scene_destroy(h)
    lib.qg_scene_destroy(h2)
    assert np.array_equal(acc[..., 3], ref_acc[..., 3])
    assert rays == ref_st.closest_rays + ref_st.shadow_rays
    assert ref_acc[..., :3].sum() > 0
    assert np.allclose(acc[..., :3], ref_acc[..., :3], rtol=1e-4, atol=1e-5 * spp)


def test_cli_gpus_n_with_photon_maps(tmp_path):
    if device_count() < 2:
        pytest.skip("needs 2 GPUs")
    pre1, pre2 = str(tmp_path) + os.sep + "one_", str(tmp_path) + os.sep + "two_"
    common = ["-batch", "-spp", "8", "-use-photon-map", "-assets", ASSETS]
    out1 = run_cli(common + ["-out", pre1, scene_xml("trc_mtl_glass")], str(tmp_path))
    out2 = run_cli(common + ["-gpus", "2", "-out", pre2, scene_xml("trc_mtl_glass")], str(tmp_path))
    assert "Photon Map: 10000 photons" in out1 and "Photon Map: 10000 photons" in out2
    a = np.asarray(Image.open(pre1 + "colorBuffer.png").convert("RGB")).astype(np.int32)
    b = np.asarray(Image.open(pre2 + "colorBuffer_MPI.png").convert("RGB")).astype(np.int32)
    assert np.abs(a - b).max() <= 1


def multi_render(lib, name, ndev, tile_split, spp, prefix=None):
    m = C.c_void_p()
    capi.check(lib, lib.qg_multi_create(None, ndev, C.byref(m)))
    capi.check(lib, lib.qg_multi_load_xml(m, os.fsencode(scene_xml(name)), os.fsencode(ASSETS)))
    info = capi.SceneInfo()
    capi.check(lib, lib.qg_scene_info(lib.qg_multi_scene(m, 0), C.byref(info)))
    W, H = info.camera.width, info.camera.height
    lin, c8 = np.zeros((H, W, 3), np.float32), np.zeros((H, W, 3), np.uint8)
    dep, sc, mask = np.zeros((H, W), np.float32), np.zeros((H, W), np.uint8), np.zeros((H, W), np.uint8)
    st = capi.RenderStats()
    p = capi.RenderParams(spp, 5, 1, 0x5EED, W, H, 0, 1, 0, 0, 0, 0)
    capi.check(lib, lib.qg_multi_render(m, C.byref(p), tile_split, os.fsencode(prefix) if prefix else None, _p(lin), _p(c8), _p(dep), _p(sc),
                                        _p(mask), C.byref(st)))
    lib.qg_multi_destroy(m)
    return lin, c8, dep, sc, mask, st


def single_render(lib, name, spp):
    h, W, H = load(lib, name)
    lin, c8 = np.zeros((H, W, 3), np.float32), np.zeros((H, W, 3), np.uint8)
    dep, sc, mask = np.zeros((H, W), np.float32), np.zeros((H, W), np.uint8), np.zeros((H, W), np.uint8)
    st = capi.RenderStats()
    p = capi.RenderParams(spp, 5, 1, 0x5EED, W, H, 0, 1, 0, 0, 0, 0)
    capi.check(lib, lib.qg_render(h, C.byref(p), _p(lin), _p(c8), _p(dep), _p(sc), _p(mask), C.byref(st)))
    lib.qg_scene_destroy(h)
    return lin, c8, dep, sc, mask, st


@pytest.mark.parametrize("ndev,tile_split", [(1, 1), (2, 1), (2, 2), (4, 2), (8, 2)])
def test_qg_multi_render_equals_qg_render(ndev, tile_split, tmp_path):
    if device_count() < ndev:
        pytest.skip("needs %d GPUs" % ndev)
    lib = capi.load()
    capi.check(lib, lib.qg_init(0))
    name, spp = "example_project12_caustics_glossy", 16
    one = single_render(lib, name, spp)
    prefix = str(tmp_path) + os.sep
    many = multi_render(lib, name, ndev, tile_split, spp, prefix)
    assert np.allclose(many[0], one[0], rtol=2e-5, atol=1e-5)
    assert np.abs(many[1].astype(np.int32) - one[1].astype(np.int32)).max() <= 1
    for k in (2, 3, 4):
        assert np.array_equal(many[k], one[k])
    assert many[5].closest_rays + many[5].shadow_rays == one[5].closest_rays + one[5].shadow_rays
    # the per-rank planes Renderer_MPI::Render saves before the gather (Renderer_MPI.cpp:132-138)
    for r in range(ndev):
        for plane in ("color", "depth", "sample"):
            assert os.path.exists("%srank_%d_%sBuffer.png" % (prefix, r, plane))
    if ndev == 1:
        # one rank: its plane IS the frame (the same accumulation buffer resolved twice; two renders differ by the order of
        # the fp32 atomic additions, hence `many`, not `one`)
        assert np.array_equal(np.asarray(Image.open(prefix + "rank_0_colorBuffer.png").convert("RGB")), many[1])


def run_cli(args, cwd):
    r = subprocess.run([CLI] + args, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout
    return r.stdout


def test_cli_writes_the_reference_output_set(tmp_path):
    """main.cpp-style CLI (qaray_b200/csrc/cli_main.cpp over renderer_b200.h — the C++ host of the boundary): its three PNGs
    decode to the planes the library returns (colour exactly; depth and sample count through the reference's normalisations,
    framebuffer.cpp:62-107)."""
    if not os.path.exists(CLI):
        pytest.fail("qaray_b200_cli is missing: run __graft_entry__.build()")
    lib = capi.load()
    capi.check(lib, lib.qg_init(0))
    name, spp = "example_project3_sphere", 8
    out = run_cli(["-batch", "-spp", str(spp), "-assets", ASSETS, "-out", str(tmp_path) + os.sep, scene_xml(name)], str(tmp_path))
    assert "Elapsed Time" in out
    one = single_render(lib, name, spp)
    col = np.asarray(Image.open(str(tmp_path / "colorBuffer.png")).convert("RGB")).astype(np.int32)
    assert np.abs(col - one[1].astype(np.int32)).max() <= 1   # two renders: the fp32 atomic additions come in a different order
    z = one[2]
    hit = z != np.float32(1e30)
    zmin, zmax = z[hit].min(), z[hit].max()
    want = np.where(hit, ((zmax - np.where(hit, z, zmax)) / (zmax - zmin) * 255).astype(np.int32), 0)
    dep = np.asarray(Image.open(str(tmp_path / "depthBuffer.png")).convert("L")).astype(np.int32)
    assert np.abs(dep - want).max() <= 1
    smp = np.asarray(Image.open(str(tmp_path / "sampleBuffer.png")).convert("L"))
    assert (smp == 0).all()   # fixed spp: every pixel took the same number of samples (framebuffer.cpp:97-98)
    # adaptive sampling: the sample-count image is stretched over [smin, smax]
    run_cli(["-batch", "-sppMin", "4", "-sppMax", "32", "-assets", ASSETS, "-out", str(tmp_path) + os.sep + "a_", scene_xml(name)], str(tmp_path))
    smp = np.asarray(Image.open(str(tmp_path / "a_sampleBuffer.png")).convert("L"))
    assert smp.min() == 0 and smp.max() == 255


def test_cli_gpus_n_composites_like_renderer_mpi(tmp_path):
    if device_count() < 2:
        pytest.skip("needs 2 GPUs")
    lib = capi.load()
    capi.check(lib, lib.qg_init(0))
    name, spp = "example_project7_object", 8
    pre = str(tmp_path) + os.sep
    run_cli(["-batch", "-spp", str(spp), "-gpus", "2", "-tilesplit", "2", "-rank-images", "-assets", ASSETS, "-out", pre, scene_xml(name)], str(tmp_path))
    one = single_render(lib, name, spp)
    col = np.asarray(Image.open(pre + "colorBuffer_MPI.png").convert("RGB")).astype(np.int32)
    assert np.abs(col - one[1].astype(np.int32)).max() <= 1
    assert os.path.exists(pre + "rank_1_colorBuffer.png") and os.path.exists(pre + "depthBuffer_MPI.png")
