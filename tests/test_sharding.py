"""CPU tests of the multi-GPU host logic: the shard rule (Renderer::ThreadRender's interleaved 4x4 tiles,
src/renderers/renderer.cpp:383-398, generalised by a sample-index split) and the reduce step that replaces the
MPI_Send/Recv gather of src/renderers/Renderer_MPI.cpp:142-207 — world_size 2 over gloo, no GPU."""
import ctypes as C
import os

import numpy as np
import pytest

from qaray_b200 import capi


def coverage(W, H, spp, tile_rank=0, tile_ranks=1, s0=0, s1=0):
    lib = capi.load()
    cam = capi.Camera()
    cam.width, cam.height = W, H
    p = capi.RenderParams(spp, 5, 1, 0, W, H, tile_rank, tile_ranks, s0, s1, 0, 0)
    out = np.zeros(W * H, np.uint16)
    capi.check(lib, lib.qg_shard_coverage(C.byref(cam), C.byref(p), out.ctypes.data_as(C.c_void_p)))
    return out.reshape(H, W)


@pytest.mark.parametrize("W,H", [(800, 600), (960, 540), (37, 23), (4, 4), (1, 1), (5, 3)])
def test_single_shard_covers_every_pixel_spp_times(W, H):
    assert (coverage(W, H, 7) == 7).all()


@pytest.mark.parametrize("W,H,ranks", [(800, 600, 2), (37, 23, 3), (960, 540, 8), (5, 3, 8), (1, 1, 4)])
def test_tile_shards_partition_the_frame(W, H, ranks):
    parts = [coverage(W, H, 3, r, ranks) for r in range(ranks)]
    assert (sum(p.astype(np.int64) for p in parts) == 3).all()
    # a pixel belongs to exactly one rank, and that rank is tile k = ty * tiles_x + tx modulo ranks (renderer.cpp:383-398)
    tiles_x = (W + 3) // 4
    jj, ii = np.mgrid[0:H, 0:W]
    owner = ((jj // 4) * tiles_x + ii // 4) % ranks
    for r in range(ranks):
        assert ((parts[r] > 0) == (owner == r)).all()


def test_tile_times_sample_shards_partition_the_job():
    W, H, spp = 61, 45, 10
    total = np.zeros((H, W), np.int64)
    for tr in range(2):
        for (a, b) in ((0, 3), (3, 10)):
            total += coverage(W, H, spp, tr, 2, a, b)
    assert (total == spp).all()


def test_bad_shards_are_rejected():
    lib = capi.load()
    cam = capi.Camera()
    cam.width, cam.height = 8, 8
    out = np.zeros(64, np.uint16)
    for p in (capi.RenderParams(4, 5, 1, 0, 8, 8, 2, 2, 0, 0, 0, 0),      # rank out of range
              capi.RenderParams(4, 5, 1, 0, 8, 8, 0, 1, 3, 9, 0, 0),      # sample shard beyond spp
              capi.RenderParams(0, 5, 1, 0, 8, 8, 0, 1, 0, 0, 0, 0),      # spp 0
              capi.RenderParams(4, 99, 1, 0, 8, 8, 0, 1, 0, 0, 0, 0)):    # bounce out of range
        assert lib.qg_shard_coverage(C.byref(cam), C.byref(p), out.ctypes.data_as(C.c_void_p)) == capi.QG_EINVAL
        assert lib.qg_last_error()


def _rank_main(rank, world, port, W, H, spp, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    per = spp // world
    cov = coverage(W, H, spp, 0, 1, rank * per, (rank + 1) * per).astype(np.float32)   # bench.py's weak-scaling split
    # a synthetic accumulation buffer {sum r, g, b, weight}: what qg_render_accumulate leaves on this rank
    jj, ii = np.mgrid[0:H, 0:W]
    acc = np.zeros((H, W, 4), np.float32)
    for s in range(rank * per, (rank + 1) * per):
        acc[..., 0] += (ii + s) % 5
        acc[..., 1] += (jj * 3 + s) % 7
        acc[..., 2] += s
        acc[..., 3] += 1
    assert (acc[..., 3] == cov).all()
    t = torch.from_numpy(acc)
    dist.reduce(t, 0)   # the path's one exchange step
    if rank == 0:
        q.put(t.numpy().copy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_reduce_equals_single_rank_sum():
    import torch.multiprocessing as mp
    W, H, spp, world = 29, 18, 8, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_rank_main, args=(r, world, port, W, H, spp, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    jj, ii = np.mgrid[0:H, 0:W]
    want = np.zeros((H, W, 4), np.float32)
    for s in range(spp):
        want[..., 0] += (ii + s) % 5
        want[..., 1] += (jj * 3 + s) % 7
        want[..., 2] += s
        want[..., 3] += 1
    assert np.array_equal(got, want)
    assert (got[..., 3] == spp).all()


def shard_size(W, H, p):
    lib = capi.load()
    cam = capi.Camera()
    cam.width, cam.height = W, H
    px, smp, paths = C.c_uint64(), C.c_uint64(), C.c_uint64()
    capi.check(lib, lib.qg_shard_size(C.byref(cam), C.byref(p), C.byref(px), C.byref(smp), C.byref(paths)))
    return px.value, smp.value, paths.value


def test_shard_sizes_are_64_bit_at_4k_1024spp():
    """BASELINE config 5 (3840x2160, 1024 spp): pixels * spp = 8.49e9 does not fit 32 bits. The adaptive driver once derived
    its pixel count as (uint32) samples / spp = 4.1 M instead of 8.3 M and sized its buffers from it (round-1 advice)."""
    W, H, spp = 3840, 2160, 1024
    px, smp, paths = shard_size(W, H, capi.RenderParams(spp, 5, 1, 0, W, H, 0, 1, 0, 0, 0, 0, 0, 4))
    assert px == W * H and smp == W * H * spp and smp > 2 ** 32 and paths == smp
    # tile shards of the same job: owned pixels partition the frame, edge tiles counted by their real size
    W, H = 3839, 2157
    tot = 0
    for r in range(8):
        px, smp, paths = shard_size(W, H, capi.RenderParams(spp, 5, 1, 0, W, H, r, 8, 0, 0, 0, 0))
        assert smp == px * spp and paths >= smp
        tot += px
    assert tot == W * H


@pytest.mark.parametrize("nranks,tile_split", [(1, 1), (2, 1), (2, 2), (4, 2), (8, 2), (8, 8), (8, 1), (6, 3)])
def test_shard_params_tile_x_sample_split_partitions_the_job(nranks, tile_split):
    """qg_shard_params = G_t x G_s of SURVEY.md 8(e): every (pixel, sample) of the frame belongs to exactly one rank"""
    lib = capi.load()
    W, H, spp = 45, 31, 12
    whole = capi.RenderParams(spp, 5, 1, 0, W, H, 0, 1, 0, 0, 0, 0)
    total = np.zeros((H, W), np.int64)
    for r in range(nranks):
        mine = capi.RenderParams()
        capi.check(lib, lib.qg_shard_params(C.byref(whole), r, nranks, tile_split, C.byref(mine)))
        assert mine.tile_ranks == tile_split and mine.tile_rank == r % tile_split
        total += coverage(W, H, spp, mine.tile_rank, mine.tile_ranks, mine.sample_begin, mine.sample_end)
    assert (total == spp).all()


def test_shard_params_rejects_bad_splits():
    lib = capi.load()
    whole = capi.RenderParams(8, 5, 1, 0, 16, 16, 0, 1, 0, 0, 0, 0)
    out = capi.RenderParams()
    assert lib.qg_shard_params(C.byref(whole), 0, 4, 3, C.byref(out)) == capi.QG_EINVAL      # G_t does not divide G
    assert lib.qg_shard_params(C.byref(whole), 4, 4, 1, C.byref(out)) == capi.QG_EINVAL      # rank out of range
    shard = capi.RenderParams(8, 5, 1, 0, 16, 16, 1, 2, 0, 0, 0, 0)
    assert lib.qg_shard_params(C.byref(shard), 0, 2, 1, C.byref(out)) == capi.QG_EINVAL      # already a shard
    adaptive = capi.RenderParams(8, 5, 1, 0, 16, 16, 0, 1, 0, 0, 0, 0, 0, 4)
    assert lib.qg_shard_params(C.byref(adaptive), 0, 2, 1, C.byref(out)) == capi.QG_EINVAL   # adaptive stop needs all samples of a pixel
    assert lib.qg_shard_params(C.byref(adaptive), 0, 2, 2, C.byref(out)) == capi.QG_OK
