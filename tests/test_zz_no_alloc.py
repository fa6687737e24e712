"""qg_trace_any_noalloc (include/qaray_b200.h): the any-hit batch for callers that cannot call cudaMalloc — device buffers in and
out, the launch's scratch in a static device array of the library. No counterpart in the reference.
Checked through the C ABI against qg_trace_any (which itself is pinned to the reference's
TraceNodeShadow, tests/test_gpu_parity.py, tests/test_gpu_cliff.py) on two calls in a row: occlusion is a byte per ray and must
be equal, as in every any-hit test here. The cliff fixture's rays go through the hand-over queue of the exact walk — the part
of the scratch whose stale contents a second call would see.

Collected last on purpose."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import ASSETS, GOLDEN_DIR, ROOT, load_golden, scene_xml
from qaray_b200 import capi

pytestmark = pytest.mark.gpu
MAX_RAYS = 1 << 20  # QG_NOALLOC_MAX_RAYS


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def open_scene(xml):
    lib = capi.load()
    capi.check(lib, lib.qg_init(0))
    h = C.c_void_p()
    capi.check(lib, lib.qg_scene_load_xml(os.fsencode(xml), os.fsencode(ASSETS), C.byref(h)))
    return lib, h


def with_alloc(lib, h, rays6, tmax):
    occ = np.zeros(len(rays6), np.uint8)
    capi.check(lib, lib.qg_trace_any(h, _p(rays6), _p(tmax), len(rays6), _p(occ)))
    return occ


def without_alloc(lib, h, torch, stream, d_rays, d_tmax, d_occ, rays6, tmax):
    """the caller owns inputs and outputs: preallocated device buffers, filled here, the call itself on `stream`"""
    n = len(rays6)
    d_rays[:6 * n].copy_(torch.from_numpy(rays6.reshape(-1)))
    d_tmax[:n].copy_(torch.from_numpy(tmax))
    d_occ.fill_(7)  # neither 0 nor 1: every byte below n must be written
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    capi.check(lib, lib.qg_trace_any_noalloc(h, C.c_void_p(d_rays.data_ptr()), C.c_void_p(d_tmax.data_ptr()), n, C.c_void_p(d_occ.data_ptr()),
                                             C.c_void_p(stream.cuda_stream)))
    stream.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    out = d_occ.cpu().numpy()
    assert (out[n:] == 7).all(), "wrote past the n rays of the call"
    return out[:n].copy(), free0, free1


def test_no_alloc_any_hit_equals_the_allocating_entry_point_twice_in_a_row():
    torch = pytest.importorskip("torch")
    g = np.load(os.path.join(GOLDEN_DIR, "cliff_directional.npz"))
    rays6, tmax, ref = np.ascontiguousarray(g["srays"], np.float32), np.ascontiguousarray(g["tmax"], np.float32), g["occ"]
    lib, h = open_scene(os.path.join(ROOT, "tests", "scenes", "cliff_directional.xml"))
    dev = torch.device("cuda:0")
    cap = len(rays6) + 64
    d_rays, d_tmax = torch.zeros(6 * cap, dtype=torch.float32, device=dev), torch.zeros(cap, dtype=torch.float32, device=dev)
    d_occ = torch.zeros(cap, dtype=torch.uint8, device=dev)
    stream = torch.cuda.Stream(device=dev)
    want = with_alloc(lib, h, rays6, tmax)
    assert np.array_equal(want, ref), "qg_trace_any itself differs from the reference fixture"
    # call 1: the whole fixture (thousands of rays through the hand-over queue)
    got, _, _ = without_alloc(lib, h, torch, stream, d_rays, d_tmax, d_occ, rays6, tmax)
    assert np.array_equal(got, want), "call 1: %d of %d rays differ" % ((got != want).sum(), len(want))
    # call 2, straight after: other rays, fewer of them — the cursor, the count and the queue hold call 1's values
    sel = np.arange(len(rays6))[::-3][:5001]
    r2, t2 = np.ascontiguousarray(rays6[sel]), np.ascontiguousarray(tmax[sel])
    got2, free0, free1 = without_alloc(lib, h, torch, stream, d_rays, d_tmax, d_occ, r2, t2)
    assert np.array_equal(got2, want[sel]), "call 2: %d of %d rays differ" % ((got2 != want[sel]).sum(), len(sel))
    assert np.array_equal(got2, with_alloc(lib, h, r2, t2))
    assert free1 >= free0, "device memory shrank by %d bytes across a call that must not allocate" % (free0 - free1)
    # call 3: the first set again through the same scratch
    got3, _, _ = without_alloc(lib, h, torch, stream, d_rays, d_tmax, d_occ, rays6, tmax)
    assert np.array_equal(got3, want)
    lib.qg_scene_destroy(h)


def test_no_alloc_on_a_reference_scene_and_the_scene_stream():
    torch = pytest.importorskip("torch")
    name = "example_project7_object"
    g = load_golden(name)
    rays6, tmax = np.ascontiguousarray(g["srays"], np.float32), np.ascontiguousarray(g["stmax"], np.float32)
    lib, h = open_scene(scene_xml(name))
    dev = torch.device("cuda:0")
    n = len(rays6)
    d_rays = torch.from_numpy(rays6.reshape(-1)).to(dev)
    d_tmax = torch.from_numpy(tmax).to(dev)
    d_occ = torch.full((n,), 7, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    for _ in range(2):  # stream NULL = the scene's own stream; nothing synchronises inside the call
        capi.check(lib, lib.qg_trace_any_noalloc(h, C.c_void_p(d_rays.data_ptr()), C.c_void_p(d_tmax.data_ptr()), n, C.c_void_p(d_occ.data_ptr()), None))
        torch.cuda.synchronize()
        got = d_occ.cpu().numpy()
        assert np.array_equal(got, g["socc"]), "%d of %d shadow rays differ from the reference" % ((got != g["socc"]).sum(), n)
        d_occ.fill_(7)
        torch.cuda.synchronize()
    assert np.array_equal(with_alloc(lib, h, rays6, tmax), g["socc"])
    # n = 0 is a no-op; a batch beyond the static scratch is refused with a message that names the limit
    capi.check(lib, lib.qg_trace_any_noalloc(h, None, None, 0, None, None))
    rc = lib.qg_trace_any_noalloc(h, C.c_void_p(d_rays.data_ptr()), C.c_void_p(d_tmax.data_ptr()), MAX_RAYS + 1, C.c_void_p(d_occ.data_ptr()), None)
    assert rc == capi.QG_EINVAL and b"QG_NOALLOC_MAX_RAYS" in lib.qg_last_error()
    lib.qg_scene_destroy(h)
