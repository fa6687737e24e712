#!/usr/bin/env python3
"""Small invocations of every kernel family for compute-sanitizer (memcheck): low-resolution renders of a many-object scene (the
ancestor-ray cache on), of the tower, the photon-map path, the ray-buffer entry points and qg_trace_any_noalloc.
    compute-sanitizer --tool memcheck python tools/sanitize_small.py"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import ASSETS, load_golden, rays18_to_struct, scene_xml  # noqa: E402
from qaray_b200 import capi  # noqa: E402
from qaray_b200.renderer import Scene  # noqa: E402

lib = capi.load()
capi.check(lib, lib.qg_init(0))
vp = lambda a: a.ctypes.data_as(C.c_void_p)
for name, cls, sort in (("trc_scene_xmas", 0, 0), ("trc_scene_tower", 0, 0), ("trc_scene_tower", 0, 1), ("example_project7_object", 1, 0),
                        ("example_project7_object", 3, 0), ("trc_mtl_glass", 0, 0)):
    h = C.c_void_p()
    capi.check(lib, lib.qg_scene_load_xml(os.fsencode(scene_xml(name)), os.fsencode(ASSETS), C.byref(h)))
    W, H, spp = 96, 54, 2
    p = capi.RenderParams(spp, 3, 1, 0x5EED, W, H, 0, 1, 0, 0, 0, sort, cls)
    if name == "trc_mtl_glass":
        pp = capi.PhotonParams.defaults(0x5EED)
        pp.photon_map_size, pp.caustics_map_size = 2000, 200
        capi.check(lib, lib.qg_photon_maps_build(h, C.byref(pp), None))
        p.use_photon_map = 1
    lin = np.zeros((H, W, 3), np.float32)
    st = capi.RenderStats()
    capi.check(lib, lib.qg_render(h, C.byref(p), vp(lin), None, None, None, None, C.byref(st)))
    print("%-28s class %d sort %d: %d rays, %d launches, finite %s" % (name, cls, sort, st.closest_rays + st.shadow_rays, st.kernel_launches,
                                                                 bool(np.isfinite(lin).all())))
    lib.qg_scene_destroy(h)
g = load_golden("example_project7_object")
sc = Scene.load(scene_xml("example_project7_object"), ASSETS)
out = sc.trace_closest(rays18_to_struct(g["rays"][:512], g["has_diff"][:512]))
occ = sc.trace_any(g["srays"][:512], g["stmax"][:512])
print("ray buffers: %d hits, %d occluded" % (int(out["hit"].sum()), int(occ.sum())))
import torch  # noqa: E402
d_r = torch.from_numpy(np.ascontiguousarray(g["srays"][:512], np.float32).reshape(-1)).cuda()
d_t = torch.from_numpy(np.ascontiguousarray(g["stmax"][:512], np.float32)).cuda()
d_o = torch.zeros(512, dtype=torch.uint8, device="cuda")
torch.cuda.synchronize()
capi.check(lib, lib.qg_trace_any_noalloc(sc.handle, C.c_void_p(d_r.data_ptr()), C.c_void_p(d_t.data_ptr()), 512, C.c_void_p(d_o.data_ptr()), None))
torch.cuda.synchronize()
assert np.array_equal(d_o.cpu().numpy(), occ.astype(np.uint8))
print("noalloc: equal")
sc.close()
