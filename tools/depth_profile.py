#!/usr/bin/env python3
"""Per-depth ray / BVH-node / triangle counts and kernel time of a scene (counting build + timing build).
Depth d's figures are the difference between renders with max_bounce = d and d-1 (paths are a pure function of
(pixel, sample, seed), so the shallower render is a prefix of the deeper one).   python tools/depth_profile.py [scene] [spp]"""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from qaray_b200 import capi
scene = sys.argv[1] if len(sys.argv) > 1 else "trc_scene_tower"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 8
xml = os.path.join(ROOT, "oracle", "_ref", "inputs", scene + ".xml")
res = {}
for counting in (True, False):
    lib = capi.load(counting)
    capi.check(lib, lib.qg_init(0))
    h = C.c_void_p()
    capi.check(lib, lib.qg_scene_load_xml(os.fsencode(xml), os.fsencode(os.path.join(ROOT, "assets")), C.byref(h)))
    info = capi.SceneInfo(); lib.qg_scene_info(h, C.byref(info))
    W, H = info.camera.width, info.camera.height
    acc = torch.zeros(W * H * 4, dtype=torch.float32, device="cuda")
    for mb in range(0, 6):
        for rep in range(2):
            p = capi.RenderParams(spp, mb, 1, 0x5EED, W, H, 0, 1, 0, 0, 0, 0)
            st = capi.RenderStats()
            capi.check(lib, lib.qg_render_accumulate(h, C.byref(p), C.c_void_p(acc.data_ptr()), None, None, C.byref(st)))
        res[(counting, mb)] = (st.closest_rays, st.shadow_rays, st.nodes_visited, st.tris_tested, st.objects_tested, st.shadow_nodes_visited,
                               st.shadow_tris_tested, st.trace_ms, st.shade_ms, st.shadow_ms, st.device_ms)
    lib.qg_scene_destroy(h)
print("%s %dx%d %d spp" % (scene, W, H, spp))
print("depth  closest_rays  nodes/ray  tris/ray  trace_ms  ns/ray | shadow_rays nodes/ray tris/ray shadow_ms ns/ray | shade_ms")
prev_c = prev_t = (0,) * 11
for mb in range(6):
    c, t = res[(True, mb)], res[(False, mb)]
    dc = [a - b for a, b in zip(c, prev_c)]
    dt = [a - b for a, b in zip(t, prev_t)]
    cr, sr = max(1, dc[0]), max(1, dc[1])
    print("%5d  %12d  %9.1f  %8.1f  %8.2f  %6.1f | %11d %9.1f %8.1f %9.2f %6.1f | %8.2f" % (
        mb, dc[0], dc[2] / cr, dc[3] / cr, dt[7], dt[7] * 1e6 / cr, dc[1], dc[5] / sr, dc[6] / sr, dt[9], dt[9] * 1e6 / sr, dt[8]))
    prev_c, prev_t = c, t
