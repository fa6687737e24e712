"""Device time of one sample shard [s0, s1) of an S-spp frame on one GPU (the work one rank of bench.py --gpus N does):
python tools/shard_time.py trc_scene_tower 64 8   ->  one line per shard of a 512-spp frame."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from qaray_b200 import capi  # noqa: E402

scene, spp, shards = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
lib = capi.load()
capi.check(lib, lib.qg_init(0))
h = C.c_void_p()
xml = os.path.join(ROOT, "oracle", "_ref", "inputs", scene + ".xml")
capi.check(lib, lib.qg_scene_load_xml(os.fsencode(xml), os.fsencode(os.path.join(ROOT, "assets")), C.byref(h)))
info = capi.SceneInfo()
capi.check(lib, lib.qg_scene_info(h, C.byref(info)))
W, H = info.camera.width, info.camera.height
acc = torch.zeros(W * H * 4, dtype=torch.float32, device="cuda")
dep = torch.zeros(W * H, dtype=torch.float32, device="cuda")
st = capi.RenderStats()
for r in list(range(shards)) * 2:
    p = capi.RenderParams(spp * shards, 5, 1, 0x5EED, W, H, 0, 1, r * spp, (r + 1) * spp, 0, 0)
    acc.zero_(); dep.zero_(); torch.cuda.synchronize()
    capi.check(lib, lib.qg_render_accumulate(h, C.byref(p), C.c_void_p(acc.data_ptr()), C.c_void_p(dep.data_ptr()), None, C.byref(st)))
    print("shard %d [%d,%d): %.2f ms  trace %.2f shade %.2f shadow %.2f  rays %d" % (
        r, r * spp, (r + 1) * spp, st.device_ms, st.trace_ms, st.shade_ms, st.shadow_ms, st.closest_rays + st.shadow_rays), flush=True)
