#!/usr/bin/env python3
"""debug helper: render (class, scene) at several bounce limits. On the GPU box: `gpu` mode -> gpurun_out/dbg_*.npy;
here: `ref` mode -> the same from the oracle variant, then `cmp`."""
import os, re, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
mode, cls, name, W, H, spp = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]), int(sys.argv[5]), int(sys.argv[6])
OUT = os.path.join(ROOT, "gpurun_out")
CLASS_ID = {"photonmap": 0, "pathtracing": 1, "montecarlogi": 2, "blinn_basic": 3, "phong_basic": 4}
for b in (0, 1, 2, 5):
    if mode == "gpu":
        from qaray_b200.renderer import Renderer_B200, RendererParam
        p = RendererParam(spp=spp, max_bounce=b); p.width, p.height, p.material_class = W, H, CLASS_ID[cls]
        r = Renderer_B200(p); r.Init(0)
        r.LoadScene(os.path.join(ROOT, "oracle", "_ref", "inputs", name + ".xml"), os.path.join(ROOT, "assets"))
        r.ComputeScene(); r.Render()
        np.save(os.path.join(OUT, "dbg_gpu_%s_%s_b%d.npy" % (cls, name, b)), r.linearBuffer); r.Terminate()
    elif mode == "ref":
        pid = os.fork()
        if pid == 0:
            from oracle import refapi
            xml = open(os.path.join(ROOT, "oracle", "_ref", "inputs", name + ".xml")).read()
            xml = re.sub(r'<width\s+value="\d+"\s*/>', '<width value="%d"/>' % W, xml)
            xml = re.sub(r'<height\s+value="\d+"\s*/>', '<height value="%d"/>' % H, xml)
            fp = tempfile.NamedTemporaryFile("w", suffix=".xml", delete=False); fp.write(xml); fp.close()
            ref = refapi.Ref("" if cls == "photonmap" else cls); ref.set_params(spp, spp, b, True, 1); ref.load(fp.name, os.path.join(ROOT, "assets"))
            ref.seed(1); rgb, _ = ref.render_linear_serial(spp)
            np.save(os.path.join(OUT, "dbg_ref_%s_%s_b%d.npy" % (cls, name, b)), rgb); os._exit(0)
        os.waitpid(pid, 0)
    else:
        g = np.load(os.path.join(OUT, "dbg_gpu_%s_%s_b%d.npy" % (cls, name, b))); r = np.load(os.path.join(OUT, "dbg_ref_%s_%s_b%d.npy" % (cls, name, b)))
        bad = (np.abs(g - r) > 2e-3 + 2e-3 * np.abs(r)).any(axis=2)
        print("bounce", b, "bad", bad.sum(), "of", bad.size, "max", np.abs(g - r).max(), "mean g/r", g.mean(), r.mean())
        if bad.any():
            ys, xs = np.nonzero(bad)
            print("  rows", ys.min(), ys.max(), "cols", xs.min(), xs.max(), " sample", [(int(y), int(x), g[y, x].round(4).tolist(), r[y, x].round(4).tolist()) for y, x in list(zip(ys, xs))[:: max(1, len(ys) // 4)][:4]])
