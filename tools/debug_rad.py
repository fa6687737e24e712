import os, sys, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import ASSETS, REF_INPUTS, load_golden
from qaray_b200.renderer import RendererParam, Scene
for name in sys.argv[1:]:
    g = load_golden(name)
    sc = Scene.load(os.path.join(REF_INPUTS, name + ".xml"), ASSETS)
    for mb in (0, 5):
        rgb, z = sc.sample_radiance(RendererParam(spp=64, max_bounce=mb, seed=int(g["rad_seed"][0])), g["rad_pix"])
        ref = g["rad_rgb"]
        bad = (np.abs(rgb - ref) > 2e-5 + 2e-4 * np.abs(ref)).any(axis=1)
        print(name, "max_bounce", mb, "bad", int(bad.sum()), "of", len(ref), "mean", rgb.mean(), ref.mean())
    idx = np.nonzero(bad)[0][:12]
    for k in idx:
        print(k, g["rad_pix"][k], "hit", g["rad_hit"][k], "gpu", rgb[k], "ref", ref[k], "diff", rgb[k] - ref[k])
    sc.close()
