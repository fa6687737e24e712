"""CPU tests that pin the golden vectors to the reference: oracle/_ref/libqaref_parity.so (the reference's own
sources compiled by oracle/Makefile) must reproduce every committed fixture bit for bit. One fresh process per
scene, because the reference keeps its scene in a global (src/scene/scene.cpp:76)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ASSETS, GOLDEN_DIR, ROOT, class_golden_cases, golden_names, scene_xml

WORKER = r"""
import sys, numpy as np
sys.path.insert(0, %(root)r)
from oracle import refapi
g = np.load(%(npz)r)
ref = refapi.Ref(); ref.set_params(1, 1, 5, True, 1); ref.load(%(xml)r, %(assets)r)
pp = g["prim_pix"]; n = len(pp) // 3
for s in range(3):
    r = ref.primary_rays(pp[s * n:(s + 1) * n, :2], s)
    assert r.view(np.uint32).tolist() == g["rays"][s * n:(s + 1) * n].view(np.uint32).tolist(), "primary rays"
h = ref.trace_closest(g["rays"], g["has_diff"])
assert h.tobytes() == g["hits"].tobytes(), "closest hits"
if "srays" in g.files:
    assert np.array_equal(ref.trace_shadow(g["srays"], g["stmax"]), g["socc"]), "shadow"
if "rad_pix" in g.files:
    rgb, z, hit = ref.samples(g["rad_pix"], int(g["rad_seed"][0]))
    assert rgb.tobytes() == g["rad_rgb"].tobytes() and z.tobytes() == g["rad_z"].tobytes(), "radiance"
print("OK")
"""


@pytest.mark.parametrize("name", golden_names())
def test_reference_reproduces_golden(name):
    from oracle import refapi
    if not refapi.available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    xml = scene_xml(name)
    code = WORKER % dict(root=ROOT, npz=os.path.join(GOLDEN_DIR, name + ".npz"), xml=xml, assets=ASSETS)
    r = subprocess.run([sys.executable, "-c", code], cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and r.stdout.strip().endswith("OK"), r.stdout[-2000:]


CLASS_WORKER = r"""
import sys, re, os, tempfile, numpy as np
sys.path.insert(0, %(root)r)
from oracle import refapi
g = np.load(%(npz)r)
W, H, spp = [int(v) for v in g["size"]]
xml = open(%(xml)r).read()
xml = re.sub(r'<width\s+value="\d+"\s*/>', '<width value="%%d"/>' %% W, xml)
xml = re.sub(r'<height\s+value="\d+"\s*/>', '<height value="%%d"/>' %% H, xml)
fp = tempfile.NamedTemporaryFile("w", suffix=".xml", delete=False); fp.write(xml); fp.close()
ref = refapi.Ref(%(cls)r); ref.set_params(spp, spp, 5, True, 1); ref.load(fp.name, %(assets)r); os.unlink(fp.name)
if "ref" in g.files:
    ref.seed(4242)
    rgb, depth = ref.render_linear_serial(spp)
    assert rgb.tobytes() == g["ref"].tobytes() and depth.tobytes() == g["depth"].tobytes(), "deterministic image"
if "rad_pix" in g.files:
    rgb, z, hit = ref.samples(g["rad_pix"], int(g["rad_seed"][0]))
    assert rgb.tobytes() == g["rad_rgb"].tobytes(), "radiance"
print("OK")
"""


@pytest.mark.parametrize("cls,name", [c for c in class_golden_cases() if c[0] in ("blinn_basic", "phong_basic", "pathtracing")])
def test_reference_class_variant_reproduces_golden(cls, name):
    """the reference compiled with `using MtlBlinn = <class>` (oracle/_ref/libqaref_parity_<class>.so) reproduces the
    deterministic class fixtures bit for bit (stochastic ref_a/ref_b images are seed pairs, checked by their statistics)"""
    from oracle import refapi
    if not refapi.available(cls):
        pytest.skip("oracle/_ref variant %s not built (needs /root/reference)" % cls)
    g = np.load(os.path.join(GOLDEN_DIR, "cls_%s_%s.npz" % (cls, name)))
    if "ref" not in g.files and "rad_pix" not in g.files:
        pytest.skip("stochastic fixture")
    code = CLASS_WORKER % dict(root=ROOT, npz=os.path.join(GOLDEN_DIR, "cls_%s_%s.npz" % (cls, name)), xml=scene_xml(name), assets=ASSETS,
                               cls=cls)
    r = subprocess.run([sys.executable, "-c", code], cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and r.stdout.strip().endswith("OK"), r.stdout[-2000:]
