"""CPU tests that pin the golden vectors to the reference: oracle/_ref/libqaref_parity.so (the reference's own
sources compiled by oracle/Makefile) must reproduce every committed fixture bit for bit. One fresh process per
scene, because the reference keeps its scene in a global (src/scene/scene.cpp:76)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ASSETS, GOLDEN_DIR, ROOT, golden_names, scene_xml

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
