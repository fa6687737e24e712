"""The OBJ / MTL reader (qaray_b200/csrc/objload.cpp) against the reference's own copy of tinyobjloader v1.1.0
(external/tinyobjloader, called as src/mesh/TriMesh.cpp:76 calls it; oracle/ref_driver.cpp qaref_tinyobj_load): random
decimal literals in every shape the number parser distinguishes, relative and absolute indices, polygons, material switches.
Vertex / normal / texcoord floats must be bit-equal, triangles and their material ids equal in file order. CPU only; skipped
where the oracle is not built."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import ROOT
from qaray_b200 import capi

REF_SO = os.path.join(ROOT, "oracle", "_ref", "libqaref_parity.so")


def literals(rng, n):
    out = []
    forms = ["{s}{i}", "{s}{i}.{f}", "{s}{i}.", "{s}{i}e{es}{e}", "{s}{i}.{f}e{es}{e}", "{s}{i}.{f}E{es}{e}", "{s}.{f}", "{s}{i}e", "{s}{i}e{es}",
             "{s}{i}.{f}e", "{s}{i}x", "{s}{i}.{f}f", "{s}0.{z}{f}", "{s}{i}{i2}.{f}{f2}", "{s}0", "{s}0.0", "{s}{i}e0", "{s}{i}.{f}e-0"]
    for _ in range(n):
        form = forms[rng.integers(0, len(forms))]
        d = lambda k: "".join(str(int(x)) for x in rng.integers(0, 10, k))
        out.append(form.format(s=["", "-", "+"][rng.integers(0, 3)], i=d(rng.integers(1, 8)), i2=d(rng.integers(1, 12)), f=d(rng.integers(1, 12)),
                               f2=d(rng.integers(1, 14)), z="0" * int(rng.integers(0, 9)), es=["", "-", "+"][rng.integers(0, 3)], e=d(rng.integers(1, 3))))
    return out


def write_case(tmp_path, rng, k):
    lit = literals(rng, 9 * 40 + 2 * 30)
    lines = ["mtllib missing.mtl m%d.mtl" % k, "# comment", ""]
    it = iter(lit)
    nv, nn, nt = 40, 40, 30
    for _ in range(nv):
        lines.append("v %s %s  %s" % (next(it), next(it), next(it)))
    for _ in range(nn):
        lines.append("vn\t%s %s %s " % (next(it), next(it), next(it)))
    for _ in range(nt):
        lines.append("vt %s %s" % (next(it), next(it)))
    mats = ["red", "green", "blue two"]
    for f in range(60):
        if rng.random() < 0.3:
            lines.append("usemtl %s" % (mats + ["nosuch"])[rng.integers(0, 4)])
        nc = int(rng.integers(3, 7))
        style = rng.integers(0, 4)
        corners = []
        for _ in range(nc):
            v = int(rng.integers(1, nv + 1)) if rng.random() < 0.7 else -int(rng.integers(1, nv + 1))
            n = int(rng.integers(1, nn + 1)) if rng.random() < 0.7 else -int(rng.integers(1, nn + 1))
            t = int(rng.integers(1, nt + 1)) if rng.random() < 0.7 else -int(rng.integers(1, nt + 1))
            corners.append([str(v), "%d/%d" % (v, t), "%d//%d" % (v, n), "%d/%d/%d" % (v, t, n)][style])
        lines.append("f " + " ".join(corners) + ("\r" if rng.random() < 0.2 else ""))
    obj = tmp_path / ("c%d.obj" % k)
    obj.write_text("\n".join(lines) + "\n")
    ml = []
    for m in mats:
        ml += ["newmtl %s" % m, "Kd %s %s %s" % tuple(literals(rng, 3)), "Ks %s %s %s  " % tuple(literals(rng, 3)), "Ns %s" % literals(rng, 1)[0],
               "illum %d" % rng.integers(0, 8), "Ni %s" % literals(rng, 1)[0], ""]
    (tmp_path / ("m%d.mtl" % k)).write_text("\n".join(ml) + "\n")
    return obj


def reference_load(path, mtl_dir):
    ref = C.CDLL(REF_SO)
    counts = (C.c_int64 * 5)()
    cap, capt = 1 << 16, 1 << 14
    v, n, t = (np.zeros(cap, np.float32) for _ in range(3))
    tri = np.zeros((capt, 10), np.int32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    ref.qaref_tinyobj_load.argtypes = [C.c_char_p, C.c_char_p, C.c_void_p] + [C.c_void_p] * 4 + [C.c_int64, C.c_int64]
    assert ref.qaref_tinyobj_load(os.fsencode(str(path)), os.fsencode(str(mtl_dir) + os.sep), counts, p(v), p(n), p(t), p(tri), cap, capt) == 1
    return v[:counts[0]], n[:counts[1]], t[:counts[2]], tri[:counts[3]], counts[4]


def own_load(tmp_path, obj):
    lib = capi.load()
    xml = tmp_path / "s.xml"
    xml.write_text('<xml><scene><object type="obj" name="%s"/></scene><camera/></xml>' % obj.name)
    h = C.c_void_p()
    capi.check(lib, lib.qg_host_scene_load_xml(os.fsencode(str(xml)), os.fsencode(str(tmp_path)), C.byref(h)))
    m = lib.qg_host_scene_desc(h).contents.meshes[0]
    arr = lambda ptr, k, dt: np.ctypeslib.as_array(ptr, shape=(k,)).astype(dt).copy() if k else np.zeros(0, dt)
    out = (arr(m.vertices, 3 * m.num_vertices, np.float32), arr(m.normals, 3 * m.num_normals, np.float32), arr(m.texcoords, 2 * m.num_texcoords, np.float32),
           arr(m.face_indices, 9 * m.num_faces, np.int32).reshape(-1, 9), arr(m.face_material, m.num_faces, np.int32))
    lib.qg_host_scene_destroy(h)
    return out


@pytest.mark.parametrize("k", range(6))
def test_reader_is_bit_equal_to_the_references_tinyobjloader(tmp_path, k):
    if not os.path.exists(REF_SO):
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(1000 + k)
    obj = write_case(tmp_path, rng, k)
    rv, rn, rt, rtri, nm = reference_load(obj, tmp_path)
    v, n, t, fi, fm = own_load(tmp_path, obj)
    assert nm == 3
    for a, b, what in ((v, rv, "vertices"), (n, rn, "normals"), (t, rt, "texcoords")):
        assert len(a) == len(b) and np.array_equal(a.view(np.uint32), b.view(np.uint32)), "%s differ in bits" % what
    # the loader sorts faces by material afterwards (TriMesh.cpp:107-114, checked on the fixture meshes by tests/test_binding.py);
    # here: the same multiset of (corner indices, material) triangles
    key = lambda idx, mat: sorted(map(tuple, np.concatenate([idx, mat[:, None]], 1).tolist()))
    assert len(fi) == len(rtri) and key(fi, fm) == key(rtri[:, :9], rtri[:, 9])
