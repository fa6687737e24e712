"""CPU tests of the host half of the boundary: scene XML / OBJ / PNG loading against the golden node tables dumped
from the reference (tools/make_golden.py), and that the C-ABI library exports what include/qaray_b200.h declares."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ASSETS, ROOT, golden_names, load_golden, scene_xml
from qaray_b200 import capi


def host_scene(xml):
    lib = capi.load()
    h = C.c_void_p()
    capi.check(lib, lib.qg_host_scene_load_xml(os.fsencode(xml), os.fsencode(ASSETS), C.byref(h)))
    return lib, h, lib.qg_host_scene_desc(h).contents


def test_library_exports_every_declared_symbol():
    lib = capi.load()
    header = open(os.path.join(ROOT, "include", "qaray_b200.h")).read()
    declared = set(re.findall(r"\b(qg_[a-z0-9_]+)\s*\(", header))
    assert declared, "no prototypes found in the header"
    assert declared == set(capi.PROTOTYPES), "capi.PROTOTYPES out of sync with the header: %s" % (declared ^ set(capi.PROTOTYPES))
    for name in declared:
        assert hasattr(lib, name), name


def test_struct_sizes_match_the_header():
    # sizes the C compiler gives the header's structs (computed once with gcc; see tests/golden/README)
    assert C.sizeof(capi.Xform) == 84
    assert C.sizeof(capi.Node) == 100
    assert C.sizeof(capi.TexturedColor) == 100
    assert C.sizeof(capi.Light) == 52
    assert C.sizeof(capi.Camera) == 56
    assert C.sizeof(capi.RenderParams) == 60
    assert C.sizeof(capi.PhotonParams) == 28 and C.sizeof(capi.PhotonStats) == 72
    assert capi.RAY_DTYPE.itemsize == 76 and capi.HIT_DTYPE.itemsize == 88


def test_compute_entry_points_fail_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = capi.load()
    h = C.c_void_p()
    xml = scene_xml("example_project3_sphere")
    rc = lib.qg_scene_load_xml(os.fsencode(xml), b"", C.byref(h))
    assert rc == capi.QG_ENODEVICE
    assert b"no CPU path" in lib.qg_last_error()


def test_missing_scene_file_is_eio():
    lib = capi.load()
    h = C.c_void_p()
    assert lib.qg_host_scene_load_xml(b"/nonexistent/scene.xml", b"", C.byref(h)) == capi.QG_EIO


@pytest.mark.parametrize("name", golden_names())
def test_node_table_matches_reference_bit_for_bit(name):
    g = load_golden(name)
    lib, h, d = host_scene(scene_xml(name))
    try:
        meta, xf = g["node_meta"], g["node_xf"]
        assert d.num_nodes == len(meta)
        for i in range(d.num_nodes):
            n = d.nodes[i]
            assert n.parent == meta[i, 0], (i, "parent")
            assert n.object_type == meta[i, 1], (i, "type")
            if n.object_type == capi.QG_OBJ_MESH:
                assert d.meshes[n.mesh].num_faces == meta[i, 3], (i, "faces")
            mine = np.array(list(n.xf.tm) + list(n.xf.itm) + list(n.xf.pos), np.float32)
            assert mine.view(np.uint32).tolist() == xf[i].view(np.uint32).tolist(), (i, mine, xf[i])
        cam = g["camera"]
        c = d.camera
        mine = np.array(list(c.pos) + list(c.dir) + list(c.up) + [c.fovy, c.focal, c.dof, c.width, c.height], np.float32)
        assert mine.view(np.uint32).tolist() == cam.view(np.uint32).tolist()
    finally:
        lib.qg_host_scene_destroy(h)


def test_tinyobj_number_parser_quirks(tmp_path):
    # values whose tinyobj parse differs from strtod must load as tinyobj would give them: digit-by-digit sums
    obj = tmp_path / "t.obj"
    obj.write_text("v 0.1 0.123456789 1e-3\nv 1 0 0\nv 0 1 0\nvn 0 0 1\nf 1//1 2//1 3//1\nf -3//-1 -2//-1 -1//-1\n")
    xml = tmp_path / "s.xml"
    xml.write_text('<xml><scene><object type="obj" name="t.obj"/></scene><camera/></xml>')
    lib = capi.load()
    h = C.c_void_p()
    capi.check(lib, lib.qg_host_scene_load_xml(os.fsencode(str(xml)), os.fsencode(str(tmp_path)), C.byref(h)))
    d = lib.qg_host_scene_desc(h).contents
    m = d.meshes[0]
    assert m.num_faces == 2 and m.num_vertices == 3
    v = [m.vertices[k] for k in range(3)]
    # 0.123456789 = sum d_k * 10^-k accumulated in double, then cast
    acc = 0.0
    for k, ch in enumerate("123456789", 1):
        acc += int(ch) * ([1.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001][k] if k < 8 else pow(10.0, -k))
    assert v[1] == np.float32(acc)
    assert v[2] == np.float32(np.ldexp(1.0 * pow(5.0, -3), -3))
    # negative indices are relative
    assert [m.face_indices[9 + 3 * k] for k in range(3)] == [0, 1, 2]
    lib.qg_host_scene_destroy(h)
