"""The drop-in proven in the reference's own language: integration/Renderer_B200.h (Flattened + Renderer_B200 : public
qaray::Renderer, what INTEGRATION.md tells a maintainer to add) compiled against the reference's unmodified headers and linked
with its objects (oracle/Makefile -> oracle/_ref/libqaref_binding.so, driver oracle/binding_driver.cpp), then driven as
src/main.cpp:55-59 drives a renderer.

CPU: the qgSceneDesc Flattened makes of the reference's in-memory qaray::Scene describes the same scene as this repository's own
loader (qg_host_scene_load_xml) on every fixture scene — node tables, meshes, lights and camera byte for byte, materials and
textures with their references resolved (the two sides number textures in a different order).
GPU (-m gpu): the subclass renders through qg_scene_create; its hits equal the reference fixtures bit for bit and its frame
equals the frame of the loader path."""
import ctypes as C
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ASSETS, GOLDEN_DIR, REF_INPUTS, ROOT, golden_names

BINDING = os.path.join(ROOT, "oracle", "_ref", "libqaref_binding.so")

WORKER = r'''
import ctypes as C, json, os, sys
import numpy as np
sys.path.insert(0, %(root)r)
sys.path.insert(0, os.path.join(%(root)r, "tests"))
from qaray_b200 import capi
name, mode = sys.argv[1], sys.argv[2]
xml = os.path.join(%(inputs)r, name + ".xml")
lib = capi.load()
bind = C.CDLL(%(binding)r, mode=C.RTLD_GLOBAL)
bind.qabind_load.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int]
bind.qabind_desc.restype = C.POINTER(capi.SceneDesc)
bind.qabind_scene.restype = C.c_void_p


def f32(ptr, n):
    return np.ctypeslib.as_array(ptr, shape=(n,)).copy() if n else np.zeros(0, np.float32)


def texcolor(d, tc):
    """a TexturedColor with its texture resolved to content"""
    out = {"color": list(tc.color), "xf": bytes(tc.xf)}
    if tc.texture >= 0:
        t = d.textures[tc.texture]
        px = bytes(C.cast(t.rgb8, C.POINTER(C.c_uint8 * (t.width * t.height * 3))).contents) if t.type == 2 and t.width * t.height else b""
        out["tex"] = (t.type, list(t.color1), list(t.color2), t.width, t.height, px)
        if t.type != 1:  # only the checker reads its colours
            out["tex"] = (t.type, t.width, t.height, px)
    return out


def material(d, m):
    out = {"kind": m.kind, "abs": list(m.absorption), "ior": m.ior, "gloss": (m.glossiness, m.reflection_glossiness, m.refraction_glossiness), "kill": m.kill}
    if m.kind == 1:
        out = {"kind": 1, "subs": [material(d, d.materials[d.multi[m.sub_first + k]]) for k in range(m.sub_count)]}
    else:
        for f in ("diffuse", "specular", "emission", "reflection", "refraction"):
            out[f] = texcolor(d, getattr(m, f))
    return out


def describe(d):
    out = {"nodes": [], "meshes": [], "lights": [bytes(d.lights[i]) for i in range(d.num_lights)], "camera": bytes(d.camera)}
    for i in range(d.num_nodes):
        n = d.nodes[i]
        out["nodes"].append((n.parent, n.object_type, n.mesh, bytes(n.xf), material(d, d.materials[n.material]) if n.material >= 0 else None))
    for i in range(d.num_meshes):
        m = d.meshes[i]
        out["meshes"].append((m.num_vertices, m.num_normals, m.num_texcoords, m.num_faces, f32(m.vertices, 3 * m.num_vertices).tobytes(),
                              f32(m.normals, 3 * m.num_normals).tobytes(), f32(m.texcoords, 2 * m.num_texcoords).tobytes() if m.num_texcoords else b"",
                              np.ctypeslib.as_array(m.face_indices, shape=(9 * m.num_faces,)).tobytes() if m.num_faces else b"",
                              np.ctypeslib.as_array(m.face_material, shape=(m.num_faces,)).tobytes() if m.num_faces else b"",
                              list(m.bound_min), list(m.bound_max)))
    out["background"] = texcolor(d, d.background)
    out["environment"] = texcolor(d, d.environment)
    return out


if mode == "desc":
    hs = C.c_void_p()
    capi.check(lib, lib.qg_host_scene_load_xml(os.fsencode(xml), os.fsencode(%(assets)r), C.byref(hs)))
    mine = describe(lib.qg_host_scene_desc(hs).contents)
    assert bind.qabind_load(os.fsencode(xml), os.fsencode(%(assets)r), 4, 5, 1) == 1, "reference LoadScene failed"
    theirs = describe(bind.qabind_desc().contents)
    diff = [k for k in mine if mine[k] != theirs[k]]
    detail = ""
    if "nodes" in diff:
        bad = [i for i, (a, b) in enumerate(zip(mine["nodes"], theirs["nodes"])) if a != b]
        detail = "nodes %%s differ (of %%d / %%d)" %% (bad[:5], len(mine["nodes"]), len(theirs["nodes"]))
        if bad:
            a, b = mine["nodes"][bad[0]], theirs["nodes"][bad[0]]
            detail += "; fields " + str([k for k in range(5) if a[k] != b[k]])
            if a[4] != b[4] and a[4] and b[4]:
                detail += " material keys " + str([k for k in a[4] if a[4].get(k) != b[4].get(k)])
    if "meshes" in diff:
        bad = [i for i, (a, b) in enumerate(zip(mine["meshes"], theirs["meshes"])) if a != b]
        detail += " meshes %%s differ" %% bad[:5]
        if bad:
            detail += " fields " + str([k for k in range(11) if mine["meshes"][bad[0]][k] != theirs["meshes"][bad[0]][k]])
    print("RESULT " + json.dumps({"diff": diff, "detail": detail, "nodes": len(mine["nodes"]), "meshes": len(mine["meshes"])}))
elif mode == "gpu":
    g = np.load(os.path.join(%(golden)r, name + ".npz"))
    from conftest import rays18_to_struct
    spp = 4
    assert bind.qabind_load(os.fsencode(xml), os.fsencode(%(assets)r), spp, 5, 1) == 1
    W, H = bind.qabind_desc().contents.camera.width, bind.qabind_desc().contents.camera.height
    rgb8, depth = np.zeros((H, W, 3), np.uint8), np.zeros((H, W), np.float32)
    sc, mask = np.zeros((H, W), np.uint8), np.zeros((H, W), np.uint8)
    rays = C.c_uint64()
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    assert bind.qabind_render(vp(rgb8), vp(depth), vp(sc), vp(mask), C.byref(rays)) == 1
    h = C.c_void_p(bind.qabind_scene())
    r = rays18_to_struct(g["rays"], g["has_diff"])
    out = np.zeros(len(r), capi.HIT_DTYPE)
    capi.check(lib, lib.qg_trace_closest(h, vp(r), len(r), vp(out)))
    ref = g["hits"]
    m = ref["hit"] == 1
    hits_ok = bool(np.array_equal(out["hit"], ref["hit"]) and np.array_equal(out["node"][m], ref["node"][m]) and np.array_equal(out["face"][m], ref["face"][m])
                   and np.array_equal(out["z"][m].view(np.uint32), ref["z"][m].view(np.uint32)))
    # the same frame through the library's own loader
    h2 = C.c_void_p()
    capi.check(lib, lib.qg_scene_load_xml(os.fsencode(xml), os.fsencode(%(assets)r), C.byref(h2)))
    rp = capi.RenderParams(spp, 5, 1, 0x5EED, W, H, 0, 1, 0, 0, 0, 0, 0, 0, 0)
    rgb8b, depthb = np.zeros((H, W, 3), np.uint8), np.zeros((H, W), np.float32)
    st = capi.RenderStats()
    capi.check(lib, lib.qg_render(h2, C.byref(rp), None, vp(rgb8b), vp(depthb), None, None, C.byref(st)))
    lib.qg_scene_destroy(h2)
    bind.qabind_terminate()
    print("RESULT " + json.dumps({"hits_ok": hits_ok, "mask": bool(mask.all()), "depth_equal": bool(np.array_equal(depth, depthb)),
                                  "rgb_maxdiff": int(np.abs(rgb8.astype(np.int32) - rgb8b.astype(np.int32)).max()),
                                  "rays_equal": bool(rays.value == st.closest_rays + st.shadow_rays), "sc": int(sc.min())}))
'''


def run_worker(name, mode, timeout=600):
    code = WORKER % {"root": ROOT, "inputs": REF_INPUTS, "binding": BINDING, "assets": ASSETS, "golden": GOLDEN_DIR}
    r = subprocess.run([sys.executable, "-c", code, name, mode], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=timeout, cwd=ROOT)
    lines = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
    assert r.returncode == 0 and lines, "worker failed: %s\n%s" % (r.stdout[-1500:], r.stderr[-3000:])
    return json.loads(lines[-1][7:])


def need_binding():
    if not os.path.exists(BINDING):
        pytest.skip("oracle/_ref/libqaref_binding.so missing (run `make -C oracle ref` where /root/reference exists, after building the library)")
    if not os.path.exists(os.path.join(ASSETS, "STANDIN_ASSETS.txt")):
        pytest.skip("stand-in assets missing")


@pytest.mark.parametrize("name", golden_names())
def test_flattened_reference_scene_equals_the_loader_tables(name):
    need_binding()
    res = run_worker(name, "desc")
    assert res["diff"] == [], "%s: %s" % (res["diff"], res["detail"])
    assert res["nodes"] >= 1


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["example_project3_sphere", "example_project7_object", "example_project12_caustics_glossy", "trc_scene_xmas"])
def test_renderer_subclass_renders_through_the_c_abi(name):
    need_binding()
    res = run_worker(name, "gpu")
    assert res["hits_ok"], "hits of the binding's scene differ from the reference fixtures"
    assert res["mask"] and res["depth_equal"] and res["rays_equal"] and res["sc"] == 255
    assert res["rgb_maxdiff"] <= 1  # two renders: the fp32 atomic additions come in a different order
