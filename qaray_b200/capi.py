"""ctypes view of include/qaray_b200.h (the C ABI built into qaray_b200/libqaray_b200.so).

Nothing here computes: it declares the structs and prototypes of the header one to one and loads the library.
Loading fails loudly when the library has not been built (`python -c "import __graft_entry__ as g; g.build()"`).
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libqaray_b200.so")
LIB_COUNT_PATH = os.path.join(HERE, "libqaray_b200_count.so")

QG_OK, QG_EINVAL, QG_EIO, QG_ENODEVICE, QG_ECUDA, QG_ENOMEM = 0, -1, -2, -3, -4, -5
QG_OBJ_NONE, QG_OBJ_SPHERE, QG_OBJ_PLANE, QG_OBJ_MESH = 0, 1, 2, 3
QG_LIGHT_AMBIENT, QG_LIGHT_DIRECT, QG_LIGHT_POINT, QG_LIGHT_SPOT = 0, 1, 2, 3
QG_TEX_NULL, QG_TEX_CHECKER, QG_TEX_IMAGE = 0, 1, 2
QG_MTL_BLINN, QG_MTL_MULTI = 0, 1

f32, i32, u32, u8, i64, u64 = C.c_float, C.c_int32, C.c_uint32, C.c_uint8, C.c_int64, C.c_uint64
fp = C.POINTER(f32)


class Xform(C.Structure):
    _fields_ = [("tm", f32 * 9), ("itm", f32 * 9), ("pos", f32 * 3)]


class Node(C.Structure):
    _fields_ = [("parent", i32), ("object_type", i32), ("mesh", i32), ("material", i32), ("xf", Xform)]


class Mesh(C.Structure):
    _fields_ = [("vertices", fp), ("normals", fp), ("texcoords", fp), ("face_indices", C.POINTER(i32)),
                ("face_material", C.POINTER(i32)), ("num_vertices", i32), ("num_normals", i32), ("num_texcoords", i32),
                ("num_faces", i32), ("bound_min", f32 * 3), ("bound_max", f32 * 3)]


class TexturedColor(C.Structure):
    _fields_ = [("color", f32 * 3), ("texture", i32), ("xf", Xform)]


class Texture(C.Structure):
    _fields_ = [("type", i32), ("color1", f32 * 3), ("color2", f32 * 3), ("width", i32), ("height", i32),
                ("rgb8", C.POINTER(u8))]


class Material(C.Structure):
    _fields_ = [("kind", i32), ("diffuse", TexturedColor), ("specular", TexturedColor), ("emission", TexturedColor),
                ("reflection", TexturedColor), ("refraction", TexturedColor), ("absorption", f32 * 3), ("ior", f32),
                ("glossiness", f32), ("reflection_glossiness", f32), ("refraction_glossiness", f32), ("kill", f32),
                ("sub_first", i32), ("sub_count", i32)]


class Light(C.Structure):
    _fields_ = [("type", i32), ("intensity", f32 * 3), ("position", f32 * 3), ("direction", f32 * 3), ("size", f32),
                ("inner", f32), ("outer", f32)]


class Camera(C.Structure):
    _fields_ = [("pos", f32 * 3), ("dir", f32 * 3), ("up", f32 * 3), ("fovy", f32), ("focal", f32), ("dof", f32),
                ("width", i32), ("height", i32)]


class SceneDesc(C.Structure):
    _fields_ = [("nodes", C.POINTER(Node)), ("meshes", C.POINTER(Mesh)), ("materials", C.POINTER(Material)),
                ("multi", C.POINTER(i32)), ("lights", C.POINTER(Light)), ("textures", C.POINTER(Texture)),
                ("num_nodes", i32), ("num_meshes", i32), ("num_materials", i32), ("num_multi", i32), ("num_lights", i32),
                ("num_textures", i32), ("background", TexturedColor), ("environment", TexturedColor), ("camera", Camera)]


class RenderParams(C.Structure):
    _fields_ = [("spp", i32), ("max_bounce", i32), ("srgb", i32), ("seed", u32), ("width", i32), ("height", i32),
                ("tile_rank", i32), ("tile_ranks", i32), ("sample_begin", i32), ("sample_end", i32), ("batch_paths", i32),
                ("sort_by_material", i32), ("material_class", i32), ("spp_min", i32), ("use_photon_map", i32)]


class PhotonParams(C.Structure):
    _fields_ = [("photon_map_size", i32), ("photon_map_bounce", i32), ("photon_map_radius", f32), ("caustics_map_size", i32),
                ("caustics_map_bounce", i32), ("caustics_map_radius", f32), ("seed", u32)]

    @classmethod
    def defaults(cls, seed=0x5EED):  # RendererParam (src/renderers/renderer.h:52-57)
        return cls(10000, 20, 0.2, 1000, 20, 1.0, seed)


class PhotonStats(C.Structure):
    _fields_ = [("photons", u64 * 2), ("emitted", u64 * 2), ("emissions", u64 * 2), ("rays", u64 * 2), ("device_ms", C.c_double)]


class RenderStats(C.Structure):
    _fields_ = [("samples", u64), ("closest_rays", u64), ("shadow_rays", u64), ("kernel_launches", u64),
                ("nodes_visited", u64), ("tris_tested", u64), ("objects_tested", u64), ("shadow_nodes_visited", u64), ("shadow_tris_tested", u64),
                ("shadow_objects_tested", u64), ("device_ms", C.c_double),
                ("trace_ms", C.c_double), ("shade_ms", C.c_double), ("shadow_ms", C.c_double)]


class SceneInfo(C.Structure):
    _fields_ = [("num_nodes", i32), ("num_meshes", i32), ("num_materials", i32), ("num_lights", i32), ("num_textures", i32),
                ("num_object_nodes", i32), ("num_faces", i64), ("num_bvh_nodes", i64), ("device_bytes", i64),
                ("camera", Camera), ("screen_u", f32 * 3), ("screen_v", f32 * 3), ("screen_a", f32 * 3)]


# numpy mirrors of qgRay / qgHit for bulk buffers
RAY_DTYPE = np.dtype([("p", "<f4", 3), ("dir", "<f4", 3), ("xp", "<f4", 3), ("xdir", "<f4", 3), ("yp", "<f4", 3),
                      ("ydir", "<f4", 3), ("has_diff", "<i4")])
HIT_DTYPE = np.dtype([("hit", "<i4"), ("node", "<i4"), ("face", "<i4"), ("mtl_id", "<i4"), ("front", "<i4"),
                      ("has_texture", "<i4"), ("z", "<f4"), ("p", "<f4", 3), ("N", "<f4", 3), ("uvw", "<f4", 3),
                      ("duvw", "<f4", 6)])
assert RAY_DTYPE.itemsize == 76 and HIT_DTYPE.itemsize == 88

# every symbol include/qaray_b200.h declares: name -> (restype, argtypes)
vp = C.c_void_p
PROTOTYPES = {
    "qg_init": (C.c_int, [C.c_int]),
    "qg_shutdown": (None, []),
    "qg_last_error": (C.c_char_p, []),
    "qg_scene_load_xml": (C.c_int, [C.c_char_p, C.c_char_p, C.POINTER(vp)]),
    "qg_host_scene_load_xml": (C.c_int, [C.c_char_p, C.c_char_p, C.POINTER(vp)]),
    "qg_host_scene_desc": (C.POINTER(SceneDesc), [vp]),
    "qg_host_scene_log": (C.c_char_p, [vp]),
    "qg_host_scene_destroy": (None, [vp]),
    "qg_scene_create": (C.c_int, [C.POINTER(SceneDesc), C.POINTER(vp)]),
    "qg_scene_destroy": (None, [vp]),
    "qg_scene_info": (C.c_int, [vp, C.POINTER(SceneInfo)]),
    "qg_scene_nodes": (C.c_int, [vp, C.POINTER(Node), i32]),
    "qg_trace_closest": (C.c_int, [vp, vp, C.c_size_t, vp]),
    "qg_mesh_bvh": (C.c_int, [vp, vp, vp, C.c_int64, vp, vp, C.c_int64, vp]),
    "qg_trace_any": (C.c_int, [vp, vp, vp, C.c_size_t, vp]),
    "qg_trace_any_noalloc": (C.c_int, [vp, vp, vp, C.c_size_t, vp, vp]),  # device pointers, stream; static scratch
    "qg_primary_rays": (C.c_int, [vp, vp, C.c_size_t, i32, i32, vp]),
    "qg_sample_radiance": (C.c_int, [vp, C.POINTER(RenderParams), vp, C.c_size_t, vp, vp]),
    "qg_render": (C.c_int, [vp, C.POINTER(RenderParams), vp, vp, vp, vp, vp, C.POINTER(RenderStats)]),
    "qg_render_accumulate": (C.c_int, [vp, C.POINTER(RenderParams), vp, vp, vp, C.POINTER(RenderStats)]),
    "qg_resolve": (C.c_int, [vp, C.POINTER(RenderParams), vp, vp, vp, vp, vp, vp, vp, vp]),
    "qg_shard_coverage": (C.c_int, [C.POINTER(Camera), C.POINTER(RenderParams), vp]),
    "qg_photon_maps_build": (C.c_int, [vp, C.POINTER(PhotonParams), C.POINTER(PhotonStats)]),
    "qg_photon_maps_clear": (None, [vp]),
    "qg_photon_map_read": (C.c_int, [vp, i32, vp, vp, vp, C.c_size_t, C.POINTER(C.c_size_t)]),
    "qg_photon_emit": (C.c_int, [vp, C.POINTER(PhotonParams), i32, u32, C.c_size_t, vp, vp]),
    "qg_photon_estimate": (C.c_int, [vp, i32, vp, vp, C.c_size_t, vp, vp]),
    "qg_photon_map_set": (C.c_int, [vp, i32, vp, vp, vp, C.c_size_t, f32, f32]),
    "qg_shard_size": (C.c_int, [C.POINTER(Camera), C.POINTER(RenderParams), C.POINTER(u64), C.POINTER(u64), C.POINTER(u64)]),
    "qg_save_png_rgb8": (C.c_int, [C.c_char_p, vp, i32, i32]),
    "qg_save_png_depth": (C.c_int, [C.c_char_p, vp, i32, i32]),
    "qg_save_png_gray8": (C.c_int, [C.c_char_p, vp, i32, i32]),
    "qg_save_png_sample_count": (C.c_int, [C.c_char_p, vp, i32, i32]),
    "qg_shard_params": (C.c_int, [C.POINTER(RenderParams), i32, i32, i32, C.POINTER(RenderParams)]),
    "qg_comm_unique_id": (C.c_int, [vp]),
    "qg_comm_create": (C.c_int, [vp, C.c_int, C.c_int, C.POINTER(vp)]),
    "qg_comm_info": (C.c_int, [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]),
    "qg_comm_destroy": (None, [vp]),
    "qg_comm_render": (C.c_int, [vp, vp, C.POINTER(RenderParams), i32, vp, vp, vp, C.POINTER(RenderStats)]),
    "qg_multi_create": (C.c_int, [C.POINTER(i32), i32, C.POINTER(vp)]),
    "qg_multi_load_xml": (C.c_int, [vp, C.c_char_p, C.c_char_p]),
    "qg_multi_create_scene": (C.c_int, [vp, C.POINTER(SceneDesc)]),
    "qg_multi_scene": (vp, [vp, i32]),
    "qg_multi_render": (C.c_int, [vp, C.POINTER(RenderParams), i32, C.c_char_p, vp, vp, vp, vp, vp, C.POINTER(RenderStats)]),
    "qg_multi_destroy": (None, [vp]),
}
QG_COMM_ID_BYTES = 128

_libs = {}


def load(counting=False):
    """dlopen the C-ABI library (the counting build when counting=True) and attach the prototypes."""
    path = LIB_COUNT_PATH if counting else LIB_PATH
    if path in _libs:
        return _libs[path]
    if not os.path.exists(path):
        raise RuntimeError("%s is missing: the CUDA extension has not been built (run __graft_entry__.build()); "
                           "qaray_b200 has no CPU fallback" % path)
    lib = C.CDLL(path)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _libs[path] = lib
    return lib


class QarayError(RuntimeError):
    def __init__(self, code, msg):
        RuntimeError.__init__(self, "qaray_b200 error %d: %s" % (code, msg))
        self.code = code


def check(lib, rc):
    if rc != QG_OK:
        raise QarayError(rc, (lib.qg_last_error() or b"").decode("utf-8", "replace"))
