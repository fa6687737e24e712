"""ctypes binding of oracle/_ref/libqaref_parity.so — TEST INFRASTRUCTURE, never imported by the product.

The shared object is the reference renderer itself (wilsonCernWq/qaray) compiled by oracle/Makefile from the
sources under /root/reference plus oracle/ref_driver.cpp. It keeps global state (the reference's `qaray::scene`
singleton, src/scene/scene.cpp:76), so one process holds one scene at a time; use `run_isolated` to get a fresh
process per scene.
"""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
REF_SO = os.path.join(REF_DIR, "libqaref_parity.so")
REF_PERF = os.path.join(REF_DIR, "qaref_perf")
REF_INPUTS = os.path.join(REF_DIR, "inputs")

HIT_DTYPE = np.dtype([
    ("hit", "<i4"), ("node", "<i4"), ("face", "<i4"), ("mtlID", "<i4"), ("front", "<i4"), ("hasTexture", "<i4"),
    ("z", "<f4"), ("p", "<f4", 3), ("N", "<f4", 3), ("uvw", "<f4", 3), ("duvw", "<f4", 6),
])
assert HIT_DTYPE.itemsize == 88


# qgRenderParams.material_class -> oracle build variant
VARIANTS = {0: "", 1: "pathtracing", 2: "montecarlogi", 3: "blinn_basic", 4: "phong_basic"}


def variant_so(variant=""):
    return REF_SO if not variant else os.path.join(REF_DIR, "libqaref_parity_%s.so" % variant)


def available(variant=""):
    return os.path.exists(variant_so(variant))


class Ref:
    """One loaded reference scene (process-global)."""

    def __init__(self, variant=""):
        """variant: "" = the reference as shipped (MtlBlinn = MtlBlinn_PhotonMap), or one of VARIANTS = the same build with
        the alias at src/materials/materials.h:57 switched to that class"""
        so = variant_so(variant)
        if not os.path.exists(so):
            raise RuntimeError("%s missing: run `make -C oracle ref` where /root/reference exists" % so)
        self.lib = C.CDLL(so)
        L = self.lib
        L.qaref_load.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        L.qaref_load.restype = C.c_int
        L.qaref_set_params.argtypes = [C.c_int] * 5
        L.qaref_seed.argtypes = [C.c_uint32]
        L.qaref_rng_state.restype = C.c_uint32
        L.qaref_render.restype = C.c_double
        L.qaref_render.argtypes = [C.c_void_p] * 3
        L.qaref_trace_closest.argtypes = [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.qaref_trace_shadow.argtypes = [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.qaref_sample.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.qaref_sample.restype = C.c_int
        L.qaref_render_linear_serial.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        L.qaref_primary_ray.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.qaref_counters.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.qaref_samples.argtypes = [C.c_int64, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
        L.qaref_seed_path.argtypes = [C.c_uint32]
        L.qaref_path_key.argtypes = [C.c_uint32] * 3
        L.qaref_path_key.restype = C.c_uint32
        self.W = self.H = 0

    def set_params(self, spp_min, spp_max, bounce=5, srgb=True, threads=0):
        self.lib.qaref_set_params(int(spp_min), int(spp_max), int(bounce), int(bool(srgb)), int(threads))

    def load(self, xml_path, asset_dir, silent=True):
        ok = self.lib.qaref_load(os.path.abspath(xml_path).encode(), os.path.abspath(asset_dir).encode(), int(silent))
        if not ok:
            raise RuntimeError("reference LoadScene failed: %s" % xml_path)
        cam = self.camera()
        self.W, self.H = int(cam[12]), int(cam[13])
        return self

    def seed(self, s):
        self.lib.qaref_seed(C.c_uint32(s))

    def camera(self):
        out = np.zeros(14, np.float32)
        self.lib.qaref_camera(out.ctypes.data_as(C.c_void_p))
        return out

    def screen(self):
        out = np.zeros(9, np.float32)
        self.lib.qaref_screen(out.ctypes.data_as(C.c_void_p))
        return out

    def nodes(self):
        n = self.lib.qaref_num_nodes()
        meta = np.zeros((n, 4), np.int32)
        xf = np.zeros((n, 21), np.float32)
        self.lib.qaref_dump_nodes(meta.ctypes.data_as(C.c_void_p), xf.ctypes.data_as(C.c_void_p))
        return meta, xf

    def primary_rays(self, pixels, s):
        """pixels: (n,2) int (i,j) -> (n,18) float32 c/x/y rays"""
        pixels = np.asarray(pixels, np.int32)
        out = np.zeros((len(pixels), 18), np.float32)
        for k, (i, j) in enumerate(pixels):
            self.lib.qaref_primary_ray(int(i), int(j), int(s), out[k].ctypes.data_as(C.c_void_p))
        return out

    def trace_closest(self, rays18, has_diff=None):
        rays18 = np.ascontiguousarray(rays18, np.float32).reshape(-1, 18)
        n = len(rays18)
        hd = None
        if has_diff is not None:
            hd = np.ascontiguousarray(has_diff, np.uint8)
        out = np.zeros(n, HIT_DTYPE)
        self.lib.qaref_trace_closest(n, rays18.ctypes.data_as(C.c_void_p),
                                     hd.ctypes.data_as(C.c_void_p) if hd is not None else None,
                                     out.ctypes.data_as(C.c_void_p))
        return out

    def trace_shadow(self, rays6, tmax):
        rays6 = np.ascontiguousarray(rays6, np.float32).reshape(-1, 6)
        tmax = np.ascontiguousarray(tmax, np.float32)
        out = np.zeros(len(rays6), np.uint8)
        self.lib.qaref_trace_shadow(len(rays6), rays6.ctypes.data_as(C.c_void_p), tmax.ctypes.data_as(C.c_void_p),
                                    out.ctypes.data_as(C.c_void_p))
        return out

    def sample(self, i, j, s):
        rgb = np.zeros(3, np.float32)
        z = C.c_float(0)
        hit = self.lib.qaref_sample(int(i), int(j), int(s), rgb.ctypes.data_as(C.c_void_p), C.byref(z))
        return hit, rgb, z.value

    def samples(self, pix_ijs, seed):
        """radiance of single samples (i,j,s) with the counter-based stream of (pixel, s, seed) (P4 patch)"""
        pix = np.ascontiguousarray(pix_ijs, np.int32).reshape(-1, 3)
        rgb = np.zeros((len(pix), 3), np.float32)
        z = np.zeros(len(pix), np.float32)
        hit = np.zeros(len(pix), np.int32)
        self.lib.qaref_samples(len(pix), pix.ctypes.data_as(C.c_void_p), C.c_uint32(seed), rgb.ctypes.data_as(C.c_void_p),
                               z.ctypes.data_as(C.c_void_p), hit.ctypes.data_as(C.c_void_p))
        return rgb, z, hit

    def render_linear_serial(self, spp):
        rgb = np.zeros((self.H, self.W, 3), np.float32)
        depth = np.zeros((self.H, self.W), np.float32)
        self.lib.qaref_render_linear_serial(int(spp), rgb.ctypes.data_as(C.c_void_p), depth.ctypes.data_as(C.c_void_p))
        return rgb, depth

    def render_adaptive_serial(self, spp_min, spp_max):
        """PixelRender's loop with SuperSamplerHalton(sppMin, sppMax): linear means, depth, samples taken per pixel"""
        rgb = np.zeros((self.H, self.W, 3), np.float32)
        depth = np.zeros((self.H, self.W), np.float32)
        count = np.zeros((self.H, self.W), np.int32)
        self.lib.qaref_render_adaptive_serial.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        self.lib.qaref_render_adaptive_serial(int(spp_min), int(spp_max), rgb.ctypes.data_as(C.c_void_p),
                                              depth.ctypes.data_as(C.c_void_p), count.ctypes.data_as(C.c_void_p))
        return rgb, depth, count

    def render(self):
        """The reference's real ThreadRender. Returns (rgb8, depth, sampleCount, seconds)."""
        rgb8 = np.zeros((self.H, self.W, 3), np.uint8)
        depth = np.zeros((self.H, self.W), np.float32)
        sc = np.zeros((self.H, self.W), np.uint8)
        dt = self.lib.qaref_render(rgb8.ctypes.data_as(C.c_void_p), depth.ctypes.data_as(C.c_void_p),
                                   sc.ctypes.data_as(C.c_void_p))
        return rgb8, depth, sc, dt

    def counters(self, reset=True):
        a = C.c_uint64(0)
        b = C.c_uint64(0)
        self.lib.qaref_counters(C.byref(a), C.byref(b), int(reset))
        return a.value, b.value
