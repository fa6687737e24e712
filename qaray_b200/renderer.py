"""Host-side mirror of the reference's renderer interface over the C ABI.

The reference drives a render as (src/main.cpp:46-59)

    renderer->Init(); LoadScene(file); renderer->ComputeScene(renderImage, scene); renderer->Render(); renderer->Terminate();

with `RendererParam` (src/renderers/renderer.h:47-68) carrying spp / sRGB flags and `Material::maxBounce` the bounce
budget. `Renderer_B200` keeps those names and argument meanings; every compute call goes through
include/qaray_b200.h into the CUDA library. There is no CPU path here: without the built library or without a GPU
the calls raise.
"""
import ctypes as C
import os

import numpy as np

from . import capi


class RendererParam(object):
    """RendererParam (src/renderers/renderer.h:47-68) + Material::maxBounce (src/core/material.cpp:31)."""

    def __init__(self, spp=16, max_bounce=5, srgb=True, seed=0x5EED):
        self.sppMin = self.sppMax = int(spp)
        self.useSRGB = bool(srgb)
        self.maxBounce = int(max_bounce)
        self.seed = int(seed)
        self.width = 0   # 0 = the scene camera's <width>/<height>
        self.height = 0
        self.batch_paths = 0
        self.sort_by_material = 0
        self.material_class = 0  # QG_CLASS_*: the class the reference's `using MtlBlinn = ...` alias selects (materials.h:57-61)

    def SetSPPMin(self, n):
        self.sppMin = int(n)

    def SetSPPMax(self, n):
        self.sppMax = int(n)

    def SetSRGBFlag(self, f):
        self.useSRGB = bool(f)

    def to_c(self, tile_rank=0, tile_ranks=1, sample_begin=0, sample_end=0):
        if self.sppMin > self.sppMax:
            raise ValueError("sppMin > sppMax")
        return capi.RenderParams(self.sppMax, self.maxBounce, int(self.useSRGB), self.seed & 0xFFFFFFFF, self.width,
                                 self.height, tile_rank, tile_ranks, sample_begin, sample_end, self.batch_paths,
                                 self.sort_by_material, self.material_class, self.sppMin if self.sppMin < self.sppMax else 0)


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Scene(object):
    """A scene resident in HBM (qgScene). Construct with Scene.load(xml, asset_dir)."""

    def __init__(self, lib, handle):
        self.lib = lib
        self.handle = handle

    @classmethod
    def load(cls, xml_path, asset_dir="", counting=False):
        lib = capi.load(counting)
        h = C.c_void_p()
        capi.check(lib, lib.qg_scene_load_xml(os.fsencode(xml_path), os.fsencode(asset_dir or ""), C.byref(h)))
        return cls(lib, h)

    def close(self):
        if self.handle:
            self.lib.qg_scene_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self):
        o = capi.SceneInfo()
        capi.check(self.lib, self.lib.qg_scene_info(self.handle, C.byref(o)))
        return o

    def nodes(self):
        n = self.info().num_nodes
        arr = (capi.Node * n)()
        capi.check(self.lib, self.lib.qg_scene_nodes(self.handle, arr, n))
        return arr

    # Scene::TraceNodeNormal over a ray buffer
    def trace_closest(self, rays):
        rays = np.ascontiguousarray(rays, capi.RAY_DTYPE)
        out = np.zeros(len(rays), capi.HIT_DTYPE)
        capi.check(self.lib, self.lib.qg_trace_closest(self.handle, _ptr(rays), len(rays), _ptr(out)))
        return out

    # Scene::TraceNodeShadow over a ray buffer
    def trace_any(self, rays6, tmax):
        rays6 = np.ascontiguousarray(rays6, np.float32).reshape(-1, 6)
        tmax = np.ascontiguousarray(tmax, np.float32)
        out = np.zeros(len(rays6), np.uint8)
        capi.check(self.lib, self.lib.qg_trace_any(self.handle, _ptr(rays6), _ptr(tmax), len(rays6), _ptr(out)))
        return out

    def primary_rays(self, pix_ijs, width=0, height=0):
        pix = np.ascontiguousarray(pix_ijs, np.int32).reshape(-1, 3)
        out = np.zeros(len(pix), capi.RAY_DTYPE)
        capi.check(self.lib, self.lib.qg_primary_rays(self.handle, _ptr(pix), len(pix), width, height, _ptr(out)))
        return out

    def sample_radiance(self, param, pix_ijs):
        """radiance of single samples (i, j, s): the body of PixelRender's loop, before accumulation"""
        pix = np.ascontiguousarray(pix_ijs, np.int32).reshape(-1, 3)
        rgb = np.zeros((len(pix), 3), np.float32)
        z = np.zeros(len(pix), np.float32)
        p = param.to_c()
        capi.check(self.lib, self.lib.qg_sample_radiance(self.handle, C.byref(p), _ptr(pix), len(pix), _ptr(rgb), _ptr(z)))
        return rgb, z


class Renderer_B200(object):
    """Drop-in for the reference's batch renderer (Renderer_MPI without MPI): same call sequence, same outputs."""

    def __init__(self, param=None):
        self.param = param or RendererParam()
        self.scene = None
        self.mpiRank, self.mpiSize = 0, 1   # Renderer::mpiRank / mpiSize (renderer.h:85-86)
        self.colorBuffer = self.depthBuffer = self.sampleCountBuffer = self.maskBuffer = self.linearBuffer = None
        self.stats = None
        self.pixelW = self.pixelH = 0
        self._asset_dir = ""

    def Init(self, device=0):
        lib = capi.load()
        capi.check(lib, lib.qg_init(int(device)))

    def LoadScene(self, filename, asset_dir=""):
        """LoadScene (src/parser/xmlload.cpp:71): returns 1 on success, 0 on failure like the reference."""
        try:
            self.scene = Scene.load(filename, asset_dir)
        except capi.QarayError as e:
            if e.code == capi.QG_EIO:
                print(str(e))
                return 0
            raise
        return 1

    def ComputeScene(self):
        cam = self.scene.info().camera
        self.pixelW = self.param.width or cam.width
        self.pixelH = self.param.height or cam.height

    def Render(self, want_linear=True):
        """Renderer::Render: fills colorBuffer (RGB8), depthBuffer, sampleCountBuffer, maskBuffer (+ fp32 linear means)."""
        W, H = self.pixelW, self.pixelH
        self.colorBuffer = np.zeros((H, W, 3), np.uint8)
        self.depthBuffer = np.zeros((H, W), np.float32)
        self.sampleCountBuffer = np.zeros((H, W), np.uint8)
        self.maskBuffer = np.zeros((H, W), np.uint8)
        self.linearBuffer = np.zeros((H, W, 3), np.float32) if want_linear else None
        p = self.param.to_c(self.mpiRank, self.mpiSize)
        st = capi.RenderStats()
        lib = self.scene.lib
        capi.check(lib, lib.qg_render(self.scene.handle, C.byref(p), _ptr(self.linearBuffer), _ptr(self.colorBuffer),
                                      _ptr(self.depthBuffer), _ptr(self.sampleCountBuffer), _ptr(self.maskBuffer), C.byref(st)))
        self.stats = st
        return st

    def SaveImages(self, prefix=""):
        """the PNG set of Renderer_MPI::Render (src/renderers/Renderer_MPI.cpp:134-138,188-193)"""
        lib = self.scene.lib
        W, H = self.pixelW, self.pixelH
        capi.check(lib, lib.qg_save_png_rgb8(os.fsencode(prefix + "colorBuffer.png"), _ptr(self.colorBuffer), W, H))
        capi.check(lib, lib.qg_save_png_depth(os.fsencode(prefix + "depthBuffer.png"), _ptr(self.depthBuffer), W, H))
        capi.check(lib, lib.qg_save_png_sample_count(os.fsencode(prefix + "sampleBuffer.png"), _ptr(self.sampleCountBuffer), W, H))

    def Terminate(self):
        if self.scene is not None:
            self.scene.close()
            self.scene = None
