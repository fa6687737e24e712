"""The reference's output set (src/fb/framebuffer.cpp:62-143, written by Renderer_MPI::Render, Renderer_MPI.cpp:132-138 and
:188-193): colour PNG, depth PNG (normalised, far = black) and the sample-count PNG stretched to 0..255. The planes are
written by the library's own PNG encoder and decoded here with PIL — no GPU needed."""
import ctypes as C
import os

import numpy as np
import pytest
from PIL import Image

from qaray_b200 import capi

BIGFLOAT = 1.0e30


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def test_colour_png_round_trip(tmp_path):
    lib = capi.load()
    rng = np.random.default_rng(7)
    for (W, H) in ((37, 23), (1, 1), (256, 3)):
        img = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
        path = str(tmp_path / ("c_%dx%d.png" % (W, H)))
        capi.check(lib, lib.qg_save_png_rgb8(os.fsencode(path), _p(img), W, H))
        back = np.asarray(Image.open(path).convert("RGB"))
        assert back.shape == (H, W, 3) and np.array_equal(back, img)


def test_sample_count_png_is_stretched_like_the_reference(tmp_path):
    """FrameBuffer::ComputeSampleCountImage (framebuffer.cpp:88-107): 255 * (s - smin) / (smax - smin) when the counts differ,
    all zero when every pixel took the same number of samples"""
    lib = capi.load()
    W, H = 16, 8
    sc = (np.arange(W * H, dtype=np.int64) % 7 * 10 + 40).astype(np.uint8).reshape(H, W)   # 40 .. 100
    path = str(tmp_path / "s.png")
    capi.check(lib, lib.qg_save_png_sample_count(os.fsencode(path), _p(sc), W, H))
    back = np.asarray(Image.open(path).convert("L"))
    smin, smax = int(sc.min()), int(sc.max())
    want = ((sc.astype(np.int32) - smin) * 255 // (smax - smin)).astype(np.uint8)
    assert np.array_equal(back, want)
    flat = np.full((H, W), 77, np.uint8)
    capi.check(lib, lib.qg_save_png_sample_count(os.fsencode(path), _p(flat), W, H))
    assert (np.asarray(Image.open(path).convert("L")) == 0).all()


def test_depth_png_is_normalised_like_the_reference(tmp_path):
    """FrameBuffer::ComputeZBufferImage (framebuffer.cpp:62-86): zmin / zmax over the pixels that hit something
    (z != BIGFLOAT); hit pixels map to (zmax - z) / (zmax - zmin) * 255, misses to 0"""
    lib = capi.load()
    W, H = 8, 4
    z = np.linspace(2.0, 9.0, W * H).astype(np.float32).reshape(H, W)
    z[0, 0] = BIGFLOAT
    z[3, 7] = BIGFLOAT
    path = str(tmp_path / "z.png")
    capi.check(lib, lib.qg_save_png_depth(os.fsencode(path), _p(z), W, H))
    back = np.asarray(Image.open(path).convert("L")).astype(np.int32)
    hit = z != BIGFLOAT
    zmin, zmax = z[hit].min(), z[hit].max()
    want = np.where(hit, ((zmax - np.where(hit, z, zmax)) / (zmax - zmin) * 255).astype(np.int32), 0)
    assert np.abs(back - want).max() <= 1
    assert back[0, 0] == 0 and back[3, 7] == 0


def test_png_writers_reject_bad_arguments(tmp_path):
    lib = capi.load()
    a = np.zeros(12, np.uint8)
    assert lib.qg_save_png_rgb8(None, _p(a), 2, 2) == capi.QG_EINVAL
    assert lib.qg_save_png_rgb8(os.fsencode(str(tmp_path / "x.png")), _p(a), 0, 2) == capi.QG_EINVAL
    assert lib.qg_save_png_rgb8(os.fsencode(str(tmp_path / "no_such_dir" / "x.png")), _p(a), 2, 2) == capi.QG_EIO
