"""qaray_b200 — B200-native drop-in for the per-pixel Monte-Carlo loop of wilsonCernWq/qaray.

The product is the CUDA library behind include/qaray_b200.h (qaray_b200/libqaray_b200.so, sources in csrc/);
this package is the thin host-side mirror of the reference's renderer interface over that C ABI.
"""
from .capi import QarayError, load  # noqa: F401
from .renderer import Renderer_B200, RendererParam, Scene  # noqa: F401
