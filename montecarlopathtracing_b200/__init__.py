"""mcpt-b200: B200-native hot path of Arieys/MonteCarloPathTracing behind its render_scene entry point.

The product is libmcpt.so (hand-written sm_100a CUDA kernels + the C++ host shell, C ABI in
include/mcpt.h).  This package is the thin Python mirror used by the tests and bench.py.
"""
from ._capi import (FLAG_ANYHIT_SHADOW, FLAG_COUNT, FLAG_JITTER, FLAG_MIS_BALANCE, FLAG_NO_PRIMARY_REUSE, FLAG_NO_SHADOW_EARLY_OUT, MODE_MEGAKERNEL, MODE_WAVEFRONT, NUM_CHUNKS, Camera, McptError,
                    SceneDesc, Stats)
from .api import Context, HostScene, MultiContext, comm_unique_id, encode_png_static, png_size, render_scene, write_png

__all__ = ["Context", "MultiContext", "comm_unique_id", "HostScene", "render_scene", "write_png", "png_size", "encode_png_static", "Camera", "SceneDesc", "Stats", "McptError", "NUM_CHUNKS",
           "MODE_MEGAKERNEL", "MODE_WAVEFRONT", "FLAG_COUNT", "FLAG_NO_PRIMARY_REUSE", "FLAG_NO_SHADOW_EARLY_OUT", "FLAG_JITTER", "FLAG_ANYHIT_SHADOW", "FLAG_MIS_BALANCE"]
