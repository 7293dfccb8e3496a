#!/usr/bin/env python
"""Launches every kernel of the path a few times on one scene, for `ncu --set full` (profiles/r2_roofline.md): scene ingest
(K16-K18), LBVH build (K1-K4b + CUB), seam 2 (K5), raygen (K7), a short wavefront render (K8, K9), the megakernel (K6), the
plane sum (K12), the PNG stage (K13-K15) and the bandwidth sweep.  wf_trace / wf_shade are better judged on a mid-render launch
of a full-size render (tools/profile_render.py with --launch-skip); here they only appear with their first iterations.

    ncu --set full --import-source on --clock-control none -f -o gpurun_out/r2_all_<scene> python tools/profile_all_kernels.py --scene cornell-box
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import montecarlopathtracing_b200 as M  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--scene", default="cornell-box")
ap.add_argument("--spp", type=int, default=2)
a = ap.parse_args()
ctx = M.Context(0)
scenes = os.path.join(ROOT, "scenes") + "/"
if a.scene.startswith("synthetic"):
    from montecarlopathtracing_b200.synthetic import make_scene
    desc, keep, cam = make_scene({"synthetic-5M": 5_000_000, "synthetic-1M": 1_000_000}[a.scene], width=1920, height=1080)
    ctx.upload_scene(desc)
    w, h = 1920, 1080
else:
    cam, _ = ctx.load_scene_gpu(scenes, a.scene)          # K16-K18
    w, h = cam.width, cam.height
ctx.build_bvh()                                          # K1, sort, K2, K3a, K3b, PLOC, K4b
from make_golden import camera_rays
rays = camera_rays(cam, w, h, 8).astype(np.float32)
ctx.trace(rays)                                          # K5
ctx.set_seed(1)
ctx.render(cam, a.spp, w, h, mode=M.MODE_WAVEFRONT)      # K7, K8, K9, K12
ctx.render(cam, a.spp, w, h, mode=M.MODE_MEGAKERNEL)     # K6
import tempfile
ctx.render_png(cam, 1, os.path.join(tempfile.mkdtemp(), "x.png"), w, h)   # K13-K15
ctx.measure_read_bandwidth(32 << 20, 4)
print("done", a.scene, ctx.n_faces)
ctx.close()
