#!/usr/bin/env python
"""One build + one render of a bench workload through the C ABI, for runs under ncu (no warm-up, nothing else launched).
Prints one JSON line with the device-counted rays and samples, which tools/ncu_to_metrics.py divides the summed ncu counters by.

    ncu --metrics ... --csv --log-file gpurun_out/x.csv python tools/profile_render.py --workload cornell-box-1024x1024 --spp 128 \
        > gpurun_out/x.json
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import montecarlopathtracing_b200 as M  # noqa: E402

SHAPES = {"cornell-box-1024x1024": ("cornell-box", 0, 0), "cornell-box-3840x2160": ("cornell-box", 3840, 2160),
          "veach-mis-1200x900": ("veach-mis", 0, 0), "synthetic-5M-1920x1080": ("synthetic-5M", 1920, 1080),
          "synthetic-1M-1920x1080": ("synthetic-1M", 1920, 1080)}

ap = argparse.ArgumentParser()
ap.add_argument("--workload", required=True, choices=sorted(SHAPES))
ap.add_argument("--spp", type=int, required=True)
ap.add_argument("--mode", default="wavefront")
ap.add_argument("--renders", type=int, default=1)
ap.add_argument("--builds", type=int, default=1)
a = ap.parse_args()
scene, w, h = SHAPES[a.workload]
ctx = M.Context(0)
if scene.startswith("synthetic"):
    from montecarlopathtracing_b200.synthetic import make_scene
    desc, keep, cam = make_scene({"synthetic-5M": 5_000_000, "synthetic-1M": 1_000_000}[scene], width=w, height=h)
    ctx.upload_scene(desc)
else:
    s = M.HostScene(os.path.join(ROOT, "scenes") + "/", scene)
    cam = s.camera
    ctx.upload_scene(s)
w, h = w or cam.width, h or cam.height
ctx.set_seed(2026)
for _ in range(a.builds):
    ms = ctx.build_bvh()
ctx.comm_init(1, 0, bytes(128))
mode = M.MODE_WAVEFRONT if a.mode == "wavefront" else M.MODE_MEGAKERNEL
rays = samples = 0.0
for _ in range(a.renders):
    ctx.render_dist(cam, a.spp, w, h, mode=mode, to_host=False)
    st = ctx.last_stats
    rays += st.rays; samples += st.samples
print(json.dumps({"workload": a.workload, "spp": a.spp, "mode": a.mode, "rays": rays, "samples": samples, "build_ms": ms,
                  "render_ms": st.render_ms, "trace_launches": st.trace_launches, "kernel_launches": st.kernel_launches,
                  "build_launches": st.build_launches}))
ctx.close()
