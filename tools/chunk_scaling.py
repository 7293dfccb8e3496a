"""Developer probe (GPU box): what one rank of an R-GPU strong-scaling run costs, measured on ONE GPU.
Rank 0 of R integrates chunks [0, 64/R) of cornell-box 3840x2160 x 1024 spp; efficiency estimate = t(64 chunks) / (R * t(64/R chunks)).
    python tools/chunk_scaling.py [width height spp]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import montecarlopathtracing_b200 as M  # noqa: E402

w, h, spp = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (3840, 2160, 1024)
s = M.HostScene(os.path.join(ROOT, "scenes") + "/", "cornell-box")
ctx = M.Context(0)
ctx.upload_scene(s)
ctx.build_bvh()
ctx.set_seed(2026)
part = torch.empty(h * w * 3, dtype=torch.float32, device="cuda")
ctx.render_chunks(s.camera, spp, 0, 4, part.data_ptr(), w, h)   # warm-up: allocations
t64 = None
for R in (1, 2, 4, 8):
    n = 64 // R
    best = 1e30
    for _ in range(2):
        st = ctx.render_chunks(s.camera, spp, 0, n, part.data_ptr(), w, h)
        best = min(best, st.render_ms)
    if R == 1:
        t64 = best
    print("R=%d: chunks [0,%d) %.1f ms, %d trace launches, %.1f Msamples/s per GPU, strong-scaling efficiency estimate %.3f" %
          (R, n, best, st.trace_launches, st.samples / best * 1e-3, t64 / (R * best)), flush=True)
ctx.close()
