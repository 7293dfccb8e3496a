"""Developer diagnostics (run on the GPU box): dumps GPU renders / hit batches into gpurun_out/ for offline comparison
with the oracle in the build container.  Not part of the product or the tests."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import montecarlopathtracing_b200 as M  # noqa: E402
from make_golden import camera_rays  # noqa: E402

OUT = os.path.join(ROOT, "gpurun_out")
os.makedirs(OUT, exist_ok=True)
ctx = M.Context(0)
res = {}
for name, w, h, spp in (("cornell-box", 96, 96, 8), ("veach-mis", 96, 72, 8)):
    s = M.HostScene(os.path.join(ROOT, "scenes") + "/", name)
    ctx.upload_scene(s)
    ms = ctx.build_bvh()
    ctx.set_seed(1234)
    res[name + "_img"] = ctx.render(s.camera, spp, w, h)
    rays = camera_rays(s.camera, s.camera.width, s.camera.height, 16)
    res[name + "_hits"] = ctx.trace(rays.astype(np.float32))
    # throughput table per mode/flag at the .camera resolution
    for spp2 in (16, 64):
        ctx.render(s.camera, spp2)
        st = ctx.last_stats
        print("%s spp %d: render %.2f ms, %.1f Msamples/s, %.1f Mrays/s, build %.3f ms" %
              (name, spp2, st.render_ms, st.samples / st.render_ms * 1e-3, st.rays / st.render_ms * 1e-3, ms), flush=True)
    ctx.render(s.camera, 16, flags=M.FLAG_COUNT)
    st = ctx.last_stats
    print("%s counts: %.2f nodes/ray %.2f tris/ray %.3f rays/sample" %
          (name, st.node_visits / st.rays, st.tri_tests / st.rays, st.rays / st.samples), flush=True)
np.savez_compressed(os.path.join(OUT, "diag.npz"), **res)
ctx.close()
