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
    for mode in (M.MODE_MEGAKERNEL, M.MODE_WAVEFRONT):
        for spp2 in (16, 64):
            ctx.render(s.camera, spp2, mode=mode)
            st = ctx.last_stats
            print("%s mode %d spp %d: render %.2f ms, %.1f Msamples/s, %.1f Mrays/s, build %.3f ms, launches %d" %
                  (name, mode, spp2, st.render_ms, st.samples / st.render_ms * 1e-3, st.rays / st.render_ms * 1e-3, ms,
                   st.kernel_launches), flush=True)
    ctx.render(s.camera, 16, flags=M.FLAG_COUNT)
    st = ctx.last_stats
    print("%s counts: %.2f nodes/ray %.2f tris/ray %.3f rays/sample" %
          (name, st.node_visits / st.rays, st.tri_tests / st.rays, st.rays / st.samples), flush=True)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from scene_util import random_soup  # noqa: E402
from montecarlopathtracing_b200.api import make_desc  # noqa: E402
desc0, keep = random_soup(400, seed=21, extent=1.2, size=0.35)
mat = np.arange(400, dtype=np.int32) % 4
mats = [{"kd": (0.6, 0.5, 0.4)}, {"kd": (0.1, 0.15, 0.4), "ks": (0.8, 0.8, 1.0), "Ns": 30.0},
        {"kd": (0.2, 0.2, 0.2), "ks": (0.5, 0.5, 0.5), "Ns": 5.0, "Ni": 1.5}, {"kd": (0, 0, 0)}]
desc, keep2 = make_desc(keep[0], keep[1], keep[2], mat, mats, [{"mat_id": 3, "radiance": (15.0, 12.0, 9.0)}])
cam = M.Camera()
cam.eye[:] = (0.3, 0.3, 6.0); cam.lookat[:] = (0.3, 0.3, 5.0); cam.up[:] = (0, 1, 0)
cam.fovy = 35.0; cam.width, cam.height = 80, 64
ctx.upload_scene(desc); ctx.build_bvh(); ctx.set_seed(77)
res["soup_img"] = ctx.render(cam, 8)
np.savez_compressed(os.path.join(OUT, "diag.npz"), **res)
ctx.close()
