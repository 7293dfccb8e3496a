"""Smallest run that touches every kernel (for compute-sanitizer): build (radix tree + PLOC incl. the global iterations:
> 1024 triangles), batch trace, megakernel and wavefront renders with several lights."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import montecarlopathtracing_b200 as M  # noqa: E402
from make_golden import camera_rays  # noqa: E402

ctx = M.Context(0)
s = M.HostScene(os.path.join(ROOT, "scenes") + "/", "veach-mis")      # 3812 triangles, 5 lights
ctx.upload_scene(s)
ctx.build_bvh()
hits = ctx.trace(camera_rays(s.camera, 64, 48, 4).astype(np.float32))
a = ctx.render(s.camera, 3, 40, 30, mode=M.MODE_MEGAKERNEL)
b = ctx.render(s.camera, 3, 40, 30, mode=M.MODE_WAVEFRONT, flags=M.FLAG_COUNT)
print("san_small ok:", int((hits["tri"] >= 0).sum()), "hits; modes identical:", bool(np.array_equal(a, b)))
ctx.close()
