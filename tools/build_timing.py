"""Developer probe (GPU box): steady-state LBVH build time (all launches, CUDA events inside mcpt_build_bvh) of the shipped scenes."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import montecarlopathtracing_b200 as M  # noqa: E402

ctx = M.Context(0)
for name in sys.argv[1:] or ["cornell-box", "veach-mis", "glass-box"]:
    if name.startswith("synthetic-"):
        from montecarlopathtracing_b200.synthetic import make_scene
        desc, keep, cam = make_scene(int(name.split("-")[1]))
        ctx.upload_scene(desc)
    else:
        s = M.HostScene(os.path.join(ROOT, "scenes") + "/", name)
        ctx.upload_scene(s)
    ms = [ctx.build_bvh() for _ in range(30)][5:]
    nodes, _, _ = ctx.get_wide_tree()
    print("%s: %d triangles, %d wide nodes, build min %.3f ms, median %.3f ms" % (name, ctx.n_faces, len(nodes), min(ms), float(np.median(ms))), flush=True)
ctx.close()
