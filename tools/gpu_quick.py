"""Developer loop (GPU box): quick throughput table for both integrators + a bit-equality check between them."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import montecarlopathtracing_b200 as M  # noqa: E402

ctx = M.Context(0)
modes = [int(a) for a in sys.argv[1:]] or [M.MODE_MEGAKERNEL, M.MODE_WAVEFRONT]
for name, spp in (("cornell-box", 64), ("veach-mis", 32)):
    s = M.HostScene(os.path.join(ROOT, "scenes") + "/", name)
    ctx.upload_scene(s)
    ctx.build_bvh()
    ctx.set_seed(3)
    imgs = {}
    for mode in modes:
        ctx.render(s.camera, 8, mode=mode)                       # warm-up (allocations)
        imgs[mode] = ctx.render(s.camera, spp, mode=mode)
        st = ctx.last_stats
        print("%s mode %d spp %d: integrate %.2f ms (trace %.2f ms in %d launches), %.1f Msamples/s, %.1f Mrays/s" %
              (name, mode, spp, st.integrate_ms, st.trace_ms, st.trace_launches, st.samples / st.render_ms * 1e-3,
               st.rays / st.render_ms * 1e-3), flush=True)
    ctx.render(s.camera, 8, mode=modes[-1], flags=M.FLAG_COUNT)
    st = ctx.last_stats
    print("  counts: %.2f nodes/ray %.2f tris/ray %.3f rays/sample" % (st.node_visits / st.rays, st.tri_tests / st.rays, st.rays / st.samples))
    if len(imgs) == 2:
        a, b = imgs.values()
        print("  modes bit-identical:", bool(np.array_equal(a, b)), "mean", a.mean(axis=(0, 1)))
ctx.close()
