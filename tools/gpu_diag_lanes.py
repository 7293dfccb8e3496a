"""Warp-fill diagnostics of wf_trace on a full workload (MCPT_DIAG=1 prints lanes per issue; needs FLAG_COUNT).
    python tools/gpu_diag_lanes.py <cornell-box|veach-mis|synthetic-N> <spp> [width height]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import montecarlopathtracing_b200 as M
name, spp = sys.argv[1], int(sys.argv[2])
w, h = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (0, 0)
ctx = M.Context(0)
if name.startswith("synthetic-"):
    from montecarlopathtracing_b200.synthetic import make_scene
    desc, keep, cam = make_scene(int(name.split("-")[1]), width=w or 1920, height=h or 1080)
    ctx.upload_scene(desc)
else:
    s = M.HostScene(os.path.join(ROOT, "scenes") + "/", name)
    ctx.upload_scene(s); cam = s.camera
ms = ctx.build_bvh(); ctx.set_seed(3)
ctx.render(cam, 8, w, h)
ctx.render(cam, spp, w, h)
st = ctx.last_stats
print("%s spp %d: build %.2f ms, integrate %.2f ms, %.1f Msamples/s, %.1f Mrays/s, %d trace launches" % (name, spp, ms, st.integrate_ms, st.samples / st.render_ms * 1e-3, st.rays / st.render_ms * 1e-3, st.trace_launches), flush=True)
if os.environ.get("MCPT_DIAG"):
    ctx.render(cam, spp, w, h, flags=M.FLAG_COUNT)
    st = ctx.last_stats
    print("  %.2f nodes/ray %.2f tris/ray %.3f rays/sample" % (st.node_visits / st.rays, st.tri_tests / st.rays, st.rays / st.samples))
