"""Warp-fill diagnostics of wf_trace on a full workload (MCPT_DIAG=1 prints lanes per issue; needs FLAG_COUNT)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import montecarlopathtracing_b200 as M
name, spp = sys.argv[1], int(sys.argv[2])
w, h = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (0, 0)
ctx = M.Context(0)
s = M.HostScene(os.path.join(ROOT, "scenes") + "/", name)
ctx.upload_scene(s); ctx.build_bvh(); ctx.set_seed(3)
ctx.render(s.camera, 8, w, h)
ctx.render(s.camera, spp, w, h)
st = ctx.last_stats
print("%s spp %d: integrate %.2f ms, %.1f Msamples/s, %.1f Mrays/s, %d trace launches" % (name, spp, st.integrate_ms, st.samples / st.render_ms * 1e-3, st.rays / st.render_ms * 1e-3, st.trace_launches), flush=True)
if os.environ.get("MCPT_DIAG"):
    ctx.render(s.camera, spp, w, h, flags=M.FLAG_COUNT)
