"""Phase timing of the GPU ingest on a shipped scene (MCPT_INGEST_TIMING=1 prints the phases), in bench.py's setting:
torch loaded, the context on torch's stream, wavefront state allocated by a render."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import montecarlopathtracing_b200 as M
name = sys.argv[1] if len(sys.argv) > 1 else "cornell-box"
scenes = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scenes") + "/"
ctx = M.Context(0)
if "--torch-stream" in sys.argv:
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
s = M.HostScene(scenes, name)
ctx.upload_scene(s); ctx.build_bvh(); ctx.render(s.camera, 4)
for i in range(4):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    cam, ms = ctx.load_scene_gpu(scenes, name)
    torch.cuda.synchronize()
    print("load_scene_gpu %.3f ms wall, %.3f ms device" % ((time.perf_counter() - t0) * 1e3, ms), flush=True)
t0 = time.perf_counter(); s = M.HostScene(scenes, name); print("host loader %.3f ms" % ((time.perf_counter() - t0) * 1e3))
