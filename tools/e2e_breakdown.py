"""Wall-clock breakdown of bench.py's e2e step (upload + build + render_split + D2H) on cornell-box 1024x1024x256."""
import ctypes, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import montecarlopathtracing_b200 as M
from montecarlopathtracing_b200 import _capi
from montecarlopathtracing_b200.multigpu import gpu_backend, render_split
ctx = M.Context(0)
s = M.HostScene(os.path.join(ROOT, "scenes") + "/", "cornell-box")
pinned = [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in (s.pos, s.nrm, s.uv, s.mat_id)]
desc = _capi.SceneDesc()
ctypes.memmove(ctypes.byref(desc), ctypes.byref(s.desc), ctypes.sizeof(desc))
desc.pos = ctypes.cast(pinned[0].data_ptr(), ctypes.POINTER(ctypes.c_double))
desc.nrm = ctypes.cast(pinned[1].data_ptr(), ctypes.POINTER(ctypes.c_double))
desc.uv = ctypes.cast(pinned[2].data_ptr(), ctypes.POINTER(ctypes.c_double))
desc.mat_id = ctypes.cast(pinned[3].data_ptr(), ctypes.POINTER(ctypes.c_int32))
host_img = torch.empty(1024 * 1024 * 3, dtype=torch.float32).pin_memory()
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
ctx.upload_scene(desc)
rp, sp, stats = gpu_backend(ctx, s.camera, 256, 1024, 1024)
def t(fn):
    torch.cuda.synchronize(); t0 = time.perf_counter(); r = fn(); torch.cuda.synchronize(); return (time.perf_counter() - t0) * 1e3, r
for i in range(4):
    a, _ = t(lambda: ctx.upload_scene(desc)); b, _ = t(lambda: ctx.build_bvh())
    c, part = t(lambda: rp(0, 64)); d, img = t(lambda: sp(part, 1, 1.0 / 256)); e, _ = t(lambda: host_img.copy_(img))
    print("upload %.2f  build %.2f  render_chunks %.2f (render_ms %.2f integrate_ms %.2f)  sum %.2f  d2h %.2f" % (a, b, c, stats[-1].render_ms, stats[-1].integrate_ms, d, e), flush=True)
