import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import montecarlopathtracing_b200 as M
for name in ("cornell-box", "veach-mis"):
    s = M.HostScene(os.path.join(ROOT, "scenes") + "/", name)
    ctx = M.Context(0); ctx.upload_scene(s); ctx.build_bvh(); ctx.set_seed(77)
    for spp in (1, 8):
        a = ctx.render(s.camera, spp, 96, 72, mode=M.MODE_WAVEFRONT, flags=32)
        b = ctx.render(s.camera, spp, 96, 72, mode=M.MODE_MEGAKERNEL, flags=32)
        d = np.abs(a - b).max(axis=2)
        idx = np.argwhere(d > 0)
        print(name, "spp", spp, "differing pixels", len(idx), "max abs diff", d.max(), "rays wf/mega", ctx.last_stats.rays)
        for y, x in idx[:5]:
            print("   pixel", y, x, "wf", a[y, x], "mega", b[y, x])
    ctx.close()
