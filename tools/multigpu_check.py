"""Run under torchrun on N GPUs: the SPP-split render gathered to rank 0 must be BIT-identical to rank 0's own
single-GPU mcpt_render of all 64 chunks.  Prints one line; exit code 1 on mismatch."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import montecarlopathtracing_b200 as M  # noqa: E402
from montecarlopathtracing_b200.multigpu import gpu_backend, render_split  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = M.Context(local)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
ok = True
for name, w, h, spp, mode in (("cornell-box", 640, 360, 40, M.MODE_WAVEFRONT), ("veach-mis", 400, 300, 19, M.MODE_WAVEFRONT),
                              ("cornell-box", 320, 200, 16, M.MODE_MEGAKERNEL)):
    s = M.HostScene(os.path.join(ROOT, "scenes") + "/", name)
    ctx.upload_scene(s); ctx.build_bvh(); ctx.set_seed(7)
    rp, sp, _ = gpu_backend(ctx, s.camera, spp, w, h, mode=mode)
    img = render_split(rp, sp, spp, rank, world)
    torch.cuda.synchronize()
    if rank == 0:
        single = ctx.render(s.camera, spp, w, h, mode=mode)
        same = bool(np.array_equal(img.cpu().numpy().reshape(h, w, 3), single))
        ok = ok and same
        print("multigpu_check world=%d %s %dx%d spp %d mode %d: bit-identical to the 1-GPU image: %s" % (world, name, w, h, spp, mode, same), flush=True)
dist.barrier()
ctx.close()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
