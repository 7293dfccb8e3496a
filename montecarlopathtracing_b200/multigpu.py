"""SPP-split rendering over R ranks (one process per GPU, torch.distributed for the plumbing).

The spp samples are always cut into NUM_CHUNKS = 64 chunks in sample-index order.  Rank r of R in {1,2,4,8} integrates
chunks [r*64/R, (r+1)*64/R) and leaves their fixed pairwise-tree SUM in one float buffer (mcpt_render_chunks); the R
buffers are gathered to rank 0 — the path's only exchange — and summed there by the same pairwise tree
(mcpt_sum_partials), then scaled by 1/spp.  The association of the float additions is therefore identical for every
R, and since Philox is keyed by (pixel, sample) the 1-, 2-, 4- and 8-GPU images are bit-identical.  ncclReduce is NOT
used: its summation order is unspecified.

`render_partial` and `sum_partials` are injected so the split/gather/sum logic can be exercised on CPU (gloo) with a
stand-in backend in tests; on a GPU box they are bound to the CUDA context by `gpu_backend`.
"""
import torch
import torch.distributed as dist

from ._capi import NUM_CHUNKS


def chunk_range(rank, world):
    if world not in (1, 2, 4, 8):
        raise ValueError("world size must be 1, 2, 4 or 8 (divides the %d sample chunks)" % NUM_CHUNKS)
    per = NUM_CHUNKS // world
    return rank * per, (rank + 1) * per


def render_split(render_partial, sum_partials, spp, rank=0, world=1, group=None):
    """Returns the final mean image tensor on rank 0 (None elsewhere).

    render_partial(chunk_begin, chunk_end) -> 1-D float32 tensor: tree SUM of this rank's chunks.
    sum_partials(stacked[(R*count)], R, scale) -> 1-D float32 tensor.
    """
    cb, ce = chunk_range(rank, world)
    part = render_partial(cb, ce)
    if world == 1:
        return sum_partials(part, 1, 1.0 / spp)
    gathered = [torch.empty_like(part) for _ in range(world)] if rank == 0 else None
    dist.gather(part, gather_list=gathered, dst=0, group=group)  # rank order = chunk order
    if rank != 0:
        return None
    return sum_partials(torch.cat(gathered), world, 1.0 / spp)


def gpu_backend(ctx, camera, spp, width, height, mode=1, flags=0):
    """Binds render_partial / sum_partials to a CUDA Context; tensors are torch CUDA tensors (device memory only)."""
    w = width if width > 0 else camera.width
    h = height if height > 0 else camera.height
    count = w * h * 3
    dev = torch.device("cuda", torch.cuda.current_device())
    stats = []

    def render_partial(cb, ce):
        out = torch.empty(count, dtype=torch.float32, device=dev)
        stats.append(ctx.render_chunks(camera, spp, cb, ce, out.data_ptr(), width=w, height=h, mode=mode, flags=flags))
        return out

    def sum_partials(stacked, n, scale):
        out = torch.empty(count, dtype=torch.float32, device=dev)
        ctx.sum_partials(stacked.data_ptr(), n, count, scale, out.data_ptr())
        return out

    return render_partial, sum_partials, stats
