"""SPP-split rendering over R ranks, one process per GPU (torch.distributed is the launcher's plumbing only).

The spp samples are always cut into NUM_CHUNKS = 64 chunks in sample-index order.  Rank r of R in {1,2,4,8} integrates
chunks [r*64/R, (r+1)*64/R) and leaves their fixed pairwise-tree SUM in one float buffer; the R buffers go to rank 0 — the
path's only exchange — and are summed there by the same pairwise tree, then scaled by 1/spp.  The association of the float
additions is therefore identical for every R, and since Philox is keyed by (pixel, sample) the 1-, 2-, 4- and 8-GPU images
are bit-identical.  ncclReduce is NOT used: its summation order is unspecified.

On a GPU box the exchange and the sum live in libmcpt.so (mcpt_render_dist, csrc/multigpu.cu: grouped ncclSend/ncclRecv on
the library's own communicator straight into one preallocated R x count buffer, then K12): `init_comm` only carries the
128-byte NCCL id from rank 0 to the other ranks.  `render_split` is the same split/gather/sum logic with the two compute
steps injected, so it can be exercised on CPU (gloo) with a stand-in backend in tests/test_multigpu_cpu.py.
"""
import torch
import torch.distributed as dist

from ._capi import NUM_CHUNKS


def chunk_range(rank, world):
    if world not in (1, 2, 4, 8):
        raise ValueError("world size must be 1, 2, 4 or 8 (divides the %d sample chunks)" % NUM_CHUNKS)
    per = NUM_CHUNKS // world
    return rank * per, (rank + 1) * per


def init_comm(ctx, rank, world, group=None):
    """Joins `ctx` (api.Context) to the render's communicator: rank 0 makes the NCCL id inside libmcpt.so, the launcher's
    process group broadcasts its 128 bytes, every rank calls mcpt_comm_init."""
    from .api import comm_unique_id
    chunk_range(rank, world)
    if world == 1:
        ctx.comm_init(1, 0, bytes(128))
        return
    box = [comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0, group=group)
    ctx.comm_init(world, rank, box[0])


def render_split(render_partial, sum_partials, spp, rank=0, world=1, group=None):
    """Returns the final mean image tensor on rank 0 (None elsewhere).

    render_partial(chunk_begin, chunk_end) -> 1-D float32 tensor: tree SUM of this rank's chunks.
    sum_partials(stacked[(R*count)], R, scale) -> 1-D float32 tensor.
    """
    cb, ce = chunk_range(rank, world)
    part = render_partial(cb, ce)
    if world == 1:
        return sum_partials(part, 1, 1.0 / spp)
    # one preallocated R x count buffer; the gather receives straight into its rows (rank order = chunk order)
    stacked = torch.empty((world, part.numel()), dtype=part.dtype, device=part.device) if rank == 0 else None
    dist.gather(part, gather_list=[stacked[r] for r in range(world)] if rank == 0 else None, dst=0, group=group)
    if rank != 0:
        return None
    return sum_partials(stacked.view(-1), world, 1.0 / spp)


def gpu_backend(ctx, camera, spp, width, height, mode=1, flags=0):
    """Binds render_partial / sum_partials to a CUDA Context; tensors are torch CUDA tensors (device memory only).

    The context is switched to torch's current stream (mcpt_set_stream), so the library's kernels and whatever torch
    enqueues around them (torch.cat, dist.gather, copies) are ordered on one stream; call ctx.use_own_stream() to undo."""
    w = width if width > 0 else camera.width
    h = height if height > 0 else camera.height
    count = w * h * 3
    dev = torch.device("cuda", torch.cuda.current_device())
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
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
