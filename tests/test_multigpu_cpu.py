"""CPU (gloo, world_size 2 and 4): the SPP-split / gather / fixed-tree-sum host logic of multigpu.render_split with a
stand-in backend.  The stand-in "renders" a chunk as a deterministic pseudo-random float plane and sums planes with the
same pairwise tree as mcpt_sum_partials (K12), so the check is exactly the one the GPU path must satisfy: the image is
BIT-identical for every world size."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from montecarlopathtracing_b200.multigpu import chunk_range, render_split
from montecarlopathtracing_b200._capi import NUM_CHUNKS

COUNT = 4099  # floats per plane (odd on purpose)
SPP = 24


def chunk_plane(c):
    """Stand-in for the per-chunk sample sum: values of very different magnitude so the association of + matters."""
    rng = np.random.default_rng(1000 + c)
    return (rng.standard_normal(COUNT) * 10.0 ** rng.integers(-3, 4, COUNT)).astype(np.float32)


def tree_sum(planes):
    """((p0+p1)+(p2+p3))+... in float32: the association of sum_planes_kernel (render.cu K12)."""
    v = [np.asarray(p, np.float32) for p in planes]
    while len(v) > 1:
        v = [(v[i] + v[i + 1]).astype(np.float32) for i in range(0, len(v), 2)]
    return v[0]


def render_partial(cb, ce):
    return torch.from_numpy(tree_sum([chunk_plane(c) for c in range(cb, ce)]).copy())


def sum_partials(stacked, n, scale):
    parts = stacked.numpy().reshape(n, COUNT)
    return torch.from_numpy((tree_sum(list(parts)) * np.float32(scale)).astype(np.float32))


def _worker(rank, world, port, out_path):
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    img = render_split(render_partial, sum_partials, SPP, rank, world)
    assert (img is None) == (rank != 0)
    if rank == 0:
        np.save(out_path, img.numpy())
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_chunk_ranges_partition_the_chunks():
    for world in (1, 2, 4, 8):
        ranges = [chunk_range(r, world) for r in range(world)]
        assert ranges[0][0] == 0 and ranges[-1][1] == NUM_CHUNKS
        assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
        assert all((e - b) == NUM_CHUNKS // world and b % (e - b) == 0 for b, e in ranges)   # aligned power-of-two ranges
    with pytest.raises(ValueError):
        chunk_range(0, 3)


@pytest.mark.parametrize("world", [2, 4])
def test_split_gather_sum_is_world_size_independent(tmp_path, world):
    single = render_split(render_partial, sum_partials, SPP, 0, 1).numpy()
    want = (tree_sum([chunk_plane(c) for c in range(NUM_CHUNKS)]) * np.float32(1.0 / SPP)).astype(np.float32)
    assert np.array_equal(single, want)
    out = str(tmp_path / "img.npy")
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    got = np.load(out)
    assert np.array_equal(got, single), "world=%d image differs from the single-rank image" % world
    # a different association would NOT be identical: the test has teeth
    naive = np.zeros(COUNT, np.float32)
    for c in range(NUM_CHUNKS):
        naive = (naive + chunk_plane(c)).astype(np.float32)
    assert not np.array_equal((naive * np.float32(1.0 / SPP)).astype(np.float32), single)
