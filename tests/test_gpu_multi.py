"""GPU, >= 2 devices (skipped on a one-GPU box; run with `gpurun --gpus 2|8`): the SPP split behind the C ABI over real NCCL —
one process driving all GPUs (mcpt_multi_*, ncclCommInitAll) and one process per GPU (mcpt_comm_init / mcpt_render_dist,
ncclCommInitRank) — must give the one-GPU image bit for bit, and mcpt_cli --gpus N the same PNG bytes as --gpus 1."""
import os
import subprocess

import numpy as np
import pytest

import montecarlopathtracing_b200 as M
from conftest import ROOT, SCENES

pytestmark = pytest.mark.gpu


def _n_gpus():
    import torch
    return torch.cuda.device_count()


def _worlds():
    return [r for r in (2, 4, 8) if r <= _n_gpus()]


@pytest.fixture(scope="module")
def single(scenes):
    ctx = M.Context(0)
    out = {}
    for name, (w, h, spp) in {"cornell-box": (640, 360, 40), "veach-mis": (300, 225, 17)}.items():
        s = scenes[name]
        ctx.upload_scene(s); ctx.build_bvh(); ctx.set_seed(2026)
        out[name] = (w, h, spp, ctx.render(s.camera, spp, w, h), ctx.last_stats.rays)
    ctx.close()
    return out


def test_one_process_all_gpus_is_bit_identical(scenes, single, tmp_path):
    if _n_gpus() < 2:
        pytest.skip("needs >= 2 GPUs")
    for R in _worlds():
        m = M.MultiContext(R)
        m.set_seed(2026)
        for name in ("cornell-box", "veach-mis"):
            w, h, spp, base, rays = single[name]
            s = scenes[name]
            m.upload_scene(s); m.build_bvh()
            for mode in (M.MODE_WAVEFRONT, M.MODE_MEGAKERNEL):
                img = m.render(s.camera, spp, w, h, mode=mode)
                assert np.array_equal(img.view(np.uint32), base.view(np.uint32)), "%s R=%d mode %d differs" % (name, R, mode)
                # the ray total is not an invariant: every path slot / megakernel thread traces its own primary ray once (the cached
                # primary hit lives in it), and their number depends on the GPU count and on the integrator
                assert 0.5 * rays < m.last_stats.rays < 2.0 * rays and m.last_stats.samples == w * h * spp and m.last_stats.exchange_ms > 0
        # GPU-side ingest on every GPU + device PNG stage
        cam = m.load_scene_gpu(SCENES, "cornell-box"); m.build_bvh()
        w, h, spp, base, _ = single["cornell-box"]
        f, g = str(tmp_path / "multi.png"), str(tmp_path / "one.png")
        m.render_png(cam, spp, f, w, h)
        M.write_png(g, base)
        assert open(f, "rb").read() == open(g, "rb").read()
        m.close()


def _rank_main(rank, world, port, name, w, h, spp, out_file):
    import torch
    import torch.distributed as dist
    from montecarlopathtracing_b200.multigpu import init_comm
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)     # the launcher's plumbing: carries the 128-byte id
    s = M.HostScene(SCENES, name)
    ctx = M.Context(rank)
    ctx.upload_scene(s); ctx.build_bvh(); ctx.set_seed(2026)
    init_comm(ctx, rank, world)
    for _ in range(2):                                               # twice: the exchange buffers are reused
        img = ctx.render_dist(s.camera, spp, w, h)
    if rank == 0:
        np.save(out_file, img)
    else:
        assert img is None
    ctx.close()
    dist.destroy_process_group()


def test_one_process_per_gpu_is_bit_identical(single, tmp_path):
    if _n_gpus() < 2:
        pytest.skip("needs >= 2 GPUs")
    import socket
    import torch.multiprocessing as mp
    for R in _worlds():
        for name in ("cornell-box", "veach-mis"):
            w, h, spp, base, _ = single[name]
            with socket.socket() as sk:
                sk.bind(("127.0.0.1", 0)); port = sk.getsockname()[1]
            out = str(tmp_path / ("dist_%s_%d.npy" % (name, R)))
            mp.spawn(_rank_main, args=(R, port, name, w, h, spp, out), nprocs=R, join=True)
            assert np.array_equal(np.load(out).view(np.uint32), base.view(np.uint32)), "%s R=%d differs" % (name, R)


def test_cli_gpus_writes_the_same_png(tmp_path):
    if _n_gpus() < 2:
        pytest.skip("needs >= 2 GPUs")
    cli = os.path.join(ROOT, "montecarlopathtracing_b200", "mcpt_cli")
    files = []
    for R in [1] + _worlds():
        out = tmp_path / ("g%d" % R)
        out.mkdir()
        subprocess.check_call([cli, "--scene-dir", SCENES, "--name", "cornell-box", "--spp", "32", "--width", "480", "--height", "270",
                               "--out", str(out) + "/", "--gpus", str(R), "--seed", "7"], stdout=subprocess.DEVNULL)
        files.append(open(out / "cornell-box-SPP32.png", "rb").read())
    assert all(f == files[0] for f in files) and len(files[0]) == M.png_size(480, 270)
