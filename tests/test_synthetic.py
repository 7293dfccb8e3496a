"""The synthetic multi-million-triangle scene of BASELINE.json configs[3] (montecarlopathtracing_b200/synthetic.py):
generator contract on CPU, and GPU parity against the oracle on a small instance + size-independent properties at 5M."""
import numpy as np
import pytest

from montecarlopathtracing_b200.synthetic import make_scene


def _pos(keep, n):
    return keep[0].reshape(n, 3, 3)


@pytest.mark.parametrize("n", [14, 2000, 20000, 123456])
def test_generator_contract(oracle, n):
    desc, keep, cam = make_scene(n, width=64, height=36)
    assert desc.n_faces == n and n % 2 == 0 and desc.n_lights == 1 and desc.n_materials == 3
    pos = _pos(keep, n)
    assert np.array_equal(pos, pos.astype(np.float32).astype(np.float64))          # float-representable coordinates
    assert pos.min() >= -0.9001 and pos.max() <= 3.9001
    # deterministic
    desc2, keep2, _ = make_scene(n, width=64, height=36)
    assert np.array_equal(keep[0], keep2[0]) and np.array_equal(keep[1], keep2[1])
    # inside the reference's Morton cube [-1,4]^3: no centroid clamps to the first or last cell on any axis
    codes = oracle.morton(keep[0])
    def compact(c):          # inverse of the 10->30-bit spread (MTPC/morton code.cpp:3-10)
        c = c & 0x49249249
        c = (c | (c >> 2)) & 0xC30C30C3; c = (c | (c >> 4)) & 0x0F00F00F; c = (c | (c >> 8)) & 0xFF0000FF; c = (c | (c >> 16)) & 0x3FF
        return c
    for shift in (2, 1, 0):
        q = compact(codes.astype(np.int64) >> shift)
        assert q.min() >= 1 and q.max() <= 1022
    nrm = keep[1].reshape(n, 3, 3)
    assert np.allclose(np.linalg.norm(nrm, axis=2), 1.0, atol=1e-12)


@pytest.mark.gpu
def test_small_instance_matches_oracle(gpu_ctx, oracle):
    import montecarlopathtracing_b200 as M
    n = 20000
    desc, keep, cam = make_scene(n, width=96, height=54)
    gpu_ctx.upload_scene(desc); gpu_ctx.build_bvh(); gpu_ctx.set_seed(5)
    osc = oracle.OracleScene(desc)
    codes_obj, _, face_sorted = gpu_ctx.get_morton()
    ocodes, operm = osc.order()
    assert np.array_equal(codes_obj, ocodes) and np.array_equal(face_sorted, operm)
    for mode in (M.MODE_WAVEFRONT, M.MODE_MEGAKERNEL):
        img = gpu_ctx.render(cam, 8, mode=mode)
        ref, _ = osc.render(cam, 8, seed=5)
        close = (np.abs(img - ref) <= 2e-3 * np.abs(ref) + 1e-4).all(axis=2)
        assert close.mean() >= 0.99, "mode %d: only %.4f of pixels match" % (mode, close.mean())
        assert np.all(np.abs(img.mean(axis=(0, 1)) - ref.mean(axis=(0, 1))) <= 5e-3 * ref.mean(axis=(0, 1)))


@pytest.mark.gpu
def test_five_million_triangles_properties(gpu_ctx, oracle):
    """Full BASELINE size: properties that do not need the CPU to trace 5M triangles — sorted keys, valid permutation,
    every Morton code bit-exact vs the CPU restatement, camera rays hit the floor mesh where geometry says they must,
    and the rendered image is finite, deterministic and lit."""
    import montecarlopathtracing_b200 as M
    n = 5_000_000
    desc, keep, cam = make_scene(n, width=480, height=270)
    gpu_ctx.upload_scene(desc)
    ms = gpu_ctx.build_bvh()
    c, s, f = gpu_ctx.get_morton()
    assert np.all(np.diff(s.astype(np.int64)) >= 0) and np.array_equal(np.sort(f), np.arange(n)) and np.array_equal(c[f], s)
    assert np.array_equal(c, oracle.morton(keep[0]))
    # straight-down rays over the interior of the height field hit a floor triangle below them at the analytic height
    rng = np.random.default_rng(3)
    xz = rng.uniform(-0.8, 3.7, size=(20000, 2))
    rays = np.concatenate([np.stack([xz[:, 0], np.full(len(xz), 3.0), xz[:, 1]], axis=1),
                           np.tile(np.array([[1e-4, -1.0, 2e-4]]), (len(xz), 1))], axis=1).astype(np.float32)
    hits = gpu_ctx.trace(rays)
    assert np.all(hits["tri"] >= 0) and np.all(hits["tri"] < n - 12)
    pos = _pos(keep, n)[hits["tri"]]
    w = np.stack([1 - hits["u"] - hits["v"], hits["u"], hits["v"]], axis=1).astype(np.float64)
    p = (pos * w[:, :, None]).sum(axis=1)
    want = rays[:, :3].astype(np.float64) + rays[:, 3:].astype(np.float64) * hits["t"][:, None].astype(np.float64)
    assert np.abs(p - want).max() < 1e-4
    gpu_ctx.set_seed(9)
    a = gpu_ctx.render(cam, 8, mode=M.MODE_WAVEFRONT)
    b = gpu_ctx.render(cam, 8, mode=M.MODE_MEGAKERNEL)
    assert np.all(np.isfinite(a)) and np.array_equal(a, b) and 0.05 < a.mean() < 5.0
    assert ms > 0
