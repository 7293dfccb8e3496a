"""GPU: seam 2 (closest hit) through the C ABI against the reference's exported ray batches and the oracle."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from scene_util import random_soup

pytestmark = pytest.mark.gpu
SCENE_NAMES = ("cornell-box", "veach-mis")

# Tolerances (BASELINE.json north_star): hit ids exact, t and barycentrics within 1e-5 relative.
T_RTOL = 1e-5
BARY_ATOL = 1e-5


def compare_hits(hits, ref_hit, ref_face, ref_t, rays, scene_pos, ref_p):
    """Returns (n_id_mismatch, worst relative t error on id-matching hits, worst barycentric error)."""
    got_hit = hits["tri"] >= 0
    same = (got_hit == (ref_hit == 1)) & (hits["tri"] == ref_face)
    m = same & got_hit
    t_err = np.abs(hits["t"][m].astype(np.float64) - ref_t[m]) / np.abs(ref_t[m])
    # reference barycentrics from its hit point p (findGarCor, MTPC/pathTracing.cpp:416-428) in float64
    tri = scene_pos[ref_face[m]]
    v1, v2, v3 = tri[:, 0], tri[:, 1], tri[:, 2]
    nrm = np.cross(v3 - v2, v1 - v3)
    an = (nrm * nrm).sum(axis=1)
    p = ref_p[m]
    b2 = (np.cross(v1 - v3, p - v1) * nrm).sum(axis=1) / an
    b3 = (np.cross(v2 - v1, p - v2) * nrm).sum(axis=1) / an
    bary_err = np.maximum(np.abs(hits["u"][m] - b2), np.abs(hits["v"][m] - b3))
    return int((~same).sum()), t_err, bary_err, same


@pytest.mark.parametrize("name", SCENE_NAMES)
def test_golden_ray_batches(gpu_ctx, scenes, name):
    """Camera, bounce, light-directed and arbitrary rays exported from the UNMODIFIED reference's ray_intersect."""
    s = scenes[name]
    g = np.load(os.path.join(GOLDEN, name + "_rays.npz"))
    gpu_ctx.upload_scene(s)
    gpu_ctx.build_bvh()
    hits = gpu_ctx.trace(g["rays"].astype(np.float32))
    n_bad, t_err, bary_err, same = compare_hits(hits, g["hit"], g["face"], g["t"], g["rays"], s.pos, g["p"])
    n = len(g["hit"])
    # K5 filters candidates in FP32 and decides them with the reference's own FP64 plane / same-side test and tie rule
    # (render.cu reference_tri_test_fp64): the ids are the reference's on EVERY ray, including the camera rays that run exactly
    # along the cornell-box wall edges on the image diagonals
    assert n_bad == 0, "hit ids must match the reference (%d of %d differ, first at ray %d)" % (n_bad, n, int(np.nonzero(~same)[0][0]))
    # every id-matching hit, not a quantile: the record is recomputed in FP64 from the FP64 vertices (render.cu)
    assert t_err.max() <= T_RTOL and bary_err.max() <= BARY_ATOL
    st = gpu_ctx.last_stats
    assert st.rays == n and st.node_visits > 0 and st.tri_tests > 0


def test_random_soup_vs_oracle(gpu_ctx, oracle):
    """20k random rays through a 20k-triangle soup, ids vs the CPU restatement (even count: reference quirk Q1)."""
    n = 20000
    desc, keep = random_soup(n, seed=11, extent=2.0, size=0.15)
    gpu_ctx.upload_scene(desc)
    gpu_ctx.build_bvh()
    rng = np.random.default_rng(1)
    o = rng.uniform(-1, 3.2, size=(20000, 3)); d = rng.normal(size=(20000, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.concatenate([o, d], axis=1).astype(np.float32)
    hits = gpu_ctx.trace(rays)
    ref = oracle.OracleScene(desc).trace(rays.astype(np.float64))
    n_bad, t_err, bary_err, same = compare_hits(hits, ref["hit"], ref["face"], ref["t"], rays, keep[0].reshape(n, 3, 3), ref["p"])
    assert n_bad == 0
    assert t_err.max() <= T_RTOL and bary_err.max() <= BARY_ATOL
    assert np.array_equal(hits["leaf"][same & (hits["tri"] >= 0)], ref["leaf"][same & (hits["tri"] >= 0)])


def test_axis_parallel_and_degenerate_rays(gpu_ctx, oracle, scenes):
    """d.x == 0 hits nothing (quirk Q2: t = (p.x-o.x)/d.x is NaN, MTPC/pathTracing.cpp:347); d.y or d.z == 0 is fine;
    rays starting on a surface need t > 0; a ray pointing away misses."""
    s = scenes["cornell-box"]
    gpu_ctx.upload_scene(s)
    gpu_ctx.build_bvh()
    rays = np.array([[0.1, 1.0, 6.8, 0, 0, -1],        # d.x == 0 -> miss by Q2
                     [0.1, 1.0, 6.8, 1e-3, 0, -1],     # d.y == 0 ok
                     [0.1, 1.0, 0.2, 1e-3, -1, 0],     # d.z == 0 ok, straight down
                     [0.1, 1.0, 6.8, 0.01, 0.02, 1],   # away from the box
                     [0.3, 0.7, 0.1, 0.5, 0.5, 0.5]], np.float64)
    hits = gpu_ctx.trace(rays.astype(np.float32))
    ref = oracle.OracleScene(s.desc).trace(rays)
    assert list(ref["hit"]) == [0, 1, 1, 0, 1]
    assert np.array_equal(hits["tri"] >= 0, ref["hit"] == 1)
    assert np.array_equal(hits["tri"], ref["face"])


def test_single_triangle_scene(gpu_ctx):
    desc, keep = random_soup(1, seed=3)
    gpu_ctx.upload_scene(desc)
    gpu_ctx.build_bvh()
    tri = keep[0].reshape(3, 3)
    c = tri.mean(axis=0)
    nrm = np.cross(tri[1] - tri[0], tri[2] - tri[0]); nrm /= np.linalg.norm(nrm)
    target = 0.2 * tri[0] + 0.3 * tri[1] + 0.5 * tri[2]          # barycentric weights of v2, v3 = (0.3, 0.5)
    o = c + 2 * nrm + np.array([0.05, 0.02, 0.01])
    d = target - o
    rays = np.array([np.concatenate([o, d]), np.concatenate([o, -d])], np.float32)
    hits = gpu_ctx.trace(rays)
    assert hits["tri"][0] == 0 and hits["tri"][1] == -1
    assert abs(hits["u"][0] - 0.3) < 1e-5 and abs(hits["v"][0] - 0.5) < 1e-5 and abs(hits["t"][0] - 1.0) < 1e-5


def test_empty_batch_and_call_order(gpu_ctx, built_lib):
    import montecarlopathtracing_b200 as M
    desc, _ = random_soup(8, seed=1)
    gpu_ctx.upload_scene(desc)
    gpu_ctx.build_bvh()
    assert len(gpu_ctx.trace(np.zeros((0, 6), np.float32))) == 0
    fresh = M.Context(0)
    with pytest.raises(M.McptError, match="upload a scene first"):
        fresh.build_bvh()
    fresh.upload_scene(desc)
    with pytest.raises(M.McptError, match="build the BVH first"):
        fresh.trace(np.zeros((1, 6), np.float32))
    fresh.close()
