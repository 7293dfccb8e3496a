"""GPU: seam 1 (Morton -> sort -> LBVH) through the C ABI, against the oracle and the reference's golden vectors."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from scene_util import random_soup

pytestmark = pytest.mark.gpu
SCENE_NAMES = ("cornell-box", "veach-mis")


@pytest.mark.parametrize("name", SCENE_NAMES)
def test_morton_codes_and_leaf_order_bit_exact(gpu_ctx, scenes, name):
    g = np.load(os.path.join(GOLDEN, name + "_morton.npz"))
    gpu_ctx.upload_scene(scenes[name])
    ms = gpu_ctx.build_bvh()
    codes_obj, codes_sorted, face_sorted = gpu_ctx.get_morton()
    assert np.array_equal(codes_obj, g["codes"])                       # every code, bit for bit, vs the reference
    assert np.array_equal(face_sorted, g["perm"])                      # canonical leaf order (ties by face index)
    assert np.array_equal(codes_sorted, g["codes"][g["perm"]])
    assert ms > 0


def check_tree(nodes, n, codes_sorted, tri_lo, tri_hi):
    """Structural invariants of the radix tree + refit: every leaf reachable exactly once, leaf ranges contiguous and
    nested, child boxes contain their subtree's triangles, parents recorded."""
    child = nodes[:, 12:14].view(np.int32)
    parent = nodes[:, 14].view(np.int32)
    assert parent[0] == -1
    seen = np.zeros(n, np.int32)
    # iterative post-order: returns (lo_leaf, hi_leaf, box_lo, box_hi) per subtree
    def box_of(i, slot):
        q = nodes[i]
        return np.array([q[4 * slot + 0], q[4 * slot + 2], q[8 + 2 * slot]]), np.array([q[4 * slot + 1], q[4 * slot + 3], q[9 + 2 * slot]])
    result = {}
    stack = [(0, False)]
    while stack:
        i, done = stack.pop()
        if not done:
            stack.append((i, True))
            for c in child[i]:
                if c >= 0:
                    assert parent[c] == i
                    stack.append((int(c), False))
            continue
        rng, lo_all, hi_all = [], np.full(3, np.inf), np.full(3, -np.inf)
        for slot, c in enumerate(child[i]):
            blo, bhi = box_of(i, slot)
            if c < 0:
                k = ~int(c); seen[k] += 1
                lo, hi, klo, khi = tri_lo[k], tri_hi[k], k, k
            else:
                klo, khi, lo, hi = result.pop(int(c))
            assert np.all(blo <= lo) and np.all(bhi >= hi), "child box does not contain its subtree"
            assert np.all(blo >= lo - 1e-5 * (1 + np.abs(lo))) and np.all(bhi <= hi + 1e-5 * (1 + np.abs(hi))), "box not tight"
            rng.append((klo, khi)); lo_all = np.minimum(lo_all, lo); hi_all = np.maximum(hi_all, hi)
        assert rng[0][1] + 1 == rng[1][0], "children must cover adjacent leaf ranges"
        result[i] = (rng[0][0], rng[1][1], lo_all, hi_all)
    assert np.all(seen == 1)
    lo, hi, _, _ = result[0]
    assert (lo, hi) == (0, n - 1)


@pytest.mark.parametrize("name", SCENE_NAMES)
def test_tree_invariants_on_shipped_scenes(gpu_ctx, scenes, name):
    s = scenes[name]
    gpu_ctx.upload_scene(s)
    gpu_ctx.build_bvh()
    _, codes_sorted, face_sorted = gpu_ctx.get_morton()
    nodes, root = gpu_ctx.get_nodes()
    assert root == 0 and nodes.shape == (s.n_faces - 1, 16)
    p = s.pos[face_sorted].astype(np.float32).astype(np.float64)
    check_tree(nodes, s.n_faces, codes_sorted, p.min(axis=1), p.max(axis=1))


@pytest.mark.parametrize("n", [1, 2, 3, 7, 64, 1001])
def test_small_and_odd_triangle_counts(gpu_ctx, oracle, n):
    """Edge cases the reference cannot even run (odd counts over-read, MTPC/BVH.cpp Q1): 1 triangle, odd counts,
    many equal codes (all centroids clamp to the same cell when far outside [-1,4]^3)."""
    desc, keep = random_soup(n, seed=n, far_outside=(n == 64))
    gpu_ctx.upload_scene(desc)
    gpu_ctx.build_bvh()
    codes_obj, codes_sorted, face_sorted = gpu_ctx.get_morton()
    want = oracle.morton(keep[0])
    assert np.array_equal(codes_obj, want)
    assert np.array_equal(face_sorted, oracle.stable_order(want))
    if n == 64:
        assert len(np.unique(codes_obj)) == 1
    if n > 1:
        nodes, _ = gpu_ctx.get_nodes()
        p = keep[0].reshape(n, 3, 3)[face_sorted].astype(np.float32).astype(np.float64)
        check_tree(nodes, n, codes_sorted, p.min(axis=1), p.max(axis=1))


def test_large_mesh_sorted_and_idempotent(gpu_ctx, oracle):
    """Size-independent properties at 1M triangles: keys sorted, permutation valid, rebuild gives the same order."""
    n = 1 << 20
    desc, keep = random_soup(n, seed=5, extent=2.4)
    gpu_ctx.upload_scene(desc)
    gpu_ctx.build_bvh()
    c1, s1, f1 = gpu_ctx.get_morton()
    assert np.all(np.diff(s1.astype(np.int64)) >= 0)
    assert np.array_equal(np.sort(f1), np.arange(n))
    assert np.array_equal(c1[f1], s1)
    ties = np.diff(s1.astype(np.int64)) == 0
    assert np.all(np.diff(f1)[ties] > 0)                   # equal codes keep face-index order
    assert np.array_equal(c1, oracle.morton(keep[0]))      # all 1M codes bit-exact vs the CPU restatement
    gpu_ctx.build_bvh()
    c2, s2, f2 = gpu_ctx.get_morton()
    assert np.array_equal(f1, f2) and np.array_equal(s1, s2)


def check_traversal_tree(nodes, root, n, tri_lo, tri_hi):
    """The tree traversal walks (PLOC over the Morton-ordered leaves): every leaf reachable exactly once from the root,
    n-1 internal nodes all used, every child box contains its subtree's triangles and is tight."""
    child = nodes[:, 12:14].view(np.int32)
    seen_leaf, seen_node = np.zeros(n, np.int32), np.zeros(len(nodes), np.int32)
    def box_of(i, slot):
        q = nodes[i]
        return np.array([q[4 * slot + 0], q[4 * slot + 2], q[8 + 2 * slot]]), np.array([q[4 * slot + 1], q[4 * slot + 3], q[9 + 2 * slot]])
    result, depth_max = {}, 0
    stack = [(root, False, 1)]
    while stack:
        i, done, depth = stack.pop()
        depth_max = max(depth_max, depth)
        if not done:
            seen_node[i] += 1
            stack.append((i, True, depth))
            for c in child[i]:
                if c >= 0:
                    stack.append((int(c), False, depth + 1))
            continue
        lo_all, hi_all = np.full(3, np.inf), np.full(3, -np.inf)
        for slot, c in enumerate(child[i]):
            blo, bhi = box_of(i, slot)
            if c < 0:
                k = ~int(c); seen_leaf[k] += 1
                lo, hi = tri_lo[k], tri_hi[k]
            else:
                lo, hi = result.pop(int(c))
            assert np.all(blo <= lo) and np.all(bhi >= hi), "child box does not contain its subtree"
            assert np.all(blo >= lo - 1e-5 * (1 + np.abs(lo))) and np.all(bhi <= hi + 1e-5 * (1 + np.abs(hi))), "box not tight"
            lo_all = np.minimum(lo_all, lo); hi_all = np.maximum(hi_all, hi)
        result[i] = (lo_all, hi_all)
    assert np.all(seen_leaf == 1) and np.all(seen_node == 1)
    assert depth_max < 64, "traversal stack is 64 entries deep"
    return depth_max


@pytest.mark.parametrize("name", SCENE_NAMES)
def test_traversal_tree_invariants_on_shipped_scenes(gpu_ctx, scenes, name):
    s = scenes[name]
    gpu_ctx.upload_scene(s)
    gpu_ctx.build_bvh()
    _, _, face_sorted = gpu_ctx.get_morton()
    nodes, root = gpu_ctx.get_traversal_tree()
    assert root == s.n_faces - 2 and nodes.shape == (s.n_faces - 1, 16)
    p = s.pos[face_sorted].astype(np.float32).astype(np.float64)
    check_traversal_tree(nodes, root, s.n_faces, p.min(axis=1), p.max(axis=1))
    # rebuilding gives the same tree bit for bit (node numbering comes from scans, not from atomics)
    gpu_ctx.build_bvh()
    nodes2, _ = gpu_ctx.get_traversal_tree()
    assert np.array_equal(nodes.view(np.uint32), nodes2.view(np.uint32))


@pytest.mark.parametrize("n", [2, 3, 7, 64, 1001, 1025, 5000])
def test_traversal_tree_small_counts(gpu_ctx, n):
    """Sizes around the single-CTA finish threshold (1024) and many equal boxes (far_outside: all codes equal)."""
    desc, keep = random_soup(n, seed=n, far_outside=(n == 64))
    gpu_ctx.upload_scene(desc)
    gpu_ctx.build_bvh()
    _, _, face_sorted = gpu_ctx.get_morton()
    nodes, root = gpu_ctx.get_traversal_tree()
    p = keep[0].reshape(n, 3, 3)[face_sorted].astype(np.float32).astype(np.float64)
    check_traversal_tree(nodes, root, n, p.min(axis=1), p.max(axis=1))


def check_wide_tree(nodes, origin, cell, n, tri_lo, tri_hi, binary_nodes=None):
    """The 4-wide tree the kernels walk (wide.cu): node 0 is the root, every node 0..n_wide-1 is referenced exactly once, every
    sorted triangle sits in exactly one leaf reference, every child's dequantised box contains its subtree, unused slots
    can never be hit, and a leaf pair is at most 15 apart."""
    n_wide = len(nodes)
    q = nodes[:, :, :3]
    lo = origin + (q & 0xFFFF).astype(np.float64) * cell          # (n_wide, 4, 3)
    hi = origin + (q >> 16).astype(np.float64) * cell
    ref = nodes[:, :, 3].view(np.int32)
    unused = np.all((q & 0xFFFF) == 0xFFFF, axis=2) & np.all((q >> 16) == 0, axis=2)
    assert np.all(ref[unused] == np.int32(-2**31))
    assert np.all(np.sum(~unused, axis=1) >= 2), "a wide node has at least two children"
    inner = (ref >= 0) & ~unused
    leaf = (ref < 0) & ~unused
    # internal references: a permutation of 1..n_wide-1
    assert np.array_equal(np.sort(ref[inner]), np.arange(1, n_wide))
    # leaves: every triangle exactly once
    bits = ~ref[leaf]
    k0, d = bits & ((1 << 26) - 1), (bits >> 26) & 15
    assert np.all((bits >> 30) == 0)
    tris = np.concatenate([k0, (k0 + d)[d > 0]])
    if n > 1:
        assert np.array_equal(np.sort(tris), np.arange(n)), "triangles are not covered exactly once"
    # leaf boxes contain their triangles (on the grid, pushed outward)
    llo, lhi = lo[leaf], hi[leaf]
    assert np.all(llo <= tri_lo[k0]) and np.all(lhi >= tri_hi[k0])
    assert np.all(llo[d > 0] <= tri_lo[(k0 + d)[d > 0]]) and np.all(lhi[d > 0] >= tri_hi[(k0 + d)[d > 0]])
    # a child box contains the boxes of that child's children; breadth-first numbering => children have larger indices
    pi, ps = np.nonzero(inner)
    ci = ref[pi, ps]
    assert np.all(ci > pi)
    used = ~unused[ci]                                            # (m, 4)
    clo = np.where(used[:, :, None], lo[ci], np.inf).min(axis=1)
    chi = np.where(used[:, :, None], hi[ci], -np.inf).max(axis=1)
    assert np.all(lo[pi, ps] <= clo + 1e-9) and np.all(hi[pi, ps] >= chi - 1e-9)
    # grid boxes are tight: at most 3 cells (+ float rounding) beyond the FP32 leaf box
    slack = 3.5 * cell + 1e-6 * (1 + np.abs(tri_lo[k0]))
    single = d == 0
    assert np.all(llo[single] >= tri_lo[k0][single] - slack[single]) and np.all(lhi[single] <= tri_hi[k0][single] + slack[single])
    return n_wide


@pytest.mark.parametrize("name", SCENE_NAMES)
def test_wide_tree_invariants_on_shipped_scenes(gpu_ctx, scenes, name):
    s = scenes[name]
    gpu_ctx.upload_scene(s)
    gpu_ctx.build_bvh()
    _, _, face_sorted = gpu_ctx.get_morton()
    nodes, origin, cell = gpu_ctx.get_wide_tree()
    p = s.pos[face_sorted].astype(np.float32).astype(np.float64)
    n_wide = check_wide_tree(nodes, origin.astype(np.float64), cell.astype(np.float64), s.n_faces, p.min(axis=1), p.max(axis=1))
    assert n_wide < 0.55 * s.n_faces, "the collapse should leave well under n/2 nodes"
    # a rebuild gives the same TREE (numbering inside a level may differ: it comes from atomics): same multiset of leaf sets
    gpu_ctx.build_bvh()
    nodes2, _, _ = gpu_ctx.get_wide_tree()
    assert len(nodes2) == n_wide
    def leaf_sets(nd):
        r = nd[:, :, 3].view(np.int32)
        return sorted(tuple(sorted(int(x) for x in row if x < 0 and x != -2**31)) for row in r)
    assert leaf_sets(nodes) == leaf_sets(nodes2)


# 40 000: levels wider than the 4 096 threads of the small-tree cluster (several trips per level); 65 537 / 65 538: the last size on
# the cluster path (candidates + level pass) and the first on the polling kernel
@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 7, 64, 1001, 1025, 5000, 40000, 65537, 65538])
def test_wide_tree_small_counts(gpu_ctx, n):
    desc, keep = random_soup(n, seed=n, far_outside=(n == 64))
    gpu_ctx.upload_scene(desc)
    gpu_ctx.build_bvh()
    _, _, face_sorted = gpu_ctx.get_morton()
    nodes, origin, cell = gpu_ctx.get_wide_tree()
    p = keep[0].reshape(n, 3, 3)[face_sorted].astype(np.float32).astype(np.float64)
    if n == 1:   # degenerate: the root names the only triangle twice
        r = nodes[0, :, 3].view(np.int32)
        assert len(nodes) == 1 and r[0] == ~0 and r[1] == ~0
        return
    check_wide_tree(nodes, origin.astype(np.float64), cell.astype(np.float64), n, p.min(axis=1), p.max(axis=1))


def test_wide_tree_large_mesh(gpu_ctx):
    """1 M triangles: more wide nodes than the collapse kernel has resident threads, so its threads walk several entries
    each (the multi-round path of wide.cu); same invariants, plus traversal against brute force on a few rays."""
    n = 1 << 20
    desc, keep = random_soup(n, seed=11, extent=2.4)
    gpu_ctx.upload_scene(desc)
    gpu_ctx.build_bvh()
    _, _, face_sorted = gpu_ctx.get_morton()
    nodes, origin, cell = gpu_ctx.get_wide_tree()
    p = keep[0].reshape(n, 3, 3)[face_sorted].astype(np.float32).astype(np.float64)
    n_wide = check_wide_tree(nodes, origin.astype(np.float64), cell.astype(np.float64), n, p.min(axis=1), p.max(axis=1))
    assert 0.3 * n < n_wide < 0.55 * n
    # closest hit of 64 rays through the cloud vs a brute-force Möller–Trumbore in FP64 over all triangles
    rng = np.random.default_rng(3)
    o = rng.uniform(0.0, 3.0, size=(64, 3)); d = rng.normal(size=(64, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.concatenate([o, d], axis=1).astype(np.float32)
    hits = gpu_ctx.trace(rays)
    tri = keep[0].reshape(n, 3, 3)
    v0, e1, e2 = tri[:, 0], tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0]
    for i in range(64):
        oo, dd = rays[i, :3].astype(np.float64), rays[i, 3:].astype(np.float64)
        pv = np.cross(dd, e2); det = np.einsum("ij,ij->i", e1, pv)
        with np.errstate(divide="ignore", invalid="ignore"):
            inv = 1.0 / det
            tv = oo - v0
            u = np.einsum("ij,ij->i", tv, pv) * inv
            qv = np.cross(tv, e1)
            v = (qv @ dd) * inv
            t = np.einsum("ij,ij->i", e2, qv) * inv
        ok = (u >= 0) & (v >= 0) & (u + v <= 1) & (t > 0) & np.isfinite(t)
        if not ok.any():
            assert hits["tri"][i] < 0 or hits["t"][i] > 0     # a grazing edge hit the FP64 test excludes is tolerated
            continue
        best = np.flatnonzero(ok)[np.argmin(t[ok])]
        if hits["tri"][i] != best:                            # only a near-tie may differ
            assert hits["tri"][i] >= 0 and abs(hits["t"][i] - t[best]) <= 1e-4 * t[best], (i, hits["tri"][i], best)
        else:
            assert abs(hits["t"][i] - t[best]) <= 1e-5 * max(t[best], 1e-3)


def test_tie_heavy_input_builds_quickly_and_correctly(gpu_ctx):
    """Identical and regularly spaced boxes make every merged area equal.  The quality pass breaks ties toward the buddy i^1 (then
    the lowest index), so a round still merges ~half of the clusters instead of ONE (ploc.cu ploc_nearest); the tree must stay
    shallow enough for the traversal stack and every triangle must be found."""
    from montecarlopathtracing_b200.api import make_desc
    n = 40000
    base = np.array([[0.0, 0.0, 0.0], [0.01, 0.0, 0.0], [0.0, 0.01, 0.0]])
    pos = np.repeat(base[None], n, axis=0)
    pos[n // 2:, :, 0] += (np.arange(n - n // 2) * (1.0 / 64))[:, None] % 3.0      # second half: a regular strip, equal spacing
    nrm = np.tile(np.array([0.0, 0.0, 1.0]), (n, 3, 1)); uv = np.zeros((n, 3, 2))
    desc, keep = make_desc(pos, nrm, uv, np.zeros(n, np.int32), [{"kd": (0.5, 0.5, 0.5)}], [])
    gpu_ctx.upload_scene(desc)
    ms = gpu_ctx.build_bvh()
    assert ms < 50.0, "tie-heavy build took %.1f ms" % ms
    nodes, origin, cell = gpu_ctx.get_wide_tree()
    _, _, face_sorted = gpu_ctx.get_morton()
    p = pos[face_sorted].astype(np.float32).astype(np.float64)
    check_wide_tree(nodes, origin.astype(np.float64), cell.astype(np.float64), n, p.min(axis=1), p.max(axis=1))
    # a ray into the stack of n/2 identical triangles: the reference's tie rule picks the lowest sorted leaf
    hits = gpu_ctx.trace(np.array([[0.002, 0.002, 1.0, 1e-4, 1e-4, -1.0]], np.float32))
    assert hits["leaf"][0] >= 0 and hits["tri"][0] == face_sorted[hits["leaf"][0]]
    tri_lo = pos.min(axis=1)
    same_box = np.nonzero((tri_lo == tri_lo[hits["tri"][0]]).all(axis=1))[0]
    leaf_of = np.empty(n, np.int64); leaf_of[face_sorted] = np.arange(n)
    assert hits["leaf"][0] == leaf_of[same_box].min()


def test_non_finite_vertices_are_rejected(gpu_ctx):
    import montecarlopathtracing_b200 as M
    for bad in (np.nan, np.inf, 1e19):
        desc, keep = random_soup(64, seed=5)
        keep[0][100] = bad
        gpu_ctx.upload_scene(desc)
        with pytest.raises(M.McptError, match="not finite"):
            gpu_ctx.build_bvh()
        with pytest.raises(M.McptError, match="build the BVH first"):
            gpu_ctx.trace(np.zeros((1, 6), np.float32))
    desc, _ = random_soup(64, seed=5)
    gpu_ctx.upload_scene(desc); gpu_ctx.build_bvh()


def test_build_succeeds_while_another_stream_keeps_the_gpu_busy(gpu_ctx, scenes):
    """The collapse and the PLOC rounds need every CTA of their grid resident; they are launched cooperatively, so a busy
    device delays them instead of deadlocking them (round 1 sized an ordinary launch from an occupancy query)."""
    import torch
    s = scenes["cornell-box"]
    gpu_ctx.upload_scene(s)
    gpu_ctx.build_bvh()
    want = gpu_ctx.get_wide_tree()[0].shape[0]
    side = torch.cuda.Stream()
    a = torch.randn(8192, 8192, device="cuda")
    with torch.cuda.stream(side):
        for _ in range(40):
            a = torch.tanh(a @ a * 1e-4)
    for _ in range(5):
        gpu_ctx.build_bvh()
        assert gpu_ctx.get_wide_tree()[0].shape[0] == want
    torch.cuda.synchronize()
