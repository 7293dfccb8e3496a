"""Deterministic synthetic scene of BASELINE.json configs[3] ("synthetic 5M-triangle subdivided-mesh scene, LBVH build +
traversal stress, 1920x1080, SPP 64"), generated in memory as a mcpt_scene_desc.

A closed room [-0.9, 3.9]^3 — inside the reference's fixed Morton cube [-1, 4]^3 (MTPC/morton code.h:6-7), so no code
clamps — whose floor is a finely tessellated, displaced height field (two triangles per grid cell), plus five walls
(10 triangles) and a one-by-one area light under the ceiling (2 triangles).  The triangle count is EXACTLY `n_tris`
(even, as the reference's BVH requires): the last cells of the grid are dropped to make room for the
12 wall/light triangles.  Vertex coordinates are float32-representable (the FP64 -> FP32 narrowing of the Morton centroid
and of the traversal layout is then exact per vertex); heights carry a seeded 1e-3 perturbation.
"""
import numpy as np

from . import _capi as capi
from .api import make_desc

LO, HI = -0.9, 3.9


def _height(x, z, rng_noise):
    h = 0.45 + 0.25 * np.sin(2.1 * x + 0.3) * np.cos(1.7 * z - 0.2) + 0.12 * np.sin(7.3 * x - 1.1 * z) + 0.06 * np.cos(13.0 * z + 2.0 * x)
    return LO + h + rng_noise


def _grad(x, z):
    dhdx = 0.25 * 2.1 * np.cos(2.1 * x + 0.3) * np.cos(1.7 * z - 0.2) + 0.12 * 7.3 * np.cos(7.3 * x - 1.1 * z) - 0.06 * 2.0 * np.sin(13.0 * z + 2.0 * x)
    dhdz = -0.25 * 1.7 * np.sin(2.1 * x + 0.3) * np.sin(1.7 * z - 0.2) - 0.12 * 1.1 * np.cos(7.3 * x - 1.1 * z) - 0.06 * 13.0 * np.sin(13.0 * z + 2.0 * x)
    return dhdx, dhdz


def grid_shape(n_tris):
    """(nx, nz) cells with 2*nx*nz >= n_tris - 12 and nx/nz ~ 2.5 (2500 x 1000 for 5M)."""
    cells = (n_tris - 12 + 1) // 2
    nz = max(1, int(np.sqrt(cells / 2.5)))
    nx = (cells + nz - 1) // nz
    return nx, nz


def make_scene(n_tris=5_000_000, seed=1234, width=1920, height=1080):
    """Returns (desc, keepalive, camera).  n_tris must be even and >= 14."""
    if n_tris % 2 or n_tris < 14:
        raise ValueError("n_tris must be even and >= 14")
    nx, nz = grid_shape(n_tris)
    n_grid = n_tris - 12
    rng = np.random.default_rng(seed)
    xs = np.linspace(LO, HI, nx + 1).astype(np.float32).astype(np.float64)
    zs = np.linspace(LO, HI, nz + 1).astype(np.float32).astype(np.float64)
    X, Z = np.meshgrid(xs, zs, indexing="ij")                                    # (nx+1, nz+1)
    Y = _height(X, Z, rng.uniform(-1e-3, 1e-3, X.shape)).astype(np.float32).astype(np.float64)
    gx, gz = _grad(X, Z)
    N = np.stack([-gx, np.ones_like(gx), -gz], axis=-1)
    N /= np.linalg.norm(N, axis=-1, keepdims=True)
    P = np.stack([X, Y, Z], axis=-1)
    # two triangles per cell, counter-clockwise seen from above: (p00, p01, p11) and (p00, p11, p10)
    def corners(A):
        return A[:-1, :-1], A[:-1, 1:], A[1:, 1:], A[1:, :-1]
    p00, p01, p11, p10 = corners(P)
    n00, n01, n11, n10 = corners(N)
    pos = np.stack([np.stack([p00, p01, p11], axis=2), np.stack([p00, p11, p10], axis=2)], axis=2).reshape(-1, 3, 3)[:n_grid]
    nrm = np.stack([np.stack([n00, n01, n11], axis=2), np.stack([n00, n11, n10], axis=2)], axis=2).reshape(-1, 3, 3)[:n_grid]
    # walls (normals point into the room) and the light
    def quad(a, b, c, d, n):
        a, b, c, d = (np.array(v, np.float64) for v in (a, b, c, d))
        return np.array([[a, b, c], [a, c, d]]), np.broadcast_to(np.array(n, np.float64), (2, 3, 3)).copy()
    walls = [quad((LO, LO, LO), (HI, LO, LO), (HI, HI, LO), (LO, HI, LO), (0, 0, 1)),      # back  z = LO
             quad((LO, LO, HI), (LO, HI, HI), (HI, HI, HI), (HI, LO, HI), (0, 0, -1)),     # front z = HI
             quad((LO, LO, LO), (LO, HI, LO), (LO, HI, HI), (LO, LO, HI), (1, 0, 0)),      # left  x = LO
             quad((HI, LO, LO), (HI, LO, HI), (HI, HI, HI), (HI, HI, LO), (-1, 0, 0)),     # right x = HI
             quad((LO, HI, LO), (HI, HI, LO), (HI, HI, HI), (LO, HI, HI), (0, -1, 0))]     # ceiling
    light = quad((1.0, 3.85, 1.0), (2.0, 3.85, 1.0), (2.0, 3.85, 2.0), (1.0, 3.85, 2.0), (0, -1, 0))
    pos = np.concatenate([pos] + [w[0] for w in walls] + [light[0]], axis=0).astype(np.float32).astype(np.float64)
    nrm = np.concatenate([nrm] + [w[1] for w in walls] + [light[1]], axis=0)
    n = pos.shape[0]
    assert n == n_tris
    mat = np.zeros(n, np.int32)
    mat[n_grid:n_grid + 10] = 1
    mat[n_grid + 10:] = 2
    uv = np.zeros((n, 3, 2))
    materials = [{"kd": (0.7, 0.7, 0.7)}, {"kd": (0.55, 0.6, 0.7)}, {"kd": (0, 0, 0)}]
    lights = [{"mat_id": 2, "radiance": (30.0, 30.0, 30.0)}]
    desc, keep = make_desc(pos, nrm, uv, mat, materials, lights)
    cam = capi.Camera()
    cam.eye[:] = (1.37, 2.6, 3.7); cam.lookat[:] = (1.45, 1.9, 3.0); cam.up[:] = (0.0, 1.0, 0.0)
    cam.fovy = 50.0
    cam.width, cam.height = width, height
    return desc, keep, cam
