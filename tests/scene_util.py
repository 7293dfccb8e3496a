"""Synthetic scenes for tests and benchmarks (triangle soups and tessellated boxes) as mcpt_scene_desc."""
import numpy as np

from montecarlopathtracing_b200.api import make_desc


def random_soup(n, seed=0, extent=2.0, size=0.2, far_outside=False):
    """n random small triangles with centroids in [-0.9, -0.9+2*extent)^3 (inside the Morton cube unless far_outside),
    one diffuse material, plus the first triangle doubling as a light when n >= 2."""
    rng = np.random.default_rng(seed)
    c = rng.uniform(-0.9, -0.9 + 2 * extent, size=(n, 1, 3))
    if far_outside:
        c = c + 50.0
    pos = (c + rng.uniform(-size, size, size=(n, 3, 3))).astype(np.float32).astype(np.float64)  # float-representable
    nrm = np.cross(pos[:, 1] - pos[:, 0], pos[:, 2] - pos[:, 0])
    nrm /= np.maximum(np.linalg.norm(nrm, axis=1, keepdims=True), 1e-30)
    nrm = np.repeat(nrm[:, None, :], 3, axis=1)
    uv = np.zeros((n, 3, 2))
    mat = np.zeros(n, np.int32)
    materials = [{"kd": (0.7, 0.6, 0.5)}, {"kd": (0, 0, 0)}]
    lights = []
    if n >= 2:
        mat[0] = 1
        lights = [{"mat_id": 1, "radiance": (20.0, 20.0, 20.0)}]
    return make_desc(pos, nrm, uv, mat, materials, lights)
