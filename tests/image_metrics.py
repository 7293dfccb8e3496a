"""Image comparison helpers shared by the CPU and GPU parity tests."""
import numpy as np


def block_mean(img, b, mask=None):
    """Mean over b x b blocks; with `mask` (H x W bool) only the selected pixels count (blocks without any are dropped)."""
    h, w = img.shape[0] // b * b, img.shape[1] // b * b
    x = np.asarray(img, np.float64)[:h, :w]
    if mask is None:
        return x.reshape(h // b, b, w // b, b, -1).mean(axis=(1, 3))
    m = np.asarray(mask, np.float64)[:h, :w, None]
    s = (x * m).reshape(h // b, b, w // b, b, -1).sum(axis=(1, 3))
    n = m.reshape(h // b, b, w // b, b, 1).sum(axis=(1, 3))
    keep = n[..., 0] > 0
    return s[keep] / n[keep]


def rel_mse(a, b, block=8, clamp=1.0, eps=1e-2, mask=None):
    """Relative MSE of block-averaged, display-clamped images: mean((A-B)^2 / (B^2 + eps)).
    The reference's output stage clamps to [0,1] (MTPC/MTPC.cpp:26-28), so emitters (radiance 50) do not dominate."""
    A = block_mean(np.minimum(np.asarray(a, np.float64), clamp), block, mask)
    B = block_mean(np.minimum(np.asarray(b, np.float64), clamp), block, mask)
    return float(np.mean((A - B) ** 2 / (B ** 2 + eps)))


def rel_mse_log(a, b, block=16, eps=1e-2, mask=None):
    """The same on UNCLAMPED radiance, compressed by log(1 + x) after block averaging: the display clamp throws away everything
    above 1.0 (46 % of veach-mis' pixels, its glossy highlights), this metric keeps it; `mask` leaves out the pixels that see an
    emitter directly (radiance 50-900, no estimator involved).  Blocks of 16 x 16: a single specular hit on veach-mis' Light1 is
    worth 631 / spp in one pixel, and with 8 x 8 blocks the metric's own run-to-run scatter (0.0024-0.0083 between pairs of
    independent 100-spp renders, reference and restatement alike) is as large as the tolerance; at 16 x 16 it is 0.0008-0.0017."""
    A = np.log1p(np.maximum(block_mean(a, block, mask), 0.0))
    B = np.log1p(np.maximum(block_mean(b, block, mask), 0.0))
    return float(np.mean((A - B) ** 2 / (B ** 2 + eps)))


def mean_rgb(a, clamp=1.0, mask=None):
    x = np.minimum(np.asarray(a, np.float64), clamp)
    return x.mean(axis=(0, 1)) if mask is None else x[np.asarray(mask, bool)].mean(axis=0)
