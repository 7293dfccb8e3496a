"""Image comparison helpers shared by the CPU and GPU parity tests."""
import numpy as np


def block_mean(img, b):
    h, w = img.shape[0] // b * b, img.shape[1] // b * b
    return img[:h, :w].reshape(h // b, b, w // b, b, -1).mean(axis=(1, 3))


def rel_mse(a, b, block=8, clamp=1.0, eps=1e-2):
    """Relative MSE of block-averaged, display-clamped images: mean((A-B)^2 / (B^2 + eps)).
    The reference's output stage clamps to [0,1] (MTPC/MTPC.cpp:26-28), so emitters (radiance 50) do not dominate."""
    A = block_mean(np.minimum(np.asarray(a, np.float64), clamp), block)
    B = block_mean(np.minimum(np.asarray(b, np.float64), clamp), block)
    return float(np.mean((A - B) ** 2 / (B ** 2 + eps)))


def mean_rgb(a, clamp=1.0):
    return np.minimum(np.asarray(a, np.float64), clamp).mean(axis=(0, 1))
