"""Image metrics of the parity gate.

SSIM: the reference's eval.py:25-37 shells out to ImageMagick `compare -metric SSIM`; ImageMagick is
not installed here, so this restates its defaults (Wang et al. 2004: 11x11 Gaussian window, sigma 1.5,
K1 = 0.01, K2 = 0.03, L = 1, per channel, mean over the image, mean over R, G, B).  A restatement,
not the `magick` binary.
"""
import numpy as np
from scipy.ndimage import correlate1d


def _gauss_kernel(radius=5, sigma=1.5):
    x = np.arange(-radius, radius + 1, dtype=np.float64)
    k = np.exp(-(x * x) / (2.0 * sigma * sigma))
    return k / k.sum()


def _blur(img, k):
    # ImageMagick uses edge-replicating virtual pixels for the compare window
    return correlate1d(correlate1d(img, k, axis=0, mode="nearest"), k, axis=1, mode="nearest")


def ssim(a, b, radius=5, sigma=1.5, k1=0.01, k2=0.03):
    """a, b: HxWxC arrays, uint8 (scaled by 1/255) or float in [0,1]. Returns the mean SSIM over channels."""
    def prep(x):
        x = np.asarray(x)
        return x.astype(np.float64) / 255.0 if x.dtype == np.uint8 else x.astype(np.float64)
    a, b = prep(a), prep(b)
    if a.ndim == 2:
        a, b = a[..., None], b[..., None]
    k = _gauss_kernel(radius, sigma)
    c1, c2 = (k1 * 1.0) ** 2, (k2 * 1.0) ** 2
    vals = []
    for c in range(a.shape[2]):
        x, y = a[..., c], b[..., c]
        mx, my = _blur(x, k), _blur(y, k)
        sxx = _blur(x * x, k) - mx * mx
        syy = _blur(y * y, k) - my * my
        sxy = _blur(x * y, k) - mx * my
        s = ((2 * mx * my + c1) * (2 * sxy + c2)) / ((mx * mx + my * my + c1) * (sxx + syy + c2))
        vals.append(s.mean())
    return float(np.mean(vals))


def max_abs_delta(a, b):
    """Per-channel max |a-b| in units of 1/255 for uint8 inputs, or of the float range otherwise."""
    a, b = np.asarray(a), np.asarray(b)
    if a.dtype == np.uint8:
        return float(np.abs(a.astype(np.int32) - b.astype(np.int32)).max())
    return float(np.abs(a.astype(np.float64) - b.astype(np.float64)).max())
