"""Image metrics of the reference's evaluate driver (src/image_utils.h:20-99), on the host.

  psnr8      computePsnr (:84-99): MSE over the three 8-bit sRGB channels, 100 dB for identical images.
  flip_ldr   computeFlip (:26-81): mean LDR-FLIP (Andersson et al., HPG 2020) of two 8-bit sRGB images at
             the reference's viewing conditions (0.7 m from a 0.7 m wide 3840-pixel monitor, :20-22).

The reference calls NVIDIA's CUDA implementation (ext/flip-cuda); this is an independent numpy restatement of
the published algorithm with that library's constants (common/sharedflip.h:10-24: qc 0.7, pc 0.4, pt 0.95,
w 0.082, qf 0.5; CSF Gaussians a1/b1/a2/b2) and its clamp-to-edge filtering.  Test infrastructure and the
evaluate-equivalent driver use it; nothing on the render path does.
"""
import numpy as np

try:  # scipy is in the image; fall back to a plain separable loop if it is not
    from scipy.ndimage import correlate as _correlate
except Exception:  # pragma: no cover
    _correlate = None

_QC, _PC, _PT, _W, _QF = 0.7, 0.4, 0.95, 0.082, 0.5
_CSF = {"A": (1.0, 0.0047, 0.0, 1e-5), "RG": (1.0, 0.0053, 0.0, 1e-5), "BY": (34.1, 0.04, 13.5, 0.025)}
_RGB2XYZ = np.array([[0.4124564391, 0.3575760776, 0.1804374833],
                     [0.2126728514, 0.7151579067, 0.0721749933],
                     [0.0193338956, 0.1191920335, 0.9503040785]])
_XYZ2RGB = np.linalg.inv(_RGB2XYZ)
_WHITE = _RGB2XYZ @ np.ones(3)


def pixels_per_degree(distance_m=0.7, resolution_x=3840.0, width_m=0.7):
    """calculatePPD, src/image_utils.h:20-22."""
    return distance_m * (resolution_x / width_m) * (np.pi / 180.0)


def psnr8(ref, test):
    d = ref[..., :3].astype(np.float64) - test[..., :3].astype(np.float64)
    mse = float(np.mean(d * d))
    return 100.0 if mse < 1e-10 else 10.0 * np.log10(255.0 * 255.0 / mse)


def _srgb_to_linear(c):
    return np.where(c <= 0.04045, c / 12.92, ((c + 0.055) / 1.055) ** 2.4)


def _linear_to_ycxcz(rgb):
    xyz = rgb @ _RGB2XYZ.T / _WHITE
    y = 116.0 * xyz[..., 1] - 16.0
    return np.stack([y, 500.0 * (xyz[..., 0] - xyz[..., 1]), 200.0 * (xyz[..., 1] - xyz[..., 2])], -1)


def _ycxcz_to_linear(ycc):
    y = (ycc[..., 0] + 16.0) / 116.0
    xyz = np.stack([y + ycc[..., 1] / 500.0, y, y - ycc[..., 2] / 200.0], -1) * _WHITE
    return xyz @ _XYZ2RGB.T


def _linear_to_lab(rgb):
    xyz = rgb @ _RGB2XYZ.T / _WHITE
    d = 6.0 / 29.0
    f = np.where(xyz > d ** 3, np.cbrt(xyz), xyz / (3.0 * d * d) + 4.0 / 29.0)
    return np.stack([116.0 * f[..., 1] - 16.0, 500.0 * (f[..., 0] - f[..., 1]), 200.0 * (f[..., 1] - f[..., 2])], -1)


def _hunt(lab):
    out = lab.copy()
    out[..., 1] *= 0.01 * lab[..., 0]
    out[..., 2] *= 0.01 * lab[..., 0]
    return out


def _hyab(a, b):
    d = a - b
    return np.abs(d[..., 0]) + np.sqrt(d[..., 1] ** 2 + d[..., 2] ** 2)


def _filter2d(img, kernel):
    if _correlate is not None:
        return _correlate(img, kernel, mode="nearest")
    r = kernel.shape[0] // 2
    p = np.pad(img, r, mode="edge")
    out = np.zeros_like(img)
    for dy in range(kernel.shape[0]):
        for dx in range(kernel.shape[1]):
            out += kernel[dy, dx] * p[dy:dy + img.shape[0], dx:dx + img.shape[1]]
    return out


def _csf_kernels(ppd):
    bmax = max(max(v[1], v[3]) for v in _CSF.values())
    radius = int(np.ceil(3.0 * np.sqrt(bmax / (2.0 * np.pi ** 2)) * ppd))
    ax = np.arange(-radius, radius + 1) / ppd
    d2 = ax[None, :] ** 2 + ax[:, None] ** 2
    ks = []
    for name in ("A", "RG", "BY"):
        a1, b1, a2, b2 = _CSF[name]
        k = a1 * np.sqrt(np.pi / b1) * np.exp(-np.pi ** 2 * d2 / b1) + a2 * np.sqrt(np.pi / b2) * np.exp(-np.pi ** 2 * d2 / b2)
        ks.append(k / k.sum())
    return ks


def _feature_kernels(ppd):
    sd = 0.5 * _W * ppd
    radius = int(np.ceil(3.0 * sd))
    x = np.arange(-radius, radius + 1, dtype=np.float64)
    g = np.exp(-(x * x) / (2.0 * sd * sd))
    dg = -x * g
    ddg = (x * x / (sd * sd) - 1.0) * g
    norm = lambda v: np.where(v > 0, v / v[v > 0].sum(), v / -v[v < 0].sum())
    return g / g.sum(), norm(dg), norm(ddg)


def _features(y, ppd):
    g, dg, ddg = _feature_kernels(ppd)
    row = lambda k: k[None, :]
    col = lambda k: k[:, None]
    dx = _filter2d(_filter2d(y, row(dg)), col(g))
    dy = _filter2d(_filter2d(y, row(g)), col(dg))
    ddx = _filter2d(_filter2d(y, row(ddg)), col(g))
    ddy = _filter2d(_filter2d(y, row(g)), col(ddg))
    return np.sqrt(dx * dx + dy * dy), np.sqrt(ddx * ddx + ddy * ddy)


def flip_ldr_map(ref, test, ppd=None):
    """Per-pixel LDR-FLIP error in [0, 1] of two uint8 sRGB images (H, W, >=3)."""
    ppd = pixels_per_degree() if ppd is None else ppd
    ycc = [_linear_to_ycxcz(_srgb_to_linear(im[..., :3].astype(np.float64) / 255.0)) for im in (ref, test)]
    # ---- colour pipeline: CSF filtering, Hunt-adjusted Lab, HyAB, redistribution ----
    ks = _csf_kernels(ppd)
    labs = []
    for y in ycc:
        f = np.stack([_filter2d(np.ascontiguousarray(y[..., c]), ks[c]) for c in range(3)], -1)
        labs.append(_hunt(_linear_to_lab(np.clip(_ycxcz_to_linear(f), 0.0, 1.0))))
    green = _hunt(_linear_to_lab(np.array([0.0, 1.0, 0.0])))
    blue = _hunt(_linear_to_lab(np.array([0.0, 0.0, 1.0])))
    cmax = _hyab(green, blue) ** _QC
    de = _hyab(labs[0], labs[1]) ** _QC
    pccmax = _PC * cmax
    de = np.where(de < pccmax, de * (_PT / pccmax), _PT + (de - pccmax) / (cmax - pccmax) * (1.0 - _PT))
    # ---- feature pipeline: edges and points on the achromatic channel ----
    er, pr = _features((ycc[0][..., 0] + 16.0) / 116.0, ppd)
    et, pt = _features((ycc[1][..., 0] + 16.0) / 116.0, ppd)
    df = (np.maximum(np.abs(er - et), np.abs(pr - pt)) / np.sqrt(2.0)) ** _QF
    return np.clip(de, 0.0, 1.0) ** (1.0 - np.clip(df, 0.0, 1.0))


def flip_ldr(ref, test, ppd=None):
    """Mean LDR-FLIP (what computeFlip returns)."""
    return float(np.mean(flip_ldr_map(ref, test, ppd)))
