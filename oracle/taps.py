"""Oracle (test infrastructure): the exact fp32 coordinate sequence of the hot path, in numpy.

"Bit-exact lookup indices and coordinates" (BASELINE.json north_star) needs an oracle for the integer
tap indices and the fp32 fractional weights, which torch's ``grid_sample`` does not expose.  This file
restates, scalar step by scalar step in ``np.float32``:

* lookup (RAFT ``bilinear_sampler`` / torchvision ``_utils.py:8-19`` + ATen ``grid_sampler_unnormalize``
  with ``align_corners=True``):
      c_l = c / 2**l ; p = c_l + d ; n = 2*p/(S-1) - 1 ; u = ((n + 1) / 2) * (S - 1)
* warp (``/root/reference/nnunet/network_architecture/integration.py:63-79`` + the same ATen step):
      p = g + f ; n = 2 * (p/(S-1) - 0.5) ; u = ((n + 1) / 2) * (S - 1)
* taps: i0 = floor(u) ; w_hi = u - i0 ; w_lo = (i0 + 1) - u ; out-of-image taps contribute zero.

The normalise -> un-normalise round trip is NOT the identity in fp32, so the CUDA kernels
perform the same sequence with FMA contraction disabled and are compared bit-for-bit against this file.
The grid_sample-free ``lookup_from_taps`` / ``warp_from_taps`` are a second, independent restatement of the
sampling itself (ATen accumulation order nw, ne, sw, se).
"""
import numpy as np

F32 = np.float32


def _unnormalize(n, size):
    return ((n + F32(1)) / F32(2)) * F32(size - 1)


def lookup_coord(p, size):
    """Absolute pixel coordinate p (fp32 array) -> coordinate actually sampled by grid_sample."""
    p = np.asarray(p, dtype=F32)
    n = (F32(2) * p) / F32(size - 1) - F32(1)
    return _unnormalize(n, size).astype(F32)


def warp_coord(p, size):
    p = np.asarray(p, dtype=F32)
    n = F32(2) * (p / F32(size - 1) - F32(0.5))
    return _unnormalize(n, size).astype(F32)


def split_taps(u):
    """u -> (i0 int32, w_lo fp32 weight of tap i0, w_hi fp32 weight of tap i0+1)."""
    fl = np.floor(u).astype(F32)
    i0 = fl.astype(np.int64).astype(np.int32)
    w_hi = (u - fl).astype(F32)
    w_lo = ((fl + F32(1)) - u).astype(F32)
    return i0, w_lo, w_hi


def pyramid_sizes(h, w, num_levels):
    out = [(h, w)]
    for _ in range(num_levels - 1):
        h, w = h // 2, w // 2
        out.append((h, w))
    return out


def lookup_taps(coords, level_sizes, radius=4):
    """coords [B,2,H,W] fp32 (ch0 = x, ch1 = y).

    Returns per level a dict with ``ux, uy`` fp32 [N, 2r+1] (N = B*H*W; axis 1 = window offset index),
    ``x0, y0`` int32 and ``wx_lo, wx_hi, wy_lo, wy_hi`` fp32 of the same shape.
    """
    coords = np.asarray(coords, dtype=F32)
    b, _, h, w = coords.shape
    cx = coords[:, 0].reshape(-1, 1)
    cy = coords[:, 1].reshape(-1, 1)
    d = np.linspace(-radius, radius, 2 * radius + 1, dtype=F32).reshape(1, -1)
    levels = []
    for lvl, (hl, wl) in enumerate(level_sizes):
        scale = F32(2 ** lvl)
        px = (cx / scale).astype(F32) + d
        py = (cy / scale).astype(F32) + d
        ux = lookup_coord(px, wl)
        uy = lookup_coord(py, hl)
        x0, wx_lo, wx_hi = split_taps(ux)
        y0, wy_lo, wy_hi = split_taps(uy)
        levels.append(dict(ux=ux, uy=uy, x0=x0, y0=y0, wx_lo=wx_lo, wx_hi=wx_hi, wy_lo=wy_lo, wy_hi=wy_hi))
    return levels


def _gather(img, yy, xx):
    """img [N,h,w]; yy, xx int arrays broadcastable to [N, ...]; zero outside the image."""
    n, h, w = img.shape
    ok = (yy >= 0) & (yy < h) & (xx >= 0) & (xx < w)
    yc = np.clip(yy, 0, h - 1)
    xc = np.clip(xx, 0, w - 1)
    idx = np.arange(n).reshape((n,) + (1,) * (yy.ndim - 1))
    return np.where(ok, img[idx, yc, xc], F32(0)).astype(F32)


def lookup_from_taps(pyramid, coords, radius=4):
    """pyramid: list of [N,1,h_l,w_l] (or [N,h_l,w_l]) fp32 arrays.  Returns [B, L*(2r+1)^2, H, W] fp32.

    Channel l*K + i*(2r+1) + j samples level l at (x + (i - r), y + (j - r)).
    """
    coords = np.asarray(coords, dtype=F32)
    b, _, h, w = coords.shape
    side = 2 * radius + 1
    imgs = [np.asarray(p, dtype=F32).reshape(p.shape[0], p.shape[-2], p.shape[-1]) for p in pyramid]
    taps = lookup_taps(coords, [im.shape[-2:] for im in imgs], radius)
    outs = []
    for im, t in zip(imgs, taps):
        x0 = t["x0"][:, :, None]
        y0 = t["y0"][:, None, :]
        wxl, wxh = t["wx_lo"][:, :, None], t["wx_hi"][:, :, None]
        wyl, wyh = t["wy_lo"][:, None, :], t["wy_hi"][:, None, :]
        x0b, y0b = np.broadcast_arrays(x0, y0)
        acc = _gather(im, y0b, x0b) * (wxl * wyl)
        acc = acc + _gather(im, y0b, x0b + 1) * (wxh * wyl)
        acc = acc + _gather(im, y0b + 1, x0b) * (wxl * wyh)
        acc = acc + _gather(im, y0b + 1, x0b + 1) * (wxh * wyh)
        outs.append(acc.reshape(b, h, w, side * side).astype(F32))
    return np.ascontiguousarray(np.concatenate(outs, axis=-1).transpose(0, 3, 1, 2))


def warp_taps(flow):
    """flow [B,2,H,W] (ch0 = row displacement, ch1 = column displacement)."""
    flow = np.asarray(flow, dtype=F32)
    b, _, h, w = flow.shape
    gy = np.arange(h, dtype=F32).reshape(1, h, 1)
    gx = np.arange(w, dtype=F32).reshape(1, 1, w)
    uy = warp_coord(gy + flow[:, 0], h)
    ux = warp_coord(gx + flow[:, 1], w)
    x0, wx_lo, wx_hi = split_taps(ux)
    y0, wy_lo, wy_hi = split_taps(uy)
    return dict(ux=ux, uy=uy, x0=x0, y0=y0, wx_lo=wx_lo, wx_hi=wx_hi, wy_lo=wy_lo, wy_hi=wy_hi)


def warp_from_taps(flow, src):
    """Bilinear, zero padding.  src [B,C,H,W] -> [B,C,H,W]."""
    src = np.asarray(src, dtype=F32)
    b, c, h, w = src.shape
    t = warp_taps(flow)
    out = np.zeros_like(src)
    for ch in range(c):
        im = src[:, ch]
        acc = _gather(im, t["y0"], t["x0"]) * (t["wx_lo"] * t["wy_lo"])
        acc = acc + _gather(im, t["y0"], t["x0"] + 1) * (t["wx_hi"] * t["wy_lo"])
        acc = acc + _gather(im, t["y0"] + 1, t["x0"]) * (t["wx_lo"] * t["wy_hi"])
        acc = acc + _gather(im, t["y0"] + 1, t["x0"] + 1) * (t["wx_hi"] * t["wy_hi"])
        out[:, ch] = acc
    return out


def avg_pool2x2(x):
    """ATen avg_pool2d(2, stride 2, floor): ((a + b) + c) + d, then / 4.  x [N,h,w] -> [N,h//2,w//2]."""
    x = np.asarray(x, dtype=F32)
    h2, w2 = x.shape[-2] // 2, x.shape[-1] // 2
    a = x[..., 0:2 * h2:2, 0:2 * w2:2]
    b = x[..., 0:2 * h2:2, 1:2 * w2:2]
    c = x[..., 1:2 * h2:2, 0:2 * w2:2]
    d = x[..., 1:2 * h2:2, 1:2 * w2:2]
    return ((((a + b) + c) + d) / F32(4)).astype(F32)
