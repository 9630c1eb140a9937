"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the preprocess arithmetic.

The reference's preprocess (``alexnet_resnet.py:57-66``) is
``Compose[Resize(256), CenterCrop(224), ToTensor(), Normalize(mean, std)]`` on a PIL image.
All of its arithmetic lives in third-party code that is not vendored under /root/reference:

* Pillow (unpinned in ``requirements.txt:2``; 12.2.0 in this image) ``Image.resize(size, BILINEAR)``
  -> ``src/libImaging/Resample.c`` (``precompute_coeffs``, ``normalize_coeffs_8bpc``,
  ``ImagingResampleHorizontal_8bpc`` / ``...Vertical_8bpc``): separable, antialiased triangle
  filter in 22-bit fixed point with a uint8 intermediate after the horizontal pass.  Restated
  here from the published algorithm; pinned against Pillow itself in tests/test_oracle.py.
* torchvision (unpinned, ``requirements.txt:3``; 0.26.0 here):
  output size rule   tv:transforms/functional.py:353-384 (``_compute_resized_output_size``),
  centre crop        tv:transforms/functional.py:556-594,
  ToTensor           tv:transforms/functional.py:127-178 (u8 -> f32, ``div(255)``),
  Normalize          tv:transforms/_functional_tensor.py:916-928 (``sub_(mean).div_(std)``).
"""
from __future__ import annotations

import math

import numpy as np

PRECISION_BITS = 32 - 8 - 2  # Resample.c: 8 bits for the result, 2 guard bits
MEAN = (0.485, 0.456, 0.406)   # alexnet_resnet.py:61
STD = (0.229, 0.224, 0.225)    # alexnet_resnet.py:61
RESIZE_SHORT = 256             # alexnet_resnet.py:58
CROP = 224                     # alexnet_resnet.py:59


def resized_size(h: int, w: int, short: int = RESIZE_SHORT) -> tuple[int, int]:
    """(new_h, new_w) of ``transforms.Resize(short)``: the shorter side becomes ``short``, the
    longer ``int(short * long / short_side)`` (tv:transforms/functional.py:368-370)."""
    s, l = (w, h) if w <= h else (h, w)
    new_s, new_l = short, int(short * l / s)
    return (new_l, new_s) if w <= h else (new_s, new_l)


def crop_offsets(h: int, w: int, crop: int = CROP) -> tuple[int, int]:
    """(top, left) of ``center_crop`` (tv:transforms/functional.py:592-593); Python ``round`` is
    banker's rounding.  Images smaller than the crop are padded by torchvision; that cannot
    happen after Resize(256) and is rejected here."""
    if h < crop or w < crop:
        raise ValueError("center_crop padding path is unreachable after Resize(256)")
    return int(round((h - crop) / 2.0)), int(round((w - crop) / 2.0))


def precompute_coeffs(in_size: int, out_size: int):
    """Pillow ``precompute_coeffs`` + ``normalize_coeffs_8bpc`` for the BILINEAR filter
    (support 1.0) over the full input box.  Returns (xmin[out], xcnt[out], k[out, ksize] int32)."""
    scale = in_size / out_size
    filterscale = max(scale, 1.0)
    support = 1.0 * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    xmin = np.zeros(out_size, dtype=np.int32)
    xcnt = np.zeros(out_size, dtype=np.int32)
    kk = np.zeros((out_size, ksize), dtype=np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        x0 = int(center - support + 0.5)
        if x0 < 0:
            x0 = 0
        x1 = int(center + support + 0.5)
        if x1 > in_size:
            x1 = in_size
        n = x1 - x0
        w = np.zeros(n, dtype=np.float64)
        ww = 0.0
        for x in range(n):
            a = (x + x0 - center + 0.5) * ss
            a = -a if a < 0.0 else a
            v = 1.0 - a if a < 1.0 else 0.0
            w[x] = v
            ww += v
        for x in range(n):
            if ww != 0.0:
                w[x] /= ww
        xmin[xx] = x0
        xcnt[xx] = n
        for x in range(n):
            v = w[x] * (1 << PRECISION_BITS)
            kk[xx, x] = int(v - 0.5) if w[x] < 0 else int(v + 0.5)
    return xmin, xcnt, kk


def _clip8(acc: np.ndarray) -> np.ndarray:
    return np.clip(acc >> PRECISION_BITS, 0, 255).astype(np.uint8)


def resize_bilinear_u8(img: np.ndarray, out_h: int, out_w: int) -> np.ndarray:
    """``Image.fromarray(img).resize((out_w, out_h), BILINEAR)`` for uint8 HWC input: horizontal
    pass -> uint8 -> vertical pass -> uint8.  Pillow skips a pass whose size is unchanged."""
    h, w, c = img.shape
    cur = img
    if out_w != w:
        xmin, xcnt, kk = precompute_coeffs(w, out_w)
        out = np.empty((h, out_w, c), dtype=np.uint8)
        src = cur.astype(np.int64)
        for xx in range(out_w):
            n = int(xcnt[xx]); x0 = int(xmin[xx])
            acc = (src[:, x0:x0 + n, :] * kk[xx, :n].astype(np.int64)[None, :, None]).sum(axis=1)
            out[:, xx, :] = _clip8(acc + (1 << (PRECISION_BITS - 1)))
        cur = out
    if out_h != h:
        ymin, ycnt, kk = precompute_coeffs(h, out_h)
        out = np.empty((out_h, cur.shape[1], c), dtype=np.uint8)
        src = cur.astype(np.int64)
        for yy in range(out_h):
            n = int(ycnt[yy]); y0 = int(ymin[yy])
            acc = (src[y0:y0 + n, :, :] * kk[yy, :n].astype(np.int64)[:, None, None]).sum(axis=0)
            out[yy, :, :] = _clip8(acc + (1 << (PRECISION_BITS - 1)))
        cur = out
    return cur


def preprocess_u8(img: np.ndarray) -> np.ndarray:
    """Resize(256) + CenterCrop(224) in the uint8 domain -> uint8 [224, 224, 3]."""
    h, w, _ = img.shape
    nh, nw = resized_size(h, w)
    r = img if (nh, nw) == (h, w) else resize_bilinear_u8(img, nh, nw)
    top, left = crop_offsets(nh, nw)
    return r[top:top + CROP, left:left + CROP, :]


def preprocess(img: np.ndarray) -> np.ndarray:
    """Full preprocess -> float32 CHW [3, 224, 224], fp32 arithmetic in torchvision's order:
    ``x = u8 / 255`` then ``(x - mean) / std`` with mean/std as float32 tensors."""
    u8 = preprocess_u8(img)
    x = u8.astype(np.float32).transpose(2, 0, 1) / np.float32(255.0)
    mean = np.asarray(MEAN, dtype=np.float32)[:, None, None]
    std = np.asarray(STD, dtype=np.float32)[:, None, None]
    return ((x - mean) / std).astype(np.float32)
