"""CPU restatement of the wrappers' image pre-processing (test infrastructure; see oracle/__init__.py).

Both wrappers turn a PIL image into the network input the same way
(inst_model/yolact/yolact.py:137-145, inst_model/Mask_RCNN/mask_rcnn.py:140-148):

    image_data = resize_image(image, (W_in, H_in))                      # utils/utils.py:19-22, Image.BICUBIC
    image_data = np.transpose(preprocess_input(np.array(image_data, np.float32)), (2, 0, 1))[None]   # utils.py:32-36
    image_data = torch.from_numpy(image_data).type(torch.FloatTensor)

`Image.resize(..., Image.BICUBIC)` is Pillow, a third-party dependency that is not vendored under /root/reference and
that the reference does not pin (no requirements file); the installed version here is Pillow 12.2.0.  `resize_u8`
restates the published algorithm of Pillow's src/libImaging/Resample.c for 8-bit images (precompute_coeffs,
normalize_coeffs_8bpc, ImagingResampleHorizontal_8bpc / Vertical_8bpc: a horizontal pass with an 8-bit intermediate
image, then a vertical pass; fixed-point coefficients with 22 fractional bits).  Parity pin: tests/golden/preprocess*.npz
hold outputs of Pillow itself (oracle/gen_golden_preprocess.py) and tests/test_oracle_preprocess.py checks this
restatement against them bit for bit; when Pillow is importable the test also compares directly on fresh inputs.

`preprocess_input` divides in float64 (numpy promotes the float32 image against the Python-tuple mean / std) and the
wrapper rounds to float32 afterwards; `normalize_chw` does exactly that.
"""
from __future__ import annotations

import math

import numpy as np

PRECISION_BITS = 32 - 8 - 2          # Resample.c: coefficients carry 22 fractional bits
BICUBIC_SUPPORT = 2.0
MEAN = (123.68, 116.78, 103.94)      # utils/utils.py:33-34
STD = (58.40, 57.12, 57.38)


def bicubic_filter(x):
    """Resample.c bicubic_filter (Keys kernel, a = -0.5), in double."""
    a = -0.5
    if x < 0.0:
        x = -x
    if x < 1.0:
        return ((a + 2.0) * x - (a + 3.0)) * x * x + 1
    if x < 2.0:
        return (((x - 5) * x + 8) * x - 4) * a
    return 0.0


def precompute_coeffs(in_size, out_size):
    """Resample.c precompute_coeffs + normalize_coeffs_8bpc for the whole axis (box = (0, in_size)):
    returns ksize, bounds int32 [out, 2] (first tap, tap count) and coefficients int32 [out, ksize]."""
    scale = float(in_size) / float(out_size)
    filterscale = scale if scale >= 1.0 else 1.0
    support = BICUBIC_SUPPORT * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    bounds = np.zeros((out_size, 2), np.int32)
    kk = np.zeros((out_size, ksize), np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        xmin = int(center - support + 0.5)          # C cast: truncation toward zero
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > in_size:
            xmax = in_size
        xmax -= xmin
        w = [bicubic_filter((x + xmin - center + 0.5) * ss) for x in range(xmax)]
        ww = 0.0
        for v in w:
            ww += v
        if ww != 0.0:
            w = [v / ww for v in w]
        for x, v in enumerate(w):
            q = v * (1 << PRECISION_BITS)
            kk[xx, x] = int(-0.5 + q) if v < 0 else int(0.5 + q)
        bounds[xx] = (xmin, xmax)
    return ksize, bounds, kk


def _clip8(acc):
    return np.clip(acc >> PRECISION_BITS, 0, 255).astype(np.uint8)


def _resample_axis0(img, out_size):
    """One pass along axis 0 of a uint8 [n, ...] array (ImagingResampleVertical_8bpc; the horizontal pass is the same
    arithmetic along the other axis)."""
    in_size = img.shape[0]
    _, bounds, kk = precompute_coeffs(in_size, out_size)
    out = np.empty((out_size,) + img.shape[1:], np.uint8)
    src = img.astype(np.int64)
    for xx in range(out_size):
        xmin, xmax = int(bounds[xx, 0]), int(bounds[xx, 1])
        acc = np.full(img.shape[1:], 1 << (PRECISION_BITS - 1), np.int64)
        for x in range(xmax):
            acc += src[xmin + x] * int(kk[xx, x])
        out[xx] = _clip8(acc)
    return out


def resize_u8(img, out_w, out_h):
    """`PIL.Image.fromarray(img).resize((out_w, out_h), Image.BICUBIC)` for a uint8 [H, W, C] array
    (utils/utils.py:19-22): horizontal pass first (skipped when the width does not change), then vertical."""
    img = np.ascontiguousarray(img, dtype=np.uint8)
    h, w = img.shape[:2]
    if w != out_w:
        img = np.swapaxes(_resample_axis0(np.swapaxes(img, 0, 1), out_w), 0, 1)
    if h != out_h:
        img = _resample_axis0(img, out_h)
    return np.ascontiguousarray(img)


def normalize_chw(img_u8, mean=MEAN, std=STD):
    """utils/utils.py:32-36 + yolact.py:145-148: float32 [1, 3, H, W] = float32((float64(pixel) - mean) / std)."""
    x = np.array(img_u8, np.float32)
    x = (x - np.asarray(mean, np.float64)) / np.asarray(std, np.float64)
    return np.transpose(x, (2, 0, 1))[None].astype(np.float32)


def image_to_input(img_u8, input_shape):
    """The wrappers' whole pre-processing: uint8 [H, W, 3] -> float32 [1, 3, H_in, W_in]."""
    return normalize_chw(resize_u8(img_u8, int(input_shape[1]), int(input_shape[0])))
