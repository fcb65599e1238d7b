"""YOLACT priors: drop-in for `inst_model/yolact/utils/anchors.py` (`get_anchors` :39-48, `make_anchors` :7-20,
`get_img_output_length` :25-37).

The reference builds the [A,4] table with nested Python loops over every cell (50 ms at 550 x 550); it is a batch constant,
so here it is one vectorised float64 numpy expression per level (same operations in the same order, hence the same bits),
cached per (input_shape, anchors_size) and, for `anchors_tensor`, uploaded once per device as the float32 tensor the wrapper
keeps (`yolact.py:81-83`).  Layout: levels P3 -> P7, cells row-major, aspect ratios [1, 1/2, 2] innermost;
(cx, cy, w, h) normalised by the input size.
"""
from __future__ import annotations

import functools
from math import sqrt

import numpy as np

ASPECT_RATIOS = (1, 1 / 2, 2)          # anchors.py:7


def get_img_output_length(height, width):
    """anchors.py:25-37: sizes of the five prediction levels (seven stride-2 stages, the last five are P3..P7)."""
    filter_sizes = [7, 3, 3, 3, 3, 3, 3]
    padding = [3, 1, 1, 1, 1, 1, 1]
    hs, ws = [], []
    for k, p in zip(filter_sizes, padding):
        height = (height + 2 * p - k) // 2 + 1
        width = (width + 2 * p - k) // 2 + 1
        hs.append(height)
        ws.append(width)
    return np.array(hs)[-5:], np.array(ws)[-5:]


def make_anchors(conv_h, conv_w, scale, input_shape=(550, 550), aspect_ratios=ASPECT_RATIOS):
    """anchors.py:7-20 for one level, as a float64 [conv_h * conv_w * len(aspect_ratios), 4] array."""
    x = (np.arange(conv_w, dtype=np.float64) + 0.5) / conv_w            # :10
    y = (np.arange(conv_h, dtype=np.float64) + 0.5) / conv_h            # :11
    roots = [sqrt(ar) for ar in aspect_ratios]                           # :14
    w = np.array([scale * r / input_shape[1] for r in roots])            # :15
    h = np.array([scale / r / input_shape[0] for r in roots])            # :16
    out = np.empty((conv_h, conv_w, len(roots), 4), np.float64)
    out[..., 0] = x[None, :, None]
    out[..., 1] = y[:, None, None]
    out[..., 2] = w[None, None, :]
    out[..., 3] = h[None, None, :]
    return out.reshape(-1, 4)


@functools.lru_cache(maxsize=16)
def _anchors_cached(input_shape, anchors_size):
    fh, fw = get_img_output_length(input_shape[0], input_shape[1])
    levels = [make_anchors(int(fh[i]), int(fw[i]), anchors_size[i], input_shape) for i in range(len(fh))]
    out = np.concatenate(levels, axis=0)
    out.setflags(write=False)
    return out


def get_anchors(input_shape=[550, 550], anchors_size=[24, 48, 96, 192, 384]):
    """anchors.py:39-48: float64 [A, 4] priors (19248 rows at 550 x 550, 18525 at 544 x 544)."""
    return np.array(_anchors_cached(tuple(int(v) for v in input_shape), tuple(anchors_size)))


_device_cache = {}


def anchors_tensor(input_shape=(550, 550), anchors_size=(24, 48, 96, 192, 384), device="cuda"):
    """The float32 tensor the wrapper keeps on the GPU (`torch.from_numpy(get_anchors(..)).type(torch.FloatTensor).cuda()`,
    yolact.py:81-83), built once per (shape, sizes, device)."""
    import torch
    dev = torch.device(device)
    key = (tuple(int(v) for v in input_shape), tuple(anchors_size), str(dev))
    t = _device_cache.get(key)
    if t is None:
        t = torch.from_numpy(get_anchors(key[0], key[1])).type(torch.FloatTensor).to(dev)
        _device_cache[key] = t
    return t


def proto_size(input_shape=(550, 550)):
    """Side lengths of the prototype map: ProtoNet up-samples P3 by two (nets/yolact.py:70-93): 138 at 550, 136 at 544."""
    fh, fw = get_img_output_length(input_shape[0], input_shape[1])
    return 2 * int(fh[0]), 2 * int(fw[0])
