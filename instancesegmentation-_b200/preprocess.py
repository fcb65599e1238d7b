"""Drop-ins for the wrappers' image pre-processing (inst_model/yolact/utils/utils.py:19-36 as used by
inst_model/yolact/yolact.py:137-148 and inst_model/Mask_RCNN/mask_rcnn.py:140-148), computed by libisb.so on the GPU:
Pillow's 8-bit bicubic resize reproduced bit for bit, the float64 normalisation, HWC -> CHW, float32.

    image_data = isb.preprocess.image_to_input(image, self.input_shape)      # replaces yolact.py:141-150

There is no CPU path: the library must be present and a CUDA device available.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import get_handle

MEAN = (123.68, 116.78, 103.94)   # utils/utils.py:33
STD = (58.40, 57.12, 57.38)      # utils/utils.py:34


def _as_u8_batch(images, device):
    """PIL image / uint8 ndarray / uint8 tensor, [H,W,3] or [B,H,W,3] -> (cuda uint8 [B,H,W,3], was_single)."""
    def rgb(im):  # utils/utils.py:9-14 cvtColor: anything that is not an RGB picture is converted by PIL
        return im.convert("RGB") if hasattr(im, "convert") and getattr(im, "mode", "RGB") != "RGB" else im
    if isinstance(images, (list, tuple)):
        images = np.stack([np.asarray(rgb(im), dtype=np.uint8) for im in images])
    images = rgb(images)
    if not torch.is_tensor(images):
        images = torch.from_numpy(np.array(images, dtype=np.uint8, order="C"))   # a writable copy (PIL buffers are read-only)
    if images.dtype != torch.uint8:
        raise TypeError(f"pictures must be uint8, got {images.dtype}")
    single = images.dim() == 3
    if single:
        images = images[None]
    if images.dim() != 4 or images.shape[-1] != 3:
        raise ValueError(f"expected [H,W,3] or [B,H,W,3] RGB pictures, got {tuple(images.shape)}")
    if not images.is_cuda:
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device())
        images = images.pin_memory().to(device, non_blocking=True) if torch.cuda.is_available() else images.to(device)
    return images.contiguous(), single


def _run(images, out_h, out_w, want_resized, want_input, mean, std):
    dev = images.device
    b, in_h, in_w = images.shape[0], images.shape[1], images.shape[2]
    resized = torch.empty((b, out_h, out_w, 3), dtype=torch.uint8, device=dev) if want_resized else None
    net_in = torch.empty((b, 3, out_h, out_w), dtype=torch.float32, device=dev) if want_input else None
    m3, s3 = (C.c_double * 3)(*[float(v) for v in mean]), (C.c_double * 3)(*[float(v) for v in std])
    h = get_handle(dev.index)
    with _lib.on_device(dev):
        h.check(h.lib.isb_preprocess(h.ptr, C.c_void_p(images.data_ptr()), b, in_h, in_w, out_h, out_w, m3, s3,
                                     C.c_void_p(resized.data_ptr()) if want_resized else None,
                                     C.c_void_p(net_in.data_ptr()) if want_input else None,
                                     C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)))
    return resized, net_in


def resize_image(image, size, device=None, as_pil=True):
    """utils/utils.py:19-22: `image.resize((w, h), Image.BICUBIC)`; `size` is (w, h) like the reference's.  Returns a PIL
    image (the reference's type) or, with `as_pil=False`, the uint8 [h,w,3] CUDA tensor."""
    w, h = int(size[0]), int(size[1])
    imgs, _ = _as_u8_batch(image, device)
    resized, _ = _run(imgs, h, w, True, False, MEAN, STD)
    if not as_pil:
        return resized[0]
    from PIL import Image
    return Image.fromarray(resized[0].cpu().numpy())


def preprocess_input(image_u8, device=None, mean=MEAN, std=STD):
    """utils/utils.py:32-36 + the wrapper's transpose / FloatTensor cast (yolact.py:145-148) for an already resized
    uint8 [H,W,3] picture: float32 [1,3,H,W] on the GPU, float32((float64(v) - mean) / std)."""
    imgs, _ = _as_u8_batch(image_u8, device)
    _, net_in = _run(imgs, imgs.shape[1], imgs.shape[2], False, True, mean, std)
    return net_in


def image_to_input(image, input_shape, device=None, mean=MEAN, std=STD, return_resized=False):
    """yolact.py:141-150 / mask_rcnn.py:144-152 in one call: PIL image, uint8 [H,W,3] array or a batch [B,H,W,3] of
    same-sized pictures -> float32 [B,3,H_in,W_in] CUDA tensor (B = 1 for a single picture), `input_shape` = (H_in, W_in)
    like the wrappers' attribute.  `return_resized` also returns the uint8 [B,H_in,W_in,3] resized pictures."""
    imgs, _ = _as_u8_batch(image, device)
    resized, net_in = _run(imgs, int(input_shape[0]), int(input_shape[1]), return_resized, True, mean, std)
    return (net_in, resized) if return_resized else net_in
