"""Generate tests/golden/preprocess.npz from Pillow itself and the reference's own helpers.

Run in the build container only (needs /root/reference and Pillow):

    PYTHONDONTWRITEBYTECODE=1 python -m oracle.gen_golden_preprocess

Small cases store the input image and the output of `resize_image` (utils/utils.py:19-22, i.e. Pillow's
`Image.resize(..., Image.BICUBIC)`); the full-size case (1080 x 1920 -> 544 x 544, the wrappers' configuration on the
repo's test images) stores the seed of the synthetic picture, a SHA-1 of the resized image and of the final float32
network input, and a sub-sampled copy of both.
"""
from __future__ import annotations

import hashlib
import os
import sys

import numpy as np
import PIL
import torch
from PIL import Image

sys.dont_write_bytecode = True
REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")

SMALL = [((97, 131), (56, 40)), ((20, 30), (48, 64)), ((50, 50), (50, 37)), ((33, 77), (91, 33)), ((7, 5), (3, 2)),
         ((3, 3), (40, 24)), ((64, 64), (64, 64)), ((120, 200), (61, 35))]


def reference_input(img_u8, input_shape):
    """yolact.py:137-148 with the reference's own helpers."""
    if REF not in sys.path:
        sys.path.insert(0, REF)
    from inst_model.yolact.utils.utils import preprocess_input, resize_image
    resized = resize_image(Image.fromarray(img_u8), (input_shape[1], input_shape[0]))
    data = np.expand_dims(np.transpose(preprocess_input(np.array(resized, np.float32)), (2, 0, 1)), 0)
    return np.array(resized), torch.from_numpy(data).type(torch.FloatTensor).numpy()


def sha(a):
    return hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    from . import synth
    kw = dict(pillow_version=PIL.__version__, torch_version=torch.__version__)
    rng = np.random.RandomState(20260501)
    for i, ((h, w), (ow, oh)) in enumerate(SMALL):
        img = rng.randint(0, 256, (h, w, 3), dtype=np.uint8) if i % 2 == 0 else synth.picture(100 + i, h, w)
        resized, net_in = reference_input(img, (oh, ow))
        kw[f"s{i}_in"] = img
        kw[f"s{i}_resized"] = resized
        kw[f"s{i}_input"] = net_in
    kw["num_small"] = len(SMALL)
    for tag, seed, (h, w), shape in (("full544", 7001, (1080, 1920), (544, 544)), ("full550", 7002, (1080, 1920), (550, 550)),
                                     ("up", 7003, (300, 400), (544, 544))):
        img = synth.picture(seed, h, w)
        resized, net_in = reference_input(img, shape)
        kw[f"{tag}_seed"] = seed
        kw[f"{tag}_hw"] = np.array([h, w])
        kw[f"{tag}_shape"] = np.array(shape)
        kw[f"{tag}_in_sha"] = sha(img)
        kw[f"{tag}_resized_sha"] = sha(resized)
        kw[f"{tag}_input_sha"] = sha(net_in)
        kw[f"{tag}_resized_sub"] = resized[::9, ::7]
        kw[f"{tag}_input_sub"] = net_in[:, :, ::9, ::7]
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, "preprocess.npz")
    np.savez_compressed(path, **kw)
    print(f"preprocess: {os.path.getsize(path) / 1024:.1f} KiB")


if __name__ == "__main__":
    main()
