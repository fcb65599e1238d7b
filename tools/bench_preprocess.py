"""Device timing of the pre-processing row: 64 pictures 1080 x 1920 -> network input 544 x 544 (the wrappers' setting on
the repo's test pictures) and -> 550 x 550 (the benchmark size)."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from isb import preprocess as pre  # noqa: E402


def timeit(fn, warm=5, iters=30):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


def main():
    res = {}
    for b, (h, w), shape in ((64, (1080, 1920), (544, 544)), (64, (1080, 1920), (550, 550)), (1, (1080, 1920), (544, 544))):
        imgs = torch.randint(0, 256, (b, h, w, 3), dtype=torch.uint8, device="cuda")
        us = timeit(lambda: pre.image_to_input(imgs, shape))
        algo = b * (h * w * 3 + shape[0] * shape[1] * 12)
        res[f"{b}x{h}x{w}_to_{shape[0]}x{shape[1]}"] = {"us": us, "us_per_picture": us / b, "pictures_per_s": b / us * 1e6,
                                                       "algorithmic_GBs": algo / us / 1e3}
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
