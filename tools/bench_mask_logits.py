"""Dense mask-logit product on the tensor cores (isb_mask_logits: tcgen05.mma kind::tf32, 3-way exact split, TMEM
accumulators) against cuBLAS sgemm, for 64 pictures x 100 detections at the benchmark shape (138 x 138 x 32 prototypes).
The fused path does NOT use it (K3a computes only the prototype pixels under each detection's box); this tool gives the
number that decision rests on: the dense product writes 19044 x 100 floats per picture whatever computes it."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from isb.yolact import BBoxUtility  # noqa: E402


def timeit(fn, warm=5, iters=20):
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
    g = torch.Generator().manual_seed(1)
    b, m = 64, 100
    protos = [torch.relu(torch.randn(138, 138, 32, generator=g)).cuda() for _ in range(b)]
    coefs = [torch.tanh(torch.randn(m, 32, generator=g)).cuda() for _ in range(b)]
    bu = BBoxUtility()
    torch.backends.cuda.matmul.allow_tf32 = False
    us_tc = timeit(lambda: [bu.mask_logits(p, c) for p, c in zip(protos, coefs)])
    us_blas = timeit(lambda: [p.reshape(-1, 32) @ c.t() for p, c in zip(protos, coefs)])
    pb, cb = torch.stack(protos).reshape(b, -1, 32), torch.stack(coefs).transpose(1, 2).contiguous()
    us_bmm = timeit(lambda: torch.bmm(pb, cb))
    pstack, cstack = torch.stack(protos), torch.stack(coefs)
    us_tc_batch = timeit(lambda: bu.mask_logits_batch(pstack, cstack))
    us_tc_batch_sig = timeit(lambda: bu.mask_logits_batch(pstack, cstack, sigmoid=True))
    err = float((bu.mask_logits_batch(pstack, cstack).double() - torch.bmm(pb.double(), cb.double()).reshape(b, 138, 138, m)).abs().max())
    out_bytes = b * 138 * 138 * m * 4
    in_bytes = b * (138 * 138 * 32 + m * 32) * 4
    print(json.dumps({"pictures": b, "detections": m, "output_MB": out_bytes / 1e6, "input_MB": in_bytes / 1e6,
                      "isb_mask_logits_tcgen05_us": us_tc, "cublas_sgemm_loop_us": us_blas, "cublas_bmm_us": us_bmm,
                      "isb_mask_logits_batch_tcgen05_us": us_tc_batch, "isb_mask_logits_batch_sigmoid_us": us_tc_batch_sig,
                      "batch_max_abs_err_vs_fp64": err,
                      "batch_out_GBs": out_bytes / us_tc_batch / 1e3,
                      "hbm_floor_us_at_6538GBs": (out_bytes + in_bytes) / 6538.3e3,
                      "note": "K3a (sigmoid maps on the box footprints only) takes ~38 us for the same 64 pictures inside the step"},
                     indent=1))


if __name__ == "__main__":
    main()
