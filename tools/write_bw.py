"""Measure write-only / read-only / copy HBM bandwidth with torch ops (context for the roofline)."""
import torch, json
n = 2_200_000_000
x = torch.empty(n, dtype=torch.uint8, device="cuda")
y = torch.empty(n, dtype=torch.uint8, device="cuda")
def timeit(f, reps=10):
    for _ in range(3): f()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best
res = {}
t = timeit(lambda: x.zero_()); res["memset_write_only_GBs"] = n / t / 1e6
t = timeit(lambda: x.fill_(1)); res["fill_write_only_GBs"] = n / t / 1e6
xi = x.view(torch.int32)
t = timeit(lambda: xi.sum()); res["sum_read_only_GBs"] = n / t / 1e6
t = timeit(lambda: y.copy_(x)); res["copy_read_plus_write_GBs"] = 2 * n / t / 1e6
print(json.dumps(res))
