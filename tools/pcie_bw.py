"""PCIe ceiling for the e2e path: pinned D2H, H2D and both at once (two streams)."""
import json, time, torch
n = 1 << 30
d = torch.empty(n, dtype=torch.uint8, device="cuda"); hdst = torch.empty(n, dtype=torch.uint8).pin_memory()
d2 = torch.empty(n, dtype=torch.uint8, device="cuda"); hsrc = torch.empty(n, dtype=torch.uint8).pin_memory()
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(do_d2h, do_h2d, reps=5):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps):
        if do_d2h:
            with torch.cuda.stream(s1): hdst.copy_(d, non_blocking=True)
        if do_h2d:
            with torch.cuda.stream(s2): d2.copy_(hsrc, non_blocking=True)
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps
run(True, True, 2)
r = {"d2h_GBs": n / run(True, False) / 1e9, "h2d_GBs": n / run(False, True) / 1e9}
t = run(True, True); r["both_each_GBs"] = n / t / 1e9
print(json.dumps(r))
