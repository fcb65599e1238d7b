"""SASS-level view of an .ncu-rep (read in the build container): executed-instruction and stall-sample totals per opcode and
per 32-instruction block, and the hottest straight-line runs.  usage: python tools/ncu_sass.py x.ncu-rep [kernel-substring]"""
import collections
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    want = sys.argv[2] if len(sys.argv) > 2 else ""
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "sass", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    kern, head, data = "", None, []
    blocks = []
    for r in rows:
        if r and r[0] in ("Function Name", "Kernel Name"):
            if data:
                blocks.append((kern, head, data))
            kern, head, data = r[1], None, []
        elif r and r[0] == "Address":
            head = r
        elif head and len(r) == len(head) and r[0].startswith("0x"):
            data.append(r)
    if data:
        blocks.append((kern, head, data))
    for kern, head, data in blocks:
        if want not in kern:
            continue
        ci, si, sm = head.index("Instructions Executed"), head.index("Source"), head.index("# Samples")
        f = lambda v: float(v.replace(",", "")) if v not in ("", "-") else 0.0
        d = [(r[si].strip(), f(r[ci]), f(r[sm])) for r in data]
        tot, ts = sum(x[1] for x in d), sum(x[2] for x in d)
        print(f"== {kern[:120]}\n   warp instructions {tot:.0f}, stall samples {ts:.0f}, SASS lines {len(d)}")
        ops = collections.Counter()
        for ins, n, s in d:
            parts = ins.split()
            op = parts[1] if parts[0].startswith("@") else parts[0]
            ops[op.split(".")[0]] += n
        print("   opcodes: " + ", ".join(f"{k} {100 * v / tot:.1f}%" for k, v in ops.most_common(14)))
        print("   blocks of 32 SASS lines with >= 2 % of the samples or instructions:  first line | samples | instructions")
        for i in range(0, len(d), 32):
            s = sum(x[2] for x in d[i:i + 32])
            n = sum(x[1] for x in d[i:i + 32])
            if s >= 0.02 * ts or n >= 0.02 * tot:
                print(f"   {i:6d} | {s:6.0f} ({100 * s / max(ts, 1):4.1f}%) | {n:10.0f} ({100 * n / tot:4.1f}%) | {d[i][0][:70]}")


if __name__ == "__main__":
    main()
