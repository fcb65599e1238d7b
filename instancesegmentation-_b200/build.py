"""Compile libisb.so in-tree with nvcc (sm_100a).  Used by __graft_entry__.build()."""
import os
import subprocess

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libisb.so")


def build(verbose=False):
    cmd = ["make", "-j", str(os.cpu_count() or 4), "-C", CSRC, f"OUT={LIB}"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout)
        print(res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc build of libisb.so failed")
    return LIB


if __name__ == "__main__":
    print(build(verbose=True))
