"""Build libraytracing_b200.so in-tree (nvcc, sm_100a).  `python -m raytracing_cuda_b200.build`"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def build(force=False, verbose=False):
    csrc = os.path.join(HERE, "csrc")
    if force:
        subprocess.check_call(["make", "-C", csrc, "clean"], stdout=subprocess.DEVNULL)
    jobs = str(min(8, os.cpu_count() or 1))
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-C", csrc, "-j", jobs], stdout=out)
    return os.path.join(HERE, "libraytracing_b200.so")


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
