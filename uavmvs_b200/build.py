"""In-tree build of libuavmvs_b200.so (sm_100a only).

    python -m uavmvs_b200.build [--force]

nvcc cross-compiles without a GPU.  --fmad=false plus the default IEEE div/sqrt and
no FTZ are part of the numerical contract (bit-exact visibility and counts).
"""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "libuavmvs_b200.so")
NVCC_FLAGS = [
    "-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "--fmad=false", "-shared", "-Xcompiler", "-fPIC", "-ccbin", "g++",
]


def sources():
    return sorted(glob.glob(os.path.join(HERE, "csrc", "*.cu")))


def stale():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = sources() + glob.glob(os.path.join(HERE, "csrc", "*.cuh")) + \
        glob.glob(os.path.join(HERE, "..", "include", "*.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not stale():
        return SO
    nvcc = os.environ.get("NVCC", "nvcc")
    extra = os.environ.get("UAV_NVCC_EXTRA", "").split()
    out = os.environ.get("UAV_SO", SO)
    cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", out] + sources()
    subprocess.check_call(cmd)
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
