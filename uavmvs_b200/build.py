"""In-tree build of libuavmvs_b200.so (sm_100a only).

    python -m uavmvs_b200.build [--force] [-v]

nvcc cross-compiles without a GPU.  --fmad=false plus the default IEEE div/sqrt and
no FTZ are part of the numerical contract (bit-exact visibility and counts).
Every csrc/*.cu becomes one object under uavmvs_b200/build/ (compiled in parallel, only
when it or a header changed), then the objects are linked into the shared library.
"""
import glob
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "libuavmvs_b200.so")
OBJ_DIR = os.path.join(HERE, "build")
NVCC_FLAGS = [
    "-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "--fmad=false", "-Xcompiler", "-fPIC", "-ccbin", "g++",
]


def sources():
    return sorted(glob.glob(os.path.join(HERE, "csrc", "*.cu")))


def headers():
    return sorted(glob.glob(os.path.join(HERE, "csrc", "*.cuh")) + glob.glob(os.path.join(HERE, "..", "include", "*.h")))


def _flags_tag(extra):
    return hashlib.sha1(" ".join(NVCC_FLAGS + extra).encode()).hexdigest()[:10]


def _obj(src, tag):
    return os.path.join(OBJ_DIR, "%s.%s.o" % (os.path.splitext(os.path.basename(src))[0], tag))


def stale():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    return any(os.path.getmtime(d) > t for d in sources() + headers())


def build(force=False, verbose=False):
    if not force and not stale():
        return SO
    nvcc = os.environ.get("NVCC", "nvcc")
    extra = os.environ.get("UAV_NVCC_EXTRA", "").split()
    out = os.environ.get("UAV_SO", SO)
    tag = _flags_tag(extra)
    os.makedirs(OBJ_DIR, exist_ok=True)
    hdr_t = max(os.path.getmtime(h) for h in headers())
    jobs = []
    for src in sources():
        o = _obj(src, tag)
        if force or not os.path.exists(o) or os.path.getmtime(o) < max(os.path.getmtime(src), hdr_t):
            jobs.append([nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", o, src])
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as pool:
        for rc, cmd in zip(pool.map(lambda c: subprocess.call(c), jobs), jobs):
            if rc != 0:
                raise subprocess.CalledProcessError(rc, cmd)
    subprocess.check_call([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC", "-ccbin", "g++",
                           "-o", out] + [_obj(s, tag) for s in sources()] + ["-ldl"])
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
