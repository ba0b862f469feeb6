"""Builds libb2train.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

Usage: python build.py [--force]   or   from build import build; build()
The shared object lands next to the Python host package (b200_train/libb2train.so) so that it
travels with the tree and is visible to the driver's loaded-library check.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
OUT = os.path.join(HERE, "b200_train", "libb2train.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr", "-I", CSRC, "-I", os.path.join(os.path.dirname(HERE), "include"),
]


def _newer(src, dst, deps):
    if not os.path.exists(dst):
        return True
    t = os.path.getmtime(dst)
    return any(os.path.getmtime(d) > t for d in [src] + deps)


def build(force=False, verbose=False):
    """B2_BUILD_PROBES=1 in the environment adds -DB2_BUILD_PROBES: the measurement aids (B2_DEBUG_STORE / B2_DEBUG_NOEPI
    branches, b2_debug_peer_copy) are compiled in. The default build, the one that ships, has none of them."""
    global OBJ, OUT
    flags = list(FLAGS)
    if os.environ.get("B2_BUILD_PROBES") == "1":
        flags += ["-DB2_BUILD_PROBES"]
        OBJ = os.path.join(HERE, "build", "probes")
        OUT = os.path.join(HERE, "build", "ab", "libb2train_probes.so")
        os.makedirs(os.path.dirname(OUT), exist_ok=True)
    os.makedirs(OBJ, exist_ok=True)
    srcs = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hdrs += [os.path.join(os.path.dirname(HERE), "include", "b2train.h")]
    hdrs = [h for h in hdrs if os.path.exists(h)]
    jobs = []
    objs = []
    for f in srcs:
        src = os.path.join(CSRC, f)
        obj = os.path.join(OBJ, f[:-3] + ".o")
        objs.append(obj)
        if force or _newer(src, obj, hdrs):
            cmd = [NVCC] + flags + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
            jobs.append(cmd)

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        return cmd, r

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for cmd, r in ex.map(run, jobs):
            if verbose or r.returncode != 0:
                sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
            if r.returncode != 0:
                raise RuntimeError("nvcc failed for " + cmd[-3])
    if jobs or force or not os.path.exists(OUT):
        cmd = [NVCC, "-shared", "-o", OUT] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
