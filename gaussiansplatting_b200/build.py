"""In-tree build of libgs_b200.so (CUDA kernels + C ABI) for sm_100a.

Run as `python -m gaussiansplatting_b200.build` or through __graft_entry__.build().  nvcc cross-compiles
without a GPU; the .so travels to the GPU box with the repo snapshot.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "_obj")
LIB = os.path.join(HERE, "libgs_b200.so")
SOURCES = ["preprocess.cu", "pairs.cu", "binning.cu", "radix_sort.cu", "blend_fwd.cu", "blend_bwd.cu", "layout.cu", "loss.cu", "viewer.cu", "adam.cu",
           "densify.cu", "peer_reduce.cu", "api.cu"]
HEADERS = ["gs_common.cuh", "gs_kernels.h", os.path.join("..", "..", "include", "gs_abi.h")]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
         "-ccbin", "/usr/bin/g++", "--expt-relaxed-constexpr"] + os.environ.get("GS_NVCC_EXTRA", "").split()
LIB = os.environ.get("GS_LIB_OUT", LIB)          # experiments: build a variant library next to the default one


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    hdrs = [os.path.join(CSRC, h) for h in HEADERS]
    jobs = []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(OBJ, s.replace(".cu", ".o"))
        if force or _stale(obj, [src] + hdrs):
            cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
            jobs.append(cmd)

    def run(cmd):
        p = subprocess.run(cmd, capture_output=True, text=True)
        return cmd, p.returncode, p.stdout + p.stderr

    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for cmd, rc, out in ex.map(run, jobs):
            if verbose or rc != 0:
                sys.stderr.write(" ".join(cmd[-3:]) + "\n" + out + "\n")
            if rc != 0:
                raise RuntimeError("nvcc failed: " + " ".join(cmd))
    objs = [os.path.join(OBJ, s.replace(".cu", ".o")) for s in SOURCES]
    if force or jobs or _stale(LIB, objs):
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-ccbin", "/usr/bin/g++", "-gencode", "arch=compute_100a,code=sm_100a"]
        p = subprocess.run(cmd, capture_output=True, text=True)
        if p.returncode != 0:
            raise RuntimeError("link failed:\n" + p.stdout + p.stderr)
    build_host(force or bool(jobs))
    return LIB


HOST_LIB = os.path.join(HERE, "libgs_host.so")
HOST_SRCS = [os.path.join(HERE, "host", f) for f in ("gs_host_capi.cpp", "engine.hpp", "gs_host.hpp", "metal_shim.hpp")]


def build_host(force=False):
    """libgs_host.so: the C++ host layer (the reference's class interfaces over the C ABI) + its C entry points"""
    if not (force or _stale(HOST_LIB, HOST_SRCS + [LIB, os.path.join(HERE, "..", "include", "gs_abi.h")])):
        return HOST_LIB
    cmd = ["/usr/bin/g++", "-std=c++17", "-O2", "-fPIC", "-Wall", "-Wextra", "-shared", "-o", HOST_LIB, HOST_SRCS[0],
           "-L" + HERE, "-l:" + os.path.basename(LIB), "-Wl,-rpath,$ORIGIN"]
    p = subprocess.run(cmd, capture_output=True, text=True)
    if p.returncode != 0:
        raise RuntimeError("host layer build failed:\n" + p.stdout + p.stderr)
    return HOST_LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
