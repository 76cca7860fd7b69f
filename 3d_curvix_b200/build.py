"""In-tree build of the CUDA extension (libcurvix_b200.so) for sm_100a, and of the C++ host drop-in.

nvcc cross-compiles without a GPU.  Flags that matter:
  -gencode arch=compute_100a,code=sm_100a   B200 only, no other targets
  -fmad=false                               un-fused FP64 so accept/reject decisions match an un-fused CPU build
  -lineinfo                                 ncu source page
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libcurvix_b200.so")
HOST_BIN = os.path.join(HERE, "host", "curvix_main")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-fmad=false",
    "-Xcompiler", "-fPIC,-O2,-ffp-contract=off,-Wall", "--expt-relaxed-constexpr",
]


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _cxx():
    # $CXX in this image points at a g++ without OpenMP specs; /usr/bin/g++ is complete
    return "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"


def build_cuda(force=False, verbose=False):
    srcs = [os.path.join(CSRC, f) for f in ("cvx_kernels.cu", "cvx_api.cu", "cvx_comm.cu", "cvx_graph.cu")]
    deps = srcs + [os.path.join(CSRC, f) for f in ("cvx_kernels.cuh", "cvx_math.cuh", "cvx_graph_math.cuh", "cvx_internal.h")] + [os.path.join(ROOT, "include", "curvix_b200.h")]
    out = os.environ.get("CVX_LIB_OUT", LIB)            # experiments: alternative output name + extra -D flags
    if not force and not _stale(out, deps):
        return out
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    extra = os.environ.get("CVX_NVCC_EXTRA", "").split()
    tmp = out + f".tmp{os.getpid()}"                    # written aside and renamed: another process never dlopens a half-written library
    cmd = [nvcc] + NVCC_FLAGS + extra + ["-ccbin", _cxx(), "-shared", "-o", tmp] + srcs + ["-ldl"]   # NCCL is dlopen-ed (cvx_comm.cu)
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd))
    try:
        subprocess.run(cmd, check=True)
        os.replace(tmp, out)
    finally:
        if os.path.exists(tmp):
            os.remove(tmp)
    return out


def build_host(force=False):
    """C++ host drop-in (EdgeSketch_Core shim + curvix_main) linked against libcurvix_b200.so."""
    hdir = os.path.join(HERE, "host")
    srcs = [os.path.join(hdir, f) for f in ("EdgeSketch_Core.cpp", "edge_mapping.cpp", "main.cpp")]
    if not all(os.path.exists(s) for s in srcs):
        return None
    deps = srcs + [os.path.join(hdir, f) for f in os.listdir(hdir) if f.endswith(".hpp")]
    if not force and not _stale(HOST_BIN, deps + [LIB]):
        return HOST_BIN
    cmd = [_cxx(), "-O2", "-std=c++17", "-ffp-contract=off", "-Wall", "-I", os.path.join(ROOT, "include"), "-o", HOST_BIN] + srcs + \
          ["-L", HERE, "-lcurvix_b200", "-Wl,-rpath," + HERE, "-Wl,-rpath,/usr/local/cuda/lib64"]
    subprocess.run(cmd, check=True)
    return HOST_BIN


if __name__ == "__main__":
    build_cuda(force="--force" in sys.argv, verbose="-v" in sys.argv)
    build_host(force="--force" in sys.argv)
    print("built", LIB)
