"""Builds libfdtd3d_b200.so (CUDA kernels + C ABI) in-tree with nvcc for sm_100a.

nvcc cross-compiles without a GPU, so this also runs in the CPU-only build container; the resulting
.so travels to the GPU box with the repository snapshot.  -fmad=false: exact mode must not contract
a*b+c into FMA (fast-mode code uses explicit fma intrinsics); host code gets -ffp-contract=off for
the same reason (coefficient tables).
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
BUILD = os.path.join(CSRC, "build")
LIB = os.path.join(HERE, "libfdtd3d_b200.so")
SOURCES = ["kernels3d_tma.cu", "kernels3d_fused.cu", "kernels2d.cu", "kernels2d_fused.cu", "aux_kernels.cu", "solver.cu", "scheme_init.cpp", "noalloc.cu"]
HEADERS = ["fdtd_device.cuh", "kernels3d_common.cuh", "launch.h", "nccl_dl.h", "tma_util.cuh", "kernels2d_common.cuh", os.path.join("..", "..", "include", "fdtd3d_b200.h")]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
    "-ccbin", "/usr/bin/g++", "-Xcompiler", "-fPIC,-ffp-contract=off,-O2", "-Xptxas", "-v",
]


def _newer(src, dst):
    return not os.path.exists(dst) or os.path.getmtime(src) > os.path.getmtime(dst)


def build(force=False, verbose=False, defines=(), lib=None, build_dir=None):
    """defines: extra -D flags (kernel tuning knobs); lib / build_dir: where a variant build goes."""
    global BUILD, LIB
    saved = (BUILD, LIB)
    if lib:
        LIB = lib
    if build_dir:
        BUILD = build_dir
    try:
        return _build(force, verbose, list(defines))
    finally:
        BUILD, LIB = saved


def _build(force, verbose, defines):
    os.makedirs(BUILD, exist_ok=True)
    hdr_paths = [os.path.normpath(os.path.join(CSRC, h)) for h in HEADERS] + [os.path.abspath(__file__)]
    jobs = []
    for src in SOURCES:
        sp = os.path.join(CSRC, src)
        obj = os.path.join(BUILD, src + ".o")
        if force or _newer(sp, obj) or any(_newer(h, obj) for h in hdr_paths):
            jobs.append((sp, obj))

    def compile_one(job):
        sp, obj = job
        cmd = [NVCC] + NVCC_FLAGS + ["-D" + d for d in defines] + ["-c", sp, "-o", obj]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        with open(obj + ".log", "w") as f:
            f.write(" ".join(cmd) + "\n" + r.stdout)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s" % (sp, r.stdout[-8000:]))
        return r.stdout

    if jobs:
        with ThreadPoolExecutor(max_workers=4) as ex:
            outs = list(ex.map(compile_one, jobs))
        if verbose:
            for o in outs:
                sys.stdout.write(o)
    objs = [os.path.join(BUILD, s + ".o") for s in SOURCES]
    if jobs or not os.path.exists(LIB):
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-ccbin", "/usr/bin/g++", "-Xcompiler", "-fPIC", "-ldl"]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stdout)
    _build_cli(LIB)
    return LIB


def _build_cli(lib):
    """fdtd3d_b200/bin/fdtd3d: the solver binary with the reference's command line (host C++, links the C ABI)."""
    host = os.path.join(HERE, "host")
    srcs = [os.path.join(host, f) for f in ("main.cpp", "Settings.cpp", "Scheme.cpp")]
    hdrs = [os.path.join(host, f) for f in ("Settings.h", "Scheme.h")]
    exe = os.path.join(HERE, "bin", "fdtd3d")
    if os.path.basename(lib) != "libfdtd3d_b200.so":
        return
    if os.path.exists(exe) and not any(_newer(f, exe) for f in srcs + hdrs + [lib]):
        return
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    cmd = ["/usr/bin/g++", "-std=c++14", "-O2", "-ffp-contract=off", "-Wall"] + srcs + \
          ["-L" + HERE, "-lfdtd3d_b200", "-Wl,-rpath,$ORIGIN/..", "-pthread", "-o", exe]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("building the fdtd3d binary failed:\n" + r.stdout)


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
