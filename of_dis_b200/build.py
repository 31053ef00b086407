"""Builds of_dis_b200/libdisb200.so (CUDA kernels + C-ABI) and the run_* CLIs in-tree with nvcc.

sm_100a only; -fmad=false keeps the fp32 operation order of the reference (no FMA contraction),
which is what makes the variational refinement bit-comparable with the CPU reference;
-lineinfo so ncu source pages map to these files.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libdisb200.so")
KERNEL_SOURCES = ["pyramid.cu", "patch.cu", "densify.cu", "varref.cu", "upsample.cu", "engine.cu", "tiled.cu", "flowio.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
    "-Xcompiler", "-fPIC", "-Xptxas", "-v",
]
# Files whose results are compared with a tolerance (the inverse search differs from the CPU code by the
# order of its reductions anyway) may contract a*b+c into FMA; everything bit-compared may not.
FMA_OK = set()   # measured: FMA in patch.cu gains 2 % of that kernel; not worth the drift from the CPU code
CLI_VARIANTS = {"run_OF_INT": (1, 1), "run_OF_RGB": (1, 3), "run_DE_INT": (2, 1), "run_DE_RGB": (2, 3),
                # SELECTCHANNEL == 2 (gradient-magnitude input): in the reference's sources, without a CMake target
                "run_OF_GRAD": (1, 2), "run_DE_GRAD": (2, 2)}


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.sep not in cand or os.path.exists(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


# "tolerance" variant (libdisb200_tol.so): FMA contraction and approximate division / square root allowed.  NOT the
# parity build: reported next to it with its own EPE against the CPU reference ("what bit-exactness costs").
TOL_LIB = os.path.join(HERE, "libdisb200_tol.so")
TOL_FLAGS = ["-prec-div=false", "-prec-sqrt=false"]


def build(force: bool = False, verbose: bool = False, variant: str = "parity") -> str:
    nvcc = _nvcc()
    tol = variant == "tol"
    tag = os.environ.get("DISB200_BUILD_TAG", "")          # A/B builds: libdisb200_<tag>.so with DISB200_NVCC_EXTRA flags
    LIB = TOL_LIB if tol else (os.path.join(HERE, "libdisb200_%s.so" % tag) if tag else globals()["LIB"])
    objdir = os.path.join(HERE, "build_tol" if tol else ("build_" + tag if tag else "build"))
    os.makedirs(objdir, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(HERE, "..", "include", "disb200.h"))
    logs = {}

    def compile_one(src):
        s = os.path.join(CSRC, src)
        o = os.path.join(objdir, src.replace(".cu", ".o"))
        if force or _stale(o, [s] + headers):
            flags = [f for f in NVCC_FLAGS if not ((tol or src in FMA_OK) and f == "-fmad=false")] + (TOL_FLAGS if tol else []) + os.environ.get("DISB200_NVCC_EXTRA", "").split()
            r = subprocess.run([nvcc] + flags + ["-c", s, "-o", o], capture_output=True, text=True)
            logs[src] = r.stderr
            if r.returncode != 0:
                raise RuntimeError("nvcc failed on %s:\n%s" % (src, r.stderr))
        return o

    with ThreadPoolExecutor(max_workers=8) as ex:
        objs = list(ex.map(compile_one, KERNEL_SOURCES))
    if force or _stale(LIB, objs):
        r = subprocess.run([nvcc, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lz"],
                           capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stderr)
    if tol or tag:
        return LIB
    # the four CLIs of the reference (CMakeLists.txt:35-61), thin mains over the library
    cli_src = os.path.join(CSRC, "run_dense.cpp")
    if os.path.exists(cli_src):
        bindir = os.path.join(HERE, "bin")
        os.makedirs(bindir, exist_ok=True)
        for name, (mode, ch) in CLI_VARIANTS.items():
            exe = os.path.join(bindir, name)
            if force or _stale(exe, [cli_src, LIB] + headers):
                r = subprocess.run(["g++", "-O2", "-std=c++17", "-DSELECTMODE=%d" % mode, "-DSELECTCHANNEL=%d" % ch,
                                    cli_src, "-o", exe, "-L" + HERE, "-ldisb200", "-Wl,-rpath,$ORIGIN/.."],
                                   capture_output=True, text=True)
                if r.returncode != 0:
                    raise RuntimeError("CLI build failed:\n" + r.stderr)
    batch_src = os.path.join(CSRC, "run_batch.cpp")
    if os.path.exists(batch_src):
        exe = os.path.join(HERE, "bin", "disb200_batch")
        os.makedirs(os.path.dirname(exe), exist_ok=True)
        if force or _stale(exe, [batch_src, LIB] + headers):
            r = subprocess.run(["g++", "-O2", "-std=c++17", "-pthread", batch_src, "-o", exe, "-L" + HERE, "-ldisb200",
                                "-Wl,-rpath,$ORIGIN/.."], capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError("disb200_batch build failed:\n" + r.stderr)
    if verbose:
        for k, v in logs.items():
            sys.stderr.write("== %s ==\n%s\n" % (k, v))
    with open(os.path.join(objdir, "ptxas.log"), "a") as f:
        for k, v in logs.items():
            f.write("== %s ==\n%s\n" % (k, v))
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, variant="tol" if "--tol" in sys.argv else "parity"))
