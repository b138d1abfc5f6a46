"""Build libspchol.so (sm_100a) in-tree with nvcc.  Usage: python -m sparsecholesky_b200.build"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libspchol.so")
SOURCES = ["symbolic.cpp", "ordering.cpp", "plan.cpp", "tree_plan.cpp", "capi_host.cpp", "engine.cu"]
HEADERS = ["symbolic.h", "ordering.h", "plan.h", "capi_internal.h", "kernels.cuh", "gemm_tma.cuh", "ingest.cuh", "tree_plan.h", "tree_kernel.cuh", "residual.cuh", "solve.cuh", "../../include/spchol.h"]


def needs_build() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return OUT
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc, "-std=c++17", "-O3", "-lineinfo", "-gencode", "arch=compute_100a,code=sm_100a",
           "-Xcompiler", "-fPIC,-O2,-Wall", "-shared", "--use_fast_math=false" if False else "-prec-div=true",
           "-prec-sqrt=true", "-fmad=true", "-x", "cu"]
    cmd += os.environ.get("SPCHOL_NVCC_DEFS", "").split()      # e.g. -DPK_OPT=5: staged kernel variants (csrc/path_kernel.cuh)
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += [os.path.join(CSRC, s) for s in SOURCES]
    cmd += ["-o", OUT, "-lcudart", "-ldl"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed building libspchol.so")
    if verbose:
        sys.stderr.write(r.stderr)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
