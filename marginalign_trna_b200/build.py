"""Builds libmargin_b200.so in-tree (nvcc cross-compiles sm_100a without a GPU)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_PATH = os.path.join(HERE, "libmargin_b200.so")
SOURCES = ["kernels.cu", "plan_device.cu", "realign.cu", "columns.cu", "plan.cpp", "model.cpp"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    # the DP must not contract a*b+c into FMA (the reference is built without FMA); kernels also use explicit
    # __dadd_rn/__dmul_rn, this flag covers everything else in the translation unit
    "-fmad=false",
    "-Xcompiler", "-fPIC,-O2,-ffp-contract=off,-Wall,-pthread",
    "-shared",
]


def needs_build():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [
        os.path.join(os.path.dirname(HERE), "include", "margin_b200.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force=False, verbose=False):
    if not force and not needs_build():
        return LIB_PATH
    cmd = ["nvcc"] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH] + [
        os.path.join(CSRC, s) for s in SOURCES]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n%s\n%s" % (res.stdout, res.stderr))
    if verbose:
        print(res.stderr)
    return LIB_PATH


if __name__ == "__main__":
    import sys
    build_library(force=True, verbose="-v" in sys.argv)
    print(LIB_PATH)
