"""In-tree build of the CUDA engine (sm_100a only) into cnf2freq_b200/lib/libcnf2_engine.so."""
import os
import subprocess

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
LIB = os.path.join(PKG, "lib", "libcnf2_engine.so")
SOURCES = [os.path.join(PKG, "csrc", "cnf2_engine.cu")]
HEADERS = [os.path.join(PKG, "csrc", h) for h in ("cnf2_device.cuh", "cnf2_estep_kernel.cuh", "cnf2_tables.cuh", "cnf2_warp_kernel.cuh",
                                                   "cnf2_mstep_kernel.cuh", "cnf2_preprocess.cuh")] + [
    os.path.join(ROOT, "include", "cnf2_engine.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-shared", "-Xcompiler",
              "-fPIC"]


def stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.exists(f) and os.path.getmtime(f) > t for f in SOURCES + HEADERS)


def build_engine(force=False, verbose=False):
    """Compiles the engine when the library is missing or older than its sources; returns the path."""
    if force or stale():
        os.makedirs(os.path.dirname(LIB), exist_ok=True)
        nvcc = os.environ.get("NVCC", "nvcc")
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + SOURCES
        subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build_engine(force=True, verbose=True))
