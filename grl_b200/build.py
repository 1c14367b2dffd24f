"""In-tree build of libgrlcu.so (hand-written sm_100a CUDA behind the C ABI of include/grlcu.h)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(CSRC, "libgrlcu.so")
SOURCES = ["grlcu.cu"]
HEADERS = ["grlcu_device.cuh", "grlcu_agent.cuh", "grlcu_cached.cuh", "grlcu_kernels.cuh", "grlcu_ac.cuh", "grlcu_layout.h", os.path.join("..", "..", "include", "grlcu.h")]

# -fmad=false: the reference binary has no FMA contraction (CMakeLists.txt:19), see DESIGN.md.
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-fmad=false", "-lineinfo",
              "-std=c++17", "-Xcompiler", "-fPIC", "-shared"]


def nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.exists(cand) or cand == "nvcc"):
            return cand
    return "nvcc"


def stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force=False, verbose=False):
    """Compile libgrlcu.so for sm_100a if it is missing or older than its sources."""
    if not force and not stale():
        return LIB
    cmd = [nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + SOURCES
    subprocess.run(cmd, cwd=CSRC, check=True)
    return LIB


if __name__ == "__main__":
    print(build(force=True, verbose=True))
