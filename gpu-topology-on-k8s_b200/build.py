"""Builds libgputopo.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "libgputopo.so")
SOURCES = [os.path.join(HERE, "csrc", f) for f in ("api.cu", "host.cc", "extender.cc", "gaia_tree.cc", "wire.cc")]
DEPS = SOURCES + [os.path.join(HERE, "csrc", "kernels.cuh"), os.path.join(HERE, "csrc", "gaia_tree.h"), os.path.join(HERE, "csrc", "wire.h"),
                  os.path.join(os.path.dirname(HERE), "include", "gputopo.h")]
NVCC_FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "-Xcompiler", "-fPIC", "-shared"]


def stale():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    return any(os.path.getmtime(d) > t for d in DEPS if os.path.exists(d))


def build(force=False, verbose=False):
    if not force and not stale():
        return SO
    nvcc = os.environ.get("NVCC", "nvcc")
    extra = os.environ.get("GT_NVCC_EXTRA", "").split()
    cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-o", SO] + SOURCES + ["-ldl"]
    subprocess.run(cmd, check=True)
    return SO


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
