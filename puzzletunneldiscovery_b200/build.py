"""Builds libosrgpu.so (sm_100a) in-tree with nvcc.  No JIT cache: the .so travels with the repo."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libosrgpu.so")
SOURCES = ["engine.cu", "geometry.cc"]
DEPS = SOURCES + ["kernels.cuh", "kernels_dfs.cuh", "kernels_seg.cuh", "kernels_surface.cuh", "kernels_knn.cuh",
                  "kernels_raster.cuh", "kernels_clearance.cuh", "kernels_keyconf.cuh", "kernels_contacts.cuh", "qnode.h", "sat17.h", "se3.h", "geometry.h",
                  os.path.join("..", "..", "include", "osrgpu.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-fmad=false",            # fp64 code must round like the reference (no FMA contraction)
    "-Xcompiler", "-fPIC,-O3,-ffp-contract=off,-Wall",
    "-shared", "-Xptxas", "-v",
]


def stale():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(os.path.join(CSRC, d)) > t for d in DEPS)


def build_variant(tag, defines):
    """Development builds for A/B runs (scripts/k1_variants.sh): libosrgpu_<tag>.so with extra -D flags, selected at
    run time with OSR_LIB=<path>.  Not part of the product build."""
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    out = os.path.join(HERE, "libosrgpu_%s.so" % tag)
    cmd = [nvcc] + NVCC_FLAGS + ["-D" + d for d in defines] + ["-ccbin", "g++", "-o", out] + \
          [os.path.join(CSRC, s) for s in SOURCES]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("nvcc failed building " + out)
    return out


def build(force=False, verbose=False):
    if not force and not stale():
        return OUT
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + ["-ccbin", "g++", "-o", OUT] + [os.path.join(CSRC, s) for s in SOURCES]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    with open(os.path.join(HERE, "build.log"), "w") as f:
        f.write(" ".join(cmd) + "\n" + r.stdout)
    if verbose or r.returncode != 0:
        sys.stderr.write(r.stdout)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed building libosrgpu.so (see puzzletunneldiscovery_b200/build.log)")
    return OUT


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--variant":   # build.py --variant TAG DEFINE [DEFINE ...]
        print(build_variant(sys.argv[2], sys.argv[3:]))
    else:
        build(force=True, verbose=True)
