"""Build libfisher_b200.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

nvcc cross-compiles without a GPU.  The .so is git-ignored but travels with the gpurun snapshot.
"""
import os
import subprocess

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libfisher_b200.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


def _sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC)) + \
        [os.path.join(os.path.dirname(PKG), "include", "fisher_b200.h")]


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(s) > t for s in _sources())


def build(force=False, verbose=False):
    if not force and not needs_build():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + ["-o", LIB, os.path.join(CSRC, "fb_api.cu")]
    if verbose:
        print(" ".join(cmd))
    subprocess.run(cmd, check=True)
    return LIB


# The library keeps earlier forms of its kernels behind build switches (the 8x4 one-pixel sweep, the per-tile 8x8 sweep
# without the segmented heavy lists, the sort without the split, the two-direction H_train kernel, the five-pass C_H = 11
# coefficients).  They are built next to the product so that tests/test_gpu_alt_forms.py can hold the shipped kernels
# against them on the GPU box; nothing in the package loads them.
ALT_DIR = os.path.join(PKG, "alt_forms")
ALT_FORMS = {
    "round1_forms": ["-DFB_SCORE_PX2=0", "-DFB_SORT_SPLIT=0", "-DFB_HESS_PX2=0", "-DFB_COEF11_DIRECT=0"],
    "per_tile_sweep": ["-DFB_SCORE_MIX=0"],
}


def alt_lib(name):
    return os.path.join(ALT_DIR, f"libfisher_b200_{name}.so")


def build_alt_forms(force=False, verbose=False):
    os.makedirs(ALT_DIR, exist_ok=True)
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    newest = max(os.path.getmtime(s) for s in _sources())
    for name, flags in ALT_FORMS.items():
        out = alt_lib(name)
        if not force and os.path.exists(out) and os.path.getmtime(out) >= newest:
            continue
        cmd = [nvcc] + NVCC_FLAGS + flags + ["-o", out, os.path.join(CSRC, "fb_api.cu")]
        if verbose:
            print(" ".join(cmd))
        subprocess.run(cmd, check=True)


if __name__ == "__main__":
    print(build(force=True, verbose=True))
    build_alt_forms(force=True, verbose=True)
