"""Build the UNMODIFIED reference CUDA extensions for sm_100a into oracle/_ref/.

TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is imported by the product
package; only tests/, __graft_entry__.smoke() and bench.py's reference /
cpu_baseline arms may load what this script builds.

The sources are compiled where they lie under /root/reference (never copied):
  thirdparty/diff-gaussian-rasterization-modified/{ext.cpp,rasterize_points.cu,
      cuda_rasterizer/{forward,backward,rasterizer_impl}.cu}   -> oracle/_ref/ref_dgr/ref_dgr.so
  thirdparty/simple-knn/{ext.cpp,spatial.cu,simple_knn.cu}      -> oracle/_ref/ref_knn/ref_knn.so
The only deviation from the reference's own setup.py is the forced include of
<cstdint>/<cfloat> (gcc 13 no longer pulls them in transitively;
cuda_rasterizer/rasterizer_impl.h:24,40-61 and simple_knn.cu:23 need them) and
the explicit sm_100a gencode.  /root/reference does not exist on the GPU box, so
this runs here; the built .so files travel with the gpurun snapshot.
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("FB_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(HERE, "_ref")

CUDA_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3",
    "-include", "cstdint", "-include", "cfloat",
]
CXX_FLAGS = ["-O3", "-include", "cstdint", "-include", "cfloat"]


def _load(name, sources, include_dirs, build=True):
    from torch.utils import cpp_extension
    bdir = os.path.join(OUT, name)
    so = os.path.join(bdir, name + ".so")
    if not build:
        return so if os.path.exists(so) else None
    os.makedirs(bdir, exist_ok=True)
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    cpp_extension.load(
        name=name, sources=sources, extra_include_paths=include_dirs,
        extra_cflags=CXX_FLAGS, extra_cuda_cflags=CUDA_FLAGS,
        build_directory=bdir, verbose=False, is_python_module=True,
    )
    return so


def build_all():
    if not os.path.isdir(REF):
        return False
    dgr = os.path.join(REF, "thirdparty", "diff-gaussian-rasterization-modified")
    knn = os.path.join(REF, "thirdparty", "simple-knn")
    _load("ref_dgr",
          [os.path.join(dgr, "ext.cpp"), os.path.join(dgr, "rasterize_points.cu"),
           os.path.join(dgr, "cuda_rasterizer", "forward.cu"),
           os.path.join(dgr, "cuda_rasterizer", "backward.cu"),
           os.path.join(dgr, "cuda_rasterizer", "rasterizer_impl.cu")],
          [os.path.join(dgr, "third_party", "glm"), dgr])
    _load("ref_knn",
          [os.path.join(knn, "ext.cpp"), os.path.join(knn, "spatial.cu"),
           os.path.join(knn, "simple_knn.cu")],
          [knn])
    return True


def import_ref(name):
    """Import a prebuilt reference extension (ref_dgr / ref_knn) or return None."""
    import importlib.util
    import torch  # noqa: F401  (libtorch must be loaded before the extension)
    so = os.path.join(OUT, name, name + ".so")
    if not os.path.exists(so):
        return None
    spec = importlib.util.spec_from_file_location(name, so)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    ok = build_all()
    print("built" if ok else "reference tree absent; nothing built")
    sys.exit(0)
