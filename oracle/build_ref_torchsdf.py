"""TEST TOOLING (build container only): compiles the REFERENCE's own TorchSDF extension for sm_100a from the sources where
they lie under /root/reference (thirdparty/TorchSDF/torchsdf/csrc: bindings.cpp, unbatched_triangle_distance.cpp,
unbatched_triangle_distance_cuda.cu - unmodified) into oracle/_ref/torchsdf_ref.so.  Only an empty <THC/THCAtomics.cuh>
stand-in (oracle/shim) is added to the include path; the reference's setup.py is not run.  The .so is git-ignored but travels
to the GPU box, where tests/test_sdf_gpu.py uses it as the checker of cdx_sdf_* (never the product path).
Run:  python -m oracle.build_ref_torchsdf"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("CDX_REFERENCE_ROOT", "/root/reference")
SRC = os.path.join(REF, "thirdparty", "TorchSDF", "torchsdf", "csrc")
OUT = os.path.join(HERE, "_ref")


def build(verbose=False):
    if not os.path.isdir(SRC):
        return None
    so = os.path.join(OUT, "torchsdf_ref.so")
    if os.path.isfile(so):
        return so
    os.makedirs(os.path.join(OUT, "build_torchsdf"), exist_ok=True)
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0a")
    os.environ.setdefault("MAX_JOBS", "4")
    from torch.utils import cpp_extension
    cpp_extension.load(
        name="torchsdf_ref", sources=[os.path.join(SRC, f) for f in ("bindings.cpp", "unbatched_triangle_distance.cpp",
                                                                   "unbatched_triangle_distance_cuda.cu")],
        extra_include_paths=[os.path.join(HERE, "shim"), SRC],
        extra_cflags=["-O2", "-DWITH_CUDA"],
        extra_cuda_cflags=["-O3", "-DWITH_CUDA", "-gencode", "arch=compute_100a,code=sm_100a"],
        build_directory=os.path.join(OUT, "build_torchsdf"), with_cuda=True, is_python_module=False, verbose=verbose)
    shutil.copyfile(os.path.join(OUT, "build_torchsdf", "torchsdf_ref.so"), so)
    return so


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv))
