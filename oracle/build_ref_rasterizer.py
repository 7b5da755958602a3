"""Test infrastructure: compile the REFERENCE's own soft-rasteriser extension, from the sources where they lie under
/root/reference/renderutils/cuda (never copied), into oracle/_ref/ (git-ignored; travels to the GPU box with gpurun).

    python oracle/build_ref_rasterizer.py

The result, oracle/_ref/soft_rasterize_ref.so, is a torch extension module exposing the reference's
forward_soft_rasterize / backward_soft_rasterize (soft_rasterize_cuda.cpp:152).  tests/test_gpu_render.py loads it,
when present, to pin both oracle/raster_oracle.py and the CUDA rasteriser against the reference kernel itself.
The one shim (oracle/ref_rasterizer_shim.h, force-included) maps the pre-1.5 `tensor.type()` dispatch to torch 2.x.
Same flags as the reference's JIT load (renderutils/soft_rasterize.py:18-23: --use_fast_math) plus the sm_100a target.
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/renderutils/cuda"
OUT = os.path.join(HERE, "_ref")
NAME = "soft_rasterize_ref"


def build(verbose=False):
    if not os.path.isdir(REF):
        return None
    so = os.path.join(OUT, NAME + ".so")
    if os.path.exists(so):
        return so
    os.makedirs(OUT, exist_ok=True)
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    from torch.utils.cpp_extension import load
    load(name=NAME, sources=[os.path.join(REF, "soft_rasterize_cuda.cpp"), os.path.join(REF, "soft_rasterize_cuda_kernel.cu")],
         build_directory=OUT, with_cuda=True, verbose=verbose, is_python_module=False,
         extra_cuda_cflags=["--use_fast_math", "-gencode", "arch=compute_100a,code=sm_100a",
                            "-include", os.path.join(HERE, "ref_rasterizer_shim.h")])
    return so if os.path.exists(so) else None


if __name__ == "__main__":
    print(build(verbose=True))
