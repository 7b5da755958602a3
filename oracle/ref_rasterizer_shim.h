// Build shim for oracle/build_ref_rasterizer.py (test infrastructure): force-included ahead of the reference's
// soft_rasterize_cuda_kernel.cu, which dispatches on the pre-1.5 `tensor.type()` (soft_rasterize_cuda_kernel.cu:683,
// 698,760); torch 2.x's AT_DISPATCH_FLOATING_TYPES wants a ScalarType.  The reference source is compiled unmodified.
#pragma once
#include <ATen/ATen.h>
#include <ATen/Dispatch.h>
#undef AT_DISPATCH_FLOATING_TYPES
#define AT_DISPATCH_FLOATING_TYPES(TYPE, NAME, ...) \
  AT_DISPATCH_SWITCH((TYPE).scalarType(), NAME, AT_DISPATCH_CASE_FLOATING_TYPES(__VA_ARGS__))
