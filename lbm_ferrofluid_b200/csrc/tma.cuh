// TMA (cp.async.bulk.tensor) and mbarrier helpers for the tile-staged stencil kernels (sm_100a).
//
// Host side: tensor maps are encoded with the driver's cuTensorMapEncodeTiled, reached through the runtime's
// cudaGetDriverEntryPoint -- the library keeps linking libcudart only (tests/test_cabi.py).  A tiled tensor map takes
// ARBITRARY element start coordinates: the 16-byte rules are on the base address, on the global strides and on the
// box's inner extent, not on where a box starts, and elements outside the tensor are zero-filled.  That is what lets
// a pulled population tile, shifted by -e_q, arrive as one box load.
// Device side: one elected lane arms an mbarrier with the expected byte count, the lanes of warp 0 issue the box
// loads, every consumer waits on the barrier's phase parity.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace lbmff {
namespace tma {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}

// Map of a contiguous fp32 tensor dims[rank-1] x ... x dims[0] (dims[0] fastest) with the given box; zero OOB fill.
inline bool make_map_f32(CUtensorMap* m, const void* base, int rank, const unsigned long long* dims,
                         const unsigned* box) {
  EncodeTiledFn fn = encode_fn();
  if (!fn || rank < 1 || rank > 5) return false;
  if (((uintptr_t)base & 15) != 0) return false;
  cuuint64_t gd[5], gs[4];
  cuuint32_t bx[5], es[5];
  unsigned long long stride = 4;
  for (int a = 0; a < rank; ++a) {
    gd[a] = dims[a];
    bx[a] = box[a];
    es[a] = 1;
    stride *= dims[a];
    if (a < rank - 1) {
      if (stride % 16) return false;
      gs[a] = stride;
    }
  }
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

#ifdef __CUDACC__
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
  unsigned done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
  } while (!done);
}
__device__ __forceinline__ void load_3d(unsigned dst, const CUtensorMap* m, unsigned bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(dst), "l"(m), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void load_4d(unsigned dst, const CUtensorMap* m, unsigned bar, int c0, int c1, int c2,
                                        int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], "
      "[%2];" ::"r"(dst),
      "l"(m), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
#endif

}  // namespace tma
}  // namespace lbmff
