// Drop-in tier, part 4: LBMMagnetic{2,3}d.get_H_int (a16).
// k_mag_collide: psi, source, collision of h (one pass over h).  k_neg_grad: H_int = -grad(psi).
#include <mutex>

#include "capi.cuh"
#include "magnetic.cuh"

namespace lbmff {

struct ChiOfPhi {
  const float* __restrict__ phi;
  Smooth sm;
  Grid g;
  __device__ __forceinline__ float operator()(int x, int y, int z) const {
    return susceptibility(__ldg(phi + lin(g, x, y, z)), sm);
  }
};

template <int DIM>
__global__ void __launch_bounds__(kThreads) k_mag_collide(const float* __restrict__ phi,
                                                          const uint8_t* __restrict__ flags, HextFaces Hf,
                                                          const float* __restrict__ h, float* __restrict__ h_out,
                                                          float* __restrict__ psi_out, Consts k, Grid g) {
  using L = Lat<DIM>;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.N) return;
  const long long b = blockIdx.y;
  float hv[L::Q];
#pragma unroll
  for (int q = 0; q < L::Q; ++q) hv[q] = h[(b * L::Q + q) * g.N + i];
  const float psi = psi_of<DIM>(hv);
  psi_out[b * g.N + i] = psi;
  const uint8_t fl = flags[b * g.N + i];
  if (fl != LBMFF_OBSTACLE) {
    float rhs = 0.0f;
    if (fl == LBMFF_FLUID) {
      const Cell c = cell_of(i, g);
      HextFaces Hb = Hf;
      Hb.x += b * (long long)g.D * g.H * (g.W + 1);
      Hb.y += b * (long long)g.D * (g.H + 1) * g.W;
      if (DIM == 3) Hb.z += b * (long long)(g.D + 1) * g.H * g.W;
      rhs = magnetic_rhs_at<DIM>(ChiOfPhi{phi + b * g.N, k.sm, g}, Hb, c.x, c.y, c.z, g, k.dx);
    }
#pragma unroll
    for (int q = 0; q < L::Q; ++q) hv[q] = mag_collide_q<DIM>(q, hv[q], psi, rhs, k);
  }
#pragma unroll
  for (int q = 0; q < L::Q; ++q) h_out[(b * L::Q + q) * g.N + i] = hv[q];
}

template <int DIM>
__global__ void __launch_bounds__(kThreads) k_neg_grad(const float* __restrict__ in,
                                                       const uint8_t* __restrict__ flags, float* __restrict__ out,
                                                       float dx, Grid g) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.N) return;
  const long long b = blockIdx.y;
  const Cell c = cell_of(i, g);
  float gr[3];
  grad_at<DIM>(ObstacleClampedField{in + b * g.N, flags + b * g.N, g}, interior(c.x, g.W), interior(c.y, g.H),
               DIM == 3 ? interior(c.z, g.D) : 0, dx, gr);
#pragma unroll
  for (int a = 0; a < DIM; ++a) out[(b * DIM + a) * g.N + i] = -gr[a];
}

template <int DIM>
static int h_int_impl(const float* phi, const uint8_t* flags, const float* Hx, const float* Hy, const float* Hz,
                      const float* h, double tau, double dx, double dt, float* H_int, float* h_out, float* psi_ws,
                      int B, int D, int H, int W, cudaStream_t st, const char* where) {
  if (!(phi && flags && Hx && Hy && (DIM == 2 || Hz) && h && H_int && h_out && psi_ws))
    return bad_arg(where, "null pointer");
  if (W < 3 || H < 3 || (DIM == 3 && D < 3)) return bad_arg(where, "every extent must be >= 3");
  if (!(tau > 0)) return bad_arg(where, "tau must be positive");
  const Grid g = make_grid(B, D, H, W);
  const Consts k = make_consts(dx, dt, tau);
  const dim3 grid(blocks_for(g.N), (unsigned)B);
  k_mag_collide<DIM><<<grid, kThreads, 0, st>>>(phi, flags, HextFaces{Hx, Hy, Hz}, h, h_out, psi_ws, k, g);
  k_neg_grad<DIM><<<grid, kThreads, 0, st>>>(psi_ws, flags, H_int, k.dx, g);
  return check_launch(where);
}

// ---- allocation-free variant (lbmff_get_H_int_d2q9_static) ---------------------------------------------------------
// The psi scratch lives in the library image: every GPU that loads liblbmff.so holds these 4 MiB until exit.
// In host code `g_psi_static` is only a stand-in for the symbol, not a device address; the device address is looked
// up once per device (with that device current) by cudaGetSymbolAddress and cached.
static __device__ __align__(256) float g_psi_static[LBMFF_STATIC_PSI_MAX_FLOATS];

static int static_psi_address(float** out, const char* where) {
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static float* cached[kMaxDevices] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    set_last_error(where, e);
    return (int)e;
  }
  if (dev < 0 || dev >= kMaxDevices) return bad_arg(where, "device ordinal out of range for the static scratch table");
  std::lock_guard<std::mutex> lock(mu);
  if (!cached[dev]) {
    float* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_psi_static);
    if (e != cudaSuccess || !p) {
      set_last_error(where, e != cudaSuccess ? e : cudaErrorInvalidSymbol);
      return (int)(e != cudaSuccess ? e : cudaErrorInvalidSymbol);
    }
    cached[dev] = p;
  }
  *out = cached[dev];
  return 0;
}

}  // namespace lbmff

using namespace lbmff;

extern "C" {

int lbmff_get_H_int_d2q9(const float* phi, const uint8_t* flags, const float* Hx, const float* Hy, const float* h,
                         double tau, double dx, double dt, float* H_int, float* h_out, float* psi_ws, int B, int H,
                         int W, lbmff_stream_t stream) {
  return h_int_impl<2>(phi, flags, Hx, Hy, nullptr, h, tau, dx, dt, H_int, h_out, psi_ws, B, 1, H, W,
                       (cudaStream_t)stream, __func__);
}
int lbmff_get_H_int_d3q19(const float* phi, const uint8_t* flags, const float* Hx, const float* Hy,
                          const float* Hz, const float* h, double tau, double dx, double dt, float* H_int,
                          float* h_out, float* psi_ws, int B, int D, int H, int W, lbmff_stream_t stream) {
  return h_int_impl<3>(phi, flags, Hx, Hy, Hz, h, tau, dx, dt, H_int, h_out, psi_ws, B, D, H, W,
                       (cudaStream_t)stream, __func__);
}

int lbmff_static_psi_max_floats(void) { return LBMFF_STATIC_PSI_MAX_FLOATS; }

int lbmff_get_H_int_d2q9_static(const float* phi, const uint8_t* flags, const float* Hx, const float* Hy,
                                const float* h, double tau, double dx, double dt, float* H_int, float* h_out, int B,
                                int H, int W, lbmff_stream_t stream) {
  if (B < 1 || H < 3 || W < 3) return bad_arg(__func__, "B must be >= 1 and every extent >= 3");
  if ((long long)B * H * W > (long long)LBMFF_STATIC_PSI_MAX_FLOATS)
    return bad_arg(__func__, "B*H*W exceeds LBMFF_STATIC_PSI_MAX_FLOATS (1048576 cells): use lbmff_get_H_int_d2q9 "
                             "with caller-provided psi_ws");
  float* psi_ws = nullptr;
  if (int rc = static_psi_address(&psi_ws, __func__)) return rc;
  // k_mag_collide writes psi for every one of the B*H*W cells before k_neg_grad reads any: nothing of the previous
  // call's contents is read, so the array needs no clearing.
  return h_int_impl<2>(phi, flags, Hx, Hy, nullptr, h, tau, dx, dt, H_int, h_out, psi_ws, B, 1, H, W,
                       (cudaStream_t)stream, __func__);
}

}  // extern "C"
