// Drop-in tier, part 2: collision operators (a9, a10, a11, a15).
#include "capi.cuh"
#include "collide.cuh"
#include "sc.cuh"

namespace lbmff {

// ---- a9 BGK with velocity-shift forcing on non-OBSTACLE cells (LBM_collision_3d.py:320-346)
// kbc >= 0 selects the entropic KBC relaxation instead (D2Q9 only, LBM_collision_MRT_2d.py:184-226).
template <int DIM, bool KBC>
__global__ void __launch_bounds__(kThreads) k_collide_single(const float* __restrict__ f,
                                                             const float* __restrict__ rho_,
                                                             const float* __restrict__ vel,
                                                             const uint8_t* __restrict__ flags,
                                                             const float* __restrict__ force,
                                                             float* __restrict__ out, int kbc, Consts k,
                                                             KbcScalars K, Grid g) {
  using L = Lat<DIM>;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.N) return;
  const long long b = blockIdx.y;
  const float* fb = f + b * L::Q * g.N;
  float* ob = out + b * L::Q * g.N;
  if (flags[b * g.N + i] == LBMFF_OBSTACLE) {
#pragma unroll
    for (int q = 0; q < L::Q; ++q) ob[q * g.N + i] = fb[q * g.N + i];
    return;
  }
  const float rho = rho_[b * g.N + i];
  float u[DIM];
#pragma unroll
  for (int a = 0; a < DIM; ++a) {
    u[a] = vel[(b * DIM + a) * g.N + i];
    if (force) u[a] = u[a] + k.tau * force[(b * DIM + a) * g.N + i] / rho;
  }
  const FeqFactors<DIM> F = feq_factors<DIM>(u, k.c);
  if constexpr (KBC) {
    float fv[9], fe[9], o[9];
#pragma unroll
    for (int q = 0; q < 9; ++q) {
      fv[q] = fb[q * g.N + i];
      fe[q] = feq_q<DIM>(F, rho, q);
    }
    // the s-part is evaluated with the *unshifted* velocity (MRT_2d.py:213-215 pass vel, not vel + tau F / rho)
    float u0[2] = {vel[(b * 2 + 0) * g.N + i], vel[(b * 2 + 1) * g.N + i]};
    kbc_relax(fv, fe, rho, u0, k.c, kbc, K, o);
#pragma unroll
    for (int q = 0; q < 9; ++q) ob[q * g.N + i] = o[q];
  } else {
#pragma unroll
    for (int q = 0; q < L::Q; ++q)
      ob[q * g.N + i] = k.keep * fb[q * g.N + i] + feq_q<DIM>(F, rho, q) / k.tau;
  }
}

// ---- a10 get_s_by_KBC as an op
__global__ void __launch_bounds__(kThreads) k_kbc_s(const float* __restrict__ f, const float* __restrict__ rho_,
                                                    const float* __restrict__ vel, float* __restrict__ out, int kbc,
                                                    float c, Grid g) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.N) return;
  const long long b = blockIdx.y;
  const float rho = rho_[b * g.N + i];
  float fr[9], s[9], u[2] = {vel[(b * 2 + 0) * g.N + i], vel[(b * 2 + 1) * g.N + i]};
#pragma unroll
  for (int q = 0; q < 9; ++q) fr[q] = f[(b * 9 + q) * g.N + i] / rho;
  kbc_s_part(fr, rho, u, c, kbc, s);
#pragma unroll
  for (int q = 0; q < 9; ++q) out[(b * 9 + q) * g.N + i] = s[q];
}

// ---- a11 Shan-Chen force (LBM_collision_SC_2d.py:22-97); ScConsts / sc_psi / make_sc: sc.cuh
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_sc_force(const float* __restrict__ density,
                                                       const uint8_t* __restrict__ flags, float* __restrict__ out,
                                                       ScConsts S, Grid g) {
  // DIM = 3 (D3Q19) is the SURVEY 8f N4 extension: the reference's Shan-Chen model is 2-D only
  using L = Lat<DIM>;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.N) return;
  const long long b = blockIdx.y;
  const float* d = density + b * g.N;
  const uint8_t* fl = flags + b * g.N;
  float fv[DIM];
#pragma unroll
  for (int a = 0; a < DIM; ++a) fv[a] = 0.f;
  if (fl[i] == LBMFF_FLUID) {
    const Cell c = cell_of(i, g);
    const float psx = sc_psi(d[i], S);
    float sv[3] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int q = 1; q < L::Q; ++q) {
      const int x = c.x + L::ex(q), y = c.y + L::ey(q), z = c.z + L::ez(q);
      float nb = 0.0f;  // zero padding outside the array (:62-67)
      if (x >= 0 && x < g.W && y >= 0 && y < g.H && z >= 0 && z < g.D) {
        const long long j = lin(g, x, y, z);
        nb = fl[j] == LBMFF_OBSTACLE ? S.psx_wall : sc_psi(d[j], S);
      }
      if (L::ex(q) != 0) sv[0] += (L::w(q) * (float)L::ex(q)) * nb;
      if (L::ey(q) != 0) sv[1] += (L::w(q) * (float)L::ey(q)) * nb;
      if (DIM == 3 && L::ez(q) != 0) sv[2] += (L::w(q) * (float)L::ez(q)) * nb;
    }
#pragma unroll
    for (int a = 0; a < DIM; ++a) fv[a] = S.negG1 * psx * S.c * sv[a];
  }
#pragma unroll
  for (int a = 0; a < DIM; ++a) out[(b * DIM + a) * g.N + i] = fv[a];
}

// ---- a15 HCZ collision on FLUID cells
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_hcz(const float* __restrict__ f, const float* __restrict__ gq,
                                                  const float* __restrict__ rho_, const float* __restrict__ vel,
                                                  const float* __restrict__ density,
                                                  const uint8_t* __restrict__ flags,
                                                  const float* __restrict__ force,
                                                  const float* __restrict__ pressure,
                                                  const float* __restrict__ dfai, const float* __restrict__ dprho,
                                                  float* __restrict__ f_out, float* __restrict__ g_out, Consts k,
                                                  int kbc_on_g, KbcScalars K, Grid g) {
  using L = Lat<DIM>;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.N) return;
  const long long b = blockIdx.y;
  float fv[L::Q], gv[L::Q];
#pragma unroll
  for (int q = 0; q < L::Q; ++q) {
    fv[q] = f[(b * L::Q + q) * g.N + i];
    gv[q] = gq[(b * L::Q + q) * g.N + i];
  }
  if (flags[b * g.N + i] == LBMFF_FLUID) {
    float u[DIM], F[DIM], df[DIM], dp[DIM];
#pragma unroll
    for (int a = 0; a < DIM; ++a) {
      const long long j = (b * DIM + a) * g.N + i;
      u[a] = vel[j]; F[a] = force[j]; df[a] = dfai[j]; dp[a] = dprho[j];
    }
    hcz_collide_cell<DIM>(fv, gv, rho_[b * g.N + i], u, density[b * g.N + i], pressure[b * g.N + i], F, df, dp, k,
                          kbc_on_g, K);
  }
#pragma unroll
  for (int q = 0; q < L::Q; ++q) {
    f_out[(b * L::Q + q) * g.N + i] = fv[q];
    g_out[(b * L::Q + q) * g.N + i] = gv[q];
  }
}

}  // namespace lbmff

using namespace lbmff;

#define GRID_ARGS(B, D, H, W)                    \
  const Grid g = make_grid(B, D, H, W);          \
  const dim3 grid(blocks_for(g.N), (unsigned)B); \
  cudaStream_t st = (cudaStream_t)stream

static Consts consts_of(const lbmff_fluid_params* p) {
  return make_consts(p->dx, p->dt, p->tau, p->tau_f, p->tau_g, p->density_liquid, p->density_gas, p->rho_liquid,
                     p->rho_gas, p->kappa, p->gravity, p->cot_theta);
}

extern "C" {

int lbmff_collision_bgk_d2q9(const float* f, const float* rho, const float* vel, const uint8_t* flags,
                             const float* force, double tau, double dx, double dt, float* f_out, int B, int H, int W,
                             lbmff_stream_t stream) {
  LBMFF_REQUIRE(f && rho && vel && flags && f_out && tau > 0, "bad arguments");
  GRID_ARGS(B, 1, H, W);
  k_collide_single<2, false><<<grid, kThreads, 0, st>>>(f, rho, vel, flags, force, f_out, 0,
                                                        make_consts(dx, dt, tau), make_kbc(tau), g);
  return check_launch(__func__);
}
int lbmff_collision_bgk_d3q19(const float* f, const float* rho, const float* vel, const uint8_t* flags,
                              const float* force, double tau, double dx, double dt, float* f_out, int B, int D, int H,
                              int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(f && rho && vel && flags && f_out && tau > 0, "bad arguments");
  GRID_ARGS(B, D, H, W);
  k_collide_single<3, false><<<grid, kThreads, 0, st>>>(f, rho, vel, flags, force, f_out, 0,
                                                        make_consts(dx, dt, tau), make_kbc(tau), g);
  return check_launch(__func__);
}
int lbmff_get_s_by_kbc_d2q9(const float* f, const float* rho, const float* vel, int kbc, double dx, double dt,
                            float* s_out, int B, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(f && rho && vel && s_out, "null pointer");
  GRID_ARGS(B, 1, H, W);
  k_kbc_s<<<grid, kThreads, 0, st>>>(f, rho, vel, s_out, kbc, (float)(dx / dt), g);
  return check_launch(__func__);
}
int lbmff_collision_kbc_d2q9(const float* f, const float* rho, const float* vel, const uint8_t* flags,
                             const float* force, double tau, int kbc, double dx, double dt, float* f_out, int B, int H,
                             int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(f && rho && vel && flags && f_out && tau > 0, "bad arguments");
  GRID_ARGS(B, 1, H, W);
  k_collide_single<2, true><<<grid, kThreads, 0, st>>>(f, rho, vel, flags, force, f_out, kbc,
                                                       make_consts(dx, dt, tau), make_kbc(tau), g);
  return check_launch(__func__);
}

int lbmff_sc_force_d2q9(const float* density, const uint8_t* flags, double density_liquid, double density_gas,
                        double dx, double dt, float* force_out, int B, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(density && flags && force_out, "null pointer");
  GRID_ARGS(B, 1, H, W);
  k_sc_force<2><<<grid, kThreads, 0, st>>>(density, flags, force_out, make_sc(density_liquid, density_gas, dx, dt), g);
  return check_launch(__func__);
}
int lbmff_sc_force_d3q19(const float* density, const uint8_t* flags, double density_liquid, double density_gas,
                         double dx, double dt, float* force_out, int B, int D, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(density && flags && force_out, "null pointer");
  GRID_ARGS(B, D, H, W);
  k_sc_force<3><<<grid, kThreads, 0, st>>>(density, flags, force_out, make_sc(density_liquid, density_gas, dx, dt), g);
  return check_launch(__func__);
}

int lbmff_collision_hcz_d2q9(const float* f, const float* g_, const float* rho, const float* vel,
                             const float* density, const uint8_t* flags, const float* force, const float* pressure,
                             const float* dfai, const float* dprho, const lbmff_fluid_params* p, int kbc_on_g,
                             float* f_out, float* g_out, int B, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(f && g_ && rho && vel && density && flags && force && pressure && dfai && dprho && p && f_out && g_out,
                "null pointer");
  GRID_ARGS(B, 1, H, W);
  k_hcz<2><<<grid, kThreads, 0, st>>>(f, g_, rho, vel, density, flags, force, pressure, dfai, dprho, f_out, g_out,
                                      consts_of(p), kbc_on_g, make_kbc(p->tau_g), g);
  return check_launch(__func__);
}
int lbmff_collision_hcz_d3q19(const float* f, const float* g_, const float* rho, const float* vel,
                              const float* density, const uint8_t* flags, const float* force,
                              const float* pressure, const float* dfai, const float* dprho,
                              const lbmff_fluid_params* p, float* f_out, float* g_out, int B, int D, int H, int W,
                              lbmff_stream_t stream) {
  LBMFF_REQUIRE(f && g_ && rho && vel && density && flags && force && pressure && dfai && dprho && p && f_out && g_out,
                "null pointer");
  GRID_ARGS(B, D, H, W);
  k_hcz<3><<<grid, kThreads, 0, st>>>(f, g_, rho, vel, density, flags, force, pressure, dfai, dprho, f_out, g_out,
                                      consts_of(p), 0, make_kbc(p->tau_g), g);
  return check_launch(__func__);
}

}  // extern "C"
