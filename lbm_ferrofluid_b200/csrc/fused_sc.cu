// Fused tier, Shan-Chen pseudopotential multiphase step on D2Q9 (demo_2d_LBM_multiphase_SC.py:113-131):
//
//     propagation -> macro_compute (rho, vel, density) -> rebounce_obstacle -> calculate_force -> BGK collision
//
// in two passes instead of ~150 eager launches.  The interaction force at x needs psi(density) at the eight
// neighbours, and density is a sum of PULLED populations, so density must exist everywhere before any cell collides
// (the same reason the HCZ step has k_hcz_moments):
//
//  k_sc_moments         pull f; rho = sum f on non-OBSTACLE cells (old rho kept on OBSTACLE,
//                       LBM_macro_compute_2d.py:85-90); density = affine(rho) (:99-106); psi = pseudopotential of the
//                       C-S equation of state, the wall value on OBSTACLE cells (LBM_collision_SC_2d.py:20, 40-59)
//  k_sc_collide_stream  pull f again (the rows are L2 / DRAM streams either way); OBSTACLE: full-way bounce-back
//                       = permutation of the pulled values; otherwise vel = sum f e c / rho,
//                       F = -G psi(x) c sum_q w_q e_q psi(x + e_q) on FLUID cells (zero outside the array, :62-97),
//                       BGK with the velocity shift u + tau F / rho (LBM_collision_2d.py:258-287).
//
// Same layout, launch geometry and speed-ordered BGK arithmetic as k_stream_collide (fused_hcz.cu); the eight psi
// neighbours are L1 hits of rows the block's other threads touch (64 x 4 blocks).  Held to the traj_sc2d golden
// (tests/test_gpu_fused.py).  Algorithmic bytes: 72 B / cell (one read + one write of f) + the re-read of f in the
// first pass (36 B) + 12 B of scalars.
#include "capi.cuh"
#include "collide.cuh"
#include "sc.cuh"

namespace lbmff {

template <bool IDX32>
__global__ void __launch_bounds__(kThreads) k_sc_moments(const float* __restrict__ f,
                                                         const uint8_t* __restrict__ flags, float* __restrict__ rho,
                                                         float* __restrict__ density, float* __restrict__ psi,
                                                         Consts k, ScConsts S, Grid g) {
  using L = Lat<2>;
  const PlaneCell c = plane_cell(g);
  if (!c.live) return;
  const int i = c.i;
  const Nbr nb(c, g);
  const PopIndex<IDX32> at(g);
  float fv[L::Q];
#pragma unroll
  for (int q = 0; q < L::Q; ++q) fv[q] = ld_spec(f + at(q, i + nb.at<2>(q, 1)));
  const bool obst = flags[i] == LBMFF_OBSTACLE;
  float s;
  if (obst) {
    s = rho[i];
  } else {
    s = 0.0f;
#pragma unroll
    for (int q = 0; q < L::Q; ++q) s += fv[q];
    rho[i] = s;
  }
  const float d = affine_density(s, k.aff);
  density[i] = d;
  psi[i] = obst ? S.psx_wall : sc_psi(d, S);
}

template <bool IDX32>
__global__ void __launch_bounds__(kThreads) k_sc_collide_stream(const float* __restrict__ fi, float* __restrict__ fo,
                                                                const uint8_t* __restrict__ flags,
                                                                const float* __restrict__ psi,
                                                                const float* __restrict__ rho_,
                                                                float* __restrict__ vel, float* __restrict__ force_out,
                                                                Consts k, ScConsts S, Grid g) {
  using L = Lat<2>;
  const PlaneCell c = plane_cell(g);
  if (!c.live) return;
  const int i = c.i;
  const Nbr nb(c, g);
  const PopIndex<IDX32> at(g);
  float fv[L::Q];   // pulls first, flag second (see k_mag_sweep)
#pragma unroll
  for (int q = 0; q < L::Q; ++q) fv[q] = ld_spec(fi + at(q, i + nb.at<2>(q, 1)));
  const uint8_t fl = flags[i];
  if (fl == LBMFF_OBSTACLE) {   // bounce-back = permutation of the pulled values; vel keeps its old value
#pragma unroll
    for (int q = 0; q < L::Q; ++q) fo[at(q, i)] = fv[L::opp(q)];
    if (force_out) { force_out[i] = 0.f; force_out[g.N + i] = 0.f; }
    return;
  }
  float m[2] = {0.f, 0.f};
#pragma unroll
  for (int q = 0; q < L::Q; ++q) {
    if (L::ex(q) != 0) m[0] += fv[q] * (float)L::ex(q);
    if (L::ey(q) != 0) m[1] += fv[q] * (float)L::ey(q);
  }
  const float rho = rho_[i];   // the first pass's sum of the same nine values
  // interaction force on FLUID cells; psi outside the array is zero (the reference pads with zeros: not periodic)
  float F[2] = {0.f, 0.f};
  if (fl == LBMFF_FLUID) {
    float sv[2] = {0.f, 0.f};
#pragma unroll
    for (int q = 1; q < L::Q; ++q) {
      const int x = c.x + L::ex(q), y = c.y + L::ey(q);
      float p = 0.0f;
      if (x >= 0 && x < g.W && y >= 0 && y < g.H) p = __ldg(psi + (i + L::ex(q) + L::ey(q) * g.W));
      if (L::ex(q) != 0) sv[0] += (L::w(q) * (float)L::ex(q)) * p;
      if (L::ey(q) != 0) sv[1] += (L::w(q) * (float)L::ey(q)) * p;
    }
    const float px = __ldg(psi + i);
    F[0] = S.negG1 * px * S.c * sv[0];
    F[1] = S.negG1 * px * S.c * sv[1];
  }
  const float inv_rho = 1.0f / rho, scale = k.c * inv_rho;
  float u[2];
#pragma unroll
  for (int a = 0; a < 2; ++a) {
    const float u0 = m[a] * scale;
    vel[a * g.N + i] = u0;
    if (force_out) force_out[a * g.N + i] = F[a];
    u[a] = fmaf(k.tau * F[a], inv_rho, u0);
  }
  const FeqFactors<2> E = feq_factors_fast<2>(u, k.inv_c);
#pragma unroll
  for (int q = 0; q < L::Q; ++q) fo[at(q, i)] = fmaf(k.keep, fv[q], feq_q<2>(E, rho, q) * k.relax);
}

}  // namespace lbmff

using namespace lbmff;

extern "C" {

int lbmff_sc_moments_d2q9(const float* f, const uint8_t* flags, float* rho, float* density, float* psi,
                          const lbmff_fluid_params* p, int B, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(f && flags && rho && density && psi && p, "null pointer");
  const Grid g = make_grid(B, 1, H, W);
  if (g.N > 0x7fffffffLL / 3) return bad_arg(__func__, "more than (2^31-1)/3 cells per batch item");
  const Consts k = make_consts(p->dx, p->dt, p->tau, p->tau_f, p->tau_g, p->density_liquid, p->density_gas,
                               p->rho_liquid, p->rho_gas, p->kappa, p->gravity, p->cot_theta);
  const ScConsts S = make_sc(p->density_liquid, p->density_gas, p->dx, p->dt);
  const PlaneLaunch pl = plane_launch(g);
  const long long N = g.N;
  for (long long b = 0; b < B; ++b) {
    if (9 * N < 0xffffffffLL)
      k_sc_moments<true><<<pl.grid, pl.block, 0, (cudaStream_t)stream>>>(f + b * 9 * N, flags + b * N, rho + b * N,
                                                                        density + b * N, psi + b * N, k, S, g);
    else
      k_sc_moments<false><<<pl.grid, pl.block, 0, (cudaStream_t)stream>>>(f + b * 9 * N, flags + b * N, rho + b * N,
                                                                         density + b * N, psi + b * N, k, S, g);
  }
  return check_launch(__func__);
}

int lbmff_sc_collide_stream_d2q9(const float* f_in, float* f_out, const uint8_t* flags, const float* psi,
                                 const float* rho, float* vel, float* force_out, const lbmff_fluid_params* p, int B,
                                 int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(f_in && f_out && f_in != f_out && flags && psi && rho && vel && p, "null or aliased pointer");
  LBMFF_REQUIRE(p->tau > 0, "tau must be positive");
  const Grid g = make_grid(B, 1, H, W);
  if (g.N > 0x7fffffffLL / 3) return bad_arg(__func__, "more than (2^31-1)/3 cells per batch item");
  const Consts k = make_consts(p->dx, p->dt, p->tau);
  const ScConsts S = make_sc(p->density_liquid, p->density_gas, p->dx, p->dt);
  const PlaneLaunch pl = plane_launch(g, 64);
  const long long N = g.N;
  for (long long b = 0; b < B; ++b) {
    float* fo = force_out ? force_out + b * 2 * N : nullptr;
    if (9 * N < 0xffffffffLL)
      k_sc_collide_stream<true><<<pl.grid, pl.block, 0, (cudaStream_t)stream>>>(
          f_in + b * 9 * N, f_out + b * 9 * N, flags + b * N, psi + b * N, rho + b * N, vel + b * 2 * N, fo, k, S, g);
    else
      k_sc_collide_stream<false><<<pl.grid, pl.block, 0, (cudaStream_t)stream>>>(
          f_in + b * 9 * N, f_out + b * 9 * N, flags + b * N, psi + b * N, rho + b * N, vel + b * 2 * N, fo, k, S, g);
  }
  return check_launch(__func__);
}

}  // extern "C"
