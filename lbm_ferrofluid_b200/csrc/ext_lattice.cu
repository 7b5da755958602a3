// SURVEY 8(f) row N4 -- extensions named by the north star that the reference does NOT contain:
//   * the D3Q27 lattice next to D3Q19,
//   * Guo forcing next to the reference's velocity-shift forcing,
//   * esoteric-pull IN-PLACE streaming (one population buffer instead of an A/B pair) next to A/B pull.
// One single-phase kernel: stream + macro_compute + bounce-back + BGK collision with the reference's entropic
// product-form equilibrium (LBM_collision_3d.py:133-187, which is lattice-generic: its weights factorise over
// the axes exactly like the D3Q27 weights do), templated on <lattice, collision, forcing, streaming>;
// plus a 3-D entropic KBC collision on D3Q27 (the reference's KBC is D2Q9 only, simulation_runner.py:107-108).
//
// No reference oracle exists for these rows.  What pins them instead (tests/test_gpu_ext.py):
//   D3Q19 / shift / A-B   == the parity-pinned k_stream_collide<3> of fused_hcz.cu,
//   esoteric-pull         == A/B pull bit for bit (same per-cell arithmetic, different addresses),
//   D3Q27, Guo            == oracle/ext_oracle.py (fp64 restatement) + mass / momentum balances.
//
// Layout: the reference's SoA [q][z][y][x] fp32.  D3Q27 keeps the reference's D3Q19 order for q < 19
// (LBM_collision_3d.py:76-103) and appends the eight corners as opposite pairs.
#include <cuda_fp16.h>

#include "capi.cuh"
#include "physics.cuh"

namespace lbmff {

struct Q19 {
  static constexpr int Q = 19;
  __host__ __device__ static constexpr int ex(int q) { return Lat<3>::ex(q); }
  __host__ __device__ static constexpr int ey(int q) { return Lat<3>::ey(q); }
  __host__ __device__ static constexpr int ez(int q) { return Lat<3>::ez(q); }
  __host__ __device__ static constexpr int opp(int q) { return Lat<3>::opp(q); }
  __host__ __device__ static constexpr float w(int q) { return Lat<3>::w(q); }
  static constexpr double wd(int q) {
    return q == 0 ? 1.0 / 3.0 : ((q < 5 || q == 9 || q == 14) ? 1.0 / 18.0 : 1.0 / 36.0);
  }
};

struct Q27 {
  static constexpr int Q = 27;
  __host__ __device__ static constexpr int ex(int q) {
    constexpr int t[8] = {1, -1, 1, -1, 1, -1, -1, 1};
    return q < 19 ? Lat<3>::ex(q) : t[q - 19];
  }
  __host__ __device__ static constexpr int ey(int q) {
    constexpr int t[8] = {1, -1, 1, -1, -1, 1, 1, -1};
    return q < 19 ? Lat<3>::ey(q) : t[q - 19];
  }
  __host__ __device__ static constexpr int ez(int q) {
    constexpr int t[8] = {1, -1, -1, 1, 1, -1, 1, -1};
    return q < 19 ? Lat<3>::ez(q) : t[q - 19];
  }
  __host__ __device__ static constexpr int opp(int q) { return q < 19 ? Lat<3>::opp(q) : (((q - 19) ^ 1) + 19); }
  static constexpr double wd(int q) {
    return q == 0 ? 8.0 / 27.0
                  : (q >= 19 ? 1.0 / 216.0 : ((q < 5 || q == 9 || q == 14) ? 2.0 / 27.0 : 1.0 / 54.0));
  }
  // fp32 roundings of the double quotients, like the reference's torch.Tensor([...]) weights
  __host__ __device__ static constexpr float w(int q) {
    return q == 0 ? (float)(8.0 / 27.0)
                  : (q >= 19 ? (float)(1.0 / 216.0)
                             : ((q < 5 || q == 9 || q == 14) ? (float)(2.0 / 27.0) : (float)(1.0 / 54.0)));
  }
};

// delta to the cell at x - s e_q (s = +1: pull source; s = -1: the cell x + e_q)
template <class LT>
__device__ __forceinline__ int nbr_delta(const Nbr& nb, int q, int s) {
  const int ex = s * LT::ex(q), ey = s * LT::ey(q), ez = s * LT::ez(q);
  return (ex == 1 ? nb.xm : (ex == -1 ? nb.xp : 0)) + (ey == 1 ? nb.ym : (ey == -1 ? nb.yp : 0)) +
         (ez == 1 ? nb.zm : (ez == -1 ? nb.zp : 0));
}

template <class LT>
__device__ __forceinline__ float feq_lt(const FeqFactors<3>& F, float rho, int q) {
  float v = F.base;
  const int e0 = LT::ex(q), e1 = LT::ey(q), e2 = LT::ez(q);
  if (e0 == 1) v *= F.A[0]; else if (e0 == -1) v *= F.invA[0];
  if (e1 == 1) v *= F.A[1]; else if (e1 == -1) v *= F.invA[1];
  if (e2 == 1) v *= F.A[2]; else if (e2 == -1) v *= F.invA[2];
  return rho * LT::w(q) * v;
}

// 3-D entropic multi-relaxation (KBC) on D3Q27: the reference's 2-D scheme (LBM_collision_MRT_2d.py:113-151,
// 184-226: centred second-order moments, s owns {T, N, Pi}; h = everything else) carried to three dimensions.
// In the D3Q27 moment representation (Karlin, Boesch & Chikatamarla 2014) the second-order moments enter
//   rest: -rho T;  (s,0,0): rho/6 (2 Nxz - Nyz + T);  (0,s,0): rho/6 (-Nxz + 2 Nyz + T);  (0,0,s): rho/6 (-Nxz - Nyz + T);
//   (s,l,0): rho/4 s l Pxy;  (s,0,d): rho/4 s d Pxz;  (0,l,d): rho/4 l d Pyz;  corners: 0,
// T = Mxx + Myy + Mzz, Nxz = Mxx - Mzz, Nyz = Myy - Mzz.  ds is linear in the moments, so it is formed once from
// the moments of f - feq about the velocity uc (lattice units; these are rho M, so rho drops out of ds).
// gamma = 2 (BGK) where dh vanishes.
//
// stress_moments: d_q = f_q - feq_q and the six second moments of d about uc = rho M_ab of the oracle's statement.  Raw moments with the integer lattice vectors are signed adds; the centring
// C_ab = S_ab - u_a S_b - u_b S_a + u_a u_b S is done once.
struct Stress { float xx, yy, zz, xy, xz, yz; };
template <class LT>
__device__ __forceinline__ Stress stress_moments(const float* fv, const float* fe, const float* uc, float* d) {
  float S = 0.f, Sx = 0.f, Sy = 0.f, Sz = 0.f;
  Stress m{0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int q = 0; q < LT::Q; ++q) {
    const int ex = LT::ex(q), ey = LT::ey(q), ez = LT::ez(q);
    d[q] = fv[q] - fe[q];
    S += d[q];
    if (ex != 0) { Sx += (float)ex * d[q]; m.xx += d[q]; }
    if (ey != 0) { Sy += (float)ey * d[q]; m.yy += d[q]; }
    if (ez != 0) { Sz += (float)ez * d[q]; m.zz += d[q]; }
    if (ex * ey != 0) m.xy += (float)(ex * ey) * d[q];
    if (ex * ez != 0) m.xz += (float)(ex * ez) * d[q];
    if (ey * ez != 0) m.yz += (float)(ey * ez) * d[q];
  }
  const float ux = uc[0], uy = uc[1], uz = uc[2];
  m.xx += ux * fmaf(ux, S, -2.0f * Sx); m.yy += uy * fmaf(uy, S, -2.0f * Sy); m.zz += uz * fmaf(uz, S, -2.0f * Sz);
  m.xy += fmaf(ux * uy, S, -fmaf(ux, Sy, uy * Sx));
  m.xz += fmaf(ux * uz, S, -fmaf(ux, Sz, uz * Sx));
  m.yz += fmaf(uy * uz, S, -fmaf(uy, Sz, uz * Sy));
  return m;
}

template <class LT>
__device__ __forceinline__ void kbc3_relax(const float* fv, const float* fe, float rho, const float* uc, float beta,
                                           float inv_beta, float two_m, float* out) {
  constexpr int Q = LT::Q;
  float d[Q];
  const Stress m = stress_moments<LT>(fv, fe, uc, d);
  (void)rho;
  const float xy = m.xy, xz = m.xz, yz = m.yz;
  const float T = m.xx + m.yy + m.zz, Nxz = m.xx - m.zz, Nyz = m.yy - m.zz;
  const float r6 = 1.0f / 6.0f, r4 = 0.25f;
  const float ax = r6 * (2.0f * Nxz - Nyz + T), ay = r6 * (2.0f * Nyz - Nxz + T), az = r6 * (T - Nxz - Nyz);
  float num = 0.f, den = 0.f, ds[Q];
#pragma unroll
  for (int q = 0; q < Q; ++q) {
    const int ex = LT::ex(q), ey = LT::ey(q), ez = LT::ez(q);
    const int n = (ex != 0) + (ey != 0) + (ez != 0);
    float v = 0.f;
    if (n == 0) v = -T;
    else if (n == 1) v = ex != 0 ? ax : (ey != 0 ? ay : az);
    else if (n == 2) v = ez == 0 ? r4 * (float)(ex * ey) * xy : (ey == 0 ? r4 * (float)(ex * ez) * xz : r4 * (float)(ey * ez) * yz);
    ds[q] = v;
    const float dh = d[q] - v;
    const float inv_fe = 1.0f / fe[q];
    num = fmaf(v * dh, inv_fe, num);
    den = fmaf(dh * dh, inv_fe, den);
  }
  const float gamma = den > 0.f ? inv_beta - two_m * num / den : 2.0f;
#pragma unroll
  for (int q = 0; q < Q; ++q) {
    const float dh = d[q] - ds[q];
    out[q] = fv[q] - beta * fmaf(gamma, dh, 2.0f * ds[q]);
  }
}

// Three-rate MRT on the same moment split (both lattices: only rest, axis and planar-diagonal populations carry
// the stress part): out = f - w_s ds_dev - w_b ds_bulk - w_h dh, with ds_dev the deviatoric (N, Pi) and ds_bulk
// the trace (T) piece of kbc3_relax's ds.  w_s = 1/tau sets the shear viscosity exactly as BGK does, w_b the bulk
// viscosity, w_h damps the non-hydrodynamic moments; all equal is BGK (oracle/ext_oracle.py::mrt3_relax).
template <class LT>
__device__ __forceinline__ void mrt3_relax(const float* fv, const float* fe, float rho, const float* uc, float w_s,
                                           float w_b, float w_h, float* out) {
  constexpr int Q = LT::Q;
  float d[Q];
  const Stress m = stress_moments<LT>(fv, fe, uc, d);
  const float xx = m.xx, yy = m.yy, zz = m.zz, xy = m.xy, xz = m.xz, yz = m.yz;
  (void)rho;
  const float T = xx + yy + zz, Nxz = xx - zz, Nyz = yy - zz;
  // (w_s - w_h) ds_dev + (w_b - w_h) ds_bulk per population class; the rest is w_h (f - feq)
  const float ds6 = (w_s - w_h) * (1.0f / 6.0f), ds4 = (w_s - w_h) * 0.25f, db = w_b - w_h;
  const float bulk_axis = db * (1.0f / 6.0f) * T, bulk_rest = -db * T;
  const float ax = fmaf(ds6, 2.0f * Nxz - Nyz, bulk_axis), ay = fmaf(ds6, 2.0f * Nyz - Nxz, bulk_axis),
              az = fmaf(ds6, -Nxz - Nyz, bulk_axis);
  const float pxy = ds4 * xy, pxz = ds4 * xz, pyz = ds4 * yz;
#pragma unroll
  for (int q = 0; q < Q; ++q) {
    const int ex = LT::ex(q), ey = LT::ey(q), ez = LT::ez(q);
    const int n = (ex != 0) + (ey != 0) + (ez != 0);
    float v = 0.f;
    if (n == 0) v = bulk_rest;
    else if (n == 1) v = ex != 0 ? ax : (ey != 0 ? ay : az);
    else if (n == 2) v = ez == 0 ? (float)(ex * ey) * pxy : (ey == 0 ? (float)(ex * ez) * pxz : (float)(ey * ez) * pyz);
    out[q] = fv[q] - fmaf(w_h, d[q], v);
  }
}

// Population storage.  float: the reference's fp32 populations.  __half: fp16 STORAGE with fp32 arithmetic
// (SURVEY 8f N4 "mixed-precision population storage"): what is stored is the deviation f_q - w_q rho0 from the
// rest equilibrium of a reference density rho0 (the shifted-DDF form of Lehmann et al. 2022), which keeps the
// 10-bit mantissa on the part of f that carries the flow; bytes per cell update halve.
// ld_nc: read-only path, issued ahead of the flag test (A/B pull).  ld: plain coherent load for the in-place
// esoteric buffers, which the same kernel rewrites.  Both volatile so they are not merged or dropped.
template <class ST>
struct Pop;
template <>
struct Pop<float> {
  __device__ static __forceinline__ float ld_nc(const float* p, float) { return ld_spec(p); }
  __device__ static __forceinline__ float ld(const float* p, float) {
    float v;
    asm volatile("ld.global.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
  }
  __device__ static __forceinline__ void st(float* p, float v, float) { *p = v; }
};
template <>
struct Pop<__half> {
  __device__ static __forceinline__ float ld_nc(const __half* p, float shift) {
    unsigned short r;
    asm volatile("ld.global.nc.u16 %0, [%1];" : "=h"(r) : "l"(p));
    return __half2float(__ushort_as_half(r)) + shift;
  }
  __device__ static __forceinline__ float ld(const __half* p, float shift) {
    unsigned short r;
    asm volatile("ld.global.u16 %0, [%1];" : "=h"(r) : "l"(p));
    return __half2float(__ushort_as_half(r)) + shift;
  }
  __device__ static __forceinline__ void st(__half* p, float v, float shift) { *p = __float2half_rn(v - shift); }
};

enum { FORCE_NONE = 0, FORCE_SHIFT = 1, FORCE_GUO = 2 };
enum { STREAM_AB = 0, STREAM_EP_EVEN = 1, STREAM_EP_ODD = 2 };
enum { COLL_BGK = 0, COLL_KBC = 1, COLL_MRT = 2 };

struct ExtForce {
  float u[3];       // uniform body force (used when the per-cell pointer is NULL)
  float guo_scale;  // dt (1 - 1/(2 tau)) c / cs2
  float half_dt;
  float inv_c;
  float beta, inv_beta, two_m;  // KBC: 1/(2 tau), 2 tau, 2 - 2 tau
  float rho0;                   // fp16 storage: reference density of the shifted populations
  float w_s, w_b, w_h;          // MRT: 1/tau, 1/tau_bulk, 1/tau_high
};

// Esoteric-pull addressing (Lehmann 2022).  For each opposite pair (p, o = opp(p)), p < o, a cell owns two
// memory slots: one at itself and one at its neighbour x + e_p, and it reads and rewrites exactly those two --
// so the update is in place and race-free without any ordering between threads.
//   even step: load  f_p <- slot o @ x        f_o <- slot p @ x+e_p
//              store f_p -> slot p @ x+e_p    f_o -> slot o @ x
//   odd step:  load  f_p <- slot p @ x        f_o <- slot o @ x+e_p
//              store f_p -> slot o @ x+e_p    f_o -> slot p @ x
// (loaded values are the populations ARRIVING at x, stored ones are post-collision.)
template <class LT, class ST, int COLL, int FORCING, int STREAM, bool IDX32>
__global__ void __launch_bounds__(kThreads, (COLL == 2 && LT::Q == 19) ? 4 : 1) k_ext_stream_collide(const ST* fi, ST* fo,
                                                                 const uint8_t* __restrict__ flags,
                                                                 const float* __restrict__ force,
                                                                 float* __restrict__ rho_, float* __restrict__ vel,
                                                                 Consts k, ExtForce XF, Grid g) {
  constexpr int Q = LT::Q;
  const PlaneCell c = plane_cell(g);
  if (!c.live) return;
  const int i = c.i;
  const Nbr nb(c, g);
  const PopIndex<IDX32> at(g);
  float fv[Q];
  if constexpr (STREAM == STREAM_AB) {
#pragma unroll
    for (int q = 0; q < Q; ++q) fv[q] = Pop<ST>::ld_nc(fi + at(q, i + nbr_delta<LT>(nb, q, 1)), LT::w(q) * XF.rho0);
  } else {
    fv[0] = Pop<ST>::ld(fi + at(0, i), LT::w(0) * XF.rho0);
#pragma unroll
    for (int p = 1; p < Q; ++p) {
      const int o = LT::opp(p);
      if (p > o) continue;
      const int j = i + nbr_delta<LT>(nb, p, -1);
      if constexpr (STREAM == STREAM_EP_EVEN) {
        fv[p] = Pop<ST>::ld(fi + at(o, i), LT::w(p) * XF.rho0);
        fv[o] = Pop<ST>::ld(fi + at(p, j), LT::w(o) * XF.rho0);
      } else {
        fv[p] = Pop<ST>::ld(fi + at(p, i), LT::w(p) * XF.rho0);
        fv[o] = Pop<ST>::ld(fi + at(o, j), LT::w(o) * XF.rho0);
      }
    }
  }
  float out[Q];
  if (flags[i] == LBMFF_OBSTACLE) {  // full-way bounce-back; rho, vel keep their old values
#pragma unroll
    for (int q = 0; q < Q; ++q) out[q] = fv[LT::opp(q)];
  } else {
    float s = 0.f, m[3] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      s += fv[q];
      if (LT::ex(q) != 0) m[0] += fv[q] * (float)LT::ex(q);
      if (LT::ey(q) != 0) m[1] += fv[q] * (float)LT::ey(q);
      if (LT::ez(q) != 0) m[2] += fv[q] * (float)LT::ez(q);
    }
    const float rho = s, inv_rho = 1.0f / rho, scale = k.c * inv_rho;
    rho_[i] = rho;
    float Fv[3] = {0.f, 0.f, 0.f}, u[3];
    if constexpr (FORCING != FORCE_NONE) {
#pragma unroll
      for (int a = 0; a < 3; ++a) Fv[a] = force ? force[a * g.N + i] : XF.u[a];
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const float u0 = m[a] * scale;
      if constexpr (FORCING == FORCE_GUO) {
        u[a] = fmaf(XF.half_dt * Fv[a], inv_rho, u0);  // physical velocity = what is stored and what feq sees
        vel[a * g.N + i] = u[a];
      } else {
        vel[a * g.N + i] = u0;  // the reference stores the bare moment (LBM_macro_compute_3d.py:100-105)
        u[a] = FORCING == FORCE_SHIFT ? fmaf(k.tau * Fv[a], inv_rho, u0) : u0;  // LBM_collision_3d.py:141-143
      }
    }
    const FeqFactors<3> F = feq_factors_fast<3>(u, XF.inv_c);  // physics.cuh: divisions shared
    if constexpr (COLL == COLL_KBC || COLL == COLL_MRT) {
      float fe[Q];
#pragma unroll
      for (int q = 0; q < Q; ++q) fe[q] = feq_lt<LT>(F, rho, q);
      // moments are centred on the velocity that is stored (bare moment for none / shift, like
      // LBM_collision_MRT_2d.py:205-209; physical velocity for Guo)
      float uc[3];
#pragma unroll
      for (int a = 0; a < 3; ++a) uc[a] = (FORCING == FORCE_GUO ? u[a] : m[a] * scale) * XF.inv_c;
      if constexpr (COLL == COLL_KBC) kbc3_relax<LT>(fv, fe, rho, uc, XF.beta, XF.inv_beta, XF.two_m, out);
      else mrt3_relax<LT>(fv, fe, rho, uc, XF.w_s, XF.w_b, XF.w_h, out);
    } else {
#pragma unroll
      for (int q = 0; q < Q; ++q) out[q] = fmaf(k.keep, fv[q], feq_lt<LT>(F, rho, q) * k.relax);
    }
    if constexpr (FORCING == FORCE_GUO) {
      // S_q dt = dt (1 - 1/2tau) w_q [ (e_q - u).F / cs2 + (e_q.u)(e_q.F) / cs2^2 ]   (e_q in units of c)
      //        = w_q (e.a - u.a / c + (e.b)(e.a)),  a = dt (1 - 1/2tau) c F / cs2,  b = c u / cs2
      float av[3], bv[3], ua = 0.f;
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        av[a] = XF.guo_scale * Fv[a];
        bv[a] = k.c * k.inv_cs2 * u[a];
        ua = fmaf(u[a] * XF.inv_c, av[a], ua);
      }
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        float ea = 0.f, eb = 0.f;
        if (LT::ex(q) != 0) { ea += (float)LT::ex(q) * av[0]; eb += (float)LT::ex(q) * bv[0]; }
        if (LT::ey(q) != 0) { ea += (float)LT::ey(q) * av[1]; eb += (float)LT::ey(q) * bv[1]; }
        if (LT::ez(q) != 0) { ea += (float)LT::ez(q) * av[2]; eb += (float)LT::ez(q) * bv[2]; }
        out[q] = fmaf(LT::w(q), fmaf(eb, ea, ea - ua), out[q]);
      }
    }
  }
  if constexpr (STREAM == STREAM_AB) {
#pragma unroll
    for (int q = 0; q < Q; ++q) Pop<ST>::st(fo + at(q, i), out[q], LT::w(q) * XF.rho0);
  } else {
    Pop<ST>::st(fo + at(0, i), out[0], LT::w(0) * XF.rho0);
#pragma unroll
    for (int p = 1; p < Q; ++p) {
      const int o = LT::opp(p);
      if (p > o) continue;
      const int j = i + nbr_delta<LT>(nb, p, -1);
      if constexpr (STREAM == STREAM_EP_EVEN) {
        Pop<ST>::st(fo + at(p, j), out[p], LT::w(p) * XF.rho0);
        Pop<ST>::st(fo + at(o, i), out[o], LT::w(o) * XF.rho0);
      } else {
        Pop<ST>::st(fo + at(o, j), out[p], LT::w(p) * XF.rho0);
        Pop<ST>::st(fo + at(p, i), out[o], LT::w(o) * XF.rho0);
      }
    }
  }
}

// Conversion between the state every other solver of this library carries (fp32, post-collision population q
// stored at its own cell: "standard") and a packed buffer of storage type ST in one of three layouts:
//   layout 0: standard slots (A/B pull)
//   layout 1: esoteric pull before an even step:  slot p @ x = f_o(x),  slot o @ x+e_p = f_p(x)
//   layout 2: esoteric pull before an odd step:   slot o @ x = f_o(x),  slot p @ x+e_p = f_p(x)
// to_packed: std -> packed, else packed -> std.  Out of place.  fp16 storage holds f_q - w_q rho0.
template <class LT, class ST, bool IDX32>
__global__ void __launch_bounds__(kThreads) k_ext_pack(const float* __restrict__ fstd_in, float* __restrict__ fstd_out,
                                                       const ST* __restrict__ pk_in, ST* __restrict__ pk_out,
                                                       int layout, int to_packed, float rho0, Grid g) {
  constexpr int Q = LT::Q;
  const PlaneCell c = plane_cell(g);
  if (!c.live) return;
  const int i = c.i;
  const Nbr nb(c, g);
  const PopIndex<IDX32> at(g);
#pragma unroll
  for (int p = 0; p < Q; ++p) {
    const int o = LT::opp(p);
    if (p > o) continue;
    const int j = i + nbr_delta<LT>(nb, p, -1);
    // (slot, cell) holding f_o(x) and f_p(x)
    const int so = layout == 0 ? o : (layout == 1 ? p : o), co = i;
    const int sp = layout == 0 ? p : (layout == 1 ? o : p), cp = layout == 0 ? i : j;
    const float wo = LT::w(o) * rho0, wp = LT::w(p) * rho0;
    if (to_packed) {
      Pop<ST>::st(pk_out + at(so, co), fstd_in[at(o, i)], wo);
      if (p != 0) Pop<ST>::st(pk_out + at(sp, cp), fstd_in[at(p, i)], wp);
    } else {
      fstd_out[at(o, i)] = Pop<ST>::ld(pk_in + at(so, co), wo);
      if (p != 0) fstd_out[at(p, i)] = Pop<ST>::ld(pk_in + at(sp, cp), wp);
    }
  }
}

template <class LT, class ST, int COLL, int FORCING, int STREAM>
static void launch_sc(const PlaneLaunch& pl, bool ix, cudaStream_t st, const ST* fi, ST* fo,
                      const uint8_t* flags, const float* force, float* rho, float* vel, const Consts& k,
                      const ExtForce& XF, const Grid& g) {
  if (ix) k_ext_stream_collide<LT, ST, COLL, FORCING, STREAM, true><<<pl.grid, pl.block, 0, st>>>(fi, fo, flags, force, rho, vel, k, XF, g);
  else k_ext_stream_collide<LT, ST, COLL, FORCING, STREAM, false><<<pl.grid, pl.block, 0, st>>>(fi, fo, flags, force, rho, vel, k, XF, g);
}

template <class LT, class ST, int COLL, int FORCING>
static void launch_sc_stream(int stream_mode, const PlaneLaunch& pl, bool ix, cudaStream_t st, const ST* fi,
                             ST* fo, const uint8_t* flags, const float* force, float* rho, float* vel,
                             const Consts& k, const ExtForce& XF, const Grid& g) {
  if (stream_mode == STREAM_AB) launch_sc<LT, ST, COLL, FORCING, STREAM_AB>(pl, ix, st, fi, fo, flags, force, rho, vel, k, XF, g);
  else if (stream_mode == STREAM_EP_EVEN) launch_sc<LT, ST, COLL, FORCING, STREAM_EP_EVEN>(pl, ix, st, fi, fo, flags, force, rho, vel, k, XF, g);
  else launch_sc<LT, ST, COLL, FORCING, STREAM_EP_ODD>(pl, ix, st, fi, fo, flags, force, rho, vel, k, XF, g);
}

template <class LT, class ST, int COLL>
static void launch_sc_force(int forcing, int stream_mode, const PlaneLaunch& pl, bool ix, cudaStream_t st,
                            const ST* fi, ST* fo, const uint8_t* flags, const float* force, float* rho,
                            float* vel, const Consts& k, const ExtForce& XF, const Grid& g) {
  if (forcing == FORCE_NONE) launch_sc_stream<LT, ST, COLL, FORCE_NONE>(stream_mode, pl, ix, st, fi, fo, flags, force, rho, vel, k, XF, g);
  else if (forcing == FORCE_SHIFT) launch_sc_stream<LT, ST, COLL, FORCE_SHIFT>(stream_mode, pl, ix, st, fi, fo, flags, force, rho, vel, k, XF, g);
  else launch_sc_stream<LT, ST, COLL, FORCE_GUO>(stream_mode, pl, ix, st, fi, fo, flags, force, rho, vel, k, XF, g);
}

// collision: 0 = BGK, 1 = KBC (D3Q27, fp32 storage only).  streaming: 0 = A/B pull (f_in != f_out), 1 = esoteric
// pull in place (f_in == f_out), parity = steps done & 1.  rho0: reference density of fp16-stored populations.
template <class LT, class ST>
static int ext_stream_collide_impl(const ST* f_in, ST* f_out, const uint8_t* flags, const float* force,
                                   const double* force_uniform, float* rho, float* vel, double tau, int collision,
                                   int forcing, int streaming, int parity, double rho0, double dx, double dt, int B,
                                   int D, int H, int W, cudaStream_t st, const char* where, double tau_bulk = 0.0,
                                   double tau_high = 0.0) {
  constexpr bool kFp32 = sizeof(ST) == 4;
  if (!(f_in && f_out && flags && rho && vel)) return bad_arg(where, "null pointer");
  if (streaming == 0 && f_in == f_out) return bad_arg(where, "A/B pull streaming needs f_in != f_out");
  if (streaming == 1 && f_in != f_out) return bad_arg(where, "esoteric-pull streaming is in place: f_in must equal f_out");
  if (streaming != 0 && streaming != 1) return bad_arg(where, "streaming must be 0 (A/B pull) or 1 (esoteric pull)");
  if (collision < COLL_BGK || collision > COLL_MRT) return bad_arg(where, "collision must be 0 (BGK), 1 (KBC) or 2 (MRT)");
  if (collision == COLL_MRT && !kFp32) return bad_arg(where, "the MRT collision needs fp32 storage");
  if (collision == COLL_MRT && !(tau_bulk > 0 && tau_high > 0)) return bad_arg(where, "tau_bulk and tau_high must be positive");
  if (collision == COLL_KBC && (LT::Q != 27 || !kFp32)) return bad_arg(where, "3-D KBC collision needs the D3Q27 lattice and fp32 storage");
  if (forcing < FORCE_NONE || forcing > FORCE_GUO) return bad_arg(where, "forcing must be 0 (none), 1 (shift) or 2 (Guo)");
  if (forcing != FORCE_NONE && !force && !force_uniform) return bad_arg(where, "forcing requested without a force");
  if (!(tau > 0)) return bad_arg(where, "tau must be positive");
  if (B < 1 || D < 1 || H < 1 || W < 1) return bad_arg(where, "non-positive extent");
  const Grid g = make_grid(B, D, H, W);
  if (g.N > 0x7fffffffLL / 3) return bad_arg(where, "more than (2^31-1)/3 cells per batch item");
  const Consts k = make_consts(dx, dt, tau);
  const double c = dx / dt, cs2 = c * c / 3.0;
  ExtForce XF;
  for (int a = 0; a < 3; ++a) XF.u[a] = force_uniform ? (float)force_uniform[a] : 0.f;
  XF.guo_scale = (float)(dt * (1.0 - 0.5 / tau) * c / cs2);
  XF.half_dt = (float)(0.5 * dt);
  XF.inv_c = (float)(1.0 / c);
  XF.beta = (float)(0.5 / tau); XF.inv_beta = (float)(2.0 * tau); XF.two_m = (float)(2.0 - 2.0 * tau);
  XF.rho0 = (float)rho0;
  XF.w_s = (float)(1.0 / tau);
  XF.w_b = collision == COLL_MRT ? (float)(1.0 / tau_bulk) : XF.w_s;
  XF.w_h = collision == COLL_MRT ? (float)(1.0 / tau_high) : XF.w_s;
  const PlaneLaunch pl = plane_launch(g);
  const long long Q = LT::Q, N = g.N;
  const bool ix = Q * N < 0xffffffffLL;
  const int mode = streaming == 0 ? STREAM_AB : ((parity & 1) ? STREAM_EP_ODD : STREAM_EP_EVEN);
  for (long long b = 0; b < B; ++b) {
    const ST* fi = f_in + b * Q * N;
    ST* fo = f_out + b * Q * N;
    const float* fc = force ? force + b * 3 * N : nullptr;
    if constexpr (LT::Q == 27 && kFp32) {
      if (collision == COLL_KBC) {
        launch_sc_force<LT, ST, COLL_KBC>(forcing, mode, pl, ix, st, fi, fo, flags + b * N, fc, rho + b * N, vel + b * 3 * N, k, XF, g);
        continue;
      }
    }
    if constexpr (kFp32) {
      if (collision == COLL_MRT) {
        launch_sc_force<LT, ST, COLL_MRT>(forcing, mode, pl, ix, st, fi, fo, flags + b * N, fc, rho + b * N, vel + b * 3 * N, k, XF, g);
        continue;
      }
    }
    launch_sc_force<LT, ST, COLL_BGK>(forcing, mode, pl, ix, st, fi, fo, flags + b * N, fc, rho + b * N, vel + b * 3 * N, k, XF, g);
  }
  return check_launch(where);
}

// layout: 0 standard, 1 / 2 esoteric pull before an even / odd step
template <class LT, class ST>
static int ext_pack_impl(const float* fstd_in, float* fstd_out, const ST* pk_in, ST* pk_out, int layout, int to_packed,
                         double rho0, int B, int D, int H, int W, cudaStream_t st, const char* where) {
  const void* src = to_packed ? (const void*)fstd_in : (const void*)pk_in;
  const void* dst = to_packed ? (const void*)pk_out : (const void*)fstd_out;
  if (!(src && dst && src != dst)) return bad_arg(where, "null or aliased pointer");
  if (layout < 0 || layout > 2) return bad_arg(where, "layout must be 0 (standard), 1 or 2 (esoteric, even / odd step)");
  if (B < 1 || D < 1 || H < 1 || W < 1) return bad_arg(where, "non-positive extent");
  const Grid g = make_grid(B, D, H, W);
  if (g.N > 0x7fffffffLL / 3) return bad_arg(where, "more than (2^31-1)/3 cells per batch item");
  const PlaneLaunch pl = plane_launch(g);
  const long long Q = LT::Q, N = g.N;
  for (long long b = 0; b < B; ++b) {
    const float* si = fstd_in ? fstd_in + b * Q * N : nullptr;
    float* so = fstd_out ? fstd_out + b * Q * N : nullptr;
    const ST* pi = pk_in ? pk_in + b * Q * N : nullptr;
    ST* po = pk_out ? pk_out + b * Q * N : nullptr;
    if (Q * N < 0xffffffffLL)
      k_ext_pack<LT, ST, true><<<pl.grid, pl.block, 0, st>>>(si, so, pi, po, layout, to_packed, (float)rho0, g);
    else
      k_ext_pack<LT, ST, false><<<pl.grid, pl.block, 0, st>>>(si, so, pi, po, layout, to_packed, (float)rho0, g);
  }
  return check_launch(where);
}

template <class LT>
static int lattice_info(int* e_out, int* opp_out, double* w_out) {
  for (int q = 0; q < LT::Q; ++q) {
    if (e_out) { e_out[3 * q] = LT::ex(q); e_out[3 * q + 1] = LT::ey(q); e_out[3 * q + 2] = LT::ez(q); }
    if (opp_out) opp_out[q] = LT::opp(q);
    if (w_out) w_out[q] = (double)LT::w(q);
  }
  return LT::Q;
}

}  // namespace lbmff

using namespace lbmff;

extern "C" {

int lbmff_ext_lattice_info(int Q, int* e_out, int* opp_out, double* w_out) {
  if (Q == 19) return lattice_info<Q19>(e_out, opp_out, w_out);
  if (Q == 27) return lattice_info<Q27>(e_out, opp_out, w_out);
  return -1;
}

int lbmff_ext_stream_collide_d3q19(const float* f_in, float* f_out, const uint8_t* flags, const float* force,
                                   const double* force_uniform, float* rho, float* vel, double tau, int collision,
                                   int forcing, int streaming, int parity, double dx, double dt, int B, int D, int H,
                                   int W, lbmff_stream_t stream) {
  return ext_stream_collide_impl<Q19, float>(f_in, f_out, flags, force, force_uniform, rho, vel, tau, collision,
                                             forcing, streaming, parity, 0.0, dx, dt, B, D, H, W,
                                             (cudaStream_t)stream, __func__);
}
int lbmff_ext_stream_collide_d3q27(const float* f_in, float* f_out, const uint8_t* flags, const float* force,
                                   const double* force_uniform, float* rho, float* vel, double tau, int collision,
                                   int forcing, int streaming, int parity, double dx, double dt, int B, int D, int H,
                                   int W, lbmff_stream_t stream) {
  return ext_stream_collide_impl<Q27, float>(f_in, f_out, flags, force, force_uniform, rho, vel, tau, collision,
                                             forcing, streaming, parity, 0.0, dx, dt, B, D, H, W,
                                             (cudaStream_t)stream, __func__);
}
int lbmff_ext_stream_collide_mrt_d3q19(const float* f_in, float* f_out, const uint8_t* flags, const float* force,
                                       const double* force_uniform, float* rho, float* vel, double tau,
                                       double tau_bulk, double tau_high, int forcing, int streaming, int parity,
                                       double dx, double dt, int B, int D, int H, int W, lbmff_stream_t stream) {
  return ext_stream_collide_impl<Q19, float>(f_in, f_out, flags, force, force_uniform, rho, vel, tau, COLL_MRT,
                                             forcing, streaming, parity, 0.0, dx, dt, B, D, H, W,
                                             (cudaStream_t)stream, __func__, tau_bulk, tau_high);
}
int lbmff_ext_stream_collide_mrt_d3q27(const float* f_in, float* f_out, const uint8_t* flags, const float* force,
                                       const double* force_uniform, float* rho, float* vel, double tau,
                                       double tau_bulk, double tau_high, int forcing, int streaming, int parity,
                                       double dx, double dt, int B, int D, int H, int W, lbmff_stream_t stream) {
  return ext_stream_collide_impl<Q27, float>(f_in, f_out, flags, force, force_uniform, rho, vel, tau, COLL_MRT,
                                             forcing, streaming, parity, 0.0, dx, dt, B, D, H, W,
                                             (cudaStream_t)stream, __func__, tau_bulk, tau_high);
}
int lbmff_ext_stream_collide_fp16_d3q19(const void* f_in, void* f_out, const uint8_t* flags, const float* force,
                                        const double* force_uniform, float* rho, float* vel, double tau,
                                        int forcing, int streaming, int parity, double rho0, double dx, double dt,
                                        int B, int D, int H, int W, lbmff_stream_t stream) {
  return ext_stream_collide_impl<Q19, __half>((const __half*)f_in, (__half*)f_out, flags, force, force_uniform, rho,
                                              vel, tau, COLL_BGK, forcing, streaming, parity, rho0, dx, dt, B, D, H,
                                              W, (cudaStream_t)stream, __func__);
}
int lbmff_ext_stream_collide_fp16_d3q27(const void* f_in, void* f_out, const uint8_t* flags, const float* force,
                                        const double* force_uniform, float* rho, float* vel, double tau,
                                        int forcing, int streaming, int parity, double rho0, double dx, double dt,
                                        int B, int D, int H, int W, lbmff_stream_t stream) {
  return ext_stream_collide_impl<Q27, __half>((const __half*)f_in, (__half*)f_out, flags, force, force_uniform, rho,
                                              vel, tau, COLL_BGK, forcing, streaming, parity, rho0, dx, dt, B, D, H,
                                              W, (cudaStream_t)stream, __func__);
}

int lbmff_ext_esoteric_convert_d3q19(const float* src, float* dst, int parity, int to_esoteric, int B, int D, int H,
                                     int W, lbmff_stream_t stream) {
  return to_esoteric ? ext_pack_impl<Q19, float>(src, nullptr, nullptr, dst, 1 + (parity & 1), 1, 0.0, B, D, H, W, (cudaStream_t)stream, __func__)
                     : ext_pack_impl<Q19, float>(nullptr, dst, src, nullptr, 1 + (parity & 1), 0, 0.0, B, D, H, W, (cudaStream_t)stream, __func__);
}
int lbmff_ext_esoteric_convert_d3q27(const float* src, float* dst, int parity, int to_esoteric, int B, int D, int H,
                                     int W, lbmff_stream_t stream) {
  return to_esoteric ? ext_pack_impl<Q27, float>(src, nullptr, nullptr, dst, 1 + (parity & 1), 1, 0.0, B, D, H, W, (cudaStream_t)stream, __func__)
                     : ext_pack_impl<Q27, float>(nullptr, dst, src, nullptr, 1 + (parity & 1), 0, 0.0, B, D, H, W, (cudaStream_t)stream, __func__);
}
int lbmff_ext_pack_fp16_d3q19(const float* f_std, void* f_packed, int layout, double rho0, int B, int D, int H, int W,
                              lbmff_stream_t stream) {
  return ext_pack_impl<Q19, __half>(f_std, nullptr, nullptr, (__half*)f_packed, layout, 1, rho0, B, D, H, W, (cudaStream_t)stream, __func__);
}
int lbmff_ext_pack_fp16_d3q27(const float* f_std, void* f_packed, int layout, double rho0, int B, int D, int H, int W,
                              lbmff_stream_t stream) {
  return ext_pack_impl<Q27, __half>(f_std, nullptr, nullptr, (__half*)f_packed, layout, 1, rho0, B, D, H, W, (cudaStream_t)stream, __func__);
}
int lbmff_ext_unpack_fp16_d3q19(const void* f_packed, float* f_std, int layout, double rho0, int B, int D, int H, int W,
                                lbmff_stream_t stream) {
  return ext_pack_impl<Q19, __half>(nullptr, f_std, (const __half*)f_packed, nullptr, layout, 0, rho0, B, D, H, W, (cudaStream_t)stream, __func__);
}
int lbmff_ext_unpack_fp16_d3q27(const void* f_packed, float* f_std, int layout, double rho0, int B, int D, int H, int W,
                                lbmff_stream_t stream) {
  return ext_pack_impl<Q27, __half>(nullptr, f_std, (const __half*)f_packed, nullptr, layout, 0, rho0, B, D, H, W, (cudaStream_t)stream, __func__);
}

}  // extern "C"
