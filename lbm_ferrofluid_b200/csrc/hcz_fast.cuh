// Speed-ordered HCZ collision for the fused step kernel.
//
// Same formulas as hcz_collide_cell (collide.cuh; HCZ_3d.py:283-361, HCZ_2d.py:193-286) but regrouped so
// that everything independent of q is hoisted and divisions by constants become multiplications by
// host-formed reciprocals:  (e c - u).v = c (e.v) - u.v ;  Gamma_q = w (A0 + t + t^2/2), t = (e.u) c / cs2,
// A0 = 1 - u^2 / (2 cs2).  Results differ from the reference-ordered version by fp32 rounding only
// (checked against the same parity bars); the instruction count per cell drops ~6x.
#pragma once
#include "collide.cuh"

namespace lbmff {

template <int DIM>
struct HczCell {
  float rb;        // rho * base of the entropic feq
  float X[3], Y[3], Z[3];  // per-axis feq factor for e = -1, 0, +1
  float A0, pd, dr;
  float uF, ufai, uprho;   // u . F, u . grad(fai), u . grad(prho)
  float u[3], F[3], gfai[3], gprho[3];
};

template <int DIM>
__device__ __forceinline__ HczCell<DIM> hcz_prepare(float rho, const float* u, float dn, float p, const float* F,
                                                    const float* gfai, const float* gprho, const Consts& k) {
  HczCell<DIM> C;
#ifdef LBMFF_HCZ_FEQ_REF   // A/B switch for tuning rebuilds (build.py: LBMFF_NVCC_EXTRA)
  const FeqFactors<DIM> f = feq_factors<DIM>(u, k.c);
#else
  const FeqFactors<DIM> f = feq_factors_fast<DIM>(u, k.inv_c);   // divisions shared: 2 per axis instead of 6
#endif
  C.rb = rho * f.base;
  C.X[0] = f.invA[0]; C.X[1] = 1.0f; C.X[2] = f.A[0];
  C.Y[0] = f.invA[1]; C.Y[1] = 1.0f; C.Y[2] = f.A[1];
  C.Z[0] = DIM == 3 ? f.invA[DIM - 1] : 1.0f; C.Z[1] = 1.0f; C.Z[2] = DIM == 3 ? f.A[DIM - 1] : 1.0f;
  float uv = 0.f;
  C.uF = 0.f; C.ufai = 0.f; C.uprho = 0.f;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const bool on = a < DIM;
    C.u[a] = on ? u[a] : 0.f; C.F[a] = on ? F[a] : 0.f;
    C.gfai[a] = on ? gfai[a] : 0.f; C.gprho[a] = on ? gprho[a] : 0.f;
    uv += C.u[a] * C.u[a];
    C.uF += C.u[a] * C.F[a]; C.ufai += C.u[a] * C.gfai[a]; C.uprho += C.u[a] * C.gprho[a];
  }
  C.A0 = 1.0f - 0.5f * uv * k.inv_cs2;
  C.pd = p - k.cs2 * dn;
  C.dr = k.cs2 * dn / rho;
  return C;
}

// e_q . v with compile-time e (adds only)
template <int DIM>
__device__ __forceinline__ float e_dot_v(int q, const float* v) {
  using L = Lat<DIM>;
  float s = 0.f;
  if (L::ex(q) == 1) s += v[0]; else if (L::ex(q) == -1) s -= v[0];
  if (L::ey(q) == 1) s += v[1]; else if (L::ey(q) == -1) s -= v[1];
  if (DIM == 3) { if (L::ez(q) == 1) s += v[2]; else if (L::ez(q) == -1) s -= v[2]; }
  return s;
}

// post-collision (f_q, g_q) from pre-collision values; KBC-on-g (2-D) is handled by the caller
template <int DIM>
__device__ __forceinline__ void hcz_collide_q(int q, const HczCell<DIM>& C, const Consts& k, float& fq, float& gq,
                                              bool g_relaxed_elsewhere = false) {
  using L = Lat<DIM>;
  const float w = L::w(q);
  const float fe = C.rb * w * C.X[1 + L::ex(q)] * C.Y[1 + L::ey(q)] * C.Z[1 + L::ez(q)];
  const float ge = w * C.pd + C.dr * fe;
  const float t = k.c * e_dot_v<DIM>(q, C.u) * k.inv_cs2;
  const float Gm = w * (C.A0 + t + 0.5f * t * t);
  const float ff = k.c_f * Gm * (C.ufai - k.c * e_dot_v<DIM>(q, C.gfai));
  const float gf = k.c_g * (Gm * (k.c * e_dot_v<DIM>(q, C.F) - C.uF) + (Gm - w) * (C.uprho - k.c * e_dot_v<DIM>(q, C.gprho)));
  if (DIM == 3) {
    fq = fq + (ff - fq * k.inv_tau_f + fe * k.inv_tau_f);
    gq = gq + (gf - gq * k.inv_tau_g + ge * k.inv_tau_g);
  } else {
    fq = fq + k.inv_tau_f * (fe - fq) + ff;
    gq = (g_relaxed_elsewhere ? gq : gq + k.inv_tau_g * (ge - gq)) + gf;
  }
}

template <int DIM>
__device__ __forceinline__ float hcz_geq_q(int q, const HczCell<DIM>& C) {
  using L = Lat<DIM>;
  const float fe = C.rb * L::w(q) * C.X[1 + L::ex(q)] * C.Y[1 + L::ey(q)] * C.Z[1 + L::ez(q)];
  return L::w(q) * C.pd + C.dr * fe;
}

}  // namespace lbmff
