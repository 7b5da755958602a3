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
  // the same vectors pre-multiplied by what every population would otherwise multiply them with: t_q = e_q . uc,
  // c e_q . F = e_q . Fc, ... become signed adds (hcz_collide_pair)
  float uc[3], Fc[3], fc[3], pc[3];
};

template <int DIM>
__device__ __forceinline__ HczCell<DIM> hcz_prepare(float rho, const float* u, float dn, float p, const float* F,
                                                    const float* gfai, const float* gprho, const Consts& k) {
  HczCell<DIM> C;
#if defined(LBMFF_HCZ_FEQ_REF)    // A/B switches for tuning rebuilds (build.py: LBMFF_NVCC_EXTRA)
  const FeqFactors<DIM> f = feq_factors<DIM>(u, k.c);
#elif defined(LBMFF_HCZ_FEQ_IEEE)
  const FeqFactors<DIM> f = feq_factors_fast<DIM>(u, k.inv_c);   // divisions shared: 2 per axis instead of 6
#else
  const FeqFactors<DIM> f = feq_factors_nr<DIM>(u, k.inv_c);     // the same, Newton-refined MUFU instead of IEEE `/`
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
    C.uc[a] = (k.c * k.inv_cs2) * C.u[a]; C.Fc[a] = k.c * C.F[a];
    C.fc[a] = k.c * C.gfai[a]; C.pc[a] = k.c * C.gprho[a];
  }
  C.A0 = 1.0f - 0.5f * uv * k.inv_cs2;
  C.pd = p - k.cs2 * dn;
#ifdef LBMFF_HCZ_FEQ_IEEE
  C.dr = k.cs2 * dn / rho;
#else
  C.dr = nr_div(k.cs2 * dn, rho);
#endif
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

// The same collision for a population q and its opposite p = opp(q) at once: e_p = -e_q, so t, the three
// e . (vector) sums and A0 + t^2 / 2 are formed once per pair, and with the pre-multiplied vectors of HczCell no
// population multiplies by c or c / cs2 any more (~22 instead of ~30 instructions per population).
template <int DIM>
__device__ __forceinline__ void hcz_collide_pair(int q, const HczCell<DIM>& C, const Consts& k, float& fq, float& gq,
                                                 float& fp, float& gp, bool g_relaxed_elsewhere = false) {
  using L = Lat<DIM>;
  const float w = L::w(q);
  const float feq_ = C.rb * w * C.X[1 + L::ex(q)] * C.Y[1 + L::ey(q)] * C.Z[1 + L::ez(q)];
  const float fep_ = C.rb * w * C.X[1 - L::ex(q)] * C.Y[1 - L::ey(q)] * C.Z[1 - L::ez(q)];
  const float wpd = w * C.pd;
  const float geq_ = fmaf(C.dr, feq_, wpd), gep_ = fmaf(C.dr, fep_, wpd);
  const float t = e_dot_v<DIM>(q, C.uc);
  const float h = fmaf(0.5f * t, t, C.A0);
  const float Gq = w * (h + t), Gp = w * (h - t);
  const float ef = e_dot_v<DIM>(q, C.fc), eF = e_dot_v<DIM>(q, C.Fc), ep = e_dot_v<DIM>(q, C.pc);
  const float ffq = k.c_f * Gq * (C.ufai - ef), ffp = k.c_f * Gp * (C.ufai + ef);
  const float gfq = k.c_g * (Gq * (eF - C.uF) + (Gq - w) * (C.uprho - ep));
  const float gfp = k.c_g * (Gp * (-eF - C.uF) + (Gp - w) * (C.uprho + ep));
  if (DIM == 3) {
    fq = fq + (ffq - fq * k.inv_tau_f + feq_ * k.inv_tau_f);
    gq = gq + (gfq - gq * k.inv_tau_g + geq_ * k.inv_tau_g);
    fp = fp + (ffp - fp * k.inv_tau_f + fep_ * k.inv_tau_f);
    gp = gp + (gfp - gp * k.inv_tau_g + gep_ * k.inv_tau_g);
  } else {
    fq = fq + k.inv_tau_f * (feq_ - fq) + ffq;
    gq = (g_relaxed_elsewhere ? gq : gq + k.inv_tau_g * (geq_ - gq)) + gfq;
    fp = fp + k.inv_tau_f * (fep_ - fp) + ffp;
    gp = (g_relaxed_elsewhere ? gp : gp + k.inv_tau_g * (gep_ - gp)) + gfp;
  }
}

template <int DIM>
__device__ __forceinline__ float hcz_geq_q(int q, const HczCell<DIM>& C) {
  using L = Lat<DIM>;
  const float fe = C.rb * L::w(q) * C.X[1 + L::ex(q)] * C.Y[1 + L::ey(q)] * C.Z[1 + L::ez(q)];
  return L::w(q) * C.pd + C.dr * fe;
}

}  // namespace lbmff
