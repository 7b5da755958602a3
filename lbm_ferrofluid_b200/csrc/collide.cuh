// Per-cell collision operators (device functions) shared by the drop-in and fused kernels.
#pragma once
#include "physics.cuh"

namespace lbmff {

// ---- KBC 's' part for D2Q9 (LBM_collision_MRT_2d.py:62-182).  fr = f / rho.
// kbc != 0 : centred moments (c (e-u)^p)(c (e-u)^q), s owns {T, N, Pi_xy}
// kbc == 0 : raw moments C_mat (preset_KBC, :23-60),  s owns {T, Pi_xy, Q_xxy, Q_xyy, A}
__device__ __forceinline__ void kbc_s_part(const float* fr, float rho, const float* u, float c, int kbc, float* s) {
  using L = Lat<2>;
  float M20 = 0.f, M02 = 0.f, M11 = 0.f, M21 = 0.f, M12 = 0.f, M22 = 0.f;
  if (kbc) {
#pragma unroll
    for (int i = 0; i < 9; ++i) {
      const float dx = (float)L::ex(i) - u[0], dy = (float)L::ey(i) - u[1];
      const float x0 = c, x1 = c * dx, x2 = c * (dx * dx);
      const float y0 = c, y1 = c * dy, y2 = c * (dy * dy);
      M20 += (x2 * y0) * fr[i];
      M02 += (x0 * y2) * fr[i];
      M11 += (x1 * y1) * fr[i];
    }
  } else {
    const float c2 = c * c, c3 = c * c2, c4 = c2 * c2;
#pragma unroll
    for (int i = 0; i < 9; ++i) {
      const float ex = (float)L::ex(i), ey = (float)L::ey(i);
      M20 += (c2 * (ex * ex)) * fr[i];
      M02 += (c2 * (ey * ey)) * fr[i];
      M11 += (c2 * (ex * ey)) * fr[i];
      M21 += (c3 * (ex * ex * ey)) * fr[i];
      M12 += (c3 * (ex * ey * ey)) * fr[i];
      M22 += (c4 * (ex * ex * ey * ey)) * fr[i];
    }
  }
  const float T = M20 + M02, N = M20 - M02;
  s[0] = rho * (1.0f - T);
  const float st = 0.5f * rho * (0.5f * T);
  s[1] = st; s[2] = st; s[3] = st; s[4] = st;
  const float sp = 0.25f * rho * M11, sm = 0.25f * rho * (-M11);
  s[5] = sp; s[6] = sm; s[7] = sp; s[8] = sm;
  if (kbc) {
    const float a = 0.5f * rho * (0.5f * N), b = 0.5f * rho * (0.5f * (-N));
    s[1] += a; s[3] += a; s[2] += b; s[4] += b;
  } else {
    s[0] += rho * M22;
    s[1] += 0.5f * rho * ((u[0] - M12) - M22);
    s[3] += 0.5f * rho * (-(u[0] - M12) - M22);
    s[2] += 0.5f * rho * ((u[1] - M21) - M22);
    s[4] += 0.5f * rho * (-(u[1] - M21) - M22);
    s[5] += 0.25f * rho * (M22 + M12 + M21);
    s[6] += 0.25f * rho * (M22 - M12 + M21);
    s[7] += 0.25f * rho * (M22 - M12 - M21);
    s[8] += 0.25f * rho * (M22 + M12 - M21);
  }
}

struct KbcScalars {
  float beta;      // 0.5 / tau
  float inv_beta;  // 1 / beta
  float two_m;     // 2 - 1/beta
};
inline KbcScalars make_kbc(double tau) {
  KbcScalars k;
  const double beta = 0.5 / tau;
  k.beta = (float)beta; k.inv_beta = (float)(1.0 / beta); k.two_m = (float)(2.0 - 1.0 / beta);
  return k;
}

// Entropic KBC relaxation f + beta (-2 ds - gamma dh)  (LBM_collision_MRT_2d.py:211-223)
__device__ __forceinline__ void kbc_relax(const float* f, const float* feq, float rho, const float* u, float c,
                                          int kbc, const KbcScalars& K, float* out) {
  float fr[9], s1[9], s2[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) fr[i] = f[i] / rho;
  kbc_s_part(fr, rho, u, c, kbc, s1);
#pragma unroll
  for (int i = 0; i < 9; ++i) fr[i] = feq[i] / rho;
  kbc_s_part(fr, rho, u, c, kbc, s2);
  float num = 0.f, den = 0.f, ds[9], dh[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) {
    ds[i] = s1[i] - s2[i];
    dh[i] = (f[i] - feq[i]) - ds[i];
    num += ds[i] * dh[i] / feq[i];
    den += dh[i] * dh[i] / feq[i];
  }
  const float gamma = K.inv_beta - K.two_m * num / den;
#pragma unroll
  for (int i = 0; i < 9; ++i) out[i] = f[i] + K.beta * (-2.0f * ds[i] - gamma * dh[i]);
}

// ---- HCZ collision of one cell (HCZ_3d.py:283-361, HCZ_2d.py:193-286).
// f, g hold the pre-collision populations and receive the post-collision ones.
template <int DIM>
__device__ __forceinline__ void hcz_collide_cell(float* f, float* g, float rho, const float* u, float density,
                                                 float pressure, const float* force, const float* dfai,
                                                 const float* dprho, const Consts& k, int kbc_on_g,
                                                 const KbcScalars& K) {
  using L = Lat<DIM>;
  const FeqFactors<DIM> F = feq_factors<DIM>(u, k.c);
  float uv = 0.f, ndfai[DIM], ndprho[DIM];
#pragma unroll
  for (int a = 0; a < DIM; ++a) {
    uv += u[a] * u[a];
    ndfai[a] = -dfai[a];
    ndprho[a] = -dprho[a];
  }
  const float pd = pressure - k.cs2 * density;
  const float dr = k.cs2 * density / rho;
  float geq[L::Q];
  float gk[L::Q];
  if constexpr (DIM == 2) {
    if (kbc_on_g) {
#pragma unroll
    for (int q = 0; q < L::Q; ++q) geq[q] = L::w(q) * pd + dr * feq_q<DIM>(F, rho, q);
    kbc_relax(g, geq, rho, u, k.c, kbc_on_g == 1 ? 1 : 0, K, gk);  // 2: raw-moment split (KBC_type = LBGK)
    }
  }
#pragma unroll
  for (int q = 0; q < L::Q; ++q) {
    const float fe = feq_q<DIM>(F, rho, q);
    const float ge = L::w(q) * pd + dr * fe;
    const float Gm = gamma_q<DIM>(q, u, uv, k.c, k.cs2);
    const float ff = k.k_f * Gm / k.cs2 * emu_dot<DIM>(q, u, ndfai, k.c) * k.dt;
    const float gf = k.k_g * (Gm * emu_dot<DIM>(q, u, force, k.c) + (Gm - L::w(q)) * emu_dot<DIM>(q, u, ndprho, k.c)) *
                     k.dt;
    if (DIM == 3) {
      f[q] = f[q] + (ff - f[q] / k.tau_f + fe / k.tau_f);   // HCZ_3d.py:333-342
      g[q] = g[q] + (gf - g[q] / k.tau_g + ge / k.tau_g);   // HCZ_3d.py:344-355
    } else {
      const float inv_tf = k.inv_tau_f, inv_tg = k.inv_tau_g;  // HCZ_2d.py:242,260 (1.0 / tau in Python)
      float cg = (kbc_on_g) ? gk[q] : g[q] + inv_tg * (ge - g[q]);
      f[q] = f[q] + inv_tf * (fe - f[q]) + ff;
      g[q] = cg + gf;
    }
  }
}

}  // namespace lbmff
