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

// The same relaxation, speed-ordered, for the fused single-phase kernel (demo_2d_LBM_KBC.py at 4096^2).
// kbc_relax spends its time in 47 IEEE divisions per cell (f / rho and feq / rho for two moment sets, 1 / feq in
// both sums).  None of them is needed:
//  * s is linear in the populations and the rho of "rho * M(f / rho)" cancels, so ds = s(f) - s(feq) is formed from
//    signed adds of d_i = f_i - feq_i over the lattice vectors (raw sums), centred once per cell:
//    sum d (e-u)_a (e-u)_b = D_ab - u_a D_b - u_b D_a + u_a u_b D_0;
//  * 1 / feq_i = 1 / (rho base) * (1 / w_i) * X_i^-1 Y_i^-1 and the inverse axis factors are already there
//    (F.invA = 1 / F.A), so ONE division serves the nine weights of <.|.>.
// feq itself is still evaluated in the reference's operation order: the reference's stabiliser is gamma = 0/0 = NaN
// exactly where the pulled populations equal their equilibrium bit for bit (LBM_collision_MRT_2d.py:213-222), and
// that pattern is pinned cell by cell (test_kbc_zero_over_zero_quirk_is_reproduced): with d == 0 every raw sum, ds
// and dh vanish and num / den is the same 0/0.  Elsewhere the result differs from kbc_relax by fp32 rounding
// (gamma is a ratio of two small numbers; it multiplies dh, which is as small), held to the trajectory bars.
template <class FE>
__device__ __forceinline__ void kbc_relax_fast(const float* f, const float* feq, const FE& F, float rho,
                                               const float* u, float c, int kbc, const KbcScalars& K, float* out) {
  using L = Lat<2>;
  float d[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) d[i] = __fsub_rn(f[i], feq[i]);   // never fused with the product that formed feq:
                                                                 // fma(-rho w, v, f) would be its rounding error, not 0
  // raw sums of d over the lattice vectors (signed adds; e is a compile-time constant in the unrolled loops)
  float D0 = 0.f, Dx = 0.f, Dy = 0.f, Dxx = 0.f, Dyy = 0.f, Dxy = 0.f, Dxxy = 0.f, Dxyy = 0.f, Dxxyy = 0.f;
#pragma unroll
  for (int i = 0; i < 9; ++i) {
    const int ex = L::ex(i), ey = L::ey(i);
    D0 += d[i];
    if (ex == 1) Dx += d[i]; else if (ex == -1) Dx -= d[i];
    if (ey == 1) Dy += d[i]; else if (ey == -1) Dy -= d[i];
    if (ex != 0) Dxx += d[i];
    if (ey != 0) Dyy += d[i];
    if (ex * ey == 1) Dxy += d[i]; else if (ex * ey == -1) Dxy -= d[i];
    if (ex != 0 && ey == 1) Dxxy += d[i]; else if (ex != 0 && ey == -1) Dxxy -= d[i];
    if (ey != 0 && ex == 1) Dxyy += d[i]; else if (ey != 0 && ex == -1) Dxyy -= d[i];
    if (ex != 0 && ey != 0) Dxxyy += d[i];
  }
  float ds[9];
  const float c2 = c * c;
  if (kbc) {   // centred second moments; s owns {T, N, Pi_xy}
    const float m20 = c2 * (Dxx - 2.0f * u[0] * Dx + u[0] * u[0] * D0);
    const float m02 = c2 * (Dyy - 2.0f * u[1] * Dy + u[1] * u[1] * D0);
    const float m11 = c2 * (Dxy - u[0] * Dy - u[1] * Dx + u[0] * u[1] * D0);
    const float T = m20 + m02, N = m20 - m02;
    ds[0] = -T;
    const float t4 = 0.25f * T, n4 = 0.25f * N, p4 = 0.25f * m11;
    ds[1] = t4 + n4; ds[3] = t4 + n4; ds[2] = t4 - n4; ds[4] = t4 - n4;
    ds[5] = p4; ds[7] = p4; ds[6] = -p4; ds[8] = -p4;
  } else {     // raw moments (preset_KBC); s owns {T, Pi_xy, Q_xxy, Q_xyy, A}; the u terms cancel in the difference
    const float c3 = c * c2, c4 = c2 * c2;
    const float m20 = c2 * Dxx, m02 = c2 * Dyy, m11 = c2 * Dxy, m21 = c3 * Dxxy, m12 = c3 * Dxyy, m22 = c4 * Dxxyy;
    const float T = m20 + m02, t4 = 0.25f * T, p4 = 0.25f * m11;
    ds[0] = -T + m22;
    ds[1] = t4 + 0.5f * (-m12 - m22); ds[3] = t4 + 0.5f * (m12 - m22);
    ds[2] = t4 + 0.5f * (-m21 - m22); ds[4] = t4 + 0.5f * (m21 - m22);
    ds[5] = p4 + 0.25f * (m22 + m12 + m21);  ds[6] = -p4 + 0.25f * (m22 - m12 + m21);
    ds[7] = p4 + 0.25f * (m22 - m12 - m21);  ds[8] = -p4 + 0.25f * (m22 + m12 - m21);
  }
  const float irb = 1.0f / (rho * F.base);
  float num = 0.f, den = 0.f, dh[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) {
    const int ex = L::ex(i), ey = L::ey(i);
    float iw = irb * (1.0f / L::w(i));            // 1 / w_i is a compile-time constant
    if (ex == 1) iw *= F.invA[0]; else if (ex == -1) iw *= F.A[0];
    if (ey == 1) iw *= F.invA[1]; else if (ey == -1) iw *= F.A[1];
    dh[i] = d[i] - ds[i];
    num += ds[i] * dh[i] * iw;
    den += dh[i] * dh[i] * iw;
  }
  const float gamma = K.inv_beta - K.two_m * num / den;   // 0/0 = NaN where f == feq, like the reference
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
