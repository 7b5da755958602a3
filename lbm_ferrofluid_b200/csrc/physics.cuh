// Per-cell physics shared by the drop-in kernels and the fused step kernels.
// Every formula cites the reference line it restates; arithmetic is fp32 with
// IEEE division / sqrt (no fast-math) and follows the reference's operation
// order where that is cheap, so results track the reference's fp32 rounding.
#pragma once
#include "common.cuh"

namespace lbmff {

// Carnahan-Starling EOS, a = 12 RT, b = 4 (so b*rho/4 == rho exactly).
// order 0: LBM_macro_compute_3d.py:59-71 (get_pressure)
// order 1: LBM_collision_3d.py:105-122  (equation_of_states)
__device__ __forceinline__ float eos_cs(float d, float RT, float a, int order) {
  const float t = d;
  const float om = 1.0f - t;
  const float den = om * om * om;
  if (order == 0) {
    return d * RT * t * (4.0f - 2.0f * t) / den - a * d * d + d * RT;
  }
  return d * RT * (4.0f * t - 2.0f * t * t) / den + d * RT - a * d * d;
}

// rho -> density affine map (LBM_macro_compute_3d.py:107-114, HCZ_3d.py:213-215)
struct Affine {
  float dg;     // density_gas
  float ddiff;  // density_liquid - density_gas
  float rg;     // rho_gas
  float rdiff;  // rho_liquid - rho_gas
  float inv_rdiff;  // 1 / (rho_liquid - rho_gas), formed in double (fused kernels only)
};
__device__ __forceinline__ float affine_density(float rho, const Affine& A) {
  return A.dg + A.ddiff * ((rho - A.rg) / A.rdiff);
}

// phi = -(2 (density - dg)/(dl - dg) - 1)   (demo_3d_LBM_Rosensweig_instability.py:162)
__device__ __forceinline__ float phase_field(float density, const Affine& A) {
  return -(2.0f * (density - A.dg) / A.ddiff - 1.0f);
}

// smoothed Heaviside (HCZ_3d.py:265-269); half_inv = 0.5/eps, pi_inv = pi/eps
struct Smooth {
  float eps, half_inv, pi_inv;
};
__device__ __forceinline__ float smooth_phi(float phi, const Smooth& S) {
  float r = phi > S.eps ? 1.0f : 0.0f;
  if (fabsf(phi) <= S.eps) {
    r += 0.5f + S.half_inv * phi + 0.15915494309189535f * sinf(S.pi_inv * phi);
  }
  return r;
}
// chi = k (1 - H_eps(phi)), k = 0.33 hard-coded (LBM_magnetic_3d.py:153,162)
__device__ __forceinline__ float susceptibility(float phi, const Smooth& S) {
  return 0.33f * (1.0f - smooth_phi(phi, S));
}

// Entropic product-form equilibrium (LBM_collision_3d.py:133-187):
//   feq_q = rho w_q prod_a (2 - s_a) * prod_a A_a^{e_qa},
//   s_a = sqrt(1 + 3 u_a^2 / c^2),  A_a = (2 u_a / c + s_a) / (1 - u_a / c)
template <int DIM>
struct FeqFactors {
  float base;
  float A[DIM];
  float invA[DIM];
};

template <int DIM>
__device__ __forceinline__ FeqFactors<DIM> feq_factors(const float* u, float c) {
  FeqFactors<DIM> F;
  F.base = 1.0f;
#pragma unroll
  for (int a = 0; a < DIM; ++a) {
    const float s = sqrtf(1.0f + 3.0f * u[a] * u[a] / c / c);
    F.base = (a == 0) ? (2.0f - s) : F.base * (2.0f - s);
    F.A[a] = (2.0f * u[a] / c + s) / (1.0f - u[a] / c);
    F.invA[a] = 1.0f / F.A[a];
  }
  return F;
}

// feq_factors for c == 1 (every demo: dx = dt = 1): x / 1.0f == x exactly, so dropping the divisions by c leaves
// every bit of the reference-ordered result in place; 2 IEEE divisions per axis instead of 4.
template <int DIM>
__device__ __forceinline__ FeqFactors<DIM> feq_factors_c1(const float* u) {
  FeqFactors<DIM> F;
  F.base = 1.0f;
#pragma unroll
  for (int a = 0; a < DIM; ++a) {
    // explicit round-to-nearest multiplies / adds: in feq_factors the divisions by c sit between the multiply and
    // the add, so nothing there can contract into an FMA -- and nothing may here
    const float s = sqrtf(__fadd_rn(1.0f, __fmul_rn(__fmul_rn(3.0f, u[a]), u[a])));
    F.base = (a == 0) ? (2.0f - s) : F.base * (2.0f - s);
    F.A[a] = __fadd_rn(__fmul_rn(2.0f, u[a]), s) / (1.0f - u[a]);
    F.invA[a] = 1.0f / F.A[a];
  }
  return F;
}

// Speed-ordered factors for the fused single-phase kernels: the same formula with the divisions shared, 2 per
// axis instead of the reference sequence's 4 (u / c / c, 1 / (1 - u / c), 1 / A).  Differs from feq_factors by
// fp32 rounding only; the kernels that use it are held to the trajectory parity bars, not to per-op bit parity.
template <int DIM>
__device__ __forceinline__ FeqFactors<DIM> feq_factors_fast(const float* u, float inv_c) {
  FeqFactors<DIM> F;
  F.base = 1.0f;
#pragma unroll
  for (int a = 0; a < DIM; ++a) {
    const float ul = u[a] * inv_c;
    const float s = sqrtf(fmaf(3.0f * ul, ul, 1.0f));
    const float num = fmaf(2.0f, ul, s), den = 1.0f - ul;
    F.base *= 2.0f - s;
    F.A[a] = num / den;
    F.invA[a] = den / num;
  }
  return F;
}

// Division and square root for the speed-ordered fused HCZ kernels: the MUFU approximation refined by one Newton
// step and a residual correction -- faithfully rounded (<= 1 ulp) for the O(1) operands that occur here (1 - u,
// 2u + s, density, rho), 6 instructions, and none of the slow-path checks (FCHK + branch + call) that an IEEE `/` or
// sqrtf drags along: 9 divisions + 3 square roots were ~100 of the ~1400 instructions per cell of
// k_hcz_collide_stream_tma.  Only the fused tier uses them; the drop-in operators keep IEEE division.
__device__ __forceinline__ float nr_rcp(float b) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(b));
  return fmaf(fmaf(-b, r, 1.0f), r, r);
}
__device__ __forceinline__ float nr_div(float a, float b) {
  const float r = nr_rcp(b), q = a * r;
  return fmaf(fmaf(-b, q, a), r, q);
}
__device__ __forceinline__ float nr_sqrt(float x) {
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  const float s = x * r;
  return fmaf(fmaf(-s, s, x), 0.5f * r, s);
}
template <int DIM>
__device__ __forceinline__ FeqFactors<DIM> feq_factors_nr(const float* u, float inv_c) {
  FeqFactors<DIM> F;
  F.base = 1.0f;
#pragma unroll
  for (int a = 0; a < DIM; ++a) {
    const float ul = u[a] * inv_c;
    const float s = nr_sqrt(fmaf(3.0f * ul, ul, 1.0f));
    const float num = fmaf(2.0f, ul, s), den = 1.0f - ul;
    F.base *= 2.0f - s;
    F.A[a] = nr_div(num, den);
    F.invA[a] = nr_div(den, num);
  }
  return F;
}

template <int DIM>
__device__ __forceinline__ float feq_q(const FeqFactors<DIM>& F, float rho, int q) {
  using L = Lat<DIM>;
  float v = F.base;
  const int e0 = L::ex(q), e1 = L::ey(q);
  if (e0 == 1) v *= F.A[0]; else if (e0 == -1) v *= F.invA[0];
  if (e1 == 1) v *= F.A[1]; else if (e1 == -1) v *= F.invA[1];
  if (DIM == 3) {
    const int e2 = L::ez(q);
    if (e2 == 1) v *= F.A[DIM - 1]; else if (e2 == -1) v *= F.invA[DIM - 1];
  }
  return rho * L::w(q) * v;
}

// e_q . v (lattice units, times c)
template <int DIM>
__device__ __forceinline__ float e_dot(int q, const float* v, float c) {
  using L = Lat<DIM>;
  float s = 0.0f;
  if (L::ex(q) != 0) s += v[0] * ((float)L::ex(q) * c);
  if (L::ey(q) != 0) s += v[1] * ((float)L::ey(q) * c);
  if (DIM == 3 && L::ez(q) != 0) s += v[DIM - 1] * ((float)L::ez(q) * c);
  return s;
}

// Gamma_q(u) = w (1 + eu/cs2 + eu^2/(2 cs2^2) - u^2/(2 cs2))   (HCZ_3d.py:271-281)
template <int DIM>
__device__ __forceinline__ float gamma_q(int q, const float* u, float uv, float c, float cs2) {
  const float eu = e_dot<DIM>(q, u, c);
  return Lat<DIM>::w(q) * (1.0f + eu / cs2 + 0.5f * eu * eu / cs2 / cs2 - 0.5f * uv / cs2);
}

// sum_a (e_qa c - u_a) v_a     (HCZ_3d.py:338,348,350)
template <int DIM>
__device__ __forceinline__ float emu_dot(int q, const float* u, const float* v, float c) {
  using L = Lat<DIM>;
  float s = ((float)L::ex(q) * c - u[0]) * v[0];
  s += ((float)L::ey(q) * c - u[1]) * v[1];
  if (DIM == 3) s += ((float)L::ez(q) * c - u[DIM - 1]) * v[DIM - 1];
  return s;
}

// ---------------------------------------------------------------------------
// Stencils.  `Field` is any callable (x, y, z) -> float.
// ---------------------------------------------------------------------------
// Scalar field sampler with the get_grad shell rule (LBM_collision_3d.py:217-218):
// an OBSTACLE cell reads the value of the nearest interior cell instead.
struct ObstacleClampedField {
  const float* __restrict__ p;
  const uint8_t* __restrict__ flags;
  Grid g;
  __device__ __forceinline__ float operator()(int x, int y, int z) const {
    long long i = lin(g, x, y, z);
    // the interior clamp is the identity off the shell, so only shell cells need their flag read
    const bool shell = x == 0 || x == g.W - 1 || y == 0 || y == g.H - 1 || z_on_shell(g, z);
    if (shell && flags[i] == LBMFF_OBSTACLE) {
      i = lin(g, interior(x, g.W), interior(y, g.H), interior_z(g, z));
    }
    return __ldg(p + i);
  }
};


// Field sampled at constant offsets from a centre index whose 18 neighbours are all inside the array
// (centre clamped to the interior): no wrap, no flag test, address = centre + immediate-like offsets.
struct CentredField {
  const float* __restrict__ p;  // already offset to the centre cell
  int W;
  long long HW;
  int cx, cy, cz;               // centre coordinates, so grad_at / laplacian_at can pass absolute ones
  __device__ __forceinline__ float operator()(int x, int y, int z) const {
    return __ldg(p + (x - cx) + (long long)(y - cy) * W + (long long)(z - cz) * HW);
  }
};

// The get_grad shell rule folded into the DATA: value a shell cell contributes to a gradient stencil.
// OBSTACLE shell cells contribute the value of the nearest interior cell (LBM_collision_3d.py:217-218).
__device__ __forceinline__ bool on_shell(const Cell& c, const Grid& g) {
  return c.x == 0 || c.x == g.W - 1 || c.y == 0 || c.y == g.H - 1 || z_on_shell(g, c.z);
}

// Isotropic central gradient at an INTERIOR cell (LBM_collision_3d.py:224-275,
// LBM_collision_2d.py:197-223): (k*axis + diagonals) / 12 / dx, k = 2 (3-D), 4 (2-D).
template <int DIM, class Field>
__device__ __forceinline__ void grad_at(const Field& X, int x, int y, int z, float dx, float* out) {
  const float k = Lat<DIM>::grad_axis_w();
  if (DIM == 3) {
    const float xp = X(x + 1, y, z), xm = X(x - 1, y, z);
    const float yp = X(x, y + 1, z), ym = X(x, y - 1, z);
    const float zp = X(x, y, z + 1), zm = X(x, y, z - 1);
    const float pp0 = X(x + 1, y + 1, z), mm0 = X(x - 1, y - 1, z);
    const float pm0 = X(x + 1, y - 1, z), mp0 = X(x - 1, y + 1, z);
    const float p0p = X(x + 1, y, z + 1), m0m = X(x - 1, y, z - 1);
    const float p0m = X(x + 1, y, z - 1), m0p = X(x - 1, y, z + 1);
    const float zpp = X(x, y + 1, z + 1), zmm = X(x, y - 1, z - 1);
    const float zpm = X(x, y + 1, z - 1), zmp = X(x, y - 1, z + 1);
    // reference term order kept (x: z-diagonals first, then y-diagonals)
    out[0] = (k * (xp - xm) + (p0p - m0m + p0m - m0p + pp0 - mm0 + pm0 - mp0)) / 12.0f / dx;
    out[1] = (k * (yp - ym) + (zpp - zmm + zpm - zmp + pp0 - mm0 + mp0 - pm0)) / 12.0f / dx;
    out[2] = (k * (zp - zm) + (zpp - zmm + zmp - zpm + p0p - m0m + m0p - p0m)) / 12.0f / dx;
  } else {
    const float xp = X(x + 1, y, z), xm = X(x - 1, y, z);
    const float yp = X(x, y + 1, z), ym = X(x, y - 1, z);
    const float pp = X(x + 1, y + 1, z), mm = X(x - 1, y - 1, z);
    const float pm = X(x + 1, y - 1, z), mp = X(x - 1, y + 1, z);
    out[0] = (k * (xp - xm) + (pp - mm + pm - mp)) / 12.0f / dx;
    out[1] = (k * (yp - ym) + (pp - mm + mp - pm)) / 12.0f / dx;
  }
}


// Same stencil as grad_at with the two divisions replaced by one multiplication by inv12dx = 1/(12 dx)
// (host-formed in double).  Used by the fused kernels; differs from grad_at by fp32 rounding only.
template <int DIM, class Field>
__device__ __forceinline__ void grad_at_fast(const Field& X, int x, int y, int z, float inv12dx, float* out) {
  const float k = Lat<DIM>::grad_axis_w();
  if (DIM == 3) {
    const float xp = X(x + 1, y, z), xm = X(x - 1, y, z);
    const float yp = X(x, y + 1, z), ym = X(x, y - 1, z);
    const float zp = X(x, y, z + 1), zm = X(x, y, z - 1);
    const float dxy = X(x + 1, y + 1, z) - X(x - 1, y - 1, z), sxy = X(x + 1, y - 1, z) - X(x - 1, y + 1, z);
    const float dxz = X(x + 1, y, z + 1) - X(x - 1, y, z - 1), sxz = X(x + 1, y, z - 1) - X(x - 1, y, z + 1);
    const float dyz = X(x, y + 1, z + 1) - X(x, y - 1, z - 1), syz = X(x, y + 1, z - 1) - X(x, y - 1, z + 1);
    out[0] = (k * (xp - xm) + (dxz + sxz + dxy + sxy)) * inv12dx;
    out[1] = (k * (yp - ym) + (dyz + syz + dxy - sxy)) * inv12dx;
    out[2] = (k * (zp - zm) + (dyz - syz + dxz - sxz)) * inv12dx;
  } else {
    const float xp = X(x + 1, y, z), xm = X(x - 1, y, z);
    const float yp = X(x, y + 1, z), ym = X(x, y - 1, z);
    const float d = X(x + 1, y + 1, z) - X(x - 1, y - 1, z), s = X(x + 1, y - 1, z) - X(x - 1, y + 1, z);
    out[0] = (k * (xp - xm) + (d + s)) * inv12dx;
    out[1] = (k * (yp - ym) + (d - s)) * inv12dx;
  }
}

// Isotropic Laplacian at an INTERIOR cell (LBM_collision_3d.py:281-318,
// LBM_collision_2d.py:229-256).
template <int DIM, class Field>
__device__ __forceinline__ float laplacian_at(const Field& X, int x, int y, int z, float dx2) {
  if (DIM == 3) {
    const float near = X(x + 1, y, z) + X(x - 1, y, z) + X(x, y + 1, z) + X(x, y - 1, z) + X(x, y, z + 1) +
                       X(x, y, z - 1);
    const float far = X(x + 1, y + 1, z) + X(x - 1, y + 1, z) + X(x + 1, y - 1, z) + X(x - 1, y - 1, z) +
                      X(x + 1, y, z + 1) + X(x - 1, y, z + 1) + X(x + 1, y, z - 1) + X(x - 1, y, z - 1) +
                      X(x, y + 1, z + 1) + X(x, y - 1, z + 1) + X(x, y + 1, z - 1) + X(x, y - 1, z - 1);
    return (2.0f * near + far - 24.0f * X(x, y, z)) / 6.0f / dx2;
  } else {
    const float near = X(x + 1, y, z) + X(x - 1, y, z) + X(x, y + 1, z) + X(x, y - 1, z);
    const float far = X(x + 1, y + 1, z) + X(x - 1, y + 1, z) + X(x + 1, y - 1, z) + X(x - 1, y - 1, z);
    return (4.0f * near + far - 20.0f * X(x, y, z)) / 5.0f / dx2;
  }
}

// ---------------------------------------------------------------------------
// Host-side constants: formed in double like the reference's Python scalars and
// rounded to fp32 once, where the reference's scalar (x) tensor op rounds them.
// ---------------------------------------------------------------------------
struct Consts {
  float dx, dt, c, cs2, a_eos, dx2;
  Affine aff;
  Smooth sm;
  float kappa, tau, tau_f, tau_g;
  float inv_tau_f, inv_tau_g;  // 1/tau_f, 1/tau_g (HCZ_2d.py:242,260)
  float keep;        // 1 - 1/tau
  float relax;       // 1/tau
  float k_f;         // dt (1 - 0.5/tau_f)        HCZ_3d.py:334-335
  float k_g;         // 1 - 0.5/tau_g             HCZ_3d.py:345
  float half_dt_RT;  // 0.5 dt RT                 HCZ_3d.py:248
  float half_dt;     // 0.5 dt                    HCZ_3d.py:256
  float kelvin;      // -0.5 mu0                  HCZ_3d.py:236-237
  float grav_y;      // -gravity                  LBM_collision_3d.py:124-131
  float cot;         // tan(pi/2 - contact_angle) HCZ_3d.py:97
  float mag_src;     // cs2 (0.5 - tau) dt        LBM_magnetic_3d.py:180
  // combined constants of the fused (speed-ordered) collision, formed in double
  float inv_cs2;     // 1 / cs2
  float inv_c;       // 1 / c
  float c_f;         // dt (1 - 0.5/tau_f) dt / cs2
  float c_g;         // (1 - 0.5/tau_g) dt
  float vel_scale;   // 1 / cs2 (velocity from g-moments)
  float inv12dx;     // 1 / (12 dx)
};

inline Consts make_consts(double dx, double dt, double tau = 1.0, double tau_f = 1.0, double tau_g = 1.0,
                          double dl = 1.0, double dg = 0.0, double rl = 1.0, double rg = 0.0, double kappa = 0.0,
                          double gravity = 0.0, double cot = 0.0) {
  Consts k;
  const double c = dx / dt, cs2 = c * c / 3.0;
  k.dx = (float)dx; k.dt = (float)dt; k.c = (float)c; k.cs2 = (float)cs2;
  k.a_eos = (float)(12.0 * cs2); k.dx2 = (float)(dx * dx);
  k.aff.dg = (float)dg; k.aff.ddiff = (float)(dl - dg); k.aff.rg = (float)rg; k.aff.rdiff = (float)(rl - rg);
  k.aff.inv_rdiff = (float)(1.0 / (rl - rg));
  const double eps = 0.1 * dx;
  k.sm.eps = (float)eps; k.sm.half_inv = (float)(0.5 / eps); k.sm.pi_inv = (float)(3.141592653589793 / eps);
  k.kappa = (float)kappa; k.tau = (float)tau; k.tau_f = (float)tau_f; k.tau_g = (float)tau_g;
  k.inv_tau_f = (float)(1.0 / tau_f); k.inv_tau_g = (float)(1.0 / tau_g);
  k.keep = (float)(1.0 - 1.0 / tau); k.relax = (float)(1.0 / tau);
  k.k_f = (float)(dt * (1.0 - 0.5 / tau_f)); k.k_g = (float)(1.0 - 0.5 / tau_g);
  k.half_dt_RT = (float)(0.5 * dt * cs2); k.half_dt = (float)(0.5 * dt);
  k.kelvin = (float)(-0.5 * (4 * 3.141592653589793 * 1e-7));
  k.grav_y = -(float)gravity; k.cot = (float)cot;
  k.mag_src = (float)(cs2 * (0.5 - tau) * dt);
  k.inv_cs2 = (float)(1.0 / cs2);
  k.inv_c = (float)(1.0 / c);
  k.c_f = (float)(dt * (1.0 - 0.5 / tau_f) * dt / cs2);
  k.c_g = (float)((1.0 - 0.5 / tau_g) * dt);
  k.vel_scale = (float)(1.0 / cs2);
  k.inv12dx = (float)(1.0 / (12.0 * dx));
  return k;
}

}  // namespace lbmff
