// Drop-in tier, part 1: one kernel per reference op.
// Thread mapping: one thread per cell, linear index with x fastest -> every
// population / scalar access is unit-stride and coalesced along x.
#include "capi.cuh"
#include "physics.cuh"

namespace lbmff {

thread_local char g_err[512] = "";
void set_last_error(const char* where, cudaError_t e) {
  snprintf(g_err, sizeof(g_err), "%s: %s (%s)", where, cudaGetErrorString(e), cudaGetErrorName(e));
}
void set_last_error_msg(const char* where, const char* msg) { snprintf(g_err, sizeof(g_err), "%s: %s", where, msg); }

// ---- a1 periodic pull streaming: f'_q(x) = f_q(x - e_q)  (LBM_propagation_3d.py:18-111)
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_propagation(const float* __restrict__ f, float* __restrict__ out,
                                                          Grid g) {
  using L = Lat<DIM>;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.N) return;
  const long long boff = (long long)blockIdx.y * L::Q * g.N;
  const Cell c = cell_of(i, g);
  const int xm = wrap(c.x - 1, g.W), xp = wrap(c.x + 1, g.W);
  const int ym = wrap(c.y - 1, g.H), yp = wrap(c.y + 1, g.H);
  const int zm = DIM == 3 ? wrap(c.z - 1, g.D) : 0, zp = DIM == 3 ? wrap(c.z + 1, g.D) : 0;
#pragma unroll
  for (int q = 0; q < L::Q; ++q) {
    const int sx = L::ex(q) == 1 ? xm : (L::ex(q) == -1 ? xp : c.x);
    const int sy = L::ey(q) == 1 ? ym : (L::ey(q) == -1 ? yp : c.y);
    const int sz = L::ez(q) == 1 ? zm : (L::ez(q) == -1 ? zp : c.z);
    out[boff + q * g.N + i] = __ldg(f + boff + q * g.N + lin(g, sx, sy, sz));
  }
}

// ---- a2 full-way bounce-back on OBSTACLE cells (LBM_propagation_3d.py:113-142)
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_rebounce(const float* __restrict__ f,
                                                       const uint8_t* __restrict__ flags, float* __restrict__ out,
                                                       Grid g) {
  using L = Lat<DIM>;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.N) return;
  const long long boff = (long long)blockIdx.y * L::Q * g.N;
  const bool obst = flags[(long long)blockIdx.y * g.N + i] == LBMFF_OBSTACLE;
#pragma unroll
  for (int q = 0; q < L::Q; ++q) {
    const int s = obst ? L::opp(q) : q;
    out[boff + q * g.N + i] = __ldg(f + boff + s * g.N + i);
  }
}

// ---- a3 moments (LBM_macro_compute_3d.py:73-120)
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_macro(const float* __restrict__ f, const float* __restrict__ rho_old,
                                                    const float* __restrict__ vel_old,
                                                    const uint8_t* __restrict__ flags, float* __restrict__ rho_out,
                                                    float* __restrict__ vel_out, float* __restrict__ density_out,
                                                    Consts k, Grid g) {
  using L = Lat<DIM>;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.N) return;
  const long long b = blockIdx.y;
  const float* fb = f + b * L::Q * g.N;
  const bool obst = flags[b * g.N + i] == LBMFF_OBSTACLE;
  float s = 0.0f, m[3] = {0.0f, 0.0f, 0.0f};
#pragma unroll
  for (int q = 0; q < L::Q; ++q) {
    const float v = __ldg(fb + q * g.N + i);
    s += v;
    if (L::ex(q) != 0) m[0] += v * (float)L::ex(q);
    if (L::ey(q) != 0) m[1] += v * (float)L::ey(q);
    if (DIM == 3 && L::ez(q) != 0) m[2] += v * (float)L::ez(q);
  }
  const float rho = obst ? rho_old[b * g.N + i] : s;
  rho_out[b * g.N + i] = rho;
  const float scale = k.c / rho;
#pragma unroll
  for (int a = 0; a < DIM; ++a) {
    const long long j = (b * DIM + a) * g.N + i;
    vel_out[j] = obst ? vel_old[j] : m[a] * scale;
  }
  if (density_out) density_out[b * g.N + i] = affine_density(rho, k.aff);
}

// ---- a4 EOS (pointwise)
__global__ void __launch_bounds__(kThreads) k_eos(const float* __restrict__ in, float* __restrict__ out, long long n,
                                                  float RT, float a, int order) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = eos_cs(in[i], RT, a, order);
}

// ---- a13 smooth_phi (pointwise)
__global__ void __launch_bounds__(kThreads) k_smooth(const float* __restrict__ in, float* __restrict__ out,
                                                     long long n, Smooth S) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = smooth_phi(in[i], S);
}

// ---- a5/a6 equilibria.  mode 0: feq; mode 1: geq (feq from feq_in if given)
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_equilibrium(const float* __restrict__ rho_,
                                                          const float* __restrict__ vel,
                                                          const float* __restrict__ force,
                                                          const float* __restrict__ density,
                                                          const float* __restrict__ pressure,
                                                          const float* __restrict__ feq_in, float* __restrict__ out,
                                                          int mode, Consts k, Grid g) {
  using L = Lat<DIM>;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.N) return;
  const long long b = blockIdx.y;
  const float rho = rho_[b * g.N + i];
  float u[DIM];
#pragma unroll
  for (int a = 0; a < DIM; ++a) {
    u[a] = vel[(b * DIM + a) * g.N + i];
    if (force) u[a] = u[a] + k.tau * force[(b * DIM + a) * g.N + i] / rho;  // LBM_collision_3d.py:141-143
  }
  FeqFactors<DIM> F;
  if (!(mode == 1 && feq_in)) F = feq_factors<DIM>(u, k.c);
  float pd = 0.0f, dr = 0.0f;
  if (mode == 1) {
    const float d = density[b * g.N + i];
    pd = pressure[b * g.N + i] - k.cs2 * d;  // LBM_collision_3d.py:205
    dr = k.cs2 * d / rho;
  }
#pragma unroll
  for (int q = 0; q < L::Q; ++q) {
    const long long j = (b * L::Q + q) * g.N + i;
    float fe = (mode == 1 && feq_in) ? feq_in[j] : feq_q<DIM>(F, rho, q);
    out[j] = mode == 0 ? fe : L::w(q) * pd + dr * fe;
  }
}

// ---- a14 Gamma(u)
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_gamma(const float* __restrict__ vel, float* __restrict__ out, Consts k,
                                                    Grid g) {
  using L = Lat<DIM>;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.N) return;
  const long long b = blockIdx.y;
  float u[DIM], uv = 0.0f;
#pragma unroll
  for (int a = 0; a < DIM; ++a) {
    u[a] = vel[(b * DIM + a) * g.N + i];
    uv += u[a] * u[a];
  }
#pragma unroll
  for (int q = 0; q < L::Q; ++q) out[(b * L::Q + q) * g.N + i] = gamma_q<DIM>(q, u, uv, k.c, k.cs2);
}

// ---- a7 get_grad: interior gradient, replicate-padded (LBM_collision_3d.py:209-279)
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_grad(const float* __restrict__ in, const uint8_t* __restrict__ flags,
                                                   float* __restrict__ out, float dx, float sign, Grid g) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.N) return;
  const long long b = blockIdx.y;
  const Cell c = cell_of(i, g);
  ObstacleClampedField X{in + b * g.N, flags + b * g.N, g};
  float gr[3];
  grad_at<DIM>(X, interior(c.x, g.W), interior(c.y, g.H), DIM == 3 ? interior(c.z, g.D) : 0, dx, gr);
#pragma unroll
  for (int a = 0; a < DIM; ++a) out[(b * DIM + a) * g.N + i] = sign * gr[a];
}

struct PlainField {
  const float* __restrict__ p;
  Grid g;
  __device__ __forceinline__ float operator()(int x, int y, int z) const { return __ldg(p + lin(g, x, y, z)); }
};

// ---- a8 get_laplacian: zero on the shell (LBM_collision_3d.py:281-318)
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_laplacian(const float* __restrict__ in, float* __restrict__ out,
                                                        float dx2, Grid g) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.N) return;
  const long long b = blockIdx.y;
  const Cell c = cell_of(i, g);
  const bool shell = c.x == 0 || c.x == g.W - 1 || c.y == 0 || c.y == g.H - 1 ||
                     (DIM == 3 && (c.z == 0 || c.z == g.D - 1));
  PlainField X{in + b * g.N, g};
  out[b * g.N + i] = shell ? 0.0f : laplacian_at<DIM>(X, c.x, c.y, c.z, dx2);
}

// ---- a17 staggered faces (grid.py:7-49)
__global__ void __launch_bounds__(kThreads) k_staggered(const float* __restrict__ in, float* __restrict__ out,
                                                        int axis, int D, int H, int W) {
  const int Wo = W + (axis == 0), Ho = H + (axis == 1), Do = D + (axis == 2);
  const long long No = (long long)Do * Ho * Wo;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= No) return;
  const long long b = blockIdx.y;
  int t = (int)(i / Wo);
  const int x = (int)(i - (long long)t * Wo);
  const int z = t / Ho;
  const int y = t - z * Ho;
  const int n = axis == 0 ? W : (axis == 1 ? H : D);
  int fidx = axis == 0 ? x : (axis == 1 ? y : z);
  fidx = clampi(fidx, 1, n - 1);  // face j in 1..n-1 averages cells j-1 and j; ends replicate
  const int x0 = axis == 0 ? fidx : x, y0 = axis == 1 ? fidx : y, z0 = axis == 2 ? fidx : z;
  const int x1 = axis == 0 ? fidx - 1 : x, y1 = axis == 1 ? fidx - 1 : y, z1 = axis == 2 ? fidx - 1 : z;
  const float* p = in + b * (long long)D * H * W;
  out[b * No + i] = (p[((long long)z0 * H + y0) * W + x0] + p[((long long)z1 * H + y1) * W + x1]) * 0.5f;
}

// ---- get_vort 2-D (LBM_macro_compute_2d.py:109-117)
__global__ void __launch_bounds__(kThreads) k_vort(const float* __restrict__ vel, float* __restrict__ out,
                                                   float two_dx, Grid g) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.N) return;
  const long long b = blockIdx.y;
  const Cell c = cell_of(i, g);
  const int x = interior(c.x, g.W), y = interior(c.y, g.H);
  const float* ux = vel + (b * 2 + 0) * g.N;
  const float* uy = vel + (b * 2 + 1) * g.N;
  out[b * g.N + i] = ((ux[lin(g, x, y + 1, 0)] - ux[lin(g, x, y - 1, 0)]) -
                      (uy[lin(g, x + 1, y, 0)] - uy[lin(g, x - 1, y, 0)])) / two_dx;
}

}  // namespace lbmff

using namespace lbmff;

#define GRID_ARGS(B, D, H, W)                 \
  const Grid g = make_grid(B, D, H, W);       \
  const dim3 grid(blocks_for(g.N), (unsigned)B); \
  cudaStream_t st = (cudaStream_t)stream

extern "C" {

const char* lbmff_last_error_string(void) { return g_err; }
int lbmff_version(void) { return 100; }

int lbmff_propagation_d2q9(const float* f, float* f_out, int B, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(f && f_out && f != f_out && B > 0 && H > 0 && W > 0, "bad arguments");
  GRID_ARGS(B, 1, H, W);
  k_propagation<2><<<grid, kThreads, 0, st>>>(f, f_out, g);
  return check_launch(__func__);
}
int lbmff_propagation_d3q19(const float* f, float* f_out, int B, int D, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(f && f_out && f != f_out && B > 0 && D > 0 && H > 0 && W > 0, "bad arguments");
  GRID_ARGS(B, D, H, W);
  k_propagation<3><<<grid, kThreads, 0, st>>>(f, f_out, g);
  return check_launch(__func__);
}
int lbmff_rebounce_obstacle_d2q9(const float* f, const uint8_t* flags, float* f_out, int B, int H, int W,
                                 lbmff_stream_t stream) {
  LBMFF_REQUIRE(f && flags && f_out && f != f_out && B > 0 && H > 0 && W > 0, "bad arguments");
  GRID_ARGS(B, 1, H, W);
  k_rebounce<2><<<grid, kThreads, 0, st>>>(f, flags, f_out, g);
  return check_launch(__func__);
}
int lbmff_rebounce_obstacle_d3q19(const float* f, const uint8_t* flags, float* f_out, int B, int D, int H, int W,
                                  lbmff_stream_t stream) {
  LBMFF_REQUIRE(f && flags && f_out && f != f_out && B > 0 && D > 0 && H > 0 && W > 0, "bad arguments");
  GRID_ARGS(B, D, H, W);
  k_rebounce<3><<<grid, kThreads, 0, st>>>(f, flags, f_out, g);
  return check_launch(__func__);
}

static Consts consts_of(const lbmff_fluid_params* p) {
  return make_consts(p->dx, p->dt, p->tau, p->tau_f, p->tau_g, p->density_liquid, p->density_gas, p->rho_liquid,
                     p->rho_gas, p->kappa, p->gravity, p->cot_theta);
}

int lbmff_macro_compute_d2q9(const float* f, const float* rho, const float* vel, const uint8_t* flags,
                             float* rho_out, float* vel_out, float* density_out, const lbmff_fluid_params* p, int B,
                             int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(f && rho && vel && flags && rho_out && vel_out && p, "null pointer");
  GRID_ARGS(B, 1, H, W);
  k_macro<2><<<grid, kThreads, 0, st>>>(f, rho, vel, flags, rho_out, vel_out, density_out, consts_of(p), g);
  return check_launch(__func__);
}
int lbmff_macro_compute_d3q19(const float* f, const float* rho, const float* vel, const uint8_t* flags,
                              float* rho_out, float* vel_out, float* density_out, const lbmff_fluid_params* p, int B,
                              int D, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(f && rho && vel && flags && rho_out && vel_out && p, "null pointer");
  GRID_ARGS(B, D, H, W);
  k_macro<3><<<grid, kThreads, 0, st>>>(f, rho, vel, flags, rho_out, vel_out, density_out, consts_of(p), g);
  return check_launch(__func__);
}

int lbmff_equation_of_states(const float* rho, float* out, long long n, double dx, double dt, int order,
                             lbmff_stream_t stream) {
  LBMFF_REQUIRE(rho && out && n >= 0, "bad arguments");
  if (n == 0) return 0;
  const Consts k = make_consts(dx, dt);
  k_eos<<<blocks_for(n), kThreads, 0, (cudaStream_t)stream>>>(rho, out, n, k.cs2, k.a_eos, order);
  return check_launch(__func__);
}

// phi = -(2 (density - dg) / (dl - dg) - 1), density = dg + (dl - dg) ((rho - rg) / (rl - rg)): the nine elementwise torch
// kernels of the demos' order-parameter line (demo_3d_LBM_Rosensweig_instability.py:162 on top of the affine density
// map, HCZ_3d.py:213-215) in ONE pass, each operation rounded separately and the two divisions by Python scalars taken
// as multiplications by the reciprocal (formed in double, rounded to fp32) -- which is how torch evaluates
// tensor / python_scalar on CUDA (tools/probe/torch_scalar_ops_probe.py, phi_probe.py) -- so the result
// equals the torch expression bit for bit (tests/test_gpu_ops.py).
struct PhiConsts {
  float rg, inv_rdiff, ddiff, dg, inv_ddiff;
};
__global__ void __launch_bounds__(kThreads) k_order_parameter(const float* __restrict__ rho, float* __restrict__ out,
                                                              long long n, PhiConsts c) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float t = __fmul_rn(__fsub_rn(rho[i], c.rg), c.inv_rdiff);
    const float density = __fadd_rn(__fmul_rn(c.ddiff, t), c.dg);
    const float u = __fmul_rn(__fmul_rn(2.0f, __fsub_rn(density, c.dg)), c.inv_ddiff);
    out[i] = -__fsub_rn(u, 1.0f);
  }
}

int lbmff_order_parameter(const float* rho, double rho_gas, double rho_liquid, double density_gas,
                          double density_liquid, float* phi_out, long long n, lbmff_stream_t stream) {
  LBMFF_REQUIRE(rho && phi_out && n >= 0, "bad arguments");
  LBMFF_REQUIRE(rho_liquid != rho_gas && density_liquid != density_gas, "liquid and gas values must differ");
  if (n == 0) return 0;
  PhiConsts c;
  c.rg = (float)rho_gas;
  c.inv_rdiff = (float)(1.0 / (rho_liquid - rho_gas));   // torch forms the reciprocal of a Python scalar in double
  c.ddiff = (float)(density_liquid - density_gas);
  c.dg = (float)density_gas;
  c.inv_ddiff = (float)(1.0 / (density_liquid - density_gas));
  const unsigned nb = blocks_for(n) > 16384u ? 16384u : blocks_for(n);
  k_order_parameter<<<nb, kThreads, 0, (cudaStream_t)stream>>>(rho, phi_out, n, c);
  return check_launch(__func__);
}

int lbmff_smooth_phi(const float* phi, float* out, long long n, double eps, lbmff_stream_t stream) {
  LBMFF_REQUIRE(phi && out && n >= 0 && eps > 0, "bad arguments");
  if (n == 0) return 0;
  Smooth S;
  S.eps = (float)eps; S.half_inv = (float)(0.5 / eps); S.pi_inv = (float)(3.141592653589793 / eps);
  k_smooth<<<blocks_for(n), kThreads, 0, (cudaStream_t)stream>>>(phi, out, n, S);
  return check_launch(__func__);
}

int lbmff_get_feq_d2q9(const float* rho, const float* vel, const float* force, double tau, double dx, double dt,
                       float* feq, int B, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(rho && vel && feq, "null pointer");
  GRID_ARGS(B, 1, H, W);
  k_equilibrium<2><<<grid, kThreads, 0, st>>>(rho, vel, force, nullptr, nullptr, nullptr, feq, 0,
                                              make_consts(dx, dt, tau), g);
  return check_launch(__func__);
}
int lbmff_get_feq_d3q19(const float* rho, const float* vel, const float* force, double tau, double dx, double dt,
                        float* feq, int B, int D, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(rho && vel && feq, "null pointer");
  GRID_ARGS(B, D, H, W);
  k_equilibrium<3><<<grid, kThreads, 0, st>>>(rho, vel, force, nullptr, nullptr, nullptr, feq, 0,
                                              make_consts(dx, dt, tau), g);
  return check_launch(__func__);
}
int lbmff_get_geq_d2q9(const float* rho, const float* vel, const float* density, const float* pressure,
                       const float* force, const float* feq_in, double tau, double dx, double dt, float* geq, int B,
                       int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(rho && vel && density && pressure && geq, "null pointer");
  GRID_ARGS(B, 1, H, W);
  k_equilibrium<2><<<grid, kThreads, 0, st>>>(rho, vel, force, density, pressure, feq_in, geq, 1,
                                              make_consts(dx, dt, tau), g);
  return check_launch(__func__);
}
int lbmff_get_geq_d3q19(const float* rho, const float* vel, const float* density, const float* pressure,
                        const float* force, const float* feq_in, double tau, double dx, double dt, float* geq, int B,
                        int D, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(rho && vel && density && pressure && geq, "null pointer");
  GRID_ARGS(B, D, H, W);
  k_equilibrium<3><<<grid, kThreads, 0, st>>>(rho, vel, force, density, pressure, feq_in, geq, 1,
                                              make_consts(dx, dt, tau), g);
  return check_launch(__func__);
}
int lbmff_compute_gamma_d2q9(const float* vel, double dx, double dt, float* gamma, int B, int H, int W,
                             lbmff_stream_t stream) {
  LBMFF_REQUIRE(vel && gamma, "null pointer");
  GRID_ARGS(B, 1, H, W);
  k_gamma<2><<<grid, kThreads, 0, st>>>(vel, gamma, make_consts(dx, dt), g);
  return check_launch(__func__);
}
int lbmff_compute_gamma_d3q19(const float* vel, double dx, double dt, float* gamma, int B, int D, int H, int W,
                              lbmff_stream_t stream) {
  LBMFF_REQUIRE(vel && gamma, "null pointer");
  GRID_ARGS(B, D, H, W);
  k_gamma<3><<<grid, kThreads, 0, st>>>(vel, gamma, make_consts(dx, dt), g);
  return check_launch(__func__);
}

int lbmff_get_grad_2d(const float* in, const uint8_t* flags, float* out, double dx, int B, int H, int W,
                      lbmff_stream_t stream) {
  LBMFF_REQUIRE(in && flags && out && H >= 3 && W >= 3, "bad arguments (needs H, W >= 3)");
  GRID_ARGS(B, 1, H, W);
  k_grad<2><<<grid, kThreads, 0, st>>>(in, flags, out, (float)dx, 1.0f, g);
  return check_launch(__func__);
}
int lbmff_get_grad_3d(const float* in, const uint8_t* flags, float* out, double dx, int B, int D, int H, int W,
                      lbmff_stream_t stream) {
  LBMFF_REQUIRE(in && flags && out && D >= 3 && H >= 3 && W >= 3, "bad arguments (needs D, H, W >= 3)");
  GRID_ARGS(B, D, H, W);
  k_grad<3><<<grid, kThreads, 0, st>>>(in, flags, out, (float)dx, 1.0f, g);
  return check_launch(__func__);
}
int lbmff_get_laplacian_2d(const float* in, float* out, double dx, int B, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(in && out && H >= 3 && W >= 3, "bad arguments (needs H, W >= 3)");
  GRID_ARGS(B, 1, H, W);
  k_laplacian<2><<<grid, kThreads, 0, st>>>(in, out, (float)(dx * dx), g);
  return check_launch(__func__);
}
int lbmff_get_laplacian_3d(const float* in, float* out, double dx, int B, int D, int H, int W,
                           lbmff_stream_t stream) {
  LBMFF_REQUIRE(in && out && D >= 3 && H >= 3 && W >= 3, "bad arguments (needs D, H, W >= 3)");
  GRID_ARGS(B, D, H, W);
  k_laplacian<3><<<grid, kThreads, 0, st>>>(in, out, (float)(dx * dx), g);
  return check_launch(__func__);
}

int lbmff_get_staggered(const float* in, float* out, int axis, int BC, int D, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(in && out && axis >= 0 && axis <= 2 && BC > 0, "bad arguments");
  const int n = axis == 0 ? W : (axis == 1 ? H : D);
  LBMFF_REQUIRE(n >= 2, "axis extent must be >= 2");
  const long long No = (long long)(D + (axis == 2)) * (H + (axis == 1)) * (W + (axis == 0));
  const dim3 grid(blocks_for(No), (unsigned)BC);
  k_staggered<<<grid, kThreads, 0, (cudaStream_t)stream>>>(in, out, axis, D, H, W);
  return check_launch(__func__);
}

int lbmff_get_vort_2d(const float* vel, float* out, double dx, int B, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(vel && out && H >= 3 && W >= 3, "bad arguments");
  GRID_ARGS(B, 1, H, W);
  k_vort<<<grid, kThreads, 0, st>>>(vel, out, (float)(2.0 * dx), g);
  return check_launch(__func__);
}

}  // extern "C"
