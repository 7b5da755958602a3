// Converged alternative to the 30 relaxation sweeps of the magnetic potential (north_star (3)).
//
// The reference never converges its Poisson lattice: it runs 30 sweeps per outer step, warm-started
// (demo_3d_LBM_Rosensweig_instability.py:163-173).  The FIXED POINT of those sweeps is the solution of a linear
// system.  At tau = 1 (the reference's value, LBM_magnetic_3d.py:12-20; the only one this entry accepts -- for
// tau != 1 the stationary populations keep a non-equilibrium part and the fixed point is not a compact stencil) a
// post-collision population is w_q * A with A = psi + S, S = dt*rhs*cs2*(1/2 - tau)*dt (LBM_magnetic_3d.py:180-181),
// a population that comes back from a wall is the cell's own (full-way bounce-back), and a non-wall cell obeys
//
//     psi(x) = 1/(1 - w0) * sum_{q >= 1} w_q * ( wall(x - e_q) ? A(x) : A(x - e_q) )          (periodic indices)
//
// i.e.  L psi = b  with the weighted graph Laplacian  (L psi)(x) = 1/(1-w0) * sum_{q: fluid nbr} w_q (psi(x) - psi(nbr))
// and b(x) = 1/(1-w0) * sum_q w_q S(nbr or x) .  L is symmetric positive semi-definite with the constants in its
// null space (only grad(psi) is ever used: H_int = -grad(psi)).  When sum(b) != 0 the sweeps have no fixed point but
// a stationary state that gains a constant c per sweep, psi_n = psi* + n c; a population returned by a wall is one
// sweep older than one received from a neighbour, so psi* solves  L psi* = b - c (1 + omega),  omega(x) =
// sum_{q: wall nbr} w_q / (1-w0),  with c = sum b / sum (1 + omega) (tests/test_poisson_math.py checks this against
// the oracle's sweeps).  That compatible system is solved by Jacobi-preconditioned conjugate gradients.  Stencil kernels on the division-free plane geometry
// (common.cuh); every dot product / residual norm is a warp-shuffle reduction to one partial per block followed by a
// fixed-order single-block finish in double precision, so the iteration is deterministic and never leaves the device:
// alpha, beta and the residual history live in device memory.
//
// This path has NO counterpart in the reference and changes results (a converged potential instead of 30 sweeps):
// it is opt-in (FerroSolver(magnetic_path="converged")) and validated against the sweeps run to convergence.
#include "capi.cuh"
#include "magnetic.cuh"

namespace lbmff {

// device scalars of the iteration
struct CgScalars {
  double iters, done, rr, bb, mean;  // the five doubles `info` exposes (include/lbmff.h); mean = drift c
  double rz, pAp, alpha, beta;
};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}

// one partial per block: warp shuffles, then the first warp adds the per-warp sums
template <int NV>
__device__ __forceinline__ void block_partials(const double (&v)[NV], double* __restrict__ partials, int nblocks) {
  __shared__ double sm[NV][8];
  const int tid = threadIdx.y * blockDim.x + threadIdx.x, lane = tid & 31, warp = tid >> 5;
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    const double s = warp_sum(v[k]);
    if (lane == 0) sm[k][warp] = s;
  }
  __syncthreads();
  if (warp == 0) {
    const int nw = (blockDim.x * blockDim.y + 31) >> 5;
    const int bid = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      double s = lane < nw ? sm[k][lane] : 0.0;
      s = warp_sum(s);
      if (lane == 0) partials[(size_t)k * nblocks + bid] = s;
    }
  }
}

// fixed-order finish of the 3 partial arrays by ONE block, then the scalar update selected by `what`
//   0: drift c = v0 / v1 (v0 = sum b, v1 = sum (1 + omega))    1: rz = v0, rr = v1, bb = |b - c(1+omega)|^2 = v2 (start)
//   2: pAp = v0 -> alpha                                          3: rz_new = v0, rr = v1 -> beta, history, stop test
__global__ void __launch_bounds__(1024) k_cg_finish(const double* __restrict__ partials, int nblocks, int what,
                                                    CgScalars* __restrict__ sc, float* __restrict__ hist,
                                                    double tol2) {
  __shared__ double sm[3][32];
  double a = 0.0, b = 0.0, c = 0.0;
  for (int i = threadIdx.x; i < nblocks; i += blockDim.x) {
    a += partials[i];
    b += partials[(size_t)nblocks + i];
    c += partials[2 * (size_t)nblocks + i];
  }
  a = warp_sum(a);
  b = warp_sum(b);
  c = warp_sum(c);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { sm[0][warp] = a; sm[1][warp] = b; sm[2][warp] = c; }
  __syncthreads();
  if (warp != 0) return;
  a = warp_sum(sm[0][lane]);
  b = warp_sum(sm[1][lane]);
  c = warp_sum(sm[2][lane]);
  if (lane != 0) return;
  if (what == 0) {
    sc->mean = b > 0.0 ? a / b : 0.0;
  } else if (what == 1) {
    sc->rz = a; sc->rr = b; sc->bb = c;
    sc->iters = 0.0;
    if (hist) hist[0] = (float)sqrt(b);
    sc->done = (b <= tol2 * c) ? 1.0 : 0.0;  // the warm start may already be good enough
  } else if (what == 2) {
    if (sc->done != 0.0) return;
    sc->pAp = a;
    sc->alpha = a > 0.0 ? sc->rz / a : 0.0;
  } else {
    if (sc->done != 0.0) return;
    sc->rr = b;
    sc->beta = sc->rz > 0.0 ? a / sc->rz : 0.0;
    sc->rz = a;
    sc->iters += 1.0;
    if (hist) hist[(int)sc->iters] = (float)sqrt(b);
    if (b <= tol2 * sc->bb) sc->done = 1.0;
  }
}

// weights of the wall-aware stencil at a non-wall cell: sum over fluid neighbours of w_q v(nbr), and their weight sum
template <int DIM>
__device__ __forceinline__ void nbr_sums(const float* __restrict__ v, const uint8_t* __restrict__ fl, int i,
                                         const Nbr& nb, float& sum_wv, float& sum_w) {
  using L = Lat<DIM>;
  sum_wv = 0.f; sum_w = 0.f;
#pragma unroll
  for (int q = 1; q < L::Q; ++q) {
    const int j = i + nb.at<DIM>(q, 1);
    const float x = __ldg(v + j);
    if (fl[j] != LBMFF_OBSTACLE) { sum_wv += L::w(q) * x; sum_w += L::w(q); }
  }
}

// b = 1/(1-w0) sum_q w_q S(nbr or self), S = dt*rhs*mag_src; partials: sum b, sum (1 + omega) (fluid cells)
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_cg_rhs(const float* __restrict__ rhs, const uint8_t* __restrict__ fl,
                                                     float* __restrict__ b, double* __restrict__ partials,
                                                     int nblocks, Consts k, Grid g) {
  using L = Lat<DIM>;
  const PlaneCell c = plane_cell(g);
  double acc[3] = {0.0, 0.0, 0.0};
  if (c.live) {
    float out = 0.f;
    if (fl[c.i] != LBMFF_OBSTACLE) {
      const Nbr nb(c, g);
      float swv, sw;
      nbr_sums<DIM>(rhs, fl, c.i, nb, swv, sw);
      const float wall_w = (1.0f - L::w(0)) - sw;
      out = (k.dt * k.mag_src) * (swv + wall_w * __ldg(rhs + c.i)) / (1.0f - L::w(0));
      acc[0] = out; acc[1] = 1.0 + (double)(wall_w / (1.0f - L::w(0)));
    }
    b[c.i] = out;
  }
  block_partials<3>(acc, partials, nblocks);
}

// r = (b - c(1+omega)) - L psi on fluid cells; z = r / diag; p = z; partials: r.z, r.r, |b - c(1+omega)|^2
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_cg_start(const float* __restrict__ b, const float* __restrict__ psi,
                                                       const uint8_t* __restrict__ fl, float* __restrict__ r,
                                                       float* __restrict__ p, double* __restrict__ partials,
                                                       int nblocks, const CgScalars* __restrict__ sc, Grid g) {
  using L = Lat<DIM>;
  const PlaneCell c = plane_cell(g);
  double acc[3] = {0.0, 0.0, 0.0};
  if (c.live) {
    float rv = 0.f, zv = 0.f;
    if (fl[c.i] != LBMFF_OBSTACLE) {
      const Nbr nb(c, g);
      float swv, sw;
      nbr_sums<DIM>(psi, fl, c.i, nb, swv, sw);
      const float inv = 1.0f / (1.0f - L::w(0));
      const float Lpsi = (sw * __ldg(psi + c.i) - swv) * inv;
      const float diag = sw * inv;
      const float bz = b[c.i] - (float)sc->mean * (2.0f - diag);  // 1 + omega = 2 - diag
      rv = bz - Lpsi;
      zv = diag > 0.f ? rv / diag : 0.f;
      acc[0] = (double)rv * zv; acc[1] = (double)rv * rv; acc[2] = (double)bz * bz;
    }
    r[c.i] = rv; p[c.i] = zv;
  }
  block_partials<3>(acc, partials, nblocks);
}

// Ap = L p; partial: p.Ap
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_cg_apply(const float* __restrict__ p, const uint8_t* __restrict__ fl,
                                                       float* __restrict__ Ap, double* __restrict__ partials,
                                                       int nblocks, const CgScalars* __restrict__ sc, Grid g) {
  using L = Lat<DIM>;
  if (sc->done != 0.0) return;
  const PlaneCell c = plane_cell(g);
  double acc[3] = {0.0, 0.0, 0.0};
  if (c.live) {
    float out = 0.f;
    if (fl[c.i] != LBMFF_OBSTACLE) {
      const Nbr nb(c, g);
      float swv, sw;
      nbr_sums<DIM>(p, fl, c.i, nb, swv, sw);
      const float pc = __ldg(p + c.i);
      out = (sw * pc - swv) / (1.0f - L::w(0));
      acc[0] = (double)pc * out;
    }
    Ap[c.i] = out;
  }
  block_partials<3>(acc, partials, nblocks);
}

// psi += alpha p; r -= alpha Ap; z = r / diag (stored in Ap); partials: r.z, r.r
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_cg_update(float* __restrict__ psi, float* __restrict__ r,
                                                        const float* __restrict__ p, float* __restrict__ Ap,
                                                        const uint8_t* __restrict__ fl, double* __restrict__ partials,
                                                        int nblocks, const CgScalars* __restrict__ sc, Grid g) {
  using L = Lat<DIM>;
  if (sc->done != 0.0) return;
  const PlaneCell c = plane_cell(g);
  double acc[3] = {0.0, 0.0, 0.0};
  if (c.live && fl[c.i] != LBMFF_OBSTACLE) {
    const float alpha = (float)sc->alpha;
    psi[c.i] += alpha * p[c.i];
    const float rv = r[c.i] - alpha * Ap[c.i];
    r[c.i] = rv;
    const Nbr nb(c, g);
    float sw = 0.f;
#pragma unroll
    for (int q = 1; q < L::Q; ++q)
      if (fl[c.i + nb.at<DIM>(q, 1)] != LBMFF_OBSTACLE) sw += L::w(q);
    const float diag = sw / (1.0f - L::w(0));
    const float zv = diag > 0.f ? rv / diag : 0.f;
    Ap[c.i] = zv;
    acc[0] = (double)rv * zv; acc[1] = (double)rv * rv;
  }
  block_partials<3>(acc, partials, nblocks);
}

// p = z + beta p
__global__ void __launch_bounds__(kThreads) k_cg_direction(float* __restrict__ p, const float* __restrict__ z,
                                                           const uint8_t* __restrict__ fl,
                                                           const CgScalars* __restrict__ sc, Grid g) {
  if (sc->done != 0.0) return;
  const PlaneCell c = plane_cell(g);
  if (c.live && fl[c.i] != LBMFF_OBSTACLE) p[c.i] = z[c.i] + (float)sc->beta * p[c.i];
}


// ---- host side ---------------------------------------------------------------------------------------------
static int cg_blocks(const PlaneLaunch& pl) { return (int)(pl.grid.x * pl.grid.y * pl.grid.z); }
static long long cg_scratch_doubles(const PlaneLaunch& pl) { return 3LL * cg_blocks(pl) + 16; }

template <int DIM>
static int poisson_cg_impl(const float* rhs, const uint8_t* mflags, float* psi, float* work, double* scratch,
                           int max_iters, double tol, double tau, double dx, double dt, float* res_hist, double* info,
                           const Grid& g, cudaStream_t st, const char* where) {
  if (!(rhs && mflags && psi && work && scratch)) return bad_arg(where, "null pointer");
  if (max_iters < 0) return bad_arg(where, "max_iters must be >= 0");
  if (!(tol >= 0)) return bad_arg(where, "tol must be >= 0");
  if (tau != 1.0) return bad_arg(where, "the converged solve is the fixed point of the tau == 1 sweeps only");
  if (g.N > 0x7fffffffLL) return bad_arg(where, "more than 2^31-1 cells per batch item");
  const Consts k = make_consts(dx, dt, tau);
  const PlaneLaunch pl = plane_launch(g, 64);
  const int nb = cg_blocks(pl);
  const long long stride = cg_scratch_doubles(pl);
  for (long long b = 0; b < g.B; ++b) {
    const float* rb = rhs + b * g.N;
    const uint8_t* fl = mflags + b * g.N;
    float* x = psi + b * g.N;
    float* bv = work + (4 * b + 0) * g.N;
    float* r = work + (4 * b + 1) * g.N;
    float* pv = work + (4 * b + 2) * g.N;
    float* Ap = work + (4 * b + 3) * g.N;
    CgScalars* sc = reinterpret_cast<CgScalars*>(scratch + b * stride);
    double* part = scratch + b * stride + 16;
    float* hist = res_hist ? res_hist + b * (max_iters + 1LL) : nullptr;
    k_cg_rhs<DIM><<<pl.grid, pl.block, 0, st>>>(rb, fl, bv, part, nb, k, g);
    k_cg_finish<<<1, 1024, 0, st>>>(part, nb, 0, sc, nullptr, 0.0);
    k_cg_start<DIM><<<pl.grid, pl.block, 0, st>>>(bv, x, fl, r, pv, part, nb, sc, g);
    k_cg_finish<<<1, 1024, 0, st>>>(part, nb, 1, sc, hist, tol * tol);
    for (int it = 0; it < max_iters; ++it) {
      k_cg_apply<DIM><<<pl.grid, pl.block, 0, st>>>(pv, fl, Ap, part, nb, sc, g);
      k_cg_finish<<<1, 1024, 0, st>>>(part, nb, 2, sc, nullptr, 0.0);
      k_cg_update<DIM><<<pl.grid, pl.block, 0, st>>>(x, r, pv, Ap, fl, part, nb, sc, g);
      k_cg_finish<<<1, 1024, 0, st>>>(part, nb, 3, sc, hist, tol * tol);
      k_cg_direction<<<pl.grid, pl.block, 0, st>>>(pv, Ap, fl, sc, g);
    }
    if (info) cudaMemcpyAsync(info + 5 * b, sc, 5 * sizeof(double), cudaMemcpyDeviceToDevice, st);
  }
  return check_launch(where);
}

}  // namespace lbmff

using namespace lbmff;

extern "C" {

int lbmff_poisson_cg_scratch_doubles(int D, int H, int W) {
  if (D < 1 || H < 1 || W < 1) return -1;
  const long long n = cg_scratch_doubles(plane_launch(make_grid(1, D, H, W), 64));
  return n > 0x7fffffffLL ? -1 : (int)n;
}

int lbmff_poisson_cg_2d(const float* rhs, const uint8_t* mflags, float* psi, float* work, double* scratch,
                        int max_iters, double tol, double tau, double dx, double dt, float* res_hist, double* info,
                        int B, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(B >= 1 && H >= 1 && W >= 1, "non-positive extent");
  return poisson_cg_impl<2>(rhs, mflags, psi, work, scratch, max_iters, tol, tau, dx, dt, res_hist, info,
                            make_grid(B, 1, H, W), (cudaStream_t)stream, __func__);
}

int lbmff_poisson_cg_3d(const float* rhs, const uint8_t* mflags, float* psi, float* work, double* scratch,
                        int max_iters, double tol, double tau, double dx, double dt, float* res_hist, double* info,
                        int B, int D, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(B >= 1 && D >= 1 && H >= 1 && W >= 1, "non-positive extent");
  return poisson_cg_impl<3>(rhs, mflags, psi, work, scratch, max_iters, tol, tau, dx, dt, res_hist, info,
                            make_grid(B, D, H, W), (cudaStream_t)stream, __func__);
}

}  // extern "C"
