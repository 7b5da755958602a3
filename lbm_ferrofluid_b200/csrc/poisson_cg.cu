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
// the oracle's sweeps).  That compatible system is solved by conjugate gradients (no preconditioner: the diagonal of
// L is 1 on every cell without a wall neighbour).  Stencil kernels on the division-free plane geometry
// (common.cuh); every dot product / residual norm is a warp-shuffle reduction to one partial per block followed by a
// fixed-order single-block finish in double precision, so the iteration is deterministic and never leaves the device:
// alpha, beta and the residual history live in device memory, and the iteration itself is a CUDA-graph WHILE node
// whose condition the reduction kernel sets (cudaGraphSetConditional): one graph launch per solve, no launches
// after convergence, no host synchronisation.
//
// This path has NO counterpart in the reference and changes results (a converged potential instead of 30 sweeps):
// it is opt-in (FerroSolver(magnetic_path="converged")) and validated against the sweeps run to convergence.
#include "capi.cuh"
#include "magnetic.cuh"
#include "poisson_mg.cuh"

#include <stdlib.h>

namespace lbmff {

// device scalars of the iteration
struct CgScalars {
  double iters, done, rr, bb, mean;  // the five doubles `info` exposes (include/lbmff.h); mean = drift c
  double rz, pAp, alpha, beta;
};

constexpr unsigned kWallSelf = 1u;  // bit 0 of a wall mask: the cell itself is a wall; bit q: neighbour x - e_q is

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
// USE_GRAPH: what == 3 also decides whether the WHILE node runs its body again.
template <bool USE_GRAPH>
__global__ void __launch_bounds__(1024) k_cg_finish(const double* __restrict__ partials, int nblocks, int what,
                                                    CgScalars* __restrict__ sc, float* __restrict__ hist,
                                                    double tol2, int max_iters, cudaGraphConditionalHandle handle) {
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
    if (sc->done == 0.0) {
      sc->rr = b;
      sc->beta = sc->rz > 0.0 ? a / sc->rz : 0.0;
      sc->rz = a;
      sc->iters += 1.0;
      if (hist) hist[(int)sc->iters] = (float)sqrt(b);
      if (b <= tol2 * sc->bb) sc->done = 1.0;
    }
    if (USE_GRAPH) cudaGraphSetConditional(handle, (sc->done == 0.0 && (int)sc->iters < max_iters) ? 1u : 0u);
  }
}

// diagonal of L from a wall mask: (sum of the weights of the non-wall neighbours) / (1 - w0)
template <int DIM>
__device__ __forceinline__ float diag_of(unsigned m) {
  using L = Lat<DIM>;
  if (m == 0) return 1.0f;
  float sw = 0.f;
#pragma unroll
  for (int q = 1; q < L::Q; ++q)
    if (!((m >> q) & 1u)) sw += L::w(q);
  return sw / (1.0f - L::w(0));
}

// sum over the non-wall neighbours of w_q v(x - e_q); mask == 0 (no wall around) takes the flag-free path
template <int DIM>
__device__ __forceinline__ float nbr_sum(const float* __restrict__ v, int i, const Nbr& nb, unsigned m) {
  using L = Lat<DIM>;
  float s_axis = 0.f, s_diag = 0.f;
#pragma unroll
  for (int q = 1; q < L::Q; ++q) {
    const bool axis = (L::ex(q) != 0) + (L::ey(q) != 0) + (L::ez(q) != 0) == 1;
    const float x = (m != 0 && ((m >> q) & 1u)) ? 0.f : __ldg(v + i + nb.at<DIM>(q, 1));
    if (axis) s_axis += x; else s_diag += x;
  }
  return L::w(1) * s_axis + L::w(5) * s_diag;  // q = 1 / q = 5: the axis / diagonal weight of both lattices
}

// (L v)(x) (1 - w0) = sum over the non-wall neighbours of w_q (v(x) - v(x - e_q)).  The DIFFERENCES are formed
// first: neighbouring values of a smooth field are within a factor 2, so each subtraction is exact, and the
// rounding error of L v scales with |grad v| instead of |v| -- without this the fp32 residual stagnates near
// 1e-4 |b| at 512^3 (diag * v - sum w v loses log2(n^2) bits to cancellation).
template <int DIM>
__device__ __forceinline__ float laplace(const float* __restrict__ v, float vc, int i, const Nbr& nb, unsigned m) {
  using L = Lat<DIM>;
  float s_axis = 0.f, s_diag = 0.f;
#pragma unroll
  for (int q = 1; q < L::Q; ++q) {
    const bool axis = (L::ex(q) != 0) + (L::ey(q) != 0) + (L::ez(q) != 0) == 1;
    const float d = (m != 0 && ((m >> q) & 1u)) ? 0.f : vc - __ldg(v + i + nb.at<DIM>(q, 1));
    if (axis) s_axis += d; else s_diag += d;
  }
  return (L::w(1) * s_axis + L::w(5) * s_diag) / (1.0f - L::w(0));
}

// the z planes a block walks: blockIdx.z, blockIdx.z + gridDim.z, ...
#define LBMFF_CG_PLANES(c)                                               \
  for (int zz_ = blockIdx.z; zz_ < g.zn; zz_ += gridDim.z)               \
    if (PlaneCell c = plane_cell_at(g, g.zb + zz_); c.live)

__device__ __forceinline__ PlaneCell plane_cell_at(const Grid& g, int z) {
  PlaneCell c;
  c.x = blockIdx.x * blockDim.x + threadIdx.x;
  c.y = blockIdx.y * blockDim.y + threadIdx.y;
  c.z = z;
  c.live = c.x < g.W && c.y < g.H;
  c.i = (c.z * g.H + c.y) * g.W + c.x;
  return c;
}

// wall mask, b = 1/(1-w0) sum_q w_q S(nbr or self) with S = dt*rhs*mag_src; partials: sum b, sum (1 + omega)
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_cg_rhs(const float* __restrict__ rhs, const uint8_t* __restrict__ fl,
                                                     float* __restrict__ b, unsigned* __restrict__ mask,
                                                     double* __restrict__ partials, int nblocks, Consts k, Grid g) {
  using L = Lat<DIM>;
  double acc[3] = {0.0, 0.0, 0.0};
  LBMFF_CG_PLANES(c) {
    float out = 0.f;
    unsigned m = kWallSelf;
    if (fl[c.i] != LBMFF_OBSTACLE) {
      const Nbr nb(c, g);
      m = 0;
#pragma unroll
      for (int q = 1; q < L::Q; ++q)
        if (fl[c.i + nb.at<DIM>(q, 1)] == LBMFF_OBSTACLE) m |= 1u << q;
      const float diag = diag_of<DIM>(m);
      const float swv = nbr_sum<DIM>(rhs, c.i, nb, m) / (1.0f - L::w(0));
      out = (k.dt * k.mag_src) * (swv + (1.0f - diag) * __ldg(rhs + c.i));
      acc[0] += out;
      acc[1] += 2.0 - (double)diag;  // 1 + omega, omega = 1 - diag
    }
    b[c.i] = out;
    mask[c.i] = m;
  }
  block_partials<3>(acc, partials, nblocks);
}

// r = (b - c(1+omega)) - L psi on fluid cells, 0 on walls; p = r; partials: r.r (twice), |b - c(1+omega)|^2
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_cg_start(const float* __restrict__ b, const float* __restrict__ psi,
                                                       const unsigned* __restrict__ mask, float* __restrict__ r,
                                                       float* __restrict__ p, double* __restrict__ partials,
                                                       int nblocks, const CgScalars* __restrict__ sc, Grid g) {
  using L = Lat<DIM>;
  double acc[3] = {0.0, 0.0, 0.0};
  const float drift = (float)sc->mean;
  LBMFF_CG_PLANES(c) {
    float rv = 0.f;
    const unsigned m = mask[c.i];
    if (!(m & kWallSelf)) {
      const Nbr nb(c, g);
      const float diag = diag_of<DIM>(m);
      const float Lpsi = laplace<DIM>(psi, __ldg(psi + c.i), c.i, nb, m);
      const float bz = b[c.i] - drift * (2.0f - diag);
      rv = bz - Lpsi;
      acc[0] += (double)rv * rv; acc[1] += (double)rv * rv; acc[2] += (double)bz * bz;
    }
    r[c.i] = rv; p[c.i] = rv;
  }
  block_partials<3>(acc, partials, nblocks);
}

// Ap = L p; partial: p.Ap
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_cg_apply(const float* __restrict__ p, const unsigned* __restrict__ mask,
                                                       float* __restrict__ Ap, double* __restrict__ partials,
                                                       int nblocks, const CgScalars* __restrict__ sc, Grid g) {
  using L = Lat<DIM>;
  if (sc->done != 0.0) return;
  double acc[3] = {0.0, 0.0, 0.0};
  LBMFF_CG_PLANES(c) {
    float out = 0.f;
    const unsigned m = mask[c.i];
    if (!(m & kWallSelf)) {
      const Nbr nb(c, g);
      const float pc = __ldg(p + c.i);
      out = laplace<DIM>(p, pc, c.i, nb, m);
      acc[0] += (double)pc * out;
    }
    Ap[c.i] = out;
  }
  block_partials<3>(acc, partials, nblocks);
}

// Ap = L p in 3-D, shared-memory tiled and marching along z: a 32 x 16 column of cells (32 x 8 threads, two rows each)
// walks zc planes; every plane of p is staged once with its periodic rim (34 x 18) into a four-slot ring (planes z-1, z,
// z+1 readable while z+2 lands), two planes ahead in two register sets.  The 18 neighbour values of a cell are then
// shared-memory reads instead of L1/L2 round trips behind per-load address arithmetic.  The arithmetic is laplace()'s:
// differences first, accumulated in the same q order, so Ap is bit-identical to k_cg_apply's (only the block partition
// of the p.Ap partials differs).  Cells with a wall among their neighbours (mask != 0) take laplace() itself.
constexpr int kApTX = 32, kApTY = 8, kApR = 2, kApRing = 4;
constexpr int kApRows = kApTY * kApR, kApTW = kApTX + 2, kApTH = kApRows + 2, kApTile = kApTW * kApTH;
constexpr int kApStage = (kApTile + kApTX * kApTY - 1) / (kApTX * kApTY);

__device__ __noinline__ float cg_apply_wall_adjacent(const float* __restrict__ p, float pc, int x, int y, int z, int i,
                                                     unsigned m, const Grid& g) {
  PlaneCell c;
  c.x = x; c.y = y; c.z = z; c.i = i; c.live = true;
  const Nbr nb(c, g);
  return laplace<3>(p, pc, i, nb, m);
}

// MODE 0: Ap = L p, partial p.Ap (k_cg_apply).  MODE 1: the start of a solve (k_cg_start): `p` is psi, r = p_out =
// (b - c (1 + omega)) - L psi, partials r.r (twice) and |b - c (1 + omega)|^2; Ap is r, bvec is b, out2 is the direction p.
template <int MODE>
__global__ void __launch_bounds__(kApTX* kApTY, 3) k_cg_apply_tiled(const float* __restrict__ p,
                                                                   const unsigned* __restrict__ mask,
                                                                   float* __restrict__ Ap, double* __restrict__ partials,
                                                                   int nblocks, const CgScalars* __restrict__ sc,
                                                                   const __grid_constant__ Grid g, int zc,
                                                                   const float* __restrict__ bvec = nullptr,
                                                                   float* __restrict__ out2 = nullptr) {
  using L = Lat<3>;
  if (MODE == 0 && sc->done != 0.0) return;
  const float drift = MODE == 1 ? (float)sc->mean : 0.f;
  __shared__ float sP[kApRing][kApTile];
  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * kApTX + tx;
  const int x0 = blockIdx.x * kApTX, y0 = blockIdx.y * kApRows;
  const int zs = blockIdx.z * zc, ze = min(zs + zc, g.D);
  const int HW = g.H * g.W;
  int off[kApStage];
#pragma unroll
  for (int t = 0; t < kApStage; ++t) {
    const int e = tid + t * kApTX * kApTY;
    const int sy = e / kApTW, sx = e - sy * kApTW;
    int gx = x0 - 1 + sx, gy = y0 - 1 + sy;
    gx = gx < 0 ? gx + g.W : gx; gy = gy < 0 ? gy + g.H : gy;
    while (gx >= g.W) gx -= g.W;
    while (gy >= g.H) gy -= g.H;
    off[t] = e < kApTile ? gy * g.W + gx : -1;
  }
  float rv[2][kApStage];
  auto fetch = [&](int z, auto SET_) {
    constexpr int SET = decltype(SET_)::value;
    const int zz = z < 0 ? z + g.D : (z >= g.D ? z - g.D : z);
    const int zo = zz * HW;
#pragma unroll
    for (int t = 0; t < kApStage; ++t)
      if (t < kApStage - 1 || off[t] >= 0) rv[SET][t] = __ldg(p + zo + off[t]);
  };
  auto put_slot = [&](int slot, auto SET_) {
    constexpr int SET = decltype(SET_)::value;
#pragma unroll
    for (int t = 0; t < kApStage; ++t)
      if (t < kApStage - 1 || off[t] >= 0) sP[slot][tid + t * kApTX * kApTY] = rv[SET][t];
  };
  using Set0 = std::integral_constant<int, 0>;
  using Set1 = std::integral_constant<int, 1>;
  fetch(zs - 1, Set0{}); put_slot(0, Set0{});
  fetch(zs, Set0{}); put_slot(1, Set0{});
  fetch(zs + 1, Set1{});
  fetch(zs + 2, Set0{});
  const int x = x0 + tx;
  bool live[kApR];
  unsigned m_next[kApR];
#pragma unroll
  for (int r = 0; r < kApR; ++r) {
    const int y = y0 + ty + kApTY * r;
    live[r] = x < g.W && y < g.H;
    m_next[r] = live[r] ? mask[zs * HW + y * g.W + x] : kWallSelf;
  }
  int i0 = zs * HW + (y0 + ty) * g.W + x;
  const int rowstep = kApTY * g.W;
  double acc[3] = {0.0, 0.0, 0.0};
  auto plane = [&](int z, auto S_) {
    constexpr int S = decltype(S_)::value;
    using Set = std::integral_constant<int, (S + 1) & 1>;
    put_slot((S + 2) & (kApRing - 1), Set{});   // plane z + 1
    if (z + 2 < ze) fetch(z + 3, Set{});
    unsigned m[kApR];
#pragma unroll
    for (int r = 0; r < kApR; ++r) {
      m[r] = m_next[r];
      if (live[r] && z + 1 < ze) m_next[r] = mask[i0 + HW + r * rowstep];
    }
    __syncthreads();
    constexpr int sl[3] = {S & (kApRing - 1), (S + 1) & (kApRing - 1), (S + 2) & (kApRing - 1)};  // planes z-1, z, z+1
#pragma unroll
    for (int r = 0; r < kApR; ++r) {
      if (!live[r]) continue;
      const int i = i0 + r * rowstep;
      float out = 0.f;
      if (!(m[r] & kWallSelf)) {
        const int ec = (ty + kApTY * r + 1) * kApTW + (tx + 1);
        const float pc = sP[sl[1]][ec];
        if (m[r] == 0) {
          float s_axis = 0.f, s_diag = 0.f;
#pragma unroll
          for (int q = 1; q < L::Q; ++q) {
            const bool axis = (L::ex(q) != 0) + (L::ey(q) != 0) + (L::ez(q) != 0) == 1;
            const float d = pc - sP[sl[1 - L::ez(q)]][ec - L::ey(q) * kApTW - L::ex(q)];
            if (axis) s_axis += d; else s_diag += d;
          }
          out = (L::w(1) * s_axis + L::w(5) * s_diag) / (1.0f - L::w(0));
        } else {
          out = cg_apply_wall_adjacent(p, pc, x, y0 + ty + kApTY * r, z, i, m[r], g);
        }
        if (MODE == 0) {
          acc[0] += (double)pc * out;
        } else {
          const float bz = bvec[i] - drift * (2.0f - diag_of<3>(m[r]));
          out = bz - out;
          acc[0] += (double)out * out; acc[1] += (double)out * out; acc[2] += (double)bz * bz;
        }
      }
      Ap[i] = out;
      if (MODE == 1) out2[i] = out;
    }
    i0 += HW;
  };
  for (int z = zs; z < ze; z += kApRing) {
    plane(z, std::integral_constant<int, 0>{});
    if (z + 1 < ze) plane(z + 1, std::integral_constant<int, 1>{});
    if (z + 2 < ze) plane(z + 2, std::integral_constant<int, 2>{});
    if (z + 3 < ze) plane(z + 3, std::integral_constant<int, 3>{});
  }
  block_partials<3>(acc, partials, nblocks);
}

// The two vector updates are flat streaming kernels (no stencil, no flags: r, p and Ap are kept at exactly 0 on wall
// cells, so psi is untouched there), float4 when the item is 16-byte aligned, grid-stride over `nblocks` blocks.
// psi += alpha p; r -= alpha Ap; partials: r.r (twice: the unpreconditioned r.z is r.r)
template <bool VEC4>
__global__ void __launch_bounds__(kThreads) k_cg_update(float* __restrict__ psi, float* __restrict__ r,
                                                        const float* __restrict__ p, const float* __restrict__ Ap,
                                                        long long n, double* __restrict__ partials, int nblocks,
                                                        const CgScalars* __restrict__ sc) {
  if (sc->done != 0.0) return;
  double acc[3] = {0.0, 0.0, 0.0};
  const float alpha = (float)sc->alpha;
  const long long stride = (long long)gridDim.x * blockDim.x, t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (VEC4) {
    for (long long i = t0; i < n / 4; i += stride) {
      float4 x = reinterpret_cast<float4*>(psi)[i], rv = reinterpret_cast<float4*>(r)[i];
      const float4 pv = reinterpret_cast<const float4*>(p)[i], av = reinterpret_cast<const float4*>(Ap)[i];
      x.x += alpha * pv.x; x.y += alpha * pv.y; x.z += alpha * pv.z; x.w += alpha * pv.w;
      rv.x -= alpha * av.x; rv.y -= alpha * av.y; rv.z -= alpha * av.z; rv.w -= alpha * av.w;
      reinterpret_cast<float4*>(psi)[i] = x;
      reinterpret_cast<float4*>(r)[i] = rv;
      acc[0] += (double)rv.x * rv.x + (double)rv.y * rv.y + (double)rv.z * rv.z + (double)rv.w * rv.w;
    }
  } else {
    for (long long i = t0; i < n; i += stride) {
      psi[i] += alpha * p[i];
      const float rv = r[i] - alpha * Ap[i];
      r[i] = rv;
      acc[0] += (double)rv * rv;
    }
  }
  acc[1] = acc[0];
  block_partials<3>(acc, partials, nblocks);
}

// p = r + beta p
template <bool VEC4>
__global__ void __launch_bounds__(kThreads) k_cg_direction(float* __restrict__ p, const float* __restrict__ r,
                                                           long long n, const CgScalars* __restrict__ sc) {
  if (sc->done != 0.0) return;
  const float beta = (float)sc->beta;
  const long long stride = (long long)gridDim.x * blockDim.x, t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (VEC4) {
    for (long long i = t0; i < n / 4; i += stride) {
      float4 pv = reinterpret_cast<float4*>(p)[i];
      const float4 rv = reinterpret_cast<const float4*>(r)[i];
      pv.x = rv.x + beta * pv.x; pv.y = rv.y + beta * pv.y; pv.z = rv.z + beta * pv.z; pv.w = rv.w + beta * pv.w;
      reinterpret_cast<float4*>(p)[i] = pv;
    }
  } else {
    for (long long i = t0; i < n; i += stride) p[i] = r[i] + beta * p[i];
  }
}

// ---- host side ---------------------------------------------------------------------------------------------
// blocks walk z: at most kZGroups plane groups, so a reduction leaves <= a few thousand partials per array
constexpr int kZGroups = 8;
static PlaneLaunch cg_launch(const Grid& g) {
  PlaneLaunch pl = plane_launch(g, 64);
  if (pl.grid.z > (unsigned)kZGroups) pl.grid.z = kZGroups;
  return pl;
}
static int cg_blocks(const PlaneLaunch& pl) { return (int)(pl.grid.x * pl.grid.y * pl.grid.z); }
// launch of the operator kernel: k_cg_apply_tiled in 3-D (LBMFF_CG_TILED=0 keeps k_cg_apply), with its own block count
struct ApplyLaunch {
  bool tiled;
  dim3 grid, block;
  int zc, nb;
};
static ApplyLaunch apply_launch(const Grid& g, const PlaneLaunch& pl, bool allow_env = true) {
  ApplyLaunch a;
  a.tiled = false; a.grid = pl.grid; a.block = pl.block; a.zc = 0; a.nb = cg_blocks(pl);
  const char* e = allow_env ? getenv("LBMFF_CG_TILED") : nullptr;
  if (g.D > 1 && g.W >= 4 && g.H >= 4 && g.D >= 3 && !(e && e[0] == '0')) {
    a.tiled = true;
    a.zc = g.D >= 64 ? 32 : (g.D >= 16 ? 8 : g.D);
    a.block = dim3(kApTX, kApTY, 1);
    a.grid = dim3((unsigned)((g.W + kApTX - 1) / kApTX), (unsigned)((g.H + kApRows - 1) / kApRows),
                  (unsigned)((g.D + a.zc - 1) / a.zc));
    a.nb = (int)(a.grid.x * a.grid.y * a.grid.z);
  }
  return a;
}
static long long cg_scratch_doubles(const PlaneLaunch& pl, const Grid& g) {
  const int nt = apply_launch(g, pl, false).nb, np = cg_blocks(pl);
  return 3LL * (nt > np ? nt : np) + 16;
}

struct CgBuffers {
  float *b, *r, *p, *Ap, *psi;
  unsigned* mask;
  CgScalars* sc;
  double* part;
  float* hist;
};

static bool cg_vec4(const CgBuffers& B, const Grid& g) {
  const auto al = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
  return g.N % 4 == 0 && al(B.psi) && al(B.r) && al(B.p) && al(B.Ap);
}

// One CG iteration = 5 kernels.  With `graph` they become the body of a WHILE conditional node: k_cg_finish<true>
// (what = 3) re-arms the loop while the solve is neither converged nor at max_iters.
template <int DIM>
static cudaError_t iterate_plain(const CgBuffers& B, const PlaneLaunch& pl, int nb, int max_iters, double tol2,
                                 const Grid& g, cudaStream_t st) {
  const bool vec4 = cg_vec4(B, g);
  const ApplyLaunch al = apply_launch(g, pl);
  for (int it = 0; it < max_iters; ++it) {
    if (DIM == 3 && al.tiled)
      k_cg_apply_tiled<0><<<al.grid, al.block, 0, st>>>(B.p, B.mask, B.Ap, B.part, al.nb, B.sc, g, al.zc, nullptr, nullptr);
    else
      k_cg_apply<DIM><<<pl.grid, pl.block, 0, st>>>(B.p, B.mask, B.Ap, B.part, al.nb, B.sc, g);
    k_cg_finish<false><<<1, 1024, 0, st>>>(B.part, al.nb, 2, B.sc, nullptr, 0.0, max_iters, 0);
    if (vec4) k_cg_update<true><<<nb, kThreads, 0, st>>>(B.psi, B.r, B.p, B.Ap, g.N, B.part, nb, B.sc);
    else k_cg_update<false><<<nb, kThreads, 0, st>>>(B.psi, B.r, B.p, B.Ap, g.N, B.part, nb, B.sc);
    k_cg_finish<false><<<1, 1024, 0, st>>>(B.part, nb, 3, B.sc, B.hist, tol2, max_iters, 0);
    if (vec4) k_cg_direction<true><<<nb, kThreads, 0, st>>>(B.p, B.r, g.N, B.sc);
    else k_cg_direction<false><<<nb, kThreads, 0, st>>>(B.p, B.r, g.N, B.sc);
  }
  return cudaGetLastError();
}

template <int DIM>
static cudaError_t iterate_graph(const CgBuffers& B, const PlaneLaunch& pl, int nb, int max_iters, double tol2,
                                 const Grid& g, cudaStream_t st) {
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  cudaError_t e = cudaGraphCreate(&graph, 0);
  if (e != cudaSuccess) return e;
  auto fail = [&](cudaError_t err) {
    if (exec) cudaGraphExecDestroy(exec);
    if (graph) cudaGraphDestroy(graph);
    return err;
  };
  cudaGraphConditionalHandle handle;
  if ((e = cudaGraphConditionalHandleCreate(&handle, graph, 1, cudaGraphCondAssignDefault)) != cudaSuccess) return fail(e);
  cudaGraphNodeParams np = {};
  np.type = cudaGraphNodeTypeConditional;
  np.conditional.handle = handle;
  np.conditional.type = cudaGraphCondTypeWhile;
  np.conditional.size = 1;
  cudaGraphNode_t loop;
  if ((e = cudaGraphAddNode(&loop, graph, nullptr, 0, &np)) != cudaSuccess) return fail(e);
  cudaGraph_t body = np.conditional.phGraph_out[0];

  cudaGraphNode_t prev = nullptr;
  auto add = [&](void* fn, dim3 grid, dim3 block, void** args) {
    cudaKernelNodeParams kp = {};
    kp.func = fn; kp.gridDim = grid; kp.blockDim = block; kp.sharedMemBytes = 0; kp.kernelParams = args;
    cudaGraphNode_t n;
    const cudaError_t err = cudaGraphAddKernelNode(&n, body, prev ? &prev : nullptr, prev ? 1 : 0, &kp);
    prev = n;
    return err;
  };
  // argument storage must outlive the AddKernelNode calls only (values are copied into the node)
  int nbv = nb, what2 = 2, what3 = 3, mi = max_iters;
  double zero = 0.0, t2 = tol2;
  float* nohist = nullptr;
  Grid gv = g;
  CgBuffers b = B;
  const ApplyLaunch al = apply_launch(g, pl);
  int nba = al.nb, zcv = al.zc;
  void* a_apply[] = {&b.p, &b.mask, &b.Ap, &b.part, &nba, &b.sc, &gv};
  const float* nob = nullptr;
  float* noo = nullptr;
  void* a_apply_t[] = {&b.p, &b.mask, &b.Ap, &b.part, &nba, &b.sc, &gv, &zcv, &nob, &noo};
  void* a_fin2[] = {&b.part, &nba, &what2, &b.sc, &nohist, &zero, &mi, &handle};
  long long nv = g.N;
  const bool vec4 = cg_vec4(B, g);
  void* a_update[] = {&b.psi, &b.r, &b.p, &b.Ap, &nv, &b.part, &nbv, &b.sc};
  void* a_fin3[] = {&b.part, &nbv, &what3, &b.sc, &b.hist, &t2, &mi, &handle};
  void* a_dir[] = {&b.p, &b.r, &nv, &b.sc};
  if (DIM == 3 && al.tiled) {
    if ((e = add((void*)k_cg_apply_tiled<0>, al.grid, al.block, a_apply_t)) != cudaSuccess) return fail(e);
  } else if ((e = add((void*)k_cg_apply<DIM>, pl.grid, pl.block, a_apply)) != cudaSuccess) {
    return fail(e);
  }
  if ((e = add((void*)k_cg_finish<true>, dim3(1), dim3(1024), a_fin2)) != cudaSuccess) return fail(e);
  if ((e = add(vec4 ? (void*)k_cg_update<true> : (void*)k_cg_update<false>, dim3(nb), dim3(kThreads), a_update)) !=
      cudaSuccess)
    return fail(e);
  if ((e = add((void*)k_cg_finish<true>, dim3(1), dim3(1024), a_fin3)) != cudaSuccess) return fail(e);
  if ((e = add(vec4 ? (void*)k_cg_direction<true> : (void*)k_cg_direction<false>, dim3(nb), dim3(kThreads), a_dir)) !=
      cudaSuccess)
    return fail(e);
  if ((e = cudaGraphInstantiate(&exec, graph, 0)) != cudaSuccess) return fail(e);
  if ((e = cudaGraphLaunch(exec, st)) != cudaSuccess) return fail(e);
  // an in-flight launch keeps what it needs; the handles can go now
  cudaGraphExecDestroy(exec);
  cudaGraphDestroy(graph);
  return cudaSuccess;
}

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
  const PlaneLaunch pl = cg_launch(g);
  const int nb = cg_blocks(pl);
  const long long stride = cg_scratch_doubles(pl, g);
  const char* env = getenv("LBMFF_CG_NO_GRAPH");
  const bool use_graph = !(env && env[0] == '1');
  for (long long b = 0; b < g.B; ++b) {
    const float* rb = rhs + b * g.N;
    const uint8_t* fl = mflags + b * g.N;
    CgBuffers B;
    B.psi = psi + b * g.N;
    B.b = work + (5 * b + 0) * g.N;
    B.r = work + (5 * b + 1) * g.N;
    B.p = work + (5 * b + 2) * g.N;
    B.Ap = work + (5 * b + 3) * g.N;
    B.mask = reinterpret_cast<unsigned*>(work + (5 * b + 4) * g.N);
    B.sc = reinterpret_cast<CgScalars*>(scratch + b * stride);
    B.part = scratch + b * stride + 16;
    B.hist = res_hist ? res_hist + b * (max_iters + 1LL) : nullptr;
    k_cg_rhs<DIM><<<pl.grid, pl.block, 0, st>>>(rb, fl, B.b, B.mask, B.part, nb, k, g);
    k_cg_finish<false><<<1, 1024, 0, st>>>(B.part, nb, 0, B.sc, nullptr, 0.0, max_iters, 0);
    {
      const ApplyLaunch al = apply_launch(g, pl);
      if (DIM == 3 && al.tiled)
        k_cg_apply_tiled<1><<<al.grid, al.block, 0, st>>>(B.psi, B.mask, B.r, B.part, al.nb, B.sc, g, al.zc, B.b, B.p);
      else
        k_cg_start<DIM><<<pl.grid, pl.block, 0, st>>>(B.b, B.psi, B.mask, B.r, B.p, B.part, al.nb, B.sc, g);
      k_cg_finish<false><<<1, 1024, 0, st>>>(B.part, al.nb, 1, B.sc, B.hist, tol * tol, max_iters, 0);
    }
    if (max_iters > 0) {
      const cudaError_t e = use_graph ? iterate_graph<DIM>(B, pl, nb, max_iters, tol * tol, g, st)
                                      : iterate_plain<DIM>(B, pl, nb, max_iters, tol * tol, g, st);
      if (e != cudaSuccess) {
        set_last_error(where, e);
        return (int)e;
      }
    }
    if (info) cudaMemcpyAsync(info + 5 * b, B.sc, 5 * sizeof(double), cudaMemcpyDeviceToDevice, st);
  }
  return check_launch(where);
}

// ================================================================================================================
// Preconditioned conjugate gradients: the same system, M^-1 = one multigrid V-cycle (poisson_mg.cuh).
// ================================================================================================================

// level-0 smoother / residual on the mask-based operator.  MODE 0: xo = xi + omega / diag (b - L xi); 1: xo = b - L xi
template <int DIM, int MODE>
__global__ void __launch_bounds__(kThreads) k_mg0_sweep(const unsigned* __restrict__ mask, const float* __restrict__ b,
                                                        const float* __restrict__ xi, float* __restrict__ xo, Grid g) {
  LBMFF_CG_PLANES(c) {
    const unsigned m = mask[c.i];
    float out = 0.f;
    if (!(m & kWallSelf)) {
      const Nbr nb(c, g);
      const float xc = __ldg(xi + c.i);
      const float r = b[c.i] - laplace<DIM>(xi, xc, c.i, nb, m);
      const float d = diag_of<DIM>(m);
      out = MODE == 0 ? (d > 0.f ? fmaf(mg::kOmega / d, r, xc) : xc) : r;
    }
    xo[c.i] = out;
  }
}
// The same smoother / residual in 3-D, shared-memory tiled and marching along z (the level-0 passes are 60 % of a
// V-cycle).  A 32 x 32 column of cells (32 x 8 threads, four consecutive rows each) walks zc planes; every plane of
// xi is staged ONCE, with its one-cell periodic rim (34 x 34), into a two-slot ring.  From the staged plane a thread forms, for its own (x, y), the centre value, the
// sum A of the 4 in-plane axis neighbours and the sum C of the 4 in-plane corners, and keeps them in registers for
// the three live planes; then
//     sum over the 6 axis neighbours      = A(z) + centre(z-1) + centre(z+1)
//     sum over the 12 diagonal neighbours = C(z) + A(z-1) + A(z+1)
// i.e. 9 shared-memory reads and ~20 adds per cell instead of 19 global loads.  The sum form (diag * v - sum w v) loses
// the exact-difference property laplace() has, which matters for the CG residual at 512^3 but not for a
// preconditioner: M^-1 only has to be one fixed symmetric operator, and it is (every cell uses the same formula in
// every cycle).  Cells with a wall among their neighbours (mask != 0: the shell-adjacent planes / rows only) take the
// mask-aware global path of k_mg0_sweep; wall cells get 0.
constexpr int kMgTX = 32, kMgTY = 8;   // 32 x 8 threads; kMgR consecutive rows per thread (template parameter)

// the rare cell with a wall among its neighbours: mask-aware global path, kept out of line so that the marching loop
// keeps its registers
__device__ __noinline__ float mg0_wall_adjacent_cell(const float* __restrict__ xi, float xc, float bc, int x, int y,
                                                     int z, int i, unsigned m, const Grid& g, int mode) {
  PlaneCell c;
  c.x = x; c.y = y; c.z = z; c.i = i; c.live = true;
  const Nbr nb(c, g);
  const float res = bc - laplace<3>(xi, xc, i, nb, m);
  const float d = diag_of<3>(m);
  return mode == 0 ? (d > 0.f ? fmaf(mg::kOmega / d, res, xc) : xc) : res;
}

template <int MODE, int kMgR>
__global__ void __launch_bounds__(kMgTX* kMgTY, kMgR == 4 ? 2 : 3) k_mg0_sweep_tiled(const unsigned* __restrict__ mask,
                                                                    const float* __restrict__ b,
                                                                    const float* __restrict__ xi,
                                                                    float* __restrict__ xo,
                                                                    const __grid_constant__ Grid g, int zc) {
  using L = Lat<3>;
  constexpr int kMgRows = kMgTY * kMgR;                         // 32 x (8 kMgR) cells per block and plane
  constexpr int kMgTW = kMgTX + 2, kMgTH = kMgRows + 2;         // staged tile with its rim
  constexpr int kMgTile = kMgTW * kMgTH;
  constexpr int kMgStage = (kMgTile + kMgTX * kMgTY - 1) / (kMgTX * kMgTY);   // staged elements per thread
  __shared__ float sv[2][kMgTile];
  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * kMgTX + tx;
  const int x0 = blockIdx.x * kMgTX, y0 = blockIdx.y * kMgRows;
  const int zs = blockIdx.z * zc, ze = min(zs + zc, g.D);
  const int HW = g.H * g.W;
  // the (up to) kMgStage rim-tile elements this thread stages per plane: tile element e = tid + t * 256 lives at
  // sv[slot][e]; its in-plane global position is periodic in x and y.  Only the last one can fall outside the tile.
  // Addresses: one 64-bit pointer per staged element (plane 0), plus a block-uniform 32-bit plane offset.
  const float* __restrict__ src[kMgStage];
  bool last_ok = true;
#pragma unroll
  for (int t = 0; t < kMgStage; ++t) {
    const int e = tid + t * kMgTX * kMgTY;
    const int sy = e / kMgTW, sx = e - sy * kMgTW;
    int gx = x0 - 1 + sx, gy = y0 - 1 + sy;
    gx = gx < 0 ? gx + g.W : gx; gy = gy < 0 ? gy + g.H : gy;
    while (gx >= g.W) gx -= g.W;
    while (gy >= g.H) gy -= g.H;
    const bool ok = e < kMgTile;
    if (t == kMgStage - 1) last_ok = ok;
    src[t] = xi + (ok ? gy * g.W + gx : 0);
  }
  float rv[kMgStage];
  auto fetch = [&](int z) {   // plane z (periodic) -> registers
    const int zz = z < 0 ? z + g.D : (z >= g.D ? z - g.D : z);
    const int zo = zz * HW;
#pragma unroll
    for (int t = 0; t < kMgStage; ++t) rv[t] = __ldg(src[t] + zo);
  };
  auto put = [&](int slot) {  // registers -> ring slot
#pragma unroll
    for (int t = 0; t < kMgStage; ++t)
      if (t < kMgStage - 1 || last_ok) sv[slot][tid + t * kMgTX * kMgTY] = rv[t];
  };
  // per row r of this thread: centre value, A = sum of the 4 in-plane axis neighbours, C = sum of the 4 corners.
  // Row sums h(j) = left + right of tile row j are shared between the rows: 3 reads per tile row, kMgR + 2 tile rows.
  struct PlaneSums { float c[kMgR], A[kMgR], C[kMgR]; };
  auto sums = [&](int slot, PlaneSums& P) {
    const float* s = sv[slot] + (ty * kMgR) * kMgTW + tx;
    float h[kMgR + 2], m[kMgR + 2];
#pragma unroll
    for (int j = 0; j < kMgR + 2; ++j) {
      h[j] = s[j * kMgTW] + s[j * kMgTW + 2];
      m[j] = s[j * kMgTW + 1];
    }
#pragma unroll
    for (int r = 0; r < kMgR; ++r) {
      P.c[r] = m[r + 1];
      P.A[r] = h[r + 1] + (m[r] + m[r + 2]);
      P.C[r] = h[r] + h[r + 2];
    }
  };
  const int x = x0 + tx, yb = y0 + ty * kMgR;
  bool live[kMgR];
#pragma unroll
  for (int r = 0; r < kMgR; ++r) live[r] = x < g.W && yb + r < g.H;
  PlaneSums pm, pc, pp;
  fetch(zs - 1); put(0);
  __syncthreads();
  sums(0, pm);
  fetch(zs); put(1);
  __syncthreads();
  sums(1, pc);
  fetch(zs + 1);
  int i0 = zs * HW + yb * g.W + x;          // cell of row 0 in plane z (only dereferenced where live)
  unsigned m_next[kMgR];
  float b_next[kMgR];
#pragma unroll
  for (int r = 0; r < kMgR; ++r) {
    m_next[r] = live[r] ? mask[i0 + r * g.W] : kWallSelf;
    b_next[r] = live[r] ? b[i0 + r * g.W] : 0.f;
  }
  constexpr float wa = L::w(1) / (1.0f - L::w(0)), wd = L::w(5) / (1.0f - L::w(0));
  // plane z; SLOT = ring slot of plane z + 1 (compile-time: the z loop is unrolled by the ring size)
  auto plane = [&](int z, auto SLOT_) {
    constexpr int SLOT = decltype(SLOT_)::value;
    put(SLOT);
    if (z + 1 < ze) fetch(z + 2);
    unsigned m[kMgR];
    float bc[kMgR];
#pragma unroll
    for (int r = 0; r < kMgR; ++r) {
      m[r] = m_next[r];
      bc[r] = b_next[r];
      if (live[r] && z + 1 < ze) {
        m_next[r] = mask[i0 + HW + r * g.W];
        b_next[r] = b[i0 + HW + r * g.W];
      }
    }
    __syncthreads();
    sums(SLOT, pp);
#pragma unroll
    for (int r = 0; r < kMgR; ++r) {
      if (!live[r]) continue;
      const int i = i0 + r * g.W;
      float out = 0.f;
      if (m[r] == 0) {
        const float s_axis = pc.A[r] + (pm.c[r] + pp.c[r]), s_diag = pc.C[r] + (pm.A[r] + pp.A[r]);
        const float Lx = wa * (6.0f * pc.c[r] - s_axis) + wd * (12.0f * pc.c[r] - s_diag);
        const float res = bc[r] - Lx;
        out = MODE == 0 ? fmaf(mg::kOmega, res, pc.c[r]) : res;
      } else if (!(m[r] & kWallSelf)) {
        out = mg0_wall_adjacent_cell(xi, pc.c[r], bc[r], x, yb + r, z, i, m[r], g, MODE);
      }
      xo[i] = out;
    }
    pm = pc;
    pc = pp;
    i0 += HW;
  };
  for (int z = zs; z < ze; z += 2) {   // plane zs - 1 sits in slot 0, zs in slot 1, zs + 1 in slot 0, ...
    plane(z, std::integral_constant<int, 0>{});
    if (z + 1 < ze) plane(z + 1, std::integral_constant<int, 1>{});
  }
}
struct TiledLaunch {
  dim3 grid, block;
  int zc, rows;
};
static TiledLaunch mg0_tiled_launch(const Grid& g) {
  TiledLaunch t;
  const char* e = getenv("LBMFF_MG_ROWS");
  t.rows = (e && e[0] == '4') ? 4 : 2;
  t.zc = g.D >= 64 ? 32 : (g.D >= 16 ? 8 : g.D);
  if (const char* z = getenv("LBMFF_MG_ZC")) t.zc = atoi(z) > 0 ? atoi(z) : t.zc;
  t.block = dim3(kMgTX, kMgTY, 1);
  const int rows = kMgTY * t.rows;
  t.grid = dim3((unsigned)((g.W + kMgTX - 1) / kMgTX), (unsigned)((g.H + rows - 1) / rows),
                (unsigned)((g.D + t.zc - 1) / t.zc));
  return t;
}

template <int DIM>
__global__ void __launch_bounds__(kThreads) k_mg0_jacobi0(const unsigned* __restrict__ mask, const float* __restrict__ b,
                                                          float* __restrict__ x, Grid g) {
  LBMFF_CG_PLANES(c) {
    const unsigned m = mask[c.i];
    float out = 0.f;
    if (!(m & kWallSelf)) {
      const float d = diag_of<DIM>(m);
      out = d > 0.f ? (mg::kOmega / d) * b[c.i] : 0.f;
    }
    x[c.i] = out;
  }
}
// partials: r.z, r.r
__global__ void __launch_bounds__(kThreads) k_cg_dot2(const float* __restrict__ r, const float* __restrict__ z,
                                                      long long n, double* __restrict__ partials, int nblocks) {
  double acc[3] = {0.0, 0.0, 0.0};
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float rv = r[i];
    acc[0] += (double)rv * z[i];
    acc[1] += (double)rv * rv;
  }
  block_partials<3>(acc, partials, nblocks);
}
// what = 4 of the finish step: rz = v0 (start of PCG: p = z)
__global__ void k_pcg_set_rz(const double* __restrict__ partials, int nblocks, CgScalars* __restrict__ sc) {
  __shared__ double sm[32];
  double a = 0.0;
  for (int i = threadIdx.x; i < nblocks; i += blockDim.x) a += partials[i];
  a = warp_sum(a);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x < 32) {
    a = warp_sum(threadIdx.x < (blockDim.x >> 5) ? sm[threadIdx.x] : 0.0);
    if (threadIdx.x == 0) sc->rz = a;
  }
}
// p = z + beta p  (beta = 0 at the start)
__global__ void __launch_bounds__(kThreads) k_pcg_direction(float* __restrict__ p, const float* __restrict__ z,
                                                            long long n, const CgScalars* __restrict__ sc, int first) {
  if (sc->done != 0.0) return;
  const float beta = first ? 0.f : (float)sc->beta;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    p[i] = first ? z[i] : fmaf(beta, p[i], z[i]);
}

// Launches kernels directly, or records them as a chain of kernel nodes of a graph (the WHILE body).
struct Launcher {
  cudaStream_t st;
  cudaGraph_t body = nullptr;
  cudaGraphNode_t prev = nullptr;
  cudaError_t err = cudaSuccess;
  template <class... A>
  void operator()(const void* fn, dim3 grid, dim3 block, A... a) {
    if (err != cudaSuccess) return;
    void* args[] = {(void*)&a...};
    if (!body) {
      err = cudaLaunchKernel(fn, grid, block, args, 0, st);
      return;
    }
    cudaKernelNodeParams kp = {};
    kp.func = const_cast<void*>(fn); kp.gridDim = grid; kp.blockDim = block; kp.sharedMemBytes = 0; kp.kernelParams = args;
    cudaGraphNode_t n;
    err = cudaGraphAddKernelNode(&n, body, prev ? &prev : nullptr, prev ? 1 : 0, &kp);
    prev = n;
  }
};

static PlaneLaunch mg_launch(const Grid& g) {
  PlaneLaunch pl = plane_launch(g, 64);
  pl.grid.z = pl.grid.z > 16u ? 16u : pl.grid.z;
  return pl;
}
static unsigned flat_blocks(long long n) {
  const long long b = (n + kThreads - 1) / kThreads;
  return (unsigned)(b > 4096 ? 4096 : (b < 1 ? 1 : b));
}

// extents of the hierarchy below a D x H x W fine grid; returns the number of coarse levels
template <int DIM>
static int mg_levels(const Grid& g0, Grid* out) {
  int n = 0;
  Grid g = g0;
  while (n < mg::kMaxLevels - 1) {
    const int W = (g.W + 1) / 2, H = (g.H + 1) / 2, D = DIM == 3 ? (g.D + 1) / 2 : 1;
    if (W < 3 || H < 3 || (DIM == 3 && D < 3)) break;
    g = make_grid(1, D, H, W);
    out[n++] = g;
  }
  return n;
}
template <int DIM>
static long long mg_workspace_floats(const Grid& g0) {
  Grid lv[mg::kMaxLevels];
  const int n = mg_levels<DIM>(g0, lv);
  long long total = 0;
  for (int l = 0; l < n; ++l) total += (long long)(Lat<DIM>::Q - 1 + 4) * ((lv[l].N + 3) / 4 * 4);
  return total;
}
template <int DIM>
static int mg_carve(const Grid& g0, float* ws, mg::Level* L) {
  Grid lv[mg::kMaxLevels];
  const int n = mg_levels<DIM>(g0, lv);
  for (int l = 0; l < n; ++l) {
    const long long N = (lv[l].N + 3) / 4 * 4;
    L[l].g = lv[l];
    L[l].w = ws; ws += (Lat<DIM>::Q - 1) * N;
    L[l].invd = ws; ws += N;
    L[l].x = ws; ws += N;
    L[l].t = ws; ws += N;
    L[l].b = ws; ws += N;
  }
  return n;
}
// the Galerkin hierarchy depends on the flags only: built once per solve (cheap: sum over levels of N / 7 cells)
template <int DIM>
static void mg_setup(const unsigned* mask, const Grid& g0, mg::Level* L, int n, cudaStream_t st) {
  for (int l = 0; l < n; ++l) {
    const PlaneLaunch pl = mg_launch(L[l].g);
    if (l == 0)
      mg::k_mg_coarsen<DIM, mg::MaskLinks<DIM>><<<pl.grid, pl.block, 0, st>>>(mg::MaskLinks<DIM>{mask}, g0, L[0].w,
                                                                             L[0].invd, L[0].g);
    else
      mg::k_mg_coarsen<DIM, mg::ArrayLinks><<<pl.grid, pl.block, 0, st>>>(mg::ArrayLinks{L[l - 1].w}, L[l - 1].g,
                                                                         L[l].w, L[l].invd, L[l].g);
  }
}
static bool mg_tiled_enabled() {
  const char* e = getenv("LBMFF_MG_TILED");
  return !(e && e[0] == '0');
}
// z = M^-1 r: one V(2,2) cycle.  Level 0 works in (z, t0) with `res0` as residual scratch; result in z.
template <int DIM>
static void mg_vcycle(Launcher& K, const unsigned* mask, const Grid& g0, const float* r, float* z, float* t0,
                      float* res0, mg::Level* L, int n) {
  const PlaneLaunch p0 = cg_launch(g0);
  K((const void*)k_mg0_jacobi0<DIM>, p0.grid, p0.block, mask, r, t0, g0);
  // level-0 Jacobi sweep / residual: the tiled kernel in 3-D (LBMFF_MG_TILED=0 selects the generic one)
  const TiledLaunch tl = mg0_tiled_launch(g0);
  auto sweep0 = [&](int mode, const float* xi, float* xo) {
    if (DIM == 3 && mg_tiled_enabled() && g0.zb == 0 && g0.zn == g0.D) {
      const void* fn = tl.rows == 4 ? (mode == 0 ? (const void*)k_mg0_sweep_tiled<0, 4> : (const void*)k_mg0_sweep_tiled<1, 4>)
                                    : (mode == 0 ? (const void*)k_mg0_sweep_tiled<0, 2> : (const void*)k_mg0_sweep_tiled<1, 2>);
      K(fn, tl.grid, tl.block, mask, r, xi, xo, g0, tl.zc);
    } else {
      K(mode == 0 ? (const void*)k_mg0_sweep<DIM, 0> : (const void*)k_mg0_sweep<DIM, 1>, p0.grid, p0.block, mask, r,
        xi, xo, g0);
    }
  };
  sweep0(0, t0, z);
  if (n > 0) {
    sweep0(1, z, res0);
    {
      const PlaneLaunch pc = mg_launch(L[0].g);
      K((const void*)mg::k_mg_restrict<DIM>, pc.grid, pc.block, (const float*)res0, g0, L[0].b, L[0].g);
    }
    for (int l = 0; l < n; ++l) {   // down
      const Grid g = L[l].g;
      const PlaneLaunch pl = mg_launch(g);
      const float *w = L[l].w, *invd = L[l].invd, *b = L[l].b;
      if (l == n - 1) {             // coarsest: Jacobi from zero, ending in x
        K((const void*)mg::k_mg_jacobi0, dim3(flat_blocks(g.N)), dim3(kThreads), b, invd, L[l].t, g.N);
        for (int s = 1; s < mg::kCoarsestSweeps; ++s) {
          float* xi = (s & 1) ? L[l].t : L[l].x;
          float* xo = (s & 1) ? L[l].x : L[l].t;
          K((const void*)mg::k_mg_sweep<DIM, 0>, pl.grid, pl.block, w, invd, b, (const float*)xi, xo, g);
        }
        static_assert(mg::kCoarsestSweeps % 2 == 0, "the coarsest solve must end in x");
        break;
      }
      K((const void*)mg::k_mg_jacobi0, dim3(flat_blocks(g.N)), dim3(kThreads), b, invd, L[l].t, g.N);
      K((const void*)mg::k_mg_sweep<DIM, 0>, pl.grid, pl.block, w, invd, b, (const float*)L[l].t, L[l].x, g);
      K((const void*)mg::k_mg_sweep<DIM, 1>, pl.grid, pl.block, w, invd, b, (const float*)L[l].x, L[l].t, g);
      const PlaneLaunch pc = mg_launch(L[l + 1].g);
      K((const void*)mg::k_mg_restrict<DIM>, pc.grid, pc.block, (const float*)L[l].t, g, L[l + 1].b, L[l + 1].g);
    }
    for (int l = n - 2; l >= 0; --l) {   // up
      const Grid g = L[l].g;
      const PlaneLaunch pl = mg_launch(g);
      const float *w = L[l].w, *invd = L[l].invd, *b = L[l].b;
      K((const void*)mg::k_mg_prolong_add<DIM>, pl.grid, pl.block, L[l].x, g, (const float*)L[l + 1].x, L[l + 1].g,
        (const unsigned*)nullptr);
      K((const void*)mg::k_mg_sweep<DIM, 0>, pl.grid, pl.block, w, invd, b, (const float*)L[l].x, L[l].t, g);
      K((const void*)mg::k_mg_sweep<DIM, 0>, pl.grid, pl.block, w, invd, b, (const float*)L[l].t, L[l].x, g);
    }
    const PlaneLaunch pf = mg_launch(g0);
    K((const void*)mg::k_mg_prolong_add<DIM>, pf.grid, pf.block, z, g0, (const float*)L[0].x, L[0].g, mask);
  }
  sweep0(0, z, t0);
  sweep0(0, t0, z);
}

struct PcgExtra {
  float *z, *t;          // preconditioned residual, Jacobi scratch of the fine level
  mg::Level L[mg::kMaxLevels];
  int nlev;
};

// body of one PCG iteration, recorded into K (direct launches or graph nodes)
template <int DIM, bool GRAPH>
static void pcg_body(Launcher& K, const CgBuffers& B, PcgExtra& X, const PlaneLaunch& pl, int nb, int max_iters,
                     double tol2, const Grid& g, cudaGraphConditionalHandle handle) {
  const bool vec4 = cg_vec4(B, g);
  float* nohist = nullptr;
  const ApplyLaunch al = apply_launch(g, pl);
  if (DIM == 3 && al.tiled)
    K((const void*)k_cg_apply_tiled<0>, al.grid, al.block, (const float*)B.p, (const unsigned*)B.mask, B.Ap, B.part, al.nb,
      (const CgScalars*)B.sc, g, al.zc, (const float*)nullptr, (float*)nullptr);
  else
    K((const void*)k_cg_apply<DIM>, pl.grid, pl.block, (const float*)B.p, (const unsigned*)B.mask, B.Ap, B.part, al.nb,
      (const CgScalars*)B.sc, g);
  K((const void*)k_cg_finish<GRAPH>, dim3(1), dim3(1024), (const double*)B.part, al.nb, 2, B.sc, nohist, 0.0, max_iters,
    handle);
  K(vec4 ? (const void*)k_cg_update<true> : (const void*)k_cg_update<false>, dim3(nb), dim3(kThreads), B.psi, B.r,
    (const float*)B.p, (const float*)B.Ap, (long long)g.N, B.part, nb, (const CgScalars*)B.sc);
  mg_vcycle<DIM>(K, B.mask, g, B.r, X.z, X.t, B.Ap, X.L, X.nlev);
  K((const void*)k_cg_dot2, dim3(nb), dim3(kThreads), (const float*)B.r, (const float*)X.z, (long long)g.N, B.part, nb);
  K((const void*)k_cg_finish<GRAPH>, dim3(1), dim3(1024), (const double*)B.part, nb, 3, B.sc, B.hist, tol2, max_iters,
    handle);
  K((const void*)k_pcg_direction, dim3(flat_blocks(g.N)), dim3(kThreads), B.p, (const float*)X.z, (long long)g.N,
    (const CgScalars*)B.sc, 0);
}

template <int DIM>
static cudaError_t iterate_pcg(bool use_graph, const CgBuffers& B, PcgExtra& X, const PlaneLaunch& pl, int nb,
                               int max_iters, double tol2, const Grid& g, cudaStream_t st) {
  if (!use_graph) {
    Launcher K{st};
    for (int it = 0; it < max_iters && K.err == cudaSuccess; ++it)
      pcg_body<DIM, false>(K, B, X, pl, nb, max_iters, tol2, g, 0);
    return K.err != cudaSuccess ? K.err : cudaGetLastError();
  }
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  cudaError_t e = cudaGraphCreate(&graph, 0);
  if (e != cudaSuccess) return e;
  auto fail = [&](cudaError_t err) {
    if (exec) cudaGraphExecDestroy(exec);
    if (graph) cudaGraphDestroy(graph);
    return err;
  };
  cudaGraphConditionalHandle handle;
  if ((e = cudaGraphConditionalHandleCreate(&handle, graph, 1, cudaGraphCondAssignDefault)) != cudaSuccess) return fail(e);
  cudaGraphNodeParams np = {};
  np.type = cudaGraphNodeTypeConditional;
  np.conditional.handle = handle;
  np.conditional.type = cudaGraphCondTypeWhile;
  np.conditional.size = 1;
  cudaGraphNode_t loop;
  if ((e = cudaGraphAddNode(&loop, graph, nullptr, 0, &np)) != cudaSuccess) return fail(e);
  Launcher K{st};
  K.body = np.conditional.phGraph_out[0];
  pcg_body<DIM, true>(K, B, X, pl, nb, max_iters, tol2, g, handle);
  if (K.err != cudaSuccess) return fail(K.err);
  if ((e = cudaGraphInstantiate(&exec, graph, 0)) != cudaSuccess) return fail(e);
  if ((e = cudaGraphLaunch(exec, st)) != cudaSuccess) return fail(e);
  cudaGraphExecDestroy(exec);
  cudaGraphDestroy(graph);
  return cudaSuccess;
}

template <int DIM>
static int poisson_pcg_impl(const float* rhs, const uint8_t* mflags, float* psi, float* work, float* mg_work,
                            double* scratch, int max_iters, double tol, double tau, double dx, double dt,
                            float* res_hist, double* info, int reuse_hierarchy, const Grid& g, cudaStream_t st,
                            const char* where) {
  if (!(rhs && mflags && psi && work && mg_work && scratch)) return bad_arg(where, "null pointer");
  if (max_iters < 0) return bad_arg(where, "max_iters must be >= 0");
  if (!(tol >= 0)) return bad_arg(where, "tol must be >= 0");
  if (tau != 1.0) return bad_arg(where, "the converged solve is the fixed point of the tau == 1 sweeps only");
  if (g.N > 0x7fffffffLL) return bad_arg(where, "more than 2^31-1 cells per batch item");
  const Consts k = make_consts(dx, dt, tau);
  const PlaneLaunch pl = cg_launch(g);
  const int nb = cg_blocks(pl);
  const long long stride = cg_scratch_doubles(pl, g);
  const long long mgf = mg_workspace_floats<DIM>(g);
  const char* env = getenv("LBMFF_CG_NO_GRAPH");
  const bool use_graph = !(env && env[0] == '1');
  for (long long b = 0; b < g.B; ++b) {
    const float* rb = rhs + b * g.N;
    const uint8_t* fl = mflags + b * g.N;
    CgBuffers B;
    B.psi = psi + b * g.N;
    B.b = work + (7 * b + 0) * g.N;
    B.r = work + (7 * b + 1) * g.N;
    B.p = work + (7 * b + 2) * g.N;
    B.Ap = work + (7 * b + 3) * g.N;
    B.mask = reinterpret_cast<unsigned*>(work + (7 * b + 4) * g.N);
    PcgExtra X;
    X.z = work + (7 * b + 5) * g.N;
    X.t = work + (7 * b + 6) * g.N;
    X.nlev = mg_carve<DIM>(g, mg_work + b * mgf, X.L);
    B.sc = reinterpret_cast<CgScalars*>(scratch + b * stride);
    B.part = scratch + b * stride + 16;
    B.hist = res_hist ? res_hist + b * (max_iters + 1LL) : nullptr;
    k_cg_rhs<DIM><<<pl.grid, pl.block, 0, st>>>(rb, fl, B.b, B.mask, B.part, nb, k, g);
    k_cg_finish<false><<<1, 1024, 0, st>>>(B.part, nb, 0, B.sc, nullptr, 0.0, max_iters, 0);
    if (!reuse_hierarchy) mg_setup<DIM>(B.mask, g, X.L, X.nlev, st);   // depends on the flags only
    {
      const ApplyLaunch al = apply_launch(g, pl);
      if (DIM == 3 && al.tiled)
        k_cg_apply_tiled<1><<<al.grid, al.block, 0, st>>>(B.psi, B.mask, B.r, B.part, al.nb, B.sc, g, al.zc, B.b, B.p);
      else
        k_cg_start<DIM><<<pl.grid, pl.block, 0, st>>>(B.b, B.psi, B.mask, B.r, B.p, B.part, al.nb, B.sc, g);
      k_cg_finish<false><<<1, 1024, 0, st>>>(B.part, al.nb, 1, B.sc, B.hist, tol * tol, max_iters, 0);
    }
    if (max_iters > 0) {
      Launcher K{st};
      mg_vcycle<DIM>(K, B.mask, g, B.r, X.z, X.t, B.Ap, X.L, X.nlev);
      K((const void*)k_cg_dot2, dim3(nb), dim3(kThreads), (const float*)B.r, (const float*)X.z, (long long)g.N, B.part, nb);
      K((const void*)k_pcg_set_rz, dim3(1), dim3(1024), (const double*)B.part, nb, B.sc);
      K((const void*)k_pcg_direction, dim3(flat_blocks(g.N)), dim3(kThreads), B.p, (const float*)X.z, (long long)g.N,
        (const CgScalars*)B.sc, 1);
      cudaError_t e = K.err;
      if (e == cudaSuccess) e = iterate_pcg<DIM>(use_graph, B, X, pl, nb, max_iters, tol * tol, g, st);
      if (e != cudaSuccess) {
        set_last_error(where, e);
        return (int)e;
      }
    }
    if (info) cudaMemcpyAsync(info + 5 * b, B.sc, 5 * sizeof(double), cudaMemcpyDeviceToDevice, st);
  }
  return check_launch(where);
}

}  // namespace lbmff

using namespace lbmff;

extern "C" {

int lbmff_poisson_cg_scratch_doubles(int D, int H, int W) {
  if (D < 1 || H < 1 || W < 1) return -1;
  const Grid g1 = make_grid(1, D, H, W);
  const long long n = cg_scratch_doubles(cg_launch(g1), g1);
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

long long lbmff_poisson_mg_workspace_floats(int D, int H, int W) {
  if (D < 1 || H < 1 || W < 1) return -1;
  return D == 1 ? mg_workspace_floats<2>(make_grid(1, 1, H, W)) : mg_workspace_floats<3>(make_grid(1, D, H, W));
}

int lbmff_poisson_pcg_2d(const float* rhs, const uint8_t* mflags, float* psi, float* work, float* mg_work,
                         double* scratch, int max_iters, double tol, double tau, double dx, double dt, float* res_hist,
                         double* info, int reuse_hierarchy, int B, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(B >= 1 && H >= 1 && W >= 1, "non-positive extent");
  return poisson_pcg_impl<2>(rhs, mflags, psi, work, mg_work, scratch, max_iters, tol, tau, dx, dt, res_hist, info,
                             reuse_hierarchy, make_grid(B, 1, H, W), (cudaStream_t)stream, __func__);
}

int lbmff_poisson_pcg_3d(const float* rhs, const uint8_t* mflags, float* psi, float* work, float* mg_work,
                         double* scratch, int max_iters, double tol, double tau, double dx, double dt, float* res_hist,
                         double* info, int reuse_hierarchy, int B, int D, int H, int W, lbmff_stream_t stream) {
  LBMFF_REQUIRE(B >= 1 && D >= 1 && H >= 1 && W >= 1, "non-positive extent");
  return poisson_pcg_impl<3>(rhs, mflags, psi, work, mg_work, scratch, max_iters, tol, tau, dx, dt, res_hist, info,
                             reuse_hierarchy, make_grid(B, D, H, W), (cudaStream_t)stream, __func__);
}

}  // extern "C"
