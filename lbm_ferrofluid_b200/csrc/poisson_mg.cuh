// Geometric multigrid V-cycle for the wall-aware weighted graph Laplacian of poisson_cg.cu, used as the
// preconditioner of its conjugate gradients (north_star item 3).
//
// Hierarchy: 2 x 2 (x 2) cell aggregation, coarse extents ceil(n / 2), periodic like the fine lattice.  Transfer:
// piecewise-constant prolongation P and restriction R = P^T.  Coarse operators: the Galerkin products P^T A P (scaled
// by kGalerkinScale = 1/2 per level, see below), which
// for piecewise-constant P are again nearest-neighbour stencils on the same 8 / 18 lattice directions -- the weight
// of the coarse link I -> I - e_Q is the sum of the fine link weights that cross from aggregate I into aggregate
// I - e_Q (k_mg_coarsen).  Walls need no special case: a link into a wall cell has weight 0 on the fine level (full-way
// bounce-back returns the population: poisson_cg.cu), so it contributes nothing; an aggregate of walls only is a
// dead cell (diagonal 0, correction 0).  Rows sum to zero on every level, as on the fine one.
// Smoother: damped Jacobi (omega = 0.9), two sweeps before and two after the coarse correction, the first from a zero
// guess (pointwise); coarsest level: 24 Jacobi sweeps from zero.  Every component is symmetric (R = P^T, equal pre / post
// smoothing, polynomial coarse solve), so M^-1 is symmetric positive semi-definite with the constants in its null
// space, as CG needs.  Stencil kernels use the division-free plane geometry of common.cuh; the per-level work is one
// coalesced SoA weight array [Q-1][N_l].
#pragma once
#include "common.cuh"

namespace lbmff {
namespace mg {

constexpr float kOmega = 0.9f;
// Piecewise-constant aggregation makes the Galerkin operator of a Laplacian too stiff by the coarsening ratio: the
// face between two 2 x 2 (x 2) aggregates carries the sum of the fine links (4 / h^2 per face in 3-D) where the
// rediscretised operator on cells of size 2h has 8 / (2h)^2 = 2 / h^2, so the coarse correction comes out half as
// large as it should and the cycle converges slowly and mesh-dependently (27 / 38 PCG iterations at 256^3 / 512^3).
// Scaling every coarse operator by 1/2 (the over-correction of Braess 1995 for unsmoothed aggregation) keeps the
// hierarchy symmetric and gives mesh-independent convergence (5-6 iterations to 1e-4).
constexpr float kGalerkinScale = 0.5f;
constexpr int kMaxLevels = 10;
constexpr int kCoarsestSweeps = 24;

struct Level {
  Grid g;        // extents of this level (B = 1)
  float* w;      // [Q-1][N] link weights, w[(q-1) N + i] = weight of the link i -> i - e_q   (levels >= 1)
  float* invd;   // [N] 1 / diagonal (0 on dead cells)                                          (levels >= 1)
  float* x;      // [N] correction
  float* t;      // [N] second Jacobi buffer / residual
  float* b;      // [N] right-hand side (restricted residual)
};

template <int DIM>
__host__ __device__ constexpr int dir_of(int dx, int dy, int dz) {
  using L = Lat<DIM>;
  for (int q = 1; q < L::Q; ++q)
    if (L::ex(q) == dx && L::ey(q) == dy && L::ez(q) == dz) return q;
  return 0;
}

__device__ __forceinline__ PlaneCell cell_at(const Grid& g, int z) {
  PlaneCell c;
  c.x = blockIdx.x * blockDim.x + threadIdx.x;
  c.y = blockIdx.y * blockDim.y + threadIdx.y;
  c.z = z;
  c.live = c.x < g.W && c.y < g.H;
  c.i = (c.z * g.H + c.y) * g.W + c.x;
  return c;
}
#define LBMFF_MG_PLANES(c, g) \
  for (int zz_ = blockIdx.z; zz_ < (g).D; zz_ += gridDim.z) \
    if (PlaneCell c = cell_at(g, zz_); c.live)

// wrapped difference of two coordinates on a periodic axis of n cells, in {-1, 0, +1}
__device__ __forceinline__ int wrap_diff(int a, int b, int n) {
  int d = a - b;
  if (d > 1) d -= n; else if (d < -1) d += n;
  return d;
}

// Fine-link accessors for k_mg_coarsen: weight of the link i -> i - e_q on the finer level.
template <int DIM>
struct MaskLinks {   // level 0: from the wall masks of poisson_cg.cu (bit 0: the cell is a wall, bit q: x - e_q is)
  const unsigned* mask;
  __device__ __forceinline__ float operator()(int i, int q, long long) const {
    const unsigned m = mask[i];
    if ((m & 1u) || ((m >> q) & 1u)) return 0.f;
    return Lat<DIM>::w(q) / (1.0f - Lat<DIM>::w(0));
  }
};
struct ArrayLinks {  // levels >= 1
  const float* w;
  __device__ __forceinline__ float operator()(int i, int q, long long N) const { return w[(size_t)(q - 1) * N + i]; }
};

// Galerkin coarse weights: one thread per COARSE cell walks its children and their links.
template <int DIM, class LINKS>
__global__ void __launch_bounds__(kThreads) k_mg_coarsen(LINKS fine, Grid gf, float* __restrict__ wc,
                                                         float* __restrict__ invd, Grid gc) {
  using L = Lat<DIM>;
  LBMFF_MG_PLANES(c, gc) {
    float acc[L::Q];
#pragma unroll
    for (int q = 0; q < L::Q; ++q) acc[q] = 0.f;
    for (int cz = 0; cz < (DIM == 3 ? 2 : 1); ++cz)
      for (int cy = 0; cy < 2; ++cy)
        for (int cx = 0; cx < 2; ++cx) {
          const int x = 2 * c.x + cx, y = 2 * c.y + cy, z = DIM == 3 ? 2 * c.z + cz : 0;
          if (x >= gf.W || y >= gf.H || z >= gf.D) continue;
          const int i = (z * gf.H + y) * gf.W + x;
#pragma unroll
          for (int q = 1; q < L::Q; ++q) {
            const float wq = fine(i, q, gf.N);
            if (wq == 0.f) continue;
            const int nx = wrap(x - L::ex(q), gf.W), ny = wrap(y - L::ey(q), gf.H);
            const int nz = DIM == 3 ? wrap(z - L::ez(q), gf.D) : 0;
            // aggregate of the neighbour, relative to this one (the link weight is stored for direction e_Q with
            // I - e_Q = J, i.e. e_Q = I - J)
            const int dx = wrap_diff(c.x, nx >> 1, gc.W), dy = wrap_diff(c.y, ny >> 1, gc.H);
            const int dz = DIM == 3 ? wrap_diff(c.z, nz >> 1, gc.D) : 0;
            if (dx == 0 && dy == 0 && dz == 0) continue;   // inside the aggregate: cancels in P^T A P
            // (dx, dy, dz) has at most two non-zero components because e_q has; find its lattice direction
#pragma unroll
            for (int Q = 1; Q < L::Q; ++Q)
              if (L::ex(Q) == dx && L::ey(Q) == dy && L::ez(Q) == dz) acc[Q] += wq;
          }
        }
    float d = 0.f;
#pragma unroll
    for (int q = 1; q < L::Q; ++q) {
      const float a = kGalerkinScale * acc[q];
      wc[(size_t)(q - 1) * gc.N + c.i] = a;
      d += a;
    }
    invd[c.i] = d > 0.f ? 1.0f / d : 0.f;
  }
}

// (A v)(i) on a level >= 1: sum_q w_q(i) (v(i) - v(i - e_q)), differences first (see laplace() in poisson_cg.cu)
template <int DIM>
__device__ __forceinline__ float apply_w(const float* __restrict__ w, long long N, const float* __restrict__ v, float vc,
                                         int i, const Nbr& nb) {
  using L = Lat<DIM>;
  float s = 0.f;
#pragma unroll
  for (int q = 1; q < L::Q; ++q) s = fmaf(__ldg(w + (size_t)(q - 1) * N + i), vc - __ldg(v + i + nb.at<DIM>(q, 1)), s);
  return s;
}

// x = omega * invd * b   (a Jacobi sweep from a zero guess)
__global__ void __launch_bounds__(kThreads) k_mg_jacobi0(const float* __restrict__ b, const float* __restrict__ invd,
                                                         float* __restrict__ x, long long n) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    x[i] = kOmega * invd[i] * b[i];
}

// MODE 0: xo = xi + omega invd (b - A xi)    (Jacobi sweep)      MODE 1: xo = b - A xi   (residual)
template <int DIM, int MODE>
__global__ void __launch_bounds__(kThreads) k_mg_sweep(const float* __restrict__ w, const float* __restrict__ invd,
                                                       const float* __restrict__ b, const float* __restrict__ xi,
                                                       float* __restrict__ xo, Grid g) {
  LBMFF_MG_PLANES(c, g) {
    const Nbr nb(c, g);
    const float xc = __ldg(xi + c.i);
    const float r = b[c.i] - apply_w<DIM>(w, g.N, xi, xc, c.i, nb);
    xo[c.i] = MODE == 0 ? fmaf(kOmega * invd[c.i], r, xc) : (invd[c.i] != 0.f ? r : 0.f);
  }
}

// coarse b = sum of the fine residual over the children (R = P^T)
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_mg_restrict(const float* __restrict__ rf, Grid gf, float* __restrict__ bc,
                                                          Grid gc) {
  LBMFF_MG_PLANES(c, gc) {
    float s = 0.f;
    for (int cz = 0; cz < (DIM == 3 ? 2 : 1); ++cz)
      for (int cy = 0; cy < 2; ++cy)
        for (int cx = 0; cx < 2; ++cx) {
          const int x = 2 * c.x + cx, y = 2 * c.y + cy, z = DIM == 3 ? 2 * c.z + cz : 0;
          if (x < gf.W && y < gf.H && z < gf.D) s += rf[(z * gf.H + y) * gf.W + x];
        }
    bc[c.i] = s;
  }
}

// fine x += coarse x of the aggregate (piecewise-constant prolongation); `live` (optional, level 0): 1 / diagonal or
// any array that is 0 exactly on the cells that must stay untouched (walls)
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_mg_prolong_add(float* __restrict__ xf, Grid gf,
                                                             const float* __restrict__ xc, Grid gc,
                                                             const unsigned* __restrict__ mask0) {
  LBMFF_MG_PLANES(c, gf) {
    if (mask0 && (mask0[c.i] & 1u)) continue;
    xf[c.i] += xc[(((DIM == 3 ? c.z >> 1 : 0) * gc.H) + (c.y >> 1)) * gc.W + (c.x >> 1)];
  }
}

}  // namespace mg
}  // namespace lbmff
