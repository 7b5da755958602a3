// Shared device helpers for the LBM-Ferrofluid sm_100a kernels.
//
// Layout (identical to the reference's tensors):
//   populations  f[b][q][z][y][x]   fp32, x fastest  (SoA: one contiguous
//   D*H*W block per population -> unit-stride, coalesced access along x)
//   scalars      s[b][0][z][y][x],  vectors v[b][a][z][y][x] (a = x,y,z)
//   flags        uint8, FLUID = 1, OBSTACLE = 2 (reference: src/LBM/utils/types.py:7-15)
// 2-D grids are handled by the same code with D = 1 and the D2Q9 lattice.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define LBMFF_FLUID 1
#define LBMFF_OBSTACLE 2

namespace lbmff {

// A Grid describes the LOCAL array a kernel works on.  On one GPU it is the whole domain.  Under z-slab
// decomposition (multi-GPU, include/lbmff.h "slab tier") the local array holds D planes = owned planes plus
// halo planes of a periodic global domain of Dg planes; local plane z is global plane (z + zoff) mod Dg.
// Everything that depends on WHERE a cell sits in the domain (outer shell, replicate clamps, wetting faces)
// uses the global z; addresses use the local z.  Fused kernels process cells [i0, i1) of the local array
// (whole planes), so a step can be issued boundary planes first and interior planes while halos travel.
struct Grid {
  int B, D, H, W;
  long long N;  // cells per batch item = D*H*W (also the channel stride)
  int Dg;       // global number of z planes (== D on one GPU)
  int zoff;     // global z of local plane 0, in (-Dg, Dg)
  long long i0, i1;  // cell range processed by fused kernels, default [0, N)
  int zb, zn;        // the same range in planes: [zb, zb + zn)
};

__host__ __device__ inline Grid make_grid(int B, int D, int H, int W) {
  Grid g;
  g.B = B; g.D = D; g.H = H; g.W = W;
  g.N = (long long)D * H * W;
  g.Dg = D; g.zoff = 0;
  g.i0 = 0; g.i1 = g.N;
  g.zb = 0; g.zn = D;
  return g;
}

// Slab-local grid: planes [zb, zb + zn) of a local array of D planes are processed.
inline Grid make_slab_grid(int B, int D, int H, int W, int Dg, int zoff, int zb, int zn) {
  Grid g = make_grid(B, D, H, W);
  g.Dg = Dg; g.zoff = zoff;
  g.i0 = (long long)zb * H * W; g.i1 = (long long)(zb + zn) * H * W;
  g.zb = zb; g.zn = zn;
  return g;
}

// D2Q9 / D3Q19 in the reference's population order
// (LBM_collision_2d.py:66-83, LBM_collision_3d.py:76-103).
template <int DIM> struct Lat;

template <> struct Lat<2> {
  static constexpr int Q = 9;
  static constexpr int DIM = 2;
  __host__ __device__ static constexpr int ex(int q) { constexpr int t[9] = {0, 1, 0, -1, 0, 1, -1, -1, 1}; return t[q]; }
  __host__ __device__ static constexpr int ey(int q) { constexpr int t[9] = {0, 0, 1, 0, -1, 1, 1, -1, -1}; return t[q]; }
  __host__ __device__ static constexpr int ez(int) { return 0; }
  __host__ __device__ static constexpr int opp(int q) { constexpr int t[9] = {0, 3, 4, 1, 2, 7, 8, 5, 6}; return t[q]; }
  // weights are fp32 roundings of the double quotients, as torch.Tensor([...]) produces
  __host__ __device__ static constexpr float w(int q) {
    return q == 0 ? (float)(4.0 / 9.0) : (q < 5 ? (float)(1.0 / 9.0) : (float)(1.0 / 36.0));
  }
  __host__ __device__ static constexpr float grad_axis_w() { return 4.0f; }
};

template <> struct Lat<3> {
  static constexpr int Q = 19;
  static constexpr int DIM = 3;
  __host__ __device__ static constexpr int ex(int q) {
    constexpr int t[19] = {0, 1, 0, -1, 0, 1, -1, -1, 1, 0, 1, 0, -1, 0, 0, 1, 0, -1, 0};
    return t[q];
  }
  __host__ __device__ static constexpr int ey(int q) {
    constexpr int t[19] = {0, 0, 1, 0, -1, 1, 1, -1, -1, 0, 0, 1, 0, -1, 0, 0, 1, 0, -1};
    return t[q];
  }
  __host__ __device__ static constexpr int ez(int q) {
    constexpr int t[19] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, -1, -1, -1, -1, -1};
    return t[q];
  }
  __host__ __device__ static constexpr int opp(int q) {
    constexpr int t[19] = {0, 3, 4, 1, 2, 7, 8, 5, 6, 14, 17, 18, 15, 16, 9, 12, 13, 10, 11};
    return t[q];
  }
  __host__ __device__ static constexpr float w(int q) {
    return q == 0 ? (float)(1.0 / 3.0)
                  : ((q < 5 || q == 9 || q == 14) ? (float)(1.0 / 18.0) : (float)(1.0 / 36.0));
  }
  __host__ __device__ static constexpr float grad_axis_w() { return 2.0f; }
};

// cell coordinates of a linear index (x fastest)
struct Cell {
  int x, y, z;
};

__device__ __forceinline__ Cell cell_of(long long i, const Grid& g) {
  Cell c;
  int t = (int)(i / g.W);
  c.x = (int)(i - (long long)t * g.W);
  c.z = t / g.H;
  c.y = t - c.z * g.H;
  return c;
}

__device__ __forceinline__ long long lin(const Grid& g, int x, int y, int z) {
  return ((long long)z * g.H + y) * g.W + x;
}

__device__ __forceinline__ int wrap(int v, int n) { return v < 0 ? v + n : (v >= n ? v - n : v); }
__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

// clamp a coordinate into the interior [1, n-2] (identity for a size-1 axis)
__device__ __forceinline__ int interior(int v, int n) { return n < 3 ? v : clampi(v, 1, n - 2); }

// global z of local plane z
__device__ __forceinline__ int global_z(const Grid& g, int z) {
  int v = z + g.zoff;
  return v < 0 ? v + g.Dg : (v >= g.Dg ? v - g.Dg : v);
}
// is local plane z one of the two outer z planes of the global domain?
__device__ __forceinline__ bool z_on_shell(const Grid& g, int z) {
  if (g.Dg <= 1) return false;
  const int v = global_z(g, z);
  return v == 0 || v == g.Dg - 1;
}
// local plane of the interior-clamped global z (moves by at most one plane; kept inside the local array)
__device__ __forceinline__ int interior_z(const Grid& g, int z) {
  if (g.Dg < 3) return z;
  const int v = global_z(g, z);
  return clampi(z + (clampi(v, 1, g.Dg - 2) - v), 0, g.D - 1);
}
// within one plane of the global z shell (tau == 1 scalar sweep: only such cells can have a wall neighbour)
__device__ __forceinline__ bool near_z_shell(const Grid& g, int z) {
  const int v = global_z(g, z);
  return v <= 1 || v >= g.Dg - 2;
}
// local plane holding global plane gz, or -1 when this slab does not hold it
__device__ __forceinline__ int local_z(const Grid& g, int gz) {
  int l = gz - g.zoff;
  if (l < 0) l += g.Dg; else if (l >= g.Dg) l -= g.Dg;
  return l < g.D ? l : -1;
}

// ---- 32-bit neighbour addressing -----------------------------------------------------------------------
// Cells per batch item fit in an int (checked on the host: N < 2^31), so a neighbour is "my cell index plus a
// signed 32-bit delta" and a population value is base_q[cell], base_q = lattice + q * N formed once per q in
// 64 bits (uniform across the grid).  The periodic wrap lives in the six per-axis deltas; a pulled or pushed
// population then costs one or two integer adds instead of a 64-bit multiply-add chain per load.
struct Nbr {
  int xm, xp, ym, yp, zm, zp;  // index deltas to x-1, x+1, y-1, y+1, z-1, z+1 (periodic in the local array)
  template <class C>
  __device__ __forceinline__ Nbr(const C& c, const Grid& g) {
    const int HW = g.H * g.W;
    xm = c.x == 0 ? g.W - 1 : -1;
    xp = c.x == g.W - 1 ? 1 - g.W : 1;
    ym = c.y == 0 ? (g.H - 1) * g.W : -g.W;
    yp = c.y == g.H - 1 ? -(g.H - 1) * g.W : g.W;
    zm = g.D > 1 ? (c.z == 0 ? (g.D - 1) * HW : -HW) : 0;
    zp = g.D > 1 ? (c.z == g.D - 1 ? -(g.D - 1) * HW : HW) : 0;
  }
  // delta to the cell at x - s*e_q (s = +1: pull source, s = -1: the cell population q moves to)
  template <int DIM>
  __device__ __forceinline__ int at(int q, int s) const {
    using L = Lat<DIM>;
    const int ex = s * L::ex(q), ey = s * L::ey(q), ez = s * L::ez(q);
    return (ex == 1 ? xm : (ex == -1 ? xp : 0)) + (ey == 1 ? ym : (ey == -1 ? yp : 0)) +
           (DIM == 3 ? (ez == 1 ? zm : (ez == -1 ? zp : 0)) : 0);
  }
};

// Element index of population q at cell i (+ delta): 32-bit unsigned when the whole lattice has fewer than
// 2^32 elements (every named configuration: 19 * 512^3 = 2.55e9), so an address is one IMAD.WIDE.U32.
template <bool IDX32>
struct PopIndex;
template <>
struct PopIndex<true> {
  unsigned N;
  __device__ __forceinline__ explicit PopIndex(const Grid& g) : N((unsigned)g.N) {}
  __device__ __forceinline__ unsigned operator()(int q, int cell) const { return (unsigned)q * N + (unsigned)cell; }
};
template <>
struct PopIndex<false> {
  long long N;
  __device__ __forceinline__ explicit PopIndex(const Grid& g) : N(g.N) {}
  __device__ __forceinline__ long long operator()(int q, int cell) const { return q * N + cell; }
};

// Read-only global load the compiler may neither drop nor merge with a later conditional reload of the same
// variable: the fused kernels issue the common-path (FLUID) loads before the cell's flag byte has arrived.
__device__ __forceinline__ float ld_spec(const float* p) {
  float v;
  asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}

constexpr int kThreads = 256;

// ---- division-free launch geometry of the fused kernels ---------------------------------------------------
// threadIdx.x/blockIdx.x run along x, threadIdx.y/blockIdx.y along y, blockIdx.z over the planes (batch items
// are separate launches with offset pointers, so no address carries a batch term): the cell
// coordinates come straight from the hardware indices -- no integer division (there is no divide unit: the two
// divisions of cell_of cost ~90 of the ~700 instructions of the sweep kernel).  Rows narrower than 256 cells
// pack several rows into one block.
struct PlaneLaunch {
  dim3 grid, block;
};
// max_bx < 256 gives a 2-D block (e.g. 64 x 4): stencil kernels then share their y neighbours inside a block.
inline PlaneLaunch plane_launch(const Grid& g, int max_bx = kThreads) {
  int bx = ((g.W + 31) / 32) * 32;
  if (bx > max_bx) bx = max_bx;
  int by = kThreads / bx;
  if (by > g.H) by = g.H;
  if (by < 1) by = 1;
  PlaneLaunch L;
  L.block = dim3((unsigned)bx, (unsigned)by, 1);
  L.grid = dim3((unsigned)((g.W + bx - 1) / bx), (unsigned)((g.H + by - 1) / by), (unsigned)g.zn);
  return L;
}
struct PlaneCell {
  int x, y, z, i;  // i = cell index inside one batch item (N < 2^31, checked on the host)
  bool live;
};
__device__ __forceinline__ PlaneCell plane_cell(const Grid& g) {
  PlaneCell c;
  c.x = blockIdx.x * blockDim.x + threadIdx.x;
  c.y = blockIdx.y * blockDim.y + threadIdx.y;
  c.z = g.zb + blockIdx.z;
  c.live = c.x < g.W && c.y < g.H;
  c.i = (c.z * g.H + c.y) * g.W + c.x;
  return c;
}

inline unsigned blocks_for(long long n, int threads = kThreads) {
  return (unsigned)((n + threads - 1) / threads);
}

}  // namespace lbmff
