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

struct Grid {
  int B, D, H, W;
  long long N;  // cells per batch item = D*H*W
};

__host__ __device__ inline Grid make_grid(int B, int D, int H, int W) {
  Grid g;
  g.B = B; g.D = D; g.H = H; g.W = W;
  g.N = (long long)D * H * W;
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

constexpr int kThreads = 256;

inline unsigned blocks_for(long long n, int threads = kThreads) {
  return (unsigned)((n + threads - 1) / threads);
}

}  // namespace lbmff
