// Wetting boundary + capillary scalar fields, shared by the drop-in op and the fused step.
// Reference: LBM_collision_HCZ_3d.py:84-215, LBM_collision_HCZ_2d.py:84-125.
#pragma once
#include "physics.cuh"

namespace lbmff {

// ---- 3-D wetting, faces of one axis (both sides).  Reads planes 1,2 / n-2,n-3 of `axis`,
// writes planes 0 / n-1 at tangential-interior cells only -> race-free in place.
// axis 0 = x (steps 1-2, HCZ_3d.py:88-116), 1 = y (steps 3-4, :117-146), 2 = z (steps 5-6, :147-168,
// no tan term).  Later axes see the shell values written by earlier ones: launch in order x, y, z.
static __global__ void __launch_bounds__(kThreads) k_wet_faces_3d(float* __restrict__ rho,
                                                           const uint8_t* __restrict__ flags, int axis, float cot,
                                                           Grid g) {
  const int n[3] = {g.W, g.H, g.D};
  const int t1 = axis == 2 ? 1 : 2;  // slower tangential axis (z unless axis is z, then y)
  const int t2 = axis == 0 ? 1 : 0;  // faster tangential axis
  const int n1 = n[t1] - 2, n2 = n[t2] - 2;
  const long long cnt = 2LL * n1 * n2;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= cnt) return;
  const long long b = blockIdx.y;
  float* r = rho + b * g.N;
  const uint8_t* fl = flags + b * g.N;
  const int side = (int)(i / ((long long)n1 * n2));
  const long long rem = i - (long long)side * n1 * n2;
  const int a1 = (int)(rem / n2) + 1, a2 = (int)(rem % n2) + 1;
  const int wall = side ? n[axis] - 1 : 0, in1 = side ? n[axis] - 2 : 1, in2 = side ? n[axis] - 3 : 2;
  int c[3];
  auto at = [&](int va, int v1, int v2) {
    c[axis] = va; c[t1] = v1; c[t2] = v2;
    return lin(g, c[0], c[1], c[2]);
  };
  const long long iw = at(wall, a1, a2);
  if (fl[iw] != LBMFF_OBSTACLE) return;
  float v = r[at(in2, a1, a2)];
  if (axis != 2) {
    const float d1 = r[at(in1, a1 + 1, a2)] - r[at(in1, a1 - 1, a2)];
    const float d2 = r[at(in1, a1, a2 + 1)] - r[at(in1, a1, a2 - 1)];
    v = v + cot * sqrtf(1e-6f + d1 * d1 + d2 * d2);
  }
  r[iw] = v;
}

// ---- 3-D wetting, the 12 edge lines (HCZ_3d.py:171-184): mean of the two adjacent face cells.
static __global__ void __launch_bounds__(kThreads) k_wet_edges_3d(float* __restrict__ rho, Grid g) {
  const int n[3] = {g.W, g.H, g.D};
  // edges along axis `fa`: 4 per axis, free coordinate in [1, n-2]
  const long long cnt = 4LL * ((g.W - 2) + (g.H - 2) + (g.D - 2));
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= cnt) return;
  float* r = rho + (long long)blockIdx.y * g.N;
  int fa = 0;
  if (i >= 4LL * (g.W - 2)) { i -= 4LL * (g.W - 2); fa = 1; }
  if (fa == 1 && i >= 4LL * (g.H - 2)) { i -= 4LL * (g.H - 2); fa = 2; }
  const int m = (int)(i / 4) + 1, corner = (int)(i % 4);
  const int wa = fa == 0 ? 1 : 0, wb = fa == 2 ? 1 : 2;  // the two wall axes
  int c[3];
  c[fa] = m;
  c[wa] = (corner & 1) ? n[wa] - 1 : 0;
  c[wb] = (corner & 2) ? n[wb] - 1 : 0;
  const long long iw = lin(g, c[0], c[1], c[2]);
  int ca[3] = {c[0], c[1], c[2]}, cb[3] = {c[0], c[1], c[2]};
  ca[wa] = c[wa] == 0 ? 1 : n[wa] - 2;
  cb[wb] = c[wb] == 0 ? 1 : n[wb] - 2;
  r[iw] = 0.5f * (r[lin(g, ca[0], ca[1], ca[2])] + r[lin(g, cb[0], cb[1], cb[2])]);
}

// ---- 3-D wetting, the 8 corners (HCZ_3d.py:187-211): (x-inner + y-inner + z-inner) / 3
static __global__ void k_wet_corners_3d(float* __restrict__ rho, Grid g) {
  const int i = threadIdx.x;
  if (i >= 8) return;
  float* r = rho + (long long)blockIdx.y * g.N;
  const int x = (i & 1) ? g.W - 1 : 0, y = (i & 2) ? g.H - 1 : 0, z = (i & 4) ? g.D - 1 : 0;
  const int xi = x == 0 ? 1 : g.W - 2, yi = y == 0 ? 1 : g.H - 2, zi = z == 0 ? 1 : g.D - 2;
  r[lin(g, x, y, z)] = (r[lin(g, xi, y, z)] + r[lin(g, x, yi, z)] + r[lin(g, x, y, zi)]) / 3.0f;
}

// ---- 2-D wetting: axis 0 = columns x = 0 / W-1 (HCZ_2d.py:88-101), axis 1 = rows (:102-115)
static __global__ void __launch_bounds__(kThreads) k_wet_faces_2d(float* __restrict__ rho,
                                                           const uint8_t* __restrict__ flags, int axis, float cot,
                                                           Grid g) {
  const int nt = (axis == 0 ? g.H : g.W) - 2, na = axis == 0 ? g.W : g.H;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 2LL * nt) return;
  float* r = rho + (long long)blockIdx.y * g.N;
  const uint8_t* fl = flags + (long long)blockIdx.y * g.N;
  const int side = (int)(i / nt), t = (int)(i % nt) + 1;
  const int wall = side ? na - 1 : 0, in1 = side ? na - 2 : 1, in2 = side ? na - 3 : 2;
  auto at = [&](int va, int vt) { return axis == 0 ? lin(g, va, vt, 0) : lin(g, vt, va, 0); };
  const long long iw = at(wall, t);
  if (fl[iw] != LBMFF_OBSTACLE) return;
  r[iw] = r[at(in2, t)] + cot * fabsf(r[at(in1, t + 1)] - r[at(in1, t - 1)]);
}

// ---- 2-D wetting corners (HCZ_2d.py:118-121; the last one averages rho[-2,-2] with itself)
static __global__ void k_wet_corners_2d(float* __restrict__ rho, Grid g) {
  const int i = threadIdx.x;
  if (i >= 4) return;
  float* r = rho + (long long)blockIdx.y * g.N;
  const int W = g.W, H = g.H;
  if (i == 0) r[lin(g, 0, 0, 0)] = 0.5f * (r[lin(g, 0, 1, 0)] + r[lin(g, 1, 0, 0)]);
  if (i == 1) r[lin(g, 0, H - 1, 0)] = 0.5f * (r[lin(g, 0, H - 2, 0)] + r[lin(g, 1, H - 1, 0)]);
  if (i == 2) r[lin(g, W - 1, 0, 0)] = 0.5f * (r[lin(g, W - 2, 0, 0)] + r[lin(g, W - 1, 1, 0)]);
  if (i == 3) r[lin(g, W - 1, H - 1, 0)] = 0.5f * (r[lin(g, W - 2, H - 2, 0)] + r[lin(g, W - 2, H - 2, 0)]);
}

// Launch the ordered wetting sequence on `st`.
template <int DIM>
inline void launch_wetting(float* rho, const uint8_t* flags, float cot, const Grid& g, cudaStream_t st) {
  if (DIM == 3) {
    for (int axis = 0; axis < 3; ++axis) {
      const int n[3] = {g.W, g.H, g.D};
      const int t1 = axis == 2 ? 1 : 2, t2 = axis == 0 ? 1 : 0;
      const long long cnt = 2LL * (n[t1] - 2) * (n[t2] - 2);
      k_wet_faces_3d<<<dim3(blocks_for(cnt), g.B), kThreads, 0, st>>>(rho, flags, axis, cot, g);
    }
    const long long ecnt = 4LL * ((g.W - 2) + (g.H - 2) + (g.D - 2));
    k_wet_edges_3d<<<dim3(blocks_for(ecnt), g.B), kThreads, 0, st>>>(rho, g);
    k_wet_corners_3d<<<dim3(1, g.B), 32, 0, st>>>(rho, g);
  } else {
    k_wet_faces_2d<<<dim3(blocks_for(2LL * (g.H - 2)), g.B), kThreads, 0, st>>>(rho, flags, 0, cot, g);
    k_wet_faces_2d<<<dim3(blocks_for(2LL * (g.W - 2)), g.B), kThreads, 0, st>>>(rho, flags, 1, cot, g);
    k_wet_corners_2d<<<dim3(1, g.B), 32, 0, st>>>(rho, g);
  }
}

}  // namespace lbmff
