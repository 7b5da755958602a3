// Wetting boundary + capillary scalar fields, shared by the drop-in op and the fused step.
// Reference: LBM_collision_HCZ_3d.py:84-215, LBM_collision_HCZ_2d.py:84-125.
#pragma once
#include "physics.cuh"

namespace lbmff {

// ---- 3-D wetting, faces of one axis (both sides).  Reads planes 1,2 / n-2,n-3 of `axis`,
// writes planes 0 / n-1 at tangential-interior cells only -> race-free in place.
// axis 0 = x (steps 1-2, HCZ_3d.py:88-116), 1 = y (steps 3-4, :117-146), 2 = z (steps 5-6, :147-168,
// no tan term).  Later axes see the shell values written by earlier ones: launch in order x, y, z.
// Slab decomposition: x / y faces are written on local planes [zlo, zhi) that are not on the global z shell;
// z faces on whichever of those planes IS global plane 0 / Dg-1 (common.cuh Grid).
static __global__ void __launch_bounds__(kThreads) k_wet_faces_3d(float* __restrict__ rho,
                                                           const uint8_t* __restrict__ flags, int axis, float cot,
                                                           Grid g, int zlo, int zhi) {
  const long long b = blockIdx.y;
  float* r = rho + b * g.N;
  const uint8_t* fl = flags + b * g.N;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (axis == 2) {
    const int n1 = g.H - 2, n2 = g.W - 2;
    if (i >= 2LL * n1 * n2) return;
    const int side = (int)(i / ((long long)n1 * n2));
    const long long rem = i - (long long)side * n1 * n2;
    const int y = (int)(rem / n2) + 1, x = (int)(rem % n2) + 1;
    const int zw = local_z(g, side ? g.Dg - 1 : 0), z2 = local_z(g, side ? g.Dg - 3 : 2);
    if (zw < zlo || zw >= zhi || z2 < 0) return;
    const long long iw = lin(g, x, y, zw);
    if (fl[iw] != LBMFF_OBSTACLE) return;
    r[iw] = r[lin(g, x, y, z2)];
    return;
  }
  // axis 0 / 1: slower tangential axis is z (all local planes of the range), faster one is the other in-plane axis
  const int na = axis == 0 ? g.W : g.H;
  const int n2 = (axis == 0 ? g.H : g.W) - 2, n1 = zhi - zlo;
  if (i >= 2LL * n1 * n2) return;
  const int side = (int)(i / ((long long)n1 * n2));
  const long long rem = i - (long long)side * n1 * n2;
  const int z = (int)(rem / n2) + zlo, a2 = (int)(rem % n2) + 1;
  if (z_on_shell(g, z) || z < 1 || z > g.D - 2) return;
  const int wall = side ? na - 1 : 0, in1 = side ? na - 2 : 1, in2 = side ? na - 3 : 2;
  auto at = [&](int va, int vz, int v2) { return axis == 0 ? lin(g, va, v2, vz) : lin(g, v2, va, vz); };
  const long long iw = at(wall, z, a2);
  if (fl[iw] != LBMFF_OBSTACLE) return;
  const float d1 = r[at(in1, z + 1, a2)] - r[at(in1, z - 1, a2)];
  const float d2 = r[at(in1, z, a2 + 1)] - r[at(in1, z, a2 - 1)];
  r[iw] = r[at(in2, z, a2)] + cot * sqrtf(1e-6f + d1 * d1 + d2 * d2);
}

// ---- 3-D wetting, the 12 edge lines (HCZ_3d.py:171-184): mean of the two adjacent face cells.
static __global__ void __launch_bounds__(kThreads) k_wet_edges_3d(float* __restrict__ rho, Grid g, int zlo,
                                                                  int zhi) {
  const int n[3] = {g.W, g.H, g.Dg};
  // edges along axis `fa`: 4 per axis; free coordinate in [1, n-2] (x, y) or the local planes [zlo, zhi) (z)
  const long long cnt = 4LL * ((g.W - 2) + (g.H - 2) + (zhi - zlo));
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= cnt) return;
  float* r = rho + (long long)blockIdx.y * g.N;
  int fa = 0;
  if (i >= 4LL * (g.W - 2)) { i -= 4LL * (g.W - 2); fa = 1; }
  if (fa == 1 && i >= 4LL * (g.H - 2)) { i -= 4LL * (g.H - 2); fa = 2; }
  const int corner = (int)(i % 4);
  const int wa = fa == 0 ? 1 : 0, wb = fa == 2 ? 1 : 2;  // the two wall axes
  int c[3];   // x, y local; z kept LOCAL in c[2]
  if (fa == 2) {
    c[2] = (int)(i / 4) + zlo;
    if (z_on_shell(g, c[2])) return;
  } else {
    c[fa] = (int)(i / 4) + 1;
  }
  c[wa] = (corner & 1) ? n[wa] - 1 : 0;
  int ca[3], cb[3];
  if (wb == 2) {
    const int gzw = (corner & 2) ? g.Dg - 1 : 0;
    c[2] = local_z(g, gzw);
    if (c[2] < zlo || c[2] >= zhi) return;
    ca[0] = cb[0] = c[0]; ca[1] = cb[1] = c[1]; ca[2] = cb[2] = c[2];
    cb[2] = local_z(g, gzw == 0 ? 1 : g.Dg - 2);
    if (cb[2] < 0) return;
  } else {
    c[wb] = (corner & 2) ? n[wb] - 1 : 0;
    ca[0] = cb[0] = c[0]; ca[1] = cb[1] = c[1]; ca[2] = cb[2] = c[2];
    cb[wb] = c[wb] == 0 ? 1 : n[wb] - 2;
  }
  ca[wa] = c[wa] == 0 ? 1 : n[wa] - 2;
  r[lin(g, c[0], c[1], c[2])] = 0.5f * (r[lin(g, ca[0], ca[1], ca[2])] + r[lin(g, cb[0], cb[1], cb[2])]);
}

// ---- 3-D wetting, the 8 corners (HCZ_3d.py:187-211): (x-inner + y-inner + z-inner) / 3
static __global__ void k_wet_corners_3d(float* __restrict__ rho, Grid g, int zlo, int zhi) {
  const int i = threadIdx.x;
  if (i >= 8) return;
  float* r = rho + (long long)blockIdx.y * g.N;
  const int x = (i & 1) ? g.W - 1 : 0, y = (i & 2) ? g.H - 1 : 0;
  const int z = local_z(g, (i & 4) ? g.Dg - 1 : 0), zi = local_z(g, (i & 4) ? g.Dg - 2 : 1);
  if (z < zlo || z >= zhi || zi < 0) return;
  const int xi = x == 0 ? 1 : g.W - 2, yi = y == 0 ? 1 : g.H - 2;
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

// Launch the ordered wetting sequence on `st`.  3-D, slab decomposition: `stage` selects what is issued so
// that halo planes can be refreshed in between -- 0: everything (one GPU), 1: x faces only, 2: the rest;
// the cell range [g.i0, g.i1) gives the local planes written.
template <int DIM>
inline void launch_wetting(float* rho, const uint8_t* flags, float cot, const Grid& g, cudaStream_t st,
                           int stage = 0) {
  if (DIM == 3) {
    const long long HW = (long long)g.H * g.W;
    const int zlo = (int)(g.i0 / HW), zhi = (int)(g.i1 / HW);
    for (int axis = 0; axis < 3; ++axis) {
      if ((stage == 1 && axis != 0) || (stage == 2 && axis == 0)) continue;
      const long long cnt = axis == 2 ? 2LL * (g.H - 2) * (g.W - 2)
                                      : 2LL * (zhi - zlo) * ((axis == 0 ? g.H : g.W) - 2);
      if (cnt <= 0) continue;
      k_wet_faces_3d<<<dim3(blocks_for(cnt), g.B), kThreads, 0, st>>>(rho, flags, axis, cot, g, zlo, zhi);
    }
    if (stage == 1) return;
    const long long ecnt = 4LL * ((g.W - 2) + (g.H - 2) + (zhi - zlo));
    k_wet_edges_3d<<<dim3(blocks_for(ecnt), g.B), kThreads, 0, st>>>(rho, g, zlo, zhi);
    k_wet_corners_3d<<<dim3(1, g.B), 32, 0, st>>>(rho, g, zlo, zhi);
  } else {
    k_wet_faces_2d<<<dim3(blocks_for(2LL * (g.H - 2)), g.B), kThreads, 0, st>>>(rho, flags, 0, cot, g);
    k_wet_faces_2d<<<dim3(blocks_for(2LL * (g.W - 2)), g.B), kThreads, 0, st>>>(rho, flags, 1, cot, g);
    k_wet_corners_2d<<<dim3(1, g.B), 32, 0, st>>>(rho, g);
  }
}

}  // namespace lbmff
