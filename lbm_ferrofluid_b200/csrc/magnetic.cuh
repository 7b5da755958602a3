// Magnetic Poisson-lattice pieces shared by the drop-in get_H_int and the fused sweeps.
// Reference: LBM_magnetic_3d.py:125-188, LBM_magnetic_2d.py:105-164.
#pragma once
#include "physics.cuh"

namespace lbmff {

// Face arrays of the applied field (get_staggered(H_ext), demo_3d...Rosensweig...py:147-149):
// Hx [D][H][W+1], Hy [D][H+1][W], Hz [D+1][H][W].
struct HextFaces {
  const float* __restrict__ x;
  const float* __restrict__ y;
  const float* __restrict__ z;
};

// Sweep-invariant source: rhs = div(chi_face H_face) dx / (1 + chi) on FLUID cells, else 0
// (LBM_magnetic_3d.py:161-179).  `Chi` is a callable (x, y, z) -> chi at a cell.
template <int DIM, class Chi>
__device__ __forceinline__ float magnetic_rhs_at(const Chi& chi, const HextFaces& Hf, int x, int y, int z,
                                                 const Grid& g, float dx) {
  const float c0 = chi(x, y, z);
  // face j (0..n) averages cells clamp(j,1,n-1)-1 and clamp(j,1,n-1) (replicate-padded, grid.py:10-13)
  const int xl = clampi(x, 1, g.W - 1), xh = clampi(x + 1, 1, g.W - 1);
  const int yl = clampi(y, 1, g.H - 1), yh = clampi(y + 1, 1, g.H - 1);
  const float fxl = (chi(xl, y, z) + chi(xl - 1, y, z)) * 0.5f, fxh = (chi(xh, y, z) + chi(xh - 1, y, z)) * 0.5f;
  const float fyl = (chi(x, yl, z) + chi(x, yl - 1, z)) * 0.5f, fyh = (chi(x, yh, z) + chi(x, yh - 1, z)) * 0.5f;
  // 32-bit face indices: (D+1)(H+1)(W+1) < 2^31 is checked on the host
  const int ix = (z * g.H + y) * (g.W + 1) + x;
  const int iy = (z * (g.H + 1) + y) * g.W + x;
  float div = (fxh * __ldg(Hf.x + ix + 1) - fxl * __ldg(Hf.x + ix)) +
              (fyh * __ldg(Hf.y + iy + g.W) - fyl * __ldg(Hf.y + iy));
  if (DIM == 3) {
    // replicate clamp on the GLOBAL face index, addresses local (slab decomposition: common.cuh Grid)
    const int gz = global_z(g, z);
    const int zl = clampi(z + (clampi(gz, 1, g.Dg - 1) - gz), 1, g.D - 1);
    const int zh = clampi(z + (clampi(gz + 1, 1, g.Dg - 1) - gz), 1, g.D - 1);
    const float fzl = (chi(x, y, zl) + chi(x, y, zl - 1)) * 0.5f, fzh = (chi(x, y, zh) + chi(x, y, zh - 1)) * 0.5f;
    const int iz = (z * g.H + y) * g.W + x;
    div = div + (fzh * __ldg(Hf.z + iz + g.H * g.W) - fzl * __ldg(Hf.z + iz));
  }
  return div * dx / (1.0f + c0);
}

// psi = sum_{q>=1} h_q / (1 - w0)   (LBM_magnetic_3d.py:155-157); fixed summation order q = 1..Q-1
template <int DIM>
__device__ __forceinline__ float psi_of(const float* h) {
  float s = h[1];
#pragma unroll
  for (int q = 2; q < Lat<DIM>::Q; ++q) s = __fadd_rn(s, h[q]);
  return s / (1.0f - Lat<DIM>::w(0));
}

// post-collision population q (LBM_magnetic_3d.py:160,180-181), uncontracted so that the
// population-faithful sweep and the tau = 1 scalar sweep round identically.
template <int DIM>
__device__ __forceinline__ float mag_collide_q(int q, float hq, float psi, float rhs, const Consts& k) {
  const float w = Lat<DIM>::w(q);
  float heq = __fmul_rn(psi, w);
  if (q == 0) heq = __fsub_rn(heq, psi);
  const float add = __fmul_rn(__fmul_rn(__fmul_rn(k.dt, w), rhs), k.mag_src);
  return __fadd_rn(__fadd_rn(__fmul_rn(k.keep, hq), __fmul_rn(k.relax, heq)), add);
}

}  // namespace lbmff
