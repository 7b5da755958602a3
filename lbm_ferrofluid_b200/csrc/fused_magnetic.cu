// Fused tier, magnetic subsystem: the 30 Poisson-lattice relaxation sweeps per outer step
// (demo_3d_LBM_Rosensweig_instability.py:163-173 -> LBM_propagation_3d.py:18-142 + LBM_magnetic_3d.py:125-188).
//
//  k_mag_source      sweep-invariant source rhs and chi from rho, once per outer step
//  k_mag_sweep       population-faithful sweep: stream + bounce + psi + collide in ONE pass over h
//                    (A/B buffers; 2*Q*4 B per cell per sweep = 152 B D3Q19, 72 B D2Q9)
//  k_mag_sweep_scalar  tau == 1 collapse: the post-collision state of a cell is (w' psi + w S), so only the
//                    two distinct outgoing values per cell (axis / diagonal weight class) are stored:
//                    8 B written + 8 B read (unique) per cell per sweep.  Arithmetic is the same uncontracted
//                    sequence as k_mag_sweep -> bit-identical psi on non-OBSTACLE cells.
#include <stdlib.h>

#include "capi.cuh"
#include "magnetic.cuh"

namespace lbmff {

struct ChiArray {
  const float* __restrict__ chi;
  int W, HW;
  __device__ __forceinline__ float operator()(int x, int y, int z) const { return __ldg(chi + (z * HW + y * W + x)); }
};

// rhs (FLUID cells of mflags, else 0) from the cell-centred susceptibility chi (raw, no shell rule)
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_mag_source(const float* __restrict__ chi,
                                                         const uint8_t* __restrict__ mflags, HextFaces Hf,
                                                         float* __restrict__ rhs, float dx, Grid g) {
  const PlaneCell c = plane_cell(g);
  if (!c.live) return;
  float r = 0.0f;
  if (mflags[c.i] == LBMFF_FLUID)
    r = magnetic_rhs_at<DIM>(ChiArray{chi, g.W, g.H * g.W}, Hf, c.x, c.y, c.z, g, dx);
  rhs[c.i] = r;
}

// ---- population-faithful fused sweep ----------------------------------------------------------
// OBSTACLE cell x: out_q(x) = in_opp(q)(x + e_q)   (stream, then bounce; walls never collide)
// other cells:     h_q = in_q(x - e_q); psi = sum_{q>=1} h_q / (1 - w0); out_q = collide(h_q, psi, rhs)
// HINT: 0 default caching, 1 streaming stores (st.global.cs)
template <int DIM, bool WRITE_PSI, int HINT = 0, bool IDX32 = true, int MB = 1>
__global__ void __launch_bounds__(kThreads, MB) k_mag_sweep(const float* __restrict__ hin, float* __restrict__ hout,
                                                        const uint8_t* __restrict__ mflags,
                                                        const float* __restrict__ rhs, float* __restrict__ psi_out,
                                                        Consts k, Grid g) {
  using L = Lat<DIM>;
  const PlaneCell c = plane_cell(g);
  if (!c.live) return;
  const int i = c.i;
  const float* in = hin;
  float* out = hout;
  const Nbr nb(c, g);
  const PopIndex<IDX32> at(g);
  // The 19 pulls are issued BEFORE the flag byte is known (a load that waits for the flag costs a second full
  // memory round trip per thread).  A wall cell needs out_q(x) = in_opp(q)(x + e_q), which is exactly the value
  // pulled for population opp(q): bounce-back is a permutation of the same 19 loads.
  float hv[L::Q];
#pragma unroll
  for (int q = 0; q < L::Q; ++q) hv[q] = ld_spec(in + at(q, i + nb.at<DIM>(q, 1)));
  const bool obst = mflags[i] == LBMFF_OBSTACLE;
  const float r = __ldg(rhs + i);
  if (obst) {  // divergence only in warps that touch a wall
#pragma unroll
    for (int q = 1; q < L::Q; ++q) {
      if (q < L::opp(q)) {
        const float t = hv[q];
        hv[q] = hv[L::opp(q)];
        hv[L::opp(q)] = t;
      }
    }
  }
  const float psi = psi_of<DIM>(hv);
  if (WRITE_PSI) psi_out[i] = psi;
  if (!obst) {
#pragma unroll
    for (int q = 0; q < L::Q; ++q) hv[q] = mag_collide_q<DIM>(q, hv[q], psi, r, k);
  }
#pragma unroll
  for (int q = 0; q < L::Q; ++q) {
    if (HINT >= 1) __stcs(out + at(q, i), hv[q]);
    else out[at(q, i)] = hv[q];
  }
}

// ---- population-faithful sweep, four cells per thread -------------------------------------------------
// Same arithmetic per cell as k_mag_sweep (bit-identical results); W % 4 == 0.  A thread owns cells
// x0..x0+3 of one row: populations with e_x = 0 are one aligned 128-bit load, those with e_x = +-1 one aligned
// 128-bit load plus the neighbouring lane's edge value (warp shuffle; the lane at a warp or row end loads
// it), every store is an aligned 128-bit store.  A quarter of the address arithmetic per cell and 16 B
// requests on every stream.  OBSTACLE cells (the x shell of every row, whole z-shell planes) are redone by
// the one-cell routine after the vector store.
template <int DIM>
__device__ __forceinline__ void mag_sweep_cell(const float* __restrict__ in, float* __restrict__ out,
                                               const uint8_t* __restrict__ fl, const float* __restrict__ rhs,
                                               float* __restrict__ psi_out, bool write_psi, long long i,
                                               const Consts& k, const Grid& g) {
  using L = Lat<DIM>;
  const Cell c = cell_of(i, g);
  const int xm = wrap(c.x - 1, g.W), xp = wrap(c.x + 1, g.W);
  const int ym = wrap(c.y - 1, g.H), yp = wrap(c.y + 1, g.H);
  const int zm = DIM == 3 ? wrap(c.z - 1, g.D) : 0, zp = DIM == 3 ? wrap(c.z + 1, g.D) : 0;
  const bool obst = fl[i] == LBMFF_OBSTACLE;
  float hv[L::Q];
#pragma unroll
  for (int q = 0; q < L::Q; ++q) {
    const int s = obst ? -1 : 1;
    const int sx = L::ex(q) * s == 1 ? xm : (L::ex(q) * s == -1 ? xp : c.x);
    const int sy = L::ey(q) * s == 1 ? ym : (L::ey(q) * s == -1 ? yp : c.y);
    const int sz = L::ez(q) * s == 1 ? zm : (L::ez(q) * s == -1 ? zp : c.z);
    hv[q] = __ldg(in + (obst ? L::opp(q) : q) * g.N + lin(g, sx, sy, sz));
  }
  const float psi = psi_of<DIM>(hv);
  if (write_psi) psi_out[i] = psi;
  if (!obst) {
    const float r = rhs[i];
#pragma unroll
    for (int q = 0; q < L::Q; ++q) hv[q] = mag_collide_q<DIM>(q, hv[q], psi, r, k);
  }
#pragma unroll
  for (int q = 0; q < L::Q; ++q) out[q * g.N + i] = hv[q];
}

template <int DIM, bool WRITE_PSI>
__global__ void __launch_bounds__(kThreads) k_mag_sweep_v4(const float* __restrict__ hin, float* __restrict__ hout,
                                                           const uint8_t* __restrict__ mflags,
                                                           const float* __restrict__ rhs,
                                                           float* __restrict__ psi_out, Consts k, Grid g) {
  using L = Lat<DIM>;
  const long long i = g.i0 + 4 * ((long long)blockIdx.x * blockDim.x + threadIdx.x);
  const bool live = i < g.i1;
  const long long b = blockIdx.y;
  const float* in = hin + b * L::Q * g.N;
  float* out = hout + b * L::Q * g.N;
  const uint8_t* fl = mflags + b * g.N;
  const long long ic = live ? i : g.i0;   // idle lanes shadow a valid cell so the shuffles stay full-warp
  const Cell c = cell_of(ic, g);
  const int lane = threadIdx.x & 31;
  const long long HW = (long long)g.H * g.W;
  const long long yo[3] = {(long long)wrap(c.y - 1, g.H) * g.W, (long long)c.y * g.W,
                           (long long)wrap(c.y + 1, g.H) * g.W};
  const long long zo[3] = {DIM == 3 ? wrap(c.z - 1, g.D) * HW : 0, c.z * HW, DIM == 3 ? wrap(c.z + 1, g.D) * HW : 0};
  const bool row_start = c.x == 0, row_end = c.x + 4 == g.W;
  const int xl = row_start ? g.W - 1 : c.x - 1, xr = row_end ? 0 : c.x + 4;
  float4 hv[L::Q];
#pragma unroll
  for (int q = 0; q < L::Q; ++q) {
    const float* row = in + q * g.N + zo[1 - L::ez(q)] + yo[1 - L::ey(q)];
    const float4 a = __ldg(reinterpret_cast<const float4*>(row + c.x));
    if (L::ex(q) == 0) {
      hv[q] = a;
    } else if (L::ex(q) == 1) {   // cell x pulls from x - 1
      float left = __shfl_up_sync(0xffffffffu, a.w, 1);
      if (lane == 0 || row_start) left = __ldg(row + xl);
      hv[q] = make_float4(left, a.x, a.y, a.z);
    } else {                       // cell x pulls from x + 1
      float right = __shfl_down_sync(0xffffffffu, a.x, 1);
      if (lane == 31 || row_end) right = __ldg(row + xr);
      hv[q] = make_float4(a.y, a.z, a.w, right);
    }
  }
  if (!live) return;
  float4 psi;
  {
    float4 s = hv[1];
#pragma unroll
    for (int q = 2; q < L::Q; ++q) {
      s.x = __fadd_rn(s.x, hv[q].x); s.y = __fadd_rn(s.y, hv[q].y);
      s.z = __fadd_rn(s.z, hv[q].z); s.w = __fadd_rn(s.w, hv[q].w);
    }
    const float d = 1.0f - L::w(0);
    psi = make_float4(s.x / d, s.y / d, s.z / d, s.w / d);
  }
  if (WRITE_PSI) *reinterpret_cast<float4*>(psi_out + b * g.N + i) = psi;
  const float4 r = __ldg(reinterpret_cast<const float4*>(rhs + b * g.N + i));
#pragma unroll
  for (int q = 0; q < L::Q; ++q) {
    float4 o;
    o.x = mag_collide_q<DIM>(q, hv[q].x, psi.x, r.x, k);
    o.y = mag_collide_q<DIM>(q, hv[q].y, psi.y, r.y, k);
    o.z = mag_collide_q<DIM>(q, hv[q].z, psi.z, r.z, k);
    o.w = mag_collide_q<DIM>(q, hv[q].w, psi.w, r.w, k);
    *reinterpret_cast<float4*>(out + q * g.N + i) = o;
  }
  const uchar4 f4 = *reinterpret_cast<const uchar4*>(fl + i);
  if (f4.x == LBMFF_OBSTACLE || f4.y == LBMFF_OBSTACLE || f4.z == LBMFF_OBSTACLE || f4.w == LBMFF_OBSTACLE) {
    const uint8_t fv[4] = {f4.x, f4.y, f4.z, f4.w};
#pragma unroll
    for (int t = 0; t < 4; ++t)
      if (fv[t] == LBMFF_OBSTACLE)
        mag_sweep_cell<DIM>(in, out, fl, rhs + b * g.N, psi_out + b * g.N, WRITE_PSI, i + t, k, g);
  }
}

// ---- tau == 1 scalar sweep ---------------------------------------------------------------------
// a[0] = outgoing value of the axis-weight populations, a[1] = of the diagonal-weight populations.
// cur = state after the previous sweep, prev = state one sweep earlier AND the output buffer: a population
// that comes back from a wall cell is this cell's own outgoing value of two sweeps ago (full-way bounce-back
// = one sweep of delay), i.e. prev at the SAME cell, read by the same thread before it overwrites it.
template <int DIM, bool WRITE_PSI>
__global__ void __launch_bounds__(kThreads) k_mag_sweep_scalar(const float* __restrict__ ca, float* pa,
                                                               const uint8_t* __restrict__ fl,
                                                               const float* __restrict__ rhs,
                                                               float* __restrict__ psi_out, Consts k, Grid g) {
  using L = Lat<DIM>;
  const PlaneCell c = plane_cell(g);
  if (!c.live) return;
  const int i = c.i, N = (int)g.N;
  const Nbr nb(c, g);
  // the 18 neighbour values are requested before any flag is looked at
  float hv[L::Q];
  hv[0] = 0.0f;
#pragma unroll
  for (int q = 1; q < L::Q; ++q) {
    const bool axis = (L::ex(q) != 0) + (L::ey(q) != 0) + (L::ez(q) != 0) == 1;
    hv[q] = ld_spec(ca + (axis ? 0 : N) + (i + nb.at<DIM>(q, 1)));
  }
  const float r = __ldg(rhs + i);
  if (fl[i] == LBMFF_OBSTACLE) return;  // wall state is implicit (never observed: get_grad masks it)
  // walls lie on the outer shell only (checked by the host), so only shell-adjacent cells test flags
  const bool near_shell = c.x <= 1 || c.x >= g.W - 2 || c.y <= 1 || c.y >= g.H - 2 ||
                          (DIM == 3 && near_z_shell(g, c.z));
  if (near_shell) {
    unsigned wall_mask = 0;
#pragma unroll
    for (int q = 1; q < L::Q; ++q)
      if (fl[i + nb.at<DIM>(q, 1)] == LBMFF_OBSTACLE) wall_mask |= 1u << q;
    if (wall_mask) {
      const float p_axis = pa[i], p_diag = pa[N + i];
#pragma unroll
      for (int q = 1; q < L::Q; ++q) {
        const bool axis = (L::ex(q) != 0) + (L::ey(q) != 0) + (L::ez(q) != 0) == 1;
        if ((wall_mask >> q) & 1u) hv[q] = axis ? p_axis : p_diag;
      }
    }
  }
  const float psi = psi_of<DIM>(hv);
  if (WRITE_PSI) psi_out[i] = psi;
  // representative populations of the two weight classes: q = 1 (axis), q = 5 (diagonal) in both lattices
  pa[i] = mag_collide_q<DIM>(1, 0.0f, psi, r, k);
  pa[N + i] = mag_collide_q<DIM>(5, 0.0f, psi, r, k);
}

// ---- tau == 1 scalar sweep, shared-memory tiled, marching along z (3-D) --------------------------------------
// One block owns a 32 x 8 column of cells and walks zc planes upwards.  Each plane of the two state arrays is
// staged once (with its one-cell x/y rim: 34 x 10 values) into a ring of four shared-memory slots, so a value is
// fetched from L2 1.33 times instead of 9 times, and the 18 neighbour values of a cell are shared-memory reads.
// Arithmetic and summation order are those of k_mag_sweep_scalar / k_mag_sweep: bit-identical psi.
constexpr int kTX = 32, kTY = 8, kRing = 4;
constexpr int kTileW = kTX + 2, kTileH = kTY + 2;

template <bool WRITE_PSI>
__global__ void __launch_bounds__(kTX* kTY) k_mag_sweep_scalar_tiled(const float* __restrict__ ca, float* pa,
                                                                      const uint8_t* __restrict__ fl,
                                                                      const float* __restrict__ rhs,
                                                                      float* __restrict__ psi_out, Consts k, Grid g,
                                                                      int zc) {
  using L = Lat<3>;
  __shared__ float sA[kRing][kTileH][kTileW];  // axis-class value
  __shared__ float sD[kRing][kTileH][kTileW];  // diagonal-class value
  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * kTX + tx;
  const int x0 = blockIdx.x * kTX, y0 = blockIdx.y * kTY;
  const int zs = g.zb + blockIdx.z * zc;
  const int ze = min(zs + zc, g.zb + g.zn);
  const int HW = g.H * g.W, N = (int)g.N;
  // the (up to) two rim-tile elements this thread stages per plane: in-plane offsets, periodic in x and y
  int off[2];
  int sy[2], sx[2];
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    const int e = tid + t * kTX * kTY;
    sy[t] = e / kTileW; sx[t] = e - sy[t] * kTileW;
    int gx = x0 - 1 + sx[t], gy = y0 - 1 + sy[t];
    gx = gx < 0 ? gx + g.W : gx; gy = gy < 0 ? gy + g.H : gy;
    while (gx >= g.W) gx -= g.W;
    while (gy >= g.H) gy -= g.H;
    off[t] = e < kTileW * kTileH ? gy * g.W + gx : -1;
  }
  // software pipeline: plane z+2 is fetched into registers while plane z is computed, and put into its ring
  // slot at the top of the next iteration -- no iteration waits for a global load it has just issued
  float ra[2], rd[2];
  auto fetch = [&](int z) {  // local plane z (periodic in the local array) -> registers
    const int zz = z < 0 ? z + g.D : (z >= g.D ? z - g.D : z);
#pragma unroll
    for (int t = 0; t < 2; ++t)
      if (off[t] >= 0) {
        const int j = zz * HW + off[t];
        ra[t] = __ldg(ca + j);
        rd[t] = __ldg(ca + N + j);
      }
  };
  auto put = [&](int z) {  // registers -> ring slot of plane z
    const int slot = (z - zs + 1) & (kRing - 1);
#pragma unroll
    for (int t = 0; t < 2; ++t)
      if (off[t] >= 0) {
        sA[slot][sy[t]][sx[t]] = ra[t];
        sD[slot][sy[t]][sx[t]] = rd[t];
      }
  };
  fetch(zs - 1); put(zs - 1);
  fetch(zs); put(zs);
  fetch(zs + 1);
  const int x = x0 + tx, y = y0 + ty;
  const bool live = x < g.W && y < g.H;
  const Nbr nb0(PlaneCell{live ? x : 0, live ? y : 0, 0, 0, true}, g);  // x / y deltas do not depend on z
  // the cell's own source and flag are prefetched one plane ahead as well
  const int i0 = live ? zs * HW + y * g.W + x : 0;
  float r_next = __ldg(rhs + i0);
  uint8_t f_next = fl[i0];
  for (int z = zs; z < ze; ++z) {
    put(z + 1);
    if (z + 1 < ze) fetch(z + 2);
    const float r = r_next;
    const uint8_t fcell = f_next;
    if (live && z + 1 < ze) {
      r_next = __ldg(rhs + (z + 1) * HW + y * g.W + x);
      f_next = fl[(z + 1) * HW + y * g.W + x];
    }
    __syncthreads();
    if (live) {
      const int i = z * HW + y * g.W + x;
      if (fcell != LBMFF_OBSTACLE) {
        const int s0 = (z - zs) & (kRing - 1), s1 = (z - zs + 1) & (kRing - 1), s2 = (z - zs + 2) & (kRing - 1);
        const int sl[3] = {s0, s1, s2};  // planes z-1, z, z+1
        float hv[L::Q];
        hv[0] = 0.0f;
#pragma unroll
        for (int q = 1; q < L::Q; ++q) {
          const bool axis = (L::ex(q) != 0) + (L::ey(q) != 0) + (L::ez(q) != 0) == 1;
          const int px = tx + 1 - L::ex(q), py = ty + 1 - L::ey(q), ps = sl[1 - L::ez(q)];
          hv[q] = axis ? sA[ps][py][px] : sD[ps][py][px];
        }
        // walls lie on the outer shell only (checked by the host), so only shell-adjacent cells test flags
        const bool near_shell = x <= 1 || x >= g.W - 2 || y <= 1 || y >= g.H - 2 || near_z_shell(g, z);
        if (near_shell) {
          Nbr nb = nb0;
          nb.zm = z == 0 ? (g.D - 1) * HW : -HW;
          nb.zp = z == g.D - 1 ? -(g.D - 1) * HW : HW;
          unsigned wall_mask = 0;
#pragma unroll
          for (int q = 1; q < L::Q; ++q)
            if (fl[i + nb.at<3>(q, 1)] == LBMFF_OBSTACLE) wall_mask |= 1u << q;
          if (wall_mask) {
            const float p_axis = pa[i], p_diag = pa[N + i];
#pragma unroll
            for (int q = 1; q < L::Q; ++q) {
              const bool axis = (L::ex(q) != 0) + (L::ey(q) != 0) + (L::ez(q) != 0) == 1;
              if ((wall_mask >> q) & 1u) hv[q] = axis ? p_axis : p_diag;
            }
          }
        }
        const float psi = psi_of<3>(hv);
        if (WRITE_PSI) psi_out[i] = psi;
        pa[i] = mag_collide_q<3>(1, 0.0f, psi, r, k);
        pa[N + i] = mag_collide_q<3>(5, 0.0f, psi, r, k);
      }
    }
  }
}

// Rebuild the population-faithful h from the scalar state (for API users that ask for h).
// cur = state after the last sweep, prev = one sweep earlier, psi = psi of the last sweep.
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_mag_materialize(const float* __restrict__ cur,
                                                              const float* __restrict__ prev,
                                                              const float* __restrict__ psi,
                                                              const float* __restrict__ rhs,
                                                              const uint8_t* __restrict__ mflags,
                                                              float* __restrict__ h, Consts k, Grid g) {
  using L = Lat<DIM>;
  const long long i = g.i0 + (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.i1) return;
  const long long b = blockIdx.y;
  const float* ca = cur + b * 2 * g.N;
  const float* pa = prev + b * 2 * g.N;
  const uint8_t* fl = mflags + b * g.N;
  float* out = h + b * L::Q * g.N;
  const Cell c = cell_of(i, g);
  if (fl[i] != LBMFF_OBSTACLE) {
    out[i] = mag_collide_q<DIM>(0, 0.0f, psi[b * g.N + i], rhs[b * g.N + i], k);
#pragma unroll
    for (int q = 1; q < L::Q; ++q) {
      const bool axis = (L::ex(q) != 0) + (L::ey(q) != 0) + (L::ez(q) != 0) == 1;
      out[q * g.N + i] = ca[(axis ? 0 : g.N) + i];
    }
  } else {
    // wall cell s: h_q(s) = (h of one sweep earlier)_{opp q}(s + e_q) when that neighbour is not a wall;
    // chains between wall cells carry values older than the two kept levels and are left at zero
    out[i] = 0.0f;
#pragma unroll
    for (int q = 1; q < L::Q; ++q) {
      const bool axis = (L::ex(q) != 0) + (L::ey(q) != 0) + (L::ez(q) != 0) == 1;
      const long long j = lin(g, wrap(c.x + L::ex(q), g.W), wrap(c.y + L::ey(q), g.H),
                              DIM == 3 ? wrap(c.z + L::ez(q), g.D) : 0);
      out[q * g.N + i] = fl[j] != LBMFF_OBSTACLE ? pa[(axis ? 0 : g.N) + j] : 0.0f;
    }
  }
}

template <int DIM>
static void launch_sweep(bool wp, int hint, bool idx32, const PlaneLaunch& pl, cudaStream_t st, const float* in,
                         float* out, const uint8_t* mflags, const float* rhs, float* psi_out, const Consts& k,
                         const Grid& g) {
#define LBMFF_SW(WP, HI, IX)                                                                                \
  k_mag_sweep<DIM, WP, HI, IX><<<pl.grid, pl.block, 0, st>>>(in + b * Lat<DIM>::Q * g.N, out + b * Lat<DIM>::Q * g.N, \
                                                             mflags + b * g.N, rhs + b * g.N,                    \
                                                             psi_out ? psi_out + b * g.N : nullptr, k, g)
  // minimum resident blocks per SM (register cap): 5 -> 48 registers, 1280 threads/SM: 3.17 ms per 512^3 sweep
  // (6.43 TB/s); uncapped (64 registers, 4 blocks) 3.29 ms; 6 -> 40 registers with spills 3.38 ms; 8 -> 4.84 ms
  static const int mb = [] { const char* e = getenv("LBMFF_SWEEP_MB"); return e ? atoi(e) : 5; }();
  for (long long b = 0; b < g.B; ++b) {
    if (idx32 && !hint && mb >= 5) {
#define LBMFF_SWMB(WP, M)                                                                                          \
  k_mag_sweep<DIM, WP, 0, true, M><<<pl.grid, pl.block, 0, st>>>(in + b * Lat<DIM>::Q * g.N, out + b * Lat<DIM>::Q * g.N, \
                                                                 mflags + b * g.N, rhs + b * g.N,                 \
                                                                 psi_out ? psi_out + b * g.N : nullptr, k, g)
      if (mb == 5) { if (wp) LBMFF_SWMB(true, 5); else LBMFF_SWMB(false, 5); }
      else if (mb == 6) { if (wp) LBMFF_SWMB(true, 6); else LBMFF_SWMB(false, 6); }
      else { if (wp) LBMFF_SWMB(true, 8); else LBMFF_SWMB(false, 8); }
#undef LBMFF_SWMB
      continue;
    }
    if (idx32) {
      if (hint) { if (wp) LBMFF_SW(true, 1, true); else LBMFF_SW(false, 1, true); }
      else { if (wp) LBMFF_SW(true, 0, true); else LBMFF_SW(false, 0, true); }
    } else {
      if (wp) LBMFF_SW(true, 0, false); else LBMFF_SW(false, 0, false);
    }
  }
#undef LBMFF_SW
}

// ---- TWO population-faithful sweeps in one pass over h (temporal blocking, 3-D) -----------------------------------
// A block owns a 32 x 8 column of cells and marches upwards in z.  Per plane it (A) runs sweep 1 on the column
// extended by a one-cell rim (34 x 10 cells, one thread each) straight from global memory and parks the 19
// post-sweep-1 populations of that plane in a four-slot shared-memory ring, then (B) runs sweep 2 for the 32 x 8
// cells of the plane below from the ring and stores the result.  h is read 1.33x and written once per TWO sweeps
// (about 180 B per cell instead of 304 B); the rim is recomputed by the neighbouring block with identical
// arithmetic, so the result is bit-identical to two k_mag_sweep launches.  The 19 pulls of the next plane are in
// flight in registers while (B) runs.
constexpr int kPX = 32, kPY = 8, kPW = kPX + 2, kPH = kPY + 2, kPCells = kPW * kPH, kPThreads = 352, kPRing = 4;
constexpr size_t kPairSmem = (size_t)kPRing * 19 * kPCells * sizeof(float);

template <bool WRITE_PSI, bool IDX32>
__global__ void __launch_bounds__(kPThreads, 2)
k_mag_sweep_pair(const float* __restrict__ in, float* __restrict__ out, const uint8_t* __restrict__ fl,
                 const float* __restrict__ rhs, float* __restrict__ psi_out, Consts k, Grid g, int zc) {
  using L = Lat<3>;
  extern __shared__ float s1[];   // [ring][q][kPH][kPW]
  const int tid = threadIdx.x;
  const int x0 = blockIdx.x * kPX, y0 = blockIdx.y * kPY;
  const int zs = g.zb + blockIdx.z * zc;
  const int ze = min(zs + zc, g.zb + g.zn);   // sweep-2 output planes [zs, ze)
  const int HW = g.H * g.W;
  const PopIndex<IDX32> at(g);
  // ---- role A: one cell of the rim-extended tile (periodic in x, y)
  const bool a_on = tid < kPCells;
  const int ay = a_on ? tid / kPW : 0, ax = a_on ? tid - ay * kPW : 0;
  int gx = x0 - 1 + ax, gy = y0 - 1 + ay;
  gx = gx < 0 ? gx + g.W : gx; gy = gy < 0 ? gy + g.H : gy;
  while (gx >= g.W) gx -= g.W;
  while (gy >= g.H) gy -= g.H;
  const int a_xy = gy * g.W + gx;
  Nbr nba(PlaneCell{gx, gy, 0, 0, true}, g);
  // ---- role B: one output cell of the tile
  const bool b_on = tid < kPX * kPY;
  const int bx = tid & (kPX - 1), by = tid / kPX;
  const int ox = x0 + bx, oy = y0 + by;
  const bool b_live = b_on && ox < g.W && oy < g.H;
  const int b_xy = b_live ? oy * g.W + ox : 0;

  float P[L::Q];      // pulls of the plane role A works on next (in flight while role B runs)
  float p_rhs = 0.f;
  uint8_t p_fl = LBMFF_FLUID;
  auto fetch = [&](int z) {   // z: any integer plane index, periodic in the local array
    if (!a_on) return;
    const int zz = z < 0 ? z + g.D : (z >= g.D ? z - g.D : z);
    nba.zm = zz == 0 ? (g.D - 1) * HW : -HW;
    nba.zp = zz == g.D - 1 ? -(g.D - 1) * HW : HW;
    const int i = zz * HW + a_xy;
#pragma unroll
    for (int q = 0; q < L::Q; ++q) P[q] = ld_spec(in + at(q, i + nba.at<3>(q, 1)));
    p_rhs = __ldg(rhs + i);
    p_fl = fl[i];
  };
  auto sweep1_to_ring = [&](int z) {   // consume P: h1 of plane z -> ring slot
    if (!a_on) return;
    if (p_fl == LBMFF_OBSTACLE) {
#pragma unroll
      for (int q = 1; q < L::Q; ++q)
        if (q < L::opp(q)) { const float t = P[q]; P[q] = P[L::opp(q)]; P[L::opp(q)] = t; }
    } else {
      const float psi = psi_of<3>(P);
#pragma unroll
      for (int q = 0; q < L::Q; ++q) P[q] = mag_collide_q<3>(q, P[q], psi, p_rhs, k);
    }
    float* slot = s1 + ((z - zs + 1) & (kPRing - 1)) * (L::Q * kPCells) + ay * kPW + ax;
#pragma unroll
    for (int q = 0; q < L::Q; ++q) slot[q * kPCells] = P[q];
  };

  fetch(zs - 1);
  sweep1_to_ring(zs - 1);
  fetch(zs);
  sweep1_to_ring(zs);
  fetch(zs + 1);
  // own flag / source of the first output plane
  float b_rhs = b_live ? __ldg(rhs + zs * HW + b_xy) : 0.f;
  uint8_t b_fl = b_live ? fl[zs * HW + b_xy] : (uint8_t)LBMFF_FLUID;
  for (int z = zs; z < ze; ++z) {
    sweep1_to_ring(z + 1);
    if (z + 1 < ze) fetch(z + 2);
    const float r = b_rhs;
    const uint8_t f0 = b_fl;
    if (b_live && z + 1 < ze) {
      b_rhs = __ldg(rhs + (z + 1) * HW + b_xy);
      b_fl = fl[(z + 1) * HW + b_xy];
    }
    __syncthreads();
    if (b_live) {
      const int i = z * HW + b_xy;
      const float* ring[3] = {s1 + ((z - zs) & (kPRing - 1)) * (L::Q * kPCells),
                              s1 + ((z - zs + 1) & (kPRing - 1)) * (L::Q * kPCells),
                              s1 + ((z - zs + 2) & (kPRing - 1)) * (L::Q * kPCells)};   // planes z-1, z, z+1
      float hv[L::Q];
#pragma unroll
      for (int q = 0; q < L::Q; ++q)
        hv[q] = ring[1 - L::ez(q)][q * kPCells + (by + 1 - L::ey(q)) * kPW + (bx + 1 - L::ex(q))];
      const bool obst = f0 == LBMFF_OBSTACLE;
      if (obst) {
#pragma unroll
        for (int q = 1; q < L::Q; ++q)
          if (q < L::opp(q)) { const float t = hv[q]; hv[q] = hv[L::opp(q)]; hv[L::opp(q)] = t; }
      }
      const float psi = psi_of<3>(hv);
      if (WRITE_PSI) psi_out[i] = psi;
      if (!obst) {
#pragma unroll
        for (int q = 0; q < L::Q; ++q) hv[q] = mag_collide_q<3>(q, hv[q], psi, r, k);
      }
#pragma unroll
      for (int q = 0; q < L::Q; ++q) out[at(q, i)] = hv[q];
    }
  }
}

// ---- the same two-sweep pass, warp-specialised ---------------------------------------------------------------------
// Producer warps (352 threads, one per cell of the 34 x 10 extended tile) run sweep 1 plane after plane into the
// ring; consumer warps (256 threads, one per output cell) run sweep 2 out of it.  The two groups are coupled only by
// one full/empty named barrier pair per ring slot (bar.arrive / bar.sync), so producers run up to three planes ahead
// and neither group ever waits for the other's memory latency at a block-wide barrier.
constexpr int kWsProd = 352, kWsCons = kPX * kPY, kWsThreads = kWsProd + kWsCons;

__device__ __forceinline__ void named_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void named_arrive(int id, int n) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory");
}

template <bool WRITE_PSI, bool IDX32>
__global__ void __launch_bounds__(kWsThreads, 2)
k_mag_sweep_pair_ws(const float* __restrict__ in, float* __restrict__ out, const uint8_t* __restrict__ fl,
                    const float* __restrict__ rhs, float* __restrict__ psi_out, Consts k, Grid g, int zc) {
  using L = Lat<3>;
  extern __shared__ float s1[];   // [ring][q][kPH][kPW]
  const int x0 = blockIdx.x * kPX, y0 = blockIdx.y * kPY;
  const int zs = g.zb + blockIdx.z * zc;
  const int ze = min(zs + zc, g.zb + g.zn);   // sweep-2 output planes [zs, ze)
  const int HW = g.H * g.W;
  const PopIndex<IDX32> at(g);
  // barrier ids: 1 + slot = "slot full", 1 + kPRing + slot = "slot empty"
  if (threadIdx.x >= kWsCons) {
    // ================= producers: sweep 1 on the extended tile =================
    const int tid = threadIdx.x - kWsCons;
    const bool on = tid < kPCells;
    const int ay = on ? tid / kPW : 0, ax = on ? tid - ay * kPW : 0;
    int gx = x0 - 1 + ax, gy = y0 - 1 + ay;
    gx = gx < 0 ? gx + g.W : gx; gy = gy < 0 ? gy + g.H : gy;
    while (gx >= g.W) gx -= g.W;
    while (gy >= g.H) gy -= g.H;
    const int a_xy = gy * g.W + gx;
    Nbr nba(PlaneCell{gx, gy, 0, 0, true}, g);
    float P[L::Q];
    float p_rhs = 0.f;
    uint8_t p_fl = LBMFF_FLUID;
    auto fetch = [&](int z) {
      const int zz = z < 0 ? z + g.D : (z >= g.D ? z - g.D : z);
      nba.zm = zz == 0 ? (g.D - 1) * HW : -HW;
      nba.zp = zz == g.D - 1 ? -(g.D - 1) * HW : HW;
      const int i = zz * HW + a_xy;
#pragma unroll
      for (int q = 0; q < L::Q; ++q) P[q] = ld_spec(in + at(q, i + nba.at<3>(q, 1)));
      p_rhs = __ldg(rhs + i);
      p_fl = fl[i];
    };
    if (on) fetch(zs - 1);
    for (int pz = zs - 1, it = 0; pz <= ze; ++pz, ++it) {
      const int slot = (pz - zs + 1) & (kPRing - 1);
      if (it >= kPRing) named_sync(1 + kPRing + slot, kWsThreads);   // consumers are done with plane pz - 4
      if (on) {
        if (p_fl == LBMFF_OBSTACLE) {
#pragma unroll
          for (int q = 1; q < L::Q; ++q)
            if (q < L::opp(q)) { const float t = P[q]; P[q] = P[L::opp(q)]; P[L::opp(q)] = t; }
        } else {
          const float psi = psi_of<3>(P);
#pragma unroll
          for (int q = 0; q < L::Q; ++q) P[q] = mag_collide_q<3>(q, P[q], psi, p_rhs, k);
        }
        float* dst = s1 + slot * (L::Q * kPCells) + ay * kPW + ax;
#pragma unroll
        for (int q = 0; q < L::Q; ++q) dst[q * kPCells] = P[q];
        if (pz < ze) fetch(pz + 1);
      }
      __threadfence_block();
      named_arrive(1 + slot, kWsThreads);
    }
  } else {
    // ================= consumers: sweep 2 on the tile =================
    const int tid = threadIdx.x;
    const int bx = tid & (kPX - 1), by = tid / kPX;
    const int ox = x0 + bx, oy = y0 + by;
    const bool live = ox < g.W && oy < g.H;
    const int b_xy = live ? oy * g.W + ox : 0;
    float b_rhs = live ? __ldg(rhs + zs * HW + b_xy) : 0.f;
    uint8_t b_fl = live ? fl[zs * HW + b_xy] : (uint8_t)LBMFF_FLUID;
    named_sync(1 + 0, kWsThreads);   // plane zs - 1 (slot 0)
    named_sync(1 + 1, kWsThreads);   // plane zs     (slot 1)
    for (int z = zs; z < ze; ++z) {
      const int s0 = (z - zs) & (kPRing - 1), s1i = (z - zs + 1) & (kPRing - 1), s2 = (z - zs + 2) & (kPRing - 1);
      const float r = b_rhs;
      const uint8_t f0 = b_fl;
      if (live && z + 1 < ze) {
        b_rhs = __ldg(rhs + (z + 1) * HW + b_xy);
        b_fl = fl[(z + 1) * HW + b_xy];
      }
      named_sync(1 + s2, kWsThreads);   // plane z + 1
      if (live) {
        const int i = z * HW + b_xy;
        const float* ring[3] = {s1 + s0 * (L::Q * kPCells), s1 + s1i * (L::Q * kPCells), s1 + s2 * (L::Q * kPCells)};
        float hv[L::Q];
#pragma unroll
        for (int q = 0; q < L::Q; ++q)
          hv[q] = ring[1 - L::ez(q)][q * kPCells + (by + 1 - L::ey(q)) * kPW + (bx + 1 - L::ex(q))];
        const bool obst = f0 == LBMFF_OBSTACLE;
        if (obst) {
#pragma unroll
          for (int q = 1; q < L::Q; ++q)
            if (q < L::opp(q)) { const float t = hv[q]; hv[q] = hv[L::opp(q)]; hv[L::opp(q)] = t; }
        }
        const float psi = psi_of<3>(hv);
        if (WRITE_PSI) psi_out[i] = psi;
        if (!obst) {
#pragma unroll
          for (int q = 0; q < L::Q; ++q) hv[q] = mag_collide_q<3>(q, hv[q], psi, r, k);
        }
#pragma unroll
        for (int q = 0; q < L::Q; ++q) out[at(q, i)] = hv[q];
      }
      named_arrive(1 + kPRing + s0, kWsThreads);   // plane z - 1 is free
    }
  }
}

// LBMFF_SWEEP_PAIR = 1: issue the population-faithful sweeps two at a time (k_mag_sweep_pair) | 0: one per launch
static int pair_variant() {
  static const int v = [] {
    const char* e = getenv("LBMFF_SWEEP_PAIR");
    return e ? atoi(e) : 0;
  }();
  return v;
}

// LBMFF_SWEEP_VARIANT = 1 (one cell per thread, default: 3.80 ms per 512^3 sweep) | 4 (four cells per thread: 3.87 ms)
static int sweep_variant() {
  static const int v = [] {
    const char* e = getenv("LBMFF_SWEEP_VARIANT");
    return e ? atoi(e) : 1;
  }();
  return v;
}

// LBMFF_SCALAR_VARIANT = 1 (shared-memory tiled, default) | 0 (one thread per cell, L1/L2 served)
static int scalar_variant() {
  static const int v = [] {
    const char* e = getenv("LBMFF_SCALAR_VARIANT");
    return e ? atoi(e) : 1;
  }();
  return v;
}

template <int DIM>
static int sweeps_impl(float* hA, float* hB, const uint8_t* mflags, const float* rhs, float* psi_out, int n,
                       double tau, double dx, double dt, const Grid& g, cudaStream_t st, const char* where) {
  if (!(hA && hB && hA != hB && mflags && rhs)) return bad_arg(where, "null or aliased pointer");
  if (n < 1) return bad_arg(where, "n_sweeps must be >= 1");
  if (!(tau > 0)) return bad_arg(where, "tau must be positive");
  if (g.N > 0x7fffffffLL) return bad_arg(where, "more than 2^31-1 cells per batch item");
  const Consts k = make_consts(dx, dt, tau);
  const dim3 grid(blocks_for(g.i1 - g.i0), (unsigned)g.B);
  // four cells per thread needs 16-byte aligned rows and plane ranges
  const bool vec4 = sweep_variant() == 4 && g.W % 4 == 0 && g.i0 % 4 == 0 && g.i1 % 4 == 0 &&
                    ((uintptr_t)hA | (uintptr_t)hB | (uintptr_t)rhs | (uintptr_t)psi_out | (uintptr_t)mflags) % 16 == 0;
  const dim3 grid4(blocks_for((g.i1 - g.i0) / 4), (unsigned)g.B);
  const PlaneLaunch pl = plane_launch(g);
  // Temporal blocking: sweeps are issued two at a time where that keeps the documented buffer parity (the result
  // must land in hA for even n, hB for odd n, and every launch flips the buffer): n - 2*np singles, np pairs,
  // with (n - np) of the same parity as n, i.e. np even.
  int np = 0;
  if (DIM == 3 && pair_variant() && g.W >= 8 && g.H >= 8 && g.zn >= 4) {
    np = (n / 2) & ~1;
    static bool attr_set = false;
    if (!attr_set) {
      cudaFuncSetAttribute(k_mag_sweep_pair<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPairSmem);
      cudaFuncSetAttribute(k_mag_sweep_pair<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPairSmem);
      cudaFuncSetAttribute(k_mag_sweep_pair<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPairSmem);
      cudaFuncSetAttribute(k_mag_sweep_pair<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPairSmem);
      cudaFuncSetAttribute(k_mag_sweep_pair_ws<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPairSmem);
      cudaFuncSetAttribute(k_mag_sweep_pair_ws<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPairSmem);
      attr_set = true;
    }
  }
  const int launches = n - np;
  int done = 0;   // sweeps completed
  for (int s = 0; s < launches; ++s) {
    const float* in = (s & 1) ? hB : hA;
    float* out = (s & 1) ? hA : hB;
    const bool pair = s < np;
    done += pair ? 2 : 1;
    const bool wp = done == n && psi_out;
    if (pair) {
      const int zc = g.zn < 32 ? g.zn : 32;
      const dim3 tg((g.W + kPX - 1) / kPX, (g.H + kPY - 1) / kPY, (g.zn + zc - 1) / zc);
      const bool ix = (long long)Lat<DIM>::Q * g.N < 0xffffffffLL;
      for (long long b = 0; b < g.B; ++b) {
        const float* i_ = in + b * 19 * g.N;
        float* o_ = out + b * 19 * g.N;
        const uint8_t* f_ = mflags + b * g.N;
        const float* r_ = rhs + b * g.N;
        float* p_ = psi_out ? psi_out + b * g.N : nullptr;
        if (ix && pair_variant() == 2) {
          if (wp) k_mag_sweep_pair_ws<true, true><<<tg, kWsThreads, kPairSmem, st>>>(i_, o_, f_, r_, p_, k, g, zc);
          else k_mag_sweep_pair_ws<false, true><<<tg, kWsThreads, kPairSmem, st>>>(i_, o_, f_, r_, p_, k, g, zc);
        } else if (ix) {
          if (wp) k_mag_sweep_pair<true, true><<<tg, kPThreads, kPairSmem, st>>>(i_, o_, f_, r_, p_, k, g, zc);
          else k_mag_sweep_pair<false, true><<<tg, kPThreads, kPairSmem, st>>>(i_, o_, f_, r_, p_, k, g, zc);
        } else {
          if (wp) k_mag_sweep_pair<true, false><<<tg, kPThreads, kPairSmem, st>>>(i_, o_, f_, r_, p_, k, g, zc);
          else k_mag_sweep_pair<false, false><<<tg, kPThreads, kPairSmem, st>>>(i_, o_, f_, r_, p_, k, g, zc);
        }
      }
      continue;
    }
    if (vec4) {
      if (wp) k_mag_sweep_v4<DIM, true><<<grid4, kThreads, 0, st>>>(in, out, mflags, rhs, psi_out, k, g);
      else k_mag_sweep_v4<DIM, false><<<grid4, kThreads, 0, st>>>(in, out, mflags, rhs, psi_out, k, g);
    } else {
      launch_sweep<DIM>(wp, sweep_variant() == 2 ? 1 : 0, (long long)Lat<DIM>::Q * g.N < 0xffffffffLL, pl, st, in, out,
                        mflags, rhs, psi_out, k, g);
    }
  }
  return check_launch(where);
}

template <int DIM>
static int sweeps_scalar_impl(float* aA, float* aB, const uint8_t* mflags, const float* rhs, float* psi_out, int n,
                              double dx, double dt, const Grid& g, cudaStream_t st, const char* where) {
  if (!(aA && aB && aA != aB && mflags && rhs)) return bad_arg(where, "null or aliased pointer");
  if (n < 1) return bad_arg(where, "n_sweeps must be >= 1");
  if (g.N > 0x7fffffffLL / 2) return bad_arg(where, "more than (2^31-1)/2 cells per batch item");
  const Consts k = make_consts(dx, dt, 1.0);
  const PlaneLaunch pl = plane_launch(g);
  for (int s = 0; s < n; ++s) {
    const float* cur = (s & 1) ? aB : aA;
    float* prev = (s & 1) ? aA : aB;
    for (long long b = 0; b < g.B; ++b) {
      if (DIM == 3 && scalar_variant() == 1 && g.W >= 4 && g.H >= 4) {
        const int zc = g.zn < 32 ? g.zn : 32;
        const dim3 tg((g.W + kTX - 1) / kTX, (g.H + kTY - 1) / kTY, (g.zn + zc - 1) / zc);
        if (s == n - 1 && psi_out)
          k_mag_sweep_scalar_tiled<true><<<tg, dim3(kTX, kTY), 0, st>>>(cur + b * 2 * g.N, prev + b * 2 * g.N,
                                                                        mflags + b * g.N, rhs + b * g.N,
                                                                        psi_out + b * g.N, k, g, zc);
        else
          k_mag_sweep_scalar_tiled<false><<<tg, dim3(kTX, kTY), 0, st>>>(cur + b * 2 * g.N, prev + b * 2 * g.N,
                                                                         mflags + b * g.N, rhs + b * g.N, nullptr,
                                                                         k, g, zc);
        continue;
      }
      if (s == n - 1 && psi_out)
        k_mag_sweep_scalar<DIM, true><<<pl.grid, pl.block, 0, st>>>(cur + b * 2 * g.N, prev + b * 2 * g.N,
                                                                    mflags + b * g.N, rhs + b * g.N,
                                                                    psi_out + b * g.N, k, g);
      else
        k_mag_sweep_scalar<DIM, false><<<pl.grid, pl.block, 0, st>>>(cur + b * 2 * g.N, prev + b * 2 * g.N,
                                                                     mflags + b * g.N, rhs + b * g.N, nullptr, k, g);
    }
  }
  return check_launch(where);
}

template <int DIM>
static int source_impl(const float* chi, const uint8_t* mflags, const float* Hx, const float* Hy, const float* Hz,
                       double dx, float* rhs, const Grid& g, cudaStream_t st, const char* where) {
  if (!(chi && mflags && Hx && Hy && (DIM == 2 || Hz) && rhs)) return bad_arg(where, "null pointer");
  if (g.W < 3 || g.H < 3 || (DIM == 3 && g.D < 3)) return bad_arg(where, "every extent must be >= 3");
  if ((g.D + 1LL) * (g.H + 1LL) * (g.W + 1LL) > 0x7fffffffLL) return bad_arg(where, "grid too large for 32-bit cell indices");
  const PlaneLaunch pl = plane_launch(g, 64);
  const long long D = g.D, H = g.H, W = g.W;
  for (long long b = 0; b < g.B; ++b) {
    const HextFaces Hf{Hx + b * D * H * (W + 1), Hy + b * D * (H + 1) * W, DIM == 3 ? Hz + b * (D + 1) * H * W : nullptr};
    k_mag_source<DIM><<<pl.grid, pl.block, 0, st>>>(chi + b * g.N, mflags + b * g.N, Hf, rhs + b * g.N, (float)dx, g);
  }
  return check_launch(where);
}

template <int DIM>
static int materialize_impl(const float* cur, const float* prev, const float* psi, const float* rhs,
                            const uint8_t* mflags, float* h, double dx, double dt, int B, int D, int H, int W,
                            cudaStream_t st, const char* where) {
  if (!(cur && prev && psi && rhs && mflags && h)) return bad_arg(where, "null pointer");
  const Grid g = make_grid(B, D, H, W);
  k_mag_materialize<DIM><<<dim3(blocks_for(g.i1 - g.i0), (unsigned)B), kThreads, 0, st>>>(cur, prev, psi, rhs, mflags, h,
                                                                                  make_consts(dx, dt, 1.0), g);
  return check_launch(where);
}

}  // namespace lbmff

using namespace lbmff;

extern "C" {

int lbmff_magnetic_source_2d(const float* chi, const uint8_t* mflags, const float* Hx, const float* Hy, double dx,
                             float* rhs_out, int B, int H, int W, lbmff_stream_t stream) {
  return source_impl<2>(chi, mflags, Hx, Hy, nullptr, dx, rhs_out, make_grid(B, 1, H, W), (cudaStream_t)stream,
                        __func__);
}
int lbmff_magnetic_source_3d(const float* chi, const uint8_t* mflags, const float* Hx, const float* Hy,
                             const float* Hz, double dx, float* rhs_out, int B, int D, int H, int W,
                             lbmff_stream_t stream) {
  return source_impl<3>(chi, mflags, Hx, Hy, Hz, dx, rhs_out, make_grid(B, D, H, W), (cudaStream_t)stream,
                        __func__);
}
int lbmff_magnetic_sweeps_d2q9(float* hA, float* hB, const uint8_t* mflags, const float* rhs, float* psi_out,
                               int n_sweeps, double tau, double dx, double dt, int B, int H, int W,
                               lbmff_stream_t stream) {
  if (!psi_out) return bad_arg(__func__, "null psi_out");
  return sweeps_impl<2>(hA, hB, mflags, rhs, psi_out, n_sweeps, tau, dx, dt, make_grid(B, 1, H, W),
                        (cudaStream_t)stream, __func__);
}
int lbmff_magnetic_sweeps_d3q19(float* hA, float* hB, const uint8_t* mflags, const float* rhs, float* psi_out,
                                int n_sweeps, double tau, double dx, double dt, int B, int D, int H, int W,
                                lbmff_stream_t stream) {
  if (!psi_out) return bad_arg(__func__, "null psi_out");
  return sweeps_impl<3>(hA, hB, mflags, rhs, psi_out, n_sweeps, tau, dx, dt, make_grid(B, D, H, W),
                        (cudaStream_t)stream, __func__);
}
int lbmff_magnetic_sweeps_scalar_2d(float* aA, float* aB, const uint8_t* mflags, const float* rhs, float* psi_out,
                                    int n_sweeps, double dx, double dt, int B, int H, int W,
                                    lbmff_stream_t stream) {
  if (!psi_out) return bad_arg(__func__, "null psi_out");
  return sweeps_scalar_impl<2>(aA, aB, mflags, rhs, psi_out, n_sweeps, dx, dt, make_grid(B, 1, H, W),
                               (cudaStream_t)stream, __func__);
}
int lbmff_magnetic_sweeps_scalar_3d(float* aA, float* aB, const uint8_t* mflags, const float* rhs, float* psi_out,
                                    int n_sweeps, double dx, double dt, int B, int D, int H, int W,
                                    lbmff_stream_t stream) {
  if (!psi_out) return bad_arg(__func__, "null psi_out");
  return sweeps_scalar_impl<3>(aA, aB, mflags, rhs, psi_out, n_sweeps, dx, dt, make_grid(B, D, H, W),
                               (cudaStream_t)stream, __func__);
}
/* ---- slab tier: one sweep over local planes [zb, zb + zn); psi_out may be NULL ---------------------- */
static int slab_grid_m(const lbmff_slab* s, int B, int D, int H, int W, Grid* g, const char* where) {
  if (!s) return bad_arg(where, "null slab descriptor");
  if (s->Dg < D || s->zoff <= -s->Dg || s->zoff >= s->Dg) return bad_arg(where, "slab: need D <= Dg and |zoff| < Dg");
  if (s->zb < 0 || s->zn < 1 || s->zb + s->zn > D) return bad_arg(where, "slab: plane range outside the local array");
  *g = make_slab_grid(B, D, H, W, s->Dg, s->zoff, s->zb, s->zn);
  return 0;
}
int lbmff_slab_magnetic_source_3d(const float* chi, const uint8_t* mflags, const float* Hx, const float* Hy,
                                  const float* Hz, double dx, float* rhs_out, const lbmff_slab* slab, int B, int D,
                                  int H, int W, lbmff_stream_t stream) {
  Grid g;
  if (int rc = slab_grid_m(slab, B, D, H, W, &g, __func__)) return rc;
  return source_impl<3>(chi, mflags, Hx, Hy, Hz, dx, rhs_out, g, (cudaStream_t)stream, __func__);
}
int lbmff_slab_magnetic_sweep_d3q19(const float* h_in, float* h_out, const uint8_t* mflags, const float* rhs,
                                    float* psi_out, double tau, double dx, double dt, const lbmff_slab* slab, int B,
                                    int D, int H, int W, lbmff_stream_t stream) {
  Grid g;
  if (int rc = slab_grid_m(slab, B, D, H, W, &g, __func__)) return rc;
  return sweeps_impl<3>(const_cast<float*>(h_in), h_out, mflags, rhs, psi_out, 1, tau, dx, dt, g,
                        (cudaStream_t)stream, __func__);
}
int lbmff_slab_magnetic_sweep_scalar_3d(const float* a_cur, float* a_prev_out, const uint8_t* mflags,
                                        const float* rhs, float* psi_out, double dx, double dt,
                                        const lbmff_slab* slab, int B, int D, int H, int W, lbmff_stream_t stream) {
  Grid g;
  if (int rc = slab_grid_m(slab, B, D, H, W, &g, __func__)) return rc;
  return sweeps_scalar_impl<3>(const_cast<float*>(a_cur), a_prev_out, mflags, rhs, psi_out, 1, dx, dt, g,
                               (cudaStream_t)stream, __func__);
}

int lbmff_magnetic_materialize_h_2d(const float* a_cur, const float* a_prev, const float* psi, const float* rhs,
                                    const uint8_t* mflags, float* h_out, double dx, double dt, int B, int H, int W,
                                    lbmff_stream_t stream) {
  return materialize_impl<2>(a_cur, a_prev, psi, rhs, mflags, h_out, dx, dt, B, 1, H, W, (cudaStream_t)stream,
                             __func__);
}
int lbmff_magnetic_materialize_h_3d(const float* a_cur, const float* a_prev, const float* psi, const float* rhs,
                                    const uint8_t* mflags, float* h_out, double dx, double dt, int B, int D, int H,
                                    int W, lbmff_stream_t stream) {
  return materialize_impl<3>(a_cur, a_prev, psi, rhs, mflags, h_out, dx, dt, B, D, H, W, (cudaStream_t)stream,
                             __func__);
}

}  // extern "C"
