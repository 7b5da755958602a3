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

#include <type_traits>
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
// Register cap: five resident blocks per SM in 3-D (48 registers, 1280 threads/SM) measured fastest -- 3.17 ms per
// 512^3 sweep against 3.29 ms uncapped (64 registers) and 3.38 ms at six blocks (40 registers, spills).
// PUSH = +1 / -1 (slab boundary planes, multi-GPU): the populations with e_z = PUSH are ALSO stored into `peer`, the
// neighbour rank's output array mapped over NVLink (symmetric memory), `peer_shift` cells away -- the halo plane that
// neighbour pulls them from.  The sweep of a boundary plane and its halo "send" are then one kernel: no pack, no
// staging buffer, no communication kernel (distributed.SymmPushExchanger).
template <int DIM, bool WRITE_PSI, bool IDX32, int PUSH = 0>
__global__ void __launch_bounds__(kThreads, DIM == 3 ? 5 : 1) k_mag_sweep(const float* __restrict__ hin, float* __restrict__ hout,
                                                        const uint8_t* __restrict__ mflags,
                                                        const float* __restrict__ rhs, float* __restrict__ psi_out,
                                                        Consts k, Grid g, float* __restrict__ peer = nullptr,
                                                        int peer_shift = 0) {
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
  for (int q = 0; q < L::Q; ++q) out[at(q, i)] = hv[q];
  if constexpr (PUSH != 0 && DIM == 3) {
#pragma unroll
    for (int q = 0; q < L::Q; ++q)
      if (L::ez(q) == PUSH) peer[at(q, i + peer_shift)] = hv[q];
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
// The z loop is unrolled by the ring size so that ring slots are compile-time constants (1.09 -> 1.02 ms per 512^3
// sweep) and the registers are capped at 48 for five blocks per SM (0.98 ms; three blocks 1.42 ms, six would spill):
// the kernel is bound by the latency its resident warps can cover -- a two-plane-deep prefetch changed nothing.
constexpr int kTX = 32, kTY = 8, kRing = 4;
constexpr int kTileW = kTX + 2, kTileH = kTY + 2;

template <bool WRITE_PSI>
__global__ void __launch_bounds__(kTX* kTY, 5) k_mag_sweep_scalar_tiled(const float* __restrict__ ca, float* pa,
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
  // running index of this thread's cell (one add per plane) and the loop-invariant part of the shell test
  int i = i0;
  const bool shell_xy = x <= 1 || x >= g.W - 2 || y <= 1 || y >= g.H - 2;
  // One plane.  S = (z - zs) mod kRing is a COMPILE-TIME constant (the z loop below is unrolled by the ring size), so
  // the ring slots of planes z-1, z, z+1 are constants and every stencil read is base + immediate offset.
  auto plane = [&](int z, auto S_) {
    constexpr int S = decltype(S_)::value;
    {  // put(z + 1): slot (S + 2) mod kRing
      constexpr int slot = (S + 2) & (kRing - 1);
#pragma unroll
      for (int t = 0; t < 2; ++t)
        if (off[t] >= 0) {
          sA[slot][sy[t]][sx[t]] = ra[t];
          sD[slot][sy[t]][sx[t]] = rd[t];
        }
    }
    if (z + 1 < ze) fetch(z + 2);
    const float r = r_next;
    const uint8_t fcell = f_next;
    if (live && z + 1 < ze) {
      r_next = __ldg(rhs + i + HW);
      f_next = fl[i + HW];
    }
    __syncthreads();
    if (live) {
      if (fcell != LBMFF_OBSTACLE) {
        constexpr int sl[3] = {S & (kRing - 1), (S + 1) & (kRing - 1), (S + 2) & (kRing - 1)};  // planes z-1, z, z+1
        float hv[L::Q];
        hv[0] = 0.0f;
#pragma unroll
        for (int q = 1; q < L::Q; ++q) {
          const bool axis = (L::ex(q) != 0) + (L::ey(q) != 0) + (L::ez(q) != 0) == 1;
          const int px = tx + 1 - L::ex(q), py = ty + 1 - L::ey(q), ps = sl[1 - L::ez(q)];
          hv[q] = axis ? sA[ps][py][px] : sD[ps][py][px];
        }
        // walls lie on the outer shell only (checked by the host), so only shell-adjacent cells test flags
        if (shell_xy || near_z_shell(g, z)) {
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
    i += HW;
  };
  for (int z = zs; z < ze; z += kRing) {  // zs + multiple of kRing: plane z + S has ring phase S
    plane(z, std::integral_constant<int, 0>{});
    if (z + 1 < ze) plane(z + 1, std::integral_constant<int, 1>{});
    if (z + 2 < ze) plane(z + 2, std::integral_constant<int, 2>{});
    if (z + 3 < ze) plane(z + 3, std::integral_constant<int, 3>{});
  }
}

// ---- tau == 1 scalar sweep in psi-only form (3-D, inside one multi-sweep call) ------------------------------------
// The two values a cell keeps are both functions of its psi and its source: a = mag_collide_q(1 | 5, 0, psi, rhs).
// Inside a call that runs n >= 3 sweeps the intermediate states therefore need not be stored as two floats per cell:
// the sweep reads psi (+ the sweep-invariant rhs) of the neighbours, rebuilds their two values WHILE STAGING the plane
// into shared memory (once per staged element, not once per use), and writes psi only -- 12 B per cell and sweep
// (psi read, rhs read, psi written) instead of 20, and no per-cell collision / second store.  The arithmetic is the
// same uncontracted mag_collide_q / psi_of sequence, so every psi is bit-identical to k_mag_sweep_scalar_tiled's.
//   NBR  = 0: neighbour values from the two-float state arrays `src` = [2][N]       (first sweep of a call)
//        = 1: rebuilt from psi `src` = [N] and rhs
//   WALL = 0: the cell's own values of two sweeps ago from the state arrays `dst` = [2][N] (first two sweeps)
//        = 1: rebuilt from psi `dst` = [N] (the buffer this sweep overwrites) and rhs
// Output: psi -> dst[i] (same-thread read-before-write, as in k_mag_sweep_scalar).  sweeps_scalar_impl chains
// (0,0) -> (1,0) -> (1,1)... and k_mag_psi_expand turns the last two psi fields back into the state arrays the next
// call, the slab exchange and materialize_h expect -- bit-identical to what n old-style sweeps leave behind.
//   MASK = 0: the wall mask of a shell-adjacent cell is built from 18 flag loads (~250 instructions executed by the
//             whole warp for its one or two shell-adjacent lanes: 40 % of all instructions of a 512^3 sweep) and
//             parked, as a bit pattern, in wmask[i] -- the diagonal half of the older state array, dead from the
//             first sweep of the call until k_mag_psi_expand rewrites it;
//        = 1: the mask is read back from wmask[i] (one load).
template <int NBR, int WALL, int MASK, int R>
__global__ void __launch_bounds__(kTX* kTY, R == 1 ? 4 : 3) k_mag_sweep_psi_tiled(const float* __restrict__ src, float* dst,
                                                                               const uint8_t* __restrict__ fl,
                                                                               const float* __restrict__ rhs,
                                                                               float* wmask, Consts k, Grid g, int zc,
                                                                               float* peer = nullptr, int peer_shift = 0) {
  using L = Lat<3>;
  // peer != nullptr (slab boundary planes, multi-GPU): psi is also stored into the neighbour rank's array, peer_shift
  // cells away -- its halo plane (the fused halo push of distributed.SymmPushExchanger)
  // 32 x 8 threads, R rows per thread (rows ty, ty + 8, ...): the per-plane bookkeeping (staging addresses, loop
  // control, barrier) is shared by R cells -- it was two thirds of the instructions with one row per thread
  constexpr int kRows = kTY * R, kTH = kRows + 2, kTile = kTH * kTileW;
  constexpr int kStage = (kTile + kTX * kTY - 1) / (kTX * kTY);   // staged tile elements per thread
  __shared__ float sA[kRing][kTile];  // axis-class value
  __shared__ float sD[kRing][kTile];  // diagonal-class value
  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * kTX + tx;
  const int x0 = blockIdx.x * kTX, y0 = blockIdx.y * kRows;
  const int zs = g.zb + blockIdx.z * zc;
  const int ze = min(zs + zc, g.zb + g.zn);
  const int HW = g.H * g.W, N = (int)g.N;
  // tile element e = tid + t * 256 of the row-major rim-extended tile: in-plane offset, periodic in x and y
  int off[kStage];
#pragma unroll
  for (int t = 0; t < kStage; ++t) {
    const int e = tid + t * kTX * kTY;
    const int sy = e / kTileW, sx = e - sy * kTileW;
    int gx = x0 - 1 + sx, gy = y0 - 1 + sy;
    gx = gx < 0 ? gx + g.W : gx; gy = gy < 0 ? gy + g.H : gy;
    while (gx >= g.W) gx -= g.W;
    while (gy >= g.H) gy -= g.H;
    off[t] = e < kTile ? gy * g.W + gx : -1;
  }
  // raw registers of the planes in flight, TWO register sets (planes z + 1 and z + 2 are on their way while plane z is
  // computed: with one set the top stall was the wait for the plane fetched one iteration earlier): (axis, diagonal)
  // values when NBR == 0, (psi, rhs) when NBR == 1.  The set is a compile-time function of the ring phase.
  float ra[2][kStage], rd[2][kStage];
  auto fetch = [&](int z, auto SET_) {
    constexpr int SET = decltype(SET_)::value;
    const int zz = z < 0 ? z + g.D : (z >= g.D ? z - g.D : z);
    const int zo = zz * HW;
#pragma unroll
    for (int t = 0; t < kStage; ++t)
      if (t < kStage - 1 || off[t] >= 0) {
        const int j = zo + off[t];
        ra[SET][t] = __ldg(src + j);
        rd[SET][t] = NBR == 0 ? __ldg(src + N + j) : __ldg(rhs + j);
      }
  };
  auto put_slot = [&](int slot, auto SET_) {
    constexpr int SET = decltype(SET_)::value;
#pragma unroll
    for (int t = 0; t < kStage; ++t)
      if (t < kStage - 1 || off[t] >= 0) {
        const int e = tid + t * kTX * kTY;
        sA[slot][e] = NBR == 0 ? ra[SET][t] : mag_collide_q<3>(1, 0.0f, ra[SET][t], rd[SET][t], k);
        sD[slot][e] = NBR == 0 ? rd[SET][t] : mag_collide_q<3>(5, 0.0f, ra[SET][t], rd[SET][t], k);
      }
  };
  using Set0 = std::integral_constant<int, 0>;
  using Set1 = std::integral_constant<int, 1>;
  fetch(zs - 1, Set0{}); put_slot(0, Set0{});
  fetch(zs, Set0{}); put_slot(1, Set0{});
  fetch(zs + 1, Set1{});   // phase S puts plane z + 1 from set (S + 1) & 1 and refills it with plane z + 3
  fetch(zs + 2, Set0{});
  const int x = x0 + tx;
  bool live[R], shell_xy[R];
  uint8_t f_next[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int y = y0 + ty + kTY * r;
    live[r] = x < g.W && y < g.H;
    shell_xy[r] = x <= 1 || x >= g.W - 2 || y <= 1 || y >= g.H - 2;
    f_next[r] = live[r] ? fl[zs * HW + y * g.W + x] : (uint8_t)LBMFF_OBSTACLE;
  }
  int i0 = zs * HW + (y0 + ty) * g.W + x;   // cell of row 0 in plane z (dereferenced only where live)
  const int rowstep = kTY * g.W;
  // rows next to the x / y shell (thread-constant) need, on every plane, their parked wall mask and their own state of
  // two sweeps ago: three dependent DRAM round trips if loaded at the point of use, and the blocks on the x rim (one
  // in eight) consist of such warps only.  They ride one plane ahead like the flag.
  float wm_next[R], wa_next[R], wb_next[R];
  auto wall_prefetch = [&](int r, int i) {
    if (MASK == 1) wm_next[r] = wmask[i];
    wa_next[r] = dst[i];
    wb_next[r] = WALL == 0 ? dst[N + i] : __ldg(rhs + i);
  };
#pragma unroll
  for (int r = 0; r < R; ++r) {
    wm_next[r] = 0.f; wa_next[r] = 0.f; wb_next[r] = 0.f;
    if (live[r] && shell_xy[r]) wall_prefetch(r, i0 + r * rowstep);
  }
  auto plane = [&](int z, auto S_) {
    constexpr int S = decltype(S_)::value;
    using Set = std::integral_constant<int, (S + 1) & 1>;
    put_slot((S + 2) & (kRing - 1), Set{});   // plane z + 1
    if (z + 2 < ze) fetch(z + 3, Set{});
    uint8_t fcell[R];
    float wm[R], wa[R], wb[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      fcell[r] = f_next[r];
      wm[r] = wm_next[r]; wa[r] = wa_next[r]; wb[r] = wb_next[r];
      if (live[r] && z + 1 < ze) {
        f_next[r] = fl[i0 + HW + r * rowstep];
        if (shell_xy[r]) wall_prefetch(r, i0 + HW + r * rowstep);
      }
    }
    __syncthreads();
    constexpr int sl[3] = {S & (kRing - 1), (S + 1) & (kRing - 1), (S + 2) & (kRing - 1)};  // planes z-1, z, z+1
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (fcell[r] == LBMFF_OBSTACLE) continue;   // also rows that are not live
      const int i = i0 + r * rowstep;
      float hv[L::Q];
      hv[0] = 0.0f;
#pragma unroll
      for (int q = 1; q < L::Q; ++q) {
        const bool axis = (L::ex(q) != 0) + (L::ey(q) != 0) + (L::ez(q) != 0) == 1;
        const int e = (ty + kTY * r + 1 - L::ey(q)) * kTileW + (tx + 1 - L::ex(q));
        const int ps = sl[1 - L::ez(q)];
        hv[q] = axis ? sA[ps][e] : sD[ps][e];
      }
      if (shell_xy[r] || near_z_shell(g, z)) {
        unsigned wall_mask = 0;
        // own state of two sweeps ago: prefetched for rows next to the x / y shell, loaded here on the two planes
        // next to the z shell
        const float va = shell_xy[r] ? wa[r] : dst[i];
        const float vb = shell_xy[r] ? wb[r] : (WALL == 0 ? dst[N + i] : __ldg(rhs + i));
        if (MASK == 0) {
          const Nbr nb(PlaneCell{x, y0 + ty + kTY * r, z, i, true}, g);
#pragma unroll
          for (int q = 1; q < L::Q; ++q)
            if (fl[i + nb.at<3>(q, 1)] == LBMFF_OBSTACLE) wall_mask |= 1u << q;
          wmask[i] = __uint_as_float(wall_mask);   // aliases dst + N in the first sweep: vb was read above
        } else {
          wall_mask = __float_as_uint(shell_xy[r] ? wm[r] : wmask[i]);
        }
        if (wall_mask) {
          float p_axis, p_diag;
          if (WALL == 0) {
            p_axis = va; p_diag = vb;
          } else {
            p_axis = mag_collide_q<3>(1, 0.0f, va, vb, k);
            p_diag = mag_collide_q<3>(5, 0.0f, va, vb, k);
          }
#pragma unroll
          for (int q = 1; q < L::Q; ++q) {
            const bool axis = (L::ex(q) != 0) + (L::ey(q) != 0) + (L::ez(q) != 0) == 1;
            if ((wall_mask >> q) & 1u) hv[q] = axis ? p_axis : p_diag;
          }
        }
      }
      const float psi_new = psi_of<3>(hv);
      dst[i] = psi_new;
      if (peer) peer[i + peer_shift] = psi_new;
    }
    i0 += HW;
  };
  for (int z = zs; z < ze; z += kRing) {
    plane(z, std::integral_constant<int, 0>{});
    if (z + 1 < ze) plane(z + 1, std::integral_constant<int, 1>{});
    if (z + 2 < ze) plane(z + 2, std::integral_constant<int, 2>{});
    if (z + 3 < ze) plane(z + 3, std::integral_constant<int, 3>{});
  }
}

// psi of the last two sweeps (newest in cur[0:N], the one before in prev[0:N]) -> the two-float state arrays, psi_out.
// OBSTACLE cells keep what the arrays held (the old-style sweeps never write them either).
__global__ void __launch_bounds__(kThreads) k_mag_psi_expand(float* __restrict__ cur, float* __restrict__ prev,
                                                             const uint8_t* __restrict__ fl,
                                                             const float* __restrict__ rhs,
                                                             float* __restrict__ psi_out, Consts k, long long i0,
                                                             long long i1, long long N) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = i0 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < i1; i += stride) {
    if (fl[i] == LBMFF_OBSTACLE) continue;
    const float pn = cur[i], pm = prev[i], r = rhs[i];
    cur[i] = mag_collide_q<3>(1, 0.0f, pn, r, k);
    cur[N + i] = mag_collide_q<3>(5, 0.0f, pn, r, k);
    prev[i] = mag_collide_q<3>(1, 0.0f, pm, r, k);
    prev[N + i] = mag_collide_q<3>(5, 0.0f, pm, r, k);
    if (psi_out) psi_out[i] = pn;
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
static void launch_sweep(bool wp, const PlaneLaunch& pl, cudaStream_t st, const float* in, float* out,
                         const uint8_t* mflags, const float* rhs, float* psi_out, const Consts& k, const Grid& g) {
  const long long Q = Lat<DIM>::Q;
  const bool idx32 = Q * g.N < 0xffffffffLL;
#define LBMFF_SW(WP, IX)                                                                                    \
  k_mag_sweep<DIM, WP, IX><<<pl.grid, pl.block, 0, st>>>(in + b * Q * g.N, out + b * Q * g.N, mflags + b * g.N, \
                                                         rhs + b * g.N, psi_out ? psi_out + b * g.N : nullptr, k, g)
  for (long long b = 0; b < g.B; ++b) {
    if (idx32) { if (wp) LBMFF_SW(true, true); else LBMFF_SW(false, true); }
    else { if (wp) LBMFF_SW(true, false); else LBMFF_SW(false, false); }
  }
#undef LBMFF_SW
}

template <int DIM>
static int sweeps_impl(float* hA, float* hB, const uint8_t* mflags, const float* rhs, float* psi_out, int n,
                       double tau, double dx, double dt, const Grid& g, cudaStream_t st, const char* where) {
  if (!(hA && hB && hA != hB && mflags && rhs)) return bad_arg(where, "null or aliased pointer");
  if (n < 1) return bad_arg(where, "n_sweeps must be >= 1");
  if (!(tau > 0)) return bad_arg(where, "tau must be positive");
  if (g.N > 0x7fffffffLL) return bad_arg(where, "more than 2^31-1 cells per batch item");
  const Consts k = make_consts(dx, dt, tau);
  const PlaneLaunch pl = plane_launch(g);
  for (int s = 0; s < n; ++s) {
    const float* in = (s & 1) ? hB : hA;
    float* out = (s & 1) ? hA : hB;
    launch_sweep<DIM>(s == n - 1 && psi_out, pl, st, in, out, mflags, rhs, psi_out, k, g);
  }
  return check_launch(where);
}

// one psi-only sweep (3-D) on the plane range of g; nbr / wall / mask: the template modes of k_mag_sweep_psi_tiled
static int psi_sweep_launch(const float* src, float* dst, const uint8_t* fl, const float* rhs, float* wmask, int nbr,
                            int wall, int mask, const Consts& k, const Grid& g, cudaStream_t st, const char* where,
                            float* peer = nullptr, int peer_shift = 0) {
  const int zc = g.zn < 32 ? g.zn : 32;
  const char* er = getenv("LBMFF_MAG_PSI_ROWS");
  const int R = er ? atoi(er) : 2;
  const int rows = kTY * (R == 1 || R == 4 ? R : 2);
  const dim3 tg((g.W + kTX - 1) / kTX, (g.H + rows - 1) / rows, (g.zn + zc - 1) / zc), tb(kTX, kTY);
#define LBMFF_PSI_R(NB, WL, MK)                                                                              \
  do {                                                                                                      \
    if (R == 1) k_mag_sweep_psi_tiled<NB, WL, MK, 1><<<tg, tb, 0, st>>>(src, dst, fl, rhs, wmask, k, g, zc, peer, peer_shift); \
    else if (R == 4) k_mag_sweep_psi_tiled<NB, WL, MK, 4><<<tg, tb, 0, st>>>(src, dst, fl, rhs, wmask, k, g, zc, peer, peer_shift); \
    else k_mag_sweep_psi_tiled<NB, WL, MK, 2><<<tg, tb, 0, st>>>(src, dst, fl, rhs, wmask, k, g, zc, peer, peer_shift);      \
  } while (0)
  if (nbr == 0 && wall == 0 && mask == 0) LBMFF_PSI_R(0, 0, 0);
  else if (nbr == 1 && wall == 0 && mask == 1) LBMFF_PSI_R(1, 0, 1);
  else if (nbr == 1 && wall == 1 && mask == 1) LBMFF_PSI_R(1, 1, 1);
  else return bad_arg(where, "psi sweep: (nbr, wall, mask) must be (0,0,0), (1,0,1) or (1,1,1)");
#undef LBMFF_PSI_R
  return 0;
}

template <int DIM>
static int sweeps_scalar_impl(float* aA, float* aB, const uint8_t* mflags, const float* rhs, float* psi_out, int n,
                              double dx, double dt, const Grid& g, cudaStream_t st, const char* where) {
  if (!(aA && aB && aA != aB && mflags && rhs)) return bad_arg(where, "null or aliased pointer");
  if (n < 1) return bad_arg(where, "n_sweeps must be >= 1");
  if (g.N > 0x7fffffffLL / 2) return bad_arg(where, "more than (2^31-1)/2 cells per batch item");
  const Consts k = make_consts(dx, dt, 1.0);
  const PlaneLaunch pl = plane_launch(g);
  if constexpr (DIM == 3) {
    // n >= 3 sweeps of a whole (non-slab) lattice in one call: psi-only intermediate states (k_mag_sweep_psi_tiled),
    // 12 instead of 20 B per cell and sweep; LBMFF_MAG_PSI=0 keeps the two-float sweeps
    const char* e = getenv("LBMFF_MAG_PSI");
    if (n >= 3 && g.W >= 4 && g.H >= 4 && g.zb == 0 && g.zn == g.D && g.Dg == g.D && !(e && e[0] == '0')) {
      for (long long b = 0; b < g.B; ++b) {
        float* A = aA + b * 2 * g.N;
        float* B = aB + b * 2 * g.N;
        const uint8_t* fl = mflags + b * g.N;
        const float* r = rhs + b * g.N;
        float* wm = B + g.N;   // diagonal half of the older state array: dead after the first sweep reads it
        psi_sweep_launch(A, B, fl, r, wm, 0, 0, 0, k, g, st, where);   // state in A (B: one older) -> psi_1 in B
        psi_sweep_launch(B, A, fl, r, wm, 1, 0, 1, k, g, st, where);   // psi_1, walls from state 0 in A -> psi_2 in A
        for (int s = 2; s < n; ++s) {                                  // psi_{s+1} over psi_{s-1}
          if (s & 1) psi_sweep_launch(B, A, fl, r, wm, 1, 1, 1, k, g, st, where);
          else psi_sweep_launch(A, B, fl, r, wm, 1, 1, 1, k, g, st, where);
        }
        float* cur = (n & 1) ? B : A;       // newest psi: sweep 1 wrote B, sweep 2 A, ...
        float* prev = (n & 1) ? A : B;
        const unsigned eb = blocks_for(g.N) > 8192u ? 8192u : blocks_for(g.N);
        k_mag_psi_expand<<<eb, kThreads, 0, st>>>(
            cur, prev, fl, r, psi_out ? psi_out + b * g.N : nullptr, k, 0, g.N, g.N);
      }
      return check_launch(where);
    }
  }
  for (int s = 0; s < n; ++s) {
    const float* cur = (s & 1) ? aB : aA;
    float* prev = (s & 1) ? aA : aB;
    for (long long b = 0; b < g.B; ++b) {
      if (DIM == 3 && g.W >= 4 && g.H >= 4) {  // tiled kernel: 0.98 ms per 512^3 sweep; one thread per cell: 2.50 ms
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
int lbmff_slab_magnetic_sweep_push_d3q19(const float* h_in, float* h_out, const uint8_t* mflags, const float* rhs,
                                         float* psi_out, double tau, double dx, double dt, float* peer_out,
                                         int direction, long long peer_shift_cells, const lbmff_slab* slab, int B,
                                         int D, int H, int W, lbmff_stream_t stream) {
  Grid g;
  if (int rc = slab_grid_m(slab, B, D, H, W, &g, __func__)) return rc;
  if (!(h_in && h_out && h_in != h_out && mflags && rhs && peer_out)) return bad_arg(__func__, "null or aliased pointer");
  if (B != 1) return bad_arg(__func__, "B must be 1");
  if (direction != 1 && direction != -1) return bad_arg(__func__, "direction must be +1 or -1");
  if (!(tau > 0)) return bad_arg(__func__, "tau must be positive");
  if (19 * g.N >= 0xffffffffLL) return bad_arg(__func__, "lattice too large for 32-bit population indices");
  if (peer_shift_cells <= -g.N || peer_shift_cells >= g.N) return bad_arg(__func__, "peer shift outside the array");
  const Consts k = make_consts(dx, dt, tau);
  const PlaneLaunch pl = plane_launch(g);
  cudaStream_t st = (cudaStream_t)stream;
  const int sh = (int)peer_shift_cells;
#define LBMFF_PUSH(WP, DIR) \
  k_mag_sweep<3, WP, true, DIR><<<pl.grid, pl.block, 0, st>>>(h_in, h_out, mflags, rhs, psi_out, k, g, peer_out, sh)
  if (psi_out) { if (direction > 0) LBMFF_PUSH(true, 1); else LBMFF_PUSH(true, -1); }
  else { if (direction > 0) LBMFF_PUSH(false, 1); else LBMFF_PUSH(false, -1); }
#undef LBMFF_PUSH
  return check_launch(__func__);
}
int lbmff_slab_magnetic_sweep_scalar_3d(const float* a_cur, float* a_prev_out, const uint8_t* mflags,
                                        const float* rhs, float* psi_out, double dx, double dt,
                                        const lbmff_slab* slab, int B, int D, int H, int W, lbmff_stream_t stream) {
  Grid g;
  if (int rc = slab_grid_m(slab, B, D, H, W, &g, __func__)) return rc;
  return sweeps_scalar_impl<3>(const_cast<float*>(a_cur), a_prev_out, mflags, rhs, psi_out, 1, dx, dt, g,
                               (cudaStream_t)stream, __func__);
}

/* psi-only form of the tau == 1 sweeps on a plane range (see k_mag_sweep_psi_tiled): ONE sweep; the caller chains the
 * modes (0,0,0) -> (1,0,1) -> (1,1,1)... and exchanges the psi halves between sweeps */
int lbmff_slab_magnetic_sweep_psi_3d(const float* src, float* dst, const uint8_t* mflags, const float* rhs,
                                     float* wmask, int nbr_mode, int wall_mode, int mask_mode, double dx, double dt,
                                     float* peer_dst, long long peer_shift_cells, const lbmff_slab* slab, int B, int D,
                                     int H, int W, lbmff_stream_t stream) {
  Grid g;
  if (int rc = slab_grid_m(slab, B, D, H, W, &g, __func__)) return rc;
  if (!(src && dst && src != dst && mflags && rhs && wmask)) return bad_arg(__func__, "null or aliased pointer");
  if (B != 1) return bad_arg(__func__, "B must be 1");
  if (W < 4 || H < 4) return bad_arg(__func__, "W and H must be >= 4");
  if (g.N > 0x7fffffffLL / 2) return bad_arg(__func__, "more than (2^31-1)/2 cells per batch item");
  if (peer_dst && (peer_shift_cells <= -g.N || peer_shift_cells >= g.N)) return bad_arg(__func__, "peer shift outside the array");
  if (int rc = psi_sweep_launch(src, dst, mflags, rhs, wmask, nbr_mode, wall_mode, mask_mode,
                                make_consts(dx, dt, 1.0), g, (cudaStream_t)stream, __func__, peer_dst,
                                (int)peer_shift_cells))
    return rc;
  return check_launch(__func__);
}
int lbmff_slab_magnetic_psi_expand_3d(float* a_cur, float* a_prev, const uint8_t* mflags, const float* rhs,
                                      float* psi_out, double dx, double dt, const lbmff_slab* slab, int B, int D,
                                      int H, int W, lbmff_stream_t stream) {
  Grid g;
  if (int rc = slab_grid_m(slab, B, D, H, W, &g, __func__)) return rc;
  if (!(a_cur && a_prev && a_cur != a_prev && mflags && rhs)) return bad_arg(__func__, "null or aliased pointer");
  if (B != 1) return bad_arg(__func__, "B must be 1");
  const unsigned eb = blocks_for(g.i1 - g.i0) > 8192u ? 8192u : blocks_for(g.i1 - g.i0);
  k_mag_psi_expand<<<eb, kThreads, 0, (cudaStream_t)stream>>>(a_cur, a_prev, mflags, rhs, psi_out,
                                                            make_consts(dx, dt, 1.0), g.i0, g.i1, g.N);
  return check_launch(__func__);
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
