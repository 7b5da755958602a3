// Fused tier, hydrodynamic part of one outer step (demo_3d_LBM_Rosensweig_instability.py:151-209):
//
//  k_hcz_moments        rho = sum_q f_q(x - e_q) on non-OBSTACLE cells          (stream + macro_compute)
//  k_hcz_fields         after the ordered wetting kernels: lap(density), fai, prho  (capillary_process, part 1)
//  k_hcz_collide_stream pull f,g; bounce on OBSTACLE; on FLUID: forces, vel, pressure, f and g collisions;
//                       write the other f,g buffers (stream + bounce + capillary_process part 2 + collision)
//  k_stream_collide     single-phase D2Q9 / D3Q19: stream + macro + bounce + BGK / KBC in one pass
//
// Populations stay in the reference's [Q][z][y][x] SoA layout: the thread index runs along x, so every
// population load (shifted by at most one cell) and store is unit-stride.  State between steps is the
// post-collision, pre-stream f,g -- the same state the reference holds at the end of a step.
#include <stdlib.h>

#include "capi.cuh"
#include "capillary.cuh"
#include "collide.cuh"
#include "hcz_fast.cuh"
#include "tma.cuh"

namespace lbmff {

// Additive periodic-neighbour offsets: index(x - s e_q) = zo[...] + yo[...] + xo[...]
template <int DIM>
struct PullIndex {
  long long zo[3], yo[3];
  int xo[3];
  __device__ __forceinline__ PullIndex(const Cell& c, const Grid& g) {
    const long long HW = (long long)g.H * g.W;
    xo[0] = wrap(c.x - 1, g.W); xo[1] = c.x; xo[2] = wrap(c.x + 1, g.W);
    yo[0] = (long long)wrap(c.y - 1, g.H) * g.W; yo[1] = (long long)c.y * g.W; yo[2] = (long long)wrap(c.y + 1, g.H) * g.W;
    zo[0] = DIM == 3 ? wrap(c.z - 1, g.D) * HW : 0; zo[1] = c.z * HW; zo[2] = DIM == 3 ? wrap(c.z + 1, g.D) * HW : 0;
  }
  __device__ __forceinline__ long long at(int q, int s) const {
    using L = Lat<DIM>;
    return zo[1 - s * L::ez(q)] + yo[1 - s * L::ey(q)] + xo[1 - s * L::ex(q)];
  }
};

// rho = sum of the pulled f on non-OBSTACLE cells (old rho kept on OBSTACLE, LBM_macro_compute_3d.py:93-98) and,
// for the magnetic subsystem, chi = k (1 - H_eps(phi(density(rho)))) on EVERY cell (demo :162, LBM_magnetic_3d.py:162)
// Launch geometry / addressing: common.cuh (plane_launch, Nbr, PopIndex); pulls are issued before the flag.
template <int DIM, bool IDX32>
__global__ void __launch_bounds__(kThreads) k_hcz_moments(const float* __restrict__ f,
                                                          const uint8_t* __restrict__ flags, float* __restrict__ rho,
                                                          float* __restrict__ chi, Consts k, Grid g) {
  using L = Lat<DIM>;
  const PlaneCell c = plane_cell(g);
  if (!c.live) return;
  const int i = c.i;
  const Nbr nb(c, g);
  const PopIndex<IDX32> at(g);
  float fv[L::Q];
#pragma unroll
  for (int q = 0; q < L::Q; ++q) fv[q] = ld_spec(f + at(q, i + nb.at<DIM>(q, 1)));
  float s;
  if (flags[i] == LBMFF_OBSTACLE) {
    s = rho[i];
  } else {
    s = 0.0f;
#pragma unroll
    for (int q = 0; q < L::Q; ++q) s += fv[q];
    rho[i] = s;
  }
  if (chi) chi[i] = susceptibility(phase_field(affine_density(s, k.aff), k.aff), k.sm);
}

// density = affine(rho) sampled at constant offsets from an interior centre cell; the division by
// (rho_liquid - rho_gas) is a multiplication by the host-formed reciprocal (differs from affine_density by
// one rounding; held to the fp32-floor parity bar like the rest of the fused step).
struct CentredDensity {
  const float* __restrict__ p;  // rho, offset to the centre cell
  int W, HW;
  int cx, cy, cz;
  Affine A;
  __device__ __forceinline__ float operator()(int x, int y, int z) const {
    const float r = __ldg(p + ((x - cx) + (y - cy) * W + (z - cz) * HW));
    return A.dg + A.ddiff * ((r - A.rg) * A.inv_rdiff);
  }
};

// ws = [lap | fai | prho], N floats each, every field with the get_grad shell rule folded into the data:
// a shell cell holds what it contributes to a gradient stencil (OBSTACLE: nearest interior value; other
// shell cells: their own value, which is 0 for lap and the replicate-padded interior value for fai / prho).
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_hcz_fields(const float* __restrict__ rho,
                                                         const uint8_t* __restrict__ flags,
                                                         const float* __restrict__ pressure, float* __restrict__ ws,
                                                         Consts k, Grid g) {
  const PlaneCell c = plane_cell(g);
  if (!c.live) return;
  const int i = c.i;
  const int x = interior(c.x, g.W), y = interior(c.y, g.H), z = DIM == 3 ? interior_z(g, c.z) : 0;
  const int HW = g.H * g.W;
  const int j = i + (x - c.x) + (y - c.y) * g.W + (z - c.z) * HW;   // nearest interior cell
  const bool shell = j != i;
  float lap = 0.0f;
  // (slab decomposition: a halo plane that is a global z-shell plane clamps onto the outermost local plane,
  //  whose stencil would leave the array; its lap is never consumed, so it is not evaluated)
  const bool inside = DIM == 2 || (z >= 1 && z <= g.D - 2);
  if (inside && (!shell || flags[i] == LBMFF_OBSTACLE))
    lap = laplacian_at<DIM>(CentredDensity{rho + j, g.W, HW, x, y, z, k.aff}, x, y, z, k.dx2);
  ws[i] = lap;
  const float rj = rho[j];
  ws[g.N + i] = eos_cs(rj, k.cs2, k.a_eos, 1) - rj * k.cs2;
  // density of the previous macro_compute at an interior cell = affine(rho) there (wetting touches the shell only)
  ws[2 * g.N + i] = pressure[j] - k.cs2 * affine_density(rj, k.aff);
}

// psi on OBSTACLE shell cells of mflags <- nearest interior value (what get_grad(psi) reads there);
// wall-cell psi is never observed otherwise.  Surface kernels: k_shell_rule_rim visits the x / y rim of every
// plane of the range (planes on the global z shell are skipped), k_shell_rule_plane one whole z-shell plane.
template <int DIM>
__device__ __forceinline__ void shell_rule_cell(float* __restrict__ field, const uint8_t* __restrict__ flags,
                                                int x, int y, int z, const Grid& g) {
  const long long i = lin(g, x, y, z);
  if (flags[i] != LBMFF_OBSTACLE) return;
  field[i] = field[lin(g, interior(x, g.W), interior(y, g.H), DIM == 3 ? interior_z(g, z) : 0)];
}

template <int DIM>
__global__ void __launch_bounds__(kThreads) k_shell_rule_rim(float* __restrict__ field,
                                                             const uint8_t* __restrict__ flags, int zlo, int zn,
                                                             Grid g) {
  const int rim = 2 * g.W + 2 * (g.H - 2);
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)zn * rim) return;
  const int z = zlo + (int)(t / rim), k = (int)(t % rim);
  if (DIM == 3 && z_on_shell(g, z)) return;
  int x, y;
  if (k < g.W) { x = k; y = 0; }
  else if (k < 2 * g.W) { x = k - g.W; y = g.H - 1; }
  else { y = 1 + (k - 2 * g.W) / 2; x = ((k - 2 * g.W) & 1) ? g.W - 1 : 0; }
  const long long b = blockIdx.y;
  shell_rule_cell<DIM>(field + b * g.N, flags + b * g.N, x, y, z, g);
}

template <int DIM>
__global__ void __launch_bounds__(kThreads) k_shell_rule_plane(float* __restrict__ field,
                                                               const uint8_t* __restrict__ flags, int z, Grid g) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)g.H * g.W) return;
  const long long b = blockIdx.y;
  shell_rule_cell<DIM>(field + b * g.N, flags + b * g.N, (int)(t % g.W), (int)(t / g.W), z, g);
}

template <int DIM>
static void launch_shell_rule(float* field, const uint8_t* flags, const Grid& g, cudaStream_t st) {
  const long long HW = (long long)g.H * g.W;
  const int zlo = (int)(g.i0 / HW), zn = (int)((g.i1 - g.i0) / HW);
  const long long rim = 2LL * g.W + 2LL * (g.H - 2);
  k_shell_rule_rim<DIM><<<dim3(blocks_for(zn * rim), (unsigned)g.B), kThreads, 0, st>>>(field, flags, zlo, zn, g);
  if (DIM == 3 && g.Dg > 1) {
    for (int side = 0; side < 2; ++side) {   // local plane of global plane 0 / Dg-1, when this range holds it
      int l = (side ? g.Dg - 1 : 0) - g.zoff;
      l = l < 0 ? l + g.Dg : (l >= g.Dg ? l - g.Dg : l);
      if (l >= zlo && l < zlo + zn)
        k_shell_rule_plane<DIM><<<dim3(blocks_for(HW), (unsigned)g.B), kThreads, 0, st>>>(field, flags, l, g);
    }
  }
}

struct HextUniform {
  float v[3];
};

// 32-bit constant-offset sampler around an interior centre cell (see CentredField in physics.cuh)
struct Centred32 {
  const float* __restrict__ p;  // already offset to the centre cell
  int W, HW;
  int cx, cy, cz;
  __device__ __forceinline__ float operator()(int x, int y, int z) const {
    return __ldg(p + ((x - cx) + (y - cy) * W + (z - cz) * HW));
  }
};

// Everything of k_hcz_collide_stream after the 2Q pulls, for one cell.  gv: the pulled g populations (registers);
// fv(q): the pulled f population q (registers, or the thread's shared-memory slots in the prefetching kernel,
// where f is first needed in the final loop).
struct CsArgs {
  float* __restrict__ fo; float* __restrict__ go;
  const float* __restrict__ rho_; const uint8_t* __restrict__ flags; const float* __restrict__ ws;
  const float* __restrict__ chi; const float* __restrict__ psi; const float* __restrict__ Hext;
  HextUniform Hu;
  float* __restrict__ vel; float* __restrict__ pressure; float* __restrict__ force_out;
  int kbc_on_g;
};
struct RegPop {
  const float* v;
  __device__ __forceinline__ float operator()(int q) const { return v[q]; }
};
struct SmemPop {   // slot q of this thread: s[q * kThreads]
  const float* s;
  __device__ __forceinline__ float operator()(int q) const { return s[q * kThreads]; }
};

// The five gradients of capillary_process part 2 at a FLUID cell (HCZ_3d.py:228-242): lap, fai, prho and, with the
// magnetic subsystem, chi and psi.  All stencil inputs carry the get_grad shell rule in their data, so the 18-point
// stencils are plain constant-offset reads around the interior-clamped centre.  GlobalStencil reads them from
// global memory (L1/L2); the TMA kernel below supplies a shared-memory sampler with the same arithmetic.
template <int DIM, bool MAGNETIC>
struct GlobalStencil {
  __device__ __forceinline__ void operator()(const PlaneCell& c, const Grid& g, const float* __restrict__ ws,
                                             const float* __restrict__ chi, const float* __restrict__ psi,
                                             float inv12dx, float* glap, float* gfai, float* gprho, float* gchi,
                                             float* gpsi) const {
    const int x = interior(c.x, g.W), y = interior(c.y, g.H), z = DIM == 3 ? interior_z(g, c.z) : 0;
    const int HW = g.H * g.W;
    const int j0 = c.i + (x - c.x) + (y - c.y) * g.W + (z - c.z) * HW;
    const float* w = ws + j0;
    grad_at_fast<DIM>(Centred32{w, g.W, HW, x, y, z}, x, y, z, inv12dx, glap);
    grad_at_fast<DIM>(Centred32{w + g.N, g.W, HW, x, y, z}, x, y, z, inv12dx, gfai);
    grad_at_fast<DIM>(Centred32{w + 2 * g.N, g.W, HW, x, y, z}, x, y, z, inv12dx, gprho);
    if (MAGNETIC) {
      grad_at_fast<DIM>(Centred32{chi + j0, g.W, HW, x, y, z}, x, y, z, inv12dx, gchi);
      grad_at_fast<DIM>(Centred32{psi + j0, g.W, HW, x, y, z}, x, y, z, inv12dx, gpsi);
    }
  }
};

template <int DIM, bool MAGNETIC, bool IDX32, class GV, class FV, class ST>
__device__ __forceinline__ void hcz_cs_cell(const PlaneCell& c, const PopIndex<IDX32>& at, const GV gv, const FV fv,
                                            const uint8_t fl, const float rho, const CsArgs& A, const Consts& k,
                                            const KbcScalars& K, const Grid& g, const ST& stencil) {
  using L = Lat<DIM>;
  const int i = c.i;
  float* __restrict__ fo = A.fo; float* __restrict__ go = A.go;
  const float* __restrict__ ws = A.ws; const float* __restrict__ chi = A.chi; const float* __restrict__ psi = A.psi;
  const float* __restrict__ Hext = A.Hext; const HextUniform& Hu = A.Hu;
  float* __restrict__ vel = A.vel; float* __restrict__ pressure = A.pressure; float* __restrict__ force_out = A.force_out;
  const int kbc_on_g = A.kbc_on_g;
  if (fl != LBMFF_FLUID) {
    // OBSTACLE: stream then full-way bounce, out_q(x) = in_opp(q)(x + e_q) -- the value pulled for opp(q).
    // Other non-FLUID cells are only streamed (HCZ collides FLUID cells only, HCZ_3d.py:357-359).
    const bool obst = fl == LBMFF_OBSTACLE;
#pragma unroll
    for (int q = 0; q < L::Q; ++q) {
      fo[at(q, i)] = obst ? fv(L::opp(q)) : fv(q);
      go[at(q, i)] = obst ? gv(L::opp(q)) : gv(q);
    }
    return;
  }
  // ---- capillary_process part 2 at a FLUID cell (HCZ_3d.py:217-261); all stencil inputs carry the
  // get_grad shell rule in their data, so the 18-point stencils are plain constant-offset loads
  float glap[3], gfai[3], gprho[3], gchi[3], gpsi[3];
  stencil(c, g, ws, chi, psi, k.inv12dx, glap, gfai, gprho, gchi, gpsi);
#ifdef LBMFF_HCZ_FEQ_IEEE
  const float dn = affine_density(rho, k.aff);
#else
  const float dn = k.aff.dg + k.aff.ddiff * ((rho - k.aff.rg) * k.aff.inv_rdiff);   // as CentredDensity above
#endif
  float F[DIM];
#pragma unroll
  for (int a = 0; a < DIM; ++a) {
    F[a] = k.kappa * dn * glap[a];
    if (a == 1) F[a] += k.grav_y * dn;
  }
  if (MAGNETIC) {
    float H2 = 0.0f;
#pragma unroll
    for (int a = 0; a < DIM; ++a) {
      const float he = Hext ? __ldg(Hext + a * g.N + i) : Hu.v[a];
      const float ht = he - gpsi[a];  // H_ext + H_int, H_int = -grad(psi) (LBM_magnetic_3d.py:184)
      H2 += ht * ht;
    }
#pragma unroll
    for (int a = 0; a < DIM; ++a) F[a] += k.kelvin * H2 * gchi[a];
  }
  float s = 0.f, m[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int q = 0; q < L::Q; ++q) {
    const float gq = gv(q);
    s += gq;
    if (L::ex(q) == 1) m[0] += gq; else if (L::ex(q) == -1) m[0] -= gq;
    if (L::ey(q) == 1) m[1] += gq; else if (L::ey(q) == -1) m[1] -= gq;
    if (DIM == 3) { if (L::ez(q) == 1) m[2] += gq; else if (L::ez(q) == -1) m[2] -= gq; }
  }
  float u[DIM], dot = 0.f;
#ifdef LBMFF_HCZ_FEQ_IEEE
  const float inv_dn = k.vel_scale / dn;  // 1 / (cs2 density)
#else
  const float inv_dn = k.vel_scale * nr_rcp(dn);
#endif
#pragma unroll
  for (int a = 0; a < DIM; ++a) {
    u[a] = (m[a] * k.c + k.half_dt_RT * F[a]) * inv_dn;
    vel[a * g.N + i] = u[a];
    if (force_out) force_out[a * g.N + i] = F[a];
    dot += u[a] * gprho[a];
  }
  const float p = s - k.half_dt * dot;
  pressure[i] = p;
  // ---- HCZ collision (HCZ_3d.py:283-361)
  const HczCell<DIM> C = hcz_prepare<DIM>(rho, u, dn, p, F, gfai, gprho, k);
  bool relaxed = false;
  float gk[DIM == 2 ? 9 : 1];
  if constexpr (DIM == 2) {
    if (kbc_on_g) {  // entropic KBC relaxation of g (HCZ_2d.py:244-256)
      float geq[9], gl[9];
#pragma unroll
      for (int q = 0; q < 9; ++q) { geq[q] = hcz_geq_q<2>(q, C); gl[q] = gv(q); }
      kbc_relax(gl, geq, rho, u, k.c, kbc_on_g == 1 ? 1 : 0, K, gk);
      relaxed = true;
    }
  }
  {   // the rest population, then opposite pairs (hcz_collide_pair shares what e_p = -e_q makes common)
    float f0 = fv(0);
    float g0 = (DIM == 2 && relaxed) ? gk[0] : gv(0);
    hcz_collide_q<DIM>(0, C, k, f0, g0, relaxed);
    fo[at(0, i)] = f0;
    go[at(0, i)] = g0;
  }
#pragma unroll
  for (int q = 1; q < L::Q; ++q) {
    const int p = L::opp(q);
    if (q > p) continue;
    float fq = fv(q), fp = fv(p);
    float gq = (DIM == 2 && relaxed) ? gk[DIM == 2 ? q : 0] : gv(q);
    float gp = (DIM == 2 && relaxed) ? gk[DIM == 2 ? p : 0] : gv(p);
    hcz_collide_pair<DIM>(q, C, k, fq, gq, fp, gp, relaxed);
    fo[at(q, i)] = fq;
    go[at(q, i)] = gq;
    fo[at(p, i)] = fp;
    go[at(p, i)] = gp;
  }
}

template <int DIM, bool MAGNETIC, bool IDX32>
__global__ void __launch_bounds__(kThreads, 2)
k_hcz_collide_stream(const float* __restrict__ fi, const float* __restrict__ gi, CsArgs A, Consts k, KbcScalars K,
                     Grid g) {
  using L = Lat<DIM>;
  const PlaneCell c = plane_cell(g);
  if (!c.live) return;
  const int i = c.i;
  const Nbr nb(c, g);
  const PopIndex<IDX32> at(g);
  // All 2Q pulls are issued first, before the flag byte is known: they are the only HBM-latency loads of the
  // kernel (the stencil fields below mostly hit L1/L2); at 512 threads/SM this keeps ~78 KB per SM in flight.
  float gv[L::Q], fv[L::Q];
#pragma unroll
  for (int q = 0; q < L::Q; ++q) {
    const auto j = at(q, i + nb.at<DIM>(q, 1));
    gv[q] = ld_spec(gi + j);
    fv[q] = ld_spec(fi + j);
  }
  const float rho = A.rho_[i];   // (harmless on non-FLUID cells, and not behind the flag's round trip)
  hcz_cs_cell<DIM, MAGNETIC, IDX32>(c, at, RegPop{gv}, RegPop{fv}, A.flags[i], rho, A, k, K, g,
                                    GlobalStencil<DIM, MAGNETIC>{});
}

// The same step with the 2Q pulls of the NEXT cell in flight while the current one is computed: a block marches
// along z over `kz` planes; each thread copies the 38 populations its next cell pulls into its own shared-memory
// slots with cp.async (4-byte LDGSTS: the pulls are shifted by one cell, so wider or bulk copies would be
// misaligned) and waits only for its own copies -- no block barrier, the slots are thread-private.  Two stages of
// 2Q x 256 floats = 76 KB per block, two blocks per SM.  Marching along z also turns the z-1 / z / z+1 planes of
// the five stencil fields into L1 re-reads of the same block.
__device__ __forceinline__ void cp_async4(float* smem, const float* gmem) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <bool MAGNETIC, bool IDX32>
__global__ void __launch_bounds__(kThreads, 2)
k_hcz_collide_stream_pf(const float* __restrict__ fi, const float* __restrict__ gi, CsArgs A, Consts k, KbcScalars K,
                        Grid g, int kz) {
  constexpr int DIM = 3, Q = 19;
  extern __shared__ float pf_smem[];   // [2 stages][g: Q | f: Q][kThreads]
  PlaneCell c;
  c.x = blockIdx.x * blockDim.x + threadIdx.x;
  c.y = blockIdx.y * blockDim.y + threadIdx.y;
  c.live = c.x < g.W && c.y < g.H;
  if (!c.live) return;
  const int z0 = g.zb + blockIdx.z * kz;
  const int z1 = min(z0 + kz, g.zb + g.zn);
  const int tid = threadIdx.y * blockDim.x + threadIdx.x;
  const PopIndex<IDX32> at(g);
  float* slot = pf_smem + tid;
  auto prefetch = [&](int z, int stage) {
    PlaneCell n = c;
    n.z = z;
    n.i = (z * g.H + c.y) * g.W + c.x;
    const Nbr nb(n, g);
    float* s = slot + stage * (2 * Q * kThreads);
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      const auto j = at(q, n.i + nb.at<DIM>(q, 1));
      cp_async4(s + q * kThreads, gi + j);
      cp_async4(s + (Q + q) * kThreads, fi + j);
    }
    cp_async_commit();
  };
  prefetch(z0, 0);
  // flag byte and rho of the next cell ride one iteration ahead too (the flag decides the branch every other
  // load sits behind: 14 % of the stall samples of the first version were the wait for it)
  uint8_t fl_next = A.flags[(z0 * g.H + c.y) * g.W + c.x];
  float rho_next = A.rho_[(z0 * g.H + c.y) * g.W + c.x];
  for (int z = z0; z < z1; ++z) {
    const int stage = (z - z0) & 1;
    const uint8_t fl = fl_next;
    const float rho = rho_next;
    if (z + 1 < z1) {
      fl_next = A.flags[((z + 1) * g.H + c.y) * g.W + c.x];
      rho_next = A.rho_[((z + 1) * g.H + c.y) * g.W + c.x];
    }
    if (z + 1 < z1) {
      prefetch(z + 1, stage ^ 1);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    c.z = z;
    c.i = (z * g.H + c.y) * g.W + c.x;
    const float* s = slot + stage * (2 * Q * kThreads);
    // g into registers once (its moments come first and it is used again in the final loop; reading it from the
    // slots at both places instead measured the same: 12.19 vs 12.23 ms), f straight from the slots
    float gv[Q];
#pragma unroll
    for (int q = 0; q < Q; ++q) gv[q] = s[q * kThreads];
    hcz_cs_cell<DIM, MAGNETIC, IDX32>(c, at, RegPop{gv}, SmemPop{s + Q * kThreads}, fl, rho, A, k, K, g,
                                      GlobalStencil<DIM, MAGNETIC>{});
  }
}

// ---- the same step with the tiles that CAN be brought in by the TMA unit brought in by it -------------------------
// 64 x 4 blocks march along z like k_hcz_collide_stream_pf.  Measured rule (tools/probe/tma_probe.cu on this B200,
// profiles/tma_probe_r02.txt): a tiled tensor-map box may start at any coordinate in the outer dimensions, negative
// ones included (zero fill), but its INNERMOST start coordinate must give a 16-byte aligned global address -- an x
// start of -1, 2 or 7 floats raises "illegal instruction".  So:
//  * the five stencil fields (lap | fai | prho of ws, chi, psi) arrive as rim-extended boxes that start at the aligned
//    x0 - 4 and are 72 wide (the stencil needs x0 - 1 .. x0 + 64), 6 rows, one z plane per step, into a four-slot
//    ring holding planes z-1, z, z+1 (+ the one in flight): the 90 stencil reads per cell are LDS with immediate
//    offsets instead of L1/L2 round trips behind per-load address arithmetic;
//  * the 18 pulled populations with e_x = 0 (f and g) arrive as 64 x 4 boxes at (x0, y0 - e_y, z - e_z, q): aligned,
//    because tiles start at multiples of 64;
//  * the 20 populations with e_x = +-1 are shifted by one float and cannot be a box without widening every tile to
//    72 (which no longer fits two blocks per SM): they keep the 4-byte cp.async copies of the _pf kernel.
// Both kinds land in the same [2Q][4][64] stage (a thread's own slot is [q][ty][tx] either way).  Completion: two
// mbarriers (expect_tx) for the boxes, cp.async.wait_group for the thread's own copies; one __syncthreads per plane
// frees the stage and the ring slot for refill.  (Measured and not kept: replacing that barrier by an arrival counter
// whose last warp issues the refill -- the barrier shows as 11 % of the stall samples in the 512^3 capture, but with
// two blocks per SM the other block fills those slots: 10.46 ms against 10.39 ms.)
// TMA zero-fills outside the tensor while the lattice is periodic: a cell whose box-loaded pull wraps around the
// array (y = 0 / H-1, local z = 0 / D-1; wall cells in the named configurations) re-reads those populations
// directly.  Stencil centres are clamped to the interior exactly as in GlobalStencil: in x / y the clamp is an index
// shift inside the rim-extended tile, planes on the global z shell (block-uniform test) take the global sampler.
// The cell arithmetic is the same hcz_cs_cell, so results are bit-identical to the other two kernels.
constexpr int kTileX = 64, kTileY = 4;
constexpr int kStW = 72, kStH = 6, kStX0 = 4;         // rim-extended stencil tile: starts at x0 - 4, y0 - 1
constexpr int kStField = kStW * kStH;                 // 432 floats = 1728 B
constexpr int kStWs = 1312;                           // [lap | fai | prho] = 3 * 432 floats, padded to 5248 B
constexpr int kStSc = 448;                            // one scalar field, padded to 1792 B (multiples of 128 B)
constexpr int kStSlot = kStWs + 2 * kStSc;            // 2208 floats = 8832 B per z plane
constexpr int kPullStage = 2 * 19 * kThreads;         // 9728 floats = 38912 B
constexpr int kTmaSmemFloats = 2 * kPullStage + 4 * kStSlot;
constexpr int kTmaSmemBytes = kTmaSmemFloats * 4 + 16 /* two mbarriers */ + 128 /* alignment slack */;

struct SmemField {   // sampler around the thread's (clamped) centre: offsets are compile-time after inlining
  const float* c;    // centre element in the slot of plane z
  int dm, dp;        // float offsets from the slot of plane z to the slots of planes z-1 / z+1
  __device__ __forceinline__ float operator()(int dx, int dy, int dz) const {
    const float* p = c + (dz < 0 ? dm : (dz > 0 ? dp : 0));
    return p[dy * kStW + dx];
  }
};
template <bool MAGNETIC>
struct SmemStencil {
  const float* c;
  int dm, dp;
  __device__ __forceinline__ void operator()(const PlaneCell&, const Grid&, const float*, const float*, const float*,
                                             float inv12dx, float* glap, float* gfai, float* gprho, float* gchi,
                                             float* gpsi) const {
    grad_at_fast<3>(SmemField{c, dm, dp}, 0, 0, 0, inv12dx, glap);
    grad_at_fast<3>(SmemField{c + kStField, dm, dp}, 0, 0, 0, inv12dx, gfai);
    grad_at_fast<3>(SmemField{c + 2 * kStField, dm, dp}, 0, 0, 0, inv12dx, gprho);
    if (MAGNETIC) {
      grad_at_fast<3>(SmemField{c + kStWs, dm, dp}, 0, 0, 0, inv12dx, gchi);
      grad_at_fast<3>(SmemField{c + kStWs + kStSc, dm, dp}, 0, 0, 0, inv12dx, gpsi);
    }
  }
};

struct TmaMaps {
  CUtensorMap f, g, ws, chi, psi;
};

template <bool MAGNETIC, bool IDX32>
__global__ void __launch_bounds__(kThreads, 2)
k_hcz_collide_stream_tma(const __grid_constant__ TmaMaps M, const float* __restrict__ fi,
                         const float* __restrict__ gi, CsArgs A, Consts k, KbcScalars K, Grid g, int kz) {
  constexpr int DIM = 3, Q = 19;
  extern __shared__ float tma_raw[];
  // 128-byte aligned start, derived from the shared-space address so that accesses stay LDS/STS
  float* sm = tma_raw + (((128u - (tma::smem_u32(tma_raw) & 127u)) & 127u) >> 2);
  float* pull = sm;
  float* ring = sm + 2 * kPullStage;
  const unsigned bar0 = tma::smem_u32(ring + 4 * kStSlot);
  const int tx = threadIdx.x, ty = threadIdx.y, tid = ty * kTileX + tx;
  const int x0 = blockIdx.x * kTileX, y0 = blockIdx.y * kTileY;
  PlaneCell c;
  c.x = x0 + tx;
  c.y = y0 + ty;
  c.live = c.x < g.W && c.y < g.H;
  const int z0 = g.zb + blockIdx.z * kz;
  const int z1 = min(z0 + kz, g.zb + g.zn);
  if (tid == 0) {
    tma::mbar_init(bar0, 1);
    tma::mbar_init(bar0 + 8, 1);
    tma::fence_barrier_init();
  }
  __syncthreads();
  const PopIndex<IDX32> at(g);
  constexpr unsigned kStBytes = (3 * kStField + (MAGNETIC ? 2 * kStField : 0)) * 4;
  constexpr int kBoxPops = 9;   // D3Q19 populations with e_x = 0
  // one elected thread arms the barrier and issues the boxes (TMA instructions are warp-uniform: issuing them from
  // several lanes only makes the compiler serialise the lanes): the e_x = 0 pull boxes of plane zpull and the
  // stencil planes [zs, zs + ns) of ws / chi / psi
  auto issue = [&](int zpull, int zs, int ns) {
    const unsigned bar = bar0 + 8 * (zpull & 1);
    tma::mbar_expect_tx(bar, 2 * kBoxPops * kThreads * 4 + ns * kStBytes);
    const unsigned dst = tma::smem_u32(pull + (zpull & 1) * kPullStage);
#pragma unroll
    for (int q = 0; q < Q; ++q) {
      if (Lat<3>::ex(q) != 0) continue;
      const int cy = y0 - Lat<3>::ey(q), cz = zpull - Lat<3>::ez(q);
      tma::load_4d(dst + q * kThreads * 4, &M.g, bar, x0, cy, cz, q);
      tma::load_4d(dst + (Q + q) * kThreads * 4, &M.f, bar, x0, cy, cz, q);
    }
    for (int n = 0; n < ns; ++n) {
      const int z = zs + n;
      const unsigned slot = tma::smem_u32(ring + (z & 3) * kStSlot);
      tma::load_4d(slot, &M.ws, bar, x0 - kStX0, y0 - 1, z, 0);
      if (MAGNETIC) {
        tma::load_3d(slot + kStWs * 4, &M.chi, bar, x0 - kStX0, y0 - 1, z);
        tma::load_3d(slot + (kStWs + kStSc) * 4, &M.psi, bar, x0 - kStX0, y0 - 1, z);
      }
    }
  };
  // every thread: 4-byte cp.async copies of its own e_x = +-1 pulls of plane z into its slots of the stage
  auto prefetch = [&](int z) {
    if (c.live) {
      PlaneCell n = c;
      n.z = z;
      n.i = (z * g.H + c.y) * g.W + c.x;
      const Nbr nb(n, g);
      float* s = pull + (z & 1) * kPullStage + tid;
#pragma unroll
      for (int q = 0; q < Q; ++q) {
        if (Lat<3>::ex(q) == 0) continue;
        const auto j = at(q, n.i + nb.at<DIM>(q, 1));
        cp_async4(s + q * kThreads, gi + j);
        cp_async4(s + (Q + q) * kThreads, fi + j);
      }
    }
    cp_async_commit();
  };
  if (tid == 0) issue(z0, z0 - 1, 3);
  prefetch(z0);
  // stencil centre of this thread inside the rim-extended tile (x / y clamp = index shift, see above)
  const int ccx = tx + kStX0 + (interior(c.x, g.W) - c.x), ccy = ty + 1 + (interior(c.y, g.H) - c.y);
  const int cell0 = (z0 * g.H + c.y) * g.W + c.x;
  uint8_t fl_next = c.live ? A.flags[cell0] : (uint8_t)0;
  float rho_next = c.live ? A.rho_[cell0] : 0.f;
  const bool sy0 = c.y == 0, sy1 = c.y == g.H - 1;
  for (int z = z0; z < z1; ++z) {
    __syncthreads();   // every thread is done with the stage and the ring slot about to be refilled
    if (z + 1 < z1) {
      if (tid == 0) issue(z + 1, z + 2, 1);
      prefetch(z + 1);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    const uint8_t fl = fl_next;
    const float rho = rho_next;
    if (c.live && z + 1 < z1) {
      fl_next = A.flags[((z + 1) * g.H + c.y) * g.W + c.x];
      rho_next = A.rho_[((z + 1) * g.H + c.y) * g.W + c.x];
    }
    if (!c.live) continue;
    tma::mbar_wait(bar0 + 8 * (z & 1), ((z - z0) >> 1) & 1);
    c.z = z;
    c.i = (z * g.H + c.y) * g.W + c.x;
    float* s = pull + (z & 1) * kPullStage + tid;
    float gv[Q];
#pragma unroll
    for (int q = 0; q < Q; ++q) gv[q] = s[q * kThreads];
    const bool sz0 = z == 0, sz1 = z == g.D - 1;
    if (sy0 | sy1 | sz0 | sz1) {   // periodic seam: the box was zero-filled where the pull wraps
      const Nbr nb(c, g);
#pragma unroll
      for (int q = 1; q < Q; ++q) {
        const int ex = Lat<3>::ex(q), ey = Lat<3>::ey(q), ez = Lat<3>::ez(q);
        if (ex != 0) continue;
        const bool wraps = (ey == 1 && sy0) || (ey == -1 && sy1) || (ez == 1 && sz0) || (ez == -1 && sz1);
        if (wraps) {
          const auto j = at(q, c.i + nb.at<DIM>(q, 1));
          gv[q] = ld_spec(gi + j);
          s[(Q + q) * kThreads] = ld_spec(fi + j);
        }
      }
      tma::fence_proxy_async();   // generic writes to a stage the TMA unit refills later
    }
    const SmemPop fv{s + Q * kThreads};
    if (interior_z(g, z) == z) {
      const float* slot = ring + (z & 3) * kStSlot;
      const SmemStencil<MAGNETIC> st{slot + ccy * kStW + ccx, (((z - 1) & 3) - (z & 3)) * kStSlot,
                                     (((z + 1) & 3) - (z & 3)) * kStSlot};
      hcz_cs_cell<DIM, MAGNETIC, IDX32>(c, at, RegPop{gv}, fv, fl, rho, A, k, K, g, st);
    } else {
      hcz_cs_cell<DIM, MAGNETIC, IDX32>(c, at, RegPop{gv}, fv, fl, rho, A, k, K, g, GlobalStencil<DIM, MAGNETIC>{});
    }
  }
}

// ---- single-phase: stream + macro + bounce + BGK/KBC (demo_2d_LBM.py:101-121) -------------------
template <int DIM, int MODE, bool IDX32>  // MODE 0: BGK, 1: KBC (D2Q9)
__global__ void __launch_bounds__(kThreads) k_stream_collide(const float* __restrict__ fi, float* __restrict__ fo,
                                                             const uint8_t* __restrict__ flags,
                                                             const float* __restrict__ force, float* __restrict__ rho_,
                                                             float* __restrict__ vel, int kbc, Consts k,
                                                             KbcScalars K, Grid g) {
  using L = Lat<DIM>;
  const PlaneCell c = plane_cell(g);
  if (!c.live) return;
  const int i = c.i;
  const Nbr nb(c, g);
  const PopIndex<IDX32> at(g);
  float fv[L::Q];   // pulls first, flag second (see k_mag_sweep)
#pragma unroll
  for (int q = 0; q < L::Q; ++q) fv[q] = ld_spec(fi + at(q, i + nb.at<DIM>(q, 1)));
  if (flags[i] == LBMFF_OBSTACLE) {  // bounce-back = permutation of the pulled values
#pragma unroll
    for (int q = 0; q < L::Q; ++q) fo[at(q, i)] = fv[L::opp(q)];
    return;  // rho, vel keep their old values on OBSTACLE cells
  }
  float s = 0.f, m[3] = {0.f, 0.f, 0.f};
#pragma unroll
  for (int q = 0; q < L::Q; ++q) {
    s += fv[q];
    if (L::ex(q) != 0) m[0] += fv[q] * (float)L::ex(q);
    if (L::ey(q) != 0) m[1] += fv[q] * (float)L::ey(q);
    if (DIM == 3 && L::ez(q) != 0) m[2] += fv[q] * (float)L::ez(q);
  }
  const float rho = s;
  rho_[i] = rho;
  if constexpr (MODE == 0) {
    // BGK, speed-ordered: one reciprocal of rho, shared equilibrium divisions, 1/tau as a multiplication
    // (29 IEEE divisions fewer per D3Q19 cell: 4.64 -> 3.43 ms per 512^3 step).  Rounding-level differences
    // from the reference order; held to the trajectory bars of tests/test_gpu_fused.py.
    const float inv_rho = 1.0f / rho, scale = k.c * inv_rho;
    float u[DIM];
#pragma unroll
    for (int a = 0; a < DIM; ++a) {
      const float u0 = m[a] * scale;
      vel[a * g.N + i] = u0;
      u[a] = force ? fmaf(k.tau * force[a * g.N + i], inv_rho, u0) : u0;
    }
    const FeqFactors<DIM> F = feq_factors_fast<DIM>(u, k.inv_c);
#pragma unroll
    for (int q = 0; q < L::Q; ++q) fo[at(q, i)] = fmaf(k.keep, fv[q], feq_q<DIM>(F, rho, q) * k.relax);
    return;
  }
  const float scale = k.c / rho;
  float u0[DIM], u[DIM];
#pragma unroll
  for (int a = 0; a < DIM; ++a) {
    u0[a] = m[a] * scale;
    vel[a * g.N + i] = u0[a];
    u[a] = force ? u0[a] + k.tau * force[a * g.N + i] / rho : u0[a];
  }
  if constexpr (MODE == 1) {
    // feq in the reference's operation order (bit for bit: the 0/0 stabiliser pattern is pinned cell by cell), the
    // relaxation itself speed-ordered (kbc_relax_fast: 3 divisions instead of 47).  LBMFF_KBC_REF restores kbc_relax.
    const FeqFactors<DIM> F = k.c == 1.0f ? feq_factors_c1<DIM>(u) : feq_factors<DIM>(u, k.c);
    float fe[9], o[9];
#pragma unroll
    for (int q = 0; q < 9; ++q) fe[q] = feq_q<DIM>(F, rho, q);
#ifdef LBMFF_KBC_REF
    kbc_relax(fv, fe, rho, u0, k.c, kbc, K, o);
#else
    kbc_relax_fast(fv, fe, F, rho, u0, k.c, kbc, K, o);
#endif
#pragma unroll
    for (int q = 0; q < 9; ++q) fo[at(q, i)] = o[q];
  } else {
    const FeqFactors<DIM> F = feq_factors<DIM>(u, k.c);
#pragma unroll
    for (int q = 0; q < L::Q; ++q) fo[at(q, i)] = k.keep * fv[q] + feq_q<DIM>(F, rho, q) / k.tau;
  }
}

static Consts consts_of(const lbmff_fluid_params* p) {
  return make_consts(p->dx, p->dt, p->tau, p->tau_f, p->tau_g, p->density_liquid, p->density_gas, p->rho_liquid,
                     p->rho_gas, p->kappa, p->gravity, p->cot_theta);
}

template <int DIM>
static int moments_impl(const float* f, const uint8_t* flags, float* rho, float* chi, const lbmff_fluid_params* p,
                        const Grid& g, cudaStream_t st, const char* where) {
  if (!(f && flags && rho && p)) return bad_arg(where, "null pointer");
  if (g.N > 0x7fffffffLL) return bad_arg(where, "more than 2^31-1 cells per batch item");
  const PlaneLaunch pl = plane_launch(g);
  const Consts k = consts_of(p);
  const long long Q = Lat<DIM>::Q;
  for (long long b = 0; b < g.B; ++b) {
    if (Q * g.N < 0xffffffffLL)
      k_hcz_moments<DIM, true><<<pl.grid, pl.block, 0, st>>>(f + b * Q * g.N, flags + b * g.N, rho + b * g.N,
                                                             chi ? chi + b * g.N : nullptr, k, g);
    else
      k_hcz_moments<DIM, false><<<pl.grid, pl.block, 0, st>>>(f + b * Q * g.N, flags + b * g.N, rho + b * g.N,
                                                              chi ? chi + b * g.N : nullptr, k, g);
  }
  return check_launch(where);
}

template <int DIM>
static int fields_impl(float* rho, const uint8_t* flags, const float* pressure, const lbmff_fluid_params* p,
                       float* ws, const Grid& g, bool wetting, bool fields, int stage, cudaStream_t st,
                       const char* where) {
  if (!(rho && flags && p && (!fields || (pressure && ws)))) return bad_arg(where, "null pointer");
  if (g.W < 4 || g.H < 4 || (DIM == 3 && g.D < 4)) return bad_arg(where, "every extent must be >= 4");
  const Consts k = consts_of(p);
  if (wetting) launch_wetting<DIM>(rho, flags, k.cot, g, st, stage);
  if (fields) {
    if (g.N > 0x7fffffffLL / 3) return bad_arg(where, "more than (2^31-1)/3 cells per batch item");
    const PlaneLaunch pl = plane_launch(g, 64);
    for (long long b = 0; b < g.B; ++b)
      k_hcz_fields<DIM><<<pl.grid, pl.block, 0, st>>>(rho + b * g.N, flags + b * g.N, pressure + b * g.N,
                                                      ws + b * 3 * g.N, k, g);
  }
  return check_launch(where);
}

template <int DIM>
static int shell_rule_impl(float* field, const uint8_t* flags, const Grid& g, cudaStream_t st, const char* where) {
  if (!(field && flags)) return bad_arg(where, "null pointer");
  launch_shell_rule<DIM>(field, flags, g, st);
  return check_launch(where);
}

// LBMFF_HCZ_PREFETCH=0/1 and LBMFF_HCZ_KZ=<planes per block> override the defaults (tuning, and the test that
// holds the two kernels bit-identical)
#ifndef LBMFF_HCZ_TMA_DEFAULT
#define LBMFF_HCZ_TMA_DEFAULT true    /* green on the box: bit-identical to the cp.async kernel, 11.7 -> 10.4 ms at 512^3 */
#endif
struct CsPrefetch { bool on; int kz; bool tma, tma_required; };
static CsPrefetch cs_prefetch_config() {
  const char* e = getenv("LBMFF_HCZ_PREFETCH");
  const char* z = getenv("LBMFF_HCZ_KZ");
  const char* t = getenv("LBMFF_HCZ_TMA");   // 0: keep the cp.async kernel even where the TMA kernel applies;
                                             // 2: fail instead of falling back (the bit-identity tests)
  CsPrefetch c{e ? atoi(e) != 0 : true, z ? atoi(z) : 16, t ? atoi(t) != 0 : LBMFF_HCZ_TMA_DEFAULT, t ? atoi(t) == 2 : false};
  if (c.kz < 1) c.kz = 1;
  return c;
}

// Tensor maps of one batch item for k_hcz_collide_stream_tma; false when the layout does not meet the TMA rules
// (row pitch W * 4 B must be a multiple of 16, bases 16-byte aligned) or the driver entry point is unavailable.
static bool make_tma_maps(TmaMaps* M, const float* fi, const float* gi, const float* ws, const float* chi,
                          const float* psi, const Grid& g) {
  if (g.W % 4 || g.W < kStW || g.H < kStH || g.D < 4) return false;
  const unsigned long long d4[4] = {(unsigned long long)g.W, (unsigned long long)g.H, (unsigned long long)g.D, 19};
  const unsigned long long w4[4] = {d4[0], d4[1], d4[2], 3};
  const unsigned bpop[4] = {kTileX, kTileY, 1, 1}, bws[4] = {kStW, kStH, 1, 3}, bsc[3] = {kStW, kStH, 1};
  if (!tma::make_map_f32(&M->f, fi, 4, d4, bpop) || !tma::make_map_f32(&M->g, gi, 4, d4, bpop) ||
      !tma::make_map_f32(&M->ws, ws, 4, w4, bws))
    return false;
  if (chi && psi) return tma::make_map_f32(&M->chi, chi, 3, d4, bsc) && tma::make_map_f32(&M->psi, psi, 3, d4, bsc);
  M->chi = M->ws;
  M->psi = M->ws;
  return true;
}

template <int DIM>
static int collide_stream_impl(const float* f_in, const float* g_in, float* f_out, float* g_out, const float* rho,
                               const uint8_t* flags, const uint8_t* mflags, const float* ws, float* chi,
                               float* psi, const float* Hext, const double* Hu, const lbmff_fluid_params* p,
                               int kbc_on_g, float* vel, float* pressure, float* force_out, const Grid& g,
                               bool fold_shell_rule, cudaStream_t st, const char* where) {
  if (!(f_in && g_in && f_out && g_out && rho && flags && ws && p && vel && pressure))
    return bad_arg(where, "null pointer");
  if (f_in == f_out || g_in == g_out) return bad_arg(where, "in and out population buffers must differ");
  const bool magnetic = chi != nullptr;
  if (magnetic && !(psi && mflags && (Hext || Hu))) return bad_arg(where, "magnetic inputs incomplete");
  const Consts k = consts_of(p);
  HextUniform hu{{0.f, 0.f, 0.f}};
  if (Hu) for (int a = 0; a < DIM; ++a) hu.v[a] = (float)Hu[a];
  if (g.N > 0x7fffffffLL / 3) return bad_arg(where, "more than (2^31-1)/3 cells per batch item");
  if (magnetic && fold_shell_rule) {
    launch_shell_rule<DIM>(psi, mflags, g, st);
    launch_shell_rule<DIM>(chi, flags, g, st);
  }
  const CsPrefetch pf = cs_prefetch_config();
  const PlaneLaunch pl = plane_launch(g, 64)  /* 64 x 4 blocks share y neighbours in L1 (one-cell kernel: 14.3 ms vs 15.2 ms for 256 x 1 at 512^3; the marching kernel is insensitive: 12.2-12.5 ms for 32 ... 256) */;
  const long long Q = Lat<DIM>::Q, N = g.N;
  const KbcScalars K = make_kbc(p->tau_g);
  const bool ix = Q * N < 0xffffffffLL;
  for (long long b = 0; b < g.B; ++b) {
    const CsArgs A{f_out + b * Q * N, g_out + b * Q * N, rho + b * N, flags + b * N, ws + b * 3 * N,
                   magnetic ? chi + b * N : nullptr, magnetic ? psi + b * N : nullptr,
                   Hext ? Hext + b * DIM * N : nullptr, hu, vel + b * DIM * N, pressure + b * N,
                   force_out ? force_out + b * DIM * N : nullptr, kbc_on_g};
    const float* fi = f_in + b * Q * N;
    const float* gi = g_in + b * Q * N;
    if constexpr (DIM == 3) {
      TmaMaps maps;
      if (pf.on && pf.tma && g.zn >= 2 && make_tma_maps(&maps, fi, gi, A.ws, A.chi, A.psi, g)) {
        // z-marching 64 x 4 blocks, every tile staged by the TMA unit
        const int kz = g.zn < pf.kz ? g.zn : pf.kz;
        const dim3 grid((unsigned)((g.W + kTileX - 1) / kTileX), (unsigned)((g.H + kTileY - 1) / kTileY),
                        (unsigned)((g.zn + kz - 1) / kz));
        const dim3 block(kTileX, kTileY, 1);
#define LBMFF_CSTMA(MAG, IX)                                                                                        \
  do {                                                                                                              \
    static bool done[64] = {};                                                                                      \
    int dev = 0;                                                                                                    \
    cudaGetDevice(&dev);                                                                                            \
    if (dev < 0 || dev >= 64 || !done[dev]) {                                                                       \
      cudaFuncSetAttribute(k_hcz_collide_stream_tma<MAG, IX>, cudaFuncAttributeMaxDynamicSharedMemorySize,          \
                           kTmaSmemBytes);                                                                          \
      if (dev >= 0 && dev < 64) done[dev] = true;                                                                   \
    }                                                                                                               \
    k_hcz_collide_stream_tma<MAG, IX><<<grid, block, kTmaSmemBytes, st>>>(maps, fi, gi, A, k, K, g, kz);            \
  } while (0)
        if (magnetic) { if (ix) LBMFF_CSTMA(true, true); else LBMFF_CSTMA(true, false); }
        else { if (ix) LBMFF_CSTMA(false, true); else LBMFF_CSTMA(false, false); }
#undef LBMFF_CSTMA
        continue;
      }
      if (pf.tma_required && g.zn >= 2)
        return bad_arg(where, "LBMFF_HCZ_TMA=2 but the TMA kernel does not apply (needs W % 4 == 0, W >= 72, H >= 6)");
      if (pf.on && g.zn >= 2) {   // z-marching blocks with the next cell's pulls in flight (cp.async)
        const int kz = g.zn < pf.kz ? g.zn : pf.kz;
        const dim3 grid(pl.grid.x, pl.grid.y, (unsigned)((g.zn + kz - 1) / kz));
        constexpr int smem = 2 * 2 * 19 * kThreads * (int)sizeof(float);
#define LBMFF_CSPF(MAG, IX)                                                                                         \
  do {                                                                                                              \
    static bool done[64] = {};   /* the attribute is per device; set once per device, outside any graph capture */ \
    int dev = 0;                                                                                                    \
    cudaGetDevice(&dev);                                                                                            \
    if (dev < 0 || dev >= 64 || !done[dev]) {                                                                       \
      cudaFuncSetAttribute(k_hcz_collide_stream_pf<MAG, IX>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);    \
      if (dev >= 0 && dev < 64) done[dev] = true;                                                                   \
    }                                                                                                               \
    k_hcz_collide_stream_pf<MAG, IX><<<grid, pl.block, smem, st>>>(fi, gi, A, k, K, g, kz);                         \
  } while (0)
        if (magnetic) { if (ix) LBMFF_CSPF(true, true); else LBMFF_CSPF(true, false); }
        else { if (ix) LBMFF_CSPF(false, true); else LBMFF_CSPF(false, false); }
#undef LBMFF_CSPF
        continue;
      }
    }
#define LBMFF_CS(MAG, IX) k_hcz_collide_stream<DIM, MAG, IX><<<pl.grid, pl.block, 0, st>>>(fi, gi, A, k, K, g)
    if (magnetic) { if (ix) LBMFF_CS(true, true); else LBMFF_CS(true, false); }
    else { if (ix) LBMFF_CS(false, true); else LBMFF_CS(false, false); }
#undef LBMFF_CS
  }
  return check_launch(where);
}

template <int DIM>
static int stream_collide_impl(const float* f_in, float* f_out, const uint8_t* flags, const float* force, float* rho,
                               float* vel, double tau, int kbc, double dx, double dt, int B, int D, int H, int W,
                               cudaStream_t st, const char* where) {
  if (!(f_in && f_out && f_in != f_out && flags && rho && vel)) return bad_arg(where, "null or aliased pointer");
  if (!(tau > 0)) return bad_arg(where, "tau must be positive");
  if (DIM == 3 && kbc >= 0) return bad_arg(where, "KBC collision exists for D2Q9 only");
  const Grid g = make_grid(B, D, H, W);
  if (g.N > 0x7fffffffLL / 3) return bad_arg(where, "more than (2^31-1)/3 cells per batch item");
  const Consts k = make_consts(dx, dt, tau);
  const PlaneLaunch pl = plane_launch(g);
  const long long Q = Lat<DIM>::Q, N = g.N;
  const KbcScalars K = make_kbc(tau);
  const bool ix = Q * N < 0xffffffffLL;
  for (long long b = 0; b < B; ++b) {
#define LBMFF_SC(MODE, IX)                                                                                        \
  k_stream_collide<DIM, MODE, IX><<<pl.grid, pl.block, 0, st>>>(f_in + b * Q * N, f_out + b * Q * N, flags + b * N, \
                                                                force ? force + b * DIM * N : nullptr, rho + b * N, \
                                                                vel + b * DIM * N, kbc, k, K, g)
    if constexpr (DIM == 2) {
      if (kbc >= 0) {
        if (ix) LBMFF_SC(1, true); else LBMFF_SC(1, false);
        continue;
      }
    }
    if (ix) LBMFF_SC(0, true); else LBMFF_SC(0, false);
#undef LBMFF_SC
  }
  return check_launch(where);
}

}  // namespace lbmff

using namespace lbmff;

extern "C" {

int lbmff_hcz_moments_d2q9(const float* f, const uint8_t* flags, float* rho, float* chi_out,
                           const lbmff_fluid_params* p, int B, int H, int W, lbmff_stream_t stream) {
  return moments_impl<2>(f, flags, rho, chi_out, p, make_grid(B, 1, H, W), (cudaStream_t)stream, __func__);
}
int lbmff_hcz_moments_d3q19(const float* f, const uint8_t* flags, float* rho, float* chi_out,
                            const lbmff_fluid_params* p, int B, int D, int H, int W, lbmff_stream_t stream) {
  return moments_impl<3>(f, flags, rho, chi_out, p, make_grid(B, D, H, W), (cudaStream_t)stream, __func__);
}
int lbmff_hcz_capillary_fields_2d(float* rho, const uint8_t* flags, const float* pressure,
                                  const lbmff_fluid_params* p, float* ws, int B, int H, int W,
                                  lbmff_stream_t stream) {
  return fields_impl<2>(rho, flags, pressure, p, ws, make_grid(B, 1, H, W), true, true, 0, (cudaStream_t)stream, __func__);
}
int lbmff_hcz_capillary_fields_3d(float* rho, const uint8_t* flags, const float* pressure,
                                  const lbmff_fluid_params* p, float* ws, int B, int D, int H, int W,
                                  lbmff_stream_t stream) {
  return fields_impl<3>(rho, flags, pressure, p, ws, make_grid(B, D, H, W), true, true, 0, (cudaStream_t)stream, __func__);
}
int lbmff_hcz_collide_stream_d2q9(const float* f_in, const float* g_in, float* f_out, float* g_out,
                                  const float* rho, const uint8_t* flags, const uint8_t* mflags, const float* ws,
                                  float* chi, float* psi, const float* H_ext, const double* H_uniform,
                                  const lbmff_fluid_params* p, int kbc_on_g, float* vel, float* pressure,
                                  float* force_out, int B, int H, int W, lbmff_stream_t stream) {
  return collide_stream_impl<2>(f_in, g_in, f_out, g_out, rho, flags, mflags, ws, chi, psi, H_ext, H_uniform, p,
                                kbc_on_g, vel, pressure, force_out, make_grid(B, 1, H, W), true,
                                (cudaStream_t)stream, __func__);
}
int lbmff_hcz_collide_stream_d3q19(const float* f_in, const float* g_in, float* f_out, float* g_out,
                                   const float* rho, const uint8_t* flags, const uint8_t* mflags, const float* ws,
                                   float* chi, float* psi, const float* H_ext, const double* H_uniform,
                                   const lbmff_fluid_params* p, float* vel, float* pressure, float* force_out,
                                   int B, int D, int H, int W, lbmff_stream_t stream) {
  return collide_stream_impl<3>(f_in, g_in, f_out, g_out, rho, flags, mflags, ws, chi, psi, H_ext, H_uniform, p, 0,
                                vel, pressure, force_out, make_grid(B, D, H, W), true, (cudaStream_t)stream,
                                __func__);
}
/* ---- slab tier (z-slab decomposition; D = local planes, halos included) ------------------------ */
static int slab_grid(const lbmff_slab* s, int B, int D, int H, int W, Grid* g, const char* where) {
  if (!s) return bad_arg(where, "null slab descriptor");
  if (s->Dg < D || s->zoff <= -s->Dg || s->zoff >= s->Dg) return bad_arg(where, "slab: need D <= Dg and |zoff| < Dg");
  if (s->zb < 0 || s->zn < 1 || s->zb + s->zn > D) return bad_arg(where, "slab: plane range outside the local array");
  *g = make_slab_grid(B, D, H, W, s->Dg, s->zoff, s->zb, s->zn);
  return 0;
}
#define LBMFF_SLAB_GRID                                              \
  Grid g;                                                            \
  if (int rc = slab_grid(slab, B, D, H, W, &g, __func__)) return rc

int lbmff_slab_hcz_moments_d3q19(const float* f, const uint8_t* flags, float* rho, float* chi_out,
                                 const lbmff_fluid_params* p, const lbmff_slab* slab, int B, int D, int H, int W,
                                 lbmff_stream_t stream) {
  LBMFF_SLAB_GRID;
  return moments_impl<3>(f, flags, rho, chi_out, p, g, (cudaStream_t)stream, __func__);
}
int lbmff_slab_hcz_wetting_3d(float* rho, const uint8_t* flags, const lbmff_fluid_params* p, int stage,
                              const lbmff_slab* slab, int B, int D, int H, int W, lbmff_stream_t stream) {
  LBMFF_SLAB_GRID;
  if (stage < 0 || stage > 2) return bad_arg(__func__, "stage must be 0, 1 or 2");
  return fields_impl<3>(rho, flags, nullptr, p, nullptr, g, true, false, stage, (cudaStream_t)stream, __func__);
}
int lbmff_slab_hcz_fields_3d(const float* rho, const uint8_t* flags, const float* pressure,
                             const lbmff_fluid_params* p, float* ws, const lbmff_slab* slab, int B, int D, int H,
                             int W, lbmff_stream_t stream) {
  LBMFF_SLAB_GRID;
  return fields_impl<3>(const_cast<float*>(rho), flags, pressure, p, ws, g, false, true, 0, (cudaStream_t)stream,
                        __func__);
}
int lbmff_slab_shell_rule_3d(float* field, const uint8_t* flags, const lbmff_slab* slab, int B, int D, int H, int W,
                             lbmff_stream_t stream) {
  LBMFF_SLAB_GRID;
  return shell_rule_impl<3>(field, flags, g, (cudaStream_t)stream, __func__);
}
int lbmff_slab_hcz_collide_stream_d3q19(const float* f_in, const float* g_in, float* f_out, float* g_out,
                                        const float* rho, const uint8_t* flags, const uint8_t* mflags,
                                        const float* ws, const float* chi, const float* psi,
                                        const double* H_uniform, const lbmff_fluid_params* p, float* vel,
                                        float* pressure, float* force_out, const lbmff_slab* slab, int B, int D,
                                        int H, int W, lbmff_stream_t stream) {
  LBMFF_SLAB_GRID;
  return collide_stream_impl<3>(f_in, g_in, f_out, g_out, rho, flags, mflags, ws, const_cast<float*>(chi),
                                const_cast<float*>(psi), nullptr, H_uniform, p, 0, vel, pressure, force_out, g, false,
                                (cudaStream_t)stream, __func__);
}

int lbmff_stream_collide_d2q9(const float* f_in, float* f_out, const uint8_t* flags, const float* force, float* rho,
                              float* vel, double tau, int kbc, double dx, double dt, int B, int H, int W,
                              lbmff_stream_t stream) {
  return stream_collide_impl<2>(f_in, f_out, flags, force, rho, vel, tau, kbc, dx, dt, B, 1, H, W,
                                (cudaStream_t)stream, __func__);
}
int lbmff_stream_collide_d3q19(const float* f_in, float* f_out, const uint8_t* flags, const float* force,
                               float* rho, float* vel, double tau, double dx, double dt, int B, int D, int H, int W,
                               lbmff_stream_t stream) {
  return stream_collide_impl<3>(f_in, f_out, flags, force, rho, vel, tau, -1, dx, dt, B, D, H, W,
                                (cudaStream_t)stream, __func__);
}

}  // extern "C"
