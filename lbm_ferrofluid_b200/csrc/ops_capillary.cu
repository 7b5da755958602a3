// Drop-in tier, part 3: LBMCollisionHCZ{2,3}d.capillary_process (a12).
// Sequence on one stream: ordered wetting kernels (in place on the rho shell) ->
// k_cap_fields (density, lap, fai, prho, chi) -> k_cap_forces (force, dfai, dprho, vel, pressure).
#include "capi.cuh"
#include "capillary.cuh"

namespace lbmff {

struct AffineRhoField {  // density = affine(rho) sampled on the fly
  const float* __restrict__ rho;
  Affine A;
  Grid g;
  __device__ __forceinline__ float operator()(int x, int y, int z) const {
    return affine_density(__ldg(rho + lin(g, x, y, z)), A);
  }
};

// ws layout: [lap | fai | prho | chi], each N floats per batch item
template <int DIM>
__global__ void __launch_bounds__(kThreads) k_cap_fields(const float* __restrict__ rho,
                                                         const float* __restrict__ density_old,
                                                         const float* __restrict__ pressure_old,
                                                         const float* __restrict__ phi,
                                                         float* __restrict__ density_out, float* __restrict__ ws,
                                                         Consts k, Grid g) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.N) return;
  const long long b = blockIdx.y;
  const Cell c = cell_of(i, g);
  const float* r = rho + b * g.N;
  float* w = ws + b * 4 * g.N;
  const bool shell = c.x == 0 || c.x == g.W - 1 || c.y == 0 || c.y == g.H - 1 ||
                     (DIM == 3 && (c.z == 0 || c.z == g.D - 1));
  // density from the (post-wetting) rho, HCZ_3d.py:213-215
  density_out[b * g.N + i] = affine_density(r[i], k.aff);
  // Laplacian of density, zero on the shell, HCZ_3d.py:220
  AffineRhoField Dn{r, k.aff, g};
  w[i] = shell ? 0.0f : laplacian_at<DIM>(Dn, c.x, c.y, c.z, k.dx2);
  // fai / prho: interior values replicate-padded, HCZ_3d.py:73-82
  const long long j = lin(g, interior(c.x, g.W), interior(c.y, g.H), DIM == 3 ? interior(c.z, g.D) : 0);
  const float rj = r[j];
  w[g.N + i] = eos_cs(rj, k.cs2, k.a_eos, 1) - rj * k.cs2;
  w[2 * g.N + i] = pressure_old[b * g.N + j] - k.cs2 * density_old[b * g.N + j];
  if (phi) w[3 * g.N + i] = susceptibility(phi[b * g.N + i], k.sm);  // HCZ_3d.py:233-234
}

template <int DIM>
__global__ void __launch_bounds__(kThreads) k_cap_forces(const float* __restrict__ gq,
                                                         const float* __restrict__ vel_old,
                                                         const float* __restrict__ pressure_old,
                                                         const uint8_t* __restrict__ flags,
                                                         const float* __restrict__ density,
                                                         const float* __restrict__ H2, const float* __restrict__ ws,
                                                         float* __restrict__ vel_out, float* __restrict__ pressure_out,
                                                         float* __restrict__ force_out, float* __restrict__ dfai_out,
                                                         float* __restrict__ dprho_out, Consts k, Grid g) {
  using L = Lat<DIM>;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.N) return;
  const long long b = blockIdx.y;
  const Cell c = cell_of(i, g);
  const uint8_t* fl = flags + b * g.N;
  const float* w = ws + b * 4 * g.N;
  const int x = interior(c.x, g.W), y = interior(c.y, g.H), z = DIM == 3 ? interior(c.z, g.D) : 0;
  float glap[3], gchi[3] = {0.f, 0.f, 0.f}, gfai[3], gprho[3];
  grad_at<DIM>(ObstacleClampedField{w, fl, g}, x, y, z, k.dx, glap);
  grad_at<DIM>(ObstacleClampedField{w + g.N, fl, g}, x, y, z, k.dx, gfai);
  grad_at<DIM>(ObstacleClampedField{w + 2 * g.N, fl, g}, x, y, z, k.dx, gprho);
  if (H2) grad_at<DIM>(ObstacleClampedField{w + 3 * g.N, fl, g}, x, y, z, k.dx, gchi);
  const float dn = density[b * g.N + i];
  float F[DIM];
#pragma unroll
  for (int a = 0; a < DIM; ++a) {
    float v = k.kappa * dn * glap[a];          // HCZ_3d.py:225-229
    if (a == 1) v += k.grav_y * dn;            // :230  (gravity acts along y)
    if (H2) v += k.kelvin * H2[b * g.N + i] * gchi[a];  // :231-240
    F[a] = v;
    force_out[(b * DIM + a) * g.N + i] = v;
    dfai_out[(b * DIM + a) * g.N + i] = gfai[a];
    dprho_out[(b * DIM + a) * g.N + i] = gprho[a];
  }
  if (fl[i] == LBMFF_FLUID) {
    float s = 0.f, m[3] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int q = 0; q < L::Q; ++q) {
      const float v = __ldg(gq + (b * L::Q + q) * g.N + i);
      s += v;
      if (L::ex(q) != 0) m[0] += v * (float)L::ex(q);
      if (L::ey(q) != 0) m[1] += v * (float)L::ey(q);
      if (DIM == 3 && L::ez(q) != 0) m[2] += v * (float)L::ez(q);
    }
    float dot = 0.f;
#pragma unroll
    for (int a = 0; a < DIM; ++a) {
      const float u = (m[a] * k.c + k.half_dt_RT * F[a]) / k.cs2 / dn;  // :247-251
      vel_out[(b * DIM + a) * g.N + i] = u;
      dot += u * gprho[a];
    }
    pressure_out[b * g.N + i] = s - k.half_dt * dot;  // :256-260
  } else {
#pragma unroll
    for (int a = 0; a < DIM; ++a) vel_out[(b * DIM + a) * g.N + i] = vel_old[(b * DIM + a) * g.N + i];
    pressure_out[b * g.N + i] = pressure_old[b * g.N + i];
  }
}

template <int DIM>
static int capillary_impl(float* rho, const float* vel, const uint8_t* flags, const float* gq, const float* density,
                          const float* pressure, const float* H2, const float* phi, const lbmff_fluid_params* p,
                          float* vel_out, float* density_out, float* pressure_out, float* force_out,
                          float* dfai_out, float* dprho_out, float* ws, int B, int D, int H, int W,
                          cudaStream_t st, const char* where) {
  if (!(rho && vel && flags && gq && density && pressure && p && vel_out && density_out && pressure_out &&
        force_out && dfai_out && dprho_out && ws))
    return bad_arg(where, "null pointer");
  if ((H2 == nullptr) != (phi == nullptr)) return bad_arg(where, "H2 and phi must be given together");
  if (W < 4 || H < 4 || (DIM == 3 && D < 4)) return bad_arg(where, "every extent must be >= 4");
  const Grid g = make_grid(B, D, H, W);
  const Consts k = make_consts(p->dx, p->dt, p->tau, p->tau_f, p->tau_g, p->density_liquid, p->density_gas,
                               p->rho_liquid, p->rho_gas, p->kappa, p->gravity, p->cot_theta);
  launch_wetting<DIM>(rho, flags, k.cot, g, st);
  const dim3 grid(blocks_for(g.N), (unsigned)B);
  k_cap_fields<DIM><<<grid, kThreads, 0, st>>>(rho, density, pressure, phi, density_out, ws, k, g);
  k_cap_forces<DIM><<<grid, kThreads, 0, st>>>(gq, vel, pressure, flags, density_out, H2, ws, vel_out, pressure_out,
                                               force_out, dfai_out, dprho_out, k, g);
  return check_launch(where);
}

}  // namespace lbmff

using namespace lbmff;

extern "C" {

int lbmff_capillary_process_d2q9(float* rho, const float* vel, const uint8_t* flags, const float* g,
                                 const float* density, const float* pressure, const float* H2, const float* phi,
                                 const lbmff_fluid_params* p, float* vel_out, float* density_out,
                                 float* pressure_out, float* force_out, float* dfai_out, float* dprho_out,
                                 float* workspace, int B, int H, int W, lbmff_stream_t stream) {
  return capillary_impl<2>(rho, vel, flags, g, density, pressure, H2, phi, p, vel_out, density_out, pressure_out,
                           force_out, dfai_out, dprho_out, workspace, B, 1, H, W, (cudaStream_t)stream, __func__);
}
int lbmff_capillary_process_d3q19(float* rho, const float* vel, const uint8_t* flags, const float* g,
                                  const float* density, const float* pressure, const float* H2, const float* phi,
                                  const lbmff_fluid_params* p, float* vel_out, float* density_out,
                                  float* pressure_out, float* force_out, float* dfai_out, float* dprho_out,
                                  float* workspace, int B, int D, int H, int W, lbmff_stream_t stream) {
  return capillary_impl<3>(rho, vel, flags, g, density, pressure, H2, phi, p, vel_out, density_out, pressure_out,
                           force_out, dfai_out, dprho_out, workspace, B, D, H, W, (cudaStream_t)stream, __func__);
}

}  // extern "C"
