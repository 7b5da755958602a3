/* lbmff.h -- C ABI of the B200-native LBM-Ferrofluid solver path (liblbmff.so).
 *
 * Plain pointers and sizes only: no torch / ATen types cross this boundary.
 * Every pointer is a DEVICE pointer (fp32 unless stated; flags are uint8 with
 * FLUID = 1, OBSTACLE = 2, reference src/LBM/utils/types.py:7-15).  Tensors use
 * the reference's contiguous layout [B][C][D][H][W], x (W) fastest; 2-D grids
 * pass D = 1 and use the *_d2q9 entry points.  `stream` is a cudaStream_t.
 * Every function returns 0 on success or a cudaError_t value;
 * lbmff_last_error_string() describes the last failure of the calling thread.
 * Outputs never alias inputs unless a parameter is documented "in place".
 *
 * Each entry point names the reference interface (file:line under
 * /root/reference) that a binding would replace with it.  The reference is
 * pure Python/PyTorch, so "the FFI a maintainer would add" is a ctypes call
 * from the body of the named method -- see INTEGRATION.md.
 */
#ifndef LBMFF_H
#define LBMFF_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* lbmff_stream_t; /* cudaStream_t */

/* rho -> density affine map parameters and the lattice scales, shared by many
 * calls (reference: constructor kwargs of every op, LBM_collision_3d.py:12-28). */
/* Scalars are doubles, like the Python floats of the reference: derived constants
 * (c = dx/dt, cs2, 1 - 0.5/tau, ...) are formed in double on the host and rounded
 * to fp32 once, exactly where the reference's scalar-times-tensor ops round them. */
typedef struct lbmff_fluid_params {
  double dx, dt;
  double density_liquid, density_gas, rho_liquid, rho_gas;
  double kappa;
  double tau;          /* single-phase / magnetic relaxation time (self._tau) */
  double tau_f, tau_g; /* HCZ relaxation times */
  double gravity;      /* scalar given to set_gravity(); acts along -y */
  double cot_theta;    /* tan(pi/2 - contact_angle) as the reference evaluates it in fp32, HCZ_3d.py:97 */
} lbmff_fluid_params;

const char* lbmff_last_error_string(void);
int lbmff_version(void);

/* ---- a1/a2  LBMPropagation{2,3}d.propagation / rebounce_obstacle
 *      (LBM_propagation_3d.py:18-111,113-142; LBM_propagation_2d.py:18-68,70-86) */
int lbmff_propagation_d2q9(const float* f, float* f_out, int B, int H, int W, lbmff_stream_t stream);
int lbmff_propagation_d3q19(const float* f, float* f_out, int B, int D, int H, int W, lbmff_stream_t stream);
int lbmff_rebounce_obstacle_d2q9(const float* f, const uint8_t* flags, float* f_out, int B, int H, int W,
                                 lbmff_stream_t stream);
int lbmff_rebounce_obstacle_d3q19(const float* f, const uint8_t* flags, float* f_out, int B, int D, int H, int W,
                                  lbmff_stream_t stream);

/* ---- a3  LBMMacroCompute{2,3}d.macro_compute (LBM_macro_compute_3d.py:73-120, _2d.py:65-107)
 *      density_out may be NULL (single-phase form). */
int lbmff_macro_compute_d2q9(const float* f, const float* rho, const float* vel, const uint8_t* flags,
                             float* rho_out, float* vel_out, float* density_out, const lbmff_fluid_params* p, int B,
                             int H, int W, lbmff_stream_t stream);
int lbmff_macro_compute_d3q19(const float* f, const float* rho, const float* vel, const uint8_t* flags,
                              float* rho_out, float* vel_out, float* density_out, const lbmff_fluid_params* p, int B,
                              int D, int H, int W, lbmff_stream_t stream);

/* ---- a4  get_pressure (order 0: LBM_macro_compute_3d.py:59-71) /
 *          equation_of_states (order 1: LBM_collision_3d.py:105-122); n = element count */
int lbmff_equation_of_states(const float* rho, float* out, long long n, double dx, double dt, int order,
                             lbmff_stream_t stream);

/* ---- a5/a6/a14  get_feq_ / get_geq_ / compute_Gamma
 *      (LBM_collision_3d.py:133-207, LBM_collision_2d.py:113-181, HCZ_3d.py:271-281)
 *      force may be NULL; feq_in may be NULL (then geq builds feq itself from rho, vel, force, tau). */
int lbmff_get_feq_d2q9(const float* rho, const float* vel, const float* force, double tau, double dx, double dt,
                       float* feq, int B, int H, int W, lbmff_stream_t stream);
int lbmff_get_feq_d3q19(const float* rho, const float* vel, const float* force, double tau, double dx, double dt,
                        float* feq, int B, int D, int H, int W, lbmff_stream_t stream);
int lbmff_get_geq_d2q9(const float* rho, const float* vel, const float* density, const float* pressure,
                       const float* force, const float* feq_in, double tau, double dx, double dt, float* geq, int B,
                       int H, int W, lbmff_stream_t stream);
int lbmff_get_geq_d3q19(const float* rho, const float* vel, const float* density, const float* pressure,
                        const float* force, const float* feq_in, double tau, double dx, double dt, float* geq, int B,
                        int D, int H, int W, lbmff_stream_t stream);
int lbmff_compute_gamma_d2q9(const float* vel, double dx, double dt, float* gamma, int B, int H, int W,
                             lbmff_stream_t stream);
int lbmff_compute_gamma_d3q19(const float* vel, double dx, double dt, float* gamma, int B, int D, int H, int W,
                              lbmff_stream_t stream);

/* ---- a13  smooth_phi (HCZ_3d.py:265-269, LBM_magnetic_3d.py:119-123) */
int lbmff_smooth_phi(const float* phi, float* out, long long n, double eps, lbmff_stream_t stream);

/* ---- a19  the demos' order-parameter line phi = -(2 (density - dg) / (dl - dg) - 1) on density = affine(rho)
 *      (demo_3d_LBM_Rosensweig_instability.py:162, HCZ_3d.py:213-215): nine elementwise torch kernels in one pass,
 *      bit-identical to the torch expression evaluated on CUDA. */
int lbmff_order_parameter(const float* rho, double rho_gas, double rho_liquid, double density_gas,
                          double density_liquid, float* phi_out, long long n, lbmff_stream_t stream);

/* ---- a7/a8  get_grad (static) / get_laplacian
 *      (LBM_collision_3d.py:209-318, LBM_collision_2d.py:183-256); out of grad is [B][dim][...] */
int lbmff_get_grad_2d(const float* in, const uint8_t* flags, float* out, double dx, int B, int H, int W,
                      lbmff_stream_t stream);
int lbmff_get_grad_3d(const float* in, const uint8_t* flags, float* out, double dx, int B, int D, int H, int W,
                      lbmff_stream_t stream);
int lbmff_get_laplacian_2d(const float* in, float* out, double dx, int B, int H, int W, lbmff_stream_t stream);
int lbmff_get_laplacian_3d(const float* in, float* out, double dx, int B, int D, int H, int W,
                           lbmff_stream_t stream);

/* ---- a17  get_staggered_x/_y/_z (src/LBM/utils/grid.py:7-49); axis 0 = x, 1 = y, 2 = z.
 *      in: [BC][D][H][W] -> out with extent + 1 along `axis`. */
int lbmff_get_staggered(const float* in, float* out, int axis, int BC, int D, int H, int W, lbmff_stream_t stream);

/* ---- get_vort (LBM_macro_compute_2d.py:109-117) */
int lbmff_get_vort_2d(const float* vel, float* out, double dx, int B, int H, int W, lbmff_stream_t stream);

/* ---- a9  LBMCollision{2,3}d.collision: BGK with velocity-shift forcing
 *      (LBM_collision_3d.py:320-346, LBM_collision_2d.py:258-287) */
int lbmff_collision_bgk_d2q9(const float* f, const float* rho, const float* vel, const uint8_t* flags,
                             const float* force, double tau, double dx, double dt, float* f_out, int B, int H, int W,
                             lbmff_stream_t stream);
int lbmff_collision_bgk_d3q19(const float* f, const float* rho, const float* vel, const uint8_t* flags,
                              const float* force, double tau, double dx, double dt, float* f_out, int B, int D, int H,
                              int W, lbmff_stream_t stream);

/* ---- a10  LBMCollisionMRT2d.get_s_by_KBC / collision (LBM_collision_MRT_2d.py:62-182,184-226)
 *      kbc != 0: centred moments, s = {T, N, Pi_xy} (every KBC_* type, types.py:72-107);
 *      kbc == 0: raw moments, s = {T, Pi_xy, Q_xxy, Q_xyy, A} (KBC_type None / LBGK). */
int lbmff_get_s_by_kbc_d2q9(const float* f, const float* rho, const float* vel, int kbc, double dx, double dt,
                            float* s_out, int B, int H, int W, lbmff_stream_t stream);
int lbmff_collision_kbc_d2q9(const float* f, const float* rho, const float* vel, const uint8_t* flags,
                             const float* force, double tau, int kbc, double dx, double dt, float* f_out, int B, int H,
                             int W, lbmff_stream_t stream);

/* ---- a11  LBMCollisionSC2d.calculate_force / collision (LBM_collision_SC_2d.py:22-97,99-137) */
int lbmff_sc_force_d2q9(const float* density, const uint8_t* flags, double density_liquid, double density_gas,
                        double dx, double dt, float* force_out, int B, int H, int W, lbmff_stream_t stream);

/* ---- a12  LBMCollisionHCZ{2,3}d.capillary_process (HCZ_3d.py:21-263, HCZ_2d.py:21-173)
 *      rho: IN PLACE (wetting boundary on the shell).  H2 / phi may both be NULL (no Kelvin force).
 *      workspace: >= 4 * B*D*H*W floats of scratch.  Outputs: vel_out [dim], density_out, pressure_out,
 *      force_out [dim], dfai_out [dim], dprho_out [dim]. */
int lbmff_capillary_process_d2q9(float* rho, const float* vel, const uint8_t* flags, const float* g,
                                 const float* density, const float* pressure, const float* H2, const float* phi,
                                 const lbmff_fluid_params* p, float* vel_out, float* density_out,
                                 float* pressure_out, float* force_out, float* dfai_out, float* dprho_out,
                                 float* workspace, int B, int H, int W, lbmff_stream_t stream);
int lbmff_capillary_process_d3q19(float* rho, const float* vel, const uint8_t* flags, const float* g,
                                  const float* density, const float* pressure, const float* H2, const float* phi,
                                  const lbmff_fluid_params* p, float* vel_out, float* density_out,
                                  float* pressure_out, float* force_out, float* dfai_out, float* dprho_out,
                                  float* workspace, int B, int D, int H, int W, lbmff_stream_t stream);

/* ---- a15  LBMCollisionHCZ{2,3}d.collision (HCZ_3d.py:283-361, HCZ_2d.py:193-286)
 *      kbc_on_g: 2-D only, entropic KBC relaxation of g (KBC_type is not None). */
int lbmff_collision_hcz_d2q9(const float* f, const float* g, const float* rho, const float* vel,
                             const float* density, const uint8_t* flags, const float* force, const float* pressure,
                             const float* dfai, const float* dprho, const lbmff_fluid_params* p, int kbc_on_g,
                             float* f_out, float* g_out, int B, int H, int W, lbmff_stream_t stream);
int lbmff_collision_hcz_d3q19(const float* f, const float* g, const float* rho, const float* vel,
                              const float* density, const uint8_t* flags, const float* force,
                              const float* pressure, const float* dfai, const float* dprho,
                              const lbmff_fluid_params* p, float* f_out, float* g_out, int B, int D, int H, int W,
                              lbmff_stream_t stream);

/* ---- a16  LBMMagnetic{2,3}d.get_H_int (LBM_magnetic_3d.py:125-188, LBM_magnetic_2d.py:105-164)
 *      H_ext_mac_*: face arrays (W+1 / H+1 / D+1 along their axis); hz may be NULL in 2-D.
 *      psi_ws: B*D*H*W floats of scratch (receives psi).  Outputs H_int [dim], h_out [Q]. */
int lbmff_get_H_int_d2q9(const float* phi, const uint8_t* flags, const float* Hx, const float* Hy, const float* h,
                         double tau, double dx, double dt, float* H_int, float* h_out, float* psi_ws, int B, int H,
                         int W, lbmff_stream_t stream);
int lbmff_get_H_int_d3q19(const float* phi, const uint8_t* flags, const float* Hx, const float* Hy,
                          const float* Hz, const float* h, double tau, double dx, double dt, float* H_int,
                          float* h_out, float* psi_ws, int B, int D, int H, int W, lbmff_stream_t stream);

/* ---- allocation-free variant of lbmff_get_H_int_d2q9.  It has NO counterpart in the reference: it was added
 *      for callers that cannot allocate (CUDA graph capture, tight serving loops,
 *      harnesses that forbid cudaMalloc).  Same arguments and results as lbmff_get_H_int_d2q9 minus psi_ws: the psi
 *      scratch is a file-scope static __device__ array of LBMFF_STATIC_PSI_MAX_FLOATS floats (4 MiB, held by every
 *      GPU that loads the library until exit).  B*H*W beyond that limit is rejected (cudaErrorInvalidValue, message
 *      in lbmff_last_error_string).  The call performs no cudaMalloc / cudaFree / allocating library routine.  All
 *      calls on one device share the array: calls that overlap on the GPU (different streams) race; calls on one
 *      stream are safe.  Nothing in it has to start at zero (every cell's psi is written before it is read). */
#define LBMFF_STATIC_PSI_MAX_FLOATS (1 << 20)
int lbmff_static_psi_max_floats(void);
int lbmff_get_H_int_d2q9_static(const float* phi, const uint8_t* flags, const float* Hx, const float* Hy,
                                const float* h, double tau, double dx, double dt, float* H_int, float* h_out, int B,
                                int H, int W, lbmff_stream_t stream);

/* =====================================================================================
 * Fused step tier.  The reference has no counterpart function: its time loop lives in the
 * demo scripts (demo_3d_LBM_Rosensweig_instability.py:151-209, demo_2d_LBM.py:101-121).
 * These entry points execute the same sequence of reference ops with far fewer passes over
 * HBM; lbm_ferrofluid_b200.solver drives them.  State between steps: post-collision f, g
 * (pre-stream), rho (shell values persist), pressure, vel, and the magnetic lattice.
 * ===================================================================================== */

/* rho <- sum_q f_q(x - e_q) on non-OBSTACLE cells, in place (propagation + macro_compute); when chi_out is
 * not NULL also chi = 0.33 (1 - H_eps(phi)) on every cell, phi from density(rho) (demo :162,
 * LBM_magnetic_3d.py:162). */
int lbmff_hcz_moments_d2q9(const float* f, const uint8_t* flags, float* rho, float* chi_out,
                           const lbmff_fluid_params* p, int B, int H, int W, lbmff_stream_t stream);
int lbmff_hcz_moments_d3q19(const float* f, const uint8_t* flags, float* rho, float* chi_out,
                            const lbmff_fluid_params* p, int B, int D, int H, int W, lbmff_stream_t stream);

/* Sweep-invariant magnetic source rhs = div(chi_face H_face) dx / (1 + chi) on FLUID cells of mflags, else 0
 * (LBM_magnetic_3d.py:161-179), from the cell-centred chi. */
int lbmff_magnetic_source_2d(const float* chi, const uint8_t* mflags, const float* Hx, const float* Hy, double dx,
                             float* rhs_out, int B, int H, int W, lbmff_stream_t stream);
int lbmff_magnetic_source_3d(const float* chi, const uint8_t* mflags, const float* Hx, const float* Hy,
                             const float* Hz, double dx, float* rhs_out, int B, int D, int H, int W,
                             lbmff_stream_t stream);

/* n_sweeps population-faithful sweeps (propagation + rebounce_obstacle + get_H_int collision per sweep,
 * one pass over h each).  Ping-pong hA -> hB -> hA ...: the result is in hA if n_sweeps is even, else hB.
 * psi_out receives psi of the LAST sweep (the one H_int is taken from, demo :166-178). */
int lbmff_magnetic_sweeps_d2q9(float* hA, float* hB, const uint8_t* mflags, const float* rhs, float* psi_out,
                               int n_sweeps, double tau, double dx, double dt, int B, int H, int W,
                               lbmff_stream_t stream);
int lbmff_magnetic_sweeps_d3q19(float* hA, float* hB, const uint8_t* mflags, const float* rhs, float* psi_out,
                                int n_sweeps, double tau, double dx, double dt, int B, int D, int H, int W,
                                lbmff_stream_t stream);

/* tau == 1 scalar form of the same sweeps (valid when every OBSTACLE cell of mflags is on the outer
 * shell): state a[2][N] = outgoing value of the axis / diagonal weight class.  aA holds the latest state,
 * aB the one before; result in aA if n_sweeps is even, else aB (the other buffer then holds "one before").
 * Bit-identical psi to lbmff_magnetic_sweeps_* on non-OBSTACLE cells. */
int lbmff_magnetic_sweeps_scalar_2d(float* aA, float* aB, const uint8_t* mflags, const float* rhs, float* psi_out,
                                    int n_sweeps, double dx, double dt, int B, int H, int W,
                                    lbmff_stream_t stream);
int lbmff_magnetic_sweeps_scalar_3d(float* aA, float* aB, const uint8_t* mflags, const float* rhs, float* psi_out,
                                    int n_sweeps, double dx, double dt, int B, int D, int H, int W,
                                    lbmff_stream_t stream);
/* Rebuild the Q-population h from the scalar state (a_cur latest, a_prev one sweep earlier). */
int lbmff_magnetic_materialize_h_2d(const float* a_cur, const float* a_prev, const float* psi, const float* rhs,
                                    const uint8_t* mflags, float* h_out, double dx, double dt, int B, int H, int W,
                                    lbmff_stream_t stream);
int lbmff_magnetic_materialize_h_3d(const float* a_cur, const float* a_prev, const float* psi, const float* rhs,
                                    const uint8_t* mflags, float* h_out, double dx, double dt, int B, int D, int H,
                                    int W, lbmff_stream_t stream);

/* capillary_process part 1: ordered wetting boundary (rho IN PLACE) then ws = [lap(density) | fai | prho],
 * 3 * B*D*H*W floats (HCZ_3d.py:73-220).  `pressure` is the previous step's. */
int lbmff_hcz_capillary_fields_2d(float* rho, const uint8_t* flags, const float* pressure,
                                  const lbmff_fluid_params* p, float* ws, int B, int H, int W,
                                  lbmff_stream_t stream);
int lbmff_hcz_capillary_fields_3d(float* rho, const uint8_t* flags, const float* pressure,
                                  const lbmff_fluid_params* p, float* ws, int B, int D, int H, int W,
                                  lbmff_stream_t stream);

/* propagation + rebounce_obstacle of f and g, capillary_process part 2 (forces, vel, pressure) and the
 * HCZ collision, in one pass: reads f_in/g_in (pulled), writes f_out/g_out.  chi/psi/mflags NULL = no
 * magnetic force.  H_ext: cell-centred [B][dim][...] or NULL with H_uniform[dim].  vel, pressure are
 * updated IN PLACE on FLUID cells; force_out may be NULL.  chi and psi get the get_grad shell rule folded in
 * IN PLACE first (OBSTACLE shell cells of flags / mflags take the nearest interior value, the value
 * get_grad would read there; those cells are never observed otherwise). */
int lbmff_hcz_collide_stream_d2q9(const float* f_in, const float* g_in, float* f_out, float* g_out,
                                  const float* rho, const uint8_t* flags, const uint8_t* mflags, const float* ws,
                                  float* chi, float* psi, const float* H_ext, const double* H_uniform,
                                  const lbmff_fluid_params* p, int kbc_on_g, float* vel, float* pressure,
                                  float* force_out, int B, int H, int W, lbmff_stream_t stream);
int lbmff_hcz_collide_stream_d3q19(const float* f_in, const float* g_in, float* f_out, float* g_out,
                                   const float* rho, const uint8_t* flags, const uint8_t* mflags, const float* ws,
                                   float* chi, float* psi, const float* H_ext, const double* H_uniform,
                                   const lbmff_fluid_params* p, float* vel, float* pressure, float* force_out,
                                   int B, int D, int H, int W, lbmff_stream_t stream);

/* Single-phase step in one pass: propagation + macro_compute + rebounce_obstacle + collision
 * (demo_2d_LBM.py:101-121).  kbc < 0: BGK; kbc = 0 / 1: LBGK-split / KBC entropic collision (D2Q9).
 * rho, vel updated IN PLACE on non-OBSTACLE cells; force may be NULL. */
int lbmff_stream_collide_d2q9(const float* f_in, float* f_out, const uint8_t* flags, const float* force, float* rho,
                              float* vel, double tau, int kbc, double dx, double dt, int B, int H, int W,
                              lbmff_stream_t stream);
int lbmff_stream_collide_d3q19(const float* f_in, float* f_out, const uint8_t* flags, const float* force,
                               float* rho, float* vel, double tau, double dx, double dt, int B, int D, int H, int W,
                               lbmff_stream_t stream);

/* Shan-Chen multiphase step on D2Q9 in two passes (demo_2d_LBM_multiphase_SC.py:113-131: propagation, macro_compute,
 * rebounce_obstacle, LBMCollisionSC2d.collision = calculate_force + BGK, LBM_collision_SC_2d.py:22-137).
 * lbmff_sc_moments_d2q9: rho <- sum of the pulled f on non-OBSTACLE cells (in place), density = affine(rho)
 * (LBM_macro_compute_2d.py:85-106), psi = C-S pseudopotential of density, the wall value on OBSTACLE cells.
 * lbmff_sc_collide_stream_d2q9: pull, full-way bounce-back on OBSTACLE, vel (in place on non-OBSTACLE cells),
 * interaction force on FLUID cells from the psi field (zero outside the array), BGK with p->tau.  force_out may be
 * NULL.  p supplies dx, dt, tau and the density map (density_liquid / density_gas / rho_liquid / rho_gas). */
int lbmff_sc_moments_d2q9(const float* f, const uint8_t* flags, float* rho, float* density, float* psi,
                          const lbmff_fluid_params* p, int B, int H, int W, lbmff_stream_t stream);
int lbmff_sc_collide_stream_d2q9(const float* f_in, float* f_out, const uint8_t* flags, const float* psi,
                                 const float* rho, float* vel, float* force_out, const lbmff_fluid_params* p, int B,
                                 int H, int W, lbmff_stream_t stream);

/* =====================================================================================
 * SLAB TIER -- z-slab domain decomposition of the fused 3-D step (one process per GPU).
 * No reference counterpart: the reference is single-device ("Parallelism").
 * The local arrays hold D planes = halo planes + owned planes + halo planes of a globally periodic domain
 * of Dg planes; local plane z is global plane (z + zoff) mod Dg.  Each call processes local planes
 * [zb, zb + zn) only, so a step can be issued "boundary planes, start the halo exchange, interior planes".
 * Position-dependent rules (outer shell, replicate clamps, wetting faces) use the global z, so a P-slab run
 * performs bit-for-bit the per-cell arithmetic of the one-GPU run.  Halo planes are filled by the caller
 * (lbm_ferrofluid_b200.distributed: NCCL point-to-point over NVLink); these entries never communicate.
 * ===================================================================================== */
typedef struct lbmff_slab {
  int Dg;   /* global number of z planes */
  int zoff; /* global z of local plane 0; negative for the lower halo of the first slab (wraps) */
  int zb;   /* first local plane processed */
  int zn;   /* number of local planes processed */
} lbmff_slab;

/* lbmff_hcz_moments_d3q19 on a plane range. */
int lbmff_slab_hcz_moments_d3q19(const float* f, const uint8_t* flags, float* rho, float* chi_out,
                                 const lbmff_fluid_params* p, const lbmff_slab* slab, int B, int D, int H, int W,
                                 lbmff_stream_t stream);
/* lbmff_magnetic_source_3d on a plane range (face arrays are slab-local: Hx [D][H][W+1], Hy [D][H+1][W],
 * Hz [D+1][H][W]). */
int lbmff_slab_magnetic_source_3d(const float* chi, const uint8_t* mflags, const float* Hx, const float* Hy,
                                  const float* Hz, double dx, float* rhs_out, const lbmff_slab* slab, int B, int D,
                                  int H, int W, lbmff_stream_t stream);
/* ONE population-faithful sweep h_in -> h_out on a plane range; psi_out NULL = psi not stored. */
int lbmff_slab_magnetic_sweep_d3q19(const float* h_in, float* h_out, const uint8_t* mflags, const float* rhs,
                                    float* psi_out, double tau, double dx, double dt, const lbmff_slab* slab, int B,
                                    int D, int H, int W, lbmff_stream_t stream);
/* The same sweep for a BOUNDARY plane range with the halo "send" fused in: the populations with e_z = direction (+1:
 * top owned plane -> upper neighbour's lower halo, -1: bottom owned plane -> lower neighbour's upper halo) are also
 * stored into peer_out, the neighbour's h_out array mapped into this process over NVLink (e.g. torch symmetric
 * memory), at cell index + peer_shift_cells.  No reference counterpart; distributed.SymmPushExchanger is the caller. */
int lbmff_slab_magnetic_sweep_push_d3q19(const float* h_in, float* h_out, const uint8_t* mflags, const float* rhs,
                                         float* psi_out, double tau, double dx, double dt, float* peer_out,
                                         int direction, long long peer_shift_cells, const lbmff_slab* slab, int B,
                                         int D, int H, int W, lbmff_stream_t stream);
/* ONE tau == 1 scalar sweep (a_cur latest state, a_prev_out the one before, overwritten with the new state). */
int lbmff_slab_magnetic_sweep_scalar_3d(const float* a_cur, float* a_prev_out, const uint8_t* mflags,
                                        const float* rhs, float* psi_out, double dx, double dt,
                                        const lbmff_slab* slab, int B, int D, int H, int W, lbmff_stream_t stream);
/* The same sweeps in psi-only form: the intermediate states of a step are psi alone (first half of
 * the two-float arrays), the neighbours' two values being rebuilt from psi and rhs while a plane is staged.  ONE sweep
 * on a plane range; src / dst alternate between the two state arrays like a_cur / a_prev_out above.  The caller chains
 *   sweep 1: (nbr, wall, mask) = (0, 0, 0)   reads the two-float state src, wall returns from the two-float dst
 *   sweep 2: (1, 0, 1)                       reads psi + rhs, wall returns from the two-float dst
 *   sweep 3..n: (1, 1, 1)                    reads psi + rhs, wall returns rebuilt from the psi in dst
 * and needs psi (first half of dst) and rhs valid one halo plane around the range.  wmask: the second half of the
 * array sweep 1 writes (dead until the expand call); sweep 1 parks the wall masks of shell-adjacent cells there.
 * lbmff_slab_magnetic_psi_expand_3d turns the last two psi fields (a_cur newest) back into the two-float state and
 * writes psi_out (may be NULL) -- bit-identical to n calls of lbmff_slab_magnetic_sweep_scalar_3d.
 * peer_dst (may be NULL): psi is also stored at cell + peer_shift_cells of the neighbour rank's dst array mapped into
 * this process (boundary planes; the fused halo push of lbmff_slab_magnetic_sweep_push_d3q19 for this form). */
int lbmff_slab_magnetic_sweep_psi_3d(const float* src, float* dst, const uint8_t* mflags, const float* rhs,
                                     float* wmask, int nbr_mode, int wall_mode, int mask_mode, double dx, double dt,
                                     float* peer_dst, long long peer_shift_cells, const lbmff_slab* slab, int B, int D,
                                     int H, int W, lbmff_stream_t stream);
int lbmff_slab_magnetic_psi_expand_3d(float* a_cur, float* a_prev, const uint8_t* mflags, const float* rhs,
                                      float* psi_out, double dx, double dt, const lbmff_slab* slab, int B, int D,
                                      int H, int W, lbmff_stream_t stream);
/* get_grad shell rule folded into the data, in place: OBSTACLE shell cells take the nearest interior value
 * (LBM_collision_3d.py:217-218). */
int lbmff_slab_shell_rule_3d(float* field, const uint8_t* flags, const lbmff_slab* slab, int B, int D, int H, int W,
                             lbmff_stream_t stream);
/* Ordered wetting boundary on rho IN PLACE (HCZ_3d.py:84-211) for the shell cells of the plane range.
 * stage 0: everything; 1: x faces only; 2: y faces, z faces, edges, corners (y faces read x-face cells of the
 * planes above and below, so stage 1 must cover one more plane on each side than stage 2). */
int lbmff_slab_hcz_wetting_3d(float* rho, const uint8_t* flags, const lbmff_fluid_params* p, int stage,
                              const lbmff_slab* slab, int B, int D, int H, int W, lbmff_stream_t stream);
/* ws = [lap(density) | fai | prho] on a plane range (second half of lbmff_hcz_capillary_fields_3d). */
int lbmff_slab_hcz_fields_3d(const float* rho, const uint8_t* flags, const float* pressure,
                             const lbmff_fluid_params* p, float* ws, const lbmff_slab* slab, int B, int D, int H,
                             int W, lbmff_stream_t stream);
/* lbmff_hcz_collide_stream_d3q19 on a plane range, uniform applied field; chi and psi must already carry the
 * shell rule (lbmff_slab_shell_rule_3d) and valid halo planes. */
int lbmff_slab_hcz_collide_stream_d3q19(const float* f_in, const float* g_in, float* f_out, float* g_out,
                                        const float* rho, const uint8_t* flags, const uint8_t* mflags,
                                        const float* ws, const float* chi, const float* psi,
                                        const double* H_uniform, const lbmff_fluid_params* p, float* vel,
                                        float* pressure, float* force_out, const lbmff_slab* slab, int B, int D,
                                        int H, int W, lbmff_stream_t stream);

/* ---- converged magnetic potential (opt-in; NOT a drop-in: the reference never converges) ----------------
 * The reference relaxes the Poisson lattice with a fixed 30 warm-started sweeps per outer step
 * (demo_3d_LBM_Rosensweig_instability.py:163-173; LBM_magnetic_3d.py:125-188).  These entries solve for the FIXED
 * POINT of those sweeps at tau == 1 instead:  L psi = b  with the wall-aware D2Q9 / D3Q19 weighted graph Laplacian
 * (full-way bounce-back = mirror), by conjugate gradients entirely on the device (dot
 * products reduced in double in a fixed order: deterministic; no host synchronisation).  psi is defined up to a
 * constant -- only its gradient is used (H_int = -grad psi, LBM_magnetic_3d.py:183-186).  If sum(b) != 0 the
 * sweeps' stationary state gains a constant c per sweep; the entry solves for that state (b is replaced by
 * b - c (1 + omega), omega = wall-weight fraction of the cell; csrc/poisson_cg.cu) and reports c.
 *   rhs      [B][N]  sweep-invariant source of lbmff_magnetic_source_*            (in)
 *   psi      [B][N]  warm start in, solution out (OBSTACLE cells of mflags untouched)
 *   work     [B][5][N] floats, scratch [B][lbmff_poisson_cg_scratch_doubles(D,H,W)] doubles    (device workspace)
 *   max_iters, tol   stop when |r| <= tol * |b - c(1+omega)| or after max_iters iterations.  The iterations run as
 *                    a CUDA-graph WHILE node whose condition the reduction kernel sets on the device: one graph
 *                    launch per batch item, nothing is launched after convergence, no host synchronisation
 *                    (LBMFF_CG_NO_GRAPH=1 in the environment launches max_iters iterations one by one instead)
 *   res_hist [B][max_iters + 1] floats or NULL: |r| after each iteration (entry 0 = warm start)
 *   info     [B][5] doubles or NULL: iterations done, converged (0/1), |r|^2, |b - c(1+omega)|^2, c
 * tau must be 1.0 (for tau != 1 the fixed point is not a compact stencil): anything else is rejected. */
int lbmff_poisson_cg_scratch_doubles(int D, int H, int W);
int lbmff_poisson_cg_2d(const float* rhs, const uint8_t* mflags, float* psi, float* work, double* scratch,
                        int max_iters, double tol, double tau, double dx, double dt, float* res_hist, double* info,
                        int B, int H, int W, lbmff_stream_t stream);
int lbmff_poisson_cg_3d(const float* rhs, const uint8_t* mflags, float* psi, float* work, double* scratch,
                        int max_iters, double tol, double tau, double dx, double dt, float* res_hist, double* info,
                        int B, int D, int H, int W, lbmff_stream_t stream);

/* The same solve with a geometric-multigrid preconditioner (csrc/poisson_mg.cuh): one V(2,2) cycle per iteration --
 * 2 x 2 (x 2) cell aggregation, piecewise-constant transfer, Galerkin coarse operators on the same lattice directions
 * (walls = zero-weight links), damped-Jacobi smoothing, the whole cycle inside the same device-side WHILE loop.
 *   work     [B][7][N] floats (the five of lbmff_poisson_cg_* + z + one smoother buffer)
 *   mg_work  [B][lbmff_poisson_mg_workspace_floats(D,H,W)] floats (levels >= 1: weights, diagonal, x, t, b)
 *   reuse_hierarchy: the coarse operators depend on mflags only; non-zero = mg_work still holds the hierarchy an
 *            earlier call built for the SAME mflags and extents (skips the set-up kernels), 0 = build it.
 * Coarse operators carry the over-correction factor 1/2 per level (unsmoothed aggregation); the level-0 smoother in
 * 3-D is a shared-memory tiled kernel marching along z (LBMFF_MG_TILED=0 selects the generic one).
 * Everything else as lbmff_poisson_cg_*; the stopping test is on the unpreconditioned residual |r|. */
long long lbmff_poisson_mg_workspace_floats(int D, int H, int W);
int lbmff_poisson_pcg_2d(const float* rhs, const uint8_t* mflags, float* psi, float* work, float* mg_work,
                         double* scratch, int max_iters, double tol, double tau, double dx, double dt, float* res_hist,
                         double* info, int reuse_hierarchy, int B, int H, int W, lbmff_stream_t stream);
int lbmff_poisson_pcg_3d(const float* rhs, const uint8_t* mflags, float* psi, float* work, float* mg_work,
                         double* scratch, int max_iters, double tol, double tau, double dx, double dt, float* res_hist,
                         double* info, int reuse_hierarchy, int B, int D, int H, int W, lbmff_stream_t stream);

/* ---- iso-surface of the order parameter on the device (replaces the host round trip
 *      mcubes.marching_cubes((-phi * s).cpu().numpy()[0, 0], 0) of src/LBM/utils/io.py:50-63).
 * u [D][H][W] fp32 (device).  Marching cubes with a case table derived at load time (csrc/iso_surface.cu; PyMCubes is
 * not vendored by the reference), triangle soup out: triangles [F][3 vertices][3] in array-index coordinates
 * (axis 0, axis 1, axis 2), the order PyMCubes returns.  Two passes, deterministic triangle order, no atomics:
 *   count: block_offsets [lbmff_iso_surface_blocks(D,H,W)] ints receive the exclusive scan of the per-block triangle
 *          counts, *total (device int) the number of triangles F;
 *   emit : writes min(F, capacity) triangles.  flip = 0: counter-clockwise normal towards DECREASING u (csrc/iso_surface.cu explains the choice).
 * lbmff_mc_case_table is host-only (no GPU): ntri [256], edges [256][24] cube-edge ids (-1 padded), for tests. */
int lbmff_mc_case_table(int* ntri, int* edges);
int lbmff_iso_surface_blocks(int D, int H, int W);
int lbmff_iso_surface_count_3d(const float* u, double iso, int* block_offsets, int* total, int D, int H, int W,
                               lbmff_stream_t stream);
int lbmff_iso_surface_emit_3d(const float* u, double iso, int flip, const int* block_offsets, float* triangles,
                              int capacity, int D, int H, int W, lbmff_stream_t stream);

/* ---- soft rasteriser, forward (renderutils/cuda/soft_rasterize_cuda.cpp:72-110 ->
 *      soft_rasterize_cuda_kernel.cu:236-475, as driven by renderutils/soft_rasterize.py:26-110).
 * Same arguments as the reference's forward_soft_rasterize: face_vertices [B][F][3][3] projected vertices (x, y in
 * [-1, 1], z = camera depth), textures [B][F][texture_size][3], soft_colors [B][4][is][is] PRE-FILLED by the caller
 * with the background colour (alpha plane 1), aggrs_info [B][2][is][is] or NULL, the scalar options with the
 * reference's meaning (func_id_dist 0 hard / 1 barycentric / 2 euclidean, func_id_rgb 0 hard / 1 softmax,
 * func_id_alpha 0 hard / 1 sum / 2 prod, texture_sample_type 0 surface / 1 vertex; dist_eps is the already
 * transformed log(1/dist_eps - 1) of soft_rasterize.py:54).  Instead of every pixel walking all faces, faces are
 * binned into 8 x 8 pixel tiles first (ordered lists, so a pixel folds its faces in the reference's order):
 *   bin    : fills face_info [B][F][24], face_box [B * F + B * ceil(F / 32)][4] (face boxes, then the union boxes
 *            of groups of 32 consecutive faces) and tile_offsets [lbmff_soft_rasterize_offsets(B, is)] (exclusive
 *            scan of the face counts per tile and binning segment, last entry = total list length L);
 *   forward: tile_faces [L] ints of scratch; writes soft_colors (and aggrs_info). */
int lbmff_soft_rasterize_tiles(int image_size);
int lbmff_soft_rasterize_offsets(int B, int image_size);
int lbmff_soft_rasterize_bin(const float* face_vertices, float* face_info, float* face_box, int* tile_offsets,
                             int B, int F, int image_size, double sigma_val, double dist_eps,
                             lbmff_stream_t stream);
int lbmff_soft_rasterize_forward(const float* face_vertices, const float* textures, const float* face_info,
                                 const float* face_box, const int* tile_offsets, int* tile_faces, float* soft_colors,
                                 float* aggrs_info, int B, int F, int image_size, int texture_size, double near_,
                                 double far_, double eps, double sigma_val, int func_id_dist, double dist_eps,
                                 double gamma_val, int func_id_rgb, int func_id_alpha, int texture_sample_type,
                                 int double_side, lbmff_stream_t stream);

/* =====================================================================================
 * EXTENSION TIER -- named by the north star, absent from the reference (its only 3-D
 * lattice is D3Q19, LBM_collision_3d.py:76-103; its only forcing is the velocity shift of
 * LBM_collision_3d.py:141-143; its streaming is out of place, LBM_propagation_3d.py:18-111).
 * Single-phase step in one pass (propagation + macro_compute + rebounce_obstacle + BGK collision with the
 * reference's entropic product equilibrium) on D3Q19 or D3Q27, in the reference's [B][Q][D][H][W] layout.
 * D3Q27 keeps the reference's D3Q19 order for q < 19 and appends the corners as opposite pairs
 * (lbmff_ext_lattice_info returns the tables; host-only, needs no device).
 *   collision 0 BGK | 1 entropic KBC (D3Q27 only): the reference's D2Q9 scheme (LBM_collision_MRT_2d.py:113-151,
 *             184-226; s = {T, N, Pi} from centred second moments, stabiliser gamma per cell) carried to 3-D;
 *             the reference's own factory has no 3-D KBC (simulation_runner.py:107-108)
 *   forcing   0 none | 1 velocity shift u + tau F / rho (reference) | 2 Guo (half-force velocity + source term;
 *             vel then holds the physical velocity (sum f e + F dt / 2) / rho)
 *   force     per-cell [B][3][D][H][W], or NULL to use force_uniform[3] (x, y, z)
 *   streaming 0 A/B pull (f_in != f_out) | 1 esoteric pull, IN PLACE (f_in == f_out): one buffer instead of
 *             two; parity = number of steps already taken on this buffer (& 1).  The slot layout of an
 *             esoteric buffer differs from the standard one; lbmff_ext_esoteric_convert_* translates
 *             (to_esoteric = 1: standard -> layout expected by a step of that parity; 0: back), out of place.
 * rho, vel are updated IN PLACE on non-OBSTACLE cells. */
int lbmff_ext_lattice_info(int Q, int* e_out, int* opp_out, double* w_out);
/* Shan-Chen interaction force on D3Q19: lbmff_sc_force_d2q9 (LBM_collision_SC_2d.py:22-97: C-S pseudopotential,
 * G = -1/3, wall psi from the mean density, zero outside the array, force on FLUID cells) with the D3Q19
 * weights; the reference's Shan-Chen model is 2-D only.  force_out: [B][3][D][H][W]. */
int lbmff_sc_force_d3q19(const float* density, const uint8_t* flags, double density_liquid, double density_gas,
                         double dx, double dt, float* force_out, int B, int D, int H, int W, lbmff_stream_t stream);
int lbmff_ext_stream_collide_d3q19(const float* f_in, float* f_out, const uint8_t* flags, const float* force,
                                   const double* force_uniform, float* rho, float* vel, double tau, int collision,
                                   int forcing, int streaming, int parity, double dx, double dt, int B, int D, int H,
                                   int W, lbmff_stream_t stream);
int lbmff_ext_stream_collide_d3q27(const float* f_in, float* f_out, const uint8_t* flags, const float* force,
                                   const double* force_uniform, float* rho, float* vel, double tau, int collision,
                                   int forcing, int streaming, int parity, double dx, double dt, int B, int D, int H,
                                   int W, lbmff_stream_t stream);
/* Three-rate MRT collision on the same step (D3Q19 and D3Q27, fp32 storage): the non-equilibrium part is split
 * into its deviatoric second-order moments (relaxed with tau: shear viscosity cs2 (tau - 1/2) dt, as in BGK), the
 * trace of the second-order moments (tau_bulk: bulk viscosity) and everything of higher order (tau_high).  All
 * three equal reproduces BGK.  No reference counterpart (the reference's "MRT" class is the 2-D KBC scheme,
 * LBM_collision_MRT_2d.py); north-star item (1). */
int lbmff_ext_stream_collide_mrt_d3q19(const float* f_in, float* f_out, const uint8_t* flags, const float* force,
                                       const double* force_uniform, float* rho, float* vel, double tau,
                                       double tau_bulk, double tau_high, int forcing, int streaming, int parity,
                                       double dx, double dt, int B, int D, int H, int W, lbmff_stream_t stream);
int lbmff_ext_stream_collide_mrt_d3q27(const float* f_in, float* f_out, const uint8_t* flags, const float* force,
                                       const double* force_uniform, float* rho, float* vel, double tau,
                                       double tau_bulk, double tau_high, int forcing, int streaming, int parity,
                                       double dx, double dt, int B, int D, int H, int W, lbmff_stream_t stream);
int lbmff_ext_esoteric_convert_d3q19(const float* src, float* dst, int parity, int to_esoteric, int B, int D, int H,
                                     int W, lbmff_stream_t stream);
int lbmff_ext_esoteric_convert_d3q27(const float* src, float* dst, int parity, int to_esoteric, int B, int D, int H,
                                     int W, lbmff_stream_t stream);

/* Mixed-precision population storage: the same step (BGK only) with populations STORED as fp16 and all arithmetic
 * in fp32.  The buffers hold the shifted populations f_q - w_q rho0 (rho0: a reference density near the mean), so
 * the 10-bit mantissa is spent on the part of f that carries the flow; bytes per cell update halve.
 * f_in / f_out: [B][Q][D][H][W] fp16.  pack / unpack translate from / to the standard fp32 populations;
 * layout 0 = standard slots (A/B pull), 1 / 2 = esoteric pull before an even / odd step. */
int lbmff_ext_stream_collide_fp16_d3q19(const void* f_in, void* f_out, const uint8_t* flags, const float* force,
                                        const double* force_uniform, float* rho, float* vel, double tau,
                                        int forcing, int streaming, int parity, double rho0, double dx, double dt,
                                        int B, int D, int H, int W, lbmff_stream_t stream);
int lbmff_ext_stream_collide_fp16_d3q27(const void* f_in, void* f_out, const uint8_t* flags, const float* force,
                                        const double* force_uniform, float* rho, float* vel, double tau,
                                        int forcing, int streaming, int parity, double rho0, double dx, double dt,
                                        int B, int D, int H, int W, lbmff_stream_t stream);
int lbmff_ext_pack_fp16_d3q19(const float* f_std, void* f_packed, int layout, double rho0, int B, int D, int H, int W,
                              lbmff_stream_t stream);
int lbmff_ext_pack_fp16_d3q27(const float* f_std, void* f_packed, int layout, double rho0, int B, int D, int H, int W,
                              lbmff_stream_t stream);
int lbmff_ext_unpack_fp16_d3q19(const void* f_packed, float* f_std, int layout, double rho0, int B, int D, int H, int W,
                                lbmff_stream_t stream);
int lbmff_ext_unpack_fp16_d3q27(const void* f_packed, float* f_std, int layout, double rho0, int B, int D, int H, int W,
                                lbmff_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* LBMFF_H */
