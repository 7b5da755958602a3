// Shan-Chen pseudopotential pieces shared by the drop-in force kernel (ops_collide.cu) and the fused step
// (fused_sc.cu): C-S equation of state pseudopotential psi = sqrt(2 (p_EOS - rho cs2) / (G cs2)), G = -1/3, wall psi
// from the mean of the liquid and gas densities (LBM_collision_SC_2d.py:20, 22-59).
#pragma once
#include <math.h>

#include "common.cuh"

namespace lbmff {

struct ScConsts {
  float RT, a, cs2, G1, negG1, c, psx_wall;
};
__device__ __forceinline__ float sc_psi(float d, const ScConsts& S) {
  const float t = d;  // density * b / 4 with b = 4
  const float om = 1.0f - t;
  const float v = S.RT * (1.0f + (4.0f * t - 2.0f * t * t) / (om * om * om)) - S.a * d - S.cs2;
  return sqrtf(2.0f * d * v / S.G1 / S.cs2);
}
inline ScConsts make_sc(double density_liquid, double density_gas, double dx, double dt) {
  const double c = dx / dt, cs2 = c * c / 3.0, RT = cs2, a = 12.0 * cs2, bb = 4.0, G1 = -1.0 / 3.0;
  const double dw = 0.5 * (density_gas + density_liquid);
  double pw = dw * bb / 4.0;
  pw = RT * (1.0 + (4.0 * pw - 2.0 * pw * pw) / ((1.0 - pw) * (1.0 - pw) * (1.0 - pw))) - a * dw - cs2;
  pw = sqrt(2.0 * dw * pw / G1 / cs2);
  ScConsts S;
  S.RT = (float)RT; S.a = (float)a; S.cs2 = (float)cs2; S.G1 = (float)G1; S.negG1 = (float)(-G1); S.c = (float)c;
  S.psx_wall = (float)pw;
  return S;
}


}  // namespace lbmff
