// Device functors of the flow path: the B200 restatement of SU2's CNumericsSIMD chain
// (SU2_CFD/include/numerics_simd/flow/convection/{roe,centered,common}.hpp,
//  flow/diffusion/{common,viscous_fluxes}.hpp, flow/variables.hpp) with Double -> double,
// one thread per edge. Shared by the edge kernels and the boundary kernels.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace su2b200 {

// Primitive layout (CEulerVariable::CIndices, CEulerVariable.hpp:51-72; variables.hpp:37-63):
// [T, v(nDim), P, rho, H, c, mu_lam, mu_t, k, cp]
template <int ND> struct PrimIdx {
  static constexpr int T = 0, VEL = 1, P = ND + 1, RHO = ND + 2, H = ND + 3, C = ND + 4,
                       MU = ND + 5, MUT = ND + 6, K = ND + 7, CP = ND + 8, NPRIM = ND + 9;
};

struct FlowParams {
  double gamma, gm1, R;          // gas
  double roe_kappa, entropy_fix; // ROE_KAPPA, ENTROPY_FIX_COEFF
  double muscl_kappa;            // MUSCL_KAPPA_FLOW
  double prandtl_turb;
  double jst_k2, jst_k4, jst_fix;
  double lax_k0;                 // LAX_SENSOR_COEFF
  int muscl, limiter, viscous, implicit;
  int lax;                       // centred scheme of an agglomerated multigrid level: CLaxScheme instead of CJSTScheme
};

template <int N>
__device__ __forceinline__ double dotN(const double* a, const double* b) {
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < N; ++i) s += a[i] * b[i];
  return s;
}

// inviscidProjFlux, common.hpp:397-410 (V = primitives, momentum = rho*v)
template <int ND>
__device__ __forceinline__ void inviscid_proj_flux(const double* V, const double* normal, double* flux) {
  using I = PrimIdx<ND>;
  double mdot = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) mdot += (V[I::RHO] * V[I::VEL + d]) * normal[d];
  flux[0] = mdot;
#pragma unroll
  for (int d = 0; d < ND; ++d) flux[d + 1] = mdot * V[I::VEL + d] + normal[d] * V[I::P];
  flux[ND + 1] = mdot * V[I::H];
}

// inviscidProjJac, common.hpp:415-449. jac row-major [NV][NV].
template <int ND, class JT>
__device__ __forceinline__ void inviscid_proj_jac(double gamma, const double* vel, double energy,
                                                  const double* normal, double scale, JT jac) {
  constexpr int NV = ND + 2;
  double projVel = dotN<ND>(vel, normal);
  double gm1 = gamma - 1;
  double phi = 0.5 * gm1 * dotN<ND>(vel, vel);
  double a1 = gamma * energy - phi;
  jac[0] = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) jac[d + 1] = scale * normal[d];
  jac[ND + 1] = 0.0;
#pragma unroll
  for (int i = 0; i < ND; ++i) {
    jac[(i + 1) * NV] = scale * (normal[i] * phi - vel[i] * projVel);
#pragma unroll
    for (int j = 0; j < ND; ++j)
      jac[(i + 1) * NV + j + 1] = scale * (normal[j] * vel[i] - gm1 * normal[i] * vel[j]);
    jac[(i + 1) * NV + i + 1] += scale * projVel;
    jac[(i + 1) * NV + ND + 1] = scale * gm1 * normal[i];
  }
  jac[(ND + 1) * NV] = scale * projVel * (phi - a1);
#pragma unroll
  for (int d = 0; d < ND; ++d) jac[(ND + 1) * NV + d + 1] = scale * (normal[d] * a1 - gm1 * vel[d] * projVel);
  jac[(ND + 1) * NV + ND + 1] = scale * gamma * projVel;
}

// pMatrix, common.hpp:281-336 (n = unit normal)
template <int ND>
__device__ __forceinline__ void p_matrix(double gamma, double rho, const double* v, double projVel, double c,
                                         const double* n, double* P) {
  constexpr int NV = ND + 2;
  const double vel2 = 0.5 * dotN<ND>(v, v);
  if (ND == 2) {
    P[0 * NV + 0] = 1.0;
    P[0 * NV + 1] = 0.0;
    P[1 * NV + 0] = v[0];
    P[1 * NV + 1] = rho * n[1];
    P[2 * NV + 0] = v[1];
    P[2 * NV + 1] = -rho * n[0];
    P[3 * NV + 0] = vel2;
    P[3 * NV + 1] = rho * (v[0] * n[1] - v[1] * n[0]);
  } else {
    P[0 * NV + 0] = n[0];
    P[0 * NV + 1] = n[1];
    P[0 * NV + 2] = n[ND - 1];
    P[1 * NV + 0] = v[0] * n[0];
    P[1 * NV + 1] = v[0] * n[1] - rho * n[ND - 1];
    P[1 * NV + 2] = v[0] * n[ND - 1] + rho * n[1];
    P[2 * NV + 0] = v[1] * n[0] + rho * n[ND - 1];
    P[2 * NV + 1] = v[1] * n[1];
    P[2 * NV + 2] = v[1] * n[ND - 1] - rho * n[0];
    P[3 * NV + 0] = v[ND - 1] * n[0] - rho * n[1];
    P[3 * NV + 1] = v[ND - 1] * n[1] + rho * n[0];
    P[3 * NV + 2] = v[ND - 1] * n[ND - 1];
    P[(NV - 1) * NV + 0] = vel2 * n[0] + rho * (v[1] * n[ND - 1] - v[ND - 1] * n[1]);
    P[(NV - 1) * NV + 1] = vel2 * n[1] - rho * (v[0] * n[ND - 1] - v[ND - 1] * n[0]);
    P[(NV - 1) * NV + 2] = vel2 * n[ND - 1] + rho * (v[0] * n[1] - v[1] * n[0]);
  }
  const double rhoOn2 = 0.5 * rho;
  const double rhoOnTwoC = rhoOn2 / c;
  P[0 * NV + ND] = rhoOnTwoC;
  P[0 * NV + ND + 1] = rhoOnTwoC;
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    P[(d + 1) * NV + ND] = rhoOnTwoC * v[d] + rhoOn2 * n[d];
    P[(d + 1) * NV + ND + 1] = rhoOnTwoC * v[d] - rhoOn2 * n[d];
  }
  P[(ND + 1) * NV + ND] = rhoOnTwoC * vel2 + rhoOn2 * (projVel + c / (gamma - 1));
  P[(ND + 1) * NV + ND + 1] = rhoOnTwoC * vel2 - rhoOn2 * (projVel - c / (gamma - 1));
}

// pMatrixInv, common.hpp:341-392
template <int ND>
__device__ __forceinline__ void p_matrix_inv(double gamma, double rho, const double* v, double projVel, double c,
                                             const double* n, double* Pi) {
  constexpr int NV = ND + 2;
  const double c2 = c * c;
  const double vel2 = 0.5 * dotN<ND>(v, v);
  const double oneOnRho = 1 / rho;
  if (ND == 2) {
    double tmp = (gamma - 1) / c2;
    Pi[0] = 1.0 - tmp * vel2;
    Pi[1] = tmp * v[0];
    Pi[2] = tmp * v[1];
    Pi[3] = -tmp;
    Pi[NV + 0] = (n[0] * v[1] - n[1] * v[0]) * oneOnRho;
    Pi[NV + 1] = n[1] * oneOnRho;
    Pi[NV + 2] = -n[0] * oneOnRho;
    Pi[NV + 3] = 0.0;
  } else {
    constexpr int Z = ND - 1;  // == 2 in 3D (kept symbolic so the 2D instantiation compiles)
    double tmp = (gamma - 1) / c2 * n[0];
    Pi[0] = n[0] - tmp * vel2 - (n[Z] * v[1] - n[1] * v[Z]) * oneOnRho;
    Pi[1] = tmp * v[0];
    Pi[2] = tmp * v[1] + n[Z] * oneOnRho;
    Pi[3] = tmp * v[Z] - n[1] * oneOnRho;
    Pi[NV - 1] = -tmp;
    tmp = (gamma - 1) / c2 * n[1];
    Pi[NV + 0] = n[1] - tmp * vel2 + (n[Z] * v[0] - n[0] * v[Z]) * oneOnRho;
    Pi[NV + 1] = tmp * v[0] - n[Z] * oneOnRho;
    Pi[NV + 2] = tmp * v[1];
    Pi[NV + 3] = tmp * v[Z] + n[0] * oneOnRho;
    Pi[NV + NV - 1] = -tmp;
    tmp = (gamma - 1) / c2 * n[Z];
    Pi[2 * NV + 0] = n[Z] - tmp * vel2 - (n[1] * v[0] - n[0] * v[1]) * oneOnRho;
    Pi[2 * NV + 1] = tmp * v[0] + n[1] * oneOnRho;
    Pi[2 * NV + 2] = tmp * v[1] - n[0] * oneOnRho;
    Pi[2 * NV + 3] = tmp * v[Z];
    Pi[2 * NV + NV - 1] = -tmp;
  }
  const double g = (gamma - 1) / (rho * c);
#pragma unroll
  for (int r = ND; r < ND + 2; ++r) {
    const double sign = (r == ND) ? 1.0 : -1.0;
    Pi[r * NV + 0] = -sign * projVel * oneOnRho + g * vel2;
#pragma unroll
    for (int d = 0; d < ND; ++d) Pi[r * NV + d + 1] = sign * n[d] * oneOnRho - g * v[d];
    Pi[r * NV + ND + 1] = g;
  }
}

// Roe flux + approximate Jacobians: CRoeBase::ComputeFlux + CRoeScheme::finalizeFlux
// (roe.hpp:87-214, :244-303), static grid, dissipation = 1 (NO_ROELOWDISS).
// Vi/Vj: (reconstructed) primitives, entries [0, ND+4) used. normal = area vector i->j.
// Outputs flux[NV]; if IMPLICIT also jac_i, jac_j (row-major NVxNV).
template <int ND, bool IMPLICIT, class JT>
__device__ __forceinline__ void roe_flux(const FlowParams& p, const double* Vi, const double* Vj,
                                         const double* normal, double* flux, JT jac_i, JT jac_j) {
  using I = PrimIdx<ND>;
  constexpr int NV = ND + 2;
  const double area = sqrt(dotN<ND>(normal, normal));
  double un[ND];
#pragma unroll
  for (int d = 0; d < ND; ++d) un[d] = normal[d] / area;

  // conservative variables (compressibleConservatives, variables.hpp:91-101)
  double Ui[NV], Uj[NV];
  Ui[0] = Vi[I::RHO];
  Uj[0] = Vj[I::RHO];
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    Ui[d + 1] = Vi[I::RHO] * Vi[I::VEL + d];
    Uj[d + 1] = Vj[I::RHO] * Vj[I::VEL + d];
  }
  Ui[ND + 1] = Vi[I::RHO] * Vi[I::H] - Vi[I::P];
  Uj[ND + 1] = Vj[I::RHO] * Vj[I::H] - Vj[I::P];

  // Roe averages (roeAveragedVariables, variables.hpp:116-130)
  const double R = sqrt(Vj[I::RHO] / Vi[I::RHO]);
  const double D = 1 / (R + 1);
  const double rrho = R * Vi[I::RHO];
  double rv[ND];
#pragma unroll
  for (int d = 0; d < ND; ++d) rv[d] = (R * Vj[I::VEL + d] + Vi[I::VEL + d]) * D;
  const double rH = (R * Vj[I::H] + Vi[I::H]) * D;
  const double rc = sqrt((p.gamma - 1) * (rH - 0.5 * dotN<ND>(rv, rv)));
  const double rq = dotN<ND>(rv, un);

  double lambda[NV];
#pragma unroll
  for (int d = 0; d < ND; ++d) lambda[d] = rq;
  lambda[ND] = rq + rc;
  lambda[ND + 1] = rq - rc;
  const double maxLambda = fabs(rq) + rc;
#pragma unroll
  for (int k = 0; k < NV; ++k) lambda[k] = fmax(fabs(lambda[k]), p.entropy_fix * maxLambda);

  double fi[NV], fj[NV];
  inviscid_proj_flux<ND>(Vi, normal, fi);
  inviscid_proj_flux<ND>(Vj, normal, fj);
#pragma unroll
  for (int k = 0; k < NV; ++k) flux[k] = 0.5 * (fi[k] + fj[k]);

  if (IMPLICIT) {
    inviscid_proj_jac<ND>(p.gamma, &Vi[I::VEL], Ui[ND + 1] / Ui[0], normal, p.roe_kappa, jac_i);
    inviscid_proj_jac<ND>(p.gamma, &Vj[I::VEL], Uj[ND + 1] / Uj[0], normal, p.roe_kappa, jac_j);
  }

  double dU[NV];
#pragma unroll
  for (int k = 0; k < NV; ++k) dU[k] = Uj[k] - Ui[k];

  // D = P |Lambda| P^-1, accumulated one eigenvalue at a time: m_ij = sum_k P_ik lambda_k Pinv_kj in the reference's
  // order of k. Column k of P and row k of P^-1 are regenerated for every k (the optimisation barrier keeps the
  // compiler from merging the five evaluations back into two resident 5x5 matrices = 100 registers).
  double m[NV * NV];
#pragma unroll
  for (int q = 0; q < NV * NV; ++q) m[q] = 0.0;
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    double rho_k = rrho, q_k = rq, c_k = rc;
#ifdef __CUDA_ARCH__
    asm volatile("" : "+d"(rho_k), "+d"(q_k), "+d"(c_k));
#endif
    double P[NV * NV], Pi[NV * NV];
    p_matrix<ND>(p.gamma, rho_k, rv, q_k, c_k, un, P);
    p_matrix_inv<ND>(p.gamma, rho_k, rv, q_k, c_k, un, Pi);
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const double pl = P[i * NV + k] * lambda[k];
#pragma unroll
      for (int j = 0; j < NV; ++j) m[i * NV + j] += pl * Pi[k * NV + j];
    }
  }
  const double scale = (1 - p.roe_kappa) * area;  // * dissipation (== 1)
#pragma unroll
  for (int i = 0; i < NV; ++i) {
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const double dDdU = m[i * NV + j] * scale;
      flux[i] -= dDdU * dU[j];
      if (IMPLICIT) {
        jac_i[i * NV + j] += dDdU;
        jac_j[i * NV + j] -= dDdU;
      }
    }
  }
}

// scalarDissipationJacobian, common.hpp:533-549
template <int ND, class JT>
__device__ __forceinline__ void scalar_dissipation_jacobian(const double* V, double gamma, double dc, JT jac) {
  using I = PrimIdx<ND>;
  constexpr int NV = ND + 2;
#pragma unroll
  for (int k = 0; k < NV - 1; ++k) jac[k * NV + k] += dc;
  jac[(NV - 1) * NV + NV - 1] += dc * gamma;
  dc *= gamma - 1.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    jac[(NV - 1) * NV + d + 1] -= dc * V[I::VEL + d];
    jac[(NV - 1) * NV + 0] += dc * V[I::VEL + d] * V[I::VEL + d];
  }
}

// Viscous terms, CCompressibleViscousFluxBase::viscousTerms (viscous_fluxes.hpp:114-202) with
// CCompressibleViscousFlux (ideal gas, :253-298). Vi/Vj are the un-reconstructed primitives (all ND+9),
// gi/gj the primitive gradients [(ND+1)][ND] of (T, v). Subtracts the viscous flux from `flux` and
// updates jac_i/jac_j. `correct` = edge-correction of the mean gradient (finest grid).
// LEGACY = the boundary functor CAvgGrad_Flow (flow_diffusion.cpp:368-532): same flux, but its heat-flux Jacobian
// uses cp/R * (k/cp + mu_t/Pr_t) (SetHeatFluxJacobian :489-532), which differs from the SIMD form when the mean cp
// is not the gas cp (far-field state carries cp = k = 0, CEulerSolver.cpp:4990-4993).
// `gmean` (optional): the mean gradient 0.5*gi + 0.5*gj already formed by the caller (the staged edge kernel keeps it
// in shared memory); used instead of gi/gj when `have_mean`.
template <int ND, bool IMPLICIT, bool LEGACY = false, class JT = double*, class GM = const double*>
__device__ __forceinline__ void viscous_terms(const FlowParams& p, const double* Vi, const double* Vj,
                                              const double* gi, const double* gj, const double* vec_ij,
                                              double area, const double* un, bool correct,
                                              double* flux, JT jac_i, JT jac_j, bool have_mean = false,
                                              GM gmean = GM()) {
  using I = PrimIdx<ND>;
  constexpr int NV = ND + 2;
  constexpr int NG = ND + 1;
  constexpr double EPS = 1.0e-16;
  double avgV[I::NPRIM];
#pragma unroll
  for (int k = 0; k < I::NPRIM; ++k) avgV[k] = 0.5 * (Vi[k] + Vj[k]);

  double dist2 = dotN<ND>(vec_ij, vec_ij);
  const double mask = (dist2 < EPS * EPS) ? 1.0 : 0.0;
  dist2 += mask / (EPS * EPS);

  double g[NG * ND];
#pragma unroll
  for (int k = 0; k < NG * ND; ++k) {
    if (have_mean) {
      g[k] = gmean[k];
    } else {
      double a = gi[k];
      a *= 0.5;
      a += 0.5 * gj[k];
      g[k] = a;
    }
  }
  if (correct) {
#pragma unroll
    for (int v = 0; v < NG; ++v) {
      const double corr = (dotN<ND>(&g[v * ND], vec_ij) - Vj[v] + Vi[v]) / dist2;
#pragma unroll
      for (int d = 0; d < ND; ++d) g[v * ND + d] -= corr * vec_ij[d];
    }
  }
  // stress tensor (diffusion/common.hpp:71-90)
  const double visc = avgV[I::MU] + avgV[I::MUT];
  double div = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) div += g[(d + 1) * ND + d];
  const double pTerm = 2.0 / 3.0 * visc * div;
  double tau[ND * ND];
#pragma unroll
  for (int i = 0; i < ND; ++i) {
#pragma unroll
    for (int j = 0; j < ND; ++j) tau[i * ND + j] = visc * (g[(j + 1) * ND + i] + g[(i + 1) * ND + j]);
    tau[i * ND + i] -= pTerm;
  }
  const double cond = avgV[I::K] + avgV[I::CP] * avgV[I::MUT] / p.prandtl_turb;
  double q[ND];
#pragma unroll
  for (int d = 0; d < ND; ++d) q[d] = cond * g[d];

  // projected flux (viscousFlux, diffusion/common.hpp:208-226)
  double vf[NV];
  vf[0] = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) vf[d + 1] = dotN<ND>(&tau[d * ND], un);
  vf[ND + 1] = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    const double work = dotN<ND>(&tau[d * ND], &avgV[I::VEL]);
    vf[ND + 1] += un[d] * (q[d] + work);
  }
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    vf[k] *= area;
    flux[k] -= vf[k];
  }
  if (!IMPLICIT) return;

  const double dist = sqrt(dist2);
  // stressTensorJacobian (diffusion/common.hpp:183-203)
  const double xi = visc / (avgV[I::RHO] * dist);
  double dtau[ND * NV];
#pragma unroll
  for (int i = 0; i < ND; ++i) {
#pragma unroll
    for (int j = 0; j < ND; ++j) dtau[i * NV + j + 1] = (-1 / 3.0) * xi * un[i] * un[j];
    dtau[i * NV + i + 1] -= xi;
    dtau[i * NV + 0] = -dotN<ND>(&dtau[i * NV + 1], &avgV[I::VEL]);
    dtau[i * NV + ND + 1] = 0.0;
  }
  // energyJacobian (viscous_fluxes.hpp:277-298)
  const double vel2 = 0.5 * dotN<ND>(&avgV[I::VEL], &avgV[I::VEL]);
  const double phi = (p.gamma - 1) / avgV[I::RHO];
  const double RdTdrho = phi * vel2 - avgV[I::P] / (avgV[I::RHO] * avgV[I::RHO]);
  const double condOnRd = LEGACY ? (p.gamma / (p.gamma - 1)) * (avgV[I::K] / avgV[I::CP] + avgV[I::MUT] / p.prandtl_turb) / dist
                                 : cond / (p.R * dist);
  double contraction = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) contraction += dtau[d * NV + 0] * avgV[I::VEL + d];
  double dEdU[NV];
  dEdU[0] = area * (condOnRd * RdTdrho - contraction);
#pragma unroll
  for (int d = 0; d < ND; ++d) dEdU[d + 1] = area * (dtau[d * NV + 0] - condOnRd * phi * avgV[I::VEL + d]);
  dEdU[ND + 1] = area * condOnRd * phi;

#pragma unroll
  for (int d = 0; d < ND; ++d) {
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      jac_i[(d + 1) * NV + k] -= area * dtau[d * NV + k];
      jac_j[(d + 1) * NV + k] += area * dtau[d * NV + k];
    }
  }
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    jac_i[(ND + 1) * NV + k] += dEdU[k];
    jac_j[(ND + 1) * NV + k] -= dEdU[k];
  }
  const double proj = dotN<ND>(&vf[1], &avgV[I::VEL]);
  const double halfOnRho = 0.5 / avgV[I::RHO];
  jac_i[(ND + 1) * NV + 0] += halfOnRho * proj;
  jac_j[(ND + 1) * NV + 0] += halfOnRho * proj;
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    jac_i[(ND + 1) * NV + d + 1] -= halfOnRho * vf[d + 1];
    jac_j[(ND + 1) * NV + d + 1] -= halfOnRho * vf[d + 1];
  }
}

}  // namespace su2b200
