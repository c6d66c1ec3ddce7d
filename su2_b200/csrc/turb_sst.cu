// f4. Menter SST (V2003m, the reference's default: option_structure.hpp:1036-1100) on the device, two variables (k, omega).
//   CTurbSSTSolver ctor / Postprocessing / Source_Residual / BC_HeatFlux_Wall (CTurbSSTSolver.cpp:40-200, :206-246, :304-392,
//       :403-514), CTurbSSTVariable::SetBlendingFunc (CTurbSSTVariable.cpp:56-97)
//   CScalarSolver::Upwind_Residual (CScalarSolver.inl:129-343) with CUpwSca_TurbSST (turb_convection.hpp:83-125)
//   CScalarSolver::Viscous_Residual_NonCons (CScalarSolver.hpp:160-240) with CAvgGrad_Scalar / CAvgGrad_TurbSST
//       (scalar_diffusion.hpp:119-154, turb_diffusion.hpp:207-346: the cross diffusion lives in the diffusive flux)
//   CSourcePieceWise_TurbSST (turb_sources.hpp:631-930), production NONE, no sustaining terms
//   CScalarSolver::BC_Far_Field (CScalarSolver.inl:382-441), Prepare/CompleteImplicitIteration (:456-544, Conservative:
//       AddClippedSolution with the current density on both sides), CTurbSolver::ComputeUnderRelaxationFactorHelper
// Same atomics-free assembly as turb_sa.cu: per-edge fluxes and 2x2 diagonal contributions, ordered point sums.
#include "internal.hpp"
#include "device_math.cuh"
#include "kernels.hpp"

namespace su2b200 {

static inline int nblk(long n, int bs) { return (int)((n + bs - 1) / bs); }

struct SSTConst {
  double sigma_k1, sigma_k2, sigma_om1, sigma_om2, beta_1, beta_2, beta_star, a1, alfa_1, alfa_2, prod_lim;
};
__host__ __device__ inline SSTConst sst_constants() {
  SSTConst k;
  k.sigma_k1 = 0.85; k.sigma_k2 = 1.0; k.sigma_om1 = 0.5; k.sigma_om2 = 0.856;
  k.beta_1 = 0.075; k.beta_2 = 0.0828; k.beta_star = 0.09; k.a1 = 0.31;
  k.alfa_1 = 5.0 / 9.0; k.alfa_2 = 0.44; k.prod_lim = 10.0;  // V2003
  return k;
}
static constexpr int SST_EF = 18;  // per-edge doubles: F[2], Rij[2], Rji[2], a0, a1, Di[4] (first node), Dj[4] (second node), pad

// ---- edge kernel -------------------------------------------------------------------------------------------------------
template <int ND>
__global__ void __launch_bounds__(128)
k_sst_edge(int E, int reducer, const int* __restrict__ ei, const int* __restrict__ ej, const double* __restrict__ enormal,
           const double* __restrict__ coord, const double* __restrict__ V, const double* __restrict__ tk,
           const double* __restrict__ gtk, const double* __restrict__ F1, const int* __restrict__ eij,
           const int* __restrict__ eji, double* __restrict__ A, double* __restrict__ ef) {
  using I = PrimIdx<ND>;
  const SSTConst K = sst_constants();
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  const int i = ei[e], j = ej[e];
  double n[ND], ev[ND];
  double q = 0.0, d2 = 0.0, pv = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    n[d] = enormal[(size_t)d * E + e];
    ev[d] = coord[(size_t)j * ND + d] - coord[(size_t)i * ND + d];
    q += 0.5 * (V[(size_t)i * I::NPRIM + I::VEL + d] + V[(size_t)j * I::NPRIM + I::VEL + d]) * n[d];
    d2 += ev[d] * ev[d];
    pv += ev[d] * n[d];
  }
  const double rho_i = V[(size_t)i * I::NPRIM + I::RHO], rho_j = V[(size_t)j * I::NPRIM + I::RHO];
  const double mu_i = V[(size_t)i * I::NPRIM + I::MU], mu_j = V[(size_t)j * I::NPRIM + I::MU];
  const double mut_i = V[(size_t)i * I::NPRIM + I::MUT], mut_j = V[(size_t)j * I::NPRIM + I::MUT];
  const double si[2] = {tk[(size_t)2 * i], tk[(size_t)2 * i + 1]}, sj[2] = {tk[(size_t)2 * j], tk[(size_t)2 * j + 1]};
  // CUpwSca_TurbSST
  const double a0 = fmax(0.0, q), a1 = fmin(0.0, q);
  const double Fc[2] = {a0 * rho_i * si[0] + a1 * rho_j * sj[0], a0 * rho_i * si[1] + a1 * rho_j * sj[1]};
  // ComputeProjectedGradient with correction, both evaluations (see turb_sa.cu for the sign symmetry)
  const double EPS = 1.0e-16;
  const double pvij = pv / fmax(d2, EPS);
  double pc_ij[2], pc_ji[2];
#pragma unroll
  for (int v = 0; v < 2; ++v) {
    double projN = 0.0, edgeProj = 0.0;
#pragma unroll
    for (int d = 0; d < ND; ++d) {
      const double mg = 0.5 * (gtk[((size_t)i * 2 + v) * ND + d] + gtk[((size_t)j * 2 + v) * ND + d]);
      projN += mg * n[d];
      edgeProj += mg * ev[d];
    }
    pc_ij[v] = projN - (edgeProj - (sj[v] - si[v])) * pvij;
    pc_ji[v] = (-projN) - ((-edgeProj) - (si[v] - sj[v])) * pvij;
  }
  // CAvgGrad_TurbSST::FinishResidualCalc for the evaluation a -> b ("a" is the cell the flux belongs to)
  double R[2][2], Ja[2][4], Jb[2][4];  // [evaluation][...]: residual, Jacobian w.r.t. a and b (row-major 2x2)
  auto diffusion = [&](int w, double F1a, double F1b, double rho_a, double rho_b, double mu_a, double mu_b, double mut_a,
                       double mut_b, double om_a, double om_b, const double* pc) {
    const double sk_a = F1a * K.sigma_k1 + (1.0 - F1a) * K.sigma_k2, sk_b = F1b * K.sigma_k1 + (1.0 - F1b) * K.sigma_k2;
    const double so_a = F1a * K.sigma_om1 + (1.0 - F1a) * K.sigma_om2, so_b = F1b * K.sigma_om1 + (1.0 - F1b) * K.sigma_om2;
    const double dk_a = mu_a + sk_a * mut_a, dk_b = mu_b + sk_b * mut_b;
    const double do_a = mu_a + so_a * mut_a, do_b = mu_b + so_b * mut_b;
    const double diff_kine = 0.5 * (dk_a + dk_b);
    const double diff_omega_T1 = 0.5 * (do_a + do_b);
    const double lambda_a = 2 * (1 - F1a) * rho_a * so_a, lambda_b = 2 * (1 - F1b) * rho_b * so_b;
    const double lambda_ij = 0.5 * (lambda_a + lambda_b);
    const double w_ij = 0.5 * (om_a + om_b);
    const double T2 = lambda_ij;
    const double T3 = -om_a * lambda_ij / w_ij;
    R[w][0] = diff_kine * pc[0];
    R[w][1] = diff_omega_T1 * pc[1] + (T2 + T3) * pc[0];
    const double por_a = pvij / rho_a, por_b = pvij / rho_b;
    Ja[w][0] = -diff_kine * por_a; Ja[w][1] = 0.0; Ja[w][2] = (T2 + T3) * -por_a; Ja[w][3] = -diff_omega_T1 * por_a;
    Jb[w][0] = diff_kine * por_b;  Jb[w][1] = 0.0; Jb[w][2] = (T2 + T3) * por_b;  Jb[w][3] = diff_omega_T1 * por_b;
  };
  const double F1i = F1[i], F1j = F1[j];
  diffusion(0, F1i, F1j, rho_i, rho_j, mu_i, mu_j, mut_i, mut_j, si[1], sj[1], pc_ij);
  diffusion(1, F1j, F1i, rho_j, rho_i, mu_j, mu_i, mut_j, mut_i, sj[1], si[1], pc_ji);
  double* o = ef + (size_t)e;  // SoA: field f at ef[f * E + e]
  o[(size_t)0 * E] = Fc[0]; o[(size_t)1 * E] = Fc[1];
  o[(size_t)2 * E] = R[0][0]; o[(size_t)3 * E] = R[0][1];
  o[(size_t)4 * E] = R[1][0]; o[(size_t)5 * E] = R[1][1];
  double* Bij = A + (size_t)eij[e] * 4;
  double* Bji = A + (size_t)eji[e] * 4;
  if (reducer) {
    // averaged one-sided Jacobians in the off-diagonal blocks, diagonal = minus the column sum (CScalarSolver.hpp:196-226)
#pragma unroll
    for (int m = 0; m < 4; ++m) {
      const double conv_ij = (m == 0 || m == 3) ? a1 : 0.0, conv_ji = (m == 0 || m == 3) ? -a0 : 0.0;
      const double bij = conv_ij - 0.5 * Jb[0][m] + 0.5 * Ja[1][m];
      const double bji = conv_ji + 0.5 * Ja[0][m] - 0.5 * Jb[1][m];
      Bij[m] = bij;
      Bji[m] = bji;
      o[(size_t)(8 + m) * E] = bji;    // first node: D_i -= B_ji (column i)
      o[(size_t)(12 + m) * E] = bij;   // second node: D_j -= B_ij (column j)
    }
    o[(size_t)6 * E] = 0.0; o[(size_t)7 * E] = 0.0;
    return;
  }
#pragma unroll
  for (int m = 0; m < 4; ++m) {
    const double dg = (m == 0 || m == 3) ? 1.0 : 0.0;
    Bij[m] = a1 * dg - Jb[0][m];     // UpdateBlocks: B_ij = a1 I, then -= jacobian_j of the i -> j evaluation
    Bji[m] = -a0 * dg - Jb[1][m];    // B_ji = -a0 I, then -= jacobian_j of the j -> i evaluation (w.r.t. i)
    o[(size_t)(8 + m) * E] = Ja[0][m];   // first node:  D_i += a0 I; D_i -= this
    o[(size_t)(12 + m) * E] = Ja[1][m];  // second node: D_j -= a1 I; D_j -= this
  }
  o[(size_t)6 * E] = a0;
  o[(size_t)7 * E] = a1;
}

// ---- CSourcePieceWise_TurbSST::ComputeResidual for one point (V2003, production NONE, no sustaining terms) ---------------
__device__ __forceinline__ void sst_source(double rho, double mut, double k, double om, double strain, double dist,
                                           double F1, double volume, double* res, double* jac) {
  const SSTConst K = sst_constants();
  res[0] = res[1] = 0.0;
  jac[0] = jac[1] = jac[2] = jac[3] = 0.0;
  const double alfa_blended = F1 * K.alfa_1 + (1.0 - F1) * K.alfa_2;
  const double beta_blended = F1 * K.beta_1 + (1.0 - F1) * K.beta_2;
  if (!(dist > 1e-10)) return;
  const double zetaFMt = 0.0;
  const double P_Base = strain;
  const double prod_limit = K.prod_lim * K.beta_star * rho * om * k;
  const double P = mut * (P_Base * P_Base);
  const double pk = fmax(0.0, fmin(P, prod_limit));
  const double pw = (alfa_blended * rho / mut) * pk;
  const double dk = K.beta_star * rho * om * k * (1.0 + zetaFMt);
  const double dw = beta_blended * rho * om * om * (1.0 - 0.09 / beta_blended * zetaFMt);
  res[0] += pk * volume;
  res[1] += pw * volume;
  res[0] -= dk * volume;
  res[1] -= dw * volume;
  jac[0] = -K.beta_star * om * volume * (1.0 + zetaFMt);
  jac[1] = -K.beta_star * k * volume * (1.0 + zetaFMt);
  jac[2] = 0.0;
  jac[3] = -2.0 * beta_blended * om * volume * (1.0 - 0.09 / beta_blended * zetaFMt);
}

// ---- point kernel: ordered sum of the edge contributions + source term ------------------------------------------------
template <int ND>
__global__ void k_sst_point(int P, int Pd, int E, const int* __restrict__ nbr_ptr, const int* __restrict__ gat_edge,
                            const int* __restrict__ dia_ptr, const double* __restrict__ ef, const double* __restrict__ V,
                            const double* __restrict__ tk, const double* __restrict__ strain,
                            const double* __restrict__ wall_dist, const double* __restrict__ F1,
                            const double* __restrict__ volume, double* __restrict__ A, double* __restrict__ res) {
  using I = PrimIdx<ND>;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  double r[2] = {0.0, 0.0}, D[4] = {0.0, 0.0, 0.0, 0.0};
  for (int k = nbr_ptr[i]; k < nbr_ptr[i + 1]; ++k) {
    const int se = gat_edge[k];
    const bool first = se >= 0;
    const size_t e = (size_t)(first ? se : ~se);
    if (first) {
      const double a0 = ef[(size_t)6 * E + e];
      r[0] += ef[(size_t)0 * E + e]; r[1] += ef[(size_t)1 * E + e];
      D[0] += a0; D[3] += a0;
      r[0] -= ef[(size_t)2 * E + e]; r[1] -= ef[(size_t)3 * E + e];
#pragma unroll
      for (int m = 0; m < 4; ++m) D[m] -= ef[(size_t)(8 + m) * E + e];
    } else {
      const double a1 = ef[(size_t)7 * E + e];
      r[0] -= ef[(size_t)0 * E + e]; r[1] -= ef[(size_t)1 * E + e];
      D[0] -= a1; D[3] -= a1;
      r[0] -= ef[(size_t)4 * E + e]; r[1] -= ef[(size_t)5 * E + e];
#pragma unroll
      for (int m = 0; m < 4; ++m) D[m] -= ef[(size_t)(12 + m) * E + e];
    }
  }
  if (i < Pd) {
    double s[2], js[4];
    sst_source(V[(size_t)i * I::NPRIM + I::RHO], V[(size_t)i * I::NPRIM + I::MUT], tk[(size_t)2 * i], tk[(size_t)2 * i + 1],
               strain[i], wall_dist[i], F1[i], volume[i], s, js);
    r[0] -= s[0]; r[1] -= s[1];
#pragma unroll
    for (int m = 0; m < 4; ++m) D[m] -= js[m];
  }
  res[(size_t)2 * i] = r[0];
  res[(size_t)2 * i + 1] = r[1];
  double* Dg = A + (size_t)dia_ptr[i] * 4;
#pragma unroll
  for (int m = 0; m < 4; ++m) Dg[m] = D[m];
}

// ---- boundaries ----------------------------------------------------------------------------------------------------------
// CScalarSolver::BC_Far_Field with CUpwSca_TurbSST: upwind flux between the node and (V_infty of the flow BC, Solution_Inf)
template <int ND>
__global__ void k_sst_bc_far(int nV, int Pd, double k_inf, double om_inf, const int* __restrict__ vnode,
                             const double* __restrict__ vnormal, const double* __restrict__ V,
                             const double* __restrict__ charac, const double* __restrict__ tk,
                             const int* __restrict__ dia_ptr, double* __restrict__ A, double* __restrict__ res) {
  using I = PrimIdx<ND>;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nV) return;
  const int i = vnode[t];
  if (i >= Pd) return;
  double q = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d)
    q += 0.5 * (V[(size_t)i * I::NPRIM + I::VEL + d] + charac[(size_t)t * I::NPRIM + I::VEL + d]) * (-vnormal[(size_t)t * ND + d]);
  const double a0 = fmax(0.0, q), a1 = fmin(0.0, q);
  const double rho_i = V[(size_t)i * I::NPRIM + I::RHO], rho_inf = charac[(size_t)t * I::NPRIM + I::RHO];
  res[(size_t)2 * i] += a0 * rho_i * tk[(size_t)2 * i] + a1 * rho_inf * k_inf;
  res[(size_t)2 * i + 1] += a0 * rho_i * tk[(size_t)2 * i + 1] + a1 * rho_inf * om_inf;
  double* Dg = A + (size_t)dia_ptr[i] * 4;
  Dg[0] += a0;
  Dg[3] += a0;
}

// CGeometry::ComputeWallDistance -> NearestNeighborDistance (CGeometry.cpp:4087-4113): the smallest wall distance among the
// neighbours that are not on a viscous wall; 2 V / |n| if every neighbour is on one
__global__ void k_mark_wall(int nV, const int* __restrict__ vnode, int* __restrict__ is_wall) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < nV) is_wall[vnode[t]] = 1;
}
template <int ND>
__global__ void k_wall_neighbor_distance(int nV, const int* __restrict__ vnode, const double* __restrict__ vnormal,
                                         const int* __restrict__ row_ptr, const int* __restrict__ col_ind,
                                         const int* __restrict__ is_wall, const double* __restrict__ wall_dist,
                                         const double* __restrict__ volume, double* __restrict__ out) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nV) return;
  const int i = vnode[t];
  const double big = 1.7976931348623157e308;
  double distance = big;
  for (int k = row_ptr[i]; k < row_ptr[i + 1]; ++k) {
    const int j = col_ind[k];
    if (j == i) continue;
    const double dist = is_wall[j] ? 0.0 : wall_dist[j];
    if (dist > 1.0e-16) distance = fmin(distance, dist);
  }
  if (!(distance > 0 && distance < big)) {
    double n2 = 0.0;
#pragma unroll
    for (int d = 0; d < ND; ++d) n2 += vnormal[(size_t)t * ND + d] * vnormal[(size_t)t * ND + d];
    distance = 2 * volume[i] / sqrt(n2);
  }
  out[t] = distance;
}

// CTurbSSTSolver::BC_HeatFlux_Wall (smooth wall): k = 0, omega = 60 mu / (rho beta_1 d^2) imposed strongly
template <int ND>
__global__ void k_sst_bc_wall(int nV, int Pd, const int* __restrict__ vnode, const double* __restrict__ nn_dist,
                              const double* __restrict__ V, const int* __restrict__ row_ptr,
                              const int* __restrict__ dia_ptr, double* __restrict__ tk, double* __restrict__ tk_old,
                              double* __restrict__ A, double* __restrict__ res) {
  using I = PrimIdx<ND>;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nV) return;
  const int i = vnode[t];
  if (i >= Pd) return;
  const double rho = V[(size_t)i * I::NPRIM + I::RHO], mu = V[(size_t)i * I::NPRIM + I::MU];
  const double d = nn_dist[t];
  const double om = 60.0 * mu / (rho * 0.075 * (d * d));  // pow(wall_dist, 2)
  tk_old[(size_t)2 * i] = 0.0; tk_old[(size_t)2 * i + 1] = om;
  tk[(size_t)2 * i] = 0.0; tk[(size_t)2 * i + 1] = om;
  res[(size_t)2 * i] = 0.0; res[(size_t)2 * i + 1] = 0.0;
  // DeleteValsRowi(iPoint*nVar), (iPoint*nVar+1) (CSysMatrix.cpp:604-614): both rows of every block zero, 1 on the diagonal
  for (int k = row_ptr[i]; k < row_ptr[i + 1]; ++k)
    for (int m = 0; m < 4; ++m) A[(size_t)k * 4 + m] = 0.0;
  A[(size_t)dia_ptr[i] * 4] = 1.0;
  A[(size_t)dia_ptr[i] * 4 + 3] = 1.0;
}

// ComputeUnderRelaxationFactorHelper (CTurbSolver.cpp:250-282) + AddClippedSolution with Conservative = true
// (CScalarSolver.inl:502-517, CVariable.hpp:468-474): val = (old * rho + ur * dU) / rho, clipped to [lower, upper]
template <int ND>
__global__ void k_sst_complete(int Pd, double allowable, double lo0, double lo1, double up0, double up1,
                               const double* __restrict__ V, const double* __restrict__ sol,
                               const double* __restrict__ tk_old, double* __restrict__ tk,
                               double* __restrict__ under_relax) {
  using I = PrimIdx<ND>;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Pd) return;
  double ur = 1.0;
#pragma unroll
  for (int v = 0; v < 2; ++v) {
    const double ratio = fabs(sol[(size_t)2 * i + v]) / (fabs(tk[(size_t)2 * i + v]) + 1.0e-16);
    if (ratio > allowable) ur = fmin(allowable / ratio, ur);
  }
  if (ur < 1e-10) ur = 0.0;
  under_relax[i] = ur;
  const double rho = V[(size_t)i * I::NPRIM + I::RHO];
  const double lo[2] = {lo0, lo1}, up[2] = {up0, up1};
#pragma unroll
  for (int v = 0; v < 2; ++v) {
    const double val = (tk_old[(size_t)2 * i + v] * rho + ur * sol[(size_t)2 * i + v]) / rho;
    tk[(size_t)2 * i + v] = fmin(fmax(val, lo[v]), up[v]);
  }
}

// CTurbSSTSolver::Postprocessing (after the gradient): CTurbSSTVariable::SetBlendingFunc, then muT with the stress limiter
template <int ND>
__global__ void k_sst_blend_mut(int P, const double* __restrict__ V, const double* __restrict__ tk,
                                const double* __restrict__ gtk, const double* __restrict__ strain,
                                const double* __restrict__ wall_dist, double* __restrict__ F1o, double* __restrict__ F2o,
                                double* __restrict__ CDo, double* __restrict__ mut) {
  using I = PrimIdx<ND>;
  const SSTConst K = sst_constants();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const double rho = V[(size_t)i * I::NPRIM + I::RHO], mu = V[(size_t)i * I::NPRIM + I::MU];
  const double dist = wall_dist[i];
  const double k = tk[(size_t)2 * i], om = tk[(size_t)2 * i + 1];
  const double EPS2 = 1.0e-16 * 1.0e-16;
  double cd = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) cd += gtk[((size_t)i * 2 + 0) * ND + d] * gtk[((size_t)i * 2 + 1) * ND + d];
  cd *= 2.0 * rho * K.sigma_om2 / om;
  cd = fmax(cd, 1.0e-10);  // pow(10.0, -prod_lim_const)
  const double arg2A = sqrt(k) / (K.beta_star * om * dist + EPS2);
  const double arg2B = 500.0 * mu / (rho * dist * dist * om + EPS2);
  double arg2 = fmax(arg2A, arg2B);
  const double arg1 = fmin(arg2, 4.0 * rho * K.sigma_om2 * k / (cd * dist * dist + EPS2));
  const double F1 = tanh(pow(arg1, 4.0));
  arg2 = fmax(2.0 * arg2A, arg2B);
  const double F2 = tanh(pow(arg2, 2.0));
  F1o[i] = F1;
  F2o[i] = F2;
  CDo[i] = cd;
  const double S = fmax(strain[i], 1e-12);
  mut[i] = fmax(0.0, rho * K.a1 * k / fmax(K.a1 * om, S * F2));
}

// ---- host wrappers (dispatched from turb_sa.cu's entry points when KIND_TURB_MODEL= SST) -------------------------------
void sst_preprocessing(su2b200_ctx* c) {
  auto& s = c->fs;
  scalar2_gradient(c, c->cfg.grad_method, s.nu.p, s.nu_grad.p);
  if (c->nranks > 1) halo_exchange(c, s.nu_grad.p, 2 * c->geo.nDim);
}

static void sst_wall_distances(su2b200_ctx* c) {
  auto& g = c->geo;
  for (auto& m : g.markers) {
    if (m.kind != SU2B200_HEAT_FLUX || m.nVertex == 0 || m.nn_dist.n == (size_t)m.nVertex) continue;
    dvec<int> is_wall;
    is_wall.alloc(g.P);
    SU2B_CUDA(cudaMemsetAsync(is_wall.p, 0, (size_t)g.P * sizeof(int), c->stream));
    for (auto& w : g.markers)
      if (w.kind == SU2B200_HEAT_FLUX && w.nVertex > 0)
        SU2B_LAUNCH(c, k_mark_wall, nblk(w.nVertex, 128), 128, w.nVertex, w.node.p, is_wall.p);
    m.nn_dist.alloc(m.nVertex);
    if (g.nDim == 2)
      SU2B_LAUNCH(c, k_wall_neighbor_distance<2>, nblk(m.nVertex, 64), 64, m.nVertex, m.node.p, m.normal.p, g.row_ptr.p,
                  g.col_ind.p, is_wall.p, g.wall_dist.p, g.volume.p, m.nn_dist.p);
    else
      SU2B_LAUNCH(c, k_wall_neighbor_distance<3>, nblk(m.nVertex, 64), 64, m.nVertex, m.node.p, m.normal.p, g.row_ptr.p,
                  g.col_ind.p, is_wall.p, g.wall_dist.p, g.volume.p, m.nn_dist.p);
    SU2B_CUDA(cudaStreamSynchronize(c->stream));  // is_wall goes out of scope
  }
}

void sst_space_integration(su2b200_ctx* c) {
  ++c->sys[1].A_gen;
  auto& g = c->geo;
  auto& s = c->fs;
  auto& ls = c->sys[1];
  SU2B_REQUIRE(g.wall_dist.n == (size_t)g.P, "RANS needs CPoint::Wall_Distance (geometry_desc.wall_distance)");
  const int reducer = c->cfg.turb_reducer >= 0
                          ? c->cfg.turb_reducer
                          : ((double)g.E / (512.0 * (double)((g.E + 511) / 512)) < 0.875 ? 1 : 0);
  if (s.turb_edge_flux.n < (size_t)SST_EF * g.E) s.turb_edge_flux.alloc((size_t)SST_EF * g.E);
  sst_wall_distances(c);
  {
    ProfScope prof_(c, PROF_TURB_EDGE);
    if (g.nDim == 2)
      SU2B_LAUNCH(c, k_sst_edge<2>, nblk(g.E, 128), 128, g.E, reducer, g.edge_i.p, g.edge_j.p, g.edge_normal.p, g.coord.p, s.V.p,
                  s.nu.p, s.nu_grad.p, s.sst_F1.p, g.edge_ij.p, g.edge_ji.p, ls.A.p, s.turb_edge_flux.p);
    else
      SU2B_LAUNCH(c, k_sst_edge<3>, nblk(g.E, 128), 128, g.E, reducer, g.edge_i.p, g.edge_j.p, g.edge_normal.p, g.coord.p, s.V.p,
                  s.nu.p, s.nu_grad.p, s.sst_F1.p, g.edge_ij.p, g.edge_ji.p, ls.A.p, s.turb_edge_flux.p);
  }
  {
    ProfScope prof_(c, PROF_TURB_SOURCE);
    if (g.nDim == 2)
      SU2B_LAUNCH(c, k_sst_point<2>, nblk(g.P, 128), 128, g.P, g.Pd, g.E, g.nbr_ptr.p, g.gat_edge.p, g.dia_ptr.p,
                  s.turb_edge_flux.p, s.V.p, s.nu.p, s.strain_mag.p, g.wall_dist.p, s.sst_F1.p, g.volume.p, ls.A.p, s.turb_res.p);
    else
      SU2B_LAUNCH(c, k_sst_point<3>, nblk(g.P, 128), 128, g.P, g.Pd, g.E, g.nbr_ptr.p, g.gat_edge.p, g.dia_ptr.p,
                  s.turb_edge_flux.p, s.V.p, s.nu.p, s.strain_mag.p, g.wall_dist.p, s.sst_F1.p, g.volume.p, ls.A.p, s.turb_res.p);
  }
  ProfScope prof_(c, PROF_BC);
  for (auto& m : g.markers) {
    if (m.kind != SU2B200_FAR_FIELD || m.nVertex == 0) continue;
    if (g.nDim == 2)
      SU2B_LAUNCH(c, k_sst_bc_far<2>, nblk(m.nVertex, 64), 64, m.nVertex, g.Pd, c->cfg.kine_inf, c->cfg.omega_inf, m.node.p,
                  m.normal.p, s.V.p, m.charac.p, s.nu.p, g.dia_ptr.p, ls.A.p, s.turb_res.p);
    else
      SU2B_LAUNCH(c, k_sst_bc_far<3>, nblk(m.nVertex, 64), 64, m.nVertex, g.Pd, c->cfg.kine_inf, c->cfg.omega_inf, m.node.p,
                  m.normal.p, s.V.p, m.charac.p, s.nu.p, g.dia_ptr.p, ls.A.p, s.turb_res.p);
  }
  for (auto& m : g.markers) {
    if (m.kind != SU2B200_HEAT_FLUX || m.nVertex == 0) continue;
    if (g.nDim == 2)
      SU2B_LAUNCH(c, k_sst_bc_wall<2>, nblk(m.nVertex, 64), 64, m.nVertex, g.Pd, m.node.p, m.nn_dist.p, s.V.p, g.row_ptr.p,
                  g.dia_ptr.p, s.nu.p, s.nu_old.p, ls.A.p, s.turb_res.p);
    else
      SU2B_LAUNCH(c, k_sst_bc_wall<3>, nblk(m.nVertex, 64), 64, m.nVertex, g.Pd, m.node.p, m.nn_dist.p, s.V.p, g.row_ptr.p,
                  g.dia_ptr.p, s.nu.p, s.nu_old.p, ls.A.p, s.turb_res.p);
  }
}

void sst_complete_implicit(su2b200_ctx* c) {
  ProfScope prof_(c, PROF_UPDATE);
  auto& g = c->geo;
  auto& s = c->fs;
  const auto& f = c->cfg;
  if (g.nDim == 2)
    SU2B_LAUNCH(c, k_sst_complete<2>, nblk(g.Pd, 256), 256, g.Pd, f.max_update_sst, f.sst_lower[0], f.sst_lower[1],
                f.sst_upper[0], f.sst_upper[1], s.V.p, s.turb_sol.p, s.nu_old.p, s.nu.p, s.turb_under_relax.p);
  else
    SU2B_LAUNCH(c, k_sst_complete<3>, nblk(g.Pd, 256), 256, g.Pd, f.max_update_sst, f.sst_lower[0], f.sst_lower[1],
                f.sst_upper[0], f.sst_upper[1], s.V.p, s.turb_sol.p, s.nu_old.p, s.nu.p, s.turb_under_relax.p);
  if (c->nranks > 1) halo_exchange(c, s.nu.p, 2);
}

void sst_postprocessing(su2b200_ctx* c) {
  auto& g = c->geo;
  auto& s = c->fs;
  sst_preprocessing(c);  // CTurbSSTSolver::Postprocessing starts with the gradient of the updated (k, omega)
  ProfScope prof_(c, PROF_TURB_OTHER);
  if (g.nDim == 2)
    SU2B_LAUNCH(c, k_sst_blend_mut<2>, nblk(g.P, 256), 256, g.P, s.V.p, s.nu.p, s.nu_grad.p, s.strain_mag.p, g.wall_dist.p,
                s.sst_F1.p, s.sst_F2.p, s.sst_CDkw.p, s.mu_t.p);
  else
    SU2B_LAUNCH(c, k_sst_blend_mut<3>, nblk(g.P, 256), 256, g.P, s.V.p, s.nu.p, s.nu_grad.p, s.strain_mag.p, g.wall_dist.p,
                s.sst_F1.p, s.sst_F2.p, s.sst_CDkw.p, s.mu_t.p);
}

}  // namespace su2b200
