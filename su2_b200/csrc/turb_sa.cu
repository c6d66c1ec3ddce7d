// a25-a27. Spalart-Allmaras transport on the device (standard SA / SA-neg, no ft2, MUSCL_TURB= NO).
//   CTurbSASolver::Preprocessing -> CScalarSolver::CommonPreprocessing (CScalarSolver.inl:88-127)
//   CScalarSolver::Upwind_Residual (CScalarSolver.inl:129-343) with CUpwScalar/CUpwSca_TurbSA
//       (scalar_convection.hpp:106-149, turb_convection.hpp:61-65)
//   CScalarSolver::Viscous_Residual_NonCons (CScalarSolver.hpp:160-240) with CAvgGrad_Scalar/CAvgGrad_TurbSA
//       (scalar_diffusion.hpp:119-154, turb_diffusion.hpp:68-112, _Neg :158-200), CNumerics::ComputeProjectedGradient
//   CTurbSASolver::Source_Residual (CTurbSASolver.cpp:310-438) with CSourceBase_TurbSA (turb_sources.hpp:138-262)
//   CTurbSASolver::BC_HeatFlux_Wall (:444-517), CScalarSolver::BC_Far_Field (CScalarSolver.inl:382-441)
//   CScalarSolver::SetTime_Step (:443-454), Prepare/CompleteImplicitIteration (:456-544),
//   CTurbSolver::ComputeUnderRelaxationFactorHelper (CTurbSolver.cpp:250-282), CTurbSASolver::Postprocessing (:210-247)
//
// Assembly without atomics: the edge kernel stores, per edge, the convective flux F, the two one-sided diffusive
// fluxes R_ij / R_ji, and the four diagonal contributions; off-diagonal 1x1 blocks are written in place. A point
// kernel then walks the incident edges in ascending edge index and applies the same operations, in the same order,
// as the reference's serial edge loop (res_i += F; res_i -= R_ij; ... ), adds the source term and its Jacobian.
#include "internal.hpp"
#include "device_math.cuh"
#include "kernels.hpp"

namespace su2b200 {

static inline int nblk(long n, int bs) { return (int)((n + bs - 1) / bs); }

struct SAConst {
  double cv1_3, k2, cb1, cw2, ct3, ct4, cw3_6, sigma, cb2, cw1;
};
__host__ __device__ inline SAConst sa_constants() {
  SAConst k;
  k.cv1_3 = 7.1 * 7.1 * 7.1;  // pow(7.1, 3) evaluates to the same double
  k.k2 = 0.41 * 0.41;
  k.cb1 = 0.1355;
  k.cw2 = 0.3;
  k.ct3 = 1.2;
  k.ct4 = 0.5;
  k.cw3_6 = 64.0;
  k.sigma = 2.0 / 3.0;
  k.cb2 = 0.622;
  k.cw1 = k.cb1 / k.k2 + (1 + k.cb2) / k.sigma;
  return k;
}

// ---- edge kernel: convection (first-order upwind) + non-conservative diffusion evaluated from both sides ---------
template <int ND>
__global__ void __launch_bounds__(128)
k_turb_edge(int E, int neg, int reducer, const int* __restrict__ ei, const int* __restrict__ ej, const double* __restrict__ enormal,
            const double* __restrict__ coord, const double* __restrict__ V, const double* __restrict__ nu,
            const double* __restrict__ gnu, const int* __restrict__ eij, const int* __restrict__ eji,
            double* __restrict__ A, double* __restrict__ ef /* SoA [7][E]: F, Rij, Rji, a0, Jij_i, a1, Jji_i */) {
  using I = PrimIdx<ND>;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= E) return;
  const int i = ei[e], j = ej[e];
  const double sigma = 2.0 / 3.0, cb2 = 0.622;
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
  const double nui = nu[i], nuj = nu[j];
  // CUpwSca_TurbSA
  const double a0 = fmax(0.0, q), a1 = fmin(0.0, q);
  const double F = a0 * nui + a1 * nuj;
  // ComputeProjectedGradient (correct = true on the finest grid). The reverse evaluation (j -> i, -normal) has the
  // same proj_vector_ij and the negated projected gradient, up to the sign symmetry of every operation involved.
  const double EPS = 1.0e-16;
  const double proj_vector_ij = pv / fmax(d2, EPS);
  double projN = 0.0, edgeProj = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    const double mg = 0.5 * (gnu[(size_t)i * ND + d] + gnu[(size_t)j * ND + d]);
    projN += mg * n[d];
    edgeProj += mg * ev[d];
  }
  const double projC_ij = projN - (edgeProj - (nuj - nui)) * proj_vector_ij;
  // j -> i: normal and edge vector flip sign, var_j - var_i flips
  const double projC_ji = (-projN) - ((-edgeProj) - (nui - nuj)) * proj_vector_ij;
  const double lam_i = V[(size_t)i * I::NPRIM + I::MU] / V[(size_t)i * I::NPRIM + I::RHO];
  const double lam_j = V[(size_t)j * I::NPRIM + I::MU] / V[(size_t)j * I::NPRIM + I::RHO];
  double Rij, Rji, Jij_i, Jij_j, Jji_i, Jji_j;  // Jxy_i: derivative w.r.t. the FIRST argument of that evaluation
  auto diffusion = [&](double la, double lb, double na, double nb, double projC, double& R, double& Ja, double& Jb) {
    if (neg) {
      // CAvgGrad_TurbSA_Neg::FinishResidualCalc (turb_diffusion.hpp:158-200)
      const double cn1 = 16.0;
      const double nu_ij = 0.5 * (la + lb), nu_tilde_ij = 0.5 * (na + nb);
      const double zeta = ((1 + cb2) * nu_tilde_ij - cb2 * na) / nu_ij;
      double fn = 1.0;
      if (zeta < 0.0) fn = (cn1 + zeta * zeta * zeta) / (cn1 - zeta * zeta * zeta);
      const double dcn = (nu_ij + (1 + cb2) * nu_tilde_ij * fn) - cb2 * na * fn;
      R = dcn * projC / sigma;
      Ja = dcn * (-proj_vector_ij) / sigma;
      Jb = dcn * proj_vector_ij / sigma;
      return;
    }
    const double nu_e = 0.5 * (la + lb + (1 + cb2) * (na + nb));
    const double dc = nu_e - cb2 * na;
    R = dc * projC / sigma;
    Ja = -dc * proj_vector_ij / sigma;
    Jb = dc * proj_vector_ij / sigma;
  };
  diffusion(lam_i, lam_j, nui, nuj, projC_ij, Rij, Jij_i, Jij_j);
  diffusion(lam_j, lam_i, nuj, nui, projC_ji, Rji, Jji_i /* w.r.t. j */, Jji_j /* w.r.t. i */);
  ef[(size_t)0 * E + e] = F;
  ef[(size_t)1 * E + e] = Rij;
  ef[(size_t)2 * E + e] = Rji;
  if (reducer) {
    // ReducerStrategy (CScalarSolver.hpp:196-203, :219-226): the Jacobians of the two one-sided evaluations are averaged
    // into the off-diagonal blocks, the diagonal is minus the column sum (SetDiagonalAsColumnSum, CScalarSolver.inl:320):
    // point i receives -B_ji, point j receives -B_ij (k_turb_point adds ef[3] - ef[4] for the first node of an edge and
    // subtracts ef[5] + ef[6] for the second)
    const double Bij = a1 - 0.5 * Jij_j + 0.5 * Jji_i;
    const double Bji = -a0 + 0.5 * Jij_i - 0.5 * Jji_j;
    A[eij[e]] = Bij;
    A[eji[e]] = Bji;
    ef[(size_t)3 * E + e] = -Bji;
    ef[(size_t)4 * E + e] = 0.0;
    ef[(size_t)5 * E + e] = Bij;
    ef[(size_t)6 * E + e] = 0.0;
    return;
  }
  // off-diagonals: UpdateBlocks overwrites (B_ij = a1, B_ji = -a0), then the diffusion terms are subtracted
  A[eij[e]] = a1 - Jij_j;
  A[eji[e]] = -a0 - Jji_j;
  ef[(size_t)3 * E + e] = a0;
  ef[(size_t)4 * E + e] = Jij_i;
  ef[(size_t)5 * E + e] = a1;
  ef[(size_t)6 * E + e] = Jji_i;
}

// ---- CSourceBase_TurbSA<Bsl / Neg, ft2::Zero>::ComputeResidual for one point -----------------------------------------
__device__ __forceinline__ void sa_source(int neg, double density, double mu, double nue, const double* vort, double dist,
                                          double volume, double& residual, double& jacobian) {
  const SAConst k = sa_constants();
  residual = 0.0;
  jacobian = 0.0;
  if (!(dist > 1e-10)) return;
  const double dist2 = dist * dist;
  const double nu = mu / density;
  const double inv_k2_d2 = 1.0 / (k.k2 * dist2);
  const double Ji = nue / nu;  // roughness = 0
  const double Ji_2 = Ji * Ji, Ji_3 = Ji_2 * Ji;
  const double fv1 = Ji_3 / (Ji_3 + k.cv1_3);
  const double den = Ji_3 + k.cv1_3;
  const double d_fv1 = 3 * Ji_2 * k.cv1_3 / (nu * (den * den));
  const double fv2 = 1 - nue / (nu + nue * fv1);
  const double t1 = 1 + Ji * fv1;
  const double d_fv2 = -(1 / nu - Ji_2 * d_fv1) / (t1 * t1);
  const double Omega = sqrt(vort[0] * vort[0] + vort[1] * vort[1] + vort[2] * vort[2]);
  double Shat, d_Shat;
  if (neg && !(nue > 0.0)) {
    Shat = Omega;
    d_Shat = 0.0;
  } else {
    const double Sbar = nue * fv2 * inv_k2_d2;
    const double c2 = 0.7, c3 = 0.9;
    const double d_Sbar = (fv2 + nue * d_fv2) * inv_k2_d2;
    if (Sbar >= -c2 * Omega) {
      Shat = Omega + Sbar;
      d_Shat = d_Sbar;
    } else {
      const double Num = Omega * (c2 * c2 * Omega + c3 * Sbar);
      const double Den = (c3 - 2 * c2) * Omega - Sbar;
      Shat = Omega + Num / Den;
      d_Shat = d_Sbar * (c3 * Omega + Num / Den) / Den;
    }
    if (Shat <= 1e-10) {
      Shat = 1e-10;
      d_Shat = 0.0;
    }
  }
  const double inv_Shat = 1.0 / Shat;
  const double Prod = Shat;
  const double ft2 = 0.0, d_ft2 = 0.0;
  double r = fmin(nue * inv_Shat * inv_k2_d2, 10.0);
  double d_r = (Shat - nue * d_Shat) * (inv_Shat * inv_Shat) * inv_k2_d2;
  if (r >= 10.0) d_r = 0.0;
  const double r2 = r * r, r5 = r2 * r2 * r, r6 = r5 * r;
  const double g = r + k.cw2 * (r6 - r);
  const double g2 = g * g, g_6 = g2 * g2 * g2;
  const double glim = pow((1 + k.cw3_6) / (g_6 + k.cw3_6), 1.0 / 6.0);
  const double fw = g * glim;
  const double d_g = d_r * (1 + k.cw2 * (6 * r5 - 1));
  const double d_fw = d_g * glim * (1 - g_6 / (g_6 + k.cw3_6));
  double Production, Destruction;
  if (neg && !(nue > 0.0)) {
    const double dP_dnu = k.cb1 * (1.0 - k.ct3) * Prod;
    Production = dP_dnu * nue;
    jacobian += dP_dnu;
    const double dD_dnu = -k.cw1 * nue / dist2;
    Destruction = dD_dnu * nue;
    jacobian -= 2 * dD_dnu;
  } else {
    const double factor = k.cb1;
    Production = factor * (1.0 - ft2) * Prod * nue;
    jacobian += factor * (-Prod * nue * d_ft2 + (1.0 - ft2) * (nue * d_Shat + Prod));
    const double cb1_k2 = k.cb1 / k.k2;
    const double f2 = k.cw1 * fw - cb1_k2 * ft2;
    Destruction = f2 * (nue * nue) / dist2;
    jacobian -= ((k.cw1 * d_fw - cb1_k2 * d_ft2) * (nue * nue) + f2 * 2 * nue) / dist2;
  }
  residual = (Production - Destruction) * volume;
  jacobian *= volume;
}

// ---- point kernel: ordered sum of the edge contributions + source term ---------------------------------------------
template <int ND>
__global__ void k_turb_point(int P, int Pd, int E, int neg, const int* __restrict__ nbr_ptr,
                             const int* __restrict__ gat_edge, const int* __restrict__ dia_ptr,
                             const double* __restrict__ ef, const double* __restrict__ V, const double* __restrict__ nu,
                             const double* __restrict__ vort, const double* __restrict__ wall_dist,
                             const double* __restrict__ volume, double* __restrict__ A, double* __restrict__ res) {
  using I = PrimIdx<ND>;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  double r = 0.0, D = 0.0;
  for (int k = nbr_ptr[i]; k < nbr_ptr[i + 1]; ++k) {
    const int se = gat_edge[k];
    const bool first = se >= 0;
    const int e = first ? se : ~se;
    if (first) {
      r += ef[(size_t)0 * E + e];
      D += ef[(size_t)3 * E + e];
      r -= ef[(size_t)1 * E + e];
      D -= ef[(size_t)4 * E + e];
    } else {
      r -= ef[(size_t)0 * E + e];
      D -= ef[(size_t)5 * E + e];
      r -= ef[(size_t)2 * E + e];
      D -= ef[(size_t)6 * E + e];
    }
  }
  if (i < Pd) {
    double s, js;
    sa_source(neg, V[(size_t)i * I::NPRIM + I::RHO], V[(size_t)i * I::NPRIM + I::MU], nu[i], vort + (size_t)i * 3,
              wall_dist[i], volume[i], s, js);
    r -= s;
    D -= js;
  }
  res[i] = r;
  A[dia_ptr[i]] = D;
}

// ---- boundaries ------------------------------------------------------------------------------------------------
// CScalarSolver::BC_Far_Field: upwind flux between the node and (V_infty of the flow far-field BC, nu_tilde_inf)
template <int ND>
__global__ void k_turb_bc_far(int nV, int Pd, double nu_inf, const int* __restrict__ vnode,
                              const double* __restrict__ vnormal, const double* __restrict__ V,
                              const double* __restrict__ charac, const double* __restrict__ nu,
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
  res[i] += a0 * nu[i] + a1 * nu_inf;
  A[dia_ptr[i]] += a0;
}

// CTurbSASolver::BC_HeatFlux_Wall (smooth wall): nu_tilde = 0 imposed strongly
__global__ void k_turb_bc_wall(int nV, int Pd, const int* __restrict__ vnode, const int* __restrict__ row_ptr,
                               const int* __restrict__ dia_ptr, double* __restrict__ nu_old, double* __restrict__ A,
                               double* __restrict__ res) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nV) return;
  const int i = vnode[t];
  if (i >= Pd) return;
  nu_old[i] = 0.0;
  res[i] = 0.0;
  for (int k = row_ptr[i]; k < row_ptr[i + 1]; ++k) A[k] = 0.0;  // DeleteValsRowi (CSysMatrix.cpp:604-614)
  A[dia_ptr[i]] = 1.0;
}

__global__ void k_turb_time_step(int Pd, double ratio, const double* __restrict__ dt, double* __restrict__ tdt) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < Pd) tdt[i] = ratio * dt[i];
}

// ComputeUnderRelaxationFactorHelper + AddClippedSolution (no clipping limits for SA)
__global__ void k_turb_complete(int Pd, int neg, double allowable, const double* __restrict__ sol,
                                const double* __restrict__ nu_old, double* __restrict__ nu,
                                double* __restrict__ under_relax) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Pd) return;
  double ur = 1.0;
  if (!neg) {
    const double ratio = fabs(sol[i]) / (fabs(nu[i]) + 1.0e-16);
    if (ratio > allowable) ur = fmin(allowable / ratio, ur);
    if (ur < 1e-10) ur = 0.0;
    under_relax[i] = ur;
  } else {
    under_relax[i] = 1.0;  // SA-neg returns early: the factor keeps its initial value (1)
  }
  nu[i] = nu_old[i] + ur * sol[i];
}

// CTurbSASolver::Postprocessing: mu_t = rho fv1 nu_hat
template <int ND>
__global__ void k_turb_mut(int P, int neg, const double* __restrict__ V, const double* __restrict__ nu,
                           double* __restrict__ mut) {
  using I = PrimIdx<ND>;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const double rho = V[(size_t)i * I::NPRIM + I::RHO], mu = V[(size_t)i * I::NPRIM + I::MU];
  const double nul = mu / rho, nu_hat = nu[i];
  const double Ji = nu_hat / nul, Ji_3 = Ji * Ji * Ji;
  const double fv1 = Ji_3 / (Ji_3 + 7.1 * 7.1 * 7.1);
  double m = rho * fv1 * nu_hat;
  if (neg && nu_hat < 0) m = 0.0;
  mut[i] = m;
}

// ---- host wrappers ------------------------------------------------------------------------------------------------------
// SST counterparts (turb_sst.cu)
void sst_preprocessing(su2b200_ctx* c);
void sst_space_integration(su2b200_ctx* c);
void sst_complete_implicit(su2b200_ctx* c);
void sst_postprocessing(su2b200_ctx* c);

void turb_preprocessing(su2b200_ctx* c) {
  auto& s = c->fs;
  if (c->cfg.rans_sst) return sst_preprocessing(c);
  SU2B_REQUIRE(c->cfg.rans_sa, "turbulence entry point called without KIND_TURB_MODEL= SA or SST");
  // residual and Jacobian are fully overwritten by the point kernel / edge kernel; halo rows are zeroed here
  scalar_gradient(c, c->cfg.grad_method, s.nu.p, s.nu_grad.p);
  if (c->nranks > 1) halo_exchange(c, s.nu_grad.p, c->geo.nDim);
}

void turb_set_time_step(su2b200_ctx* c) {
  ProfScope prof_(c, PROF_TURB_OTHER);
  auto& g = c->geo;
  auto& s = c->fs;
  SU2B_LAUNCH(c, k_turb_time_step, nblk(g.Pd, 256), 256, g.Pd, c->cfg.cfl_turb_ratio, s.dt.p, s.turb_dt.p);
}

void turb_space_integration(su2b200_ctx* c) {
  if (c->cfg.rans_sst) return sst_space_integration(c);
  ++c->sys[1].A_gen;
  auto& g = c->geo;
  auto& s = c->fs;
  auto& ls = c->sys[1];
  SU2B_REQUIRE(g.wall_dist.n == (size_t)g.P, "RANS needs CPoint::Wall_Distance (geometry_desc.wall_distance)");
  const int neg = c->cfg.sa_negative;
  // the reference's serial rule for CScalarSolver::ReducerStrategy (see su2b200_flow_config::turb_reducer)
  const int reducer = c->cfg.turb_reducer >= 0
                          ? c->cfg.turb_reducer
                          : ((double)g.E / (512.0 * (double)((g.E + 511) / 512)) < 0.875 ? 1 : 0);
  if (s.turb_edge_flux.n < (size_t)7 * g.E) s.turb_edge_flux.alloc((size_t)7 * g.E);
  {
    ProfScope prof_(c, PROF_TURB_EDGE);
    if (g.nDim == 2)
      SU2B_LAUNCH(c, k_turb_edge<2>, nblk(g.E, 128), 128, g.E, neg, reducer, g.edge_i.p, g.edge_j.p, g.edge_normal.p, g.coord.p,
                  s.V.p, s.nu.p, s.nu_grad.p, g.edge_ij.p, g.edge_ji.p, ls.A.p, s.turb_edge_flux.p);
    else
      SU2B_LAUNCH(c, k_turb_edge<3>, nblk(g.E, 128), 128, g.E, neg, reducer, g.edge_i.p, g.edge_j.p, g.edge_normal.p, g.coord.p,
                  s.V.p, s.nu.p, s.nu_grad.p, g.edge_ij.p, g.edge_ji.p, ls.A.p, s.turb_edge_flux.p);
  }
  {
    ProfScope prof_(c, PROF_TURB_SOURCE);
    if (g.nDim == 2)
      SU2B_LAUNCH(c, k_turb_point<2>, nblk(g.P, 128), 128, g.P, g.Pd, g.E, neg, g.nbr_ptr.p, g.gat_edge.p, g.dia_ptr.p,
                  s.turb_edge_flux.p, s.V.p, s.nu.p, s.vorticity.p, g.wall_dist.p, g.volume.p, ls.A.p, s.turb_res.p);
    else
      SU2B_LAUNCH(c, k_turb_point<3>, nblk(g.P, 128), 128, g.P, g.Pd, g.E, neg, g.nbr_ptr.p, g.gat_edge.p, g.dia_ptr.p,
                  s.turb_edge_flux.p, s.V.p, s.nu.p, s.vorticity.p, g.wall_dist.p, g.volume.p, ls.A.p, s.turb_res.p);
  }
  ProfScope prof_(c, PROF_BC);
  // Space_Integration order: weak BCs (far-field), strong BCs (walls); symmetry / Euler walls are no-ops for scalars
  for (auto& m : g.markers) {
    if (m.kind != SU2B200_FAR_FIELD || m.nVertex == 0) continue;
    if (g.nDim == 2)
      SU2B_LAUNCH(c, k_turb_bc_far<2>, nblk(m.nVertex, 64), 64, m.nVertex, g.Pd, c->cfg.nu_tilde_inf, m.node.p, m.normal.p,
                  s.V.p, m.charac.p, s.nu.p, g.dia_ptr.p, ls.A.p, s.turb_res.p);
    else
      SU2B_LAUNCH(c, k_turb_bc_far<3>, nblk(m.nVertex, 64), 64, m.nVertex, g.Pd, c->cfg.nu_tilde_inf, m.node.p, m.normal.p,
                  s.V.p, m.charac.p, s.nu.p, g.dia_ptr.p, ls.A.p, s.turb_res.p);
  }
  for (auto& m : g.markers) {
    if (m.kind != SU2B200_HEAT_FLUX || m.nVertex == 0) continue;
    SU2B_LAUNCH(c, k_turb_bc_wall, nblk(m.nVertex, 64), 64, m.nVertex, g.Pd, m.node.p, g.row_ptr.p, g.dia_ptr.p, s.nu_old.p,
                ls.A.p, s.turb_res.p);
  }
}

void turb_complete_implicit(su2b200_ctx* c) {
  if (c->cfg.rans_sst) return sst_complete_implicit(c);
  ProfScope prof_(c, PROF_UPDATE);
  auto& g = c->geo;
  auto& s = c->fs;
  SU2B_LAUNCH(c, k_turb_complete, nblk(g.Pd, 256), 256, g.Pd, c->cfg.sa_negative, c->cfg.max_update_sa, s.turb_sol.p,
              s.nu_old.p, s.nu.p, s.turb_under_relax.p);
  if (c->nranks > 1) halo_exchange(c, s.nu.p, 1);
}

void turb_postprocessing(su2b200_ctx* c) {
  if (c->cfg.rans_sst) return sst_postprocessing(c);
  ProfScope prof_(c, PROF_TURB_OTHER);
  auto& g = c->geo;
  auto& s = c->fs;
  if (g.nDim == 2) SU2B_LAUNCH(c, k_turb_mut<2>, nblk(g.P, 256), 256, g.P, c->cfg.sa_negative, s.V.p, s.nu.p, s.mu_t.p);
  else SU2B_LAUNCH(c, k_turb_mut<3>, nblk(g.P, 256), 256, g.P, c->cfg.sa_negative, s.V.p, s.nu.p, s.mu_t.p);
}

}  // namespace su2b200
