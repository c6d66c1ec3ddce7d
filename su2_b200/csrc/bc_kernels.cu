// a14. Boundary-vertex kernels (one thread per vertex, one launch per marker, markers in the order of
// CIntegration::Space_Integration, CIntegration.cpp:105-200). A point appears at most once per marker, so
// the read-modify-write of its residual / Jacobian row is race free without atomics.
#include "internal.hpp"
#include "device_math.cuh"
#include "kernels.hpp"

namespace su2b200 {

static inline int nblk(long n, int bs) { return (int)((n + bs - 1) / bs); }
FlowParams flow_params(const su2b200_ctx* c);

struct FarFieldInf {
  double rho, p, vel[3];
};

// CEulerSolver::BC_Far_Field (CEulerSolver.cpp:4780-5034): Riemann-invariant far-field state, then the
// upwind (Roe) boundary flux of CUpwRoe_Flow (roe.cpp:84-249) added to residual and diagonal block;
// viscous part through CAvgGrad_Flow (flow_diffusion.cpp:368-484) with the reflected coordinate.
template <int ND, bool VISC>
__global__ void k_bc_far_field(int nV, int Pd, FlowParams p, FarFieldInf inf, const int* __restrict__ vnode,
                               const double* __restrict__ vnormal, const int* __restrict__ vnbr,
                               const double* __restrict__ coord, const double* __restrict__ V,
                               const double* __restrict__ grad_prim, int ngs, double* __restrict__ charac,
                               const int* __restrict__ dia_ptr, double* __restrict__ A, double* __restrict__ res) {
  using I = PrimIdx<ND>;
  constexpr int NV = ND + 2;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nV) return;
  const int i = vnode[t];
  if (i >= Pd) return;
  double normal[ND], un[ND];
#pragma unroll
  for (int d = 0; d < ND; ++d) normal[d] = -vnormal[(size_t)t * ND + d];
  const double area = sqrt(dotN<ND>(normal, normal));
#pragma unroll
  for (int d = 0; d < ND; ++d) un[d] = normal[d] / area;
  double Vd[I::NPRIM];
#pragma unroll
  for (int k = 0; k < I::NPRIM; ++k) Vd[k] = V[(size_t)i * I::NPRIM + k];

  const double gamma = p.gamma, gm1 = p.gamma - 1.0;
  const double rho_b = Vd[I::RHO];
  double vn_b = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) vn_b += Vd[I::VEL + d] * un[d];
  const double p_b = Vd[I::P];
  const double c_b = sqrt(gamma * p_b / rho_b);
  const double s_b = pow(rho_b, gamma) / p_b;

  double vn_inf = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) vn_inf += inf.vel[d] * un[d];
  const double c_inf = sqrt(gamma * inf.p / inf.rho);
  const double s_inf = pow(inf.rho, gamma) / inf.p;
  const double qn_inf = vn_inf;

  double Rp, Rm;
  if (qn_inf > -c_inf) Rp = vn_b + 2.0 * c_b / gm1; else Rp = vn_inf + 2.0 * c_inf / gm1;
  if (qn_inf > c_inf) Rm = vn_b - 2.0 * c_b / gm1; else Rm = vn_inf - 2.0 * c_inf / gm1;
  const double vn = 0.5 * (Rp + Rm);
  const double c = 0.25 * (Rp - Rm) * gm1;
  double vel[ND], entropy;
  if (qn_inf > 0.0) {
#pragma unroll
    for (int d = 0; d < ND; ++d) vel[d] = Vd[I::VEL + d] + (vn - vn_b) * un[d];
    entropy = s_b;
  } else {
#pragma unroll
    for (int d = 0; d < ND; ++d) vel[d] = inf.vel[d] + (vn - vn_inf) * un[d];
    entropy = s_inf;
  }
  const double rho = pow(entropy * c * c / gamma, 1.0 / gm1);
  double v2 = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) v2 += vel[d] * vel[d];
  const double pres = rho * c * c / gamma;
  const double energy = pres / (gm1 * rho) + 0.5 * v2;

  double* Vinf = charac + (size_t)t * I::NPRIM;  // persistent CharacPrimVar: entries not set here keep their value
  double Vj[I::NPRIM];
#pragma unroll
  for (int k = 0; k < I::NPRIM; ++k) Vj[k] = Vinf[k];
  Vj[I::T] = pres / (p.R * rho);
#pragma unroll
  for (int d = 0; d < ND; ++d) Vj[I::VEL + d] = vel[d];
  Vj[I::P] = pres;
  Vj[I::RHO] = rho;
  Vj[I::H] = energy + pres / rho;

  double flux[NV], jac_i[NV * NV], jac_j[NV * NV];
  // legacy CUpwRoeBase_Flow guards against a negative Roe sound speed (roe.cpp:139-155): flux = 0
  {
    const double R = sqrt(fabs(Vj[I::RHO] / Vd[I::RHO]));
    double sq = 0.0;
#pragma unroll
    for (int d = 0; d < ND; ++d) { const double rv = (R * Vj[I::VEL + d] + Vd[I::VEL + d]) / (R + 1); sq += rv * rv; }
    const double rH = (R * Vj[I::H] + Vd[I::H]) / (R + 1);
    if ((gamma - 1) * (rH - 0.5 * sq) <= 0.0) {
#pragma unroll
      for (int k = 0; k < NV; ++k) flux[k] = 0.0;
#pragma unroll
      for (int k = 0; k < NV * NV; ++k) jac_i[k] = 0.0;
    } else {
      FlowParams q = p;
      q.roe_kappa = 0.5;  // legacy functor: kappa hard-wired through GetRoe_Kappa too; same value
      q.roe_kappa = p.roe_kappa;
      roe_flux<ND, true>(q, Vd, Vj, normal, flux, jac_i, jac_j);
    }
  }
  double* D = A + (size_t)dia_ptr[i] * NV * NV;
#pragma unroll
  for (int k = 0; k < NV; ++k) res[(size_t)i * NV + k] += flux[k];
#pragma unroll
  for (int k = 0; k < NV * NV; ++k) D[k] += jac_i[k];

  if (VISC) {
    Vj[I::MU] = Vd[I::MU];
    Vj[I::MUT] = Vd[I::MUT];
    const int nb = vnbr[t];
    double vij[ND];
#pragma unroll
    for (int d = 0; d < ND; ++d) {
      // PointPointReflect(Point_Normal, iPoint): reflected = 2*x_i - x_nbr ; edge vector = reflected - x_i
      const double xi = coord[(size_t)i * ND + d], xn = coord[(size_t)nb * ND + d];
      const double refl = xi + (xi - xn);
      vij[d] = refl - xi;
    }
    double vflux[NV], vjac_i[NV * NV], vjac_j[NV * NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) vflux[k] = 0.0;
#pragma unroll
    for (int k = 0; k < NV * NV; ++k) { vjac_i[k] = 0.0; vjac_j[k] = 0.0; }
    const double* g = grad_prim + (size_t)i * ngs * ND;
    viscous_terms<ND, true>(p, Vd, Vj, g, g, vij, area, un, true, vflux, vjac_i, vjac_j);
    // viscous_terms subtracts the viscous flux and its Jacobian from (flux, jac): exactly
    // LinSysRes.SubtractBlock / Jacobian.SubtractBlock2Diag of the reference.
#pragma unroll
    for (int k = 0; k < NV; ++k) res[(size_t)i * NV + k] += vflux[k];
#pragma unroll
    for (int k = 0; k < NV * NV; ++k) D[k] += vjac_i[k];
  }
#pragma unroll
  for (int k = 0; k < I::NPRIM; ++k) Vinf[k] = Vj[k];
}

// CFVMFlowSolverBase::BC_Sym_Plane == BC_Euler_Wall (CFVMFlowSolverBase.inl:1114-1272), static grid.
template <int ND>
__global__ void k_bc_sym_plane(int nV, int Pd, const int* __restrict__ vnode, const double* __restrict__ unit,
                               const int* __restrict__ row_ptr, double* __restrict__ U_old, double* __restrict__ res,
                               double* __restrict__ A) {
  constexpr int NV = ND + 2;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nV) return;
  const int i = vnode[t];
  if (i >= Pd) return;
  double n[ND];
#pragma unroll
  for (int d = 0; d < ND; ++d) n[d] = unit[(size_t)t * ND + d];
  // remove normal momentum of Solution_Old
  double vp = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) vp += U_old[(size_t)i * NV + 1 + d] * n[d];
#pragma unroll
  for (int d = 0; d < ND; ++d) U_old[(size_t)i * NV + 1 + d] -= vp * n[d];
  // tangential momentum residual only
  double nr = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) nr += res[(size_t)i * NV + 1 + d] * n[d];
#pragma unroll
  for (int d = 0; d < ND; ++d) res[(size_t)i * NV + 1 + d] -= nr * n[d];
  // J_row <- (I - n n^T) J_row for every block of row i (diagonal + all neighbours == whole CSR row)
  double mat[NV * NV];
#pragma unroll
  for (int k = 0; k < NV * NV; ++k) mat[k] = 0.0;
#pragma unroll
  for (int k = 0; k < NV; ++k) mat[k * NV + k] = 1.0;
#pragma unroll
  for (int a = 0; a < ND; ++a)
#pragma unroll
    for (int b = 0; b < ND; ++b) mat[(a + 1) * NV + b + 1] -= n[a] * n[b];
  for (int k = row_ptr[i]; k < row_ptr[i + 1]; ++k) {
    double* B = A + (size_t)k * NV * NV;
    double jac[NV * NV];
#pragma unroll
    for (int q = 0; q < NV * NV; ++q) jac[q] = B[q];
#pragma unroll
    for (int r = 0; r < NV; ++r)
#pragma unroll
      for (int cidx = 0; cidx < NV; ++cidx) {
        double s = 0.0;
#pragma unroll
        for (int m = 0; m < NV; ++m) s += mat[r * NV + m] * jac[m * NV + cidx];
        B[r * NV + cidx] = s;
      }
  }
}

// CNSSolver::BC_HeatFlux_Wall_Generic (CNSSolver.cpp:423-578), static grid, HEAT_FLUX kind.
template <int ND>
__global__ void k_bc_heatflux_wall(int nV, int Pd, double wall_heat_flux, const int* __restrict__ vnode,
                                   const double* __restrict__ vnormal, const int* __restrict__ row_ptr,
                                   const int* __restrict__ dia_ptr, double* __restrict__ U_old,
                                   double* __restrict__ res, double* __restrict__ A) {
  constexpr int NV = ND + 2;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nV) return;
  const int i = vnode[t];
  if (i >= Pd) return;
  double area2 = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) { const double x = vnormal[(size_t)t * ND + d]; area2 += x * x; }
  const double area = sqrt(area2);
  const double res_visc = wall_heat_flux * area;
  // SetVelocity_Old(zero): Solution_Old momentum = 0 * Solution density
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    U_old[(size_t)i * NV + 1 + d] = 0.0;
    res[(size_t)i * NV + 1 + d] = 0.0;
  }
  res[(size_t)i * NV + ND + 1] += 0.0 - res_visc;
  // DeleteValsRowi for the momentum rows (CSysMatrix.cpp:604-614): zero the row in every block, 1 on the diagonal
  for (int k = row_ptr[i]; k < row_ptr[i + 1]; ++k) {
    double* B = A + (size_t)k * NV * NV;
#pragma unroll
    for (int r = 1; r <= ND; ++r)
#pragma unroll
      for (int cidx = 0; cidx < NV; ++cidx) B[r * NV + cidx] = 0.0;
  }
  double* D = A + (size_t)dia_ptr[i] * NV * NV;
#pragma unroll
  for (int r = 1; r <= ND; ++r) D[r * NV + r] = 1.0;
}

void flow_bc_far_field(su2b200_ctx* c, int iMarker) {
  auto& g = c->geo;
  auto& s = c->fs;
  auto& m = g.markers.at(iMarker);
  if (m.nVertex == 0) return;
  const FlowParams p = flow_params(c);
  FarFieldInf inf;
  inf.rho = c->cfg.density_inf;
  inf.p = c->cfg.pressure_inf;
  for (int d = 0; d < 3; ++d) inf.vel[d] = c->cfg.velocity_inf[d];
  const int nb = nblk(m.nVertex, 64);
  double* A = c->sys[0].A.p;
  if (g.nDim == 2) {
    if (c->cfg.viscous)
      SU2B_LAUNCH(c, (k_bc_far_field<2, true>), nb, 64, m.nVertex, g.Pd, p, inf, m.node.p, m.normal.p, m.normal_nbr.p,
                  g.coord.p, s.V.p, s.grad_prim.p, s.nG, m.charac.p, g.dia_ptr.p, A, s.res.p);
    else
      SU2B_LAUNCH(c, (k_bc_far_field<2, false>), nb, 64, m.nVertex, g.Pd, p, inf, m.node.p, m.normal.p, m.normal_nbr.p,
                  g.coord.p, s.V.p, s.grad_prim.p, s.nG, m.charac.p, g.dia_ptr.p, A, s.res.p);
  } else {
    if (c->cfg.viscous)
      SU2B_LAUNCH(c, (k_bc_far_field<3, true>), nb, 64, m.nVertex, g.Pd, p, inf, m.node.p, m.normal.p, m.normal_nbr.p,
                  g.coord.p, s.V.p, s.grad_prim.p, s.nG, m.charac.p, g.dia_ptr.p, A, s.res.p);
    else
      SU2B_LAUNCH(c, (k_bc_far_field<3, false>), nb, 64, m.nVertex, g.Pd, p, inf, m.node.p, m.normal.p, m.normal_nbr.p,
                  g.coord.p, s.V.p, s.grad_prim.p, s.nG, m.charac.p, g.dia_ptr.p, A, s.res.p);
  }
}

void flow_bc_sym_plane(su2b200_ctx* c, int iMarker) {
  auto& g = c->geo;
  auto& s = c->fs;
  auto& m = g.markers.at(iMarker);
  if (m.nVertex == 0) return;
  const int nb = nblk(m.nVertex, 64);
  if (g.nDim == 2)
    SU2B_LAUNCH(c, k_bc_sym_plane<2>, nb, 64, m.nVertex, g.Pd, m.node.p, m.unit_sym.p, g.row_ptr.p, s.U_old.p, s.res.p, c->sys[0].A.p);
  else
    SU2B_LAUNCH(c, k_bc_sym_plane<3>, nb, 64, m.nVertex, g.Pd, m.node.p, m.unit_sym.p, g.row_ptr.p, s.U_old.p, s.res.p, c->sys[0].A.p);
}

void flow_bc_heatflux_wall(su2b200_ctx* c, int iMarker) {
  auto& g = c->geo;
  auto& s = c->fs;
  auto& m = g.markers.at(iMarker);
  if (m.nVertex == 0) return;
  const int nb = nblk(m.nVertex, 64);
  if (g.nDim == 2)
    SU2B_LAUNCH(c, k_bc_heatflux_wall<2>, nb, 64, m.nVertex, g.Pd, m.heat_flux, m.node.p, m.normal.p, g.row_ptr.p, g.dia_ptr.p, s.U_old.p, s.res.p, c->sys[0].A.p);
  else
    SU2B_LAUNCH(c, k_bc_heatflux_wall<3>, nb, 64, m.nVertex, g.Pd, m.heat_flux, m.node.p, m.normal.p, g.row_ptr.p, g.dia_ptr.p, s.U_old.p, s.res.p, c->sys[0].A.p);
}

}  // namespace su2b200
