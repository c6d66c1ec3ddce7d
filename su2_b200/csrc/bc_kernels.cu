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
  double tke;  // GetTke_Inf() with SST (tkeNeeded, CEulerSolver.cpp:4801, :4951), else 0
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
  double energy = pres / (gm1 * rho) + 0.5 * v2;
  energy += inf.tke;  // if (tkeNeeded) Energy += GetTke_Inf(); x + 0.0 == x otherwise

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
      roe_flux<ND, true>(p, Vd, Vj, normal, flux, jac_i, jac_j);
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
    // visc_bound_term = CAvgGrad_Flow(nDim, nVar, correct_gradient=false) (CDriver.cpp:1810)
    viscous_terms<ND, true, true>(p, Vd, Vj, g, g, vij, area, un, false, vflux, vjac_i, vjac_j);
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
                               double* __restrict__ A, double* __restrict__ trunc) {
  constexpr int NV = ND + 2;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nV) return;
  const int i = vnode[t];
  if (i >= Pd) return;
  double n[ND];
#pragma unroll
  for (int d = 0; d < ND; ++d) n[d] = unit[(size_t)t * ND + d];
  if (trunc) {  // "Correction for multigrid" (CFVMFlowSolverBase.inl:1260-1268): the forcing term loses its normal momentum
    double nt = 0.0;
#pragma unroll
    for (int d = 0; d < ND; ++d) nt += trunc[(size_t)i * NV + 1 + d] * n[d];
#pragma unroll
    for (int d = 0; d < ND; ++d) trunc[(size_t)i * NV + 1 + d] -= nt * n[d];
  }
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

// momentum components of a [P][nDim+2] array := 0 at the vertices of a marker
__global__ void k_zero_velocity(int nV, int Pd, int nDim, const int* __restrict__ vnode, double* __restrict__ arr) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nV) return;
  const int i = vnode[t];
  if (i >= Pd) return;
  for (int d = 0; d < nDim; ++d) arr[(size_t)i * (nDim + 2) + 1 + d] = 0.0;
}

void flow_bc_far_field(su2b200_ctx* c, int iMarker) {
  ++c->sys[0].A_gen;
  ProfScope prof_(c, PROF_BC);
  auto& g = c->geo;
  auto& s = c->fs;
  auto& m = g.markers.at(iMarker);
  if (m.nVertex == 0) return;
  const FlowParams p = flow_params(c);
  FarFieldInf inf;
  inf.rho = c->cfg.density_inf;
  inf.p = c->cfg.pressure_inf;
  inf.tke = c->cfg.rans_sst ? c->cfg.kine_inf : 0.0;
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
  ++c->sys[0].A_gen;
  ProfScope prof_(c, PROF_BC);
  auto& g = c->geo;
  auto& s = c->fs;
  auto& m = g.markers.at(iMarker);
  if (m.nVertex == 0) return;
  const int nb = nblk(m.nVertex, 64);
  if (g.nDim == 2)
    SU2B_LAUNCH(c, k_bc_sym_plane<2>, nb, 64, m.nVertex, g.Pd, m.node.p, m.unit_sym.p, g.row_ptr.p, s.U_old.p, s.res.p, c->sys[0].A.p, s.res_trunc.p);
  else
    SU2B_LAUNCH(c, k_bc_sym_plane<3>, nb, 64, m.nVertex, g.Pd, m.node.p, m.unit_sym.p, g.row_ptr.p, s.U_old.p, s.res.p, c->sys[0].A.p, s.res_trunc.p);
}

void flow_bc_heatflux_wall(su2b200_ctx* c, int iMarker) {
  ++c->sys[0].A_gen;
  ProfScope prof_(c, PROF_BC);
  auto& g = c->geo;
  auto& s = c->fs;
  auto& m = g.markers.at(iMarker);
  if (m.nVertex == 0) return;
  const int nb = nblk(m.nVertex, 64);
  if (g.nDim == 2)
    SU2B_LAUNCH(c, k_bc_heatflux_wall<2>, nb, 64, m.nVertex, g.Pd, m.heat_flux, m.node.p, m.normal.p, g.row_ptr.p, g.dia_ptr.p, s.U_old.p, s.res.p, c->sys[0].A.p);
  else
    SU2B_LAUNCH(c, k_bc_heatflux_wall<3>, nb, 64, m.nVertex, g.Pd, m.heat_flux, m.node.p, m.normal.p, g.row_ptr.p, g.dia_ptr.p, s.U_old.p, s.res.p, c->sys[0].A.p);
  if (s.res_trunc.p)  // SetVel_ResTruncError_Zero (CNSSolver.cpp:511): no forcing on the momentum of a no-slip wall
    SU2B_LAUNCH(c, k_zero_velocity, nb, 64, m.nVertex, g.Pd, g.nDim, m.node.p, s.res_trunc.p);
}

}  // namespace su2b200

// ------------------------------------------------------------------------------------------------
// a29. CFVMFlowSolverBase::Pressure_Forces (CFVMFlowSolverBase.inl:1784-2100): per-vertex force and moment,
// then a deterministic single-block reduction. out6 = (Fx,Fy,Fz,Mx,My,Mz) already scaled by 1/AeroCoeffForceRef.
// ------------------------------------------------------------------------------------------------
namespace su2b200 {

template <int ND>
__global__ void k_pressure_force(int nV, int Pd, double p_inf, double factor, double ref_len, double ox, double oy,
                                 double oz, const int* __restrict__ vnode, const double* __restrict__ vnormal,
                                 const double* __restrict__ coord, const double* __restrict__ V,
                                 double* __restrict__ out /*[nV][6]*/) {
  using I = PrimIdx<ND>;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nV) return;
  const int i = vnode[t];
  double f[3] = {0, 0, 0}, m[3] = {0, 0, 0};
  if (i < Pd) {
    const double pres = V[(size_t)i * I::NPRIM + I::P];
    double x[3] = {0, 0, 0};
    const double o[3] = {ox, oy, oz};
    double md[3] = {0, 0, 0};
#pragma unroll
    for (int d = 0; d < ND; ++d) {
      x[d] = coord[(size_t)i * ND + d];
      md[d] = x[d] - o[d];
      f[d] = -(pres - p_inf) * vnormal[(size_t)t * ND + d] * factor;
    }
    if (ND == 3) {
      m[0] = (f[2] * md[1] - f[1] * md[2]) / ref_len;
      m[1] = (f[0] * md[2] - f[2] * md[0]) / ref_len;
    }
    m[2] = (f[1] * md[0] - f[0] * md[1]) / ref_len;
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) { out[(size_t)t * 6 + k] = f[k]; out[(size_t)t * 6 + 3 + k] = m[k]; }
}

// CFVMFlowSolverBase::Friction_Forces (CFVMFlowSolverBase.inl:2385-2720): tau from the laminar viscosity and the
// primitive gradient at the wall point (CNumerics::ComputeStressTensor), force = tau.n * Area / AeroCoeffForceRef.
template <int ND>
__global__ void k_friction_force(int nV, int Pd, int ngs, double factor, double ref_len, double ox, double oy, double oz,
                                 const int* __restrict__ vnode, const double* __restrict__ vnormal,
                                 const double* __restrict__ coord, const double* __restrict__ V,
                                 const double* __restrict__ grad_prim, double* __restrict__ out /*[nV][6]*/) {
  using I = PrimIdx<ND>;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nV) return;
  const int i = vnode[t];
  double f[3] = {0, 0, 0}, m[3] = {0, 0, 0};
  if (i < Pd) {
    const double mu = V[(size_t)i * I::NPRIM + I::MU];
    const double* g = grad_prim + (size_t)i * ngs * ND;  // g[(1+a)*ND+b] = d v_a / d x_b
    double n[ND], area = 0.0;
#pragma unroll
    for (int d = 0; d < ND; ++d) { n[d] = vnormal[(size_t)t * ND + d]; area += n[d] * n[d]; }
    area = sqrt(area);
#pragma unroll
    for (int d = 0; d < ND; ++d) n[d] /= area;
    double div = 0.0;
#pragma unroll
    for (int d = 0; d < ND; ++d) div += g[(1 + d) * ND + d];
    const double pTerm = 2.0 / 3.0 * mu * div;
    const double o[3] = {ox, oy, oz};
    double md[3] = {0, 0, 0};
#pragma unroll
    for (int a = 0; a < ND; ++a) {
      double te = 0.0;
#pragma unroll
      for (int b = 0; b < ND; ++b) {
        double tau = mu * (g[(1 + b) * ND + a] + g[(1 + a) * ND + b]);
        if (a == b) tau -= pTerm;
        te += tau * n[b];
      }
      f[a] = te * area * factor;
      md[a] = coord[(size_t)i * ND + a] - o[a];
    }
    if (ND == 3) {
      m[0] = (f[2] * md[1] - f[1] * md[2]) / ref_len;
      m[1] = (f[0] * md[2] - f[2] * md[0]) / ref_len;
    }
    m[2] = (f[1] * md[0] - f[0] * md[1]) / ref_len;
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) { out[(size_t)t * 6 + k] = f[k]; out[(size_t)t * 6 + 3 + k] = m[k]; }
}

// sum over the vertices of a marker, fixed order (the reference adds vertex by vertex): blocks of 256 consecutive
// vertices are summed by a binary tree in shared memory, the block sums then added in ascending order by one thread per
// component (a single block walking the whole marker took 0.34 ms per call on a 185^2-vertex wall)
__global__ void __launch_bounds__(256) k_force_partial(int nV, const double* __restrict__ in, double* __restrict__ part) {
  __shared__ double sm[6][256];
  const int t = threadIdx.x;
  const size_t v = (size_t)blockIdx.x * 256 + t;
#pragma unroll
  for (int k = 0; k < 6; ++k) sm[k][t] = v < (size_t)nV ? in[v * 6 + k] : 0.0;
  __syncthreads();
  for (int h = 128; h > 0; h >>= 1) {
    if (t < h) {
#pragma unroll
      for (int k = 0; k < 6; ++k) sm[k][t] += sm[k][t + h];
    }
    __syncthreads();
  }
  if (t < 6) part[(size_t)blockIdx.x * 6 + t] = sm[t][0];
}

__global__ void k_force_sum(int nB, const double* __restrict__ part, double* __restrict__ acc6) {
  const int k = threadIdx.x;
  if (k >= 6) return;
  double s = 0.0;
  for (int b = 0; b < nB; ++b) s += part[(size_t)b * 6 + k];
  acc6[k] = s;
}

void flow_forces(su2b200_ctx* c, double* coeffs) {
  ProfScope prof_(c, PROF_FORCES);
  auto& g = c->geo;
  auto& s = c->fs;
  const auto& cfg = c->cfg;
  double v2 = 0.0;
  for (int d = 0; d < g.nDim; ++d) v2 += cfg.velocity_inf[d] * cfg.velocity_inf[d];
  const double force_ref = 0.5 * cfg.density_inf * cfg.ref_area * v2;  // SetReferenceValues
  const double factor = 1.0 / force_ref;
  const double PI = 3.14159265358979323846;
  const double alpha = cfg.aoa_deg * PI / 180.0, beta = cfg.aos_deg * PI / 180.0;
  double CL = 0, CD = 0, CSF = 0, CF[3] = {0, 0, 0}, CM[3] = {0, 0, 0};
  // pass 0: Pressure_Forces on every monitored marker; pass 1: Friction_Forces on monitored viscous walls
  for (int pass = 0; pass < (cfg.viscous ? 2 : 1); ++pass)
  for (auto& m : g.markers) {
    if (!m.monitoring || m.nVertex == 0) continue;
    if (pass == 1 && m.kind != SU2B200_HEAT_FLUX) continue;
    const int nPart = nblk(m.nVertex, 256);
    const size_t need = (size_t)m.nVertex * 6 + 8 + (size_t)nPart * 6;
    if (s.red.n < need) s.red.alloc(need);
    double* buf = s.red.p;
    double* acc = s.red.p + (size_t)m.nVertex * 6;
    double* part = acc + 8;
    const int nb = nblk(m.nVertex, 128);
    if (pass == 0) {
      if (g.nDim == 2)
        SU2B_LAUNCH(c, k_pressure_force<2>, nb, 128, m.nVertex, g.Pd, cfg.pressure_inf, factor, cfg.ref_length,
                    cfg.ref_origin[0], cfg.ref_origin[1], cfg.ref_origin[2], m.node.p, m.normal.p, g.coord.p, s.V.p, buf);
      else
        SU2B_LAUNCH(c, k_pressure_force<3>, nb, 128, m.nVertex, g.Pd, cfg.pressure_inf, factor, cfg.ref_length,
                    cfg.ref_origin[0], cfg.ref_origin[1], cfg.ref_origin[2], m.node.p, m.normal.p, g.coord.p, s.V.p, buf);
    } else {
      if (g.nDim == 2)
        SU2B_LAUNCH(c, k_friction_force<2>, nb, 128, m.nVertex, g.Pd, s.nG, factor, cfg.ref_length, cfg.ref_origin[0],
                    cfg.ref_origin[1], cfg.ref_origin[2], m.node.p, m.normal.p, g.coord.p, s.V.p, s.grad_prim.p, buf);
      else
        SU2B_LAUNCH(c, k_friction_force<3>, nb, 128, m.nVertex, g.Pd, s.nG, factor, cfg.ref_length, cfg.ref_origin[0],
                    cfg.ref_origin[1], cfg.ref_origin[2], m.node.p, m.normal.p, g.coord.p, s.V.p, s.grad_prim.p, buf);
    }
    SU2B_LAUNCH(c, k_force_partial, nPart, 256, m.nVertex, buf, part);
    SU2B_LAUNCH(c, k_force_sum, 1, 32, nPart, part, acc);
    SU2B_CUDA(cudaMemcpyAsync(s.h_red, acc, 6 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    SU2B_CUDA(cudaStreamSynchronize(c->stream));
    const double* F = s.h_red;
    if (g.nDim == 2) {
      CD += F[0] * cos(alpha) + F[1] * sin(alpha);
      CL += -F[0] * sin(alpha) + F[1] * cos(alpha);
    } else {
      CD += F[0] * cos(alpha) * cos(beta) + F[1] * sin(beta) + F[2] * sin(alpha) * cos(beta);
      CL += -F[0] * sin(alpha) + F[2] * cos(alpha);
      CSF += -F[0] * sin(beta) * cos(alpha) + F[1] * cos(beta) - F[2] * sin(beta) * sin(alpha);
    }
    for (int k = 0; k < 3; ++k) { CF[k] += F[k]; CM[k] += F[3 + k]; }
  }
  if (c->nranks > 1) {  // every coefficient is linear in the surface integrals: one sum over the ranks
    double v[9] = {CL, CD, CSF, CF[0], CF[1], CF[2], CM[0], CM[1], CM[2]};
    allreduce_vec(c, v, 9, 0);
    CL = v[0]; CD = v[1]; CSF = v[2];
    for (int k = 0; k < 3; ++k) { CF[k] = v[3 + k]; CM[k] = v[6 + k]; }
  }
  if (coeffs) {
    coeffs[0] = CL; coeffs[1] = CD; coeffs[2] = CSF;
    coeffs[3] = CF[0]; coeffs[4] = CF[1]; coeffs[5] = CF[2];
    coeffs[6] = CM[0]; coeffs[7] = CM[1]; coeffs[8] = CM[2];
  }
}

}  // namespace su2b200
