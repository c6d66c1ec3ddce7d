// CB200FlowBackend -- the reference-side binding of libsu2b200.so (include/su2b200.h).
//
// This header is compiled INTO a scratch copy of SU2 (integration/build_shim.sh copies /root/reference to /tmp, applies
// integration/apply_patch.py and builds SU2_CFD with -DHAVE_SU2B200 -lsu2b200); nothing of it belongs to the reference
// and nothing of the reference is copied here. It is what a maintainer would add to run the implicit flow iteration of
// SU2_CFD on a B200 under the existing driver, config system, mesh reader and output: ENABLE_CUDA= YES in the cfg
// selects it at the three places the patch touches
//   (1) CSysMatrixVectorProduct::operator()     (Common/include/linear_algebra/CMatrixVectorProduct.hpp:98-109, the
//       reference's own GPU hook)                                     -> su2b200_linsys_spmv
//   (2) CFVMFlowSolverBase::ImplicitEuler_Iteration's System.Solve    (SU2_CFD/include/solvers/CFVMFlowSolverBase.inl:608)
//                                                                     -> su2b200_linsys_solve       [SU2B200_MODE=linear]
//   (3) CMultiGridIntegration::MultiGrid_Iteration's MultiGrid_Cycle  (SU2_CFD/src/integration/CMultiGridIntegration.cpp:109)
//       for the flow system at MGLEVEL= 0: the CSolver virtuals one by one
//       (Preprocessing, Set_OldSolution, SetTime_Step, Space_Integration, PrepareImplicitIteration, System.Solve,
//       CompleteImplicitIteration)                                    -> su2b200_flow_*             [default mode]
// Geometry, markers and configuration come from the reference's own CGeometry / CConfig objects; the solution lives in
// the reference's CVariable between iterations (uploaded before, downloaded after every iteration), so output, forces,
// the turbulence solver (SA or SST: its eddy viscosity, and for SST its k, are uploaded every iteration) and convergence
// monitoring stay the reference's.
#pragma once
#ifdef HAVE_SU2B200

#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

/*--- lives in Common/include/linear_algebra/ of the patched tree (integration/apply_patch.py copies it there) ---*/
#include "su2b200.h"
#include "../CConfig.hpp"
#include "../geometry/CGeometry.hpp"
#include "CSysMatrix.hpp"
#include "CSysVector.hpp"

#define B200_CHECK(call)                                                       \
  do {                                                                         \
    if ((call) != 0) SU2_MPI::Error(su2b200_last_error(), CURRENT_FUNCTION);   \
  } while (0)

class CB200FlowBackend {
 public:
  static CB200FlowBackend& Get() {
    static CB200FlowBackend instance;
    return instance;
  }

  /*--- SU2B200_MODE=linear keeps the residual assembly on the CPU and moves only System.Solve to the GPU. ---*/
  static bool WholeIteration() {
    const char* m = std::getenv("SU2B200_MODE");
    return !(m && std::string(m) == "linear");
  }

  ~CB200FlowBackend() {
    if (flowCtx) su2b200_destroy(flowCtx);
    if (linCtx) su2b200_destroy(linCtx);
  }

  /*--- (1) y = A x through the library, host buffers (the call CSysMatrix::GPUMatrixVectorProduct stands for). ---*/
  template <class ScalarType>
  void MatrixVectorProduct(const CSysMatrix<ScalarType>& A, const CSysVector<ScalarType>& u, CSysVector<ScalarType>& v) {
    BindLinearSystem(A, 1);
    UploadMatrix(A, 1);
    std::vector<double> x(u.GetLocSize()), y(u.GetLocSize(), 0.0);
    for (size_t k = 0; k < x.size(); ++k) x[k] = SU2_TYPE::GetValue(u[k]);
    B200_CHECK(su2b200_linsys_spmv(linCtx, 1, x.data(), y.data()));
    for (size_t k = 0; k < y.size(); ++k) v[k] = y[k];
  }

  /*--- (2) System.Solve(Jacobian, LinSysRes, LinSysSol): preconditioner build + Krylov solve on the GPU. ---*/
  template <class ScalarType>
  unsigned long Solve(const CSysMatrix<ScalarType>& A, const CSysVector<su2double>& b, CSysVector<su2double>& x,
                      const CConfig* config, su2double* residual) {
    BindLinearSystem(A, 0);
    UploadMatrix(A, 0);
    std::vector<double> hb(b.GetLocSize()), hx(b.GetLocSize(), 0.0);
    for (size_t k = 0; k < hb.size(); ++k) hb[k] = SU2_TYPE::GetValue(b[k]);
    unsigned long iters = 0;
    double res = 0.0;
    B200_CHECK(su2b200_linsys_solve(linCtx, 0, SolverKind(config), PrecKind(config),
                                    SU2_TYPE::GetValue(config->GetLinear_Solver_Error()),
                                    (int)config->GetLinear_Solver_Iter(), hb.data(), hx.data(), &iters, &res));
    for (size_t k = 0; k < hx.size(); ++k) x[k] = hx[k];
    if (residual) *residual = res;
    return iters;
  }

  /*--- (3) one implicit iteration of the flow solver on the device. `Solver` is CSolver (kept a template parameter so
   *        that this header does not pull the solver headers into Common). Returns the Krylov iteration count. ---*/
  template <class Solver>
  unsigned long FlowIteration(CGeometry* geometry, Solver** solver_container, CConfig* config, int flowSol, int turbSol,
                              std::vector<double>& rms, su2double* linRes) {
    Solver* flow = solver_container[flowSol];
    const unsigned long nPoint = geometry->GetnPoint();
    const unsigned short nVar = flow->GetnVar(), nDim = geometry->GetnDim();
    const bool rans = config->GetKind_Turb_Model() != TURB_MODEL::NONE;
    const bool sst = config->GetKind_Turb_Model() == TURB_MODEL::SST;
    if (!flowCtx) SetupFlow(geometry, flow, rans ? solver_container[turbSol] : nullptr, config);
    /*--- state in: Solution (and the eddy viscosity of the reference's turbulence solver; with SST also (k, omega):
     *    k enters the static energy of the flow primitives, CNSVariable::SetPrimVar) ---*/
    std::vector<double> buf((size_t)nPoint * nVar);
    for (unsigned long i = 0; i < nPoint; ++i)
      for (unsigned short v = 0; v < nVar; ++v) buf[i * nVar + v] = SU2_TYPE::GetValue(flow->GetNodes()->GetSolution(i, v));
    B200_CHECK(su2b200_array_upload(flowCtx, SU2B200_ARR_SOLUTION, buf.data()));
    if (rans) {
      std::vector<double> mut(nPoint);
      for (unsigned long i = 0; i < nPoint; ++i) mut[i] = SU2_TYPE::GetValue(solver_container[turbSol]->GetNodes()->GetmuT(i));
      B200_CHECK(su2b200_array_upload(flowCtx, SU2B200_ARR_EDDY_VISC, mut.data()));
    }
    if (sst) {
      std::vector<double> kw((size_t)nPoint * 2);
      for (unsigned long i = 0; i < nPoint; ++i)
        for (unsigned short v = 0; v < 2; ++v) kw[i * 2 + v] = SU2_TYPE::GetValue(solver_container[turbSol]->GetNodes()->GetSolution(i, v));
      B200_CHECK(su2b200_array_upload(flowCtx, SU2B200_ARR_TURB_SOLUTION, kw.data()));
    }
    /*--- CMultiGridIntegration::MultiGrid_Cycle at MGLEVEL 0 (CMultiGridIntegration.cpp:163-206), call by call ---*/
    unsigned long nonPhys = 0, iters = 0;
    double minDt = 0, maxDt = 0, res = 0;
    rms.assign(nVar, 0.0);
    std::vector<double> mx(nVar, 0.0);
    std::vector<unsigned long> idx(nVar, 0);
    B200_CHECK(su2b200_flow_preprocessing(flowCtx, 0, &nonPhys));
    B200_CHECK(su2b200_flow_set_old_solution(flowCtx));
    B200_CHECK(su2b200_flow_set_time_step(flowCtx, &minDt, &maxDt));
    B200_CHECK(su2b200_flow_space_integration(flowCtx));
    B200_CHECK(su2b200_flow_prepare_implicit(flowCtx, rms.data(), mx.data(), idx.data()));
    B200_CHECK(su2b200_flow_linear_solve(flowCtx, &iters, &res));
    B200_CHECK(su2b200_flow_complete_implicit(flowCtx));
    /*--- state out ---*/
    B200_CHECK(su2b200_array_download(flowCtx, SU2B200_ARR_SOLUTION, buf.data()));
    for (unsigned long i = 0; i < nPoint; ++i)
      for (unsigned short v = 0; v < nVar; ++v) flow->GetNodes()->SetSolution(i, v, buf[i * nVar + v]);
    /*--- the scalar solvers take their time step from the flow solver's (CScalarSolver::SetTime_Step) ---*/
    std::vector<double> dt(nPoint);
    B200_CHECK(su2b200_array_download(flowCtx, SU2B200_ARR_DELTA_TIME, dt.data()));
    for (unsigned long i = 0; i < nPoint; ++i) flow->GetNodes()->SetDelta_Time(i, dt[i]);
    /*--- ... and their far-field fluxes from the characteristic state of the flow far-field BC ---*/
    for (unsigned short m = 0; m < geometry->GetnMarker(); ++m) {
      if (config->GetMarker_All_KindBC(m) != FAR_FIELD) continue;
      const unsigned long nV = geometry->GetnVertex(m);
      const unsigned short nPrim = nDim + 9;
      std::vector<double> ch((size_t)nV * nPrim);
      B200_CHECK(su2b200_marker_charac_download(flowCtx, m, ch.data()));
      for (unsigned long v = 0; v < nV; ++v)
        for (unsigned short k = 0; k < nPrim; ++k) flow->SetCharacPrimVar(m, v, k, ch[v * nPrim + k]);
    }
    if (linRes) *linRes = res;
    return iters;
  }

 private:
  su2b200_ctx* flowCtx = nullptr;
  su2b200_ctx* linCtx = nullptr;
  const void* bound[2] = {nullptr, nullptr};

  static int Device() {
    const char* d = std::getenv("SU2B200_DEVICE");
    return d ? std::atoi(d) : 0;
  }
  static int SolverKind(const CConfig* c) {
    switch (c->GetKind_Linear_Solver()) {
      case BCGSTAB: return SU2B200_SOLVER_BCGSTAB;
      case RESTARTED_FGMRES: return SU2B200_SOLVER_RESTARTED_FGMRES;
      default: return SU2B200_SOLVER_FGMRES;
    }
  }
  static int PrecKind(const CConfig* c) {
    switch (c->GetKind_Linear_Solver_Prec()) {
      case JACOBI: return SU2B200_PREC_JACOBI;
      case LU_SGS: return SU2B200_PREC_LU_SGS;
      default: return SU2B200_PREC_ILU;
    }
  }

  template <class ScalarType>
  void BindLinearSystem(const CSysMatrix<ScalarType>& A, int which) {
    if (!linCtx) B200_CHECK(su2b200_create(Device(), &linCtx));
    if (bound[which] == (const void*)&A) return;
    B200_CHECK(su2b200_linsys_create(linCtx, which, (int)A.B200_nVar(), A.B200_nPoint(), A.B200_nPointDomain(),
                                     A.B200_RowPtr(), A.B200_ColInd()));
    bound[which] = (const void*)&A;
  }
  template <class ScalarType>
  void UploadMatrix(const CSysMatrix<ScalarType>& A, int which) {
    const size_t n = (size_t)A.B200_nnz() * A.B200_nVar() * A.B200_nVar();
    std::vector<double> vals(n);
    const ScalarType* m = A.B200_Values();
    for (size_t k = 0; k < n; ++k) vals[k] = (double)m[k];
    B200_CHECK(su2b200_linsys_set_matrix(linCtx, which, vals.data()));
  }

  /*--- geometry + markers + configuration of the flow problem, once (after CSolver construction) ---*/
  template <class Solver>
  void SetupFlow(CGeometry* g, Solver* flow, Solver* turb, CConfig* c) {
    B200_CHECK(su2b200_create(Device(), &flowCtx));
    const unsigned long nDim = g->GetnDim(), P = g->GetnPoint(), Pd = g->GetnPointDomain(), E = g->GetnEdge();
    std::vector<unsigned long> e0(E), e1(E);
    std::vector<double> nrm(E * nDim), coord(P * nDim), vol(P), wd(P);
    std::vector<unsigned short> nnb(P);
    std::vector<unsigned char> pb(P);
    for (unsigned long e = 0; e < E; ++e) {
      e0[e] = g->edges->GetNode(e, 0);
      e1[e] = g->edges->GetNode(e, 1);
      for (unsigned long d = 0; d < nDim; ++d) nrm[e * nDim + d] = SU2_TYPE::GetValue(g->edges->GetNormal(e)[d]);
    }
    std::vector<unsigned long> pptr(P + 1, 0), pnbr, pedge;
    for (unsigned long i = 0; i < P; ++i) {
      for (unsigned long d = 0; d < nDim; ++d) coord[i * nDim + d] = SU2_TYPE::GetValue(g->nodes->GetCoord(i, d));
      vol[i] = SU2_TYPE::GetValue(g->nodes->GetVolume(i));
      wd[i] = SU2_TYPE::GetValue(g->nodes->GetWall_Distance(i));
      nnb[i] = g->nodes->GetnNeighbor(i);
      pb[i] = g->nodes->GetPhysicalBoundary(i);
      for (unsigned short k = 0; k < g->nodes->GetnPoint(i); ++k) {
        pnbr.push_back(g->nodes->GetPoint(i, k));
        pedge.push_back(g->nodes->GetEdge(i, k));
      }
      pptr[i + 1] = pnbr.size();
    }
    const auto& pat = g->GetSparsePattern(ConnectivityType::FiniteVolume, 0);
    const auto& emap = g->GetEdgeToSparsePatternMap();
    std::vector<unsigned long> rp(pat.outerPtr(), pat.outerPtr() + P + 1), ci(pat.innerIdx(), pat.innerIdx() + pat.getNumNonZeros()),
        dp(pat.diagPtr(), pat.diagPtr() + P), ep(2 * E);
    for (unsigned long e = 0; e < E; ++e) { ep[2 * e] = emap(e, 0); ep[2 * e + 1] = emap(e, 1); }
    su2b200_geometry_desc d{};
    d.nDim = (int)nDim; d.nPoint = P; d.nPointDomain = Pd; d.nEdge = E;
    d.edge_node0 = e0.data(); d.edge_node1 = e1.data(); d.edge_normal = nrm.data(); d.coord = coord.data(); d.volume = vol.data();
    d.point_ptr = pptr.data(); d.point_nbr = pnbr.data(); d.point_edge = pedge.data();
    d.row_ptr = rp.data(); d.col_ind = ci.data(); d.dia_ptr = dp.data(); d.edge_ptr = ep.data();
    d.nNeighbor = nnb.data(); d.physical_boundary = pb.data(); d.wall_distance = wd.data();
    B200_CHECK(su2b200_geometry_upload(flowCtx, &d));
    for (unsigned short m = 0; m < g->GetnMarker(); ++m) {
      const unsigned long nV = g->GetnVertex(m);
      std::vector<unsigned long> node(nV), nb(nV);
      std::vector<double> nr(nV * nDim), un(nV * nDim);
      for (unsigned long v = 0; v < nV; ++v) {
        node[v] = g->vertex[m][v]->GetNode();
        nb[v] = g->vertex[m][v]->GetNormal_Neighbor();
        double a = 0;
        for (unsigned long k = 0; k < nDim; ++k) { nr[v * nDim + k] = SU2_TYPE::GetValue(g->vertex[m][v]->GetNormal()[k]); a += nr[v * nDim + k] * nr[v * nDim + k]; }
        a = std::sqrt(a);
        for (unsigned long k = 0; k < nDim; ++k) un[v * nDim + k] = nr[v * nDim + k] / a;
        if (m < g->symmetryNormals.size()) {  /*--- CGeometry::ComputeModifiedSymmetryNormals ---*/
          auto it = g->symmetryNormals[m].find(v);
          if (it != g->symmetryNormals[m].end())
            for (unsigned long k = 0; k < nDim; ++k) un[v * nDim + k] = SU2_TYPE::GetValue(it->second[k]);
        }
      }
      const std::string tag = c->GetMarker_All_TagBound(m);
      double hf = 0.0;
      if (c->GetMarker_All_KindBC(m) == HEAT_FLUX) hf = SU2_TYPE::GetValue(c->GetWall_HeatFlux(tag) / c->GetHeat_Flux_Ref());
      B200_CHECK(su2b200_marker_upload(flowCtx, m, c->GetMarker_All_KindBC(m), nV, node.data(), nr.data(), nb.data(), un.data(),
                                       c->GetMarker_All_Monitoring(m), hf));
    }
    /*--- the CConfig getters the path reads (non-dimensional values) ---*/
    su2b200_flow_config f{};
    auto val = [](su2double x) { return SU2_TYPE::GetValue(x); };
    f.gamma = val(c->GetGamma()); f.gas_constant = val(c->GetGas_ConstantND());
    f.viscous = c->GetViscous(); f.rans_sa = c->GetKind_Turb_Model() == TURB_MODEL::SA;
    f.sa_negative = c->GetSAParsedOptions().version == SA_OPTIONS::NEG;
    f.mu_ref = val(c->GetMu_RefND()); f.mu_temperature_ref = val(c->GetMu_Temperature_RefND());
    f.mu_sutherland = c->GetKind_ViscosityModel() == VISCOSITYMODEL::SUTHERLAND ? val(c->GetMu_SND()) : -1.0;
    f.mu_constant = val(c->GetMu_ConstantND());
    f.prandtl_lam = val(c->GetPrandtl_Lam()); f.prandtl_turb = val(c->GetPrandtl_Turb());
    f.density_inf = val(flow->GetDensity_Inf()); f.pressure_inf = val(flow->GetPressure_Inf());
    for (unsigned long k = 0; k < nDim; ++k) f.velocity_inf[k] = val(flow->GetVelocity_Inf(k));
    f.energy_inf = val(c->GetEnergy_FreeStreamND()); f.temperature_inf = val(c->GetTemperature_FreeStreamND());
    f.nu_tilde_inf = val(c->GetNuFactor_FreeStream() * c->GetViscosity_FreeStreamND() / c->GetDensity_FreeStreamND());
    f.conv_scheme = c->GetKind_ConvNumScheme_Flow() == SPACE_CENTERED ? SU2B200_CONV_JST : SU2B200_CONV_ROE;
    f.muscl = c->GetMUSCL_Flow(); f.muscl_kappa = val(c->GetMUSCL_Kappa_Flow());
    f.limiter = c->GetKind_SlopeLimit_Flow() == LIMITER::VENKATAKRISHNAN ? SU2B200_LIMITER_VENKATAKRISHNAN : SU2B200_LIMITER_NONE;
    f.venkat_coeff = val(c->GetVenkat_LimiterCoeff()); f.ref_elem_length = val(c->GetRefElemLength());
    auto grad = [](unsigned short k) {
      return k == LEAST_SQUARES ? SU2B200_GRAD_LEAST_SQUARES
                                : (k == WEIGHTED_LEAST_SQUARES ? SU2B200_GRAD_WEIGHTED_LEAST_SQUARES : SU2B200_GRAD_GREEN_GAUSS);
    };
    f.grad_method = grad(c->GetKind_Gradient_Method()); f.grad_method_recon = grad(c->GetKind_Gradient_Method_Recon());
    f.roe_kappa = val(c->GetRoe_Kappa()); f.entropy_fix = val(c->GetEntropyFix_Coeff());
    f.jst_kappa2 = val(c->GetKappa_2nd_Flow()); f.jst_kappa4 = val(c->GetKappa_4th_Flow());
    f.cent_jac_fix_factor = val(c->GetCent_Jac_Fix_Factor());
    f.cfl = val(c->GetCFL(MESH_0)); f.cfl_turb_ratio = val(c->GetCFLRedCoeff_Turb()); f.max_delta_time = val(c->GetMax_DeltaTime());
    f.max_update_flow = val(c->GetMaxUpdateFractionFlow()); f.max_update_sa = val(c->GetMaxUpdateFractionSA());
    f.linear_solver = f.turb_linear_solver = SolverKind(c); f.linear_prec = f.turb_linear_prec = PrecKind(c);
    f.linear_tol = f.turb_linear_tol = val(c->GetLinear_Solver_Error());
    f.linear_iter = f.turb_linear_iter = (int)c->GetLinear_Solver_Iter();
    f.linear_restart = (int)c->GetLinear_Solver_Restart_Frequency();
    f.aoa_deg = val(c->GetAoA()); f.aos_deg = val(c->GetAoS()); f.ref_area = val(c->GetRefArea()); f.ref_length = val(c->GetRefLength());
    f.ref_origin[0] = val(c->GetRefOriginMoment_X(0)); f.ref_origin[1] = val(c->GetRefOriginMoment_Y(0)); f.ref_origin[2] = val(c->GetRefOriginMoment_Z(0));
    f.assembly = SU2B200_ASSEMBLY_GATHER; f.prec_partitions = 1;
    f.turb_reducer = -1; f.mg_level = 0; f.lax_kappa1 = val(c->GetKappa_1st_Flow());
    if (f.rans_sa) {  /*--- free-stream eddy viscosity of the CTurbSASolver constructor (CTurbSASolver.cpp:100-134) ---*/
      const double mu = val(c->GetViscosity_FreeStreamND());
      const double chi = f.nu_tilde_inf / (mu / f.density_inf);
      f.eddy_visc_inf = f.density_inf * chi * chi * chi / (chi * chi * chi + 7.1 * 7.1 * 7.1) * f.nu_tilde_inf;
    }
    if (turb && c->GetKind_Turb_Model() == TURB_MODEL::SST) {
      /*--- free stream of the CTurbSSTSolver constructor (CTurbSSTSolver.cpp:118-147); energy_inf above already holds k
       *    (CEulerSolver::SetNondimensionalization adds Tke_FreeStreamND when the model needs it) ---*/
      f.rans_sst = 1;
      f.kine_inf = val(turb->GetTke_Inf()); f.omega_inf = val(turb->GetOmega_Inf());
      f.eddy_visc_inf = f.density_inf * f.kine_inf / f.omega_inf;
      f.sst_lower[0] = 1.0e-10; f.sst_lower[1] = 1.0e-4; f.sst_upper[0] = 1.0e10; f.sst_upper[1] = 1.0e15;
      f.max_update_sst = val(c->GetMaxUpdateFractionSST());
    }
    B200_CHECK(su2b200_flow_configure(flowCtx, &f));
    B200_CHECK(su2b200_flow_allocate(flowCtx));
  }
};

#endif  // HAVE_SU2B200
