// oracle_dump — TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Drives the UNMODIFIED reference (su2code/SU2, /root/reference) through its own classes and writes every
// intermediate array of one implicit iteration to a binary file, so that the CUDA kernels and the numpy
// restatement under oracle/ can be pinned against the reference itself.
// It includes the reference headers from where they lie and links libSU2core.a / libSU2Common.a built by
// oracle/build_ref.sh; no reference source is copied into this repository.
//
// Sequence mirrors CMultiGridIntegration::MultiGrid_Cycle at MGLEVEL=0 (CMultiGridIntegration.cpp:163-206),
// CIntegration::Space_Integration (CIntegration.cpp:37-201), CFVMFlowSolverBase::ImplicitEuler_Iteration
// (CFVMFlowSolverBase.inl:595-617) and CSingleGridIntegration::SingleGrid_Iteration for the SA solver.
//
// usage: oracle_dump <case.cfg> <out.bin> [nIter=1] [dumpIter=0,...]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <set>
#include <sstream>
#include <iostream>
#include <fstream>
#include <map>
#include <unordered_map>
#include <algorithm>
#include <memory>
#include <functional>
#include <iomanip>
#include <array>
#include <cmath>
#include <complex>
#include <list>
#include <limits>
#include <numeric>
#include <random>
#include <tuple>
#include <type_traits>
#include <cassert>
#include <climits>
#include <omp.h>

#define private public
#define protected public
#include "SU2_CFD/include/drivers/CSinglezoneDriver.hpp"
#include "SU2_CFD/include/solvers/CEulerSolver.hpp"
#include "SU2_CFD/include/solvers/CNSSolver.hpp"
#include "SU2_CFD/include/solvers/CTurbSASolver.hpp"
#include "SU2_CFD/include/solvers/CTurbSSTSolver.hpp"
#include "SU2_CFD/include/integration/CIntegration.hpp"
#include "SU2_CFD/include/integration/CMultiGridIntegration.hpp"
#include "SU2_CFD/include/variables/CEulerVariable.hpp"
#include "SU2_CFD/include/variables/CNSVariable.hpp"
#include "SU2_CFD/include/variables/CTurbSAVariable.hpp"
#include "SU2_CFD/include/output/COutput.hpp"
#undef private
#undef protected

struct Writer {
  FILE* f;
  explicit Writer(const char* path) { f = fopen(path, "wb"); if (!f) { perror(path); exit(2); } fwrite("SU2DUMP1", 1, 8, f); }
  ~Writer() { fclose(f); }
  // record: u32 name_len, name, u8 dtype (0=f64,1=i64,2=i8,3=u8), u32 ndim, u64 dims[], data
  void rec(const std::string& name, unsigned char dtype, const std::vector<unsigned long>& dims, const void* data, size_t esize) {
    unsigned nl = name.size();
    fwrite(&nl, 4, 1, f); fwrite(name.data(), 1, nl, f);
    fwrite(&dtype, 1, 1, f);
    unsigned nd = dims.size();
    fwrite(&nd, 4, 1, f);
    size_t n = 1;
    for (auto d : dims) { unsigned long long dd = d; fwrite(&dd, 8, 1, f); n *= d; }
    if (n) fwrite(data, esize, n, f);
  }
  void f64(const std::string& name, const std::vector<unsigned long>& dims, const double* d) { rec(name, 0, dims, d, 8); }
  void i64(const std::string& name, const std::vector<unsigned long>& dims, const long* d) { rec(name, 1, dims, d, 8); }
  void scalar(const std::string& name, double v) { f64(name, {1}, &v); }
  void iscalar(const std::string& name, long v) { i64(name, {1}, &v); }
  void text(const std::string& name, const std::string& s) { rec(name, 3, {s.size()}, s.data(), 1); }
};

template <class M>
static void dump_matrix(Writer& w, const std::string& name, const M& m, unsigned long rows, unsigned long cols) {
  std::vector<double> buf(rows * cols);
  for (unsigned long i = 0; i < rows; ++i)
    for (unsigned long j = 0; j < cols; ++j) buf[i * cols + j] = m(i, j);
  w.f64(name, {rows, cols}, buf.data());
}
template <class G>
static void dump_grad(Writer& w, const std::string& name, const G& g, unsigned long n, unsigned long nv, unsigned long nd) {
  std::vector<double> buf(n * nv * nd);
  for (unsigned long i = 0; i < n; ++i)
    for (unsigned long v = 0; v < nv; ++v)
      for (unsigned long d = 0; d < nd; ++d) buf[(i * nv + v) * nd + d] = g(i, v, d);
  w.f64(name, {n, nv, nd}, buf.data());
}
static void dump_sysvec(Writer& w, const std::string& name, const CSysVector<su2double>& v, unsigned long n, unsigned long nv) {
  std::vector<double> buf(n * nv);
  for (unsigned long i = 0; i < n * nv; ++i) buf[i] = v[i];
  w.f64(name, {n, nv}, buf.data());
}

class Dumper : public CSinglezoneDriver {
 public:
  Dumper(char* cfg) : CSinglezoneDriver(cfg, 1, SU2_MPI::GetComm()) {}

  void DumpGeometry(Writer& w) { DumpGeometryLevel(w, geometry_container[ZONE_0][INST_0][MESH_0], ""); }

  // geometry of one (multigrid) level; names prefixed with G ("" on the finest grid, "mg<k>_" on coarse level k)
  void DumpGeometryLevel(Writer& w, CGeometry* g, const std::string& G) {
    CConfig* cfg = config_container[ZONE_0];
    const unsigned long nDim = g->GetnDim(), P = g->GetnPoint(), Pd = g->GetnPointDomain(), E = g->GetnEdge();
    w.iscalar(G + "nDim", nDim); w.iscalar(G + "nPoint", P); w.iscalar(G + "nPointDomain", Pd); w.iscalar(G + "nEdge", E);
    std::vector<long> en(2 * E);
    std::vector<double> nrm(E * nDim);
    for (unsigned long e = 0; e < E; ++e) {
      en[e] = g->edges->GetNode(e, 0); en[E + e] = g->edges->GetNode(e, 1);
      for (unsigned long d = 0; d < nDim; ++d) nrm[e * nDim + d] = g->edges->GetNormal(e)[d];
    }
    w.i64(G + "edge_nodes", {2, E}, en.data());
    w.f64(G + "edge_normal", {E, nDim}, nrm.data());
    std::vector<double> coord(P * nDim), vol(P), wd(P);
    std::vector<long> gidx(P), nnb(P), pb(P);
    for (unsigned long i = 0; i < P; ++i) {
      for (unsigned long d = 0; d < nDim; ++d) coord[i * nDim + d] = g->nodes->GetCoord(i, d);
      vol[i] = g->nodes->GetVolume(i);
      wd[i] = g->nodes->GetWall_Distance(i);
      gidx[i] = g->nodes->GetGlobalIndex(i);
      nnb[i] = g->nodes->GetnNeighbor(i);
      pb[i] = g->nodes->GetPhysicalBoundary(i);
    }
    w.f64(G + "coord", {P, nDim}, coord.data()); w.f64(G + "volume", {P}, vol.data()); w.f64(G + "wall_distance", {P}, wd.data());
    w.i64(G + "global_index", {P}, gidx.data()); w.i64(G + "nNeighbor", {P}, nnb.data()); w.i64(G + "physical_boundary", {P}, pb.data());
    std::vector<long> pptr(P + 1, 0), pnbr, pedge;
    for (unsigned long i = 0; i < P; ++i) {
      for (unsigned short k = 0; k < g->nodes->GetnPoint(i); ++k) { pnbr.push_back(g->nodes->GetPoint(i, k)); pedge.push_back(g->nodes->GetEdge(i, k)); }
      pptr[i + 1] = pnbr.size();
    }
    w.i64(G + "point_ptr", {P + 1}, pptr.data()); w.i64(G + "point_nbr", {pnbr.size()}, pnbr.data()); w.i64(G + "point_edge", {pedge.size()}, pedge.data());
    const auto& pat = g->GetSparsePattern(ConnectivityType::FiniteVolume, 0);
    const auto& emap = g->GetEdgeToSparsePatternMap();
    const unsigned long nnz = pat.getNumNonZeros();
    std::vector<long> rp(P + 1), ci(nnz), dp(P), ep(2 * E);
    for (unsigned long i = 0; i <= P; ++i) rp[i] = pat.outerPtr()[i];
    for (unsigned long k = 0; k < nnz; ++k) ci[k] = pat.innerIdx()[k];
    for (unsigned long i = 0; i < P; ++i) dp[i] = pat.diagPtr()[i];
    for (unsigned long e = 0; e < E; ++e) { ep[2 * e] = emap(e, 0); ep[2 * e + 1] = emap(e, 1); }
    w.i64(G + "row_ptr", {P + 1}, rp.data()); w.i64(G + "col_ind", {nnz}, ci.data()); w.i64(G + "dia_ptr", {P}, dp.data()); w.i64(G + "edge_ptr", {E, 2}, ep.data());
    // markers
    const unsigned short nM = g->GetnMarker();
    w.iscalar(G + "nMarker", nM);
    for (unsigned short m = 0; m < nM; ++m) {
      const std::string tag = cfg->GetMarker_All_TagBound(m);
      const std::string pre = G + "marker" + std::to_string(m) + "_";
      w.text(pre + "tag", tag);
      w.iscalar(pre + "kind", cfg->GetMarker_All_KindBC(m));
      w.iscalar(pre + "monitoring", cfg->GetMarker_All_Monitoring(m));
      const unsigned long nV = g->GetnVertex(m);
      std::vector<long> node(nV), nb(nV);
      std::vector<double> nr(nV * nDim), un(nV * nDim);
      for (unsigned long v = 0; v < nV; ++v) {
        node[v] = g->vertex[m][v]->GetNode();
        nb[v] = g->vertex[m][v]->GetNormal_Neighbor();
        double a = 0;
        for (unsigned long d = 0; d < nDim; ++d) { nr[v * nDim + d] = g->vertex[m][v]->GetNormal()[d]; a += nr[v * nDim + d] * nr[v * nDim + d]; }
        a = std::sqrt(a);
        for (unsigned long d = 0; d < nDim; ++d) un[v * nDim + d] = nr[v * nDim + d] / a;
        if (m < g->symmetryNormals.size()) {
          auto it = g->symmetryNormals[m].find(v);
          if (it != g->symmetryNormals[m].end()) for (unsigned long d = 0; d < nDim; ++d) un[v * nDim + d] = it->second[d];
        }
      }
      w.i64(pre + "node", {nV}, node.data()); w.i64(pre + "normal_neighbor", {nV}, nb.data());
      w.f64(pre + "normal", {nV, nDim}, nr.data()); w.f64(pre + "sym_unit_normal", {nV, nDim}, un.data());
      double hf = 0.0;
      if (cfg->GetMarker_All_KindBC(m) == HEAT_FLUX) hf = cfg->GetWall_HeatFlux(tag) / cfg->GetHeat_Flux_Ref();
      w.scalar(pre + "heat_flux", hf);
    }
  }

  void DumpConfig(Writer& w) {
    CConfig* c = config_container[ZONE_0];
    auto* flow = solver_container[ZONE_0][INST_0][MESH_0][FLOW_SOL];
    const unsigned short nDim = geometry_container[ZONE_0][INST_0][MESH_0]->GetnDim();
    w.scalar("cfg_gamma", c->GetGamma()); w.scalar("cfg_gas_constant", c->GetGas_ConstantND());
    w.iscalar("cfg_viscous", c->GetViscous());
    w.iscalar("cfg_rans_sa", c->GetKind_Turb_Model() == TURB_MODEL::SA);
    w.iscalar("cfg_rans_sst", c->GetKind_Turb_Model() == TURB_MODEL::SST);
    if (c->GetKind_Turb_Model() == TURB_MODEL::SST) {
      auto* tsol = solver_container[ZONE_0][INST_0][MESH_0][TURB_SOL];
      w.iscalar("cfg_turb_reducer", static_cast<CTurbSSTSolver*>(tsol)->ReducerStrategy);
      w.scalar("cfg_kine_inf", tsol->GetTke_Inf());
      w.scalar("cfg_omega_inf", tsol->GetOmega_Inf());
      w.scalar("cfg_max_update_sst", c->GetMaxUpdateFractionSST());
      w.iscalar("cfg_sst_v2003", c->GetSSTParsedOptions().version == SST_OPTIONS::V2003);
      w.iscalar("cfg_sst_dll", c->GetSSTParsedOptions().dll);
    }
    w.iscalar("cfg_sa_negative", c->GetSAParsedOptions().version == SA_OPTIONS::NEG);
    if (c->GetKind_Turb_Model() == TURB_MODEL::SA)  // averaged SA diffusion Jacobians (CScalarSolver.inl:59)
      w.iscalar("cfg_turb_reducer", static_cast<CTurbSASolver*>(solver_container[ZONE_0][INST_0][MESH_0][TURB_SOL])->ReducerStrategy);
    w.scalar("cfg_mu_ref", c->GetMu_RefND()); w.scalar("cfg_mu_temperature_ref", c->GetMu_Temperature_RefND());
    w.scalar("cfg_mu_sutherland", c->GetMu_SND()); w.scalar("cfg_mu_constant", c->GetMu_ConstantND());
    w.iscalar("cfg_viscosity_model_sutherland", c->GetKind_ViscosityModel() == VISCOSITYMODEL::SUTHERLAND);
    w.scalar("cfg_prandtl_lam", c->GetPrandtl_Lam()); w.scalar("cfg_prandtl_turb", c->GetPrandtl_Turb());
    w.scalar("cfg_density_inf", flow->GetDensity_Inf()); w.scalar("cfg_pressure_inf", flow->GetPressure_Inf());
    double vel[3] = {0, 0, 0};
    for (unsigned short d = 0; d < nDim; ++d) vel[d] = flow->GetVelocity_Inf(d);
    w.f64("cfg_velocity_inf", {3}, vel);
    w.scalar("cfg_energy_inf", c->GetEnergy_FreeStreamND()); w.scalar("cfg_temperature_inf", c->GetTemperature_FreeStreamND());
    w.scalar("cfg_nu_tilde_inf", c->GetNuFactor_FreeStream() * c->GetViscosity_FreeStreamND() / c->GetDensity_FreeStreamND());
    w.iscalar("cfg_conv_centered", c->GetKind_ConvNumScheme_Flow() == SPACE_CENTERED);
    w.iscalar("cfg_muscl", c->GetMUSCL_Flow()); w.scalar("cfg_muscl_kappa", c->GetMUSCL_Kappa_Flow());
    w.iscalar("cfg_limiter", (long)c->GetKind_SlopeLimit_Flow());
    w.scalar("cfg_venkat_coeff", c->GetVenkat_LimiterCoeff()); w.scalar("cfg_ref_elem_length", c->GetRefElemLength());
    w.iscalar("cfg_grad_method", c->GetKind_Gradient_Method()); w.iscalar("cfg_grad_method_recon", c->GetKind_Gradient_Method_Recon());
    w.scalar("cfg_roe_kappa", c->GetRoe_Kappa()); w.scalar("cfg_entropy_fix", c->GetEntropyFix_Coeff());
    w.scalar("cfg_jst_kappa2", c->GetKappa_2nd_Flow()); w.scalar("cfg_jst_kappa4", c->GetKappa_4th_Flow());
    w.scalar("cfg_cent_jac_fix_factor", c->GetCent_Jac_Fix_Factor());
    w.scalar("cfg_cfl", c->GetCFL(MESH_0)); w.scalar("cfg_cfl_turb_ratio", c->GetCFLRedCoeff_Turb());
    w.scalar("cfg_max_delta_time", c->GetMax_DeltaTime());
    w.iscalar("cfg_cfl_adapt", c->GetCFL_Adapt() == YES);
    {
      double ap[6];
      for (int k = 0; k < 6; ++k) ap[k] = c->GetCFL_AdaptParam(k);
      w.f64("cfg_cfl_adapt_param", {6}, ap);
    }
    w.iscalar("cfg_n_inner_iter", c->GetnInner_Iter());
    w.scalar("cfg_max_update_flow", c->GetMaxUpdateFractionFlow()); w.scalar("cfg_max_update_sa", c->GetMaxUpdateFractionSA());
    w.iscalar("cfg_linear_solver", (long)c->GetKind_Linear_Solver()); w.iscalar("cfg_linear_prec", (long)c->GetKind_Linear_Solver_Prec());
    w.scalar("cfg_linear_tol", c->GetLinear_Solver_Error()); w.iscalar("cfg_linear_iter", c->GetLinear_Solver_Iter());
    w.iscalar("cfg_linear_restart", c->GetLinear_Solver_Restart_Frequency());
    w.scalar("cfg_aoa", c->GetAoA()); w.scalar("cfg_aos", c->GetAoS());
    w.scalar("cfg_ref_area", c->GetRefArea()); w.scalar("cfg_ref_length", c->GetRefLength());
    double org[3] = {c->GetRefOriginMoment_X(0), c->GetRefOriginMoment_Y(0), c->GetRefOriginMoment_Z(0)};
    w.f64("cfg_ref_origin", {3}, org);
    w.iscalar("enum_GREEN_GAUSS", GREEN_GAUSS); w.iscalar("enum_LEAST_SQUARES", LEAST_SQUARES); w.iscalar("enum_WEIGHTED_LEAST_SQUARES", WEIGHTED_LEAST_SQUARES);
    w.iscalar("enum_LIMITER_NONE", (long)LIMITER::NONE); w.iscalar("enum_LIMITER_VENKATAKRISHNAN", (long)LIMITER::VENKATAKRISHNAN);
    w.iscalar("enum_FGMRES", FGMRES); w.iscalar("enum_BCGSTAB", BCGSTAB); w.iscalar("enum_RESTARTED_FGMRES", RESTARTED_FGMRES);
    w.iscalar("enum_ILU", ILU); w.iscalar("enum_LU_SGS", LU_SGS); w.iscalar("enum_JACOBI", JACOBI);
  }

  void DumpFlowState(Writer& w, const std::string& pre) {
    CGeometry* g = geometry_container[ZONE_0][INST_0][MESH_0];
    auto* flow = static_cast<CEulerSolver*>(solver_container[ZONE_0][INST_0][MESH_0][FLOW_SOL]);
    auto* nodes = flow->nodes;
    const unsigned long P = g->GetnPoint(), nDim = g->GetnDim(), nVar = flow->nVar, nPrim = flow->nPrimVar;
    dump_matrix(w, pre + "solution", nodes->GetSolution(), P, nVar);
    dump_matrix(w, pre + "solution_old", nodes->Solution_Old, P, nVar);
    dump_matrix(w, pre + "primitive", nodes->GetPrimitive(), P, nPrim);
    // gradient containers hold nPrimVarGrad variables (nDim+1 for centred schemes, CEulerVariable.cpp:31-42)
    const unsigned long ng = nodes->GetGradient_Primitive().rows();
    CConfig* c = config_container[ZONE_0];
    const bool centered = c->GetKind_ConvNumScheme_Flow() == SPACE_CENTERED;
    if (c->GetMUSCL_Flow() && !centered) {
      dump_grad(w, pre + "grad_recon", nodes->GetGradient_Reconstruction(), P, ng, nDim);
      if (c->GetKind_SlopeLimit_Flow() != LIMITER::NONE) dump_matrix(w, pre + "limiter", nodes->GetLimiter_Primitive(), P, ng);
    }
    if (c->GetViscous()) dump_grad(w, pre + "grad_prim", nodes->GetGradient_Primitive(), P, ng, nDim);
    if (c->GetKind_ConvNumScheme_Flow() == SPACE_CENTERED) {
      std::vector<double> lam(P), sen(P);
      for (unsigned long i = 0; i < P; ++i) { lam[i] = nodes->GetLambda(i); sen[i] = nodes->GetSensor(i); }
      w.f64(pre + "lambda", {P}, lam.data()); w.f64(pre + "sensor", {P}, sen.data());
      dump_matrix(w, pre + "und_lapl", nodes->GetUndivided_Laplacian(), P, nVar);
    }
    if (c->GetViscous()) {
      std::vector<double> vort(P * 3), sm(P);
      for (unsigned long i = 0; i < P; ++i) { for (int k = 0; k < 3; ++k) vort[i * 3 + k] = nodes->GetVorticity(i)[k]; sm[i] = nodes->GetStrainMag(i); }
      w.f64(pre + "vorticity", {P, 3}, vort.data()); w.f64(pre + "strain_mag", {P}, sm.data());
    }
  }

  void DumpSystem(Writer& w, const std::string& pre, CSolver* s, bool with_matrix) {
    CGeometry* g = geometry_container[ZONE_0][INST_0][MESH_0];
    const unsigned long P = g->GetnPoint(), nVar = s->GetnVar();
    dump_sysvec(w, pre + "linsysres", s->LinSysRes, P, nVar);
    if (with_matrix) {
      const unsigned long nnz = s->Jacobian.nnz;
      w.f64(pre + "jacobian", {nnz, nVar, nVar}, s->Jacobian.matrix);
    }
  }

  // ---- FAS multigrid (MGLEVEL > 0): CMultiGridIntegration::MultiGrid_Cycle (CMultiGridIntegration.cpp:129-300) replayed
  //      call by call through the reference's own member functions, with the state of every level written after each
  //      transfer step of the dumped iterations ("it<i>_ev<n>_<what>_L<level>", n = running event number of the cycle)
  int mgEvent = 0;
  void MGDumpLevel(Writer& w, const std::string& pre, const char* what, unsigned short iMesh, bool sol, bool res, bool trunc) {
    CGeometry* g = geometry_container[ZONE_0][INST_0][iMesh];
    CSolver* s = solver_container[ZONE_0][INST_0][iMesh][FLOW_SOL];
    const unsigned long P = g->GetnPoint(), nVar = s->GetnVar();
    const std::string name = pre + "ev" + std::to_string(mgEvent++) + "_" + what + "_L" + std::to_string(iMesh);
    if (sol) dump_matrix(w, name + "_solution", s->GetNodes()->GetSolution(), P, nVar);
    if (res) dump_sysvec(w, name + "_linsysres", s->LinSysRes, P, nVar);
    if (trunc) {
      std::vector<double> t(P * nVar);
      for (unsigned long i = 0; i < P; ++i) for (unsigned long v = 0; v < nVar; ++v) t[i * nVar + v] = s->GetNodes()->GetResTruncError(i)[v];
      w.f64(name + "_trunc", {P, nVar}, t.data());
    }
  }
  void MGCycle(Writer& w, bool D, const std::string& pre, unsigned short iMesh, unsigned short RecursiveParam) {
    CConfig* config = config_container[ZONE_0];
    auto* mg = static_cast<CMultiGridIntegration*>(integration_container[ZONE_0][INST_0][FLOW_SOL]);
    CGeometry* gf = geometry_container[ZONE_0][INST_0][iMesh];
    CSolver** scf = solver_container[ZONE_0][INST_0][iMesh];
    CSolver* sf = scf[FLOW_SOL];
    CNumerics** nf = numerics_container[ZONE_0][INST_0][iMesh][FLOW_SOL];
    auto smooth = [&](unsigned short n) {
      for (unsigned short k = 0; k < n; ++k) {
        sf->Preprocessing(gf, scf, config, iMesh, 0, RUNTIME_FLOW_SYS, false);
        sf->Set_OldSolution();
        sf->SetTime_Step(gf, scf, config, iMesh, config->GetTimeIter());
        mg->Space_Integration(gf, scf, nf, config, iMesh, 0, RUNTIME_FLOW_SYS);
        mg->Time_Integration(gf, scf, config, 0, RUNTIME_FLOW_SYS);
        sf->Postprocessing(gf, scf, config, iMesh);
      }
    };
    smooth(config->GetMG_PreSmooth(iMesh));
    if (D) MGDumpLevel(w, pre, "presmooth", iMesh, true, false, false);
    if (iMesh < config->GetnMGLevels()) {
      CGeometry* gc = geometry_container[ZONE_0][INST_0][iMesh + 1];
      CSolver** scc = solver_container[ZONE_0][INST_0][iMesh + 1];
      CSolver* sc = scc[FLOW_SOL];
      CNumerics** nc = numerics_container[ZONE_0][INST_0][iMesh + 1][FLOW_SOL];
      config->SetKind_TimeIntScheme(EULER_EXPLICIT);
      sf->Preprocessing(gf, scf, config, iMesh, NO_RK_ITER, RUNTIME_FLOW_SYS, false);
      mg->Space_Integration(gf, scf, nf, config, iMesh, NO_RK_ITER, RUNTIME_FLOW_SYS);
      mg->SetResidual_Term(gf, sf);
      if (D) MGDumpLevel(w, pre, "fineres", iMesh, false, true, false);
      mg->SetRestricted_Solution(RUNTIME_FLOW_SYS, sf, sc, gf, gc, config);
      sc->Preprocessing(gc, scc, config, iMesh + 1, NO_RK_ITER, RUNTIME_FLOW_SYS, false);
      mg->Space_Integration(gc, scc, nc, config, iMesh + 1, NO_RK_ITER, RUNTIME_FLOW_SYS);
      if (D) MGDumpLevel(w, pre, "coarseres", iMesh + 1, true, true, false);
      mg->SetForcing_Term(sf, sc, gf, gc, config, iMesh + 1);
      if (D) MGDumpLevel(w, pre, "forcing", iMesh + 1, false, false, true);
      config->SetKind_TimeIntScheme(EULER_IMPLICIT);
      for (unsigned short imu = 0; imu <= RecursiveParam; imu++) {
        unsigned short next = RecursiveParam;
        if (iMesh == config->GetnMGLevels() - 2) next = 0;
        MGCycle(w, D, pre, iMesh + 1, next);
      }
      mg->GetProlongated_Correction(RUNTIME_FLOW_SYS, sf, sc, gf, gc, config);
      mg->SmoothProlongated_Correction(RUNTIME_FLOW_SYS, sf, gf, config->GetMG_CorrecSmooth(iMesh), 1.25, config);
      if (D) MGDumpLevel(w, pre, "correction", iMesh, false, true, false);
      mg->SetProlongated_Correction(sf, gf, config, iMesh);
      if (D) MGDumpLevel(w, pre, "corrected", iMesh, true, false, false);
      smooth(config->GetMG_PostSmooth(iMesh));
    }
  }
  void RunMG(const char* out, int nIter, const std::set<int>& dumpIters) {
    Writer w(out);
    CConfig* config = config_container[ZONE_0];
    CGeometry* g = geometry_container[ZONE_0][INST_0][MESH_0];
    CSolver** solvers = solver_container[ZONE_0][INST_0][MESH_0];
    auto* flow = static_cast<CEulerSolver*>(solvers[FLOW_SOL]);
    auto* mg = static_cast<CMultiGridIntegration*>(integration_container[ZONE_0][INST_0][FLOW_SOL]);
    const unsigned long nVar = flow->nVar;
    const unsigned short nLev = config->GetnMGLevels();
    Preprocess(0);
    DumpGeometry(w);
    DumpConfig(w);
    w.iscalar("nIter", nIter);
    w.iscalar("cfg_mg_levels", nLev);
    w.iscalar("cfg_mg_cycle", config->GetMGCycle());
    w.scalar("cfg_mg_damp_restriction", config->GetDamp_Res_Restric());
    w.scalar("cfg_mg_damp_prolongation", config->GetDamp_Correc_Prolong());
    w.scalar("cfg_lax_kappa1", config->GetKappa_1st_Flow());
    std::vector<long> pre(nLev + 1), post(nLev + 1), corr(nLev + 1);
    std::vector<double> cfl(nLev + 1);
    for (unsigned short l = 0; l <= nLev; ++l) { pre[l] = config->GetMG_PreSmooth(l); post[l] = config->GetMG_PostSmooth(l); corr[l] = config->GetMG_CorrecSmooth(l); cfl[l] = config->GetCFL(l); }
    w.i64("cfg_mg_pre_smooth", {(unsigned long)nLev + 1}, pre.data()); w.i64("cfg_mg_post_smooth", {(unsigned long)nLev + 1}, post.data());
    w.i64("cfg_mg_correction_smooth", {(unsigned long)nLev + 1}, corr.data()); w.f64("cfg_mg_cfl", {(unsigned long)nLev + 1}, cfl.data());
    for (unsigned short l = 1; l <= nLev; ++l) {
      CGeometry* gc = geometry_container[ZONE_0][INST_0][l];
      const std::string G = "mg" + std::to_string(l) + "_";
      DumpGeometryLevel(w, gc, G);
      // agglomeration: children (points of level l-1) of every coarse point, CSR
      const unsigned long Pc = gc->GetnPoint();
      std::vector<long> cptr(Pc + 1, 0), child;
      for (unsigned long i = 0; i < Pc; ++i) {
        for (unsigned short k = 0; k < gc->nodes->GetnChildren_CV(i); ++k) child.push_back(gc->nodes->GetChildren_CV(i, k));
        cptr[i + 1] = child.size();
      }
      w.i64(G + "children_ptr", {Pc + 1}, cptr.data()); w.i64(G + "children", {child.size()}, child.data());
    }
    std::vector<double> hist;
    for (int it = 0; it < nIter; ++it) {
      config->SetInnerIter(it);
      const bool D = dumpIters.count(it) > 0;
      const std::string p = "it" + std::to_string(it) + "_";
      config->SetGlobalParam(config->GetKind_Solver(), RUNTIME_FLOW_SYS);
      if (D) dump_matrix(w, p + "pre_solution", flow->nodes->GetSolution(), g->GetnPoint(), nVar);
      mgEvent = 0;
      unsigned short rec = config->GetMGCycle();
      if (rec == FULLMG_CYCLE) rec = V_CYCLE;
      MGCycle(w, D, p, MESH_0, rec);
      flow->Preprocessing(g, solvers, config, MESH_0, NO_RK_ITER, RUNTIME_FLOW_SYS, true);
      flow->Pressure_Forces(g, config);
      flow->Momentum_Forces(g, config);
      flow->Friction_Forces(g, config);
      if (D) dump_matrix(w, p + "post_solution", flow->nodes->GetSolution(), g->GetnPoint(), nVar);
      for (unsigned v = 0; v < 5; ++v) hist.push_back(v < nVar ? flow->GetRes_RMS(v) : 0.0);
      hist.push_back(0.0);
      hist.push_back(flow->GetTotal_CL()); hist.push_back(flow->GetTotal_CD());
      hist.push_back((double)flow->GetIterLinSolver()); hist.push_back(flow->GetResLinSolver());
      hist.push_back(0.0); hist.push_back(0.0);
      fprintf(stderr, "[oracle_dump mg] it %d rms0 %.9e CL %.9f CD %.9f linit %lu\n", it, flow->GetRes_RMS(0), flow->GetTotal_CL(), flow->GetTotal_CD(), flow->GetIterLinSolver());
    }
    (void)mg;
    w.f64("history", {(unsigned long)nIter, 12}, hist.data());
    dump_matrix(w, "final_solution", flow->nodes->GetSolution(), g->GetnPoint(), nVar);
  }

  void Run(const char* out, int nIter, const std::set<int>& dumpIters) {
    Writer w(out);
    CGeometry* g = geometry_container[ZONE_0][INST_0][MESH_0];
    CConfig* config = config_container[ZONE_0];
    CSolver** solvers = solver_container[ZONE_0][INST_0][MESH_0];
    auto* flow = static_cast<CEulerSolver*>(solvers[FLOW_SOL]);
    CNumerics** numerics = numerics_container[ZONE_0][INST_0][MESH_0][FLOW_SOL];
    const bool rans = config->GetKind_Turb_Model() != TURB_MODEL::NONE;
    const unsigned long P = g->GetnPoint(), Pd = g->GetnPointDomain(), nVar = flow->nVar;
    Preprocess(0);
    DumpGeometry(w);
    DumpConfig(w);
    w.iscalar("nIter", nIter);
    std::vector<double> hist;
    for (int it = 0; it < nIter; ++it) {
      config->SetInnerIter(it);
      const bool D = dumpIters.count(it) > 0;
      const std::string pre = "it" + std::to_string(it) + "_";
      // ---- flow: MultiGrid_Cycle on MESH_0 ----
      config->SetGlobalParam(config->GetKind_Solver(), RUNTIME_FLOW_SYS);  // CFluidIteration.cpp:74-75
      flow->Preprocessing(g, solvers, config, MESH_0, NO_RK_ITER, RUNTIME_FLOW_SYS, false);
      if (D) DumpFlowState(w, pre + "pre_");
      flow->Set_OldSolution();
      flow->SetTime_Step(g, solvers, config, MESH_0, config->GetTimeIter());
      if (D) {
        std::vector<double> dt(P), li(P), lv(P);
        for (unsigned long i = 0; i < P; ++i) { dt[i] = flow->nodes->GetDelta_Time(i); li[i] = flow->nodes->GetMax_Lambda_Inv(i); lv[i] = config->GetViscous() ? flow->nodes->GetMax_Lambda_Visc(i) : 0.0; }
        w.f64(pre + "delta_time", {P}, dt.data()); w.f64(pre + "max_lambda_inv", {P}, li.data()); w.f64(pre + "max_lambda_visc", {P}, lv.data());
        w.scalar(pre + "min_dt", flow->Min_Delta_Time); w.scalar(pre + "max_dt", flow->Max_Delta_Time);
      }
      // Space_Integration, split so the pre-BC state can be dumped
      if (config->GetKind_ConvNumScheme() == SPACE_CENTERED) flow->Centered_Residual(g, solvers, numerics, config, MESH_0, NO_RK_ITER);
      else flow->Upwind_Residual(g, solvers, numerics, config, MESH_0);
      flow->Viscous_Residual(g, solvers, numerics, config, MESH_0, NO_RK_ITER);
      flow->Source_Residual(g, solvers, numerics, config, MESH_0);
      if (D) DumpSystem(w, pre + "edge_", flow, true);
      CNumerics* conv_b = numerics[CONV_BOUND_TERM];
      CNumerics* visc_b = numerics[VISC_BOUND_TERM];
      for (unsigned short m = 0; m < config->GetnMarker_All(); ++m)
        if (config->GetMarker_All_KindBC(m) == FAR_FIELD) flow->BC_Far_Field(g, solvers, conv_b, visc_b, config, m);
      if (D) DumpSystem(w, pre + "far_", flow, true);
      for (unsigned short m = 0; m < config->GetnMarker_All(); ++m)
        if (config->GetMarker_All_KindBC(m) == HEAT_FLUX) flow->BC_HeatFlux_Wall(g, solvers, conv_b, visc_b, config, m);
      for (unsigned short m = 0; m < config->GetnMarker_All(); ++m) {
        if (config->GetMarker_All_KindBC(m) == SYMMETRY_PLANE) flow->BC_Sym_Plane(g, solvers, conv_b, visc_b, config, m);
        else if (config->GetMarker_All_KindBC(m) == EULER_WALL) flow->BC_Euler_Wall(g, solvers, conv_b, visc_b, config, m);
      }
      if (D) {
        DumpSystem(w, pre + "bc_", flow, true);
        dump_matrix(w, pre + "bc_solution_old", flow->nodes->Solution_Old, P, nVar);
        std::vector<long> cnt(g->GetnEdge());
        for (unsigned long e = 0; e < g->GetnEdge(); ++e) cnt[e] = flow->nodes->NonPhysicalEdgeCounter[e];
        w.i64(pre + "edge_counter", {g->GetnEdge()}, cnt.data());
      }
      // Time_Integration -> ImplicitEuler_Iteration
      flow->PrepareImplicitIteration(g, nullptr, config);
      for (unsigned long i = Pd; i < P; ++i) { flow->LinSysRes.SetBlock_Zero(i); flow->LinSysSol.SetBlock_Zero(i); }
      if (D) {
        DumpSystem(w, pre + "sys_", flow, true);
        std::vector<double> rms(nVar);
        for (unsigned long v = 0; v < nVar; ++v) rms[v] = flow->GetRes_RMS(v);
        w.f64(pre + "res_rms", {nVar}, rms.data());
      }
      auto iter = flow->System.Solve(flow->Jacobian, flow->LinSysRes, flow->LinSysSol, g, config);
      flow->SetIterLinSolver(iter);
      flow->SetResLinSolver(flow->System.GetResidual());
      if (D) {
        dump_sysvec(w, pre + "linsyssol", flow->LinSysSol, P, nVar);
        w.iscalar(pre + "linsol_iter", iter); w.scalar(pre + "linsol_res", flow->System.GetResidual());
        if (config->GetKind_Linear_Solver_Prec() == ILU) {
          w.f64(pre + "ilu", {flow->Jacobian.nnz_ilu, nVar, nVar}, flow->Jacobian.ILU_matrix);
          w.f64(pre + "invM", {Pd, nVar, nVar}, flow->Jacobian.invM);
        }
      }
      flow->CompleteImplicitIteration(g, nullptr, config);
      flow->Postprocessing(g, solvers, config, MESH_0);
      if (D) {
        dump_matrix(w, pre + "post_solution", flow->nodes->GetSolution(), P, nVar);
        std::vector<double> ur(P);
        for (unsigned long i = 0; i < P; ++i) ur[i] = flow->nodes->GetUnderRelaxation(i);
        w.f64(pre + "under_relax", {P}, ur.data());
      }
      // MultiGrid_Iteration tail
      flow->Preprocessing(g, solvers, config, MESH_0, NO_RK_ITER, RUNTIME_FLOW_SYS, true);
      flow->Pressure_Forces(g, config);
      flow->Momentum_Forces(g, config);
      flow->Friction_Forces(g, config);
      double rms_turb = 0.0, rms_turb2 = 0.0;
      long turb_iter = 0;
      double turb_linres = 0.0;
      if (rans) {
        auto* turb = solvers[TURB_SOL];
        CNumerics** tnum = numerics_container[ZONE_0][INST_0][MESH_0][TURB_SOL];
        const std::string tp = pre + "turb_";
        config->SetGlobalParam(config->GetKind_Solver(), RUNTIME_TURB_SYS);  // CFluidIteration.cpp:96
        turb->Preprocessing(g, solvers, config, MESH_0, 0, RUNTIME_TURB_SYS, false);
        const unsigned long tv = turb->GetnVar();
        auto dump_sst = [&](const std::string& name) {
          if (tv != 2) return;
          std::vector<double> f1(P), f2(P), cd(P), mut(P);
          for (unsigned long i = 0; i < P; ++i) {
            f1[i] = turb->GetNodes()->GetF1blending(i); f2[i] = turb->GetNodes()->GetF2blending(i);
            cd[i] = turb->GetNodes()->GetCrossDiff(i); mut[i] = turb->GetNodes()->GetmuT(i);
          }
          w.f64(name + "F1", {P}, f1.data()); w.f64(name + "F2", {P}, f2.data());
          w.f64(name + "CDkw", {P}, cd.data()); w.f64(name + "mut", {P}, mut.data());
        };
        if (D) {
          dump_matrix(w, tp + "pre_solution", turb->GetNodes()->GetSolution(), P, tv);
          dump_grad(w, tp + "pre_gradient", turb->GetNodes()->GetGradient(), P, tv, g->GetnDim());
          dump_sst(tp + "pre_");
          DumpFlowState(w, pre + "flowout_");
        }
        turb->Set_OldSolution();
        turb->SetTime_Step(g, solvers, config, MESH_0, config->GetTimeIter());
        if (D) {
          std::vector<double> dt(P);
          for (unsigned long i = 0; i < P; ++i) dt[i] = turb->GetNodes()->GetDelta_Time(i);
          w.f64(tp + "delta_time", {P}, dt.data());
        }
        turb->Upwind_Residual(g, solvers, tnum, config, MESH_0);
        if (D) DumpSystem(w, tp + "conv_", turb, true);
        turb->Viscous_Residual(g, solvers, tnum, config, MESH_0, 0);
        if (D) DumpSystem(w, tp + "edge_", turb, true);
        turb->Source_Residual(g, solvers, tnum, config, MESH_0);
        if (D) DumpSystem(w, tp + "src_", turb, true);
        CNumerics* tconv_b = tnum[CONV_BOUND_TERM];
        CNumerics* tvisc_b = tnum[VISC_BOUND_TERM];
        for (unsigned short m = 0; m < config->GetnMarker_All(); ++m)
          if (config->GetMarker_All_KindBC(m) == FAR_FIELD) turb->BC_Far_Field(g, solvers, tconv_b, tvisc_b, config, m);
        for (unsigned short m = 0; m < config->GetnMarker_All(); ++m)
          if (config->GetMarker_All_KindBC(m) == HEAT_FLUX) turb->BC_HeatFlux_Wall(g, solvers, tconv_b, tvisc_b, config, m);
        for (unsigned short m = 0; m < config->GetnMarker_All(); ++m) {
          if (config->GetMarker_All_KindBC(m) == SYMMETRY_PLANE) turb->BC_Sym_Plane(g, solvers, tconv_b, tvisc_b, config, m);
          else if (config->GetMarker_All_KindBC(m) == EULER_WALL) turb->BC_Euler_Wall(g, solvers, tconv_b, tvisc_b, config, m);
        }
        if (D) DumpSystem(w, tp + "bc_", turb, true);
        turb->ImplicitEuler_Iteration(g, solvers, config);
        turb->Postprocessing(g, solvers, config, MESH_0);
        if (D) {
          dump_sysvec(w, tp + "linsyssol", turb->LinSysSol, P, tv);
          dump_matrix(w, tp + "post_solution", turb->GetNodes()->GetSolution(), P, tv);
          dump_sst(tp + "post_");
          dump_grad(w, tp + "post_gradient", turb->GetNodes()->GetGradient(), P, tv, g->GetnDim());
          std::vector<double> mut(P);
          for (unsigned long i = 0; i < P; ++i) mut[i] = turb->GetNodes()->GetmuT(i);
          w.f64(tp + "post_mut", {P}, mut.data());
        }
        rms_turb = turb->GetRes_RMS(0);
        if (tv > 1) rms_turb2 = turb->GetRes_RMS(1);
        turb_iter = turb->GetIterLinSolver();
        turb_linres = turb->GetResLinSolver();
      }
      // CFluidIteration::Iterate (CFluidIteration.cpp:131-137)
      if (config->GetCFL_Adapt() == YES) flow->AdaptCFLNumber(geometry_container[ZONE_0][INST_0], solver_container[ZONE_0][INST_0], config);
      if (D && config->GetCFL_Adapt() == YES) {
        std::vector<double> lc(P);
        for (unsigned long i = 0; i < P; ++i) lc[i] = flow->nodes->GetLocalCFL(i);
        w.f64(pre + "local_cfl", {P}, lc.data());
      }
      // history row: rms[nVar] (5 slots), rms_turb, CL, CD, lin iters/res flow, lin iters/res turb
      for (unsigned v = 0; v < 5; ++v) hist.push_back(v < nVar ? flow->GetRes_RMS(v) : 0.0);
      hist.push_back(rms_turb);
      hist.push_back(flow->GetTotal_CL()); hist.push_back(flow->GetTotal_CD());
      hist.push_back((double)flow->GetIterLinSolver()); hist.push_back(flow->GetResLinSolver());
      hist.push_back((double)turb_iter); hist.push_back(turb_linres);
      hist.push_back(rms_turb2);  // second turbulence variable (SST: omega), 0 otherwise
      fprintf(stderr, "[oracle_dump] it %d rms0 %.9e CL %.9f CD %.9f linit %lu\n", it, flow->GetRes_RMS(0), flow->GetTotal_CL(), flow->GetTotal_CD(), flow->GetIterLinSolver());
    }
    w.f64("history", {(unsigned long)nIter, 13}, hist.data());
    dump_matrix(w, "final_solution", flow->nodes->GetSolution(), P, nVar);
  }
};

int main(int argc, char** argv) {
  if (argc < 3) { fprintf(stderr, "usage: %s case.cfg out.bin [nIter] [dumpIter,...]\n", argv[0]); return 1; }
  omp_set_num_threads(1);
  SU2_MPI::Init(&argc, &argv);
  int nIter = argc > 3 ? atoi(argv[3]) : 1;
  std::set<int> dumpIters;
  if (argc > 4) { std::stringstream ss(argv[4]); std::string tok; while (std::getline(ss, tok, ',')) dumpIters.insert(atoi(tok.c_str())); }
  else dumpIters.insert(0);
  {
    std::streambuf* orig = std::cout.rdbuf();
    std::ofstream devnull("/dev/null");
    if (!getenv("ORACLE_DUMP_VERBOSE")) std::cout.rdbuf(devnull.rdbuf());
    Dumper d(argv[1]);
    if (d.config_container[ZONE_0]->GetnMGLevels() > 0) d.RunMG(argv[2], nIter, dumpIters);
    else d.Run(argv[2], nIter, dumpIters);
    std::cout.rdbuf(orig);
  }
  return 0;
}
