// extern "C" entry points that mirror the CSolver virtuals and CSysMatrix/CSysSolve calls.
#include "internal.hpp"
#include "kernels.hpp"
#include <algorithm>
#include <cmath>

namespace su2b200 {
void bench_edge_kernel(su2b200_ctx* c, int reps, float* ms_avg);
}
using namespace su2b200;

#define API_BEGIN try {
#define API_END                                   \
  return 0;                                       \
  }                                               \
  catch (const std::exception& e) {               \
    su2b200::set_error(e.what());                 \
    return 1;                                     \
  }                                               \
  catch (...) {                                   \
    su2b200::set_error("unknown C++ exception");  \
    return 2;                                     \
  }

static void require_flow(su2b200_ctx* c) {
  if (!c) throw Error("null context");
  SU2B_CUDA(cudaSetDevice(c->device));
  SU2B_REQUIRE(c->fs.allocated, "call su2b200_flow_allocate first");
}

// CEulerSolver::CommonPreprocessing + Preprocessing (CEulerSolver.cpp:1596-1698) and CNSSolver::Preprocessing
// (CNSSolver.cpp:70-126), MGLEVEL = 0.
static void flow_preprocessing(su2b200_ctx* c, bool output, unsigned long* nonphys) {
  auto& s = c->fs;
  const auto& cfg = c->cfg;
  const bool center = cfg.conv_scheme == SU2B200_CONV_JST;
  const bool muscl = cfg.muscl != 0 && cfg.mg_level == 0;  // MUSCL only on MESH_0 (CEulerSolver.cpp:1662)
  const bool limiter = cfg.limiter != SU2B200_LIMITER_NONE;
  const int nd = c->geo.nDim;
  flow_primitives(c);
  if (nonphys) {
    SU2B_CUDA(cudaMemcpyAsync(s.h_counters, s.counters.p, sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    SU2B_CUDA(cudaStreamSynchronize(c->stream));
    *nonphys = (unsigned long)s.h_counters[0];
  }
  if (center && !output) {
    flow_jst_preprocess(c);
    if (c->nranks > 1) {
      halo_exchange(c, s.lambda.p, 1);
      if (cfg.mg_level == 0) {
        halo_exchange(c, s.sensor.p, 1);
        halo_exchange(c, s.und_lapl.p, s.nVar);
      }
    }
  }
  if (!output) {
    s.res.zero(c->stream);
    // Jacobian.SetValZero(): needed only where blocks are accumulated. The gather assembly OVERWRITES every block (each
    // off-diagonal block belongs to exactly one edge, k_point_sum_rows stores the diagonal ones), so the 8.8 GB memset of
    // the 185^3 case (1.2 ms per iteration) is left out; the coloured strategy adds into the diagonal blocks and keeps it.
    if (cfg.assembly == SU2B200_ASSEMBLY_COLORED) c->sys[0].A.zero(c->stream);
    ++c->sys[0].A_gen;
  }
  if (!cfg.viscous) {
    if (!output && muscl && !center) {
      flow_gradient(c, 1, nd + 2);
      if (c->nranks > 1) halo_exchange(c, s.recon_aliases_prim ? s.grad_prim.p : s.grad_recon.p, s.nG * nd);
      if (limiter) {
        flow_limiter(c);
        if (c->nranks > 1) halo_exchange(c, s.limiter.p, s.nG);
      }
    }
  } else {
    // nPrimVarGrad: nDim+1 for centred schemes (EulerNPrimVarGrad, CEulerVariable.cpp:31-42) and in the Output pass
    // (CNSSolver.cpp:87-88), else nDim+2
    const int nvars = (output || center) ? nd + 1 : nd + 2;
    if (!s.recon_aliases_prim && muscl && !center) {
      flow_gradient(c, 1, nvars);
      if (c->nranks > 1) halo_exchange(c, s.grad_recon.p, s.nG * nd);
    }
    flow_gradient(c, 0, nvars);
    if (c->nranks > 1) halo_exchange(c, s.grad_prim.p, s.nG * nd);
    if (muscl && !center && limiter && !output) {
      flow_limiter(c);
      if (c->nranks > 1) halo_exchange(c, s.limiter.p, s.nG);
    }
    flow_vorticity_strain(c);
  }
}

// CIntegration::Space_Integration (CIntegration.cpp:37-201): edge loop, weak BCs (far-field) in marker order,
// strong BCs (heat-flux walls), then symmetry planes / Euler walls.
static void flow_space_integration(su2b200_ctx* c, unsigned long* nonphys_edges) {
  flow_edge_residual(c, nonphys_edges);
  auto& mk = c->geo.markers;
  for (size_t m = 0; m < mk.size(); ++m)
    if (mk[m].kind == SU2B200_FAR_FIELD) flow_bc_far_field(c, (int)m);
  for (size_t m = 0; m < mk.size(); ++m)
    if (mk[m].kind == SU2B200_HEAT_FLUX) flow_bc_heatflux_wall(c, (int)m);
  for (size_t m = 0; m < mk.size(); ++m)
    if (mk[m].kind == SU2B200_SYMMETRY_PLANE || mk[m].kind == SU2B200_EULER_WALL) flow_bc_sym_plane(c, (int)m);
}

// f2. The reference's mixed-precision mode (-Denable-mixedprec: su2mixedfloat = float, code_config.hpp:133-139): the
// Jacobian, the preconditioner and every Krylov vector are single precision; CSysSolve::HandleTemporariesIn/Out
// (CSysSolve.hpp:242-286) copy LinSysRes / LinSysSol to and from float vectors around the solve.  Here the Jacobian is
// assembled in double as always, rounded once into the float system (the reference rounds every contribution as it adds
// it), and the whole solve runs on the f32 instantiation of the linear algebra (linalg/precond*/krylov *_impl.inc).
template <class Tin, class Tout>
__global__ void k_convert(size_t n, const Tin* __restrict__ a, Tout* __restrict__ b) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) b[i] = (Tout)a[i];
}
static unsigned long mixed_solve(su2b200_ctx* c, int which, int solver, int prec, double tol, int max_iter, int restart,
                                 const double* b, double* x, double* residual) {
  LinSys& ls = c->sys[which];
  LinSysF& lf = c->sysf[which];
  if (lf.nVar != ls.nVar || lf.nnz != ls.nnz || lf.P != ls.P) {  // bind to the same pattern
    lf.nVar = ls.nVar; lf.P = ls.P; lf.Pd = ls.Pd; lf.nnz = ls.nnz;
    lf.row_ptr = ls.row_ptr; lf.col_ind = ls.col_ind; lf.dia_ptr = ls.dia_ptr;
    lf.h_row_ptr = ls.h_row_ptr; lf.h_col_ind = ls.h_col_ind; lf.h_dia_ptr = ls.h_dia_ptr;
    lf.levels_ready = false;
    lf.A.alloc((size_t)ls.nnz * ls.nVar * ls.nVar);
  }
  lf.nparts = ls.nparts; lf.h_parts = ls.h_parts; lf.parts_explicit = ls.parts_explicit;
  lf.sweep_mode = ls.sweep_mode; lf.build_mode = ls.build_mode;
  lf.wide_sweep = ls.wide_sweep; lf.wide_factor = ls.wide_factor; lf.factor_rows_mode = ls.factor_rows_mode;
  const size_t nA = lf.A.n, nV = ls.vecLen();
  const int grid = 148 * 8;
  SU2B_LAUNCH(c, (k_convert<double, float>), grid, 256, nA, (const double*)ls.A.p, lf.A.p);
  ++lf.A_gen;
  const int top = 2 * (std::max(max_iter, 6) + 1) + 2;  // above the solver's own workspace
  float* fb = lf.vec(top).p;
  float* fx = lf.vec(top + 1).p;
  SU2B_LAUNCH(c, (k_convert<double, float>), grid, 256, nV, b, fb);
  SU2B_CUDA(cudaMemsetAsync(fx, 0, nV * sizeof(float), c->stream));
  float res = 0.0f;
  const unsigned long it = linsys_solve(c, lf, solver, prec, (float)tol, max_iter, restart, (const float*)fb, fx, &res);
  SU2B_LAUNCH(c, (k_convert<float, double>), grid, 256, nV, (const float*)fx, x);
  if (residual) *residual = (double)res;
  return it;
}

static void flow_linear_solve(su2b200_ctx* c) {
  auto& s = c->fs;
  const auto& cfg = c->cfg;
  double res = 0.0;
  if (cfg.mixed_precision)
    c->lin_iters[0] = mixed_solve(c, 0, cfg.linear_solver, cfg.linear_prec, cfg.linear_tol, cfg.linear_iter,
                                  cfg.linear_restart, s.res.p, s.sol.p, &res);
  else
    c->lin_iters[0] = linsys_solve(c, c->sys[0], cfg.linear_solver, cfg.linear_prec, cfg.linear_tol, cfg.linear_iter,
                                   cfg.linear_restart, s.res.p, s.sol.p, &res);
  c->lin_res[0] = res;
}
static unsigned long turb_linear_solve(su2b200_ctx* c, double* res) {
  const auto& cfg = c->cfg;
  if (cfg.mixed_precision)
    return mixed_solve(c, 1, cfg.turb_linear_solver, cfg.turb_linear_prec, cfg.turb_linear_tol, cfg.turb_linear_iter,
                       cfg.linear_restart, c->fs.turb_res.p, c->fs.turb_sol.p, res);
  return linsys_solve(c, c->sys[1], cfg.turb_linear_solver, cfg.turb_linear_prec, cfg.turb_linear_tol, cfg.turb_linear_iter,
                      cfg.linear_restart, c->fs.turb_res.p, c->fs.turb_sol.p, res);
}

extern "C" {

int su2b200_flow_preprocessing(su2b200_ctx* c, int output, unsigned long* nonphysical_points) {
  API_BEGIN
  require_flow(c);
  flow_preprocessing(c, output != 0, nonphysical_points);
  API_END
}

int su2b200_flow_set_old_solution(su2b200_ctx* c) {
  API_BEGIN
  require_flow(c);
  vec_copy(c, c->fs.U.n, c->fs.U.p, c->fs.U_old.p);
  API_END
}

int su2b200_flow_set_time_step(su2b200_ctx* c, double* min_dt, double* max_dt) {
  API_BEGIN
  require_flow(c);
  flow_time_step(c);
  if (min_dt) *min_dt = c->min_dt;
  if (max_dt) *max_dt = c->max_dt;
  API_END
}

int su2b200_flow_edge_residual(su2b200_ctx* c, unsigned long* nonphysical_edges) {
  API_BEGIN
  require_flow(c);
  flow_edge_residual(c, nonphysical_edges);
  API_END
}

int su2b200_flow_bc_far_field(su2b200_ctx* c, int m) {
  API_BEGIN
  require_flow(c);
  flow_bc_far_field(c, m);
  API_END
}
int su2b200_flow_bc_sym_plane(su2b200_ctx* c, int m) {
  API_BEGIN
  require_flow(c);
  flow_bc_sym_plane(c, m);
  API_END
}
int su2b200_flow_bc_heatflux_wall(su2b200_ctx* c, int m) {
  API_BEGIN
  require_flow(c);
  flow_bc_heatflux_wall(c, m);
  API_END
}

int su2b200_flow_space_integration(su2b200_ctx* c) {
  API_BEGIN
  require_flow(c);
  flow_space_integration(c, nullptr);
  API_END
}

int su2b200_flow_prepare_implicit(su2b200_ctx* c, double* rms, double* mx, unsigned long* idx) {
  API_BEGIN
  require_flow(c);
  double r[5] = {0, 0, 0, 0, 0};
  flow_prepare_implicit(c, r, mx, idx);
  for (int v = 0; v < c->fs.nVar; ++v) { c->fs.last_rms[v] = r[v]; if (rms) rms[v] = r[v]; }
  API_END
}

int su2b200_flow_linear_solve(su2b200_ctx* c, unsigned long* iterations, double* residual) {
  API_BEGIN
  require_flow(c);
  flow_linear_solve(c);
  if (iterations) *iterations = c->lin_iters[0];
  if (residual) *residual = c->lin_res[0];
  API_END
}

int su2b200_flow_complete_implicit(su2b200_ctx* c) {
  API_BEGIN
  require_flow(c);
  flow_complete_implicit(c);
  if (c->nranks > 1) halo_exchange(c, c->fs.U.p, c->fs.nVar);
  API_END
}

int su2b200_flow_forces(su2b200_ctx* c, double* coeffs) {
  API_BEGIN
  require_flow(c);
  flow_forces(c, coeffs);
  API_END
}

// ---- turbulence ------------------------------------------------------------------------------------
int su2b200_turb_preprocessing(su2b200_ctx* c) {
  API_BEGIN
  require_flow(c);
  turb_preprocessing(c);
  API_END
}
int su2b200_turb_set_time_step(su2b200_ctx* c) {
  API_BEGIN
  require_flow(c);
  turb_set_time_step(c);
  API_END
}
int su2b200_turb_space_integration(su2b200_ctx* c) {
  API_BEGIN
  require_flow(c);
  turb_space_integration(c);
  API_END
}
int su2b200_turb_prepare_implicit(su2b200_ctx* c, double* rms, double* mx, unsigned long* idx) {
  API_BEGIN
  require_flow(c);
  double r[2] = {0.0, 0.0};  // rms/mx/idx: [1] for SA, [2] for SST
  turb_prepare_implicit_generic(c, r, mx, idx);
  c->fs.last_rms_turb = r[0];
  c->fs.last_rms_turb2 = r[1];
  if (rms) for (int v = 0; v < c->fs.nTurb; ++v) rms[v] = r[v];
  API_END
}
int su2b200_turb_linear_solve(su2b200_ctx* c, unsigned long* iterations, double* residual) {
  API_BEGIN
  require_flow(c);
  const auto& cfg = c->cfg;
  double res = 0.0;
  c->lin_iters[1] = turb_linear_solve(c, &res);
  c->lin_res[1] = res;
  if (iterations) *iterations = c->lin_iters[1];
  if (residual) *residual = res;
  API_END
}
int su2b200_turb_complete_implicit(su2b200_ctx* c) {
  API_BEGIN
  require_flow(c);
  turb_complete_implicit(c);
  API_END
}
int su2b200_turb_postprocessing(su2b200_ctx* c) {
  API_BEGIN
  require_flow(c);
  turb_postprocessing(c);
  API_END
}

// ---- whole iterations: CMultiGridIntegration::MultiGrid_Iteration (MGLEVEL 0) + SingleGrid_Iteration(TURB) ----
int su2b200_iterate(su2b200_ctx* c, int nIter, su2b200_iter_report* reports) {
  API_BEGIN
  require_flow(c);
  auto& s = c->fs;
  const int nv = s.nVar;
  for (int it = 0; it < nIter; ++it) {
    su2b200_iter_report rep{};
    unsigned long np = 0, ne = 0;
    // MultiGrid_Cycle, one pre-smoothing step on the finest grid (CMultiGridIntegration.cpp:163-206)
    flow_preprocessing(c, false, &np);
    vec_copy(c, s.U.n, s.U.p, s.U_old.p);
    flow_time_step(c);
    flow_space_integration(c, &ne);
    double rms[5] = {0, 0, 0, 0, 0};
    flow_prepare_implicit(c, rms, nullptr, nullptr);
    for (int v = 0; v < nv; ++v) {
      if (rms[v] != rms[v]) throw Error("SU2 has diverged (NaN detected).");
      if (std::log10(rms[v]) > 20.0) throw Error("SU2 has diverged (Residual > 10^20 detected).");
    }
    flow_linear_solve(c);
    flow_complete_implicit(c);
    if (c->nranks > 1) halo_exchange(c, s.U.p, nv);
    // MultiGrid_Iteration: Preprocessing(Output=true) then forces (CMultiGridIntegration.cpp:116-127)
    flow_preprocessing(c, true, nullptr);
    double coeffs[9] = {0};
    flow_forces(c, coeffs);
    for (int v = 0; v < nv; ++v) rep.rms_flow[v] = rms[v];
    rep.CL = coeffs[0];
    rep.CD = coeffs[1];
    rep.linsol_iter_flow = c->lin_iters[0];
    rep.linsol_res_flow = c->lin_res[0];
    rep.nonphysical_points = np;
    rep.nonphysical_edges = ne;
    if (c->cfg.rans_sa || c->cfg.rans_sst) {
      // CSingleGridIntegration::SingleGrid_Iteration (CSingleGridIntegration.cpp:40-105)
      turb_preprocessing(c);
      vec_copy(c, s.nu.n, s.nu.p, s.nu_old.p);
      turb_set_time_step(c);
      turb_space_integration(c);
      double trms2[2] = {0.0, 0.0};
      turb_prepare_implicit_generic(c, trms2, nullptr, nullptr);
      const double trms = trms2[0];
      const auto& cfg = c->cfg;
      double res = 0.0;
      c->lin_iters[1] = turb_linear_solve(c, &res);
      c->lin_res[1] = res;
      turb_complete_implicit(c);
      turb_postprocessing(c);
      rep.rms_turb = trms;
      rep.rms_turb2 = trms2[1];
      s.last_rms_turb2 = trms2[1];
      rep.linsol_iter_turb = c->lin_iters[1];
      rep.linsol_res_turb = res;
      s.last_rms_turb = trms;
    }
    for (int v = 0; v < nv; ++v) s.last_rms[v] = rms[v];
    // CFluidIteration::Iterate (CFluidIteration.cpp:131-137): CFL adaptation closes the iteration
    if (c->cfg.cfl_adapt) flow_adapt_cfl(c, c->inner_iter, nullptr, nullptr, nullptr);
    ++c->inner_iter;
    if (reports) reports[it] = rep;
  }
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  API_END
}

int su2b200_flow_adapt_cfl(su2b200_ctx* c, unsigned long inner_iter, double* cfl_min, double* cfl_max, double* cfl_avg) {
  API_BEGIN
  require_flow(c);
  flow_adapt_cfl(c, inner_iter, cfl_min, cfl_max, cfl_avg);
  API_END
}

// ---- f1. FAS multigrid: CMultiGridIntegration (SU2_CFD/src/integration/CMultiGridIntegration.cpp) -------------------
// Every level is a context of its own (geometry of the agglomerated grid, its CFL, mg_level > 0 => first-order / Lax
// scheme); the transfer operators below are gather kernels over the Children_CV lists of the coarse level, on the
// finest level's stream. The agglomeration itself (CMultiGridGeometry) is host preprocessing outside the hot path: the
// coarse grids arrive through su2b200_geometry_upload like any other grid.
}  // extern "C"

namespace {
inline int nblk_(long n, int bs) { return (int)((n + bs - 1) / bs); }

// CSolver::MultigridRestriction (CSolver.hpp:4343-4362): volume-weighted average of the children
template <int NV>
__global__ void k_mg_restrict(int Pc, const int* __restrict__ cptr, const int* __restrict__ child,
                              const double* __restrict__ volF, const double* __restrict__ volC,
                              const double* __restrict__ UF, double* __restrict__ UC) {
  const int ic = blockIdx.x * blockDim.x + threadIdx.x;
  if (ic >= Pc) return;
  double acc[NV];
#pragma unroll
  for (int v = 0; v < NV; ++v) acc[v] = 0.0;
  const double scale = 1 / volC[ic];
  for (int k = cptr[ic]; k < cptr[ic + 1]; ++k) {
    const int f = child[k];
    const double w = volF[f] * scale;
#pragma unroll
    for (int v = 0; v < NV; ++v) acc[v] += w * UF[(size_t)f * NV + v];
  }
#pragma unroll
  for (int v = 0; v < NV; ++v) UC[(size_t)ic * NV + v] = acc[v];
}
// SetForcing_Term, first loop (:516-531): trunc_c = sum over children of damp * fine residual
template <int NV>
__global__ void k_mg_sum_children(int Pc, double factor, const int* __restrict__ cptr, const int* __restrict__ child,
                                  const double* __restrict__ resF, double* __restrict__ truncC) {
  const int ic = blockIdx.x * blockDim.x + threadIdx.x;
  if (ic >= Pc) return;
  double acc[NV];
#pragma unroll
  for (int v = 0; v < NV; ++v) acc[v] = 0.0;
  for (int k = cptr[ic]; k < cptr[ic + 1]; ++k) {
    const int f = child[k];
#pragma unroll
    for (int v = 0; v < NV; ++v) acc[v] += factor * resF[(size_t)f * NV + v];
  }
#pragma unroll
  for (int v = 0; v < NV; ++v) truncC[(size_t)ic * NV + v] = acc[v];
}
__global__ void k_mg_axpy(size_t n, double alpha, const double* __restrict__ x, double* __restrict__ y) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) y[i] += alpha * x[i];
}
// GetProlongated_Correction, first loop (:319-342): Solution_Old_c = u_c - restriction(u_f), in the reference's order
template <int NV>
__global__ void k_mg_correction(int Pc, const int* __restrict__ cptr, const int* __restrict__ child,
                                const double* __restrict__ volF, const double* __restrict__ volC,
                                const double* __restrict__ UF, const double* __restrict__ UC, double* __restrict__ UoldC) {
  const int ic = blockIdx.x * blockDim.x + threadIdx.x;
  if (ic >= Pc) return;
  double acc[NV];
#pragma unroll
  for (int v = 0; v < NV; ++v) acc[v] = 0.0;
  const double areaParent = volC[ic];
  for (int k = cptr[ic]; k < cptr[ic + 1]; ++k) {
    const int f = child[k];
    const double areaChild = volF[f];
#pragma unroll
    for (int v = 0; v < NV; ++v) acc[v] -= UF[(size_t)f * NV + v] * areaChild / areaParent;
  }
#pragma unroll
  for (int v = 0; v < NV; ++v) UoldC[(size_t)ic * NV + v] = acc[v] + UC[(size_t)ic * NV + v];
}
// second loop (:370-377): the correction of a coarse point goes to each of its children (fine LinSysRes)
template <int NV>
__global__ void k_mg_prolong(int Pc, const int* __restrict__ cptr, const int* __restrict__ child,
                             const double* __restrict__ UoldC, double* __restrict__ resF) {
  const int ic = blockIdx.x * blockDim.x + threadIdx.x;
  if (ic >= Pc) return;
  for (int k = cptr[ic]; k < cptr[ic + 1]; ++k) {
    const int f = child[k];
#pragma unroll
    for (int v = 0; v < NV; ++v) resF[(size_t)f * NV + v] = UoldC[(size_t)ic * NV + v];
  }
}
// SmoothProlongated_Correction (:383-456): one Jacobi sweep res = (old + c * sum of neighbours) / (1 + c * nNeighbours)
template <int NV>
__global__ void k_mg_smooth(int P, double coeff, const int* __restrict__ nptr, const int* __restrict__ nbr,
                            const double* __restrict__ old, const double* __restrict__ cur, double* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  double sum[NV];
#pragma unroll
  for (int v = 0; v < NV; ++v) sum[v] = 0.0;
  for (int k = nptr[i]; k < nptr[i + 1]; ++k) {
    const int j = nbr[k];
#pragma unroll
    for (int v = 0; v < NV; ++v) sum[v] += cur[(size_t)j * NV + v];
  }
  const double factor = 1.0 / (1.0 + coeff * (double)(nptr[i + 1] - nptr[i]));
#pragma unroll
  for (int v = 0; v < NV; ++v) out[(size_t)i * NV + v] = (old[(size_t)i * NV + v] + coeff * sum[v]) * factor;
}
template <int NV>
__global__ void k_mg_restore_vertices(int nV, const int* __restrict__ vnode, const double* __restrict__ old,
                                      double* __restrict__ res) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nV) return;
  const int i = vnode[t];
#pragma unroll
  for (int v = 0; v < NV; ++v) res[(size_t)i * NV + v] = old[(size_t)i * NV + v];
}
// SetProlongated_Correction (:458-484): u_f += damp * correction, a NaN correction counts as zero
__global__ void k_mg_apply(size_t n, double factor, double* __restrict__ res, double* __restrict__ U) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    double r = res[i];
    if (r != r) { r = 0.0; res[i] = 0.0; }
    U[i] += factor * r;
  }
}
__global__ void k_mg_zero_velocity(int nV, int Pd, int nDim, const int* __restrict__ vnode, double* __restrict__ arr) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nV) return;
  const int i = vnode[t];
  if (i >= Pd) return;
  for (int d = 0; d < nDim; ++d) arr[(size_t)i * (nDim + 2) + 1 + d] = 0.0;
}

void require_link(su2b200_ctx* fine, su2b200_ctx* coarse) {
  if (!fine || !coarse) throw Error("null context");
  SU2B_CUDA(cudaSetDevice(fine->device));
  SU2B_REQUIRE(fine->fs.allocated && coarse->fs.allocated, "allocate both levels first");
  SU2B_REQUIRE(coarse->mg_fine == fine && coarse->mg_child_ptr.n == (size_t)coarse->geo.P + 1,
               "levels are not linked (su2b200_mg_link)");
  SU2B_REQUIRE(fine->nranks == 1, "multigrid runs on a single rank for now");
}
// no-slip walls of the COARSE level: momentum components of `arr` := 0 (SetVelSolutionVector / SetVelocity_Old /
// SetVel_ResTruncError_Zero with a zero vector)
void zero_wall_velocity(su2b200_ctx* coarse, double* arr) {
  auto& g = coarse->geo;
  for (auto& m : g.markers)
    if (m.kind == SU2B200_HEAT_FLUX && m.nVertex > 0)
      SU2B_LAUNCH(coarse, k_mg_zero_velocity, nblk_(m.nVertex, 64), 64, m.nVertex, g.Pd, g.nDim, m.node.p, arr);
}
#define MG_NV(nd, K, ...)                        \
  do {                                           \
    if ((nd) == 2) K<4> __VA_ARGS__; else K<5> __VA_ARGS__; \
  } while (0)

void mg_set_restricted_solution(su2b200_ctx* fine, su2b200_ctx* coarse) {
  require_link(fine, coarse);
  auto &gf = fine->geo, &gc = coarse->geo;
  const int nb = nblk_(gc.Pd, 128);
  if (gf.nDim == 2) k_mg_restrict<4><<<nb, 128, 0, fine->stream>>>(gc.Pd, coarse->mg_child_ptr.p, coarse->mg_child.p, gf.volume.p, gc.volume.p, fine->fs.U.p, coarse->fs.U.p);
  else k_mg_restrict<5><<<nb, 128, 0, fine->stream>>>(gc.Pd, coarse->mg_child_ptr.p, coarse->mg_child.p, gf.volume.p, gc.volume.p, fine->fs.U.p, coarse->fs.U.p);
  fine->launches++;
  SU2B_CUDA(cudaGetLastError());
  zero_wall_velocity(coarse, coarse->fs.U.p);
}
void mg_set_residual_term(su2b200_ctx* c) {
  SU2B_REQUIRE(c->fs.res_trunc.n == c->fs.res.n, "no multigrid forcing term on this level");
  const size_t n = (size_t)c->geo.Pd * c->fs.nVar;
  k_mg_axpy<<<nblk_((long)n, 256), 256, 0, c->stream>>>(n, 1.0, c->fs.res_trunc.p, c->fs.res.p);
  c->launches++;
  SU2B_CUDA(cudaGetLastError());
}
void mg_set_forcing_term(su2b200_ctx* fine, su2b200_ctx* coarse, double damp) {
  require_link(fine, coarse);
  auto& gc = coarse->geo;
  const int nb = nblk_(gc.Pd, 128);
  if (gc.nDim == 2) k_mg_sum_children<4><<<nb, 128, 0, fine->stream>>>(gc.Pd, damp, coarse->mg_child_ptr.p, coarse->mg_child.p, fine->fs.res.p, coarse->fs.res_trunc.p);
  else k_mg_sum_children<5><<<nb, 128, 0, fine->stream>>>(gc.Pd, damp, coarse->mg_child_ptr.p, coarse->mg_child.p, fine->fs.res.p, coarse->fs.res_trunc.p);
  fine->launches++;
  SU2B_CUDA(cudaGetLastError());
  zero_wall_velocity(coarse, coarse->fs.res_trunc.p);
  const size_t n = (size_t)gc.Pd * coarse->fs.nVar;
  k_mg_axpy<<<nblk_((long)n, 256), 256, 0, fine->stream>>>(n, -1.0, coarse->fs.res.p, coarse->fs.res_trunc.p);
  fine->launches++;
  SU2B_CUDA(cudaGetLastError());
}
void mg_get_prolongated_correction(su2b200_ctx* fine, su2b200_ctx* coarse) {
  require_link(fine, coarse);
  auto &gf = fine->geo, &gc = coarse->geo;
  const int nb = nblk_(gc.Pd, 128);
  if (gc.nDim == 2) k_mg_correction<4><<<nb, 128, 0, fine->stream>>>(gc.Pd, coarse->mg_child_ptr.p, coarse->mg_child.p, gf.volume.p, gc.volume.p, fine->fs.U.p, coarse->fs.U.p, coarse->fs.U_old.p);
  else k_mg_correction<5><<<nb, 128, 0, fine->stream>>>(gc.Pd, coarse->mg_child_ptr.p, coarse->mg_child.p, gf.volume.p, gc.volume.p, fine->fs.U.p, coarse->fs.U.p, coarse->fs.U_old.p);
  fine->launches++;
  SU2B_CUDA(cudaGetLastError());
  zero_wall_velocity(coarse, coarse->fs.U_old.p);
  if (gc.nDim == 2) k_mg_prolong<4><<<nb, 128, 0, fine->stream>>>(gc.Pd, coarse->mg_child_ptr.p, coarse->mg_child.p, coarse->fs.U_old.p, fine->fs.res.p);
  else k_mg_prolong<5><<<nb, 128, 0, fine->stream>>>(gc.Pd, coarse->mg_child_ptr.p, coarse->mg_child.p, coarse->fs.U_old.p, fine->fs.res.p);
  fine->launches++;
  SU2B_CUDA(cudaGetLastError());
}
void mg_smooth_prolongated_correction(su2b200_ctx* c, int nSmooth, double coeff) {
  if (nSmooth <= 0) return;
  auto& g = c->geo;
  auto& s = c->fs;
  const size_t n = s.res.n;
  dvec<double>& old = c->sys[0].vec(0);   // Residual_Old
  dvec<double>& tmp = c->sys[0].vec(1);
  SU2B_CUDA(cudaMemcpyAsync(old.p, s.res.p, n * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
  for (int it = 0; it < nSmooth; ++it) {
    if (g.nDim == 2) k_mg_smooth<4><<<nblk_(g.P, 128), 128, 0, c->stream>>>(g.P, coeff, g.nbr_ptr.p, g.nbr_pt.p, old.p, s.res.p, tmp.p);
    else k_mg_smooth<5><<<nblk_(g.P, 128), 128, 0, c->stream>>>(g.P, coeff, g.nbr_ptr.p, g.nbr_pt.p, old.p, s.res.p, tmp.p);
    c->launches++;
    SU2B_CUDA(cudaGetLastError());
    SU2B_CUDA(cudaMemcpyAsync(s.res.p, tmp.p, n * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
    for (auto& m : g.markers) {  // boundary points keep the unsmoothed correction
      if (m.nVertex == 0 || m.kind == SU2B200_SEND_RECEIVE) continue;
      if (g.nDim == 2) k_mg_restore_vertices<4><<<nblk_(m.nVertex, 64), 64, 0, c->stream>>>(m.nVertex, m.node.p, old.p, s.res.p);
      else k_mg_restore_vertices<5><<<nblk_(m.nVertex, 64), 64, 0, c->stream>>>(m.nVertex, m.node.p, old.p, s.res.p);
      c->launches++;
    }
  }
}
void mg_set_prolongated_correction(su2b200_ctx* c, double damp) {
  const size_t n = (size_t)c->geo.Pd * c->fs.nVar;
  k_mg_apply<<<nblk_((long)n, 256), 256, 0, c->stream>>>(n, damp, c->fs.res.p, c->fs.U.p);
  c->launches++;
  SU2B_CUDA(cudaGetLastError());
}

// one smoothing step on a level == the body of MultiGrid_Cycle's pre/post-smoothing loops (:168-206) for EULER_IMPLICIT
void mg_smooth_level(su2b200_ctx* c, double* rms, unsigned long* np, unsigned long* ne) {
  auto& s = c->fs;
  flow_preprocessing(c, false, np);
  vec_copy(c, s.U.n, s.U.p, s.U_old.p);
  flow_time_step(c);
  flow_space_integration(c, ne);
  flow_prepare_implicit(c, rms, nullptr, nullptr);
  flow_linear_solve(c);
  flow_complete_implicit(c);
}

// MultiGrid_Cycle (:129-300), recursion over the levels
void mg_cycle(su2b200_ctx** lv, const su2b200_mg_config& mc, int iMesh, int recursive, double* rms_fine,
              unsigned long* np, unsigned long* ne) {
  su2b200_ctx* fine = lv[iMesh];
  double rms[5] = {0, 0, 0, 0, 0};
  for (int k = 0; k < mc.pre_smooth[iMesh]; ++k) {
    mg_smooth_level(fine, rms, iMesh == 0 ? np : nullptr, iMesh == 0 ? ne : nullptr);
    if (iMesh == 0) for (int v = 0; v < 5; ++v) rms_fine[v] = rms[v];
  }
  if (iMesh < mc.nLevels) {
    su2b200_ctx* coarse = lv[iMesh + 1];
    // r_k = P_k + F_k(u_k)   (the reference switches to EULER_EXPLICIT here only to skip the Jacobian)
    flow_preprocessing(fine, false, nullptr);
    flow_space_integration(fine, nullptr);
    if (iMesh > 0) mg_set_residual_term(fine);  // the finest level carries no forcing term
    // r_(k+1) = F_(k+1)(I u_k)
    mg_set_restricted_solution(fine, coarse);
    flow_preprocessing(coarse, false, nullptr);
    flow_space_integration(coarse, nullptr);
    // P_(k+1) = I(r_k) - r_(k+1)
    mg_set_forcing_term(fine, coarse, mc.damp_restriction);
    for (int imu = 0; imu <= recursive; ++imu) {
      int next = recursive;
      if (iMesh == mc.nLevels - 2) next = 0;
      mg_cycle(lv, mc, iMesh + 1, next, rms_fine, np, ne);
    }
    mg_get_prolongated_correction(fine, coarse);
    mg_smooth_prolongated_correction(fine, mc.correction_smooth[iMesh], 1.25);
    mg_set_prolongated_correction(fine, mc.damp_prolongation);
    for (int k = 0; k < mc.post_smooth[iMesh]; ++k) {
      mg_smooth_level(fine, rms, nullptr, nullptr);
      if (iMesh == 0) for (int v = 0; v < 5; ++v) rms_fine[v] = rms[v];
    }
  }
}
}  // namespace

extern "C" {

int su2b200_mg_link(su2b200_ctx* fine, su2b200_ctx* coarse, const unsigned long* children_ptr,
                    const unsigned long* children) {
  API_BEGIN
  if (!fine || !coarse) throw Error("null context");
  SU2B_CUDA(cudaSetDevice(fine->device));
  SU2B_REQUIRE(fine->device == coarse->device, "multigrid levels must live on one device");
  SU2B_REQUIRE(fine->fs.allocated && coarse->fs.allocated, "allocate both levels first");
  SU2B_REQUIRE(children_ptr && children, "null children arrays");
  SU2B_REQUIRE(fine->geo.nDim == coarse->geo.nDim && fine->fs.nVar == coarse->fs.nVar, "levels differ in nDim / nVar");
  const int Pc = coarse->geo.P;
  std::vector<int> cp(Pc + 1), ch(children_ptr[Pc]);
  for (int i = 0; i <= Pc; ++i) cp[i] = (int)children_ptr[i];
  std::vector<char> seen(fine->geo.P, 0);
  for (size_t k = 0; k < ch.size(); ++k) {
    SU2B_REQUIRE(children[k] < (unsigned long)fine->geo.P, "child index outside the fine grid");
    SU2B_REQUIRE(!seen[children[k]], "a fine point has two parents");
    seen[children[k]] = 1;
    ch[k] = (int)children[k];
  }
  SU2B_CUDA(cudaStreamSynchronize(coarse->stream));
  SU2B_CUDA(cudaStreamSynchronize(fine->stream));
  coarse->mg_child_ptr.upload(cp);
  coarse->mg_child.upload(ch);
  coarse->mg_fine = fine;
  if (!coarse->stream_borrowed) {  // every level on the finest level's stream: the cycle is one ordered sequence
    cudaStreamDestroy(coarse->stream);
    coarse->stream_borrowed = true;
  }
  coarse->stream = fine->stream;
  for (su2b200_ctx* c : {fine, coarse}) {
    if (c->fs.res_trunc.n != c->fs.res.n) {
      c->fs.res_trunc.alloc(c->fs.res.n);
      c->fs.res_trunc.zero(fine->stream);
    }
  }
  SU2B_CUDA(cudaStreamSynchronize(fine->stream));
  API_END
}

int su2b200_mg_set_restricted_solution(su2b200_ctx* fine, su2b200_ctx* coarse) {
  API_BEGIN
  mg_set_restricted_solution(fine, coarse);
  API_END
}
int su2b200_mg_set_residual_term(su2b200_ctx* c) {
  API_BEGIN
  require_flow(c);
  mg_set_residual_term(c);
  API_END
}
int su2b200_mg_set_forcing_term(su2b200_ctx* fine, su2b200_ctx* coarse, double damp_restriction) {
  API_BEGIN
  mg_set_forcing_term(fine, coarse, damp_restriction);
  API_END
}
int su2b200_mg_get_prolongated_correction(su2b200_ctx* fine, su2b200_ctx* coarse) {
  API_BEGIN
  mg_get_prolongated_correction(fine, coarse);
  API_END
}
int su2b200_mg_smooth_prolongated_correction(su2b200_ctx* c, int nSmooth, double coeff) {
  API_BEGIN
  require_flow(c);
  mg_smooth_prolongated_correction(c, nSmooth, coeff);
  API_END
}
int su2b200_mg_set_prolongated_correction(su2b200_ctx* c, double damp_prolongation) {
  API_BEGIN
  require_flow(c);
  mg_set_prolongated_correction(c, damp_prolongation);
  API_END
}

// CMultiGridIntegration::MultiGrid_Iteration (:37-127): nIter FAS cycles, then Preprocessing(Output) + forces on the finest grid
int su2b200_mg_iterate(su2b200_ctx** levels, const su2b200_mg_config* mc, int nIter, su2b200_iter_report* reports) {
  API_BEGIN
  SU2B_REQUIRE(levels && mc, "null argument");
  SU2B_REQUIRE(mc->nLevels >= 1 && mc->nLevels <= 7, "nLevels (coarse levels) must be in 1..7");
  for (int l = 0; l <= mc->nLevels; ++l) {
    require_flow(levels[l]);
    SU2B_REQUIRE(levels[l]->cfg.mg_level == l, "level contexts must be configured with mg_level = their position");
    SU2B_REQUIRE(!levels[l]->cfg.rans_sa, "multigrid drives the mean-flow system only");
    if (l > 0) SU2B_REQUIRE(levels[l]->mg_fine == levels[l - 1], "levels are not linked in order (su2b200_mg_link)");
  }
  su2b200_ctx* c = levels[0];
  const int nv = c->fs.nVar;
  for (int it = 0; it < nIter; ++it) {
    su2b200_iter_report rep{};
    unsigned long np = 0, ne = 0;
    double rms[5] = {0, 0, 0, 0, 0};
    int recursive = mc->cycle;  // 0 = V, 1 = W (FULLMG starts as V)
    if (recursive > 1) recursive = 0;
    mg_cycle(levels, *mc, 0, recursive, rms, &np, &ne);
    for (int v = 0; v < nv; ++v) {
      if (rms[v] != rms[v]) throw Error("SU2 has diverged (NaN detected).");
      if (std::log10(rms[v]) > 20.0) throw Error("SU2 has diverged (Residual > 10^20 detected).");
    }
    flow_preprocessing(c, true, nullptr);
    double coeffs[9] = {0};
    flow_forces(c, coeffs);
    for (int v = 0; v < nv; ++v) rep.rms_flow[v] = rms[v];
    rep.CL = coeffs[0];
    rep.CD = coeffs[1];
    rep.linsol_iter_flow = c->lin_iters[0];
    rep.linsol_res_flow = c->lin_res[0];
    rep.nonphysical_points = np;
    rep.nonphysical_edges = ne;
    if (reports) reports[it] = rep;
  }
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  API_END
}

// ---- CSysMatrix / CSysSolve level ----------------------------------------------------------------------
int su2b200_linsys_create(su2b200_ctx* c, int which, int nVar, unsigned long nPoint, unsigned long nPointDomain,
                          const unsigned long* row_ptr, const unsigned long* col_ind) {
  API_BEGIN
  if (!c) throw Error("null context");
  SU2B_CUDA(cudaSetDevice(c->device));
  SU2B_REQUIRE(which == 0 || which == 1, "system index must be 0 or 1");
  SU2B_REQUIRE(nVar == 1 || nVar == 2 || nVar == 4 || nVar == 5, "nVar must be 1, 2, 4 or 5");
  SU2B_REQUIRE(row_ptr && col_ind, "null pattern");
  auto& ls = c->sys[which];
  ls.nVar = nVar;
  ls.P = (int)nPoint;
  ls.Pd = (int)nPointDomain;
  ls.h_row_ptr.resize(nPoint + 1);
  for (size_t i = 0; i <= nPoint; ++i) ls.h_row_ptr[i] = (int)row_ptr[i];
  ls.nnz = ls.h_row_ptr[nPoint];
  ls.h_col_ind.resize(ls.nnz);
  for (long k = 0; k < ls.nnz; ++k) ls.h_col_ind[k] = (int)col_ind[k];
  ls.h_dia_ptr.assign(nPoint, -1);
  for (int i = 0; i < ls.P; ++i) {
    for (int k = ls.h_row_ptr[i]; k < ls.h_row_ptr[i + 1]; ++k) {
      if (k > ls.h_row_ptr[i]) SU2B_REQUIRE(ls.h_col_ind[k] > ls.h_col_ind[k - 1], "columns must ascend within a row");
      if (ls.h_col_ind[k] == i) ls.h_dia_ptr[i] = k;
    }
    SU2B_REQUIRE(ls.h_dia_ptr[i] >= 0, "pattern lacks a diagonal entry");
  }
  ls.own_row_ptr.upload(ls.h_row_ptr);
  ls.own_col_ind.upload(ls.h_col_ind);
  ls.own_dia_ptr.upload(ls.h_dia_ptr);
  ls.row_ptr = ls.own_row_ptr.p;
  ls.col_ind = ls.own_col_ind.p;
  ls.dia_ptr = ls.own_dia_ptr.p;
  ls.levels_ready = false;
  ls.h_parts.clear();
  ls.A.alloc((size_t)ls.nnz * nVar * nVar);
  ls.A.zero();
  if (c->configured) ls.nparts = std::max(1, c->cfg.prec_partitions);
  API_END
}

static LinSys& get_sys(su2b200_ctx* c, int which) {
  if (!c) throw Error("null context");
  SU2B_CUDA(cudaSetDevice(c->device));
  SU2B_REQUIRE(which == 0 || which == 1, "system index must be 0 or 1");
  SU2B_REQUIRE(c->sys[which].nVar > 0, "linear system not created");
  return c->sys[which];
}

int su2b200_linsys_set_partitions(su2b200_ctx* c, int which, int nparts) {
  API_BEGIN
  auto& ls = get_sys(c, which);
  ls.nparts = std::max(1, nparts);
  ls.h_parts.clear();
  ls.parts_explicit = false;
  API_END
}

int su2b200_linsys_set_schedule(su2b200_ctx* c, int which, int apply_mode, int build_mode) {
  API_BEGIN
  auto& ls = get_sys(c, which);
  if (apply_mode >= 0) { SU2B_REQUIRE(apply_mode <= 2, "apply_mode: 0 levels, 1 flag sweeps, 2 polling sweeps"); ls.sweep_mode = apply_mode; }
  if (build_mode >= 0) { SU2B_REQUIRE(build_mode <= 2, "build_mode: 0 levels, 1 flag sweep, 2 polling factorisation"); ls.build_mode = build_mode; }
  API_END
}

int su2b200_linsys_set_wide_threshold(su2b200_ctx* c, int which, int sweep_rows, int factor_rows, int factor_mode) {
  API_BEGIN
  auto& ls = get_sys(c, which);
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  if (sweep_rows > 0 && sweep_rows != ls.wide_sweep) {
    ls.wide_sweep = sweep_rows;
    ls.sweep_meta_lower.release();  // build_meta() regenerates the hybrid split
    ls.sweep_meta_upper.release();
    ls.sweep_metar_lower.release();
    ls.sweep_metar_upper.release();
  }
  if ((factor_rows > 0 && factor_rows != ls.wide_factor) || (factor_mode >= 0 && factor_mode != ls.factor_rows_mode)) {
    if (factor_rows > 0) ls.wide_factor = factor_rows;
    if (factor_mode >= 0) { SU2B_REQUIRE(factor_mode <= 2, "factor_mode: 0 warp per row, 1 hybrid, 2 rows kernel"); ls.factor_rows_mode = factor_mode; }
    ls.factor_ready = false;  // factor_symbolic() regenerates descriptors and split
  }
  API_END
}

int su2b200_linsys_set_partition_rows(su2b200_ctx* c, int which, int nparts, const unsigned long* first_row) {
  API_BEGIN
  auto& ls = get_sys(c, which);
  SU2B_REQUIRE(nparts >= 1 && first_row, "need nparts >= 1 and nparts+1 row offsets");
  SU2B_REQUIRE(first_row[0] == 0 && (long)first_row[nparts] == ls.Pd, "partition offsets must span [0, nPointDomain]");
  std::vector<int> h(nparts + 1);
  for (int k = 0; k <= nparts; ++k) {
    h[k] = (int)first_row[k];
    if (k) SU2B_REQUIRE(h[k] >= h[k - 1], "partition offsets must ascend");
  }
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  ls.nparts = nparts;
  ls.h_parts = h;
  ls.parts.release();
  ls.parts_explicit = true;
  API_END
}

int su2b200_linsys_set_matrix(su2b200_ctx* c, int which, const double* blocks) {
  API_BEGIN
  auto& ls = get_sys(c, which);
  SU2B_REQUIRE(blocks, "null matrix");
  SU2B_CUDA(cudaMemcpyAsync(ls.A.p, blocks, ls.A.n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  ++ls.A_gen;
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  API_END
}

int su2b200_linsys_spmv(su2b200_ctx* c, int which, const double* x, double* y) {
  API_BEGIN
  auto& ls = get_sys(c, which);
  const size_t n = ls.vecLen();
  double* dx = ls.vec(0).p;
  double* dy = ls.vec(1).p;
  SU2B_CUDA(cudaMemcpyAsync(dx, x, n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  SU2B_CUDA(cudaMemsetAsync(dy, 0, n * sizeof(double), c->stream));
  linsys_spmv(c, ls, dx, dy);
  SU2B_CUDA(cudaMemcpyAsync(y, dy, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  API_END
}

int su2b200_linsys_build_preconditioner(su2b200_ctx* c, int which, int prec) {
  API_BEGIN
  auto& ls = get_sys(c, which);
  linsys_build_prec(c, ls, prec);
  linsys_check_sweeps(c, ls);
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  API_END
}

int su2b200_linsys_apply_preconditioner(su2b200_ctx* c, int which, int prec, const double* x, double* y) {
  API_BEGIN
  auto& ls = get_sys(c, which);
  const size_t n = ls.vecLen();
  double* dx = ls.vec(0).p;
  double* dy = ls.vec(1).p;
  SU2B_CUDA(cudaMemcpyAsync(dx, x, n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  SU2B_CUDA(cudaMemsetAsync(dy, 0, n * sizeof(double), c->stream));
  linsys_apply_prec(c, ls, prec, dx, dy);
  linsys_check_sweeps(c, ls);
  SU2B_CUDA(cudaMemcpyAsync(y, dy, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  API_END
}

int su2b200_linsys_precondition_and_multiply(su2b200_ctx* c, int which, int prec, const double* v, double* z, double* w) {
  API_BEGIN
  auto& ls = get_sys(c, which);
  const size_t n = ls.vecLen();
  double* dv = ls.vec(0).p;
  double* dz = ls.vec(1).p;
  double* dw = ls.vec(2).p;
  SU2B_CUDA(cudaMemcpyAsync(dv, v, n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  SU2B_CUDA(cudaMemsetAsync(dz, 0, n * sizeof(double), c->stream));
  SU2B_CUDA(cudaMemsetAsync(dw, 0, n * sizeof(double), c->stream));
  linsys_prec_then_product(c, ls, prec, dv, dz, dw);
  linsys_check_sweeps(c, ls);
  SU2B_CUDA(cudaMemcpyAsync(z, dz, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  SU2B_CUDA(cudaMemcpyAsync(w, dw, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  API_END
}

int su2b200_linsys_solve(su2b200_ctx* c, int which, int solver, int prec, double tol, int max_iter, const double* b,
                         double* x, unsigned long* iterations, double* residual) {
  API_BEGIN
  auto& ls = get_sys(c, which);
  const size_t n = ls.vecLen();
  // the solver's own workspace occupies the low pool slots; keep b and x at the top
  const int top = 2 * (std::max(max_iter, 6) + 1) + 2;
  double* db = ls.vec(top).p;
  double* dx = ls.vec(top + 1).p;
  SU2B_CUDA(cudaMemcpyAsync(db, b, n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  SU2B_CUDA(cudaMemsetAsync(dx, 0, n * sizeof(double), c->stream));
  double res = 0.0;
  const int restart = c->configured ? c->cfg.linear_restart : max_iter;
  const unsigned long it = linsys_solve(c, ls, solver, prec, tol, max_iter, restart, db, dx, &res);
  SU2B_CUDA(cudaMemcpyAsync(x, dx, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  if (iterations) *iterations = it;
  if (residual) *residual = res;
  API_END
}

int su2b200_vector_dot(su2b200_ctx* c, int which, const double* a, const double* b, double* result) {
  API_BEGIN
  auto& ls = get_sys(c, which);
  const size_t n = ls.vecLen();
  double* da = ls.vec(0).p;
  double* db = ls.vec(1).p;
  SU2B_CUDA(cudaMemcpyAsync(da, a, n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  SU2B_CUDA(cudaMemcpyAsync(db, b, n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  *result = vec_dot(c, ls, da, db);
  API_END
}

// ---- timing -----------------------------------------------------------------------------------------------
int su2b200_timer_start(su2b200_ctx* c) {
  API_BEGIN
  if (!c) throw Error("null context");
  SU2B_CUDA(cudaEventRecord(c->ev0, c->stream));
  API_END
}
int su2b200_timer_stop(su2b200_ctx* c, float* ms) {
  API_BEGIN
  if (!c) throw Error("null context");
  SU2B_CUDA(cudaEventRecord(c->ev1, c->stream));
  SU2B_CUDA(cudaEventSynchronize(c->ev1));
  SU2B_CUDA(cudaEventElapsedTime(ms, c->ev0, c->ev1));
  API_END
}
unsigned long su2b200_launch_count(su2b200_ctx* c) { return c ? c->launches : 0; }

int su2b200_profile_enable(su2b200_ctx* c, int on) {
  API_BEGIN
  if (!c) throw Error("null context");
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  for (auto& r : c->prof.recs) { c->prof.pool.push_back(r.a); c->prof.pool.push_back(r.b); }
  c->prof.recs.clear();
  c->prof.on = on != 0;
  API_END
}
int su2b200_profile_ncat(void) { return PROF_NCAT; }
const char* su2b200_profile_name(int cat) {
  static const char* names[PROF_NCAT] = {"primitives", "gradient", "limiter", "time_step", "jst_preprocess", "edge_flux",
                                         "point_sum", "boundary", "prepare_implicit", "prec_build", "prec_apply", "spmv",
                                         "dot", "vector_ops", "update", "forces", "turb_edge", "turb_source",
                                         "turb_other", "halo"};
  return (cat >= 0 && cat < PROF_NCAT) ? names[cat] : "";
}
int su2b200_profile_report(su2b200_ctx* c, double* ms, unsigned long* calls) {
  API_BEGIN
  if (!c) throw Error("null context");
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  for (int k = 0; k < PROF_NCAT; ++k) { if (ms) ms[k] = 0.0; if (calls) calls[k] = 0; }
  for (auto& r : c->prof.recs) {
    float t = 0;
    SU2B_CUDA(cudaEventElapsedTime(&t, r.a, r.b));
    if (ms) ms[r.cat] += t;
    if (calls) calls[r.cat]++;
  }
  // SU2B200_PROFILE_GAPS=1: time between the end of one scope and the start of the next (records are in stream order;
  // nested scopes give negative gaps and are skipped), summed per pair of categories, to stderr: what the categories miss
  if (getenv("SU2B200_PROFILE_GAPS") && atoi(getenv("SU2B200_PROFILE_GAPS")) && c->prof.recs.size() > 1) {
    std::vector<double> gap((size_t)PROF_NCAT * PROF_NCAT, 0.0);
    std::vector<int> cnt((size_t)PROF_NCAT * PROF_NCAT, 0);
    double total = 0.0;
    for (size_t k = 0; k + 1 < c->prof.recs.size(); ++k) {
      float t = 0;
      if (cudaEventElapsedTime(&t, c->prof.recs[k].b, c->prof.recs[k + 1].a) != cudaSuccess || t <= 0.0f) continue;
      const size_t key = (size_t)c->prof.recs[k].cat * PROF_NCAT + c->prof.recs[k + 1].cat;
      gap[key] += t;
      cnt[key]++;
      total += t;
    }
    fprintf(stderr, "profile gaps: %.3f ms in total\n", total);
    for (size_t key = 0; key < gap.size(); ++key)
      if (gap[key] > 0.02)
        fprintf(stderr, "  after %-16s before %-16s %8.3f ms in %d gaps\n", su2b200_profile_name((int)(key / PROF_NCAT)),
                su2b200_profile_name((int)(key % PROF_NCAT)), gap[key], cnt[key]);
  }
  API_END
}

int su2b200_bench_edge_kernel(su2b200_ctx* c, int reps, float* ms_avg) {
  API_BEGIN
  require_flow(c);
  bench_edge_kernel(c, reps, ms_avg);
  API_END
}
int su2b200_bench_spmv(su2b200_ctx* c, int which, int reps, float* ms_avg) {
  API_BEGIN
  auto& ls = get_sys(c, which);
  double* dx = ls.vec(0).p;
  double* dy = ls.vec(1).p;
  SU2B_CUDA(cudaEventRecord(c->ev0, c->stream));
  const int nr = c->nranks;
  c->nranks = 1;  // kernel only
  for (int r = 0; r < reps; ++r) linsys_spmv(c, ls, dx, dy);
  c->nranks = nr;
  SU2B_CUDA(cudaEventRecord(c->ev1, c->stream));
  SU2B_CUDA(cudaEventSynchronize(c->ev1));
  float ms = 0;
  SU2B_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
  *ms_avg = ms / reps;
  API_END
}

// preconditioner micro-benchmark on the resident matrix of system `which`: average ms of `reps` builds and applies
int su2b200_bench_prec(su2b200_ctx* c, int which, int prec, int reps, float* ms_build, float* ms_apply) {
  API_BEGIN
  auto& ls = get_sys(c, which);
  double* dx = ls.vec(0).p;
  double* dy = ls.vec(1).p;
  const int nr = c->nranks;
  c->nranks = 1;  // kernels only
  linsys_build_prec(c, ls, prec);  // warm-up (level schedule, allocations)
  linsys_apply_prec(c, ls, prec, dx, dy);
  float ms = 0;
  SU2B_CUDA(cudaEventRecord(c->ev0, c->stream));
  for (int r = 0; r < reps; ++r) linsys_build_prec(c, ls, prec);
  SU2B_CUDA(cudaEventRecord(c->ev1, c->stream));
  SU2B_CUDA(cudaEventSynchronize(c->ev1));
  SU2B_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
  if (ms_build) *ms_build = ms / reps;
  SU2B_CUDA(cudaEventRecord(c->ev0, c->stream));
  for (int r = 0; r < reps; ++r) linsys_apply_prec(c, ls, prec, dx, dy);
  SU2B_CUDA(cudaEventRecord(c->ev1, c->stream));
  SU2B_CUDA(cudaEventSynchronize(c->ev1));
  SU2B_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
  if (ms_apply) *ms_apply = ms / reps;
  c->nranks = nr;
  linsys_check_sweeps(c, ls);
  API_END
}

}  // extern "C"
