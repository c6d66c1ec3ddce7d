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
  const bool muscl = cfg.muscl != 0;
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
      halo_exchange(c, s.sensor.p, 1);
      halo_exchange(c, s.und_lapl.p, s.nVar);
    }
  }
  if (!output) {
    s.res.zero(c->stream);
    c->sys[0].A.zero(c->stream);
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

static void flow_linear_solve(su2b200_ctx* c) {
  auto& s = c->fs;
  const auto& cfg = c->cfg;
  double res = 0.0;
  c->lin_iters[0] = linsys_solve(c, c->sys[0], cfg.linear_solver, cfg.linear_prec, cfg.linear_tol, cfg.linear_iter,
                                 cfg.linear_restart, s.res.p, s.sol.p, &res);
  c->lin_res[0] = res;
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
  flow_prepare_implicit(c, rms, mx, idx);
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
  turb_prepare_implicit_generic(c, rms, mx, idx);
  API_END
}
int su2b200_turb_linear_solve(su2b200_ctx* c, unsigned long* iterations, double* residual) {
  API_BEGIN
  require_flow(c);
  const auto& cfg = c->cfg;
  double res = 0.0;
  c->lin_iters[1] = linsys_solve(c, c->sys[1], cfg.turb_linear_solver, cfg.turb_linear_prec, cfg.turb_linear_tol,
                                 cfg.turb_linear_iter, cfg.linear_restart, c->fs.turb_res.p, c->fs.turb_sol.p, &res);
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
    if (c->cfg.rans_sa) {
      // CSingleGridIntegration::SingleGrid_Iteration (CSingleGridIntegration.cpp:40-105)
      turb_preprocessing(c);
      vec_copy(c, s.nu.n, s.nu.p, s.nu_old.p);
      turb_set_time_step(c);
      turb_space_integration(c);
      double trms = 0;
      turb_prepare_implicit_generic(c, &trms, nullptr, nullptr);
      const auto& cfg = c->cfg;
      double res = 0.0;
      c->lin_iters[1] = linsys_solve(c, c->sys[1], cfg.turb_linear_solver, cfg.turb_linear_prec, cfg.turb_linear_tol,
                                     cfg.turb_linear_iter, cfg.linear_restart, s.turb_res.p, s.turb_sol.p, &res);
      c->lin_res[1] = res;
      turb_complete_implicit(c);
      turb_postprocessing(c);
      rep.rms_turb = trms;
      rep.linsol_iter_turb = c->lin_iters[1];
      rep.linsol_res_turb = res;
    }
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
  SU2B_REQUIRE(nVar == 1 || nVar == 4 || nVar == 5, "nVar must be 1, 4 or 5");
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
