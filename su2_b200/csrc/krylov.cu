// a21-a22: Krylov solvers on device-resident vectors.
//   CSysSolve<T>::FGMRES_LinSolver (Common/src/linear_algebra/CSysSolve.cpp:418-613), ModGramSchmidt (:179-240),
//   Generate/ApplyGivens (:132-162), SolveReduced (:165-176), LinearCombination (:58-119)
//   CSysSolve<T>::BCGSTAB_LinSolver (:989-1159), RFGMRES_LinSolver (:616-639), entry CSysSolve<T>::Solve (:1318-1531)
// The Hessenberg/Givens recurrences are host scalars exactly as in the reference; every vector operation,
// preconditioner application and SpMV is a kernel; dot products come back through one pinned double.
#include "internal.hpp"
#include "kernels.hpp"
#include <cmath>
#include <algorithm>

namespace su2b200 {

__global__ void k_div(size_t n, double d, double* __restrict__ x) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    x[i] /= d;
}
// v += sum_{k<cnt} w[k] * vs[k] (LinearCombination, up to 4 terms per pass like the reference's unrolling)
struct LinComb4 {
  const double* v[4];
  double w[4];
  int cnt;
};
__global__ void k_lincomb_inc(size_t n, LinComb4 lc, double* __restrict__ out) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    double s = lc.v[0][i] * lc.w[0];
    for (int k = 1; k < lc.cnt; ++k) s += lc.v[k][i] * lc.w[k];
    out[i] += s;
  }
}
// p = beta * (p - omega * v) + r   (BCGSTAB)
__global__ void k_bcg_p(size_t n, double beta, double omega, const double* __restrict__ v, const double* __restrict__ r,
                        double* __restrict__ p) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    p[i] = beta * (p[i] - omega * v[i]) + r[i];
}
static inline int vgrid(size_t n) { return (int)std::min<size_t>((n + 255) / 256, 148 * 8); }

static inline double sign_of(double x, double y) { return y < 0.0 ? -fabs(x) : fabs(x); }  // CSysSolve::Sign

static void apply_givens(double s, double c, double& h1, double& h2) {
  const double temp = c * h1 + s * h2;
  h2 = c * h2 - s * h1;
  h1 = temp;
}
static void generate_givens(double& dx, double& dy, double& s, double& c) {
  if ((dx == 0.0) && (dy == 0.0)) {
    c = 1.0;
    s = 0.0;
  } else if (fabs(dy) > fabs(dx)) {
    const double tmp = dx / dy;
    dx = sqrt(1.0 + tmp * tmp);
    s = sign_of(1.0 / dx, dy);
    c = tmp * s;
  } else if (fabs(dy) <= fabs(dx)) {
    const double tmp = dy / dx;
    dy = sqrt(1.0 + tmp * tmp);
    c = sign_of(1.0 / dy, dx);
    s = tmp * c;
  } else {
    dx = 0.0;
    dy = 0.0;
    c = 1.0;
    s = 0.0;
  }
  dx = fabs(dx * dy);
  dy = 0.0;
}

static unsigned long fgmres(su2b200_ctx* c, LinSys& ls, int prec, double tol, int m, const double* b, double* x,
                            bool xIsZero, double* residual) {
  SU2B_REQUIRE(m >= 1, "Number of linear solver iterations must be greater than 0.");
  SU2B_REQUIRE(m <= 1000, "FGMRES subspace is too large (>1000).");
  const double eps = 1e-14;  // linSolEpsilon<double>
  const size_t n = ls.vecLen();
  // workspace: V[0..m] at pool slots [0, m], Z[0..m] at [m+1, 2m+1]
  auto V = [&](int i) -> double* { return ls.vec(i).p; };
  auto Z = [&](int i) -> double* { return ls.vec(m + 1 + i).p; };
  std::vector<double> g(m + 1, 0.0), sn(m + 1, 0.0), cs(m + 1, 0.0), y(m, 0.0);
  std::vector<double> H((size_t)(m + 1) * m, 0.0);
  auto Hm = [&](int r, int col) -> double& { return H[(size_t)r * m + col]; };

  double norm0 = sqrt(vec_dot(c, ls, b, b));
  if (!xIsZero) {
    linsys_spmv(c, ls, x, V(0));
    vec_axpy(c, n, -1.0, b, V(0));
  } else {
    vec_set_scaled(c, n, -1.0, b, V(0));
  }
  double beta = xIsZero ? norm0 : sqrt(vec_dot(c, ls, V(0), V(0)));
  if (beta < tol * norm0 || beta < eps) {
    *residual = beta;
    return 0;
  }
  SU2B_LAUNCH(c, k_div, vgrid(n), 256, n, -beta, V(0));
  g[0] = beta;
  int i = 0;
  for (i = 0; i < m; i++) {
    if (beta < tol * norm0) break;
    linsys_apply_prec(c, ls, prec, V(i), Z(i));
    linsys_spmv(c, ls, Z(i), V(i + 1));
    // ModGramSchmidt(i, H, V)
    {
      const double reorth = 0.98;
      double* w = V(i + 1);
      double nrm = vec_dot(c, ls, w, w);
      double thr = nrm * reorth;
      if ((nrm <= 0.0) || (nrm != nrm)) throw Error("FGMRES orthogonalization failed, linear solver diverged.");
      for (int k = 0; k < i + 1; k++) {
        double prod = vec_dot(c, ls, w, V(k));
        double h_ki = prod;
        vec_axpy(c, n, -prod, V(k), w);
        if (prod * prod > thr) {
          prod = vec_dot(c, ls, w, V(k));
          h_ki += prod;
          vec_axpy(c, n, -prod, V(k), w);
        }
        Hm(k, i) = h_ki;
        nrm -= h_ki * h_ki;
        nrm = std::max(nrm, 0.0);
        thr = nrm * reorth;
      }
      nrm = sqrt(vec_dot(c, ls, w, w));
      Hm(i + 1, i) = nrm;
      SU2B_LAUNCH(c, k_div, vgrid(n), 256, n, nrm, w);
    }
    for (int k = 0; k < i; k++) apply_givens(sn[k], cs[k], Hm(k, i), Hm(k + 1, i));
    generate_givens(Hm(i, i), Hm(i + 1, i), sn[i], cs[i]);
    apply_givens(sn[i], cs[i], g[i], g[i + 1]);
    beta = fabs(g[i + 1]);
  }
  // SolveReduced(i, H, g, y)
  for (int k = 0; k < i; k++) y[k] = g[k];
  for (int k = i - 1; k >= 0; k--) {
    y[k] /= Hm(k, k);
    for (int j = k - 1; j >= 0; j--) y[j] -= Hm(j, k) * y[k];
  }
  // x += sum y_k Z_k, four terms per pass (LinearCombination with inc=true)
  for (int k = 0; k < i; k += 4) {
    LinComb4 lc;
    lc.cnt = std::min(4, i - k);
    for (int t = 0; t < 4; ++t) {
      lc.v[t] = Z(std::min(k + t, i - 1));
      lc.w[t] = (t < lc.cnt) ? y[k + t] : 0.0;
    }
    SU2B_LAUNCH(c, k_lincomb_inc, vgrid(n), 256, n, lc, x);
  }
  *residual = beta / norm0;
  return (unsigned long)i;
}

static unsigned long bcgstab(su2b200_ctx* c, LinSys& ls, int prec, double tol, int m, const double* b, double* x,
                             bool xIsZero, double* residual) {
  SU2B_REQUIRE(m >= 1, "Number of linear solver iterations must be greater than 0.");
  const double eps = 1e-14;
  const size_t n = ls.vecLen();
  double* A_x = ls.vec(0).p;
  double* r_0 = ls.vec(1).p;
  double* r = ls.vec(2).p;
  double* p = ls.vec(3).p;
  double* v = ls.vec(4).p;
  double* z = ls.vec(5).p;
  if (!xIsZero) {
    linsys_spmv(c, ls, x, A_x);
    vec_copy(c, n, b, r);
    vec_axpy(c, n, -1.0, A_x, r);
  } else {
    vec_copy(c, n, b, r);
  }
  double norm0 = sqrt(vec_dot(c, ls, b, b));
  double norm_r = xIsZero ? norm0 : sqrt(vec_dot(c, ls, r, r));
  if ((norm_r < tol * norm0) || (norm_r < eps)) {
    *residual = norm_r;  // reference leaves `residual` untouched here; report the norm
    return 0;
  }
  double alpha = 1.0, omega = 1.0, rho = 1.0, rho_prime = 1.0;
  SU2B_CUDA(cudaMemsetAsync(p, 0, n * sizeof(double), c->stream));
  SU2B_CUDA(cudaMemsetAsync(v, 0, n * sizeof(double), c->stream));
  vec_copy(c, n, r, r_0);
  int i = 0;
  for (i = 0; i < m; i++) {
    rho_prime = rho;
    rho = vec_dot(c, ls, r, r_0);
    const double beta = (rho / rho_prime) * (alpha / omega);
    SU2B_LAUNCH(c, k_bcg_p, vgrid(n), 256, n, beta, omega, v, r, p);
    linsys_apply_prec(c, ls, prec, p, z);
    linsys_spmv(c, ls, z, v);
    const double r_0_v = vec_dot(c, ls, r_0, v);
    alpha = rho / r_0_v;
    vec_axpy(c, n, alpha, z, x);
    vec_axpy(c, n, -alpha, v, r);
    linsys_apply_prec(c, ls, prec, r, z);
    linsys_spmv(c, ls, z, A_x);
    omega = vec_dot(c, ls, A_x, A_x);
    if (omega == 0.0) break;
    omega = vec_dot(c, ls, A_x, r) / omega;
    vec_axpy(c, n, omega, z, x);
    vec_axpy(c, n, -omega, A_x, r);
    norm_r = sqrt(vec_dot(c, ls, r, r));
    if (norm_r < tol * norm0) break;
  }
  *residual = norm_r / norm0;
  return (unsigned long)i;
}

// CSysSolve::Solve (CSysSolve.cpp:1318-1531): builds the preconditioner, runs the chosen solver with x0 = 0.
unsigned long linsys_solve(su2b200_ctx* c, LinSys& ls, int solver, int prec, double tol, int max_iter, int restart,
                           const double* b, double* x, double* residual) {
  linsys_build_prec(c, ls, prec);
  unsigned long it = 0;
  double res = 0.0;
  switch (solver) {
    case SU2B200_SOLVER_FGMRES:
      it = fgmres(c, ls, prec, tol, max_iter, b, x, true, &res);
      break;
    case SU2B200_SOLVER_RESTARTED_FGMRES: {
      // RFGMRES_LinSolver (CSysSolve.cpp:616-639)
      const int rf = std::max(1, restart);
      bool xz = true;
      res = 1.0;
      while (it < (unsigned long)max_iter) {
        const int mm = std::min<long>(rf, max_iter - (long)it);
        const unsigned long k = fgmres(c, ls, prec, tol, mm, b, x, xz, &res);
        it += k;
        if (res <= tol || k < (unsigned long)mm || k == 0) break;
        xz = false;
      }
    } break;
    case SU2B200_SOLVER_BCGSTAB:
      it = bcgstab(c, ls, prec, tol, max_iter, b, x, true, &res);
      break;
    default:
      throw Error("unknown linear solver");
  }
  linsys_check_sweeps(c, ls);
  if (residual) *residual = res;
  return it;
}

}  // namespace su2b200
