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


// ---- ModGramSchmidt (CSysSolve.cpp:179-240) with the scalars kept on the device -----------------------------------
// The reference alternates dot(w, v_k) and w -= h v_k, each dot a global reduction the host waits for. Here the axpy of
// step k-1 is fused into the pass that computes the dot of step k, the coefficients stay in device memory (sc[]), the
// 0.98 re-orthogonalisation test is evaluated on the device (the second projection's kernels exit at once when it is
// not needed), and the host reads the finished Hessenberg column once per Krylov iteration. Values are bit-identical
// to the unfused sequence (same elementwise fma, same two-stage reduction order).
//   sc[0] coefficient of the pending axpy, sc[1] running norm^2, sc[2] threshold, sc[3] re-orthogonalise flag,
//   sc[4] failure flag, sc[5] final norm, sc[6] reduction result, sc[8+k] = H(k, i)
constexpr int MGS_BLOCKS = 148 * 4;
__global__ void __launch_bounds__(256) k_mgs_dot(size_t n_owned, size_t n_all, double* __restrict__ w,
                                                  const double* __restrict__ vprev, const double* __restrict__ vk,
                                                  const double* __restrict__ sc, int need_flag,
                                                  double* __restrict__ part) {
  if (need_flag && sc[3] == 0.0) return;
  const double coef = vprev ? sc[0] : 0.0;
  // four independent element streams per thread (loads of a round issued together), fixed summation order
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  double acc[4] = {0.0, 0.0, 0.0, 0.0};
  for (; i + 3 * stride < n_all; i += 4 * stride) {
    double wi[4], pv[4], kv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) wi[u] = w[i + u * stride];
#pragma unroll
    for (int u = 0; u < 4; ++u) pv[u] = vprev ? vprev[i + u * stride] : 0.0;
#pragma unroll
    for (int u = 0; u < 4; ++u) kv[u] = vk ? vk[i + u * stride] : 0.0;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (vprev) { wi[u] += -coef * pv[u]; w[i + u * stride] = wi[u]; }
      if (i + u * stride < n_owned) acc[u] += wi[u] * (vk ? kv[u] : wi[u]);
    }
  }
  for (; i < n_all; i += stride) {
    double wi = w[i];
    if (vprev) { wi += -coef * vprev[i]; w[i] = wi; }
    if (i < n_owned) acc[0] += wi * (vk ? vk[i] : wi);
  }
  double s = (acc[0] + acc[1]) + (acc[2] + acc[3]);
  __shared__ double sm[32];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const int wp = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) sm[wp] = s;
  __syncthreads();
  if (wp == 0) {
    s = (l < (blockDim.x >> 5)) ? sm[l] : 0.0;
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (l == 0) part[blockIdx.x] = s;
  }
}
__global__ void k_mgs_stage2(int nb, const double* __restrict__ part, double* __restrict__ sc, int need_flag) {
  if (need_flag && sc[3] == 0.0) { if (threadIdx.x == 0) sc[6] = 0.0; return; }
  double s = 0.0;
  for (int i = threadIdx.x; i < nb; i += blockDim.x) s += part[i];
  __shared__ double sm[32];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const int wp = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) sm[wp] = s;
  __syncthreads();
  if (wp == 0) {
    s = (l < (blockDim.x >> 5)) ? sm[l] : 0.0;
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (l == 0) sc[6] = s;
  }
}
// mode 0: nrm = w.w ; 1: first projection on v_k ; 2: second projection (if flagged) ; 3: final norm
__device__ __forceinline__ void mgs_logic(double* __restrict__ sc, int mode, int k, double prod);
__global__ void k_mgs_logic(double* __restrict__ sc, int mode, int k) { mgs_logic(sc, mode, k, sc[6]); }
// Multi-rank form of stage 2: block sum, allreduce through the peer mailboxes, and the scalar logic, in ONE launch
// (instead of k_mgs_stage2 + ncclAllReduce + k_mgs_logic).  The skipped second projection skips the exchange too: the
// flag sc[3] is the same on every rank, because every rank adds the ranks' sums in the same order.
__global__ void __launch_bounds__(256) k_mgs_stage2_mbox(int nb, const double* __restrict__ part, double* __restrict__ sc,
                                                         int need_flag, MailboxView mb, int mode, int k) {
  if (need_flag && sc[3] == 0.0) {
    if (threadIdx.x == 0) mgs_logic(sc, mode, k, 0.0);
    return;
  }
  double s = 0.0;
  for (int i = threadIdx.x; i < nb; i += blockDim.x) s += part[i];
  __shared__ double sm[32];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const int wp = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) sm[wp] = s;
  __syncthreads();
  if (wp == 0) {
    s = (l < (blockDim.x >> 5)) ? sm[l] : 0.0;
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (l == 0) sm[0] = s;
  }
  __syncthreads();
  bool timed_out = false;
  const double total = mb.nranks > 1 ? mailbox_allreduce_block(mb, sm[0], &timed_out) : sm[0];
  if (threadIdx.x == 0) {
    sc[6] = total;
    mgs_logic(sc, mode, k, timed_out ? __longlong_as_double(0x7ff8000000000000LL) : total);
    if (timed_out) sc[4] = 2.0;
  }
}
__device__ __forceinline__ void mgs_logic(double* __restrict__ sc, int mode, int k, double prod) {
  const double reorth = 0.98;
  if (mode == 0) {
    sc[1] = prod; sc[2] = prod * reorth; sc[0] = 0.0; sc[3] = 0.0;
    sc[4] = ((prod <= 0.0) || (prod != prod)) ? 1.0 : 0.0;
  } else if (mode == 1) {
    sc[8 + k] = prod;
    sc[0] = prod;
    if (prod * prod > sc[2]) {
      sc[3] = 1.0;
    } else {
      sc[3] = 0.0;
      double nrm = sc[1] - prod * prod;
      nrm = fmax(nrm, 0.0);
      sc[1] = nrm; sc[2] = nrm * reorth;
    }
  } else if (mode == 2) {
    if (sc[3] == 0.0) return;
    const double h = sc[8 + k] + prod;
    sc[8 + k] = h;
    sc[0] = prod;
    sc[3] = 0.0;
    double nrm = sc[1] - h * h;
    nrm = fmax(nrm, 0.0);
    sc[1] = nrm; sc[2] = nrm * reorth;
  } else {
    const double nrm = sqrt(prod);
    sc[5] = nrm;
    sc[8 + k] = nrm;
    // a NaN here (a poisoned basis, e.g. the sentinel of a sweep whose watchdog fired) must stop the solve now: the
    // reference would only notice at the start of the next orthogonalisation, after scaling the vector by it
    if (prod != prod) sc[4] = 1.0;
  }
}
__global__ void k_div_dev(size_t n, const double* __restrict__ sc, double* __restrict__ x) {
  const double d = sc[5];
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    x[i] /= d;
}

// orthogonalise w = V(i+1) against V(0..i); fills H(0..i+1, i) through `hcol`
template <class VecFn>
static void mod_gram_schmidt_device(su2b200_ctx* c, LinSys& ls, int i, VecFn V, double* hcol) {
  const size_t n_all = ls.vecLen(), n_owned = (size_t)ls.Pd * ls.nVar;
  if (ls.scalars.n < 4096) {
    ls.scalars.alloc(4096);
    ls.scalars.zero(c->stream);
  }
  if (!ls.h_mgs) SU2B_CUDA(cudaMallocHost(&ls.h_mgs, 1100 * sizeof(double)));
  double* part = ls.scalars.p + 1024;  // [MGS_BLOCKS] (the first DOT_BLOCKS + 64 entries belong to vec_dot)
  double* sc = ls.scalars.p + 2048;    // [8 + m + 2]
  const int nb = (int)std::min<size_t>((n_all + 255) / 256, MGS_BLOCKS);
  double* w = V(i + 1);
  auto dot = [&](const double* vprev, const double* vk, int need_flag, int mode, int k) {
    {
      ProfScope prof_(c, PROF_DOT);
      SU2B_LAUNCH(c, k_mgs_dot, nb, 256, n_owned, n_all, w, vprev, vk, (const double*)sc, need_flag, part);
      MailboxView mb{nullptr, nullptr, 0, 1};
      if (c->nranks <= 1 || mailbox_view(c, &mb)) {  // one rank: the same kernel without the exchange
        SU2B_LAUNCH(c, k_mgs_stage2_mbox, 1, 256, nb, (const double*)part, sc, need_flag, mb, mode, k);
      } else {
        SU2B_LAUNCH(c, k_mgs_stage2, 1, 256, nb, (const double*)part, sc, need_flag);
        allreduce_dev_sum(c, sc + 6, 1);
        SU2B_LAUNCH(c, k_mgs_logic, 1, 1, sc, mode, k);
      }
    }
  };
  dot(nullptr, nullptr, 0, 0, 0);  // nrm = w.w
  for (int k = 0; k <= i; ++k) {
    dot(k ? V(k - 1) : nullptr, V(k), 0, 1, k);  // pending axpy with v_{k-1}, then w.v_k
    dot(V(k), V(k), 1, 2, k);                    // second projection on v_k, only if the device flag is set
  }
  dot(V(i), nullptr, 0, 3, i + 1);               // last axpy, then |w|
  {
    ProfScope prof_(c, PROF_VECOPS);
    SU2B_LAUNCH(c, k_div_dev, (int)std::min<size_t>((n_all + 255) / 256, 148 * 8), 256, n_all, (const double*)sc, w);
  }
  SU2B_CUDA(cudaMemcpyAsync(ls.h_mgs, sc, (size_t)(8 + i + 2) * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  if (ls.h_mgs[4] == 2.0)
    throw Error("mailbox allreduce timed out: a peer rank stopped or the ranks disagree on the sequence of reductions");
  if (ls.h_mgs[4] != 0.0) throw Error("FGMRES orthogonalization failed, linear solver diverged.");
  for (int k = 0; k <= i + 1; ++k) hcol[k] = ls.h_mgs[8 + k];
}

// z = M^-1 v, w = A z: the pair every Krylov iteration runs.  The ILU backward sweep is asked for the upper part of the
// product on its way (LinSys::upper_prod; used when the upper factor is the Jacobian's own upper triangle), the product
// then reads 4 of the 7 blocks of a hexahedral row.  Summation order of (A z)_i: lower, diagonal, (upper sum), halo
// columns -- the reference's RowProduct adds the upper products one by one; the difference is one rounding per row.
void linsys_prec_then_product(su2b200_ctx* c, LinSys& ls, int prec, const double* v, double* z, double* w) {
  ls.want_upper_prod = (prec == SU2B200_PREC_ILU);
  try {
    linsys_apply_prec(c, ls, prec, v, z);
  } catch (...) {
    ls.want_upper_prod = false;
    throw;
  }
  ls.want_upper_prod = false;
  linsys_spmv_exchange_x(c, ls, z, w);
  ls.upper_prod_of = nullptr;
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
    linsys_prec_then_product(c, ls, prec, V(i), Z(i), V(i + 1));
    // ModGramSchmidt(i, H, V)
    {
      std::vector<double> hcol(i + 2, 0.0);
      mod_gram_schmidt_device(c, ls, i, V, hcol.data());
      for (int k = 0; k <= i + 1; ++k) Hm(k, i) = hcol[k];
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
    linsys_prec_then_product(c, ls, prec, p, z, v);
    const double r_0_v = vec_dot(c, ls, r_0, v);
    alpha = rho / r_0_v;
    vec_axpy(c, n, alpha, z, x);
    vec_axpy(c, n, -alpha, v, r);
    linsys_prec_then_product(c, ls, prec, r, z, A_x);
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
  // the whole Krylov workspace now: growing the pool inside the loop, the first time a solve needs one more iteration
  // than any before it, costs two cudaMallocs + fills in the middle of a nonlinear iteration (a 10-20 ms hiccup)
  {
    const int mm = solver == SU2B200_SOLVER_RESTARTED_FGMRES ? std::max(1, std::min(restart, max_iter)) : max_iter;
    const int nvec = solver == SU2B200_SOLVER_BCGSTAB ? 6 : 2 * (std::min(mm, 1000) + 1);
    for (int k = 0; k < nvec; ++k) (void)ls.vec((size_t)k);
  }
  // e3: inside the solve the preconditioners and products leave their trailing halo exchange to the overlapped product
  // (SU2B200_OVERLAP=0: every operator exchanges its own output in-stream, as the reference's CSysMatrixComms does)
  static const bool overlap = !(getenv("SU2B200_OVERLAP") && atoi(getenv("SU2B200_OVERLAP")) == 0);
  struct DeferGuard {
    su2b200_ctx* c;
    ~DeferGuard() { c->defer_halo = false; }
  } guard_{c};
  c->defer_halo = overlap && c->nranks > 1;
  try {
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
  } catch (...) {
    // a fired sweep watchdog is the root cause of whatever failed downstream (NaN norms): report THAT, and clear the
    // abort flag so that the next build/apply on this system does not stop at once
    linsys_check_sweeps(c, ls);
    throw;
  }
  linsys_check_sweeps(c, ls);
  if (residual) *residual = res;
  return it;
}

}  // namespace su2b200
