// a16-a20, a23: block-CSR linear algebra on the device.
//   CSysMatrix<T>::MatrixVectorProduct (Common/src/linear_algebra/CSysMatrix.cpp:617-645, RowProduct CSysMatrix.inl:173-181)
//   Build/ComputeJacobiPreconditioner (:648-670), MatrixInverse (:548-601)
//   Build/ComputeILUPreconditioner (:673-823), ComputeLU_SGSPreconditioner (:826-888)
//   CSysVector<T> BLAS-1 (Common/include/linear_algebra/CSysVector.hpp:292-378)
// Layout: blocks row-major [nnz][nVar][nVar] fp64 exactly like CSysMatrix::matrix; int32 row_ptr/col_ind/dia_ptr.
//
// Warp mapping used by every block kernel: LPB = nVar*nVar lanes own one scalar of a block (q = r*nVar+c),
// BPW = 32/LPB blocks are processed per warp step. Block loads are therefore contiguous 8*LPB-byte reads.
#include "internal.hpp"
#include "kernels.hpp"
#include <algorithm>
#include <mutex>

namespace su2b200 {

static inline int nblk(long n, int bs) { return (int)((n + bs - 1) / bs); }
static constexpr unsigned FULL = 0xffffffffu;

LinSys::~LinSys() {
  for (auto* v : work) delete v;
  if (h_scalars) cudaFreeHost(h_scalars);
  if (h_mgs) cudaFreeHost(h_mgs);
}
dvec<double>& LinSys::vec(size_t i) {
  while (work.size() <= i) work.push_back(new dvec<double>());
  if (work[i]->n != vecLen()) {
    work[i]->alloc(vecLen());
    work[i]->zero();
  }
  return *work[i];
}

// ------------------------------------------------------------------------------------------------
// BLAS-1
// ------------------------------------------------------------------------------------------------
__global__ void k_axpy(size_t n, double alpha, const double* __restrict__ x, double* __restrict__ y) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    y[i] += alpha * x[i];
}
__global__ void k_axpby(size_t n, double alpha, const double* __restrict__ x, double beta, double* __restrict__ y) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    y[i] = alpha * x[i] + beta * y[i];
}
__global__ void k_scale(size_t n, double alpha, double* __restrict__ x) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    x[i] *= alpha;
}
__global__ void k_set_scaled(size_t n, double alpha, const double* __restrict__ x, double* __restrict__ y) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    y[i] = alpha * x[i];
}

static inline int vec_grid(size_t n) { return (int)std::min<size_t>((n + 255) / 256, 148 * 8); }

void vec_axpy(su2b200_ctx* c, size_t n, double alpha, const double* x, double* y) {
  ProfScope prof_(c, PROF_VECOPS);
  SU2B_LAUNCH(c, k_axpy, vec_grid(n), 256, n, alpha, x, y);
}
void vec_axpby(su2b200_ctx* c, size_t n, double alpha, const double* x, double beta, double* y) {
  ProfScope prof_(c, PROF_VECOPS);
  SU2B_LAUNCH(c, k_axpby, vec_grid(n), 256, n, alpha, x, beta, y);
}
void vec_scale(su2b200_ctx* c, size_t n, double alpha, double* x) { SU2B_LAUNCH(c, k_scale, vec_grid(n), 256, n, alpha, x); }
void vec_set_scaled(su2b200_ctx* c, size_t n, double alpha, const double* x, double* y) {
  ProfScope prof_(c, PROF_VECOPS);
  SU2B_LAUNCH(c, k_set_scaled, vec_grid(n), 256, n, alpha, x, y);
}
void vec_copy(su2b200_ctx* c, size_t n, const double* x, double* y) {
  ProfScope prof_(c, PROF_VECOPS);
  SU2B_CUDA(cudaMemcpyAsync(y, x, n * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
}

// CSysVector::dot (CSysVector.hpp:335-366): sum over the owned (domain) entries, deterministic two-stage
// reduction with warp shuffles; stage 2 is a single block so the result does not depend on scheduling.
constexpr int DOT_BLOCKS = 148 * 4;
__global__ void k_dot_stage1(size_t n, const double* __restrict__ a, const double* __restrict__ b,
                             double* __restrict__ part) {
  double s = 0.0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    s += a[i] * b[i];
  __shared__ double sm[32];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(FULL, s, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) sm[w] = s;
  __syncthreads();
  if (w == 0) {
    s = (l < (blockDim.x >> 5)) ? sm[l] : 0.0;
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(FULL, s, o);
    if (l == 0) part[blockIdx.x] = s;
  }
}
__global__ void k_dot_stage2(int nb, const double* __restrict__ part, double* __restrict__ out) {
  double s = 0.0;
  for (int i = threadIdx.x; i < nb; i += blockDim.x) s += part[i];
  __shared__ double sm[32];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(FULL, s, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) sm[w] = s;
  __syncthreads();
  if (w == 0) {
    s = (l < (blockDim.x >> 5)) ? sm[l] : 0.0;
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(FULL, s, o);
    if (l == 0) out[0] = s;
  }
}

static void ensure_scalars(LinSys& ls) {
  if (ls.scalars.n < 4096) ls.scalars.alloc(4096);  // [0,1024) vec_dot, [1024,4096) device Gram-Schmidt (krylov.cu)
  if (!ls.h_scalars) SU2B_CUDA(cudaMallocHost(&ls.h_scalars, 64 * sizeof(double)));
}

double vec_dot(su2b200_ctx* c, LinSys& ls, const double* a, const double* b) {
  ProfScope prof_(c, PROF_DOT);
  ensure_scalars(ls);
  const size_t n = (size_t)ls.Pd * ls.nVar;
  const int nb = (int)std::min<size_t>((n + 255) / 256, DOT_BLOCKS);
  SU2B_LAUNCH(c, k_dot_stage1, nb, 256, n, a, b, ls.scalars.p);
  SU2B_LAUNCH(c, k_dot_stage2, 1, 256, nb, ls.scalars.p, ls.scalars.p + DOT_BLOCKS);
  SU2B_CUDA(cudaMemcpyAsync(ls.h_scalars, ls.scalars.p + DOT_BLOCKS, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  double v = ls.h_scalars[0];
  if (c->nranks > 1) v = allreduce_sum(c, v);
  return v;
}

// ------------------------------------------------------------------------------------------------
// SpMV: one warp per block-row.
// ------------------------------------------------------------------------------------------------
#ifndef SPMV_UNROLL
#define SPMV_UNROLL 1
#endif
template <int NV>
__global__ void __launch_bounds__(256) k_spmv(int Pd, const int* __restrict__ row_ptr, const int* __restrict__ col_ind,
                                              const double* __restrict__ A, const double* __restrict__ x,
                                              double* __restrict__ y) {
  constexpr int LPB = NV * NV;
  constexpr int BPW = 32 / LPB;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= Pd) return;
  const int g = lane / LPB, q = lane % LPB, cc = q % NV;
  const bool active = g < BPW;
  const int kb = row_ptr[warp], ke = row_ptr[warp + 1];
  double acc = 0.0;
  if (active) {
#if SPMV_UNROLL > 1
    // SPMV_UNROLL entries per trip with all their loads issued before the first use: twice (or four times) the bytes in
    // flight per warp; the accumulation order is unchanged
    int k = kb + g;
    for (; k + (SPMV_UNROLL - 1) * BPW < ke; k += SPMV_UNROLL * BPW) {
      int col[SPMV_UNROLL];
      double a[SPMV_UNROLL], xv[SPMV_UNROLL];
#pragma unroll
      for (int u = 0; u < SPMV_UNROLL; ++u) col[u] = col_ind[k + u * BPW];
#pragma unroll
      for (int u = 0; u < SPMV_UNROLL; ++u) a[u] = A[(size_t)(k + u * BPW) * LPB + q];
#pragma unroll
      for (int u = 0; u < SPMV_UNROLL; ++u) xv[u] = x[(size_t)col[u] * NV + cc];
#pragma unroll
      for (int u = 0; u < SPMV_UNROLL; ++u) acc += a[u] * xv[u];
    }
    for (; k < ke; k += BPW) {
#else
    for (int k = kb + g; k < ke; k += BPW) {
#endif
#ifdef SPMV_STREAM
      // matrix entries are used once per product: evict-first, the L2 is kept for the gathered x
      const int col = __ldcs(col_ind + k);
      acc += __ldcs(A + (size_t)k * LPB + q) * x[(size_t)col * NV + cc];
#else
      const int col = col_ind[k];
      acc += A[(size_t)k * LPB + q] * x[(size_t)col * NV + cc];
#endif
    }
  }
  // reduce over the column index within a block row (lanes q = r*NV .. r*NV+NV-1 are adjacent)
  double s = acc;
#pragma unroll
  for (int o = 1; o < NV; ++o) {
    const double t = __shfl_down_sync(FULL, acc, o);
    s += t;
  }
  // s is valid on lanes with cc == 0; now add the BPW sub-groups
#pragma unroll
  for (int gg = 1; gg < BPW; ++gg) {
    const double t = __shfl_down_sync(FULL, s, gg * LPB);
    if (g == 0) s += t;
  }
  if (active && g == 0 && cc == 0) y[(size_t)warp * NV + q / NV] = s;
}

// nVar lanes per block row, 32/nVar rows per warp: lane r owns component r of its row's product and walks the row's
// blocks in the reference's order (RowProduct, CSysMatrix.inl:173-181: blocks outer, columns inner, one accumulation
// chain per component). A warp reads the blocks of 6
// consecutive rows = one contiguous 8 KB stretch of the matrix; nothing is reduced across lanes. This is the mapping
// that took the wide levels of the ILU sweeps to 85 % of the HBM peak (precond_poll.cu: k_poll_sweep_rows).
// Measured on 185^3 (ms): blocks per step / resident blocks per SM = 7/2 2.02, 4/4 1.83, 3/5 1.77, 2/6 1.58, 1/8 1.52,
// 1/12 1.49 — occupancy beats loads in flight per lane; the old warp-per-row kernel: 2.31.
#ifndef SPMV_ROWS_CH
#define SPMV_ROWS_CH 1
#endif
#ifndef SPMV_ROWS_MIN_BLOCKS
#define SPMV_ROWS_MIN_BLOCKS 12
#endif
template <int NV>
__global__ void __launch_bounds__(128, SPMV_ROWS_MIN_BLOCKS) k_spmv_rows(int Pd, const int* __restrict__ row_ptr,
                                                                         const int* __restrict__ col_ind,
                                                                         const double* __restrict__ A,
                                                                         const double* __restrict__ x, double* __restrict__ y) {
  constexpr int LPB = NV * NV, RPW = 32 / NV, CH = SPMV_ROWS_CH;
  const int lane = threadIdx.x & 31, sub = lane / NV, r = lane % NV;
  const long wg = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long row = wg * RPW + sub;
  const bool on = sub < RPW && row < Pd;
  const int kb = on ? row_ptr[row] : 0, ke = on ? row_ptr[row + 1] : 0;
  double acc = 0.0;
  for (int k0 = kb; k0 < ke; k0 += CH) {
    int col[CH];
    double a[CH][NV], xv[CH][NV];
#pragma unroll
    for (int u = 0; u < CH; ++u) col[u] = k0 + u < ke ? col_ind[k0 + u] : 0;
#pragma unroll
    for (int u = 0; u < CH; ++u)
#pragma unroll
      for (int c = 0; c < NV; ++c) a[u][c] = k0 + u < ke ? A[(size_t)(k0 + u) * LPB + r * NV + c] : 0.0;
#pragma unroll
    for (int u = 0; u < CH; ++u)
#pragma unroll
      for (int c = 0; c < NV; ++c) xv[u][c] = k0 + u < ke ? x[(size_t)col[u] * NV + c] : 0.0;
#pragma unroll
    for (int u = 0; u < CH; ++u)
      if (k0 + u < ke) {
#pragma unroll
        for (int c = 0; c < NV; ++c) acc += a[u][c] * xv[u][c];
      }
  }
  if (on) y[(size_t)row * NV + r] = acc;
}

// scalar CSR (nVar == 1): 8 lanes per row
__global__ void __launch_bounds__(256) k_spmv_scalar(int Pd, const int* __restrict__ row_ptr,
                                                     const int* __restrict__ col_ind, const double* __restrict__ A,
                                                     const double* __restrict__ x, double* __restrict__ y) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int row = t >> 3, l = t & 7;
  double acc = 0.0;
  if (row < Pd) {
    for (int k = row_ptr[row] + l; k < row_ptr[row + 1]; k += 8) acc += A[k] * x[col_ind[k]];
  }
  acc += __shfl_down_sync(FULL, acc, 4, 8);
  acc += __shfl_down_sync(FULL, acc, 2, 8);
  acc += __shfl_down_sync(FULL, acc, 1, 8);
  if (row < Pd && l == 0) y[row] = acc;
}

// y = A x for an x that just left the ILU backward sweep together with u_i = sum_{j>i, owned} A_ij x_j (LinSys::upper_prod):
// only the lower triangle, the diagonal and the halo columns are multiplied here.  Same lane mapping as k_spmv_rows.
template <int NV>
__global__ void __launch_bounds__(128, SPMV_ROWS_MIN_BLOCKS) k_spmv_rows_rest(int Pd, const int* __restrict__ row_ptr,
                                                                              const int* __restrict__ dia_ptr,
                                                                              const int* __restrict__ halo_ptr,
                                                                              const int* __restrict__ col_ind,
                                                                              const double* __restrict__ A,
                                                                              const double* __restrict__ x,
                                                                              const double* __restrict__ upper,
                                                                              double* __restrict__ y) {
  constexpr int LPB = NV * NV, RPW = 32 / NV;
  const int lane = threadIdx.x & 31, sub = lane / NV, r = lane % NV;
  const long wg = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long row = wg * RPW + sub;
  const bool on = sub < RPW && row < Pd;
  const int kb = on ? row_ptr[row] : 0, kd = on ? dia_ptr[row] + 1 : 0;
  const int ke = on ? row_ptr[row + 1] : 0, kh = (on && halo_ptr) ? halo_ptr[row] : ke;
  const double up = on ? upper[(size_t)row * NV + r] : 0.0;
  double acc = 0.0;
  for (int k = kb; k < kd; ++k) {
    const int col = col_ind[k];
    double a[NV], xv[NV];
#pragma unroll
    for (int c = 0; c < NV; ++c) a[c] = A[(size_t)k * LPB + r * NV + c];
#pragma unroll
    for (int c = 0; c < NV; ++c) xv[c] = x[(size_t)col * NV + c];
#pragma unroll
    for (int c = 0; c < NV; ++c) acc += a[c] * xv[c];
  }
  acc += up;
  for (int k = kh; k < ke; ++k) {
    const int col = col_ind[k];
#pragma unroll
    for (int c = 0; c < NV; ++c) acc += A[(size_t)k * LPB + r * NV + c] * x[(size_t)col * NV + c];
  }
  if (on) y[(size_t)row * NV + r] = acc;
}

static bool spmv_fused_upper(su2b200_ctx* c, LinSys& ls, const double* x, double* y) {
  static const bool enabled = !(getenv("SU2B200_FUSED_UPPER") && atoi(getenv("SU2B200_FUSED_UPPER")) == 0);
  if (!enabled || ls.nVar == 1 || ls.upper_prod_of != x || x == nullptr) return false;
  ls.upper_prod_of = nullptr;  // one product per apply: whoever changes x afterwards is not tracked
  const int* hp = nullptr;
  if (ls.P > ls.Pd) {
    if (!ls.rest_halo_ready) {
      std::vector<int> h(ls.Pd);
      for (int i = 0; i < ls.Pd; ++i) {
        int ke = ls.h_row_ptr[i + 1];
        while (ke > ls.h_dia_ptr[i] + 1 && ls.h_col_ind[ke - 1] >= ls.Pd) --ke;
        h[i] = ke;
      }
      ls.rest_halo_ptr.upload(h);
      ls.rest_halo_ready = true;
    }
    hp = ls.rest_halo_ptr.p;
  }
  const long warps = ((long)ls.Pd + 32 / ls.nVar - 1) / (32 / ls.nVar);
  if (ls.nVar == 2)
    SU2B_LAUNCH(c, k_spmv_rows_rest<2>, nblk(warps * 32, 128), 128, ls.Pd, ls.row_ptr, ls.dia_ptr, hp, ls.col_ind, ls.A.p, x,
                (const double*)ls.upper_prod.p, y);
  else if (ls.nVar == 4)
    SU2B_LAUNCH(c, k_spmv_rows_rest<4>, nblk(warps * 32, 128), 128, ls.Pd, ls.row_ptr, ls.dia_ptr, hp, ls.col_ind, ls.A.p, x,
                (const double*)ls.upper_prod.p, y);
  else if (ls.nVar == 5)
    SU2B_LAUNCH(c, k_spmv_rows_rest<5>, nblk(warps * 32, 128), 128, ls.Pd, ls.row_ptr, ls.dia_ptr, hp, ls.col_ind, ls.A.p, x,
                (const double*)ls.upper_prod.p, y);
  else
    return false;
  return true;
}

void linsys_spmv(su2b200_ctx* c, LinSys& ls, const double* x, double* y) {
  ProfScope prof_(c, PROF_SPMV);
  if (spmv_fused_upper(c, ls, x, y)) {
    if (c->nranks > 1 && !c->defer_halo) halo_exchange(c, y, ls.nVar);
    return;
  }
  const long threads = (long)ls.Pd * 32;
  static const int rowsK = getenv("SU2B200_SPMV_ROWS") ? atoi(getenv("SU2B200_SPMV_ROWS")) : 1;
  if (rowsK && ls.nVar > 1) {
    const long warps = ((long)ls.Pd + 32 / ls.nVar - 1) / (32 / ls.nVar);
    if (ls.nVar == 2) SU2B_LAUNCH(c, k_spmv_rows<2>, nblk(warps * 32, 128), 128, ls.Pd, ls.row_ptr, ls.col_ind, ls.A.p, x, y);
    else if (ls.nVar == 4) SU2B_LAUNCH(c, k_spmv_rows<4>, nblk(warps * 32, 128), 128, ls.Pd, ls.row_ptr, ls.col_ind, ls.A.p, x, y);
    else if (ls.nVar == 5) SU2B_LAUNCH(c, k_spmv_rows<5>, nblk(warps * 32, 128), 128, ls.Pd, ls.row_ptr, ls.col_ind, ls.A.p, x, y);
    else throw Error("linsys_spmv: unsupported nVar");
    if (c->nranks > 1 && !c->defer_halo) halo_exchange(c, y, ls.nVar);
    return;
  }
  switch (ls.nVar) {
    case 1: SU2B_LAUNCH(c, k_spmv_scalar, nblk((long)ls.Pd * 8, 256), 256, ls.Pd, ls.row_ptr, ls.col_ind, ls.A.p, x, y); break;
    case 2: SU2B_LAUNCH(c, k_spmv<2>, nblk(threads, 256), 256, ls.Pd, ls.row_ptr, ls.col_ind, ls.A.p, x, y); break;
    case 4: SU2B_LAUNCH(c, k_spmv<4>, nblk(threads, 256), 256, ls.Pd, ls.row_ptr, ls.col_ind, ls.A.p, x, y); break;
    case 5: SU2B_LAUNCH(c, k_spmv<5>, nblk(threads, 256), 256, ls.Pd, ls.row_ptr, ls.col_ind, ls.A.p, x, y); break;
    default: throw Error("linsys_spmv: unsupported nVar");
  }
  if (c->nranks > 1 && !c->defer_halo) halo_exchange(c, y, ls.nVar);
}

// e3. Product of a Krylov iteration, y = A x with x fresh from the preconditioner (owned entries only): the halos of x travel
// on comm_stream while EVERY owned row is multiplied; the rows with a halo column (a thin layer next to the rank
// boundary) read stale halos in that pass and are multiplied again once the exchange has landed.  The halos of y are not
// exchanged: no Krylov operation reads them (dots run over owned entries, the preconditioners over owned rows and columns).
template <int NV>
__global__ void __launch_bounds__(128) k_spmv_row_list(int n, const int* __restrict__ rows, const int* __restrict__ row_ptr,
                                                       const int* __restrict__ col_ind, const double* __restrict__ A,
                                                       const double* __restrict__ x, double* __restrict__ y) {
  // one warp per listed row; lane = (k, r): block column k of the row (strided), block row r... kept simple: the list is
  // a few per cent of the rows
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (w >= n) return;
  const int row = rows[w];
  const int kb = row_ptr[row], ke = row_ptr[row + 1];
  constexpr int LPB = NV * NV;
  double acc[NV];
#pragma unroll
  for (int r = 0; r < NV; ++r) acc[r] = 0.0;
  for (int k = kb + lane; k < ke; k += 32) {
    const int col = col_ind[k];
#pragma unroll
    for (int r = 0; r < NV; ++r)
#pragma unroll
      for (int cc = 0; cc < NV; ++cc) acc[r] += A[(size_t)k * LPB + r * NV + cc] * x[(size_t)col * NV + cc];
  }
#pragma unroll
  for (int r = 0; r < NV; ++r) {
    double v = acc[r];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(FULL, v, o);
    if (lane == 0) y[(size_t)row * NV + r] = v;
  }
}

static void build_halo_rows(LinSys& ls) {
  std::vector<int> rows;
  for (int i = 0; i < ls.Pd; ++i)
    for (int k = ls.h_row_ptr[i]; k < ls.h_row_ptr[i + 1]; ++k)
      if (ls.h_col_ind[k] >= ls.Pd) { rows.push_back(i); break; }
  ls.n_halo_rows = (int)rows.size();
  if (!rows.empty()) ls.halo_rows.upload(rows);
}

void linsys_spmv_exchange_x(su2b200_ctx* c, LinSys& ls, double* x, double* y) {
  if (c->nranks <= 1 || !c->defer_halo) {
    // the operator that produced x has exchanged it already
    linsys_spmv(c, ls, x, y);
    return;
  }
  if (ls.n_halo_rows < 0) build_halo_rows(ls);
  halo_exchange_start(c, x, ls.nVar);
  linsys_spmv(c, ls, x, y);
  halo_exchange_finish(c);
  if (ls.n_halo_rows > 0) {
    ProfScope prof_(c, PROF_SPMV);
    const int n = ls.n_halo_rows;
    switch (ls.nVar) {
      case 1: SU2B_LAUNCH(c, k_spmv_row_list<1>, nblk((long)n * 32, 128), 128, n, ls.halo_rows.p, ls.row_ptr, ls.col_ind, ls.A.p, x, y); break;
      case 2: SU2B_LAUNCH(c, k_spmv_row_list<2>, nblk((long)n * 32, 128), 128, n, ls.halo_rows.p, ls.row_ptr, ls.col_ind, ls.A.p, x, y); break;
      case 4: SU2B_LAUNCH(c, k_spmv_row_list<4>, nblk((long)n * 32, 128), 128, n, ls.halo_rows.p, ls.row_ptr, ls.col_ind, ls.A.p, x, y); break;
      case 5: SU2B_LAUNCH(c, k_spmv_row_list<5>, nblk((long)n * 32, 128), 128, n, ls.halo_rows.p, ls.row_ptr, ls.col_ind, ls.A.p, x, y); break;
      default: throw Error("linsys_spmv: unsupported nVar");
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Dense nVar x nVar helpers (per thread, registers/local memory)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void regularize_pivot(double& pivot) {
  const double eps = (double)1e-12f;  // `const float eps = 1e-12` in RegularizePivot (CSysMatrix.cpp:500-508)
  if (fabs(pivot) < eps) pivot = copysign(eps, pivot);
}

// CSysMatrix::MatrixInverse (CSysMatrix.cpp:548-601); A is destroyed.
template <int NV>
__device__ __forceinline__ void matrix_inverse(double* A, double* M) {
#pragma unroll
  for (int i = 0; i < NV; ++i)
#pragma unroll
    for (int j = 0; j < NV; ++j) M[i * NV + j] = (i == j) ? 1.0 : 0.0;
#pragma unroll
  for (int i = 1; i < NV; ++i) {
#pragma unroll
    for (int j = 0; j < i; ++j) {
      regularize_pivot(A[j * NV + j]);
      const double w = A[i * NV + j] / A[j * NV + j];
#pragma unroll
      for (int k = j; k < NV; ++k) A[i * NV + k] -= w * A[j * NV + k];
#pragma unroll
      for (int k = 0; k <= j; ++k) M[i * NV + k] -= w * M[j * NV + k];
    }
  }
#pragma unroll
  for (int i = NV - 1; i >= 0; --i) {
#pragma unroll
    for (int j = i + 1; j < NV; ++j)
#pragma unroll
      for (int k = 0; k < NV; ++k) M[i * NV + k] -= A[i * NV + j] * M[j * NV + k];
    regularize_pivot(A[i * NV + i]);
#pragma unroll
    for (int k = 0; k < NV; ++k) M[i * NV + k] /= A[i * NV + i];
  }
}

// CSysMatrix::Gauss_Elimination (CSysMatrix.cpp:510-545); A and v are modified, v holds the solution.
template <int NV>
__device__ __forceinline__ void gauss_elimination(double* A, double* v) {
#pragma unroll
  for (int i = 1; i < NV; ++i) {
#pragma unroll
    for (int j = 0; j < i; ++j) {
      regularize_pivot(A[j * NV + j]);
      const double w = A[i * NV + j] / A[j * NV + j];
#pragma unroll
      for (int k = j; k < NV; ++k) A[i * NV + k] -= w * A[j * NV + k];
      v[i] -= w * v[j];
    }
  }
#pragma unroll
  for (int i = NV - 1; i >= 0; --i) {
#pragma unroll
    for (int j = i + 1; j < NV; ++j) v[i] -= A[i * NV + j] * v[j];
    regularize_pivot(A[i * NV + i]);
    v[i] /= A[i * NV + i];
  }
}

// ------------------------------------------------------------------------------------------------
// Jacobi
// ------------------------------------------------------------------------------------------------
template <int NV>
__global__ void k_jacobi_build(int Pd, const int* __restrict__ dia_ptr, const double* __restrict__ A,
                               double* __restrict__ invM) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Pd) return;
  double B[NV * NV], M[NV * NV];
  const double* D = A + (size_t)dia_ptr[i] * NV * NV;
#pragma unroll
  for (int k = 0; k < NV * NV; ++k) B[k] = D[k];
  matrix_inverse<NV>(B, M);
#pragma unroll
  for (int k = 0; k < NV * NV; ++k) invM[(size_t)i * NV * NV + k] = M[k];
}

template <int NV>
__global__ void k_jacobi_apply(int Pd, const double* __restrict__ invM, const double* __restrict__ x,
                               double* __restrict__ y) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Pd) return;
  double v[NV];
#pragma unroll
  for (int k = 0; k < NV; ++k) v[k] = x[(size_t)i * NV + k];
#pragma unroll
  for (int r = 0; r < NV; ++r) {
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < NV; ++k) s += invM[(size_t)i * NV * NV + r * NV + k] * v[k];
    y[(size_t)i * NV + r] = s;
  }
}

void linsys_build_jacobi(su2b200_ctx* c, LinSys& ls) {
  ls.invM.alloc((size_t)ls.Pd * ls.nVar * ls.nVar);
  switch (ls.nVar) {
    case 1: SU2B_LAUNCH(c, k_jacobi_build<1>, nblk(ls.Pd, 128), 128, ls.Pd, ls.dia_ptr, ls.A.p, ls.invM.p); break;
    case 2: SU2B_LAUNCH(c, k_jacobi_build<2>, nblk(ls.Pd, 128), 128, ls.Pd, ls.dia_ptr, ls.A.p, ls.invM.p); break;
    case 4: SU2B_LAUNCH(c, k_jacobi_build<4>, nblk(ls.Pd, 128), 128, ls.Pd, ls.dia_ptr, ls.A.p, ls.invM.p); break;
    case 5: SU2B_LAUNCH(c, k_jacobi_build<5>, nblk(ls.Pd, 128), 128, ls.Pd, ls.dia_ptr, ls.A.p, ls.invM.p); break;
    default: throw Error("jacobi: unsupported nVar");
  }
}
void linsys_apply_jacobi(su2b200_ctx* c, LinSys& ls, const double* x, double* y) {
  switch (ls.nVar) {
    case 1: SU2B_LAUNCH(c, k_jacobi_apply<1>, nblk(ls.Pd, 128), 128, ls.Pd, ls.invM.p, x, y); break;
    case 2: SU2B_LAUNCH(c, k_jacobi_apply<2>, nblk(ls.Pd, 128), 128, ls.Pd, ls.invM.p, x, y); break;
    case 4: SU2B_LAUNCH(c, k_jacobi_apply<4>, nblk(ls.Pd, 128), 128, ls.Pd, ls.invM.p, x, y); break;
    case 5: SU2B_LAUNCH(c, k_jacobi_apply<5>, nblk(ls.Pd, 128), 128, ls.Pd, ls.invM.p, x, y); break;
    default: throw Error("jacobi: unsupported nVar");
  }
  if (c->nranks > 1 && !c->defer_halo) halo_exchange(c, y, ls.nVar);
}
}  // namespace su2b200

// ------------------------------------------------------------------------------------------------
// Allocation-free dot product (su2b200_vector_dot_noalloc).  No counterpart in the reference: added
// for callers that may not call cudaMalloc (graph capture, tight serving loops, harnesses that
// forbid it).  Same two kernels as vec_dot; the staging copies of the two host vectors and the block partials live in
// ONE file-scope device array instead of the pooled vectors / ls.scalars that vec_dot allocates on first use.
// Every GPU that loads the library holds the array (16 MiB + 8 KiB) until exit.  All calls share it: calls whose GPU
// work overlaps (two streams, two contexts on one device, two host threads) race on it.  Nothing in it has to start
// at zero: each call overwrites the n staged entries and the nb partials it then reads, so the previous call's
// contents are never observed.
// ------------------------------------------------------------------------------------------------
namespace su2b200 {
constexpr unsigned long NOALLOC_DOT_MAX = SU2B200_NOALLOC_DOT_MAX;  // doubles per vector
constexpr size_t NOALLOC_PART = 2 * (size_t)NOALLOC_DOT_MAX;        // offset of the partials: [DOT_BLOCKS] + result
static __device__ __align__(256) double g_noalloc_dot_scratch[NOALLOC_PART + 1024];
static_assert(DOT_BLOCKS + 1 <= 1024, "partials and the result must fit the tail of the scratch array");

// In host code `g_noalloc_dot_scratch` is only a stand-in for the symbol, not a device address: the address is taken
// once per device with cudaGetSymbolAddress, with that device current.
static double* noalloc_dot_scratch(int device) {
  constexpr int MAX_DEV = 64;
  static double* ptr[MAX_DEV] = {};
  static std::mutex mu;
  SU2B_REQUIRE(device >= 0 && device < MAX_DEV, "su2b200_vector_dot_noalloc: device ordinal out of range");
  std::lock_guard<std::mutex> lk(mu);
  if (!ptr[device]) {
    int cur = -1;
    SU2B_CUDA(cudaGetDevice(&cur));
    SU2B_REQUIRE(cur == device, "su2b200_vector_dot_noalloc: the context's device is not current");
    double* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_dot_scratch);
    if (e != cudaSuccess || !p)
      throw Error(std::string("su2b200_vector_dot_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
    ptr[device] = p;
  }
  return ptr[device];
}
}  // namespace su2b200

extern "C" int su2b200_vector_dot_noalloc(su2b200_ctx* c, const double* a, const double* b, unsigned long n,
                                          double* result) {
  using namespace su2b200;
  try {
    if (!c) throw Error("null context");
    SU2B_REQUIRE(a && b && result, "su2b200_vector_dot_noalloc: null pointer");
    if (n > NOALLOC_DOT_MAX)
      throw Error("su2b200_vector_dot_noalloc: n = " + std::to_string(n) + " exceeds the static scratch (SU2B200_NOALLOC_DOT_MAX = " +
                  std::to_string(NOALLOC_DOT_MAX) + " doubles per vector); use su2b200_vector_dot");
    SU2B_CUDA(cudaSetDevice(c->device));
    if (n == 0) { *result = 0.0; return 0; }
    double* s = noalloc_dot_scratch(c->device);
    double* da = s;
    double* db = s + NOALLOC_DOT_MAX;
    double* part = s + NOALLOC_PART;
    SU2B_CUDA(cudaMemcpyAsync(da, a, n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    SU2B_CUDA(cudaMemcpyAsync(db, b, n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    const int nb = (int)std::min<size_t>(((size_t)n + 255) / 256, DOT_BLOCKS);
    SU2B_LAUNCH(c, k_dot_stage1, nb, 256, (size_t)n, (const double*)da, (const double*)db, part);
    SU2B_LAUNCH(c, k_dot_stage2, 1, 256, nb, (const double*)part, part + DOT_BLOCKS);
    SU2B_CUDA(cudaMemcpyAsync(result, part + DOT_BLOCKS, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    SU2B_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
  } catch (const std::exception& e) {
    set_error(e.what());
    return 1;
  } catch (...) {
    set_error("unknown C++ exception");
    return 2;
  }
}
