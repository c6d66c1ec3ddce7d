// a17-a19: ILU(0) and LU-SGS on the device.
//   CSysMatrix<T>::BuildILUPreconditioner   (Common/src/linear_algebra/CSysMatrix.cpp:673-767)
//   CSysMatrix<T>::ComputeILUPreconditioner (:770-823)
//   CSysMatrix<T>::ComputeLU_SGSPreconditioner (:826-888), Lower/Upper/DiagonalProduct (CSysMatrix.inl:184-211)
//
// The reference factorises the owned sub-matrix [0, nPointDomain) per MPI rank and, with OpenMP, per
// nnz-balanced row partition [begin, end) (block-Jacobi between partitions, CSysMatrix.cpp:199-227, :701-711).
// Two schedules are provided for the same arithmetic:
//   exact (ilu_exact=1): one partition == the serial reference; rows are grouped by dependency level
//                        (level-scheduled triangular sweeps), one warp per row, one launch per level;
//   partitioned        : K partitions as the reference does with K threads; one warp walks one partition
//                        sequentially, all partitions in one launch.
// In both, a warp owns a row: lane q = r*nVar + c holds one scalar of a block, products use shuffles.
#include "internal.hpp"
#include "kernels.hpp"
#include <algorithm>
#include <cstdlib>

namespace su2b200 {

static inline int nblk(long n, int bs) { return (int)((n + bs - 1) / bs); }
static constexpr unsigned FULL = 0xffffffffu;

template <int NV>
__device__ __forceinline__ void matrix_inverse_d(double* A, double* M);  // defined below (copy of linalg.cu's)

__device__ __forceinline__ void reg_pivot(double& pivot) {
  const double eps = (double)1e-12f;
  if (fabs(pivot) < eps) pivot = copysign(eps, pivot);
}
template <int NV>
__device__ __forceinline__ void matrix_inverse_d(double* A, double* M) {
#pragma unroll
  for (int i = 0; i < NV; ++i)
#pragma unroll
    for (int j = 0; j < NV; ++j) M[i * NV + j] = (i == j) ? 1.0 : 0.0;
#pragma unroll
  for (int i = 1; i < NV; ++i) {
#pragma unroll
    for (int j = 0; j < i; ++j) {
      reg_pivot(A[j * NV + j]);
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
    reg_pivot(A[i * NV + i]);
#pragma unroll
    for (int k = 0; k < NV; ++k) M[i * NV + k] /= A[i * NV + i];
  }
}
template <int NV>
__device__ __forceinline__ void gauss_elim_d(double* A, double* v) {
#pragma unroll
  for (int i = 1; i < NV; ++i) {
#pragma unroll
    for (int j = 0; j < i; ++j) {
      reg_pivot(A[j * NV + j]);
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
    reg_pivot(A[i * NV + i]);
    v[i] /= A[i * NV + i];
  }
}

// C[q] = sum_m A[r][m] * B[m][c] with A,B distributed one scalar per lane (lanes 0..NV*NV-1)
template <int NV>
__device__ __forceinline__ double warp_matmul(double a, double b, int r, int cc) {
  double s = 0.0;
#pragma unroll
  for (int m = 0; m < NV; ++m) {
    const double am = __shfl_sync(FULL, a, r * NV + m);
    const double bm = __shfl_sync(FULL, b, m * NV + cc);
    s += am * bm;
  }
  return s;
}

// ---- one ILU(0) row, warp-cooperative (IKJ variant, CSysMatrix.cpp:714-760) -----------------------
template <int NV>
__device__ __forceinline__ void ilu_factor_row(int i, int begin, int end, const int* __restrict__ row_ptr,
                                               const int* __restrict__ col_ind, const int* __restrict__ dia_ptr,
                                               double* __restrict__ ILU, double* __restrict__ invM, int lane) {
  constexpr int LPB = NV * NV;
  const int q = lane < LPB ? lane : LPB - 1;  // idle lanes mirror the last scalar (never store)
  const bool act = lane < LPB;
  const int r = q / NV, cc = q % NV;
  const int rb = row_ptr[i], re = row_ptr[i + 1], di = dia_ptr[i];
  for (int idx = rb; idx < di; ++idx) {
    const int j = col_ind[idx];
    if (j < begin) continue;
    const double aij = ILU[(size_t)idx * LPB + q];
    const double inv = invM[(size_t)j * LPB + q];
    const double w = warp_matmul<NV>(aij, inv, r, cc);  // weight = A_ij * inv(A_jj)
    for (int idx2 = dia_ptr[j] + 1; idx2 < row_ptr[j + 1]; ++idx2) {
      const int k = col_ind[idx2];
      if (k >= end) break;
      // find (i,k) in row i
      int pos = -1;
      for (int s = rb; s < re; ++s)
        if (col_ind[s] == k) { pos = s; break; }
      if (pos < 0) continue;
      const double ajk = ILU[(size_t)idx2 * LPB + q];
      const double aux = warp_matmul<NV>(w, ajk, r, cc);
      if (act) ILU[(size_t)pos * LPB + q] -= aux;
    }
    if (act) ILU[(size_t)idx * LPB + q] = w;
    __syncwarp();
  }
  // invM_i = inverse of the (updated) diagonal block; every lane inverts redundantly, lane q stores entry q
  double B[NV * NV], M[NV * NV];
#pragma unroll
  for (int k = 0; k < NV * NV; ++k) B[k] = ILU[(size_t)di * LPB + k];
  matrix_inverse_d<NV>(B, M);
  if (act) invM[(size_t)i * LPB + q] = M[q];
}

// ---- forward / backward substitution rows (CSysMatrix.cpp:791-817) ------------------------------------
template <int NV>
__device__ __forceinline__ void ilu_forward_row(int i, int begin, const int* __restrict__ row_ptr,
                                                const int* __restrict__ col_ind, const int* __restrict__ dia_ptr,
                                                const double* __restrict__ ILU, double* __restrict__ y, int lane) {
  constexpr int LPB = NV * NV;
  const int q = lane < LPB ? lane : LPB - 1;
  const int cc = q % NV;
  double acc = 0.0;
  for (int idx = row_ptr[i]; idx < dia_ptr[i]; ++idx) {
    const int j = col_ind[idx];
    if (j < begin) continue;
    acc += ILU[(size_t)idx * LPB + q] * y[(size_t)j * NV + cc];
  }
  double s = acc;
#pragma unroll
  for (int o = 1; o < NV; ++o) s += __shfl_down_sync(FULL, acc, o);
  if (lane < LPB && cc == 0) y[(size_t)i * NV + q / NV] -= s;
}

template <int NV>
__device__ __forceinline__ void ilu_backward_row(int i, int end, const int* __restrict__ row_ptr,
                                                 const int* __restrict__ col_ind, const int* __restrict__ dia_ptr,
                                                 const double* __restrict__ ILU, const double* __restrict__ invM,
                                                 double* __restrict__ y, int lane) {
  constexpr int LPB = NV * NV;
  const int q = lane < LPB ? lane : LPB - 1;
  const int r = q / NV, cc = q % NV;
  double acc = 0.0;
  for (int idx = dia_ptr[i] + 1; idx < row_ptr[i + 1]; ++idx) {
    const int j = col_ind[idx];
    if (j >= end) break;
    acc += ILU[(size_t)idx * LPB + q] * y[(size_t)j * NV + cc];
  }
  double s = acc;
#pragma unroll
  for (int o = 1; o < NV; ++o) s += __shfl_down_sync(FULL, acc, o);
  // aux[r] = y_i[r] - s (valid on lanes with cc == 0, i.e. lane r*NV)
  const double aux_r = y[(size_t)i * NV + r] - s;
  // x_i = invM_i * aux : lane q needs aux[cc], held by lane cc*NV
  const double aux_c = __shfl_sync(FULL, aux_r, cc * NV);
  const double prod = invM[(size_t)i * LPB + q] * aux_c;
  double t = prod;
#pragma unroll
  for (int o = 1; o < NV; ++o) t += __shfl_down_sync(FULL, prod, o);
  __syncwarp();
  if (lane < LPB && cc == 0) y[(size_t)i * NV + r] = t;
}

// ---- level-scheduled kernels: one warp per row of the level ----------------------------------------------
template <int NV>
__global__ void k_ilu_factor_level(int n, const int* __restrict__ rows, int begin, int end,
                                   const int* __restrict__ row_ptr, const int* __restrict__ col_ind,
                                   const int* __restrict__ dia_ptr, double* __restrict__ ILU, double* __restrict__ invM) {
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (w >= n) return;
  ilu_factor_row<NV>(rows[w], begin, end, row_ptr, col_ind, dia_ptr, ILU, invM, threadIdx.x & 31);
}
template <int NV>
__global__ void k_ilu_forward_level(int n, const int* __restrict__ rows, int begin, const int* __restrict__ row_ptr,
                                    const int* __restrict__ col_ind, const int* __restrict__ dia_ptr,
                                    const double* __restrict__ ILU, double* __restrict__ y) {
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (w >= n) return;
  ilu_forward_row<NV>(rows[w], begin, row_ptr, col_ind, dia_ptr, ILU, y, threadIdx.x & 31);
}
template <int NV>
__global__ void k_ilu_backward_level(int n, const int* __restrict__ rows, int end, const int* __restrict__ row_ptr,
                                     const int* __restrict__ col_ind, const int* __restrict__ dia_ptr,
                                     const double* __restrict__ ILU, const double* __restrict__ invM,
                                     double* __restrict__ y) {
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (w >= n) return;
  ilu_backward_row<NV>(rows[w], end, row_ptr, col_ind, dia_ptr, ILU, invM, y, threadIdx.x & 31);
}

// ---- partitioned kernels: one warp per partition, rows in sequence ---------------------------------------
template <int NV>
__global__ void k_ilu_factor_part(int nparts, const int* __restrict__ part, const int* __restrict__ row_ptr,
                                  const int* __restrict__ col_ind, const int* __restrict__ dia_ptr,
                                  double* __restrict__ ILU, double* __restrict__ invM) {
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (w >= nparts) return;
  const int begin = part[w], end = part[w + 1];
  for (int i = begin; i < end; ++i) {
    ilu_factor_row<NV>(i, begin, end, row_ptr, col_ind, dia_ptr, ILU, invM, threadIdx.x & 31);
    __syncwarp();
  }
}
template <int NV>
__global__ void k_ilu_apply_part(int nparts, const int* __restrict__ part, const int* __restrict__ row_ptr,
                                 const int* __restrict__ col_ind, const int* __restrict__ dia_ptr,
                                 const double* __restrict__ ILU, const double* __restrict__ invM,
                                 double* __restrict__ y) {
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (w >= nparts) return;
  const int begin = part[w], end = part[w + 1];
  const int lane = threadIdx.x & 31;
  for (int i = begin + 1; i < end; ++i) {
    ilu_forward_row<NV>(i, begin, row_ptr, col_ind, dia_ptr, ILU, y, lane);
    __syncwarp();
  }
  for (int i = end - 1; i >= begin; --i) {
    ilu_backward_row<NV>(i, end, row_ptr, col_ind, dia_ptr, ILU, invM, y, lane);
    __syncwarp();
  }
}

// ---- LU-SGS (CSysMatrix.cpp:826-888): thread per row within a level ----------------------------------------
template <int NV>
__global__ void k_lusgs_forward_level(int n, const int* __restrict__ rows, int begin, const int* __restrict__ row_ptr,
                                      const int* __restrict__ col_ind, const int* __restrict__ dia_ptr,
                                      const double* __restrict__ A, const double* __restrict__ b,
                                      double* __restrict__ x) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int i = rows[t];
  double low[NV];
#pragma unroll
  for (int r = 0; r < NV; ++r) low[r] = 0.0;
  for (int idx = row_ptr[i]; idx < dia_ptr[i]; ++idx) {
    const int j = col_ind[idx];
    if (j < begin) continue;
    const double* B = A + (size_t)idx * NV * NV;
#pragma unroll
    for (int r = 0; r < NV; ++r)
#pragma unroll
      for (int k = 0; k < NV; ++k) low[r] += B[r * NV + k] * x[(size_t)j * NV + k];
  }
  double v[NV], D[NV * NV];
#pragma unroll
  for (int r = 0; r < NV; ++r) v[r] = b[(size_t)i * NV + r] - low[r];
  const double* Dg = A + (size_t)dia_ptr[i] * NV * NV;
#pragma unroll
  for (int k = 0; k < NV * NV; ++k) D[k] = Dg[k];
  gauss_elim_d<NV>(D, v);
#pragma unroll
  for (int r = 0; r < NV; ++r) x[(size_t)i * NV + r] = v[r];
}

template <int NV>
__global__ void k_lusgs_backward_level(int n, const int* __restrict__ rows, int col_ub, int Pd,
                                       const int* __restrict__ row_ptr, const int* __restrict__ col_ind,
                                       const int* __restrict__ dia_ptr, const double* __restrict__ A,
                                       double* __restrict__ x) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int i = rows[t];
  double up[NV], dia[NV], xi[NV], D[NV * NV];
  const double* Dg = A + (size_t)dia_ptr[i] * NV * NV;
#pragma unroll
  for (int k = 0; k < NV * NV; ++k) D[k] = Dg[k];
#pragma unroll
  for (int r = 0; r < NV; ++r) { up[r] = 0.0; xi[r] = x[(size_t)i * NV + r]; }
#pragma unroll
  for (int r = 0; r < NV; ++r) {
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < NV; ++k) s += D[r * NV + k] * xi[k];
    dia[r] = s;
  }
  for (int idx = dia_ptr[i] + 1; idx < row_ptr[i + 1]; ++idx) {
    const int j = col_ind[idx];
    if (j < col_ub || j >= Pd) {
      const double* B = A + (size_t)idx * NV * NV;
#pragma unroll
      for (int r = 0; r < NV; ++r)
#pragma unroll
        for (int k = 0; k < NV; ++k) up[r] += B[r * NV + k] * x[(size_t)j * NV + k];
    }
  }
  double v[NV];
#pragma unroll
  for (int r = 0; r < NV; ++r) v[r] = dia[r] - up[r];
  gauss_elim_d<NV>(D, v);
#pragma unroll
  for (int r = 0; r < NV; ++r) x[(size_t)i * NV + r] = v[r];
}

// ------------------------------------------------------------------------------------------------
// Single-launch dependency-driven sweeps (the default schedule of the exact ILU0).
// One launch performs a whole triangular sweep (or the whole factorisation): warps draw tickets from a global
// counter and process the rows of the level-ordered list in ticket order; a row spins (acquire loads) on the
// `ready` word of each row it depends on and publishes its own with a release store. Because the list is in level
// order, every dependency of ticket t has a smaller ticket, i.e. was drawn by a warp that is already running, so the
// scheme cannot deadlock whatever the residency of the grid (no cooperative launch, no cross-launch waiting).
// A spin watchdog sets ctrl[1] and releases every waiter if a flag never arrives; the host turns that into an error.
// Arithmetic and summation order per row are those of the level kernels above (== the reference's serial sweeps).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release(int* p, int v) {
  asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ bool wait_ready(const int* ready, int j, int epoch, int* ctrl) {
  unsigned spins = 0;
  while (ld_acquire(ready + j) != epoch) {
    if ((++spins & 1023u) == 0) {
      if (spins > (1u << 24)) atomicExch(ctrl + 1, 1);
      if (ld_acquire(ctrl + 1) != 0) return false;
    }
  }
  return true;
}
__device__ __forceinline__ int draw_ticket(int* ctrl, int lane) {
  int t = 0;
  if (lane == 0) t = atomicAdd(ctrl, 1);
  return __shfl_sync(FULL, t, 0);
}

// ---- block systems (nVar 4/5): one warp per row -----------------------------------------------------
template <int NV>
__global__ void __launch_bounds__(128) k_sweep_forward(int n, const int* __restrict__ rows, int epoch, int* ready, int* ctrl,
                                                       const int* __restrict__ row_ptr, const int* __restrict__ col_ind,
                                                       const int* __restrict__ dia_ptr, const double* __restrict__ ILU,
                                                       double* y) {
  constexpr int LPB = NV * NV;
  constexpr int CH = 4;
  const int lane = threadIdx.x & 31;
  const int q = lane < LPB ? lane : LPB - 1;
  const int cc = q % NV;
  for (;;) {
    const int t = draw_ticket(ctrl, lane);
    if (t >= n) return;
    const int i = rows[t];
    const int rb = row_ptr[i], di = dia_ptr[i];
    double acc = 0.0;
    for (int base = rb; base < di; base += CH) {
      double a[CH];
      int jj[CH];
#pragma unroll
      for (int k = 0; k < CH; ++k) {  // matrix entries do not depend on the sweep: fetch them before waiting
        const int idx = base + k < di ? base + k : di - 1;
        jj[k] = col_ind[idx];
        a[k] = ILU[(size_t)idx * LPB + q];
      }
#pragma unroll
      for (int k = 0; k < CH; ++k) {
        if (base + k < di) {
          if (!wait_ready(ready, jj[k], epoch, ctrl)) return;
          acc += a[k] * __ldcg(y + (size_t)jj[k] * NV + cc);
        }
      }
    }
    double s = acc;
#pragma unroll
    for (int o = 1; o < NV; ++o) s += __shfl_down_sync(FULL, acc, o);
    if (lane < LPB && cc == 0 && di > rb) {
      y[(size_t)i * NV + q / NV] -= s;
      __threadfence();
    }
    __syncwarp();
    if (lane == 0) st_release(ready + i, epoch);
  }
}

template <int NV>
__global__ void __launch_bounds__(128) k_sweep_backward(int n, const int* __restrict__ rows, int end, int epoch, int* ready,
                                                        int* ctrl, const int* __restrict__ row_ptr,
                                                        const int* __restrict__ col_ind, const int* __restrict__ dia_ptr,
                                                        const double* __restrict__ ILU, const double* __restrict__ invM,
                                                        double* y) {
  constexpr int LPB = NV * NV;
  constexpr int CH = 4;
  const int lane = threadIdx.x & 31;
  const int q = lane < LPB ? lane : LPB - 1;
  const int r = q / NV, cc = q % NV;
  for (;;) {
    const int t = draw_ticket(ctrl, lane);
    if (t >= n) return;
    const int i = rows[t];
    const int ub = dia_ptr[i] + 1, re = row_ptr[i + 1];
    const double minv = invM[(size_t)i * LPB + q];
    const double yi = y[(size_t)i * NV + r];
    double acc = 0.0;
    for (int base = ub; base < re; base += CH) {
      double a[CH];
      int jj[CH];
#pragma unroll
      for (int k = 0; k < CH; ++k) {
        const int idx = base + k < re ? base + k : re - 1;
        jj[k] = col_ind[idx];
        a[k] = ILU[(size_t)idx * LPB + q];
      }
#pragma unroll
      for (int k = 0; k < CH; ++k) {
        if (base + k < re && jj[k] < end) {
          if (!wait_ready(ready, jj[k], epoch, ctrl)) return;
          acc += a[k] * __ldcg(y + (size_t)jj[k] * NV + cc);
        }
      }
    }
    double s = acc;
#pragma unroll
    for (int o = 1; o < NV; ++o) s += __shfl_down_sync(FULL, acc, o);
    const double aux_r = yi - s;  // valid on lanes r*NV
    const double aux_c = __shfl_sync(FULL, aux_r, cc * NV);
    const double prod = minv * aux_c;
    double tsum = prod;
#pragma unroll
    for (int o = 1; o < NV; ++o) tsum += __shfl_down_sync(FULL, prod, o);
    if (lane < LPB && cc == 0) {
      y[(size_t)i * NV + r] = tsum;
      __threadfence();
    }
    __syncwarp();
    if (lane == 0) st_release(ready + i, epoch);
  }
}

template <int NV>
__global__ void __launch_bounds__(128) k_sweep_factor(int n, const int* __restrict__ rows, int end, int epoch, int* ready,
                                                      int* ctrl, const int* __restrict__ row_ptr,
                                                      const int* __restrict__ col_ind, const int* __restrict__ dia_ptr,
                                                      double* ILU, double* invM) {
  constexpr int LPB = NV * NV;
  const int lane = threadIdx.x & 31;
  const int q = lane < LPB ? lane : LPB - 1;
  const bool act = lane < LPB;
  const int r = q / NV, cc = q % NV;
  for (;;) {
    const int t = draw_ticket(ctrl, lane);
    if (t >= n) return;
    const int i = rows[t];
    const int rb = row_ptr[i], re = row_ptr[i + 1], di = dia_ptr[i];
    for (int idx = rb; idx < di; ++idx) {
      const int j = col_ind[idx];
      const double aij = __ldcg(ILU + (size_t)idx * LPB + q);  // own element (this lane wrote it, if at all)
      if (!wait_ready(ready, j, epoch, ctrl)) return;
      const double inv = __ldcg(invM + (size_t)j * LPB + q);
      const double w = warp_matmul<NV>(aij, inv, r, cc);
      for (int idx2 = dia_ptr[j] + 1; idx2 < row_ptr[j + 1]; ++idx2) {
        const int k = col_ind[idx2];
        if (k >= end) break;
        int pos = -1;
        for (int s = rb; s < re; ++s)
          if (col_ind[s] == k) { pos = s; break; }
        if (pos < 0) continue;
        const double ajk = __ldcg(ILU + (size_t)idx2 * LPB + q);
        const double aux = warp_matmul<NV>(w, ajk, r, cc);
        if (act) ILU[(size_t)pos * LPB + q] -= aux;
      }
      if (act) ILU[(size_t)idx * LPB + q] = w;
    }
    // inverse of the updated diagonal block: gather it through the lanes that own its scalars
    const double myd = __ldcg(ILU + (size_t)di * LPB + q);
    double B[NV * NV], M[NV * NV];
#pragma unroll
    for (int k = 0; k < NV * NV; ++k) B[k] = __shfl_sync(FULL, myd, k);
    matrix_inverse_d<NV>(B, M);
    if (act) {
      invM[(size_t)i * LPB + q] = M[q];
      __threadfence();
    }
    __syncwarp();
    if (lane == 0) st_release(ready + i, epoch);
  }
}

// ---- scalar systems (nVar 1): one thread per row; `rows` is padded per level to a multiple of 32 with -1, so the
//      rows of a warp never depend on each other ---------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_sweep_forward_s(int nWarpTickets, const int* __restrict__ rows, int epoch,
                                                         int* ready, int* ctrl, const int* __restrict__ row_ptr,
                                                         const int* __restrict__ col_ind, const int* __restrict__ dia_ptr,
                                                         const double* __restrict__ ILU, double* y) {
  const int lane = threadIdx.x & 31;
  for (;;) {
    const int t = draw_ticket(ctrl, lane);
    if (t >= nWarpTickets) return;
    const int i = rows[(size_t)t * 32 + lane];
    if (i < 0) continue;
    double acc = 0.0;
    bool ok = true;  // on a watchdog abort keep drawing tickets (the warp-wide shuffle needs every lane) but do no work
    const int rb = row_ptr[i], di = dia_ptr[i];
    for (int idx = rb; idx < di && ok; ++idx) {
      const int j = col_ind[idx];
      const double a = ILU[idx];
      ok = wait_ready(ready, j, epoch, ctrl);
      if (ok) acc += a * __ldcg(y + j);
    }
    if (di > rb && ok) {
      y[i] -= acc;
      __threadfence();
    }
    st_release(ready + i, epoch);
  }
}
__global__ void __launch_bounds__(128) k_sweep_backward_s(int nWarpTickets, const int* __restrict__ rows, int end, int epoch,
                                                          int* ready, int* ctrl, const int* __restrict__ row_ptr,
                                                          const int* __restrict__ col_ind, const int* __restrict__ dia_ptr,
                                                          const double* __restrict__ ILU, const double* __restrict__ invM,
                                                          double* y) {
  const int lane = threadIdx.x & 31;
  for (;;) {
    const int t = draw_ticket(ctrl, lane);
    if (t >= nWarpTickets) return;
    const int i = rows[(size_t)t * 32 + lane];
    if (i < 0) continue;
    double acc = 0.0;
    const double minv = invM[i], yi = y[i];
    bool ok = true;
    for (int idx = dia_ptr[i] + 1; idx < row_ptr[i + 1] && ok; ++idx) {
      const int j = col_ind[idx];
      if (j >= end) break;
      const double a = ILU[idx];
      ok = wait_ready(ready, j, epoch, ctrl);
      if (ok) acc += a * __ldcg(y + j);
    }
    if (ok) y[i] = minv * (yi - acc);
    __threadfence();
    st_release(ready + i, epoch);
  }
}
__global__ void __launch_bounds__(128) k_sweep_factor_s(int nWarpTickets, const int* __restrict__ rows, int end, int epoch,
                                                        int* ready, int* ctrl, const int* __restrict__ row_ptr,
                                                        const int* __restrict__ col_ind, const int* __restrict__ dia_ptr,
                                                        double* ILU, double* invM) {
  const int lane = threadIdx.x & 31;
  for (;;) {
    const int t = draw_ticket(ctrl, lane);
    if (t >= nWarpTickets) return;
    const int i = rows[(size_t)t * 32 + lane];
    if (i < 0) continue;
    const int rb = row_ptr[i], re = row_ptr[i + 1], di = dia_ptr[i];
    bool ok = true;
    for (int idx = rb; idx < di && ok; ++idx) {
      const int j = col_ind[idx];
      ok = wait_ready(ready, j, epoch, ctrl);
      if (!ok) break;
      const double w = ILU[idx] * __ldcg(invM + j);
      for (int idx2 = dia_ptr[j] + 1; idx2 < row_ptr[j + 1]; ++idx2) {
        const int k = col_ind[idx2];
        if (k >= end) break;
        int pos = -1;
        for (int s = rb; s < re; ++s)
          if (col_ind[s] == k) { pos = s; break; }
        if (pos < 0) continue;
        ILU[pos] -= w * __ldcg(ILU + idx2);
      }
      ILU[idx] = w;
    }
    double d = ILU[di], one = 1.0;
    reg_pivot(d);
    invM[i] = one / d;
    __threadfence();
    st_release(ready + i, epoch);
  }
}

// ------------------------------------------------------------------------------------------------
// Host: dependency levels of the owned sub-matrix (lower: deps j<i ; upper: deps j>i, j<Pd)
// ------------------------------------------------------------------------------------------------
static void build_levels(LinSys& ls) {
  if (ls.levels_ready) return;
  // A/B switches: SU2B200_ILU_APPLY = 0 (launch per level) | 1 (flag sweeps) | 2 (data-polling sweeps, default)
  //               SU2B200_ILU_BUILD = 0 (launch per level) | 1 (flag sweep) | 2 (polling factorisation, default)
  if (const char* e = getenv("SU2B200_ILU_APPLY")) ls.sweep_mode = atoi(e);
  if (const char* e = getenv("SU2B200_ILU_BUILD")) ls.build_mode = atoi(e);
  const int n = ls.Pd;
  const auto& rp = ls.h_row_ptr;
  const auto& ci = ls.h_col_ind;
  const auto& dp = ls.h_dia_ptr;
  std::vector<int> lev(n, 0);
  int nl = 0;
  for (int i = 0; i < n; ++i) {
    int l = 0;
    for (int k = rp[i]; k < dp[i]; ++k) l = std::max(l, lev[ci[k]] + 1);
    lev[i] = l;
    nl = std::max(nl, l + 1);
  }
  auto bucket = [&](int nlev, std::vector<int>& ptr, dvec<int>& rows_d) {
    ptr.assign(nlev + 1, 0);
    for (int i = 0; i < n; ++i) ptr[lev[i] + 1]++;
    for (int l = 0; l < nlev; ++l) ptr[l + 1] += ptr[l];
    std::vector<int> rows(n), pos(ptr.begin(), ptr.end() - 1);
    for (int i = 0; i < n; ++i) rows[pos[lev[i]]++] = i;
    rows_d.upload(rows);
  };
  ls.nlev_lower = nl;
  bucket(nl, ls.lvl_ptr_lower, ls.lvl_rows_lower);
  nl = 0;
  for (int i = n - 1; i >= 0; --i) {
    int l = 0;
    for (int k = dp[i] + 1; k < rp[i + 1]; ++k) {
      if (ci[k] >= n) break;
      l = std::max(l, lev[ci[k]] + 1);
    }
    lev[i] = l;
    nl = std::max(nl, l + 1);
  }
  ls.nlev_upper = nl;
  bucket(nl, ls.lvl_ptr_upper, ls.lvl_rows_upper);
  // ticket lists of the single-launch sweeps: block systems use the level-ordered list as is (a warp per row);
  // scalar systems pad every level to a multiple of 32 so that the rows of one warp are mutually independent
  auto padded = [&](const std::vector<int>& ptr, const dvec<int>& rows_d, dvec<int>& out, int& count) {
    if (ls.nVar != 1) { count = n; return; }
    const std::vector<int> rows = rows_d.to_host();
    std::vector<int> pad;
    pad.reserve(rows.size() + 32 * (ptr.size() - 1));
    for (size_t l = 0; l + 1 < ptr.size(); ++l) {
      for (int k = ptr[l]; k < ptr[l + 1]; ++k) pad.push_back(rows[k]);
      while (pad.size() % 32) pad.push_back(-1);
    }
    count = (int)(pad.size() / 32);
    out.upload(pad);
  };
  padded(ls.lvl_ptr_lower, ls.lvl_rows_lower, ls.sweep_rows_lower, ls.sweep_n_lower);
  padded(ls.lvl_ptr_upper, ls.lvl_rows_upper, ls.sweep_rows_upper, ls.sweep_n_upper);
  ls.sweep_ready.alloc(n > 0 ? n : 1);
  ls.sweep_ready.zero();
  ls.sweep_ctrl.alloc(2);
  ls.sweep_ctrl.zero();
  ls.sweep_epoch = 0;
  if (!ls.h_sweep_ctrl) SU2B_CUDA(cudaMallocHost(&ls.h_sweep_ctrl, 2 * sizeof(int)));
  SU2B_CUDA(cudaDeviceSynchronize());
  ls.levels_ready = true;
}

// grid of the persistent sweep kernels: enough warps to fill the machine, never more than there are tickets
static inline int sweep_grid(int tickets) {
  const int warpsPerBlock = 4;
  return std::max(1, std::min((tickets + warpsPerBlock - 1) / warpsPerBlock, 148 * 16));
}
// new epoch + ticket reset (stream ordered); abort word is sticky until checked
static inline int sweep_begin(su2b200_ctx* c, LinSys& ls) {
  SU2B_CUDA(cudaMemsetAsync(ls.sweep_ctrl.p, 0, sizeof(int), c->stream));
  return ++ls.sweep_epoch;
}
void linsys_check_sweeps(su2b200_ctx* c, LinSys& ls) {
  if (!ls.sweep_ctrl.p || !ls.h_sweep_ctrl) return;
  SU2B_CUDA(cudaMemcpyAsync(ls.h_sweep_ctrl, ls.sweep_ctrl.p, 2 * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  if (ls.h_sweep_ctrl[1] != 0) {
    SU2B_CUDA(cudaMemsetAsync(ls.sweep_ctrl.p, 0, 2 * sizeof(int), c->stream));
    throw Error("triangular sweep watchdog fired: a row dependency never completed (corrupt level schedule?)");
  }
}

// nnz-balanced partitions, CSysMatrix::Initialize (CSysMatrix.cpp:199-227)
static void build_partitions(LinSys& ls, int nparts) {
  if ((int)ls.h_parts.size() == nparts + 1) {
    if (ls.parts.n != ls.h_parts.size()) ls.parts.upload(ls.h_parts);  // explicit boundaries (set_partition_rows)
    return;
  }
  const auto& rp = ls.h_row_ptr;
  const int n = ls.Pd;
  ls.h_parts.assign(nparts + 1, n);
  ls.h_parts[0] = 0;
  const long nnz_per = (long)((rp[n] + nparts - 1) / nparts);  // roundUpDiv
  int part = 0;
  for (int i = 0; i < n && part < nparts; ++i) {
    if (rp[i + 1] >= nnz_per * (part + 1)) ls.h_parts[++part] = i + 1;
  }
  for (int p = part + 1; p <= nparts; ++p) ls.h_parts[p] = n;
  ls.parts.upload(ls.h_parts);
}

template <int NV>
static void ilu_build_impl(su2b200_ctx* c, LinSys& ls) {
  const size_t nb = (size_t)ls.nnz * NV * NV;
  ls.ilu.alloc(nb);
  ls.invM.alloc((size_t)ls.Pd * NV * NV);
  if (ls.nparts <= 1) build_levels(ls);
  ls.ilu_slim = false;
  ls.ilu_gen = ls.A_gen;
  if (ls.nparts <= 1 && ls.build_mode == 2 && factor_symbolic(ls)) {
    ilu_factor_poll(c, ls);  // reads A, writes ILU: no copy
    return;
  }
  SU2B_CUDA(cudaMemcpyAsync(ls.ilu.p, ls.A.p, nb * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
  if (ls.nparts <= 1 && ls.build_mode == 1) {
    const int epoch = sweep_begin(c, ls);
    if (NV == 1) {
      SU2B_LAUNCH(c, k_sweep_factor_s, sweep_grid(ls.sweep_n_lower), 128, ls.sweep_n_lower, ls.sweep_rows_lower.p, ls.Pd,
                  epoch, ls.sweep_ready.p, ls.sweep_ctrl.p, ls.row_ptr, ls.col_ind, ls.dia_ptr, ls.ilu.p, ls.invM.p);
    } else {
      SU2B_LAUNCH(c, k_sweep_factor<(NV == 1 ? 4 : NV)>, sweep_grid(ls.Pd), 128, ls.Pd, ls.lvl_rows_lower.p, ls.Pd, epoch,
                  ls.sweep_ready.p, ls.sweep_ctrl.p, ls.row_ptr, ls.col_ind, ls.dia_ptr, ls.ilu.p, ls.invM.p);
    }
  } else if (ls.nparts <= 1) {
    build_levels(ls);
    for (int l = 0; l < ls.nlev_lower; ++l) {
      const int n = ls.lvl_ptr_lower[l + 1] - ls.lvl_ptr_lower[l];
      SU2B_LAUNCH(c, k_ilu_factor_level<NV>, nblk((long)n * 32, 128), 128, n, ls.lvl_rows_lower.p + ls.lvl_ptr_lower[l], 0,
                  ls.Pd, ls.row_ptr, ls.col_ind, ls.dia_ptr, ls.ilu.p, ls.invM.p);
    }
  } else {
    build_partitions(ls, ls.nparts);
    SU2B_LAUNCH(c, k_ilu_factor_part<NV>, nblk((long)ls.nparts * 32, 64), 64, ls.nparts, ls.parts.p, ls.row_ptr,
                ls.col_ind, ls.dia_ptr, ls.ilu.p, ls.invM.p);
  }
}

template <int NV>
static void ilu_apply_impl(su2b200_ctx* c, LinSys& ls, const double* x, double* y) {
  const size_t n = (size_t)ls.Pd * NV;
  if (ls.nparts <= 1) build_levels(ls);
  if (ls.nparts <= 1 && ls.sweep_mode == 2) {
    ilu_apply_poll(c, ls, x, y);
    return;
  }
  SU2B_REQUIRE(!ls.ilu_slim, "the ILU factors were built for the polling sweeps (upper blocks left in the Jacobian): "
                             "rebuild the preconditioner after changing the schedule");
  if (x != y) SU2B_CUDA(cudaMemcpyAsync(y, x, n * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
  if (ls.nparts <= 1 && ls.sweep_mode == 1) {
    int epoch = sweep_begin(c, ls);
    if (NV == 1) {
      SU2B_LAUNCH(c, k_sweep_forward_s, sweep_grid(ls.sweep_n_lower), 128, ls.sweep_n_lower, ls.sweep_rows_lower.p, epoch,
                  ls.sweep_ready.p, ls.sweep_ctrl.p, ls.row_ptr, ls.col_ind, ls.dia_ptr, ls.ilu.p, y);
      epoch = sweep_begin(c, ls);
      SU2B_LAUNCH(c, k_sweep_backward_s, sweep_grid(ls.sweep_n_upper), 128, ls.sweep_n_upper, ls.sweep_rows_upper.p, ls.Pd,
                  epoch, ls.sweep_ready.p, ls.sweep_ctrl.p, ls.row_ptr, ls.col_ind, ls.dia_ptr, ls.ilu.p, ls.invM.p, y);
    } else {
      constexpr int NB = NV == 1 ? 4 : NV;
      SU2B_LAUNCH(c, k_sweep_forward<NB>, sweep_grid(ls.Pd), 128, ls.Pd, ls.lvl_rows_lower.p, epoch, ls.sweep_ready.p,
                  ls.sweep_ctrl.p, ls.row_ptr, ls.col_ind, ls.dia_ptr, ls.ilu.p, y);
      epoch = sweep_begin(c, ls);
      SU2B_LAUNCH(c, k_sweep_backward<NB>, sweep_grid(ls.Pd), 128, ls.Pd, ls.lvl_rows_upper.p, ls.Pd, epoch,
                  ls.sweep_ready.p, ls.sweep_ctrl.p, ls.row_ptr, ls.col_ind, ls.dia_ptr, ls.ilu.p, ls.invM.p, y);
    }
  } else if (ls.nparts <= 1) {
    for (int l = 1; l < ls.nlev_lower; ++l) {
      const int m = ls.lvl_ptr_lower[l + 1] - ls.lvl_ptr_lower[l];
      SU2B_LAUNCH(c, k_ilu_forward_level<NV>, nblk((long)m * 32, 128), 128, m, ls.lvl_rows_lower.p + ls.lvl_ptr_lower[l], 0,
                  ls.row_ptr, ls.col_ind, ls.dia_ptr, ls.ilu.p, y);
    }
    for (int l = 0; l < ls.nlev_upper; ++l) {
      const int m = ls.lvl_ptr_upper[l + 1] - ls.lvl_ptr_upper[l];
      SU2B_LAUNCH(c, k_ilu_backward_level<NV>, nblk((long)m * 32, 128), 128, m, ls.lvl_rows_upper.p + ls.lvl_ptr_upper[l],
                  ls.Pd, ls.row_ptr, ls.col_ind, ls.dia_ptr, ls.ilu.p, ls.invM.p, y);
    }
  } else {
    SU2B_LAUNCH(c, k_ilu_apply_part<NV>, nblk((long)ls.nparts * 32, 64), 64, ls.nparts, ls.parts.p, ls.row_ptr, ls.col_ind,
                ls.dia_ptr, ls.ilu.p, ls.invM.p, y);
  }
}

void linsys_build_ilu(su2b200_ctx* c, LinSys& ls) {
  switch (ls.nVar) {
    case 1: ilu_build_impl<1>(c, ls); break;
    case 2: ilu_build_impl<2>(c, ls); break;
    case 4: ilu_build_impl<4>(c, ls); break;
    case 5: ilu_build_impl<5>(c, ls); break;
    default: throw Error("ilu: unsupported nVar");
  }
}
void linsys_apply_ilu(su2b200_ctx* c, LinSys& ls, const double* x, double* y) {
  switch (ls.nVar) {
    case 1: ilu_apply_impl<1>(c, ls, x, y); break;
    case 2: ilu_apply_impl<2>(c, ls, x, y); break;
    case 4: ilu_apply_impl<4>(c, ls, x, y); break;
    case 5: ilu_apply_impl<5>(c, ls, x, y); break;
    default: throw Error("ilu: unsupported nVar");
  }
  if (c->nranks > 1 && !c->defer_halo) halo_exchange(c, y, ls.nVar);
}

template <int NV>
static void lusgs_impl(su2b200_ctx* c, LinSys& ls, const double* b, double* x) {
  build_levels(ls);
  for (int l = 0; l < ls.nlev_lower; ++l) {
    const int m = ls.lvl_ptr_lower[l + 1] - ls.lvl_ptr_lower[l];
    SU2B_LAUNCH(c, k_lusgs_forward_level<NV>, nblk(m, 128), 128, m, ls.lvl_rows_lower.p + ls.lvl_ptr_lower[l], 0, ls.row_ptr,
                ls.col_ind, ls.dia_ptr, ls.A.p, b, x);
  }
  if (c->nranks > 1) halo_exchange(c, x, NV);
  for (int l = 0; l < ls.nlev_upper; ++l) {
    const int m = ls.lvl_ptr_upper[l + 1] - ls.lvl_ptr_upper[l];
    SU2B_LAUNCH(c, k_lusgs_backward_level<NV>, nblk(m, 128), 128, m, ls.lvl_rows_upper.p + ls.lvl_ptr_upper[l], ls.Pd, ls.Pd,
                ls.row_ptr, ls.col_ind, ls.dia_ptr, ls.A.p, x);
  }
  if (c->nranks > 1 && !c->defer_halo) halo_exchange(c, x, NV);
}

void linsys_apply_lusgs(su2b200_ctx* c, LinSys& ls, const double* b, double* x) {
  SU2B_REQUIRE(b != x, "LU-SGS needs distinct input and output vectors");
  switch (ls.nVar) {
    case 1: lusgs_impl<1>(c, ls, b, x); break;
    case 2: lusgs_impl<2>(c, ls, b, x); break;
    case 4: lusgs_impl<4>(c, ls, b, x); break;
    case 5: lusgs_impl<5>(c, ls, b, x); break;
    default: throw Error("lusgs: unsupported nVar");
  }
}

void linsys_build_prec(su2b200_ctx* c, LinSys& ls, int prec) {
  ProfScope prof_(c, PROF_PREC_BUILD);
  switch (prec) {
    case SU2B200_PREC_JACOBI: linsys_build_jacobi(c, ls); break;
    case SU2B200_PREC_ILU: linsys_build_ilu(c, ls); break;
    case SU2B200_PREC_LU_SGS: break;
    default: throw Error("unknown preconditioner");
  }
}
void linsys_apply_prec(su2b200_ctx* c, LinSys& ls, int prec, const double* x, double* y) {
  ProfScope prof_(c, PROF_PREC_APPLY);
  switch (prec) {
    case SU2B200_PREC_JACOBI: linsys_apply_jacobi(c, ls, x, y); break;
    case SU2B200_PREC_ILU: linsys_apply_ilu(c, ls, x, y); break;
    case SU2B200_PREC_LU_SGS: linsys_apply_lusgs(c, ls, x, y); break;
    default: throw Error("unknown preconditioner");
  }
}

}  // namespace su2b200
