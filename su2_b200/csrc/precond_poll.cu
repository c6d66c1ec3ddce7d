// a18. ILU0 application (CSysMatrix<T>::ComputeILUPreconditioner, CSysMatrix.cpp:770-823) as two single-launch,
// data-driven triangular sweeps.
//
// The level-scheduled form needs one launch per dependency level (553 levels forward + 553 backward on a 185^3 hex
// box) and is launch-latency bound; a flag-per-row form pays store -> fence -> flag -> acquire per level. Here the
// DATA is the flag: the output vector is pre-filled with a sentinel bit pattern (a NaN payload arithmetic never
// produces), a producer simply stores its results, and each consumer lane polls exactly the one scalar it needs
// (y[j][c] for lane q = r*nVar + c) until it differs from the sentinel. One L2 round trip per level, no fences.
//   forward : w_i = x_i - sum_{j<i} L_ij w_j          (x read-only, w = sentinel-filled scratch)
//   backward: y_i = U_ii^-1 (w_i - sum_{j>i} U_ij y_j)  (y sentinel-filled output)
// Rows are taken in level order, round-robin over the warps of a grid that the cooperative launch guarantees to be
// co-resident: the smallest unfinished row is always owned by a running warp whose earlier rows are done, so the
// sweeps cannot deadlock. A watchdog turns a dependency that never arrives into an error instead of a hang.
// Arithmetic per row and its summation order are those of the level kernels (== the reference's serial sweeps).
#include "internal.hpp"
#include "kernels.hpp"
#include <algorithm>

namespace su2b200 {

static constexpr unsigned FULLM = 0xffffffffu;
#ifndef POLL_SLEEP_NS
#define POLL_SLEEP_NS 0  // measured on 129^3: any __nanosleep in the poll loop doubles the sweep time (1.69 -> 3.4 ms)
#endif
static constexpr unsigned long long SENTINEL = 0xFFF8C0DEC0DE5A5AULL;  // quiet NaN with a private payload

__device__ __forceinline__ double ld_relaxed(const double* p) {
  double v;
  asm volatile("ld.relaxed.gpu.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed(double* p, double v) {
  asm volatile("st.relaxed.gpu.global.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}
__device__ __forceinline__ int ld_relaxed_i(const int* p) {
  int v;
  asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// spin until *p is no longer the sentinel; false if the watchdog fired (ctrl[1] set)
__device__ __forceinline__ bool poll(const double* p, int* ctrl, double& out) {
  unsigned spins = 0;
  for (;;) {
    const double v = ld_relaxed(p);
    if ((unsigned long long)__double_as_longlong(v) != SENTINEL) { out = v; return true; }
#if POLL_SLEEP_NS > 0
    __nanosleep(POLL_SLEEP_NS);  // a waiting warp must not eat the issue slots of the warps that have work
#endif
    if ((++spins & 1023u) == 0) {
      if (spins > (1u << 24)) atomicExch(ctrl + 1, 1);
      if (ld_relaxed_i(ctrl + 1) != 0) return false;
    }
  }
}

__global__ void k_fill_sentinel(size_t n, double* __restrict__ a, double* __restrict__ b) {
  const double s = __longlong_as_double((long long)SENTINEL);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    a[i] = s;
    if (b) b[i] = s;
  }
}

// ---- block systems: one warp per row ---------------------------------------------------------------------------------
// Row t of the level-ordered list belongs to warp t mod W (W = every warp of the co-resident grid), so a warp knows its
// future rows: while it polls for row t it already holds the matrix entries of row t+W and the descriptor of row t+2W.
// Only the poll -> fma -> store hop of a row is on the dependency chain.
// factor entries are read once per sweep: streaming loads (evict-first) leave the L2 to the vector being polled
#ifdef POLL_STREAM_LOADS
#define POLL_LD(p) __ldcs(p)
#else
#define POLL_LD(p) (*(p))
#endif
template <int NV>
struct RowData {
  static constexpr int CH = 4;  // entries fetched ahead (3 lower/upper neighbours on a hex box, more on tets: tail loop)
  double a[CH];
  int jj[CH];
  double rhs, minv;
};
template <int NV, bool FORWARD>
__device__ __forceinline__ void fetch_row(RowData<NV>& d, const int4 m, int q, int r, const int* __restrict__ col_ind,
                                          const double* __restrict__ ILU, const double* __restrict__ invM,
                                          const double* __restrict__ in) {
  constexpr int LPB = NV * NV;
#pragma unroll
  for (int k = 0; k < RowData<NV>::CH; ++k) {
    const int idx = m.y + k < m.z ? m.y + k : (m.z > m.y ? m.z - 1 : m.y);
    d.jj[k] = m.z > m.y ? POLL_LD(col_ind + idx) : 0;
    d.a[k] = m.z > m.y ? POLL_LD(ILU + (size_t)idx * LPB + q) : 0.0;
  }
  d.rhs = in[(size_t)m.x * NV + r];
  d.minv = FORWARD ? 0.0 : invM[(size_t)m.x * LPB + q];
}

template <int NV, bool FORWARD>
__global__ void __launch_bounds__(128) k_poll_sweep(int n, const int4* __restrict__ meta, int* ctrl,
                                                    const int* __restrict__ col_ind, const double* __restrict__ ILU,
                                                    const double* __restrict__ invM, const double* __restrict__ in,
                                                    double* out) {
  constexpr int LPB = NV * NV;
  constexpr int CH = RowData<NV>::CH;
  const int lane = threadIdx.x & 31;
  const int q = lane < LPB ? lane : LPB - 1;
  const int r = q / NV, cc = q % NV;
  const int W = gridDim.x * (blockDim.x >> 5);
  int t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (t >= n) return;
  int4 m_cur = meta[t];
  int4 m_nxt = t + W < n ? meta[t + W] : m_cur;
  RowData<NV> cur, nxt;
  fetch_row<NV, FORWARD>(cur, m_cur, q, r, col_ind, ILU, invM, in);
  for (; t < n; t += W) {
    const int4 m_nn = t + 2 * W < n ? meta[t + 2 * W] : m_nxt;
    if (t + W < n) fetch_row<NV, FORWARD>(nxt, m_nxt, q, r, col_ind, ILU, invM, in);
    const int i = m_cur.x, kb = m_cur.y, ke = m_cur.z;
    double acc = 0.0;
    double yv[CH];
#pragma unroll
    for (int k = 0; k < CH; ++k)  // all first looks in flight together: one L2 round trip when the row is ready
      yv[k] = kb + k < ke ? ld_relaxed(out + (size_t)cur.jj[k] * NV + cc) : 0.0;
#pragma unroll
    for (int k = 0; k < CH; ++k) {
      if (kb + k < ke) {
        if ((unsigned long long)__double_as_longlong(yv[k]) == SENTINEL &&
            !poll(out + (size_t)cur.jj[k] * NV + cc, ctrl, yv[k]))
          return;  // watchdog
        acc += cur.a[k] * yv[k];
      }
    }
    for (int idx = kb + CH; idx < ke; ++idx) {  // rows with more than CH entries on this side
      const int j = col_ind[idx];
      const double a = ILU[(size_t)idx * LPB + q];
      double yj;
      if (!poll(out + (size_t)j * NV + cc, ctrl, yj)) return;
      acc += a * yj;
    }
    double s = acc;
#pragma unroll
    for (int o = 1; o < NV; ++o) s += __shfl_down_sync(FULLM, acc, o);
    if (FORWARD) {
      if (lane < LPB && cc == 0) st_relaxed(out + (size_t)i * NV + r, cur.rhs - s);
    } else {
      const double aux_r = cur.rhs - s;  // valid on lanes r*NV
      const double aux_c = __shfl_sync(FULLM, aux_r, cc * NV);
      const double prod = cur.minv * aux_c;
      double tsum = prod;
#pragma unroll
      for (int o = 1; o < NV; ++o) tsum += __shfl_down_sync(FULLM, prod, o);
      if (lane < LPB && cc == 0) st_relaxed(out + (size_t)i * NV + r, tsum);
    }
    cur = nxt;
    m_cur = m_nxt;
    m_nxt = m_nn;
  }
}

// ---- scalar systems: 8 lanes per row (one lane per entry on this side of the diagonal), 4 rows per warp ----------------
// Few rows per warp on purpose: a warp advances when the slowest of its rows has its dependencies, and the sweep moves
// at the pace of the most coupled wavefront.
static constexpr int SROWS = 4, SLANES = 8;
template <bool FORWARD>
__global__ void __launch_bounds__(128) k_poll_sweep_s(int nGroups, const int4* __restrict__ meta, int* ctrl,
                                                      const int* __restrict__ col_ind, const double* __restrict__ ILU,
                                                      const double* __restrict__ invM, const double* __restrict__ in,
                                                      double* out) {
  const int lane = threadIdx.x & 31;
  const int sub = lane / SLANES, e = lane % SLANES, base = sub * SLANES;
  const int W = gridDim.x * (blockDim.x >> 5);
  int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (g >= nGroups) return;
  struct Row { double a, rhs, minv; int j; };
  auto fetch = [&](const int4 m, Row& d) {
    const bool has = m.x >= 0 && m.y + e < m.z;
    d.j = has ? col_ind[m.y + e] : -1;
    d.a = has ? ILU[m.y + e] : 0.0;
    d.rhs = m.x >= 0 ? in[m.x] : 0.0;
    d.minv = (!FORWARD && m.x >= 0) ? invM[m.x] : 1.0;
  };
  int4 m_cur = meta[(size_t)g * SROWS + sub];
  int4 m_nxt = g + W < nGroups ? meta[(size_t)(g + W) * SROWS + sub] : m_cur;
  Row cur, nxt;
  fetch(m_cur, cur);
  nxt = cur;
  bool ok = true;  // after the watchdog fired the lanes idle through their remaining rows (shuffles need them all)
  for (; g < nGroups; g += W) {
    const int4 m_nn = g + 2 * W < nGroups ? meta[(size_t)(g + 2 * W) * SROWS + sub] : m_nxt;
    if (g + W < nGroups) fetch(m_nxt, nxt);
    double prod = 0.0;
    if (ok && cur.j >= 0) {
      double yj;
      ok = poll(out + cur.j, ctrl, yj);
      if (ok) prod = cur.a * yj;
    }
    double v = cur.rhs;
#pragma unroll
    for (int k = 0; k < SLANES; ++k) v -= __shfl_sync(FULLM, prod, base + k);  // the reference's order of subtraction
    if (m_cur.x >= 0 && m_cur.z - m_cur.y > SLANES && e == 0 && ok) {  // rows with more than 8 entries on this side
      for (int idx = m_cur.y + SLANES; idx < m_cur.z && ok; ++idx) {
        double yj;
        ok = poll(out + col_ind[idx], ctrl, yj);
        if (ok) v -= ILU[idx] * yj;
      }
    }
    if (ok && e == 0 && m_cur.x >= 0) st_relaxed(out + m_cur.x, FORWARD ? v : cur.minv * v);
    ok = __all_sync(FULLM, ok);
    cur = nxt;
    m_cur = m_nxt;
    m_nxt = m_nn;
  }
}

// Every warp of the grid must be resident (rows are assigned round-robin and a warp may wait for any smaller row):
// cooperative launch makes the driver enforce that, the occupancy query sizes the grid.
template <class K, class... Args>
static void launch_resident(su2b200_ctx* c, K kernel, int maxBlocksUseful, int blocksPerSmCap, Args... args) {
  int dev = 0, nsm = 0, perSm = 0;
  SU2B_CUDA(cudaGetDevice(&dev));
  SU2B_CUDA(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
  SU2B_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, kernel, 128, 0));
  if (blocksPerSmCap > 0) perSm = std::min(perSm, blocksPerSmCap);
  const int grid = std::max(1, std::min(maxBlocksUseful, nsm * perSm));
  void* argv[] = {(void*)&args...};
  SU2B_CUDA(cudaLaunchCooperativeKernel((const void*)kernel, dim3(grid), dim3(128), argv, 0, c->stream));
  c->launches++;
}

// level-ordered row descriptors {row, first entry, one past the last entry on this side of the diagonal, 0}
static void build_meta(LinSys& ls) {
  if (ls.sweep_meta_lower.n) return;
  const int n = ls.Pd;
  const auto& rp = ls.h_row_ptr;
  const auto& ci = ls.h_col_ind;
  const auto& dp = ls.h_dia_ptr;
  auto make = [&](const std::vector<int>& rows, bool lower, dvec<int>& out) {
    std::vector<int> m(rows.size() * 4);
    for (size_t t = 0; t < rows.size(); ++t) {
      const int i = rows[t];
      int kb = 0, ke = 0;
      if (i >= 0) {
        if (lower) { kb = rp[i]; ke = dp[i]; }
        else {
          kb = dp[i] + 1; ke = rp[i + 1];
          while (ke > kb && ci[ke - 1] >= n) --ke;  // halo columns take no part in the owned factorisation
        }
      }
      m[4 * t] = i; m[4 * t + 1] = kb; m[4 * t + 2] = ke; m[4 * t + 3] = 0;
    }
    out.upload(m);
  };
  if (ls.nVar == 1) {
    // scalar sweeps take SROWS rows per warp: pad every level so that the rows of a warp never depend on one another
    auto pad = [&](const std::vector<int>& ptr, const std::vector<int>& rows, int& groups) {
      std::vector<int> p;
      p.reserve(rows.size() + SROWS * ptr.size());
      for (size_t l = 0; l + 1 < ptr.size(); ++l) {
        for (int k = ptr[l]; k < ptr[l + 1]; ++k) p.push_back(rows[k]);
        while (p.size() % SROWS) p.push_back(-1);
      }
      groups = (int)(p.size() / SROWS);
      return p;
    };
    make(pad(ls.lvl_ptr_lower, ls.lvl_rows_lower.to_host(), ls.poll_groups_lower), true, ls.sweep_meta_lower);
    make(pad(ls.lvl_ptr_upper, ls.lvl_rows_upper.to_host(), ls.poll_groups_upper), false, ls.sweep_meta_upper);
  } else {
    make(ls.lvl_rows_lower.to_host(), true, ls.sweep_meta_lower);
    make(ls.lvl_rows_upper.to_host(), false, ls.sweep_meta_upper);
  }
}

// x -> y through the scratch vector w; all three distinct (x may alias y only through the copy made by the caller)
template <int NV>
static void ilu_apply_poll_impl(su2b200_ctx* c, LinSys& ls, const double* x, double* y, double* w) {
  const size_t n = (size_t)ls.Pd * NV;
  build_meta(ls);
  SU2B_LAUNCH(c, k_fill_sentinel, (int)std::min<size_t>((n + 255) / 256, 148 * 8), 256, n, w, y);
  const int4* ml = reinterpret_cast<const int4*>(ls.sweep_meta_lower.p);
  const int4* mu = reinterpret_cast<const int4*>(ls.sweep_meta_upper.p);
  const int* ci = ls.col_ind;
  const double* ilu = ls.ilu.p;
  const double* invM = ls.invM.p;
  int* ctrl = ls.sweep_ctrl.p;
  // resident blocks per SM: every waiting lane of a scalar sweep polls its own sector, so too many waiting warps
  // saturate L2 with polls and slow the rows on the critical path; block rows poll one or two sectors per warp
  static const int capS = getenv("SU2B200_SWEEP_BLOCKS_S") ? atoi(getenv("SU2B200_SWEEP_BLOCKS_S")) : 4;
  static const int capB = getenv("SU2B200_SWEEP_BLOCKS") ? atoi(getenv("SU2B200_SWEEP_BLOCKS")) : 0;
  if constexpr (NV == 1) {
    launch_resident(c, k_poll_sweep_s<true>, (ls.poll_groups_lower + 3) / 4, capS, ls.poll_groups_lower, ml, ctrl, ci, ilu, invM, x, w);
    launch_resident(c, k_poll_sweep_s<false>, (ls.poll_groups_upper + 3) / 4, capS, ls.poll_groups_upper, mu, ctrl, ci, ilu, invM,
                    (const double*)w, y);
  } else {
    launch_resident(c, k_poll_sweep<NV, true>, (ls.Pd + 3) / 4, capB, ls.Pd, ml, ctrl, ci, ilu, invM, x, w);
    launch_resident(c, k_poll_sweep<NV, false>, (ls.Pd + 3) / 4, capB, ls.Pd, mu, ctrl, ci, ilu, invM, (const double*)w, y);
  }
}

void ilu_apply_poll(su2b200_ctx* c, LinSys& ls, const double* x, double* y) {
  const size_t n = ls.vecLen();
  if (ls.sweep_tmp.n != n) { ls.sweep_tmp.alloc(n); ls.sweep_tmp.zero(c->stream); }
  if (ls.sweep_tmp2.n != n && x == y) { ls.sweep_tmp2.alloc(n); ls.sweep_tmp2.zero(c->stream); }
  const double* xin = x;
  if (x == y) {  // in-place request: the rhs must survive the sentinel fill of y
    SU2B_CUDA(cudaMemcpyAsync(ls.sweep_tmp2.p, x, (size_t)ls.Pd * ls.nVar * sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
    xin = ls.sweep_tmp2.p;
  }
  switch (ls.nVar) {
    case 1: ilu_apply_poll_impl<1>(c, ls, xin, y, ls.sweep_tmp.p); break;
    case 4: ilu_apply_poll_impl<4>(c, ls, xin, y, ls.sweep_tmp.p); break;
    case 5: ilu_apply_poll_impl<5>(c, ls, xin, y, ls.sweep_tmp.p); break;
    default: throw Error("ilu: unsupported nVar");
  }
}

}  // namespace su2b200
