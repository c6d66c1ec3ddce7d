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

// Warp-converged form (every lane of the warp calls it; lanes whose v holds the sentinel re-load it; all leave
// together). Per-lane spin loops split the warp into single-lane paths that take turns on the load pipe: 2.6 us per
// level instead of 0.87 us in the scalar sweeps. false = the watchdog fired (for all lanes).
__device__ __forceinline__ bool poll_warp(const double* p, int* ctrl, double& v) {
  for (unsigned spins = 0;; ++spins) {
    const bool miss = (unsigned long long)__double_as_longlong(v) == SENTINEL;
    if (!__any_sync(FULLM, miss)) return true;
    if (miss) v = ld_relaxed(p);
    if ((spins & 1023u) == 1023u) {
      if (spins > (1u << 24)) atomicExch(ctrl + 1, 1);
      if (__any_sync(FULLM, ld_relaxed_i(ctrl + 1) != 0)) return false;
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
                                                    double* out, double* sout) {
  // sout (backward sweep only, may be null): receives sum_{j>i} U_ij x_j of every row, the upper part of the product
  // A x when U_ij == A_ij (linalg.cu: linsys_spmv with the fused upper product)
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
    // whatever is missing is looked at again, all of it per round (one L2 round trip per round, not per dependency)
    for (unsigned spins = 0;; ++spins) {
      bool miss = false;
#pragma unroll
      for (int k = 0; k < CH; ++k)
        if ((unsigned long long)__double_as_longlong(yv[k]) == SENTINEL) {
          miss = true;
          yv[k] = ld_relaxed(out + (size_t)cur.jj[k] * NV + cc);
        }
      if (!__any_sync(FULLM, miss)) break;
      if ((spins & 1023u) == 1023u) {
        if (spins > (1u << 22)) atomicExch(ctrl + 1, 1);
        if (ld_relaxed_i(ctrl + 1) != 0) return;  // watchdog (warp-uniform: the vote above keeps the lanes together)
      }
    }
#pragma unroll
    for (int k = 0; k < CH; ++k)
      if (kb + k < ke) acc += cur.a[k] * yv[k];
    for (int idx = kb + CH; idx < ke; ++idx) {  // rows with more than CH entries on this side
      const int j = col_ind[idx];
      const double a = ILU[(size_t)idx * LPB + q];
      double yj = __longlong_as_double((long long)SENTINEL);
      if (!poll_warp(out + (size_t)j * NV + cc, ctrl, yj)) return;
      acc += a * yj;
    }
    double s = acc;
#pragma unroll
    for (int o = 1; o < NV; ++o) s += __shfl_down_sync(FULLM, acc, o);
    if (FORWARD) {
      if (lane < LPB && cc == 0) st_relaxed(out + (size_t)i * NV + r, cur.rhs - s);
    } else {
      if (sout && lane < LPB && cc == 0) sout[(size_t)i * NV + r] = s;
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

// ---- the same sweep with a deeper, shared-memory pipeline --------------------------------------------------------------
// In the wide levels every row is ready when its warp reaches it and the sweep is bound by how many bytes of the factor
// are in flight: with a one-row register prefetch a warp waits for DRAM at every row. Here the entries, columns, rhs and
// inverse diagonal of the next PIPE-1 rows are streamed into a ring of PIPE shared-memory stages with cp.async.
static constexpr int PIPE = 4;
template <int NV>
struct SweepStage {  // one row
  static constexpr int CH = 4;
  double a[CH][NV * NV];
  double minv[NV * NV];
  double rhs[NV];
  int col[CH];
};
__device__ __forceinline__ void cp_async_8(void* smem, const void* gmem) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_4(void* smem, const void* gmem) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s), "l"(gmem) : "memory");
}

template <int NV, bool FORWARD>
__global__ void __launch_bounds__(128) k_poll_sweep_pipe(int n, const int4* __restrict__ meta, int* ctrl,
                                                         const int* __restrict__ col_ind, const double* __restrict__ ILU,
                                                         const double* __restrict__ invM, const double* __restrict__ in,
                                                         double* out) {
  constexpr int LPB = NV * NV;
  constexpr int CH = SweepStage<NV>::CH;
  extern __shared__ double smem_raw[];
  SweepStage<NV>* ring = reinterpret_cast<SweepStage<NV>*>(smem_raw) + (threadIdx.x >> 5) * PIPE;
  const int lane = threadIdx.x & 31;
  const int q = lane < LPB ? lane : LPB - 1;
  const int r = q / NV, cc = q % NV;
  const int W = gridDim.x * (blockDim.x >> 5);
  int t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (t >= n) return;

  auto stage = [&](const int4 m, SweepStage<NV>& s) {
    if (lane < LPB) {
#pragma unroll
      for (int k = 0; k < CH; ++k)
        if (m.y + k < m.z) cp_async_8(&s.a[k][lane], ILU + (size_t)(m.y + k) * LPB + lane);
      if (!FORWARD) cp_async_8(&s.minv[lane], invM + (size_t)m.x * LPB + lane);
    }
    if (lane < NV) cp_async_8(&s.rhs[lane], in + (size_t)m.x * NV + lane);
    if (lane < CH && m.y + lane < m.z) cp_async_4(&s.col[lane], col_ind + m.y + lane);
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  // descriptors PIPE rows ahead, data PIPE-1 rows ahead
  int4 m[PIPE];
#pragma unroll
  for (int p = 0; p < PIPE; ++p) m[p] = t + p * W < n ? meta[t + p * W] : make_int4(0, 0, 0, 0);
#pragma unroll
  for (int p = 0; p < PIPE - 1; ++p) {
    if (t + p * W < n) stage(m[p], ring[p]);
    else asm volatile("cp.async.commit_group;" ::: "memory");
  }
  int slot = 0;
  // first looks at the dependencies are issued one row ahead as well: in the wide levels the values are there long
  // before the warp arrives, and the L2 round trip of the look would otherwise be paid once per row per warp
  double yv[CH], yvN[CH];
  int jj[CH], jjN[CH];
  auto look = [&](const int4 mm, const SweepStage<NV>& s, int (&cols)[CH], double (&vals)[CH]) {
#pragma unroll
    for (int k = 0; k < CH; ++k) {
      cols[k] = mm.y + k < mm.z ? s.col[k] : 0;
      vals[k] = mm.y + k < mm.z ? ld_relaxed(out + (size_t)cols[k] * NV + cc) : 0.0;
    }
  };
  asm volatile("cp.async.wait_group %0;" ::"n"(PIPE - 2) : "memory");
  __syncwarp();
  look(m[0], ring[0], jj, yv);
  for (; t < n; t += W) {
    const int4 m_new = t + PIPE * W < n ? meta[t + PIPE * W] : make_int4(0, 0, 0, 0);
    if (t + (PIPE - 1) * W < n) stage(m[PIPE - 1], ring[(slot + PIPE - 1) % PIPE]);
    else asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group %0;" ::"n"(PIPE - 2) : "memory");  // stages of this row and of the next one
    __syncwarp();
    if (t + W < n) look(m[1], ring[(slot + 1) % PIPE], jjN, yvN);
    const SweepStage<NV>& s = ring[slot];
    const int i = m[0].x, kb = m[0].y, ke = m[0].z;
    double acc = 0.0;
#pragma unroll
    for (int k = 0; k < CH; ++k) {
      if (kb + k < ke) {
        if ((unsigned long long)__double_as_longlong(yv[k]) == SENTINEL &&
            !poll(out + (size_t)jj[k] * NV + cc, ctrl, yv[k]))
          return;  // watchdog
        acc += s.a[k][q] * yv[k];
      }
    }
    for (int idx = kb + CH; idx < ke; ++idx) {  // rows with more than CH entries on this side
      const int j = col_ind[idx];
      const double a = ILU[(size_t)idx * LPB + q];
      double yj = __longlong_as_double((long long)SENTINEL);
      if (!poll_warp(out + (size_t)j * NV + cc, ctrl, yj)) return;
      acc += a * yj;
    }
    double sum = acc;
#pragma unroll
    for (int o = 1; o < NV; ++o) sum += __shfl_down_sync(FULLM, acc, o);
    const double rhs = s.rhs[r];
    if (FORWARD) {
      if (lane < LPB && cc == 0) st_relaxed(out + (size_t)i * NV + r, rhs - sum);
    } else {
      const double aux_r = rhs - sum;  // valid on lanes r*NV
      const double aux_c = __shfl_sync(FULLM, aux_r, cc * NV);
      const double prod = s.minv[q] * aux_c;
      double tsum = prod;
#pragma unroll
      for (int o = 1; o < NV; ++o) tsum += __shfl_down_sync(FULLM, prod, o);
      if (lane < LPB && cc == 0) st_relaxed(out + (size_t)i * NV + r, tsum);
    }
    __syncwarp();  // every lane is done with this stage before it is refilled
#pragma unroll
    for (int p = 0; p < PIPE - 1; ++p) m[p] = m[p + 1];
    m[PIPE - 1] = m_new;
    slot = (slot + 1) % PIPE;
#pragma unroll
    for (int k = 0; k < CH; ++k) { yv[k] = yvN[k]; jj[k] = jjN[k]; }
  }
}

// ---- block systems, nVar lanes per row ----------------------------------------------------------------------------------
// ncu showed the warp-per-row sweeps to be bound by instruction issue in the wide levels (every resident warp has a
// ready row; ~210 warp instructions per row, 25 of 32 lanes busy). Here lane r of a group of nVar lanes owns block row
// r: it loads the nVar entries of its row of each block (the group reads the 200 contiguous bytes of the block), the
// nVar values of the dependency, and accumulates exactly as the reference does (product of block and vector summed
// over the column, then subtracted from the row). 32/nVar rows per warp and ~30 warp instructions per row. The rows of
// a warp belong to one level (levels padded), so they never wait for one another.
#ifndef ROWS_CH
#define ROWS_CH 3
#endif
#ifndef ROWS_U
#define ROWS_U 1  // row groups per warp and iteration of the wide-level kernel (k_poll_sweep_rowsU); levels are padded to it
#endif
#ifndef ROWS_MIN_BLOCKS
#define ROWS_MIN_BLOCKS 5
#endif
template <int NV, bool FORWARD>
__global__ void __launch_bounds__(128, ROWS_MIN_BLOCKS) k_poll_sweep_rows(int nGroups, const int4* __restrict__ meta, int* ctrl,
                                                         const int* __restrict__ col_ind, const double* __restrict__ ILU,
                                                         const double* __restrict__ invM, const double* __restrict__ in,
                                                         double* out, double* sout) {
  constexpr int LPB = NV * NV;
  constexpr int RPW = 32 / NV;  // rows per warp
  constexpr int CH = ROWS_CH;   // entries held in registers (a hex row has 3 on each side); longer rows: tail loop
  const int lane = threadIdx.x & 31;
  const int sub = lane / NV, r = lane % NV;
  const bool live = sub < RPW;
  const int W = gridDim.x * (blockDim.x >> 5);
  bool ok = true;  // after the watchdog fired the lanes idle through their remaining rows (shuffles need them all)
  for (int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); g < nGroups; g += W) {
    const int4 m = live ? meta[(size_t)g * RPW + sub] : make_int4(-1, 0, 0, 0);
    const int i = m.x, kb = m.y, ke = m.z;
    const bool on = ok && i >= 0;
    // everything that does not depend on the sweep first: columns, block rows, rhs (and the inverse diagonal)
    int jj[CH];
    double a[CH][NV];
#pragma unroll
    for (int k = 0; k < CH; ++k) {
      const bool has = on && kb + k < ke;
      jj[k] = has ? col_ind[kb + k] : 0;
#pragma unroll
      for (int c = 0; c < NV; ++c) a[k][c] = has ? ILU[(size_t)(kb + k) * LPB + r * NV + c] : 0.0;
    }
    double v = on ? in[(size_t)i * NV + r] : 0.0;
    double dinv[NV];
    if (!FORWARD) {
#pragma unroll
      for (int c = 0; c < NV; ++c) dinv[c] = on ? invM[(size_t)i * LPB + r * NV + c] : 0.0;
    }
    // lane r watches component r of every dependency (one or two sectors per row and dependency: with each lane
    // re-reading all nVar components the waiting warps saturated the L2 and the hop from level to level took 2.4 us);
    // first look at all of them, then all that is still missing again per round, the warp as one instruction stream
    double mine[CH];
#pragma unroll
    for (int k = 0; k < CH; ++k) mine[k] = (on && kb + k < ke) ? ld_relaxed(out + (size_t)jj[k] * NV + r) : 0.0;
    for (unsigned spins = 0;; ++spins) {
      bool miss = false;
#pragma unroll
      for (int k = 0; k < CH; ++k)
        if ((unsigned long long)__double_as_longlong(mine[k]) == SENTINEL) {
          miss = true;
          mine[k] = ld_relaxed(out + (size_t)jj[k] * NV + r);
        }
      if (!__any_sync(FULLM, miss)) break;
      if ((spins & 1023u) == 1023u) {
        if (spins > (1u << 22)) atomicExch(ctrl + 1, 1);
        if (__any_sync(FULLM, ld_relaxed_i(ctrl + 1) != 0)) { ok = false; break; }
      }
    }
    double y[CH][NV];
#pragma unroll
    for (int k = 0; k < CH; ++k)
#pragma unroll
      for (int c = 0; c < NV; ++c) y[k][c] = __shfl_sync(FULLM, mine[k], sub * NV + c);
    double su = 0.0;
#pragma unroll
    for (int k = 0; k < CH; ++k) {
      if (on && ok && kb + k < ke) {
        double p = 0.0;
#pragma unroll
        for (int c = 0; c < NV; ++c) p += a[k][c] * y[k][c];
        v -= p;
        if (!FORWARD) su += p;
      }
    }
    if (on) {
      for (int idx = kb + CH; idx < ke && ok; ++idx) {  // rows with more than CH entries on this side
        const int j = col_ind[idx];
        double p = 0.0;
        for (int c = 0; c < NV && ok; ++c) {
          double yj;
          ok = poll(out + (size_t)j * NV + c, ctrl, yj);
          p += ILU[(size_t)idx * LPB + r * NV + c] * yj;
        }
        v -= p;
        if (!FORWARD) su += p;
      }
    }
    if (FORWARD) {
      if (on && ok) st_relaxed(out + (size_t)i * NV + r, v);
    } else {
      if (sout && on && ok) sout[(size_t)i * NV + r] = su;
      // x_i = invM_i * aux (MatrixVectorProduct): lane r needs the whole aux vector of its row
      double x = 0.0;
#pragma unroll
      for (int c = 0; c < NV; ++c) x += dinv[c] * __shfl_sync(FULLM, v, sub * NV + c);
      if (on && ok) st_relaxed(out + (size_t)i * NV + r, x);
    }
    ok = __all_sync(FULLM, ok);
  }
}

// ---- the same mapping with descriptors and columns prefetched, U row groups per warp and iteration -------------------------
// What bounds the wide levels (measured on the 185^3 RANS matrix, RCM order; profiles/r02_ilu_apply_ab.txt):
//   * run WITHOUT the sentinel fill (every dependency "there" at the first look, i.e. as pure streaming kernels) one
//     application takes 2.10 ms = 5.2 TB/s on the algorithmic bytes (1.99 ms with every level in this kernel): that is
//     the bandwidth bound of this data layout (600 of every 1400 bytes of a block row are read per sweep);
//   * with the dependencies it takes 2.78 ms: the chain of 2 x 553 levels adds 0.68 ms (0.6 us per level, one L2 round
//     trip), and nothing tried moves it: a cp.async ring two/three groups deep (3.0-3.6 ms, fewer resident warps),
//     looking at the next group's dependencies one iteration early (3.2 ms: a warp's next group sits a whole wave
//     further on, mostly in the NEXT level), U = 2 or 3 groups per iteration as one instruction stream (2.86-3.5 ms),
//     20 instead of 16 resident warps (2.77 ms), level-width thresholds 256...4096 (2.94-2.97 ms at U = 2).
// Shipped: U = 1, descriptors two iterations and columns one iteration ahead in registers (the descriptor -> columns
// part of the latency chain of an iteration is gone), 16 resident warps per SM. Same arithmetic in the same order as
// k_poll_sweep_rows: bit-identical results.
#ifndef ROWSU_MIN_BLOCKS
#define ROWSU_MIN_BLOCKS 4
#endif
template <int NV, bool FORWARD, int U>
__global__ void __launch_bounds__(128, ROWSU_MIN_BLOCKS) k_poll_sweep_rowsU(int nIters, const int4* __restrict__ meta, int* ctrl,
                                                          const int* __restrict__ col_ind, const double* __restrict__ ILU,
                                                          const double* __restrict__ invM, const double* __restrict__ in,
                                                          double* out, double* sout) {
  constexpr int LPB = NV * NV, RPW = 32 / NV, CH = ROWS_CH;
  const int lane = threadIdx.x & 31;
  const int sub = lane / NV, r = lane % NV;
  const bool live = sub < RPW;
  const int W = gridDim.x * (blockDim.x >> 5);
  const int q0 = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (q0 >= nIters) return;
  auto load_meta = [&](int q, int u) {
    return (live && q < nIters) ? meta[((size_t)q * U + u) * RPW + sub] : make_int4(-1, 0, 0, 0);
  };
  int4 m0[U], m1[U];  // descriptors of iteration q and q+W
  int jj[U][CH];      // columns of iteration q
#pragma unroll
  for (int u = 0; u < U; ++u) {
    m0[u] = load_meta(q0, u);
    m1[u] = load_meta(q0 + W, u);
  }
#pragma unroll
  for (int u = 0; u < U; ++u)
#pragma unroll
    for (int k = 0; k < CH; ++k) jj[u][k] = (m0[u].x >= 0 && m0[u].y + k < m0[u].z) ? col_ind[m0[u].y + k] : 0;
  bool ok = true;  // after the watchdog fired the lanes idle through their remaining rows (shuffles need them all)
  for (int q = q0; q < nIters; q += W) {
    // prefetch: descriptors two iterations ahead, columns one iteration ahead
    int4 m2[U];
    int jn[U][CH];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      m2[u] = load_meta(q + 2 * W, u);
#pragma unroll
      for (int k = 0; k < CH; ++k) jn[u][k] = (m1[u].x >= 0 && m1[u].y + k < m1[u].z) ? col_ind[m1[u].y + k] : 0;
    }
    // everything that does not depend on the sweep: block rows, rhs (and the inverse diagonal), for all U groups
    bool on[U];
    double a[U][CH][NV], v[U], dinv[U][NV], mine[U][CH], su[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      on[u] = ok && m0[u].x >= 0;
#pragma unroll
      for (int k = 0; k < CH; ++k) {
        const bool has = on[u] && m0[u].y + k < m0[u].z;
        mine[u][k] = has ? ld_relaxed(out + (size_t)jj[u][k] * NV + r) : 0.0;
#pragma unroll
        for (int c = 0; c < NV; ++c) a[u][k][c] = has ? ILU[(size_t)(m0[u].y + k) * LPB + r * NV + c] : 0.0;
      }
      v[u] = on[u] ? in[(size_t)m0[u].x * NV + r] : 0.0;
      su[u] = 0.0;
      if (!FORWARD) {
#pragma unroll
        for (int c = 0; c < NV; ++c) dinv[u][c] = on[u] ? invM[(size_t)m0[u].x * LPB + r * NV + c] : 0.0;
      }
    }
    // lane r watches component r of every dependency of its rows; whatever is missing is looked at again, all of it
    // per round, the warp as one instruction stream
    for (unsigned spins = 0;; ++spins) {
      bool miss = false;
#pragma unroll
      for (int u = 0; u < U; ++u)
#pragma unroll
        for (int k = 0; k < CH; ++k)
          if ((unsigned long long)__double_as_longlong(mine[u][k]) == SENTINEL) {
            miss = true;
            mine[u][k] = ld_relaxed(out + (size_t)jj[u][k] * NV + r);
          }
      if (!__any_sync(FULLM, miss)) break;
      if ((spins & 1023u) == 1023u) {
        if (spins > (1u << 22)) atomicExch(ctrl + 1, 1);
        if (__any_sync(FULLM, ld_relaxed_i(ctrl + 1) != 0)) { ok = false; break; }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
#pragma unroll
      for (int k = 0; k < CH; ++k) {
        double y[NV];
#pragma unroll
        for (int c = 0; c < NV; ++c) y[c] = __shfl_sync(FULLM, mine[u][k], sub * NV + c);
        if (on[u] && ok && m0[u].y + k < m0[u].z) {
          double p = 0.0;
#pragma unroll
          for (int c = 0; c < NV; ++c) p += a[u][k][c] * y[c];
          v[u] -= p;
          if (!FORWARD) su[u] += p;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int i = m0[u].x, kb = m0[u].y, ke = m0[u].z;
      // rows with more than CH entries on this side (tetrahedra, prisms: up to ~8): further chunks of CH entries, each
      // with the same warp-converged look-and-poll as the first one (the serial per-entry, per-component polls this
      // replaces took 17-24 ms per application on a tetrahedral box, against 5.8 ms for the warp-per-row kernel).
      // Same products, same order of subtraction: bit-identical to the one-by-one loop.
      const int longest = __reduce_max_sync(FULLM, on[u] ? ke - kb : 0);
      for (int k0 = CH; k0 < longest && ok; k0 += CH) {
        int jc[CH];
        double ac[CH][NV], mc[CH];
#pragma unroll
        for (int k = 0; k < CH; ++k) {
          const bool has = on[u] && kb + k0 + k < ke;
          jc[k] = has ? col_ind[kb + k0 + k] : 0;
#pragma unroll
          for (int c = 0; c < NV; ++c) ac[k][c] = has ? ILU[(size_t)(kb + k0 + k) * LPB + r * NV + c] : 0.0;
        }
#pragma unroll
        for (int k = 0; k < CH; ++k)
          mc[k] = (on[u] && kb + k0 + k < ke) ? ld_relaxed(out + (size_t)jc[k] * NV + r) : 0.0;
        for (unsigned spins = 0;; ++spins) {
          bool miss = false;
#pragma unroll
          for (int k = 0; k < CH; ++k)
            if ((unsigned long long)__double_as_longlong(mc[k]) == SENTINEL) {
              miss = true;
              mc[k] = ld_relaxed(out + (size_t)jc[k] * NV + r);
            }
          if (!__any_sync(FULLM, miss)) break;
          if ((spins & 1023u) == 1023u) {
            if (spins > (1u << 22)) atomicExch(ctrl + 1, 1);
            if (__any_sync(FULLM, ld_relaxed_i(ctrl + 1) != 0)) { ok = false; break; }
          }
        }
#pragma unroll
        for (int k = 0; k < CH; ++k) {
          double y[NV];
#pragma unroll
          for (int c = 0; c < NV; ++c) y[c] = __shfl_sync(FULLM, mc[k], sub * NV + c);
          if (on[u] && ok && kb + k0 + k < ke) {
            double p = 0.0;
#pragma unroll
            for (int c = 0; c < NV; ++c) p += ac[k][c] * y[c];
            v[u] -= p;
            if (!FORWARD) su[u] += p;
          }
        }
      }
      if (FORWARD) {
        if (on[u] && ok) st_relaxed(out + (size_t)i * NV + r, v[u]);
      } else {
        if (sout && on[u] && ok) sout[(size_t)i * NV + r] = su[u];
        double x = 0.0;
#pragma unroll
        for (int c = 0; c < NV; ++c) x += dinv[u][c] * __shfl_sync(FULLM, v[u], sub * NV + c);
        if (on[u] && ok) st_relaxed(out + (size_t)i * NV + r, x);
      }
    }
    ok = __all_sync(FULLM, ok);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      m0[u] = m1[u];
      m1[u] = m2[u];
#pragma unroll
      for (int k = 0; k < CH; ++k) jj[u][k] = jn[u][k];
    }
  }
}

// ---- scalar systems: 8 lanes per row (one lane per entry on this side of the diagonal), 4 rows per warp ----------------
// Few rows per warp on purpose: a warp advances when the slowest of its rows has its dependencies, and the sweep moves
// at the pace of the most coupled wavefront.
#ifndef SWEEP_SROWS
#define SWEEP_SROWS 4
#endif
static constexpr int SROWS = SWEEP_SROWS, SLANES = 8;
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
#ifdef SWEEP_S_LANE_POLL
    if (ok && cur.j >= 0) {
      double yj;
      ok = poll(out + cur.j, ctrl, yj);
      if (ok) prod = cur.a * yj;
    }
#else
    {  // one converged loop for the whole warp: every round re-issues the loads of all lanes still waiting
      const bool need = ok && cur.j >= 0;
      bool miss = need;
      double yj = 0.0;
      for (unsigned spins = 0;; ++spins) {
        if (miss) {
          yj = ld_relaxed(out + cur.j);
          miss = (unsigned long long)__double_as_longlong(yj) == SENTINEL;
        }
        if (!__any_sync(FULLM, miss)) break;
        if ((spins & 1023u) == 1023u) {
          if (spins > (1u << 24)) atomicExch(ctrl + 1, 1);
          if (__any_sync(FULLM, ld_relaxed_i(ctrl + 1) != 0)) { ok = false; break; }
        }
      }
      if (need && ok) prod = cur.a * yj;
    }
#endif
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
static void launch_resident_smem(su2b200_ctx* c, K kernel, size_t smem, int maxBlocksUseful, int blocksPerSmCap,
                                 Args... args) {
  int dev = 0, nsm = 0, perSm = 0;
  SU2B_CUDA(cudaGetDevice(&dev));
  SU2B_CUDA(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
  SU2B_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, kernel, 128, smem));
  SU2B_REQUIRE(perSm > 0, "sweep kernel does not fit an SM");
  if (blocksPerSmCap > 0) perSm = std::min(perSm, blocksPerSmCap);
  const int grid = std::max(1, std::min(maxBlocksUseful, nsm * perSm));
  void* argv[] = {(void*)&args...};
  SU2B_CUDA(cudaLaunchCooperativeKernel((const void*)kernel, dim3(grid), dim3(128), argv, smem, c->stream));
  c->launches++;
}
template <class K, class... Args>
static void launch_resident(su2b200_ctx* c, K kernel, int maxBlocksUseful, int blocksPerSmCap, Args... args) {
  launch_resident_smem(c, kernel, 0, maxBlocksUseful, blocksPerSmCap, args...);
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
    // k_poll_sweep_rowsU: ROWS_U groups of 32/nVar rows per warp and iteration, every level padded to that
    const int rpw = ROWS_U * (32 / ls.nVar);
    auto pad = [&](const std::vector<int>& ptr, const std::vector<int>& rows, int& groups) {
      std::vector<int> p;
      p.reserve(rows.size() + (size_t)rpw * ptr.size());
      for (size_t l = 0; l + 1 < ptr.size(); ++l) {
        for (int k = ptr[l]; k < ptr[l + 1]; ++k) p.push_back(rows[k]);
        while (p.size() % rpw) p.push_back(-1);
      }
      groups = (int)(p.size() / rpw);
      return p;
    };
    make(pad(ls.lvl_ptr_lower, ls.lvl_rows_lower.to_host(), ls.poll_groups_lower), true, ls.sweep_metar_lower);
    make(pad(ls.lvl_ptr_upper, ls.lvl_rows_upper.to_host(), ls.poll_groups_upper), false, ls.sweep_metar_upper);
    // hybrid schedule: the contiguous run of wide levels [first, last] goes to the rows kernel, the narrow head and tail
    // to the warp-per-row kernel. split = {rows before the run, groups before the run, rows at its end, groups at its end}
    // (185^3, ms per application: threshold 1024 -> 3.75, 4096 -> 3.53, 16384 -> 3.46; warp-per-row only 3.97)
    const int wide = getenv("SU2B200_SWEEP_WIDE") ? atoi(getenv("SU2B200_SWEEP_WIDE")) : ls.wide_sweep;
    auto split = [&](const std::vector<int>& ptr, int (&sp)[4]) {
      const int nl = (int)ptr.size() - 1;
      int first = -1, last = -1;
      for (int l = 0; l < nl; ++l)
        if (ptr[l + 1] - ptr[l] >= wide) { if (first < 0) first = l; last = l; }
      auto groups_before = [&](int lev) {
        long g = 0;
        for (int l = 0; l < lev; ++l) g += (ptr[l + 1] - ptr[l] + rpw - 1) / rpw;
        return (int)g;
      };
      if (first < 0) { sp[0] = sp[2] = ptr[nl]; sp[1] = sp[3] = groups_before(nl); return; }
      sp[0] = ptr[first]; sp[1] = groups_before(first);
      sp[2] = ptr[last + 1]; sp[3] = groups_before(last + 1);
    };
    split(ls.lvl_ptr_lower, ls.poll_split_lower);
    split(ls.lvl_ptr_upper, ls.poll_split_upper);
  }
}

// x -> y through the scratch vector w; all three distinct (x may alias y only through the copy made by the caller)
template <int NV>
static void ilu_apply_poll_impl(su2b200_ctx* c, LinSys& ls, const double* x, double* y, double* w) {
  const size_t n = (size_t)ls.Pd * NV;
  build_meta(ls);
  // SU2B200_SWEEP_NOFILL=1 (timing experiment only, wrong results): without the sentinel fill every dependency is
  // "there" at the first look (the values of the previous application), i.e. the sweeps run as pure streaming kernels
  static const int nofill = getenv("SU2B200_SWEEP_NOFILL") ? atoi(getenv("SU2B200_SWEEP_NOFILL")) : 0;
  if (!nofill) SU2B_LAUNCH(c, k_fill_sentinel, (int)std::min<size_t>((n + 255) / 256, 148 * 8), 256, n, w, y);
  const int4* ml = reinterpret_cast<const int4*>(ls.sweep_meta_lower.p);
  const int4* mu = reinterpret_cast<const int4*>(ls.sweep_meta_upper.p);
  const int* ci = ls.col_ind;
  const double* ilu = ls.ilu.p;
  // slim ILU (precond_factor.cu): the factorisation left the upper blocks where they were, in the Jacobian
  const double* iluU = (NV > 1 && ls.ilu_slim) ? ls.A.p : ls.ilu.p;
  const double* invM = ls.invM.p;
  int* ctrl = ls.sweep_ctrl.p;
  // resident blocks per SM (0 = what fits): with warp-converged polling more resident warps only help
  static const int capS = getenv("SU2B200_SWEEP_BLOCKS_S") ? atoi(getenv("SU2B200_SWEEP_BLOCKS_S")) : 0;
  static const int capB = getenv("SU2B200_SWEEP_BLOCKS") ? atoi(getenv("SU2B200_SWEEP_BLOCKS")) : 0;
  if constexpr (NV == 1) {
    launch_resident(c, k_poll_sweep_s<true>, (ls.poll_groups_lower + 3) / 4, capS, ls.poll_groups_lower, ml, ctrl, ci, ilu, invM, x, w);
    launch_resident(c, k_poll_sweep_s<false>, (ls.poll_groups_upper + 3) / 4, capS, ls.poll_groups_upper, mu, ctrl, ci, ilu, invM,
                    (const double*)w, y);
  } else {
    // 2: nVar lanes per row (default); 1: warp per row with the shared-memory pipeline; 0: warp per row, register prefetch
    static const int pipe = getenv("SU2B200_SWEEP_PIPE") ? atoi(getenv("SU2B200_SWEEP_PIPE")) : 2;
    if (pipe == 2) {
      // hybrid: narrow levels (bound by the hop from level to level: 0.8 us with a warp per row, 2 us with 32/nVar
      // coupled rows per warp) -> k_poll_sweep; wide levels (bound by instruction issue: 210 vs 30 warp instructions
      // per row) -> k_poll_sweep_rows. Up to three launches per sweep, split at level boundaries.
      const int4* mrl = reinterpret_cast<const int4*>(ls.sweep_metar_lower.p);
      const int4* mru = reinterpret_cast<const int4*>(ls.sweep_metar_upper.p);
      // sp[1], sp[3] count padded units of ROWS_U groups (one consumer iteration of the wide kernel)
      constexpr int UNIT = ROWS_U * (32 / NV);
      static const int capU = getenv("SU2B200_SWEEP_BLOCKS_U") ? atoi(getenv("SU2B200_SWEEP_BLOCKS_U")) : 0;
      static const int useU = getenv("SU2B200_SWEEP_U") ? atoi(getenv("SU2B200_SWEEP_U")) : ROWS_U;
      // fused upper product (linalg.cu): the backward sweep leaves sum_{j>i} U_ij y_j per row when the caller asked for it
      // and the upper factor is the Jacobian's own upper triangle
      double* sout = nullptr;
      ls.upper_prod_of = nullptr;
      if (ls.want_upper_prod && ls.upper_stable && ls.nparts <= 1 && ls.ilu_gen == ls.A_gen) {
        if (ls.upper_prod.n != (size_t)ls.Pd * NV) ls.upper_prod.alloc((size_t)ls.Pd * NV);
        sout = ls.upper_prod.p;
      }
      auto sweep = [&](auto narrowK, auto wideK, auto wideK1, const int4* mnar, const int4* mwid, const int (&sp)[4],
                       const double* ilu, const double* src, double* dst, double* sout) {
        const int nRows = ls.Pd;
        if (sp[0] > 0) launch_resident(c, narrowK, (sp[0] + 3) / 4, capB, sp[0], mnar, ctrl, ci, ilu, invM, src, dst, sout);
        if (sp[3] > sp[1]) {
          if (useU == 1)  // A/B: one group per iteration over the same padded list (every unit is ROWS_U groups)
            launch_resident(c, wideK1, ((sp[3] - sp[1]) * ROWS_U + 3) / 4, capU, (sp[3] - sp[1]) * ROWS_U,
                            mwid + (size_t)sp[1] * UNIT, ctrl, ci, ilu, invM, src, dst, sout);
          else
            launch_resident(c, wideK, (sp[3] - sp[1] + 3) / 4, capU, sp[3] - sp[1], mwid + (size_t)sp[1] * UNIT, ctrl, ci,
                            ilu, invM, src, dst, sout);
        }
        if (nRows > sp[2])
          launch_resident(c, narrowK, (nRows - sp[2] + 3) / 4, capB, nRows - sp[2], mnar + sp[2], ctrl, ci, ilu, invM, src, dst,
                          sout);
      };
      sweep(k_poll_sweep<NV, true>, k_poll_sweep_rowsU<NV, true, ROWS_U>, k_poll_sweep_rowsU<NV, true, 1>, ml, mrl,
            ls.poll_split_lower, ilu, x, w, (double*)nullptr);
      sweep(k_poll_sweep<NV, false>, k_poll_sweep_rowsU<NV, false, ROWS_U>, k_poll_sweep_rowsU<NV, false, 1>, mu, mru,
            ls.poll_split_upper, iluU, (const double*)w, y, sout);
      if (sout) ls.upper_prod_of = y;
    } else if (pipe) {
      const size_t smem = (size_t)4 * PIPE * sizeof(SweepStage<NV>);
      launch_resident_smem(c, k_poll_sweep_pipe<NV, true>, smem, (ls.Pd + 3) / 4, capB, ls.Pd, ml, ctrl, ci, ilu, invM, x, w);
      launch_resident_smem(c, k_poll_sweep_pipe<NV, false>, smem, (ls.Pd + 3) / 4, capB, ls.Pd, mu, ctrl, ci, iluU, invM,
                           (const double*)w, y);
    } else {
      launch_resident(c, k_poll_sweep<NV, true>, (ls.Pd + 3) / 4, capB, ls.Pd, ml, ctrl, ci, ilu, invM, x, w, (double*)nullptr);
      launch_resident(c, k_poll_sweep<NV, false>, (ls.Pd + 3) / 4, capB, ls.Pd, mu, ctrl, ci, iluU, invM, (const double*)w, y,
                      (double*)nullptr);
    }
  }
}

void ilu_apply_poll(su2b200_ctx* c, LinSys& ls, const double* x, double* y) {
  // slim factors keep U in the Jacobian: L, D and invM of one matrix must not be combined with the U of another
  SU2B_REQUIRE(!(ls.ilu_slim && ls.ilu_gen != ls.A_gen),
               "ILU preconditioner is stale: the matrix changed after BuildILUPreconditioner (rebuild before applying)");
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
    case 2: ilu_apply_poll_impl<2>(c, ls, xin, y, ls.sweep_tmp.p); break;
    case 4: ilu_apply_poll_impl<4>(c, ls, xin, y, ls.sweep_tmp.p); break;
    case 5: ilu_apply_poll_impl<5>(c, ls, xin, y, ls.sweep_tmp.p); break;
    default: throw Error("ilu: unsupported nVar");
  }
}

}  // namespace su2b200
