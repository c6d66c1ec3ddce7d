// a17. ILU0 factorisation (CSysMatrix<T>::BuildILUPreconditioner, CSysMatrix.cpp:714-760) as ONE launch.
//
// Row i of the IKJ factorisation only ever writes row i, and needs, from every lower neighbour j, the inverse of the
// factored diagonal (invM_j) and the upper blocks U_jk that meet row i's pattern. The host does the symbolic work once
// per sparsity pattern: for every row the list of updates {source block (j,k), target slot (i,k), lower slot (i,j)}
// in the reference's loop order, and which blocks the factorisation never modifies ("stable": on a hexahedral box all
// off-diagonal upper blocks). The numeric kernel then
//   * takes rows in dependency-level order, round-robin over the warps of a co-resident (cooperative) grid;
//   * streams the row's own blocks and its stable sources from the input matrix A into shared memory with cp.async,
//     one row ahead (descriptors two and three rows ahead), so that none of that is on the dependency chain;
//   * waits for invM_j by polling the data itself (invM is pre-filled with a sentinel NaN payload);
//   * does the 5x5 (4x4) products and the Gauss inversion cooperatively (one scalar per lane, shuffles);
//   * writes the finished row to the ILU matrix and publishes invM_i (after a fence only if some block of the row
//     is an unstable source for a later row).
// A is read once and ILU written once: the A -> ILU copy of the level-scheduled path disappears.
// The arithmetic per block, and its order, are those of ilu_factor_row (precond.cu) == the reference's.
#include "internal.hpp"
#include "kernels.hpp"
#include <algorithm>

namespace su2b200 {

static constexpr unsigned FULLM = 0xffffffffu;
#ifndef POLL_SLEEP_NS
#define POLL_SLEEP_NS 0  // measured on 129^3: any __nanosleep in the poll loop doubles the sweep time (1.69 -> 3.4 ms)
#endif
static constexpr unsigned long long SENTINEL = 0xFFF8C0DEC0DE5A5AULL;
static constexpr int MAXR = 16;  // blocks of a row held in shared memory (longer rows: level-scheduled fallback)
static constexpr int MAXS = 12;  // stable sources staged ahead per row (the others are fetched when needed)
static constexpr unsigned STABLE = 0x80000000u;
#ifndef FACTOR_MIN_BLOCKS
#define FACTOR_MIN_BLOCKS 6
#endif

__device__ __forceinline__ double ldr(const double* p) {
  double v;
  asm volatile("ld.relaxed.gpu.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void str(double* p, double v) {
  asm volatile("st.relaxed.gpu.global.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}
__device__ __forceinline__ int ldr_i(const int* p) {
  int v;
  asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ bool poll_f(const double* p, int* ctrl, double& out) {
  unsigned spins = 0;
  for (;;) {
    const double v = ldr(p);
    if ((unsigned long long)__double_as_longlong(v) != SENTINEL) { out = v; return true; }
#if POLL_SLEEP_NS > 0
    __nanosleep(POLL_SLEEP_NS);  // a waiting warp must not eat the issue slots of the warps that have work
#endif
    if ((++spins & 1023u) == 0) {
      if (spins > (1u << 24)) atomicExch(ctrl + 1, 1);
      if (ldr_i(ctrl + 1) != 0) return false;
    }
  }
}
// Warp-converged form: EVERY lane of the warp calls it; lanes whose v still holds the sentinel re-load it, all lanes
// leave together. (Per-lane spin loops split the warp into single-lane paths that take turns on the load pipe: the
// scalar sweeps measured 2.6 us per level that way and 0.87 us with the converged loop.) false = watchdog, for all lanes.
__device__ __forceinline__ bool poll_warp(const double* p, int* ctrl, double& v) {
  for (unsigned spins = 0;; ++spins) {
    const bool miss = (unsigned long long)__double_as_longlong(v) == SENTINEL;
    if (!__any_sync(FULLM, miss)) return true;
    if (miss) v = ldr(p);
    if ((spins & 1023u) == 1023u) {
      if (spins > (1u << 24)) atomicExch(ctrl + 1, 1);
      if (__any_sync(FULLM, ldr_i(ctrl + 1) != 0)) return false;
    }
  }
}
__device__ __forceinline__ void cp_async8(double* smem, const double* gmem) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void fix_pivot(double& pivot) {
  const double eps = (double)1e-12f;
  if (fabs(pivot) < eps) pivot = copysign(eps, pivot);
}

template <int NV>
__device__ __forceinline__ double wmatmul(double a, double b, int r, int cc) {
  double s = 0.0;
#pragma unroll
  for (int m = 0; m < NV; ++m) s += __shfl_sync(FULLM, a, r * NV + m) * __shfl_sync(FULLM, b, m * NV + cc);
  return s;
}

// Gauss elimination without pivoting + back substitution (CSysMatrix::MatrixInverse, CSysMatrix.cpp:1000-1058), one
// scalar of A and of the inverse per lane; per element the same operations in the same order as the serial loops.
template <int NV>
__device__ __forceinline__ double winverse(double a, int r, int c, int q) {
  double m = (r == c) ? 1.0 : 0.0;
#pragma unroll
  for (int j = 0; j < NV - 1; ++j) {
    double pj = __shfl_sync(FULLM, a, j * NV + j);
    fix_pivot(pj);
    if (q == j * NV + j) a = pj;
    const double arj = __shfl_sync(FULLM, a, r * NV + j);
    const double w = arj / pj;
    const double ajc = __shfl_sync(FULLM, a, j * NV + c);
    const double mjc = __shfl_sync(FULLM, m, j * NV + c);
    if (r > j) {
      if (c >= j) a -= w * ajc;
      if (c <= j) m -= w * mjc;
    }
  }
#pragma unroll
  for (int i = NV - 1; i >= 0; --i) {
#pragma unroll
    for (int j = i + 1; j < NV; ++j) {
      const double aij = __shfl_sync(FULLM, a, i * NV + j);
      const double mjc = __shfl_sync(FULLM, m, j * NV + c);
      if (r == i) m -= aij * mjc;
    }
    double pii = __shfl_sync(FULLM, a, i * NV + i);
    fix_pivot(pii);
    if (r == i) m /= pii;
  }
  return m;
}

// row descriptor: a = {row, first entry, diagonal entry, one past the last entry}
//                 b = {first op, op count, write-back mask (bit s: slot s changes), 1 if a fence must precede invM_i}
struct RowDesc { int4 a, b; };

template <int NV>
__global__ void __launch_bounds__(128, FACTOR_MIN_BLOCKS) k_factor_poll(int n, int slim, int ownCap, int srcCap, const RowDesc* __restrict__ desc,
                                                     const int2* __restrict__ ops,
                                                     int* ctrl, const int* __restrict__ col_ind,
                                                     const double* __restrict__ A, double* ILU, double* invM) {
  constexpr int LPB = NV * NV;
  extern __shared__ double smem[];  // [4 warps][2 buffers][ownCap blocks] then [4][2][srcCap]
  double* sOwn = smem;
  double* sSrc = smem + (size_t)4 * 2 * ownCap * LPB;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int q = lane < LPB ? lane : LPB - 1;
  const bool act = lane < LPB;
  const int r = q / NV, cc = q % NV;
  const int W = gridDim.x * (blockDim.x >> 5);
  int t = blockIdx.x * (blockDim.x >> 5) + wib;
  if (t >= n) return;

  auto load_ops = [&](const RowDesc& d) -> int2 {  // lane u holds op u of the row (rows with more: fetched when used)
    return lane < d.b.y ? ops[d.b.x + lane] : make_int2(0, 0);
  };
  auto load_cols = [&](const RowDesc& d) -> int {  // lane s holds the column of lower entry s
    return d.a.y + lane < d.a.z ? col_ind[d.a.y + lane] : 0;
  };
  auto stage = [&](const RowDesc& d, int2 myop, int buf) {
    const int nOwn = slim ? d.a.z - d.a.y + 1 : d.a.w - d.a.y;  // slim: the upper blocks are neither changed nor copied
    double* own = sOwn + (size_t)(wib * 2 + buf) * ownCap * LPB;
    double* src = sSrc + (size_t)(wib * 2 + buf) * srcCap * LPB;
    if (act)
      for (int s = 0; s < nOwn; ++s) cp_async8(own + s * LPB + q, A + (size_t)(d.a.y + s) * LPB + q);
    const int ns = min(d.b.y, srcCap);
    for (int u = 0; u < ns; ++u) {
      const unsigned sidx = (unsigned)__shfl_sync(FULLM, myop.x, u);
      if (act && (sidx & STABLE)) cp_async8(src + u * LPB + q, A + (size_t)(sidx & ~STABLE) * LPB + q);
    }
    cp_async_commit();
  };

  // pipeline: descriptor of row t+3W, ops/columns of row t+2W, staged data of row t+W, arithmetic of row t
  RowDesc d0 = desc[t];
  RowDesc d1 = t + W < n ? desc[t + W] : d0;
  RowDesc d2 = t + 2 * W < n ? desc[t + 2 * W] : d1;
  int2 op0 = load_ops(d0), op1 = load_ops(d1);
  int col0 = load_cols(d0), col1 = load_cols(d1);
  int buf = 0;
  stage(d0, op0, buf);
  for (; t < n; t += W) {
    const RowDesc d3 = t + 3 * W < n ? desc[t + 3 * W] : d2;
    int2 op2 = make_int2(0, 0);
    int col2 = 0;
    if (t + 2 * W < n) { op2 = load_ops(d2); col2 = load_cols(d2); }
    if (t + W < n) stage(d1, op1, buf ^ 1); else cp_async_commit();
    cp_async_wait<1>();
    __syncwarp();

    double* own = sOwn + (size_t)(wib * 2 + buf) * ownCap * LPB;
    const double* src = sSrc + (size_t)(wib * 2 + buf) * srcCap * LPB;
    const int i = d0.a.x, rb = d0.a.y, di = d0.a.z, re = d0.a.w;
    const int nl = di - rb, nOps = d0.b.y;
    int u = 0;
    bool ok = true;
    // first look at the inverse diagonals of the first lower neighbours, all in flight together
    constexpr int LOOK = 4;
    double look[LOOK];
#pragma unroll
    for (int s = 0; s < LOOK; ++s) {
      const int j = __shfl_sync(FULLM, col0, s);
      look[s] = s < nl ? ldr(invM + (size_t)j * LPB + q) : 0.0;
    }
    for (int s = 0; s < nl; ++s) {
      const int j = s < 32 ? __shfl_sync(FULLM, col0, s) : col_ind[rb + s];
      double inv = __longlong_as_double((long long)SENTINEL);
#pragma unroll
      for (int k = 0; k < LOOK; ++k)
        if (s == k) inv = look[k];
      ok = poll_warp(invM + (size_t)j * LPB + q, ctrl, inv);
      if (!ok) return;  // watchdog
      const double aij = own[s * LPB + q];
      const double w = wmatmul<NV>(aij, inv, r, cc);  // A_ij * inv(U_jj)
      bool fenced = false;
      for (;;) {
        if (u >= nOps) break;
        int2 op = u < 32 ? make_int2(__shfl_sync(FULLM, op0.x, u), __shfl_sync(FULLM, op0.y, u)) : ops[d0.b.x + u];
        if ((op.y >> 16) != s) break;
        const unsigned sidx = (unsigned)op.x;
        double ajk;
        if ((sidx & STABLE) && u < srcCap) ajk = src[u * LPB + q];
        else if (sidx & STABLE) ajk = A[(size_t)(sidx & ~STABLE) * LPB + q];
        else {  // block (j,k) was modified by row j: ordered after the invM_j we have just seen
          if (!fenced) { __threadfence(); fenced = true; }
          ajk = ldr(ILU + (size_t)sidx * LPB + q);
        }
        const double aux = wmatmul<NV>(w, ajk, r, cc);
        if (act) own[(op.y & 0xffff) * LPB + q] -= aux;
        ++u;
      }
      if (act) own[s * LPB + q] = w;
      __syncwarp();
    }
    const double minv = winverse<NV>(own[nl * LPB + q], r, cc, q);
    if (act) {
      // every block goes to the ILU matrix (it is not a copy of A any more); slim: only the lower ones and the
      // diagonal, the upper blocks are the Jacobian's own (ls.upper_stable)
      const int nw = slim ? nl + 1 : re - rb;
      for (int s = 0; s < nw; ++s) ILU[(size_t)(rb + s) * LPB + q] = own[s * LPB + q];
      if (d0.b.w) __threadfence();
      str(invM + (size_t)i * LPB + q, minv);
    }
    __syncwarp();
    d0 = d1; d1 = d2; d2 = d3;
    op0 = op1; op1 = op2;
    col0 = col1; col1 = col2;
    buf ^= 1;
  }
}

// ---- scalar systems (SA): SROWS_F rows per warp, SGRP lanes per row all doing the same arithmetic (lane 0 stores) -------
// The block kernel with nVar = 1 keeps one lane of 32 busy and is bound by its own instruction stream; here a row is a
// handful of registers: own[8] and up to four stable sources are prefetched one row ahead, updates select their target
// with predicates (no dynamic register index). Rows that do not fit (more than 8 entries, 4 lower neighbours or 4
// updates, or an unstable source) take the in-memory path below, same arithmetic.
static constexpr int SGRP = 4, SROWS_F = 32 / SGRP, SOWN = 8, SLOW = 4;
struct ScalarRow {
  double own[SOWN], src[SLOW];
  int col[SLOW];
  int2 op[SLOW];
};
__device__ __forceinline__ void fetch_scalar_row(ScalarRow& r, const RowDesc& d, const int2* __restrict__ ops,
                                                 const int* __restrict__ col_ind, const double* __restrict__ A) {
  const bool ok = d.a.x >= 0 && d.b.z != 0;
#pragma unroll
  for (int s = 0; s < SOWN; ++s) r.own[s] = (ok && d.a.y + s < d.a.w) ? A[d.a.y + s] : 0.0;
#pragma unroll
  for (int s = 0; s < SLOW; ++s) {
    r.col[s] = (ok && d.a.y + s < d.a.z) ? col_ind[d.a.y + s] : 0;
    r.op[s] = (ok && s < d.b.y) ? ops[d.b.x + s] : make_int2(0, -1);
  }
#pragma unroll
  for (int s = 0; s < SLOW; ++s) r.src[s] = (ok && s < d.b.y) ? A[(unsigned)r.op[s].x & ~STABLE] : 0.0;
}

__global__ void __launch_bounds__(128) k_factor_poll_s(int nSlots, const RowDesc* __restrict__ desc,
                                                       const int2* __restrict__ ops, int* ctrl,
                                                       const int* __restrict__ col_ind, const double* __restrict__ A,
                                                       double* ILU, double* invM) {
  const int lane = threadIdx.x & 31, sub = lane / SGRP;
  const bool writer = (lane % SGRP) == 0;
  const int W = gridDim.x * (blockDim.x >> 5) * SROWS_F;
  int t = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * SROWS_F + sub;
  if (t >= nSlots) return;
  RowDesc d0 = desc[t];
  RowDesc d1 = t + W < nSlots ? desc[t + W] : d0;
  RowDesc d2 = t + 2 * W < nSlots ? desc[t + 2 * W] : d1;
  ScalarRow cur, nxt;
  fetch_scalar_row(cur, d0, ops, col_ind, A);
  nxt = cur;
  bool ok = true;
  for (; t < nSlots; t += W) {
    const RowDesc d3 = t + 3 * W < nSlots ? desc[t + 3 * W] : d2;
    if (t + W < nSlots) fetch_scalar_row(nxt, d1, ops, col_ind, A);
    const int i = d0.a.x, rb = d0.a.y, di = d0.a.z, re = d0.a.w;
    const int nl = di - rb;
    // register-path rows: the inverse diagonals of all lower neighbours, waited for by the whole warp together
    double look[SLOW];
    {
      const bool fast = i >= 0 && ok && d0.b.z != 0;
#pragma unroll
      for (int s = 0; s < SLOW; ++s) look[s] = (fast && s < nl) ? ldr(invM + cur.col[s]) : 0.0;
#pragma unroll
      for (int s = 0; s < SLOW; ++s) ok = poll_warp(invM + cur.col[s], ctrl, look[s]) && ok;
    }
    if (i >= 0 && ok) {
      if (d0.b.z != 0) {
        // ---- register path ----
#pragma unroll
        for (int s = 0; s < SLOW; ++s) {
          if (s < nl && ok) {
            const double inv = look[s];
            {
              const double w = cur.own[s] * inv;
#pragma unroll
              for (int u = 0; u < SLOW; ++u) {
                if (u < d0.b.y && (cur.op[u].y >> 16) == s) {
                  const int pos = cur.op[u].y & 0xffff;
                  const double aux = w * cur.src[u];
#pragma unroll
                  for (int p = 0; p < SOWN; ++p)
                    if (p == pos) cur.own[p] -= aux;
                }
              }
              cur.own[s] = w;
            }
          }
        }
        if (ok) {
          double dgn = 0.0;
#pragma unroll
          for (int p = 0; p < SOWN; ++p)
            if (p == nl) dgn = cur.own[p];
          fix_pivot(dgn);
          const double minv = 1.0 / dgn;
          if (writer) {
#pragma unroll
            for (int p = 0; p < SOWN; ++p)
              if (rb + p < re) ILU[rb + p] = cur.own[p];
            if (d0.b.w) __threadfence();  // a modified upper entry of this row is read by a later row
            str(invM + i, minv);
          }
        }
      } else {
        // ---- in-memory path: the row is copied to the ILU array and updated there (writer lane only) ----
        if (writer) {
          for (int s = rb; s < re; ++s) ILU[s] = A[s];
          int u = d0.b.x;
          const int uEnd = d0.b.x + d0.b.y;
          for (int s = 0; s < nl && ok; ++s) {
            double inv;
            ok = poll_f(invM + col_ind[rb + s], ctrl, inv);
            if (!ok) break;
            __threadfence();
            const double w = ILU[rb + s] * inv;
            while (u < uEnd && (ops[u].y >> 16) == s) {
              const int2 op = ops[u];
              const unsigned sidx = (unsigned)op.x;
              const double ajk = (sidx & STABLE) ? A[sidx & ~STABLE] : ldr(ILU + sidx);
              ILU[rb + (op.y & 0xffff)] -= w * ajk;
              ++u;
            }
            ILU[rb + s] = w;
          }
          if (ok) {
            double dgn = ILU[di];
            fix_pivot(dgn);
            __threadfence();
            str(invM + i, 1.0 / dgn);
          }
        }
      }
    }
    ok = __all_sync(FULLM, ok);  // after the watchdog fired every lane idles through its remaining rows
    d0 = d1; d1 = d2; d2 = d3;
    cur = nxt;
  }
}

// ---- block systems, nVar lanes per matrix row ---------------------------------------------------------------------------
// k_factor_poll spends ~1500 warp instructions per row, most of them shuffles and index arithmetic of the
// one-scalar-per-lane products (ncu: issue-bound, 6 % of its instructions are DFMA). Here lane r of a group of nVar
// lanes owns block row r of every block of its matrix row; the blocks live in shared-memory tiles (own row, staged
// stable sources, the current inverse diagonal), products read their second operand from the tile (no shuffles), the
// Gauss inversion exchanges pivot rows inside the group. 32/nVar matrix rows per warp, ~100 warp instructions per row.
// Same arithmetic, element by element, as CSysMatrix.cpp:714-760 / MatrixInverse :548-601.
static constexpr int RF_OWN = 8;   // most blocks of a row (longer rows: k_factor_poll handles the matrix); tiles follow the matrix
static constexpr int RF_SRC = 4;   // stable sources staged with the row (others are read when needed)
static constexpr int RF_LOW = 4;   // lower neighbours whose inverse diagonal is looked at ahead

__device__ __forceinline__ bool is_sentinel(double v) {
  // every double with this high word is a NaN that arithmetic cannot produce: the high word identifies the sentinel
  return (unsigned)__double2hiint(v) == (unsigned)(SENTINEL >> 32);
}

// Inverse of a block by Gauss elimination (CSysMatrix::MatrixInverse), nVar lanes per block: lane r holds ROW r of the block
// being eliminated (ar) and builds COLUMN r of the inverse (mc): the elimination weights (one division per lane and
// pivot) and the pivot rows travel by shuffle, every update of the inverse is then local to the lane that owns the
// column, and the closing divisions run on all columns at once.  All 32 lanes call it (gbase = first lane of the group).
template <int NV>
__device__ __forceinline__ void invert_block_rows(double (&ar)[NV], double (&mc)[NV], int gbase, int r) {
#pragma unroll
  for (int c = 0; c < NV; ++c) mc[c] = (r == c) ? 1.0 : 0.0;
#pragma unroll
  for (int jp = 0; jp < NV - 1; ++jp) {
    double prow[NV];
#pragma unroll
    for (int c = jp; c < NV; ++c) prow[c] = __shfl_sync(FULLM, ar[c], gbase + jp);
    double pj = prow[jp];
    fix_pivot(pj);
    if (r == jp) ar[jp] = pj;
    double wgt = 0.0;
    if (r > jp) {
      wgt = ar[jp] / pj;
#pragma unroll
      for (int c = jp; c < NV; ++c) ar[c] -= wgt * (c == jp ? pj : prow[c]);
    }
#pragma unroll
    for (int m = jp + 1; m < NV; ++m) {
      const double wm = __shfl_sync(FULLM, wgt, gbase + m);
      if (r <= jp) mc[m] -= wm * mc[jp];
    }
  }
#pragma unroll
  for (int ip = NV - 1; ip >= 0; --ip) {
#pragma unroll
    for (int jp = ip + 1; jp < NV; ++jp) {
      const double a = __shfl_sync(FULLM, ar[jp], gbase + ip);
      mc[ip] -= a * mc[jp];
    }
    double pii = __shfl_sync(FULLM, ar[ip], gbase + ip);
    fix_pivot(pii);
    mc[ip] /= pii;
  }
}

#ifndef FACTOR_ROWS_MIN_BLOCKS
#define FACTOR_ROWS_MIN_BLOCKS 3  // resident blocks per SM the register allocation aims at
#endif
template <int NV>
__global__ void __launch_bounds__(128, FACTOR_ROWS_MIN_BLOCKS) k_factor_rows(int nGroups, int slim, int dyn, int ownCap, int srcCap, const RowDesc* __restrict__ desc,
                                                     const int2* __restrict__ ops, int* ctrl,
                                                     const int* __restrict__ col_ind, const double* __restrict__ A,
                                                     double* ILU, double* invM) {
  constexpr int LPB = NV * NV;
  constexpr int RPW = 32 / NV;
  const int TILES = ownCap + srcCap + 1;  // own row, staged stable sources, the inverse diagonal in use
  extern __shared__ double smem[];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int sub = lane / NV, r = lane % NV;
  const bool live = sub < RPW;
  const unsigned gmask = live ? (((1u << NV) - 1u) << (sub * NV))
                              : (RPW * NV < 32 ? (0xffffffffu << ((RPW * NV) & 31)) : 0u);
  double* T = smem + (size_t)((wib * RPW + (live ? sub : 0)) * TILES) * LPB;
  double* invT = T + (size_t)(ownCap + srcCap) * LPB;
  const int W = gridDim.x * (blockDim.x >> 5);
  bool ok = true;
  // descriptors run two groups ahead so that the rows of the next group can be pulled into L2 while this one works
  RowDesc d1, d2;
  d1.a = d2.a = make_int4(-1, 0, 0, 0);
  d1.b = d2.b = make_int4(0, 0, 0, 0);
  // Group order: static (warp w takes w, w + W, ...) or, dyn != 0, by ticket (ctrl[0], zeroed by the host): warps differ in
  // pace (position of their SM relative to the L2 slices and the second die), and with the static order the others end
  // up polling for the rows of the slowest (6.59 -> 6.39 ms on 185^3).  Tickets are drawn three groups ahead
  // (descriptors two ahead); a warp works through its tickets in increasing order, so the lowest unfinished group is
  // always being worked on: no deadlock.
  auto draw = [&]() {
    int t = 0;
    if (lane == 0) t = atomicAdd(ctrl, 1);
    return __shfl_sync(FULLM, t, 0);
  };
  const int g0 = blockIdx.x * (blockDim.x >> 5) + wib;
  int t0 = g0, t1 = g0 + W, t2 = g0 + 2 * W, t3 = g0 + 3 * W;
  if (dyn) { t0 = draw(); t1 = draw(); t2 = draw(); t3 = draw(); }
  if (live && t0 < nGroups) d1 = desc[(size_t)t0 * RPW + sub];
  if (live && t1 < nGroups) d2 = desc[(size_t)t1 * RPW + sub];
  for (int g = t0; g < nGroups; g = t0) {
    const RowDesc d = d1;
    d1 = d2;
    d2.a = make_int4(-1, 0, 0, 0);
    d2.b = make_int4(0, 0, 0, 0);
    if (live && t2 < nGroups) d2 = desc[(size_t)t2 * RPW + sub];
    t0 = t1; t1 = t2; t2 = t3;
    t3 = dyn ? draw() : t3 + W;
    if (d1.a.x >= 0) {
      const char* nb = reinterpret_cast<const char*>(A + (size_t)d1.a.y * LPB);
      const int bytes = (slim ? d1.a.z - d1.a.y + 1 : d1.a.w - d1.a.y) * LPB * (int)sizeof(double);
      for (int off = r * 128; off < bytes; off += NV * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(nb + off));
      if (r == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(col_ind + d1.a.y));
      if (r == 1 && d1.b.y > 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(ops + d1.b.x));
    }
    const int i = d.a.x, rb = d.a.y, di = d.a.z, re = d.a.w;
    const bool on = ok && i >= 0;
    const int nOwn = on ? re - rb : 0, nl = on ? di - rb : 0, nOps = on ? d.b.y : 0;
    // ---- regular groups: straight-line code ------------------------------------------------------------------------
    // Every row of the warp has at most RG lower neighbours, exactly one update per lower neighbour, landing on the
    // diagonal block, from a stable staged source (any mesh without triangles in its graph: hexahedra, quadrilaterals;
    // flagged per row by factor_symbolic).  D_i = A_ii - sum_k (A_ik invD_k) A_ki then needs no update lists, no slot
    // searches and no loops of data-dependent length: ~2800 warp instructions per group in the general code below, of
    // which 260 are DFMA (ncu, profiles/r02_ncu_factor_rows.txt).  Same operations in the same order.
    if (slim && __all_sync(FULLM, !live || (d.b.w & 2))) {
      constexpr int RG = 3;
#pragma unroll
      for (int s = 0; s <= RG; ++s)
        if (on && s <= nl) {
#pragma unroll
          for (int c = 0; c < NV; ++c) cp_async8(T + s * LPB + r * NV + c, A + (size_t)(rb + s) * LPB + r * NV + c);
        }
      int jc[RG];
      unsigned src[RG];
#pragma unroll
      for (int k = 0; k < RG; ++k) {
        jc[k] = k < nl ? col_ind[rb + k] : 0;
        src[k] = k < nl ? ((unsigned)ops[d.b.x + k].x & ~STABLE) : 0u;
      }
      double inv[RG][NV];
#pragma unroll
      for (int k = 0; k < RG; ++k)
#pragma unroll
        for (int c = 0; c < NV; ++c) inv[k][c] = k < nl ? ldr(invM + (size_t)jc[k] * LPB + r * NV + c) : 0.0;
#pragma unroll
      for (int k = 0; k < RG; ++k)
        if (k < nl) {
#pragma unroll
          for (int c = 0; c < NV; ++c)
            cp_async8(T + (ownCap + k) * LPB + r * NV + c, A + (size_t)src[k] * LPB + r * NV + c);
        }
      cp_async_commit();
      // all awaited inverse diagonals at once: one element per row and neighbour is watched, the rest fetched on arrival
      for (unsigned spins = 0;; ++spins) {
        bool miss = false;
#pragma unroll
        for (int k = 0; k < RG; ++k)
#pragma unroll
          for (int c = 0; c < NV; ++c) miss = miss || is_sentinel(inv[k][c]);
        if (!__any_sync(FULLM, miss)) break;
#pragma unroll
        for (int k = 0; k < RG; ++k) {
          bool mk = false;
#pragma unroll
          for (int c = 0; c < NV; ++c) mk = mk || is_sentinel(inv[k][c]);
          if (mk && !is_sentinel(ldr(invM + (size_t)jc[k] * LPB + (LPB - 1)))) {
#pragma unroll
            for (int c = 0; c < NV; ++c)
              if (is_sentinel(inv[k][c])) inv[k][c] = ldr(invM + (size_t)jc[k] * LPB + r * NV + c);
          }
        }
        if ((spins & 1023u) == 1023u) {
          if (spins > (1u << 22)) atomicExch(ctrl + 1, 1);
          if (__any_sync(FULLM, ldr_i(ctrl + 1) != 0)) { ok = false; break; }
        }
      }
      // the inverses go to the tiles behind the diagonal (free in the slim form: the upper blocks are not staged)
#pragma unroll
      for (int k = 0; k < RG; ++k)
        if (k < nl) {
#pragma unroll
          for (int c = 0; c < NV; ++c) T[(nl + 1 + k) * LPB + r * NV + c] = inv[k][c];
        }
      cp_async_wait<0>();
      __syncwarp();
      const bool go = on && ok;
      double dg[NV], mc[NV];
#pragma unroll
      for (int c = 0; c < NV; ++c) dg[c] = go ? T[nl * LPB + r * NV + c] : (r == c ? 1.0 : 0.0);
#pragma unroll
      for (int k = 0; k < RG; ++k)
        if (go && k < nl) {
          double a[NV], w[NV];
#pragma unroll
          for (int m = 0; m < NV; ++m) a[m] = T[k * LPB + r * NV + m];
          const double* iv = T + (nl + 1 + k) * LPB;
#pragma unroll
          for (int c = 0; c < NV; ++c) {
            double acc = 0.0;
#pragma unroll
            for (int m = 0; m < NV; ++m) acc += a[m] * iv[m * NV + c];
            w[c] = acc;
          }
#pragma unroll
          for (int c = 0; c < NV; ++c) ILU[(size_t)(rb + k) * LPB + r * NV + c] = w[c];
          const double* sp = T + (ownCap + k) * LPB;
#pragma unroll
          for (int c = 0; c < NV; ++c) {
            double acc = 0.0;
#pragma unroll
            for (int m = 0; m < NV; ++m) acc += w[m] * sp[m * NV + c];
            dg[c] -= acc;
          }
        }
      if (go) {
#pragma unroll
        for (int c = 0; c < NV; ++c) ILU[(size_t)(rb + nl) * LPB + r * NV + c] = dg[c];
      }
      invert_block_rows<NV>(dg, mc, sub * NV, r);
      if (go) {
#pragma unroll
        for (int m = 0; m < NV; ++m) str(invM + (size_t)i * LPB + m * NV + r, mc[m]);
      }
      __syncwarp();
      ok = __all_sync(FULLM, ok);
      continue;
    }
    // the row itself (asynchronous copies into its tiles), the first update ops and the columns of the lower entries
    // in one round of loads, then the first look at the awaited inverse diagonals (row r of each) and the staged
    // sources: three dependent rounds of memory latency per group instead of one per array
    const int nLoad = slim ? min(nl + 1, nOwn) : nOwn;  // slim: the upper blocks are neither changed nor copied
    for (int s = 0; s < nLoad; ++s)
#pragma unroll
      for (int c = 0; c < NV; ++c) cp_async8(T + s * LPB + r * NV + c, A + (size_t)(rb + s) * LPB + r * NV + c);
    int2 opr[RF_SRC];
#pragma unroll
    for (int k = 0; k < RF_SRC; ++k) opr[k] = k < nOps ? ops[d.b.x + k] : make_int2(0, 0);
    int col[RF_LOW];
#pragma unroll
    for (int s = 0; s < RF_LOW; ++s) col[s] = s < nl ? col_ind[rb + s] : 0;
    double look[RF_LOW][NV];
#pragma unroll
    for (int s = 0; s < RF_LOW; ++s) {
#pragma unroll
      for (int c = 0; c < NV; ++c) look[s][c] = s < nl ? ldr(invM + (size_t)col[s] * LPB + r * NV + c) : 0.0;
    }
#pragma unroll
    for (int k = 0; k < RF_SRC; ++k) {
      const unsigned sidx = (unsigned)opr[k].x;
      if (k < min(nOps, srcCap) && (sidx & STABLE)) {
#pragma unroll
        for (int c = 0; c < NV; ++c)
          cp_async8(T + (ownCap + k) * LPB + r * NV + c, A + (size_t)(sidx & ~STABLE) * LPB + r * NV + c);
      }
    }
    cp_async_commit();
    cp_async_wait<0>();
    __syncwarp();
    int u = 0;
    // the matrix rows of a warp walk their lower entries in lockstep (one converged instruction stream: divergent
    // per-row poll loops share the warp's scoreboard and serialise, measured 40 us per level)
    const int nlMax = __reduce_max_sync(FULLM, nl);
    for (int s = 0; s < nlMax; ++s) {
      const bool act = s < nl;
      // inverse diagonal of neighbour j: row r of it, re-polled until complete
      double inv[NV];
      int jcol = (act && s >= RF_LOW) ? col_ind[rb + s] : 0;
#pragma unroll
      for (int c = 0; c < NV; ++c) inv[c] = act ? __longlong_as_double((long long)SENTINEL) : 0.0;
#pragma unroll
      for (int k = 0; k < RF_LOW; ++k)
        if (s == k) {
          jcol = col[k];
#pragma unroll
          for (int c = 0; c < NV; ++c) inv[c] = act ? look[k][c] : 0.0;
        }
      const double* pin = invM + (size_t)jcol * LPB + r * NV;
      // the lanes of a row watch ONE element of the awaited block (a single sector per row and round: 1776 warps
      // re-reading whole blocks saturate the L2 and slow the rows that everybody waits for), then fetch their own
      const double* plast = invM + (size_t)jcol * LPB + (LPB - 1);
      for (unsigned spins = 0;; ++spins) {
        bool miss = false;
#pragma unroll
        for (int c = 0; c < NV; ++c) miss = miss || is_sentinel(inv[c]);
        if (!__any_sync(FULLM, miss)) break;
        if (miss && !is_sentinel(ldr(plast))) {
#pragma unroll
          for (int c = 0; c < NV; ++c)
            if (is_sentinel(inv[c])) inv[c] = ldr(pin + c);
        }
        if ((spins & 1023u) == 1023u) {
          if (spins > (1u << 22)) atomicExch(ctrl + 1, 1);
          if (__any_sync(FULLM, ldr_i(ctrl + 1) != 0)) { ok = false; break; }
        }
      }
      if (!ok) break;
      if (act) {
#pragma unroll
        for (int c = 0; c < NV; ++c) invT[r * NV + c] = inv[c];
      }
      __syncwarp();
      // w = A_ij * inv(U_jj): row r
      double w[NV];
#pragma unroll
      for (int c = 0; c < NV; ++c) w[c] = 0.0;
      if (act) {
        double a[NV];
#pragma unroll
        for (int m = 0; m < NV; ++m) a[m] = T[s * LPB + r * NV + m];
#pragma unroll
        for (int c = 0; c < NV; ++c) {
          double acc = 0.0;
#pragma unroll
          for (int m = 0; m < NV; ++m) acc += a[m] * invT[m * NV + c];
          w[c] = acc;
        }
      }
      // A_ik -= w * U_jk for the entries k of row j that meet the pattern of row i
      bool fenced = false;
      for (;;) {
        bool has = act && u < nOps;
        int2 op = make_int2(0, 0);
        if (has) {
          if (u >= RF_SRC) op = ops[d.b.x + u];
#pragma unroll
          for (int k = 0; k < RF_SRC; ++k)
            if (u == k) op = opr[k];
          has = (op.y >> 16) == s;
        }
        if (!__any_sync(FULLM, has)) break;
        if (has) {
          const unsigned sidx = (unsigned)op.x;
          const double* sp;
          if ((sidx & STABLE) && u < srcCap) sp = T + (size_t)(ownCap + u) * LPB;
          else if (sidx & STABLE) sp = A + (size_t)(sidx & ~STABLE) * LPB;
          else {
            if (!fenced) { __threadfence(); fenced = true; }
            sp = ILU + (size_t)sidx * LPB;
          }
          const bool from_ilu = !(sidx & STABLE);
          double* tgt = T + (size_t)(op.y & 0xffff) * LPB + r * NV;
#pragma unroll
          for (int c = 0; c < NV; ++c) {
            double acc = 0.0;
#pragma unroll
            for (int m = 0; m < NV; ++m) acc += w[m] * (from_ilu ? ldr(sp + m * NV + c) : sp[m * NV + c]);
            tgt[c] -= acc;
          }
          ++u;
        }
      }
      if (act) {
#pragma unroll
        for (int c = 0; c < NV; ++c) T[s * LPB + r * NV + c] = w[c];
      }
      __syncwarp();  // invT is rewritten by the next lower entry
    }
    __syncwarp();
    double ar[NV], mc[NV];
#pragma unroll
    for (int c = 0; c < NV; ++c) ar[c] = (on && ok) ? T[nl * LPB + r * NV + c] : (r == c ? 1.0 : 0.0);
    invert_block_rows<NV>(ar, mc, sub * NV, r);
    if (on && ok) {
      const int nw = slim ? nLoad : nOwn;
      for (int s = 0; s < nw; ++s)
#pragma unroll
        for (int c = 0; c < NV; ++c) ILU[(size_t)(rb + s) * LPB + r * NV + c] = T[s * LPB + r * NV + c];
      if (d.b.w & 1) __threadfence();
#pragma unroll
      for (int m = 0; m < NV; ++m) str(invM + (size_t)i * LPB + m * NV + r, mc[m]);
    }
    __syncwarp();
    ok = __all_sync(FULLM, ok);
  }
}

__global__ void k_fill_sentinel_f(size_t n, double* __restrict__ a) {
  const double s = __longlong_as_double((long long)SENTINEL);
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) a[i] = s;
}

// ---- host: symbolic factorisation -----------------------------------------------------------------------------------
// returns false if a row does not fit the kernel's shared-memory slots (caller falls back to the level schedule)
bool factor_symbolic(LinSys& ls) {
  if (ls.factor_ready) return ls.factor_ok;
  ls.factor_ready = true;
  ls.factor_ok = false;
  const int n = ls.Pd;
  const auto& rp = ls.h_row_ptr;
  const auto& ci = ls.h_col_ind;
  const auto& dp = ls.h_dia_ptr;
  for (int i = 0; i < n; ++i)
    if (rp[i + 1] - rp[i] > MAXR || dp[i] - rp[i] > 32) return false;
  // pass 1: update list per row, in the order of the reference's loops; mark the blocks that get modified
  std::vector<long> opPtr(n + 1, 0);
  std::vector<int2> opList;
  opList.reserve((size_t)rp[n] / 2);
  std::vector<unsigned char> modified(rp[n], 0);
  for (int i = 0; i < n; ++i) {
    const int rb = rp[i], re = rp[i + 1];
    for (int idx = rb; idx < dp[i]; ++idx) {
      const int j = ci[idx];
      int s = rb;  // both rows are sorted by column: merge
      for (int idx2 = dp[j] + 1; idx2 < rp[j + 1]; ++idx2) {
        const int k = ci[idx2];
        if (k >= n) break;
        while (s < re && ci[s] < k) ++s;
        if (s < re && ci[s] == k) {
          opList.push_back(make_int2(idx2, ((idx - rb) << 16) | (s - rb)));
          modified[s] = 1;
        }
      }
    }
    opPtr[i + 1] = (long)opList.size();
  }
  if (opList.size() >= (size_t)0x7fffffff) return false;
  // pass 2: stable sources, write-back masks, fence flags; descriptors in level order
  for (auto& op : opList)
    if (!modified[op.x]) op.x = (int)((unsigned)op.x | STABLE);
  ls.upper_stable = true;
  for (int i = 0; i < n && ls.upper_stable; ++i)
    for (int s = dp[i] + 1; s < rp[i + 1]; ++s)
      if (modified[s]) { ls.upper_stable = false; break; }
  const std::vector<int> rows = ls.lvl_rows_lower.to_host();
  std::vector<RowDesc> desc(n);
  int maxOwn = 1, maxOps = 0;
  std::vector<int2> opsLevel;
  opsLevel.reserve(opList.size());
  for (int t = 0; t < n; ++t) {
    const int i = rows[t];
    RowDesc d;
    d.a = make_int4(i, rp[i], dp[i], rp[i + 1]);
    unsigned wb = 0;
    int fence = 0;
    for (int s = rp[i]; s < rp[i + 1]; ++s) {
      const bool changes = s <= dp[i] || modified[s];
      if (changes) wb |= 1u << (s - rp[i]);
      if (s > dp[i] && modified[s]) fence = 1;  // an upper block that a later row reads from the ILU matrix
    }
    d.b = make_int4((int)opsLevel.size(), (int)(opPtr[i + 1] - opPtr[i]), (int)wb, fence);
    for (long u = opPtr[i]; u < opPtr[i + 1]; ++u) opsLevel.push_back(opList[u]);
    maxOwn = std::max(maxOwn, rp[i + 1] - rp[i]);
    maxOps = std::max(maxOps, (int)(opPtr[i + 1] - opPtr[i]));
    desc[t] = d;
  }
  static_assert(sizeof(RowDesc) == 32, "RowDesc layout");
  if (ls.nVar == 1) {
    // scalar kernel: every level padded to SROWS_F slots (rows of a warp never depend on one another), b.z = 1 for
    // rows that fit the register path
    std::vector<RowDesc> padded;
    padded.reserve(desc.size() + (size_t)SROWS_F * ls.lvl_ptr_lower.size());
    RowDesc empty;
    empty.a = make_int4(-1, 0, 0, 0);
    empty.b = make_int4(0, 0, 0, 0);
    for (size_t l = 0; l + 1 < ls.lvl_ptr_lower.size(); ++l) {
      for (int t = ls.lvl_ptr_lower[l]; t < ls.lvl_ptr_lower[l + 1]; ++t) {
        RowDesc d = desc[t];
        bool fast = d.a.w - d.a.y <= SOWN && d.a.z - d.a.y <= SLOW && d.b.y <= SLOW;
        for (int u = 0; u < d.b.y && fast; ++u) fast = ((unsigned)opsLevel[d.b.x + u].x & STABLE) != 0;
        d.b.z = fast ? 1 : 0;
        padded.push_back(d);
      }
      while (padded.size() % SROWS_F) padded.push_back(empty);
    }
    desc.swap(padded);
    ls.factor_slots = (int)desc.size();
  } else {
    // k_factor_rows: 32/nVar matrix rows per warp, levels padded to that; usable if every row fits RF_OWN tiles
    const int rpw = 32 / ls.nVar;
    bool fits = true;
    for (const auto& d : desc) fits = fits && (d.a.w - d.a.y <= RF_OWN);
    ls.factor_rows_ok = fits;
    if (fits) {
      std::vector<RowDesc> padded;
      padded.reserve(desc.size() + (size_t)rpw * ls.lvl_ptr_lower.size());
      RowDesc empty;
      empty.a = make_int4(-1, 0, 0, 0);
      empty.b = make_int4(0, 0, 0, 0);
      // wide middle (first to last level of at least `wide` rows): k_factor_rows; the narrow head and tail, where the
      // latency of a level counts and not the rows in it, stay with the warp-per-row kernel
      const int wide = getenv("SU2B200_FACTOR_WIDE") ? atoi(getenv("SU2B200_FACTOR_WIDE")) : ls.wide_factor;
      int first = -1, last = -1;
      const int nLev = (int)ls.lvl_ptr_lower.size() - 1;
      for (int l = 0; l < nLev; ++l)
        if (ls.lvl_ptr_lower[l + 1] - ls.lvl_ptr_lower[l] >= wide) {
          if (first < 0) first = l;
          last = l;
        }
      ls.factor_split[0] = ls.factor_split[1] = n;
      ls.factor_split[2] = ls.factor_split[3] = 0;
      for (int l = 0; l < nLev; ++l) {
        if (l == first) {
          ls.factor_split[0] = ls.lvl_ptr_lower[l];
          ls.factor_split[2] = (int)(padded.size() / rpw);
        }
        for (int t = ls.lvl_ptr_lower[l]; t < ls.lvl_ptr_lower[l + 1]; ++t) {
          // b.w bit 1: the row qualifies for the straight-line path of k_factor_rows (see there)
          RowDesc d = desc[t];
          const int nl = d.a.z - d.a.y;
          bool reg = nl <= 3 && d.b.y == nl && d.b.w == 0 && 2 * nl + 1 <= maxOwn;
          for (int u = 0; u < d.b.y && reg; ++u) {
            const int2 op = opsLevel[d.b.x + u];
            reg = ((unsigned)op.x & STABLE) && (op.y >> 16) == u && (op.y & 0xffff) == nl;
          }
          if (reg && getenv("SU2B200_FACTOR_REGULAR") == nullptr) d.b.w |= 2;
          else if (reg && atoi(getenv("SU2B200_FACTOR_REGULAR")) != 0) d.b.w |= 2;
          padded.push_back(d);
        }
        { RowDesc e = empty; e.b.w = 2; while (padded.size() % rpw) padded.push_back(e); }
        if (l == last) {
          ls.factor_split[1] = ls.lvl_ptr_lower[l + 1];
          ls.factor_split[3] = (int)(padded.size() / rpw);
        }
      }
      ls.factor_groups = (int)(padded.size() / rpw);
      ls.factor_desc_rows.upload(reinterpret_cast<const int*>(padded.data()), padded.size() * 8);
    }
  }
  ls.factor_desc.upload(reinterpret_cast<const int*>(desc.data()), desc.size() * 8);
  if (opsLevel.empty()) opsLevel.push_back(make_int2(0, 0));
  ls.factor_ops.upload(reinterpret_cast<const int*>(opsLevel.data()), opsLevel.size() * 2);
  ls.factor_own_cap = maxOwn;
  ls.factor_src_cap = std::max(1, std::min(maxOps, MAXS));
  ls.factor_ok = true;
  return true;
}

template <int NV>
static void fill_invM(su2b200_ctx* c, LinSys& ls) {
  const size_t nInv = (size_t)ls.Pd * NV * NV;
  SU2B_LAUNCH(c, k_fill_sentinel_f, (int)std::min<size_t>((nInv + 255) / 256, 148 * 8), 256, nInv, ls.invM.p);
}

// warp-per-row kernel over the level-ordered rows [begin, end)
template <int NV>
static void factor_launch(su2b200_ctx* c, LinSys& ls, int begin, int end) {
  if (end <= begin) return;
  int dev = 0, nsm = 0, perSm = 0;
  int ownCap = ls.factor_own_cap, srcCap = ls.factor_src_cap;
  const size_t smem = (size_t)4 * 2 * (ownCap + srcCap) * NV * NV * sizeof(double);
  SU2B_CUDA(cudaGetDevice(&dev));
  SU2B_CUDA(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
  SU2B_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_factor_poll<NV>, 128, smem));
  static const int cap = getenv("SU2B200_FACTOR_BLOCKS") ? atoi(getenv("SU2B200_FACTOR_BLOCKS")) : 0;
  if (cap > 0) perSm = std::min(perSm, cap);
  SU2B_REQUIRE(perSm > 0, "ilu factor kernel does not fit an SM");
  int n = end - begin;
  const int grid = std::max(1, std::min((n + 3) / 4, nsm * perSm));
  const RowDesc* desc = reinterpret_cast<const RowDesc*>(ls.factor_desc.p) + begin;
  const int2* ops = reinterpret_cast<const int2*>(ls.factor_ops.p);
  int* ctrl = ls.sweep_ctrl.p;
  const int* ci = ls.col_ind;
  const double* A = ls.A.p;
  double* ilu = ls.ilu.p;
  double* invM = ls.invM.p;
  int slim = ls.ilu_slim ? 1 : 0;
  void* argv[] = {&n, &slim, &ownCap, &srcCap, &desc, &ops, &ctrl, &ci, &A, &ilu, &invM};
  SU2B_CUDA(cudaLaunchCooperativeKernel((const void*)k_factor_poll<NV>, dim3(grid), dim3(128), argv, smem, c->stream));
  c->launches++;
}

static void factor_launch_scalar(su2b200_ctx* c, LinSys& ls) {
  int dev = 0, nsm = 0, perSm = 0;
  SU2B_CUDA(cudaGetDevice(&dev));
  SU2B_CUDA(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
  SU2B_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_factor_poll_s, 128, 0));
  static const int cap = getenv("SU2B200_FACTOR_BLOCKS_S") ? atoi(getenv("SU2B200_FACTOR_BLOCKS_S")) : 0;
  if (cap > 0) perSm = std::min(perSm, cap);
  SU2B_REQUIRE(perSm > 0, "scalar ilu factor kernel does not fit an SM");
  int nSlots = ls.factor_slots;
  const int warps = (nSlots + SROWS_F - 1) / SROWS_F;
  const int grid = std::max(1, std::min((warps + 3) / 4, nsm * perSm));
  SU2B_LAUNCH(c, k_fill_sentinel_f, (int)std::min<size_t>(((size_t)ls.Pd + 255) / 256, 148 * 8), 256, (size_t)ls.Pd,
              ls.invM.p);
  const RowDesc* desc = reinterpret_cast<const RowDesc*>(ls.factor_desc.p);
  const int2* ops = reinterpret_cast<const int2*>(ls.factor_ops.p);
  int* ctrl = ls.sweep_ctrl.p;
  const int* ci = ls.col_ind;
  const double* A = ls.A.p;
  double* ilu = ls.ilu.p;
  double* invM = ls.invM.p;
  void* argv[] = {&nSlots, &desc, &ops, &ctrl, &ci, &A, &ilu, &invM};
  SU2B_CUDA(cudaLaunchCooperativeKernel((const void*)k_factor_poll_s, dim3(grid), dim3(128), argv, 0, c->stream));
  c->launches++;
}

// nVar-lanes-per-row kernel over the padded groups [gbeg, gend)
template <int NV>
static void factor_launch_rows(su2b200_ctx* c, LinSys& ls, int gbeg, int gend) {
  if (gend <= gbeg) return;
  int dev = 0, nsm = 0, perSm = 0;
  int ownCap = ls.factor_own_cap, srcCap = std::min(ls.factor_src_cap, RF_SRC);
  const size_t smem = (size_t)4 * (32 / NV) * (ownCap + srcCap + 1) * NV * NV * sizeof(double);
  SU2B_CUDA(cudaFuncSetAttribute((const void*)k_factor_rows<NV>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)((size_t)4 * (32 / NV) * (RF_OWN + RF_SRC + 1) * NV * NV * sizeof(double))));
  SU2B_CUDA(cudaGetDevice(&dev));
  SU2B_CUDA(cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev));
  SU2B_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_factor_rows<NV>, 128, smem));
  static const int capR =
      getenv("SU2B200_FACTOR_ROWS_BLOCKS") ? atoi(getenv("SU2B200_FACTOR_ROWS_BLOCKS")) : FACTOR_ROWS_MIN_BLOCKS;
  if (capR > 0) perSm = std::min(perSm, capR);
  SU2B_REQUIRE(perSm > 0, "ilu factor (rows) kernel does not fit an SM");
  int nGroups = gend - gbeg;
  const int grid = std::max(1, std::min((nGroups + 3) / 4, nsm * perSm));
  const RowDesc* desc = reinterpret_cast<const RowDesc*>(ls.factor_desc_rows.p) + (size_t)gbeg * (32 / NV);
  const int2* ops = reinterpret_cast<const int2*>(ls.factor_ops.p);
  int* ctrl = ls.sweep_ctrl.p;
  const int* ci = ls.col_ind;
  const double* A = ls.A.p;
  double* ilu = ls.ilu.p;
  double* invM = ls.invM.p;
  int slim = ls.ilu_slim ? 1 : 0;
  static const int dynEnv = getenv("SU2B200_FACTOR_DYN") ? atoi(getenv("SU2B200_FACTOR_DYN")) : 1;
  int dyn = dynEnv;
  if (dyn) SU2B_CUDA(cudaMemsetAsync(ctrl, 0, sizeof(int), c->stream));
  void* argv[] = {&nGroups, &slim, &dyn, &ownCap, &srcCap, &desc, &ops, &ctrl, &ci, &A, &ilu, &invM};
  SU2B_CUDA(cudaLaunchCooperativeKernel((const void*)k_factor_rows<NV>, dim3(grid), dim3(128), argv, smem, c->stream));
  c->launches++;
}

// block systems: narrow head -> k_factor_poll, wide middle -> k_factor_rows, narrow tail -> k_factor_poll (launches
// on one stream: a later launch finds everything the earlier ones produced). SU2B200_FACTOR_ROWS=0: one kernel
// (warp per row) for everything, =2: k_factor_rows for everything.
template <int NV>
static void factor_block(su2b200_ctx* c, LinSys& ls) {
  const int rows = getenv("SU2B200_FACTOR_ROWS") ? atoi(getenv("SU2B200_FACTOR_ROWS")) : ls.factor_rows_mode;
  // slim ILU: when no upper block changes (7-point hexahedral stencils: every update lands on a diagonal) and the
  // polling sweeps are the consumer, the upper blocks stay in the Jacobian: not read, not written here
  static const int slimOk = getenv("SU2B200_ILU_SLIM") ? atoi(getenv("SU2B200_ILU_SLIM")) : 1;
  ls.ilu_slim = slimOk && ls.upper_stable && ls.sweep_mode == 2;
  fill_invM<NV>(c, ls);
  if (rows == 0 || !ls.factor_rows_ok) return factor_launch<NV>(c, ls, 0, ls.Pd);
  if (rows == 2) return factor_launch_rows<NV>(c, ls, 0, ls.factor_groups);
  factor_launch<NV>(c, ls, 0, ls.factor_split[0]);
  factor_launch_rows<NV>(c, ls, ls.factor_split[2], ls.factor_split[3]);
  factor_launch<NV>(c, ls, ls.factor_split[1], ls.Pd);
}

// ILU <- factor(A); requires build_levels(ls) and factor_symbolic(ls) == true
void ilu_factor_poll(su2b200_ctx* c, LinSys& ls) {
  switch (ls.nVar) {
    case 1: factor_launch_scalar(c, ls); break;
    case 2: factor_block<2>(c, ls); break;
    case 4: factor_block<4>(c, ls); break;
    case 5: factor_block<5>(c, ls); break;
    default: throw Error("ilu_factor_poll: unsupported nVar");
  }
}

}  // namespace su2b200
