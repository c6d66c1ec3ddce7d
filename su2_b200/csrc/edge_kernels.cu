// a11-a13. The fused edge kernel: CFVMFlowSolverBase::EdgeFluxResidual (CFVMFlowSolverBase.inl:1517-1573)
// with the CNumericsSIMD functor chain: MUSCL + point limiter (common.hpp:113-279), Roe (roe.hpp:87-303)
// or JST (centered.hpp:81-268), viscous decorator (viscous_fluxes.hpp:114-202), and the linear-system
// update (util.hpp:231-263, CSysMatrix.hpp:615-655, CSysVector.hpp:466-483).
//
// Assembly without atomics (SU2B200_ASSEMBLY_GATHER, default):
//   k_edge_*   : one thread per edge. Writes flux_e (SoA [nVar][E]), J_ij = jac_j, J_ji = -jac_i.
//   k_point_sum_rows: nVar lanes per point. res_i = sum_{e in edges(i), ascending} +-flux_e and
//                J_ii = -sum_k J_ki (column sum through the edge map) -- the reference's reducer strategy
//                (SumEdgeFluxes, CFVMFlowSolverBase.inl:1575-1591; SetDiagonalAsColumnSum) which adds in the
//                same order as its serial scatter loop.
// SU2B200_ASSEMBLY_COLORED: one launch per edge colour, read-modify-write of residual and diagonal blocks.
#include "internal.hpp"
#include "device_math.cuh"
#include "kernels.hpp"

namespace su2b200 {

static inline int nblk(long n, int bs) { return (int)((n + bs - 1) / bs); }

struct EdgeArgs {
  int E;
  const int* ei;
  const int* ej;
  const double* enormal;   // SoA [ND][E]
  const double* coord;     // [P][ND]
  const double* V;         // [P][ND+9]
  const double* grad_rec;  // [P][NGS][ND]
  const double* limiter;   // [P][ND+2]
  const double* grad_prim; // [P][NGS][ND]
  int ngs;
  signed char* counter;    // [E]
  // JST
  const double* lambda;
  const double* sensor;
  const double* lapl;
  const int* nnbr;
  // outputs
  double* flux;            // SoA [NV][E]  (GATHER)
  double* A;               // Jacobian blocks
  const int* eij;
  const int* eji;
  // COLORED
  const int* color_edges;  // edge list of this colour (nullptr: identity)
  double* res;             // [P][NV]
  const int* dia_ptr;
  unsigned long long* nonphys;
};

// --- MUSCL reconstruction: reconstructPrimitives (common.hpp:194-279) with musclPointLimited (:113-141) /
//     musclUnlimited (:84-108). Returns true if the edge falls back to first order.
template <int ND>
__device__ __forceinline__ void muscl_reconstruct(const FlowParams& p, const EdgeArgs& a, int e,
                                                  const double* gi, const double* gj,  // reconstruction gradients
                                                  const double* lim_i, const double* lim_j,  // point limiters
                                                  const double* Vi1, const double* Vj1, const double* vij,
                                                  double* Vi, double* Vj) {
  using I = PrimIdx<ND>;
  constexpr int NR = ND + 4;   // reconstructed primitive set (T, v, P, rho, H)
  constexpr int NGR = ND + 2;  // variables with gradients (T, v, P)
#pragma unroll
  for (int k = 0; k < NR; ++k) { Vi[k] = Vi1[k]; Vj[k] = Vj1[k]; }
#pragma unroll
  for (int v = 0; v < NGR; ++v) {
    const double delta = Vj[v] - Vi[v];
    double pi_ = 0.0, pj_ = 0.0;
#pragma unroll
    for (int d = 0; d < ND; ++d) { pi_ += gi[v * ND + d] * vij[d]; pj_ += gj[v * ND + d] * vij[d]; }
    // musclReconstruction: umusclRamp * ((1-kappa)*proj + kappa*delta), ramp == 1
    const double proj_i = 1.0 * ((1.0 - p.muscl_kappa) * pi_ + p.muscl_kappa * delta);
    const double proj_j = 1.0 * ((1.0 - p.muscl_kappa) * pj_ + p.muscl_kappa * delta);
    if (p.limiter) {
      const double li = lim_i[v];
      const double lj = lim_j[v];
      Vi[v] += 0.5 * li * proj_i;
      Vj[v] -= 0.5 * lj * proj_j;
    } else {
      Vi[v] += 0.5 * proj_i;
      Vj[v] -= 0.5 * proj_j;
    }
  }
  Vi[I::RHO] = Vi[I::P] / (p.R * Vi[I::T]);
  Vj[I::RHO] = Vj[I::P] / (p.R * Vj[I::T]);
  const double cp = p.R * p.gamma / (p.gamma - 1);
  Vi[I::H] += cp * (Vi[I::T] - Vi1[I::T]);
  Vj[I::H] += cp * (Vj[I::T] - Vj1[I::T]);
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    Vi[I::H] += 0.5 * (Vi[I::VEL + d] * Vi[I::VEL + d] - Vi1[I::VEL + d] * Vi1[I::VEL + d]);
    Vj[I::H] += 0.5 * (Vj[I::VEL + d] * Vj[I::VEL + d] - Vj1[I::VEL + d] * Vj1[I::VEL + d]);
  }
  // non-physical reconstruction detection
  const bool neg_p_or_rho = (fmin(Vi[I::P], Vj[I::P]) < 0.0) || (fmin(Vi[I::RHO], Vj[I::RHO]) < 0.0);
  const double R = sqrt(fabs(Vj[I::RHO] / Vi[I::RHO]));
  const double enthalpy = R * Vj[I::H] + Vi[I::H];
  double v2 = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) { const double t = R * Vj[I::VEL + d] + Vi[I::VEL + d]; v2 += t * t; }
  const bool neg_c = enthalpy * (R + 1) < 0.5 * v2;
  bool bad = neg_p_or_rho || neg_c;
  // UpdateNonPhysicalEdgeCounter (CFlowVariable.hpp:80-87)
  signed char cnt = a.counter[e];
  if (bad) cnt = 21;
  cnt = (cnt - 1 > 0) ? (signed char)(cnt - 1) : (signed char)0;
  a.counter[e] = cnt;
  if (cnt > 0) {
    atomicAdd(a.nonphys, 1ull);
#pragma unroll
    for (int k = 0; k < NR; ++k) { Vi[k] = Vi1[k]; Vj[k] = Vj1[k]; }
  }
}

// The two Jacobian blocks of an edge live in shared memory while the thread builds them (entry k of thread t at
// [k * JSTRIDE + lane] of the warp's tile: conflict-free for the owning thread, and for the 25 lanes that later read one
// edge's block),
// which takes 100 registers off the thread and lets the warp write every 200-byte block as one contiguous store.
#ifndef EDGE_MIN_BLOCKS
#define EDGE_MIN_BLOCKS 3  // resident blocks per SM the register allocation aims at
#endif
static constexpr int EDGE_BLOCK = 128;
// per-edge data is touched once per sweep (indices, normals, flux, Jacobian blocks): streaming loads / stores keep the
// caches for the point data that neighbouring edges share
// (measured on 185^3: streaming block stores 5.53 -> 4.80 ms; streaming the index/normal loads and the flux stores as
//  well 5.01 ms, so only the Jacobian blocks are streamed)
#define LD_ONCE(p) (*(p))
#ifdef EDGE_NO_STREAM
#define ST_ONCE(p, v) (*(p) = (v))
#else
#define ST_ONCE(p, v) __stcs((p), (v))
#endif

// warp-major tiles: warp w owns sjac[w * rows * JSTRIDE ...], entry k of its lane l at [k * JSTRIDE + l]
static constexpr int JSTRIDE = 33;
struct SmemJac {
  double* p;
  __device__ __forceinline__ double& operator[](int k) const { return p[k * JSTRIDE]; }
};
template <int NV>
static constexpr size_t edge_smem_bytes(int extra_rows = 0) {
  return (size_t)(EDGE_BLOCK / 32) * (2 * NV * NV + extra_rows) * JSTRIDE * sizeof(double);
}

// flux (GATHER) or residual/diagonal read-modify-write (COLORED) by the owning thread, then the off-diagonal blocks
// J_ij = jac_j, J_ji = -jac_i by the whole warp: for each of its 32 edges, lanes 0..NV*NV-1 store one block.
template <int ND, bool COLORED>
__device__ __forceinline__ void store_edge_outputs(const EdgeArgs& a, bool valid, int e, int i, int j, int bij, int bji,
                                                   const double* flux, SmemJac jac_i, SmemJac jac_j,
                                                   const double* wtile) {
  constexpr int NV = ND + 2;
  if (valid) {
    if (!COLORED) {
#pragma unroll
      for (int v = 0; v < NV; ++v) a.flux[(size_t)v * a.E + e] = flux[v];
    } else {
      double* bii = a.A + (size_t)a.dia_ptr[i] * NV * NV;
      double* bjj = a.A + (size_t)a.dia_ptr[j] * NV * NV;
#pragma unroll
      for (int k = 0; k < NV * NV; ++k) { bii[k] -= -jac_i[k]; bjj[k] -= jac_j[k]; }
#pragma unroll
      for (int v = 0; v < NV; ++v) { a.res[(size_t)i * NV + v] += flux[v]; a.res[(size_t)j * NV + v] -= flux[v]; }
    }
  }
  __syncwarp();
  const int lane = threadIdx.x & 31;
  const double* si = wtile;
  const double* sj = wtile + NV * NV * JSTRIDE;
#pragma unroll 4
  for (int t = 0; t < 32; ++t) {
    const int dst_ij = __shfl_sync(0xffffffffu, bij, t);
    const int dst_ji = __shfl_sync(0xffffffffu, bji, t);
    if (dst_ij >= 0 && lane < NV * NV) {
      // write-once data: evict-first, so that 400 bytes per edge of Jacobian do not push the gathered point data out
      // of L2 (measured on 185^3: 5.53 -> 4.80 ms)
      ST_ONCE(a.A + (size_t)dst_ij * NV * NV + lane, sj[lane * JSTRIDE + t]);
      ST_ONCE(a.A + (size_t)dst_ji * NV * NV + lane, -si[lane * JSTRIDE + t]);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Staged gathers (experiment). A thread that reads its own point records (AoS, 24..120 bytes each) with scalar loads
// makes every load instruction touch up to 32 different 128-byte lines, and the L1 data pipe is the busiest unit of the
// kernel (75 %). Here the WARP loads the 32 records of its 32 edges with consecutive lanes on consecutive doubles (1-2 lines per
// record), and the records are handed to their owning lanes through shared memory. The scratch is the warp's slice of
// the Jacobian tile, which is not in use before the flux is computed.
// ------------------------------------------------------------------------------------------------
// MEASURED AND NOT ADOPTED (185^3 RANS, B200): 5.46 ms staged against 4.80 ms with the direct gathers below. ncu on 129^3
// (profiles/r01_ncu_edge_staging.txt): L1 sectors of the global loads 401 M -> 110 M and the data-pipe wavefronts per SM
// 2.57 M -> 2.28 M (75 % -> 63 % busy), but the executed instructions grow by ~40 % (index arithmetic, shuffles, STS/LDS)
// and the two load -> shared -> sync -> read phases sit on the critical path of a warp: with 12 warps per SM (168
// registers) the kernel is bound by latency per warp, not by the L1. Kept behind EDGE_STAGED=1 for the parity suite.
#ifndef EDGE_STAGED
#define EDGE_STAGED 0
#endif
// One load instruction takes G = 32 / REC whole records (lane = g * REC + c reads component c of record k * G + g), so
// that every address is a per-lane base plus a compile-time offset; a record of the scratch is S = REC | 1 doubles
// (odd: the per-lane record reads are conflict-free).
template <int REC>
struct StageRec {
  static constexpr int S = REC | 1, SIZE = 32 * S, G = 32 / REC, NI = (32 + G - 1) / G;
};
template <int REC>
__device__ __forceinline__ void stage_load(const double* __restrict__ base, int stride, int idx, double* tmp) {
  using R = StageRec<REC>;
  const int lane = threadIdx.x & 31, g = lane / REC, c = lane - g * REC;
#pragma unroll
  for (int k = 0; k < R::NI; ++k) {
    const int r = k * R::G + g;
    const int src = __shfl_sync(0xffffffffu, idx, r & 31);
    if (g < R::G && r < 32) tmp[k] = __ldg(base + (size_t)src * stride + c);  // (stage_put has the same guard)
  }
}
template <int REC>
__device__ __forceinline__ void stage_put(double* sc, const double* tmp) {
  using R = StageRec<REC>;
  const int lane = threadIdx.x & 31, g = lane / REC, c = lane - g * REC;
  double* mine = sc + g * R::S + c;
#pragma unroll
  for (int k = 0; k < R::NI; ++k)
    if (g < R::G && k * R::G + g < 32) mine[k * R::G * R::S] = tmp[k];
}
template <int REC>
__device__ __forceinline__ void stage_get(const double* sc, double* out) {
  const double* mine = sc + (threadIdx.x & 31) * StageRec<REC>::S;
#pragma unroll
  for (int c = 0; c < REC; ++c) out[c] = mine[c];
}

// ------------------------------------------------------------------------------------------------
// Roe + MUSCL (+ viscous) edge kernel
// ------------------------------------------------------------------------------------------------
// rows of a warp's tile beyond the two Jacobians: the mean viscous gradient of each edge when the gathers are staged
template <int ND, bool VISC>
static constexpr int edge_roe_extra_rows() {
  return EDGE_STAGED && VISC ? (ND + 1) * ND : 0;
}

template <int ND, bool VISC, bool COLORED>
__global__ void __launch_bounds__(EDGE_BLOCK, EDGE_MIN_BLOCKS) k_edge_roe(EdgeArgs a, FlowParams p, int n) {
  using I = PrimIdx<ND>;
  constexpr int NV = ND + 2;
  constexpr int NP1 = VISC ? ND + 9 : ND + 4;  // gathered first-order primitives
  constexpr int NGR = ND + 2;                  // reconstructed variables with gradients
  extern __shared__ double sjac[];
  constexpr int WT = (2 * NV * NV + edge_roe_extra_rows<ND, VISC>()) * JSTRIDE;
  double* const wtile = sjac + (threadIdx.x >> 5) * WT;
  const SmemJac jac_i{wtile + (threadIdx.x & 31)}, jac_j{wtile + NV * NV * JSTRIDE + (threadIdx.x & 31)};
  const int t = blockIdx.x * EDGE_BLOCK + threadIdx.x;
  const bool valid = t < n;
  int e = 0, i = 0, j = 0, bij = -1, bji = -1;
  if (valid) {
    e = COLORED ? a.color_edges[t] : t;
    i = LD_ONCE(a.ei + e); j = LD_ONCE(a.ej + e);
    bij = LD_ONCE(a.eij + e); bji = LD_ONCE(a.eji + e);  // used by the stores at the very end
  }
  double flux[NV];
  double normal[ND], vij[ND];
  double Vi1[I::NPRIM], Vj1[I::NPRIM];
#if EDGE_STAGED
  double* const sc = wtile;  // the Jacobian rows of the tile are free until the flux is computed
  const SmemJac gmean{wtile + 2 * NV * NV * JSTRIDE + (threadIdx.x & 31)};
  // the viscous part wants the mean of the same gradients the reconstruction reads: form it while they are here
  const bool mean_staged = VISC && p.muscl && a.grad_prim == a.grad_rec;
  {
    double tv[2][StageRec<NP1>::NI], tc[2][StageRec<ND>::NI];
    stage_load<NP1>(a.V, I::NPRIM, i, tv[0]);
    stage_load<NP1>(a.V, I::NPRIM, j, tv[1]);
    stage_load<ND>(a.coord, ND, i, tc[0]);
    stage_load<ND>(a.coord, ND, j, tc[1]);
    constexpr int OV = StageRec<NP1>::SIZE, OC = StageRec<ND>::SIZE;
    static_assert(2 * OV + 2 * OC <= 2 * NV * NV * JSTRIDE, "stage 1 exceeds the warp's Jacobian tile");
    stage_put<NP1>(sc, tv[0]);
    stage_put<NP1>(sc + OV, tv[1]);
    stage_put<ND>(sc + 2 * OV, tc[0]);
    stage_put<ND>(sc + 2 * OV + OC, tc[1]);
    __syncwarp();
    double ci[ND], cj[ND];
    stage_get<NP1>(sc, Vi1);
    stage_get<NP1>(sc + OV, Vj1);
    stage_get<ND>(sc + 2 * OV, ci);
    stage_get<ND>(sc + 2 * OV + OC, cj);
#pragma unroll
    for (int d = 0; d < ND; ++d) vij[d] = cj[d] - ci[d];
    __syncwarp();
  }
  double gi[NGR * ND], gj[NGR * ND], lim_i[NGR], lim_j[NGR];
  if (p.muscl) {
    double tg[2][StageRec<NGR * ND>::NI], tl[2][StageRec<NGR>::NI];
    stage_load<NGR * ND>(a.grad_rec, a.ngs * ND, i, tg[0]);
    stage_load<NGR * ND>(a.grad_rec, a.ngs * ND, j, tg[1]);
    if (p.limiter) {
      stage_load<NGR>(a.limiter, NGR, i, tl[0]);
      stage_load<NGR>(a.limiter, NGR, j, tl[1]);
    }
    constexpr int OG = StageRec<NGR * ND>::SIZE, OL = StageRec<NGR>::SIZE;
    static_assert(2 * OG + 2 * OL <= 2 * NV * NV * JSTRIDE, "stage 2 exceeds the warp's Jacobian tile");
    stage_put<NGR * ND>(sc, tg[0]);
    stage_put<NGR * ND>(sc + OG, tg[1]);
    if (p.limiter) {
      stage_put<NGR>(sc + 2 * OG, tl[0]);
      stage_put<NGR>(sc + 2 * OG + OL, tl[1]);
    }
    __syncwarp();
    stage_get<NGR * ND>(sc, gi);
    stage_get<NGR * ND>(sc + OG, gj);
    if (p.limiter) {
      stage_get<NGR>(sc + 2 * OG, lim_i);
      stage_get<NGR>(sc + 2 * OG + OL, lim_j);
    }
    __syncwarp();
    if (mean_staged) {
#pragma unroll
      for (int k = 0; k < (ND + 1) * ND; ++k) {  // the expression of viscous_terms
        double m = gi[k];
        m *= 0.5;
        m += 0.5 * gj[k];
        gmean[k] = m;
      }
    }
  }
#else
  const double *gi = nullptr, *gj = nullptr, *lim_i = nullptr, *lim_j = nullptr;
  const bool mean_staged = false;
  const SmemJac gmean{wtile};
  if (valid) {
#pragma unroll
    for (int d = 0; d < ND; ++d) vij[d] = a.coord[(size_t)j * ND + d] - a.coord[(size_t)i * ND + d];
#pragma unroll
    for (int k = 0; k < NP1; ++k) {
      Vi1[k] = a.V[(size_t)i * I::NPRIM + k];
      Vj1[k] = a.V[(size_t)j * I::NPRIM + k];
    }
    gi = a.grad_rec + (size_t)i * a.ngs * ND;
    gj = a.grad_rec + (size_t)j * a.ngs * ND;
    lim_i = a.limiter + (size_t)i * NGR;
    lim_j = a.limiter + (size_t)j * NGR;
  }
#endif
  if (valid) {
#pragma unroll
    for (int d = 0; d < ND; ++d) normal[d] = LD_ONCE(a.enormal + (size_t)d * a.E + e);
    double Vi[ND + 4], Vj[ND + 4];
    if (p.muscl) {
      muscl_reconstruct<ND>(p, a, e, gi, gj, lim_i, lim_j, Vi1, Vj1, vij, Vi, Vj);
    } else {
#pragma unroll
      for (int k = 0; k < ND + 4; ++k) { Vi[k] = Vi1[k]; Vj[k] = Vj1[k]; }
    }
    roe_flux<ND, true>(p, Vi, Vj, normal, flux, jac_i, jac_j);
    if (VISC) {
      const double area = sqrt(dotN<ND>(normal, normal));
      double un[ND];
#pragma unroll
      for (int d = 0; d < ND; ++d) un[d] = normal[d] / area;
      viscous_terms<ND, true, false>(p, Vi1, Vj1, a.grad_prim + (size_t)i * a.ngs * ND,
                                     a.grad_prim + (size_t)j * a.ngs * ND, vij, area, un, true, flux, jac_i, jac_j,
                                     mean_staged, gmean);
    }
  }  // valid
  store_edge_outputs<ND, COLORED>(a, valid, e, i, j, bij, bji, flux, jac_i, jac_j, wtile);
}

// ------------------------------------------------------------------------------------------------
// JST (+ viscous) edge kernel: CCenteredBase::ComputeFlux + CJSTScheme::finalizeFlux
// ------------------------------------------------------------------------------------------------
template <int ND, bool VISC, bool COLORED>
__global__ void __launch_bounds__(EDGE_BLOCK, EDGE_MIN_BLOCKS) k_edge_jst(EdgeArgs a, FlowParams p, int n) {
  using I = PrimIdx<ND>;
  constexpr int NV = ND + 2;
  constexpr int NP1 = VISC ? ND + 9 : ND + 5;
  extern __shared__ double sjac[];
  double* const wtile = sjac + (threadIdx.x >> 5) * (2 * NV * NV * JSTRIDE);
  const SmemJac jac_i{wtile + (threadIdx.x & 31)}, jac_j{wtile + NV * NV * JSTRIDE + (threadIdx.x & 31)};
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = t < n;
  int e = 0, i = 0, j = 0, bij = -1, bji = -1;
  double flux[NV];
  if (valid) {
  e = COLORED ? a.color_edges[t] : t;
  i = LD_ONCE(a.ei + e); j = LD_ONCE(a.ej + e);
  bij = LD_ONCE(a.eij + e); bji = LD_ONCE(a.eji + e);
  double normal[ND];
#pragma unroll
  for (int d = 0; d < ND; ++d) normal[d] = LD_ONCE(a.enormal + (size_t)d * a.E + e);
  const double area = sqrt(dotN<ND>(normal, normal));
  double un[ND];
#pragma unroll
  for (int d = 0; d < ND; ++d) un[d] = normal[d] / area;
  double Vi[I::NPRIM], Vj[I::NPRIM], avgV[I::NPRIM];
#pragma unroll
  for (int k = 0; k < NP1; ++k) {
    Vi[k] = a.V[(size_t)i * I::NPRIM + k];
    Vj[k] = a.V[(size_t)j * I::NPRIM + k];
    avgV[k] = 0.5 * (Vi[k] + Vj[k]);
  }
  // conservative variables
  double Ui[NV], Uj[NV], diffU[NV];
  Ui[0] = Vi[I::RHO];
  Uj[0] = Vj[I::RHO];
#pragma unroll
  for (int d = 0; d < ND; ++d) { Ui[d + 1] = Vi[I::RHO] * Vi[I::VEL + d]; Uj[d + 1] = Vj[I::RHO] * Vj[I::VEL + d]; }
  Ui[ND + 1] = Vi[I::RHO] * Vi[I::H] - Vi[I::P];
  Uj[ND + 1] = Vj[I::RHO] * Vj[I::H] - Vj[I::P];
#pragma unroll
  for (int k = 0; k < NV - 1; ++k) diffU[k] = Ui[k] - Uj[k];
  diffU[NV - 1] = Vi[I::RHO] * Vi[I::H] - Vj[I::RHO] * Vj[I::H];

  inviscid_proj_flux<ND>(avgV, normal, flux);
  inviscid_proj_jac<ND>(p.gamma, &Vi[I::VEL], Ui[ND + 1] / Ui[0], normal, 0.5, jac_i);
  inviscid_proj_jac<ND>(p.gamma, &Vj[I::VEL], Uj[ND + 1] / Uj[0], normal, 0.5, jac_j);
  const double projVel = dotN<ND>(&avgV[I::VEL], normal);

  // CJSTScheme::finalizeFlux (centered.hpp:212-268)
  double lambda = fabs(projVel) + avgV[I::C] * area;
  {  // correctedSpectralRadius (common.hpp:514-528), stretchParam = 0.3
    const double li = a.lambda[i], lj = a.lambda[j];
    const double phi_i = pow(0.25 * li / lambda, 0.3);
    const double phi_j = pow(0.25 * lj / lambda, 0.3);
    lambda = 4 * phi_i * phi_j / (phi_i + phi_j) * lambda;
  }
  const double ni = (double)a.nnbr[i], nj = (double)a.nnbr[j];
  if (p.lax) {
    // CLaxScheme::finalizeFlux (centered.hpp:474-530): first-order scalar dissipation on the coarse multigrid levels
    const double dissip = p.lax_k0 * ND * (ni + nj) / (ni * nj) * lambda;
#pragma unroll
    for (int v = 0; v < NV; ++v) flux[v] += dissip * diffU[v];
    scalar_dissipation_jacobian<ND>(Vi, p.gamma, p.jst_fix * dissip, jac_i);
    scalar_dissipation_jacobian<ND>(Vj, p.gamma, -p.jst_fix * dissip, jac_j);
  } else {
  const double sc2 = 3 * (ni + nj) / (ni * nj);
  const double sc4 = 0.25 * sc2 * sc2;
  const double si = a.sensor[i], sj = a.sensor[j];
  const double eps2 = p.jst_k2 * 0.5 * (si + sj) * sc2;
  const double eps4 = fmax(0.0, p.jst_k4 - eps2) * sc4;
#pragma unroll
  for (int v = 0; v < NV; ++v) {
    const double li = a.lapl[(size_t)i * NV + v], lj = a.lapl[(size_t)j * NV + v];
    flux[v] += (eps2 * diffU[v] - eps4 * (li - lj)) * lambda;
  }
  const double dissip_i = p.jst_fix * (eps2 + eps4 * (ni + 1)) * lambda;
  const double dissip_j = -p.jst_fix * (eps2 + eps4 * (nj + 1)) * lambda;
  scalar_dissipation_jacobian<ND>(Vi, p.gamma, dissip_i, jac_i);
  scalar_dissipation_jacobian<ND>(Vj, p.gamma, dissip_j, jac_j);
  }

  if (VISC) {
    double vij[ND];
#pragma unroll
    for (int d = 0; d < ND; ++d) vij[d] = a.coord[(size_t)j * ND + d] - a.coord[(size_t)i * ND + d];
    viscous_terms<ND, true, false>(p, Vi, Vj, a.grad_prim + (size_t)i * a.ngs * ND,
                                   a.grad_prim + (size_t)j * a.ngs * ND, vij, area, un, true, flux, jac_i, jac_j);
  }
  }  // valid
  store_edge_outputs<ND, COLORED>(a, valid, e, i, j, bij, bji, flux, jac_i, jac_j, wtile);
}

// ------------------------------------------------------------------------------------------------
// Point gather: residual = signed sum of edge fluxes, diagonal block = -(column sum of off-diagonals), over the
// neighbour CSR in ascending edge index (== reference order of CPoint::Edge, which is built in edge-creation order).
// gat_* hold, per point, its edges in that order: gat_edge = e if the point is the edge's first node, ~e if it is the
// second; gat_blk = block (nbr, i).
// NV lanes per point (the mapping of k_spmv_rows): lane r of a point's group accumulates elements r, r + NV, ... of
// the diagonal block (the NV lanes of a group read NV consecutive doubles of every block row) and component r of the
// residual; 32 / NV points per warp, 40 registers, full occupancy. Measured on 185^3: 1.99 ms (5.9 TB/s of DRAM traffic,
// 90 % of the measured peak) against 2.89 ms for one thread per point (25 scalar loads per block, each touching 32
// lines) and 6.3 ms for a warp walking the CSR range of 32 points with 25 lanes per block (serialised entries).
// ------------------------------------------------------------------------------------------------
template <int NV>
__global__ void __launch_bounds__(128) k_point_sum_rows(int P, int E, const int* __restrict__ nbr_ptr,
                                                        const int* __restrict__ gat_edge, const int* __restrict__ gat_blk,
                                                        const int* __restrict__ dia_ptr, const double* __restrict__ flux,
                                                        double* A, double* __restrict__ res) {
  constexpr int RPW = 32 / NV;
  const int lane = threadIdx.x & 31, sub = lane / NV, r = lane - sub * NV;
  const int i = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * RPW + sub;
  if (sub >= RPW || i >= P) return;
  double D[NV], rr = 0.0;
#pragma unroll
  for (int c = 0; c < NV; ++c) D[c] = 0.0;
  const int kend = nbr_ptr[i + 1];
#pragma unroll 2
  for (int k = nbr_ptr[i]; k < kend; ++k) {
    const int se = gat_edge[k];
    const bool first = se >= 0;
    const double sgn = first ? 1.0 : -1.0;
    rr += sgn * flux[(size_t)r * E + (first ? se : ~se)];
    const double* B = A + (size_t)gat_blk[k] * NV * NV + r;
#pragma unroll
    for (int c = 0; c < NV; ++c) D[c] -= LD_ONCE(B + c * NV);
  }
  res[(size_t)i * NV + r] = rr;
  double* Dg = A + (size_t)dia_ptr[i] * NV * NV + r;
#pragma unroll
  for (int c = 0; c < NV; ++c) Dg[c * NV] = D[c];
}

template <int NV>
static void launch_point_sum(su2b200_ctx* c) {
  auto& g = c->geo;
  auto& s = c->fs;
  auto& ls = c->sys[0];
  constexpr int PPB = 4 * (32 / NV);  // points per block of 4 warps
  SU2B_LAUNCH(c, k_point_sum_rows<NV>, nblk(g.P, PPB), 128, g.P, g.E, g.nbr_ptr.p, g.gat_edge.p, g.gat_blk.p,
              g.dia_ptr.p, s.edge_flux.p, ls.A.p, s.res.p);
}

static FlowParams make_flow_params(const su2b200_flow_config& c) {
  FlowParams p;
  p.gamma = c.gamma;
  p.gm1 = c.gamma - 1.0;
  p.R = c.gas_constant;
  p.roe_kappa = c.roe_kappa;
  p.entropy_fix = c.entropy_fix;
  p.muscl_kappa = c.muscl_kappa;
  p.prandtl_turb = c.prandtl_turb;
  p.jst_k2 = c.jst_kappa2;
  p.jst_k4 = c.jst_kappa4;
  p.jst_fix = c.cent_jac_fix_factor;
  // agglomerated multigrid levels: first order (MUSCL only on MESH_0, CEulerSolver.cpp:1662) and, for centred
  // schemes, the Lax-Friedrichs dissipation (CNumericsSIMD.cpp:70)
  p.muscl = c.muscl && c.mg_level == 0;
  p.limiter = (c.limiter != SU2B200_LIMITER_NONE) && c.mg_level == 0;
  p.viscous = c.viscous;
  p.implicit = 1;
  p.lax = c.mg_level > 0;
  p.lax_k0 = c.lax_kappa1;
  return p;
}
FlowParams flow_params(const su2b200_ctx* c) { return make_flow_params(c->cfg); }

template <class K>
static void launch_edge_kernel(su2b200_ctx* c, K kernel, size_t smem, const EdgeArgs& a, const FlowParams& p, int n,
                               int edges_per_thread) {
  if (n <= 0) return;
  // above the 48 KB default: opt in once per kernel
  static const void* configured[32] = {};
  bool seen = false;
  for (const void* k : configured) seen = seen || k == (const void*)kernel;
  if (!seen) {
    SU2B_CUDA(cudaFuncSetAttribute((const void*)kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (auto& k : configured)
      if (!k) { k = (const void*)kernel; break; }
  }
  kernel<<<nblk(n, EDGE_BLOCK * edges_per_thread), EDGE_BLOCK, smem, c->stream>>>(a, p, n);
  c->launches++;
  SU2B_CUDA(cudaGetLastError());
}

template <int ND>
static void launch_edge(su2b200_ctx* c, const EdgeArgs& a, const FlowParams& p, int n, bool colored) {
  const bool visc = c->cfg.viscous;
  const bool jst = (c->cfg.conv_scheme == SU2B200_CONV_JST);
  const size_t smem = edge_smem_bytes<ND + 2>();
#define SU2B_EDGE(K, V, C, EPT) launch_edge_kernel(c, K<ND, V, C>, smem, a, p, n, EPT)
  if (jst) {
    if (visc) { if (colored) SU2B_EDGE(k_edge_jst, true, true, 1); else SU2B_EDGE(k_edge_jst, true, false, 1); }
    else { if (colored) SU2B_EDGE(k_edge_jst, false, true, 1); else SU2B_EDGE(k_edge_jst, false, false, 1); }
  } else {
#define SU2B_EDGE_ROE(V, C) \
  launch_edge_kernel(c, k_edge_roe<ND, V, C>, edge_smem_bytes<ND + 2>(edge_roe_extra_rows<ND, V>()), a, p, n, 1)
    if (visc) { if (colored) SU2B_EDGE_ROE(true, true); else SU2B_EDGE_ROE(true, false); }
    else { if (colored) SU2B_EDGE_ROE(false, true); else SU2B_EDGE_ROE(false, false); }
#undef SU2B_EDGE_ROE
  }
#undef SU2B_EDGE
}

void flow_edge_residual(su2b200_ctx* c, unsigned long* nonphysical_edges) {
  ++c->sys[0].A_gen;
  auto& g = c->geo;
  auto& s = c->fs;
  auto& ls = c->sys[0];
  EdgeArgs a{};
  a.E = g.E;
  a.ei = g.edge_i.p;
  a.ej = g.edge_j.p;
  a.enormal = g.edge_normal.p;
  a.coord = g.coord.p;
  a.V = s.V.p;
  a.grad_rec = s.recon_aliases_prim ? s.grad_prim.p : s.grad_recon.p;
  a.limiter = s.limiter.p;
  a.grad_prim = s.grad_prim.p;
  a.ngs = s.nG;
  a.counter = s.edge_counter.p;
  a.lambda = s.lambda.p;
  a.sensor = s.sensor.p;
  a.lapl = s.und_lapl.p;
  a.nnbr = g.nnbr.p;
  a.flux = s.edge_flux.p;
  a.A = ls.A.p;
  a.eij = g.edge_ij.p;
  a.eji = g.edge_ji.p;
  a.color_edges = nullptr;
  a.res = s.res.p;
  a.dia_ptr = g.dia_ptr.p;
  a.nonphys = s.counters.p + 1;
  const FlowParams p = make_flow_params(c->cfg);
  SU2B_CUDA(cudaMemsetAsync(s.counters.p + 1, 0, sizeof(unsigned long long), c->stream));

  if (c->cfg.assembly == SU2B200_ASSEMBLY_COLORED) {
    SU2B_REQUIRE(g.color_ptr.size() >= 2, "COLORED assembly needs an edge colouring");
    const int nc = (int)g.color_ptr.size() - 1;
    ProfScope prof_(c, PROF_EDGE);
    for (int col = 0; col < nc; ++col) {
      a.color_edges = g.color_edges.p + g.color_ptr[col];
      const int n = g.color_ptr[col + 1] - g.color_ptr[col];
      if (g.nDim == 2) launch_edge<2>(c, a, p, n, true); else launch_edge<3>(c, a, p, n, true);
    }
  } else {
    if (g.nDim == 2) {
      { ProfScope prof_(c, PROF_EDGE); launch_edge<2>(c, a, p, g.E, false); }
      ProfScope prof2_(c, PROF_POINT_SUM);
      launch_point_sum<4>(c);
    } else {
      { ProfScope prof_(c, PROF_EDGE); launch_edge<3>(c, a, p, g.E, false); }
      ProfScope prof2_(c, PROF_POINT_SUM);
      launch_point_sum<5>(c);
    }
  }
  if (nonphysical_edges) {
    SU2B_CUDA(cudaMemcpyAsync(s.h_counters, s.counters.p, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    SU2B_CUDA(cudaStreamSynchronize(c->stream));
    *nonphysical_edges = (unsigned long)s.h_counters[1];
  }
}

// micro-benchmark entry: re-launch only the edge kernel (GATHER form) `reps` times
void bench_edge_kernel(su2b200_ctx* c, int reps, float* ms_avg) {
  auto& g = c->geo;
  auto& s = c->fs;
  auto& ls = c->sys[0];
  EdgeArgs a{};
  a.E = g.E; a.ei = g.edge_i.p; a.ej = g.edge_j.p; a.enormal = g.edge_normal.p; a.coord = g.coord.p; a.V = s.V.p;
  a.grad_rec = s.recon_aliases_prim ? s.grad_prim.p : s.grad_recon.p; a.limiter = s.limiter.p; a.grad_prim = s.grad_prim.p;
  a.ngs = s.nG; a.counter = s.edge_counter.p; a.lambda = s.lambda.p; a.sensor = s.sensor.p; a.lapl = s.und_lapl.p;
  a.nnbr = g.nnbr.p; a.flux = s.edge_flux.p; a.A = ls.A.p; a.eij = g.edge_ij.p; a.eji = g.edge_ji.p;
  a.res = s.res.p; a.dia_ptr = g.dia_ptr.p; a.nonphys = s.counters.p + 1;
  const FlowParams p = make_flow_params(c->cfg);
  // the kernel decrements the per-edge non-physical counters (NonPhysicalEdgeCounter is solver state) and bumps the
  // device counter: a benchmark must leave both as it found them. The off-diagonal blocks and the edge fluxes it
  // rewrites are scratch of the next assembly (the Jacobian is zeroed by the next Preprocessing); A_gen records it.
  dvec<signed char> saved;
  saved.alloc(s.edge_counter.n);
  if (saved.n) SU2B_CUDA(cudaMemcpyAsync(saved.p, s.edge_counter.p, saved.n, cudaMemcpyDeviceToDevice, c->stream));
  unsigned long long cnt[2] = {0, 0};
  SU2B_CUDA(cudaMemcpyAsync(cnt, s.counters.p, sizeof(cnt), cudaMemcpyDeviceToHost, c->stream));
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  ++ls.A_gen;
  SU2B_CUDA(cudaEventRecord(c->ev0, c->stream));
  for (int r = 0; r < reps; ++r) {
    if (g.nDim == 2) launch_edge<2>(c, a, p, g.E, false); else launch_edge<3>(c, a, p, g.E, false);
  }
  SU2B_CUDA(cudaEventRecord(c->ev1, c->stream));
  SU2B_CUDA(cudaEventSynchronize(c->ev1));
  float ms = 0;
  SU2B_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
  *ms_avg = ms / reps;
  if (saved.n) SU2B_CUDA(cudaMemcpyAsync(s.edge_counter.p, saved.p, saved.n, cudaMemcpyDeviceToDevice, c->stream));
  SU2B_CUDA(cudaMemcpyAsync(s.counters.p, cnt, sizeof(cnt), cudaMemcpyHostToDevice, c->stream));
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
}

}  // namespace su2b200
