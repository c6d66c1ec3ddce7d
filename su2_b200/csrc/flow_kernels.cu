// Point-loop kernels of the flow solver: primitives, gradients (Green-Gauss / least squares),
// Venkatakrishnan limiter, local time step, JST support fields, implicit prepare/complete.
// Every kernel cites the reference loop it restates (su2code/SU2 v8.4.0).
#include "internal.hpp"
#include "device_math.cuh"
#include "kernels.hpp"

namespace su2b200 {

static inline int nblk(long n, int bs) { return (int)((n + bs - 1) / bs); }

// =============================================================================================
// a5. Primitive variables: CEulerVariable::SetPrimVar (CEulerVariable.cpp:102-155),
//     CNSVariable::SetPrimVar (CNSVariable.cpp:138-210), CIdealGas::SetTDState_rhoe (CIdealGas.cpp:40-52),
//     CSutherland::SetViscosity (CSutherland.hpp:47-56), CConstantPrandtl (CConstantPrandtl.hpp:47-52).
// =============================================================================================
struct PrimParams {
  double gamma, gm1, R, cp;
  int viscous;
  double mu_ref, t_ref, suth, mu_const, pr_lam;
};

template <int ND>
__device__ __forceinline__ bool prim_from_cons(const PrimParams& pp, const double* U, double* V, double tke = 0.0) {
  using I = PrimIdx<ND>;
  constexpr int NV = ND + 2;
  double vel2 = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) {
    V[I::VEL + d] = U[d + 1] / U[0];
    vel2 += V[I::VEL + d] * V[I::VEL + d];
  }
  const double rho = U[0];
  // CNSVariable::SetPrimVar: staticEnergy = E - 0.5 v^2 - turb_ke (SST only; x - 0.0 == x otherwise)
  const double e = U[NV - 1] / U[0] - 0.5 * vel2 - tke;
  const double pres = pp.gm1 * rho * e;
  const double temp = pp.gm1 * e / pp.R;
  const double c2 = pp.gamma * pres / rho;
  V[I::RHO] = rho;
  V[I::P] = pres;
  V[I::T] = temp;
  bool bad = (rho <= 0.0) || (pres <= 0.0) || (temp <= 0.0);
  if (c2 < 0.0) bad = true; else V[I::C] = sqrt(c2);
  return bad;
}

template <int ND>
__global__ void k_primitives(int P, PrimParams pp, double* __restrict__ U, const double* __restrict__ U_old,
                             const double* __restrict__ mu_t, const double* __restrict__ tk, double* __restrict__ V,
                             unsigned long long* __restrict__ nonphys) {
  using I = PrimIdx<ND>;
  constexpr int NV = ND + 2;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  double u[NV], v[I::NPRIM];
#pragma unroll
  for (int k = 0; k < NV; ++k) u[k] = U[(size_t)i * NV + k];
#pragma unroll
  for (int k = 0; k < I::NPRIM; ++k) v[k] = V[(size_t)i * I::NPRIM + k];  // sound speed keeps old value if c2<0
  const double tke = tk ? tk[(size_t)2 * i] : 0.0;  // CNSSolver.cpp:135-158: tkeNeeded = SST
  bool bad = prim_from_cons<ND>(pp, u, v, tke);
  if (bad) {
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      u[k] = U_old[(size_t)i * NV + k];
      U[(size_t)i * NV + k] = u[k];
    }
    prim_from_cons<ND>(pp, u, v, tke);
    atomicAdd(nonphys, 1ull);
  }
  v[I::H] = (u[NV - 1] + v[I::P]) / u[0];
  if (pp.viscous) {
    double mu;
    if (pp.suth >= 0.0) {
      const double t_ref_inv = 1.0 / pp.t_ref;
      const double t_nondim = t_ref_inv * v[I::T];
      mu = pp.mu_ref * t_nondim * sqrt(t_nondim) * ((pp.t_ref + pp.suth) / (v[I::T] + pp.suth));
    } else {
      mu = pp.mu_const;
    }
    v[I::MU] = mu;
    v[I::MUT] = mu_t ? mu_t[i] : 0.0;
    v[I::K] = mu * pp.cp / pp.pr_lam;
    v[I::CP] = pp.cp;
  }
#pragma unroll
  for (int k = 0; k < I::NPRIM; ++k) V[(size_t)i * I::NPRIM + k] = v[k];
}

// =============================================================================================
// a6. Green-Gauss gradient: computeGradientsGreenGauss (computeGradientsGreenGauss.hpp:62-183).
//     Interior part: point loop over the neighbour CSR in the reference's order.
//     Fields 0..NG-1 of the primitive array; gradient layout [P][NGS][ND] (NGS = stored stride).
// =============================================================================================
template <int ND, int NG>
__global__ void k_grad_gg(int Pd, int nprim, int ngs, const int* __restrict__ nbr_ptr, const int* __restrict__ nbr_pt,
                          const int* __restrict__ nbr_edge, const double* __restrict__ enormal, int E,
                          const double* __restrict__ volume, const double* __restrict__ V,
                          double* __restrict__ grad) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Pd) return;
  double g[NG * ND];
#pragma unroll
  for (int k = 0; k < NG * ND; ++k) g[k] = 0.0;
  double fi[NG];
#pragma unroll
  for (int v = 0; v < NG; ++v) fi[v] = V[(size_t)i * nprim + v];
  const double halfOnVol = 0.5 / volume[i];
  for (int k = nbr_ptr[i]; k < nbr_ptr[i + 1]; ++k) {
    const int j = nbr_pt[k];
    const int e = nbr_edge[k];
    const double dir = (i < j) ? 1.0 : -1.0;
    const double weight = dir * halfOnVol;
    double area[ND];
#pragma unroll
    for (int d = 0; d < ND; ++d) area[d] = enormal[(size_t)d * E + e];
#pragma unroll
    for (int v = 0; v < NG; ++v) {
      const double flux = weight * (fi[v] + V[(size_t)j * nprim + v]);
#pragma unroll
      for (int d = 0; d < ND; ++d) g[v * ND + d] += flux * area[d];
    }
  }
#pragma unroll
  for (int v = 0; v < NG; ++v)
#pragma unroll
    for (int d = 0; d < ND; ++d) grad[((size_t)i * ngs + v) * ND + d] = g[v * ND + d];
}

// The same sums with NG lanes per point (the mapping of k_spmv_rows / k_point_sum_rows): lane v of a point's group
// owns variable v, so the NG lanes read NG consecutive doubles of every neighbour record instead of each thread
// walking its own records (32 lines per load instruction). Same operations in the same order per (variable, dimension).
// Measured on 185^3 RANS: the gradient calls of a step 2.26 -> 1.88 ms.
#ifndef GRAD_GG_ROWS
#define GRAD_GG_ROWS 1
#endif
template <int ND, int NG>
__global__ void __launch_bounds__(128) k_grad_gg_rows(int Pd, int nprim, int ngs, const int* __restrict__ nbr_ptr,
                                                      const int* __restrict__ nbr_pt, const int* __restrict__ nbr_edge,
                                                      const double* __restrict__ enormal, int E,
                                                      const double* __restrict__ volume, const double* __restrict__ V,
                                                      double* __restrict__ grad) {
  constexpr int RPW = 32 / NG;
  const int lane = threadIdx.x & 31, sub = lane / NG, v = lane - sub * NG;
  const int i = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * RPW + sub;
  if (sub >= RPW || i >= Pd) return;
  double g[ND];
#pragma unroll
  for (int d = 0; d < ND; ++d) g[d] = 0.0;
  const double fi = V[(size_t)i * nprim + v];
  const double halfOnVol = 0.5 / volume[i];
  const int kend = nbr_ptr[i + 1];
#pragma unroll 2
  for (int k = nbr_ptr[i]; k < kend; ++k) {
    const int j = nbr_pt[k];
    const int e = nbr_edge[k];
    const double dir = (i < j) ? 1.0 : -1.0;
    const double weight = dir * halfOnVol;
    const double flux = weight * (fi + V[(size_t)j * nprim + v]);
#pragma unroll
    for (int d = 0; d < ND; ++d) g[d] += flux * enormal[(size_t)d * E + e];
  }
#pragma unroll
  for (int d = 0; d < ND; ++d) grad[((size_t)i * ngs + v) * ND + d] = g[d];
}

// boundary part (computeGradientsGreenGauss.hpp:131-160): one launch per marker, a point appears
// at most once per marker so the read-modify-write is race free; markers run in reference order.
template <int ND, int NG>
__global__ void k_grad_gg_bnd(int nV, int Pd, int nprim, int ngs, const int* __restrict__ vnode,
                              const double* __restrict__ vnormal, const double* __restrict__ volume,
                              const double* __restrict__ V, double* __restrict__ grad) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nV) return;
  const int i = vnode[t];
  if (i >= Pd) return;
  const double vol = volume[i];
#pragma unroll
  for (int v = 0; v < NG; ++v) {
    const double flux = V[(size_t)i * nprim + v] / vol;
#pragma unroll
    for (int d = 0; d < ND; ++d) grad[((size_t)i * ngs + v) * ND + d] -= flux * vnormal[(size_t)t * ND + d];
  }
}

// correctGradientsSymmetry (correctGradientsSymmetry.hpp:49-146), idxVel = 1, one launch per sym marker.
template <int ND, int NG>
__global__ void k_grad_sym(int nV, int ngs, const int* __restrict__ vnode, const double* __restrict__ unit,
                           double* __restrict__ grad) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nV) return;
  const int i = vnode[t];
  double n[ND];
#pragma unroll
  for (int d = 0; d < ND; ++d) n[d] = unit[(size_t)t * ND + d];
  double* gp = grad + (size_t)i * ngs * ND;
  double g[NG * ND];
#pragma unroll
  for (int k = 0; k < NG * ND; ++k) g[k] = gp[k];
  constexpr int iv = 1;
  {
    double normalGrad[ND], gradNormalVel[ND];
#pragma unroll
    for (int d = 0; d < ND; ++d) gradNormalVel[d] = 0.0;
#pragma unroll
    for (int a = 0; a < ND; ++a) {
      normalGrad[a] = dotN<ND>(&g[(iv + a) * ND], n);
#pragma unroll
      for (int b = 0; b < ND; ++b) gradNormalVel[b] += n[a] * g[(iv + a) * ND + b];
    }
    const double nGnV = dotN<ND>(n, gradNormalVel);
#pragma unroll
    for (int a = 0; a < ND; ++a)
#pragma unroll
      for (int b = 0; b < ND; ++b) {
        g[(iv + a) * ND + b] -= normalGrad[a] * n[b] + n[a] * gradNormalVel[b];
        g[(iv + a) * ND + b] += 2 * n[a] * nGnV * n[b];
      }
  }
#pragma unroll
  for (int v = 0; v < NG; ++v) {
    if (v >= iv && v < iv + ND) continue;
    const double ng = dotN<ND>(n, &g[v * ND]);
#pragma unroll
    for (int d = 0; d < ND; ++d) g[v * ND + d] -= ng * n[d];
  }
#pragma unroll
  for (int k = 0; k < NG * ND; ++k) gp[k] = g[k];
}

// =============================================================================================
// a7. (Weighted) least squares: computeGradientsLeastSquares (computeGradientsLeastSquares.hpp:185-325)
//     + solveLeastSquares (:76-164).
// =============================================================================================
template <int ND, int NG>
__global__ void k_grad_ls(int Pd, int nprim, int ngs, int weighted, const int* __restrict__ nbr_ptr,
                          const int* __restrict__ nbr_pt, const double* __restrict__ coord,
                          const double* __restrict__ V, double* __restrict__ grad) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Pd) return;
  double ci[ND];
#pragma unroll
  for (int d = 0; d < ND; ++d) ci[d] = coord[(size_t)i * ND + d];
  double fi[NG];
#pragma unroll
  for (int v = 0; v < NG; ++v) fi[v] = V[(size_t)i * nprim + v];
  double g[NG * ND];
#pragma unroll
  for (int k = 0; k < NG * ND; ++k) g[k] = 0.0;
  double Rm[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
  for (int k = nbr_ptr[i]; k < nbr_ptr[i + 1]; ++k) {
    const int j = nbr_pt[k];
    double dij[ND];
#pragma unroll
    for (int d = 0; d < ND; ++d) dij[d] = coord[(size_t)j * ND + d] - ci[d];
    double weight = 1.0;
    if (weighted) weight = dotN<ND>(dij, dij);
    if (weight > 0.0) {
      weight = 1.0 / weight;
#pragma unroll
      for (int a = 0; a < ND; ++a)
#pragma unroll
        for (int b = a; b < ND; ++b) Rm[a][b] += dij[a] * dij[b] * weight;
      if (ND == 3) Rm[2][1] += dij[0] * dij[ND - 1] * weight;
#pragma unroll
      for (int v = 0; v < NG; ++v) {
        const double delta = weight * (V[(size_t)j * nprim + v] - fi[v]);
#pragma unroll
        for (int d = 0; d < ND; ++d) g[v * ND + d] += dij[d] * delta;
      }
    }
  }
  // solveLeastSquares
  const double eps = 2.220446049250313e-16 * 2.220446049250313e-16;
  double r11 = Rm[0][0], r12 = Rm[0][1], r22 = Rm[1][1];
  double r13 = 0.0, r23 = 0.0, r33 = 1.0;
  r11 = sqrt(fmax(r11, eps));
  r12 /= r11;
  r22 = sqrt(fmax(r22 - r12 * r12, eps));
  if (ND == 3) {
    r13 = Rm[0][2];
    r33 = Rm[2][2];
    const double r23_a = Rm[1][2];
    const double r23_b = Rm[2][1];
    r13 /= r11;
    r23 = r23_a / r22 - r23_b * r12 / (r11 * r22);
    r33 = sqrt(fmax(r33 - r23 * r23 - r13 * r13, eps));
  }
  const double det = r11 * r22 * r33;
  const double detR2 = det * det;
  double S[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
  if (detR2 > eps) {
    if (ND == 2) {
      S[0][0] = (r12 * r12 + r22 * r22) / detR2;
      S[0][1] = -r11 * r12 / detR2;
      S[1][1] = r11 * r11 / detR2;
    } else {
      const double z11 = r22 * r33, z12 = -r12 * r33, z13 = r12 * r23 - r13 * r22;
      const double z22 = r11 * r33, z23 = -r11 * r23, z33 = r11 * r22;
      S[0][0] = (z11 * z11 + z12 * z12 + z13 * z13) / detR2;
      S[0][1] = (z12 * z22 + z13 * z23) / detR2;
      S[0][2] = (z13 * z33) / detR2;
      S[1][1] = (z22 * z22 + z23 * z23) / detR2;
      S[1][2] = (z23 * z33) / detR2;
      S[2][2] = (z33 * z33) / detR2;
    }
  }
#pragma unroll
  for (int v = 0; v < NG; ++v) {
    double c[ND];
#pragma unroll
    for (int a = 0; a < ND; ++a) {
      c[a] = 0.0;
#pragma unroll
      for (int b = 0; b < ND; ++b) c[a] += S[a < b ? a : b][a < b ? b : a] * g[v * ND + b];
    }
#pragma unroll
    for (int d = 0; d < ND; ++d) grad[((size_t)i * ngs + v) * ND + d] = c[d];
  }
}

// =============================================================================================
// a8. Venkatakrishnan / Barth-Jespersen limiter: computeLimiters_impl (computeLimiters_impl.hpp:65-247),
//     CLimiterDetails (CLimiterDetails.hpp:136-207), venkatFunction (:91-95).
// =============================================================================================
// (One thread per point stays: the NG-lanes-per-point mapping that pays for the gradient, k_grad_gg_rows, was measured
//  slower here -- 0.89 against 0.71 ms on 185^3 -- every lane repeats the coordinate differences and the point's
//  15 gradient entries are read once either way.)
template <int ND, int NG>
__global__ void k_limiter(int Pd, int nprim, int ngs, double eps2, double kappa, const int* __restrict__ nbr_ptr,
                          const int* __restrict__ nbr_pt, const double* __restrict__ coord,
                          const double* __restrict__ V, const double* __restrict__ grad,
                          double* __restrict__ limiter) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Pd) return;
  double ci[ND], fi[NG], fmx[NG], fmn[NG], pmx[NG], pmn[NG], g[NG * ND];
#pragma unroll
  for (int d = 0; d < ND; ++d) ci[d] = coord[(size_t)i * ND + d];
#pragma unroll
  for (int v = 0; v < NG; ++v) {
    fi[v] = V[(size_t)i * nprim + v];
    fmx[v] = fmn[v] = fi[v];
    pmx[v] = pmn[v] = 0.0;
#pragma unroll
    for (int d = 0; d < ND; ++d) g[v * ND + d] = grad[((size_t)i * ngs + v) * ND + d];
  }
  for (int k = nbr_ptr[i]; k < nbr_ptr[i + 1]; ++k) {
    const int j = nbr_pt[k];
    double dij[ND];
#pragma unroll
    for (int d = 0; d < ND; ++d) dij[d] = 0.5 * (coord[(size_t)j * ND + d] - ci[d]);
#pragma unroll
    for (int v = 0; v < NG; ++v) {
      double proj = 0.0;
#pragma unroll
      for (int d = 0; d < ND; ++d) proj += dij[d] * g[v * ND + d];
      const double fj = V[(size_t)j * nprim + v];
      const double cent0 = 0.5 * (fj - fi[v]);
      // LimiterHelpers::umusclProjection(proj, cent0, kappa)
      const double cent = 0.5 * cent0;
      const double upw = proj - cent;
      proj = (1.0 - kappa) * upw + (1.0 + kappa) * cent;
      pmx[v] = fmax(pmx[v], proj);
      pmn[v] = fmin(pmn[v], proj);
      fmx[v] = fmax(fmx[v], fj);
      fmn[v] = fmin(fmn[v], fj);
    }
  }
#pragma unroll
  for (int v = 0; v < NG; ++v) {
    double delta = fmx[v] - fi[v];
    double y = delta * (delta + pmx[v]) + eps2;
    const double limMax = (y + delta * pmx[v]) / (y + 2 * pmx[v] * pmx[v]);
    delta = fmn[v] - fi[v];
    y = delta * (delta + pmn[v]) + eps2;
    const double limMin = (y + delta * pmn[v]) / (y + 2 * pmn[v] * pmn[v]);
    limiter[(size_t)i * NG + v] = fmin(limMax, limMin);
  }
}

// =============================================================================================
// a9. Local time step: CFVMFlowSolverBase::SetTime_Step_impl (CFVMFlowSolverBase.hpp:377-620) with the
//     functors of CEulerSolver::SetTime_Step (CEulerSolver.cpp:1727-1779). Also used for JST's
//     SetMax_Eigenvalue_impl (:628-716) with mode==1 (Area instead of sqrt(Area2), no viscous part).
// =============================================================================================
template <int ND>
__device__ __forceinline__ double lambda_visc(const double* Vi, const double* Vj, double gamma, double prt) {
  using I = PrimIdx<ND>;
  const double mu = 0.5 * (Vi[I::MU] + Vj[I::MU]);
  const double mut = 0.5 * (Vi[I::MUT] + Vj[I::MUT]);
  const double rho = 0.5 * (Vi[I::RHO] + Vj[I::RHO]);
  const double k = 0.5 * (Vi[I::K] + Vj[I::K]);
  const double cp = 0.5 * (Vi[I::CP] + Vj[I::CP]);
  const double L1 = (4.0 / 3.0) * (mu + mut);
  const double L2 = (k / cp + mut / prt) * gamma;
  return (L1 + L2) / rho;
}

template <int ND>
__global__ void k_lambda_points(int Pd, int viscous, double gamma, double prt, const int* __restrict__ nbr_ptr,
                                const int* __restrict__ nbr_pt, const int* __restrict__ nbr_edge,
                                const double* __restrict__ enormal, int E, const double* __restrict__ V,
                                double* __restrict__ lam_inv, double* __restrict__ lam_visc) {
  using I = PrimIdx<ND>;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Pd) return;
  const double* Vi = V + (size_t)i * I::NPRIM;
  double li = 0.0, lv = 0.0;
  for (int k = nbr_ptr[i]; k < nbr_ptr[i + 1]; ++k) {
    const int j = nbr_pt[k];
    const int e = nbr_edge[k];
    const double* Vj = V + (size_t)j * I::NPRIM;
    double n[ND];
#pragma unroll
    for (int d = 0; d < ND; ++d) n[d] = enormal[(size_t)d * E + e];
    const double area2 = dotN<ND>(n, n);
    const double pv = 0.5 * (dotN<ND>(&Vi[I::VEL], n) + dotN<ND>(&Vj[I::VEL], n));
    const double c = 0.5 * (Vi[I::C] + Vj[I::C]) * sqrt(area2);
    li += fabs(pv) + c;
    if (viscous) lv += lambda_visc<ND>(Vi, Vj, gamma, prt) * area2;
  }
  lam_inv[i] = li;
  if (viscous) lam_visc[i] = lv;
}

template <int ND>
__global__ void k_lambda_bnd(int nV, int Pd, int viscous, double gamma, double prt, const int* __restrict__ vnode,
                             const double* __restrict__ vnormal, const double* __restrict__ V,
                             double* __restrict__ lam_inv, double* __restrict__ lam_visc) {
  using I = PrimIdx<ND>;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nV) return;
  const int i = vnode[t];
  if (i >= Pd) return;
  const double* Vi = V + (size_t)i * I::NPRIM;
  double n[ND];
#pragma unroll
  for (int d = 0; d < ND; ++d) n[d] = vnormal[(size_t)t * ND + d];
  const double area2 = dotN<ND>(n, n);
  const double pv = dotN<ND>(&Vi[I::VEL], n);
  const double c = Vi[I::C] * sqrt(area2);
  lam_inv[i] += fabs(pv) + c;
  if (viscous) lam_visc[i] += lambda_visc<ND>(Vi, Vi, gamma, prt) * area2;
}

// final dt + min/max block reduction (CFVMFlowSolverBase.hpp:497-532); out[0]=min, out[1]=max per block
__global__ void k_delta_time(int Pd, int viscous, double cfl_all, const double* __restrict__ local_cfl, double max_dt,
                             const double* __restrict__ volume, const double* __restrict__ lam_inv,
                             const double* __restrict__ lam_visc, double* __restrict__ dt, double* __restrict__ blk_minmax) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  double mn = 1e30, mx = 0.0;
  if (i < Pd) {
    const double vol = volume[i];
    const double cfl = local_cfl ? local_cfl[i] : cfl_all;  // CVariable::LocalCFL
    if (vol != 0.0) {
      double ldt = cfl * vol / lam_inv[i];
      if (viscous) {
        const double dtv = cfl * 0.25 * vol * vol / lam_visc[i];
        ldt = fmin(ldt, dtv);
      }
      mn = ldt;
      mx = ldt;
      dt[i] = fmin(ldt, max_dt);
    } else {
      dt[i] = 0.0;
    }
  }
  __shared__ double smn[32], smx[32];
  for (int o = 16; o > 0; o >>= 1) {
    mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) { smn[w] = mn; smx[w] = mx; }
  __syncthreads();
  if (w == 0) {
    const int nw = blockDim.x >> 5;
    mn = (l < nw) ? smn[l] : 1e30;
    mx = (l < nw) ? smx[l] : 0.0;
    for (int o = 16; o > 0; o >>= 1) {
      mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
      mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if (l == 0) { blk_minmax[2 * blockIdx.x] = mn; blk_minmax[2 * blockIdx.x + 1] = mx; }
  }
}

__global__ void k_minmax_final(int n, const double* __restrict__ blk, double* __restrict__ out) {
  double mn = 1e30, mx = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) { mn = fmin(mn, blk[2 * i]); mx = fmax(mx, blk[2 * i + 1]); }
  __shared__ double smn[32], smx[32];
  for (int o = 16; o > 0; o >>= 1) {
    mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) { smn[w] = mn; smx[w] = mx; }
  __syncthreads();
  if (threadIdx.x == 0) {
    const int nw = blockDim.x >> 5;
    for (int k = 1; k < nw; ++k) { mn = fmin(mn, smn[k]); mx = fmax(mx, smx[k]); }
    out[0] = mn; out[1] = mx;
  }
}

// =============================================================================================
// a10. JST support: SetCentered_Dissipation_Sensor_impl (CFVMFlowSolverBase.hpp:729-783) and
//      CEulerSolver::SetUndivided_Laplacian (CEulerSolver.cpp:2342-2387), fused in one neighbour loop.
// =============================================================================================
template <int ND>
__global__ void k_jst_sensor_lapl(int Pd, const int* __restrict__ nbr_ptr, const int* __restrict__ nbr_pt,
                                  const unsigned char* __restrict__ phys_bnd, const double* __restrict__ U,
                                  const double* __restrict__ V, double* __restrict__ sensor,
                                  double* __restrict__ lapl) {
  using I = PrimIdx<ND>;
  constexpr int NV = ND + 2;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Pd) return;
  const bool bi = phys_bnd[i];
  const double pi = V[(size_t)i * I::NPRIM + I::P];
  double ui[NV], L[NV];
#pragma unroll
  for (int k = 0; k < NV; ++k) { ui[k] = U[(size_t)i * NV + k]; L[k] = 0.0; }
  double sdiff = 0.0, ssum = 0.0;
  for (int k = nbr_ptr[i]; k < nbr_ptr[i + 1]; ++k) {
    const int j = nbr_pt[k];
    const bool bj = phys_bnd[j];
    if (bi && !bj) continue;
    const double pj = V[(size_t)j * I::NPRIM + I::P];
    sdiff += pj - pi;
    ssum += pj + pi;
#pragma unroll
    for (int v = 0; v < NV; ++v) L[v] += U[(size_t)j * NV + v] - ui[v];
    L[NV - 1] += pj - pi;
  }
  sensor[i] = fabs(sdiff) / ssum;
#pragma unroll
  for (int v = 0; v < NV; ++v) lapl[(size_t)i * NV + v] = L[v];
}

// =============================================================================================
// ComputeVorticityAndStrainMag (CFVMFlowSolverBase.inl:620-703)
// =============================================================================================
template <int ND>
__global__ void k_vorticity_strain(int P, int ngs, const double* __restrict__ grad, double* __restrict__ vort,
                                   double* __restrict__ strain) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P) return;
  const double* g = grad + (size_t)i * ngs * ND;  // g[(1+a)*ND+b] = d v_a / d x_b
  double w[3] = {0, 0, 0};
  w[2] = g[(1 + 1) * ND + 0] - g[(1 + 0) * ND + 1];
  if (ND == 3) {
    w[0] = g[(1 + ND - 1) * ND + 1] - g[(1 + 1) * ND + ND - 1];
    w[1] = -(g[(1 + ND - 1) * ND + 0] - g[(1 + 0) * ND + ND - 1]);
  }
  vort[(size_t)i * 3 + 0] = w[0];
  vort[(size_t)i * 3 + 1] = w[1];
  vort[(size_t)i * 3 + 2] = w[2];
  double s = 0.0;
  // diagonal part
#pragma unroll
  for (int d = 0; d < ND; ++d) { const double t = g[(1 + d) * ND + d]; s += t * t; }
  // off-diagonal
  { const double t = g[(1 + 0) * ND + 1] + g[(1 + 1) * ND + 0]; s += 2.0 * (0.5 * t) * (0.5 * t); }
  if (ND == 3) {
    { const double t = g[(1 + 0) * ND + ND - 1] + g[(1 + ND - 1) * ND + 0]; s += 2.0 * (0.5 * t) * (0.5 * t); }
    { const double t = g[(1 + 1) * ND + ND - 1] + g[(1 + ND - 1) * ND + 1]; s += 2.0 * (0.5 * t) * (0.5 * t); }
  }
  strain[i] = sqrt(2.0 * s);
}

// =============================================================================================
// a15. PrepareImplicitIteration_impl (CFVMFlowSolverBase.hpp:912-973) + ResidualReductions (CSolver.hpp:4403-4430).
//      Generic in NV (flow 4/5, SA 1). Block partial sums go to `part` ([nblk][NV] sum r^2, [nblk][NV] max, idx).
// =============================================================================================
template <int NV>
__global__ void k_prepare_implicit(int Pd, const int* __restrict__ dia_ptr, const double* __restrict__ volume,
                                   const double* __restrict__ dt, double* __restrict__ A, double* __restrict__ res,
                                   const double* __restrict__ trunc, double* __restrict__ sol,
                                   double* __restrict__ part_rms, double* __restrict__ part_max,
                                   int* __restrict__ part_idx) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  double r2[NV], rm[NV];
  int ri[NV];
#pragma unroll
  for (int v = 0; v < NV; ++v) { r2[v] = 0.0; rm[v] = 0.0; ri[v] = 0; }
  if (i < Pd) {
    double* D = A + (size_t)dia_ptr[i] * NV * NV;
    const double dti = dt[i];
    if (dti != 0.0) {
      const double delta = volume[i] / dti;
#pragma unroll
      for (int v = 0; v < NV; ++v) D[v * NV + v] += delta;
    } else {
#pragma unroll
      for (int k = 0; k < NV * NV; ++k) D[k] = 0.0;
#pragma unroll
      for (int v = 0; v < NV; ++v) D[v * NV + v] = 1.0;
    }
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      double r = res[(size_t)i * NV + v];
      double te = trunc ? trunc[(size_t)i * NV + v] : 0.0;
      if (dti == 0.0) { r = 0.0; te = 0.0; }
      r = -(r + te);
      res[(size_t)i * NV + v] = r;
      sol[(size_t)i * NV + v] = 0.0;
      const double a = fabs(r);
      r2[v] = a * a;
      rm[v] = a;
      ri[v] = i;
    }
  }
  // block reduction (deterministic tree)
  __shared__ double s2[NV][32], sm[NV][32];
  __shared__ int si[NV][32];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
#pragma unroll
  for (int v = 0; v < NV; ++v) {
    double a = r2[v], m = rm[v];
    int id = ri[v];
    for (int o = 16; o > 0; o >>= 1) {
      a += __shfl_xor_sync(0xffffffffu, a, o);
      const double m2 = __shfl_xor_sync(0xffffffffu, m, o);
      const int id2 = __shfl_xor_sync(0xffffffffu, id, o);
      if (m2 > m || (m2 == m && id2 < id)) { m = m2; id = id2; }
    }
    if (l == 0) { s2[v][w] = a; sm[v][w] = m; si[v][w] = id; }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const int nw = blockDim.x >> 5;
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      double a = 0.0, m = 0.0;
      int id = 0;
      for (int k = 0; k < nw; ++k) {
        a += s2[v][k];
        if (sm[v][k] > m || (sm[v][k] == m && si[v][k] < id)) { m = sm[v][k]; id = si[v][k]; }
      }
      part_rms[(size_t)blockIdx.x * NV + v] = a;
      part_max[(size_t)blockIdx.x * NV + v] = m;
      part_idx[(size_t)blockIdx.x * NV + v] = id;
    }
  }
}

template <int NV>
__global__ void k_residual_final(int nb, const double* __restrict__ part_rms, const double* __restrict__ part_max,
                                 const int* __restrict__ part_idx, double* __restrict__ out /*[3*NV]*/) {
  // one block of 256 threads per variable: thread t reduces a contiguous chunk of the block partials in order, thread 0
  // then combines the chunks in ascending order (fixed summation order whatever the launch; first maximum wins)
  __shared__ double sa[256], sm[256];
  __shared__ int si[256];
  const int t = threadIdx.x, v = blockIdx.x;
  const int per = (nb + 255) / 256;
  const int b0 = min(t * per, nb), b1 = min(b0 + per, nb);
  double a = 0.0, m = 0.0;
  int id = 0;
  for (int b = b0; b < b1; ++b) {
    a += part_rms[(size_t)b * NV + v];
    const double mb = part_max[(size_t)b * NV + v];
    if (mb > m) { m = mb; id = part_idx[(size_t)b * NV + v]; }
  }
  sa[t] = a; sm[t] = m; si[t] = id;
  __syncthreads();
  if (t == 0) {
    a = 0.0; m = 0.0; id = 0;
    for (int k = 0; k < 256; ++k) {
      a += sa[k];
      if (sm[k] > m) { m = sm[k]; id = si[k]; }
    }
    out[v] = a;
    out[NV + v] = m;
    out[2 * NV + v] = (double)id;
  }
}

// zero halo entries of rhs and solution (CFVMFlowSolverBase.inl:601-606)
__global__ void k_zero_range(double* __restrict__ a, double* __restrict__ b, size_t begin, size_t end) {
  const size_t i = begin + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < end) { a[i] = 0.0; b[i] = 0.0; }
}

// =============================================================================================
// a24. ComputeUnderRelaxationFactor (CFVMFlowSolverBase.inl:558-592) + CompleteImplicitIteration_impl
//      (CFVMFlowSolverBase.hpp:983-1008)
// =============================================================================================
template <int NV>
__global__ void k_complete_implicit(int Pd, double allowable, const double* __restrict__ sol,
                                    const double* __restrict__ U_old, double* __restrict__ U,
                                    double* __restrict__ under_relax) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Pd) return;
  double ur = 1.0;
  double dx[NV], u[NV];
#pragma unroll
  for (int v = 0; v < NV; ++v) { dx[v] = sol[(size_t)i * NV + v]; u[v] = U[(size_t)i * NV + v]; }
#pragma unroll
  for (int v = 0; v < NV; ++v) {
    if (v == 0 || v == NV - 1) {
      const double ratio = fabs(dx[v]) / (fabs(u[v]) + 1.0e-16);
      if (ratio > allowable) ur = fmin(allowable / ratio, ur);
    }
  }
  if (ur < 1e-10) ur = 0.0;
  under_relax[i] = ur;
#pragma unroll
  // CVariable::AddSolution (CVariable.hpp:341-343): Solution = Solution_Old + value (Solution_Old carries the
  // strong-BC corrections made by BC_Sym_Plane / BC_HeatFlux_Wall)
  for (int v = 0; v < NV; ++v) U[(size_t)i * NV + v] = U_old[(size_t)i * NV + v] + ur * dx[v];
}

// =============================================================================================
// Host wrappers
// =============================================================================================
#define LAUNCH(ctx, kern, grid, block, ...)                     \
  do {                                                          \
    if ((grid) > 0) {                                           \
      kern<<<(grid), (block), 0, (ctx)->stream>>>(__VA_ARGS__); \
      (ctx)->launches++;                                        \
      SU2B_CUDA(cudaGetLastError());                            \
    }                                                           \
  } while (0)

static PrimParams make_prim_params(const su2b200_flow_config& c) {
  PrimParams pp;
  pp.gamma = c.gamma;
  pp.gm1 = c.gamma - 1.0;
  pp.R = c.gas_constant;
  pp.cp = c.gamma / (c.gamma - 1.0) * c.gas_constant;
  pp.viscous = c.viscous;
  pp.mu_ref = c.mu_ref;
  pp.t_ref = c.mu_temperature_ref;
  pp.suth = c.mu_sutherland;
  pp.mu_const = c.mu_constant;
  pp.pr_lam = c.prandtl_lam;
  return pp;
}

void flow_primitives(su2b200_ctx* c) {
  ProfScope prof_(c, PROF_PRIMITIVES);
  auto& g = c->geo;
  auto& s = c->fs;
  const PrimParams pp = make_prim_params(c->cfg);
  const double* mut = (c->cfg.viscous && (c->cfg.rans_sa || c->cfg.rans_sst)) ? s.mu_t.p : nullptr;
  const double* tk = (c->cfg.viscous && c->cfg.rans_sst) ? s.nu.p : nullptr;
  SU2B_CUDA(cudaMemsetAsync(s.counters.p, 0, sizeof(unsigned long long), c->stream));
  if (g.nDim == 2) LAUNCH(c, k_primitives<2>, nblk(g.P, 256), 256, g.P, pp, s.U.p, s.U_old.p, mut, tk, s.V.p, s.counters.p);
  else LAUNCH(c, k_primitives<3>, nblk(g.P, 256), 256, g.P, pp, s.U.p, s.U_old.p, mut, tk, s.V.p, s.counters.p);
}

// field = primitives (first NG = nDim+2 or nDim+1 variables); `into` selects the gradient array.
template <int ND, int NG>
static void gradient_impl(su2b200_ctx* c, int method, double* grad, int ngs) {
  auto& g = c->geo;
  auto& s = c->fs;
  const int nprim = s.nPrim;
  if (method == SU2B200_GRAD_GREEN_GAUSS) {
#if GRAD_GG_ROWS
    LAUNCH(c, (k_grad_gg_rows<ND, NG>), nblk(g.Pd, 4 * (32 / NG)), 128, g.Pd, nprim, ngs, g.nbr_ptr.p, g.nbr_pt.p,
           g.nbr_edge.p, g.edge_normal.p, g.E, g.volume.p, s.V.p, grad);
#else
    LAUNCH(c, (k_grad_gg<ND, NG>), nblk(g.Pd, 128), 128, g.Pd, nprim, ngs, g.nbr_ptr.p, g.nbr_pt.p, g.nbr_edge.p,
           g.edge_normal.p, g.E, g.volume.p, s.V.p, grad);
#endif
    for (auto& m : g.markers) {
      if (m.kind == SU2B200_SEND_RECEIVE || m.nVertex == 0) continue;
      LAUNCH(c, (k_grad_gg_bnd<ND, NG>), nblk(m.nVertex, 128), 128, m.nVertex, g.Pd, nprim, ngs, m.node.p,
             m.normal.p, g.volume.p, s.V.p, grad);
    }
  } else {
    const int weighted = (method == SU2B200_GRAD_WEIGHTED_LEAST_SQUARES);
    LAUNCH(c, (k_grad_ls<ND, NG>), nblk(g.Pd, 128), 128, g.Pd, nprim, ngs, weighted, g.nbr_ptr.p, g.nbr_pt.p,
           g.coord.p, s.V.p, grad);
  }
  for (auto& m : g.markers) {
    if ((m.kind == SU2B200_SYMMETRY_PLANE || m.kind == SU2B200_EULER_WALL) && m.nVertex > 0)
      LAUNCH(c, (k_grad_sym<ND, NG>), nblk(m.nVertex, 128), 128, m.nVertex, ngs, m.node.p, m.unit_sym.p, grad);
  }
}

// reconstruction != 0 -> Gradient_Reconstruction with NUM_METHOD_GRAD_RECON, else Gradient_Primitive.
// nvars: nDim+2 (all reconstructed variables) or nDim+1 (Output pass of CNSSolver::Preprocessing).
void flow_gradient(su2b200_ctx* c, int reconstruction, int nvars) {
  ProfScope prof_(c, PROF_GRADIENT);
  auto& s = c->fs;
  const int nd = c->geo.nDim;
  const int method = reconstruction ? c->cfg.grad_method_recon : c->cfg.grad_method;
  double* grad = (reconstruction && !s.recon_aliases_prim) ? s.grad_recon.p : s.grad_prim.p;
  const int ngs = s.nG;
  if (nd == 2) {
    if (nvars == 4) gradient_impl<2, 4>(c, method, grad, ngs); else gradient_impl<2, 3>(c, method, grad, ngs);
  } else {
    if (nvars == 5) gradient_impl<3, 5>(c, method, grad, ngs); else gradient_impl<3, 4>(c, method, grad, ngs);
  }
}

// scalar variant of the symmetry correction (idxVel = -1: only n.grad(phi) = 0), one launch per sym marker
template <int ND>
__global__ void k_grad_sym_scalar(int nV, const int* __restrict__ vnode, const double* __restrict__ unit,
                                  double* __restrict__ grad) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nV) return;
  const int i = vnode[t];
  double ng = 0.0;
#pragma unroll
  for (int d = 0; d < ND; ++d) ng += unit[(size_t)t * ND + d] * grad[(size_t)i * ND + d];
#pragma unroll
  for (int d = 0; d < ND; ++d) grad[(size_t)i * ND + d] -= ng * unit[(size_t)t * ND + d];
}

// CScalarSolver::SetSolution_Gradient_GG / _LS (CScalarSolver.inl:88-127 via CSolver::SetSolution_Gradient_*):
// gradient of a one-variable field [P] -> [P][ND], idxVel = -1.
template <int ND>
static void scalar_gradient_impl(su2b200_ctx* c, int method, const double* field, double* grad) {
  auto& g = c->geo;
  if (method == SU2B200_GRAD_GREEN_GAUSS) {
    LAUNCH(c, (k_grad_gg<ND, 1>), nblk(g.Pd, 128), 128, g.Pd, 1, 1, g.nbr_ptr.p, g.nbr_pt.p, g.nbr_edge.p,
           g.edge_normal.p, g.E, g.volume.p, field, grad);
    for (auto& m : g.markers) {
      if (m.kind == SU2B200_SEND_RECEIVE || m.nVertex == 0) continue;
      LAUNCH(c, (k_grad_gg_bnd<ND, 1>), nblk(m.nVertex, 128), 128, m.nVertex, g.Pd, 1, 1, m.node.p, m.normal.p,
             g.volume.p, field, grad);
    }
  } else {
    const int weighted = (method == SU2B200_GRAD_WEIGHTED_LEAST_SQUARES);
    LAUNCH(c, (k_grad_ls<ND, 1>), nblk(g.Pd, 128), 128, g.Pd, 1, 1, weighted, g.nbr_ptr.p, g.nbr_pt.p, g.coord.p, field,
           grad);
  }
  for (auto& m : g.markers)
    if ((m.kind == SU2B200_SYMMETRY_PLANE || m.kind == SU2B200_EULER_WALL) && m.nVertex > 0)
      LAUNCH(c, k_grad_sym_scalar<ND>, nblk(m.nVertex, 128), 128, m.nVertex, m.node.p, m.unit_sym.p, grad);
}
void scalar_gradient(su2b200_ctx* c, int method, const double* field, double* grad) {
  ProfScope prof_(c, PROF_GRADIENT);
  if (c->geo.nDim == 2) scalar_gradient_impl<2>(c, method, field, grad);
  else scalar_gradient_impl<3>(c, method, field, grad);
}

// the same for a two-variable field [P][2] -> [P][2][ND] (SST: k, omega)
template <int ND>
__global__ void k_grad_sym_scalar2(int nV, const int* __restrict__ vnode, const double* __restrict__ unit,
                                   double* __restrict__ grad) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nV) return;
  const int i = vnode[t];
#pragma unroll
  for (int v = 0; v < 2; ++v) {
    double* gv = grad + ((size_t)i * 2 + v) * ND;
    double ng = 0.0;
#pragma unroll
    for (int d = 0; d < ND; ++d) ng += unit[(size_t)t * ND + d] * gv[d];
#pragma unroll
    for (int d = 0; d < ND; ++d) gv[d] -= ng * unit[(size_t)t * ND + d];
  }
}
template <int ND>
static void scalar2_gradient_impl(su2b200_ctx* c, int method, const double* field, double* grad) {
  auto& g = c->geo;
  if (method == SU2B200_GRAD_GREEN_GAUSS) {
    LAUNCH(c, (k_grad_gg<ND, 2>), nblk(g.Pd, 128), 128, g.Pd, 2, 2, g.nbr_ptr.p, g.nbr_pt.p, g.nbr_edge.p,
           g.edge_normal.p, g.E, g.volume.p, field, grad);
    for (auto& m : g.markers) {
      if (m.kind == SU2B200_SEND_RECEIVE || m.nVertex == 0) continue;
      LAUNCH(c, (k_grad_gg_bnd<ND, 2>), nblk(m.nVertex, 128), 128, m.nVertex, g.Pd, 2, 2, m.node.p, m.normal.p,
             g.volume.p, field, grad);
    }
  } else {
    const int weighted = (method == SU2B200_GRAD_WEIGHTED_LEAST_SQUARES);
    LAUNCH(c, (k_grad_ls<ND, 2>), nblk(g.Pd, 128), 128, g.Pd, 2, 2, weighted, g.nbr_ptr.p, g.nbr_pt.p, g.coord.p, field,
           grad);
  }
  for (auto& m : g.markers)
    if ((m.kind == SU2B200_SYMMETRY_PLANE || m.kind == SU2B200_EULER_WALL) && m.nVertex > 0)
      LAUNCH(c, k_grad_sym_scalar2<ND>, nblk(m.nVertex, 128), 128, m.nVertex, m.node.p, m.unit_sym.p, grad);
}
void scalar2_gradient(su2b200_ctx* c, int method, const double* field, double* grad) {
  ProfScope prof_(c, PROF_GRADIENT);
  if (c->geo.nDim == 2) scalar2_gradient_impl<2>(c, method, field, grad);
  else scalar2_gradient_impl<3>(c, method, field, grad);
}

void flow_limiter(su2b200_ctx* c) {
  ProfScope prof_(c, PROF_LIMITER);
  auto& g = c->geo;
  auto& s = c->fs;
  const double macheps = 2.220446049250313e-16;
  double eps2 = macheps;
  if (c->cfg.limiter == SU2B200_LIMITER_VENKATAKRISHNAN) {
    const double eps1 = fabs(c->cfg.ref_elem_length * c->cfg.venkat_coeff);
    eps2 = fmax(eps1 * eps1 * eps1, macheps);
  }
  const double* grad = s.recon_aliases_prim ? s.grad_prim.p : s.grad_recon.p;
  if (g.nDim == 2)
    LAUNCH(c, (k_limiter<2, 4>), nblk(g.Pd, 128), 128, g.Pd, s.nPrim, s.nG, eps2, c->cfg.muscl_kappa, g.nbr_ptr.p,
           g.nbr_pt.p, g.coord.p, s.V.p, grad, s.limiter.p);
  else
    LAUNCH(c, (k_limiter<3, 5>), nblk(g.Pd, 128), 128, g.Pd, s.nPrim, s.nG, eps2, c->cfg.muscl_kappa, g.nbr_ptr.p,
           g.nbr_pt.p, g.coord.p, s.V.p, grad, s.limiter.p);
}

static bool marker_in_spectral_radius(int kind) {
  // CFVMFlowSolverBase.hpp:445-450: all but INTERNAL, SYMMETRY_PLANE, EULER_WALL, NEARFIELD, PERIODIC (and halos)
  return kind != SU2B200_SYMMETRY_PLANE && kind != SU2B200_EULER_WALL && kind != SU2B200_SEND_RECEIVE;
}

template <int ND>
static void spectral_radius_impl(su2b200_ctx* c, int viscous, double* lam_inv, double* lam_visc) {
  auto& g = c->geo;
  auto& s = c->fs;
  LAUNCH(c, k_lambda_points<ND>, nblk(g.Pd, 128), 128, g.Pd, viscous, c->cfg.gamma, c->cfg.prandtl_turb, g.nbr_ptr.p,
         g.nbr_pt.p, g.nbr_edge.p, g.edge_normal.p, g.E, s.V.p, lam_inv, lam_visc);
  for (auto& m : g.markers) {
    if (!marker_in_spectral_radius(m.kind) || m.nVertex == 0) continue;
    LAUNCH(c, k_lambda_bnd<ND>, nblk(m.nVertex, 128), 128, m.nVertex, g.Pd, viscous, c->cfg.gamma,
           c->cfg.prandtl_turb, m.node.p, m.normal.p, s.V.p, lam_inv, lam_visc);
  }
}

void flow_time_step(su2b200_ctx* c) {
  ProfScope prof_(c, PROF_TIMESTEP);
  auto& g = c->geo;
  auto& s = c->fs;
  const int visc = c->cfg.viscous;
  if (g.nDim == 2) spectral_radius_impl<2>(c, visc, s.lam_inv.p, s.lam_visc.p);
  else spectral_radius_impl<3>(c, visc, s.lam_inv.p, s.lam_visc.p);
  const int nb = nblk(g.Pd, 256);
  if (s.red.n < (size_t)2 * nb + 16) s.red.alloc((size_t)2 * nb + 16);
  LAUNCH(c, k_delta_time, nb, 256, g.Pd, visc, c->cfg.cfl, (const double*)s.local_cfl.p, c->cfg.max_delta_time,
         g.volume.p, s.lam_inv.p, s.lam_visc.p, s.dt.p, s.red.p);
  LAUNCH(c, k_minmax_final, 1, 256, nb, s.red.p, s.red.p + 2 * nb);
  SU2B_CUDA(cudaMemcpyAsync(s.h_red, s.red.p + 2 * nb, 2 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  c->min_dt = s.h_red[0];
  c->max_dt = s.h_red[1];
}

// CSolver::AdaptCFLNumber, the point loop (CSolver.cpp:1846-1915): blk = {min, max, sum} of the new CFL per block
__global__ void k_adapt_cfl(int Pd, int reduce, int reset, int can_increase, double fdec, double finc, double cfl_min,
                            double cfl_max, double mg_factor, const double* __restrict__ ur_flow,
                            const double* __restrict__ ur_turb, double* __restrict__ local_cfl,
                            double* __restrict__ blk) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  double mn = 1e30, mx = 0.0, sum = 0.0;
  if (i < Pd) {
    double CFL = local_cfl[i];
    const double ur = fmin(ur_flow[i], ur_turb ? ur_turb[i] : 1.0);
    double factor;
    if (ur < 0.1 || reduce) factor = fdec;
    else if ((ur >= 0.1 && ur < 1.0) || !can_increase) factor = 1.0;
    else factor = finc;
    if (CFL * factor <= cfl_min) { CFL = cfl_min; factor = mg_factor; }
    else if (CFL * factor >= cfl_max) { CFL = cfl_max; factor = mg_factor; }
    if (reset) { CFL = cfl_min; factor = mg_factor; }
    CFL *= factor;
    local_cfl[i] = CFL;
    mn = mx = sum = CFL;
  }
  __shared__ double smn[32], smx[32], ssm[32];
  for (int o = 16; o > 0; o >>= 1) {
    mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    sum += __shfl_xor_sync(0xffffffffu, sum, o);
  }
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) { smn[w] = mn; smx[w] = mx; ssm[w] = sum; }
  __syncthreads();
  if (threadIdx.x == 0) {
    const int nw = blockDim.x >> 5;
    for (int k = 1; k < nw; ++k) { mn = fmin(mn, smn[k]); mx = fmax(mx, smx[k]); sum += ssm[k]; }
    blk[(size_t)blockIdx.x * 3 + 0] = mn;
    blk[(size_t)blockIdx.x * 3 + 1] = mx;
    blk[(size_t)blockIdx.x * 3 + 2] = sum;
  }
}

// CSolver::AdaptCFLNumber (CSolver.cpp:1707-1949), finest grid: the decisions on the host from the linear residuals and
// the nonlinear residual series of this context, the per-point update on the device
void flow_adapt_cfl(su2b200_ctx* c, unsigned long inner_iter, double* out_min, double* out_max, double* out_avg) {
  ProfScope prof_(c, PROF_UPDATE);
  auto& g = c->geo;
  auto& s = c->fs;
  const auto& cfg = c->cfg;
  SU2B_REQUIRE(s.local_cfl.n == (size_t)g.P, "CFL adaptation is off (su2b200_flow_config::cfl_adapt)");
  const double fdec = cfg.cfl_adapt_param[0], finc = cfg.cfl_adapt_param[1], cmin = cfg.cfl_adapt_param[2],
               cmax = cfg.cfl_adapt_param[3], acceptable = cfg.cfl_adapt_param[4], starting = cfg.cfl_adapt_param[5];
  const unsigned long res_count = std::min(100ul, cfg.n_inner_iter > 0 ? cfg.n_inner_iter - 1 : 0ul);
  const bool rans = cfg.rans_sa != 0 || cfg.rans_sst != 0;
  const double linRes = std::max(c->lin_res[0], rans ? c->lin_res[1] : 0.0);
  const double linTol = std::max(acceptable, cfg.linear_tol);
  bool reset = linRes > 0.99;
  bool reduce = (linRes > 1.2 * linTol) && ((double)inner_iter >= starting);
  const bool can_increase = (linRes < linTol) && ((double)inner_iter >= starting);
  if (res_count > 0) {
    s.adapt_old_func = s.adapt_new_func;
    if (s.adapt_series.empty()) s.adapt_series.assign(res_count, 0.0);
    double f = 0.0;
    int total = 0;
    for (int v = 0; v < s.nVar; ++v) { f += std::log10(s.last_rms[v]); ++total; }
    if (rans) { f += std::log10(s.last_rms_turb); ++total; }
    if (cfg.rans_sst) { f += std::log10(s.last_rms_turb2); ++total; }
    s.adapt_new_func = f / total;
    if (inner_iter == 1 && s.adapt_new_func - s.adapt_old_func > 10) s.adapt_old_func = s.adapt_new_func;
    s.adapt_series[s.adapt_counter] = s.adapt_new_func - s.adapt_old_func;
    if (++s.adapt_counter == res_count) s.adapt_counter = 0;
    if (inner_iter >= res_count) {
      unsigned long sign_changes = 0;
      double total_change = 0.0;
      double prev = s.adapt_series.front();
      for (double val : s.adapt_series) {
        total_change += val;
        sign_changes += (prev > 0) ^ (val > 0);
        prev = val;
      }
      reduce = reduce || ((sign_changes > res_count / 4) && (total_change > -0.5));
      if (total_change > 2.0) {
        reset = true;
        s.adapt_counter = 0;
        for (auto& val : s.adapt_series) val = 0.0;
      }
    }
  }
  const int nb = nblk(g.Pd, 256);
  if (s.red.n < (size_t)3 * nb + 16) s.red.alloc((size_t)3 * nb + 16);
  LAUNCH(c, k_adapt_cfl, nb, 256, g.Pd, (int)reduce, (int)reset, (int)can_increase, fdec, finc, cmin, cmax, 1.0,
         (const double*)s.under_relax.p, rans ? (const double*)s.turb_under_relax.p : (const double*)nullptr,
         s.local_cfl.p, s.red.p);
  std::vector<double> blk((size_t)3 * nb);
  SU2B_CUDA(cudaMemcpyAsync(blk.data(), s.red.p, blk.size() * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  double v[3] = {1e30, 0.0, 0.0};
  for (int b = 0; b < nb; ++b) { v[0] = std::min(v[0], blk[3 * b]); v[1] = std::max(v[1], blk[3 * b + 1]); v[2] += blk[3 * b + 2]; }
  double npts = (double)g.Pd;
  if (c->nranks > 1) {
    allreduce_vec(c, &v[0], 1, 3);
    allreduce_vec(c, &v[1], 1, 2);
    double sums[2] = {v[2], npts};
    allreduce_vec(c, sums, 2, 0);
    v[2] = sums[0];
    npts = sums[1];
  }
  if (out_min) *out_min = v[0];
  if (out_max) *out_max = v[1];
  if (out_avg) *out_avg = v[2] / npts;
}

void flow_jst_preprocess(su2b200_ctx* c) {
  ProfScope prof_(c, PROF_JST_PRE);
  auto& g = c->geo;
  auto& s = c->fs;
  if (c->cfg.mg_level > 0) {
    // agglomerated levels: SetMax_Eigenvalue only; sensor and undivided Laplacian belong to JST on MESH_0
    // (center_jst = ... && iMesh == MESH_0, CEulerSolver.cpp:1584)
    if (g.nDim == 2) spectral_radius_impl<2>(c, 0, s.lambda.p, nullptr);
    else spectral_radius_impl<3>(c, 0, s.lambda.p, nullptr);
    return;
  }
  if (g.nDim == 2) {
    spectral_radius_impl<2>(c, 0, s.lambda.p, nullptr);
    LAUNCH(c, k_jst_sensor_lapl<2>, nblk(g.Pd, 128), 128, g.Pd, g.nbr_ptr.p, g.nbr_pt.p, g.phys_bnd.p, s.U.p, s.V.p,
           s.sensor.p, s.und_lapl.p);
  } else {
    spectral_radius_impl<3>(c, 0, s.lambda.p, nullptr);
    LAUNCH(c, k_jst_sensor_lapl<3>, nblk(g.Pd, 128), 128, g.Pd, g.nbr_ptr.p, g.nbr_pt.p, g.phys_bnd.p, s.U.p, s.V.p,
           s.sensor.p, s.und_lapl.p);
  }
}

void flow_vorticity_strain(su2b200_ctx* c) {
  ProfScope prof_(c, PROF_GRADIENT);
  auto& g = c->geo;
  auto& s = c->fs;
  if (g.nDim == 2) LAUNCH(c, k_vorticity_strain<2>, nblk(g.P, 256), 256, g.P, s.nG, s.grad_prim.p, s.vorticity.p, s.strain_mag.p);
  else LAUNCH(c, k_vorticity_strain<3>, nblk(g.P, 256), 256, g.P, s.nG, s.grad_prim.p, s.vorticity.p, s.strain_mag.p);
}

// generic prepare-implicit for a system with NV variables
template <int NV>
static void prepare_implicit_impl(su2b200_ctx* c, LinSys& ls, const double* dt, double* res, double* sol,
                                  const double* trunc, double* out_rms, double* out_max, unsigned long* out_idx) {
  ++ls.A_gen;  // V/dt lands on the diagonal blocks
  auto& g = c->geo;
  auto& s = c->fs;
  const int nb = nblk(g.Pd, 256);
  const size_t need = (size_t)nb * NV * 2 + 3 * NV + 16;
  if (s.red.n < need + (size_t)nb * NV) s.red.alloc(need + (size_t)nb * NV);
  double* part_rms = s.red.p;
  double* part_max = s.red.p + (size_t)nb * NV;
  int* part_idx = reinterpret_cast<int*>(s.red.p + (size_t)2 * nb * NV);
  double* out = s.red.p + (size_t)3 * nb * NV;
  LAUNCH(c, k_prepare_implicit<NV>, nb, 256, g.Pd, ls.dia_ptr, g.volume.p, dt, ls.A.p, res, trunc, sol, part_rms,
         part_max, part_idx);
  LAUNCH(c, k_residual_final<NV>, NV, 256, nb, part_rms, part_max, part_idx, out);
  const size_t b = (size_t)g.Pd * NV, e = (size_t)g.P * NV;
  if (e > b) LAUNCH(c, k_zero_range, nblk(e - b, 256), 256, res, sol, b, e);
  SU2B_CUDA(cudaMemcpyAsync(s.h_red, out, 3 * NV * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  double npts = (double)g.Pd;
  if (c->nranks > 1) {  // CSolver::SetResidual_RMS: sums and point count over all ranks, maxima by MPI_MAX
    double sums[NV + 1];
    for (int v = 0; v < NV; ++v) sums[v] = s.h_red[v];
    sums[NV] = npts;
    allreduce_vec(c, sums, NV + 1, 0);
    double mx[NV];
    for (int v = 0; v < NV; ++v) mx[v] = s.h_red[NV + v];
    allreduce_vec(c, mx, NV, 2);
    for (int v = 0; v < NV; ++v) {
      s.h_red[v] = sums[v];
      if (mx[v] != s.h_red[NV + v]) s.h_red[2 * NV + v] = 0.0;  // the maximum is on another rank
      s.h_red[NV + v] = mx[v];
    }
    npts = sums[NV];
  }
  for (int v = 0; v < NV; ++v) {
    // CSolver::SetResidual_RMS (CSolver.cpp:1951-1985): N = global nPointDomain
    double rms = sqrt(s.h_red[v] / npts);
    const double eps = 1.0e-16;
    if (!(rms > eps * eps)) rms = (rms != rms) ? rms : eps * eps;
    if (out_rms) out_rms[v] = rms;
    if (out_max) out_max[v] = s.h_red[NV + v];
    if (out_idx) out_idx[v] = (unsigned long)s.h_red[2 * NV + v];
  }
}

void flow_prepare_implicit(su2b200_ctx* c, double* rms, double* mx, unsigned long* idx) {
  ProfScope prof_(c, PROF_PREPARE);
  auto& s = c->fs;
  // rhs = -(R + Res_TruncError): the multigrid forcing term of a coarse level (null without multigrid)
  const double* trunc = s.res_trunc.p;
  if (c->geo.nDim == 2) prepare_implicit_impl<4>(c, c->sys[0], s.dt.p, s.res.p, s.sol.p, trunc, rms, mx, idx);
  else prepare_implicit_impl<5>(c, c->sys[0], s.dt.p, s.res.p, s.sol.p, trunc, rms, mx, idx);
}

void turb_prepare_implicit_generic(su2b200_ctx* c, double* rms, double* mx, unsigned long* idx) {
  ProfScope prof_(c, PROF_PREPARE);
  auto& s = c->fs;
  if (s.nTurb == 2) prepare_implicit_impl<2>(c, c->sys[1], s.turb_dt.p, s.turb_res.p, s.turb_sol.p, nullptr, rms, mx, idx);
  else prepare_implicit_impl<1>(c, c->sys[1], s.turb_dt.p, s.turb_res.p, s.turb_sol.p, nullptr, rms, mx, idx);
}

void flow_complete_implicit(su2b200_ctx* c) {
  ProfScope prof_(c, PROF_UPDATE);
  auto& g = c->geo;
  auto& s = c->fs;
  if (g.nDim == 2) LAUNCH(c, k_complete_implicit<4>, nblk(g.Pd, 256), 256, g.Pd, c->cfg.max_update_flow, s.sol.p, s.U_old.p, s.U.p, s.under_relax.p);
  else LAUNCH(c, k_complete_implicit<5>, nblk(g.Pd, 256), 256, g.Pd, c->cfg.max_update_flow, s.sol.p, s.U_old.p, s.U.p, s.under_relax.p);
}

}  // namespace su2b200
