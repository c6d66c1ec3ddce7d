// Internal structures of libsu2b200 (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include <algorithm>
#include <stdexcept>
#include "../../include/su2b200.h"

namespace su2b200 {

// ---- errors -------------------------------------------------------------------------------
void set_error(const std::string& msg);
struct Error : std::runtime_error { using std::runtime_error::runtime_error; };

#define SU2B_CUDA(call)                                                                      \
  do {                                                                                       \
    cudaError_t e_ = (call);                                                                 \
    if (e_ != cudaSuccess)                                                                   \
      throw su2b200::Error(std::string(#call) + " failed: " + cudaGetErrorString(e_) + " at " + \
                           __FILE__ + ":" + std::to_string(__LINE__));                       \
  } while (0)

#define SU2B_REQUIRE(cond, msg)                                      \
  do {                                                               \
    if (!(cond)) throw su2b200::Error(std::string(msg));             \
  } while (0)

// ---- device vector --------------------------------------------------------------------------
template <class T>
struct dvec {
  T* p = nullptr;
  size_t n = 0;
  dvec() = default;
  dvec(const dvec&) = delete;
  dvec& operator=(const dvec&) = delete;
  dvec(dvec&& o) noexcept : p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
  dvec& operator=(dvec&& o) noexcept {
    if (this != &o) { release(); p = o.p; n = o.n; o.p = nullptr; o.n = 0; }
    return *this;
  }
  ~dvec() { release(); }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    n = 0;
  }
  void alloc(size_t count) {
    if (count == n && p) return;
    release();
    n = count;
    if (count) SU2B_CUDA(cudaMalloc(&p, count * sizeof(T)));
  }
  void zero(cudaStream_t s = 0) {
    if (n) SU2B_CUDA(cudaMemsetAsync(p, 0, n * sizeof(T), s));
  }
  void upload(const T* h, size_t count, cudaStream_t s = 0) {
    alloc(count);
    if (count) {
      SU2B_CUDA(cudaMemcpyAsync(p, h, count * sizeof(T), cudaMemcpyHostToDevice, s));
      SU2B_CUDA(cudaStreamSynchronize(s));
    }
  }
  void upload(const std::vector<T>& h, cudaStream_t s = 0) { upload(h.data(), h.size(), s); }
  void download(T* h, cudaStream_t s = 0) const {
    if (n) {
      SU2B_CUDA(cudaMemcpyAsync(h, p, n * sizeof(T), cudaMemcpyDeviceToHost, s));
      SU2B_CUDA(cudaStreamSynchronize(s));
    }
  }
  std::vector<T> to_host(cudaStream_t s = 0) const {
    std::vector<T> h(n);
    download(h.data(), s);
    return h;
  }
};

// ---- geometry ---------------------------------------------------------------------------------
struct Marker {
  int kind = 0;
  int nVertex = 0;
  int monitoring = 0;
  double heat_flux = 0.0;
  dvec<int> node;          // [nV]
  dvec<double> normal;     // [nV][nDim]
  dvec<int> normal_nbr;    // [nV]
  dvec<double> unit_sym;   // [nV][nDim] unit normal used by symmetry/Euler-wall treatment
  dvec<double> charac;     // [nV][nDim+9] CharacPrimVar (CFVMFlowSolverBase.hpp:176), zero-initialised
  dvec<double> nn_dist;    // [nV] CVertex::NearestNeighborDistance of viscous walls (SST wall omega; built on first use)
};

struct Geometry {
  int nDim = 0;
  int P = 0, Pd = 0, E = 0;
  long nnz = 0;
  dvec<int> edge_i, edge_j;       // [E]
  dvec<double> edge_normal;       // SoA [nDim][E]
  dvec<double> coord;             // [P][nDim]
  dvec<double> volume;            // [P]
  dvec<int> nbr_ptr, nbr_pt, nbr_edge;  // neighbour CSR in the reference's order
  dvec<int> row_ptr, col_ind, dia_ptr;  // Jacobian pattern
  dvec<int> edge_ij, edge_ji;     // [E] block index of (i,j) and (j,i)
  dvec<int> gat_edge, gat_blk;    // per-point edge list sorted by edge index (see k_point_sum)
  dvec<int> nnbr;                 // CPoint::nNeighbor
  bool nnbr_global = false;       // halo entries of nnbr hold the owner's count (set once comm is up)
  dvec<unsigned char> phys_bnd;   // CPoint::PhysicalBoundary
  dvec<double> wall_dist;
  // edge colouring (COLORED assembly)
  std::vector<int> color_ptr;     // host
  dvec<int> color_edges;
  std::vector<Marker> markers;
  bool uploaded = false;
  bool have_phys_bnd = false;
  // host copies needed by host-side setup (ILU level schedule)
  std::vector<int> h_row_ptr, h_col_ind, h_dia_ptr;
};

// ---- block-CSR linear system (CSysMatrix + CSysVector + CSysSolve workspace) -----------------
template <class T>
struct LinSysT {
  int nVar = 0;
  int P = 0, Pd = 0;
  long nnz = 0;
  const int* row_ptr = nullptr;  // device, borrowed from Geometry or owned below
  const int* col_ind = nullptr;
  const int* dia_ptr = nullptr;
  dvec<int> own_row_ptr, own_col_ind, own_dia_ptr;
  std::vector<int> h_row_ptr, h_col_ind, h_dia_ptr;
  dvec<T> A;      // [nnz][nVar*nVar]
  dvec<T> ilu;    // ILU_matrix
  dvec<T> invM;   // [Pd][nVar*nVar] inverse diagonal (Jacobi or ILU's U_ii^-1)
  // level schedule for exact ILU0 build/apply (rows grouped by dependency level)
  bool levels_ready = false;
  std::vector<int> lvl_ptr_lower, lvl_ptr_upper;  // host offsets into the row lists
  dvec<int> lvl_rows_lower, lvl_rows_upper;
  int nlev_lower = 0, nlev_upper = 0;
  // single-launch dependency-driven sweeps (precond.cu): rows in level order, padded per level to the group size;
  // ready[i] == epoch marks row i as finished in the current sweep; ctrl = {ticket, abort}
  dvec<int> sweep_rows_lower, sweep_rows_upper;
  int sweep_n_lower = 0, sweep_n_upper = 0;
  dvec<int> sweep_ready, sweep_ctrl;
  int sweep_epoch = 0;
  int* h_sweep_ctrl = nullptr;       // pinned
  int sweep_mode = 2;                // apply: 2 data-polling sweeps (precond_poll.cu), 1 flag sweeps, 0 launch per level
  int build_mode = 2;                // factor: 2 single-launch polling factorisation (precond_factor.cu), 1 flag sweep,
                                     //         0 launch per level
  // level width from which the nVar-lanes-per-row kernels take over; 0 = by pattern: 256 when no row has more than three
  // entries on either side of the diagonal (hexahedra, quadrilaterals: 185^3 apply 2.80 -> 2.67 ms, build 6.41 -> 6.25 ms
  // against 4096), else 4096 (rows in chunks: tetrahedral box 5.7 ms at 4096, 6.3 ms at 512)
  int wide_sweep = 0, wide_factor = 0;
  int auto_wide() const {
    int longest = 0;
    for (int i = 0; i < Pd; ++i) {
      longest = std::max(longest, h_dia_ptr[i] - h_row_ptr[i]);
      int up = 0;
      for (int k = h_dia_ptr[i] + 1; k < h_row_ptr[i + 1]; ++k) up += h_col_ind[k] < Pd;
      longest = std::max(longest, up);
    }
    return longest <= 3 ? 256 : 4096;
  }
  int factor_rows_mode = 1;                   // 0 warp-per-row only, 1 hybrid, 2 rows kernel only (k_factor_rows)
  bool parts_explicit = false;                // h_parts given by su2b200_linsys_set_partition_rows
  // generation of the matrix values (bumped by every writer of A) and the one the ILU factors were built from: the
  // slim factors read the upper blocks from A at apply time, so an apply after A changed must fail loudly
  unsigned long A_gen = 1, ilu_gen = 0;
  bool factor_ready = false, factor_ok = false;
  dvec<int> factor_desc, factor_ops;  // symbolic factorisation: row descriptors (level order) and update lists
  int factor_own_cap = 0, factor_src_cap = 0;
  bool factor_rows_ok = false;        // every row fits k_factor_rows' tiles
  int factor_groups = 0;              // groups of 32/nVar level-padded rows (k_factor_rows)
  dvec<int> factor_desc_rows;
  bool upper_stable = false;  // no upper block is ever modified by the factorisation: U_ik == A_ik for k > i
  bool ilu_slim = false;      // the last build wrote only the lower blocks: the backward sweep reads the upper ones from A
  int factor_split[4] = {0, 0, 0, 0};  // wide middle: level-ordered rows [0],[1]) = padded groups [2],[3])
  int factor_slots = 0;               // scalar systems: level-padded row slots of k_factor_poll_s
  dvec<int> sweep_metar_lower, sweep_metar_upper;  // same, levels padded to 32/nVar rows (k_poll_sweep_rows)
  dvec<int> sweep_meta_lower, sweep_meta_upper;  // int4 {row, first, last+1, 0} per level-ordered position (mode 2)
  int poll_split_lower[4] = {0, 0, 0, 0}, poll_split_upper[4] = {0, 0, 0, 0};  // hybrid sweep: see build_meta
  int poll_groups_lower = 0, poll_groups_upper = 0;  // scalar systems: groups of 4 level-padded rows
  dvec<T> sweep_tmp, sweep_tmp2;  // forward-sweep result / copy of an in-place rhs (mode 2)
  // reference-style row partitions (block-Jacobi ILU across partitions); nparts <= 1: exact serial factor
  int nparts = 1;
  std::vector<int> h_parts;
  dvec<int> parts;
  // Krylov workspace
  std::vector<dvec<T>*> work;   // lazily grown pool of [P*nVar] vectors
  dvec<T> scalars;              // device scalars for reductions
  T* h_scalars = nullptr;       // pinned host mirror
  T* h_mgs = nullptr;           // pinned: Hessenberg column of the device Gram-Schmidt
  // fused upper product: a backward ILU sweep whose upper factor is the Jacobian's upper triangle (upper_stable) computes
  // sum_{j>i} A_ij z_j on its way; the product A z that follows in a Krylov iteration then only reads the lower triangle,
  // the diagonal and the halo columns (4 of 7 blocks on a hexahedral mesh) and adds it
  dvec<T> upper_prod;               // [Pd][nVar]
  const T* upper_prod_of = nullptr; // the vector z it belongs to (set by the apply, cleared by the next apply)
  bool want_upper_prod = false;          // set by the Krylov solver around apply + product
  dvec<int> rest_halo_ptr;               // per owned row: first entry with a halo column (multi-rank only)
  bool rest_halo_ready = false;
  // owned rows with at least one halo column (multi-rank): redone by the overlapped product once the halos arrived
  dvec<int> halo_rows;
  int n_halo_rows = -1;              // -1: not built yet
  ~LinSysT();
  dvec<T>& vec(size_t i);
  size_t vecLen() const { return (size_t)P * nVar; }
};
template <class T>
LinSysT<T>::~LinSysT() {
  for (auto* v : work) delete v;
  if (h_scalars) cudaFreeHost(h_scalars);
  if (h_mgs) cudaFreeHost(h_mgs);
}
template <class T>
dvec<T>& LinSysT<T>::vec(size_t i) {
  while (work.size() <= i) work.push_back(new dvec<T>());
  if (work[i]->n != vecLen()) {
    work[i]->alloc(vecLen());
    work[i]->zero();
  }
  return *work[i];
}
using LinSys = LinSysT<double>;
using LinSysF = LinSysT<float>;  // the reference's mixed-precision mode (code_config.hpp:133-139)

// ---- flow state ---------------------------------------------------------------------------------
struct FlowState {
  bool allocated = false;
  int nVar = 0, nPrim = 0, nG = 0;
  dvec<double> U, U_old, V;
  dvec<double> grad_recon, grad_prim, limiter;
  dvec<double> dt, lam_inv, lam_visc, under_relax;
  dvec<double> lambda, sensor, und_lapl;     // JST
  dvec<double> res, sol;                     // LinSysRes, LinSysSol
  dvec<double> res_trunc;                    // Res_TruncError (zero without multigrid)
  dvec<double> local_cfl;                    // CVariable::LocalCFL (CFL_ADAPT= YES only; else the scalar cfg.cfl)
  // CSolver::AdaptCFLNumber bookkeeping (CSolver.hpp: Old_Func, New_Func, NonLinRes_Series, NonLinRes_Counter)
  double adapt_old_func = 0.0, adapt_new_func = 0.0;
  std::vector<double> adapt_series;
  unsigned long adapt_counter = 0;
  double last_rms[5] = {0, 0, 0, 0, 0}, last_rms_turb = 0.0;
  dvec<double> edge_flux;                    // SoA [nVar][E] (GATHER assembly)
  dvec<signed char> edge_counter;            // NonPhysicalEdgeCounter
  dvec<double> vorticity, strain_mag;
  // turbulence (SA)
  dvec<double> nu, nu_old, nu_grad, mu_t, turb_res, turb_sol, turb_dt, turb_under_relax, turb_edge_flux;
  double last_rms_turb2 = 0.0;
  int nTurb = 1;                             // variables of the turbulence model: 1 (SA: nu_tilde), 2 (SST: k, omega)
  dvec<double> sst_F1, sst_F2, sst_CDkw;     // CTurbSSTVariable blending functions and cross diffusion [P]
  dvec<unsigned long long> counters;         // device counters (non-physical points/edges)
  dvec<double> red;                          // reduction scratch
  double* h_red = nullptr;                   // pinned
  unsigned long long* h_counters = nullptr;  // pinned
  bool recon_aliases_prim = false;           // Gradient_Reconstruction is Gradient_Primitive
};

// ---- optional per-category device timing (bench.py's kernel shares; off in timed regions) ------------------
enum ProfCat { PROF_PRIMITIVES = 0, PROF_GRADIENT, PROF_LIMITER, PROF_TIMESTEP, PROF_JST_PRE, PROF_EDGE, PROF_POINT_SUM,
               PROF_BC, PROF_PREPARE, PROF_PREC_BUILD, PROF_PREC_APPLY, PROF_SPMV, PROF_DOT, PROF_VECOPS, PROF_UPDATE,
               PROF_FORCES, PROF_TURB_EDGE, PROF_TURB_SOURCE, PROF_TURB_OTHER, PROF_HALO, PROF_NCAT };
struct Profiler {
  bool on = false;
  struct Rec { int cat; cudaEvent_t a, b; };
  std::vector<Rec> recs;
  std::vector<cudaEvent_t> pool;
  cudaEvent_t get() {
    if (!pool.empty()) { cudaEvent_t e = pool.back(); pool.pop_back(); return e; }
    cudaEvent_t e;
    cudaEventCreate(&e);
    return e;
  }
};

}  // namespace su2b200

struct su2b200_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t comm_stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  unsigned long launches = 0;
  su2b200::Geometry geo;
  su2b200_flow_config cfg{};
  bool configured = false;
  su2b200::FlowState fs;
  su2b200::LinSys sys[2];
  su2b200::LinSysF sysf[2];  // single-precision copies for the mixed-precision linear solve (empty unless cfg.mixed_precision)
  // comm
  void* nccl = nullptr;
  int rank = 0, nranks = 1;
  std::vector<int> p2p_ranks, p2p_send_ptr, p2p_recv_ptr;
  su2b200::dvec<int> p2p_send_idx, p2p_recv_idx;
  su2b200::dvec<double> p2p_send_buf, p2p_recv_buf;
  // small-message allreduce over peer memory (comm.cu): every rank owns a mailbox that its peers write into over NVLink
  static constexpr int MBOX_MAX_RANKS = 16, MBOX_SLOTS = 4;
  unsigned long long* mbox_local = nullptr;          // device, cudaMalloc'ed (exported through cudaIpc)
  unsigned long long* mbox_peer_host[MBOX_MAX_RANKS] = {};
  su2b200::dvec<unsigned long long> mbox_peers;      // device table of the nranks mailbox addresses (own one included)
  bool mbox_connected = false;  // peers' handles opened
  bool mbox_ready = false;      // agreed by all ranks (su2b200_comm_mailbox_enable)
  // overlapped exchange (e3): pack on the compute stream, send/recv + unpack on comm_stream, joined by events
  cudaEvent_t ev_packed = nullptr, ev_unpacked = nullptr;
  bool exchange_in_flight = false;
  bool defer_halo = false;  // inside a Krylov solve: the operators skip their trailing exchange, the product does it overlapped
  // FAS multigrid: this context as a COARSE level (children = points of the next finer level, CPoint::Children_CV)
  su2b200_ctx* mg_fine = nullptr;
  su2b200::dvec<int> mg_child_ptr, mg_child;
  bool stream_borrowed = false;  // a coarse level runs on the stream of the finest one
  unsigned long inner_iter = 0;  // CConfig::GetInnerIter of the next su2b200_iterate iteration
  // last linear solve info
  unsigned long lin_iters[2] = {0, 0};
  double lin_res[2] = {0, 0};
  double min_dt = 0, max_dt = 0;
  su2b200::Profiler prof;
};

namespace su2b200 {
// RAII scope: records two events on the compute stream around a category when profiling is enabled
struct ProfScope {
  su2b200_ctx* c;
  cudaEvent_t a = nullptr;
  int cat;
  ProfScope(su2b200_ctx* ctx, int category) : c(ctx), cat(category) {
    if (c->prof.on) { a = c->prof.get(); cudaEventRecord(a, c->stream); }
  }
  ~ProfScope() {
    if (a) { cudaEvent_t b = c->prof.get(); cudaEventRecord(b, c->stream); c->prof.recs.push_back({cat, a, b}); }
  }
};
}  // namespace su2b200
