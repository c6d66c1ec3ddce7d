// Host-side launch wrappers (one per reference loop). All run on ctx->stream.
#pragma once
#include "internal.hpp"

namespace su2b200 {

// flow_kernels.cu
void flow_primitives(su2b200_ctx* c);
void flow_gradient(su2b200_ctx* c, int reconstruction, int nvars);
void flow_limiter(su2b200_ctx* c);
void scalar2_gradient(su2b200_ctx* c, int method, const double* field, double* grad);  // [P][2] -> [P][2][nDim]
void scalar_gradient(su2b200_ctx* c, int method, const double* field, double* grad);
void flow_time_step(su2b200_ctx* c);
void flow_jst_preprocess(su2b200_ctx* c);
void flow_vorticity_strain(su2b200_ctx* c);
void flow_prepare_implicit(su2b200_ctx* c, double* rms, double* mx, unsigned long* idx);
void turb_prepare_implicit_generic(su2b200_ctx* c, double* rms, double* mx, unsigned long* idx);
void flow_complete_implicit(su2b200_ctx* c);
void flow_adapt_cfl(su2b200_ctx* c, unsigned long inner_iter, double* cfl_min, double* cfl_max, double* cfl_avg);

// edge_kernels.cu
void flow_edge_residual(su2b200_ctx* c, unsigned long* nonphysical_edges);

// bc_kernels.cu
void flow_bc_far_field(su2b200_ctx* c, int iMarker);
void flow_bc_sym_plane(su2b200_ctx* c, int iMarker);
void flow_bc_heatflux_wall(su2b200_ctx* c, int iMarker);
void flow_forces(su2b200_ctx* c, double* coeffs);

// linalg.cu, precond.cu, precond_poll.cu, precond_factor.cu, krylov.cu are compiled twice: namespace f64 (real = double, the path
// every parity gate runs) and namespace f32 (real = float: the reference's mixed-precision mode, su2mixedfloat = float,
// code_config.hpp:133-139, where the Jacobian, the preconditioner and every Krylov vector are single precision).  Both sets
// of overloads are visible in su2b200; the LinSysT<double> / LinSysT<float> argument selects.
#define real double
#define LinSysR LinSysT<double>
namespace f64 {
#include "linalg_decl.inc"
}
#undef real
#undef LinSysR
#define real float
#define LinSysR LinSysT<float>
namespace f32 {
#include "linalg_decl.inc"
}
#undef real
#undef LinSysR
using namespace f64;
using namespace f32;

// turb_sa.cu
void turb_preprocessing(su2b200_ctx* c);
void turb_set_time_step(su2b200_ctx* c);
void turb_space_integration(su2b200_ctx* c);
void turb_complete_implicit(su2b200_ctx* c);
void turb_postprocessing(su2b200_ctx* c);

// comm.cu
void halo_exchange(su2b200_ctx* c, double* arr, int count_per_point);
void halo_exchange(su2b200_ctx* c, float* arr, int count_per_point);
void halo_exchange_start(su2b200_ctx* c, float* arr, int count_per_point);
void halo_exchange_start(su2b200_ctx* c, double* arr, int count_per_point);  // returns at once; the halos of arr are in flight
void halo_exchange_finish(su2b200_ctx* c);                                   // the compute stream waits for them
double allreduce_sum(su2b200_ctx* c, double v);
void allreduce_dev_sum(su2b200_ctx* c, double* d, int n);       // device buffer, in stream
void allreduce_dev_sum(su2b200_ctx* c, float* d, int n);
// Mailbox allreduce of ONE double inside a kernel (one block; every thread of it must call): see comm.cu
struct MailboxView {
  unsigned long long* local;
  unsigned long long* const* peers;  // device table [nranks]
  int rank, nranks;
};
bool mailbox_view(su2b200_ctx* c, MailboxView* v);  // false: not connected (use NCCL)
void mailbox_release(su2b200_ctx* c);
void allreduce_vec(su2b200_ctx* c, double* v, int n, int op);  // op: 0 sum, 2 max, 3 min

}  // namespace su2b200

#define SU2B_LAUNCH(ctx, kern, grid, block, ...)                \
  do {                                                          \
    if ((grid) > 0) {                                           \
      kern<<<(grid), (block), 0, (ctx)->stream>>>(__VA_ARGS__); \
      (ctx)->launches++;                                        \
      SU2B_CUDA(cudaGetLastError());                            \
    }                                                           \
  } while (0)

// ---- device side of the mailbox allreduce (header: called from krylov.cu's reduction kernel) ---------------------------
// Layout of a mailbox (unsigned long long words): [0] next sequence number (local use), [16 + (slot * 16 + r) * 2 + {0, 1}]
// = {value bits, tag} written by rank r for the operation whose sequence number has that slot.  A rank stores its value
// and then, release at system scope, the tag seq + 1 into EVERY rank's mailbox (posted NVLink stores), and reads only
// its own memory: acquire-polls the nranks tags of the slot, then sums the values in rank order (the same order on every
// rank: bit-identical results everywhere, which the device-side branch decisions of the Gram-Schmidt rely on).
// Slot reuse: a rank starts operation k + 1 only after it has read every value of operation k, and sees a peer's k + 1
// contribution only after that peer has done the same; with 4 slots a store for k + 4 can never overtake a read of k.
#ifdef __CUDACC__
namespace su2b200 {
__device__ __forceinline__ double mailbox_allreduce_block(const MailboxView mb, double v, bool* timed_out) {
  constexpr int SLOTS = 4, MAXR = 16;
  __shared__ double mb_vals[MAXR];
  __shared__ int mb_fail;
  const int t = threadIdx.x;
  if (t == 0) mb_fail = 0;
  __syncthreads();
  const unsigned long long seq = mb.local[0];
  const unsigned long long tag = seq + 1;
  const size_t base = 16 + (size_t)((seq % SLOTS) * MAXR) * 2;
  if (t < mb.nranks) {
    unsigned long long* dst = mb.peers[t] + base + (size_t)mb.rank * 2;
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(dst), "l"((unsigned long long)__double_as_longlong(v)) : "memory");
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(dst + 1), "l"(tag) : "memory");
    const unsigned long long* src = mb.local + base + (size_t)t * 2;
    unsigned long long got = 0;
    for (unsigned spins = 0;; ++spins) {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(got) : "l"(src + 1) : "memory");
      if (got == tag) break;
      if (spins > (1u << 24)) { mb_fail = 1; break; }  // seconds: a peer died or the ranks disagree on the sequence
      if (spins > 64) __nanosleep(64);
    }
    unsigned long long bits = 0;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(bits) : "l"(src) : "memory");
    mb_vals[t] = __longlong_as_double((long long)bits);
  }
  __syncthreads();
  double s = 0.0;
  for (int r = 0; r < mb.nranks; ++r) s += mb_vals[r];
  if (timed_out) *timed_out = mb_fail != 0;
  __syncthreads();
  if (t == 0) mb.local[0] = seq + 1;
  return s;
}
}  // namespace su2b200
#endif
