// a28. Halo exchange and global reductions: CSolver::InitiateComms/CompleteComms (SU2_CFD/src/solvers/CSolver.cpp:1406-1697),
// CSysMatrixComms (Common/src/linear_algebra/CSysMatrix.cpp:251-477), tables of CGeometry::PreprocessP2PComms
// (Common/src/geometry/CGeometry.cpp:139-332). Device pack -> ncclSend/ncclRecv group -> device unpack.
// NCCL is resolved at run time with dlopen (libnccl.so.2: the copy torch already loaded, or the system one), so
// the single-GPU path has no NCCL dependency.
#include "internal.hpp"
#include "kernels.hpp"
#include <dlfcn.h>

namespace su2b200 {

static inline int nblk(long n, int bs) { return (int)((n + bs - 1) / bs); }

// minimal NCCL ABI (nccl.h 2.x): opaque comm, 128-byte unique id, ncclFloat64 = 8, ncclSum = 0
struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(void*) = nullptr;
  int (*CommInitRank)(void**, int, unsigned char[128], int) = nullptr;  // id passed by value (128-byte struct)
  int (*CommDestroy)(void*) = nullptr;
  int (*Send)(const void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*Recv)(void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
struct NcclId { unsigned char b[128]; };

static NcclApi& nccl() {
  static NcclApi api;
  if (api.lib) return api;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (api.lib) break;
  }
  if (!api.lib) throw Error("NCCL not found (dlopen libnccl.so.2 failed): multi-GPU runs need it");
  auto sym = [&](const char* s) {
    void* p = dlsym(api.lib, s);
    if (!p) throw Error(std::string("NCCL symbol missing: ") + s);
    return p;
  };
  api.GetUniqueId = (int (*)(void*))sym("ncclGetUniqueId");
  api.CommInitRank = (int (*)(void**, int, unsigned char[128], int))sym("ncclCommInitRank");
  api.CommDestroy = (int (*)(void*))sym("ncclCommDestroy");
  api.Send = (int (*)(const void*, size_t, int, int, void*, cudaStream_t))sym("ncclSend");
  api.Recv = (int (*)(void*, size_t, int, int, void*, cudaStream_t))sym("ncclRecv");
  api.AllReduce = (int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t))sym("ncclAllReduce");
  api.GroupStart = (int (*)())sym("ncclGroupStart");
  api.GroupEnd = (int (*)())sym("ncclGroupEnd");
  api.GetErrorString = (const char* (*)(int))sym("ncclGetErrorString");
  return api;
}
#define SU2B_NCCL(call)                                                                         \
  do {                                                                                          \
    int r_ = (call);                                                                            \
    if (r_ != 0) throw Error(std::string(#call) + " failed: " + nccl().GetErrorString(r_));     \
  } while (0)

template <class T>
__global__ void k_pack(int n, int cnt, const int* __restrict__ idx, const T* __restrict__ arr, T* __restrict__ buf) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * cnt) return;
  const int p = t / cnt, k = t % cnt;
  buf[t] = arr[(size_t)idx[p] * cnt + k];
}
template <class T>
__global__ void k_unpack(int n, int cnt, const int* __restrict__ idx, const T* __restrict__ buf, T* __restrict__ arr) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * cnt) return;
  const int p = t / cnt, k = t % cnt;
  arr[(size_t)idx[p] * cnt + k] = buf[t];
}

// ncclDataType_t: ncclFloat32 = 7, ncclFloat64 = 8; the staging buffers are sized in doubles and reused for floats
template <class T> struct NcclType;
template <> struct NcclType<double> { static constexpr int id = 8; };
template <> struct NcclType<float> { static constexpr int id = 7; };

template <class T>
static void halo_exchange_impl(su2b200_ctx* c, T* arr, int cnt) {
  if (c->nranks <= 1 || c->p2p_ranks.empty()) return;
  SU2B_REQUIRE(c->nccl, "halo exchange requested without a communicator (su2b200_comm_init)");
  const int nn = (int)c->p2p_ranks.size();
  const int nsend = c->p2p_send_ptr[nn], nrecv = c->p2p_recv_ptr[nn];
  if (c->p2p_send_buf.n < (size_t)nsend * cnt) c->p2p_send_buf.alloc((size_t)nsend * cnt);
  if (c->p2p_recv_buf.n < (size_t)nrecv * cnt) c->p2p_recv_buf.alloc((size_t)nrecv * cnt);
  T* sendb = reinterpret_cast<T*>(c->p2p_send_buf.p);
  T* recvb = reinterpret_cast<T*>(c->p2p_recv_buf.p);
  SU2B_LAUNCH(c, k_pack<T>, nblk((long)nsend * cnt, 256), 256, nsend, cnt, c->p2p_send_idx.p, arr, sendb);
  auto& api = nccl();
  SU2B_NCCL(api.GroupStart());
  for (int k = 0; k < nn; ++k) {
    const int ns = c->p2p_send_ptr[k + 1] - c->p2p_send_ptr[k], nr = c->p2p_recv_ptr[k + 1] - c->p2p_recv_ptr[k];
    if (nr) SU2B_NCCL(api.Recv(recvb + (size_t)c->p2p_recv_ptr[k] * cnt, (size_t)nr * cnt, NcclType<T>::id, c->p2p_ranks[k], c->nccl, c->stream));
    if (ns) SU2B_NCCL(api.Send(sendb + (size_t)c->p2p_send_ptr[k] * cnt, (size_t)ns * cnt, NcclType<T>::id, c->p2p_ranks[k], c->nccl, c->stream));
  }
  SU2B_NCCL(api.GroupEnd());
  SU2B_LAUNCH(c, k_unpack<T>, nblk((long)nrecv * cnt, 256), 256, nrecv, cnt, c->p2p_recv_idx.p, recvb, arr);
}
void halo_exchange(su2b200_ctx* c, double* arr, int cnt) { halo_exchange_impl<double>(c, arr, cnt); }
void halo_exchange(su2b200_ctx* c, float* arr, int cnt) { halo_exchange_impl<float>(c, arr, cnt); }

// e3. The same exchange split in two so that interior work overlaps it: the pack runs on the compute stream (it reads what
// the preceding kernel wrote), send/recv and the unpack run on comm_stream, and halo_exchange_finish makes the compute
// stream wait for the unpack.  Between start and finish the compute stream may read the owned entries of `arr` and must
// not trust (nor write) its halo entries.  One exchange in flight at a time (the staging buffers are shared).
template <class T>
static void halo_exchange_start_impl(su2b200_ctx* c, T* arr, int cnt) {
  if (c->nranks <= 1 || c->p2p_ranks.empty()) return;
  SU2B_REQUIRE(c->nccl, "halo exchange requested without a communicator (su2b200_comm_init)");
  SU2B_REQUIRE(!c->exchange_in_flight, "halo_exchange_start: the previous exchange was not finished");
  if (!c->comm_stream) SU2B_CUDA(cudaStreamCreateWithFlags(&c->comm_stream, cudaStreamNonBlocking));
  if (!c->ev_packed) {
    SU2B_CUDA(cudaEventCreateWithFlags(&c->ev_packed, cudaEventDisableTiming));
    SU2B_CUDA(cudaEventCreateWithFlags(&c->ev_unpacked, cudaEventDisableTiming));
  }
  const int nn = (int)c->p2p_ranks.size();
  const int nsend = c->p2p_send_ptr[nn], nrecv = c->p2p_recv_ptr[nn];
  if (c->p2p_send_buf.n < (size_t)nsend * cnt) c->p2p_send_buf.alloc((size_t)nsend * cnt);
  if (c->p2p_recv_buf.n < (size_t)nrecv * cnt) c->p2p_recv_buf.alloc((size_t)nrecv * cnt);
  T* sendb = reinterpret_cast<T*>(c->p2p_send_buf.p);
  T* recvb = reinterpret_cast<T*>(c->p2p_recv_buf.p);
  SU2B_LAUNCH(c, k_pack<T>, nblk((long)nsend * cnt, 256), 256, nsend, cnt, c->p2p_send_idx.p, arr, sendb);
  SU2B_CUDA(cudaEventRecord(c->ev_packed, c->stream));
  SU2B_CUDA(cudaStreamWaitEvent(c->comm_stream, c->ev_packed, 0));
  auto& api = nccl();
  SU2B_NCCL(api.GroupStart());
  for (int k = 0; k < nn; ++k) {
    const int ns = c->p2p_send_ptr[k + 1] - c->p2p_send_ptr[k], nr = c->p2p_recv_ptr[k + 1] - c->p2p_recv_ptr[k];
    if (nr) SU2B_NCCL(api.Recv(recvb + (size_t)c->p2p_recv_ptr[k] * cnt, (size_t)nr * cnt, NcclType<T>::id, c->p2p_ranks[k], c->nccl, c->comm_stream));
    if (ns) SU2B_NCCL(api.Send(sendb + (size_t)c->p2p_send_ptr[k] * cnt, (size_t)ns * cnt, NcclType<T>::id, c->p2p_ranks[k], c->nccl, c->comm_stream));
  }
  SU2B_NCCL(api.GroupEnd());
  k_unpack<T><<<nblk((long)nrecv * cnt, 256), 256, 0, c->comm_stream>>>(nrecv, cnt, c->p2p_recv_idx.p, recvb, arr);
  SU2B_CUDA(cudaGetLastError());
  ++c->launches;
  SU2B_CUDA(cudaEventRecord(c->ev_unpacked, c->comm_stream));
  c->exchange_in_flight = true;
}
void halo_exchange_start(su2b200_ctx* c, double* arr, int cnt) { halo_exchange_start_impl<double>(c, arr, cnt); }
void halo_exchange_start(su2b200_ctx* c, float* arr, int cnt) { halo_exchange_start_impl<float>(c, arr, cnt); }
void halo_exchange_finish(su2b200_ctx* c) {
  if (!c->exchange_in_flight) return;
  SU2B_CUDA(cudaStreamWaitEvent(c->stream, c->ev_unpacked, 0));
  c->exchange_in_flight = false;
}

// ---- mailbox allreduce over peer memory ----------------------------------------------------------------------------------
// The Gram-Schmidt of FGMRES needs ~100 single-double allreduces per nonlinear iteration, each on the critical path
// (ncclAllReduce in-stream: ~26 us each on 2 GPUs, 2.6 ms of a 70 ms step).  Instead the reduction kernel itself writes
// its partial sum into every peer's mailbox over NVLink/NVSwitch and reads the peers' sums from its own (kernels.hpp:
// mailbox_allreduce_block).  The mailboxes are plain cudaMalloc memory shared with cudaIpc handles, which the host side
// allgathers with whatever it has (torch.distributed in bench.py, MPI in the reference-side shim).
static constexpr size_t MBOX_WORDS = 16 + (size_t)su2b200_ctx::MBOX_SLOTS * su2b200_ctx::MBOX_MAX_RANKS * 2;

bool mailbox_view(su2b200_ctx* c, MailboxView* v) {
  if (!c->mbox_ready) return false;
  v->local = c->mbox_local;
  v->peers = reinterpret_cast<unsigned long long* const*>(c->mbox_peers.p);
  v->rank = c->rank;
  v->nranks = c->nranks;
  return true;
}

void mailbox_release(su2b200_ctx* c) {
  c->mbox_ready = false;
  c->mbox_connected = false;
  for (int r = 0; r < su2b200_ctx::MBOX_MAX_RANKS; ++r) {
    if (c->mbox_peer_host[r] && r != c->rank) cudaIpcCloseMemHandle(c->mbox_peer_host[r]);
    c->mbox_peer_host[r] = nullptr;
  }
  if (c->mbox_local) cudaFree(c->mbox_local);
  c->mbox_local = nullptr;
}

// in-stream, in-place sum of n device doubles over the ranks (no host synchronisation)
void allreduce_dev_sum(su2b200_ctx* c, double* d, int n) {
  if (c->nranks <= 1) return;
  SU2B_REQUIRE(c->nccl, "allreduce requested without a communicator");
  SU2B_NCCL(nccl().AllReduce(d, d, (size_t)n, 8, 0, c->nccl, c->stream));
}
void allreduce_dev_sum(su2b200_ctx* c, float* d, int n) {
  if (c->nranks <= 1) return;
  SU2B_REQUIRE(c->nccl, "allreduce requested without a communicator");
  SU2B_NCCL(nccl().AllReduce(d, d, (size_t)n, 7, 0, c->nccl, c->stream));
}

// in-place allreduce of n (<= 32) host doubles; op: 0 sum, 2 max, 3 min (ncclRedOp_t)
void allreduce_vec(su2b200_ctx* c, double* v, int n, int op) {
  if (c->nranks <= 1) return;
  SU2B_REQUIRE(c->nccl, "allreduce requested without a communicator");
  SU2B_REQUIRE(n <= 32, "allreduce_vec: at most 32 values");
  if (c->p2p_send_buf.n < 64) c->p2p_send_buf.alloc(64);
  double* d = c->p2p_send_buf.p;
  SU2B_CUDA(cudaMemcpyAsync(d, v, n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  SU2B_NCCL(nccl().AllReduce(d, d, (size_t)n, 8, op, c->nccl, c->stream));
  SU2B_CUDA(cudaMemcpyAsync(v, d, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
}

double allreduce_sum(su2b200_ctx* c, double v) {
  if (c->nranks <= 1) return v;
  SU2B_REQUIRE(c->nccl, "allreduce requested without a communicator");
  if (c->p2p_send_buf.n < 2) c->p2p_send_buf.alloc(64);
  double* d = c->p2p_send_buf.p;
  SU2B_CUDA(cudaMemcpyAsync(d, &v, sizeof(double), cudaMemcpyHostToDevice, c->stream));
  SU2B_NCCL(nccl().AllReduce(d, d, 1, 8, 0, c->nccl, c->stream));
  double out = 0;
  SU2B_CUDA(cudaMemcpyAsync(&out, d, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  return out;
}

}  // namespace su2b200

using namespace su2b200;
extern "C" {
int su2b200_comm_unique_id(unsigned char id[128]) {
  try {
    SU2B_NCCL(nccl().GetUniqueId(id));
    return 0;
  } catch (const std::exception& e) { set_error(e.what()); return 1; }
}
int su2b200_comm_mailbox_create(su2b200_ctx* c, unsigned char handle[64]) {
  try {
    if (!c) throw Error("null context");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    SU2B_CUDA(cudaSetDevice(c->device));
    SU2B_REQUIRE(c->nranks > 1, "mailboxes are for multi-rank runs (call su2b200_comm_init first)");
    SU2B_REQUIRE(c->nranks <= su2b200_ctx::MBOX_MAX_RANKS, "mailbox allreduce: more ranks than mailbox entries");
    if (!c->mbox_local) {
      SU2B_CUDA(cudaMalloc((void**)&c->mbox_local, MBOX_WORDS * sizeof(unsigned long long)));
      SU2B_CUDA(cudaMemset(c->mbox_local, 0, MBOX_WORDS * sizeof(unsigned long long)));
    }
    cudaIpcMemHandle_t h;
    SU2B_CUDA(cudaIpcGetMemHandle(&h, c->mbox_local));
    memcpy(handle, &h, 64);
    return 0;
  } catch (const std::exception& e) { set_error(e.what()); return 1; }
}
int su2b200_comm_mailbox_connect(su2b200_ctx* c, const unsigned char* handles) {
  try {
    if (!c) throw Error("null context");
    SU2B_CUDA(cudaSetDevice(c->device));
    SU2B_REQUIRE(c->mbox_local, "su2b200_comm_mailbox_create first");
    std::vector<unsigned long long> table(c->nranks, 0);
    for (int r = 0; r < c->nranks; ++r) {
      if (r == c->rank) { c->mbox_peer_host[r] = c->mbox_local; }
      else {
        cudaIpcMemHandle_t h;
        memcpy(&h, handles + (size_t)r * 64, 64);
        void* p = nullptr;
        const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
          cudaGetLastError();
          throw Error(std::string("cudaIpcOpenMemHandle failed (") + cudaGetErrorString(e) +
                      "): the ranks must be processes on one node with peer access; NCCL allreduces stay in use");
        }
        c->mbox_peer_host[r] = (unsigned long long*)p;
      }
      table[r] = (unsigned long long)(uintptr_t)c->mbox_peer_host[r];
    }
    c->mbox_peers.upload(table);
    SU2B_CUDA(cudaDeviceSynchronize());
    c->mbox_connected = true;  // in use only after su2b200_comm_mailbox_enable: every rank must have got this far
    return 0;
  } catch (const std::exception& e) { set_error(e.what()); return 1; }
}
int su2b200_comm_mailbox_enable(su2b200_ctx* c, int on) {
  try {
    if (!c) throw Error("null context");
    static const bool allowed = !(getenv("SU2B200_MAILBOX") && atoi(getenv("SU2B200_MAILBOX")) == 0);
    c->mbox_ready = on != 0 && c->mbox_connected && allowed;
    return 0;
  } catch (const std::exception& e) { set_error(e.what()); return 1; }
}
int su2b200_comm_init(su2b200_ctx* c, int rank, int size, const unsigned char id[128]) {
  try {
    if (!c) throw Error("null context");
    SU2B_CUDA(cudaSetDevice(c->device));
    c->rank = rank;
    c->nranks = size;
    if (size > 1) {
      NcclId nid;
      memcpy(nid.b, id, 128);
      // ncclCommInitRank(ncclComm_t*, int nranks, ncclUniqueId commId (by value), int rank)
      auto fn = (int (*)(void**, int, NcclId, int))dlsym(nccl().lib, "ncclCommInitRank");
      SU2B_NCCL(fn(&c->nccl, size, nid, rank));
    }
    return 0;
  } catch (const std::exception& e) { set_error(e.what()); return 1; }
}
}
