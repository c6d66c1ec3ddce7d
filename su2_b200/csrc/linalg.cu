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
#define real double
#define LinSysR LinSysT<double>
#define SU2B_F64 1
namespace f64 {
#include "linalg_impl.inc"
}  // namespace f64
#undef real
#undef LinSysR
#undef SU2B_F64
#define real float
#define LinSysR LinSysT<float>
#define SU2B_F64 0
namespace f32 {
#include "linalg_impl.inc"
}  // namespace f32
#undef real
#undef LinSysR
#undef SU2B_F64
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
static_assert(f64::DOT_BLOCKS + 1 <= 1024, "partials and the result must fit the tail of the scratch array");

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
    const int nb = (int)std::min<size_t>(((size_t)n + 255) / 256, f64::DOT_BLOCKS);
    SU2B_LAUNCH(c, f64::k_dot_stage1, nb, 256, (size_t)n, (const double*)da, (const double*)db, part);
    SU2B_LAUNCH(c, f64::k_dot_stage2, 1, 256, nb, (const double*)part, part + f64::DOT_BLOCKS);
    SU2B_CUDA(cudaMemcpyAsync(result, part + f64::DOT_BLOCKS, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
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
