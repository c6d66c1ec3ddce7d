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
#include "real_traits.cuh"
#include <algorithm>

namespace su2b200 {
#define real double
#define LinSysR LinSysT<double>
#define SU2B_F64 1
namespace f64 {
#include "precond_poll_impl.inc"
}  // namespace f64
#undef real
#undef LinSysR
#undef SU2B_F64
#define real float
#define LinSysR LinSysT<float>
#define SU2B_F64 0
namespace f32 {
#include "precond_poll_impl.inc"
}  // namespace f32
#undef real
#undef LinSysR
#undef SU2B_F64
}  // namespace su2b200
