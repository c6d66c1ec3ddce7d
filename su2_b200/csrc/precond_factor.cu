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
#include "real_traits.cuh"
#include <algorithm>

namespace su2b200 {
#define real double
#define LinSysR LinSysT<double>
#define SU2B_F64 1
namespace f64 {
#include "precond_factor_impl.inc"
}  // namespace f64
#undef real
#undef LinSysR
#undef SU2B_F64
#define real float
#define LinSysR LinSysT<float>
#define SU2B_F64 0
namespace f32 {
#include "precond_factor_impl.inc"
}  // namespace f32
#undef real
#undef LinSysR
#undef SU2B_F64
}  // namespace su2b200
