// a17-a19: ILU(0) and LU-SGS on the device.
//   CSysMatrix<T>::BuildILUPreconditioner   (Common/src/linear_algebra/CSysMatrix.cpp:673-767)
//   CSysMatrix<T>::ComputeILUPreconditioner (:770-823)
//   CSysMatrix<T>::ComputeLU_SGSPreconditioner (:826-888), Lower/Upper/DiagonalProduct (CSysMatrix.inl:184-211)
//
// The reference factorises the owned sub-matrix [0, nPointDomain) per MPI rank and, with OpenMP, per
// nnz-balanced row partition [begin, end) (block-Jacobi between partitions, CSysMatrix.cpp:199-227, :701-711).
// Two schedules are provided for the same arithmetic:
//   exact (ilu_exact=1): one partition == the serial reference; rows are grouped by dependency level
//                        (level-scheduled triangular sweeps), one warp per row, one launch per level;
//   partitioned        : K partitions as the reference does with K threads; one warp walks one partition
//                        sequentially, all partitions in one launch.
// In both, a warp owns a row: lane q = r*nVar + c holds one scalar of a block, products use shuffles.
#include "internal.hpp"
#include "kernels.hpp"
#include <algorithm>
#include <cstdlib>

namespace su2b200 {
#define real double
#define LinSysR LinSysT<double>
#define SU2B_F64 1
namespace f64 {
#include "precond_impl.inc"
}  // namespace f64
#undef real
#undef LinSysR
#undef SU2B_F64
#define real float
#define LinSysR LinSysT<float>
#define SU2B_F64 0
namespace f32 {
#include "precond_impl.inc"
}  // namespace f32
#undef real
#undef LinSysR
#undef SU2B_F64
}  // namespace su2b200
