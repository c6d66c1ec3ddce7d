// a21-a22: Krylov solvers on device-resident vectors.
//   CSysSolve<T>::FGMRES_LinSolver (Common/src/linear_algebra/CSysSolve.cpp:418-613), ModGramSchmidt (:179-240),
//   Generate/ApplyGivens (:132-162), SolveReduced (:165-176), LinearCombination (:58-119)
//   CSysSolve<T>::BCGSTAB_LinSolver (:989-1159), RFGMRES_LinSolver (:616-639), entry CSysSolve<T>::Solve (:1318-1531)
// The Hessenberg/Givens recurrences are host scalars exactly as in the reference; every vector operation,
// preconditioner application and SpMV is a kernel; dot products come back through one pinned double.
#include "internal.hpp"
#include "kernels.hpp"
#include <cmath>
#include <algorithm>

namespace su2b200 {
#define real double
#define LinSysR LinSysT<double>
#define SU2B_F64 1
namespace f64 {
#include "krylov_impl.inc"
}  // namespace f64
#undef real
#undef LinSysR
#undef SU2B_F64
#define real float
#define LinSysR LinSysT<float>
#define SU2B_F64 0
namespace f32 {
#include "krylov_impl.inc"
}  // namespace f32
#undef real
#undef LinSysR
#undef SU2B_F64
}  // namespace su2b200
