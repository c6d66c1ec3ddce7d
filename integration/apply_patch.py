#!/usr/bin/env python
"""Patches a SCRATCH COPY of the reference tree (never /root/reference, never this repository) so that SU2_CFD calls
libsu2b200 when the cfg says ENABLE_CUDA= YES.  usage: apply_patch.py <scratch SU2 root> <repo root>

Five small edits, each guarded by #ifdef HAVE_SU2B200 and each anchored on a line of the reference that must be found
exactly once (the script stops otherwise, so a changed reference cannot be patched silently wrong):
  Common/include/linear_algebra/CSysMatrix.hpp        raw read-only accessors of the block-CSR arrays
  Common/include/linear_algebra/CMatrixVectorProduct.hpp   the existing `if (config->GetCUDA())` branch -> backend SpMV
  SU2_CFD/include/solvers/CSolver.hpp                 a setter for Residual_RMS (the device computes the norms)
  SU2_CFD/include/solvers/CFVMFlowSolverBase.inl      System.Solve of ImplicitEuler_Iteration -> backend solve
  SU2_CFD/src/integration/CMultiGridIntegration.cpp   MultiGrid_Cycle of the flow system at MGLEVEL 0 -> backend iteration
plus the backend header itself (integration/CB200FlowBackend.hpp, this repository's code) copied next to CSysMatrix.hpp.
"""
import os
import shutil
import sys


def edit(path, old, new, count=1):
    s = open(path).read()
    if new in s:
        return  # already patched
    n = s.count(old)
    if n != count:
        raise SystemExit("%s: anchor found %d times, expected %d:\n%s" % (path, n, count, old))
    open(path, "w").write(s.replace(old, new))


def main(root, repo):
    la = os.path.join(root, "Common", "include", "linear_algebra")
    shutil.copy(os.path.join(repo, "integration", "CB200FlowBackend.hpp"), os.path.join(la, "CB200FlowBackend.hpp"))
    shutil.copy(os.path.join(repo, "include", "su2b200.h"), os.path.join(la, "su2b200.h"))

    edit(os.path.join(la, "CSysMatrix.hpp"),
         "  FORCEINLINE const ScalarType* GetBlock(unsigned long block_i, unsigned long block_j) const {",
         "#ifdef HAVE_SU2B200\n"
         "  /*--- read-only views for the libsu2b200 binding (CB200FlowBackend.hpp) ---*/\n"
         "  const ScalarType* B200_Values() const { return matrix; }\n"
         "  const unsigned long* B200_RowPtr() const { return row_ptr; }\n"
         "  const unsigned long* B200_ColInd() const { return col_ind; }\n"
         "  unsigned long B200_nnz() const { return nnz; }\n"
         "  unsigned long B200_nVar() const { return nVar; }\n"
         "  unsigned long B200_nPoint() const { return nPoint; }\n"
         "  unsigned long B200_nPointDomain() const { return nPointDomain; }\n"
         "#endif\n"
         "  FORCEINLINE const ScalarType* GetBlock(unsigned long block_i, unsigned long block_j) const {")

    mv = os.path.join(la, "CMatrixVectorProduct.hpp")
    edit(mv, "#ifdef HAVE_CUDA\n      matrix.GPUMatrixVectorProduct(u, v, geometry, config);",
         "#if defined(HAVE_SU2B200)\n      CB200FlowBackend::Get().MatrixVectorProduct(matrix, u, v);\n"
         "#elif defined(HAVE_CUDA)\n      matrix.GPUMatrixVectorProduct(u, v, geometry, config);")
    edit(mv, '#include "CSysMatrix.hpp"', '#include "CSysMatrix.hpp"\n#ifdef HAVE_SU2B200\n#include "CB200FlowBackend.hpp"\n#endif')

    edit(os.path.join(root, "SU2_CFD", "include", "solvers", "CSolver.hpp"),
         "  inline void SetIterLinSolver(unsigned short val_iterlinsolver) { IterLinSolver = val_iterlinsolver; }",
         "#ifdef HAVE_SU2B200\n"
         "  /*--- residual norms computed on the device (su2b200_flow_prepare_implicit) ---*/\n"
         "  inline void B200_SetRes_RMS(unsigned short iVar, su2double val) { Residual_RMS[iVar] = val; }\n"
         "#endif\n"
         "  inline void SetIterLinSolver(unsigned short val_iterlinsolver) { IterLinSolver = val_iterlinsolver; }")

    edit(os.path.join(root, "SU2_CFD", "include", "solvers", "CFVMFlowSolverBase.inl"),
         "  auto iter = System.Solve(Jacobian, LinSysRes, LinSysSol, geometry, config);\n\n"
         "  BEGIN_SU2_OMP_SAFE_GLOBAL_ACCESS {\n    SetIterLinSolver(iter);\n    SetResLinSolver(System.GetResidual());",
         "#ifdef HAVE_SU2B200\n"
         "  unsigned long iter = 0;\n"
         "  su2double b200Res = 0.0;\n"
         "  const bool b200 = config->GetCUDA();\n"
         "  if (b200) {\n"
         "    BEGIN_SU2_OMP_SAFE_GLOBAL_ACCESS { iter = CB200FlowBackend::Get().Solve(Jacobian, LinSysRes, LinSysSol, config, &b200Res); }\n"
         "    END_SU2_OMP_SAFE_GLOBAL_ACCESS\n"
         "  } else {\n"
         "    iter = System.Solve(Jacobian, LinSysRes, LinSysSol, geometry, config);\n"
         "  }\n"
         "#else\n"
         "  auto iter = System.Solve(Jacobian, LinSysRes, LinSysSol, geometry, config);\n"
         "#endif\n\n"
         "  BEGIN_SU2_OMP_SAFE_GLOBAL_ACCESS {\n    SetIterLinSolver(iter);\n"
         "#ifdef HAVE_SU2B200\n    SetResLinSolver(b200 ? b200Res : System.GetResidual());\n#else\n"
         "    SetResLinSolver(System.GetResidual());\n#endif")

    edit(os.path.join(root, "SU2_CFD", "include", "solvers", "CFVMFlowSolverBase.inl"),
         '#include "CFVMFlowSolverBase.hpp"',
         '#include "CFVMFlowSolverBase.hpp"\n#ifdef HAVE_SU2B200\n'
         '#include "../../../Common/include/linear_algebra/CB200FlowBackend.hpp"\n#endif')

    mg = os.path.join(root, "SU2_CFD", "src", "integration", "CMultiGridIntegration.cpp")
    edit(mg,
         "  MultiGrid_Cycle(geometry, solver_container, numerics_container, config,\n"
         "                  FinestMesh, RecursiveParam, RunTime_EqSystem, iZone, iInst);\n\n\n"
         "  /*--- Computes primitive variables and gradients in the finest mesh",
         "#ifdef HAVE_SU2B200\n"
         "  if (config[iZone]->GetCUDA() && CB200FlowBackend::WholeIteration() && RunTime_EqSystem == RUNTIME_FLOW_SYS &&\n"
         "      config[iZone]->GetnMGLevels() == 0) {\n"
         "    /*--- the implicit iteration of the flow solver on the device, CSolver virtual by CSolver virtual ---*/\n"
         "    BEGIN_SU2_OMP_SAFE_GLOBAL_ACCESS {\n"
         "      CSolver** solvers = solver_container[iZone][iInst][MESH_0];\n"
         "      std::vector<double> rms;\n"
         "      su2double linRes = 0.0;\n"
         "      const auto iters = CB200FlowBackend::Get().FlowIteration(geometry[iZone][iInst][MESH_0], solvers, config[iZone],\n"
         "                                                              FLOW_SOL, TURB_SOL, rms, &linRes);\n"
         "      for (unsigned short iVar = 0; iVar < rms.size(); ++iVar) solvers[FLOW_SOL]->B200_SetRes_RMS(iVar, rms[iVar]);\n"
         "      solvers[FLOW_SOL]->SetIterLinSolver(iters);\n"
         "      solvers[FLOW_SOL]->SetResLinSolver(linRes);\n"
         "    }\n"
         "    END_SU2_OMP_SAFE_GLOBAL_ACCESS\n"
         "  } else\n"
         "#endif\n"
         "  MultiGrid_Cycle(geometry, solver_container, numerics_container, config,\n"
         "                  FinestMesh, RecursiveParam, RunTime_EqSystem, iZone, iInst);\n\n\n"
         "  /*--- Computes primitive variables and gradients in the finest mesh")
    edit(mg, '#include "../../include/integration/CMultiGridIntegration.hpp"',
         '#include "../../include/integration/CMultiGridIntegration.hpp"\n'
         '#ifdef HAVE_SU2B200\n#include "../../../Common/include/linear_algebra/CB200FlowBackend.hpp"\n#endif')
    print("patched", root)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
