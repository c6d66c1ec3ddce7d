/*
 * su2b200.h — C ABI of libsu2b200.so: the B200-native (sm_100a) implementation of SU2's
 * implicit finite-volume flow iteration (edge residual/Jacobian assembly + preconditioned
 * Krylov solve). This is the drop-in boundary: every entry point names the reference
 * interface (su2code/SU2 v8.4.0, paths relative to the reference tree) it replaces.
 *
 * Conventions
 *  - plain C types only; index arrays use `unsigned long` exactly like the reference
 *    (CGeometry / CSysMatrix store `unsigned long`), converted to int32 on upload;
 *  - every function returns 0 on success, non-zero on error; su2b200_last_error() gives
 *    the message (the C++ shim turns it into SU2_MPI::Error, mpi_structure.hpp:111);
 *  - the library owns all device memory behind the opaque handle; host pointers are read
 *    or written during the call only;
 *  - fp64 everywhere (su2double = double, code_config.hpp:128);
 *  - one context == one MPI rank == one GPU; single caller thread per context.
 *  - there is NO CPU fallback: every compute call runs CUDA kernels or fails.
 */
#ifndef SU2B200_H
#define SU2B200_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct su2b200_ctx su2b200_ctx;

/* ---- enums (values follow Common/include/option_structure.hpp where one exists) ---- */

/* BC_TYPE, option_structure.hpp:1654-1688 */
enum { SU2B200_EULER_WALL = 1, SU2B200_FAR_FIELD = 2, SU2B200_SYMMETRY_PLANE = 3,
       SU2B200_HEAT_FLUX = 26, SU2B200_SEND_RECEIVE = 99 };

/* convective scheme: CONV_NUM_METHOD_FLOW */
enum { SU2B200_CONV_ROE = 0, SU2B200_CONV_JST = 1 };
/* NUM_METHOD_GRAD / NUM_METHOD_GRAD_RECON */
enum { SU2B200_GRAD_GREEN_GAUSS = 0, SU2B200_GRAD_LEAST_SQUARES = 1, SU2B200_GRAD_WEIGHTED_LEAST_SQUARES = 2 };
/* SLOPE_LIMITER_FLOW */
enum { SU2B200_LIMITER_NONE = 0, SU2B200_LIMITER_VENKATAKRISHNAN = 1, SU2B200_LIMITER_BARTH_JESPERSEN = 2 };
/* LINEAR_SOLVER */
enum { SU2B200_SOLVER_FGMRES = 0, SU2B200_SOLVER_BCGSTAB = 1, SU2B200_SOLVER_RESTARTED_FGMRES = 2 };
/* LINEAR_SOLVER_PREC */
enum { SU2B200_PREC_JACOBI = 0, SU2B200_PREC_LU_SGS = 1, SU2B200_PREC_ILU = 2 };
/* which linear system of the context */
enum { SU2B200_SYS_FLOW = 0, SU2B200_SYS_TURB = 1 };
/* how the edge loop accumulates into points (both are atomics-free and deterministic):
 * GATHER  = edge kernel stores fluxes + off-diagonal blocks, a point kernel sums them in
 *           ascending edge order (the reference's "reducer strategy",
 *           CFVMFlowSolverBase.inl:1575-1591 + CSysMatrix::SetDiagonalAsColumnSum);
 * COLORED = one launch per edge colour with read-modify-write of residual/diagonal
 *           (CFVMFlowSolverBase.inl:1543-1570 with CGeometry::GetEdgeColoring). */
enum { SU2B200_ASSEMBLY_GATHER = 0, SU2B200_ASSEMBLY_COLORED = 1 };

/* named device arrays that can be read back / overwritten (parity tests, restart, output) */
enum {
  SU2B200_ARR_SOLUTION = 0,      /* CVariable::Solution            [P][nVar]            */
  SU2B200_ARR_SOLUTION_OLD = 1,  /* CVariable::Solution_Old        [P][nVar]            */
  SU2B200_ARR_PRIMITIVE = 2,     /* CFlowVariable::Primitive       [P][nDim+9]          */
  SU2B200_ARR_GRAD_RECON = 3,    /* Gradient_Reconstruction        [P][nDim+2][nDim]    */
  SU2B200_ARR_GRAD_PRIM = 4,     /* Gradient_Primitive             [P][nDim+2][nDim]    */
  SU2B200_ARR_LIMITER = 5,       /* Limiter_Primitive              [P][nDim+2]          */
  SU2B200_ARR_DELTA_TIME = 6,    /* CVariable::Delta_Time          [P]                  */
  SU2B200_ARR_LINSYSRES = 7,     /* CSolver::LinSysRes             [P][nVar]            */
  SU2B200_ARR_LINSYSSOL = 8,     /* CSolver::LinSysSol             [P][nVar]            */
  SU2B200_ARR_JACOBIAN = 9,      /* CSysMatrix::matrix             [nnz][nVar][nVar]    */
  SU2B200_ARR_LAMBDA = 10,       /* CEulerVariable::Lambda         [P]                  */
  SU2B200_ARR_SENSOR = 11,       /* CEulerVariable::Sensor         [P]                  */
  SU2B200_ARR_UND_LAPL = 12,     /* CEulerVariable::Undivided_Laplacian [P][nVar]       */
  SU2B200_ARR_MAX_LAMBDA_INV = 13,
  SU2B200_ARR_MAX_LAMBDA_VISC = 14,
  SU2B200_ARR_ILU = 15,          /* CSysMatrix::ILU_matrix         [nnz][nVar][nVar]    */
  SU2B200_ARR_INVM = 16,         /* CSysMatrix::invM               [Pd][nVar][nVar]     */
  SU2B200_ARR_UNDER_RELAX = 17,  /* CVariable::UnderRelaxation     [P]                  */
  SU2B200_ARR_LOCAL_CFL = 19,    /* CVariable::LocalCFL            [P] (empty unless cfl_adapt)                         */
  SU2B200_ARR_RES_TRUNC = 18,    /* CVariable::Res_TruncError      [P][nVar] (multigrid forcing term; empty without) */
  SU2B200_ARR_TURB_SOLUTION = 20,/* CTurbSAVariable::Solution      [P][1]               */
  SU2B200_ARR_TURB_GRADIENT = 21,/* CScalarVariable::Gradient      [P][1][nDim]         */
  SU2B200_ARR_TURB_LINSYSRES = 22,
  SU2B200_ARR_TURB_LINSYSSOL = 23,
  SU2B200_ARR_TURB_JACOBIAN = 24,
  SU2B200_ARR_EDDY_VISC = 25,    /* CTurbVariable::muT             [P]                  */
  SU2B200_ARR_VORTICITY = 26,    /* CFlowVariable::Vorticity       [P][3]               */
  SU2B200_ARR_STRAIN_MAG = 27,   /* CFlowVariable::StrainMag       [P]                  */
  SU2B200_ARR_WALL_DIST = 28,    /* CPoint::Wall_Distance          [P]                  */
  SU2B200_ARR_TURB_DELTA_TIME = 29,
  SU2B200_ARR_TURB_SOLUTION_OLD = 30,
  SU2B200_ARR_SST_F1 = 31,       /* CTurbSSTVariable::F1, F2, CDkw [P] each (SST only) */
  SU2B200_ARR_SST_F2 = 32,
  SU2B200_ARR_SST_CDKW = 33
};

/* ---- lifecycle --------------------------------------------------------------------- */

/* cudaSetDevice(device) + streams. Replaces the gpu_alloc calls in CSysMatrix::Initialize
 * (Common/src/linear_algebra/CSysMatrix.cpp:142-160) and CSysVector::Initialize (CSysVector.cpp:55). */
int su2b200_create(int device, su2b200_ctx** out);
int su2b200_destroy(su2b200_ctx* ctx);
/* message of the last failing call on this thread (GPUComms.cuh:46-53 gpuErrChk prints+exits instead) */
const char* su2b200_last_error(void);
int su2b200_version(void);
int su2b200_synchronize(su2b200_ctx* ctx);
/* sizeof() of the three POD structs of this header, so that FFI bindings can verify their layout */
int su2b200_struct_sizes(int* geometry_desc, int* flow_config, int* iter_report);

/* ---- geometry: CGeometry dual grid + sparse pattern --------------------------------- */

typedef struct {
  int nDim;                          /* CGeometry::GetnDim()                                   */
  unsigned long nPoint;              /* GetnPoint() (owned + halo)                             */
  unsigned long nPointDomain;        /* GetnPointDomain()                                      */
  unsigned long nEdge;               /* GetnEdge()                                             */
  const unsigned long* edge_node0;   /* CEdge::Nodes column 0 [nEdge]  (CEdge.hpp:44, col-major)*/
  const unsigned long* edge_node1;   /* CEdge::Nodes column 1 [nEdge]                          */
  const double* edge_normal;         /* CEdge::Normal [nEdge][nDim] row-major (CEdge.hpp:46)   */
  const double* coord;               /* CPoint::Coord [nPoint][nDim]  (CPoint.hpp)             */
  const double* volume;              /* CPoint::Volume [nPoint]                                */
  const unsigned long* point_ptr;    /* CPoint::Point outer ptr [nPoint+1] (neighbour CSR)     */
  const unsigned long* point_nbr;    /* CPoint::Point inner idx [2*nEdge], reference order     */
  const unsigned long* point_edge;   /* CPoint::Edge  inner idx [2*nEdge]                      */
  const unsigned long* row_ptr;      /* CGeometry::GetSparsePattern outer ptr [nPoint+1] (CGeometry.cpp:3702) */
  const unsigned long* col_ind;      /* inner idx [nnz], ascending, diagonal included          */
  const unsigned long* dia_ptr;      /* diagonal position [nPoint]                             */
  const unsigned long* edge_ptr;     /* GetEdgeToSparsePatternMap [nEdge][2] row-major (CGeometry.cpp:3720) */
  const unsigned short* nNeighbor;   /* CPoint::nNeighbor [nPoint] (JST), may be NULL => from point_ptr */
  const unsigned char* physical_boundary; /* CPoint::PhysicalBoundary [nPoint] (JST), may be NULL => derived from markers */
  const double* wall_distance;       /* CPoint::Wall_Distance [nPoint], NULL unless RANS        */
  const unsigned long* color_ptr;    /* GetEdgeColoring outer ptr [nColor+1], NULL => natural/own colouring */
  const unsigned long* color_edges;  /* edge indices by colour [nEdge]                          */
  unsigned long nColor;
} su2b200_geometry_desc;

/* Upload once after CSolver construction. */
int su2b200_geometry_upload(su2b200_ctx* ctx, const su2b200_geometry_desc* g);

/* One boundary marker: geometry->vertex[iMarker][iVertex]->{GetNode,GetNormal,GetNormal_Neighbor}
 * and, for SYMMETRY_PLANE/EULER_WALL, the (possibly modified) unit normal of
 * CGeometry::symmetryNormals (CGeometry.cpp:2469-2585); sym_unit_normal may be NULL => normal/|normal|.
 * monitoring != 0 marks the marker as part of MARKER_MONITORING (force integration).
 * heat_flux is MARKER_HEATFLUX's value / Heat_Flux_Ref (CNSSolver.cpp:440). */
int su2b200_marker_upload(su2b200_ctx* ctx, int iMarker, int kindBC, unsigned long nVertex,
                          const unsigned long* vertex_node, const double* vertex_normal,
                          const unsigned long* vertex_normal_neighbor, const double* sym_unit_normal,
                          int monitoring, double heat_flux);

/* CFVMFlowSolverBase::CharacPrimVar of a marker (CFVMFlowSolverBase.hpp:176): the far-field state BC_Far_Field computed
 * for every vertex, [nVertex][nDim+9]; the scalar solvers' far-field fluxes read it (CScalarSolver.inl:382-441) */
int su2b200_marker_charac_download(su2b200_ctx* ctx, int iMarker, double* host);

/* Halo tables of CGeometry::PreprocessP2PComms (CGeometry.cpp:139-332): neighbour ranks, cumulative
 * point counts and the local point lists for send/recv. */
int su2b200_p2p_upload(su2b200_ctx* ctx, int nNeighborRanks, const int* neighbor_rank,
                       const int* nPoint_P2PSend /*[n+1] cumulative*/, const unsigned long* Local_Point_P2PSend,
                       const int* nPoint_P2PRecv /*[n+1] cumulative*/, const unsigned long* Local_Point_P2PRecv);

/* ---- host-side geometry preprocessing: the part of CGeometry/CPhysicalGeometry the path consumes (host C++, no GPU needed) ----
 * Produces exactly the arrays of su2b200_geometry_desc / su2b200_marker_upload / su2b200_p2p_upload. */
typedef struct su2b200_mesh su2b200_mesh;
/* MESH_FORMAT= BOX / RECTANGLE (CBoxMeshReaderFVM.cpp:31-330, CRectangularMeshReaderFVM.cpp): n points per direction,
 * markers x_minus, x_plus, y_minus, y_plus[, z_minus, z_plus]; the local part of `rank` in a slab decomposition along the
 * last axis (owned points first, one halo plane per neighbour, halo tables as CGeometry::PreprocessP2PComms builds them).
 * tets != 0 splits every hexahedron into 6 tetrahedra (quadrilateral into 2 triangles). */
int su2b200_mesh_create_box(int nDim, const unsigned long n[3], const double length[3], const double offset[3],
                            int rank, int nranks, int tets, su2b200_mesh** out);
/* MESH_FORMAT= SU2 (CSU2ASCIIMeshReaderBase.cpp): NDIME / NELEM / NPOIN / NMARK blocks, VTK element ids */
int su2b200_mesh_read_su2(const char* path, su2b200_mesh** out);
/* .su2 ASCII writer, the inverse of su2b200_mesh_read_su2 (same blocks; layout of QuickStart/mesh_NACA0012_inv.su2:1-5):
 * how the tetrahedral and the stretched synthetic boxes reach the reference for the parity fixtures (SURVEY 8(d)(3,4)) */
int su2b200_mesh_write_su2(su2b200_mesh* mesh, const char* path);
/* geometric stretching of a BOX/RECTANGLE mesh along `axis` before preprocessing: cell sizes first_cell, first_cell*ratio,
 * ... from the minus side, rescaled to the box length (the stretched wall-normal spacing of the "wing-box" input) */
int su2b200_mesh_stretch(su2b200_mesh* mesh, int axis, double first_cell, double ratio);
/* Decomposition of an undivided, not yet preprocessed mesh over nranks (SetColorGrid_Parallel + SetSendReceive,
 * CPhysicalGeometry.cpp:7675-7755, :2956): point colouring by weighted recursive coordinate bisection with the reference's
 * vertex weights (wp + we * degree), then per rank the owned points, the one-layer halo, every element touching an owned
 * point, and the send / receive lists of CGeometry::PreprocessP2PComms. rank >= 0: that rank's local mesh;
 * rank = -1: the whole mesh renumbered colour by colour, part_ptr[nranks+1] = first point of every colour (the row
 * partitions with which one GPU reproduces the nranks-rank block-Jacobi preconditioner). Marker settings are inherited. */
int su2b200_mesh_partition(su2b200_mesh* whole, int nranks, int rank, su2b200_mesh** out, unsigned long* part_ptr);
int su2b200_mesh_destroy(su2b200_mesh* mesh);
/* MARKER_EULER / MARKER_FAR / MARKER_SYM / MARKER_HEATFLUX + MARKER_MONITORING of the cfg file */
int su2b200_mesh_set_marker(su2b200_mesh* mesh, const char* tag, int kindBC, int monitoring, double heat_flux);
/* SetRCM_Ordering (CPhysicalGeometry.cpp:4478, if rcm != 0), SetEdges (CGeometry.cpp:1263), SetControlVolume
 * (CPhysicalGeometry.cpp:7059), SetBoundControlVolume (:7212), FindNormal_Neighbor (:8036), ComputeModifiedSymmetryNormals
 * (CGeometry.cpp:2469), ComputeWall_Distance, GetSparsePattern (:3702), GetEdgeToSparsePatternMap (:3720), GetEdgeColoring (:3737) */
int su2b200_mesh_preprocess(su2b200_mesh* mesh, int rcm);
int su2b200_mesh_sizes(su2b200_mesh* mesh, int* nDim, unsigned long* nPoint, unsigned long* nPointDomain,
                       unsigned long* nEdge, unsigned long* nElem, unsigned long* nnz, int* nMarker);
/* read-only view of a named array ("coord", "volume", "edge_normal", "edge_node0", "row_ptr", "marker0_normal", ...);
 * dtype: 0 = double, 1 = unsigned long, 2 = unsigned short, 3 = unsigned char, 4 = int */
int su2b200_mesh_array(su2b200_mesh* mesh, const char* name, const void** data, unsigned long* count, int* dtype);
int su2b200_mesh_marker_info(su2b200_mesh* mesh, int iMarker, char tag[64], int* kindBC, int* monitoring,
                             unsigned long* nVertex, unsigned long* nElemBound);
/* geometry_upload + marker_upload for every marker + p2p_upload */
int su2b200_mesh_upload(su2b200_ctx* ctx, su2b200_mesh* mesh);

/* ---- configuration: the CConfig getters the hot path reads, as a POD ------------------ */

typedef struct {
  /* fluid (CIdealGas.cpp:40-52, CSutherland.hpp:47-56, CConstantPrandtl) -- non-dimensional values */
  double gamma, gas_constant;
  int viscous;                       /* NAVIER_STOKES / RANS                                  */
  int rans_sa;                       /* KIND_TURB_MODEL= SA                                   */
  int sa_negative;                   /* SA_OPTIONS= NEGATIVE                                  */
  double mu_ref, mu_temperature_ref, mu_sutherland; /* Sutherland non-dim; mu_constant if mu_sutherland<0 */
  double mu_constant;
  double prandtl_lam, prandtl_turb;
  /* free stream (CEulerSolver::GetDensity_Inf etc.) */
  double density_inf, pressure_inf, velocity_inf[3], energy_inf, temperature_inf;
  double nu_tilde_inf, eddy_visc_inf;
  /* numerics */
  int conv_scheme;                   /* CONV_NUM_METHOD_FLOW                                  */
  int muscl;                         /* MUSCL_FLOW                                            */
  double muscl_kappa;                /* MUSCL_KAPPA_FLOW (CConfig.cpp:2015)                   */
  int limiter;                       /* SLOPE_LIMITER_FLOW                                    */
  double venkat_coeff;               /* VENKAT_LIMITER_COEFF                                  */
  double ref_elem_length;            /* REF_ELEM_LENGTH                                       */
  int grad_method, grad_method_recon;/* NUM_METHOD_GRAD, NUM_METHOD_GRAD_RECON                */
  double roe_kappa;                  /* ROE_KAPPA (0.5)                                       */
  double entropy_fix;                /* ENTROPY_FIX_COEFF                                     */
  double jst_kappa2, jst_kappa4;     /* JST_SENSOR_COEFF                                      */
  double cent_jac_fix_factor;        /* CENTRAL_JACOBIAN_FIX_FACTOR                           */
  /* time integration */
  double cfl;                        /* CFL_NUMBER (LocalCFL of every point)                  */
  double cfl_turb_ratio;             /* CFL_REDUCTION_TURB                                    */
  double max_delta_time;             /* MAX_DELTA_TIME                                        */
  double max_update_flow;            /* MAX_UPDATE_FLOW (0.2)                                 */
  double max_update_sa;              /* MAX_UPDATE_SA (0.99)                                  */
  /* linear solver */
  int linear_solver, linear_prec;    /* LINEAR_SOLVER, LINEAR_SOLVER_PREC                     */
  double linear_tol;                 /* LINEAR_SOLVER_ERROR                                   */
  int linear_iter;                   /* LINEAR_SOLVER_ITER                                    */
  int linear_restart;                /* LINEAR_SOLVER_RESTART_FREQUENCY                       */
  int turb_linear_solver, turb_linear_prec, turb_linear_iter;
  double turb_linear_tol;
  /* forces (CFVMFlowSolverBase::Pressure_Forces) */
  double aoa_deg, aos_deg, ref_area, ref_length, ref_origin[3];
  /* implementation choices (not CConfig) */
  int assembly;                      /* SU2B200_ASSEMBLY_*                                    */
  int prec_partitions;               /* <=1: level-scheduled ILU0/LU-SGS == the reference's serial sweeps;
                                        K>1: K nnz-balanced row partitions, block-Jacobi between them, exactly what the
                                        reference does with LINEAR_SOLVER_PREC_THREADS=K (CSysMatrix.cpp:199-227,701-711) */
  int turb_reducer;                  /* CScalarSolver::ReducerStrategy (CScalarSolver.inl:59): with it the two one-sided
                                        SA diffusion Jacobians of an edge are AVERAGED into the off-diagonal blocks and
                                        the diagonal is minus the column sum (CScalarSolver.hpp:196-203, .inl:318-320).
                                        The reference switches it on when the "colouring efficiency" it computes for
                                        the scalar solver is below 0.875 -- in a serial run that is
                                        nEdge / (512 ceil(nEdge/512)) (CGeometry.cpp:3826, graph_toolbox.hpp:661), so
                                        small meshes with an unlucky edge count get it. 1 / 0 = on / off, -1 = that rule */
  /* CFL adaptation (CSolver::AdaptCFLNumber, CSolver.cpp:1707-1949) */
  int cfl_adapt;                     /* CFL_ADAPT                                                                      */
  double cfl_adapt_param[6];         /* CFL_ADAPT_PARAM: factor down, factor up, CFL min, CFL max, acceptable linear
                                        tolerance (0.001), starting iteration (0)                                      */
  unsigned long n_inner_iter;        /* ITER: the stagnation window is min(100, ITER - 1) iterations                    */
  /* FAS multigrid (CMultiGridIntegration): the level this context represents */
  int mg_level;                      /* 0 = finest grid; > 0: agglomerated level (no MUSCL, no JST sensor/Laplacian:
                                        CEulerSolver.cpp:1584-1586, :1662; centred schemes use CLaxScheme, CNumericsSIMD.cpp:70) */
  double lax_kappa1;                 /* LAX_SENSOR_COEFF (CLaxScheme, centered.hpp:474-530): 1st-order dissipation of the coarse levels */
  /* Menter SST, version V2003m = the reference's default (CTurbSSTSolver.cpp:40-200, option_structure.hpp:1036-1100).
   * The turbulence arrays then hold two variables per point (k, omega): TURB_SOLUTION [P][2], TURB_GRADIENT [P][2][nDim],
   * TURB_LINSYSRES/SOL [P][2], TURB_JACOBIAN [nnz][2][2]. */
  int rans_sst;                      /* KIND_TURB_MODEL= SST (exclusive with rans_sa)                                  */
  double kine_inf, omega_inf;        /* Solution_Inf: 3/2 |v_inf|^2 I^2 and rho_inf k_inf / (mu_inf ratio)             */
  double sst_lower[2], sst_upper[2]; /* lowerlimit / upperlimit of (k, omega): 1e-10, 1e-4 / 1e10, 1e15 without DLL    */
  double max_update_sst;             /* MAX_UPDATE_SST (1.0)                                                           */
  /* The reference's mixed-precision build (meson -Denable-mixedprec=true: su2mixedfloat = float, code_config.hpp:133-139;
   * CSysSolve::HandleTemporariesIn/Out, CSysSolve.hpp:242-286): Jacobian, preconditioner and all Krylov vectors of the
   * flow and turbulence solves in single precision, residual assembly and update in double. */
  int mixed_precision;
} su2b200_flow_config;

int su2b200_flow_configure(su2b200_ctx* ctx, const su2b200_flow_config* cfg);

/* ---- state ---------------------------------------------------------------------------- */

/* Allocates CVariable-equivalent device arrays and sets every point to the free stream
 * (CEulerVariable ctor, CEulerSolver.cpp:293-297). Needs geometry + config. */
int su2b200_flow_allocate(su2b200_ctx* ctx);
/* element count of a named array (doubles) */
long su2b200_array_size(su2b200_ctx* ctx, int which);
int su2b200_array_download(su2b200_ctx* ctx, int which, double* host);
int su2b200_array_upload(su2b200_ctx* ctx, int which, const double* host);
/* NonPhysicalEdgeCounter (CFlowVariable.hpp:80-87), int8 per edge */
int su2b200_edge_counter_download(su2b200_ctx* ctx, signed char* host);

/* ---- CSolver virtuals of the flow solver (one call == one reference call) --------------- */

/* CEulerSolver::Preprocessing (CEulerSolver.cpp:1667) / CNSSolver::Preprocessing (CNSSolver.cpp:70):
 * primitives, [JST] lambda/sensor/laplacian, zero residual+Jacobian, [MUSCL] recon gradient + limiter,
 * [viscous] primitive gradient, vorticity/strain. nonphysical = CEulerSolver::ErrorCounter. */
int su2b200_flow_preprocessing(su2b200_ctx* ctx, int output, unsigned long* nonphysical_points);
/* CSolver::Set_OldSolution */
int su2b200_flow_set_old_solution(su2b200_ctx* ctx);
/* CEulerSolver::SetTime_Step (CFVMFlowSolverBase.hpp:377-620) */
int su2b200_flow_set_time_step(su2b200_ctx* ctx, double* min_dt, double* max_dt);
/* CEulerSolver::Upwind_Residual / Centered_Residual -> EdgeFluxResidual (CFVMFlowSolverBase.inl:1517-1573)
 * with the CNumericsSIMD functor chain (roe.hpp, centered.hpp, viscous_fluxes.hpp). */
int su2b200_flow_edge_residual(su2b200_ctx* ctx, unsigned long* nonphysical_edges);
/* CEulerSolver::BC_Far_Field (CEulerSolver.cpp:4780-5034) */
int su2b200_flow_bc_far_field(su2b200_ctx* ctx, int iMarker);
/* CFVMFlowSolverBase::BC_Sym_Plane == BC_Euler_Wall (CFVMFlowSolverBase.inl:1114-1272) */
int su2b200_flow_bc_sym_plane(su2b200_ctx* ctx, int iMarker);
/* CNSSolver::BC_HeatFlux_Wall (CNSSolver.cpp:423-578) */
int su2b200_flow_bc_heatflux_wall(su2b200_ctx* ctx, int iMarker);
/* CIntegration::Space_Integration (CIntegration.cpp:37-201): edge residual + all BCs in the reference's order */
int su2b200_flow_space_integration(su2b200_ctx* ctx);
/* CFVMFlowSolverBase::PrepareImplicitIteration_impl (CFVMFlowSolverBase.hpp:912-973);
 * rms[nVar] = CSolver::Residual_RMS (sqrt(sum r^2 / N)), max[nVar], idx[nVar] = local point of the max */
int su2b200_flow_prepare_implicit(su2b200_ctx* ctx, double* rms, double* max, unsigned long* idx);
/* System.Solve(Jacobian, LinSysRes, LinSysSol) (CSysSolve.cpp:1318-1531, call site CFVMFlowSolverBase.inl:608) */
int su2b200_flow_linear_solve(su2b200_ctx* ctx, unsigned long* iterations, double* residual);
/* CFVMFlowSolverBase::CompleteImplicitIteration_impl (CFVMFlowSolverBase.hpp:983-1008) incl. under-relaxation */
int su2b200_flow_complete_implicit(su2b200_ctx* ctx);
/* CSolver::AdaptCFLNumber (CSolver.cpp:1707-1949): per-point CVariable::LocalCFL scaled down / up from the linear solver
 * residuals, the under-relaxation factors of the last update and the nonlinear residual series (flip-flop / stall
 * detection); su2b200_iterate calls it after the turbulence iteration when cfl_adapt is set. min/max/avg of the new CFL. */
int su2b200_flow_adapt_cfl(su2b200_ctx* ctx, unsigned long inner_iter, double* cfl_min, double* cfl_max, double* cfl_avg);
/* CFVMFlowSolverBase::Pressure_Forces (+Friction_Forces) on MARKER_MONITORING: CL, CD, CSF, CMx..z, CFx..z */
int su2b200_flow_forces(su2b200_ctx* ctx, double* coeffs /*[9]: CL CD CSF CFx CFy CFz CMx CMy CMz*/);

/* ---- CTurbSASolver virtuals -------------------------------------------------------------- */
int su2b200_turb_preprocessing(su2b200_ctx* ctx);        /* CTurbSASolver::Preprocessing (CTurbSASolver.cpp:176) */
int su2b200_turb_set_time_step(su2b200_ctx* ctx);        /* CScalarSolver::SetTime_Step (CScalarSolver.inl:443) */
int su2b200_turb_space_integration(su2b200_ctx* ctx);    /* Upwind_/Viscous_/Source_Residual + BCs */
int su2b200_turb_prepare_implicit(su2b200_ctx* ctx, double* rms, double* max, unsigned long* idx);
int su2b200_turb_linear_solve(su2b200_ctx* ctx, unsigned long* iterations, double* residual);
int su2b200_turb_complete_implicit(su2b200_ctx* ctx);
int su2b200_turb_postprocessing(su2b200_ctx* ctx);       /* CTurbSASolver::Postprocessing (mu_t) */

/* ---- whole iterations (CMultiGridIntegration::MultiGrid_Iteration at MGLEVEL=0 +
 *      CSingleGridIntegration::SingleGrid_Iteration(TURB_SOL)), state stays on the device --------- */
typedef struct {
  double rms_flow[5];          /* log10 not applied: CSolver::GetRes_RMS                       */
  double rms_turb;
  double CL, CD;
  unsigned long linsol_iter_flow, linsol_iter_turb;
  double linsol_res_flow, linsol_res_turb;
  unsigned long nonphysical_points, nonphysical_edges;
  double rms_turb2;            /* second turbulence variable (SST: omega); 0 for SA */
} su2b200_iter_report;
int su2b200_iterate(su2b200_ctx* ctx, int nIter, su2b200_iter_report* reports /*[nIter], may be NULL*/);

/* ---- FAS multigrid: CMultiGridIntegration (SU2_CFD/src/integration/CMultiGridIntegration.cpp) ---------------------
 * One context per level: level 0 is the finest grid, level k > 0 an agglomerated grid uploaded like any other geometry
 * (CMultiGridGeometry is host preprocessing outside the hot path) and configured with mg_level = k (first-order upwind /
 * CLaxScheme, no JST sensor, its own CFL: CConfig::GetCFL(iMesh)). All levels on one device, single rank. */
/* CPoint::Children_CV of the coarse level: children_ptr[nPointCoarse+1], children[] = points of `fine` */
int su2b200_mg_link(su2b200_ctx* fine, su2b200_ctx* coarse, const unsigned long* children_ptr,
                    const unsigned long* children);
/* SetRestricted_Solution (:563-608): coarse Solution = volume-weighted average of the children, zero velocity on no-slip walls */
int su2b200_mg_set_restricted_solution(su2b200_ctx* fine, su2b200_ctx* coarse);
/* SetResidual_Term (:552-561): LinSysRes += Res_TruncError */
int su2b200_mg_set_residual_term(su2b200_ctx* ctx);
/* SetForcing_Term (:502-550): coarse Res_TruncError = damp * sum of the children's LinSysRes - coarse LinSysRes */
int su2b200_mg_set_forcing_term(su2b200_ctx* fine, su2b200_ctx* coarse, double damp_restriction);
/* GetProlongated_Correction (:308-381): coarse Solution_Old = u_c - I(u_f); fine LinSysRes(child) = that correction */
int su2b200_mg_get_prolongated_correction(su2b200_ctx* fine, su2b200_ctx* coarse);
/* SmoothProlongated_Correction (:383-456): nSmooth Jacobi sweeps over the neighbours, boundary points restored */
int su2b200_mg_smooth_prolongated_correction(su2b200_ctx* ctx, int nSmooth, double coeff);
/* SetProlongated_Correction (:458-484): Solution += damp * LinSysRes over the owned points */
int su2b200_mg_set_prolongated_correction(su2b200_ctx* ctx, double damp_prolongation);
typedef struct {
  int nLevels;                 /* number of COARSE levels (CConfig::GetnMGLevels)                                */
  int cycle;                   /* MGCYCLE: 0 V_CYCLE, 1 W_CYCLE, 2 FULLMG_CYCLE (runs as V)                      */
  int pre_smooth[8];           /* MG_PRE_SMOOTH per level                                                        */
  int post_smooth[8];          /* MG_POST_SMOOTH                                                                 */
  int correction_smooth[8];    /* MG_CORRECTION_SMOOTH                                                           */
  double damp_restriction;     /* MG_DAMP_RESTRICTION                                                            */
  double damp_prolongation;    /* MG_DAMP_PROLONGATION                                                           */
} su2b200_mg_config;
/* CMultiGridIntegration::MultiGrid_Iteration (:37-127) + MultiGrid_Cycle (:129-300): nIter FAS cycles over
 * levels[0..nLevels], forces and residual norms of the finest grid in the reports */
int su2b200_mg_iterate(su2b200_ctx** levels, const su2b200_mg_config* cfg, int nIter, su2b200_iter_report* reports);

/* ---- CSysMatrix / CSysVector / CSysSolve level (host-buffer entry points, HtD/DtH inside) ---- */

/* CSysMatrix::Initialize (CSysMatrix.cpp:86-247) for a system not tied to a flow state:
 * creates system `which` of the context from a pattern. Geometry upload does this implicitly. */
int su2b200_linsys_create(su2b200_ctx* ctx, int which, int nVar, unsigned long nPoint, unsigned long nPointDomain,
                          const unsigned long* row_ptr, const unsigned long* col_ind);
/* number of row partitions of the preconditioner sweeps: LINEAR_SOLVER_PREC_THREADS / omp_partitions
 * (CSysMatrix.cpp:199-227); 1 = the serial reference order (level-scheduled on the device) */
int su2b200_linsys_set_partitions(su2b200_ctx* ctx, int which, int nparts);
/* schedule of the exact (nparts == 1) ILU0 kernels, for A/B tests; -1 keeps the current value.
 * apply_mode: 0 one launch per dependency level, 1 single-launch flag sweeps, 2 single-launch data-polling sweeps
 * (default); build_mode: 0 one launch per level, 1 flag sweep, 2 single-launch polling factorisation (default).
 * All schedules perform the arithmetic of CSysMatrix.cpp:673-823 in the same order. */
int su2b200_linsys_set_schedule(su2b200_ctx* ctx, int which, int apply_mode, int build_mode);
/* level width (rows) from which the hybrid ILU0 schedules hand a level to the nVar-lanes-per-row kernels (default 4096
 * for both): sweep_rows for the application, factor_rows for the factorisation; factor_mode 0 = warp-per-row kernel only,
 * 1 = hybrid, 2 = rows kernel only; values <= 0 (factor_mode < 0) keep the current setting. Same arithmetic either way:
 * the tests run the wide-level kernels on small meshes through this. */
int su2b200_linsys_set_wide_threshold(su2b200_ctx* ctx, int which, int sweep_rows, int factor_rows, int factor_mode);
/* explicit row partitions of the block-Jacobi preconditioner: first_row[nparts+1], first_row[0] = 0,
 * first_row[nparts] = nPointDomain. The reference's omp_partitions (CSysMatrix.cpp:199-227) with caller-chosen
 * boundaries; used to reproduce on one GPU the operator N ranks build (partitions = the ranks' row ranges). */
int su2b200_linsys_set_partition_rows(su2b200_ctx* ctx, int which, int nparts, const unsigned long* first_row);
/* CSysMatrix::HtDTransfer (CSysMatrixGPU.cu:59-64) */
int su2b200_linsys_set_matrix(su2b200_ctx* ctx, int which, const double* blocks);
/* CSysMatrix::GPUMatrixVectorProduct (CSysMatrixGPU.cu:66-86) == MatrixVectorProduct (CSysMatrix.cpp:617-645):
 * y = A x over owned rows; x,y host [nPoint][nVar]. */
int su2b200_linsys_spmv(su2b200_ctx* ctx, int which, const double* x, double* y);
/* CPreconditioner::Build() (CPreconditioner.hpp:49-75): BuildJacobiPreconditioner / BuildILUPreconditioner */
int su2b200_linsys_build_preconditioner(su2b200_ctx* ctx, int which, int prec);
/* CPreconditioner::operator()(u,v): ComputeJacobi/ILU/LU_SGSPreconditioner (CSysMatrix.cpp:660-888) */
int su2b200_linsys_apply_preconditioner(su2b200_ctx* ctx, int which, int prec, const double* x, double* y);
/* The pair every Krylov iteration runs, `precond(V[i], Z[i]); mat_vec(Z[i], V[i+1]);` (CSysSolve.cpp:533-537 FGMRES,
 * :342-351 / :376 BCGSTAB): z = M^-1 v, w = A z; v, z, w host [nPoint][nVar]. With ILU0 on a pattern whose upper factor is
 * the Jacobian's own upper triangle (hexahedral / quadrilateral meshes) the backward sweep hands its sum_{j>i} A_ij z_j to
 * the product, which then reads only the lower triangle, the diagonal and the halo columns; otherwise the two
 * operators run one after the other. su2b200_linsys_solve uses the same routine. SU2B200_FUSED_UPPER=0 disables it. */
int su2b200_linsys_precondition_and_multiply(su2b200_ctx* ctx, int which, int prec, const double* v, double* z, double* w);
/* CSysSolve::Solve (CSysSolve.cpp:1318): b,x host [nPoint][nVar]; x0 = 0 */
int su2b200_linsys_solve(su2b200_ctx* ctx, int which, int solver, int prec, double tol, int max_iter,
                         const double* b, double* x, unsigned long* iterations, double* residual);
/* CSysVector::dot (CSysVector.hpp:335-366) over owned entries, on the device */
int su2b200_vector_dot(su2b200_ctx* ctx, int which, const double* a, const double* b, double* result);
/* Allocation-free variant of su2b200_vector_dot. It has NO counterpart in the reference: it was added
 * for callers that may not call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it). a, b: host [n]; result: host; the sum runs over all n entries on the context's device and
 * stream, rank-local (no allreduce). The staging copies and the block partials live in a static device array that
 * every GPU loading the library holds until exit (16 MiB + 8 KiB), so n is limited to SU2B200_NOALLOC_DOT_MAX doubles
 * per vector; larger n is rejected with an error. The call makes no cudaMalloc/cudaFree/cudaMallocHost and uses no
 * allocating library routine. All calls share the array: calls whose GPU work overlaps race on it. */
#define SU2B200_NOALLOC_DOT_MAX 1048576ul
int su2b200_vector_dot_noalloc(su2b200_ctx* ctx, const double* a, const double* b, unsigned long n, double* result);

/* ---- restart / solution files in the SU2 native formats ------------------------------------------
 * binary ".dat" (CSU2BinaryFileWriter.cpp:39-94, CSolver::Read_SU2_Restart_Binary CSolver.cpp:2916) and CSV
 * (CSolver::Read_SU2_Restart_ASCII :2754); columns x,y[,z],Density,Momentum_*,Energy[,Nu_Tilde]; rows by GLOBAL point index
 * (global_index[nPoint] = CPoint::GlobalIndex, NULL = identity), so files are independent of RCM and partitioning.
 * A restart written here is read by the reference (RESTART_SOL= YES) and vice versa. */
/* COutput::SetResultFiles (RESTART / RESTART_ASCII): the owned points of this rank; in parallel every rank writes its rows
 * into the one binary file (rank 0 the header), nPointGlobal = global point count (0 = single rank) */
int su2b200_restart_write(su2b200_ctx* ctx, const char* path, int binary, const unsigned long* global_index,
                          unsigned long nPointGlobal);
/* CFVMFlowSolverBase::LoadRestart_impl (CFVMFlowSolverBase.inl:824-1008) + CTurbSASolver::LoadRestart: Solution, Solution_Old
 * (and nu~) of every local point (halos included) from the file; the format is detected from the magic number */
int su2b200_restart_read(su2b200_ctx* ctx, const char* path, const unsigned long* global_index);
/* the same on host arrays (no GPU needed): solution [nPoint][nDim+2], turb [nPoint] when nTurb = 1 */
int su2b200_restart_write_arrays(const char* path, int binary, int nDim, int nTurb, unsigned long nPoint,
                                 const unsigned long* global_index, unsigned long nPointGlobal, const double* coord,
                                 const double* solution, const double* turb, int write_header);
int su2b200_restart_read_arrays(const char* path, int nDim, int nTurb, unsigned long nPoint,
                                const unsigned long* global_index, double* solution, double* turb);

/* ---- communication: CSolver::InitiateComms/CompleteComms (CSolver.cpp:1406-1697) ------------ */
/* 128-byte ncclUniqueId: rank 0 creates it, the host side broadcasts it (MPI_Bcast / torch.distributed) */
int su2b200_comm_unique_id(unsigned char id[128]);
int su2b200_comm_init(su2b200_ctx* ctx, int rank, int size, const unsigned char id[128]);
/* Optional, after su2b200_comm_init, ranks = processes on one node: allreduces of single doubles inside the Krylov
 * reduction kernels over peer memory (NVLink/NVSwitch) instead of ncclAllReduce (the role of SU2_MPI::Allreduce in
 * CSysVector::dot, CSysVector.hpp:335-366, called ~100 times per nonlinear iteration by ModGramSchmidt,
 * CSysSolve.cpp:179-240). _create returns this rank's 64-byte cudaIpc handle; the caller allgathers the handles
 * (MPI_Allgather in the reference, torch.distributed here) and passes all of them, in rank order, to _connect.
 * Without it, or if _connect fails, the NCCL path stays in use. SU2B200_MAILBOX=0 disables it. */
int su2b200_comm_mailbox_create(su2b200_ctx* ctx, unsigned char handle[64]);
int su2b200_comm_mailbox_connect(su2b200_ctx* ctx, const unsigned char* handles /*[size][64]*/);
/* The mailboxes are used only after EVERY rank has connected: the caller allgathers the return codes of _connect and
 * calls _enable(ctx, all ranks succeeded) on every rank (a rank on NCCL and a rank on mailboxes would wait for each
 * other forever). */
int su2b200_comm_mailbox_enable(su2b200_ctx* ctx, int on);
/* halo exchange of a named per-point array (MPI_QUANTITIES: SOLUTION, PRIMITIVE_GRADIENT, PRIMITIVE_LIMITER, ...):
 * InitiateComms + CompleteComms in one call; count per point = su2b200_array_size / nPoint; no-op on a single rank */
int su2b200_halo_exchange(su2b200_ctx* ctx, int which_array);

/* ---- timing helpers used by bench.py (CUDA events on the context's compute stream) ----------- */
int su2b200_timer_start(su2b200_ctx* ctx);
int su2b200_timer_stop(su2b200_ctx* ctx, float* ms);
/* kernels launched by this context since creation (bench.py's gpu_launches) */
unsigned long su2b200_launch_count(su2b200_ctx* ctx);
/* per-category device time of the calls made while enabled (CUDA events on the compute stream): bench.py's kernel shares.
 * ms/calls have su2b200_profile_ncat() entries, named by su2b200_profile_name(). Keep it off inside timed regions. */
int su2b200_profile_enable(su2b200_ctx* ctx, int on);
int su2b200_profile_ncat(void);
const char* su2b200_profile_name(int cat);
int su2b200_profile_report(su2b200_ctx* ctx, double* ms, unsigned long* calls);
/* micro-benchmarks of the two roofline kernels on the resident state: average ms over `reps` launches */
int su2b200_bench_edge_kernel(su2b200_ctx* ctx, int reps, float* ms_avg);
int su2b200_bench_spmv(su2b200_ctx* ctx, int which, int reps, float* ms_avg);
int su2b200_bench_prec(su2b200_ctx* ctx, int which, int prec, int reps, float* ms_build, float* ms_apply);

#ifdef __cplusplus
}
#endif
#endif
