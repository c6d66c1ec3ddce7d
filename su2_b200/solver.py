"""Host-side mirror of the reference's solver interface for the hot path, over the C ABI.

Method names follow the reference classes they stand in for (CSolver virtuals of CEulerSolver /
CNSSolver / CTurbSASolver, CSysMatrix, CSysSolve), so a parity test reads like the reference's
call sequence in CMultiGridIntegration::MultiGrid_Cycle (SU2_CFD/src/integration/
CMultiGridIntegration.cpp:163-206) and CIntegration::Space_Integration (CIntegration.cpp:37-201).
All state lives on the GPU; this class only marshals arguments.
"""
import ctypes as C
import numpy as np
from . import capi
from .capi import check, as_ulong, as_f64, ptr_ul, ptr_d


class FlowSolver:
    """One rank == one GPU: geometry + flow state + linear system(s) on `device`."""

    def __init__(self, device=0):
        self.lib = capi.load()
        self.ctx = C.c_void_p()
        check(self.lib.su2b200_create(int(device), C.byref(self.ctx)))
        self.nDim = self.nPoint = self.nPointDomain = self.nEdge = 0
        self.nnz = 0
        self.markers = []
        self.cfg = None
        self._keep = []

    def close(self):
        if self.ctx:
            self.lib.su2b200_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- geometry (CGeometry arrays) -----------------------------------------------------------
    def set_geometry(self, geo):
        """geo: dict with the CGeometry arrays (names as in oracle_dump / su2b200_geometry_desc)."""
        d = capi.GeometryDesc()
        nDim = int(geo["nDim"])
        P, Pd, E = int(geo["nPoint"]), int(geo["nPointDomain"]), int(geo["nEdge"])
        en = as_ulong(geo["edge_nodes"]).reshape(2, E)
        keep = dict(
            e0=as_ulong(en[0]), e1=as_ulong(en[1]), en=as_f64(geo["edge_normal"]).reshape(E, nDim),
            coord=as_f64(geo["coord"]).reshape(P, nDim), vol=as_f64(geo["volume"]),
            pp=as_ulong(geo["point_ptr"]), pn=as_ulong(geo["point_nbr"]), pe=as_ulong(geo["point_edge"]),
            rp=as_ulong(geo["row_ptr"]), ci=as_ulong(geo["col_ind"]), dp=as_ulong(geo["dia_ptr"]),
            ep=as_ulong(geo["edge_ptr"]).reshape(E, 2))
        d.nDim, d.nPoint, d.nPointDomain, d.nEdge = nDim, P, Pd, E
        d.edge_node0, d.edge_node1 = ptr_ul(keep["e0"]), ptr_ul(keep["e1"])
        d.edge_normal, d.coord, d.volume = ptr_d(keep["en"]), ptr_d(keep["coord"]), ptr_d(keep["vol"])
        d.point_ptr, d.point_nbr, d.point_edge = ptr_ul(keep["pp"]), ptr_ul(keep["pn"]), ptr_ul(keep["pe"])
        d.row_ptr, d.col_ind, d.dia_ptr, d.edge_ptr = ptr_ul(keep["rp"]), ptr_ul(keep["ci"]), ptr_ul(keep["dp"]), ptr_ul(keep["ep"])
        if geo.get("nNeighbor") is not None:
            keep["nn"] = np.ascontiguousarray(geo["nNeighbor"], dtype=np.uint16)
            d.nNeighbor = keep["nn"].ctypes.data_as(C.POINTER(C.c_ushort))
        if geo.get("physical_boundary") is not None:
            keep["pb"] = np.ascontiguousarray(geo["physical_boundary"], dtype=np.uint8)
            d.physical_boundary = keep["pb"].ctypes.data_as(C.POINTER(C.c_ubyte))
        if geo.get("wall_distance") is not None:
            keep["wd"] = as_f64(geo["wall_distance"])
            d.wall_distance = ptr_d(keep["wd"])
        if geo.get("color_ptr") is not None:
            keep["cp"] = as_ulong(geo["color_ptr"])
            keep["ce"] = as_ulong(geo["color_edges"])
            d.color_ptr, d.color_edges, d.nColor = ptr_ul(keep["cp"]), ptr_ul(keep["ce"]), len(keep["cp"]) - 1
        check(self.lib.su2b200_geometry_upload(self.ctx, C.byref(d)))
        self.nDim, self.nPoint, self.nPointDomain, self.nEdge = nDim, P, Pd, E
        self.nnz = int(keep["rp"][-1])
        self.markers = []
        for m in geo.get("markers", []):
            self.add_marker(**m)

    def add_marker(self, kind, node, normal, normal_neighbor=None, sym_unit_normal=None, monitoring=0,
                   heat_flux=0.0, tag=""):
        i = len(self.markers)
        node = as_ulong(node)
        nV = len(node)
        normal = as_f64(normal).reshape(nV, self.nDim)
        nb = as_ulong(normal_neighbor) if normal_neighbor is not None else None
        un = as_f64(sym_unit_normal).reshape(nV, self.nDim) if sym_unit_normal is not None else None
        check(self.lib.su2b200_marker_upload(self.ctx, i, int(kind), C.c_ulong(nV), ptr_ul(node), ptr_d(normal),
                                             ptr_ul(nb), ptr_d(un), int(monitoring), C.c_double(heat_flux)))
        self.markers.append(dict(kind=int(kind), tag=tag, nVertex=nV, monitoring=int(monitoring)))
        return i

    # ---- configuration ---------------------------------------------------------------------------
    def configure(self, **kw):
        cfg = default_config()
        for k, v in kw.items():
            if k in ("velocity_inf", "ref_origin", "cfl_adapt_param", "sst_lower", "sst_upper"):
                arr = getattr(cfg, k)
                for i, x in enumerate(v):
                    arr[i] = float(x)
            else:
                if not hasattr(cfg, k):
                    raise KeyError(f"unknown config field {k}")
                setattr(cfg, k, v)
        check(self.lib.su2b200_flow_configure(self.ctx, C.byref(cfg)))
        self.cfg = cfg

    def allocate(self):
        check(self.lib.su2b200_flow_allocate(self.ctx))

    @property
    def nVar(self):
        return self.nDim + 2

    # ---- arrays --------------------------------------------------------------------------------------
    def get(self, name):
        wid = capi.ARR[name]
        n = self.lib.su2b200_array_size(self.ctx, wid)
        if n < 0:
            raise KeyError(name)
        out = np.empty(n, dtype=np.float64)
        if n:
            check(self.lib.su2b200_array_download(self.ctx, wid, ptr_d(out)))
        return out

    def set(self, name, values):
        wid = capi.ARR[name]
        n = self.lib.su2b200_array_size(self.ctx, wid)
        v = as_f64(values).ravel()
        if v.size != n:
            raise ValueError(f"{name}: expected {n} values, got {v.size}")
        check(self.lib.su2b200_array_upload(self.ctx, wid, ptr_d(v)))

    def download_into(self, name, host):
        """D2H copy of a named device array into a caller-owned (ideally pinned) float64 buffer."""
        wid = capi.ARR[name]
        assert host.dtype == np.float64 and host.size == self.lib.su2b200_array_size(self.ctx, wid)
        check(self.lib.su2b200_array_download(self.ctx, wid, ptr_d(host)))

    def upload_from(self, name, host):
        wid = capi.ARR[name]
        assert host.dtype == np.float64 and host.size == self.lib.su2b200_array_size(self.ctx, wid)
        check(self.lib.su2b200_array_upload(self.ctx, wid, ptr_d(host)))

    # ---- communication (one context == one rank == one GPU) -----------------------------------------------
    def comm_unique_id(self):
        buf = (C.c_ubyte * 128)()
        check(self.lib.su2b200_comm_unique_id(buf))
        return np.frombuffer(buf, dtype=np.uint8).copy()

    def comm_init(self, rank, size, unique_id):
        uid = np.ascontiguousarray(unique_id, dtype=np.uint8)
        assert uid.size == 128
        check(self.lib.su2b200_comm_init(self.ctx, int(rank), int(size), uid.ctypes.data_as(C.POINTER(C.c_ubyte))))

    def mailbox_connect(self, allgather):
        """Peer-memory allreduce of the Krylov scalars (su2b200_comm_mailbox_*). `allgather(bytes) -> [bytes per rank]` is
        the caller's collective (torch.distributed, MPI...). Returns False (NCCL stays in use) if the handles cannot be opened."""
        h = (C.c_ubyte * 64)()
        ok = self.lib.su2b200_comm_mailbox_create(self.ctx, h) == 0
        allh = allgather(bytes(h))  # collective: every rank takes part whatever its own outcome
        if ok:
            buf = (C.c_ubyte * (64 * len(allh))).from_buffer_copy(b"".join(allh))
            ok = self.lib.su2b200_comm_mailbox_connect(self.ctx, buf) == 0
        everyone = all(f == b"\x01" for f in allgather(b"\x01" if ok else b"\x00"))
        self.lib.su2b200_comm_mailbox_enable(self.ctx, int(everyone))
        return everyone

    def halo_exchange(self, name):
        """InitiateComms/CompleteComms of a named per-point array (su2b200_halo_exchange)."""
        check(self.lib.su2b200_halo_exchange(self.ctx, capi.ARR[name]))

    def set_partition_rows(self, which, first_row):
        """Explicit block-Jacobi partitions of system `which` (su2b200_linsys_set_partition_rows)."""
        fr = as_ulong(first_row)
        check(self.lib.su2b200_linsys_set_partition_rows(self.ctx, int(which), len(fr) - 1, ptr_ul(fr)))

    def set_wide_threshold(self, which, sweep_rows=0, factor_rows=0, factor_mode=-1):
        check(self.lib.su2b200_linsys_set_wide_threshold(self.ctx, int(which), int(sweep_rows), int(factor_rows),
                                                         int(factor_mode)))

    # ---- restart files (SU2 native formats) -----------------------------------------------------------------
    def write_restart(self, path, binary=True, global_index=None, n_global=0):
        """COutput::SetResultFiles(RESTART / RESTART_ASCII) of the owned points (su2b200_restart_write)."""
        gi = as_ulong(global_index) if global_index is not None else None
        check(self.lib.su2b200_restart_write(self.ctx, str(path).encode(), int(bool(binary)), ptr_ul(gi),
                                             C.c_ulong(int(n_global))))

    def load_restart(self, path, global_index=None):
        """CSolver::LoadRestart: Solution (+ nu~) from a .dat / .csv restart file (su2b200_restart_read)."""
        gi = as_ulong(global_index) if global_index is not None else None
        check(self.lib.su2b200_restart_read(self.ctx, str(path).encode(), ptr_ul(gi)))

    def edge_counter(self):
        out = np.empty(self.nEdge, dtype=np.int8)
        check(self.lib.su2b200_edge_counter_download(self.ctx, out.ctypes.data_as(C.POINTER(C.c_byte))))
        return out

    # ---- CSolver virtuals (flow) --------------------------------------------------------------------------
    def Preprocessing(self, Output=False):
        n = C.c_ulong(0)
        check(self.lib.su2b200_flow_preprocessing(self.ctx, int(bool(Output)), C.byref(n)))
        return n.value

    def Set_OldSolution(self):
        check(self.lib.su2b200_flow_set_old_solution(self.ctx))

    def SetTime_Step(self):
        mn, mx = C.c_double(), C.c_double()
        check(self.lib.su2b200_flow_set_time_step(self.ctx, C.byref(mn), C.byref(mx)))
        return mn.value, mx.value

    def EdgeFluxResidual(self):
        """Upwind_Residual / Centered_Residual (+ fused viscous terms)."""
        n = C.c_ulong(0)
        check(self.lib.su2b200_flow_edge_residual(self.ctx, C.byref(n)))
        return n.value

    Upwind_Residual = EdgeFluxResidual
    Centered_Residual = EdgeFluxResidual

    def BC_Far_Field(self, iMarker):
        check(self.lib.su2b200_flow_bc_far_field(self.ctx, int(iMarker)))

    def BC_Sym_Plane(self, iMarker):
        check(self.lib.su2b200_flow_bc_sym_plane(self.ctx, int(iMarker)))

    BC_Euler_Wall = BC_Sym_Plane

    def BC_HeatFlux_Wall(self, iMarker):
        check(self.lib.su2b200_flow_bc_heatflux_wall(self.ctx, int(iMarker)))

    def Space_Integration(self):
        check(self.lib.su2b200_flow_space_integration(self.ctx))

    def PrepareImplicitIteration(self):
        nv = self.nVar
        rms = (C.c_double * nv)()
        mx = (C.c_double * nv)()
        idx = (C.c_ulong * nv)()
        check(self.lib.su2b200_flow_prepare_implicit(self.ctx, rms, mx, idx))
        return np.array(rms[:]), np.array(mx[:]), np.array(idx[:])

    def Solve(self):
        it, res = C.c_ulong(0), C.c_double(0)
        check(self.lib.su2b200_flow_linear_solve(self.ctx, C.byref(it), C.byref(res)))
        return it.value, res.value

    def CompleteImplicitIteration(self):
        check(self.lib.su2b200_flow_complete_implicit(self.ctx))

    def Pressure_Forces(self):
        out = (C.c_double * 9)()
        check(self.lib.su2b200_flow_forces(self.ctx, out))
        return dict(zip(["CL", "CD", "CSF", "CFx", "CFy", "CFz", "CMx", "CMy", "CMz"], out[:]))

    # ---- CSolver virtuals (SA) ---------------------------------------------------------------------------
    def Turb_Preprocessing(self):
        check(self.lib.su2b200_turb_preprocessing(self.ctx))

    def Turb_SetTime_Step(self):
        check(self.lib.su2b200_turb_set_time_step(self.ctx))

    def Turb_Space_Integration(self):
        check(self.lib.su2b200_turb_space_integration(self.ctx))

    def Turb_PrepareImplicitIteration(self):
        rms, mx, idx = (C.c_double * 2)(), (C.c_double * 2)(), (C.c_ulong * 2)()  # one entry per turbulence variable
        check(self.lib.su2b200_turb_prepare_implicit(self.ctx, rms, mx, idx))
        return rms[0] if not self.cfg.rans_sst else (rms[0], rms[1])

    def Turb_Solve(self):
        it, res = C.c_ulong(0), C.c_double(0)
        check(self.lib.su2b200_turb_linear_solve(self.ctx, C.byref(it), C.byref(res)))
        return it.value, res.value

    def Turb_CompleteImplicitIteration(self):
        check(self.lib.su2b200_turb_complete_implicit(self.ctx))

    def Turb_Postprocessing(self):
        check(self.lib.su2b200_turb_postprocessing(self.ctx))

    # ---- whole iterations ------------------------------------------------------------------------------------
    def iterate(self, n=1):
        reps = (capi.IterReport * n)()
        check(self.lib.su2b200_iterate(self.ctx, int(n), reps))
        out = []
        for r in reps:
            out.append(dict(rms_flow=np.array(r.rms_flow[:self.nVar]), rms_turb=r.rms_turb, rms_turb2=r.rms_turb2, CL=r.CL, CD=r.CD,
                            linsol_iter_flow=r.linsol_iter_flow, linsol_res_flow=r.linsol_res_flow,
                            linsol_iter_turb=r.linsol_iter_turb, linsol_res_turb=r.linsol_res_turb,
                            nonphysical_points=r.nonphysical_points, nonphysical_edges=r.nonphysical_edges))
        return out

    # ---- timing ----------------------------------------------------------------------------------------------
    def synchronize(self):
        check(self.lib.su2b200_synchronize(self.ctx))

    def timer_start(self):
        check(self.lib.su2b200_timer_start(self.ctx))

    def timer_stop(self):
        ms = C.c_float()
        check(self.lib.su2b200_timer_stop(self.ctx, C.byref(ms)))
        return ms.value

    def launch_count(self):
        return int(self.lib.su2b200_launch_count(self.ctx))

    def profile_enable(self, on=True):
        check(self.lib.su2b200_profile_enable(self.ctx, int(bool(on))))

    def profile_report(self):
        n = self.lib.su2b200_profile_ncat()
        ms = (C.c_double * n)()
        calls = (C.c_ulong * n)()
        check(self.lib.su2b200_profile_report(self.ctx, ms, calls))
        return {self.lib.su2b200_profile_name(k).decode(): (ms[k], calls[k]) for k in range(n) if calls[k]}

    def bench_edge_kernel(self, reps=10):
        ms = C.c_float()
        check(self.lib.su2b200_bench_edge_kernel(self.ctx, int(reps), C.byref(ms)))
        return ms.value

    def bench_spmv(self, which=0, reps=10):
        ms = C.c_float()
        check(self.lib.su2b200_bench_spmv(self.ctx, int(which), int(reps), C.byref(ms)))
        return ms.value

    def bench_prec(self, which=0, prec=capi.PREC_ILU, reps=5):
        """(ms per build, ms per apply) of the preconditioner on the resident matrix."""
        b, a = C.c_float(), C.c_float()
        check(self.lib.su2b200_bench_prec(self.ctx, int(which), int(prec), int(reps), C.byref(b), C.byref(a)))
        return b.value, a.value


class MultiGrid:
    """CMultiGridIntegration stand-in: a list of FlowSolver levels (0 = finest) linked by their Children_CV lists.

    levels[k] must be configured with mg_level = k (and the CFL of that level, CConfig::GetCFL(iMesh)); children[k-1]
    = (children_ptr, children) of level k.  Method names follow CMultiGridIntegration.cpp."""

    def __init__(self, levels, children, cycle=1, pre_smooth=(1, 2, 3, 3), post_smooth=(0, 0, 0, 0),
                 correction_smooth=(0, 0, 0, 0), damp_restriction=1.0, damp_prolongation=1.0):
        self.lib = capi.load()
        self.levels = list(levels)
        assert len(children) == len(self.levels) - 1
        self._keep = []
        for k, (cp, ch) in enumerate(children):
            cp, ch = as_ulong(cp), as_ulong(ch)
            self._keep.append((cp, ch))
            check(self.lib.su2b200_mg_link(self.levels[k].ctx, self.levels[k + 1].ctx, ptr_ul(cp), ptr_ul(ch)))
        mc = capi.MgConfig()
        mc.nLevels, mc.cycle = len(self.levels) - 1, int(cycle)
        for k in range(len(self.levels)):
            mc.pre_smooth[k] = int(pre_smooth[k])
            mc.post_smooth[k] = int(post_smooth[k])
            mc.correction_smooth[k] = int(correction_smooth[k])
        mc.damp_restriction, mc.damp_prolongation = float(damp_restriction), float(damp_prolongation)
        self.cfg = mc

    def SetRestricted_Solution(self, k):
        check(self.lib.su2b200_mg_set_restricted_solution(self.levels[k].ctx, self.levels[k + 1].ctx))

    def SetResidual_Term(self, k):
        check(self.lib.su2b200_mg_set_residual_term(self.levels[k].ctx))

    def SetForcing_Term(self, k):
        check(self.lib.su2b200_mg_set_forcing_term(self.levels[k].ctx, self.levels[k + 1].ctx,
                                                   C.c_double(self.cfg.damp_restriction)))

    def GetProlongated_Correction(self, k):
        check(self.lib.su2b200_mg_get_prolongated_correction(self.levels[k].ctx, self.levels[k + 1].ctx))

    def SmoothProlongated_Correction(self, k, coeff=1.25):
        check(self.lib.su2b200_mg_smooth_prolongated_correction(self.levels[k].ctx, int(self.cfg.correction_smooth[k]),
                                                               C.c_double(coeff)))

    def SetProlongated_Correction(self, k):
        check(self.lib.su2b200_mg_set_prolongated_correction(self.levels[k].ctx, C.c_double(self.cfg.damp_prolongation)))

    def smooth(self, k):
        """One implicit smoothing step on level k (the pre/post-smoothing body of MultiGrid_Cycle)."""
        s = self.levels[k]
        s.Preprocessing(False)
        s.Set_OldSolution()
        s.SetTime_Step()
        s.Space_Integration()
        rms, _, _ = s.PrepareImplicitIteration()
        s.Solve()
        s.CompleteImplicitIteration()
        return rms

    def iterate(self, n=1):
        """n FAS cycles on the device (su2b200_mg_iterate)."""
        reps = (capi.IterReport * n)()
        ctxs = (C.c_void_p * len(self.levels))(*[s.ctx for s in self.levels])
        check(self.lib.su2b200_mg_iterate(ctxs, C.byref(self.cfg), int(n), reps))
        nv = self.levels[0].nVar
        return [dict(rms_flow=np.array(r.rms_flow[:nv]), CL=r.CL, CD=r.CD, linsol_iter_flow=r.linsol_iter_flow,
                     linsol_res_flow=r.linsol_res_flow) for r in reps]

    def close(self):
        for s in reversed(self.levels):
            s.close()


class LinearSystem:
    """CSysMatrix + CSysSolve stand-in with host-buffer entry points (HtD/DtH inside every call)."""

    def __init__(self, nVar, nPoint, nPointDomain, row_ptr, col_ind, device=0, which=0, solver=None):
        self.lib = capi.load()
        self.owner = solver
        if solver is None:
            self.ctx = C.c_void_p()
            check(self.lib.su2b200_create(int(device), C.byref(self.ctx)))
        else:
            self.ctx = solver.ctx
        self.which = which
        self.nVar, self.nPoint, self.nPointDomain = int(nVar), int(nPoint), int(nPointDomain)
        rp, ci = as_ulong(row_ptr), as_ulong(col_ind)
        check(self.lib.su2b200_linsys_create(self.ctx, which, self.nVar, C.c_ulong(self.nPoint),
                                             C.c_ulong(self.nPointDomain), ptr_ul(rp), ptr_ul(ci)))
        self.nnz = int(rp[-1])

    def close(self):
        if self.owner is None and self.ctx:
            self.lib.su2b200_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def set_partitions(self, n):
        check(self.lib.su2b200_linsys_set_partitions(self.ctx, self.which, int(n)))

    def set_schedule(self, apply_mode=-1, build_mode=-1):
        """ILU0 kernel schedules (see su2b200_linsys_set_schedule); -1 keeps the current one."""
        check(self.lib.su2b200_linsys_set_schedule(self.ctx, self.which, int(apply_mode), int(build_mode)))

    def set_wide_threshold(self, sweep_rows=0, factor_rows=0, factor_mode=-1):
        """Level width from which the nVar-lanes-per-row ILU kernels take over (su2b200_linsys_set_wide_threshold)."""
        check(self.lib.su2b200_linsys_set_wide_threshold(self.ctx, self.which, int(sweep_rows), int(factor_rows),
                                                         int(factor_mode)))

    def set_partition_rows(self, first_row):
        fr = as_ulong(first_row)
        check(self.lib.su2b200_linsys_set_partition_rows(self.ctx, self.which, len(fr) - 1, ptr_ul(fr)))

    def set_matrix(self, blocks):
        b = as_f64(blocks).ravel()
        assert b.size == self.nnz * self.nVar * self.nVar
        check(self.lib.su2b200_linsys_set_matrix(self.ctx, self.which, ptr_d(b)))

    def MatrixVectorProduct(self, x):
        x = as_f64(x).ravel()
        y = np.zeros_like(x)
        check(self.lib.su2b200_linsys_spmv(self.ctx, self.which, ptr_d(x), ptr_d(y)))
        return y

    def BuildPreconditioner(self, prec):
        check(self.lib.su2b200_linsys_build_preconditioner(self.ctx, self.which, int(prec)))

    def ApplyPreconditioner(self, prec, x):
        x = as_f64(x).ravel()
        y = np.zeros_like(x)
        check(self.lib.su2b200_linsys_apply_preconditioner(self.ctx, self.which, int(prec), ptr_d(x), ptr_d(y)))
        return y

    def Solve(self, b, solver=capi.SOLVER_FGMRES, prec=capi.PREC_ILU, tol=1e-10, max_iter=10):
        b = as_f64(b).ravel()
        x = np.zeros_like(b)
        it, res = C.c_ulong(0), C.c_double(0)
        check(self.lib.su2b200_linsys_solve(self.ctx, self.which, int(solver), int(prec), C.c_double(tol),
                                            int(max_iter), ptr_d(b), ptr_d(x), C.byref(it), C.byref(res)))
        return x, it.value, res.value

    def PreconditionAndMultiply(self, prec, v):
        """z = M^-1 v, w = A z (one Krylov iteration's operator pair; fused upper product where the pattern allows)."""
        v = as_f64(v).ravel()
        z, w = np.empty_like(v), np.empty_like(v)
        check(self.lib.su2b200_linsys_precondition_and_multiply(self.ctx, self.which, int(prec), ptr_d(v), ptr_d(z), ptr_d(w)))
        return z, w

    def dot(self, a, b):
        a, b = as_f64(a).ravel(), as_f64(b).ravel()
        r = C.c_double()
        check(self.lib.su2b200_vector_dot(self.ctx, self.which, ptr_d(a), ptr_d(b), C.byref(r)))
        return r.value


def default_config():
    """Defaults of the CConfig options the path reads (Common/src/CConfig.cpp option registrations)."""
    c = capi.FlowConfig()
    c.gamma, c.gas_constant = 1.4, 287.058
    c.mu_sutherland = -1.0
    c.mu_ref, c.mu_temperature_ref, c.mu_constant = 1.716e-5, 273.15, 1.716e-5
    c.prandtl_lam, c.prandtl_turb = 0.72, 0.90
    c.conv_scheme, c.muscl, c.muscl_kappa = capi.CONV_ROE, 1, 0.0
    c.limiter, c.venkat_coeff, c.ref_elem_length = capi.LIMITER_VENKATAKRISHNAN, 0.05, 0.1
    c.grad_method = c.grad_method_recon = capi.GRAD_GREEN_GAUSS
    c.roe_kappa, c.entropy_fix = 0.5, 0.001
    c.jst_kappa2, c.jst_kappa4, c.cent_jac_fix_factor = 0.5, 0.02, 4.0
    c.cfl, c.cfl_turb_ratio, c.max_delta_time = 1.25, 1.0, 1e6
    c.max_update_flow, c.max_update_sa = 0.2, 0.99
    c.linear_solver, c.linear_prec, c.linear_tol, c.linear_iter, c.linear_restart = capi.SOLVER_FGMRES, capi.PREC_ILU, 1e-6, 10, 10
    c.turb_linear_solver, c.turb_linear_prec, c.turb_linear_tol, c.turb_linear_iter = capi.SOLVER_FGMRES, capi.PREC_ILU, 1e-6, 10
    c.ref_area, c.ref_length = 1.0, 1.0
    c.assembly, c.prec_partitions = capi.ASSEMBLY_GATHER, 1
    c.turb_reducer, c.mg_level, c.lax_kappa1 = -1, 0, 0.15
    c.cfl_adapt, c.n_inner_iter = 0, 1000
    c.rans_sst, c.max_update_sst = 0, 1.0   # MAX_UPDATE_SST
    c.sst_lower[0], c.sst_lower[1], c.sst_upper[0], c.sst_upper[1] = 1.0e-10, 1.0e-4, 1.0e10, 1.0e15
    for k, v in enumerate((0.1, 1.2, 10.0, 100.0, 0.001, 0.0)):  # default_cfl_adapt (CConfig.cpp:1831-1836)
        c.cfl_adapt_param[k] = v
    return c
