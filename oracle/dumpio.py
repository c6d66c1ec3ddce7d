"""TEST INFRASTRUCTURE ONLY. Reader for the files written by oracle/oracle_dump.cpp (the unmodified
reference driven through its own classes) and converters to the argument dictionaries of
su2_b200.solver.FlowSolver.  Golden fixtures under tests/golden/ are these dumps re-packed as .npz
by oracle/make_golden.py.
"""
import struct
import numpy as np

_DT = {0: np.float64, 1: np.int64, 2: np.int8, 3: np.uint8}


def load_dump(path):
    out = {}
    with open(path, "rb") as f:
        assert f.read(8) == b"SU2DUMP1", "not an oracle dump"
        while True:
            h = f.read(4)
            if len(h) < 4:
                break
            (nl,) = struct.unpack("<I", h)
            name = f.read(nl).decode()
            (dt,) = struct.unpack("<B", f.read(1))
            (nd,) = struct.unpack("<I", f.read(4))
            dims = struct.unpack("<%dQ" % nd, f.read(8 * nd))
            n = int(np.prod(dims)) if nd else 1
            dtype = np.dtype(_DT[dt])
            a = np.frombuffer(f.read(n * dtype.itemsize), dtype=dtype).reshape(dims)
            if dt == 3:
                out[name] = a.tobytes().decode()
            elif a.size == 1 and nd == 1 and (name.startswith("cfg_") or name.startswith("enum_") or name.split("_", 1)[-1] in (
                    "nDim", "nPoint", "nPointDomain", "nEdge", "nMarker") or name in (
                    "nDim", "nPoint", "nPointDomain", "nEdge", "nMarker", "nIter") or name.endswith(("_kind", "_monitoring", "_heat_flux", "_iter", "_res", "_dt"))):
                out[name] = a.reshape(()).item()
            else:
                out[name] = a.copy()
    return out


def load_golden(path):
    z = np.load(path, allow_pickle=False)
    out = {}
    for k in z.files:
        a = z[k]
        if a.dtype.kind == "U":
            out[k] = str(a)
        elif a.shape == ():
            out[k] = a.item()
        else:
            out[k] = a
    return out


def geometry_from_dump(d, level=0):
    """CGeometry arrays of the finest grid (level 0) or of agglomerated multigrid level `level` ("mg<level>_" keys)."""
    G = "" if level == 0 else "mg%d_" % level
    g = {k: d[G + k] for k in ("nDim", "nPoint", "nPointDomain", "nEdge", "edge_nodes", "edge_normal", "coord", "volume",
                               "point_ptr", "point_nbr", "point_edge", "row_ptr", "col_ind", "dia_ptr", "edge_ptr",
                               "nNeighbor", "physical_boundary", "wall_distance")}
    markers = []
    for m in range(int(d[G + "nMarker"])):
        p = G + "marker%d_" % m
        markers.append(dict(kind=int(d[p + "kind"]), node=d[p + "node"], normal=d[p + "normal"],
                            normal_neighbor=d[p + "normal_neighbor"], sym_unit_normal=d[p + "sym_unit_normal"],
                            monitoring=int(d[p + "monitoring"]), heat_flux=float(d[p + "heat_flux"]), tag=d[p + "tag"]))
    g["markers"] = markers
    return g


def config_from_dump(d):
    """Map the dumped CConfig getters to su2b200_flow_config fields."""
    from su2_b200 import capi
    grad = {d["enum_GREEN_GAUSS"]: capi.GRAD_GREEN_GAUSS, d["enum_LEAST_SQUARES"]: capi.GRAD_LEAST_SQUARES,
            d["enum_WEIGHTED_LEAST_SQUARES"]: capi.GRAD_WEIGHTED_LEAST_SQUARES}
    lim = {d["enum_LIMITER_NONE"]: capi.LIMITER_NONE, d["enum_LIMITER_VENKATAKRISHNAN"]: capi.LIMITER_VENKATAKRISHNAN}
    sol = {d["enum_FGMRES"]: capi.SOLVER_FGMRES, d["enum_BCGSTAB"]: capi.SOLVER_BCGSTAB,
           d["enum_RESTARTED_FGMRES"]: capi.SOLVER_RESTARTED_FGMRES}
    pre = {d["enum_ILU"]: capi.PREC_ILU, d["enum_LU_SGS"]: capi.PREC_LU_SGS, d["enum_JACOBI"]: capi.PREC_JACOBI}
    gm = grad[d["cfg_grad_method"]]
    gr = grad.get(d["cfg_grad_method_recon"], gm)
    kw = dict(
        gamma=d["cfg_gamma"], gas_constant=d["cfg_gas_constant"], viscous=int(d["cfg_viscous"]),
        rans_sa=int(d["cfg_rans_sa"]), sa_negative=int(d["cfg_sa_negative"]),
        mu_ref=d["cfg_mu_ref"], mu_temperature_ref=d["cfg_mu_temperature_ref"],
        mu_sutherland=d["cfg_mu_sutherland"] if d["cfg_viscosity_model_sutherland"] else -1.0,
        mu_constant=d["cfg_mu_constant"], prandtl_lam=d["cfg_prandtl_lam"], prandtl_turb=d["cfg_prandtl_turb"],
        density_inf=d["cfg_density_inf"], pressure_inf=d["cfg_pressure_inf"], velocity_inf=list(d["cfg_velocity_inf"]),
        energy_inf=d["cfg_energy_inf"], temperature_inf=d["cfg_temperature_inf"], nu_tilde_inf=d["cfg_nu_tilde_inf"],
        conv_scheme=capi.CONV_JST if d["cfg_conv_centered"] else capi.CONV_ROE,
        muscl=int(d["cfg_muscl"]), muscl_kappa=d["cfg_muscl_kappa"], limiter=lim[d["cfg_limiter"]],
        venkat_coeff=d["cfg_venkat_coeff"], ref_elem_length=d["cfg_ref_elem_length"],
        grad_method=gm, grad_method_recon=gr, roe_kappa=d["cfg_roe_kappa"], entropy_fix=d["cfg_entropy_fix"],
        jst_kappa2=d["cfg_jst_kappa2"], jst_kappa4=d["cfg_jst_kappa4"], cent_jac_fix_factor=d["cfg_cent_jac_fix_factor"],
        cfl=d["cfg_cfl"], cfl_turb_ratio=d["cfg_cfl_turb_ratio"], max_delta_time=d["cfg_max_delta_time"],
        max_update_flow=d["cfg_max_update_flow"], max_update_sa=d["cfg_max_update_sa"],
        linear_solver=sol[d["cfg_linear_solver"]], linear_prec=pre[d["cfg_linear_prec"]],
        linear_tol=d["cfg_linear_tol"], linear_iter=int(d["cfg_linear_iter"]), linear_restart=int(d["cfg_linear_restart"]),
        turb_linear_solver=sol[d["cfg_linear_solver"]], turb_linear_prec=pre[d["cfg_linear_prec"]],
        turb_linear_tol=d["cfg_linear_tol"], turb_linear_iter=int(d["cfg_linear_iter"]),
        aoa_deg=d["cfg_aoa"], aos_deg=d["cfg_aos"], ref_area=d["cfg_ref_area"], ref_length=d["cfg_ref_length"],
        ref_origin=list(d["cfg_ref_origin"]),
        # CScalarSolver::ReducerStrategy of the run that made the fixture (dumped by newer fixtures; else the rule applies)
        turb_reducer=int(d.get("cfg_turb_reducer", -1)))
    if "cfg_cfl_adapt" in d:  # fixtures made since CFL adaptation exists
        kw.update(cfl_adapt=int(d["cfg_cfl_adapt"]), n_inner_iter=int(d["cfg_n_inner_iter"]),
                  cfl_adapt_param=[float(x) for x in d["cfg_cfl_adapt_param"]])
    if int(d.get("cfg_rans_sst", 0)):
        assert int(d["cfg_sst_v2003"]) == 1 and int(d["cfg_sst_dll"]) == 0, "fixture made with SST options outside V2003m"
        kw.update(rans_sst=1, kine_inf=float(d["cfg_kine_inf"]), omega_inf=float(d["cfg_omega_inf"]),
                  max_update_sst=float(d["cfg_max_update_sst"]))
        kw["eddy_visc_inf"] = kw["density_inf"] * kw["kine_inf"] / kw["omega_inf"]
    if kw["rans_sa"]:
        # free-stream eddy viscosity of the CTurbSASolver constructor (CTurbSASolver.cpp:100-134)
        t = kw["temperature_inf"]
        mu = kw["mu_ref"] * (t / kw["mu_temperature_ref"]) ** 1.5 * (kw["mu_temperature_ref"] + kw["mu_sutherland"]) / (t + kw["mu_sutherland"])
        chi = kw["nu_tilde_inf"] / (mu / kw["density_inf"])
        kw["eddy_visc_inf"] = kw["density_inf"] * chi ** 3 / (chi ** 3 + 7.1 ** 3) * kw["nu_tilde_inf"]
    return kw
