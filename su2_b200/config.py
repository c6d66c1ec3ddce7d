"""SU2 .cfg reader for the options the hot path consumes, and the non-dimensionalisation that turns them into
su2b200_flow_config values.

Mirrors CConfig::SetConfig_Parsing (Common/src/CConfig.cpp, `KEY= value` lines, `%` comments) and
CConfig::SetNondimensionalization (CConfig.cpp) / CEulerSolver::SetNondimensionalization
(SU2_CFD/src/solvers/CEulerSolver.cpp:799-1200) for the compressible, ideal-gas (STANDARD_AIR), SI-units case with
REF_DIMENSIONALIZATION = DIMENSIONAL or FREESTREAM_PRESS_EQ_ONE, and the free-stream SA values of the CTurbSASolver
constructor (CTurbSASolver.cpp:100-134).  Anything else raises: this is not a general CConfig.
"""
import math
import re
from . import capi

_DEFAULTS = {
    "SOLVER": "EULER", "FLUID_MODEL": "STANDARD_AIR", "KIND_TURB_MODEL": "NONE", "MACH_NUMBER": "0.0", "AOA": "0.0", "SIDESLIP_ANGLE": "0.0",
    "FREESTREAM_PRESSURE": "101325.0", "FREESTREAM_TEMPERATURE": "288.15", "REYNOLDS_NUMBER": "0.0",
    "REYNOLDS_LENGTH": "1.0", "REF_DIMENSIONALIZATION": "DIMENSIONAL", "GAMMA_VALUE": "1.4", "GAS_CONSTANT": "287.058",
    "MU_REF": "1.716E-5", "MU_T_REF": "273.15", "SUTHERLAND_CONSTANT": "110.4", "PRANDTL_LAM": "0.72",
    "PRANDTL_TURB": "0.90", "FREESTREAM_NU_FACTOR": "3.0", "REF_AREA": "1.0", "REF_LENGTH": "1.0",
    "REF_ORIGIN_MOMENT_X": "0.0", "REF_ORIGIN_MOMENT_Y": "0.0", "REF_ORIGIN_MOMENT_Z": "0.0",
    "NUM_METHOD_GRAD": "GREEN_GAUSS", "NUM_METHOD_GRAD_RECON": "", "CFL_NUMBER": "1.25", "CFL_REDUCTION_TURB": "1.0",
    "MAX_DELTA_TIME": "1000000", "CONV_NUM_METHOD_FLOW": "ROE", "MUSCL_FLOW": "YES", "MUSCL_KAPPA_FLOW": "0.0",
    "SLOPE_LIMITER_FLOW": "VENKATAKRISHNAN", "VENKAT_LIMITER_COEFF": "0.05", "JST_SENSOR_COEFF": "( 0.5, 0.02 )",
    "ROE_KAPPA": "0.5", "ENTROPY_FIX_COEFF": "0.001", "CENTRAL_JACOBIAN_FIX_FACTOR": "4.0",
    "LINEAR_SOLVER": "FGMRES", "LINEAR_SOLVER_PREC": "ILU", "LINEAR_SOLVER_ERROR": "1E-6", "LINEAR_SOLVER_ITER": "10",
    "LINEAR_SOLVER_RESTART_FREQUENCY": "10", "MAX_UPDATE_FLOW": "0.2", "MAX_UPDATE_SA": "0.99", "MAX_UPDATE_SST": "1.0",
    "SST_OPTIONS": "NONE", "FREESTREAM_TURBULENCEINTENSITY": "0.05", "FREESTREAM_TURB2LAMVISCRATIO": "10.0",
    "TIME_DISCRE_FLOW": "EULER_IMPLICIT", "MGLEVEL": "0", "SA_OPTIONS": "NONE", "MESH_FORMAT": "SU2",
    "MESH_FILENAME": "mesh.su2", "MESH_BOX_SIZE": "33, 33, 33", "MESH_BOX_LENGTH": "1.0, 1.0, 1.0",
    "MESH_BOX_OFFSET": "0.0, 0.0, 0.0", "ITER": "1", "CFL_ADAPT": "NO", "CFL_ADAPT_PARAM": "",
}
_MARKER_KEYS = {"MARKER_EULER": capi.EULER_WALL, "MARKER_FAR": capi.FAR_FIELD, "MARKER_SYM": capi.SYMMETRY_PLANE,
                "MARKER_HEATFLUX": capi.HEAT_FLUX}


def parse_cfg(text):
    """`KEY= value` lines -> dict of upper-case keys to raw strings (later lines win), `%` starts a comment."""
    opts = {}
    for line in text.splitlines():
        line = line.split("%", 1)[0].strip()
        if not line or "=" not in line:
            continue
        k, v = line.split("=", 1)
        opts[k.strip().upper()] = v.strip()
    return opts


def _list(s):
    return [x.strip() for x in re.sub(r"[()]", " ", s).split(",") if x.strip() and x.strip().upper() != "NONE"]


def markers_from_cfg(opts):
    """tag -> (kindBC, monitoring, heat_flux) from MARKER_EULER / _FAR / _SYM / _HEATFLUX / _MONITORING."""
    mon = set(_list(opts.get("MARKER_MONITORING", "")))
    out = {}
    for key, kind in _MARKER_KEYS.items():
        items = _list(opts.get(key, ""))
        if key == "MARKER_HEATFLUX":
            for tag, q in zip(items[0::2], items[1::2]):
                out[tag] = (kind, int(tag in mon), float(q))
        else:
            for tag in items:
                out[tag] = (kind, int(tag in mon), 0.0)
    return out


def _ref_elem_length(o):
    if o.get("SYSTEM_MEASUREMENTS", "SI").upper() != "SI":
        raise ValueError("SYSTEM_MEASUREMENTS= %s not supported (SI only)" % o["SYSTEM_MEASUREMENTS"])
    if "REF_ELEM_LENGTH" in o:
        raise ValueError("REF_ELEM_LENGTH is not an option of the reference (v8.4: RefElemLength is fixed to 1.0)")
    return 1.0


def flow_config_from_cfg(opts, nDim):
    """Values of su2b200_flow_config (keyword dict for FlowSolver.configure) from parsed cfg options."""
    o = dict(_DEFAULTS)
    o.update(opts)
    solver = o["SOLVER"].upper()
    if solver not in ("EULER", "NAVIER_STOKES", "RANS"):
        raise ValueError("SOLVER= %s is outside the hot path" % solver)
    viscous = solver != "EULER"
    rans = solver == "RANS"
    turb = o["KIND_TURB_MODEL"].upper()
    if rans and turb not in ("SA", "SST"):
        raise ValueError("KIND_TURB_MODEL= %s not supported (SA, SST)" % turb)
    sst = rans and turb == "SST"
    if sst and o["SST_OPTIONS"].upper() not in ("NONE", "V2003M"):
        raise ValueError("SST_OPTIONS= %s not supported (the default V2003m only)" % o["SST_OPTIONS"])
    rans = rans and not sst  # `rans` below means the SA model
    if o["TIME_DISCRE_FLOW"].upper() != "EULER_IMPLICIT":
        raise ValueError("only TIME_DISCRE_FLOW= EULER_IMPLICIT is on the path")
    if int(o["MGLEVEL"]) != 0:
        raise ValueError("MGLEVEL must be 0 (multigrid is a 'next' row)")
    fluid = o["FLUID_MODEL"].upper()
    if fluid == "STANDARD_AIR":  # CFluidModel selection in CEulerSolver::SetNondimensionalization: R = 287.058, gamma = 1.4
        gamma, R = 1.4, 287.058
    elif fluid == "IDEAL_GAS":
        gamma, R = float(o["GAMMA_VALUE"]), float(o["GAS_CONSTANT"])
    else:
        raise ValueError("FLUID_MODEL= %s not supported" % fluid)
    mach, T = float(o["MACH_NUMBER"]), float(o["FREESTREAM_TEMPERATURE"])
    alpha, beta = math.radians(float(o["AOA"])), math.radians(float(o["SIDESLIP_ANGLE"]))
    mu_ref, mu_tref, mu_s = float(o["MU_REF"]), float(o["MU_T_REF"]), float(o["SUTHERLAND_CONSTANT"])
    c = math.sqrt(gamma * R * T)
    vmag = mach * c
    if nDim == 2:
        vel = [math.cos(alpha) * vmag, math.sin(alpha) * vmag, 0.0]
    else:
        vel = [math.cos(alpha) * math.cos(beta) * vmag, math.sin(beta) * vmag, math.sin(alpha) * math.cos(beta) * vmag]
    mu_inf = 0.0
    if viscous:
        # INIT_OPTION= REYNOLDS: density from the Reynolds number, pressure from the gas law
        mu_inf = mu_ref * (T / mu_tref) ** 1.5 * (mu_tref + mu_s) / (T + mu_s)
        rho = float(o["REYNOLDS_NUMBER"]) * mu_inf / (vmag * float(o["REYNOLDS_LENGTH"]))
        p = rho * R * T
    else:
        p = float(o["FREESTREAM_PRESSURE"])
        rho = p / (R * T)
    refdim = o["REF_DIMENSIONALIZATION"].upper()
    if refdim == "DIMENSIONAL":
        p_ref = rho_ref = t_ref = 1.0
    elif refdim == "FREESTREAM_PRESS_EQ_ONE":
        p_ref, rho_ref, t_ref = p, rho, T
    else:
        raise ValueError("REF_DIMENSIONALIZATION= %s not supported" % refdim)
    v_ref = math.sqrt(p_ref / rho_ref)
    r_ref = v_ref * v_ref / t_ref
    visc_ref = rho_ref * v_ref * 1.0
    rho_nd, p_nd, t_nd = rho / rho_ref, p / p_ref, T / t_ref
    vel_nd = [v / v_ref for v in vel]
    R_nd = R / r_ref
    energy_nd = p_nd / (rho_nd * (gamma - 1.0)) + 0.5 * sum(v * v for v in vel_nd)
    mu_inf_nd = mu_inf / visc_ref
    nu_tilde = eddy = 0.0
    sa_neg = "NEGATIVE" in o["SA_OPTIONS"].upper()
    if rans:
        fac = float(o["FREESTREAM_NU_FACTOR"])
        nu_tilde = fac * mu_inf_nd / rho_nd
        if sa_neg:
            nu_tilde = 0.005 * fac * mu_inf_nd / rho_nd
        chi = nu_tilde / (mu_inf_nd / rho_nd)
        fv1 = chi ** 3 / (chi ** 3 + 7.1 ** 3)
        eddy = rho_nd * fv1 * nu_tilde
    kine_inf = omega_inf = 0.0
    if sst:
        # CTurbSSTSolver ctor (CTurbSSTSolver.cpp:118-131): k = 3/2 |v|^2 I^2, omega = rho k / (mu ratio); muT = rho k / omega
        inten, ratio = float(o["FREESTREAM_TURBULENCEINTENSITY"]), float(o["FREESTREAM_TURB2LAMVISCRATIO"])
        vel2 = sum(v * v for v in vel_nd[:nDim])
        kine_inf = 3.0 / 2.0 * (vel2 * inten * inten)
        omega_inf = rho_nd * kine_inf / (mu_inf_nd * ratio)
        eddy = rho_nd * kine_inf / omega_inf
        energy_nd += kine_inf  # if (tkeNeeded) Energy_FreeStreamND += Tke_FreeStreamND (CEulerSolver.cpp:1153)
    grad = {"GREEN_GAUSS": capi.GRAD_GREEN_GAUSS, "LEAST_SQUARES": capi.GRAD_LEAST_SQUARES,
            "WEIGHTED_LEAST_SQUARES": capi.GRAD_WEIGHTED_LEAST_SQUARES}
    gm = grad[o["NUM_METHOD_GRAD"].upper()]
    gr = grad[o["NUM_METHOD_GRAD_RECON"].upper()] if o["NUM_METHOD_GRAD_RECON"] else gm
    conv = o["CONV_NUM_METHOD_FLOW"].upper()
    if conv not in ("ROE", "JST"):
        raise ValueError("CONV_NUM_METHOD_FLOW= %s not supported (ROE, JST)" % conv)
    lim = {"NONE": capi.LIMITER_NONE, "VENKATAKRISHNAN": capi.LIMITER_VENKATAKRISHNAN}[o["SLOPE_LIMITER_FLOW"].upper()]
    sol = {"FGMRES": capi.SOLVER_FGMRES, "BCGSTAB": capi.SOLVER_BCGSTAB, "RESTARTED_FGMRES": capi.SOLVER_RESTARTED_FGMRES}
    pre = {"ILU": capi.PREC_ILU, "LU_SGS": capi.PREC_LU_SGS, "JACOBI": capi.PREC_JACOBI}
    jst = [float(x) for x in _list(o["JST_SENSOR_COEFF"])]
    muscl = int(o["MUSCL_FLOW"].upper() == "YES")
    if conv == "JST":  # centred schemes neither reconstruct nor limit (CConfig::SetPostprocessing)
        muscl, lim = 0, capi.LIMITER_NONE
    adapt = [0.1, 1.2, 10.0, 100.0, 0.001, 0.0]  # default_cfl_adapt (CConfig.cpp:1831-1836)
    for k, x in enumerate(_list(o["CFL_ADAPT_PARAM"])[:6]):
        adapt[k] = float(x)
    kw = dict(
        cfl_adapt=int(o["CFL_ADAPT"].upper() == "YES"), cfl_adapt_param=adapt, n_inner_iter=int(o["ITER"]),
        gamma=gamma, gas_constant=R_nd, viscous=int(viscous), rans_sa=int(rans), sa_negative=int(sa_neg),
        mu_ref=mu_ref / visc_ref, mu_temperature_ref=mu_tref / t_ref, mu_sutherland=mu_s / t_ref,
        mu_constant=mu_ref / visc_ref, prandtl_lam=float(o["PRANDTL_LAM"]), prandtl_turb=float(o["PRANDTL_TURB"]),
        density_inf=rho_nd, pressure_inf=p_nd, velocity_inf=vel_nd, energy_inf=energy_nd, temperature_inf=t_nd,
        nu_tilde_inf=nu_tilde, eddy_visc_inf=eddy, rans_sst=int(sst), kine_inf=kine_inf, omega_inf=omega_inf,
        max_update_sst=float(o["MAX_UPDATE_SST"]),
        conv_scheme=capi.CONV_JST if conv == "JST" else capi.CONV_ROE,
        muscl=muscl, muscl_kappa=float(o["MUSCL_KAPPA_FLOW"]), limiter=lim,
        venkat_coeff=float(o["VENKAT_LIMITER_COEFF"]),
        # this version of the reference has no REF_ELEM_LENGTH option any more: CConfig::SetPostprocessing sets
        # RefElemLength = 1.0 (CConfig.cpp:5061; /0.3048 under SYSTEM_MEASUREMENTS= US, which this front end rejects)
        ref_elem_length=_ref_elem_length(o), grad_method=gm, grad_method_recon=gr,
        roe_kappa=float(o["ROE_KAPPA"]), entropy_fix=float(o["ENTROPY_FIX_COEFF"]), jst_kappa2=jst[0], jst_kappa4=jst[1],
        cent_jac_fix_factor=float(o["CENTRAL_JACOBIAN_FIX_FACTOR"]), cfl=float(o["CFL_NUMBER"]),
        cfl_turb_ratio=float(o["CFL_REDUCTION_TURB"]), max_delta_time=float(o["MAX_DELTA_TIME"]),
        max_update_flow=float(o["MAX_UPDATE_FLOW"]), max_update_sa=float(o["MAX_UPDATE_SA"]),
        linear_solver=sol[o["LINEAR_SOLVER"].upper()], linear_prec=pre[o["LINEAR_SOLVER_PREC"].upper()],
        linear_tol=float(o["LINEAR_SOLVER_ERROR"]), linear_iter=int(o["LINEAR_SOLVER_ITER"]),
        linear_restart=int(o["LINEAR_SOLVER_RESTART_FREQUENCY"]),
        turb_linear_solver=sol[o["LINEAR_SOLVER"].upper()], turb_linear_prec=pre[o["LINEAR_SOLVER_PREC"].upper()],
        turb_linear_tol=float(o["LINEAR_SOLVER_ERROR"]), turb_linear_iter=int(o["LINEAR_SOLVER_ITER"]),
        aoa_deg=float(o["AOA"]), aos_deg=float(o["SIDESLIP_ANGLE"]), ref_area=float(o["REF_AREA"]),
        ref_length=float(o["REF_LENGTH"]),
        ref_origin=[float(o["REF_ORIGIN_MOMENT_X"]), float(o["REF_ORIGIN_MOMENT_Y"]), float(o["REF_ORIGIN_MOMENT_Z"])])
    return kw


def mesh_from_cfg(opts, rank=0, nranks=1, base_dir=".", rcm=True, tets=False, stretch=None):
    """MESH_FORMAT= BOX / RECTANGLE / SU2 -> preprocessed su2_b200.mesh.Mesh with the cfg's boundary conditions."""
    import os
    from .mesh import Mesh
    o = dict(_DEFAULTS)
    o.update(opts)
    fmt = o["MESH_FORMAT"].upper()
    if fmt in ("BOX", "RECTANGLE"):
        nd = 3 if fmt == "BOX" else 2
        n = [int(float(x)) for x in _list(o["MESH_BOX_SIZE"])][:nd]
        ln = [float(x) for x in _list(o["MESH_BOX_LENGTH"])][:nd]
        of = [float(x) for x in _list(o["MESH_BOX_OFFSET"])][:nd]
        m = Mesh.box(n, ln, of, rank=rank, nranks=nranks, tets=tets)
        if stretch:
            m.stretch(*stretch)
    elif fmt == "SU2":
        m = Mesh.read_su2(os.path.join(base_dir, o["MESH_FILENAME"]))
        if nranks != 1:  # every rank reads the file and keeps its part (SetColorGrid_Parallel + SetSendReceive)
            for tag, (kind, mon, q) in markers_from_cfg(o).items():
                m.set_marker(tag, kind, mon, q)
            whole, m = m, m.partition(nranks, rank)
            whole.close()
    else:
        raise ValueError("MESH_FORMAT= %s not supported" % fmt)
    for tag, (kind, mon, q) in markers_from_cfg(o).items():
        m.set_marker(tag, kind, mon, q)
    # the reference RCM-orders every mesh, generated boxes included (CPhysicalGeometry.cpp:4478, called from
    # CDriver::InitializeGeometryFVM): the ILU(0) factors depend on the ordering, so the default follows it
    m.preprocess(rcm=rcm)
    return m
