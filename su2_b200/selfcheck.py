"""Multi-rank self-check: N slabs over NCCL against the SAME operator on one GPU, whole iterations.

Across ranks the reference's preconditioner is block-Jacobi: every rank factors and sweeps only its own rows and columns
(CSysMatrix.cpp:795,809 skip columns outside the owned range; :701-711 for thread partitions). One GPU reproduces exactly
that operator with `prec_partitions` = N whose boundaries are the ranks' row ranges (su2b200_linsys_set_partition_rows),
provided both runs number the points alike -- hence the generator's lexicographic order (rcm=False) here: slab r then
owns the contiguous global range [s0, s1) * plane in the same relative order.  With the same operator, the N-rank run
must follow the one-GPU run iteration by iteration: equal Krylov iteration counts, equal rms residuals, LinSysSol and
Solution to 1e-9 (the only difference left is the summation order of the allreduced dot products), CL/CD to 1e-6.
This covers what a pre-solve comparison cannot: the halo exchange of Krylov vectors after the preconditioner, the
in-stream allreduce of the Gram-Schmidt scalars, the rank-local ILU, CompleteImplicitIteration + the SOLUTION exchange
and the force reduction.

Used by tools/check_multigpu.py (pytest, needs 2 GPUs) and by bench.py --gpus N (before the timed region, so that the
driver's scaling run records `multigpu_parity` in every N > 1 line).
"""
import numpy as np

from .config import parse_cfg, flow_config_from_cfg, mesh_from_cfg
from .solver import FlowSolver


def _perturb(s, xyz, rans):
    # a smooth, position-dependent state (same on every rank, halos included): the points next to a slab boundary then
    # depend on exchanged data from the first gradient on
    bump = 1.0 + 0.02 * np.sin(3 * xyz[:, 0] + 1.0) * np.cos(2 * xyz[:, 1]) * np.sin(xyz[:, 2] + 0.5)
    nv = s.nVar
    for name in ("SOLUTION", "SOLUTION_OLD"):
        s.set(name, (s.get(name).reshape(-1, nv) * bump[:, None]).ravel())
    if rans:
        for name in ("TURB_SOLUTION", "TURB_SOLUTION_OLD"):  # [P] (SA) or [P][2] (SST)
            t = s.get(name)
            s.set(name, (t.reshape(len(bump), -1) * bump[:, None]).ravel())


def unstructured_mesh(opts, path, rank, world, nparts):
    """Local part `rank` of `world` of the .su2 mesh at `path` (su2b200_mesh_partition), or -- world == 1 -- the whole mesh
    renumbered colour by colour for `nparts` partitions.  Returns (preprocessed mesh, partition row offsets or None).
    No RCM: the local order of a rank's owned points is then the order of its colour group in the renumbered whole."""
    from .config import markers_from_cfg, _DEFAULTS
    from .mesh import Mesh
    o = dict(_DEFAULTS)
    o.update(opts)
    whole = Mesh.read_su2(path)
    for tag, (kind, mon, q) in markers_from_cfg(o).items():
        whole.set_marker(tag, kind, mon, q)
    rows = None
    if world > 1:
        m = whole.partition(world, rank)
    else:
        m, rows = whole.partition(nparts, -1)
    whole.close()
    m.preprocess(rcm=False)
    return m, rows


def run_iterations(cfg_text, n_iter, rank, world, device, comm_id=None, partition_planes=None, overrides=None,
                   su2_path=None, nparts=1):
    """`n_iter` iterations of the cfg's case on slab `rank` of `world` (or on the undivided mesh with the block-Jacobi
    partitions given by `partition_planes` = cumulative plane counts).  With `su2_path` the mesh is that unstructured
    file, decomposed by su2b200_mesh_partition.  Returns (global ids of the owned points, dict)."""
    opts = parse_cfg(cfg_text)
    part_rows = None
    if su2_path:
        mesh, part_rows = unstructured_mesh(opts, su2_path, rank, world, nparts)
    else:
        mesh = mesh_from_cfg(opts, rank=rank, nranks=world, rcm=False)
    s = FlowSolver(device)
    try:
        if comm_id is not None:
            uid = comm_id(s)
            allgather = None
            if isinstance(uid, tuple):  # (NCCL id, allgather of bytes): also connect the peer-memory mailboxes
                uid, allgather = uid
            s.comm_init(rank, world, uid)
            if allgather is not None and world > 1:
                s.mailbox_connect(allgather)
        mesh.upload(s)
        z = mesh.sizes()
        P, Pd = z["nPoint"], z["nPointDomain"]
        gid = mesh.array("global_index").astype(np.int64)
        nd = z["nDim"]
        xyz = mesh.array("coord").reshape(P, nd)
        if nd == 2:
            xyz = np.hstack([xyz, np.zeros((P, 1))])
        kw = flow_config_from_cfg(opts, nd)
        kw.update(overrides or {})
        s.configure(**kw)
        s.allocate()
        mesh.close()
        rans = bool(kw["rans_sa"]) or bool(kw.get("rans_sst"))
        if partition_planes is not None:
            plane = Pd // partition_planes[-1]
            part_rows = [p * plane for p in partition_planes]
        if part_rows is not None and len(part_rows) > 2:
            s.set_partition_rows(0, part_rows)
            if rans:
                s.set_partition_rows(1, part_rows)
        _perturb(s, xyz, rans)
        reps = s.iterate(n_iter)
        nv = s.nVar
        out = dict(
            solution=s.get("SOLUTION").reshape(-1, nv)[:Pd].copy(),
            linsyssol=s.get("LINSYSSOL").reshape(-1, nv)[:Pd].copy(),
            rms=np.array([r["rms_flow"] for r in reps]),
            iters=np.array([r["linsol_iter_flow"] for r in reps]),
            CL=np.array([r["CL"] for r in reps]), CD=np.array([r["CD"] for r in reps]))
        if rans:
            out["nu"] = s.get("TURB_SOLUTION").reshape(P, -1)[:Pd].copy()
            out["rms_turb"] = np.array([r["rms_turb"] for r in reps])
            out["iters_turb"] = np.array([r["linsol_iter_turb"] for r in reps])
        return gid[:Pd], out
    finally:
        s.close()


def slab_planes(n_planes, world):
    """Cumulative plane counts of the slab decomposition along the last axis (mesh_host.cu: make_box)."""
    return [n_planes * r // world for r in range(world + 1)]


def compare(parts, whole_gid, whole, single_precision=False):
    """parts: [(gid, out)] per rank.  Returns the summary dict that bench.py prints as `multigpu_parity`."""
    pos = {int(g): k for k, g in enumerate(whole_gid)}
    worst = 0.0
    detail = {}
    for g, part in parts:
        idx = np.array([pos[int(x)] for x in g])
        for key in ("solution", "linsyssol", "nu"):
            if key not in part:
                continue
            ref = whole[key][idx]
            scale = np.maximum(np.max(np.abs(whole[key]), axis=0), 1e-300)
            d = float(np.max(np.abs(part[key] - ref) / scale))
            detail[key] = max(detail.get(key, 0.0), d)
            worst = max(worst, d)
    p0 = parts[0][1]
    rms_rel = float(np.max(np.abs(p0["rms"] - whole["rms"]) / np.maximum(np.abs(whole["rms"]), 1e-300)))
    iters_equal = bool(np.array_equal(p0["iters"], whole["iters"]))
    if "iters_turb" in whole:
        iters_equal = iters_equal and bool(np.array_equal(p0["iters_turb"], whole["iters_turb"]))
        rms_rel = max(rms_rel, float(np.max(np.abs(p0["rms_turb"] - whole["rms_turb"]) / np.abs(whole["rms_turb"]))))
    coef = lambda k: float(np.max(np.abs(p0[k] - whole[k]) / np.maximum(np.abs(whole[k]), 1e-3)))
    return {"max_rel": worst, "rms_rel": rms_rel, "iters_equal": iters_equal, "cl_rel": coef("CL"), "cd_rel": coef("CD"),
            "iters": [int(x) for x in p0["iters"]], "iters_one_gpu": [int(x) for x in whole["iters"]],
            "detail": detail, "steps": int(len(p0["iters"])),
            # mixed precision: the ranks' dot products are summed in another order than on one GPU, in single precision
            "ok": bool(worst <= (1e-3 if single_precision else 1e-9) and rms_rel <= (1e-2 if single_precision else 1e-8)
                       and (iters_equal or single_precision) and coef("CL") <= (1e-4 if single_precision else 1e-6)
                       and coef("CD") <= (1e-4 if single_precision else 1e-6))}


def whole_iteration_parity(cfg_text, n_planes, n_iter, rank, world, device, comm_id, gather, overrides=None,
                           su2_path=None):
    """Collective over the ranks: every rank runs its slab, rank 0 also runs the undivided mesh with `world` block-Jacobi
    partitions cut at the slab boundaries and compares.  `gather(obj)` returns the list of every rank's obj on rank 0
    (None elsewhere); `comm_id(solver)` returns the broadcast 128-byte NCCL id.  Returns the summary on rank 0."""
    gid, mine = run_iterations(cfg_text, n_iter, rank, world, device, comm_id=comm_id, overrides=overrides,
                               su2_path=su2_path)
    parts = gather((gid, mine))
    if rank != 0:
        return None
    gw, whole = run_iterations(cfg_text, n_iter, 0, 1, device, overrides=overrides, su2_path=su2_path, nparts=world,
                               partition_planes=None if su2_path else slab_planes(n_planes, world))
    return compare(parts, gw, whole, single_precision=bool((overrides or {}).get("mixed_precision")))
