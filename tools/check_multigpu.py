"""Multi-GPU parity of the first implicit iteration: N slabs over NCCL against the undivided mesh on one GPU.

torchrun --nproc-per-node N tools/check_multigpu.py [EULER|NAVIER_STOKES|RANS|EULER_JST|NAVIER_STOKES_JST]
Every rank assembles the residual of iteration 0 on its slab (halo exchange of solution, gradients, limiters, nu~);
rank 0 also runs the undivided mesh and compares, by global point index, the rhs vector, the local time steps, the
limiters and the rms residuals: they must agree to round-off (the preconditioner is the only rank-dependent part of an
iteration, and it comes after). Prints `multigpu parity: max rel diff = X` and exits 0 iff X <= 1e-12.

With ITER_<kind> (ITER_RANS, ITER_EULER, ITER_NAVIER_STOKES) the check covers WHOLE iterations instead
(su2_b200/selfcheck.py): 5 iterations on N slabs against the undivided mesh with N block-Jacobi partitions cut at the
slab boundaries -- equal Krylov iteration counts, rms 1e-8, LinSysSol / Solution 1e-9, CL / CD 1e-6."""
import os, sys
import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (cfg template of the benchmark)
from su2_b200.config import parse_cfg, flow_config_from_cfg, mesh_from_cfg  # noqa: E402
from su2_b200.solver import FlowSolver  # noqa: E402


def first_iteration(opts, rank, world, device, comm=None):
    mesh = mesh_from_cfg(opts, rank=rank, nranks=world)
    s = FlowSolver(device)
    if comm is not None:
        s.comm_init(rank, world, comm(s))
    mesh.upload(s)
    gid = mesh.array("global_index").astype(np.int64)
    xyz = mesh.array("coord").reshape(len(gid), 3)
    Pd = mesh.sizes()["nPointDomain"]
    s.configure(**flow_config_from_cfg(opts, 3))
    s.allocate()
    mesh.close()
    rans = bool(opts.get("KIND_TURB_MODEL", "NONE") not in ("NONE", ""))
    out = {}
    # a smooth, position-dependent state (same on every rank, halos included): gradients, limiters and fluxes of the
    # points next to a slab boundary then depend on exchanged data
    bump = 1.0 + 0.02 * np.sin(3 * xyz[:, 0] + 1.0) * np.cos(2 * xyz[:, 1]) * np.sin(xyz[:, 2] + 0.5)
    nv = s.nVar
    for name in ("SOLUTION", "SOLUTION_OLD"):
        s.set(name, (s.get(name).reshape(-1, nv) * bump[:, None]).ravel())
    if rans:
        for name in ("TURB_SOLUTION", "TURB_SOLUTION_OLD"):
            s.set(name, s.get(name) * bump)
    s.Preprocessing()
    s.SetTime_Step()
    s.Space_Integration()
    rms, _, _ = s.PrepareImplicitIteration()
    out["rhs"] = s.get("LINSYSRES").reshape(-1, nv)[:Pd]
    out["dt"] = s.get("DELTA_TIME")[:Pd]
    lim = s.get("LIMITER")
    if lim.size:
        out["lim"] = lim.reshape(len(gid), -1)[:Pd]
    for name, key in (("SENSOR", "sensor"), ("UND_LAPL", "lapl"), ("LAMBDA", "lambda")):
        a = s.get(name)
        if a.size:
            out[key] = a.reshape(len(gid), -1)[:Pd]
    out["rms"] = rms
    if rans:
        s.Turb_Preprocessing()
        s.Turb_SetTime_Step()
        s.Turb_Space_Integration()
        out["rms_turb"] = np.array([s.Turb_PrepareImplicitIteration()])
        out["turb_rhs"] = s.get("TURB_LINSYSRES")[:Pd]
    s.close()
    return gid[:Pd], out


def whole_iterations(kind, rank, world, local):
    from su2_b200.selfcheck import whole_iteration_parity
    n_planes = 6 * world + 1
    su2_path = None
    if kind.startswith("TET_"):
        # unstructured path: a tetrahedral box written as .su2 by rank 0, decomposed by su2b200_mesh_partition
        kind = kind[4:]
        su2_path = "/tmp/su2b200_check_tet_%s.su2" % os.environ.get("MASTER_PORT", "0")
        if rank == 0:
            from su2_b200.mesh import Mesh
            m = Mesh.box([11, 9, 8], [2.0, 1.0, 1.0], tets=True)
            m.write_su2(su2_path)
            m.close()
        dist.barrier()
    sst = kind == "SST"
    mixed = kind == "MIXED"  # RANS-SA with the single-precision linear solve (float halo exchange, float allreduces)
    text = bench.make_cfg("RANS" if (sst or mixed) else kind, [12, 10, n_planes], 1)
    if sst:
        text = text.replace("KIND_TURB_MODEL= SA", "KIND_TURB_MODEL= SST")
    if su2_path:
        text = text.replace("MESH_FORMAT= BOX", "MESH_FORMAT= SU2\nMESH_FILENAME= " + su2_path)

    def allgather_bytes(b):
        t = torch.frombuffer(bytearray(b), dtype=torch.uint8).cuda()
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return [o.cpu().numpy().tobytes() for o in out]

    def comm(s):
        idbuf = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idbuf.copy_(torch.from_numpy(s.comm_unique_id()).cuda())
        dist.broadcast(idbuf, 0)
        return idbuf.cpu().numpy(), allgather_bytes

    def gather(obj):
        parts = [None] * world
        dist.all_gather_object(parts, obj)
        return parts

    res = whole_iteration_parity(text, n_planes, 5, rank, world, local, comm, gather, su2_path=su2_path,
                                 overrides={"mixed_precision": 1} if mixed else None)
    ok = 1.0
    if rank == 0:
        print("multigpu parity, whole iterations (%s, %d ranks): %s" % (kind, world, res))
        ok = 1.0 if res["ok"] else 0.0
    flag = torch.tensor([ok], device="cuda")
    dist.broadcast(flag, 0)
    return flag.item() == 1.0


def main():
    kind = sys.argv[1] if len(sys.argv) > 1 else "RANS"
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if kind.startswith("ITER_"):
        ok = whole_iterations(kind[5:], rank, world, local)
        dist.destroy_process_group()
        sys.exit(0 if ok else 1)
    n = [12, 10, 8 * world + 1]
    jst = kind.endswith("_JST")
    text = bench.make_cfg(kind[:-4] if jst else kind, n, 1)
    if jst:  # centred scheme: max eigenvalue, pressure sensor and undivided Laplacian are exchanged as well
        text = text.replace("CONV_NUM_METHOD_FLOW= ROE", "CONV_NUM_METHOD_FLOW= JST").replace("MUSCL_FLOW= YES", "MUSCL_FLOW= NO")
    opts = parse_cfg(text)

    def comm(s):
        idbuf = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idbuf.copy_(torch.from_numpy(s.comm_unique_id()).cuda())
        dist.broadcast(idbuf, 0)
        return idbuf.cpu().numpy()

    gid, mine = first_iteration(opts, rank, world, local, comm)
    # gather the slabs on rank 0 (object gather over the process group: small arrays)
    parts = [None] * world
    dist.all_gather_object(parts, (gid, {k: v for k, v in mine.items()}))
    worst = 0.0
    if rank == 0:
        gw, whole = first_iteration(opts, 0, 1, local)
        pos = {int(g): k for k, g in enumerate(gw)}
        for g, part in parts:
            idx = np.array([pos[int(x)] for x in g])
            for key in ("rhs", "dt", "lim", "sensor", "lapl", "lambda", "turb_rhs"):
                if key not in part:
                    continue
                a, b = part[key], whole[key][idx]
                d = np.max(np.abs(a - b)) / max(np.max(np.abs(whole[key])), 1e-300)
                print("  rank slab %-9s max rel diff %.3e" % (key, d))
                worst = max(worst, d)
        for key in ("rms", "rms_turb"):
            if key in whole:
                d = np.max(np.abs(parts[0][1][key] - whole[key]) / np.abs(whole[key]))
                print("  %-19s max rel diff %.3e" % (key, d))
                worst = max(worst, d)
        print("multigpu parity (%s, %d ranks): max rel diff = %.3e" % (kind, world, worst))
    flag = torch.tensor([1.0 if worst <= 1e-12 else 0.0], device="cuda")
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1.0 else 1)


if __name__ == "__main__":
    main()
