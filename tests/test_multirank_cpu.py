"""world_size-2/3 `gloo` tests of the host logic behind the multi-GPU path: every rank builds its slab of the box
(su2b200_mesh_create_box with rank/nranks), and the send/receive tables it would hand to su2b200_p2p_upload are
exercised with real point-to-point messages.  The device path uses the same tables over NCCL (csrc/comm.cu)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from su2_b200 import capi
from su2_b200.mesh import Mesh

N, LENGTH = [5, 4, 11], [1.0, 0.8, 2.2]
TAGS = ("x_minus", "x_plus", "y_minus", "y_plus", "z_minus", "z_plus")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _field(gid):
    return np.stack([np.sin(0.37 * gid) + 2.0, 1e-3 * gid, np.cos(0.11 * gid)], axis=1)


def _worker(rank, world, port, errs):
    try:
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        m = Mesh.box(N, LENGTH, rank=rank, nranks=world)
        for t in TAGS:
            m.set_marker(t, capi.FAR_FIELD)
        m.preprocess()
        z = m.sizes()
        P, Pd = z["nPoint"], z["nPointDomain"]
        gid = m.array("global_index").astype(np.int64)
        f = _field(gid.astype(np.float64))
        f[Pd:] = np.nan  # halo values must come from the owners
        ranks = m.array("p2p_ranks")
        sp, snd = m.array("p2p_send_ptr"), m.array("p2p_send").astype(np.int64)
        rp, rcv = m.array("p2p_recv_ptr"), m.array("p2p_recv").astype(np.int64)
        # InitiateComms / CompleteComms (CSolver.cpp:1376, 1582): pack owned values, post, unpack into the halo
        reqs, bufs = [], []
        for k, q in enumerate(ranks):
            out = torch.from_numpy(np.ascontiguousarray(f[snd[sp[k]:sp[k + 1]]]))
            inc = torch.empty((rp[k + 1] - rp[k], 3), dtype=torch.float64)
            reqs.append(dist.isend(out, int(q)))
            reqs.append(dist.irecv(inc, int(q)))
            bufs.append((k, inc, out))
        for r in reqs:
            r.wait()
        for k, inc, _ in bufs:
            f[rcv[rp[k]:rp[k + 1]]] = inc.numpy()
        assert not np.isnan(f).any(), "halo point not covered by a receive list"
        assert np.array_equal(f, _field(gid.astype(np.float64)))
        # global reductions the iteration needs (rms over the global owned count, summed volume)
        vol = torch.tensor([float(m.array("volume")[:Pd].sum()), float(Pd)], dtype=torch.float64)
        dist.all_reduce(vol)
        assert abs(vol[0].item() - LENGTH[0] * LENGTH[1] * LENGTH[2]) < 1e-12
        assert int(vol[1].item()) == N[0] * N[1] * N[2]
        # every edge a rank holds touches at least one owned point; halo-halo edges are dropped
        e0, e1 = m.array("edge_node0").astype(np.int64), m.array("edge_node1").astype(np.int64)
        assert np.all((e0 < Pd) | (e1 < Pd))
        m.close()
        dist.barrier()
        dist.destroy_process_group()
    except Exception as e:  # pragma: no cover - reported to the parent
        import traceback
        errs.put("rank %d: %s\n%s" % (rank, e, traceback.format_exc()))


@pytest.mark.parametrize("world", [2, 3])
def test_halo_tables_exchange_over_gloo(world):
    ctx = mp.get_context("spawn")
    errs = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, errs)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
    msgs = []
    while not errs.empty():
        msgs.append(errs.get())
    for p in procs:
        if p.is_alive():
            p.terminate()
            msgs.append("worker timed out")
    assert not msgs, "\n".join(msgs)
    assert all(p.exitcode == 0 for p in procs)


def test_selfcheck_compare_flags_a_divergent_rank():
    """su2_b200.selfcheck.compare (the `multigpu_parity` summary of bench.py): slabs reassembled by global index."""
    import numpy as np
    from su2_b200.selfcheck import compare, slab_planes
    assert slab_planes(13, 2) == [0, 6, 13] and slab_planes(25, 4) == [0, 6, 12, 18, 25]
    rng = np.random.default_rng(3)
    n, nv, k = 40, 5, 4
    whole = dict(solution=rng.standard_normal((n, nv)), linsyssol=rng.standard_normal((n, nv)),
                 rms=np.abs(rng.standard_normal((k, nv))), iters=np.array([5, 6, 6, 7]),
                 CL=rng.standard_normal(k), CD=rng.standard_normal(k))
    gid = rng.permutation(n)
    owned = [gid[:17], gid[17:]]
    pos = {int(g): i for i, g in enumerate(np.arange(n))}
    def part(g, eps=0.0):
        idx = np.array([pos[int(x)] for x in g])
        return g, dict(solution=whole["solution"][idx] * (1 + eps), linsyssol=whole["linsyssol"][idx], rms=whole["rms"],
                       iters=whole["iters"].copy(), CL=whole["CL"], CD=whole["CD"])
    good = compare([part(owned[0]), part(owned[1])], np.arange(n), whole)
    assert good["ok"] and good["max_rel"] == 0.0 and good["iters_equal"]
    bad = compare([part(owned[0]), part(owned[1], 1e-6)], np.arange(n), whole)
    assert not bad["ok"] and bad["max_rel"] > 1e-9
    p0 = part(owned[0])
    p0[1]["iters"][2] += 1
    assert not compare([p0, part(owned[1])], np.arange(n), whole)["ok"]
