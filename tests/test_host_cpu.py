"""CPU-side tests (no GPU): the C-ABI library loads and exports what include/su2b200.h declares, the host geometry
preprocessing reproduces the reference's dual grid (golden fixtures dumped from the unmodified reference), the cfg
front end reproduces the reference's CConfig values, and the slab decomposition is self-consistent."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from oracle.dumpio import load_golden, config_from_dump
from su2_b200 import capi
from su2_b200.config import parse_cfg, flow_config_from_cfg, mesh_from_cfg
from su2_b200.mesh import Mesh

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = ["naca0012_roe", "naca0012_jst", "box_euler", "rect_ns", "box_rans_sa", "rect_rans_sa", "naca0012_jacobi",
         "naca0012_fgmres_lusgs", "naca0012_bcgstab_ilu", "naca0012_rfgmres", "naca0012_ls", "box_rans_sst", "rect_rans_sst", "rect_rans_sst_wls"]
GENERATED = ["box_euler", "rect_ns", "box_rans_sa"]


def _header_symbols():
    txt = open(os.path.join(ROOT, "include", "su2b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    txt = re.sub(r"//[^\n]*", "", txt)
    return sorted(set(re.findall(r"\b(su2b200_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    lib = capi.load()
    declared = _header_symbols()
    assert len(declared) > 40
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, missing


def test_symbol_list_matches_header():
    assert sorted(capi.EXPORTED_SYMBOLS) == _header_symbols()


def test_compute_entry_points_fail_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from su2_b200.solver import FlowSolver
    with pytest.raises(Exception):
        FlowSolver(0)


def test_errors_are_reported_through_the_abi():
    lib = capi.load()
    lib.su2b200_last_error.restype = C.c_char_p
    h = C.c_void_p()
    rc = lib.su2b200_mesh_read_su2(b"/nonexistent/mesh.su2", C.byref(h))
    assert rc != 0
    assert b"mesh" in lib.su2b200_last_error().lower() or lib.su2b200_last_error()


@pytest.mark.parametrize("case", CASES)
def test_cfg_front_end_matches_reference_config(case):
    gold = load_golden(os.path.join(ROOT, "tests", "golden", case + ".npz"))
    ref = config_from_dump(gold)
    opts = parse_cfg(open(os.path.join(ROOT, "tests", "cases", case + ".cfg")).read())
    got = flow_config_from_cfg(opts, int(gold["nDim"]))
    skip = {"turb_linear_solver", "turb_linear_prec", "turb_linear_tol", "turb_linear_iter"}
    if not ref["rans_sa"]:
        skip |= {"nu_tilde_inf"}  # set by the reference whatever the solver, unused without SA
        if not ref.get("rans_sst"):
            skip |= {"eddy_visc_inf"}  # set by the reference whatever the solver, unused without SA
    for k, v in ref.items():
        if k in skip or k not in got:
            continue
        a, b = np.asarray(got[k], dtype=np.float64).ravel(), np.asarray(v, dtype=np.float64).ravel()
        n = min(a.size, b.size)
        assert np.allclose(a[:n], b[:n], rtol=1e-12, atol=1e-14), (k, got[k], v)


def _key(coord):
    return [tuple(np.round(c, 9)) for c in coord]


@pytest.mark.parametrize("case", GENERATED)
def test_generated_mesh_matches_reference_dual_grid(case):
    gold = load_golden(os.path.join(ROOT, "tests", "golden", case + ".npz"))
    opts = parse_cfg(open(os.path.join(ROOT, "tests", "cases", case + ".cfg")).read())
    m = mesh_from_cfg(opts)
    z = m.sizes()
    nd, P, E = int(gold["nDim"]), int(gold["nPoint"]), int(gold["nEdge"])
    assert (z["nDim"], z["nPoint"], z["nPointDomain"], z["nEdge"]) == (nd, P, int(gold["nPointDomain"]), E)
    gc = np.asarray(gold["coord"]).reshape(P, nd)
    mc = m.array("coord").reshape(P, nd)
    gidx = {k: i for i, k in enumerate(_key(gc))}
    to_gold = np.array([gidx[k] for k in _key(mc)])  # my local index -> reference index
    assert len(set(to_gold.tolist())) == P
    # control volumes, neighbour counts, boundary flags, wall distance
    assert np.allclose(m.array("volume"), np.asarray(gold["volume"])[to_gold], rtol=1e-13, atol=0)
    assert np.array_equal(m.array("nNeighbor"), np.asarray(gold["nNeighbor"])[to_gold])
    assert np.array_equal(m.array("physical_boundary") != 0, np.asarray(gold["physical_boundary"])[to_gold] != 0)
    if int(gold["cfg_rans_sa"]):
        assert np.allclose(m.array("wall_distance"), np.asarray(gold["wall_distance"])[to_gold], rtol=1e-12, atol=1e-14)
    # edges: same set of node pairs, normals equal up to the orientation of the pair
    gen = np.asarray(gold["edge_nodes"]).reshape(2, E).T
    gnorm = np.asarray(gold["edge_normal"]).reshape(E, nd)
    gmap = {(int(a), int(b)): k for k, (a, b) in enumerate(gen)}
    e0, e1 = m.array("edge_node0"), m.array("edge_node1")
    mnorm = m.array("edge_normal").reshape(E, nd)
    assert np.all(e0 < e1)
    for k in range(E):
        a, b = int(to_gold[e0[k]]), int(to_gold[e1[k]])
        if (a, b) in gmap:
            ref = gnorm[gmap[(a, b)]]
        else:
            ref = -gnorm[gmap[(b, a)]]
        assert np.allclose(mnorm[k], ref, rtol=1e-12, atol=1e-15), (k, mnorm[k], ref)
    # closed control volumes: the dual faces of an interior point sum to zero
    div = np.zeros((P, nd))
    np.add.at(div, e0.astype(np.int64), mnorm)
    np.add.at(div, e1.astype(np.int64), -mnorm)
    interior = m.array("physical_boundary") == 0
    assert np.max(np.abs(div[interior])) < 1e-13
    # markers: vertex normals (stored pointing into the fluid like the reference's), normal neighbours
    tags = {str(gold["marker%d_tag" % i]): i for i in range(int(gold["nMarker"]))}
    for im in range(z["nMarker"]):
        info = m.marker_info(im)
        ig = tags[info["tag"]]
        assert info["kind"] == int(gold["marker%d_kind" % ig])
        assert info["monitoring"] == int(gold["marker%d_monitoring" % ig])
        gnode = np.asarray(gold["marker%d_node" % ig]).astype(np.int64)
        gn = np.asarray(gold["marker%d_normal" % ig]).reshape(-1, nd)
        gnb = np.asarray(gold["marker%d_normal_neighbor" % ig]).astype(np.int64)
        gu = np.asarray(gold["marker%d_sym_unit_normal" % ig]).reshape(-1, nd)
        pos = {int(p): k for k, p in enumerate(gnode)}
        node = m.array("marker%d_node" % im).astype(np.int64)
        nrm = m.array("marker%d_normal" % im).reshape(-1, nd)
        nbr = m.array("marker%d_normal_neighbor" % im).astype(np.int64)
        unit = m.array("marker%d_sym_unit_normal" % im).reshape(-1, nd)
        assert node.size == gnode.size
        for k, p in enumerate(node):
            g = pos[int(to_gold[p])]
            assert np.allclose(nrm[k], gn[g], rtol=1e-12, atol=1e-15)
            assert to_gold[nbr[k]] == gnb[g]
            if gu.size and unit.size:
                assert np.allclose(unit[k], gu[g], rtol=1e-12, atol=1e-14)
    m.close()


PINNED = ["naca0012_roe", "box_euler", "rect_ns", "box_rans_sa"]


@pytest.mark.parametrize("case", PINNED)
def test_own_geometry_path_reproduces_the_reference_numbering(case):
    """Mesh reader / generator + RCM + dual grid against the reference's dump INDEX BY INDEX: the point permutation of
    SetRCM_Ordering (CPhysicalGeometry.cpp:4478-4621), the edge numbering of SetEdges (CGeometry.cpp:1263), CPoint::Point
    / CPoint::Edge in the reference's order, the sparse pattern and the edge -> block map, normals and volumes."""
    gold = load_golden(os.path.join(ROOT, "tests", "golden", case + ".npz"))
    opts = parse_cfg(open(os.path.join(ROOT, "tests", "cases", case + ".cfg")).read())
    m = mesh_from_cfg(opts, base_dir="/root/reference/QuickStart" if case.startswith("naca") else ".") \
        if (not case.startswith("naca") or os.path.isdir("/root/reference/QuickStart")) else None
    if m is None:
        pytest.skip("QuickStart mesh not on this machine")
    E = int(gold["nEdge"])
    assert np.array_equal(m.array("global_index").astype(np.int64), np.asarray(gold["global_index"]).astype(np.int64))
    en = np.asarray(gold["edge_nodes"]).reshape(2, E)
    assert np.array_equal(m.array("edge_node0"), en[0]) and np.array_equal(m.array("edge_node1"), en[1])
    for k in ("point_ptr", "point_nbr", "point_edge", "row_ptr", "col_ind", "dia_ptr", "edge_ptr", "nNeighbor"):
        assert np.array_equal(m.array(k).astype(np.int64), np.asarray(gold[k]).ravel().astype(np.int64)), k
    scale = np.max(np.abs(gold["edge_normal"]))
    assert np.max(np.abs(m.array("edge_normal") - np.asarray(gold["edge_normal"]).ravel())) <= 1e-13 * scale
    assert np.allclose(m.array("volume"), gold["volume"], rtol=1e-11, atol=0)  # sums of up to 8 sub-volumes, other order
    assert np.allclose(m.array("coord"), np.asarray(gold["coord"]).ravel(), rtol=0, atol=1e-15)
    m.close()


def test_su2_writer_round_trip(tmp_path):
    """su2b200_mesh_write_su2 -> su2b200_mesh_read_su2 gives the same dual grid (tetrahedral and stretched boxes)."""
    tags = ("x_minus", "x_plus", "y_minus", "y_plus", "z_minus", "z_plus")
    for tets, stretch in ((True, None), (False, (1, 1e-3, 1.2))):
        a = Mesh.box([5, 6, 4], [2.0, 1.0, 1.0], tets=tets)
        if stretch:
            a.stretch(*stretch)
        path = tmp_path / ("box_%d.su2" % tets)
        a.write_su2(path)
        b = Mesh.read_su2(path)
        for m in (a, b):
            for t in tags:
                m.set_marker(t, capi.FAR_FIELD)
            m.preprocess(rcm=True)
        assert a.sizes() == b.sizes()
        for k in ("coord", "volume", "edge_normal", "edge_node0", "edge_node1", "col_ind", "edge_ptr", "global_index"):
            assert np.array_equal(a.array(k), b.array(k)), k
        if stretch:
            y = np.unique(np.round(a.array("coord").reshape(-1, 3)[:, 1], 12))
            d = np.diff(y)
            assert np.allclose(d[1:] / d[:-1], 1.2, rtol=1e-9) and abs(y[-1] - 1.0) < 1e-12
        a.close()
        b.close()


def test_sparse_pattern_and_edge_map_are_consistent():
    m = Mesh.box([7, 6, 5], [1.0, 1.0, 1.0])
    for tag in ("x_minus", "x_plus", "y_minus", "y_plus", "z_minus", "z_plus"):
        m.set_marker(tag, capi.FAR_FIELD)
    m.preprocess()
    z = m.sizes()
    P, E = z["nPoint"], z["nEdge"]
    rp, ci, dp, ep = (m.array(k).astype(np.int64) for k in ("row_ptr", "col_ind", "dia_ptr", "edge_ptr"))
    e0, e1 = m.array("edge_node0").astype(np.int64), m.array("edge_node1").astype(np.int64)
    assert rp[-1] == z["nnz"] == P + 2 * E
    for i in range(P):
        cols = ci[rp[i]:rp[i + 1]]
        assert np.all(np.diff(cols) > 0)
        assert ci[dp[i]] == i
    ep = ep.reshape(E, 2)
    assert np.array_equal(ci[ep[:, 0]], e1) and np.array_equal(ci[ep[:, 1]], e0)
    for k in range(E):
        assert rp[e0[k]] <= ep[k, 0] < rp[e0[k] + 1]
        assert rp[e1[k]] <= ep[k, 1] < rp[e1[k] + 1]
    # point -> edge adjacency in ascending edge order (the fixed summation order of the gather assembly)
    pp, pe, pn = (m.array(k).astype(np.int64) for k in ("point_ptr", "point_edge", "point_nbr"))
    for i in range(P):
        edges = pe[pp[i]:pp[i + 1]]
        nbrs = pn[pp[i]:pp[i + 1]]
        for e, j in zip(edges, nbrs):
            assert {int(e0[e]), int(e1[e])} == {i, int(j)}
    # edge colouring: no two edges of one colour share a point
    cp, ce = m.array("color_ptr").astype(np.int64), m.array("color_edges").astype(np.int64)
    assert sorted(ce.tolist()) == list(range(E))
    for c in range(len(cp) - 1):
        es = ce[cp[c]:cp[c + 1]]
        pts = np.concatenate([e0[es], e1[es]])
        assert len(set(pts.tolist())) == pts.size
    m.close()


@pytest.mark.parametrize("nranks", [2, 3])
def test_slab_decomposition_matches_the_undivided_mesh(nranks):
    n, length = [6, 5, 9], [1.0, 0.7, 1.3]
    tags = ("x_minus", "x_plus", "y_minus", "y_plus", "z_minus", "z_plus")

    def build(rank, nr):
        m = Mesh.box(n, length, rank=rank, nranks=nr)
        for t in tags:
            m.set_marker(t, capi.EULER_WALL if t.startswith("y") else capi.FAR_FIELD)
        m.preprocess()
        return m

    whole = build(0, 1)
    wv = whole.array("volume")
    wg = whole.array("global_index").astype(np.int64)
    vol_by_gid = dict(zip(wg.tolist(), wv.tolist()))
    parts = [build(r, nranks) for r in range(nranks)]
    owned = []
    for r, m in enumerate(parts):
        z = m.sizes()
        gid = m.array("global_index").astype(np.int64)
        vol = m.array("volume")
        Pd = z["nPointDomain"]
        owned += gid[:Pd].tolist()
        # owned control volumes are complete (halo layer closes them) and equal the undivided ones
        for g, v in zip(gid[:Pd], vol[:Pd]):
            assert abs(v - vol_by_gid[int(g)]) <= 1e-14 * abs(v)
    assert sorted(owned) == sorted(wg.tolist())
    # send list of r towards q names the same global points, in the same order, as q's receive list from r
    for r, m in enumerate(parts):
        ranks = m.array("p2p_ranks")
        sp, snd = m.array("p2p_send_ptr"), m.array("p2p_send").astype(np.int64)
        gid = m.array("global_index").astype(np.int64)
        for k, q in enumerate(ranks):
            mq = parts[int(q)]
            rq = mq.array("p2p_ranks").tolist().index(r)
            rp_, rcv = mq.array("p2p_recv_ptr"), mq.array("p2p_recv").astype(np.int64)
            gq = mq.array("global_index").astype(np.int64)
            mine = gid[snd[sp[k]:sp[k + 1]]]
            theirs = gq[rcv[rp_[rq]:rp_[rq + 1]]]
            assert np.array_equal(mine, theirs)
            assert np.all(snd[sp[k]:sp[k + 1]] < m.sizes()["nPointDomain"])
            assert np.all(rcv[rp_[rq]:rp_[rq + 1]] >= mq.sizes()["nPointDomain"])
    for m in parts + [whole]:
        m.close()


def _partition_checks(make_whole, nranks):
    """Unstructured decomposition (su2b200_mesh_partition): owned sets partition the points, owned control volumes are
    complete, send and receive lists of a pair name the same global points in the same order, weights are balanced."""
    whole = make_whole()
    whole.preprocess(rcm=True)
    wv = dict(zip(whole.array("global_index").astype(np.int64).tolist(), whole.array("volume").tolist()))
    deg = dict(zip(whole.array("global_index").astype(np.int64).tolist(), np.diff(whole.array("point_ptr").astype(np.int64)).tolist()))
    wd = dict(zip(whole.array("global_index").astype(np.int64).tolist(), whole.array("wall_distance").tolist()))
    nE_whole = whole.sizes()["nEdge"]
    parts = []
    for r in range(nranks):
        w = make_whole()
        p = w.partition(nranks, r)
        w.close()
        p.preprocess(rcm=True)
        parts.append(p)
    owned, edges_owned = [], 0
    for r, m in enumerate(parts):
        z = m.sizes()
        gid = m.array("global_index").astype(np.int64)
        vol = m.array("volume")
        Pd = z["nPointDomain"]
        owned += gid[:Pd].tolist()
        for g, v in zip(gid[:Pd], vol[:Pd]):
            assert abs(v - wv[int(g)]) <= 1e-13 * abs(v), (r, g)
        # the nearest wall may be on another rank: distances come from the global wall surface
        assert np.array_equal(m.array("wall_distance"), np.array([wd[int(g)] for g in gid]))
        e0, e1 = m.array("edge_node0").astype(np.int64), m.array("edge_node1").astype(np.int64)
        assert not np.any((e0 >= Pd) & (e1 >= Pd))  # no halo-halo edges
        edges_owned += int(np.sum((e0 < Pd) & (e1 < Pd))) + 0.5 * int(np.sum((e0 < Pd) ^ (e1 < Pd)))
    assert sorted(owned) == sorted(wv.keys())
    assert edges_owned == nE_whole  # every edge once: interior edges on one rank, cut edges half on each side
    sizes = [m.sizes()["nPointDomain"] for m in parts]
    # balanced in the reference's vertex weight (wp + we * degree = the number of neighbours), to within a few points
    load = [sum(deg[int(g)] for g in m.array("global_index").astype(np.int64)[:m.sizes()["nPointDomain"]]) for m in parts]
    assert max(load) - min(load) <= 0.05 * sum(load) / nranks + 3 * max(deg.values()), load
    for r, m in enumerate(parts):
        ranks = m.array("p2p_ranks")
        sp, snd = m.array("p2p_send_ptr"), m.array("p2p_send").astype(np.int64)
        gid = m.array("global_index").astype(np.int64)
        assert len(ranks) > 0
        for k, q in enumerate(ranks):
            mq = parts[int(q)]
            rq = mq.array("p2p_ranks").tolist().index(r)
            rp_, rcv = mq.array("p2p_recv_ptr"), mq.array("p2p_recv").astype(np.int64)
            gq = mq.array("global_index").astype(np.int64)
            mine, theirs = gid[snd[sp[k]:sp[k + 1]]], gq[rcv[rp_[rq]:rp_[rq + 1]]]
            assert np.array_equal(mine, theirs) and len(mine) > 0
            assert np.all(snd[sp[k]:sp[k + 1]] < m.sizes()["nPointDomain"])
            assert np.all(rcv[rp_[rq]:rp_[rq + 1]] >= mq.sizes()["nPointDomain"])
    # the whole mesh renumbered colour by colour: group sizes = the ranks' owned counts
    w = make_whole()
    ren, pp = w.partition(nranks, -1)
    w.close()
    assert [pp[k + 1] - pp[k] for k in range(nranks)] == sizes
    ren.close()
    for m in parts + [whole]:
        m.close()


@pytest.mark.parametrize("nranks", [2, 3, 4])
def test_unstructured_partition_of_a_tetrahedral_box(nranks, tmp_path):
    path = tmp_path / "tet.su2"
    a = Mesh.box([7, 6, 5], [2.0, 1.0, 1.0], tets=True)
    a.write_su2(path)
    a.close()

    def make():
        m = Mesh.read_su2(path)
        for t in ("x_minus", "x_plus", "y_minus", "y_plus", "z_minus", "z_plus"):
            m.set_marker(t, capi.HEAT_FLUX if t == "y_minus" else (capi.EULER_WALL if t == "y_plus" else capi.FAR_FIELD))
        return m

    _partition_checks(make, nranks)


@pytest.mark.parametrize("nranks", [2, 4])
def test_unstructured_partition_of_the_quickstart_mesh(nranks):
    path = "/root/reference/QuickStart/mesh_NACA0012_inv.su2"
    if not os.path.exists(path):
        pytest.skip("QuickStart mesh not on this machine")

    def make():
        m = Mesh.read_su2(path)
        m.set_marker("airfoil", capi.EULER_WALL, 1)
        m.set_marker("farfield", capi.FAR_FIELD)
        return m

    _partition_checks(make, nranks)


def test_shim_patch_anchors_match_the_reference(tmp_path):
    """integration/apply_patch.py on a scratch copy of the five reference files it edits: every anchor must be found
    exactly once (the script stops otherwise), the edits must sit behind HAVE_SU2B200, and a second run must be a no-op.
    Needs the reference tree (this container); the GPU box only uses the prebuilt integration/_ref/SU2_CFD_b200."""
    ref = os.environ.get("SU2_REFERENCE", "/root/reference")
    files = ["Common/include/linear_algebra/CSysMatrix.hpp", "Common/include/linear_algebra/CMatrixVectorProduct.hpp",
             "SU2_CFD/include/solvers/CSolver.hpp", "SU2_CFD/include/solvers/CFVMFlowSolverBase.inl",
             "SU2_CFD/src/integration/CMultiGridIntegration.cpp"]
    if not all(os.path.exists(os.path.join(ref, f)) for f in files):
        pytest.skip("no reference tree at %s" % ref)
    import importlib.util
    import shutil
    for f in files:
        dst = tmp_path / f
        dst.parent.mkdir(parents=True, exist_ok=True)
        shutil.copy(os.path.join(ref, f), dst)
        os.chmod(dst, 0o644)
    spec = importlib.util.spec_from_file_location("apply_patch", os.path.join(ROOT, "integration", "apply_patch.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.main(str(tmp_path), ROOT)
    patched = {f: (tmp_path / f).read_text() for f in files}
    for f, text in patched.items():
        assert "HAVE_SU2B200" in text, f
        assert text.count("#ifdef HAVE_SU2B200") + text.count("#if defined(HAVE_SU2B200)") >= 1, f
        orig = open(os.path.join(ref, f)).read()
        assert len(text) > len(orig)  # additions only; every one of them sits behind the macro
    assert (tmp_path / "Common/include/linear_algebra/CB200FlowBackend.hpp").exists()
    assert (tmp_path / "Common/include/linear_algebra/su2b200.h").exists()
    mod.main(str(tmp_path), ROOT)  # idempotent
    for f in files:
        assert (tmp_path / f).read_text() == patched[f], f
