"""Host-side mirror of the CGeometry preprocessing the path consumes (su2b200_mesh_* of include/su2b200.h).

The work is done by the C++ host code in csrc/mesh_host.cu; this class only marshals arguments.  It runs without a
GPU (mesh generation, dual grid, sparse pattern, halo tables are host work); `upload()` needs one.
"""
import ctypes as C
import numpy as np
from . import capi
from .capi import check

_DT = {0: np.float64, 1: np.uint64, 2: np.uint16, 3: np.uint8, 4: np.int32}


class Mesh:
    def __init__(self, handle):
        self.lib = capi.load()
        self.h = handle

    @classmethod
    def box(cls, n, length, offset=(0.0, 0.0, 0.0), rank=0, nranks=1, tets=False):
        """MESH_FORMAT= BOX (len(n) == 3) / RECTANGLE (len(n) == 2), local part of `rank` of a slab decomposition."""
        lib = capi.load()
        nd = len(n)
        nn = (C.c_ulong * 3)(*(list(n) + [1] * (3 - nd)))
        ll = (C.c_double * 3)(*(list(length) + [0.0] * (3 - len(length))))
        oo = (C.c_double * 3)(*(list(offset) + [0.0] * (3 - len(offset))))
        h = C.c_void_p()
        check(lib.su2b200_mesh_create_box(nd, nn, ll, oo, int(rank), int(nranks), int(bool(tets)), C.byref(h)))
        return cls(h)

    @classmethod
    def read_su2(cls, path):
        lib = capi.load()
        h = C.c_void_p()
        check(lib.su2b200_mesh_read_su2(str(path).encode(), C.byref(h)))
        return cls(h)

    def partition(self, nranks, rank):
        """Local mesh of `rank` out of `nranks` (su2b200_mesh_partition); rank = -1: the whole mesh renumbered colour by
        colour, returned together with the first point of every colour."""
        h = C.c_void_p()
        pp = (C.c_ulong * (int(nranks) + 1))()
        check(self.lib.su2b200_mesh_partition(self.h, int(nranks), int(rank), C.byref(h), pp))
        m = Mesh(h)
        return (m, [int(x) for x in pp]) if rank < 0 else m

    def write_su2(self, path):
        """ASCII .su2 file of the mesh in its current numbering (su2b200_mesh_write_su2)."""
        check(self.lib.su2b200_mesh_write_su2(self.h, str(path).encode()))

    def stretch(self, axis, first_cell, ratio):
        """Geometric wall-normal stretching of a box mesh before preprocess() (su2b200_mesh_stretch)."""
        check(self.lib.su2b200_mesh_stretch(self.h, int(axis), C.c_double(first_cell), C.c_double(ratio)))

    def close(self):
        if self.h:
            self.lib.su2b200_mesh_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_marker(self, tag, kind, monitoring=0, heat_flux=0.0):
        check(self.lib.su2b200_mesh_set_marker(self.h, tag.encode(), int(kind), int(monitoring), C.c_double(heat_flux)))

    def preprocess(self, rcm=False):
        check(self.lib.su2b200_mesh_preprocess(self.h, int(bool(rcm))))

    def sizes(self):
        nd, nm = C.c_int(), C.c_int()
        P, Pd, E, Ne, nnz = (C.c_ulong() for _ in range(5))
        check(self.lib.su2b200_mesh_sizes(self.h, C.byref(nd), C.byref(P), C.byref(Pd), C.byref(E), C.byref(Ne),
                                          C.byref(nnz), C.byref(nm)))
        return dict(nDim=nd.value, nPoint=P.value, nPointDomain=Pd.value, nEdge=E.value, nElem=Ne.value,
                    nnz=nnz.value, nMarker=nm.value)

    def array(self, name):
        """Copy of a named mesh array (see su2b200_mesh_array)."""
        data, cnt, dt = C.c_void_p(), C.c_ulong(), C.c_int()
        check(self.lib.su2b200_mesh_array(self.h, name.encode(), C.byref(data), C.byref(cnt), C.byref(dt)))
        dtype = np.dtype(_DT[dt.value])
        if cnt.value == 0:
            return np.empty(0, dtype=dtype)
        buf = (C.c_char * (cnt.value * dtype.itemsize)).from_address(data.value)
        return np.frombuffer(buf, dtype=dtype).copy()

    def marker_info(self, i):
        tag = C.create_string_buffer(64)
        kind, mon = C.c_int(), C.c_int()
        nv, ne = C.c_ulong(), C.c_ulong()
        check(self.lib.su2b200_mesh_marker_info(self.h, int(i), tag, C.byref(kind), C.byref(mon), C.byref(nv), C.byref(ne)))
        return dict(tag=tag.value.decode(), kind=kind.value, monitoring=mon.value, nVertex=nv.value, nElemBound=ne.value)

    def upload(self, solver):
        """Hand geometry, markers and halo tables to a FlowSolver (su2b200_mesh_upload)."""
        check(self.lib.su2b200_mesh_upload(solver.ctx, self.h))
        z = self.sizes()
        solver.nDim, solver.nPoint, solver.nPointDomain, solver.nEdge, solver.nnz = (
            z["nDim"], z["nPoint"], z["nPointDomain"], z["nEdge"], z["nnz"])
        solver.markers = []
        for i in range(z["nMarker"]):
            mi = self.marker_info(i)
            solver.markers.append(dict(kind=mi["kind"], tag=mi["tag"], nVertex=mi["nVertex"], monitoring=mi["monitoring"]))
