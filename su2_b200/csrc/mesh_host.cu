// a1-a4. Host-side geometry preprocessing: the part of CGeometry / CPhysicalGeometry the hot path consumes,
// producing exactly the arrays su2b200_geometry_upload / su2b200_marker_upload / su2b200_p2p_upload take.
//   mesh sources : CBoxMeshReaderFVM (Common/src/geometry/meshreader/CBoxMeshReaderFVM.cpp:31-330),
//                  CRectangularMeshReaderFVM, CSU2ASCIIMeshReaderBase (.su2 ASCII: NDIME/NELEM/NPOIN/NMARK)
//   ordering     : CPhysicalGeometry::SetRCM_Ordering (CPhysicalGeometry.cpp:4478)
//   dual grid    : SetEdges (CGeometry.cpp:1263), SetControlVolume (CPhysicalGeometry.cpp:7059-7209),
//                  SetBoundControlVolume (:7212-7296), FindNormal_Neighbor (:8036-8078),
//                  ComputeModifiedSymmetryNormals (CGeometry.cpp:2469-2585, straight markers)
//   pattern      : GetSparsePattern (CGeometry.cpp:3702), GetEdgeToSparsePatternMap (:3720)
//   halo tables  : PreprocessP2PComms (CGeometry.cpp:139-332) for the slab decomposition of BOX/RECTANGLE meshes
// Everything here is plain host C++ (no device code); it is in a .cu file only to share the build rule.
#include "internal.hpp"
#include <algorithm>
#include <array>
#include <cmath>
#include <fstream>
#include <map>
#include <numeric>
#include <queue>
#include <sstream>

using namespace su2b200;
typedef unsigned long ulong_t;

namespace {

// VTK element types used by the .su2 format
enum { VTK_LINE = 3, VTK_TRI = 5, VTK_QUAD = 9, VTK_TET = 10, VTK_HEX = 12, VTK_PRISM = 13, VTK_PYRAMID = 14 };

struct ElemTopo {
  int nNodes, nFaces;
  int nNodesFace[6];
  int faces[6][4];
};
// face node lists (outward-oriented, VTK node numbering), same convention as the reference's primal-grid classes
static const ElemTopo& topo(int vtk) {
  static const ElemTopo tri = {3, 3, {2, 2, 2}, {{0, 1}, {1, 2}, {2, 0}}};
  static const ElemTopo quad = {4, 4, {2, 2, 2, 2}, {{0, 1}, {1, 2}, {2, 3}, {3, 0}}};
  static const ElemTopo tet = {4, 4, {3, 3, 3, 3}, {{0, 2, 1}, {0, 1, 3}, {0, 3, 2}, {1, 2, 3}}};
  static const ElemTopo hex = {8, 6, {4, 4, 4, 4, 4, 4},
                               {{0, 1, 5, 4}, {1, 2, 6, 5}, {2, 3, 7, 6}, {3, 0, 4, 7}, {0, 3, 2, 1}, {4, 5, 6, 7}}};
  static const ElemTopo prism = {6, 5, {4, 4, 4, 3, 3}, {{3, 4, 1, 0}, {5, 2, 1, 4}, {2, 5, 3, 0}, {0, 1, 2}, {5, 4, 3}}};
  static const ElemTopo pyra = {5, 5, {4, 3, 3, 3, 3}, {{0, 3, 2, 1}, {4, 3, 0}, {4, 0, 1}, {2, 4, 1}, {3, 4, 2}}};
  switch (vtk) {
    case VTK_TRI: return tri;
    case VTK_QUAD: return quad;
    case VTK_TET: return tet;
    case VTK_HEX: return hex;
    case VTK_PRISM: return prism;
    case VTK_PYRAMID: return pyra;
    default: throw Error("unsupported volume element type " + std::to_string(vtk));
  }
}
static int bound_nodes(int vtk) {
  switch (vtk) {
    case VTK_LINE: return 2;
    case VTK_TRI: return 3;
    case VTK_QUAD: return 4;
    default: throw Error("unsupported boundary element type " + std::to_string(vtk));
  }
}

struct HostMarker {
  std::string tag;
  int kind = 0, monitoring = 0;
  double heat_flux = 0.0;
  std::vector<int> etype;
  std::vector<long> eptr{0};
  std::vector<int> enode;
  // after preprocessing
  std::vector<ulong_t> vnode, vnbr;
  std::vector<double> vnormal, vunit;
};

}  // namespace

struct su2b200_mesh {
  int nDim = 0;
  long P = 0, Pd = 0;
  std::vector<double> coord;
  std::vector<ulong_t> global_index;
  std::vector<int> etype;
  std::vector<long> eptr{0};
  std::vector<int> enode;
  std::vector<HostMarker> markers;
  // structured-box metadata (wall distance shortcut, slab decomposition)
  bool is_box = false;
  double box_len[3] = {0, 0, 0}, box_off[3] = {0, 0, 0};
  int rank = 0, nranks = 1;
  std::vector<int> p2p_ranks, p2p_send_ptr{0}, p2p_recv_ptr{0};
  std::vector<ulong_t> p2p_send, p2p_recv;
  // dual grid
  bool preprocessed = false;
  long E = 0, nnz = 0;
  std::vector<ulong_t> edge_n0, edge_n1, point_ptr, point_nbr, point_edge, row_ptr, col_ind, dia_ptr, edge_ptr;
  std::vector<double> edge_normal, volume, wall_distance;
  std::vector<unsigned short> nNeighbor;
  std::vector<unsigned char> physical_boundary;
  std::vector<ulong_t> color_ptr, color_edges;
  std::vector<ulong_t> perm;  // new index of the point that had index i before reordering
  std::vector<double> wall_distance_global;  // per ORIGINAL (global) point, set by the partitioner: the wall surface is global
  long nElem() const { return (long)etype.size(); }
};

namespace {

inline const double* X(const su2b200_mesh& m, long i) { return &m.coord[(size_t)i * m.nDim]; }

void add_elem(su2b200_mesh& m, int type, const int* nodes, int n) {
  m.etype.push_back(type);
  for (int k = 0; k < n; ++k) m.enode.push_back(nodes[k]);
  m.eptr.push_back((long)m.enode.size());
}
void add_belem(HostMarker& mk, int type, const int* nodes, int n) {
  mk.etype.push_back(type);
  for (int k = 0; k < n; ++k) mk.enode.push_back(nodes[k]);
  mk.eptr.push_back((long)mk.enode.size());
}

// ---- structured BOX / RECTANGLE mesh, slab-decomposed along the last axis -------------------------------------
// Global point numbering is the reference's: index = k*m*n + j*n + i (CBoxMeshReaderFVM.cpp:84-106), hexahedra
// with the connectivity of :131-138, markers x_minus, x_plus, y_minus, y_plus[, z_minus, z_plus] (:165-292).
// Rank r owns the planes [s0, s1) of the slab axis; it also stores the neighbouring planes s0-1 and s1 as halo
// points (owned first, then the halo of the lower rank, then of the upper rank) and every element that touches an
// owned point, so that owned control volumes are complete -- the same closure the reference's partitioner builds.
su2b200_mesh* make_box(int nDim, const ulong_t n[3], const double len[3], const double off[3], int rank, int nranks,
                       int tets) {
  SU2B_REQUIRE(nDim == 2 || nDim == 3, "nDim must be 2 or 3");
  SU2B_REQUIRE(nranks >= 1 && rank >= 0 && rank < nranks, "invalid rank");
  const long nx = (long)n[0], ny = (long)n[1], nz = nDim == 3 ? (long)n[2] : 1;
  SU2B_REQUIRE(nx >= 2 && ny >= 2 && (nDim == 2 || nz >= 2), "box needs at least 2 points per direction");
  const long ns = nDim == 3 ? nz : ny;  // planes along the slab axis
  SU2B_REQUIRE(ns >= 2L * nranks, "too many ranks for this box (need >= 2 planes per rank)");
  SU2B_REQUIRE((double)nx * ny * nz / nranks < 2.0e9, "more than 2^31 points per rank");
  auto* m = new su2b200_mesh();
  m->nDim = nDim;
  m->is_box = true;
  m->rank = rank;
  m->nranks = nranks;
  for (int d = 0; d < 3; ++d) { m->box_len[d] = len[d]; m->box_off[d] = off[d]; }
  const long s0 = ns * rank / nranks, s1 = ns * (rank + 1) / nranks;
  const bool lo = s0 > 0, hi = s1 < ns;
  const long plane = nDim == 3 ? nx * ny : nx;  // points per slab plane
  const long nOwnedPlanes = s1 - s0;
  m->Pd = plane * nOwnedPlanes;
  m->P = m->Pd + (lo ? plane : 0) + (hi ? plane : 0);
  // local index of global (i, j, k)
  auto local = [&](long i, long j, long k) -> int {
    const long s = nDim == 3 ? k : j;
    const long inplane = nDim == 3 ? j * nx + i : i;
    if (s >= s0 && s < s1) return (int)((s - s0) * plane + inplane);
    if (s == s0 - 1) return (int)(m->Pd + inplane);
    return (int)(m->Pd + (lo ? plane : 0) + inplane);  // s == s1
  };
  m->coord.resize((size_t)m->P * nDim);
  m->global_index.resize(m->P);
  const long sa = lo ? s0 - 1 : s0, sb = hi ? s1 + 1 : s1;
  const double dx = len[0] / (double)(nx - 1), dy = len[1] / (double)(ny - 1), dz = nDim == 3 ? len[2] / (double)(nz - 1) : 0.0;
  (void)dx; (void)dy; (void)dz;
  for (long s = sa; s < sb; ++s) {
    for (long q = 0; q < plane; ++q) {
      const long i = q % nx, j = nDim == 3 ? q / nx : s, k = nDim == 3 ? s : 0;
      const int l = local(i, j, k);
      m->coord[(size_t)l * nDim + 0] = len[0] * ((double)i) / ((double)(nx - 1)) + off[0];
      m->coord[(size_t)l * nDim + 1] = len[1] * ((double)j) / ((double)(ny - 1)) + off[1];
      if (nDim == 3) m->coord[(size_t)l * nDim + 2] = len[2] * ((double)k) / ((double)(nz - 1)) + off[2];
      m->global_index[l] = (ulong_t)(k * ny * nx + j * nx + i);
    }
  }
  // volume elements: cells whose slab index lies in [sa, sb-1)
  m->markers.resize(nDim == 3 ? 6 : 4);
  const char* tags3[6] = {"x_minus", "x_plus", "y_minus", "y_plus", "z_minus", "z_plus"};
  for (size_t t = 0; t < m->markers.size(); ++t) m->markers[t].tag = tags3[t];
  // Kuhn split of a hexahedron into 6 tetrahedra around the diagonal 0-6 (translation invariant, so faces match)
  static const int kuhn[6][4] = {{0, 1, 2, 6}, {0, 2, 3, 6}, {0, 3, 7, 6}, {0, 7, 4, 6}, {0, 4, 5, 6}, {0, 5, 1, 6}};
  auto add_bquad = [&](HostMarker& mk, const int q[4], bool diag02) {
    if (!tets) { add_belem(mk, VTK_QUAD, q, 4); return; }
    if (diag02) { const int a[3] = {q[0], q[1], q[2]}, b[3] = {q[0], q[2], q[3]}; add_belem(mk, VTK_TRI, a, 3); add_belem(mk, VTK_TRI, b, 3); }
    else { const int a[3] = {q[0], q[1], q[3]}, b[3] = {q[1], q[2], q[3]}; add_belem(mk, VTK_TRI, a, 3); add_belem(mk, VTK_TRI, b, 3); }
  };
  if (nDim == 3) {
    for (long k = sa; k < sb - 1; ++k)
      for (long j = 0; j < ny - 1; ++j)
        for (long i = 0; i < nx - 1; ++i) {
          const int c[8] = {local(i, j, k), local(i + 1, j, k), local(i + 1, j + 1, k), local(i, j + 1, k),
                            local(i, j, k + 1), local(i + 1, j, k + 1), local(i + 1, j + 1, k + 1), local(i, j + 1, k + 1)};
          if (!tets) add_elem(*m, VTK_HEX, c, 8);
          else
            for (int t = 0; t < 6; ++t) { const int tt[4] = {c[kuhn[t][0]], c[kuhn[t][1]], c[kuhn[t][2]], c[kuhn[t][3]]}; add_elem(*m, VTK_TET, tt, 4); }
          // boundary quads with the reference's node order; the Kuhn diagonal of a box face joins its lowest and
          // highest corner in (i,j,k)
          if (i == 0) { const int q[4] = {c[0], c[4], c[7], c[3]}; add_bquad(m->markers[0], q, true); }
          if (i == nx - 2) { const int q[4] = {c[1], c[2], c[6], c[5]}; add_bquad(m->markers[1], q, true); }
          if (j == 0) { const int q[4] = {c[0], c[1], c[5], c[4]}; add_bquad(m->markers[2], q, true); }
          if (j == ny - 2) { const int q[4] = {c[3], c[2], c[6], c[7]}; add_bquad(m->markers[3], q, true); }
          if (k == 0) { const int q[4] = {c[0], c[1], c[2], c[3]}; add_bquad(m->markers[4], q, true); }
          if (k == nz - 2) { const int q[4] = {c[4], c[5], c[6], c[7]}; add_bquad(m->markers[5], q, true); }
        }
  } else {
    for (long j = sa; j < sb - 1; ++j)
      for (long i = 0; i < nx - 1; ++i) {
        const int c[4] = {local(i, j, 0), local(i + 1, j, 0), local(i + 1, j + 1, 0), local(i, j + 1, 0)};
        if (!tets) add_elem(*m, VTK_QUAD, c, 4);
        else { const int a[3] = {c[0], c[1], c[2]}, b[3] = {c[0], c[2], c[3]}; add_elem(*m, VTK_TRI, a, 3); add_elem(*m, VTK_TRI, b, 3); }
        if (i == 0) { const int q[2] = {c[3], c[0]}; add_belem(m->markers[0], VTK_LINE, q, 2); }
        if (i == nx - 2) { const int q[2] = {c[1], c[2]}; add_belem(m->markers[1], VTK_LINE, q, 2); }
        if (j == 0) { const int q[2] = {c[0], c[1]}; add_belem(m->markers[2], VTK_LINE, q, 2); }
        if (j == ny - 2) { const int q[2] = {c[2], c[3]}; add_belem(m->markers[3], VTK_LINE, q, 2); }
      }
  }
  // halo tables: planes s0 <-> rank-1, s1-1 <-> rank+1, both sides enumerate the plane in the same order
  if (lo) {
    m->p2p_ranks.push_back(rank - 1);
    for (long q = 0; q < plane; ++q) { m->p2p_send.push_back((ulong_t)q); m->p2p_recv.push_back((ulong_t)(m->Pd + q)); }
    m->p2p_send_ptr.push_back((int)m->p2p_send.size());
    m->p2p_recv_ptr.push_back((int)m->p2p_recv.size());
  }
  if (hi) {
    m->p2p_ranks.push_back(rank + 1);
    for (long q = 0; q < plane; ++q) {
      m->p2p_send.push_back((ulong_t)((nOwnedPlanes - 1) * plane + q));
      m->p2p_recv.push_back((ulong_t)(m->Pd + (lo ? plane : 0) + q));
    }
    m->p2p_send_ptr.push_back((int)m->p2p_send.size());
    m->p2p_recv_ptr.push_back((int)m->p2p_recv.size());
  }
  return m;
}

// ---- .su2 ASCII reader (serial meshes) ---------------------------------------------------------------------------
su2b200_mesh* read_su2(const char* path) {
  std::ifstream f(path);
  if (!f) throw Error(std::string("cannot open mesh file ") + path);
  auto* m = new su2b200_mesh();
  std::string line;
  auto keyval = [](const std::string& l, const char* key, long* v) {
    const size_t p = l.find(key);
    if (p == std::string::npos) return false;
    const size_t e = l.find('=', p);
    if (e == std::string::npos) return false;
    *v = std::atol(l.c_str() + e + 1);
    return true;
  };
  auto next_data_line = [&](std::string& l) {
    while (std::getline(f, l)) {
      const size_t p = l.find_first_not_of(" \t\r");
      if (p == std::string::npos || l[p] == '%') continue;
      return true;
    }
    return false;
  };
  long v = 0;
  while (next_data_line(line)) {
    if (keyval(line, "NDIME", &v)) {
      m->nDim = (int)v;
    } else if (keyval(line, "NELEM", &v)) {
      for (long e = 0; e < v; ++e) {
        if (!next_data_line(line)) throw Error("unexpected end of file in NELEM block");
        std::istringstream is(line);
        int type;
        is >> type;
        const int nn = topo(type).nNodes;
        int nodes[8];
        for (int k = 0; k < nn; ++k) is >> nodes[k];
        add_elem(*m, type, nodes, nn);
      }
    } else if (keyval(line, "NPOIN", &v)) {
      SU2B_REQUIRE(m->nDim == 2 || m->nDim == 3, "NDIME must precede NPOIN");
      m->P = m->Pd = v;
      m->coord.resize((size_t)v * m->nDim);
      m->global_index.resize(v);
      for (long i = 0; i < v; ++i) {
        if (!next_data_line(line)) throw Error("unexpected end of file in NPOIN block");
        std::istringstream is(line);
        for (int d = 0; d < m->nDim; ++d) is >> m->coord[(size_t)i * m->nDim + d];
        m->global_index[i] = (ulong_t)i;
      }
    } else if (keyval(line, "NMARK", &v)) {
      const long nm = v;
      for (long k = 0; k < nm; ++k) {
        HostMarker mk;
        if (!next_data_line(line)) throw Error("unexpected end of file in NMARK block");
        const size_t e = line.find('=');
        SU2B_REQUIRE(line.find("MARKER_TAG") != std::string::npos && e != std::string::npos, "expected MARKER_TAG");
        std::string tag = line.substr(e + 1);
        tag.erase(0, tag.find_first_not_of(" \t"));
        tag.erase(tag.find_last_not_of(" \t\r\n") + 1);
        mk.tag = tag;
        long ne = 0;
        if (!next_data_line(line) || !keyval(line, "MARKER_ELEMS", &ne)) throw Error("expected MARKER_ELEMS");
        for (long b = 0; b < ne; ++b) {
          if (!next_data_line(line)) throw Error("unexpected end of file in marker elements");
          std::istringstream is(line);
          int type;
          is >> type;
          const int nn = bound_nodes(type);
          int nodes[4];
          for (int q = 0; q < nn; ++q) is >> nodes[q];
          add_belem(mk, type, nodes, nn);
        }
        m->markers.push_back(std::move(mk));
      }
    }
  }
  SU2B_REQUIRE(m->P > 0 && m->nElem() > 0, "mesh file holds no points or no elements");
  return m;
}

// ---- geometry helpers ---------------------------------------------------------------------------------------------
inline void cross(const double* a, const double* b, double* c) {
  c[0] = a[1] * b[2] - a[2] * b[1];
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
}

// positive for a right-handed (outward faces) element; used to flip badly ordered elements
// (CPhysicalGeometry::Check_IntElem_Orientation)
double elem_signed_volume(const su2b200_mesh& m, int type, const int* nd) {
  const int D = m.nDim;
  if (D == 2) {
    const int nn = topo(type).nNodes;
    double a = 0.0;
    for (int k = 0; k < nn; ++k) {
      const double* p = X(m, nd[k]);
      const double* q = X(m, nd[(k + 1) % nn]);
      a += p[0] * q[1] - q[0] * p[1];
    }
    return 0.5 * a;
  }
  // 3D: divergence theorem over the triangulated faces
  const ElemTopo& t = topo(type);
  double cg[3] = {0, 0, 0};
  for (int k = 0; k < t.nNodes; ++k)
    for (int d = 0; d < 3; ++d) cg[d] += X(m, nd[k])[d] / t.nNodes;
  double vol = 0.0;
  for (int f = 0; f < t.nFaces; ++f) {
    const int nf = t.nNodesFace[f];
    double fc[3] = {0, 0, 0};
    for (int k = 0; k < nf; ++k)
      for (int d = 0; d < 3; ++d) fc[d] += X(m, nd[t.faces[f][k]])[d] / nf;
    for (int k = 0; k < nf; ++k) {
      const double* p = X(m, nd[t.faces[f][k]]);
      const double* q = X(m, nd[t.faces[f][(k + 1) % nf]]);
      double a[3], b[3], c[3], r[3];
      for (int d = 0; d < 3; ++d) { a[d] = p[d] - fc[d]; b[d] = q[d] - fc[d]; r[d] = fc[d] - cg[d]; }
      cross(a, b, c);
      vol += (c[0] * r[0] + c[1] * r[1] + c[2] * r[2]) / 6.0;
    }
  }
  return vol;
}

void fix_orientation(su2b200_mesh& m) {
  for (long e = 0; e < m.nElem(); ++e) {
    int* nd = &m.enode[m.eptr[e]];
    if (elem_signed_volume(m, m.etype[e], nd) >= 0.0) continue;
    switch (m.etype[e]) {
      case VTK_TRI: std::swap(nd[1], nd[2]); break;
      case VTK_QUAD: std::swap(nd[1], nd[3]); break;
      case VTK_TET: std::swap(nd[0], nd[1]); break;
      case VTK_HEX: std::swap(nd[1], nd[3]); std::swap(nd[5], nd[7]); break;
      case VTK_PRISM: std::swap(nd[0], nd[1]); std::swap(nd[3], nd[4]); break;
      case VTK_PYRAMID: std::swap(nd[1], nd[3]); break;
    }
  }
}

// unique undirected element edges -> neighbour CSR (ascending neighbours), dropping halo-halo pairs
void build_neighbours(const su2b200_mesh& m, std::vector<long>& ptr, std::vector<int>& nbr) {
  const long P = m.P;
  std::vector<long> cnt(P + 1, 0);
  auto for_each_edge = [&](auto&& fn) {
    for (long e = 0; e < m.nElem(); ++e) {
      const ElemTopo& t = topo(m.etype[e]);
      const int* nd = &m.enode[m.eptr[e]];
      if (m.nDim == 2) {
        for (int f = 0; f < t.nFaces; ++f) fn(nd[t.faces[f][0]], nd[t.faces[f][1]]);
      } else {
        for (int f = 0; f < t.nFaces; ++f)
          for (int k = 0; k < t.nNodesFace[f]; ++k) fn(nd[t.faces[f][k]], nd[t.faces[f][(k + 1) % t.nNodesFace[f]]]);
      }
    }
  };
  for_each_edge([&](int a, int b) {
    if (a >= m.Pd && b >= m.Pd) return;
    cnt[a + 1]++; cnt[b + 1]++;
  });
  for (long i = 0; i < P; ++i) cnt[i + 1] += cnt[i];
  std::vector<int> raw(cnt[P]);
  std::vector<long> pos(cnt.begin(), cnt.end() - 1);
  for_each_edge([&](int a, int b) {
    if (a >= m.Pd && b >= m.Pd) return;
    raw[pos[a]++] = b; raw[pos[b]++] = a;
  });
  ptr.assign(P + 1, 0);
  nbr.clear();
  nbr.reserve(raw.size() / 2);
  for (long i = 0; i < P; ++i) {
    std::sort(raw.begin() + cnt[i], raw.begin() + cnt[i + 1]);
    auto e = std::unique(raw.begin() + cnt[i], raw.begin() + cnt[i + 1]);
    nbr.insert(nbr.end(), raw.begin() + cnt[i], e);
    ptr[i + 1] = (long)nbr.size();
  }
}

// Local node adjacency of an element (the nodes that share an element edge with local node k), listed in the order of
// the reference's primal-grid classes (C<Elem>Connectivity::Neighbor_Nodes, Common/include/geometry/primal_grid/*.hpp):
// that order decides the order of CPoint::Point, hence the RCM tie-breaking and the edge numbering.
struct NodeAdj { int cnt[8]; int adj[8][4]; };
static const NodeAdj& node_adjacency(int vtk) {
  static const NodeAdj tri = {{2, 2, 2}, {{1, 2}, {2, 0}, {0, 1}}};
  static const NodeAdj quad = {{2, 2, 2, 2}, {{1, 3}, {2, 0}, {3, 1}, {0, 2}}};
  static const NodeAdj tet = {{3, 3, 3, 3}, {{1, 2, 3}, {0, 2, 3}, {0, 1, 3}, {0, 1, 2}}};
  static const NodeAdj hex = {{3, 3, 3, 3, 3, 3, 3, 3},
                              {{1, 3, 4}, {0, 2, 5}, {1, 3, 6}, {0, 2, 7}, {0, 5, 7}, {4, 6, 1}, {2, 5, 7}, {4, 3, 6}}};
  static const NodeAdj prism = {{3, 3, 3, 3, 3, 3}, {{1, 2, 3}, {0, 2, 4}, {1, 0, 5}, {0, 4, 5}, {3, 5, 1}, {4, 3, 2}}};
  static const NodeAdj pyra = {{3, 3, 3, 3, 4}, {{1, 3, 4}, {0, 2, 4}, {1, 3, 4}, {2, 0, 4}, {0, 1, 2, 3}}};
  switch (vtk) {
    case VTK_TRI: return tri;
    case VTK_QUAD: return quad;
    case VTK_TET: return tet;
    case VTK_HEX: return hex;
    case VTK_PRISM: return prism;
    case VTK_PYRAMID: return pyra;
    default: throw Error("unsupported volume element type " + std::to_string(vtk));
  }
}

// Neighbours of every point in the order CPhysicalGeometry::SetPoint_Connectivity (CPhysicalGeometry.cpp:4417-4476)
// finds them: elements of the point by ascending element index, within an element the adjacency listing above, first
// occurrence wins. Same sets (and the same halo-halo exclusion) as build_neighbours, so `ptr` is shared.
void build_neighbours_elem_order(const su2b200_mesh& m, const std::vector<long>& ptr, std::vector<int>& rnbr) {
  const long P = m.P;
  std::vector<long> eptr(P + 1, 0);
  for (int n : m.enode) eptr[n + 1]++;
  for (long i = 0; i < P; ++i) eptr[i + 1] += eptr[i];
  std::vector<int> pel(eptr[P]);
  {
    std::vector<long> pos(eptr.begin(), eptr.end() - 1);
    for (long e = 0; e < m.nElem(); ++e)
      for (long q = m.eptr[e]; q < m.eptr[e + 1]; ++q) pel[pos[m.enode[q]]++] = (int)e;
  }
  rnbr.assign(ptr[P], -1);
  for (long i = 0; i < P; ++i) {
    long w = ptr[i];
    for (long q = eptr[i]; q < eptr[i + 1]; ++q) {
      const int e = pel[q];
      const int* nd = &m.enode[m.eptr[e]];
      const NodeAdj& na = node_adjacency(m.etype[e]);
      const int nn = (int)(m.eptr[e + 1] - m.eptr[e]);
      for (int k = 0; k < nn; ++k) {
        if (nd[k] != i) continue;
        for (int a = 0; a < na.cnt[k]; ++a) {
          const int j = nd[na.adj[k][a]];
          if (i >= m.Pd && j >= m.Pd) continue;
          if (std::find(rnbr.begin() + ptr[i], rnbr.begin() + w, j) == rnbr.begin() + w) {
            SU2B_REQUIRE(w < ptr[i + 1], "inconsistent neighbour lists");
            rnbr[w++] = j;
          }
        }
      }
    }
    SU2B_REQUIRE(w == ptr[i + 1], "inconsistent neighbour lists");
  }
}

// Reverse Cuthill-McKee over the owned points (halos keep their place after them): BFS from a minimum-degree seed,
// neighbours visited by ascending degree, order reversed -- the recipe of CPhysicalGeometry::SetRCM_Ordering.
void rcm_reorder(su2b200_mesh& m) {
  std::vector<long> ptr;
  std::vector<int> nbr;
  {
    std::vector<int> asc;
    build_neighbours(m, ptr, asc);
    build_neighbours_elem_order(m, ptr, nbr);  // the reference walks CPoint::Point, which is in this order
  }
  const long Pd = m.Pd, P = m.P;
  auto degree = [&](long i) { return ptr[i + 1] - ptr[i]; };
  std::vector<char> seen(P, 0);
  std::vector<int> order;
  order.reserve(Pd);
  std::vector<int> byDeg(Pd);
  std::iota(byDeg.begin(), byDeg.end(), 0);
  std::stable_sort(byDeg.begin(), byDeg.end(), [&](int a, int b) { return degree(a) < degree(b); });
  size_t seedPos = 0;
  std::vector<int> tmp;
  while ((long)order.size() < Pd) {
    while (seen[byDeg[seedPos]]) ++seedPos;
    std::queue<int> q;
    q.push(byDeg[seedPos]);
    seen[byDeg[seedPos]] = 1;
    while (!q.empty()) {
      const int i = q.front();
      q.pop();
      order.push_back(i);
      tmp.clear();
      for (long k = ptr[i]; k < ptr[i + 1]; ++k) {
        const int j = nbr[k];
        if (j < Pd && !seen[j]) { seen[j] = 1; tmp.push_back(j); }
      }
      std::stable_sort(tmp.begin(), tmp.end(), [&](int a, int b) { return degree(a) < degree(b); });
      for (int j : tmp) q.push(j);
    }
  }
  std::reverse(order.begin(), order.end());
  std::vector<ulong_t> newIdx(P);
  for (long k = 0; k < Pd; ++k) newIdx[order[k]] = (ulong_t)k;
  for (long i = Pd; i < P; ++i) newIdx[i] = (ulong_t)i;
  // apply
  std::vector<double> c2(m.coord.size());
  std::vector<ulong_t> g2(P);
  for (long i = 0; i < P; ++i) {
    for (int d = 0; d < m.nDim; ++d) c2[(size_t)newIdx[i] * m.nDim + d] = m.coord[(size_t)i * m.nDim + d];
    g2[newIdx[i]] = m.global_index[i];
  }
  m.coord.swap(c2);
  m.global_index.swap(g2);
  for (auto& n : m.enode) n = (int)newIdx[n];
  for (auto& mk : m.markers)
    for (auto& n : mk.enode) n = (int)newIdx[n];
  for (auto& s : m.p2p_send) s = newIdx[s];
  for (auto& r : m.p2p_recv) r = newIdx[r];
  if (m.perm.empty()) m.perm = newIdx;
  else for (auto& p : m.perm) p = newIdx[p];
}

// ---- e2. decomposition of an unstructured mesh over N ranks --------------------------------------------------------------
// The reference colours the points with ParMETIS k-way on the point graph, vertex weights wp + we * degree
// (CPhysicalGeometry::SetColorGrid_Parallel, CPhysicalGeometry.cpp:7675-7755; PARMETIS_POINT_WEIGHT 0, PARMETIS_EDGE_WEIGHT 1)
// and then builds, per rank, the owned points, a one-layer halo and the send/receive lists (SetSendReceive :2956,
// PreprocessP2PComms CGeometry.cpp:139-332). ParMETIS is not part of this library; the colouring here is a weighted
// recursive coordinate bisection with the SAME vertex weights (deterministic, balanced to one point's weight): any
// colouring gives the same global answer, the partition only decides the halo volume. Everything after the colouring
// follows the reference: a rank keeps every element that touches an owned point (so owned control volumes are
// complete), owned points first, halos after them grouped by owner, both sides of a send/receive pair ordered by global
// point index.
void color_rcb(const su2b200_mesh& m, const std::vector<double>& w, std::vector<int>& idx, long b, long e, int c0, int nc,
               std::vector<int>& color) {
  if (nc == 1) {
    for (long k = b; k < e; ++k) color[idx[k]] = c0;
    return;
  }
  const int D = m.nDim;
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  for (long k = b; k < e; ++k)
    for (int d = 0; d < D; ++d) {
      const double x = m.coord[(size_t)idx[k] * D + d];
      lo[d] = std::min(lo[d], x);
      hi[d] = std::max(hi[d], x);
    }
  int axis = 0;
  for (int d = 1; d < D; ++d)
    if (hi[d] - lo[d] > hi[axis] - lo[axis]) axis = d;
  std::stable_sort(idx.begin() + b, idx.begin() + e, [&](int p, int q) {
    const double xp = m.coord[(size_t)p * D + axis], xq = m.coord[(size_t)q * D + axis];
    return xp < xq || (xp == xq && p < q);
  });
  const int ncl = nc / 2;
  double total = 0.0;
  for (long k = b; k < e; ++k) total += w[idx[k]];
  const double target = total * (double)ncl / (double)nc;
  double acc = 0.0;
  long cut = b;
  while (cut < e && acc + 0.5 * w[idx[cut]] < target) acc += w[idx[cut++]];
  cut = std::min(std::max(cut, b + 1), e - 1);
  color_rcb(m, w, idx, b, cut, c0, ncl, color);
  color_rcb(m, w, idx, cut, e, c0 + ncl, nc - ncl, color);
}

void wall_distance(su2b200_mesh& m);  // below
bool is_wall_kind(int k);

void color_points(const su2b200_mesh& m, int nranks, std::vector<int>& color) {
  std::vector<long> ptr;
  std::vector<int> nbr;
  build_neighbours(m, ptr, nbr);
  const double wp = 0.0, we = 1.0;  // PARMETIS_POINT_WEIGHT, PARMETIS_EDGE_WEIGHT defaults (CConfig.cpp:2944-2947)
  std::vector<double> w(m.P);
  for (long i = 0; i < m.P; ++i) w[i] = wp + we * (double)(ptr[i + 1] - ptr[i]);
  std::vector<int> idx(m.P);
  std::iota(idx.begin(), idx.end(), 0);
  color.assign(m.P, 0);
  color_rcb(m, w, idx, 0, m.P, 0, nranks, color);
}

// rank >= 0: the local mesh of that rank; rank == -1: the WHOLE mesh renumbered colour by colour (points of rank 0 first,
// ... each group in ascending original index), the ordering in which one GPU reproduces the N-rank block-Jacobi
// preconditioner with N row partitions (su2b200_linsys_set_partition_rows); part_ptr returns the group offsets
su2b200_mesh* partition_mesh(const su2b200_mesh& whole, int nranks, int rank, std::vector<ulong_t>* part_ptr) {
  SU2B_REQUIRE(whole.nranks == 1 && whole.P == whole.Pd, "partitioning starts from an undivided mesh");
  SU2B_REQUIRE(!whole.preprocessed, "partition before preprocessing");
  SU2B_REQUIRE(nranks >= 1 && rank >= -1 && rank < nranks, "invalid rank");
  SU2B_REQUIRE(whole.P >= 2L * nranks, "too many ranks for this mesh");
  std::vector<int> color;
  color_points(whole, nranks, color);
  const int D = whole.nDim;
  auto* m = new su2b200_mesh();
  m->nDim = D;
  std::vector<int> local(whole.P, -1);  // new index of a point of the whole mesh (-1: not on this rank)
  std::vector<int> pts;                 // points of the new mesh, in order (original indices)
  if (rank < 0) {
    if (part_ptr) part_ptr->assign(1, 0);
    for (int c = 0; c < nranks; ++c) {
      for (long i = 0; i < whole.P; ++i)
        if (color[i] == c) { local[i] = (int)pts.size(); pts.push_back((int)i); }
      if (part_ptr) part_ptr->push_back((ulong_t)pts.size());
    }
    m->Pd = whole.P;
  } else {
    m->rank = rank;
    m->nranks = nranks;
    for (long i = 0; i < whole.P; ++i)
      if (color[i] == rank) { local[i] = (int)pts.size(); pts.push_back((int)i); }
    m->Pd = (long)pts.size();
    SU2B_REQUIRE(m->Pd > 0, "a rank received no points");
    // halo layer: the other points of every element that touches an owned point, grouped by owner, ascending index
    std::vector<char> halo(whole.P, 0);
    for (long e = 0; e < whole.nElem(); ++e) {
      bool touches = false;
      for (long q = whole.eptr[e]; q < whole.eptr[e + 1]; ++q) touches = touches || color[whole.enode[q]] == rank;
      if (!touches) continue;
      for (long q = whole.eptr[e]; q < whole.eptr[e + 1]; ++q)
        if (color[whole.enode[q]] != rank) halo[whole.enode[q]] = 1;
    }
    for (int c = 0; c < nranks; ++c) {
      if (c == rank) continue;
      const size_t before = pts.size();
      for (long i = 0; i < whole.P; ++i)
        if (halo[i] && color[i] == c) { local[i] = (int)pts.size(); pts.push_back((int)i); }
      if (pts.size() > before) {
        m->p2p_ranks.push_back(c);
        for (size_t k = before; k < pts.size(); ++k) m->p2p_recv.push_back((ulong_t)k);
        m->p2p_recv_ptr.push_back((int)m->p2p_recv.size());
      }
    }
    // send lists: my owned points that are halos of rank c == share an element with a point owned by c. The halo
    // relation is symmetric (an element with points of r and c makes halos on both sides), so the neighbour ranks
    // found above are the complete list
    std::vector<std::vector<char>> sendTo(m->p2p_ranks.size(), std::vector<char>());
    std::vector<int> slot(nranks, -1);
    for (size_t k = 0; k < m->p2p_ranks.size(); ++k) { slot[m->p2p_ranks[k]] = (int)k; sendTo[k].assign(whole.P, 0); }
    for (long e = 0; e < whole.nElem(); ++e) {
      unsigned long long mask = 0;  // owner ranks present in the element (as slots)
      bool mine = false;
      for (long q = whole.eptr[e]; q < whole.eptr[e + 1]; ++q) {
        const int c = color[whole.enode[q]];
        if (c == rank) mine = true;
        else if (slot[c] >= 0 && slot[c] < 64) mask |= 1ull << slot[c];
      }
      if (!mine || !mask) continue;
      for (long q = whole.eptr[e]; q < whole.eptr[e + 1]; ++q) {
        const int p = whole.enode[q];
        if (color[p] != rank) continue;
        for (size_t k = 0; k < m->p2p_ranks.size(); ++k)
          if (mask >> k & 1ull) sendTo[k][p] = 1;
      }
    }
    SU2B_REQUIRE(m->p2p_ranks.size() <= 64, "more than 64 neighbour ranks");
    for (size_t k = 0; k < m->p2p_ranks.size(); ++k) {
      for (long i = 0; i < whole.P; ++i)
        if (sendTo[k][i]) m->p2p_send.push_back((ulong_t)local[i]);
      m->p2p_send_ptr.push_back((int)m->p2p_send.size());
    }
  }
  m->P = (long)pts.size();
  m->coord.resize((size_t)m->P * D);
  m->global_index.resize(m->P);
  for (long l = 0; l < m->P; ++l) {
    for (int d = 0; d < D; ++d) m->coord[(size_t)l * D + d] = whole.coord[(size_t)pts[l] * D + d];
    m->global_index[l] = whole.global_index[pts[l]];
  }
  auto keep = [&](const int* nd, int n) {
    bool any = false;
    for (int k = 0; k < n; ++k) {
      if (local[nd[k]] < 0) return false;  // an element reaching beyond the halo layer (cannot touch an owned point)
      any = any || local[nd[k]] < m->Pd;
    }
    return any;
  };
  for (long e = 0; e < whole.nElem(); ++e) {
    const int n = (int)(whole.eptr[e + 1] - whole.eptr[e]);
    const int* nd = &whole.enode[whole.eptr[e]];
    if (!keep(nd, n)) continue;
    int ln[8];
    for (int k = 0; k < n; ++k) ln[k] = local[nd[k]];
    add_elem(*m, whole.etype[e], ln, n);
  }
  for (const auto& wm : whole.markers) {
    HostMarker mk;
    mk.tag = wm.tag;
    mk.kind = wm.kind;
    mk.monitoring = wm.monitoring;
    mk.heat_flux = wm.heat_flux;
    for (size_t b = 0; b < wm.etype.size(); ++b) {
      const int n = (int)(wm.eptr[b + 1] - wm.eptr[b]);
      const int* nd = &wm.enode[wm.eptr[b]];
      if (!keep(nd, n)) continue;
      int ln[4];
      for (int k = 0; k < n; ++k) ln[k] = local[nd[k]];
      add_belem(mk, wm.etype[b], ln, n);
    }
    m->markers.push_back(std::move(mk));
  }
  // wall distance (RANS): the nearest wall element may lie on another rank -- computed once on the whole mesh
  bool anyWall = false;
  for (const auto& wm : whole.markers) anyWall = anyWall || is_wall_kind(wm.kind);
  if (anyWall && !whole.is_box) {
    su2b200_mesh tmp = whole;
    wall_distance(tmp);
    ulong_t maxg = 0;
    for (long i = 0; i < whole.P; ++i) maxg = std::max(maxg, whole.global_index[i]);
    m->wall_distance_global.assign(maxg + 1, 0.0);
    for (long i = 0; i < whole.P; ++i) m->wall_distance_global[whole.global_index[i]] = tmp.wall_distance[i];
  }
  return m;
}

bool is_sym_kind(int k) { return k == SU2B200_SYMMETRY_PLANE || k == SU2B200_EULER_WALL; }
bool is_wall_kind(int k) { return k == SU2B200_HEAT_FLUX || k == 25 /*ISOTHERMAL*/; }

double point_segment_dist2(const double* p, const double* a, const double* b, int D) {
  double ab[3] = {0, 0, 0}, ap[3] = {0, 0, 0}, l2 = 0, t = 0;
  for (int d = 0; d < D; ++d) { ab[d] = b[d] - a[d]; ap[d] = p[d] - a[d]; l2 += ab[d] * ab[d]; t += ab[d] * ap[d]; }
  t = l2 > 0 ? std::min(1.0, std::max(0.0, t / l2)) : 0.0;
  double d2 = 0;
  for (int d = 0; d < D; ++d) { const double r = ap[d] - t * ab[d]; d2 += r * r; }
  return d2;
}
double point_triangle_dist2(const double* p, const double* a, const double* b, const double* c) {
  // Ericson, Real-Time Collision Detection, closest point on triangle
  double ab[3], ac[3], ap[3];
  for (int d = 0; d < 3; ++d) { ab[d] = b[d] - a[d]; ac[d] = c[d] - a[d]; ap[d] = p[d] - a[d]; }
  auto dot = [](const double* x, const double* y) { return x[0] * y[0] + x[1] * y[1] + x[2] * y[2]; };
  auto d2to = [&](const double* q) { double s = 0; for (int d = 0; d < 3; ++d) s += (p[d] - q[d]) * (p[d] - q[d]); return s; };
  const double d1 = dot(ab, ap), d2 = dot(ac, ap);
  if (d1 <= 0 && d2 <= 0) return d2to(a);
  double bp[3]; for (int d = 0; d < 3; ++d) bp[d] = p[d] - b[d];
  const double d3 = dot(ab, bp), d4 = dot(ac, bp);
  if (d3 >= 0 && d4 <= d3) return d2to(b);
  const double vc = d1 * d4 - d3 * d2;
  if (vc <= 0 && d1 >= 0 && d3 <= 0) { const double v = d1 / (d1 - d3); double q[3]; for (int d = 0; d < 3; ++d) q[d] = a[d] + v * ab[d]; return d2to(q); }
  double cp[3]; for (int d = 0; d < 3; ++d) cp[d] = p[d] - c[d];
  const double d5 = dot(ab, cp), d6 = dot(ac, cp);
  if (d6 >= 0 && d5 <= d6) return d2to(c);
  const double vb = d5 * d2 - d1 * d6;
  if (vb <= 0 && d2 >= 0 && d6 <= 0) { const double w = d2 / (d2 - d6); double q[3]; for (int d = 0; d < 3; ++d) q[d] = a[d] + w * ac[d]; return d2to(q); }
  const double va = d3 * d6 - d5 * d4;
  if (va <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0) {
    const double w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
    double q[3]; for (int d = 0; d < 3; ++d) q[d] = b[d] + w * (c[d] - b[d]);
    return d2to(q);
  }
  const double den = 1.0 / (va + vb + vc), v = vb * den, w = vc * den;
  double q[3]; for (int d = 0; d < 3; ++d) q[d] = a[d] + ab[d] * v + ac[d] * w;
  return d2to(q);
}

// CGeometry::ComputeWall_Distance for the no-slip markers held by this rank. BOX meshes whose walls are whole box
// faces use the coordinate distance (exact, and valid across the slab decomposition without communication).
void wall_distance(su2b200_mesh& m) {
  const int D = m.nDim;
  m.wall_distance.assign(m.P, 0.0);
  std::vector<int> walls;
  for (size_t k = 0; k < m.markers.size(); ++k)
    if (is_wall_kind(m.markers[k].kind)) walls.push_back((int)k);
  if (walls.empty()) return;
  if (m.is_box) {
    for (long i = 0; i < m.P; ++i) {
      double best = 1e300;
      for (int w : walls) {
        const int axis = w / 2;
        const double plane = m.box_off[axis] + ((w & 1) ? m.box_len[axis] : 0.0);
        best = std::min(best, std::fabs(X(m, i)[axis] - plane));
      }
      m.wall_distance[i] = best;
    }
    return;
  }
  if (!m.wall_distance_global.empty()) {  // partitioned mesh: distances to the GLOBAL wall surface, by global index
    for (long i = 0; i < m.P; ++i) m.wall_distance[i] = m.wall_distance_global[m.global_index[i]];
    return;
  }
  SU2B_REQUIRE(m.nranks == 1, "wall distance of a partitioned unstructured mesh needs the global wall surface");
  for (long i = 0; i < m.P; ++i) {
    double best = 1e300;
    const double* p = X(m, i);
    for (int w : walls) {
      const auto& mk = m.markers[w];
      for (size_t b = 0; b < mk.etype.size(); ++b) {
        const int* nd = &mk.enode[mk.eptr[b]];
        double d2;
        if (mk.etype[b] == VTK_LINE) d2 = point_segment_dist2(p, X(m, nd[0]), X(m, nd[1]), D);
        else if (mk.etype[b] == VTK_TRI) d2 = point_triangle_dist2(p, X(m, nd[0]), X(m, nd[1]), X(m, nd[2]));
        else d2 = std::min(point_triangle_dist2(p, X(m, nd[0]), X(m, nd[1]), X(m, nd[2])),
                           point_triangle_dist2(p, X(m, nd[0]), X(m, nd[2]), X(m, nd[3])));
        best = std::min(best, d2);
      }
    }
    m.wall_distance[i] = std::sqrt(best);
  }
}

void preprocess(su2b200_mesh& m, int rcm) {
  const int D = m.nDim;
  fix_orientation(m);
  if (rcm) rcm_reorder(m);
  const long P = m.P;
  std::vector<long> ptr;
  std::vector<int> nbr;
  build_neighbours(m, ptr, nbr);
  // ---- edges: (i, j) with i < j, numbered as CGeometry::SetEdges (CGeometry.cpp:1263-1293) does: points ascending,
  // within a point its neighbours in CPoint::Point order (element order, see build_neighbours_elem_order) ----
  std::vector<int> rnbr;
  build_neighbours_elem_order(m, ptr, rnbr);
  auto asc_pos = [&](long a, long b) -> long {
    const long k = std::lower_bound(nbr.begin() + ptr[a], nbr.begin() + ptr[a + 1], (int)b) - nbr.begin();
    return (k < ptr[a + 1] && nbr[k] == b) ? k : -1;
  };
  long E = 0;
  for (long i = 0; i < P; ++i)
    for (long k = ptr[i]; k < ptr[i + 1]; ++k) E += nbr[k] > i;
  m.E = E;
  m.edge_n0.resize(E);
  m.edge_n1.resize(E);
  m.point_ptr.assign(ptr.begin(), ptr.end());
  m.point_nbr.assign(rnbr.begin(), rnbr.end());
  m.point_edge.assign(nbr.size(), 0);
  std::vector<long> asc_edge(nbr.size(), -1);  // edge of the k-th entry of the ascending neighbour lists
  {
    long e = 0;
    for (long i = 0; i < P; ++i)
      for (long k = ptr[i]; k < ptr[i + 1]; ++k) {
        const long j = rnbr[k];
        if (j < i) continue;
        m.edge_n0[e] = (ulong_t)i;
        m.edge_n1[e] = (ulong_t)j;
        asc_edge[asc_pos(i, j)] = e;
        asc_edge[asc_pos(j, i)] = e;
        ++e;
      }
    for (long i = 0; i < P; ++i)
      for (long k = ptr[i]; k < ptr[i + 1]; ++k) m.point_edge[k] = (ulong_t)asc_edge[asc_pos(i, rnbr[k])];
  }
  auto find_edge = [&](int a, int b) -> long {
    const long k = asc_pos(a, b);
    return k >= 0 ? asc_edge[k] : -1;
  };
  // ---- median-dual control volumes ----
  m.edge_normal.assign((size_t)E * D, 0.0);
  m.volume.assign(P, 0.0);
  for (long el = 0; el < m.nElem(); ++el) {
    const ElemTopo& t = topo(m.etype[el]);
    const int* nd = &m.enode[m.eptr[el]];
    double ec[3] = {0, 0, 0};
    for (int k = 0; k < t.nNodes; ++k)
      for (int d = 0; d < D; ++d) ec[d] += X(m, nd[k])[d];
    for (int d = 0; d < D; ++d) ec[d] /= t.nNodes;
    for (int f = 0; f < t.nFaces; ++f) {
      const int nf = t.nNodesFace[f];
      double fc[3] = {0, 0, 0};
      if (D == 3)
        for (int k = 0; k < nf; ++k)
          for (int d = 0; d < 3; ++d) fc[d] += X(m, nd[t.faces[f][k]])[d] / nf;
      const int nEdgesFace = D == 2 ? 1 : nf;
      for (int k = 0; k < nEdgesFace; ++k) {
        const int a = nd[t.faces[f][k]], b = nd[t.faces[f][D == 2 ? 1 : (k + 1) % nf]];
        const double *xa = X(m, a), *xb = X(m, b);
        double mid[3] = {0, 0, 0};
        for (int d = 0; d < D; ++d) mid[d] = 0.5 * (xa[d] + xb[d]);
        const double sgn = a < b ? 1.0 : -1.0;
        double va, vb, nrm[3] = {0, 0, 0};
        if (D == 2) {
          nrm[0] = ec[1] - mid[1];
          nrm[1] = -(ec[0] - mid[0]);
          auto tri_area = [&](const double* p) {
            return 0.5 * std::fabs((ec[0] - p[0]) * (mid[1] - p[1]) - (ec[1] - p[1]) * (mid[0] - p[0]));
          };
          va = tri_area(xa);
          vb = tri_area(xb);
        } else {
          double u[3], w[3];
          for (int d = 0; d < 3; ++d) { u[d] = ec[d] - mid[d]; w[d] = fc[d] - mid[d]; }
          cross(u, w, nrm);
          for (int d = 0; d < 3; ++d) nrm[d] *= 0.5;
          auto tet_vol = [&](const double* p) {
            double a1[3], a2[3], a3[3], c[3];
            for (int d = 0; d < 3; ++d) { a1[d] = mid[d] - p[d]; a2[d] = fc[d] - p[d]; a3[d] = ec[d] - p[d]; }
            cross(a1, a2, c);
            return std::fabs(a3[0] * c[0] + a3[1] * c[1] + a3[2] * c[2]) / 6.0;
          };
          va = tet_vol(xa);
          vb = tet_vol(xb);
        }
        m.volume[a] += va;
        m.volume[b] += vb;
        const long e = find_edge(std::min(a, b), std::max(a, b));
        if (e >= 0)
          for (int d = 0; d < D; ++d) m.edge_normal[(size_t)e * D + d] += sgn * nrm[d];
      }
    }
  }
  const double EPS = 1e-16;
  for (long e = 0; e < E; ++e) {
    double a2 = 0;
    for (int d = 0; d < D; ++d) a2 += m.edge_normal[(size_t)e * D + d] * m.edge_normal[(size_t)e * D + d];
    if (a2 == 0.0) m.edge_normal[(size_t)e * D] = EPS * EPS;
  }
  // ---- boundary vertices: normals point into the fluid (the reference's stored orientation) ----
  // point -> elements adjacency restricted to boundary points is avoided: orientation is fixed per boundary element
  // with the centroid of the neighbours of its nodes (interior side).
  for (auto& mk : m.markers) {
    std::vector<int> vidx(P, -1);
    mk.vnode.clear();
    for (int n : mk.enode)
      if (vidx[n] < 0) { vidx[n] = (int)mk.vnode.size(); mk.vnode.push_back((ulong_t)n); }
    const size_t nV = mk.vnode.size();
    mk.vnormal.assign(nV * D, 0.0);
    for (size_t b = 0; b < mk.etype.size(); ++b) {
      const int nn = bound_nodes(mk.etype[b]);
      const int* nd = &mk.enode[mk.eptr[b]];
      double cg[3] = {0, 0, 0};
      for (int k = 0; k < nn; ++k)
        for (int d = 0; d < D; ++d) cg[d] += X(m, nd[k])[d] / nn;
      // inward reference direction: towards the mean of all neighbours of the element's nodes
      double in[3] = {0, 0, 0};
      long cntn = 0;
      for (int k = 0; k < nn; ++k)
        for (long q = ptr[nd[k]]; q < ptr[nd[k] + 1]; ++q) {
          for (int d = 0; d < D; ++d) in[d] += X(m, nbr[q])[d] - cg[d];
          ++cntn;
        }
      double contrib[4][3] = {{0}};
      double tot[3] = {0, 0, 0};
      for (int k = 0; k < nn; ++k) {
        const double* xv = X(m, nd[k]);
        if (D == 2) {
          // half of the segment, rotated
          const double dxv = xv[0] - cg[0], dyv = xv[1] - cg[1];
          const double s = k == 0 ? 1.0 : -1.0;
          contrib[k][0] = s * dyv;
          contrib[k][1] = -s * dxv;
        } else {
          const double* xn = X(m, nd[(k + 1) % nn]);
          const double* xp = X(m, nd[(k + nn - 1) % nn]);
          double m1[3], m2[3], u[3], w[3], c1[3], c2[3];
          for (int d = 0; d < 3; ++d) { m1[d] = 0.5 * (xv[d] + xn[d]); m2[d] = 0.5 * (xv[d] + xp[d]); }
          for (int d = 0; d < 3; ++d) { u[d] = xv[d] - cg[d]; w[d] = m1[d] - cg[d]; }
          cross(u, w, c1);
          for (int d = 0; d < 3; ++d) { u[d] = xv[d] - m2[d]; w[d] = cg[d] - m2[d]; }
          cross(u, w, c2);
          for (int d = 0; d < 3; ++d) contrib[k][d] = 0.5 * (c1[d] + c2[d]);
        }
        for (int d = 0; d < D; ++d) tot[d] += contrib[k][d];
      }
      double s = 0;
      for (int d = 0; d < D; ++d) s += tot[d] * in[d];
      const double flip = (cntn > 0 && s < 0.0) ? -1.0 : 1.0;
      for (int k = 0; k < nn; ++k)
        for (int d = 0; d < D; ++d) mk.vnormal[(size_t)vidx[nd[k]] * D + d] += flip * contrib[k][d];
    }
    mk.vnbr.assign(nV, 0);
    mk.vunit.assign(nV * D, 0.0);
    for (size_t v = 0; v < nV; ++v) {
      const long i = (long)mk.vnode[v];
      double* nv = &mk.vnormal[v * D];
      double a2 = 0;
      for (int d = 0; d < D; ++d) a2 += nv[d] * nv[d];
      if (a2 == 0.0) { nv[0] = EPS * EPS; a2 = nv[0] * nv[0]; }
      const double area = std::sqrt(a2);
      for (int d = 0; d < D; ++d) mk.vunit[v * D + d] = nv[d] / area;
      // FindNormal_Neighbor: neighbour best aligned with the (inward) normal
      double cosmax = -1.0;
      long best = ptr[i + 1] > ptr[i] ? rnbr[ptr[i]] : i;
      for (long q = ptr[i]; q < ptr[i + 1]; ++q) {
        const long j = rnbr[q];  // CPoint::Point order: the last neighbour that reaches the maximum wins
        double sp = 0, nv2 = 0;
        for (int d = 0; d < D; ++d) { const double dc = X(m, j)[d] - X(m, i)[d]; sp += dc * nv[d]; nv2 += dc * dc; }
        const double ca = sp / (std::sqrt(nv2) * area);
        if (ca >= cosmax) { best = j; cosmax = ca; }
      }
      mk.vnbr[v] = (ulong_t)best;
    }
  }
  // ---- modified symmetry normals: Gram-Schmidt against the earlier symmetry / Euler-wall markers sharing the point
  {
    std::vector<int> sym;
    for (size_t k = 0; k < m.markers.size(); ++k)
      if (is_sym_kind(m.markers[k].kind)) sym.push_back((int)k);
    std::vector<std::map<long, long>> vertexOf(m.markers.size());
    for (int k : sym)
      for (size_t v = 0; v < m.markers[k].vnode.size(); ++v) vertexOf[k][(long)m.markers[k].vnode[v]] = (long)v;
    for (size_t a = 1; a < sym.size(); ++a) {
      auto& mk = m.markers[sym[a]];
      for (size_t v = 0; v < mk.vnode.size(); ++v) {
        const long i = (long)mk.vnode[v];
        if (i >= m.Pd) continue;
        double nrm[3] = {0, 0, 0};
        for (int d = 0; d < D; ++d) nrm[d] = mk.vunit[v * D + d];
        bool shared = false;
        for (size_t b = 0; b < a; ++b) {
          auto it = vertexOf[sym[b]].find(i);
          if (it == vertexOf[sym[b]].end()) continue;
          shared = true;
          const double* jn = &m.markers[sym[b]].vunit[(size_t)it->second * D];
          double proj = 0;
          for (int d = 0; d < D; ++d) proj += jn[d] * nrm[d];
          for (int d = 0; d < D; ++d) nrm[d] -= proj * jn[d];
        }
        if (!shared) continue;
        double a2 = 0;
        for (int d = 0; d < D; ++d) a2 += nrm[d] * nrm[d];
        const double area = std::sqrt(a2);
        if (area > 1e-12)
          for (int d = 0; d < D; ++d) mk.vunit[v * D + d] = nrm[d] / area;
      }
    }
  }
  // ---- sparse pattern (ascending columns, diagonal included) and edge -> block map ----
  m.row_ptr.assign(P + 1, 0);
  for (long i = 0; i < P; ++i) m.row_ptr[i + 1] = m.row_ptr[i] + (ulong_t)(ptr[i + 1] - ptr[i] + 1);
  m.nnz = (long)m.row_ptr[P];
  m.col_ind.resize(m.nnz);
  m.dia_ptr.resize(P);
  for (long i = 0; i < P; ++i) {
    ulong_t w = m.row_ptr[i];
    bool placed = false;
    for (long k = ptr[i]; k < ptr[i + 1]; ++k) {
      if (!placed && nbr[k] > i) { m.dia_ptr[i] = w; m.col_ind[w++] = (ulong_t)i; placed = true; }
      m.col_ind[w++] = (ulong_t)nbr[k];
    }
    if (!placed) { m.dia_ptr[i] = w; m.col_ind[w++] = (ulong_t)i; }
  }
  m.edge_ptr.resize((size_t)2 * E);
  for (long i = 0; i < P; ++i)
    for (long k = ptr[i]; k < ptr[i + 1]; ++k) {
      const long j = nbr[k];
      const long e = asc_edge[k];
      // block (i, j): position of j within row i = k - ptr[i] (+1 if the diagonal precedes it)
      const ulong_t posn = m.row_ptr[i] + (ulong_t)(k - ptr[i]) + (j > i ? 1 : 0);
      if (i < j) m.edge_ptr[2 * (size_t)e] = posn; else m.edge_ptr[2 * (size_t)e + 1] = posn;
    }
  // ---- point flags ----
  m.nNeighbor.resize(P);
  for (long i = 0; i < P; ++i) m.nNeighbor[i] = (unsigned short)(ptr[i + 1] - ptr[i]);
  m.physical_boundary.assign(P, 0);
  for (auto& mk : m.markers)
    if (mk.kind != SU2B200_SEND_RECEIVE)
      for (ulong_t n : mk.vnode) m.physical_boundary[n] = 1;
  wall_distance(m);
  // ---- greedy edge colouring (GetEdgeColoring, CGeometry.cpp:3737; group size 1) ----
  {
    std::vector<unsigned long long> used(P, 0ull);
    std::vector<int> color(E);
    int nc = 0;
    for (long e = 0; e < E; ++e) {
      const unsigned long long mm = used[m.edge_n0[e]] | used[m.edge_n1[e]];
      int c = 0;
      while (c < 64 && (mm >> c & 1ull)) ++c;
      SU2B_REQUIRE(c < 64, "edge colouring needs more than 64 colours");
      color[e] = c;
      used[m.edge_n0[e]] |= 1ull << c;
      used[m.edge_n1[e]] |= 1ull << c;
      nc = std::max(nc, c + 1);
    }
    m.color_ptr.assign(nc + 1, 0);
    for (long e = 0; e < E; ++e) m.color_ptr[color[e] + 1]++;
    for (int c = 0; c < nc; ++c) m.color_ptr[c + 1] += m.color_ptr[c];
    std::vector<ulong_t> pos(m.color_ptr.begin(), m.color_ptr.end() - 1);
    m.color_edges.resize(E);
    for (long e = 0; e < E; ++e) m.color_edges[pos[color[e]]++] = (ulong_t)e;
  }
  m.preprocessed = true;
}

// ---- .su2 ASCII writer (format of CSU2ASCIIMeshReaderBase.cpp: NDIME / NELEM / NPOIN / NMARK blocks, VTK type ids;
// layout as in QuickStart/mesh_NACA0012_inv.su2). Writes the mesh in its current numbering; serial meshes only. ----
void write_su2(const su2b200_mesh& m, const char* path) {
  SU2B_REQUIRE(m.nranks == 1, "only undivided meshes can be written");
  FILE* f = std::fopen(path, "w");
  if (!f) throw Error(std::string("cannot open ") + path + " for writing");
  std::fprintf(f, "NDIME= %d\nNELEM= %ld\n", m.nDim, m.nElem());
  for (long e = 0; e < m.nElem(); ++e) {
    std::fprintf(f, "%d", m.etype[e]);
    for (long q = m.eptr[e]; q < m.eptr[e + 1]; ++q) std::fprintf(f, "\t%d", m.enode[q]);
    std::fprintf(f, "\t%ld\n", e);
  }
  std::fprintf(f, "NPOIN= %ld\n", m.P);
  for (long i = 0; i < m.P; ++i) {
    for (int d = 0; d < m.nDim; ++d) std::fprintf(f, "%.17g\t", m.coord[(size_t)i * m.nDim + d]);
    std::fprintf(f, "%ld\n", i);
  }
  std::fprintf(f, "NMARK= %zu\n", m.markers.size());
  for (const auto& mk : m.markers) {
    std::fprintf(f, "MARKER_TAG= %s\nMARKER_ELEMS= %zu\n", mk.tag.c_str(), mk.etype.size());
    for (size_t b = 0; b < mk.etype.size(); ++b) {
      std::fprintf(f, "%d", mk.etype[b]);
      for (long q = mk.eptr[b]; q < mk.eptr[b + 1]; ++q) std::fprintf(f, "\t%d", mk.enode[q]);
      std::fprintf(f, "\n");
    }
  }
  std::fclose(f);
}

// Geometric stretching of a structured box along one axis (the "wing-box" input of SURVEY 8(d)(4): wall-normal spacing
// first_cell, first_cell*ratio, ... from the minus side, rescaled to the box length). Coordinates only.
void stretch_axis(su2b200_mesh& m, int axis, double first_cell, double ratio) {
  SU2B_REQUIRE(m.is_box && !m.preprocessed, "stretching applies to BOX/RECTANGLE meshes before preprocessing");
  SU2B_REQUIRE(axis >= 0 && axis < m.nDim && first_cell > 0.0 && ratio > 0.0, "invalid stretching parameters");
  const double L = m.box_len[axis], o = m.box_off[axis];
  // number of cells along the axis from the uniform spacing of the generator
  double hmin = L;
  for (long i = 0; i < m.P; ++i) {
    const double x = m.coord[(size_t)i * m.nDim + axis] - o;
    if (x > 1e-14 * L && x < hmin) hmin = x;
  }
  const long nc = std::lround(L / hmin);
  std::vector<double> y(nc + 1, 0.0);
  double h = first_cell;
  for (long k = 1; k <= nc; ++k) { y[k] = y[k - 1] + h; h *= ratio; }
  for (long k = 0; k <= nc; ++k) y[k] *= L / y[nc];
  for (long i = 0; i < m.P; ++i) {
    double& x = m.coord[(size_t)i * m.nDim + axis];
    const long k = std::lround((x - o) / hmin);
    x = o + y[std::min(std::max(k, 0L), nc)];
  }
}

}  // namespace

#define API_BEGIN try {
#define API_END                                   \
  return 0;                                       \
  }                                               \
  catch (const std::exception& e) {               \
    su2b200::set_error(e.what());                 \
    return 1;                                     \
  }                                               \
  catch (...) {                                   \
    su2b200::set_error("unknown C++ exception");  \
    return 2;                                     \
  }

extern "C" {

int su2b200_mesh_create_box(int nDim, const unsigned long n[3], const double length[3], const double offset[3],
                            int rank, int nranks, int tets, su2b200_mesh** out) {
  API_BEGIN
  SU2B_REQUIRE(out && n && length && offset, "null argument");
  *out = make_box(nDim, n, length, offset, rank, nranks, tets);
  API_END
}

int su2b200_mesh_read_su2(const char* path, su2b200_mesh** out) {
  API_BEGIN
  SU2B_REQUIRE(out && path, "null argument");
  *out = read_su2(path);
  API_END
}

int su2b200_mesh_write_su2(su2b200_mesh* m, const char* path) {
  API_BEGIN
  SU2B_REQUIRE(m && path, "null argument");
  write_su2(*m, path);
  API_END
}

int su2b200_mesh_stretch(su2b200_mesh* m, int axis, double first_cell, double ratio) {
  API_BEGIN
  SU2B_REQUIRE(m, "null mesh");
  stretch_axis(*m, axis, first_cell, ratio);
  API_END
}

int su2b200_mesh_partition(su2b200_mesh* whole, int nranks, int rank, su2b200_mesh** out, unsigned long* part_ptr) {
  API_BEGIN
  SU2B_REQUIRE(whole && out, "null argument");
  std::vector<ulong_t> pp;
  *out = partition_mesh(*whole, nranks, rank, &pp);
  if (part_ptr && rank < 0)
    for (size_t k = 0; k < pp.size(); ++k) part_ptr[k] = pp[k];
  API_END
}

int su2b200_mesh_destroy(su2b200_mesh* m) {
  delete m;
  return 0;
}

int su2b200_mesh_set_marker(su2b200_mesh* m, const char* tag, int kindBC, int monitoring, double heat_flux) {
  API_BEGIN
  SU2B_REQUIRE(m && tag, "null argument");
  for (auto& mk : m->markers)
    if (mk.tag == tag) {
      mk.kind = kindBC;
      mk.monitoring = monitoring;
      mk.heat_flux = heat_flux;
      m->preprocessed = false;
      return 0;
    }
  throw Error(std::string("marker not found in mesh: ") + tag);
  API_END
}

int su2b200_mesh_preprocess(su2b200_mesh* m, int rcm) {
  API_BEGIN
  SU2B_REQUIRE(m, "null mesh");
  for (auto& mk : m->markers) SU2B_REQUIRE(mk.kind != 0, "marker '" + mk.tag + "' has no boundary condition (su2b200_mesh_set_marker)");
  preprocess(*m, rcm);
  API_END
}

int su2b200_mesh_sizes(su2b200_mesh* m, int* nDim, unsigned long* nPoint, unsigned long* nPointDomain,
                       unsigned long* nEdge, unsigned long* nElem, unsigned long* nnz, int* nMarker) {
  API_BEGIN
  SU2B_REQUIRE(m, "null mesh");
  if (nDim) *nDim = m->nDim;
  if (nPoint) *nPoint = (ulong_t)m->P;
  if (nPointDomain) *nPointDomain = (ulong_t)m->Pd;
  if (nEdge) *nEdge = (ulong_t)m->E;
  if (nElem) *nElem = (ulong_t)m->nElem();
  if (nnz) *nnz = (ulong_t)m->nnz;
  if (nMarker) *nMarker = (int)m->markers.size();
  API_END
}

// read-only view of a named array (tests, output): dtype 0 = f64, 1 = u64, 2 = u16, 3 = u8, 4 = i32
int su2b200_mesh_array(su2b200_mesh* m, const char* name, const void** data, unsigned long* count, int* dtype) {
  API_BEGIN
  SU2B_REQUIRE(m && name && data && count && dtype, "null argument");
  const std::string s(name);
  auto f64 = [&](const std::vector<double>& v) { *data = v.data(); *count = v.size(); *dtype = 0; };
  auto u64 = [&](const std::vector<ulong_t>& v) { *data = v.data(); *count = v.size(); *dtype = 1; };
  auto i32 = [&](const std::vector<int>& v) { *data = v.data(); *count = v.size(); *dtype = 4; };
  if (s == "coord") f64(m->coord);
  else if (s == "volume") f64(m->volume);
  else if (s == "edge_normal") f64(m->edge_normal);
  else if (s == "wall_distance") f64(m->wall_distance);
  else if (s == "global_index") u64(m->global_index);
  else if (s == "edge_node0") u64(m->edge_n0);
  else if (s == "edge_node1") u64(m->edge_n1);
  else if (s == "point_ptr") u64(m->point_ptr);
  else if (s == "point_nbr") u64(m->point_nbr);
  else if (s == "point_edge") u64(m->point_edge);
  else if (s == "row_ptr") u64(m->row_ptr);
  else if (s == "col_ind") u64(m->col_ind);
  else if (s == "dia_ptr") u64(m->dia_ptr);
  else if (s == "edge_ptr") u64(m->edge_ptr);
  else if (s == "color_ptr") u64(m->color_ptr);
  else if (s == "color_edges") u64(m->color_edges);
  else if (s == "p2p_send") u64(m->p2p_send);
  else if (s == "p2p_recv") u64(m->p2p_recv);
  else if (s == "p2p_ranks") i32(m->p2p_ranks);
  else if (s == "p2p_send_ptr") i32(m->p2p_send_ptr);
  else if (s == "p2p_recv_ptr") i32(m->p2p_recv_ptr);
  else if (s == "nNeighbor") { *data = m->nNeighbor.data(); *count = m->nNeighbor.size(); *dtype = 2; }
  else if (s == "physical_boundary") { *data = m->physical_boundary.data(); *count = m->physical_boundary.size(); *dtype = 3; }
  else if (s.rfind("marker", 0) == 0) {
    // marker<k>_{node,normal,normal_neighbor,sym_unit_normal}
    const size_t us = s.find('_');
    SU2B_REQUIRE(us != std::string::npos, "bad marker array name");
    const int k = std::atoi(s.substr(6, us - 6).c_str());
    SU2B_REQUIRE(k >= 0 && k < (int)m->markers.size(), "marker index out of range");
    const std::string what = s.substr(us + 1);
    auto& mk = m->markers[k];
    if (what == "node") u64(mk.vnode);
    else if (what == "normal") f64(mk.vnormal);
    else if (what == "normal_neighbor") u64(mk.vnbr);
    else if (what == "sym_unit_normal") f64(mk.vunit);
    else throw Error("unknown marker array " + what);
  } else {
    throw Error("unknown mesh array " + s);
  }
  API_END
}

int su2b200_mesh_marker_info(su2b200_mesh* m, int iMarker, char tag[64], int* kindBC, int* monitoring,
                             unsigned long* nVertex, unsigned long* nElemBound) {
  API_BEGIN
  SU2B_REQUIRE(m && iMarker >= 0 && iMarker < (int)m->markers.size(), "marker index out of range");
  auto& mk = m->markers[iMarker];
  if (tag) { std::strncpy(tag, mk.tag.c_str(), 63); tag[63] = 0; }
  if (kindBC) *kindBC = mk.kind;
  if (monitoring) *monitoring = mk.monitoring;
  if (nVertex) *nVertex = (ulong_t)mk.vnode.size();
  if (nElemBound) *nElemBound = (ulong_t)mk.etype.size();
  API_END
}

// CSolver construction on the device: geometry + markers + halo tables of a preprocessed mesh
int su2b200_mesh_upload(su2b200_ctx* c, su2b200_mesh* m) {
  if (!c || !m) { su2b200::set_error("null argument"); return 1; }
  if (!m->preprocessed) { su2b200::set_error("call su2b200_mesh_preprocess first"); return 1; }
  su2b200_geometry_desc d{};
  d.nDim = m->nDim;
  d.nPoint = (ulong_t)m->P;
  d.nPointDomain = (ulong_t)m->Pd;
  d.nEdge = (ulong_t)m->E;
  d.edge_node0 = m->edge_n0.data();
  d.edge_node1 = m->edge_n1.data();
  d.edge_normal = m->edge_normal.data();
  d.coord = m->coord.data();
  d.volume = m->volume.data();
  d.point_ptr = m->point_ptr.data();
  d.point_nbr = m->point_nbr.data();
  d.point_edge = m->point_edge.data();
  d.row_ptr = m->row_ptr.data();
  d.col_ind = m->col_ind.data();
  d.dia_ptr = m->dia_ptr.data();
  d.edge_ptr = m->edge_ptr.data();
  d.nNeighbor = m->nNeighbor.data();
  d.physical_boundary = m->physical_boundary.data();
  d.wall_distance = m->wall_distance.data();
  d.color_ptr = m->color_ptr.data();
  d.color_edges = m->color_edges.data();
  d.nColor = m->color_ptr.empty() ? 0 : (ulong_t)m->color_ptr.size() - 1;
  int rc = su2b200_geometry_upload(c, &d);
  if (rc) return rc;
  for (size_t k = 0; k < m->markers.size(); ++k) {
    auto& mk = m->markers[k];
    rc = su2b200_marker_upload(c, (int)k, mk.kind, (ulong_t)mk.vnode.size(), mk.vnode.data(), mk.vnormal.data(),
                               mk.vnbr.data(), mk.vunit.data(), mk.monitoring, mk.heat_flux);
    if (rc) return rc;
  }
  if (!m->p2p_ranks.empty()) {
    rc = su2b200_p2p_upload(c, (int)m->p2p_ranks.size(), m->p2p_ranks.data(), m->p2p_send_ptr.data(), m->p2p_send.data(),
                            m->p2p_recv_ptr.data(), m->p2p_recv.data());
    if (rc) return rc;
  }
  return 0;
}

}  // extern "C"
