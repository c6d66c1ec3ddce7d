// The extern "C" layer (include/su2b200.h): argument checking, uploads, and the sequencing of the
// reference's solver virtuals. No numerics live here.
#include "internal.hpp"
#include "kernels.hpp"
#include <algorithm>
#include <cmath>
#include <numeric>

namespace su2b200 {
static thread_local std::string g_last_error;
void set_error(const std::string& msg) { g_last_error = msg; }
void bench_edge_kernel(su2b200_ctx* c, int reps, float* ms_avg);
}  // namespace su2b200

using namespace su2b200;

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

static void require_ctx(su2b200_ctx* c) {
  if (!c) throw Error("null context");
  SU2B_CUDA(cudaSetDevice(c->device));
}

extern "C" {

const char* su2b200_last_error(void) { return g_last_error.c_str(); }
int su2b200_version(void) { return 100; }
int su2b200_struct_sizes(int* g, int* f, int* r) {
  if (g) *g = (int)sizeof(su2b200_geometry_desc);
  if (f) *f = (int)sizeof(su2b200_flow_config);
  if (r) *r = (int)sizeof(su2b200_iter_report);
  return 0;
}

int su2b200_create(int device, su2b200_ctx** out) {
  API_BEGIN
  SU2B_REQUIRE(out, "null output pointer");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    throw Error(std::string("no CUDA device available (libsu2b200 has no CPU fallback): ") + cudaGetErrorString(e));
  SU2B_REQUIRE(device >= 0 && device < ndev, "device index out of range");
  SU2B_CUDA(cudaSetDevice(device));
  auto* c = new su2b200_ctx();
  c->device = device;
  SU2B_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  SU2B_CUDA(cudaStreamCreateWithFlags(&c->comm_stream, cudaStreamNonBlocking));
  SU2B_CUDA(cudaEventCreate(&c->ev0));
  SU2B_CUDA(cudaEventCreate(&c->ev1));
  *out = c;
  API_END
}

int su2b200_destroy(su2b200_ctx* c) {
  API_BEGIN
  if (!c) return 0;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  if (c->fs.h_red) cudaFreeHost(c->fs.h_red);
  if (c->fs.h_counters) cudaFreeHost(c->fs.h_counters);
  cudaEventDestroy(c->ev0);
  cudaEventDestroy(c->ev1);
  mailbox_release(c);
  if (c->ev_packed) cudaEventDestroy(c->ev_packed);
  if (c->ev_unpacked) cudaEventDestroy(c->ev_unpacked);
  if (c->comm_stream) cudaStreamSynchronize(c->comm_stream);
  cudaStream_t s0 = c->stream, s1 = c->comm_stream;
  const bool borrowed = c->stream_borrowed;
  delete c;
  if (!borrowed) cudaStreamDestroy(s0);
  cudaStreamDestroy(s1);
  API_END
}

int su2b200_synchronize(su2b200_ctx* c) {
  API_BEGIN
  require_ctx(c);
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  API_END
}

// ---------------------------------------------------------------------------------------------
// geometry
// ---------------------------------------------------------------------------------------------
static std::vector<int> to_int(const unsigned long* p, size_t n) {
  std::vector<int> v(n);
  for (size_t i = 0; i < n; ++i) {
    if (p[i] > 0x7fffffffUL) throw Error("index exceeds int32 range (per-GPU limit 2^31 entries)");
    v[i] = (int)p[i];
  }
  return v;
}

int su2b200_geometry_upload(su2b200_ctx* c, const su2b200_geometry_desc* d) {
  API_BEGIN
  require_ctx(c);
  SU2B_REQUIRE(d, "null geometry descriptor");
  SU2B_REQUIRE(d->nDim == 2 || d->nDim == 3, "nDim must be 2 or 3");
  SU2B_REQUIRE(d->edge_node0 && d->edge_node1 && d->edge_normal && d->coord && d->volume, "missing dual-grid arrays");
  SU2B_REQUIRE(d->point_ptr && d->point_nbr && d->point_edge, "missing neighbour CSR");
  SU2B_REQUIRE(d->row_ptr && d->col_ind && d->edge_ptr, "missing sparse pattern");
  auto& g = c->geo;
  g.nDim = d->nDim;
  g.P = (int)d->nPoint;
  g.Pd = (int)d->nPointDomain;
  g.E = (int)d->nEdge;
  const int nd = g.nDim, P = g.P, E = g.E;
  auto ei = to_int(d->edge_node0, E), ej = to_int(d->edge_node1, E);
  for (int e = 0; e < E; ++e) SU2B_REQUIRE(ei[e] < ej[e] && ej[e] < P, "edges must satisfy iPoint < jPoint < nPoint");
  g.edge_i.upload(ei);
  g.edge_j.upload(ej);
  std::vector<double> nsoa((size_t)nd * E);
  for (int e = 0; e < E; ++e)
    for (int k = 0; k < nd; ++k) nsoa[(size_t)k * E + e] = d->edge_normal[(size_t)e * nd + k];
  g.edge_normal.upload(nsoa);
  g.coord.upload(d->coord, (size_t)P * nd);
  g.volume.upload(d->volume, P);
  auto nptr = to_int(d->point_ptr, P + 1);
  SU2B_REQUIRE(nptr[P] == 2 * E, "neighbour CSR must hold 2*nEdge entries");
  auto npt = to_int(d->point_nbr, 2 * (size_t)E), ned = to_int(d->point_edge, 2 * (size_t)E);
  g.nbr_ptr.upload(nptr);
  g.nbr_pt.upload(npt);
  g.nbr_edge.upload(ned);
  g.h_row_ptr = to_int(d->row_ptr, P + 1);
  g.nnz = g.h_row_ptr[P];
  g.h_col_ind = to_int(d->col_ind, g.nnz);
  g.h_dia_ptr.resize(P);
  for (int i = 0; i < P; ++i) {
    int dp = -1;
    for (int k = g.h_row_ptr[i]; k < g.h_row_ptr[i + 1]; ++k)
      if (g.h_col_ind[k] == i) dp = k;
    SU2B_REQUIRE(dp >= 0, "sparse pattern lacks a diagonal entry");
    if (d->dia_ptr) SU2B_REQUIRE((int)d->dia_ptr[i] == dp, "dia_ptr inconsistent with col_ind");
    g.h_dia_ptr[i] = dp;
  }
  g.row_ptr.upload(g.h_row_ptr);
  g.col_ind.upload(g.h_col_ind);
  g.dia_ptr.upload(g.h_dia_ptr);
  std::vector<int> eij(E), eji(E);
  for (int e = 0; e < E; ++e) {
    eij[e] = (int)d->edge_ptr[2 * (size_t)e];
    eji[e] = (int)d->edge_ptr[2 * (size_t)e + 1];
    SU2B_REQUIRE(g.h_col_ind[eij[e]] == ej[e] && g.h_col_ind[eji[e]] == ei[e], "edge_ptr inconsistent with pattern");
  }
  g.edge_ij.upload(eij);
  g.edge_ji.upload(eji);
  // gather lists: per point, incident edges in ascending edge index with the transposed block
  {
    std::vector<int> ge(2 * (size_t)E), gb(2 * (size_t)E);
    for (int i = 0; i < P; ++i) {
      const int b = nptr[i], en = nptr[i + 1];
      std::vector<int> ord(en - b);
      std::iota(ord.begin(), ord.end(), b);
      std::sort(ord.begin(), ord.end(), [&](int a, int bb) { return ned[a] < ned[bb]; });
      for (int k = 0; k < en - b; ++k) {
        const int e = ned[ord[k]];
        const bool first = (ei[e] == i);
        SU2B_REQUIRE(first || ej[e] == i, "CPoint::Edge inconsistent with CEdge::Nodes");
        ge[b + k] = first ? e : ~e;
        gb[b + k] = first ? eji[e] : eij[e];
      }
    }
    g.gat_edge.upload(ge);
    g.gat_blk.upload(gb);
  }
  std::vector<int> nn(P);
  for (int i = 0; i < P; ++i) nn[i] = d->nNeighbor ? (int)d->nNeighbor[i] : nptr[i + 1] - nptr[i];
  g.nnbr.upload(nn);
  g.nnbr_global = false;
  std::vector<unsigned char> pb(P, 0);
  if (d->physical_boundary) pb.assign(d->physical_boundary, d->physical_boundary + P);
  g.phys_bnd.upload(pb);
  g.have_phys_bnd = d->physical_boundary != nullptr;
  if (d->wall_distance) g.wall_dist.upload(d->wall_distance, P);
  if (d->color_ptr && d->color_edges && d->nColor > 0) {
    g.color_ptr = to_int(d->color_ptr, d->nColor + 1);
    g.color_edges.upload(to_int(d->color_edges, E));
  } else {
    // greedy edge colouring (same goal as colorSparsePattern, graph_toolbox.hpp:491-602, group size 1)
    std::vector<int> color(E, -1);
    std::vector<unsigned long long> used(P, 0ull);
    int nc = 0;
    for (int e = 0; e < E; ++e) {
      unsigned long long m = used[ei[e]] | used[ej[e]];
      int col = 0;
      while (col < 64 && (m >> col & 1ull)) ++col;
      SU2B_REQUIRE(col < 64, "edge colouring needs more than 64 colours");
      color[e] = col;
      used[ei[e]] |= 1ull << col;
      used[ej[e]] |= 1ull << col;
      nc = std::max(nc, col + 1);
    }
    g.color_ptr.assign(nc + 1, 0);
    for (int e = 0; e < E; ++e) g.color_ptr[color[e] + 1]++;
    for (int k = 0; k < nc; ++k) g.color_ptr[k + 1] += g.color_ptr[k];
    std::vector<int> pos(g.color_ptr.begin(), g.color_ptr.end() - 1), ce(E);
    for (int e = 0; e < E; ++e) ce[pos[color[e]]++] = e;
    g.color_edges.upload(ce);
  }
  g.markers.clear();
  g.uploaded = true;
  c->fs.allocated = false;
  API_END
}

int su2b200_marker_upload(su2b200_ctx* c, int iMarker, int kindBC, unsigned long nVertex,
                          const unsigned long* vertex_node, const double* vertex_normal,
                          const unsigned long* vertex_normal_neighbor, const double* sym_unit_normal,
                          int monitoring, double heat_flux) {
  API_BEGIN
  require_ctx(c);
  auto& g = c->geo;
  SU2B_REQUIRE(g.uploaded, "upload the geometry before its markers");
  SU2B_REQUIRE(iMarker >= 0 && iMarker < 4096, "marker index out of range");
  if ((int)g.markers.size() <= iMarker) g.markers.resize(iMarker + 1);
  auto& m = g.markers[iMarker];
  const int nd = g.nDim, nV = (int)nVertex;
  m.kind = kindBC;
  m.nVertex = nV;
  m.monitoring = monitoring;
  m.heat_flux = heat_flux;
  if (nV == 0) return 0;
  SU2B_REQUIRE(vertex_node && vertex_normal, "missing vertex arrays");
  auto vn = to_int(vertex_node, nV);
  {
    std::vector<char> seen(g.P, 0);
    for (int v : vn) {
      SU2B_REQUIRE(v < g.P, "vertex node out of range");
      SU2B_REQUIRE(!seen[v], "a point appears twice in one marker");
      seen[v] = 1;
    }
  }
  m.node.upload(vn);
  m.normal.upload(vertex_normal, (size_t)nV * nd);
  std::vector<int> nb(nV, 0);
  if (vertex_normal_neighbor) nb = to_int(vertex_normal_neighbor, nV);
  m.normal_nbr.upload(nb);
  std::vector<double> un((size_t)nV * nd);
  for (int v = 0; v < nV; ++v) {
    if (sym_unit_normal) {
      for (int k = 0; k < nd; ++k) un[(size_t)v * nd + k] = sym_unit_normal[(size_t)v * nd + k];
    } else {
      double a = 0;
      for (int k = 0; k < nd; ++k) a += vertex_normal[(size_t)v * nd + k] * vertex_normal[(size_t)v * nd + k];
      a = std::sqrt(a);
      for (int k = 0; k < nd; ++k) un[(size_t)v * nd + k] = vertex_normal[(size_t)v * nd + k] / a;
    }
  }
  m.unit_sym.upload(un);
  m.charac.alloc((size_t)nV * (nd + 9));
  m.charac.zero();
  SU2B_CUDA(cudaDeviceSynchronize());
  // CPoint::PhysicalBoundary derived from markers unless given (CPhysicalGeometry.cpp:3402-3408)
  if (!g.have_phys_bnd && kindBC != SU2B200_SEND_RECEIVE) {
    auto pb = g.phys_bnd.to_host();
    for (int v : vn) pb[v] = 1;
    g.phys_bnd.upload(pb);
  }
  API_END
}

int su2b200_marker_charac_download(su2b200_ctx* c, int iMarker, double* host) {
  API_BEGIN
  require_ctx(c);
  SU2B_REQUIRE(iMarker >= 0 && iMarker < (int)c->geo.markers.size() && host, "marker index out of range");
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  c->geo.markers[iMarker].charac.download(host);
  API_END
}

int su2b200_p2p_upload(su2b200_ctx* c, int n, const int* ranks, const int* send_ptr, const unsigned long* send_idx,
                       const int* recv_ptr, const unsigned long* recv_idx) {
  API_BEGIN
  require_ctx(c);
  c->p2p_ranks.assign(ranks, ranks + n);
  c->p2p_send_ptr.assign(send_ptr, send_ptr + n + 1);
  c->p2p_recv_ptr.assign(recv_ptr, recv_ptr + n + 1);
  c->p2p_send_idx.upload(to_int(send_idx, send_ptr[n]));
  c->p2p_recv_idx.upload(to_int(recv_idx, recv_ptr[n]));
  API_END
}

// ---------------------------------------------------------------------------------------------
// configuration + state
// ---------------------------------------------------------------------------------------------
int su2b200_flow_configure(su2b200_ctx* c, const su2b200_flow_config* cfg) {
  API_BEGIN
  require_ctx(c);
  SU2B_REQUIRE(cfg, "null config");
  SU2B_REQUIRE(cfg->gamma > 1.0 && cfg->gas_constant > 0.0, "invalid gas model");
  SU2B_REQUIRE(cfg->conv_scheme == SU2B200_CONV_ROE || cfg->conv_scheme == SU2B200_CONV_JST, "unsupported CONV_NUM_METHOD_FLOW");
  c->cfg = *cfg;
  c->configured = true;
  for (int s = 0; s < 2; ++s) c->sys[s].nparts = std::max(1, cfg->prec_partitions);
  API_END
}

static void linsys_bind(su2b200_ctx* c, int which, int nVar) {
  auto& g = c->geo;
  auto& ls = c->sys[which];
  ls.nVar = nVar;
  ls.P = g.P;
  ls.Pd = g.Pd;
  ls.nnz = g.nnz;
  ls.row_ptr = g.row_ptr.p;
  ls.col_ind = g.col_ind.p;
  ls.dia_ptr = g.dia_ptr.p;
  ls.h_row_ptr = g.h_row_ptr;
  ls.h_col_ind = g.h_col_ind;
  ls.h_dia_ptr = g.h_dia_ptr;
  ls.levels_ready = false;
  ls.n_halo_rows = -1;
  ls.rest_halo_ready = false;
  ls.upper_prod_of = nullptr;
  ls.h_parts.clear();
  ls.A.alloc((size_t)g.nnz * nVar * nVar);
  ls.A.zero();
}

int su2b200_flow_allocate(su2b200_ctx* c) {
  API_BEGIN
  require_ctx(c);
  auto& g = c->geo;
  auto& s = c->fs;
  SU2B_REQUIRE(g.uploaded && c->configured, "geometry and configuration must be set first");
  const int nd = g.nDim, P = g.P, E = g.E;
  s.nVar = nd + 2;
  s.nPrim = nd + 9;
  s.nG = nd + 2;
  const int nv = s.nVar;
  auto alloc0 = [](dvec<double>& v, size_t n) { v.alloc(n); v.zero(); };
  if (c->nranks > 1 && !g.nnbr_global) {
    // CPoint::nNeighbor of a halo point is the count on its owner (the local neighbour list of a halo point is
    // incomplete); the JST coefficients of an edge use it for both end points (centered.hpp:123-127)
    std::vector<int> nn = g.nnbr.to_host();
    std::vector<double> t(nn.begin(), nn.end());
    dvec<double> d;
    d.upload(t);
    SU2B_CUDA(cudaDeviceSynchronize());
    halo_exchange(c, d.p, 1);
    SU2B_CUDA(cudaDeviceSynchronize());
    t = d.to_host();
    for (size_t i = 0; i < nn.size(); ++i) nn[i] = (int)(t[i] + 0.5);
    g.nnbr.upload(nn);
    g.nnbr_global = true;
  }
  alloc0(s.U, (size_t)P * nv);
  alloc0(s.U_old, (size_t)P * nv);
  alloc0(s.V, (size_t)P * s.nPrim);
  alloc0(s.grad_prim, (size_t)P * s.nG * nd);
  s.recon_aliases_prim = (c->cfg.grad_method_recon == c->cfg.grad_method);
  if (!s.recon_aliases_prim) alloc0(s.grad_recon, (size_t)P * s.nG * nd);
  s.limiter.alloc((size_t)P * s.nG);
  {
    std::vector<double> one((size_t)P * s.nG, 1.0);  // limiters start at 1 (CFlowVariable ctor)
    s.limiter.upload(one);
  }
  alloc0(s.dt, P);
  alloc0(s.lam_inv, P);
  alloc0(s.lam_visc, P);
  alloc0(s.under_relax, P);
  if (c->cfg.cfl_adapt) {  // CVariable::LocalCFL starts at CFL_NUMBER everywhere (CVariable ctor / CSolver::SetInitialCondition)
    std::vector<double> lc(P, c->cfg.cfl);
    s.local_cfl.upload(lc);
  } else {
    s.local_cfl.release();
  }
  s.adapt_series.clear();
  s.adapt_counter = 0;
  s.adapt_old_func = s.adapt_new_func = 0.0;
  c->inner_iter = 0;
  alloc0(s.lambda, P);
  alloc0(s.sensor, P);
  alloc0(s.und_lapl, (size_t)P * nv);
  alloc0(s.res, (size_t)P * nv);
  alloc0(s.sol, (size_t)P * nv);
  alloc0(s.edge_flux, (size_t)E * nv);
  s.edge_counter.alloc(E);
  s.edge_counter.zero();
  alloc0(s.vorticity, (size_t)P * 3);
  alloc0(s.strain_mag, P);
  s.counters.alloc(8);
  s.counters.zero();
  s.red.alloc(4096);
  if (!s.h_red) SU2B_CUDA(cudaMallocHost(&s.h_red, 64 * sizeof(double)));
  if (!s.h_counters) SU2B_CUDA(cudaMallocHost(&s.h_counters, 8 * sizeof(unsigned long long)));
  linsys_bind(c, 0, nv);
  // free stream everywhere (CEulerVariable ctor, CEulerSolver.cpp:293-297)
  std::vector<double> U((size_t)P * nv);
  for (int i = 0; i < P; ++i) {
    U[(size_t)i * nv] = c->cfg.density_inf;
    for (int k = 0; k < nd; ++k) U[(size_t)i * nv + 1 + k] = c->cfg.density_inf * c->cfg.velocity_inf[k];
    U[(size_t)i * nv + nv - 1] = c->cfg.density_inf * c->cfg.energy_inf;
  }
  s.U.upload(U);
  s.U_old.upload(U);
  SU2B_REQUIRE(!(c->cfg.rans_sa && c->cfg.rans_sst), "one turbulence model at a time");
  if (c->cfg.rans_sst) {
    // CTurbSSTSolver ctor (CTurbSSTSolver.cpp:118-150): (k, omega) = Solution_Inf, muT = rho k / omega everywhere
    SU2B_REQUIRE(c->cfg.viscous, "SST needs the viscous flow solver");
    s.nTurb = 2;
    alloc0(s.nu_old, (size_t)P * 2);
    alloc0(s.nu_grad, (size_t)P * 2 * nd);
    alloc0(s.turb_res, (size_t)P * 2);
    alloc0(s.turb_sol, (size_t)P * 2);
    alloc0(s.turb_dt, P);
    alloc0(s.turb_under_relax, P);
    alloc0(s.sst_F1, P);
    alloc0(s.sst_F2, P);
    alloc0(s.sst_CDkw, P);
    std::vector<double> tk((size_t)P * 2), mut(P, c->cfg.density_inf * c->cfg.kine_inf / c->cfg.omega_inf);
    for (int i = 0; i < P; ++i) { tk[(size_t)2 * i] = c->cfg.kine_inf; tk[(size_t)2 * i + 1] = c->cfg.omega_inf; }
    s.nu.upload(tk);
    s.nu_old.upload(tk);
    s.mu_t.upload(mut);
    linsys_bind(c, 1, 2);
  }
  if (c->cfg.rans_sa) {
    s.nTurb = 1;
    alloc0(s.nu_old, P);
    alloc0(s.nu_grad, (size_t)P * nd);
    alloc0(s.turb_res, P);
    alloc0(s.turb_sol, P);
    alloc0(s.turb_dt, P);
    alloc0(s.turb_under_relax, P);
    alloc0(s.turb_edge_flux, E);
    std::vector<double> nu(P, c->cfg.nu_tilde_inf), mut(P, c->cfg.eddy_visc_inf);
    s.nu.upload(nu);
    s.nu_old.upload(nu);
    s.mu_t.upload(mut);
    linsys_bind(c, 1, 1);
  }
  SU2B_CUDA(cudaDeviceSynchronize());
  s.allocated = true;
  if (c->cfg.rans_sst) {
    // the driver's solver preprocessing runs CTurbSSTSolver::Postprocessing once on the initial state: F1, F2, CDkw and
    // the limited muT of the free stream (zero gradients, zero strain)
    flow_primitives(c);
    turb_postprocessing(c);
    SU2B_CUDA(cudaDeviceSynchronize());
  }
  API_END
}

static bool locate_array(su2b200_ctx* c, int which, double** p, size_t* n) {
  auto& s = c->fs;
  auto set = [&](dvec<double>& v) { *p = v.p; *n = v.n; return true; };
  switch (which) {
    case SU2B200_ARR_SOLUTION: return set(s.U);
    case SU2B200_ARR_SOLUTION_OLD: return set(s.U_old);
    case SU2B200_ARR_PRIMITIVE: return set(s.V);
    case SU2B200_ARR_GRAD_RECON: return set(s.recon_aliases_prim ? s.grad_prim : s.grad_recon);
    case SU2B200_ARR_GRAD_PRIM: return set(s.grad_prim);
    case SU2B200_ARR_LIMITER: return set(s.limiter);
    case SU2B200_ARR_DELTA_TIME: return set(s.dt);
    case SU2B200_ARR_LINSYSRES: return set(s.res);
    case SU2B200_ARR_LINSYSSOL: return set(s.sol);
    case SU2B200_ARR_JACOBIAN: return set(c->sys[0].A);
    case SU2B200_ARR_LAMBDA: return set(s.lambda);
    case SU2B200_ARR_SENSOR: return set(s.sensor);
    case SU2B200_ARR_UND_LAPL: return set(s.und_lapl);
    case SU2B200_ARR_MAX_LAMBDA_INV: return set(s.lam_inv);
    case SU2B200_ARR_MAX_LAMBDA_VISC: return set(s.lam_visc);
    case SU2B200_ARR_ILU: return set(c->sys[0].ilu);
    case SU2B200_ARR_INVM: return set(c->sys[0].invM);
    case SU2B200_ARR_UNDER_RELAX: return set(s.under_relax);
    case SU2B200_ARR_RES_TRUNC: return set(s.res_trunc);
    case SU2B200_ARR_LOCAL_CFL: return set(s.local_cfl);
    case SU2B200_ARR_TURB_SOLUTION: return set(s.nu);
    case SU2B200_ARR_TURB_SOLUTION_OLD: return set(s.nu_old);
    case SU2B200_ARR_TURB_GRADIENT: return set(s.nu_grad);
    case SU2B200_ARR_TURB_LINSYSRES: return set(s.turb_res);
    case SU2B200_ARR_TURB_LINSYSSOL: return set(s.turb_sol);
    case SU2B200_ARR_TURB_JACOBIAN: return set(c->sys[1].A);
    case SU2B200_ARR_TURB_DELTA_TIME: return set(s.turb_dt);
    case SU2B200_ARR_EDDY_VISC: return set(s.mu_t);
    case SU2B200_ARR_VORTICITY: return set(s.vorticity);
    case SU2B200_ARR_STRAIN_MAG: return set(s.strain_mag);
    case SU2B200_ARR_WALL_DIST: return set(c->geo.wall_dist);
    case SU2B200_ARR_SST_F1: return set(s.sst_F1);
    case SU2B200_ARR_SST_F2: return set(s.sst_F2);
    case SU2B200_ARR_SST_CDKW: return set(s.sst_CDkw);
    default: return false;
  }
}

long su2b200_array_size(su2b200_ctx* c, int which) {
  double* p;
  size_t n;
  if (!c || !locate_array(c, which, &p, &n)) return -1;
  return (long)n;
}
int su2b200_array_download(su2b200_ctx* c, int which, double* host) {
  API_BEGIN
  require_ctx(c);
  double* p;
  size_t n;
  SU2B_REQUIRE(locate_array(c, which, &p, &n), "unknown array id");
  SU2B_REQUIRE(host, "null host buffer");
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  if (n) SU2B_CUDA(cudaMemcpy(host, p, n * sizeof(double), cudaMemcpyDeviceToHost));
  if (which == SU2B200_ARR_ILU && c->sys[0].ilu_slim) {
    // slim factors (precond_factor.cu): the upper blocks were never copied out of the Jacobian
    const LinSys& ls = c->sys[0];
    SU2B_REQUIRE(ls.ilu_gen == ls.A_gen, "ILU factors are stale: the Jacobian changed after the preconditioner was built");
    const size_t bs = (size_t)ls.nVar * ls.nVar;
    std::vector<double> a = ls.A.to_host();
    for (int i = 0; i < ls.Pd; ++i)
      for (size_t k = (size_t)ls.h_dia_ptr[i] + 1; k < (size_t)ls.h_row_ptr[i + 1]; ++k)
        std::memcpy(host + k * bs, a.data() + k * bs, bs * sizeof(double));
  }
  API_END
}
int su2b200_array_upload(su2b200_ctx* c, int which, const double* host) {
  API_BEGIN
  require_ctx(c);
  double* p;
  size_t n;
  SU2B_REQUIRE(locate_array(c, which, &p, &n), "unknown array id");
  SU2B_REQUIRE(host, "null host buffer");
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  if (n) SU2B_CUDA(cudaMemcpy(p, host, n * sizeof(double), cudaMemcpyHostToDevice));
  if (which == SU2B200_ARR_JACOBIAN) ++c->sys[0].A_gen;
  if (which == SU2B200_ARR_TURB_JACOBIAN) ++c->sys[1].A_gen;
  API_END
}
// CSolver::InitiateComms + CompleteComms (CSolver.cpp:1406-1697) for one named per-point array: the values of the points
// in Local_Point_P2PSend go to the neighbour ranks and land in their Local_Point_P2PRecv points (count per point =
// array size / nPoint, the reference's COUNT_PER_POINT of CSolver.cpp:1324-1385). A single rank has nothing to exchange.
int su2b200_halo_exchange(su2b200_ctx* c, int which) {
  API_BEGIN
  require_ctx(c);
  double* p;
  size_t n;
  SU2B_REQUIRE(locate_array(c, which, &p, &n), "unknown array id");
  SU2B_REQUIRE(c->geo.uploaded && c->geo.P > 0 && n % (size_t)c->geo.P == 0 && n > 0,
               "halo exchange needs a per-point array ([nPoint][count])");
  SU2B_REQUIRE(which != SU2B200_ARR_JACOBIAN && which != SU2B200_ARR_TURB_JACOBIAN && which != SU2B200_ARR_ILU &&
                   which != SU2B200_ARR_INVM, "matrices are not exchanged");
  halo_exchange(c, p, (int)(n / (size_t)c->geo.P));
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  API_END
}
int su2b200_edge_counter_download(su2b200_ctx* c, signed char* host) {
  API_BEGIN
  require_ctx(c);
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  c->fs.edge_counter.download(host);
  API_END
}

}  // extern "C"
