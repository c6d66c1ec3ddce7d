// f3. Restart / solution I/O boundary: the SU2 native restart formats, read and written from the device state.
//   binary ".dat" : CSU2BinaryFileWriter::WriteData (SU2_CFD/src/output/filewriter/CSU2BinaryFileWriter.cpp:39-94) /
//                   CSolver::Read_SU2_Restart_Binary (SU2_CFD/src/solvers/CSolver.cpp:2916-3145): 5 ints {535532 ("SU2"),
//                   nFields, nPointGlobal, 0, 0}, nFields names of 33 chars, then [nPointGlobal][nFields] doubles by
//                   GLOBAL point index;
//   ASCII ".csv"  : CSU2FileWriter / CSolver::Read_SU2_Restart_ASCII (:2754-2914): "PointID","x",...,"Energy" header, one
//                   row per global point.
// The reader takes what CFVMFlowSolverBase::LoadRestart_impl (CFVMFlowSolverBase.inl:824-1008) takes: the nDim coordinate
// columns are skipped, the next nVar columns are the conservative solution, then (RANS) the turbulence variable
// (CTurbSASolver::LoadRestart: skipVars = nDim + nVar of the flow solver). Points are matched by global index, so the
// file is independent of the RCM permutation and of the partitioning: every rank reads / writes the rows of its own
// points (binary: at their offsets of ONE shared file, rank 0 writes the header -- a parallel checkpoint without MPI-IO).
// Plain host C++ apart from the two copies to / from the device arrays.
#include "internal.hpp"
#include "kernels.hpp"
#include <algorithm>
#include <cerrno>
#include <cmath>
#include <fstream>
#include <sstream>

using namespace su2b200;
typedef unsigned long ulong_t;

namespace {
constexpr int SU2_MAGIC = 535532;
constexpr int NAME_LEN = 33;  // CGNS_STRING_SIZE

std::vector<std::string> field_names(int nDim, bool turb) {
  std::vector<std::string> f = {"x", "y"};
  if (nDim == 3) f.push_back("z");
  f.push_back("Density");
  f.push_back("Momentum_x");
  f.push_back("Momentum_y");
  if (nDim == 3) f.push_back("Momentum_z");
  f.push_back("Energy");
  if (turb) f.push_back("Nu_Tilde");
  return f;
}

std::string trim_quotes(std::string s) {
  const char* ws = " \t\r\n\"";
  const size_t b = s.find_first_not_of(ws), e = s.find_last_not_of(ws);
  return b == std::string::npos ? std::string() : s.substr(b, e - b + 1);
}

// rows [nPoint][nFields] of the owned points in the file's column order
void write_restart(const char* path, bool binary, const std::vector<std::string>& names, const std::vector<double>& rows,
                   const ulong_t* gid, ulong_t nLocal, ulong_t nGlobal, bool header) {
  const size_t nF = names.size();
  if (binary) {
    FILE* f = std::fopen(path, header ? "wb" : "r+b");
    if (!f && !header) f = std::fopen(path, "wb");
    if (!f) throw Error(std::string("cannot open restart file ") + path + " for writing: " + std::strerror(errno));
    const size_t head = 5 * sizeof(int) + nF * NAME_LEN;
    if (header) {
      const int var_buf[5] = {SU2_MAGIC, (int)nF, (int)nGlobal, 0, 0};
      std::fwrite(var_buf, sizeof(int), 5, f);
      for (const auto& n : names) {
        char buf[NAME_LEN] = {0};
        std::strncpy(buf, n.c_str(), NAME_LEN - 1);
        std::fwrite(buf, 1, NAME_LEN, f);
      }
    }
    // consecutive global indices go out in one write
    ulong_t k = 0;
    while (k < nLocal) {
      ulong_t e = k + 1;
      while (e < nLocal && gid[e] == gid[e - 1] + 1) ++e;
      if (std::fseek(f, (long)(head + gid[k] * nF * sizeof(double)), SEEK_SET) != 0)
        throw Error("seek failed in restart file");
      std::fwrite(rows.data() + k * nF, sizeof(double), (e - k) * nF, f);
      k = e;
    }
    std::fclose(f);
    return;
  }
  SU2B_REQUIRE(nLocal == nGlobal, "ASCII restart files are written by a single rank (use the binary format in parallel)");
  std::vector<ulong_t> order(nLocal);
  for (ulong_t i = 0; i < nLocal; ++i) order[i] = i;
  std::sort(order.begin(), order.end(), [&](ulong_t a, ulong_t b) { return gid[a] < gid[b]; });
  FILE* f = std::fopen(path, "w");
  if (!f) throw Error(std::string("cannot open restart file ") + path + " for writing");
  std::fprintf(f, "\"PointID\"");
  for (const auto& n : names) std::fprintf(f, ",\"%s\"", n.c_str());
  std::fprintf(f, "\n");
  for (ulong_t q = 0; q < nLocal; ++q) {
    const ulong_t i = order[q];
    std::fprintf(f, "%lu", gid[i]);
    for (size_t v = 0; v < nF; ++v) std::fprintf(f, ", %.15e", rows[i * nF + v]);
    std::fprintf(f, "\n");
  }
  std::fclose(f);
}

// fills rows[nLocal][nWanted] with the columns [skip, skip + nWanted) of the file for the given global indices
void read_restart(const char* path, const ulong_t* gid, ulong_t nLocal, int skip, int nWanted, std::vector<double>& rows,
                  std::vector<std::string>* names_out) {
  FILE* f = std::fopen(path, "rb");
  if (!f) throw Error(std::string("Unable to open SU2 restart file ") + path);
  int var_buf[5] = {0, 0, 0, 0, 0};
  const size_t got = std::fread(var_buf, sizeof(int), 5, f);
  rows.assign((size_t)nLocal * nWanted, 0.0);
  if (got == 5 && var_buf[0] == SU2_MAGIC) {
    const ulong_t nF = (ulong_t)var_buf[1], nP = (ulong_t)var_buf[2];
    SU2B_REQUIRE((int)nF >= skip + nWanted, "restart file holds fewer fields than the solver needs");
    std::vector<std::string> names(nF);
    for (ulong_t k = 0; k < nF; ++k) {
      char buf[NAME_LEN + 1] = {0};
      if (std::fread(buf, 1, NAME_LEN, f) != (size_t)NAME_LEN) { std::fclose(f); throw Error("Error reading restart file."); }
      names[k] = buf;
    }
    if (names_out) *names_out = names;
    const size_t head = 5 * sizeof(int) + nF * NAME_LEN;
    std::vector<double> line(nF);
    for (ulong_t i = 0; i < nLocal; ++i) {
      if (gid[i] >= nP) { std::fclose(f); throw Error("The solution file does not match the mesh (point index beyond the file)."); }
      std::fseek(f, (long)(head + gid[i] * nF * sizeof(double)), SEEK_SET);
      if (std::fread(line.data(), sizeof(double), nF, f) != nF) { std::fclose(f); throw Error("Error reading restart file."); }
      for (int v = 0; v < nWanted; ++v) rows[i * nWanted + v] = line[skip + v];
    }
    std::fclose(f);
    return;
  }
  std::fclose(f);
  // CSV
  std::ifstream in(path);
  std::string text;
  if (!std::getline(in, text)) throw Error("Restart file is empty.");
  std::vector<std::string> fields;
  {
    std::stringstream ss(text);
    std::string tok;
    while (std::getline(ss, tok, ',')) fields.push_back(trim_quotes(tok));
  }
  SU2B_REQUIRE(fields.size() > 1, "Restart file does not seem to be a CSV file.");
  const int nF = (int)fields.size() - 1;  // without PointID
  SU2B_REQUIRE(nF >= skip + nWanted, "restart file holds fewer fields than the solver needs");
  if (names_out) names_out->assign(fields.begin() + 1, fields.end());
  ulong_t maxg = 0;
  for (ulong_t i = 0; i < nLocal; ++i) maxg = std::max(maxg, gid[i]);
  std::vector<long> local(maxg + 1, -1);
  for (ulong_t i = 0; i < nLocal; ++i) local[gid[i]] = (long)i;
  ulong_t found = 0, row = 0;
  while (std::getline(in, text)) {
    if (text.find_first_not_of(" \t\r\n") == std::string::npos) continue;
    if (row <= maxg && local[row] >= 0) {
      std::stringstream ss(text);
      std::string tok;
      std::getline(ss, tok, ',');  // PointID
      for (int v = 0; v < skip + nWanted; ++v) {
        if (!std::getline(ss, tok, ',')) throw Error("short line in the CSV restart file");
        if (v >= skip) rows[(size_t)local[row] * nWanted + (v - skip)] = std::stod(tok);
      }
      ++found;
    }
    ++row;
  }
  if (found != nLocal) throw Error("The solution file does not match the mesh, currently only binary files can be interpolated.");
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

int su2b200_restart_write_arrays(const char* path, int binary, int nDim, int nTurb, unsigned long nPoint,
                                 const unsigned long* global_index, unsigned long nPointGlobal, const double* coord,
                                 const double* solution, const double* turb, int write_header) {
  API_BEGIN
  SU2B_REQUIRE(path && coord && solution && (nDim == 2 || nDim == 3), "invalid argument");
  SU2B_REQUIRE(nTurb == 0 || turb, "turbulence variable missing");
  const int nv = nDim + 2;
  const auto names = field_names(nDim, nTurb > 0);
  const size_t nF = names.size();
  std::vector<double> rows((size_t)nPoint * nF);
  std::vector<ulong_t> gid(nPoint);
  for (ulong_t i = 0; i < nPoint; ++i) {
    gid[i] = global_index ? global_index[i] : i;
    double* r = &rows[i * nF];
    for (int d = 0; d < nDim; ++d) r[d] = coord[i * nDim + d];
    for (int v = 0; v < nv; ++v) r[nDim + v] = solution[i * nv + v];
    if (nTurb) r[nDim + nv] = turb[i];
  }
  write_restart(path, binary != 0, names, rows, gid.data(), nPoint, nPointGlobal, write_header != 0);
  API_END
}

int su2b200_restart_read_arrays(const char* path, int nDim, int nTurb, unsigned long nPoint,
                                const unsigned long* global_index, double* solution, double* turb) {
  API_BEGIN
  SU2B_REQUIRE(path && solution && (nDim == 2 || nDim == 3), "invalid argument");
  const int nv = nDim + 2, want = nv + (nTurb ? 1 : 0);
  std::vector<ulong_t> gid(nPoint);
  for (ulong_t i = 0; i < nPoint; ++i) gid[i] = global_index ? global_index[i] : i;
  std::vector<double> rows;
  read_restart(path, gid.data(), nPoint, nDim, want, rows, nullptr);
  for (ulong_t i = 0; i < nPoint; ++i) {
    for (int v = 0; v < nv; ++v) solution[i * nv + v] = rows[i * want + v];
    if (nTurb && turb) turb[i] = rows[i * want + nv];
  }
  API_END
}

// COutput::SetResultFiles for the RESTART / RESTART_ASCII file of this rank's owned points
int su2b200_restart_write(su2b200_ctx* c, const char* path, int binary, const unsigned long* global_index,
                          unsigned long nPointGlobal) {
  API_BEGIN
  if (!c) throw Error("null context");
  SU2B_CUDA(cudaSetDevice(c->device));
  SU2B_REQUIRE(c->fs.allocated, "call su2b200_flow_allocate first");
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  auto& g = c->geo;
  const ulong_t Pd = (ulong_t)g.Pd;
  const int nd = g.nDim;
  std::vector<double> coord = g.coord.to_host(), U = c->fs.U.to_host(), nu;
  const int nTurb = c->cfg.rans_sa ? 1 : 0;
  SU2B_REQUIRE(!c->cfg.rans_sst, "restart files of SST runs are not supported");
  if (nTurb) nu = c->fs.nu.to_host();
  const int rc = su2b200_restart_write_arrays(path, binary, nd, nTurb, Pd, global_index, nPointGlobal ? nPointGlobal : Pd,
                                              coord.data(), U.data(), nTurb ? nu.data() : nullptr,
                                              c->rank == 0 ? 1 : 0);
  if (rc) throw Error(su2b200_last_error());
  API_END
}

// CFVMFlowSolverBase::LoadRestart_impl (+ CTurbSASolver::LoadRestart): Solution (and nu~) of every local point, halos
// included, straight from the file by global index; Solution_Old follows (Set_OldSolution)
int su2b200_restart_read(su2b200_ctx* c, const char* path, const unsigned long* global_index) {
  API_BEGIN
  if (!c) throw Error("null context");
  SU2B_CUDA(cudaSetDevice(c->device));
  SU2B_REQUIRE(c->fs.allocated, "call su2b200_flow_allocate first");
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  auto& g = c->geo;
  const ulong_t P = (ulong_t)g.P;
  const int nd = g.nDim, nv = nd + 2;
  const int nTurb = c->cfg.rans_sa ? 1 : 0;
  SU2B_REQUIRE(!c->cfg.rans_sst, "restart files of SST runs are not supported");
  std::vector<double> U((size_t)P * nv), nu(nTurb ? P : 0);
  const int rc = su2b200_restart_read_arrays(path, nd, nTurb, P, global_index, U.data(), nTurb ? nu.data() : nullptr);
  if (rc) throw Error(su2b200_last_error());
  SU2B_CUDA(cudaMemcpy(c->fs.U.p, U.data(), U.size() * sizeof(double), cudaMemcpyHostToDevice));
  SU2B_CUDA(cudaMemcpy(c->fs.U_old.p, U.data(), U.size() * sizeof(double), cudaMemcpyHostToDevice));
  if (nTurb) {
    SU2B_CUDA(cudaMemcpy(c->fs.nu.p, nu.data(), nu.size() * sizeof(double), cudaMemcpyHostToDevice));
    SU2B_CUDA(cudaMemcpy(c->fs.nu_old.p, nu.data(), nu.size() * sizeof(double), cudaMemcpyHostToDevice));
  }
  // what LoadRestart does after filling the containers: flow Preprocessing (primitives), turbulence Postprocessing (mu_t)
  flow_primitives(c);
  if (nTurb) turb_postprocessing(c);
  SU2B_CUDA(cudaStreamSynchronize(c->stream));
  API_END
}

}  // extern "C"
