// two_rank_assemble.cc -- the partitioned FEAT10 assembly driven from host C++ only: one process per GPU, each with the
// reference's class names (include/tlfea_shim.hpp) over its x-slab of a Kuhn tet block, the multi-GPU data plane of the
// library (tlfea_dist_*: NCCL communicator from a 128-byte id, halo sum over NVLink peer memory from 64-byte IPC
// handles) and NOTHING else -- the few bytes the ranks must hand each other travel through files in a directory they
// share (an MPI host would broadcast them).  New design: the reference is single-GPU.
//
//   g++ -std=c++17 -Iinclude examples/two_rank_assemble.cc -Ltotal-lagrangian-fea_b200/lib -ltlfea_b200 -o two_rank
//   ./two_rank <rank> <world> <exchange dir> [cells per edge = 8]      (one process per rank, device = rank)
// prints "rank r owned_f2 <sum of f_int^2 over the DOFs the rank owns> owned_fsum <sum of f_int>": the sums over the
// ranks must not depend on the partition (tests/test_dist_newton_gpu.py compares world = 2 with world = 1).
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <string>
#include <thread>
#include <vector>

#include "tlfea_shim.hpp"

namespace {

// ---- exchange of small blobs through files (write to a temporary name, then rename: readers never see half a file)
void put_file(const std::string &dir, const std::string &name, const void *data, size_t n) {
  const std::string tmp = dir + "/." + name + ".tmp", fin = dir + "/" + name;
  {
    std::ofstream f(tmp, std::ios::binary);
    f.write(static_cast<const char *>(data), (std::streamsize)n);
  }
  std::rename(tmp.c_str(), fin.c_str());
}
bool get_file(const std::string &dir, const std::string &name, void *data, size_t n, double timeout_s = 120.0) {
  const std::string fin = dir + "/" + name;
  const auto t0 = std::chrono::steady_clock::now();
  for (;;) {
    std::ifstream f(fin, std::ios::binary);
    if (f && f.read(static_cast<char *>(data), (std::streamsize)n)) return true;
    if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > timeout_s) return false;
    std::this_thread::sleep_for(std::chrono::milliseconds(20));
  }
}

// deterministic perturbation in (-1, 1) of component c of the lattice node gid: both holders of a shared node agree
double jitter(uint64_t gid, int c) {
  uint64_t z = gid * 3 + (uint64_t)c + 0x9e3779b97f4a7c15ull;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  z ^= z >> 31;
  return (double)(z >> 11) / 9007199254740992.0 * 2.0 - 1.0;
}

}  // namespace

int main(int argc, char **argv) {
  if (argc < 4) {
    std::fprintf(stderr, "usage: %s <rank> <world> <exchange dir> [cells per edge]\n", argv[0]);
    return 2;
  }
  const int rank = std::atoi(argv[1]), world = std::atoi(argv[2]);
  const std::string dir = argv[3];
  const int n = argc > 4 ? std::atoi(argv[4]) : 8;
  if (world < 1 || rank < 0 || rank >= world || n < world) {
    std::fprintf(stderr, "bad rank / world / n\n");
    return 2;
  }

  // ---- the rank's slab of the Kuhn block: whole cell layers [l0, l1) along x, six tets per cell (one per permutation
  // of the axes), nodes on the (2n+1)^3 half-integer lattice, local id ((i - 2 l0) N + j) N + k.  The first lattice
  // plane is shared with rank - 1, the last one with rank + 1: contiguous id ranges, as tlfea_dist_set_halo wants them.
  const int l0 = (int)((int64_t)n * rank / world), l1 = (int)((int64_t)n * (rank + 1) / world);
  const int N = 2 * n + 1, ni = 2 * (l1 - l0) + 1;
  const int nn = ni * N * N, ne = 6 * (l1 - l0) * n * n, pn = N * N;
  const bool has_left = rank > 0, has_right = rank + 1 < world;
  std::vector<double> X(nn), Y(nn), Z(nn), x(nn), y(nn), z(nn);
  const double amp = 0.05 / n;
  for (int i = 0; i < ni; i++)
    for (int j = 0; j < N; j++)
      for (int k = 0; k < N; k++) {
        const int id = (i * N + j) * N + k;
        const uint64_t gid = ((uint64_t)(i + 2 * l0) * N + j) * N + k;
        X[id] = (i + 2 * l0) / (2.0 * n), Y[id] = j / (2.0 * n), Z[id] = k / (2.0 * n);
        x[id] = X[id] + amp * jitter(gid, 0), y[id] = Y[id] + amp * jitter(gid, 1), z[id] = Z[id] + amp * jitter(gid, 2);
      }
  static const int perms[6][3] = {{0, 1, 2}, {0, 2, 1}, {1, 0, 2}, {1, 2, 0}, {2, 0, 1}, {2, 1, 0}};
  static const int edges[6][2] = {{0, 1}, {1, 2}, {0, 2}, {0, 3}, {1, 3}, {2, 3}};
  std::vector<int> conn((size_t)10 * ne);  // column-major [10][ne], like the reference's Eigen::MatrixXi
  int e = 0;
  for (int ci = 0; ci < l1 - l0; ci++)
    for (int cj = 0; cj < n; cj++)
      for (int ck = 0; ck < n; ck++)
        for (int t = 0; t < 6; t++, e++) {
          int v[4][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
          for (int s = 0; s < 3; s++) {
            for (int c = 0; c < 3; c++) v[s + 1][c] = v[s][c];
            v[s + 1][perms[t][s]] += 2;
          }
          int a[3][3];
          for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) a[r][c] = v[r + 1][c] - v[0][c];
          const int det = a[0][0] * (a[1][1] * a[2][2] - a[1][2] * a[2][1]) - a[0][1] * (a[1][0] * a[2][2] - a[1][2] * a[2][0]) +
                          a[0][2] * (a[1][0] * a[2][1] - a[1][1] * a[2][0]);
          if (det < 0)  // odd permutations are mirror images: swap two vertices so that detJ > 0
            for (int c = 0; c < 3; c++) std::swap(v[2][c], v[3][c]);
          int loc[10][3];
          for (int q = 0; q < 4; q++)
            for (int c = 0; c < 3; c++) loc[q][c] = v[q][c];
          for (int q = 0; q < 6; q++)
            for (int c = 0; c < 3; c++) loc[4 + q][c] = (v[edges[q][0]][c] + v[edges[q][1]][c]) / 2;
          for (int q = 0; q < 10; q++)
            conn[(size_t)q * ne + e] = ((2 * ci + loc[q][0]) * N + 2 * cj + loc[q][1]) * N + 2 * ck + loc[q][2];
        }

  // ---- element and solver handles on device `rank`, with the reference's call sequence
  GPU_FEAT10_Data data(ne, nn, rank);
  data.Initialize();
  const double b = 1.0 / 6.0;
  const std::vector<double> qx = {0.25, b, 0.5, b, b}, qy = {0.25, b, b, 0.5, b}, qz = {0.25, b, b, b, 0.5},
                            qw = {-4.0 / 30.0, 9.0 / 120.0, 9.0 / 120.0, 9.0 / 120.0, 9.0 / 120.0};  // Keast 5-point rule
  data.Setup(qx, qy, qz, qw, X, Y, Z, conn);
  data.SetDensity(2700.0);
  data.SetSVK(7e8, 0.33);
  data.CalcDnDuPre();
  data.CalcMassMatrix();
  SyncedNewtonSolver solver(&data, 0);
  SyncedNewtonParams prm{1e-4, 0.0, 1e-6, 1e14, 1, 1, 1e-3};
  solver.SetParameters(&prm);
  solver.Setup();
  solver.AnalyzeHessianSparsity();
  TLFEA_OR_DIE(tlfea_elem_update_positions(data.d_data, x.data(), y.data(), z.data(), nn));

  // ---- multi-GPU data plane: communicator, shared id ranges, peer-memory mailboxes
  bool p2p = false;
  if (world > 1) {
    unsigned char id[128];
    if (rank == 0) {
      TLFEA_OR_DIE(tlfea_dist_unique_id(id));
      put_file(dir, "nccl_id", id, sizeof(id));
    } else if (!get_file(dir, "nccl_id", id, sizeof(id))) {
      std::fprintf(stderr, "rank %d: no nccl_id in %s\n", rank, dir.c_str());
      return 1;
    }
    TLFEA_OR_DIE(tlfea_dist_init(solver.s, id, rank, world));
    std::vector<int> nbr;
    std::vector<int64_t> lo, hi;
    if (has_left) nbr.push_back(rank - 1), lo.push_back(0), hi.push_back(pn);
    if (has_right) nbr.push_back(rank + 1), lo.push_back(nn - pn), hi.push_back(nn);
    TLFEA_OR_DIE(tlfea_dist_set_halo(solver.s, (int)nbr.size(), nbr.data(), lo.data(), hi.data()));
    // every rank exports the mailbox each neighbour writes into and opens the neighbours' mailboxes for itself; if a
    // rank cannot, the flag files say so and everybody stays on the NCCL send/recv path
    bool ok = true;
    for (size_t k = 0; k < nbr.size() && ok; k++) {
      unsigned char h[64];
      ok = tlfea_dist_p2p_export(solver.s, (int)k, h) == TLFEA_OK;
      if (ok) put_file(dir, "mail_" + std::to_string(rank) + "_from_" + std::to_string(nbr[k]), h, sizeof(h));
    }
    for (size_t k = 0; k < nbr.size() && ok; k++) {
      unsigned char h[64];
      ok = get_file(dir, "mail_" + std::to_string(nbr[k]) + "_from_" + std::to_string(rank), h, sizeof(h), 20.0) &&
           tlfea_dist_p2p_connect(solver.s, (int)k, h) == TLFEA_OK;
    }
    const unsigned char flag = ok ? 1 : 0;
    put_file(dir, "p2p_ok_" + std::to_string(rank), &flag, 1);
    p2p = true;
    for (int r = 0; r < world; r++) {
      unsigned char f = 0;
      p2p = get_file(dir, "p2p_ok_" + std::to_string(r), &f, 1) && f && p2p;
    }
    if (!p2p) TLFEA_OR_DIE(tlfea_dist_set_halo(solver.s, (int)nbr.size(), nbr.data(), lo.data(), hi.data()));  // drops the mailboxes
  }

  // ---- one assemble of the slab, interface entries of f_int completed with the neighbours
  TLFEA_OR_DIE(tlfea_newton_assemble(solver.s));
  if (world > 1) TLFEA_OR_DIE(tlfea_dist_halo_sum(solver.s, tlfea_elem_devptr_f_int(data.d_data), 3));
  std::vector<double> f(3 * (size_t)nn);
  TLFEA_OR_DIE(tlfea_elem_get_internal_force(data.d_data, f.data(), 3 * nn));
  // every shared plane is counted once: a rank with a left neighbour skips its first plane (the neighbour owns it)
  double f2 = 0.0, fs = 0.0;
  for (size_t i = has_left ? 3 * (size_t)pn : 0; i < f.size(); i++) f2 += f[i] * f[i], fs += f[i];
  std::printf("rank %d of %d halo %s elements %d nodes %d owned_f2 %.17g owned_fsum %.17g\n", rank, world,
              world > 1 ? (p2p ? "peer-memory" : "nccl") : "none", ne, nn, f2, fs);
  if (world > 1) TLFEA_OR_DIE(tlfea_dist_finalize(solver.s));
  return 0;
}
