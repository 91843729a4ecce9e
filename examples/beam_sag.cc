// beam_sag.cc -- the reference's beam_sag scenario (lib_bin/beam_sag/test_ancf3243.cc: ANCF3243 cantilever, 30 elements of
// L = 0.5, W = H = 0.1, E = 7e8, nu = 0.33, rho0 = 2700, Kelvin-Voigt 1e5 / 1e5, tip load Fz = 3100, dt = 1e-3, Newton
// parameters {1e-4, 0, 1e-6, 1e14, 5, 10, dt}, the four coefficients of node 0 pinned) written against the reference's
// class names through include/tlfea_shim.hpp -- what a lib_bin driver looks like after the switch.  Host C++ only;
// every kernel runs inside libtlfea_b200.so.
//
//   g++ -std=c++17 -Iinclude examples/beam_sag.cc -Ltotal-lagrangian-fea_b200/lib -ltlfea_b200 -o beam_sag
//   ./beam_sag [n_steps]        prints "step k tip_z <z of the tip node>"
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "tlfea_shim.hpp"

int main(int argc, char **argv) {
  const int n_steps = argc > 1 ? std::atoi(argv[1]) : 5;
  const int n_elem = 30, n_nodes = n_elem + 1, n_coef = 4 * n_nodes;
  const double L = 0.5, W = 0.1, H = 0.1, dt = 1e-3;

  // straight beam along x: per node [r, r_u = e_x, r_v = e_y, r_w = e_z]
  std::vector<double> x(n_coef, 0.0), y(n_coef, 0.0), z(n_coef, 0.0);
  for (int i = 0; i < n_nodes; i++) {
    x[4 * i] = i * L;
    x[4 * i + 1] = 1.0;
    y[4 * i + 2] = 1.0;
    z[4 * i + 3] = 1.0;
  }
  std::vector<int> conn;
  for (int e = 0; e < n_elem; e++) {
    conn.push_back(e);
    conn.push_back(e + 1);
  }
  std::vector<double> f_ext(3 * (size_t)n_coef, 0.0);
  f_ext[3 * (4 * n_elem) + 2] = 3100.0;  // z component on the position coefficient of the tip node

  GPU_ANCF3243_Data beam(n_nodes, n_elem);
  beam.Initialize();
  beam.SetExternalForce(f_ext);
  beam.Setup(L, W, H, x, y, z, conn);
  beam.SetDensity(2700.0);
  beam.SetDamping(1e5, 1e5);
  beam.SetSVK(7e8, 0.33);
  beam.SetNodalFixed({0, 1, 2, 3});
  beam.CalcDsDuPre();
  beam.CalcMassMatrix();
  beam.CalcConstraintData();

  SyncedNewtonSolver solver(&beam, 12);
  SyncedNewtonParams prm{1e-4, 0.0, 1e-6, 1e14, 5, 10, dt};
  solver.SetParameters(&prm);
  solver.Setup();
  solver.AnalyzeHessianSparsity();
  solver.SetFixedSparsityPattern(true);
  solver.SetPCG(20000, 1e-13);  // stands in for the cuDSS solve of the reference

  std::vector<double> px, py, pz;
  for (int k = 0; k < n_steps; k++) {
    solver.Solve();
    beam.RetrievePositionToCPU(px, py, pz);
    std::printf("step %d tip_z %.17g\n", k, pz[4 * n_elem]);
  }
  return 0;
}
