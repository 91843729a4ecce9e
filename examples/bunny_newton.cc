// bunny_newton.cc -- config C2 of BASELINE.json as a host C++ driver: the call sequence of the reference's
// lib_bin/mesh_deform/test_feat10_bunny_newton.cc (read the TetGen mesh, pin z < -4, load z > 4 with -35 000 N, Setup,
// material, CalcDnDuPre, RetrieveDnDuPreToCPU / RetrieveDetJToCPU, CalcMassMatrix, CalcConstraintData,
// ConvertToCSR_ConstraintJacT, BuildConstraintJacobianCSR, CalcP, RetrievePFromFToCPU, CalcInternalForce,
// RetrieveInternalForceToCPU, solver {1e-4, 1e-6, 1e-4, 1e14, 5, 10, 1e-3}, Setup, AnalyzeHessianSparsity,
// SetFixedSparsityPattern, Solve loop, WriteOutputVTK, RetrievePositionToCPU) written against the reference's class and
// method names through include/tlfea_shim.hpp.  The Eigen types of the reference are the stand-ins of mini_matrix.hpp
// here (Eigen is absent from this image); with Eigen present the two `using` lines below are the only change.
//
//   g++ -std=c++17 -Iinclude -Iexamples examples/bunny_newton.cc -Ltotal-lagrangian-fea_b200/lib -ltlfea_b200 -o bunny
//   ./bunny <mesh.node> <mesh.ele> [svk|mr] [n_steps] [out.vtk]
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "mini_matrix.hpp"
#include "tlfea_shim.hpp"

using VectorXd = mini::VectorXd;  // Eigen::VectorXd
using VectorXi = mini::VectorXi;  // Eigen::VectorXi
using MatrixXd = mini::MatrixXd;  // Eigen::MatrixXd
using MatrixXi = mini::MatrixXi;  // Eigen::MatrixXi

// TetGen .node / .ele, 0- or 1-based ids, mid-edge nodes reordered to the element's convention
// (what ANCFCPUUtils::FEAT10_read_nodes / _elements do, lib_utils/cpu_utils.cc:607-754)
static std::vector<std::vector<double>> read_rows(const std::string &path) {
  std::ifstream in(path);
  if (!in) {
    std::fprintf(stderr, "cannot open %s\n", path.c_str());
    std::exit(EXIT_FAILURE);
  }
  std::vector<std::vector<double>> rows;
  std::string line;
  while (std::getline(in, line)) {
    if (line.empty() || line[line.find_first_not_of(" \t")] == '#') continue;
    std::istringstream ss(line);
    std::vector<double> r;
    double v;
    while (ss >> v) r.push_back(v);
    if (!r.empty()) rows.push_back(r);
  }
  return rows;
}

int main(int argc, char **argv) {
  if (argc < 3) {
    std::fprintf(stderr, "usage: %s mesh.node mesh.ele [svk|mr] [n_steps] [out.vtk]\n", argv[0]);
    return 2;
  }
  const bool mooney = argc > 3 && std::string(argv[3]) == "mr";
  const int n_steps = argc > 4 ? std::atoi(argv[4]) : 2;
  const double E = 3.0e8, nu = 0.4, rho0 = 920.0;

  auto nrows = read_rows(argv[1]), erows = read_rows(argv[2]);
  const int n_nodes = (int)nrows[0][0], n_elems = (int)erows[0][0];
  const int nbase = (int)nrows[1][0];
  MatrixXd nodes(n_nodes, 3);
  for (int i = 0; i < n_nodes; i++)
    for (int k = 0; k < 3; k++) nodes((int)nrows[1 + i][0] - nbase, k) = nrows[1 + i][1 + k];
  int ebase = 1 << 30;
  for (int e = 0; e < n_elems; e++)
    for (int a = 0; a < 10; a++) ebase = std::min(ebase, (int)erows[1 + e][1 + a]);
  ebase = ebase == 0 ? 0 : 1;
  const int order[10] = {0, 1, 2, 3, 6, 7, 9, 5, 8, 4};
  MatrixXi elements(n_elems, 10);
  for (int e = 0; e < n_elems; e++)
    for (int a = 0; a < 10; a++) elements(e, a) = (int)erows[1 + e][1 + order[a]] - ebase;

  GPU_FEAT10_Data gpu_t10_data(n_elems, n_nodes);
  gpu_t10_data.Initialize();

  VectorXd h_x12(n_nodes), h_y12(n_nodes), h_z12(n_nodes);
  for (int i = 0; i < n_nodes; i++) h_x12(i) = nodes(i, 0), h_y12(i) = nodes(i, 1), h_z12(i) = nodes(i, 2);
  std::vector<int> fixed_node_indices;
  for (int i = 0; i < n_nodes; i++)
    if (h_z12(i) < -4.0) fixed_node_indices.push_back(i);
  VectorXi h_fixed_nodes((long)fixed_node_indices.size());
  for (size_t i = 0; i < fixed_node_indices.size(); i++) h_fixed_nodes((long)i) = fixed_node_indices[i];
  gpu_t10_data.SetNodalFixed(h_fixed_nodes);

  VectorXd h_f_ext(gpu_t10_data.get_n_coef() * 3);
  h_f_ext.setZero();
  for (int i = 0; i < n_nodes; i++)
    if (h_z12(i) > 4.0) h_f_ext(3 * i + 2) = -35000.0;
  gpu_t10_data.SetExternalForce(h_f_ext);

  // Keast 5-point rule (lib_utils/quadrature_utils.h:135-158)
  VectorXd qx(5), qy(5), qz(5), qw(5);
  const double pts[5][4] = {{0.25, 0.25, 0.25, -2.0 / 15.0},
                            {1.0 / 6.0, 1.0 / 6.0, 1.0 / 6.0, 3.0 / 40.0},
                            {0.5, 1.0 / 6.0, 1.0 / 6.0, 3.0 / 40.0},
                            {1.0 / 6.0, 0.5, 1.0 / 6.0, 3.0 / 40.0},
                            {1.0 / 6.0, 1.0 / 6.0, 0.5, 3.0 / 40.0}};
  for (int q = 0; q < 5; q++) qx(q) = pts[q][0], qy(q) = pts[q][1], qz(q) = pts[q][2], qw(q) = pts[q][3];
  gpu_t10_data.Setup(qx, qy, qz, qw, h_x12, h_y12, h_z12, elements);
  gpu_t10_data.SetDensity(rho0);
  gpu_t10_data.SetDamping(0.0, 0.0);
  if (!mooney) {
    gpu_t10_data.SetSVK(E, nu);
  } else {
    const double mu = E / (2.0 * (1.0 + nu)), K = E / (3.0 * (1.0 - 2.0 * nu));
    gpu_t10_data.SetMooneyRivlin(0.30 * mu, 0.20 * mu, 1.5 * K);
  }
  gpu_t10_data.CalcDnDuPre();
  std::vector<std::vector<MatrixXd>> ref_grads;
  gpu_t10_data.RetrieveDnDuPreToCPU(ref_grads);
  std::vector<std::vector<double>> detJ;
  gpu_t10_data.RetrieveDetJToCPU(detJ);
  double vol = 0.0;
  for (auto &e : detJ)
    for (int q = 0; q < 5; q++) vol += e[(size_t)q] * qw(q);
  std::printf("elements %d nodes %d fixed %ld volume %.12g gradN(0,0)(0,0) %.12g\n", n_elems, n_nodes,
              h_fixed_nodes.size(), vol, ref_grads[0][0](0, 0));

  gpu_t10_data.CalcMassMatrix();
  gpu_t10_data.CalcConstraintData();
  gpu_t10_data.ConvertToCSR_ConstraintJacT();
  gpu_t10_data.BuildConstraintJacobianCSR();
  gpu_t10_data.CalcP();
  std::vector<std::vector<MatrixXd>> p_from_F;
  gpu_t10_data.RetrievePFromFToCPU(p_from_F);
  gpu_t10_data.CalcInternalForce();
  VectorXd f_int;
  gpu_t10_data.RetrieveInternalForceToCPU(f_int);
  double fmax = 0.0, pmax = 0.0;
  for (long i = 0; i < f_int.size(); i++) fmax = std::fmax(fmax, std::fabs(f_int(i)));
  for (auto &e : p_from_F)
    for (auto &m : e)
      for (long i = 0; i < m.size(); i++) pmax = std::fmax(pmax, std::fabs(m.data()[i]));
  std::printf("rest state: max|P| %.3e max|f_int| %.3e\n", pmax, fmax);
  MatrixXd J;
  gpu_t10_data.RetrieveConstraintJacobianToCPU(J);
  MatrixXi conn_back;
  gpu_t10_data.RetrieveConnectivityToCPU(conn_back);
  std::printf("J %ld x %ld, connectivity round trip %s\n", J.rows(), J.cols(),
              conn_back(n_elems - 1, 9) == elements(n_elems - 1, 9) ? "ok" : "MISMATCH");

  SyncedNewtonParams params = {1e-4, 1e-6, 1e-4, 1e14, 5, 10, 1e-3};
  SyncedNewtonSolver solver(&gpu_t10_data, gpu_t10_data.get_n_constraint());
  solver.Setup();
  solver.SetParameters(&params);
  solver.AnalyzeHessianSparsity();
  solver.SetFixedSparsityPattern(true);
  solver.SetPCG(50000, 1e-13);  // stands in for the cuDSS solve of the reference

  VectorXd x12, y12, z12;
  for (int i = 0; i < n_steps; i++) {
    solver.Solve();
    gpu_t10_data.RetrievePositionToCPU(x12, y12, z12);
    double zmin = 1e300;
    for (long k = 0; k < z12.size(); k++) zmin = std::fmin(zmin, z12(k));
    std::printf("step %d z_top %.17g z_min %.17g\n", i, z12(n_nodes - 1), zmin);
  }
  if (argc > 5) gpu_t10_data.WriteOutputVTK(argv[5]);
  for (long k = 0; k < x12.size(); k++) std::printf("pos %ld %.17g %.17g %.17g\n", k, x12(k), y12(k), z12(k));
  gpu_t10_data.Destroy();
  return 0;
}
