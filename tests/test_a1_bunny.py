"""Config C2 (BASELINE.json configs[1]): FEAT10 mesh_deform on data/meshes/T10/bunny_ascii_26 (1 066 tets, 2 095 nodes),
set up exactly as lib_bin/mesh_deform/test_feat10_bunny_newton.cc:26-28,37-40,70-126,202-208: E = 3e8, nu = 0.4,
rho0 = 920, nodes with z < -4 pinned, -35 000 N in z on nodes with z > 4, dt = 1e-3, Newton parameters
{1e-4, 1e-6, 1e-4, 1e14, 5, 10, 1e-3}, SVK or Mooney-Rivlin (0.3 mu, 0.2 mu, 1.5 K).  The mesh files are the
reference's own test data (copied to tests/golden/meshes).  An unstructured TetGen mesh is the input the structured
Kuhn block cannot stand in for: irregular valence, ragged patches, arbitrary numbering."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, rel_err

NODE = os.path.join(GOLDEN, "meshes", "bunny_ascii_26.1.node")
ELE = os.path.join(GOLDEN, "meshes", "bunny_ascii_26.1.ele")
E_MOD, NU, RHO, DT, ALM = 3e8, 0.4, 920.0, 1e-3, 1e14
PRM = (1e-4, 1e-6, 1e-4, ALM, 5, 10, DT)


def _mat(orc, model):
    if model == "svk":
        return orc.Material.svk(E_MOD, NU)
    mu, K = E_MOD / (2 * (1 + NU)), E_MOD / (3 * (1 - 2 * NU))
    return orc.Material.mooney_rivlin(0.3 * mu, 0.2 * mu, 1.5 * K)


def _load():
    from tlfea import meshgen

    X, E = meshgen.read_tetgen(NODE, ELE)
    fixed = np.nonzero(X[:, 2] < -4.0)[0].astype(np.int32)
    f_ext = np.zeros(3 * X.shape[0])
    f_ext[3 * np.nonzero(X[:, 2] > 4.0)[0] + 2] = -35000.0
    return X, E, fixed, f_ext


def test_tetgen_reader_on_the_reference_mesh():
    """FEAT10_read_nodes / FEAT10_read_elements (cpu_utils.cc:607-754): 0-based ids detected, mid-edge nodes
    reordered [0,1,2,3,6,7,9,5,8,4]: every mid-edge node then sits at the midpoint of its edge in the reference's
    local order (01),(12),(02),(03),(13),(23), and all elements have positive volume."""
    X, E, fixed, f_ext = _load()
    assert X.shape == (2095, 3) and E.shape == (1066, 10) and E.min() == 0 and E.max() == 2094
    raw = np.loadtxt(ELE, skiprows=1, dtype=np.int64, comments="#")[:, 1:11]
    assert np.array_equal(E, raw[:, [0, 1, 2, 3, 6, 7, 9, 5, 8, 4]])
    edges = [(0, 1), (1, 2), (0, 2), (0, 3), (1, 3), (2, 3)]
    for k, (a, b) in enumerate(edges):
        mid = 0.5 * (X[E[:, a]] + X[E[:, b]])
        assert np.abs(X[E[:, 4 + k]] - mid).max() < 1e-6 * np.abs(X).max()
    v = X[E[:, 1:4]] - X[E[:, 0]][:, None, :]
    assert (np.linalg.det(v) > 0).all()
    assert len(fixed) > 0 and (f_ext != 0).sum() > 0


def test_patch_plan_on_the_bunny():
    """The single-pass assembly plan on an unstructured mesh: every element once, colours conflict-free, the replayed
    summation order reproduces the oracle's H (tests/test_patch_plan.py machinery)."""
    from oracle import oracle as orc
    from test_patch_plan import build_plan, check_plan_invariants, mass_lookup, replay

    X, E, fixed, f_ext = _load()
    o = orc.T10Oracle(E, X)
    mat = _mat(orc, "svk")
    rng = np.random.default_rng(5)
    x = X + rng.normal(0, 0.01, X.shape)
    for W in (4, 8):
        plan = build_plan(E, X, o.off, o.cols, W)
        check_plan_invariants(plan, E, W)
        assert plan["uncovered"].size == 0
    Ke = [DT * orc.element_tangent(E[e], x, o.grad_N[e], o.detJ[e], o.qw, mat) for e in range(E.shape[0])]
    H = replay(plan, Ke, mass_lookup(o.off, o.cols, o.mass(RHO)), DT, o.cols.shape[0])
    assert rel_err(H, o.hessian(x, mat, DT, RHO)) < 1e-13


def _make(model, X, E, fixed, f_ext):
    import tlfea

    d = tlfea.GPU_FEAT10_Data(E.shape[0], X.shape[0])
    d.Initialize()
    d.SetNodalFixed(fixed)
    d.SetExternalForce(f_ext)
    d.Setup(*tlfea.tet5pt_quadrature(), X[:, 0], X[:, 1], X[:, 2], E)
    d.SetDensity(RHO)
    d.SetDamping(0.0, 0.0)
    if model == "svk":
        d.SetSVK(E_MOD, NU)
    else:
        mu, K = E_MOD / (2 * (1 + NU)), E_MOD / (3 * (1 - 2 * NU))
        d.SetMooneyRivlin(0.3 * mu, 0.2 * mu, 1.5 * K)
    d.CalcDnDuPre()
    d.CalcMassMatrix()
    d.CalcConstraintData()
    s = tlfea.SyncedNewtonSolver(d, 3 * len(fixed))
    s.Setup()
    s.SetParameters(tlfea.SyncedNewtonParams(*PRM))
    s.AnalyzeHessianSparsity()
    s.SetFixedSparsityPattern(True)
    return d, s


@pytest.mark.gpu
@pytest.mark.parametrize("model,mode", [("svk", 2), ("svk", 1), ("mr", 2), ("mr", 3)])
def test_bunny_assembly_vs_oracle(model, mode):
    """Pattern bit-exact, f_int and H = M/h + h K_t + h^2 rho J^T J within 1e-12 of the oracle on the deformed bunny."""
    from oracle import oracle as orc

    X, E, fixed, f_ext = _load()
    d, s = _make(model, X, E, fixed, f_ext)
    s.SetAssembly(mode)
    o = orc.T10Oracle(E, X)
    off, cols, mvals = d.RetrieveMassCSRToCPU()
    assert np.array_equal(off, o.off) and np.array_equal(cols, o.cols)
    assert rel_err(mvals, o.mass(RHO)) < 1e-12
    ro, ci, _ = s.RetrieveHessianCSR(want_values=False)
    ro_ref, ci_ref = orc.dof_pattern(o.off, o.cols)
    assert np.array_equal(ro, ro_ref) and np.array_equal(ci, ci_ref)
    rng = np.random.default_rng(5)
    x = X + rng.normal(0, 0.01, X.shape)
    d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
    s.Assemble()
    mat = _mat(orc, model)
    f_ref, _, _ = o.force(x, mat)
    H_ref = o.hessian(x, mat, DT, RHO, fixed, ALM)
    H, f = s.RetrieveHessianValues(), d.RetrieveInternalForceToCPU()
    big = np.abs(H_ref - o.hessian(x, mat, DT, RHO)) > 0.5 * DT**2 * ALM  # the 3 |fixed| penalty entries h^2 rho
    assert big.sum() == 3 * len(fixed)
    assert rel_err(f, f_ref) < 1e-12
    assert rel_err(H[big], H_ref[big]) < 1e-12 and rel_err(H[~big], H_ref[~big]) < 1e-12
    s.Assemble()
    assert np.array_equal(H, s.RetrieveHessianValues())


@pytest.mark.gpu
@pytest.mark.parametrize("model", ["svk", "mr"])
def test_bunny_newton_step_vs_cpu_restatement(model):
    """One Solve() (OneStepNewtonCuDSS, SyncedNewton.cu:909-1394; PCG to 1e-13 in place of cuDSS) against the same
    ALM/Newton loop written from the oracle's pieces with scipy's sparse LU: positions within 1e-8 of the largest
    displacement, ||c|| below the outer tolerance."""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl
    from oracle import oracle as orc

    X, E, fixed, f_ext = _load()
    d, s = _make(model, X, E, fixed, f_ext)
    s.SetPCG(50000, 1e-13)
    s.Solve()
    st = s.LastStats()
    gx, gy, gz = d.RetrievePositionToCPU()
    pos_gpu = np.stack([gx, gy, gz], axis=1)

    o = orc.T10Oracle(E, X)
    mat = _mat(orc, model)
    nn = X.shape[0]
    mass = o.mass(RHO)
    J, Jt = orc.fixed_node_jacobians(fixed, nn)
    ro, ci = orc.dof_pattern(o.off, o.cols)
    inner_atol, inner_rtol, outer_tol, rho, max_outer, max_inner, h = PRM
    pos, v, v_prev, lam = X.copy(), np.zeros(3 * nn), np.zeros(3 * nn), np.zeros(3 * len(fixed))
    n_newton = 0
    for _outer in range(max_outer):
        g0 = None
        for _it in range(max_inner):
            f_int, _, _ = o.force(pos, mat)
            con = (pos[fixed] - X[fixed]).reshape(-1)
            g = orc.grad_L(o.off, o.cols, mass, v, v_prev, f_int, f_ext, h, Jt, lam, con, rho)
            ng = np.linalg.norm(g)
            g0 = ng if g0 is None else g0
            if ng < inner_atol or (inner_rtol > 0 and ng <= inner_rtol * g0):
                break
            Hv = o.hessian(pos, mat, h, RHO, fixed, rho)
            A = sp.csr_matrix((Hv, ci, ro), shape=(3 * nn, 3 * nn)).tocsc()
            v = v + spl.spsolve(A, -g)
            pos = X + h * v.reshape(-1, 3)
            n_newton += 1
        v_prev = v.copy()
        con = (pos[fixed] - X[fixed]).reshape(-1)
        lam = lam + rho * con
        if np.linalg.norm(con) < outer_tol:
            break
    scale = np.abs(pos - X).max()
    assert scale > 0 and st["newton_iters"] == n_newton, (st, n_newton)
    assert st["norm_c"] < outer_tol
    assert np.abs(pos_gpu - pos).max() < 1e-8 * scale + 1e-14, (np.abs(pos_gpu - pos).max(), scale)


@pytest.mark.gpu
@pytest.mark.parametrize("model", ["svk", "mr"])
def test_cpp_bunny_driver_matches_the_python_mirror(tmp_path, model):
    """examples/bunny_newton.cc -- the call sequence of lib_bin/mesh_deform/test_feat10_bunny_newton.cc written against
    the reference's class and method names with Eigen-style containers (include/tlfea_shim.hpp templates) -- compiled
    with g++, run on the GPU: positions after two steps equal the Python mirror's bit for bit (same library, same
    calls), the VTK file has the reference's layout (FEAT10Data.cu:841-877)."""
    import shutil
    import subprocess

    gxx = shutil.which("g++")
    if gxx is None:
        pytest.skip("no g++")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib_dir = os.path.join(root, "total-lagrangian-fea_b200", "lib")
    exe, vtk = tmp_path / "bunny", tmp_path / "bunny.vtk"
    subprocess.check_call([gxx, "-std=c++17", "-O1", "-I", os.path.join(root, "include"), "-I",
                           os.path.join(root, "examples"), os.path.join(root, "examples", "bunny_newton.cc"), "-o",
                           str(exe), "-L", lib_dir, "-ltlfea_b200", "-Wl,-rpath," + lib_dir,
                           "-Wl,--unresolved-symbols=ignore-in-shared-libs"])
    out = subprocess.check_output([str(exe), NODE, ELE, model, "2", str(vtk)],
                                  env=dict(os.environ, LD_LIBRARY_PATH=lib_dir), text=True)
    pos_cpp = np.array([[float(t) for t in ln.split()[2:5]] for ln in out.splitlines() if ln.startswith("pos ")])
    head = [ln for ln in out.splitlines() if ln.startswith("elements")][0].split()
    assert int(head[1]) == 1066 and int(head[3]) == 2095
    assert "connectivity round trip ok" in out
    X, E, fixed, f_ext = _load()
    assert f"J {3 * len(fixed)} x {3 * X.shape[0]}" in out
    d, s = _make(model, X, E, fixed, f_ext)
    d.CalcP()
    d.CalcInternalForce()
    s.SetPCG(50000, 1e-13)
    for _ in range(2):
        s.Solve()
    gx, gy, gz = d.RetrievePositionToCPU()
    assert pos_cpp.shape == (X.shape[0], 3)
    assert np.array_equal(pos_cpp, np.stack([gx, gy, gz], axis=1))
    lines = vtk.read_text().splitlines()
    assert lines[3] == "DATASET UNSTRUCTURED_GRID" and lines[4] == f"POINTS {X.shape[0]} float"
    assert f"CELLS {E.shape[0]} {E.shape[0] * 11}" in lines and lines[-1] == "24"
