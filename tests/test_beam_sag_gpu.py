"""Config C1 (BASELINE.json): beam_sag -- ANCF3243 cantilever under a tip load, lib_bin/beam_sag/test_ancf3243.cc
defaults (30 elements, L=0.5, W=H=0.1, E=7e8, nu=0.33, rho0=2700, damping 1e5/1e5, tip Fz=3100, dt=1e-3, Newton
parameters {1e-4, 0, 1e-6, 1e14, 5, 10, dt}, coefficients 0..3 pinned).

The GPU step (tlfea_newton_solve_step: OneStepNewtonCuDSS, SyncedNewton.cu:909-1394, with the cuDSS solve
replaced by PCG) is checked against the same ALM/Newton loop written here on the CPU from the oracle's
pieces with a sparse direct solve.  Tolerance on positions after several steps: 1e-8 relative (the PCG is run to
1e-13; the remaining difference is rounding through the ill-conditioned ALM system, rho = 1e14)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

N_EL, L, W, H = 30, 0.5, 0.1, 0.1
E_MOD, NU, RHO, ETA, LAMD = 7e8, 0.33, 2700.0, 1e5, 1e5
FZ, DT, ALM = 3100.0, 1e-3, 1e14
PRM = (1e-4, 0.0, 1e-6, ALM, 5, 10, DT)
N_STEPS = 5


def _mesh():
    nn = N_EL + 1
    x = np.zeros(4 * nn); y = np.zeros(4 * nn); z = np.zeros(4 * nn)
    for i in range(nn):
        x[4 * i:4 * i + 4] = [i * L, 1.0, 0.0, 0.0]
        y[4 * i:4 * i + 4] = [0.0, 0.0, 1.0, 0.0]
        z[4 * i:4 * i + 4] = [0.0, 0.0, 0.0, 1.0]
    conn = np.array([[i, i + 1] for i in range(N_EL)], dtype=np.int32)
    return x, y, z, conn


def _cpu_steps(x, y, z, conn, f_ext, fixed):
    """OneStepNewtonCuDSS restated with the oracle's kernels + scipy's sparse LU."""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl
    from oracle import oracle as orc

    nc = x.shape[0]
    ec = orc.ancf_ecoef(3243, conn)
    Binv = orc.ancf_binv(3243, [L] * N_EL, [W] * N_EL, [H] * N_EL)
    LWH = np.tile([L, W, H], (N_EL, 1))
    gN, dJ = orc.ancf_precompute(3243, ec, x, y, z, LWH, Binv)
    qw = orc.ancf_stiffness_weights(3243)
    off, cols = orc.coef_pattern(ec, nc)
    mass = orc.ancf_mass(3243, ec, x, y, z, LWH, Binv, RHO, off, cols)
    mat = orc.Material.svk(E_MOD, NU, ETA, LAMD)
    (J, Jt) = orc.fixed_node_jacobians(fixed, nc)
    ro, ci = orc.dof_pattern(off, cols)
    pos = np.stack([x, y, z], axis=1).copy()
    ref = pos.copy()
    v = np.zeros(3 * nc); v_prev = np.zeros(3 * nc); lam = np.zeros(3 * len(fixed))
    inner_atol, inner_rtol, outer_tol, rho, max_outer, max_inner, h = PRM
    out = []
    for _ in range(N_STEPS):
        pos_prev = pos.copy()
        for _outer in range(max_outer):
            for _it in range(max_inner):
                _, P, _, _ = orc.compute_p(ec, pos, gN, mat, v, want_visc=True)
                f_int = orc.internal_force(ec, gN, dJ, qw, P, nc)
                con = (pos[fixed] - ref[fixed]).reshape(-1)
                g = orc.grad_L(off, cols, mass, v, v_prev, f_int, f_ext, h, Jt, lam, con, rho)
                if np.linalg.norm(g) < inner_atol:
                    break
                Hv = np.zeros(9 * cols.shape[0])
                orc.hessian_mass(off, cols, mass, h, Hv)
                orc.tangent_assemble(ec, pos, gN, dJ, qw, mat, h, off, cols, Hv)
                orc.hessian_constraints(*J, h, rho, off, cols, Hv)
                A = sp.csr_matrix((Hv, ci, ro), shape=(3 * nc, 3 * nc)).tocsc()
                dv = spl.spsolve(A, -g)
                v = v + dv
                pos = pos_prev + h * v.reshape(-1, 3)
            v_prev = v.copy()
            con = (pos[fixed] - ref[fixed]).reshape(-1)
            lam = lam + rho * con
            if np.linalg.norm(con) < outer_tol:
                break
        out.append(pos.copy())
    return out


def test_beam_sag_newton_steps_match_cpu_restatement():
    import tlfea

    x, y, z, conn = _mesh()
    nc = x.shape[0]
    fixed = np.arange(4, dtype=np.int32)
    f_ext = np.zeros(3 * nc)
    f_ext[3 * (4 * N_EL) + 2] = FZ  # position coefficient of the tip node, z

    d = tlfea.GPU_ANCF3243_Data(N_EL + 1, N_EL)
    d.Initialize()
    d.SetExternalForce(f_ext)
    d.Setup(L, W, H, x, y, z, conn)
    d.SetDensity(RHO)
    d.SetDamping(ETA, LAMD)
    d.SetSVK(E_MOD, NU)
    d.SetNodalFixed(fixed)
    d.CalcDsDuPre()
    d.CalcMassMatrix()
    d.CalcConstraintData()
    s = tlfea.SyncedNewtonSolver(d, 3 * len(fixed))
    s.SetParameters(tlfea.SyncedNewtonParams(*PRM))
    s.Setup()
    s.AnalyzeHessianSparsity()
    s.SetFixedSparsityPattern(True)
    s.SetPCG(20000, 1e-13)

    ref = _cpu_steps(x, y, z, conn, f_ext, fixed)
    tip_prev = 0.0
    for k in range(N_STEPS):
        s.Solve()
        st = s.LastStats()
        gx, gy, gz = d.RetrievePositionToCPU()
        pos = np.stack([gx, gy, gz], axis=1)
        scale = np.abs(ref[k] - np.stack([x, y, z], axis=1)).max()
        assert st["norm_c"] < 1e-6, st
        assert np.abs(pos - ref[k]).max() < 1e-8 * max(scale, 1e-12) + 1e-13, (k, np.abs(pos - ref[k]).max(), scale)
        tip = gz[4 * N_EL]
        assert tip > tip_prev  # the tip keeps moving up under the +z load
        tip_prev = tip


def test_cpp_driver_through_the_shim_matches_the_python_mirror(tmp_path):
    """examples/beam_sag.cc -- the beam_sag driver written in host C++ against the reference's class names
    (include/tlfea_shim.hpp over the C ABI) -- is compiled with g++, run on the GPU, and must print the tip
    positions the Python mirror produces for the same calls (same library, same operations: same bits)."""
    import os
    import shutil
    import subprocess

    import tlfea

    gxx = shutil.which("g++")
    if gxx is None:
        pytest.skip("no g++")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib_dir = os.path.join(root, "total-lagrangian-fea_b200", "lib")
    exe = tmp_path / "beam_sag"
    subprocess.check_call([gxx, "-std=c++17", "-O1", "-I", os.path.join(root, "include"),
                           os.path.join(root, "examples", "beam_sag.cc"), "-o", str(exe), "-L", lib_dir, "-ltlfea_b200",
                           "-Wl,-rpath," + lib_dir, "-Wl,--unresolved-symbols=ignore-in-shared-libs"])
    out = subprocess.check_output([str(exe), str(N_STEPS)], env=dict(os.environ, LD_LIBRARY_PATH=lib_dir), text=True)
    tips_cpp = [float(ln.split()[3]) for ln in out.splitlines() if ln.startswith("step")]
    assert len(tips_cpp) == N_STEPS

    x, y, z, conn = _mesh()
    f_ext = np.zeros(3 * x.shape[0])
    f_ext[3 * (4 * N_EL) + 2] = FZ
    d = tlfea.GPU_ANCF3243_Data(N_EL + 1, N_EL)
    d.Initialize()
    d.SetExternalForce(f_ext)
    d.Setup(L, W, H, x, y, z, conn)
    d.SetDensity(RHO)
    d.SetDamping(ETA, LAMD)
    d.SetSVK(E_MOD, NU)
    d.SetNodalFixed(np.arange(4, dtype=np.int32))
    d.CalcDsDuPre()
    d.CalcMassMatrix()
    d.CalcConstraintData()
    s = tlfea.SyncedNewtonSolver(d, 12)
    s.SetParameters(tlfea.SyncedNewtonParams(*PRM))
    s.Setup()
    s.AnalyzeHessianSparsity()
    s.SetFixedSparsityPattern(True)
    s.SetPCG(20000, 1e-13)
    for k in range(N_STEPS):
        s.Solve()
        tip = d.RetrievePositionToCPU()[2][4 * N_EL]
        assert tip == tips_cpp[k], (k, tip, tips_cpp[k])
    assert tips_cpp[-1] > tips_cpp[0] > 0.0
