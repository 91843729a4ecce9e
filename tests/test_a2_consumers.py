"""SURVEY 8(f) consumers of the hot path.
(f3) VBD: per-node internal force and 3x3 diagonal tangent block, pinned to the reference's own prototype
     (test-scripts/vbd_proto/alm_vbd_t10_svk.py:175-223, fixture tests/golden/vbd_t10_beam.npz made by
     tests/golden/make_golden_vbd.py) -- the CPU statement of vbd_accumulate_residual_and_hessian_diag
     (FEAT10DataFunc.cuh:295-395).
(f4) contact coupling: [x | y | z] node buffer out, f_ext = gravity + contact in, one Solve()
     (CollisionSystemBase.h:17-47, test_item_drop.cc:560-602)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, rel_err

E_MOD, NU, RHO, H_STEP = 7e8, 0.33, 2700.0, 1e-3


def _diag_blocks(off, cols, Hv):
    n = off.shape[0] - 1
    D = np.zeros((n, 3, 3))
    for i in range(n):
        deg = off[i + 1] - off[i]
        s = int(np.searchsorted(cols[off[i]:off[i + 1]], i))
        for d in range(3):
            base = 9 * off[i] + d * 3 * deg + 3 * s
            D[i, d] = Hv[base:base + 3]
    return D


def test_oracle_diag_blocks_match_vbd_prototype(golden_t10):
    from oracle import oracle as orc

    g, v = golden_t10, dict(np.load(os.path.join(GOLDEN, "vbd_t10_beam.npz")))
    o = orc.T10Oracle(g["E"], g["X"])
    mat = orc.Material.svk(E_MOD, NU)
    f, _, _ = o.force(g["x"], mat)
    assert rel_err(f.reshape(-1, 3), v["f_int_node"]) < 1e-12
    Kd = _diag_blocks(o.off, o.cols, o.tangent(g["x"], mat, 1.0))
    assert rel_err(Kd, v["Kii"]) < 1e-12
    Md = _diag_blocks(o.off, o.cols, orc.hessian_mass(o.off, o.cols, o.mass(RHO), 1.0, np.zeros(9 * o.cols.shape[0])))
    assert rel_err(Md, v["M_diag"]) < 1e-12


@pytest.mark.gpu
def test_gpu_diag_blocks_and_force_match_vbd_prototype(golden_t10):
    """tlfea_newton_get_diag_blocks = M_ii/h I + h K_ii, f_int per node: against the prototype's numbers."""
    import tlfea

    g, v = golden_t10, dict(np.load(os.path.join(GOLDEN, "vbd_t10_beam.npz")))
    X, E, x = g["X"], g["E"], g["x"]
    d = tlfea.GPU_FEAT10_Data(E.shape[0], X.shape[0])
    d.Initialize()
    d.Setup(*tlfea.tet5pt_quadrature(), X[:, 0], X[:, 1], X[:, 2], E)
    d.SetDensity(RHO)
    d.SetSVK(E_MOD, NU)
    d.CalcDnDuPre()
    d.CalcMassMatrix()
    s = tlfea.SyncedNewtonSolver(d, 0)
    s.SetParameters(tlfea.SyncedNewtonParams(1e-4, 0.0, 1e-6, 1e14, 1, 1, H_STEP))
    s.Setup()
    s.AnalyzeHessianSparsity()
    d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
    for mode in (2, 1):
        s.SetAssembly(mode)
        s.Assemble()
        D = s.RetrieveHessianDiagBlocks()
        want = v["M_diag"] / H_STEP + H_STEP * v["Kii"]
        assert rel_err(D, want) < 1e-12
        assert rel_err(d.RetrieveInternalForceToCPU().reshape(-1, 3), v["f_int_node"]) < 1e-12
        s.AssembleForce()  # the gradient-only entry gives the same nodal forces
        assert rel_err(d.RetrieveInternalForceToCPU().reshape(-1, 3), v["f_int_node"]) < 1e-12


@pytest.mark.gpu
def test_contact_coupling_step_matches_cpu():
    """A stand-in collision system (penalty floor at z = z0, computed in torch FROM the packed [x|y|z] buffer) feeds
    f_ext = gravity + contact on the device; one Solve() then matches the CPU restatement with the same f_ext."""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl
    import torch
    import tlfea
    from tlfea import meshgen
    from oracle import oracle as orc

    X, E = meshgen.kuhn_block(3)
    nn = X.shape[0]
    fixed = np.nonzero(X[:, 0] == 0.0)[0].astype(np.int32)
    d = tlfea.GPU_FEAT10_Data(E.shape[0], nn)
    d.Initialize()
    d.SetNodalFixed(fixed)
    d.Setup(*tlfea.tet5pt_quadrature(), X[:, 0], X[:, 1], X[:, 2], E)
    d.SetDensity(RHO)
    d.SetSVK(E_MOD, NU)
    d.CalcDnDuPre()
    d.CalcMassMatrix()
    d.CalcConstraintData()
    prm = (1e-6, 0.0, 1e-8, 1e12, 3, 8, H_STEP)
    s = tlfea.SyncedNewtonSolver(d, 3 * len(fixed))
    s.Setup()
    s.SetParameters(tlfea.SyncedNewtonParams(*prm))
    s.AnalyzeHessianSparsity()
    s.SetPCG(50000, 1e-13)

    lump = np.add.reduceat(d.RetrieveMassCSRToCPU()[2], d.RetrieveMassCSRToCPU()[0][:-1])
    grav = np.zeros(3 * nn)
    grav[2::3] = -9.81 * lump
    z0, kpen = 0.2, 5e4
    nodes = torch.empty(3 * nn, dtype=torch.float64, device="cuda")
    d.PackNodesXYZ(nodes.data_ptr())
    d.Sync()
    got = nodes.cpu().numpy()
    assert np.array_equal(got[:nn], X[:, 0]) and np.array_equal(got[nn:2 * nn], X[:, 1]) and np.array_equal(got[2 * nn:], X[:, 2])
    contact = torch.zeros(3 * nn, dtype=torch.float64, device="cuda")
    contact[2::3] = kpen * torch.clamp(z0 - nodes[2 * nn:], min=0.0)
    g_dev = torch.from_numpy(grav).cuda()
    torch.cuda.synchronize()
    d.SetExternalForceDevice(g_dev.data_ptr(), contact.data_ptr(), 1.0)
    f_ext = grav.copy()
    f_ext[2::3] += kpen * np.maximum(z0 - X[:, 2], 0.0)
    s.Solve()
    gx, gy, gz = d.RetrievePositionToCPU()
    pos_gpu = np.stack([gx, gy, gz], axis=1)

    o = orc.T10Oracle(E, X)
    mat = orc.Material.svk(E_MOD, NU)
    mass = o.mass(RHO)
    J, Jt = orc.fixed_node_jacobians(fixed, nn)
    ro, ci = orc.dof_pattern(o.off, o.cols)
    atol, rtol, otol, rho, max_outer, max_inner, h = prm
    pos, v, v_prev, lam = X.copy(), np.zeros(3 * nn), np.zeros(3 * nn), np.zeros(3 * len(fixed))
    for _outer in range(max_outer):
        for _it in range(max_inner):
            f_int, _, _ = o.force(pos, mat)
            con = (pos[fixed] - X[fixed]).reshape(-1)
            gL = orc.grad_L(o.off, o.cols, mass, v, v_prev, f_int, f_ext, h, Jt, lam, con, rho)
            if np.linalg.norm(gL) < atol:
                break
            A = sp.csr_matrix((o.hessian(pos, mat, h, RHO, fixed, rho), ci, ro), shape=(3 * nn, 3 * nn)).tocsc()
            v = v + spl.spsolve(A, -gL)
            pos = X + h * v.reshape(-1, 3)
        v_prev = v.copy()
        con = (pos[fixed] - X[fixed]).reshape(-1)
        lam = lam + rho * con
        if np.linalg.norm(con) < otol:
            break
    scale = np.abs(pos - X).max()
    assert scale > 0
    assert np.abs(pos_gpu - pos).max() < 1e-8 * scale + 1e-15, (np.abs(pos_gpu - pos).max(), scale)
