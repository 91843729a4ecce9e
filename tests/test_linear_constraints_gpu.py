"""General linear constraints (SetLinearConstraintsCSR, ANCF3243Data.cuh:810-940) and the constraint-aware Hessian
pattern (AnalyzeHessianSparsity, SyncedNewton.cu:556-801): two separate ANCF3243 beams welded tip to root by a
constraint row that couples coefficients of different mesh components.  Without the J^T J couplings in the pattern
the weld's off-diagonal blocks would be dropped (SyncedNewton.cu:336); the reference widens the pattern and so do we.

Checked against the oracle assembled on the widened pattern (bit-exact pattern, 1e-12 values), and the mass getter
must still hand out the element adjacency only (BuildMassCSRPattern)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

L, W, H = 0.5, 0.1, 0.1
E_MOD, NU, RHO = 7e8, 0.33, 2700.0
DT, ALM = 1e-3, 1e10
PRM = (1e-6, 0.0, 1e-8, ALM, 3, 8, DT)


def _two_beams():
    """beam A: nodes 0-1-2 along x at y=0; beam B: nodes 3-4-5 along x starting at the tip of A, lifted in y."""
    nn = 6
    x = np.zeros(4 * nn); y = np.zeros(4 * nn); z = np.zeros(4 * nn)
    for i in range(nn):
        beam, k = divmod(i, 3)
        x[4 * i:4 * i + 4] = [(2 * beam + k) * L, 1.0, 0.0, 0.0]
        y[4 * i:4 * i + 4] = [0.05 * beam, 0.0, 1.0, 0.0]
        z[4 * i:4 * i + 4] = [0.0, 0.0, 0.0, 1.0]
    conn = np.array([[0, 1], [1, 2], [3, 4], [4, 5]], dtype=np.int32)
    return x, y, z, conn


def _constraints(x, y, z):
    """clamp the 4 coefficients of node 0 (12 rows) + weld position coefficient of node 2 to that of node 3 (3 rows)."""
    pos = np.stack([x, y, z], axis=1)
    off, cols, vals, rhs = [0], [], [], []
    for c in range(4):
        for d in range(3):
            cols.append(3 * c + d); vals.append(1.0); rhs.append(pos[c, d]); off.append(len(cols))
    ca, cb = 4 * 2, 4 * 3
    for d in range(3):
        cols += [3 * ca + d, 3 * cb + d]; vals += [1.0, -1.0]; rhs.append(pos[ca, d] - pos[cb, d]); off.append(len(cols))
    return (np.array(off, dtype=np.int32), np.array(cols, dtype=np.int32), np.array(vals), np.array(rhs))


def _widened_pattern(off, cols, j_off, j_cols, n_coef):
    rows = [set(cols[off[i]:off[i + 1]].tolist()) for i in range(n_coef)]
    for r in range(j_off.shape[0] - 1):
        cs = sorted({int(c) // 3 for c in j_cols[j_off[r]:j_off[r + 1]]})
        for a in cs:
            rows[a].update(cs)
    o = np.zeros(n_coef + 1, dtype=np.int32)
    o[1:] = np.cumsum([len(r) for r in rows])
    return o, np.array([c for r in rows for c in sorted(r)], dtype=np.int32)


def _build(order_late: bool):
    import tlfea

    x, y, z, conn = _two_beams()
    j_off, j_cols, j_vals, rhs = _constraints(x, y, z)
    d = tlfea.GPU_ANCF3243_Data(6, 4)
    d.Initialize()
    d.Setup(L, W, H, x, y, z, conn)
    d.SetDensity(RHO)
    d.SetSVK(E_MOD, NU)
    f_ext = np.zeros(3 * x.shape[0])
    f_ext[3 * (4 * 5) + 2] = -50.0  # tip of beam B
    d.SetExternalForce(f_ext)
    if not order_late:
        d.SetLinearConstraintsCSR(j_off, j_cols, j_vals, rhs)
    d.CalcDsDuPre()
    d.CalcMassMatrix()
    if order_late:  # the pattern and the mass already exist: both must be rebuilt with the couplings
        d.SetLinearConstraintsCSR(j_off, j_cols, j_vals, rhs)
    d.CalcConstraintData()
    s = tlfea.SyncedNewtonSolver(d, rhs.shape[0])
    s.SetParameters(tlfea.SyncedNewtonParams(*PRM))
    s.Setup()
    s.AnalyzeHessianSparsity()
    s.SetFixedSparsityPattern(True)
    return d, s, (x, y, z, conn), (j_off, j_cols, j_vals, rhs), f_ext


@pytest.mark.parametrize("order_late", [False, True])
def test_weld_couples_components_and_matches_oracle(order_late):
    from oracle import oracle as orc

    d, s, (x, y, z, conn), (j_off, j_cols, j_vals, rhs), _ = _build(order_late)
    nc = x.shape[0]
    ec = orc.ancf_ecoef(3243, conn)
    b_off, b_cols = orc.coef_pattern(ec, nc)
    w_off, w_cols = _widened_pattern(b_off, b_cols, j_off, j_cols, nc)
    assert w_cols.shape[0] == b_cols.shape[0] + 2  # the weld adds (8,12) and (12,8)

    # mass getter: element adjacency only, values as the oracle's
    Binv = orc.ancf_binv(3243, [L] * 4, [W] * 4, [H] * 4)
    LWH = np.tile([L, W, H], (4, 1))
    m_off, m_cols, m_vals = d.RetrieveMassCSRToCPU()
    n_base = int(m_off[-1])
    assert np.array_equal(m_off, b_off) and np.array_equal(m_cols[:n_base], b_cols)
    mass_b = orc.ancf_mass(3243, ec, x, y, z, LWH, Binv, RHO, b_off, b_cols)
    assert np.allclose(m_vals[:n_base], mass_b, rtol=1e-12, atol=1e-12 * np.abs(mass_b).max())

    # Hessian on the widened pattern: pattern bit-exact, values 1e-12
    rng = np.random.default_rng(7)
    pos = np.stack([x, y, z], axis=1) + 1e-3 * rng.uniform(-1, 1, (nc, 3))
    d.UpdatePositions(pos[:, 0].copy(), pos[:, 1].copy(), pos[:, 2].copy())
    s.Assemble()
    ro, ci, va = s.RetrieveHessianCSR()
    r_ro, r_ci = orc.dof_pattern(w_off, w_cols)
    assert np.array_equal(ro, r_ro.astype(np.int64)) and np.array_equal(ci, r_ci)
    gN, dJ = orc.ancf_precompute(3243, ec, x, y, z, LWH, Binv)
    qw = orc.ancf_stiffness_weights(3243)
    mass_w = orc.ancf_mass(3243, ec, x, y, z, LWH, Binv, RHO, w_off, w_cols)
    Hv = np.zeros(9 * w_cols.shape[0])
    orc.hessian_mass(w_off, w_cols, mass_w, DT, Hv)
    orc.tangent_assemble(ec, pos, gN, dJ, qw, orc.Material.svk(E_MOD, NU), DT, w_off, w_cols, Hv)
    orc.hessian_constraints(j_off, j_cols, j_vals, DT, ALM, w_off, w_cols, Hv)
    assert np.abs(va - Hv).max() <= 1e-12 * np.abs(Hv).max()
    # the weld's off-diagonal block is really there: -h^2 rho on the (coef 8, coef 12) diagonal
    import scipy.sparse as sp

    A = sp.csr_matrix((va, ci, ro), shape=(3 * nc, 3 * nc))
    assert A[3 * 8, 3 * 12] == pytest.approx(-DT * DT * ALM, rel=1e-12)
    assert A[3 * 12 + 2, 3 * 8 + 2] == pytest.approx(-DT * DT * ALM, rel=1e-12)


def test_welded_beams_step_holds_the_weld():
    d, s, (x, y, z, conn), (j_off, j_cols, j_vals, rhs), _ = _build(False)
    s.SetPCG(20000, 1e-12)
    for _ in range(3):
        s.Solve()
    st = s.LastStats()
    gx, gy, gz = d.RetrievePositionToCPU()
    pos = np.stack([gx, gy, gz], axis=1)
    # weld: relative position of coefficient 8 and 12 unchanged; clamp: node 0 coefficients unchanged
    ref = np.stack([x, y, z], axis=1)
    assert np.abs((pos[8] - pos[12]) - (ref[8] - ref[12])).max() < 1e-6
    assert np.abs(pos[:4] - ref[:4]).max() < 1e-6  # ALM with rho = 1e10, 3 outer iterations
    assert st["norm_c"] < 1e-6
    assert gz[4 * 5] < 0.0  # the loaded tip of beam B moved down
    assert np.abs(pos[8] - ref[8]).max() > 1e-9  # beam A carries no load: it can only have moved through the weld


def _plate_3443():
    """two ANCF3443 elements side by side (6 nodes), node 0 clamped, the position coefficient of node 5 tied to node 4 by
    rows with unequal weights (a general linear row, not a selection)."""
    import tlfea
    from test_ancf_gpu import plate

    x, y, z, conn = plate(2, 1, L, 0.4)
    nc = x.shape[0]
    pos = np.stack([x, y, z], axis=1)
    off, cols, vals, rhs = [0], [], [], []
    for c in range(4):
        for dd in range(3):
            cols.append(3 * c + dd); vals.append(1.0); rhs.append(pos[c, dd]); off.append(len(cols))
    ca, cb = 4 * 4, 4 * 5
    for dd in range(3):
        cols += [3 * ca + dd, 3 * cb + dd]; vals += [1.0, -0.5]; rhs.append(pos[ca, dd] - 0.5 * pos[cb, dd]); off.append(len(cols))
    jc = (np.array(off, dtype=np.int32), np.array(cols, dtype=np.int32), np.array(vals), np.array(rhs))
    d = tlfea.GPU_ANCF3443_Data(nc // 4, conn.shape[0])
    d.Initialize()
    d.Setup(L, 0.4, H, x, y, z, conn)
    d.SetDensity(RHO)
    d.SetSVK(E_MOD, NU)
    f_ext = 30.0 * np.random.default_rng(3).uniform(-1, 1, 3 * nc)
    d.SetExternalForce(f_ext)
    d.SetLinearConstraintsCSR(*jc)
    d.CalcDsDuPre()
    d.CalcMassMatrix()
    d.CalcConstraintData()
    s = tlfea.SyncedNewtonSolver(d, jc[3].shape[0])
    s.SetParameters(tlfea.SyncedNewtonParams(*PRM))
    s.Setup()
    s.AnalyzeHessianSparsity()
    return d, s, (x, y, z, conn), jc, f_ext, (L, 0.4, H)


@pytest.mark.parametrize("kind", [3243, 3443])
def test_gradient_vs_oracle_linear_constraints(kind):
    """g = M (v - v_prev)/h + f_int - f_ext + h J^T (lambda + rho c)  (solver_grad_L, SyncedNewton.cu:343-407) for the
    ANCF elements under general linear CSR constraints, non-zero v, v_prev, f_ext and constraint violation, against
    orc.grad_L with the oracle's own f_int and mass (lambda is zero after Setup, as in the reference)."""
    import scipy.sparse as sp
    from oracle import oracle as orc

    if kind == 3243:
        d, s, (x, y, z, conn), (j_off, j_cols, j_vals, rhs), f_ext = _build(False)
        dims = (L, W, H)
    else:
        d, s, (x, y, z, conn), (j_off, j_cols, j_vals, rhs), f_ext, dims = _plate_3443()
    nc, ne = x.shape[0], conn.shape[0]
    rng = np.random.default_rng(21)
    pos = np.stack([x, y, z], axis=1) + 2e-3 * rng.uniform(-1, 1, (nc, 3))
    v, vp = 0.1 * rng.uniform(-1, 1, 3 * nc), 0.1 * rng.uniform(-1, 1, 3 * nc)
    d.UpdatePositions(pos[:, 0].copy(), pos[:, 1].copy(), pos[:, 2].copy())
    s.SetVelocityGuess(v)
    s.SetVelocityPrev(vp)
    s.Assemble()
    ng = s.Residual()
    g = s.RetrieveGradient()

    ec = orc.ancf_ecoef(kind, conn)
    b_off, b_cols = orc.coef_pattern(ec, nc)
    w_off, w_cols = _widened_pattern(b_off, b_cols, j_off, j_cols, nc)
    Binv = orc.ancf_binv(kind, [dims[0]] * ne, [dims[1]] * ne, [dims[2]] * ne)
    LWH = np.tile(list(dims), (ne, 1))
    gN, dJ = orc.ancf_precompute(kind, ec, x, y, z, LWH, Binv)
    mat = orc.Material.svk(E_MOD, NU)
    _, P = orc.compute_p(ec, pos, gN, mat)
    f_int = orc.internal_force(ec, gN, dJ, orc.ancf_stiffness_weights(kind), P, nc)
    mass_w = orc.ancf_mass(kind, ec, x, y, z, LWH, Binv, RHO, w_off, w_cols)
    J = sp.csr_matrix((j_vals, j_cols, j_off), shape=(rhs.shape[0], 3 * nc))
    con = J @ pos.reshape(-1) - rhs
    assert np.abs(con).max() > 1e-4  # the constraints really are violated
    JT = J.T.tocsr()
    JT.sort_indices()
    jt = (JT.indptr.astype(np.int32), JT.indices.astype(np.int32), JT.data)
    g_ref = orc.grad_L(w_off, w_cols, mass_w, v, vp, f_int, f_ext, DT, jt, np.zeros_like(con), con, ALM)
    assert np.abs(g - g_ref).max() <= 1e-12 * np.abs(g_ref).max()
    assert abs(ng - np.linalg.norm(g_ref)) <= 1e-12 * np.linalg.norm(g_ref)
