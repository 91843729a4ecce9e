"""GPU parity tests, ANCF3243 beam and ANCF3443 shell.

Set-up path (B^-T, reference gradients, mass) is pinned by the reference's own known-answer mass matrices
data/utest/mass_matrix_{2,3}_beam.csv (lib_utest/utest_3243.cc:74-112,159-197; tolerance 1e-4 there, 1e-9 here)
and by quadrature invariants; the force / tangent kernels for S = 8 and S = 16 are compared with the oracle's
generic restatement (compute_p, compute_internal_force, compute_hessian_assemble_csr twins:
ANCF3243DataFunc.cuh:189-454,642-940, ANCF3443DataFunc.cuh:281-535,724-1033) on the same gradients, within
1e-12 relative to the largest entry."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-12
E_MOD, NU, RHO, H_STEP = 7e8, 0.33, 2700.0, 1e-3

GL = {
    2: ([-0.57735026918962576, 0.57735026918962576], [1.0, 1.0]),
    3: ([-0.77459666924148340, 0.0, 0.77459666924148340],
        [0.55555555555555556, 0.88888888888888889, 0.55555555555555556]),
    4: ([-0.8611363115940526, -0.3399810435848563, 0.3399810435848563, 0.8611363115940526],
        [0.3478548451374538, 0.6521451548625461, 0.6521451548625461, 0.3478548451374538]),
}


def stiffness_weights(kind):
    n = (3, 2, 2) if kind == 3243 else (4, 4, 3)
    w = [GL[k][1] for k in n]
    return np.array([w[0][q // (n[1] * n[2])] * w[1][(q // n[2]) % n[1]] * w[2][q % n[2]]
                     for q in range(n[0] * n[1] * n[2])])


def straight_beam(n_elem, L):
    """GridMeshGenerator(n*L, 0, L, horizontal) of the reference's unit tests: nodes on the x axis,
    coefficients [r, r_u = e_x, r_v = e_y, r_w = e_z]."""
    nn = n_elem + 1
    x = np.zeros(4 * nn); y = np.zeros(4 * nn); z = np.zeros(4 * nn)
    for i in range(nn):
        x[4 * i:4 * i + 4] = [i * L, 1.0, 0.0, 0.0]
        y[4 * i:4 * i + 4] = [0.0, 0.0, 1.0, 0.0]
        z[4 * i:4 * i + 4] = [0.0, 0.0, 0.0, 1.0]
    conn = np.array([[i, i + 1] for i in range(n_elem)], dtype=np.int32)
    return x, y, z, conn


def plate(nx, ny, L, W):
    """Flat ANCF3443 plate in the xy plane: corner order P1..P4 = (-,-), (+,-), (+,+), (-,+) (cpu_utils.cc:214-217)."""
    nn = (nx + 1) * (ny + 1)
    x = np.zeros(4 * nn); y = np.zeros(4 * nn); z = np.zeros(4 * nn)
    for i in range(nx + 1):
        for j in range(ny + 1):
            n = i * (ny + 1) + j
            x[4 * n:4 * n + 4] = [i * L, 1.0, 0.0, 0.0]
            y[4 * n:4 * n + 4] = [j * W, 0.0, 1.0, 0.0]
            z[4 * n:4 * n + 4] = [0.0, 0.0, 0.0, 1.0]
    nid = lambda i, j: i * (ny + 1) + j
    conn = np.array([[nid(i, j), nid(i + 1, j), nid(i + 1, j + 1), nid(i, j + 1)]
                     for i in range(nx) for j in range(ny)], dtype=np.int32)
    return x, y, z, conn


def make(kind, x, y, z, conn, dims, model="svk", damp=False):
    import tlfea

    nn = x.shape[0] // 4
    cls = tlfea.GPU_ANCF3243_Data if kind == 3243 else tlfea.GPU_ANCF3443_Data
    d = cls(nn, conn.shape[0])
    d.Initialize()
    d.Setup(*dims, x, y, z, conn)
    d.SetDensity(RHO)
    if model == "svk":
        d.SetSVK(E_MOD, NU)
    else:
        mu = E_MOD / (2 * (1 + NU)); K = E_MOD / (3 * (1 - 2 * NU))
        d.SetMooneyRivlin(0.3 * mu, 0.2 * mu, 1.5 * K)
    if damp:
        d.SetDamping(1e5, 2e5)
    d.CalcDsDuPre()
    d.CalcMassMatrix()
    s = tlfea.SyncedNewtonSolver(d, 0)
    s.SetParameters(tlfea.SyncedNewtonParams(1e-4, 0.0, 1e-6, 1e14, 1, 1, H_STEP))
    s.Setup()
    s.AnalyzeHessianSparsity()
    return d, s


def dense_mass(d):
    off, cols, vals = d.RetrieveMassCSRToCPU()
    n = off.shape[0] - 1
    M = np.zeros((n, n))
    for i in range(n):
        M[i, cols[off[i]:off[i + 1]]] = vals[off[i]:off[i + 1]]
    return M


@pytest.mark.parametrize("nb", [2, 3])
def test_3243_mass_matrix_known_answer(nb):
    """MassMatrix2Beams / MassMatrix3Beams of lib_utest/utest_3243.cc."""
    L, W, H = 2.0, 1.0, 1.0
    x, y, z, conn = straight_beam(nb, L)
    d, _ = make(3243, x, y, z, conn, (L, W, H))
    M = dense_mass(d)
    ref = np.load(os.path.join(GOLDEN, f"ancf3243_mass_{nb}_beam.npy"))
    assert M.shape == ref.shape
    assert np.abs(M - ref).max() < 1e-4  # the reference's own tolerance (CSV has ~6 digits)
    assert np.abs(M - M.T).max() < 1e-9 * np.abs(M).max()
    assert np.linalg.det(M) > 0
    assert abs(M[0::4, 0::4].sum() - RHO * nb * L * W * H) < 1e-9 * RHO * nb * L * W * H  # total mass


@pytest.mark.parametrize("kind", [3243, 3443])
def test_precompute_invariants(kind):
    if kind == 3243:
        dims = (0.5, 0.1, 0.1); x, y, z, conn = straight_beam(5, dims[0]); vol = 5 * 0.5 * 0.1 * 0.1
    else:
        dims = (0.4, 0.3, 0.05); x, y, z, conn = plate(3, 2, dims[0], dims[1]); vol = 6 * 0.4 * 0.3 * 0.05
    d, s = make(kind, x, y, z, conn, dims)
    detJ = d.RetrieveDetJToCPU().reshape(conn.shape[0], -1)
    qw = stiffness_weights(kind)
    assert abs((detJ * qw).sum() - vol) < 1e-12 * vol
    M = dense_mass(d)
    assert abs(M[0::4, 0::4].sum() - RHO * vol) < 1e-10 * RHO * vol
    # at rest: F = I, P = 0, f_int = 0
    s.AssembleForce()
    f = d.RetrieveInternalForceToCPU()
    assert np.abs(f).max() < 1e-12 * E_MOD  # rounding only (forces under load are ~E * area)


@pytest.mark.parametrize("kind", [3243, 3443])
@pytest.mark.parametrize("model,damp", [("svk", False), ("svk", True), ("mr", False), ("mr", True)])
def test_force_tangent_vs_oracle(kind, model, damp):
    from oracle import oracle as orc

    if kind == 3243:
        dims = (0.5, 0.1, 0.1); x, y, z, conn = straight_beam(6, dims[0]); S = 8
    else:
        dims = (0.4, 0.3, 0.05); x, y, z, conn = plate(3, 3, dims[0], dims[1]); S = 16
    d, s = make(kind, x, y, z, conn, dims, model, damp)
    ne, nc = conn.shape[0], x.shape[0]
    rng = np.random.Generator(np.random.PCG64(2234))
    amp = np.tile([0.02 * dims[0], 0.02, 0.02, 0.02], nc // 4)
    xd = x + amp * rng.uniform(-1, 1, nc); yd = y + amp * rng.uniform(-1, 1, nc); zd = z + amp * rng.uniform(-1, 1, nc)
    v = 0.1 * rng.uniform(-1, 1, 3 * nc)
    d.UpdatePositions(xd, yd, zd)
    s.SetVelocityGuess(v)
    s.Assemble()
    f = d.RetrieveInternalForceToCPU()
    H = s.RetrieveHessianValues()

    ecoef = np.array([[4 * conn[e, a // 4] + a % 4 for a in range(S)] for e in range(ne)], dtype=np.int32)
    Q = stiffness_weights(kind).shape[0]
    gradN = d.RetrieveDsDuPreToCPU().reshape(ne, Q, 3 * S)
    detJ = d.RetrieveDetJToCPU().reshape(ne, Q)
    qw = stiffness_weights(kind)
    eta, lam_d = (1e5, 2e5) if damp else (0.0, 0.0)
    if model == "svk":
        mat = orc.Material.svk(E_MOD, NU, eta, lam_d)
    else:
        mu = E_MOD / (2 * (1 + NU)); K = E_MOD / (3 * (1 - 2 * NU))
        mat = orc.Material.mooney_rivlin(0.3 * mu, 0.2 * mu, 1.5 * K, eta, lam_d)
    coords = np.stack([xd, yd, zd], axis=1)
    if damp:
        F, P, Fd, Pv = orc.compute_p(ecoef, coords, gradN, mat, v, want_visc=True)
    else:
        F, P = orc.compute_p(ecoef, coords, gradN, mat)
    f_ref = orc.internal_force(ecoef, gradN, detJ, qw, P, nc)
    off, cols, mvals = d.RetrieveMassCSRToCPU()
    H_ref = np.zeros(9 * cols.shape[0])
    orc.hessian_mass(off, cols, mvals, H_STEP, H_ref)
    orc.tangent_assemble(ecoef, coords, gradN, detJ, qw, mat, H_STEP, off, cols, H_ref)
    assert rel_err(f, f_ref) < TOL
    assert rel_err(H, H_ref) < TOL
    # bit-reproducible
    s.Assemble()
    assert np.array_equal(H, s.RetrieveHessianValues())
    assert np.array_equal(f, d.RetrieveInternalForceToCPU())
    # the two assembly paths (element kernel + row reduce for SVK; generic row gather) agree, and so does force-only
    s.SetAssembly(0)
    s.Assemble()
    assert rel_err(s.RetrieveHessianValues(), H) < 1e-13
    assert rel_err(d.RetrieveInternalForceToCPU(), f) < 1e-13
    s.SetAssembly(1)
    s.AssembleForce()
    assert rel_err(d.RetrieveInternalForceToCPU(), f) < 1e-13


@pytest.mark.parametrize("kind", [3243, 3443])
def test_force_tangent_vs_oracle_many_elements(kind):
    """A 40 x 40 ANCF3443 plate (1 600 elements) and a 2 000-element ANCF3243 chain: full f_int and H against the oracle,
    1e-12.  Crosses the CTA-group, ticket and grid-stride boundaries the handful-of-elements meshes never reach."""
    from oracle import oracle as orc

    if kind == 3243:
        dims = (0.5, 0.1, 0.1); x, y, z, conn = straight_beam(2000, dims[0]); S = 8
    else:
        dims = (0.4, 0.3, 0.05); x, y, z, conn = plate(40, 40, dims[0], dims[1]); S = 16
    d, s = make(kind, x, y, z, conn, dims, "svk", False)
    ne, nc = conn.shape[0], x.shape[0]
    rng = np.random.Generator(np.random.PCG64(77))
    amp = np.tile([0.02 * dims[0], 0.02, 0.02, 0.02], nc // 4)
    xd = x + amp * rng.uniform(-1, 1, nc); yd = y + amp * rng.uniform(-1, 1, nc); zd = z + amp * rng.uniform(-1, 1, nc)
    d.UpdatePositions(xd, yd, zd)
    s.Assemble()
    f, H = d.RetrieveInternalForceToCPU(), s.RetrieveHessianValues()
    ecoef = orc.ancf_ecoef(kind, conn)
    LWH = np.tile(list(dims), (ne, 1))
    Binv = orc.ancf_binv(kind, [dims[0]] * ne, [dims[1]] * ne, [dims[2]] * ne)
    gradN, detJ = orc.ancf_precompute(kind, ecoef, x, y, z, LWH, Binv)  # the oracle's own set-up path
    qw = orc.ancf_stiffness_weights(kind)
    mat = orc.Material.svk(E_MOD, NU)
    coords = np.stack([xd, yd, zd], axis=1)
    _, P = orc.compute_p(ecoef, coords, gradN, mat)
    f_ref = orc.internal_force(ecoef, gradN, detJ, qw, P, nc)
    off, cols = orc.coef_pattern(ecoef, nc)
    g_off, g_cols, _ = d.RetrieveMassCSRToCPU()
    assert np.array_equal(off, g_off) and np.array_equal(cols, g_cols)
    H_ref = np.zeros(9 * cols.shape[0])
    orc.hessian_mass(off, cols, orc.ancf_mass(kind, ecoef, x, y, z, LWH, Binv, RHO, off, cols), H_STEP, H_ref)
    orc.tangent_assemble(ecoef, coords, gradN, detJ, qw, mat, H_STEP, off, cols, H_ref)
    # F = sum_a x_a grad N_a cancels positions of the order (elements along the axis) x (element length) down to O(1):
    # on the 2 000-element chain both sides lose three digits there and round differently (measured 4e-12 on f_int,
    # 5e-13 on H); the plate (40 elements along an axis) stays within the 1e-12 of the small meshes (measured 9e-14)
    tol = 10 * TOL if kind == 3243 else TOL
    assert rel_err(f, f_ref) < tol
    assert rel_err(H, H_ref) < tol


@pytest.mark.parametrize("kind", [3243, 3443])
def test_setup_path_vs_oracle(kind):
    """CalcDsDuPre / CalcMassMatrix / pattern against the oracle restatement on a curved configuration."""
    from oracle import oracle as orc

    if kind == 3243:
        dims = (0.5, 0.1, 0.2); x, y, z, conn = straight_beam(5, dims[0])
    else:
        dims = (0.4, 0.3, 0.05); x, y, z, conn = plate(3, 2, dims[0], dims[1])
    rng = np.random.Generator(np.random.PCG64(77))
    nc = x.shape[0]
    amp = np.tile([0.03 * dims[0], 0.03, 0.03, 0.03], nc // 4)
    x = x + amp * rng.uniform(-1, 1, nc); y = y + amp * rng.uniform(-1, 1, nc); z = z + amp * rng.uniform(-1, 1, nc)
    d, s = make(kind, x, y, z, conn, dims)
    ne = conn.shape[0]
    ec = orc.ancf_ecoef(kind, conn)
    Binv = orc.ancf_binv(kind, [dims[0]] * ne, [dims[1]] * ne, [dims[2]] * ne)
    LWH = np.tile(dims, (ne, 1))
    g_ref, dj_ref = orc.ancf_precompute(kind, ec, x, y, z, LWH, Binv)
    assert rel_err(d.RetrieveDsDuPreToCPU().reshape(g_ref.shape), g_ref) < TOL
    assert rel_err(d.RetrieveDetJToCPU().reshape(dj_ref.shape), dj_ref) < TOL
    off, cols, vals = d.RetrieveMassCSRToCPU()
    off_ref, cols_ref = orc.coef_pattern(ec, nc)
    assert np.array_equal(off, off_ref) and np.array_equal(cols, cols_ref)  # bit-exact
    assert rel_err(vals, orc.ancf_mass(kind, ec, x, y, z, LWH, Binv, RHO, off, cols)) < TOL
    ro, ci, _ = s.RetrieveHessianCSR(want_values=False)
    ro_ref, ci_ref = orc.dof_pattern(off_ref, cols_ref)
    assert np.array_equal(ro, ro_ref) and np.array_equal(ci, ci_ref)  # bit-exact


@pytest.mark.parametrize("kind", [3243, 3443])
def test_reference_prototype_vectors(kind):
    """GPU set-up, mass, SVK force and tangent against vectors produced by the reference's own numpy prototypes
    (test-scripts/3243-beam/f-form-3243-nesterov.py, test-scripts/3443-shell/f-form-3443-nesterov.py;
    tests/golden/make_golden_ancf.py): consistent mass, internal force of a perturbed configuration, and the exact
    derivative of the prototype's force as the tangent.  1e-11: the prototypes' quadrature tables carry 15 digits."""
    g = dict(np.load(os.path.join(GOLDEN, f"ancf{kind}_proto.npz")))
    dims = (float(g["L"]), float(g["W"]), float(g["H"]))
    conn = np.ascontiguousarray(g["conn"], dtype=np.int32)
    import tlfea

    nc = g["x"].shape[0]
    cls = tlfea.GPU_ANCF3243_Data if kind == 3243 else tlfea.GPU_ANCF3443_Data
    d = cls(nc // 4, conn.shape[0])
    d.Initialize()
    d.Setup(*dims, g["x_ref"], g["y_ref"], g["z_ref"], conn)
    d.SetDensity(float(g["rho"]))
    d.SetSVK(float(g["Emod"]), float(g["nu"]))
    d.CalcDsDuPre()
    d.CalcMassMatrix()
    h = 1.0
    s = tlfea.SyncedNewtonSolver(d, 0)
    s.SetParameters(tlfea.SyncedNewtonParams(1e-4, 0.0, 1e-6, 1e14, 1, 1, h))
    s.Setup()
    s.AnalyzeHessianSparsity()
    M = dense_mass(d)
    assert rel_err(M, g["M"]) < 1e-11
    d.UpdatePositions(g["x"], g["y"], g["z"])
    for mode in (1, 0):  # element kernel + row reduce, generic row gather
        s.SetAssembly(mode)
        s.Assemble()
        assert rel_err(d.RetrieveInternalForceToCPU(), g["f_int"]) < 1e-11
        ro, ci, Hv = s.RetrieveHessianCSR()
        Hd = np.zeros((3 * nc, 3 * nc))
        for r in range(3 * nc):
            Hd[r, ci[ro[r]:ro[r + 1]]] = Hv[ro[r]:ro[r + 1]]
        # H = M/h (x) I3 + h K_t with h = 1
        assert rel_err(Hd - np.kron(M, np.eye(3)), g["K_t"]) < 1e-11


@pytest.mark.parametrize("kind", [3243, 3443])
def test_host_pipelined_assemble_matches_the_plain_sequence(kind):
    """tlfea_newton_assemble_host_begin/_end on the ANCF elements (element kernel in chunks under the chunked H2D,
    rows reduced under the D2H): bit-identical to UpdatePositions + Assemble + Residual + Retrieve."""
    if kind == 3243:
        dims = (0.5, 0.1, 0.1); x, y, z, conn = straight_beam(37, dims[0])
    else:
        dims = (0.4, 0.3, 0.05); x, y, z, conn = plate(5, 4, dims[0], dims[1])
    d, s = make(kind, x, y, z, conn, dims, "svk", True)
    nc = x.shape[0]
    rng = np.random.Generator(np.random.PCG64(11))
    amp = np.tile([0.02 * dims[0], 0.02, 0.02, 0.02], nc // 4)
    s.SetVelocityGuess(0.1 * rng.uniform(-1, 1, 3 * nc))
    for _ in range(2):
        xd = x + amp * rng.uniform(-1, 1, nc); yd = y + amp * rng.uniform(-1, 1, nc); zd = z + amp * rng.uniform(-1, 1, nc)
        d.UpdatePositions(xd, yd, zd)
        s.Assemble()
        ng_ref, f_ref, H_ref = s.Residual(), d.RetrieveInternalForceToCPU(), s.RetrieveHessianValues()
        d.UpdatePositions(x, y, z)
        ng, f = s.AssembleFromHost(xd, yd, zd)
        assert ng == ng_ref and np.array_equal(f, f_ref)
        assert np.array_equal(s.RetrieveHessianValues(), H_ref)
