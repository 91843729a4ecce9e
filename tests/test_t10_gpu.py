"""GPU parity tests, FEAT10: the CUDA path (through the C ABI) against the oracle and the golden
vectors on the same inputs.  Integer work (CSR pattern) bit-exact; FP64 values within 1e-12
relative to the largest entry of the compared array (north_star tolerance)."""
import numpy as np
import pytest

from conftest import rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-12
E_MOD, NU, RHO, H_STEP = 7e8, 0.33, 2700.0, 1e-3


def _mat(orc, model, damp):
    eta, lam_d = (1e5, 2e5) if damp else (0.0, 0.0)
    if model == "svk":
        return orc.Material.svk(E_MOD, NU, eta, lam_d)
    mu = E_MOD / (2 * (1 + NU))
    K = E_MOD / (3 * (1 - 2 * NU))
    return orc.Material.mooney_rivlin(0.3 * mu, 0.2 * mu, 1.5 * K, eta, lam_d)


def _make(X, E, model="svk", damp=False, fixed=None, alm_rho=1e14):
    import tlfea

    d = tlfea.GPU_FEAT10_Data(E.shape[0], X.shape[0])
    d.Initialize()
    d.Setup(*tlfea.tet5pt_quadrature(), X[:, 0], X[:, 1], X[:, 2], E)
    d.SetDensity(RHO)
    if model == "svk":
        d.SetSVK(E_MOD, NU)
    else:
        mu = E_MOD / (2 * (1 + NU))
        K = E_MOD / (3 * (1 - 2 * NU))
        d.SetMooneyRivlin(0.3 * mu, 0.2 * mu, 1.5 * K)
    if damp:
        d.SetDamping(1e5, 2e5)
    nfix = 0
    if fixed is not None:
        d.SetNodalFixed(fixed)
        nfix = 3 * len(fixed)
    d.CalcDnDuPre()
    d.CalcMassMatrix()
    s = tlfea.SyncedNewtonSolver(d, nfix)
    s.SetParameters(tlfea.SyncedNewtonParams(1e-4, 0.0, 1e-6, alm_rho, 1, 1, H_STEP))
    s.Setup()
    s.AnalyzeHessianSparsity()
    return d, s


@pytest.fixture(scope="module")
def mesh():
    from tlfea import meshgen

    X, E = meshgen.kuhn_block(4)
    x = meshgen.perturb(X, 0.05 / 4, 1234)
    v = 0.1 * np.random.Generator(np.random.PCG64(1235)).uniform(-1, 1, size=X.shape)
    return X, E, x, v


def test_precompute_mass_pattern_vs_oracle(mesh):
    from oracle import oracle as orc

    X, E, x, v = mesh
    d, s = _make(X, E)
    o = orc.T10Oracle(E, X)
    assert rel_err(d.RetrieveDnDuPreToCPU().reshape(o.grad_N.shape), o.grad_N) < TOL
    assert rel_err(d.RetrieveDetJToCPU().reshape(o.detJ.shape), o.detJ) < TOL
    off, cols, vals = d.RetrieveMassCSRToCPU()
    assert np.array_equal(off, o.off) and np.array_equal(cols, o.cols)  # bit-exact
    assert rel_err(vals, o.mass(RHO)) < TOL
    ro, ci, _ = s.RetrieveHessianCSR(want_values=False)
    ro_ref, ci_ref = orc.dof_pattern(o.off, o.cols)
    assert np.array_equal(ro, ro_ref) and np.array_equal(ci, ci_ref)  # bit-exact


@pytest.mark.parametrize("model,damp", [("svk", False), ("svk", True), ("mr", False), ("mr", True)])
def test_force_tangent_hessian_vs_oracle(mesh, model, damp):
    from oracle import oracle as orc

    X, E, x, v = mesh
    fixed = np.nonzero(X[:, 0] == 0.0)[0].astype(np.int32)
    d, s = _make(X, E, model, damp, fixed)
    d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
    s.SetVelocityGuess(v.reshape(-1))
    s.Assemble()
    o = orc.T10Oracle(E, X)
    mat = _mat(orc, model, damp)
    f_ref, F_ref, P_ref = o.force(x, mat, v.reshape(-1) if damp else None)
    H_ref = o.hessian(x, mat, H_STEP, RHO, fixed, 1e14)
    assert rel_err(d.RetrieveInternalForceToCPU(), f_ref) < TOL
    # the ALM penalty h^2*rho = 1e8 sits on a few diagonal entries; compare those and the rest separately
    H = s.RetrieveHessianValues()
    big = np.abs(H_ref) > 0.5 * H_STEP**2 * 1e14
    assert big.sum() == 3 * len(fixed)
    assert rel_err(H[big], H_ref[big]) < TOL
    assert rel_err(H[~big], H_ref[~big]) < TOL


def test_calc_p_and_accumulating_force(mesh):
    """CalcP stores elastic F,P; CalcInternalForce ACCUMULATES (FEAT10Data.cu:294-333)."""
    from oracle import oracle as orc

    X, E, x, v = mesh
    d, s = _make(X, E)
    d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
    d.CalcP()
    o = orc.T10Oracle(E, X)
    f_ref, F_ref, P_ref = o.force(x, _mat(orc, "svk", False))
    assert rel_err(d.RetrieveDeformationGradientToCPU().reshape(F_ref.shape), F_ref) < TOL
    assert rel_err(d.RetrievePFromFToCPU().reshape(P_ref.shape), P_ref) < TOL
    d.ClearInternalForce()
    d.CalcInternalForce()
    assert rel_err(d.RetrieveInternalForceToCPU(), f_ref) < TOL
    d.CalcInternalForce()
    assert rel_err(d.RetrieveInternalForceToCPU(), 2 * f_ref) < TOL


def test_golden_prototype_vectors(golden_t10):
    """Straight against the reference numpy prototype's outputs (tests/golden)."""
    g = golden_t10
    d, s = _make(g["X"], g["E"])
    x = g["x"]
    d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
    s.Assemble()
    assert rel_err(d.RetrieveInternalForceToCPU(), g["f_int"]) < TOL
    ro, ci, H = s.RetrieveHessianCSR()
    n = ro.shape[0] - 1
    D = np.zeros((n, n))
    for r in range(n):
        D[r, ci[ro[r]:ro[r + 1]]] = H[ro[r]:ro[r + 1]]
    want = g["M"] * (RHO / float(g["rho"])) / H_STEP + H_STEP * g["K"]
    assert rel_err(D, want) < TOL


def test_bit_reproducible(mesh):
    X, E, x, v = mesh
    d, s = _make(X, E)
    d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
    s.Assemble()
    H1, f1 = s.RetrieveHessianValues(), d.RetrieveInternalForceToCPU()
    for _ in range(3):
        s.Assemble()
        assert np.array_equal(H1, s.RetrieveHessianValues())
        assert np.array_equal(f1, d.RetrieveInternalForceToCPU())


def test_properties_at_scale():
    """Size-independent properties on a mesh the oracle would not finish quickly (48^3 cells,
    663 552 tets): symmetry of H, H*(translation) = (M/h)*(translation), sum f_int = 0."""
    import torch
    import tlfea
    from tlfea import meshgen

    n = 48
    X, E = meshgen.kuhn_block(n)
    x = meshgen.perturb(X, 0.05 / n, 1234)
    d, s = _make(X, E)
    d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
    s.Assemble()
    f = d.RetrieveInternalForceToCPU().reshape(-1, 3)
    assert np.abs(f.sum(0)).max() < 1e-9 * np.abs(f).max()
    nd = s.n_dofs
    t = torch.zeros(nd, dtype=torch.float64, device="cuda")
    t[0::3] = 1.0
    y = torch.empty_like(t)
    tlfea._lib.check(tlfea.lib.tlfea_newton_spmv(s._h, t.data_ptr(), y.data_ptr()))
    d.Sync()
    off, cols, mvals = d.RetrieveMassCSRToCPU()
    rowsum = np.add.reduceat(mvals, off[:-1]) / H_STEP
    yy = y.cpu().numpy().reshape(-1, 3)
    assert rel_err(yy[:, 0], rowsum) < 1e-9
    assert np.abs(yy[:, 1:]).max() < 1e-9 * np.abs(rowsum).max()
    # symmetry: x^T H y == y^T H x for random vectors
    g = torch.Generator(device="cuda").manual_seed(1)
    a = torch.rand(nd, dtype=torch.float64, device="cuda", generator=g)
    b = torch.rand(nd, dtype=torch.float64, device="cuda", generator=g)
    Ha, Hb = torch.empty_like(a), torch.empty_like(a)
    tlfea._lib.check(tlfea.lib.tlfea_newton_spmv(s._h, a.data_ptr(), Ha.data_ptr()))
    tlfea._lib.check(tlfea.lib.tlfea_newton_spmv(s._h, b.data_ptr(), Hb.data_ptr()))
    d.Sync()
    lhs, rhs = torch.dot(b, Ha).item(), torch.dot(a, Hb).item()
    assert abs(lhs - rhs) < 1e-10 * abs(lhs)


def test_newton_step_converges(golden_t10):
    """One implicit step on the 36-tet beam: PCG-based Newton drives ||g|| below tolerance."""
    import tlfea

    g = golden_t10
    X = g["X"]
    fixed = np.nonzero(X[:, 0] == 0.0)[0].astype(np.int32)
    d, s = _make(X, g["E"], fixed=fixed, alm_rho=1e12)
    fext = np.zeros(3 * X.shape[0])
    fext[2::3] = -100.0
    d.SetExternalForce(fext)
    s.SetParameters(tlfea.SyncedNewtonParams(1e-4, 0.0, 1e-6, 1e12, 3, 10, H_STEP))
    s.SetPCG(2000, 1e-12)
    s.Solve()
    st = s.LastStats()
    assert st["norm_g"] < 1e-4, st
    assert st["norm_c"] < 1e-6, st


@pytest.mark.parametrize("model", ["svk", "mr"])
@pytest.mark.parametrize("damp", [False, True])
def test_assembly_kernels_agree(mesh, damp, model):
    """The two deterministic assembly paths (1: element kernel + row reduce, 0: generic warp-per-row gather)
    give the same H and f_int to rounding, and each is bit-reproducible."""
    X, E, x, v = mesh
    fixed = np.nonzero(X[:, 0] == 0.0)[0].astype(np.int32)
    d, s = _make(X, E, model, damp, fixed)
    d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
    s.SetVelocityGuess(v.reshape(-1))
    s.SetAssembly(1)
    s.Assemble()
    H1, f1 = s.RetrieveHessianValues(), d.RetrieveInternalForceToCPU()
    s.Assemble()
    assert np.array_equal(H1, s.RetrieveHessianValues())
    assert np.array_equal(f1, d.RetrieveInternalForceToCPU())
    s.SetAssembly(0)
    s.Assemble()
    H0, f0 = s.RetrieveHessianValues(), d.RetrieveInternalForceToCPU()
    s.Assemble()
    assert np.array_equal(H0, s.RetrieveHessianValues())
    big = np.abs(H0) > 0.5 * H_STEP**2 * 1e14
    tol = 1e-14 if model == "svk" else 1e-13  # the MR tensor entries cancel to ~1e-2 of their terms
    assert rel_err(H1[~big], H0[~big]) < tol
    assert rel_err(H1[big], H0[big]) < tol
    assert rel_err(f1, f0) < tol


@pytest.mark.parametrize("case", ["permuted", "odd_count", "odd_count_mr", "unused_nodes", "single_element"])
def test_general_numbering_and_ragged_meshes(case):
    """Nothing in the path may depend on the structured numbering of the synthetic block: random element and
    node renumbering, an odd element count (the element kernel works on pairs and groups), nodes that no element
    touches (empty H rows), and a single element."""
    from oracle import oracle as orc
    from tlfea import meshgen

    X, E = meshgen.kuhn_block(3)
    rng = np.random.Generator(np.random.PCG64(99))
    if case == "permuted":
        pn = rng.permutation(X.shape[0])          # new id of old node
        Xp = np.empty_like(X); Xp[pn] = X
        E = pn[E][rng.permutation(E.shape[0])].astype(np.int32)
        X = Xp
    elif case.startswith("odd_count"):
        E = E[:E.shape[0] - 5]                     # 157 elements
    elif case == "unused_nodes":
        E = E[rng.permutation(E.shape[0])[:40]]    # many nodes lose all their elements
    else:
        E = E[7:8]
    x = meshgen.perturb(X, 0.05 / 3, 4321)
    model = "mr" if case.endswith("_mr") else "svk"
    d, s = _make(X, np.ascontiguousarray(E), model)
    d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
    s.Assemble()
    o = orc.T10Oracle(np.ascontiguousarray(E), X)
    mat = _mat(orc, model, False)
    f_ref, _, _ = o.force(x, mat)
    H_ref = o.hessian(x, mat, H_STEP, RHO)
    off, cols, _ = d.RetrieveMassCSRToCPU()
    assert np.array_equal(off, o.off) and np.array_equal(cols, o.cols)
    assert rel_err(d.RetrieveInternalForceToCPU(), f_ref) < TOL
    assert rel_err(s.RetrieveHessianValues(), H_ref) < TOL


def test_full_size_properties():
    """Size-independent properties at a size the oracle cannot check in seconds (n = 40: 384 000 tets, 1.9 M Gauss
    points): internal forces self-equilibrate (sum f = 0, sum x cross f = 0: frame indifference of SVK),
    H is symmetric (y.Hz = z.Hy), K annihilates rigid translations (H t = (M/h) t), H is reproducible bit for bit."""
    import torch
    from tlfea import meshgen

    n = 40
    X, E = meshgen.kuhn_block(n)
    x = meshgen.perturb(X, 0.05 / n, 1234)
    d, s = _make(X, E)
    d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
    s.Assemble()
    f = d.RetrieveInternalForceToCPU().reshape(-1, 3)
    scale = np.abs(f).sum()
    assert np.abs(f.sum(axis=0)).max() < 1e-12 * scale
    assert np.abs(np.cross(x, f).sum(axis=0)).max() < 1e-12 * scale
    nd = 3 * X.shape[0]
    g = torch.Generator(device="cpu").manual_seed(5)
    y = torch.rand(nd, dtype=torch.float64, generator=g).cuda()
    z = torch.rand(nd, dtype=torch.float64, generator=g).cuda()
    Hy, Hz = torch.empty_like(y), torch.empty_like(z)
    torch.cuda.synchronize()
    s.SpMV(y.data_ptr(), Hy.data_ptr())
    s.SpMV(z.data_ptr(), Hz.data_ptr())
    d.Sync()
    a, b = torch.dot(z, Hy).item(), torch.dot(y, Hz).item()
    assert abs(a - b) < 1e-12 * abs(a)
    # rigid translation: the tangent part vanishes, the mass part remains
    t = torch.zeros(nd, dtype=torch.float64).cuda(); t[0::3] = 1.0
    Ht = torch.empty_like(t)
    torch.cuda.synchronize()
    s.SpMV(t.data_ptr(), Ht.data_ptr())
    d.Sync()
    off, cols, mvals = d.RetrieveMassCSRToCPU()
    row_sum = np.add.reduceat(mvals, off[:-1]) / H_STEP
    ht = Ht.cpu().numpy().reshape(-1, 3)
    hmax = np.abs(s.RetrieveHessianValues()).max()
    assert np.abs(ht[:, 0] - row_sum).max() < 1e-11 * hmax
    assert np.abs(ht[:, 1:]).max() < 1e-11 * hmax
    H1 = s.RetrieveHessianValues()
    s.Assemble()
    assert np.array_equal(H1, s.RetrieveHessianValues())


def test_update_nodal_fixed_moves_the_penalty_entries(mesh):
    """UpdateNodalFixed (FEAT10Data.cu:751-839) under a live solver: same count, different pinned set."""
    from oracle import oracle as orc

    X, E, x, v = mesh
    face0 = np.nonzero(X[:, 0] == 0.0)[0].astype(np.int32)
    face1 = np.nonzero(X[:, 0] == X[:, 0].max())[0].astype(np.int32)
    assert len(face0) == len(face1)
    d, s = _make(X, E, fixed=face0)
    d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
    o = orc.T10Oracle(E, X)
    mat = _mat(orc, "svk", False)
    for fixed in (face0, face1, face0):
        d.UpdateNodalFixed(fixed)
        s.Assemble()
        H_ref = o.hessian(x, mat, H_STEP, RHO, fixed, 1e14)
        H = s.RetrieveHessianValues()
        big = np.abs(H_ref) > 0.5 * H_STEP**2 * 1e14
        assert big.sum() == 3 * len(fixed)
        assert np.array_equal(np.abs(H) > 0.5 * H_STEP**2 * 1e14, big)
        assert rel_err(H[big], H_ref[big]) < TOL and rel_err(H[~big], H_ref[~big]) < TOL
        d.CalcConstraintData()
        c = d.RetrieveConstraintDataToCPU()
        assert rel_err(c, (x - X)[fixed].reshape(-1)) < TOL
    with pytest.raises(Exception):  # a different count cannot be absorbed by the bound solver
        d.UpdateNodalFixed(face0[:-1])
        s.Assemble()


def test_diag_blocks_match_the_assembled_hessian(mesh):
    """tlfea_newton_get_diag_blocks: the per-node 3x3 blocks a VBD sweep needs are exactly H's diagonal blocks."""
    import scipy.sparse as sp

    X, E, x, v = mesh
    fixed = np.nonzero(X[:, 0] == 0.0)[0].astype(np.int32)
    d, s = _make(X, E, "svk", False, fixed)
    d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
    s.Assemble()
    ro, ci, va = s.RetrieveHessianCSR()
    A = sp.csr_matrix((va, ci, ro), shape=(s.n_dofs, s.n_dofs)).tobsr(blocksize=(3, 3))
    D = s.RetrieveHessianDiagBlocks()
    nn = X.shape[0]
    ref = np.zeros((nn, 3, 3))
    for i in range(nn):
        for k in range(A.indptr[i], A.indptr[i + 1]):
            if A.indices[k] == i:
                ref[i] = A.data[k]
    assert np.array_equal(D, ref)


@pytest.mark.parametrize("model,damp,permute,chunks", [("svk", False, False, 4), ("svk", True, False, 4),
                                                       ("mr", False, False, 4), ("svk", False, True, 4),
                                                       ("svk", False, False, 1), ("svk", False, True, 3)])
def test_host_pipelined_assemble_matches_the_plain_sequence(model, damp, permute, chunks):
    """tlfea_newton_assemble_host_begin/_end (chunked H2D overlapped with the element kernel, H rows reduced on a
    second stream while f_int goes back to the host) against UpdatePositions + Assemble + Residual +
    RetrieveInternalForceToCPU: bit-identical H, f_int and ||g||, also when the node numbering is scrambled (every
    element chunk then needs every coordinate chunk) and when it is called repeatedly with changing positions."""
    import torch
    from tlfea import meshgen

    X, E = meshgen.kuhn_block(6)
    if permute:
        perm = np.random.Generator(np.random.PCG64(7)).permutation(X.shape[0])
        inv = np.empty_like(perm)
        inv[perm] = np.arange(X.shape[0])
        X, E = X[perm], inv[E].astype(np.int32)
    fixed = np.nonzero(X[:, 0] == 0.0)[0].astype(np.int32)
    d, s = _make(X, E, model, damp, fixed)
    s.SetPatchChunks(chunks)  # the automatic choice gives a mesh this small one chunk: force the chunked pipeline
    if damp:
        s.SetVelocityGuess(0.1 * np.random.Generator(np.random.PCG64(1235)).uniform(-1, 1, size=3 * X.shape[0]))
    pinned = [torch.empty(X.shape[0], dtype=torch.float64).pin_memory().numpy() for _ in range(3)]
    out = torch.empty(3 * X.shape[0], dtype=torch.float64).pin_memory().numpy()
    for seed in (1234, 99, 5):
        x = meshgen.perturb(X, 0.05 / 6, seed)
        d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
        s.Assemble()
        ng_ref = s.Residual()
        f_ref = d.RetrieveInternalForceToCPU()
        H_ref = s.RetrieveHessianValues()
        d.UpdatePositions(X[:, 0], X[:, 1], X[:, 2])  # make sure the pipelined call really uploads
        for k in range(3):
            pinned[k][:] = x[:, k]
        ng, f = s.AssembleFromHost(pinned[0], pinned[1], pinned[2], out)
        assert ng == ng_ref
        assert np.array_equal(f, f_ref)
        assert np.array_equal(s.RetrieveHessianValues(), H_ref)
        gx, gy, gz = d.RetrievePositionToCPU()
        assert np.array_equal(gx, x[:, 0]) and np.array_equal(gz, x[:, 2])
    # pageable host arrays work too (the copy is then staged by the driver)
    ng2, f2 = s.AssembleFromHost(x[:, 0].copy(), x[:, 1].copy(), x[:, 2].copy())
    assert ng2 == ng_ref and np.array_equal(f2, f_ref)
    # generic row-gather path: same call, no overlap, same numbers as its own plain sequence
    s.SetAssembly(0)
    s.Assemble()
    H0, f0 = s.RetrieveHessianValues(), d.RetrieveInternalForceToCPU()
    ng0, f0b = s.AssembleFromHost(pinned[0], pinned[1], pinned[2])
    assert np.array_equal(f0b, f0) and np.array_equal(s.RetrieveHessianValues(), H0)
    # misuse: end without begin
    import tlfea

    with pytest.raises(tlfea.TlfeaError):
        s.AssembleFromHostEnd()
