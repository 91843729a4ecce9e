"""tlfea_dot_noalloc (include/tlfea_b200.h): the allocation-free variant of the deterministic dot product.  It has no
counterpart in the reference (added for callers that cannot cudaMalloc); the
check is against the existing entry point that runs the same kernels with cudaMalloc'ed scratch:
tlfea_newton_residual's ||g||.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-12  # the relative tolerance of the other GPU parity tests; the two paths in fact agree bitwise


def _solver(n, seed):
    import tlfea
    from tlfea import meshgen

    X, E = meshgen.kuhn_block(n)
    x = meshgen.perturb(X, 0.05 / n, seed)
    d = tlfea.GPU_FEAT10_Data(E.shape[0], X.shape[0])
    d.Initialize()
    d.Setup(*tlfea.tet5pt_quadrature(), X[:, 0], X[:, 1], X[:, 2], E)
    d.SetDensity(2700.0)
    d.SetSVK(7e8, 0.33)
    d.CalcDnDuPre()
    d.CalcMassMatrix()
    d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
    s = tlfea.SyncedNewtonSolver(d, 0)
    s.SetParameters(tlfea.SyncedNewtonParams(1e-4, 0.0, 1e-6, 1e14, 1, 1, 1e-3))
    s.Setup()
    s.AnalyzeHessianSparsity()
    s.Assemble()
    return d, s, 3 * X.shape[0]


def test_dot_noalloc_matches_handle_path_twice_in_a_row():
    import torch
    from tlfea._lib import check, lib

    torch.cuda.set_device(0)
    out = torch.zeros(1, dtype=torch.float64, device="cuda:0")  # the caller's output buffer
    keep = []
    # two calls in a row on different data and different lengths: the second sees the first one's partials in the
    # shared static scratch and must not be affected by them
    for n_block, seed in ((4, 11), (3, 12)):
        d, s, n = _solver(n_block, seed)
        keep.append((d, s))
        norm_ref = C.c_double()
        check(lib.tlfea_newton_residual(s._h, C.byref(norm_ref)))  # existing entry point (cudaMalloc'ed scratch)
        g_ptr = lib.tlfea_newton_devptr_g(s._h)
        d.Sync()
        check(lib.tlfea_dot_noalloc(g_ptr, g_ptr, n, out.data_ptr(), None))
        torch.cuda.synchronize()
        got = float(np.sqrt(out.item()))
        assert norm_ref.value > 0.0
        assert abs(got - norm_ref.value) <= TOL * norm_ref.value, (got, norm_ref.value)
        # against a plain host sum of the same vector as well
        g = s.RetrieveGradient()
        assert abs(got - np.linalg.norm(g)) <= 1e-10 * np.linalg.norm(g)


def test_dot_noalloc_general_pair_and_limits():
    import torch
    from tlfea._lib import TlfeaError, check, lib

    torch.cuda.set_device(0)
    gen = torch.Generator(device="cpu").manual_seed(5)
    a = torch.randn(1_000_003, dtype=torch.float64, generator=gen)
    b = torch.randn(1_000_003, dtype=torch.float64, generator=gen)
    da, db = a.cuda(), b.cuda()
    out = torch.full((1,), 123.0, dtype=torch.float64, device="cuda:0")
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    vals = []
    for _ in range(2):  # a caller-owned stream; same inputs twice -> bitwise the same
        check(lib.tlfea_dot_noalloc(da.data_ptr(), db.data_ptr(), a.numel(), out.data_ptr(), st.cuda_stream))
        st.synchronize()
        vals.append(out.item())
    assert vals[0] == vals[1]
    ref = float(np.dot(a.numpy().astype(np.longdouble), b.numpy().astype(np.longdouble)))
    scale = float(np.abs(a.numpy() * b.numpy()).sum())
    assert abs(vals[0] - ref) <= TOL * scale
    # empty input: 0
    check(lib.tlfea_dot_noalloc(da.data_ptr(), db.data_ptr(), 0, out.data_ptr(), None))
    torch.cuda.synchronize()
    assert out.item() == 0.0
    # beyond the stated maximum: refused with a message, nothing launched
    with pytest.raises(TlfeaError) as ei:
        check(lib.tlfea_dot_noalloc(da.data_ptr(), db.data_ptr(), (1 << 32) + 1, out.data_ptr(), None))
    assert ei.value.code == -1 and "TLFEA_NOALLOC_DOT_MAX_N" in str(ei.value)
    with pytest.raises(TlfeaError):
        check(lib.tlfea_dot_noalloc(None, db.data_ptr(), 8, out.data_ptr(), None))
