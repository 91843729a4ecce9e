"""Error behaviour at the boundary (SURVEY.md 8b "Errors"): where the reference prints to std::cerr and returns
(call order, sizes; e.g. FEAT10Data.cuh:442-445,539-543,637-640) the C ABI returns TLFEA_ERR_INVALID with a message
and leaves the handle usable; nothing falls back to the CPU and nothing aborts the process."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _mesh():
    from tlfea import meshgen

    return meshgen.kuhn_block(2)


def test_call_order_and_bad_arguments_are_reported():
    import tlfea

    X, E = _mesh()
    with pytest.raises(tlfea.TlfeaError) as ei:
        tlfea.GPU_FEAT10_Data(0, X.shape[0])  # empty mesh
    assert ei.value.code == -1
    with pytest.raises(tlfea.TlfeaError):
        tlfea.ElementBase(99, 4, 4)  # unknown element type

    d = tlfea.GPU_FEAT10_Data(E.shape[0], X.shape[0])
    d.Initialize()
    with pytest.raises(tlfea.TlfeaError) as ei:
        d.CalcDnDuPre()  # before Setup
    assert ei.value.code == -1 and str(ei.value)
    bad = E.copy()
    bad[3, 7] = X.shape[0]  # node id out of range
    with pytest.raises(tlfea.TlfeaError):
        d.Setup(*tlfea.tet5pt_quadrature(), X[:, 0], X[:, 1], X[:, 2], bad)
    d.Setup(*tlfea.tet5pt_quadrature(), X[:, 0], X[:, 1], X[:, 2], E)
    d.SetDensity(2700.0)
    d.SetSVK(7e8, 0.33)
    with pytest.raises(tlfea.TlfeaError):
        d.SetNodalFixed(np.array([0, X.shape[0]], dtype=np.int32))  # fixed node out of range
    s = tlfea.SyncedNewtonSolver(d, 0)
    s.SetParameters(tlfea.SyncedNewtonParams(1e-4, 0.0, 1e-6, 1e14, 1, 1, 1e-3))
    with pytest.raises(tlfea.TlfeaError) as ei:
        s.Assemble()  # neither CalcDnDuPre nor AnalyzeHessianSparsity has run
    assert "must run before" in str(ei.value)
    with pytest.raises(tlfea.TlfeaError) as ei:
        s.AssembleForce()  # force only: no pattern needed, but the reference gradients are
    assert "CalcDnDuPre" in str(ei.value)
    d.CalcDnDuPre()
    with pytest.raises(tlfea.TlfeaError) as ei:
        s.Assemble()  # AnalyzeHessianSparsity has not run
    assert "AnalyzeHessianSparsity" in str(ei.value)
    d.CalcMassMatrix()
    s.Setup()
    s.AnalyzeHessianSparsity()
    s.SetParameters(tlfea.SyncedNewtonParams(1e-4, 0.0, 1e-6, 1e14, 1, 1, 0.0))
    with pytest.raises(tlfea.TlfeaError):
        s.Assemble()  # time_step must be positive
    # the handles survive every refusal: the same objects assemble once the calls are in order
    s.SetParameters(tlfea.SyncedNewtonParams(1e-4, 0.0, 1e-6, 1e14, 1, 1, 1e-3))
    s.Assemble()
    assert np.isfinite(s.RetrieveHessianValues()).all()
