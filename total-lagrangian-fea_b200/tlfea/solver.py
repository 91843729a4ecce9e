"""Host-side mirror of SyncedNewtonSolver (lib_src/solvers/SyncedNewton.{cu,cuh}) over the C ABI."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import NewtonParams, c_dp, c_ip, check, lib
from .elements import ElementBase


def SyncedNewtonParams(inner_atol, inner_rtol, outer_tol, rho, max_outer, max_inner, time_step) -> NewtonParams:
    """SyncedNewtonParams {inner_atol, inner_rtol, outer_tol, rho, max_outer, max_inner, time_step}
    (SyncedNewton.cuh:29-33)."""
    return NewtonParams(inner_atol, inner_rtol, outer_tol, rho, int(max_outer), int(max_inner), time_step)


def _d(a):
    return a.ctypes.data_as(c_dp)


class SolverBase:
    """SolverBase (lib_src/solvers/SolverBase.h:16-22)."""

    def Solve(self):
        raise NotImplementedError

    def SetParameters(self, params):
        raise NotImplementedError


class SyncedNewtonSolver(SolverBase):
    def __init__(self, data: ElementBase, n_constraints: int):
        self.data = data
        self._h = _lib.c_vp()
        check(lib.tlfea_newton_create(data._h, int(n_constraints), C.byref(self._h)))
        self.n_dofs = lib.tlfea_newton_n_dofs(self._h)

    def Destroy(self):
        """Frees H, the staging records and the Krylov vectors (the reference frees in the destructor,
        SyncedNewton.cuh:151-204)."""
        if self._h:
            lib.tlfea_newton_destroy(self._h)
            self._h = _lib.c_vp()

    def __del__(self):
        try:
            self.Destroy()
        except Exception:
            pass

    # ---- reference API ---------------------------------------------------
    def Setup(self):
        check(lib.tlfea_newton_setup(self._h))

    def SetParameters(self, params: NewtonParams):
        self._params = params
        check(lib.tlfea_newton_set_params(self._h, C.byref(params)))

    def AnalyzeHessianSparsity(self):
        check(lib.tlfea_newton_analyze_sparsity(self._h))

    def SetFixedSparsityPattern(self, fixed: bool):
        check(lib.tlfea_newton_set_fixed_pattern(self._h, int(bool(fixed))))

    def Solve(self):
        """One implicit step (OneStepNewtonCuDSS, SyncedNewton.cu:909-1394)."""
        check(lib.tlfea_newton_solve_step(self._h))

    def GetVelocityGuessDevicePtr(self):
        return lib.tlfea_newton_devptr_v(self._h)

    # ---- the benchmarked unit and its inspection hooks ---------------------
    def Assemble(self):
        check(lib.tlfea_newton_assemble(self._h))

    def AssembleForce(self):
        check(lib.tlfea_newton_assemble_force(self._h))

    def AssembleFromHostBegin(self, x, y, z, out=None):
        """UpdatePositions(x, y, z) + Assemble with the H2D copies overlapped with the element kernel
        (tlfea_newton_assemble_host_begin).  Returns without waiting; x, y, z must stay alive (and unchanged)
        until AssembleFromHostEnd returns -- pass pinned float64 arrays for a truly asynchronous copy.
        out: the (pinned, contiguous float64) array AssembleFromHostEnd will be given; known here, finished parts of
        f_int travel back under the remaining kernels (single-pass FEAT10 path; nothing may touch f_int in between except the
        library's own halo sum, whose shared ranges AssembleFromHostEnd sends again)."""
        n = self.data.n_coef
        if out is not None and not (out.dtype == np.float64 and out.flags["C_CONTIGUOUS"] and out.shape == (3 * n,)):
            out = None
        check(lib.tlfea_newton_set_host_result_buffer(self._h, _d(out) if out is not None else None, 3 * n))
        self._pipe_in = tuple(np.ascontiguousarray(a, dtype=np.float64) for a in (x, y, z))
        check(lib.tlfea_newton_assemble_host_begin(self._h, *(_d(a) for a in self._pipe_in), n))

    def AssembleFromHostEnd(self, out=None, want_norm=True):
        """Residual() + RetrieveInternalForceToCPU(out) while H is still being reduced; returns (||g||, f_int)
        once f_int, ||g|| and H are all complete."""
        n = 3 * self.data.n_coef
        if out is None:
            out = np.empty(n)
        ng = C.c_double(0.0)
        check(lib.tlfea_newton_assemble_host_end(self._h, _d(out), n, C.byref(ng) if want_norm else None))
        self._pipe_in = None
        return ng.value, out

    def AssembleFromHost(self, x, y, z, out=None):
        self.AssembleFromHostBegin(x, y, z, out)
        return self.AssembleFromHostEnd(out)

    def Residual(self) -> float:
        out = C.c_double(0.0)
        check(lib.tlfea_newton_residual(self._h, C.byref(out)))
        return out.value

    def SetPCG(self, max_iters: int, rel_tol: float):
        check(lib.tlfea_newton_set_pcg(self._h, int(max_iters), float(rel_tol)))

    def LastStats(self):
        a, b = C.c_int(0), C.c_int(0)
        g, c = C.c_double(0), C.c_double(0)
        check(lib.tlfea_newton_last_stats(self._h, C.byref(a), C.byref(b), C.byref(g), C.byref(c)))
        rel, conv = C.c_double(0), C.c_int(1)
        check(lib.tlfea_newton_last_linear_solve(self._h, C.byref(rel), C.byref(conv)))
        return {"newton_iters": a.value, "pcg_iters": b.value, "norm_g": g.value, "norm_c": c.value,
                "pcg_rel_residual": rel.value, "pcg_converged": bool(conv.value)}

    def nnz(self) -> int:
        return lib.tlfea_newton_nnz(self._h)

    def RetrieveHessianCSR(self, want_cols=True, want_values=True):
        nnz = self.nnz()
        ro = np.empty(self.n_dofs + 1, dtype=np.int64)
        ci = np.empty(nnz, dtype=np.int32) if want_cols else None
        va = np.empty(nnz) if want_values else None
        check(lib.tlfea_newton_get_hessian_csr(
            self._h, ro.ctypes.data_as(_lib.c_i64p),
            None if ci is None else ci.ctypes.data_as(c_ip), None if va is None else _d(va)))
        return ro, ci, va

    def RetrieveHessianDiagBlocks(self):
        """[n_coef, 3, 3]: the per-node diagonal blocks of H (VBD / block-Jacobi consumers)."""
        out = np.empty(3 * self.n_dofs)
        check(lib.tlfea_newton_get_diag_blocks(self._h, _d(out), out.shape[0]))
        return out.reshape(-1, 3, 3)

    def RetrieveHessianValues(self):
        va = np.empty(self.nnz())
        check(lib.tlfea_newton_get_hessian_csr(self._h, None, None, _d(va)))
        return va

    def RetrieveGradient(self):
        out = np.empty(self.n_dofs)
        check(lib.tlfea_newton_get_g(self._h, _d(out), self.n_dofs))
        return out

    def RetrieveVelocity(self):
        out = np.empty(self.n_dofs)
        check(lib.tlfea_newton_get_v(self._h, _d(out), self.n_dofs))
        return out

    def SetVelocityGuess(self, v):
        v = np.ascontiguousarray(v, dtype=np.float64)
        check(lib.tlfea_newton_set_v(self._h, _d(v), v.shape[0]))

    def SetVelocityPrev(self, v):
        v = np.ascontiguousarray(v, dtype=np.float64)
        check(lib.tlfea_newton_set_v_prev(self._h, _d(v), v.shape[0]))

    def Timing(self, reset=False):
        ms, calls = C.c_double(0), C.c_int(0)
        check(lib.tlfea_newton_timing(self._h, C.byref(ms), C.byref(calls), int(reset)))
        return ms.value, calls.value

    def SetAssembly(self, mode: int):
        """1 = element kernel + row reduce (default where supported: FEAT10 + SVK), 0 = generic warp-per-row gather."""
        check(lib.tlfea_newton_set_assembly(self._h, int(mode)))

    def SetPatchWarps(self, warps: int):
        check(lib.tlfea_newton_set_patch_warps(self._h, int(warps)))

    def SetPatchChunks(self, chunks: int):
        """id-range chunks of the patch launch order: 0 = automatic, 1..16 = exactly that many"""
        check(lib.tlfea_newton_set_patch_chunks(self._h, int(chunks)))

    def SetAffineCompression(self, on: bool):
        check(lib.tlfea_newton_set_affine_compression(self._h, int(bool(on))))

    def PatchStats(self):
        import numpy as np

        out = np.zeros(8, dtype=np.int64)
        check(lib.tlfea_newton_patch_stats(self._h, out.ctypes.data_as(C.POINTER(C.c_int64))))
        return dict(patches=int(out[0]), colours=int(out[1]), warps=int(out[2]), local_blocks=int(out[3]),
                    list_bytes=int(out[4]), uncovered=int(out[5]))

    def SpMV(self, d_x: int, d_y: int):
        """y = H x on the device with the assembled values; d_x, d_y are device pointers to 3*n_coef doubles."""
        check(lib.tlfea_newton_spmv(self._h, C.c_void_p(int(d_x)), C.c_void_p(int(d_y))))

    def SetProfile(self, on: bool):
        check(lib.tlfea_newton_set_profile(self._h, int(on)))

    def KernelMs(self):
        a, b, c = C.c_double(0), C.c_double(0), C.c_double(0)
        check(lib.tlfea_newton_kernel_ms(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return {"gp_ms": a.value, "row_ms": b.value, "other_ms": c.value}
