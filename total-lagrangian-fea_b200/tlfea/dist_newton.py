"""One implicit Newton iteration on an element-partitioned mesh (new design: the reference's OneStepNewtonCuDSS,
SyncedNewton.cu:909-1394, is single-GPU with a cuDSS solve).

Every rank assembles its own sub-assembled H and local g; the global operator is the sum over ranks.  The library
does all arithmetic (tlfea_newton_krylov_op); this module only moves data between ranks: a halo sum of interface
entries after each op that produces a vector, and an all-reduce of the 4 device-resident partial scalars.  alpha and
beta are formed on the device, so a PCG iteration has no host synchronisation.
"""
from __future__ import annotations

from . import _lib
from ._lib import check, lib
from .partition import Slab

KOP = dict(SAVE=0, RESIDUAL=1, NORM_G=2, BEGIN=3, PRECOND=4, SPMV=5, PAP=6, UPDATE_XR=7, UPDATE_P=8, APPLY=9)
KVEC = dict(Z=0, P=1, AP=2, DIAG=3, R=4, DV=5, SCAL=6)


class _Dev:
    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (int(ptr), False), "version": 2}


class DistNewton:
    """d: GPU_FEAT10_Data of the rank's slab, s: its SyncedNewtonSolver (analysed), dist: torch.distributed or None."""

    def __init__(self, d, s, slab: Slab, rank: int, world: int, dist, torch, device, halo=None):
        """halo: a partition.LibHalo -- the data plane then runs inside the library (NCCL send/recv with the neighbours,
        all-reduce of the device scalars); None keeps the torch.distributed all-reduce of the compact interface vector."""
        self.d, self.s, self.slab, self.rank, self.world, self.dist, self.torch = d, s, slab, rank, world, dist, torch
        self.lib_halo = halo
        self.h = s._h
        self.n = s.n_dofs
        check(lib.tlfea_newton_krylov_op(self.h, KOP["SAVE"]))  # also allocates the Krylov vectors
        t = lambda ptr, n: torch.as_tensor(_Dev(ptr, n), device=device)
        kp = lambda w: lib.tlfea_newton_devptr_krylov(self.h, KVEC[w])
        self.g = t(lib.tlfea_newton_devptr_g(self.h), self.n)
        self.Ap = t(kp("AP"), self.n)
        self.diag = t(kp("DIAG"), 3 * self.n)
        self.scal = t(kp("SCAL"), 4)
        pn = slab.plane_nodes
        self.halo3 = torch.zeros(max(world - 1, 1) * 3 * pn, dtype=torch.float64, device=device)
        self.halo9 = torch.zeros(max(world - 1, 1) * 9 * pn, dtype=torch.float64, device=device)
        # the left plane belongs to rank-1: dot products skip it here
        check(lib.tlfea_newton_set_owned_range(self.h, 3 * pn if slab.has_left else 0, self.n))
        self.halo_bytes = 0

    def _op(self, name):
        check(lib.tlfea_newton_krylov_op(self.h, KOP[name]))

    def _halo(self, vec, width, buf):
        if self.world == 1:
            return
        if self.lib_halo is not None:
            self.lib_halo.halo_sum(vec.data_ptr(), width)
            self.halo_bytes += self.lib_halo.bytes_per_halo * width // 3
            return
        w = width * self.slab.plane_nodes
        r = self.rank
        buf.zero_()
        if self.slab.has_left:
            buf[(r - 1) * w: r * w] = vec[:w]
        if self.slab.has_right:
            buf[r * w:(r + 1) * w] = vec[-w:]
        self.dist.all_reduce(buf)
        if self.slab.has_left:
            vec[:w] = buf[(r - 1) * w: r * w]
        if self.slab.has_right:
            vec[-w:] = buf[r * w:(r + 1) * w]
        self.halo_bytes += buf.numel() * 8

    def _reduce_scalars(self, lo, hi):
        """sum the partial scalars scal[lo:hi] an op has just written (already-reduced entries must not be re-summed)"""
        if self.world > 1:
            if self.lib_halo is not None:
                self.lib_halo.allreduce(self.scal[lo:hi].data_ptr(), hi - lo)
            else:
                self.dist.all_reduce(self.scal[lo:hi])

    def begin_step(self):
        self._op("SAVE")

    def iteration(self, pcg_iters: int, rel_tol: float = 0.0):
        """assemble, residual, PCG (block-Jacobi, at most pcg_iters iterations), update.  Returns (||g||, iterations)."""
        self.s.Assemble()
        self._op("RESIDUAL")
        self._halo(self.g, 3, self.halo3)
        self._op("NORM_G")
        self._reduce_scalars(3, 4)
        self._op("BEGIN")
        self._halo(self.diag, 9, self.halo9)
        norm_g = float(self.scal[3].item()) ** 0.5  # the one host read of the iteration (convergence test)
        self._op("PRECOND")
        self._reduce_scalars(0, 1)
        self._reduce_scalars(3, 4)
        r0 = None
        it = 0
        for it in range(1, pcg_iters + 1):
            self._op("SPMV")
            self._halo(self.Ap, 3, self.halo3)
            self._op("PAP")
            self._reduce_scalars(1, 2)
            self._op("UPDATE_XR")
            self._reduce_scalars(2, 4)
            if rel_tol > 0.0:
                rr = float(self.scal[3].item())
                r0 = rr if r0 is None else r0
                if rr <= rel_tol * rel_tol * max(norm_g * norm_g, 1e-300):
                    break
            self._op("UPDATE_P")
        self._op("APPLY")
        return norm_g, it
