"""Element partition of the structured FEAT10 block across ranks (new design: the reference is
single-GPU).  Slabs of whole cell layers along x; each rank builds ONLY its own slab with a local
node numbering in which the two interface planes are contiguous, so halo packs are plain slices.

Sub-assembled storage: each rank assembles f_int and H over its own elements only; the global
operator is the sum over ranks.  f_int (and every SpMV result) is completed by a halo SUM on
the interface planes; H itself is never exchanged.
"""
from __future__ import annotations

import itertools
from dataclasses import dataclass

import numpy as np

_EDGES = [(0, 1), (1, 2), (0, 2), (0, 3), (1, 3), (2, 3)]


@dataclass
class Slab:
    n: int                 # cells per edge of the global block
    rank: int
    n_parts: int
    layer0: int            # first cell layer (along x) owned
    layer1: int            # one past the last
    X: np.ndarray          # (n_local_nodes, 3) reference coordinates
    E: np.ndarray          # (n_local_elems, 10) local node ids
    node_gid: np.ndarray   # (n_local_nodes,) global node id (i*N+j)*N+k on the (2n+1)^3 lattice
    plane_nodes: int       # nodes per lattice plane = (2n+1)^2
    has_left: bool
    has_right: bool

    @property
    def left_plane(self) -> slice:
        """local node ids shared with rank-1"""
        return slice(0, self.plane_nodes)

    @property
    def right_plane(self) -> slice:
        """local node ids shared with rank+1"""
        return slice(self.X.shape[0] - self.plane_nodes, self.X.shape[0])


def slab_bounds(n: int, n_parts: int):
    return [(n * p) // n_parts for p in range(n_parts + 1)]


def kuhn_slab(n: int, rank: int, n_parts: int) -> Slab:
    """The rank's part of meshgen.kuhn_block(n): same tets, same orientation, same element order
    (restricted to the slab), local node id ((i-2*layer0)*N + j)*N + k."""
    b = slab_bounds(n, n_parts)
    l0, l1 = b[rank], b[rank + 1]
    N = 2 * n + 1
    ni = 2 * (l1 - l0) + 1
    lin = np.arange(N, dtype=np.float64) / (2.0 * n)
    li = (np.arange(ni, dtype=np.float64) + 2 * l0) / (2.0 * n)
    X = np.empty((ni * N * N, 3))
    X[:, 0] = np.repeat(li, N * N)
    X[:, 1] = np.tile(np.repeat(lin, N), ni)
    X[:, 2] = np.tile(lin, ni * N)
    gi = np.repeat(np.arange(ni, dtype=np.int64) + 2 * l0, N * N)
    node_gid = gi * N * N + np.tile(np.arange(N * N, dtype=np.int64), ni)

    ci, cj, ck = np.meshgrid(np.arange(l1 - l0), np.arange(n), np.arange(n), indexing="ij")
    base = np.stack([2 * ci.ravel(), 2 * cj.ravel(), 2 * ck.ravel()], axis=1).astype(np.int64)
    n_cells = base.shape[0]
    E = np.empty((n_cells, 6, 10), dtype=np.int32)
    for t, perm in enumerate(itertools.permutations(range(3))):
        v = np.zeros((4, 3), dtype=np.int64)
        for s in range(3):
            v[s + 1] = v[s]
            v[s + 1, perm[s]] += 2
        if np.linalg.det((v[1:] - v[0]).astype(np.float64)) < 0:
            v[[2, 3]] = v[[3, 2]]
        loc = [v[0], v[1], v[2], v[3]] + [(v[a] + v[c]) // 2 for a, c in _EDGES]
        for a, off in enumerate(loc):
            p = base + off[None, :]
            E[:, t, a] = ((p[:, 0] * N + p[:, 1]) * N + p[:, 2]).astype(np.int32)
    return Slab(n, rank, n_parts, l0, l1, X, E.reshape(n_cells * 6, 10), node_gid, N * N, rank > 0,
                rank < n_parts - 1)


def perturb_global(slab: Slab, amplitude: float, seed: int) -> np.ndarray:
    """Same perturbation every rank would see on the GLOBAL mesh (so interface nodes agree):
    a counter-based hash of (seed, global node id, component) -> U(-1,1)."""
    gid = slab.node_gid.astype(np.uint64)
    out = np.empty_like(slab.X)
    old = np.seterr(over="ignore")  # splitmix64 wraps modulo 2^64 by design
    for c in range(3):
        z = gid * np.uint64(3) + np.uint64(c) + np.uint64(seed) * np.uint64(0x9E3779B97F4A7C15)
        # splitmix64
        z = (z + np.uint64(0x9E3779B97F4A7C15))
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
        u = (z >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
        out[:, c] = slab.X[:, c] + amplitude * (2.0 * u - 1.0)
    np.seterr(**old)
    return out


def halo_sum(f_int, slab: Slab, rank: int, world: int, halo, dist):
    """Sum the interface-plane entries of a nodal vector (3 per node, torch tensor on any device) over the
    two ranks sharing each plane: one all-reduce of the compact interface vector; slot p of `halo`
    (length (world-1) * 3 * plane_nodes) holds the plane between rank p and p+1."""
    if world == 1:
        return
    pn3 = 3 * slab.plane_nodes
    halo.zero_()
    if slab.has_left:
        halo[(rank - 1) * pn3: rank * pn3] = f_int[:pn3]
    if slab.has_right:
        halo[rank * pn3:(rank + 1) * pn3] = f_int[-pn3:]
    dist.all_reduce(halo)
    if slab.has_left:
        f_int[:pn3] = halo[(rank - 1) * pn3: rank * pn3]
    if slab.has_right:
        f_int[-pn3:] = halo[rank * pn3:(rank + 1) * pn3]


class LibHalo:
    """The library's own multi-GPU data plane (csrc/dist.cu: NCCL send/recv with the two x-neighbours, all-reduce of
    device scalars).  torch.distributed is used ONCE, to hand rank 0's ncclUniqueId to the other ranks; a C++ host
    would use a file or MPI for that."""

    def __init__(self, solver, slab: Slab, rank: int, world: int, dist, torch, device):
        import ctypes as C

        from ._lib import check, lib

        self.s, self.lib, self.check, self.C = solver, lib, check, C
        ident = torch.zeros(128, dtype=torch.uint8, device=device)
        if rank == 0:
            raw = (C.c_ubyte * 128)()
            check(lib.tlfea_dist_unique_id(C.cast(raw, C.c_void_p)))
            ident.copy_(torch.tensor(list(raw), dtype=torch.uint8))
        if world > 1:
            dist.broadcast(ident, src=0)
        raw = (C.c_ubyte * 128)(*ident.cpu().tolist())
        check(lib.tlfea_dist_init(solver._h, C.cast(raw, C.c_void_p), rank, world))
        nn, pn = slab.X.shape[0], slab.plane_nodes
        nbr, lo, hi = [], [], []
        if slab.has_left:
            nbr.append(rank - 1), lo.append(0), hi.append(pn)
        if slab.has_right:
            nbr.append(rank + 1), lo.append(nn - pn), hi.append(nn)
        self._arrs = ((C.c_int * max(len(nbr), 1))(*nbr), (C.c_int64 * max(len(nbr), 1))(*lo),
                      (C.c_int64 * max(len(nbr), 1))(*hi))
        check(lib.tlfea_dist_set_halo(solver._h, len(nbr), self._arrs[0], self._arrs[1], self._arrs[2]))
        self.bytes_per_halo = int(lib.tlfea_dist_halo_bytes(solver._h))
        # peer-memory path of the halo sum (one node, NVLink): every rank exports the mailbox each neighbour writes
        # into, the 64-byte IPC handles travel through torch.distributed once, every rank opens its neighbours'.
        # Collective: either all ranks connect or all stay on the NCCL send/recv path.
        import os

        self.p2p = False
        if world > 1 and os.environ.get("TLFEA_HALO_P2P", "1") != "0":
            mine, ok = [], True
            try:
                for k, nb in enumerate(nbr):
                    raw = (C.c_ubyte * 64)()
                    check(lib.tlfea_dist_p2p_export(solver._h, k, C.cast(raw, C.c_void_p)))
                    mine.append((rank, nb, bytes(raw)))
            except Exception:
                ok = False
            gathered = [None] * world
            dist.all_gather_object(gathered, (ok, mine))
            if all(g[0] for g in gathered):
                table = {(owner, frm): h for g in gathered for owner, frm, h in g[1]}
                try:
                    for k, nb in enumerate(nbr):
                        raw = (C.c_ubyte * 64)(*table[(nb, rank)])
                        check(lib.tlfea_dist_p2p_connect(solver._h, k, C.cast(raw, C.c_void_p)))
                except Exception:
                    ok = False
            flags = [None] * world
            dist.all_gather_object(flags, ok)
            self.p2p = all(flags)
            if not self.p2p:  # back to NCCL everywhere: re-declaring the halo drops the mailboxes
                check(lib.tlfea_dist_set_halo(solver._h, len(nbr), self._arrs[0], self._arrs[1], self._arrs[2]))

    def halo_sum(self, ptr: int, width: int = 3):
        self.check(self.lib.tlfea_dist_halo_sum(self.s._h, self.C.c_void_p(int(ptr)), width))

    def allreduce(self, ptr: int, n: int):
        self.check(self.lib.tlfea_dist_allreduce_sum(self.s._h, self.C.c_void_p(int(ptr)), n))

    def finalize(self):
        self.check(self.lib.tlfea_dist_finalize(self.s._h))
