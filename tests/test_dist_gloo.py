"""Host-side multi-GPU logic on CPU: 2 ranks over gloo.  Each rank takes its x-slab of the Kuhn block
(tlfea/partition.py), evaluates its elements' internal force with the oracle (standing in for the device kernels,
which need a GPU), sums the interface plane with partition.halo_sum, and the result must equal the
single-domain force on every node of the slab; residual norms are all-reduced like the Newton loop does."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, out):
    for p in (ROOT, os.path.join(ROOT, "total-lagrangian-fea_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch
    import torch.distributed as dist
    from oracle import oracle as orc
    from tlfea import meshgen, partition

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    slab = partition.kuhn_slab(n, rank, world)
    x = partition.perturb_global(slab, 0.05 / n, 1234)
    mat = orc.Material.svk(7e8, 0.33)
    o = orc.T10Oracle(slab.E, slab.X)
    f_loc, _, _ = o.force(x, mat)
    f = torch.from_numpy(f_loc.copy())
    halo = torch.zeros(max(world - 1, 1) * 3 * slab.plane_nodes, dtype=torch.float64)
    partition.halo_sum(f, slab, rank, world, halo, dist)
    # single-domain reference on the global mesh with the same (global-id keyed) perturbation
    Xg, Eg = meshgen.kuhn_block(n)
    whole = partition.kuhn_slab(n, 0, 1)
    xg = partition.perturb_global(whole, 0.05 / n, 1234)
    fg, _, _ = orc.T10Oracle(Eg, Xg).force(xg, mat)
    fg = fg.reshape(-1, 3)[slab.node_gid].reshape(-1)
    err = np.abs(f.numpy() - fg).max() / np.abs(fg).max()
    # distributed norm: owner = lowest rank touching the node (the right plane belongs to this rank only if last)
    own = np.ones(slab.X.shape[0], dtype=bool)
    if slab.has_left:
        own[:slab.plane_nodes] = False
    sq = torch.tensor([float((f.numpy().reshape(-1, 3)[own] ** 2).sum())], dtype=torch.float64)
    dist.all_reduce(sq)
    nerr = abs(np.sqrt(sq.item()) - np.linalg.norm(fg_full(n, mat))) / np.linalg.norm(fg_full(n, mat))
    out[rank] = (err, nerr, int(slab.E.shape[0]))
    dist.barrier()
    dist.destroy_process_group()


def fg_full(n, mat):
    from oracle import oracle as orc
    from tlfea import meshgen, partition

    Xg, Eg = meshgen.kuhn_block(n)
    xg = partition.perturb_global(partition.kuhn_slab(n, 0, 1), 0.05 / n, 1234)
    f, _, _ = orc.T10Oracle(Eg, Xg).force(xg, mat)
    return f


def test_two_rank_halo_sum_matches_single_domain():
    import torch.multiprocessing as mp

    world, n = 2, 4
    port = _free_port()
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(world, port, n, out), nprocs=world, join=True)
        res = dict(out)
    assert sorted(res) == [0, 1]
    assert sum(r[2] for r in res.values()) == 6 * n ** 3  # every element on exactly one rank
    for err, nerr, _ in res.values():
        assert err < 1e-12 and nerr < 1e-12
