"""Distributed Newton iteration (tlfea/dist_newton.py + tlfea_newton_krylov_op) on 2 GPUs against the single-GPU
library step on the same global mesh: element-partitioned slabs, sub-assembled H, halo sums over NCCL, all-reduced
dot products.  Needs two GPUs (skipped on a one-GPU box; run with `gpurun --gpus 2`).  Tolerance 1e-9 of the largest
displacement: the two runs sum interface contributions in different orders and PCG amplifies rounding."""
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

N_CELLS, NEWTON_ITERS, PCG_ITERS = 6, 2, 40
E_MOD, NU, RHO, H_STEP, ALM = 7e8, 0.33, 2700.0, 1e-3, 1e14


def _build(slab, device, torch):
    import tlfea
    from tlfea import partition

    x = partition.perturb_global(slab, 0.05 / slab.n, 1234)
    d = tlfea.GPU_FEAT10_Data(slab.E.shape[0], slab.X.shape[0], device)
    d.Initialize()
    d.SetStream(torch.cuda.current_stream().cuda_stream)
    d.Setup(*tlfea.tet5pt_quadrature(), slab.X[:, 0], slab.X[:, 1], slab.X[:, 2], slab.E)
    d.SetDensity(RHO)
    d.SetSVK(E_MOD, NU)
    # pinned: the global face x = 0 AND a strip of every interface plane (gid-based, so both holders pin the same
    # nodes); loaded: every node, by global id -- f_ext and the constraint terms of shared nodes may count once only
    gid = slab.node_gid
    plane = gid // slab.plane_nodes
    fixed = np.nonzero((plane == 0) | ((plane == N_CELLS) & (gid % slab.plane_nodes < 9)))[0].astype(np.int32)
    if len(fixed):
        d.SetNodalFixed(fixed)
    fext = np.zeros(3 * gid.shape[0])
    fext[2::3] = -50.0 * (1.0 + (gid % 7))
    d.SetExternalForce(fext)
    d.CalcDnDuPre()
    d.CalcMassMatrix()
    s = tlfea.SyncedNewtonSolver(d, 3 * len(fixed))
    s.SetParameters(tlfea.SyncedNewtonParams(1e-30, 0.0, 1e-6, ALM, 1, NEWTON_ITERS, H_STEP))
    s.Setup()
    s.AnalyzeHessianSparsity()
    s.SetFixedSparsityPattern(True)
    d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
    return d, s, x


ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out_dir):
    for p in (ROOT, os.path.join(ROOT, "total-lagrangian-fea_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch
    import torch.distributed as dist

    from tlfea import partition
    from tlfea.dist_newton import DistNewton

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    slab = partition.kuhn_slab(N_CELLS, rank, world)
    d, s, _ = _build(slab, rank, torch)
    halo = partition.LibHalo(s, slab, rank, world, dist, torch, dev)  # the library's NCCL data plane
    dn = DistNewton(d, s, slab, rank, world, dist, torch, dev, halo=halo)
    dn.begin_step()
    norms = []
    for _ in range(NEWTON_ITERS):
        ng, it = dn.iteration(PCG_ITERS)
        norms.append(ng)
        assert it == PCG_ITERS
    torch.cuda.synchronize()
    gx, gy, gz = d.RetrievePositionToCPU()
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), pos=np.stack([gx, gy, gz], axis=1), gid=slab.node_gid,
             norms=np.array(norms))
    dist.barrier()
    halo.finalize()
    dist.destroy_process_group()


def test_two_gpu_newton_matches_single_gpu(tmp_path):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp

    from tlfea import partition
    from tlfea.dist_newton import DistNewton

    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)

    # single GPU, same global mesh: (a) the library's own step, (b) the Krylov ops driven with world = 1
    torch.cuda.set_device(0)
    whole = partition.kuhn_slab(N_CELLS, 0, 1)
    d, s, x0 = _build(whole, 0, torch)
    s.SetPCG(PCG_ITERS, 0.0)
    s.Solve()
    assert s.LastStats()["newton_iters"] == NEWTON_ITERS
    ref = np.stack(d.RetrievePositionToCPU(), axis=1)
    d2, s2, _ = _build(whole, 0, torch)
    dn = DistNewton(d2, s2, whole, 0, 1, None, torch, torch.device("cuda", 0))
    dn.begin_step()
    norms1 = [dn.iteration(PCG_ITERS)[0] for _ in range(NEWTON_ITERS)]
    torch.cuda.synchronize()
    one = np.stack(d2.RetrievePositionToCPU(), axis=1)
    scale = np.abs(ref - x0).max()
    assert scale > 0
    assert np.abs(one - ref).max() < 1e-12 * scale + 1e-15, np.abs(one - ref).max() / scale

    gid_pos = {int(g): i for i, g in enumerate(whole.node_gid)}
    for r in range(2):
        z = np.load(os.path.join(str(tmp_path), f"rank{r}.npz"))
        idx = np.array([gid_pos[int(g)] for g in z["gid"]])
        err = np.abs(z["pos"] - ref[idx]).max()
        assert err < 1e-9 * scale, (r, err / scale)
        assert np.allclose(z["norms"], norms1, rtol=1e-9)


def test_cpp_two_rank_assemble_matches_one_rank(tmp_path):
    """examples/two_rank_assemble.cc: the partitioned assembly from host C++ only (shim classes + tlfea_dist_*, ids and
    IPC handles exchanged through files).  Two processes on two GPUs must give the same owned sums of f_int as one
    process on the whole block (1e-12: the interface entries are summed in a different order)."""
    import shutil
    import subprocess

    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    gxx = shutil.which("g++")
    if gxx is None:
        pytest.skip("no g++")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib_dir = os.path.join(root, "total-lagrangian-fea_b200", "lib")
    exe = tmp_path / "two_rank"
    subprocess.check_call([gxx, "-std=c++17", "-O1", "-I", os.path.join(root, "include"),
                           os.path.join(root, "examples", "two_rank_assemble.cc"), "-L", lib_dir, "-ltlfea_b200",
                           "-o", str(exe)])
    env = dict(os.environ, LD_LIBRARY_PATH=lib_dir + os.pathsep + os.environ.get("LD_LIBRARY_PATH", ""))

    def sums(world):
        xdir = tmp_path / f"x{world}"
        xdir.mkdir()
        procs = [subprocess.Popen([str(exe), str(r), str(world), str(xdir), "8"], env=env, text=True,
                                  stdout=subprocess.PIPE, stderr=subprocess.PIPE) for r in range(world)]
        f2 = fs = 0.0
        halo = set()
        for p in procs:
            out, err = p.communicate(timeout=300)
            assert p.returncode == 0, err
            tok = out.split()
            f2 += float(tok[tok.index("owned_f2") + 1])
            fs += float(tok[tok.index("owned_fsum") + 1])
            halo.add(tok[tok.index("halo") + 1])
        return f2, fs, halo

    f2_1, fs_1, _ = sums(1)
    f2_2, fs_2, halo = sums(2)
    assert f2_1 > 0.0 and halo <= {"peer-memory", "nccl"} and len(halo) == 1
    assert abs(f2_2 - f2_1) <= 1e-12 * f2_1
    assert abs(fs_2 - fs_1) <= 1e-9 * np.sqrt(f2_1)  # the components cancel: absolute scale of the entries
