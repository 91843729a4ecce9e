"""Timing of one force+tangent assembly on ANCF meshes: python tools/time_ancf.py 3443 NX [NY] | 3243 NELEM"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "total-lagrangian-fea_b200"))
import numpy as np, tlfea
kind = int(sys.argv[1]); nx = int(sys.argv[2]); ny = int(sys.argv[3]) if len(sys.argv) > 3 else nx
t0 = time.time()
if kind == 3443:
    L, W, H = 1e-2, 1e-2, 1e-3
    ii, jj = np.meshgrid(np.arange(nx + 1), np.arange(ny + 1), indexing="ij")
    nn = (nx + 1) * (ny + 1)
    x = np.zeros((nn, 4)); y = np.zeros((nn, 4)); z = np.zeros((nn, 4))
    x[:, 0] = ii.ravel() * L; x[:, 1] = 1.0; y[:, 0] = jj.ravel() * W; y[:, 2] = 1.0; z[:, 3] = 1.0
    nid = lambda i, j: i * (ny + 1) + j
    ci, cj = np.meshgrid(np.arange(nx), np.arange(ny), indexing="ij"); ci = ci.ravel(); cj = cj.ravel()
    conn = np.stack([nid(ci, cj), nid(ci + 1, cj), nid(ci + 1, cj + 1), nid(ci, cj + 1)], axis=1).astype(np.int32)
    d = tlfea.GPU_ANCF3443_Data(nn, conn.shape[0]); Q = 48
else:
    L, W, H = 0.5, 0.1, 0.1
    nn = nx + 1
    x = np.zeros((nn, 4)); y = np.zeros((nn, 4)); z = np.zeros((nn, 4))
    x[:, 0] = np.arange(nn) * L; x[:, 1] = 1.0; y[:, 2] = 1.0; z[:, 3] = 1.0
    conn = np.stack([np.arange(nx), np.arange(nx) + 1], axis=1).astype(np.int32)
    d = tlfea.GPU_ANCF3243_Data(nn, conn.shape[0]); Q = 12
x, y, z = x.ravel(), y.ravel(), z.ravel()
rng = np.random.Generator(np.random.PCG64(2234))
amp = np.tile([0.05 * L, 0.02, 0.02, 0.02], nn)
d.Setup(L, W, H, x, y, z, conn); d.SetDensity(2700.); d.SetSVK(7e8, 0.33)
if os.environ.get("TLFEA_MAT", "svk") == "mr":  # 0.3 mu, 0.2 mu, 1.5 K
    mu, K = 7e8 / (2 * 1.33), 7e8 / (3 * (1 - 0.66)); d.SetMooneyRivlin(0.3 * mu, 0.2 * mu, 1.5 * K)
d.CalcDsDuPre(); d.CalcMassMatrix()
s = tlfea.SyncedNewtonSolver(d, 0); s.SetParameters(tlfea.SyncedNewtonParams(1e-4, 0, 1e-6, 1e14, 1, 1, 1e-3)); s.Setup(); s.AnalyzeHessianSparsity()
s.SetAssembly(int(os.environ.get("TLFEA_MODE", "1")))
d.UpdatePositions(x + amp * rng.uniform(-1, 1, 4 * nn), y + amp * rng.uniform(-1, 1, 4 * nn), z + amp * rng.uniform(-1, 1, 4 * nn))
setup_s = time.time() - t0
for i in range(2): s.Assemble()
s.Timing(reset=True); s.SetProfile(True)
for i in range(3): s.Assemble()
k = {a: b / 3 for a, b in s.KernelMs().items()}
tot = k["gp_ms"] + k["row_ms"] + k["other_ms"]
print("kind", kind, "elems", conn.shape[0], "GPs", conn.shape[0] * Q, {a: round(b, 3) for a, b in k.items()},
      "GP/s %.3e" % (conn.shape[0] * Q / (tot * 1e-3)), "nnz_H", s.nnz(), "setup_s %.1f" % setup_s, flush=True)
