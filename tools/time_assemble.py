"""Quick kernel timing of one FEAT10 assemble on a Kuhn block: python tools/time_assemble.py [n] [mode]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "total-lagrangian-fea_b200"))
import numpy as np, tlfea
from tlfea import meshgen
n = int(sys.argv[1]) if len(sys.argv) > 1 else 48
mode = int(sys.argv[2]) if len(sys.argv) > 2 else 1
X, E = meshgen.kuhn_block(n); x = meshgen.perturb(X, 0.05 / n, 1234)
d = tlfea.GPU_FEAT10_Data(E.shape[0], X.shape[0]); d.Setup(*tlfea.tet5pt_quadrature(), X[:, 0], X[:, 1], X[:, 2], E)
d.SetDensity(2700.); d.SetSVK(7e8, 0.33); d.CalcDnDuPre(); d.CalcMassMatrix()
s = tlfea.SyncedNewtonSolver(d, 0); s.SetParameters(tlfea.SyncedNewtonParams(1e-4, 0, 1e-6, 1e14, 1, 1, 1e-3)); s.Setup(); s.AnalyzeHessianSparsity()
s.SetAssembly(mode)
d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
for i in range(3): s.Assemble()
s.Timing(reset=True); s.SetProfile(True)
for i in range(5): s.Assemble()
print("n", n, "mode", mode, "elems", E.shape[0], {k: round(v / 5, 3) for k, v in s.KernelMs().items()}, s.TileStats(), flush=True)
import ctypes as C
buf = (C.c_uint64 * 8)()
if mode == 1 and tlfea.lib.tlfea_newton_tile_timing(s._h, buf, 0) == 0 and sum(buf):
    tot = sum(buf); names = ["prologue", "visits(prev)+loop", "wait+barrier", "stage", "barrier2", "prefetch", "final wait", "writeout"]
    print({nm: round(100 * v / tot, 1) for nm, v in zip(names, buf)}, "cycles/tile/warp", tot / s.TileStats()["n_tiles"] / 8 / 8)
