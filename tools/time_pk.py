"""Times the single-pass FEAT10 patch assembly on a Kuhn block: python tools/time_pk.py [n] [warps,warps,...] [modes]
Prints ms per assemble (CUDA events inside the library) per configuration, plus plan statistics and set-up time."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "total-lagrangian-fea_b200"))
import numpy as np, tlfea
from tlfea import meshgen
n = int(sys.argv[1]) if len(sys.argv) > 1 else 48
warps = [int(w) for w in (sys.argv[2] if len(sys.argv) > 2 else "4,8").split(",")]
modes = [int(m) for m in (sys.argv[3] if len(sys.argv) > 3 else "2").split(",")]
X, E = meshgen.kuhn_block(n); x = meshgen.perturb(X, 0.05 / n, 1234)
d = tlfea.GPU_FEAT10_Data(E.shape[0], X.shape[0]); d.Setup(*tlfea.tet5pt_quadrature(), X[:, 0], X[:, 1], X[:, 2], E)
d.SetDensity(2700.); d.SetSVK(7e8, 0.33)
mat, damp = os.environ.get("TLFEA_MAT", "svk"), os.environ.get("TLFEA_DAMP", "0") == "1"
if mat == "mr":
    mu, K = 7e8 / (2 * 1.33), 7e8 / (3 * (1 - 0.66))
    d.SetMooneyRivlin(0.3 * mu, 0.2 * mu, 1.5 * K)
if damp:
    d.SetDamping(1e5, 1e5)
d.CalcDnDuPre(); d.CalcMassMatrix()
s = tlfea.SyncedNewtonSolver(d, 0); s.SetParameters(tlfea.SyncedNewtonParams(1e-4, 0, 1e-6, 1e14, 1, 1, 1e-3)); s.Setup(); s.AnalyzeHessianSparsity()
d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
if damp:
    s.SetVelocityGuess(0.1 * np.random.Generator(np.random.PCG64(1235)).uniform(-1, 1, size=3 * X.shape[0]))
for mode in modes:
    s.SetAssembly(mode)
    for w in (warps if mode == 2 else [0]):
        if mode == 2:
            s.SetPatchWarps(w)
        t0 = time.time(); s.Assemble(); d.Sync(); t_first = time.time() - t0
        for i in range(2): s.Assemble()
        s.Timing(reset=True); s.SetProfile(True)
        reps = 5
        for i in range(reps): s.Assemble()
        ms, calls = s.Timing()
        st = s.PatchStats() if mode == 2 else {}
        print(mat, "damp" if damp else "", "n", n, "elems", E.shape[0], "mode", mode, "warps", w, "ms/assemble", round(ms / calls, 3),
              {k: round(v / reps, 3) for k, v in s.KernelMs().items()}, "first call s", round(t_first, 2), st, flush=True)
