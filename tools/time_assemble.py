"""Quick kernel timing of one FEAT10 assemble on a Kuhn block: python tools/time_assemble.py [n] [mode] [morton]
TLFEA_MAT=mr selects Mooney-Rivlin (0.3 mu, 0.2 mu, 1.5 K), TLFEA_DAMP=1 adds Kelvin-Voigt damping (1e5, 1e5) with v != 0.
morton=1 renumbers the elements along a Morton curve of their first vertex before Setup."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "total-lagrangian-fea_b200"))
import numpy as np, tlfea
from tlfea import meshgen
n = int(sys.argv[1]) if len(sys.argv) > 1 else 48
mode = int(sys.argv[2]) if len(sys.argv) > 2 else 1
morton = int(sys.argv[3]) if len(sys.argv) > 3 else 0
X, E = meshgen.kuhn_block(n); x = meshgen.perturb(X, 0.05 / n, 1234)
if morton:
    def spread(v):
        v = v.astype(np.uint64) & np.uint64(0x1fffff)
        v = (v | (v << np.uint64(32))) & np.uint64(0x1f00000000ffff)
        v = (v | (v << np.uint64(16))) & np.uint64(0x1f0000ff0000ff)
        v = (v | (v << np.uint64(8))) & np.uint64(0x100f00f00f00f00f)
        v = (v | (v << np.uint64(4))) & np.uint64(0x10c30c30c30c30c3)
        v = (v | (v << np.uint64(2))) & np.uint64(0x1249249249249249)
        return v
    c = X[E[:, :4]].mean(axis=1)
    q = np.floor(c * n).astype(np.int64)
    key = (spread(q[:, 0]) << np.uint64(2)) | (spread(q[:, 1]) << np.uint64(1)) | spread(q[:, 2])
    E = np.ascontiguousarray(E[np.argsort(key, kind="stable")])
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
s.SetAssembly(mode)
d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
if damp:
    s.SetVelocityGuess(0.1 * np.random.Generator(np.random.PCG64(1235)).uniform(-1, 1, size=3 * X.shape[0]))
for i in range(3): s.Assemble()
s.Timing(reset=True); s.SetProfile(True)
for i in range(5): s.Assemble()
print(mat, "damp" if damp else "", "n", n, "mode", mode, "morton", morton, "elems", E.shape[0], {k: round(v / 5, 3) for k, v in s.KernelMs().items()}, flush=True)
