"""Several bunny solvers in one process, each compared with the oracle; classifies wrong entries by plan role."""
import os, sys, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "total-lagrangian-fea_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import test_a1_bunny as tb
from test_patch_plan import build_plan, patch_sections
from oracle import oracle as orc

X, E, fixed, f_ext = tb._load()
o = orc.T10Oracle(E, X)
rng = np.random.default_rng(5)
x = X + rng.normal(0, 0.01, X.shape)
plan = build_plan(E, X, o.off, o.cols, 4, 4)
role = {}
for pi, dsc in enumerate(plan["desc"]):
    for w, (own, dep, ops) in enumerate(patch_sections(plan, dsc)):
        for kind, seg in (("own", own), ("dep", dep)):
            for first, ents in ((1, seg[0]), (0, seg[1])):
                for en in ents:
                    pos = int(dsc["hbase"]) + int(en["hoff"])
                    for dd in range(3):
                        for e in range(3):
                            role[pos + dd * int(en["rowlen"]) + e] = (kind, first, pi, w, int(dsc["n_sub"]))
keep = []
for model, mode in [("svk", 2), ("svk", 2), ("mr", 3), ("svk", 2), ("mr", 3)]:
    d, s = tb._make(model, X, E, fixed, f_ext)
    keep.append((d, s))
    s.SetAssembly(mode)
    d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
    s.Assemble()
    H = s.RetrieveHessianValues()
    H_ref = o.hessian(x, tb._mat(orc, model), tb.DT, tb.RHO, fixed, tb.ALM)
    err = np.abs(H - H_ref)
    bad = np.nonzero(err > 1e-10 * np.abs(H_ref).max())[0]
    c = collections.Counter(role.get(int(b), ("none",))[:2] for b in bad)
    pw = collections.Counter(role.get(int(b), ("none", 0, -1, -1, -1))[2:5] for b in bad)
    ro, ci, _ = s.RetrieveHessianCSR(want_values=False)
    rows = np.searchsorted(ro, bad, side="right") - 1
    mvals = o.mass(tb.RHO)
    for b, r in list(zip(bad, rows))[:9]:
        i, j = r // 3, ci[b] // 3
        mij = mvals[o.off[i] + np.searchsorted(o.cols[o.off[i]:o.off[i + 1]], j)]
        print("   dof row", r, "col", ci[b], "node", i, j, "H", H[b], "ref", H_ref[b], "diff", H[b] - H_ref[b], "M/h", mij / tb.DT,
              "fixed" if i in set(fixed.tolist()) else "", role.get(int(b)))
    print(model, mode, "bad entries", bad.size, dict(c), "patch/warp/nsub", dict(pw), "max abs err", err.max(), flush=True)
