"""Repeats the bunny assembly and classifies the H entries that differ between runs by their write-out role in the
patch plan (own / merged, first / later touch): python tools/race_hunt.py [reps] [mode]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "total-lagrangian-fea_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import test_a1_bunny as tb
from test_patch_plan import build_plan, patch_sections
from oracle import oracle as orc

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
mode = int(sys.argv[2]) if len(sys.argv) > 2 else 2
X, E, fixed, f_ext = tb._load()
d, s = tb._make("svk", X, E, fixed, f_ext)
s.SetAssembly(mode)
rng = np.random.default_rng(5)
x = X + rng.normal(0, 0.01, X.shape)
d.UpdatePositions(x[:, 0], x[:, 1], x[:, 2])
o = orc.T10Oracle(E, X)
off, cols, _ = d.RetrieveMassCSRToCPU()
plan = build_plan(E, X, off, cols, 4, 4)
role = {}
for pi, dsc in enumerate(plan["desc"]):
    for w, (own, dep, ops) in enumerate(patch_sections(plan, dsc)):
        for kind, seg in (("own", own), ("dep", dep)):
            for first, ents in ((1, seg[0]), (0, seg[1])):
                for en in ents:
                    pos = int(dsc["hbase"]) + int(en["hoff"])
                    for dd in range(3):
                        for e in range(3):
                            role[pos + dd * int(en["rowlen"]) + e] = (kind, first, pi, w, int(en["src"]) >> 15)
Hs = []
for r in range(reps):
    s.Assemble()
    Hs.append(s.RetrieveHessianValues().copy())
ref = Hs[0]
import collections
for r in range(1, reps):
    bad = np.nonzero(Hs[r] != ref)[0]
    if bad.size:
        c = collections.Counter(role.get(int(b), ("none",))[:2] for b in bad)
        pw = collections.Counter(role.get(int(b), ("none", 0, -1, -1))[2:4] for b in bad)
        print("run", r, "differs in", bad.size, "entries", dict(c), "patch/warp", dict(pw), "max abs", np.abs(Hs[r][bad] - ref[bad]).max())
print("done", reps, "runs")
