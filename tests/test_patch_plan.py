"""CPU checks of the single-pass FEAT10 assembly plan (csrc/patch_plan.cu): the plan is pure host data, so the
kernel's summation order (warp accumulators -> combine in warp order -> in-place update of H colour by colour) is
replayed here in numpy with the oracle's element tangents and compared with the oracle's assembled H."""
import ctypes as C

import numpy as np
import pytest

from conftest import rel_err

PELEM = np.dtype([("elem", "<i4"), ("pad0", "<i4"), ("node", "<i4", (10,)), ("blk4", "<u4", (15,)),
                  ("tr", "<u4", (2,)), ("pad1", "<i4", (3,))])
PROW = np.dtype([("hbase", "<i8"), ("rowlen", "<i4"), ("mass_base", "<i4"), ("ent_begin", "<i4"),
                 ("n_ent", "<i4"), ("node", "<i4"), ("pad", "<i4")])
PDESC = np.dtype([("sub_ptr", "<i4", (9,)), ("row_begin", "<i4"), ("n_rows", "<i4"), ("src_base", "<i4"),
                  ("n_sub", "<i4"), ("pad", "<i4", (3,))])
TILES = [(0, j) for j in range(5)] + [(1, j) for j in range(1, 5)] + [(2, j) for j in range(2, 5)] + \
        [(3, 3), (3, 4), (4, 4)]
CAPU = 216


def build_plan(E, X, off, cols, W):
    from tlfea import _lib

    assert PELEM.itemsize == 128 and PROW.itemsize == 32 and PDESC.itemsize == 64
    E = np.ascontiguousarray(E, dtype=np.int32)
    x, y, z = (np.ascontiguousarray(X[:, k], dtype=np.float64) for k in range(3))
    off = np.ascontiguousarray(off, dtype=np.int32)
    cols = np.ascontiguousarray(cols, dtype=np.int32)
    h = C.c_void_p()
    ip = lambda a: a.ctypes.data_as(_lib.c_ip)  # noqa: E731
    dp = lambda a: a.ctypes.data_as(_lib.c_dp)  # noqa: E731
    _lib.check(_lib.lib.tlfea_debug_patch_plan_build(E.shape[0], X.shape[0], ip(E), dp(x), dp(y), dp(z), ip(off),
                                                      ip(cols), W, C.byref(h)))
    sizes = np.zeros(8, dtype=np.int64)
    _lib.check(_lib.lib.tlfea_debug_patch_plan_sizes(h, sizes.ctypes.data_as(_lib.c_i64p)))
    npatch, ncol, nel, nrow, nent, nsrc, nunc, suml = (int(v) for v in sizes)
    plan = dict(color_ptr=np.zeros(ncol + 1, np.int32), desc=np.zeros(npatch, PDESC), elems=np.zeros(nel, PELEM),
                rows=np.zeros(nrow, PROW), ent=np.zeros(nent, np.uint32), src=np.zeros(nsrc, np.uint16),
                uncovered=np.zeros(nunc, np.int64), sum_local=suml)
    vp = lambda a: C.c_void_p(a.ctypes.data)  # noqa: E731
    _lib.check(_lib.lib.tlfea_debug_patch_plan_copy(h, vp(plan["color_ptr"]), vp(plan["desc"]), vp(plan["elems"]),
                                                     vp(plan["rows"]), vp(plan["ent"]), vp(plan["src"]),
                                                     vp(plan["uncovered"])))
    _lib.check(_lib.lib.tlfea_debug_patch_plan_free(h))
    return plan


def replay(plan, Ke, mass, h, nnz):
    """Ke[e] = h*K_e (30x30, dof = 3*a + i).  Returns H values in the scalar-CSR layout of the library."""
    H = np.full(9 * nnz, np.nan)
    cp, desc, el, rows, ent, src = (plan[k] for k in ("color_ptr", "desc", "elems", "rows", "ent", "src"))
    for c in range(len(cp) - 1):
        for p in range(cp[c], cp[c + 1]):
            d = desc[p]
            acc = np.zeros((8, CAPU, 3, 3))
            for w in range(d["n_sub"]):
                for k in range(d["sub_ptr"][w], d["sub_ptr"][w + 1]):
                    pe = el[k]
                    K = Ke[pe["elem"]]
                    for t, (I, J) in enumerate(TILES):
                        for ta in range(2):
                            for tb in range(2):
                                u = (int(pe["blk4"][t]) >> (8 * (ta * 2 + tb))) & 255
                                if I == J and ta == 1 and tb == 0:
                                    assert u == 255
                                    continue
                                a, b = 2 * I + ta, 2 * J + tb
                                bit = t * 4 + ta * 2 + tb
                                tr = (int(pe["tr"][bit >> 5]) >> (bit & 31)) & 1
                                blk = K[3 * a:3 * a + 3, 3 * b:3 * b + 3]
                                assert u < CAPU
                                acc[w, u] += blk.T if tr else blk
            for r in rows[d["row_begin"]:d["row_begin"] + d["n_rows"]]:
                for k in range(r["ent_begin"], r["ent_begin"] + r["n_ent"]):
                    w0 = int(ent[k])
                    slot, first, ns, so = w0 & 8191, (w0 >> 13) & 1, ((w0 >> 14) & 7) + 1, w0 >> 17
                    val = np.zeros((3, 3))
                    for q in range(ns):
                        s = int(src[d["src_base"] + so + q])
                        b = acc[(s >> 8) & 7, s & 255]
                        val = val + (b.T if (s >> 11) & 1 else b)
                    for dd in range(3):
                        base = r["hbase"] + dd * r["rowlen"] + 3 * slot
                        old = np.where(np.arange(3) == dd, mass[r["mass_base"] + slot] / h, 0.0) if first \
                            else H[base:base + 3]
                        assert first or not np.isnan(old).any()
                        H[base:base + 3] = old + val[dd]
    return H


def check_plan_invariants(plan, E, W):
    el, desc, cp, rows = plan["elems"], plan["desc"], plan["color_ptr"], plan["rows"]
    assert sorted(el["elem"].tolist()) == list(range(E.shape[0]))
    assert (el["node"] == E[el["elem"]]).all()
    assert cp[0] == 0 and cp[-1] == len(desc)
    for c in range(len(cp) - 1):  # patches of one colour share no coefficient
        seen = set()
        for p in range(cp[c], cp[c + 1]):
            nodes = set(rows["node"][desc[p]["row_begin"]:desc[p]["row_begin"] + desc[p]["n_rows"]].tolist())
            assert not (nodes & seen)
            seen |= nodes
    assert (desc["n_sub"] <= W).all() and (desc["n_sub"] >= 1).all()


@pytest.mark.parametrize("n,W", [(3, 8), (4, 4), (2, 8)])
def test_plan_replay_matches_oracle_kuhn(n, W):
    from tlfea import meshgen
    from oracle import oracle as orc

    X, E = meshgen.kuhn_block(n)
    x = meshgen.perturb(X, 0.05 / n, 1234)
    o = orc.T10Oracle(E, X)
    mat = orc.Material.svk(7e8, 0.33)
    h, rho = 1e-3, 2700.0
    plan = build_plan(E, X, o.off, o.cols, W)
    check_plan_invariants(plan, E, W)
    assert plan["uncovered"].size == 0
    Ke = [h * orc.element_tangent(E[e], x, o.grad_N[e], o.detJ[e], o.qw, mat) for e in range(E.shape[0])]
    H = replay(plan, Ke, o.mass(rho), h, o.cols.shape[0])
    assert not np.isnan(H).any()
    assert rel_err(H, o.hessian(x, mat, h, rho)) < 1e-13
    if n >= 4 and W == 8:  # even block: every patch is a 2x2x2 brick of cells, 8 colours
        assert len(plan["color_ptr"]) - 1 == 8
        assert (plan["desc"]["n_sub"] == 8).all()


def test_plan_unstructured_permuted():
    """Scrambled element order and node numbering: the plan may not rely on the generator's layout."""
    from tlfea import meshgen
    from oracle import oracle as orc

    rng = np.random.default_rng(7)
    X, E = meshgen.kuhn_block(3)
    perm = rng.permutation(X.shape[0])
    inv = np.empty_like(perm)
    inv[perm] = np.arange(perm.size)
    X2 = X[perm] * np.array([1.0, 0.6, 1.7]) + rng.normal(0, 0.01, X.shape)
    E2 = inv[E][rng.permutation(E.shape[0])].astype(np.int32)
    o = orc.T10Oracle(E2, X2)
    mat = orc.Material.svk(3e8, 0.4)
    h, rho = 1e-3, 920.0
    x = X2 + rng.normal(0, 0.004, X2.shape)
    for W in (4, 8):
        plan = build_plan(E2, X2, o.off, o.cols, W)
        check_plan_invariants(plan, E2, W)
        Ke = [h * orc.element_tangent(E2[e], x, o.grad_N[e], o.detJ[e], o.qw, mat) for e in range(E2.shape[0])]
        H = replay(plan, Ke, o.mass(rho), h, o.cols.shape[0])
        assert rel_err(H, o.hessian(x, mat, h, rho)) < 1e-13


def test_plan_rejects_repeated_node():
    from tlfea import meshgen, _lib
    from oracle import oracle as orc

    X, E = meshgen.kuhn_block(2)
    o = orc.T10Oracle(E, X)
    E = E.copy()
    E[5, 7] = E[5, 2]
    with pytest.raises(_lib.TlfeaError):
        build_plan(E, X, o.off, o.cols, 8)
