"""CPU checks of the single-pass FEAT10 assembly plan (csrc/patch_plan.cu): the plan is pure host data, so the
kernel's summation order (warp accumulators -> combine in warp order -> in-place update of H colour by colour) is
replayed here in numpy with the oracle's element tangents and compared with the oracle's assembled H."""
import ctypes as C

import numpy as np
import pytest

from conftest import rel_err

PELEM = np.dtype([("elem", "<i4"), ("pad0", "<i4"), ("node", "<i4", (10,)), ("blk4", "<u4", (15,)),
                  ("tr", "<u4", (2,)), ("first", "<u4", (2,)), ("pad1", "<i4")])
PENT = np.dtype([("hoff", "<u4"), ("rowlen", "<u2"), ("src", "<u2")])
PENTF = np.dtype([("hoff", "<u4"), ("rowlen", "<u2"), ("src", "<u2"), ("mass", "<f8")])
HDR_INTS, MAXW = 64, 8  # csrc/patch_plan.cuh: PK_HDR_INTS, PK_MAXW
PDESC = np.dtype([("sub_ptr", "<i4", (9,)), ("n_sub", "<i4"), ("blob_bytes", "<i4"), ("n_first", "<i4"), ("n_merge", "<i4"),
                  ("blob_off", "<u4"), ("hbase", "<i8")])
TILES = [(0, j) for j in range(5)] + [(1, j) for j in range(1, 5)] + [(2, j) for j in range(2, 5)] + \
        [(3, 3), (3, 4), (4, 4)]
CAPU = 210


def build_plan(E, X, off, cols, W, n_chunks=1):
    from tlfea import _lib

    assert PELEM.itemsize == 128 and PENT.itemsize == 8 and PENTF.itemsize == 16 and PDESC.itemsize == 64
    E = np.ascontiguousarray(E, dtype=np.int32)
    x, y, z = (np.ascontiguousarray(X[:, k], dtype=np.float64) for k in range(3))
    off = np.ascontiguousarray(off, dtype=np.int32)
    cols = np.ascontiguousarray(cols, dtype=np.int32)
    h = C.c_void_p()
    ip = lambda a: a.ctypes.data_as(_lib.c_ip)  # noqa: E731
    dp = lambda a: a.ctypes.data_as(_lib.c_dp)  # noqa: E731
    _lib.check(_lib.lib.tlfea_debug_patch_plan_build(E.shape[0], X.shape[0], ip(E), dp(x), dp(y), dp(z), ip(off),
                                                      ip(cols), W, n_chunks, C.byref(h)))
    sizes = np.zeros(8, dtype=np.int64)
    _lib.check(_lib.lib.tlfea_debug_patch_plan_sizes(h, sizes.ctypes.data_as(_lib.c_i64p)))
    npatch, ncol, nel, nrow, nblob, maxblob, nunc, suml = (int(v) for v in sizes)
    plan = dict(color_ptr=np.zeros(ncol + 1, np.int32), desc=np.zeros(npatch, PDESC), elems=np.zeros(nel, PELEM),
                row_node=np.zeros(nrow, np.int32), row_ptr=np.zeros(npatch + 1, np.int64),
                blob=np.zeros(nblob, np.uint8),
                uncovered=np.zeros(nunc, np.int64), sum_local=suml, max_blob=maxblob)
    vp = lambda a: C.c_void_p(a.ctypes.data)  # noqa: E731
    _lib.check(_lib.lib.tlfea_debug_patch_plan_copy(h, vp(plan["color_ptr"]), vp(plan["desc"]), vp(plan["elems"]),
                                                     vp(plan["row_node"]), vp(plan["row_ptr"]), vp(plan["blob"]),
                                                     vp(plan["uncovered"])))
    _lib.check(_lib.lib.tlfea_debug_patch_plan_free(h))
    return plan


def al16(b):
    return (b + 15) & ~15


def patch_sections(plan, d):
    """Chunks of one patch's blob (csrc/patch_plan.cuh): per warp w ((own first16, own later8), (merged first16,
    merged later8), ops)."""
    b0 = int(d["blob_off"]) * 16
    bl = plan["blob"][b0:b0 + int(d["blob_bytes"])]
    hdr = bl[:4 * HDR_INTS].view(np.int32)
    chunks = []
    end = 4 * HDR_INTS
    for w in range(MAXW):
        o, nf, nl, df, dl, nops = (int(v) for v in hdr[8 * w:8 * w + 6])
        assert 16 * o == end and (hdr[8 * w + 6:8 * w + 8] == 0).all()  # chunks are contiguous
        at = 16 * o
        own = (bl[at:at + 16 * nf].view(PENTF), bl[at + 16 * nf:at + 16 * nf + 8 * nl].view(PENT))
        at += 16 * nf + al16(8 * nl)
        dep = (bl[at:at + 16 * df].view(PENTF), bl[at + 16 * df:at + 16 * df + 8 * dl].view(PENT))
        at += 16 * df + al16(8 * dl)
        ops = bl[at:at + 4 * nops].view(np.uint32)
        end = at + al16(4 * nops)
        chunks.append((own, dep, ops))
    assert end == int(d["blob_bytes"])
    assert sum(len(c[0][0]) + len(c[1][0]) for c in chunks) == int(d["n_first"])
    assert sum(len(c[2]) for c in chunks) == int(d["n_merge"])
    return chunks


def replay(plan, Ke, mass_of_hpos, h, nnz):
    """Ke[e] = h*K_e (30x30, dof = 3*a + i); mass_of_hpos(p) = M_ij of the block whose (0,0) entry sits at H[p].
    Returns H values in the scalar-CSR layout of the library."""
    H = np.full(9 * nnz, np.nan)
    cp, desc, el = (plan[k] for k in ("color_ptr", "desc", "elems"))
    for c in range(len(cp) - 1):
        for p in range(cp[c], cp[c + 1]):
            d = desc[p]
            acc = np.full((8, CAPU, 3, 3), np.nan)  # never zero-filled: the first touch of a block stores
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
                                fs = (int(pe["first"][bit >> 5]) >> (bit & 31)) & 1
                                assert fs == int(np.isnan(acc[w, u]).all())
                                acc[w, u] = (0.0 if fs else acc[w, u]) + (blk.T if tr else blk)
            chunks = patch_sections(plan, d)
            acc = acc.reshape(-1)  # block addresses are offsets in doubles: warp * 9 CAPU + 9 slot
            ablk = lambda a: acc[a:a + 9].reshape(3, 3)  # noqa: E731

            def write_segment(seg):
                for first, ents in ((True, seg[0]), (False, seg[1])):
                    for en in ents:
                        s_ = int(en["src"])
                        b = ablk(s_ & 0x3fff)
                        assert not np.isnan(b).any()
                        val = b.T if s_ >> 15 else b
                        pos = int(d["hbase"]) + int(en["hoff"])
                        for dd in range(3):
                            base = pos + dd * int(en["rowlen"])
                            old = np.where(np.arange(3) == dd, mass_of_hpos(pos) / h, 0.0) if first else H[base:base + 3]
                            assert first or not np.isnan(old).any()
                            assert not first or np.isnan(H[base:base + 3]).all()
                            H[base:base + 3] = old + val[dd]

            all_ops = [int(op) for c in chunks for op in c[2]]
            merged = set(op & 0x3fff for op in all_ops)
            sources = set(op >> 16 for op in all_ops)
            for w, (own, dep, ops) in enumerate(chunks):  # (a) every warp writes the blocks it holds alone, before any merge
                for part in own:
                    for en in part:
                        a = int(en["src"]) & 0x3fff
                        assert a // (9 * CAPU) == w and a not in merged and a not in sources
                write_segment(own)
            seen_dst = set()
            for own, dep, ops in chunks:  # (b) warp w: dst += src in list order, one dst after the other ...
                mine, prev = set(), None
                for op in ops:
                    dst, cont, src = int(op) & 0x3fff, (int(op) >> 15) & 1, int(op) >> 16
                    assert cont == int(dst == prev) and (cont or dst not in seen_dst)  # the ops of a dst are consecutive
                    assert src // (9 * CAPU) > dst // (9 * CAPU)  # into the lowest warp, ascending warp order
                    ablk(dst)[:] += ablk(src)
                    seen_dst.add(dst)
                    mine.add(dst)
                    prev = dst
                for part in dep:  # (c) ... then the entries of exactly those dsts
                    for en in part:
                        assert (int(en["src"]) & 0x3fff) in mine
                write_segment(dep)
    return H


def mass_lookup(off, cols, mass):
    """H position of a block's (0,0) entry -> its mass value."""
    table = {}
    for i in range(off.shape[0] - 1):
        for s in range(off[i + 1] - off[i]):
            table[9 * int(off[i]) + 3 * s] = mass[off[i] + s]
    return table.__getitem__


def bank_conflict_factor(plan):
    """Average shared-memory wavefronts per half-warp add of the accumulation step (1.0 = conflict-free)."""
    tot = cnt = 0
    for pe in plan["elems"]:
        for tt in range(4):
            res = [((int(w) >> (8 * tt)) & 255) % 16 for t, w in enumerate(pe["blk4"]) if not (TILES[t][0] == TILES[t][1] and tt == 2)]
            tot += max(np.bincount(res, minlength=16))
            cnt += 1
    return tot / cnt


def check_plan_invariants(plan, E, W):
    el, desc, cp, rn, rp = plan["elems"], plan["desc"], plan["color_ptr"], plan["row_node"], plan["row_ptr"]
    assert sorted(el["elem"].tolist()) == list(range(E.shape[0]))
    assert (el["node"] == E[el["elem"]]).all()
    assert cp[0] == 0 and cp[-1] == len(desc)
    for c in range(len(cp) - 1):  # patches of one colour share no coefficient
        seen = set()
        for p in range(cp[c], cp[c + 1]):
            nodes = set(rn[rp[p]:rp[p + 1]].tolist())
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
    H = replay(plan, Ke, mass_lookup(o.off, o.cols, o.mass(rho)), h, o.cols.shape[0])
    assert not np.isnan(H).any()
    assert rel_err(H, o.hessian(x, mat, h, rho)) < 1e-13
    if n >= 4 and W == 8:  # even block: every patch is a 2x2x2 brick of cells, 8 colours
        assert len(plan["color_ptr"]) - 1 == 8
        assert (plan["desc"]["n_sub"] == 8).all()
    assert bank_conflict_factor(plan) < 1.35  # random placement gives ~2.3


def test_plan_chunked_launch_order():
    """Host-buffer pipelining: launches ordered (id-range chunk, colour).  Replaying that order still reproduces the
    oracle's H (the first-touch flags follow the launch order), a chunk only reads ids below the end of its range."""
    from tlfea import meshgen
    from oracle import oracle as orc

    X, E = meshgen.kuhn_block(4)
    x = meshgen.perturb(X, 0.01, 9)
    o = orc.T10Oracle(E, X)
    mat = orc.Material.svk(7e8, 0.33)
    h, rho, K = 1e-3, 2700.0, 4
    plan = build_plan(E, X, o.off, o.cols, 4, K)
    check_plan_invariants(plan, E, 4)
    cp, rn, rp = plan["color_ptr"], plan["row_node"], plan["row_ptr"]
    C_ = (len(cp) - 1) // K
    assert (len(cp) - 1) == K * C_ and C_ >= 8
    nn = X.shape[0]
    for c in range(K):
        for p in range(cp[c * C_], cp[(c + 1) * C_]):
            mx = rn[rp[p]:rp[p + 1]].max()
            tot = 4 * (K - 2) + 2  # csrc/patch_plan.cuh pk_chunk_bound: short first and last id range
            bound = [nn * (0 if r == 0 else tot if r == K else 1 + 4 * (r - 1)) // tot for r in range(K + 1)]
            assert bound[c] <= mx < bound[c + 1]
    Ke = [h * orc.element_tangent(E[e], x, o.grad_N[e], o.detJ[e], o.qw, mat) for e in range(E.shape[0])]
    H = replay(plan, Ke, mass_lookup(o.off, o.cols, o.mass(rho)), h, o.cols.shape[0])
    assert rel_err(H, o.hessian(x, mat, h, rho)) < 1e-13


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
        H = replay(plan, Ke, mass_lookup(o.off, o.cols, o.mass(rho)), h, o.cols.shape[0])
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
