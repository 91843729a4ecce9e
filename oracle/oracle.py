"""ctypes front-end for the CPU oracle (TEST INFRASTRUCTURE ONLY).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
`--impl reference` legs may import this module.  The product package never does.
Each wrapper names the reference file:line its C body restates (see
oracle/tlfea_oracle.c).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int)


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "tlfea_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "liboracle.so"])
    return _LIB_PATH


class Material(C.Structure):
    """orc_material; mirrors the setters FEAT10Data.cuh:538-634."""

    _fields_ = [
        ("model", C.c_int),
        ("lam", C.c_double),
        ("mu", C.c_double),
        ("mu10", C.c_double),
        ("mu01", C.c_double),
        ("kappa", C.c_double),
        ("eta_damp", C.c_double),
        ("lambda_damp", C.c_double),
    ]

    @staticmethod
    def svk(E: float, nu: float, eta: float = 0.0, lam_d: float = 0.0) -> "Material":
        mu = E / (2 * (1 + nu))
        lam = (E * nu) / ((1 + nu) * (1 - 2 * nu))
        return Material(0, lam, mu, 0.0, 0.0, 0.0, eta, lam_d)

    @staticmethod
    def mooney_rivlin(mu10: float, mu01: float, kappa: float, eta: float = 0.0, lam_d: float = 0.0) -> "Material":
        return Material(1, 0.0, 0.0, mu10, mu01, kappa, eta, lam_d)


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.orc_coef_pattern.restype = C.c_int64
    return _lib


def _d(a):
    return a.ctypes.data_as(c_dp)


def _i(a):
    return a.ctypes.data_as(c_ip)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def t10_quadrature():
    """Keast 5-point rule (quadrature_utils.h:135-158) -> qx,qy,qz,qw."""
    q = [np.zeros(5) for _ in range(4)]
    lib().orc_t10_quadrature(*[_d(a) for a in q])
    return q


def conn_colmajor(elements: np.ndarray) -> np.ndarray:
    """(n_elem, S) int -> flat column-major int32 as Eigen::MatrixXi stores it."""
    return _i32(np.asarray(elements).T.reshape(-1))


def t10_precompute(elements, X):
    """dn_du_pre_kernel (FEAT10Data.cu:97-204) -> grad_N[ne,5,30], detJ[ne,5]."""
    ne = elements.shape[0]
    conn = conn_colmajor(elements)
    x, y, z = (_f64(X[:, k]) for k in range(3))
    qx, qy, qz, qw = t10_quadrature()
    g = np.zeros((ne, 5, 30))
    dj = np.zeros((ne, 5))
    lib().orc_t10_precompute(ne, _i(conn), _d(x), _d(y), _d(z), _d(qx), _d(qy), _d(qz), _d(g), _d(dj))
    return g, dj


def coef_pattern(ecoef, n_coef):
    """BuildMassCSRPattern (FEAT10Data.cu:372-440) -> offsets[n_coef+1], columns[nnz]."""
    ecoef = _i32(ecoef)
    ne, S = ecoef.shape
    off = np.zeros(n_coef + 1, dtype=np.int32)
    colp = c_ip()
    nnz = lib().orc_coef_pattern(ne, S, _i(ecoef), n_coef, _i(off), C.byref(colp))
    cols = np.ctypeslib.as_array(colp, shape=(max(nnz, 1),))[:nnz].copy()
    lib().orc_free(colp)
    return off, cols


def t10_mass(elements, detJ, rho, off, cols):
    """mass_matrix_qp_kernel (FEAT10Data.cu:206-278) -> values[nnz]."""
    ne = elements.shape[0]
    conn = conn_colmajor(elements)
    qx, qy, qz, qw = t10_quadrature()
    vals = np.zeros(cols.shape[0])
    detJ = _f64(detJ)
    lib().orc_t10_mass(ne, _i(conn), _d(detJ), _d(qx), _d(qy), _d(qz), _d(qw), C.c_double(rho), _i(off), _i(cols), _d(vals))
    return vals


def compute_p(ecoef, coords, grad_N, mat: Material, v=None, want_visc=False):
    """compute_p (FEAT10DataFunc.cuh:85-293 and the ANCF twins) -> F,P[,Fdot,Pvis] as [ne,Q,9]."""
    ecoef = _i32(ecoef)
    ne, S = ecoef.shape
    Q = grad_N.shape[1]
    cx, cy, cz = (_f64(coords[:, k]) for k in range(3))
    grad_N = _f64(grad_N)
    F = np.zeros((ne, Q, 9))
    P = np.zeros((ne, Q, 9))
    Fd = np.zeros((ne, Q, 9))
    Pv = np.zeros((ne, Q, 9))
    vv = None if v is None else _f64(v)
    lib().orc_compute_p(ne, S, Q, _i(ecoef), _d(cx), _d(cy), _d(cz), _d(grad_N), C.byref(mat),
                        None if vv is None else _d(vv), _d(F), _d(P), _d(Fd), _d(Pv))
    if want_visc:
        return F, P, Fd, Pv
    return F, P


def internal_force(ecoef, grad_N, detJ, qw, P, n_coef, f_int=None):
    """compute_internal_force (FEAT10DataFunc.cuh:397-458); accumulates like the reference."""
    ecoef = _i32(ecoef)
    ne, S = ecoef.shape
    Q = grad_N.shape[1]
    f = np.zeros(3 * n_coef) if f_int is None else _f64(f_int)
    grad_N, detJ, qw, P = _f64(grad_N), _f64(detJ), _f64(qw), _f64(P)
    lib().orc_internal_force(ne, S, Q, _i(ecoef), _d(grad_N), _d(detJ), _d(qw), _d(P), _d(f))
    return f


def dof_pattern(off, cols, want_cols=True):
    """SyncedNewton.cu:162-205 -> row_offsets int64[3n+1], col_indices int32[9 nnz]."""
    n = off.shape[0] - 1
    ro = np.zeros(3 * n + 1, dtype=np.int64)
    ci = np.zeros(9 * cols.shape[0], dtype=np.int32) if want_cols else None
    lib().orc_dof_pattern(n, _i(off), _i(cols), ro.ctypes.data_as(C.POINTER(C.c_int64)), None if ci is None else _i(ci))
    return ro, ci


def tangent_assemble(ecoef, coords, grad_N, detJ, qw, mat: Material, h, off, cols, H=None):
    """assemble_sparse_hessian_tangent (SyncedNewton.cu:262-288) ->
    compute_hessian_assemble_csr (FEAT10DataFunc.cuh:512-791): H += h*K_t (+ C_visc)."""
    ecoef = _i32(ecoef)
    ne, S = ecoef.shape
    Q = grad_N.shape[1]
    cx, cy, cz = (_f64(coords[:, k]) for k in range(3))
    grad_N, detJ, qw = _f64(grad_N), _f64(detJ), _f64(qw)
    Hv = np.zeros(9 * cols.shape[0]) if H is None else H
    lib().orc_tangent_assemble(ne, S, Q, _i(ecoef), _d(cx), _d(cy), _d(cz), _d(grad_N), _d(detJ), _d(qw),
                               C.byref(mat), C.c_double(h), _i(off), _i(cols), _d(Hv))
    return Hv


def element_tangent(ec, coords, grad_N_e, detJ_e, qw, mat: Material):
    """Dense 3S x 3S tangent of one element (h=1), for comparison with the numpy prototypes."""
    ec = _i32(ec)
    S = ec.shape[0]
    Q = grad_N_e.shape[0]
    cx, cy, cz = (_f64(coords[:, k]) for k in range(3))
    K = np.zeros((3 * S, 3 * S))
    g, dj, qw = _f64(grad_N_e), _f64(detJ_e), _f64(qw)
    lib().orc_element_tangent(S, Q, _i(ec), _d(cx), _d(cy), _d(cz), _d(g), _d(dj), _d(qw), C.byref(mat), _d(K))
    return K


def hessian_mass(off, cols, mass_values, h, H):
    """assemble_sparse_hessian_mass (SyncedNewton.cu:213-259): H += M/h on block diagonals."""
    mass_values = _f64(mass_values)
    lib().orc_hessian_mass(off.shape[0] - 1, _i(off), _i(cols), _d(mass_values), C.c_double(h), _d(H))
    return H


def hessian_constraints(j_off, j_cols, j_vals, h, rho, off, cols, H):
    """assemble_sparse_hessian_constraints (SyncedNewton.cu:291-341): H += h^2 rho J^T J."""
    j_off, j_cols, j_vals = _i32(j_off), _i32(j_cols), _f64(j_vals)
    lib().orc_hessian_constraints(j_off.shape[0] - 1, _i(j_off), _i(j_cols), _d(j_vals), C.c_double(h), C.c_double(rho),
                                  _i(off), _i(cols), _d(H))
    return H


def grad_L(off, cols, mass_values, v_guess, v_prev, f_int, f_ext, h, jt=None, lam=None, con=None, rho=0.0):
    """solver_grad_L (SyncedNewton.cu:343-407)."""
    n = off.shape[0] - 1
    g = np.zeros(3 * n)
    args = [_f64(a) for a in (mass_values, v_guess, v_prev, f_int, f_ext)]
    if jt is None:
        lib().orc_grad_L(n, _i(off), _i(cols), *[_d(a) for a in args], C.c_double(h), 0, None, None, None, None, None,
                         C.c_double(0.0), _d(g))
    else:
        jo, jc, jv = _i32(jt[0]), _i32(jt[1]), _f64(jt[2])
        lam, con = _f64(lam), _f64(con)
        lib().orc_grad_L(n, _i(off), _i(cols), *[_d(a) for a in args], C.c_double(h), lam.shape[0], _i(jo), _i(jc),
                         _d(jv), _d(lam), _d(con), C.c_double(rho), _d(g))
    return g


def fixed_node_jacobians(fixed_nodes, n_coef):
    """J (rows = constraints) and J^T (rows = DOFs) for pinned nodes:
    FEAT10Data.cu:443-496 (J has one unit entry per row at DOF 3*node+d)."""
    fixed_nodes = np.asarray(fixed_nodes, dtype=np.int32)
    nc = 3 * fixed_nodes.shape[0]
    j_off = np.arange(nc + 1, dtype=np.int32)
    j_cols = (3 * np.repeat(fixed_nodes, 3) + np.tile(np.arange(3), fixed_nodes.shape[0])).astype(np.int32)
    j_vals = np.ones(nc)
    counts = np.zeros(3 * n_coef + 1, dtype=np.int32)
    np.add.at(counts, j_cols + 1, 1)
    jt_off = np.cumsum(counts).astype(np.int32)
    order = np.argsort(j_cols, kind="stable")
    jt_cols = order.astype(np.int32)
    jt_vals = np.ones(nc)
    return (j_off, j_cols, j_vals), (jt_off, jt_cols, jt_vals)


def num_threads() -> int:
    return lib().orc_num_threads()


def set_num_threads(n: int) -> None:
    lib().orc_set_num_threads(int(n))


@dataclass
class T10Oracle:
    """Convenience bundle: the FEAT10 pipeline the drivers run before and inside
    the Newton loop."""

    elements: np.ndarray
    X: np.ndarray

    def __post_init__(self):
        self.ne = self.elements.shape[0]
        self.nn = self.X.shape[0]
        self.ecoef = _i32(self.elements)
        self.qw = t10_quadrature()[3]
        self.grad_N, self.detJ = t10_precompute(self.elements, self.X)
        self.off, self.cols = coef_pattern(self.ecoef, self.nn)

    def mass(self, rho):
        return t10_mass(self.elements, self.detJ, rho, self.off, self.cols)

    def force(self, x, mat, v=None):
        F, P = compute_p(self.ecoef, x, self.grad_N, mat, v)
        return internal_force(self.ecoef, self.grad_N, self.detJ, self.qw, P, self.nn), F, P

    def tangent(self, x, mat, h, H=None):
        return tangent_assemble(self.ecoef, x, self.grad_N, self.detJ, self.qw, mat, h, self.off, self.cols, H)

    def hessian(self, x, mat, h, rho_mass, fixed_nodes=None, alm_rho=0.0):
        """H = M/h + h K_t (+C_visc) + h^2 rho J^T J in the reference's order
        (SyncedNewton.cu:1080-1097)."""
        H = np.zeros(9 * self.cols.shape[0])
        hessian_mass(self.off, self.cols, self.mass(rho_mass), h, H)
        self.tangent(x, mat, h, H)
        if fixed_nodes is not None and len(fixed_nodes):
            J, _ = fixed_node_jacobians(fixed_nodes, self.nn)
            hessian_constraints(*J, h, alm_rho, self.off, self.cols, H)
        return H


# --------------------------------------------------------------------------- ANCF set-up path
_GL = {
    2: ([-0.57735026918962576, 0.57735026918962576], [1.0, 1.0]),
    3: ([-0.77459666924148340, 0.0, 0.77459666924148340],
        [0.55555555555555556, 0.88888888888888889, 0.55555555555555556]),
    -3: ([-0.7745966692414834, 0.0, 0.7745966692414834], [0.5555555555555556, 0.8888888888888888, 0.5555555555555556]),
    4: ([-0.8611363115940526, -0.3399810435848563, 0.3399810435848563, 0.8611363115940526],
        [0.3478548451374538, 0.6521451548625461, 0.6521451548625461, 0.3478548451374538]),
    6: ([-0.93246951420315202, -0.66120938646626451, -0.23861918608319691, 0.23861918608319691,
         0.66120938646626451, 0.93246951420315202],
        [0.17132449237917034, 0.36076157304813861, 0.46791393457269104, 0.46791393457269104,
         0.36076157304813861, 0.17132449237917034]),
    7: ([-0.949107912342759, -0.741531185599394, -0.405845151377397, 0.0, 0.405845151377397, 0.741531185599394,
         0.949107912342759],
        [0.129484966168870, 0.279705391489277, 0.381830050505119, 0.417959183673469, 0.381830050505119,
         0.279705391489277, 0.129484966168870]),
}


def ancf_rule(kind: int, mass: bool):
    """(n[3], points[3][7], weights[3][7]) of lib_utils/quadrature_utils.h: stiffness 3x2x2 / 4x4x3,
    mass 6x2x2 / 7x7x3."""
    ax = {(3243, False): (3, 2, 2), (3243, True): (6, 2, 2), (3443, False): (4, 4, 3), (3443, True): (7, 7, -3)}[(kind, mass)]
    n = np.array([abs(a) for a in ax], dtype=np.int32)
    p = np.zeros((3, 7)); w = np.zeros((3, 7))
    for k, a in enumerate(ax):
        p[k, :abs(a)], w[k, :abs(a)] = _GL[a]
    return n, p, w


def ancf_stiffness_weights(kind: int):
    n, _, w = ancf_rule(kind, False)
    return np.array([w[0][q // (n[1] * n[2])] * w[1][(q // n[2]) % n[1]] * w[2][q % n[2]]
                     for q in range(int(n.prod()))])


def ancf_ecoef(kind: int, conn):
    conn = np.asarray(conn)
    S = 8 if kind == 3243 else 16
    return np.array([[4 * conn[e, a // 4] + a % 4 for a in range(S)] for e in range(conn.shape[0])], dtype=np.int32)


def ancf_binv(kind: int, L, W, H):
    S = 8 if kind == 3243 else 16
    L, W, H = (np.atleast_1d(np.asarray(a, dtype=np.float64)) for a in (L, W, H))
    out = np.zeros((L.shape[0], S * S))
    for e in range(L.shape[0]):
        if lib().orc_ancf_binv(S, C.c_double(L[e]), C.c_double(W[e]), C.c_double(H[e]), _d(out[e])) != 0:
            raise ValueError("singular B matrix")
    return out


def ancf_precompute(kind, ecoef, x, y, z, LWH, Binv):
    ecoef = _i32(ecoef)
    ne, S = ecoef.shape
    n, p, _ = ancf_rule(kind, False)
    Q = int(n.prod())
    g = np.zeros((ne, Q, 3 * S)); dj = np.zeros((ne, Q))
    x, y, z, LWH, Binv, p = _f64(x), _f64(y), _f64(z), _f64(LWH), _f64(Binv), _f64(p)
    lib().orc_ancf_precompute(ne, S, _i(ecoef), _d(x), _d(y), _d(z), _d(LWH), _d(Binv), _i(n), _d(p), _d(g), _d(dj))
    return g, dj


def ancf_mass(kind, ecoef, x, y, z, LWH, Binv, rho, off, cols):
    ecoef = _i32(ecoef)
    ne, S = ecoef.shape
    n, p, w = ancf_rule(kind, True)
    vals = np.zeros(cols.shape[0])
    x, y, z, LWH, Binv, p, w = _f64(x), _f64(y), _f64(z), _f64(LWH), _f64(Binv), _f64(p), _f64(w)
    lib().orc_ancf_mass(ne, S, _i(ecoef), _d(x), _d(y), _d(z), _d(LWH), _d(Binv), _i(n), _d(p), _d(w),
                        C.c_double(rho), _i(off), _i(cols), _d(vals))
    return vals
