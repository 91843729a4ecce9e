"""Host-side mirror of the reference's element objects, over the C ABI.

Method names, argument meaning and call order follow
lib_src/elements/{ElementBase.h, FEAT10Data.cuh, ANCF3243Data.cuh, ANCF3443Data.cuh};
numpy arrays stand where the reference takes Eigen vectors/matrices.  Misuse raises
TlfeaError where the reference prints to std::cerr and returns.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import c_dp, c_ip, check, lib

TYPE_3243, TYPE_3443, TYPE_T10 = 0, 1, 2  # ElementBase.h:19


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _d(a):
    return a.ctypes.data_as(c_dp)


def _i(a):
    return a.ctypes.data_as(c_ip)


def tet5pt_quadrature():
    """Keast 5-point rule, same numbers as Quadrature::tet5pt_* (quadrature_utils.h:135-158)."""
    a, b = 0.5, 1.0 / 6.0
    bary = np.array([[0.25] * 4, [a, b, b, b], [b, a, b, b], [b, b, a, b], [b, b, b, a]])
    w = np.array([-4.0 / 5.0, 9.0 / 20.0, 9.0 / 20.0, 9.0 / 20.0, 9.0 / 20.0]) * (1.0 / 6.0)
    return bary[:, 1].copy(), bary[:, 2].copy(), bary[:, 3].copy(), w


class ElementBase:
    """ElementBase (lib_src/elements/ElementBase.h:21-49)."""

    type: int = -1

    def __init__(self, etype: int, n_elem: int, n_nodes: int, device: int = 0):
        self.type = etype
        self._h = _lib.c_vp()
        check(lib.tlfea_elem_create(etype, int(n_elem), int(n_nodes), int(device), C.byref(self._h)))
        self.n_elem = lib.tlfea_elem_n_elem(self._h)
        self.n_coef = lib.tlfea_elem_n_coef(self._h)
        self.n_shape = lib.tlfea_elem_n_shape(self._h)
        self.n_qp = lib.tlfea_elem_n_qp(self._h)
        self.device = device

    # ---- lifecycle -----------------------------------------------------
    def Initialize(self):
        """Initialize() (FEAT10Data.cuh:382-433): allocation happens at construction here."""
        return None

    def Destroy(self):
        if self._h:
            check(lib.tlfea_elem_destroy(self._h))
            self._h = _lib.c_vp()

    def __del__(self):
        try:
            self.Destroy()
        except Exception:
            pass

    def get_n_beam(self):
        return self.n_elem

    def get_n_coef(self):
        return self.n_coef

    # ---- parameters ----------------------------------------------------
    def SetDensity(self, rho0):
        check(lib.tlfea_elem_set_density(self._h, float(rho0)))

    def SetDamping(self, eta_damp, lambda_damp):
        check(lib.tlfea_elem_set_damping(self._h, float(eta_damp), float(lambda_damp)))

    def SetSVK(self, E, nu):
        check(lib.tlfea_elem_set_svk(self._h, float(E), float(nu)))

    def SetMooneyRivlin(self, mu10, mu01, kappa):
        check(lib.tlfea_elem_set_mooney_rivlin(self._h, float(mu10), float(mu01), float(kappa)))

    def SetExternalForce(self, f_ext):
        f = _f64(f_ext)
        check(lib.tlfea_elem_set_external_force(self._h, _d(f), f.shape[0]))

    def SetNodalFixed(self, fixed_nodes):
        fx = _i32(fixed_nodes)
        check(lib.tlfea_elem_set_nodal_fixed(self._h, _i(fx), fx.shape[0]))

    def UpdateNodalFixed(self, fixed_nodes):
        f = _i32(fixed_nodes)
        check(lib.tlfea_elem_update_nodal_fixed(self._h, _i(f), f.shape[0]))

    def SetLinearConstraintsCSR(self, offsets, columns, values, rhs):
        o, c, v, r = _i32(offsets), _i32(columns), _f64(values), _f64(rhs)
        check(lib.tlfea_elem_set_linear_constraints_csr(self._h, o.shape[0] - 1, _i(o), _i(c), _d(v), _d(r)))

    def UpdatePositions(self, x, y, z):
        x, y, z = _f64(x), _f64(y), _f64(z)
        check(lib.tlfea_elem_update_positions(self._h, _d(x), _d(y), _d(z), x.shape[0]))

    def UpdateConstraintTargets(self, x, y, z):
        x, y, z = _f64(x), _f64(y), _f64(z)
        check(lib.tlfea_elem_update_constraint_targets(self._h, _d(x), _d(y), _d(z), x.shape[0]))

    # ---- computation ---------------------------------------------------
    def BuildMassCSRPattern(self):
        check(lib.tlfea_elem_build_mass_pattern(self._h))

    def CalcMassMatrix(self):
        check(lib.tlfea_elem_calc_mass(self._h))

    def CalcP(self):
        check(lib.tlfea_elem_calc_p(self._h))

    def CalcInternalForce(self):
        check(lib.tlfea_elem_calc_internal_force(self._h))

    def ClearInternalForce(self):
        check(lib.tlfea_elem_clear_internal_force(self._h))

    def CalcConstraintData(self):
        check(lib.tlfea_elem_calc_constraint(self._h))

    def SetStream(self, cuda_stream: int):
        """Run the handle's work on a caller-owned cudaStream_t (raw pointer value)."""
        check(lib.tlfea_elem_set_stream(self._h, cuda_stream))

    def Sync(self):
        check(lib.tlfea_elem_sync(self._h))

    # ---- retrieval -----------------------------------------------------
    def RetrieveInternalForceToCPU(self, out=None):
        """out: caller's buffer like the reference's VectorXd& (a pinned one makes the copy a plain DMA)."""
        if out is None:
            out = np.empty(3 * self.n_coef)
        assert out.dtype == np.float64 and out.flags.c_contiguous and out.shape[0] == 3 * self.n_coef
        check(lib.tlfea_elem_get_internal_force(self._h, _d(out), out.shape[0]))
        return out

    def RetrievePositionToCPU(self):
        x, y, z = (np.empty(self.n_coef) for _ in range(3))
        check(lib.tlfea_elem_get_positions(self._h, _d(x), _d(y), _d(z), self.n_coef))
        return x, y, z

    def RetrieveDeformationGradientToCPU(self):
        """[n_elem, Q, 9]; each 9 is a column-major 3x3 like the reference's Eigen::Map."""
        out = np.empty((self.n_elem, self.n_qp, 9))
        check(lib.tlfea_elem_get_F(self._h, _d(out), out.size))
        return out

    def RetrievePFromFToCPU(self):
        out = np.empty((self.n_elem, self.n_qp, 9))
        check(lib.tlfea_elem_get_P(self._h, _d(out), out.size))
        return out

    def RetrieveDetJToCPU(self):
        out = np.empty((self.n_elem, self.n_qp))
        check(lib.tlfea_elem_get_detJ(self._h, _d(out), out.size))
        return out

    def _retrieve_grad(self):
        out = np.empty((self.n_elem, self.n_qp, 3 * self.n_shape))
        check(lib.tlfea_elem_get_grad_N(self._h, _d(out), out.size))
        return out

    def RetrieveMassCSRToCPU(self):
        nnz = lib.tlfea_elem_mass_nnz(self._h)
        if nnz < 0:
            return np.zeros(self.n_coef + 1, dtype=np.int32), np.zeros(0, dtype=np.int32), np.zeros(0)
        off = np.empty(self.n_coef + 1, dtype=np.int32)
        cols = np.empty(nnz, dtype=np.int32)
        vals = np.empty(nnz)
        check(lib.tlfea_elem_get_mass_csr(self._h, _i(off), _i(cols), _d(vals)))
        n = int(off[-1])  # < nnz when the stored pattern also holds constraint couplings
        return off, cols[:n], vals[:n]

    def RetrieveConstraintDataToCPU(self):
        n = lib.tlfea_elem_n_constraint(self._h)
        out = np.empty(n)
        check(lib.tlfea_elem_get_constraint(self._h, _d(out), n))
        return out

    def RetrieveConnectivityToCPU(self):
        raise NotImplementedError

    def GetExternalForceDevicePtr(self):
        return lib.tlfea_elem_devptr_f_ext(self._h)

    def PackNodesXYZ(self, d_nodes_xyz: int):
        """CollisionSystemInput::d_nodes_xyz = [x | y | z] from the current positions (device pointer, 3 n doubles)."""
        check(lib.tlfea_elem_pack_nodes_xyz(self._h, C.c_void_p(int(d_nodes_xyz)), self.n_coef))

    def SetExternalForceDevice(self, d_base: int, d_contact: int = 0, alpha: float = 1.0):
        """f_ext = base + alpha * contact on the device (gravity + contact forces, test_item_drop.cc:589-598)."""
        check(lib.tlfea_elem_set_external_force_device(self._h, C.c_void_p(int(d_base)),
                                                       C.c_void_p(int(d_contact)) if d_contact else None, float(alpha),
                                                       3 * self.n_coef))

    def GetX12DevicePtr(self):
        return lib.tlfea_elem_devptr_x(self._h)

    def GetY12DevicePtr(self):
        return lib.tlfea_elem_devptr_y(self._h)

    def GetZ12DevicePtr(self):
        return lib.tlfea_elem_devptr_z(self._h)

    def GetInternalForceDevicePtr(self):
        return lib.tlfea_elem_devptr_f_int(self._h)


class GPU_FEAT10_Data(ElementBase):
    """GPU_FEAT10_Data (lib_src/elements/FEAT10Data.cuh:19-852)."""

    def __init__(self, num_elements: int, num_nodes: int, device: int = 0):
        super().__init__(TYPE_T10, num_elements, num_nodes, device)

    def Setup(self, tet5pt_x, tet5pt_y, tet5pt_z, tet5pt_weights, h_x12, h_y12, h_z12, element_connectivity):
        """Setup(...) (FEAT10Data.cuh:435-533).  element_connectivity is (n_elem, 10); it is
        passed down column-major, as Eigen::MatrixXi::data() would be."""
        conn = np.asarray(element_connectivity)
        if conn.shape != (self.n_elem, 10):
            raise _lib.TlfeaError(-1, f"connectivity must be ({self.n_elem}, 10)")
        cm = _i32(conn.T.reshape(-1))
        q = [_f64(a) for a in (tet5pt_x, tet5pt_y, tet5pt_z, tet5pt_weights)]
        x, y, z = _f64(h_x12), _f64(h_y12), _f64(h_z12)
        for a in (x, y, z):
            if a.shape[0] != self.n_coef:
                raise _lib.TlfeaError(-1, "coordinate vector size mismatch")
        check(lib.tlfea_elem_setup_t10(self._h, *[_d(a) for a in q], _d(x), _d(y), _d(z), _i(cm)))

    def CalcDnDuPre(self):
        check(lib.tlfea_elem_precompute(self._h))

    def RetrieveDnDuPreToCPU(self):
        """[n_elem, 5, 30]; entry (a, j) of a Gauss point at [j*10 + a] (FEAT10Data.cuh:41-45)."""
        return self._retrieve_grad()

    def RetrieveConnectivityToCPU(self):
        out = np.empty(self.n_elem * 10, dtype=np.int32)
        check(lib.tlfea_elem_get_connectivity(self._h, _i(out), out.size))
        return out.reshape(10, self.n_elem).T.copy()


class _ANCFBase(ElementBase):
    def CalcDsDuPre(self):
        check(lib.tlfea_elem_precompute(self._h))

    def RetrieveDsDuPreToCPU(self):
        return self._retrieve_grad()


class GPU_ANCF3243_Data(_ANCFBase):
    """GPU_ANCF3243_Data (lib_src/elements/ANCF3243Data.cuh:33-1153); ctor takes (n_nodes, n_elem)."""

    def __init__(self, num_nodes: int, num_elements: int, device: int = 0):
        super().__init__(TYPE_3243, num_elements, num_nodes, device)

    def Setup(self, L, W, H, h_x12, h_y12, h_z12, element_connectivity):
        """Per-element L, W, H (scalars broadcast), 4 coefficients per node, connectivity (n_elem, 2)."""
        L, W, H = (_f64(np.broadcast_to(np.asarray(a, dtype=np.float64), (self.n_elem,))) for a in (L, W, H))
        conn = _i32(np.asarray(element_connectivity).reshape(self.n_elem, 2))  # row-major (ANCF3243Data.cuh:310-312)
        x, y, z = _f64(h_x12), _f64(h_y12), _f64(h_z12)
        check(lib.tlfea_elem_setup_3243(self._h, _d(L), _d(W), _d(H), _d(x), _d(y), _d(z), _i(conn)))


class GPU_ANCF3443_Data(_ANCFBase):
    """GPU_ANCF3443_Data (lib_src/elements/ANCF3443Data.cuh); ctor takes (n_nodes, n_elem)."""

    def __init__(self, num_nodes: int, num_elements: int, device: int = 0):
        super().__init__(TYPE_3443, num_elements, num_nodes, device)

    def Setup(self, L, W, H, h_x12, h_y12, h_z12, element_connectivity):
        L, W, H = (_f64(np.broadcast_to(np.asarray(a, dtype=np.float64), (self.n_elem,))) for a in (L, W, H))
        conn = np.asarray(element_connectivity).reshape(self.n_elem, 4)
        cm = _i32(conn.T.reshape(-1))  # column-major (ANCF3443Data.cuh:212-213)
        x, y, z = _f64(h_x12), _f64(h_y12), _f64(h_z12)
        check(lib.tlfea_elem_setup_3443(self._h, _d(L), _d(W), _d(H), _d(x), _d(y), _d(z), _i(cm)))
