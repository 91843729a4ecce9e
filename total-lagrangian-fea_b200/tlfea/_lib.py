"""ctypes binding of libtlfea_b200.so (the C ABI declared in include/tlfea_b200.h).

The product path: there is NO CPU fallback.  If the shared library is missing this
module raises at import; if no CUDA device is visible, handle creation fails with
TLFEA_ERR_NO_DEVICE.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TLFEA_LIB", os.path.join(os.path.dirname(_HERE), "lib", "libtlfea_b200.so"))  # TLFEA_LIB: A/B builds

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int)
c_i64p = C.POINTER(C.c_int64)
c_vp = C.c_void_p


class NewtonParams(C.Structure):
    """Bit-compatible with SyncedNewtonParams (lib_src/solvers/SyncedNewton.cuh:29-33)."""

    _fields_ = [
        ("inner_atol", C.c_double),
        ("inner_rtol", C.c_double),
        ("outer_tol", C.c_double),
        ("rho", C.c_double),
        ("max_outer", C.c_int),
        ("max_inner", C.c_int),
        ("time_step", C.c_double),
    ]


class TlfeaError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"tlfea status {code}: {msg}")
        self.code = code


# name -> (restype, argtypes).  Every symbol include/tlfea_b200.h declares is listed here;
# tests/test_abi.py checks the two stay in sync.
_SIGS = {
    "tlfea_last_error": (C.c_char_p, []),
    "tlfea_version": (C.c_int, []),
    "tlfea_launch_count": (C.c_int64, []),
    "tlfea_device_count": (C.c_int, []),
    "tlfea_set_host_threads": (C.c_int, [C.c_int]),
    "tlfea_elem_create": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(c_vp)]),
    "tlfea_elem_destroy": (C.c_int, [c_vp]),
    "tlfea_elem_sync": (C.c_int, [c_vp]),
    "tlfea_elem_set_stream": (C.c_int, [c_vp, c_vp]),
    "tlfea_elem_setup_t10": (C.c_int, [c_vp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_ip]),
    "tlfea_elem_setup_3243": (C.c_int, [c_vp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_ip]),
    "tlfea_elem_setup_3443": (C.c_int, [c_vp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_ip]),
    "tlfea_elem_set_density": (C.c_int, [c_vp, C.c_double]),
    "tlfea_elem_set_damping": (C.c_int, [c_vp, C.c_double, C.c_double]),
    "tlfea_elem_set_svk": (C.c_int, [c_vp, C.c_double, C.c_double]),
    "tlfea_elem_set_mooney_rivlin": (C.c_int, [c_vp, C.c_double, C.c_double, C.c_double]),
    "tlfea_elem_set_external_force": (C.c_int, [c_vp, c_dp, C.c_int]),
    "tlfea_elem_set_nodal_fixed": (C.c_int, [c_vp, c_ip, C.c_int]),
    "tlfea_elem_update_nodal_fixed": (C.c_int, [c_vp, c_ip, C.c_int]),
    "tlfea_elem_set_linear_constraints_csr": (C.c_int, [c_vp, C.c_int, c_ip, c_ip, c_dp, c_dp]),
    "tlfea_elem_update_positions": (C.c_int, [c_vp, c_dp, c_dp, c_dp, C.c_int]),
    "tlfea_elem_update_constraint_targets": (C.c_int, [c_vp, c_dp, c_dp, c_dp, C.c_int]),
    "tlfea_elem_precompute": (C.c_int, [c_vp]),
    "tlfea_elem_build_mass_pattern": (C.c_int, [c_vp]),
    "tlfea_elem_calc_mass": (C.c_int, [c_vp]),
    "tlfea_elem_calc_p": (C.c_int, [c_vp]),
    "tlfea_elem_calc_internal_force": (C.c_int, [c_vp]),
    "tlfea_elem_clear_internal_force": (C.c_int, [c_vp]),
    "tlfea_elem_calc_constraint": (C.c_int, [c_vp]),
    "tlfea_elem_n_elem": (C.c_int, [c_vp]),
    "tlfea_elem_n_coef": (C.c_int, [c_vp]),
    "tlfea_elem_n_shape": (C.c_int, [c_vp]),
    "tlfea_elem_n_qp": (C.c_int, [c_vp]),
    "tlfea_elem_n_constraint": (C.c_int, [c_vp]),
    "tlfea_elem_mass_nnz": (C.c_int64, [c_vp]),
    "tlfea_elem_get_internal_force": (C.c_int, [c_vp, c_dp, C.c_int]),
    "tlfea_elem_get_positions": (C.c_int, [c_vp, c_dp, c_dp, c_dp, C.c_int]),
    "tlfea_elem_get_F": (C.c_int, [c_vp, c_dp, C.c_int64]),
    "tlfea_elem_get_P": (C.c_int, [c_vp, c_dp, C.c_int64]),
    "tlfea_elem_get_grad_N": (C.c_int, [c_vp, c_dp, C.c_int64]),
    "tlfea_elem_get_detJ": (C.c_int, [c_vp, c_dp, C.c_int64]),
    "tlfea_elem_get_mass_csr": (C.c_int, [c_vp, c_ip, c_ip, c_dp]),
    "tlfea_elem_get_constraint": (C.c_int, [c_vp, c_dp, C.c_int]),
    "tlfea_elem_get_connectivity": (C.c_int, [c_vp, c_ip, C.c_int64]),
    "tlfea_elem_constraint_jacobian_nnz": (C.c_int64, [c_vp]),
    "tlfea_elem_get_constraint_jacobian_csr": (C.c_int, [c_vp, c_ip, c_ip, c_dp]),
    "tlfea_elem_devptr_x": (c_vp, [c_vp]),
    "tlfea_elem_devptr_y": (c_vp, [c_vp]),
    "tlfea_elem_devptr_z": (c_vp, [c_vp]),
    "tlfea_elem_devptr_f_ext": (c_vp, [c_vp]),
    "tlfea_elem_devptr_f_int": (c_vp, [c_vp]),
    "tlfea_elem_pack_nodes_xyz": (C.c_int, [c_vp, c_vp, C.c_int]),
    "tlfea_elem_set_external_force_device": (C.c_int, [c_vp, c_vp, c_vp, C.c_double, C.c_int]),
    "tlfea_newton_create": (C.c_int, [c_vp, C.c_int, C.POINTER(c_vp)]),
    "tlfea_newton_destroy": (C.c_int, [c_vp]),
    "tlfea_newton_set_params": (C.c_int, [c_vp, C.POINTER(NewtonParams)]),
    "tlfea_newton_setup": (C.c_int, [c_vp]),
    "tlfea_newton_analyze_sparsity": (C.c_int, [c_vp]),
    "tlfea_newton_set_fixed_pattern": (C.c_int, [c_vp, C.c_int]),
    "tlfea_newton_assemble": (C.c_int, [c_vp]),
    "tlfea_newton_assemble_force": (C.c_int, [c_vp]),
    "tlfea_newton_residual": (C.c_int, [c_vp, c_dp]),
    "tlfea_newton_assemble_host_begin": (C.c_int, [c_vp, c_dp, c_dp, c_dp, C.c_int]),
    "tlfea_newton_assemble_host_end": (C.c_int, [c_vp, c_dp, C.c_int, c_dp]),
    "tlfea_newton_set_host_result_buffer": (C.c_int, [c_vp, c_dp, C.c_int]),
    "tlfea_newton_solve_step": (C.c_int, [c_vp]),
    "tlfea_newton_last_stats": (C.c_int, [c_vp, c_ip, c_ip, c_dp, c_dp]),
    "tlfea_newton_set_pcg": (C.c_int, [c_vp, C.c_int, C.c_double]),
    "tlfea_newton_last_linear_solve": (C.c_int, [c_vp, c_dp, c_ip]),
    "tlfea_newton_nnz": (C.c_int64, [c_vp]),
    "tlfea_newton_n_dofs": (C.c_int, [c_vp]),
    "tlfea_newton_get_hessian_csr": (C.c_int, [c_vp, c_i64p, c_ip, c_dp]),
    "tlfea_newton_get_g": (C.c_int, [c_vp, c_dp, C.c_int]),
    "tlfea_newton_get_v": (C.c_int, [c_vp, c_dp, C.c_int]),
    "tlfea_newton_set_v": (C.c_int, [c_vp, c_dp, C.c_int]),
    "tlfea_newton_set_v_prev": (C.c_int, [c_vp, c_dp, C.c_int]),
    "tlfea_newton_devptr_v": (c_vp, [c_vp]),
    "tlfea_newton_devptr_hessian_values": (c_vp, [c_vp]),
    "tlfea_newton_devptr_g": (c_vp, [c_vp]),
    "tlfea_newton_get_diag_blocks": (C.c_int, [c_vp, c_dp, C.c_int64]),
    "tlfea_newton_set_owned_range": (C.c_int, [c_vp, C.c_int64, C.c_int64]),
    "tlfea_newton_krylov_op": (C.c_int, [c_vp, C.c_int]),
    "tlfea_newton_devptr_krylov": (C.c_void_p, [c_vp, C.c_int]),
    "tlfea_newton_spmv": (C.c_int, [c_vp, c_vp, c_vp]),
    "tlfea_newton_timing": (C.c_int, [c_vp, c_dp, c_ip, C.c_int]),
    "tlfea_newton_set_assembly": (C.c_int, [c_vp, C.c_int]),
    "tlfea_newton_set_patch_warps": (C.c_int, [c_vp, C.c_int]),
    "tlfea_newton_set_patch_chunks": (C.c_int, [c_vp, C.c_int]),
    "tlfea_newton_set_affine_compression": (C.c_int, [c_vp, C.c_int]),
    "tlfea_newton_patch_stats": (C.c_int, [c_vp, c_i64p]),
    "tlfea_newton_set_profile": (C.c_int, [c_vp, C.c_int]),
    "tlfea_newton_kernel_ms": (C.c_int, [c_vp, c_dp, c_dp, c_dp]),
    "tlfea_dist_unique_id": (C.c_int, [c_vp]),
    "tlfea_dist_init": (C.c_int, [c_vp, c_vp, C.c_int, C.c_int]),
    "tlfea_dist_set_halo": (C.c_int, [c_vp, C.c_int, c_ip, c_i64p, c_i64p]),
    "tlfea_dist_halo_sum": (C.c_int, [c_vp, c_vp, C.c_int]),
    "tlfea_dist_allreduce_sum": (C.c_int, [c_vp, c_vp, C.c_int]),
    "tlfea_dist_halo_bytes": (C.c_int64, [c_vp]),
    "tlfea_dist_finalize": (C.c_int, [c_vp]),
    "tlfea_dist_p2p_export": (C.c_int, [c_vp, C.c_int, c_vp]),
    "tlfea_dist_p2p_connect": (C.c_int, [c_vp, C.c_int, c_vp]),
    "tlfea_dot_noalloc": (C.c_int, [c_vp, c_vp, C.c_int64, c_vp, c_vp]),
    "tlfea_fp64_peak": (C.c_int, [C.c_int, c_dp]),
    "tlfea_debug_patch_plan_build": (C.c_int, [C.c_int, C.c_int, c_ip, c_dp, c_dp, c_dp, c_ip, c_ip, C.c_int, C.c_int,
                                               C.POINTER(c_vp)]),
    "tlfea_debug_patch_plan_sizes": (C.c_int, [c_vp, c_i64p]),
    "tlfea_debug_patch_plan_copy": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "tlfea_debug_patch_plan_free": (C.c_int, [c_vp]),
}

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(there is no CPU fallback for the tlfea hot path)"
    )

lib = C.CDLL(LIB_PATH)
for _name, (_res, _args) in _SIGS.items():
    _fn = getattr(lib, _name)
    _fn.restype = _res
    _fn.argtypes = _args


def check(status: int) -> None:
    if status != 0:
        raise TlfeaError(status, lib.tlfea_last_error().decode())


def exported_symbols():
    return sorted(_SIGS)
