"""Generates tests/golden/vbd_t10_beam.npz from the REFERENCE's VBD prototype, imported in place:
test-scripts/vbd_proto/alm_vbd_t10_svk.py:175-223 (local_internal_force_and_Kii: per-node internal force and the
3x3 diagonal tangent block, the CPU statement of vbd_accumulate_residual_and_hessian_diag,
lib_src/elements/FEAT10DataFunc.cuh:295-395), on data/meshes/T10/beam_3x2x1 with the positions of
tests/golden/t10_beam_3x2x1.npz.  Run in the build container only:  python tests/golden/make_golden_vbd.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_proto  # noqa: E402


def main():
    assert ref_proto.available(), "reference not mounted"
    ref_proto.T10_DIR, keep = os.path.join(ref_proto.REF_ROOT, "test-scripts", "vbd_proto"), ref_proto.T10_DIR
    vbd = ref_proto.load("alm_vbd_t10_svk.py")
    ref_proto.T10_DIR = keep
    g = dict(np.load(os.path.join(HERE, "t10_beam_3x2x1.npz")))
    X, E, x = g["X"], g["E"], g["x"]
    Emod, nu = 7e8, 0.33
    mu = Emod / (2 * (1 + nu))
    lam = Emod * nu / ((1 + nu) * (1 - 2 * nu))
    pre = vbd.tet10_precompute_reference_mesh(X, E)
    inc = vbd.build_incidence(E, X.shape[0])
    f = np.zeros((X.shape[0], 3))
    K = np.zeros((X.shape[0], 3, 3))
    for i in range(X.shape[0]):
        f[i], K[i] = vbd.local_internal_force_and_Kii(i, x, E, pre, inc[i], lam, mu)
    Md = vbd.tet10_consistent_mass_mesh_diag_blocks(X, E, 2700.0)
    np.savez_compressed(os.path.join(HERE, "vbd_t10_beam.npz"), f_int_node=f, Kii=K, M_diag=np.asarray(Md))
    print("nodes", X.shape[0], "max|Kii|", np.abs(K).max(), "M_diag shape", np.asarray(Md).shape)


if __name__ == "__main__":
    main()
