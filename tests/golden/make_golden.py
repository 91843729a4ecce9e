"""Generates tests/golden/*.npz by running the REFERENCE's own numpy prototypes in place.

Run in the build container only (needs /root/reference):  python tests/golden/make_golden.py
The GPU box never runs this; it reads the committed .npz files.

Sources (relative to the reference root):
  test-scripts/T10-tets/f-form-T10-beam-newton.py         precompute :69-94, force :99-131,
                                                          mass :136-177, SVK tangent :221-282
  test-scripts/T10-tets/f-form-T10-beam-newton-damped.py  damped force :95-165, viscous tangent :316-383
  test-scripts/T10-tets/tet_mesh_reader.py                TetGen reader + mid-edge reorder
  data/meshes/T10/beam_3x2x1.1.{node,ele}                 36 tets, 105 nodes
  data/utest/mass_matrix_{2,3}_beam.csv                   ANCF3243 known-answer mass (utest_3243.cc:74-112,159-197)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_proto  # noqa: E402


def main():
    assert ref_proto.available(), "reference not mounted"
    ref = ref_proto.REF_ROOT
    proto = ref_proto.load("f-form-T10-beam-newton.py")
    damped = ref_proto.load("f-form-T10-beam-newton-damped.py")
    sys.path.insert(0, ref_proto.T10_DIR)
    import tet_mesh_reader as tmr

    X = tmr.read_node(os.path.join(ref, "data/meshes/T10/beam_3x2x1.1.node"))
    E = tmr.read_ele(os.path.join(ref, "data/meshes/T10/beam_3x2x1.1.ele"))
    nn = X.shape[0]
    Emod, nu, rho = 7e8, 0.33, 2700.0
    mu = Emod / (2 * (1 + nu))
    lam = Emod * nu / ((1 + nu) * (1 - 2 * nu))
    eta_d, lam_d = 1e5, 2e5

    rng = np.random.Generator(np.random.PCG64(1234))
    x = X + 0.02 * rng.uniform(-1, 1, size=X.shape)
    v = 0.1 * np.random.Generator(np.random.PCG64(1235)).uniform(-1, 1, size=X.shape)

    with ref_proto.quiet():
        pre = proto.tet10_precompute_reference_mesh(X, E)
        f_int = proto.tet10_internal_force_mesh(x, E, pre, lam, mu)
        K = proto.tet10_tangent_svk_mesh(X, x, E, pre, lam, mu)
        M = proto.tet10_consistent_mass_mesh(X, E, rho)
        pre_d = damped.tet10_precompute_reference_mesh(X, E)
        f_damped = damped.tet10_internal_force_mesh_damped(x, v, E, pre_d, lam, mu, eta_d, lam_d)
        C = damped.tet10_viscous_tangent_mesh(X, x, E, pre_d, eta_d, lam_d)
    grad_N = np.array([[q["grad_N"] for q in p] for p in pre])  # (ne,5,10,3)
    detJ = np.array([[q["detJ"] for q in p] for p in pre])
    w = np.array([q["w"] for q in pre[0]])
    np.savez_compressed(
        os.path.join(HERE, "t10_beam_3x2x1.npz"),
        X=X, E=E.astype(np.int32), x=x, v=v, lam=lam, mu=mu, rho=rho, Emod=Emod, nu=nu, eta_d=eta_d, lam_d=lam_d,
        grad_N=grad_N, detJ=detJ, w=w, f_int=f_int, K=K, M=np.asarray(M), f_damped=f_damped, C=C,
    )
    for nb in (2, 3):
        m = np.loadtxt(os.path.join(ref, f"data/utest/mass_matrix_{nb}_beam.csv"), delimiter=",")
        np.save(os.path.join(HERE, f"ancf3243_mass_{nb}_beam.npy"), m)
    print("golden vectors written")


if __name__ == "__main__":
    main()
