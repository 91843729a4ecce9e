"""Generates tests/golden/ancf{3243,3443}_proto.npz from the REFERENCE's own numpy prototypes.

Run in the build container only (needs /root/reference):  python tests/golden/make_golden_ancf.py
The GPU box never runs this; it reads the committed .npz files.

The ANCF prototypes run a whole simulation at module level, so they cannot be imported.  The script
executes the reference file IN PLACE up to its time loop (`for step in range(Nt):` -- everything before
it is set-up: quadrature tables, B^-1, the consistent mass matrix `m`, the precomputed shape-function
gradients) and then the nested `compute_internal_force` definition taken from the loop body, all in one
namespace; nothing of the reference is copied into this repository.

Sources (relative to the reference root):
  test-scripts/3243-beam/f-form-3243-nesterov.py   b_vec :135, B12_matrix :154, mass loop :340-372,
                                                  ds_du_pre / detJ_pre :418-443, compute_internal_force :488-541 (and its derivative, see central())
  test-scripts/3443-shell/f-form-3443-nesterov.py  b_vec :184, B12_matrix :299, mass loop :459-497,
                                                  ds_du_pre / detJ_pre :507-531, compute_internal_force :575-638
Both use SVK with E = 7e8, nu = 0.33, rho0 = 2700 and elements of L = 2, W = H = 1 (3 beams in a chain;
2 shells sharing an edge with connectivity [[0,1,2,3],[1,4,5,2]]).
"""
import contextlib
import io
import os
import sys
import textwrap
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("TLFEA_REFERENCE_ROOT", "/root/reference")


def _stubs():
    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.animation", "pandas"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.modules["matplotlib"].animation = sys.modules["matplotlib.animation"]

    class _DF:  # the 3443 prototype pretty-prints its mass matrix through pandas
        def __init__(self, a):
            self.a = a

        def to_string(self, **kw):
            return ""

    sys.modules["pandas"].DataFrame = _DF


def load_prototype(rel_path: str) -> dict:
    """Namespace of the prototype after its set-up section, with compute_internal_force defined."""
    _stubs()
    src = open(os.path.join(REF, rel_path)).read().split("\n")
    cut = next(i for i, ln in enumerate(src) if ln.startswith("for step in range(Nt):"))
    head = "\n".join(src[:cut])
    body = src[cut + 1:]
    first = next(i for i, ln in enumerate(body) if ln.startswith("    def compute_deformation_gradient"))
    last = next(i for i, ln in enumerate(body) if ln.startswith("    def alm_nesterov_step"))
    defs = textwrap.dedent("\n".join(body[first:last]))
    ns: dict = {"__name__": "ref_ancf_proto"}
    with contextlib.redirect_stdout(io.StringIO()):
        exec(compile(head, rel_path, "exec"), ns)
        exec(compile(defs, rel_path + ":loop-body", "exec"), ns)
    return ns


def one(kind: int, rel_path: str, seed: int):
    ns = load_prototype(rel_path)
    xr, yr, zr = (np.array(ns[k], dtype=np.float64) for k in ("x12", "y12", "z12"))
    nc = xr.shape[0]
    rng = np.random.Generator(np.random.PCG64(seed))
    amp = np.tile([0.05, 0.03, 0.03, 0.03], nc // 4)
    x = xr + amp * rng.uniform(-1, 1, nc)
    y = yr + amp * rng.uniform(-1, 1, nc)
    z = zr + amp * rng.uniform(-1, 1, nc)
    with contextlib.redirect_stdout(io.StringIO()):
        f_int = ns["compute_internal_force"](x, y, z)
        f_rest = ns["compute_internal_force"](xr, yr, zr)
    assert np.abs(f_rest).max() < 1e-6 * np.abs(f_int).max()
    # tangent of the prototype's force: SVK f_int is a cubic polynomial of the coefficients, so the Richardson
    # combination (4 D(h) - D(2h)) / 3 of central differences is exact up to rounding (~1e-14 relative here)
    def central(hh):
        K = np.zeros((3 * nc, 3 * nc))
        for c in range(nc):
            for k, arr in enumerate((x, y, z)):
                save = arr[c]
                arr[c] = save + hh
                fp = ns["compute_internal_force"](x, y, z)
                arr[c] = save - hh
                fm = ns["compute_internal_force"](x, y, z)
                arr[c] = save
                K[:, 3 * c + k] = (fp - fm) / (2 * hh)
        return K

    with contextlib.redirect_stdout(io.StringIO()):
        hh = 2.0 ** -7
        K_t = (4.0 * central(hh) - central(2 * hh)) / 3.0
    assert np.abs(K_t - K_t.T).max() < 1e-11 * np.abs(K_t).max()
    if kind == 3243:
        conn = np.array([[e, e + 1] for e in range(int(ns["n_beam"]))], dtype=np.int32)
    else:
        conn = np.array(ns["element_connectivity"], dtype=np.int32)
    out = dict(x_ref=xr, y_ref=yr, z_ref=zr, x=x, y=y, z=z, conn=conn, f_int=f_int, K_t=K_t, M=np.array(ns["m"]),
               B_inv=np.array(ns["B_inv"]), L=float(ns["L"]), W=float(ns["W"]), H=float(ns["H"]),
               Emod=float(ns["E"]), nu=float(ns["nu"]), rho=float(ns["rho0"]))
    np.savez_compressed(os.path.join(HERE, f"ancf{kind}_proto.npz"), **out)
    print(f"ancf{kind}_proto.npz: {conn.shape[0]} elements, {nc} coefficients, |f|max {np.abs(f_int).max():.3e}")


def main():
    assert os.path.isdir(REF), "reference not mounted"
    one(3243, "test-scripts/3243-beam/f-form-3243-nesterov.py", 3243)
    one(3443, "test-scripts/3443-shell/f-form-3443-nesterov.py", 3443)


if __name__ == "__main__":
    main()
