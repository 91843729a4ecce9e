"""Caller-side data formats: .ancf3243mesh / .ancf3443mesh readers with pinned / welded
constraints -> LinearConstraintCSR (lib_utils/mesh_utils.cc:165-364,446-1027), VTK writer (FEAT10Data.cu:841-877).
Fixtures are written here in the reference's format; when the reference tree is mounted (build container only) its
own mesh files are parsed too and checked against their headers."""
import os

import numpy as np
import pytest

from tlfea import meshio

NET = """ancf3243_mesh 1
grid nx 1 ny 1 L 0.5 origin 0 0 0

nodes 4
# id family x0 x1 x2 x3 y0 y1 y2 y3 z0 z1 z2 z3
0 H 0 1 0 0 0 0 1 0 0 0 0 1
1 V 0 0 -1 0 0 1 0 0 0 0 0 1   # same point, rotated frame
2 H 0.5 1 0 0 0 0 1 0 0 0 0 1
3 V 0 0 -1 0 0.5 1 0 0 0 0 0 1

elements 2
# id family n0 n1
0 H 0 2
1 V 1 3

constraints 2
pinned 0 1
welded 0 1 0 -1 0 1 0 0 0 0 1
"""

SHELL = """ancf3443_mesh 1
meta two shells
nodes 6
0 R 0 1 0 0 0 0 1 0 0 0 0 1
1 R 2 1 0 0 0 0 1 0 0 0 0 1
2 R 2 1 0 0 1 0 1 0 0 0 0 1
3 R 0 1 0 0 1 0 1 0 0 0 0 1
4 S 4 1 0 0 0 0 1 0 0 0 0 1
5 S 4 1 0 0 1 0 1 0 0 0 0 1
elements 2
# id family L W H n0 n1 n2 n3
1 S 2 1 0.5 1 4 5 2
0 R 2 1 1 0 1 2 3
"""


def test_builder_drops_zeros_and_keeps_order():
    b = meshio.LinearConstraintBuilder(12)
    assert b.AddRow([(3, 1.0), (1, 0.0), (0, -2.0)], 0.5) == 0
    assert b.AddFixedDof(7, 1.25) == 1
    c = b.ToCSR()
    assert c.offsets.tolist() == [0, 2, 3] and c.columns.tolist() == [3, 0, 7]
    assert c.values.tolist() == [1.0, -2.0, 1.0] and c.rhs.tolist() == [0.5, 1.25]
    assert c.NumRows() == 2 and c.NumNonZeros() == 3 and not c.Empty()
    with pytest.raises(IndexError):
        b.AddRow([(12, 1.0)])
    with pytest.raises(ValueError):
        b.AddRow([])


def test_read_ancf3243mesh(tmp_path):
    p = tmp_path / "net.ancf3243mesh"
    p.write_text(NET)
    m = meshio.ReadANCF3243MeshFromFile(str(p))
    assert (m.version, m.n_nodes, m.n_elements) == (1, 4, 2)
    assert m.grid["nx"] == 1 and m.grid["L"] == 0.5
    assert m.node_family == ["H", "V", "H", "V"] and m.element_family == ["H", "V"]
    assert m.element_connectivity.tolist() == [[0, 2], [1, 3]]
    assert m.x12[4:8].tolist() == [0, 0, -1, 0] and m.y12[12:16].tolist() == [0.5, 1, 0, 0]
    c = m.constraints
    # pinned: 3 rows; welded: 3 position rows + 3 slots x 3 rows
    assert c.NumRows() == 3 + 12
    A = np.zeros((c.NumRows(), 48))
    for r in range(c.NumRows()):
        A[r, c.columns[c.offsets[r]:c.offsets[r + 1]]] = c.values[c.offsets[r]:c.offsets[r + 1]]
    q = np.stack([m.x12, m.y12, m.z12], axis=1).reshape(-1)  # DOF = 3 * coefficient + component
    # the file's reference configuration satisfies its own constraints: r_b = r_a, grad_b = Q grad_a
    assert np.abs(A @ q - c.rhs).max() == 0.0
    # welded gradient row: r_u(b).x - (Q r_u(a)).x with Q = [[0,-1,0],[1,0,0],[0,0,1]] -> +1 at b, +1 at a.y
    row = 3 + 3  # first gradient row
    cols = c.columns[c.offsets[row]:c.offsets[row + 1]].tolist()
    vals = c.values[c.offsets[row]:c.offsets[row + 1]].tolist()
    assert cols == [(1 * 4 + 1) * 3 + 0, (0 * 4 + 1) * 3 + 1] and vals == [1.0, 1.0]


def test_read_ancf3443mesh_and_errors(tmp_path):
    p = tmp_path / "two.ancf3443mesh"
    p.write_text(SHELL)
    m = meshio.ReadANCF3443MeshFromFile(str(p))
    assert (m.n_nodes, m.n_elements) == (6, 2)
    assert m.element_connectivity.tolist() == [[0, 1, 2, 3], [1, 4, 5, 2]]  # ids may come in any order
    assert m.element_H.tolist() == [1.0, 0.5] and m.element_family == ["R", "S"]
    assert m.constraints.Empty()
    bad = tmp_path / "bad.ancf3443mesh"
    bad.write_text(SHELL.replace("1 S 2 1 0.5 1 4 5 2", "1 S 2 1 0.5 1 4 9 2"))
    with pytest.raises(meshio.MeshFormatError, match="out of range"):
        meshio.ReadANCF3443MeshFromFile(str(bad))
    bad.write_text(NET)
    with pytest.raises(meshio.MeshFormatError, match="expected header"):
        meshio.ReadANCF3443MeshFromFile(str(bad))
    bad.write_text(SHELL + "constraints 1\nglued 0 1\n")
    with pytest.raises(meshio.MeshFormatError, match="unknown constraint type"):
        meshio.ReadANCF3443MeshFromFile(str(bad))
    with pytest.raises(meshio.MeshFormatError, match="failed to open"):
        meshio.ReadANCF3443MeshFromFile(str(tmp_path / "missing"))


def test_vtk_writer(tmp_path):
    from tlfea import meshgen

    X, E = meshgen.kuhn_block(1)
    f = tmp_path / "o.vtk"
    meshio.write_vtk_t10(str(f), X[:, 0], X[:, 1], X[:, 2], E)
    lines = f.read_text().split("\n")
    assert lines[:4] == ["# vtk DataFile Version 3.0", "T10 mesh output", "ASCII", "DATASET UNSTRUCTURED_GRID"]
    assert lines[4] == f"POINTS {X.shape[0]} float"
    assert lines[5 + X.shape[0]] == f"CELLS {E.shape[0]} {E.shape[0] * 11}"
    assert lines[6 + X.shape[0]].split() == ["10"] + [str(v) for v in E[0]]
    assert lines[-2] == "24" and lines.count("24") == E.shape[0]


REF = "/root/reference/data/meshes"


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not mounted (GPU box)")
def test_reference_mesh_files_parse():
    m = meshio.ReadANCF3243MeshFromFile(os.path.join(REF, "ANCF3243/net_welded_nx20_ny20_L0.5.ancf3243mesh"))
    assert (m.n_nodes, m.n_elements, m.constraints.NumRows()) == (882, 840, 441 * 12)
    q = np.stack([m.x12, m.y12, m.z12], axis=1).reshape(-1)
    c = m.constraints
    r = np.add.reduceat(c.values * q[c.columns], c.offsets[:-1]) - c.rhs
    assert np.abs(r).max() < 1e-12
    m = meshio.ReadANCF3243MeshFromFile(os.path.join(REF, "ANCF3243/net_pinned_nx20_ny20_L0.5.ancf3243mesh"))
    assert m.constraints.NumRows() == 441 * 3
    m = meshio.ReadANCF3443MeshFromFile(os.path.join(REF, "ANCF3443/airless_tire.ancf3443mesh"))
    assert (m.n_nodes, m.n_elements, m.constraints.NumRows()) == (1120, 720, 160 * 12)
    q = np.stack([m.x12, m.y12, m.z12], axis=1).reshape(-1)
    c = m.constraints
    r = np.add.reduceat(c.values * q[c.columns], c.offsets[:-1]) - c.rhs
    assert np.abs(r).max() < 1e-9


def _net_text(nx, ny, L):
    """A welded net in the reference's .ancf3243mesh layout (as written by test-scripts/3243-beam/net_mesh_gen.py):
    horizontal beams with frame (e_x, e_y, e_z), vertical beams with frame (e_y, -e_x, e_z), one H and one V node
    per grid point welded with Q = Rz(90 deg)."""
    nodes, elems, cons = [], [], []
    hid, vid = {}, {}
    for j in range(ny + 1):
        for i in range(nx + 1):
            hid[i, j] = len(nodes); nodes.append(f"{len(nodes)} H {i * L} 1 0 0 {j * L} 0 1 0 0 0 0 1")
            vid[i, j] = len(nodes); nodes.append(f"{len(nodes)} V {i * L} 0 -1 0 {j * L} 1 0 0 0 0 0 1")
            cons.append(f"welded {hid[i, j]} {vid[i, j]} 0 -1 0 1 0 0 0 0 1")
    for j in range(ny + 1):
        for i in range(nx):
            elems.append(f"{len(elems)} H {hid[i, j]} {hid[i + 1, j]}")
    for i in range(nx + 1):
        for j in range(ny):
            elems.append(f"{len(elems)} V {vid[i, j]} {vid[i, j + 1]}")
    return (f"ancf3243_mesh 1\ngrid nx {nx} ny {ny} L {L} origin 0 0 0\n\nnodes {len(nodes)}\n" + "\n".join(nodes)
            + f"\n\nelements {len(elems)}\n" + "\n".join(elems) + f"\n\nconstraints {len(cons)}\n" + "\n".join(cons) + "\n")


@pytest.mark.gpu
def test_welded_net_from_file_steps_on_the_gpu(tmp_path):
    """File -> reader -> SetLinearConstraintsCSR -> constraint-aware pattern -> ALM/Newton steps: the welds and the
    clamped corners hold while the loaded centre sags (the lib_bin net demos' flow)."""
    import tlfea

    L, W, H = 0.5, 0.05, 0.05
    p = tmp_path / "net.ancf3243mesh"
    p.write_text(_net_text(2, 2, L))
    m = meshio.ReadANCF3243MeshFromFile(str(p))
    b = meshio.LinearConstraintBuilder(3 * 4 * m.n_nodes, m.constraints)
    corners = [n for n in range(m.n_nodes) if m.node_family[n] == "H"
               and m.x12[4 * n] in (0.0, 2 * L) and m.y12[4 * n] in (0.0, 2 * L)]
    assert len(corners) == 4
    for n in corners:
        for slot in range(4):
            meshio.append_fixed_coefficient(b, 4 * n + slot, m.x12, m.y12, m.z12)
    c = b.ToCSR()
    d = tlfea.GPU_ANCF3243_Data(m.n_nodes, m.n_elements)
    d.Initialize()
    d.Setup(L, W, H, m.x12, m.y12, m.z12, m.element_connectivity)
    d.SetDensity(2700.0)
    d.SetSVK(7e8, 0.33)
    centre = next(n for n in range(m.n_nodes) if m.node_family[n] == "H" and m.x12[4 * n] == L and m.y12[4 * n] == L)
    f_ext = np.zeros(3 * 4 * m.n_nodes)
    f_ext[3 * (4 * centre) + 2] = -20.0
    d.SetExternalForce(f_ext)
    d.SetLinearConstraintsCSR(c.offsets, c.columns, c.values, c.rhs)
    d.CalcDsDuPre()
    d.CalcMassMatrix()
    d.CalcConstraintData()
    s = tlfea.SyncedNewtonSolver(d, c.NumRows())
    s.SetParameters(tlfea.SyncedNewtonParams(1e-6, 0.0, 1e-8, 1e10, 3, 8, 1e-3))
    s.Setup()
    s.AnalyzeHessianSparsity()
    s.SetFixedSparsityPattern(True)
    s.SetPCG(20000, 1e-12)
    for _ in range(3):
        s.Solve()
    gx, gy, gz = d.RetrievePositionToCPU()
    q = np.stack([gx, gy, gz], axis=1).reshape(-1)
    assert np.isfinite(q).all()
    r = np.add.reduceat(c.values * q[c.columns], c.offsets[:-1]) - c.rhs
    assert np.abs(r).max() < 1e-6      # welds and clamps hold (ALM, rho = 1e10, 3 outer iterations)
    assert gz[4 * centre] < -1e-9      # the loaded crossing moved down ...
    v_centre = centre + 1              # ... and took its welded V node with it
    assert abs(gz[4 * v_centre] - gz[4 * centre]) < 1e-6


@pytest.mark.gpu
def test_shell_file_with_per_element_thickness_on_the_gpu(tmp_path):
    """.ancf3443mesh -> Setup with the per-element L, W, H arrays of the file (ANCF3443Data.cuh:534-544):
    quadrature volume and total mass follow each element's own dimensions."""
    import tlfea

    p = tmp_path / "two.ancf3443mesh"
    p.write_text(SHELL)
    m = meshio.ReadANCF3443MeshFromFile(str(p))
    d = tlfea.GPU_ANCF3443_Data(m.n_nodes, m.n_elements)
    d.Initialize()
    d.Setup(m.element_L, m.element_W, m.element_H, m.x12, m.y12, m.z12, m.element_connectivity)
    d.SetDensity(1000.0)
    d.SetSVK(7e8, 0.33)
    d.CalcDsDuPre()
    d.CalcMassMatrix()
    vol = float((m.element_L * m.element_W * m.element_H).sum())  # 2*1*1 + 2*1*0.5
    off, cols, vals = d.RetrieveMassCSRToCPU()
    rows = np.repeat(np.arange(off.shape[0] - 1), np.diff(off))
    pos = (rows % 4 == 0) & (cols % 4 == 0)  # position-position entries: partition of unity
    assert abs(vals[pos].sum() - 1000.0 * vol) < 1e-10 * 1000.0 * vol
