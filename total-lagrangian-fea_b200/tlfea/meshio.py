"""Data formats on the caller's side of the hot path: the reference's ANCF mesh files with
their linear constraints, and the VTK dump of a FEAT10 mesh.  New code written from the file formats; behaviour
follows the reference readers record by record:

  LinearConstraintBuilder / LinearConstraintCSR     lib_utils/mesh_utils.cc:165-245, mesh_utils.h:100-159
  pinned / welded rows for ANCF nodes               lib_utils/mesh_utils.cc:251-298 (3243), :316-364 (3443)
  ReadANCF3243MeshFromFile (.ancf3243mesh)          lib_utils/mesh_utils.cc:446-731
  ReadANCF3443MeshFromFile (.ancf3443mesh)          lib_utils/mesh_utils.cc:738-1027
  GPU_FEAT10_Data::WriteOutputVTK                   lib_src/elements/FEAT10Data.cu:841-877

The CSR the builder produces goes straight into SetLinearConstraintsCSR (ANCF3243Data.cuh:810-940 ->
tlfea_elem_set_linear_constraints_csr)."""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np


class MeshFormatError(ValueError):
    """The reference readers return false and fill an error string; here the message is raised."""


@dataclass
class LinearConstraintCSR:
    offsets: np.ndarray
    columns: np.ndarray
    values: np.ndarray
    rhs: np.ndarray

    def NumRows(self) -> int:
        return int(self.rhs.shape[0])

    def NumNonZeros(self) -> int:
        return int(self.columns.shape[0])

    def Empty(self) -> bool:
        return self.NumRows() == 0


class LinearConstraintBuilder:
    """Row-by-row builder of c(q) = A q - rhs in scalar-DOF space (DOF = 3 * coefficient + component).
    Zero-valued entries are dropped, rows keep their insertion order, columns keep the order given."""

    def __init__(self, n_dofs: int, initial: "LinearConstraintCSR | None" = None):
        if n_dofs < 0:
            raise ValueError("LinearConstraintBuilder: n_dofs must be >= 0")
        self.n_dofs = int(n_dofs)
        self._off = [0]
        self._col: list[int] = []
        self._val: list[float] = []
        self._rhs: list[float] = []
        if initial is not None:  # continue an existing system (mesh_utils.cc:176-207)
            if len(initial.offsets) != initial.NumRows() + 1 or len(initial.values) != initial.NumNonZeros():
                raise ValueError("LinearConstraintBuilder: initial CSR size mismatch")
            if initial.offsets[0] != 0 or initial.offsets[-1] != initial.NumNonZeros():
                raise ValueError("LinearConstraintBuilder: initial CSR offsets invalid")
            self._off = [int(v) for v in initial.offsets]
            self._col = [int(v) for v in initial.columns]
            self._val = [float(v) for v in initial.values]
            self._rhs = [float(v) for v in initial.rhs]

    def AddRow(self, entries, rhs: float = 0.0) -> int:
        entries = list(entries)
        if not entries:
            raise ValueError("LinearConstraintBuilder::AddRow: empty row")
        for col, val in entries:
            if col < 0 or col >= self.n_dofs:
                raise IndexError("LinearConstraintBuilder::AddRow: col out of range")
            if val == 0.0:
                continue
            self._col.append(int(col))
            self._val.append(float(val))
        self._rhs.append(float(rhs))
        self._off.append(len(self._col))
        return len(self._rhs) - 1

    def AddFixedDof(self, col: int, rhs: float) -> int:
        return self.AddRow([(col, 1.0)], rhs)

    def ToCSR(self) -> LinearConstraintCSR:
        return LinearConstraintCSR(np.array(self._off, dtype=np.int32), np.array(self._col, dtype=np.int32),
                                   np.array(self._val, dtype=np.float64), np.array(self._rhs, dtype=np.float64))


def _dof(node: int, slot: int, comp: int) -> int:
    return (node * 4 + slot) * 3 + comp


def append_vector_equality(builder: LinearConstraintBuilder, node_a: int, node_b: int, slot: int) -> None:
    """r(b, slot) - r(a, slot) = 0, three rows (slot 0 = position, 1..3 = r_u, r_v, r_w)."""
    if not 0 <= slot <= 3:
        raise IndexError("coef_slot out of range")
    for c in range(3):
        builder.AddRow([(_dof(node_b, slot, c), 1.0), (_dof(node_a, slot, c), -1.0)], 0.0)


def append_vector_welded(builder: LinearConstraintBuilder, node_a: int, node_b: int, slot: int, Q) -> None:
    """r(b, slot) - Q r(a, slot) = 0, three rows; zero entries of Q are left out of the row."""
    if not 0 <= slot <= 3:
        raise IndexError("coef_slot out of range")
    Q = np.asarray(Q, dtype=np.float64).reshape(3, 3)
    for row in range(3):
        ent = [(_dof(node_b, slot, row), 1.0)]
        for k in range(3):
            w = -Q[row, k]
            if w != 0.0:
                ent.append((_dof(node_a, slot, k), w))
        builder.AddRow(ent, 0.0)


def append_fixed_coefficient(builder: LinearConstraintBuilder, coef: int, x_ref, y_ref, z_ref) -> None:
    """Pins the three components of one coefficient at their reference values (mesh_utils.cc:296-313)."""
    if coef < 0 or coef >= len(x_ref):
        raise IndexError("coef_index out of range")
    builder.AddFixedDof(coef * 3 + 0, float(x_ref[coef]))
    builder.AddFixedDof(coef * 3 + 1, float(y_ref[coef]))
    builder.AddFixedDof(coef * 3 + 2, float(z_ref[coef]))


@dataclass
class ANCFMesh:
    """ANCF3243Mesh / ANCF3443Mesh (mesh_utils.h:161-205): x12, y12, z12 are [4 * n_nodes] with slot 0 the
    position and slots 1..3 the gradients."""
    kind: int
    version: int = 0
    n_nodes: int = 0
    n_elements: int = 0
    node_family: list = field(default_factory=list)
    x12: np.ndarray = None
    y12: np.ndarray = None
    z12: np.ndarray = None
    element_family: list = field(default_factory=list)
    element_connectivity: np.ndarray = None
    element_L: np.ndarray = None   # 3443 only
    element_W: np.ndarray = None
    element_H: np.ndarray = None
    grid: dict = None              # 3243: {"nx", "ny", "L", "origin"} when the file has a grid line
    meta: list = field(default_factory=list)  # 3443: 'tire' / 'meta' lines, tokenised
    constraints: LinearConstraintCSR = None


def _records(path: str):
    """Lines with '#' comments stripped, blank lines skipped (ReadNextRecord, mesh_utils.cc:392-403)."""
    with open(path) as f:
        for raw in f:
            ln = raw.split("#", 1)[0].strip()
            if ln:
                yield ln.split()


def _int(tok: str, what: str) -> int:
    try:
        return int(tok)
    except ValueError:
        raise MeshFormatError(f"{what}: '{tok}' is not an integer") from None


def _float(tok: str, what: str) -> float:
    try:
        return float(tok)
    except ValueError:
        raise MeshFormatError(f"{what}: '{tok}' is not a number") from None


def _read_ancf(path: str, kind: int) -> ANCFMesh:
    tag = f"ancf{kind}_mesh"
    who = f"ReadANCF{kind}MeshFromFile"
    try:
        rec = _records(path)
        first = next(rec, None)
    except OSError:
        raise MeshFormatError(f"{who}: failed to open {path}") from None
    if first is None:
        raise MeshFormatError(f"{who}: empty file")
    if len(first) != 2 or first[0] != tag:
        raise MeshFormatError(f"{who}: expected header '{tag} <version>'")
    m = ANCFMesh(kind=kind, version=_int(first[1], who))
    t = next(rec, None)
    # optional metadata lines before 'nodes'
    while t is not None and t[0] != "nodes":
        if kind == 3243 and t[0] == "grid":
            if len(t) != 11 or t[1] != "nx" or t[3] != "ny" or t[5] != "L" or t[7] != "origin":
                raise MeshFormatError(f"{who}: invalid grid line")
            m.grid = {"nx": _int(t[2], who), "ny": _int(t[4], who), "L": _float(t[6], who),
                      "origin": np.array([_float(v, who) for v in t[8:11]])}
        elif kind == 3443 and t[0] in ("tire", "meta"):
            m.meta.append(t)
        else:
            raise MeshFormatError(f"{who}: expected 'nodes <N>'")
        t = next(rec, None)
    if t is None or len(t) != 2:
        raise MeshFormatError(f"{who}: expected 'nodes <N>'")
    nn = _int(t[1], who)
    if nn <= 0:
        raise MeshFormatError(f"{who}: invalid node count")
    m.n_nodes = nn
    m.node_family = [""] * nn
    m.x12 = np.zeros(4 * nn); m.y12 = np.zeros(4 * nn); m.z12 = np.zeros(4 * nn)
    seen = np.zeros(nn, dtype=bool)
    for _ in range(nn):
        t = next(rec, None)
        if t is None:
            raise MeshFormatError(f"{who}: unexpected EOF in nodes")
        if len(t) != 14:
            raise MeshFormatError(f"{who}: node line expects 14 fields")
        i = _int(t[0], who)
        if i < 0 or i >= nn or seen[i]:
            raise MeshFormatError(f"{who}: node id out of range or duplicated")
        seen[i] = True
        m.node_family[i] = t[1]
        vals = [_float(v, who) for v in t[2:14]]
        m.x12[4 * i:4 * i + 4] = vals[0:4]
        m.y12[4 * i:4 * i + 4] = vals[4:8]
        m.z12[4 * i:4 * i + 4] = vals[8:12]
    t = next(rec, None)
    if t is None or len(t) != 2 or t[0] != "elements":
        raise MeshFormatError(f"{who}: expected 'elements <M>'")
    ne = _int(t[1], who)
    if ne <= 0:
        raise MeshFormatError(f"{who}: invalid element count")
    m.n_elements = ne
    npe = 2 if kind == 3243 else 4
    nfld = 2 + npe if kind == 3243 else 5 + npe
    m.element_family = [""] * ne
    m.element_connectivity = np.zeros((ne, npe), dtype=np.int32)
    if kind == 3443:
        m.element_L = np.zeros(ne); m.element_W = np.zeros(ne); m.element_H = np.zeros(ne)
    seen = np.zeros(ne, dtype=bool)
    for _ in range(ne):
        t = next(rec, None)
        if t is None:
            raise MeshFormatError(f"{who}: unexpected EOF in elements")
        if len(t) != nfld:
            raise MeshFormatError(f"{who}: element line expects {nfld} fields")
        e = _int(t[0], who)
        if e < 0 or e >= ne or seen[e]:
            raise MeshFormatError(f"{who}: element id out of range or duplicated")
        seen[e] = True
        m.element_family[e] = t[1]
        k = 2
        if kind == 3443:
            m.element_L[e], m.element_W[e], m.element_H[e] = (_float(v, who) for v in t[2:5])
            k = 5
        nodes = [_int(v, who) for v in t[k:k + npe]]
        if min(nodes) < 0 or max(nodes) >= nn:
            raise MeshFormatError(f"{who}: element node id out of range")
        m.element_connectivity[e] = nodes
    builder = LinearConstraintBuilder(3 * 4 * nn)
    t = next(rec, None)
    if t is not None:
        if len(t) != 2 or t[0] != "constraints":
            raise MeshFormatError(f"{who}: invalid constraints header")
        nc = _int(t[1], who)
        if nc < 0:
            raise MeshFormatError(f"{who}: invalid constraints header")
        for _ in range(nc):
            t = next(rec, None)
            if t is None:
                raise MeshFormatError(f"{who}: unexpected EOF in constraints")
            if t[0] == "pinned":
                if len(t) != 3:
                    raise MeshFormatError(f"{who}: pinned expects 'pinned a b'")
                a, b = _int(t[1], who), _int(t[2], who)
                if not (0 <= a < nn and 0 <= b < nn):
                    raise MeshFormatError(f"{who}: pinned node id out of range")
                append_vector_equality(builder, a, b, 0)
            elif t[0] == "welded":
                if len(t) != 12:
                    raise MeshFormatError(f"{who}: welded expects 'welded a b q00..q22'")
                a, b = _int(t[1], who), _int(t[2], who)
                if not (0 <= a < nn and 0 <= b < nn):
                    raise MeshFormatError(f"{who}: welded node id out of range")
                Q = np.array([_float(v, who) for v in t[3:12]]).reshape(3, 3)
                append_vector_equality(builder, a, b, 0)  # position continuity
                for slot in (1, 2, 3):                     # gradient continuity through Q
                    append_vector_welded(builder, a, b, slot, Q)
            else:
                raise MeshFormatError(f"{who}: unknown constraint type '{t[0]}'")
    m.constraints = builder.ToCSR()
    return m


def ReadANCF3243MeshFromFile(path: str) -> ANCFMesh:
    return _read_ancf(path, 3243)


def ReadANCF3443MeshFromFile(path: str) -> ANCFMesh:
    return _read_ancf(path, 3443)


def write_vtk_t10(filename: str, x, y, z, connectivity) -> None:
    """WriteOutputVTK: legacy ASCII unstructured grid, one VTK_QUADRATIC_TETRA (type 24) per element, node order
    as stored."""
    x, y, z = (np.asarray(a, dtype=np.float64) for a in (x, y, z))
    conn = np.asarray(connectivity)
    with open(filename, "w") as out:
        out.write("# vtk DataFile Version 3.0\nT10 mesh output\nASCII\nDATASET UNSTRUCTURED_GRID\n")
        out.write(f"POINTS {x.shape[0]} float\n")
        for i in range(x.shape[0]):
            out.write(f"{x[i]:.6g} {y[i]:.6g} {z[i]:.6g}\n")
        out.write(f"CELLS {conn.shape[0]} {conn.shape[0] * 11}\n")
        for row in conn:
            out.write("10 " + " ".join(str(int(v)) for v in row) + " \n")
        out.write(f"CELL_TYPES {conn.shape[0]}\n")
        out.write("24\n" * conn.shape[0])
