"""Synthetic structured meshes for the benchmarks (new code: the reference ships only file readers).

kuhn_block(n): n^3 hexahedral cells on the unit cube, each split into 6 Kuhn tets (vertex paths along
the axis permutations), T10 mid-edge nodes on the doubled lattice (2n+1)^3, node id (i*N+j)*N+k,
local node order [v0..v3,(01),(12),(02),(03),(13),(23)] (cpu_utils.cc:609-611), detJ > 0.
"""
from __future__ import annotations

import itertools

import numpy as np

_EDGES = [(0, 1), (1, 2), (0, 2), (0, 3), (1, 3), (2, 3)]


def kuhn_block(n: int, dtype=np.int32):
    """Returns X (Nn,3) float64 and elements (6 n^3, 10) int32."""
    N = 2 * n + 1
    lin = np.arange(N, dtype=np.float64) / (2.0 * n)
    X = np.empty((N * N * N, 3))
    X[:, 0] = np.repeat(lin, N * N)
    X[:, 1] = np.tile(np.repeat(lin, N), N)
    X[:, 2] = np.tile(lin, N * N)

    ci, cj, ck = np.meshgrid(np.arange(n), np.arange(n), np.arange(n), indexing="ij")
    base = np.stack([2 * ci.ravel(), 2 * cj.ravel(), 2 * ck.ravel()], axis=1).astype(np.int64)  # (cells,3)
    n_cells = base.shape[0]
    elements = np.empty((n_cells, 6, 10), dtype=np.int64)
    for t, perm in enumerate(itertools.permutations(range(3))):
        # vertex lattice offsets along the path 0 -> e_p0 -> e_p0+e_p1 -> (1,1,1)
        v = np.zeros((4, 3), dtype=np.int64)
        for s in range(3):
            v[s + 1] = v[s]
            v[s + 1, perm[s]] += 2
        # odd permutations are mirror images: swap two vertices so detJ > 0
        sign = np.linalg.det((v[1:] - v[0]).astype(np.float64))
        if sign < 0:
            v[[2, 3]] = v[[3, 2]]
        loc = [v[0], v[1], v[2], v[3]] + [(v[a] + v[b]) // 2 for a, b in _EDGES]
        for a, off in enumerate(loc):
            p = base + off[None, :]
            elements[:, t, a] = (p[:, 0] * N + p[:, 1]) * N + p[:, 2]
    return X, elements.reshape(n_cells * 6, 10).astype(dtype)


def perturb(X: np.ndarray, amplitude: float, seed: int) -> np.ndarray:
    """x = X + amplitude * U(-1,1) per component, PCG64(seed) (SURVEY.md 8d synthetic inputs)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    return X + amplitude * rng.uniform(-1.0, 1.0, size=X.shape)


def slab_partition(n: int, n_parts: int):
    """Element partition of kuhn_block(n) into n_parts slabs of whole cell layers along x.
    Returns a list of element-index arrays (ascending)."""
    bounds = [(n * p) // n_parts for p in range(n_parts + 1)]
    per_layer = 6 * n * n
    return [np.arange(bounds[p] * per_layer, bounds[p + 1] * per_layer, dtype=np.int64) for p in range(n_parts)]


def read_tetgen(node_file: str, ele_file: str):
    """TetGen .node/.ele reader with 0/1-based auto-detect and the mid-edge reorder
    [0,1,2,3,6,7,9,5,8,4] (cpu_utils.cc:607-754)."""
    with open(node_file) as f:
        head = f.readline().split()
        nn = int(head[0])
        rows = [ln.split() for ln in f if ln.strip() and not ln.lstrip().startswith("#")][:nn]
    ids = np.array([int(r[0]) for r in rows])
    base = 0 if ids.min() == 0 else 1
    X = np.zeros((nn, 3))
    X[ids - base] = np.array([[float(r[1]), float(r[2]), float(r[3])] for r in rows])
    with open(ele_file) as f:
        head = f.readline().split()
        ne, k = int(head[0]), int(head[1])
        if k != 10:
            raise ValueError("only T10 elements are supported")
        rows = [ln.split() for ln in f if ln.strip() and not ln.lstrip().startswith("#")][:ne]
    eid = np.array([int(r[0]) for r in rows])
    raw = np.array([[int(v) for v in r[1:11]] for r in rows])
    ebase = 0 if eid.min() == 0 else 1
    nbase = 0 if raw.min() == 0 else 1
    E = np.zeros((ne, 10), dtype=np.int32)
    E[eid - ebase] = (raw - nbase)[:, [0, 1, 2, 3, 6, 7, 9, 5, 8, 4]]
    return X, E


def ancf_plate(nx: int, ny: int, L: float, W: float):
    """Flat ANCF3443 plate of nx x ny elements in the xy plane (config C3: 1000 x 1000, L = W = 1e-2, H = 1e-3).
    Node (i, j) -> id i*(ny+1)+j with coefficients [r, r_u = e_x, r_v = e_y, r_w = e_z]; element corners in the
    reference's order P1..P4 = (-,-), (+,-), (+,+), (-,+) (cpu_utils.cc:214-217).  Returns x, y, z ([4 nn]), conn."""
    ii, jj = np.meshgrid(np.arange(nx + 1), np.arange(ny + 1), indexing="ij")
    nn = (nx + 1) * (ny + 1)
    x = np.zeros((nn, 4)); y = np.zeros((nn, 4)); z = np.zeros((nn, 4))
    x[:, 0] = ii.ravel() * L; x[:, 1] = 1.0
    y[:, 0] = jj.ravel() * W; y[:, 2] = 1.0
    z[:, 3] = 1.0
    ci, cj = (a.ravel() for a in np.meshgrid(np.arange(nx), np.arange(ny), indexing="ij"))
    nid = lambda i, j: i * (ny + 1) + j  # noqa: E731
    conn = np.stack([nid(ci, cj), nid(ci + 1, cj), nid(ci + 1, cj + 1), nid(ci, cj + 1)], axis=1).astype(np.int32)
    return x.ravel(), y.ravel(), z.ravel(), conn


def ancf_chain(n_elem: int, L: float):
    """Straight ANCF3243 beam of n_elem elements along x (GridMeshGenerator's horizontal line, beam_sag layout)."""
    nn = n_elem + 1
    x = np.zeros((nn, 4)); y = np.zeros((nn, 4)); z = np.zeros((nn, 4))
    x[:, 0] = np.arange(nn) * L; x[:, 1] = 1.0
    y[:, 2] = 1.0
    z[:, 3] = 1.0
    conn = np.stack([np.arange(n_elem), np.arange(n_elem) + 1], axis=1).astype(np.int32)
    return x.ravel(), y.ravel(), z.ravel(), conn


def ancf_perturb(x, y, z, L: float, seed: int, pos: float = 0.05, grad: float = 0.02):
    """SURVEY.md 8d recipe: position slots moved by pos*L*U(-1,1), gradient slots by grad*U(-1,1), PCG64(seed)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    n = x.shape[0]
    amp = np.tile([pos * L, grad, grad, grad], n // 4)
    return tuple(a + amp * rng.uniform(-1.0, 1.0, n) for a in (x, y, z))
