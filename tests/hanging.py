"""Test utility: a HEX8 cube mesh with one level of local refinement and the hanging-node constraint rows libMesh would
compute for it (compute_proj_constraints on a level-one-conforming mesh: a node hanging on a coarse edge is the mean of the
edge's two vertices, a node hanging on a coarse face the mean of the face's four)."""
import numpy as np

# HEX8 local node -> lattice offset (src/geom/cell_hex8.C node order)
HEX8 = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0], [0, 0, 1], [1, 0, 1], [1, 1, 1], [0, 1, 1]])
EDGES = [(0, 1), (1, 2), (2, 3), (3, 0), (0, 4), (1, 5), (2, 6), (3, 7), (4, 5), (5, 6), (6, 7), (7, 4)]
FACES = [(0, 3, 2, 1), (0, 1, 5, 4), (1, 2, 6, 5), (2, 3, 7, 6), (3, 0, 4, 7), (4, 5, 6, 7)]


def refined_cube(n, refine, bounds=(-1.0, 1.0)):
    """n^3 HEX8 elements on [lo, hi]^3; the elements with refine(i, j, k) true are replaced by their 8 children.
    Returns (xyz, conn, boundary node mask, hanging {node: [(node, coef), ...]}).  Active elements: the unrefined coarse ones
    in id order, then the children (Elem::refine appends them)."""
    lo, hi = bounds
    key_of, xyz = {}, []

    def node(a, b, c):  # coordinates on the 2n lattice
        k = (a, b, c)
        if k not in key_of:
            key_of[k] = len(xyz)
            xyz.append([lo + (hi - lo) * a / (2.0 * n), lo + (hi - lo) * b / (2.0 * n), lo + (hi - lo) * c / (2.0 * n)])
        return key_of[k]

    coarse, fine = [], []
    for k in range(n):
        for j in range(n):
            for i in range(n):
                if not refine(i, j, k):
                    coarse.append([node(2 * (i + o[0]), 2 * (j + o[1]), 2 * (k + o[2])) for o in HEX8])
    for k in range(n):
        for j in range(n):
            for i in range(n):
                if refine(i, j, k):
                    for ck in range(2):
                        for cj in range(2):
                            for c_i in range(2):
                                fine.append([node(2 * i + c_i + o[0], 2 * j + cj + o[1], 2 * k + ck + o[2]) for o in HEX8])
    conn = np.array(coarse + fine, dtype=np.uint32)
    inv = {v: k for k, v in key_of.items()}
    hanging = {}
    for el in coarse:  # a node at the middle of an edge / face of an ACTIVE coarse element hangs there
        for a, b in EDGES:
            ka, kb = inv[el[a]], inv[el[b]]
            mid = tuple((x + y) // 2 for x, y in zip(ka, kb))
            if mid in key_of:
                hanging[key_of[mid]] = [(el[a], 0.5), (el[b], 0.5)]
        for f in FACES:
            ks = [inv[el[v]] for v in f]
            mid = tuple(sum(c) // 4 for c in zip(*ks))
            if mid in key_of and key_of[mid] not in (el[v] for v in f):
                hanging[key_of[mid]] = [(el[v], 0.25) for v in f]
    xyz = np.array(xyz)
    mask = np.zeros(len(xyz), dtype=np.uint8)
    for (a, b, c), idx in key_of.items():
        mask[idx] = ((c == 0) << 0) | ((b == 0) << 1) | ((a == 2 * n) << 2) | ((b == 2 * n) << 3) | ((a == 0) << 4) | ((c == 2 * n) << 5)
    return xyz, conn, mask, hanging


# ---- HEX27 ------------------------------------------------------------------------------------------------------------
# HEX27 local node -> offset in half-element units (src/geom/cell_hex27.C node order: corners, mid-edges, face centres, centre)
_I0 = [0, 1, 1, 0, 0, 1, 1, 0, 2, 1, 2, 0, 0, 1, 1, 0, 2, 1, 2, 0, 2, 2, 1, 2, 0, 2, 2]
_I1 = [0, 0, 1, 1, 0, 0, 1, 1, 0, 2, 1, 2, 0, 0, 1, 1, 0, 2, 1, 2, 2, 0, 2, 1, 2, 2, 2]
_I2 = [0, 0, 0, 0, 1, 1, 1, 1, 0, 0, 0, 0, 2, 2, 2, 2, 1, 1, 1, 1, 0, 2, 2, 2, 2, 1, 2]
_OFF = {0: 0, 1: 2, 2: 1}
HEX27 = np.array([[_OFF[a], _OFF[b], _OFF[c]] for a, b, c in zip(_I0, _I1, _I2)])


def _quad_weights(rel):
    """1D quadratic Lagrange interpolation from the coarse nodes at rel = 0, 2, 4 (parameters -1, 0, 1) to the lattice point
    rel in 0..4 (fe_lagrange_shape_1D.h): {coarse rel: weight}, zero weights dropped."""
    p = (rel - 2) / 2.0
    w = {0: 0.5 * p * (p - 1.0), 2: 1.0 - p * p, 4: 0.5 * p * (p + 1.0)}
    return {k: v for k, v in w.items() if v != 0.0}


def refined_cube27(n, refine, bounds=(-1.0, 1.0)):
    """The HEX27 analogue of refined_cube: n^3 HEX27 elements, those with refine(i, j, k) true replaced by their 8 children.
    A child node on a face (or edge) of an ACTIVE coarse element that is not one of that element's own nodes hangs: its value
    is the coarse side's quadratic interpolant there — for LAGRANGE what compute_proj_constraints produces on a
    level-one-conforming mesh — i.e. the tensor product of the 1D weights 3/8, 3/4, -1/8 at the quarter points."""
    lo, hi = bounds
    key_of, xyz = {}, []

    def node(a, b, c):  # coordinates on the 4n lattice (a coarse element spans 4 units, a child 2)
        k = (a, b, c)
        if k not in key_of:
            key_of[k] = len(xyz)
            xyz.append([lo + (hi - lo) * a / (4.0 * n), lo + (hi - lo) * b / (4.0 * n), lo + (hi - lo) * c / (4.0 * n)])
        return key_of[k]

    coarse, coarse_at, fine = [], [], []
    for k in range(n):
        for j in range(n):
            for i in range(n):
                if not refine(i, j, k):
                    coarse.append([node(4 * i + 2 * o[0], 4 * j + 2 * o[1], 4 * k + 2 * o[2]) for o in HEX27])
                    coarse_at.append((4 * i, 4 * j, 4 * k))
    for k in range(n):
        for j in range(n):
            for i in range(n):
                if refine(i, j, k):
                    for ck in range(2):
                        for cj in range(2):
                            for c_i in range(2):
                                fine.append([node(4 * i + 2 * c_i + o[0], 4 * j + 2 * cj + o[1], 4 * k + 2 * ck + o[2]) for o in HEX27])
    conn = np.array(coarse + fine, dtype=np.uint32)
    hanging = {}
    for (x0, y0, z0) in coarse_at:
        base = (x0, y0, z0)
        for axis in range(3):
            s_ax, t_ax = [d for d in range(3) if d != axis]
            for side in (0, 4):
                for rs in range(5):
                    for rt in range(5):
                        if rs % 2 == 0 and rt % 2 == 0:
                            continue  # one of the coarse element's own nodes
                        rel = [0, 0, 0]
                        rel[axis], rel[s_ax], rel[t_ax] = side, rs, rt
                        key = tuple(b + r for b, r in zip(base, rel))
                        if key not in key_of:
                            continue
                        row = []
                        for ms, ws in _quad_weights(rs).items():
                            for mt, wt in _quad_weights(rt).items():
                                mrel = [0, 0, 0]
                                mrel[axis], mrel[s_ax], mrel[t_ax] = side, ms, mt
                                row.append((key_of[tuple(b + r for b, r in zip(base, mrel))], ws * wt))
                        row.sort()
                        h = key_of[key]
                        assert h not in hanging or hanging[h] == row, "two coarse neighbours disagree on a hanging node"
                        hanging[h] = row
    xyz = np.array(xyz)
    mask = np.zeros(len(xyz), dtype=np.uint8)
    for (a, b, c), idx in key_of.items():
        mask[idx] = ((c == 0) << 0) | ((b == 0) << 1) | ((a == 4 * n) << 2) | ((b == 4 * n) << 3) | ((a == 0) << 4) | ((c == 4 * n) << 5)
    return xyz, conn, mask, hanging
