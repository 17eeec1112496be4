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
