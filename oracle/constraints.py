"""TEST INFRASTRUCTURE (see fe_oracle.h): general linear dof constraints — hanging nodes — at element level, restated from
libMesh in plain Python for small meshes.  Nothing in the product imports this.

    u_i = sum_j coef_ij u_j + g_i          one DofConstraintRow per constrained dof i (include/base/dof_map.h:97-123);
                                           an empty row is a Dirichlet value

  build_constraint_matrix_and_vector   src/base/dof_map_constraints.C:3543-3673 (and :3433-3541 without H)
  constrain (heterogeneously_...)      src/base/dof_map_constraints.C:2519-2614: K <- C^T K C, F <- C^T (F - K H), then the
                                       constrained rows: zero, 1 on the diagonal, -coef at the constraining dofs
                                       (asymmetric_constraint_rows, the default), rhs = g_i
  connected_dofs                       DofMap::find_connected_dofs, src/base/dof_map.C:2915-2975: what SparsityPattern::Build
                                       couples for an element (src/base/sparsity_pattern.C:139)
  assemble                             the per-element loop of introduction_ex4.C:393-563 with these pieces: Ke/Fe from the C
                                       oracle (unconstrained), constrained here, added in element order

Pinned against the reference's own functions: the four definitions are cut out of src/base/dof_map_constraints.C at build time
and compiled over the reference's DenseMatrix sources into oracle/_ref/libref_app.so; tests/test_oracle_ref_app.py compares
constrain() with them on hand-made rows (recursion, Dirichlet masters, self-constrained elements, both flags) and on every
element of locally refined HEX8 / HEX27 meshes — same widened dof sets, values to rounding (numpy's matrix products sum in
another order).  And by construction: with empty rows only it equals the C oracle's Dirichlet path entry for entry, and a
hanging-node mesh reproduces a linear (HEX27: quadratic) solution exactly (tests/test_oracle_constraints.py).
"""
import numpy as np


class ConstraintRows:
    """rows: {dof: [(dof_j, coef), ...]} (possibly empty lists), values: {dof: g} (missing = 0)."""

    def __init__(self, rows, values=None):
        self.rows = {int(k): [(int(j), float(c)) for j, c in v] for k, v in rows.items()}
        self.values = {int(k): float(v) for k, v in (values or {}).items()}

    def is_constrained(self, d):
        return int(d) in self.rows

    def arrays(self):
        """(dofs, row_ptr, cols, coefs, values) in ascending dof order: the argument of feasm_set_constraint_rows."""
        dofs = np.array(sorted(self.rows), dtype=np.uint32)
        ptr, cols, coefs, vals = [0], [], [], []
        for d in dofs:
            for j, c in self.rows[int(d)]:
                cols.append(j); coefs.append(c)
            ptr.append(len(cols))
            vals.append(self.values.get(int(d), 0.0))
        return dofs, np.array(ptr, dtype=np.uint64), np.array(cols, dtype=np.uint32), np.array(coefs, dtype=np.float64), np.array(vals, dtype=np.float64)


def build_constraint_matrix_and_vector(con, elem_dofs, called_recursively=False):
    """-> (C or None, H or None, elem_dofs extended in place like the reference).  None = the element is not constrained."""
    dof_set = set()
    have = False
    for d in elem_dofs:
        if con.is_constrained(d):
            have = True
            for j, _ in con.rows[d]:
                dof_set.add(j)
    if not have:
        return None, None
    dof_set -= set(elem_dofs)
    if not dof_set and called_recursively:
        return None, None
    old = len(elem_dofs)
    elem_dofs.extend(sorted(dof_set))  # std::set iteration order
    n = len(elem_dofs)
    Cm, H = np.zeros((old, n)), np.zeros(old)
    for i in range(old):
        d = elem_dofs[i]
        if con.is_constrained(d):
            for j, c in con.rows[d]:
                for k in range(n):
                    if elem_dofs[k] == j:
                        Cm[i, k] = c
            H[i] = con.values.get(d, 0.0)
        else:
            Cm[i, i] = 1.0
    Cn, Hn = build_constraint_matrix_and_vector(con, elem_dofs, True)
    if Cn is not None and Cm.shape[1] == Cn.shape[0] and Cn.shape[1] == len(elem_dofs):
        H = H + Cm @ Hn       # x = C y + h, y = D z + g  ->  x = (C D) z + (C g + h)
        Cm = Cm @ Cn
    return Cm, H


def constrain(con, Ke, Fe, elem_dofs, hetero=True, asymmetric=True):
    """-> (K', F', dofs') of (heterogeneously_)constrain_element_matrix_and_vector; unchanged inputs if nothing is constrained."""
    dofs = [int(d) for d in elem_dofs]
    Cm, H = build_constraint_matrix_and_vector(con, dofs)
    if Cm is None or Cm.shape[0] != Ke.shape[0] or Cm.shape[1] != len(dofs):
        return Ke, Fe, dofs
    F = Cm.T @ ((Fe - Ke @ H) if hetero else Fe)
    K = Cm.T @ Ke @ Cm
    for i, d in enumerate(dofs):
        if con.is_constrained(d):
            K[i, :] = 0.0
            K[i, i] = 1.0
            if asymmetric:
                for j, c in con.rows[d]:
                    for k, dk in enumerate(dofs):
                        if dk == j:
                            K[i, k] = -c
                if hetero:
                    F[i] = con.values.get(d, 0.0)
            elif hetero:
                F[i] = 0.0
            # constrain_element_matrix_and_vector (:2439-2514) leaves rhs(i) = (C^T F)(i): zero for rows as
            # DofMap::process_constraints leaves them (no constraining dof is itself constrained, so column i of C is zero)
    return K, F, dofs


def connected_dofs(con, elem_dofs):
    s = set(int(d) for d in elem_dofs)
    while True:
        add = set()
        for d in s:
            if con.is_constrained(d):
                for j, _ in con.rows[d]:
                    if j not in s:
                        add.add(j)
        if not add:
            return sorted(s)
        s |= add


def sparsity(con, elem_dofs, n_dofs):
    """CSR graph: every element couples all of its connected dofs."""
    rows = [set() for _ in range(n_dofs)]
    for ed in elem_dofs:
        cd = connected_dofs(con, ed)
        for r in cd:
            rows[r].update(cd)
    rp = np.zeros(n_dofs + 1, dtype=np.uint64)
    for r in range(n_dofs):
        rp[r + 1] = rp[r] + len(rows[r])
    ci = np.array([c for r in rows for c in sorted(r)], dtype=np.uint32)
    return rp, ci


def assemble(con, elem_dofs, Ke, Fe, rp, ci, hetero=True):
    """(values, rhs): constrained element matrices added in element order into the pattern (rp, ci)."""
    n = len(rp) - 1
    vals, rhs = np.zeros(int(rp[-1])), np.zeros(n)
    rp = rp.astype(np.int64)
    for e in range(len(elem_dofs)):
        K, F, dofs = constrain(con, Ke[e], Fe[e], elem_dofs[e], hetero=hetero)
        for i, r in enumerate(dofs):
            seg = ci[rp[r]:rp[r + 1]]
            pos = np.searchsorted(seg, dofs)
            if (pos >= len(seg)).any() or (seg[np.minimum(pos, len(seg) - 1)] != np.asarray(dofs)).any():
                raise ValueError("entry outside the sparsity pattern")
            np.add.at(vals, rp[r] + pos, K[i])
            rhs[r] += F[i]
    return vals, rhs
