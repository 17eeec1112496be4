"""The oracle's build_cube restatement (oracle/orc_mesh.cpp) against the REFERENCE's own MeshTools::Generation::build_cube
(src/mesh/mesh_generation.C:317-1938 with its idx() helpers :81-157): the definitions are cut out of the file where it lies and
run over a mesh that records what the generator does (oracle/ref_mesh/ref_mesh_world.h -> oracle/_ref/libref_mesh.so), followed
by the reference's own ReplicatedMesh::renumber_nodes_and_elements (src/mesh/replicated_mesh.C:700-857), the numbering step of
prepare_for_use.  Compared BEFORE that step — lattice indexing, node coordinates (the reference's own scaling arithmetic —
bitwise), connectivity in element order, the boundary id of every boundary side, for TET10 the 24-tet split of every cell — and
AFTER it: the final node numbering and coordinates of HEX8 / HEX20 / HEX27 / PRISM18 / TET10 meshes are the reference's — for
TET10 through the reference's own UnstructuredMesh::all_second_order_range with all_increased_order_range / transfer_elem /
map_hi_order_node (src/mesh/unstructured_mesh.C), Elem::second_order_equivalent_type and Tet10's mid-edge table, which raise the
TET4 split and renumber.  Runs where /root/reference exists."""
import numpy as np
import pytest

from oracle import oracle as O
from oracle import ref as R

pytestmark = pytest.mark.skipif(not R.mesh_available(), reason="oracle/_ref/libref_mesh.so exists only where /root/reference does")
NAME = {O.HEX8: "HEX8", O.HEX20: "HEX20", O.HEX27: "HEX27", O.TET10: "TET10", O.PRISM18: "PRISM18"}
BOUNDS = [(-1.0, 1.0, -1.0, 1.0, -1.0, 1.0), (0.0, 1.0, 0.0, 0.3, 0.0, 0.1), (-1.0, 1.0, -0.5, 0.25, 0.0, 3.0)]


def _compacted(r):
    """The generator's record with the lattice nodes no element uses taken out (HEX20: face and cell centres; the tet split:
    mid-edge lattice nodes) and the gaps closed in ascending id order — the oracle's renumber_nodes=False form."""
    xb, cb = r["before"]
    used = np.unique(cb)
    rank = np.full(xb.shape[0], -1, dtype=np.int64)
    rank[used] = np.arange(len(used))
    return rank[cb], xb[used]


@pytest.mark.parametrize("etype", [O.HEX8, O.HEX20, O.HEX27, O.PRISM18, O.TET10], ids=lambda e: NAME[e])
@pytest.mark.parametrize("dims", [(3, 2, 4), (1, 1, 1), (5, 5, 5)], ids=lambda d: "x".join(map(str, d)))
@pytest.mark.parametrize("bounds", BOUNDS, ids=["ex4_cube", "ex6_bar", "ragged_box"])
def test_build_cube_before_renumbering_is_the_references(etype, dims, bounds):
    r = R.build_cube(NAME[etype], *dims, bounds)
    assert r["prepared"] and r["second_order_requests"] == (1 if etype == O.TET10 else 0)
    m = O.Mesh(etype, *dims, bounds=bounds, renumber_nodes=False)
    xyz, conn, _ = m.arrays()
    rc, rx = _compacted(r)
    nv = rc.shape[1]  # TET10: the four vertices of the TET4 the reference builds before all_second_order
    assert rc.shape[0] == conn.shape[0] and np.array_equal(r["elem_id"], np.arange(conn.shape[0]))
    assert np.array_equal(rc, conn[:, :nv]), "connectivity differs"
    assert np.array_equal(rx, xyz[: len(rx)]), f"node coordinates differ by {np.abs(rx - xyz[:len(rx)]).max()}"
    if etype != O.TET10:
        assert len(rx) == len(xyz)


@pytest.mark.parametrize("etype", [O.HEX8, O.HEX20, O.HEX27, O.PRISM18, O.TET10], ids=lambda e: NAME[e])
@pytest.mark.parametrize("dims", [(3, 2, 4), (1, 1, 1), (6, 6, 6)], ids=lambda d: "x".join(map(str, d)))
def test_final_numbering_is_the_references_renumber_nodes_and_elements(etype, dims):
    """The mesh as an assembly sees it: nodes renumbered in order of first touch by the elements, orphans removed; for TET10 after
    the reference's all_second_order created the mid-edge nodes (coordinates: the mean of the two vertices, bitwise)."""
    for bounds in BOUNDS[:2]:
        r = R.build_cube(NAME[etype], *dims, bounds)
        xyz, conn, _ = O.Mesh(etype, *dims, bounds=bounds).arrays()
        assert np.array_equal(r["conn"], conn) and np.array_equal(r["xyz"], xyz)
        assert np.array_equal(r["elem_id"], np.arange(len(conn)))


@pytest.mark.parametrize("etype", [O.HEX8, O.HEX27], ids=lambda e: NAME[e])
def test_boundary_sides_are_the_references(etype):
    """BoundaryInfo::add_side calls of build_cube: 0 back (min z), 1 bottom (min y), 2 right (max x), 3 top (max y), 4 left (min x),
    5 front (max z)."""
    dims = (3, 2, 4)
    r = R.build_cube(NAME[etype], *dims, BOUNDS[1])
    se, sn, sb = r["sides"]
    m = O.Mesh(etype, *dims, bounds=BOUNDS[1], renumber_nodes=False)
    total = 0
    for bid in range(6):
        e, s = m.boundary_sides(bid)
        want = sorted(zip(e.tolist(), s.tolist()))
        got = sorted((int(a), int(b)) for a, b, c in zip(se, sn, sb) if c == bid)
        assert got == want, f"boundary id {bid}"
        total += len(want)
    assert total == len(se) == 2 * (3 * 2 + 3 * 4 + 2 * 4)


def test_reference_rejects_what_it_rejects():
    with pytest.raises(RuntimeError, match="Unrecognized 3D element type"):
        R.build_cube("QUAD9", 2, 2, 2, BOUNDS[0])


@pytest.mark.parametrize("etype", [O.HEX8, O.HEX20, O.HEX27, O.PRISM18, O.TET10], ids=lambda e: NAME[e])
@pytest.mark.parametrize("n_vars", [1, 3])
def test_dof_numbering_is_the_references_distribute_local_dofs_var_major(etype, n_vars):
    """DofMap::distribute_local_dofs_var_major (src/base/dof_map.C:1417-1535, the reference's own definition) on the reference's
    own mesh: nodes get their dof base in order of first touch by the elements; one variable group of n_vars LAGRANGE variables
    (what add_variable makes of u, v, w in ex6), so variable v of a node sits at base + v.  The oracle's node bases, n_dofs and
    the element dof lists DofMap::dof_indices returns (variable-major) are those."""
    for dims in [(3, 2, 4), (1, 1, 1), (5, 4, 3)]:
        r = R.build_cube(NAME[etype], *dims, BOUNDS[1])
        n_dofs, base = R.distribute_dofs(n_vars, len(r["xyz"]))
        m = O.Mesh(etype, *dims, bounds=BOUNDS[1])
        s = O.System(m, n_vars)
        ed, nb, fd = s.dofs()
        assert n_dofs == s.n_dofs == n_vars * m.n_nodes and fd.tolist() == [0, n_dofs]
        assert np.array_equal(nb, base)
        assert np.array_equal(ed, np.concatenate([base[r["conn"]] + v for v in range(n_vars)], axis=1))


@pytest.mark.parametrize("dims,n_parts", [((3, 2, 4), 2), ((3, 2, 4), 5), ((5, 5, 5), 8), ((2, 2, 2), 3), ((3, 3, 3), 7), ((4, 4, 4), 64)])
def test_linear_partitioner_is_the_references(dims, n_parts):
    """LinearPartitioner::partition_range (src/partitioning/linear_partitioner.C:36-117, the reference's own definition): blocks of
    n_elem / n_parts elements in id order, the remainder to rank 0 — the element shards of the multi-GPU runs.  (More parts than
    elements divides by zero in the reference; not asked.)"""
    r = R.build_cube("HEX27", *dims, BOUNDS[0])
    pid = R.partition_linear(n_parts, len(r["conn"]))
    m = O.Mesh(O.HEX27, *dims, bounds=BOUNDS[0])
    m.partition_linear(n_parts)
    assert np.array_equal(pid, m.partition()[0])


@pytest.mark.skipif(not R.app_available(), reason="oracle/_ref/libref_app.so not built")
@pytest.mark.parametrize("etype,n,order,through_main", [(O.HEX27, 4, 2, True), (O.HEX8, 5, 1, True), (O.TET10, 2, 2, False), (O.HEX20, 3, 2, False),
                                                        (O.PRISM18, 3, 2, False)], ids=lambda v: NAME.get(v, str(v)) if isinstance(v, int) and v in NAME else str(v))
def test_the_whole_chain_by_the_references_own_code(etype, n, order, through_main):
    """Nothing of the oracle goes IN: the mesh is the reference's build_cube (+ all_second_order for TET10) after its own
    renumbering, the dof numbering its distribute_local_dofs_var_major, the sparsity graph the element-couples-its-dofs definition
    written out in numpy, the boundary nodes a coordinate test; then main() / assemble_poisson of introduction_ex4.C assemble over
    them.  What comes OUT — the global matrix, the right-hand side, the Dirichlet set — is the oracle's, bit for bit."""
    bounds = BOUNDS[0]
    r = R.build_cube(NAME[etype], n, n, n, bounds)
    xyz, conn = r["xyz"], r["conn"]
    n_dofs, base = R.distribute_dofs(1, len(xyz))
    ed = base[conn]
    # SparsityPattern::Build for one variable without constraints: every element couples all of its dofs
    pairs = np.unique((ed[:, :, None].astype(np.int64) * n_dofs + ed[:, None, :]).reshape(-1))
    rows, cols = pairs // n_dofs, pairs % n_dofs
    rp = np.zeros(n_dofs + 1, dtype=np.uint64)
    rp[1:] = np.cumsum(np.bincount(rows, minlength=n_dofs))
    mask = np.zeros(len(xyz), dtype=np.uint8)  # build_cube's ids: 0 min z, 1 min y, 2 max x, 3 max y, 4 min x, 5 max z
    for b, (axis, value) in enumerate([(2, bounds[4]), (1, bounds[2]), (0, bounds[1]), (1, bounds[3]), (0, bounds[0]), (2, bounds[5])]):
        mask[xyz[:, axis] == value] |= 1 << b
    node_dof = base[None, :]
    nnz, nd = R.load_world_arrays(NAME[etype], (n, n, n), bounds, xyz, conn, mask, node_dof, rp, cols.astype(np.uint32))
    out = R.run_ex4(["-d", 3, "-n", n] + (["-o", "FIRST"] if order == 1 else []), nnz, nd) if through_main else R.assemble_ex4(order, nnz, nd)
    assert out["untouched_entries"] == 0
    m = O.Mesh(etype, n, n, n, bounds=bounds)
    s = O.System(m, 1)
    s.add_dirichlet((0, 1, 2, 3, 4, 5), 0)
    rp0, ci0 = s.sparsity()
    assert np.array_equal(rp0, rp) and np.array_equal(ci0, cols)
    v, rhs, _ = s.assemble(O.POISSON_EX4)
    cd, cv = s.constraints()
    assert np.array_equal(cd, out["con_dof"]) and np.array_equal(cv, out["con_val"])
    assert np.array_equal(v, out["values"]) and np.array_equal(rhs, out["rhs"])


# ---- PRODUCT host code against the reference's own generator and numbering, directly (no GPU: host entry points of libfeasm.so) -----
@pytest.mark.parametrize("etype", [O.HEX8, O.HEX20, O.HEX27, O.PRISM18, O.TET10], ids=lambda e: NAME[e])
def test_product_build_cube_is_the_references_generator(etype):
    """feasm_build_cube (libmesh_b200/csrc/feasm_host.cpp, written independently of the oracle) against the record of the
    reference's own build_cube: the lattice nodes an element uses in ascending id order (the product does not renumber nodes by
    first touch — its dof numbering, below, does not depend on that), coordinates bitwise, connectivity; TET10: the four vertices
    of every tet of the reference's 24-split."""
    from libmesh_b200 import capi as F

    for dims, bounds in [((3, 2, 4), BOUNDS[1]), ((2, 2, 2), BOUNDS[0])]:
        xyz, conn, _ = F.build_cube(etype, *dims, bounds)
        rc, rx = _compacted(R.build_cube(NAME[etype], *dims, bounds))
        nv = rc.shape[1]
        assert np.array_equal(conn[:, :nv], rc) and np.array_equal(xyz[: len(rx)], rx)


@pytest.mark.parametrize("etype", [O.HEX8, O.HEX27, O.TET10], ids=lambda e: NAME[e])
@pytest.mark.parametrize("n_vars", [1, 3])
def test_product_dof_numbering_is_the_references(etype, n_vars):
    """feasm_distribute_dofs (host) on the reference's own finished mesh against the reference's own
    DofMap::distribute_local_dofs_var_major."""
    from libmesh_b200 import capi as F

    r = R.build_cube(NAME[etype], 3, 2, 2, BOUNDS[0])
    n_dofs, base = R.distribute_dofs(n_vars, len(r["xyz"]))
    n, pbase, first, _ = F.distribute_dofs(r["conn"], len(r["xyz"]), n_vars, 1, np.zeros(len(r["conn"]), dtype=np.int32))
    assert n == n_dofs and np.array_equal(pbase, base) and first.tolist() == [0, n_dofs]
    assert np.array_equal(F.elem_dofs_from_nodes(r["conn"], pbase, n_vars), np.concatenate([base[r["conn"]] + v for v in range(n_vars)], axis=1))
