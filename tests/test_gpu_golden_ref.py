"""GPU parity against outputs of the reference itself: the CUDA path, through the C ABI, against the global matrix and right-hand
side that the reference's own introduction_ex4.C and systems_of_equations_ex6.C produced (tests/golden/reference_ex4_runs.npz,
reference_ex6_runs.npz, written by tests/golden/make_ref_app_golden.py from the reference's sources compiled where they lie).  Same bar as test_gpu_parity.py:
pattern bit-exact, values within 1e-12 per entry; the in-kernel finite-difference forcing at 1e-11 of max|rhs|,
the tabulated forcing at 1e-12."""
import numpy as np
import pytest

from conftest import PARITY_TOL, csr_parity_error, vec_parity_error
from test_golden_ref_app import G, G6, GH, golden_case, golden_case6, golden_case_h
from test_gpu_parity import make_ctx

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", [str(n) for n in G["names"]])
def test_cuda_path_against_the_references_own_output(name):
    from libmesh_b200 import capi as F

    m, s = golden_case(name)
    ctx = make_ctx(m, s)
    rp, ci = s.sparsity()
    rp2, ci2 = ctx.get_pattern()
    assert np.array_equal(rp, rp2) and np.array_equal(ci, ci2)
    vg, rg = G[name + "/values"], G[name + "/rhs"]
    assert len(vg) == ctx.nnz and len(rg) == s.n_dofs
    # as the reference runs it: forcing by finite differences of the exact solution, inside the kernel
    ctx.assemble(F.make_params(F.POISSON, 0, F.FORCING_EX4_FD))
    v, r = ctx.get_values()
    assert csr_parity_error(v, vg, rp) < PARITY_TOL
    assert np.abs(r - rg).max() < 1e-11 * np.abs(rg).max()
    # forcing tabulated at the quadrature points (the oracle's table): the right-hand side at the 1e-12 bar as well
    f_tab, _ = s.forcing_at_qps(0)
    ctx.assemble(F.make_params(F.POISSON, 0, F.FORCING_TABULATED, f_host=f_tab))
    v, r = ctx.get_values()
    assert csr_parity_error(v, vg, rp) < PARITY_TOL
    assert vec_parity_error(r, rg) < PARITY_TOL
    ctx.close()


@pytest.mark.parametrize("name", [str(n) for n in G6["names"]])
def test_cuda_elasticity_against_the_references_own_output(name):
    """systems_of_equations_ex6.C: three variables, the traction on the max-x sides, homogeneous Dirichlet on the min-x side."""
    from libmesh_b200 import capi as F

    m, s, e, sd = golden_case6(name)
    ctx = make_ctx(m, s)
    ctx.set_side_loads(e, sd, (0.0, 0.0, -1.0))
    rp, ci = s.sparsity()
    rp2, ci2 = ctx.get_pattern()
    assert np.array_equal(rp, rp2) and np.array_equal(ci, ci2)
    ctx.assemble(F.make_params(F.ELASTICITY, 0, F.FORCING_NONE, F.CONSTRAIN_HOMOGENEOUS))
    v, r = ctx.get_values()
    assert csr_parity_error(v, G6[name + "/values"], rp) < PARITY_TOL
    assert vec_parity_error(r, G6[name + "/rhs"]) < PARITY_TOL
    ctx.close()


@pytest.mark.parametrize("name", [str(n) for n in GH["names"]])
def test_cuda_hanging_node_assembly_against_the_references_own_output(name):
    """feasm_set_constraint_rows: locally refined HEX8 / HEX27 meshes against the system the reference's own element loop and
    DofMap::heterogeneously_constrain_element_matrix_and_vector produced (tests/golden/reference_hanging_runs.npz)."""
    from libmesh_b200 import capi as F

    xyz, conn, m, s, ed, con, (rp, ci) = golden_case_h(name)
    ctx = F.Context(0)
    ctx.set_mesh(m.elem_type, xyz, conn)
    ctx.set_dofs(ed, s.n_dofs)
    ctx.set_constraint_rows(*con.arrays())
    ctx.build_pattern()
    rp2, ci2 = ctx.get_pattern()
    assert np.array_equal(rp, rp2) and np.array_equal(ci, ci2)
    f_tab, _ = s.forcing_at_qps(0)
    ctx.assemble(F.make_params(F.POISSON, 0, F.FORCING_TABULATED, f_host=f_tab))
    v, r = ctx.get_values()
    assert csr_parity_error(v, GH[name + "/values"], rp) < PARITY_TOL
    assert vec_parity_error(r, GH[name + "/rhs"]) < PARITY_TOL
    ctx.close()
