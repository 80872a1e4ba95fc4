"""Solver hand-off format: the structural CSC pattern of the plan must equal
``scipy.sparse.csc_matrix(dense)`` of the reference's own matrices (int32
indices / indptr, bit-exact) -- the example feeds exactly that to the solver
(biped_mpc_loop.py:57-58)."""
import numpy as np
import pytest
import scipy.sparse as sp

import _cases as cases
from _numpy_backend import NumpyBackend
from mpc_interface_b200 import plan as plan_mod


@pytest.mark.parametrize("name", ["biped_n24_batch", "biped_n100_batch", "biped_n24_terminal"])
def test_pattern_equals_scipy_csc_of_the_reference(cpu_extend, name):
    g = cases.golden(name)
    ops = NumpyBackend()
    form, params = cases.biped_batched(g, ops)
    compiled = form._compiled()
    pP, pG, cP, cG = plan_mod.csc_patterns(compiled, ops)
    for b in range(g["Q"].shape[0]):
        for pattern, dense in ((pP, g["Q"][b]), (pG, g["A"][b])):
            ref = sp.csc_matrix(dense)
            assert pattern.indices.dtype == ref.indices.dtype == np.int32
            assert np.array_equal(pattern.indptr, ref.indptr)
            assert np.array_equal(pattern.indices, ref.indices)
            assert pattern.nnz == ref.nnz and pattern.padded % 2 == 0
    # the compact constants are the dense ones gathered in CSC order
    assert np.array_equal(cP[:, : pP.nnz], compiled.DP.reshape(len(compiled.P_terms), -1)[:, pP.flat])
    # upper triangle (OSQP convention)
    uP = plan_mod.csc_patterns(compiled, ops, upper_P=True)[0]
    ref = sp.csc_matrix(np.triu(g["Q"][0]))
    assert np.array_equal(uP.indptr, ref.indptr) and np.array_equal(uP.indices, ref.indices)


def test_pattern_generic_dense_blocks(cpu_extend):
    g = cases.golden("generic_small")
    ops = NumpyBackend()
    form, params, given = cases.generic_batched(g, ops)
    pP, pG, _, _ = plan_mod.csc_patterns(form._compiled(), ops)
    refP, refG = sp.csc_matrix(g["Q"][0]), sp.csc_matrix(g["A"][0])
    assert np.array_equal(pP.indices, refP.indices) and np.array_equal(pP.indptr, refP.indptr)
    assert np.array_equal(pG.indices, refG.indices) and np.array_equal(pG.indptr, refG.indptr)


@pytest.mark.parametrize("seed", range(660, 672))
def test_pattern_of_random_structures_equals_scipy_csc_of_the_reference(cpu_extend, seed):
    """The batch-wide CSC pattern (indices / indptr, int32) of random formulations around a shared
    plant -- auxiliary systems, scheduled / crossed / state-input costs, ``L`` matrices, any variable
    order (tests/_embedded_forms.py) -- is bit for bit scipy.sparse.csc_matrix of what the unmodified
    reference assembles for an instance (biped_mpc_loop.py:57-58)."""
    import _embedded_forms as forms
    from mpc_interface_b200 import Param
    from oracle import reference

    if not reference.available():
        pytest.skip("oracle/_ref is not staged")
    spec = forms.draw(seed, scheduled=True, general=True, with_L=True)
    omega, tau = 3.4, 0.1 * min(1.0, 30.0 / spec["N"])
    ops = NumpyBackend()
    form = forms.build(spec, forms.own_api(), omega, tau, Param("vel"), Param("gain"), backend=ops)
    pP, pG, _, _ = plan_mod.csc_patterns(form._compiled(), ops)
    vals = forms.inputs(spec, form, 1, seed + 1000)
    one = forms.build(spec, forms.reference_api(reference.load()), omega, tau, float(vals["vel"][0]), float(vals["gain"][0]))
    A, _h, Q, _q = one.generate_all_qp_matrices(one.arrange_given(
        {name: np.asarray(vals[name][0]).reshape([-1, 1]) for name in one.given_ID}))
    for pattern, dense in ((pP, Q), (pG, A)):
        ref = sp.csc_matrix(dense)
        assert pattern.indices.dtype == ref.indices.dtype == np.int32
        assert np.array_equal(pattern.indptr, ref.indptr) and np.array_equal(pattern.indices, ref.indices)
