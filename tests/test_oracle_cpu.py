"""The numpy oracle against the reference's golden vectors (CPU).

The goldens were written by the UNMODIFIED reference (tests/golden/make_goldens.py);
lip_matrices.npz holds the arrays of the reference's own pickle
python/tests/LIP_matrices (test_dynamics.py:139-152).  The oracle runs over this
repo's host objects, so these tests pin both the oracle's arithmetic and the
host mirror's bookkeeping (index layout, row order)."""
import numpy as np
import pytest

from oracle import qp_oracle as orc
import _cases as cases
from _numpy_backend import NumpyBackend


def test_lip_pickle_golden():
    g = cases.golden("lip_matrices")
    S, U = orc.extend_matrices(36, g["A"], g["B"])
    # same tolerance as the reference's test (np.isclose defaults)
    assert np.isclose(U[0], g["U"]).all()
    assert np.isclose(S, g["S"]).all()
    assert cases.rel_err(U[0], g["U"]) < 1e-13 and cases.rel_err(S, g["S"]) < 1e-13


@pytest.mark.parametrize("tag", list("abcde"))
def test_extend_cases(tag):
    g = cases.golden("extend_cases")
    A, B = g[tag + "_A"], g[tag + "_B"]
    S, U = orc.extend_matrices(g[tag + "_S"].shape[0], A, B)
    assert S.shape == g[tag + "_S"].shape and len(U) == B.shape[1]
    assert cases.rel_err(S, g[tag + "_S"]) < 1e-14
    assert cases.rel_err(np.stack(U), g[tag + "_U"]) < 1e-14


def test_plan_steps_is_integer_exact():
    E = orc.plan_steps(24, 0, regular_time=12)
    assert E.dtype.kind == "i" and E.shape == (24, 2)
    assert E[:, 0].tolist() == [0] + [1] * 23 and E[:, 1].tolist() == [0] * 13 + [1] * 11


def test_biped_c1(cpu_extend):
    g = cases.golden("biped_c1")
    form = cases.formulate_biped(cases.default_configuration(), backend=NumpyBackend())
    cases.check_layout(form, g)
    got = orc.all_qp_matrices(form, g["given"])
    cases.assert_qp_close(got, (g["A"], g["h"], g["Q"], g["q"]), 1e-14)
    PM = orc.preview_matrices(form)
    assert cases.rel_err(np.hstack(PM["(b+n-s)_x"]), g["PM_bns_x"]) < 1e-14


@pytest.mark.parametrize("name", ["biped_n24_batch", "biped_n24_terminal", "biped_n100_batch"])
def test_biped_per_instance_omega(cpu_extend, name):
    g = cases.golden(name)
    form, params = cases.biped_batched(g, NumpyBackend())
    cases.check_layout(form, g)
    for b in range(g["given"].shape[0]):
        one = {k: float(v[b]) for k, v in params.items()}
        got = orc.all_qp_matrices(form, g["given"][b][:, None], one)
        cases.assert_qp_close(got, (g["A"][b], g["h"][b], g["Q"][b], g["q"][b]), 1e-13)


def test_reference_test_body_case(cpu_extend):
    g = cases.golden("test_body")
    body = cases.body_test_formulation(NumpyBackend())
    cases.check_layout(body, g)
    got = orc.all_qp_matrices(body, g["given"])
    assert got[0].shape == (54, body.optim_len)  # test_body.py:145-147
    cases.assert_qp_close(got, (g["A"], g["h"], g["Q"], g["q"]), 1e-13)
    PM = orc.preview_matrices(body)
    cQ, cq = orc.qp_cost(body.goals["crossed"], PM, g["given"])
    assert cases.rel_err(cQ, g["cQ"]) < 1e-13 and cases.rel_err(cq, g["cq"]) < 1e-13
    A1, h1 = orc.qp_constraint(body.constraints["steppingArea"][0], PM, g["given"])
    assert cases.rel_err(A1, g["A1"]) < 1e-13 and cases.rel_err(h1, g["h1"]) < 1e-13


@pytest.mark.parametrize("name", ["generic_small", "generic_c4_shape"])
def test_generic(cpu_extend, name):
    g = cases.golden(name)
    form, params, given = cases.generic_batched(g, NumpyBackend())
    cases.check_layout(form, g)
    for b in range(given.shape[0]):
        one = {k: float(v[b]) for k, v in params.items()}
        got = orc.all_qp_matrices(form, given[b][:, None], one)
        cases.assert_qp_close(got, (g["A"][b], g["h"][b], g["Q"][b], g["q"][b]), 1e-12)
