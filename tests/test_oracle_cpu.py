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


def test_port_against_digests_at_the_baseline_shapes(cpu_extend):
    """The port at the C4 shape with per-instance plants and the host mirror + plan
    compiler at a long biped horizon, against digests of reference output."""
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import synthetic

    g = cases.golden("generic_c4_perinst_digest")
    for b in (0, 7):
        form = cases.formulate_generic(g["A_sys"][b], g["B_sys"][b], g["Q_w"], g["R_w"], int(g["N"]),
                                       g["rows"], refs=list(g["ref"][b]), backend=NumpyBackend())
        got = orc.all_qp_matrices(form, g["x0"][b][:, None])
        cases.assert_digest_close(got, g, "i%d_" % b, int(g["seed"]) + b, 1e-12)
    g = cases.golden("biped_long_digest")
    N, seed = 200, int(g["seed"]) + 200
    form = cases.formulate_biped(cases.default_configuration(N), omega=Param("omega"),
                                 target_vel=[Param("vel_x"), Param("vel_y")], backend=NumpyBackend())
    vals = synthetic.biped_inputs(seed, 1, N)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, 1)
    out = form.generate_all_qp_matrices_batched(given, {k: vals[k] for k in ("omega", "vel_x", "vel_y")})
    cases.assert_digest_close((out.G[0], out.h[0], out.P[0], out.q[0]), g, "n200_", seed, 1e-12)


def test_live_reference_pins_the_port_and_the_generic_recipes():
    """Where a reference tree is present: the port equals the unmodified reference on fresh
    random instances (not only on the stored goldens), and the two ways of driving the
    reference for per-instance plants (fresh formulation / re-incorporated dynamics) give
    the same bits."""
    from oracle import reference
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import synthetic
    import mpc_interface_b200.tools as tools

    if not reference.available():
        pytest.skip("no reference tree here")
    old, tools.extend_matrices = tools.extend_matrices, orc.extend_matrices
    try:
        N, B = 24, 5
        form = cases.formulate_biped(cases.default_configuration(N), omega=Param("omega"),
                                     target_vel=[Param("vel_x"), Param("vel_y")], backend=NumpyBackend())
        vals = synthetic.biped_inputs(77, B, N)
        given = synthetic.pack_given(form.given_ID, form.given_len, vals, B)
        ref_form = reference.biped_form(N)
        for b in range(B):
            one = {k: float(vals[k][b]) for k in ("omega", "vel_x", "vel_y")}
            cases.assert_qp_close(orc.all_qp_matrices(form, given[b][:, None], one),
                                  reference.biped_instance(ref_form, vals, b), 1e-13)
    finally:
        tools.extend_matrices = old
    nx, nu, H = 4, 2, 9
    _, _, Q, R = synthetic.generic_system(5, nx, nu)
    A, Bm = synthetic.generic_plants(5, 3, nx, nu)
    v = synthetic.generic_inputs(5, 3, nx)
    rows = np.random.default_rng(6).normal(size=[2, nx])
    plants = reference.GenericPlants(A[0], Bm[0], Q, R, H, rows)
    for b in range(3):
        fresh = reference.generic_instance(A[b], Bm[b], Q, R, H, rows, v["ref"][b], v["x0"][b])
        again = plants.instance(A[b], Bm[b], v["ref"][b], v["x0"][b])
        assert all(np.array_equal(x, y) for x, y in zip(fresh, again))
