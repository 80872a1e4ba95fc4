"""The plan compiler (symbolic host logic) against the goldens, on CPU.

``NumpyBackend`` (tests/_numpy_backend.py) stands in for the CUDA backend so
the monomial bookkeeping, row offsets and constant layout of plan.py can be
checked without a GPU; the same comparisons run on the real kernels in
test_gpu_parity.py."""
import numpy as np
import pytest

import _cases as cases
from _numpy_backend import NumpyBackend
from mpc_interface_b200 import Cost, Param, plan as planmod


def test_biped_c1_single_instance(cpu_extend):
    g = cases.golden("biped_c1")
    form = cases.formulate_biped(cases.default_configuration(), backend=NumpyBackend())
    A, h, Q, q = form.generate_all_qp_matrices(g["given"])
    assert A.shape == (108, 52) and h.shape == (108, 1) and Q.shape == (52, 52) and q.shape == (52, 1)
    cases.assert_qp_close((A, h, Q, q), (g["A"], g["h"], g["Q"], g["q"]))
    Mg, Mo = form.PM["(b+n-s)_x"]
    assert cases.rel_err(np.hstack([Mg, Mo]), g["PM_bns_x"]) < 1e-14
    assert len(form.PM) == len(form.definitions) == 30


@pytest.mark.parametrize("name", ["biped_n24_batch", "biped_n24_terminal", "biped_n100_batch"])
def test_biped_batched(cpu_extend, name):
    g = cases.golden(name)
    form, params = cases.biped_batched(g, NumpyBackend())
    out = form.generate_all_qp_matrices_batched(g["given"], params)
    cases.assert_qp_close((out.G, out.h, out.P, out.q), (g["A"], g["h"], g["Q"], g["q"]))
    compiled = form._compiled()
    assert form.parameter_names() == ["omega", "vel_x", "vel_y"]
    assert max(len(m) for m in compiled.monomials) <= planmod.MAX_FACTORS
    assert compiled.row_offsets[0] == 0 and sum(compiled.limit_rows) == g["A"].shape[1]


def test_plan_structure_biped_n100(cpu_extend):
    g = cases.golden("biped_n100_batch")
    form, _ = cases.biped_batched(g, NumpyBackend())
    p = form._compiled()
    assert (p.nv, p.ng, p.nc) == (218, 208, 440)
    # stepping area 4 x 9, support polygon 4 x 100, terminal 4 x 1
    assert p.limit_rows == [9] * 4 + [100] * 4 + [1] * 4
    assert p.row_offsets == list(np.cumsum([0] + p.limit_rows[:-1]))
    # P is a polynomial of degree 2 in -1/omega^2: 3 constant matrices
    assert len(p.P_terms) == 3 and p.DP.shape == (3, 218, 218)
    assert p.DG.shape[1:] == (440, 218) and p.DY.shape[1:] == (658, 209)
    # integer tables the kernels index with
    segs = p.Y_segs
    assert segs[0, 2] == 0 and segs[-1, 3] == len(p.Y_val) == len(p.Y_col)
    assert (segs[:, 2] < segs[:, 3]).all() and (segs[1:, 2] == segs[:-1, 3]).all()
    assert (np.diff(segs[:, 0]) >= 0).all() and p.Y_col.max() <= p.ng
    # the sparse rows are exactly the non-zeros of the stacked constants
    dense = np.zeros_like(p.DY)
    for row, term, e0, e1 in segs:
        dense[term, row, p.Y_col[e0:e1]] = p.Y_val[e0:e1]
    assert np.array_equal(dense, p.DY)


def test_reference_test_body_case(cpu_extend):
    g = cases.golden("test_body")
    body = cases.body_test_formulation(NumpyBackend())
    got = body.generate_all_qp_matrices(g["given"])
    cases.assert_qp_close(got, (g["A"], g["h"], g["Q"], g["q"]))
    cQ, cq = body.generate_qp_cost(body.goals["crossed"], g["given"])
    assert cases.rel_err(cQ, g["cQ"]) < 1e-12 and cases.rel_err(cq, g["cq"]) < 1e-12
    assert np.abs(cQ - cQ.T).max() > 0  # crossed Hessians are not symmetrised (SURVEY B#3)
    limit = body.constraints["steppingArea"][0]
    A1, h1 = body.generate_qp_constraint(limit, g["given"])
    assert cases.rel_err(A1, g["A1"]) < 1e-12 and cases.rel_err(h1, g["h1"]) < 1e-12
    # the identity the reference's test asserts (test_body.py:138-142)
    kin = body.constraints["kinematics"][0]
    A0, h0 = body.generate_qp_constraint(kin, g["given"])
    assert h0.shape == (9, 1) and A0.shape == (9, body.optim_len)
    A2, h2 = body.generate_all_qp_constraints(g["given"])
    assert A2.shape == (54, body.optim_len) and h2.shape == (54, 1)
    Q, q = body.generate_all_qp_costs(g["given"])
    assert cases.rel_err(Q, g["Q"]) < 1e-12 and cases.rel_err(q, g["q"]) < 1e-12


@pytest.mark.parametrize("name", ["generic_small", "generic_c4_shape"])
def test_generic_batched(cpu_extend, name):
    g = cases.golden(name)
    form, params, given = cases.generic_batched(g, NumpyBackend())
    out = form.generate_all_qp_matrices_batched(given, params)
    cases.assert_qp_close((out.G, out.h, out.P, out.q), (g["A"], g["h"], g["Q"], g["q"]))
    # shared dynamics: P and G do not depend on the instance -> one constant term
    assert form._compiled().P_terms == [0] and form._compiled().G_terms == [0]


def test_biped_ticks_ragged_sizes(cpu_extend):
    """Formulation.update changes nv / nc between ticks."""
    g = cases.golden("biped_ticks")
    conf = cases.default_configuration(24)
    form = cases.formulate_biped(conf, backend=NumpyBackend())
    sizes = set()
    for tick in (0, 11, 12, 13):
        form.update(step_times=g["t%d_step_times" % tick], step_count=int(g["t%d_step_count" % tick]))
        got = form.generate_all_qp_matrices(g["t%d_given" % tick])
        ref = tuple(g["t%d_%s" % (tick, k)] for k in "AhQq")
        assert got[0].shape == ref[0].shape
        cases.assert_qp_close(got, ref)
        sizes.add(got[0].shape)
    assert len(sizes) > 1


def test_update_invalidates_plan(cpu_extend):
    form = cases.formulate_biped(cases.default_configuration(), backend=NumpyBackend())
    given = np.zeros([form.given_len, 1])
    q0 = form.generate_all_qp_matrices(given)[3]
    form.goals["track vel_x"].update(aim=0.1)
    q1 = form.generate_all_qp_matrices(given)[3]
    assert np.abs(q0 - q1).max() > 0


def test_per_instance_weight_and_errors(cpu_extend):
    form = cases.formulate_biped(cases.default_configuration(), backend=NumpyBackend())
    form.incorporate_goal("extra", Cost("CoM", Param("w_extra"), aim=[0.0, 0.0], axes=["_x", "_y"]))
    given = np.zeros([2, form.given_len])
    out = form.generate_all_qp_matrices_batched(given, {"w_extra": np.array([0.0, 2.0])})
    base = cases.formulate_biped(cases.default_configuration(), backend=NumpyBackend())
    ref = base.generate_all_qp_matrices_batched(given[:1])
    assert cases.rel_err(out.P[0], ref.P[0]) < 1e-14
    assert np.abs(out.P[1] - ref.P[0]).max() > 0
    with pytest.raises(KeyError):
        form.generate_all_qp_matrices_batched(given, {})
    with pytest.raises(ValueError):  # multi-row aims do not work in the reference either
        bad = cases.formulate_biped(cases.default_configuration(), backend=NumpyBackend())
        bad.goals["track vel_x"].update(aim=np.ones([3, 1]))
        bad.generate_all_qp_matrices(np.zeros([bad.given_len, 1]))


# ------------------------------------------------- per-instance constraint geometry
@pytest.mark.parametrize("name", ["biped_n24_yaw", "biped_n100_yaw"])
def test_biped_per_instance_yaw(cpu_extend, name):
    """Box.rotate_in_TS with a Param angle: G becomes per-instance through the
    monomials cos(yaw), sin(yaw) (x the omega ones), checked against the
    reference's rotate_in_TS run instance by instance."""
    g = cases.golden(name)
    form, params = cases.biped_yaw(g, NumpyBackend())
    cases.check_layout(form, g)
    out = form.generate_all_qp_matrices_batched(g["given"], params)
    cases.assert_qp_close((out.G, out.h, out.P, out.q), (g["A"], g["h"], g["Q"], g["q"]))
    p = form._compiled()
    assert sorted(form.parameter_names()) == ["omega", "vel_x", "vel_y", "yaw"]
    # G: {1, -1/w^2, 1/w} for the straight boxes, {cos, sin} x {1, -1/w^2} and {cos, sin} x 1/w
    assert len(p.G_terms) <= planmod.MAX_TERMS and len(p.P_terms) == 3


def test_geometry_expressions_and_errors(cpu_extend):
    from mpc_interface_b200 import Box, Constraint, tools as use
    from mpc_interface_b200.params import cos, sin, split_array, evaluate_array

    yaw, d = Param("yaw"), Param("d")
    R = use.rotation2D(yaw)
    assert R.dtype == object and use.rotation3D(yaw, "y").dtype == object
    for axis in "xyz":  # numeric rotations are unchanged
        R3 = use.rotation3D(0.3, axis)
        assert np.allclose(R3 @ R3.T, np.eye(3)) and np.isclose(np.linalg.det(R3), 1)
    assert np.allclose(use.rotation3D(0.3, "y")[0, 2], np.sin(0.3))  # right-handed about y
    parts = dict(split_array(np.array([[1.0, 2.0]]) @ R.T))
    assert set(len(m) for m in parts) == {1} and len(parts) == 2
    # a polynomial keeps products of atoms together: (1 + d) * cos(yaw)
    poly = ((1 + d) * cos(yaw)).polynomial()
    assert sorted(len(m) for m in poly) == [1, 2]
    assert np.isclose(evaluate_array(np.array([(1 + d) * sin(yaw)]), {"d": 2.0, "yaw": 0.5})[0], 3 * np.sin(0.5))

    box = Box.task_space("CoM", use.make_simetric_vertices([0.1, 0.05]), ["_x", "_y"])
    box.rotate_in_TS(R)
    assert all(lim.arrow.dtype == object for lim in box.constraints)
    box.translate_in_TS(np.array([d, 0.0]))      # per-instance centre -> per-instance bound
    assert box.constraints[0].bound().dtype == object
    lim = Constraint("CoM_x", d)                  # a Param extreme is not sign-normalised
    assert lim.extreme.dtype == object and lim.arrow[0, 0] == 1.0
    with pytest.raises(TypeError):
        box.set_safety_margin(0.01)               # needs numeric arrows (norm of the arrow)


def test_per_instance_center_and_extreme(cpu_extend):
    """recenter / translate / scale with Params against the oracle (h only changes)."""
    from mpc_interface_b200.examples import synthetic
    from oracle import qp_oracle as orc

    N, batch = 24, 4
    form = cases.formulate_biped(cases.default_configuration(N), backend=NumpyBackend())
    box = form.constraint_boxes["support_polygon"]
    box.translate_in_TS(np.array([Param("cx"), Param("cy")]))
    box.scale_box(Param("scale"))
    vals = synthetic.biped_inputs(5, batch, N)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    rng = np.random.default_rng(9)
    params = {"cx": rng.normal(size=batch) * 0.01, "cy": rng.normal(size=batch) * 0.01,
              "scale": rng.uniform(0.8, 1.2, batch)}
    out = form.generate_all_qp_matrices_batched(given, params)
    for b in range(batch):
        one = {k: float(v[b]) for k, v in params.items()}
        ref = orc.all_qp_matrices(form, given[b][:, None], one)
        cases.assert_qp_close((out.G[b], out.h[b], out.P[b], out.q[b]), ref)
    assert form._compiled().G_terms == [0]        # G itself stays shared


def test_more_than_eight_streamed_terms(cpu_extend):
    """Three boxes with three per-instance angles: 10 G monomials (> 8, the generic
    stream kernel on the device), still within plan.MAX_TERMS; against the oracle."""
    from oracle import qp_oracle as orc

    N, batch = 24, 3
    form = cases.biped_three_yaws(N, NumpyBackend())
    given, params = cases.three_yaw_inputs(form, 12, batch, N)
    out = form.generate_all_qp_matrices_batched(given, params)
    p = form._compiled()
    assert 8 < len(p.G_terms) <= planmod.MAX_TERMS
    for b in range(batch):
        one = {k: float(v[b]) for k, v in params.items()}
        ref = orc.all_qp_matrices(form, given[b][:, None], one)
        cases.assert_qp_close((out.G[b], out.h[b], out.P[b], out.q[b]), ref)


# ------------------------------------------------------------- randomised formulations
@pytest.mark.parametrize("seed", range(12))
def test_random_formulations_against_reference(cpu_extend, seed):
    """tests/_random_forms.py: the reference built each instance with floats
    (golden random_forms), here one formulation with nine Params -- axes or not,
    matrix coefficients, nested definitions, a flipped limit, per-row geometry,
    schedules, L matrices, polygons, crossed costs, mixed optimisation domains."""
    g = cases.golden("random_forms")
    form, given, params, ref = cases.random_form(g, seed, NumpyBackend())
    out = form.generate_all_qp_matrices_batched(given, params)
    cases.assert_qp_close((out.G, out.h, out.P, out.q), ref)


@pytest.mark.parametrize("seed", range(640, 652))
def test_random_structures_around_a_shared_plant_against_the_live_reference(cpu_extend, seed):
    """tests/_embedded_forms.py in mode A (the compiled affine plan, evaluated by the NumPy
    backend): named plants, slack / footsteps / disturbance, scheduled, crossed and state-input
    costs, ``L`` on costs and limits; per-instance aims / bounds and weights.  Every instance
    formulated by the unmodified reference."""
    import _embedded_forms as forms
    from mpc_interface_b200 import Param
    from oracle import reference

    if not reference.available():
        pytest.skip("oracle/_ref is not staged")
    spec = forms.draw(seed, scheduled=True, general=True, with_L=True)
    omega, tau = 3.4, 0.1 * min(1.0, 30.0 / spec["N"])
    form = forms.build(spec, forms.own_api(), omega, tau, Param("vel"), Param("gain"), backend=NumpyBackend())
    batch = 3
    vals = forms.inputs(spec, form, batch, seed + 1000)
    given = np.zeros([batch, form.given_len])
    for name, ids in form.given_ID.items():
        given[:, ids.start:ids.stop] = vals[name]
    out = form.generate_all_qp_matrices_batched(given, {"vel": vals["vel"], "gain": vals["gain"]})
    api = forms.reference_api(reference.load())
    for b in range(batch):
        one = forms.build(spec, api, omega, tau, float(vals["vel"][b]), float(vals["gain"][b]))
        want = one.generate_all_qp_matrices(one.arrange_given(
            {name: np.asarray(vals[name][b]).reshape([-1, 1]) for name in one.given_ID}))
        cases.assert_qp_close((out.G[b], out.h[b], out.P[b], out.q[b]), want)


def test_box_transforms_per_instance_against_the_live_reference(cpu_extend):
    """rotate_in_TS, translate_in_TS, scale_box and recenter_in_TS
    (restrictions.py:380-486) with one value per robot -- the boxes' arrows, centres and extremes
    hold Param expressions, the plan splits them into monomials.  The reference applies the same
    calls with numbers to a formulation of its own per instance."""
    import types

    from mpc_interface_b200 import Param, tools
    from mpc_interface_b200.examples import biped, synthetic
    from oracle import reference

    if not reference.available():
        pytest.skip("oracle/_ref is not staged")
    ref = reference.load()
    api = types.SimpleNamespace(ControlSystem=ref.ControlSystem, ExtendedSystem=ref.ExtendedSystem,
                                DomainVariable=ref.DomainVariable, Formulation=ref.Formulation,
                                LineCombo=ref.LineCombo, Cost=ref.Cost, Box=ref.Box, tools=ref.tools)
    N, batch = 16, 4
    conf = biped.default_configuration(N)
    form = biped.formulate_biped(conf, omega=Param("omega"), target_vel=[Param("vel_x"), Param("vel_y")],
                                 backend=NumpyBackend())
    cases.transform_boxes(form, tools.rotation2D(Param("yaw")), (Param("dx"), Param("dy")), Param("sc"), Param("rc"))
    vals = synthetic.biped_inputs(3, batch, N)
    vals.update(cases.transform_values(batch, 4))
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y", "yaw", "dx", "dy", "sc", "rc")}
    out = form.generate_all_qp_matrices_batched(given, params)
    for b in range(batch):
        one = biped.formulate_biped(conf, omega=float(vals["omega"][b]),
                                    target_vel=[float(vals["vel_x"][b]), float(vals["vel_y"][b])], api=api)
        cases.transform_boxes(one, ref.tools.rotation2D(float(vals["yaw"][b])),
                              (float(vals["dx"][b]), float(vals["dy"][b])), float(vals["sc"][b]), float(vals["rc"][b]))
        want = one.generate_all_qp_matrices(one.arrange_given(
            {v: np.asarray(vals[v][b]).reshape([-1, 1]) for v in one.given_ID}))
        cases.assert_qp_close((out.G[b], out.h[b], out.P[b], out.q[b]), want)


def test_negative_per_instance_extreme_is_refused(cpu_extend):
    """The reference flips a limit with a negative extreme (restrictions.py:190-194);
    for a Param extreme the batch is checked and refused instead of differing."""
    from mpc_interface_b200 import Constraint

    form = cases.tiny_plant(NumpyBackend(), optim=("u",))
    form.incorporate_constraint("cap", Constraint("a", Param("e")))
    form.make_preview_matrices()
    given = np.zeros([2, form.given_len])
    values = {"w": np.ones(2), "aim": np.zeros(2)}
    form.generate_all_qp_matrices_batched(given, dict(values, e=np.array([0.5, 0.0])))
    with pytest.raises(ValueError, match="negative for instance 1"):
        form.generate_all_qp_matrices_batched(given, dict(values, e=np.array([0.5, -0.1])))


def test_plan_cache_is_keyed_on_content(cpu_extend):
    """Regression (round-1 advisor finding): plans were cached under (id(obj), stamp); a
    fresh Constraint reusing the address of a dead one got the stale plan, and a field
    changed by hand (no .update()) was ignored.  The reference re-reads the objects on
    every generate_* call (body.py:236-302)."""
    from mpc_interface_b200 import Constraint

    form = cases.tiny_plant(backend=NumpyBackend(), optim=("u",))
    given = np.zeros([form.given_len, 1])
    seen = []
    for extreme in (1.0, 2.0, 3.0):  # temporaries: CPython hands out the same id again
        A, h = form.generate_qp_constraint(Constraint("b", extreme), given)
        seen.append(float(h[0, 0]))
    assert seen == [1.0, 2.0, 3.0]
    # direct attribute mutation, no update()
    limit = form.constraints["lim"][0]
    _, h0 = form.generate_qp_constraint(limit, given)
    limit.extreme = limit.extreme * 2
    _, h1 = form.generate_qp_constraint(limit, given)
    assert np.allclose(h1, 2 * h0) and float(h0[0, 0]) == 3.0
    cost = form.goals["g2"]
    Q0, _ = form.generate_qp_cost(cost, given)
    cost.weight = 10 * cost.weight
    Q1, _ = form.generate_qp_cost(cost, given)
    assert np.allclose(Q1, 10 * Q0) and np.abs(Q0).max() > 0
    _, h_all = form.generate_all_qp_constraints(given)
    assert float(h_all[0, 0]) == 6.0  # the full plan sees the mutated limit as well
    # equal content -> one plan; the cache does not grow without bound
    n_before = len(form._plans)
    for _ in range(5):
        form.generate_qp_constraint(Constraint("b", 7.0), given)
    assert len(form._plans) == n_before + 1
    for k in range(50):
        form.generate_qp_constraint(Constraint("b", 10.0 + k), given)
    assert len(form._plans) <= 32


def test_box_rows_of_the_biped_are_the_stepping_area(cpu_extend):
    """plan.BoxRows (lb / ub extraction): at the biped the rows that bound
    one variable with an instance-independent coefficient are exactly the stepping area's
    (Box on Ds itself: 4 rows per step); the support polygon and the terminal box mix
    variables and carry omega."""
    from mpc_interface_b200 import Param, plan as _plan

    g = cases.golden("biped_n24_batch")
    form = cases.formulate_biped(cases.default_configuration(24), omega=Param("omega"),
                                 target_vel=[Param("vel_x"), Param("vel_y")], backend=NumpyBackend())
    box = _plan.BoxRows(form._compiled(), form._ops())
    k = form.domain["Ds_x"]
    assert list(box.rows) == list(range(4 * k)) and list(box.rest) == list(range(4 * k, form._compiled().nc))
    A = g["A"][0]  # the reference's G: those rows have ONE non-zero, the value BoxRows reports
    for r, c, a in zip(box.row, np.repeat(np.arange(form.optim_len), np.diff(box.col_ptr)), box.coef):
        assert np.count_nonzero(A[r]) == 1 and A[r, c] == a
    assert box.col_ptr[-1] == 4 * k and set(np.nonzero(np.diff(box.col_ptr))[0]) == \
        set(form.optim_ID["Ds_x"]) | set(form.optim_ID["Ds_y"])
