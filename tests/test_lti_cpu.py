"""Host lowering of per-instance-dynamics formulations (lti.py) against golden
vectors written by the unmodified reference (tests/golden/make_goldens.py
per_instance): the lowered stage tables, evaluated by a plain NumPy statement
of the kernel's formulas, must reproduce the reference's (A, h, Q, q)."""
import numpy as np
import pytest

import _cases as cases
import _lti_numpy
from _numpy_backend import NumpyBackend
from mpc_interface_b200 import Constraint, Cost, LineCombo, Param, lti


def _formulation(g, cpu_extend, per_instance):
    nx = g["Q_w"].shape[0]
    refs = [Param("ref%d" % i) for i in range(nx)]
    A0 = g["A_sys"][0] if per_instance else g["A_sys"]
    B0 = g["B_sys"][0] if per_instance else g["B_sys"]
    return cases.formulate_generic(A0, B0, g["Q_w"], g["R_w"], int(g["N"]), g["rows"], refs=refs,
                                   backend=NumpyBackend())


@pytest.mark.parametrize("name,per_instance", [
    ("generic_perinst_small", True), ("generic_perinst_odd", True), ("generic_perinst_mid", True),
    ("generic_small", False), ("generic_c4_shape", False)])
def test_lowering_matches_reference(cpu_extend, name, per_instance):
    g = cases.golden(name)
    form = _formulation(g, cpu_extend, per_instance)
    cases.check_layout(form, g)
    plan = form.lti_plan("plant")
    nx = g["Q_w"].shape[0]
    assert (plan.nv, plan.nc) == g["A"].shape[:0:-1]
    assert plan.row_offsets == [i * plan.N for i in range(len(plan.row_offsets))]
    names = [repr(e) for e in plan.slots.exprs]
    for b in range(g["x0"].shape[0]):
        A = g["A_sys"][b] if per_instance else g["A_sys"]
        B = g["B_sys"][b] if per_instance else g["B_sys"]
        scalars = np.array([g["ref"][b, int(nm[3:])] for nm in names])
        got = _lti_numpy.evaluate(plan, A, B, g["x0"][b], scalars)
        cases.assert_qp_close(got, (g["A"][b], g["h"][b], g["Q"][b], g["q"][b]))


def test_lowering_refuses_unsupported(cpu_extend):
    g = cases.golden("generic_small")
    form = _formulation(g, cpu_extend, False)
    form.incorporate_goal("sched", Cost("z0", 1.0, schedule=range(2, int(g["N"]) + 1)))
    with pytest.raises(IndexError):  # (the reference's fancy index fails the same way, body.py:283)
        form.lti_plan("plant")
    form = _formulation(g, cpu_extend, False)
    form.incorporate_goal("short L", Cost("z0", 1.0, L=np.eye(int(g["N"]) - 1)))
    with pytest.raises(ValueError, match="matmul"):  # (as the reference's L @ Mo[schedule], body.py:280)
        form.lti_plan("plant")
    form = _formulation(g, cpu_extend, False)
    form.incorporate_constraint("Param in L", Constraint("z1", 3.0, L=np.array(
        [[Param("ref0")] * int(g["N"])], dtype=object)))
    with pytest.raises((NotImplementedError, TypeError, ValueError)):
        form.lti_plan("plant")
    form = _formulation(g, cpu_extend, False)
    with pytest.raises(KeyError):
        form.generate_all_qp_matrices_batched(g["x0"], {}, dynamics={"nope": (g["A_sys"], g["B_sys"])})


def test_lowering_input_forms_and_derived_definitions(cpu_extend):
    """A derived point-wise definition of states in a limit and a cost on an input."""
    g = cases.golden("generic_small")
    form = _formulation(g, cpu_extend, False)
    form.incorporate_definition("sum01", LineCombo({"z0": 1.0, "z1": -2.0}))
    form.incorporate_definition("deep", LineCombo({"sum01": 0.5, "z2": 1.0}))
    form.incorporate_constraint("deep", Constraint("deep", 7.0))
    form.incorporate_goal("u_aim", Cost("u1", 0.3, aim=[Param("ua")]))
    form.make_preview_matrices()
    plan = form.lti_plan("plant")
    from oracle import qp_oracle as orc

    nx = 4
    names = [repr(e) for e in plan.slots.exprs]
    values = {"ref%d" % i: float(g["ref"][0, i]) for i in range(nx)}
    values["ua"] = 0.7
    scalars = np.array([values[nm] for nm in names])
    got = _lti_numpy.evaluate(plan, g["A_sys"], g["B_sys"], g["x0"][0], scalars)
    ref = orc.all_qp_matrices(form, g["x0"][0][:, None], values)
    cases.assert_qp_close(got, ref)


def _lip(g, cpu_extend):
    from mpc_interface_b200.examples import lip

    conf = lip.default_configuration(int(g["N"]))
    form = lip.formulate_lip(conf, omega=Param("omega"), tau=Param("tau"),
                             target_vel=[Param("vel_x"), Param("vel_y")], backend=NumpyBackend())
    return form


@pytest.mark.parametrize("name", ["lip_perinst", "lip_perinst_n9"])
def test_lip_two_axes_param_coefficients_terminal_schedule(cpu_extend, name):
    """Mode B widened (round 2): the plant on two axes (block-diagonal lowering), Param
    coefficients in the definitions (b = CoM - CoM_ddot / omega^2, DCM = CoM + CoM_dot /
    omega), multi-axis box facets, a scheduled terminal limit, multi-axis costs -- against
    the reference, which formulated every instance with its own
    ControlSystem.from_name("dP->CCC", tau_b, omega_b)."""
    g = cases.golden(name)
    form = _lip(g, cpu_extend)
    cases.check_layout(form, g)
    plan = form.lti_plan("LIP")
    N = int(g["N"])
    assert plan.per_instance and (plan.n, plan.m, plan.n0, plan.m0) == (6, 2, 3, 1)
    assert (plan.nv, plan.nc) == (2 * N, 4 * N + 4) and plan.limit_rows == [N] * 4 + [1] * 4
    batch = g["A"].shape[0]
    params = {k: g[k] for k in ("omega", "tau", "vel_x", "vel_y")}
    assert sorted(plan.parameter_names()) == ["omega", "vel_x", "vel_y"]  # tau lives in the plant
    tables = plan.instance_tables(params, batch)
    given = np.hstack([g["x0_x"], g["x0_y"]])
    for b in range(batch):
        one = {k: v[b] for k, v in tables.items()}
        got = _lti_numpy.evaluate(plan, g["A_sys"][b], g["B_sys"][b], given[b], None, one)
        cases.assert_qp_close(got, (g["A"][b], g["h"][b], g["Q"][b], g["q"][b]))
    # the cross blocks of P between the axes are exact zeros, as in the reference
    assert np.all(g["Q"][0][:N, N:] == 0.0) and np.all(got[2][:N, N:] == 0.0)


def test_named_systems_closed_forms_match_sympy():
    """the closed forms the device kernel evaluates (restated here in
    NumPy, the kernel itself is checked on the GPU) against the SymPy-derived lambdas of
    tools.get_system_matrices for every named system."""
    from mpc_interface_b200 import tools as use

    T, w = 0.1, 3.3445
    c, s = np.cosh(w * T), np.sinh(w * T)
    closed = {
        "J->CCC": ([[1, T, T * T / 2], [0, 1, T], [0, 0, 1]], [T**3 / 6, T * T / 2, T]),
        "P->CC": ([[c, s / w], [w * s, c]], [1 - c, -w * s]),
        "P->X": ([[np.exp(w * T)]], [1 - np.exp(w * T)]),
        "dP->CCC": ([[1, s / w, (c - 1) / w**2], [0, c, s / w], [0, w * s, c]], [T - s / w, 1 - c, -w * s]),
        "dP->CCP": ([[c, s / w, 1 - c], [w * s, c, -w * s], [0, 0, 1]], [T - s / w, 1 - c, T]),
    }
    for name, (A, B) in closed.items():
        get_A, get_B, _ = use.get_system_matrices(name)
        assert np.allclose(np.array(get_A(tau=T, omega=w), dtype=float), np.array(A, dtype=float), rtol=1e-13, atol=1e-15), name
        assert np.allclose(np.array(get_B(tau=T, omega=w), dtype=float).ravel(), np.array(B, dtype=float), rtol=1e-13, atol=1e-15), name


def test_lowering_permuted_inputs_and_scheduled_limit(cpu_extend):
    """The optimisation variables in another order than the plant's inputs, and a limit on
    a window of the horizon."""
    g = cases.golden("generic_small")
    form = _formulation(g, cpu_extend, False)
    form.incorporate_constraint("late", Constraint("z2", 1.5, schedule=range(3, 7)))
    form.identify_qp_domain(["u1", "u0"])
    form.make_preview_matrices()
    plan = form.lti_plan("plant")
    assert plan.inputs == [(0, 1), (0, 0)] and 4 in plan.limit_rows and plan.nc == sum(plan.limit_rows)
    from oracle import qp_oracle as orc

    names = [repr(e) for e in plan.slots.exprs]
    values = {"ref%d" % i: float(g["ref"][0, i]) for i in range(4)}
    scalars = np.array([values[nm] for nm in names])
    got = _lti_numpy.evaluate(plan, g["A_sys"], g["B_sys"], g["x0"][0], scalars)
    ref = orc.all_qp_matrices(form, g["x0"][0][:, None], values)
    cases.assert_qp_close(got, ref)
    assert got[0].shape[0] == ref[0].shape[0]


# ---------------------------------------------------------------------------
# the walking biped with one plant per instance (embedded layout + coupling stage)
# ---------------------------------------------------------------------------
def _biped_plants(g, system):
    from mpc_interface_b200.examples import biped

    conf = biped.default_configuration(int(g["N"]))
    conf.system = system
    form = biped.formulate_biped(conf, omega=Param("omega"), tau=Param("tau"),
                                 target_vel=[Param("vel_x"), Param("vel_y")], backend=NumpyBackend())
    if "tick_step_times" in g:
        form.update(step_times=g["tick_step_times"], step_count=int(g["tick_step_count"]))
    return form


@pytest.mark.parametrize("name,system", [("biped_plants_n24", "dP->CCC"), ("biped_plants_n30_tick", "dP->CCC"),
                                         ("biped_plants_jccc_n16", "J->CCC")])
def test_biped_with_per_instance_plants_lowering(cpu_extend, name, system):
    """the biped's whole structure (steps system, bias through a matrix
    coefficient, boxes on (b+n-s), Ds and (DCM-s)) lowered around a per-instance plant; the
    tables of lti.lower -- fused-kernel part in the embedded layout plus the coupling stage,
    both restated in NumPy (tests/_lti_numpy.py) -- against the reference, which formulated
    every instance with its own ControlSystem.from_name(system, tau_b, omega_b)."""
    from mpc_interface_b200.examples import synthetic

    g = cases.golden(name)
    form = _biped_plants(g, system)
    cases.check_layout(form, g)
    plan = form.lti_plan("LIP")
    N, batch = int(g["N"]), g["A"].shape[0]
    k = form.domain["Ds_x"]
    assert plan.embedded and plan.per_instance and (plan.n, plan.m, plan.nz) == (6, 2, 2 * k)
    assert (plan.nv, plan.nc) == g["A"].shape[:0:-1] == (2 * N + 2 * k, 4 * k + 4 * N + 4)
    assert plan.var_off == [0, N + k] and plan.z_col == list(range(N, N + k)) + list(range(2 * N + k, 2 * N + 2 * k))
    assert plan.varying == {"Qxx", "cx"}  # omega reaches the CoP definitions only
    assert list(plan.row_kind[:4 * k]) == [1] * (4 * k) and not plan.row_kind[4 * k:].any()
    params = {key: g[key] for key in ("omega", "tau", "vel_x", "vel_y")}
    assert sorted(plan.parameter_names()) == ["omega", "vel_x", "vel_y"]  # tau lives in the plant
    tables = plan.instance_tables(params, batch)
    ccx, ccu = plan.coupled_tables(params, batch)
    given = synthetic.pack_given(form.given_ID, form.given_len, {key: g[key] for key in form.given_ID}, batch)
    for b in range(batch):
        one = {key: v[b] for key, v in tables.items()}
        got = _lti_numpy.evaluate_embedded(plan, g["A_sys"][b], g["B_sys"][b], given[b], None, one, (ccx[b], ccu[b]))
        cases.assert_qp_close(got, (g["A"][b], g["h"][b], g["Q"][b], g["q"][b]))


@pytest.mark.parametrize("name", ["soft_lip_n12", "soft_lip_n17_slack_first"])
def test_soft_constrained_lip_lowering(cpu_extend, name):
    """The embedded layout away from the biped (examples/soft_lip.py): an optimised slack, a
    disturbance through a shift matrix, a cost with a NON-ZERO aim whose signal has a plant part
    and an auxiliary part (its aim must reach q[u] and q[z] once each), row-wise arrows, rows
    of the slack alone; the plant's input first or second in the QP."""
    from mpc_interface_b200.examples import soft_lip, synthetic

    g = cases.golden(name)
    N, batch, order = int(g["N"]), g["A"].shape[0], tuple(str(x) for x in g["order"])
    form = soft_lip.formulate_soft_lip(N, Param("omega"), Param("tau"), Param("vel"), order=order,
                                       backend=NumpyBackend())
    cases.check_layout(form, g)
    plan = form.lti_plan("LIP")
    assert plan.embedded and plan.per_instance and (plan.n, plan.m, plan.nz, plan.nv) == (3, 1, N, 2 * N)
    first = order[0] == "cCoP_dot_x"
    assert plan.var_off == [0 if first else N] and plan.z_col == list(range(N, 2 * N) if first else range(N))
    assert len(plan.coupled) == 2 and plan.varying == {"Qxx", "cx"}
    params = {key: g[key] for key in ("omega", "vel")}
    tables = plan.instance_tables(params, batch)
    ccx, ccu = plan.coupled_tables(params, batch)
    given = synthetic.pack_given(form.given_ID, form.given_len, {key: g[key] for key in form.given_ID}, batch)
    for b in range(batch):
        one = {key: v[b] for key, v in tables.items()}
        got = _lti_numpy.evaluate_embedded(plan, g["A_sys"][b], g["B_sys"][b], given[b], None, one, (ccx[b], ccu[b]))
        cases.assert_qp_close(got, (g["A"][b], g["h"][b], g["Q"][b], g["q"][b]))


@pytest.mark.parametrize("seed,scheduled,general,turned", [(s, False, False, False) for s in range(200, 212)]
                         + [(s, True, False, False) for s in range(220, 232)]
                         + [(s, True, True, False) for s in range(240, 256)]
                         + [(s, True, True, True) for s in range(260, 272)]
                         + [(s, True, True, "L") for s in range(280, 290)]
                         + [(s, True, True, "LD") for s in range(300, 310)]
                         + [(s, True, True, "LDP") for s in range(320, 332)]
                         + [(s, True, True, "rows") for s in range(340, 350)])
def test_random_embedded_formulations_lowering(cpu_extend, seed, scheduled, general, turned):
    """tests/_embedded_forms.py: random structures around a per-instance plant (four named
    plants, one or two axes, slack / footsteps / disturbance, definitions of definitions, costs
    with aims and on signals shorter than the horizon, facets, row-wise arrows, centres,
    schedules, any variable order), lowered and evaluated by the NumPy restatement of the
    kernels' arithmetic, against the unmodified reference formulating every instance itself.
    ``scheduled``: some costs act on a few samples only (terminal, a window, every 2nd / 3rd
    sample; per-instance weights and aims) -- the stage of csrc/mpci_sched.cu.  ``general``: also
    crossed costs between any two signals, per-instance weights / aims on signals with a part outside
    the plant and a definition that mixes states and the input (whole-horizon terms of that stage).
    ``turned``: limits turned / scaled per instance, on signals with parts outside the plant too
    (per-instance Gz and h map of the coupling stage); "L": also costs and limits with a matrix ``L``
    over their samples; "LD": also definitions with a matrix coefficient on a plant signal (a rate
    ``(I - shift) @ u1``, a delayed state), whose costs and limits go row by row through mpci_sched.cu;
    "LDP": also per-instance coefficients (1 / omega, omega) on the variables outside the plant;
    "rows": also arrows that turn from row to row per instance."""
    import _embedded_forms as forms
    from oracle import reference

    if not reference.available():
        pytest.skip("oracle/_ref is not staged")
    from mpc_interface_b200 import dynamics

    spec = forms.draw(seed, scheduled, general, with_L=turned in ("L", "LD", "LDP", "rows"),
                      turned="rows" if turned == "rows" else bool(turned),
                      delayed=turned in ("LD", "LDP", "rows"), paramaux=turned in ("LDP", "rows"))
    form = forms.build(spec, forms.own_api(), Param("omega"), Param("tau"), Param("vel"), Param("gain"),
                       backend=NumpyBackend())
    plan = form.lti_plan("LIP")
    assert plan.embedded
    if not general:
        assert len(plan.sched_terms) == sum(len(spec["axes"]) for c in spec["costs"] if c["schedule"])
    batch = 2
    vals = forms.inputs(spec, form, batch, seed + 1000)
    given = np.zeros([batch, form.given_len])
    for name, ids in form.given_ID.items():
        given[:, ids.start:ids.stop] = vals[name]
    params = {key: vals[key] for key in plan.parameter_names()}
    tables = plan.instance_tables(params, batch)
    ccx, ccu = plan.coupled_tables(params, batch)
    api = forms.reference_api(reference.load())
    for b in range(batch):
        one = forms.build(spec, api, *(float(vals[key][b]) for key in ("omega", "tau", "vel", "gain")))
        assert list(one.optim_ID.items()) == list(form.optim_ID.items())
        assert list(one.given_ID.items()) == list(form.given_ID.items())
        want = one.generate_all_qp_matrices(one.arrange_given(
            {name: np.asarray(vals[name][b]).reshape([-1, 1]) for name in one.given_ID}))
        plant = dynamics.ControlSystem.from_name(spec["system"], ["_x"], tau=float(vals["tau"][b]),
                                                 omega=float(vals["omega"][b]))
        extras = tuple(None if t is None else t[b] for t in plan.coupled_extras(params, batch))
        got = _lti_numpy.evaluate_embedded(plan, plant.A, plant.B, given[b], None,
                                           {key: v[b] for key, v in tables.items()}, (ccx[b], ccu[b]), extras)
        if plan.sched_terms:
            terms = plan.sched_tables(params, batch)
            got = got[:2] + _lti_numpy.add_scheduled(plan, plant.A, plant.B, given[b],
                                                     terms[b if terms.shape[0] > 1 else 0], got[2], got[3])
        if plan.lrows:
            rows = plan.lrow_tables(params, batch)
            got = _lti_numpy.add_limit_rows(plan, plant.A, plant.B, given[b], rows[b if rows.shape[0] > 1 else 0],
                                            got[0], got[1]) + tuple(got[2:])
        cases.assert_qp_close(got, want)


@pytest.mark.parametrize("per_instance", [False, True])
def test_generic_mpc_with_a_terminal_cost_lowering(cpu_extend, per_instance):
    """The textbook terminal cost (x_N - ref)' Qf (x_N - ref) on the generic per-instance MPC:
    nx^2 crossed costs with schedule = range(N - 1, N) (examples/generic.py), in the plain layout;
    with ``Param`` references and, in the per-instance variant, a ``Param`` weight on one more
    scheduled cost.  Against the oracle's restatement of body.py:266-302."""
    from oracle import qp_oracle

    rng = np.random.default_rng(77)
    nx, nu, N, batch = 4, 2, 9, 3
    A = np.eye(nx) + 0.1 * rng.normal(size=[batch, nx, nx])
    B = 0.3 * rng.normal(size=[batch, nx, nu])
    M = rng.normal(size=[nx, nx])
    Q, R, Qf = M @ M.T / nx + np.eye(nx), np.eye(nu), 3.0 * (M.T @ M) / nx + np.eye(nx)
    refs = [Param("ref%d" % i) for i in range(nx)]
    form = cases.formulate_generic(A[0], B[0], Q, R, N, rng.normal(size=[2, nx]), refs=refs,
                                   backend=NumpyBackend(), terminal=Qf)
    if per_instance:
        form.incorporate_goal("window", Cost("z1", Param("gain"), aim=[Param("ref0") * 2], schedule=range(2, 7, 2)))
    plan = form.lti_plan("plant")
    assert not plan.embedded and len(plan.sched_terms) == nx * nx + (1 if per_instance else 0)
    # the nx^2 crossed costs share their sample: nx merged terms (T_N' Qf T_N), the window stays one
    assert [e[0] for e in plan.sched_layout()].count("merged") == nx
    assert len(plan.sched_layout()) == nx + (1 if per_instance else 0)
    x0 = rng.normal(size=[batch, nx])
    params = {"ref%d" % i: rng.normal(size=batch) for i in range(nx)}
    params["gain"] = rng.uniform(0.5, 2.0, batch)
    names = [repr(e) for e in plan.slots.exprs]
    terms = plan.sched_tables(params, batch)
    assert terms.shape[0] == batch
    for b in range(batch):
        one = {k: float(v[b]) for k, v in params.items()}
        scalars = np.array([one[nm] for nm in names])
        G, h, P, q = _lti_numpy.evaluate(plan, A[b], B[b], x0[b], scalars)
        P, q = _lti_numpy.add_scheduled(plan, A[b], B[b], x0[b], terms[b], P, q)
        single = cases.formulate_generic(A[b], B[b], Q, R, N, form_rows(form), refs=refs,
                                         backend=NumpyBackend(), terminal=Qf)
        if per_instance:
            single.incorporate_goal("window", Cost("z1", Param("gain"), aim=[Param("ref0") * 2], schedule=range(2, 7, 2)))
        want = qp_oracle.all_qp_matrices(single, x0[b][:, None], one)
        cases.assert_qp_close((G, h, P, q), want)


def form_rows(form):
    """The linear rows a generic formulation was built with (its ``lin<r>`` definitions)."""
    rows = []
    r = 0
    while "lin%d" % r in form.definitions:
        combo = form.definitions["lin%d" % r]
        rows.append([float(combo.matrices[i]) for i in range(len(combo.variables))])
        r += 1
    return np.array(rows)


def test_per_instance_tables_built_with_torch_equal_the_numpy_ones(cpu_extend):
    """The GPU path evaluates the Param expressions on torch tensors (lti._Device); on the CPU
    device the same code must give the NumPy tables bit for bit."""
    import torch

    g = cases.golden("biped_plants_n24")
    plan = _biped_plants(g, "dP->CCC").lti_plan("LIP")
    batch = g["A"].shape[0]
    params = {key: g[key] for key in ("omega", "tau", "vel_x", "vel_y")}
    host = plan.instance_tables(params, batch)
    there = plan.instance_tables(params, batch, torch.device("cpu"))
    assert sorted(host) == sorted(there) == ["Qxx", "cx", "tu", "tx"]
    for key in host:
        assert np.array_equal(host[key], there[key].numpy()), key
    for a, b in zip(plan.coupled_tables(params, batch), plan.coupled_tables(params, batch, torch.device("cpu"))):
        assert np.array_equal(a, b.numpy())


def test_embedded_lowering_refuses_what_the_stages_cannot_do(cpu_extend):
    from mpc_interface_b200.examples import biped

    conf = biped.default_configuration(16)
    form = biped.formulate_biped(conf, backend=NumpyBackend())
    form.incorporate_definition("odd_x", LineCombo({"CoM_x": 1.0, "Ds_x": Param("k")}))
    form.incorporate_goal("odd", Cost("odd", 1.0, axes=["_x"]))
    with pytest.raises(ValueError, match="broadcast"):  # (N samples + fewer footsteps: the reference fails too)
        form.lti_plan("LIP")
    form = biped.formulate_biped(conf, backend=NumpyBackend())
    form.incorporate_definition("rate_x", LineCombo({"CoM_x": np.eye(16) - np.eye(16, k=-1)}))
    form.incorporate_definition("scaled_x", LineCombo({"rate_x": Param("k")}))
    form.incorporate_constraint("short arrows", Constraint("scaled", 1.0, axes=["_x"], arrow=np.array(
        [[Param("k") * (1 + r)] for r in range(5)], dtype=object)))
    with pytest.raises(ValueError, match="broadcast"):  # (5 arrows for 16 samples: the reference fails too)
        form.lti_plan("LIP")


def test_rate_limits_and_rate_costs_on_plant_signals(cpu_extend):
    """A matrix coefficient ON a plant signal -- the CoP rate ``(I - shift) @ b`` of a jerk-driven
    LIP, limited and penalised: such a signal has no stage form, its cost becomes crossed terms of
    the scheduled-cost stage over the pairs of its parts, its limit rows of mpci_lti_sched_rows.
    With the biped's footsteps in the same signal; against the oracle."""
    from mpc_interface_b200 import dynamics
    from mpc_interface_b200.examples import biped, synthetic
    from oracle import qp_oracle

    N, batch = 12, 3
    conf = biped.default_configuration(N)
    form = biped.formulate_biped(conf, omega=Param("omega"), backend=NumpyBackend())
    D = np.eye(N) - np.eye(N, k=-1)
    for ax in ("_x", "_y"):
        form.incorporate_definition("raw" + ax, LineCombo({"b" + ax: D, "s" + ax: -0.5}))
        form.incorporate_definition("db" + ax, LineCombo({"raw" + ax: 1 / Param("omega")}))  # (per-instance scaling)
    form.incorporate_goal("CoP rate", Cost("db", 0.3, aim=[0.01, 0.0], axes=["_x", "_y"]))
    form.incorporate_constraint("CoP rate limit", Constraint("db", 0.2, axes=["_x"]))
    form.incorporate_constraint("late rate", Constraint("db", 0.1, axes=["_y"], arrow=[[-1.0]], schedule=range(N - 3, N)))
    form.make_preview_matrices()
    plan = form.lti_plan("LIP")
    # parts of db: (empty direct part), (I - shift) @ b / omega, -0.5 s / omega: 3 x 3 crossed terms per axis
    assert plan.embedded and len(plan.lrows) == 2 and len(plan.sched_terms) == 2 * 9
    vals = synthetic.biped_inputs(8, batch, N)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    params = {"omega": vals["omega"]}
    tables = plan.instance_tables(params, batch)
    ccx, ccu = plan.coupled_tables(params, batch)
    terms, rows = plan.sched_tables(params, batch), plan.lrow_tables(params, batch)
    plant = dynamics.ControlSystem.from_name(conf.system, ["_x"], tau=conf.mpc_period, omega=conf.omega)
    for b in range(batch):
        one = {"omega": float(vals["omega"][b])}
        extras = tuple(None if t is None else t[b] for t in plan.coupled_extras(params, batch))
        G, h, P, q = _lti_numpy.evaluate_embedded(plan, plant.A, plant.B, given[b], np.zeros([0]),
                                                   {k: v[b] for k, v in tables.items()}, (ccx[b], ccu[b]), extras)
        P, q = _lti_numpy.add_scheduled(plan, plant.A, plant.B, given[b], terms[b if terms.shape[0] > 1 else 0], P, q)
        G, h = _lti_numpy.add_limit_rows(plan, plant.A, plant.B, given[b], rows[b if rows.shape[0] > 1 else 0], G, h)
        cases.assert_qp_close((G, h, P, q), qp_oracle.all_qp_matrices(form, given[b][:, None], one))


def test_costs_beyond_the_stage_weights_take_the_general_stage(cpu_extend):
    """What T' Qxx T + Quu x I and the coupling stage's constant tables cannot express is lowered
    to whole-horizon terms of the scheduled-cost stage instead of being refused: a cost that
    couples states and inputs (generic plant), a per-instance weight and aim on a signal with a
    footstep part and a crossed cost on it (biped).  Against the oracle's body.py:266-302."""
    from mpc_interface_b200.examples import biped, synthetic
    from oracle import qp_oracle

    g = cases.golden("generic_small")
    form = _formulation(g, cpu_extend, False)
    form.incorporate_definition("mix", LineCombo({"z0": 1.0, "u0": 2.0}))
    form.incorporate_goal("mixed", Cost("mix", 1.5, aim=[Param("ref0")]))
    form.make_preview_matrices()
    plan = form.lti_plan("plant")
    N = int(g["N"])
    assert len(plan.sched_terms) == 1 and plan.sched_terms[0][5] == list(range(N))
    names = [repr(e) for e in plan.slots.exprs]
    batch = g["x0"].shape[0]
    params = {"ref%d" % i: g["ref"][:, i] for i in range(g["ref"].shape[1])}
    terms = plan.sched_tables(params, batch)
    for b in range(batch):
        one = {k: float(v[b]) for k, v in params.items()}
        G, h, P, q = _lti_numpy.evaluate(plan, g["A_sys"], g["B_sys"], g["x0"][b], np.array([one[nm] for nm in names]))
        P, q = _lti_numpy.add_scheduled(plan, g["A_sys"], g["B_sys"], g["x0"][b], terms[b if terms.shape[0] > 1 else 0], P, q)
        cases.assert_qp_close((G, h, P, q), qp_oracle.all_qp_matrices(form, g["x0"][b][:, None], one))

    N, batch = 16, 3
    conf = biped.default_configuration(N)
    form = biped.formulate_biped(conf, backend=NumpyBackend())
    form.incorporate_goal("steps", Cost("(c-s)", Param("w_steps"), aim=[Param("lead"), 0], axes=["_x", "_y"]))
    form.incorporate_goal("crossed", Cost("(c-s)", 0.7, aim=[0.05], cross="CoM", cross_aim=[0.1], axes=["_x"]))
    plan = form.lti_plan("LIP")
    # the per-instance weight / aim stay with the coupling stage (per-instance cost_w, cost_aim, Pzz);
    # the crossed cost is one whole-horizon term of the general stage
    assert plan.embedded and len(plan.sched_terms) == 1 and plan.per_instance
    assert [a is not None for a in plan.coupled_aims[-2:]] == [True, False] and len(plan.Pzz_terms) == 2
    vals = synthetic.biped_inputs(5, batch, N)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    params = {"w_steps": np.array([0.5, 1.0, 2.0]), "lead": np.array([0.0, 0.1, -0.2])}
    terms = plan.sched_tables(params, batch)
    from mpc_interface_b200 import dynamics
    plant = dynamics.ControlSystem.from_name(conf.system, ["_x"], tau=conf.mpc_period, omega=conf.omega)
    for b in range(batch):
        one = {k: float(v[b]) for k, v in params.items()}
        tables = {key: v[b] for key, v in plan.instance_tables(params, batch).items()}
        ccx, ccu = plan.coupled_tables(params, batch)
        extras = tuple(None if t is None else t[b] for t in plan.coupled_extras(params, batch))
        got = _lti_numpy.evaluate_embedded(plan, plant.A, plant.B, given[b], np.zeros([0]), tables,
                                           (ccx[b], ccu[b]), extras)
        P, q = _lti_numpy.add_scheduled(plan, plant.A, plant.B, given[b], terms[b if terms.shape[0] > 1 else 0],
                                        got[2], got[3])
        cases.assert_qp_close((got[0], got[1], P, q), qp_oracle.all_qp_matrices(form, given[b][:, None], one))


def test_lip_per_instance_plants_and_yaw_lowering(cpu_extend):
    """A box turned per instance (rotate_in_TS by a Param) on plant signals: the arrow's
    expressions join the limit's form (per-instance cx).  NumPy restatement of the kernel against
    the live reference."""
    import types

    from mpc_interface_b200 import dynamics
    from mpc_interface_b200.examples import lip
    from oracle import reference

    if not reference.available():
        pytest.skip("oracle/_ref is not staged")
    ref = reference.load()
    api = types.SimpleNamespace(ControlSystem=ref.ControlSystem, ExtendedSystem=ref.ExtendedSystem,
                                Formulation=ref.Formulation, LineCombo=ref.LineCombo, Cost=ref.Cost, Box=ref.Box,
                                make_simetric_vertices=ref.tools.make_simetric_vertices,
                                rotation2D=ref.tools.rotation2D)
    conf = lip.default_configuration(11)
    form = lip.formulate_lip(conf, omega=Param("omega"), tau=Param("tau"),
                             target_vel=[Param("vel_x"), Param("vel_y")], yaw=Param("yaw"), backend=NumpyBackend())
    plan = form.lti_plan("LIP")
    assert not plan.embedded and "cx" in plan.varying and "yaw" in plan.parameter_names()
    batch = 3
    vals = lip.lip_inputs(5, batch)
    vals["yaw"] = np.array([0.4, -2.0, 3.0])
    given = np.hstack([vals["x0_x"], vals["x0_y"]])
    tables = plan.instance_tables({k: vals[k] for k in plan.parameter_names()}, batch)
    for b in range(batch):
        one = lip.formulate_lip(conf, omega=float(vals["omega"][b]), tau=float(vals["tau"][b]),
                                target_vel=[float(vals["vel_x"][b]), float(vals["vel_y"][b])],
                                yaw=float(vals["yaw"][b]), api=api)
        want = one.generate_all_qp_matrices(one.arrange_given(
            {v: np.asarray(vals[v][b]).reshape([-1, 1]) for v in one.given_ID}))
        plant = dynamics.ControlSystem.from_name("dP->CCC", ["_x"], tau=float(vals["tau"][b]),
                                                 omega=float(vals["omega"][b]))
        got = _lti_numpy.evaluate(plan, plant.A, plant.B, given[b], np.zeros([0]), {k: v[b] for k, v in tables.items()})
        cases.assert_qp_close(got, want)


@pytest.mark.parametrize("system,N", [("dP->CCC", 14), ("J->CCC", 24)])
def test_biped_with_per_instance_plants_and_yaw_lowering(cpu_extend, system, N):
    """The walking biped with EVERYTHING per robot: its plant (tau_b, omega_b) and its heading --
    support polygon and terminal box turned by `rotate_in_TS(rotation2D(Param("yaw")))` on
    (b + n - s) and (DCM - s), signals with footstep and bias parts: the turned rows' entries under
    the footsteps (Gz) and their given parts (the h map) become per-instance tables of the
    coupling stage.  NumPy restatement of the kernels against the live reference."""
    import types

    from mpc_interface_b200 import dynamics
    from mpc_interface_b200.examples import biped, synthetic
    from oracle import reference

    if not reference.available():
        pytest.skip("oracle/_ref is not staged")
    ref = reference.load()
    api = types.SimpleNamespace(ControlSystem=ref.ControlSystem, ExtendedSystem=ref.ExtendedSystem,
                                DomainVariable=ref.DomainVariable, Formulation=ref.Formulation,
                                LineCombo=ref.LineCombo, Cost=ref.Cost, Box=ref.Box, tools=ref.tools)
    conf = biped.default_configuration(N)
    conf.system = system
    form = biped.formulate_biped(conf, omega=Param("omega"), tau=Param("tau"), yaw=Param("yaw"),
                                 target_vel=[Param("vel_x"), Param("vel_y")], backend=NumpyBackend())
    plan = form.lti_plan("LIP")
    assert plan.embedded and plan.Gz_terms and plan.Hg_terms and "yaw" in plan.parameter_names()
    batch = 3
    vals = synthetic.biped_inputs(3 + N, batch, N)
    vals["tau"] = np.array([0.09, 0.1, 0.11])
    vals["yaw"] = np.array([0.5, -1.3, 2.8])
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    params = {k: vals[k] for k in plan.parameter_names()}
    tables = plan.instance_tables(params, batch)
    ccx, ccu = plan.coupled_tables(params, batch)
    extras = plan.coupled_extras(params, batch)
    assert extras[3].shape == (batch, plan.nc, plan.nz) and extras[4].shape == (batch, int(plan.h_mask().sum()))
    for b in range(batch):
        one = biped.formulate_biped(conf, omega=float(vals["omega"][b]), tau=float(vals["tau"][b]),
                                    yaw=float(vals["yaw"][b]),
                                    target_vel=[float(vals["vel_x"][b]), float(vals["vel_y"][b])], api=api)
        want = one.generate_all_qp_matrices(one.arrange_given(
            {v: np.asarray(vals[v][b]).reshape([-1, 1]) for v in one.given_ID}))
        plant = dynamics.ControlSystem.from_name(system, ["_x"], tau=float(vals["tau"][b]),
                                                 omega=float(vals["omega"][b]))
        got = _lti_numpy.evaluate_embedded(plan, plant.A, plant.B, given[b], None,
                                           {k: v[b] for k, v in tables.items()}, (ccx[b], ccu[b]),
                                           tuple(None if t is None else t[b] for t in extras))
        cases.assert_qp_close(got, want)


@pytest.mark.parametrize("per_instance", [False, True])
def test_costs_and_limits_with_L_on_generic_plants(cpu_extend, per_instance):
    """``L`` (goal.py:88-104, restrictions.py:75-90): rows that combine the samples of the horizon.
    Costs go through the scheduled-cost stage with every row a combination (also crossed, with its
    own ``cross_L``); a limit becomes rows of the embedded layout -- bound in the h map, plant part
    added by mpci_lti_sched_rows.  A plain generic plant: the layout is the embedded one with no
    auxiliary variable at all.  Against the oracle's body.py:236-302."""
    from oracle import qp_oracle

    rng = np.random.default_rng(11)
    nx, nu, N, batch = 3, 2, 8, 3
    A = np.eye(nx) + 0.1 * rng.normal(size=[batch, nx, nx])
    B = 0.3 * rng.normal(size=[batch, nx, nu])
    Q, R = np.eye(nx), np.eye(nu)
    refs = [Param("ref%d" % i) for i in range(nx)]
    L1, L2, L3 = rng.normal(size=[3, N]), rng.normal(size=[3, N]), rng.normal(size=[2, 4])
    gain = Param("gain") if per_instance else 1.7

    def build(b):
        f = cases.formulate_generic(A[b], B[b], Q, R, N, rng.normal(size=[0, nx]), refs=refs, backend=NumpyBackend())
        f.incorporate_goal("smooth", Cost("z0", gain, aim=[refs[0]], L=L1))
        f.incorporate_goal("crossed", Cost("z1", 0.4, aim=[0.1], L=L1, cross="u0", cross_aim=[refs[1]], cross_L=L2))
        f.incorporate_goal("window", Cost("u1", 2.0, L=L3, schedule=range(3, 7)))
        f.incorporate_constraint("mean", Constraint("z2", 1.5, L=L1, arrow=[[-1.0]]))
        f.incorporate_constraint("late", Constraint("u0", np.array([0.5, 0.7]), L=L3, schedule=range(2, 6), center=[[0.1]]))
        f.make_preview_matrices()
        return f

    form = build(0)
    plan = form.lti_plan("plant")
    assert plan.embedded and plan.nz == 0 and len(plan.lrows) == 2 and sorted(plan.sched_L) == [0, 1, 2]
    assert plan.nc == form.lti_plan("plant").nc == 2 * nu * N + 3 + 2
    x0 = rng.normal(size=[batch, nx])
    params = {"ref%d" % i: rng.normal(size=batch) for i in range(nx)}
    params["gain"] = rng.uniform(0.5, 2.0, batch)
    terms, rows = plan.sched_tables(params, batch), plan.lrow_tables(params, batch)
    tables = plan.instance_tables(params, batch) if plan.per_instance else None
    names = [repr(e) for e in plan.slots.exprs]
    for b in range(batch):
        one = {k: float(v[b]) for k, v in params.items()}
        scalars = np.array([one[nm] for nm in names])
        inst = {k: v[b] for k, v in tables.items()} if tables is not None else None
        G, h, P, q = _lti_numpy.evaluate_embedded(plan, A[b], B[b], x0[b], scalars, inst)
        P, q = _lti_numpy.add_scheduled(plan, A[b], B[b], x0[b], terms[b if terms.shape[0] > 1 else 0], P, q)
        G, h = _lti_numpy.add_limit_rows(plan, A[b], B[b], x0[b], rows[b if rows.shape[0] > 1 else 0], G, h)
        cases.assert_qp_close((G, h, P, q), qp_oracle.all_qp_matrices(build(b), x0[b][:, None], one))
