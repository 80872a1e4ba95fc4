"""The embedded per-instance path on a formulation that is NOT the biped: a soft-constrained LIP.

Plant dP->CCC on one axis with one (tau_b, omega_b) per instance; next to it an OPTIMISED slack
``e`` (DomainVariable, N samples), a GIVEN disturbance ``d`` entering one sample late (matrix
coefficient), the signal ``soft = b + S d - e`` with ``b = CoM - CoM_ddot / omega_b^2``; costs on
the slack alone, on ``soft`` and on the CoM velocity (per-instance aim); limits on ``soft`` with
one arrow, with an arrow per row (the rows leave through the warps, not the bulk copies), on the
slack alone (rows without a plant part) and a scheduled one.  Every instance is formulated
from scratch by the unmodified reference.
"""
import os
import types

import numpy as np
import pytest

import _cases as cases
from oracle import reference

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["soft_lip_n12", "soft_lip_n17_slack_first"])
def test_soft_constrained_lip_goldens(name):
    """The committed vectors (tests/golden/make_goldens.py soft_lip: every instance formulated by
    the reference with its own plant), both Hessian variants."""
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import soft_lip, synthetic

    g = cases.golden(name)
    N, batch, order = int(g["N"]), g["A"].shape[0], tuple(str(x) for x in g["order"])
    form = soft_lip.formulate_soft_lip(N, Param("omega"), Param("tau"), Param("vel"), order=order)
    cases.check_layout(form, g)
    given = synthetic.pack_given(form.given_ID, form.given_len, {key: g[key] for key in form.given_ID}, batch)
    for mode in ("recursion", "dmma"):
        out = form.generate_all_qp_matrices_batched(
            given, {"omega": g["omega"], "vel": g["vel"]},
            dynamics={"LIP": ("dP->CCC", g["tau"], g["omega"])}, hessian=mode)
        got = tuple(x.cpu().numpy() for x in (out.G, out.h, out.P, out.q))
        for b in range(batch):
            cases.assert_qp_close(tuple(x[b] for x in got), (g["A"][b], g["h"][b], g["Q"][b], g["q"][b]))


@pytest.mark.skipif(not reference.available(), reason="oracle/_ref is not staged")
@pytest.mark.parametrize("N,order", [(20, ("cCoP_dot_x", "e_x")), (33, ("e_x", "cCoP_dot_x")),
                                     (70, ("cCoP_dot_x", "e_x"))])
def test_soft_constrained_lip_against_the_live_reference(N, order):
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import soft_lip

    batch = 6
    seed = int.from_bytes(os.urandom(4), "little")
    rng = np.random.default_rng(seed)
    vals = {"x0_x": rng.normal(size=[batch, 3]) * [0.03, 0.1, 0.2], "d_x": rng.normal(size=[batch, N]) * 0.01,
            "omega": np.sqrt(9.81 / rng.uniform(0.75, 1.0, batch)), "tau": rng.uniform(0.08, 0.12, batch) * min(1.0, 35.0 / N),  # (same preview time: A^N stays ~1e5)
            "vel": rng.uniform(0.0, 0.4, batch)}
    form = soft_lip.formulate_soft_lip(N, Param("omega"), Param("tau"), Param("vel"), order=order)
    plan = form.lti_plan("LIP")
    assert plan.embedded and plan.nz == N and plan.nv == 2 * N
    given = np.zeros([batch, form.given_len])
    for name, ids in form.given_ID.items():
        given[:, ids.start:ids.stop] = vals[name]
    ref = reference.load()
    api = types.SimpleNamespace(ControlSystem=ref.ControlSystem, ExtendedSystem=ref.ExtendedSystem,
                                DomainVariable=ref.DomainVariable, Formulation=ref.Formulation,
                                LineCombo=ref.LineCombo, Cost=ref.Cost, Constraint=ref.Constraint)
    for mode in ("recursion", "dmma"):
        out = form.generate_all_qp_matrices_batched(
            given, {"omega": vals["omega"], "vel": vals["vel"]},
            dynamics={"LIP": ("dP->CCC", vals["tau"], vals["omega"])}, hessian=mode)
        got = [x.cpu().numpy() for x in (out.G, out.h, out.P, out.q)]
        for b in range(batch):
            one = soft_lip.formulate_soft_lip(N, float(vals["omega"][b]), float(vals["tau"][b]), float(vals["vel"][b]),
                                              order=order, api=api)
            if b == 0:
                assert list(one.optim_ID.items()) == list(form.optim_ID.items())
                assert list(one.given_ID.items()) == list(form.given_ID.items())
            want = one.generate_all_qp_matrices(one.arrange_given(
                {name: np.asarray(vals[name][b]).reshape([-1, 1]) for name in one.given_ID}))
            try:
                cases.assert_qp_close(tuple(x[b] for x in got), want)
            except AssertionError as e:
                raise AssertionError("seed {} N {} instance {} ({}): {}".format(seed, N, b, mode, e))


@pytest.mark.skipif(not reference.available(), reason="oracle/_ref is not staged")
@pytest.mark.parametrize("seed,scheduled,general,turned", [(s, False, False, False) for s in range(300, 324)]
                         + [(s, True, False, False) for s in range(400, 424)]
                         + [(s, True, True, False) for s in range(500, 524)]
                         + [(s, True, True, True) for s in range(700, 720)]
                         + [(s, True, True, "L") for s in range(800, 820)]
                         + [(s, True, True, "LD") for s in range(900, 920)]
                         + [(s, True, True, "LDP") for s in range(1000, 1024)]
                         + [(s, True, True, "rows") for s in range(1100, 1116)])
def test_random_embedded_formulations_against_the_live_reference(seed, scheduled, general, turned):
    """tests/_embedded_forms.py through the kernels: random structures around a per-instance plant
    (four named plants, one or two axes, slack / footsteps / disturbance, costs with aims and on
    signals shorter than the horizon, facets, row-wise arrows, centres, schedules, any variable
    order), both Hessian variants, every instance checked against the unmodified reference.
    ``scheduled``: some costs act on a few samples only (terminal, a window, every 2nd / 3rd sample;
    per-instance weights and aims) -- the stage of csrc/mpci_sched.cu after the two kernels.
    ``general``: also crossed costs between any two signals, per-instance weights / aims on signals with
    a part outside the plant, a definition that mixes states and the input.  ``turned``: limits turned /
    scaled per instance, on signals with parts outside the plant too; "L": also costs and limits that
    combine their samples through a matrix ``L`` (mpci_lti_sched_rows for the limits' plant part);
    "LD": also definitions with a matrix coefficient on a plant signal (a rate, a delayed state);
    "LDP": also per-instance coefficients (1 / omega, omega) on the variables outside the plant;
    "rows": also arrows that turn from row to row per instance."""
    import _embedded_forms as forms
    from mpc_interface_b200 import Param

    spec = forms.draw(seed, scheduled, general, with_L=turned in ("L", "LD", "LDP", "rows"),
                      turned="rows" if turned == "rows" else bool(turned),
                      delayed=turned in ("LD", "LDP", "rows"), paramaux=turned in ("LDP", "rows"))
    form = forms.build(spec, forms.own_api(), Param("omega"), Param("tau"), Param("vel"), Param("gain"))
    plan = form.lti_plan("LIP")
    assert plan.embedded
    batch = 5
    vals = forms.inputs(spec, form, batch, seed + 1000)
    given = np.zeros([batch, form.given_len])
    for name, ids in form.given_ID.items():
        given[:, ids.start:ids.stop] = vals[name]
    params = {key: vals[key] for key in plan.parameter_names()}
    api = forms.reference_api(reference.load())
    wants = []
    for b in range(batch):
        one = forms.build(spec, api, *(float(vals[key][b]) for key in ("omega", "tau", "vel", "gain")))
        assert list(one.optim_ID.items()) == list(form.optim_ID.items())
        wants.append(one.generate_all_qp_matrices(one.arrange_given(
            {name: np.asarray(vals[name][b]).reshape([-1, 1]) for name in one.given_ID})))
    for mode in ("recursion", "dmma"):
        out = form.generate_all_qp_matrices_batched(
            given, params, dynamics={"LIP": (spec["system"], vals["tau"], vals["omega"])}, hessian=mode)
        got = [x.cpu().numpy() for x in (out.G, out.h, out.P, out.q)]
        for b in range(batch):
            try:
                cases.assert_qp_close(tuple(x[b] for x in got), wants[b])
            except AssertionError as e:
                raise AssertionError("seed {} instance {} ({}; {} {} N {}): {}".format(
                    seed, b, mode, spec["system"], spec["axes"], spec["N"], e))


@pytest.mark.skipif(not reference.available(), reason="oracle/_ref is not staged")
@pytest.mark.parametrize("seed", range(600, 632))
def test_random_formulations_with_a_shared_plant_against_the_live_reference(seed):
    """The same random structures through MODE A (the headline path: shared plant, affine plan,
    `stream_terms` / `qh` kernels): omega and tau are numbers, the per-instance parameters are the
    aims / bounds (`vel`) and the weights (`gain`); scheduled, crossed and state-input costs and
    costs / limits with a matrix ``L`` over their samples included -- mode A takes them all.  Every
    instance against the unmodified reference."""
    import _embedded_forms as forms
    from mpc_interface_b200 import Param

    spec = forms.draw(seed, scheduled=True, general=True, with_L=True)
    omega, tau = 3.4, 0.1 * min(1.0, 30.0 / spec["N"])
    form = forms.build(spec, forms.own_api(), omega, tau, Param("vel"), Param("gain"))
    batch = 5
    vals = forms.inputs(spec, form, batch, seed + 1000)
    given = np.zeros([batch, form.given_len])
    for name, ids in form.given_ID.items():
        given[:, ids.start:ids.stop] = vals[name]
    import scipy.sparse as sp

    params = {"vel": vals["vel"], "gain": vals["gain"]}
    out = form.generate_all_qp_matrices_batched(given, params)
    compact = form.generate_all_qp_matrices_batched(given, params, layout="csc")  # the solver hand-off
    got = [x.cpu().numpy() for x in (out.G, out.h, out.P, out.q)]
    # the step after a solve (body.py:209-234): every definition previewed from ONE launch, the goals' distances
    optim = np.random.default_rng(seed + 7).normal(size=[batch, form.optim_len]) * 0.1
    previews = {k: v.cpu().numpy() for k, v in form.preview_all_batched(given, optim, params=params).items()}
    distance = form.full_goal_distance_batched(given, optim, params=params).cpu().numpy()
    api = forms.reference_api(reference.load())
    for b in range(batch):
        one = forms.build(spec, api, omega, tau, float(vals["vel"][b]), float(vals["gain"][b]))
        assert list(one.optim_ID.items()) == list(form.optim_ID.items())
        g_b = one.arrange_given({name: np.asarray(vals[name][b]).reshape([-1, 1]) for name in one.given_ID})
        for name in one.definitions:
            want_v = one.preview(g_b, optim[b][:, None], name)[:, 0]
            assert cases.rel_err(previews[name][b], want_v) <= cases.TOL, (seed, b, name)
        want_d = one.full_goal_distance(g_b, optim[b][:, None])
        assert abs(distance[b] - want_d) <= cases.TOL * max(1.0, abs(want_d)), (seed, b)
        want = one.generate_all_qp_matrices(one.arrange_given(
            {name: np.asarray(vals[name][b]).reshape([-1, 1]) for name in one.given_ID}))
        try:
            cases.assert_qp_close(tuple(x[b] for x in got), want)
            for mine, dense in zip(compact.csc(b), (want[2], want[0])):  # biped_mpc_loop.py:57-58
                ref = sp.csc_matrix(dense)
                assert np.array_equal(mine.indptr, ref.indptr) and np.array_equal(mine.indices, ref.indices)
                assert cases.rel_err(mine.data, ref.data) <= cases.TOL
        except AssertionError as e:
            raise AssertionError("seed {} instance {} ({} {} N {}): {}".format(
                seed, b, spec["system"], spec["axes"], spec["N"], e))
