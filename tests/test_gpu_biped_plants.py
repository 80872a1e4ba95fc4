"""The walking biped with ONE PLANT PER INSTANCE, on the GPU, against
the unmodified reference.

``formulate_biped`` with ``conf.system = "dP->CCC"``: the plant (CoP rate -> CoM, CoM_dot,
CoM_ddot, python/mpc_interface/tools.py:253-280) depends on every robot's pendulum frequency
and sampling period, so nothing of the prediction can be shared.  The QP keeps the whole
structure of the reference's example (biped_formulation.py:23-191): footstep variables Ds of
the shared ``steps`` system, the bias ``n`` entering one sample late through a matrix
coefficient, the support-polygon box on ``(b+n-s)``, the stepping area on ``Ds`` alone, the
terminal box on ``(DCM-s)``.  One ``generate_all_qp_matrices_batched(..., dynamics=...)`` call
assembles the batch (fused per-instance kernel in the embedded layout + the coupling kernel);
the reference formulates every instance from scratch with its own classes.
"""
import os

import numpy as np
import pytest

import _cases as cases
from oracle import reference

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not reference.available(), reason="oracle/_ref is not staged")]


def _inputs(seed, batch, N):
    from mpc_interface_b200.examples import synthetic

    vals = synthetic.biped_inputs(seed, batch, N)
    vals["tau"] = np.random.default_rng(seed + 5).uniform(0.08, 0.12, batch)
    return vals


def _conf(N, system):
    from mpc_interface_b200.examples import biped

    conf = biped.default_configuration(N)
    conf.system = system
    return conf


def _reference_instance(conf, vals, b, tick=None):
    from mpc_interface_b200.examples import biped

    yaw = {"yaw": float(vals["yaw"][b])} if "yaw" in vals else {}
    one = biped.formulate_biped(conf, omega=float(vals["omega"][b]), tau=float(vals["tau"][b]),
                                target_vel=[float(vals["vel_x"][b]), float(vals["vel_y"][b])],
                                api=reference.load(), **yaw)
    if tick is not None:
        one.update(**tick)
    collector = {name: np.asarray(vals[name][b]).reshape([-1, 1]) for name in one.given_ID}
    return one, one.generate_all_qp_matrices(one.arrange_given(collector))


@pytest.mark.parametrize("system,N,batch", [("dP->CCC", 24, 16), ("dP->CCC", 100, 12), ("J->CCC", 100, 6),
                                            ("dP->CCC", 37, 8)])
def test_biped_with_per_instance_plants_against_the_live_reference(system, N, batch):
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import biped, synthetic

    seed = int.from_bytes(os.urandom(4), "little")
    vals = _inputs(seed, batch, N)
    conf = _conf(N, system)
    form = biped.formulate_biped(conf, omega=Param("omega"), tau=Param("tau"),
                                 target_vel=[Param("vel_x"), Param("vel_y")])
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    got = {}
    for mode in ("recursion", "dmma"):
        out = form.generate_all_qp_matrices_batched(
            given, params, dynamics={"LIP": (system, vals["tau"], vals["omega"])}, hessian=mode)
        got[mode] = [x.cpu().numpy() for x in (out.G, out.h, out.P, out.q)]
    for b in range(batch):
        one, want = _reference_instance(conf, vals, b)
        if b == 0:
            assert list(one.optim_ID.items()) == list(form.optim_ID.items()), "integer layout"
            assert list(one.given_ID.items()) == list(form.given_ID.items())
        for mode in got:
            try:
                cases.assert_qp_close(tuple(x[b] for x in got[mode]), want)
            except AssertionError as e:
                raise AssertionError("seed {} instance {} ({}): {}".format(seed, b, mode, e))


def test_biped_plants_follow_the_walking_cycle():
    """Ticks of the receding horizon (biped_formulation.py:156-187): the step matrix E, the
    number of Ds variables and the stepping-area centres change with the phase; every tick is
    lowered again and must match the reference updated the same way."""
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import biped, synthetic

    N, batch, system = 30, 4, "dP->CCC"
    seed = int.from_bytes(os.urandom(4), "little")
    conf = _conf(N, system)
    form = biped.formulate_biped(conf, omega=Param("omega"), tau=Param("tau"),
                                 target_vel=[Param("vel_x"), Param("vel_y")])
    sizes = set()
    for tick in (dict(step_times=np.array([5, 17, 29]), step_count=1),
                 dict(step_times=np.array([0, 12, 24]), step_count=2),
                 dict(step_times=np.array([11, 23]), step_count=3)):
        form.update(**tick)
        vals = _inputs(seed, batch, N)
        given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
        out = form.generate_all_qp_matrices_batched(
            given, {k: vals[k] for k in ("omega", "vel_x", "vel_y")},
            dynamics={"LIP": (system, vals["tau"], vals["omega"])}, hessian="recursion")
        got = [x.cpu().numpy() for x in (out.G, out.h, out.P, out.q)]
        sizes.add(got[0].shape[1:])
        for b in range(batch):
            one, want = _reference_instance(conf, vals, b, tick)
            assert list(one.optim_ID.items()) == list(form.optim_ID.items())
            try:
                cases.assert_qp_close(tuple(x[b] for x in got), want)
            except AssertionError as e:
                raise AssertionError("seed {} tick {} instance {}: {}".format(seed, tick, b, e))
    assert len(sizes) > 1, "the ticks were meant to change the QP's size"


def test_embedded_plan_refuses_the_plain_entry_points():
    from mpc_interface_b200 import Param, device, lti
    from mpc_interface_b200.examples import biped

    form = biped.formulate_biped(_conf(24, "dP->CCC"), omega=Param("omega"), tau=Param("tau"))
    plan = form.lti_plan("LIP")
    assert plan.embedded and plan.nz == 2 * form.dynamics["steps"].domain["Ds_x"]
    with pytest.raises(ValueError, match="mpci_lti_assemble_coupled"):
        device.check(device.lib().mpci_lti_assemble(plan.device_plan().handle, None, 0, None, 0, None, None,
                                                    None, None, None, None, 1, lti.HESSIAN_RECURSION, None))


@pytest.mark.parametrize("name,system", [("biped_plants_n24", "dP->CCC"), ("biped_plants_n30_tick", "dP->CCC"),
                                         ("biped_plants_jccc_n16", "J->CCC")])
def test_biped_with_per_instance_plants_against_the_committed_golden(name, system):
    """The same against the reference's output stored under tests/golden (written by
    make_goldens.py biped_plants), explicit (A_b, B_b) instead of a named system."""
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import biped, synthetic

    g = cases.golden(name)
    conf = _conf(int(g["N"]), system)
    form = biped.formulate_biped(conf, omega=Param("omega"), tau=Param("tau"),
                                 target_vel=[Param("vel_x"), Param("vel_y")])
    if "tick_step_times" in g:
        form.update(step_times=g["tick_step_times"], step_count=int(g["tick_step_count"]))
    cases.check_layout(form, g)
    batch = g["A"].shape[0]
    given = synthetic.pack_given(form.given_ID, form.given_len, {k: g[k] for k in form.given_ID}, batch)
    for mode in ("recursion", "dmma"):
        out = form.generate_all_qp_matrices_batched(
            given, {k: g[k] for k in ("omega", "vel_x", "vel_y")},
            dynamics={"LIP": (g["A_sys"], g["B_sys"])}, hessian=mode)
        got = [x.cpu().numpy() for x in (out.G, out.h, out.P, out.q)]
        for b in range(batch):
            cases.assert_qp_close(tuple(x[b] for x in got), (g["A"][b], g["h"][b], g["Q"][b], g["q"][b]))


def test_biped_plants_large_batch_properties_and_samples():
    """8,192 robots at N = 100 in one call (9.4 GB of matrices): instances with equal inputs get
    bitwise equal matrices wherever they sit in the batch (the persistent CTAs walk the batch
    in strides), P is symmetric, the rows of the stepping area do not depend on the instance,
    and sampled instances match the live reference."""
    import torch

    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import biped, synthetic

    N, batch, system = 100, 8192, "dP->CCC"
    seed = int.from_bytes(os.urandom(4), "little")
    vals = _inputs(seed, batch, N)
    twins = [(0, 4097), (17, 8191), (1000, 1001)]
    for a, b in twins:
        for k in vals:
            vals[k][b] = vals[k][a]
    conf = _conf(N, system)
    form = biped.formulate_biped(conf, omega=Param("omega"), tau=Param("tau"),
                                 target_vel=[Param("vel_x"), Param("vel_y")])
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    out = form.generate_all_qp_matrices_batched(
        given, {k: vals[k] for k in ("omega", "vel_x", "vel_y")},
        dynamics={"LIP": (system, vals["tau"], vals["omega"])}, hessian="recursion")
    torch.cuda.synchronize()
    for a, b in twins:
        for x in (out.P, out.q, out.G, out.h):
            assert torch.equal(x[a], x[b]), "seed {} twins {} {}".format(seed, a, b)
    asym = (out.P - out.P.transpose(1, 2)).abs().amax()
    assert float(asym) <= 1e-10 * float(out.P.abs().amax())
    k = form.domain["Ds_x"]
    area = out.G[:, :4 * k, :]  # the stepping area: constants of the formulation
    assert torch.equal(area, area[:1].expand_as(area))
    assert bool(torch.isfinite(out.h).all()) and bool(torch.isfinite(out.q).all())
    for b in (0, 4097, 2048, 8191):
        one, want = _reference_instance(conf, vals, b)
        got = tuple(x[b].cpu().numpy() for x in (out.G, out.h, out.P, out.q))
        try:
            cases.assert_qp_close(got, want)
        except AssertionError as e:
            raise AssertionError("seed {} instance {}: {}".format(seed, b, e))


def test_embedded_plan_without_auxiliary_variables():
    """Edge of the embedded layout: the footsteps GIVEN instead of optimised (nz = 0: no
    auxiliary columns, only auxiliary givens -- s0, Ds, the bias); the stepping area then has
    no optimisation part at all (zero rows of G, h from the given steps)."""
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import biped, synthetic

    N, batch, system = 24, 5, "dP->CCC"
    seed = int.from_bytes(os.urandom(4), "little")
    vals = _inputs(seed, batch, N)
    conf = _conf(N, system)
    form = biped.formulate_biped(conf, omega=Param("omega"), tau=Param("tau"),
                                 target_vel=[Param("vel_x"), Param("vel_y")])
    drive = ["cCoP_dot_x", "cCoP_dot_y"]
    form.identify_qp_domain(drive)
    form.make_preview_matrices()
    rng = np.random.default_rng(seed)
    for ax in ("_x", "_y"):
        vals["Ds" + ax] = rng.normal(size=[batch, form.domain["Ds" + ax]]) * 0.05
    plan = form.lti_plan("LIP")
    assert plan.embedded and plan.nz == 0 and plan.nv == 2 * N
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    for mode in ("recursion", "dmma"):
        out = form.generate_all_qp_matrices_batched(
            given, {k: vals[k] for k in ("omega", "vel_x", "vel_y")},
            dynamics={"LIP": (system, vals["tau"], vals["omega"])}, hessian=mode)
        got = [x.cpu().numpy() for x in (out.G, out.h, out.P, out.q)]
        for b in range(batch):
            one = biped.formulate_biped(conf, omega=float(vals["omega"][b]), tau=float(vals["tau"][b]),
                                        target_vel=[float(vals["vel_x"][b]), float(vals["vel_y"][b])],
                                        api=reference.load())
            one.identify_qp_domain(drive)
            one.make_preview_matrices()
            assert list(one.given_ID.items()) == list(form.given_ID.items())
            want = one.generate_all_qp_matrices(one.arrange_given(
                {name: np.asarray(vals[name][b]).reshape([-1, 1]) for name in one.given_ID}))
            try:
                cases.assert_qp_close(tuple(x[b] for x in got), want)
            except AssertionError as e:
                raise AssertionError("seed {} instance {} ({}): {}".format(seed, b, mode, e))


@pytest.mark.parametrize("everything", [False, True])
def test_receding_tick_with_per_instance_plants(everything):
    """StructureBuckets.assemble_device(..., dynamics=...): robots at different walking phases
    (different QP sizes) AND with one plant each -- every bucket through the fused per-instance
    kernel; bitwise the matrices of a direct batched call on the bucket's instances, and the
    live reference on a sample.  ``everything``: also one heading per robot (turned boxes) and the
    terminal cost switched on (the scheduled-cost stage inside the tick)."""
    import torch

    from mpc_interface_b200 import Param, lti, receding
    from mpc_interface_b200.examples import biped, synthetic

    N, batch, system = 30, 61, "dP->CCC"
    seed = int.from_bytes(os.urandom(4), "little")
    conf = _conf(N, system)
    extra = {}
    if everything:
        conf.cost_weights["terminal"] = 25.0
        extra = {"yaw": Param("yaw")}

    def factory():
        return biped.formulate_biped(conf, omega=Param("omega"), tau=Param("tau"),
                                     target_vel=[Param("vel_x"), Param("vel_y")], **extra)

    rng = np.random.default_rng(seed)
    ticks = [dict(step_times=np.array([5, 17, 29]), step_count=1), dict(step_times=np.array([0, 12, 24]), step_count=2),
             dict(step_times=np.array([11, 23, 35]), step_count=3)]
    phase = rng.integers(0, len(ticks), size=batch)
    times = np.stack([ticks[t]["step_times"] for t in phase])
    count = np.array([ticks[t]["step_count"] for t in phase])
    vals = _inputs(seed, batch, N)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    if everything:
        vals["yaw"] = params["yaw"] = rng.uniform(-np.pi, np.pi, batch)
    collector = {k: torch.from_numpy(np.ascontiguousarray(vals[k])).cuda()
                 for k in ("s0_x", "s0_y", "x0_x", "x0_y", "n_x", "n_y")}
    A, B = lti.named_system_batched(system, vals["tau"], vals["omega"])
    buckets = receding.StructureBuckets(factory, biped.structure_key(N))
    done = buckets.assemble_device(biped.structure_keys(N, times, count),
                                   lambda i: dict(step_times=times[i], step_count=int(count[i])), collector,
                                   params=params, dynamics={"LIP": (A, B)})
    assert len(done) == len(set(phase)) and len({(qp.nv, qp.nc) for *_x, qp in done}) > 1
    seen = 0
    for key, index, form, qp in done:
        idx = index.cpu().numpy()
        seen += len(idx)
        given = synthetic.pack_given(form.given_ID, form.given_len, {k: vals[k][idx] for k in form.given_ID}, len(idx))
        direct = form.generate_all_qp_matrices_batched(given, {k: v[idx] for k, v in params.items()},
                                                       dynamics={"LIP": (A[index], B[index])}, hessian="recursion")
        for x, y in ((qp.P, direct.P), (qp.q, direct.q), (qp.G, direct.G), (qp.h, direct.h)):
            assert torch.equal(x, y)
        b = int(idx[0])
        _one, want = _reference_instance(conf, vals, b, dict(step_times=times[b], step_count=int(count[b])))
        cases.assert_qp_close(tuple(x[0].cpu().numpy() for x in (qp.G, qp.h, qp.P, qp.q)), want)
    assert seen == batch


def test_biped_plants_host_results():
    """``lti.assemble_host`` on an embedded plan with per-instance tables: chunks of instances
    through the two kernels on two streams, results in pinned host memory, bit-identical with
    the device-resident call (ragged last chunk; the named plants as CUDA tensors)."""
    import torch

    from mpc_interface_b200 import Param, lti
    from mpc_interface_b200.examples import biped, synthetic

    N, batch = 24, 11
    conf = biped.default_configuration(N)
    conf.system = "dP->CCC"
    form = biped.formulate_biped(conf, omega=Param("omega"), tau=Param("tau"),
                                 target_vel=[Param("vel_x"), Param("vel_y")])
    vals = synthetic.biped_inputs(31, batch, N)
    vals["tau"] = np.random.default_rng(2).uniform(0.08, 0.12, batch)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    plan = form.lti_plan("LIP")
    A, B = lti.named_system_batched("dP->CCC", vals["tau"], vals["omega"])
    want = lti.assemble(plan, A, B, given, params, hessian=lti.HESSIAN_RECURSION)
    host = lti.assemble_host(plan, A, B, given, params, hessian=lti.HESSIAN_RECURSION, chunk=4)
    for got, ref in zip(host, (want.P, want.q, want.G, want.h)):
        assert got.is_pinned() and torch.equal(got, ref.cpu())


@pytest.mark.skipif(not reference.available(), reason="oracle/_ref is not staged")
@pytest.mark.parametrize("N,batch", [(24, 6), (100, 3)])
def test_biped_plants_with_per_instance_cost_weights(N, batch):
    """One weight per robot on EVERY cost of the biped (biped_configuration.py:44-50): jerk and
    velocity tracking stay inside the plant (per-instance Qxx / Quu), `relax ankles` acts on
    (b + n - s) -- a signal with footstep and bias parts: per-instance cost_w and Pzz of the
    coupling stage, mpci_lti_coupled_tables --, and the terminal cost is a scheduled one.  Every
    instance formulated by the unmodified reference with its own plant and its own numbers."""
    import copy
    import types

    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import biped, synthetic

    ref = reference.load()
    api = types.SimpleNamespace(ControlSystem=ref.ControlSystem, ExtendedSystem=ref.ExtendedSystem,
                                DomainVariable=ref.DomainVariable, Formulation=ref.Formulation,
                                LineCombo=ref.LineCombo, Cost=ref.Cost, Box=ref.Box, tools=ref.tools)
    conf = biped.default_configuration(N)
    conf.system = "dP->CCC"
    names = {"minimize jerk": "w_jerk", "track velocity": "w_vel", "relax ankles": "w_ankles", "terminal": "w_end"}
    conf.cost_weights = {k: Param(v) for k, v in names.items()}
    form = biped.formulate_biped(conf, omega=Param("omega"), tau=Param("tau"),
                                 target_vel=[Param("vel_x"), Param("vel_y")])
    plan = form.lti_plan("LIP")
    assert plan.embedded and len(plan.Pzz_terms) == 2 and len(plan.sched_terms) == 2
    rng = np.random.default_rng(N)
    vals = synthetic.biped_inputs(9 + N, batch, N)
    vals["tau"] = rng.uniform(0.08, 0.12, batch) * min(1.0, 30.0 / N)
    for v, (lo, hi) in zip(names.values(), ((1e-4, 1e-2), (0.005, 0.05), (0.2, 3.0), (1.0, 50.0))):
        vals[v] = rng.uniform(lo, hi, batch)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    params = {k: vals[k] for k in ["omega", "vel_x", "vel_y"] + list(names.values())}
    wants = []
    for b in range(batch):
        one_conf = copy.copy(conf)
        one_conf.cost_weights = {k: float(vals[v][b]) for k, v in names.items()}
        one = biped.formulate_biped(one_conf, omega=float(vals["omega"][b]), tau=float(vals["tau"][b]),
                                    target_vel=[float(vals["vel_x"][b]), float(vals["vel_y"][b])], api=api)
        wants.append(one.generate_all_qp_matrices(one.arrange_given(
            {v: np.asarray(vals[v][b]).reshape([-1, 1]) for v in one.given_ID})))
    for mode in ("recursion", "dmma"):
        out = form.generate_all_qp_matrices_batched(
            given, params, dynamics={"LIP": ("dP->CCC", vals["tau"], vals["omega"])}, hessian=mode)
        got = [x.cpu().numpy() for x in (out.G, out.h, out.P, out.q)]
        for b in range(batch):
            cases.assert_qp_close(tuple(x[b] for x in got), wants[b])


@pytest.mark.skipif(not reference.available(), reason="oracle/_ref is not staged")
@pytest.mark.parametrize("system,N,batch", [("dP->CCC", 24, 6), ("J->CCC", 37, 4), ("dP->CCC", 100, 3)])
def test_biped_with_per_instance_plants_and_per_instance_yaw(system, N, batch):
    """Every robot with its own plant (tau_b, omega_b) AND its own heading: support polygon and
    terminal box turned by `rotate_in_TS(rotation2D(Param("yaw")))` (restrictions.py:380-401) on
    (b + n - s) and (DCM - s) -- the rows' entries under the footsteps and their given parts travel
    as the per-instance Gz / h_val of mpci_lti_coupled_tables.  Every instance formulated, turned
    and assembled by the unmodified reference."""
    import types

    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import biped, synthetic

    ref = reference.load()
    api = types.SimpleNamespace(ControlSystem=ref.ControlSystem, ExtendedSystem=ref.ExtendedSystem,
                                DomainVariable=ref.DomainVariable, Formulation=ref.Formulation,
                                LineCombo=ref.LineCombo, Cost=ref.Cost, Box=ref.Box, tools=ref.tools)
    conf = biped.default_configuration(N)
    conf.system = system
    form = biped.formulate_biped(conf, omega=Param("omega"), tau=Param("tau"), yaw=Param("yaw"),
                                 target_vel=[Param("vel_x"), Param("vel_y")])
    vals = synthetic.biped_inputs(50 + N, batch, N)
    rng = np.random.default_rng(N)
    vals["tau"] = rng.uniform(0.08, 0.12, batch) * min(1.0, 30.0 / N)
    vals["yaw"] = rng.uniform(-np.pi, np.pi, batch)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y", "yaw")}
    wants = []
    for b in range(batch):
        one = biped.formulate_biped(conf, omega=float(vals["omega"][b]), tau=float(vals["tau"][b]),
                                    yaw=float(vals["yaw"][b]),
                                    target_vel=[float(vals["vel_x"][b]), float(vals["vel_y"][b])], api=api)
        wants.append(one.generate_all_qp_matrices(one.arrange_given(
            {v: np.asarray(vals[v][b]).reshape([-1, 1]) for v in one.given_ID})))
    for mode in ("recursion", "dmma"):
        out = form.generate_all_qp_matrices_batched(
            given, params, dynamics={"LIP": (system, vals["tau"], vals["omega"])}, hessian=mode)
        got = [x.cpu().numpy() for x in (out.G, out.h, out.P, out.q)]
        for b in range(batch):
            cases.assert_qp_close(tuple(x[b] for x in got), wants[b])


@pytest.mark.skipif(not reference.available(), reason="oracle/_ref is not staged")
@pytest.mark.parametrize("plants", [False, True])
def test_box_transforms_per_instance(plants):
    """SURVEY 8(f) row 4 through the kernels: the support polygon turned, shifted and scaled and the
    terminal box re-centred with one value per robot (restrictions.py:380-486 with Params) -- in
    mode A (shared plant) and on the per-instance path (one plant per robot as well).  Every
    instance formulated and transformed by the unmodified reference."""
    import types

    from mpc_interface_b200 import Param, tools
    from mpc_interface_b200.examples import biped, synthetic

    ref = reference.load()
    api = types.SimpleNamespace(ControlSystem=ref.ControlSystem, ExtendedSystem=ref.ExtendedSystem,
                                DomainVariable=ref.DomainVariable, Formulation=ref.Formulation,
                                LineCombo=ref.LineCombo, Cost=ref.Cost, Box=ref.Box, tools=ref.tools)
    N, batch = 30, 5
    conf = biped.default_configuration(N)
    conf.system = "dP->CCC" if plants else conf.system
    kw = {"tau": Param("tau")} if plants else {}
    form = biped.formulate_biped(conf, omega=Param("omega"), target_vel=[Param("vel_x"), Param("vel_y")], **kw)
    cases.transform_boxes(form, tools.rotation2D(Param("yaw")), (Param("dx"), Param("dy")), Param("sc"), Param("rc"))
    vals = synthetic.biped_inputs(13, batch, N)
    vals.update(cases.transform_values(batch, 14))
    vals["tau"] = np.random.default_rng(15).uniform(0.08, 0.12, batch) if plants else np.full(batch, conf.mpc_period)
    if not plants and conf.system != "J->CCC":
        pytest.skip("the shared plant must not depend on omega")
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y", "yaw", "dx", "dy", "sc", "rc")}
    dynamics = {"dynamics": {"LIP": ("dP->CCC", vals["tau"], vals["omega"])}} if plants else {}
    out = form.generate_all_qp_matrices_batched(given, params, **dynamics)
    got = [x.cpu().numpy() for x in (out.G, out.h, out.P, out.q)]
    for b in range(batch):
        one = biped.formulate_biped(conf, omega=float(vals["omega"][b]), tau=float(vals["tau"][b]),
                                    target_vel=[float(vals["vel_x"][b]), float(vals["vel_y"][b])], api=api)
        cases.transform_boxes(one, ref.tools.rotation2D(float(vals["yaw"][b])),
                              (float(vals["dx"][b]), float(vals["dy"][b])), float(vals["sc"][b]), float(vals["rc"][b]))
        want = one.generate_all_qp_matrices(one.arrange_given(
            {v: np.asarray(vals[v][b]).reshape([-1, 1]) for v in one.given_ID}))
        cases.assert_qp_close(tuple(x[b] for x in got), want)


@pytest.mark.skipif(not reference.available(), reason="oracle/_ref is not staged")
@pytest.mark.parametrize("N,batch", [(24, 5), (60, 3)])
def test_biped_plants_with_cop_rate_limits_and_costs(N, batch):
    """A matrix coefficient ON a plant signal: the CoP rate ``(I - shift) @ b`` (b with its
    per-instance 1 / omega_b^2) next to the footsteps, penalised and limited -- crossed terms of
    the scheduled-cost stage over the pairs of the signal's parts, rows of mpci_lti_sched_rows.
    One plant per robot; every instance formulated by the unmodified reference."""
    import types

    from mpc_interface_b200 import Constraint, Cost, LineCombo, Param
    from mpc_interface_b200.examples import biped, synthetic

    ref = reference.load()
    api = types.SimpleNamespace(ControlSystem=ref.ControlSystem, ExtendedSystem=ref.ExtendedSystem,
                                DomainVariable=ref.DomainVariable, Formulation=ref.Formulation,
                                LineCombo=ref.LineCombo, Cost=ref.Cost, Box=ref.Box, tools=ref.tools)
    D = np.eye(N) - np.eye(N, k=-1)

    def with_rates(form, combo, cost, constraint, omega):
        for ax in ("_x", "_y"):
            form.incorporate_definition("raw" + ax, combo({"b" + ax: D, "s" + ax: -0.5}))
            form.incorporate_definition("db" + ax, combo({"raw" + ax: 1 / omega}))  # (a per-instance scaling on top)
        form.incorporate_goal("CoP rate", cost("db", 0.3, aim=[0.01, 0.0], axes=["_x", "_y"]))
        form.incorporate_constraint("CoP rate limit", constraint("db", 0.2, axes=["_x"]))
        form.incorporate_constraint("late rate", constraint("db", 0.1, axes=["_y"], arrow=[[-1.0]],
                                                            schedule=range(N - 3, N)))
        form.make_preview_matrices()
        return form

    conf = biped.default_configuration(N)
    conf.system = "dP->CCC"
    form = with_rates(biped.formulate_biped(conf, omega=Param("omega"), tau=Param("tau"),
                                            target_vel=[Param("vel_x"), Param("vel_y")]), LineCombo, Cost, Constraint,
                      Param("omega"))
    vals = synthetic.biped_inputs(21 + N, batch, N)
    vals["tau"] = np.random.default_rng(N).uniform(0.08, 0.12, batch) * min(1.0, 30.0 / N)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    wants = []
    for b in range(batch):
        one = with_rates(biped.formulate_biped(conf, omega=float(vals["omega"][b]), tau=float(vals["tau"][b]),
                                               target_vel=[float(vals["vel_x"][b]), float(vals["vel_y"][b])], api=api),
                         ref.LineCombo, ref.Cost, ref.Constraint, float(vals["omega"][b]))
        wants.append(one.generate_all_qp_matrices(one.arrange_given(
            {v: np.asarray(vals[v][b]).reshape([-1, 1]) for v in one.given_ID})))
    for mode in ("recursion", "dmma"):
        out = form.generate_all_qp_matrices_batched(
            given, params, dynamics={"LIP": ("dP->CCC", vals["tau"], vals["omega"])}, hessian=mode)
        got = [x.cpu().numpy() for x in (out.G, out.h, out.P, out.q)]
        for b in range(batch):
            cases.assert_qp_close(tuple(x[b] for x in got), wants[b])
