"""Costs with a ``schedule`` on the per-instance path: the stage of csrc/mpci_sched.cu
(``mpci_lti_sched_add``, through the C ABI) adds them to the P and q of the fused kernel.

Reference: generate_qp_cost with ``cost.schedule`` (body.py:266-302, the rows sigma of :272).
Checked against the oracle's restatement on generic plants (plain layout) and against the
unmodified reference on the biped and on random formulations around a per-instance plant
(embedded layout).  Tolerance: relative error <= 1e-10 per array (cases.TOL), normwise."""
import types

import numpy as np
import pytest

import _cases as cases
from oracle import qp_oracle as orc
from oracle import reference

pytestmark = pytest.mark.gpu


def _terminal_weight(seed, nx):
    M = np.random.default_rng(seed).normal(size=[nx, nx])
    return 3.0 * (M.T @ M) / nx + np.eye(nx)


@pytest.mark.parametrize("nx,nu,N,batch,window", [(4, 2, 9, 5, True), (12, 4, 50, 300, False), (2, 1, 5, 3, True),
                                                  (1, 1, 1, 2, False), (6, 2, 400, 3, False), (5, 5, 13, 4, True)])
def test_generic_mpc_with_a_terminal_cost(monkeypatch, nx, nu, N, batch, window):
    """(x_N - ref)' Qf (x_N - ref): nx^2 crossed costs on the last sample (examples/generic.py
    ``terminal=``), more rows than one chunk of shared memory holds at (12, 4, 50) and (6, 2, 400);
    ``window``: one more cost on scattered samples with a per-instance weight and aim (the plan
    then carries per-instance term tables).  Both Hessian variants; a few instances against the
    oracle."""
    import mpc_interface_b200.tools as tools
    from _numpy_backend import NumpyBackend
    from mpc_interface_b200 import Cost, Param
    from mpc_interface_b200.examples import synthetic

    seed = 500 + nx
    _, _, Q, R = synthetic.generic_system(seed, nx, nu)
    A, B = synthetic.generic_plants(seed, batch, nx, nu)
    rows = np.random.default_rng(seed + 1).normal(size=[2, nx])
    vals = synthetic.generic_inputs(seed, batch, nx)
    Qf = _terminal_weight(seed + 2, nx)
    gain = np.random.default_rng(seed + 3).uniform(0.5, 2.0, batch)
    samples = range(min(1, N - 1), N, 2)

    def build(b, refs, weight, backend=None):
        f = cases.formulate_generic(A[b], B[b], Q, R, N, rows, refs=refs, backend=backend, terminal=Qf)
        if window:
            f.incorporate_goal("window", Cost("z0", weight, aim=[refs[0] * 2], schedule=samples))
        return f

    form = build(0, [Param("ref%d" % i) for i in range(nx)], Param("gain"))
    plan = form.lti_plan("plant")
    assert len(plan.sched_terms) == nx * nx + (1 if window else 0) and not plan.embedded
    params = {"ref%d" % i: vals["ref"][:, i] for i in range(nx)}
    params["gain"] = gain
    outs = {}
    for mode in ("dmma", "recursion"):
        out = form.generate_all_qp_matrices_batched(
            vals["x0"], params, dynamics={"plant": (A, B)}, hessian=mode)
        outs[mode] = (out.G_numpy(), out.h_numpy(), out.P_numpy(), out.q_numpy())
    monkeypatch.setattr(tools, "extend_matrices", orc.extend_matrices)
    for b in sorted({0, batch // 2, batch - 1}):
        ref_form = build(b, [float(x) for x in vals["ref"][b]], float(gain[b]), backend=NumpyBackend())
        ref = orc.all_qp_matrices(ref_form, vals["x0"][b][:, None])
        for mode in outs:
            cases.assert_qp_close([x[b] for x in outs[mode]], ref)


def test_scheduled_costs_partial_outputs_and_refusals():
    """``parts``: the stage adds to the outputs that were asked for and leaves the others alone;
    ``assemble_host`` returns the same in host memory; the ring entry point refuses such a plan (no
    silent omission of the terminal cost)."""
    import torch

    from mpc_interface_b200 import Param, device, lti
    from mpc_interface_b200.examples import synthetic

    nx, nu, N, batch = 3, 2, 8, 6
    _, _, Q, R = synthetic.generic_system(9, nx, nu)
    A, B = synthetic.generic_plants(9, batch, nx, nu)
    vals = synthetic.generic_inputs(9, batch, nx)
    refs = [Param("ref%d" % i) for i in range(nx)]
    form = cases.formulate_generic(A[0], B[0], Q, R, N, np.ones([1, nx]), refs=refs, terminal=_terminal_weight(3, nx))
    plain = cases.formulate_generic(A[0], B[0], Q, R, N, np.ones([1, nx]), refs=refs)
    params = {"ref%d" % i: vals["ref"][:, i] for i in range(nx)}
    plan = form.lti_plan("plant")
    job = lti.LtiJob(plan, A, B, vals["x0"], params)
    full = job.launch()
    base = plain.generate_all_qp_matrices_batched(vals["x0"], params, dynamics={"plant": (A, B)})
    assert cases.rel_err(full.P.cpu().numpy(), base.P.cpu().numpy()) > 1e-3  # the terminal cost is there
    for parts in ("P", "q", "Gh", "Pq"):
        out = device.QPBatch(batch, plan.nv, plan.nc, torch.device("cuda", 0))
        for t in (out.P, out.q, out.G, out.h):
            t.fill_(-7.0)
        job.launch(out, parts=parts)
        for tag, got, want in zip("PqGh", (out.P, out.q, out.G, out.h), (full.P, full.q, full.G, full.h)):
            if tag in parts:
                assert torch.equal(got, want), (parts, tag)
            else:
                assert bool((got == -7.0).all()), (parts, tag)
    # HOST results: chunks through the same two launches (6 instances, chunks of 4 + 2)
    host = lti.assemble_host(plan, A, B, vals["x0"], params, chunk=4)
    for got, want in zip(host, (full.P, full.q, full.G, full.h)):
        assert torch.equal(got, want.cpu())
    with pytest.raises(NotImplementedError, match="schedule"):
        lti.assemble_host_stream(plan, A, B, vals["x0"], host, 1, params)
    with pytest.raises(KeyError, match="ref1"):
        lti.LtiJob(plan, A, B, vals["x0"], {"ref0": vals["ref"][:, 0]})


@pytest.mark.skipif(not reference.available(), reason="oracle/_ref is not staged")
@pytest.mark.parametrize("system,N,batch", [("dP->CCC", 24, 6), ("J->CCC", 30, 4), ("dP->CCC", 100, 3)])
def test_biped_with_per_instance_plants_and_a_terminal_cost(system, N, batch):
    """The biped's own terminal cost on (DCM - s), biped_formulation.py:118-126, switched on
    (its default weight is 0, biped_configuration.py:49): a scheduled cost on a signal with a
    per-instance plant part (1 / omega_b), a footstep part and the last sample only; every
    instance formulated by the unmodified reference with its own plant."""
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import biped, synthetic

    ref = reference.load()
    api = types.SimpleNamespace(ControlSystem=ref.ControlSystem, ExtendedSystem=ref.ExtendedSystem,
                                DomainVariable=ref.DomainVariable, Formulation=ref.Formulation,
                                LineCombo=ref.LineCombo, Cost=ref.Cost, Box=ref.Box, tools=ref.tools)
    conf = biped.default_configuration(N)
    conf.system = system
    conf.cost_weights["terminal"] = 40.0
    form = biped.formulate_biped(conf, omega=Param("omega"), tau=Param("tau"),
                                 target_vel=[Param("vel_x"), Param("vel_y")])
    plan = form.lti_plan("LIP")
    assert plan.embedded and len(plan.sched_terms) == 2 and plan.sched_terms[0][5] == [N - 1]
    vals = synthetic.biped_inputs(77 + N, batch, N)
    vals["tau"] = np.random.default_rng(N).uniform(0.08, 0.12, batch) * min(1.0, 30.0 / N)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    wants = []
    for b in range(batch):
        one = biped.formulate_biped(conf, omega=float(vals["omega"][b]), tau=float(vals["tau"][b]),
                                    target_vel=[float(vals["vel_x"][b]), float(vals["vel_y"][b])], api=api)
        wants.append(one.generate_all_qp_matrices(one.arrange_given(
            {v: np.asarray(vals[v][b]).reshape([-1, 1]) for v in one.given_ID})))
    for mode in ("recursion", "dmma"):
        out = form.generate_all_qp_matrices_batched(
            given, params, dynamics={"LIP": (system, vals["tau"], vals["omega"])}, hessian=mode)
        got = [x.cpu().numpy() for x in (out.G, out.h, out.P, out.q)]
        for b in range(batch):
            cases.assert_qp_close(tuple(x[b] for x in got), wants[b])


def test_costs_beyond_the_stage_weights_on_generic_plants(monkeypatch):
    """A cost that couples states and inputs (a LineCombo of a state and an input: no Qxu block in
    T' Qxx T + Quu x I) and a crossed state x input cost are lowered to whole-horizon terms of the
    scheduled-cost stage -- N rows each, several chunks of shared memory -- instead of being refused."""
    import mpc_interface_b200.tools as tools
    from _numpy_backend import NumpyBackend
    from mpc_interface_b200 import Cost, LineCombo, Param
    from mpc_interface_b200.examples import synthetic

    nx, nu, N, batch = 5, 2, 120, 7
    _, _, Q, R = synthetic.generic_system(61, nx, nu)
    A, B = synthetic.generic_plants(61, batch, nx, nu)
    A = np.eye(nx) + 0.2 * (A - np.eye(nx))  # (N = 120: keep A^N moderate)
    rows = np.random.default_rng(62).normal(size=[2, nx])
    vals = synthetic.generic_inputs(61, batch, nx)
    gain = np.random.default_rng(63).uniform(0.5, 2.0, batch)

    def build(b, refs, weight, backend=None):
        f = cases.formulate_generic(A[b], B[b], Q, R, N, rows, refs=refs, backend=backend)
        f.incorporate_definition("mix", LineCombo({"z0": 1.0, "u1": 2.0, "z3": -0.5}))
        f.incorporate_goal("mixed", Cost("mix", weight, aim=[refs[1]]))
        f.incorporate_goal("state x input", Cost("z2", 0.3, aim=[0.1], cross="u0", cross_aim=[refs[0]]))
        f.make_preview_matrices()
        return f

    form = build(0, [Param("ref%d" % i) for i in range(nx)], Param("gain"))
    plan = form.lti_plan("plant")
    assert [t[5] for t in plan.sched_terms] == [list(range(N))] * 2
    params = {"ref%d" % i: vals["ref"][:, i] for i in range(nx)}
    params["gain"] = gain
    outs = {}
    for mode in ("dmma", "recursion"):
        out = form.generate_all_qp_matrices_batched(vals["x0"], params, dynamics={"plant": (A, B)}, hessian=mode)
        outs[mode] = (out.G_numpy(), out.h_numpy(), out.P_numpy(), out.q_numpy())
    monkeypatch.setattr(tools, "extend_matrices", orc.extend_matrices)
    for b in (0, 3, batch - 1):
        ref_form = build(b, [float(x) for x in vals["ref"][b]], float(gain[b]), backend=NumpyBackend())
        ref = orc.all_qp_matrices(ref_form, vals["x0"][b][:, None])
        for mode in outs:
            cases.assert_qp_close([x[b] for x in outs[mode]], ref)


@pytest.mark.parametrize("nx,nu,N,batch", [(3, 2, 8, 5), (6, 2, 60, 9), (12, 4, 50, 4)])
def test_costs_and_limits_with_L_on_generic_plants(monkeypatch, nx, nu, N, batch):
    """``L`` on the per-instance path (goal.py:88-104, restrictions.py:75-90): costs whose rows combine
    the samples of the horizon (also crossed with its own ``cross_L``, also on a window) through the
    scheduled-cost stage; limits as rows of the embedded layout whose plant part mpci_lti_sched_rows
    adds.  Generic plants: the embedded layout without any auxiliary variable.  Against the oracle."""
    import mpc_interface_b200.tools as tools
    from _numpy_backend import NumpyBackend
    from mpc_interface_b200 import Constraint, Cost, Param
    from mpc_interface_b200.examples import synthetic

    seed = 900 + nx
    rng = np.random.default_rng(seed)
    _, _, Q, R = synthetic.generic_system(seed, nx, nu)
    A, B = synthetic.generic_plants(seed, batch, nx, nu)
    A = np.eye(nx) + 0.3 * (A - np.eye(nx))
    vals = synthetic.generic_inputs(seed, batch, nx)
    gain = rng.uniform(0.5, 2.0, batch)
    L1, L2, L3 = rng.normal(size=[3, N]), rng.normal(size=[3, N]), rng.normal(size=[2, 4])
    lin = rng.normal(size=[1, nx])

    def build(b, refs, weight, backend=None):
        f = cases.formulate_generic(A[b], B[b], Q, R, N, lin, refs=refs, backend=backend)
        f.incorporate_goal("smooth", Cost("z0", weight, aim=[refs[0]], L=L1))
        f.incorporate_goal("crossed", Cost("z1", 0.4, aim=[0.1], L=L1, cross="u0", cross_aim=[refs[1]], cross_L=L2))
        f.incorporate_goal("window", Cost("u1", 2.0, L=L3, schedule=range(3, 7)))
        f.incorporate_constraint("mean", Constraint("z2", 1.5, L=L1, arrow=[[-1.0]]))
        f.incorporate_constraint("late", Constraint("u0", np.array([0.5, 0.7]), L=L3, schedule=range(2, 6), center=[[0.1]]))
        f.make_preview_matrices()
        return f

    form = build(0, [Param("ref%d" % i) for i in range(nx)], Param("gain"))
    plan = form.lti_plan("plant")
    assert plan.embedded and plan.nz == 0 and len(plan.lrows) == 2
    params = {"ref%d" % i: vals["ref"][:, i] for i in range(nx)}
    params["gain"] = gain
    outs = {}
    for mode in ("dmma", "recursion"):
        out = form.generate_all_qp_matrices_batched(vals["x0"], params, dynamics={"plant": (A, B)}, hessian=mode)
        outs[mode] = (out.G_numpy(), out.h_numpy(), out.P_numpy(), out.q_numpy())
    monkeypatch.setattr(tools, "extend_matrices", orc.extend_matrices)
    for b in sorted({0, batch // 2, batch - 1}):
        ref_form = build(b, [float(x) for x in vals["ref"][b]], float(gain[b]), backend=NumpyBackend())
        ref = orc.all_qp_matrices(ref_form, vals["x0"][b][:, None])
        for mode in outs:
            cases.assert_qp_close([x[b] for x in outs[mode]], ref)
