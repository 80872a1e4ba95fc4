#!/usr/bin/env python3
"""Generates the golden vectors under tests/golden/ from the UNMODIFIED
reference (imported from /root/reference, build container only).

    python tests/golden/make_goldens.py

Inputs come from mpc_interface_b200.examples.synthetic (pure numpy RNG); every
matrix in the .npz files is produced by the reference's own code:
``formulate_biped`` + ``Formulation.generate_all_qp_matrices`` (body.py:333),
``tools.extend_matrices`` (tools.py:14) and the pickle the reference's test
suite pins (python/tests/LIP_matrices, test_dynamics.py:139-152).
"""
import os
import pickle
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference/python"
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)
sys.path.insert(0, REF + "/use_examples/simple_functional_example")
sys.dont_write_bytecode = True

import mpc_interface.body as ref_body  # noqa: E402
import mpc_interface.combinations as ref_comb  # noqa: E402
import mpc_interface.dynamics as ref_dyn  # noqa: E402
import mpc_interface.goal as ref_goal  # noqa: E402
import mpc_interface.restrictions as ref_res  # noqa: E402
import mpc_interface.tools as ref_tools  # noqa: E402
import biped_configuration  # noqa: E402
from biped_formulation import formulate_biped  # noqa: E402

from mpc_interface_b200.examples import synthetic  # noqa: E402
from mpc_interface_b200.examples.generic import formulate_generic  # noqa: E402


class RefAPI:
    ControlSystem, ExtendedSystem = ref_dyn.ControlSystem, ref_dyn.ExtendedSystem
    DomainVariable = ref_dyn.DomainVariable
    Formulation, LineCombo = ref_body.Formulation, ref_comb.LineCombo
    Cost, Box, Constraint = ref_goal.Cost, ref_res.Box, ref_res.Constraint


def save(name, **arrays):
    path = os.path.join(HERE, name)
    np.savez_compressed(path, **arrays)
    print("{:28s} {:9.1f} kB".format(name, os.path.getsize(path) / 1e3))


def layout(form):
    """Integer layout as arrays: [start, stop] per variable, in dict order."""
    out = {}
    for kind in ("optim", "given"):
        ids = getattr(form, kind + "_ID")
        out[kind + "_names"] = np.array(list(ids.keys()))
        out[kind + "_ranges"] = np.array([[r.start, r.stop] for r in ids.values()], dtype=np.int64)
    return out


def lip_pickle():
    with open(REF + "/tests/LIP_matrices", "rb") as f:
        data = pickle.load(f)
    U, S = data["matrices"]
    A = np.array([[1, 0.1, 0.1**2 / 2], [0, 1, 0.1], [0, 0, 1]])
    LIP = ref_dyn.ControlSystem.from_name("J->CCC", tau=0.1, omega=3.3445)
    save("lip_matrices.npz", U=U, S=S, A=np.array(LIP.A, dtype=float), B=np.array(LIP.B, dtype=float))
    assert np.allclose(LIP.A, A)


def extend_cases():
    rng = np.random.default_rng(11)
    out = {}
    for tag, (n, m, N) in {"a": (3, 1, 9), "b": (8, 6, 20), "c": (12, 4, 50), "d": (1, 1, 5), "e": (6, 2, 33)}.items():
        A = np.eye(n) + 0.1 * rng.normal(size=[n, n])
        B = rng.normal(size=[n, m])
        S, U = ref_tools.extend_matrices(N, A, B)
        out.update({tag + "_A": A, tag + "_B": B, tag + "_S": S, tag + "_U": np.stack(U)})
    save("extend_cases.npz", **out)


def biped_conf(N):
    conf = types.SimpleNamespace(
        **{k: v for k, v in vars(biped_configuration).items() if not k.startswith("__")}
    )
    conf.horizon_lenght = N
    return conf


def set_omega(form, w):
    """SURVEY Appendix C probe8: per-instance omega through the reference."""
    for ax in ("_x", "_y"):
        d = form.definitions["b" + ax]
        d.matrices[d.variables.index("CoM_ddot" + ax)] = -1 / w**2
        d = form.definitions["DCM" + ax]
        d.matrices[d.variables.index("CoM_dot" + ax)] = 1 / w
    form.make_preview_matrices()


def biped_c1():
    form = formulate_biped(biped_configuration)
    collector = {v: np.zeros([s, 1]) for v, s in form.domain.items()}
    collector["x0_y"][0] = biped_configuration.strt_y
    collector["s0_y"][0] = biped_configuration.strt_y
    given = form.arrange_given(collector)
    A, h, Q, q = form.generate_all_qp_matrices(given)
    PM_b = np.hstack(form.PM["(b+n-s)_x"])
    save("biped_c1.npz", given=given, A=A, h=h, Q=Q, q=q, PM_bns_x=PM_b, **layout(form))


def biped_batch(name, N, batch, seed, terminal_weight=None):
    conf = biped_conf(N)
    if terminal_weight is not None:
        conf.cost_weights = dict(conf.cost_weights, terminal=terminal_weight)
    form = formulate_biped(conf)
    vals = synthetic.biped_inputs(seed, batch, N)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    As, hs, Qs, qs = [], [], [], []
    for b in range(batch):
        set_omega(form, vals["omega"][b])
        form.goals["track vel_x"].update(aim=vals["vel_x"][b])
        form.goals["track vel_y"].update(aim=vals["vel_y"][b])
        A, h, Q, q = form.generate_all_qp_matrices(given[b][:, None])
        As.append(A), hs.append(h[:, 0]), Qs.append(Q), qs.append(q[:, 0])
    save(
        name,
        seed=seed, N=N, terminal_weight=-1.0 if terminal_weight is None else terminal_weight,
        given=given, omega=vals["omega"], vel_x=vals["vel_x"], vel_y=vals["vel_y"],
        A=np.stack(As), h=np.stack(hs), Q=np.stack(Qs), q=np.stack(qs), **layout(form),
    )


def biped_yaw(name, N, batch, seed):
    """Per-instance feet orientation: the support polygon and the terminal box
    turned by yaw_b through the reference's Box.rotate_in_TS
    (restrictions.py:437-457) on top of per-instance omega / aims."""
    form = formulate_biped(biped_conf(N))
    vals = synthetic.biped_inputs(seed, batch, N)
    yaw = np.random.default_rng(seed + 7).uniform(-1.2, 1.2, batch)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    boxes = [form.constraint_boxes[k] for k in ("support_polygon", "terminal_Constraint")]
    straight = [[lim.arrow.copy() for lim in box.constraints] for box in boxes]
    As, hs, Qs, qs = [], [], [], []
    for b in range(batch):
        set_omega(form, vals["omega"][b])
        form.goals["track vel_x"].update(aim=vals["vel_x"][b])
        form.goals["track vel_y"].update(aim=vals["vel_y"][b])
        for box, arrows in zip(boxes, straight):
            for lim, arrow in zip(box.constraints, arrows):
                lim.update(arrow=arrow.copy())  # rotate_in_TS turns the CURRENT arrows
            box.rotate_in_TS(ref_tools.rotation2D(yaw[b]))
        A, h, Q, q = form.generate_all_qp_matrices(given[b][:, None])
        As.append(A), hs.append(h[:, 0]), Qs.append(Q), qs.append(q[:, 0])
    save(name, seed=seed, N=N, given=given, omega=vals["omega"], vel_x=vals["vel_x"],
         vel_y=vals["vel_y"], yaw=yaw, terminal_weight=-1.0,
         A=np.stack(As), h=np.stack(hs), Q=np.stack(Qs), q=np.stack(qs), **layout(form))


def biped_tick(N=24):
    """Structure after Formulation.update (ragged sizes)."""
    conf = biped_conf(N)
    form = formulate_biped(conf)
    n = conf.step_samples
    out = {}
    step_times = np.array([(i + 1) * n - 1 for i in range(form.domain["Ds_x"])])
    step_count = 0
    rng = np.random.default_rng(5)
    for tick in range(14):
        form.update(step_times=step_times, step_count=step_count)
        if tick in (0, 11, 12, 13):
            given = rng.normal(size=[form.given_len, 1]) * 0.05
            A, h, Q, q = form.generate_all_qp_matrices(given)
            out.update({"t%d_given" % tick: given, "t%d_A" % tick: A, "t%d_h" % tick: h,
                        "t%d_Q" % tick: Q, "t%d_q" % tick: q,
                        "t%d_step_times" % tick: step_times.copy(),
                        "t%d_step_count" % tick: step_count})
        step_times = step_times - 1
        if step_times[0] == -1:
            step_times = step_times + n
            step_count += 1
    save("biped_ticks.npz", **out)


def test_body_case():
    """The formulation of the reference's own test (python/tests/test_body.py:23-98):
    L matrices, a crossed cost, x0 among the optimisation variables."""
    LIP = ref_dyn.ControlSystem.from_name("J->CCC", tau=0.1, omega=3.5, axes=["_x", "_y"])
    LIP_ext = ref_dyn.ExtendedSystem.from_cotrol_system(LIP, "x", 9)
    E = ref_tools.plan_steps(9, 1, step_times=np.array([2, 5, 8]))[:, :, None]
    steps = ref_dyn.ExtendedSystem(["Ds"], ["s"], "s", np.ones([9, 1, 1]), [E], ["_x", "_y"])
    bias = ref_dyn.DomainVariable("n", 9, ["_x", "_y"])
    body = ref_body.Formulation()
    body.incorporate_dynamics("steps", steps)
    body.incorporate_dynamics("LIP", LIP_ext)
    body.incorporate_dynamics("n", bias)
    body.incorporate_definitions({
        "DCM_x": ref_comb.LineCombo({"CoM_x": 1, "CoM_dot_x": 1 / 3.5}),
        "DCM_y": ref_comb.LineCombo({"CoM_y": 1, "CoM_dot_y": 1 / 3.5}),
    })
    body.incorporate_constraint("kinematics", ref_res.Constraint("CoM_x", 4))
    body.incorporate_constraint(
        "steppingArea", ref_res.Constraint("s", 10, arrow=[1, 1], axes=["_x", "_y"], L=np.eye(9)))
    body.incorporate_goal("velocity", ref_goal.Cost("CoM_dot", aim=[1, 2], weight=10, axes=["_x", "_y"]))
    body.incorporate_goal("stability", ref_goal.Cost("DCM_x", aim=2, weight=1))
    body.incorporate_goal(
        "terminal", ref_goal.Cost("CoM", aim=[50, 50], weight=100, axes=["_x", "_y"], schedule=range(8, 9)))
    body.incorporate_goal("crossed", ref_goal.Cost("DCM_x", aim=2, weight=1, cross="s_y", cross_L=np.eye(9)))
    vertices = np.array([[0, 1], [1, 0], [0, -1], [-1, 0]])
    body.incorporate_box("kine", ref_res.Box.task_space("CoM", vertices, ["_x", "_y"]))
    body.identify_qp_domain(["x0_x", "x0_y", "Ds_x", "Ds_y", "CoM_dddot_x", "CoM_dddot_y"])
    body.make_preview_matrices()
    given = np.random.default_rng(3).normal(size=[body.given_len, 1])
    A, h, Q, q = body.generate_all_qp_matrices(given)
    cQ, cq = body.generate_qp_cost(body.goals["crossed"], given)
    A1, h1 = body.generate_qp_constraint(body.constraints["steppingArea"][0], given)
    save("test_body.npz", given=given, A=A, h=h, Q=Q, q=q, cQ=cQ, cq=cq, A1=A1, h1=h1, **layout(body))


def generic_case(name, seed, nx, nu, N, n_rows, batch):
    A, B, Q, R = synthetic.generic_system(seed, nx, nu)
    rows = np.random.default_rng(seed + 1).normal(size=[n_rows, nx])
    vals = synthetic.generic_inputs(seed, batch, nx)
    As, hs, Qs, qs = [], [], [], []
    for b in range(batch):
        form = formulate_generic(A, B, Q, R, N, rows, refs=list(vals["ref"][b]), api=RefAPI)
        given = form.arrange_given({"x0": vals["x0"][b][:, None]})
        Am, h, Qm, q = form.generate_all_qp_matrices(given)
        As.append(Am), hs.append(h[:, 0]), Qs.append(Qm), qs.append(q[:, 0])
    save(name, seed=seed, A_sys=A, B_sys=B, Q_w=Q, R_w=R, rows=rows, x0=vals["x0"], ref=vals["ref"],
         A=np.stack(As), h=np.stack(hs), Q=np.stack(Qs), q=np.stack(qs), N=N, **layout(form))


def generic_per_instance(name, seed, nx, nu, N, n_rows, batch):
    """Every instance has its own plant (A_b, B_b): mode B of DESIGN.md."""
    _, _, Q, R = synthetic.generic_system(seed, nx, nu)
    A, B = synthetic.generic_plants(seed, batch, nx, nu)
    rows = np.random.default_rng(seed + 1).normal(size=[n_rows, nx])
    vals = synthetic.generic_inputs(seed, batch, nx)
    As, hs, Qs, qs = [], [], [], []
    for b in range(batch):
        form = formulate_generic(A[b], B[b], Q, R, N, rows, refs=list(vals["ref"][b]), api=RefAPI)
        given = form.arrange_given({"x0": vals["x0"][b][:, None]})
        Am, h, Qm, q = form.generate_all_qp_matrices(given)
        As.append(Am), hs.append(h[:, 0]), Qs.append(Qm), qs.append(q[:, 0])
    save(name, seed=seed, A_sys=A, B_sys=B, Q_w=Q, R_w=R, rows=rows, x0=vals["x0"], ref=vals["ref"],
         A=np.stack(As), h=np.stack(hs), Q=np.stack(Qs), q=np.stack(qs), N=N, **layout(form))


def _digests(out, tag, qp, seed):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _cases

    A, h, Q, q = qp
    out.update({tag + k: v for k, v in _cases.digest(A, h, Q, q, seed).items()})


def biped_long(name, horizons, seed):
    """Long horizons (given_len = 8 + 2N = 408 / 808): one instance each with its own
    omega and aims; stored as digests (tests/_cases.py: digest) because A is 12 MB at
    N = 400."""
    out = {"horizons": np.array(horizons), "seed": seed}
    for N in horizons:
        form = formulate_biped(biped_conf(N))
        vals = synthetic.biped_inputs(seed + N, 1, N)
        given = synthetic.pack_given(form.given_ID, form.given_len, vals, 1)
        set_omega(form, vals["omega"][0])
        form.goals["track vel_x"].update(aim=vals["vel_x"][0])
        form.goals["track vel_y"].update(aim=vals["vel_y"][0])
        _digests(out, "n%d_" % N, form.generate_all_qp_matrices(given[0][:, None]), seed + N)
        out.update({"n%d_%s" % (N, k): v for k, v in layout(form).items()})
    save(name, **out)


def generic_per_instance_digest(name, seed, nx, nu, N, n_rows, batch):
    """Per-instance plants at the BASELINE shapes C4 / C5, stored as digests."""
    _, _, Q, R = synthetic.generic_system(seed, nx, nu)
    A, B = synthetic.generic_plants(seed, batch, nx, nu)
    rows = np.random.default_rng(seed + 1).normal(size=[n_rows, nx])
    vals = synthetic.generic_inputs(seed, batch, nx)
    out = dict(seed=seed, A_sys=A, B_sys=B, Q_w=Q, R_w=R, rows=rows, x0=vals["x0"], ref=vals["ref"], N=N)
    for b in range(batch):
        form = formulate_generic(A[b], B[b], Q, R, N, rows, refs=list(vals["ref"][b]), api=RefAPI)
        given = form.arrange_given({"x0": vals["x0"][b][:, None]})
        _digests(out, "i%d_" % b, form.generate_all_qp_matrices(given), seed + b)
    out.update(layout(form))
    save(name, **out)


def lip_per_instance(name, N, batch, seed):
    """examples/lip.py on the reference's classes, one formulation per instance: its own
    pendulum frequency and sampling period enter the PLANT
    (ControlSystem.from_name("dP->CCC", tau=..., omega=...), tools.py:253-280) and the
    definitions b, DCM."""
    from mpc_interface_b200.examples import lip

    api = types.SimpleNamespace(ControlSystem=ref_dyn.ControlSystem, ExtendedSystem=ref_dyn.ExtendedSystem,
                                Formulation=ref_body.Formulation, LineCombo=ref_comb.LineCombo,
                                Cost=ref_goal.Cost, Box=ref_res.Box,
                                make_simetric_vertices=ref_tools.make_simetric_vertices)
    vals = lip.lip_inputs(seed, batch)
    conf = lip.default_configuration(N)
    As, hs, Qs, qs, Asys, Bsys = [], [], [], [], [], []
    for b in range(batch):
        form = lip.formulate_lip(conf, omega=float(vals["omega"][b]), tau=float(vals["tau"][b]),
                                 target_vel=[float(vals["vel_x"][b]), float(vals["vel_y"][b])], api=api)
        given = form.arrange_given({"x0_x": vals["x0_x"][b][:, None], "x0_y": vals["x0_y"][b][:, None]})
        A, h, Q, q = form.generate_all_qp_matrices(given)
        As.append(A), hs.append(h[:, 0]), Qs.append(Q), qs.append(q[:, 0])
        system = ref_dyn.ControlSystem.from_name("dP->CCC", tau=float(vals["tau"][b]), omega=float(vals["omega"][b]))
        Asys.append(np.array(system.A, dtype=float)), Bsys.append(np.array(system.B, dtype=float))
    save(name, seed=seed, N=N, A=np.stack(As), h=np.stack(hs), Q=np.stack(Qs), q=np.stack(qs),
         A_sys=np.stack(Asys), B_sys=np.stack(Bsys), **vals, **layout(form))


def biped_plants(name, N, batch, seed, system="dP->CCC", tick=None):
    """examples/biped.py (the structure of biped_formulation.py:23-191: steps system, bias
    through its matrix coefficient, three boxes) with ONE PLANT PER INSTANCE -- its own
    pendulum frequency and sampling period in ControlSystem.from_name(system, tau_b, omega_b)
    -- formulated per instance on the reference's classes; ``tick`` = kargs of
    Formulation.update (another phase of the walking cycle)."""
    from mpc_interface_b200.examples import biped

    api = types.SimpleNamespace(ControlSystem=ref_dyn.ControlSystem, ExtendedSystem=ref_dyn.ExtendedSystem,
                                DomainVariable=ref_dyn.DomainVariable, Formulation=ref_body.Formulation,
                                LineCombo=ref_comb.LineCombo, Cost=ref_goal.Cost, Box=ref_res.Box,
                                tools=ref_tools)
    vals = synthetic.biped_inputs(seed, batch, N)
    vals["tau"] = np.random.default_rng(seed + 5).uniform(0.08, 0.12, batch)
    conf = biped.default_configuration(N)
    conf.system = system
    As, hs, Qs, qs, Asys, Bsys = [], [], [], [], [], []
    for b in range(batch):
        form = biped.formulate_biped(conf, omega=float(vals["omega"][b]), tau=float(vals["tau"][b]),
                                     target_vel=[float(vals["vel_x"][b]), float(vals["vel_y"][b])], api=api)
        if tick is not None:
            form.update(**tick)
        collector = {v: np.asarray(vals[v][b]).reshape([-1, 1]) for v in form.given_ID}
        A, h, Q, q = form.generate_all_qp_matrices(form.arrange_given(collector))
        As.append(A), hs.append(h[:, 0]), Qs.append(Q), qs.append(q[:, 0])
        system_b = ref_dyn.ControlSystem.from_name(system, tau=float(vals["tau"][b]), omega=float(vals["omega"][b]))
        Asys.append(np.array(system_b.A, dtype=float)), Bsys.append(np.array(system_b.B, dtype=float))
    extra = {} if tick is None else {"tick_" + k: np.asarray(v) for k, v in tick.items()}
    save(name, seed=seed, N=N, A=np.stack(As), h=np.stack(hs), Q=np.stack(Qs), q=np.stack(qs),
         A_sys=np.stack(Asys), B_sys=np.stack(Bsys), **vals, **extra, **layout(form))


def soft_lip_plants(name, N, batch, seed, order):
    """examples/soft_lip.py (slack variable, late disturbance, a cost with a non-zero aim over
    a plant part and an auxiliary part, row-wise arrows) with one dP->CCC plant per instance,
    every instance formulated on the reference's classes."""
    from mpc_interface_b200.examples import soft_lip

    api = types.SimpleNamespace(ControlSystem=ref_dyn.ControlSystem, ExtendedSystem=ref_dyn.ExtendedSystem,
                                DomainVariable=ref_dyn.DomainVariable, Formulation=ref_body.Formulation,
                                LineCombo=ref_comb.LineCombo, Cost=ref_goal.Cost, Constraint=ref_res.Constraint)
    rng = np.random.default_rng(seed)
    vals = {"x0_x": rng.normal(size=[batch, 3]) * [0.03, 0.1, 0.2], "d_x": rng.normal(size=[batch, N]) * 0.01,
            "omega": np.sqrt(9.81 / rng.uniform(0.75, 1.0, batch)), "tau": rng.uniform(0.08, 0.12, batch),
            "vel": rng.uniform(0.0, 0.4, batch)}
    As, hs, Qs, qs, Asys, Bsys = [], [], [], [], [], []
    for b in range(batch):
        form = soft_lip.formulate_soft_lip(N, float(vals["omega"][b]), float(vals["tau"][b]),
                                           float(vals["vel"][b]), order=order, api=api)
        collector = {v: np.asarray(vals[v][b]).reshape([-1, 1]) for v in form.given_ID}
        A, h, Q, q = form.generate_all_qp_matrices(form.arrange_given(collector))
        As.append(A), hs.append(h[:, 0]), Qs.append(Q), qs.append(q[:, 0])
        system_b = ref_dyn.ControlSystem.from_name("dP->CCC", ["_x"], tau=float(vals["tau"][b]),
                                                   omega=float(vals["omega"][b]))
        Asys.append(np.array(system_b.A, dtype=float)), Bsys.append(np.array(system_b.B, dtype=float))
    save(name, seed=seed, N=N, order=np.array(order), A=np.stack(As), h=np.stack(hs), Q=np.stack(Qs),
         q=np.stack(qs), A_sys=np.stack(Asys), B_sys=np.stack(Bsys), **vals, **layout(form))


def random_forms(name, seeds, batch):
    """tests/_random_forms.py built by the reference, one formulation per instance."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _random_forms as rf

    out = {"seeds": np.array(seeds), "batch": batch}
    for seed in seeds:
        vals = rf.random_scalars(seed, batch)
        As, hs, Qs, qs = [], [], [], []
        for b in range(batch):
            form = rf.build(RefAPI, seed, {k: float(v[b]) for k, v in vals.items()})
            given = form.arrange_given(rf.given_collector(form, seed))
            A, h, Q, q = form.generate_all_qp_matrices(given)
            As.append(A), hs.append(h[:, 0]), Qs.append(Q), qs.append(q[:, 0])
        tag = "s%d_" % seed
        out.update({tag + "given": given[:, 0], tag + "A": np.stack(As), tag + "h": np.stack(hs),
                    tag + "Q": np.stack(Qs), tag + "q": np.stack(qs)})
        out.update({tag + k: v for k, v in layout(form).items()})
    save(name, **out)


if __name__ == "__main__":
    if sys.argv[1:] == ["random"]:
        random_forms("random_forms.npz", list(range(12)), 3)
        sys.exit(0)
    if sys.argv[1:] == ["yaw"]:
        biped_yaw("biped_n24_yaw.npz", 24, 5, 1011)
        biped_yaw("biped_n100_yaw.npz", 100, 2, 1012)
        sys.exit(0)
    if sys.argv[1:] == ["lip"]:
        lip_per_instance("lip_perinst.npz", 40, 6, 1030)
        lip_per_instance("lip_perinst_n9.npz", 9, 3, 1031)
        sys.exit(0)
    if sys.argv[1:] == ["biped_plants"]:
        biped_plants("biped_plants_n24.npz", 24, 4, 1040)
        biped_plants("biped_plants_n30_tick.npz", 30, 3, 1041,
                     tick=dict(step_times=np.array([5, 17, 29]), step_count=1))
        biped_plants("biped_plants_jccc_n16.npz", 16, 2, 1042, system="J->CCC")
        sys.exit(0)
    if sys.argv[1:] == ["soft_lip"]:
        soft_lip_plants("soft_lip_n12.npz", 12, 3, 1050, ("cCoP_dot_x", "e_x"))
        soft_lip_plants("soft_lip_n17_slack_first.npz", 17, 2, 1051, ("e_x", "cCoP_dot_x"))
        sys.exit(0)
    if sys.argv[1:] == ["big"]:
        biped_long("biped_long_digest.npz", [200, 400], 1020)
        generic_per_instance_digest("generic_c4_perinst_digest.npz", 1004, 12, 4, 50, 4, 8)
        generic_per_instance_digest("generic_c5_perinst_digest.npz", 1005, 6, 2, 400, 2, 4)
        sys.exit(0)
    if sys.argv[1:] == ["per_instance"]:
        generic_per_instance("generic_perinst_small.npz", 31, 4, 2, 10, 2, 4)
        generic_per_instance("generic_perinst_odd.npz", 32, 3, 3, 7, 1, 3)
        generic_per_instance("generic_perinst_mid.npz", 33, 6, 2, 40, 2, 2)
        sys.exit(0)
    lip_pickle()
    extend_cases()
    biped_c1()
    biped_batch("biped_n24_batch.npz", 24, 6, 1001)
    biped_batch("biped_n24_terminal.npz", 24, 3, 1006, terminal_weight=0.5)
    biped_batch("biped_n100_batch.npz", 100, 2, 1002)
    biped_tick()
    test_body_case()
    generic_case("generic_small.npz", 21, 4, 2, 10, 2, 3)
    generic_case("generic_c4_shape.npz", 1004, 12, 4, 50, 4, 1)
