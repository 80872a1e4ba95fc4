"""Biped walking MPC: LIP centre of mass driven by jerk + footsteps.

The same problem as the reference's example
(python/use_examples/simple_functional_example/biped_formulation.py:23-191
with biped_configuration.py:11-52), written against this package.  ``omega``
may be a float (one instance, as in the reference) or a ``Param`` so that the
pendulum frequency differs between the instances of a batch; the velocity aims
may be ``Param``s as well.

``conf.system`` may also be "dP->CCC" (tools.py:253-280: the same states CoM, CoM_dot,
CoM_ddot driven by the CoP rate): that plant depends on the pendulum frequency, so a batch
with one ``omega`` / ``tau`` per instance needs one plant per instance --
``generate_all_qp_matrices_batched(..., dynamics={"LIP": ("dP->CCC", tau_b, omega_b)})``,
with ``Param``s for ``omega`` / ``tau`` here (``nominal`` numbers then build the structure).
``api=`` a namespace of the reference's own classes builds the identical formulation with
them (the oracle of the parity tests).
"""
from types import SimpleNamespace

import numpy as np

from .. import tools as now
from ..body import Formulation
from ..combinations import LineCombo
from ..dynamics import ControlSystem, DomainVariable, ExtendedSystem
from ..goal import Cost
from ..params import is_expr
from ..restrictions import Box

AXES = ["_x", "_y"]
OPTIM = ["CoM_dddot_x", "Ds_x", "CoM_dddot_y", "Ds_y"]


def _own_api():
    return SimpleNamespace(ControlSystem=ControlSystem, ExtendedSystem=ExtendedSystem,
                           DomainVariable=DomainVariable, Formulation=Formulation, LineCombo=LineCombo,
                           Cost=Cost, Box=Box, tools=now)


def default_configuration(horizon_lenght=None):
    """Values of biped_configuration.py:11-52; ``horizon_lenght`` overrides
    the default of num_steps * step_samples = 24."""
    c = SimpleNamespace()
    c.system = "J->CCC"
    c.num_steps = 2
    c.mpc_period = 0.1
    c.ss_duration = round(11 * c.mpc_period, 5)
    c.ds_duration = c.mpc_period
    c.regular_step_duration = round(c.ss_duration + c.ds_duration, 5)
    c.step_samples = int(round(c.regular_step_duration / c.mpc_period))
    c.horizon_lenght = c.num_steps * c.step_samples if horizon_lenght is None else horizon_lenght
    c.feet_sep = 0.08
    c.foot_corner = np.array([0.1, 0.05])
    c.stepping_corner = np.array([0.3, 0.1])
    c.stepping_center = np.array([0, 2 * c.foot_corner[1] + c.stepping_corner[1] + c.feet_sep])
    c.strt_y = c.feet_sep / 2 + c.foot_corner[1]
    c.com_height = 0.877
    c.gravity = [0, 0, -9.81]
    c.omega = np.sqrt(-c.gravity[2] / c.com_height)
    c.cop_safety_margin = 0.02
    c.target_vel = np.array([0.6, 0, 0])
    c.cost_weights = {
        "minimize jerk": 0.001,
        "track velocity": 0.01,
        "relax ankles": 1,
        "terminal": 0,
    }
    return c


def formulate_biped(conf, omega=None, target_vel=None, backend=None, preview=True, yaw=None,
                    tau=None, nominal=None, api=None):
    """Builds the Formulation.  ``omega`` / ``target_vel`` default to the
    configuration's numbers; pass ``Param``s to make them per-instance.
    ``yaw`` (float or ``Param``) turns the support polygon and the terminal box
    by that angle with ``Box.rotate_in_TS`` (restrictions.py:437-457): the feet of
    every instance of a batch may point somewhere else.
    ``tau``: the sampling period of the plant (default ``conf.mpc_period``; a ``Param`` for a
    per-instance plant -- ``nominal = (tau, omega)`` numbers then build the structure)."""
    now, ExtendedSystem, ControlSystem, DomainVariable, LineCombo, Box, Cost, Formulation = (
        lambda a: (a.tools, a.ExtendedSystem, a.ControlSystem, a.DomainVariable, a.LineCombo, a.Box,
                   a.Cost, a.Formulation))(_own_api() if api is None else api)
    w = conf.omega if omega is None else omega
    period = conf.mpc_period if tau is None else tau
    vel = conf.target_vel if target_vel is None else target_vel
    N = conf.horizon_lenght
    tau_n, omega_n = nominal if nominal is not None else (conf.mpc_period, conf.omega)

    # footsteps: s = 1*s0 + E*Ds with a 0/1 step-indicator matrix
    steps = ExtendedSystem(
        "Ds",
        "s",
        "s",
        S=np.ones([N, 1]),
        U=now.plan_steps(N, 0, regular_time=conf.step_samples),
        axes=AXES,
        how_to_update_matrices=now.update_step_matrices,
        time_variant=True,
    )
    steps.define_output("stamps", {"s0": 1, "Ds": 1}, time_variant=True, how_to_update=now.adapt_size)

    # centre of mass
    lip = ControlSystem.from_name(conf.system, AXES, tau=tau_n if is_expr(period) else period,
                                  omega=omega_n if is_expr(w) else (conf.omega if omega is None else w))
    lip_ext = ExtendedSystem.from_cotrol_system(lip, "x", N)
    drive = lip.input_names[0]  # "CoM_dddot" (J->CCC) or "cCoP_dot" (dP->CCC)

    # non-linearity bias, one sample per horizon step
    bias = DomainVariable("n", N, AXES)

    shift = np.diag(np.ones([N - 1]), 1)  # n enters one sample later
    extra = {}
    if conf.system in ("J->CCC", "dP->CCC"):  # both carry CoM, CoM_dot, CoM_ddot
        for ax in AXES:
            extra["b" + ax] = LineCombo({"CoM" + ax: 1, "CoM_ddot" + ax: -1 / w**2})
            extra["(b+n-s)" + ax] = LineCombo({"b" + ax: 1, "n" + ax: shift, "s" + ax: -1})
            extra["(c-s)" + ax] = LineCombo({"CoM" + ax: 1, "s" + ax: -1})
            extra["DCM" + ax] = LineCombo({"CoM" + ax: 1, "CoM_dot" + ax: 1 / w})
            extra["(DCM-s)" + ax] = LineCombo({"DCM" + ax: 1, "s" + ax: -1})

    foot = now.make_simetric_vertices(conf.foot_corner)
    support_polygon = Box.task_space("(b+n-s)", foot, AXES)
    stepping_area = Box.task_space(
        "Ds",
        now.make_simetric_vertices(conf.stepping_corner),
        AXES,
        how_to_update=now.update_stepping_area,
        time_variant=True,
    )
    stepping_area.update(
        step_count=0, n_next_steps=steps.domain["Ds_x"], xy_lenght=conf.stepping_center
    )
    last = range(N - 1, N)
    terminal_constraint = Box.task_space("(DCM-s)", foot, AXES, schedule=last)
    support_polygon.set_safety_margin(conf.cop_safety_margin)
    terminal_constraint.set_safety_margin(conf.cop_safety_margin)
    if yaw is not None:
        support_polygon.rotate_in_TS(now.rotation2D(yaw))
        terminal_constraint.rotate_in_TS(now.rotation2D(yaw))

    weights = conf.cost_weights
    goals = {
        "relax ankles x": Cost("(b+n-s)", weights["relax ankles"], aim=[0], axes=["_x"]),
        "relax ankles y": Cost("(b+n-s)", weights["relax ankles"], aim=[0], axes=["_y"]),
        "minimize jerk": Cost(drive, weights["minimize jerk"], aim=[0, 0], axes=AXES),
        "track vel_x": Cost("CoM_dot", weights["track velocity"], aim=vel[0], axes=["_x"]),
        "track vel_y": Cost("CoM_dot", weights["track velocity"], aim=vel[1], axes=["_y"]),
        "terminal_cost": Cost("(DCM-s)", weights["terminal"], aim=[0, 0], axes=AXES, schedule=last),
    }

    form = Formulation() if (api is not None or backend is None) else Formulation(backend=backend)
    form.incorporate_dynamics("steps", steps)
    form.incorporate_dynamics("LIP", lip_ext)
    form.incorporate_dynamics("bias", bias)
    form.incorporate_definitions(extra)
    for name, goal in goals.items():
        form.incorporate_goal(name, goal)
    form.incorporate_box("stepping area", stepping_area)
    form.incorporate_box("support_polygon", support_polygon)
    form.incorporate_box("terminal_Constraint", terminal_constraint)
    form.identify_qp_domain([drive + "_x", "Ds_x", drive + "_y", "Ds_y"])
    if preview:  # builds the device plan; False leaves a host-only description
        form.make_preview_matrices()

    def tick(body, **kargs):
        """Needs ``step_times`` and ``step_count`` (biped_formulation.py:156-187)."""
        body.dynamics["steps"].update(step_times=kargs["step_times"], N=N)
        body.constraint_boxes["stepping area"].update(
            step_count=kargs["step_count"],
            n_next_steps=body.dynamics["steps"].domain["Ds_x"],
            xy_lenght=conf.stepping_center,
        )

    form.set_updating_rule(tick)
    return form


def structure_key(horizon_lenght):
    """Structural key of one biped instance for ``receding.StructureBuckets``:
    the steps inside the preview window (they fix E and the number of ``Ds``
    variables, tools.py:79-101) and which foot steps next (the side of the
    stepping-area centres, tools.py:158-165)."""
    def key(step_times, step_count, **_):
        times = np.asarray(step_times)
        inside = times[(times >= 0) & (times < horizon_lenght - 1)]
        return (tuple(int(t) for t in inside), int(step_count) % 2)

    return key


def structure_keys(horizon_lenght, step_times, step_count):
    """``structure_key`` for a whole batch at once: step_times int [B, n_steps] (ascending
    per row), step_count int [B] -> int64 [B, n_steps + 1] key rows for
    ``receding.StructureBuckets.group_keys``: the step times inside the preview window
    (the others masked to -1) and the parity of the step count.  Two instances get equal
    rows exactly when ``structure_key`` gives them equal keys."""
    if hasattr(step_times, "is_cuda"):  # CUDA tensors in, a CUDA tensor out (same rows)
        import torch

        times = step_times.to(torch.int64)
        inside = (times >= 0) & (times < horizon_lenght - 1)
        big = torch.iinfo(torch.int64).max
        rows = torch.where(inside, times, torch.full_like(times, big))
        rows = torch.sort(rows, dim=1).values
        rows = torch.where(rows == big, torch.full_like(rows, -1), rows)
        parity = (torch.as_tensor(step_count, device=times.device).to(torch.int64) % 2).reshape(-1, 1)
        return torch.cat([rows, parity], dim=1)
    times = np.asarray(step_times).astype(np.int64)
    inside = (times >= 0) & (times < horizon_lenght - 1)
    # the inside times left-aligned (sorted, the masked ones last), so that a row depends on
    # the SET of inside times alone, as the tuple of structure_key does
    big = np.iinfo(np.int64).max
    rows = np.where(inside, times, big)
    if rows.shape[1] > 1 and not np.all(rows[:, :-1] <= rows[:, 1:]):
        rows = np.sort(rows, axis=1)  # (ascending step times: already in place)
    rows = np.where(rows == big, -1, rows)
    parity = (np.asarray(step_count).astype(np.int64) % 2).reshape(-1, 1)
    return np.concatenate([rows, parity], axis=1)
