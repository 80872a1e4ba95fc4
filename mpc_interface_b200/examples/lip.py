"""LIP balance / walking-in-place MPC with the CoP rate as input -- a formulation whose
PLANT depends on the per-instance pendulum frequency and sampling period.

``ControlSystem.from_name("dP->CCC", axes, tau, omega)`` (tools.py:253-280: states CoM,
CoM_dot, CoM_ddot, input cCoP_dot) on both axes; per axis the centre of pressure
``b = CoM - CoM_ddot / omega^2`` and the divergent component ``DCM = CoM + CoM_dot / omega``
(the definitions of biped_formulation.py:62,65); the CoP is kept inside the foot polygon
over the whole horizon (``Box.task_space`` with a safety margin), the last DCM sample inside
it too (``schedule=range(N - 1, N)``, as the biped's terminal constraint), the CoM velocity
tracks a target and the CoP rate and the CoP offset are penalised.  Unlike the biped of the
reference's example there are no footsteps: the support foot stays where it is.

With numbers for ``omega`` / ``tau`` this is ONE instance (``api=`` the reference's own
classes writes the golden vectors).  For a batch pass ``Param``s: the definitions then
carry per-instance coefficients and the plant itself comes per instance through
``generate_all_qp_matrices_batched(..., dynamics={"LIP": ("dP->CCC", tau_b, omega_b)})``.
"""
from types import SimpleNamespace

import numpy as np

from ..params import is_expr

AXES = ["_x", "_y"]


def default_configuration(horizon_lenght=40):
    c = SimpleNamespace()
    c.system = "dP->CCC"
    c.horizon_lenght = horizon_lenght
    c.mpc_period = 0.1
    c.omega = float(np.sqrt(9.81 / 0.877))
    c.foot_corner = np.array([0.1, 0.05])
    c.cop_safety_margin = 0.02
    c.target_vel = np.array([0.2, 0.0])
    c.cost_weights = {"track velocity": 1.0, "minimize CoP rate": 0.001, "center CoP": 0.1}
    return c


def formulate_lip(conf, omega=None, tau=None, target_vel=None, backend=None, api=None, nominal=None, yaw=None):
    """``omega`` / ``tau`` / ``target_vel`` default to the configuration's numbers; ``Param``s
    make them per-instance (then ``nominal = (tau, omega)`` numbers build the structure-only
    plant the formulation is created with).  ``yaw`` (a number or a ``Param``) turns the foot
    polygon and the terminal box in the task space (restrictions.py:380-401 ``rotate_in_TS``)."""
    if api is None:
        from .. import body, combinations, dynamics, goal, restrictions, tools

        api = SimpleNamespace(ControlSystem=dynamics.ControlSystem, ExtendedSystem=dynamics.ExtendedSystem,
                              Formulation=body.Formulation, LineCombo=combinations.LineCombo,
                              Cost=goal.Cost, Box=restrictions.Box,
                              make_simetric_vertices=tools.make_simetric_vertices, rotation2D=tools.rotation2D)
    w = conf.omega if omega is None else omega
    T = conf.mpc_period if tau is None else tau
    vel = conf.target_vel if target_vel is None else target_vel
    N = conf.horizon_lenght
    tau_n, omega_n = nominal if nominal is not None else (conf.mpc_period, conf.omega)
    lip = api.ControlSystem.from_name(conf.system, AXES, tau=tau_n if is_expr(T) else T,
                                      omega=omega_n if is_expr(w) else w)
    lip_ext = api.ExtendedSystem.from_cotrol_system(lip, "x", N)

    extra = {}
    for ax in AXES:
        extra["b" + ax] = api.LineCombo({"CoM" + ax: 1, "CoM_ddot" + ax: -1 / w**2})
        extra["DCM" + ax] = api.LineCombo({"CoM" + ax: 1, "CoM_dot" + ax: 1 / w})

    foot = api.make_simetric_vertices(conf.foot_corner)
    support = api.Box.task_space("b", foot, AXES)
    terminal = api.Box.task_space("DCM", foot, AXES, schedule=range(N - 1, N))
    support.set_safety_margin(conf.cop_safety_margin)
    terminal.set_safety_margin(conf.cop_safety_margin)

    if yaw is not None:
        support.rotate_in_TS(api.rotation2D(yaw))
        terminal.rotate_in_TS(api.rotation2D(yaw))

    weights = conf.cost_weights
    goals = {
        "track velocity": api.Cost("CoM_dot", weights["track velocity"], aim=[vel[0], vel[1]], axes=AXES),
        "minimize CoP rate": api.Cost("cCoP_dot", weights["minimize CoP rate"], aim=[0, 0], axes=AXES),
        "center CoP": api.Cost("b", weights["center CoP"], aim=[0, 0], axes=AXES),
    }
    form = api.Formulation() if backend is None else api.Formulation(backend=backend)
    form.incorporate_dynamics("LIP", lip_ext)
    form.incorporate_definitions(extra)
    for name, goal in goals.items():
        form.incorporate_goal(name, goal)
    form.incorporate_box("support_polygon", support)
    form.incorporate_box("terminal_Constraint", terminal)
    form.identify_qp_domain(["cCoP_dot_x", "cCoP_dot_y"])
    form.make_preview_matrices()
    return form


def lip_inputs(seed, batch):
    """Per-instance values: initial states (CoM, CoM_dot, CoM_ddot per axis), pendulum
    frequency, sampling period and velocity targets."""
    rng = np.random.default_rng(seed)
    sigma = np.array([0.03, 0.1, 0.2])
    return {
        "x0_x": rng.normal(size=[batch, 3]) * sigma,
        "x0_y": rng.normal(size=[batch, 3]) * sigma,
        "omega": np.sqrt(9.81 / rng.uniform(0.75, 1.0, batch)),
        "tau": rng.uniform(0.08, 0.12, batch),
        "vel_x": rng.uniform(0.0, 0.4, batch),
        "vel_y": rng.uniform(-0.1, 0.1, batch),
    }
