"""A soft-constrained LIP -- an embedded per-instance formulation that is NOT the biped.

Plant ``ControlSystem.from_name("dP->CCC", ["_x"], tau, omega)`` (tools.py:253-280) on one axis,
one (tau_b, omega_b) per instance when ``Param``s are passed; next to it an OPTIMISED slack ``e``
(a ``DomainVariable`` of N samples, dynamics.py:320-345), a GIVEN disturbance ``d`` that enters
one sample late through a matrix coefficient, and the signal ``soft = b + S d - e`` with the
centre of pressure ``b = CoM - CoM_ddot / omega^2``.  Costs on the slack alone, on ``soft`` with
a NON-ZERO aim (a cost with a plant part and an auxiliary part), on the CoM velocity
(per-instance aim) and on the input; limits on ``soft`` with one arrow, with an arrow per row,
on the slack alone (rows without a plant part) and a scheduled one on ``b``.
``api=`` the reference's own classes formulates one instance the reference's way (golden
vectors, live checks).
"""
from types import SimpleNamespace

import numpy as np

from ..params import is_expr


def _own_api():
    from .. import body, combinations, dynamics, goal, restrictions

    return SimpleNamespace(ControlSystem=dynamics.ControlSystem, ExtendedSystem=dynamics.ExtendedSystem,
                           DomainVariable=dynamics.DomainVariable, Formulation=body.Formulation,
                           LineCombo=combinations.LineCombo, Cost=goal.Cost, Constraint=restrictions.Constraint)


def formulate_soft_lip(N, omega, tau, vel, nominal=(0.1, 3.3), order=("cCoP_dot_x", "e_x"), api=None, backend=None):
    """``order``: the optimisation variables' order in the QP (the plant's input first, or the slack)."""
    api = _own_api() if api is None else api
    axes = ["_x"]
    lip = api.ControlSystem.from_name("dP->CCC", axes, tau=nominal[0] if is_expr(tau) else tau,
                                      omega=nominal[1] if is_expr(omega) else omega)
    ext = api.ExtendedSystem.from_cotrol_system(lip, "x", N)
    slack = api.DomainVariable("e", N, axes)
    dist = api.DomainVariable("d", N, axes)
    shift = np.diag(np.ones([N - 1]), 1)
    arrows = np.linspace(0.5, 1.5, N).reshape([N, 1])
    form = api.Formulation() if backend is None else api.Formulation(backend=backend)
    form.incorporate_dynamics("LIP", ext)
    form.incorporate_dynamics("slack", slack)
    form.incorporate_dynamics("dist", dist)
    form.incorporate_definitions({
        "b_x": api.LineCombo({"CoM_x": 1, "CoM_ddot_x": -1 / omega**2}),
        "soft_x": api.LineCombo({"b_x": 1, "d_x": shift, "e_x": -1}),
    })
    form.incorporate_goal("slack", api.Cost("e", 10.0, aim=[0], axes=axes))
    form.incorporate_goal("soft", api.Cost("soft", 0.5, aim=[0.1], axes=axes))
    form.incorporate_goal("vel", api.Cost("CoM_dot", 1.0, aim=[vel], axes=axes))
    form.incorporate_goal("rate", api.Cost("cCoP_dot", 0.01, aim=[0], axes=axes))
    form.incorporate_constraint("upper", [api.Constraint("soft", 0.05, axes=axes)])
    form.incorporate_constraint("lower", [api.Constraint("soft", 0.07, axes=axes, arrow=[[-1.0]])])
    form.incorporate_constraint("rowwise", [api.Constraint("soft", 0.2, axes=axes, arrow=arrows)])
    form.incorporate_constraint("slack box", [api.Constraint("e", 1.0, axes=axes),
                                              api.Constraint("e", 0.0, axes=axes, arrow=[[-1.0]])])
    form.incorporate_constraint("terminal", [api.Constraint("b", 0.03, axes=axes, schedule=range(N - 2, N))])
    form.identify_qp_domain(list(order))
    form.make_preview_matrices()
    return form
