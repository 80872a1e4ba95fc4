"""Generic linear MPC ``x+ = A x + B u`` with dense stage weights Q, R, an
input box and linear state rows, expressed through the formulation API the way
SURVEY.md 3.3 describes: one ``Cost`` per (i, j) entry of Q and R using
``cross=`` (goal.py:75-80, body.py:292), ``Box.task_space`` on every input
(restrictions.py:498-500) and ``LineCombo`` + ``Constraint`` rows."""
import numpy as np

from ..body import Formulation
from ..combinations import LineCombo
from ..dynamics import ControlSystem, ExtendedSystem
from ..goal import Cost
from ..restrictions import Box, Constraint


class _Own:
    ControlSystem, ExtendedSystem, Formulation = ControlSystem, ExtendedSystem, Formulation
    LineCombo, Cost, Box, Constraint = LineCombo, Cost, Box, Constraint


def formulate_generic(A, B, Q, R, horizon, rows, refs=None, u_box=(-1.0, 2.0), row_bound=5.0,
                      backend=None, api=None, terminal=None):
    """``rows``: [n_rows, nx] coefficients of the linear state constraints
    ``c' x_k <= row_bound``; ``refs``: per-state aims (numbers or Params).
    ``api`` is any namespace with the formulation classes (default: this
    package) -- the golden-vector script passes the reference's own classes so
    both sides are built by the same recipe.  ``terminal``: [nx, nx] weight of the terminal
    cost ``(x_N - ref)' Qf (x_N - ref)``, the same crossed costs on the last sample only
    (``schedule=range(horizon - 1, horizon)``, goal.py:46-60)."""
    api = _Own if api is None else api
    ControlSystem, ExtendedSystem = api.ControlSystem, api.ExtendedSystem
    LineCombo, Cost, Box, Constraint = api.LineCombo, api.Cost, api.Box, api.Constraint
    nx, nu = B.shape
    inputs = ["u%d" % j for j in range(nu)]
    states = ["z%d" % i for i in range(nx)]
    refs = [0.0] * nx if refs is None else refs

    system = ControlSystem(inputs, states, A, B)
    extended = ExtendedSystem.from_cotrol_system(system, "x", horizon)
    form = api.Formulation() if backend is None else api.Formulation(backend=backend)
    form.incorporate_dynamics("plant", extended)

    for i in range(nx):
        for j in range(nx):
            if Q[i, j] == 0.0:
                continue
            if i == j:
                cost = Cost(states[i], Q[i, i], aim=[refs[i]])
            else:
                cost = Cost(states[i], Q[i, j], aim=[refs[i]], cross=states[j], cross_aim=[refs[j]])
            form.incorporate_goal("Q_%d_%d" % (i, j), cost)
    last = range(horizon - 1, horizon)
    for i in range(nx if terminal is not None else 0):
        for j in range(nx):
            if terminal[i, j] == 0.0:
                continue
            if i == j:
                cost = Cost(states[i], terminal[i, i], aim=[refs[i]], schedule=last)
            else:
                cost = Cost(states[i], terminal[i, j], aim=[refs[i]], cross=states[j], cross_aim=[refs[j]],
                            schedule=last)
            form.incorporate_goal("Qf_%d_%d" % (i, j), cost)
    for i in range(nu):
        for j in range(nu):
            if R[i, j] == 0.0:
                continue
            cost = Cost(inputs[i], R[i, j]) if i == j else Cost(inputs[i], R[i, j], cross=inputs[j])
            form.incorporate_goal("R_%d_%d" % (i, j), cost)

    for j in range(nu):
        form.incorporate_box("box_" + inputs[j], Box.task_space(inputs[j], np.array([[u_box[0]], [u_box[1]]])))
    for r, coefficients in enumerate(np.atleast_2d(rows)):
        name = "lin%d" % r
        form.incorporate_definition(
            name, LineCombo({states[i]: float(c) for i, c in enumerate(coefficients)})
        )
        form.incorporate_constraint(name, Constraint(name, row_bound))

    form.identify_qp_domain(inputs)
    form.make_preview_matrices()
    return form
