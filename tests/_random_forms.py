"""Randomised formulations: one recipe, driven by a seed, that builds the same
problem through ANY namespace holding the formulation classes.  The golden
script runs it with the reference's classes (tests/golden/make_goldens.py
random), the tests with this package's; per-instance scalars arrive in
``vals`` -- plain floats on the reference side (one formulation per instance),
``Param`` objects on ours (one formulation, one batched call).

The recipe walks through the argument space of the reference's API rather than
one application: with and without axes, matrix-valued LineCombo coefficients
and definitions of definitions (combinations.py:24-60), limits with a negative
extreme (flipped on construction, restrictions.py:181-194; numeric -- the sign of
a per-instance extreme is not known to the host), per-row arrows and centres,
schedules and L matrices (restrictions.py:16-66), task-space boxes from vertex
lists (restrictions.py:459-520), costs with aims, schedules, L and crossed terms
(goal.py:11-110), and optimisation domains that mix inputs with other variables
(body.py:213-246)."""
import numpy as np

SCALARS = ("a", "c", "w0", "w1", "e0", "e1", "aim0", "aim1", "ctr")


def random_scalars(seed, batch):
    """Per-instance scalar values, [batch] each."""
    rng = np.random.default_rng(1000 + seed)
    vals = {k: rng.uniform(0.4, 1.6, batch) for k in SCALARS}
    vals["aim1"] = vals["aim1"] - 1.0
    return vals


def structure(seed):
    """Everything that is shared by the instances of one seed."""
    rng = np.random.default_rng(seed)
    s = {}
    s["n"] = int(rng.integers(2, 4))
    s["m"] = int(rng.integers(1, 3))
    s["N"] = int(rng.integers(4, 9))
    s["axes"] = ["_x", "_y"] if rng.random() < 0.6 else None
    n, m, N = s["n"], s["m"], s["N"]
    s["A"] = np.eye(n) + 0.2 * rng.normal(size=[n, n])
    s["B"] = rng.normal(size=[n, m])
    s["M"] = rng.normal(size=[N, N])  # a matrix coefficient inside a LineCombo
    lo = int(rng.integers(0, N - 2))
    s["window"] = range(lo, int(rng.integers(lo + 2, N + 1)))
    t = len(s["window"])
    s["L_rows"] = int(rng.integers(1, 4))
    s["L"] = rng.normal(size=[s["L_rows"], t])
    s["L_cost"] = rng.normal(size=[2, t])
    s["arrow_rows"] = rng.normal(size=[N, 2 if s["axes"] else 1])
    s["row_extremes"] = rng.uniform(0.5, 2.0, size=[N, 1])
    angles = np.sort(rng.uniform(0, 2 * np.pi, int(rng.integers(3, 6))))
    s["polygon"] = np.stack([np.cos(angles), np.sin(angles)], axis=1) * rng.uniform(0.5, 1.5)
    s["segment"] = np.array([[-rng.uniform(0.5, 1.5)], [rng.uniform(0.5, 1.5)]])
    s["bias_optim"] = bool(rng.random() < 0.5)
    # never both: the reference's arrange_given returns a shapeless array for an
    # empty given vector (body.py:205-206) and its row stacking then fails
    s["x0_optim"] = bool(rng.random() < 0.4) and not s["bias_optim"]
    s["x0"] = rng.normal(size=[n])
    s["bias"] = rng.normal(size=[N]) * 0.3
    return s


def build(api, seed, vals):
    """``vals``: {name: float | Param} for SCALARS."""
    s = structure(seed)
    n, m, N, axes = s["n"], s["m"], s["N"], s["axes"]
    ax = axes if axes else [""]
    inputs = ["u%d" % j for j in range(m)]
    states = ["z%d" % i for i in range(n)]

    plant = api.ControlSystem(inputs, states, s["A"], s["B"], axes=axes)
    extended = api.ExtendedSystem.from_cotrol_system(plant, "x", N)
    bias = api.DomainVariable("n", N, axes)
    form = api.Formulation()
    form.incorporate_dynamics("plant", extended)
    form.incorporate_dynamics("bias", bias)

    for x in ax:
        form.incorporate_definition(
            "d0" + x, api.LineCombo({"z0" + x: vals["a"], "z1" + x: -0.5, "n" + x: vals["c"]}))
        # a definition of a definition, with a matrix coefficient on an input
        form.incorporate_definition("d1" + x, api.LineCombo({"d0" + x: 1.0, "u0" + x: s["M"]}))

    k = len(ax)
    # limits ------------------------------------------------------------------
    form.incorporate_constraint(
        "plain", api.Constraint("d0", vals["e0"], axes=axes, arrow=[1.0, -0.7][:k]))
    form.incorporate_constraint(
        "flipped", api.Constraint("z1", -0.8, axes=axes, arrow=[0.3, 1.0][:k],
                                  center=[vals["ctr"], -0.2][:k]))
    form.incorporate_constraint(
        "rows", api.Constraint("d1", s["row_extremes"], axes=axes, arrow=s["arrow_rows"]))
    form.incorporate_constraint(
        "window", api.Constraint("z0", vals["e1"], axes=axes, arrow=[1.0, 1.0][:k], schedule=s["window"]))
    form.incorporate_constraint(
        "mixed", api.Constraint("d0", 2.5, axes=axes, arrow=[-1.0, 0.4][:k], L=s["L"],
                                schedule=s["window"]))
    if axes:
        form.incorporate_box("polygon", api.Box.task_space("z0", s["polygon"], axes))
    else:
        form.incorporate_box("segment", api.Box.task_space("u0", s["segment"]))

    # goals -------------------------------------------------------------------
    form.incorporate_goal("track", api.Cost("d1", vals["w0"], aim=[vals["aim0"], vals["aim1"]][:k], axes=axes))
    form.incorporate_goal("effort", api.Cost(inputs[-1], 0.01, axes=axes))
    form.incorporate_goal(
        "late", api.Cost("z0", vals["w1"], aim=[vals["aim1"], 0.3][:k], axes=axes, schedule=s["window"]))
    form.incorporate_goal(
        "shaped", api.Cost("d0", 0.7, aim=[0.2, vals["aim0"]][:k], axes=axes, L=s["L_cost"],
                           schedule=s["window"]))
    form.incorporate_goal(
        "crossed", api.Cost("d0" + ax[0], 0.25, aim=vals["aim0"], cross="z1" + ax[-1], cross_aim=0.1))

    optim = [u + x for u in inputs for x in ax]
    if s["bias_optim"]:
        optim += ["n" + x for x in ax]
    if s["x0_optim"]:
        optim = ["x0" + x for x in ax] + optim
    form.identify_qp_domain(optim)
    form.make_preview_matrices()
    return form


def given_collector(form, seed):
    """Values of the given variables of one seed (shared by the instances)."""
    s = structure(seed)
    ax = s["axes"] if s["axes"] else [""]
    out = {}
    for i, x in enumerate(ax):
        out["x0" + x] = (s["x0"] * (1 + 0.5 * i))[:, None]
        out["n" + x] = (s["bias"] * (1 - 0.3 * i))[:, None]
    return {k: v for k, v in out.items() if k in form.given_ID}
