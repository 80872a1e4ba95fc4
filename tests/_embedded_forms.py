"""Random formulations AROUND a per-instance plant (the embedded layout of lti.lower), drawn
from a seed as plain data first and then built on either API -- this package's classes with
``Param``s (one batch) or the reference's with numbers (one instance) -- so both sides formulate
the same structure.  Used by the CPU lowering fuzz (tests/test_lti_cpu.py) and by the GPU one
(tests/test_gpu_embedded_general.py).

What varies: the named plant (tools.py:240-285), one or two axes, the horizon; an optimised
slack ``e`` (DomainVariable), a footstep-like ExtendedSystem ``Ds -> s`` (S = 1, U a 0/1 matrix,
fewer actions than samples), a given disturbance ``d``; definitions with per-instance
coefficients on plant variables (1/omega, -1/omega^2, ...) and scalar / matrix coefficients on
the others, definitions of definitions; costs with non-zero aims (per-instance ones and
per-instance weights where the cost stays inside the plant); limits with one arrow, an arrow
per row, a facet over both axes, a numeric centre, a per-instance bound, schedules; the order of
the optimisation variables in the QP.
"""
import numpy as np

SYSTEMS = ("dP->CCC", "J->CCC", "dP->CCP", "P->CC")
STATES = {"dP->CCC": ["CoM", "CoM_dot", "CoM_ddot"], "J->CCC": ["CoM", "CoM_dot", "CoM_ddot"],
          "dP->CCP": ["CoM", "CoM_dot", "cCoP"], "P->CC": ["CoM", "CoM_dot"]}
INPUT = {"dP->CCC": "cCoP_dot", "J->CCC": "CoM_dddot", "dP->CCP": "cCoP_dot", "P->CC": "cCoP"}
OMEGA_COEFS = ("inv", "neg_inv2", "times", "num")


def draw(seed, scheduled=False, general=False, with_L=False, turned=False, delayed=False, paramaux=False):
    """The structure as plain data (a dict of lists and numbers).  ``scheduled``: some of the
    costs act on a few samples only (goal.py `schedule`).  ``general``: also costs that the stage
    weights of the fused kernel cannot express -- crossed costs between any two signals, per-instance
    weights and aims on signals with a part outside the plant, a definition that mixes states and
    the input.  ``with_L``: some uncrossed costs and some limits combine their samples through a
    matrix ``L`` (goal.py:88-104, restrictions.py:75-90) -- mode A only, the per-instance path
    refuses ``L``.  ``turned``: the facets over both axes and some one-axis limits are turned / scaled
    per instance (their one-row arrows are rotated by the angle ``vel``, as rotate_in_TS does).
    ``delayed``: definitions with a MATRIX coefficient on a plant signal (a rate ``(I - shift) @ u1``,
    a delayed state).  ``paramaux``: per-instance coefficients (1 / omega, omega) on the variables
    OUTSIDE the plant too."""
    rng = np.random.default_rng(seed)
    system = SYSTEMS[rng.integers(len(SYSTEMS))]
    axes = ["_x", "_y"] if rng.random() < 0.4 else ["_x"]
    N = int(rng.integers(7, 27))
    spec = {"seed": seed, "system": system, "axes": axes, "N": N}
    has_e, has_d, has_s = rng.random() < 0.75, rng.random() < 0.7, rng.random() < 0.5
    if not (has_e or has_d or has_s):
        has_e = True
    spec["has"] = {"e": bool(has_e), "d": bool(has_d), "s": bool(has_s)}
    if has_s:
        k = int(rng.integers(1, max(2, N // 4) + 1))
        starts = np.sort(rng.choice(np.arange(1, N), size=k, replace=False))
        U = np.zeros([N, k])
        for j, t in enumerate(starts):
            U[t:, j] = 1.0
        spec["steps_U"] = U
    plant_vars = STATES[system] + [INPUT[system]]
    aux_vars = [v for v, on in (("e", has_e), ("d", has_d), ("s", has_s)) if on]

    def plant_terms(count):
        names = list(rng.choice(STATES[system], size=min(count, len(STATES[system])), replace=False))
        return [(str(v), OMEGA_COEFS[rng.integers(len(OMEGA_COEFS))], float(rng.uniform(0.3, 1.5))) for v in names]

    def aux_terms():
        terms = []
        for v in aux_vars:
            if rng.random() < 0.7:
                kind = "num"
                if v != "s" and rng.random() < 0.4:
                    kind = ("shift", "diag")[rng.integers(2)]
                elif paramaux and rng.random() < 0.6:
                    kind = ("inv", "times")[rng.integers(2)]
                terms.append((v, kind, float(rng.uniform(-1.5, 1.5))))
        return terms or [(aux_vars[0], "num", -1.0)]

    defs = [("u1", plant_terms(2))]
    defs.append(("mix", [("u1", "num", 1.0)] + aux_terms()))
    if rng.random() < 0.6:
        defs.append(("side", aux_terms()))
    if rng.random() < 0.5:
        defs.append(("mix2", [("mix", "num", float(rng.uniform(0.5, 2.0)))] + plant_terms(1)))
    if delayed:
        defs.append(("rate", [("u1", "rate", float(rng.uniform(0.5, 2.0)))]
                     + ([(aux_vars[0], "num", float(rng.uniform(-1.0, 1.0)))] if rng.random() < 0.5 else [])))
        if rng.random() < 0.5:
            defs.append(("late", [(str(STATES[system][0]), "shift", 1.0), ("rate", "num", float(rng.uniform(-1.0, 1.0)))]))
    if general and rng.random() < 0.7:
        defs.append(("ux", plant_terms(1) + [(INPUT[system], "num", float(rng.uniform(-0.5, 0.5)))]))
    spec["defs"] = defs
    names = plant_vars + [d[0] for d in defs] + aux_vars + (["Ds"] if has_s else [])
    inside = set(plant_vars + ["u1"])

    costs = []
    for i in range(int(rng.integers(2, 6))):
        v = str(names[rng.integers(len(names))])
        per_instance = v in inside and rng.random() < 0.5
        costs.append({"name": "cost{}".format(i), "var": v, "weight": float(rng.uniform(0.01, 5.0)),
                      "weight_param": bool(per_instance and rng.random() < 0.3),
                      "aim": [float(rng.uniform(-0.3, 0.3)) if rng.random() < 0.7 else 0.0 for _ in axes],
                      "aim_param": bool(per_instance)})
    for c in costs:  # some of them on a few samples only (a terminal cost, a window, scattered samples)
        c["schedule"] = None
        rows = spec["steps_U"].shape[1] if c["var"] == "Ds" else N
        if scheduled and rng.random() < 0.45:
            kind = rng.random()
            if kind < 0.4:
                c["schedule"] = (rows - 1, rows, 1)
            elif kind < 0.8:
                a0 = int(rng.integers(0, rows))
                c["schedule"] = (a0, int(rng.integers(a0 + 1, rows + 1)), 1)
            else:
                c["schedule"] = (int(rng.integers(0, min(3, rows))), rows, int(rng.integers(2, 4)))
            if c["var"] not in ("e", "d", "s", "Ds", "side") and rng.random() < 0.5:
                c["weight_param"] = c["aim_param"] = True  # (allowed on any signal with a plant part)
    for c in costs:
        c["cross"] = None
        if general and c["var"] != "Ds":
            if rng.random() < 0.4:
                others = [v for v in names if v not in ("Ds", c["var"])]
                c["cross"] = str(others[rng.integers(len(others))])
                c["cross_aim"] = [float(rng.uniform(-0.2, 0.2)) for _ in axes]
            if rng.random() < 0.4:
                c["weight_param"] = c["aim_param"] = True
    if not any(c["var"] == INPUT[system] for c in costs):
        costs.append({"name": "effort", "var": INPUT[system], "weight": 0.01, "weight_param": False,
                      "aim": [0.0 for _ in axes], "aim_param": False, "schedule": None, "cross": None})
    spec["costs"] = costs

    limits = []
    for i in range(int(rng.integers(2, 7))):
        v = str(names[rng.integers(len(names))])
        rows = spec["steps_U"].shape[1] if v == "Ds" else N
        lim = {"name": "limit{}".format(i), "var": v, "extreme": float(rng.uniform(0.02, 1.0)),
               "extreme_param": bool(v in inside and rng.random() < 0.3), "schedule": None, "center": None}
        if rng.random() < 0.4:
            a = int(rng.integers(0, rows))
            lim["schedule"] = (a, int(rng.integers(a + 1, rows + 1)))
        cnt = rows if lim["schedule"] is None else lim["schedule"][1] - lim["schedule"][0]
        shape = rng.random()
        if len(axes) == 2 and shape < 0.4:
            lim["axes"], lim["arrow"] = list(axes), rng.normal(size=[1, 2])
        else:
            lim["axes"] = [axes[rng.integers(len(axes))]]
            lim["arrow"] = rng.uniform(0.5, 1.5, size=[cnt, 1]) * rng.choice([-1.0, 1.0]) if shape > 0.7 \
                else np.array([[float(rng.choice([-1.0, 1.0]))]])
        if rng.random() < 0.3:
            lim["center"] = rng.uniform(-0.2, 0.2, size=[1, len(lim["axes"])])
        limits.append(lim)
    spec["limits"] = limits

    optim = [INPUT[system] + ax for ax in axes]
    optim += ["e" + ax for ax in axes] if has_e else []
    optim += ["Ds" + ax for ax in axes] if has_s else []
    spec["order"] = [optim[i] for i in rng.permutation(len(optim))]
    for lim in limits:
        lim["turned"] = bool(turned and np.asarray(lim["arrow"]).shape[0] == 1 and rng.random() < 0.7)
        lim["rowwise"] = bool(turned == "rows" and np.asarray(lim["arrow"]).shape[0] > 1)
    if with_L:
        def samples_of(var, schedule):
            rows = spec["steps_U"].shape[1] if var == "Ds" else N
            return rows if schedule is None else len(range(*schedule))

        for c in costs:
            c["L"] = None
            stepped = c["schedule"] is not None and c["schedule"][2] != 1  # (L counts stop - start columns)
            if c["cross"] is None and not stepped and rng.random() < 0.5:
                c["L"] = rng.normal(size=[int(rng.integers(1, 5)), samples_of(c["var"], c["schedule"])])
        for lim in limits:
            lim["L"] = None
            if rng.random() < 0.5:
                schedule = None if lim["schedule"] is None else lim["schedule"] + (1,)
                lines = int(rng.integers(1, 5))
                lim["L"] = rng.normal(size=[lines, samples_of(lim["var"], schedule)])
                lim["arrow"] = np.array(lim["arrow"])[:1]  # (one arrow for all lines)
                lim["rowwise"] = False
    return spec


def build(spec, api, omega, tau, vel, gain, nominal=(0.1, 3.3), backend=None):
    """The formulation of ``spec`` on ``api`` (ControlSystem, ExtendedSystem, DomainVariable,
    Formulation, LineCombo, Cost, Constraint).  ``omega`` / ``tau``: numbers, or ``Param``s (then
    the structure-only plant is built with ``nominal``); ``vel`` shifts the per-instance aims and
    bounds, ``gain`` scales the per-instance weights."""
    from mpc_interface_b200.params import is_expr

    axes, N, system = spec["axes"], spec["N"], spec["system"]
    lip = api.ControlSystem.from_name(system, axes, tau=nominal[0] if is_expr(tau) else tau,
                                      omega=nominal[1] if is_expr(omega) else omega)
    form = api.Formulation() if backend is None else api.Formulation(backend=backend)
    form.incorporate_dynamics("LIP", api.ExtendedSystem.from_cotrol_system(lip, "x", N))
    if spec["has"]["e"]:
        form.incorporate_dynamics("slack", api.DomainVariable("e", N, axes))
    if spec["has"]["d"]:
        form.incorporate_dynamics("dist", api.DomainVariable("d", N, axes))
    if spec["has"]["s"]:
        form.incorporate_dynamics("steps", api.ExtendedSystem("Ds", "s", "s", S=np.ones([N, 1]),
                                                              U=spec["steps_U"].copy(), axes=axes))
    shift = np.diag(np.ones([N - 1]), 1)
    ramp = np.diag(np.linspace(0.5, 1.5, N))

    def coefficient(kind, value):
        if kind == "inv":
            return value / omega
        if kind == "neg_inv2":
            return -value / omega**2
        if kind == "times":
            return value * omega
        if kind == "rate":
            return value * (np.eye(N) - shift.T)
        if kind == "shift":
            return value * shift
        if kind == "diag":
            return value * ramp
        return value

    definitions = {}
    for name, terms in spec["defs"]:
        for ax in axes:
            definitions[name + ax] = api.LineCombo({v + ax: coefficient(kind, value) for v, kind, value in terms})
    form.incorporate_definitions(definitions)
    for c in spec["costs"]:
        weight = c["weight"] * gain if c["weight_param"] else c["weight"]
        aim = [a + vel if c["aim_param"] else a for a in c["aim"]]
        kargs = {"schedule": range(*c["schedule"])} if c.get("schedule") else {}
        if c.get("L") is not None:
            kargs["L"] = [np.array(c["L"]) for _ in axes]
        if c.get("cross"):
            kargs.update(cross=c["cross"], cross_aim=[x + vel if c["aim_param"] else x for x in c["cross_aim"]])
        form.incorporate_goal(c["name"], api.Cost(c["var"], weight, aim=aim, axes=axes, **kargs))
    for lim in spec["limits"]:
        extreme = lim["extreme"] + vel if lim["extreme_param"] else lim["extreme"]
        kargs = {"axes": lim["axes"], "arrow": np.array(lim["arrow"])}
        if lim.get("turned"):
            from mpc_interface_b200.params import cos, sin

            a = np.array(lim["arrow"])[0]
            if len(a) == 2:  # the facet's normal turned by the angle `vel`
                kargs["arrow"] = np.array([[a[0] * cos(vel) - a[1] * sin(vel), a[0] * sin(vel) + a[1] * cos(vel)]])
            else:            # one axis: the arrow scaled by cos(vel) > 0
                kargs["arrow"] = np.array([[a[0] * cos(vel)]])
        elif lim.get("rowwise"):  # every row its own per-instance scaling (the 2nd one numeric)
            from mpc_interface_b200.params import cos

            column = np.array(lim["arrow"])[:, 0]
            kargs["arrow"] = np.array([[c * cos(vel) * (1 + 0.1 * r) if r != 1 else c] for r, c in enumerate(column)],
                                      dtype=object if hasattr(cos(vel), "evaluate") else float)
        if lim["schedule"] is not None:
            kargs["schedule"] = range(*lim["schedule"])
        if lim["center"] is not None:
            kargs["center"] = np.array(lim["center"])
        if lim.get("L") is not None:
            kargs["L"] = [np.array(lim["L"]) for _ in lim["axes"]]
        form.incorporate_constraint(lim["name"], [api.Constraint(lim["var"], extreme, **kargs)])
    form.identify_qp_domain(list(spec["order"]))
    form.make_preview_matrices()
    return form


def inputs(spec, form, batch, seed):
    """Per-instance values: the given variables of ``form`` (by name) and the parameters
    omega, tau (same preview time for every horizon, so that A^N stays moderate), vel >= 0
    (bounds stay positive), gain."""
    rng = np.random.default_rng(seed)
    vals = {name: rng.normal(size=[batch, len(ids)]) * 0.05 for name, ids in form.given_ID.items()}
    vals["omega"] = np.sqrt(9.81 / rng.uniform(0.75, 1.0, batch))
    vals["tau"] = rng.uniform(0.08, 0.12, batch) * min(1.0, 30.0 / spec["N"])
    vals["vel"] = rng.uniform(0.0, 0.4, batch)
    vals["gain"] = rng.uniform(0.5, 2.0, batch)
    return vals


def own_api():
    from mpc_interface_b200 import body, combinations, dynamics, goal, restrictions
    from types import SimpleNamespace

    return SimpleNamespace(ControlSystem=dynamics.ControlSystem, ExtendedSystem=dynamics.ExtendedSystem,
                           DomainVariable=dynamics.DomainVariable, Formulation=body.Formulation,
                           LineCombo=combinations.LineCombo, Cost=goal.Cost, Constraint=restrictions.Constraint)


def reference_api(ref):
    from types import SimpleNamespace

    return SimpleNamespace(ControlSystem=ref.ControlSystem, ExtendedSystem=ref.ExtendedSystem,
                           DomainVariable=ref.DomainVariable, Formulation=ref.Formulation,
                           LineCombo=ref.LineCombo, Cost=ref.Cost, Constraint=ref.Constraint)
