"""The UNMODIFIED reference as a live oracle and as the timed CPU arm -- TEST INFRASTRUCTURE.

Finds a staged copy of the reference's Python tree (oracle/make_ref.py), imports its
modules as they are and drives them through the reference's own public API, in the way
its example does (python/use_examples/simple_functional_example/biped_formulation.py:23-191
for the biped; SURVEY.md 3.3 / Appendix C probe6 for the generic plants).  Nothing here
restates reference arithmetic -- every number comes out of the reference's own functions.

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.
"""
import importlib
import importlib.util
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
EXAMPLE = os.path.join("use_examples", "simple_functional_example")

_loaded = None


def find():
    """(path of the reference's `python/` directory, where it was found) or (None, None).
    Probe order: the staged copy that travels with the repo, the driver's install slot,
    the read-only checkout of the build container."""
    for path, kind in ((os.path.join(HERE, "_ref", "python"), "oracle/_ref"),
                       (os.path.join(ROOT, "baseline", "_ref", "python"), "baseline/_ref"),
                       ("/root/reference/python", "/root/reference")):
        if os.path.isfile(os.path.join(path, "mpc_interface", "body.py")):
            return path, kind
    return None, None


def available():
    return find()[0] is not None


def synthetic():
    """mpc_interface_b200/examples/synthetic.py (pure NumPy input generation) loaded by
    file path, so that the reference arm imports nothing else of the product package."""
    name = "_mpci_synthetic_inputs"
    if name not in sys.modules:
        path = os.path.join(ROOT, "mpc_interface_b200", "examples", "synthetic.py")
        spec = importlib.util.spec_from_file_location(name, path)
        module = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(module)
        sys.modules[name] = module
    return sys.modules[name]


def load():
    """Namespace of the reference's modules and classes (imported once per process)."""
    global _loaded
    if _loaded is not None:
        return _loaded
    path, kind = find()
    if path is None:
        raise RuntimeError("the reference is not staged: run `python oracle/make_ref.py` in the "
                           "build container (oracle/_ref travels to the GPU box)")
    present = sys.modules.get("mpc_interface")
    if present is not None and not getattr(present, "__file__", "").startswith(path):
        raise RuntimeError("another module named mpc_interface is already imported")
    sys.dont_write_bytecode = True  # the staged tree stays byte-identical to its manifest
    sys.path.insert(0, path)
    sys.path.insert(0, os.path.join(path, EXAMPLE))
    ns = types.SimpleNamespace(path=path, kind=kind)
    for sub in ("tools", "combinations", "dynamics", "goal", "restrictions", "body"):
        setattr(ns, sub, importlib.import_module("mpc_interface." + sub))
    ns.biped_configuration = importlib.import_module("biped_configuration")
    ns.formulate_biped = importlib.import_module("biped_formulation").formulate_biped
    ns.ControlSystem, ns.ExtendedSystem = ns.dynamics.ControlSystem, ns.dynamics.ExtendedSystem
    ns.DomainVariable = ns.dynamics.DomainVariable
    ns.Formulation, ns.LineCombo = ns.body.Formulation, ns.combinations.LineCombo
    ns.Cost, ns.Box, ns.Constraint = ns.goal.Cost, ns.restrictions.Box, ns.restrictions.Constraint
    _loaded = ns
    return ns


# ---------------------------------------------------------------------------
# biped (BASELINE configs[0..2])
# ---------------------------------------------------------------------------
def biped_conf(N, **overrides):
    """The example's configuration module as a namespace with another horizon."""
    ref = load()
    conf = types.SimpleNamespace(
        **{k: v for k, v in vars(ref.biped_configuration).items() if not k.startswith("__")})
    conf.horizon_lenght = N
    for k, v in overrides.items():
        setattr(conf, k, v)
    return conf


def biped_form(N, **overrides):
    return load().formulate_biped(biped_conf(N, **overrides))


def set_omega(form, w):
    """Per-instance pendulum frequency through the reference (SURVEY Appendix C probe8):
    overwrite the two omega-dependent LineCombo coefficients per axis
    (biped_formulation.py:62,65) and rebuild the preview matrices (body.py:149-156)."""
    for ax in ("_x", "_y"):
        d = form.definitions["b" + ax]
        d.matrices[d.variables.index("CoM_ddot" + ax)] = -1 / w**2
        d = form.definitions["DCM" + ax]
        d.matrices[d.variables.index("CoM_dot" + ax)] = 1 / w
    form.make_preview_matrices()


def biped_instance(form, vals, b):
    """(A, h, Q, q) of instance ``b`` of ``vals`` (synthetic.biped_inputs): per-instance
    omega, velocity aims (Cost.update, goal.py:106), arrange_given (body.py:195-207),
    generate_all_qp_matrices (body.py:333-348)."""
    set_omega(form, float(vals["omega"][b]))
    form.goals["track vel_x"].update(aim=float(vals["vel_x"][b]))
    form.goals["track vel_y"].update(aim=float(vals["vel_y"][b]))
    collector = {name: np.asarray(vals[name][b]).reshape([-1, 1]) for name in form.given_ID}
    given = form.arrange_given(collector)
    return form.generate_all_qp_matrices(given)


# ---------------------------------------------------------------------------
# generic plants (BASELINE configs[3..4]) through the reference API
# ---------------------------------------------------------------------------
def generic_form(A, B, Q, R, horizon, rows, refs, u_box=(-1.0, 2.0), row_bound=5.0):
    """x+ = A x + B u with dense stage weights as one crossed Cost per entry of Q and R
    (goal.py:75-80, body.py:292), an input box per input (restrictions.py:498-500) and
    LineCombo + Constraint rows -- SURVEY.md 3.3, Appendix C probe6 -- on the reference's
    own classes."""
    ref = load()
    nx, nu = B.shape
    inputs = ["u%d" % j for j in range(nu)]
    states = ["z%d" % i for i in range(nx)]
    system = ref.ControlSystem(inputs, states, A, B)
    extended = ref.ExtendedSystem.from_cotrol_system(system, "x", horizon)
    form = ref.Formulation()
    form.incorporate_dynamics("plant", extended)
    for i in range(nx):
        for j in range(nx):
            if Q[i, j] == 0.0:
                continue
            if i == j:
                cost = ref.Cost(states[i], Q[i, i], aim=[refs[i]])
            else:
                cost = ref.Cost(states[i], Q[i, j], aim=[refs[i]], cross=states[j], cross_aim=[refs[j]])
            form.incorporate_goal("Q_%d_%d" % (i, j), cost)
    for i in range(nu):
        for j in range(nu):
            if R[i, j] == 0.0:
                continue
            cost = ref.Cost(inputs[i], R[i, j]) if i == j else ref.Cost(inputs[i], R[i, j], cross=inputs[j])
            form.incorporate_goal("R_%d_%d" % (i, j), cost)
    for j in range(nu):
        form.incorporate_box("box_" + inputs[j],
                             ref.Box.task_space(inputs[j], np.array([[u_box[0]], [u_box[1]]])))
    for r, coefficients in enumerate(np.atleast_2d(rows)):
        name = "lin%d" % r
        form.incorporate_definition(
            name, ref.LineCombo({states[i]: float(c) for i, c in enumerate(coefficients)}))
        form.incorporate_constraint(name, ref.Constraint(name, row_bound))
    form.identify_qp_domain(inputs)
    form.make_preview_matrices()
    return form


def generic_instance(A, B, Q, R, horizon, rows, ref_b, x0_b):
    """(A, h, Q, q) of one instance with its own plant, formulated from scratch."""
    form = generic_form(A, B, Q, R, horizon, rows, [float(x) for x in ref_b])
    given = form.arrange_given({"x0": np.asarray(x0_b, dtype=float).reshape([-1, 1])})
    return form.generate_all_qp_matrices(given)


class GenericPlants:
    """The cheapest way the reference's API offers to assemble a stream of instances that
    each have their own plant (A_b, B_b): keep ONE Formulation with its 160 / 40 costs,
    boxes and limits; per instance build the plant's ExtendedSystem
    (dynamics.py:216-246 -> tools.extend_matrices, tools.py:14-33), re-incorporate it
    under the same name (body.py:38-47 overwrites its definitions in place), set the aims
    (goal.py:106-136), rebuild the preview matrices (body.py:149-156) and generate
    (body.py:333-348).  Bit-identical to ``generic_instance`` (tests/test_oracle_cpu.py)."""

    def __init__(self, A, B, Q, R, horizon, rows):
        self.Q, self.horizon = Q, horizon
        nx, nu = B.shape
        self.inputs = ["u%d" % j for j in range(nu)]
        self.states = ["z%d" % i for i in range(nx)]
        self.form = generic_form(A, B, Q, R, horizon, rows, [0.0] * nx)

    def instance(self, A, B, ref_b, x0_b):
        ref, form, Q = load(), self.form, self.Q
        system = ref.ControlSystem(self.inputs, self.states, A, B)
        form.incorporate_dynamics(
            "plant", ref.ExtendedSystem.from_cotrol_system(system, "x", self.horizon))
        nx = len(self.states)
        for i in range(nx):
            for j in range(nx):
                if Q[i, j] == 0.0:
                    continue
                goal = form.goals["Q_%d_%d" % (i, j)]
                if i == j:
                    goal.update(aim=[float(ref_b[i])])
                else:
                    goal.update(aim=[float(ref_b[i])], cross_aim=np.array([[float(ref_b[j])]]))
        form.make_preview_matrices()
        given = form.arrange_given({"x0": np.asarray(x0_b, dtype=float).reshape([-1, 1])})
        return form.generate_all_qp_matrices(given)
