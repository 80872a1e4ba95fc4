"""Builders shared by the CPU and GPU tests: formulations that correspond to
the golden files of tests/golden/ (see make_goldens.py for the reference side)."""
import os

import numpy as np

from mpc_interface_b200 import (
    Box, Constraint, ControlSystem, Cost, DomainVariable, ExtendedSystem, Formulation,
    LineCombo, Param,
)
from mpc_interface_b200 import tools as use
from mpc_interface_b200.examples.biped import default_configuration, formulate_biped
from mpc_interface_b200.examples.generic import formulate_generic

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = 1e-10  # north_star: relative error <= 1e-10 on P, q, G, h (normwise, per array)


def golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def rel_err(x, ref):
    x, ref = np.asarray(x, dtype=float), np.asarray(ref, dtype=float)
    assert x.shape == ref.shape, (x.shape, ref.shape)
    if ref.size == 0:
        return 0.0
    scale = np.max(np.abs(ref))
    return float(np.max(np.abs(x - ref)) / (scale if scale > 0 else 1.0))


def assert_qp_close(got, ref, tol=TOL):
    """got / ref: (A, h, Q, q) tuples (any of them batched)."""
    for name, x, r in zip("AhQq", got, ref):
        err = rel_err(np.squeeze(x), np.squeeze(r))
        assert err <= tol, "{}: relative error {:.3e} > {:.1e}".format(name, err, tol)


def digest(A, h, Q, q, seed, rows=8):
    """Small fingerprint of one large QP for the committed goldens (a C5 instance is
    20 MB): h and q in full, ``rows`` sampled rows of A and of Q, and the products of A
    and Q with a seeded random vector -- every entry of both matrices enters those with
    its own random weight, so a wrong entry anywhere moves them."""
    rng = np.random.default_rng(seed)
    rA = np.sort(rng.choice(A.shape[0], min(rows, A.shape[0]), replace=False))
    rQ = np.sort(rng.choice(Q.shape[0], min(rows, Q.shape[0]), replace=False))
    v = rng.normal(size=A.shape[1])
    return {"h": np.ravel(h), "q": np.ravel(q), "A_idx": rA, "A_rows": A[rA], "Q_idx": rQ,
            "Q_rows": Q[rQ], "A_v": A @ v, "Q_v": Q @ v, "Qt_v": Q.T @ v,
            "shape": np.array([A.shape[0], A.shape[1]])}


def assert_digest_close(got, g, prefix, seed, tol=TOL):
    """got = (A, h, Q, q) of one instance against the digest stored under ``prefix``."""
    A, h, Q, q = (np.asarray(x, dtype=float) for x in got)
    assert tuple(g[prefix + "shape"]) == A.shape and Q.shape == (A.shape[1], A.shape[1])
    mine = digest(A, h, Q, q, seed, rows=len(g[prefix + "A_idx"]))
    for key, value in mine.items():
        if key.endswith("idx") or key == "shape":
            assert np.array_equal(value, g[prefix + key]), key
        else:
            err = rel_err(value, g[prefix + key])
            assert err <= tol, "{}{}: relative error {:.3e} > {:.1e}".format(prefix, key, err, tol)


def check_layout(form, g):
    for kind in ("optim", "given"):
        ids = getattr(form, kind + "_ID")
        assert list(ids.keys()) == [str(s) for s in g[kind + "_names"]]
        got = np.array([[r.start, r.stop] for r in ids.values()], dtype=np.int64)
        assert np.array_equal(got, g[kind + "_ranges"])  # bit-exact integer layout


def biped_batched(g, backend=None):
    conf = default_configuration(int(g["N"]))
    if float(g["terminal_weight"]) >= 0:
        conf.cost_weights["terminal"] = float(g["terminal_weight"])
    form = formulate_biped(
        conf, omega=Param("omega"), target_vel=[Param("vel_x"), Param("vel_y")], backend=backend
    )
    params = {k: g[k] for k in ("omega", "vel_x", "vel_y")}
    return form, params


def biped_yaw(g, backend=None):
    """Per-instance feet orientation (golden biped_n*_yaw: the reference's
    Box.rotate_in_TS applied instance by instance)."""
    form = formulate_biped(
        default_configuration(int(g["N"])), omega=Param("omega"),
        target_vel=[Param("vel_x"), Param("vel_y")], yaw=Param("yaw"), backend=backend,
    )
    params = {k: g[k] for k in ("omega", "vel_x", "vel_y", "yaw")}
    return form, params


def body_test_formulation(backend=None):
    """The formulation of the reference's python/tests/test_body.py:23-98."""
    axes = ["_x", "_y"]
    LIP = ControlSystem.from_name("J->CCC", tau=0.1, omega=3.5, axes=axes)
    LIP_ext = ExtendedSystem.from_cotrol_system(LIP, "x", 9)
    E = use.plan_steps(9, 1, step_times=np.array([2, 5, 8]))[:, :, None]
    steps = ExtendedSystem(["Ds"], ["s"], "s", np.ones([9, 1, 1]), [E], axes)
    bias = DomainVariable("n", 9, axes)
    body = Formulation(backend=backend)
    body.incorporate_dynamics("steps", steps)
    body.incorporate_dynamics("LIP", LIP_ext)
    body.incorporate_dynamics("n", bias)
    body.incorporate_definitions(
        {
            "DCM_x": LineCombo({"CoM_x": 1, "CoM_dot_x": 1 / 3.5}),
            "DCM_y": LineCombo({"CoM_y": 1, "CoM_dot_y": 1 / 3.5}),
        }
    )
    body.incorporate_constraint("kinematics", Constraint("CoM_x", 4))
    body.incorporate_constraint(
        "steppingArea", Constraint("s", 10, arrow=[1, 1], axes=axes, L=np.eye(9))
    )
    body.incorporate_goal("velocity", Cost("CoM_dot", aim=[1, 2], weight=10, axes=axes))
    body.incorporate_goal("stability", Cost("DCM_x", aim=2, weight=1))
    body.incorporate_goal(
        "terminal", Cost("CoM", aim=[50, 50], weight=100, axes=axes, schedule=range(8, 9))
    )
    body.incorporate_goal("crossed", Cost("DCM_x", aim=2, weight=1, cross="s_y", cross_L=np.eye(9)))
    vertices = np.array([[0, 1], [1, 0], [0, -1], [-1, 0]])
    body.incorporate_box("kine", Box.task_space("CoM", vertices, axes))
    body.identify_qp_domain(["x0_x", "x0_y", "Ds_x", "Ds_y", "CoM_dddot_x", "CoM_dddot_y"])
    body.make_preview_matrices()
    return body


def generic_batched(g, backend=None):
    nx = g["A_sys"].shape[0]
    refs = [Param("ref%d" % i) for i in range(nx)]
    form = formulate_generic(
        g["A_sys"], g["B_sys"], g["Q_w"], g["R_w"], int(g["N"]), g["rows"], refs=refs,
        backend=backend,
    )
    params = {"ref%d" % i: g["ref"][:, i] for i in range(nx)}
    given = g["x0"]  # the only given variable is x0
    return form, params, given


def tiny_plant(backend=None, optim=("x0", "u"), with_limit=True, N=5):
    """Edge shapes: ng = 0 when x0 is optimised, nc = 0 without limits, odd nv."""
    rng = np.random.default_rng(0)
    A = np.eye(3) + 0.1 * rng.normal(size=[3, 3])
    B = rng.normal(size=[3, 1])
    ext = ExtendedSystem.from_cotrol_system(ControlSystem(["u"], ["a", "b", "c"], A, B), "x", N)
    form = Formulation(backend=backend)
    form.incorporate_dynamics("plant", ext)
    form.incorporate_goal("g1", Cost("a", Param("w"), aim=[Param("aim")]))
    form.incorporate_goal("g2", Cost("u", 0.1, aim=[0.0]))
    if with_limit:
        form.incorporate_constraint("lim", Constraint("b", 3.0))
    form.identify_qp_domain(list(optim))
    form.make_preview_matrices()
    return form


def biped_three_yaws(N, backend=None):
    """Every box of the biped turned by its own per-instance angle: more than 8 streamed G
    monomials (the kernels' generic path)."""
    form = formulate_biped(default_configuration(N), omega=Param("omega"),
                           target_vel=[Param("vel_x"), Param("vel_y")], backend=backend, preview=False)
    for name, angle in (("support_polygon", "yaw_a"), ("terminal_Constraint", "yaw_b"),
                        ("stepping area", "yaw_c")):
        form.constraint_boxes[name].rotate_in_TS(use.rotation2D(Param(angle)))
    form.make_preview_matrices()
    return form


def three_yaw_inputs(form, seed, batch, N):
    from mpc_interface_b200.examples import synthetic

    vals = synthetic.biped_inputs(seed, batch, N)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    rng = np.random.default_rng(seed + 3)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    for k in ("yaw_a", "yaw_b", "yaw_c"):
        params[k] = rng.uniform(-1.0, 1.0, batch)
    return given, params


# ------------------------------------------------------------ randomised formulations
class _OwnAPI:
    ControlSystem, ExtendedSystem, DomainVariable = ControlSystem, ExtendedSystem, DomainVariable
    LineCombo, Cost, Box, Constraint = LineCombo, Cost, Box, Constraint

    def __init__(self, backend):
        self.backend = backend

    def Formulation(self):
        return Formulation() if self.backend is None else Formulation(backend=self.backend)


def random_form(g, seed, backend=None):
    """Seed ``seed`` of tests/_random_forms.py with every scalar a Param; returns
    (form, given [B, ng], params, reference (A, h, Q, q) of golden random_forms)."""
    import _random_forms as rf

    batch = int(g["batch"])
    vals = rf.random_scalars(seed, batch)
    form = rf.build(_OwnAPI(backend), seed, {k: Param(k) for k in vals})
    tag = "s%d_" % seed
    layout = {k: g[tag + k] for k in ("optim_names", "optim_ranges", "given_names", "given_ranges")}
    check_layout(form, layout)
    given = np.repeat(g[tag + "given"][None, :], batch, 0)
    ref = tuple(g[tag + k] for k in "AhQq")
    return form, given, vals, ref


def transform_boxes(form, rotation, shift, scale, recenter):
    """The Box transforms of restrictions.py:380-486 on the biped's boxes, the same calls for this
    package's formulation (arguments may hold Params: one value per robot) and for the
    reference's (numbers): the support polygon turned, shifted and scaled, the terminal box
    re-centred."""
    box = form.constraint_boxes["support_polygon"]
    box.rotate_in_TS(rotation)
    box.translate_in_TS(np.array(list(shift)))
    box.scale_box(scale)
    form.constraint_boxes["terminal_Constraint"].recenter_in_TS(np.array([recenter, 0.0]))
    return form


def transform_values(batch, seed):
    rng = np.random.default_rng(seed)
    return {"yaw": rng.uniform(-np.pi, np.pi, batch), "dx": rng.uniform(-0.03, 0.03, batch),
            "dy": rng.uniform(-0.02, 0.02, batch), "sc": rng.uniform(0.7, 1.4, batch),
            "rc": rng.uniform(-0.02, 0.02, batch)}
