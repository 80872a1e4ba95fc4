"""Parity of the per-instance-dynamics kernel (mpci_lti_assemble, through the
C ABI) with the reference's golden vectors and with the oracle.

Tolerance: relative error <= 1e-10 per array (cases.TOL), normwise."""
import numpy as np
import pytest

import _cases as cases
from oracle import qp_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    from mpc_interface_b200 import device

    device.require_gpu()
    return device


def _form(g, per_instance):
    from mpc_interface_b200 import Param

    nx = g["Q_w"].shape[0]
    refs = [Param("ref%d" % i) for i in range(nx)]
    A0 = g["A_sys"][0] if per_instance else g["A_sys"]
    B0 = g["B_sys"][0] if per_instance else g["B_sys"]
    form = cases.formulate_generic(A0, B0, g["Q_w"], g["R_w"], int(g["N"]), g["rows"], refs=refs)
    params = {"ref%d" % i: g["ref"][:, i] for i in range(nx)}
    return form, params


@pytest.mark.parametrize("hessian", ["dmma", "recursion"])
@pytest.mark.parametrize("name,per_instance", [
    ("generic_perinst_small", True), ("generic_perinst_odd", True), ("generic_perinst_mid", True),
    ("generic_small", False), ("generic_c4_shape", False)])
def test_lti_matches_reference_goldens(dev, name, per_instance, hessian):
    g = cases.golden(name)
    form, params = _form(g, per_instance)
    cases.check_layout(form, g)
    out = form.generate_all_qp_matrices_batched(
        g["x0"], params, dynamics={"plant": (g["A_sys"], g["B_sys"])}, hessian=hessian)
    got = (out.G_numpy(), out.h_numpy(), out.P_numpy(), out.q_numpy())
    cases.assert_qp_close(got, (g["A"], g["h"], g["Q"], g["q"]))


def test_lti_agrees_with_affine_path_when_dynamics_are_shared(dev):
    """Shared (A, B): mode A (plan constants + streams) and mode B must agree."""
    g = cases.golden("generic_small")
    form, params = _form(g, False)
    a = form.generate_all_qp_matrices_batched(g["x0"], params)
    b = form.generate_all_qp_matrices_batched(
        g["x0"], params, dynamics={"plant": (g["A_sys"], g["B_sys"])})
    for x, y in ((a.P, b.P), (a.q, b.q), (a.G, b.G), (a.h, b.h)):
        assert cases.rel_err(y.cpu().numpy(), x.cpu().numpy()) <= cases.TOL


@pytest.mark.parametrize("nx,nu,N,batch", [(12, 4, 50, 300), (6, 2, 100, 40), (2, 1, 5, 7), (5, 5, 13, 9), (1, 1, 1, 3), (3, 2, 33, 20),
                          (8, 6, 20, 10), (6, 2, 400, 5), (5, 3, 260, 4)])
def test_lti_random_batches_against_oracle(dev, monkeypatch, nx, nu, N, batch):
    """More instances than resident CTAs (persistent loop), odd sizes, m = 5
    (general recursion kernel), long horizons (N >= 256: 384-thread CTAs, one lane per
    warp issues the row copies); checked against the oracle on a few instances."""
    import mpc_interface_b200.tools as tools
    from _numpy_backend import NumpyBackend
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import synthetic

    seed = 100 + nx
    _, _, Q, R = synthetic.generic_system(seed, nx, nu)
    A, B = synthetic.generic_plants(seed, batch, nx, nu)
    rows = np.random.default_rng(seed + 1).normal(size=[2, nx])
    vals = synthetic.generic_inputs(seed, batch, nx)
    refs = [Param("ref%d" % i) for i in range(nx)]
    from mpc_interface_b200 import Constraint

    def with_row_arrows(f):
        # one arrow / extreme per horizon sample: these rows cannot be bulk copies
        signs = np.where(np.arange(N) % 3 == 0, -1.0, 1.0)
        f.incorporate_constraint(
            "per_row", Constraint("z0", np.linspace(1.0, 2.0, N), arrow=signs, center=np.linspace(-0.1, 0.1, N)))
        f.incorporate_constraint("per_row_u", Constraint("u0", np.linspace(0.5, 1.5, N), arrow=-signs))
        f.make_preview_matrices()
        return f

    form = with_row_arrows(cases.formulate_generic(A[0], B[0], Q, R, N, rows, refs=refs))
    params = {"ref%d" % i: vals["ref"][:, i] for i in range(nx)}
    outs = {}
    for mode in ("dmma", "recursion"):
        out = form.generate_all_qp_matrices_batched(
            vals["x0"], params, dynamics={"plant": (A, B)}, hessian=mode)
        outs[mode] = (out.G_numpy(), out.h_numpy(), out.P_numpy(), out.q_numpy())
    monkeypatch.setattr(tools, "extend_matrices", orc.extend_matrices)
    for b in sorted({0, batch // 2, batch - 1}):
        ref_form = with_row_arrows(cases.formulate_generic(
            A[b], B[b], Q, R, N, rows, refs=list(vals["ref"][b]), backend=NumpyBackend()))
        ref = orc.all_qp_matrices(ref_form, vals["x0"][b][:, None])
        for mode in outs:
            cases.assert_qp_close([x[b] for x in outs[mode]], ref)


def test_lti_partial_outputs_and_errors(dev):
    import ctypes

    import torch
    from mpc_interface_b200 import lti

    g = cases.golden("generic_perinst_small")
    form, params = _form(g, True)
    plan = form.lti_plan("plant")
    full = lti.assemble(plan, g["A_sys"], g["B_sys"], g["x0"], params)
    # only q and h requested
    dp = plan.device_plan()
    A, B, x0 = (torch.from_numpy(np.ascontiguousarray(g[k])).cuda() for k in ("A_sys", "B_sys", "x0"))
    scal = torch.from_numpy(plan.slots.evaluate(params, x0.shape[0])).cuda()
    q = torch.zeros_like(full.q)
    h = torch.zeros_like(full.h)
    null = ctypes.c_void_p(0)
    rc = dev.lib().mpci_lti_assemble(dp.handle, A.data_ptr(), 16, B.data_ptr(), 8, x0.data_ptr(),
                                     scal.data_ptr(), null, q.data_ptr(), null, h.data_ptr(), 4, 0, null)
    assert rc == 0
    torch.cuda.synchronize()
    assert torch.equal(q, full.q) and torch.equal(h, full.h)
    # invalid arguments come back as negative codes, not crashes
    assert dev.lib().mpci_lti_assemble(dp.handle, A.data_ptr(), 3, B.data_ptr(), 8, x0.data_ptr(),
                                       scal.data_ptr(), null, null, null, null, 4, 0, null) < 0
    assert dev.lib().mpci_lti_assemble(dp.handle, A.data_ptr(), 16, B.data_ptr(), 8, x0.data_ptr(),
                                       scal.data_ptr(), null, null, null, null, 4, 7, null) < 0
    with pytest.raises(ValueError):
        lti.assemble(plan, g["A_sys"], g["B_sys"], g["x0"][:, :3], params)


def test_lti_launch_info(dev):
    """The plan reports how it maps onto an SM: the recursion kernel fits more CTAs than the
    DMMA one (no zero padding, fewer registers) and uses the DMMA patch scan when its
    unchecked loads are safe (not for a 1 x 1 x 1 problem)."""
    from mpc_interface_b200 import lti

    g = cases.golden("generic_c4_shape")
    form, params, _ = cases.generic_batched(g)
    plan = form.lti_plan("plant")
    dm, rec = plan.launch_info(lti.HESSIAN_DMMA), plan.launch_info(lti.HESSIAN_RECURSION)
    # N = 50: whole rows of a block per warp (scan 2); long horizons keep the band scan (1)
    assert dm["scan"] == 0 and rec["scan"] == 2
    assert rec["smem_bytes"] < dm["smem_bytes"] <= 227 * 1024
    assert rec["ctas_per_sm"] >= dm["ctas_per_sm"] >= 1
    # skewed rows: (rows * LD) mod 16 in {4, 12} for the strided DMMA operand loads
    assert (5 * rec["ld_gz"]) % 16 in (4, 12) and (4 * rec["ld_wz"]) % 16 in (4, 12)


@pytest.mark.parametrize("per_instance", [True, False])
def test_lti_host_buffer_entry_point(dev, per_instance):
    """mpci_lti_assemble_host (chunks on two streams, ragged last chunk) returns exactly
    what the device-resident call computes, for per-instance and for shared dynamics."""
    import torch
    from mpc_interface_b200 import lti

    g = cases.golden("generic_perinst_small")
    form, params = _form(g, True)
    plan = form.lti_plan("plant")
    reps = 5  # 20 instances, chunk 6 -> chunks of 6, 6, 6, 2
    A = np.tile(g["A_sys"], [reps, 1, 1]) if per_instance else g["A_sys"][0]
    B = np.tile(g["B_sys"], [reps, 1, 1]) if per_instance else g["B_sys"][0]
    x0 = np.tile(g["x0"], [reps, 1])
    big = {k: np.tile(v, reps) for k, v in params.items()}
    for mode in (lti.HESSIAN_DMMA, lti.HESSIAN_RECURSION):
        dev_out = lti.assemble(plan, A, B, x0, big, hessian=mode)
        P, q, G, h = lti.assemble_host(plan, A, B, x0, big, hessian=mode, chunk=6)
        assert torch.equal(P, dev_out.P.cpu()) and torch.equal(q, dev_out.q.cpu())
        assert torch.equal(G, dev_out.G.cpu()) and torch.equal(h, dev_out.h.cpu())
    null = ctypes_null()
    assert dev.lib().mpci_lti_assemble_host(plan.device_plan().handle, null, 16, null, 8, null, null,
                                            null, null, null, null, 4, 0, 2) < 0


def ctypes_null():
    import ctypes

    return ctypes.c_void_p(0)


@pytest.mark.parametrize("hessian", ["dmma", "recursion"])
@pytest.mark.parametrize("name", ["generic_c4_perinst_digest", "generic_c5_perinst_digest"])
def test_lti_at_the_baseline_shapes_against_reference_digests(dev, name, hessian):
    """BASELINE configs[3] / configs[4] shapes with one plant per instance (8 plants at
    nx=12, nu=4, N=50; 4 at nx=6, nu=2, N=400), both Hessian modes, against digests of the
    matrices the unmodified reference produced (make_goldens.py big)."""
    g = cases.golden(name)
    form, params = _form(g, True)
    cases.check_layout(form, g)
    out = form.generate_all_qp_matrices_batched(
        g["x0"], params, dynamics={"plant": (g["A_sys"], g["B_sys"])}, hessian=hessian)
    got = [x.cpu().numpy() for x in (out.G, out.h, out.P, out.q)]
    for b in range(g["x0"].shape[0]):
        cases.assert_digest_close(tuple(x[b] for x in got), g, "i%d_" % b, int(g["seed"]) + b)


@pytest.mark.parametrize("hessian", ["dmma", "recursion"])
@pytest.mark.parametrize("name", ["lip_perinst", "lip_perinst_n9"])
def test_lip_named_system_per_instance_omega_and_tau(dev, name, hessian):
    """The LIP with CoP-rate input on two axes, one pendulum frequency AND one sampling period
    per instance: the plants come from the device-side closed forms of "dP->CCC"
    (mpci_named_system), the definitions b / DCM carry Param coefficients (per-instance
    tables, mpci_lti_assemble_ex), the terminal DCM box is a scheduled limit.  Golden: the
    reference formulated every instance with ControlSystem.from_name("dP->CCC", tau_b,
    omega_b) (make_goldens.py lip).  ONE batched call."""
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import lip

    g = cases.golden(name)
    conf = lip.default_configuration(int(g["N"]))
    form = lip.formulate_lip(conf, omega=Param("omega"), tau=Param("tau"),
                             target_vel=[Param("vel_x"), Param("vel_y")])
    cases.check_layout(form, g)
    params = {k: g[k] for k in ("omega", "vel_x", "vel_y")}
    given = np.hstack([g["x0_x"], g["x0_y"]])
    out = form.generate_all_qp_matrices_batched(
        given, params, dynamics={"LIP": ("dP->CCC", g["tau"], g["omega"])}, hessian=hessian)
    got = (out.G_numpy(), out.h_numpy(), out.P_numpy(), out.q_numpy())
    cases.assert_qp_close(got, (g["A"], g["h"], g["Q"], g["q"]))
    N = int(g["N"])
    assert np.all(got[2][:, :N, N:] == 0.0)  # the axes do not couple: exact zeros, as in the reference


def test_named_systems_on_device_match_the_sympy_lambdas(dev):
    """mpci_named_system against tools.get_system_matrices (SymPy exp / integrate,
    tools.py:240-285) for every named system and a batch of (tau, omega)."""
    from mpc_interface_b200 import lti
    from mpc_interface_b200 import tools as use

    rng = np.random.default_rng(3)
    tau, omega = rng.uniform(0.05, 0.2, 9), rng.uniform(2.5, 4.0, 9)
    for name in ("J->CCC", "P->CC", "P->X", "dP->CCC", "dP->CCP"):
        A, B = lti.named_system_batched(name, tau, omega)
        get_A, get_B, _ = use.get_system_matrices(name)
        for b in range(9):
            refA = np.array(get_A(tau=tau[b], omega=omega[b]), dtype=float)
            refB = np.array(get_B(tau=tau[b], omega=omega[b]), dtype=float)
            assert cases.rel_err(A[b].cpu().numpy(), refA) <= 1e-13, name
            assert cases.rel_err(B[b].cpu().numpy(), refB.reshape(B[b].shape)) <= 1e-13, name
    A1, _ = lti.named_system_batched("J->CCC", 0.1, None, batch=4)  # shared tau, no omega
    assert tuple(A1.shape) == (4, 3, 3) and float(A1[3, 0, 1]) == 0.1


@pytest.mark.parametrize("hessian", ["dmma", "recursion"])
def test_permuted_inputs_scheduled_limit_and_param_weight(dev, monkeypatch, hessian):
    """Optimisation variables in another order than the plant inputs, a limit on a window of
    the horizon, a per-instance cost weight and a per-instance bound: against the oracle."""
    import mpc_interface_b200.tools as tools
    from _numpy_backend import NumpyBackend
    from mpc_interface_b200 import Constraint, Cost, Param

    g = cases.golden("generic_small")

    def build(backend, refs, weight, extreme):
        form = cases.formulate_generic(g["A_sys"], g["B_sys"], g["Q_w"], g["R_w"], int(g["N"]), g["rows"],
                                       refs=refs, backend=backend)
        form.incorporate_constraint("late", Constraint("z2", extreme, schedule=range(3, 7)))
        form.incorporate_goal("extra", Cost("z1", weight, aim=[0.25]))
        form.identify_qp_domain(["u1", "u0"])
        form.make_preview_matrices()
        return form

    nx, batch = 4, g["x0"].shape[0]
    form = build(None, [Param("ref%d" % i) for i in range(nx)], Param("w"), Param("lim"))
    params = {"ref%d" % i: g["ref"][:, i] for i in range(nx)}
    params["w"] = np.array([0.5, 2.0, 3.5])[:batch]
    params["lim"] = np.array([1.0, 1.5, 0.2])[:batch]
    out = form.generate_all_qp_matrices_batched(
        g["x0"], params, dynamics={"plant": (g["A_sys"], g["B_sys"])}, hessian=hessian)
    got = (out.G_numpy(), out.h_numpy(), out.P_numpy(), out.q_numpy())
    monkeypatch.setattr(tools, "extend_matrices", orc.extend_matrices)
    for b in range(batch):
        ref_form = build(NumpyBackend(), list(g["ref"][b]), float(params["w"][b]), float(params["lim"][b]))
        ref = orc.all_qp_matrices(ref_form, g["x0"][b][:, None])
        cases.assert_qp_close(tuple(x[b] for x in got), ref)
    with pytest.raises(ValueError):  # a negative per-instance extreme would be flipped by the reference
        form.generate_all_qp_matrices_batched(
            g["x0"], dict(params, lim=-params["lim"]), dynamics={"plant": (g["A_sys"], g["B_sys"])})


@pytest.mark.parametrize("N,batch", [(12, 7), (40, 5)])
def test_lip_per_instance_plants_and_per_instance_yaw(dev, N, batch):
    """SURVEY 8(f) row 4 on the per-instance path: the foot polygon and the terminal DCM box turned
    by one yaw per robot (`rotate_in_TS(rotation2D(Param("yaw")))`, restrictions.py:380-401) while
    every robot also has its own plant (tau_b, omega_b).  The turned arrow joins the per-instance
    coefficients of the limit's form.  Every instance formulated, rotated and assembled by the
    unmodified reference."""
    import types

    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import lip
    from oracle import reference

    if not reference.available():
        pytest.skip("oracle/_ref is not staged")
    ref = reference.load()
    api = types.SimpleNamespace(ControlSystem=ref.ControlSystem, ExtendedSystem=ref.ExtendedSystem,
                                Formulation=ref.Formulation, LineCombo=ref.LineCombo, Cost=ref.Cost, Box=ref.Box,
                                make_simetric_vertices=ref.tools.make_simetric_vertices,
                                rotation2D=ref.tools.rotation2D)
    conf = lip.default_configuration(N)
    form = lip.formulate_lip(conf, omega=Param("omega"), tau=Param("tau"),
                             target_vel=[Param("vel_x"), Param("vel_y")], yaw=Param("yaw"))
    vals = lip.lip_inputs(40 + N, batch)
    vals["tau"] = vals["tau"] * min(1.0, 30.0 / N)
    vals["yaw"] = np.random.default_rng(N).uniform(-np.pi, np.pi, batch)
    given = np.hstack([vals["x0_x"], vals["x0_y"]])
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y", "yaw")}
    wants = []
    for b in range(batch):
        one = lip.formulate_lip(conf, omega=float(vals["omega"][b]), tau=float(vals["tau"][b]),
                                target_vel=[float(vals["vel_x"][b]), float(vals["vel_y"][b])],
                                yaw=float(vals["yaw"][b]), api=api)
        wants.append(one.generate_all_qp_matrices(one.arrange_given(
            {v: np.asarray(vals[v][b]).reshape([-1, 1]) for v in one.given_ID})))
    for hessian in ("dmma", "recursion"):
        out = form.generate_all_qp_matrices_batched(
            given, params, dynamics={"LIP": ("dP->CCC", vals["tau"], vals["omega"])}, hessian=hessian)
        got = (out.G_numpy(), out.h_numpy(), out.P_numpy(), out.q_numpy())
        for b in range(batch):
            cases.assert_qp_close(tuple(x[b] for x in got), wants[b])
