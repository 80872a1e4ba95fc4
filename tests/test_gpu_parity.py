"""Parity of the CUDA path (through the C ABI of include/mpci.h) with the
oracle and with the golden vectors written by the unmodified reference.

Tolerance: north_star asks for relative error <= 1e-10 on P, q, G, h; the
metric is normwise per array, max|x - ref| / max|ref| (cases.TOL).  Integer
layouts (optim_ID / given_ID, row offsets) must be equal bit for bit."""
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


# ---------------------------------------------------------------- K1
def test_lip_pickle_golden(dev):
    """The reference's only value-level pin of K1 (test_dynamics.py:139-152)."""
    from mpc_interface_b200 import ControlSystem, ExtendedSystem

    g = cases.golden("lip_matrices")
    LIP = ControlSystem.from_name("J->CCC", tau=0.1, omega=3.3445)
    ext = ExtendedSystem.from_cotrol_system(LIP, "x", 36)
    assert np.isclose(ext.matrices[0], g["U"]).all()
    assert np.isclose(ext.matrices[1], g["S"]).all()
    assert cases.rel_err(ext.matrices[0], g["U"]) < 1e-13


@pytest.mark.parametrize("tag", list("abcde"))
def test_extend_matrices_cases(dev, tag):
    from mpc_interface_b200 import tools

    g = cases.golden("extend_cases")
    A, B = g[tag + "_A"], g[tag + "_B"]
    N = g[tag + "_S"].shape[0]
    S, U = tools.extend_matrices(N, A, B)
    assert S.shape == (N,) + A.shape and len(U) == B.shape[1] and U[0].shape == (N, N, A.shape[0])
    assert cases.rel_err(S, g[tag + "_S"]) < 1e-12
    assert cases.rel_err(np.stack(U), g[tag + "_U"]) < 1e-12
    assert (np.stack(U)[:, np.triu_indices(N, 1)[0], np.triu_indices(N, 1)[1]] == 0).all()


def test_extend_matrices_batched_device(dev):
    import torch

    rng = np.random.default_rng(0)
    batch, n, m, N = 37, 6, 2, 40
    A = np.eye(n) + 0.05 * rng.normal(size=[batch, n, n])
    B = rng.normal(size=[batch, n, m])
    S, U = dev.extend_matrices_device(N, torch.from_numpy(A).cuda(), torch.from_numpy(B).cuda())
    Phi, Gam = dev.extend_matrices_device(
        N, torch.from_numpy(A).cuda(), torch.from_numpy(B).cuda(), dense=False
    )
    S, U, Phi, Gam = (x.cpu().numpy() for x in (S, U, Phi, Gam))
    for b in (0, 17, 36):
        So, Uo = orc.extend_matrices(N, A[b], B[b])
        assert cases.rel_err(S[b], So) < 1e-12 and cases.rel_err(U[b], np.stack(Uo)) < 1e-12
        assert cases.rel_err(Phi[b], So.transpose(0, 2, 1)) < 1e-12
        assert cases.rel_err(Gam[b][:, :, 0], Uo[0][:, 0, :]) < 1e-12


def test_gecko_cpp_extend_matrices(dev):
    """The reference's C++ test cases (cpp/tests/tools.cpp) through the C++ header
    include/mpci_gecko.hh: U(j,k,p,i) / S(k,c,i) index order, exact zeros above the
    diagonal, values within 1e-10 of a plain-loop recomputation."""
    import _gecko_build

    done = _gecko_build.run()
    assert done.returncode == 0, done.stdout + done.stderr
    assert "gecko::tools::extend_matrices ok" in done.stdout
    # dynamics.hh's ControlSystem / ExtendedSystem incl. the from_control_system the
    # reference left as a TODO, name -> index maps as in dynamics.py:248-267
    assert "ControlSystem / ExtendedSystem ok" in done.stdout


def test_extend_matrices_errors(dev):
    from mpc_interface_b200 import tools

    with pytest.raises(ValueError):
        tools.extend_matrices(5, np.eye(3), np.ones([2, 1]))
    with pytest.raises(ValueError):
        tools.extend_matrices(0, np.eye(3), np.ones([3, 1]))


# ---------------------------------------------------------------- plan-time kernels
@pytest.mark.parametrize("shape", [(1, 1, 1), (7, 5, 3), (64, 64, 16), (100, 218, 100), (218, 426, 100), (130, 70, 33)])
@pytest.mark.parametrize("transA", [False, True])
def test_dmma_gemm(dev, shape, transA):
    ops = dev.DeviceBackend()
    M, N, K = shape
    rng = np.random.default_rng(M * 1000 + N)
    A = rng.normal(size=[K, M] if transA else [M, K])
    B = rng.normal(size=[K, N])
    C = ops.download(ops.gemm(ops.upload(A), ops.upload(B), transA=transA))
    ref = (A.T if transA else A) @ B
    assert cases.rel_err(C, ref) < 1e-13
    # strided views (what the plan compiler passes)
    big = ops.upload(rng.normal(size=[K + 3, N + 9]))
    Cv = ops.download(ops.gemm(ops.upload(A), ops.sub(big, 2, K + 2, 4, N + 4), transA=transA))
    assert cases.rel_err(Cv, (A.T if transA else A) @ ops.download(big)[2 : K + 2, 4 : N + 4]) < 1e-13


def test_axpy2d(dev):
    ops = dev.DeviceBackend()
    rng = np.random.default_rng(4)
    dst0, src = rng.normal(size=[9, 13]), rng.normal(size=[9, 13])
    scale = rng.normal(size=9)
    d = ops.upload(dst0)
    ops.axpy(d, ops.upload(src), 0.5, scale)
    assert np.allclose(ops.download(d), dst0 + 0.5 * scale[:, None] * src, rtol=1e-15, atol=0)
    d = ops.upload(dst0)
    ops.axpy(d, ops.upload(src[:1]), -2.0, None)
    assert np.allclose(ops.download(d), dst0 - 2.0 * src[:1], rtol=1e-15, atol=0)
    with pytest.raises(ValueError):
        ops.axpy(d, ops.upload(src[:4]), 1.0, None)


# ---------------------------------------------------------------- the assembly path
def test_biped_c1_single_instance(dev):
    """Config 1: the repo example at its default horizon, one instance."""
    g = cases.golden("biped_c1")
    form = cases.formulate_biped(cases.default_configuration())
    cases.check_layout(form, g)
    A, h, Q, q = form.generate_all_qp_matrices(g["given"])
    assert A.shape == (108, 52) and h.shape == (108, 1) and Q.shape == (52, 52) and q.shape == (52, 1)
    cases.assert_qp_close((A, h, Q, q), (g["A"], g["h"], g["Q"], g["q"]))
    Mg, Mo = form.PM["(b+n-s)_x"]
    assert cases.rel_err(np.hstack([Mg, Mo]), g["PM_bns_x"]) < 1e-13


@pytest.mark.parametrize("name", ["biped_n24_batch", "biped_n24_terminal", "biped_n100_batch"])
def test_biped_batched_goldens(dev, name):
    g = cases.golden(name)
    form, params = cases.biped_batched(g)
    cases.check_layout(form, g)
    out = form.generate_all_qp_matrices_batched(g["given"], params)
    assert out.P.is_cuda and out.P.shape == g["Q"].shape and out.G.shape == g["A"].shape
    cases.assert_qp_close(
        (out.G_numpy(), out.h_numpy(), out.P_numpy(), out.q_numpy()),
        (g["A"], g["h"], g["Q"], g["q"]),
    )


@pytest.mark.parametrize("N,batch", [(24, 67), (100, 9)])
def test_biped_vs_oracle_random(dev, N, batch):
    """Seeded inputs of SURVEY 8(d), ragged batch sizes (not multiples of the
    32 / 16 instance groups of the kernels)."""
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import synthetic

    conf = cases.default_configuration(N)
    form = cases.formulate_biped(conf, omega=Param("omega"), target_vel=[Param("vel_x"), Param("vel_y")])
    vals = synthetic.biped_inputs(77 + N, batch, N)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    out = form.generate_all_qp_matrices_batched(given, params)
    P, q, G, h = out.P_numpy(), out.q_numpy(), out.G_numpy(), out.h_numpy()
    for b in sorted({0, 1, batch // 2, batch - 2, batch - 1}):
        one = {k: float(v[b]) for k, v in params.items()}
        ref = orc.all_qp_matrices(form, given[b][:, None], one)
        cases.assert_qp_close((G[b], h[b], P[b], q[b]), ref)


@pytest.mark.parametrize("name", ["biped_n24_yaw", "biped_n100_yaw"])
def test_biped_per_instance_yaw_goldens(dev, name):
    """Box.rotate_in_TS with a per-instance angle; the golden is the
    reference's rotate_in_TS run instance by instance (make_goldens.py yaw)."""
    g = cases.golden(name)
    form, params = cases.biped_yaw(g)
    cases.check_layout(form, g)
    out = form.generate_all_qp_matrices_batched(g["given"], params)
    cases.assert_qp_close(
        (out.G_numpy(), out.h_numpy(), out.P_numpy(), out.q_numpy()),
        (g["A"], g["h"], g["Q"], g["q"]),
    )


@pytest.mark.parametrize("batch", [1, 63, 700, 2600, 10000])
def test_qh_block_splits_over_batch_sizes(dev, batch):
    """The [q;h] kernel spreads a small batch over up to 8 blocks per group of 64
    instances (launch_qh: 8 at 1..576 instances, 4 / 2 up to 4736, 1 above): every
    setting against the oracle, first / middle / last instance."""
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import synthetic

    N = 24
    form = cases.formulate_biped(cases.default_configuration(N), omega=Param("omega"),
                                 target_vel=[Param("vel_x"), Param("vel_y")])
    vals = synthetic.biped_inputs(77, batch, N)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    out = form.generate_all_qp_matrices_batched(given, params)
    q, h = out.q_numpy(), out.h_numpy()
    assert np.isfinite(q).all() and np.isfinite(h).all()
    for b in sorted({0, batch // 2, batch - 1}):
        one = {k: float(v[b]) for k, v in params.items()}
        ref = orc.all_qp_matrices(form, given[b][:, None], one)
        assert cases.rel_err(q[b], np.squeeze(ref[3])) <= cases.TOL
        assert cases.rel_err(h[b], np.squeeze(ref[1])) <= cases.TOL


def test_biped_per_instance_geometry_vs_oracle(dev):
    """Rotation + per-instance centre and scale of the support polygon, batch 45
    at N=100 (7 streamed G terms), against the oracle."""
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import synthetic

    N, batch = 100, 45
    form = cases.formulate_biped(
        cases.default_configuration(N), omega=Param("omega"),
        target_vel=[Param("vel_x"), Param("vel_y")], yaw=Param("yaw"), preview=False)
    box = form.constraint_boxes["support_polygon"]
    box.translate_in_TS(np.array([Param("cx"), Param("cy")]))
    box.scale_box(Param("scale"))
    form.make_preview_matrices()
    vals = synthetic.biped_inputs(321, batch, N)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    rng = np.random.default_rng(4)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    params.update(yaw=rng.uniform(-3, 3, batch), cx=rng.normal(size=batch) * 0.01,
                  cy=rng.normal(size=batch) * 0.01, scale=rng.uniform(0.8, 1.2, batch))
    out = form.generate_all_qp_matrices_batched(given, params)
    P, q, G, h = out.P_numpy(), out.q_numpy(), out.G_numpy(), out.h_numpy()
    assert np.abs(G[0] - G[1]).max() > 1e-3  # G really differs between instances
    for b in (0, 1, 22, 44):
        one = {k: float(v[b]) for k, v in params.items()}
        ref = orc.all_qp_matrices(form, given[b][:, None], one)
        cases.assert_qp_close((G[b], h[b], P[b], q[b]), ref)


@pytest.mark.parametrize("optim,with_limit", [(("x0", "u"), True), (("u",), False), (("x0", "u"), False)])
@pytest.mark.parametrize("batch", [0, 1, 33])
def test_edge_shapes(dev, optim, with_limit, batch):
    """Empty given vector (x0 optimised), no constraint rows, odd nv (scalar stores),
    empty / single / ragged batches -- against the oracle."""
    form = cases.tiny_plant(optim=optim, with_limit=with_limit)
    ng, nv = form.given_len, form.optim_len
    assert ng == (0 if "x0" in optim else 3) and nv == (8 if "x0" in optim else 5)
    rng = np.random.default_rng(batch)
    given = rng.normal(size=[batch, ng])
    params = {"w": rng.uniform(0.5, 2.0, batch), "aim": rng.normal(size=batch)}
    out = form.generate_all_qp_matrices_batched(given, params)
    nc = 5 if with_limit else 0
    assert out.P.shape == (batch, nv, nv) and out.q.shape == (batch, nv)
    assert out.G.shape == (batch, nc, nv) and out.h.shape == (batch, nc)
    P, q, G, h = out.P_numpy(), out.q_numpy(), out.G_numpy(), out.h_numpy()
    for b in range(min(batch, 3)):
        one = {k: float(v[b]) for k, v in params.items()}
        PM = orc.preview_matrices(form, one)
        g = given[b][:, None]
        costs = [orc.qp_cost(c, PM, g, one) for c in form.goals.values()]
        assert cases.rel_err(P[b], np.add.reduce([c[0] for c in costs])) <= cases.TOL
        assert cases.rel_err(q[b], np.add.reduce([c[1] for c in costs])[:, 0]) <= cases.TOL
        if with_limit:
            A_ref, h_ref = orc.qp_constraint(form.constraints["lim"][0], PM, g, one)
            assert cases.rel_err(G[b], A_ref) <= cases.TOL and cases.rel_err(h[b], h_ref[:, 0]) <= cases.TOL
    if not with_limit:  # the single-instance API fails like the reference (np.vstack of nothing)
        with pytest.raises(ValueError):
            form.generate_all_qp_matrices(np.zeros([ng, 1]))


@pytest.mark.parametrize("N,batch", [(24, 37), (100, 5)])
def test_more_than_eight_streamed_terms(dev, N, batch):
    """10 G monomials: stream_terms_generic (constants through L1 / L2 instead of
    registers) against the oracle."""
    form = cases.biped_three_yaws(N)
    given, params = cases.three_yaw_inputs(form, 40 + N, batch, N)
    assert len(form._compiled().G_terms) > 8
    out = form.generate_all_qp_matrices_batched(given, params)
    P, q, G, h = out.P_numpy(), out.q_numpy(), out.G_numpy(), out.h_numpy()
    for b in sorted({0, batch // 2, batch - 1}):
        one = {k: float(v[b]) for k, v in params.items()}
        ref = orc.all_qp_matrices(form, given[b][:, None], one)
        cases.assert_qp_close((G[b], h[b], P[b], q[b]), ref)


@pytest.mark.parametrize("seed", range(12))
def test_random_formulations_against_reference(dev, seed):
    """tests/_random_forms.py through the CUDA path against what the reference
    produced for the same recipe (golden random_forms, make_goldens.py random):
    13 P monomials (generic stream kernel), nine per-instance scalars, odd sizes.
    The CSC hand-off of the same batch is checked against the dense one."""
    g = cases.golden("random_forms")
    form, given, params, ref = cases.random_form(g, seed)
    out = form.generate_all_qp_matrices_batched(given, params)
    got = (out.G_numpy(), out.h_numpy(), out.P_numpy(), out.q_numpy())
    cases.assert_qp_close(got, ref)
    csc = form.generate_all_qp_matrices_batched(given, params, layout="csc")
    for b in range(given.shape[0]):
        Pb, Gb = csc.csc(b)
        assert np.array_equal(Pb.toarray(), got[2][b]) and np.array_equal(Gb.toarray(), got[0][b])


def test_negative_per_instance_extreme_is_refused(dev):
    from mpc_interface_b200 import Constraint, Param

    form = cases.tiny_plant(optim=("u",))
    form.incorporate_constraint("cap", Constraint("a", Param("e")))
    form.make_preview_matrices()
    given = np.zeros([2, form.given_len])
    values = {"w": np.ones(2), "aim": np.zeros(2)}
    with pytest.raises(ValueError, match="negative for instance 1"):
        form.generate_all_qp_matrices_batched(given, dict(values, e=np.array([0.5, -0.1])))


def test_reference_test_body_case(dev):
    g = cases.golden("test_body")
    body = cases.body_test_formulation()
    cases.check_layout(body, g)
    cases.assert_qp_close(body.generate_all_qp_matrices(g["given"]), (g["A"], g["h"], g["Q"], g["q"]))
    cQ, cq = body.generate_qp_cost(body.goals["crossed"], g["given"])
    assert cases.rel_err(cQ, g["cQ"]) < cases.TOL and cases.rel_err(cq, g["cq"]) < cases.TOL
    kin = body.constraints["kinematics"][0]
    A0, h0 = body.generate_qp_constraint(kin, g["given"])
    assert (h0 == kin.bound()).all()  # given does not affect this limit (test_body.py:141-142)
    A2, h2 = body.generate_all_qp_constraints(g["given"])
    assert A2.shape == (54, body.optim_len) and h2.shape == (54, 1)


@pytest.mark.parametrize("name", ["generic_small", "generic_c4_shape"])
def test_generic_goldens(dev, name):
    g = cases.golden(name)
    form, params, given = cases.generic_batched(g)
    out = form.generate_all_qp_matrices_batched(given, params)
    cases.assert_qp_close(
        (out.G_numpy(), out.h_numpy(), out.P_numpy(), out.q_numpy()),
        (g["A"], g["h"], g["Q"], g["q"]),
    )


def test_biped_ticks_ragged_sizes(dev):
    g = cases.golden("biped_ticks")
    form = cases.formulate_biped(cases.default_configuration(24))
    for tick in (0, 11, 12, 13):
        form.update(step_times=g["t%d_step_times" % tick], step_count=int(g["t%d_step_count" % tick]))
        got = form.generate_all_qp_matrices(g["t%d_given" % tick])
        cases.assert_qp_close(got, tuple(g["t%d_%s" % (tick, k)] for k in "AhQq"))


def test_host_buffer_entry_point_and_reuse(dev):
    """mpci_assemble_host (chunked, two streams) equals the device-resident call;
    an output batch can be reused as a ring slot."""
    import torch
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import synthetic

    N, batch = 24, 150
    form = cases.formulate_biped(
        cases.default_configuration(N), omega=Param("omega"), target_vel=[Param("vel_x"), Param("vel_y")]
    )
    vals = synthetic.biped_inputs(5, batch, N)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    out = form.generate_all_qp_matrices_batched(given, params)
    again = form.generate_all_qp_matrices_batched(given, params, out=out)
    assert again is out
    plan = form._compiled()
    scalars = torch.from_numpy(plan.slots.evaluate(params, batch)).pin_memory()
    hp = torch.empty([batch, plan.nv, plan.nv], dtype=torch.float64).pin_memory()
    hq = torch.empty([batch, plan.nv], dtype=torch.float64).pin_memory()
    hG = torch.empty([batch, plan.nc, plan.nv], dtype=torch.float64).pin_memory()
    hh = torch.empty([batch, plan.nc], dtype=torch.float64).pin_memory()
    form._ops().assemble_host(plan, scalars, torch.from_numpy(given).pin_memory(), hp, hq, hG, hh, batch, 64)
    assert torch.equal(hp, out.P.cpu()) and torch.equal(hG, out.G.cpu())
    assert torch.equal(hq, out.q.cpu()) and torch.equal(hh, out.h.cpu())


def test_full_width_properties(dev):
    """Size-independent checks on a larger batch at N=100 (config 2 shape):
    P symmetric, [q; h] affine in given, a checksum of checksums of G."""
    import torch
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import synthetic

    N, batch = 100, 2048
    form = cases.formulate_biped(
        cases.default_configuration(N), omega=Param("omega"), target_vel=[Param("vel_x"), Param("vel_y")]
    )
    vals = synthetic.biped_inputs(1002, batch, N)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    out = form.generate_all_qp_matrices_batched(given, params)
    assert out.P.shape == (batch, 218, 218) and out.G.shape == (batch, 440, 218)
    sym = (out.P - out.P.transpose(1, 2)).abs().amax().item()
    assert sym <= 1e-12 * out.P.abs().amax().item()
    # affine in given: f(g1) + f(g2) - f(0) == f(g1 + g2) with the same parameters
    half = batch // 2
    p2 = {k: np.concatenate([v[:half], v[:half]]) for k, v in params.items()}
    g_sum = np.concatenate([given[:half] + given[half:], np.zeros_like(given[:half])])
    o2 = form.generate_all_qp_matrices_batched(g_sum, p2)
    p1 = {k: np.concatenate([v[:half], v[:half]]) for k, v in params.items()}
    o1 = form.generate_all_qp_matrices_batched(given, p1)
    lhs = o1.q[:half] + o1.q[half:] - o2.q[half:]
    assert (lhs - o2.q[:half]).abs().amax().item() <= 1e-11 * o2.q.abs().amax().item()
    lhs = o1.h[:half] + o1.h[half:] - o2.h[half:]
    assert (lhs - o2.h[:half]).abs().amax().item() <= 1e-11 * o2.h.abs().amax().item()
    # checksum of checksums: sum_b sum(G_b) == sum_t (sum_b c_t(b)) * sum(DG_t)
    plan = form._compiled()
    scal = plan.slots.evaluate(params, batch)
    table = plan.mono_table()
    coef = np.ones([batch, table.shape[0]])
    for j in range(table.shape[0]):
        for f in table[j]:
            if f >= 0:
                coef[:, j] *= scal[:, f]
    sums = plan.DG.sum(dim=(1, 2)).cpu().numpy()
    expect = float((coef[:, plan.G_terms].sum(axis=0) * sums).sum())
    got = float(out.G.sum().item())
    assert abs(got - expect) <= 1e-9 * max(1.0, abs(expect))


@pytest.mark.parametrize("name", ["biped_n24_batch", "biped_n100_batch"])
def test_csc_hand_off_layout(dev, name):
    """layout="csc": values in scipy's CSC order with one pattern per batch; the
    integer arrays equal scipy.sparse.csc_matrix of the reference's matrices bit
    for bit, the values equal the dense path exactly."""
    import scipy.sparse as sp
    import torch

    g = cases.golden(name)
    form, params = cases.biped_batched(g)
    dense = form.generate_all_qp_matrices_batched(g["given"], params)
    out = form.generate_all_qp_matrices_batched(g["given"], params, layout="csc")
    assert torch.equal(out.q, dense.q) and torch.equal(out.h, dense.h)
    for b in range(g["Q"].shape[0]):
        P, G = out.csc(b)
        for got, ref_dense, mine in ((P, g["Q"][b], dense.P[b]), (G, g["A"][b], dense.G[b])):
            ref = sp.csc_matrix(ref_dense)
            assert np.array_equal(got.indices, ref.indices) and np.array_equal(got.indptr, ref.indptr)
            assert got.indices.dtype == np.int32 and got.indptr.dtype == np.int32
            assert cases.rel_err(got.data, ref.data) <= cases.TOL
            assert np.array_equal(got.toarray(), mine.cpu().numpy())
    up = form.generate_all_qp_matrices_batched(g["given"], params, layout="csc_upper")
    assert np.array_equal(up.csc(0)[0].toarray(), np.triu(dense.P[0].cpu().numpy()))
    # host-buffer entry point (what the e2e bench times)
    plan = form._compiled()
    ops = form._ops()
    pP, pG = ops.compact(plan)
    batch = g["given"].shape[0]
    scalars = torch.from_numpy(plan.slots.evaluate(params, batch)).pin_memory()
    hP = torch.empty([batch, pP.padded], dtype=torch.float64).pin_memory()
    hG = torch.empty([batch, pG.padded], dtype=torch.float64).pin_memory()
    hq = torch.empty([batch, plan.nv], dtype=torch.float64).pin_memory()
    hh = torch.empty([batch, plan.nc], dtype=torch.float64).pin_memory()
    ops.assemble_csc_host(plan, scalars, torch.from_numpy(g["given"]).pin_memory(), hP, hq, hG, hh, batch, 2)
    assert torch.equal(hP[:, : pP.nnz], out.P_data.cpu()) and torch.equal(hG[:, : pG.nnz], out.G_data.cpu())
    assert torch.equal(hq, out.q.cpu()) and torch.equal(hh, out.h.cpu())


@pytest.mark.parametrize("N", [24, 100])
def test_preview_and_goal_distance_batched(dev, N):
    """Formulation.preview / goal_distance (body.py:209-234)
    for a batch with per-instance omega, against the oracle's preview matrices."""
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import synthetic

    batch = 70
    form = cases.formulate_biped(
        cases.default_configuration(N), omega=Param("omega"), target_vel=[Param("vel_x"), Param("vel_y")])
    vals = synthetic.biped_inputs(9, batch, N)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    optim = np.random.default_rng(2).normal(size=[batch, form.optim_len])
    dcm = form.preview_batched(given, optim, "DCM", ["_x", "_y"], params)
    com = form.preview_batched(given, optim, "CoM_x", None, params)
    assert tuple(dcm.shape) == (batch, N, 2) and tuple(com.shape) == (batch, N)
    dist = form.goal_distance_batched(given, optim, "track vel_x", params)
    total = form.full_goal_distance_batched(given, optim, params)
    dcm, com, dist, total = (x.cpu().numpy() for x in (dcm, com, dist, total))
    for b in (0, 31, 32, 63, 64, batch - 1):
        one = {k: float(v[b]) for k, v in params.items()}
        PM = orc.preview_matrices(form, one)
        g, z = given[b][:, None], optim[b][:, None]
        for a, axis in enumerate(("_x", "_y")):
            ref = PM["DCM" + axis][0] @ g + PM["DCM" + axis][1] @ z
            assert cases.rel_err(dcm[b, :, a], ref[:, 0]) <= cases.TOL
        assert cases.rel_err(com[b], (PM["CoM_x"][0] @ g + PM["CoM_x"][1] @ z)[:, 0]) <= cases.TOL
        ref_total = 0.0
        for name, goal in form.goals.items():
            d = 0.0
            for i, axis in enumerate(goal.axes):
                Mg, Mo = PM[goal.variable + axis]
                aim = goal.aim[0, i]
                aim = aim.evaluate(one) if hasattr(aim, "evaluate") else float(aim)
                gap = Mg @ g + Mo @ z - aim
                d += float((gap.T @ gap)[0, 0])
            ref_total += d
            if name == "track vel_x":
                assert abs(dist[b] - d) <= cases.TOL * max(1.0, abs(d))
        assert abs(total[b] - ref_total) <= cases.TOL * max(1.0, abs(ref_total))
    # one instance through the reference-named single-instance calls
    single = cases.formulate_biped(cases.default_configuration(N))
    g, z = given[0][:, None], optim[0][:, None]
    v = single.preview(g, z, "CoM", ["_x", "_y"])
    assert v.shape == (N, 2)
    PM = orc.preview_matrices(single)
    want = np.hstack([PM["CoM" + a][0] @ g + PM["CoM" + a][1] @ z for a in ("_x", "_y")])
    assert cases.rel_err(v, want) <= cases.TOL
    col = single.preview(g, z, "DCM_y")
    assert col.shape == (N, 1) and cases.rel_err(col, PM["DCM_y"][0] @ g + PM["DCM_y"][1] @ z) <= cases.TOL
    gap = PM["CoM_dot_x"][0] @ g + PM["CoM_dot_x"][1] @ z - single.goals["track vel_x"].aim[:, 0]
    d = single.goal_distance(g, z, "track vel_x")
    assert abs(d - float(gap.T @ gap)) <= cases.TOL * max(1.0, float(gap.T @ gap))
    assert single.full_goal_distance(g, z) >= d


@pytest.mark.parametrize("N", [200, 400])
def test_biped_long_horizons_in_mode_a(dev, N):
    """given_len = 8 + 2N = 408 / 808: [given; 1] no longer fits the shared memory of the
    [q; h] kernel in one piece (round 1 refused it above ~340); the columns are staged in
    passes.  Against digests written by the unmodified reference (make_goldens.py big),
    and the same instance inside a batch of 70 (more than one group of 64)."""
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import synthetic

    g = cases.golden("biped_long_digest")
    seed = int(g["seed"]) + N
    form = cases.formulate_biped(cases.default_configuration(N), omega=Param("omega"),
                                 target_vel=[Param("vel_x"), Param("vel_y")])
    cases.check_layout(form, {k: g["n%d_%s" % (N, k)] for k in
                              ("optim_names", "optim_ranges", "given_names", "given_ranges")})
    one = synthetic.biped_inputs(seed, 1, N)
    batch = 70
    vals = synthetic.biped_inputs(seed + 1, batch, N)
    for k in vals:  # the golden instance sits at positions 3 and 66
        vals[k][3] = one[k][0]
        vals[k][66] = one[k][0]
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    assert form.given_len == 8 + 2 * N
    out = form.generate_all_qp_matrices_batched(given, {k: vals[k] for k in ("omega", "vel_x", "vel_y")})
    for b in (3, 66):
        got = tuple(x[b].cpu().numpy() for x in (out.G, out.h, out.P, out.q))
        cases.assert_digest_close(got, g, "n%d_" % N, seed)
    other = tuple(x[5].cpu().numpy() for x in (out.G, out.h, out.P, out.q))
    params5 = {k: float(vals[k][5]) for k in ("omega", "vel_x", "vel_y")}
    cases.assert_qp_close(other, orc.all_qp_matrices(form, given[5][:, None], params5))


def test_integration_snippet_runs(dev):
    """the ctypes call sequence a maintainer would write (a
    hand-made mpci_plan_desc, mpci_plan_create, mpci_assemble_host) is executed as it is
    printed there, asserts included."""
    import os
    import re

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(root, "INTEGRATION.md")).read()
    block = re.search(r"<!-- snippet:plan -->\n```python\n(.*?)\n```", text, flags=re.S).group(1)
    here = os.getcwd()
    os.chdir(root)  # the snippet loads the library by its path inside the repo
    try:
        exec(compile(block, "INTEGRATION.md:snippet:plan", "exec"), {"__name__": "snippet"})
    finally:
        os.chdir(here)


def test_large_and_small_batches_give_the_same_q_and_h(dev):
    """[q; h] of a large batch (one persistent block per SM, every warp on its own lists) and
    of a small one (several blocks share a group of instances); N = 200 adds a second pass
    over the columns of [given; 1] to both.  Same instances -> bitwise the same q and h.
    Only q and h are requested, so the 43 GB of P and G are not allocated."""
    import ctypes

    import torch
    from mpc_interface_b200 import Param, device
    from mpc_interface_b200.examples import synthetic

    N, small, large = 200, 70, 9600
    form = cases.formulate_biped(cases.default_configuration(N), omega=Param("omega"),
                                 target_vel=[Param("vel_x"), Param("vel_y")])
    plan, ops = form._compiled(), form._ops()
    vals = synthetic.biped_inputs(77, large, N)
    given = torch.from_numpy(synthetic.pack_given(form.given_ID, form.given_len, vals, large)).cuda()
    scal = torch.from_numpy(plan.slots.evaluate({k: vals[k] for k in ("omega", "vel_x", "vel_y")}, large)).cuda()
    handle = ops.device_plan(plan).handle
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def qh(batch):
        q = torch.full([batch, plan.nv], float("nan"), dtype=torch.float64, device="cuda")
        h = torch.full([batch, plan.nc], float("nan"), dtype=torch.float64, device="cuda")
        device.check(device.lib().mpci_assemble(handle, scal.data_ptr(), given.data_ptr(), None,
                                                q.data_ptr(), None, h.data_ptr(), batch, stream))
        return q, h

    q_small, h_small = qh(small)
    q_large, h_large = qh(large)
    assert torch.equal(q_large[:small], q_small) and torch.equal(h_large[:small], h_small)
    assert bool(torch.isfinite(q_large).all()) and bool(torch.isfinite(h_large).all())
    b = large - 1  # and the last instance against the oracle
    one = {k: float(vals[k][b]) for k in ("omega", "vel_x", "vel_y")}
    ref = orc.all_qp_matrices(form, given[b].cpu().numpy()[:, None], one)
    assert cases.rel_err(h_large[b].cpu().numpy(), ref[1][:, 0]) <= cases.TOL
    assert cases.rel_err(q_large[b].cpu().numpy(), ref[3][:, 0]) <= cases.TOL


@pytest.mark.parametrize("batch,n,m,N", [(1000, 3, 1, 24), (7, 3, 1, 24), (1, 2, 1, 10), (513, 4, 2, 8),
                                         (260, 3, 1, 30), (33, 5, 1, 20), (64, 1, 1, 6)])
def test_k1_short_rows_warp_per_system(batch, n, m, N):
    """extend_small_kernel (a warp per system, plain 16-byte stores; rows shorter than 1 kB)
    against the three-kernel path (mpci_extend_matrices_static: another implementation of
    tools.py:14-33) bit for bit, and against NumPy powers on a few entries."""
    import ctypes

    import torch

    from mpc_interface_b200 import device

    rng = np.random.default_rng(batch + N)
    A = np.eye(n) + 0.1 * rng.normal(size=[batch, n, n])
    B = rng.normal(size=[batch, n, m])
    dA, dB = torch.from_numpy(A).cuda(), torch.from_numpy(B).cuda()
    got = {}
    for entry in ("mpci_extend_matrices", "mpci_extend_matrices_static"):
        S = torch.full([batch, N, n, n], float("nan"), dtype=torch.float64, device="cuda")
        U = torch.full([batch, m, N, N, n], float("nan"), dtype=torch.float64, device="cuda")
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        device.check(getattr(device.lib(), entry)(device._ptr(dA), device._ptr(dB), device._ptr(S),
                                                  device._ptr(U), batch, n, m, N, stream))
        torch.cuda.synchronize()
        got[entry] = (S.cpu().numpy(), U.cpu().numpy())
    S, U = got["mpci_extend_matrices"]
    assert np.array_equal(S, got["mpci_extend_matrices_static"][0])
    assert np.array_equal(U, got["mpci_extend_matrices_static"][1])
    b = batch - 1
    k, p = N - 1, 1
    assert np.allclose(S[b, k], np.linalg.matrix_power(A[b], k + 1).T, rtol=1e-12, atol=1e-13)
    assert np.allclose(U[b, m - 1, k, p], (np.linalg.matrix_power(A[b], k - p) @ B[b])[:, m - 1], rtol=1e-12, atol=1e-13)
    assert np.all(U[b, 0, 2, 3:] == 0.0)
    # S alone / U alone
    Sx = torch.full([batch, N, n, n], float("nan"), dtype=torch.float64, device="cuda")
    device.check(device.lib().mpci_extend_matrices(device._ptr(dA), device._ptr(dB), device._ptr(Sx), None,
                                                   batch, n, m, N, None))
    torch.cuda.synchronize()
    assert np.array_equal(Sx.cpu().numpy(), S)


@pytest.mark.parametrize("name,layout", [("biped_n24_batch", "dense"), ("biped_n100_batch", "csc")])
def test_box_bounds_batched_against_the_reference_matrices(name, layout):
    """lb / ub extraction: for every instance the bounds equal what the
    single-entry rows of the REFERENCE's A and h say, and the remaining rows are the
    reference's other rows -- together the same feasible set as A z <= h."""
    from mpc_interface_b200 import Param

    g = cases.golden(name)
    N, batch = int(g["N"]), g["A"].shape[0]
    form = cases.formulate_biped(cases.default_configuration(N), omega=Param("omega"),
                                 target_vel=[Param("vel_x"), Param("vel_y")])
    qp = form.generate_all_qp_matrices_batched(g["given"], {k: g[k] for k in ("omega", "vel_x", "vel_y")},
                                               layout=layout)
    lb, ub, rest = form.box_bounds_batched(qp)
    lb, ub = lb.cpu().numpy(), ub.cpu().numpy()
    nv = form.optim_len
    assert lb.shape == ub.shape == (batch, nv)
    for b in range(batch):
        A, h = g["A"][b], g["h"][b]
        want_lb, want_ub, general = np.full(nv, -np.inf), np.full(nv, np.inf), []
        for r in range(A.shape[0]):
            nz = np.nonzero(A[r])[0]
            if len(nz) == 1 and r < 4 * form.domain["Ds_x"]:
                v = h[r] / A[r, nz[0]]
                if A[r, nz[0]] > 0:
                    want_ub[nz[0]] = min(want_ub[nz[0]], v)
                else:
                    want_lb[nz[0]] = max(want_lb[nz[0]], v)
            else:
                general.append(r)
        assert list(rest) == general
        finite = np.isfinite(want_lb)
        assert np.array_equal(np.isfinite(lb[b]), finite) and np.array_equal(np.isfinite(ub[b]), np.isfinite(want_ub))
        assert np.abs(lb[b][finite] - want_lb[finite]).max() <= cases.TOL * np.abs(want_lb[finite]).max()
        assert np.abs(ub[b][finite] - want_ub[finite]).max() <= cases.TOL * np.abs(want_ub[finite]).max()
        assert np.all(lb[b][finite] <= ub[b][finite])
