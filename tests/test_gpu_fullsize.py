"""BASELINE.json's full sizes (configs[1..4]) through size-independent
properties: sampled instances against the oracle, bitwise determinism, a
checksum of checksums against an independent host formula, structure that must
not depend on the per-instance values.  Logical batches that do not fit in HBM
(C3: 1.21 TB, C4 / C5: 337 GB) go through a ring of two output slots.

Tolerance: relative error <= 1e-10 per array against the oracle (cases.TOL);
checksums are sums of up to 1e6 terms and are compared to 1e-9."""
import numpy as np
import pytest

import _cases as cases
from oracle import qp_oracle as orc
from oracle import reference

pytestmark = pytest.mark.gpu
SAMPLES = 64  # SURVEY.md 8(c) "Parity metric": >= 64 sampled instances per config


class _BipedExpect:
    """Expected (A, h, Q, q) of sampled biped instances: the unmodified reference where
    its staged copy is present (the GPU box: oracle/_ref), else the NumPy port."""

    def __init__(self, form, N):
        self.form = form
        self.ref_form = reference.biped_form(N) if reference.available() else None

    def __call__(self, vals, given, b):
        if self.ref_form is not None:
            return reference.biped_instance(self.ref_form, vals, b)
        one = {k: float(vals[k][b]) for k in ("omega", "vel_x", "vel_y")}
        return orc.all_qp_matrices(self.form, given[b][:, None], one)


@pytest.fixture(scope="module")
def biped100():
    from mpc_interface_b200 import Param, device

    device.require_gpu()
    form = cases.formulate_biped(
        cases.default_configuration(100), omega=Param("omega"), target_vel=[Param("vel_x"), Param("vel_y")])
    return form


def _monomial_values(plan, scal):
    table = plan.mono_table()
    coef = np.ones([scal.shape[0], table.shape[0]])
    for j in range(table.shape[0]):
        for f in table[j]:
            if f >= 0:
                coef[:, j] *= scal[:, f]
    return coef


def _expected_sums(plan, coef, given):
    """sum_b q_b, sum_b h_b, sum(P), sum(G) from the plan constants (host formula,
    independent of the kernels: sums commute with the affine maps)."""
    DY = plan.DY.cpu().numpy()
    ghat = np.hstack([given, np.ones([given.shape[0], 1])])
    y = np.zeros(DY.shape[1])
    for t, mono in enumerate(plan.Y_terms):
        y += DY[t] @ (coef[:, mono] @ ghat)
    sP = float((coef[:, plan.P_terms].sum(axis=0) * plan.DP.sum(dim=(1, 2)).cpu().numpy()).sum())
    sG = float((coef[:, plan.G_terms].sum(axis=0) * plan.DG.sum(dim=(1, 2)).cpu().numpy()).sum())
    return y[: plan.nv], y[plan.nv:], sP, sG


def _close(got, want, tol=1e-9):
    scale = max(1.0, float(np.max(np.abs(want))))
    return float(np.max(np.abs(np.asarray(got) - np.asarray(want)))) <= tol * scale


def test_c2_full_batch_65536(biped100):
    """configs[1]: 65,536 biped instances at N=100 in one call (75.6 GB)."""
    import torch
    from mpc_interface_b200.examples import synthetic

    form, B = biped100, 65536
    vals = synthetic.biped_inputs(1002, B, 100)
    vals["omega"][B // 2:] = vals["omega"][: B // 2]  # pairs of instances share omega
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, B)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    out = form.generate_all_qp_matrices_batched(given, params)
    assert out.P.shape == (B, 218, 218) and out.G.shape == (B, 440, 218)
    # sampled instances against the reference: first / last of the batch and of 32-instance
    # groups, the rest spread over the batch
    expect = _BipedExpect(form, 100)
    picks = sorted(set([0, 31, 32, 40000, B - 33, B - 1]) |
                   set(int(x) for x in np.random.default_rng(5).integers(0, B, SAMPLES - 6)))
    assert len(picks) >= SAMPLES - 2
    for b in picks:
        got = tuple(x[b].cpu().numpy() for x in (out.G, out.h, out.P, out.q))
        cases.assert_qp_close(got, expect(vals, given, b))
    # P and G depend on omega only: equal omega -> bitwise equal matrices
    assert torch.equal(out.P[: B // 2], out.P[B // 2:]) and torch.equal(out.G[: B // 2], out.G[B // 2:])
    # the 36 stepping-area rows do not depend on any per-instance value (probe8)
    assert bool((out.G[:, :36] == out.G[0, :36]).all())
    # checksum of checksums
    plan = form._compiled()
    coef = _monomial_values(plan, plan.slots.evaluate(params, B))
    sq, sh, sP, sG = _expected_sums(plan, coef, given)
    assert _close(out.q.sum(dim=0).cpu().numpy(), sq) and _close(out.h.sum(dim=0).cpu().numpy(), sh)
    assert _close(out.P.sum().item(), sP) and _close(out.G.sum().item(), sG)
    # determinism: a second pass into the same slot is bitwise identical
    keep = (out.q.clone(), out.h.clone(), out.P[1000:1016].clone(), out.G[-16:].clone())
    form.generate_all_qp_matrices_batched(given, params, out=out)
    assert torch.equal(keep[0], out.q) and torch.equal(keep[1], out.h)
    assert torch.equal(keep[2], out.P[1000:1016]) and torch.equal(keep[3], out.G[-16:])
    del out
    torch.cuda.empty_cache()


def test_c3_logical_batch_1048576_through_a_ring(biped100):
    """configs[2]: 1,048,576 instances; this GPU's shard of an 8-GPU job is
    131,072, the whole logical batch is streamed here through a ring of two
    32,768-instance slots (chunk checksums against the host formula)."""
    import torch
    from mpc_interface_b200 import device, sharding
    from mpc_interface_b200.examples import synthetic

    form, total, chunk = biped100, 1048576, 32768
    plan = form._compiled()
    ring = [device.QPBatch(chunk, plan.nv, plan.nc) for _ in range(2)]
    assert sharding.shard_range(total, 7, 8) == (917504, 1048576)
    rng_chunks = [0, 1, 15, 31]  # checksums are verified on these chunks, all chunks are assembled
    sums = torch.zeros(4, dtype=torch.float64, device="cuda")
    expect = np.zeros(4)
    reference_of = _BipedExpect(form, 100)
    for c in range(total // chunk):
        vals = synthetic.biped_inputs(1003 + c, chunk, 100)  # chunk c of the logical batch
        given = synthetic.pack_given(form.given_ID, form.given_len, vals, chunk)
        params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
        out = form.generate_all_qp_matrices_batched(given, params, out=ring[c % 2])
        if c in rng_chunks:
            sums += sharding.checksums(out)
            coef = _monomial_values(plan, plan.slots.evaluate(params, chunk))
            sq, sh, sP, sG = _expected_sums(plan, coef, given)
            expect += np.array([sq.sum(), sh.sum(), sP, sG])
            for b in [chunk - 1] + [int(x) for x in np.random.default_rng(c).integers(0, chunk, SAMPLES // 4 - 1)]:
                cases.assert_qp_close(tuple(x[b].cpu().numpy() for x in (out.G, out.h, out.P, out.q)),
                                      reference_of(vals, given, b))
    assert _close(sums.cpu().numpy(), expect)
    del ring
    torch.cuda.empty_cache()


@pytest.mark.parametrize("nx,nu,N,n_rows,total,chunk,seed,hessian,samples", [
    (12, 4, 50, 4, 262144, 16384, 1004, "recursion", 32),   # configs[3]: 2 x 32 samples
    (12, 4, 50, 4, 32768, 16384, 1004, "dmma", 32),         # same shape, tensor-core Hessian
    (6, 2, 400, 2, 16384, 1024, 1005, "dmma", 2),           # configs[4]: 2 x 2 samples (0.7 s each)
    (6, 2, 400, 2, 2048, 1024, 1005, "recursion", 2),       # same shape, displacement recursion
])
def test_generic_full_batches_per_instance_plants(monkeypatch, nx, nu, N, n_rows, total, chunk, seed,
                                                   hessian, samples):
    """configs[3] / configs[4] with one plant per instance (mode B), streamed
    through a ring of two slots; ``samples`` instances of the first and of the last chunk
    (their first / last instance and random ones) against the unmodified reference (the
    NumPy port where oracle/_ref is absent), P symmetric, the input-box rows independent
    of the plant."""
    import torch
    import mpc_interface_b200.tools as tools
    from _numpy_backend import NumpyBackend
    from mpc_interface_b200 import Param, device
    from mpc_interface_b200.examples import synthetic

    _, _, Q, R = synthetic.generic_system(seed, nx, nu)
    rows = np.random.default_rng(seed + 1).normal(size=[n_rows, nx])
    refs = [Param("ref%d" % i) for i in range(nx)]
    A0, B0 = synthetic.generic_plants(seed, 1, nx, nu)
    form = cases.formulate_generic(A0[0], B0[0], Q, R, N, rows, refs=refs)
    plan = form.lti_plan("plant")
    ring = [device.QPBatch(chunk, plan.nv, plan.nc) for _ in range(2)]
    n_chunks = total // chunk
    checked = []
    for c in range(n_chunks):
        A, B = synthetic.generic_plants(seed + 100 * c, chunk, nx, nu)
        vals = synthetic.generic_inputs(seed + 100 * c, chunk, nx)
        params = {"ref%d" % i: vals["ref"][:, i] for i in range(nx)}
        out = form.generate_all_qp_matrices_batched(
            vals["x0"], params, out=ring[c % 2], dynamics={"plant": (A, B)}, hessian=hessian)
        if c in (0, n_chunks - 1):
            sym = (out.P - out.P.transpose(1, 2)).abs().amax().item()
            assert sym <= 1e-11 * out.P.abs().amax().item()
            box = out.G[:, n_rows * N:]  # the 2 m N input-box rows: +-1 on the diagonal, plant-free
            assert bool((box == box[0]).all()) and float(box.abs().sum(dim=(1, 2)).max()) == 2 * nu * N
            picks = [0, chunk - 1]
            for x in np.random.default_rng(c).permutation(chunk):
                if len(picks) < samples and int(x) not in picks:
                    picks.append(int(x))
            for b in picks:
                checked.append((c, b, A[b], B[b], vals["x0"][b], vals["ref"][b],
                                tuple(x[b].cpu().numpy() for x in (out.G, out.h, out.P, out.q))))
    del ring
    torch.cuda.empty_cache()
    assert len(checked) == 2 * samples
    if reference.available():
        plants = reference.GenericPlants(A0[0], B0[0], Q, R, N, rows)
        for c, b, Ab, Bb, x0, ref_b, got in checked:
            cases.assert_qp_close(got, plants.instance(Ab, Bb, ref_b, x0))
        return
    monkeypatch.setattr(tools, "extend_matrices", orc.extend_matrices)
    for c, b, Ab, Bb, x0, ref_b, got in checked:
        ref_form = cases.formulate_generic(Ab, Bb, Q, R, N, rows, refs=list(ref_b), backend=NumpyBackend())
        cases.assert_qp_close(got, orc.all_qp_matrices(ref_form, x0[:, None]))
