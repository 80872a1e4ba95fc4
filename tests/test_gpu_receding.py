"""Receding-horizon batches: robots at different walking
phases have different QP structures (nv, nc change);
StructureBuckets groups them, every bucket is one batched assembly."""
import numpy as np
import pytest

import _cases as cases
from oracle import qp_oracle as orc

pytestmark = pytest.mark.gpu


def _tick_sequence(n_ticks, step_samples, n_steps):
    """(step_times, step_count) per tick: the count_steps logic of the example
    loop (biped_mpc_loop.py:41-45), as in tests/golden/make_goldens.py."""
    step_times = np.array([(i + 1) * step_samples - 1 for i in range(n_steps)])
    step_count, out = 0, []
    for _ in range(n_ticks):
        out.append((step_times.copy(), step_count))
        step_times = step_times - 1
        if step_times[0] == -1:
            step_times = step_times + step_samples
            step_count += 1
    return out


@pytest.mark.parametrize("N", [24, 100])
def test_structure_buckets_against_oracle(N):
    from mpc_interface_b200 import Param, device
    from mpc_interface_b200.examples import synthetic
    from mpc_interface_b200.examples.biped import default_configuration, formulate_biped, structure_key
    from mpc_interface_b200.receding import StructureBuckets

    device.require_gpu()
    conf = default_configuration(N)

    def factory():
        return formulate_biped(conf, omega=Param("omega"), target_vel=[Param("vel_x"), Param("vel_y")])

    n_steps = factory().domain["Ds_x"]
    ticks = _tick_sequence(30, conf.step_samples, n_steps)
    batch = 50
    rng = np.random.default_rng(N)
    phase = rng.integers(0, len(ticks), size=batch)
    kargs = [dict(step_times=ticks[t][0], step_count=ticks[t][1]) for t in phase]
    vals = synthetic.biped_inputs(3, batch, N)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    buckets = StructureBuckets(factory, structure_key(N))
    results = buckets.assemble(kargs, vals, params)
    sizes = {(form.optim_len, out.nc) for _, _, form, out in results}
    assert len(sizes) >= 2  # the structure really is ragged across the batch
    assert sorted(np.concatenate([idx for _, idx, _, _ in results]).tolist()) == list(range(batch))
    seen = len(buckets)
    for _, idx, form, out in results:
        assert out.P.shape == (len(idx), form.optim_len, form.optim_len)
        for pos in (0, len(idx) - 1):
            b = int(idx[pos])
            ref_form = factory()
            ref_form.update(**kargs[b])  # an independent formulation at that instance's phase
            cases.check_layout(form, {  # same integer layout as the instance's own formulation
                "optim_names": np.array(list(ref_form.optim_ID)), "given_names": np.array(list(ref_form.given_ID)),
                "optim_ranges": np.array([[r.start, r.stop] for r in ref_form.optim_ID.values()]),
                "given_ranges": np.array([[r.start, r.stop] for r in ref_form.given_ID.values()])})
            given = ref_form.arrange_given({k: vals[k][b][:, None] for k in ref_form.given_ID})
            ref = orc.all_qp_matrices(ref_form, given, {k: float(v[b]) for k, v in params.items()})
            got = tuple(x[pos].cpu().numpy() for x in (out.G, out.h, out.P, out.q))
            cases.assert_qp_close(got, ref)
    # a second tick with the same phases reuses the compiled structures
    buckets.assemble(kargs, vals, params)
    assert len(buckets) == seen


@pytest.mark.parametrize("N", [24, 100])
def test_preview_all_and_advance_collector(N):
    """N = 100 is the size the bench runs at: [given; optim; 1] has 427 columns, which
    the sparse-row kernel stages in two column passes (round-1 advisor finding).
    The rest of the example's tick (biped_mpc_loop.py:62-65, :81-92) for a batch:
    every definition previewed in one launch, the next initial state read off
    the horizon (row 0 of the LIP states, row 1 of the step positions), the
    bias passed through; against the oracle's preview matrices and the loop's
    own per-instance assignments."""
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import synthetic
    from mpc_interface_b200.examples.biped import default_configuration, formulate_biped
    from mpc_interface_b200.receding import advance_collector

    batch = 45
    form = formulate_biped(default_configuration(N), omega=Param("omega"),
                           target_vel=[Param("vel_x"), Param("vel_y")])
    vals = synthetic.biped_inputs(11, batch, N)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, batch)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    optim = np.random.default_rng(4).normal(size=[batch, form.optim_len])
    motion = form.preview_all_batched(given, optim, params=params)
    assert list(motion) == list(form.definitions)
    collector = {name: vals[name] for name in form.given_ID}
    nxt = advance_collector(form, given, optim, {"LIP": 0, "steps": 1}, params, collector)
    assert set(nxt) == set(form.given_ID)
    for b in (0, 17, 32, batch - 1):
        one = {k: float(v[b]) for k, v in params.items()}
        PM = orc.preview_matrices(form, one)
        g, z = given[b][:, None], optim[b][:, None]
        ref = {name: Mg @ g + Mo @ z for name, (Mg, Mo) in PM.items()}
        for name, value in motion.items():
            assert tuple(value.shape) == (batch, ref[name].shape[0])
            assert cases.rel_err(value[b].cpu().numpy(), ref[name][:, 0]) <= cases.TOL, name
        # update_given_collector, restated on the reference's per-instance arrays
        want = {name: np.array(vals[name][b], dtype=float).reshape(-1) for name in form.given_ID}
        for state, ident in form.dynamics["LIP"].state_ID.items():
            want["x0" + state[-2:]][ident] = ref[state][0, 0]
        for state, ident in form.dynamics["steps"].state_ID.items():
            want["s0" + state[-2:]][ident] = ref[state][1, 0]
        for name in form.given_ID:
            got = nxt[name][b].cpu().numpy()
            if name.startswith("n_"):
                assert np.array_equal(got, want[name])  # passed through untouched
            else:
                assert cases.rel_err(got, want[name]) <= cases.TOL, name
    # the caller's collector is left as it was
    assert all(np.array_equal(collector[name], vals[name]) for name in collector)
    with pytest.raises(IndexError):
        advance_collector(form, given, optim, {"LIP": N}, params)


def test_device_tick_equals_the_per_instance_grouping():
    """StructureBuckets.assemble_device (np.unique grouping, inputs gathered on the GPU,
    output slots reused across ticks) gives bit-identical matrices to assemble()."""
    import torch
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import synthetic
    from mpc_interface_b200.examples.biped import (default_configuration, formulate_biped,
                                                   structure_key, structure_keys)
    from mpc_interface_b200.receding import StructureBuckets

    N, batch = 24, 333
    conf = default_configuration(N)

    def factory():
        return formulate_biped(conf, omega=Param("omega"), target_vel=[Param("vel_x"), Param("vel_y")])

    n_steps = factory().domain["Ds_x"]
    ticks = _tick_sequence(30, conf.step_samples, n_steps)
    phase = np.random.default_rng(1).integers(0, len(ticks), size=batch)
    times = np.stack([ticks[t][0] for t in phase])
    count = np.array([ticks[t][1] for t in phase])
    kargs = [dict(step_times=times[b], step_count=count[b]) for b in range(batch)]
    vals = synthetic.biped_inputs(5, batch, N)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    a = StructureBuckets(factory, structure_key(N))
    want = a.assemble(kargs, vals, params)
    b = StructureBuckets(factory, structure_key(N))
    collector = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in vals.items()
                 if k not in params}
    outs = {}
    got = b.assemble_device(structure_keys(N, times, count), lambda i: kargs[i], collector,
                            params=params, outs=outs)
    assert len(got) == len(want) and len(outs) == len(got)
    for (_, idx_w, _, qp_w), (_, idx_g, _, qp_g) in zip(want, got):
        assert np.array_equal(idx_w, idx_g.cpu().numpy())
        for x, y in ((qp_w.P, qp_g.P), (qp_w.q, qp_g.q), (qp_w.G, qp_g.G), (qp_w.h, qp_g.h)):
            assert torch.equal(x, y)
    # key rows computed and grouped on the device: the same buckets (ordered by key there)
    dev_keys = structure_keys(N, torch.from_numpy(times).cuda(), torch.from_numpy(count).cuda())
    assert np.array_equal(dev_keys.cpu().numpy(), structure_keys(N, times, count))
    on_dev = b.assemble_device(dev_keys, lambda i: kargs[i], collector, params=params)
    by_key = {k: (idx, qp) for k, idx, _, qp in got}
    assert len(on_dev) == len(got)
    for k, idx, _, qp in on_dev:
        assert np.array_equal(idx.cpu().numpy(), by_key[k][0].cpu().numpy())
        assert torch.equal(qp.P, by_key[k][1].P) and torch.equal(qp.h, by_key[k][1].h)
    # a second tick reuses plans and output slots
    slots = {k: v.P.data_ptr() for k, v in outs.items()}
    again = b.assemble_device(structure_keys(N, times, count), lambda i: kargs[i], collector,
                              params=params, outs=outs)
    assert {k: v.P.data_ptr() for k, v in outs.items()} == slots and len(again) == len(got)


@pytest.mark.parametrize("N,batch", [(24, 333), (100, 97)])
def test_advance_device_equals_advance_collector_per_bucket(N, batch):
    """StructureBuckets.advance_device (one launch per bucket over the 8 sampled rows, on the
    inputs the tick gathered in structure order, one scatter per variable) against
    advance_collector bucket by bucket (full previews) -- and so against the oracle's preview
    matrices, which test_preview_all_and_advance_collector pins."""
    import torch
    from mpc_interface_b200 import Param, receding
    from mpc_interface_b200.examples import synthetic
    from mpc_interface_b200.examples.biped import default_configuration, formulate_biped, structure_key, structure_keys

    conf = default_configuration(N)

    def factory():
        return formulate_biped(conf, omega=Param("omega"), target_vel=[Param("vel_x"), Param("vel_y")])

    n_steps = factory().domain["Ds_x"]
    ticks = _tick_sequence(3 * conf.step_samples, conf.step_samples, n_steps)
    phase = np.random.default_rng(2).integers(0, len(ticks), size=batch)
    times = np.stack([ticks[t][0] for t in phase])
    count = np.array([ticks[t][1] for t in phase])
    vals = synthetic.biped_inputs(6, batch, N)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    collector = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in vals.items() if k not in params}
    buckets = receding.StructureBuckets(factory, structure_key(N))
    keys = structure_keys(N, times, count)
    done = buckets.assemble_device(keys, lambda i: dict(step_times=times[i], step_count=count[i]), collector,
                                   params=params)
    plan0 = done[0][2]._compiled()
    scalars = torch.from_numpy(plan0.slots.evaluate(params, batch)).cuda()
    done = buckets.assemble_device(keys, None, collector, scalars=scalars)  # warm, shared scalar slots
    rng = np.random.default_rng(3)
    optims = [torch.from_numpy(rng.normal(size=[int(idx.shape[0]), form.optim_len]) * 0.1).cuda()
              for _, idx, form, _ in done]
    take = {"LIP": 0, "steps": 1}
    got = buckets.advance_device(done, optims, take, collector, scalars=scalars)
    want = buckets._advance_by_bucket(done, optims, take, collector, scalars, None)
    assert sorted(got) == sorted(want) == sorted(collector)
    moved = 0
    for name in want:
        a, b = got[name].cpu().numpy(), want[name].cpu().numpy()
        assert a.shape == b.shape
        assert np.abs(a - b).max() <= cases.TOL * max(1.0, np.abs(b).max()), name
        moved += int(np.any(b != collector[name].cpu().numpy().reshape(b.shape)))
        assert got[name].data_ptr() != collector[name].data_ptr()  # fresh tensors
    assert moved >= 4  # x0_x, x0_y, s0_x, s0_y changed; the bias passes through
    assert np.array_equal(got["n_x"].cpu().numpy(), vals["n_x"])
    # a stale `done` (not the last tick's) takes the per-bucket route and agrees as well
    stale = list(done)
    again = buckets.advance_device(stale, optims, take, collector, scalars=scalars)
    for name in want:
        assert torch.allclose(again[name], want[name], rtol=0, atol=1e-12)
    with pytest.raises(IndexError):
        buckets.advance_device(done, optims, {"LIP": N}, collector, scalars=scalars)
