"""The N>1 host logic on CPU: world_size-2 gloo processes assemble their shards
of one logical batch (NumPy stand-in backend: no GPU here) and the gathered
result must equal the single-process result, bit for bit."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mpc_interface_b200 import sharding

HERE = os.path.dirname(os.path.abspath(__file__))
BATCH, N = 11, 12  # odd batch: ragged shards


def test_shard_ranges_cover_the_batch():
    for total in (0, 1, 7, 64, 1048576):
        for world in (1, 2, 3, 8):
            spans = [sharding.shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sharding.shard_range(10, 2, 2)


def _assemble(rank, world):
    sys.path.insert(0, HERE)
    from _numpy_backend import NumpyBackend
    from oracle import qp_oracle
    import mpc_interface_b200.tools as tools
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import synthetic
    from mpc_interface_b200.examples.biped import default_configuration, formulate_biped

    tools.extend_matrices = qp_oracle.extend_matrices  # CPU test only
    form = formulate_biped(default_configuration(N), omega=Param("omega"),
                           target_vel=[Param("vel_x"), Param("vel_y")], backend=NumpyBackend())
    vals = sharding.shard_inputs(synthetic.biped_inputs(7, BATCH, N), rank, world)
    count = len(vals["omega"])
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, count)
    out = form.generate_all_qp_matrices_batched(
        given, {k: vals[k] for k in ("omega", "vel_x", "vel_y")})
    return out


def _worker(rank, world, port, result_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    out = _assemble(rank, world)
    q = sharding.gather_rows(torch.from_numpy(out.q), BATCH, rank, world)
    h = sharding.gather_rows(torch.from_numpy(out.h), BATCH, rank, world)
    sums = torch.tensor([out.q.sum(), out.h.sum(), out.P.sum(), out.G.sum()], dtype=torch.float64)
    parts = sharding.gather_small(sums, world)
    slowest = sharding.max_over_ranks(1.0 + rank, world)
    dist.barrier()
    if rank == 0:
        np.savez(os.path.join(result_dir, "gathered.npz"), q=q.numpy(), h=h.numpy(),
                 sums=torch.stack(parts).numpy(), slowest=slowest)
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_two_gloo_ranks_reproduce_the_single_process_batch(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    got = np.load(tmp_path / "gathered.npz")
    whole = _assemble(0, 1)
    assert np.array_equal(got["q"], whole.q) and np.array_equal(got["h"], whole.h)
    start, stop = sharding.shard_range(BATCH, 1, world)
    assert np.isclose(got["sums"][1, 2], whole.P[start:stop].sum(), rtol=1e-12)
    assert np.isclose(got["sums"][:, 3].sum(), whole.G.sum(), rtol=1e-12)
    assert float(got["slowest"]) == 2.0


def test_count_steps_batched_follows_the_example_loop():
    """receding.count_steps against the scalar logic of biped_mpc_loop.py:41-45,
    for instances at different phases."""
    from mpc_interface_b200.receding import count_steps

    n, n_steps, batch = 12, 4, 9
    times = np.stack([np.array([(i + 1) * n - 1 for i in range(n_steps)]) - b for b in range(batch)])
    count = np.arange(batch)
    t0, c0 = times.copy(), count.copy()
    for _ in range(30):
        new_t, new_c = count_steps(times, count, n)
        for b in range(batch):
            st, sc = times[b] - 1, count[b]
            if st[0] == -1:
                st, sc = st + n, sc + 1
            assert np.array_equal(new_t[b], st) and new_c[b] == sc
        times, count = new_t, new_c
    assert np.array_equal(t0[0], [(i + 1) * n - 1 for i in range(n_steps)]) and np.array_equal(c0, np.arange(batch))


def test_bind_host_near_gpu_on_a_fake_sysfs(tmp_path):
    """sharding.bind_host_near_gpu: reads the GPU's NUMA node and that node's CPU
    list from sysfs, keeps the CPUs this process may use; unknown node or missing
    files leave everything alone.  The binding itself runs in a child process."""
    import subprocess

    assert sharding._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert sharding.bind_host_near_gpu(0, 0x1B, 0, sysfs=str(tmp_path)) == {
        "numa_node": -1, "cpus": 0, "mempolicy": None}
    dev = tmp_path / "bus/pci/devices/0000:1b:00.0"
    dev.mkdir(parents=True)
    (dev / "numa_node").write_text("-1\n")
    assert sharding.bind_host_near_gpu(0, 0x1B, 0, sysfs=str(tmp_path))["numa_node"] == -1
    allowed = sorted(os.sched_getaffinity(0))
    keep = allowed[: max(1, len(allowed) // 2)]
    (dev / "numa_node").write_text("0\n")
    node = tmp_path / "devices/system/node/node0"
    node.mkdir(parents=True)
    (node / "cpulist").write_text(",".join(str(c) for c in keep) + ",4094-4095\n")
    code = ("import os, sys, json; sys.path.insert(0, {!r}); from mpc_interface_b200 import sharding; "
            "info = sharding.bind_host_near_gpu(0, 0x1B, 0, sysfs={!r}); "
            "print(json.dumps([info, sorted(os.sched_getaffinity(0))]))").format(
                os.path.dirname(os.path.dirname(os.path.abspath(__file__))), str(tmp_path))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, check=True).stdout
    import json

    info, bound = json.loads(out.strip().splitlines()[-1])
    assert info["numa_node"] == 0 and info["cpus"] == len(keep) and bound == keep
    assert sorted(os.sched_getaffinity(0)) == allowed  # the parent is untouched


def test_structure_keys_group_like_the_per_instance_key():
    """receding.StructureBuckets.group_keys (np.unique over integer key rows) forms the same
    buckets, in the same order, as the per-instance Python loop over structure_key."""
    from mpc_interface_b200.examples.biped import structure_key, structure_keys
    from mpc_interface_b200.receding import StructureBuckets

    rng = np.random.default_rng(0)
    B, N = 3000, 100
    first = rng.integers(0, 12, B)
    count = rng.integers(0, 50, B)
    times = first[:, None] + 12 * np.arange(9)[None, :]
    groups = StructureBuckets.group_keys(structure_keys(N, times, count))
    key = structure_key(N)
    loop = {}
    for b in range(B):
        loop.setdefault(key(step_times=times[b], step_count=count[b]), []).append(b)
    assert len(groups) == len(loop) == 24  # 12 phases x 2 feet
    assert [list(idx) for _, idx in groups] == list(loop.values())
    assert StructureBuckets.group_keys(np.zeros([0, 3], dtype=np.int64)) == []
