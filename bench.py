#!/usr/bin/env python3
"""Benchmark: assembled QPs/s (P, q, G, h) of batched linear-MPC condensing.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl own|reference]
                    [--config c2|c3|c4|c5]

One step = one pass of the assembly over the configuration's batch of synthetic instances
(SURVEY.md 8(d)):

  c2  BASELINE configs[1]  biped N=100, 65,536 instances per GPU, per-instance omega / x0 /
                            bias / velocity aims                       (default at 1 GPU)
  c3  BASELINE configs[2]  the same formulation, ONE logical batch of 1,048,576 instances
                            in contiguous shards over the GPUs, streamed through a ring of
                            two 32,768-instance output slots          (default at N > 1)
  c4  BASELINE configs[3]  generic nx=12, nu=4, N=50, one plant per instance, 262,144
  c5  BASELINE configs[4]  generic nx=6, nu=2, N=400, one plant per instance, 16,384

Prints ONE JSON line (rank 0).  Under torchrun every rank drives its own GPU on its own
shard; there is no collective on the data path.  At one GPU the default line also carries
the c4 and c5 lines (`other_configs`), each with its own roofline / cpu_baseline / e2e.
`--impl reference` times the UNMODIFIED reference (oracle/_ref, staged by
oracle/make_ref.py) on the box's host cores.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

# stdout carries the ONE JSON line of the contract.  Libraries write to file descriptor 1 too
# (NCCL prints its version banner there under NCCL_DEBUG=VERSION, whatever NCCL_DEBUG_FILE says),
# so descriptor 1 is pointed at stderr for the whole run and the JSON line goes to a private
# copy of the original stdout.
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
_RESULT_OUT = os.fdopen(os.dup(1), "w")
sys.stdout.flush()
os.dup2(2, 1)


def emit(line):
    _RESULT_OUT.write(json.dumps(line) + "\n")
    _RESULT_OUT.flush()


ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

HORIZON = 100
C3_TOTAL, C3_CHUNK = 1048576, 32768
CONFIGS = {
    "c2": dict(kind="biped", seed=1002, batch=65536,
               metric="assembled QPs/sec (P,q,G,h) batched biped MPC N=100",
               workload="configs[1]: biped formulation (LIP jerk input + footsteps), N=100, "
                        "65,536 instances per GPU, per-instance omega / x0 / bias / velocity aims"),
    "c3": dict(kind="biped", seed=1003, batch=C3_TOTAL,
               metric="assembled QPs/sec (P,q,G,h) batched biped MPC N=100",
               workload="configs[2]: biped formulation, N=100, ONE logical batch of 1,048,576 "
                        "instances in contiguous shards over the GPUs, each shard streamed through "
                        "a ring of two 32,768-instance output slots"),
    "c4": dict(kind="lti", seed=1004, batch=262144, nx=12, nu=4, N=50, rows=4, chunk=16384, cpu_sample=2,
               metric="assembled QPs/sec (P,q,G,h) generic LTI MPC nx=12 nu=4 N=50, one plant per instance",
               workload="configs[3]: generic linear system nx=12, nu=4, N=50, dense Q/R (160 crossed "
                        "costs), input box + 4 linear rows per step, 262,144 instances, one plant "
                        "(A_b, B_b) per instance, streamed through a ring of two 16,384-instance slots"),
    "c5": dict(kind="lti", seed=1005, batch=16384, nx=6, nu=2, N=400, rows=2, chunk=1024, cpu_sample=1,
               metric="assembled QPs/sec (P,q,G,h) generic LTI MPC nx=6 nu=2 N=400, one plant per instance",
               workload="configs[4]: long-horizon stress test nx=6, nu=2, N=400, dense Q/R (40 crossed "
                        "costs), input box + 2 linear rows per step, 16,384 instances, one plant per "
                        "instance, streamed through a ring of two 1,024-instance slots"),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--config", default=None, choices=sorted(CONFIGS),
                    help="default: c2 on one GPU, c3 (the sharded logical batch) on several")
    ap.add_argument("--batch", type=int, default=0, help="override the instances per step (profiling runs)")
    ap.add_argument("--e2e-batch", type=int, default=65536, help="instances per GPU per e2e step")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-lti", action="store_true", help="skip the c4 / c5 lines and the extras")
    ap.add_argument("--no-numa-bind", action="store_true",
                    help="leave the rank's CPU affinity / memory policy alone")
    return ap.parse_args()


def measured_peaks():
    """HBM copy peak of the driver (MEASURED_PEAKS.json) plus this repo's own measurements of
    the write-only (fill) HBM peak and the FP64 DMMA peak (profiles/r02_peaks.json, written by
    scripts/peaks.cu on this pool's B200)."""
    out = {"hbm_gbs": 6650.0, "hbm_source": "fallback (B200_PROFILING.md)"}
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            out["hbm_gbs"] = json.load(f)["hbm_gbs"]
        out["hbm_source"] = "MEASURED_PEAKS.json (device copy, read + write)"
    out.update({"hbm_fill_gbs": None, "fp64_dmma_tflops": 36.9,
                "fp64_source": "profiles/r01_fp64_peak.txt (scripts/fp64_peak.cu)"})
    path = os.path.join(ROOT, "profiles", "r02_peaks.json")
    if os.path.exists(path):
        with open(path) as f:
            own = json.load(f)
        out["hbm_fill_gbs"] = own.get("hbm_fill_gbs")
        out["fp64_dmma_tflops"] = own.get("fp64_dmma_tflops", out["fp64_dmma_tflops"])
        out["fp64_source"] = "profiles/r02_peaks.json (scripts/peaks.cu)"
    return out


def measured_traffic(name):
    """dram read + write bytes per instance of the dominant kernel from the committed ncu
    --set full capture (a launch is linear in the batch: nothing is re-read), or None."""
    for path in (os.path.join(ROOT, "profiles", "r02_traffic.json"),
                 os.path.join(ROOT, "profiles", "r01_traffic.json")):
        if os.path.exists(path):
            with open(path) as f:
                data = json.load(f)
            if name in data:
                return data[name]["dram_bytes_per_instance"], os.path.relpath(path, ROOT)
            if name == "c2" and "dram_bytes_per_instance" in data:
                return data["dram_bytes_per_instance"], os.path.relpath(path, ROOT)
    return None, None


# ---------------------------------------------------------------------------
# CPU arm: the unmodified reference on the host cores
# ---------------------------------------------------------------------------
def _cpu_worker(job):
    """Times `count` instances of `config` starting at `start` in ONE process, one BLAS
    thread.  Returns (seconds, kind)."""
    name, start, count, total = job
    from threadpoolctl import threadpool_limits

    threadpool_limits(1)  # OPENBLAS_NUM_THREADS=1, one process per core
    cfg = CONFIGS[name]
    from oracle import reference

    if not reference.available():
        return _cpu_worker_port(job), "port"
    syn = reference.synthetic()
    if cfg["kind"] == "biped":
        # formulate once per process (SymPy discretisation etc. is set-up, not the path); per
        # instance: omega-dependent LineCombo coefficients + make_preview_matrices(),
        # Cost.update(aim=...), arrange_given, generate_all_qp_matrices (SURVEY App. C probe8)
        form = reference.biped_form(HORIZON)
        vals = syn.biped_inputs(cfg["seed"], total, HORIZON)
        run = lambda b: reference.biped_instance(form, vals, b)  # noqa: E731
    else:
        nx, nu, N = cfg["nx"], cfg["nu"], cfg["N"]
        _, _, Q, R = syn.generic_system(cfg["seed"], nx, nu)
        A, B = syn.generic_plants(cfg["seed"], total, nx, nu)
        rows = np.random.default_rng(cfg["seed"] + 1).normal(size=[cfg["rows"], nx])
        vals = syn.generic_inputs(cfg["seed"], total, nx)
        plants = reference.GenericPlants(A[0], B[0], Q, R, N, rows)
        run = lambda b: plants.instance(A[b], B[b], vals["ref"][b], vals["x0"][b])  # noqa: E731
    for b in range(start, start + min(3 if cfg["kind"] == "biped" else 1, count)):  # warm-up
        run(b)
    t0 = time.perf_counter()
    for b in range(start, start + count):
        run(b)
    return time.perf_counter() - t0, "reference"


def _cpu_worker_port(job):
    """Fallback when no reference tree is present: the NumPy port (oracle/qp_oracle.py) over
    this repo's host mirror -- biped only."""
    name, start, count, total = job
    cfg = CONFIGS[name]
    if cfg["kind"] != "biped":
        raise RuntimeError("no reference tree (oracle/_ref) for the generic configurations")
    from oracle import qp_oracle as orc
    import mpc_interface_b200.tools as tools
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import biped, synthetic

    tools.extend_matrices = orc.extend_matrices
    form = biped.formulate_biped(biped.default_configuration(HORIZON), omega=Param("omega"),
                                 target_vel=[Param("vel_x"), Param("vel_y")], preview=False)
    vals = synthetic.biped_inputs(cfg["seed"], total, HORIZON)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, total)
    names = ("omega", "vel_x", "vel_y")
    for b in range(start, start + min(3, count)):
        orc.all_qp_matrices(form, given[b][:, None], {k: float(vals[k][b]) for k in names})
    t0 = time.perf_counter()
    for b in range(start, start + count):
        orc.all_qp_matrices(form, given[b][:, None], {k: float(vals[k][b]) for k in names})
    return time.perf_counter() - t0


def cpu_baseline(name, per_core, cores):
    """QPs/s of the reference over `per_core` instances on each of `cores` processes."""
    import multiprocessing as mp

    os.environ["OPENBLAS_NUM_THREADS"] = "1"  # inherited by the spawned workers
    os.environ["OMP_NUM_THREADS"] = "1"
    cores, per_core = max(1, cores), max(1, per_core)
    jobs = [(name, i * per_core, per_core, per_core * cores) for i in range(cores)]
    if cores == 1:
        results = [_cpu_worker(jobs[0])]
    else:
        with mp.get_context("spawn").Pool(cores) as pool:
            results = pool.map(_cpu_worker, jobs)
    inner = max(t for t, _ in results)
    return per_core * cores / inner, per_core * cores, results[0][1]


def cpu_baseline_object(name, cores, per_core_all, per_core_one):
    """The `cpu_baseline` object of one line: all host cores (the value) and one core."""
    v_all, n_all, kind = cpu_baseline(name, per_core_all, cores)
    v_one, n_one, _ = cpu_baseline(name, per_core_one, 1)
    what = "unmodified reference (oracle/_ref)" if kind == "reference" else "NumPy port (oracle/qp_oracle.py)"
    return {"value": v_all, "unit": "QPs/s", "cores": cores, "kind": kind,
            "sample": "%d instances of the same workload over %d processes, 1 BLAS thread each: %s "
                      "driven through its own API, formulated once per process" % (n_all, cores, what),
            "one_core": {"value": v_one, "instances": n_one}}


# ---------------------------------------------------------------------------
# clocks during the timed region (B200_PROFILING.md)
# ---------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.samples, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS,
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(s[0]) for s in self.samples if len(s) >= 6 and s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if len(s) >= 6 and s[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names)
                   if any(len(s) >= 6 and s[2 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


class Env:
    """Rank / world / device of this process and the distributed plumbing."""

    def __init__(self, args):
        import torch

        self.torch = torch
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist

            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist
        from mpc_interface_b200 import sharding

        self.sharding = sharding
        props = torch.cuda.get_device_properties(self.local)
        self.numa = (sharding.bind_host_near_gpu(props.pci_domain_id, props.pci_bus_id, props.pci_device_id)
                     if not args.no_numa_bind else None)

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_ms(self, ms):
        return self.sharding.max_over_ranks(ms, self.world, "cuda")

    def timed_steps(self, step, warmup, steps, sampler=None):
        """W untimed steps, then exactly K steps between barriers, CUDA events on the
        launching stream; returns ms for the K steps, max over ranks."""
        torch = self.torch
        for _ in range(warmup):
            step()
        if sampler is not None:
            sampler.start()
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            step()
        e1.record()
        self.barrier()
        return self.max_ms(e0.elapsed_time(e1))

    def timed(self, fn, reps):
        torch = self.torch
        fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    def wall_rate(self, step, units_per_step, warmup, steps):
        """Units/s of `step` by wall clock with a sync on both sides (host-buffer paths: the
        call returns when the results are in host memory), whole job, max over ranks."""
        torch = self.torch
        for _ in range(warmup):
            step()
        self.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            step()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if self.dist is not None:
            t = torch.tensor([dt], device="cuda", dtype=torch.float64)
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
            dt = float(t.item())
        return units_per_step * self.world * steps / dt


# ---------------------------------------------------------------------------
# GPU arm, biped (mode A: shared dynamics, per-instance scalars)  -- c2 / c3
# ---------------------------------------------------------------------------
def c3_chunk_inputs(synthetic, chunk_index):
    """Chunk `chunk_index` of the 1,048,576-instance logical batch: instance i's inputs depend
    on i only (seed 1003 + chunk index), so results do not depend on the number of GPUs."""
    return synthetic.biped_inputs(CONFIGS["c3"]["seed"] + chunk_index, C3_CHUNK, HORIZON)


def run_biped(args, env, name):
    torch = env.torch
    from mpc_interface_b200 import Param, device
    from mpc_interface_b200.examples import synthetic
    from mpc_interface_b200.examples.biped import default_configuration, formulate_biped

    cfg = CONFIGS[name]
    rank, world = env.rank, env.world
    form = formulate_biped(default_configuration(HORIZON), omega=Param("omega"),
                           target_vel=[Param("vel_x"), Param("vel_y")])
    plan, ops = form._compiled(), form._ops()
    nv, ng, nc = plan.nv, plan.ng, plan.nc
    names = ("omega", "vel_x", "vel_y")
    if name == "c2":  # weak: every GPU its own 65,536 (one launch per array)
        B = args.batch or cfg["batch"]
        vals = env.sharding.shard_inputs(synthetic.biped_inputs(cfg["seed"], B * world, HORIZON), rank, world)
        chunks = [vals]
        per_gpu, total, chunk = B, B * world, B
    else:  # c3, strong: the rank's contiguous shard of the logical batch, chunk by chunk
        total = args.batch or cfg["batch"]
        n_chunks = total // C3_CHUNK
        if n_chunks % world:
            raise SystemExit("c3: %d chunks do not split over %d ranks" % (n_chunks, world))
        mine = range(rank * n_chunks // world, (rank + 1) * n_chunks // world)
        chunks = [c3_chunk_inputs(synthetic, c) for c in mine]
        per_gpu, chunk = len(mine) * C3_CHUNK, C3_CHUNK
    given, scalars = [], []
    for vals in chunks:
        given.append(torch.from_numpy(synthetic.pack_given(form.given_ID, form.given_len, vals, chunk)).cuda())
        scalars.append(torch.from_numpy(plan.slots.evaluate({k: vals[k] for k in names}, chunk)).cuda())
    ring = [device.QPBatch(chunk, nv, nc) for _ in range(min(2, len(chunks)))]
    lib, handle = device.lib(), ops.device_plan(plan).handle
    stream = torch.cuda.current_stream().cuda_stream
    null = None

    def call(c, P, q, G, h):
        device.check(lib.mpci_assemble(handle, scalars[c].data_ptr(), given[c].data_ptr(),
                                       P, q, G, h, chunk, stream))

    def step():
        for c in range(len(chunks)):
            out = ring[c % len(ring)]
            call(c, out.P.data_ptr(), out.q.data_ptr(), out.G.data_ptr(), out.h.data_ptr())

    sampler = ClockSampler(env.local) if rank == 0 else None
    for _ in range(args.warmup):
        step()
    if sampler:
        sampler.start()
    env.barrier()
    # the library's own kernel timers: CUDA events around each launch, on the stream it
    # runs on, inside the timed region (the roofline figure is read from these)
    device.check(lib.mpci_plan_set_timing(handle, 1))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    env.barrier()
    ms = e0.elapsed_time(e1)
    in_step = (ctypes.c_double * 3)()
    pieces = ctypes.c_long(0)
    device.check(lib.mpci_plan_read_timing(handle, ctypes.byref(in_step), ctypes.byref(pieces)))
    device.check(lib.mpci_plan_set_timing(handle, 0))
    per_launch = [v / max(1, pieces.value) for v in in_step]  # P stream, G stream, [q;h]
    clocks = sampler.stop() if sampler else None
    ms = env.max_ms(ms)

    out = ring[0]
    reps = max(3, min(args.steps, 10))
    t_P = env.timed(lambda: call(0, out.P.data_ptr(), null, null, null), reps)
    t_G = env.timed(lambda: call(0, null, null, out.G.data_ptr(), null), reps)
    t_qh = env.timed(lambda: call(0, null, out.q.data_ptr(), null, out.h.data_ptr()), reps)

    # solver hand-off layout (CSC values, one pattern per batch): same kernels, fewer bytes
    pP, pG = ops.compact(plan)
    csc = device.QPBatchCSC(chunk, nv, nc, pP, pG)
    t_csc = env.timed(lambda: device.check(lib.mpci_assemble_compact(
        handle, scalars[0].data_ptr(), given[0].data_ptr(), csc._P.data_ptr(), csc.q.data_ptr(),
        csc._G.data_ptr(), csc.h.data_ptr(), chunk, stream)), reps)
    csc_bytes = device.QPBatchCSC.bytes_per_instance(nv, nc, pP, pG)
    del csc

    # SURVEY 7.3#2 asks for the other mode beside the headline: omega shared by the batch, so
    # P and G are the same for every instance (one constant term each) and only q, h differ.
    # The same kernels still WRITE all four arrays per instance; reported separately.
    shared = None
    if rank == 0 and not args.no_lti:
        form_s = formulate_biped(default_configuration(HORIZON), target_vel=[Param("vel_x"), Param("vel_y")])
        plan_s = form_s._compiled()
        handle_s = form_s._ops().device_plan(plan_s).handle
        scal_s = torch.from_numpy(plan_s.slots.evaluate({k: chunks[0][k] for k in ("vel_x", "vel_y")}, chunk)).cuda()
        t_shared = env.timed(lambda: device.check(lib.mpci_assemble(
            handle_s, scal_s.data_ptr(), given[0].data_ptr(), out.P.data_ptr(), out.q.data_ptr(),
            out.G.data_ptr(), out.h.data_ptr(), chunk, stream)), reps)
        shared = {"value": chunk / (t_shared * 1e-3), "unit": "QPs/s per GPU", "ms_per_step": t_shared,
                  "P_terms": len(plan_s.P_terms), "G_terms": len(plan_s.G_terms),
                  "note": "omega shared by the batch (P and G equal for all instances, still written per "
                          "instance); reported separately from the headline, which has per-instance omega"}
        step()  # leave the headline mode's results in the ring for the checksums below
        torch.cuda.synchronize()

    # optional final gather (not on the data path): per-shard checksums over NCCL
    gathered = env.sharding.gather_small(env.sharding.checksums(ring[-1]), world)
    assert len(gathered) == world and all(bool(torch.isfinite(g).all()) for g in gathered)

    launches = lib.mpci_assemble_launches(handle) * len(chunks) * world
    del ring, out
    torch.cuda.empty_cache()

    e2e = None if args.no_e2e else e2e_biped(args, env, form, plan, ops, name, per_gpu)
    if rank != 0:
        return None
    peaks = measured_peaks()
    steps = args.steps
    bytes_G, bytes_P = 8.0 * nc * nv * chunk, 8.0 * nv * nv * chunk
    bytes_all = 8.0 * (nv * nv + nv + nc * nv + nc) * per_gpu
    achieved = bytes_G / (per_launch[1] * 1e-3) / 1e9
    per_inst, traffic_src = measured_traffic("c2")
    line = {
        "metric": cfg["metric"], "value": total * steps / (ms * 1e-3), "unit": "QPs/s",
        "n_gpus": world, "steps": steps, "warmup": args.warmup, "ms_per_step": ms / steps,
        "higher_is_better": True, "scaling": "weak" if name == "c2" else "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": cfg["workload"], "name": name, "horizon": HORIZON,
                   "instances_per_step": total, "instances_per_gpu": per_gpu,
                   "instances_per_launch": chunk, "nv": nv, "ng": ng, "nc": nc, "seed": cfg["seed"],
                   "sharding": "contiguous shards, no data-path collective",
                   "l2": "every launch writes %.1f GB, far beyond the 126 MB L2 (no flush needed)"
                         % (8e-9 * (nv * nv + nv + nc * nv + nc) * chunk)},
        "clocks": clocks, "e2e": e2e, "host_binding_rank0": env.numa,
        "gpu_launches": launches * steps,
        "roofline": {
            "bound": "hbm", "kernel": "stream_terms_kernel<3,double2> (G stack)",
            "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
            "frac": achieved / peaks["hbm_gbs"],
            "traffic": per_inst * chunk if per_inst else None,
            "traffic_note": "ncu dram bytes per instance (%s) x the instances of a launch" % traffic_src,
            "peak_source": peaks["hbm_source"],
            "peak_fill": peaks["hbm_fill_gbs"],
            "frac_of_fill_peak": achieved / peaks["hbm_fill_gbs"] if peaks["hbm_fill_gbs"] else None,
            "peak_fill_source": "profiles/r02_peaks.json: write-only double2 fill of 8 GiB (scripts/peaks.cu)",
            "algorithmic_bytes_per_launch": bytes_G, "ms_per_launch": per_launch[1],
            "timing": "CUDA events around the launch on its stream, inside the timed region "
                      "([q;h] runs beside it on a side stream)",
            "note": "peak = a device copy (read + write); the kernel only writes, so the fraction "
                    "against the write-only fill peak is the honest one"},
        "kernels_ms_in_step": {"P_stream": per_launch[0], "G_stream": per_launch[1],
                               "qh_side_stream": per_launch[2], "launches_of_each": pieces.value // max(1, steps)},
        "kernels_ms": {"P_stream": t_P, "G_stream": t_G, "qh": t_qh,
                       "note": "each kernel launched alone, after the timed region"},
        "kernels_gbs": {"P_stream": bytes_P / t_P / 1e6, "G_stream": bytes_G / t_G / 1e6,
                        "step_all_outputs": bytes_all / (ms / steps) / 1e6},
        "csc_layout": {"value": chunk / (t_csc * 1e-3), "unit": "QPs/s per GPU", "ms_per_launch_set": t_csc,
                       "bytes_per_qp": csc_bytes, "nnz_P": pP.nnz, "nnz_G": pG.nnz,
                       "gbs": csc_bytes * chunk / t_csc / 1e6,
                       "note": "same step writing only the CSC values of P and G (solver hand-off "
                               "layout); not the headline"},
    }
    if shared:
        line["shared_omega_mode"] = shared
    return line


def d2h_ceiling(env, bytes_per_chunk, n_chunks, host_views):
    """Plain-copy ceiling of the host-buffer path: the same bytes per chunk from a device
    buffer into the same pinned ring, alternating two streams, no kernels."""
    torch = env.torch
    dev = torch.empty([bytes_per_chunk // 8], dtype=torch.float64, device="cuda")
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]

    def sweep():
        for c in range(n_chunks):
            with torch.cuda.stream(streams[c & 1]):
                at = 0
                for view in host_views(c):
                    flat = view.reshape(-1)
                    flat.copy_(dev[at:at + flat.numel()], non_blocking=True)
                    at += flat.numel()
        torch.cuda.synchronize()

    sweep()
    env.barrier()
    t0 = time.perf_counter()
    sweep()
    dt = time.perf_counter() - t0
    if env.dist is not None:
        t = torch.tensor([dt], device="cuda", dtype=torch.float64)
        env.dist.all_reduce(t, op=env.dist.ReduceOp.MAX)
        dt = float(t.item())
    return bytes_per_chunk * n_chunks * env.world / dt / 1e9


def e2e_biped(args, env, form, plan, ops, name, per_gpu):
    """The metric through the C-ABI entry point with HOST buffers (mpci_assemble_host_stream):
    every step copies that step's inputs from pinned host memory to the device and all of
    P, q, G, h of every instance back into a ring of pinned host slots."""
    torch = env.torch
    from mpc_interface_b200.examples import synthetic

    Be = min(args.e2e_batch, per_gpu)
    chunk, ring = min(512, Be), 8
    ring = max(2, min(ring, (Be + chunk - 1) // chunk))
    vals = synthetic.biped_inputs(CONFIGS[name]["seed"] + 1000 + env.rank, Be, HORIZON)
    given = torch.from_numpy(synthetic.pack_given(form.given_ID, form.given_len, vals, Be)).pin_memory()
    scalars = torch.from_numpy(
        plan.slots.evaluate({k: vals[k] for k in ("omega", "vel_x", "vel_y")}, Be)).pin_memory()
    nv, nc = plan.nv, plan.nc
    slots = ring * chunk
    host = [torch.empty(shape, dtype=torch.float64).pin_memory()
            for shape in ([slots, nv, nv], [slots, nv], [slots, nc, nv], [slots, nc])]
    steps, warm = max(1, min(args.e2e_steps, args.steps)), 1
    dense = env.wall_rate(lambda: ops.assemble_host_stream(
        plan, scalars, given, host[0], host[1], host[2], host[3], Be, chunk, ring), Be, warm, steps)
    h2d = int(scalars.numel() + given.numel()) * 8
    per_qp = 8 * (nv * nv + nv + nc * nv + nc)
    n_chunks = (Be + chunk - 1) // chunk
    ceiling = d2h_ceiling(env, per_qp * chunk, n_chunks,
                          lambda c: [x[(c % ring) * chunk:(c % ring + 1) * chunk] for x in host])
    # the solver hand-off layout through the same kind of call: CSC values only
    pP, pG = ops.compact(plan)
    hostc = [torch.empty(shape, dtype=torch.float64).pin_memory()
             for shape in ([slots, pP.padded], [slots, nv], [slots, pG.padded], [slots, nc])]
    sparse = env.wall_rate(lambda: ops.assemble_host_stream(
        plan, scalars, given, hostc[0], hostc[1], hostc[2], hostc[3], Be, chunk, ring, layout="csc"),
        Be, warm, steps)
    csc_per_qp = 8 * (pP.padded + nv + pG.padded + nc)
    return {"value": dense, "unit": "QPs/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": per_qp * Be,
            "batch_per_gpu": Be, "steps": steps, "chunk": chunk, "host_ring_slots": ring,
            "d2h_gbs": dense * per_qp / 1e9,
            "d2h_ceiling_gbs": ceiling, "frac_of_d2h_ceiling": dense * per_qp / 1e9 / ceiling,
            "note": "mpci_assemble_host_stream: all of P,q,G,h (dense, as the reference returns them) "
                    "copied back to pinned host memory, chunk by chunk into a ring of host slots; "
                    "bound by the PCIe link.  d2h_ceiling_gbs = plain cudaMemcpyAsync of the same bytes "
                    "into the same ring, all ranks at once, no kernels (aggregate GB/s)",
            "csc": {"value": sparse, "unit": "QPs/s", "d2h_bytes_per_step": csc_per_qp * Be,
                    "d2h_gbs": sparse * csc_per_qp / 1e9,
                    "note": "mpci_assemble_compact_host_stream: CSC values of P and G (pattern shared "
                            "by the batch: what biped_mpc_loop.py:57-58 hands the solver) + dense q, h"}}


# ---------------------------------------------------------------------------
# GPU arm, per-instance plants (mode B: lti_assemble_kernel)  -- c4 / c5
# ---------------------------------------------------------------------------
def run_lti(args, env, name, nested=False):
    torch = env.torch
    from mpc_interface_b200 import Param, device, lti
    from mpc_interface_b200.examples import synthetic
    from mpc_interface_b200.examples.generic import formulate_generic

    cfg = CONFIGS[name]
    rank, world = env.rank, env.world
    nx, nu, N, chunk = cfg["nx"], cfg["nu"], cfg["N"], cfg["chunk"]
    total = (args.batch if (args.batch and not nested) else cfg["batch"])
    chunk = min(chunk, total)
    n_chunks = total // chunk
    if n_chunks % world:
        raise SystemExit("%s: %d chunks do not split over %d ranks" % (name, n_chunks, world))
    mine = range(rank * n_chunks // world, (rank + 1) * n_chunks // world)
    _, _, Q, R = synthetic.generic_system(cfg["seed"], nx, nu)
    rows = np.random.default_rng(cfg["seed"] + 1).normal(size=[cfg["rows"], nx])
    refs = [Param("ref%d" % i) for i in range(nx)]
    A0, B0 = synthetic.generic_plants(cfg["seed"], 1, nx, nu)
    form = formulate_generic(A0[0], B0[0], Q, R, N, rows, refs=refs)
    plan = form.lti_plan("plant")
    dp = plan.device_plan()
    nv, nc = plan.nv, plan.nc
    host_in, dev_in = [], []
    for c in mine:  # chunk c of the logical batch: the plants and inputs depend on c only
        A, B = synthetic.generic_plants(cfg["seed"] + 100 * c, chunk, nx, nu)
        vals = synthetic.generic_inputs(cfg["seed"] + 100 * c, chunk, nx)
        scal = plan.slots.evaluate({"ref%d" % i: vals["ref"][:, i] for i in range(nx)}, chunk)
        host_in.append((A, B, vals["x0"], scal))
        dev_in.append(tuple(torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in host_in[-1]))
    ring = [device.QPBatch(chunk, nv, nc) for _ in range(min(2, len(mine)))]
    lib = device.lib()
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    null = ctypes.c_void_p(0)

    def call(c, P, q, G, h, mode):
        dA, dB, dx, ds = dev_in[c]
        device.check(lib.mpci_lti_assemble(dp.handle, dA.data_ptr(), nx * nx, dB.data_ptr(), nx * nu,
                                           dx.data_ptr(), ds.data_ptr(), P, q, G, h, chunk, mode, stream))

    events = []

    def make_step(mode, record):
        def step():
            for c in range(len(mine)):
                out = ring[c % len(ring)]
                if record:
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record()
                call(c, out.P.data_ptr(), out.q.data_ptr(), out.G.data_ptr(), out.h.data_ptr(), mode)
                if record:
                    b.record()
                    events.append((a, b))
        return step

    per_gpu = len(mine) * chunk
    results = {}
    sampler = ClockSampler(env.local) if (rank == 0 and not nested) else None
    for label, mode in (("recursion", lti.HESSIAN_RECURSION), ("dmma", lti.HESSIAN_DMMA)):
        for _ in range(args.warmup):
            make_step(mode, False)()
        del events[:]
        ms = env.timed_steps(make_step(mode, True), 0, args.steps, sampler if label == "recursion" else None)
        launch_ms = float(np.mean([a.elapsed_time(b) for a, b in events]))
        if label == "recursion" and sampler:
            clocks = sampler.stop()
        out = ring[0]
        reps = max(2, min(args.steps, 5))
        t_P = env.timed(lambda: call(0, out.P.data_ptr(), null, null, null, mode), reps)
        t_G = env.timed(lambda: call(0, null, null, out.G.data_ptr(), out.h.data_ptr(), mode), reps)
        t_0 = env.timed(lambda: call(0, null, null, null, null, mode), reps)
        results[label] = dict(ms=ms, launch_ms=launch_ms, t_P=t_P, t_G=t_G, t_0=t_0,
                              flops=plan.hessian_flops(mode), info=plan.launch_info(mode))
    if sampler is None:
        clocks = None
    del ring
    torch.cuda.empty_cache()

    e2e = None
    if not args.no_e2e:
        e2e = e2e_lti(args, env, plan, lti, host_in, chunk, nv, nc)
    if rank != 0:
        return None
    peaks = measured_peaks()
    steps = args.steps
    per_qp = 8.0 * (nv * nv + nv + nc * nv + nc)
    structural = 2.0 * (nx * N) * (nu * N) ** 2 + 2.0 * nx * nx * N * (nu * N)
    rec, dm = results["recursion"], results["dmma"]
    achieved = per_qp * chunk / (rec["launch_ms"] * 1e-3) / 1e9
    dmma_tflops = dm["flops"] * chunk / (dm["t_P"] * 1e-3) / 1e12
    per_inst, traffic_src = measured_traffic(name)
    line = {
        "metric": cfg["metric"], "value": total * steps / (rec["ms"] * 1e-3), "unit": "QPs/s",
        "n_gpus": world, "steps": steps, "warmup": args.warmup, "ms_per_step": rec["ms"] / steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": cfg["workload"], "name": name, "nx": nx, "nu": nu, "horizon": N,
                   "instances_per_step": total, "instances_per_gpu": per_gpu,
                   "instances_per_launch": chunk, "nv": nv, "ng": nx, "nc": nc, "seed": cfg["seed"],
                   "hessian": "displacement recursion (the headline); the dense DMMA contraction is "
                              "in dmma_mode",
                   "l2": "every launch writes %.1f GB, far beyond the 126 MB L2" % (per_qp * chunk / 1e9)},
        "clocks": clocks, "e2e": e2e, "gpu_launches": len(mine) * world * steps,
        "roofline": {
            "bound": "hbm", "kernel": "lti_assemble_kernel<recursion> (prediction, P, q, G, h fused)",
            "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
            "frac": achieved / peaks["hbm_gbs"], "peak_source": peaks["hbm_source"],
            "peak_fill": peaks["hbm_fill_gbs"],
            "frac_of_fill_peak": achieved / peaks["hbm_fill_gbs"] if peaks["hbm_fill_gbs"] else None,
            "traffic": per_inst * chunk if per_inst else None,
            "traffic_note": ("ncu dram bytes per instance (%s) x the instances of a launch" % traffic_src)
                            if per_inst else None,
            "algorithmic_bytes_per_launch": per_qp * chunk, "ms_per_launch": rec["launch_ms"],
            "timing": "CUDA events around every launch on its stream, inside the timed region"},
        "kernels_ms": {"P_only": rec["t_P"], "Gh_only": rec["t_G"], "prediction_only": rec["t_0"],
                       "all": rec["launch_ms"], "note": "one launch of %d instances; phases alone after "
                       "the timed region" % chunk,
                       "gbs_P": 8.0 * nv * nv * chunk / rec["t_P"] / 1e6,
                       "gbs_Gh": 8.0 * (nc * nv + nc) * chunk / rec["t_G"] / 1e6},
        "launch": rec["info"],
        "dmma_mode": {
            "value": total * steps / (dm["ms"] * 1e-3), "unit": "QPs/s", "ms_per_step": dm["ms"] / steps,
            "hbm_gbs_all_outputs": per_qp * chunk / (dm["launch_ms"] * 1e-3) / 1e9,
            "roofline": {"bound": "tensor", "kernel": "lti_assemble_kernel<dmma>, Hessian phase (P-only launch)",
                         "achieved": dmma_tflops, "peak": peaks["fp64_dmma_tflops"], "unit": "TFLOP/s",
                         "frac": dmma_tflops / peaks["fp64_dmma_tflops"], "peak_source": peaks["fp64_source"],
                         "flops_per_qp_executed": dm["flops"], "flops_per_qp_structural": structural,
                         "ms_per_launch": dm["t_P"], "traffic": None,
                         "note": "FP64 DMMA.8x8x4 (tcgen05 has no FP64 kind); executed flops after "
                                 "block-triangular skipping"},
            "kernels_ms": {"P_only": dm["t_P"], "Gh_only": dm["t_G"], "prediction_only": dm["t_0"],
                           "all": dm["launch_ms"]},
            "launch": dm["info"]},
    }
    return line


def e2e_lti(args, env, plan, lti, host_in, chunk, nv, nc):
    """mpci_lti_assemble_host_stream over the rank's shard: plants and inputs from pinned host
    memory, all of P, q, G, h back into a ring of pinned host slots (one pass, recursion mode)."""
    torch = env.torch
    A = np.concatenate([x[0] for x in host_in])
    B = np.concatenate([x[1] for x in host_in])
    x0 = np.concatenate([x[2] for x in host_in])
    scal = np.concatenate([x[3] for x in host_in])
    Be = x0.shape[0]
    per_qp = 8 * (nv * nv + nv + nc * nv + nc)
    sub = max(1, min(chunk, (512 << 20) // per_qp))  # ~0.5 GB per host slot
    ring = 4
    slots = ring * sub
    host = [torch.empty(shape, dtype=torch.float64).pin_memory()
            for shape in ([slots, nv, nv], [slots, nv], [slots, nc, nv], [slots, nc])]
    pin = [torch.from_numpy(np.ascontiguousarray(x)).pin_memory().numpy() for x in (A, B, x0)]
    params_scal = torch.from_numpy(np.ascontiguousarray(scal)).pin_memory().numpy()
    from mpc_interface_b200 import device

    S = scal.shape[1]
    hp = lambda a: ctypes.c_void_p(a.ctypes.data)  # noqa: E731

    def run(count):
        device.check(device.lib().mpci_lti_assemble_host_stream(
            plan.device_plan().handle, hp(pin[0]), plan.n * plan.n, hp(pin[1]), plan.n * plan.m,
            hp(pin[2]), hp(params_scal) if S else ctypes.c_void_p(0),
            *(ctypes.c_void_p(t.data_ptr()) for t in host), count, lti.HESSIAN_RECURSION, sub, ring,
            device.ChunkFn(0), None))

    run(min(Be, 2 * sub))  # warm-up: allocates the device slots
    rate = env.wall_rate(lambda: run(Be), Be, 0, 1)
    n_chunks = (Be + sub - 1) // sub
    ceiling = d2h_ceiling(env, per_qp * sub, n_chunks,
                          lambda c: [x[(c % ring) * sub:(c % ring + 1) * sub] for x in host])
    h2d = int(A.size + B.size + x0.size + scal.size) * 8
    return {"value": rate, "unit": "QPs/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": per_qp * Be,
            "batch_per_gpu": Be, "steps": 1, "chunk": sub, "host_ring_slots": ring,
            "d2h_gbs": rate * per_qp / 1e9, "d2h_ceiling_gbs": ceiling,
            "frac_of_d2h_ceiling": rate * per_qp / 1e9 / ceiling,
            "note": "mpci_lti_assemble_host_stream: per-instance A, B, x0, refs uploaded, all of "
                    "P,q,G,h copied back to a ring of pinned host slots; bound by the PCIe link"}


def run_single():
    """scripts/latency.py: ONE instance through the reference's own call
    (Formulation.generate_all_qp_matrices, NumPy in / NumPy out), wall clock per call."""
    import importlib.util

    spec = importlib.util.spec_from_file_location("latency", os.path.join(ROOT, "scripts", "latency.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return [mod.measure(N, 200) for N in (24, 100)]


def run_extras():
    """Builder-side extras that ride along in the default line (scripts/bench_extras.py)."""
    import importlib.util

    path = os.path.join(ROOT, "scripts", "bench_extras.py")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location("bench_extras", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.run()


# ---------------------------------------------------------------------------
def reference_arm(args, name, cores):
    """--impl reference: the unmodified reference on all host cores, exactly K timed steps; a
    step = a bounded sample of the workload (about 2 s of work per process, less when K is
    large so that the whole run stays within minutes); every worker warms up first."""
    cfg = CONFIGS[name]
    steps = max(1, args.steps)
    if cfg["kind"] == "biped":
        per_core = 200 if steps <= 10 else max(20, 2000 // steps)
    else:
        per_core = max(1, cfg["cpu_sample"] * (8 if steps <= 4 else 2) // 2)
    total_t0 = time.perf_counter()
    seconds, n, kind = 0.0, 0, "reference"
    for _ in range(steps):
        v, n, kind = cpu_baseline(name, per_core, cores)
        seconds += n / v
    value = n * steps / seconds
    what = "the unmodified reference (oracle/_ref)" if kind == "reference" else "the NumPy port (reference tree absent)"
    desc = {"value": value, "unit": "QPs/s", "cores": cores, "kind": kind,
            "sample": "%d instances of the same workload per step over %d processes, 1 BLAS thread "
                      "each: %s through its own API" % (n, cores, what)}
    return {"impl": "reference", "metric": cfg["metric"], "value": value, "unit": "QPs/s",
            "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup, "ms_per_step": 1e3 * n / value,
            "higher_is_better": True, "scaling": "weak" if name == "c2" else "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": cfg["workload"], "name": name}, "cpu_baseline": desc,
            "e2e": {"value": value, "unit": "QPs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "wall_s": time.perf_counter() - total_t0}


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cores = len(os.sched_getaffinity(0))
    name = args.config or ("c2" if max(world, args.gpus) == 1 else "c3")

    if args.impl == "reference":
        if rank == 0:
            emit(reference_arm(args, name, cores))
        return

    env = Env(args)
    if CONFIGS[name]["kind"] == "biped":
        line = run_biped(args, env, name)
    else:
        line = run_lti(args, env, name)
    others = []
    if world == 1 and CONFIGS[name]["kind"] == "biped" and not args.no_lti:
        for other in ("c4", "c5"):
            sub = run_lti(args, env, other, nested=True)
            if not args.no_cpu_baseline:
                per = CONFIGS[other]["cpu_sample"]
                sub["cpu_baseline"] = cpu_baseline_object(other, cores, per, 2 * per)
            others.append(sub)
        line["other_configs"] = others
        line["single_instance_latency"] = {
            "calls": run_single(),
            "note": "one instance per call through generate_all_qp_matrices(given) (the reference's own "
                    "usage, NumPy in / NumPy out), wall clock; oracle_ms_per_call = the CPU port on one core"}
        extras = run_extras()
        if extras:
            line["extras"] = extras
    if env.dist is not None:
        env.dist.destroy_process_group()
    if rank != 0:
        return
    if not args.no_cpu_baseline and world == 1:
        if CONFIGS[name]["kind"] == "biped":
            line["cpu_baseline"] = cpu_baseline_object(name, cores, 96, 384)
        else:
            per = CONFIGS[name]["cpu_sample"]
            line["cpu_baseline"] = cpu_baseline_object(name, cores, per, 2 * per)
    emit(line)


if __name__ == "__main__":
    main()
