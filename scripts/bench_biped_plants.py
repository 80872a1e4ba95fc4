#!/usr/bin/env python3
"""Throughput of the walking biped with ONE PLANT PER INSTANCE (dP->CCC, per-instance omega
and tau): the fused per-instance kernel in the embedded layout + the coupling kernel, outputs
resident in HBM.  One JSON line:  QPs/s, ms per batch, GB/s over all outputs (the algorithmic
bytes 8 (nv^2 + nv + nc nv + nc) per QP), per-output timings, the plan's sizes.
    python scripts/bench_biped_plants.py [--batch 16384] [--horizon 100] [--steps 10]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


PARTS = tuple(os.environ.get("PARTS", "P q Gh").split(" "))  # outputs timed alone ("-" = none: prediction only)


def run(batch=16384, horizon=100, steps=10, warmup=3, system="dP->CCC"):
    import torch

    from mpc_interface_b200 import Param, device, lti
    from mpc_interface_b200.examples import biped, synthetic

    conf = biped.default_configuration(horizon)
    conf.system = system
    form = biped.formulate_biped(conf, omega=Param("omega"), tau=Param("tau"),
                                 target_vel=[Param("vel_x"), Param("vel_y")])
    vals = synthetic.biped_inputs(2024, batch, horizon)
    vals["tau"] = np.random.default_rng(7).uniform(0.08, 0.12, batch)
    dev = torch.device("cuda", 0)
    given = torch.from_numpy(synthetic.pack_given(form.given_ID, form.given_len, vals, batch)).to(dev)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    plan = form.lti_plan("LIP")
    A, B = lti.named_system_batched(system, vals["tau"], vals["omega"])
    out = device.QPBatch(batch, plan.nv, plan.nc, dev)
    bytes_per_qp = 8 * (plan.nv * plan.nv + plan.nv + plan.nc * plan.nv + plan.nc)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    line = {"workload": "biped N={} with per-instance {} plants (omega_b, tau_b), {} instances".format(
                horizon, system, batch),
            "nv": plan.nv, "nc": plan.nc, "nz": plan.nz, "var_off": plan.var_off,
            "bytes_per_qp": bytes_per_qp, "steps": steps, "flush": "256 MiB fill between steps"}
    t0 = time.perf_counter()
    job = lti.LtiJob(plan, A, B, given, params)
    torch.cuda.synchronize()
    line["job_build_ms"] = (time.perf_counter() - t0) * 1e3  # tables evaluated on the GPU, first call

    def timed(fn):
        times = []
        for it in range(warmup + steps):
            flush.fill_(it & 255)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            if it >= warmup:
                times.append(e0.elapsed_time(e1))
        return float(np.mean(times)), float(np.min(times))

    for mode, code in (("recursion", lti.HESSIAN_RECURSION), ("dmma", lti.HESSIAN_DMMA)):
        ms, ms_min = timed(lambda: job.launch(out, code))
        line[mode] = {"ms": ms, "ms_min": ms_min, "qps": batch / ms * 1e3,
                      "hbm_gbs_all": bytes_per_qp * batch / ms / 1e6,
                      "parts_ms": {parts: timed(lambda: job.launch(out, code, parts))[0]
                                   for parts in PARTS},
                      "launch": plan.launch_info(code)}
    line["with_job_build_ms"] = timed(lambda: lti.assemble(plan, A, B, given, params, out, lti.HESSIAN_RECURSION))[0]
    return line


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=16384)
    ap.add_argument("--horizon", type=int, default=100)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--system", default="dP->CCC")
    a = ap.parse_args()
    print(json.dumps(run(a.batch, a.horizon, a.steps, system=a.system)))
