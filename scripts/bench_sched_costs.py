#!/usr/bin/env python3
"""What a terminal cost costs on the per-instance path: the stage of csrc/mpci_sched.cu
(mpci_lti_sched_add) reads and writes P once more per chunk of scheduled rows.  Two workloads,
each timed without the terminal cost, with it, and the stage's share (difference), plus the
stage's HBM rate over its algorithmic bytes 2 * 8 nv^2 per QP and pass:
  biped   N = 100, per-instance dP->CCC plants, terminal cost on (DCM - s)   (embedded layout)
  generic nx = 12, nu = 4, N = 50 per-instance plants, dense terminal weight (144 crossed costs,
          merged into 12 rows by LtiPlan.sched_layout)                        (plain layout)
One JSON line.   python scripts/bench_sched_costs.py [--batch 16384] [--steps 10]
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def timed(torch, flush, fn, steps, warmup=3):
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
    return float(np.mean(times))


def biped_conf(N, biped):
    conf = biped.default_configuration(N)
    conf.system = "dP->CCC"
    return conf


def run(batch, steps):
    import torch

    from mpc_interface_b200 import Param, device, lti
    from mpc_interface_b200.examples import biped, synthetic
    from mpc_interface_b200.examples.generic import formulate_generic

    dev = torch.device("cuda", 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    line = {"batch": batch, "steps": steps, "flush": "256 MiB fill between steps", "hessian": "recursion"}

    def measure(tag, plans, A, B, given, params):
        res = {}
        out = None
        for key, plan in plans.items():
            job = lti.LtiJob(plan, A, B, given, params)
            if out is None:
                out = device.QPBatch(batch, plan.nv, plan.nc, dev)
            res[key + "_ms"] = timed(torch, flush, lambda: job.launch(out, lti.HESSIAN_RECURSION), steps)
        plan = plans["with"]
        tab = plan.sched_desc_tables()
        res["stage_ms"] = res["with_ms"] - res["without_ms"]
        res["nv"], res["scheduled_rows"] = plan.nv, int(tab["row_ptr"][-1])
        res["stage_hbm_gbs"] = 2 * 8 * plan.nv * plan.nv * batch / max(res["stage_ms"], 1e-9) / 1e6
        res["qps_with"] = batch / res["with_ms"] * 1e3
        line[tag] = res

    N = 100
    vals = synthetic.biped_inputs(2024, batch, N)
    vals["tau"] = np.random.default_rng(7).uniform(0.08, 0.12, batch)
    plans = {}
    for key, weight in (("without", 0.0), ("with", 40.0)):
        conf = biped.default_configuration(N)
        conf.system = "dP->CCC"
        conf.cost_weights["terminal"] = weight
        form = biped.formulate_biped(conf, omega=Param("omega"), tau=Param("tau"),
                                     target_vel=[Param("vel_x"), Param("vel_y")])
        plans[key] = form.lti_plan("LIP")
    given = torch.from_numpy(synthetic.pack_given(form.given_ID, form.given_len, vals, batch)).to(dev)
    A, B = lti.named_system_batched("dP->CCC", vals["tau"], vals["omega"])
    measure("biped_n100_plants", plans, A, B, given, {k: vals[k] for k in ("omega", "vel_x", "vel_y")})

    # one weight per robot on `relax ankles`, a cost on (b + n - s): per-instance cost_w / Pzz of the
    # coupling stage ("with") against the constant weight ("without"); no extra pass over P
    plans = {}
    for key, weight in (("without", 1.0), ("with", Param("w_ankles"))):
        conf = biped.default_configuration(N)
        conf.system = "dP->CCC"
        conf.cost_weights["relax ankles"] = weight
        form = biped.formulate_biped(conf, omega=Param("omega"), tau=Param("tau"),
                                     target_vel=[Param("vel_x"), Param("vel_y")])
        plans[key] = form.lti_plan("LIP")
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    params["w_ankles"] = np.random.default_rng(8).uniform(0.5, 2.0, batch)
    res = {}
    out = None
    for key, plan in plans.items():
        job = lti.LtiJob(plan, A, B, given, params)
        out = device.QPBatch(batch, plan.nv, plan.nc, dev) if out is None else out
        res[key + "_ms"] = timed(torch, flush, lambda: job.launch(out, lti.HESSIAN_RECURSION), steps)
    line["biped_n100_per_instance_ankle_weight"] = res

    # one heading per robot: support polygon and terminal box turned by Param("yaw") -- the turned rows'
    # entries under the footsteps and their given parts are per-instance tables (Gz 63 kB, h map 7 kB per QP)
    form = biped.formulate_biped(biped_conf(N, biped), omega=Param("omega"), tau=Param("tau"), yaw=Param("yaw"),
                                 target_vel=[Param("vel_x"), Param("vel_y")])
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    params["yaw"] = np.random.default_rng(9).uniform(-np.pi, np.pi, batch)
    import time

    t0 = time.perf_counter()
    job = lti.LtiJob(form.lti_plan("LIP"), A, B, given, params)
    torch.cuda.synchronize()
    build_ms = (time.perf_counter() - t0) * 1e3
    line["biped_n100_plants_yaw"] = {
        "ms": timed(torch, flush, lambda: job.launch(out, lti.HESSIAN_RECURSION), steps), "job_build_ms": build_ms,
        "per_instance_table_bytes": int(sum(t.numel() * 8 for t in job.cextra if t is not None) // batch)}

    nx, nu, N = 12, 4, 50
    _, _, Q, R = synthetic.generic_system(1004, nx, nu)
    A, B = synthetic.generic_plants(1004, batch, nx, nu)
    rows = np.random.default_rng(1005).normal(size=[4, nx])
    gv = synthetic.generic_inputs(1004, batch, nx)
    M = np.random.default_rng(3).normal(size=[nx, nx])
    Qf = 3.0 * (M.T @ M) / nx + np.eye(nx)
    refs = [Param("ref%d" % i) for i in range(nx)]
    plans = {key: formulate_generic(A[0], B[0], Q, R, N, rows, refs=refs, terminal=T).lti_plan("plant")
             for key, T in (("without", None), ("with", Qf))}
    params = {"ref%d" % i: gv["ref"][:, i] for i in range(nx)}
    measure("generic_c4_shape", plans, torch.from_numpy(A).to(dev), torch.from_numpy(B).to(dev), gv["x0"], params)
    return line


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=16384)
    ap.add_argument("--steps", type=int, default=10)
    a = ap.parse_args()
    print(json.dumps(run(a.batch, a.steps)))
