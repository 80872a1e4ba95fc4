#!/usr/bin/env python3
"""Times the per-instance-dynamics kernel (mode B) on the generic configurations
C4 (nx=12, nu=4, N=50) and C5 (nx=6, nu=2, N=400) of BASELINE.json, in both
Hessian modes, phase by phase (P only / q only / G,h only / all).

    python scripts/bench_lti.py [--config c4|c5|all] [--batch B] [--reps R]

Prints one JSON line per (config, mode) with QPs/s, achieved FP64 TFLOP/s of the
Hessian phase (executed and structural flops) and achieved HBM GB/s.  Batches
are bounded (the full C4/C5 batches are 337 GB): the kernel is persistent, one
CTA per instance, so throughput does not depend on the batch beyond a few waves.
"""
import argparse
import ctypes
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CONFIGS = {"c4": dict(nx=12, nu=4, N=50, rows=4, seed=1004, batch=16384),
           "c5": dict(nx=6, nu=2, N=400, rows=2, seed=1005, batch=1024),
           "small": dict(nx=4, nu=2, N=24, rows=2, seed=1007, batch=65536)}
DMMA_PEAK_TFLOPS = 36.9  # scripts/fp64_peak.cu on this pool's B200 (profiles/r01_fp64_peak.txt)


def run_config(name, batch=0, reps=5, peaks=None):
    """Lines (one per Hessian mode) for one configuration; needs a CUDA device."""
    import torch
    from mpc_interface_b200 import Param, device, lti
    from mpc_interface_b200.examples import synthetic
    from mpc_interface_b200.examples.generic import formulate_generic

    peaks = peaks or {"hbm_gbs": 6650.0}
    c = CONFIGS[name]
    nx, nu, N = c["nx"], c["nu"], c["N"]
    B = batch or c["batch"]
    _, _, Q, R = synthetic.generic_system(c["seed"], nx, nu)
    A, Bm = synthetic.generic_plants(c["seed"], B, nx, nu)
    rows = np.random.default_rng(c["seed"] + 1).normal(size=[c["rows"], nx])
    vals = synthetic.generic_inputs(c["seed"], B, nx)
    refs = [Param("ref%d" % i) for i in range(nx)]
    form = formulate_generic(A[0], Bm[0], Q, R, N, rows, refs=refs)
    plan = form.lti_plan("plant")
    dp = plan.device_plan()
    params = {"ref%d" % i: vals["ref"][:, i] for i in range(nx)}
    dA, dB, dx = (torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in (A, Bm, vals["x0"]))
    scal = torch.from_numpy(plan.slots.evaluate(params, B)).cuda()
    out = device.QPBatch(B, plan.nv, plan.nc)
    lib = device.lib()
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    null = ctypes.c_void_p(0)

    def call(P, q, G, h, mode):
        device.check(lib.mpci_lti_assemble(
            dp.handle, dA.data_ptr(), nx * nx, dB.data_ptr(), nx * nu, dx.data_ptr(),
            scal.data_ptr(), P, q, G, h, B, mode, stream))

    def timed(fn):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    nv, nc = plan.nv, plan.nc
    bytes_qp = 8.0 * (nv * nv + nv + nc * nv + nc)
    structural = 2.0 * (nx * N) * (nu * N) ** 2 + 2.0 * nx * nx * N * (nu * N)
    lines = []
    for mode, label in ((lti.HESSIAN_DMMA, "dmma"), (lti.HESSIAN_RECURSION, "recursion")):
        P, q, G, h = (x.data_ptr() for x in (out.P, out.q, out.G, out.h))
        t_all = timed(lambda: call(P, q, G, h, mode))
        t_P = timed(lambda: call(P, null, null, null, mode))
        t_q = timed(lambda: call(null, q, null, null, mode))
        t_G = timed(lambda: call(null, null, G, h, mode))
        t_0 = timed(lambda: call(null, null, null, null, mode))
        executed = plan.hessian_flops(mode)
        lines.append({
            "config": name, "hessian": label, "batch": B, "n": nx, "m": nu, "N": N, "nv": nv, "nc": nc,
            "qps": B / (t_all * 1e-3), "ms": {"all": t_all, "P": t_P, "q": t_q, "Gh": t_G, "prediction_only": t_0},
            "hessian_tflops_executed": executed * B / (t_P * 1e-3) / 1e12,
            "hessian_tflops_structural": structural * B / (t_P * 1e-3) / 1e12,
            "flops_per_qp": {"executed": executed, "structural": structural},
            "fp64_peak_tflops": DMMA_PEAK_TFLOPS,
            "fp64_frac_executed": executed * B / (t_P * 1e-3) / 1e12 / DMMA_PEAK_TFLOPS,
            "hbm_gbs_all": bytes_qp * B / (t_all * 1e-3) / 1e9,
            "hbm_gbs_Gh": 8.0 * (nc * nv + nc) * B / (t_G * 1e-3) / 1e9,
            "hbm_gbs_P": 8.0 * nv * nv * B / (t_P * 1e-3) / 1e9,
            "hbm_peak_gbs": peaks["hbm_gbs"], "launch": plan.launch_info(mode),
        })
    del out
    return lines


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="all")
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peaks = json.load(open(path)) if os.path.exists(path) else None
    names = ["c4", "c5", "small"] if args.config == "all" else [args.config]
    for name in names:
        for line in run_config(name, args.batch, args.reps, peaks):
            print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
