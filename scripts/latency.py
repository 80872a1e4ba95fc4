#!/usr/bin/env python3
"""Latency of ONE instance through the reference's own call,
``Formulation.generate_all_qp_matrices(given)`` (body.py:333) -- NumPy in, NumPy out --
for the biped at N = 24 (C1) and N = 100 (C2 structure), next to the oracle on one core.

    python scripts/latency.py [--calls 200]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def measure(N, calls):
    import torch
    from mpc_interface_b200.examples.biped import default_configuration, formulate_biped
    from oracle import qp_oracle

    form = formulate_biped(default_configuration(N))
    given = np.random.default_rng(N).normal(size=[form.given_len, 1]) * 0.05
    for _ in range(5):
        out = form.generate_all_qp_matrices(given)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(calls):
        out = form.generate_all_qp_matrices(given)
    t_gpu = (time.perf_counter() - t0) / calls
    ref = qp_oracle.all_qp_matrices(form, given)
    t0 = time.perf_counter()
    for _ in range(5):
        ref = qp_oracle.all_qp_matrices(form, given)
    t_cpu = (time.perf_counter() - t0) / 5
    err = max(float(np.max(np.abs(np.squeeze(x) - np.squeeze(r))) / max(np.max(np.abs(r)), 1e-300))
              for x, r in zip(out, ref))
    return {"N": N, "nv": form.optim_len, "nc": int(out[0].shape[0]), "ms_per_call": t_gpu * 1e3,
            "oracle_ms_per_call": t_cpu * 1e3, "worst_rel_err": err}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--calls", type=int, default=200)
    args = ap.parse_args()
    for N in (24, 100):
        print(json.dumps(measure(N, args.calls)), flush=True)
