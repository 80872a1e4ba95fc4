#!/usr/bin/env python3
"""K1 alone: batched ``mpci_extend_matrices`` (S [B,N,n,n] and U [B,m,N,N,n] in the
reference's layouts, tools.py:14-33) on the sizes of C2 / C4 / C5, CUDA events, against
the measured HBM peak.  Algorithmic bytes per system: 8 (N n n + m N N n) written.

    python scripts/bench_extend.py [--reps 5]
"""
import argparse
import ctypes
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CONFIGS = {"c2 (LIP, n=3, m=1, N=100)": (3, 1, 100, 32768),
           "c4 (n=12, m=4, N=50)": (12, 4, 50, 8192),
           "c5 (n=6, m=2, N=400)": (6, 2, 400, 1184),
           "c1 (LIP, n=3, m=1, N=24)": (3, 1, 24, 262144)}


def run(reps=5):
    """One line (dict) per configuration; needs a CUDA device."""
    import torch
    from mpc_interface_b200 import device

    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    # MEASURED_PEAKS.json (driver-written); the fallback is the B200_PROFILING.md figure
    peak = json.load(open(path))["hbm_gbs"] if os.path.exists(path) else 6650.0
    lib = device.lib()
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    lines = []
    for name, (n, m, N, batch) in CONFIGS.items():
        rng = np.random.default_rng(n * 100 + N)
        A = torch.from_numpy(np.eye(n) + 0.05 * rng.normal(size=[batch, n, n])).cuda()
        B = torch.from_numpy(rng.normal(size=[batch, n, m])).cuda()
        S = torch.empty([batch, N, n, n], dtype=torch.float64, device="cuda")
        U = torch.empty([batch, m, N, N, n], dtype=torch.float64, device="cuda")

        def call():
            device.check(lib.mpci_extend_matrices(A.data_ptr(), B.data_ptr(), S.data_ptr(), U.data_ptr(),
                                                  batch, n, m, N, stream))
        call()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            call()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        nbytes = 8.0 * batch * (N * n * n + m * N * N * n)
        lines.append({"config": name, "batch": batch, "ms": ms, "systems_per_s": batch / ms * 1e3,
                      "bytes_per_system": nbytes / batch, "gbs": nbytes / ms / 1e6,
                      "hbm_peak_gbs": peak, "frac": nbytes / ms / 1e6 / peak})
        del S, U
    return lines


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    for line in run(args.reps):
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
