#!/usr/bin/env python3
"""Extras of the bench line (builder-side measurements that ride along in BENCH):

  receding_tick     65,536 biped instances at N=100 spread over the walking phases (the
                    example loop's count_steps sequence, biped_mpc_loop.py:41-45): one warm
                    tick = group by structure (np.unique), per structure gather the given
                    rows on the device and assemble; then the rest of the loop without the
                    solve: preview_all_batched + advance_collector, count_steps
  plan_compile_ms   Formulation.update + plan compile of ONE new structure (cold)
  generic_stream    stream_terms_generic (9-32 monomials) at 10 and 32 G terms
  k1_small_horizon  K1 at N = 24 (short rows)
  biped_per_instance_plants   the biped at N = 100 with one plant per instance (embedded layout)

    python scripts/bench_extras.py
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _events(torch, fn, reps):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def receding_tick(batch=65536, N=100):
    import torch
    from mpc_interface_b200 import Param, receding
    from mpc_interface_b200.examples import synthetic
    from mpc_interface_b200.examples.biped import (default_configuration, formulate_biped,
                                                   structure_key, structure_keys)

    conf = default_configuration(N)

    def factory():
        return formulate_biped(conf, omega=Param("omega"), target_vel=[Param("vel_x"), Param("vel_y")])

    n_steps = factory().domain["Ds_x"]
    rng = np.random.default_rng(7)
    first = rng.integers(0, conf.step_samples, batch)
    times = first[:, None] + conf.step_samples * np.arange(n_steps)[None, :]
    count = rng.integers(0, 40, batch)
    vals = synthetic.biped_inputs(1007, batch, N)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    collector = {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in vals.items() if k not in params}
    buckets = receding.StructureBuckets(factory, structure_key(N))
    outs = {}

    def kargs_of(b):
        return dict(step_times=times[b], step_count=count[b])

    # cold: every structure compiles its plan
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    done = buckets.assemble_device(structure_keys(N, times, count), kargs_of, collector, params=params, outs=outs)
    torch.cuda.synchronize()
    cold = time.perf_counter() - t0
    n_struct = len(done)
    # one plan compile alone: Formulation.update to the structure of instance 0 on a fresh
    # formulation (the set-up of the formulation itself -- SymPy discretisation, qhull -- is
    # not part of a tick and is left out)
    form = factory()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    form.update(**kargs_of(0))
    torch.cuda.synchronize()
    compile_ms = (time.perf_counter() - t0) * 1e3
    # all scalars of the batch evaluated once (the structures share the slot order)
    plan0 = done[0][2]._compiled()
    same_slots = all([e.key() for e in f._compiled().slots.exprs] == [e.key() for e in plan0.slots.exprs]
                     for _, _, f, _ in done)
    scalars = torch.from_numpy(plan0.slots.evaluate(params, batch)).cuda() if same_slots else None

    times_dev, count_dev = torch.from_numpy(times).cuda(), torch.from_numpy(count).cuda()

    def assemble_tick():
        keys = structure_keys(N, times_dev, count_dev)  # key rows, grouping, gathers: on the device
        return buckets.assemble_device(keys, kargs_of, collector, scalars=scalars,
                                       params=None if same_slots else params, outs=outs)

    for _ in range(2):
        assemble_tick()
    torch.cuda.synchronize()
    if os.environ.get("MPCI_PROFILE_TICK"):
        import cProfile
        import pstats

        prof = cProfile.Profile()
        prof.enable()
        assemble_tick()
        torch.cuda.synchronize()
        prof.disable()
        pstats.Stats(prof, stream=sys.stderr).sort_stats("cumulative").print_stats(35)
    reps = 5
    t0 = time.perf_counter()
    for _ in range(reps):
        done = assemble_tick()
    torch.cuda.synchronize()
    warm = (time.perf_counter() - t0) / reps
    # device time of the same tick (events), to separate the host work from the kernels
    dev_ms = _events(torch, assemble_tick, 3)

    # the rest of the loop (no solve: the optimum is stood in for by zeros)
    done = assemble_tick()  # (advance_device works on the inputs of the LAST tick)
    torch.cuda.synchronize()
    optims = [torch.zeros([index.shape[0], form.optim_len], dtype=torch.float64, device="cuda")
              for _, index, form, _ in done]

    def rest():
        moved = buckets.advance_device(done, optims, {"LIP": 0, "steps": 1}, collector, scalars=scalars,
                                       params=None if scalars is not None else params)
        receding.count_steps(times, count, conf.step_samples)
        return moved

    rest()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        rest()
    torch.cuda.synchronize()
    rest_s = (time.perf_counter() - t0) / 3
    sizes = sorted({(f.optim_len, qp.nc) for _, _, f, qp in done})
    del outs, done
    torch.cuda.empty_cache()
    return {"instances": batch, "horizon": N, "structures": n_struct, "qp_sizes_nv_nc": sizes,
            "warm_tick_ms": warm * 1e3, "warm_tick_device_ms": dev_ms,
            "qps_per_s": batch / warm, "ticks_per_s": 1.0 / warm,
            "cold_first_tick_s": cold, "plan_compile_ms": compile_ms,
            "rest_of_loop_ms": rest_s * 1e3,
            "full_tick_qps_per_s": batch / (warm + rest_s),
            "single_structure_step_ms_reference": "ms_per_step of the headline line (same 65,536 instances, one structure)",
            "note": "warm tick = structure keys + grouping on the device (torch.unique over packed key "
                    "rows) + per structure: device gather of the given rows, "
                    "mpci_assemble into a reused slot (all 65,536 instances, 24 structures); "
                    "rest_of_loop = StructureBuckets.advance_device (per structure one launch over the 8 sampled "
                    "state rows on the tick's sorted inputs, one scatter per given variable) + count_steps "
                    "(biped_mpc_loop.py:41-45, 62-65, 81-92), the solve itself left out"}


def generic_stream(batch=8192, N=100):
    """stream_terms_generic: the biped with its three boxes turned by three per-instance
    angles (more than 8 streamed G monomials), and a synthetic 32-term plan."""
    import ctypes

    import torch
    from mpc_interface_b200 import Param, device
    from mpc_interface_b200 import tools as use
    from mpc_interface_b200.examples import synthetic
    from mpc_interface_b200.examples.biped import default_configuration, formulate_biped

    form = formulate_biped(default_configuration(N), omega=Param("omega"),
                           target_vel=[Param("vel_x"), Param("vel_y")], preview=False)
    for name, angle in (("support_polygon", "yaw_a"), ("terminal_Constraint", "yaw_b"), ("stepping area", "yaw_c")):
        form.constraint_boxes[name].rotate_in_TS(use.rotation2D(Param(angle)))
    form.make_preview_matrices()
    plan, ops = form._compiled(), form._ops()
    vals = synthetic.biped_inputs(3, batch, N)
    rng = np.random.default_rng(4)
    params = {k: vals[k] for k in ("omega", "vel_x", "vel_y")}
    params.update({k: rng.uniform(-1, 1, batch) for k in ("yaw_a", "yaw_b", "yaw_c")})
    given = torch.from_numpy(synthetic.pack_given(form.given_ID, form.given_len, vals, batch)).cuda()
    scal = torch.from_numpy(plan.slots.evaluate(params, batch)).cuda()
    out = device.QPBatch(batch, plan.nv, plan.nc)
    lib, handle = device.lib(), ops.device_plan(plan).handle
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    t_G = _events(torch, lambda: device.check(lib.mpci_assemble(
        handle, scal.data_ptr(), given.data_ptr(), None, None, out.G.data_ptr(), None, batch, stream)), 5)
    lines = [{"G_terms": len(plan.G_terms), "batch": batch, "ms": t_G,
              "gbs": 8.0 * plan.nc * plan.nv * batch / t_G / 1e6, "what": "biped N=100, three per-instance yaws"}]
    # a synthetic plan with the maximum of 32 terms: G = sum_t s_t * DG_t
    K, nv, nc = 32, plan.nv, plan.nc
    DG = torch.from_numpy(rng.normal(size=[K, nc, nv])).cuda()
    table = np.array([[t, -1, -1, -1] for t in range(K)], dtype=np.int32)
    terms = np.arange(K, dtype=np.int32)
    segs = np.zeros([0, 4], dtype=np.int32)
    desc = device.PlanDesc(nv, 0, nc, K, K, table.ctypes.data, 1, terms.ctypes.data, 0, K, terms.ctypes.data,
                           DG.data_ptr(), 1, terms.ctypes.data, 0, segs.ctypes.data, 0, 0, 0)
    h32 = ctypes.c_void_p()
    device.check(lib.mpci_plan_create(ctypes.byref(desc), ctypes.byref(h32)))
    s32 = torch.from_numpy(rng.normal(size=[batch, K])).cuda()
    t32 = _events(torch, lambda: device.check(lib.mpci_assemble(
        h32, s32.data_ptr(), None, None, None, out.G.data_ptr(), None, batch, stream)), 5)
    check = float((out.G[5] - torch.einsum("k,kij->ij", s32[5], DG)).abs().max())
    lib.mpci_plan_destroy(h32)
    lines.append({"G_terms": K, "batch": batch, "ms": t32, "gbs": 8.0 * nc * nv * batch / t32 / 1e6,
                  "max_abs_err_vs_einsum": check, "what": "synthetic 32-term plan of the same shape"})
    del out
    torch.cuda.empty_cache()
    return lines


def k1_extend():
    """scripts/bench_extend.py: K1 alone (batched extend_matrices, S and U in the reference's
    layouts) at the sizes of C2 / C4 / C5 / C1."""
    import importlib.util

    spec = importlib.util.spec_from_file_location("bench_extend", os.path.join(ROOT, "scripts", "bench_extend.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.run(3)


def biped_plants(batch=16384, N=100):
    """The walking biped with one dP->CCC plant per instance (per-instance omega and tau): the
    fused per-instance kernel in the embedded layout + the coupling kernel (scripts/
    bench_biped_plants.py); recursion mode is the headline, DMMA mode beside it."""
    import importlib.util

    spec = importlib.util.spec_from_file_location("bench_biped_plants", os.path.join(ROOT, "scripts", "bench_biped_plants.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    line = mod.run(batch=batch, horizon=N, steps=5)
    keep = {k: line[k] for k in ("workload", "nv", "nc", "nz", "bytes_per_qp", "flush", "job_build_ms")}
    for mode in ("recursion", "dmma"):
        keep[mode] = {k: line[mode][k] for k in ("ms", "qps", "hbm_gbs_all", "parts_ms")}
    keep["launches_per_call"] = 2
    return keep


def biped_plants_variants(batch=16384):
    """scripts/bench_sched_costs.py: the same path with the biped's terminal cost switched on (the
    scheduled-cost stage), one `relax ankles` weight per robot, one heading per robot (turned limits),
    and a dense terminal weight on generic plants at C4's shape."""
    import importlib.util

    spec = importlib.util.spec_from_file_location("bench_sched_costs", os.path.join(ROOT, "scripts", "bench_sched_costs.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.run(batch, 5)


def run():
    out = {}
    for name, fn in (("receding_tick", receding_tick), ("generic_stream", generic_stream),
                     ("k1_extend_matrices", k1_extend), ("biped_per_instance_plants", biped_plants),
                     ("per_instance_plants_variants", biped_plants_variants)):
        try:
            out[name] = fn()
        except Exception as e:  # an extra must never cost the headline line
            import traceback

            out[name] = {"error": "%s: %s" % (type(e).__name__, e), "where": traceback.format_exc()[-600:]}
    return out


if __name__ == "__main__":
    print(json.dumps(run(), indent=1))
