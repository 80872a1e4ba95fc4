"""The CUDA path against the UNMODIFIED reference, live, on the GPU box.

oracle/_ref (staged by oracle/make_ref.py, verified against oracle/REF_MANIFEST.json)
travels with the repo, so here the reference itself -- not the NumPy port, not a stored
vector -- produces the expected matrices on fresh random instances:

  * the reference's own 16 unit tests + its example formulation on the CUDA backend,
  * configs[1] (biped N=100, per-instance omega / x0 / bias / aims): 64 instances,
  * configs[3] (nx=12, nu=4, N=50, one plant per instance): 64 instances, both Hessian modes,
  * configs[4] (nx=6, nu=2, N=400): 4 instances (0.7 s each on the CPU), both Hessian modes.

SURVEY.md 8(c) "Parity metric": max|X - X_ref| / max|X_ref| <= 1e-10 per array
(cases.TOL), shapes and integer layouts exact.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

import _cases as cases
from oracle import make_ref, reference

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not reference.available(), reason="oracle/_ref is not staged")]
HERE = os.path.dirname(os.path.abspath(__file__))


def test_staged_tree_is_the_reference():
    assert make_ref.verify(reference.find()[0])


def test_reference_unittests_on_the_cuda_backend():
    """python/tests/test_*.py of the reference (16 tests) and its biped_formulation.py,
    with ``mpc_interface`` aliased to this package and NOTHING replaced: K1, the preview
    matrices and every QP matrix they look at come from libmpci.so."""
    tests = os.path.join(reference.find()[0], "tests")
    done = subprocess.run(
        [sys.executable, os.path.join(HERE, "_run_reference_suite.py"), "--device", tests],
        capture_output=True, text=True, timeout=900)
    lines = done.stdout.strip().splitlines()
    assert done.returncode == 0 and lines and lines[-1].split()[1:] == ["0", "0"], done.stderr[-3000:]
    assert int(lines[-1].split()[0]) >= 16
    assert "example ok" in done.stdout and "backend libmpci" in done.stdout


def test_c2_sixty_four_fresh_instances_against_the_live_reference():
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import synthetic

    N, B = 100, 64
    form = cases.formulate_biped(cases.default_configuration(N), omega=Param("omega"),
                                 target_vel=[Param("vel_x"), Param("vel_y")])
    seed = int.from_bytes(os.urandom(4), "little")  # fresh every run; printed on failure
    vals = synthetic.biped_inputs(seed, B, N)
    given = synthetic.pack_given(form.given_ID, form.given_len, vals, B)
    out = form.generate_all_qp_matrices_batched(given, {k: vals[k] for k in ("omega", "vel_x", "vel_y")})
    got = [x.cpu().numpy() for x in (out.G, out.h, out.P, out.q)]
    ref_form = reference.biped_form(N)
    assert list(ref_form.optim_ID.items()) == list(form.optim_ID.items()), "integer layout"
    assert list(ref_form.given_ID.items()) == list(form.given_ID.items())
    for b in range(B):
        ref = reference.biped_instance(ref_form, vals, b)
        try:
            cases.assert_qp_close(tuple(x[b] for x in got), ref)
        except AssertionError as e:
            raise AssertionError("seed {} instance {}: {}".format(seed, b, e))


@pytest.mark.parametrize("nx,nu,N,n_rows,count,seed", [
    (12, 4, 50, 4, 64, 1004),   # configs[3]
    (6, 2, 400, 2, 4, 1005),    # configs[4]
])
def test_per_instance_plants_against_the_live_reference(nx, nu, N, n_rows, count, seed):
    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import synthetic

    _, _, Q, R = synthetic.generic_system(seed, nx, nu)
    fresh = int.from_bytes(os.urandom(4), "little")
    A, B = synthetic.generic_plants(fresh, count, nx, nu)
    vals = synthetic.generic_inputs(fresh, count, nx)
    rows = np.random.default_rng(seed + 1).normal(size=[n_rows, nx])
    refs = [Param("ref%d" % i) for i in range(nx)]
    form = cases.formulate_generic(A[0], B[0], Q, R, N, rows, refs=refs)
    params = {"ref%d" % i: vals["ref"][:, i] for i in range(nx)}
    got = {}
    for mode in ("recursion", "dmma"):
        out = form.generate_all_qp_matrices_batched(vals["x0"], params, dynamics={"plant": (A, B)},
                                                    hessian=mode)
        got[mode] = [x.cpu().numpy() for x in (out.G, out.h, out.P, out.q)]
        del out
    plants = reference.GenericPlants(A[0], B[0], Q, R, N, rows)
    assert list(plants.form.optim_ID.items()) == list(form.optim_ID.items())
    for b in range(count):
        ref = plants.instance(A[b], B[b], vals["ref"][b], vals["x0"][b])
        for mode in got:
            try:
                cases.assert_qp_close(tuple(x[b] for x in got[mode]), ref)
            except AssertionError as e:
                raise AssertionError("seed {} instance {} ({}): {}".format(fresh, b, mode, e))


def test_lip_with_per_instance_plants_against_the_live_reference():
    """examples/lip.py, 32 fresh instances: every one its own omega, tau (so its own
    "dP->CCC" plant), initial state and velocity target; the reference formulates each
    instance from scratch with its own classes, the GPU does the batch in one call."""
    import types

    from mpc_interface_b200 import Param
    from mpc_interface_b200.examples import lip

    ref = reference.load()
    api = types.SimpleNamespace(ControlSystem=ref.ControlSystem, ExtendedSystem=ref.ExtendedSystem,
                                Formulation=ref.Formulation, LineCombo=ref.LineCombo, Cost=ref.Cost,
                                Box=ref.Box, make_simetric_vertices=ref.tools.make_simetric_vertices)
    N, B = 50, 32
    seed = int.from_bytes(os.urandom(4), "little")
    vals = lip.lip_inputs(seed, B)
    conf = lip.default_configuration(N)
    form = lip.formulate_lip(conf, omega=Param("omega"), tau=Param("tau"),
                             target_vel=[Param("vel_x"), Param("vel_y")])
    given = np.hstack([vals["x0_x"], vals["x0_y"]])
    got = {}
    for mode in ("recursion", "dmma"):
        out = form.generate_all_qp_matrices_batched(
            given, {k: vals[k] for k in ("omega", "vel_x", "vel_y")},
            dynamics={"LIP": ("dP->CCC", vals["tau"], vals["omega"])}, hessian=mode)
        got[mode] = [x.cpu().numpy() for x in (out.G, out.h, out.P, out.q)]
    for b in range(B):
        one = lip.formulate_lip(conf, omega=float(vals["omega"][b]), tau=float(vals["tau"][b]),
                                target_vel=[float(vals["vel_x"][b]), float(vals["vel_y"][b])], api=api)
        g = one.arrange_given({"x0_x": vals["x0_x"][b][:, None], "x0_y": vals["x0_y"][b][:, None]})
        want = one.generate_all_qp_matrices(g)
        for mode in got:
            try:
                cases.assert_qp_close(tuple(x[b] for x in got[mode]), want)
            except AssertionError as e:
                raise AssertionError("seed {} instance {} ({}): {}".format(seed, b, mode, e))
