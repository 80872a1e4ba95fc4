"""Host lowering of per-instance-dynamics formulations (lti.py) against golden
vectors written by the unmodified reference (tests/golden/make_goldens.py
per_instance): the lowered stage tables, evaluated by a plain NumPy statement
of the kernel's formulas, must reproduce the reference's (A, h, Q, q)."""
import numpy as np
import pytest

import _cases as cases
import _lti_numpy
from _numpy_backend import NumpyBackend
from mpc_interface_b200 import Constraint, Cost, LineCombo, Param, lti


def _formulation(g, cpu_extend, per_instance):
    nx = g["Q_w"].shape[0]
    refs = [Param("ref%d" % i) for i in range(nx)]
    A0 = g["A_sys"][0] if per_instance else g["A_sys"]
    B0 = g["B_sys"][0] if per_instance else g["B_sys"]
    return cases.formulate_generic(A0, B0, g["Q_w"], g["R_w"], int(g["N"]), g["rows"], refs=refs,
                                   backend=NumpyBackend())


@pytest.mark.parametrize("name,per_instance", [
    ("generic_perinst_small", True), ("generic_perinst_odd", True), ("generic_perinst_mid", True),
    ("generic_small", False), ("generic_c4_shape", False)])
def test_lowering_matches_reference(cpu_extend, name, per_instance):
    g = cases.golden(name)
    form = _formulation(g, cpu_extend, per_instance)
    cases.check_layout(form, g)
    plan = form.lti_plan("plant")
    nx = g["Q_w"].shape[0]
    assert (plan.nv, plan.nc) == g["A"].shape[:0:-1]
    assert plan.row_offsets == [i * plan.N for i in range(len(plan.row_offsets))]
    names = [repr(e) for e in plan.slots.exprs]
    for b in range(g["x0"].shape[0]):
        A = g["A_sys"][b] if per_instance else g["A_sys"]
        B = g["B_sys"][b] if per_instance else g["B_sys"]
        scalars = np.array([g["ref"][b, int(nm[3:])] for nm in names])
        got = _lti_numpy.evaluate(plan, A, B, g["x0"][b], scalars)
        cases.assert_qp_close(got, (g["A"][b], g["h"][b], g["Q"][b], g["q"][b]))


def test_lowering_refuses_unsupported(cpu_extend):
    g = cases.golden("generic_small")
    form = _formulation(g, cpu_extend, False)
    form.incorporate_goal("sched", Cost("z0", 1.0, schedule=range(2, 4)))
    with pytest.raises(NotImplementedError):
        form.lti_plan("plant")
    form = _formulation(g, cpu_extend, False)
    form.incorporate_definition("mix", LineCombo({"z0": 1.0, "u0": 2.0}))
    form.incorporate_goal("mixed", Cost("mix", 1.0))
    with pytest.raises(NotImplementedError):
        form.lti_plan("plant")
    form = _formulation(g, cpu_extend, False)
    form.incorporate_constraint("L", Constraint("z1", 3.0, L=np.eye(int(g["N"]))))
    with pytest.raises(NotImplementedError):
        form.lti_plan("plant")
    form = _formulation(g, cpu_extend, False)
    with pytest.raises(KeyError):
        form.generate_all_qp_matrices_batched(g["x0"], {}, dynamics={"nope": (g["A_sys"], g["B_sys"])})


def test_lowering_input_forms_and_derived_definitions(cpu_extend):
    """A derived point-wise definition of states in a limit and a cost on an input."""
    g = cases.golden("generic_small")
    form = _formulation(g, cpu_extend, False)
    form.incorporate_definition("sum01", LineCombo({"z0": 1.0, "z1": -2.0}))
    form.incorporate_definition("deep", LineCombo({"sum01": 0.5, "z2": 1.0}))
    form.incorporate_constraint("deep", Constraint("deep", 7.0))
    form.incorporate_goal("u_aim", Cost("u1", 0.3, aim=[Param("ua")]))
    form.make_preview_matrices()
    plan = form.lti_plan("plant")
    from oracle import qp_oracle as orc

    nx = 4
    names = [repr(e) for e in plan.slots.exprs]
    values = {"ref%d" % i: float(g["ref"][0, i]) for i in range(nx)}
    values["ua"] = 0.7
    scalars = np.array([values[nm] for nm in names])
    got = _lti_numpy.evaluate(plan, g["A_sys"], g["B_sys"], g["x0"][0], scalars)
    ref = orc.all_qp_matrices(form, g["x0"][0][:, None], values)
    cases.assert_qp_close(got, ref)
