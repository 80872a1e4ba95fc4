"""Per-instance dynamics: lowering a Formulation to the condensed LTI plan.

When every instance of a batch has its own plant ``x+ = A_b x + B_b u`` nothing
of the reference's path can be shared: ``extend_matrices`` (tools.py:14-33),
``make_preview_matrices`` (body.py:149-193), ``generate_qp_cost``
(body.py:266-302) and ``generate_qp_constraint`` (body.py:236-264) all depend
on ``(A_b, B_b)``.  The fused kernel behind ``mpci_lti_assemble``
(csrc/mpci_lti.cu) does all of it per instance; this module turns the
formulation objects into the small stage tables that kernel needs
(include/mpci.h: ``mpci_lti_desc``).

Supported structure (anything else raises ``NotImplementedError`` -- there is
no slower fallback):
  * one ``ExtendedSystem`` built by ``from_cotrol_system`` is the plant; its
    inputs are the optimisation variables (in the plant's order), its initial
    state is the only given variable;
  * every other definition is a scalar-coefficient ``LineCombo`` of plant states
    / inputs / such definitions (a point-wise linear form);
  * costs: scalar weight, no ``L`` / ``schedule``, single axis, aims that are
    numbers or ``Param`` expressions, optional ``cross``; a cost may not mix
    states and inputs;
  * limits: no ``L`` / ``schedule``, single axis, ``arrow`` / ``center`` /
    ``extreme`` with one row or one row per horizon sample.

The tables are formulation constants (n x n and smaller); they are accumulated
here in float64 in the order the reference visits costs and limits.
"""
import ctypes
import numbers

import numpy as np

from . import plan as _plan
from .params import is_expr

HESSIAN_DMMA = 0
HESSIAN_RECURSION = 1


class LtiDesc(ctypes.Structure):
    """mpci_lti_desc of include/mpci.h"""

    _fields_ = [
        ("n", ctypes.c_int),
        ("m", ctypes.c_int),
        ("N", ctypes.c_int),
        ("n_scalars", ctypes.c_int),
        ("n_groups", ctypes.c_int),
        ("Qxx", ctypes.c_void_p),
        ("Quu", ctypes.c_void_p),
        ("Tx", ctypes.c_void_p),
        ("Tu", ctypes.c_void_p),
        ("grp_cx", ctypes.c_void_p),
        ("grp_cu", ctypes.c_void_p),
        ("grp_arrow", ctypes.c_void_p),
        ("grp_bound", ctypes.c_void_p),
    ]


class LtiPlan:
    """Host tables of one lowered formulation (see module docstring)."""

    def __init__(self, n, m, N):
        self.n, self.m, self.N = n, m, N
        self.nv, self.ng, self.nc = m * N, n, 0
        self.slots = _plan.SlotTable()
        self.Qxx = np.zeros([n, n])
        self.Quu = np.zeros([m, m])
        self._tx = {}  # slot index (or None for the constant) -> vector [n]
        self._tu = {}
        self.grp_cx, self.grp_cu, self.grp_arrow, self.grp_bound = [], [], [], []
        self.row_offsets, self.limit_rows = [], []
        self.optim_ID, self.given_ID = {}, {}
        self._device = None

    # ---- tables as contiguous arrays --------------------------------------
    def _affine(self, table, rows):
        S = len(self.slots.exprs)
        out = np.zeros([rows, S + 1])
        for slot, vec in table.items():
            out[:, S if slot is None else slot] += vec
        return out

    def tables(self):
        G, n, m, N = len(self.grp_cx), self.n, self.m, self.N
        return dict(
            Qxx=np.ascontiguousarray(self.Qxx),
            Quu=np.ascontiguousarray(self.Quu),
            Tx=self._affine(self._tx, n),
            Tu=self._affine(self._tu, m),
            grp_cx=np.array(self.grp_cx, dtype=np.float64).reshape([G, n]),
            grp_cu=np.array(self.grp_cu, dtype=np.float64).reshape([G, m]),
            grp_arrow=np.array(self.grp_arrow, dtype=np.float64).reshape([G, N]),
            grp_bound=np.array(self.grp_bound, dtype=np.float64).reshape([G, N]),
        )

    def hessian_flops(self, mode=HESSIAN_DMMA):
        from . import device

        return float(device.lib().mpci_lti_hessian_flops(self.device_plan().handle, mode))

    def launch_info(self, mode=HESSIAN_DMMA):
        """{smem_bytes, ctas_per_sm, scan, ld_gz, ld_wz, threads} of the kernel in ``mode``."""
        import ctypes

        from . import device

        info = (ctypes.c_long * 6)()
        device.check(device.lib().mpci_lti_plan_info(self.device_plan().handle, mode, info))
        return dict(zip(("smem_bytes", "ctas_per_sm", "scan", "ld_gz", "ld_wz", "threads"), list(info)))

    def device_plan(self):
        if self._device is None:
            self._device = _DeviceLtiPlan(self)
        return self._device


class _DeviceLtiPlan:
    def __init__(self, plan):
        from . import device

        device.require_gpu()
        t = plan.tables()
        self._keep = t
        desc = LtiDesc(
            plan.n, plan.m, plan.N, len(plan.slots.exprs), len(plan.grp_cx),
            t["Qxx"].ctypes.data, t["Quu"].ctypes.data, t["Tx"].ctypes.data, t["Tu"].ctypes.data,
            t["grp_cx"].ctypes.data, t["grp_cu"].ctypes.data, t["grp_arrow"].ctypes.data,
            t["grp_bound"].ctypes.data,
        )
        handle = ctypes.c_void_p()
        device.check(device.lib().mpci_lti_plan_create(ctypes.byref(desc), ctypes.byref(handle)))
        self.handle = handle
        self._lib = device.lib()

    def __del__(self):
        handle, self.handle = getattr(self, "handle", None), None
        if handle:
            self._lib.mpci_lti_plan_destroy(handle)


# ---------------------------------------------------------------------------
# lowering
# ---------------------------------------------------------------------------
def _unsupported(what):
    raise NotImplementedError(
        "per-instance dynamics: {} is not supported by the condensed LTI path".format(what)
    )


def _linear_forms(form, plant_name):
    """{definition name: (cx [n], cu [m])} for every point-wise linear form."""
    plant = form.dynamics[plant_name]
    axes = plant.axes
    if axes != [""]:
        _unsupported("a plant with several axes")
    n_inputs = len(plant.matrices) - 1
    inputs = [None] * n_inputs
    x0_name = None
    for name, ID in plant.domain_ID.items():
        if ID < n_inputs:
            inputs[ID] = name
        else:
            x0_name = name
    states = [None] * len(plant.state_ID)
    for name, ID in plant.state_ID.items():
        states[ID] = name
    n, m = len(states), n_inputs
    forms = {}
    for i, name in enumerate(states):
        forms[name] = (np.eye(n)[i], np.zeros(m))
    for j, name in enumerate(inputs):
        forms[name] = (np.zeros(n), np.eye(m)[j])
    for name, combo in form.definitions.items():
        if name in forms or name == x0_name:
            continue
        if name in form.of:
            _unsupported("a second dynamics ('{}')".format(form.of[name]))
        cx, cu = np.zeros(n), np.zeros(m)
        for var, coef in combo.items():
            if var not in forms:
                _unsupported("the definition '{}' (depends on '{}')".format(name, var))
            if is_expr(coef) or not isinstance(coef, numbers.Real):
                arr = None if is_expr(coef) else np.asarray(coef)
                if arr is None or arr.ndim != 0:
                    _unsupported("a non-scalar coefficient in the definition '{}'".format(name))
                coef = float(arr)
            cx = cx + float(coef) * forms[var][0]
            cu = cu + float(coef) * forms[var][1]
        forms[name] = (cx, cu)
    return forms, inputs, states, x0_name


def _aim_terms(plan, aim, scale, vec, table):
    """table[slot] += scale * aim * vec  for a single-row aim entry."""
    column = np.asarray(aim, dtype=object).reshape([-1])
    if column.shape[0] != 1:
        raise ValueError(
            "operands could not be broadcast together: an aim must have one row "
            "per axis, got {}".format(column.shape[0])
        )
    value = column[0]
    if is_expr(value):
        slot, number = plan.slots.slot(value), 1.0
    else:
        slot, number = None, float(value)
        if number == 0.0:
            return
    table[slot] = table.get(slot, 0.0) + scale * number * vec


def lower(form, plant_name):
    """Formulation -> LtiPlan (raises NotImplementedError outside the supported
    structure, and the reference's own exceptions for malformed objects)."""
    plant = form.dynamics[plant_name]
    forms, inputs, states, x0_name = _linear_forms(form, plant_name)
    n, m = len(states), len(inputs)
    N = plant.matrices[-1].shape[0]
    if list(form.optim_variables) != inputs:
        _unsupported("an optimisation domain other than the plant inputs in their order")
    if list(form.given_variables) != [x0_name]:
        _unsupported("given variables other than the initial state")
    plan = LtiPlan(n, m, N)
    plan.optim_ID = dict(form.optim_ID)
    plan.given_ID = dict(form.given_ID)

    # ---- limits, in the reference's stacking order (body.py:306-318) -------
    for limit in _plan.ordered_limits(form):
        if limit.L or limit.schedule:
            _unsupported("a limit with L or a schedule")
        if len(limit.axes) != 1 or limit.axes != [""]:
            _unsupported("a limit over several axes")
        if limit.variable not in forms:
            _unsupported("a limit on '{}'".format(limit.variable))
        nlines = limit.nlines
        if nlines is not None and int(nlines) != N:
            raise ValueError(
                "operands could not be broadcast together with shapes ({},1) ({},{})".format(
                    int(nlines), N, plan.nv
                )
            )
        cx, cu = forms[limit.variable]
        arrow = np.asarray(limit.arrow, dtype=np.float64)[:, 0]
        bound = np.asarray(limit.bound(), dtype=np.float64).reshape([-1])
        plan.row_offsets.append(plan.nc)
        plan.limit_rows.append(N)
        plan.nc += N
        plan.grp_cx.append(cx)
        plan.grp_cu.append(cu)
        plan.grp_arrow.append(np.broadcast_to(arrow, [N]) if arrow.shape[0] == 1 else arrow)
        plan.grp_bound.append(np.broadcast_to(bound, [N]) if bound.shape[0] == 1 else bound)

    # ---- costs, in dict order (body.py:322-329) -----------------------------
    for cost in form.goals.values():
        if cost.L or cost.cross_L or cost.schedule:
            _unsupported("a cost with L or a schedule")
        if cost.axes != [""]:
            _unsupported("a cost over several axes")
        if is_expr(cost.weight):
            _unsupported("a per-instance cost weight")
        w = float(cost.weight)
        if cost.variable not in forms or cost.cross not in forms:
            _unsupported("a cost on '{}'".format(cost.variable))
        cv, dv = forms[cost.variable]
        cc, dc = forms[cost.cross]
        if (np.any(cv) and np.any(dc)) or (np.any(dv) and np.any(cc)):
            _unsupported("a cost that couples states and inputs")
        plan.Qxx += w * np.outer(cv, cc)  # Q += w Vo^T Co   (body.py:292)
        plan.Quu += w * np.outer(dv, dc)
        # q += w (Vo^T (.. - caim) + Co^T (.. - aim)) / 2    (body.py:293-300)
        for vec_x, vec_u, aim in ((cv, dv, cost.cross_aim), (cc, dc, cost.aim)):
            _aim_terms(plan, np.asarray(aim, dtype=object)[:, 0], w / 2, vec_x, plan._tx)
            _aim_terms(plan, np.asarray(aim, dtype=object)[:, 0], w / 2, vec_u, plan._tu)
    return plan


# ---------------------------------------------------------------------------
# evaluation
# ---------------------------------------------------------------------------
def assemble(plan, A, B, x0, params=None, out=None, hessian=HESSIAN_DMMA):
    """Runs ``mpci_lti_assemble``.  ``A`` [batch,n,n] or [n,n] (shared), ``B``
    likewise, ``x0`` [batch,n]: numpy arrays or CUDA tensors (float64)."""
    import torch

    from . import device

    dp = plan.device_plan()
    dev = torch.device("cuda", torch.cuda.current_device())

    def to_dev(x):
        if torch.is_tensor(x):
            return x.to(device=dev, dtype=torch.float64).contiguous()
        return torch.from_numpy(np.ascontiguousarray(x, dtype=np.float64)).to(dev)

    A, B, x0 = to_dev(A), to_dev(B), to_dev(x0)
    n, m = plan.n, plan.m
    if x0.dim() != 2 or x0.shape[1] != n:
        raise ValueError(
            "matmul: the given vector has {} elements per instance, the formulation "
            "expects {}".format(x0.shape[-1], n)
        )
    batch = x0.shape[0]
    if tuple(A.shape[-2:]) != (n, n) or tuple(B.shape[-2:]) != (n, m):
        raise ValueError("A must be [..., {0}, {0}] and B [..., {0}, {1}]".format(n, m))
    for M in (A, B):
        if M.dim() == 3 and M.shape[0] != batch:
            raise ValueError("the dynamics hold {} instances, x0 {}".format(M.shape[0], batch))
    strideA = n * n if A.dim() == 3 else 0
    strideB = n * m if B.dim() == 3 else 0
    S = len(plan.slots.exprs)
    scalars = to_dev(plan.slots.evaluate(params or {}, batch)) if S else None
    if out is None:
        out = device.QPBatch(batch, plan.nv, plan.nc, dev)
    elif (out.batch, out.nv, out.nc) != (batch, plan.nv, plan.nc):
        raise ValueError("the output batch does not match the plan")
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    ptr = device._ptr
    device.check(
        device.lib().mpci_lti_assemble(
            dp.handle, ptr(A), strideA, ptr(B), strideB, ptr(x0), ptr(scalars),
            ptr(out.P), ptr(out.q), ptr(out.G), ptr(out.h), batch, int(hessian), stream,
        )
    )
    return out


def assemble_host(plan, A, B, x0, params=None, hessian=HESSIAN_DMMA, chunk=256, out=None):
    """Runs ``mpci_lti_assemble_host``: numpy inputs (``A`` [batch,n,n] or [n,n], ``B``
    likewise, ``x0`` [batch,n]) and HOST results.  ``out`` = (P, q, G, h) pinned CPU
    tensors to fill (allocated when None); returns that tuple."""
    import torch

    from . import device

    dp = plan.device_plan()
    n, m = plan.n, plan.m
    A = np.ascontiguousarray(A, dtype=np.float64)
    B = np.ascontiguousarray(B, dtype=np.float64)
    x0 = np.ascontiguousarray(x0, dtype=np.float64)
    if x0.ndim != 2 or x0.shape[1] != n:
        raise ValueError(
            "matmul: the given vector has {} elements per instance, the formulation "
            "expects {}".format(x0.shape[-1], n)
        )
    batch = x0.shape[0]
    if A.shape[-2:] != (n, n) or B.shape[-2:] != (n, m):
        raise ValueError("A must be [..., {0}, {0}] and B [..., {0}, {1}]".format(n, m))
    for M in (A, B):
        if M.ndim == 3 and M.shape[0] != batch:
            raise ValueError("the dynamics hold {} instances, x0 {}".format(M.shape[0], batch))
    S = len(plan.slots.exprs)
    scalars = plan.slots.evaluate(params or {}, batch) if S else None
    if out is None:
        shapes = ([batch, plan.nv, plan.nv], [batch, plan.nv], [batch, plan.nc, plan.nv], [batch, plan.nc])
        out = tuple(torch.empty(shape, dtype=torch.float64).pin_memory() for shape in shapes)
    hp = lambda a: ctypes.c_void_p(a.ctypes.data) if a is not None else ctypes.c_void_p(0)  # noqa: E731
    device.check(
        device.lib().mpci_lti_assemble_host(
            dp.handle, hp(A), n * n if A.ndim == 3 else 0, hp(B), n * m if B.ndim == 3 else 0,
            hp(x0), hp(scalars), *(ctypes.c_void_p(t.data_ptr()) for t in out),
            batch, int(hessian), int(chunk),
        )
    )
    return out


def assemble_host_stream(plan, A, B, x0, out, ring, params=None, hessian=HESSIAN_RECURSION, chunk=256,
                         on_chunk=None):
    """Runs ``mpci_lti_assemble_host_stream`` -- logical batches larger than host memory.
    ``out`` = (P, q, G, h) pinned CPU tensors holding ``ring`` chunks of ``chunk``
    instances; chunk c lands in slot ``c % ring`` and ``on_chunk(first, count, slot)`` is
    called when it is complete, before the slot is written again."""
    from . import device

    dp = plan.device_plan()
    n, m = plan.n, plan.m
    A = np.ascontiguousarray(A, dtype=np.float64)
    B = np.ascontiguousarray(B, dtype=np.float64)
    x0 = np.ascontiguousarray(x0, dtype=np.float64)
    if x0.ndim != 2 or x0.shape[1] != n:
        raise ValueError(
            "matmul: the given vector has {} elements per instance, the formulation "
            "expects {}".format(x0.shape[-1], n)
        )
    batch = x0.shape[0]
    if A.shape[-2:] != (n, n) or B.shape[-2:] != (n, m):
        raise ValueError("A must be [..., {0}, {0}] and B [..., {0}, {1}]".format(n, m))
    S = len(plan.slots.exprs)
    scalars = plan.slots.evaluate(params or {}, batch) if S else None
    hp = lambda a: ctypes.c_void_p(a.ctypes.data) if a is not None else ctypes.c_void_p(0)  # noqa: E731
    callback = device.ChunkFn(lambda _u, first, count, slot: on_chunk(first, count, slot)) \
        if on_chunk is not None else device.ChunkFn(0)
    device.check(
        device.lib().mpci_lti_assemble_host_stream(
            dp.handle, hp(A), n * n if A.ndim == 3 else 0, hp(B), n * m if B.ndim == 3 else 0,
            hp(x0), hp(scalars), *(ctypes.c_void_p(t.data_ptr()) for t in out),
            batch, int(hessian), int(chunk), int(ring), callback, None,
        )
    )
    return out
