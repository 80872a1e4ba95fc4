"""Per-instance dynamics: lowering a Formulation to the condensed LTI plan.

When every instance of a batch has its own plant ``x+ = A_b x + B_b u`` nothing
of the reference's path can be shared: ``extend_matrices`` (tools.py:14-33),
``make_preview_matrices`` (body.py:149-193), ``generate_qp_cost``
(body.py:266-302) and ``generate_qp_constraint`` (body.py:236-264) all depend
on ``(A_b, B_b)``.  The fused kernel behind ``mpci_lti_assemble``
(csrc/mpci_lti.cu) does all of it per instance; this module turns the
formulation objects into the small stage tables that kernel needs
(include/mpci.h: ``mpci_lti_desc`` / ``mpci_lti_tables``).

Supported structure (anything else raises ``NotImplementedError`` -- there is
no slower fallback):
  * one ``ExtendedSystem`` built by ``from_cotrol_system`` is the plant, over one or
    several axes (the same ``(A_b, B_b)`` on ``_x`` and ``_y``: the kernel sees the
    block-diagonal plant); the plant's inputs are the optimisation variables, in any
    order; the initial states ``<vector>0<axis>`` are the only given variables;
  * every other definition is a scalar-coefficient ``LineCombo`` of plant states /
    inputs / such definitions (a point-wise linear form).  The coefficient may be a
    ``Param`` expression: the LIP's centre of pressure ``b = CoM - CoM_ddot / omega_b^2``
    with one pendulum frequency per instance (biped_formulation.py:62);
  * costs: scalar weight (number or ``Param``), no ``L`` / ``schedule``, aims that are
    numbers or ``Param`` expressions, optional ``cross``, one or several axes; a cost
    may not mix states and inputs;
  * limits: no ``L``; ``schedule`` a contiguous ``range`` (the terminal constraint of the
    biped, biped_formulation.py:97-102); one axis with ``arrow`` / ``center`` / ``extreme``
    of one row or one row per sample, or several axes with a single-row ``arrow`` (a facet
    of a ``Box.task_space``); ``center`` / ``extreme`` may hold ``Param`` expressions.

Without ``Param`` coefficients the tables are formulation constants (per-instance aims
travel as the plan's scalar slots).  With them the tables are evaluated per instance --
vectorised over the batch, n x n and smaller -- and handed to ``mpci_lti_assemble_ex``.
Everything is accumulated in float64 in the order the reference visits costs and limits.
"""
import ctypes
import numbers

import numpy as np

from . import plan as _plan
from .params import is_expr

HESSIAN_DMMA = 0
HESSIAN_RECURSION = 1

NAMED_SYSTEMS = {"J->CCC": (0, 3), "P->CC": (1, 2), "P->X": (2, 1), "dP->CCC": (3, 3), "dP->CCP": (4, 3)}


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
        ("grp_k0", ctypes.c_void_p),
        ("grp_rows", ctypes.c_void_p),
    ]


class LtiTables(ctypes.Structure):
    """mpci_lti_tables of include/mpci.h (device pointers, NULL = the plan's table)"""

    _fields_ = [(name, ctypes.c_void_p) for name in ("Qxx", "Quu", "tx", "tu", "cx", "cu", "bound")]


# ---------------------------------------------------------------------------
# point-wise linear forms with (possibly per-instance) scalar coefficients
# ---------------------------------------------------------------------------
class _Form:
    """``sum_i cx[i] * state_i + sum_j cu[j] * input_j`` as sparse maps index -> coefficient
    (a float or a ``Param`` expression)."""

    def __init__(self, cx=None, cu=None):
        self.cx, self.cu = dict(cx or {}), dict(cu or {})

    def add(self, other, coef):
        for mine, theirs in ((self.cx, other.cx), (self.cu, other.cu)):
            for idx, c in theirs.items():
                term = coef * c
                mine[idx] = mine[idx] + term if idx in mine else term
        return self

    def symbolic(self):
        return any(is_expr(c) for c in list(self.cx.values()) + list(self.cu.values()))

    def touches_states(self):
        return any(is_expr(c) or c != 0.0 for c in self.cx.values())

    def touches_inputs(self):
        return any(is_expr(c) or c != 0.0 for c in self.cu.values())


def _value(coef, params, batch):
    """A coefficient for the whole batch: float, or [B] for a ``Param`` expression."""
    if is_expr(coef):
        return np.broadcast_to(np.asarray(coef.evaluate(params), dtype=np.float64), [batch])
    return float(coef)


def _dense(mapping, size, params, batch):
    """[B, size] (or [1, size] when nothing in it is per-instance)."""
    rows = batch if any(is_expr(c) for c in mapping.values()) else 1
    out = np.zeros([rows, size])
    for idx, c in mapping.items():
        out[:, idx] += _value(c, params, batch) if is_expr(c) else float(c)
    return out


class LtiPlan:
    """Host description of one lowered formulation (see module docstring)."""

    def __init__(self, n, m, N):
        self.n, self.m, self.N = n, m, N
        self.nv, self.ng, self.nc = m * N, n, 0
        self.slots = _plan.SlotTable()
        # costs: (weight, form of the variable, form of its cross, aim, cross aim)
        self.cost_terms = []
        # limits: (form, arrow [N], bound parts [(coef or Expr, [N] array)], k0, rows)
        self.groups = []
        self.row_offsets, self.limit_rows = [], []
        self.optim_ID, self.given_ID = {}, {}
        # block-diagonal expansion of a multi-axis plant: kernel state block -> axis,
        # kernel input k -> (state block, plant input j)
        self.n0, self.m0 = n, m
        self.blocks, self.inputs = [0], [(0, j) for j in range(m)]
        self.per_instance = False  # any Param in weights / forms / bounds
        self._device = None

    # ---- numeric tables -----------------------------------------------------
    def _cost_tables(self, params, batch):
        """Qxx [R, n, n], Quu [R, m, m], and the aim parts; R = batch when per-instance."""
        n, m = self.n, self.m
        Qxx, Quu = np.zeros([1, n, n]), np.zeros([1, m, m])
        for w, fv, fc, _aim, _caim in self.cost_terms:
            wv = np.reshape(_value(w, params, batch), [-1, 1, 1])
            cvx, ccx = _dense(fv.cx, n, params, batch), _dense(fc.cx, n, params, batch)
            cvu, ccu = _dense(fv.cu, m, params, batch), _dense(fc.cu, m, params, batch)
            Qxx = Qxx + wv * cvx[:, :, None] * ccx[:, None, :]  # Q += w Vo^T Co   (body.py:292)
            Quu = Quu + wv * cvu[:, :, None] * ccu[:, None, :]
        return Qxx, Quu

    def _offsets(self, params, batch):
        """tx [R, n], tu [R, m]: (w/2)(cv * caim + cc * aim) summed over the costs
        (body.py:293-300), every factor evaluated per instance."""
        n, m = self.n, self.m
        tx, tu = np.zeros([1, n]), np.zeros([1, m])
        for w, fv, fc, aim, caim in self.cost_terms:
            half = np.reshape(_value(w, params, batch), [-1, 1]) / 2
            for form, target in ((fv, caim), (fc, aim)):
                a = np.reshape(_value(target, params, batch), [-1, 1])
                tx = tx + half * a * _dense(form.cx, n, params, batch)
                tu = tu + half * a * _dense(form.cu, m, params, batch)
        return tx, tu

    def _group_tables(self, params, batch):
        G, n, m, N = len(self.groups), self.n, self.m, self.N
        rows = batch if self.per_instance else 1
        cx, cu = np.zeros([rows, G, n]), np.zeros([rows, G, m])
        bound = np.zeros([rows, G, N])
        for g, (form, _arrow, parts, _k0, _cnt) in enumerate(self.groups):
            cx[:, g] = _dense(form.cx, n, params, batch)
            cu[:, g] = _dense(form.cu, m, params, batch)
            for coef, col in parts:
                bound[:, g] += np.reshape(_value(coef, params, batch), [-1, 1]) * col[None, :]
        return cx, cu, bound

    def tables(self):
        """Constant tables for ``mpci_lti_plan_create``.  A per-instance plan gets
        placeholders whose only job is the structure: ones wherever a form may touch a
        state (the kernel skips the state part of limits whose ``cx`` is all zero)."""
        G, n, m, N, S = len(self.groups), self.n, self.m, self.N, len(self.slots.exprs)
        arrow = np.array([g[1] for g in self.groups], dtype=np.float64).reshape([G, N])
        k0 = np.array([g[3] for g in self.groups], dtype=np.int32).reshape([G])
        cnt = np.array([g[4] for g in self.groups], dtype=np.int32).reshape([G])
        if self.per_instance:
            cx = np.zeros([G, n])
            for g, (form, *_rest) in enumerate(self.groups):
                for idx, c in form.cx.items():
                    cx[g, idx] = 1.0 if (is_expr(c) or c != 0.0) else 0.0
            return dict(Qxx=np.zeros([n, n]), Quu=np.zeros([m, m]), Tx=np.zeros([n, 1]), Tu=np.zeros([m, 1]),
                        grp_cx=cx, grp_cu=np.zeros([G, m]), grp_arrow=arrow, grp_bound=np.zeros([G, N]),
                        grp_k0=k0, grp_rows=cnt)
        Qxx, Quu = self._cost_tables({}, 1)
        cx, cu, bound = self._group_tables({}, 1)
        # aims: affine in the scalar slots (last column = the constant part)
        Tx, Tu = np.zeros([n, S + 1]), np.zeros([m, S + 1])
        for w, fv, fc, aim, caim in self.cost_terms:
            for form, target in ((fv, caim), (fc, aim)):
                column = self.slots.slot(target) if is_expr(target) else S
                scale = float(w) / 2 * (1.0 if is_expr(target) else float(target))
                for idx, c in form.cx.items():
                    Tx[idx, column] += scale * float(c)
                for idx, c in form.cu.items():
                    Tu[idx, column] += scale * float(c)
        return dict(Qxx=np.ascontiguousarray(Qxx[0]), Quu=np.ascontiguousarray(Quu[0]), Tx=Tx, Tu=Tu,
                    grp_cx=np.ascontiguousarray(cx[0]), grp_cu=np.ascontiguousarray(cu[0]),
                    grp_arrow=arrow, grp_bound=np.ascontiguousarray(bound[0]), grp_k0=k0, grp_rows=cnt)

    def instance_tables(self, params, batch):
        """Per-instance tables of a plan with ``Param`` coefficients: ``{name: [B, ...]}``
        for mpci_lti_tables (float64, C order)."""
        for e, what in self.slots.nonnegative:
            v = np.broadcast_to(np.asarray(e.evaluate(params), dtype=np.float64), [batch])
            if not np.all(v >= 0):
                bad = int(np.argmax(~(v >= 0)))
                raise ValueError(
                    "{}: the per-instance extreme '{}' is negative for instance {} ({:g}); the "
                    "reference would flip that limit (restrictions.py:190-194), which needs a "
                    "numeric extreme".format(what, e, bad, float(v[bad])))
        Qxx, Quu = self._cost_tables(params, batch)
        tx, tu = self._offsets(params, batch)
        cx, cu, bound = self._group_tables(params, batch)
        full = lambda a: np.ascontiguousarray(np.broadcast_to(a, (batch,) + a.shape[1:]))  # noqa: E731
        return dict(Qxx=full(Qxx), Quu=full(Quu), tx=full(tx), tu=full(tu), cx=full(cx), cu=full(cu),
                    bound=full(bound))

    def parameter_names(self):
        names = []
        for e in self._exprs():
            for nm in e.names():
                if nm not in names:
                    names.append(nm)
        return names

    def _exprs(self):
        out = list(self.slots.exprs)
        for w, fv, fc, aim, caim in self.cost_terms:
            out += [x for x in (w, aim, caim) if is_expr(x)]
            for form in (fv, fc):
                out += [c for c in list(form.cx.values()) + list(form.cu.values()) if is_expr(c)]
        for form, _arrow, parts, _k0, _cnt in self.groups:
            out += [c for c in list(form.cx.values()) + list(form.cu.values()) if is_expr(c)]
            out += [coef for coef, _ in parts if is_expr(coef)]
        return out

    # ---- the plant as the kernel sees it --------------------------------------
    def expand_numpy(self, A, B):
        """(A_b, B_b) of the plant ([.., n0, n0], [.., n0, m0]) -> the block-diagonal
        system over the axes with the inputs in optimisation order ([.., n, n], [.., n, m])."""
        A, B = np.asarray(A, dtype=np.float64), np.asarray(B, dtype=np.float64)
        n0, n, m = self.n0, self.n, self.m
        if len(self.blocks) == 1 and [j for _, j in self.inputs] == list(range(self.m0)):
            return A, B
        Ae = np.zeros(A.shape[:-2] + (n, n))
        Be = np.zeros(B.shape[:-2] + (n, m))
        for blk in range(len(self.blocks)):
            Ae[..., blk * n0:(blk + 1) * n0, blk * n0:(blk + 1) * n0] = A
        for k, (blk, j) in enumerate(self.inputs):
            Be[..., blk * n0:(blk + 1) * n0, k] = B[..., :, j]
        return Ae, Be

    def expand_device(self, A, B):
        """The same on CUDA tensors (placement of blocks: no arithmetic)."""
        import torch

        n0, n, m = self.n0, self.n, self.m
        if len(self.blocks) == 1 and [j for _, j in self.inputs] == list(range(self.m0)):
            return A, B
        Ae = torch.zeros(tuple(A.shape[:-2]) + (n, n), dtype=torch.float64, device=A.device)
        Be = torch.zeros(tuple(B.shape[:-2]) + (n, m), dtype=torch.float64, device=B.device)
        for blk in range(len(self.blocks)):
            Ae[..., blk * n0:(blk + 1) * n0, blk * n0:(blk + 1) * n0] = A
        for k, (blk, j) in enumerate(self.inputs):
            Be[..., blk * n0:(blk + 1) * n0, k] = B[..., :, j]
        return Ae, Be

    def hessian_flops(self, mode=HESSIAN_DMMA):
        from . import device

        return float(device.lib().mpci_lti_hessian_flops(self.device_plan().handle, mode))

    def launch_info(self, mode=HESSIAN_DMMA):
        """{smem_bytes, ctas_per_sm, scan, ld_gz, ld_wz, threads} of the kernel in ``mode``."""
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
            plan.n, plan.m, plan.N, 0 if plan.per_instance else len(plan.slots.exprs), len(plan.groups),
            t["Qxx"].ctypes.data, t["Quu"].ctypes.data, t["Tx"].ctypes.data, t["Tu"].ctypes.data,
            t["grp_cx"].ctypes.data, t["grp_cu"].ctypes.data, t["grp_arrow"].ctypes.data,
            t["grp_bound"].ctypes.data, t["grp_k0"].ctypes.data, t["grp_rows"].ctypes.data,
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


def _scalar(coef, where):
    """A LineCombo coefficient as a float or a Param expression."""
    if is_expr(coef):
        return coef
    if isinstance(coef, numbers.Real):
        return float(coef)
    arr = np.asarray(coef)
    if arr.dtype == object or arr.ndim != 0:
        _unsupported("a non-scalar coefficient in the definition '{}'".format(where))
    return float(arr)


def _plant_layout(form, plant_name):
    """Names of the plant per axis and their places in the kernel's state / input order."""
    plant = form.dynamics[plant_name]
    axes = list(plant.axes)
    n_inputs = len(plant.matrices) - 1
    base_inputs, base_states, x0_base = [None] * n_inputs, [None] * 0, plant.state_vector_name + "0"
    first = axes[0]
    states = {}
    for name, ident in plant.state_ID.items():
        if name.endswith(first) or first == "":
            states[ident] = name[: len(name) - len(first)] if first else name
    base_states = [states[i] for i in range(len(states))]
    for name, ident in plant.domain_ID.items():
        if ident < n_inputs and (name.endswith(first) or first == ""):
            base_inputs[ident] = name[: len(name) - len(first)] if first else name
    n0, m0 = len(base_states), n_inputs
    # state blocks follow the order of the given variables, inputs the optimisation order
    given = list(form.given_variables)
    if sorted(given) != sorted(x0_base + ax for ax in axes):
        _unsupported("given variables other than the plant's initial states")
    blocks = [axes.index(g[len(x0_base):]) for g in given]  # block -> axis index
    block_of_axis = {axis: blk for blk, axis in enumerate(blocks)}
    wanted = {inp + ax: (block_of_axis[a], j) for a, ax in enumerate(axes) for j, inp in enumerate(base_inputs)}
    optim = list(form.optim_variables)
    if sorted(optim) != sorted(wanted):
        _unsupported("an optimisation domain other than the plant inputs")
    inputs = [wanted[v] for v in optim]  # kernel input k -> (block, plant input j)
    return plant, axes, base_states, base_inputs, x0_base, n0, m0, blocks, inputs, block_of_axis


def _linear_forms(form, plant_name, layout):
    """{definition name: _Form} for every point-wise linear form."""
    plant, axes, base_states, base_inputs, x0_base, n0, m0, blocks, inputs, block_of_axis = layout
    forms = {}
    for a, ax in enumerate(axes):
        for i, name in enumerate(base_states):
            forms[name + ax] = _Form(cx={block_of_axis[a] * n0 + i: 1.0})
    for k, name in enumerate(form.optim_variables):
        forms[name] = _Form(cu={k: 1.0})
    skip = set(x0_base + ax for ax in axes)
    for name, combo in form.definitions.items():
        if name in forms or name in skip:
            continue
        if name in form.of:
            _unsupported("a second dynamics ('{}')".format(form.of[name]))
        out = _Form()
        for var, coef in combo.items():
            if var not in forms:
                _unsupported("the definition '{}' (depends on '{}')".format(name, var))
            out.add(forms[var], _scalar(coef, name))
        forms[name] = out
    return forms


def _single(entry, what):
    column = np.asarray(entry, dtype=object).reshape([-1])
    if column.shape[0] != 1:
        raise ValueError(
            "operands could not be broadcast together: {} must have one row per axis, got {}".format(
                what, column.shape[0]))
    value = column[0]
    return value if is_expr(value) else float(value)


def lower(form, plant_name):
    """Formulation -> LtiPlan (raises NotImplementedError outside the supported
    structure, and the reference's own exceptions for malformed objects)."""
    layout = _plant_layout(form, plant_name)
    plant, axes, base_states, base_inputs, x0_base, n0, m0, blocks, inputs, block_of_axis = layout
    forms = _linear_forms(form, plant_name, layout)
    n, m = n0 * len(axes), len(inputs)
    N = plant.matrices[-1].shape[0]
    plan = LtiPlan(n, m, N)
    plan.n0, plan.m0, plan.blocks, plan.inputs = n0, m0, blocks, inputs
    plan.optim_ID = dict(form.optim_ID)
    plan.given_ID = dict(form.given_ID)

    # ---- limits, in the reference's stacking order (body.py:306-318) -------
    for limit in _plan.ordered_limits(form):
        if limit.L:
            _unsupported("a limit with L")
        k0, cnt = 0, N
        if limit.schedule:
            sched = limit.schedule
            if not isinstance(sched, range) or sched.step != 1 or sched.start < 0 or sched.stop > N:
                _unsupported("a limit whose schedule is not a contiguous range inside the horizon")
            k0, cnt = sched.start, len(sched)
        names = [limit.variable + ax for ax in limit.axes]
        for nm in names:
            if nm not in forms:
                _unsupported("a limit on '{}'".format(nm))
        arrow = np.asarray(limit.arrow)
        if arrow.dtype == object:
            _unsupported("a per-instance arrow (a limit turned by a Param)")
        arrow = arrow.astype(np.float64)
        nlines = limit.nlines
        if nlines is not None and int(nlines) != cnt:
            raise ValueError(
                "operands could not be broadcast together with shapes ({},1) ({},{})".format(
                    int(nlines), cnt, plan.nv))
        rows_form, rows_arrow = _Form(), np.ones([N])
        if len(names) == 1:
            rows_form.add(forms[names[0]], 1.0)
            column = arrow[:, 0]
            rows_arrow = np.zeros([N])
            rows_arrow[k0:k0 + cnt] = np.broadcast_to(column, [cnt]) if column.shape[0] == 1 else column
        else:
            if arrow.shape[0] != 1:
                _unsupported("a limit over several axes whose arrow changes from row to row")
            for a, nm in enumerate(names):  # one facet: the axes' forms weighted by the normal
                rows_form.add(forms[nm], float(arrow[0, a]))
        # bound = extreme + sum_a arrow_a * center_a   (restrictions.py:198-199), rows k0..
        parts = []
        for coef, col in _bound_parts(limit, cnt):
            full = np.zeros([N])
            full[k0:k0 + cnt] = col
            parts.append((coef, full))
        for e in np.asarray(limit.extreme, dtype=object).ravel():
            if is_expr(e):  # the reference would flip such a limit (restrictions.py:190-194)
                plan.slots.require_nonnegative(e, "limit on '{}'".format(limit.variable))
        plan.row_offsets.append(plan.nc)
        plan.limit_rows.append(cnt)
        plan.nc += cnt
        plan.groups.append((rows_form, rows_arrow, parts, k0, cnt))

    # ---- costs, in dict order (body.py:322-329) -----------------------------
    for cost in form.goals.values():
        if cost.L or cost.cross_L or cost.schedule:
            _unsupported("a cost with L or a schedule")
        if isinstance(cost.weight, np.ndarray) and cost.weight.ndim:
            raise TypeError("a cost weight must be a scalar")
        w = cost.weight if is_expr(cost.weight) else float(cost.weight)
        for a, ax in enumerate(cost.axes):
            vname, cname = cost.variable + ax, cost.cross + ax
            if vname not in forms or cname not in forms:
                _unsupported("a cost on '{}'".format(vname))
            fv, fc = forms[vname], forms[cname]
            if (fv.touches_states() and fc.touches_inputs()) or (fv.touches_inputs() and fc.touches_states()):
                _unsupported("a cost that couples states and inputs")
            aim = _single(np.asarray(cost.aim, dtype=object)[:, a], "an aim")
            caim = _single(np.asarray(cost.cross_aim, dtype=object)[:, a], "an aim")
            plan.cost_terms.append((w, fv, fc, aim, caim))

    plan.per_instance = (
        any(is_expr(w) or fv.symbolic() or fc.symbolic() for w, fv, fc, _, _ in plan.cost_terms)
        or any(f.symbolic() or any(is_expr(c) for c, _ in parts) for f, _, parts, _, _ in plan.groups))
    if not plan.per_instance:
        for _, _, _, aim, caim in plan.cost_terms:  # fix the slot order (first use, as before)
            for target in (caim, aim):
                if is_expr(target):
                    plan.slots.slot(target)
    return plan


def _bound_parts(limit, cnt):
    """``limit.bound()`` as ``[(coefficient, [cnt] array)]``: the numeric part, and one
    constant column per ``Param`` monomial (params.split_array)."""
    from .params import split_array

    parts = []
    for atoms, col in split_array(limit.bound()):
        col = np.asarray(col, dtype=np.float64).reshape([-1])
        if col.shape[0] not in (1, cnt):
            raise ValueError("the bound has {} rows for {} lines".format(col.shape[0], cnt))
        coef = 1.0
        for atom in atoms:
            coef = atom if not is_expr(coef) else coef * atom
        parts.append((coef, np.array(np.broadcast_to(col, [cnt]))))
    return parts


# ---------------------------------------------------------------------------
# evaluation
# ---------------------------------------------------------------------------
def named_system_batched(name, tau, omega=None, batch=None):
    """(A [B, n, n], B [B, n, 1]) of a named LIP system (tools.get_system_matrices,
    tools.py:240-285) with one ``tau`` / ``omega`` per instance, built on the GPU from the
    closed forms (``mpci_named_system``).  Scalars are shared by the batch."""
    import torch

    from . import device

    if name not in NAMED_SYSTEMS:
        raise ValueError("unknown system '{}'".format(name))
    ident, n = NAMED_SYSTEMS[name]
    dev = torch.device("cuda", torch.cuda.current_device())

    def column(x):
        if x is None:
            return None, 0
        t = torch.as_tensor(np.asarray(x, dtype=np.float64) if not torch.is_tensor(x) else x,
                            dtype=torch.float64).reshape(-1).to(dev).contiguous()
        return t, (1 if t.numel() > 1 else 0)

    tau_t, st = column(tau)
    om_t, sw = column(omega)
    if ident != 0 and om_t is None:
        raise ValueError("the system '{}' takes the parameters tau and omega".format(name))
    sizes = [t.numel() for t in (tau_t, om_t) if t is not None and t.numel() > 1]
    count = int(batch if batch is not None else (sizes[0] if sizes else 1))
    if any(s != count for s in sizes):
        raise ValueError("tau and omega hold different numbers of instances")
    A = torch.empty([count, n, n], dtype=torch.float64, device=dev)
    B = torch.empty([count, n, 1], dtype=torch.float64, device=dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    ptr = device._ptr
    device.check(device.lib().mpci_named_system(ident, ptr(tau_t), st, ptr(om_t), sw, ptr(A), ptr(B),
                                                count, stream))
    return A, B


def assemble(plan, A, B, x0, params=None, out=None, hessian=HESSIAN_DMMA):
    """Runs ``mpci_lti_assemble`` (``mpci_lti_assemble_ex`` for a plan with per-instance
    tables).  ``A`` [batch, n0, n0] or [n0, n0] (shared), ``B`` likewise -- the plant of ONE
    axis -- and ``x0`` [batch, given_len] in the formulation's given order: numpy arrays or
    CUDA tensors (float64)."""
    import torch

    from . import device

    dp = plan.device_plan()
    dev = torch.device("cuda", torch.cuda.current_device())

    def to_dev(x):
        if torch.is_tensor(x):
            return x.to(device=dev, dtype=torch.float64).contiguous()
        return torch.from_numpy(np.ascontiguousarray(x, dtype=np.float64)).to(dev)

    A, B, x0 = to_dev(A), to_dev(B), to_dev(x0)
    n, m, n0, m0 = plan.n, plan.m, plan.n0, plan.m0
    if x0.dim() != 2 or x0.shape[1] != n:
        raise ValueError(
            "matmul: the given vector has {} elements per instance, the formulation "
            "expects {}".format(x0.shape[-1], n)
        )
    batch = x0.shape[0]
    if tuple(A.shape[-2:]) != (n0, n0) or tuple(B.shape[-2:]) != (n0, m0):
        raise ValueError("A must be [..., {0}, {0}] and B [..., {0}, {1}]".format(n0, m0))
    for M in (A, B):
        if M.dim() == 3 and M.shape[0] != batch:
            raise ValueError("the dynamics hold {} instances, x0 {}".format(M.shape[0], batch))
    A, B = plan.expand_device(A, B)
    A, B = A.contiguous(), B.contiguous()
    strideA = n * n if A.dim() == 3 else 0
    strideB = n * m if B.dim() == 3 else 0
    if out is None:
        out = device.QPBatch(batch, plan.nv, plan.nc, dev)
    elif (out.batch, out.nv, out.nc) != (batch, plan.nv, plan.nc):
        raise ValueError("the output batch does not match the plan")
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    ptr = device._ptr
    if plan.per_instance:
        missing = [nm for nm in plan.parameter_names() if nm not in (params or {})]
        if missing:
            raise KeyError("no value given for the parameter '{}'".format(missing[0]))
        host = plan.instance_tables(params, batch)
        keep = {k: to_dev(v) for k, v in host.items()}
        tables = LtiTables(*(keep[k].data_ptr() for k in ("Qxx", "Quu", "tx", "tu", "cx", "cu", "bound")))
        device.check(
            device.lib().mpci_lti_assemble_ex(
                dp.handle, ptr(A), strideA, ptr(B), strideB, ptr(x0), None, ctypes.byref(tables),
                ptr(out.P), ptr(out.q), ptr(out.G), ptr(out.h), batch, int(hessian), stream,
            )
        )
        out._keep = keep  # the launch is asynchronous: the tables live as long as the result
        return out
    S = len(plan.slots.exprs)
    scalars = to_dev(plan.slots.evaluate(params or {}, batch)) if S else None
    device.check(
        device.lib().mpci_lti_assemble(
            dp.handle, ptr(A), strideA, ptr(B), strideB, ptr(x0), ptr(scalars),
            ptr(out.P), ptr(out.q), ptr(out.G), ptr(out.h), batch, int(hessian), stream,
        )
    )
    return out


def _host_inputs(plan, A, B, x0, params):
    n, m, n0, m0 = plan.n, plan.m, plan.n0, plan.m0
    if plan.per_instance:
        _unsupported("Param coefficients through the host-buffer entry points")
    A = np.asarray(A, dtype=np.float64)
    B = np.asarray(B, dtype=np.float64)
    x0 = np.ascontiguousarray(x0, dtype=np.float64)
    if x0.ndim != 2 or x0.shape[1] != n:
        raise ValueError(
            "matmul: the given vector has {} elements per instance, the formulation "
            "expects {}".format(x0.shape[-1], n)
        )
    batch = x0.shape[0]
    if A.shape[-2:] != (n0, n0) or B.shape[-2:] != (n0, m0):
        raise ValueError("A must be [..., {0}, {0}] and B [..., {0}, {1}]".format(n0, m0))
    for M in (A, B):
        if M.ndim == 3 and M.shape[0] != batch:
            raise ValueError("the dynamics hold {} instances, x0 {}".format(M.shape[0], batch))
    A, B = plan.expand_numpy(A, B)
    A, B = np.ascontiguousarray(A), np.ascontiguousarray(B)
    S = len(plan.slots.exprs)
    scalars = plan.slots.evaluate(params or {}, batch) if S else None
    return A, B, x0, scalars, batch


def assemble_host(plan, A, B, x0, params=None, hessian=HESSIAN_DMMA, chunk=256, out=None):
    """Runs ``mpci_lti_assemble_host``: numpy inputs (``A`` [batch,n,n] or [n,n], ``B``
    likewise, ``x0`` [batch,n]) and HOST results.  ``out`` = (P, q, G, h) pinned CPU
    tensors to fill (allocated when None); returns that tuple."""
    import torch

    from . import device

    dp = plan.device_plan()
    n, m = plan.n, plan.m
    A, B, x0, scalars, batch = _host_inputs(plan, A, B, x0, params)
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
    A, B, x0, scalars, batch = _host_inputs(plan, A, B, x0, params)
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
