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
  * every other definition is a ``LineCombo`` of plant states / inputs / such definitions with
    scalar coefficients (a point-wise linear form), or with numeric matrix coefficients (a rate
    ``(I - shift) @ b``: no stage form, handled row by row by mpci_sched.cu).  The coefficient may be a
    ``Param`` expression: the LIP's centre of pressure ``b = CoM - CoM_ddot / omega_b^2``
    with one pendulum frequency per instance (biped_formulation.py:62);
  * costs: scalar weight (number or ``Param``), aims that are numbers or ``Param`` expressions,
    optional ``cross``, one or several axes; a cost with a ``schedule`` (a terminal cost:
    ``range(N - 1, N)``; any list of samples) or with ``L`` is added to P and q by a second stage
    (csrc/mpci_sched.cu), and so is -- row by row over the whole horizon -- a cost that couples
    states and inputs;
  * limits: ``schedule`` a contiguous ``range`` (the terminal constraint of the
    biped, biped_formulation.py:97-102); one axis with ``arrow`` / ``center`` / ``extreme``
    of one row or one row per sample, or several axes with a single-row ``arrow`` (a facet
    of a ``Box.task_space``); ``center`` / ``extreme`` may hold ``Param`` expressions, and so may
    a single-row ``arrow`` (a box turned per instance); a limit with ``L`` keeps its auxiliary
    and given parts in the coupling stage and gets its plant part from ``mpci_lti_sched_rows``.

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


class LtiCoupledTables(ctypes.Structure):
    """mpci_lti_coupled_tables of include/mpci.h"""

    _fields_ = [(name, ctypes.c_void_p) for name in ("cost_cx", "cost_cu", "cost_w", "cost_aim", "Pzz", "Gz", "h_val")]


class LtiSchedDesc(ctypes.Structure):
    """mpci_lti_sched_desc of include/mpci.h"""

    _fields_ = [
        ("n", ctypes.c_int),
        ("m", ctypes.c_int),
        ("N", ctypes.c_int),
        ("nv", ctypes.c_int),
        ("var_off", ctypes.c_void_p),
        ("nz", ctypes.c_int),
        ("z_col", ctypes.c_void_p),
        ("given_len", ctypes.c_int),
        ("n_terms", ctypes.c_int),
        ("row_ptr", ctypes.c_void_p),
        ("row_k", ctypes.c_void_p),
        ("v_Mz", ctypes.c_void_p),
        ("v_Mg", ctypes.c_void_p),
        ("c_Mz", ctypes.c_void_p),
        ("c_Mg", ctypes.c_void_p),
        ("ent_ptr", ctypes.c_void_p),
        ("ent_k", ctypes.c_void_p),
        ("ent_val", ctypes.c_void_p),
        ("c_ent_ptr", ctypes.c_void_p),
        ("c_ent_k", ctypes.c_void_p),
        ("c_ent_val", ctypes.c_void_p),
        ("row_out", ctypes.c_void_p),
    ]


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


class LtiEmbed(ctypes.Structure):
    """mpci_lti_embed of include/mpci.h (host pointers)"""

    _fields_ = [
        ("nv_total", ctypes.c_int),
        ("nc_total", ctypes.c_int),
        ("var_off", ctypes.c_void_p),
        ("grp_row0", ctypes.c_void_p),
        ("nz", ctypes.c_int),
        ("z_col", ctypes.c_void_p),
        ("given_len", ctypes.c_int),
        ("n_costs", ctypes.c_int),
        ("cost_cx", ctypes.c_void_p),
        ("cost_cu", ctypes.c_void_p),
        ("cost_w", ctypes.c_void_p),
        ("cost_MzT", ctypes.c_void_p),
        ("cost_e_ptr", ctypes.c_void_p),
        ("cost_e_col", ctypes.c_void_p),
        ("cost_e_val", ctypes.c_void_p),
        ("Pzz", ctypes.c_void_p),
        ("Gz", ctypes.c_void_p),
        ("h_ptr", ctypes.c_void_p),
        ("h_col", ctypes.c_void_p),
        ("h_val", ctypes.c_void_p),
        ("row_kind", ctypes.c_void_p),
    ]


# ---------------------------------------------------------------------------
# point-wise linear forms with (possibly per-instance) scalar coefficients
# ---------------------------------------------------------------------------
class _Form:
    """``sum_i cx[i] * state_i + sum_j cu[j] * input_j`` as sparse maps index -> coefficient
    (a float or a ``Param`` expression), plus an auxiliary part ``Mz @ z + Mg @ given`` over
    the optimisation variables that are not plant inputs and the given variables that are not
    the plant's initial state (dense numeric [rows, nz] / [rows, given_len], or None).
    ``rows``: samples of the signal (None: not fixed yet; a plant part means the horizon).
    ``mix``: plant parts that enter through a MATRIX coefficient -- ``M @ (cx . x + cu . u)`` with M
    numeric [rows, N] (a delayed plant signal, a rate ``(I - shift) @ b``): [(M, cx, cu)].  Such a
    signal has no stage form; its costs and limits go row by row through mpci_sched.cu.
    ``sym_aux``: auxiliary parts under a ``Param`` coefficient -- ``coef * (Mz @ z + Mg @ given)``:
    [(coef, Mz or None, Mg or None)]; limits keep them as per-instance Gz / h-map terms, costs take
    the row-by-row stage with the coefficient in the terms' weights."""

    def __init__(self, cx=None, cu=None, rows=None, Mz=None, Mg=None):
        self.cx, self.cu = dict(cx or {}), dict(cu or {})
        self.rows, self.Mz, self.Mg = rows, Mz, Mg
        self.mix, self.sym_aux = [], []

    def add(self, other, coef):
        """self += coef * other, ``coef`` a float or a ``Param`` expression (the latter on a
        form without auxiliary part only)."""
        self._fit_rows(other.rows)
        if is_expr(coef):  # (a per-instance coefficient joins the part's form, the matrix stays numeric)
            self.mix += [(M, {i: coef * c for i, c in cx.items()}, {j: coef * c for j, c in cu.items()})
                         for M, cx, cu in other.mix]
        else:
            self.mix += [(coef * M, cx, cu) for M, cx, cu in other.mix]
        self.sym_aux += [(coef * c2, Mz, Mg) for c2, Mz, Mg in other.sym_aux]
        for mine, theirs in ((self.cx, other.cx), (self.cu, other.cu)):
            for idx, c in theirs.items():
                term = coef * c
                mine[idx] = mine[idx] + term if idx in mine else term
        if is_expr(coef):  # the auxiliary part under a per-instance coefficient
            if any(M is not None and np.any(M) for M in (other.Mz, other.Mg)):
                self.sym_aux.append((coef, other.Mz, other.Mg))
            return self
        for name in ("Mz", "Mg"):
            theirs = getattr(other, name)
            if theirs is not None:
                mine = getattr(self, name)
                setattr(self, name, coef * theirs if mine is None else mine + coef * theirs)
        return self

    def add_matrix(self, other, matrix, where):
        """self += matrix @ other for a numeric [rows, other.rows] coefficient (the shifted
        bias of biped_formulation.py:57,63); a plant part of ``other`` joins ``mix``."""
        if other.rows is not None and matrix.shape[1] != other.rows:
            raise ValueError("matmul: a [{}x{}] coefficient on a variable of {} samples in '{}'".format(
                matrix.shape[0], matrix.shape[1], other.rows, where))
        self._fit_rows(matrix.shape[0])
        if other.cx or other.cu:
            self.mix.append((np.array(matrix, dtype=np.float64), dict(other.cx), dict(other.cu)))
        self.mix += [(matrix @ M, cx, cu) for M, cx, cu in other.mix]
        self.sym_aux += [(c2, None if Mz is None else matrix @ Mz, None if Mg is None else matrix @ Mg)
                         for c2, Mz, Mg in other.sym_aux]
        for name in ("Mz", "Mg"):
            theirs = getattr(other, name)
            if theirs is not None:
                mine = getattr(self, name)
                setattr(self, name, matrix @ theirs if mine is None else mine + matrix @ theirs)
        return self

    def _fit_rows(self, rows):
        if rows is None:
            return
        if self.rows is not None and self.rows != rows:
            raise ValueError("operands could not be broadcast together with shapes ({},) ({},)".format(
                self.rows, rows))
        self.rows = rows

    def has_aux(self):
        return bool(self.sym_aux) or any(M is not None and np.any(M) for M in (self.Mz, self.Mg))

    def has_plant(self):
        return self.touches_states() or self.touches_inputs()

    def _maps(self, which):
        return [getattr(self, which)] + [part[1 if which == "cx" else 2] for part in self.mix]

    def symbolic(self):
        return any(is_expr(c) for which in ("cx", "cu") for mp in self._maps(which) for c in mp.values())

    def touches_states(self):
        return any(is_expr(c) or c != 0.0 for mp in self._maps("cx") for c in mp.values())

    def touches_inputs(self):
        return any(is_expr(c) or c != 0.0 for mp in self._maps("cu") for c in mp.values())

    def parts(self, N):
        """The signal as ``sum_a coef_a * M_a @ (form_a)``: [(form without mix / sym_aux, M_a [rows, N] or
        None = identity, coef_a a Param expression or None = 1)], the first one carrying the direct plant
        part and the numeric auxiliary part."""
        out = [(_Form(self.cx, self.cu, self.rows, self.Mz, self.Mg), None, None)]
        out += [(_Form(cx, cu, N), M, None) for M, cx, cu in self.mix]
        return out + [(_Form(rows=self.rows, Mz=Mz, Mg=Mg), None, coef) for coef, Mz, Mg in self.sym_aux]


class _Host:
    """Array namespace of the table builders on the host (NumPy)."""

    @staticmethod
    def zeros(shape):
        return np.zeros(shape)

    @staticmethod
    def const(array):
        return np.asarray(array, dtype=np.float64)

    @staticmethod
    def full(a, batch):
        return np.ascontiguousarray(np.broadcast_to(a, (batch,) + tuple(a.shape[1:])))

    @staticmethod
    def params(values, batch):
        return {k: np.broadcast_to(np.asarray(v, dtype=np.float64), [batch]) for k, v in (values or {}).items()}


class _Device:
    """The same on the GPU (torch tensors): the per-instance tables of a large batch are built
    where the kernel reads them -- a few dozen elementwise launches instead of NumPy temporaries
    of the batch and an upload from pageable memory."""

    def __init__(self, dev):
        import torch

        self.torch, self.dev = torch, dev

    def zeros(self, shape):
        return self.torch.zeros(list(shape), dtype=self.torch.float64, device=self.dev)

    def const(self, array):
        return self.torch.from_numpy(np.require(array, dtype=np.float64, requirements=["C", "W"])).to(self.dev)

    def full(self, a, batch):
        return a.expand(batch, *a.shape[1:]).contiguous()

    def params(self, values, batch):
        out = {}
        for k, v in (values or {}).items():
            t = v if self.torch.is_tensor(v) else self.torch.from_numpy(
                np.require(np.asarray(v, dtype=np.float64).reshape([-1]), requirements=["C", "W"]))
            out[k] = t.to(device=self.dev, dtype=self.torch.float64).reshape(-1).expand(batch)
        return out


def _value(coef, params):
    """A coefficient for the whole batch: float, or [B] (array / tensor) for a ``Param`` expression."""
    return coef.evaluate(params) if is_expr(coef) else float(coef)


def _col(v, trailing):
    """A per-instance value shaped to broadcast against [B, ...] with ``trailing`` more axes."""
    return v.reshape([-1] + [1] * trailing) if hasattr(v, "reshape") else v


def _dense(mapping, size, params, batch, xp=_Host):
    """[B, size] (or [1, size] when nothing in it is per-instance)."""
    rows = batch if any(is_expr(c) for c in mapping.values()) else 1
    out = xp.zeros([rows, size])
    for idx, c in mapping.items():
        out[:, idx] += _value(c, params)
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
        self.varying = set()       # ... and which of the kernel's tables they reach
        self._device = None
        # embedded layout: the QP has variables / givens / rows outside the plant (lower())
        self.embedded = False
        self.nz, self.z_col, self.var_off, self.x0_index = 0, [], None, None
        self.group_row0 = []  # first row of every plant group in G / h
        self.coupled = []     # costs on signals with an auxiliary part: (w, form, Mz [N,nz], E [N,ng+1])
        self.Pzz = self.Gz = self.Hg = self.row_kind = None
        # limits turned per instance on signals with auxiliary parts: (arrow expression, first row,
        # matrix) to be added to Gz [nc, nz] / to Hg [nc, ng + 1] per instance
        self.Gz_terms, self.Hg_terms = [], []
        self.coupled_aims = []  # per coupled cost: its aim when that is a Param expression, else None
        self.Pzz_terms = []     # (Param weight, Mz' Mz) of the coupled costs with a per-instance weight
        # costs with a schedule (the stage of csrc/mpci_sched.cu, added to P and q afterwards):
        # (weight, form, cross form, aim, cross aim, samples [list of k])
        self.sched_terms = []
        self.sched_L = {}   # index into sched_terms -> (L, cross L): the cost combines its samples (goal.py:88-104)
        # limits with L: the plant part of their rows, added by mpci_lti_sched_rows --
        # (form without auxiliary part, arrow * L [lines, samples], samples, first row in G / h,
        #  the arrow's Param expression when the limit is turned per instance -- then L alone -- or None)
        self.lrows = []
        self._sched = self._lrows = None

    # ---- numeric tables -----------------------------------------------------
    def _cost_tables(self, params, batch, xp=_Host):
        """Qxx [R, n, n], Quu [R, m, m], and the aim parts; R = batch when per-instance."""
        n, m = self.n, self.m
        Qxx, Quu = xp.zeros([1, n, n]), xp.zeros([1, m, m])
        for w, fv, fc, _aim, _caim in self.cost_terms:
            wv = _col(_value(w, params), 2)
            cvx, ccx = _dense(fv.cx, n, params, batch, xp), _dense(fc.cx, n, params, batch, xp)
            cvu, ccu = _dense(fv.cu, m, params, batch, xp), _dense(fc.cu, m, params, batch, xp)
            Qxx = Qxx + wv * cvx[:, :, None] * ccx[:, None, :]  # Q += w Vo^T Co   (body.py:292)
            Quu = Quu + wv * cvu[:, :, None] * ccu[:, None, :]
        return Qxx, Quu

    def _offsets(self, params, batch, xp=_Host):
        """tx [R, n], tu [R, m]: (w/2)(cv * caim + cc * aim) summed over the costs
        (body.py:293-300), every factor evaluated per instance."""
        n, m = self.n, self.m
        tx, tu = xp.zeros([1, n]), xp.zeros([1, m])
        for w, fv, fc, aim, caim in self.cost_terms:
            half = _col(_value(w, params), 1) / 2
            for form, target in ((fv, caim), (fc, aim)):
                a = _col(_value(target, params), 1)
                tx = tx + half * a * _dense(form.cx, n, params, batch, xp)
                tu = tu + half * a * _dense(form.cu, m, params, batch, xp)
        return tx, tu

    def _group_tables(self, params, batch, xp=_Host, which=("cx", "cu", "bound")):
        G, n, m, N = len(self.groups), self.n, self.m, self.N
        rows = {k: batch if (self.per_instance and k in self.varying) else 1 for k in ("cx", "cu", "bound")}
        cx = xp.zeros([rows["cx"], G, n]) if "cx" in which else None
        cu = xp.zeros([rows["cu"], G, m]) if "cu" in which else None
        bound = xp.zeros([rows["bound"], G, N]) if "bound" in which else None
        for g, (form, _arrow, parts, _k0, _cnt) in enumerate(self.groups):
            if cx is not None:
                cx[:, g] = _dense(form.cx, n, params, batch, xp)
            if cu is not None:
                cu[:, g] = _dense(form.cu, m, params, batch, xp)
            for coef, col in (parts if bound is not None else ()):
                bound[:, g] += _col(_value(coef, params), 1) * xp.const(col)[None, :]
        return cx, cu, bound

    def _find_varying(self):
        """Which tables differ between the instances of a per-instance plan: only those travel
        per instance (mpci_lti_tables), the others stay formulation constants of the plan."""
        v = set()
        for w, fv, fc, _aim, _caim in self.cost_terms:
            sym_x = any(is_expr(c) for f in (fv, fc) for c in f.cx.values())
            sym_u = any(is_expr(c) for f in (fv, fc) for c in f.cu.values())
            if (is_expr(w) or sym_x) and (fv.cx and fc.cx):
                v.add("Qxx")
            if (is_expr(w) or sym_u) and (fv.cu and fc.cu):
                v.add("Quu")
        for form, _arrow, parts, _k0, _cnt in self.groups:
            if any(is_expr(c) for c in form.cx.values()):
                v.add("cx")
            if any(is_expr(c) for c in form.cu.values()):
                v.add("cu")
            if any(is_expr(coef) for coef, _ in parts):
                v.add("bound")
        return v

    def tables(self):
        """Constant tables for ``mpci_lti_plan_create``.  In a per-instance plan the tables
        that vary get placeholders whose only job is the structure: ones wherever a form may
        touch a state (the kernel skips the state part of limits whose ``cx`` is all zero)."""
        G, n, m, N, S = len(self.groups), self.n, self.m, self.N, len(self.slots.exprs)
        arrow = np.array([g[1] for g in self.groups], dtype=np.float64).reshape([G, N])
        k0 = np.array([g[3] for g in self.groups], dtype=np.int32).reshape([G])
        cnt = np.array([g[4] for g in self.groups], dtype=np.int32).reshape([G])
        if self.per_instance:
            fixed = [k for k in ("cx", "cu", "bound") if k not in self.varying]
            cx, cu, bound = self._group_tables({}, 1, which=fixed)
            if cx is None:
                cx = np.zeros([1, G, n])
                for g, (form, *_rest) in enumerate(self.groups):
                    for idx, c in form.cx.items():
                        cx[0, g, idx] = 1.0 if (is_expr(c) or c != 0.0) else 0.0
            cu = np.zeros([1, G, m]) if cu is None else cu
            bound = np.zeros([1, G, N]) if bound is None else bound
            Qxx, Quu = np.zeros([1, n, n]), np.zeros([1, m, m])
            if "Qxx" not in self.varying or "Quu" not in self.varying:
                # (a table that does not vary has no Param in it: any values do for the others)
                Qc, Rc = self._cost_tables({k: 1.0 for k in self.parameter_names()}, 1)
                Qxx = Qc if "Qxx" not in self.varying else Qxx
                Quu = Rc if "Quu" not in self.varying else Quu
            c = np.ascontiguousarray
            return dict(Qxx=c(Qxx[0]), Quu=c(Quu[0]), Tx=np.zeros([n, 1]), Tu=np.zeros([m, 1]),
                        grp_cx=c(cx[0]), grp_cu=c(cu[0]), grp_arrow=arrow, grp_bound=c(bound[0]),
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

    def embed_tables(self):
        """The constant tables of mpci_lti_embed (include/mpci.h) for an embedded plan."""
        n, m, N, nz, nC = self.n, self.m, self.N, self.nz, len(self.coupled)
        i32 = lambda x: np.ascontiguousarray(x, dtype=np.int32)  # noqa: E731
        f64 = lambda x: np.ascontiguousarray(x, dtype=np.float64)  # noqa: E731

        def csr(dense):  # rows of a dense [R, ng + 1] map, exact zeros dropped
            ptr, col, val = [0], [], []
            for row in dense:
                nzc = np.nonzero(row)[0]
                col += list(nzc)
                val += list(row[nzc])
                ptr.append(len(col))
            return i32(ptr), i32(col), f64(val)

        cx, cu = np.zeros([nC, n]), np.zeros([nC, m])
        if not self.per_instance:
            for c, (_w, fv, _Mz, _E) in enumerate(self.coupled):
                cx[c], cu[c] = _dense(fv.cx, n, {}, 1)[0], _dense(fv.cu, m, {}, 1)[0]
        e_ptr, e_col, e_val = csr(np.vstack([E for *_r, E in self.coupled]) if nC else np.zeros([0, self.ng + 1]))
        mask = self.h_mask()  # (the entries a turned limit may fill are part of the pattern)
        h_ptr = i32(np.concatenate([[0], np.cumsum(mask.sum(axis=1))]))
        h_col, h_val = i32(np.nonzero(mask)[1]), f64(self.Hg[mask])
        MzT = np.stack([Mz.T for _w, _f, Mz, _E in self.coupled]) if nC else np.zeros([0, nz, N])
        return dict(var_off=i32(self.var_off), grp_row0=i32(self.group_row0), z_col=i32(self.z_col),
                    cost_cx=f64(cx), cost_cu=f64(cu), cost_w=f64([0.0 if is_expr(w) else w for w, *_r in self.coupled]),
                    cost_MzT=f64(MzT), cost_e_ptr=e_ptr, cost_e_col=e_col, cost_e_val=e_val,
                    Pzz=f64(self.Pzz), Gz=f64(self.Gz), h_ptr=h_ptr, h_col=h_col, h_val=h_val,
                    row_kind=i32(self.row_kind))

    def coupled_tables(self, params, batch, dev=None):
        """Per-instance signal tables of the coupled costs: cost_cx [B, nC, n], cost_cu [B, nC, m]
        (NumPy, or CUDA tensors when ``dev`` is given)."""
        xp = _Host if dev is None else _Device(dev)
        params = xp.params(params, batch)
        nC = len(self.coupled)
        cx, cu = xp.zeros([batch, nC, self.n]), xp.zeros([batch, nC, self.m])
        for c, (_w, fv, _Mz, _E) in enumerate(self.coupled):
            cx[:, c] = _dense(fv.cx, self.n, params, batch, xp)
            cu[:, c] = _dense(fv.cu, self.m, params, batch, xp)
        return cx, cu

    def h_mask(self):
        """[nc, ng + 1] pattern of the h map: the non-zeros of Hg and of every per-instance term."""
        mask = self.Hg != 0.0
        for _expr, row0, M in self.Hg_terms:
            mask[row0:row0 + len(M)] |= M != 0.0
        return mask

    def coupled_symbolic(self):
        """Whether anything about the coupling stage's tables differs between the instances."""
        return bool(self.Pzz_terms or self.Gz_terms or self.Hg_terms) or any(
            a is not None for a in self.coupled_aims) or any(
            is_expr(w) or fv.symbolic() for w, fv, _Mz, _E in self.coupled)

    def coupled_extras(self, params, batch, dev=None):
        """(cost_w [B, nC], cost_aim [B, nC], Pzz [B, nz, nz], Gz [B, nc, nz], h_val [B, nnz]) of
        mpci_lti_coupled_tables: per-instance weights / aims of the coupled costs, and the auxiliary
        parts of limits turned per instance -- each None when nothing in it varies."""
        xp = _Host if dev is None else _Device(dev)
        params = xp.params(params, batch)
        nC, nz = len(self.coupled), self.nz
        w = aim = Pzz = None
        if any(is_expr(wc) for wc, *_r in self.coupled):
            w = xp.zeros([batch, nC])
            for c, (wc, *_r) in enumerate(self.coupled):
                w[:, c] = _value(wc, params)
            Pzz = xp.full(xp.const(self.Pzz)[None], batch)
            for wc, MM in self.Pzz_terms:
                Pzz = Pzz + _col(_value(wc, params), 2) * xp.const(MM)[None]
        if any(a is not None for a in self.coupled_aims):
            aim = xp.zeros([batch, nC])
            for c, a in enumerate(self.coupled_aims):
                if a is not None:
                    aim[:, c] = _value(a, params)
        Gz = hval = None
        if self.Gz_terms:
            Gz = xp.full(xp.const(self.Gz)[None], batch)
            for expr, row0, M in self.Gz_terms:
                Gz[:, row0:row0 + len(M)] += _col(_value(expr, params), 2) * xp.const(M)[None]
        if self.Hg_terms:
            mask = self.h_mask()
            hval = xp.full(xp.const(self.Hg[mask])[None], batch)
            for expr, row0, M in self.Hg_terms:
                full = np.zeros_like(self.Hg)
                full[row0:row0 + len(M)] = M
                hval = hval + _col(_value(expr, params), 1) * xp.const(full[mask])[None]
        return w, aim, Pzz, Gz, hval

    def instance_tables(self, params, batch, dev=None):
        """Per-instance tables of a plan with ``Param`` coefficients: ``{name: [B, ...]}`` for
        mpci_lti_tables (float64, C order) -- ``tx`` / ``tu`` and whichever of ``Qxx``, ``Quu``,
        ``cx``, ``cu``, ``bound`` vary (``self.varying``).  NumPy arrays, or CUDA tensors built
        on the device when ``dev`` is given."""
        xp = _Host if dev is None else _Device(dev)
        params = xp.params(params, batch)
        for e, what in self.slots.nonnegative:
            v = e.evaluate(params)
            if not bool((v >= 0).all()):
                v = np.asarray(v.cpu() if hasattr(v, "cpu") else v)
                bad = int(np.argmax(~(v >= 0)))
                raise ValueError(
                    "{}: the per-instance extreme '{}' is negative for instance {} ({:g}); the "
                    "reference would flip that limit (restrictions.py:190-194), which needs a "
                    "numeric extreme".format(what, e, bad, float(v[bad])))
        out = {}
        if "Qxx" in self.varying or "Quu" in self.varying:
            Qxx, Quu = self._cost_tables(params, batch, xp)
            if "Qxx" in self.varying:
                out["Qxx"] = Qxx
            if "Quu" in self.varying:
                out["Quu"] = Quu
        out["tx"], out["tu"] = self._offsets(params, batch, xp)
        moving = [k for k in ("cx", "cu", "bound") if k in self.varying]
        if moving:
            for k, t in zip(("cx", "cu", "bound"), self._group_tables(params, batch, xp, which=moving)):
                if t is not None:
                    out[k] = t
        return {k: xp.full(a, batch) for k, a in out.items()}

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
        for w, fv, fc, aim, caim, _rows in self.sched_terms:
            out += [x for x in (w, aim, caim) if is_expr(x)]
            for form in (fv, fc):
                out += [c for c in list(form.cx.values()) + list(form.cu.values()) if is_expr(c)]
        out += [expr for expr, _r, _M in self.Gz_terms + self.Hg_terms]
        for form, *_r, coef in self.lrows:
            out += [c for c in list(form.cx.values()) + list(form.cu.values()) if is_expr(c)]
            out += [coef] if coef is not None else []
        for (w, fv, _Mz, _E), aim in zip(self.coupled, self.coupled_aims):
            out += [x for x in (w, aim) if is_expr(x)]
            out += [c for c in list(fv.cx.values()) + list(fv.cu.values()) if is_expr(c)]
        return out

    # ---- costs with a schedule ---------------------------------------------------
    def sched_layout(self):
        """The stage's terms: one per scheduled cost and axis, except that costs on plant
        STATES alone which share their samples -- the nx^2 crossed costs of a dense terminal weight
        (x_N - r)' Qf (x_N - r) -- are merged into n terms per group: with Qk = sum_t w_t cx_t cc_t^T
        and tx = sum_t w_t / 2 (cx_t caim_t + cc_t aim_t), the group adds T_k^T Qk T_k to P and
        T_k^T ((Qk + Qk^T) / 2 x_k - tx) to q, which are n outer products per sample (term a: the
        variable's form e_a, the cross form Qk[a, :], weight 1, aims (0, 2 tx[a])) instead of one
        per cost.  Entries: ("one", index into sched_terms) / ("merged", [indices], a)."""
        groups, layout = {}, []
        for i, (_w, fv, fc, _aim, _caim, samples) in enumerate(self.sched_terms):
            states_only = not (fv.touches_inputs() or fc.touches_inputs() or fv.has_aux() or fc.has_aux()
                               or i in self.sched_L)
            if states_only:
                groups.setdefault(tuple(samples), []).append(i)
            else:
                layout.append(("one", i))
        for members in groups.values():
            if len(members) > self.n:
                layout += [("merged", members, a) for a in range(self.n)]
            else:
                layout += [("one", i) for i in members]
        return layout

    def sched_tables(self, params, batch, dev=None):
        """terms [R, T, 3 + 2n + 2m] of mpci_lti_sched_add -- per term of ``sched_layout`` (w, aim,
        caim, v_cx, v_cu, c_cx, c_cu); R = batch when anything in them is per-instance, else 1."""
        xp = _Host if dev is None else _Device(dev)
        params = xp.params(params, batch)
        n, m = self.n, self.m
        layout = self.sched_layout()
        symbolic = any(is_expr(x) for x in self._sched_exprs())
        out = xp.zeros([batch if symbolic else 1, len(layout), 3 + 2 * n + 2 * m])
        merged = {}
        for t, entry in enumerate(layout):
            if entry[0] == "one":
                w, fv, fc, aim, caim, _rows = self.sched_terms[entry[1]]
                for col, x in enumerate((w, aim, caim)):
                    out[:, t, col] = _value(x, params)
                out[:, t, 3:3 + n] = _dense(fv.cx, n, params, batch, xp)
                out[:, t, 3 + n:3 + n + m] = _dense(fv.cu, m, params, batch, xp)
                out[:, t, 3 + n + m:3 + 2 * n + m] = _dense(fc.cx, n, params, batch, xp)
                out[:, t, 3 + 2 * n + m:] = _dense(fc.cu, m, params, batch, xp)
                continue
            _kind, members, a = entry
            key = tuple(members)
            if key not in merged:
                Qk, tx = xp.zeros([1, n, n]), xp.zeros([1, n])
                for i in members:
                    w, fv, fc, aim, caim, _rows = self.sched_terms[i]
                    wv = _value(w, params)
                    cv, cc = _dense(fv.cx, n, params, batch, xp), _dense(fc.cx, n, params, batch, xp)
                    Qk = Qk + _col(wv, 2) * cv[:, :, None] * cc[:, None, :]
                    tx = tx + _col(wv, 1) / 2 * (_col(_value(caim, params), 1) * cv + _col(_value(aim, params), 1) * cc)
                merged[key] = (Qk, tx)
            Qk, tx = merged[key]
            out[:, t, 0] = 1.0
            out[:, t, 2] = 2.0 * tx[:, a]
            out[:, t, 3 + a] = 1.0
            out[:, t, 3 + n + m:3 + 2 * n + m] = Qk[:, a, :]
        return out

    def _sched_exprs(self):
        out = []
        for w, fv, fc, aim, caim, _rows in self.sched_terms:
            out += [w, aim, caim]
            for form in (fv, fc):
                out += list(form.cx.values()) + list(form.cu.values())
        return out

    def sched_desc_tables(self):
        """The constant tables of mpci_lti_sched_desc (include/mpci.h), terms as ``sched_layout``."""
        i32 = lambda x: np.ascontiguousarray(x, dtype=np.int32)  # noqa: E731
        f64 = lambda x: np.ascontiguousarray(x, dtype=np.float64)  # noqa: E731
        nz, ng = self.nz, (self.ng if self.embedded else 0)
        layout = self.sched_layout()
        combos = []  # per term: (samples, L or None, cross L or None)
        for entry in layout:
            if entry[0] == "one":
                Lv, Lc = self.sched_L.get(entry[1], (None, None))
                combos.append((list(self.sched_terms[entry[1]][5]), Lv, Lc))
            else:
                combos.append((list(self.sched_terms[entry[1][0]][5]), None, None))
        ptr = [0]
        for samples, Lv, _Lc in combos:
            ptr.append(ptr[-1] + (len(samples) if Lv is None else Lv.shape[0]))

        def stack(which, attr, width):
            singles = [self.sched_terms[e[1]] for e in layout if e[0] == "one"]
            if not width or not any(getattr(term[which], attr) is not None for term in singles):
                return None
            blocks = []
            for t, entry in enumerate(layout):
                term = self.sched_terms[entry[1]] if entry[0] == "one" else None
                M = getattr(term[which], attr) if term is not None else None
                if M is None:
                    blocks.append(np.zeros([ptr[t + 1] - ptr[t], width]))
                    continue
                picked = np.asarray(M)[combos[t][0]]
                L = combos[t][which]
                blocks.append(picked if L is None else L @ picked)  # vMo = L @ Mo[sigma]  (body.py:279-290)
            return f64(np.vstack(blocks))

        crossed = any(e[0] == "one" and self.sched_terms[e[1]][1] is not self.sched_terms[e[1]][2] for e in layout)
        t = dict(row_ptr=i32(ptr), row_k=None, v_Mz=stack(1, "Mz", nz), v_Mg=stack(1, "Mg", ng),
                 c_Mz=stack(2, "Mz", nz) if crossed else None, c_Mg=stack(2, "Mg", ng) if crossed else None,
                 var_off=i32(self.var_off) if self.embedded else None,
                 z_col=i32(self.z_col) if nz else None, row_out=None)
        t.update(_combination_tables(combos))
        for own, cross in (("v_Mz", "c_Mz"), ("v_Mg", "c_Mg")):
            if t[cross] is not None and t[own] is None:  # (the C side wants the own table beside a crossed one)
                t[own] = np.zeros_like(t[cross])
            if crossed and t[own] is not None and t[cross] is None:  # (NULL would mean "the variable's own")
                t[cross] = np.zeros_like(t[own])
        return t

    # ---- limits with L ------------------------------------------------------------
    def lrow_desc_tables(self):
        """mpci_lti_sched_desc of the limits with ``L``: one term per (limit, axis with a plant part),
        its rows the lines of ``arrow * L`` (restrictions.py:175-179), each added to its row of G / h."""
        i32 = lambda x: np.ascontiguousarray(x, dtype=np.int32)  # noqa: E731
        combos = [(list(samples), C, None) for _form, C, samples, _row0, _coef in self.lrows]
        ptr, out = [0], []
        for _form, C, _samples, row0, _coef in self.lrows:
            ptr.append(ptr[-1] + C.shape[0])
            out += list(range(row0, row0 + C.shape[0]))
        t = dict(row_ptr=i32(ptr), row_k=None, v_Mz=None, v_Mg=None, c_Mz=None, c_Mg=None,
                 var_off=i32(self.var_off), z_col=i32(self.z_col) if self.nz else None, row_out=i32(out))
        t.update(_combination_tables(combos))
        return t

    def lrow_tables(self, params, batch, dev=None):
        """terms [R, T, 3 + 2n + 2m] of mpci_lti_sched_rows: weight 1 and the plant form of every term."""
        xp = _Host if dev is None else _Device(dev)
        params = xp.params(params, batch)
        n, m = self.n, self.m
        symbolic = any(form.symbolic() or coef is not None for form, *_r, coef in self.lrows)
        out = xp.zeros([batch if symbolic else 1, len(self.lrows), 3 + 2 * n + 2 * m])
        for t, (form, *_r, coef) in enumerate(self.lrows):
            out[:, t, 0] = 1.0 if coef is None else _value(coef, params)
            out[:, t, 3:3 + n] = _dense(form.cx, n, params, batch, xp)
            out[:, t, 3 + n:3 + n + m] = _dense(form.cu, m, params, batch, xp)
        return out

    def device_lrows(self):
        if self._lrows is None:
            self._lrows = _DeviceLtiSched(self, self.lrow_desc_tables())
        return self._lrows

    def device_sched(self):
        if self._sched is None:
            self._sched = _DeviceLtiSched(self, self.sched_desc_tables())
        return self._sched

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
        if plan.embedded:
            e = plan.embed_tables()
            self._keep_embed = e
            at = lambda k: e[k].ctypes.data  # noqa: E731
            embed = LtiEmbed(
                plan.nv, plan.nc, at("var_off"), at("grp_row0"), plan.nz, at("z_col"), plan.ng,
                len(plan.coupled), at("cost_cx"), at("cost_cu"), at("cost_w"), at("cost_MzT"),
                at("cost_e_ptr"), at("cost_e_col"), at("cost_e_val"), at("Pzz"), at("Gz"),
                at("h_ptr"), at("h_col"), at("h_val"), at("row_kind"))
            device.check(device.lib().mpci_lti_plan_create_embedded(
                ctypes.byref(desc), ctypes.byref(embed), ctypes.byref(handle)))
        else:
            device.check(device.lib().mpci_lti_plan_create(ctypes.byref(desc), ctypes.byref(handle)))
        self.handle = handle
        self._lib = device.lib()

    def __del__(self):
        handle, self.handle = getattr(self, "handle", None), None
        if handle:
            self._lib.mpci_lti_plan_destroy(handle)


class _DeviceLtiSched:
    """mpci_lti_sched of a plan with scheduled costs."""

    def __init__(self, plan, tables):
        from . import device

        device.require_gpu()
        t = self._keep = tables
        at = lambda k: t[k].ctypes.data if t.get(k) is not None else None  # noqa: E731
        desc = LtiSchedDesc(plan.n, plan.m, plan.N, plan.nv, at("var_off"), plan.nz, at("z_col"),
                            plan.ng if plan.embedded else 0, len(t["row_ptr"]) - 1, at("row_ptr"), at("row_k"),
                            at("v_Mz"), at("v_Mg"), at("c_Mz"), at("c_Mg"), at("ent_ptr"), at("ent_k"),
                            at("ent_val"), at("c_ent_ptr"), at("c_ent_k"), at("c_ent_val"), at("row_out"))
        handle = ctypes.c_void_p()
        device.check(device.lib().mpci_lti_sched_create(ctypes.byref(desc), ctypes.byref(handle)))
        self.handle = handle
        self._lib = device.lib()

    def __del__(self):
        handle, self.handle = getattr(self, "handle", None), None
        if handle:
            self._lib.mpci_lti_sched_destroy(handle)


def _combination_tables(combos):
    """``combos``: per term (samples, L or None, cross L or None).  The rows as combinations of
    samples for mpci_lti_sched_desc: CSR ``ent_ptr / ent_k / ent_val`` over all rows (a term
    without L has one entry of 1 per row), ``c_ent_*`` when some cross L differs from its L."""
    i32 = lambda x: np.ascontiguousarray(x, dtype=np.int32)  # noqa: E731

    def csr(which):
        ptr, ks, vals = [0], [], []
        for samples, Lv, Lc in combos:
            L = (Lv, Lc if Lc is not None else Lv)[which]
            if L is None:
                ks += list(samples)
                vals += [1.0] * len(samples)
                ptr += list(range(ptr[-1] + 1, ptr[-1] + len(samples) + 1))
                continue
            for line in np.asarray(L, dtype=np.float64):
                cols = np.nonzero(line)[0]
                ks += [samples[c] for c in cols]
                vals += list(line[cols])
                ptr.append(len(ks))
        return i32(ptr), i32(ks), np.ascontiguousarray(vals, dtype=np.float64)

    out = dict(zip(("ent_ptr", "ent_k", "ent_val"), csr(0)))
    out.update(c_ent_ptr=None, c_ent_k=None, c_ent_val=None)
    if any(Lc is not None and Lc is not Lv for _s, Lv, Lc in combos):
        out.update(zip(("c_ent_ptr", "c_ent_k", "c_ent_val"), csr(1)))
    return out


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
    # state blocks follow the order of the given variables, inputs the optimisation order;
    # whatever else is given / optimised is auxiliary (another dynamics' variables, a bias)
    x0_names = [x0_base + ax for ax in axes]
    given = [g for g in form.given_variables if g in x0_names]
    if sorted(given) != sorted(x0_names):
        _unsupported("a plant whose initial states are not all given")
    blocks = [axes.index(g[len(x0_base):]) for g in given]  # block -> axis index
    block_of_axis = {axis: blk for blk, axis in enumerate(blocks)}
    wanted = {inp + ax: (block_of_axis[a], j) for a, ax in enumerate(axes) for j, inp in enumerate(base_inputs)}
    optim = [v for v in form.optim_variables if v in wanted]
    if sorted(optim) != sorted(wanted):
        _unsupported("a plant whose inputs are not all optimisation variables")
    inputs = [wanted[v] for v in optim]  # kernel input k -> (block, plant input j)
    return plant, axes, base_states, base_inputs, x0_base, n0, m0, blocks, inputs, block_of_axis


def _aux_layout(form, layout):
    """The variables outside the plant: ``(z_of, nz, z_col)`` -- auxiliary optimisation
    variable -> its range of z indices, their number and their columns in the QP -- and the
    set of auxiliary given variables (their columns are ``form.given_ID``)."""
    plant, axes, base_states, base_inputs, x0_base = layout[:5]
    plant_inputs = set(inp + ax for ax in axes for inp in base_inputs)
    z_of, z_col = {}, []
    for v in form.optim_variables:
        if v not in plant_inputs:
            ids = form.optim_ID[v]
            z_of[v] = range(len(z_col), len(z_col) + len(ids))
            z_col += list(ids)
    x0_names = set(x0_base + ax for ax in axes)
    aux_given = [g for g in form.given_variables if g not in x0_names]
    return z_of, len(z_col), z_col, aux_given


def _linear_forms(form, plant_name, layout):
    """{definition name: _Form} for every point-wise linear form."""
    plant, axes, base_states, base_inputs, x0_base, n0, m0, blocks, inputs, block_of_axis = layout
    z_of, nz, _z_col, aux_given = _aux_layout(form, layout)
    N = plant.matrices[-1].shape[0]
    forms = {}
    for a, ax in enumerate(axes):
        for i, name in enumerate(base_states):
            forms[name + ax] = _Form(cx={block_of_axis[a] * n0 + i: 1.0}, rows=N)
    k = 0
    for name in form.optim_variables:
        if name in z_of:  # an auxiliary optimisation variable: its own slots of z
            sel = np.zeros([len(z_of[name]), nz])
            sel[np.arange(len(z_of[name])), list(z_of[name])] = 1.0
            forms[name] = _Form(rows=sel.shape[0], Mz=sel)
        else:
            forms[name] = _Form(cu={k: 1.0}, rows=N)
            k += 1
    for name in aux_given:
        ids = form.given_ID[name]
        sel = np.zeros([len(ids), form.given_len])
        sel[np.arange(len(ids)), list(ids)] = 1.0
        forms[name] = _Form(rows=sel.shape[0], Mg=sel)
    skip = set(x0_base + ax for ax in axes)
    for name, combo in form.definitions.items():
        if name in forms or name in skip:
            continue
        out = _Form()
        try:
            for var, coef in combo.items():
                if var not in forms:
                    _unsupported("the definition '{}' (depends on '{}')".format(name, var))
                if isinstance(forms[var], Exception):
                    raise forms[var]
                kind, value = _plan._scalar_kind(coef)
                if kind == "mat":  # blocks of another dynamics (tools.py:79-101), shifted biases
                    out.add_matrix(forms[var], value if value.ndim == 2 else value.reshape([value.shape[0], -1]), name)
                else:
                    out.add(forms[var], value)
            forms[name] = out
        except (NotImplementedError, ValueError, TypeError) as problem:
            forms[name] = problem  # raised when a cost or a limit asks for this variable
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
    z_of, nz, z_col, aux_given = _aux_layout(form, layout)
    forms = _linear_forms(form, plant_name, layout)
    n, m = n0 * len(axes), len(inputs)
    N = plant.matrices[-1].shape[0]
    plan = LtiPlan(n, m, N)
    plan.n0, plan.m0, plan.blocks, plan.inputs = n0, m0, blocks, inputs
    plan.optim_ID = dict(form.optim_ID)
    plan.given_ID = dict(form.given_ID)
    # (limits with L are rows of the embedded layout: the fused kernel leaves them to the later stages)
    def rows_stage(limit):  # L, a signal with a matrix coefficient on a plant variable, row-wise Param arrows
        members = [forms.get(limit.variable + ax) for ax in limit.axes]
        arrow = np.asarray(limit.arrow)
        rowwise = arrow.dtype == object and arrow.shape[0] != 1 and any(is_expr(x) for x in arrow.ravel())
        return bool(limit.L) or rowwise or any(isinstance(f, _Form) and f.mix for f in members)

    plan.embedded = bool(nz or aux_given or any(rows_stage(limit) for limit in _plan.ordered_limits(form)))
    if plan.embedded:
        plan.nv, plan.ng, plan.nz, plan.z_col = form.optim_len, form.given_len, nz, list(z_col)
        plant_optim = [v for v in form.optim_variables if v not in z_of]
        for v in plant_optim:
            if len(form.optim_ID[v]) != N:
                _unsupported("a plant input with {} actions over {} samples".format(len(form.optim_ID[v]), N))
        plan.var_off = [form.optim_ID[v].start for v in plant_optim]
        plan.x0_index = [i for blk in blocks for i in form.given_ID[x0_base + axes[blk]]]

    def form_of(name, what):
        if name not in forms:
            _unsupported("{} on '{}'".format(what, name))
        if isinstance(forms[name], Exception):
            raise forms[name]
        return forms[name]

    # ---- limits, in the reference's stacking order (body.py:306-318) -------
    Gz_rows, Hg_rows, kinds = [], [], []
    for limit in _plan.ordered_limits(form):
        names = [limit.variable + ax for ax in limit.axes]
        members = [form_of(nm, "a limit") for nm in names]
        if rows_stage(limit):  # rows that combine samples: c @ Mo[sigma] with c = arrow * L  (body.py:253-255)
            rows = members[0].rows if members[0].rows is not None else N
            if rows > N or any(f.has_plant() and f.rows != N for f in members):
                _unsupported("a limit on '{}' ({} samples, the horizon has {})".format(names[0], rows, N))
            samples = [int(k) for k in limit.schedule] if limit.schedule else list(range(rows))
            if any(k < 0 or k >= rows for k in samples):
                raise IndexError("index {} is out of bounds for axis 0 with size {}".format(
                    next(k for k in samples if k < 0 or k >= rows), rows))
            arrow = np.asarray(limit.arrow)
            turned = arrow.dtype == object and any(is_expr(x) for x in arrow.ravel())
            lines = int(np.asarray(limit.L[0]).shape[0]) if limit.L else len(samples)
            if not limit.L and limit.nlines is not None and int(limit.nlines) != lines:
                raise ValueError("operands could not be broadcast together with shapes ({},1) ({},{})".format(
                    int(limit.nlines), lines, plan.nv))
            if arrow.shape[0] not in (1, lines):
                raise ValueError("operands could not be broadcast together with shapes ({},1) ({},{})".format(
                    arrow.shape[0], lines, plan.nv))
            Gz_new, Hg_new = np.zeros([lines, nz]), np.zeros([lines, form.given_len + 1])

            def emit(f, r0, C, coef):
                """Rows r0 .. r0 + len(C) of this limit: C [k, samples] numeric, times ``coef`` (a Param
                expression: it travels as the terms' weight / per-instance Gz and h-map terms) or None."""
                k = C.shape[0]
                for c2, Mz2, Mg2 in [(None, f.Mz, f.Mg)] + list(f.sym_aux):
                    both = coef if c2 is None else (c2 if coef is None else coef * c2)
                    if Mz2 is not None and np.any(Mz2):
                        if both is not None:
                            plan.Gz_terms.append((both, plan.nc + r0, C @ Mz2[samples]))
                        else:
                            Gz_new[r0:r0 + k] += C @ Mz2[samples]
                    if Mg2 is not None and np.any(Mg2):
                        if both is not None:
                            term = np.zeros([k, form.given_len + 1])
                            term[:, :-1] = -(C @ Mg2[samples])
                            plan.Hg_terms.append((both, plan.nc + r0, term))
                        else:
                            Hg_new[r0:r0 + k, :-1] -= C @ Mg2[samples]
                for part, M, _c in f.parts(N):  # the plant parts: direct, and through matrix coefficients
                    if not part.has_plant():
                        continue
                    if M is None:
                        full = np.zeros([k, N])
                        full[:, samples] = C
                    else:
                        full = C @ M[samples, :]
                    plan.lrows.append((_Form(part.cx, part.cu, N), full, list(range(N)), plan.nc + r0, coef))

            for a, f in enumerate(members):
                La = np.asarray(limit.L[a], dtype=np.float64) if limit.L else np.eye(lines)
                if La.shape != (lines, len(samples)):
                    raise ValueError("matmul: L is {} for {} scheduled samples and {} lines".format(
                        La.shape, len(samples), lines))
                # c = arrow * L (restrictions.py:175-179)
                column = arrow[:, a]
                if not any(is_expr(x) for x in column):
                    emit(f, 0, np.asarray(column, dtype=np.float64)[:, None] * La, None)
                elif column.shape[0] == 1:
                    emit(f, 0, La, column[0])
                else:  # an arrow that turns from row to row: every row its own weight
                    for r in range(lines):
                        if is_expr(column[r]):
                            emit(f, r, La[r:r + 1], column[r])
                        else:
                            emit(f, r, float(column[r]) * La[r:r + 1], None)
            for coef, col in _bound_parts(limit, lines):
                if is_expr(coef):
                    term = np.zeros([lines, form.given_len + 1])
                    term[:, -1] = col
                    plan.Hg_terms.append((coef, plan.nc, term))
                else:
                    Hg_new[:, -1] += float(coef) * col
            for e in np.asarray(limit.extreme, dtype=object).ravel():
                if is_expr(e):
                    plan.slots.require_nonnegative(e, "limit on '{}'".format(limit.variable))
            plan.row_offsets.append(plan.nc)
            plan.limit_rows.append(lines)
            Gz_rows.append(Gz_new)
            Hg_rows.append(Hg_new)
            kinds += [1] * lines
            plan.nc += lines
            continue
        on_plant = any(f.has_plant() for f in members) or not plan.embedded
        rows = N if on_plant else members[0].rows
        if rows is None:
            _unsupported("a limit on '{}' (no samples)".format(names[0]))
        k0, cnt = 0, rows
        if limit.schedule:
            sched = limit.schedule
            if not isinstance(sched, range) or sched.step != 1 or sched.start < 0 or sched.stop > rows:
                _unsupported("a limit whose schedule is not a contiguous range inside the horizon")
            k0, cnt = sched.start, len(sched)
        arrow = np.asarray(limit.arrow)
        turned = arrow.dtype == object and any(is_expr(x) for x in arrow.ravel())
        if turned:  # a limit turned per instance (rotate_in_TS by a Param): its arrow joins the form's coefficients
            if arrow.shape[0] != 1:
                _unsupported("a per-instance arrow that changes from row to row")
        else:
            arrow = arrow.astype(np.float64)
        nlines = limit.nlines
        if nlines is not None and int(nlines) != cnt:
            raise ValueError(
                "operands could not be broadcast together with shapes ({},1) ({},{})".format(
                    int(nlines), cnt, plan.nv))
        rows_form, rows_arrow = _Form(), np.ones([rows])
        turned_aux = []  # (arrow expression, Mz [rows, nz] or None, Mg [rows, ng] or None) per axis
        if turned:
            for a, f in enumerate(members):
                if is_expr(arrow[0, a]):  # the plant part joins the per-instance coefficients of the form,
                    rows_form.add(_Form(f.cx, f.cu, f.rows), arrow[0, a])
                    if f.has_aux():       # the part outside the plant the per-instance Gz / h map
                        turned_aux.append((arrow[0, a], f.Mz, f.Mg))
                        turned_aux += [(arrow[0, a] * c2, Mz2, Mg2) for c2, Mz2, Mg2 in f.sym_aux]
                else:
                    rows_form.add(f, float(arrow[0, a]))
            rows_arrow = np.zeros([rows])
            rows_arrow[k0:k0 + cnt] = 1.0
        elif len(names) == 1:
            rows_form.add(members[0], 1.0)
            column = arrow[:, 0]
            rows_arrow = np.zeros([rows])
            rows_arrow[k0:k0 + cnt] = np.broadcast_to(column, [cnt]) if column.shape[0] == 1 else column
        else:
            if arrow.shape[0] != 1:
                _unsupported("a limit over several axes whose arrow changes from row to row")
            for a, f in enumerate(members):  # one facet: the axes' forms weighted by the normal
                rows_form.add(f, float(arrow[0, a]))
        # bound = extreme + sum_a arrow_a * center_a   (restrictions.py:198-199), rows k0..
        parts = []
        for coef, col in _bound_parts(limit, cnt):
            full = np.zeros([rows])
            full[k0:k0 + cnt] = col
            parts.append((coef, full))
        for e in np.asarray(limit.extreme, dtype=object).ravel():
            if is_expr(e):  # the reference would flip such a limit (restrictions.py:190-194)
                plan.slots.require_nonnegative(e, "limit on '{}'".format(limit.variable))
        plan.row_offsets.append(plan.nc)
        plan.limit_rows.append(cnt)
        if plan.embedded:  # the rows' part over z and over the auxiliary givens (+ constant)
            Mz = rows_form.Mz if rows_form.Mz is not None else np.zeros([rows, nz])
            Mg = rows_form.Mg if rows_form.Mg is not None else np.zeros([rows, form.given_len])
            Gz_rows.append((rows_arrow[:, None] * Mz)[k0:k0 + cnt])
            Hg = np.zeros([cnt, form.given_len + 1])
            Hg[:, :-1] = -(rows_arrow[:, None] * Mg)[k0:k0 + cnt]
            if not on_plant:  # (the bounds of the plant's groups travel in the group tables)
                for coef, col in parts:
                    if is_expr(coef):
                        term = np.zeros([cnt, form.given_len + 1])
                        term[:, -1] = col[k0:k0 + cnt]
                        plan.Hg_terms.append((coef, plan.nc, term))
                    else:
                        Hg[:, -1] += float(coef) * col[k0:k0 + cnt]
            # (auxiliary parts under a per-instance coefficient of a definition: scaled by the rows' arrows)
            turned_aux += [(c2, None if Mz2 is None else rows_arrow[:, None] * Mz2,
                            None if Mg2 is None else rows_arrow[:, None] * Mg2) for c2, Mz2, Mg2 in rows_form.sym_aux]
            for coef, Mz_a, Mg_a in turned_aux:
                if Mz_a is not None and np.any(Mz_a):
                    plan.Gz_terms.append((coef, plan.nc, np.array(Mz_a[k0:k0 + cnt])))
                if Mg_a is not None and np.any(Mg_a):
                    term = np.zeros([cnt, form.given_len + 1])
                    term[:, :-1] = -Mg_a[k0:k0 + cnt]
                    plan.Hg_terms.append((coef, plan.nc, term))
            Hg_rows.append(Hg)
            kinds += [0 if on_plant else 1] * cnt
        if on_plant:
            plan.group_row0.append(plan.nc)
            plan.groups.append((rows_form, rows_arrow, parts, k0, cnt))
        plan.nc += cnt

    # ---- costs, in dict order (body.py:322-329) -----------------------------
    Pzz = np.zeros([nz, nz])
    for cost in form.goals.values():
        if isinstance(cost.weight, np.ndarray) and cost.weight.ndim:
            raise TypeError("a cost weight must be a scalar")
        w = cost.weight if is_expr(cost.weight) else float(cost.weight)
        if not is_expr(w) and w == 0.0 and (plan.embedded or cost.L or cost.cross_L or cost.schedule):
            continue  # adds exact zeros (the biped's terminal cost, biped_configuration.py:49)
        for a, ax in enumerate(cost.axes):
            vname, cname = cost.variable + ax, cost.cross + ax
            fv, fc = form_of(vname, "a cost"), form_of(cname, "a cost")
            aim = _single(np.asarray(cost.aim, dtype=object)[:, a], "an aim")
            caim = _single(np.asarray(cost.cross_aim, dtype=object)[:, a], "an aim")
            # What the fused kernel's stage weights (and the coupling stage's constant tables) cannot
            # express goes through the general stage of mpci_sched.cu, row by row: a schedule; a cost
            # that couples states and inputs (no Qxu block in T' Qxx T + Quu x I); a crossed cost or a
            # per-instance aim on a signal outside the plant that has fewer samples than the horizon.
            mixes = (fv.touches_states() and fc.touches_inputs()) or (fv.touches_inputs() and fc.touches_states())
            outside = fv.has_aux() or fc.has_aux()
            short = not fv.has_plant() and fv.rows != N  # (e.g. a cost on the footsteps themselves)
            # a matrix coefficient on a plant variable, a Param coefficient on an auxiliary one: no stage form
            mixed = bool(fv.mix or fc.mix or fv.sym_aux or fc.sym_aux)
            general = mixes or mixed or (outside and (fv is not fc or ((is_expr(aim) or is_expr(caim)) and short)))
            if cost.schedule or general or cost.L or cost.cross_L:
                rows = N if (fv.has_plant() or fc.has_plant() or not plan.embedded) else fv.rows
                if rows is None or rows > N or (fc.rows is not None and not fc.has_plant() and fc.rows < rows):
                    _unsupported("a cost on '{}' ({} samples, the horizon has {})".format(vname, rows, N))
                samples = [int(k) for k in cost.schedule] if cost.schedule else list(range(rows))
                if any(k < 0 or k >= rows for k in samples):
                    raise IndexError("index {} is out of bounds for axis 0 with size {}".format(
                        next(k for k in samples if k < 0 or k >= rows), rows))
                Lv = Lc = None
                if cost.L or cost.cross_L or mixed:  # vMo = L @ Mo[sigma], cMo = cross_L @ Mo_c[sigma]  (body.py:279-290)
                    Lv = np.asarray(cost.L[a], dtype=np.float64) if cost.L else np.eye(len(samples))
                    same = (not cost.cross_L and not cost.L) or cost.cross_L is cost.L
                    Lc = Lv if same else (np.asarray(cost.cross_L[a], dtype=np.float64) if cost.cross_L
                                          else np.eye(len(samples)))
                    for M in (Lv, Lc):
                        if M.ndim != 2 or M.shape[1] != len(samples) or M.shape[0] != Lv.shape[0]:
                            raise ValueError("matmul: L is {} for {} scheduled samples ({} lines)".format(
                                M.shape, len(samples), Lv.shape[0]))
                if mixed:  # V = sum_a L_a T_a, C = sum_b L_b T_b: one crossed term per pair of parts
                    if rows != N:
                        _unsupported("a cost on '{}' ({} samples, the horizon has {})".format(vname, rows, N))

                    def over_horizon(L0, M):
                        if M is not None:
                            return L0 @ M[samples, :]
                        full = np.zeros([L0.shape[0], N])
                        full[:, samples] = L0
                        return full

                    vparts = fv.parts(N)
                    cparts = vparts if fc is fv else fc.parts(N)
                    for ia, (pa, Ma, ka) in enumerate(vparts):
                        for ib, (pb, Mb, kb) in enumerate(cparts):
                            weight = w  # w coef_a coef_b: the parts' per-instance coefficients ride in the weight
                            for k in (ka, kb):
                                weight = weight if k is None else (k * weight if is_expr(k) or is_expr(weight)
                                                                   else float(k) * weight)
                            plan.sched_L[len(plan.sched_terms)] = (over_horizon(Lv, Ma), over_horizon(Lc, Mb))
                            plan.sched_terms.append((weight, pa, pb, aim if ia == 0 else 0.0,
                                                     caim if ib == 0 else 0.0, list(range(N))))
                    continue
                if Lv is not None:
                    plan.sched_L[len(plan.sched_terms)] = (Lv, Lc)
                plan.sched_terms.append((w, fv, fc, aim, caim, samples))
                continue
            if outside:
                rows = fv.rows
                if rows != N and (fv.has_plant() or rows > N):
                    _unsupported("a cost on '{}' ({} samples, the horizon has {})".format(vname, rows, N))
                # (a signal of fewer samples, e.g. the footsteps: the missing rows are exact zeros)
                Mz, Mg = np.zeros([N, nz]), np.zeros([N, form.given_len])
                if fv.Mz is not None:
                    Mz[:rows] = fv.Mz
                if fv.Mg is not None:
                    Mg[:rows] = fv.Mg
                E = np.zeros([N, form.given_len + 1])  # e = Mg given - aim
                E[:, :-1], E[:rows, -1] = Mg, (0.0 if is_expr(aim) else -float(aim))
                plan.coupled.append((w, fv, Mz, E))
                plan.coupled_aims.append(aim if is_expr(aim) else None)  # (then subtracted per instance)
                if is_expr(w):  # body.py:292 with Vo = Co = [T | Mz]
                    plan.Pzz_terms.append((w, Mz.T @ Mz))
                else:
                    Pzz += w * (Mz.T @ Mz)
            if fv.has_plant() or fc.has_plant() or not plan.embedded:
                # (a coupled cost's aim is carried by E alone: the plant part must not subtract it again)
                coupled_here = fv.has_aux() or fc.has_aux()
                plan.cost_terms.append((w, fv, fc, 0.0 if coupled_here else aim, 0.0 if coupled_here else caim))
    if plan.embedded:
        plan.Pzz = Pzz
        plan.Gz = np.vstack(Gz_rows) if Gz_rows else np.zeros([0, nz])
        plan.Hg = np.vstack(Hg_rows) if Hg_rows else np.zeros([0, form.given_len + 1])
        plan.row_kind = np.array(kinds, dtype=np.int32)

    plan.per_instance = (
        any(is_expr(w) or fv.symbolic() or fc.symbolic() for w, fv, fc, _, _ in plan.cost_terms)
        or any(f.symbolic() or any(is_expr(c) for c, _ in parts) for f, _, parts, _, _ in plan.groups))
    plan.varying = plan._find_varying() if plan.per_instance else set()
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


def _lti_tables(tensors):
    """mpci_lti_tables of the per-instance CUDA tensors present in ``tensors`` (NULL = the plan's)."""
    return LtiTables(*(tensors[k].data_ptr() if k in tensors else None
                       for k in ("Qxx", "Quu", "tx", "tu", "cx", "cu", "bound")))


class LtiJob:
    """The device-resident inputs of one batch for a lowered plan: the expanded plants, the
    given vectors, the scalar slots or per-instance tables (evaluated on the GPU) -- everything
    ``launch`` hands to ``mpci_lti_assemble`` / ``_ex`` / ``_coupled``.  Building the job and
    launching it are separate so that a receding-horizon loop (or a benchmark) can launch the
    same inputs again without re-evaluating the tables."""

    def __init__(self, plan, A, B, x0, params=None):
        import torch

        self.plan = plan
        self.handle = plan.device_plan().handle
        dev = self.dev = torch.device("cuda", torch.cuda.current_device())

        def to_dev(x):
            if torch.is_tensor(x):
                return x.to(device=dev, dtype=torch.float64).contiguous()
            return torch.from_numpy(np.require(x, dtype=np.float64, requirements=["C", "W"])).to(dev)

        A, B, x0 = to_dev(A), to_dev(B), to_dev(x0)
        n, m, n0, m0 = plan.n, plan.m, plan.n0, plan.m0
        if x0.dim() != 2 or x0.shape[1] != plan.ng:
            raise ValueError(
                "matmul: the given vector has {} elements per instance, the formulation "
                "expects {}".format(x0.shape[-1], plan.ng)
            )
        self.given = x0
        if plan.embedded:  # the plant's initial state, in the kernel's block order
            x0 = x0[:, torch.as_tensor(plan.x0_index, device=dev)].contiguous()
        self.x0 = x0
        batch = self.batch = x0.shape[0]
        if tuple(A.shape[-2:]) != (n0, n0) or tuple(B.shape[-2:]) != (n0, m0):
            raise ValueError("A must be [..., {0}, {0}] and B [..., {0}, {1}]".format(n0, m0))
        for M in (A, B):
            if M.dim() == 3 and M.shape[0] != batch:
                raise ValueError("the dynamics hold {} instances, x0 {}".format(M.shape[0], batch))
        A, B = plan.expand_device(A, B)
        self.A, self.B = A.contiguous(), B.contiguous()
        self.strideA = n * n if A.dim() == 3 else 0
        self.strideB = n * m if B.dim() == 3 else 0
        self.tables = self.scalars = self.ccx = self.ccu = self.coupled = None
        if plan.per_instance or plan.embedded or plan.sched_terms:
            missing = [nm for nm in plan.parameter_names() if nm not in (params or {})]
            if missing:
                raise KeyError("no value given for the parameter '{}'".format(missing[0]))
        if plan.per_instance:
            self.tensors = plan.instance_tables(params, batch, dev)
            self.tables = _lti_tables(self.tensors)
        elif len(plan.slots.exprs):
            self.scalars = to_dev(plan.slots.evaluate(params or {}, batch))
        if plan.embedded and plan.coupled_symbolic():
            from . import device

            self.ccx, self.ccu = (t.contiguous() for t in plan.coupled_tables(params, batch, dev))
            self.cextra = tuple(t.contiguous() if t is not None else None
                                for t in plan.coupled_extras(params, batch, dev))
            at = lambda t: t.data_ptr() if t is not None else None  # noqa: E731
            self.coupled = LtiCoupledTables(at(self.ccx), at(self.ccu), *(at(t) for t in self.cextra))
        self.lrow_terms = None
        if plan.lrows:
            self.lrows = plan.device_lrows()
            self.lrow_terms = plan.lrow_tables(params, batch, dev).contiguous()
        self.sched_terms = None
        if plan.sched_terms:
            self.sched = plan.device_sched()
            self.sched_terms = plan.sched_tables(params, batch, dev).contiguous()

    def launch(self, out=None, hessian=HESSIAN_DMMA, parts="PqGh"):
        """One assembly of the batch into ``out`` (a ``QPBatch``, allocated when None) on the
        current stream; ``parts`` picks the outputs (the others are not computed)."""
        import torch

        from . import device

        plan, batch = self.plan, self.batch
        if out is None:
            out = device.QPBatch(batch, plan.nv, plan.nc, self.dev)
        elif (out.batch, out.nv, out.nc) != (batch, plan.nv, plan.nc):
            raise ValueError("the output batch does not match the plan")
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        ptr = device._ptr
        outs = [ptr(t) if tag in parts else None for tag, t in zip("PqGh", (out.P, out.q, out.G, out.h))]
        tables = ctypes.byref(self.tables) if self.tables is not None else None
        head = (self.handle, ptr(self.A), self.strideA, ptr(self.B), self.strideB, ptr(self.x0))
        lib = device.lib()
        if plan.embedded:
            coupled = ctypes.byref(self.coupled) if self.coupled is not None else None
            rc = lib.mpci_lti_assemble_coupled_ex(*head, ptr(self.given), ptr(self.scalars), tables, coupled,
                                                  *outs, batch, int(hessian), stream)
        elif plan.per_instance:
            rc = lib.mpci_lti_assemble_ex(*head, None, tables, *outs, batch, int(hessian), stream)
        else:
            rc = lib.mpci_lti_assemble(*head, ptr(self.scalars), *outs, batch, int(hessian), stream)
        device.check(rc)
        if self.lrow_terms is not None and (outs[2] is not None or outs[3] is not None):
            terms = self.lrow_terms
            stride = terms.shape[1] * terms.shape[2] if terms.shape[0] == batch and batch > 1 else 0
            device.check(lib.mpci_lti_sched_rows(
                self.lrows.handle, ptr(self.A), self.strideA, ptr(self.B), self.strideB, ptr(self.x0),
                ptr(terms), stride, outs[2], outs[3], plan.nc, batch, stream))
        if self.sched_terms is not None and (outs[0] is not None or outs[1] is not None):
            terms = self.sched_terms
            stride = terms.shape[1] * terms.shape[2] if terms.shape[0] == batch and batch > 1 else 0
            device.check(lib.mpci_lti_sched_add(
                self.sched.handle, ptr(self.A), self.strideA, ptr(self.B), self.strideB, ptr(self.x0),
                ptr(self.given) if plan.embedded else None, ptr(terms), stride, outs[0], outs[1], batch, stream))
        out._keep = self  # the launches are asynchronous: the inputs live as long as the result
        return out


def assemble(plan, A, B, x0, params=None, out=None, hessian=HESSIAN_DMMA):
    """Runs ``mpci_lti_assemble`` (``mpci_lti_assemble_ex`` for a plan with per-instance
    tables, ``mpci_lti_assemble_coupled`` for an embedded one).  ``A`` [batch, n0, n0] or
    [n0, n0] (shared), ``B`` likewise -- the plant of ONE axis -- and ``x0`` [batch, given_len]
    in the formulation's given order: numpy arrays or CUDA tensors (float64)."""
    return LtiJob(plan, A, B, x0, params).launch(out, hessian)


def _host_inputs(plan, A, B, x0, params):
    n, m, n0, m0 = plan.n, plan.m, plan.n0, plan.m0
    if plan.per_instance:
        _unsupported("Param coefficients through the host-buffer entry points")
    if plan.sched_terms:
        _unsupported("costs with a schedule through the host-buffer entry points")
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

    if plan.per_instance or plan.sched_terms or plan.embedded:
        return _assemble_host_chunks(plan, A, B, x0, params, hessian, chunk, out)
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


def _assemble_host_chunks(plan, A, B, x0, params, hessian, chunk, out):
    """HOST results for the plans ``mpci_lti_assemble_host`` does not take (per-instance tables,
    an embedded layout, scheduled costs): chunks of instances through ``LtiJob`` on two streams
    with two device slots -- chunk c's four device-to-host copies (into pinned memory) run behind
    chunk c + 1's kernels.  ``x0`` [batch, given_len] in the formulation's given order."""
    import torch

    from . import device

    device.require_gpu()
    x0 = np.ascontiguousarray(x0, dtype=np.float64)
    if x0.ndim != 2:
        raise ValueError("the given vectors must be [batch, {}]".format(plan.ng))
    batch = x0.shape[0]
    if not torch.is_tensor(A):
        A = np.asarray(A, dtype=np.float64)
    if not torch.is_tensor(B):
        B = np.asarray(B, dtype=np.float64)
    if out is None:
        shapes = ([batch, plan.nv, plan.nv], [batch, plan.nv], [batch, plan.nc, plan.nv], [batch, plan.nc])
        out = tuple(torch.empty(shape, dtype=torch.float64).pin_memory() for shape in shapes)
    dev = torch.device("cuda", torch.cuda.current_device())
    chunk = max(1, min(int(chunk), batch))
    slots = [device.QPBatch(chunk, plan.nv, plan.nc, dev) for _ in range(2 if batch > chunk else 1)]
    streams = [torch.cuda.Stream(device=dev) for _ in slots]
    part = lambda v, lo, hi: v[lo:hi] if (hasattr(v, "__len__") and len(v) == batch) else v  # noqa: E731
    jobs = []
    for c, lo in enumerate(range(0, batch, chunk)):
        hi = min(lo + chunk, batch)
        slot, stream = slots[c % len(slots)], streams[c % len(slots)]
        view = slot if hi - lo == chunk else _slot_view(slot, hi - lo)
        with torch.cuda.stream(stream):
            job = LtiJob(plan, part(A, lo, hi) if A.ndim == 3 else A, part(B, lo, hi) if B.ndim == 3 else B,
                         x0[lo:hi], {k: part(v, lo, hi) for k, v in (params or {}).items()})
            job.launch(view, hessian)
            for host, devt in zip(out, (view.P, view.q, view.G, view.h)):
                host[lo:hi].copy_(devt, non_blocking=True)
        jobs.append(job)  # (inputs stay alive until the streams are drained)
    for stream in streams:
        stream.synchronize()
    return out


def _slot_view(slot, count):
    """The first ``count`` instances of a device slot as a QPBatch."""
    from . import device

    view = device.QPBatch.__new__(device.QPBatch)
    view.batch, view.nv, view.nc = count, slot.nv, slot.nc
    view.P, view.q, view.G, view.h = slot.P[:count], slot.q[:count], slot.G[:count], slot.h[:count]
    return view


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
