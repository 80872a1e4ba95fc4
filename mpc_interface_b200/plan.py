"""Compiler: Formulation -> affine assembly plan for the batched CUDA path.

What varies between the instances of a batch is a vector of per-instance
scalars ``s_b`` (Param expressions of LineCombo coefficients, cost weights and
aims, params.py) and the given vector ``g_b``.  Every preview matrix of the
reference (body.py:149-193) is then a polynomial in ``s_b`` whose coefficients
are constant matrices, and so are the outputs of body.py:236-348:

    P_b = sum_m  c_m(s_b) * DP_m                      (nv x nv)
    G_b = sum_m  c_m(s_b) * DG_m                      (nc x nv)
    [q_b; h_b] = sum_m c_m(s_b) * DY_m @ [g_b; 1]     (nv + nc)

with monomials ``c_m(s) = prod_f s[idx_mf]``.  This module walks the
formulation in the reference's order (definitions in insertion order, limits:
constraints dict then boxes dict, costs in dict order), keeps the symbolic
part (monomials, index layout, row offsets) on the host and sends every dense
product / accumulation to ``ops`` -- the device backend (device.py: DMMA GEMM
and 2-D axpy kernels of csrc/).  The result, an ``AffinePlan``, is what the
streaming kernels of csrc/mpci_assemble.cu consume.

No floating-point reduction happens in this file; numpy is used for
placement of blocks into zero matrices (the scatter of body.py:158-177) and
for structural scans (which entries of DY are non-zero).
"""
import numbers

import numpy as np

from .params import Expr, is_expr, split_array

MAX_FACTORS = 4  # factors per monomial the kernels accept (mpci.h: MPCI_MAX_FACTORS)
MAX_TERMS = 32  # monomials per streamed array (mpci.h: MPCI_MAX_TERMS; fast path up to 8)


def _array_token(x):
    """Hashable content of an array of numbers and / or Param expressions."""
    arr = np.asarray(x)
    if arr.dtype == object:
        return ("o", arr.shape, tuple(e.key() if is_expr(e) else float(e) for e in arr.ravel()))
    arr = np.ascontiguousarray(arr)
    return (arr.shape, arr.dtype.str, arr.tobytes())


def _schedule_token(schedule):
    if isinstance(schedule, range):
        return (schedule.start, schedule.stop, schedule.step)
    return tuple(schedule) if schedule is not None else None


def cost_token(cost):
    """Everything of a Cost that generate_qp_cost reads (body.py:266-302)."""
    return ("cost", cost.variable, cost.cross, tuple(cost.axes), _array_token(cost.weight),
            _array_token(cost.aim), _array_token(cost.cross_aim),
            tuple(_array_token(m) for m in cost.L), tuple(_array_token(m) for m in cost.cross_L),
            _schedule_token(cost.schedule))


def limit_token(limit):
    """Everything of a Constraint that generate_qp_constraint reads (body.py:236-264)."""
    return ("limit", limit.variable, tuple(limit.axes), _array_token(limit.arrow),
            _array_token(limit.center), _array_token(limit.extreme),
            tuple(_array_token(m) for m in limit.L), _schedule_token(limit.schedule))


class SlotTable:
    """Distinct per-instance scalar expressions, in order of first use."""

    def __init__(self):
        self.exprs = []
        self._index = {}
        self.nonnegative = []  # (expression, description): checked at every evaluate

    def require_nonnegative(self, expr, what):
        """The reference flips a limit whose extreme is negative
        (restrictions.py:190-194); the host cannot do that for a per-instance
        extreme, so every batch is checked instead of silently differing."""
        if all(expr.key() != e.key() for e, _ in self.nonnegative):
            self.nonnegative.append((expr, what))

    def slot(self, expr):
        key = expr.key()
        if key not in self._index:
            self._index[key] = len(self.exprs)
            self.exprs.append(expr)
        return self._index[key]

    def names(self):
        out = []
        for e in self.exprs:
            for n in e.names():
                if n not in out:
                    out.append(n)
        return out

    def evaluate(self, values, batch):
        """[batch, n_slots] float64 (at least one column so pointers stay valid)."""
        out = np.ones([batch, max(1, len(self.exprs))])
        for i, e in enumerate(self.exprs):
            out[:, i] = e.evaluate(values)
        for e, what in self.nonnegative:
            v = np.broadcast_to(np.asarray(e.evaluate(values), dtype=np.float64), [batch])
            if not np.all(v >= 0):
                raise ValueError(
                    "{}: the per-instance extreme '{}' is negative for instance {} ({:g}); "
                    "the reference would flip that limit (restrictions.py:190-194), which "
                    "needs a numeric extreme".format(what, e, int(np.argmax(~(v >= 0))), float(v[np.argmax(~(v >= 0))]))
                )
        return out


def _mono(*parts):
    """Merge slot tuples / None into one sorted monomial."""
    slots = []
    for p in parts:
        if p is None:
            continue
        slots.extend(p if isinstance(p, tuple) else (p,))
    return tuple(sorted(slots))


class Poly:
    """{monomial: matrix handle}; all matrices share one shape."""

    def __init__(self, ops, rows, cols):
        self.ops, self.rows, self.cols = ops, rows, cols
        self.terms = {}

    def acc(self, mono, src, alpha=1.0, rowscale=None):
        """terms[mono] += alpha * rowscale * src  (rows of src / rowscale broadcast)."""
        if mono not in self.terms:
            self.terms[mono] = self.ops.zeros(self.rows, self.cols)
        self.ops.axpy(self.terms[mono], src, float(alpha), rowscale)

    def take_rows(self, index):
        out = Poly(self.ops, len(index), self.cols)
        for mono, H in self.terms.items():
            out.terms[mono] = self.ops.rows(H, index)
        return out

    def left_multiply(self, matrix):
        matrix = np.atleast_2d(np.asarray(matrix, dtype=np.float64))
        out = Poly(self.ops, matrix.shape[0], self.cols)
        M = self.ops.upload(matrix)
        for mono, H in self.terms.items():
            out.terms[mono] = self.ops.gemm(M, H)
        return out


def _scalar_kind(coef):
    """('num', float) | ('expr', Expr) | ('mat', ndarray)"""
    if is_expr(coef):
        return "expr", coef
    if isinstance(coef, numbers.Real):
        return "num", float(coef)
    arr = np.asarray(coef)
    if arr.dtype == object:
        raise TypeError("array coefficients holding Param expressions are not supported")
    if arr.ndim == 0:
        return "num", float(arr)
    return "mat", arr.astype(np.float64)


class AffinePlan:
    """Host description + constant device arrays of one formulation structure."""

    def __init__(self):
        self.nv = self.ng = self.nc = 0
        self.slots = None
        self.monomials = []  # list of slot tuples (global table)
        self.P_terms = []  # indices into monomials
        self.G_terms = []
        self.Y_terms = []
        self.DP = self.DG = self.DY = None  # stacked constants [K, rows, cols]
        self.row_offsets = []  # first row of every limit in G / h
        self.limit_rows = []
        self.Y_segs = None  # int32 [n_segs, 4]: (row, term, nz_begin, nz_end) of DY
        self.Y_val = None  # float64 [n_nz]
        self.Y_col = None  # int32 [n_nz]
        self.optim_ID = {}
        self.given_ID = {}

    def mono_table(self):
        table = -np.ones([max(1, len(self.monomials)), MAX_FACTORS], dtype=np.int32)
        for i, mono in enumerate(self.monomials):
            table[i, : len(mono)] = mono
        return table


def _preview_polys(form, ops, slots):
    """PM as polynomials: {variable: Poly of [rows x (ng + nv)]}  (body.py:149-193)."""
    ng, nv = form.given_len, form.optim_len
    width = ng + nv
    PM = {}
    for name, combo in form.definitions.items():
        if name in form.of:  # body.py:158-177 -- scatter into zero matrices
            rows = form.dynamics[form.of[name]].all_variables[name]
            dense = np.zeros([rows, width])
            for var, block in combo.items():
                if is_expr(block):
                    raise TypeError("dynamics blocks cannot be Param expressions")
                if var in form.given_variables:
                    ids = form.given_ID[var]
                    dense[:, ids.start : ids.stop] = block
                elif var in form.optim_variables:
                    ids = form.optim_ID[var]
                    dense[:, ng + ids.start : ng + ids.stop] = block
                else:
                    raise ValueError(
                        "The variable {} in the definition of {} seems to not be "
                        "given nor optimal.".format(var, name)
                    )
            poly = Poly(ops, rows, width)
            poly.terms[()] = ops.upload(dense)
        else:  # body.py:179-193 -- sum of coefficient * dependency
            poly = None
            for var, coef in combo.items():
                kind, value = _scalar_kind(coef)
                dep = PM[var]
                if kind == "mat":
                    term = dep.left_multiply(value)
                    factor, alpha = None, 1.0
                elif kind == "expr":
                    term, factor, alpha = dep, slots.slot(value), 1.0
                else:
                    term, factor, alpha = dep, None, value
                if poly is None:
                    poly = Poly(ops, term.rows, width)
                for mono, H in term.terms.items():
                    poly.acc(_mono(mono, factor), H, alpha)
        PM[name] = poly
    return PM


def ordered_limits(form):
    """Row order of G/h: constraints dict, then boxes dict (body.py:306-318)."""
    limits = [lim for area in form.constraints.values() for lim in area]
    limits += [lim for box in form.constraint_boxes.values() for lim in box.constraints]
    return limits


def _atoms_mono(slots, atoms):
    """Slot indices of the atoms of one geometry monomial (params.split_array)."""
    return tuple(slots.slot(a) for a in atoms)


def _limit_poly(form, ops, PM, limit, slots):
    """(Poly [c_rows x (ng+nv)], bound parts) of one limit (body.py:236-264).

    ``arrow`` / ``center`` / ``extreme`` may hold per-instance expressions (a box
    rotated by each instance's yaw, restrictions.py:437-457): every multiplier
    array is split into ``sum_m prod(atoms_m) * C_m`` with numeric ``C_m`` and each
    part goes through the numeric path with its atoms joined to the monomial.
    The bound comes back as ``[(monomial, [c_rows x 1] array)]``."""
    width = form.given_len + form.optim_len
    rows = PM[limit.variable + limit.axes[0]].rows
    nlines = limit.nlines
    c_rows = rows if nlines is None else int(nlines)
    schedule = range(rows) if not limit.schedule else limit.schedule
    if limit.schedule and not limit.L and len(schedule) != c_rows:
        # nlines is stop - start (restrictions.py:152-153); a stepped range selects fewer
        # rows and the reference fails in the broadcast of body.py:258-259
        raise ValueError(
            "operands could not be broadcast together: the schedule {} selects {} rows "
            "but the limit counts {} lines (stop - start)".format(limit.schedule, len(schedule), c_rows)
        )
    out = Poly(ops, c_rows, width)
    multipliers = limit.matrices()
    for a, axis in enumerate(limit.axes):
        picked = PM[limit.variable + axis].take_rows(list(schedule))
        for atoms, c in split_array(multipliers[a]):
            extra = _atoms_mono(slots, atoms)
            if limit.L:
                term = picked.left_multiply(c)
                for mono, H in term.terms.items():
                    out.acc(_mono(mono, extra), H)
            else:
                k = c.shape[0]
                if k != 1 and picked.rows != 1 and k != picked.rows:
                    raise ValueError(
                        "operands could not be broadcast together with shapes "
                        "({},1) ({},{})".format(k, picked.rows, width)
                    )
                if max(k, picked.rows) not in (1, c_rows):
                    raise ValueError(
                        "non-broadcastable output operand with shape ({},{})".format(c_rows, width)
                    )
                for mono, H in picked.terms.items():
                    out.acc(_mono(mono, extra), H, 1.0, rowscale=c[:, 0])
    for e in np.asarray(limit.extreme, dtype=object).ravel():
        if is_expr(e):
            slots.require_nonnegative(e, "limit on '{}'".format(limit.variable))
    bound_parts = []
    for atoms, col in split_array(limit.bound()):
        col = col.reshape([-1, 1])
        if col.shape[0] not in (1, c_rows):
            raise ValueError("the bound has {} rows for {} lines".format(col.shape[0], c_rows))
        bound_parts.append((_mono(_atoms_mono(slots, atoms)), np.broadcast_to(col, [c_rows, 1])))
    return out, bound_parts


def _aim_entry(aim, axis_index):
    """The single aim value of one axis: ('num', x) or ('expr', e).
    Multi-row aims do not work in the reference either (body.py:296-299
    broadcasts (m,) against (rows,1)); they are refused here."""
    column = aim[:, axis_index]
    if column.shape[0] != 1:
        raise ValueError(
            "operands could not be broadcast together: an aim must have one row "
            "per axis, got {}".format(column.shape[0])
        )
    value = column[0]
    if is_expr(value):
        return "expr", value
    return "num", float(value)


def _block_acc(poly, mono, view, src, alpha):
    """poly.terms[mono][view] += alpha * src  (view = (r0, r1, c0, c1))."""
    if mono not in poly.terms:
        poly.terms[mono] = poly.ops.zeros(poly.rows, poly.cols)
    dst = poly.ops.sub(poly.terms[mono], *view)
    poly.ops.axpy(dst, src, float(alpha), None)


def _cost_into(form, ops, PM, slots, cost, Pacc, Yacc):
    """Adds one cost to the P and [q; h] accumulators (body.py:266-302)."""
    ng, nv = form.given_len, form.optim_len
    kind, w = _scalar_kind(cost.weight)
    if kind == "mat":
        raise TypeError("a cost weight must be a scalar")
    w_slot, w_num = (slots.slot(w), 1.0) if kind == "expr" else (None, w)
    if kind == "num" and w_num == 0.0:
        return  # contributes exact zeros in the reference
    rows = PM[cost.variable + cost.axes[0]].rows
    schedule = list(range(rows) if not cost.schedule else cost.schedule)

    for i, axis in enumerate(cost.axes):
        V = PM[cost.variable + axis].take_rows(schedule)
        if cost.L:
            V = V.left_multiply(cost.L[i])
        same = (not cost.crossed) or (
            cost.cross == cost.variable and cost.cross_L is cost.L
        )
        if same:
            C = V
        else:
            C = PM[cost.cross + axis].take_rows(schedule)
            if cost.cross_L:
                C = C.left_multiply(cost.cross_L[i])
        if C.rows != V.rows:
            raise ValueError(
                "matmul: the cost variable has {} rows and its cross {}".format(V.rows, C.rows)
            )
        ones = ops.upload(np.ones([V.rows, 1]))
        for m1, HV in V.terms.items():
            Vo = ops.sub(HV, 0, V.rows, ng, ng + nv)
            Vg = ops.sub(HV, 0, V.rows, 0, ng)
            for m2, HC in C.terms.items():
                mono = _mono(w_slot, m1, m2)
                Co = ops.sub(HC, 0, C.rows, ng, ng + nv)
                # Vo^T [Cg | Co]: the right block is the Hessian term (body.py:292),
                # the left one multiplies given in the gradient (body.py:296)
                T1 = ops.gemm(Vo, HC, transA=True)
                _block_acc(Pacc, mono, (0, nv, 0, nv), ops.sub(T1, 0, nv, ng, ng + nv), w_num)
                if ng:
                    _block_acc(Yacc, mono, (0, nv, 0, ng), ops.sub(T1, 0, nv, 0, ng), w_num / 2)
                    # Co^T Vg  (body.py:298)
                    T2 = ops.gemm(Co, Vg, transA=True)
                    _block_acc(Yacc, mono, (0, nv, 0, ng), T2, w_num / 2)
        # aims: -(w/2) (Vo^T 1 caim + Co^T 1 aim)   (body.py:296-299)
        for poly, aim in ((V, cost.cross_aim), (C, cost.aim)):
            a_kind, a_val = _aim_entry(np.asarray(aim, dtype=object), i)
            if a_kind == "num" and a_val == 0.0:
                continue
            a_slot, a_num = (slots.slot(a_val), 1.0) if a_kind == "expr" else (None, a_val)
            for m, H in poly.terms.items():
                colsum = ops.gemm(ops.sub(H, 0, poly.rows, ng, ng + nv), ones, transA=True)
                _block_acc(
                    Yacc, _mono(w_slot, m, a_slot), (0, nv, ng, ng + 1), colsum, -w_num * a_num / 2
                )


def _segments(counts, K):
    """(row, term, nz_begin, nz_end) of the non-empty (row, term) pairs; counts int [rows, K]."""
    rows = counts.shape[0]
    flat = counts.reshape(-1)
    ends = np.cumsum(flat)
    begins = ends - flat
    keep = flat > 0
    rr, tt = np.divmod(np.arange(rows * K), K)
    return np.stack([rr[keep], tt[keep], begins[keep], ends[keep]], axis=1).astype(np.int32).reshape([-1, 4])


def _scan_qh_rows(DY):
    """Sparse rows of the stacked DY constants for the [q; h] kernel.

    Returns ``segs`` int32 [n_segs, 4] = (row, term, nz_begin, nz_end) sorted by
    row then term, ``val`` float64 [n_nz] and ``col`` int32 [n_nz].  Structural
    scan on the host (which entries are non-zero) -- values are copied, not
    computed."""
    K, rows, cols = DY.shape
    mask = DY != 0.0
    counts = mask.sum(axis=2).T  # [rows, K]
    # entries ordered by (row, term, column)
    by_row = np.ascontiguousarray(DY.transpose(1, 0, 2))
    row_idx, term_idx, col_idx = np.nonzero(np.ascontiguousarray(mask.transpose(1, 0, 2)))
    val = by_row[row_idx, term_idx, col_idx].astype(np.float64)
    ends = np.cumsum(counts.reshape(-1))
    begins = ends - counts.reshape(-1)
    keep = counts.reshape(-1) > 0
    rr, tt = np.divmod(np.arange(rows * K), K)
    segs = np.stack([rr[keep], tt[keep], begins[keep], ends[keep]], axis=1).astype(np.int32)
    return segs.reshape([-1, 4]), val, col_idx.astype(np.int32)


class PreviewPlan:
    """The rows of some named variables as a plan for the [q; h] kernel:
    ``value = sum_m c_m(s_b) * D_m @ [given_b; optim_b; 1]`` (Formulation.preview,
    body.py:209-219).  Shares the scalar slots of the assembly plan it is cut from."""

    def __init__(self, compiled, ops, names, picks=None):
        """``picks``: per name, the rows of that variable to keep (a list), or None for all --
        the receding-horizon loop needs ONE sample of every state (biped_mpc_loop.py:81-88)."""
        polys = [compiled.PM[name] for name in names]
        width = compiled.ng + compiled.nv
        self.names = list(names)
        picks = [None] * len(polys) if picks is None else list(picks)
        for p, rows in zip(polys, picks):
            if rows is not None and not all(0 <= r < p.rows for r in rows):
                raise IndexError("a picked row is outside the variable's {} rows".format(p.rows))
        self.rows = [p.rows if rows is None else len(rows) for p, rows in zip(polys, picks)]
        self.nv, self.ng, self.nc = sum(self.rows), width, 0  # the kernel's view of the sizes
        self.slots = compiled.slots
        monos = []
        for p in polys:
            for mono in p.terms:
                if mono not in monos:
                    monos.append(mono)
        monos = sorted(monos, key=lambda m: (len(m), m)) or [()]
        if any(len(m) > MAX_FACTORS for m in monos):
            raise NotImplementedError("a monomial exceeds the kernel limit of factors")
        self.monomials = monos
        self.Y_terms = list(range(len(monos)))
        self.P_terms, self.G_terms = [0], [0]
        self.DP = self.DG = None
        DY = np.zeros([len(monos), self.nv, width + 1])
        row0 = 0
        for p, rows, count in zip(polys, picks, self.rows):
            for mono, H in p.terms.items():
                block = ops.download(H)
                DY[monos.index(mono), row0 : row0 + count, :width] = block if rows is None else block[list(rows)]
            row0 += count
        self.DY = DY
        self.Y_segs, self.Y_val, self.Y_col = _scan_qh_rows(DY)

    def mono_table(self):
        table = -np.ones([max(1, len(self.monomials)), MAX_FACTORS], dtype=np.int32)
        for i, mono in enumerate(self.monomials):
            table[i, : len(mono)] = mono
        return table


class CscPattern:
    """Sparsity pattern shared by a batch, in scipy's CSC convention
    (``scipy.sparse.csc_matrix(dense)``: int32 ``indices`` sorted by row inside
    each column, int32 ``indptr``) -- what the reference's example hands to the
    solver (biped_mpc_loop.py:57-58).  Derived STRUCTURALLY: an entry belongs to
    the pattern when it is non-zero in any constant matrix of the affine plan, so
    it is the same for every instance; an instance whose monomials happen to
    cancel an entry exactly keeps an explicit 0.0 there (scipy would drop it)."""

    def __init__(self, mask, upper=False):
        rows, cols = mask.shape
        if upper:
            mask = np.triu(mask)
        col_idx, row_idx = np.nonzero(mask.T)  # column-major order, rows ascending
        self.shape = (rows, cols)
        self.indices = row_idx.astype(np.int32)
        self.indptr = np.concatenate([[0], np.cumsum(mask.sum(axis=0))]).astype(np.int32)
        self.nnz = int(self.indices.shape[0])
        self.flat = (row_idx.astype(np.int64) * cols + col_idx.astype(np.int64))  # gather index
        self.padded = self.nnz + (self.nnz % 2)  # even, for 16-byte stores


def csc_patterns(plan, ops, upper_P=False):
    """(P pattern, G pattern, compact constants DPc [K, padded], DGc [K, padded])."""
    out = []
    for D, upper in ((plan.DP, upper_P), (plan.DG, False)):
        host = ops.download3(D)
        pattern = CscPattern(np.any(host != 0.0, axis=0), upper)
        index = np.concatenate([pattern.flat, np.zeros(pattern.padded - pattern.nnz, dtype=np.int64)])
        out.append((pattern, ops.gather_flat(D, index)))
    (pP, cP), (pG, cG) = out
    return pP, pG, cP, cG


class BoxRows:
    """The rows of G that bound ONE optimisation variable with a coefficient that does not
    depend on the instance -- ``a * z_c <= h_r`` -- for solvers that take ``lb <= z <= ub``
    beside ``G z <= h`` (qpsolvers' ``lb`` / ``ub``).  Structural, like
    the CSC patterns: a row qualifies when, over ALL constant matrices of the plan, it has one
    non-zero column, and that entry lives in the constant term only.

    rows   : int32, the qualifying rows of G (ascending)
    rest   : int32, the other rows (what stays in ``G z <= h``)
    col_ptr, row, coef : CSR over the variables: variable c is bounded by the rows
             ``row[col_ptr[c]:col_ptr[c+1]]`` with the coefficients ``coef[...]``"""

    def __init__(self, plan, ops):
        nc, nv = plan.nc, plan.nv
        self.rows = np.zeros([0], dtype=np.int32)
        rows, cols, coefs = [], [], []
        if plan.DG is not None and nc:
            DG = ops.download3(plan.DG)  # [K, nc, nv]
            const = [t for t, mono in enumerate(plan.G_terms) if plan.monomials[mono] == ()]
            mask = DG != 0.0
            single = mask.any(axis=0).sum(axis=1) == 1
            varying = np.zeros([nc], dtype=bool)
            for t in range(DG.shape[0]):
                if t not in const:
                    varying |= mask[t].any(axis=1)
            for r in np.nonzero(single & ~varying)[0]:
                c = int(np.nonzero(mask[:, r, :].any(axis=0))[0][0])
                rows.append(int(r)), cols.append(c), coefs.append(float(sum(DG[t, r, c] for t in const)))
        order = np.argsort(np.asarray(cols, dtype=np.int64), kind="stable")
        self.rows = np.asarray(rows, dtype=np.int32)
        self.rest = np.setdiff1d(np.arange(nc, dtype=np.int32), self.rows).astype(np.int32)
        self.row = self.rows[order].astype(np.int32)
        self.coef = np.asarray(coefs, dtype=np.float64)[order]
        counts = np.bincount(np.asarray(cols, dtype=np.int64), minlength=nv) if cols else np.zeros([nv], dtype=np.int64)
        self.col_ptr = np.concatenate([[0], np.cumsum(counts)]).astype(np.int32)
        self._device = None


def compile_plan(form, ops, costs=None, limits=None):
    """Builds the AffinePlan of ``form`` with backend ``ops``.

    ``costs`` / ``limits`` restrict the plan to some goals / limits (used by
    ``generate_qp_cost`` / ``generate_qp_constraint``); default is everything,
    in the reference's order."""
    plan = AffinePlan()
    slots = SlotTable()
    ng, nv = form.given_len, form.optim_len
    PM = _preview_polys(form, ops, slots)

    limits = ordered_limits(form) if limits is None else list(limits)
    costs = list(form.goals.values()) if costs is None else list(costs)

    # ---- constraint rows -------------------------------------------------
    pieces, bounds = [], []
    for limit in limits:
        poly, bound = _limit_poly(form, ops, PM, limit, slots)
        plan.row_offsets.append(sum(plan.limit_rows))
        plan.limit_rows.append(poly.rows)
        pieces.append(poly)
        bounds.append(bound)
    nc = sum(plan.limit_rows)

    Gacc = Poly(ops, nc, nv)
    Pacc = Poly(ops, nv, nv)
    Yacc = Poly(ops, nv + nc, ng + 1)
    if nc:
        Yacc.terms[()] = ops.zeros(nv + nc, ng + 1)
    for row0, parts in zip(plan.row_offsets, bounds):  # the constant column of h
        for mono, col in parts:
            _block_acc(Yacc, mono, (nv + row0, nv + row0 + col.shape[0], ng, ng + 1),
                       ops.upload(np.array(col)), 1.0)
    for row0, poly in zip(plan.row_offsets, pieces):
        for mono, H in poly.terms.items():
            _block_acc(Gacc, mono, (row0, row0 + poly.rows, 0, nv),
                       ops.sub(H, 0, poly.rows, ng, ng + nv), 1.0)
            if ng:  # h = bound - Gg @ given  (body.py:262)
                _block_acc(Yacc, mono, (nv + row0, nv + row0 + poly.rows, 0, ng),
                           ops.sub(H, 0, poly.rows, 0, ng), -1.0)

    # ---- costs -----------------------------------------------------------
    for cost in costs:
        _cost_into(form, ops, PM, slots, cost, Pacc, Yacc)
    if not Pacc.terms:
        Pacc.terms[()] = ops.zeros(nv, nv)
    if not Gacc.terms:
        Gacc.terms[()] = ops.zeros(nc, nv)
    if not Yacc.terms:
        Yacc.terms[()] = ops.zeros(nv + nc, ng + 1)

    # ---- global monomial table and stacked constants ----------------------
    table = {}

    def index_of(mono):
        if len(mono) > MAX_FACTORS:
            raise NotImplementedError(
                "a monomial of {} per-instance factors exceeds the kernel limit {}".format(
                    len(mono), MAX_FACTORS
                )
            )
        if mono not in table:
            table[mono] = len(table)
        return table[mono]

    def finish(acc, what):
        monos = sorted(acc.terms, key=lambda m: (len(m), m))
        if len(monos) > MAX_TERMS and what != "Y":
            raise NotImplementedError(
                "{} needs {} monomials, the kernel limit is {}".format(what, len(monos), MAX_TERMS)
            )
        return [index_of(m) for m in monos], ops.stack([acc.terms[m] for m in monos])

    plan.P_terms, plan.DP = finish(Pacc, "P")
    plan.G_terms, plan.DG = finish(Gacc, "G")
    plan.Y_terms, plan.DY = finish(Yacc, "Y")
    plan.monomials = [m for m, _ in sorted(table.items(), key=lambda kv: kv[1])]
    # (structural scan where the constants live: only the non-zeros come back to the host)
    plan.Y_segs, plan.Y_val, plan.Y_col = (ops.scan_rows(plan.DY) if hasattr(ops, "scan_rows")
                                           else _scan_qh_rows(ops.download3(plan.DY)))

    plan.nv, plan.ng, plan.nc = nv, ng, nc
    plan.slots = slots
    plan.optim_ID = dict(form.optim_ID)
    plan.given_ID = dict(form.given_ID)
    plan.PM = PM
    return plan
