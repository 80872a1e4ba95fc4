"""NumPy restatement of the reference's QP-assembly path (TEST INFRASTRUCTURE).

Every function restates one reference routine (cited as file:line relative to
the reference checkout) with the same operations in the same association
order, so that it (1) reproduces the reference's float64 results and (2) costs
about what the reference costs on a CPU -- it is also the timed CPU baseline.

The functions are duck-typed: ``form`` may be the reference's own
``mpc_interface.body.Formulation`` or this repo's host mirror; only plain
attributes are read (``definitions``, ``of``, ``dynamics``, ``goals``,
``constraints``, ``constraint_boxes``, ``given_ID`` ...).  Methods of the
objects (``limit.matrices()``, ``limit.bound()`` ...) are deliberately NOT
called; their arithmetic is restated here from the object fields.

A LineCombo coefficient may be a per-instance parameter expression of the host
mirror (anything with ``.evaluate(params)``); the oracle evaluates it to a
plain float / ndarray for ONE instance before use.
"""
import numpy as np


# --------------------------------------------------------------------------
# prediction matrices                                    tools.py:14-33
# --------------------------------------------------------------------------
def extend_matrices(N, A, B):
    """S (N,n,n), U list[m] of (N,N,n) with S[k,c,i] = (A^(k+1))[i,c] and
    U_j[k,p,i] = (A^(k-p) B)[i,j] for p<=k (0 above the diagonal).

    Same recursion as python/mpc_interface/tools.py:23-29 (left multiplication
    by A once per horizon step), written on the impulse response instead of on
    the stacked working arrays."""
    A = np.asarray(A, dtype=float)
    B = np.asarray(B, dtype=float)
    n, m = B.shape
    S = np.zeros((N, n, n))
    U = [np.zeros((N, N, n)) for _ in range(m)]
    Ak = A.copy()       # A^(k+1)
    blk = np.zeros((n, N, m))    # block row k of the convolution: blk[:,p,:] = A^(k-p) B
    blk[:, 0, :] = B
    for k in range(N):
        if k:
            for j in range(m):
                blk[:, 0:k, j] = A.dot(blk[:, 0:k, j])   # tools.py:24-26
            blk[:, k, :] = B                              # tools.py:27
            Ak = A.dot(Ak)                                # tools.py:29
        S[k] = Ak.T                                       # tools.py:31
        for j in range(m):
            U[j][k] = blk[:, :, j].T                      # tools.py:32
    return S, U


def plan_steps(N, count=0, step_times=None, regular_time=None, phase=0):
    """0/1 step-indicator matrix E[k,j] = (count+k) > t_j   (tools.py:79-101)."""
    t = count + np.arange(N)
    if step_times is not None:
        step_times = np.asarray(step_times)
        nxt = step_times[(step_times >= count) & (step_times < count + N - 1)]
    elif regular_time is not None:
        nxt = np.array([x for x in t
                        if (x + phase) % regular_time == 0 and x < count + N - 1])
    else:
        raise KeyError("step_times or regular_time needed")
    return (t[:, None] > nxt).astype(int)


# --------------------------------------------------------------------------
# preview matrices                                       body.py:149-193
# --------------------------------------------------------------------------
def _value(coef, params):
    if hasattr(coef, "evaluate"):
        return coef.evaluate(params or {})
    return coef


def preview_matrices(form, params=None):
    """PM[v] = (Mg, Mo) with v = Mg*given + Mo*optim, for every definition in
    insertion order (body.py:149-156)."""
    PM = {}
    ng, nv = form.given_len, form.optim_len
    for name, combo in form.definitions.items():
        if name in form.of:                               # body.py:158-177
            rows = form.dynamics[form.of[name]].all_variables[name]
            Mg = np.zeros((rows, ng))
            Mo = np.zeros((rows, nv))
            for var, mat in combo.items():
                mat = _value(mat, params)
                if var in form.given_variables:
                    Mg[:, form.given_ID[var]] = mat
                elif var in form.optim_variables:
                    Mo[:, form.optim_ID[var]] = mat
                else:
                    raise ValueError(var + " is neither given nor optim")
        else:                                             # body.py:179-193
            Mg = Mo = None
            for var, coef in combo.items():
                c = np.array(_value(coef, params))
                tg, to = c.dot(PM[var][0]), c.dot(PM[var][1])
                if Mg is None:
                    Mg, Mo = tg, to
                    if Mg.ndim == 1:
                        Mg, Mo = Mg[None, :], Mo[None, :]
                else:
                    Mg += tg
                    Mo += to
        PM[name] = (Mg, Mo)
    return PM


def arrange_given(form, collector):
    """body.py:195-207"""
    g = np.zeros((form.given_len, 1))
    for var, idx in form.given_ID.items():
        g[idx] = collector[var]
    return g


# --------------------------------------------------------------------------
# constraint rows                     body.py:236-264, restrictions.py:147-199
# --------------------------------------------------------------------------
def _nlines(limit):
    """restrictions.py:147-162"""
    if limit.L:
        return limit.L[0].shape[0]
    if limit.schedule:
        return limit.schedule.stop - limit.schedule.start
    sizes = np.array([limit.arrow.shape[0], limit.center.shape[0], limit.extreme.shape[0]])
    if np.all(sizes == 1):
        return None
    return sizes[sizes != 1][0]


def _numbers(x, params):
    """Geometry of ONE instance as float64: entries that are per-instance
    expressions of the host mirror are evaluated with ``params``."""
    arr = np.asarray(x)
    if arr.dtype != object:
        return arr
    out = np.empty(arr.shape)
    for idx, e in np.ndenumerate(arr):
        out[idx] = _value(e, params)
    return out


def qp_constraint(limit, PM, given, params=None):
    """(A, h) of one limit: A = sum_a c_a (*|@) Mo[sched], h = bound - cMg@given."""
    arrow, center = _numbers(limit.arrow, params), _numbers(limit.center, params)
    extreme = _numbers(limit.extreme, params)
    ng = PM[limit.variable + limit.axes[0]][0].shape[1]
    nv = PM[limit.variable + limit.axes[0]][1].shape[1]
    rows = PM[limit.variable + limit.axes[0]][0].shape[0]
    nl = _nlines(limit)
    c_rows = rows if nl is None else nl
    cMg = np.zeros([c_rows, ng])
    cMo = np.zeros([c_rows, nv])
    sched = range(rows) if not limit.schedule else limit.schedule
    # restrictions.py:198-199
    bound = extreme + np.sum(arrow * center, axis=1).reshape([-1, 1])
    for i, axis in enumerate(limit.axes):
        Mg, Mo = PM[limit.variable + axis]
        c = arrow[:, i][:, None]                          # restrictions.py:175-179
        if limit.L:
            c = c * limit.L[i]
            cMg += c @ Mg[sched]
            cMo += c @ Mo[sched]
        else:
            cMg += c * Mg[sched]
            cMo += c * Mo[sched]
    return cMo, bound - cMg @ given


def all_limits(form):
    """body.py:306-318: constraints dict first, then the boxes."""
    out = [lim for area in form.constraints.values() for lim in area]
    out += [lim for box in form.constraint_boxes.values() for lim in box.constraints]
    return out


# --------------------------------------------------------------------------
# cost                                                   body.py:266-302
# --------------------------------------------------------------------------
def _number(x, params):
    x = _value(x, params)
    return x


def qp_cost(cost, PM, given, params=None):
    first = PM[cost.variable + cost.axes[0]]
    nv = first[1].shape[1]
    Q = np.zeros([nv, nv])
    q = np.zeros([nv, 1])
    rows = first[0].shape[0]
    sched = range(rows) if not cost.schedule else cost.schedule
    w = _number(cost.weight, params)
    for i, axis in enumerate(cost.axes):
        vMg, vMo = PM[cost.variable + axis]
        cMg, cMo = PM[cost.cross + axis]
        vMg, vMo, cMg, cMo = vMg[sched], vMo[sched], cMg[sched], cMo[sched]
        if cost.L:
            vMg, vMo = cost.L[i] @ vMg, cost.L[i] @ vMo
        if cost.cross_L:
            cMg, cMo = cost.cross_L[i] @ cMg, cost.cross_L[i] @ cMo
        aim = np.array([_number(a, params) for a in np.asarray(cost.aim, dtype=object)[:, i]], dtype=float)
        caim = np.array([_number(a, params) for a in np.asarray(cost.cross_aim, dtype=object)[:, i]], dtype=float)
        Q += w * vMo.T @ cMo                              # body.py:292
        q += w * (vMo.T @ (cMg @ given - caim) + cMo.T @ (vMg @ given - aim)) / 2
    return Q, q


# --------------------------------------------------------------------------
# entry point                                            body.py:304-348
# --------------------------------------------------------------------------
def all_qp_matrices(form, given, params=None, PM=None):
    """(A, h, Q, q) of one instance; ``given`` is a (ng, 1) column."""
    if PM is None:
        PM = preview_matrices(form, params)
    parts = [qp_constraint(lim, PM, given, params) for lim in all_limits(form)]
    A = np.vstack([p[0] for p in parts])
    h = np.vstack([p[1] for p in parts])
    costs = [qp_cost(c, PM, given, params) for c in form.goals.values()]
    Q = np.add.reduce([c[0] for c in costs])
    q = np.add.reduce([c[1] for c in costs])
    return A, h, Q, q
