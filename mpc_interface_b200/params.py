"""Per-instance scalar parameters for batched formulations.

The reference (Gepetto/mpc-interface) bakes every coefficient of a
``LineCombo`` / ``Cost`` into a Python float, so one ``Formulation`` describes
ONE instance (e.g. ``-1 / w**2`` in
python/use_examples/simple_functional_example/biped_formulation.py:62).  A
batch of independent instances shares the structure but not those numbers.
``Param("omega")`` is a named per-instance scalar; ordinary arithmetic on it
builds a small expression tree, so the same user code

    LineCombo({"CoM_x": 1, "CoM_ddot_x": -1 / w**2})

works with ``w`` a float (one instance, as in the reference) or a ``Param``
(one value per instance of the batch).  Every distinct expression becomes one
*scalar slot* of the device plan; the host evaluates the slots vectorised over
the batch (input preparation) and the kernels form the monomials.
"""
import numbers

import numpy as np


class Expr:
    """A scalar expression over named per-instance parameters."""

    # make ``ndarray (op) Expr`` defer to the reflected Expr operator
    __array_ufunc__ = None
    __slots__ = ("op", "args")

    def __init__(self, op, args):
        self.op = op
        self.args = tuple(args)

    # -- structure ---------------------------------------------------------
    def key(self):
        """Hashable structural identity (two equal keys -> one scalar slot)."""
        return (self.op,) + tuple(
            a.key() if isinstance(a, Expr) else ("const", float(a)) for a in self.args
        )

    def names(self):
        found = []
        for a in self.args:
            if isinstance(a, Expr):
                for n in a.names():
                    if n not in found:
                        found.append(n)
        return found

    def __hash__(self):
        return hash(self.key())

    def __eq__(self, other):
        return isinstance(other, Expr) and self.key() == other.key()

    # -- evaluation --------------------------------------------------------
    def evaluate(self, values):
        """Value for ``values = {name: float | ndarray[B]}`` (float64)."""
        v = [a.evaluate(values) if isinstance(a, Expr) else a for a in self.args]
        if self.op == "add":
            return v[0] + v[1]
        if self.op == "sub":
            return v[0] - v[1]
        if self.op == "mul":
            return v[0] * v[1]
        if self.op == "div":
            return v[0] / v[1]
        if self.op == "pow":
            return v[0] ** v[1]
        if self.op == "neg":
            return -v[0]
        if self.op in ("sqrt", "cos", "sin"):  # ndarray / float, or a tensor (its own method)
            own = getattr(v[0], self.op, None)
            return own() if callable(own) else getattr(np, self.op)(v[0])
        raise ValueError("unknown operation " + self.op)

    # -- polynomial form ---------------------------------------------------
    def polynomial(self):
        """``{atoms: coefficient}`` with ``self == sum coefficient * prod atoms``.

        ``atoms`` is a tuple of Expr (sorted by structural key, repeated for
        powers); an atom is a Param or a sub-expression that is not polynomial
        (division by an expression, sqrt, cos, sin, non-integer powers).  Used
        where an ARRAY of expressions multiplies matrices (constraint geometry,
        restrictions.py): entries that share atoms then share one constant
        matrix of the assembly plan instead of costing one each."""
        op = self.op
        if op in ("add", "sub"):
            out = dict(_poly(self.args[0]))
            sign = 1.0 if op == "add" else -1.0
            for mono, c in _poly(self.args[1]).items():
                out[mono] = out.get(mono, 0.0) + sign * c
            return out
        if op == "neg":
            return {mono: -c for mono, c in _poly(self.args[0]).items()}
        if op == "mul":
            return _poly_mul(_poly(self.args[0]), _poly(self.args[1]))
        if op == "div" and not isinstance(self.args[1], Expr):
            return {mono: c / float(self.args[1]) for mono, c in _poly(self.args[0]).items()}
        if op == "div":
            inverse = {(Expr("div", (1.0, self.args[1])),): 1.0}
            return _poly_mul(_poly(self.args[0]), inverse)
        if op == "pow" and isinstance(self.args[0], Expr) and not isinstance(self.args[1], Expr):
            e = float(self.args[1])
            if e == int(e) and 0 <= e <= 4:
                out = {(): 1.0}
                for _ in range(int(e)):
                    out = _poly_mul(out, _poly(self.args[0]))
                return out
        return {(self,): 1.0}

    # -- arithmetic --------------------------------------------------------
    @staticmethod
    def _ok(x):
        return isinstance(x, (Expr, numbers.Real))

    def _bin(self, op, other, swap=False):
        if isinstance(other, np.ndarray):  # array (op) expression: element by element
            out = np.empty(other.shape, dtype=object)
            for idx, x in np.ndenumerate(other):
                out[idx] = self._bin(op, x, swap)
            return out
        if not Expr._ok(other):
            return NotImplemented
        return Expr(op, (other, self) if swap else (self, other))

    def __add__(self, o):
        return self._bin("add", o)

    def __radd__(self, o):
        return self._bin("add", o, True)

    def __sub__(self, o):
        return self._bin("sub", o)

    def __rsub__(self, o):
        return self._bin("sub", o, True)

    def __mul__(self, o):
        return self._bin("mul", o)

    def __rmul__(self, o):
        return self._bin("mul", o, True)

    def __truediv__(self, o):
        return self._bin("div", o)

    def __rtruediv__(self, o):
        return self._bin("div", o, True)

    def __pow__(self, o):
        return self._bin("pow", o)

    def __neg__(self):
        return Expr("neg", (self,))

    def __pos__(self):
        return self

    def sqrt(self):
        return Expr("sqrt", (self,))

    # numpy calls these on the elements of object arrays (np.cos(array_of_Expr))
    def cos(self):
        return Expr("cos", (self,))

    def sin(self):
        return Expr("sin", (self,))

    def __repr__(self):
        if self.op == "neg":
            return "(-{!r})".format(self.args[0])
        if self.op in ("sqrt", "cos", "sin"):
            return "{}({!r})".format(self.op, self.args[0])
        sym = {"add": "+", "sub": "-", "mul": "*", "div": "/", "pow": "**"}[self.op]
        return "({!r} {} {!r})".format(self.args[0], sym, self.args[1])


class Param(Expr):
    """Leaf: one named float64 value per instance of the batch."""

    __slots__ = ("name",)

    def __init__(self, name):
        if not isinstance(name, str):
            raise TypeError("a parameter name must be a string")
        Expr.__init__(self, "param", ())
        self.name = name

    def key(self):
        return ("param", self.name)

    def names(self):
        return [self.name]

    def evaluate(self, values):
        if self.name not in values:
            raise KeyError("no value given for the parameter '{}'".format(self.name))
        return values[self.name]

    def __hash__(self):
        return hash(self.key())

    def __eq__(self, other):
        return isinstance(other, Expr) and self.key() == other.key()

    def __repr__(self):
        return self.name


def _poly(x):
    return x.polynomial() if isinstance(x, Expr) else {(): float(x)}


def _poly_mul(a, b):
    out = {}
    for m1, c1 in a.items():
        for m2, c2 in b.items():
            mono = tuple(sorted(m1 + m2, key=lambda e: repr(e.key())))
            out[mono] = out.get(mono, 0.0) + c1 * c2
    return out


def cos(x):
    """cos of a float or of a per-instance expression."""
    return x.cos() if isinstance(x, Expr) else np.cos(x)


def sin(x):
    return x.sin() if isinstance(x, Expr) else np.sin(x)


def is_expr(x):
    return isinstance(x, Expr)


def evaluate_array(x, values):
    """float64 copy of an array whose entries may be expressions (ONE instance:
    every parameter value a float)."""
    arr = np.asarray(x)
    if arr.dtype != object:
        return arr.astype(np.float64)
    out = np.empty(arr.shape, dtype=np.float64)
    for idx, e in np.ndenumerate(arr):
        out[idx] = e.evaluate(values) if isinstance(e, Expr) else e
    return out


def split_array(x):
    """An array of numbers / expressions as ``[(atoms, float64 array)]`` with
    ``x == sum prod(atoms) * array`` (atoms as in Expr.polynomial); purely numeric
    arrays come back as one part with ``atoms == ()``."""
    arr = np.asarray(x)
    if arr.dtype != object:
        return [((), arr.astype(np.float64))]
    parts = {}
    for idx, e in np.ndenumerate(arr):
        for mono, c in _poly(e).items():
            if c == 0.0:
                continue
            if mono not in parts:
                parts[mono] = np.zeros(arr.shape)
            parts[mono][idx] += c
    if not parts:
        return [((), np.zeros(arr.shape))]
    order = sorted(parts, key=lambda m: (len(m), [repr(a.key()) for a in m]))
    return [(m, parts[m]) for m in order]


def contains_expr(x):
    """True for an Expr or a (nested) list / object array holding one."""
    if isinstance(x, Expr):
        return True
    if isinstance(x, np.ndarray):
        return x.dtype == object and any(contains_expr(e) for e in x.flat)
    if isinstance(x, (list, tuple)):
        return any(contains_expr(e) for e in x)
    return False
