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
        if self.op == "sqrt":
            return np.sqrt(v[0])
        raise ValueError("unknown operation " + self.op)

    # -- arithmetic --------------------------------------------------------
    @staticmethod
    def _ok(x):
        return isinstance(x, (Expr, numbers.Real))

    def _bin(self, op, other, swap=False):
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

    def __repr__(self):
        if self.op == "neg":
            return "(-{!r})".format(self.args[0])
        if self.op == "sqrt":
            return "sqrt({!r})".format(self.args[0])
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


def is_expr(x):
    return isinstance(x, Expr)


def contains_expr(x):
    """True for an Expr or a (nested) list / object array holding one."""
    if isinstance(x, Expr):
        return True
    if isinstance(x, np.ndarray):
        return x.dtype == object and any(contains_expr(e) for e in x.flat)
    if isinstance(x, (list, tuple)):
        return any(contains_expr(e) for e in x)
    return False
