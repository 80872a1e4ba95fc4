"""Quadratic cost descriptor -- the ``Cost`` of python/mpc_interface/goal.py:11-161.

``Cost(variable, weight, aim, axes, L, schedule, cross, cross_aim, cross_L)`` asks for
``V = L @ v[schedule]`` to approach ``aim`` with some weight; with ``cross`` the product
``(V - aim) * (C - cross_aim)`` is driven to zero instead of the square.  The object only
describes the term; body.py turns it into blocks of P and q (body.py:266-302 of the
reference, plan.py here).

New here: ``weight`` and the entries of ``aim`` / ``cross_aim`` may be per-instance
``Param`` expressions (params.py); they then become scalar slots of the batched plan
instead of constants.

Every cost has the attributes ``variable, weight, aim, L, schedule, axes, axes_len,
crossed, cross, cross_aim, cross_L``.  A cost without ``cross`` is crossed with itself:
its three ``cross*`` attributes are the very objects ``variable / aim / L`` (the reference
aliases them, goal.py:82-86, and its tests rely on the identity), and trying to set them
separately is a KeyError.
"""
import numpy as np

from ._shared import axes_list, one_matrix_per_axis, window_length
from .params import contains_expr


def _aim_table(values, n_axes):
    """``values`` as an array with one column per axis (one row per aimed sample);
    object dtype when Param expressions sit in it."""
    if contains_expr(values):
        return np.ravel(np.array(values, dtype=object)).reshape([-1, n_axes])
    return np.array(values).reshape([-1, n_axes])


class Cost:
    def __init__(self, variable, weight, aim=None, axes=None, L=None, schedule=None,
                 cross=None, cross_aim=None, cross_L=None):
        self.axes = axes_list(axes, empty=True)
        self.axes_len = len(self.axes)
        self.variable, self.weight = variable, weight
        self.schedule = schedule or range(0)
        self.crossed = bool(cross)
        self.cross = cross if self.crossed else variable

        origin = np.zeros([1, self.axes_len])
        self._put("aim", _aim_table(origin if aim is None else aim, self.axes_len))
        self._put("L", [] if L is None else self.arrange_L(L))
        if self.crossed:
            # the reference tests the truth value of cross_aim (goal.py:78): 0 / [] / None
            # all mean "aim the crossed variable at the origin"
            unset = not contains_expr(cross_aim) and not cross_aim
            self.cross_aim = _aim_table(origin if unset else cross_aim, self.axes_len)
            self.cross_L = [] if cross_L is None else self.arrange_L(cross_L)

    # -- storage: own fields drag their cross twins along while the cost is not crossed
    def _put(self, field, value):
        setattr(self, field, value)
        if not self.crossed:
            setattr(self, "cross_" + field, value)

    def _put_cross(self, field, value):
        if not self.crossed:
            raise KeyError("Trying to set cross_{} in a non-crossed cost".format(field))
        setattr(self, "cross_" + field, value)

    @property
    def t(self):
        """Samples the schedule selects, None without a schedule (goal.py:138-142)."""
        return window_length(self.schedule)

    def arrange_L(self, L):
        """One matrix per axis, checked against the CURRENT schedule (goal.py:88-104)."""
        return one_matrix_per_axis(L, self.axes_len, self.t)

    def update(self, aim=None, weight=None, L=None, schedule=None, cross_aim=None, cross_L=None):
        """Replaces the given fields (goal.py:106-136).  The schedule goes first so that
        new L matrices are checked against it; ``cross_aim`` is stored as handed in, like
        the reference does."""
        self._stamp = getattr(self, "_stamp", 0) + 1  # compiled plans look at this
        if schedule is not None:
            self.schedule = schedule
        if L is not None:
            self._put("L", self.arrange_L(L))
        if cross_L is not None:
            self._put_cross("L", self.arrange_L(cross_L) if self.crossed else cross_L)
        if aim is not None:
            self._put("aim", _aim_table(aim, self.axes_len))
        if weight is not None:
            self.weight = weight
        if cross_aim is not None:
            self._put_cross("aim", cross_aim)

    def __repr__(self):
        rows = []
        for column, axis in enumerate(self.axes):
            targets = ", ".join(str(x) for x in self.aim[:, column])
            rows.append("{}{} --> {}".format(self.variable, axis, targets))
        tail = "   (weight {})".format(self.weight)
        if self.crossed:
            tail += "   crossed with {}".format(self.cross)
        return "\n".join(rows) + tail + "\n"

    __str__ = __repr__
