"""Quadratic cost descriptor (mirror of python/mpc_interface/goal.py:11-161).

``Cost(variable, weight, aim, axes, L, schedule, cross, cross_aim, cross_L)``
asks for ``V = L @ v[schedule]`` to approach ``aim`` with some weight; with
``cross`` the product ``(V - aim) * (C - cross_aim)`` is driven to zero.
New here: ``weight`` and the entries of ``aim`` / ``cross_aim`` may be
per-instance ``Param`` expressions (params.py); they then become scalar slots
of the batched plan instead of constants.
"""
import numpy as np

from .params import contains_expr


def _aim_array(aim, n_axes):
    """[m, len(axes)] array; object dtype when it holds Param expressions."""
    if contains_expr(aim):
        return np.ravel(np.array(aim, dtype=object)).reshape([-1, n_axes])
    return np.array(aim).reshape([-1, n_axes])


class Cost:
    def __init__(
        self,
        variable,
        weight,
        aim=None,
        axes=None,
        L=None,
        schedule=None,
        cross=None,
        cross_aim=None,
        cross_L=None,
    ):
        self.axes = axes if axes else [""]
        if not isinstance(self.axes, list):
            raise TypeError("The axes must be a list of strings")
        self.axes_len = len(self.axes)
        self.variable = variable
        self.weight = weight
        self.aim = _aim_array(np.zeros([1, self.axes_len]) if aim is None else aim, self.axes_len)
        self.schedule = schedule if schedule else range(0)
        self.L = [] if L is None else self.arrange_L(L)

        self.crossed = bool(cross)
        if self.crossed:
            self.cross = cross
            # goal.py:78 tests the truth of cross_aim itself
            given = cross_aim if contains_expr(cross_aim) else (cross_aim if cross_aim else None)
            self.cross_aim = _aim_array(
                np.zeros([1, self.axes_len]) if given is None else given, self.axes_len
            )
            self.cross_L = [] if cross_L is None else self.arrange_L(cross_L)
        else:
            # a plain cost is crossed with itself: the fields are aliases (goal.py:82-86)
            self.cross = self.variable
            self.cross_aim = self.aim
            self.cross_L = self.L

    @property
    def t(self):
        if self.schedule:
            return self.schedule.stop - self.schedule.start
        return None

    def arrange_L(self, L):
        """One matrix per axis; needs an up-to-date schedule (goal.py:88-104)."""
        matrices = L if isinstance(L, list) else [L]
        if len(matrices) == 1:
            matrices = matrices * self.axes_len
        elif len(matrices) not in (self.axes_len, 0):
            raise IndexError(
                "'L' must have 0, 1 or len(axes) = {} elements".format(self.axes_len)
            )
        if self.schedule and any(mat.shape[-1] != self.t for mat in matrices):
            raise ValueError(
                "arrays in L must have {} columns, which is given by the 'schedule'.".format(
                    self.t
                )
            )
        return matrices

    def update(self, aim=None, weight=None, L=None, schedule=None, cross_aim=None, cross_L=None):
        self._stamp = getattr(self, "_stamp", 0) + 1  # invalidates compiled plans
        if schedule is not None:
            self.schedule = schedule
        if L is not None:
            self.L = self.arrange_L(L)
            if not self.crossed:
                self.cross_L = self.L
        if cross_L is not None:
            if not self.crossed:
                raise KeyError("Trying to set cross_aim in a non-crossed cost")
            self.cross_L = self.arrange_L(cross_L)
        if aim is not None:
            self.aim = _aim_array(aim, self.axes_len)
            if not self.crossed:
                self.cross_aim = self.aim
        if weight is not None:
            self.weight = weight
        if cross_aim is not None:
            if not self.crossed:
                raise KeyError("Trying to set cross_aim in a non-crossed cost")
            self.cross_aim = cross_aim

    def __repr__(self):
        lines = []
        for i, axis in enumerate(self.axes):
            name = self.variable + axis
            for j, target in enumerate(self.aim[:, i]):
                if j == 0:
                    lines.append(
                        "\n{} --> {}\t\t~ weight = {} ~".format(name, target, self.weight)
                    )
                else:
                    lines.append("{} --> {}".format(" " * len(name), target))
        return "\n".join(lines) + "\n\n"

    __str__ = __repr__
