"""Named linear combinations (mirror of python/mpc_interface/combinations.py).

``LineCombo({"CoM_x": 1, "CoM_ddot_x": -1 / w**2})`` states
``value = 1*CoM_x + (-1/w^2)*CoM_ddot_x``.  A coefficient is a Python scalar, a
1-D / 2-D array (combinations.py:4-12) or -- new here -- a per-instance
``Param`` expression (params.py).  Order of the variables is kept: it is the
order in which the preview matrices are accumulated (body.py:179-193).
"""
from . import tools as use


class LineCombo:
    def __init__(self, combination=None, data=None, how_to_update=None, time_variant=False):
        self.variables = []
        self.matrices = []
        if combination is not None:
            for variable, coefficient in combination.items():
                self.variables.append(variable)
                self.matrices.append(coefficient)
        self.data = [] if data is None else data
        self.time_variant = time_variant
        # the rule is only kept for time-variant combinations (combinations.py:17-25)
        self._rule = how_to_update if (how_to_update and time_variant) else use.do_not_update
        self._coefficients = ["C%d" % i for i in range(len(self.variables))]

    def update(self, **kargs):
        self._rule(self, **kargs)

    def __getitem__(self, variable):
        return self.matrices[self.variables.index(variable)]

    def items(self):
        return zip(self.variables, self.matrices)

    def keys(self):
        return self.variables

    def values(self):
        return self.matrices

    def __repr__(self):
        return " + ".join(
            "{} ( {} )".format(c, v) for c, v in zip(self._coefficients, self.variables)
        )

    __str__ = __repr__
