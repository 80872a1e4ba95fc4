"""Named linear combinations -- ``LineCombo`` of python/mpc_interface/combinations.py.

``LineCombo({"CoM_x": 1, "CoM_ddot_x": -1 / w**2})`` states
``value = 1*CoM_x + (-1/w^2)*CoM_ddot_x``.  A coefficient is a Python scalar, a
1-D / 2-D array (combinations.py:4-12) or -- new here -- a per-instance ``Param``
expression (params.py).

The two parallel lists ``variables`` and ``matrices`` are the state of the object and are
public: formulations edit them in place (``combo.matrices[i] = ...`` inside update rules,
use_examples/.../biped_formulation.py), and their order is the order in which body.py
accumulates the preview matrices (body.py:179-193), so it is kept as given.
"""
from . import tools as use


class LineCombo:
    def __init__(self, combination=None, data=None, how_to_update=None, time_variant=False):
        terms = list(combination.items()) if combination is not None else []
        self.variables = [name for name, _ in terms]
        self.matrices = [coefficient for _, coefficient in terms]
        self.data = data if data is not None else []
        self.time_variant = time_variant
        # an update rule only counts for a time-variant combination (combinations.py:17-25)
        self._rule = how_to_update if (time_variant and how_to_update is not None) else use.do_not_update
        self._coefficients = ["C%d" % position for position in range(len(terms))]

    def update(self, **kargs):
        """Hands the combination and the tick's arguments to its update rule."""
        self._rule(self, **kargs)

    # -- read access by variable name, dict style ---------------------------------
    def __getitem__(self, variable):
        position = self.variables.index(variable)  # ValueError for an unknown name, as a list gives
        return self.matrices[position]

    def keys(self):
        return self.variables

    def values(self):
        return self.matrices

    def items(self):
        return zip(self.keys(), self.values())

    def __repr__(self):
        return " + ".join("%s ( %s )" % pair for pair in zip(self._coefficients, self.variables))

    __str__ = __repr__
