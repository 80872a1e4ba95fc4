"""Dynamics objects of the formulation API (host bookkeeping).

Mirror of python/mpc_interface/dynamics.py: ``DomainVariable`` (:29-123),
``ExtendedSystem`` (:131-344) and ``ControlSystem`` (:347-407) with the same
constructor arguments, attributes (``domain``, ``domain_ID``, ``state_ID``,
``all_variables``, ``definitions``, ``matrices``, ``outputs``) and update
protocol.  These classes only keep names, sizes and views of the prediction
matrices; the matrices themselves come from the CUDA kernel behind
``tools.extend_matrices``.
"""
from collections.abc import Iterable

import numpy as np

from . import tools as use
from .combinations import LineCombo


def _as_name_list(names, what):
    names = [names] if isinstance(names, str) else names
    for name in names:
        if not isinstance(name, str):
            raise TypeError("all {} names must be strings.".format(what))
    return names


def _as_axes(axes):
    if axes is None:
        return [""]
    if isinstance(axes, str):
        return [axes]
    if not isinstance(axes, Iterable):
        return [str(axes)]
    return axes


def _pick_rule(rule, time_variant):
    return rule if (rule is not None and time_variant) else use.do_not_update


def _identity_combo(variable, size):
    combo = LineCombo({variable: np.eye(size)})
    combo._coefficients = ["I"]
    return combo


class _WithOutputs:
    """Shared by DomainVariable and ExtendedSystem: user-defined outputs."""

    _owner_key = None  # name of the karg the output's update rule receives

    def define_output(self, name, combination, time_variant=False, how_to_update=None):
        """Adds ``name+axis = sum coef * (var+axis)`` for every axis
        (dynamics.py:92-113 and :300-322)."""
        for axis in self.axes:
            combo = LineCombo(
                {var + axis: coef for var, coef in combination.items()},
                time_variant=time_variant,
                how_to_update=how_to_update,
            )
            self.definitions[name + axis] = combo
            if time_variant and self._owner_key == "extSyst":
                combo.update(extSyst=self)  # dynamics.py:320-321
            self.outputs.append(name + axis)


class DomainVariable(_WithOutputs):
    """Free named vectors of the QP domain, e.g. a bias of N samples per axis."""

    _owner_key = "domVar"

    def __init__(self, names, sizes, axes=None, time_variant=False, how_to_update_size=None):
        names = _as_name_list(names, "variable")
        sizes = list(sizes) if isinstance(sizes, Iterable) else [sizes]
        if len(sizes) != len(names):
            raise IndexError("one size per name is needed: {} names, {} sizes".format(len(names), len(sizes)))
        self.axes = _as_axes(axes)
        self.time_variant = time_variant
        self._size_rule = _pick_rule(how_to_update_size, time_variant)
        self.outputs, self.definitions = [], {}
        self.identify_domain(names)
        self.set_sizes(names, sizes)
        self.make_definitions()

    def identify_domain(self, names):
        """``domain_ID[name + axis]`` = position of ``name`` in ``names`` (dynamics.py:71-77)."""
        self.domain_ID = {name + axis: position
                          for axis in self.axes for position, name in enumerate(names)}

    def set_sizes(self, names, sizes):
        """``domain`` maps every variable to its length; ``all_variables`` is the same dict
        object, as in the reference (dynamics.py:79-81)."""
        self.domain = {}
        for variable, position in self.domain_ID.items():
            self.domain[variable] = sizes[position]
        self.all_variables = self.domain

    def make_definitions(self):
        self.definitions.update(
            (variable, _identity_combo(variable, length)) for variable, length in self.domain.items())

    def update_sizes(self, **kargs):
        self._size_rule(self, **kargs)

    def update_definitions(self):
        """Identities of the (possibly new) sizes, then the outputs' own rules."""
        for variable, length in self.domain.items():
            self.definitions[variable].matrices[0] = np.eye(length)
        for name in self.outputs:
            self.definitions[name].update(domVar=self)

    def update(self, **kargs):
        """One tick: sizes first, then everything that depends on them."""
        self.update_sizes(**kargs)
        self.update_definitions()


class ExtendedSystem(_WithOutputs):
    """Horizon-extended dynamics ``x = S x0 + sum_j U_j u_j``.

    ``S`` has shape [N, n, n] and ``U`` is a list of m arrays [N, p_u, n]
    (dynamics.py:143-159).  With one input an array is accepted for ``U``; with
    one state 2-D arrays are accepted."""

    _owner_key = "extSyst"

    def __init__(self, input_names, state_names, state_vector_name, S, U, axes=None,
                 time_variant=False, how_to_update_matrices=None):
        input_names = _as_name_list(input_names, "input")
        state_names = _as_name_list(state_names, "variable")
        if not isinstance(state_vector_name, str):
            raise TypeError("the 'state_vector_name' must be a single string.")
        if not isinstance(U, list) and len(input_names) == 1:
            U = [U]
        if len(state_names) == 1:
            if S.ndim == 2:
                S = S[:, :, None]
            U = [u[:, :, None] if u.ndim == 2 else u for u in U]

        self.axes = _as_axes(axes)
        self.matrices = list(U) + [S]  # inputs first, x0 last (dynamics.py:199)
        self.state_vector_name = state_vector_name
        self.identify_domain(input_names, state_names)
        self.set_sizes()
        self.outputs = []
        self.definitions = {}
        self.make_definitions()
        self.time_variant = time_variant
        self._matrix_rule = _pick_rule(how_to_update_matrices, time_variant)

    @classmethod
    def from_cotrol_system(cls, control_system, state_vector_name, horizon_lenght):
        """Extends a ControlSystem over ``horizon_lenght`` samples
        (dynamics.py:216-246; the spelling is the reference's API)."""
        def extend(system):
            S, U = use.extend_matrices(horizon_lenght, system.A, system.B)
            return S, U

        def follow_the_control_system(extended, **kargs):
            # the tick hands the one-step system over as ``control_system=``; it refreshes
            # its own (A, B) first, then the horizon matrices are rebuilt from them
            system = kargs["control_system"]
            system.update_matrices(**kargs)
            S, U = extend(system)
            extended.matrices = U + [S]

        S, U = extend(control_system)
        varies = control_system.time_variant
        return cls(input_names=control_system.input_names, state_names=control_system.state_names,
                   state_vector_name=state_vector_name, S=S, U=U, axes=control_system.axes,
                   time_variant=varies,
                   how_to_update_matrices=follow_the_control_system if varies else None)

    def identify_domain(self, input_names, state_names):
        """Positions in ``matrices``: input j at j, the initial state ``<vector>0`` after the
        inputs; ``state_ID`` numbers the states (dynamics.py:248-258)."""
        slots = list(input_names) + [self.state_vector_name + "0"]
        self.domain_ID, self.state_ID = {}, {}
        for axis in self.axes:
            self.domain_ID.update((name + axis, slot) for slot, name in enumerate(slots))
            self.state_ID.update((name + axis, row) for row, name in enumerate(state_names))

    def _domain_lengths(self):
        return {variable: self.matrices[slot].shape[1] for variable, slot in self.domain_ID.items()}

    def set_sizes(self):
        """Domain variables are as long as their matrix has columns (actions of an input,
        entries of x0); every state spans the horizon (dynamics.py:260-267)."""
        samples = self.matrices[-1].shape[0]
        self.domain = self._domain_lengths()
        self.all_variables = {**self.domain, **dict.fromkeys(self.state_ID, samples)}

    def update_sizes(self):
        if not self.time_variant:
            return
        self.domain.update(self._domain_lengths())
        self.all_variables.update(self.domain)

    def _same_axis(self, a, b):
        # the reference compares the last two characters (dynamics.py:293)
        return self.axes == [""] or a[-2:] == b[-2:]

    def make_definitions(self):
        for variable, size in self.domain.items():
            self.definitions[variable] = _identity_combo(variable, size)
        n_inputs = len(self.matrices) - 1
        for state, sID in self.state_ID.items():
            combo = LineCombo(
                {
                    base: self.matrices[dID][..., sID]
                    for base, dID in self.domain_ID.items()
                    if self._same_axis(base, state)
                }
            )
            combo._coefficients = n_inputs * ["U"] + ["S"]
            self.definitions[state] = combo

    def update_definitions(self):
        """Re-slices the state definitions out of the current matrices (in place: other
        objects hold these LineCombos), then lets the outputs follow."""
        for variable, length in self.domain.items():
            self.definitions[variable].matrices[0] = np.eye(length)
        for state, row in self.state_ID.items():
            combo = self.definitions[state]
            combo.matrices[:] = [self.matrices[self.domain_ID[name]][..., row] for name in combo.variables]
        for name in self.outputs:
            self.definitions[name].update(extSyst=self)

    def update_matrices(self, **kargs):
        self._matrix_rule(self, **kargs)

    def update(self, **kargs):
        """One tick of a time-variant system: matrices, sizes, definitions, in that order."""
        if not self.time_variant:
            return
        self.update_matrices(**kargs)
        self.update_sizes()
        self.update_definitions()


class ControlSystem:
    """One-step system ``x+ = A x + B u`` with named inputs and states."""

    def __init__(self, input_names, state_names, A, B, axes=None, time_variant=False,
                 how_to_update_matrices=None):
        self.input_names, self.state_names = input_names, state_names
        self.A, self.B = A, B
        self.axes = axes
        self.time_variant = time_variant
        self._matrix_rule = _pick_rule(how_to_update_matrices, time_variant)

    def update_matrices(self, **kargs):
        self._matrix_rule(self, **kargs)

    @staticmethod
    def check_system_parameters(system_name):
        return use.get_system_matrices(system_name)[2]

    @classmethod
    def from_name(cls, system_name, axes=None, time_variant=False, how_to_update_matrices=None,
                  **kargs):
        """A named LIP-family system, discretised exactly (dynamics.py:380-407).  ``kargs``
        are the physical parameters (``tau``, ``omega``); a missing one is a ValueError
        that lists what the system needs.  The matrix builders stay on the object
        (``get_A``, ``get_B``) for update rules that re-evaluate them."""
        builders = use.get_system_matrices(system_name)
        try:
            A, B = builders[0](**kargs), builders[1](**kargs)
        except Exception as problem:
            raise ValueError("the system '{}' takes the parameters {}".format(
                system_name, builders[2])) from problem
        inputs, states = use.get_system_variables(system_name)
        system = cls(inputs, states, A, B, axes=axes, time_variant=time_variant,
                     how_to_update_matrices=how_to_update_matrices)
        system.get_A, system.get_B = builders[0], builders[1]
        system.parameters, system.system_name = kargs, system_name
        return system
