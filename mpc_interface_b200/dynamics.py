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
        if not isinstance(sizes, Iterable):
            sizes = [sizes]
        if len(names) != len(sizes):
            raise IndexError("'names' and 'sizes' must have the same amount of elements.")
        self.axes = _as_axes(axes)
        self.identify_domain(names)
        self.set_sizes(names, sizes)
        self.outputs = []
        self.definitions = {}
        self.make_definitions()
        self.time_variant = time_variant
        self._size_rule = _pick_rule(how_to_update_size, time_variant)

    def identify_domain(self, names):
        self.domain_ID = {}
        for axis in self.axes:
            for index, name in enumerate(names):
                self.domain_ID[name + axis] = index

    def set_sizes(self, names, sizes):
        self.domain = {name: sizes[ID] for name, ID in self.domain_ID.items()}
        self.all_variables = self.domain

    def update_sizes(self, **kargs):
        self._size_rule(self, **kargs)

    def make_definitions(self):
        for variable, size in self.domain.items():
            self.definitions[variable] = _identity_combo(variable, size)

    def update_definitions(self):
        for variable, size in self.domain.items():
            self.definitions[variable].matrices[0] = np.eye(size)
        for output in self.outputs:
            self.definitions[output].update(domVar=self)

    def update(self, **kargs):
        self.update_sizes(**kargs)
        self.update_definitions()


class ExtendedSystem(_WithOutputs):
    """Horizon-extended dynamics ``x = S x0 + sum_j U_j u_j``.

    ``S`` has shape [N, n, n] and ``U`` is a list of m arrays [N, p_u, n]
    (dynamics.py:143-159).  With one input an array is accepted for ``U``; with
    one state 2-D arrays are accepted."""

    _owner_key = "extSyst"

    def __init__(
        self,
        input_names,
        state_names,
        state_vector_name,
        S,
        U,
        axes=None,
        time_variant=False,
        how_to_update_matrices=None,
    ):
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
        S, U = use.extend_matrices(horizon_lenght, control_system.A, control_system.B)
        rule = None
        if control_system.time_variant:

            def rule(ext_syst, **kargs):
                source = kargs["control_system"]
                source.update_matrices(**kargs)
                S, U = use.extend_matrices(horizon_lenght, source.A, source.B)
                ext_syst.matrices = U + [S]

        return cls(
            control_system.input_names,
            control_system.state_names,
            state_vector_name,
            S,
            U,
            control_system.axes,
            control_system.time_variant,
            rule,
        )

    def identify_domain(self, input_names, state_names):
        base_domain = {name: i for i, name in enumerate(input_names)}
        base_domain[self.state_vector_name + "0"] = len(input_names)
        self.domain_ID = {}
        self.state_ID = {}
        for axis in self.axes:
            for name, ID in base_domain.items():
                self.domain_ID[name + axis] = ID
            for i, name in enumerate(state_names):
                self.state_ID[name + axis] = i

    def set_sizes(self):
        horizon = self.matrices[-1].shape[0]
        self.domain = {v: self.matrices[ID].shape[1] for v, ID in self.domain_ID.items()}
        self.all_variables = dict(self.domain)
        self.all_variables.update({state: horizon for state in self.state_ID})

    def update_sizes(self):
        if self.time_variant:
            for variable, ID in self.domain_ID.items():
                self.domain[variable] = self.matrices[ID].shape[1]
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
        for variable, size in self.domain.items():
            self.definitions[variable].matrices[0] = np.eye(size)
        for state, sID in self.state_ID.items():
            combo = self.definitions[state]
            for k, var in enumerate(combo.variables):
                combo.matrices[k] = self.matrices[self.domain_ID[var]][..., sID]
        for output in self.outputs:
            self.definitions[output].update(extSyst=self)

    def update_matrices(self, **kargs):
        self._matrix_rule(self, **kargs)

    def update(self, **kargs):
        if self.time_variant:
            self.update_matrices(**kargs)
            self.update_sizes()
            self.update_definitions()


class ControlSystem:
    """One-step system ``x+ = A x + B u`` with named inputs and states."""

    def __init__(
        self,
        input_names,
        state_names,
        A,
        B,
        axes=None,
        time_variant=False,
        how_to_update_matrices=None,
    ):
        self.state_names = state_names
        self.input_names = input_names
        self.A = A
        self.B = B
        self.axes = axes
        self.time_variant = time_variant
        self._matrix_rule = _pick_rule(how_to_update_matrices, time_variant)

    def update_matrices(self, **kargs):
        self._matrix_rule(self, **kargs)

    @staticmethod
    def check_system_parameters(system_name):
        return use.get_system_matrices(system_name)[2]

    @classmethod
    def from_name(
        cls, system_name, axes=None, time_variant=False, how_to_update_matrices=None, **kargs
    ):
        """A named LIP-family system, discretised exactly (dynamics.py:380-407)."""
        input_names, state_names = use.get_system_variables(system_name)
        get_A, get_B, parameters = use.get_system_matrices(system_name)
        try:
            A = get_A(**kargs)
            B = get_B(**kargs)
        except Exception:
            raise ValueError("This system formulation requires: " + str(parameters))
        self = cls(input_names, state_names, A, B, axes, time_variant, how_to_update_matrices)
        self.get_A = get_A
        self.get_B = get_B
        self.parameters = kargs
        self.system_name = system_name
        return self
