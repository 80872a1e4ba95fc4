"""The Formulation: named dynamics, definitions, costs and constraints -> QP.

Mirror of python/mpc_interface/body.py:14-348 with the same method names,
argument meaning, return order ``(A, h, Q, q)`` (= qpsolvers ``G, h, P, q``)
and error behaviour.  The bookkeeping (dict insertion orders, ``optim_ID`` /
``given_ID`` ranges) is host code because it defines the bit-exact index
layout; every matrix is produced on the GPU: the formulation is compiled into
an affine plan (plan.py) and evaluated by the streaming kernels behind
``mpci_assemble`` (include/mpci.h).  A single instance is a batch of one;
``generate_all_qp_matrices_batched`` is the batched entry point.
"""
import numpy as np

from . import tools as use
from . import plan as _plan


class _PreviewMatrices:
    """``form.PM[var] -> (Mg, Mo)`` like the reference's dict (body.py:149-156),
    downloaded lazily from the device-resident plan."""

    def __init__(self, form):
        self._form = form
        self._cache = {}

    def _polys(self):
        return self._form._compiled().PM

    def keys(self):
        return self._polys().keys()

    def __len__(self):
        return len(self._polys())

    def __iter__(self):
        return iter(self._polys())

    def __contains__(self, name):
        return name in self._polys()

    def __getitem__(self, name):
        if name not in self._cache:
            poly = self._polys()[name]
            if set(poly.terms) - {()}:
                raise ValueError(
                    "'{}' depends on per-instance parameters; it has no single "
                    "preview matrix".format(name)
                )
            ng = self._form.given_len
            ops = self._form._ops()
            dense = ops.download(poly.terms[()]) if () in poly.terms else np.zeros(
                [poly.rows, poly.cols]
            )
            self._cache[name] = (dense[:, :ng].copy(), dense[:, ng:].copy())
        return self._cache[name]

    def items(self):
        return ((k, self[k]) for k in self.keys())


class Formulation:
    def __init__(self, backend=None):
        self.domain = {}
        self.optim_variables = []
        self.given_variables = []
        self.optim_sizes = []
        self.given_sizes = []
        self.optim_len = 0
        self.given_len = 0
        self.optim_ID = {}
        self.given_ID = {}
        self.domain_ID = {"optim_ID": self.optim_ID, "given_ID": self.given_ID}

        self.definitions = {}
        self.dynamics = {}
        self.of = {}
        self.constraint_boxes = {}
        self.constraints = {}
        self.goals = {}
        self.update_incorporations = use.do_not_update

        self._backend = backend  # None -> the CUDA device backend (device.py)
        self._plans = {}

    # ------------------------------------------------------------------
    # incorporations (body.py:38-94)
    # ------------------------------------------------------------------
    def _dirty(self):
        self._plans = {}

    def _require_defined(self, names, message):
        for name in names:
            assert name in self.definitions, message.format(name)

    def incorporate_dynamics(self, name, new_dynamics):
        self.domain.update(new_dynamics.domain)
        self.dynamics[name] = new_dynamics
        for variable in new_dynamics.all_variables:
            self.of[variable] = name
        self.definitions.update(new_dynamics.definitions)
        self._dirty()

    def incorporate_definition(self, name, new_definition):
        self._require_defined(
            new_definition.keys(),
            "The definition must depend only on defined variables, but the "
            "this definition depends on {}",
        )
        self.definitions[name] = new_definition
        self._dirty()

    def incorporate_definitions(self, dict_of_defs):
        for name, combination in dict_of_defs.items():
            self.incorporate_definition(name, combination)

    def _require_limits_defined(self, limits):
        for limit in limits:
            self._require_defined(
                [limit.variable + axis for axis in limit.axes],
                "All constrained variables must be previously defined, "
                "but '{}' was not defined.",
            )

    def incorporate_constraint(self, name, new_limits):
        if not isinstance(new_limits, list):
            new_limits = [new_limits]
        self._require_limits_defined(new_limits)
        self.constraints[name] = new_limits
        self._dirty()

    def incorporate_box(self, name, new_box):
        self._require_limits_defined(new_box.constraints)
        self.constraint_boxes[name] = new_box
        self._dirty()

    def incorporate_goal(self, name, new_goal):
        self._require_defined(
            [new_goal.variable + axis for axis in new_goal.axes],
            "All constrained variables must be previously defined, "
            "but '{}' was not defined.",
        )
        self.goals[name] = new_goal
        self._dirty()

    # ------------------------------------------------------------------
    # index layout of the QP (body.py:97-136) -- integers, host side
    # ------------------------------------------------------------------
    def identify_qp_domain(self, optimization_domain):
        self.optim_variables = optimization_domain
        self.given_variables = [v for v in self.domain if v not in optimization_domain]
        self.update_qp_sizes()
        self.update_qp_IDs()

    def update_qp_domain(self):
        for behavior in self.dynamics.values():
            if behavior.time_variant:
                self.domain.update(behavior.domain)

    def update_qp_sizes(self):
        self.optim_sizes = [self.domain[v] for v in self.optim_variables]
        self.given_sizes = [self.domain[v] for v in self.given_variables]
        self.optim_len = sum(self.optim_sizes)
        self.given_len = sum(self.given_sizes)

    @staticmethod
    def _consecutive_ranges(names, sizes):
        out, start = {}, 0
        for name, size in zip(names, sizes):
            out[name] = range(start, start + size)
            start += size
        return out

    def update_qp_IDs(self):
        # only adds / overwrites keys, like body.py:135-136
        self.optim_ID.update(self._consecutive_ranges(self.optim_variables, self.optim_sizes))
        self.given_ID.update(self._consecutive_ranges(self.given_variables, self.given_sizes))
        self._dirty()

    def set_updating_rule(self, how_to_update=None):
        self.update_incorporations = how_to_update

    def update(self, **kargs):
        self.update_incorporations(self, **kargs)
        self.update_qp_domain()
        self.update_qp_sizes()
        self.update_qp_IDs()
        self.make_preview_matrices()

    # ------------------------------------------------------------------
    # device plan
    # ------------------------------------------------------------------
    def _ops(self):
        if self._backend is None:
            from . import device

            self._backend = device.DeviceBackend()
        return self._backend

    def _compiled(self, costs=None, limits=None):
        key = self._signature(
            list(self.goals.values()) if costs is None else costs,
            _plan.ordered_limits(self) if limits is None else limits,
        )
        if key not in self._plans:
            self._trim_plans()
            self._plans[key] = _plan.compile_plan(self, self._ops(), costs, limits)
        return self._plans[key]

    def _trim_plans(self, keep=32):
        """Plans (and their device constants) of costs / limits that were passed once and
        never again do not pile up: the oldest go first (dicts keep insertion order)."""
        while len(self._plans) >= keep:
            self._plans.pop(next(iter(self._plans)))

    def make_preview_matrices(self):
        """Rebuilds the preview matrices ``PM[var] = (Mg, Mo)`` with
        ``var = Mg @ given + Mo @ optim`` (body.py:149-156) -- on the device,
        as part of the plan -- after any change of the incorporated objects."""
        self._dirty()
        self.PM = _PreviewMatrices(self)
        self._compiled()

    def arrange_given(self, collector):
        """Packs named (k, 1) arrays into the given column (body.py:195-207)."""
        if not self.given_len:
            return np.array([])
        given = np.zeros([self.given_len, 1])
        for variable, indices in self.given_ID.items():
            given[indices] = collector[variable]
        return given

    def arrange_given_batched(self, collector):
        """Batched ``arrange_given``: ``{variable: array [B, size]}`` -> [B, given_len]
        in the order of ``given_ID`` (body.py:195-207 for every instance)."""
        batch = len(next(iter(collector.values()))) if collector else 0
        given = np.zeros([batch, self.given_len])
        for variable, indices in self.given_ID.items():
            if variable in self.given_variables:
                given[:, indices.start : indices.stop] = np.asarray(collector[variable]).reshape(
                    batch, len(indices))
        return given

    # ------------------------------------------------------------------
    # evaluation helpers
    # ------------------------------------------------------------------
    @staticmethod
    def _signature(costs, limits):
        """Cache key of a compiled plan: the CONTENT of the costs and limits it was built
        from (plan.cost_token / limit_token).  The reference re-reads the objects on every
        ``generate_*`` call (body.py:236-302), so a field changed by hand, or a fresh
        object that happens to reuse the address of a dead one, must never meet a plan
        compiled from other numbers."""
        return (tuple(_plan.cost_token(c) for c in costs),
                tuple(_plan.limit_token(lim) for lim in limits))

    def _run(self, given, params, costs=None, limits=None, out=None, layout="dense"):
        """Evaluates a (sub)plan for a batch: given [B, ng], params {name: [B]}."""
        compiled = self._compiled(costs, limits)
        if isinstance(given, (np.ndarray, list, tuple)):  # anything else: a device tensor
            given = np.asarray(given, dtype=np.float64)
        batch = given.shape[0]
        scalars = compiled.slots.evaluate(params or {}, batch)
        if layout == "dense":
            return self._ops().assemble(compiled, scalars, given, batch, out)
        if layout not in ("csc", "csc_upper"):
            raise ValueError("layout must be 'dense', 'csc' or 'csc_upper'")
        return self._ops().assemble_csc(compiled, scalars, given, batch, out, layout == "csc_upper")

    def _single(self, given):
        given = np.asarray(given, dtype=np.float64)
        if given.size != self.given_len:
            raise ValueError(
                "matmul: the given vector has {} elements, the formulation expects "
                "{}".format(given.size, self.given_len)
            )
        return given.reshape([1, self.given_len])

    def preview(self, given, optim, variable, axes=None):
        """Value of a variable along the horizon (body.py:209-219): ``Mg @ given + Mo @ optim``
        -- a batch of one through the device preview plan (the sparse-row kernel), returned
        like the reference's: a column [rows, 1] (``axes`` None) or [rows, len(axes)]."""
        given = np.asarray(given, dtype=np.float64).reshape(1, self.given_len)
        optim = np.asarray(optim, dtype=np.float64).reshape(1, self.optim_len)
        values = self.preview_batched(given, optim, variable, axes)
        values = values.cpu().numpy()[0]
        return values.reshape(-1, 1) if axes is None else values

    def goal_distance(self, given, optim, goal_name):
        """Squared distance to the aim of one goal (body.py:221-229), a batch of one through
        ``goal_distance_batched`` (preview rows + the reduction kernel)."""
        given = np.asarray(given, dtype=np.float64).reshape(1, self.given_len)
        optim = np.asarray(optim, dtype=np.float64).reshape(1, self.optim_len)
        return float(self.goal_distance_batched(given, optim, goal_name).cpu().numpy()[0])

    def full_goal_distance(self, given, optim):
        return sum(self.goal_distance(given, optim, name) for name in self.goals)

    # ------------------------------------------------------------------
    # QP matrices (body.py:236-348) -- one instance = a batch of one
    # ------------------------------------------------------------------
    def generate_qp_constraint(self, limit, given):
        """Rows ``A z <= h`` of one limit (body.py:236-264)."""
        out = self._run(self._single(given), None, costs=[], limits=[limit])
        return out.G_numpy()[0], out.h_numpy()[0][:, None]

    def generate_qp_cost(self, cost, given):
        """``Q, q`` of one cost (body.py:266-302)."""
        out = self._run(self._single(given), None, costs=[cost], limits=[])
        return out.P_numpy()[0], out.q_numpy()[0][:, None]

    def generate_all_qp_constraints(self, given):
        """All limits stacked: constraints first, then boxes (body.py:304-320)."""
        limits = _plan.ordered_limits(self)
        if not limits:
            raise ValueError("need at least one array to concatenate")
        out = self._run(self._single(given), None, costs=[], limits=limits)
        return out.G_numpy()[0], out.h_numpy()[0][:, None]

    def generate_all_qp_costs(self, given):
        """Sum over all goals (body.py:322-329)."""
        out = self._run(self._single(given), None, costs=list(self.goals.values()), limits=[])
        return out.P_numpy()[0], out.q_numpy()[0][:, None]

    def generate_all_qp_matrices(self, given):
        """Returns ``A, h, Q, q`` for ``A z <= h`` and ``0.5 z'Qz + q'z``
        (body.py:333-348; the sign of q is the one the code and the example
        use, see biped_mpc_loop.py:60)."""
        if not _plan.ordered_limits(self):
            raise ValueError("need at least one array to concatenate")
        compiled, ops = self._compiled(), self._ops()
        if not compiled.slots.exprs and hasattr(ops, "assemble_single"):
            return ops.assemble_single(compiled, self._single(given))
        out = self._run(self._single(given), None)
        return (
            out.G_numpy()[0],
            out.h_numpy()[0][:, None],
            out.P_numpy()[0],
            out.q_numpy()[0][:, None],
        )

    # ------------------------------------------------------------------
    # batched entry point (new): device-resident outputs
    # ------------------------------------------------------------------
    def parameter_names(self):
        """Names of the per-instance Param values the compiled plan needs."""
        return self._compiled().slots.names()

    def generate_all_qp_matrices_batched(self, given, params=None, out=None, dynamics=None,
                                         hessian="dmma", layout="dense"):
        """Assembles a whole batch.

        given    : [B, given_len] float64 (numpy, or a CUDA torch tensor)
        params   : {name: [B] array} for every ``Param`` used in the formulation
        out      : optional ``QPBatch`` to overwrite (ring-buffer slot)
        dynamics : ``{name: (A, B)}`` with ``A [B,n,n]``, ``B [B,n,m]`` -- one plant
                   per instance replacing the matrices the dynamics ``name`` was
                   built with (the formulation then only fixes the structure);
                   everything, prediction matrices included, is computed per
                   instance by the fused kernel of csrc/mpci_lti.cu (lti.py).
                   ``{name: (system_name, tau, omega)}`` builds the plants of a named
                   LIP system on the GPU from one sampling period / pendulum
                   frequency per instance (``ControlSystem.from_name``,
                   dynamics.py:380-407, for a whole batch).
        hessian  : with ``dynamics``: "dmma" (dense contraction on the FP64 tensor
                   cores) or "recursion" (displacement recursion, O(N^2) flops)

        layout   : "dense" (default), or the solver hand-off "csc" / "csc_upper": only
                   the CSC values of P and G are written (``QPBatchCSC``; one sparsity
                   pattern for the batch, the int32 ``indices`` / ``indptr`` of
                   ``scipy.sparse.csc_matrix`` as in biped_mpc_loop.py:57-58;
                   "csc_upper" keeps the upper triangle of P for OSQP-style solvers)

        Returns a ``QPBatch`` with device tensors ``P [B,nv,nv]``, ``q [B,nv]``,
        ``G [B,nc,nv]``, ``h [B,nc]`` (row-major, unpadded)."""
        if dynamics is None:
            return self._run(given, params, out=out, layout=layout)
        if layout != "dense":
            raise NotImplementedError("per-instance dynamics write the dense layout only")
        from . import lti

        if len(dynamics) != 1:
            raise NotImplementedError("exactly one dynamics can vary per instance")
        (name, spec), = dynamics.items()
        if name not in self.dynamics:
            raise KeyError("no dynamics named '{}' was incorporated".format(name))
        if isinstance(spec[0], str):
            A, B = lti.named_system_batched(spec[0], *spec[1:])
        else:
            A, B = spec
        modes = {"dmma": lti.HESSIAN_DMMA, "recursion": lti.HESSIAN_RECURSION}
        if hessian not in modes:
            raise ValueError("hessian must be 'dmma' or 'recursion'")
        return lti.assemble(self.lti_plan(name), A, B, given, params, out, modes[hessian])

    # ------------------------------------------------------------------
    # after the solve, batched (body.py:209-234)
    # ------------------------------------------------------------------
    def _preview_plan(self, names):
        compiled = self._compiled()
        cache = compiled.__dict__.setdefault("_previews", {})
        key = tuple(names)
        if key not in cache:
            cache[key] = _plan.PreviewPlan(compiled, self._ops(), names)
        return compiled, cache[key]

    def _given_optim(self, given, optim):
        ops = self._ops()
        g, z = ops._to_device(given), ops._to_device(optim)
        batch = z.shape[0]
        g = g.reshape(batch, self.given_len)
        z = z.reshape(batch, self.optim_len)
        return ops.torch.cat([g, z], dim=1).contiguous(), batch

    def preview_batched(self, given, optim, variable, axes=None, params=None):
        """Batched ``preview`` (body.py:209-219): the value of ``variable`` along
        the horizon for every instance, ``Mg @ given + Mo @ optim`` with the
        per-instance preview matrices.  given [B, given_len], optim [B, optim_len]
        (numpy or CUDA tensors).  Returns a CUDA tensor [B, rows] (``axes`` None)
        or [B, rows, len(axes)] -- the reference's hstack over the axes."""
        names = [variable] if axes is None else [variable + axis for axis in axes]
        compiled, pplan = self._preview_plan(names)
        if len(set(pplan.rows)) != 1:
            raise ValueError("all the arrays must have the same number of rows")
        gz, batch = self._given_optim(given, optim)
        scalars = compiled.slots.evaluate(params or {}, batch)
        values = self._ops().preview(pplan, scalars, gz, batch)
        if axes is None:
            return values
        return values.reshape(batch, len(names), pplan.rows[0]).transpose(1, 2)

    def preview_all_batched(self, given, optim, names=None, params=None, scalars=None):
        """Batched form of the example loop's ``preview_all`` (biped_mpc_loop.py:62-65):
        ``{name: CUDA tensor [B, rows]}`` for every definition (or for ``names``),
        all of them from ONE launch of the sparse-row kernel (the variables may have
        different numbers of rows)."""
        names = list(self.definitions) if names is None else list(names)
        compiled, pplan = self._preview_plan(names)
        gz, batch = self._given_optim(given, optim)
        if scalars is None:  # else: a CUDA tensor [B, n_slots] evaluated by the caller
            scalars = compiled.slots.evaluate(params or {}, batch)
        values = self._ops().preview(pplan, scalars, gz, batch)
        out, at = {}, 0
        for name, rows in zip(names, pplan.rows):
            out[name] = values[:, at : at + rows]
            at += rows
        return out

    def preview_rows_batched(self, given, optim, picks, params=None, scalars=None, out=None):
        """Single samples of several variables for a batch: ``picks = [(name, row), ...]`` ->
        CUDA tensor [B, len(picks)], column k = row ``row_k`` of ``name_k`` (``preview``,
        body.py:209-219, restricted to the rows the receding-horizon loop reads,
        biped_mpc_loop.py:81-88).  One launch of the sparse-row kernel over exactly those rows."""
        picks = [(str(name), int(row)) for name, row in picks]
        compiled = self._compiled()
        cache = compiled.__dict__.setdefault("_previews", {})
        key = ("rows",) + tuple(picks)
        if key not in cache:
            cache[key] = _plan.PreviewPlan(compiled, self._ops(), [n for n, _ in picks], [[r] for _, r in picks])
        gz, batch = self._given_optim(given, optim)
        if scalars is None:
            scalars = compiled.slots.evaluate(params or {}, batch)
        return self._ops().preview(cache[key], scalars, gz, batch, out=out)

    def box_rows(self):
        """``plan.BoxRows`` of the compiled formulation (cached with the plan)."""
        compiled = self._compiled()
        if "_box_rows" not in compiled.__dict__:
            compiled._box_rows = _plan.BoxRows(compiled, self._ops())
        return compiled._box_rows

    def box_bounds_batched(self, qp):
        """``(lb, ub, rest)`` for a batch assembled by ``generate_all_qp_matrices_batched``
        (dense or CSC layout): CUDA tensors ``lb, ub [B, nv]`` (-inf / +inf where a variable
        has no bound row) from the rows of G that bound one variable each, and ``rest``, the
        int32 indices of the rows that remain general constraints -- ``G[:, rest] z <= h[:, rest]``
        with ``lb <= z <= ub`` is the same feasible set as ``G z <= h``."""
        box = self.box_rows()
        return self._ops().box_bounds(box, qp.h, self.optim_len) + (box.rest,)

    def goal_distance_batched(self, given, optim, goal_name, params=None):
        """Batched ``goal_distance`` (body.py:221-229): squared distance to the aim
        of one goal for every instance; a CUDA tensor [B]."""
        goal = self.goals[goal_name]
        names = [goal.variable + axis for axis in goal.axes]
        compiled, pplan = self._preview_plan(names)
        gz, batch = self._given_optim(given, optim)
        scalars = compiled.slots.evaluate(params or {}, batch)
        values = self._ops().preview(pplan, scalars, gz, batch)
        aim = np.asarray(goal.aim, dtype=object)
        if aim.shape[0] != 1:
            raise ValueError("operands could not be broadcast together: an aim must have one row")
        columns = []
        for entry in aim[0]:
            value = entry.evaluate(params or {}) if hasattr(entry, "evaluate") else float(entry)
            columns.append(np.broadcast_to(np.asarray(value, dtype=np.float64), [batch]))
        aims = np.stack(columns, axis=1)
        return self._ops().goal_distance(values, aims, pplan.rows[0], len(names))

    def full_goal_distance_batched(self, given, optim, params=None):
        total = None
        for name in self.goals:
            d = self.goal_distance_batched(given, optim, name, params)
            total = d if total is None else total + d
        return total

    def lti_plan(self, name):
        """The condensed LTI plan of this formulation with ``name`` as the
        per-instance plant (cached until the formulation changes)."""
        from . import lti

        key = ("lti", name) + self._signature(
            list(self.goals.values()), _plan.ordered_limits(self)
        )
        if key not in self._plans:
            self._trim_plans()
            self._plans[key] = lti.lower(self, name)
        return self._plans[key]
