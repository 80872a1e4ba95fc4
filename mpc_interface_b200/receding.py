"""Receding-horizon batches: instances at different phases of a time-variant
formulation.

``Formulation.update(**kargs)`` (body.py:142-147) may change the STRUCTURE of
the QP between ticks: the biped's ``plan_steps`` keeps only the steps that fall
inside the preview window (tools.py:84), so the number of ``Ds`` variables -- and
with it nv, nc, the index layout and the 0/1 step matrix E -- differs between
walking phases (nv 216/218, nc 436/440 at N=100).  A
batch can only share one structure, so the instances of a tick are grouped by a
structural key; every key owns one updated ``Formulation`` and its compiled
device plan (built once, then reused on every later tick with that structure),
and each bucket is assembled by one batched call.  Integer work only on the host:
the grouping, the index layouts; all matrices come from the kernels.
"""
import numpy as np


class StructureBuckets:
    def __init__(self, factory, key):
        """``factory()`` builds a fresh Formulation with an updating rule;
        ``key(**kargs)`` maps the update arguments of one instance to a hashable
        structural key (equal keys <=> same formulation after ``update``)."""
        self._factory = factory
        self._key = key
        self._forms = {}

    def formulation(self, **kargs):
        """The (cached) formulation updated to the structure of ``kargs``."""
        k = self._key(**kargs)
        if k not in self._forms:
            form = self._factory()
            form.update(**kargs)
            self._forms[k] = form
        return self._forms[k]

    def __len__(self):
        return len(self._forms)

    def group(self, instance_kargs):
        """{key: (representative kargs, int64 indices of its instances)} in order of
        first appearance."""
        groups = {}
        for b, kargs in enumerate(instance_kargs):
            k = self._key(**kargs)
            if k not in groups:
                groups[k] = (kargs, [])
            groups[k][1].append(b)
        return {k: (kargs, np.asarray(idx, dtype=np.int64)) for k, (kargs, idx) in groups.items()}

    @staticmethod
    def _partition(keys):
        """(unique key rows in order of first appearance, permutation of the instances that
        lists the groups one after the other, group boundaries [n_groups + 1]) -- integer
        work only, no Python loop over the instances.  Rows whose columns (shifted by their
        minima) fit 64 bits together are packed into one uint64 each, so the grouping is
        ``np.unique`` over a 1-D array instead of over rows (65,536 x 10: 95 ms on the bench
        box when done over rows)."""
        keys = np.ascontiguousarray(keys, dtype=np.int64)
        if keys.ndim != 2:
            raise ValueError("keys must be [B, K]")
        count, width = keys.shape
        if count == 0:
            return keys, np.zeros([0], dtype=np.int64), np.zeros([1], dtype=np.int64)
        low = keys.min(axis=0)
        bits = [max(1, int(x).bit_length()) for x in (keys.max(axis=0) - low)]
        if sum(bits) <= 64:
            packed = np.zeros([count], dtype=np.uint64)
            for column in range(width):
                packed = (packed << np.uint64(bits[column])) | (keys[:, column] - low[column]).astype(np.uint64)
            _, first, inverse = np.unique(packed, return_index=True, return_inverse=True)
        else:
            _, first, inverse = np.unique(keys, axis=0, return_index=True, return_inverse=True)
        inverse = inverse.reshape(-1)
        by_appearance = np.argsort(first, kind="stable")          # group rank -> unique index
        rank = np.empty_like(by_appearance)
        rank[by_appearance] = np.arange(len(by_appearance))
        group_of = rank[inverse]                                   # instance -> group rank
        order = np.argsort(group_of, kind="stable")
        bounds = np.concatenate([[0], np.cumsum(np.bincount(group_of, minlength=len(first)))])
        return keys[first[by_appearance]], order, bounds

    @staticmethod
    def _partition_device(keys):
        """``_partition`` for key rows that live on the GPU (a CUDA int64 tensor [B, K]): the
        packing, the grouping and the permutation stay on the device (integer tensor
        plumbing, no float arithmetic); only the handful of unique rows and the group sizes
        come back.  Groups are ordered by key, their indices ascending."""
        import torch

        low = keys.min(dim=0).values
        span = (keys.max(dim=0).values - low).tolist()
        bits = [max(1, int(x).bit_length()) for x in span]
        if sum(bits) > 64:
            rows, order, bounds = StructureBuckets._partition(keys.cpu().numpy())
            return rows, torch.from_numpy(order).to(keys.device), bounds
        packed = torch.zeros_like(keys[:, 0])
        for column, width in enumerate(bits):  # (wrapping into the sign bit keeps rows distinct)
            packed = (packed << width) | (keys[:, column] - low[column])
        _, inverse = torch.unique(packed, return_inverse=True)
        order = torch.argsort(inverse, stable=True)
        counts = torch.bincount(inverse)
        bounds = np.concatenate([[0], np.cumsum(counts.cpu().numpy())])
        first = order[torch.from_numpy(bounds[:-1]).to(keys.device)]
        return keys[first].cpu().numpy(), order, bounds

    @staticmethod
    def group_keys(keys):
        """Vectorised grouping: ``keys`` int64 [B, K], one structural key row per instance
        (e.g. ``biped.structure_keys``).  Returns ``[(key tuple, int64 indices)]`` ordered by
        first appearance, the indices of a group ascending."""
        rows, order, bounds = StructureBuckets._partition(keys)
        return [(tuple(int(x) for x in rows[g]), order[bounds[g]:bounds[g + 1]])
                for g in range(len(rows))]

    def assemble_device(self, keys, kargs_of, collector, scalars=None, params=None, outs=None,
                        dynamics=None, hessian="recursion"):
        """One tick for a batch whose inputs live on the GPU (a receding-horizon stream: the
        collector of tick t+1 is produced on the device by ``advance_collector``).

        keys      : int64 [B, K] structural key rows (``group_keys``): a numpy array, or a
                    CUDA tensor -- then the grouping runs on the device as well
        kargs_of  : ``kargs_of(b)`` -> the update arguments of instance b; called once per
                    structure that has not been seen before (for its representative)
        collector : {given variable: CUDA tensor [B, size]}
        params    : {name: array [B]} per-instance ``Param`` values (host), evaluated per
                    structure; or ``scalars``: a CUDA tensor [B, n_slots] already evaluated
                    in the slot order the structures share (``slots_of``)
        outs      : optional dict reused across ticks: {(key, count): QPBatch} ring slots
        dynamics  : ``{name: (A [B, n, n], B [B, n, m])}`` CUDA tensors (or numpy): one plant per
                    robot for the dynamics ``name`` (``generate_all_qp_matrices_batched(...,
                    dynamics=...)`` per bucket: fused per-instance kernel, embedded layout when
                    the structure has other variables); ``params`` then holds the per-instance
                    ``Param`` values ([B] arrays or CUDA tensors), ``hessian`` picks the mode

        Returns ``[(key, indices (CUDA int64 tensor), formulation, QPBatch)]``; the rows of a
        bucket's inputs are gathered on the device (no host copy of ``given``)."""
        on_device = hasattr(keys, "is_cuda") and keys.is_cuda
        rows, order, bounds = self._partition_device(keys) if on_device else self._partition(keys)
        groups = []
        for g in range(len(rows)):
            key = tuple(int(x) for x in rows[g])
            if key not in self._forms:
                form = self._factory()
                form.update(**kargs_of(int(order[int(bounds[g])])))
                self._forms[key] = form
            groups.append((key, self._forms[key], int(bounds[g]), int(bounds[g + 1])))
        if not groups:
            return []
        ops = groups[0][1]._ops()
        torch = ops.torch
        # ONE upload of the permutation; a bucket's indices are a slice of it
        order_dev = order if on_device else torch.from_numpy(order).to(ops.device)
        # The permutation sorts the batch by structure.  When the structures agree on the layout
        # of the given vector (the biped's do: only the optimisation side is ragged) the inputs
        # are gathered ONCE for the whole batch and a bucket's rows are a contiguous slice.
        first = groups[0][1]
        shared = all(list(f.given_ID.items()) == list(first.given_ID.items()) for _, f, _, _ in groups)
        given_all = scal_all = None
        if shared and first.given_variables:
            given_all = torch.cat([collector[v].reshape(collector[v].shape[0], -1).index_select(0, order_dev)
                                   for v in first.given_variables], dim=1)
        if scalars is not None:
            scal_all = scalars.index_select(0, order_dev)
        plant = None
        if dynamics is not None:
            from . import lti

            if len(dynamics) != 1:
                raise NotImplementedError("exactly one dynamics can vary per instance")
            (plant_name, (A_b, B_b)), = dynamics.items()
            to_dev = lambda x: x if torch.is_tensor(x) else torch.from_numpy(  # noqa: E731
                np.require(x, dtype=np.float64, requirements=["C", "W"])).to(ops.device)
            plant = (plant_name, to_dev(A_b).index_select(0, order_dev), to_dev(B_b).index_select(0, order_dev),
                     {k: to_dev(np.asarray(v) if not torch.is_tensor(v) else v).reshape(-1).index_select(0, order_dev)
                      for k, v in (params or {}).items()},
                     {"dmma": lti.HESSIAN_DMMA, "recursion": lti.HESSIAN_RECURSION}[hessian])
        # The buckets are independent: their launches go round-robin over a few streams, so that
        # the tail of one bucket's kernels overlaps the head of the next one's.
        main = torch.cuda.current_stream()
        streams = self._streams(torch, ops.device) if len(groups) > 1 else [main]
        ready = torch.cuda.Event()
        ready.record(main)
        out = []
        for g, (key, form, b0, b1) in enumerate(groups):
            count = b1 - b0
            index = order_dev[b0:b1]
            side = streams[g % len(streams)]
            if side is not main and g < len(streams):
                side.wait_event(ready)
            with torch.cuda.stream(side):
                if given_all is not None:
                    given = given_all[b0:b1]
                elif form.given_variables:
                    given = torch.cat([collector[v].reshape(collector[v].shape[0], -1).index_select(0, index)
                                       for v in form.given_variables], dim=1)
                else:
                    given = torch.zeros([count, 0], dtype=torch.float64, device=ops.device)
                slot = None if outs is None else outs.get((key, count))
                if plant is not None:  # one plant per robot: the bucket through the fused kernel
                    plant_name, A_all, B_all, par_all, mode = plant
                    job = lti.LtiJob(form.lti_plan(plant_name), A_all[b0:b1], B_all[b0:b1], given,
                                     {k: v[b0:b1] for k, v in par_all.items()})
                    qp = job.launch(slot, mode)
                else:
                    compiled = form._compiled()
                    if scal_all is not None:
                        scal = scal_all[b0:b1]
                    else:
                        idx = index.cpu().numpy() if on_device else order[b0:b1]
                        sub = {name: np.asarray(v)[idx] for name, v in (params or {}).items()}
                        scal = compiled.slots.evaluate(sub, count)
                    qp = ops.assemble(compiled, scal, given, count, slot)
            if outs is not None:
                outs[(key, count)] = qp
            out.append((key, index, form, qp))
        for side in streams:
            if side is not main:
                main.wait_stream(side)
        # what advance_device needs again: the batch in structure order
        self._last = dict(order=order_dev, given=given_all, scalars=scal_all,
                          bounds=[(b0, b1) for _, _, b0, b1 in groups], done=out)
        return out

    def advance_device(self, done, optims, take, collector, scalars=None, params=None):
        """The example loop's ``update_given_collector`` (biped_mpc_loop.py:81-88) for the whole
        batch of the last ``assemble_device`` tick: the initial states of the next tick are
        single samples of the previewed states.

        done      : what ``assemble_device`` returned
        optims    : per bucket of ``done``, the optimum [count, optim_len] (CUDA tensors, rows in
                    the bucket's order -- the order of its QPBatch)
        take      : {dynamics name: horizon row} as in ``advance_collector``
        collector : {given variable: CUDA tensor [B, size]}, batch order

        Per bucket ONE launch over exactly the rows that are read (``preview_rows_batched``: 8
        rows for the biped instead of the 800 of the full previews), on the inputs the tick
        gathered in structure order; the results go back to batch order with one
        ``index_copy_`` per given variable.  Returns the new collector (fresh tensors; entries
        that are not produced here are copied through)."""
        last = getattr(self, "_last", None)
        if last is None or last["done"] is not done or last["given"] is None:
            return self._advance_by_bucket(done, optims, take, collector, scalars, params)
        ops = done[0][2]._ops()
        torch = ops.torch
        order, given_all, scal_all = last["order"], last["given"], last["scalars"]
        if scal_all is None and scalars is not None:
            scal_all = scalars.index_select(0, order)
        batch = given_all.shape[0]
        first = done[0][2]
        picks = _state_picks(first, take)
        sampled = torch.empty([batch, len(picks)], dtype=torch.float64, device=ops.device)
        main = torch.cuda.current_stream()
        streams = self._streams(torch, ops.device) if len(done) > 1 else [main]
        ready = torch.cuda.Event()
        ready.record(main)
        for g, ((key, index, form, _qp), (b0, b1)) in enumerate(zip(done, last["bounds"])):
            side = streams[g % len(streams)]
            if side is not main and g < len(streams):
                side.wait_event(ready)
            with torch.cuda.stream(side):
                if scal_all is not None:
                    scal, sub = scal_all[b0:b1], None
                else:
                    idx = index.cpu().numpy()
                    scal, sub = None, {name: np.asarray(v)[idx] for name, v in (params or {}).items()}
                form.preview_rows_batched(given_all[b0:b1], optims[g], [(s_, r) for s_, _, _, r in picks],
                                          params=sub, scalars=scal, out=sampled[b0:b1])
        for side in streams:
            if side is not main:
                main.wait_stream(side)
        out = {name: ops._to_device(value).reshape(batch, -1).clone() for name, value in collector.items()}
        by_variable = {}
        for k, (_state, variable, ident, _row) in enumerate(picks):
            by_variable.setdefault(variable, []).append((ident, k))
        for variable, pairs in by_variable.items():
            ids = first.given_ID[variable]
            fresh = given_all[:, ids.start:ids.stop].clone()  # structure order
            fresh[:, [i for i, _ in pairs]] = sampled[:, [k for _, k in pairs]]
            if variable not in out:
                out[variable] = torch.empty_like(fresh)
            out[variable].index_copy_(0, order, fresh)
        return out

    def _advance_by_bucket(self, done, optims, take, collector, scalars, params):
        """``advance_device`` without the sorted inputs of the tick: one ``advance_collector``
        per bucket, scattered back."""
        ops = done[0][2]._ops()
        torch = ops.torch
        batch = sum(int(index.shape[0]) for _, index, _, _ in done)
        out = {name: ops._to_device(value).reshape(batch, -1).clone() for name, value in collector.items()}
        for (key, index, form, _qp), optim in zip(done, optims):
            given = torch.cat([out[v].index_select(0, index) for v in form.given_variables], dim=1)
            sub = scalars.index_select(0, index) if scalars is not None else None
            sub_params = None if sub is not None else {k: np.asarray(v)[index.cpu().numpy()]
                                                       for k, v in (params or {}).items()}
            moved = advance_collector(form, given, optim, take, sub_params,
                                      {k: v.index_select(0, index) for k, v in out.items()}, scalars=sub)
            for name, value in moved.items():
                if name not in out:
                    out[name] = torch.zeros([batch, value.shape[1]], dtype=torch.float64, device=ops.device)
                out[name].index_copy_(0, index, value)
        return out

    def _streams(self, torch, device, count=4):
        if getattr(self, "_side_streams", None) is None:
            self._side_streams = [torch.cuda.Stream(device=device) for _ in range(count)]
        return self._side_streams

    def assemble(self, instance_kargs, collector, params=None, layout="dense"):
        """One tick for a batch.

        instance_kargs : per instance, the keyword arguments its ``update`` needs
        collector      : {given variable: array [B, size]} (the batched form of the
                         example's ``given_collector``, biped_mpc_loop.py:81-92)
        params         : {name: array [B]} per-instance ``Param`` values

        Returns a list of ``(key, indices, formulation, QPBatch)``: the instances
        ``indices`` (positions in the batch) share the structure of ``formulation``
        and their matrices are the rows of that ``QPBatch``, in the same order."""
        out = []
        for k, (kargs, idx) in self.group(instance_kargs).items():
            form = self.formulation(**kargs)
            given = form.arrange_given_batched({name: np.asarray(v)[idx] for name, v in collector.items()})
            sub = {name: np.asarray(v)[idx] for name, v in (params or {}).items()}
            out.append((k, idx, form, form.generate_all_qp_matrices_batched(given, sub, layout=layout)))
        return out


def _state_picks(form, take):
    """[(state, given variable, entry of that variable, horizon row)] for ``take``."""
    picks = []
    for name, row in take.items():
        dyn = form.dynamics[name]
        for state, ident in dyn.state_ID.items():
            if not 0 <= row < dyn.all_variables[state]:
                raise IndexError("row {} is outside the horizon of {}".format(row, state))
            picks.append((state, dyn.state_vector_name + "0" + state[-2:], ident, row))
    return picks


def count_steps(step_times, step_count, step_samples):
    """The example loop's ``count_steps`` (biped_mpc_loop.py:41-45) for a batch, integer
    work on the host: step_times int [B, n_steps], step_count int [B].  Every
    instance's times go down by one; where the first one reaches -1 the step has
    been taken: the times move on by ``step_samples`` and the count by one.
    Returns new arrays (the inputs are not modified)."""
    times = np.asarray(step_times).astype(np.int64) - 1
    count = np.asarray(step_count).astype(np.int64).copy()
    landed = times[:, 0] == -1
    times[landed] += step_samples
    count[landed] += 1
    return times, count


def advance_collector(form, given, optim, take, params=None, collector=None, scalars=None):
    """The example loop's ``update_given_collector`` (biped_mpc_loop.py:81-88) for a batch:
    the initial state of the next tick is read off the previewed horizon.

    take : {dynamics name: horizon row}; for every state of that ``ExtendedSystem``
           entry ``state_ID[state]`` of its ``<state_vector_name>0<axis>`` given
           variable becomes row ``take[...]`` of the state's preview (the example takes
           row 0 of the LIP states and row 1 of the step positions).

    The previews of all those states come from one kernel launch
    (``Formulation.preview_all_batched``).  Returns ``{given variable: CUDA tensor
    [B, size]}``; entries of ``collector`` (``{name: [B, size]}``) that are not
    produced here -- the bias of the example, which the loop resets itself -- are
    passed through, moved to the device."""
    ops = form._ops()
    states = []
    for name, row in take.items():
        dyn = form.dynamics[name]
        for state, ident in dyn.state_ID.items():
            if not 0 <= row < dyn.all_variables[state]:
                raise IndexError("row {} is outside the horizon of {}".format(row, state))
            states.append((state, dyn.state_vector_name + "0" + state[-2:], ident, row))
    motion = form.preview_all_batched(given, optim, [s[0] for s in states], params, scalars=scalars)
    batch = next(iter(motion.values())).shape[0] if motion else 0
    out = {}
    for name, value in (collector or {}).items():
        out[name] = ops._to_device(value).reshape(batch, -1)
    fresh = set()
    for state, variable, ident, row in states:
        if variable not in fresh:
            # a copy: the caller's collector is not written to
            size = form.domain[variable]
            base = out.get(variable)
            out[variable] = (ops.torch.zeros([batch, size], dtype=ops.torch.float64, device=ops.device)
                             if base is None else base.clone())
            fresh.add(variable)
        out[variable][:, ident] = motion[state][:, row]
    return out
