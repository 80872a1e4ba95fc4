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
