"""Argument handling shared by ``Cost`` (goal.py) and ``Constraint``
(restrictions.py): both take ``axes``, a ``schedule`` and per-axis ``L`` matrices
with the same rules (reference: goal.py:62-66, :88-104 and restrictions.py:58-62,
:75-90, where the two classes each carry their own copy of them)."""


def axes_list(axes, empty):
    """The list of axis suffixes; ``empty`` tells whether a missing / empty value
    means the single unnamed axis.  Anything but a list is a TypeError."""
    chosen = [""] if (axes is None or (empty and not axes)) else axes
    if not isinstance(chosen, list):
        raise TypeError("The axes must be a list of strings")
    return chosen


def window_length(schedule):
    """Samples a non-empty ``range`` schedule covers (its stop minus its start, as the
    reference counts them), ``None`` for the empty schedule that means "all"."""
    return (schedule.stop - schedule.start) if schedule else None


def one_matrix_per_axis(L, n_axes, columns):
    """``L`` as a list with one matrix per axis.  A single matrix (bare or in a list)
    serves every axis; an empty list means "no L"; any other count than ``n_axes`` is
    an IndexError.  With a schedule in force (``columns`` not None) every matrix must
    have that many columns, else ValueError."""
    matrices = list(L) if isinstance(L, list) else [L]
    if len(matrices) == 1:
        matrices = matrices * n_axes
    elif matrices and len(matrices) != n_axes:
        raise IndexError("'L' takes no matrix, one matrix, or one per axis ({})".format(n_axes))
    if columns is not None:
        for matrix in matrices:
            if matrix.shape[-1] != columns:
                raise ValueError(
                    "the schedule selects {} samples, so every matrix of 'L' needs {} columns "
                    "(got {})".format(columns, columns, matrix.shape[-1]))
    return matrices
