"""Half-space constraints and convex boxes (host descriptors).

Mirror of python/mpc_interface/restrictions.py: ``Constraint`` (:15-284),
``Box`` (:287-489) and ``box_boundaries`` (:492-531).  A constraint asks for

    arrow * ( V - center ) < extreme,    V = [L_x @ v_x[schedule], L_y @ ...]

and yields m (or t, or horizon) rows of the QP inequality ``G z <= h``.  The
facet ORDER of a box comes from qhull (``scipy.spatial.ConvexHull``) exactly as
in the reference, because it fixes the row order of G.
"""
import numpy as np

from . import tools as use
from ._shared import axes_list, one_matrix_per_axis, window_length
from .params import contains_expr, is_expr


def _geometry(value, shape, as_float=False):
    """``value`` as a 2-D array of ``shape``; object dtype when it holds
    per-instance ``Param`` expressions (kept symbolic for the plan compiler,
    which splits it into monomials x constant arrays)."""
    arr = np.array(value, dtype=object if contains_expr(value) else None).reshape(shape)
    return arr.astype(float) if (as_float and arr.dtype != object) else arr


class Constraint:
    def __init__(self, variable, extreme, axes=None, arrow=None, center=None, L=None, schedule=None):
        self.variable = variable
        self.axes = axes_list(axes, empty=False)
        self.axes_len = len(self.axes)
        self.schedule = schedule if schedule is not None else range(0)
        self.L = []
        if L is not None:
            self.arrange_L(L)
        self.extreme = _geometry(extreme, [-1, 1], as_float=True)
        self.initialize_geometry(arrow, center)
        self.check_geometry()
        self.normalize()

    # -- sizes -------------------------------------------------------------
    @property
    def m(self):
        """Rows of the L matrices, None without L."""
        return self.L[0].shape[0] if self.L else None

    @property
    def t(self):
        """Samples the schedule selects, None without a schedule."""
        return window_length(self.schedule)

    def _geometry_rows(self):
        return np.array([part.shape[0] for part in (self.arrow, self.center, self.extreme)])

    @property
    def nlines(self):
        """Number of rows this limit produces, None when it follows the size
        of the constrained variable (restrictions.py:147-162): L decides first, then the
        schedule, then whichever of arrow / center / extreme has more than one row."""
        if self.L:
            return self.m
        if self.schedule:
            return self.t
        several = [rows for rows in self._geometry_rows() if rows != 1]
        return several[0] if several else None

    # -- set-up ------------------------------------------------------------
    def arrange_L(self, L):
        """One matrix per axis, checked against the CURRENT schedule
        (restrictions.py:75-90); on a refusal the limit keeps the L it had."""
        self.L = one_matrix_per_axis(L, self.axes_len, self.t)

    def initialize_geometry(self, arrow, center):
        """Defaults: with one axis the arrow is +1 on every row of ``extreme``; with
        several axes a direction must be given (ValueError).  The centre defaults to the
        origin (restrictions.py:92-112)."""
        k = self.axes_len
        if arrow is None and k > 1:
            raise ValueError("a limit over {} axes needs its normal direction 'arrow'".format(k))
        self.arrow = np.ones(self.extreme.shape) if arrow is None else _geometry(arrow, [-1, k])
        self.center = np.zeros([1, k]) if center is None else _geometry(center, [-1, k])

    def check_geometry(self):
        """arrow / center / extreme have one row or a common number of rows (ValueError);
        arrow and center have one column per axis (IndexError) -- restrictions.py:114-133."""
        counts = self._geometry_rows()
        several = set(int(c) for c in counts if c != 1)
        if len(several) > 1:
            raise ValueError(
                "'arrow', 'center' and 'extreme' have {} rows: apart from single rows they "
                "must agree".format(list(counts)))
        if self.axes_len > 1:
            widths = [self.arrow.shape[1], self.center.shape[1]]
            if any(w != self.axes_len for w in widths):
                raise IndexError(
                    "'arrow' / 'center' have {} columns for {} axes".format(widths, self.axes_len))

    def normalize(self):
        """Makes every extreme non-negative by flipping the row's arrow
        (restrictions.py:181-194).  The sign of a per-instance (``Param``)
        extreme is not known here: such a row is left as written, the caller
        promises it is non-negative for every instance."""
        n_arrow, n_extreme = self.arrow.shape[0], self.extreme.shape[0]
        if n_arrow != n_extreme:
            if n_arrow == 1:
                self.arrow = np.resize(self.arrow, [n_extreme, self.axes_len])
            elif n_extreme == 1:
                self.extreme = np.resize(self.extreme, [n_arrow, 1])
        for i in range(self.extreme.shape[0]):
            if is_expr(self.extreme[i, 0]):
                continue
            if self.extreme[i] < 0:
                self.extreme[i] = -self.extreme[i]
                self.arrow[i] = -self.arrow[i]

    def broadcast(self):
        n = self.nlines
        if n:
            if self.extreme.shape[0] == 1:
                self.extreme = np.resize(self.extreme, [n, 1])
            if self.arrow.shape[0] == 1:
                self.arrow = np.resize(self.arrow, [n, self.axes_len])
            if self.center.shape[0] == 1:
                self.center = np.resize(self.center, [n, self.axes_len])

    # -- what the QP assembly reads ---------------------------------------
    def matrices(self):
        """Per axis, the row multipliers ``arrow[:, a, None]`` (times ``L_a``
        when L is set)  (restrictions.py:175-179)."""
        columns = [self.arrow[:, a][:, None] for a in range(self.axes_len)]
        if self.L:
            return [col * mat for col, mat in zip(columns, self.L)]
        return columns

    def bound(self):
        """Right-hand side in task space: ``extreme + arrow . center`` row by row, a column
        (restrictions.py:198-199)."""
        offset = (self.arrow * self.center).sum(axis=1)
        return self.extreme + offset.reshape([-1, 1])

    def update(self, extreme=None, arrow=None, center=None, L=None, schedule=None):
        """Replaces the given fields (restrictions.py:201-219).  The schedule goes first so
        that a new L is checked against it; new geometry is re-checked, and re-normalised
        when the extreme or the arrow changed."""
        self._stamp = getattr(self, "_stamp", 0) + 1  # compiled plans look at this
        if schedule is not None:
            self.schedule = schedule
        if L is not None:
            self.arrange_L(L)
        k = self.axes_len
        fresh = {"extreme": (extreme, [-1, 1]), "arrow": (arrow, [-1, k]), "center": (center, [-1, k])}
        changed = [name for name, (value, _) in fresh.items() if value is not None]
        for name in changed:
            value, shape = fresh[name]
            setattr(self, name, _geometry(value, shape))
        if changed:
            self.check_geometry()
        if "extreme" in changed or "arrow" in changed:
            self.normalize()

    # -- diagnostics (restrictions.py:221-257) -----------------------------
    def SS_to_TS(self, ss_point):
        if self.L:
            return np.vstack(
                [mat @ p.T for mat, p in zip(self.L, np.transpose(ss_point))]
            ).T
        return ss_point

    def _is_feasible_TS(self, ts_point):
        return np.sum(self.arrow * (ts_point - self.center), axis=1) < self.extreme

    def _is_feasible_SS(self, ss_point):
        return self._is_feasible_TS(self.SS_to_TS(ss_point))

    def is_feasible(self, points, space="SS"):
        check = {"SS": self._is_feasible_SS, "TS": self._is_feasible_TS}.get(space)
        if check is None:
            return None
        if isinstance(points, list):
            return [check(p) for p in points]
        return check(points)

    def __repr__(self):
        suffix = "" if self.axes == [""] else "_" + "".join(a[1:] for a in self.axes)
        text = "\nvariable: " + self.variable + suffix
        if self.L:
            text += "\nwith L = " + ",\n".join(str(mat) for mat in self.L)
        for label, rows, pad in (
            ("arrow", self.arrow, 7),
            ("center", self.center, 8),
            ("extreme", self.extreme, 9),
        ):
            text += "\n\t\t\t\t\t{}: ".format(label) + (" " * pad).join(str(r) for r in rows)
        return text + "\n"

    __str__ = __repr__


class Box:
    """A convex polygon / polyhedron expressed as one Constraint per facet.

    ``ts_*`` describe the box in task space (the space of ``[v_x, v_y, ...]``), ``ss_*`` in
    state space (consecutive samples of one variable); which of them are filled depends on
    the constructor used.  The attribute names, ``ss_dimention`` included, are the
    reference's (restrictions.py:287-311)."""

    # what a box looks like before a constructor has described it
    _BLANK = dict(ts_orientation=list, ss_orientation=list, schedule=lambda: range(0),
                  safety_margin=int, ss_dimention=int, ts_dimention=int, axes=lambda: [""],
                  scale_factor=lambda: np.array([1.0]))

    def __init__(self, time_variant=None, how_to_update=None):
        self.constraints = []
        for name in ("ts_vertices", "ss_vertices", "ts_center", "ss_center"):
            setattr(self, name, np.array([]))
        for name, make in self._BLANK.items():
            setattr(self, name, make())
        self.time_variant = time_variant
        self._rule = how_to_update if (time_variant and how_to_update is not None) else use.do_not_update

    def _facets(self, limits, schedule, **description):
        """Takes over the facet limits and the description of the box they form."""
        self.constraints.extend(limits)
        self.axes = self.constraints[0].axes
        self.schedule = schedule
        for name, value in description.items():
            setattr(self, name, value)
        return self

    @classmethod
    def task_space(cls, variable, vertices, axes=None, L=None, schedule=None, time_variant=None,
                   how_to_update=None):
        """Box around the task-space variable ``[v_x, v_y, ...]``: one limit per facet of the
        hull of ``vertices`` (one vertex per row), all sharing the centroid as centre
        (restrictions.py:313-340)."""
        normals, offsets, centroid = box_boundaries(vertices)
        limits = [Constraint(variable, offset, axes, normal, centroid, L, schedule)
                  for normal, offset in zip(normals, offsets)]
        dim = vertices.shape[1]
        return cls(time_variant, how_to_update)._facets(
            limits, schedule, ts_vertices=vertices, ts_center=centroid,
            ss_center=np.zeros(centroid.shape), ts_dimention=dim, ss_dimention=dim,
            ts_orientation=[np.eye(dim)], ss_orientation=[np.eye(dim)])

    @classmethod
    def state_space(cls, variable, vertices, axes=None, schedule=None, time_variant=None,
                    how_to_update=None):
        """Box in the space of consecutive samples of ONE variable: facet f becomes the
        scalar limit ``normal_f . v[schedule] < ...`` with the normal as a one-row ``L``
        and the centroid's projection on it as centre (restrictions.py:342-378)."""
        normals, offsets, centroid = box_boundaries(vertices)
        projections = (normals * centroid).sum(axis=1).reshape([-1, 1])
        limits = [Constraint(variable, offset, axes=axes, center=projection, L=normal, schedule=schedule)
                  for normal, offset, projection in zip(normals, offsets, projections)]
        dim = vertices.shape[1]
        return cls(time_variant, how_to_update)._facets(
            limits, schedule, ss_vertices=vertices, ss_center=centroid.reshape([-1, 1]), ts_center=0,
            ss_dimention=dim, ts_dimention=1, ss_orientation=[np.eye(dim)], ts_orientation=[1])

    # -- transforms (restrictions.py:380-486) -----------------------------
    def _check_ss_shape(self, value):
        shape = np.shape(value)
        if shape not in ((self.ss_dimention, self.ts_dimention), (self.ts_dimention,)):
            raise ValueError(
                "The 'new_center' must have {} rows and {} columns, but its shape "
                "is {}".format(self.ss_dimention, self.ts_dimention, shape)
            )

    def recenter_in_TS(self, new_center):
        self.ts_center = np.array(new_center)
        for limit in self.constraints:
            limit.update(center=self.ts_center)

    def recenter_in_SS(self, new_center):
        self._check_ss_shape(new_center)
        self.ss_center = np.array(new_center)
        for limit in self.constraints:
            limit.update(center=limit.SS_to_TS(new_center))

    def reschedule(self, new_schedule):
        self.schedule = new_schedule
        for limit in self.constraints:
            limit.update(schedule=new_schedule)

    def translate_in_TS(self, translation):
        self.ts_center = self.ts_center + translation  # (not +=: the translation may be a Param)
        for limit in self.constraints:
            limit.update(center=limit.center + translation)

    def translate_in_SS(self, translation):
        self._check_ss_shape(translation)
        self.ss_center += np.array(translation)
        for limit in self.constraints:
            limit.update(center=limit.center + limit.SS_to_TS(translation))

    def rotate_in_TS(self, rotations):
        lines = self.constraints[0].nlines
        count = 1 if lines is None else lines
        if not isinstance(rotations, list):
            rotations = [rotations] * count
        if len(rotations) not in (count, 1):
            raise IndexError("'rotations' must contain 1 or {} rotation matrices".format(count))
        if len(rotations) == 1:
            rotations = rotations * count
        for limit in self.constraints:
            arrows = limit.arrow
            if arrows.shape[0] == 1 and count > 1:
                arrows = np.resize(arrows, (count, limit.axes_len))
            limit.update(arrow=np.vstack([n @ R.T for R, n in zip(rotations, arrows)]))

    def rotate_in_SS(self, rotations):
        raise NotImplementedError("Maybe later.")

    def scale_box(self, scale_factor):
        for limit in self.constraints:
            limit.update(extreme=limit.extreme * scale_factor / self.scale_factor)
        self.scale_factor = scale_factor

    def set_safety_margin(self, margin):
        # Frobenius norm of the whole arrow array, as in restrictions.py:484
        for limit in self.constraints:
            limit.update(extreme=limit.extreme - margin * np.linalg.norm(limit.arrow))
        self.safety_margin = margin

    def is_feasible(self, points, space="SS"):
        if not isinstance(points, list):
            points = [points]
        return [all(limit.is_feasible(p, space) for limit in self.constraints) for p in points]

    def update(self, **kargs):
        self._rule(self, **kargs)


def box_boundaries(vertices):
    """Facet normals (unit length), facet offsets from the centroid and the centroid of the
    convex hull of ``vertices`` (one vertex per row) -- what restrictions.py:492-531 hands to
    ``Box``.  The facets come in qhull's simplex order, because that order becomes the row
    order of G; the normal of a facet is the 90-degree turn of its edge (2-D) or the cross
    product of two of its edges (3-D), as qhull lists the vertices, and is NOT re-oriented
    here: ``Constraint.normalize`` flips a limit whose offset comes out negative."""
    import scipy.spatial

    points = np.asarray(vertices, dtype="float64")
    dim = points.shape[1]
    centroid = points.sum(axis=0) / points.shape[0]
    if dim == 1:  # a segment: its two ends, both with the normal +1
        facets = np.array([[0], [1]])
        normals = np.ones([2, 1])
    elif dim in (2, 3):
        facets = scipy.spatial.ConvexHull(points).simplices
        corner = points[facets]  # [facet, vertex of the facet, coordinate]
        if dim == 2:
            quarter_turn = np.array([[0.0, -1.0], [1.0, 0.0]])  # (x, y) -> (y, -x)
            normals = (corner[:, 0] - corner[:, 1]) @ quarter_turn
        else:
            normals = np.cross(corner[:, 0] - corner[:, 2], corner[:, 1] - corner[:, 2])
    else:
        raise ValueError("boxes are built in 1, 2 or 3 dimensions, not {}".format(dim))
    normals = normals / np.linalg.norm(normals, axis=1, keepdims=True)
    offsets = ((points[facets[:, 0]] - centroid) * normals).sum(axis=1, keepdims=True)
    return normals, offsets, centroid
