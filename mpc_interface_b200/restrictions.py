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
from .params import contains_expr, is_expr


def _geometry(value, shape, as_float=False):
    """``value`` as a 2-D array of ``shape``; object dtype when it holds
    per-instance ``Param`` expressions (kept symbolic for the plan compiler,
    which splits it into monomials x constant arrays)."""
    arr = np.array(value, dtype=object if contains_expr(value) else None).reshape(shape)
    return arr.astype(float) if (as_float and arr.dtype != object) else arr


def _span(schedule):
    return (schedule.stop - schedule.start) if schedule else None


class Constraint:
    def __init__(
        self, variable, extreme, axes=None, arrow=None, center=None, L=None, schedule=None
    ):
        self.variable = variable
        self.axes = [""] if axes is None else axes
        if not isinstance(self.axes, list):
            raise TypeError("The axes must be a list of string")
        self.axes_len = len(self.axes)
        self.schedule = range(0) if schedule is None else schedule
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
        return self.L[0].shape[0] if self.L else None

    @property
    def t(self):
        return _span(self.schedule)

    def _geometry_rows(self):
        return np.array([self.arrow.shape[0], self.center.shape[0], self.extreme.shape[0]])

    @property
    def nlines(self):
        """Number of rows this limit produces, None when it follows the size
        of the constrained variable (restrictions.py:147-162)."""
        if self.L:
            return self.m
        if self.schedule:
            return self.t
        rows = self._geometry_rows()
        many = rows[rows != 1]
        return None if many.size == 0 else many[0]

    # -- set-up ------------------------------------------------------------
    def arrange_L(self, L):
        """One matrix per axis; needs an up-to-date schedule (restrictions.py:75-90)."""
        matrices = L if isinstance(L, list) else [L]
        if len(matrices) == 1:
            matrices = matrices * self.axes_len
        elif len(matrices) not in (self.axes_len, 0):
            self.L = matrices
            raise IndexError(
                "'L' must have 0, 1 or len(axes) = {} elements".format(self.axes_len)
            )
        self.L = matrices
        if self.schedule and any(mat.shape[-1] != self.t for mat in matrices):
            raise ValueError(
                "arrays in L must have {} columns, which is given by the 'schedule'.".format(
                    self.t
                )
            )

    def initialize_geometry(self, arrow, center):
        if arrow is not None:
            self.arrow = _geometry(arrow, [-1, self.axes_len])
        elif self.axes_len == 1:
            self.arrow = np.ones(self.extreme.shape)
        else:
            raise ValueError(
                "When using multiple axes, some normal direction 'arrow' must be provided"
            )
        if center is None:
            self.center = np.zeros([1, self.axes_len])
        else:
            self.center = _geometry(center, [-1, self.axes_len])

    def check_geometry(self):
        rows = self._geometry_rows()
        many = rows[rows != 1]
        if many.size and np.any(many != many[0]):
            raise ValueError(
                "The number of rows in 'arrow', 'center' and 'extreme' must be equal "
                "or 1, but they are {} respectively".format(rows)
            )
        if self.axes_len > 1:
            cols = np.array([self.arrow.shape[1], self.center.shape[1]])
            if np.any(cols != self.axes_len):
                raise IndexError(
                    "'arrow' and 'center' have {} columns but they must have {}, "
                    "one per axis.".format(cols, self.axes_len)
                )

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
        """``extreme + sum_a arrow_a * center_a``  (restrictions.py:198-199)."""
        return self.extreme + np.sum(self.arrow * self.center, axis=1).reshape([-1, 1])

    def update(self, extreme=None, arrow=None, center=None, L=None, schedule=None):
        self._stamp = getattr(self, "_stamp", 0) + 1  # invalidates compiled plans
        if schedule is not None:
            self.schedule = schedule
        if L is not None:
            self.arrange_L(L)
        if extreme is not None:
            self.extreme = _geometry(extreme, [-1, 1])
        if arrow is not None:
            self.arrow = _geometry(arrow, [-1, self.axes_len])
        if center is not None:
            self.center = _geometry(center, [-1, self.axes_len])
        if extreme is not None or arrow is not None:
            self.check_geometry()
            self.normalize()
        elif center is not None:
            self.check_geometry()

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
    """A convex polygon / polyhedron expressed as one Constraint per facet."""

    def __init__(self, time_variant=None, how_to_update=None):
        self.constraints = []
        self.ts_vertices = np.array([])
        self.ss_vertices = np.array([])
        self.ts_center = np.array([])
        self.ts_orientation = []
        self.ss_center = np.array([])
        self.ss_orientation = []
        self.scale_factor = np.array([1.0])
        self.schedule = range(0)
        self.safety_margin = 0
        self.axes = [""]
        self.ss_dimention = 0
        self.ts_dimention = 0
        self.time_variant = time_variant
        self._rule = how_to_update if (how_to_update and time_variant) else use.do_not_update

    @classmethod
    def task_space(
        cls, variable, vertices, axes=None, L=None, schedule=None, time_variant=None, how_to_update=None
    ):
        """Box on the task-space variable ``[v_x, v_y, ...]`` (restrictions.py:313-340)."""
        box = cls(time_variant, how_to_update)
        arrows, extremes, center = box_boundaries(vertices)
        for arrow, extreme in zip(arrows, extremes):
            box.constraints.append(Constraint(variable, extreme, axes, arrow, center, L, schedule))
        dim = vertices.shape[1]
        box.ss_dimention = box.ts_dimention = dim
        box.axes = box.constraints[0].axes
        box.ts_center = center
        box.ss_center = np.zeros(center.shape)
        box.ts_orientation = [np.eye(dim)]
        box.ss_orientation = [np.eye(dim)]
        box.ts_vertices = vertices
        box.schedule = schedule
        return box

    @classmethod
    def state_space(
        cls, variable, vertices, axes=None, schedule=None, time_variant=None, how_to_update=None
    ):
        """Box on consecutive samples of one variable: every facet normal is
        used as an ``L`` row (restrictions.py:342-378)."""
        box = cls(time_variant, how_to_update)
        normals, extremes, centroid = box_boundaries(vertices)
        offsets = np.sum(normals * centroid, axis=1).reshape([-1, 1])
        for normal, extreme, offset in zip(normals, extremes, offsets):
            box.constraints.append(
                Constraint(variable, extreme, axes=axes, center=offset, L=normal, schedule=schedule)
            )
        box.ss_dimention = vertices.shape[1]
        box.ts_dimention = 1
        box.axes = box.constraints[0].axes
        box.ss_center = centroid.reshape([-1, 1])
        box.ts_center = 0
        box.ss_orientation = [np.eye(box.ss_dimention)]
        box.ts_orientation = [1]
        box.ss_vertices = vertices
        box.schedule = schedule
        return box

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
    """Outward unit normals, offsets and centroid of the convex hull of
    ``vertices`` (one vertex per row)  (restrictions.py:492-531).  Facets are
    listed in qhull's simplex order."""
    import scipy.spatial as sp

    vertices = vertices.astype("float64")
    count, dim = vertices.shape
    center = vertices.sum(axis=0) / count
    if dim == 1:
        simplices = np.array([[0], [1]])
        arrows = np.ones([2, 1])
    elif dim == 2:
        simplices = sp.ConvexHull(vertices).simplices
        edge = vertices[simplices[:, 0]] - vertices[simplices[:, 1]]
        arrows = np.column_stack([edge[:, 1], -edge[:, 0]])
    elif dim == 3:
        simplices = sp.ConvexHull(vertices).simplices
        e0 = vertices[simplices[:, 0]] - vertices[simplices[:, 2]]
        e1 = vertices[simplices[:, 1]] - vertices[simplices[:, 2]]
        arrows = np.column_stack(
            [
                e0[:, 1] * e1[:, 2] - e0[:, 2] * e1[:, 1],
                e0[:, 2] * e1[:, 0] - e0[:, 0] * e1[:, 2],
                e0[:, 0] * e1[:, 1] - e0[:, 1] * e1[:, 0],
            ]
        )
    for i in range(arrows.shape[0]):
        arrows[i] = arrows[i] / np.linalg.norm(arrows[i])
    anchor = vertices[simplices[:, 0]]
    extremes = np.sum(arrows * (anchor - center), axis=1).reshape([-1, 1])
    return arrows, extremes, center
