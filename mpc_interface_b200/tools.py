"""Numerical tools of the formulation API (host side).

Mirror of the reference module python/mpc_interface/tools.py: same function
names, arguments and error behaviour for everything the QP-assembly path uses.
The floating-point work of ``extend_matrices`` runs on the GPU (kernel K1,
csrc/mpci_extend.cu) through the C-ABI library; there is no CPU fallback.
Integer / structural helpers (step plans, stepping-area centres) stay on the
host exactly as in the reference, because they define index layouts.
"""
import numpy as np

from . import params as _params

# ---------------------------------------------------------------------------
# prediction matrices (reference: tools.py:14-33)
# ---------------------------------------------------------------------------


def extend_matrices(N, A, B):
    """Prediction matrices of ``x+ = A x + B u`` over a horizon of N samples.

    Returns ``S`` with shape (N, n, n), ``S[k, c, i] = (A^(k+1))[i, c]`` and
    ``U``, a list of m arrays (N, N, n), ``U[j][k, p, i] = (A^(k-p) B)[i, j]``
    for p <= k and 0 above the diagonal -- the layout of the reference
    (tools.py:31-32).  Computed by the CUDA kernel behind
    ``mpci_extend_matrices`` (include/mpci.h).
    """
    from . import device

    A = np.ascontiguousarray(np.asarray(A, dtype=np.float64))
    B = np.ascontiguousarray(np.asarray(B, dtype=np.float64))
    n, m = B.shape
    if A.shape != (n, n):
        raise ValueError(
            "A must be square with as many rows as B, got {} and {}".format(
                A.shape, B.shape
            )
        )
    S, U = device.extend_matrices_host(int(N), A[None], B[None])
    return S[0], [U[0, j] for j in range(m)]


# ---------------------------------------------------------------------------
# footstep "dynamics" (reference: tools.py:36-101)
# ---------------------------------------------------------------------------


def plan_steps(N, count=0, step_times=None, regular_time=None, phase=0):
    """0/1 matrix ``E[k, j] = (count + k) > t_j`` for the steps ``t_j`` that
    fall inside the preview window [count, count+N-1)  (tools.py:79-101).
    Integer output: it fixes how many ``Ds`` variables the QP has."""
    ticks = count + np.arange(N)
    last = count + N - 1
    if step_times is not None:
        step_times = np.asarray(step_times)
        upcoming = step_times[(step_times >= count) & (step_times < last)]
    elif regular_time is not None:
        upcoming = np.array(
            [t for t in ticks if (t + phase) % regular_time == 0 and t < last]
        )
    else:
        raise KeyError(
            "either the step_times or some regular_time for steps must be provided"
        )
    return (ticks[:, None] > upcoming).astype(int)


def update_step_matrices(extSyst, **kargs):
    """``how_to_update_matrices`` rule of the steps system (tools.py:36-76).

    Needs ``count`` (defaults to 0) and one of ``step_times`` /
    ``regular_time``; optional ``w_phase``."""
    horizon = extSyst.matrices[-1].shape[0]
    if "step_times" in kargs:
        times, period = kargs["step_times"], None
    elif "regular_time" in kargs:
        times, period = None, kargs["regular_time"]
    else:
        raise KeyError(
            "This funtion needs either 'step_times' or 'regular_time', but the "
            "kargs introduced are {}".format(kargs.keys())
        )
    E = plan_steps(
        horizon, kargs.get("count", 0), times, period, kargs.get("w_phase", 0)
    )
    extSyst.matrices[0] = E[:, :, None]


def n_predicted_steps(count, N, step_times):
    step_times = np.asarray(step_times)
    return int(np.count_nonzero((step_times >= count) & (step_times < count + N - 1)))


def adapt_size(stamps, **kargs):
    """Update rule of the ``stamps`` output of the steps system
    (tools.py:204-215): stamps = 1*s0 + strictly-lower-ones * Ds."""
    system = kargs["extSyst"]
    axis = stamps.variables[0][2:]
    n_steps = system.domain["Ds" + axis]
    stamps.matrices[stamps.variables.index("Ds" + axis)] = np.tril(
        np.ones([n_steps + 1, n_steps]), -1
    )
    stamps.matrices[stamps.variables.index("s0" + axis)] = np.ones([n_steps + 1, 1])


def find_step_centers(step_count, n_next_steps, xy_lenght):
    """Centres of the next stepping areas, alternating sides (tools.py:158-165)."""
    first_side = (-1) ** (step_count + 1)
    signs = np.where(np.arange(n_next_steps) % 2 == 0, 1, -1)[:, None]
    return np.hstack(
        [
            np.full([n_next_steps, 1], 1.0) * xy_lenght[0],
            first_side * signs * xy_lenght[1],
        ]
    )


def update_stepping_area(box, **kargs):
    """tools.py:149-155"""
    centers = find_step_centers(
        kargs["step_count"], kargs["n_next_steps"], kargs["xy_lenght"]
    )
    for limit in box.constraints:
        limit.update(center=centers)


def make_simetric_vertices(xy_corner):
    """Four corners (+,+), (-,+), (-,-), (+,-) of a centred rectangle
    (tools.py:197-201)."""
    sx = np.array([1, -1, -1, 1])[:, None] * xy_corner[0]
    sy = np.array([1, 1, -1, -1])[:, None] * xy_corner[1]
    return np.hstack([sx, sy])


def recenter_on_real_state_x(box, **kargs):
    box.recenter_in_SS(new_center=kargs["x0_x"])


def recenter_on_real_state_y(box, **kargs):
    box.recenter_in_SS(new_center=kargs["x0_y"])


def recenter_support(box, **kargs):
    box.recenter_in_TS(new_center=kargs["s0"])


def rotation2D(angle):
    """[[c, -s], [s, c]]  (tools.py:217-220); ``angle`` may be a per-instance
    ``Param`` expression, the matrix then holds expressions."""
    c, s = _params.cos(angle), _params.sin(angle)
    return np.array([[c, -s], [s, c]])


def rotation3D(angle, axis="z"):
    """Rotation about one coordinate axis (tools.py:223-237); about "y" the 2-D
    block lands transposed on rows/columns (0, 2), i.e. the right-handed
    rotation, exactly as the reference fills it."""
    keep = {"x": 0, "y": 1, "z": 2}[axis]
    plane = [i for i in range(3) if i != keep]
    block = rotation2D(angle)
    if axis == "y":
        block = block.T
    R = np.zeros([3, 3], dtype=block.dtype)
    R[keep, keep] = 1
    R[np.ix_(plane, plane)] = block
    return R


def do_not_update(sys, **kargs):
    return None


# ---------------------------------------------------------------------------
# named LIP-family systems (reference: tools.py:240-358) -- host-side setup
# ---------------------------------------------------------------------------
# name -> (continuous G rows, continuous input column h, inputs, states, needs omega)
_W = "w"
_SYSTEMS = {
    "P->CC": ([[0, 1], [(_W, 2), 0]], [0, (_W, -2)], ["cCoP"], ["CoM", "CoM_dot"], True),
    "P->X": ([[(_W, 1)]], [(_W, -1)], ["cCoP"], ["DCM"], True),
    "dP->CCC": (
        [[0, 1, 0], [0, 0, 1], [0, (_W, 2), 0]],
        [0, 0, (_W, -2)],
        ["cCoP_dot"],
        ["CoM", "CoM_dot", "CoM_ddot"],
        True,
    ),
    "dP->CCP": (
        [[0, 1, 0], [(_W, 2), 0, (_W, -2)], [0, 0, 0]],
        [0, 0, 1],
        ["cCoP_dot"],
        ["CoM", "CoM_dot", "cCoP"],
        True,
    ),
    "J->CCC": (
        [[0, 1, 0], [0, 0, 1], [0, 0, 0]],
        [0, 0, 1],
        ["CoM_dddot"],
        ["CoM", "CoM_dot", "CoM_ddot"],
        False,
    ),
}
_LAMBDA_CACHE = {}


def get_system_variables(system):
    """(input_variables, state_variables) of a named system (tools.py:288-325)."""
    _, _, inputs, states, _ = _SYSTEMS[system]
    return list(inputs), list(states)


def discretize(G, *H):
    """Exact zero-order-hold discretisation, symbolically (tools.py:328-358):
    ``A = exp(G tau)``, ``b_i = int_0^tau exp(G t) dt * h_i``."""
    import sympy as sy

    tau, t = sy.symbols("tau, t")
    A = sy.exp(G * tau)
    Bs = tuple(sy.integrate(sy.exp(G * t), (t, 0, tau), conds="none") * h for h in H)
    return (A,) + Bs


def get_system_matrices(system):
    """Callables ``get_A(**params)``, ``get_B(**params)`` and the parameter
    list of a named system (tools.py:240-285).  The lambdas take ``tau`` (and
    ``omega``) as keywords and swallow any other keyword, as in the reference."""
    if system in _LAMBDA_CACHE:
        return _LAMBDA_CACHE[system]
    import sympy as sy

    rows, col, _, _, needs_omega = _SYSTEMS[system]
    w = sy.symbols("omega")

    def entry(e):
        if isinstance(e, tuple):  # (_W, p): sign(p) * omega**|p|
            return sy.sign(e[1]) * w ** abs(e[1])
        return e

    G = sy.Matrix([[entry(e) for e in r] for r in rows])
    h = sy.Matrix([entry(e) for e in col])
    A_, B_ = sy.simplify(discretize(G, h))
    parameters = (
        ("tau=None", "omega=None", "**kwargs") if needs_omega else ("tau=None", "**kwargs")
    )
    out = (
        sy.lambdify(parameters, A_, "numpy"),
        sy.lambdify(parameters, B_, "numpy"),
        parameters,
    )
    _LAMBDA_CACHE[system] = out
    return out
