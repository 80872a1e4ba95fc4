"""Synthetic per-instance inputs of the benchmark configurations (SURVEY.md
section 8(d)).  Everything derives from ``numpy.random.default_rng(seed)`` over
the whole logical batch, so a shard is a slice and results do not depend on the
number of GPUs."""
import numpy as np


def biped_inputs(seed, batch, horizon):
    """Named per-instance values for the biped formulation.

    Returns ``{name: array[batch, size]}`` for the given variables
    (s0_x, s0_y, x0_x, x0_y, n_x, n_y) plus the per-instance scalars
    ``omega``, ``vel_x``, ``vel_y`` (shape [batch])."""
    rng = np.random.default_rng(seed)
    s0_x = rng.uniform(-0.1, 0.1, [batch, 1])
    s0_y = 0.09 * np.where(rng.random([batch, 1]) < 0.5, -1.0, 1.0)
    sigma = np.array([0.05, 0.2, 0.5])
    x0_x = rng.normal(0.0, 1.0, [batch, 3]) * sigma
    x0_y = rng.normal(0.0, 1.0, [batch, 3]) * sigma
    x0_x[:, 0] += s0_x[:, 0]
    x0_y[:, 0] += s0_y[:, 0]
    n_x = rng.normal(0.0, 0.005, [batch, horizon])
    n_y = rng.normal(0.0, 0.005, [batch, horizon])
    height = rng.uniform(0.75, 1.0, [batch])
    return {
        "s0_x": s0_x,
        "s0_y": s0_y,
        "x0_x": x0_x,
        "x0_y": x0_y,
        "n_x": n_x,
        "n_y": n_y,
        "omega": np.sqrt(9.81 / height),
        "vel_x": rng.uniform(0.0, 0.8, [batch]),
        "vel_y": rng.uniform(-0.1, 0.1, [batch]),
    }


def pack_given(given_ID, given_len, values, batch):
    """[batch, given_len] matrix in the formulation's given order
    (the batched form of Formulation.arrange_given, body.py:195-207)."""
    given = np.zeros([batch, given_len])
    for name, ids in given_ID.items():
        given[:, ids.start : ids.stop] = values[name]
    return given


def generic_system(seed, nx, nu):
    """Shared dynamics, dense SPD weights and constraint rows of the generic
    configurations C4 / C5."""
    rng = np.random.default_rng(seed)
    A = np.eye(nx) + 0.05 * rng.normal(size=[nx, nx])
    B = 0.1 * rng.normal(size=[nx, nu])
    M = rng.normal(size=[nx, nx])
    Q = M @ M.T / nx + np.eye(nx)
    M = rng.normal(size=[nu, nu])
    R = M @ M.T / nu + np.eye(nu)
    return A, B, Q, R


def generic_plants(seed, batch, nx, nu):
    """Per-instance plants (A_b, B_b) of the generic configurations in
    per-instance mode: the same distribution as ``generic_system``."""
    rng = np.random.default_rng(seed + 13)
    A = np.eye(nx) + 0.05 * rng.normal(size=[batch, nx, nx])
    B = 0.1 * rng.normal(size=[batch, nx, nu])
    return A, B


def generic_inputs(seed, batch, nx):
    rng = np.random.default_rng(seed + 7)
    return {"x0": rng.normal(size=[batch, nx]), "ref": rng.normal(size=[batch, nx])}
