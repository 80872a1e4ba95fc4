"""mpci_extend_matrices_static: K1 through an entry point that allocates nothing (its scratch
is a static device array, include/mpci.h) against the existing mpci_extend_matrices, through
the C ABI, on two calls in a row that share the array.

Collected last on purpose; passes alone in a fresh process
and after the other tests.  No counterpart in the
reference.
"""
import ctypes

import numpy as np
import pytest

import _cases as cases

pytestmark = pytest.mark.gpu


def _run(entry, A, B, N, want_S=True, want_U=True):
    import torch

    from mpc_interface_b200 import device

    dev = torch.device("cuda", torch.cuda.current_device())
    batch, n, m = B.shape
    dA, dB = torch.from_numpy(A).to(dev), torch.from_numpy(B).to(dev)
    # poisoned outputs: every element must be written by the call
    S = torch.full([batch, N, n, n], float("nan"), dtype=torch.float64, device=dev) if want_S else None
    U = torch.full([batch, m, N, N, n], float("nan"), dtype=torch.float64, device=dev) if want_U else None
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    device.check(getattr(device.lib(), entry)(device._ptr(dA), device._ptr(dB), device._ptr(S), device._ptr(U),
                                              batch, n, m, N, stream))
    torch.cuda.synchronize()
    return (S.cpu().numpy() if want_S else None), (U.cpu().numpy() if want_U else None)


def _close(got, want, what):
    assert got.shape == want.shape and np.isfinite(got).all(), what
    assert np.abs(got - want).max() <= cases.TOL * max(1.0, np.abs(want).max()), what


def _systems(seed, batch, n, m):
    rng = np.random.default_rng(seed)
    return np.eye(n) + 0.1 * rng.normal(size=[batch, n, n]), rng.normal(size=[batch, n, m])


def test_static_scratch_entry_point_matches_the_allocating_one_twice_in_a_row():
    from mpc_interface_b200 import device

    cap = device.lib().mpci_extend_static_capacity()
    assert cap == 524288
    # first call: close to the limit (64 * 50 * 12 * 16 = 614,400 would not fit; 48 does);
    # second call: a smaller, odd shape right after it -- stale contents of the shared array
    # from the first call must not reach its result
    shapes = [(48, 12, 4, 50), (5, 3, 1, 7), (48, 12, 4, 50)]
    for call, (batch, n, m, N) in enumerate(shapes):
        assert batch * N * n * (n + m) <= cap
        A, B = _systems(100 + call, batch, n, m)
        want_S, want_U = _run("mpci_extend_matrices", A, B, N)
        got_S, got_U = _run("mpci_extend_matrices_static", A, B, N)
        _close(got_S, want_S, "S, call {}".format(call))
        _close(got_U, want_U, "U, call {}".format(call))
        # the reference's layout itself (tools.py:31-32), instance 0: U[j][k][p] = A^(k-p) B[:, j]
        Ak = np.linalg.matrix_power(A[0], 3)
        _close(got_S[0, 2], Ak.T, "S[2] = (A^3)^T")
        _close(got_U[0, 0, 4, 2], (A[0] @ A[0] @ B[0])[:, 0], "U[0][4][2] = A^2 B[:, 0]")


def test_static_scratch_entry_point_single_outputs_and_limits():
    from mpc_interface_b200 import device

    A, B = _systems(7, 6, 6, 2)
    want_S, want_U = _run("mpci_extend_matrices", A, B, 33)
    got_S, _ = _run("mpci_extend_matrices_static", A, B, 33, want_U=False)
    _, got_U = _run("mpci_extend_matrices_static", A, B, 33, want_S=False)
    _close(got_S, want_S, "S alone")
    _close(got_U, want_U, "U alone")
    # beyond the static array: refused with a message that names the need and the limit
    A, B = _systems(8, 64, 12, 4)
    with pytest.raises(ValueError, match=r"614400 doubles of scratch.*holds 524288"):
        _run("mpci_extend_matrices_static", A, B, 50)
    with pytest.raises(ValueError, match="NULL"):
        device.check(device.lib().mpci_extend_matrices_static(None, None, None, None, 1, 2, 1, 3, None))
    assert device.lib().mpci_extend_matrices_static(None, None, None, None, 0, 2, 1, 3, None) in (0, -3)
