"""Out-of-bounds writes would land in the guard zones placed around every output
array here (compute-sanitizer is not available on the GPU pool): the outputs of the
shared-dynamics path and of the per-instance-dynamics path are carved out of one
buffer filled with a sentinel, also at addresses that are only 8-byte aligned."""
import numpy as np
import pytest

import _cases as cases

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    from mpc_interface_b200 import device

    device.require_gpu()
    return device

SENTINEL = -1.2345e300
GUARD = 64  # doubles before and after each array


class Guarded:
    """Same attributes as device.QPBatch, arrays inside one guarded buffer."""

    def __init__(self, batch, nv, nc, shift):
        import torch

        self.batch, self.nv, self.nc = batch, nv, nc
        sizes = {"P": batch * nv * nv, "q": batch * nv, "G": batch * nc * nv, "h": batch * nc}
        total = sum(GUARD + n + (n & 1) for n in sizes.values()) + GUARD + 2
        self.buf = torch.full([total], SENTINEL, dtype=torch.float64, device="cuda")
        self.spans, pos = {}, shift  # shift = 1: every array starts at an odd double
        for name, n in sizes.items():
            pos += GUARD
            self.spans[name] = (pos, pos + n)
            pos += n + (n & 1)
        self.P = self.buf[slice(*self.spans["P"])].view(batch, nv, nv)
        self.q = self.buf[slice(*self.spans["q"])].view(batch, nv)
        self.G = self.buf[slice(*self.spans["G"])].view(batch, nc, nv)
        self.h = self.buf[slice(*self.spans["h"])].view(batch, nc)

    def check(self):
        import torch

        inside = torch.zeros_like(self.buf, dtype=torch.bool)
        for a, b in self.spans.values():
            inside[a:b] = True
        outside = self.buf[~inside]
        assert bool((outside == SENTINEL).all()), "a kernel wrote outside its output arrays"
        assert not bool((self.buf[inside] == SENTINEL).any()), "an output entry was left unwritten"


@pytest.mark.parametrize("shift", [0, 1])
@pytest.mark.parametrize("seed", [0, 2, 5])
def test_shared_dynamics_outputs_stay_inside(dev, seed, shift):
    g = cases.golden("random_forms")
    form, given, params, ref = cases.random_form(g, seed)
    plan = form._compiled()
    out = Guarded(given.shape[0], plan.nv, plan.nc, shift)
    form.generate_all_qp_matrices_batched(given, params, out=out)
    out.check()
    got = tuple(x.cpu().numpy() for x in (out.G, out.h, out.P, out.q))
    cases.assert_qp_close(got, ref)


@pytest.mark.parametrize("shift", [0, 1])
@pytest.mark.parametrize("hessian", ["dmma", "recursion"])
@pytest.mark.parametrize("name", ["generic_perinst_small", "generic_perinst_odd"])
def test_per_instance_dynamics_outputs_stay_inside(dev, name, hessian, shift):
    g = cases.golden(name)
    form = cases.formulate_generic(g["A_sys"][0], g["B_sys"][0], g["Q_w"], g["R_w"], int(g["N"]), g["rows"],
                                   refs=[cases.Param("ref%d" % i) for i in range(g["ref"].shape[1])])
    params = {"ref%d" % i: g["ref"][:, i] for i in range(g["ref"].shape[1])}
    plan = form.lti_plan("plant")
    batch = g["x0"].shape[0]
    out = Guarded(batch, plan.nv, plan.nc, shift)
    form.generate_all_qp_matrices_batched(g["x0"], params, out=out, hessian=hessian,
                                          dynamics={"plant": (g["A_sys"], g["B_sys"])})
    out.check()
    got = tuple(x.cpu().numpy() for x in (out.G, out.h, out.P, out.q))
    cases.assert_qp_close(got, (g["A"], g["h"], g["Q"], g["q"]))


@pytest.mark.parametrize("n,m,N,batch", [(3, 1, 10, 5), (3, 1, 9, 4), (4, 2, 7, 300), (5, 3, 12, 7), (1, 1, 2, 3)])
def test_prediction_matrices_stay_inside(dev, n, m, N, batch):
    """K1 in the dense layouts: fused kernel with TMA row copies (N n even; shifted
    source for odd n) and the three-kernel path (N n odd), outputs between guard zones,
    against the oracle."""
    import ctypes

    import torch
    from oracle import qp_oracle as orc

    rng = np.random.default_rng(n * 10 + N)
    A = np.eye(n) + 0.1 * rng.normal(size=[batch, n, n])
    B = rng.normal(size=[batch, n, m])
    s_len, u_len = batch * N * n * n, batch * m * N * N * n
    buf = torch.full([3 * GUARD + s_len + (s_len & 1) + u_len + 2], SENTINEL, dtype=torch.float64, device="cuda")
    s0 = GUARD
    u0 = s0 + s_len + (s_len & 1) + GUARD
    dA, dB = torch.from_numpy(A).cuda(), torch.from_numpy(B).cuda()
    ptr = lambda t, off=0: ctypes.c_void_p(t.data_ptr() + 8 * off)
    dev.check(dev.lib().mpci_extend_matrices(ptr(dA), ptr(dB), ptr(buf, s0), ptr(buf, u0), batch, n, m, N,
                                             ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    host = buf.cpu().numpy()
    inside = np.zeros(host.shape, dtype=bool)
    inside[s0:s0 + s_len] = True
    inside[u0:u0 + u_len] = True
    assert (host[~inside] == SENTINEL).all() and not (host[inside] == SENTINEL).any()
    S = host[s0:s0 + s_len].reshape(batch, N, n, n)
    U = host[u0:u0 + u_len].reshape(batch, m, N, N, n)
    for b in sorted({0, batch // 2, batch - 1}):
        So, Uo = orc.extend_matrices(N, A[b], B[b])
        assert cases.rel_err(S[b], So) < 1e-12 and cases.rel_err(U[b], np.stack(Uo)) < 1e-12
        assert (U[b][:, np.triu_indices(N, 1)[0], np.triu_indices(N, 1)[1]] == 0.0).all()  # exact zeros


def test_kernel_timers_inside_a_call(dev):
    """mpci_plan_set_timing / mpci_plan_read_timing: one record of three positive
    durations per call while on, nothing collected while off, results unchanged."""
    import ctypes

    g = cases.golden("random_forms")
    form, given, params, ref = cases.random_form(g, 0)
    plan = form._compiled()
    handle = form._ops().device_plan(plan).handle
    lib = dev.lib()
    ms, pieces = (ctypes.c_double * 3)(), ctypes.c_long(-1)
    dev.check(lib.mpci_plan_set_timing(handle, 1))
    for _ in range(3):
        out = form.generate_all_qp_matrices_batched(given, params)
    dev.check(lib.mpci_plan_read_timing(handle, ctypes.byref(ms), ctypes.byref(pieces)))
    assert pieces.value == 3
    assert all(0.0 <= t < 1e3 for t in ms) and sum(ms) > 0.0
    dev.check(lib.mpci_plan_set_timing(handle, 0))
    form.generate_all_qp_matrices_batched(given, params)
    dev.check(lib.mpci_plan_read_timing(handle, ctypes.byref(ms), ctypes.byref(pieces)))
    assert pieces.value == 0 and list(ms) == [0.0, 0.0, 0.0]
    cases.assert_qp_close((out.G_numpy(), out.h_numpy(), out.P_numpy(), out.q_numpy()), ref)
