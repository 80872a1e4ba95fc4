"""ctypes binding of libmpci.so and the device backend of the plan compiler.

PyTorch is used for device memory and streams only (allocation, slicing views,
host<->device copies); every floating-point operation goes through the C ABI
of include/mpci.h into the hand-written sm_100a kernels of csrc/.  There is no
CPU path: a missing library or GPU raises ``RuntimeError``.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libmpci.so")

MAX_FACTORS = 4
_c_double_p = ctypes.c_void_p
_lib = None


class PlanDesc(ctypes.Structure):
    """mpci_plan_desc of include/mpci.h"""

    _fields_ = [
        ("nv", ctypes.c_int),
        ("ng", ctypes.c_int),
        ("nc", ctypes.c_int),
        ("n_scalars", ctypes.c_int),
        ("n_monomials", ctypes.c_int),
        ("monomials", ctypes.c_void_p),
        ("n_P_terms", ctypes.c_int),
        ("P_terms", ctypes.c_void_p),
        ("DP_dev", ctypes.c_void_p),
        ("n_G_terms", ctypes.c_int),
        ("G_terms", ctypes.c_void_p),
        ("DG_dev", ctypes.c_void_p),
        ("n_Y_terms", ctypes.c_int),
        ("Y_terms", ctypes.c_void_p),
        ("n_Y_segs", ctypes.c_int),
        ("Y_segs", ctypes.c_void_p),
        ("n_Y_nz", ctypes.c_int),
        ("Y_val", ctypes.c_void_p),
        ("Y_col", ctypes.c_void_p),
    ]


# mpci_chunk_fn: void (*)(void* user, long first_instance, long count, long slot)
ChunkFn = ctypes.CFUNCTYPE(None, ctypes.c_void_p, ctypes.c_long, ctypes.c_long, ctypes.c_long)

_SIGNATURES = {
    "mpci_version": (ctypes.c_int, []),
    "mpci_last_error": (ctypes.c_char_p, []),
    "mpci_device_count": (ctypes.c_int, []),
    "mpci_extend_matrices": (
        ctypes.c_int,
        [_c_double_p] * 4 + [ctypes.c_int] * 4 + [ctypes.c_void_p],
    ),
    "mpci_extend_compact": (
        ctypes.c_int,
        [_c_double_p] * 4 + [ctypes.c_int] * 4 + [ctypes.c_void_p],
    ),
    "mpci_extend_matrices_host": (ctypes.c_int, [_c_double_p] * 4 + [ctypes.c_int] * 4),
    "mpci_gemm": (
        ctypes.c_int,
        [ctypes.c_int] * 4
        + [ctypes.c_double, _c_double_p, ctypes.c_int, ctypes.c_long]
        + [_c_double_p, ctypes.c_int, ctypes.c_long]
        + [ctypes.c_double, _c_double_p, ctypes.c_int, ctypes.c_long, ctypes.c_int]
        + [ctypes.c_void_p],
    ),
    "mpci_axpy2d": (
        ctypes.c_int,
        [ctypes.c_int, ctypes.c_int, ctypes.c_double, _c_double_p, ctypes.c_int]
        + [_c_double_p, ctypes.c_int, ctypes.c_int, _c_double_p, ctypes.c_int, ctypes.c_void_p],
    ),
    "mpci_plan_create": (ctypes.c_int, [ctypes.POINTER(PlanDesc), ctypes.POINTER(ctypes.c_void_p)]),
    "mpci_plan_destroy": (ctypes.c_int, [ctypes.c_void_p]),
    "mpci_assemble": (
        ctypes.c_int,
        [ctypes.c_void_p] + [_c_double_p] * 6 + [ctypes.c_long, ctypes.c_void_p],
    ),
    "mpci_assemble_host": (
        ctypes.c_int,
        [ctypes.c_void_p] + [_c_double_p] * 6 + [ctypes.c_long, ctypes.c_long],
    ),
    "mpci_assemble_host_stream": (
        ctypes.c_int,
        [ctypes.c_void_p] + [_c_double_p] * 6 + [ctypes.c_long] * 3 + [ChunkFn, ctypes.c_void_p],
    ),
    "mpci_assemble_compact_host_stream": (
        ctypes.c_int,
        [ctypes.c_void_p] + [_c_double_p] * 6 + [ctypes.c_long] * 3 + [ChunkFn, ctypes.c_void_p],
    ),
    "mpci_assemble_launches": (ctypes.c_int, [ctypes.c_void_p]),
    "mpci_plan_set_timing": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int]),
    "mpci_plan_read_timing": (
        ctypes.c_int,
        [ctypes.c_void_p, ctypes.POINTER(ctypes.c_double * 3), ctypes.POINTER(ctypes.c_long)],
    ),
    "mpci_goal_distance": (
        ctypes.c_int,
        [_c_double_p, _c_double_p, ctypes.c_long, ctypes.c_int, ctypes.c_int, _c_double_p,
         ctypes.c_long, ctypes.c_void_p],
    ),
    "mpci_box_bounds": (
        ctypes.c_int,
        [_c_double_p, ctypes.c_long, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, _c_double_p,
         _c_double_p, _c_double_p, ctypes.c_long, ctypes.c_void_p],
    ),
    "mpci_plan_set_compact": (
        ctypes.c_int, [ctypes.c_void_p, _c_double_p, ctypes.c_long, _c_double_p, ctypes.c_long]),
    "mpci_assemble_compact": (
        ctypes.c_int,
        [ctypes.c_void_p] + [_c_double_p] * 6 + [ctypes.c_long, ctypes.c_void_p],
    ),
    "mpci_assemble_compact_host": (
        ctypes.c_int,
        [ctypes.c_void_p] + [_c_double_p] * 6 + [ctypes.c_long, ctypes.c_long],
    ),
    "mpci_lti_plan_create": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]),
    "mpci_lti_plan_destroy": (ctypes.c_int, [ctypes.c_void_p]),
    "mpci_lti_assemble_host": (
        ctypes.c_int,
        [ctypes.c_void_p, _c_double_p, ctypes.c_long, _c_double_p, ctypes.c_long, _c_double_p,
         _c_double_p, _c_double_p, _c_double_p, _c_double_p, _c_double_p, ctypes.c_long,
         ctypes.c_int, ctypes.c_long],
    ),
    "mpci_lti_assemble_host_stream": (
        ctypes.c_int,
        [ctypes.c_void_p, _c_double_p, ctypes.c_long, _c_double_p, ctypes.c_long, _c_double_p,
         _c_double_p, _c_double_p, _c_double_p, _c_double_p, _c_double_p, ctypes.c_long,
         ctypes.c_int, ctypes.c_long, ctypes.c_long, ChunkFn, ctypes.c_void_p],
    ),
    "mpci_lti_assemble_ex": (
        ctypes.c_int,
        [ctypes.c_void_p, _c_double_p, ctypes.c_long, _c_double_p, ctypes.c_long, _c_double_p,
         _c_double_p, ctypes.c_void_p]
        + [_c_double_p] * 4
        + [ctypes.c_long, ctypes.c_int, ctypes.c_void_p],
    ),
    "mpci_lti_plan_create_embedded": (
        ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]),
    "mpci_lti_assemble_coupled": (
        ctypes.c_int,
        [ctypes.c_void_p, _c_double_p, ctypes.c_long, _c_double_p, ctypes.c_long, _c_double_p,
         _c_double_p, _c_double_p, ctypes.c_void_p, _c_double_p, _c_double_p]
        + [_c_double_p] * 4
        + [ctypes.c_long, ctypes.c_int, ctypes.c_void_p],
    ),
    "mpci_lti_assemble_coupled_ex": (
        ctypes.c_int,
        [ctypes.c_void_p, _c_double_p, ctypes.c_long, _c_double_p, ctypes.c_long, _c_double_p,
         _c_double_p, _c_double_p, ctypes.c_void_p, ctypes.c_void_p]
        + [_c_double_p] * 4
        + [ctypes.c_long, ctypes.c_int, ctypes.c_void_p],
    ),
    "mpci_lti_sched_create": (ctypes.c_int, [ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)]),
    "mpci_lti_sched_destroy": (ctypes.c_int, [ctypes.c_void_p]),
    "mpci_lti_sched_add": (
        ctypes.c_int,
        [ctypes.c_void_p, _c_double_p, ctypes.c_long, _c_double_p, ctypes.c_long, _c_double_p, _c_double_p,
         _c_double_p, ctypes.c_long, _c_double_p, _c_double_p, ctypes.c_long, ctypes.c_void_p],
    ),
    "mpci_lti_sched_rows": (
        ctypes.c_int,
        [ctypes.c_void_p, _c_double_p, ctypes.c_long, _c_double_p, ctypes.c_long, _c_double_p,
         _c_double_p, ctypes.c_long, _c_double_p, _c_double_p, ctypes.c_long, ctypes.c_long, ctypes.c_void_p],
    ),
    "mpci_extend_matrices_static": (
        ctypes.c_int,
        [_c_double_p, _c_double_p, _c_double_p, _c_double_p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
         ctypes.c_int, ctypes.c_void_p],
    ),
    "mpci_extend_static_capacity": (ctypes.c_long, []),
    "mpci_named_system": (
        ctypes.c_int,
        [ctypes.c_int, _c_double_p, ctypes.c_long, _c_double_p, ctypes.c_long, _c_double_p,
         _c_double_p, ctypes.c_long, ctypes.c_void_p],
    ),
    "mpci_lti_hessian_flops": (ctypes.c_double, [ctypes.c_void_p, ctypes.c_int]),
    "mpci_lti_plan_info": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]),
    "mpci_lti_assemble": (
        ctypes.c_int,
        [ctypes.c_void_p, _c_double_p, ctypes.c_long, _c_double_p, ctypes.c_long]
        + [_c_double_p] * 6
        + [ctypes.c_long, ctypes.c_int, ctypes.c_void_p],
    ),
}


def exported_symbols():
    return sorted(_SIGNATURES)


def lib():
    """Loads libmpci.so (built in-tree by __graft_entry__.build / csrc/Makefile)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "the CUDA library {} is missing: run `make -C {}` (there is no CPU "
                "fallback)".format(LIB_PATH, os.path.dirname(LIB_PATH))
            )
        loaded = ctypes.CDLL(LIB_PATH)
        for name, (restype, argtypes) in _SIGNATURES.items():
            fn = getattr(loaded, name)
            fn.restype = restype
            fn.argtypes = argtypes
        _lib = loaded
    return _lib


def check(rc):
    if rc != 0:
        message = lib().mpci_last_error().decode("utf-8", "replace")
        if rc < 0:
            raise ValueError("libmpci: {} (code {})".format(message, rc))
        raise RuntimeError("libmpci: CUDA error {}: {}".format(rc, message))


def require_gpu():
    if lib().mpci_device_count() < 1:
        raise RuntimeError(
            "no CUDA device is visible: mpc_interface_b200 computes every matrix on "
            "the GPU and has no CPU fallback"
        )


def _torch():
    import torch

    return torch


def _stream_ptr():
    # (torch._C._cuda_getCurrentRawStream: what torch.cuda.current_stream() wraps, without
    # building a Stream object -- a plan compile asks ~200 times)
    torch = _torch()
    return ctypes.c_void_p(torch._C._cuda_getCurrentRawStream(torch.cuda.current_device()))


def _ptr(tensor):
    return ctypes.c_void_p(tensor.data_ptr()) if tensor is not None else ctypes.c_void_p(0)


# ---------------------------------------------------------------------------
# K1
# ---------------------------------------------------------------------------
def extend_matrices_host(N, A, B):
    """A [batch,n,n], B [batch,n,m] (numpy) -> S [batch,N,n,n], U [batch,m,N,N,n]."""
    require_gpu()
    batch, n, m = B.shape
    S = np.empty([batch, N, n, n])
    U = np.empty([batch, m, N, N, n])
    check(
        lib().mpci_extend_matrices_host(
            A.ctypes.data, B.ctypes.data, S.ctypes.data, U.ctypes.data, batch, n, m, N
        )
    )
    return S, U


def extend_matrices_device(N, A, B, dense=True):
    """CUDA tensors in, CUDA tensors out.  dense=False returns the compact
    impulse response (Phi [batch,N,n,n], Gam [batch,N,n,m])."""
    require_gpu()
    torch = _torch()
    batch, n, m = B.shape
    A = A.contiguous()
    B = B.contiguous()
    if dense:
        S = torch.empty([batch, N, n, n], dtype=torch.float64, device=A.device)
        U = torch.empty([batch, m, N, N, n], dtype=torch.float64, device=A.device)
        check(lib().mpci_extend_matrices(_ptr(A), _ptr(B), _ptr(S), _ptr(U), batch, n, m, N, _stream_ptr()))
        return S, U
    Phi = torch.empty([batch, N, n, n], dtype=torch.float64, device=A.device)
    Gam = torch.empty([batch, N, n, m], dtype=torch.float64, device=A.device)
    check(lib().mpci_extend_compact(_ptr(A), _ptr(B), _ptr(Phi), _ptr(Gam), batch, n, m, N, _stream_ptr()))
    return Phi, Gam


# ---------------------------------------------------------------------------
# batched outputs
# ---------------------------------------------------------------------------
class QPBatch:
    """Device-resident QP matrices of a batch: ``P [B,nv,nv]``, ``q [B,nv]``,
    ``G [B,nc,nv]``, ``h [B,nc]`` -- row-major and unpadded, i.e. instance b is
    exactly the reference's C-ordered arrays (body.py:345-348)."""

    def __init__(self, batch, nv, nc, device="cuda"):
        torch = _torch()
        kw = dict(dtype=torch.float64, device=device)
        self.batch, self.nv, self.nc = batch, nv, nc
        self.P = torch.empty([batch, nv, nv], **kw)
        self.q = torch.empty([batch, nv], **kw)
        self.G = torch.empty([batch, nc, nv], **kw)
        self.h = torch.empty([batch, nc], **kw)

    @staticmethod
    def bytes_per_instance(nv, nc):
        return 8 * (nv * nv + nv + nc * nv + nc)

    def P_numpy(self):
        return self.P.cpu().numpy()

    def q_numpy(self):
        return self.q.cpu().numpy()

    def G_numpy(self):
        return self.G.cpu().numpy()

    def h_numpy(self):
        return self.h.cpu().numpy()


class QPBatchCSC:
    """Solver hand-off layout: the values of ``P`` and ``G`` in CSC order with one
    sparsity pattern for the whole batch (``P_indices`` / ``P_indptr`` ... are the
    int32 arrays of ``scipy.sparse.csc_matrix``, see plan.CscPattern); ``q`` and
    ``h`` stay dense.  ``csc(b)`` rebuilds scipy matrices of instance ``b``."""

    def __init__(self, batch, nv, nc, P_pattern, G_pattern, device="cuda"):
        torch = _torch()
        kw = dict(dtype=torch.float64, device=device)
        self.batch, self.nv, self.nc = batch, nv, nc
        self.P_pattern, self.G_pattern = P_pattern, G_pattern
        self._P = torch.empty([batch, P_pattern.padded], **kw)
        self._G = torch.empty([batch, G_pattern.padded], **kw)
        self.P_data = self._P[:, : P_pattern.nnz]
        self.G_data = self._G[:, : G_pattern.nnz]
        self.q = torch.empty([batch, nv], **kw)
        self.h = torch.empty([batch, nc], **kw)

    @staticmethod
    def bytes_per_instance(nv, nc, P_pattern, G_pattern):
        return 8 * (P_pattern.padded + G_pattern.padded + nv + nc)

    def csc(self, b):
        import scipy.sparse as sp

        mats = []
        for data, pat in ((self.P_data, self.P_pattern), (self.G_data, self.G_pattern)):
            mats.append(sp.csc_matrix((data[b].cpu().numpy(), pat.indices, pat.indptr), shape=pat.shape))
        return mats[0], mats[1]


class DevicePlan:
    """Owns the mpci_plan handle and keeps the constant tensors alive."""

    def __init__(self, plan):
        self.keep = (plan.DP, plan.DG, plan.DY)
        table = np.ascontiguousarray(plan.mono_table(), dtype=np.int32)
        P_terms = np.ascontiguousarray(plan.P_terms, dtype=np.int32)
        G_terms = np.ascontiguousarray(plan.G_terms, dtype=np.int32)
        Y_terms = np.ascontiguousarray(plan.Y_terms, dtype=np.int32)
        segs = np.ascontiguousarray(plan.Y_segs, dtype=np.int32)
        val = np.ascontiguousarray(plan.Y_val, dtype=np.float64)
        col = np.ascontiguousarray(plan.Y_col, dtype=np.int32)
        desc = PlanDesc(
            plan.nv, plan.ng, plan.nc, max(1, len(plan.slots.exprs)),
            table.shape[0], table.ctypes.data,
            len(P_terms), P_terms.ctypes.data, plan.DP.data_ptr() if plan.DP is not None else 0,
            len(G_terms), G_terms.ctypes.data, plan.DG.data_ptr() if plan.DG is not None else 0,
            len(Y_terms), Y_terms.ctypes.data,
            segs.shape[0], segs.ctypes.data, val.shape[0], val.ctypes.data, col.ctypes.data,
        )
        handle = ctypes.c_void_p()
        check(lib().mpci_plan_create(ctypes.byref(desc), ctypes.byref(handle)))
        self.handle = handle
        self.launches = lib().mpci_assemble_launches(handle)

    def __del__(self):
        handle, self.handle = getattr(self, "handle", None), None
        if handle and _lib is not None:
            _lib.mpci_plan_destroy(handle)


class DeviceBackend:
    """The ``ops`` interface of plan.py on CUDA tensors (float64)."""

    def __init__(self, device=None):
        require_gpu()
        torch = _torch()
        self.torch = torch
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)

    # -- memory (plumbing) ---------------------------------------------------
    def upload(self, array):
        return self.torch.from_numpy(np.ascontiguousarray(array, dtype=np.float64)).to(self.device)

    def zeros(self, rows, cols):
        return self.torch.zeros([rows, cols], dtype=self.torch.float64, device=self.device)

    def download(self, H):
        return H.cpu().numpy()

    download3 = download

    def scan_rows(self, DY):
        """plan._scan_qh_rows on the device: the sparse rows (segs, val, col) of the stacked
        constants DY [K, rows, cols] -- which entries are non-zero is found where they live,
        only those (about 1 % for the biped) are downloaded."""
        from . import plan as _plan

        by_row = DY.permute(1, 0, 2)  # entries ordered by (row, term, column)
        mask = by_row != 0
        where = mask.nonzero()
        val = by_row[mask].cpu().numpy().astype(np.float64)
        col = where[:, 2].to(self.torch.int32).cpu().numpy()
        counts = mask.sum(dim=2).cpu().numpy()
        return _plan._segments(counts, DY.shape[0]), val, col

    def rows(self, H, index):
        index = list(index)
        if index and index == list(range(index[0], index[0] + len(index))):
            return H[index[0] : index[0] + len(index)]
        idx = self.torch.tensor(index, dtype=self.torch.long, device=self.device)
        return H.index_select(0, idx)

    def sub(self, H, r0, r1, c0, c1):
        return H[r0:r1, c0:c1]

    def stack(self, handles):
        return self.torch.stack([h.contiguous() for h in handles]).contiguous()

    def gather_flat(self, D, index):
        """D [K, rows, cols] -> [K, len(index)] entries at flat positions (a copy, no arithmetic)."""
        idx = self.torch.from_numpy(np.ascontiguousarray(index, dtype=np.int64)).to(self.device)
        return D.reshape(D.shape[0], -1).index_select(1, idx).contiguous()

    # -- arithmetic (kernels of csrc/mpci_linalg.cu) ---------------------------
    @staticmethod
    def _ld(H):
        if H.dim() != 2 or (H.shape[1] > 1 and H.stride(1) != 1):
            raise ValueError("expected a row-major 2-D view")
        return max(H.stride(0), H.shape[1], 1) if H.shape[0] > 1 else max(H.shape[1], 1)

    def gemm(self, A, B, transA=False):
        M, K = (A.shape[1], A.shape[0]) if transA else (A.shape[0], A.shape[1])
        if B.shape[0] != K:
            raise ValueError(
                "matmul: mismatch in the contracted dimension ({} and {})".format(K, B.shape[0])
            )
        N = B.shape[1]
        C = self.torch.empty([M, N], dtype=self.torch.float64, device=self.device)
        check(
            lib().mpci_gemm(
                1 if transA else 0, M, N, K, 1.0, _ptr(A), self._ld(A), 0, _ptr(B), self._ld(B), 0,
                0.0, _ptr(C), max(N, 1), 0, 1, _stream_ptr(),
            )
        )
        return C

    def axpy(self, dst, src, alpha, rowscale=None):
        rows, cols = dst.shape
        if src.shape[1] != cols or src.shape[0] not in (1, rows):
            raise ValueError(
                "operands could not be broadcast together with shapes {} {}".format(
                    tuple(dst.shape), tuple(src.shape)
                )
            )
        scale = None
        if rowscale is not None:
            scale = self.upload(np.asarray(rowscale, dtype=np.float64).reshape(-1))
            if scale.shape[0] not in (1, rows):
                raise ValueError("row multipliers do not match the {} rows".format(rows))
        check(
            lib().mpci_axpy2d(
                rows, cols, float(alpha), _ptr(scale), 0 if scale is None else scale.shape[0],
                _ptr(src), self._ld(src), src.shape[0], _ptr(dst), self._ld(dst), _stream_ptr(),
            )
        )

    # -- the hot path ----------------------------------------------------------
    def device_plan(self, plan):
        if getattr(plan, "_device_plan", None) is None:
            plan._device_plan = DevicePlan(plan)
        return plan._device_plan

    def assemble(self, plan, scalars, given, batch, out=None):
        torch = self.torch
        dp = self.device_plan(plan)
        if not torch.is_tensor(scalars):
            scalars = torch.from_numpy(np.ascontiguousarray(scalars, dtype=np.float64)).to(self.device)
        if not torch.is_tensor(given):
            given = torch.from_numpy(np.ascontiguousarray(given, dtype=np.float64)).to(self.device)
        scalars = scalars.contiguous()
        given = given.contiguous().reshape(batch, plan.ng)
        if out is None:
            out = QPBatch(batch, plan.nv, plan.nc, self.device)
        elif (out.batch, out.nv, out.nc) != (batch, plan.nv, plan.nc):
            raise ValueError("the output batch does not match the plan")
        check(
            lib().mpci_assemble(
                dp.handle, _ptr(scalars), _ptr(given), _ptr(out.P), _ptr(out.q), _ptr(out.G),
                _ptr(out.h), batch, _stream_ptr(),
            )
        )
        return out

    def assemble_single(self, plan, given):
        """One instance, NumPy in and NumPy out -- the reference's own call pattern
        (body.py:333-348 once per control tick).  The plan keeps one device buffer
        [given | P | q | G | h] and its pinned host mirror, so a call is one upload,
        the three kernels and ONE download; the four arrays returned are fresh copies,
        as the reference's are."""
        torch = self.torch
        dp = self.device_plan(plan)
        nv, nc, ng = plan.nv, plan.nc, plan.ng
        ws = getattr(plan, "_single_ws", None)
        if ws is None:
            sizes = [ng, nv * nv, nv, nc * nv, nc]
            starts = np.concatenate([[0], np.cumsum([n + (n & 1) for n in sizes])])  # 16-byte aligned
            dev = torch.empty([int(starts[-1]) + 2], dtype=torch.float64, device=self.device)
            host = torch.empty([int(starts[-1]) + 2], dtype=torch.float64).pin_memory()
            ones = torch.ones([1, max(1, len(plan.slots.exprs))], dtype=torch.float64, device=self.device)
            ws = plan._single_ws = (dev, host, [int(x) for x in starts], sizes, ones)
        dev, host, starts, sizes, ones = ws
        part = [dev[a:a + n] for a, n in zip(starts, sizes)]
        host[:ng] = torch.from_numpy(np.ascontiguousarray(given, dtype=np.float64).reshape(ng))
        part[0].copy_(host[:ng], non_blocking=True)
        check(lib().mpci_assemble(dp.handle, _ptr(ones), _ptr(part[0]), _ptr(part[1]), _ptr(part[2]),
                                  _ptr(part[3]), _ptr(part[4]), 1, _stream_ptr()))
        host[starts[1]:].copy_(dev[starts[1]:], non_blocking=True)
        torch.cuda.current_stream().synchronize()
        arr = host.numpy()
        P, q, G, h = (np.array(arr[a:a + n]) for a, n in zip(starts[1:], sizes[1:]))
        return G.reshape(nc, nv), h.reshape(nc, 1), P.reshape(nv, nv), q.reshape(nv, 1)

    def preview(self, pplan, scalars, given_optim, batch, out=None):
        """Rows of a PreviewPlan for a batch: [batch, total rows] (device tensor; written
        into ``out`` when given -- a contiguous [batch, total rows] CUDA tensor)."""
        torch = self.torch
        dp = self.device_plan(pplan)
        scalars = self._to_device(scalars)
        if out is not None and (tuple(out.shape) != (batch, pplan.nv) or not out.is_contiguous()):
            raise ValueError("out must be a contiguous [{}, {}] tensor".format(batch, pplan.nv))
        values = torch.empty([batch, pplan.nv], dtype=torch.float64, device=self.device) if out is None else out
        null = ctypes.c_void_p(0)
        check(lib().mpci_assemble(dp.handle, _ptr(scalars), _ptr(given_optim), null, _ptr(values),
                                  null, null, batch, _stream_ptr()))
        return values

    def box_bounds(self, box, h, nv):
        """mpci_box_bounds for a ``plan.BoxRows`` and the h of a batch ([B, nc] CUDA tensor)."""
        torch = self.torch
        if box._device is None:
            box._device = (torch.from_numpy(box.col_ptr).to(self.device), torch.from_numpy(box.row).to(self.device),
                           torch.from_numpy(box.coef).to(self.device))
        col_ptr, row, coef = box._device
        batch = h.shape[0]
        h = h.contiguous()
        lb = torch.empty([batch, nv], dtype=torch.float64, device=self.device)
        ub = torch.empty([batch, nv], dtype=torch.float64, device=self.device)
        check(lib().mpci_box_bounds(_ptr(h), h.shape[1], nv, ctypes.c_void_p(col_ptr.data_ptr()),
                                    ctypes.c_void_p(row.data_ptr()), _ptr(coef), _ptr(lb), _ptr(ub), batch,
                                    _stream_ptr()))
        return lb, ub

    def goal_distance(self, values, aims, rows, n_axes):
        """values [batch, n_axes * rows], aims [batch, n_axes] or [1, n_axes] -> [batch]."""
        torch = self.torch
        batch = values.shape[0]
        aims = self._to_device(aims).reshape(-1, n_axes)
        out = torch.empty([batch], dtype=torch.float64, device=self.device)
        stride = n_axes if aims.shape[0] == batch and batch > 1 else 0
        if aims.shape[0] not in (1, batch):
            raise ValueError("aims must have 1 or {} rows".format(batch))
        check(lib().mpci_goal_distance(_ptr(values), _ptr(aims), stride, rows, n_axes, _ptr(out),
                                       batch, _stream_ptr()))
        return out

    def _to_device(self, x):
        torch = self.torch
        if not torch.is_tensor(x):
            x = torch.from_numpy(np.ascontiguousarray(x, dtype=np.float64))
        return x.to(device=self.device, dtype=torch.float64).contiguous()

    def compact(self, plan, upper_P=False):
        """Registers the CSC constant sets with the device plan (once per layout)."""
        from . import plan as _plan

        key = "_csc_upper" if upper_P else "_csc"
        if getattr(plan, key, None) is None:
            setattr(plan, key, _plan.csc_patterns(plan, self, upper_P))
        pP, pG, cP, cG = getattr(plan, key)
        dp = self.device_plan(plan)
        if getattr(dp, "compact_key", None) != key:
            check(lib().mpci_plan_set_compact(dp.handle, _ptr(cP), pP.padded, _ptr(cG), pG.padded))
            dp.compact_key = key
            dp.keep_compact = (cP, cG)
        return pP, pG

    def assemble_csc(self, plan, scalars, given, batch, out=None, upper_P=False):
        torch = self.torch
        pP, pG = self.compact(plan, upper_P)
        dp = self.device_plan(plan)
        if not torch.is_tensor(scalars):
            scalars = torch.from_numpy(np.ascontiguousarray(scalars, dtype=np.float64)).to(self.device)
        if not torch.is_tensor(given):
            given = torch.from_numpy(np.ascontiguousarray(given, dtype=np.float64)).to(self.device)
        scalars = scalars.contiguous()
        given = given.contiguous().reshape(batch, plan.ng)
        if out is None:
            out = QPBatchCSC(batch, plan.nv, plan.nc, pP, pG, self.device)
        elif (out.batch, out.nv, out.nc, out.P_pattern.padded, out.G_pattern.padded) != (
                batch, plan.nv, plan.nc, pP.padded, pG.padded):
            raise ValueError("the output batch does not match the plan")
        check(
            lib().mpci_assemble_compact(
                dp.handle, _ptr(scalars), _ptr(given), _ptr(out._P), _ptr(out.q), _ptr(out._G),
                _ptr(out.h), batch, _stream_ptr(),
            )
        )
        return out

    def assemble_csc_host(self, plan, scalars, given, P, q, G, h, batch, chunk, upper_P=False):
        """Host buffers; ``P`` [batch, P_pattern.padded], ``G`` [batch, G_pattern.padded]."""
        self.compact(plan, upper_P)
        dp = self.device_plan(plan)

        def p(x):
            if x is None:
                return ctypes.c_void_p(0)
            return ctypes.c_void_p(x.data_ptr() if hasattr(x, "data_ptr") else x.ctypes.data)

        check(lib().mpci_assemble_compact_host(dp.handle, p(scalars), p(given), p(P), p(q), p(G), p(h), batch, chunk))

    def assemble_host(self, plan, scalars, given, P, q, G, h, batch, chunk):
        """Host (ideally pinned) numpy / CPU-tensor buffers in and out."""
        dp = self.device_plan(plan)
        p = _host_ptr
        check(lib().mpci_assemble_host(dp.handle, p(scalars), p(given), p(P), p(q), p(G), p(h), batch, chunk))

    def assemble_host_stream(self, plan, scalars, given, P, q, G, h, batch, chunk, ring,
                             on_chunk=None, layout="dense"):
        """Logical batches larger than host memory: ``P, q, G, h`` hold ``ring`` chunks of
        ``chunk`` instances, chunk c lands in slot ``c % ring`` and
        ``on_chunk(first_instance, count, slot)`` is called when it is complete in host
        memory, before the slot is written again (mpci_assemble_host_stream).
        ``layout="csc"``: ``P`` / ``G`` take the CSC values (compact())."""
        dp = self.device_plan(plan)
        p = _host_ptr
        callback = ChunkFn(lambda _user, first, count, slot: on_chunk(first, count, slot)) \
            if on_chunk is not None else ChunkFn(0)
        fn = lib().mpci_assemble_host_stream
        if layout != "dense":
            self.compact(plan, layout == "csc_upper")
            fn = lib().mpci_assemble_compact_host_stream
        check(fn(dp.handle, p(scalars), p(given), p(P), p(q), p(G), p(h), batch, chunk, ring,
                 callback, None))


def _host_ptr(x):
    if x is None:
        return ctypes.c_void_p(0)
    return ctypes.c_void_p(x.data_ptr() if hasattr(x, "data_ptr") else x.ctypes.data)
