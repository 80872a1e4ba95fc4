"""NumPy stand-in for the device backend -- TEST INFRASTRUCTURE ONLY.

Implements the small ``ops`` interface plan.py compiles against (upload,
zeros, rows, sub, gemm, axpy, stack, assemble) so that the plan compiler -- the
symbolic host logic -- can be checked against the oracle on a machine without
a GPU.  ``assemble`` restates what the three streaming kernels compute
(monomials, P/G streams, [q; h] products).  The product package never imports
this module; its only backend is the CUDA one (mpc_interface_b200/device.py).
"""
import numpy as np


class _Out:
    def __init__(self, P, q, G, h):
        self.P, self.q, self.G, self.h = P, q, G, h

    def P_numpy(self):
        return self.P

    def q_numpy(self):
        return self.q

    def G_numpy(self):
        return self.G

    def h_numpy(self):
        return self.h


class NumpyBackend:
    def upload(self, array):
        return np.array(array, dtype=np.float64)

    def zeros(self, rows, cols):
        return np.zeros([rows, cols])

    def download(self, H):
        return np.array(H)

    def download3(self, H):
        return np.array(H)

    def rows(self, H, index):
        return H[list(index)]

    def sub(self, H, r0, r1, c0, c1):
        return H[r0:r1, c0:c1]

    def gemm(self, A, B, transA=False):
        return (A.T if transA else A) @ B

    def axpy(self, dst, src, alpha, rowscale=None):
        term = alpha * src
        if rowscale is not None:
            term = np.asarray(rowscale)[:, None] * term
        dst += term

    def stack(self, handles):
        return np.stack(handles)

    def gather_flat(self, D, index):
        return D.reshape(D.shape[0], -1)[:, index].copy()

    # what mpci_assemble computes ------------------------------------------
    def assemble(self, plan, scalars, given, batch, out=None):
        table = plan.mono_table()
        coef = np.ones([batch, table.shape[0]])
        for j in range(table.shape[0]):
            for f in table[j]:
                if f >= 0:
                    coef[:, j] *= scalars[:, f]
        P = np.einsum("bk,kij->bij", coef[:, plan.P_terms], plan.DP)
        G = np.einsum("bk,kij->bij", coef[:, plan.G_terms], plan.DG)
        ghat = np.hstack([given.reshape(batch, plan.ng), np.ones([batch, 1])])
        y = np.einsum("bk,kij,bj->bi", coef[:, plan.Y_terms], plan.DY, ghat)
        return _Out(P, y[:, : plan.nv].copy(), G, y[:, plan.nv :].copy())
