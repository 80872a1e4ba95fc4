"""NumPy statement of what ``mpci_lti_assemble`` / ``mpci_lti_assemble_ex`` compute from
the lowered stage tables -- TEST INFRASTRUCTURE ONLY (checks the host lowering of
mpc_interface_b200/lti.py on a machine without a GPU).  Formulas: include/mpci.h."""
import numpy as np


def evaluate(plan, A, B, x0, scalars, instance=None):
    """One instance: the PLANT's A [n0,n0], B [n0,m0] (one axis), x0 [n] in the given order,
    scalars [S] -> (G, h, P, q).  ``instance`` = the per-instance tables of this instance
    (``plan.instance_tables`` row: Qxx, Quu, tx, tu, cx, cu, bound) for a plan with Param
    coefficients."""
    t = plan.tables()
    A, B = plan.expand_numpy(A, B)
    n, m, N = plan.n, plan.m, plan.N
    nv = m * N
    Gam = [B]
    for _ in range(N - 1):
        Gam.append(A @ Gam[-1])
    f, x = [], x0
    for _ in range(N):
        x = A @ x
        f.append(x)
    if instance is None:
        s1 = np.append(scalars, 1.0)
        tx, tu = t["Tx"] @ s1, t["Tu"] @ s1
        Qxx, Quu, cxs, cus, bounds = t["Qxx"], t["Quu"], t["grp_cx"], t["grp_cu"], t["grp_bound"]
    else:
        tx, tu, Qxx, Quu = instance["tx"], instance["tu"], instance["Qxx"], instance["Quu"]
        cxs, cus, bounds = instance["cx"], instance["cu"], instance["bound"]
    Qs = 0.5 * (Qxx + Qxx.T)
    # dense Toeplitz operator: rows (k, i), columns (j, p)
    T = np.zeros([N, n, m, N])
    for k in range(N):
        for p in range(k + 1):
            T[k, :, :, p] = Gam[k - p]
    T2 = T.reshape([N * n, nv])
    P = T2.T @ np.kron(np.eye(N), Qxx) @ T2 + np.kron(Quu, np.eye(N))
    r = np.stack([Qs @ f[k] - tx for k in range(N)]).reshape([N * n])
    q = T2.T @ r - np.repeat(tu, N)
    Gs, hs = [], []
    for g in range(cxs.shape[0]):
        cx, cu = cxs[g], cus[g]
        k0, cnt = int(t["grp_k0"][g]), int(t["grp_rows"][g])
        rows = np.einsum("i,kijp->kjp", cx, T).reshape([N, nv])
        rows = rows + np.kron(cu[None, :], np.eye(N))
        arrow = t["grp_arrow"][g]
        Gs.append((arrow[:, None] * rows)[k0:k0 + cnt])
        hs.append((bounds[g] - arrow * np.array([cx @ f[k] for k in range(N)]))[k0:k0 + cnt])
    G = np.vstack(Gs) if Gs else np.zeros([0, nv])
    h = np.concatenate(hs) if hs else np.zeros([0])
    return G, h, P, q
