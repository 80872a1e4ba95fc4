"""NumPy statement of what ``mpci_lti_assemble`` computes from the lowered stage
tables -- TEST INFRASTRUCTURE ONLY (checks the host lowering of
mpc_interface_b200/lti.py on a machine without a GPU).  Formulas: include/mpci.h."""
import numpy as np


def evaluate(plan, A, B, x0, scalars):
    """One instance: A [n,n], B [n,m], x0 [n], scalars [S] -> (G, h, P, q)."""
    t = plan.tables()
    n, m, N = plan.n, plan.m, plan.N
    nv = m * N
    Gam = [B]
    for _ in range(N - 1):
        Gam.append(A @ Gam[-1])
    f, x = [], x0
    for _ in range(N):
        x = A @ x
        f.append(x)
    s1 = np.append(scalars, 1.0)
    tx, tu = t["Tx"] @ s1, t["Tu"] @ s1
    Qxx, Quu = t["Qxx"], t["Quu"]
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
    for g in range(t["grp_cx"].shape[0]):
        cx, cu = t["grp_cx"][g], t["grp_cu"][g]
        rows = np.einsum("i,kijp->kjp", cx, T).reshape([N, nv])
        rows = rows + np.kron(cu[None, :], np.eye(N))
        Gs.append(t["grp_arrow"][g][:, None] * rows)
        hs.append(t["grp_bound"][g] - t["grp_arrow"][g] * np.array([cx @ f[k] for k in range(N)]))
    G = np.vstack(Gs) if Gs else np.zeros([0, nv])
    h = np.concatenate(hs) if hs else np.zeros([0])
    return G, h, P, q
