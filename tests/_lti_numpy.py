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
    else:  # whatever does not vary between the instances stays in the plan's tables
        tx, tu = instance["tx"], instance["tu"]
        Qxx, Quu = instance.get("Qxx", t["Qxx"]), instance.get("Quu", t["Quu"])
        cxs, cus = instance.get("cx", t["grp_cx"]), instance.get("cu", t["grp_cu"])
        bounds = instance.get("bound", t["grp_bound"])
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


def evaluate_embedded(plan, A, B, given, scalars, instance=None, coupled=None, extras=None):
    """An embedded plan (variables / givens / rows outside the plant, lti.lower): what the
    fused kernel writes at its place in the larger matrices plus what the coupling kernel
    (csrc/mpci_couple.cu) adds, from the plan's tables.  ``given`` [given_len];
    ``coupled`` = (cost_cx [nC, n], cost_cu [nC, m]) of this instance for a per-instance plan;
    ``extras`` = this instance's row of ``plan.coupled_extras`` (cost_w [nC], cost_aim [nC],
    Pzz [nz, nz], Gz [nc, nz], h_val [nnz of plan.h_mask()]; each may be None)."""
    given = np.asarray(given, dtype=np.float64)
    x0 = given[plan.x0_index]
    Gc, hc, Pc, qc = evaluate(plan, A, B, x0, scalars, instance)
    n, m, N, nv, nc = plan.n, plan.m, plan.N, plan.nv, plan.nc
    cols = np.concatenate([np.arange(o, o + N) for o in plan.var_off])
    z = np.array(plan.z_col, dtype=int)
    P, q = np.zeros([nv, nv]), np.zeros([nv])
    G, h = np.zeros([nc, nv]), np.zeros([nc])
    P[np.ix_(cols, cols)], q[cols] = Pc, qc
    at = 0
    for g, (_form, _arrow, _parts, _k0, cnt) in enumerate(plan.groups):
        rows = plan.group_row0[g] + np.arange(cnt)
        G[np.ix_(rows, cols)] = Gc[at:at + cnt]
        h[rows] = hc[at:at + cnt]
        at += cnt
    g1 = np.append(given, 1.0)
    w_b, aim_b, Pzz_b, Gz_b, hval_b = (tuple(extras) + (None,) * 5)[:5] if extras is not None else (None,) * 5
    G[:, z] = plan.Gz if Gz_b is None else Gz_b
    Hg = plan.Hg
    if hval_b is not None:
        Hg = np.zeros_like(plan.Hg)
        Hg[plan.h_mask()] = hval_b
    extra = Hg @ g1
    h = np.where(plan.row_kind == 1, extra, h + extra)
    P[np.ix_(z, z)] = plan.Pzz if Pzz_b is None else Pzz_b
    Ae, Be = plan.expand_numpy(A, B)
    Gam = [Be]
    for _ in range(N - 1):
        Gam.append(Ae @ Gam[-1])
    f, x = [], x0
    for _ in range(N):
        x = Ae @ x
        f.append(x)
    tab = plan.embed_tables()
    for c, (w, _fv, Mz, E) in enumerate(plan.coupled):
        cx = tab["cost_cx"][c] if coupled is None else coupled[0][c]
        cu = tab["cost_cu"][c] if coupled is None else coupled[1][c]
        T = np.zeros([N, m, N])
        for k in range(N):
            for p in range(k + 1):
                T[k, :, p] = cx @ Gam[k - p]
            T[k, :, k] += cu
        T = T.reshape([N, m * N])
        fr = np.array([cx @ f[k] for k in range(N)])
        ee = E @ g1 - (aim_b[c] if aim_b is not None else 0.0)
        w = tab["cost_w"][c] if w_b is None else w_b[c]
        P[np.ix_(cols, z)] += w * T.T @ Mz
        P[np.ix_(z, cols)] += w * Mz.T @ T
        q[cols] += w * T.T @ ee
        q[z] += w * Mz.T @ (fr + ee)
    return G, h, P, q


def add_scheduled(plan, A, B, given, terms, P, q):
    """What the stage of csrc/mpci_sched.cu adds for the plan's scheduled costs (include/mpci.h:
    mpci_lti_sched_desc), from the tables the stage gets: ``terms`` [T, 3 + 2n + 2m] of this
    instance (``plan.sched_tables`` row) and ``plan.sched_desc_tables``, ``given`` the formulation's given vector (the plant's x0 alone in the plain layout).
    Returns (P, q) with the terms added."""
    n, m, N, nv = plan.n, plan.m, plan.N, plan.nv
    given = np.asarray(given, dtype=np.float64)
    x0 = given[plan.x0_index] if plan.embedded else given
    Ae, Be = plan.expand_numpy(A, B)
    W, xs, x = [Be], [], x0
    for _ in range(N - 1):
        W.append(Ae @ W[-1])
    for _ in range(N):
        x = Ae @ x
        xs.append(x)
    voff = plan.var_off if plan.embedded else [j * N for j in range(m)]
    tab = plan.sched_desc_tables()
    P, q = P.copy(), q.copy()
    for t in range(len(tab["row_ptr"]) - 1):
        w, aim, caim = terms[t, :3]
        vcx, vcu = terms[t, 3:3 + n], terms[t, 3 + n:3 + n + m]
        ccx, ccu = terms[t, 3 + n + m:3 + 2 * n + m], terms[t, 3 + 2 * n + m:]
        for r in range(tab["row_ptr"][t], tab["row_ptr"][t + 1]):
            rows = []
            own = tab["c_ent_ptr"] is None
            for cx, cu, Mz, Mg, ent in ((vcx, vcu, tab["v_Mz"], tab["v_Mg"], ("ent_ptr", "ent_k", "ent_val")),
                                        (ccx, ccu, tab["c_Mz"] if tab["c_Mz"] is not None else tab["v_Mz"],
                                         tab["c_Mg"] if tab["c_Mg"] is not None else tab["v_Mg"],
                                         ("ent_ptr", "ent_k", "ent_val") if own else ("c_ent_ptr", "c_ent_k", "c_ent_val"))):
                v, free = np.zeros([nv]), 0.0
                for e in range(tab[ent[0]][r], tab[ent[0]][r + 1]):
                    k, val = int(tab[ent[1]][e]), tab[ent[2]][e]
                    for j in range(m):
                        for p in range(k + 1):
                            v[voff[j] + p] += val * (cx @ W[k - p][:, j] + (cu[j] if p == k else 0.0))
                    free += val * (cx @ xs[k])
                if Mz is not None:
                    v[np.array(plan.z_col, dtype=int)] = Mz[r]
                free += Mg[r] @ given if Mg is not None else 0.0
                rows.append((v, free))
            (v, fv), (c, fc) = rows
            P += w * np.outer(v, c)
            q += w / 2 * (v * (fc - caim) + c * (fv - aim))
    return P, q


def add_limit_rows(plan, A, B, given, terms, G, h):
    """What mpci_lti_sched_rows adds for the limits with ``L`` (``plan.lrows``): the plant part of
    their rows of G and of h, from ``terms`` [T, 3 + 2n + 2m] of this instance
    (``plan.lrow_tables`` row) and ``plan.lrow_desc_tables``.  Returns (G, h)."""
    n, m, N, nv = plan.n, plan.m, plan.N, plan.nv
    given = np.asarray(given, dtype=np.float64)
    x0 = given[plan.x0_index]
    Ae, Be = plan.expand_numpy(A, B)
    W, xs, x = [Be], [], x0
    for _ in range(N - 1):
        W.append(Ae @ W[-1])
    for _ in range(N):
        x = Ae @ x
        xs.append(x)
    tab = plan.lrow_desc_tables()
    G, h = G.copy(), h.copy()
    for t in range(len(tab["row_ptr"]) - 1):
        w, cx, cu = terms[t, 0], terms[t, 3:3 + n], terms[t, 3 + n:3 + n + m]
        for r in range(tab["row_ptr"][t], tab["row_ptr"][t + 1]):
            v, free = np.zeros([nv]), 0.0
            for e in range(tab["ent_ptr"][r], tab["ent_ptr"][r + 1]):
                k, val = int(tab["ent_k"][e]), tab["ent_val"][e]
                for j in range(m):
                    for p in range(k + 1):
                        v[plan.var_off[j] + p] += val * (cx @ W[k - p][:, j] + (cu[j] if p == k else 0.0))
                free += val * (cx @ xs[k])
            G[tab["row_out"][r]] += w * v
            h[tab["row_out"][r]] -= w * free
    return G, h
