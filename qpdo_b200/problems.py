"""Seeded synthetic QP generators for the five BASELINE.json configs.

Every generator returns a dict(Q, q, A, l, u[, extra]) with scipy CSC matrices (Q full symmetric;
the ABI layer keeps the lower triangle, stype=-1, like interfaces/mex/qpdo_mex.c:150) and float64
vectors, `min 1/2 x'Qx + q'x  s.t.  l <= Ax <= u` (README.md:3-10 of the reference).  The same
arrays are handed to the reference build and to the CUDA library, so both see identical bits.
"""
import numpy as np
import scipy.sparse as sp

INF = 1e20


def random_qp(n=1000, m=500, density=0.01, seed=1):
    """C1: random sparse convex QP (BASELINE.json configs[0])."""
    rng = np.random.default_rng(seed)
    noff = int(round(0.5 * density * n * n))
    r = rng.integers(0, n, noff)
    c = rng.integers(0, n, noff)
    keep = r != c
    r, c = r[keep], c[keep]
    lo = np.maximum(r, c)
    hi = np.minimum(r, c)
    L = sp.coo_matrix((rng.standard_normal(lo.size), (lo, hi)), shape=(n, n)).tocsc()
    L.sum_duplicates()
    S = L + L.T
    d = np.asarray(abs(S).sum(axis=1)).ravel() + rng.random(n)
    Q = (S + sp.diags(d)).tocsc()
    nA = int(round(density * m * n))
    A = sp.coo_matrix((rng.standard_normal(nA), (rng.integers(0, m, nA), rng.integers(0, n, nA))),
                      shape=(m, n)).tocsc()
    A.sum_duplicates()
    q = rng.standard_normal(n)
    l = -rng.random(m)
    u = rng.random(m)
    return dict(Q=Q, q=q, A=A, l=l, u=u)


def infeasibility_trio():
    """C2: the three known-answer QPs of examples/infeasibility_tests.m:9-75.
    Returns [(name, problem, expected_status)]."""
    out = []
    for name, (a, b, c), st in (("degenerate", (0.0, 3.0, 0.0), 1),
                                 ("primal_infeasible", (1.0, 3.0, 0.0), -3),
                                 ("dual_infeasible", (0.0, INF, -1.0), -4)):
        Q = sp.csc_matrix(np.array([[1.0, 0.0], [0.0, 0.0]]))
        A = sp.csc_matrix(np.array([[a, a], [1.0, 0.0], [0.0, 1.0]]))
        q = np.array([1.0, c])
        l = np.array([-INF, 1.0, 1.0])
        u = np.array([0.0, 3.0, b])
        out.append((name, dict(Q=Q, q=q, A=A, l=l, u=u), st))
    return out


def replicate_block_diag(p, k):
    """k block-diagonal copies of a QP (C2 'at size' variants)."""
    return dict(Q=sp.block_diag([p["Q"]] * k, format="csc"), q=np.tile(p["q"], k),
                A=sp.block_diag([p["A"]] * k, format="csc"), l=np.tile(p["l"], k),
                u=np.tile(p["u"], k))


def cont_like(N=300):
    """C3: Maros-Meszaros CONT-300-shaped boundary-control QP: N^2 states on a
    grid, N controls on one edge, 5-point-Laplacian equality rows + identity box rows."""
    h = 1.0 / (N + 1)
    ns = N * N
    n = ns + N
    idx = np.arange(ns).reshape(N, N)          # idx[i, j], i = row of grid
    rows, cols, vals = [], [], []

    def add(r, c, v):
        rows.append(np.asarray(r).ravel()); cols.append(np.asarray(c).ravel())
        vals.append(np.full(np.asarray(r).size, v))
    eq = idx  # equality row id == state id
    add(eq, idx, 4.0)
    add(eq[1:, :], idx[:-1, :], -1.0)
    add(eq[:-1, :], idx[1:, :], -1.0)
    add(eq[:, 1:], idx[:, :-1], -1.0)
    add(eq[:, :-1], idx[:, 1:], -1.0)
    add(eq[0, :], ns + np.arange(N), -1.0)     # control couples on the i=0 edge
    add(ns + np.arange(n), np.arange(n), 1.0)  # identity box rows
    m = ns + n
    A = sp.coo_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))),
                      shape=(m, n)).tocsc()
    l = np.empty(m); u = np.empty(m)
    l[:ns] = h * h; u[:ns] = h * h
    l[ns:ns + ns] = -INF; u[ns:ns + ns] = 0.371
    l[ns + ns:] = 0.0; u[ns + ns:] = 1.0
    qd = np.concatenate([np.full(ns, h * h), np.full(N, 1e-2 * h)])
    Q = sp.diags(qd).tocsc()
    g = (np.arange(N) + 1) * h
    X, Y = np.meshgrid(g, g, indexing="ij")
    yd = 1.0 + 2.0 * (X * (X - 1.0) + Y * (Y - 1.0))
    q = np.concatenate([-(h * h) * yd.ravel(), np.zeros(N)])
    return dict(Q=Q, q=q, A=A, l=l, u=u)


def lasso_dual(n=200000, m=100000, per_feature=10, window=500, seed=4, npath=20, shrink=0.8):
    """C4: Lasso as a QP in its dual form  min 1/2|v|^2 - b'v  s.t. -lam <= D'v <= lam, with a
    banded-random design D (n samples x m features).  Returns the problem at lam_max plus the
    regularisation path `lams`; each path step is update_bounds(+-lam) + warm_start + solve."""
    rng = np.random.default_rng(seed)
    stride = n / float(m)
    k = np.repeat(np.arange(m), per_feature)
    centre = (k * stride).astype(np.int64)
    off = rng.integers(-window // 2, window // 2, k.size)
    s = np.clip(centre + off, 0, n - 1)
    v = rng.standard_normal(k.size) / np.sqrt(per_feature)
    D = sp.coo_matrix((v, (s, k)), shape=(n, m)).tocsc()
    D.sum_duplicates()
    x0 = np.zeros(m)
    nz = rng.choice(m, max(1, m // 100), replace=False)
    x0[nz] = rng.standard_normal(nz.size)
    b = D @ x0 + 0.01 * rng.standard_normal(n)
    A = D.T.tocsc()
    lam_max = float(np.abs(A @ b).max())
    lams = lam_max * shrink ** np.arange(npath)
    Q = sp.identity(n, format="csc")
    return dict(Q=Q, q=-b, A=A, l=np.full(m, -lams[0]), u=np.full(m, lams[0]), lams=lams)


# ----------------------------------------------------------------------------- C5: MPC batch

MPC_T, MPC_NX, MPC_NU, MPC_NXB = 10, 8, 2, 5     # horizon, states, inputs, boxed states/stage


def mpc_pattern():
    """Common sparsity of the C5 MPC QPs: n = T(nu+nx) = 100, m = T*nx + T*nu + T*5 = 150.
    Variable order (u_0, x_1, u_1, x_2, ..., u_{T-1}, x_T).  Returns (rows, cols, kind, blk) of
    A's nonzeros in CSC order, where kind: 0 = +1 (x_{t+1}), 1 = -A_d[i,j], 2 = -B_d[i,j],
    3 = +1 (box row)."""
    T, nx, nu, nb = MPC_T, MPC_NX, MPC_NU, MPC_NXB
    ns = nx + nu
    n = T * ns
    ent = []
    for t in range(T):
        ucol = t * ns
        xcol = t * ns + nu            # x_{t+1}
        for i in range(nx):
            r = t * nx + i
            ent.append((r, xcol + i, 0, 0))
            for j in range(nu):
                ent.append((r, ucol + j, 2, i * nu + j))
            if t > 0:
                pcol = (t - 1) * ns + nu
                for j in range(nx):
                    ent.append((r, pcol + j, 1, i * nx + j))
    r0 = T * nx
    for t in range(T):
        for j in range(nu):
            ent.append((r0 + t * nu + j, t * ns + j, 3, 0))
    r1 = r0 + T * nu
    for t in range(T):
        for j in range(nb):
            ent.append((r1 + t * nb + j, t * ns + nu + j, 3, 0))
    m = r1 + T * nb
    ent.sort(key=lambda e: (e[1], e[0]))
    a = np.array(ent, dtype=np.int64)
    return n, m, a[:, 0], a[:, 1], a[:, 2], a[:, 3]


def mpc_nominal(seed=5):
    rng = np.random.default_rng(seed)
    nx, nu = MPC_NX, MPC_NU
    Ad = rng.standard_normal((nx, nx))
    Ad *= 1.02 / np.abs(np.linalg.eigvals(Ad)).max()
    Bd = rng.standard_normal((nx, nu))
    return Ad, Bd


def mpc_batch(B, first=0, seed=5):
    """C5: instances first..first+B-1 of the MPC family (seed = instance id).  Returns the common
    CSC pattern of A and per-instance value arrays:
       Ap(n+1), Ai(nnz)  int32;  Ax (B,nnz);  Qd (n,) common diagonal;  q,(B,n); l,u (B,m)."""
    n, m, rows, cols, kind, blk = mpc_pattern()
    Ad0, Bd0 = mpc_nominal(seed)
    nx, nu, T, nb = MPC_NX, MPC_NU, MPC_T, MPC_NXB
    ns = nx + nu
    nnz = rows.size
    Ap = np.zeros(n + 1, dtype=np.int32)
    np.add.at(Ap, cols + 1, 1)
    Ap = np.cumsum(Ap).astype(np.int32)
    Ai = rows.astype(np.int32)
    Ax = np.empty((B, nnz)); q = np.empty((B, n)); l = np.empty((B, m)); u = np.empty((B, m))
    qd = np.tile(np.concatenate([np.full(nu, 0.1), np.ones(nx)]), T)
    for b in range(B):
        rng = np.random.default_rng(1000003 * seed + first + b)
        Ad = Ad0 * (1.0 + 0.05 * rng.uniform(-1, 1, Ad0.shape))
        Bd = Bd0 * (1.0 + 0.05 * rng.uniform(-1, 1, Bd0.shape))
        x0 = rng.standard_normal(nx)
        v = np.where(kind == 0, 1.0, np.where(kind == 3, 1.0, 0.0))
        k1 = kind == 1
        v[k1] = -Ad.ravel()[blk[k1]]
        k2 = kind == 2
        v[k2] = -Bd.ravel()[blk[k2]]
        Ax[b] = v
        q[b] = 0.1 * rng.standard_normal(n)
        rhs = np.zeros(T * nx)
        rhs[:nx] = Ad @ x0
        l[b, :T * nx] = rhs; u[b, :T * nx] = rhs
        l[b, T * nx:T * nx + T * nu] = -1.0; u[b, T * nx:T * nx + T * nu] = 1.0
        l[b, T * nx + T * nu:] = -2.0; u[b, T * nx + T * nu:] = 2.0
    return dict(n=n, m=m, Ap=Ap, Ai=Ai, Ax=Ax, Qd=qd, q=q, l=l, u=u)


def mpc_instance(batch, b):
    """One C5 instance as a standalone problem dict (for driving it through qpdo.h)."""
    n, m = batch["n"], batch["m"]
    A = sp.csc_matrix((batch["Ax"][b], batch["Ai"], batch["Ap"]), shape=(m, n))
    return dict(Q=sp.diags(batch["Qd"]).tocsc(), q=batch["q"][b].copy(), A=A,
                l=batch["l"][b].copy(), u=batch["u"][b].copy())
