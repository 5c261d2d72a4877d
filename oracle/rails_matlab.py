"""TEST INFRASTRUCTURE -- numpy restatement of the reference's MATLAB solver,
/root/reference/matlab/RAILSsolver.m, for small dense problems.

Used by tests/ only, as the checker for the MATLAB-only features the product
loop offers as options: M-orthogonalisation, null-space
projection, the relative restart cut, restart_upon_start, and orthogonalising
the expansion candidates before they are selected.  The product never imports
this file.

PARITY UNPINNED: neither MATLAB nor Octave exists on this image, and the
reference's MATLAB tests hold inequalities only (matlab/test/test_opts.m,
test_Laplace.m: res < tol, explicit residual < tol, iter < n - 10, size(V,2)
after a restart).  Those inequalities are what tests/test_matlab_oracle.py pins
this restatement to; agreement between it and the product is then a cross-check
of two independent implementations, not a pin against MATLAB output.

Deliberate restatement choices (each marked below):
  * `eigs(@R, n, nev, 'lm')` (ARPACK, RAILSsolver.m:388-392) is replaced by the
    exact nev largest-magnitude eigenpairs of the same operator, computed from
    its low-rank factor R = F S F'.  ARPACK converges to these; its start vector
    is random, so no MATLAB run is reproducible beyond that anyway.
  * `lyap` (matlab/mex/lyap.c:95-150: SLICOT sb03md / sg03ad) is scipy's
    Bartels-Stewart solver; the generalised equation is reduced with E^-1.
  * A may be a matrix or a function handle (RAILSsolver.m:262-270: `A(x)`), as
    RAILSschur returns one; opts['projection_method'] and opts['Ainv'] follow
    :7-24,280-308,520-524.
"""
import numpy as np
import scipy.linalg


def lyap(A, C, E=None):
    """X with A X + X A' + C = 0, or A X E' + E X A' + C = 0 (lyap.c:118-137)."""
    if E is None:
        return scipy.linalg.solve_continuous_lyapunov(A, -C)
    Ei = np.linalg.inv(E)
    return scipy.linalg.solve_continuous_lyapunov(Ei @ A, -(Ei @ C @ Ei.T))


def orth(W):
    """MATLAB orth: left singular vectors above the rank tolerance."""
    if W.shape[1] == 0:
        return W.copy()
    U, s, _ = np.linalg.svd(W, full_matrices=False)
    if s.size == 0:
        return U
    tol = max(W.shape) * np.spacing(s[0])
    return U[:, :int(np.sum(s > tol))]


def morth(W, M=None, nullspace=None, V=None, fast=False):
    """Morth, RAILSsolver.m:538-617: orthonormalise the columns of W against
    `nullspace` and V (Euclidean or M inner product) and append them to V."""
    m = W.shape[0]
    V = np.zeros((m, 0)) if V is None else V
    has_null = nullspace is not None and nullspace.shape[1] > 0
    num_iter, tol = 1, 1e-8                                   # :553-554

    def nullspace_op(x):                                      # :599-616
        y = x
        if has_null:
            for ii in range(nullspace.shape[1]):
                n_i = nullspace[:, ii]
                y = y - n_i * (n_i @ y) if M is None else y - n_i * (n_i @ (M @ y))
        return y

    if M is None:
        if fast:                                              # :557-565
            if has_null:
                W = W - nullspace @ (nullspace.T @ W)
            if V.shape[1] > 0:
                W = W - V @ (V.T @ W)
            return np.hstack([V, orth(W)])
        for i in range(W.shape[1]):                           # :567-582 modified GS
            v1 = W[:, i] / np.linalg.norm(W[:, i])
            for _ in range(num_iter):
                v1 = nullspace_op(v1)
                for j in range(V.shape[1]):
                    v1 = v1 - V[:, j] * (V[:, j] @ v1)
            nrm = np.linalg.norm(v1)
            if nrm < tol:
                continue
            V = np.hstack([V, (v1 / nrm)[:, None]])
        return V
    for i in range(W.shape[1]):                               # :584-597 M inner product
        v1 = W[:, i] / np.linalg.norm(W[:, i])
        for _ in range(num_iter):
            v1 = nullspace_op(v1)
            for j in range(V.shape[1]):
                v1 = v1 - V[:, j] * (V[:, j] @ (M @ v1))
        nrm = np.sqrt(v1 @ (M @ v1))
        if nrm < tol:
            continue
        V = np.hstack([V, (v1 / nrm)[:, None]])
    return V


def residual_eigs(AV, MV, T, B, nev):
    """The nev largest-magnitude eigenpairs of
    R = AV T MV' + MV T AV' + B B'  (RAILSsolver.m:388-392), exactly: R = F S F'
    with F = [MV, AV, B]."""
    k, n_b = AV.shape[1], B.shape[1]
    F = np.hstack([MV, AV, B])
    c = 2 * k + n_b
    S = np.zeros((c, c))
    S[:k, k:2 * k] = T
    S[k:2 * k, :k] = T
    S[2 * k:, 2 * k:] = np.eye(n_b)
    Q, R = np.linalg.qr(F)
    E = R @ S @ R.T
    w, Y = np.linalg.eigh(0.5 * (E + E.T))
    order = np.argsort(-np.abs(w), kind="stable")[:nev]
    return Q @ Y[:, order], w[order]


def rails_solver(A, M, B, maxit=100, tol=1e-4, opts=None, V0=None):
    """RAILSsolver.m:1-531.  A, M (or None), B dense arrays.  V0 replaces the
    random start vector of :277-280.  Returns a dict: V, T, res, iter, resvec,
    space (columns of V when each residual was estimated)."""
    opts = dict(opts or {})
    m = B.shape[0]
    if callable(A):                                            # :262-270 function handle
        A_apply = A
    else:
        A_mat = A
        A_apply = lambda x: A_mat @ x
    has_m = M is not None
    restart_size = opts.get("restart_size", -1)                # :118-131 defaults
    restart_iterations = opts.get("restart_iterations", -1)
    expand = opts.get("expand", min(3, B.shape[1]))
    mortho = opts.get("ortho", None) == "M"
    ortho = not opts.get("space_is_orthogonalized", False)
    fast = opts.get("fast_orthogonalization", True)
    restart_tolerance = opts.get("restart_tolerance", 1e-3 * tol)
    restart_upon_convergence = opts.get("restart_upon_convergence", True)
    restart_upon_start = opts.get("restart_upon_start", False)
    nev = opts.get("lanczos_vectors", None)
    if "reduced_size" in opts:                                 # :238-246
        reduced_size = opts["reduced_size"]
    elif restart_size > 0:
        reduced_size = restart_size / 2
    else:
        reduced_size = -1
    nullspace = None
    if opts.get("nullspace", None) is not None:                # :218-220
        nullspace = morth(np.asarray(opts["nullspace"], dtype=float), None)

    projection_method = opts.get("projection_method", 1)       # :131,224-226
    invA = opts.get("Ainv", opts.get("invA", None))            # :155-160
    if invA is None:
        invA = lambda x: np.linalg.solve(A_mat, x)             # :284-286
    frac = projection_method - np.floor(projection_method)
    eps = np.sqrt(np.finfo(float).eps)

    V = opts.get("space", V0)
    if V is None:
        V = (np.random.rand(m, 1) - 0.5) * 2                   # :277-280
    V = np.array(V, dtype=float).reshape(m, -1)
    W = V
    if abs(frac - 0.1) < eps:                                  # :294-296
        W = invA(V)
    if abs(frac - 0.2) < eps:                                  # :297-300
        V = np.array(B, dtype=float)
        W = invA(V)
    if np.floor(projection_method) == 2 and abs(frac - 0.3) > eps:     # :302-307
        V = np.hstack([V, W])
    elif np.floor(projection_method) == 1 and abs(frac - 0.3) > eps:
        V = W
    if mortho:                                                 # :303-307
        V = morth(V, M, nullspace, None, fast)
    elif ortho:
        V = morth(V, None, nullspace, None, fast)
    if nev is None:
        nev = expand                                           # :309-311

    AV = np.zeros((m, 0)); MV = np.zeros((m, 0)); BV = np.zeros((B.shape[1], 0))
    VAV = np.zeros((0, 0)); VMV = np.zeros((0, 0)); VBV = np.zeros((0, 0))
    T = np.zeros((0, 0))
    resvec, space = [], []
    it = 0
    iter_since_restart = 0
    converged = False
    reduced = False
    res = np.inf
    r0 = np.linalg.norm(B.T @ B, 2)                            # :333

    for i in range(1, maxit + 1):                              # :336
        it = i
        iter_since_restart += 1
        k_old = VAV.shape[1]
        new = slice(k_old, V.shape[1])
        AVnew = A_apply(V[:, new])                             # :346-353
        VAV = V.T @ AVnew if k_old == 0 else np.block([[VAV], [V[:, new].T @ AV]])
        if k_old > 0:
            VAV = np.hstack([VAV, V.T @ AVnew])
        AV = np.hstack([AV, AVnew])
        BVnew = B.T @ V[:, new]                                # :356-364
        if k_old == 0:
            VBV = BVnew.T @ BVnew
        else:
            t = BVnew.T @ BV
            VBV = np.block([[VBV, t.T], [t, BVnew.T @ BVnew]])
        BV = np.hstack([BV, BVnew])
        if has_m:                                              # :368-373
            MVnew = M @ V[:, new]
            MV_old = MV
            MV = np.hstack([MV, MVnew])
        if has_m and not mortho:                               # :375-385
            if k_old == 0:
                VMV = V.T @ MVnew
            else:
                VMV = np.hstack([np.vstack([VMV, V[:, new].T @ MV_old]), V.T @ MVnew])
            T = lyap(VAV, VBV, VMV)
        else:
            T = lyap(VAV, VBV)

        # eigs of the residual operator (:388-401), largest magnitude first
        V2, d2 = residual_eigs(AV, MV if has_m else V, T, B, min(nev, 2 * V.shape[1] + B.shape[1]))
        # candidates orthogonalised against the space BEFORE selection (:405-409)
        V2 = morth(V2, M if mortho else None, V, None, fast)

        res = np.max(np.abs(d2)) / r0                          # :411-416
        if iter_since_restart > 1 or i == 1:                   # :418-421
            resvec.append(res)
            space.append(V.shape[1])

        if abs(res) < tol:                                     # :423-436
            if converged or not restart_upon_convergence:
                break
            converged = True
        if i >= maxit or V.shape[1] >= m:                      # :438-454
            break

        if ((i == 1 and restart_upon_start)
                or (restart_iterations > 0 and iter_since_restart == restart_iterations)
                or (restart_size > 0 and T.shape[0] > restart_size)
                or (converged and not reduced)):               # :456-460
            w, Y = np.linalg.eigh(0.5 * (T + T.T))
            if 0 < reduced_size < V.shape[1]:                  # :462-474 eigs(T, reduced_size)
                order = np.argsort(-np.abs(w), kind="stable")[:int(reduced_size)]
                w, Y = w[order], Y[:, order]
            d = np.abs(w)
            keep = np.where(d / np.max(d) > restart_tolerance)[0]   # :476-477 RELATIVE cut
            keep = keep[np.argsort(-d[keep], kind="stable")]        # :480-481
            V3 = Y[:, keep]
            V = V @ V3
            reduced = converged
            VAV = V3.T @ VAV @ V3                              # :498-509
            AV = AV @ V3
            if has_m:
                MV = MV @ V3
            if VMV.size:
                VMV = V3.T @ VMV @ V3
            VBV = V3.T @ VBV @ V3
            VBV = 0.5 * (VBV + VBV.T)
            BV = BV @ V3
            iter_since_restart = 0
            continue

        num = min(expand, V2.shape[1], m - V.shape[1])         # :517
        W = V2[:, :num]
        if 1 < projection_method < 2:                          # :520-524
            W = invA(W)
        elif 2 < projection_method < 3:
            W = np.hstack([W, invA(W)])
        V = morth(W, M if mortho else None, nullspace, V, fast)   # :527-531

    return dict(V=V, T=T, res=res, iter=it, resvec=np.array(resvec), space=np.array(space))
