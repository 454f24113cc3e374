import numpy as np
def eig_tql(A):
    A = A.astype(np.float32).copy(); n = A.shape[0]
    f32 = np.float32
    d = np.zeros(n, f32); e = np.zeros(n, f32); hh = np.zeros(n, f32)
    for i in range(n-1, 1, -1):
        x = A[i, :i].copy()
        scale = f32(np.abs(x).sum())
        if scale == 0:
            e[i] = x[i-1]; hh[i] = 0; A[:i, i] = 0; d[i] = A[i, i]; continue
        u = (x / scale).astype(f32)
        h = f32((u*u).sum()); f = u[i-1]
        g = f32(-np.copysign(np.sqrt(h), f))
        e[i] = scale * g; h = f32(h - f*g); u[i-1] = f - g
        A[:i, i] = u
        p = (A[:i, :i] @ u / h).astype(f32)
        K = f32((u*p).sum() / (2*h))
        q = (p - K*u).astype(f32)
        A[:i, :i] -= np.outer(u, q) + np.outer(q, u)
        hh[i] = h; d[i] = A[i, i]
    d[0] = A[0,0]; d[1] = A[1,1]; e[1] = A[1,0]; e[0] = 0
    A[0,0] = 1
    for i in range(2, n+1):
        A[i-1, :i-1] = 0; A[:i-1, i-1] = 0; A[i-1, i-1] = 1   # column i-1 held u^(i-1): consumed
        if i < n and hh[i] != 0:
            u = A[:i, i].copy()
            g = u @ A[:i, :i]
            A[:i, :i] -= np.outer(u, g / hh[i])
    Z = A
    # tql2
    e[:-1] = e[1:]; e[n-1] = 0
    f = f32(0); tst1 = f32(0); eps = f32(2.0**-23)
    for l in range(n):
        tst1 = max(tst1, abs(d[l]) + abs(e[l]))
        m = l
        while m < n and abs(e[m]) > eps*tst1: m += 1
        if m > l:
            it = 0
            while True:
                it += 1
                g = d[l]; p = (d[l+1]-g)/(f32(2)*e[l]); r = f32(np.hypot(p, f32(1)))
                if p < 0: r = -r
                d[l] = e[l]/(p+r); d[l+1] = e[l]*(p+r); dl1 = d[l+1]; h = g - d[l]
                d[l+2:] -= h; f += h
                p = d[m]; c = f32(1); c2 = c; c3 = c; el1 = e[l+1]; s = f32(0); s2 = f32(0)
                for i in range(m-1, l-1, -1):
                    c3 = c2; c2 = c; s2 = s
                    g = c*e[i]; h = c*p; r = f32(np.hypot(p, e[i]))
                    e[i+1] = s*r; s = e[i]/r; c = p/r
                    p = c*d[i] - s*g
                    d[i+1] = h + s*(c*g + s*d[i])
                    hcol = Z[:, i+1].copy()
                    Z[:, i+1] = s*Z[:, i] + c*hcol
                    Z[:, i] = c*Z[:, i] - s*hcol
                p = -s*s2*c3*el1*e[l]/dl1
                e[l] = s*p; d[l] = c*p
                if not (abs(e[l]) > eps*tst1) or it > 60: break
        d[l] = d[l] + f; e[l] = 0
    return d, Z
rng = np.random.default_rng(0)
for trial in range(5):
    X = rng.standard_normal((512, 12)) @ rng.standard_normal((12, 48)) * 3 + 0.3*rng.standard_normal((512,48))
    G = (X.T @ X).astype(np.float32)
    # bf16-round like the reference
    Gi = G.view(np.uint32); Gi = ((Gi + 0x7fff + ((Gi >> 16) & 1)) & 0xffff0000).astype(np.uint32); G = Gi.view(np.float32).copy()
    G = np.triu(G) + np.triu(G,1).T + np.float32(5e-3)*np.eye(48, dtype=np.float32)
    d, Z = eig_tql(G)
    w, V = np.linalg.eigh(G.astype(np.float64))
    print("eig err", np.abs(np.sort(d) - w).max() / np.abs(w).max(), "orth", np.abs(Z.T@Z - np.eye(48)).max(),
          "resid", np.abs(G@Z - Z*d).max()/np.abs(w).max(), "min|w|", np.abs(w).min())
    Minv = (Z / np.abs(d)) @ Z.T
    Mref = (V / np.abs(w)) @ V.T
    xs = X[:64].astype(np.float64)
    s1 = np.einsum('ti,ij,tj->t', xs, Minv.astype(np.float64), xs); s2 = np.einsum('ti,ij,tj->t', xs, Mref, xs)
    print("  score rel err", np.abs(s1-s2).max()/np.abs(s2).max())
