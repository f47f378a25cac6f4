"""numpy model of the two-stage tridiagonalisation of csrc/eigh2.cuh.  TEST INFRASTRUCTURE ONLY.

It restates, step for step and in the same data layouts, what the CUDA kernels do, so that the index logic (band
storage, bulge-chasing step order, reflector storage, order of application in the back-transformation) can be checked
on the CPU against numpy's eigh before any GPU time is spent:

  stage 1  sy2sb   dense symmetric -> band of half-width B: per panel of B rows a Householder QR of the part right of
                   the band (panel kernel: one fused dot-product pass + ONE reduction per column), compact-WY factor
                   T, W = Y - 1/2 V S, and the rank-2B update  A22 -= V W^T + W V^T  (two tensor-core GEMMs).
  stage 2  sb2st   band -> tridiagonal by bulge chasing, one sweep per column, each sweep a chain of steps on
                   B x B blocks (right-apply the previous reflector to the sub-diagonal block, annihilate its first
                   column, two-sided update of the next diagonal block).
  back     Q2, Q1  eigenvectors of T -> eigenvectors of A: the stage-2 reflectors in groups of G sweeps over row
                   tiles (later sweeps first, tiles top to bottom), then the stage-1 reflectors as compact-WY blocks.

Replaces (together with the divide & conquer of eigh.cuh) torch.linalg.eigh at kfac/layers/eigen.py:309-311,330-332.
"""
from __future__ import annotations

import numpy as np

B = 32   # band half-width = panel width = bulge-chasing block size
G = 32   # sweeps per back-transformation group


def house(alpha, s_full, dtype):
    """Householder from the pivot alpha and s_full = alpha^2 + |tail|^2: returns (beta, tau, scal) with
    v = [1, tail * scal].  Matches the device code: no reflector (tau = 0) only when the whole column is zero."""
    if s_full <= 0:
        return dtype(alpha), dtype(0), dtype(0)
    nb = np.sqrt(dtype(s_full))
    beta = dtype(-nb if alpha >= 0 else nb)
    tau = dtype((beta - alpha) / beta)
    scal = dtype(1) / dtype(alpha - beta)
    return beta, tau, scal


def sy2sb(a: np.ndarray, b: int = B):
    """Returns (W, tau, AB): W rows k < nref hold v_k (explicit 1 at column k + b, zeros before), tau per reflector,
    AB[j, d] = band[j + d, j] for d = 0..2b-1 (lower band storage with room for the bulge)."""
    dtype = a.dtype.type
    n = a.shape[0]
    w = a.copy()
    nref = max(0, n - b - 1)
    tau_all = np.zeros(n, dtype=a.dtype)
    k0 = 0
    while n - k0 - b > 1:
        m = n - k0 - b
        nb = min(b, m - 1)                      # reflectors of this panel
        tile = w[k0:k0 + b, k0 + b:n].copy()    # tile[l, c]  (l = panel column = row of W, c = position along m)
        tmat = np.zeros((b, b), dtype=a.dtype)
        for j in range(nb):
            # one fused pass: s[l] = sum_{c >= j} tile[j, c] * tile[l, c] for ALL l (l < j: rows hold earlier v's)
            s = tile[:, j:] @ tile[j, j:]
            piv = tile[:, j].copy()             # pivot row (c = j) of every panel column
            alpha = piv[j]
            beta, tau, scal = house(alpha, s[j], dtype)
            # w_l = v^T y_l = (s_l - beta * y_l[j]) * scal  for l > j   (v = (x - beta e1) * scal)
            wl = (s - beta * piv) * scal
            # T column: z_l = V[:, l]^T v_j = piv_l + (s_l - piv_l * alpha) * scal  for l < j
            z = piv[:j] + (s[:j] - piv[:j] * alpha) * scal
            if tau != 0:
                tmat[:j, j] = -tau * (tmat[:j, :j] @ z)
            tmat[j, j] = tau
            # v tail in place, R in place
            v = tile[j, j:].copy() * scal
            v[0] = 1
            for l in range(j + 1, b):
                tile[l, j:] -= tau * wl[l] * v
            tile[j, j + 1:] = v[1:]
            tile[j, j] = beta
            tau_all[k0 + j] = tau
        # explicit V (m x b), zeros above the unit diagonal, for the GEMMs
        vt = np.zeros((b, m), dtype=a.dtype)
        for j in range(nb):
            vt[j, j] = 1
            vt[j, j + 1:] = tile[j, j + 1:]
        a22 = w[k0 + b:, k0 + b:]
        y0 = a22 @ vt.T                          # GEMM 1 (m x b)
        y = y0 @ tmat                            # W kernel
        gm = vt @ y
        smat = tmat.T @ gm
        wm = y - dtype(0.5) * (vt.T @ smat)
        a22 -= vt.T @ wm.T + wm @ vt             # GEMM 2 (rank-2b, symmetric)
        # write the panel back LAPACK style: R (incl. diagonal) in place, v tails beyond
        w[k0:k0 + b, k0 + b:n] = tile
        k0 += b
    # band extraction + V finalisation (one elementwise kernel)
    ab = np.zeros((n, 2 * b), dtype=a.dtype)
    for j in range(n):
        for d in range(0, b + 1):
            if j + d < n:
                ab[j, d] = w[j, j + d]
    for k in range(n):
        if k < nref:
            w[k, :k + b] = 0
            w[k, k + b] = 1
        else:
            w[k, :] = 0
    return w, tau_all, ab


def tiles_of_group(n: int, g: int, b: int = B, gs: int = G) -> int:
    """Number of row tiles that hold reflectors of sweep group g (first sweep g * gs)."""
    s0 = g * gs
    if s0 > n - 3:
        return 0
    return (n - 2 - s0) // b + 1


def sb2st(ab: np.ndarray, b: int = B, gs: int = G):
    """Bulge chasing on the lower band storage.  Returns (d, e, V2, tau2, goff) with the reflector of (sweep s,
    step t) at V2[goff[s // gs] + t, s % gs, :] (zero padded to b) and tau2 likewise."""
    dtype = ab.dtype.type
    n = ab.shape[0]
    ab = ab.copy()
    ngroups = (max(n - 2, 0) + gs - 1) // gs
    goff = np.zeros(ngroups + 1, dtype=np.int64)
    for g in range(ngroups):
        goff[g + 1] = goff[g] + tiles_of_group(n, g, b, gs)
    v2 = np.zeros((goff[-1], gs, b), dtype=ab.dtype)
    tau2 = np.zeros((goff[-1], gs), dtype=ab.dtype)

    def get(i, j):            # element (i, j), i >= j
        return ab[j, i - j]

    def put(i, j, val):
        ab[j, i - j] = val

    for s in range(n - 2):
        g, sl = s // gs, s % gs
        vprev = None
        tauprev = dtype(0)
        t = 0
        while True:
            r0 = s + 1 + t * b
            if r0 > n - 1:
                break
            ln = min(b, n - r0)                 # rows of this step
            if t == 0:
                x = np.array([get(r0 + r, s) for r in range(ln)], dtype=ab.dtype)
            else:
                c0 = r0 - b                     # first column of the previous block (lp == b always here)
                lp = b
                e = np.zeros((ln, lp), dtype=ab.dtype)
                for r in range(ln):
                    for c in range(lp):
                        e[r, c] = get(r0 + r, c0 + c)
                # right-apply H_{t-1}
                yv = e @ vprev
                e -= tauprev * np.outer(yv, vprev)
                x = e[:, 0].copy()
            alpha = x[0]
            s_full = dtype(x @ x)
            if ln > 1:
                beta, tau, scal = house(alpha, s_full, dtype)
                if not np.any(x[1:] != 0):
                    beta, tau, scal = dtype(alpha), dtype(0), dtype(0)   # nothing to annihilate: keep the sign
            else:
                beta, tau, scal = dtype(alpha), dtype(0), dtype(0)
            v = np.zeros(b, dtype=ab.dtype)
            v[0] = 1
            v[1:ln] = x[1:] * scal
            if t == 0:
                put(r0, s, beta)
                for r in range(1, ln):
                    put(r0 + r, s, 0)
            else:
                e[0, 0] = beta
                e[1:, 0] = 0
                # left-apply H_t to the remaining columns
                zc = v[:ln] @ e[:, 1:]
                e[:, 1:] -= tau * np.outer(v[:ln], zc)
                for r in range(ln):
                    for c in range(lp):
                        put(r0 + r, c0 + c, e[r, c])
            # two-sided update of the diagonal block
            dblk = np.zeros((ln, ln), dtype=ab.dtype)
            for r in range(ln):
                for c in range(r + 1):
                    dblk[r, c] = get(r0 + r, r0 + c)
                    dblk[c, r] = dblk[r, c]
            u = dblk @ v[:ln]
            vdv = v[:ln] @ u
            wv = tau * u - dtype(0.5) * tau * tau * vdv * v[:ln]
            dblk -= np.outer(v[:ln], wv) + np.outer(wv, v[:ln])
            for r in range(ln):
                for c in range(r + 1):
                    put(r0 + r, r0 + c, dblk[r, c])
            v2[goff[g] + t, sl, :] = v
            tau2[goff[g] + t, sl] = tau
            vprev, tauprev = v, tau
            if ln < b:
                break
            t += 1
    d = ab[:, 0].copy()
    e = np.zeros(n, dtype=ab.dtype)
    e[:n - 1] = ab[:n - 1, 1]
    return d, e, v2, tau2, goff


def q2_apply(z: np.ndarray, v2, tau2, goff, b: int = B, gs: int = G) -> np.ndarray:
    """Z <- Q2 Z with Q2 = H_{0,*} H_{1,*} ... : groups from the last to the first, inside a group row tiles top to
    bottom, inside a tile the sweeps of the group from the last to the first."""
    n = z.shape[0]
    z = z.copy()
    ngroups = len(goff) - 1
    for g in range(ngroups - 1, -1, -1):
        nt = goff[g + 1] - goff[g]
        for t in range(nt):
            base = g * gs + 1 + t * b           # first row of the tile (window of b + gs - 1 rows)
            for sl in range(gs - 1, -1, -1):
                tau = tau2[goff[g] + t, sl]
                if tau == 0:
                    continue
                r0 = base + sl
                ln = min(b, n - r0)
                if ln <= 0:
                    continue
                v = v2[goff[g] + t, sl, :ln]
                zz = z[r0:r0 + ln, :]
                zz -= tau * np.outer(v, v @ zz)
    return z


def q1_apply(z: np.ndarray, w: np.ndarray, tau: np.ndarray) -> np.ndarray:
    """Z <- H_0 H_1 ... H_{nref-1} Z with v_k = row k of W (what stage 3 of eigh.cuh does in compact-WY blocks)."""
    z = z.copy()
    for k in range(len(tau) - 1, -1, -1):
        if tau[k] != 0:
            v = w[k]
            z -= tau[k] * np.outer(v, v @ z)
    return z


def eigh_two_stage(a: np.ndarray, b: int = B, gs: int = G):
    w, tau1, ab = sy2sb(a, b)
    d, e, v2, tau2, goff = sb2st(ab, b, gs)
    n = a.shape[0]
    t = np.diag(d.astype(np.float64)) + np.diag(e[:n - 1].astype(np.float64), 1) + np.diag(e[:n - 1].astype(np.float64), -1)
    lam, zt = np.linalg.eigh(t)
    z = q2_apply(zt.astype(a.dtype), v2, tau2, goff, b, gs)
    z = q1_apply(z, w, tau1)
    return lam, z


if __name__ == '__main__':
    import sys
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    dt = np.float32 if (len(sys.argv) > 2 and sys.argv[2] == 'f32') else np.float64
    rng = np.random.default_rng(0)
    x = np.maximum(rng.standard_normal((4 * n, n)), 0)
    a = (0.95 * np.eye(n) + 0.05 * x.T @ x / (4 * n)).astype(dt)
    lam, q = eigh_two_stage(a)
    ad, qd = a.astype(np.float64), q.astype(np.float64)
    rec = np.linalg.norm(qd @ np.diag(lam) @ qd.T - ad) / np.linalg.norm(ad)
    orth = np.abs(qd.T @ qd - np.eye(n)).max()
    ev = np.abs(lam - np.linalg.eigvalsh(ad)).max() / np.abs(lam).max()
    print(f'n={n} dtype={dt.__name__} reconstruction {rec:.2e} orthogonality {orth:.2e} eigenvalues {ev:.2e}')
