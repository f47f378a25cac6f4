"""numpy model of the batched GPU eigensolver (csrc/eigh.cuh).  TEST INFRASTRUCTURE ONLY.

Same algorithm, same data flow, written for clarity so the CUDA kernels can be checked stage by stage:
  1. Householder tridiagonalisation  A = H T H^T               (fp32 on the GPU)
  2. Cuppen divide & conquer on T with 1x1 leaves; every merge solves the secular equation in fp64,
     recomputes z by the Gu-Eisenstat / Loewner formula (orthogonal eigenvectors without extended
     precision), folds deflation rotations into the merge matrix U and updates Q <- Q U with a GEMM
  3. back-transformation  Q <- H Q  by blocked compact-WY reflectors (GEMMs)
Replaces torch.linalg.eigh at kfac/layers/eigen.py:309-311,330-332.
"""
from __future__ import annotations

import numpy as np


def tridiagonalize(a: np.ndarray):
    """Unblocked Householder reduction (lower variant). Returns d, e, V (rows k hold v_k), tau."""
    a = a.copy()
    n = a.shape[0]
    d = np.zeros(n, a.dtype)
    e = np.zeros(max(n - 1, 0), a.dtype)
    tau = np.zeros(max(n - 1, 0), a.dtype)
    vs = np.zeros((n, n), a.dtype)
    for k in range(n - 1):
        x = a[k + 1:, k].copy()
        alpha = x[0]
        xnorm = np.sqrt(np.sum(x[1:].astype(np.float64) ** 2)).astype(a.dtype)
        if xnorm == 0:
            tau[k] = 0
            e[k] = alpha
            v = np.zeros_like(x)
            v[0] = 1
        else:
            beta = -np.copysign(np.sqrt(alpha.astype(np.float64) ** 2 + xnorm.astype(np.float64) ** 2), alpha).astype(a.dtype)
            tau[k] = (beta - alpha) / beta
            v = x / (alpha - beta)
            v[0] = 1
            e[k] = beta
        vs[k, k + 1:] = v
        d[k] = a[k, k]
        if tau[k] != 0:
            a22 = a[k + 1:, k + 1:]
            p = tau[k] * (a22 @ v)
            w = p - (tau[k] / 2) * (p @ v) * v
            a22 -= np.outer(v, w) + np.outer(w, v)
    d[n - 1] = a[n - 1, n - 1]
    return d, e, vs, tau


def back_transform(vs: np.ndarray, tau: np.ndarray, z: np.ndarray) -> np.ndarray:
    """Q = H_0 H_1 ... H_{n-2} Z."""
    q = z.copy()
    n = z.shape[0]
    for k in range(n - 2, -1, -1):
        if tau[k] != 0:
            v = vs[k, k + 1:]
            q[k + 1:, :] -= tau[k] * np.outer(v, v @ q[k + 1:, :])
    return q


def build_tree(n: int):
    """Levels of merge nodes (lo, mid, hi), deepest level first."""
    levels: dict[int, list[tuple[int, int, int]]] = {}

    def rec(lo, hi, depth):
        if hi - lo <= 1:
            return
        mid = lo + (hi - lo) // 2
        levels.setdefault(depth, []).append((lo, mid, hi))
        rec(lo, mid, depth + 1)
        rec(mid, hi, depth + 1)

    rec(0, n, 0)
    return [levels[k] for k in sorted(levels, reverse=True)]


def secular_root(dk: np.ndarray, z2: np.ndarray, rho: float, j: int):
    """Root j of 1 + rho * sum z_i^2 / (d_i - lam) = 0 as (origin, mu): lam = d[origin] + mu."""
    k = dk.shape[0]
    if j < k - 1:
        gap = dk[j + 1] - dk[j]
        half = 0.5 * gap
        delta = dk - dk[j]
        fmid = 1.0 + rho * np.sum(z2 / (delta - half))
        if fmid > 0:
            origin, sgn = j, 1.0
            tlo = rho * z2[j] / (1.0 + 2.0 * rho / gap)
        else:
            origin, sgn = j + 1, -1.0
            delta = dk - dk[j + 1]
            tlo = z2[j + 1] * half
        thi = half
    else:
        origin, sgn = k - 1, 1.0
        delta = dk - dk[k - 1]
        thi = rho * np.sum(z2)
        tlo = rho * z2[k - 1]

    def g(t):
        den = delta - sgn * t
        return 1.0 + rho * np.sum(z2 / den), rho * np.sum(z2 / den ** 2)

    # g(mu) is increasing in mu; in t = |mu| it is increasing for sgn=+1 and decreasing for sgn=-1
    tlo = min(tlo, thi)
    for _ in range(64):
        val, _d = g(tlo)
        if sgn * val <= 0 or tlo == 0:
            break
        tlo *= 1.0 / 16
    t = 0.5 * (tlo + thi) if tlo > 0 else 0.5 * thi
    t = np.sqrt(tlo * thi) if tlo > 0 and thi / tlo > 16 else t
    for _ in range(200):
        val, der = g(t)
        if sgn * val > 0:
            thi = t
        else:
            tlo = t
        if thi - tlo <= 4e-16 * thi:
            break
        tn = t - sgn * val / der if der > 0 else -1.0
        if not (tlo < tn < thi):
            tn = np.sqrt(tlo * thi) if tlo > 0 and thi / tlo > 16 else 0.5 * (tlo + thi)
        if tn == t:
            break
        t = tn
    return origin, sgn * t


def merge(dcur: np.ndarray, q: np.ndarray, lo: int, mid: int, hi: int, beta: float, tol_scale: float = 7.5e-9):
    """One Cuppen merge on the node [lo, hi): returns the (m x m) matrix U and the new sorted eigenvalues."""
    m = hi - lo
    rho = abs(beta)
    sgn = 1.0 if beta >= 0 else -1.0
    z = np.concatenate([q[mid - 1, lo:mid], sgn * q[mid, mid:hi]]).astype(np.float64)
    d = dcur[lo:hi].astype(np.float64).copy()
    nrm = np.linalg.norm(z)
    if nrm == 0 or rho == 0:
        order = np.argsort(d, kind='stable')
        u = np.zeros((m, m))
        u[order, np.arange(m)] = 1.0
        return u, d[order]
    z /= nrm
    rho = rho * nrm * nrm
    order = np.argsort(d, kind='stable')
    ds, zs = d[order], z[order]
    tol = tol_scale * max(np.abs(ds).max(), rho)
    # deflation scan over sorted entries
    keep: list[int] = []          # positions (in sorted order) that stay in the secular problem
    deflated: list[int] = []
    rotations: list[tuple[int, int, float, float]] = []   # (sorted pos i, sorted pos j, c, s)
    prev = -1
    for i in range(m):
        if rho * abs(zs[i]) <= tol:
            deflated.append(i)
            continue
        if prev >= 0:
            # try to rotate (prev, i) so that z[prev] becomes 0
            s_, c_ = zs[prev], zs[i]
            tau_ = np.hypot(c_, s_)
            t_ = ds[i] - ds[prev]
            c_, s_ = c_ / tau_, -s_ / tau_
            if abs(t_ * c_ * s_) <= tol:
                zs[i] = tau_
                zs[prev] = 0.0
                rotations.append((prev, i, c_, s_))
                dp = ds[prev] * c_ * c_ + ds[i] * s_ * s_
                di = ds[prev] * s_ * s_ + ds[i] * c_ * c_
                ds[prev], ds[i] = dp, di
                deflated.append(prev)
                # keep deflated value ordering loosely; final sort handles it
                prev = i
                continue
            keep.append(prev)
        prev = i
    if prev >= 0:
        keep.append(prev)
    k = len(keep)
    u = np.zeros((m, m))          # rows: sorted positions (mapped back through `order` at the end)
    newvals = np.zeros(m)
    col = 0
    for i in deflated:
        u[i, col] = 1.0
        newvals[col] = ds[i]
        col += 1
    if k > 0:
        dk, zk = ds[keep], zs[keep]
        if k == 1:
            lam_origin, lam_mu = [0], [rho * zk[0] ** 2]
        else:
            z2 = zk * zk
            lam_origin, lam_mu = [], []
            for j in range(k):
                o, mu = secular_root(dk, z2, rho, j)
                lam_origin.append(o)
                lam_mu.append(mu)
        lam_origin = np.array(lam_origin)
        lam_mu = np.array(lam_mu)
        # Loewner: zhat_i^2 = prod_j (lam_j - d_i) / (rho * prod_{j != i} (d_j - d_i))
        zhat = np.zeros(k)
        for i in range(k):
            diff_l = (dk[lam_origin] - dk[i]) + lam_mu          # lam_j - d_i
            prod = diff_l[i] / rho
            for j in range(k):
                if j != i:
                    prod *= diff_l[j] / (dk[j] - dk[i])
            zhat[i] = np.copysign(np.sqrt(abs(prod)), zk[i])
        for j in range(k):
            den = (dk - dk[lam_origin[j]]) - lam_mu[j]          # d_i - lam_j
            vec = zhat / den
            vec /= np.linalg.norm(vec)
            u[keep, col] = vec
            newvals[col] = dk[lam_origin[j]] + lam_mu[j]
            col += 1
    # undo the deflation rotations (Q' = Q G1 G2 ... => U <- G1 (G2 (... U)))
    for (i, j, c_, s_) in reversed(rotations):
        ri, rj = u[i].copy(), u[j].copy()
        # columns (i, j) of Q were replaced by (c*qi + s*qj, -s*qi + c*qj)
        u[i] = c_ * ri - s_ * rj
        u[j] = s_ * ri + c_ * rj
    # sort columns by new eigenvalue, map rows back to original positions
    corder = np.argsort(newvals, kind='stable')
    ufull = np.zeros((m, m))
    ufull[order, :] = u[:, corder]
    return ufull, newvals[corder]


def tridiag_dc(d: np.ndarray, e: np.ndarray, gemm_dtype=np.float32):
    """Eigen-decomposition of the symmetric tridiagonal (d, e): returns (w, Z), Z in gemm_dtype."""
    n = d.shape[0]
    ea = np.abs(e.astype(np.float64))
    dcur = d.astype(np.float64).copy()
    dcur[:-1] -= ea
    dcur[1:] -= ea
    q = np.eye(n, dtype=gemm_dtype)
    for level in build_tree(n):
        for (lo, mid, hi) in level:
            u, newd = merge(dcur, q, lo, mid, hi, float(e[mid - 1]))
            q[lo:hi, lo:hi] = q[lo:hi, lo:hi] @ u.astype(gemm_dtype)
            dcur[lo:hi] = newd
    return dcur, q


def eigh(a: np.ndarray, dtype=np.float32):
    """Full model: A (symmetric) -> (w ascending, Q with eigenvectors in columns)."""
    a = np.asarray(a, dtype=dtype)
    d, e, vs, tau = tridiagonalize(a)
    w, z = tridiag_dc(d, e, gemm_dtype=dtype)
    q = back_transform(vs, tau, z)
    return w.astype(dtype), q
