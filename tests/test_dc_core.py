"""CPU unit tests of the divide-and-conquer scalar core (csrc/dc_core.h), compiled for the host with g++,
against numpy and the numpy model of the eigensolver (oracle/eigh_model.py)."""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np
import pytest

from oracle import eigh_model as em

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope='module')
def lib():
    out = os.path.join(HERE, '_build', 'libdc_core_test.so')
    os.makedirs(os.path.dirname(out), exist_ok=True)
    subprocess.check_call(['g++', '-O2', '-shared', '-fPIC', '-o', out, os.path.join(HERE, 'dc_core_harness.cpp')])
    return ctypes.CDLL(out)


def _dp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


def _ip(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int))


def _secular(lib, dk, z, rho):
    k = dk.shape[0]
    origin = np.zeros(k, np.int32)
    mu = np.zeros(k)
    z2 = z * z
    lib.dc_secular(_dp(dk), _dp(z2), ctypes.c_double(rho), k, _ip(origin), _dp(mu))
    return origin, mu


@pytest.mark.parametrize('k,rho,spread', [(2, 1.0, 1.0), (5, 0.3, 1.0), (40, 2.0, 1e-3), (200, 1e-4, 1.0), (64, 5.0, 1e6)])
def test_secular_roots_match_dense_eigenvalues(lib, k, rho, spread):
    rng = np.random.default_rng(k)
    dk = np.sort(rng.uniform(0, spread, k))
    z = rng.standard_normal(k)
    z /= np.linalg.norm(z)
    origin, mu = _secular(lib, dk, z, rho)
    lam = dk[origin] + mu
    ref = np.linalg.eigvalsh(np.diag(dk) + rho * np.outer(z, z))
    assert np.all(np.diff(lam) > 0)
    assert np.allclose(lam, ref, rtol=0, atol=1e-12 * max(spread, rho))
    # interlacing: d_j < lam_j < d_{j+1}
    assert np.all(lam > dk) and np.all(lam[:-1] < dk[1:])
    # the offset form keeps the differences accurate: residual of the secular equation is tiny
    for j in range(k):
        den = (dk - dk[origin[j]]) - mu[j]
        terms = rho * z * z / den
        assert abs(1.0 + terms.sum()) <= 1e-9 * (1.0 + np.abs(terms).sum())


def test_secular_tiny_weights(lib):
    """Poles with weights just above the deflation threshold: roots hug the poles."""
    dk = np.array([0.0, 1.0, 2.0, 3.0])
    z = np.array([1e-7, 1.0, 1e-6, 1e-7])
    z /= np.linalg.norm(z)
    origin, mu = _secular(lib, dk, z, 1.0)
    lam = dk[origin] + mu
    ref = np.linalg.eigvalsh(np.diag(dk) + np.outer(z, z))
    assert np.allclose(lam, ref, atol=1e-13)
    assert mu[0] > 0 and abs(mu[0]) < 1e-12


def _model_deflate(ds, zs, rho):
    """The scan of oracle/eigh_model.merge, isolated."""
    m = ds.shape[0]
    ds, zs = ds.copy(), zs.copy()
    tol = 7.5e-9 * max(np.abs(ds).max(), rho)
    keep, defl, rots = [], [], []
    prev = -1
    for i in range(m):
        if rho * abs(zs[i]) <= tol:
            defl.append(i)
            continue
        if prev >= 0:
            s_, c_ = zs[prev], zs[i]
            tau = np.hypot(c_, s_)
            t = ds[i] - ds[prev]
            c_, s_ = c_ / tau, -s_ / tau
            if abs(t * c_ * s_) <= tol:
                zs[i], zs[prev] = tau, 0.0
                rots.append((prev, i, c_, s_))
                dp = ds[prev] * c_ * c_ + ds[i] * s_ * s_
                di = ds[prev] * s_ * s_ + ds[i] * c_ * c_
                ds[prev], ds[i] = dp, di
                defl.append(prev)
                prev = i
                continue
            keep.append(prev)
        prev = i
    if prev >= 0:
        keep.append(prev)
    return keep, defl, rots, ds, zs


@pytest.mark.parametrize('m,kind', [(1, 'rand'), (7, 'rand'), (50, 'clustered'), (33, 'tiny_z'), (20, 'all_equal')])
def test_deflate_scan_matches_model(lib, m, kind):
    rng = np.random.default_rng(m)
    ds = np.sort(rng.uniform(0, 1, m))
    zs = rng.standard_normal(m)
    if kind == 'clustered':
        ds[10:30] = ds[10]
    if kind == 'tiny_z':
        zs[::3] *= 1e-12
    if kind == 'all_equal':
        ds[:] = 0.95
    zs /= np.linalg.norm(zs)
    keep_m, defl_m, rots_m, ds_m, zs_m = _model_deflate(ds, zs, 0.7)
    ds_c, zs_c = ds.copy(), zs.copy()
    keep = np.zeros(m, np.int32)
    defl = np.zeros(m, np.int32)
    ri = np.zeros(m, np.int32)
    rj = np.zeros(m, np.int32)
    rc = np.zeros(m)
    rs = np.zeros(m)
    counts = np.zeros(3, np.int32)
    lib.dc_deflate(m, ctypes.c_double(0.7), _dp(ds_c), _dp(zs_c), _ip(keep), _ip(defl), _ip(ri), _ip(rj), _dp(rc),
                   _dp(rs), _ip(counts))
    k, nd, nr = counts
    assert list(keep[:k]) == keep_m and list(defl[:nd]) == defl_m and nr == len(rots_m)
    assert k + nd == m
    for t, (i, j, c, s) in enumerate(rots_m):
        assert (ri[t], rj[t]) == (i, j) and np.isclose(rc[t], c) and np.isclose(rs[t], s)
    assert np.allclose(ds_c, ds_m) and np.allclose(zs_c, zs_m)
    if kind == 'all_equal':
        assert k == 1  # a multiple pole collapses to a single secular unknown


def test_model_eigh_against_numpy():
    """The numpy model the CUDA solver mirrors is itself a correct eigensolver."""
    rng = np.random.default_rng(3)
    for n in (1, 2, 3, 17, 64, 90):
        x = np.maximum(rng.standard_normal((3 * n, n)), 0)
        a = (0.95 * np.eye(n) + 0.05 * x.T @ x / (3 * n)).astype(np.float32)
        w, q = em.eigh(a, dtype=np.float64)
        assert np.all(np.diff(w) >= 0)
        assert np.abs(q.T @ q - np.eye(n)).max() < 1e-12
        assert np.linalg.norm(q @ np.diag(w) @ q.T - a) / np.linalg.norm(a) < 1e-7
