"""oracle/twostage_model.py (the numpy model of the two-stage tridiagonalisation in csrc/eigh2.cuh) against
numpy.linalg.eigh: the model is what the GPU tests compare the kernels' intermediate results with, so it is pinned
here first, on the CPU, in fp64 (index logic) and fp32 (error growth of the arithmetic the kernels use).
Reference call site replaced: torch.linalg.eigh in kfac/layers/eigen.py:309-311,330-332."""
from __future__ import annotations

import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'oracle'))
import twostage_model as tm  # noqa: E402


def factor(n: int, dtype) -> np.ndarray:
    rng = np.random.default_rng(n)
    x = np.maximum(rng.standard_normal((4 * n, n)), 0)
    return (0.95 * np.eye(n) + 0.05 * x.T @ x / (4 * n)).astype(dtype)


def errors(a, lam, q):
    ad, qd = a.astype(np.float64), q.astype(np.float64)
    n = a.shape[0]
    rec = np.linalg.norm(qd @ np.diag(lam) @ qd.T - ad) / np.linalg.norm(ad)
    orth = np.abs(qd.T @ qd - np.eye(n)).max()
    ev = np.abs(lam - np.linalg.eigvalsh(ad)).max() / np.abs(lam).max()
    return rec, orth, ev


@pytest.mark.parametrize('n', [3, 34, 35, 66, 100, 161])
def test_model_is_exact_in_fp64(n):
    a = factor(n, np.float64)
    rec, orth, ev = errors(a, *tm.eigh_two_stage(a))
    assert rec < 1e-13 and orth < 1e-13 and ev < 1e-13, (rec, orth, ev)


def test_model_band_and_reflector_layout():
    n = 100
    a = factor(n, np.float64)
    w, tau1, ab = tm.sy2sb(a)
    assert ab.shape == (n, 64) and not ab[:, 33:].any()          # half-width 32, bulge room zero
    nref = n - 33
    for k in range(nref):
        assert w[k, k + 32] == 1 and not w[k, :k + 32].any()      # v_k: unit entry at column k + 32
    assert not w[nref:].any() and not tau1[nref:].any()
    band = np.zeros((n, n))
    for d in range(33):
        idx = np.arange(n - d)
        band[idx + d, idx] = ab[:n - d, d]
        band[idx, idx + d] = ab[:n - d, d]
    assert np.abs(np.linalg.eigvalsh(band) - np.linalg.eigvalsh(a)).max() < 1e-13
    d, e, v2, tau2, goff = tm.sb2st(ab)
    assert goff[-1] == sum(tm.tiles_of_group(n, g) for g in range(len(goff) - 1)) and v2.shape[1:] == (32, 32)
    t = np.diag(d) + np.diag(e[:n - 1], 1) + np.diag(e[:n - 1], -1)
    assert np.abs(np.linalg.eigvalsh(t) - np.linalg.eigvalsh(a)).max() < 1e-13


def test_model_in_fp32_stays_inside_the_gpu_tolerances():
    a = factor(300, np.float32)
    rec, orth, ev = errors(a, *tm.eigh_two_stage(a))
    assert rec < 1e-5 and orth < 1e-5 and ev < 2e-6, (rec, orth, ev)
