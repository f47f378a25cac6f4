"""Batched, asynchronous Newton-Schulz inverse (kfac_inverse_newton_schulz_batched) against torch.linalg.inv, the call it
replaces (kfac/layers/inverse.py:185-212), and the grouped inverse-method precondition (inverse.py:214-233).
Tolerance: 1e-4 relative Frobenius on the preconditioned gradient and on the inverse itself for K-FAC-conditioned
factors (north star); the residual ||X (F + damping I) - I||_F / sqrt(n) is reported as SURVEY.md section 8d asks."""
from __future__ import annotations

import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 1e-4


@pytest.fixture(scope='module')
def ops():
    from kfac_b200 import ops as _ops
    return _ops


def relerr(a, b):
    a, b = a.double(), b.double()
    return (torch.linalg.norm(a - b) / torch.linalg.norm(b)).item()


def factor(n, seed, m=None):
    g = torch.Generator(device='cuda').manual_seed(seed)
    x = torch.relu(torch.randn(m or 2 * n, n, device='cuda', generator=g))
    return (0.95 * torch.eye(n, device='cuda') + 0.05 * (x.t() @ x) / x.shape[0]).contiguous()


def test_batched_matches_linalg_inv(ops):
    ns = (1, 7, 64, 147, 513, 1000, 2049)
    fs = [factor(n, seed=n) for n in ns]
    damping = 0.001
    invs, info = ops.inverse_newton_schulz_batched(fs, damping)
    host = info.cpu()
    for f, x, row in zip(fs, invs, host):
        n = f.shape[0]
        m = f.double() + damping * torch.eye(n, device='cuda', dtype=torch.double)
        ref = torch.linalg.inv(m)
        assert int(row[1]) == 0 and 1 <= int(row[0]) <= 32, row
        assert relerr(x, ref) < TOL, (n, relerr(x, ref))
        resid = (torch.linalg.norm(x.double() @ m - torch.eye(n, device='cuda', dtype=torch.double)) / n ** 0.5).item()
        assert resid < TOL, (n, resid)
        assert relerr(x, x.t()) < 1e-5   # the inverse of a symmetric matrix


def test_batched_leaves_inputs_alone_and_reports_failures(ops):
    good = factor(96, seed=1)
    keep = good.clone()
    bad = factor(64, seed=2)
    bad[3, 5] = float('nan')
    slow = factor(300, seed=3)
    slow *= 1e4            # eigenvalues 1e4 .. against damping 1e-3: far beyond 3 iterations
    invs, info = ops.inverse_newton_schulz_batched([good, bad, slow], 0.001, max_iters=3)
    host = info.cpu()
    assert torch.equal(good, keep)
    assert int(host[1, 1]) == 2                      # non-finite input is reported, not masked
    assert int(host[2, 1]) == 1 and int(host[2, 0]) == 3   # stopped at max_iters without convergence
    # the single-matrix entry point reports both through its return code (torch.linalg.inv raises there)
    with pytest.raises(RuntimeError, match='non-finite'):
        ops.inverse_newton_schulz(bad, 0.001)
    with pytest.raises(RuntimeError, match='no convergence'):
        ops.inverse_newton_schulz(slow, 0.001, max_iters=3)
    x, iters = ops.inverse_newton_schulz(good, 0.001)
    assert 1 <= iters <= 32
    ref = torch.linalg.inv(good.double() + 0.001 * torch.eye(96, device='cuda', dtype=torch.double))
    assert relerr(x, ref) < TOL


def test_largest_resnet50_factor(ops):
    """n = 4608 (config 4 of BASELINE.json): parity with torch.linalg.inv <= 1e-4."""
    f = factor(4608, seed=4608, m=1568)
    (x,), info = ops.inverse_newton_schulz_batched([f], 0.001)
    assert int(info.cpu()[0, 1]) == 0
    m = f.double() + 0.001 * torch.eye(4608, device='cuda', dtype=torch.double)
    assert relerr(x, torch.linalg.inv(m)) < TOL


def test_grouped_inverse_precondition_matches_per_layer_and_fp64(ops):
    torch.manual_seed(9)
    items, refs = [], []
    vg = torch.zeros(1, dtype=torch.float64, device='cuda')
    expect_vg = 0.0
    for (g, a_w, bias) in [(64, 147, False), (128, 1152, False), (10, 64, True), (1000, 2048, True)]:
        a = a_w + int(bias)
        fa, fg = factor(a, seed=a), factor(g, seed=g + 1)
        (a_inv, g_inv), info = ops.inverse_newton_schulz_batched([fa, fg], 0.001)
        wg = torch.randn(g, a_w, device='cuda')
        bg = torch.randn(g, device='cuda') if bias else None
        out = torch.empty(g, a, device='cuda')
        items.append({'w_grad': wg, 'b_grad': bg, 'a_inv': a_inv, 'g_inv': g_inv, 'out': out})
        grad = torch.cat([wg, bg[:, None]], 1) if bias else wg
        eye_a = torch.eye(a, device='cuda', dtype=torch.double)
        eye_g = torch.eye(g, device='cuda', dtype=torch.double)
        ref = torch.linalg.inv(fg.double() + 0.001 * eye_g) @ grad.double() @ torch.linalg.inv(fa.double() + 0.001 * eye_a)
        refs.append(ref)
        expect_vg += (ref * grad.double()).sum().item()
    ops.precondition_inverse_grouped(items, vg_accum=vg)
    for it, ref in zip(items, refs):
        assert relerr(it['out'], ref) < TOL
        single = ops.precondition_inverse(it['w_grad'], it['b_grad'], it['a_inv'], it['g_inv'])
        assert relerr(it['out'], single) < 1e-6
    assert abs(vg.item() - expect_vg) <= 1e-4 * abs(expect_vg)
