"""Batched eigensolver on the matrix sizes that dominate the benchmark (n >= 2560, 4608, mixed rounds, 8193).

The small-size tests in test_kernels_gpu.py stop at n = 1537; the kernels that run the big ResNet-50 factors are
chosen by size inside kfac_eigh_batched, so these tests pin both the numbers and -- through kfac_eigh_last_paths --
WHICH tridiagonalisation path ran, so that a change of the dispatch thresholds cannot silently drop coverage
(reference call site: kfac/layers/eigen.py:309-311,330-332; results compared after reconstruction because
eigenvectors are not unique).
"""
from __future__ import annotations

import pytest
import torch

pytestmark = pytest.mark.gpu

REC_TOL = 1e-5     # ||Q diag(d) Q^T - F|| / ||F||   (north star: factors 1e-5)
ORTH_TOL = 1e-5    # max |Q^T Q - I|
EVAL_TOL = 2e-6    # max |d - eigvalsh_fp64(F)| / max |d|  (32 fp32 ulps of the largest eigenvalue)


@pytest.fixture(scope='module')
def ops():
    from kfac_b200 import ops as _ops
    return _ops


def spd_factor(n: int, seed: int, decay: float = 0.95) -> torch.Tensor:
    """SURVEY 8d config-5 matrix: F = decay I + (1 - decay) X^T X / M, X = relu(N(0,1)), M = 4n."""
    g = torch.Generator(device='cuda').manual_seed(seed)
    x = torch.relu(torch.randn(4 * n, n, device='cuda', generator=g))
    return (decay * torch.eye(n, device='cuda') + (1 - decay) * (x.t() @ x) / (4 * n)).contiguous()


def check_eig(f: torch.Tensor, q: torch.Tensor, d: torch.Tensor) -> dict:
    n = f.shape[0]
    fd, qd, dd = f.double(), q.double(), d.double()
    assert torch.isfinite(q).all() and torch.isfinite(d).all()
    assert (d[1:] >= d[:-1]).all(), 'eigenvalues not ascending'
    rec = (torch.linalg.norm(qd @ torch.diag(dd) @ qd.t() - fd) / torch.linalg.norm(fd)).item()
    orth = (qd.t() @ qd - torch.eye(n, device='cuda', dtype=torch.double)).abs().max().item()
    ref = torch.linalg.eigvalsh(fd)
    ev = ((dd - ref).abs().max() / ref.abs().max()).item()
    assert rec < REC_TOL, ('reconstruction', n, rec)
    assert orth < ORTH_TOL, ('orthogonality', n, orth)
    assert ev < EVAL_TOL, ('eigenvalues', n, ev)
    return {'rec': rec, 'orth': orth, 'ev': ev}


def large_path_count(ops) -> int:
    """Matrices of the last call that went through one of the large-matrix paths (anything but the full-matrix
    pass used below n = 1024)."""
    p = ops.eigh_last_paths()
    return p['sym_ldg'] + p['sym_bulk'] + p['two_stage']


@pytest.mark.parametrize('n', [2560, 2561, 4608])
def test_single_large_matrix(ops, n):
    """One big factor alone (the N=8 critical path is a lone 4608): even and odd sizes."""
    f = spd_factor(n, seed=n)
    (q,), (d,) = ops.eigh_batched([f], clamp_nonneg=True)
    check_eig(f, q, d)
    assert large_path_count(ops) == 1, ops.eigh_last_paths()


def test_round_led_by_big_matrices_with_small_ones_riding_along(ops):
    """The shape of the ResNet-50 batch: a round led by 4608s, smaller factors in the same call."""
    ns = (4608, 4608, 2304, 1152, 64)
    fs = [spd_factor(n, seed=100 + i) for i, n in enumerate(ns)]
    qs, ds = ops.eigh_batched(fs, clamp_nonneg=True)
    for f, q, d in zip(fs, qs, ds):
        check_eig(f, q, d)
    paths = ops.eigh_last_paths()
    assert sum(paths.values()) == len(ns), paths
    assert large_path_count(ops) >= 2, paths
    # the inputs are not modified (F is the running average the next step keeps updating)
    assert torch.equal(fs[0], spd_factor(4608, seed=100))


def test_low_rank_plus_identity_factor(ops):
    """A conv A factor early in training: M < n samples, so F = decay^k I + (rank-M part): a huge cluster of equal
    eigenvalues (everything deflates there) next to a few large ones."""
    n, m = 2600, 300
    g = torch.Generator(device='cuda').manual_seed(7)
    x = torch.relu(torch.randn(m, n, device='cuda', generator=g))
    f = (0.95 * torch.eye(n, device='cuda') + 0.05 * (x.t() @ x) / m).contiguous()
    (q,), (d,) = ops.eigh_batched([f], clamp_nonneg=True)
    check_eig(f, q, d)


def test_largest_gpt_neox_factor(ops):
    """n = 8193 (GPT-NeoX 4h->h A factor): 0.8 GB of scratch."""
    if torch.cuda.mem_get_info()[0] < 6 << 30:
        pytest.skip('needs ~3 GB of free device memory')
    n = 8193
    f = spd_factor(n, seed=8193)
    (q,), (d,) = ops.eigh_batched([f], clamp_nonneg=True)
    check_eig(f, q, d)
    assert large_path_count(ops) == 1, ops.eigh_last_paths()
