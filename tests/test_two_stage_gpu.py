"""Two-stage tridiagonalisation (csrc/eigh2.cuh) on the GPU, stage by stage and end to end.

* kfac_debug_two_stage exposes the intermediate results: the band after stage 1 and the tridiagonal matrix after
  stage 2 must have the eigenvalues of the input, and the storage layouts must be the ones the numpy model
  (oracle/twostage_model.py, pinned on the CPU by tests/test_twostage_model.py) defines.
* kfac_eigh_batched with the two-stage path forced for every n >= 96: reconstruction, orthogonality, eigenvalues,
  for sizes around every boundary of the algorithm (one panel, partial last panel, several sweep groups, window wrap).
Reference call site replaced: torch.linalg.eigh in kfac/layers/eigen.py:309-311,330-332; results are compared after
reconstruction because eigenvectors are not unique.  Tolerances: 1e-5 reconstruction / orthogonality (north star),
eigenvalues 2e-6 of the largest (32 fp32 ulps)."""
from __future__ import annotations

import pytest
import torch

pytestmark = pytest.mark.gpu

REC_TOL = 1e-5
ORTH_TOL = 1e-5
EVAL_TOL = 2e-6


@pytest.fixture(scope='module')
def ops():
    from kfac_b200 import ops as _ops
    return _ops


@pytest.fixture()
def two_stage_mode():
    from kfac_b200 import _native
    lib = _native.load()
    _native.check(lib.kfac_eigh_set_mode(2))
    yield lib
    _native.check(lib.kfac_eigh_set_mode(0))


def factor(n: int, seed: int) -> torch.Tensor:
    g = torch.Generator(device='cuda').manual_seed(seed)
    x = torch.relu(torch.randn(4 * n, n, device='cuda', generator=g))
    return (0.95 * torch.eye(n, device='cuda') + 0.05 * (x.t() @ x) / (4 * n)).contiguous()


def check(f, q, d):
    n = f.shape[0]
    fd, qd, dd = f.double(), q.double(), d.double()
    assert torch.isfinite(q).all() and torch.isfinite(d).all()
    assert (d[1:] >= d[:-1]).all()
    rec = (torch.linalg.norm(qd @ torch.diag(dd) @ qd.t() - fd) / torch.linalg.norm(fd)).item()
    orth = (qd.t() @ qd - torch.eye(n, device='cuda', dtype=torch.double)).abs().max().item()
    ref = torch.linalg.eigvalsh(fd)
    ev = ((dd - ref).abs().max() / ref.abs().max()).item()
    assert rec < REC_TOL and orth < ORTH_TOL and ev < EVAL_TOL, (n, rec, orth, ev)


def band_to_dense(band: torch.Tensor, n: int) -> torch.Tensor:
    out = torch.zeros(n, n, dtype=torch.float64, device=band.device)
    for d in range(33):
        idx = torch.arange(n - d, device=band.device)
        out[idx + d, idx] = band[:n - d, d].double()
        out[idx, idx + d] = band[:n - d, d].double()
    return out


@pytest.mark.parametrize('n', [97, 200, 700])
def test_stages_preserve_the_spectrum_and_use_the_model_layouts(n):
    from kfac_b200 import _native
    lib = _native.load()
    f = factor(n, seed=n)
    tiles = lib.kfac_debug_two_stage_tiles(n)
    w = torch.empty(n, n, device='cuda')
    tau1 = torch.empty(n, device='cuda')
    band = torch.empty(n, 64, device='cuda')
    d = torch.empty(n, device='cuda')
    e = torch.empty(n, device='cuda')
    v2 = torch.zeros(max(tiles, 1), 32, 32, device='cuda')
    tau2 = torch.zeros(max(tiles, 1), 32, device='cuda')
    need = lib.kfac_debug_two_stage_workspace_bytes(n)
    ws = torch.empty(need, dtype=torch.uint8, device='cuda')
    _native.check(lib.kfac_debug_two_stage(f.data_ptr(), n, w.data_ptr(), tau1.data_ptr(), band.data_ptr(),
                                           d.data_ptr(), e.data_ptr(), v2.data_ptr(), tau2.data_ptr(), ws.data_ptr(),
                                           need, torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    ref = torch.linalg.eigvalsh(f.double())
    # stage 1: a band of half-width 32 with the spectrum of F; nothing stored beyond the band
    assert not band[:, 33:].any()
    ev_band = torch.linalg.eigvalsh(band_to_dense(band, n))
    assert ((ev_band - ref).abs().max() / ref.abs().max()).item() < EVAL_TOL
    # stage-1 reflectors: row k holds v_k with its unit entry at column k + 32, nothing between the diagonal and it
    nref = n - 33
    rows = torch.arange(nref, device='cuda')
    assert torch.equal(w[rows, rows + 32], torch.ones(nref, device='cuda'))
    assert not torch.triu(w[:nref], 1).tril(31).any()
    assert not torch.triu(w[nref:], 1 + nref).any() and not tau1[nref:].any()
    assert ((tau1 >= 0) & (tau1 <= 2)).all()
    # stage 2: a tridiagonal matrix with the spectrum of F; one reflector per (sweep, step), tau in [0, 2]
    tm = torch.diag(d.double()) + torch.diag(e[:n - 1].double(), 1) + torch.diag(e[:n - 1].double(), -1)
    ev_t = torch.linalg.eigvalsh(tm)
    assert ((ev_t - ref).abs().max() / ref.abs().max()).item() < EVAL_TOL
    assert ((tau2 >= 0) & (tau2 <= 2)).all() and torch.isfinite(v2).all()
    used = tau2 != 0
    assert torch.equal(v2[..., 0][used], torch.ones_like(v2[..., 0][used]))   # unit leading entry


@pytest.mark.parametrize('n', [96, 97, 129, 161, 257, 545, 600, 1024, 1537, 2561])
def test_two_stage_eigh(ops, two_stage_mode, n):
    f = factor(n, seed=1000 + n)
    (q,), (d,) = ops.eigh_batched([f], clamp_nonneg=True)
    check(f, q, d)
    assert ops.eigh_last_paths()['two_stage'] == 1


def test_two_stage_mixed_batch(ops, two_stage_mode):
    """Big and small matrices in one call: the ones below n = 96 stay on the one-stage kernels."""
    ns = (2304, 1152, 640, 200, 64, 7, 1)
    fs = [factor(n, seed=50 + i) for i, n in enumerate(ns)]
    qs, ds = ops.eigh_batched(fs, clamp_nonneg=True)
    for f, q, d in zip(fs, qs, ds):
        check(f, q, d)
    paths = ops.eigh_last_paths()
    assert paths['two_stage'] == 4 and paths['full'] == 3, paths


def test_default_dispatch_by_size_and_batch(ops):
    """Default mode: a call with four or more matrices of n >= 1000 sends those through the two-stage path; a matrix
    that is (almost) alone only from n = 4096 (the measured crossovers, kfac_eigh_set_mode in include/kfac_b200.h)."""
    from kfac_b200 import _native
    lib = _native.load()
    _native.check(lib.kfac_eigh_set_mode(0))
    fs = [factor(n, seed=n + i) for i, n in enumerate((1000, 1000, 1024, 1100, 999))]
    qs, ds = ops.eigh_batched(fs)
    for f, q, d in zip(fs, qs, ds):
        check(f, q, d)
    paths = ops.eigh_last_paths()
    assert paths['two_stage'] == 4 and paths['full'] == 1, paths
    (q,), (d,) = ops.eigh_batched([fs[0]])
    assert ops.eigh_last_paths()['two_stage'] == 0
    f = factor(4096, seed=4096)
    (q,), (d,) = ops.eigh_batched([f])
    check(f, q, d)
    assert ops.eigh_last_paths()['two_stage'] == 1
    assert lib.kfac_eigh_set_mode(50) != 0 and b'mode must be' in lib.kfac_last_error()


def test_low_rank_plus_identity_two_stage(ops, two_stage_mode):
    """Early-training conv factor: a huge cluster of equal eigenvalues (zero columns in the panels, tau = 0 paths)."""
    n, m = 1200, 150
    g = torch.Generator(device='cuda').manual_seed(3)
    x = torch.relu(torch.randn(m, n, device='cuda', generator=g))
    f = (0.95 * torch.eye(n, device='cuda') + 0.05 * (x.t() @ x) / m).contiguous()
    (q,), (d,) = ops.eigh_batched([f], clamp_nonneg=True)
    check(f, q, d)
    eye = torch.eye(300, device='cuda')   # everything deflates, every reflector is trivial
    (q,), (d,) = ops.eigh_batched([eye])
    assert torch.allclose(d, torch.ones(300, device='cuda'), atol=1e-6)
    assert (q.double().t() @ q.double() - torch.eye(300, device='cuda', dtype=torch.double)).abs().max() < 1e-5
