"""GPU parity tests, through the C ABI (kfac_b200.ops -> ctypes -> libkfac_b200.so), against
  * the committed outputs of the unmodified reference (tests/golden/*.pt),
  * the CPU oracle (oracle/kfac_oracle.py) on seeded inputs,
  * size-independent properties at BASELINE.json's full sizes (symmetry, linearity, reconstruction).
Tolerances are the north star's: factors <= 1e-5, preconditioned gradients <= 1e-4 relative Frobenius error."""
from __future__ import annotations

import os

import pytest
import torch

from oracle import kfac_oracle as orc

pytestmark = pytest.mark.gpu

FACTOR_TOL = 1e-5
GRAD_TOL = 1e-4


def relerr(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return (torch.linalg.norm(a - b) / torch.linalg.norm(b).clamp_min(1e-300)).item()


def _load(golden_dir, name):
    return torch.load(os.path.join(golden_dir, name), weights_only=False)


@pytest.fixture(scope='module')
def ops():
    from kfac_b200 import ops as _ops
    return _ops


def test_get_cov_known_answers(ops, golden_dir):
    """The reference's exact vectors (tests/layers/utils_test.py:24-56) through the GPU kernel. The b=None cases
    are the hot path (A = a^T a / scale); integers this small are exact in the 3xTF32 split."""
    for c in _load(golden_dir, 'kat_get_cov.pt'):
        if c['b'] is not None:
            continue
        a = c['a'].cuda()
        m = a.shape[0]
        scale = c['scale'] if c['scale'] is not None else m
        out = torch.empty(a.shape[1], a.shape[1], device='cuda')
        ops.factor_linear(a, False, out, alpha=m / scale)
        assert torch.equal(out.cpu(), c['expected'])
        assert torch.equal(out, out.t())


def test_factors_match_reference_goldens(ops, golden_dir):
    for c in _load(golden_dir, 'factors.pt'):
        a, g = c['a'].cuda(), c['g'].cuda()
        A = torch.empty_like(c['A'], device='cuda')
        G = torch.empty_like(c['G'], device='cuda')
        if c['kind'] == 'linear':
            ops.factor_linear(a, c['bias'], A)
            ops.factor_linear(g, False, G)
        else:
            ops.factor_conv2d_a(a, c['kernel'], c['stride'], c['padding'], c['bias'], A)
            ops.factor_conv2d_g(g, G)
        assert relerr(A, c['A']) < FACTOR_TOL, (c['kind'], relerr(A, c['A']))
        assert relerr(G, c['G']) < FACTOR_TOL
        assert torch.equal(A, A.t()) and torch.equal(G, G.t())


@pytest.mark.parametrize('shape', [
    # (N, C, H, W, k, stride, pad, bias): ragged sizes, 1x1, strided 7x7, non-square kernels are in the goldens
    (1, 1, 1, 1, 1, 1, 0, True),
    (2, 3, 5, 5, 5, 1, 0, False),
    (3, 7, 13, 11, 3, 2, 1, True),
    (8, 64, 28, 28, 3, 1, 1, False),
    (4, 256, 14, 14, 1, 1, 0, False),
    (2, 3, 64, 64, 7, 2, 3, False),
])
def test_conv_factors_match_oracle(ops, shape):
    nb, c, h, w, k, s, p, bias = shape
    torch.manual_seed(sum(shape))
    x = torch.relu(torch.randn(nb, c, h, w))
    ref = orc.conv2d_a_factor(x.double(), (k, k), (s, s), (p, p), bias)
    out = torch.empty(ref.shape, device='cuda')
    ops.factor_conv2d_a(x.cuda(), (k, k), (s, s), (p, p), bias, out)
    assert relerr(out, ref) < FACTOR_TOL
    oh = (h + 2 * p - k) // s + 1
    g = torch.randn(nb, 5, oh, oh) / (nb * oh * oh)
    refg = orc.conv2d_g_factor(g.double())
    outg = torch.empty(5, 5, device='cuda')
    ops.factor_conv2d_g(g.cuda(), outg)
    assert relerr(outg, refg) < FACTOR_TOL


def test_running_average_fused_and_accumulated(ops):
    """alpha/beta epilogue == update_a_factor (base.py:374-388); accumulate flag == micro-batch sum (:351-356)."""
    torch.manual_seed(0)
    x1, x2 = torch.randn(40, 30), torch.randn(40, 30)
    a1, a2 = orc.linear_a_factor(x1.double(), True), orc.linear_a_factor(x2.double(), True)
    ref = orc.running_average(None, a1 + a2, 2, 0.9)
    ref = orc.running_average(ref, a1, 1, 0.9)
    batch = torch.empty(31, 31, device='cuda')
    ops.factor_linear(x1.cuda(), True, batch)
    ops.factor_linear(x2.cuda(), True, batch, beta=1.0)
    f = torch.empty(31, 31, device='cuda')
    ops.set_identity(f)
    ops.axpby(f, batch, 0.1 / 2, 0.9)
    ops.factor_linear(x1.cuda(), True, f, alpha=0.1, beta=0.9)  # fused second update
    assert relerr(f, ref) < FACTOR_TOL
    # AMP: grad-output divided by the scaler before the factor (base.py:363-365)
    g = torch.randn(16, 12)
    out = torch.empty(12, 12, device='cuda')
    ops.factor_linear((g * 1024).cuda(), False, out, in_scale=1 / 1024)
    assert relerr(out, orc.linear_g_factor(g.double())) < FACTOR_TOL


def test_full_size_factor_properties(ops):
    """ResNet-50 shapes (BASELINE.json configs[1]): conv1 A (147, M=401408), layer1 3x3 A (576, M=100352),
    layer4 3x3 A (4608, M=1568), fc A (2049, M=32): symmetry, linearity in alpha, PSD diagonal, and a checksum
    identity: sum(A) == |colsum of patches|^2 / (M S^2) computed independently with torch ops."""
    torch.manual_seed(1)
    for (nb, c, h, k, s, p) in [(32, 3, 224, 7, 2, 3), (32, 64, 56, 3, 1, 1), (32, 512, 7, 3, 1, 1)]:
        x = torch.relu(torch.randn(nb, c, h, h, device='cuda'))
        n = c * k * k
        a1 = torch.empty(n, n, device='cuda')
        a2 = torch.empty(n, n, device='cuda')
        ops.factor_conv2d_a(x, (k, k), (s, s), (p, p), False, a1)
        ops.factor_conv2d_a(x, (k, k), (s, s), (p, p), False, a2, alpha=3.0)
        assert torch.equal(a1, a1.t())
        assert relerr(a2, 3.0 * a1) < 1e-6
        assert (a1.diagonal() >= 0).all()
        oh = (h + 2 * p - k) // s + 1
        cols = torch.nn.functional.unfold(x.double(), k, padding=p, stride=s)      # (N, n, S)
        checksum = (cols.sum(dim=1) ** 2).sum() / (nb * oh * oh) / (oh * oh) ** 2   # sum_m (row sum)^2 / (M S^2)
        assert abs(a1.double().sum().item() - checksum.item()) / checksum.item() < 1e-5
    x = torch.randn(32, 2048, device='cuda')
    a = torch.empty(2049, 2049, device='cuda')
    ops.factor_linear(x, True, a)
    assert relerr(a, orc.linear_a_factor(x.double().cpu(), True)) < FACTOR_TOL
    assert a[-1, -1].item() == 1.0


def test_gemm_matches_fp64(ops):
    torch.manual_seed(2)
    for (m, n, k, ta, tb) in [(1, 1, 1, False, False), (33, 65, 17, True, False), (257, 300, 1000, False, True),
                              (128, 256, 4096, True, True)]:
        a = torch.randn((k, m) if ta else (m, k), device='cuda')
        b = torch.randn((n, k) if tb else (k, n), device='cuda')
        ref = (a.double().t() if ta else a.double()) @ (b.double().t() if tb else b.double())
        assert relerr(ops.gemm(a, b, transa=ta, transb=tb), ref) < 5e-6


def test_elementwise(ops, golden_dir):
    for c in _load(golden_dir, 'triu.pt'):
        x = c['x'].cuda()
        packed = ops.triu_pack(x)
        assert torch.equal(packed.cpu(), c['triu'])
        assert torch.equal(ops.triu_unpack(packed, x.shape[0]).cpu(), c['filled'])
    dg, da = torch.rand(7, device='cuda'), torch.rand(9, device='cuda')
    assert relerr(ops.eigen_outer(dg, da, 0.01), orc.eigen_outer(dg.cpu().double(), da.cpu().double(), 0.01)) < 1e-6
    f = torch.randn(6, 6, device='cuda')
    ref = f.clone()
    ref.diagonal().add_(0.25)
    assert torch.allclose(ops.add_diag(f, 0.25), ref)


def test_layer_pipeline_matches_reference_goldens(ops, golden_dir):
    """factors -> eigendecomposition / inverse -> preconditioned gradient -> scaled write-back, per layer,
    against what the unmodified reference produced (tests/golden/layers.pt)."""
    for c in _load(golden_dir, 'layers.pt'):
        last = c['updates'][-1]
        A, G = last['A'].cuda(), last['G'].cuda()
        w = c['w_grad'].cuda()
        wg = w.reshape(w.shape[0], -1).contiguous()
        bg = c['b_grad'].cuda() if c['b_grad'] is not None else None
        vg = torch.zeros(1, dtype=torch.float64, device='cuda')
        if c['method'] == 'inverse':
            a_inv, _ = ops.inverse_newton_schulz(A, 0.01)
            g_inv, _ = ops.inverse_newton_schulz(G, 0.01)
            assert relerr(a_inv, c['a_inv']) < GRAD_TOL and relerr(g_inv, c['g_inv']) < GRAD_TOL
            pg = ops.precondition_inverse(wg, bg, a_inv, g_inv, vg_accum=vg)
        else:
            (qa, qg), (da, dg) = ops.eigh_batched([A, G])
            assert (da >= 0).all() and (dg >= 0).all()
            if c['method'] == 'eigen_prediv':
                pg = ops.precondition_eigen(wg, bg, qa, qg, dgda=ops.eigen_outer(dg, da, 0.01), vg_accum=vg)
            else:
                pg = ops.precondition_eigen(wg, bg, qa, qg, da=da, dg=dg, damping=0.01, vg_accum=vg)
        assert relerr(pg, c['precond']) < GRAD_TOL, (c['kind'], c['method'], relerr(pg, c['precond']))
        grad = orc.stacked_grad(c['w_grad'], c['b_grad']).double()
        assert abs(vg.item() - (c['precond'].double() * grad).sum().item()) <= 1e-4 * abs(vg.item()) + 1e-12
        half = torch.full((1,), 0.5, device='cuda')
        w_out = torch.empty_like(wg)
        b_out = torch.empty_like(bg) if bg is not None else None
        ops.scale_write_grads(pg, half, w_out, b_out)
        assert relerr(w_out.reshape(c['w_out'].shape), c['w_out']) < GRAD_TOL
        if bg is not None:
            assert relerr(b_out, c['b_out']) < GRAD_TOL


@pytest.mark.parametrize('n', [1, 2, 3, 31, 64, 147, 513, 1024])
def test_eigh_reconstruction(ops, n):
    """Eigenvectors are not unique: compare after reconstruction (north star) + orthogonality + ordering."""
    torch.manual_seed(n)
    x = torch.relu(torch.randn(4 * n, n, device='cuda'))
    f = (0.95 * torch.eye(n, device='cuda') + 0.05 * x.t() @ x / (4 * n)).contiguous()
    (q,), (d,) = ops.eigh_batched([f], clamp_nonneg=True)
    qd, dd, fd = q.double(), d.double(), f.double()
    assert (d[1:] >= d[:-1]).all() and (d >= 0).all()
    assert relerr(qd @ torch.diag(dd) @ qd.t(), fd) < 1e-5
    assert (qd.t() @ qd - torch.eye(n, device='cuda', dtype=torch.double)).abs().max().item() < 1e-5
    ref = torch.linalg.eigvalsh(fd)
    assert ((dd - ref).abs().max() / ref.abs().max()).item() < 1e-6


def test_eigh_edge_cases(ops):
    """Repeated eigenvalues (identity: everything deflates), negative eigenvalues clamped (eigen.py:320), a
    heterogeneous batch in one call, and a matrix with zero off-diagonals."""
    eye = torch.eye(40, device='cuda')
    (q,), (d,) = ops.eigh_batched([eye])
    assert torch.equal(d, torch.ones(40, device='cuda')) and relerr(q @ q.t(), eye) < 1e-6
    m = torch.diag(torch.tensor([-2.0, -1.0, 0.5, 3.0], device='cuda'))
    (q,), (d,) = ops.eigh_batched([m], clamp_nonneg=True)
    assert torch.equal(d.cpu(), torch.tensor([0.0, 0.0, 0.5, 3.0]))
    (q,), (d,) = ops.eigh_batched([m], clamp_nonneg=False)
    assert relerr(q @ torch.diag(d) @ q.t(), m) < 1e-6
    torch.manual_seed(5)
    fs = []
    for n in (5, 64, 1, 200, 33):
        x = torch.randn(n, n, device='cuda')
        fs.append(((x + x.t()) / 2).contiguous())
    qs, ds = ops.eigh_batched(fs, clamp_nonneg=False)
    for f, q, d in zip(fs, qs, ds):
        assert relerr(q.double() @ torch.diag(d.double()) @ q.double().t(), f) < 1e-5


def test_precondition_matches_oracle_resnet_shapes(ops):
    """(g, a) of ResNet-50 layers incl. fc (1000 x 2049 with bias): native pipeline vs the fp64 arbiter."""
    torch.manual_seed(3)
    for (g, a_w, bias) in [(64, 147, False), (256, 64, False), (128, 1152, False), (1000, 2048, True)]:
        a = a_w + int(bias)
        xa = torch.relu(torch.randn(4 * a, a, device='cuda'))
        A = (0.95 * torch.eye(a, device='cuda') + 0.05 * xa.t() @ xa / (4 * a)).contiguous()
        xg = torch.randn(4 * g, g, device='cuda')
        G = (0.95 * torch.eye(g, device='cuda') + 0.05 * xg.t() @ xg / (4 * g)).contiguous()
        wg = torch.randn(g, a_w, device='cuda')
        bg = torch.randn(g, device='cuda') if bias else None
        (qa, qg), (da, dg) = ops.eigh_batched([A, G])
        pg = ops.precondition_eigen(wg, bg, qa, qg, da=da, dg=dg, damping=0.001)
        grad = orc.stacked_grad(wg.cpu(), bg.cpu() if bias else None)
        ref = orc.precondition_eigen_fp64(grad, A.cpu(), G.cpu(), 0.001)
        assert relerr(pg, ref) < GRAD_TOL, (g, a, relerr(pg, ref))


def test_klclip_scale_and_grad_dot(ops):
    torch.manual_seed(4)
    v, w, b = torch.randn(5, 8, device='cuda'), torch.randn(5, 7, device='cuda'), torch.randn(5, device='cuda')
    vg = torch.zeros(1, dtype=torch.float64, device='cuda')
    ops.grad_dot(v, w, b, vg)
    ref = (v[:, :7] * w).sum().item() + (v[:, 7] * b).sum().item()
    assert abs(vg.item() - ref) < 1e-5 * abs(ref) + 1e-6
    scale = torch.empty(1, device='cuda')
    ops.klclip_scale(vg, 0.1, 0.001, scale)
    assert abs(scale.item() - orc.kl_clip_scale([(v[:, :7].cpu(), w.cpu()), (v[:, 7].cpu(), b.cpu())], 0.1, 0.001)) < 1e-6
    ops.klclip_scale(torch.zeros(1, dtype=torch.float64, device='cuda'), 0.1, 0.001, scale)
    assert scale.item() == 1.0   # vg == 0 -> 1.0 (base_preconditioner.py:432)


def test_workspace_too_small_is_an_error():
    from kfac_b200 import _native
    lib = _native.load()
    x = torch.randn(100000, 64, device='cuda')
    f = torch.empty(64, 64, device='cuda')
    need = lib.kfac_factor_workspace_bytes(64, 100000)
    assert need > 0
    rc = lib.kfac_factor_linear(x.data_ptr(), 100000, 64, 0, 1.0, 1.0, 0.0, f.data_ptr(), None, 0,
                                torch.cuda.current_stream().cuda_stream)
    assert rc != 0 and b'workspace too small' in lib.kfac_last_error()


def test_eigh_symmetric_path_odd_size_and_mixed_batch(ops):
    """n >= 1024 takes the lower-triangle tridiagonalisation; 1537 is odd (padded leading dimension, ragged last
    column block) and shares one call with small matrices that take the full-matrix path in the same rounds."""
    torch.manual_seed(11)
    fs = []
    for n in (1537, 96, 1100, 7):
        x = torch.relu(torch.randn(2 * n, n, device='cuda'))
        fs.append((0.9 * torch.eye(n, device='cuda') + 0.1 * x.t() @ x / (2 * n)).contiguous())
    qs, ds = ops.eigh_batched(fs, clamp_nonneg=True)
    for f, q, d in zip(fs, qs, ds):
        n = f.shape[0]
        qd, dd = q.double(), d.double()
        assert (d[1:] >= d[:-1]).all()
        assert relerr(qd @ torch.diag(dd) @ qd.t(), f.double()) < 1e-5
        assert (qd.t() @ qd - torch.eye(n, device='cuda', dtype=torch.double)).abs().max().item() < 1e-5


def test_stage_timers(ops):
    """kfac_profile_enable / kfac_profile_last_ms: refused before a profiled call, three positive stage times and
    the number of tridiagonalisation launches after one; no effect on the results."""
    import ctypes
    from kfac_b200 import _native
    lib = _native.load()
    out = (ctypes.c_float * 4)()
    torch.manual_seed(3)
    x = torch.randn(600, 300, device='cuda')
    f = (x.t() @ x / 600).contiguous()
    (q0,), (d0,) = ops.eigh_batched([f])
    lib.kfac_profile_enable(1)
    try:
        (q1,), (d1,) = ops.eigh_batched([f])
        assert lib.kfac_profile_last_ms(out) == 0
        assert all(out[i] > 0 for i in range(3)) and out[3] >= 1
        # the finer split exists only when a matrix of the call took the two-stage path
        det = (ctypes.c_float * 6)()
        assert lib.kfac_profile_detail_ms(det) != 0
        assert lib.kfac_eigh_set_mode(2) == 0
        try:
            (q2,), (d2,) = ops.eigh_batched([f])
        finally:
            lib.kfac_eigh_set_mode(0)
        assert lib.kfac_profile_detail_ms(det) == 0
        assert all(det[i] > 0 for i in range(4)) and det[4] >= 1 and det[5] >= 1
        assert lib.kfac_profile_last_ms(out) == 0
        # band reduction + chasing lie inside the tridiagonalisation stage, Q2 + Q1 make up the back-transformation
        assert det[0] + det[1] <= out[0] * 1.05 + 0.05
        assert abs(det[2] + det[3] - out[2]) <= 0.05 * out[2] + 0.05
    finally:
        lib.kfac_profile_enable(0)
    assert torch.equal(d0, d1) and torch.equal(q0, q1)
    assert (d2 - d0).abs().max().item() <= 2e-6 * d0.abs().max().item()


@pytest.mark.parametrize('dtype', [torch.float16, torch.bfloat16])
def test_factors_from_half_precision_inputs(ops, dtype):
    """fp16 / bf16 activations and grad-outputs (torch.autocast) are read in place: the factor must equal the fp32
    factor of the same (already rounded) values -- 16-bit floats are exact in tf32, so there is no extra error."""
    torch.manual_seed(21)
    # linear, with and without the bias column; 3-D input (batch, seq, features)
    x = torch.randn(6, 50, 96, device='cuda').to(dtype)
    for bias in (True, False):
        n = 96 + int(bias)
        f_h = ops.factor_linear(x, bias, torch.empty(n, n, device='cuda'))
        f_f = ops.factor_linear(x.float(), bias, torch.empty(n, n, device='cuda'))
        assert relerr(f_h, f_f) < 2e-6
    # running average into an existing factor
    f0 = torch.eye(97, device='cuda')
    f_h = ops.factor_linear(x, True, f0.clone(), alpha=0.05, beta=0.95)
    f_f = ops.factor_linear(x.float(), True, f0.clone(), alpha=0.05, beta=0.95)
    assert relerr(f_h, f_f) < 2e-6
    # conv A: 3x3 stride 1 pad 1 with bias, 1x1 (rerouted to the (N, C, S) loader), strided 3x3
    xc = torch.relu(torch.randn(4, 16, 14, 14, device='cuda')).to(dtype)
    for (k, s, p, bias) in [((3, 3), (1, 1), (1, 1), True), ((1, 1), (1, 1), (0, 0), False), ((3, 3), (2, 2), (1, 1), False)]:
        n = 16 * k[0] * k[1] + int(bias)
        f_h = ops.factor_conv2d_a(xc, k, s, p, bias, torch.empty(n, n, device='cuda'))
        f_f = ops.factor_conv2d_a(xc.float(), k, s, p, bias, torch.empty(n, n, device='cuda'))
        assert relerr(f_h, f_f) < 2e-6, (k, s, p, bias)
    # conv G with the AMP unscale factor
    g = (torch.randn(4, 24, 7, 7, device='cuda') * 64).to(dtype)
    f_h = ops.factor_conv2d_g(g, torch.empty(24, 24, device='cuda'), in_scale=1.0 / 64)
    f_f = ops.factor_conv2d_g(g.float(), torch.empty(24, 24, device='cuda'), in_scale=1.0 / 64)
    assert relerr(f_h, f_f) < 2e-6
    with pytest.raises(RuntimeError):
        ops.factor_linear(x.double(), True, torch.empty(97, 97, device='cuda'))
