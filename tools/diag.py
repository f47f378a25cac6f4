"""GPU bring-up diagnostics: runs kernel checks in child processes with timeouts and prints error stats.

Usage (on a GPU box):  python tools/diag.py [group ...]      -> writes gpurun_out/diag.log
Not part of the test-suite; used while developing kernels to get many data points from one gpurun call.
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'distributed-kfac-pytorch_b200'))


def rel(a, b):
    import torch
    return (torch.linalg.norm((a.double() - b.double())) / torch.linalg.norm(b.double())).item()


def emit(**kw):
    print(json.dumps(kw), flush=True)


def time_it(fn, iters=10):
    import torch
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def g_gemm_basic():
    import torch
    from kfac_b200 import ops
    torch.manual_seed(0)
    for (m, n, k) in [(128, 256, 32), (128, 256, 256), (128, 32, 64), (64, 64, 64), (200, 147, 100),
                      (1024, 1024, 1024), (512, 4608, 512)]:
        for ta in (False, True):
            for tb in (False, True):
                a = torch.randn((k, m) if ta else (m, k), device='cuda')
                b = torch.randn((n, k) if tb else (k, n), device='cuda')
                ref = (a.double().t() if ta else a.double()) @ (b.double().t() if tb else b.double())
                out = ops.gemm(a, b, transa=ta, transb=tb)
                torch.cuda.synchronize()
                tf = (a.t() if ta else a) @ (b.t() if tb else b)
                emit(case='gemm', m=m, n=n, k=k, ta=ta, tb=tb, rel=rel(out, ref), rel_torch_fp32=rel(tf, ref))


def g_gemm_beta():
    import torch
    from kfac_b200 import ops
    torch.manual_seed(1)
    a = torch.randn(300, 5000, device='cuda')
    b = torch.randn(5000, 200, device='cuda')
    c = torch.randn(300, 200, device='cuda')
    ref = 0.5 * a.double() @ b.double() + 2.0 * c.double()
    out = ops.gemm(a, b, alpha=0.5, beta=2.0, out=c.clone())
    emit(case='gemm_beta_splitk', rel=rel(out, ref))


def g_factor():
    import torch
    import torch.nn.functional as F
    from kfac_b200 import ops
    torch.manual_seed(2)
    # linear
    for (m, n_in, bias) in [(32, 64, True), (128, 65, False), (4096, 1024, True), (32, 2048, True), (7, 10, True)]:
        x = torch.randn(m, n_in, device='cuda')
        xd = x.double()
        if bias:
            xd = torch.cat([xd, torch.ones(m, 1, device='cuda', dtype=torch.double)], 1)
        ref = xd.t() @ xd / m
        n = n_in + int(bias)
        out = torch.empty(n, n, device='cuda')
        ops.factor_linear(x, bias, out)
        emit(case='factor_linear', m=m, n_in=n_in, bias=bias, rel=rel(out, ref), asym=(out - out.t()).abs().max().item())
    # ema
    x = torch.randn(256, 100, device='cuda')
    f = torch.eye(101, device='cuda')
    ref = 0.95 * torch.eye(101, device='cuda', dtype=torch.double) + 0.05 * (
        torch.cat([x.double(), torch.ones(256, 1, device='cuda', dtype=torch.double)], 1).t()
        @ torch.cat([x.double(), torch.ones(256, 1, device='cuda', dtype=torch.double)], 1) / 256)
    ops.factor_linear(x, True, f, alpha=0.05, beta=0.95)
    emit(case='factor_linear_ema', rel=rel(f, ref))
    # conv A
    for (nb, c, h, w, k, s, p, bias) in [(2, 3, 8, 8, 3, 1, 1, True), (4, 16, 14, 14, 3, 2, 1, False),
                                        (2, 3, 32, 32, 7, 2, 3, False), (8, 64, 28, 28, 1, 1, 0, False),
                                        (4, 64, 28, 28, 3, 1, 1, True), (32, 128, 28, 28, 3, 1, 1, False), (32, 3, 224, 224, 7, 2, 3, False), (32, 64, 56, 56, 3, 1, 1, False),
                                        (2, 5, 9, 11, 2, 1, 0, True)]:
        x = torch.relu(torch.randn(nb, c, h, w, device='cuda'))
        cols = F.unfold(x.double(), k, padding=p, stride=s)  # (N, C*k*k, L)
        L = cols.shape[-1]
        pm = cols.transpose(1, 2).reshape(-1, cols.shape[1])
        if bias:
            pm = torch.cat([pm, torch.ones(pm.shape[0], 1, device='cuda', dtype=torch.double)], 1)
        pm = pm / L
        ref = pm.t() @ pm / pm.shape[0]
        n = c * k * k + int(bias)
        out = torch.empty(n, n, device='cuda')
        ops.factor_conv2d_a(x, (k, k), (s, s), (p, p), bias, out)
        emit(case='factor_conv_a', shape=[nb, c, h, w, k, s, p, bias], n=n, M=pm.shape[0], rel=rel(out, ref))
    # conv G
    for (nb, c, oh, ow) in [(2, 16, 8, 8), (32, 64, 56, 56), (32, 512, 7, 7), (3, 10, 5, 7)]:
        g = torch.randn(nb, c, oh, ow, device='cuda') / (nb * oh * ow)
        gm = g.double().permute(0, 2, 3, 1).reshape(-1, c) / (oh * ow)
        ref = gm.t() @ gm / gm.shape[0]
        out = torch.empty(c, c, device='cuda')
        ops.factor_conv2d_g(g, out)
        emit(case='factor_conv_g', shape=[nb, c, oh, ow], rel=rel(out, ref))


def g_precond():
    import torch
    from kfac_b200 import ops
    torch.manual_seed(3)
    for (g, a_w, bias) in [(10, 20, True), (64, 576, False), (256, 1024, True), (1000, 2048, True)]:
        a = a_w + int(bias)
        wg = torch.randn(g, a_w, device='cuda')
        bg = torch.randn(g, device='cuda') if bias else None
        qa = torch.linalg.qr(torch.randn(a, a, device='cuda'))[0].contiguous()
        qg = torch.linalg.qr(torch.randn(g, g, device='cuda'))[0].contiguous()
        da = torch.rand(a, device='cuda')
        dg = torch.rand(g, device='cuda')
        grad = torch.cat([wg, bg.view(-1, 1)], 1) if bias else wg
        gd, qad, qgd = grad.double(), qa.double(), qg.double()
        v1 = qgd.t() @ gd @ qad
        v2 = v1 / (torch.outer(dg.double(), da.double()) + 0.001)
        ref = qgd @ v2 @ qad.t()
        vg = torch.zeros(1, dtype=torch.double, device='cuda')
        out = ops.precondition_eigen(wg, bg, qa, qg, da=da, dg=dg, damping=0.001, vg_accum=vg)
        dgda = ops.eigen_outer(dg, da, 0.001)
        out2 = ops.precondition_eigen(wg, bg, qa, qg, dgda=dgda)
        emit(case='precond_eigen', g=g, a=a, rel=rel(out, ref), rel_prediv=rel(out2, ref),
             vg=vg.item(), vg_ref=(ref * gd).sum().item())
        ainv = torch.randn(a, a, device='cuda')
        ginv = torch.randn(g, g, device='cuda')
        refi = ginv.double() @ gd @ ainv.double()
        outi = ops.precondition_inverse(wg, bg, ainv, ginv)
        emit(case='precond_inverse', g=g, a=a, rel=rel(outi, refi))
        sc = torch.empty(1, device='cuda')
        ops.klclip_scale(vg, 0.1, 0.001, sc)
        wg2, bg2 = wg.clone(), (bg.clone() if bias else None)
        ops.scale_write_grads(out, sc, wg2, bg2)
        emit(case='scale_write', scale=sc.item(), rel_w=rel(wg2, sc.item() * out[:, :a_w]),
             rel_b=(rel(bg2, sc.item() * out[:, a_w]) if bias else 0.0))


def g_elementwise():
    import torch
    from kfac_b200 import ops
    torch.manual_seed(4)
    for n in (5, 64, 147, 1000):
        f = torch.randn(n, n, device='cuda')
        f = f + f.t()
        p = ops.triu_pack(f)
        idx = torch.triu_indices(n, n, device='cuda')
        ok_pack = torch.equal(p, f[idx[0], idx[1]])
        f2 = ops.triu_unpack(p, n)
        emit(case='triu', n=n, pack_equal=ok_pack, roundtrip_equal=torch.equal(f, f2))
    y = torch.randn(1000003, device='cuda')
    x = torch.randn(1000003, device='cuda')
    ref = 0.3 * x + 0.7 * y
    ops.axpby(y, x, 0.3, 0.7)
    emit(case='axpby', maxerr=(y - ref).abs().max().item())
    e = torch.empty(33, 33, device='cuda')
    ops.set_identity(e)
    ops.add_diag(e, 0.5)
    emit(case='identity', ok=torch.equal(e, 1.5 * torch.eye(33, device='cuda')))


def g_perf():
    import torch
    from kfac_b200 import ops
    torch.manual_seed(5)
    for (m, n, k) in [(4096, 4096, 4096), (8192, 8192, 8192)]:
        a = torch.randn(m, k, device='cuda')
        b = torch.randn(n, k, device='cuda')
        out = torch.empty(m, n, device='cuda')
        ms = time_it(lambda: ops.gemm(a, b, transb=True, out=out))
        emit(case='perf_gemm_nt', m=m, n=n, k=k, ms=ms, tflops=2 * m * n * k / ms / 1e9)
        torch.backends.cuda.matmul.allow_tf32 = False
        ms = time_it(lambda: torch.matmul(a, b.t(), out=out))
        emit(case='perf_torch_fp32', m=m, n=n, k=k, ms=ms, tflops=2 * m * n * k / ms / 1e9)
        torch.backends.cuda.matmul.allow_tf32 = True
        ms = time_it(lambda: torch.matmul(a, b.t(), out=out))
        emit(case='perf_torch_tf32', m=m, n=n, k=k, ms=ms, tflops=2 * m * n * k / ms / 1e9)
        torch.backends.cuda.matmul.allow_tf32 = False
    for (nb, c, h, w, k, s, p) in [(32, 64, 56, 56, 3, 1, 1), (32, 256, 56, 56, 1, 1, 0), (32, 3, 224, 224, 7, 2, 3),
                                   (32, 512, 7, 7, 3, 1, 1), (32, 128, 28, 28, 3, 1, 1)]:
        x = torch.relu(torch.randn(nb, c, h, w, device='cuda'))
        n = c * k * k
        out = torch.empty(n, n, device='cuda')
        ms = time_it(lambda: ops.factor_conv2d_a(x, (k, k), (s, s), (p, p), False, out))
        oh = (h + 2 * p - k) // s + 1
        M = nb * oh * oh
        emit(case='perf_conv_a', shape=[nb, c, h, w, k, s, p], n=n, M=M, ms=ms, tflops_sym=M * n * (n + 1) / ms / 1e9,
             gbs=x.numel() * 4 / ms / 1e6)


def _kfac_like(n, seed, kind='relu'):
    import torch
    g = torch.Generator(device='cuda').manual_seed(seed)
    if kind == 'relu':
        x = torch.relu(torch.randn(4 * n, n, device='cuda', generator=g))
        return (0.95 * torch.eye(n, device='cuda') + 0.05 * x.t() @ x / (4 * n)).contiguous()
    if kind == 'lowrank':
        x = torch.randn(max(1, n // 8), n, device='cuda', generator=g)
        return (0.95 * torch.eye(n, device='cuda') + 0.05 * x.t() @ x / max(1, n // 8)).contiguous()
    if kind == 'rand':
        x = torch.randn(n, n, device='cuda', generator=g)
        return ((x + x.t()) / 2).contiguous()
    if kind == 'eye':
        return torch.eye(n, device='cuda')
    raise KeyError(kind)


def _eigh_check(fs, label):
    import torch
    from kfac_b200 import ops
    qs, ds = ops.eigh_batched(fs, clamp_nonneg=False)
    torch.cuda.synchronize()
    for f, q, d in zip(fs, qs, ds):
        n = f.shape[0]
        fd, qd, dd = f.double(), q.double(), d.double()
        rec = (torch.linalg.norm(qd @ torch.diag(dd) @ qd.t() - fd) / torch.linalg.norm(fd).clamp_min(1e-30)).item()
        orth = (qd.t() @ qd - torch.eye(n, device='cuda', dtype=torch.double)).abs().max().item()
        wref = torch.linalg.eigvalsh(fd)
        verr = ((dd - wref).abs().max() / wref.abs().max().clamp_min(1e-30)).item()
        emit(case='eigh', label=label, n=n, rec=rec, orth=orth, valerr=verr, sorted=bool((d[1:] >= d[:-1]).all().item()),
             finite=bool(torch.isfinite(q).all().item()))


def g_eigh_small():
    for n in (1, 2, 3, 5, 17, 64, 97, 100, 147):
        for kind in ('relu', 'rand', 'eye', 'lowrank'):
            _eigh_check([_kfac_like(n, n, kind)], kind)
    _eigh_check([_kfac_like(n, 7 + n, 'relu') for n in (33, 64, 5, 128, 100)], 'batch')


def g_eigh_mid():
    for n in (256, 300, 512, 1000, 1024):
        for kind in ('relu', 'lowrank'):
            _eigh_check([_kfac_like(n, n, kind)], kind)
    _eigh_check([_kfac_like(n, 3 * n, 'rand') for n in (257, 640)], 'rand')


def g_eigh_big():
    import torch
    from kfac_b200 import ops
    for n in (2048, 4608):
        f = _kfac_like(n, n, 'relu')
        _eigh_check([f], 'relu')
        ms = time_it(lambda: ops.eigh_batched([f]), iters=2)
        torch.backends.cuda.preferred_linalg_library('default')
        ms_t = time_it(lambda: torch.linalg.eigh(f), iters=2)
        emit(case='perf_eigh', n=n, ms=ms, torch_ms=ms_t)


def g_eigh_perf():
    import torch
    from kfac_b200 import ops
    for ns in ([64] * 16, [256] * 16, [512] * 8, [1024] * 8, [1024], [2304], [64, 64, 256, 64, 576, 64, 128, 512, 1152, 128, 256, 1024, 2304, 512, 2048]):
        fs = [_kfac_like(n, i + n, 'relu') for i, n in enumerate(ns)]
        ms = time_it(lambda: ops.eigh_batched(fs), iters=3)
        ms_t = time_it(lambda: [torch.linalg.eigh(f) for f in fs], iters=2)
        emit(case='perf_eigh_batch', ns=ns if len(ns) < 6 else [ns[0], len(ns)], ms=ms, torch_ms=ms_t)


def g_inverse():
    import torch
    from kfac_b200 import ops
    for n in (5, 64, 147, 512, 1024):
        f = _kfac_like(n, n, 'relu')
        out, iters = ops.inverse_newton_schulz(f, 0.001)
        ref = torch.linalg.inv(f.double() + 0.001 * torch.eye(n, device='cuda', dtype=torch.double))
        t32 = torch.linalg.inv(f + 0.001 * torch.eye(n, device='cuda'))
        emit(case='inverse', n=n, iters=iters, rel=rel(out, ref), rel_torch_fp32=rel(t32, ref))


GROUPS = {
    'eigh_small': g_eigh_small, 'eigh_mid': g_eigh_mid, 'eigh_big': g_eigh_big, 'eigh_perf': g_eigh_perf,
    'inverse': g_inverse,
    'gemm_basic': g_gemm_basic, 'gemm_beta': g_gemm_beta, 'factor': g_factor, 'precond': g_precond,
    'elementwise': g_elementwise, 'perf': g_perf,
}


def main():
    if len(sys.argv) >= 3 and sys.argv[1] == '--child':
        GROUPS[sys.argv[2]]()
        return
    groups = sys.argv[1:] or list(GROUPS)
    os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
    with open(os.path.join(ROOT, 'gpurun_out', 'diag.log'), 'a') as log:
        for g in groups:
            t0 = time.time()
            try:
                r = subprocess.run([sys.executable, __file__, '--child', g], capture_output=True, text=True, timeout=240)
                out, err, rc = r.stdout, r.stderr[-3000:], r.returncode
            except subprocess.TimeoutExpired as ex:
                out, err, rc = (ex.stdout or b'').decode() if isinstance(ex.stdout, bytes) else (ex.stdout or ''), 'TIMEOUT', -9
            msg = f'=== {g} rc={rc} {time.time() - t0:.1f}s\n{out}\n{err if rc != 0 else ""}\n'
            print(msg, flush=True)
            log.write(msg)
            log.flush()
            if rc == -9:
                print('stopping after timeout (possible hang)')
                break


if __name__ == '__main__':
    main()
