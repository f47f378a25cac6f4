"""Stage-by-stage check of the two-stage tridiagonalisation on the GPU (builder tool; the tests in
tests/test_two_stage_gpu.py are the driver-run version of the same checks).

    python tools/ts_check.py --n 257 [--model] [--full] [--time]

--model  compare W / tau1 / band / d / e / V2 / tau2 element-wise with oracle/twostage_model.py (small n only)
--full   run kfac_eigh_batched in one-stage and two-stage mode, report reconstruction / orthogonality / eigenvalues
--time   stage timers of both modes
"""
import argparse
import ctypes
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, 'distributed-kfac-pytorch_b200'), ROOT):
    sys.path.insert(0, p)
import numpy as np
import torch

from kfac_b200 import _native
from kfac_b200 import ops


def spd(n, seed=0):
    g = torch.Generator(device='cuda').manual_seed(seed)
    x = torch.relu(torch.randn(4 * n, n, device='cuda', generator=g))
    return (0.95 * torch.eye(n, device='cuda') + 0.05 * (x.t() @ x) / (4 * n)).contiguous()


def debug_two_stage(f):
    lib = _native.load()
    n = f.shape[0]
    tiles = lib.kfac_debug_two_stage_tiles(n)
    out = {
        'W': torch.empty(n, n, device='cuda'), 'tau1': torch.empty(n, device='cuda'),
        'band': torch.empty(n, 64, device='cuda'), 'd': torch.empty(n, device='cuda'),
        'e': torch.empty(n, device='cuda'), 'V2': torch.zeros(max(tiles, 1), 32, 32, device='cuda'),
        'tau2': torch.zeros(max(tiles, 1), 32, device='cuda'),
    }
    need = lib.kfac_debug_two_stage_workspace_bytes(n)
    ws = torch.empty(need, dtype=torch.uint8, device='cuda')
    _native.check(lib.kfac_debug_two_stage(f.data_ptr(), n, out['W'].data_ptr(), out['tau1'].data_ptr(),
                                           out['band'].data_ptr(), out['d'].data_ptr(), out['e'].data_ptr(),
                                           out['V2'].data_ptr(), out['tau2'].data_ptr(), ws.data_ptr(), need,
                                           torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return out


def band_to_dense(band, n, b=32):
    bd = torch.zeros(n, n, dtype=torch.float64, device=band.device)
    for d in range(b + 1):
        v = band[:n - d, d]
        idx = torch.arange(n - d, device=band.device)
        bd[idx + d, idx] = v
        bd[idx, idx + d] = v
    return bd


def rel(a, b):
    return (torch.linalg.norm(a.double() - b.double()) / max(torch.linalg.norm(b.double()).item(), 1e-300)).item()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--n', type=int, default=257)
    ap.add_argument('--model', action='store_true')
    ap.add_argument('--full', action='store_true')
    ap.add_argument('--time', action='store_true')
    ap.add_argument('--q', action='store_true', help='check Q1 / Q2 explicitly (O(n^3) torch work)')
    args = ap.parse_args()
    n = args.n
    lib = _native.load()
    f = spd(n, seed=n)
    fd = f.double()
    ev_ref = torch.linalg.eigvalsh(fd)
    t0 = time.time()
    out = debug_two_stage(f)
    print(f'[n={n}] debug_two_stage ran in {time.time() - t0:.2f}s (incl. lazy init)', flush=True)
    for k, v in out.items():
        if not torch.isfinite(v).all():
            print(f'[n={n}] NON-FINITE values in {k}')
    bd = band_to_dense(out['band'].double(), n)
    ev_b = torch.linalg.eigvalsh(bd)
    print(f'[n={n}] stage 1: eigenvalues of the band vs A: {((ev_b - ev_ref).abs().max() / ev_ref.abs().max()).item():.2e}')
    d, e = out['d'].double(), out['e'].double()
    tmat = torch.diag(d) + torch.diag(e[:n - 1], 1) + torch.diag(e[:n - 1], -1)
    ev_t = torch.linalg.eigvalsh(tmat)
    print(f'[n={n}] stage 2: eigenvalues of T vs A:        {((ev_t - ev_ref).abs().max() / ev_ref.abs().max()).item():.2e}')
    if args.q:
        w, tau1 = out['W'].double(), out['tau1'].double()
        q1 = torch.eye(n, dtype=torch.float64, device='cuda')
        for k in range(n - 1, -1, -1):
            if tau1[k] != 0:
                v = w[k]
                q1 -= tau1[k] * torch.outer(v, v @ q1)
        print(f'[n={n}] stage 1: |Q1^T A Q1 - B| / |A| = {rel(q1.t() @ fd @ q1, bd):.2e} (relative to B), '
              f'orthogonality {(q1.t() @ q1 - torch.eye(n, dtype=torch.float64, device="cuda")).abs().max().item():.2e}')
    if args.model:
        sys.path.insert(0, os.path.join(ROOT, 'oracle'))
        import twostage_model as tm
        a = f.cpu().numpy()
        w_m, tau1_m, ab_m = tm.sy2sb(a)
        d_m, e_m, v2_m, tau2_m, goff = tm.sb2st(ab_m)
        cmp = [('W', np.triu(out['W'].cpu().numpy(), 1), np.triu(w_m, 1)), ('tau1', out['tau1'].cpu().numpy(), tau1_m),
               ('band', out['band'].cpu().numpy()[:, :33], ab_m[:, :33]), ('d', out['d'].cpu().numpy(), d_m),
               ('e', out['e'].cpu().numpy()[:n - 1], e_m[:n - 1])]
        if goff[-1] > 0:
            cmp += [('tau2', out['tau2'].cpu().numpy(), tau2_m), ('V2', out['V2'].cpu().numpy(), v2_m)]
        for name, got, want in cmp:
            err = np.abs(got.astype(np.float64) - want.astype(np.float64)).max() / max(np.abs(want).max(), 1e-300)
            print(f'[n={n}] model: {name:5s} max abs err / max |ref| = {err:.2e}')
    if args.full or args.time:
        for mode, name in ((1, 'one-stage'), (2, 'two-stage')):
            _native.check(lib.kfac_eigh_set_mode(mode))
            try:
                (q,), (dd,) = ops.eigh_batched([f], clamp_nonneg=True)
                torch.cuda.synchronize()
                qd, ddd = q.double(), dd.double()
                rec = rel(qd @ torch.diag(ddd) @ qd.t(), fd)
                orth = (qd.t() @ qd - torch.eye(n, dtype=torch.float64, device='cuda')).abs().max().item()
                ev = ((ddd - ev_ref).abs().max() / ev_ref.abs().max()).item()
                msg = f'[n={n}] {name}: reconstruction {rec:.2e} orthogonality {orth:.2e} eigenvalues {ev:.2e} paths {ops.eigh_last_paths()}'
                if args.time:
                    lib.kfac_profile_enable(1)
                    o4 = (ctypes.c_float * 4)()
                    best = None
                    for _ in range(3):
                        ops.eigh_batched([f], clamp_nonneg=True, qs=[q], ds=[dd])
                        torch.cuda.synchronize()
                        lib.kfac_profile_last_ms(o4)
                        cur = (o4[0], o4[1], o4[2])
                        best = cur if best is None or sum(cur) < sum(best) else best
                    lib.kfac_profile_enable(0)
                    msg += f' | ms: tridiag {best[0]:.2f} d&c {best[1]:.2f} back {best[2]:.2f} total {sum(best):.2f}'
                print(msg, flush=True)
            finally:
                lib.kfac_eigh_set_mode(0)


if __name__ == '__main__':
    main()
