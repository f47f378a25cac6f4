"""Profile helper: representative gemm3x launches (big square NT, back-transform update shape, im2col factor)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'distributed-kfac-pytorch_b200'))
import torch
from kfac_b200 import ops
torch.manual_seed(0)
which = sys.argv[1] if len(sys.argv) > 1 else 'all'
if which in ('all', 'square'):
    a = torch.randn(4096, 4096, device='cuda'); b = torch.randn(4096, 4096, device='cuda'); c = torch.empty(4096, 4096, device='cuda')
    for _ in range(2): ops.gemm(a, b, transb=True, out=c)
    for _ in range(2): ops.gemm(a, b, out=c)
if which in ('all', 'update'):
    n = 4608
    w = torch.randn(n, n, device='cuda'); z = torch.randn(n, n, device='cuda'); c2 = torch.randn(128, n, device='cuda')
    vb = w[128:256, 129:]            # 128 x kk, unaligned start like the WY blocks
    zb = z[129:, :]
    for _ in range(2): ops.gemm(vb, c2, transa=True, alpha=-1.0, beta=1.0, out=zb)
if which in ('all', 'conv'):
    x = torch.relu(torch.randn(32, 128, 28, 28, device='cuda')); f = torch.empty(1152, 1152, device='cuda')
    for _ in range(2): ops.factor_conv2d_a(x, (3, 3), (1, 1), (1, 1), False, f)
torch.cuda.synchronize()
print('done')
