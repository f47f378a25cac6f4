"""Profile helper: one batched eigendecomposition of the 108 ResNet-50 factor sizes (or a subset)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'distributed-kfac-pytorch_b200'))
sys.path.insert(0, ROOT)
import torch
import bench
from kfac_b200 import ops

which = sys.argv[1] if len(sys.argv) > 1 else 'all'
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
layers = bench.resnet50_layers()
ns = [L['a'] for L in layers] + [L['g'] for L in layers]
if which == 'big':
    ns = [4608]
elif which == 'mid':
    ns = [1024, 1152, 2304]
fs = []
for i, n in enumerate(ns):
    g = torch.Generator(device='cuda').manual_seed(i)
    x = torch.relu(torch.randn(max(64, min(2 * n, 4096)), n, device='cuda', generator=g))
    fs.append((0.95 * torch.eye(n, device='cuda') + 0.05 * x.t() @ x / x.shape[0]).contiguous())
qs = [torch.empty_like(f) for f in fs]
ds = [torch.empty(f.shape[0], device='cuda') for f in fs]
torch.cuda.synchronize()
for r in range(reps):
    t0 = time.perf_counter()
    ops.eigh_batched(fs, qs=qs, ds=ds)
    torch.cuda.synchronize()
    print(f'eigh_batched[{which}] {len(ns)} matrices: {1e3 * (time.perf_counter() - t0):.1f} ms', flush=True)
