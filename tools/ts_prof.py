"""Builder tool: one two-stage eigendecomposition of a single n x n matrix, for `ncu --metrics gpu__time_duration.sum`
launch lists (per-kernel breakdown of the two-stage path)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, 'distributed-kfac-pytorch_b200'), ROOT):
    sys.path.insert(0, p)
import torch

from kfac_b200 import _native
from kfac_b200 import ops

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4608
mode = int(sys.argv[2]) if len(sys.argv) > 2 else 2
g = torch.Generator(device='cuda').manual_seed(n)
x = torch.relu(torch.randn(4 * n, n, device='cuda', generator=g))
f = (0.95 * torch.eye(n, device='cuda') + 0.05 * (x.t() @ x) / (4 * n)).contiguous()
lib = _native.load()
_native.check(lib.kfac_eigh_set_mode(mode))
(q,), (d,) = ops.eigh_batched([f])
torch.cuda.synchronize()
print('done', d[-1].item())
