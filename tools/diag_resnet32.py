"""Builder tool: per-layer comparison of one K-FAC step on CIFAR ResNet-32 (GPU engine vs CPU oracle)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, 'distributed-kfac-pytorch_b200'), ROOT, os.path.join(ROOT, 'tests')):
    sys.path.insert(0, p)
import torch

import models
from kfac_b200 import KFACPreconditioner
from oracle import kfac_oracle as orc


def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return (torch.linalg.norm(a - b) / max(torch.linalg.norm(b).item(), 1e-300)).item()


torch.manual_seed(0)
state0 = models.MODELS['resnet32']().state_dict()
res = {}
for impl in ('b200', 'oracle'):
    net = models.MODELS['resnet32']()
    net.load_state_dict(state0)
    dev = 'cuda' if impl == 'b200' else 'cpu'
    net = net.to(dev).train()
    pre = KFACPreconditioner(net, loglevel=0) if impl == 'b200' else orc.OracleKFAC(net)
    x, y = models.make_batch('resnet32', 0)
    models.loss_fn('resnet32', net(x.to(dev)), y.to(dev)).backward()
    raw = {n: p.grad.detach().clone() for n, p in net.named_parameters()}
    pre.step()
    out = {n: p.grad.detach().clone() for n, p in net.named_parameters()}
    if impl == 'b200':
        fac = {name: (layer.a_factor.clone(), layer.g_factor.clone()) for name, layer in pre._layers.values()}
        scale = float(pre._scale.item())
    else:
        fac = {st['name']: (st['A'], st['G']) for st in pre.layers.values()}
        scale = pre.last_scale
    res[impl] = (raw, out, fac, scale)
print('scale', res['b200'][3], res['oracle'][3])
for n in res['oracle'][0]:
    print(f'{n:32s} raw {rel(res["b200"][0][n], res["oracle"][0][n]):.2e}  out {rel(res["b200"][1][n], res["oracle"][1][n]):.2e}'
          f'  out/scale {rel(res["b200"][1][n] / res["b200"][3], res["oracle"][1][n] / res["oracle"][3]):.2e}')
for name in res['oracle'][2]:
    a0, g0 = res['oracle'][2][name]
    a1, g1 = res['b200'][2][name]
    da = (a1.cpu().double() - a0.double()); da.fill_diagonal_(0); oa = a0.double().clone(); oa.fill_diagonal_(0)
    dg = (g1.cpu().double() - g0.double()); dg.fill_diagonal_(0); og = g0.double().clone(); og.fill_diagonal_(0)
    print(f'{name:24s} A {tuple(a0.shape)} rel {rel(a1, a0):.2e} offdiag {(da.norm() / oa.norm()).item():.2e}   '
          f'G {tuple(g0.shape)} rel {rel(g1, g0):.2e} offdiag {(dg.norm() / og.norm()).item():.2e}')
