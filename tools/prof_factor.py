"""Per-layer timing of the ResNet-50 factor stage (A and G kernels), grouped by layer shape."""
import os, sys, json, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'distributed-kfac-pytorch_b200')); sys.path.insert(0, ROOT)
import torch
import bench
from kfac_b200 import ops

def timed(fn, iters=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / iters

layers = bench.resnet50_layers()
seen = collections.OrderedDict()
for L in layers:
    key = (L['kind'], L['cin'], L['cout'], L.get('k'), L.get('s'), L.get('h'))
    seen.setdefault(key, [L, 0]); seen[key][1] += 1
tot_a = tot_g = 0.0
for key, (L, cnt) in seen.items():
    d = bench.synth(L, 0, torch.device('cuda'))
    fa = torch.empty(L['a'], L['a'], device='cuda'); fg = torch.empty(L['g'], L['g'], device='cuda')
    if L['kind'] == 'conv':
        ta = timed(lambda: ops.factor_conv2d_a(d['x'], (L['k'],)*2, (L['s'],)*2, (L['p'],)*2, False, fa, alpha=0.05, beta=0.95))
        tg = timed(lambda: ops.factor_conv2d_g(d['go'], fg, alpha=0.05, beta=0.95))
    else:
        ta = timed(lambda: ops.factor_linear(d['x'], True, fa, alpha=0.05, beta=0.95))
        tg = timed(lambda: ops.factor_linear(d['go'], False, fg, alpha=0.05, beta=0.95))
    fl_a = L['M'] * L['a'] * (L['a'] + 1); fl_g = L['M'] * L['g'] * (L['g'] + 1)
    tot_a += cnt * ta; tot_g += cnt * tg
    print(json.dumps(dict(name=L['name'], count=cnt, a=L['a'], g=L['g'], M=L['M'], k=L.get('k'), s=L.get('s'), a_ms=round(ta, 3), g_ms=round(tg, 3),
                          a_tflops=round(fl_a / ta / 1e9, 1), g_tflops=round(fl_g / tg / 1e9, 1), a_total_ms=round(cnt * ta, 2), g_total_ms=round(cnt * tg, 2))), flush=True)
print(json.dumps(dict(total_a_ms=tot_a, total_g_ms=tot_g)))
