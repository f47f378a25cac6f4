"""BASELINE.json configs 3-5 on one GPU: parity against the fp64 arbiter / torch and stage timings.

  config 3  BERT-large linear layers (A in {1025, 4097}, G in {1024, 4096}, M = 8 x 512 tokens), eigen
  config 4  ResNet-50 shapes, inverse method (Newton-Schulz)
  config 5  GPT-NeoX-style eigensolver sweep, n in {512 ... 8193}: F = 0.95 I + 0.05 X^T X / M, X = relu(N(0,1)), M = 4n
Writes one JSON line per measurement (collected into profiles/ by hand).
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'distributed-kfac-pytorch_b200'))
sys.path.insert(0, ROOT)
import torch

import bench
from kfac_b200 import KFACPreconditioner, ops
from oracle import kfac_oracle as orc


def emit(**kw):
    print(json.dumps(kw), flush=True)


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return (torch.linalg.norm(a - b) / torch.linalg.norm(b)).item()


def timed(fn, iters=3):
    fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def config3():
    """24 x {4 x (1024->1024), 1024->4096, 4096->1024}; only one encoder layer's six Linear are instantiated and
    timed (the 24 layers are identical shapes); activations N(0,1), b = 8, seq 512 -> M = 4096."""
    torch.manual_seed(0)
    dims = [(1024, 1024)] * 4 + [(1024, 4096), (4096, 1024)]
    model = torch.nn.ModuleList([torch.nn.Linear(i, o) for i, o in dims]).cuda().train()
    pre = KFACPreconditioner(model, loglevel=0)
    xs = [torch.randn(8, 512, i, device='cuda') for i, _ in dims]
    gos = [torch.randn(8, 512, o, device='cuda') / 4096 for _, o in dims]

    def step():
        for m, x in zip(model, xs):
            pre._save_input(m, (x,))
        for m, g in zip(reversed(list(model)), reversed(gos)):
            pre._save_grad_output(m, (None,), (g,))
        for m in model:
            m.weight.grad = torch.randn_like(m.weight)
            m.bias.grad = torch.randn_like(m.bias)
        pre.step()

    ms = timed(step, 2)
    # parity of one layer against the fp64 arbiter built from the same factors
    name, layer = list(pre._layers.values())[4]
    m = model[4]
    w, b = torch.randn_like(m.weight), torch.randn_like(m.bias)
    m.weight.grad, m.bias.grad = w.clone(), b.clone()
    layer.preconditioned_grad(damping=0.001)
    ref = orc.precondition_eigen_fp64(orc.stacked_grad(w.cpu(), b.cpu()), layer.a_factor.cpu(), layer.g_factor.cpu(), 0.001)
    x = xs[4].reshape(-1, 1024)
    a_ref = orc.running_average(None, orc.linear_a_factor(x.double().cpu(), True), 1, 0.95)
    emit(config=3, what='BERT-large encoder layer (6 linears), full K-FAC step', ms_per_encoder_layer=ms, ms_x24=24 * ms,
         precond_rel_vs_fp64=rel(layer.grad, ref), note='factor parity checked on first update only in tests')


def config4():
    layers = bench.resnet50_layers()
    model = bench.make_modules(layers).cuda().train()
    pre = KFACPreconditioner(model, compute_method='inverse', loglevel=0)
    mods = [model[L['name'].replace('.', '_')] for L in layers]
    data = [bench.synth(L, 0, torch.device('cuda')) for L in layers]

    def step():
        for m, d in zip(mods, data):
            pre._save_input(m, (d['x'],))
        for m, d in zip(reversed(mods), reversed(data)):
            pre._save_grad_output(m, (None,), (d['go'],))
        for m, d in zip(mods, data):
            m.weight.grad = d['wg'].clone()
            if d['bg'] is not None:
                m.bias.grad = d['bg'].clone()
        pre.step()

    ms = timed(step, 2)
    iters = [l.ns_iterations for _, l in pre._layers.values()]
    name, layer = list(pre._layers.values())[10]
    a = layer.a_factor
    n = a.shape[0]
    inv_ref = torch.linalg.inv(a.double() + 0.001 * torch.eye(n, device='cuda', dtype=torch.double))
    resid = (layer.a_inv.double() @ (a.double() + 0.001 * torch.eye(n, device='cuda', dtype=torch.double))
             - torch.eye(n, device='cuda', dtype=torch.double)).norm().item() / n ** 0.5
    emit(config=4, what='ResNet-50 inverse method (Newton-Schulz), factor+inverse step', ms=ms,
         ns_iterations_max=max(max(i.values()) for i in iters), a_inv_rel_vs_fp64=rel(layer.a_inv, inv_ref),
         residual_frob_over_sqrt_n=resid)


def config5():
    torch.manual_seed(0)
    for n in (512, 1024, 2048, 2049, 4096, 6144, 8192, 8193):
        x = torch.relu(torch.randn(4 * n, n, device='cuda'))
        f = (0.95 * torch.eye(n, device='cuda') + 0.05 * (x.t() @ x) / (4 * n)).contiguous()
        del x
        q = torch.empty_like(f)
        d = torch.empty(n, device='cuda')
        ms = timed(lambda: ops.eigh_batched([f], qs=[q], ds=[d]), 1)
        t0 = time.perf_counter()
        torch.linalg.eigh(f)
        torch.cuda.synchronize()
        ms_torch = 1e3 * (time.perf_counter() - t0)
        qd = q.double()
        rec = ((qd * d.double()) @ qd.t() - f.double()).norm().item() / f.double().norm().item()
        orth = (qd.t() @ qd - torch.eye(n, device='cuda', dtype=torch.double)).abs().max().item()
        emit(config=5, n=n, ms=ms, cusolver_ms=ms_torch, declared_tflops=14 / 3 * n ** 3 / ms / 1e9, rec=rec, orth_max=orth)
        del q, qd
        torch.cuda.empty_cache()


if __name__ == '__main__':
    for fn in (config3, config4, config5):
        try:
            fn()
        except Exception as ex:  # keep going: one config must not hide the others
            emit(error=f'{fn.__name__}: {type(ex).__name__}: {ex}')
