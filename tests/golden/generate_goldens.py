"""Generate the committed golden fixtures by running the UNMODIFIED reference (kfac_pytorch 0.4.1).

Run in the build container only (the reference tree does not exist on the GPU box):

    PYTHONPATH=/root/reference python tests/golden/generate_goldens.py

Outputs (all small, committed):
  tests/golden/kat_get_cov.pt        the reference's own known-answer vectors (tests/layers/utils_test.py:24-56)
  tests/golden/assignment.json       the reference's own placement goldens (tests/assignment_test.py:60-141,
                                     193-428) + KAISAAssignment tables computed by the reference class
  tests/golden/factors.pt            Linear / Conv2d helper factors on seeded inputs (kfac/layers/modules.py)
  tests/golden/layers.pt             per-layer pipeline: factors -> eigen / inverse -> preconditioned gradient
  tests/golden/e2e.pt                KFACPreconditioner end-to-end on tests/models.py nets, several cadences
  tests/golden/triu.pt               get_triu / fill_triu (kfac/distributed.py:416-459)
"""
from __future__ import annotations

import json
import os
import sys
import warnings

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))           # tests/ (for models.py)
sys.path.insert(0, '/root/reference')
sys.path.insert(0, '/root/reference/tests')

warnings.filterwarnings('ignore')

import kfac  # noqa: E402
from kfac.assignment import KAISAAssignment  # noqa: E402
from kfac.distributed import fill_triu, get_triu, TorchDistributedCommunicator  # noqa: E402
from kfac.layers.eigen import KFACEigenLayer  # noqa: E402
from kfac.layers.inverse import KFACInverseLayer  # noqa: E402
from kfac.layers.modules import Conv2dModuleHelper, LinearModuleHelper  # noqa: E402
from kfac.layers.utils import get_cov  # noqa: E402
from kfac.preconditioner import KFACPreconditioner  # noqa: E402

import models  # noqa: E402

assert kfac.__version__ == '0.4.1'
torch.set_num_threads(4)


def gen_kat_get_cov() -> None:
    import importlib.util
    spec = importlib.util.spec_from_file_location('ref_utils_test', '/root/reference/tests/layers/utils_test.py')
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    cases = []
    for mark in mod.test_get_cov.pytestmark:
        if mark.name == 'parametrize':
            for a, b, scale, expected in mark.args[1]:
                if len(a.shape) != 2 or (b is not None and a.shape != b.shape):
                    continue
                # re-run the reference to make sure the table is what the reference computes
                assert torch.equal(get_cov(a, b, scale), expected)
                cases.append({'a': a, 'b': b, 'scale': scale, 'expected': expected})
    torch.save(cases, os.path.join(HERE, 'kat_get_cov.pt'))
    print('kat_get_cov', len(cases))


def gen_assignment() -> None:
    import importlib.util
    spec = importlib.util.spec_from_file_location('ref_assignment_test', '/root/reference/tests/assignment_test.py')
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)

    def params(fn):
        for mark in fn.pytestmark:
            if mark.name == 'parametrize':
                return mark.args[1]
        raise KeyError

    out = {
        'partition_grad_workers': [list(c) for c in params(mod.test_partition_grad_workers)],
        'partition_grad_receivers': [list(c) for c in params(mod.test_partition_grad_receivers)],
        'greedy_assignment': [list(c) for c in params(mod.test_kaisa_assignment_greedy_assignment)],
        'group_sizes': [list(c) for c in params(mod.test_kaisa_assignment_group_sizes)],
    }
    # Full KAISA tables from the reference class on a ResNet-like work dict (cost = n^3)
    dims = [(147, 64), (64, 64), (64, 256), (256, 64), (576, 64), (256, 128), (1152, 128), (128, 512),
            (2304, 256), (1024, 256), (512, 2048), (4608, 512), (2049, 1000)]
    work = {f'layer{i}': {'A': float(a) ** 3, 'G': float(g) ** 3} for i, (a, g) in enumerate(dims)}
    tables = []
    for world, frac in [(1, 1.0), (2, 0.5), (2, 1.0), (4, 0.25), (4, 0.5), (4, 1.0), (8, 0.125), (8, 0.25), (8, 0.5),
                        (8, 1.0)]:
        for colocate in (True, False):
            per_rank = []
            for rank in range(world):
                asg = KAISAAssignment(work, local_rank=rank, world_size=world, grad_worker_fraction=frac,
                                      group_func=lambda ranks: tuple(sorted(ranks)), colocate_factors=colocate)
                per_rank.append({
                    'inv_worker': {layer: {f: asg.inv_worker(layer, f) for f in asg.get_factors(layer)}
                                   for layer in asg.get_layers()},
                    'is_grad_worker': {layer: asg.is_grad_worker(layer) for layer in asg.get_layers()},
                    'src_grad_worker': {layer: asg.src_grad_worker(layer) for layer in asg.get_layers()},
                    'grad_worker_group': {layer: list(asg.grad_worker_group(layer)) for layer in asg.get_layers()},
                    'grad_receiver_group': {layer: list(asg.grad_receiver_group(layer)) for layer in asg.get_layers()},
                    'broadcast_gradients': asg.broadcast_gradients(),
                    'broadcast_inverses': asg.broadcast_inverses(),
                })
            tables.append({'world': world, 'frac': frac, 'colocate': colocate, 'ranks': per_rank})
    out['kaisa_work'] = work
    out['kaisa_tables'] = tables
    with open(os.path.join(HERE, 'assignment.json'), 'w') as f:
        json.dump(out, f)
    print('assignment', {k: len(v) for k, v in out.items()})


def gen_factors() -> None:
    torch.manual_seed(11)
    cases = []
    for (m_shape, n_in, n_out, bias) in [((5,), 6, 4, True), ((3, 7), 9, 5, False), ((16,), 33, 10, True)]:
        lin = torch.nn.Linear(n_in, n_out, bias=bias)
        h = LinearModuleHelper(lin)
        a = torch.randn(*m_shape, n_in)
        g = torch.randn(*m_shape, n_out)
        cases.append({'kind': 'linear', 'bias': bias, 'a': a, 'g': g, 'A': h.get_a_factor(a.clone()),
                      'G': h.get_g_factor(g.clone())})
    for (nb, c, hh, ww, cout, k, s, p, bias) in [
        (2, 3, 8, 8, 4, (3, 3), (1, 1), (1, 1), True),
        (3, 4, 9, 7, 6, (3, 3), (2, 2), (1, 1), False),
        (2, 2, 12, 12, 5, (5, 5), (2, 2), (2, 2), True),
        (4, 8, 6, 6, 8, (1, 1), (1, 1), (0, 0), False),
        (2, 3, 7, 9, 4, (2, 3), (1, 2), (0, 1), True),
        (2, 3, 16, 16, 8, (7, 7), (2, 2), (3, 3), False),
    ]:
        conv = torch.nn.Conv2d(c, cout, k, stride=s, padding=p, bias=bias)
        h = Conv2dModuleHelper(conv)
        a = torch.relu(torch.randn(nb, c, hh, ww))
        out = conv(a)
        g = torch.randn_like(out) / out[0].numel()
        cases.append({'kind': 'conv2d', 'bias': bias, 'kernel': k, 'stride': s, 'padding': p, 'a': a, 'g': g,
                      'A': h.get_a_factor(a.clone()), 'G': h.get_g_factor(g.clone())})
    torch.save(cases, os.path.join(HERE, 'factors.pt'))
    print('factors', len(cases))


def gen_layers() -> None:
    """Per-layer pipeline of tests/layers/layers_test.py:85-213 with values dumped at every stage."""
    torch.manual_seed(12)
    cases = []
    tdc = TorchDistributedCommunicator()
    specs = [
        ('linear', torch.nn.Linear(10, 6, bias=True), (8, 10)),
        ('linear', torch.nn.Linear(20, 12, bias=False), (4, 3, 20)),
        ('conv2d', torch.nn.Conv2d(3, 5, 3, padding=1, bias=True), (4, 3, 6, 6)),
        ('conv2d', torch.nn.Conv2d(4, 6, 3, stride=2, bias=False), (3, 4, 9, 9)),
    ]
    for kind, module, in_shape in specs:
        helper = LinearModuleHelper(module) if kind == 'linear' else Conv2dModuleHelper(module)
        for method in ('eigen', 'eigen_prediv', 'inverse'):
            for accumulate in (1, 2):
                if method == 'inverse':
                    layer = KFACInverseLayer(helper, tdc=tdc)
                else:
                    layer = KFACEigenLayer(helper, tdc=tdc, prediv_eigenvalues=(method == 'eigen_prediv'))
                record = {'kind': kind, 'method': method, 'accumulate': accumulate,
                          'module_state': {k: v.clone() for k, v in module.state_dict().items()},
                          'module_args': _module_args(module), 'updates': []}
                for update in range(2):  # two running-average updates
                    ins, gouts = [], []
                    for _ in range(accumulate):
                        x = torch.randn(*in_shape)
                        out = module(x)
                        go = torch.randn_like(out) / out.shape[0]
                        layer.save_layer_input([x])
                        layer.save_layer_grad_output((go,))
                        ins.append(x)
                        gouts.append(go)
                    layer.update_a_factor(alpha=0.9)
                    layer.update_g_factor(alpha=0.9)
                    record['updates'].append({'inputs': ins, 'grad_outputs': gouts,
                                              'A': layer.a_factor.clone(), 'G': layer.g_factor.clone()})
                layer.compute_a_inv(damping=0.01)
                layer.compute_g_inv(damping=0.01)
                module.zero_grad()
                wgrad = torch.randn_like(module.weight)
                module.weight.grad = wgrad.clone()
                bgrad = None
                if module.bias is not None:
                    bgrad = torch.randn_like(module.bias)
                    module.bias.grad = bgrad.clone()
                layer.preconditioned_grad(damping=0.01)
                record['w_grad'] = wgrad
                record['b_grad'] = bgrad
                record['precond'] = layer.grad.clone()
                if method == 'inverse':
                    record['a_inv'] = layer.a_inv.clone()
                    record['g_inv'] = layer.g_inv.clone()
                layer.update_grad(scale=0.5)
                record['w_out'] = module.weight.grad.clone()
                record['b_out'] = module.bias.grad.clone() if module.bias is not None else None
                cases.append(record)
    torch.save(cases, os.path.join(HERE, 'layers.pt'))
    print('layers', len(cases))


def _module_args(module):
    if isinstance(module, torch.nn.Linear):
        return {'in': module.in_features, 'out': module.out_features, 'bias': module.bias is not None}
    return {'in': module.in_channels, 'out': module.out_channels, 'kernel': module.kernel_size,
            'stride': module.stride, 'padding': module.padding, 'bias': module.bias is not None}


def gen_e2e() -> None:
    """Reference KFACPreconditioner on tests/models.py nets; dump grads after step() for several configs."""
    cases = []
    configs = [
        dict(model='convnet', kwargs=dict(), steps=3),
        dict(model='convnet', kwargs=dict(compute_eigenvalue_outer_product=False), steps=2),
        dict(model='convnet', kwargs=dict(factor_update_steps=2, inv_update_steps=2), steps=4),
        dict(model='convnet', kwargs=dict(compute_method='inverse'), steps=3),
        dict(model='convnet', kwargs=dict(update_factors_in_hook=False, kl_clip=0.01, lr=0.5), steps=2),
        dict(model='convnet', kwargs=dict(accumulation_steps=2), steps=2),
        dict(model='convnet', kwargs=dict(skip_layers=['conv3'], damping=0.01, factor_decay=0.8), steps=2),
        dict(model='mlp', kwargs=dict(), steps=3),
        dict(model='mlp', kwargs=dict(compute_method='inverse', damping=0.003), steps=2),
    ]
    for cfg in configs:
        torch.manual_seed(21)
        name = cfg['model']
        net = models.MODELS[name]()
        state0 = {k: v.clone() for k, v in net.state_dict().items()}
        pre = KFACPreconditioner(net, **cfg['kwargs'])
        acc = cfg['kwargs'].get('accumulation_steps', 1)
        per_step = []
        mb = 0
        for step in range(cfg['steps']):
            net.zero_grad()
            for micro in range(acc):
                x, y = models.make_batch(name, mb)
                mb += 1
                loss = models.loss_fn(name, net(x), y) / acc
                loss.backward()
            raw = {n: p.grad.clone() for n, p in net.named_parameters()}
            pre.step()
            out = {n: p.grad.clone() for n, p in net.named_parameters()}
            per_step.append({'raw_grads': raw, 'grads': out})
            # plain SGD update so later steps see different weights (kept outside K-FAC)
            with torch.no_grad():
                for p in net.parameters():
                    p -= 0.05 * p.grad
        factors = {lname: {'A': layer.a_factor.clone(), 'G': layer.g_factor.clone()}
                   for lname, layer in pre._layers.values()}
        cases.append({'model': name, 'kwargs': cfg['kwargs'], 'steps': cfg['steps'], 'state0': state0,
                      'per_step': per_step, 'factors': factors,
                      'state_dict_keys': sorted(pre.state_dict().keys())})
    torch.save(cases, os.path.join(HERE, 'e2e.pt'))
    print('e2e', len(cases))


def gen_triu() -> None:
    torch.manual_seed(31)
    cases = []
    for n in (1, 2, 5, 16):
        x = torch.randn(n, n)
        x = x + x.t()
        t = get_triu(x)
        cases.append({'x': x, 'triu': t, 'filled': fill_triu((n, n), t)})
    torch.save(cases, os.path.join(HERE, 'triu.pt'))
    print('triu', len(cases))


if __name__ == '__main__':
    gen_kat_get_cov()
    gen_assignment()
    gen_factors()
    gen_layers()
    gen_e2e()
    gen_triu()
