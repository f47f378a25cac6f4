"""End-to-end parity on the GPU: kfac_b200.KFACPreconditioner vs the gradients the UNMODIFIED reference produced
(tests/golden/e2e.pt) for every configuration, and vs the CPU oracle for the remaining options."""
from __future__ import annotations

import os

import pytest
import torch

import models
from oracle import kfac_oracle as orc

pytestmark = pytest.mark.gpu

GRAD_TOL = 1e-4


def relerr(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return (torch.linalg.norm(a - b) / torch.linalg.norm(b).clamp_min(1e-300)).item()


def _run(case, golden=True):
    from kfac_b200 import KFACPreconditioner
    name = case['model']
    net = models.MODELS[name]()
    net.load_state_dict(case['state0'])
    net = net.cuda()
    pre = KFACPreconditioner(net, **case['kwargs'])
    acc = case['kwargs'].get('accumulation_steps', 1)
    mb = 0
    worst = 0.0
    for step in range(case['steps']):
        net.zero_grad()
        for _ in range(acc):
            x, y = models.make_batch(name, mb)
            mb += 1
            (models.loss_fn(name, net(x.cuda()), y.cuda()) / acc).backward()
        pre.step()
        ref = case['per_step'][step]['grads']
        for n, p in net.named_parameters():
            err = relerr(p.grad, ref[n])
            worst = max(worst, err)
            assert err < GRAD_TOL, (case['kwargs'], step, n, err)
        with torch.no_grad():
            for p in net.parameters():
                p -= 0.05 * p.grad
    for lname, layer in pre._layers.values():
        assert relerr(layer.a_factor, case['factors'][lname]['A']) < 1e-5
        assert relerr(layer.g_factor, case['factors'][lname]['G']) < 1e-5
    return pre, net, worst


def test_matches_reference_on_every_golden_config(golden_dir):
    cases = torch.load(os.path.join(golden_dir, 'e2e.pt'), weights_only=False)
    assert len(cases) == 9
    for case in cases:
        pre, _, _ = _run(case)
        assert sorted(pre.state_dict().keys()) == case['state_dict_keys']
        assert pre.steps == case['steps']


def test_state_dict_roundtrip_recomputes_second_order(golden_dir):
    """state_dict -> load_state_dict(compute_inverses=True) on a fresh preconditioner gives the same next step
    (kfac/base_preconditioner.py:247-306)."""
    from kfac_b200 import KFACPreconditioner
    case = torch.load(os.path.join(golden_dir, 'e2e.pt'), weights_only=False)[0]
    pre, net, _ = _run(case)
    sd = pre.state_dict()
    net2 = models.MODELS['convnet']().cuda()
    net2.load_state_dict(net.state_dict())
    pre2 = KFACPreconditioner(net2, inv_update_steps=100)  # second-order data must come from the load
    pre2.load_state_dict(sd)
    pre._inv_update_steps = 100
    pre2._inv_update_steps = 100   # the loaded state_dict carried the saved value (1) back in
    x, y = models.make_batch('convnet', 99)
    for n_, p_ in ((net, pre), (net2, pre2)):
        n_.zero_grad()
        models.loss_fn('convnet', n_(x.cuda()), y.cuda()).backward()
        p_.step()
    for (n1, p1), (_, p2) in zip(net.named_parameters(), net2.named_parameters()):
        assert relerr(p2.grad, p1.grad) < GRAD_TOL, n1
    mem = pre.memory_usage()
    assert set(mem) >= {'a_factors', 'g_factors', 'a_batch', 'g_batch', 'a_inverses', 'g_inverses', 'total'}
    assert mem['total'] > 0


def test_eval_mode_and_cadence_gates():
    """Hooks are inert in eval mode and on non-factor steps (base_preconditioner.py:442-444, 465-467)."""
    from kfac_b200 import KFACPreconditioner
    torch.manual_seed(0)
    net = models.MODELS['mlp']().cuda()
    pre = KFACPreconditioner(net, factor_update_steps=2, inv_update_steps=2)
    x, y = models.make_batch('mlp', 0)
    net.eval()
    net(x.cuda())
    assert all(layer.a_factor is None for _, layer in pre._layers.values())
    net.train()
    models.loss_fn('mlp', net(x.cuda()), y.cuda()).backward()
    pre.step()
    a0 = [layer.a_factor.clone() for _, layer in pre._layers.values()]
    net.zero_grad()
    models.loss_fn('mlp', net(x.cuda()), y.cuda()).backward()   # step 1: not a factor step
    pre.step()
    for a, (_, layer) in zip(a0, pre._layers.values()):
        assert torch.equal(a, layer.a_factor)


def test_kl_clip_scale_matches_oracle_and_no_clip():
    from kfac_b200 import KFACPreconditioner
    torch.manual_seed(1)
    state0 = models.MODELS['convnet']().state_dict()
    grads = {}
    for impl in ('b200', 'oracle'):
        net = models.MODELS['convnet']()
        net.load_state_dict(state0)
        if impl == 'b200':
            net = net.cuda()
            pre = KFACPreconditioner(net, kl_clip=0.05, lr=0.3, damping=0.02)
        else:
            pre = orc.OracleKFAC(net, kl_clip=0.05, lr=0.3, damping=0.02)
        x, y = models.make_batch('convnet', 3)
        dev = 'cuda' if impl == 'b200' else 'cpu'
        models.loss_fn('convnet', net(x.to(dev)), y.to(dev)).backward()
        pre.step()
        grads[impl] = {n: p.grad.detach().cpu() for n, p in net.named_parameters()}
        if impl == 'b200':
            scale = pre._scale.item()
        else:
            assert abs(scale - pre.last_scale) < 1e-4 * pre.last_scale
    for n in grads['oracle']:
        assert relerr(grads['b200'][n], grads['oracle'][n]) < GRAD_TOL


def test_grad_scaler_unscales_g_factor():
    """AMP: the backward hook sees grad-outputs of the SCALED loss and divides by grad_scaler() before the G factor
    (kfac/layers/base.py:363-365).  A run with loss * 1024 and grad_scaler = 1024 must reproduce the unscaled factors
    and, once the user has unscaled the gradients (scaler.unscale_), the same preconditioned gradients."""
    from kfac_b200 import KFACPreconditioner
    torch.manual_seed(2)
    state0 = models.MODELS['convnet']().state_dict()
    out = {}
    for scale in (1.0, 1024.0):
        net = models.MODELS['convnet']()
        net.load_state_dict(state0)
        net = net.cuda()
        pre = KFACPreconditioner(net, grad_scaler=(lambda s=scale: s) if scale != 1.0 else None)
        x, y = models.make_batch('convnet', 1)
        (models.loss_fn('convnet', net(x.cuda()), y.cuda()) * scale).backward()
        with torch.no_grad():
            for p in net.parameters():
                p.grad /= scale                      # what GradScaler.unscale_ does before preconditioner.step()
        pre.step()
        out[scale] = (
            {n: (l.a_factor.clone(), l.g_factor.clone()) for n, l in pre._layers.values()},
            {n: p.grad.clone() for n, p in net.named_parameters()},
        )
    for n in out[1.0][0]:
        assert relerr(out[1024.0][0][n][0], out[1.0][0][n][0]) < 1e-6
        assert relerr(out[1024.0][0][n][1], out[1.0][0][n][1]) < 1e-5
    for n in out[1.0][1]:
        assert relerr(out[1024.0][1][n], out[1.0][1][n]) < GRAD_TOL


def test_autocast_half_activations_match_fp32_of_the_same_values():
    """Under torch.autocast the hooks receive bf16 activations and grad-outputs.  The factors are computed in fp32
    from exactly those values (the reference computes the covariance in the tensors' own half dtype unless
    factor_dtype is given, kfac/layers/base.py:349,382-383: its result is the rounded one), so they must equal the
    oracle's factors of the up-cast tensors to the fp32 tolerance."""
    from kfac_b200 import KFACPreconditioner
    torch.manual_seed(4)
    net = models.MODELS['mlp']().cuda()
    pre = KFACPreconditioner(net)
    seen = {}

    def grab_in(mod, inp):
        seen.setdefault(mod, {})['x'] = inp[0].detach().clone()

    def grab_out(mod, gin, gout):
        seen.setdefault(mod, {})['g'] = gout[0].detach().clone()

    handles = []
    for mod in pre._layers:
        handles.append(mod.register_forward_pre_hook(grab_in))
        handles.append(mod.register_full_backward_hook(grab_out))
    x, y = models.make_batch('mlp', 2)
    with torch.autocast('cuda', dtype=torch.bfloat16):
        loss = models.loss_fn('mlp', net(x.cuda()), y.cuda())
    loss.backward()
    for h in handles:
        h.remove()
    assert any(v['x'].dtype == torch.bfloat16 for v in seen.values())   # the test exercises half inputs
    for mod, (name, layer) in pre._layers.items():
        xs = seen[mod]['x'].float().cpu().reshape(-1, seen[mod]['x'].shape[-1])
        gs = seen[mod]['g'].float().cpu().reshape(-1, seen[mod]['g'].shape[-1])
        if mod.bias is not None:
            xs = torch.cat([xs, torch.ones(xs.shape[0], 1)], dim=1)
        # first factor step: running average of the identity and the batch covariance (base.py:374-404, decay 0.95)
        a_ref = 0.95 * torch.eye(xs.shape[1], dtype=torch.double) + 0.05 * (xs.double().t() @ xs.double() / xs.shape[0])
        g_ref = 0.95 * torch.eye(gs.shape[1], dtype=torch.double) + 0.05 * (gs.double().t() @ gs.double() / gs.shape[0])
        assert relerr(layer.a_factor, a_ref) < 1e-5, name
        assert relerr(layer.g_factor, g_ref) < 1e-5, name
