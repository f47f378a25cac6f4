"""Multi-GPU parity check (run under torchrun on N GPUs):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/check_multigpu.py

N data-parallel ranks that each see one micro-batch are mathematically the reference with accumulation_steps = N
(factors averaged over ranks == averaged over micro-batches; DDP-averaged gradients == gradients of the mean loss).
For N = 2 the expected gradients are the UNMODIFIED reference's (tests/golden/e2e.pt, accumulation_steps=2 case);
for other N they come from the CPU oracle run with accumulation_steps = N on rank 0's host.  Every KAISA strategy
(COMM_OPT, HYBRID_OPT, MEM_OPT) x {eigen, inverse} must give the same gradients on every rank.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, 'distributed-kfac-pytorch_b200'), ROOT, os.path.join(ROOT, 'tests')):
    sys.path.insert(0, p)
import torch
import torch.distributed as dist

import models
from kfac_b200 import KFACPreconditioner
from kfac_b200.enums import DistributedStrategy
from oracle import kfac_oracle as orc


def relerr(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return (torch.linalg.norm(a - b) / torch.linalg.norm(b)).item()


def case_kwargs_ok():
    return torch.cuda.device_count() >= 2


def device_guard_check():
    """A model on cuda:1 driven by a thread whose current device is cuda:0 (no torch.cuda.set_device): the
    reference's torch ops follow the tensors' device, so the native wrappers must too (ops._on_device)."""
    case = torch.load(os.path.join(ROOT, 'tests', 'golden', 'e2e.pt'), weights_only=False)[0]
    name = case['model']
    assert torch.cuda.current_device() == 0
    dev = torch.device('cuda', 1)
    net = models.MODELS[name]()
    net.load_state_dict(case['state0'])
    net = net.to(dev)
    pre = KFACPreconditioner(net, **case['kwargs'])
    worst = 0.0
    for step in range(case['steps']):
        net.zero_grad()
        x, y = models.make_batch(name, step)
        models.loss_fn(name, net(x.to(dev)), y.to(dev)).backward()
        pre.step()
        assert torch.cuda.current_device() == 0
        for n, p in net.named_parameters():
            worst = max(worst, relerr(p.grad, case['per_step'][step]['grads'][n]))
        with torch.no_grad():
            for p in net.parameters():
                p -= 0.05 * p.grad
    assert worst < 1e-4, worst
    print(f'ok device guard: model on cuda:1, current device cuda:0, max rel err {worst:.3e} vs reference golden',
          flush=True)


def main():
    rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
    if rank == 0 and world >= 2 and case_kwargs_ok():
        device_guard_check()
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    name, steps = 'convnet', 2
    golden = None
    if world == 2:
        cases = torch.load(os.path.join(ROOT, 'tests', 'golden', 'e2e.pt'), weights_only=False)
        golden = [c for c in cases if c['kwargs'].get('accumulation_steps') == 2][0]
        state0 = golden['state0']
    else:
        torch.manual_seed(21)
        state0 = models.MODELS[name]().state_dict()
    worst = 0.0
    for method in ('eigen', 'inverse'):
        # expected gradients
        if golden is not None and method == 'eigen':
            expected = [s['grads'] for s in golden['per_step']]
        else:
            net = models.MODELS[name]()
            net.load_state_dict(state0)
            ref = orc.OracleKFAC(net, accumulation_steps=world, compute_method=method)
            expected = []
            for step in range(steps):
                net.zero_grad()
                for r in range(world):
                    x, y = models.make_batch(name, step * world + r)
                    (models.loss_fn(name, net(x), y) / world).backward()
                ref.step()
                expected.append({n: p.grad.clone() for n, p in net.named_parameters()})
                with torch.no_grad():
                    for p in net.parameters():
                        p -= 0.05 * p.grad
        for strategy in (DistributedStrategy.COMM_OPT, DistributedStrategy.HYBRID_OPT, DistributedStrategy.MEM_OPT):
            # (symmetry_aware, wrapped in DistributedDataParallel): the DDP case is how the reference's examples run
            # (examples/torch_cifar10_resnet.py wraps the model first, then builds the preconditioner on it)
            for sym, ddp in ((False, False), (True, False), (False, True)):
                net = models.MODELS[name]()
                net.load_state_dict(state0)
                net = net.cuda()
                if ddp:
                    net = torch.nn.parallel.DistributedDataParallel(net, device_ids=[local])
                pre = KFACPreconditioner(net, grad_worker_fraction=strategy, compute_method=method,
                                         symmetry_aware=sym, allreduce_bucket_cap_mb=25.0 if not sym else 0.0,
                                         loglevel=0)
                for step in range(steps):
                    net.zero_grad()
                    x, y = models.make_batch(name, step * world + rank)
                    # loss / world + SUM-reduce == DDP's mean gradient, and makes every rank's grad-outputs (hence
                    # the G factors) identical to the reference's micro-batches, which back-propagate loss / N
                    (models.loss_fn(name, net(x.cuda()), y.cuda()) / world).backward()
                    if ddp:   # DDP's reducer averaged the gradients during backward: mean x N == the sum below
                        for p in net.parameters():
                            p.grad.mul_(world)
                    else:
                        for p in net.parameters():
                            dist.all_reduce(p.grad)
                    pre.step()
                    for n, p in net.named_parameters():
                        n = n[len('module.'):] if ddp else n
                        err = relerr(p.grad, expected[step][n])
                        worst = max(worst, err)
                        if err > 1e-4:
                            raise AssertionError(f'rank {rank} {method} {strategy} sym={sym} ddp={ddp} step {step} {n}: {err}')
                    with torch.no_grad():
                        for p in net.parameters():
                            p -= 0.05 * p.grad
                # every rank ends with the same gradients
                flat = torch.cat([p.grad.flatten() for p in net.parameters()])
                ref0 = flat.clone()
                dist.broadcast(ref0, src=0)
                assert relerr(flat, ref0) < 1e-6
                if rank == 0:
                    print(f'ok world={world} {method} {strategy.name} symmetry_aware={sym} ddp={ddp} '
                          f'workers={pre._assignment.grad_workers} allreduce_calls={pre._tdc.stats["allreduce_calls"]} '
                          f'broadcast_calls={pre._tdc.stats["broadcast_calls"]}', flush=True)
    t = torch.tensor([worst], device='cuda')
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        src = 'reference golden (accumulation_steps=2)' if golden is not None else 'oracle'
        print(f'MULTIGPU PARITY OK world={world} max rel err {t.item():.3e} vs {src}', flush=True)
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
