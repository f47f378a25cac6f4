"""world_size-2 gloo tests (CPU) of the multi-rank host logic: communicator collectives, zero-copy coalescing of
adjacent tensors, Pending handles, and the KAISA collective order of a step (no compute: kernels need a GPU)."""
from __future__ import annotations

import os
import socket
import traceback

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, fn_name: str, errq) -> None:
    try:
        os.environ['MASTER_ADDR'] = '127.0.0.1'
        os.environ['MASTER_PORT'] = str(port)
        dist.init_process_group('gloo', rank=rank, world_size=world)
        globals()[fn_name](rank, world)
        dist.barrier()
    except Exception:  # pragma: no cover
        errq.put((rank, traceback.format_exc()))
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


def _run(fn_name: str, world: int = 2) -> None:
    ctx = mp.get_context('spawn')
    errq = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, fn_name, errq)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        if p.is_alive():
            p.terminate()
            pytest.fail(f'{fn_name}: rank hung')
    if not errq.empty():
        rank, tb = errq.get()
        pytest.fail(f'{fn_name} failed on rank {rank}:\n{tb}')
    assert all(p.exitcode == 0 for p in procs)


# ---------------------------------------------------------------- bodies (run inside the spawned ranks)
def _body_collectives(rank: int, world: int) -> None:
    from kfac_b200.distributed import NonSquareTensorError, Pending, TorchDistributedCommunicator
    tdc = TorchDistributedCommunicator()
    t = torch.full((4, 4), float(rank + 1))
    out = tdc.allreduce(t, average=True)
    assert isinstance(out, Pending)
    assert torch.equal(out.wait(), torch.full((4, 4), 1.5)) and out.wait() is t
    t = torch.full((3,), float(rank))
    assert torch.equal(tdc.allreduce(t).wait(), torch.full((3,), 1.0))
    b = torch.full((5,), float(rank + 7))
    assert torch.equal(tdc.broadcast(b, src=1).wait(), torch.full((5,), 8.0))
    with pytest.raises(NonSquareTensorError):
        tdc.allreduce(torch.ones(2, 3), symmetric=True)
    with pytest.raises(NonSquareTensorError):
        tdc.broadcast(torch.ones(3), src=0, symmetric=True)
    # sub-group of one rank: no communication, tensor returned as is (distributed.py:215-216)
    g = dist.new_group([rank]) if False else None
    assert g is None


def _body_bucketed(rank: int, world: int) -> None:
    from kfac_b200.distributed import TorchDistributedCommunicator
    tdc = TorchDistributedCommunicator(bucket_cap_mb=1e-4)  # 100 bytes
    arena = torch.arange(40, dtype=torch.float32) * (rank + 1)
    a, b, c = arena[0:9].view(3, 3), arena[9:18].view(3, 3), arena[18:34].view(4, 4)
    lone = torch.ones(6) * (rank + 1)
    ha = tdc.allreduce_bucketed(a, average=True)
    hb = tdc.allreduce_bucketed(b, average=True)     # adjacent: coalesced with a (72 bytes <= cap)
    hc = tdc.allreduce_bucketed(c, average=True)     # would exceed the cap -> a+b launched, c starts a new run
    hl = tdc.allreduce_bucketed(lone, average=True)  # not adjacent -> c launched alone
    assert tdc.stats['allreduce_calls'] == 2
    tdc.flush_allreduce_buckets()
    assert tdc.stats['allreduce_calls'] == 3
    expect = torch.arange(40, dtype=torch.float32) * 1.5
    assert torch.equal(ha.wait(), expect[0:9].view(3, 3)) and ha.wait().data_ptr() == a.data_ptr()
    assert torch.equal(hb.wait(), expect[9:18].view(3, 3))
    assert torch.equal(hc.wait(), expect[18:34].view(4, 4))
    assert torch.equal(hl.wait(), torch.ones(6) * 1.5)
    assert torch.equal(arena[34:], torch.arange(34, 40, dtype=torch.float32) * (rank + 1))  # untouched
    # waiting on a handle of an open run launches it (no deadlock when flush is forgotten)
    x = torch.ones(2, 2) * rank
    assert torch.equal(tdc.allreduce_bucketed(x, average=False).wait(), torch.ones(2, 2))
    # one tensor larger than the cap travels alone (distributed.py:318-321)
    big = torch.ones(100) * (rank + 1)
    hb2 = tdc.allreduce_bucketed(big, average=True)
    tdc.flush_allreduce_buckets()
    assert torch.equal(hb2.wait(), torch.ones(100) * 1.5)


def _body_kaisa_order(rank: int, world: int) -> None:
    """Both strategies available at world 2: the sequence of (collective, src) calls derived from the assignment
    is identical on the two ranks, gradients end up identical, MEM_OPT sends gradients and COMM_OPT eigenbases."""
    from kfac_b200.assignment import KAISAAssignment
    from kfac_b200.distributed import TorchDistributedCommunicator
    work = {f'l{i}': {'A': float(n) ** 3, 'G': float(m) ** 3} for i, (n, m) in enumerate([(8, 4), (6, 6), (3, 9)])}
    for frac in (1.0, 0.5):
        tdc = TorchDistributedCommunicator()
        asg = KAISAAssignment(work, local_rank=rank, world_size=world, grad_worker_fraction=frac,
                              group_func=dist.new_group)
        log = []
        grads = {}
        for layer in reversed(asg.get_layers()):
            # second-order data: inverse worker -> other gradient workers
            payload = torch.full((4,), float(asg.inv_worker(layer, 'A') + 1) if rank == asg.inv_worker(layer, 'A') else 0.0)
            if asg.broadcast_inverses() and asg.is_grad_worker(layer):
                payload = tdc.broadcast(payload, src=asg.inv_worker(layer, 'A'), group=asg.grad_worker_group(layer)).wait()
                log.append(('inv', layer, asg.inv_worker(layer, 'A')))
            g = torch.full((2,), 0.0)
            if asg.is_grad_worker(layer):
                assert payload[0].item() == asg.inv_worker(layer, 'A') + 1   # has the eigenbasis it needs
                g = torch.full((2,), 42.0 + len(layer))
            if asg.broadcast_gradients():
                g = tdc.broadcast(g, src=asg.src_grad_worker(layer), group=asg.grad_receiver_group(layer)).wait()
                log.append(('grad', layer, asg.src_grad_worker(layer)))
            grads[layer] = g
        assert all(torch.equal(g, torch.full((2,), 42.0 + len(k))) for k, g in grads.items())
        gathered = [None] * world
        dist.all_gather_object(gathered, log)
        if frac == 1.0:
            assert all(kind == 'inv' for kind, _, _ in log) and len(log) == 3
        else:
            assert all(kind == 'grad' for kind, _, _ in log) and len(log) == 3
        assert gathered[0] == gathered[1]


def _body_preconditioner_construct(rank: int, world: int) -> None:
    """KFACPreconditioner under an initialised process group: strategy resolution at world 2 and group creation."""
    import warnings
    from kfac_b200 import KFACPreconditioner
    from kfac_b200.enums import DistributedStrategy
    net = torch.nn.Sequential(torch.nn.Linear(5, 4), torch.nn.Linear(4, 3))
    p = KFACPreconditioner(net, grad_worker_fraction=DistributedStrategy.HYBRID_OPT)
    assert p.grad_worker_fraction == 0.5 and p._assignment.grad_workers == 1   # HYBRID at world 2 == MEM_OPT
    assert p._assignment.broadcast_gradients() and not p._assignment.broadcast_inverses()
    p = KFACPreconditioner(net, grad_worker_fraction=DistributedStrategy.MEM_OPT)
    assert p.grad_worker_fraction == 0.5
    p = KFACPreconditioner(net, grad_worker_fraction=1)
    assert p.distributed_strategy == DistributedStrategy.COMM_OPT and p._assignment.broadcast_inverses()
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter('always')
        p = KFACPreconditioner(net, grad_worker_fraction=0.5, colocate_factors=False,
                               compute_eigenvalue_outer_product=False)
        assert p.colocate_factors and any('MEM_OPT' in str(x.message) for x in w)
    # both ranks computed the same placement
    table = {n: p._assignment.inv_worker(n, 'A') for n in p._assignment.get_layers()}
    gathered = [None] * world
    dist.all_gather_object(gathered, table)
    assert gathered[0] == gathered[1]


# ---------------------------------------------------------------- pytest entry points
def test_collectives_world2():
    _run('_body_collectives')


def test_bucketed_coalescing_world2():
    _run('_body_bucketed')


def test_kaisa_collective_order_world2():
    _run('_body_kaisa_order')


def test_preconditioner_construction_world2():
    _run('_body_preconditioner_construct')


def test_world1_returns_tensors_untouched():
    from kfac_b200.distributed import TorchDistributedCommunicator, get_rank, get_world_size
    tdc = TorchDistributedCommunicator()
    t = torch.ones(3, 3)
    assert tdc.allreduce(t, average=True) is t
    assert tdc.allreduce_bucketed(t, average=True, symmetric=True) is t
    assert tdc.broadcast(t, src=0) is t
    tdc.flush_allreduce_buckets()
    assert get_rank() == 0 and get_world_size() == 1
    assert tdc.bucket_cap_bytes == 25_000_000


def _body_grouped_exchange_proxy(rank: int, world: int) -> None:
    """Body of tests/test_exchange_cpu.py::test_grouped_exchange_moves_the_data_world2 (the spawn helper looks
    bodies up in this module)."""
    import test_exchange_cpu
    test_exchange_cpu._body_grouped_exchange(rank, world)
