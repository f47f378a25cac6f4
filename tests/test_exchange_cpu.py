"""Grouped KAISA exchanges (kfac_b200/exchange.py): the plan is a pure function of the assignment, so it is checked on
the CPU for every strategy at world 8 -- all members of a group must derive the same arenas in the same order, every
layer must be covered exactly once, and a step must need at most `group size` broadcasts per group instead of the
reference's one per tensor (kfac/layers/eigen.py:177-292, kfac/layers/base.py:223-251).  A world-2 gloo run moves real
data through GroupedExchange with stand-in layers (the compute kernels need a GPU; the exchange does not)."""
from __future__ import annotations

import pytest
import torch

from test_distributed_cpu import _run

WORLD = 8


def _layers():
    """(a, g) of a ResNet-like stack: eigen layers with the eigenvalue outer product precomputed."""
    dims = [(147, 64), (64, 64), (576, 64), (256, 128), (1152, 128), (512, 256), (2304, 256), (1024, 512),
            (4608, 512), (2049, 1000)]
    return {f'layer{i}': ag for i, ag in enumerate(dims)}


def _plans(fraction):
    from kfac_b200.assignment import KAISAAssignment
    from kfac_b200.exchange import plan_gradients, plan_second_order
    layers = _layers()
    work = {n: {'A': float(a) ** 3, 'G': float(g) ** 3} for n, (a, g) in layers.items()}
    payloads = {n: {'A': [('qa', (a, a))], 'G': [('qg', (g, g)), ('dgda', (g, a))]} for n, (a, g) in layers.items()}
    shapes = {n: (g, a) for n, (a, g) in layers.items()}
    out = []
    for rank in range(WORLD):
        asg = KAISAAssignment(work, local_rank=rank, world_size=WORLD, grad_worker_fraction=fraction,
                              group_func=lambda ranks: tuple(ranks))
        out.append((asg, plan_second_order(payloads, asg, rank), plan_gradients(shapes, asg, rank)))
    return layers, out


def _signature(arena):
    return (arena.ranks, arena.src, arena.numel, [(e.layer, e.what, e.offset, tuple(e.buffers)) for e in arena.entries])


@pytest.mark.parametrize('fraction', [1.0, 0.5, 0.25, 0.125])
def test_plans_agree_inside_every_group_and_cover_every_layer(fraction):
    layers, plans = _plans(fraction)
    by_key = {}
    for rank, (asg, second, grads) in enumerate(plans):
        for arena in second + grads:
            assert rank in arena.ranks and arena.src in arena.ranks
            by_key.setdefault((arena.entries[0].what == 'grad', arena.key), []).append(_signature(arena))
        # per rank: every layer it is a gradient worker for appears once per factor; every layer once as a gradient
        if asg.broadcast_inverses():
            seen = sorted((e.layer, e.what) for a in second for e in a.entries)
            want = sorted((n, f) for n in layers if asg.is_grad_worker(n) for f in ('A', 'G'))
            assert seen == want
        else:
            assert second == []
        if asg.broadcast_gradients():
            assert sorted(e.layer for a in grads for e in a.entries) == sorted(layers)
        else:
            assert grads == []
        # the arenas are walked in sorted key order on every rank: identical collective order inside a group
        assert [a.key for a in second] == sorted(a.key for a in second)
        assert [a.key for a in grads] == sorted(a.key for a in grads)
        # one broadcast per (group, source): never more than the group has members
        for arenas in (second, grads):
            per_group = {}
            for a in arenas:
                per_group[a.ranks] = per_group.get(a.ranks, 0) + 1
            assert all(cnt <= len(ranks) for ranks, cnt in per_group.items())
    # every member of a group derived the same arena (same entries, same offsets)
    for key, sigs in by_key.items():
        assert len(sigs) == len(key[1][0]), key      # one signature per member rank of the group
        assert all(s == sigs[0] for s in sigs), key
    # against the reference's pattern: 3 collectives per layer and step for this rank's layers
    asg, second, grads = plans[0]
    per_tensor = 3 * sum(asg.is_grad_worker(n) for n in layers) * asg.broadcast_inverses() + \
        len(layers) * asg.broadcast_gradients()
    grouped = len(second) + len(grads)
    assert grouped <= max(per_tensor, 0) and (per_tensor == 0 or grouped <= 8 + 8)


class _FakeLayer:
    """Stands in for a K-FAC layer on the CPU: just the buffers the exchange touches."""

    def __init__(self, a, g):
        self.a, self.g = a, g
        self.bufs = {}
        self.grad_store = None
        self.adopted = []

    def second_order_payload(self):
        return {'A': [('qa', (self.a, self.a))], 'G': [('qg', (self.g, self.g)), ('dgda', (self.g, self.a))]}

    def _grad_shape(self):
        return (self.g, self.a)

    def attach_second_order_buffers(self, views):
        self.bufs.update(views)

    def attach_grad_buffer(self, view):
        self.grad_store = view

    def adopt_second_order(self, which):
        self.adopted.append(which)

    def adopt_grad(self):
        self.adopted.append('grad')


def _body_grouped_exchange(rank: int, world: int) -> None:
    import torch.distributed as dist
    from kfac_b200.assignment import KAISAAssignment
    from kfac_b200.distributed import TorchDistributedCommunicator
    from kfac_b200.exchange import GroupedExchange
    dims = {'l0': (5, 3), 'l1': (7, 4), 'l2': (3, 3), 'l3': (9, 2)}
    work = {n: {'A': float(a) ** 3, 'G': float(g) ** 3} for n, (a, g) in dims.items()}
    for fraction, expect_second, expect_grad in ((1.0, True, False), (0.5, False, True)):
        asg = KAISAAssignment(work, local_rank=rank, world_size=world, grad_worker_fraction=fraction,
                              group_func=lambda ranks: dist.new_group(ranks))
        layers = {n: _FakeLayer(a, g) for n, (a, g) in dims.items()}
        tdc = TorchDistributedCommunicator()
        ex = GroupedExchange(layers, asg, tdc, rank, torch.device('cpu'))
        assert bool(ex.second_order) == expect_second and bool(ex.gradients) == expect_grad
        # the source ranks fill their buffers with a value that names (layer, buffer); the others with garbage
        code = {n: i + 1 for i, n in enumerate(dims)}
        for n, layer in layers.items():
            for bi, (bname, buf) in enumerate(sorted(layer.bufs.items())):
                src = asg.inv_worker(n, 'A' if bname == 'qa' else 'G')
                buf.fill_(100.0 * code[n] + bi if rank == src else -1.0)
            if layer.grad_store is not None:
                layer.grad_store.fill_(7.0 * code[n] if rank == asg.src_grad_worker(n) else -1.0)
        before = tdc.stats['broadcast_calls']
        if expect_second:
            ex.broadcast_second_order()
            for n, layer in layers.items():
                for bi, (bname, buf) in enumerate(sorted(layer.bufs.items())):
                    assert torch.all(buf == 100.0 * code[n] + bi), (n, bname)
                assert sorted(layer.adopted) == ['A', 'G']
            assert tdc.stats['broadcast_calls'] - before == len(ex.second_order) <= world
        if expect_grad:
            ex.broadcast_gradients()
            for n, layer in layers.items():
                assert torch.all(layer.grad_store == 7.0 * code[n]), n
                assert 'grad' in layer.adopted
            assert tdc.stats['broadcast_calls'] - before == len(ex.gradients) <= world


def test_grouped_exchange_moves_the_data_world2():
    _run('_body_grouped_exchange_proxy')
