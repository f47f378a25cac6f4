"""Grouped KAISA exchanges: ONE broadcast per (communication group, source rank) instead of one per tensor.

The reference sends every piece of second-order data and every preconditioned gradient on its own
(kfac/layers/eigen.py:177-292, kfac/layers/inverse.py:111-183, kfac/layers/base.py:223-251): three collectives per
layer and step, 162 for ResNet-50.  Here everything one rank sends into one group lives in one flat arena -- the
eigensolver / Newton-Schulz / precondition kernels write straight into views of it, the receivers' layer attributes
are views of their copy -- so a step needs one broadcast per source rank and group, at most `group size` of them
("AllGather of per-rank arenas").  The placement itself is KAISAAssignment's, unchanged.

`plan_*` are pure host functions of the assignment (tested on the CPU); every rank derives the same plan, and walks
its arenas in the same sorted order, so all members of a group enter the same collectives in the same order.
"""
from __future__ import annotations

from dataclasses import dataclass
from dataclasses import field
from typing import Any

import torch

from kfac_b200.assignment import WorkAssignment
from kfac_b200.distributed import Pending
from kfac_b200.distributed import TorchDistributedCommunicator


@dataclass
class ArenaEntry:
    layer: str
    what: str                                  # 'A', 'G' (second-order data of that factor) or 'grad'
    buffers: list[tuple[str, tuple[int, ...]]]  # (buffer name, shape) in arena order
    offset: int = 0                            # first float of this entry in the arena


@dataclass
class Arena:
    ranks: tuple[int, ...]
    src: int
    group: Any
    entries: list[ArenaEntry] = field(default_factory=list)
    numel: int = 0
    tensor: torch.Tensor | None = None

    @property
    def key(self) -> tuple[tuple[int, ...], int]:
        return (self.ranks, self.src)


def _numel(shape: tuple[int, ...]) -> int:
    n = 1
    for s in shape:
        n *= s
    return n


def _add(arenas: dict, ranks: frozenset[int], src: int, group: Any, entry: ArenaEntry) -> None:
    key = (tuple(sorted(ranks)), src)
    arena = arenas.get(key)
    if arena is None:
        arena = arenas[key] = Arena(ranks=key[0], src=src, group=group)
    entry.offset = arena.numel
    arena.entries.append(entry)
    arena.numel += sum(_numel(shape) for _, shape in entry.buffers)


def plan_second_order(
    payloads: dict[str, dict[str, list[tuple[str, tuple[int, ...]]]]],
    assignment: WorkAssignment,
    local_rank: int,
) -> list[Arena]:
    """Arenas of second-order data this rank takes part in.  payloads[layer][factor] lists the buffers that the
    inverse worker of (layer, factor) sends to the layer's gradient-worker group (eigen.py:214-225,274-292)."""
    arenas: dict = {}
    if not assignment.broadcast_inverses():
        return []
    for layer in payloads:                      # registration order: identical on every rank
        if not assignment.is_grad_worker(layer):
            continue
        ranks = assignment.grad_worker_ranks(layer)
        group = assignment.grad_worker_group(layer)
        for factor in ('A', 'G'):
            bufs = payloads[layer].get(factor)
            if bufs:
                _add(arenas, ranks, assignment.inv_worker(layer, factor), group, ArenaEntry(layer, factor, list(bufs)))
    return [arenas[k] for k in sorted(arenas)]


def plan_gradients(
    shapes: dict[str, tuple[int, ...]],
    assignment: WorkAssignment,
    local_rank: int,
) -> list[Arena]:
    """Arenas of preconditioned gradients: one per source rank in this rank's receiver group (base.py:223-251)."""
    arenas: dict = {}
    if not assignment.broadcast_gradients():
        return []
    for layer, shape in shapes.items():
        _add(arenas, assignment.grad_receiver_ranks(layer), assignment.src_grad_worker(layer),
             assignment.grad_receiver_group(layer), ArenaEntry(layer, 'grad', [('grad', tuple(shape))]))
    return [arenas[k] for k in sorted(arenas)]


def views_of(arena: Arena, entry: ArenaEntry) -> dict[str, torch.Tensor]:
    """The entry's buffers as views into the arena's tensor."""
    assert arena.tensor is not None
    out = {}
    off = entry.offset
    for name, shape in entry.buffers:
        n = _numel(shape)
        out[name] = arena.tensor[off:off + n].view(shape)
        off += n
    return out


class GroupedExchange:
    """Owns the arenas of one preconditioner on one rank and issues the grouped broadcasts."""

    def __init__(self, layers: dict[str, Any], assignment: WorkAssignment, tdc: TorchDistributedCommunicator,
                 local_rank: int, device: torch.device) -> None:
        self.layers = layers
        self.tdc = tdc
        payloads = {name: layer.second_order_payload() for name, layer in layers.items()}
        shapes = {name: tuple(layer._grad_shape()) for name, layer in layers.items()}
        self.second_order = plan_second_order(payloads, assignment, local_rank)
        self.gradients = plan_gradients(shapes, assignment, local_rank)
        for arena in self.second_order + self.gradients:
            arena.tensor = torch.empty(arena.numel, dtype=torch.float32, device=device)
            for entry in arena.entries:
                views = views_of(arena, entry)
                if entry.what == 'grad':
                    layers[entry.layer].attach_grad_buffer(views['grad'])
                else:
                    layers[entry.layer].attach_second_order_buffers(views)

    def _broadcast(self, arenas: list[Arena]) -> None:
        pending: list[Pending | torch.Tensor] = [self.tdc.broadcast(a.tensor, src=a.src, group=a.group) for a in arenas]
        for p in pending:
            if isinstance(p, Pending):
                p.wait()

    def broadcast_second_order(self) -> None:
        """One broadcast per (worker group, inverse worker); afterwards every gradient worker's layer attributes point
        at the received (or locally computed) data."""
        self._broadcast(self.second_order)
        for arena in self.second_order:
            for entry in arena.entries:
                self.layers[entry.layer].adopt_second_order(entry.what)

    def broadcast_gradients(self) -> None:
        """One broadcast per (receiver group, source gradient worker)."""
        self._broadcast(self.gradients)
        for arena in self.gradients:
            for entry in arena.entries:
                self.layers[entry.layer].adopt_grad()

    def collectives_per_step(self) -> dict[str, int]:
        return {'second_order': len(self.second_order), 'gradients': len(self.gradients)}
