"""KAISA work placement (host-side, pure Python): which rank eigendecomposes which factor, which ranks
precondition which layer, and who receives gradients from whom.

Semantics follow kfac/assignment.py:29-470 of the reference exactly -- the placement drives the order of
collectives, so it has to be identical on every rank and identical to the reference's
(tests/test_assignment.py checks it against the reference's own golden tables).
"""
from __future__ import annotations

from abc import ABCMeta
from abc import abstractmethod
from typing import Any
from typing import Callable


class WorkAssignment(metaclass=ABCMeta):
    """Interface the preconditioner queries (kfac/assignment.py:29-117)."""

    def __repr__(self) -> str:
        rows = [
            f'  layer="{layer}": is_grad_worker={self.is_grad_worker(layer)}, '
            f'src_grad_worker={self.src_grad_worker(layer)}, '
            f'inv_workers={ {f: self.inv_worker(layer, f) for f in self.get_factors(layer)} }'
            for layer in sorted(self.get_layers())
        ]
        return f'{self.__class__.__name__}(\n' + ',\n'.join(rows) + '\n)'

    @abstractmethod
    def broadcast_gradients(self) -> bool:
        """True if preconditioned gradients have to be sent to ranks that did not compute them."""

    @abstractmethod
    def broadcast_inverses(self) -> bool:
        """True if second-order data has to be sent from the inverse worker to other gradient workers."""

    @abstractmethod
    def get_layers(self) -> tuple[str, ...]:
        """Names of all assigned layers."""

    @abstractmethod
    def get_factors(self, layer: str) -> tuple[str, ...]:
        """Names of the factors of a layer."""

    @abstractmethod
    def inv_worker(self, layer: str, factor: str) -> int:
        """Rank that computes the second-order data of (layer, factor)."""

    @abstractmethod
    def is_grad_worker(self, layer: str) -> bool:
        """Whether this rank preconditions the gradient of the layer."""

    @abstractmethod
    def src_grad_worker(self, layer: str) -> int:
        """Rank this rank gets the layer's preconditioned gradient from (itself if it is a worker)."""

    @abstractmethod
    def factor_group(self, layer: str, factor: str) -> Any:
        """Communication group over which the factor is averaged."""

    @abstractmethod
    def grad_worker_group(self, layer: str) -> Any:
        """Communication group of the layer's gradient workers (second-order broadcast)."""

    @abstractmethod
    def grad_receiver_group(self, layer: str) -> Any:
        """Communication group in which the preconditioned gradient is broadcast."""


class KAISAAssignment(WorkAssignment):
    """Ranks on a (grad_workers x world/grad_workers) grid: columns precondition together, rows share results."""

    def __init__(
        self,
        work: dict[str, dict[str, float]],
        *,
        local_rank: int,
        world_size: int,
        grad_worker_fraction: float,
        group_func: Callable[[list[int]], Any],
        colocate_factors: bool = True,
    ) -> None:
        if not 0 <= grad_worker_fraction <= 1:
            raise ValueError(f'grad_worker_fraction must be in [0, 1]. Got {grad_worker_fraction}.')
        if local_rank < 0:
            raise ValueError('local_rank must be > 0')
        if world_size < 0:
            raise ValueError('world_size must be > 0')
        workers = max(1, world_size * grad_worker_fraction)
        if workers != int(workers):
            raise ValueError(
                'world_size*grad_worker_fraction must produce an integer value. '
                f'Found {world_size}*{grad_worker_fraction}={workers}.',
            )
        if local_rank >= world_size:
            raise ValueError(f'local_rank={local_rank} larger than world_size={world_size}')
        self.local_rank = local_rank
        self.world_size = world_size
        self.grad_worker_fraction = grad_worker_fraction
        self.grad_workers = int(workers)
        self.group_func = group_func
        self.colocate_factors = colocate_factors

        worker_sets = self.partition_grad_workers(world_size, self.grad_workers)
        receiver_sets = self.partition_grad_receivers(world_size, self.grad_workers)
        # one communication group per distinct rank set; every rank walks the same set in the same order so the
        # collective group creation lines up (kfac/assignment.py:192-200)
        handles = {ranks: group_func(list(ranks)) for ranks in worker_sets | receiver_sets}

        self._inv_assignments = self.greedy_assignment(
            work, [list(ranks) for ranks in worker_sets], world_size, colocate_factors,
        )
        self._worker_group: dict[str, tuple[frozenset[int], Any]] = {}
        self._receiver_group: dict[str, tuple[frozenset[int], Any]] = {}
        for layer, placed in self._inv_assignments.items():
            anchor = list(placed.values()).pop()
            for ranks in worker_sets:
                if anchor in ranks:
                    self._worker_group[layer] = (ranks, handles[ranks])
            for ranks in receiver_sets:
                if local_rank in ranks:
                    self._receiver_group[layer] = (ranks, handles[ranks])

    # ------------------------------------------------------------------ placement
    @staticmethod
    def greedy_assignment(
        work: dict[str, dict[str, float]],
        worker_groups: list[list[int]],
        world_size: int,
        colocate_factors: bool,
    ) -> dict[str, dict[str, int]]:
        """Longest-processing-time greedy placement constrained to worker groups (kfac/assignment.py:226-318).

        Layers are visited by decreasing total cost; each goes to the least-loaded worker group, then either
        entirely to that group's least-loaded rank (colocate) or factor by factor, costliest first.
        Ties always resolve to the first candidate, which makes the result reproducible on every rank.
        """
        load = [0.0] * world_size
        placed = {layer: {factor: -1 for factor in factors} for layer, factors in work.items()}
        totals = {layer: sum(factors.values()) for layer, factors in work.items()}
        for layer, _ in sorted(totals.items(), key=lambda kv: kv[1], reverse=True):
            group_load = [sum(load[r] for r in group) for group in worker_groups]
            group = worker_groups[group_load.index(min(group_load))]
            if colocate_factors:
                member_load = [load[r] for r in group]
                rank = group[member_load.index(min(member_load))]
                load[rank] += totals[layer]
                for factor in work[layer]:
                    placed[layer][factor] = rank
            else:
                for factor, cost in sorted(work[layer].items(), key=lambda kv: (kv[1], kv[0]), reverse=True):
                    member_load = [load[r] for r in group]
                    rank = group[member_load.index(min(member_load))]
                    load[rank] += cost
                    placed[layer][factor] = rank
        for layer in placed:
            for factor in placed[layer]:
                assert placed[layer][factor] >= 0
        return placed

    @staticmethod
    def partition_grad_workers(world_size: int, grad_workers: int) -> set[frozenset[int]]:
        """Columns of the rank grid: {i, i + p, i + 2p, ...} with p = world/grad_workers (assignment.py:320-362)."""
        if not 0 < world_size:
            raise ValueError('world_size must be > 0')
        if world_size % grad_workers != 0:
            raise ValueError('world_size must be an integer multiple of the gradient worker count')
        p = world_size // grad_workers
        return {frozenset(range(i, world_size, p)) for i in range(p)}

    @staticmethod
    def partition_grad_receivers(world_size: int, grad_workers: int) -> set[frozenset[int]]:
        """Rows of the rank grid: {ip, ..., ip + p - 1} (assignment.py:364-394)."""
        if not 0 < world_size:
            raise ValueError('world_size must be > 0')
        if world_size % grad_workers != 0:
            raise ValueError('world_size must be an integer multiple of the gradient worker count')
        p = world_size // grad_workers
        return {frozenset(range(i * p, i * p + p)) for i in range(grad_workers)}

    # ------------------------------------------------------------------ queries
    def broadcast_gradients(self) -> bool:
        return self.grad_workers < self.world_size

    def broadcast_inverses(self) -> bool:
        return self.grad_workers > 1

    def get_layers(self) -> tuple[str, ...]:
        return tuple(self._inv_assignments.keys())

    def get_factors(self, layer: str) -> tuple[str, ...]:
        return tuple(self._inv_assignments[layer].keys())

    def inv_worker(self, layer: str, factor: str) -> int:
        return self._inv_assignments[layer][factor]

    def is_grad_worker(self, layer: str) -> bool:
        return self.local_rank in self._worker_group[layer][0]

    def src_grad_worker(self, layer: str) -> int:
        return set(self._worker_group[layer][0] & self._receiver_group[layer][0]).pop()

    def factor_group(self, layer: str, factor: str) -> Any:
        """Factors are averaged over the whole world: plain data parallelism (assignment.py:441-452)."""
        return None

    def grad_worker_group(self, layer: str) -> Any:
        return self._worker_group[layer][1]

    def grad_receiver_group(self, layer: str) -> Any:
        return self._receiver_group[layer][1]

    def grad_worker_ranks(self, layer: str) -> frozenset[int]:
        """Ranks of the layer's gradient-worker group (used to size grouped broadcasts)."""
        return self._worker_group[layer][0]

    def grad_receiver_ranks(self, layer: str) -> frozenset[int]:
        return self._receiver_group[layer][0]
