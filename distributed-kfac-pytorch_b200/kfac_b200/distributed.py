"""Collective communication for the K-FAC step, over torch.distributed (NCCL on NVLink/NVSwitch, gloo in tests).

Same public surface as the reference's kfac/distributed.py (TorchDistributedCommunicator with allreduce /
broadcast / allreduce_bucketed / flush_allreduce_buckets, get_rank / get_world_size, get_triu / fill_triu,
NonSquareTensorError) but built for persistent device arenas instead of flatten/unflatten copies:

* "bucketed" allreduce coalesces tensors that are ADJACENT IN MEMORY (the preconditioner lays its factors
  out in hook order inside one arena) into a single collective on a flat view -- no flatten, no unflatten,
  no per-tensor callback (reference: kfac/distributed.py:35-115 copies twice per bucket);
* averaging uses ReduceOp.AVG when the backend has it (NCCL) instead of a separate multiply pass
  (reference: kfac/distributed.py:235,363);
* results are handed back as `Pending` handles: `.wait()` orders the caller's CUDA stream after the
  collective (no host block on NCCL) and returns the tensor, like the Future the reference returns.
"""
from __future__ import annotations

from typing import Any
from typing import Callable

import torch
import torch.distributed as dist


class NonSquareTensorError(Exception):
    """Symmetric communication was requested for a tensor that is not 2-D square (distributed.py:29-33)."""


def get_rank(group: Any = None) -> int:
    """Rank in the group, 0 when torch.distributed is not initialised (distributed.py:382-396)."""
    return dist.get_rank(group) if dist.is_initialized() else 0


def get_world_size(group: Any = None) -> int:
    """Group size, 1 when torch.distributed is not initialised (distributed.py:399-413)."""
    return dist.get_world_size(group) if dist.is_initialized() else 1


def get_triu(tensor: torch.Tensor) -> torch.Tensor:
    """Row-major upper triangle as a flat vector (distributed.py:416-427), native kernel, no index tensors."""
    if tensor.dim() != 2:
        raise ValueError('triu(tensor) requires tensor to be 2 dimensional')
    if tensor.shape[0] > tensor.shape[1]:
        raise ValueError('tensor cannot have more rows than columns')
    if tensor.shape[0] != tensor.shape[1]:
        raise NonSquareTensorError('packed communication is implemented for square factors only')
    from kfac_b200 import ops
    return ops.triu_pack(tensor.contiguous())


def fill_triu(shape: tuple[int, ...], triu_tensor: torch.Tensor) -> torch.Tensor:
    """Symmetric matrix from its packed upper triangle (distributed.py:430-459)."""
    if len(shape) != 2:
        raise ValueError('shape must be 2 dimensional')
    from kfac_b200 import ops
    return ops.triu_unpack(triu_tensor.contiguous(), shape[0])


class Pending:
    """Handle to a tensor whose collective may still be in flight."""

    def __init__(self, tensor: torch.Tensor, works: list[Any] | None = None,
                 finish: Callable[[torch.Tensor], torch.Tensor] | None = None) -> None:
        self._tensor = tensor
        self._works = works or []
        self._finish = finish
        self._done = False

    def wait(self) -> torch.Tensor:
        if not self._done:
            for w in self._works:
                if w is not None:
                    w.wait()
            if self._finish is not None:
                self._tensor = self._finish(self._tensor)
            self._works = []
            self._done = True
        return self._tensor

    # torch.futures.Future compatibility for callers that used .value()
    def value(self) -> torch.Tensor:
        return self.wait()


def resolve(x: Any) -> Any:
    """Tensor behind a Pending handle (or x itself)."""
    return x.wait() if isinstance(x, Pending) else x


def _supports_avg(group: Any) -> bool:
    try:
        return dist.get_backend(group) == 'nccl'
    except Exception:  # pragma: no cover
        return False


def _same_storage_adjacent(prev: torch.Tensor, nxt: torch.Tensor) -> bool:
    """True if `nxt` starts exactly where `prev` ends inside the same storage (both contiguous, same dtype)."""
    return (
        prev.dtype == nxt.dtype
        and prev.device == nxt.device
        and prev.is_contiguous()
        and nxt.is_contiguous()
        and prev.untyped_storage().data_ptr() == nxt.untyped_storage().data_ptr()
        and prev.storage_offset() + prev.numel() == nxt.storage_offset()
    )


class _Run:
    """A run of memory-adjacent tensors that will travel as one collective."""

    def __init__(self, group: Any, average: bool) -> None:
        self.group = group
        self.average = average
        self.tensors: list[torch.Tensor] = []
        self.handles: list[Pending] = []
        self.nbytes = 0
        self.launched = False

    def flat(self) -> torch.Tensor:
        first = self.tensors[0]
        n = sum(t.numel() for t in self.tensors)
        return torch.as_strided(first, (n,), (1,), first.storage_offset())


class TorchDistributedCommunicator:
    """Collectives for factors, second-order data and gradients."""

    def __init__(self, bucket_cap_mb: float = 25.0) -> None:
        self._bucket_cap_mb = bucket_cap_mb
        self._runs: dict[Any, _Run | None] = {}
        self.stats = {'allreduce_calls': 0, 'allreduce_bytes': 0, 'broadcast_calls': 0, 'broadcast_bytes': 0}

    @property
    def bucket_cap_bytes(self) -> int:
        """Cap in bytes; MB = 10^6 bytes as in the reference (distributed.py:137-140)."""
        return int(self._bucket_cap_mb * 1000 * 1000)

    # ------------------------------------------------------------------ helpers
    def _check_square(self, tensor: torch.Tensor) -> None:
        shape = tensor.size()
        if len(shape) != 2 or shape[0] != shape[1]:
            raise NonSquareTensorError(
                f'Symmetric communication can only be done with a 2D square tensor. Got tensor with shape {shape}.',
            )

    def _all_reduce(self, flat: torch.Tensor, group: Any, average: bool) -> tuple[Any, bool]:
        """Launch the collective; returns (work, needs_scale)."""
        self.stats['allreduce_calls'] += 1
        self.stats['allreduce_bytes'] += flat.numel() * flat.element_size()
        if average and _supports_avg(group):
            return dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=group, async_op=True), False
        return dist.all_reduce(flat, group=group, async_op=True), average

    # ------------------------------------------------------------------ allreduce
    def allreduce(self, tensor: torch.Tensor, *, average: bool = False, group: Any = None,
                  symmetric: bool = False) -> Pending | torch.Tensor:
        """Asynchronous in-place allreduce of one tensor (distributed.py:184-240)."""
        if get_world_size(group) == 1:
            return tensor
        shape = tensor.size()
        if symmetric:
            self._check_square(tensor)
            payload = get_triu(tensor)
        else:
            payload = tensor if tensor.is_contiguous() else tensor.contiguous()
        work, needs_scale = self._all_reduce(payload, group, average)
        world = get_world_size(group)

        def finish(t: torch.Tensor) -> torch.Tensor:
            if needs_scale:
                t.mul_(1.0 / world)
            if symmetric:
                from kfac_b200 import ops
                return ops.triu_unpack(t, shape[0], out=tensor if tensor.is_contiguous() else None)
            if t is not tensor:
                tensor.copy_(t)
            return tensor

        return Pending(payload, [work], finish)

    def allreduce_bucketed(self, tensor: torch.Tensor, *, average: bool = False, group: Any = None,
                           symmetric: bool = False) -> Pending | torch.Tensor:
        """Allreduce that coalesces memory-adjacent tensors into one collective (distributed.py:299-368).

        A pending run is launched when the next tensor is not adjacent to it, when adding the tensor would exceed
        the bucket cap (a single tensor larger than the cap travels alone, as in the reference), or on
        flush_allreduce_buckets().
        """
        if get_world_size(group) == 1:
            return tensor
        if symmetric:
            # packed payloads are fresh buffers, never adjacent: send on their own
            return self.allreduce(tensor, average=average, group=group, symmetric=True)
        if not tensor.is_contiguous():
            return self.allreduce(tensor, average=average, group=group)
        key = id(group) if group is not None else None
        run = self._runs.get(key)
        nbytes = tensor.numel() * tensor.element_size()
        if run is not None and (
            run.average != average
            or not _same_storage_adjacent(run.tensors[-1], tensor)
            or run.nbytes + nbytes > self.bucket_cap_bytes
        ):
            self._launch(run)
            run = None
        if run is None:
            run = _Run(group, average)
            self._runs[key] = run
        handle = Pending(tensor)
        run.tensors.append(tensor)
        run.handles.append(handle)
        run.nbytes += nbytes
        handle._works = [_LazyRunWork(self, run)]
        return handle

    def _launch(self, run: _Run) -> None:
        if run.launched or not run.tensors:
            return
        run.launched = True
        flat = run.flat()
        work, needs_scale = self._all_reduce(flat, run.group, run.average)
        world = get_world_size(run.group)
        scaled = [False]

        def finish_once() -> None:
            work.wait()
            if needs_scale and not scaled[0]:
                flat.mul_(1.0 / world)
                scaled[0] = True

        run.finish_once = finish_once  # type: ignore[attr-defined]
        for key, r in list(self._runs.items()):
            if r is run:
                self._runs[key] = None

    def flush_allreduce_buckets(self) -> None:
        """Launch every pending run (distributed.py:374-379)."""
        for run in list(self._runs.values()):
            if run is not None:
                self._launch(run)

    # ------------------------------------------------------------------ broadcast
    def broadcast(self, tensor: torch.Tensor, *, src: int, group: Any = None,
                  symmetric: bool = False) -> Pending | torch.Tensor:
        """Asynchronous in-place broadcast from global rank `src` (distributed.py:242-297)."""
        if get_world_size(group) == 1:
            return tensor
        shape = tensor.size()
        if symmetric:
            self._check_square(tensor)
            payload = get_triu(tensor)
        else:
            payload = tensor if tensor.is_contiguous() else tensor.contiguous()
        self.stats['broadcast_calls'] += 1
        self.stats['broadcast_bytes'] += payload.numel() * payload.element_size()
        work = dist.broadcast(payload, src=src, group=group, async_op=True)

        def finish(t: torch.Tensor) -> torch.Tensor:
            if symmetric:
                from kfac_b200 import ops
                return ops.triu_unpack(t, shape[0], out=tensor if tensor.is_contiguous() else None)
            if t is not tensor:
                tensor.copy_(t)
            return tensor

        return Pending(payload, [work], finish)

    def group_ranks(self, group: Any) -> frozenset[int]:
        return frozenset(range(get_world_size(group)))


class _LazyRunWork:
    """`work` stand-in for tensors of a coalesced run: waiting launches the run if it is still open."""

    def __init__(self, comm: TorchDistributedCommunicator, run: _Run) -> None:
        self._comm = comm
        self._run = run

    def wait(self) -> None:
        if not self._run.launched:
            self._comm._launch(self._run)
        self._run.finish_once()  # type: ignore[attr-defined]
