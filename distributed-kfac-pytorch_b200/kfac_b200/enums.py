"""Enumerations of the public API; names and values as in the reference (kfac/enums.py:9-53)."""
from __future__ import annotations

from enum import Enum


class AllreduceMethod(Enum):
    """How factors are averaged across ranks: one collective per factor, or adjacent factors coalesced."""

    ALLREDUCE = 1
    ALLREDUCE_BUCKETED = 2


class AssignmentStrategy(Enum):
    """Load-balancing cost of a factor: n^3 (eigendecomposition time) or n^2 (memory)."""

    COMPUTE = 1
    MEMORY = 2


class ComputeMethod(Enum):
    """Second-order data kept per factor: eigendecomposition or explicit damped inverse."""

    EIGEN = 1
    INVERSE = 2


class DistributedStrategy(Enum):
    """KAISA shortcuts for grad_worker_fraction: COMM_OPT = 1, HYBRID_OPT = 1/2, MEM_OPT = 1/world_size."""

    COMM_OPT = 1
    MEM_OPT = 2
    HYBRID_OPT = 3
