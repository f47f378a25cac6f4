"""KFACPreconditioner: the public entry point, drop-in for kfac.preconditioner.KFACPreconditioner.

Keyword arguments, defaults, validation errors and the grad_worker_fraction / DistributedStrategy resolution
follow kfac/preconditioner.py:50-330 of the reference.  Example (identical to the reference's):

    >>> model = torch.nn.parallel.DistributedDataParallel(model, ...)
    >>> optimizer = optim.SGD(model.parameters(), ...)
    >>> preconditioner = KFACPreconditioner(model, ...)
    >>> for data, target in loader:
    ...     optimizer.zero_grad()
    ...     loss = criterion(model(data), target)
    ...     loss.backward()
    ...     preconditioner.step()
    ...     optimizer.step()
"""
from __future__ import annotations

import logging
import warnings
from typing import Any
from typing import Callable

import torch
import torch.distributed as dist

from kfac_b200.assignment import KAISAAssignment
from kfac_b200.base_preconditioner import BaseKFACPreconditioner
from kfac_b200.distributed import get_rank
from kfac_b200.distributed import get_world_size
from kfac_b200.distributed import TorchDistributedCommunicator
from kfac_b200.enums import AllreduceMethod
from kfac_b200.enums import AssignmentStrategy
from kfac_b200.enums import ComputeMethod
from kfac_b200.enums import DistributedStrategy
from kfac_b200.layers.base import KFACBaseLayer
from kfac_b200.layers.eigen import KFACEigenLayer
from kfac_b200.layers.inverse import KFACInverseLayer
from kfac_b200.layers.register import register_modules

logger = logging.getLogger(__name__)


class KFACPreconditioner(BaseKFACPreconditioner):
    """K-FAC preconditioner with the KAISA work distribution, running on hand-written sm_100a kernels."""

    def __init__(
        self,
        model: torch.nn.Module,
        *,
        factor_update_steps: Callable[[int], int] | int = 1,
        inv_update_steps: Callable[[int], int] | int = 1,
        damping: Callable[[int], float] | float = 0.001,
        factor_decay: Callable[[int], float] | float = 0.95,
        kl_clip: Callable[[int], float] | float = 0.001,
        lr: Callable[[int], float] | float = 0.1,
        accumulation_steps: int = 1,
        allreduce_bucket_cap_mb: float = 25.0,
        assignment_strategy: AssignmentStrategy | str = AssignmentStrategy.COMPUTE,
        colocate_factors: bool = True,
        compute_method: ComputeMethod | str = ComputeMethod.EIGEN,
        compute_eigenvalue_outer_product: bool = True,
        grad_worker_fraction: DistributedStrategy | float = DistributedStrategy.COMM_OPT,
        symmetry_aware: bool = False,
        grad_scaler: Any | Callable[[], float] | None = None,
        factor_dtype: torch.dtype | None = None,
        inv_dtype: torch.dtype = torch.float32,
        skip_layers: list[str] | None = None,
        update_factors_in_hook: bool = True,
        loglevel: int = logging.DEBUG,
    ) -> None:
        if allreduce_bucket_cap_mb < 0:
            raise ValueError('allreduce_bucket_cap_mb must be >= 0')
        if isinstance(assignment_strategy, str):
            assignment_strategy = AssignmentStrategy[assignment_strategy.upper()]
        if isinstance(compute_method, str):
            compute_method = ComputeMethod[compute_method.upper()]
        if compute_method == ComputeMethod.EIGEN and compute_eigenvalue_outer_product and not colocate_factors:
            raise ValueError('colocate_factors must be True to use compute_eigenvalue_outer_product')

        # grad_worker_fraction <-> strategy (kfac/preconditioner.py:169-207)
        size = get_world_size()
        if isinstance(grad_worker_fraction, DistributedStrategy):
            distributed_strategy = grad_worker_fraction
            if distributed_strategy == DistributedStrategy.COMM_OPT:
                grad_worker_fraction = 1.0
            elif distributed_strategy == DistributedStrategy.HYBRID_OPT:
                grad_worker_fraction = 0.5
            elif distributed_strategy == DistributedStrategy.MEM_OPT:
                grad_worker_fraction = 1.0 / size
            else:
                raise AssertionError(f'Unknown enum {grad_worker_fraction}')
        else:
            if not 0 <= grad_worker_fraction or not 1 >= grad_worker_fraction:
                raise ValueError('grad_worker_fraction must in [0, 1]')
            if grad_worker_fraction == 0:
                grad_worker_fraction = 1.0 / size
            if size % max(1, round(size * grad_worker_fraction)) != 0:
                raise ValueError('grad_worker_fraction must produce groups of equal size')
            if grad_worker_fraction == 1:
                grad_worker_fraction = 1.0
                distributed_strategy = DistributedStrategy.COMM_OPT
            elif grad_worker_fraction <= 1 / size:
                distributed_strategy = DistributedStrategy.MEM_OPT
            else:
                distributed_strategy = DistributedStrategy.HYBRID_OPT
        assert isinstance(grad_worker_fraction, float)

        if not colocate_factors and distributed_strategy is DistributedStrategy.MEM_OPT:
            warnings.warn(
                'grad_worker_frac=1/world_size (MEM_OPT) requires colocate_factors=True. '
                'Enabling colocate_factors.',
            )
            colocate_factors = True

        self.allreduce_bucket_cap_mb = allreduce_bucket_cap_mb
        self.assignment_strategy = assignment_strategy
        self.colocate_factors = colocate_factors
        self.compute_eigenvalue_outer_product = compute_eigenvalue_outer_product
        self.compute_method = compute_method
        self.distributed_strategy = distributed_strategy
        self.grad_worker_fraction = grad_worker_fraction
        self.grad_scaler = grad_scaler
        self.factor_dtype = factor_dtype
        self.inv_dtype = inv_dtype
        self.skip_layers = [] if skip_layers is None else skip_layers
        self.symmetry_aware = symmetry_aware

        self.allreduce_method = (
            AllreduceMethod.ALLREDUCE_BUCKETED if self.allreduce_bucket_cap_mb > 0 else AllreduceMethod.ALLREDUCE
        )
        self.tdc = TorchDistributedCommunicator(bucket_cap_mb=self.allreduce_bucket_cap_mb)

        layer_kwargs: dict[str, Any] = dict(
            allreduce_method=self.allreduce_method,
            grad_scaler=self.grad_scaler,
            factor_dtype=self.factor_dtype,
            inv_dtype=self.inv_dtype,
            symmetry_aware=self.symmetry_aware,
            tdc=self.tdc,
        )
        layer_type: type[KFACBaseLayer]
        if self.compute_method == ComputeMethod.EIGEN:
            layer_type = KFACEigenLayer
            layer_kwargs['prediv_eigenvalues'] = self.compute_eigenvalue_outer_product
        elif self.compute_method == ComputeMethod.INVERSE:
            layer_type = KFACInverseLayer
        else:
            raise AssertionError(f'Unknown compute_method={self.compute_method}')

        kfac_layers = register_modules(model, kfac_layer_type=layer_type, skip_layers=self.skip_layers, **layer_kwargs)
        for name, kfac_layer in kfac_layers.values():
            logger.log(loglevel, f'Registered name="{name}": {repr(kfac_layer)}')

        # load-balancing cost per factor (kfac/preconditioner.py:266-281)
        if self.assignment_strategy == AssignmentStrategy.COMPUTE:
            cost = lambda n: n ** 3  # noqa: E731
        elif self.assignment_strategy == AssignmentStrategy.MEMORY:
            cost = lambda n: n ** 2  # noqa: E731
        else:
            raise AssertionError(f'Unknown assignment_strategy={self.assignment_strategy}')
        work = {
            name: {'A': cost(layer.module.a_factor_shape[0]), 'G': cost(layer.module.g_factor_shape[0])}
            for name, layer in kfac_layers.values()
        }
        assignment = KAISAAssignment(
            work,
            local_rank=get_rank(),
            world_size=size,
            grad_worker_fraction=self.grad_worker_fraction,
            group_func=dist.new_group if dist.is_initialized() else (lambda ranks: None),
            colocate_factors=self.colocate_factors,
        )
        logger.log(loglevel, f'KFAC layer assignments: {assignment}')

        defaults = {
            'allreduce_bucket_cap_mb': self.allreduce_bucket_cap_mb,
            'allreduce_method': self.allreduce_method,
            'assignment_strategy': self.assignment_strategy,
            'colocate_factors': self.colocate_factors,
            'compute_eigenvalue_outer_product': self.compute_eigenvalue_outer_product,
            'compute_method': self.compute_method,
            'distributed_strategy': self.distributed_strategy,
            'grad_worker_fraction': self.grad_worker_fraction,
            'grad_scaler': self.grad_scaler is not None,
            'factor_dtype': self.factor_dtype,
            'inv_dtype': self.inv_dtype,
            'skip_layers': self.skip_layers,
            'symmetry_aware': self.symmetry_aware,
        }
        super().__init__(
            kfac_layers,
            factor_update_steps=factor_update_steps,
            inv_update_steps=inv_update_steps,
            factor_decay=factor_decay,
            damping=damping,
            kl_clip=kl_clip,
            lr=lr,
            accumulation_steps=accumulation_steps,
            assignment=assignment,
            update_factors_in_hook=update_factors_in_hook,
            defaults=defaults,
            tdc=self.tdc,
            loglevel=loglevel,
        )
