"""Per-layer K-FAC state machine over the native kernels.

Mirrors KFACBaseLayer of the reference (kfac/layers/base.py:18-422): same constructor arguments, same state
(`a_factor`, `g_factor`, `grad`, batch accumulators and counts), same method names and error behaviour, so
`preconditioner._layers[module] = (name, layer)` stays inspectable the way the reference's tests and
`state_dict` expect.  What changes is underneath:

* factors are fp32 buffers the layer owns for its lifetime (optionally views into the preconditioner's arena
  so that neighbouring factors coalesce into one collective); they are updated IN PLACE by the kernels;
* `save_layer_input` / `save_layer_grad_output` run one fused launch (im2col + bias column + scaling +
  accumulate) instead of clone/pad/unfold/cat/matmul/transposes;
* `fused_a_update` / `fused_g_update` fold the running average into that launch when no micro-batch
  accumulation is pending (8 B per factor element instead of 12 + temporaries);
* the preconditioned gradient is written straight into `weight.grad` / `bias.grad`.
"""
from __future__ import annotations

from typing import Any
from typing import Callable

import torch

from kfac_b200 import ops
from kfac_b200.distributed import get_rank
from kfac_b200.distributed import Pending
from kfac_b200.distributed import resolve
from kfac_b200.distributed import TorchDistributedCommunicator
from kfac_b200.enums import AllreduceMethod
from kfac_b200.layers.modules import ModuleHelper


class KFACBaseLayer:
    """State and communication shared by the eigen and inverse layers."""

    def __init__(
        self,
        module: ModuleHelper,
        *,
        tdc: TorchDistributedCommunicator,
        allreduce_method: AllreduceMethod = AllreduceMethod.ALLREDUCE,
        factor_dtype: torch.dtype | None = None,
        grad_scaler: Any | Callable[[], float] | None = None,
        inv_dtype: torch.dtype = torch.float32,
        symmetry_aware: bool = False,
    ) -> None:
        self.module = module
        self.tdc = tdc
        self.allreduce_method = allreduce_method
        self.factor_dtype = factor_dtype
        if grad_scaler is not None and hasattr(grad_scaler, 'get_scale'):
            grad_scaler = grad_scaler.get_scale
        self.grad_scaler: Callable[[], float] | None = grad_scaler
        self.inv_dtype = inv_dtype
        self.symmetry_aware = symmetry_aware
        # Low-precision modes (kfac/layers/base.py:349,363-365,382-383; kfac/layers/eigen.py:318-319): factor_dtype =
        # bf16 / fp16 rounds activations and grad-outputs to that type before the covariance, exactly as the reference's
        # `.to(self.factor_dtype)`; the products of the rounded values are then EXACT on the tensor cores (single-pass
        # tf32 UMMA, fp32 accumulation) and the factor itself stays in fp32 -- tighter than the reference, which also
        # accumulates and stores in the 16-bit type.  inv_dtype = bf16 / fp16 rounds the second-order data to that
        # type's values (they stay in fp32 storage, which is what the kernels read).  Stated tolerances against the
        # fp32 path: 2e-2 (bf16) / 3e-3 (fp16), tests/test_low_precision_gpu.py.
        if factor_dtype not in (None, torch.float32, torch.bfloat16, torch.float16):
            raise ValueError(f'factor_dtype must be None, float32, bfloat16 or float16, got {factor_dtype}')
        if inv_dtype not in (torch.float32, torch.bfloat16, torch.float16):
            raise ValueError(f'inv_dtype must be float32, bfloat16 or float16, got {inv_dtype}')

        self.eps = 1e-10
        self.symmetric_factors = self.module.has_symmetric_factors()

        # batch accumulators (sum over micro-batches) and how many were summed (base.py:73-80)
        self._a_batch: torch.Tensor | None = None
        self._g_batch: torch.Tensor | None = None
        self._a_count: int = 0
        self._g_count: int = 0
        # running averages (tensor, or a Pending handle while their allreduce is in flight)
        self._a_factor: torch.Tensor | Pending | None = None
        self._g_factor: torch.Tensor | Pending | None = None
        # preconditioned gradient, stacked (out, in [+1])
        self._grad: torch.Tensor | Pending | None = None
        # persistent storage, optionally provided by the preconditioner's arena
        self._a_store: torch.Tensor | None = None
        self._g_store: torch.Tensor | None = None
        self._a_batch_store: torch.Tensor | None = None
        self._g_batch_store: torch.Tensor | None = None
        self._grad_store: torch.Tensor | None = None

    def __repr__(self) -> str:
        return f'{self.__class__.__name__}({repr(self.module)})'

    # ------------------------------------------------------------------ storage
    def attach_storage(self, a_store: torch.Tensor, g_store: torch.Tensor) -> None:
        """Use caller-provided (arena) buffers for the running averages."""
        assert tuple(a_store.shape) == self.module.a_factor_shape
        assert tuple(g_store.shape) == self.module.g_factor_shape
        self._a_store, self._g_store = a_store, g_store

    def _store(self, which: str) -> torch.Tensor:
        attr = f'_{which}_store'
        buf = getattr(self, attr)
        if buf is None:
            shape = self.module.a_factor_shape if which in ('a', 'a_batch') else self.module.g_factor_shape
            buf = torch.empty(shape, dtype=torch.float32, device=self.module.device)
            setattr(self, attr, buf)
        return buf

    # ------------------------------------------------------------------ properties (base.py:92-127)
    @property
    def a_factor(self) -> torch.Tensor | None:
        self._a_factor = resolve(self._a_factor)
        return self._a_factor  # type: ignore[return-value]

    @a_factor.setter
    def a_factor(self, value: torch.Tensor | Pending | None) -> None:
        self._a_factor = value

    @property
    def g_factor(self) -> torch.Tensor | None:
        self._g_factor = resolve(self._g_factor)
        return self._g_factor  # type: ignore[return-value]

    @g_factor.setter
    def g_factor(self, value: torch.Tensor | Pending | None) -> None:
        self._g_factor = value

    @property
    def grad(self) -> torch.Tensor | None:
        self._grad = resolve(self._grad)
        return self._grad  # type: ignore[return-value]

    @grad.setter
    def grad(self, value: torch.Tensor | Pending | None) -> None:
        self._grad = value

    # ------------------------------------------------------------------ checkpoint (base.py:129-164)
    def state_dict(self) -> dict[str, torch.Tensor | None]:
        """Only the running averages: second-order data is recomputed on load."""
        # copies, not views: the running averages live in an arena that the factor kernels keep updating IN PLACE,
        # while the reference rebinds a_factor / g_factor to fresh tensors on every update (base.py:388,404), so a
        # state_dict held in memory there is a snapshot -- and must be one here too
        a, g = self.a_factor, self.g_factor
        return {'A': a.clone() if a is not None else None, 'G': g.clone() if g is not None else None}

    def load_state_dict(self, state_dict: dict[str, torch.Tensor | None]) -> None:
        if 'A' not in state_dict or 'G' not in state_dict:
            raise KeyError("KFACLayer state_dict must contain keys 'A' and 'G'")
        device = self.module.device
        if state_dict['A'] is not None:
            self.a_factor = self._store('a').copy_(state_dict['A'].to(device))
        if state_dict['G'] is not None:
            self.g_factor = self._store('g').copy_(state_dict['G'].to(device))

    def memory_usage(self) -> dict[str, int]:
        """Bytes held by this layer on this rank (base.py:166-183)."""
        def size(t: torch.Tensor | None) -> int:
            return t.nelement() * t.element_size() if t is not None else 0
        return {
            'a_factors': size(self.a_factor),
            'g_factors': size(self.g_factor),
            'a_batch': size(self._a_batch),
            'g_batch': size(self._g_batch),
        }

    # ------------------------------------------------------------------ abstract second-order ops
    def broadcast_a_inv(self, src: int, group: Any = None) -> None:
        raise NotImplementedError

    def broadcast_g_inv(self, src: int, group: Any = None) -> None:
        raise NotImplementedError

    def compute_a_inv(self, damping: float = 0.001) -> None:
        raise NotImplementedError

    def compute_g_inv(self, damping: float = 0.001) -> None:
        raise NotImplementedError

    def preconditioned_grad(self, damping: float = 0.001, vg_accum: torch.Tensor | None = None) -> None:
        raise NotImplementedError

    # ------------------------------------------------------------------ communication (base.py:223-335)
    def broadcast_grad(self, src: int, group: Any = None) -> None:
        """Broadcast the preconditioned gradient from `src`; every rank of the group enters."""
        if self.grad is None:
            if get_rank() == src:
                raise RuntimeError(
                    f'Attempt to broadcast gradient from src={src} but this rank has not computed the '
                    'preconditioned gradient yet.',
                )
            self.grad = self._grad_buffer()
        self.grad = self.tdc.broadcast(self.grad, src=src, group=group)  # type: ignore[arg-type]

    def _allreduce(self):
        if self.allreduce_method == AllreduceMethod.ALLREDUCE:
            return self.tdc.allreduce
        if self.allreduce_method == AllreduceMethod.ALLREDUCE_BUCKETED:
            return self.tdc.allreduce_bucketed
        raise AssertionError(f'Unknown allreduce_method={self.allreduce_method}')

    def reduce_a_factor(self, group: Any = None) -> None:
        """Start averaging A over the group; the result replaces a_factor (possibly as a pending handle)."""
        if self.a_factor is None:
            raise RuntimeError('a_factor is None, cannot reduce')
        self.a_factor = self._allreduce()(
            self.a_factor, average=True, symmetric=self.symmetric_factors and self.symmetry_aware, group=group,
        )

    def reduce_g_factor(self, group: Any = None) -> None:
        if self.g_factor is None:
            raise RuntimeError('g_factor is None, cannot reduce')
        self.g_factor = self._allreduce()(
            self.g_factor, average=True, symmetric=self.symmetric_factors and self.symmetry_aware, group=group,
        )

    # ------------------------------------------------------------------ factor stage (base.py:337-404)
    def reset_batch(self) -> None:
        self._a_batch = None
        self._a_count = 0
        self._g_batch = None
        self._g_count = 0

    def _g_in_scale(self) -> float:
        return 1.0 / self.grad_scaler() if self.grad_scaler is not None else 1.0

    def _cast_in(self, t: torch.Tensor) -> torch.Tensor:
        """`.to(self.factor_dtype)` of base.py:349 / :363 for the 16-bit factor modes (None / float32: the tensor is
        read in place in whatever of fp32 / fp16 / bf16 it arrives in, which is exact)."""
        t = t.detach()
        if self.factor_dtype in (torch.bfloat16, torch.float16) and t.dtype != self.factor_dtype:
            return t.to(self.factor_dtype)
        return t

    def _round_inv(self, t: torch.Tensor) -> torch.Tensor:
        """Round second-order data to inv_dtype's values, in place (eigen.py:318-319,336-337; inverse.py:201,212)."""
        if self.inv_dtype != torch.float32:
            t.copy_(t.to(self.inv_dtype))
        return t

    def save_layer_input(self, input: list[torch.Tensor]) -> None:
        """Accumulate the A factor of this forward pass into the batch buffer (one fused launch)."""
        first = self._a_batch is None
        buf = self._store('a_batch')
        self.module.a_factor_into(self._cast_in(input[0]), buf, alpha=1.0, beta=0.0 if first else 1.0)
        self._a_batch = buf
        self._a_count = 1 if first else self._a_count + 1

    def save_layer_grad_output(self, grad_output: tuple[torch.Tensor, ...]) -> None:
        """Accumulate the G factor of this backward pass (unscaled by the AMP grad scaler, base.py:363-365)."""
        first = self._g_batch is None
        buf = self._store('g_batch')
        self.module.g_factor_into(self._cast_in(grad_output[0]), buf, alpha=1.0, beta=0.0 if first else 1.0,
                                  in_scale=self._g_in_scale())
        self._g_batch = buf
        self._g_count = 1 if first else self._g_count + 1

    def _running_average(self, which: str, batch: torch.Tensor, count: int, alpha: float) -> torch.Tensor:
        current = self.a_factor if which == 'a' else self.g_factor
        store = self._store(which)
        if current is None:
            ops.set_identity(store)          # first update starts from I (base.py:386-387)
        elif current is not store:
            store.copy_(current)
        ops.axpby(store, batch, (1.0 - alpha) / max(count, 1), alpha)
        return store

    def update_a_factor(self, alpha: float = 0.95) -> None:
        """A <- alpha A + (1 - alpha) mean(batch) (base.py:374-388)."""
        if self._a_batch is None:
            return
        self.a_factor = self._running_average('a', self._a_batch, self._a_count, alpha)
        self._a_batch = None

    def update_g_factor(self, alpha: float = 0.95) -> None:
        if self._g_batch is None:
            return
        self.g_factor = self._running_average('g', self._g_batch, self._g_count, alpha)
        self._g_batch = None

    def fused_a_update(self, input: list[torch.Tensor], alpha: float) -> None:
        """save_layer_input + update_a_factor in ONE launch (no micro-batch accumulation pending):
        A <- alpha A + (1 - alpha) cov in the SYRK epilogue."""
        assert self._a_batch is None
        store = self._store('a')
        current = self.a_factor
        if current is None:
            ops.set_identity(store)
        elif current is not store:
            store.copy_(current)
        self.module.a_factor_into(self._cast_in(input[0]), store, alpha=1.0 - alpha, beta=alpha)
        self.a_factor = store

    def fused_g_update(self, grad_output: tuple[torch.Tensor, ...], alpha: float) -> None:
        assert self._g_batch is None
        store = self._store('g')
        current = self.g_factor
        if current is None:
            ops.set_identity(store)
        elif current is not store:
            store.copy_(current)
        self.module.g_factor_into(self._cast_in(grad_output[0]), store, alpha=1.0 - alpha, beta=alpha,
                                  in_scale=self._g_in_scale())
        self.g_factor = store

    # ------------------------------------------------------------------ gradients (base.py:406-422)
    def _grad_shape(self) -> tuple[int, int]:
        return (self.module.g_factor_shape[0], self.module.a_factor_shape[0])

    # ---- grouped exchange (exchange.py): buffers that are views into per-(group, source) arenas
    def attach_grad_buffer(self, view: torch.Tensor) -> None:
        assert tuple(view.shape) == tuple(self._grad_shape())
        self._grad_store = view

    def adopt_grad(self) -> None:
        """After the grouped gradient broadcast: the (computed or received) gradient is in the arena view."""
        self.grad = self._grad_buffer()

    def _grad_buffer(self) -> torch.Tensor:
        if self._grad_store is None:
            self._grad_store = torch.empty(self._grad_shape(), dtype=torch.float32, device=self.module.device)
        return self._grad_store

    def _wb_grads(self) -> tuple[torch.Tensor, torch.Tensor | None]:
        w = self.module.get_weight_grad()
        if w is None:
            raise RuntimeError('module has no gradient: call backward() before preconditioning')
        if w.dtype != torch.float32:
            raise RuntimeError('the B200 kernels precondition float32 gradients')
        b = self.module.get_bias_grad() if self.module.has_bias() else None
        return w.contiguous(), (b.contiguous() if b is not None else None)

    def update_grad(self, scale: float | torch.Tensor | None = None) -> None:
        """Write scale * preconditioned gradient into weight.grad / bias.grad, in place.

        `scale` may be a device tensor (the KL-clip factor computed on the GPU): no host synchronisation.
        """
        grad = self.grad
        if grad is None:
            raise RuntimeError(
                'preconditionined gradient is None. This may be because update_grad() was called before '
                'preconditioned_grad()',
            )
        w, b = self.module.get_weight_grad(), (self.module.get_bias_grad() if self.module.has_bias() else None)
        if not w.is_contiguous() or (b is not None and not b.is_contiguous()):
            stacked = grad if scale is None else grad * scale
            self.module.set_grad(stacked)
        else:
            if scale is not None and not isinstance(scale, torch.Tensor):
                scale = torch.full((1,), float(scale), dtype=torch.float32, device=grad.device)
            ops.scale_write_grads(grad, scale, w, b)
        self.grad = None
