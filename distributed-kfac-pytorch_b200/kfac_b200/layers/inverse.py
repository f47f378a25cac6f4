"""Explicit-inverse K-FAC layer: Newton-Schulz on the tensor cores instead of an LU inverse.

Mirrors KFACInverseLayer (kfac/layers/inverse.py:18-233): attributes a_inv / g_inv, same errors and broadcasts.
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
from kfac_b200.layers.base import KFACBaseLayer
from kfac_b200.layers.modules import ModuleHelper


class KFACInverseLayer(KFACBaseLayer):
    """Preconditions with (G + damping I)^-1 grad (A + damping I)^-1."""

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
        super().__init__(module=module, tdc=tdc, allreduce_method=allreduce_method, factor_dtype=factor_dtype,
                         grad_scaler=grad_scaler, inv_dtype=inv_dtype, symmetry_aware=symmetry_aware)
        self._a_inv: torch.Tensor | Pending | None = None
        self._g_inv: torch.Tensor | Pending | None = None
        self._bufs: dict[str, torch.Tensor] = {}
        self.ns_iterations = {'A': 0, 'G': 0}

    def _buf(self, name: str, n: int) -> torch.Tensor:
        b = self._bufs.get(name)
        if b is None:
            b = torch.empty(n, n, dtype=torch.float32, device=self.module.device)
            self._bufs[name] = b
        return b

    @property
    def a_inv(self) -> torch.Tensor | None:
        self._a_inv = resolve(self._a_inv)
        return self._a_inv  # type: ignore[return-value]

    @a_inv.setter
    def a_inv(self, value: torch.Tensor | Pending | None) -> None:
        self._a_inv = value

    @property
    def g_inv(self) -> torch.Tensor | None:
        self._g_inv = resolve(self._g_inv)
        return self._g_inv  # type: ignore[return-value]

    @g_inv.setter
    def g_inv(self, value: torch.Tensor | Pending | None) -> None:
        self._g_inv = value

    def memory_usage(self) -> dict[str, int]:
        def size(t):
            return t.nelement() * t.element_size() if t is not None else 0
        sizes = super().memory_usage()
        sizes['a_inverses'] = size(self.a_inv)
        sizes['g_inverses'] = size(self.g_inv)
        return sizes

    def broadcast_a_inv(self, src: int, group: Any = None) -> None:
        if self.a_inv is None:
            if get_rank() == src:
                raise RuntimeError(
                    f'Attempt to broadcast A inv from src={src} but this rank has not computed A inv yet.',
                )
            self.a_inv = self._buf('a', self.module.a_factor_shape[0])
        self.a_inv = self.tdc.broadcast(self.a_inv, src=src, group=group,
                                        symmetric=self.symmetric_factors and self.symmetry_aware)

    def broadcast_g_inv(self, src: int, group: Any = None) -> None:
        if self.g_inv is None:
            if get_rank() == src:
                raise RuntimeError(
                    f'Attempt to broadcast G inv from src={src} but this rank has not computed G inv yet.',
                )
            self.g_inv = self._buf('g', self.module.g_factor_shape[0])
        self.g_inv = self.tdc.broadcast(self.g_inv, src=src, group=group,
                                        symmetric=self.symmetric_factors and self.symmetry_aware)

    def compute_a_inv(self, damping: float = 0.001) -> None:
        """(A + damping I)^-1 (inverse.py:185-201)."""
        if self.a_factor is None:
            raise RuntimeError('Cannot invert A before A has been computed')
        out, iters = ops.inverse_newton_schulz(self.a_factor, damping, out=self._buf('a', self.a_factor.shape[0]))
        self.a_inv = self._round_inv(out)
        self.ns_iterations['A'] = iters

    def compute_g_inv(self, damping: float = 0.001) -> None:
        if self.g_factor is None:
            raise RuntimeError('Cannot invert G before G has been computed')
        out, iters = ops.inverse_newton_schulz(self.g_factor, damping, out=self._buf('g', self.g_factor.shape[0]))
        self.g_inv = self._round_inv(out)
        self.ns_iterations['G'] = iters

    # ---- grouped exchange (exchange.py)
    def second_order_payload(self) -> dict[str, list[tuple[str, tuple[int, ...]]]]:
        """A^-1 and G^-1 travel whole (inverse.py:111-183).  With symmetry_aware the reference packs the upper
        triangle per tensor; that per-tensor path is kept for it, so no grouped payload is declared."""
        if self.symmetric_factors and self.symmetry_aware:
            return {}
        a, g = self.module.a_factor_shape[0], self.module.g_factor_shape[0]
        return {'A': [('a', (a, a))], 'G': [('g', (g, g))]}

    def attach_second_order_buffers(self, views: dict[str, torch.Tensor]) -> None:
        self._bufs.update(views)

    def adopt_second_order(self, which: str) -> None:
        if which == 'A':
            self.a_inv = self._bufs['a']
        else:
            self.g_inv = self._bufs['g']

    # ---- batched path: the preconditioner collects one request per factor it owns, runs ONE asynchronous
    # Newton-Schulz call over all of them (ops.inverse_newton_schulz_batched) and hands the results back
    def inverse_request(self, which: str) -> tuple[torch.Tensor, torch.Tensor]:
        factor = self.a_factor if which == 'A' else self.g_factor
        if factor is None:
            raise RuntimeError(f'Cannot invert {which} before {which} has been computed')
        return factor, self._buf(which.lower(), factor.shape[0])

    def inverse_adopt(self, which: str) -> None:
        if which == 'A':
            self.a_inv = self._round_inv(self._bufs['a'])
        else:
            self.g_inv = self._round_inv(self._bufs['g'])

    def precondition_request(self) -> dict:
        """Pointers for the grouped precondition call (ops.precondition_inverse_grouped)."""
        if self.a_inv is None or self.g_inv is None:
            raise RuntimeError('Cannot precondition gradient before A and G have been inverted')
        w, b = self._wb_grads()
        return {'w_grad': w.view(w.shape[0], -1), 'b_grad': b, 'a_inv': self.a_inv, 'g_inv': self.g_inv,
                'out': self._grad_buffer()}

    def precondition_adopt(self) -> None:
        self.grad = self._grad_buffer()

    def preconditioned_grad(self, damping: float = 0.001, vg_accum: torch.Tensor | None = None) -> None:
        """G^-1 grad A^-1 (inverse.py:214-233)."""
        if self.a_inv is None or self.g_inv is None:
            raise RuntimeError('Cannot precondition gradient before A and G have been inverted')
        w, b = self._wb_grads()
        self.grad = ops.precondition_inverse(w, b, self.a_inv, self.g_inv, out=self._grad_buffer(),
                                             vg_accum=vg_accum)
