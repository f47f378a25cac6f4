"""Eigendecomposition-based K-FAC layer over the native eigensolver and the fused precondition pipeline.

Mirrors KFACEigenLayer (kfac/layers/eigen.py:18-384): attributes qa, qg, da, dg, dgda, `prediv_eigenvalues`,
the same RuntimeErrors when stages are called out of order, the same broadcast payloads.
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


class KFACEigenLayer(KFACBaseLayer):
    """Preconditions with F^-1 ~ Q (d + damping)^-1 Q^T on both Kronecker factors."""

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
        prediv_eigenvalues: bool = False,
    ) -> None:
        super().__init__(module=module, tdc=tdc, allreduce_method=allreduce_method, factor_dtype=factor_dtype,
                         grad_scaler=grad_scaler, inv_dtype=inv_dtype, symmetry_aware=symmetry_aware)
        self.prediv_eigenvalues = prediv_eigenvalues
        self._qa: torch.Tensor | Pending | None = None
        self._qg: torch.Tensor | Pending | None = None
        self._da: torch.Tensor | Pending | None = None
        self._dg: torch.Tensor | Pending | None = None
        self._dgda: torch.Tensor | Pending | None = None
        self._bufs: dict[str, torch.Tensor] = {}

    def _buf(self, name: str, shape: tuple[int, ...]) -> torch.Tensor:
        b = self._bufs.get(name)
        if b is None:
            b = torch.empty(shape, dtype=torch.float32, device=self.module.device)
            self._bufs[name] = b
        return b

    def _prop(name: str):  # noqa: N805
        attr = '_' + name

        def getter(self):
            setattr(self, attr, resolve(getattr(self, attr)))
            return getattr(self, attr)

        def setter(self, value):
            setattr(self, attr, value)

        return property(getter, setter)

    qa = _prop('qa')
    qg = _prop('qg')
    da = _prop('da')
    dg = _prop('dg')
    dgda = _prop('dgda')
    del _prop

    def memory_usage(self) -> dict[str, int]:
        def size(t):
            return t.nelement() * t.element_size() if t is not None else 0
        sizes = super().memory_usage()
        sizes['a_inverses'] = size(self.qa) + size(self.da)
        sizes['g_inverses'] = size(self.qg) + size(self.dg) + size(self.dgda)
        return sizes

    # ------------------------------------------------------------------ broadcasts (eigen.py:177-292)
    def broadcast_a_inv(self, src: int, group: Any = None) -> None:
        if self.qa is None or (not self.prediv_eigenvalues and self.da is None):
            if get_rank() == src:
                raise RuntimeError(
                    f'Attempt to broadcast A inv from src={src} but this rank has not computed A inv yet.',
                )
            n = self.module.a_factor_shape[0]
            self.qa = self._buf('qa', (n, n))
            self.da = self._buf('da', (n,))
        self.qa = self.tdc.broadcast(self.qa, src=src, group=group)
        if not self.prediv_eigenvalues:
            self.da = self.tdc.broadcast(self.da, src=src, group=group)

    def broadcast_g_inv(self, src: int, group: Any = None) -> None:
        if (
            self.qg is None
            or (not self.prediv_eigenvalues and self.dg is None)
            or (self.prediv_eigenvalues and self.dgda is None)
        ):
            if get_rank() == src:
                raise RuntimeError(
                    f'Attempt to broadcast G inv from src={src} but this rank has not computed G inv yet.',
                )
            g, a = self.module.g_factor_shape[0], self.module.a_factor_shape[0]
            self.qg = self._buf('qg', (g, g))
            if not self.prediv_eigenvalues:
                self.dg = self._buf('dg', (g,))
            else:
                self.dgda = self._buf('dgda', (g, a))
        self.qg = self.tdc.broadcast(self.qg, src=src, group=group)
        if not self.prediv_eigenvalues:
            self.dg = self.tdc.broadcast(self.dg, src=src, group=group)
        else:
            self.dgda = self.tdc.broadcast(self.dgda, src=src, group=group)

    # ------------------------------------------------------------------ grouped exchange (exchange.py)
    def second_order_payload(self) -> dict[str, list[tuple[str, tuple[int, ...]]]]:
        """What the inverse worker of each factor sends to the gradient workers (eigen.py:214-225, 274-292):
        A: qa (+ da unless the eigenvalue outer product is precomputed); G: qg + (dgda | dg)."""
        a, g = self.module.a_factor_shape[0], self.module.g_factor_shape[0]
        if self.prediv_eigenvalues:
            return {'A': [('qa', (a, a))], 'G': [('qg', (g, g)), ('dgda', (g, a))]}
        return {'A': [('qa', (a, a)), ('da', (a,))], 'G': [('qg', (g, g)), ('dg', (g,))]}

    def attach_second_order_buffers(self, views: dict[str, torch.Tensor]) -> None:
        self._bufs.update(views)

    def adopt_second_order(self, which: str) -> None:
        """After the grouped broadcast: point the public attributes at the arena views (source and receivers alike)."""
        if which == 'A':
            self.qa = self._bufs['qa']
            if not self.prediv_eigenvalues:
                self.da = self._bufs['da']
            return
        self.qg = self._bufs['qg']
        if self.prediv_eigenvalues:
            self.dgda = self._bufs['dgda']
        else:
            self.dg = self._bufs['dg']

    # ------------------------------------------------------------------ eigendecompositions (eigen.py:294-347)
    def eigh_request(self, which: str) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        """(factor, Q buffer, d buffer) for the batched solver; results are adopted by `eigh_adopt`."""
        factor = self.a_factor if which == 'A' else self.g_factor
        if not isinstance(factor, torch.Tensor):
            raise RuntimeError(f'Cannot eigendecompose {which} before {which} has been computed')
        n = factor.shape[0]
        low = which.lower()
        return factor, self._buf('q' + low, (n, n)), self._buf('d' + low, (n,))

    def eigh_adopt(self, which: str, damping: float) -> None:
        """Take the solver outputs written into this layer's buffers as qa/da or qg/dg (+ dgda)."""
        if which == 'A':
            self.qa, self.da = self._round_inv(self._bufs['qa']), self._round_inv(self._bufs['da'])
            return
        self.qg, self.dg = self._round_inv(self._bufs['qg']), self._round_inv(self._bufs['dg'])
        if self.prediv_eigenvalues:
            if self.da is None:
                raise AssertionError('compute_g_inv needs the eigenvalues of A (eigen.py:340)')
            g, a = self.module.g_factor_shape[0], self.module.a_factor_shape[0]
            self.dgda = self._round_inv(ops.eigen_outer(self.dg, self.da, damping, out=self._buf('dgda', (g, a))))
            self.dg = None
            self.da = None

    def compute_a_inv(self, damping: float = 0.001) -> None:
        f, q, d = self.eigh_request('A')
        ops.eigh_batched([f], clamp_nonneg=True, qs=[q], ds=[d])
        self.eigh_adopt('A', damping)

    def compute_g_inv(self, damping: float = 0.001) -> None:
        f, q, d = self.eigh_request('G')
        ops.eigh_batched([f], clamp_nonneg=True, qs=[q], ds=[d])
        self.eigh_adopt('G', damping)

    # ------------------------------------------------------------------ precondition (eigen.py:349-384)
    def precondition_request(self) -> dict:
        """Pointers for the grouped precondition call; the result lands in this layer's gradient buffer and is
        adopted by `precondition_adopt`."""
        if (
            self.qa is None
            or self.qg is None
            or (not self.prediv_eigenvalues and (self.da is None or self.dg is None))
            or (self.prediv_eigenvalues and self.dgda is None)
        ):
            raise RuntimeError('Eigendecompositions for both A and G have not been computed')
        w, b = self._wb_grads()
        return {'w_grad': w.view(w.shape[0], -1), 'b_grad': b, 'qa': self.qa, 'qg': self.qg,
                'da': None if self.prediv_eigenvalues else self.da, 'dg': None if self.prediv_eigenvalues else self.dg,
                'dgda': self.dgda if self.prediv_eigenvalues else None, 'out': self._grad_buffer()}

    def precondition_adopt(self) -> None:
        self.grad = self._grad_buffer()

    def preconditioned_grad(self, damping: float = 0.001, vg_accum: torch.Tensor | None = None) -> None:
        if (
            self.qa is None
            or self.qg is None
            or (not self.prediv_eigenvalues and (self.da is None or self.dg is None))
            or (self.prediv_eigenvalues and self.dgda is None)
        ):
            raise RuntimeError('Eigendecompositions for both A and G have not been computed')
        w, b = self._wb_grads()
        self.grad = ops.precondition_eigen(
            w, b, self.qa, self.qg, da=self.da, dg=self.dg, dgda=self.dgda if self.prediv_eigenvalues else None,
            damping=damping, out=self._grad_buffer(), vg_accum=vg_accum,
        )
