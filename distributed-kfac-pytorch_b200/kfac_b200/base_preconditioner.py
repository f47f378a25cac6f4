"""Orchestrator of the K-FAC step on B200: hooks (factor stage) and step() (second-order data, preconditioning,
KL-clip, gradient write-back).

API and control flow follow BaseKFACPreconditioner of the reference (kfac/base_preconditioner.py:21-477):
same constructor arguments and validation, same properties, state_dict / load_state_dict format,
memory_usage keys, hook points and cadence (`steps % factor_update_steps`, `steps % inv_update_steps`), same
layer order (`reversed(list(layers))`) so every rank enters every collective in the same order, and the private
attribute names (`_damping`, `_lr`, `_steps`, `_layers`, `_assignment`, `_tdc`, ...) that kfac/scheduler.py and
the reference's tests poke at.  What differs is how the work is issued:

* the factor hooks run one fused kernel per factor (and fold the running average into it when possible);
* all eigendecompositions a rank owns go to the GPU as ONE batched call;
* the KL-clip scalar stays on the device: no `.item()` per layer (the reference synchronises the host 1-2
  times per layer, base_preconditioner.py:428-431);
* preconditioned gradients are written into the existing `.grad` tensors.
"""
from __future__ import annotations

import logging
import warnings
from collections import defaultdict
from typing import Any
from typing import Callable

import torch

from kfac_b200 import ops
from kfac_b200.assignment import WorkAssignment
from kfac_b200.distributed import get_rank
from kfac_b200.distributed import TorchDistributedCommunicator
from kfac_b200.layers.base import KFACBaseLayer

logger = logging.getLogger(__name__)


class BaseKFACPreconditioner:
    """Distributed K-FAC gradient preconditioner over a dict of registered layers."""

    def __init__(
        self,
        layers: dict[torch.nn.Module, tuple[str, KFACBaseLayer]],
        *,
        assignment: WorkAssignment,
        tdc: TorchDistributedCommunicator,
        factor_update_steps: Callable[[int], int] | int = 1,
        inv_update_steps: Callable[[int], int] | int = 1,
        damping: Callable[[int], float] | float = 0.001,
        factor_decay: Callable[[int], float] | float = 0.95,
        kl_clip: Callable[[int], float] | float = 0.001,
        lr: Callable[[int], float] | float = 0.1,
        accumulation_steps: int = 1,
        update_factors_in_hook: bool = True,
        defaults: dict[str, Any] | None = None,
        loglevel: int = logging.DEBUG,
    ) -> None:
        # validation: kfac/base_preconditioner.py:86-108
        if not callable(factor_update_steps) and not 0 < factor_update_steps:
            raise ValueError('factor_update_steps must be > 0')
        if not callable(inv_update_steps) and not 0 < inv_update_steps:
            raise ValueError('inv_update_steps must be > 0')
        if not callable(damping) and not 0.0 < damping:
            raise ValueError('damping must be > 0')
        if not callable(factor_decay) and not 0.0 < factor_decay <= 1:
            raise ValueError('factor_decay must be in (0, 1]')
        if not callable(kl_clip) and not 0.0 < kl_clip:
            raise ValueError('kl_clip must be > 0')
        if not callable(lr) and not 0.0 <= lr:
            raise ValueError('lr be > 0')
        if not 0 < accumulation_steps:
            raise ValueError('accumulation_steps must be > 0')
        if (
            not callable(inv_update_steps)
            and not callable(factor_update_steps)
            and not 0 == inv_update_steps % factor_update_steps
        ):
            warnings.warn('It is suggested that inv_update_steps be an integer multiple of factor_update_steps')

        self._accumulation_steps = accumulation_steps
        self._assignment = assignment
        self._damping = damping
        self._defaults = defaults
        self._factor_decay = factor_decay
        self._factor_update_steps = factor_update_steps
        self._inv_update_steps = inv_update_steps
        self._kl_clip = kl_clip
        self._layers = layers
        self._loglevel = loglevel
        self._lr = lr
        self._tdc = tdc
        self._update_factors_in_hook = update_factors_in_hook

        self._steps = 0
        self._mini_steps: dict[str, int] = defaultdict(int)
        # optional observer called with 'second_order' / 'precondition' / 'end' at step()'s stage boundaries
        # (bench.py records CUDA events there); never set in normal use
        self._stage_hook: Callable[[str], None] | None = None
        self._vg: torch.Tensor | None = None      # device accumulator of sum(precond * grad)
        self._scale: torch.Tensor | None = None   # device KL-clip factor of the last step

        self._allocate_factor_arena()
        self._exchange = None           # grouped KAISA broadcasts (exchange.py); None = the per-tensor path
        self._factor_pg = None          # dedicated process group for the factor all-reduce
        self._setup_communication()
        for module in self._layers:
            module.register_forward_pre_hook(self._save_input)
            module.register_full_backward_hook(self._save_grad_output)

    # ------------------------------------------------------------------ storage
    def _allocate_factor_arena(self) -> None:
        """One flat fp32 arena per device for all running averages, laid out in hook order (A factors in
        registration order, then G factors in reverse order) so that consecutive reduce_*_factor calls see
        memory-adjacent tensors and travel as one collective."""
        by_device: dict[torch.device, list[tuple[KFACBaseLayer, str, int]]] = defaultdict(list)
        entries = list(self._layers.values())
        for _, layer in entries:
            by_device[layer.module.device].append((layer, 'a', layer.module.a_factor_shape[0]))
        for _, layer in reversed(entries):
            by_device[layer.module.device].append((layer, 'g', layer.module.g_factor_shape[0]))
        self._arenas: dict[torch.device, torch.Tensor] = {}
        for device, items in by_device.items():
            total = sum(n * n for _, _, n in items)
            arena = torch.empty(total, dtype=torch.float32, device=device)
            self._arenas[device] = arena
            views: dict[int, dict[str, torch.Tensor]] = defaultdict(dict)
            off = 0
            for layer, which, n in items:
                views[id(layer)][which] = arena[off:off + n * n].view(n, n)
                off += n * n
            for layer, which, _ in items:
                if which == 'g':
                    layer.attach_storage(views[id(layer)]['a'], views[id(layer)]['g'])

    def _setup_communication(self) -> None:
        """World > 1 only.  (1) The factor all-reduce gets its own process group -- with NCCL its own communicator and
        stream -- so that the chunks launched from inside the hooks (base_preconditioner.py:448-455,472-477) run
        beside DistributedDataParallel's gradient buckets, which live on the default group, instead of queueing
        behind them on one stream.  (2) Second-order data and preconditioned gradients are exchanged through
        per-(group, source) arenas: one broadcast each instead of one per tensor."""
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
            return
        self._factor_pg = dist.new_group(list(range(dist.get_world_size())))   # collective: every rank gets here
        entries = list(self._layers.values())
        layers = {name: layer for name, layer in entries}
        if not layers or not all(hasattr(l, 'second_order_payload') for l in layers.values()):
            return
        if any(not l.second_order_payload() for l in layers.values()):
            return
        devices = {l.module.device for l in layers.values()}
        if len(devices) != 1:
            return
        from kfac_b200.exchange import GroupedExchange
        self._exchange = GroupedExchange(layers, self._assignment, self._tdc, get_rank(), devices.pop())

    def _factor_group(self, name: str, factor: str) -> Any:
        group = self._assignment.factor_group(name, factor)
        return self._factor_pg if group is None and self._factor_pg is not None else group

    def __repr__(self) -> str:
        params = [
            ('accumulation_steps', self._accumulation_steps),
            ('assignment', self._assignment.__class__.__name__),
            ('damping', self._damping),
            ('factor_decay', self._factor_decay),
            ('factor_update_steps', self._factor_update_steps),
            ('inv_update_steps', self._inv_update_steps),
            ('kl_clip', self._kl_clip),
            ('layers', len(self._layers)),
            ('loglevel', self._loglevel),
            ('lr', self._lr),
            ('steps', self.steps),
            ('update_factors_in_hook', self._update_factors_in_hook),
        ]
        if self._defaults is not None:
            params.extend(list(self._defaults.items()))
        body = '\n'.join(f'  {name}={value},' for name, value in sorted(params, key=lambda x: x[0]))
        return f'{self.__class__.__name__}(\n{body}\n)'

    # ------------------------------------------------------------------ hyper-parameters (may be callables of step)
    def _value(self, v: Any) -> Any:
        return v(self.steps) if callable(v) else v

    @property
    def damping(self) -> float:
        return self._value(self._damping)

    @property
    def factor_decay(self) -> float:
        return self._value(self._factor_decay)

    @property
    def kl_clip(self) -> float:
        return self._value(self._kl_clip)

    @property
    def lr(self) -> float:
        return self._value(self._lr)

    @property
    def factor_update_steps(self) -> int:
        return self._value(self._factor_update_steps)

    @property
    def inv_update_steps(self) -> int:
        return self._value(self._inv_update_steps)

    @property
    def steps(self) -> int:
        return self._steps

    # ------------------------------------------------------------------ checkpoint (base_preconditioner.py:213-306)
    def state_dict(self, include_factors: bool = True) -> dict[str, Any]:
        """{'steps', non-callable hyper-parameters, 'layers': {name: {'A', 'G'}}} -- the reference's format."""
        state: dict[str, Any] = {'steps': self.steps}
        for key in ('factor_update_steps', 'inv_update_steps', 'damping', 'factor_decay', 'kl_clip', 'lr'):
            value = getattr(self, '_' + key)
            if not callable(value):
                state[key] = value
        if include_factors:
            self._tdc.flush_allreduce_buckets()
            state['layers'] = {name: layer.state_dict() for name, layer in self._layers.values()}
        return state

    def load_state_dict(self, state_dict: dict[str, Any], compute_inverses: bool = True) -> None:
        self._steps = state_dict['steps']
        for key in ('factor_update_steps', 'inv_update_steps', 'damping', 'factor_decay', 'kl_clip', 'lr'):
            if key in state_dict:
                setattr(self, '_' + key, state_dict[key])
        if 'layers' in state_dict:
            if len(state_dict['layers']) != len(self._layers):
                raise ValueError('loaded state dict contains a different number of layers')
            by_name = {name: layer for name, layer in self._layers.values()}
            for found_name, layer_state in state_dict['layers'].items():
                if found_name in by_name:
                    by_name[found_name].load_state_dict(layer_state)
        elif compute_inverses:
            warnings.warn(
                'Layer factors are not included in the state_dict so inverses cannot be computed. '
                'Skipping inverse computation.',
            )
            compute_inverses = False
        if compute_inverses:
            # every rank recomputes everything (base_preconditioner.py:294-306)
            self._second_order(everywhere=True)

    # ------------------------------------------------------------------ step
    def _second_order(self, everywhere: bool = False) -> None:
        """Compute (batched) and broadcast the second-order data, in the reference's per-layer order."""
        rank = get_rank()
        order = list(reversed(list(self._layers.values()))) if not everywhere else list(self._layers.values())
        damping = self.damping
        mine_a = [(n, l) for n, l in order if everywhere or rank == self._assignment.inv_worker(n, 'A')]
        mine_g = [(n, l) for n, l in order if everywhere or rank == self._assignment.inv_worker(n, 'G')]
        batched = all(hasattr(l, 'eigh_request') for _, l in order)
        if not batched and all(hasattr(l, 'inverse_request') for _, l in order):
            # inverse method: every factor this rank owns in ONE asynchronous Newton-Schulz call; its per-matrix
            # status stays on the device and is looked at when the next step starts (_check_inverse_info)
            batched = True
            reqs = [l.inverse_request('A') for _, l in mine_a] + [l.inverse_request('G') for _, l in mine_g]
            if reqs:
                _, info = ops.inverse_newton_schulz_batched([r[0] for r in reqs], damping, outs=[r[1] for r in reqs])
                names = [f'{n} (A)' for n, _ in mine_a] + [f'{n} (G)' for n, _ in mine_g]
                self._inverse_info = (info, names)
            for _, l in mine_a:
                l.inverse_adopt('A')
            for _, l in mine_g:
                l.inverse_adopt('G')
        elif batched:
            reqs = [l.eigh_request('A') for _, l in mine_a] + [l.eigh_request('G') for _, l in mine_g]
            if reqs:
                ops.eigh_batched([r[0] for r in reqs], clamp_nonneg=True, qs=[r[1] for r in reqs],
                                 ds=[r[2] for r in reqs])
            for _, l in mine_a:
                l.eigh_adopt('A', damping)
            for _, l in mine_g:
                l.eigh_adopt('G', damping)
        if batched and self._exchange is not None:
            if self._assignment.broadcast_inverses():
                self._exchange.broadcast_second_order()
            return
        for name, layer in order:
            if not batched and (everywhere or rank == self._assignment.inv_worker(name, 'A')):
                layer.compute_a_inv(damping=damping)
            if self._assignment.broadcast_inverses() and (everywhere or self._assignment.is_grad_worker(name)):
                layer.broadcast_a_inv(src=self._assignment.inv_worker(name, 'A'),
                                      group=self._assignment.grad_worker_group(name))
            if not batched and (everywhere or rank == self._assignment.inv_worker(name, 'G')):
                layer.compute_g_inv(damping=damping)
            if self._assignment.broadcast_inverses() and (everywhere or self._assignment.is_grad_worker(name)):
                layer.broadcast_g_inv(src=self._assignment.inv_worker(name, 'G'),
                                      group=self._assignment.grad_worker_group(name))

    def _check_inverse_info(self) -> None:
        """Status of the last batched Newton-Schulz call.  It was written by the device long ago (a whole training
        iteration has passed), so reading it costs no stall.  torch.linalg.inv raises for a singular factor
        (kfac/layers/inverse.py:201); the same RuntimeError surfaces here, one step late."""
        pending = getattr(self, '_inverse_info', None)
        if pending is None:
            return
        self._inverse_info = None
        info, names = pending
        host = info.cpu()
        bad = [(names[i], int(host[i, 0]), int(host[i, 1])) for i in range(host.shape[0]) if int(host[i, 1]) != 0]
        if bad:
            what = ', '.join(f'{n}: {"non-finite iterate" if st == 2 else "no convergence"} after {it} iterations'
                             for n, it, st in bad)
            raise RuntimeError(f'Newton-Schulz inverse failed for {what}')

    @torch.no_grad()
    def step(self) -> None:
        """One K-FAC step: call after backward() (gradients already averaged over ranks) and before
        optimizer.step(); rewrites the gradients of the registered modules in place."""
        self._check_inverse_info()
        if not self._update_factors_in_hook and self.steps % self.factor_update_steps == 0:
            for name, layer in reversed(list(self._layers.values())):
                self._mini_steps[name] = 0
                layer.update_a_factor(alpha=self.factor_decay)
                layer.reduce_a_factor(self._factor_group(name, 'A'))
                layer.update_g_factor(alpha=self.factor_decay)
                layer.reduce_g_factor(self._factor_group(name, 'G'))
        self._tdc.flush_allreduce_buckets()

        if self._stage_hook is not None:
            self._stage_hook('second_order')
        if self.steps % self.inv_update_steps == 0:
            self._second_order()
            self._tdc.flush_allreduce_buckets()
        if self._stage_hook is not None:
            self._stage_hook('precondition')

        kl_clip = self.kl_clip
        order = list(reversed(list(self._layers.values())))
        vg = None
        if kl_clip is not None and order:
            vg = self._vg_buffer(order[0][1].module.device)
            vg.zero_()
        received: list[KFACBaseLayer] = []
        workers = [(n, l) for n, l in order if self._assignment.is_grad_worker(n)]
        grouped = bool(workers) and all(hasattr(l, 'precondition_request') for _, l in workers)
        if grouped:
            # every layer this rank preconditions, in four grouped launches (one per GEMM stage)
            reqs = [l.precondition_request() for _, l in workers]
            if 'a_inv' in reqs[0]:
                ops.precondition_inverse_grouped(reqs, vg_accum=vg)
            else:
                ops.precondition_eigen_grouped(reqs, damping=self.damping, vg_accum=vg)
            for _, l in workers:
                l.precondition_adopt()
        for name, layer in order:
            worker = self._assignment.is_grad_worker(name)
            if worker and not grouped:
                layer.preconditioned_grad(damping=self.damping, vg_accum=vg)
            elif not worker and vg is not None:
                received.append(layer)
            if self._assignment.broadcast_gradients() and self._exchange is None:
                layer.broadcast_grad(src=self._assignment.src_grad_worker(name),
                                     group=self._assignment.grad_receiver_group(name))
        if self._assignment.broadcast_gradients() and self._exchange is not None:
            self._exchange.broadcast_gradients()
        self._tdc.flush_allreduce_buckets()

        scale = None
        if vg is not None:
            for layer in received:  # gradients that arrived by broadcast contribute to the clip too
                w = layer.module.get_weight_grad()
                b = layer.module.get_bias_grad() if layer.module.has_bias() else None
                ops.grad_dot(layer.grad, w.contiguous(), b, vg)
            scale = self._scale_buffer(vg.device)
            ops.klclip_scale(vg, float(self.lr), float(kl_clip), scale)
        self._scale = scale

        # gradient write-back: every layer whose .grad tensors are contiguous in ONE launch, the rest one by one
        grouped_items = []
        for _, layer in order:
            w = layer.module.get_weight_grad()
            b = layer.module.get_bias_grad() if layer.module.has_bias() else None
            if (isinstance(layer.grad, torch.Tensor) and w is not None and w.is_contiguous()
                    and (b is None or b.is_contiguous()) and (scale is None or isinstance(scale, torch.Tensor))):
                grouped_items.append((layer.grad, w, b))
                layer.grad = None
            else:
                layer.update_grad(scale=scale)
        ops.scale_write_grads_grouped(grouped_items, scale)
        if self._stage_hook is not None:
            self._stage_hook('end')

        self._steps += 1
        self._mini_steps = defaultdict(int)

    def _vg_buffer(self, device: torch.device) -> torch.Tensor:
        if self._vg is None or self._vg.device != device:
            self._vg = torch.zeros(1, dtype=torch.float64, device=device)
        return self._vg

    def _scale_buffer(self, device: torch.device) -> torch.Tensor:
        if self._scale is None or self._scale.device != device:
            self._scale = torch.ones(1, dtype=torch.float32, device=device)
        return self._scale

    def reset_batch(self) -> None:
        for _, layer in self._layers.values():
            layer.reset_batch()

    def memory_usage(self) -> dict[str, int]:
        """Bytes per category over all layers on this rank + 'total' (base_preconditioner.py:387-407)."""
        sizes: dict[str, int] = defaultdict(int)
        self._tdc.flush_allreduce_buckets()
        for _, layer in self._layers.values():
            for key, size in layer.memory_usage().items():
                sizes[key] += size
        sizes['total'] = sum(sizes.values())
        return sizes

    def _compute_grad_scale(self) -> float:
        """KL-clip factor of the current preconditioned gradients as a host float
        (base_preconditioner.py:409-433).  step() does not call this: it keeps the factor on the device."""
        order = list(reversed(list(self._layers.values())))
        if not order:
            return 1.0
        vg = torch.zeros(1, dtype=torch.float64, device=order[0][1].module.device)
        for _, layer in order:
            if layer.grad is None:
                raise AssertionError('layer gradient has not been preconditioned')
            w = layer.module.get_weight_grad()
            b = layer.module.get_bias_grad() if layer.module.has_bias() else None
            ops.grad_dot(layer.grad, w.contiguous(), b, vg)
        scale = torch.ones(1, dtype=torch.float32, device=vg.device)
        ops.klclip_scale(vg, float(self.lr), float(self.kl_clip), scale)
        return float(scale.item())

    # ------------------------------------------------------------------ hooks (base_preconditioner.py:435-477)
    @torch.no_grad()
    def _save_input(self, module: torch.nn.Module, input: list[torch.Tensor]) -> None:
        if not module.training:
            return
        if self.steps % self.factor_update_steps == 0:
            name, layer = self._layers[module]
            self._mini_steps[name] += 1
            in_hook = self._update_factors_in_hook and self._mini_steps[name] % self._accumulation_steps == 0
            if in_hook and layer._a_batch is None:
                layer.fused_a_update(input, alpha=self.factor_decay)
            else:
                layer.save_layer_input(input)
                if in_hook:
                    layer.update_a_factor(alpha=self.factor_decay)
            if in_hook:
                layer.reduce_a_factor(self._factor_group(name, 'A'))

    @torch.no_grad()
    def _save_grad_output(
        self,
        module: torch.nn.Module,
        grad_input: tuple[torch.Tensor, ...] | torch.Tensor,
        grad_output: tuple[torch.Tensor, ...] | torch.Tensor,
    ) -> None:
        if not module.training:
            return
        if self.steps % self.factor_update_steps == 0:
            name, layer = self._layers[module]
            if isinstance(grad_output, torch.Tensor):
                grad_output = (grad_output,)
            in_hook = self._update_factors_in_hook and self._mini_steps[name] % self._accumulation_steps == 0
            if in_hook and layer._g_batch is None:
                layer.fused_g_update(grad_output, alpha=self.factor_decay)
            else:
                layer.save_layer_grad_output(grad_output)
                if in_hook:
                    layer.update_g_factor(alpha=self.factor_decay)
            if in_hook:
                layer.reduce_g_factor(self._factor_group(name, 'G'))
