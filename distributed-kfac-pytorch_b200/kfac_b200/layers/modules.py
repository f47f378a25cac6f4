"""Module helpers: the seam between a torch.nn module and the native factor / gradient kernels.

Interface as in the reference (kfac/layers/modules.py:13-97: a_factor_shape, g_factor_shape, device,
get_a_factor, get_g_factor, get_grad, get_weight_grad, get_bias_grad, has_bias, has_symmetric_factors,
set_grad).  The arithmetic is native: one fused launch per factor (bias column, spatial scaling, im2col and
optionally the running average inside the tcgen05 SYRK kernel) instead of pad/unfold/cat/matmul/transposes.
"""
from __future__ import annotations

from typing import cast

import torch

from kfac_b200 import ops


class ModuleHelper:
    """Base helper: shapes, gradient access and the factor entry points."""

    def __init__(self, module: torch.nn.Module):
        self.module = module

    def __repr__(self) -> str:
        return f'{self.__class__.__name__}({repr(self.module)})'

    @property
    def a_factor_shape(self) -> tuple[int, int]:
        raise NotImplementedError

    @property
    def g_factor_shape(self) -> tuple[int, int]:
        raise NotImplementedError

    @property
    def device(self) -> torch.device:
        return next(self.module.parameters()).device

    # ---- factors: out <- beta * out + alpha * factor(x); returns out (allocated when None)
    def a_factor_into(self, a: torch.Tensor, out: torch.Tensor | None = None, *, alpha: float = 1.0,
                      beta: float = 0.0) -> torch.Tensor:
        raise NotImplementedError

    def g_factor_into(self, g: torch.Tensor, out: torch.Tensor | None = None, *, alpha: float = 1.0,
                      beta: float = 0.0, in_scale: float = 1.0) -> torch.Tensor:
        raise NotImplementedError

    def get_a_factor(self, a: torch.Tensor) -> torch.Tensor:
        """A factor of this module for the forward input `a` (modules.py:48-50)."""
        return self.a_factor_into(a)

    def get_g_factor(self, g: torch.Tensor) -> torch.Tensor:
        """G factor of this module for the grad w.r.t. its output (modules.py:52-54)."""
        return self.g_factor_into(g)

    # ---- gradients
    def get_grad(self) -> torch.Tensor:
        """[weight.grad viewed (out, -1) | bias.grad] (modules.py:56-69). The native precondition kernel reads
        weight.grad / bias.grad in place, so this concatenation is only for API users."""
        g = cast(torch.Tensor, self.module.weight.grad)
        g = g.view(g.size(0), -1)
        if self.has_bias():
            g = torch.cat([g, self.module.bias.grad.view(-1, 1)], 1)  # type: ignore[union-attr]
        return g

    def get_bias_grad(self) -> torch.Tensor:
        return cast(torch.Tensor, self.module.bias.grad)

    def get_weight_grad(self) -> torch.Tensor:
        return cast(torch.Tensor, self.module.weight.grad)

    def has_bias(self) -> bool:
        return hasattr(self.module, 'bias') and self.module.bias is not None

    def has_symmetric_factors(self) -> bool:
        return True

    def set_grad(self, grad: torch.Tensor) -> None:
        """Assign new contiguous .grad tensors from the stacked gradient (modules.py:87-97)."""
        if self.has_bias():
            self.module.bias.grad = grad[:, -1:].reshape(self.get_bias_grad().size()).contiguous()
            self.module.weight.grad = grad[:, :-1].reshape(self.get_weight_grad().size()).contiguous()
        else:
            self.module.weight.grad = grad.reshape(self.get_weight_grad().size()).contiguous()

    def _new_factor(self, n: int, like: torch.Tensor) -> torch.Tensor:
        return torch.empty(n, n, dtype=torch.float32, device=like.device)


class LinearModuleHelper(ModuleHelper):
    """torch.nn.Linear (modules.py:100-141)."""

    @property
    def a_factor_shape(self) -> tuple[int, int]:
        x = self.module.weight.size(1) + int(self.has_bias())  # type: ignore[union-attr]
        return (x, x)

    @property
    def g_factor_shape(self) -> tuple[int, int]:
        x = self.module.weight.size(0)  # type: ignore[union-attr]
        return (x, x)

    def a_factor_into(self, a, out=None, *, alpha=1.0, beta=0.0):
        if out is None:
            out = self._new_factor(self.a_factor_shape[0], a)
        return ops.factor_linear(a, self.has_bias(), out, alpha=alpha, beta=beta)

    def g_factor_into(self, g, out=None, *, alpha=1.0, beta=0.0, in_scale=1.0):
        if out is None:
            out = self._new_factor(self.g_factor_shape[0], g)
        return ops.factor_linear(g, False, out, alpha=alpha, beta=beta, in_scale=in_scale)


class Conv2dModuleHelper(ModuleHelper):
    """torch.nn.Conv2d (modules.py:144-237); the im2col of _extract_patches happens inside the kernel's loader."""

    @property
    def a_factor_shape(self) -> tuple[int, int]:
        m = self.module
        x = m.in_channels * m.kernel_size[0] * m.kernel_size[1] + int(self.has_bias())  # type: ignore
        return (x, x)

    @property
    def g_factor_shape(self) -> tuple[int, int]:
        return (self.module.out_channels, self.module.out_channels)  # type: ignore

    def a_factor_into(self, a, out=None, *, alpha=1.0, beta=0.0):
        m = self.module
        if out is None:
            out = self._new_factor(self.a_factor_shape[0], a)
        return ops.factor_conv2d_a(a, tuple(m.kernel_size), tuple(m.stride), tuple(m.padding),  # type: ignore
                                   self.has_bias(), out, alpha=alpha, beta=beta)

    def g_factor_into(self, g, out=None, *, alpha=1.0, beta=0.0, in_scale=1.0):
        if out is None:
            out = self._new_factor(self.g_factor_shape[0], g)
        return ops.factor_conv2d_g(g, out, alpha=alpha, beta=beta, in_scale=in_scale)

    def get_grad(self) -> torch.Tensor:
        return super().get_grad()
