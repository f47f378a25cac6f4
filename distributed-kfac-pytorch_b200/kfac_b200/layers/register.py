"""Discovery of the modules K-FAC preconditions (kfac/layers/register.py:14-94)."""
from __future__ import annotations

import re
from typing import Any

import torch

from kfac_b200.layers.base import KFACBaseLayer
from kfac_b200.layers.modules import Conv2dModuleHelper
from kfac_b200.layers.modules import LinearModuleHelper
from kfac_b200.layers.modules import ModuleHelper

KNOWN_MODULES = {'linear', 'conv2d'}
LINEAR_TYPES: tuple[type[torch.nn.Module], ...] = (torch.nn.Linear,)
CONV2D_TYPES: tuple[type[torch.nn.Module], ...] = (torch.nn.Conv2d,)


def get_flattened_modules(root: torch.nn.Module) -> list[tuple[str, torch.nn.Module]]:
    """(name, module) for every leaf of the module tree."""
    return [(name, m) for name, m in root.named_modules() if len(list(m.children())) == 0]


def requires_grad(module: torch.nn.Module) -> bool:
    """False if any parameter of the module is frozen."""
    return all(p.requires_grad for p in module.parameters())


def get_module_helper(module: torch.nn.Module) -> ModuleHelper | None:
    if isinstance(module, LINEAR_TYPES):
        return LinearModuleHelper(module)
    if isinstance(module, CONV2D_TYPES):
        return Conv2dModuleHelper(module)
    return None


def any_match(query: str, patterns: list[str]) -> bool:
    """re.search of any pattern anywhere in the query."""
    return any(re.compile(p).search(query) for p in patterns)


def register_modules(
    model: torch.nn.Module,
    kfac_layer_type: type[KFACBaseLayer],
    skip_layers: list[str],
    **layer_kwargs: Any,
) -> dict[torch.nn.Module, tuple[str, KFACBaseLayer]]:
    """One KFAC layer per supported, trainable leaf module whose name / class name matches no skip pattern."""
    layers: dict[torch.nn.Module, tuple[str, KFACBaseLayer]] = {}
    for name, module in get_flattened_modules(model):
        if any_match(name, skip_layers) or any_match(module.__class__.__name__, skip_layers):
            continue
        if not requires_grad(module):
            continue
        helper = get_module_helper(module)
        if helper is None:
            continue
        assert module not in layers
        layers[module] = (name, kfac_layer_type(helper, **layer_kwargs))
    return layers
