"""Host-side contract tests that need no GPU: the C ABI exports what the header declares, the public
constructor validates like the reference, and nothing silently falls back to the CPU."""
from __future__ import annotations

import ctypes
import os
import re
import warnings

import pytest
import torch

import kfac_b200
from kfac_b200 import _native
from kfac_b200 import ops
from kfac_b200.enums import AssignmentStrategy
from kfac_b200.enums import ComputeMethod
from kfac_b200.enums import DistributedStrategy
from kfac_b200.preconditioner import KFACPreconditioner

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, 'include', 'kfac_b200.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return set(re.findall(r'\b(kfac_\w+)\s*\(', text))


def test_library_exports_every_declared_symbol():
    """Loads libkfac_b200.so (no compute calls) and resolves every entry point of include/kfac_b200.h."""
    assert os.path.exists(_native.LIB_PATH), 'run `python __graft_entry__.py build` first'
    lib = ctypes.CDLL(_native.LIB_PATH)
    names = _header_symbols()
    assert len(names) >= 24
    for name in names:
        assert hasattr(lib, name), name
    assert names == set(_native.SIGNATURES), names ^ set(_native.SIGNATURES)
    assert _native.load().kfac_abi_version() == 1


def test_no_cpu_fallback():
    x = torch.randn(8, 4)
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        ops.factor_linear(x, True, torch.empty(5, 5))
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        ops.gemm(torch.randn(4, 4), torch.randn(4, 4))


def test_errors_come_back_as_runtime_errors():
    lib = _native.load()
    # null pointers are rejected before any launch
    assert lib.kfac_set_identity(None, 4, None) != 0
    assert b'bad arguments' in lib.kfac_last_error()
    assert lib.kfac_factor_linear(None, 4, 4, 0, 1.0, 1.0, 0.0, None, None, 0, None) != 0
    with pytest.raises(RuntimeError):
        _native.check(1)
    assert lib.kfac_factor_workspace_bytes(0, 10) == 0
    assert lib.kfac_factor_workspace_bytes(256, 100352) > 0


def _model():
    return torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.ReLU(), torch.nn.Linear(5, 3, bias=False))


@pytest.mark.parametrize('kwargs', [
    dict(allreduce_bucket_cap_mb=-1),
    dict(compute_eigenvalue_outer_product=True, colocate_factors=False),
    dict(grad_worker_fraction=2),
    dict(grad_worker_fraction=-1),
    dict(factor_update_steps=0),
    dict(inv_update_steps=-1),
    dict(damping=0),
    dict(factor_decay=1.5),
    dict(factor_decay=0),
    dict(kl_clip=0),
    dict(lr=-1),
    dict(accumulation_steps=0),
])
def test_constructor_rejects_like_reference(kwargs):
    """ValueErrors of kfac/preconditioner.py:151-161,181-189 and kfac/base_preconditioner.py:86-100."""
    with pytest.raises(ValueError):
        KFACPreconditioner(_model(), **kwargs)


def test_constructor_resolution_and_repr():
    p = KFACPreconditioner(_model(), assignment_strategy='memory', compute_method='inverse')
    assert p.assignment_strategy == AssignmentStrategy.MEMORY
    assert p.compute_method == ComputeMethod.INVERSE
    assert p.distributed_strategy == DistributedStrategy.COMM_OPT and p.grad_worker_fraction == 1.0
    p = KFACPreconditioner(_model(), grad_worker_fraction=DistributedStrategy.MEM_OPT)
    assert p.grad_worker_fraction == 1.0  # world size 1
    p = KFACPreconditioner(_model(), grad_worker_fraction=0)
    assert p.distributed_strategy == DistributedStrategy.COMM_OPT
    assert 'KFACPreconditioner(' in repr(p) and 'layers=2' in repr(p)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter('always')
        KFACPreconditioner(_model(), factor_update_steps=3, inv_update_steps=4)
        assert any('integer multiple' in str(x.message) for x in w)


def test_callable_hyperparameters_and_state_dict_keys():
    p = KFACPreconditioner(_model(), damping=lambda s: 0.1 + s, lr=lambda s: 0.5, skip_layers=['2'])
    assert len(p._layers) == 1
    assert p.damping == 0.1 and p.lr == 0.5
    p._steps = 2
    assert p.damping == 2.1
    sd = p.state_dict(include_factors=False)
    assert 'damping' not in sd and 'lr' not in sd and sd['steps'] == 2 and sd['kl_clip'] == 0.001
    assert set(p.state_dict()['layers']) == {'0'}
    # attributes the reference's LambdaParamScheduler mutates (kfac/scheduler.py:118-166)
    for attr in ('_damping', '_lr', '_kl_clip', '_factor_decay', '_factor_update_steps', '_inv_update_steps',
                 '_steps', '_layers', '_assignment', '_tdc'):
        assert hasattr(p, attr)
    with pytest.raises(ValueError):
        q = KFACPreconditioner(_model())
        q.load_state_dict({'steps': 0, 'layers': {}})
    q = KFACPreconditioner(_model())
    with pytest.warns(UserWarning, match='not included'):
        q.load_state_dict({'steps': 7, 'lr': 0.3})
    assert q.steps == 7 and q.lr == 0.3
    assert q.memory_usage()['total'] == 0


def test_layer_state_machine_errors():
    """Same RuntimeErrors as kfac/layers/base.py:292,414 and eigen.py:303,360 when stages run out of order."""
    p = KFACPreconditioner(_model())
    (_, layer) = list(p._layers.values())[0]
    with pytest.raises(RuntimeError, match='cannot reduce'):
        layer.reduce_a_factor()
    with pytest.raises(RuntimeError, match='before A has been computed'):
        layer.compute_a_inv()
    with pytest.raises(RuntimeError, match='have not been computed'):
        layer.preconditioned_grad()
    with pytest.raises(RuntimeError, match='preconditionined gradient is None'):
        layer.update_grad()
    with pytest.raises(KeyError):
        layer.load_state_dict({'A': None})
    q = KFACPreconditioner(_model(), compute_method='inverse')
    (_, layer) = list(q._layers.values())[0]
    with pytest.raises(RuntimeError, match='before A has been computed'):
        layer.compute_a_inv()
    with pytest.raises(RuntimeError, match='have been inverted'):
        layer.preconditioned_grad()
    assert kfac_b200.__version__
