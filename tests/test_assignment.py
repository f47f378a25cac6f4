"""KAISA placement against the reference's own golden tables (tests/assignment_test.py:60-141, 193-428, 431-525)
and against full tables computed by the reference's KAISAAssignment (tests/golden/assignment.json)."""
from __future__ import annotations

import json
import os

import pytest

from kfac_b200.assignment import KAISAAssignment


@pytest.fixture(scope='module')
def gold(golden_dir):
    with open(os.path.join(golden_dir, 'assignment.json')) as f:
        return json.load(f)


def test_partition_grad_workers(gold):
    for world, workers, expected in gold['partition_grad_workers']:
        assert KAISAAssignment.partition_grad_workers(world, workers) == {frozenset(r) for r in expected}


def test_partition_grad_receivers(gold):
    for world, workers, expected in gold['partition_grad_receivers']:
        assert KAISAAssignment.partition_grad_receivers(world, workers) == {frozenset(r) for r in expected}


@pytest.mark.parametrize('world,workers', [(4, 8), (4, 3), (0, 2)])
def test_partition_input_checks(world, workers):
    with pytest.raises(ValueError):
        KAISAAssignment.partition_grad_workers(world, workers)
    with pytest.raises(ValueError):
        KAISAAssignment.partition_grad_receivers(world, workers)


def test_greedy_assignment_goldens(gold):
    assert len(gold['greedy_assignment']) >= 14
    for work, groups, world, colocate, expected in gold['greedy_assignment']:
        assert KAISAAssignment.greedy_assignment(work, groups, world, colocate) == expected


def test_group_sizes(gold):
    work = {'l1': {'A': 1, 'G': 1}, 'l2': {'A': 1, 'G': 1}}
    for world, frac, colocate, worker_size, receiver_size in gold['group_sizes']:
        for rank in range(world):
            a = KAISAAssignment(work, local_rank=rank, world_size=world, grad_worker_fraction=frac,
                                group_func=lambda r: tuple(sorted(r)), colocate_factors=colocate)
            for layer in a.get_layers():
                assert len(a.grad_worker_group(layer)) == worker_size
                assert len(a.grad_receiver_group(layer)) == receiver_size
            assert a.broadcast_gradients() == (worker_size < world)
            assert a.broadcast_inverses() == (worker_size > 1)


def test_full_tables_match_reference(gold):
    work = gold['kaisa_work']
    for table in gold['kaisa_tables']:
        world, frac, colocate = table['world'], table['frac'], table['colocate']
        for rank, ref in enumerate(table['ranks']):
            a = KAISAAssignment(work, local_rank=rank, world_size=world, grad_worker_fraction=frac,
                                group_func=lambda r: tuple(sorted(r)), colocate_factors=colocate)
            assert set(a.get_layers()) == set(ref['inv_worker'])
            for layer in a.get_layers():
                assert {f: a.inv_worker(layer, f) for f in a.get_factors(layer)} == ref['inv_worker'][layer]
                assert a.is_grad_worker(layer) == ref['is_grad_worker'][layer]
                assert a.src_grad_worker(layer) == ref['src_grad_worker'][layer]
                assert list(a.grad_worker_group(layer)) == ref['grad_worker_group'][layer]
                assert list(a.grad_receiver_group(layer)) == ref['grad_receiver_group'][layer]
                assert a.factor_group(layer, 'A') is None
            assert a.broadcast_gradients() == ref['broadcast_gradients']
            assert a.broadcast_inverses() == ref['broadcast_inverses']


@pytest.mark.parametrize('kwargs', [
    dict(local_rank=0, world_size=4, grad_worker_fraction=1.5),
    dict(local_rank=-1, world_size=4, grad_worker_fraction=1),
    dict(local_rank=0, world_size=-1, grad_worker_fraction=1),
    dict(local_rank=5, world_size=4, grad_worker_fraction=1),
    dict(local_rank=0, world_size=4, grad_worker_fraction=0.33),
])
def test_input_checks(kwargs):
    with pytest.raises(ValueError):
        KAISAAssignment({}, group_func=lambda r: None, **kwargs)
