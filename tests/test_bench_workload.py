"""The bench's synthetic workload is the one BASELINE.json configs[1] / SURVEY.md section 8d name (CPU-only checks).

bench.py times the K-FAC hot path on the layer geometry of torchvision's ResNet-50; a wrong table there would make
every figure it prints a figure for another workload, so the table is pinned here against the algorithmic totals
BASELINE.md quotes (54 layers, 108 factors, sum n^3 = 452.8 G, 0.803 TFLOP of SYRK per step, 153.9 M factor
elements = 615 MB, 25.5 M gradient elements) and against the layer the reference would register
(kfac/layers/register.py: every Conv2d / Linear leaf).
"""
from __future__ import annotations

import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import bench  # noqa: E402


def test_resnet50_layer_table_matches_the_quoted_totals():
    layers = bench.resnet50_layers()
    assert len(layers) == 54
    assert sum(1 for L in layers if L['kind'] == 'conv') == 53 and layers[-1]['kind'] != 'conv'
    # factor sizes: A = cin * k * k (+ 1 with a bias), G = cout
    for L in layers:
        if L['kind'] == 'conv':
            assert L['a'] == L['cin'] * L['k'] * L['k'] and L['g'] == L['cout']
        else:
            assert L['a'] == L['cin'] + 1 and L['g'] == L['cout']
    n3 = sum(L['a'] ** 3 + L['g'] ** 3 for L in layers)
    assert abs(n3 / 1e9 - 452.8) < 0.1, n3
    n2 = sum(L['a'] ** 2 + L['g'] ** 2 for L in layers)
    assert abs(n2 / 1e6 - 153.9) < 0.1, n2                      # 615 MB of fp32 factors
    grads = sum(L['a'] * L['g'] for L in layers)
    assert abs(grads / 1e6 - 25.5) < 0.1, grads
    syrk = sum(bench.work_of(L)['syrk'] for L in layers)
    assert abs(syrk / 1e12 - 0.8036) < 0.001, syrk               # flops of one factor stage at batch 32
    assert max(L['a'] for L in layers) == 4608 and sum(1 for L in layers if L['a'] == 4608) == 3


def test_modules_and_synthetic_tensors_have_the_layers_shapes():
    layers = bench.resnet50_layers()
    mods = bench.make_modules(layers)
    assert len(list(mods.children())) == 54
    for L in (layers[0], layers[1], layers[2], layers[-1]):
        m = mods[L['name'].replace('.', '_')]
        d = bench.synth(L, 0, None)
        out = m(d['x'])
        assert out.shape == d['go'].shape, (L['name'], out.shape, d['go'].shape)
        assert d['wg'].shape == m.weight.shape
        assert (d['bg'] is None) == (m.bias is None)
        # rows of the im2col matrix / of the activation matrix
        rows = out.numel() // L['g']
        assert rows == L['M'], (L['name'], rows, L['M'])
    # data differs per rank, gradients do not (they stand for DDP-averaged gradients)
    a, b = bench.synth(layers[3], 0, None), bench.synth(layers[3], 1, None)
    assert not torch.equal(a['x'], b['x']) and torch.equal(a['wg'], b['wg'])


def test_parity_sample_covers_first_largest_and_biased_layer():
    layers = bench.resnet50_layers()
    idx = bench.parity_sample(layers)
    names = [layers[i]['name'] for i in idx]
    assert idx[0] == 0 and len(idx) == len(set(idx)) == 10
    assert names[-1] == layers[-1]['name']                       # fc: the only bias column
    assert any(layers[i]['a'] == 4608 for i in idx)              # a 4608 x 512 layer4 convolution
