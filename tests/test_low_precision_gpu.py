"""Low-precision modes of the host layer (factor_dtype / inv_dtype = bfloat16 or float16), reference semantics:
kfac/layers/base.py:349,363-365,382-383 (`.to(self.factor_dtype)` of activations and grad-outputs) and
kfac/layers/eigen.py:318-319 (second-order data cast to inv_dtype).

What is stated and tested here (north star: "any bf16 mode is stated separately with its own tolerance"):
  * the factor of 16-bit inputs is the EXACT covariance of the rounded values (single-pass tf32 UMMA, fp32
    accumulation, fp32 storage): <= 2e-6 against the fp64 covariance of the same rounded values;
  * against the fp32 factor of the unrounded inputs the mode costs the rounding of the inputs only:
    <= 2e-2 (bf16, 8 mantissa bits) / <= 3e-3 (fp16, 11 bits) relative Frobenius on the off-diagonal part;
  * inv_dtype rounds Q / d / dgda to the 16-bit type's values: preconditioned gradient within 3e-2 (bf16) / 4e-3
    (fp16) of the fp32 mode."""
from __future__ import annotations

import pytest
import torch

pytestmark = pytest.mark.gpu


def relerr(a, b):
    a, b = a.double(), b.double()
    return (torch.linalg.norm(a - b) / torch.linalg.norm(b)).item()


def offdiag(t):
    t = t.double().clone()
    t.fill_diagonal_(0)
    return t


def _net():
    import models
    torch.manual_seed(3)
    return models.MODELS['convnet']().cuda()


@pytest.mark.parametrize('dtype,tol_in', [(torch.bfloat16, 2e-2), (torch.float16, 3e-3)])
def test_factor_dtype_rounds_inputs_and_multiplies_exactly(dtype, tol_in):
    import models
    from kfac_b200 import KFACPreconditioner
    from oracle import kfac_oracle as orc
    x, y = models.make_batch('convnet', 0)
    nets = {}
    for key, kw in (('low', {'factor_dtype': dtype}), ('f32', {})):
        net = _net()
        pre = KFACPreconditioner(net, loglevel=0, **kw)
        models.loss_fn('convnet', net(x.cuda()), y.cuda()).backward()
        nets[key] = (net, pre)
    low = {name: layer for name, layer in nets['low'][1]._layers.values()}
    f32 = {name: layer for name, layer in nets['f32'][1]._layers.values()}
    # expected A factor of conv1 from the ROUNDED input, through the oracle, in fp64-grade arithmetic
    xr = x.cuda().to(dtype).double()
    conv = nets['low'][0].conv1
    expect = orc.conv2d_a_factor(xr.cpu(), tuple(conv.kernel_size), tuple(conv.stride), tuple(conv.padding), True)
    expect = 0.95 * torch.eye(expect.shape[0], dtype=torch.double) + 0.05 * expect
    got = low['conv1'].a_factor
    assert got.dtype == torch.float32
    assert relerr(offdiag(got.cpu()), offdiag(expect)) < 2e-6
    for name in low:
        for which in ('a_factor', 'g_factor'):
            a, b = getattr(low[name], which), getattr(f32[name], which)
            err = relerr(offdiag(a), offdiag(b))
            assert 0 < err < tol_in, (name, which, err)


@pytest.mark.parametrize('dtype,tol', [(torch.bfloat16, 3e-2), (torch.float16, 4e-3)])
def test_inv_dtype_rounds_second_order_data(dtype, tol):
    import models
    from kfac_b200 import KFACPreconditioner
    x, y = models.make_batch('convnet', 0)
    grads = {}
    for key, kw in (('low', {'inv_dtype': dtype}), ('f32', {})):
        net = _net()
        pre = KFACPreconditioner(net, loglevel=0, **kw)
        models.loss_fn('convnet', net(x.cuda()), y.cuda()).backward()
        pre.step()
        grads[key] = {n: p.grad.detach().clone() for n, p in net.named_parameters()}
        if key == 'low':
            for _, layer in pre._layers.values():
                q = layer.qa
                assert q.dtype == torch.float32 and torch.equal(q, q.to(dtype).float())   # values of the 16-bit type
    for n in grads['f32']:
        err = relerr(grads['low'][n], grads['f32'][n])
        assert 0 < err < tol, (n, err)


def test_unsupported_dtypes_are_rejected():
    from kfac_b200 import KFACPreconditioner
    with pytest.raises(ValueError, match='factor_dtype'):
        KFACPreconditioner(_net(), factor_dtype=torch.float64, loglevel=0)
    with pytest.raises(ValueError, match='inv_dtype'):
        KFACPreconditioner(_net(), inv_dtype=torch.float64, loglevel=0)
