"""kfac_factor_linear_noalloc: the factor entry point for hosts that may not allocate device memory mid-step
(their framework owns the whole device, or they cannot afford the device-wide stall of cudaMalloc/cudaFree).
Its scratch is a fixed 16 MiB array inside the library, so the checks here are: same numbers as the regular
entry point when called through the C ABI, twice in a row with different data (the second call finds the first
call's partial sums in the scratch and must not be disturbed by them), and a clean refusal -- not an allocation --
when a problem needs more scratch than the library carries.

Collected last on purpose (file name sorts after every other test file); passes alone in a fresh process and
after the rest of the suite."""
from __future__ import annotations

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 1e-5   # the factor tolerance used throughout tests/test_kernels_gpu.py


def _relerr(a, b):
    return (torch.linalg.norm(a.double() - b.double()) / torch.linalg.norm(b.double())).item()


def test_noalloc_matches_regular_entry_point_two_calls_in_a_row():
    from kfac_b200 import _native
    lib = _native.load()
    stream = torch.cuda.current_stream().cuda_stream
    cap = lib.kfac_noalloc_scratch_bytes()
    assert cap == 16 << 20
    torch.manual_seed(1234)
    # tall-skinny inputs: these take the split-K path, i.e. they really use the scratch
    for call, (m, n_in, bias, alpha, beta) in enumerate([(60000, 96, True, 0.05, 0.95), (40000, 64, False, 1.0, 0.0)]):
        n = n_in + int(bias)
        need = lib.kfac_factor_workspace_bytes(n, m)
        assert 0 < need <= cap, (need, cap)
        x = torch.randn(m, n_in, device='cuda')
        start = torch.eye(n, device='cuda')
        ref, out = start.clone(), start.clone()
        ws = torch.empty(need, dtype=torch.uint8, device='cuda')
        _native.check(lib.kfac_factor_linear(x.data_ptr(), m, n_in, int(bias), 1.0, alpha, beta, ref.data_ptr(),
                                             ws.data_ptr(), ws.numel(), stream))
        # Free device memory as the driver reports it (this also sees a raw cudaMalloc, which torch's own counters
        # cannot).  The very first use of the library on a device may materialise its module, built-in scratch
        # included, so the strict comparison is made around the SECOND call.
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        torch_before = torch.cuda.memory_allocated()
        _native.check(lib.kfac_factor_linear_noalloc(x.data_ptr(), m, n_in, int(bias), 1.0, alpha, beta,
                                                     out.data_ptr(), stream))
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == torch_before
        if call > 0:
            assert torch.cuda.mem_get_info()[0] == free_before
        assert _relerr(out, ref) < TOL, (call, _relerr(out, ref))
        assert torch.equal(out, out.t())
        # and against fp64 directly, so both entry points cannot be wrong together
        xd = x.double()
        if bias:
            xd = torch.cat([xd, torch.ones(m, 1, device='cuda', dtype=torch.double)], 1)
        truth = beta * torch.eye(n, device='cuda', dtype=torch.double) + alpha * xd.t() @ xd / m
        assert _relerr(out, truth) < TOL


def test_noalloc_rejects_problems_beyond_its_scratch():
    from kfac_b200 import _native
    lib = _native.load()
    m, n_in = 400000, 256
    assert lib.kfac_factor_workspace_bytes(n_in, m) > lib.kfac_noalloc_scratch_bytes()
    x = torch.zeros(8, device='cuda')   # never dereferenced: the size check comes first
    f = torch.zeros(n_in, n_in, device='cuda')
    rc = lib.kfac_factor_linear_noalloc(x.data_ptr(), m, n_in, 0, 1.0, 1.0, 0.0, f.data_ptr(),
                                        torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    msg = lib.kfac_last_error().decode()
    assert 'built into the library' in msg and 'kfac_factor_linear' in msg
    with pytest.raises(RuntimeError):
        _native.check(rc)
