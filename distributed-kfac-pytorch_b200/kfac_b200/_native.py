"""ctypes binding of libkfac_b200.so (C ABI declared in include/kfac_b200.h).

The library is the product: there is no CPU or eager-PyTorch fallback.  If the shared object is missing
or a call fails, a RuntimeError is raised (the reference raises RuntimeError for misuse of its layer state
machine, kfac/layers/base.py:239-244; errors from native code surface the same way).
"""
from __future__ import annotations

import ctypes
import os
import threading
from ctypes import c_char_p
from ctypes import c_double
from ctypes import c_float
from ctypes import c_int
from ctypes import c_longlong
from ctypes import c_size_t
from ctypes import c_void_p
from ctypes import POINTER

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# KFAC_B200_LIB: another build of the same library (A/B timing of kernel variants); the default is the in-tree build
LIB_PATH = os.environ.get('KFAC_B200_LIB') or os.path.join(os.path.dirname(_HERE), 'lib', 'libkfac_b200.so')

# name -> (restype, argtypes); must list every symbol of include/kfac_b200.h (checked by the CPU tests)
_P = c_void_p
SIGNATURES: dict[str, tuple] = {
    'kfac_last_error': (c_char_p, []),
    'kfac_abi_version': (c_int, []),
    'kfac_launch_count': (ctypes.c_ulonglong, []),
    'kfac_profile_enable': (None, [ctypes.c_int]),
    'kfac_profile_last_ms': (ctypes.c_int, [ctypes.POINTER(ctypes.c_float)]),
    'kfac_profile_detail_ms': (ctypes.c_int, [ctypes.POINTER(ctypes.c_float)]),
    'kfac_eigh_last_paths': (ctypes.c_int, [ctypes.POINTER(ctypes.c_int)]),
    'kfac_factor_workspace_bytes': (c_size_t, [c_int, c_longlong]),
    'kfac_factor_linear': (
        c_int,
        [_P, c_longlong, c_int, c_int, c_float, c_float, c_float, _P, _P, c_size_t, _P],
    ),
    'kfac_noalloc_scratch_bytes': (c_size_t, []),
    'kfac_factor_linear_noalloc': (
        c_int,
        [_P, c_longlong, c_int, c_int, c_float, c_float, c_float, _P, _P],
    ),
    'kfac_factor_conv2d_a': (
        c_int,
        [_P] + [c_int] * 11 + [c_float, c_float, _P, _P, c_size_t, _P],
    ),
    'kfac_factor_conv2d_g': (
        c_int,
        [_P, c_int, c_int, c_int, c_int, c_float, c_float, c_float, _P, _P, c_size_t, _P],
    ),
    'kfac_factor_linear_dt': (
        c_int,
        [_P, c_int, c_longlong, c_int, c_int, c_float, c_float, c_float, _P, _P, c_size_t, _P],
    ),
    'kfac_factor_conv2d_a_dt': (
        c_int,
        [_P, c_int] + [c_int] * 11 + [c_float, c_float, _P, _P, c_size_t, _P],
    ),
    'kfac_factor_conv2d_g_dt': (
        c_int,
        [_P, c_int, c_int, c_int, c_int, c_int, c_float, c_float, c_float, _P, _P, c_size_t, _P],
    ),
    'kfac_axpby': (c_int, [_P, _P, c_float, c_float, c_longlong, _P]),
    'kfac_set_identity': (c_int, [_P, c_int, _P]),
    'kfac_add_diag': (c_int, [_P, c_int, c_float, _P]),
    'kfac_triu_pack': (c_int, [_P, c_int, _P, _P]),
    'kfac_triu_unpack': (c_int, [_P, c_int, _P, _P]),
    'kfac_eigen_outer': (c_int, [_P, _P, c_int, c_int, c_float, _P, _P]),
    'kfac_gemm_workspace_bytes': (c_size_t, [c_int, c_int, c_int]),
    'kfac_gemm': (
        c_int,
        [c_int, c_int, c_int, c_int, c_int, c_float, _P, c_longlong, _P, c_longlong, c_float, _P,
         c_longlong, _P, c_size_t, _P],
    ),
    'kfac_precondition_workspace_bytes': (c_size_t, [c_int, c_int]),
    'kfac_precondition_eigen': (
        c_int,
        [_P, _P, c_int, c_int, _P, _P, _P, _P, _P, c_float, _P, _P, _P, c_size_t, _P],
    ),
    'kfac_precondition_grouped_workspace_bytes': (c_size_t, [POINTER(c_int), POINTER(c_int), c_int]),
    'kfac_precondition_eigen_grouped': (
        c_int,
        [c_int, POINTER(c_void_p), POINTER(c_void_p), POINTER(c_int), POINTER(c_int), POINTER(c_void_p),
         POINTER(c_void_p), POINTER(c_void_p), POINTER(c_void_p), POINTER(c_void_p), c_float, POINTER(c_void_p), _P,
         _P, c_size_t, _P],
    ),
    'kfac_precondition_inverse': (
        c_int,
        [_P, _P, c_int, c_int, _P, _P, _P, _P, _P, c_size_t, _P],
    ),
    'kfac_precondition_inverse_grouped': (
        c_int,
        [c_int, POINTER(c_void_p), POINTER(c_void_p), POINTER(c_int), POINTER(c_int), POINTER(c_void_p),
         POINTER(c_void_p), POINTER(c_void_p), _P, _P, c_size_t, _P],
    ),
    'kfac_grad_dot': (c_int, [_P, _P, _P, c_int, c_int, _P, _P]),
    'kfac_klclip_scale': (c_int, [_P, c_float, c_float, _P, _P]),
    'kfac_scale_write_grads': (c_int, [_P, c_int, c_int, _P, _P, _P, _P]),
    'kfac_scale_write_grouped_workspace_bytes': (c_size_t, [c_int]),
    'kfac_scale_write_grads_grouped': (
        c_int,
        [c_int, POINTER(c_void_p), POINTER(c_int), POINTER(c_int), _P, POINTER(c_void_p), POINTER(c_void_p), _P,
         c_size_t, _P],
    ),
    'kfac_eigh_workspace_bytes': (c_size_t, [POINTER(c_int), c_int]),
    'kfac_eigh_batched': (
        c_int,
        [POINTER(c_void_p), POINTER(c_int), c_int, POINTER(c_void_p), POINTER(c_void_p), c_int, _P,
         c_size_t, _P],
    ),
    'kfac_eigh_set_mode': (c_int, [c_int]),
    'kfac_debug_two_stage_workspace_bytes': (c_size_t, [c_int]),
    'kfac_debug_two_stage_tiles': (c_int, [c_int]),
    'kfac_debug_two_stage': (c_int, [_P, c_int, _P, _P, _P, _P, _P, _P, _P, _P, c_size_t, _P]),
    'kfac_inverse_batched_workspace_bytes': (c_size_t, [POINTER(c_int), c_int]),
    'kfac_inverse_newton_schulz_batched': (
        c_int,
        [POINTER(c_void_p), POINTER(c_int), c_int, c_float, POINTER(c_void_p), c_int, c_float, _P, _P, c_size_t, _P],
    ),
    'kfac_inverse_workspace_bytes': (c_size_t, [c_int]),
    'kfac_inverse_newton_schulz': (
        c_int,
        [_P, c_int, c_float, _P, c_int, c_float, POINTER(c_int), _P, c_size_t, _P],
    ),
}

_lib = None
_lock = threading.Lock()


def load() -> ctypes.CDLL:
    """Load libkfac_b200.so (once) and bind every entry point. Raises RuntimeError if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f'{LIB_PATH} is missing: build it with `python __graft_entry__.py build` '
                '(nvcc, sm_100a). There is no fallback path.',
            )
        lib = ctypes.CDLL(LIB_PATH)
        for name, (restype, argtypes) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = restype
            fn.argtypes = argtypes
        _lib = lib
    return _lib


def check(status: int) -> None:
    """Turn a non-zero status into a RuntimeError carrying kfac_last_error()."""
    if status != 0:
        msg = load().kfac_last_error()
        raise RuntimeError(msg.decode() if msg else f'kfac_b200 native call failed ({status})')


def ptr(t: torch.Tensor | None) -> int | None:
    """Device pointer of a contiguous CUDA fp32/fp64 tensor (None passes NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError(
            'kfac_b200 kernels run on CUDA tensors only (there is no CPU fallback); '
            f'got a tensor on {t.device}',
        )
    if not t.is_contiguous():
        raise RuntimeError('kfac_b200 kernels need contiguous tensors')
    return t.data_ptr()


def stream_handle(device: torch.device | None = None) -> int:
    """cudaStream_t of torch's current stream (kernels are enqueued there)."""
    return torch.cuda.current_stream(device).cuda_stream


class Workspace:
    """Grow-only scratch buffer per device, handed to the C ABI (the library itself never allocates)."""

    def __init__(self) -> None:
        self._bufs: dict[torch.device, torch.Tensor] = {}

    def get(self, nbytes: int, device: torch.device) -> tuple[int | None, int]:
        if nbytes <= 0:
            return None, 0
        buf = self._bufs.get(device)
        if buf is None or buf.numel() < nbytes:
            # the caching allocator returns >= 512-byte aligned blocks
            buf = torch.empty(int(nbytes), dtype=torch.uint8, device=device)
            self._bufs[device] = buf
        return buf.data_ptr(), buf.numel()


WORKSPACE = Workspace()
