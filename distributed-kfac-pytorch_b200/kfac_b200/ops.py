"""Tensor-level wrappers over the C ABI: unpack torch tensors into device pointers, sizes and the stream.

Every function enqueues native sm_100a kernels on torch's current CUDA stream and returns immediately.
Input tensors must be contiguous fp32 CUDA tensors; there is no CPU path.
"""
from __future__ import annotations

import ctypes
import functools

import torch

from kfac_b200 import _native
from kfac_b200._native import check
from kfac_b200._native import ptr
from kfac_b200._native import stream_handle
from kfac_b200._native import WORKSPACE


def _first_device(obj) -> torch.device | None:
    if isinstance(obj, torch.Tensor):
        return obj.device if obj.is_cuda else None
    if isinstance(obj, dict):
        obj = list(obj.values())
    if isinstance(obj, (list, tuple)):
        for o in obj:
            d = _first_device(o)
            if d is not None:
                return d
    return None


def _on_device(fn):
    """Run the wrapped op with the device of its first CUDA tensor argument current: the library launches on the
    CURRENT device, while pointers and the stream handle come from the tensors (a model on cuda:1 driven from a
    thread whose current device is cuda:0 must still work, as it does with the reference's torch ops)."""
    @functools.wraps(fn)
    def wrapped(*args, **kwargs):
        dev = _first_device(args) or _first_device(kwargs)
        if dev is None or dev.index is None or dev.index == torch.cuda.current_device():
            return fn(*args, **kwargs)
        with torch.cuda.device(dev):
            return fn(*args, **kwargs)
    return wrapped


def _f32(t: torch.Tensor, name: str) -> torch.Tensor:
    if t.dtype != torch.float32:
        raise RuntimeError(f'{name} must be float32, got {t.dtype}')
    return t


_DTYPE_CODE = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}


def _factor_input(t: torch.Tensor, name: str) -> tuple[torch.Tensor, int]:
    """Activations / grad-outputs are read in place in fp32, fp16 or bf16 (KFAC_DTYPE_*); anything else is an error."""
    code = _DTYPE_CODE.get(t.dtype)
    if code is None:
        raise RuntimeError(f'{name} must be float32, float16 or bfloat16, got {t.dtype}')
    return (t if t.is_contiguous() else t.contiguous()), code


@_on_device
def factor_linear(
    x: torch.Tensor,
    has_bias: bool,
    out: torch.Tensor,
    *,
    alpha: float = 1.0,
    beta: float = 0.0,
    in_scale: float = 1.0,
) -> torch.Tensor:
    """out <- beta*out + alpha*in_scale^2 * [x,1]^T[x,1]/M  (modules.py:123-141 + utils.py:50-56)."""
    lib = _native.load()
    x2, dt = _factor_input(x.reshape(-1, x.shape[-1]), 'x')
    m, n_in = x2.shape
    n = n_in + int(has_bias)
    assert out.shape == (n, n)
    ws, ws_bytes = WORKSPACE.get(lib.kfac_factor_workspace_bytes(n, m), x.device)
    check(lib.kfac_factor_linear_dt(ptr(x2), dt, m, n_in, int(has_bias), in_scale, alpha, beta,
                                    ptr(_f32(out, 'out')), ws, ws_bytes, stream_handle(x.device)))
    return out


@_on_device
def factor_conv2d_a(
    x: torch.Tensor,
    kernel_size: tuple[int, int],
    stride: tuple[int, int],
    padding: tuple[int, int],
    has_bias: bool,
    out: torch.Tensor,
    *,
    alpha: float = 1.0,
    beta: float = 0.0,
) -> torch.Tensor:
    """A factor of a Conv2d from its NCHW input, im2col fused (modules.py:170-178,210-237)."""
    lib = _native.load()
    x, dt = _factor_input(x, 'x')
    nb, c, h, w = x.shape
    kh, kw = kernel_size
    sh, sw = stride
    ph, pw = padding
    n = c * kh * kw + int(has_bias)
    assert out.shape == (n, n)
    oh = (h + 2 * ph - kh) // sh + 1
    ow = (w + 2 * pw - kw) // sw + 1
    ws, ws_bytes = WORKSPACE.get(lib.kfac_factor_workspace_bytes(n, nb * oh * ow), x.device)
    check(lib.kfac_factor_conv2d_a_dt(ptr(x), dt, nb, c, h, w, kh, kw, sh, sw, ph, pw, int(has_bias), alpha, beta,
                                      ptr(_f32(out, 'out')), ws, ws_bytes, stream_handle(x.device)))
    return out


@_on_device
def factor_conv2d_g(
    g: torch.Tensor,
    out: torch.Tensor,
    *,
    alpha: float = 1.0,
    beta: float = 0.0,
    in_scale: float = 1.0,
) -> torch.Tensor:
    """G factor of a Conv2d from its (N, Cout, OH, OW) grad-output (modules.py:180-192)."""
    lib = _native.load()
    g, dt = _factor_input(g, 'g')
    nb, c, oh, ow = g.shape
    assert out.shape == (c, c)
    ws, ws_bytes = WORKSPACE.get(lib.kfac_factor_workspace_bytes(c, nb * oh * ow), g.device)
    check(lib.kfac_factor_conv2d_g_dt(ptr(g), dt, nb, c, oh, ow, in_scale, alpha, beta, ptr(_f32(out, 'out')), ws,
                                      ws_bytes, stream_handle(g.device)))
    return out


@_on_device
def axpby(y: torch.Tensor, x: torch.Tensor, a: float, b: float) -> torch.Tensor:
    """y <- a*x + b*y (base.py:382-388)."""
    lib = _native.load()
    assert x.numel() == y.numel()
    check(lib.kfac_axpby(ptr(_f32(y, 'y')), ptr(_f32(x, 'x')), a, b, y.numel(), stream_handle(y.device)))
    return y


@_on_device
def set_identity(f: torch.Tensor) -> torch.Tensor:
    lib = _native.load()
    check(lib.kfac_set_identity(ptr(_f32(f, 'F')), f.shape[0], stream_handle(f.device)))
    return f


@_on_device
def add_diag(f: torch.Tensor, damping: float) -> torch.Tensor:
    lib = _native.load()
    check(lib.kfac_add_diag(ptr(_f32(f, 'F')), f.shape[0], damping, stream_handle(f.device)))
    return f


@_on_device
def triu_pack(f: torch.Tensor, packed: torch.Tensor | None = None) -> torch.Tensor:
    """get_triu (distributed.py:416-427) without building index tensors."""
    lib = _native.load()
    n = f.shape[0]
    if packed is None:
        packed = torch.empty(n * (n + 1) // 2, dtype=torch.float32, device=f.device)
    check(lib.kfac_triu_pack(ptr(_f32(f, 'F')), n, ptr(packed), stream_handle(f.device)))
    return packed


@_on_device
def triu_unpack(packed: torch.Tensor, n: int, out: torch.Tensor | None = None) -> torch.Tensor:
    """fill_triu (distributed.py:430-459)."""
    lib = _native.load()
    if out is None:
        out = torch.empty(n, n, dtype=torch.float32, device=packed.device)
    check(lib.kfac_triu_unpack(ptr(_f32(packed, 'packed')), n, ptr(out), stream_handle(packed.device)))
    return out


@_on_device
def eigen_outer(dg: torch.Tensor, da: torch.Tensor, damping: float, out: torch.Tensor | None = None) -> torch.Tensor:
    """1 / (outer(dg, da) + damping)  (eigen.py:344-345)."""
    lib = _native.load()
    g, a = dg.numel(), da.numel()
    if out is None:
        out = torch.empty(g, a, dtype=torch.float32, device=dg.device)
    check(lib.kfac_eigen_outer(ptr(_f32(dg, 'dg')), ptr(_f32(da, 'da')), g, a, damping, ptr(out),
                               stream_handle(dg.device)))
    return out


@_on_device
def gemm(
    a: torch.Tensor,
    b: torch.Tensor,
    *,
    transa: bool = False,
    transb: bool = False,
    alpha: float = 1.0,
    beta: float = 0.0,
    out: torch.Tensor | None = None,
) -> torch.Tensor:
    """out <- alpha * op(a) @ op(b) + beta * out on the tensor cores with fp32-grade accuracy."""
    lib = _native.load()
    m, k = (a.shape[1], a.shape[0]) if transa else (a.shape[0], a.shape[1])
    k2, n = (b.shape[1], b.shape[0]) if transb else (b.shape[0], b.shape[1])
    assert k == k2
    if out is None:
        out = torch.empty(m, n, dtype=torch.float32, device=a.device)
    ws, ws_bytes = WORKSPACE.get(lib.kfac_gemm_workspace_bytes(m, n, k), a.device)
    for t in (a, b, out):   # row-major views with a leading dimension are fine (sub-matrices)
        if not t.is_cuda or t.dim() != 2 or t.stride(1) != 1:
            raise RuntimeError('gemm operands must be 2-D CUDA tensors with unit column stride (no CPU fallback)')
    check(lib.kfac_gemm(int(transa), int(transb), m, n, k, alpha, _f32(a, 'a').data_ptr(), a.stride(0),
                        _f32(b, 'b').data_ptr(), b.stride(0), beta, out.data_ptr(), out.stride(0), ws, ws_bytes,
                        stream_handle(a.device)))
    return out


@_on_device
def precondition_eigen(
    w_grad: torch.Tensor,
    b_grad: torch.Tensor | None,
    qa: torch.Tensor,
    qg: torch.Tensor,
    *,
    da: torch.Tensor | None = None,
    dg: torch.Tensor | None = None,
    dgda: torch.Tensor | None = None,
    damping: float = 0.001,
    out: torch.Tensor | None = None,
    vg_accum: torch.Tensor | None = None,
) -> torch.Tensor:
    """KFACEigenLayer.preconditioned_grad (eigen.py:349-384) on weight.grad / bias.grad in place."""
    lib = _native.load()
    g = w_grad.shape[0]
    wg = _f32(w_grad, 'w_grad').reshape(g, -1)
    a_w = wg.shape[1]
    a = a_w + int(b_grad is not None)
    if out is None:
        out = torch.empty(g, a, dtype=torch.float32, device=w_grad.device)
    ws, ws_bytes = WORKSPACE.get(lib.kfac_precondition_workspace_bytes(g, a), w_grad.device)
    check(lib.kfac_precondition_eigen(ptr(wg), ptr(b_grad), g, a_w, ptr(qa), ptr(qg), ptr(da), ptr(dg), ptr(dgda),
                                      damping, ptr(out), ptr(vg_accum), ws, ws_bytes,
                                      stream_handle(w_grad.device)))
    return out


@_on_device
def precondition_eigen_grouped(
    items: list[dict],
    *,
    damping: float,
    vg_accum: torch.Tensor | None = None,
) -> None:
    """All layers of a step in four grouped launches (base_preconditioner.py:363-365 + eigen.py:349-384).

    items: dicts with w_grad (g, a_w) contiguous, b_grad or None, qa, qg, out (g, a) and either dgda or da + dg.
    """
    lib = _native.load()
    count = len(items)
    if count == 0:
        return
    dev = items[0]['w_grad'].device

    def arr(key: str):
        return (ctypes.c_void_p * count)(*[ptr(it.get(key)) for it in items])

    gs = [int(it['w_grad'].shape[0]) for it in items]
    aws = [int(it['w_grad'].numel() // it['w_grad'].shape[0]) for it in items]
    a_full = [aw + int(it.get('b_grad') is not None) for aw, it in zip(aws, items)]
    g_arr = (ctypes.c_int * count)(*gs)
    aw_arr = (ctypes.c_int * count)(*aws)
    a_arr = (ctypes.c_int * count)(*a_full)
    ws, ws_bytes = WORKSPACE.get(lib.kfac_precondition_grouped_workspace_bytes(g_arr, a_arr, count), dev)
    check(lib.kfac_precondition_eigen_grouped(count, arr('w_grad'), arr('b_grad'), g_arr, aw_arr, arr('qa'), arr('qg'),
                                              arr('da'), arr('dg'), arr('dgda'), damping, arr('out'), ptr(vg_accum),
                                              ws, ws_bytes, stream_handle(dev)))


@_on_device
def precondition_inverse_grouped(items: list[dict], *, vg_accum: torch.Tensor | None = None) -> None:
    """All layers of a step in two grouped launches (base_preconditioner.py:363-365 + inverse.py:214-233).

    items: dicts with w_grad (g, a_w) contiguous, b_grad or None, a_inv, g_inv, out (g, a)."""
    lib = _native.load()
    count = len(items)
    if count == 0:
        return
    dev = items[0]['w_grad'].device

    def arr(key: str):
        return (ctypes.c_void_p * count)(*[ptr(it.get(key)) for it in items])

    gs = [int(it['w_grad'].shape[0]) for it in items]
    aws = [int(it['w_grad'].numel() // it['w_grad'].shape[0]) for it in items]
    a_full = [aw + int(it.get('b_grad') is not None) for aw, it in zip(aws, items)]
    g_arr = (ctypes.c_int * count)(*gs)
    aw_arr = (ctypes.c_int * count)(*aws)
    a_arr = (ctypes.c_int * count)(*a_full)
    ws, ws_bytes = WORKSPACE.get(lib.kfac_precondition_grouped_workspace_bytes(g_arr, a_arr, count), dev)
    check(lib.kfac_precondition_inverse_grouped(count, arr('w_grad'), arr('b_grad'), g_arr, aw_arr, arr('a_inv'),
                                                arr('g_inv'), arr('out'), ptr(vg_accum), ws, ws_bytes,
                                                stream_handle(dev)))


@_on_device
def precondition_inverse(
    w_grad: torch.Tensor,
    b_grad: torch.Tensor | None,
    a_inv: torch.Tensor,
    g_inv: torch.Tensor,
    *,
    out: torch.Tensor | None = None,
    vg_accum: torch.Tensor | None = None,
) -> torch.Tensor:
    """KFACInverseLayer.preconditioned_grad (inverse.py:214-233)."""
    lib = _native.load()
    g = w_grad.shape[0]
    wg = _f32(w_grad, 'w_grad').reshape(g, -1)
    a_w = wg.shape[1]
    a = a_w + int(b_grad is not None)
    if out is None:
        out = torch.empty(g, a, dtype=torch.float32, device=w_grad.device)
    ws, ws_bytes = WORKSPACE.get(lib.kfac_precondition_workspace_bytes(g, a), w_grad.device)
    check(lib.kfac_precondition_inverse(ptr(wg), ptr(b_grad), g, a_w, ptr(a_inv), ptr(g_inv), ptr(out),
                                        ptr(vg_accum), ws, ws_bytes, stream_handle(w_grad.device)))
    return out


@_on_device
def grad_dot(v: torch.Tensor, w_grad: torch.Tensor, b_grad: torch.Tensor | None, vg_accum: torch.Tensor) -> None:
    """vg_accum += sum(v * [Wg | bg])  (base_preconditioner.py:415-431)."""
    lib = _native.load()
    g = w_grad.shape[0]
    wg = w_grad.reshape(g, -1)
    check(lib.kfac_grad_dot(ptr(_f32(v, 'v')), ptr(_f32(wg, 'w_grad')), ptr(b_grad), g, wg.shape[1],
                            ptr(vg_accum), stream_handle(v.device)))


@_on_device
def klclip_scale(vg_accum: torch.Tensor, lr: float, kl_clip: float, scale_out: torch.Tensor) -> torch.Tensor:
    """scale_out[0] <- min(1, sqrt(kl_clip/|vg*lr^2|)) on the device (base_preconditioner.py:432-433)."""
    lib = _native.load()
    check(lib.kfac_klclip_scale(ptr(vg_accum), lr, kl_clip, ptr(scale_out), stream_handle(vg_accum.device)))
    return scale_out


@_on_device
def scale_write_grads(
    pg: torch.Tensor,
    scale_dev: torch.Tensor | None,
    w_grad: torch.Tensor,
    b_grad: torch.Tensor | None,
) -> None:
    """weight.grad / bias.grad <- scale * pg split (base.py:406-422, modules.py:87-97), in place."""
    lib = _native.load()
    g = w_grad.shape[0]
    a_w = w_grad.numel() // g
    check(lib.kfac_scale_write_grads(ptr(_f32(pg, 'pg')), g, a_w, ptr(scale_dev), ptr(_f32(w_grad, 'w_grad')),
                                     ptr(b_grad), stream_handle(pg.device)))


@_on_device
def scale_write_grads_grouped(items: list[tuple[torch.Tensor, torch.Tensor, torch.Tensor | None]],
                              scale_dev: torch.Tensor | None) -> None:
    """update_grad of all layers in one launch (base_preconditioner.py:376-377): items = (pg, w_grad, b_grad)."""
    lib = _native.load()
    count = len(items)
    if count == 0:
        return
    dev = items[0][0].device
    gs = [int(w.shape[0]) for _, w, _ in items]
    aws = [int(w.numel() // w.shape[0]) for _, w, _ in items]
    pg_arr = (ctypes.c_void_p * count)(*[ptr(_f32(pg, 'pg')) for pg, _, _ in items])
    w_arr = (ctypes.c_void_p * count)(*[ptr(_f32(w, 'w_grad')) for _, w, _ in items])
    b_arr = (ctypes.c_void_p * count)(*[ptr(b) for _, _, b in items])
    ws, ws_bytes = WORKSPACE.get(lib.kfac_scale_write_grouped_workspace_bytes(count), dev)
    check(lib.kfac_scale_write_grads_grouped(count, pg_arr, (ctypes.c_int * count)(*gs), (ctypes.c_int * count)(*aws),
                                             ptr(scale_dev), w_arr, b_arr, ws, ws_bytes, stream_handle(dev)))


@_on_device
def eigh_batched(
    factors: list[torch.Tensor],
    *,
    clamp_nonneg: bool = True,
    qs: list[torch.Tensor] | None = None,
    ds: list[torch.Tensor] | None = None,
) -> tuple[list[torch.Tensor], list[torch.Tensor]]:
    """Batched symmetric eigendecomposition F = Q diag(d) Q^T (eigen.py:294-343)."""
    lib = _native.load()
    count = len(factors)
    if count == 0:
        return [], []
    dev = factors[0].device
    ns = [int(f.shape[0]) for f in factors]
    if qs is None:
        qs = [torch.empty(n, n, dtype=torch.float32, device=dev) for n in ns]
    if ds is None:
        ds = [torch.empty(n, dtype=torch.float32, device=dev) for n in ns]
    ns_arr = (ctypes.c_int * count)(*ns)
    f_arr = (ctypes.c_void_p * count)(*[ptr(_f32(f, 'F')) for f in factors])
    q_arr = (ctypes.c_void_p * count)(*[ptr(q) for q in qs])
    d_arr = (ctypes.c_void_p * count)(*[ptr(d) for d in ds])
    ws, ws_bytes = WORKSPACE.get(lib.kfac_eigh_workspace_bytes(ns_arr, count), dev)
    check(lib.kfac_eigh_batched(f_arr, ns_arr, count, q_arr, d_arr, int(clamp_nonneg), ws, ws_bytes,
                                stream_handle(dev)))
    return qs, ds


@_on_device
def inverse_newton_schulz(
    f: torch.Tensor,
    damping: float,
    *,
    out: torch.Tensor | None = None,
    max_iters: int = 32,
    tol: float = 1e-6,
) -> tuple[torch.Tensor, int]:
    """(F + damping I)^-1 by Newton-Schulz on the tensor cores (replaces inverse.py:197-212)."""
    lib = _native.load()
    n = f.shape[0]
    if out is None:
        out = torch.empty(n, n, dtype=torch.float32, device=f.device)
    iters = ctypes.c_int(0)
    ws, ws_bytes = WORKSPACE.get(lib.kfac_inverse_workspace_bytes(n), f.device)
    check(lib.kfac_inverse_newton_schulz(ptr(_f32(f, 'F')), n, damping, ptr(out), max_iters, tol,
                                         ctypes.byref(iters), ws, ws_bytes, stream_handle(f.device)))
    return out, iters.value


@_on_device
def inverse_newton_schulz_batched(
    factors: list[torch.Tensor],
    damping: float,
    *,
    outs: list[torch.Tensor] | None = None,
    max_iters: int = 32,
    tol: float = 1e-6,
) -> tuple[list[torch.Tensor], torch.Tensor]:
    """(F_i + damping I)^-1 for all factors in one asynchronous call (replaces the per-layer torch.linalg.inv of
    inverse.py:185-212).  Returns (inverses, info) with info an int32 device tensor (count, 4): iterations, status
    (0 ok, 1 not converged, 2 non-finite), residual bits, 0 -- valid in stream order; nothing here waits for it."""
    lib = _native.load()
    count = len(factors)
    if count == 0:
        return [], torch.empty(0, 4, dtype=torch.int32)
    dev = factors[0].device
    ns = [int(f.shape[0]) for f in factors]
    if outs is None:
        outs = [torch.empty(n, n, dtype=torch.float32, device=dev) for n in ns]
    info = torch.empty(count, 4, dtype=torch.int32, device=dev)
    ns_arr = (ctypes.c_int * count)(*ns)
    f_arr = (ctypes.c_void_p * count)(*[ptr(_f32(f, 'F')) for f in factors])
    o_arr = (ctypes.c_void_p * count)(*[ptr(o) for o in outs])
    ws, ws_bytes = WORKSPACE.get(lib.kfac_inverse_batched_workspace_bytes(ns_arr, count), dev)
    check(lib.kfac_inverse_newton_schulz_batched(f_arr, ns_arr, count, damping, o_arr, max_iters, tol, ptr(info), ws,
                                                 ws_bytes, stream_handle(dev)))
    return outs, info


def eigh_last_paths() -> dict[str, int]:
    """How many matrices of the last eigh_batched call took each tridiagonalisation path (kfac_eigh_last_paths)."""
    lib = _native.load()
    out = (ctypes.c_int * 4)()
    check(lib.kfac_eigh_last_paths(out))
    return {'full': out[0], 'sym_ldg': out[1], 'sym_bulk': out[2], 'two_stage': out[3]}
