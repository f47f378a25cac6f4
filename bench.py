"""bench.py -- K-FAC hot path on ResNet-50 ImageNet shapes (BASELINE.json configs[1]): synthetic 224x224, batch
32 per GPU, eigen method, grad_worker_fraction = HYBRID_OPT, 54 layers / 108 factors.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one full K-FAC update over one batch: the factor stage the reference runs in its hooks (108 fused
SYRK launches incl. running average; kfac/base_preconditioner.py:435-477) + preconditioner.step() on a
factor+inverse step (flush -> 108 eigendecompositions -> broadcasts -> precondition -> KL-clip -> gradient
write-back; :308-380).  The model-free workload of SURVEY.md section 8d is used: per layer X = relu(N(0,1)) of
the layer's input shape, grad-output = N(0,1)/M, gradients = N(0,1); data differs per rank (seed 1000 + rank),
gradients are identical on all ranks (already DDP-averaged).  The objects driven are the public ones
(KFACPreconditioner over real nn.Conv2d / nn.Linear modules; the hook methods are called with the synthetic
tensors, then step()).

Reported: `value` = ms per step with inputs resident in HBM (device-timed, max over ranks); `e2e` = the same
step fed from pinned HOST buffers (H2D of activations / grad-outputs / gradients and D2H of the KL-clip scale +
per-layer gradient norms inside the timed region; `e2e.value` with double-buffered device inputs so that the copies
of step k+1 run under the kernels of step k, `e2e.serial_ms` with a step's copies starting after the previous step
ended); `roofline` for the dominant stage (the tridiagonalisation) with per-kernel figures under
`stages.second_order.kernels`; `tensor_peaks` = cuBLAS TF32 / fp32 and gemm3x on 8192^3; `collectives_per_step`
(N > 1); `configs` = BASELINE configs 3, 4 and 5; `cpu_baseline` =
one full 54-layer step of the reference's own KFACPreconditioner on the host cores (oracle/_ref; the oracle port if
that is absent), measured, not extrapolated; `parity` = a fresh preconditioner's first step on the same synthetic
tensors compared, outside the timed region, with the CPU oracle on a sample of the layers (every rank's data enters
the oracle's factor average, so at N > 1 this covers the all-reduce and both KAISA broadcasts).
`--impl reference` runs only the CPU arm (rank 0 alone under torchrun), up to 3 timed steps.
`--profile-step` runs one warm-up step and --steps steps and nothing else (the command the ncu captures use).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
import zlib

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (os.path.join(ROOT, 'distributed-kfac-pytorch_b200'), ROOT):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

BATCH = 32
METRIC = 'kfac_full_step_ms (factor stage + preconditioner.step() on a factor+inverse step, ResNet-50)'


# --------------------------------------------------------------------------- workload
def resnet50_layers() -> list[dict]:
    """Geometry of the 53 Conv2d + 1 Linear that kfac registers on torchvision.models.resnet50 (v1.5: stride on the
    3x3), in registration order. Matches the (A, G, M) multiset of SURVEY.md section 8d."""
    layers = [dict(name='conv1', kind='conv', cin=3, cout=64, k=7, s=2, p=3, h=224)]
    h, cin = 56, 64
    for li, (width, blocks, stride) in enumerate([(64, 3, 1), (128, 4, 2), (256, 6, 2), (512, 3, 2)], 1):
        for b in range(blocks):
            s = stride if b == 0 else 1
            pre = f'layer{li}.{b}'
            layers.append(dict(name=f'{pre}.conv1', kind='conv', cin=cin, cout=width, k=1, s=1, p=0, h=h))
            layers.append(dict(name=f'{pre}.conv2', kind='conv', cin=width, cout=width, k=3, s=s, p=1, h=h))
            h2 = (h + 2 - 3) // s + 1
            layers.append(dict(name=f'{pre}.conv3', kind='conv', cin=width, cout=width * 4, k=1, s=1, p=0, h=h2))
            if b == 0:
                layers.append(dict(name=f'{pre}.downsample.0', kind='conv', cin=cin, cout=width * 4, k=1, s=s, p=0, h=h))
            cin, h = width * 4, h2
    layers.append(dict(name='fc', kind='linear', cin=2048, cout=1000))
    for L in layers:
        if L['kind'] == 'conv':
            L['oh'] = (L['h'] + 2 * L['p'] - L['k']) // L['s'] + 1
            L['a'] = L['cin'] * L['k'] * L['k']
            L['g'] = L['cout']
            L['M'] = BATCH * L['oh'] * L['oh']
        else:
            L['a'] = L['cin'] + 1
            L['g'] = L['cout']
            L['M'] = BATCH
    return layers


def work_of(L: dict) -> dict:
    """Algorithmic work per layer: SYRK M n (n+1) flops per factor; eigensolver (14/3) n^3;
    precondition 4 g a (g + a)."""
    a, g, m = L['a'], L['g'], L['M']
    return {'syrk': m * a * (a + 1) + m * g * (g + 1), 'eigh': 14 / 3 * (a ** 3 + g ** 3), 'precond': 4 * g * a * (g + a)}


def make_modules(layers: list[dict]) -> torch.nn.Module:
    mods = torch.nn.ModuleDict()
    for L in layers:
        key = L['name'].replace('.', '_')
        if L['kind'] == 'conv':
            mods[key] = torch.nn.Conv2d(L['cin'], L['cout'], L['k'], stride=L['s'], padding=L['p'], bias=False)
        else:
            mods[key] = torch.nn.Linear(L['cin'], L['cout'], bias=True)
    return mods


def synth(L: dict, rank: int, device, pin: bool = False) -> dict:
    """Synthetic per-layer tensors; data is seeded per rank, gradients are common to all ranks."""
    tag = zlib.crc32(L['name'].encode()) % 1000   # stable across processes (hash() is salted)
    gd = torch.Generator().manual_seed(1000 + rank + 7919 * tag)
    gg = torch.Generator().manual_seed(4242 + tag)
    if L['kind'] == 'conv':
        x = torch.relu(torch.randn(BATCH, L['cin'], L['h'], L['h'], generator=gd))
        go = torch.randn(BATCH, L['cout'], L['oh'], L['oh'], generator=gd) / L['M']
        wg = torch.randn(L['cout'], L['cin'], L['k'], L['k'], generator=gg)
        bg = None
    else:
        x = torch.relu(torch.randn(BATCH, L['cin'], generator=gd))
        go = torch.randn(BATCH, L['cout'], generator=gd) / L['M']
        wg = torch.randn(L['cout'], L['cin'], generator=gg)
        bg = torch.randn(L['cout'], generator=gg)
    host = {'x': x, 'go': go, 'wg': wg, 'bg': bg}
    if pin:
        host = {k: (v.pin_memory() if v is not None else None) for k, v in host.items()}
    if device is None:
        return host
    dev = {k: (v.to(device) if v is not None else None) for k, v in host.items()}
    dev['host'] = host if pin else None
    return dev


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi sampled every 200 ms during the timed region (B200_PROFILING.md clocks line)."""

    Q = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, index: int) -> None:
        self.index = index
        self.proc = None
        self.lines: list[str] = []

    def start(self) -> None:
        try:
            self.proc = subprocess.Popen(['nvidia-smi', f'--query-gpu={self.Q}', '--format=csv,noheader,nounits',
                                          '-lms', '200', '-i', str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self) -> None:
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for line in self.lines:
            f = [x.strip() for x in line.split(',')]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), f[5:9]):
                if val.lower().startswith('active'):
                    reasons.add(name)
        busy = [s for s in sm if s > 0.5 * max(sm)] if sm else []
        return {'sm_mhz': statistics.median(busy) if busy else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': sorted(reasons), 'samples': len(sm)}


# --------------------------------------------------------------------------- CPU arm
REF_DIR = os.path.join(ROOT, 'oracle', '_ref')
CPU_MAX_TIMED_STEPS = 3   # a full 54-layer CPU step costs seconds; --steps is honoured up to this cap


def oracle_step(layers: list[dict], data_per_rank: list[list[dict]], state: list[dict] | None = None) -> list[dict]:
    """One factor+inverse step of the reference's algorithm (oracle/kfac_oracle.py, torch CPU ops) on `layers`.
    data_per_rank[r][i] are rank r's synthetic tensors of layer i: the factors are averaged over the ranks (the
    factor all-reduce, kfac/layers/base.py:281-335).  Returns per layer {'A', 'G', 'pg', 'grad'}; the KL-clip
    scale (base_preconditioner.py:409-433) over exactly these layers is in every entry as 'scale'."""
    from oracle import kfac_oracle as orc
    if state is None:
        state = [{'A': None, 'G': None} for _ in layers]
    world = len(data_per_rank)
    for i, (L, st) in enumerate(zip(layers, state)):
        a_sum = g_sum = None
        for r in range(world):
            d = data_per_rank[r][i]
            if L['kind'] == 'conv':
                a = orc.conv2d_a_factor(d['x'], (L['k'], L['k']), (L['s'], L['s']), (L['p'], L['p']), False)
                g = orc.conv2d_g_factor(d['go'])
            else:
                a = orc.linear_a_factor(d['x'], True)
                g = orc.linear_g_factor(d['go'])
            a = orc.running_average(st['A'], a, 1, 0.95)
            g = orc.running_average(st['G'], g, 1, 0.95)
            a_sum = a if a_sum is None else a_sum + a
            g_sum = g if g_sum is None else g_sum + g
        st['A'], st['G'] = a_sum / world, g_sum / world
    pairs = []
    for i, st in enumerate(state):
        st['da'], st['qa'] = orc.eigen_decompose(st['A'])
        st['dg'], st['qg'] = orc.eigen_decompose(st['G'])
        st['dgda'] = orc.eigen_outer(st['dg'], st['da'], 0.001)
        d = data_per_rank[0][i]
        st['grad'] = orc.stacked_grad(d['wg'], d['bg'])
        st['pg'] = orc.precondition_eigen(st['grad'], st['qa'], st['qg'], dgda=st['dgda'])
        pairs.append((st['pg'], st['grad']))
    scale = orc.kl_clip_scale(pairs, 0.1, 0.001)
    for st in state:
        st['scale'] = scale
    return state


def parity_sample(layers: list[dict]) -> list[int]:
    """Indices of the layers the in-bench parity check recomputes on the CPU: every 6th layer in registration
    order (conv1 ... the 4608 x 512 layer4 convolution) and the fc layer (2049 x 1000, the only bias column)."""
    idx = list(range(0, len(layers), 6))
    if len(layers) - 1 not in idx:
        idx.append(len(layers) - 1)
    return idx


def run_reference_arm(layers: list[dict], steps: int, warmup: int) -> dict:
    """The reference's own KFACPreconditioner (unmodified kfac_pytorch 0.4.1 installed into oracle/_ref by
    __graft_entry__.build()) on CPU tensors, all host threads, ALL 54 layers, same synthetic tensors, same
    HYBRID_OPT configuration, driven exactly like the GPU arm (hook methods, then step()).  Falls back to the
    oracle port (oracle/kfac_oracle.py) when oracle/_ref is absent."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    data = [synth(L, 0, None) for L in layers]
    kind = 'port'
    if os.path.isdir(os.path.join(REF_DIR, 'kfac')):
        kind = 'reference'
        sys.path.insert(0, REF_DIR)
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            from kfac.enums import DistributedStrategy as RefStrategy
            from kfac.preconditioner import KFACPreconditioner as RefKFAC
        model = make_modules(layers).train()
        pre = RefKFAC(model, grad_worker_fraction=RefStrategy.HYBRID_OPT, loglevel=0)
        mods = [model[L['name'].replace('.', '_')] for L in layers]

        def one_step() -> None:
            for m, d in zip(mods, data):
                m.weight.grad = d['wg'].clone()
                if d['bg'] is not None:
                    m.bias.grad = d['bg'].clone()
            for m, d in zip(mods, data):
                pre._save_input(m, (d['x'],))
            for m, d in zip(reversed(mods), reversed(data)):
                pre._save_grad_output(m, (None,), (d['go'],))
            pre.step()
    else:
        state = [{'A': None, 'G': None} for _ in layers]

        def one_step() -> None:
            oracle_step(layers, [data], state)

    for _ in range(warmup):
        one_step()
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        one_step()
        times.append(time.perf_counter() - t0)
    ms = 1e3 * sum(times) / len(times)
    what = ('unmodified kfac_pytorch 0.4.1 KFACPreconditioner (oracle/_ref)' if kind == 'reference'
            else 'oracle/kfac_oracle.py (oracle/_ref not built)')
    return {
        'value': ms, 'unit': 'ms', 'cores': cores, 'kind': kind,
        'sample': (f'all {len(layers)} layers, full steps (no sampling, no scaling): {steps} timed after {warmup} '
                   f'warm-up, mean; {what}; one CPU process with {cores} threads on rank 0 -- at n_gpus > 1 this '
                   'is still ONE process on the whole host, not a gloo world'),
        'step_ms': [1e3 * t for t in times],
    }


# --------------------------------------------------------------------------- other BASELINE.json configs
def bert_large_layers(batch: int = 8, seq: int = 512) -> list[dict]:
    """BASELINE.json configs[2]: the Linear layers of BERT-large (24 x {4 x (1024 -> 1024), 1024 -> 4096,
    4096 -> 1024}, bias=True).  BASELINE.json gives no batch: b = 8 per GPU (M = 4096 tokens), SURVEY.md section 8d."""
    layers = []
    for i in range(24):
        for j in range(4):
            layers.append(dict(name=f'enc{i}.attn{j}', kind='linear', cin=1024, cout=1024))
        layers.append(dict(name=f'enc{i}.ffn_in', kind='linear', cin=1024, cout=4096))
        layers.append(dict(name=f'enc{i}.ffn_out', kind='linear', cin=4096, cout=1024))
    for L in layers:
        L['a'], L['g'], L['M'] = L['cin'] + 1, L['cout'], batch * seq
    return layers


def _drive(pre, mods, data) -> None:
    for m, d in zip(mods, data):
        pre._save_input(m, (d['x'],))
    for m, d in zip(reversed(mods), reversed(data)):
        pre._save_grad_output(m, (None,), (d['go'],))
    pre.step()


def _timed(fn, n: int, world: int, device) -> float:
    fn()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t = a.elapsed_time(b) / n
    if world > 1:
        tt = torch.tensor([t], device=device)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t = tt.item()
    return t


def config4_inverse_mem_opt(layers, mods, data, model, set_grads, world, device) -> dict:
    """BASELINE.json configs[3]: ResNet-50, inverse method (batched Newton-Schulz), MEM_OPT, factor_update_steps=10,
    inv_update_steps=100 (examples/torch_imagenet_resnet.py:156-167): the three step kinds and the 100-iteration mean."""
    from kfac_b200 import KFACPreconditioner
    from kfac_b200.enums import ComputeMethod, DistributedStrategy
    pre = KFACPreconditioner(model, compute_method=ComputeMethod.INVERSE, grad_worker_fraction=DistributedStrategy.MEM_OPT,
                             factor_update_steps=10, inv_update_steps=100, loglevel=0)

    def step_of(kind: str):
        def run():
            pre._steps = {'factor_inverse': 0, 'factor_only': 10, 'plain': 1}[kind]
            set_grads()
            _drive(pre, mods, data)
        return run

    t_fi = _timed(step_of('factor_inverse'), 2, world, device)
    t_f = _timed(step_of('factor_only'), 2, world, device)
    t_p = _timed(step_of('plain'), 3, world, device)
    # parity of the second-order data this rank owns: Newton-Schulz vs torch.linalg.inv in fp64 on the device, and the
    # preconditioned gradient of the largest layer this rank preconditions
    worst_inv = worst_pg = 0.0
    rank = int(os.environ.get('RANK', '0'))
    lay = {id(m): layer for m, (_, layer) in pre._layers.items()}
    names = {id(m): name for m, (name, _) in pre._layers.items()}
    step_of('factor_inverse')()   # the inverses checked below must belong to the factors as they are now
    set_grads()
    mine = [(m, d) for m, d in zip(mods, data) if pre._assignment.inv_worker(names[id(m)], 'A') == rank]
    mine.sort(key=lambda md: -md[0].weight.shape[0] * md[0].weight[0].numel())
    for m, d in mine[:3]:
        layer = lay[id(m)]
        a, g = layer.a_factor.double(), layer.g_factor.double()
        ai = torch.linalg.inv(a + 0.001 * torch.eye(a.shape[0], device=device, dtype=torch.double))
        gi = torch.linalg.inv(g + 0.001 * torch.eye(g.shape[0], device=device, dtype=torch.double))
        worst_inv = max(worst_inv, _rel(layer.a_inv, ai), _rel(layer.g_inv, gi))
        grad = d['wg'].reshape(d['wg'].shape[0], -1).double()
        if d['bg'] is not None:
            grad = torch.cat([grad, d['bg'].double()[:, None]], 1)
        from kfac_b200 import ops as _ops
        got = _ops.precondition_inverse(d['wg'].reshape(d['wg'].shape[0], -1).contiguous(), d['bg'], layer.a_inv, layer.g_inv)
        worst_pg = max(worst_pg, _rel(got, gi @ grad @ ai))
    return {
        'what': 'ResNet-50, compute_method=INVERSE (batched Newton-Schulz), MEM_OPT, factor_update_steps=10, '
                'inv_update_steps=100; ms per step kind, and the mean over 100 iterations (1 + 9 + 90)',
        'factor_inverse_ms': t_fi, 'factor_only_ms': t_f, 'plain_ms': t_p,
        'mean_per_iteration_ms': (t_fi + 9 * t_f + 90 * t_p) / 100,
        'parity': {'inverse_vs_linalg_inv_fp64': worst_inv, 'preconditioned_grad_vs_fp64': worst_pg,
                   'tolerance': 1e-4, 'ok': bool(worst_inv <= 1e-4 and worst_pg <= 1e-4)},
    }


def config3_bert_large(world, device, rank) -> dict:
    """BASELINE.json configs[2]: BERT-large Linear layers, eigen method, COMM_OPT, b = 8 / GPU (stated: no batch is
    given by BASELINE.json), M = 4096 tokens; one warm-up and one timed full (factor + inverse) step."""
    from kfac_b200 import KFACPreconditioner
    from kfac_b200.enums import DistributedStrategy
    layers = bert_large_layers()
    model = make_modules(layers).to(device).train()
    mods = [model[L['name'].replace('.', '_')] for L in layers]
    gen = torch.Generator(device=device).manual_seed(77 + rank)
    # one activation / grad-output / gradient tensor per distinct shape, shared by the layers of that shape
    cache = {}
    data = []
    for L in layers:
        key = (L['cin'], L['cout'])
        if key not in cache:
            cache[key] = {'x': torch.randn(8, 512, L['cin'], device=device, generator=gen),
                          'go': torch.randn(8, 512, L['cout'], device=device, generator=gen) / L['M'],
                          'wg': torch.randn(L['cout'], L['cin'], device=device, generator=torch.Generator(device=device).manual_seed(5)),
                          'bg': torch.randn(L['cout'], device=device, generator=torch.Generator(device=device).manual_seed(6))}
        data.append(cache[key])
    pre = KFACPreconditioner(model, grad_worker_fraction=DistributedStrategy.COMM_OPT, loglevel=0)

    def run():
        for m, d in zip(mods, data):
            m.weight.grad = d['wg'].clone()
            m.bias.grad = d['bg'].clone()
        _drive(pre, mods, data)

    t = _timed(run, 1, world, device)
    # parity on one layer of each shape: the fp64 formula on the device (eigen.py:349-384 without the fp32 cast)
    worst = 0.0
    lay = {id(m): layer for m, (_, layer) in pre._layers.items()}
    scale = float(pre._scale.item())
    for i in (0, 4, 5):
        m, d = mods[i], data[i]
        layer = lay[id(m)]
        a, g = layer.a_factor.double(), layer.g_factor.double()
        da, qa = torch.linalg.eigh(a)
        dg, qg = torch.linalg.eigh(g)
        grad = torch.cat([d['wg'].double(), d['bg'].double()[:, None]], 1)
        v = qg.t() @ grad @ qa / (torch.outer(dg.clamp(min=0), da.clamp(min=0)) + 0.001)
        ref = qg @ v @ qa.t()
        got = torch.cat([m.weight.grad, m.bias.grad[:, None]], 1) / scale
        worst = max(worst, _rel(got, ref))
    del pre, model
    return {
        'what': 'BERT-large Linear layers (24 x {4 x 1024->1024, 1024->4096, 4096->1024}, bias), eigen, COMM_OPT, '
                'b = 8 per GPU x seq 512 (M = 4096; BASELINE.json gives no batch), 144 layers / 288 factors; '
                'one full factor+inverse step',
        'full_step_ms': t, 'ms_per_encoder_layer': t / 24,
        'parity': {'preconditioned_grad_vs_fp64': worst, 'tolerance': 1e-4, 'ok': bool(worst <= 1e-4)},
    }


def config5_eigensolver_sweep(device) -> dict:
    """BASELINE.json configs[4]: GPT-NeoX-style factor sizes (hidden 2048): the batched eigensolver alone on
    F = 0.95 I + 0.05 X^T X / M, X = relu(N(0,1)), M = 4n, one matrix per call, against torch.linalg.eigh (cuSOLVER)
    timed the same way: one warm-up call each, then CUDA events."""
    from kfac_b200 import ops as _ops
    rows = []
    for n in (512, 1024, 2048, 2049, 4096, 6144, 8192, 8193):
        g = torch.Generator(device=device).manual_seed(n)
        x = torch.relu(torch.randn(4 * n, n, device=device, generator=g))
        f = (0.95 * torch.eye(n, device=device) + 0.05 * (x.t() @ x) / (4 * n)).contiguous()
        del x
        q = torch.empty(n, n, device=device)
        d = torch.empty(n, device=device)

        def ours():
            _ops.eigh_batched([f], qs=[q], ds=[d])

        def cusolver():
            torch.linalg.eigh(f)

        t_ours = _timed(ours, 2, 1, device)
        t_ref = _timed(cusolver, 1, 1, device)
        qd, dd, fd = q.double(), d.double(), f.double()
        rec = (torch.linalg.norm(qd @ torch.diag(dd) @ qd.t() - fd) / torch.linalg.norm(fd)).item()
        orth = (qd.t() @ qd - torch.eye(n, device=device, dtype=torch.double)).abs().max().item()
        rows.append({'n': n, 'ms': t_ours, 'cusolver_syevd_ms': t_ref, 'reconstruction': rec, 'orthogonality': orth})
        del qd, fd
    return {'what': 'eigensolver sweep, batch 1, fp32, both arms warmed and timed with CUDA events', 'sizes': rows,
            'parity_ok': bool(all(r['reconstruction'] <= 1e-5 and r['orthogonality'] <= 1e-5 for r in rows))}


def _rel(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return (torch.linalg.norm(a - b) / torch.linalg.norm(b)).item()


def _rel_offdiag(a: torch.Tensor, b: torch.Tensor) -> float:
    """Relative error of the off-diagonal part.  After one running-average step from the identity a factor is
    0.95 I + 0.05 C; the reference divides convolution factors by the spatial size twice (modules.py:176,190), so C
    is tiny next to the identity and a whole-matrix error would only measure the 0.95 on the diagonal."""
    a, b = a.detach().double().cpu().clone(), b.detach().double().cpu().clone()
    a.fill_diagonal_(0.0)
    b.fill_diagonal_(0.0)
    return (torch.linalg.norm(a - b) / torch.linalg.norm(b)).item()


# --------------------------------------------------------------------------- GPU arm
def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-e2e', action='store_true')
    ap.add_argument('--no-parity', action='store_true')
    ap.add_argument('--no-extra-configs', action='store_true', help='skip BASELINE.json configs 3, 4 and 5')
    ap.add_argument('--profile-step', action='store_true',
                    help='for ncu launch lists: 1 warm-up step, --steps steps, nothing else; prints no bench line')
    ap.add_argument('--eigh-mode', type=int, default=0, help='kfac_eigh_set_mode: 0 by size, 1 one-stage, 2 two-stage')
    args = ap.parse_args()
    warmup = max(args.warmup, 3) if args.impl == 'ours' else max(args.warmup, 1)
    if args.profile_step:
        warmup, args.no_e2e = 1, True
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    layers = resnet50_layers()
    config = {
        'workload': 'ResNet-50 ImageNet shapes, synthetic 224x224, batch 32/GPU, eigen method, '
                    'grad_worker_fraction=HYBRID_OPT (BASELINE.json configs[1]); 54 layers, 108 factors; '
                    'every step is a factor+inverse step',
        'global_batch': BATCH * world, 'parallelism': f'dp{world}',
        'l2': 'inputs per step (2.7 GB activations+grad-outputs) exceed the 126 MB L2; no explicit flush',
    }

    if args.impl == 'reference':
        if rank != 0:
            return
        steps = max(1, min(args.steps, CPU_MAX_TIMED_STEPS))
        wu = 1
        res = run_reference_arm(layers, steps, wu)
        print(json.dumps({
            'impl': 'reference', 'metric': METRIC, 'value': res['value'], 'unit': 'ms', 'n_gpus': args.gpus,
            'steps': steps, 'warmup': wu, 'ms_per_step': res['value'], 'higher_is_better': False,
            'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic', 'config': config,
            'cpu_baseline': res,
            'e2e': {'value': res['value'], 'unit': 'ms', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'gpu_launches': 0,
        }))
        return

    if not torch.cuda.is_available():
        raise RuntimeError('bench.py --impl ours needs CUDA (there is no CPU fallback); use --impl reference')
    torch.cuda.set_device(local_rank)
    device = torch.device('cuda', local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=device)
    from kfac_b200 import _native
    from kfac_b200 import KFACPreconditioner
    from kfac_b200.enums import DistributedStrategy
    lib = _native.load()
    _native.check(lib.kfac_eigh_set_mode(args.eigh_mode))

    model = make_modules(layers).to(device).train()
    pre = KFACPreconditioner(model, grad_worker_fraction=DistributedStrategy.HYBRID_OPT, loglevel=0)
    mods = [model[L['name'].replace('.', '_')] for L in layers]
    want_e2e = not args.no_e2e
    data = [synth(L, rank, device, pin=want_e2e) for L in layers]
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731

    def set_grads() -> None:
        for m, d in zip(mods, data):
            m.weight.grad = d['wg'].clone()
            if d['bg'] is not None:
                m.bias.grad = d['bg'].clone()

    stage_ms = {'factor': 0.0, 'second_order': 0.0, 'precondition': 0.0}
    eigh_ms = [0.0, 0.0, 0.0]   # tridiagonalisation / divide & conquer / back-transformation (library stage timers)
    sytrd_launches = [0]
    eigh_detail = {'band': 0.0, 'chase': 0.0, 'q2': 0.0, 'q1': 0.0}
    eigh_counts: dict[str, float] = {}
    import ctypes
    lib.kfac_profile_enable(1)

    marks: dict[str, torch.cuda.Event] = {}
    pre._stage_hook = lambda name: marks.__setitem__(name, _mark())

    def _mark() -> torch.cuda.Event:
        e = ev()
        e.record()
        return e

    def one_step(timed: bool) -> None:
        e0 = _mark()
        for m, d in zip(mods, data):                      # forward hooks: A factors (+ running average, + allreduce)
            pre._save_input(m, (d['x'],))
        for m, d in zip(reversed(mods), reversed(data)):  # backward hooks: G factors
            pre._save_grad_output(m, (None,), (d['go'],))
        pre.step()                                        # flush -> eigen -> broadcasts -> precondition -> clip -> write
        if timed:
            torch.cuda.synchronize()
            stage_ms['factor'] += e0.elapsed_time(marks['second_order'])
            stage_ms['second_order'] += marks['second_order'].elapsed_time(marks['precondition'])
            stage_ms['precondition'] += marks['precondition'].elapsed_time(marks['end'])
            out4 = (ctypes.c_float * 4)()
            if lib.kfac_profile_last_ms(out4) == 0:   # this rank ran a batched eigendecomposition in this step
                for i in range(3):
                    eigh_ms[i] += out4[i]
                sytrd_launches[0] = int(out4[3])
            out6 = (ctypes.c_float * 6)()
            if lib.kfac_profile_detail_ms(out6) == 0:   # ... and some of its matrices took the two-stage path
                for i, k in enumerate(('band', 'chase', 'q2', 'q1')):
                    eigh_detail[k] += out6[i]
                eigh_counts['chase'], eigh_counts['q2'] = out6[4], out6[5]
                p4 = (ctypes.c_int * 4)()
                lib.kfac_eigh_last_paths(p4)
                eigh_counts['two_stage'] = p4[3]

    def barrier() -> None:
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        set_grads()
        one_step(False)
    set_grads()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = lib.kfac_launch_count()
    t_begin, t_end = ev(), ev()
    t_begin.record()
    for _ in range(args.steps):
        one_step(True)
    t_end.record()
    barrier()
    ms = t_begin.elapsed_time(t_end) / args.steps
    launches = (lib.kfac_launch_count() - launches0) // args.steps
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([ms], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = t.item()

    if args.profile_step:
        if rank == 0:
            print(json.dumps({'profile_step': True, 'ms_per_step': ms, 'gpu_launches': int(launches),
                              'stage_ms': {k: v / args.steps for k, v in stage_ms.items()},
                              'note': 'run under a profiler for launch lists: not a bench value'}))
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- e2e: same step fed from pinned host buffers, result read back
    e2e = None
    if want_e2e:
        h2d = sum(v.numel() * 4 for d in data for v in d['host'].values() if v is not None)
        norms = torch.zeros(len(mods), device=device)
        host_out = torch.zeros(len(mods) + 1).pin_memory()

        copy_stream = torch.cuda.Stream(device=device)

        def e2e_step() -> None:
            # host -> device copies run on their own stream in the order the hooks consume them (activations in
            # forward order, grad-outputs in backward order, gradients last); each hook waits only for its own
            # tensor, so the factor kernels overlap the remaining copies
            main = torch.cuda.current_stream()
            copy_stream.wait_stream(main)          # the previous step no longer reads the device buffers
            ev_x, ev_go = [], []
            with torch.cuda.stream(copy_stream):
                for d in data:
                    d['x'].copy_(d['host']['x'], non_blocking=True)
                    ev_x.append(copy_stream.record_event())
                for d in reversed(data):
                    d['go'].copy_(d['host']['go'], non_blocking=True)
                    ev_go.append(copy_stream.record_event())
                for m, d in zip(mods, data):
                    if m.weight.grad is None:
                        m.weight.grad = torch.empty_like(m.weight)
                    m.weight.grad.copy_(d['host']['wg'], non_blocking=True)
                    if d['bg'] is not None:
                        if m.bias.grad is None:
                            m.bias.grad = torch.empty_like(m.bias)
                        m.bias.grad.copy_(d['host']['bg'], non_blocking=True)
                ev_grads = copy_stream.record_event()
            for m, d, e in zip(mods, data, ev_x):
                main.wait_event(e)
                pre._save_input(m, (d['x'],))
            for m, d, e in zip(reversed(mods), reversed(data), ev_go):
                main.wait_event(e)
                pre._save_grad_output(m, (None,), (d['go'],))
            main.wait_event(ev_grads)
            pre.step()
            for i, m in enumerate(mods):
                norms[i] = m.weight.grad.norm()
            host_out[:-1].copy_(norms, non_blocking=True)
            host_out[-1:].copy_(pre._scale, non_blocking=True)
            torch.cuda.current_stream().synchronize()

        def time_e2e(step_fn, n: int) -> float:
            barrier()
            t0, t1 = ev(), ev()
            t0.record()
            for _ in range(n):
                step_fn()
            t1.record()
            barrier()
            t_ms = t0.elapsed_time(t1) / n
            if world > 1:
                t = torch.tensor([t_ms], device=device)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                t_ms = t.item()
            return t_ms

        e2e_step()
        n_e2e = max(1, min(args.steps, 3))
        serial_ms = time_e2e(e2e_step, n_e2e)

        # The same step with the input pipeline a training loop has (DataLoader prefetch + non_blocking copies): TWO
        # sets of device input buffers, the copies of step k+1 are issued when step k starts and run on the copy
        # stream under step k's kernels.  Every step still copies all of its inputs from pinned host memory and reads
        # its result back inside the timed region; only the order of issue differs from `serial_ms`.
        slots = [
            [{'x': d['x'], 'go': d['go'], 'wg': m.weight.grad, 'bg': m.bias.grad if d['bg'] is not None else None}
             for m, d in zip(mods, data)],
            [{'x': torch.empty_like(d['x']), 'go': torch.empty_like(d['go']), 'wg': torch.empty_like(m.weight),
              'bg': torch.empty_like(m.bias) if d['bg'] is not None else None} for m, d in zip(mods, data)],
        ]
        ready: list[dict | None] = [None, None]
        state = {'k': 0, 'done': None}

        def issue_copies(slot: int) -> None:
            if state['done'] is not None:
                copy_stream.wait_event(state['done'])      # the step that last read this slot has finished
            evs = {'x': [], 'go': []}
            with torch.cuda.stream(copy_stream):
                for d, sl in zip(data, slots[slot]):
                    sl['x'].copy_(d['host']['x'], non_blocking=True)
                    evs['x'].append(copy_stream.record_event())
                for d, sl in zip(reversed(data), reversed(slots[slot])):
                    sl['go'].copy_(d['host']['go'], non_blocking=True)
                    evs['go'].append(copy_stream.record_event())
                for d, sl in zip(data, slots[slot]):
                    sl['wg'].copy_(d['host']['wg'], non_blocking=True)
                    if sl['bg'] is not None:
                        sl['bg'].copy_(d['host']['bg'], non_blocking=True)
                evs['grads'] = copy_stream.record_event()
            ready[slot] = evs

        def e2e_step_pipelined() -> None:
            main = torch.cuda.current_stream()
            slot = state['k'] & 1
            if ready[slot] is None:
                issue_copies(slot)
            evs, ready[slot] = ready[slot], None
            issue_copies(slot ^ 1)                          # next step's inputs, under this step's kernels
            cur = slots[slot]
            for m, sl, e in zip(mods, cur, evs['x']):
                main.wait_event(e)
                pre._save_input(m, (sl['x'],))
            for m, sl, e in zip(reversed(mods), reversed(cur), evs['go']):
                main.wait_event(e)
                pre._save_grad_output(m, (None,), (sl['go'],))
            main.wait_event(evs['grads'])
            for m, sl in zip(mods, cur):
                m.weight.grad = sl['wg']
                if sl['bg'] is not None:
                    m.bias.grad = sl['bg']
            pre.step()
            for i, m in enumerate(mods):
                norms[i] = m.weight.grad.norm()
            host_out[:-1].copy_(norms, non_blocking=True)
            host_out[-1:].copy_(pre._scale, non_blocking=True)
            state['done'] = main.record_event()
            state['k'] += 1
            main.synchronize()

        e2e_step_pipelined()
        e2e_ms = time_e2e(e2e_step_pipelined, n_e2e)
        torch.cuda.synchronize()
        for m, sl in zip(mods, slots[0]):                   # the sections below use the original buffers again
            m.weight.grad = sl['wg']
            if sl['bg'] is not None:
                m.bias.grad = sl['bg']
        del slots
        e2e = {'value': e2e_ms, 'unit': 'ms', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': host_out.numel() * 4,
               'steps': n_e2e, 'serial_ms': serial_ms,
               'how': 'value: double-buffered inputs, the host-to-device copies of step k+1 run on a copy stream under '
                      'the kernels of step k (every step copies all its inputs and reads its result back inside the '
                      'timed region); serial_ms: the copies of a step start only after the previous step has ended'}

    # ---- the other two cadences of SURVEY 8(d): factor-only step (no eigendecomposition) and plain step
    # (precondition with the existing second-order data only); the headline above is the factor+inverse step
    def timed_steps(n: int) -> float:
        set_grads()
        one_step(False)
        barrier()
        a, b = ev(), ev()
        a.record()
        for _ in range(n):
            one_step(False)
        b.record()
        barrier()
        t_ms = a.elapsed_time(b) / n
        if world > 1:
            tt = torch.tensor([t_ms], device=device)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t_ms = tt.item()
        return t_ms

    keep = (pre._factor_update_steps, pre._inv_update_steps)
    pre._inv_update_steps = 1 << 30
    factor_only_ms = timed_steps(3)
    pre._factor_update_steps = 1 << 30
    plain_ms = timed_steps(3)
    pre._factor_update_steps, pre._inv_update_steps = keep
    cadences = {'factor_inverse_ms': ms, 'factor_only_ms': factor_only_ms, 'plain_ms': plain_ms}

    # ---- memory-bound passes and the factor all-reduce, timed alone (north star: reported every run)
    def gbs_of(fn, nbytes: float, iters: int = 5) -> dict:
        fn()
        torch.cuda.synchronize()
        a, b = ev(), ev()
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        torch.cuda.synchronize()
        t_ms = a.elapsed_time(b) / iters
        return {'ms': t_ms, 'gbs': nbytes / t_ms / 1e6}

    from kfac_b200 import ops as _ops
    arena = pre._factor_arena if hasattr(pre, '_factor_arena') else None
    n_fac = sum(L['a'] ** 2 + L['g'] ** 2 for L in layers)
    ew_x = torch.empty(n_fac, device=device).normal_()
    ew_y = torch.empty(n_fac, device=device).normal_()
    n_grad = sum(L['a'] * L['g'] for L in layers)          # 25.5 M gradient floats: larger than the L2 with v
    gw = torch.randn(5000, n_grad // 5000, device=device)
    gv = torch.randn(5000, n_grad // 5000, device=device)
    acc = torch.zeros(1, dtype=torch.float64, device=device)
    one = torch.ones(1, device=device)
    elementwise = {
        # running average over all 108 factors at once: read F, read C, write F = 12 B per element
        'axpby_factor_arena': gbs_of(lambda: _ops.axpby(ew_y, ew_x, 0.05, 0.95), 12.0 * n_fac),
        # KL-clip dot and scale+write over a gradient-sized matrix (all layers' worth): 8 B per element each
        'grad_dot': gbs_of(lambda: _ops.grad_dot(gv, gw, None, acc), 8.0 * gw.numel()),
        'scale_write_grads': gbs_of(lambda: _ops.scale_write_grads(gv, one, gw, None), 8.0 * gw.numel()),
    }
    nf = 4608
    tf = torch.randn(nf, nf, device=device)
    tp = torch.empty(nf * (nf + 1) // 2, device=device)
    dgv, dav = torch.rand(4096, device=device), torch.rand(8192, device=device)
    eo = torch.empty(4096, 8192, device=device)
    elementwise.update({
        # symmetry-aware communication: read the upper triangle, write it packed (and back): 8 B per packed element
        'triu_pack_4608': gbs_of(lambda: _ops.triu_pack(tf, tp), 8.0 * tp.numel()),
        'triu_unpack_4608': gbs_of(lambda: _ops.triu_unpack(tp, nf, tf), 4.0 * tp.numel() + 4.0 * nf * nf),
        # 1 / (dg x da + damping): write-only, 4 B per element (134 MB, larger than the L2)
        'eigen_outer_4096x8192': gbs_of(lambda: _ops.eigen_outer(dgv, dav, 0.001, eo), 4.0 * eo.numel()),
        # the ceiling of a WRITE-ONLY pass over the same buffer (torch fill): the copy peak counts read + write bytes
        'fill_4096x8192_torch': gbs_of(lambda: eo.fill_(1.0), 4.0 * eo.numel()),
    })
    del ew_x, ew_y, tf, tp, eo
    # ---- tensor-core ceilings measured here (rank 0 of a 1-GPU run): cuBLAS TF32 on 8192^3 is what ONE tf32 product
    # costs at library quality; a 3xTF32 product cannot beat a third of it; gemm3x on the same problem beside it
    tensor_peaks = None
    if world == 1:
        n_g = 8192
        ga, gb2 = torch.randn(n_g, n_g, device=device), torch.randn(n_g, n_g, device=device)
        gc = torch.empty(n_g, n_g, device=device)
        keep_tf32 = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = True
        t_tf32 = gbs_of(lambda: torch.matmul(ga, gb2, out=gc), 0.0, iters=5)['ms']
        torch.backends.cuda.matmul.allow_tf32 = False
        t_fp32 = gbs_of(lambda: torch.matmul(ga, gb2, out=gc), 0.0, iters=2)['ms']
        torch.backends.cuda.matmul.allow_tf32 = keep_tf32
        t_3x = gbs_of(lambda: _ops.gemm(ga, gb2, out=gc), 0.0, iters=5)['ms']
        fl = 2.0 * n_g ** 3 / 1e9
        tensor_peaks = {'problem': '8192^3 fp32 operands', 'cublas_tf32_tflops': fl / t_tf32,
                        'cublas_fp32_simt_tflops': fl / t_fp32, 'gemm3x_tflops': fl / t_3x,
                        'gemm3x_frac_of_3xtf32_ceiling': (fl / t_3x) / (fl / t_tf32 / 3.0),
                        'note': '3xTF32 ceiling = cuBLAS TF32 / 3 (three tf32 products per fp32-grade product)'}
        del ga, gb2, gc
    allreduce = None
    if world > 1:
        buf = torch.empty(n_fac, device=device).normal_()
        dist.all_reduce(buf)
        barrier()
        a, b = ev(), ev()
        a.record()
        for _ in range(3):
            dist.all_reduce(buf, op=dist.ReduceOp.AVG)
        b.record()
        barrier()
        t_ms = a.elapsed_time(b) / 3
        tt = torch.tensor([t_ms], device=device)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_ms = tt.item()
        allreduce = {'bytes': 4 * n_fac, 'ms': t_ms, 'algbw_gbs': 4 * n_fac / t_ms / 1e6,
                     'busbw_gbs': 2 * (world - 1) / world * 4 * n_fac / t_ms / 1e6,
                     'what': 'one NCCL all-reduce (AVG) of all 108 factors (the arena size), timed alone'}
        del buf

    # ---- the two grouped KAISA exchanges, timed alone on the arenas of the last step (N > 1): one broadcast per
    # (group, source rank); bytes = what the busiest rank sends or receives, so gbs is a broadcast bus bandwidth
    exchange_bw = None
    if world > 1 and getattr(pre, '_exchange', None) is not None:
        ex = pre._exchange

        def time_exchange(fn, arenas) -> dict:
            nbytes = float(sum(a.numel for a in arenas) * 4)
            fn()
            barrier()
            a, b = ev(), ev()
            a.record()
            for _ in range(3):
                fn()
            b.record()
            barrier()
            tt = torch.tensor([a.elapsed_time(b) / 3, nbytes], device=device)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t_ms, mx = tt[0].item(), tt[1].item()
            return {'ms': t_ms, 'max_bytes_per_rank': mx, 'gbs': mx / t_ms / 1e6 if t_ms > 0 else None,
                    'collectives_on_rank0': len(arenas)}

        exchange_bw = {'second_order': time_exchange(ex.broadcast_second_order, ex.second_order),
                       'gradients': time_exchange(ex.broadcast_gradients, ex.gradients)}

    # ---- parity, outside the timed region, at every N: a FRESH preconditioner's first factor+inverse step on the
    # same synthetic tensors, against the CPU oracle on a sample of the layers.  Rank 0 regenerates every rank's
    # data on the host (the seeds are a function of the rank), so the oracle's factors are the all-reduced ones, and
    # the gradients it checks on rank 0 arrived through whatever broadcasts the KAISA placement prescribes.
    parity = None
    if not args.no_parity:
        lib.kfac_profile_enable(0)
        pre._stage_hook = None
        pre2 = KFACPreconditioner(model, grad_worker_fraction=DistributedStrategy.HYBRID_OPT, loglevel=0)
        set_grads()
        for m, d in zip(mods, data):
            pre2._save_input(m, (d['x'],))
        for m, d in zip(reversed(mods), reversed(data)):
            pre2._save_grad_output(m, (None,), (d['go'],))
        pre2.step()
        torch.cuda.synchronize()
        # every rank must end with the same gradients (bit-identical: all ranks run the same kernels or receive copies)
        flat = torch.cat([m.weight.grad.flatten() for m in mods])
        same = 0.0
        if world > 1:
            ref0 = flat.clone()
            dist.broadcast(ref0, src=0)
            diff = ((flat - ref0).norm() / ref0.norm()).reshape(1)
            dist.all_reduce(diff, op=dist.ReduceOp.MAX)
            same = diff.item()
        if rank == 0:
            t0 = time.perf_counter()
            idx = parity_sample(layers)
            sub = [layers[i] for i in idx]
            host = [[synth(L, r, None) for L in sub] for r in range(world)]
            torch.set_num_threads(os.cpu_count() or 1)
            exp = oracle_step(sub, host)
            scale_gpu = float(pre2._scale.item())
            layer_objs = {name: layer for name, layer in pre2._layers.values()}
            by_mod = {id(m): name for m, (name, _) in pre2._layers.items()}
            worst_f = worst_g = 0.0
            detail = []
            for i, L, st in zip(idx, sub, exp):
                m = mods[i]
                layer = layer_objs[by_mod[id(m)]]
                ef = max(_rel(layer.a_factor, st['A']), _rel(layer.g_factor, st['G']),
                         _rel_offdiag(layer.a_factor, st['A']), _rel_offdiag(layer.g_factor, st['G']))
                got = m.weight.grad.reshape(m.weight.shape[0], -1)
                if m.bias is not None:
                    got = torch.cat([got, m.bias.grad.reshape(-1, 1)], 1)
                eg = _rel(got / scale_gpu, st['pg'])
                worst_f, worst_g = max(worst_f, ef), max(worst_g, eg)
                detail.append({'layer': L['name'], 'a': L['a'], 'g': L['g'], 'factor_rel': ef, 'grad_rel': eg})
            # the KL-clip scale covers all 54 layers: recompute it from the device gradients with torch ops
            vg = 0.0
            for m, d in zip(mods, data):
                vg += float((m.weight.grad.double() / scale_gpu * d['wg'].double()).sum()) * 0.1 ** 2
                if d['bg'] is not None:
                    vg += float((m.bias.grad.double() / scale_gpu * d['bg'].double()).sum()) * 0.1 ** 2
            scale_ref = 1.0 if vg == 0 else min(1.0, (0.001 / abs(vg)) ** 0.5)
            parity = {
                'parity_max_rel': worst_g, 'factor_max_rel': worst_f, 'ranks_identical_rel': same,
                'klclip_scale': scale_gpu, 'klclip_scale_recomputed': scale_ref,
                'tolerance': {'grad': 1e-4, 'factor': 1e-5},
                'ok': bool(worst_g <= 1e-4 and worst_f <= 1e-5 and same <= 1e-6
                           and abs(scale_gpu - scale_ref) <= 1e-4 * scale_ref),
                'what': (f'first factor+inverse step of a fresh KFACPreconditioner (HYBRID_OPT, world {world}) vs '
                         f'oracle/kfac_oracle.py on {len(idx)} of {len(layers)} layers (every 6th + fc), factors '
                         'averaged over all ranks\' data; relative Frobenius error of the factors (whole matrix and off-diagonal part, the larger) and of the '
                         'preconditioned gradients (before the KL-clip scale)'),
                'layers': detail, 'cpu_seconds': time.perf_counter() - t0,
            }
        del pre2

    # ---- the other BASELINE.json configs, each with its own parity figure (outside the headline's timed region)
    extra = None
    if not args.no_extra_configs:
        lib.kfac_profile_enable(0)
        pre._stage_hook = None
        extra = {}
        extra['config4_resnet50_inverse_mem_opt'] = config4_inverse_mem_opt(layers, mods, data, model, set_grads,
                                                                            world, device)
        torch.cuda.empty_cache()
        extra['config3_bert_large_comm_opt'] = config3_bert_large(world, device, rank)
        torch.cuda.empty_cache()
        if world == 1:
            extra['config5_eigensolver_sweep'] = config5_eigensolver_sweep(device)
        extra['config1_resnet32_cifar'] = ('parity-only config (the reference runs it on one CPU process): '
                                           'tests/test_fullsize_gpu.py::test_resnet32_cifar_config1_end_to_end')

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant stage's kernel
    peaks = {'hbm_gbs': 6650.0, 'bf16_tflops': 1590.0, 'bf16_tflops_sustained': 1400.0, 'source': 'fallback'}
    pk_path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(pk_path):
        with open(pk_path) as f:
            peaks.update(json.load(f))
        peaks['source'] = 'measured'
    per = {k: v / args.steps for k, v in stage_ms.items()}
    # which layers this rank owns in each stage
    asg = pre._assignment
    names = [n for n, _ in pre._layers.values()]
    by_name = dict(zip(names, layers))
    syrk_flops = sum(work_of(L)['syrk'] for L in layers)
    my_eigh_n3 = sum(by_name[n]['a'] ** 3 for n in names if asg.inv_worker(n, 'A') == rank) + \
        sum(by_name[n]['g'] ** 3 for n in names if asg.inv_worker(n, 'G') == rank)
    my_precond = sum(work_of(by_name[n])['precond'] for n in names if asg.is_grad_worker(n))
    stages = {
        'factor': {'ms': per['factor'], 'kernel': 'gemm3x_kernel (tcgen05 3xTF32 SYRK, fused im2col + running average)',
                   'tflops_algorithmic': syrk_flops / per['factor'] / 1e9 if per['factor'] else None},
        'second_order': {'ms': per['second_order'], 'kernel': 'sytrd_kernel + dc_* + gemm3x (batched eigensolver)',
                         'declared_tflops': 14 / 3 * my_eigh_n3 / per['second_order'] / 1e9 if per['second_order'] else None,
                         'symv_floor_gbs': (2 / 3) * my_eigh_n3 / per['second_order'] / 1e6 if per['second_order'] else None},
        'precondition': {'ms': per['precondition'], 'kernel': 'gemm3x_kernel x4 per layer + klclip + scale_write',
                         'tflops_algorithmic': my_precond / per['precondition'] / 1e9 if per['precondition'] else None},
    }
    for v in elementwise.values():
        v['frac_of_hbm_peak'] = v['gbs'] / peaks['hbm_gbs']
    sy_ms, dc_ms, bt_ms = (v / args.steps for v in eigh_ms)
    stages['second_order'].update({'tridiagonalise_ms': sy_ms, 'divide_conquer_ms': dc_ms, 'back_transform_ms': bt_ms})
    # DRAM bytes of the dominant kernel's launches on this workload, from one `ncu --set full` capture
    # (profiles/r02_q2_traffic.json, written by tools/ncu_traffic.py): per launch, like `achieved`
    traffic = None
    tr_path = os.path.join(ROOT, 'profiles', 'r02_q2_traffic.json')
    if world == 1 and os.path.exists(tr_path):
        with open(tr_path) as f:
            traffic = json.load(f).get('dram_bytes_per_launch')
    dominant = max(per, key=per.get)
    if dominant == 'second_order':
        det = {k: v / args.steps for k, v in eigh_detail.items()}
        # matrices that took the two-stage path on this rank (what ts_* kernels process)
        mine = [by_name[n]['a'] for n in names if asg.inv_worker(n, 'A') == rank] + \
               [by_name[n]['g'] for n in names if asg.inv_worker(n, 'G') == rank]
        ts_ns = sorted(mine, reverse=True)[:int(eigh_counts.get('two_stage', 0))]
        # Q2 back-transformation (ts_q2_kernel): Z (n x n fp32) is read and written once per group of TS_G = 32
        # sweeps, n / 32 groups: 8 n^2 * n / 32 = n^3 / 4 bytes per matrix -- the kernel's own streaming requirement
        # (DESIGN.md "Q2"); launches are sized so that a launch's matrices fit the L2, so DRAM traffic is far lower.
        q2_bytes = sum(n ** 3 / 4 for n in ts_ns)
        q2_ms = det.get('q2', 0.0)
        n_q2 = max(1, int(eigh_counts.get('q2', 1)))
        kernels = {
            'band_reduction (ts_panel + ts_av + ts_g0 + ts_w + ts_syr2k + ts_band)': {
                'ms': det.get('band'), 'bound': 'hbm/L2 (A22 read once per panel by ts_av and once by ts_syr2k)',
                'algorithmic_gbs': sum(n ** 3 / 12 for n in ts_ns) / det['band'] / 1e6 if det.get('band') else None,
                'note': 'bytes = 2 passes x sum_panels 4 m^2 = n^3 / 12 per matrix'},
            'ts_chase_kernel': {
                'ms': det.get('chase'), 'launches': int(eigh_counts.get('chase', 0)),
                'bound': 'latency (dependent chain of 32 x 32 reflector steps; see DESIGN.md "Bulge chasing")',
                'flops_gf': sum(6 * 32 * n * n for n in ts_ns) / det['chase'] / 1e6 if det.get('chase') else None},
            'ts_q2_kernel': {
                'ms': q2_ms, 'launches': n_q2, 'bound': 'L2/hbm streaming',
                'achieved_gbs': q2_bytes / q2_ms / 1e6 if q2_ms else None},
            'back_transform_q1 (gemm3x grouped, compact WY)': {'ms': det.get('q1'), 'bound': 'tensor (3xTF32)'},
            'divide_conquer (dc_* + gemm3x grouped)': {'ms': dc_ms, 'bound': 'tensor (3xTF32) + secular equations'},
        }
        stages['second_order']['kernels'] = kernels
        n_launch = max(1, sytrd_launches[0] + int(eigh_counts.get('chase', 0)))
        alg_bytes = (2 / 3) * my_eigh_n3              # (2/3) n^3 bytes per factor (symmetric symv floor)
        t_kernel = sy_ms if sy_ms > 0 else per['second_order']
        achieved = alg_bytes / t_kernel / 1e6
        roofline = {'bound': 'hbm', 'kernel': 'tridiagonalisation: ts_panel/ts_av/ts_w/ts_syr2k (band reduction) + '
                                               'ts_chase_kernel (two-stage, n >= 1000) and sytrd_kernel (one-stage)',
                    'achieved': achieved, 'peak': peaks['hbm_gbs'], 'unit': 'GB/s',
                    'frac': achieved / peaks['hbm_gbs'], 'traffic': traffic,
                    'stage_ms': t_kernel, 'cooperative_launches_per_step': n_launch,
                    'algorithmic_bytes_per_step': alg_bytes,
                    'note': 'algorithmic bytes = (2/3) n^3 per factor (the one-stage symmetric symv floor, SURVEY 8d) '
                            'summed over the step\'s factors, divided by the duration of the whole tridiagonalisation '
                            'stage measured with CUDA events on the launching stream inside the library '
                            '(kfac_profile_last_ms).  The two-stage path does NOT move those bytes (its band '
                            'reduction reads the trailing matrix twice per 32 columns: n^3/12 bytes), so a fraction '
                            'above what one-stage can reach is the point of it; stages.second_order.kernels has '
                            'each kernel\'s own figure; `traffic` = DRAM bytes per ts_q2_kernel launch (ncu).  '
                            'peak ' + peaks['source']}
    else:
        achieved = stages[dominant]['tflops_algorithmic']
        roofline = {'bound': 'tensor', 'kernel': 'gemm3x_kernel', 'achieved': achieved, 'peak': peaks['bf16_tflops_sustained'],
                    'unit': 'TFLOP/s', 'frac': achieved / peaks['bf16_tflops_sustained'], 'traffic': None,
                    'note': 'peak = measured bf16 (sustained); the kernel issues 3 TF32 UMMAs per product, so its '
                            'own ceiling is bf16/6; peak ' + peaks['source']}

    # collectives one factor+inverse step issues on this rank (the reference: 1 all-reduce per factor unless bucketed,
    # 2-3 broadcasts per layer for the second-order data, 1 per layer for the gradients)
    collectives = None
    if world > 1:
        collectives = {'factor_allreduce_calls_total': pre._tdc.stats.get('allreduce_calls'),
                       'broadcast_calls_total': pre._tdc.stats.get('broadcast_calls'),
                       'steps_counted': int(pre.steps)}
        if pre._exchange is not None:
            collectives.update(pre._exchange.collectives_per_step())
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cpu = run_reference_arm(layers, 1, 0)

    line = {
        'metric': METRIC, 'value': ms, 'unit': 'ms', 'n_gpus': world, 'steps': args.steps, 'warmup': warmup,
        'ms_per_step': ms, 'higher_is_better': False, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32',
        'data': 'synthetic', 'config': config, 'clocks': clocks, 'e2e': e2e, 'gpu_launches': int(launches),
        'roofline': roofline, 'stages': stages, 'cadences': cadences, 'elementwise': elementwise,
        'allreduce': allreduce, 'tensor_peaks': tensor_peaks, 'collectives_per_step': collectives, 'exchange': exchange_bw, 'cpu_baseline': cpu, 'parity': parity, 'configs': extra,
        'parity_max_rel': parity['parity_max_rel'] if parity else None,
        'factor_tflops': stages['factor']['tflops_algorithmic'],
    }
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
