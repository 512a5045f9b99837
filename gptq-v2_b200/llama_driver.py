"""Layer-by-layer GPTQ of a Llama model on the B200 hot path  (BASELINE configs 2-4).

Follows the flow of the reference's quantize_llama (/root/reference/llama.py:65-204): for every decoder block, the
four Hessian groups are processed in dependency order

    [v, k, q] <- input_layernorm(x)     [o] <- attention output     [up, gate] <- post_attention_layernorm(h)
    [down] <- silu(gate) * up

each group accumulates its Hessian over all calibration batches (HessianHook.add_batch, llama.py:121-125 /
model_utils.py:101-103), inverts it (llama.py:138), quantises its linears (llama.py:148-163), rebuilds them
(reconstruct_nn_linear, llama.py:172) and replays the calibration batches through the QUANTISED weights so that later
groups see error-compensated activations (llama.py:176-177).  What is different from the reference:

  * the block structure of Llama is written down explicitly instead of being rediscovered per block by tracing the
    autograd graph under FakeTensorMode (model_utils.py:146-248, H9 in SURVEY.md: identical for all blocks);
  * activations stay on the GPU (the reference's save_gpu_mem=False setting) as whole tensors, not Python lists;
  * with torch.distributed initialised, calibration sequences are sharded over ranks, Hessians are all-reduced and the
    rows of every linear are sharded (sharded.py); every rank keeps its own sequences through all layers.

The transformer forward itself (RMSNorm, RoPE, SDPA, SwiGLU, dense replay GEMMs) is plain PyTorch: it is outside the
hot path ("caller of the hot path").
"""
from __future__ import annotations

import time
from concurrent.futures import Future, ThreadPoolExecutor
from dataclasses import dataclass
from typing import Callable, Optional

import torch
import torch.nn.functional as F

import sharded
from gptq_py import HessianHook, gptq_quant
from quant import reconstruct_nn_linear


@dataclass
class LlamaShape:
    hidden: int = 4096
    inter: int = 11008
    layers: int = 32
    heads: int = 32
    vocab: int = 32000
    rms_eps: float = 1e-5
    rope_theta: float = 10000.0

    @staticmethod
    def llama2_7b() -> 'LlamaShape':
        return LlamaShape()

    @staticmethod
    def llama2_13b() -> 'LlamaShape':
        return LlamaShape(hidden=5120, inter=13824, layers=40, heads=40)


LINEAR_NAMES = ('self_attn.v_proj', 'self_attn.k_proj', 'self_attn.q_proj', 'self_attn.o_proj', 'mlp.up_proj',
                'mlp.gate_proj', 'mlp.down_proj')


def random_block_weights(shape: LlamaShape, layer: int, device: torch.device, dtype=torch.float16, seed: int = 0) -> dict:
    """Random-init weights of one decoder block (normal, std 0.02 = HF initializer_range; norms = 1), generated on the
    device from (seed, layer) so that every rank builds the same model without holding all 32 blocks."""
    g = torch.Generator(device=device).manual_seed(seed * 1000 + layer)
    h, i = shape.hidden, shape.inter

    def w(r, c):
        return (torch.randn(r, c, device=device, generator=g) * 0.02).to(dtype)
    return {'self_attn.q_proj': w(h, h), 'self_attn.k_proj': w(h, h), 'self_attn.v_proj': w(h, h),
            'self_attn.o_proj': w(h, h), 'mlp.gate_proj': w(i, h), 'mlp.up_proj': w(i, h), 'mlp.down_proj': w(h, i),
            'input_layernorm': torch.ones(h, device=device, dtype=dtype),
            'post_attention_layernorm': torch.ones(h, device=device, dtype=dtype)}


def hf_block_weights(block, device: torch.device) -> dict:
    """Weights of a transformers LlamaDecoderLayer in the dict form used here."""
    out = {n: block.get_submodule(n).weight.detach().to(device) for n in LINEAR_NAMES}
    out['input_layernorm'] = block.input_layernorm.weight.detach().to(device)
    out['post_attention_layernorm'] = block.post_attention_layernorm.weight.detach().to(device)
    return out


def _fused_ok(*ts: torch.Tensor) -> bool:
    """The one-pass kernels of csrc/llama_ops.cu take contiguous fp16 / bf16 CUDA tensors; everything else (the CPU tests of
    the driver logic, fp32) stays on the PyTorch expressions below, which the kernels reproduce rounding for rounding."""
    return all(t.is_cuda and t.dtype in (torch.float16, torch.bfloat16) and t.is_contiguous() for t in ts)


def rms_norm(x: torch.Tensor, weight: torch.Tensor, eps: float, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    if _fused_ok(x, weight) and x.shape[-1] <= 8192 and x.shape[-1] % 8 == 0 and (out is None or _fused_ok(out)):
        import gptq as _ops
        return _ops.llama_rms_norm(x, weight, eps, out=out)
    xf = x.float()
    y = (weight * (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).to(x.dtype))
    if out is None:
        return y
    out.copy_(y)
    return out


def rope_tables(seqlen: int, head_dim: int, theta: float, device, dtype):
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, device=device, dtype=torch.float32) / head_dim))
    ang = torch.outer(torch.arange(seqlen, device=device, dtype=torch.float32), inv)
    emb = torch.cat([ang, ang], dim=-1)
    return emb.cos().to(dtype).contiguous(), emb.sin().to(dtype).contiguous()


def apply_rope(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    """x (B, H, N, hd); HF Llama convention (rotate_half)."""
    hd = x.shape[-1]
    rot = torch.cat([-x[..., hd // 2:], x[..., :hd // 2]], dim=-1)
    return x * cos + rot * sin


def silu_mul(gate: torch.Tensor, up: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """F.silu(gate) * up, into `out` when given."""
    if _fused_ok(gate, up) and gate.numel() % 8 == 0 and (out is None or _fused_ok(out)):
        import gptq as _ops
        return _ops.llama_silu_mul(gate, up, out=out)
    y = F.silu(gate) * up
    if out is None:
        return y
    out.copy_(y)
    return out


def linear(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    """x (..., K) @ w (N, K)^T.  The replay of a quantised linear (llama.py:172-177) is K6's large-batch route: the weights are
    dequantised once (reconstruct_nn_linear) and multiplied on the library's own tcgen05 GEMM (gq_gemm_f16_nt), not cuBLAS."""
    if x.is_cuda and x.dtype in (torch.float16, torch.bfloat16) and w.dtype == x.dtype and x.shape[-1] % 8 == 0 and w.shape[0] % 2 == 0:
        import gptq as _ops
        y = _ops.gemm_f16_nt(x.reshape(-1, x.shape[-1]).contiguous(), w.contiguous())
        return y.view(*x.shape[:-1], w.shape[0])
    return x @ w.t()


def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, heads: int, cos, sin, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """q, k, v (B, N, heads * hd), not used by the caller afterwards (the rotary embedding is applied in place on CUDA);
    the result goes into `out` (B, N, heads * hd) when given: one transposing copy instead of two."""
    B, N, Hd = q.shape
    hd = Hd // heads
    if _fused_ok(q, k, cos, sin) and hd % 16 == 0 and cos.shape == (N, hd):
        import gptq as _ops
        _ops.llama_rope_(q, cos, sin, heads)
        _ops.llama_rope_(k, cos, sin, heads)
        q, k, v = (t.view(B, N, heads, hd).transpose(1, 2) for t in (q, k, v))
    else:
        q, k, v = (t.view(B, N, heads, hd).transpose(1, 2) for t in (q, k, v))
        q, k = apply_rope(q, cos, sin), apply_rope(k, cos, sin)
    o = F.scaled_dot_product_attention(q, k, v, is_causal=True)
    if out is None:
        return o.transpose(1, 2).reshape(B, N, Hd)
    out.view(B, N, heads, hd).copy_(o.transpose(1, 2))
    return out


@torch.no_grad()
def block_forward(shape: LlamaShape, weights: dict, x: torch.Tensor) -> torch.Tensor:
    """Dense forward of one decoder block with the given (original or reconstructed) weights; x (B, N, hidden)."""
    cos, sin = rope_tables(x.shape[1], shape.hidden // shape.heads, shape.rope_theta, x.device, x.dtype)
    h = rms_norm(x, weights['input_layernorm'], shape.rms_eps)
    a = attention(h @ weights['self_attn.q_proj'].t(), h @ weights['self_attn.k_proj'].t(),
                  h @ weights['self_attn.v_proj'].t(), shape.heads, cos, sin)
    x = x + a @ weights['self_attn.o_proj'].t()
    h = rms_norm(x, weights['post_attention_layernorm'], shape.rms_eps)
    m = F.silu(h @ weights['mlp.gate_proj'].t()) * (h @ weights['mlp.up_proj'].t())
    return x + m @ weights['mlp.down_proj'].t()


class PhaseTimer:
    """CUDA-event phase timing without host syncs inside the loop; summary() resolves the events."""

    def __init__(self, enabled: bool = True):
        self.enabled = enabled and torch.cuda.is_available()
        self._spans: list[tuple[str, torch.cuda.Event, torch.cuda.Event]] = []

    class _Span:
        def __init__(self, owner, name):
            self.owner, self.name = owner, name

        def __enter__(self):
            if self.owner.enabled:
                self.t0 = torch.cuda.Event(enable_timing=True)
                self.t0.record()
            return self

        def __exit__(self, *exc):
            if self.owner.enabled:
                t1 = torch.cuda.Event(enable_timing=True)
                t1.record()
                self.owner._spans.append((self.name, self.t0, t1))
            return False

    def __call__(self, name: str) -> 'PhaseTimer._Span':
        return PhaseTimer._Span(self, name)

    def summary(self) -> dict[str, float]:
        out: dict[str, float] = {}
        if not self.enabled:
            return out
        torch.cuda.synchronize()
        for name, t0, t1 in self._spans:
            out[name] = out.get(name, 0.0) + t0.elapsed_time(t1) * 1e-3
        return out


class MetaOffloader:
    """Moves finished quant_meta dicts to the host in the background (the reference does it inline, llama.py:173).

    A few lanes work side by side; in each, the device->host copy goes through two recycled pinned staging buffers on
    the lane's stream and the lane's thread copies staging -> pageable result tensors.  The main stream never waits for it; results() joins at the end."""

    def __init__(self, device: torch.device, staging_bytes: int = 64 << 20, workers: int = 3):
        self.device = device
        self.workers = workers
        # one single-thread executor per lane: a lane owns its stream and staging buffers, lanes run side by side
        self.lanes = [ThreadPoolExecutor(max_workers=1) for _ in range(workers)]
        self.streams = [torch.cuda.Stream(device=device) for _ in range(workers)]
        self.staging_bytes = staging_bytes
        self._staging: list[list[torch.Tensor]] = [[] for _ in range(workers)]
        self._futures: dict[str, Future] = {}
        self._next = 0

    def _stage(self, lane: int, nbytes: int, which: int) -> torch.Tensor:
        st = self._staging[lane]
        while len(st) < 2:
            st.append(torch.empty(self.staging_bytes, dtype=torch.uint8, pin_memory=True))
        if nbytes > st[which].numel():
            st[which] = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
        return st[which][:nbytes]

    def _work(self, lane: int, meta: dict, ready: torch.cuda.Event) -> dict:
        out = {}
        stream = self.streams[lane]
        with torch.cuda.device(self.device), torch.cuda.stream(stream):
            stream.wait_event(ready)
            which = 0
            pending = []                                            # (key, staging view, dtype, shape, event, source)
            for k, v in meta.items():
                if not torch.is_tensor(v) or v.device.type != 'cuda':
                    out[k] = v
                    continue
                src = v.contiguous()
                if len(pending) == 2:                               # both staging buffers busy: drain the older one
                    self._drain(pending.pop(0), out)
                st = self._stage(lane, src.numel() * src.element_size(), which)
                st.copy_(src.view(-1).view(torch.uint8), non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(stream)
                pending.append((k, st, src.dtype, tuple(src.shape), ev, src))
                which ^= 1
            for item in pending:
                self._drain(item, out)
        return {k: out[k] for k in meta}                            # original key order

    @staticmethod
    def _drain(item, out: dict) -> None:
        k, st, dtype, shape, ev, _keep = item
        ev.synchronize()
        out[k] = st.clone().view(dtype).view(shape)                 # pageable copy made by this thread

    def submit(self, name: str, meta: dict) -> None:
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream(self.device))
        lane = self._next
        self._next = (self._next + 1) % self.workers
        self._futures[name] = self.lanes[lane].submit(self._work, lane, meta, ready)

    def results(self) -> dict:
        out = {name: f.result() for name, f in self._futures.items()}
        self._futures.clear()
        return out

    def close(self) -> None:
        for lane in self.lanes:
            lane.shutdown(wait=True)


def _batches(n: int, bs: int):
    for b in range(0, n, bs):
        yield slice(b, min(n, b + bs))


class _GroupQuantizer:
    """One Hessian group: accumulate -> (all-reduce) -> invert -> quantise each linear -> dense replay weights."""

    def __init__(self, n_global: int, timer: PhaseTimer, pgroup, damp: float, act_order: bool, bits: int, gsize: int,
                 use_kernel_semantics: bool, save_device: torch.device, quant_options: Optional[dict] = None):
        self.hook = HessianHook()
        # the Quantizer options llama.py:152-162 pins (asymmetric, no MSE search, no EXL2 scales) unless overridden
        self.quant_options = {'scale_bit_width': None, 'quant_symmetric': False, 'quant_mse': False}
        self.quant_options.update(quant_options or {})
        self.n_global, self.timer, self.pgroup = n_global, timer, pgroup
        self.damp, self.act_order, self.bits, self.gsize = damp, act_order, bits, gsize
        self.use_kernel_semantics, self.save_device = use_kernel_semantics, save_device

    def add(self, x: torch.Tensor) -> None:
        with self.timer('hessian'):
            self.hook.add_batch(x)                                       # llama.py:121-125 -> model_utils.py:101-103

    def finish(self) -> None:
        with self.timer('allreduce'):
            sharded.allreduce_hessian(self.hook, self.n_global, self.pgroup)
        with self.timer('invert'):
            self.hook.invert(self.damp, self.act_order, defer_check=True)   # llama.py:138; checked behind the queued solves (settle)

    def quantize(self, weight: torch.Tensor) -> tuple[dict, dict, torch.Tensor]:
        """-> (quant_meta on save_device, metrics, dense reconstructed weight on the GPU)."""
        return self.quantize_many([weight])[0]

    def quantize_many(self, weights: list) -> list[tuple[dict, dict, torch.Tensor]]:
        """All linears that share this group's Hessian (v/k/q, up/gate, or one): spread over the ranks by
        sharded.plan_group (whole linears first, rows inside a linear's own ranks), one all-gather for the group."""
        dev = weights[0].device
        G = weights[0].shape[1] // self.gsize
        gs = torch.full([G], self.gsize, dtype=torch.int32, device=dev)          # llama.py:146-147
        gb = torch.full([G], self.bits, dtype=torch.int32, device=dev)
        with self.timer('solve'):
            if self.quant_options.get('scale_bit_width') is not None:
                # EXL2 scales are a maximum over all rows (quant.py:110): row sharding refuses them, one rank's worth of
                # work on every rank instead
                from gptq_py import gptq_quant
                self.hook.settle()
                results = [gptq_quant(weight=w, hessian_hook=self.hook, group_sizes=gs, group_bit_widths=gb,
                                      gptq_use_kernel=self.use_kernel_semantics, gptq_block_sizes=gs, save_device=dev,
                                      **self.quant_options) for w in weights]
            else:
                results = sharded.gptq_quant_group_sharded(
                    weights, self.hook, group=self.pgroup, group_sizes=gs, group_bit_widths=gb,
                    gptq_use_kernel=self.use_kernel_semantics, gptq_block_sizes=gs, save_device=dev,
                    **self.quant_options)                                         # llama.py:148-163
        out = []
        for weight, res in zip(weights, results):
            with self.timer('reconstruct'):
                what = reconstruct_nn_linear(res['quant_meta'], dtype=weight.dtype, device=dev, group_size=self.gsize).weight.data   # llama.py:172
            meta = res['quant_meta']
            if self.save_device != dev:
                with self.timer('offload'):
                    meta = {k: (v.to(self.save_device) if torch.is_tensor(v) else v) for k, v in meta.items()}
            out.append((meta, res['metrics'], what))
        return out


@torch.no_grad()
def quantize_block(shape: LlamaShape, weights: dict, inps: torch.Tensor, n_global: int, batch_size: int = 16,
                   timer: Optional[PhaseTimer] = None, pgroup=None, damp: float = 1e-2, act_order: bool = True,
                   bits: int = 4, gsize: int = 128, use_kernel_semantics: bool = True,
                   save_device: torch.device = torch.device('cpu'),
                   quant_options: Optional[dict] = None) -> tuple[torch.Tensor, dict, dict]:
    """GPTQ of one decoder block.  inps (n_local, N, hidden) on the GPU -> (next inps, {name: quant_meta},
    {name: metrics}).  Group order and the order inside a group are the reference's (v, k, q / o / up, gate / down:
    the hoisted dependency list of model_utils.py:146-248 on a Llama block)."""
    timer = timer or PhaseTimer(False)
    n, N, _ = inps.shape
    dev, dt = inps.device, inps.dtype
    data: dict = {}
    metrics: dict = {}
    cos, sin = rope_tables(N, shape.hidden // shape.heads, shape.rope_theta, dev, dt)

    def make_group():
        return _GroupQuantizer(n_global, timer, pgroup, damp, act_order, bits, gsize, use_kernel_semantics, save_device,
                               quant_options)

    def quantize_into(grp: _GroupQuantizer, *names: str):
        whats = []
        for name, (meta, met, what) in zip(names, grp.quantize_many([weights[n] for n in names])):
            data[name], metrics[name] = meta, met
            whats.append(what)
        return whats[0] if len(whats) == 1 else whats

    # ---- group 1: v, k, q on input_layernorm(x)
    grp = make_group()
    h1 = torch.empty_like(inps)
    with timer('forward'):
        for b in _batches(n, batch_size):
            rms_norm(inps[b], weights['input_layernorm'], shape.rms_eps, out=h1[b])
    for b in _batches(n, batch_size):
        grp.add(h1[b])
    grp.finish()
    wv, wk, wq = quantize_into(grp, 'self_attn.v_proj', 'self_attn.k_proj', 'self_attn.q_proj')
    del grp
    # ---- group 2: o on the attention output (computed from the QUANTISED q/k/v, llama.py:176-177)
    grp = make_group()
    att = torch.empty_like(inps)
    with timer('forward'):
        for b in _batches(n, batch_size):
            x = h1[b]
            attention(linear(x, wq), linear(x, wk), linear(x, wv), shape.heads, cos, sin, out=att[b])
    del h1, wq, wk, wv
    for b in _batches(n, batch_size):
        grp.add(att[b])
    grp.finish()
    wo = quantize_into(grp, 'self_attn.o_proj')
    del grp
    # ---- group 3: up, gate on post_attention_layernorm(x + o)
    grp = make_group()
    h2 = torch.empty_like(inps)
    h3 = torch.empty_like(inps)
    with timer('forward'):
        for b in _batches(n, batch_size):
            h2[b] = inps[b] + linear(att[b], wo)
            rms_norm(h2[b], weights['post_attention_layernorm'], shape.rms_eps, out=h3[b])
    del att, wo
    for b in _batches(n, batch_size):
        grp.add(h3[b])
    grp.finish()
    wu, wg = quantize_into(grp, 'mlp.up_proj', 'mlp.gate_proj')
    del grp
    # ---- group 4: down on silu(gate) * up
    grp = make_group()
    act = torch.empty(n, N, shape.inter, dtype=dt, device=dev)
    with timer('forward'):
        for b in _batches(n, batch_size):
            x = h3[b]
            silu_mul(linear(x, wg), linear(x, wu), out=act[b])
    del h3, wu, wg
    for b in _batches(n, batch_size):
        grp.add(act[b])
    grp.finish()
    wd = quantize_into(grp, 'mlp.down_proj')
    del grp
    # ---- inputs of the next block (llama.py:181-189)
    with timer('forward'):
        for b in _batches(n, batch_size):
            h2[b] += linear(act[b], wd)
    # plain {name: float} dicts like the reference's; the deferred values are read here, once per block, behind the queued work
    metrics = {name: dict(met) for name, met in metrics.items()}
    return h2, data, metrics


@torch.no_grad()
def quantize_llama(shape: LlamaShape, block_weights: Callable[[int], dict], inps: torch.Tensor,
                   n_global: Optional[int] = None, batch_size: int = 16, n_blocks: Optional[int] = None, pgroup=None,
                   timing: bool = True, log: Optional[Callable[[str], None]] = None, **block_kw) -> dict:
    """The reference's quantize_llama (llama.py:65-204) on this build's hot path.

    block_weights(i) -> the weights of decoder block i on the GPU (random_block_weights / hf_block_weights);
    inps: this rank's embedded calibration sequences (n_local, N, hidden), fp16/bf16 on the GPU;
    n_global: total number of calibration sequences over all ranks (default: n_local x world).
    Returns {'data': {canonical name: quant_meta}, 'metrics': {canonical name: metrics}, 'timing': {phase: s},
    'seconds': wall seconds, 'block_seconds': [...]} -- 'data' / 'metrics' as the reference's results dict."""
    rank, world = sharded.world_info(pgroup)
    if n_global is None:
        n_global = inps.shape[0] * world
    timer = PhaseTimer(timing)
    results: dict = {'data': {}, 'metrics': {}, 'block_seconds': []}
    n_blocks = shape.layers if n_blocks is None else n_blocks
    dev = inps.device
    save_device = block_kw.pop('save_device', torch.device('cpu'))
    to_host = torch.device(save_device).type == 'cpu'
    # the gathered quant_meta is identical on every rank: rank 0 keeps the host copy, the others drop it
    offloader = MetaOffloader(dev) if to_host and rank == 0 else None
    torch.cuda.synchronize(dev)
    t_all = time.perf_counter()
    for i in range(n_blocks):
        t0 = time.perf_counter()
        with timer('weights'):
            w = block_weights(i)
        inps, data, metrics = quantize_block(shape, w, inps, n_global, batch_size, timer, pgroup, save_device=dev, **block_kw)
        for name in LINEAR_NAMES:
            full = f'model.layers.{i}.{name}'
            results['data'][full] = data[name] if (rank == 0 or not to_host) else None
            results['metrics'][full] = metrics[name]
            if offloader is not None:
                offloader.submit(full, data[name])            # host copy in the background (llama.py:173 does it inline)
        del w, data
        torch.cuda.synchronize(dev)
        results['block_seconds'].append(time.perf_counter() - t0)
        if log is not None and rank == 0:
            log(f'finished block {i} in {results["block_seconds"][-1]:.2f} s')
    if offloader is not None:
        t0 = time.perf_counter()
        results['data'].update(offloader.results())
        offloader.close()
        results['offload_join_seconds'] = time.perf_counter() - t0
    results['seconds'] = time.perf_counter() - t_all
    results['timing'] = timer.summary()
    results['final_hidden'] = inps
    return results


def synthetic_inputs(shape: LlamaShape, n_seq: int, seqlen: int, device: torch.device, dtype=torch.float16,
                     seed: int = 1234) -> torch.Tensor:
    """Embedded synthetic calibration batch: random token ids through a random-init embedding table (std 0.02).
    There is no network for wikitext2/c4 (data_utils.py) or checkpoints, so calibration data is synthetic."""
    g = torch.Generator(device=device).manual_seed(seed)
    emb = (torch.randn(shape.vocab, shape.hidden, device=device, generator=g) * 0.02).to(dtype)
    tok = torch.randint(0, shape.vocab, (n_seq, seqlen), device=device, generator=g)
    return emb[tok]


def _rope_theta(cfg) -> float:
    theta = getattr(cfg, 'rope_theta', None)
    if theta is None:                                   # transformers >= 5 keeps it under rope_parameters
        theta = (getattr(cfg, 'rope_parameters', None) or {}).get('rope_theta', 10000.0)
    return float(theta)


def quantize_hf_llama(model, encodings: torch.Tensor, device: torch.device, batch_size: int = 16, **kw) -> dict:
    """Same entry point shape as the reference's quantize_llama(model, encodings, device, batch_size): a transformers
    LlamaForCausalLM and token ids (n_seq, N)."""
    cfg = model.config
    shape = LlamaShape(hidden=cfg.hidden_size, inter=cfg.intermediate_size, layers=cfg.num_hidden_layers,
                       heads=cfg.num_attention_heads, vocab=cfg.vocab_size, rms_eps=cfg.rms_norm_eps,
                       rope_theta=_rope_theta(cfg))
    if getattr(cfg, 'num_key_value_heads', shape.heads) != shape.heads:
        raise NotImplementedError('grouped-query attention is not handled by this driver')
    emb = model.model.embed_tokens.weight.detach().to(device)
    inps = emb[encodings.to(device)]
    layers = model.model.layers
    return quantize_llama(shape, lambda i: hf_block_weights(layers[i], device), inps, batch_size=batch_size, **kw)


# ---------------------------------------------------------------------------------------------------------------------
# Evaluation on the quantised model  (the reference's evaluate_llama, llama.py:207-247)

def _linear(x: torch.Tensor, w) -> torch.Tensor:
    """w: dense (R, C) weight or a module (W4A16Linear / nn.Linear)."""
    return x @ w.t() if torch.is_tensor(w) else w(x)


@torch.no_grad()
def block_forward_any(shape: LlamaShape, weights: dict, x: torch.Tensor) -> torch.Tensor:
    """block_forward with each linear either a dense weight or a module (e.g. quant.W4A16Linear on packed codes)."""
    cos, sin = rope_tables(x.shape[1], shape.hidden // shape.heads, shape.rope_theta, x.device, x.dtype)
    h = rms_norm(x, weights['input_layernorm'], shape.rms_eps)
    a = attention(_linear(h, weights['self_attn.q_proj']), _linear(h, weights['self_attn.k_proj']),
                  _linear(h, weights['self_attn.v_proj']), shape.heads, cos, sin)
    x = x + _linear(a, weights['self_attn.o_proj'])
    h = rms_norm(x, weights['post_attention_layernorm'], shape.rms_eps)
    m = silu_mul(_linear(h, weights['mlp.gate_proj']), _linear(h, weights['mlp.up_proj']))
    return x + _linear(m, weights['mlp.down_proj'])


@torch.no_grad()
def evaluate_llama(shape: LlamaShape, block_weights: Callable[[int], dict], embed: torch.Tensor, final_norm: torch.Tensor,
                   lm_head: torch.Tensor, encodings: torch.Tensor, batch_size: int = 1,
                   n_blocks: Optional[int] = None) -> torch.Tensor:
    """Layer-streamed perplexity as the reference computes it (llama.py:207-247): all sequences go through block i
    before block i+1 is touched, then final norm, lm_head, and the mean over sequences of the per-batch mean token NLL,
    exponentiated.  block_weights(i) may mix dense weights and W4A16Linear modules (checkpoint.quantized_linears), so
    the quantised model is evaluated straight from its packed codes through K6."""
    dev = embed.device
    inps = embed[encodings.to(dev)]
    n_blocks = shape.layers if n_blocks is None else n_blocks
    for i in range(n_blocks):
        w = block_weights(i)
        for b in _batches(len(inps), batch_size):
            inps[b] = block_forward_any(shape, w, inps[b])
        del w
    nlls = []
    for b in _batches(len(inps), batch_size):
        h = rms_norm(inps[b], final_norm, shape.rms_eps)
        logits = h @ lm_head.t()
        shift_logits = logits[:, :-1, :]
        shift_labels = encodings[b, 1:].to(dev)
        nll = F.cross_entropy(shift_logits.flatten(end_dim=-2).float(), shift_labels.flatten())
        nlls.extend([nll] * logits.shape[0])
    return torch.stack(nlls).to(torch.float32).mean().exp()


def packed_block_weights(packed: dict, norms: Callable[[int], dict], device: torch.device, dtype: torch.dtype) -> Callable[[int], dict]:
    """block_weights(i) for evaluate_llama from a packed model (checkpoint.pack_results / load_packed): the seven
    linears of block i as W4A16Linear modules, the two norm weights from norms(i)."""
    from quant import W4A16Linear

    def get(i: int) -> dict:
        out = dict(norms(i))
        for name in LINEAR_NAMES:
            layer = packed['layers'][f'model.layers.{i}.{name}']
            out[name] = W4A16Linear({
                'qweight_packed': layer['qweight_packed'].to(device),
                'scale_t': layer['scale_t'].to(device=device, dtype=dtype), 'qzero_t': layer['qzero_t'].to(device),
                'perm': layer['perm'].to(device) if layer.get('perm') is not None else None,
                'group_size': int(layer['group_size'])})
        return out
    return get
