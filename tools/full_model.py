"""Full-model GPTQ wall time (BASELINE config 2 / north_star: Llama-2-7B, 4-bit g128, 128 x 2048 calibration tokens).

  python tools/full_model.py [--blocks 32] [--n-seq 128] [--seqlen 2048] [--model 7b|13b]
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/full_model.py ...

Random-init weights and synthetic token ids (no network for checkpoints / wikitext2).  Prints one JSON line on rank 0:
wall seconds for the whole model, seconds per block, and the CUDA-event phase breakdown of rank 0.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, 'gptq-v2_b200')):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import llama_driver as L  # noqa: E402
import sharded  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--blocks', type=int, default=None)
    ap.add_argument('--n-seq', type=int, default=128)
    ap.add_argument('--seqlen', type=int, default=2048)
    ap.add_argument('--batch', type=int, default=16)
    ap.add_argument('--model', default='7b')
    ap.add_argument('--bits', type=int, default=4)
    ap.add_argument('--pure', action='store_true', help='fp32 PyTorch-path semantics (gptq_use_kernel=False)')
    ap.add_argument('--eval-seqs', type=int, default=0, help='after quantising: pack the result (checkpoint.py) and run the layer-streamed perplexity of N sequences through K6 and through dense reconstructed weights (rank 0, N=1 only)')
    ap.add_argument('--mse', action='store_true', help='Quantizer.mse grid search (quant_mse=True)')
    ap.add_argument('--sym', action='store_true', help='symmetric grid (quant_symmetric=True)')
    ap.add_argument('--keep', default='cpu', help='where quant_meta ends up: cpu (as the reference) or cuda')
    args = ap.parse_args()

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    shape = L.LlamaShape.llama2_13b() if args.model == '13b' else L.LlamaShape.llama2_7b()
    lo, hi = sharded.shard_range(args.n_seq, world, rank)
    # every rank draws the same global batch and keeps its own sequences
    inps = L.synthetic_inputs(shape, args.n_seq, args.seqlen, dev)[lo:hi].contiguous()
    # warm-up: one block at a tiny size loads the library, builds tensor maps, primes cuBLAS / SDPA
    warm = L.LlamaShape(hidden=256, inter=512, layers=1, heads=4, vocab=64)
    L.quantize_llama(warm, lambda i: L.random_block_weights(warm, i, dev), L.synthetic_inputs(warm, 2 * world, 128, dev)[:2],
                     n_global=2 * world, batch_size=2, timing=False)
    if world > 1:
        dist.barrier()
    res = L.quantize_llama(shape, lambda i: L.random_block_weights(shape, i, dev), inps, n_global=args.n_seq,
                           batch_size=args.batch, n_blocks=args.blocks, use_kernel_semantics=not args.pure, bits=args.bits,
                           save_device=torch.device(args.keep if args.keep == 'cpu' else dev),
                           quant_options={'quant_mse': args.mse, 'quant_symmetric': args.sym},
                           log=(lambda s: print(s, file=sys.stderr, flush=True)))
    secs = torch.tensor([res['seconds']], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(secs, op=dist.ReduceOp.MAX)
    if rank == 0:
        nb = len(res['block_seconds'])
        errs = [m['l2.0_error'] / max(m['l2.0_norm'], 1e-30) for m in res['metrics'].values()]
        print(json.dumps({
            'what': 'full-model GPTQ', 'model': f'llama-2-{args.model} (random init)', 'blocks': nb, 'n_gpus': world,
            'n_seq': args.n_seq, 'seqlen': args.seqlen, 'seconds': round(float(secs.item()), 3),
            'seconds_per_block': round(float(secs.item()) / nb, 4),
            'extrapolated_full_model_s': round(float(secs.item()) / nb * shape.layers, 2),
            'phases_rank0_s': {k: round(v, 3) for k, v in sorted(res['timing'].items())},
            'mean_rel_l2_error': sum(errs) / len(errs),
            'bits': args.bits, 'quant_mse': args.mse, 'quant_symmetric': args.sym, 'semantics': 'fp32-pure' if args.pure else 'kernel-compat', 'quant_meta_on': args.keep}), flush=True)
    if args.eval_seqs > 0 and world == 1:
        import time
        import checkpoint
        from quant import reconstruct_nn_linear
        nb = len(res['block_seconds'])
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        packed = checkpoint.pack_results(res, dev)
        torch.cuda.synchronize()
        t_pack = time.perf_counter() - t0
        g = torch.Generator(device=dev).manual_seed(77)
        emb = (torch.randn(shape.vocab, shape.hidden, device=dev, generator=g) * 0.02).half()
        head = (torch.randn(shape.vocab, shape.hidden, device=dev, generator=g) * 0.02).half()
        fnorm = torch.ones(shape.hidden, device=dev, dtype=torch.float16)
        tok = torch.randint(0, shape.vocab, (args.eval_seqs, args.seqlen), device=dev, generator=g).cpu()
        ones = {'input_layernorm': fnorm, 'post_attention_layernorm': fnorm}

        def dense_block(i):
            out = dict(ones)
            for n in L.LINEAR_NAMES:
                out[n] = reconstruct_nn_linear(res['data'][f'model.layers.{i}.{n}'], torch.float16, dev).weight.data
            return out
        out = {}
        from quant import W4A16Linear
        pbw = L.packed_block_weights(packed, lambda i: ones, dev, torch.float16)
        # 'packed': W4A16Linear as shipped (2048-row batches take the dequantise-once + dense GEMM route);
        # 'packed_k6_only': the fused K6 kernel for every batch size; 'dense': the reference's reconstructed fp16 weights
        for label, bw in (('packed', pbw), ('packed_k6_only', pbw), ('dense', dense_block)):
            W4A16Linear.dense_route = label != 'packed_k6_only'
            L.evaluate_llama(shape, bw, emb, fnorm, head, tok[:1], batch_size=1, n_blocks=1)      # warm-up
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ppl = float(L.evaluate_llama(shape, bw, emb, fnorm, head, tok, batch_size=1, n_blocks=nb))
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            out[label] = {'ppl': ppl, 'seconds': round(dt, 3), 'tokens_per_s': round(args.eval_seqs * args.seqlen / dt, 1)}
        W4A16Linear.dense_route = True
        raw = sum(v.numel() * v.element_size() for m in res['data'].values() for v in m.values() if torch.is_tensor(v))
        print(json.dumps({'what': 'packed checkpoint + layer-streamed eval', 'blocks': nb, 'eval_seqs': args.eval_seqs,
                          'pack_seconds': round(t_pack, 3), 'packed_bytes': checkpoint.packed_bytes(packed),
                          'reference_format_bytes': raw, 'eval': out}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
