#!/usr/bin/env python
"""Train a stack of Transformer FFN sub-blocks on synthetic data under 1 GPU / DDP / FSDP / MP.

Same command line as the reference (``train_ffns.py:343-351``):

    python train_ffns.py --num_steps 16 --batch_size 8 --seq_len 1024 --layers 1 --model_size 8192 --method M

``M``: 0 = all methods, 1 = one GPU, 2 = DDP, 3 = FSDP, 4 = MP/TP (``:5-6``).  Extra flags select what the
reference hard-codes: ``--ffn_dim --dtype --activation --residual --bias --optimizer --lr --world_size --backend
--device --impl --comm --recompute --mp_mode --fsdp_gather --mp_input --save_path --load_path --profile --cuda_graph --data``.
"""
from __future__ import annotations

import argparse
import time

import torch

from distributed_llm_code_samples_b200.config import TrainConfig
from distributed_llm_code_samples_b200.models.ffn import init_stack, num_params
from distributed_llm_code_samples_b200.parallel import api
from distributed_llm_code_samples_b200.utils.checkpoint import load_checkpoint, save_checkpoint
from distributed_llm_code_samples_b200.utils.data import make_seeds


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    p.add_argument('-s', '--num_steps', type=int, default=1)
    p.add_argument('-bs', '--batch_size', type=int, default=8)
    p.add_argument('-n', '--seq_len', type=int, default=1024)
    p.add_argument('-l', '--layers', type=int, default=1)
    p.add_argument('-d', '--model_size', type=int, default=4)
    p.add_argument('-m', '--method', type=int, default=0)
    p.add_argument('-r', '--random_seed', type=int, default=0)   # 0 = unseeded (train_ffns.py:355-358)
    # extensions
    p.add_argument('--ffn_dim', type=int, default=None)
    p.add_argument('--dtype', default='fp32', choices=['fp32', 'bf16'])
    p.add_argument('--activation', default='relu', choices=['relu', 'gelu'])
    p.add_argument('--residual', action='store_true')
    p.add_argument('--bias', action='store_true')
    p.add_argument('--optimizer', default='sgd', choices=['sgd', 'adam'])
    p.add_argument('--lr', type=float, default=None)
    p.add_argument('--grad_scale', type=float, default=1.0)
    p.add_argument('--world_size', type=int, default=0)
    p.add_argument('--backend', default='auto', choices=['auto', 'nccl', 'gloo'])
    p.add_argument('--device', default='auto', choices=['auto', 'cuda', 'cpu'])
    p.add_argument('--impl', default='auto', choices=['auto', 'native', 'torch'])
    p.add_argument('--comm', default='auto', choices=['auto', 'fused', 'native', 'nccl'])
    p.add_argument('--recompute', action='store_true')
    p.add_argument('--mp_mode', default='ar', choices=['ar', 'sp', 'spf'])
    p.add_argument('--fsdp_gather', default='fused', choices=['dma', 'fused'], help='fused FSDP: copy-engine prefetch, or all-gather inside the consuming GEMMs')
    p.add_argument('--mp_input', default='sharded', choices=['full', 'sharded'])
    p.add_argument('--stochastic_rounding', action='store_true', help='bf16: round optimizer updates stochastically (no fp32 master weights)')
    p.add_argument('--nvls', default='auto', choices=['auto', 'off'], help='fused MP all-reduce: sum inside the NVSwitch (multimem) where available')
    p.add_argument('--master_port', type=int, default=29500)
    p.add_argument('--save_path', default=None)
    p.add_argument('--load_path', default=None)
    p.add_argument('--save_every', type=int, default=0, help='also checkpoint every N optimizer steps (needs --save_path)')
    p.add_argument('--profile', action='store_true')
    p.add_argument('--cuda_graph', action='store_true', help='replay each step from a CUDA graph (1 GPU, SGD)')
    p.add_argument('--data', default='cpu', choices=['cpu', 'device'])
    return p


def main(argv=None) -> int:
    args = build_parser().parse_args(argv)
    cfg_kwargs = {k: v for k, v in vars(args).items() if v is not None}
    cfg = TrainConfig(**cfg_kwargs)
    if cfg.dtype == 'bf16' and not cfg.stochastic_rounding and cfg.lr * cfg.grad_scale < 1e-3:
        print(f'note: bf16 weights are updated in place without an fp32 master copy; at lr={cfg.lr} most updates are below half a '
              'bf16 ulp of the weight and round away. Pass --stochastic_rounding (or use --dtype fp32, the reference\'s dtype).')
    ffn = cfg.ffn_dim

    print(f'ARGS:\n num_steps: {args.num_steps}\n BS: {args.batch_size}\n N: {args.seq_len}\n D: {args.model_size}\n FFN: {ffn}\n')

    gen = None
    if args.random_seed != 0:
        gen = torch.Generator()
        gen.manual_seed(args.random_seed)

    seeds = make_seeds(args.num_steps, gen)
    layers_params = init_stack(args.layers, args.model_size, ffn, gen, bias=args.bias)
    start_step = 0
    opt_state = None
    if args.load_path:
        ck = load_checkpoint(args.load_path)
        layers_params = ck["layers_params"]
        start_step = ck.get("step", 0)
        opt_state = ck.get("optimizer") or None       # Adam moments + step count, full layout: resumes under any strategy
        if args.optimizer == "adam" and not (opt_state and opt_state.get("kind") == "adam"):
            print("warning: resuming an Adam run from a checkpoint without Adam state - the moments restart at zero")
        if ck.get("seeds") is not None and len(ck["seeds"]) == len(seeds):
            seeds = ck["seeds"]
        print(f'resumed from {args.load_path} at step {start_step}')

    n_params = num_params(layers_params)
    print(f'PARAMS: {n_params:_} (size {4 * n_params / (1024 * 1024 * 1024)} GB)')
    print('\n')
    print('initial layers_params[0]', layers_params[0][0].shape, layers_params[0][1].shape)
    print('initial layers_params[0]', layers_params[0][0][:5, :5], layers_params[0][1][:5, :5])

    fns = [api.train_1gpu, api.train_ddp, api.train_fsdp, api.train_tp]
    fns_layers_params = []
    run_seeds = seeds[start_step:] if start_step else seeds
    for i, fn in enumerate(fns):
        if args.method == 0 or args.method == i + 1:
            t0 = time.time()
            out = fn(layers_params, run_seeds, args.batch_size * args.seq_len, args.model_size, cfg, opt_state)
            t1 = time.time()
            fns_layers_params.append(out)
            st = api.last_stats()
            print(f'\n{fn.__name__} takes {t1 - t0} seconds')
            print(f'{fn.__name__} device time {st.get("ms", 0.0):.3f} ms, {st.get("tokens_per_s", 0.0):.1f} tokens/s, '
                  f'{st.get("model_tflops", 0.0):.2f} model TFLOP/s, peak {st.get("max_mem_gib", 0.0):.2f} GiB '
                  f'on {st.get("world", 1)} rank(s) (device-timed, max over ranks)')
            print(f'final {fn.__name__} layers_params[0]', out[0][0].shape, out[0][1].shape)
            print(f'final {fn.__name__} layers_params[0]', out[0][0][:5, :5], out[0][1][:5, :5])
            if args.save_path:
                path = args.save_path if args.method != 0 else f'{args.save_path}.{fn.__name__}'
                save_checkpoint(path, out, step=len(seeds), args=cfg.to_dict(), seeds=seeds, optimizer=api.last_optimizer_state())
                print(f'saved {path}')

    if args.method == 0:   # compare DDP against FSDP (train_ffns.py:386-391)
        worst_abs, worst_ulp = 0.0, 0.0
        mant = 8 if cfg.dtype == 'bf16' else 24              # mantissa bits of the storage dtype
        for l in range(args.layers):
            for k in range(2):
                a, b = fns_layers_params[1][l][k], fns_layers_params[2][l][k]
                if not torch.allclose(a, b):
                    print(f"SoftAssertionError: L {l} ddp[{k}] {a} fsdp[{k}] {b}")
                d = (a.double() - b.double()).abs()
                worst_abs = max(worst_abs, d.max().item())
                spacing = torch.exp2(torch.floor(torch.log2(a.double().abs().clamp_min(1e-30))) - (mant - 1))
                worst_ulp = max(worst_ulp, (d / spacing).max().item())
        print(f'DDP vs FSDP: max |difference| {worst_abs:.3e} = {worst_ulp:.2f} ulp of the {cfg.dtype} weights')
    return 0


if __name__ == '__main__':
    raise SystemExit(main())
