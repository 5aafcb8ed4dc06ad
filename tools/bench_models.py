"""North-star model configurations under data parallelism: ResNet-50 (bf16, 224x224 synthetic) and
GPT-2 small (bf16, 1024-token synthetic), comparing this framework's DDP (flat symmetric-memory
arenas + NVLink all-reduce kernels overlapped with backward) with stock torch DDP + NCCL on the same
GPUs. Launch: python -m torch.distributed.run --nproc-per-node N tools/bench_models.py --model gpt2 --ddp ours
Prints one JSON line from rank 0; device-timed with CUDA events, max over ranks."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from ray_torch_distributed_checkpoint_b200.models import GPT2, GPT2Config, resnet50  # noqa: E402
from ray_torch_distributed_checkpoint_b200.parallel.ddp import DistributedDataParallel  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", choices=["resnet50", "gpt2"], required=True)
    ap.add_argument("--ddp", choices=["ours", "torch"], default="ours")
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--algo", default="auto")
    ap.add_argument("--ckpt-every", type=int, default=0, help="async checkpoint every N steps (ours) / torch.save (torch)")
    ap.add_argument("--linear", choices=["ours", "torch"], default=None, help="gpt2: nn.Linear on the repo's bf16 tcgen05 GEMM (default with --ddp ours) or cuBLAS")
    ap.add_argument("--opt", choices=["ours", "torch"], default=None, help="gpt2: the repo's fused bf16 AdamW kernel over the flat arena (default with --ddp ours) or torch's fused AdamW")
    ap.add_argument("--head", choices=["ours", "torch"], default=None, help="gpt2: LM head GEMM + cross-entropy + backward on the repo's kernels (ops.lm_head_cross_entropy; default with --linear ours) or F.linear + F.cross_entropy")
    ap.add_argument("--graph", action="store_true", help="capture the whole step (zero_grad, forward, backward, optimizer) in one CUDA graph (utils.GraphedStep); one GPU, no checkpoints")
    ap.add_argument("--norm", choices=["ours", "torch"], default=None, help="gpt2: LayerNorm forward / backward on the repo's kernels (ops.use_fused_layernorm; default with --linear ours) or torch's")
    ap.add_argument("--vocab", type=int, default=50257, help="gpt2: vocabulary size (50304 = padded to a multiple of 64, the usual work-around for cuBLAS's unaligned-N fallback)")
    args = ap.parse_args()
    args.linear = args.linear or ("ours" if args.ddp == "ours" else "torch")
    args.opt = args.opt or ("ours" if args.ddp == "ours" else "torch")
    args.head = args.head or args.linear
    args.norm = args.norm or args.linear
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(0)
    torch.backends.cudnn.benchmark = True
    if args.model == "resnet50":
        B = args.batch or 128
        model = resnet50().to(dev, dtype=torch.bfloat16, memory_format=torch.channels_last)
        x = torch.randn(B, 3, 224, 224, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
        y = torch.randint(0, 1000, (B,), device=dev)
        loss_of = lambda m: torch.nn.functional.cross_entropy(m(x).float(), y)
        opt_cls = lambda m: torch.optim.SGD(m.parameters(), lr=0.1, momentum=0.9)
        per_sample = 1
    else:
        B = args.batch or 16
        model = GPT2(GPT2Config(vocab_size=args.vocab)).to(dev, dtype=torch.bfloat16)
        if args.linear == "ours":
            from ray_torch_distributed_checkpoint_b200.ops import use_bf16_linears
            use_bf16_linears(model)                      # 48 Linear layers: forward, dgrad, wgrad on csrc/gemm_bf16.cu
        model.fused_head = args.head == "ours"           # tied head + loss: gemm_bf16 + csrc/lm_head.cu, no fp32 logits
        if args.norm == "ours":
            from ray_torch_distributed_checkpoint_b200.ops import use_fused_layernorm
            use_fused_layernorm(model)                   # 25 LayerNorms: csrc/layernorm.cu
        idx = torch.randint(0, args.vocab, (B, 1024), device=dev)
        tgt = torch.randint(0, args.vocab, (B, 1024), device=dev)
        loss_of = lambda m: m(idx, tgt)
        if args.opt == "ours":
            from ray_torch_distributed_checkpoint_b200.optim import AdamW
            opt_cls = lambda m: AdamW(m, lr=3e-4)        # adam_bf16_kernel: fp32 master + moments, one launch over the arena
        else:
            opt_cls = lambda m: torch.optim.AdamW(m.parameters(), lr=3e-4, fused=True, capturable=args.graph)
        per_sample = 1024
    if world > 1:
        if args.ddp == "ours":
            model = DistributedDataParallel(model, device=dev, algo=args.algo)
        else:
            model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local], gradient_as_bucket_view=True)
    opt = opt_cls(model)
    ckpt = None
    if args.ckpt_every and rank == 0 and args.ddp == "ours":
        from ray_torch_distributed_checkpoint_b200.ckpt import AsyncCheckpointer
        ckpt = AsyncCheckpointer(dev)
    ckpt_dir = os.path.join("/tmp", f"rtdc_bench_ckpt_{os.getpid()}")

    def step(i):
        if isinstance(model, DistributedDataParallel):
            model.zero_grad()
        else:
            opt.zero_grad(set_to_none=False)
        loss = loss_of(model)
        loss.backward()
        opt.step()
        if args.ckpt_every and rank == 0 and (i + 1) % args.ckpt_every == 0:
            state = {"step": i, "model_state_dict": model.state_dict(), "optimizer_state_dict": opt.state_dict()}
            if ckpt is not None:
                ckpt.save(state, ckpt_dir, "latest_model.pt")
            else:
                os.makedirs(ckpt_dir, exist_ok=True)
                torch.save(state, os.path.join(ckpt_dir, "latest_model.pt"))
        return loss

    if args.graph:
        if world > 1 or args.ckpt_every:
            raise SystemExit("--graph is for one GPU without checkpoints")
        from ray_torch_distributed_checkpoint_b200.utils import GraphedStep
        eager_step = step
        graphed = GraphedStep(lambda: eager_step(0), warmup=3)
        step = lambda i: graphed()
    for i in range(args.warmup):
        step(i)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        loss = step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    # time to durable: until the last checkpoint's files are published (the asynchronous writer is still busy when the last
    # step's kernels finish; torch.save in the stock arm is synchronous, so for it both numbers coincide)
    import time
    t_wall = time.perf_counter()
    if ckpt is not None:
        ckpt.wait()
    durable_extra_ms = (time.perf_counter() - t_wall) * 1e3
    if world > 1:
        t = torch.tensor([ms, durable_extra_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, durable_extra_ms = float(t[0].item()), float(t[1].item())
    if rank == 0:
        print(json.dumps({"model": args.model, "ddp": args.ddp, "n_gpus": world, "batch_per_gpu": B, "steps": args.steps,
                          "ms_per_step": round(ms / args.steps, 3), "samples_per_s": round(B * world * args.steps / (ms / 1e3), 1),
                          "tokens_per_s": round(B * world * per_sample * args.steps / (ms / 1e3), 1) if per_sample > 1 else None,
                          "ms_per_step_until_durable": round((ms + durable_extra_ms) / args.steps, 3),
                          "linear": args.linear if args.model == "gpt2" else "torch", "head": args.head if args.model == "gpt2" else "torch", "norm": args.norm if args.model == "gpt2" else "torch",
                          "vocab": args.vocab if args.model == "gpt2" else None, "optimizer": args.opt if args.model == "gpt2" else "torch",
                          "ckpt_every": args.ckpt_every, "cuda_graph": bool(args.graph), "loss": round(float(loss), 4), "dtype": "bf16"}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
