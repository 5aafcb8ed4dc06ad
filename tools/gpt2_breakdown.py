"""Where a GPT-2 small training step spends its GPU time, by kernel (CUPTI through torch.profiler; one GPU).

    python tools/gpt2_breakdown.py --arm ours|torch [--steps 3] [--top 30] [--out gpurun_out/gpt2_breakdown_ours.md]

Numbers taken under the profiler are for ATTRIBUTION only (which kernels, what share); step times come from
tools/bench_models.py. The two arms are the ones of bench_models.py: stock torch (cuBLAS Linears, F.cross_entropy, fused
AdamW) and this framework (bf16 tcgen05 Linears, fused LM head + loss, AdamW kernel)."""
import argparse
import collections
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from ray_torch_distributed_checkpoint_b200.models import GPT2, GPT2Config  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arm", choices=["ours", "torch"], default="ours")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--top", type=int, default=30)
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    model = GPT2(GPT2Config()).to(dev, dtype=torch.bfloat16)
    if args.arm == "ours":
        from ray_torch_distributed_checkpoint_b200.ops import use_bf16_linears, use_fused_layernorm
        from ray_torch_distributed_checkpoint_b200.optim import AdamW
        use_bf16_linears(model)
        model.fused_head = True
        use_fused_layernorm(model)
        opt = AdamW(model, lr=3e-4)
    else:
        opt = torch.optim.AdamW(model.parameters(), lr=3e-4, fused=True)
    idx = torch.randint(0, 50257, (args.batch, 1024), device=dev)
    tgt = torch.randint(0, 50257, (args.batch, 1024), device=dev)

    def step():
        opt.zero_grad(set_to_none=False)
        loss = model(idx, tgt)
        loss.backward()
        opt.step()

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(args.steps):
            step()
        torch.cuda.synchronize()
    per = collections.defaultdict(lambda: [0, 0.0])
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            dur = getattr(ev, "device_time", None)
            if dur is None:
                dur = getattr(ev, "cuda_time", 0.0)
            per[ev.name][0] += 1
            per[ev.name][1] += float(dur)
    total = sum(v[1] for v in per.values())
    rows = sorted(per.items(), key=lambda kv: -kv[1][1])
    lines = [f"# GPT-2 small, one GPU, arm = {args.arm}: GPU time by kernel over {args.steps} steps (under the profiler: shares, not step times)",
             "", f"sum of kernel time per step: {total / args.steps / 1e3:.2f} ms", "",
             "| kernel | launches / step | ms / step | share |", "|---|---:|---:|---:|"]
    for name, (n, us) in rows[:args.top]:
        lines.append(f"| `{name[:110]}` | {n / args.steps:.1f} | {us / args.steps / 1e3:.3f} | {100 * us / total:.1f} % |")
    rest = rows[args.top:]
    if rest:
        lines.append(f"| ({len(rest)} more kernels) | {sum(v[0] for _, v in rest) / args.steps:.1f} | {sum(v[1] for _, v in rest) / args.steps / 1e3:.3f} | "
                     f"{100 * sum(v[1] for _, v in rest) / total:.1f} % |")
    text = "\n".join(lines) + "\n"
    print(text)
    if args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as f:
            f.write(text)


if __name__ == "__main__":
    main()
