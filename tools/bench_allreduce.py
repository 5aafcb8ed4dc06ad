"""Gradient-bucket all-reduce latency: this framework's NVLink kernels (two-shot P2P, NVLS multimem) vs NCCL, on the
bucket sizes of SURVEY.md §2.5-C3. Device-timed (CUDA events around a CUDA graph of REPS back-to-back calls), max over ranks.
Launch: python -m torch.distributed.run --nproc-per-node N tools/bench_allreduce.py"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ray_torch_distributed_checkpoint_b200.parallel.symm import SymmetricArena  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
arena = SymmetricArena(272 << 20, device=dev)
REPS = 20
sizes = {"mlp_bucket0": 1071144, "mlp_bucket1": 1607680, "mlp_merged": 2678824, "resnet_bf16_27MiB": 27 << 20, "gpt2_bf16_98MiB": 98 << 20,
         "64KiB": 65536}


def timed(fn):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(3):
            fn(s)
        s.synchronize(); dist.barrier()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(REPS):
                fn(s)
        g.replay(); s.synchronize(); dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(s); g.replay(); b.record(s); s.synchronize()
    t = torch.tensor([a.elapsed_time(b) / REPS * 1e3], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


out = {"world": world, "multicast": arena.has_multicast}
# ---- the fused distributed optimizer step on the MLP's arenas (gradients / parameters / momentum, 669 728 fp32 each), next to
# what it replaces: an all-reduce of the gradient arena followed by the full-arena SGD kernel on every rank
from ray_torch_distributed_checkpoint_b200 import ops  # noqa: E402
N_ARENA = 669728
g_off, g = arena.alloc(N_ARENA)
p_off, p = arena.alloc(N_ARENA)
m_off, m = arena.alloc(N_ARENA)
g.fill_(1e-3); p.fill_(0.5); m.zero_()
sgd = ops.SgdState(dev, 1e-3, 0.9)
counter = torch.zeros(1, dtype=torch.int32, device=dev)
row = {"elements": N_ARENA}
for algo in ["two_shot"] + (["nvls"] if arena.has_multicast else []):
    for blocks in (0, 16, 32, 64):
        row[f"fused_sgd_{algo}_b{blocks}_us"] = round(timed(lambda s: arena.sgd_step(g_off, p_off, m_off, N_ARENA, sgd, counter=counter, algo=algo,
                                                                                    max_blocks=blocks, stream=s)), 2)
    def unfused(s, algo=algo):
        arena.all_reduce(g_off, N_ARENA, scale=1.0 / world, algo=algo, stream=s)
        ops.sgd_step(p, g, m, sgd, stream=s, advance=True, counter=counter)
    row[f"allreduce_{algo}_then_sgd_us"] = round(timed(unfused), 2)
row["barrier_us"] = round(timed(lambda s: arena.barrier(stream=s)), 2)
out["mlp_fused_step"] = row
if rank == 0:
    print("mlp_fused_step", json.dumps(row), flush=True)
for name, nbytes in sizes.items():
    n = (nbytes // 4 + 3) // 4 * 4
    buf = arena.tensor(16 << 20, n)
    buf.fill_(1.0)
    nccl_buf = torch.ones(n, device=dev)
    row = {"bytes": n * 4}
    row["two_shot_us"] = round(timed(lambda s: arena.all_reduce(16 << 20, n, scale=1.0 / world, algo="two_shot", stream=s)), 2)
    if arena.has_multicast:
        row["nvls_us"] = round(timed(lambda s: arena.all_reduce(16 << 20, n, scale=1.0 / world, algo="nvls", stream=s)), 2)
    row["nccl_us"] = round(timed(lambda s: dist.all_reduce(nccl_buf, op=dist.ReduceOp.AVG)), 2)
    best = min(v for k, v in row.items() if k.endswith("_us") and k != "nccl_us")
    # bus bandwidth convention: 2*(W-1)/W * bytes / time
    row["ours_busbw_GBps"] = round(2 * (world - 1) / world * n * 4 / best / 1e3, 1)
    row["nccl_busbw_GBps"] = round(2 * (world - 1) / world * n * 4 / row["nccl_us"] / 1e3, 1)
    out[name] = row
    if rank == 0:
        print(name, json.dumps(row), flush=True)
if rank == 0:
    print(json.dumps(out))
dist.barrier()
arena.close()
dist.destroy_process_group()
