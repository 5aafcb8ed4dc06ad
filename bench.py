#!/usr/bin/env python
"""Headline benchmark: train samples/sec of the FashionMNIST-shape MLP (784-512-512-10, fp32, SGD
momentum 0.9, dropout 0.25, DDP) at N B200 GPUs, max over ranks — BASELINE.json's metric on its config
"same MLP DDP on 8xB200, large global batch, synthetic 28x28 data".

    python bench.py --gpus 1 --steps 50 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
        bench.py --gpus 8 --steps 50 --warmup 5
    python bench.py --impl reference ...      # the unmodified reference (baseline/_ref) on stock torch DDP + NCCL

Weak scaling: the per-GPU batch is fixed (--batch-per-gpu), the global batch is N times that.

Two numbers per run:
* ``value``: K optimizer steps timed on the device with CUDA events (barrier + synchronize on both
  sides, max over ranks); inputs rotate over several device-resident batches so that each step's
  working set (> 126 MB L2) streams from HBM.
* ``e2e.value``: the same K steps driven through the public API a user calls — ``data.U8BatchLoader``
  (pinned host dataset -> per-step H2D of that step's uint8 pixels + int64 labels) feeding
  ``parallel.MLPTrainEngine.step`` — with the step's loss read back to the host every step.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "train samples/sec (MLP 784-512-512-10 fp32 DDP, max over ranks)"
MODEL = "FashionMNIST-shape MLP 784-512-512-10 (ReLU, Dropout 0.25, trailing ReLU), CrossEntropy, SGD lr 1e-3 momentum 0.9"


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clock + throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index: int, period_s: float = 0.05):
        self.index, self.period = index, period_s
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._stop = threading.Event()
        self._th = threading.Thread(target=self._run, daemon=True)
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception:
            pass

    _NAMES = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
              0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
              0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                bits = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for b, name in self._NAMES.items():
                    if bits & b and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        if self.ok:
            self._th.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self.ok:
            self._th.join(timeout=1.0)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        return {"sm_mhz": int(statistics.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def _dist_env():
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", rank))
    return rank, world, local


def _max_over_ranks(value: float, device) -> float:
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


# ------------------------------------------------------------------------------------------------ our arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    from ray_torch_distributed_checkpoint_b200 import data, ops
    from ray_torch_distributed_checkpoint_b200.models import NeuralNetwork
    from ray_torch_distributed_checkpoint_b200.parallel.engine import MLPTrainEngine

    rank, world, local = _dist_env()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}; launch with torch.distributed.run --nproc-per-node {args.gpus}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    from ray_torch_distributed_checkpoint_b200.utils.affinity import bind_to_gpu_numa
    numa = bind_to_gpu_numa(local)                           # pinned input buffers next to this GPU's PCIe root port
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)      # control plane + max-over-ranks only
    ops.load_library()
    B, K, W = args.batch_per_gpu, args.steps, args.warmup
    torch.manual_seed(1234)
    eng = MLPTrainEngine(NeuralNetwork(), batch_size=B, lr=1e-3, momentum=0.9, device=dev, seed=2026, algo=args.algo)

    # ---- synthetic FashionMNIST-shape data: uint8 pixels, int64 labels
    rot = args.rotate
    g = torch.Generator().manual_seed(99 + rank)
    dev_x = [torch.randint(0, 256, (B, 784), dtype=torch.uint8, generator=g).to(dev) for _ in range(rot)]
    dev_y = [torch.randint(0, 10, (B,), generator=g).to(dev) for _ in range(rot)]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- device-timed region. Batch i+1 is staged (device copy + uint8 -> fp32 conversion on the engine's staging
    # stream) before step i is launched, so the conversion overlaps that step; the timed region holds exactly K
    # stagings and K steps (the first staging is exposed, the others overlap).
    for i in range(max(W, 3)):
        eng.step(dev_x[i % rot], dev_y[i % rot])
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        barrier()
        e0.record()
        eng.stage(dev_x[0], dev_y[0])
        for i in range(K):
            if i + 1 < K:
                eng.stage(dev_x[(i + 1) % rot], dev_y[(i + 1) % rot])
            eng.step()
        e1.record()
        torch.cuda.synchronize(dev)
        barrier()
    ms = _max_over_ranks(e0.elapsed_time(e1), dev)
    launches_per_step = eng.kernel_launches_per_step.get("train_u8", 0)
    final_loss = float(eng.loss.item())

    # ---- end to end through the public API: pinned host dataset -> H2D each step -> step -> loss D2H each step
    n_host = B * (K + max(W, 3))
    hx = torch.randint(0, 256, (n_host, 784), dtype=torch.uint8, generator=g)
    hy = torch.randint(0, 10, (n_host,), generator=g)
    loader = data.U8BatchLoader(hx, hy, B, dev, shuffle=True, seed=5, lookahead=args.lookahead,
                                prefetch_next_epoch=False)   # one epoch only: no background shuffle of a next one
    loss_host = torch.zeros(4, dtype=torch.float32).pin_memory()
    it = iter(loader)
    for _ in range(max(W, 3)):
        xb, yb = next(it)
        eng.step(xb, yb)
    barrier()
    losses = []
    pending = []                                              # (slot, event) of loss copies in flight
    t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
    wall0 = time.perf_counter()
    t0.record()
    xb, yb = next(it)                                         # step 0's H2D (exposed)
    eng.stage(xb, yb)
    for i in range(K):
        if i + 1 < K:
            xb, yb = next(it)                                 # next step's H2D (issued `lookahead` steps ahead) ...
            eng.stage(xb, yb)                                 # ... and conversion, overlapping this step
        loss = eng.step()
        slot = i % 4
        loss_host[slot].copy_(loss, non_blocking=True)        # this step's D2H
        ev = torch.cuda.Event(); ev.record()
        pending.append((slot, ev))
        if len(pending) > 1:                                  # read the previous step's loss while this one runs
            s_, e_ = pending.pop(0)
            e_.synchronize()
            losses.append(float(loss_host[s_]))
    for s_, e_ in pending:
        e_.synchronize()
        losses.append(float(loss_host[s_]))
    t1.record()
    torch.cuda.synchronize(dev)
    wall = time.perf_counter() - wall0
    barrier()
    e2e_ms = _max_over_ranks(max(t0.elapsed_time(t1), wall * 1e3), dev)
    assert len(losses) == K and all(l == l for l in losses)

    global_batch = B * world
    value = global_batch * K / (ms / 1e3)
    e2e_value = global_batch * K / (e2e_ms / 1e3)
    flops = 3209216 * global_batch * K
    out = {
        "metric": METRIC, "value": round(value, 1), "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": round(ms / K, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "fp32 (3xTF32 split on tcgen05 tensor cores, fp32 accumulate in TMEM; matches the reference's fp32)",
        "data": "synthetic FashionMNIST-shape uint8 28x28 images + int64 labels, random-init weights",
        "impl": "ours",
        "config": {"model": MODEL, "global_batch": global_batch, "batch_per_gpu": B, "seq_len": None,
                   "parallelism": f"dp{world}", "l2_policy": f"no flush: inputs rotate over {rot} device batches and the "
                   f"per-step working set (~{int(B * (784 * 5 + 512 * 16 + 128) / 1e6)} MB of activations) exceeds the 126 MB L2",
                   "allreduce": f"{args.algo}/{eng.ar_mode}", "cuda_graph": True, "host": {"numa": numa, "cpu_count": os.cpu_count()}},
        "clocks": clk.summary(),
        "e2e": {"value": round(e2e_value, 1), "unit": "samples/s", "ms_per_step": round(e2e_ms / K, 4),
                "h2d_bytes_per_step": B * (784 + 8), "d2h_bytes_per_step": 4,
                "api": "data.U8BatchLoader -> parallel.MLPTrainEngine.stage/step -> loss.copy_(pinned)"},
        "gpu_launches": launches_per_step * K,
        "gpu_launches_per_step": launches_per_step,
        "tflops_fp32_equiv": round(flops / (ms / 1e3) / 1e12, 2),
        "final_loss": round(final_loss, 4),
    }
    if rank == 0:
        print(json.dumps(out), flush=True)
    eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args):
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    src = os.path.join(ref_dir, "my_ray_module.py")
    rank, world, local = _dist_env()
    if not os.path.exists(src):
        if os.path.exists("/root/reference/my_ray_module.py") and rank == 0:
            import shutil
            os.makedirs(ref_dir, exist_ok=True)
            for f in ("my_ray_module.py", "train_flow.py", "eval_flow.py", "README.md"):
                shutil.copy("/root/reference/" + f, os.path.join(ref_dir, f))
        if not os.path.exists(src):
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref/my_ray_module.py missing (the reference is "
                                  "not pip-installable: no setup.py/pyproject; its sources were not shipped to this box)"}))
            return
    try:
        import torch
        import torchvision  # noqa: F401
        import filelock  # noqa: F401
    except Exception as e:
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": f"reference dependency missing: {e!r}"}))
        return
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    B, K, W = args.batch_per_gpu, args.steps, max(args.warmup, 3)
    steps_total = K + W
    # the reference hard-codes ~/data and a download; give it a HOME with synthetic idx files of the right shape
    home = os.path.join(ROOT, ".ref_home")                    # scratch inside the repo copy, ignored by git and gpurun
    os.makedirs(home, exist_ok=True)
    os.environ["HOME"] = home
    n_train = B * world * steps_total
    marker = os.path.join(home, f"data_ready_{n_train}")
    if rank == 0 and not os.path.exists(marker):
        sys.path.insert(0, os.path.join(ROOT, "baseline"))
        import make_idx
        make_idx.write(os.path.join(home, "data"), n_train=n_train, n_test=max(256, world * 32))
        open(marker, "w").close()
    t_wait = time.time()
    while not os.path.exists(marker):
        time.sleep(0.2)
        if time.time() - t_wait > 600:
            raise SystemExit("timed out waiting for the synthetic dataset")
    sys.path.insert(0, os.path.join(ROOT, "baseline", "ray_standin"))
    sys.path.insert(0, ref_dir)
    for mod in [m for m in sys.modules if m.startswith("ray_torch_distributed_checkpoint_b200")]:
        del sys.modules[mod]                                  # nothing of ours stays importable state on this path
    import my_ray_module as ref
    from ray.train.torch import TIMING
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    storage = os.path.join(home, "storage")
    with ClockSampler(local) as clk:
        result = ref.train_fashion_mnist(num_workers=world, use_gpu=True, global_batch_size=B * world, learning_rate=1e-3,
                                         epochs=1, checkpoint_storage_path=storage)
    torch.cuda.synchronize(dev)
    rec = TIMING["records"]
    begins = [r for r in rec if r[0] == "step_begin"]
    ends = [r for r in rec if r[0] == "step_end"]
    n = min(len(begins), len(ends))
    if n < W + K:
        K = max(1, n - W)
    # kernel-side time of each timed step: begin event (recorded right before the step's launches, GPU idle
    # while the reference's CPU loader ran) to end event (after optimizer.step())
    dev_ms = sum(begins[i][2].elapsed_time(ends[i][2]) for i in range(W, W + K))
    # end to end: first timed batch handed to the loop -> last timed step finished, loader stalls included
    e2e_ms = begins[W][2].elapsed_time(ends[W + K - 1][2])
    dev_ms = _max_over_ranks(dev_ms, dev)
    e2e_ms = _max_over_ranks(e2e_ms, dev)
    gb = B * world
    out = {
        "metric": METRIC, "value": round(gb * K / (dev_ms / 1e3), 1), "unit": "samples/s", "n_gpus": world, "steps": K,
        "warmup": W, "ms_per_step": round(dev_ms / K, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "fp32", "data": "synthetic FashionMNIST-shape idx files read by torchvision.datasets.FashionMNIST, random-init weights",
        "impl": "reference",
        "config": {"model": MODEL, "global_batch": gb, "batch_per_gpu": B, "seq_len": None, "parallelism": f"dp{world}",
                   "entry": "baseline/_ref/my_ray_module.py::train_fashion_mnist (unmodified) on a stock-torch stand-in for "
                            "ray.train (torch DDP + NCCL + cuBLAS); value = sum of per-step device intervals (the "
                            "reference's CPU DataLoader stalls excluded), e2e = loader included"},
        "clocks": clk.summary(),
        "e2e": {"value": round(gb * K / (e2e_ms / 1e3), 1), "unit": "samples/s", "ms_per_step": round(e2e_ms / K, 4),
                "h2d_bytes_per_step": B * (784 * 4 + 8), "d2h_bytes_per_step": 0,
                "api": "my_ray_module.train_fashion_mnist (torchvision dataset -> DataLoader -> prepare_data_loader -> DDP step)"},
        "gpu_launches": 0,
        "final_metrics": getattr(result, "metrics", None),
    }
    if rank == 0:
        print(json.dumps(out), flush=True)
    import torch.distributed as dist
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--batch-per-gpu", type=int, default=int(os.environ.get("RTDC_BENCH_BATCH", 18944)),
                    help="per-GPU batch; 18944 = 148 SMs x 128 rows, i.e. exactly two waves of 128x256 output tiles")
    ap.add_argument("--rotate", type=int, default=4, help="distinct device-resident input batches for the device-timed loop")
    ap.add_argument("--lookahead", type=int, default=3, help="H2D batches in flight ahead of the consumer (e2e loop)")
    ap.add_argument("--algo", default="auto", choices=["auto", "two_shot", "nvls"])
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
