#!/usr/bin/env python
"""Host->device ceiling of the input path: N ranks (one per GPU) each stream a training step's worth of uint8 pixels + labels
(18 944 x 785 B = 14.9 MB) from pinned host memory to their GPU, back to back, with no compute on the GPUs.

    python tools/h2d_ceiling.py                                   # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 tools/h2d_ceiling.py

Prints one JSON line (rank 0): per-rank GB/s (min / max over ranks), aggregate GB/s, and the ms per training step that
bandwidth allows — the floor of the end-to-end step time whatever the kernels do. ``--streams 2`` splits every step's copy
over two copy streams; ``--gather`` adds the loader's host-side shuffle (index_select of the next epoch into a second
pinned buffer on a background thread) to see what it costs the copies."""
import argparse
import json
import os
import threading
import time

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=18944)
    ap.add_argument("--row-bytes", type=int, default=785)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--dataset-steps", type=int, default=24, help="pinned dataset size in steps (copies cycle through it)")
    ap.add_argument("--streams", type=int, default=1)
    ap.add_argument("--gather", action="store_true")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    step_bytes = a.batch * a.row_bytes
    host = torch.empty(a.dataset_steps * step_bytes, dtype=torch.uint8).pin_memory()
    host.random_(0, 256)
    devbuf = [torch.empty(step_bytes, dtype=torch.uint8, device=dev) for _ in range(4)]
    streams = [torch.cuda.Stream(dev) for _ in range(a.streams)]
    stop = threading.Event()
    gathered = [0]
    if a.gather:
        src = host.view(a.dataset_steps * a.batch, a.row_bytes)
        dst = torch.empty_like(src).pin_memory()

        def work():
            while not stop.is_set():
                idx = torch.randperm(src.shape[0])
                torch.index_select(src, 0, idx, out=dst)
                gathered[0] += 1
        th = threading.Thread(target=work, daemon=True)
        th.start()

    def run(n):
        part = step_bytes // a.streams // 16 * 16
        for i in range(n):
            off = (i % a.dataset_steps) * step_bytes
            d = devbuf[i % 4]
            for k, s in enumerate(streams):
                lo = k * part
                hi = step_bytes if k == a.streams - 1 else lo + part
                with torch.cuda.stream(s):
                    d[lo:hi].copy_(host[off + lo:off + hi], non_blocking=True)

    run(10)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    run(a.steps)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    stop.set()
    gbs = a.steps * step_bytes / dt / 1e9
    t = torch.tensor([gbs, -gbs, dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    mx, mn, tmax = float(t[0]), -float(t[1]), float(t[2])
    out = {"what": "pinned host -> device copy ceiling, no compute", "n_gpus": world, "bytes_per_step": step_bytes, "steps": a.steps,
           "copy_streams": a.streams, "host_gather_running": bool(a.gather), "gather_epochs_done": gathered[0],
           "per_gpu_gbs_min": round(mn, 2), "per_gpu_gbs_max": round(mx, 2), "aggregate_gbs": round(world * a.steps * step_bytes / tmax / 1e9, 2),
           "ms_per_step_floor": round(tmax / a.steps * 1e3, 4), "cpu_count": os.cpu_count()}
    if rank == 0:
        print(json.dumps(out), flush=True)
        if a.out:
            with open(a.out, "a") as f:
                f.write(json.dumps(out) + "\n")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
