#!/usr/bin/env python
"""Where does an input step's time go? (1) H2D of one step's pixels as N pieces of various sizes, no compute; (2) each
U8BatchLoader transfer mode iterated alone: host time per batch and device time per batch; (3) the engine-backed API loop
fed by each mode: host time per step (time between consecutive batches handed out) and device time per step.
One GPU: python tools/input_probe.py [--out gpurun_out/input_probe.json]"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=18944)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    from ray_torch_distributed_checkpoint_b200 import ops
    from ray_torch_distributed_checkpoint_b200.data import U8BatchLoader
    from ray_torch_distributed_checkpoint_b200.models import NeuralNetwork
    from ray_torch_distributed_checkpoint_b200.nn import CrossEntropyLoss
    from ray_torch_distributed_checkpoint_b200.optim import SGD
    from ray_torch_distributed_checkpoint_b200.train import session
    from ray_torch_distributed_checkpoint_b200.train.torch import prepare_model
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    B, S = a.batch, a.steps
    res = {"batch": B, "steps": S}
    n = B * S
    host = torch.randint(0, 256, (n, 784), dtype=torch.uint8).pin_memory()
    labels = torch.randint(0, 10, (n,))
    dst = [torch.empty(B, 784, dtype=torch.uint8, device=dev) for _ in range(4)]
    st = torch.cuda.Stream(dev)
    # ---- (1) piece size sweep
    sweep = {}
    for rows in (B, 4096, 2048, 1024, 512, 256, 128):
        pieces = -(-B // rows)
        sizes = torch.tensor([min(rows, B - i * rows) * 784 for i in range(pieces)], dtype=torch.int64)
        g = torch.Generator().manual_seed(rows)
        t_host = 0.0
        for rep in range(2):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for i in range(S):
                starts = torch.randint(0, n - rows, (pieces,), generator=g, dtype=torch.int64) * 784
                h0 = time.perf_counter()
                ops.h2d_pieces(dst[i % 4], host, starts, sizes, stream=st)
                t_host += time.perf_counter() - h0
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        sweep[rows] = {"pieces": pieces, "ms_per_step": round(dt / S * 1e3, 4), "gbs": round(B * 784 * S / dt / 1e9, 1),
                       "host_issue_us_per_step": round(t_host / (2 * S) * 1e6, 1)}
    res["h2d_piece_rows_sweep"] = sweep
    print(json.dumps({"h2d_piece_rows_sweep": sweep}), flush=True)

    # ---- (2) loaders alone
    def run_loader(mode, block):
        ld = U8BatchLoader(host, labels, B, dev, shuffle=True, transfer=mode, shuffle_block=block, lookahead=3, wrap=True)
        out = {}
        for ep in range(2):
            ld.sampler.set_epoch(ep)
            torch.cuda.synchronize()
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            e0.record()
            k = 0
            for xb, yb in ld:
                k += 1
            e1.record()
            t_host = time.perf_counter() - t0
            torch.cuda.synchronize()
            out = {"host_ms_per_batch": round(t_host / k * 1e3, 4), "device_ms_per_batch": round(e0.elapsed_time(e1) / k, 4)}
        ld.close()
        return out
    res["loader_alone"] = {}
    for mode, block in (("resident", 0), ("stream", 512), ("stream", 2048), ("host_gather", 0)):
        res["loader_alone"][f"{mode}{block or ''}"] = run_loader(mode, block)
    print(json.dumps({"loader_alone": res["loader_alone"]}), flush=True)

    # ---- (3) API loop
    session._set_session(session._Session(session.TrainContext(use_gpu=True), None, None))
    res["api_loop"] = {}
    for mode, block in (("resident", 0), ("stream", 512), ("stream", 2048)):
        torch.manual_seed(0)
        model = prepare_model(NeuralNetwork())
        loss_fn, opt = CrossEntropyLoss(), SGD(model, lr=1e-3, momentum=0.9)
        ld = U8BatchLoader(host, labels, B, dev, shuffle=True, transfer=mode, shuffle_block=block, lookahead=3, wrap=True)
        for ep in range(3):
            ld.sampler.set_epoch(ep)
            torch.cuda.synchronize()
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            stamps = []
            it = iter(ld)
            first = True
            for xb, yb in it:
                if first:
                    e0.record(); first = False
                stamps.append(time.perf_counter())
                loss = loss_fn(model(xb), yb)
                opt.zero_grad(); loss.backward(); opt.step()
            e1.record()
            host_done = time.perf_counter()
            torch.cuda.synchronize()
            gaps = sorted(b - a_ for a_, b in zip(stamps, stamps[1:]))
            out = {"host_ms_per_step_median": round(gaps[len(gaps) // 2] * 1e3, 4), "host_ms_per_step_mean": round((host_done - stamps[0]) / len(stamps) * 1e3, 4),
                   "device_ms_per_step": round(e0.elapsed_time(e1) / len(stamps), 4)}
        res["api_loop"][f"{mode}{block or ''}"] = out
        ld.close()
        model.close()
    print(json.dumps({"api_loop": res["api_loop"]}), flush=True)
    if a.out:
        with open(a.out, "w") as f:
            json.dump(res, f, indent=1)


if __name__ == "__main__":
    main()
