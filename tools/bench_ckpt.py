"""Checkpoint save / restore timings (BASELINE.md rows "MLP checkpoint save" and "restore + eval").

save:    reference style = two synchronous torch.save calls of the GPU state dict (latest + best, my_ray_module.py:179-201)
         ours            = AsyncCheckpointer.save critical path (device snapshot + enqueue), files written in the background
restore: torch.load(weights_only=True) + load_state_dict + eval of 10 000 x 784 at batch 512 (eval_flow.py:37,83-90)
Prints one JSON line."""
import json
import os
import shutil
import sys
import tempfile
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ray_torch_distributed_checkpoint_b200 import ckpt  # noqa: E402
from ray_torch_distributed_checkpoint_b200.models import NeuralNetwork, fuse_linear_chains  # noqa: E402
from ray_torch_distributed_checkpoint_b200.parallel.engine import MLPTrainEngine  # noqa: E402

dev = torch.device("cuda", 0)
torch.manual_seed(0)
B = 4096
eng = MLPTrainEngine(NeuralNetwork(), batch_size=B, device=dev, distributed=False)
x = torch.randint(0, 256, (B, 784), dtype=torch.uint8, device=dev); y = torch.randint(0, 10, (B,), device=dev)
for _ in range(4):
    eng.step(x, y)
torch.cuda.synchronize()
state = lambda: {"epoch": 0, "model_state_dict": eng.model_state_dict(), "optimizer_state_dict": eng.optimizer_state_dict(),
                 "val_losses": [1.0], "val_accuracy": [0.5]}
tmp = tempfile.mkdtemp(prefix="rtdc_ckpt_bench_")
res = {}
# reference style
ts = []
for i in range(5):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    torch.save(state(), os.path.join(tmp, "latest_model.pt"))
    torch.save(state(), os.path.join(tmp, "best_model.pt"))
    torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
res["sync_torch_save_x2_ms"] = round(sorted(ts)[len(ts) // 2], 3)
# ours
cp = ckpt.AsyncCheckpointer(dev)
ts, tot = [], []
for i in range(5):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    pending = cp.save(state(), os.path.join(tmp, f"a{i}"), ["latest_model.pt", "best_model.pt"])
    ts.append((time.perf_counter() - t0) * 1e3)
    for _ in range(3):
        eng.step(x, y)                      # training continues while the files are written
    pending.result(); tot.append((time.perf_counter() - t0) * 1e3)
res["async_critical_path_ms"] = round(sorted(ts)[len(ts) // 2], 3)
res["async_total_until_published_ms"] = round(sorted(tot)[len(tot) // 2], 3)
res["file_bytes"] = os.path.getsize(os.path.join(tmp, "a0", "latest_model.pt"))
# restore + eval
xs = torch.randn(10000, 1, 28, 28)
ys = torch.randint(0, 10, (10000,))
def restore_eval(fused):
    t0 = time.perf_counter()
    sd = torch.load(os.path.join(tmp, "a0", "best_model.pt"), map_location=dev, weights_only=True)
    m = NeuralNetwork()
    m.load_state_dict({k.replace("module.", ""): v for k, v in sd["model_state_dict"].items()})
    m.to(dev)
    if fused:
        fuse_linear_chains(m)
    m.eval()
    correct = 0
    with torch.inference_mode():
        for i in range(0, 10000, 512):
            logits = m(xs[i:i + 512].to(dev)).cpu()
            correct += int((logits.argmax(1) == ys[i:i + 512]).sum())
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) * 1e3
restore_eval(True); restore_eval(False)
res["restore_eval_stock_modules_ms"] = round(min(restore_eval(False) for _ in range(3)), 3)
res["restore_eval_fused_kernels_ms"] = round(min(restore_eval(True) for _ in range(3)), 3)
print(json.dumps(res))
shutil.rmtree(tmp, ignore_errors=True)
