"""Checkpoint writer throughput on this host: stock torch.save vs the parallel archive writer (ckpt/fast_zip.py) on a
GPT-2-small-sized state (bf16 parameters + fp32 Adam moments, 1.24 GB, 148 x 3 tensors). CPU only.

    python tools/ckpt_writer_probe.py [--dir /tmp] [--threads 2 3 5 8]"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ray_torch_distributed_checkpoint_b200.ckpt import fast_zip  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--dir", default="/tmp")
ap.add_argument("--threads", type=int, nargs="*", default=[2, 3, 5, 8])
ap.add_argument("--reps", type=int, default=2)
args = ap.parse_args()
sizes = [50257 * 768, 1024 * 768] + [s for _ in range(12) for s in (768, 768, 2304 * 768, 2304, 768 * 768, 768, 768, 768, 3072 * 768, 3072, 768 * 3072, 768)] + [768, 768]
state = {"model_state_dict": {f"p{i}": torch.randn(n).to(torch.bfloat16) for i, n in enumerate(sizes)},
         "optimizer_state_dict": {"state": {i: {"exp_avg": torch.randn(n), "exp_avg_sq": torch.rand(n), "step": torch.tensor(3.0)} for i, n in enumerate(sizes)}}}
out = {"state_gb": round(fast_zip.state_bytes(state) / 1e9, 3), "cpus": os.cpu_count(), "dir": args.dir, "torch_save_s": [], "fast_zip_s": {}}
p_fast, p_ref = os.path.join(args.dir, "rtdc_probe_fast.pt"), os.path.join(args.dir, "rtdc_probe_ref.pt")
for _ in range(args.reps):
    t0 = time.perf_counter(); torch.save(state, p_ref); out["torch_save_s"].append(round(time.perf_counter() - t0, 3))
    for thr in args.threads:
        t0 = time.perf_counter(); fast_zip.save(state, p_fast, threads=thr)
        out["fast_zip_s"].setdefault(str(thr), []).append(round(time.perf_counter() - t0, 3))
a, b = torch.load(p_fast, weights_only=True), torch.load(p_ref, weights_only=True)
out["loads_equal"] = all(torch.equal(a["model_state_dict"][k], b["model_state_dict"][k]) for k in a["model_state_dict"])
os.remove(p_fast); os.remove(p_ref)
print(json.dumps(out))
