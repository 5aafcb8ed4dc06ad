"""Multi-GPU tier (needs >= 2 GPUs on one NVSwitch box; skipped otherwise): NVLink all-reduce variants vs
NCCL, bit-identical replicas, flag-reuse stress, DDP wrapper vs torch DDP, distributed engine."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.parametrize("what", ["collectives", "ddp", "engine", "trainer"])
def test_multi_gpu(what):
    n = _ngpu()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    n = 8 if n >= 8 else (4 if n >= 4 else 2)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tests", "_mgpu_worker.py"), what]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    with open(os.path.join(out_dir, f"mgpu_tests_n{n}.log"), "a") as f:          # kept as evidence (profiles/)
        f.write(f"$ {' '.join(cmd)}\n{p.stdout}\n{p.stderr[-2000:] if p.returncode else ''}\n")
    assert p.returncode == 0, p.stdout[-4000:] + p.stderr[-4000:]
    assert "[mgpu] all ok" in p.stdout
