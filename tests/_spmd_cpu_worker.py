"""One rank of ``TorchTrainer.fit()`` in SPMD mode (the ranks were launched by torch.distributed.run, as bench.py's are at
N > 1) on CPU / gloo: asynchronous reports with checkpoints run on the report thread over the link's own gloo group while the
training thread issues collectives on the default group; every rank gets the Result."""
import os
import sys
import tempfile

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from ray_torch_distributed_checkpoint_b200 import train  # noqa: E402
from ray_torch_distributed_checkpoint_b200.train import CheckpointConfig, RunConfig, ScalingConfig, TorchTrainer  # noqa: E402


def train_fn(config):
    ctx = train.get_context()
    rank, world = ctx.get_world_rank(), ctx.get_world_size()
    assert world == 2 and dist.is_initialized() and dist.get_backend() == "gloo"
    w = torch.zeros(4)
    for epoch in range(config["epochs"]):
        for step in range(20):                                   # gradient-like traffic on the DEFAULT group while reports fly
            g = torch.full((4,), float(rank + 1))
            dist.all_reduce(g)
            w += g / world
        d = tempfile.mkdtemp(prefix="spmd_ck_")
        if rank == 0:
            torch.save({"epoch": epoch, "w": w.clone()}, os.path.join(d, "latest_model.pt"))
        ck = train.Checkpoint.from_directory(d) if rank == 0 else None
        train.report({"epoch": epoch, "w0": float(w[0]), "rank": rank}, checkpoint=ck, checkpoint_upload_mode="async")


def main():
    storage = sys.argv[1]
    trainer = TorchTrainer(train_fn, train_loop_config={"epochs": 4}, scaling_config=ScalingConfig(num_workers=2, use_gpu=False),
                           run_config=RunConfig(storage_path=storage, name="spmd", checkpoint_config=CheckpointConfig(num_to_keep=2)))
    result = trainer.fit()
    rank = int(os.environ["RANK"])
    assert result.error is None and result.metrics["epoch"] == 3 and result.metrics["rank"] == 0, result      # rank 0's row, on every rank
    assert abs(result.metrics["w0"] - 4 * 20 * 1.5) < 1e-6
    kept = sorted(d for d in os.listdir(result.path) if d.startswith("checkpoint_"))
    assert kept == ["checkpoint_000002", "checkpoint_000003"], kept
    with result.checkpoint.as_directory() as d:
        sd = torch.load(os.path.join(d, "latest_model.pt"), weights_only=True)
    assert sd["epoch"] == 3 and float(sd["w"][0]) == result.metrics["w0"]
    assert len(result.metrics_history) == 4
    assert not dist.is_initialized()                             # fit() created the group, fit() tore it down
    print(f"SPMD_OK rank {rank}", flush=True)


if __name__ == "__main__":
    main()
