"""FashionMNIST-shape MLP training on the B200-native framework — the counterpart of the reference's
``my_ray_module.py`` (same functions, constants and checkpoint format), written against
``ray_torch_distributed_checkpoint_b200``.

What differs from the reference on purpose (SURVEY.md §3.2 hot spots):
* on a GPU worker ``prepare_model`` recognises this network and backs it with the whole-step engine: the training
  loop below — the reference's, line for line — replays ONE CUDA graph of the fused sm_100a kernels per step
  (forward, loss, backward, NVLink gradient reduce-scatter + SGD + parameter multicast); ``prepare_data_loader`` keeps the
  images as bytes in pinned memory and moves uint8 pixels and labels over PCIe;
* validation keeps its loss / correct-count accumulators on the device (``nn.EvalAccumulator``): one D2H per epoch
  instead of two ``.item()`` syncs per batch (ref ``my_ray_module.py:168,170``);
* the checkpoint is snapshotted asynchronously (device copy -> pinned D2H on a side stream -> writer
  thread) and only rank 0 writes files; names, dict keys and the ``torch.save`` format are the reference's;
* ``set_weights_from_checkpoint`` falls back to ``latest_model.pt`` when ``best_model.pt`` is absent in
  the latest checkpoint directory (a latent FileNotFoundError in the reference).
"""
from __future__ import annotations

import os
import tempfile
import time
from typing import Dict

import numpy as np
import torch
from torch.utils.data import DataLoader

import ray_torch_distributed_checkpoint_b200 as rtdc
from ray_torch_distributed_checkpoint_b200 import ckpt, data, train
from ray_torch_distributed_checkpoint_b200.models import NeuralNetwork
from ray_torch_distributed_checkpoint_b200.nn import CrossEntropyLoss, EvalAccumulator
from ray_torch_distributed_checkpoint_b200.optim import SGD
from ray_torch_distributed_checkpoint_b200.train import Checkpoint, CheckpointConfig, RunConfig, ScalingConfig
from ray_torch_distributed_checkpoint_b200.train.torch import (TorchTrainer, get_device, prepare_data_loader,
                                                                prepare_model)

BEST_CHECKPOINT_FILENAME = ckpt.BEST_CHECKPOINT_FILENAME      # "best_model.pt"   (ref my_ray_module.py:27)
LATEST_CHECKPOINT_FILENAME = ckpt.LATEST_CHECKPOINT_FILENAME  # "latest_model.pt" (ref my_ray_module.py:28)

_TRAIN_SIZE = int(os.environ.get("RTDC_TRAIN_SIZE", 60000))
_TEST_SIZE = int(os.environ.get("RTDC_TEST_SIZE", 10000))


def get_dataloaders(batch_size, val_only=False, as_ray_ds=False):
    """Train/validation loaders (or ``data.Dataset`` objects with ``as_ray_ds=True``) over the synthetic
    FashionMNIST-shape data; ``val_only=True`` returns just the test split (ref my_ray_module.py:30-76)."""
    real = _real_fashion_mnist()
    if real is not None:
        training, test = real
        if val_only:
            return _as_ds(test) if as_ray_ds else DataLoader(test, batch_size=batch_size)
        if as_ray_ds:
            return _as_ds(training), _as_ds(test)
        return DataLoader(training, batch_size=batch_size, shuffle=True), DataLoader(test, batch_size=batch_size)
    test = data.FashionMNISTSynthetic(train=False, size=_TEST_SIZE)
    if val_only:
        if as_ray_ds:
            return data.from_numpy({"features": test.normalized().numpy(), "labels": test.targets.numpy()})
        return DataLoader(test, batch_size=batch_size)
    training = data.FashionMNISTSynthetic(train=True, size=_TRAIN_SIZE)
    if as_ray_ds:
        return (data.from_numpy({"features": training.normalized().numpy(), "labels": training.targets.numpy()}),
                data.from_numpy({"features": test.normalized().numpy(), "labels": test.targets.numpy()}))
    return DataLoader(training, batch_size=batch_size, shuffle=True), DataLoader(test, batch_size=batch_size)


def _real_fashion_mnist():
    """The real dataset when it is already on disk under ~/data (there is no network to download it).
    Guarded by the same file lock the reference uses so co-located workers do not race (ref :41,54)."""
    root = os.path.expanduser("~/data")
    if os.environ.get("RTDC_SYNTHETIC_DATA") or not os.path.isdir(os.path.join(root, "FashionMNIST", "raw")):
        return None
    from filelock import FileLock
    from torchvision import datasets, transforms
    tf = transforms.Compose([transforms.ToTensor(), transforms.Normalize((0.5,), (0.5,))])
    with FileLock(os.path.expanduser("~/data.lock")):
        return (datasets.FashionMNIST(root=root, train=True, download=False, transform=tf),
                datasets.FashionMNIST(root=root, train=False, download=False, transform=tf))


def _as_ds(torch_dataset):
    xs = ((torch_dataset.data.to(torch.float32) / 255.0 - 0.5) / 0.5).unsqueeze(1).numpy()
    return data.from_numpy({"features": xs, "labels": torch_dataset.targets.numpy()})


get_labels_map = data.get_labels_map


def _checkpoint_state(epoch, model, optimizer, val_losses, val_acc):
    return {"epoch": epoch, "model_state_dict": model.state_dict(), "optimizer_state_dict": optimizer.state_dict(),
            "val_losses": list(val_losses), "val_accuracy": list(val_acc)}


def train_func_per_worker(config: Dict):
    lr, epochs, batch_size = config["lr"], config["epochs"], config["batch_size_per_worker"]
    warm_start = config.get("checkpoint")
    exact_resume = bool(config.get("exact_resume", False))
    device = get_device()
    ctx = train.get_context()
    rank, world = ctx.get_world_rank(), ctx.get_world_size()

    print("[my_module] Preparing distributed data loaders...")
    train_loader, val_loader = get_dataloaders(batch_size=batch_size)
    transfer = config.get("input_transfer")          # None: "auto" (dataset resident in HBM when it fits, else streamed)
    train_loader = prepare_data_loader(train_loader, transfer=transfer)
    val_loader = prepare_data_loader(val_loader, transfer=transfer)

    model = NeuralNetwork()
    start_epoch, val_losses, val_acc, best_val_loss = 0, [], [], float("inf")
    resume = train.get_checkpoint()              # set after a worker-group restart (FailureConfig) or resume_from_checkpoint
    saved = None
    if resume is not None:
        print(f"[my_module] Restarting from the run's own checkpoint at {resume.path}.")
        saved = load_checkpoint_dict(resume, "cpu", prefer_best=False)
        model.load_state_dict(strip_module_prefix(saved["model_state_dict"]))
        start_epoch = saved["epoch"] + 1
        val_losses, val_acc = list(saved["val_losses"]), list(saved["val_accuracy"])
        best_val_loss = min(val_losses) if val_losses else float("inf")
    elif warm_start is not None:
        print(f"[my_module] Resuming from checkpoint at {warm_start.path}.")
        saved = set_weights_from_checkpoint(model, warm_start, "cpu") if exact_resume else None
        if saved is None:
            set_weights_from_checkpoint(model, warm_start, "cpu")
    model = prepare_model(model)
    print("[my_module] Model on-device. Training model...")

    loss_fn = CrossEntropyLoss()
    optimizer = SGD(model, lr=lr, momentum=0.9)
    if saved is not None and (resume is not None or exact_resume):
        optimizer.load_state_dict(saved["optimizer_state_dict"])
    checkpointer = ckpt.AsyncCheckpointer(device) if rank == 0 else None

    t0_full = time.time()
    phases = {"train": 0.0, "validate": 0.0, "checkpoint": 0.0, "report": 0.0}     # host seconds per phase (RTDC_PROFILE_EPOCHS)
    for epoch in range(start_epoch, epochs):
        t0 = time.time()
        if world > 1:
            train_loader.sampler.set_epoch(epoch)     # reshuffle the shards each epoch
        model.train()
        for X, y in train_loader:
            pred = model(X)
            loss = loss_fn(pred, y)
            optimizer.zero_grad()
            loss.backward()
            optimizer.step()

        t1 = time.time()
        model.eval()
        metrics = EvalAccumulator(model)                      # loss sum and correct count stay on the device
        with torch.no_grad():
            for X, y in val_loader:
                pred = model(X)
                loss = loss_fn(pred, y)
                metrics.update(loss, pred, y)
        val_loss, accuracy = metrics.compute()                # the epoch's only device->host sync
        val_losses.append(val_loss)
        val_acc.append(accuracy)
        t2 = time.time()

        pending = None
        if rank == 0:
            names = [LATEST_CHECKPOINT_FILENAME]
            if val_loss < best_val_loss:
                names.append(BEST_CHECKPOINT_FILENAME)
            state = _checkpoint_state(epoch, model, optimizer, val_losses, val_acc)
            t2b = time.time()
            pending = checkpointer.save(state, tempfile.mkdtemp(prefix="rtdc_ckpt_"), names)
            if os.environ.get("RTDC_PROFILE_EPOCHS"):
                print(f"[my_module] state_dict {(t2b - t2) * 1e3:.2f} ms, save() {(time.time() - t2b) * 1e3:.2f} ms", flush=True)
        best_val_loss = min(best_val_loss, val_loss)
        t3 = time.time()
        # every rank reports; the checkpoint write, its persistence and the cross-worker barrier finish in
        # the background while the next epoch trains
        train.report({"val_loss": val_loss, "accuracy": accuracy, "epoch": epoch}, checkpoint=pending,
                     checkpoint_upload_mode="async")
        t4 = time.time()
        for k, dt in zip(phases, (t1 - t0, t2 - t1, t3 - t2, t4 - t3)):
            phases[k] += dt
        print(f"[my_module] Model on-device. Last epoch took {round((time.time() - t0) / 60, 3)} minutes. Training model...")
    print(f"[my_module] Training completed in {round((time.time() - t0_full) / 60, 3)} minutes!")
    if os.environ.get("RTDC_PROFILE_EPOCHS") and epochs > start_epoch:
        n = epochs - start_epoch
        print(f"[my_module] rank {rank} host ms per epoch: " + ", ".join(f"{k} {v / n * 1e3:.2f}" for k, v in phases.items()))


def train_fashion_mnist(num_workers=1, use_gpu=False, global_batch_size=32, learning_rate=1e-3, epochs=10,
                        num_checkpoints_to_keep=2, checkpoint_storage_path=None, checkpoint=None, exact_resume=False,
                        max_failures=0, input_transfer=None):
    """Same signature and defaults as the reference's driver (my_ray_module.py:216-251), plus
    ``exact_resume`` (also restore optimizer state / epoch), ``max_failures`` (worker-group restarts) and
    ``input_transfer`` (``prepare_data_loader(transfer=...)``: "resident", "stream" or "host_gather"; default "auto")."""
    train_config = {"lr": learning_rate, "epochs": epochs, "batch_size_per_worker": global_batch_size // num_workers,
                    "exact_resume": exact_resume, "input_transfer": input_transfer}
    if checkpoint is not None:
        train_config["checkpoint"] = checkpoint
    run_config = RunConfig(checkpoint_config=CheckpointConfig(num_to_keep=num_checkpoints_to_keep),
                           storage_path=checkpoint_storage_path, verbose=1,
                           failure_config=train.FailureConfig(max_failures=max_failures))
    trainer = TorchTrainer(train_loop_per_worker=train_func_per_worker, train_loop_config=train_config,
                           scaling_config=ScalingConfig(num_workers=num_workers, use_gpu=use_gpu), run_config=run_config)
    return trainer.fit()


def strip_module_prefix(state_dict):
    """DDP-wrapped models save keys as ``module.<name>``; a bare model wants ``<name>`` (ref :260-263)."""
    return {(k[len("module."):] if k.startswith("module.") else k): v for k, v in state_dict.items()}


def load_checkpoint_dict(checkpoint: Checkpoint, device, prefer_best: bool = True):
    with checkpoint.as_directory() as d:
        order = [BEST_CHECKPOINT_FILENAME, LATEST_CHECKPOINT_FILENAME] if prefer_best else [LATEST_CHECKPOINT_FILENAME, BEST_CHECKPOINT_FILENAME]
        for name in order:
            path = os.path.join(d, name)
            if os.path.exists(path):
                return torch.load(path, map_location=device, weights_only=True)
        raise FileNotFoundError(f"neither {order[0]} nor {order[1]} in checkpoint {checkpoint.path}")


def set_weights_from_checkpoint(model_structure, checkpoint, device):
    """Load model weights (only) from a checkpoint into a bare model (ref my_ray_module.py:253-264)."""
    saved = load_checkpoint_dict(checkpoint, device)
    model_structure.load_state_dict(strip_module_prefix(saved["model_state_dict"]))
    return saved


class TorchPredictor:
    """Batch-inference callable for ``Dataset.map_batches`` (ref my_ray_module.py:266-284)."""

    def __init__(self, checkpoint: Checkpoint, cpu_only=False):
        self.device = torch.device("cpu") if cpu_only else torch.device("cuda")
        self.model = NeuralNetwork()
        set_weights_from_checkpoint(self.model, checkpoint, self.device)
        self.model.to(self.device)
        if self.device.type == "cuda":
            from ray_torch_distributed_checkpoint_b200.models import fuse_linear_chains
            fuse_linear_chains(self.model)           # inference forward on the fused kernels (K12)
        self.model.eval()

    def __call__(self, batch: Dict[str, np.ndarray]) -> Dict[str, np.ndarray]:
        features = batch["features"]
        if features.ndim == 5 and features.shape[0] == 1:       # (1, B, 1, 28, 28) from some batch formats
            features = features.squeeze(0)
        x = torch.as_tensor(features, dtype=torch.float32).to(self.device)
        with torch.inference_mode():
            logits = self.model(x).cpu().numpy().astype(np.float32)
        return {"logits": logits, "predicted_values": logits.argmax(axis=1)}


if __name__ == "__main__":
    train_fashion_mnist(num_workers=int(os.environ.get("RTDC_NUM_WORKERS", 4)), use_gpu=torch.cuda.is_available())
