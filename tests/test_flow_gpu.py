"""BASELINE.json config #3 on GPUs: ``train_flow.py run`` with one worker per GPU (1 GPU and, when the box has them, up to
8) -> per-epoch checkpoint with ``num_to_keep=2`` -> ``run --from-run`` warm start -> ``eval_flow.py run --from-run`` on one GPU.
The flow's train step goes through ``TorchTrainer.fit()`` in driver mode (native launcher, one rank per GPU) and the
engine-backed model, so this is the reference's flow (``/root/reference/train_flow.py:57-77``, ``eval_flow.py:82-94``) end to end."""
import os
import subprocess
import sys
import zipfile

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EX = os.path.join(ROOT, "examples")


def _run(cmd, env, log):
    p = subprocess.run([sys.executable] + cmd, env=env, capture_output=True, text=True, timeout=900)
    log.write(f"$ {' '.join(cmd)}\n{p.stdout}\n{p.stderr}\n")
    log.flush()
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    return p.stdout + p.stderr


def _worker_counts():
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    return [1] + ([8 if n >= 8 else (4 if n >= 4 else 2)] if n >= 2 else [])


@pytest.mark.parametrize("slot", [0, 1])
def test_reference_config_3_train_resume_eval(tmp_path, slot):
    counts = _worker_counts()
    if slot >= len(counts):
        pytest.skip("needs at least 2 GPUs")
    n = counts[slot]
    from ray_torch_distributed_checkpoint_b200 import flow
    datastore = str(tmp_path / "ds")
    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    env = dict(os.environ, RTDC_FLOW_DATASTORE=datastore, RTDC_TRAIN_SIZE="16384", RTDC_TEST_SIZE="4096", RTDC_SYNTHETIC_DATA="1",
               RTDC_FLOW_N_PARALLEL=str(n), RTDC_FLOW_GPU="1")
    old = os.environ.get("RTDC_FLOW_DATASTORE")
    os.environ["RTDC_FLOW_DATASTORE"] = datastore
    try:
        with open(os.path.join(out_dir, f"config3_n{n}.log"), "w") as log:
            out = _run([os.path.join(EX, "train_flow.py"), "run", "--epochs", "3", "--batch_size", str(256 * n), "--learning_rate", "0.05"], env, log)
            assert "Training from newly initialized" in out and "backend=nccl" in out
            run = flow.Flow("RayTorchTrain").latest_successful_run
            result = run.data.result
            assert result.metrics["training_iteration"] == 3 and result.error is None
            kept = sorted(d for d in os.listdir(result.path) if d.startswith("checkpoint_"))
            assert kept == ["checkpoint_000001", "checkpoint_000002"]             # CheckpointConfig(num_to_keep=2)
            with result.checkpoint.as_directory() as d:
                path = os.path.join(d, "latest_model.pt")
                sd = torch.load(path, weights_only=True)                          # the reference's loader call (:255-259)
                # byte layout: same archive entries and sizes as a stock torch.save of the same object
                ref_path = str(tmp_path / f"ref{n}" / "latest_model.pt")
                os.makedirs(os.path.dirname(ref_path))
                torch.save(sd, ref_path)
                za, zb = zipfile.ZipFile(path), zipfile.ZipFile(ref_path)
                entries = lambda z: sorted((i.filename, i.file_size) for i in z.infolist() if not i.filename.endswith("serialization_id"))
                assert entries(za) == entries(zb)
            prefix = "module." if n > 1 else ""
            assert list(sd["model_state_dict"]) == [f"{prefix}linear_relu_stack.{i}.{k}" for i in (0, 3, 6) for k in ("weight", "bias")]
            assert sd["epoch"] == 2 and len(sd["val_losses"]) == 3 and sd["val_losses"][-1] < sd["val_losses"][0]
            assert sd["optimizer_state_dict"]["state"][0]["momentum_buffer"].shape == (512, 784)
            assert float(sd["optimizer_state_dict"]["state"][0]["momentum_buffer"].abs().sum()) > 0      # every slice gathered

            out = _run([os.path.join(EX, "train_flow.py"), "run", "--epochs", "1", "--batch_size", str(256 * n), "--learning_rate", "0.05",
                        "--from-run", run.pathspec], env, log)
            assert "Resuming from checkpoint at" in out
            resumed = flow.Flow("RayTorchTrain").latest_successful_run.data.result
            assert resumed.metrics["val_loss"] < sd["val_losses"][0]                # started from the trained weights

            # evaluation restores the n-rank checkpoint on ONE GPU ('module.' prefix stripped, my_ray_module.py:261)
            env1 = dict(env, CUDA_VISIBLE_DEVICES="0")
            _run([os.path.join(EX, "eval_flow.py"), "run", "--from-run", run.pathspec], env1, log)
            ev = flow.Flow("RayTorchEval").latest_successful_run
            assert ev.data.predictions.shape[0] == 4096
            log.write(f"[config3] n={n} trained accuracy (rank 0 shard) {result.metrics['accuracy']:.4f}, eval flow accuracy {ev.data.accuracy:.4f}\n")
            assert abs(ev.data.accuracy - result.metrics["accuracy"]) < 0.03
    finally:
        if old is None:
            os.environ.pop("RTDC_FLOW_DATASTORE", None)
        else:
            os.environ["RTDC_FLOW_DATASTORE"] = old
