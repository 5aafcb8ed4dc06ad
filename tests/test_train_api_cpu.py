"""CPU tier: configs, Checkpoint, retention, storage layout, TorchTrainer on 2 gloo workers, retry/resume."""
import os
import pickle
import sys

import pytest
import torch

from ray_torch_distributed_checkpoint_b200 import train
from ray_torch_distributed_checkpoint_b200.train import (Checkpoint, CheckpointConfig, FailureConfig, Result, RunConfig,
                                                          ScalingConfig, TorchTrainer, TrainingFailedError)
from ray_torch_distributed_checkpoint_b200.train.storage import CheckpointManager, StorageContext, checkpoint_dir_name

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples"))


def test_configs_validate():
    assert ScalingConfig(num_workers=2, use_gpu=False).backend == "gloo"
    assert ScalingConfig(num_workers=2, use_gpu=True).backend == "nccl"
    assert ScalingConfig(num_workers=2, use_gpu=True).num_gpus_per_worker == 1
    with pytest.raises(ValueError):
        ScalingConfig(num_workers=0)
    with pytest.raises(ValueError):
        CheckpointConfig(num_to_keep=0)
    rc = RunConfig(storage_path="/tmp/x")
    assert rc.checkpoint_config.num_to_keep is None and rc.failure_config.max_failures == 0


def test_checkpoint_roundtrip(tmp_path):
    d = tmp_path / "ck"
    d.mkdir()
    (d / "latest_model.pt").write_bytes(b"abc")
    c = Checkpoint.from_directory(str(d))
    assert c.path == str(d)
    with c.as_directory() as p:
        assert p == str(d) and os.path.exists(os.path.join(p, "latest_model.pt"))
    c2 = pickle.loads(pickle.dumps(c))
    assert c2 == c and c2.path == c.path
    out = c.to_directory(str(tmp_path / "copy"))
    assert open(os.path.join(out, "latest_model.pt"), "rb").read() == b"abc"
    c.set_metadata({"a": 1})
    c.update_metadata({"b": 2})
    assert c.get_metadata() == {"a": 1, "b": 2}
    assert "Checkpoint(" in repr(c)
    c3 = Checkpoint("file://" + str(d))
    with c3.as_directory() as p:
        assert p == str(d)


def test_retention_keeps_newest(tmp_path):
    mgr = CheckpointManager(CheckpointConfig(num_to_keep=2))
    dirs = []
    for i in range(4):
        d = tmp_path / checkpoint_dir_name(i)
        d.mkdir()
        dirs.append(d)
        mgr.register(Checkpoint.from_directory(str(d)), {"i": i}, i)
    assert [d.exists() for d in dirs] == [False, False, True, True]
    assert mgr.latest.path == str(dirs[3])
    assert len(mgr.best_checkpoints) == 2


def test_retention_by_score_never_drops_latest(tmp_path):
    mgr = CheckpointManager(CheckpointConfig(num_to_keep=2, checkpoint_score_attribute="acc", checkpoint_score_order="max"))
    accs = [0.9, 0.8, 0.1]
    dirs = []
    for i, a in enumerate(accs):
        d = tmp_path / checkpoint_dir_name(i)
        d.mkdir()
        dirs.append(d)
        mgr.register(Checkpoint.from_directory(str(d)), {"acc": a}, i)
    assert dirs[0].exists() and dirs[2].exists() and not dirs[1].exists()


def test_persist_rank0_wins(tmp_path):
    st = StorageContext(str(tmp_path / "store"), "exp", "trial")
    st.make_dirs()
    a, b = tmp_path / "r0", tmp_path / "r1"
    a.mkdir(); b.mkdir()
    (a / "latest_model.pt").write_text("rank0")
    (b / "latest_model.pt").write_text("rank1")
    (b / "only_rank1.bin").write_text("x")
    st.persist_checkpoint_files(str(b), 3, rank=1)
    st.persist_checkpoint_files(str(a), 3, rank=0)
    dst = st.checkpoint_path(3)
    assert dst.endswith("checkpoint_000003")
    assert open(os.path.join(dst, "latest_model.pt")).read() == "rank0"
    assert os.path.exists(os.path.join(dst, "only_rank1.bin"))
    assert not [f for f in os.listdir(dst) if f.endswith(".tmp")]
    with pytest.raises(ValueError):
        StorageContext("nosuchscheme://bucket/x")            # remote URIs go through pyarrow.fs; unknown schemes are refused


def _loop(config):
    import tempfile
    import torch.distributed as dist
    ctx = train.get_context()
    assert ctx.get_world_size() == config["world"]
    assert dist.get_world_size() == config["world"] and dist.get_backend() == "gloo"
    start = 0
    ck = train.get_checkpoint()
    if ck is not None:
        with ck.as_directory() as d:
            start = torch.load(os.path.join(d, "state.pt"), weights_only=True)["epoch"] + 1
    t = torch.ones(1) * (ctx.get_world_rank() + 1)
    dist.all_reduce(t)
    for epoch in range(start, config["epochs"]):
        d = tempfile.mkdtemp()
        torch.save({"epoch": epoch, "sum": t}, os.path.join(d, "state.pt"))
        train.report({"epoch": epoch, "sum": float(t), "rank": ctx.get_world_rank()}, checkpoint=Checkpoint.from_directory(d))


def test_trainer_two_gloo_workers(tmp_path):
    trainer = TorchTrainer(_loop, train_loop_config={"world": 2, "epochs": 4},
                           scaling_config=ScalingConfig(num_workers=2, use_gpu=False),
                           run_config=RunConfig(storage_path=str(tmp_path), checkpoint_config=CheckpointConfig(num_to_keep=2), verbose=0))
    result = trainer.fit()
    assert isinstance(result, Result) and result.error is None
    assert result.metrics["epoch"] == 3 and result.metrics["sum"] == 3.0 and result.metrics["rank"] == 0
    assert result.metrics["training_iteration"] == 4
    kept = sorted(d for d in os.listdir(result.path) if d.startswith("checkpoint_"))
    assert kept == ["checkpoint_000002", "checkpoint_000003"]
    assert result.checkpoint.path.endswith("checkpoint_000003")
    assert len(result.metrics_dataframe) == 4
    for f in ("result.json", "progress.csv", "params.json"):
        assert os.path.exists(os.path.join(result.path, f))
    r2 = pickle.loads(pickle.dumps(result))           # rides inside a flow artifact
    assert r2.checkpoint == result.checkpoint
    assert "Result(" in repr(result)
    rebuilt = Result.from_path(result.path)
    assert rebuilt.checkpoint == result.checkpoint and len(rebuilt.metrics_history) == 4
    assert result.get_best_checkpoint("epoch", "max") == result.checkpoint


def _failing(config):
    raise ValueError("boom from the training function")


def test_worker_exception_surfaces(tmp_path):
    trainer = TorchTrainer(_failing, scaling_config=ScalingConfig(num_workers=2), run_config=RunConfig(storage_path=str(tmp_path), verbose=0))
    with pytest.raises(TrainingFailedError, match="boom from the training function"):
        trainer.fit()


def test_fault_injection_restart_resumes_from_checkpoint(tmp_path, monkeypatch):
    # rank 1 dies right after its 2nd report on attempt 0; the group restarts and get_checkpoint() feeds epoch 2
    monkeypatch.setenv("RTDC_FAULT", "1:2:0")
    trainer = TorchTrainer(_loop, train_loop_config={"world": 2, "epochs": 4},
                           scaling_config=ScalingConfig(num_workers=2, use_gpu=False),
                           run_config=RunConfig(storage_path=str(tmp_path), failure_config=FailureConfig(max_failures=1),
                                                checkpoint_config=CheckpointConfig(num_to_keep=2), verbose=0))
    result = trainer.fit()
    epochs = [r["epoch"] for r in result.metrics_history]
    assert epochs == [0, 1, 2, 3]                      # no epoch repeated, none lost
    assert result.checkpoint.path.endswith("checkpoint_000003")
    monkeypatch.setenv("RTDC_FAULT", "0:1:0")
    t2 = TorchTrainer(_loop, train_loop_config={"world": 2, "epochs": 2}, scaling_config=ScalingConfig(num_workers=2),
                      run_config=RunConfig(storage_path=str(tmp_path / "b"), verbose=0))
    with pytest.raises(TrainingFailedError):
        t2.fit()                                        # max_failures=0: the failure is fatal


def test_session_functions_outside_fit():
    assert train.get_context().get_world_size() == 1
    with pytest.raises(RuntimeError):
        train.report({"a": 1})
    from ray_torch_distributed_checkpoint_b200.train.torch import get_device
    assert get_device().type == "cpu"


def _remote_loop(config):
    import tempfile
    for i in range(3):
        d = tempfile.mkdtemp()
        torch.save({"i": i}, os.path.join(d, "latest_model.pt"))
        train.report({"i": i}, checkpoint=Checkpoint.from_directory(d))


def test_remote_storage_through_pyarrow_filesystem(tmp_path, monkeypatch):
    """A storage_path that is not a local directory (ref: a datastore URI, train_flow.py:65): staged locally, mirrored through
    pyarrow.fs after every report, retention applied on the remote side, Result / Checkpoint handles point at the remote copy."""
    import pyarrow.fs as pafs
    bucket = tmp_path / "bucket"
    bucket.mkdir()
    monkeypatch.setenv("RTDC_STAGING_DIR", str(tmp_path / "staging"))
    fs = pafs.SubTreeFileSystem(str(bucket), pafs.LocalFileSystem())
    trainer = TorchTrainer(_remote_loop, scaling_config=ScalingConfig(num_workers=1, use_gpu=False),
                           run_config=RunConfig(storage_path="runs/exp", storage_filesystem=fs, name="remote",
                                                checkpoint_config=CheckpointConfig(num_to_keep=2), verbose=0))
    result = trainer.fit()
    assert result.path.startswith("runs/exp/remote/") and result.checkpoint.filesystem is not None
    trial = bucket / result.path
    assert sorted(p.name for p in trial.iterdir() if p.name.startswith("checkpoint_")) == ["checkpoint_000001", "checkpoint_000002"]
    assert (trial / "result.json").exists() and (trial / "progress.csv").exists()
    with result.checkpoint.as_directory() as d:                      # downloaded to a temp dir, removed afterwards
        assert not d.startswith(str(bucket)) and torch.load(os.path.join(d, "latest_model.pt"))["i"] == 2
    assert pickle.loads(pickle.dumps(result)).checkpoint.path == result.checkpoint.path
    with pytest.raises(ValueError):
        StorageContext("nosuchscheme://bucket/x")


def test_heartbeat_timeout_fails_a_silent_worker(tmp_path):
    def stuck(config):
        import time
        from ray_torch_distributed_checkpoint_b200.train import session
        session._get_session("x").link._stop.set()                   # a worker whose host side went silent
        time.sleep(30)

    trainer = TorchTrainer(stuck, scaling_config=ScalingConfig(num_workers=1, use_gpu=False),
                           run_config=RunConfig(storage_path=str(tmp_path), verbose=0, worker_report_timeout_s=2.0, worker_heartbeat_s=0.2))
    with pytest.raises(TrainingFailedError, match="no worker activity"):
        trainer.fit()
    assert RunConfig().worker_report_timeout_s == 600.0               # finite by default


def test_fit_in_spmd_mode_under_torchrun_on_gloo(tmp_path):
    """fit() inside ranks that torch.distributed.run started (how bench.py runs at N > 1): reports with checkpoints on the
    report thread's own gloo group, retention, and the Result on every rank."""
    import socket
    import subprocess
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    worker = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_spmd_cpu_worker.py")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), worker, str(tmp_path / "store")]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
    assert "SPMD_OK rank 0" in out.stdout and "SPMD_OK rank 1" in out.stdout
