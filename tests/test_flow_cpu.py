"""CPU tier: the flow runner (datastore, parameters, gang step, retry, cards, trigger chain) and the
reference-shaped example flows end to end on 2 gloo workers: train -> checkpoint -> --from-run -> eval."""
import os
import subprocess
import sys

import pytest

from ray_torch_distributed_checkpoint_b200 import flow
from ray_torch_distributed_checkpoint_b200.flow import FlowSpec, Parameter, Run, Task, current, retry, step
from ray_torch_distributed_checkpoint_b200.flow.cards import Image, Markdown, Table
from ray_torch_distributed_checkpoint_b200.flow.core import run_flow

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EX = os.path.join(ROOT, "examples")


@pytest.fixture()
def datastore(tmp_path, monkeypatch):
    monkeypatch.setenv("RTDC_FLOW_DATASTORE", str(tmp_path / "ds"))
    return str(tmp_path / "ds")


class _Toy(FlowSpec):
    n = Parameter("n", default=3)
    tag = Parameter("from-run", default=None)
    attempts = []

    @step
    def start(self):
        self.base = self.n * 2
        self.next(self.work, num_parallel=2)

    @retry(times=2)
    @flow.metaflow_ray
    @flow.card(type="blank", id="c1")
    @step
    def work(self):
        _Toy.attempts.append(current.retry_count)
        if current.retry_count == 0:
            raise RuntimeError("first attempt fails")
        self.storage = current.ray_storage_path
        self.result = {"v": self.base + 1, "node": current.parallel.node_index}
        current.card["c1"].append(Markdown("# hello"))
        current.card["c1"].append(Table(headers=["img"], data=[[Image.from_array(__import__("numpy").zeros((4, 4)))]]))
        self.next(self.join)

    @step
    def join(self, inputs):
        got = [i.result for i in inputs if hasattr(i, "result")]
        assert len(got) == 1                       # only the head of the gang ran the body
        self.result = got[0]
        self.next(self.end)

    @step
    def end(self):
        self.done = True


def test_flow_runner_semantics(datastore):
    _Toy.attempts.clear()
    run = run_flow(_Toy, n=5)
    assert run.successful and run.data.result == {"v": 11, "node": 0} and run.data.done is True
    assert run.data.n == 5                                        # parameters are artifacts too
    assert _Toy.attempts == [0, 1]                                # @retry re-ran the step
    tasks = run["work"].tasks()
    assert len(tasks) == 2 and "result" in tasks[0].data and "result" not in tasks[1].data
    t = Task(tasks[0].pathspec)
    assert t.data.storage.startswith(datastore) and "attempt_1" in t.data.storage
    card = os.path.join(datastore, "_Toy", run.id, "work", tasks[0].id, "cards", "c1.html")
    assert "<h1>hello</h1>" in open(card).read() and "data:image/png;base64" in open(card).read()
    assert Run(run.pathspec).data.result["v"] == 11
    assert flow.Flow("_Toy").latest_successful_run.id == run.id
    with pytest.raises(flow.FlowError):
        Run("_Toy/does_not_exist")
    with pytest.raises(AttributeError):
        current.ray_storage_path


def _run(cmd, env):
    p = subprocess.run([sys.executable] + cmd, env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    return p.stdout + p.stderr


def test_reference_config_1_train_resume_eval_on_gloo(datastore):
    """BASELINE.json config #1: MLP, TorchTrainer num_workers=2 use_gpu=False, save + resume + eval."""
    import torch
    env = dict(os.environ, RTDC_FLOW_DATASTORE=datastore, RTDC_TRAIN_SIZE="1024", RTDC_TEST_SIZE="256", RTDC_FLOW_GPU="0")
    out = _run([os.path.join(EX, "train_flow.py"), "--environment=fast-bakery", "run", "--epochs", "3", "--batch_size", "64"], env)
    assert "Training from newly initialized" in out and "backend=gloo" in out
    run = flow.Flow("RayTorchTrain").latest_successful_run
    result = run.data.result
    assert result.metrics["training_iteration"] == 3 and result.error is None
    kept = sorted(d for d in os.listdir(result.path) if d.startswith("checkpoint_"))
    assert kept == ["checkpoint_000001", "checkpoint_000002"]             # CheckpointConfig(num_to_keep=2)
    with result.checkpoint.as_directory() as d:
        sd = torch.load(os.path.join(d, "latest_model.pt"), weights_only=True)
    assert set(sd) == {"epoch", "model_state_dict", "optimizer_state_dict", "val_losses", "val_accuracy"}
    assert sd["epoch"] == 2 and len(sd["val_losses"]) == 3
    assert list(sd["model_state_dict"]) == [f"module.linear_relu_stack.{i}.{k}" for i in (0, 3, 6) for k in ("weight", "bias")]
    pg = sd["optimizer_state_dict"]["param_groups"][0]
    assert pg["momentum"] == 0.9 and pg["params"] == [0, 1, 2, 3, 4, 5]
    assert sd["optimizer_state_dict"]["state"][0]["momentum_buffer"].shape == (512, 784)

    out = _run([os.path.join(EX, "train_flow.py"), "run", "--epochs", "1", "--batch_size", "64", "--from-run", run.pathspec], env)
    assert "Resuming from checkpoint at" in out
    out = _run([os.path.join(EX, "train_flow.py"), "run", "--epochs", "1", "--from-run", "null"], env)
    assert "Training from newly initialized" in out                      # the "null" sentinel means unset

    out = _run([os.path.join(EX, "eval_flow.py"), "run", "--from-run", run.pathspec], env)
    ev = flow.Flow("RayTorchEval").latest_successful_run
    assert ev.data.predictions.shape[0] == 256
    assert set(ev.data.predictions.columns) == {"features", "labels", "logits", "predicted_values"}
    assert 0.0 <= ev.data.accuracy <= 1.0
    p = subprocess.run([sys.executable, os.path.join(EX, "eval_flow.py"), "run"], env=env, capture_output=True, text=True)
    assert p.returncode != 0 and "you must specify an upstream run or task id" in p.stderr


def test_trigger_on_finish_chain(datastore):
    env = dict(os.environ, RTDC_FLOW_DATASTORE=datastore, RTDC_TRAIN_SIZE="512", RTDC_TEST_SIZE="128", RTDC_FLOW_GPU="0")
    _run([os.path.join(EX, "train_flow.py"), "argo-workflows", "create"], env)
    _run([os.path.join(EX, "eval_flow.py"), "argo-workflows", "create"], env)
    out = _run([os.path.join(EX, "train_flow.py"), "argo-workflows", "trigger", "--epochs", "1", "--batch_size", "64"], env)
    assert "triggering RayTorchEval" in out
    ev = flow.Flow("RayTorchEval").latest_successful_run
    assert ev is not None and ev._meta()["triggered_by"].startswith("RayTorchTrain/")


def test_cron_schedule_and_serve_loop(datastore):
    import datetime as dt
    from ray_torch_distributed_checkpoint_b200.flow import CronSchedule
    every5 = CronSchedule("*/5 * * * *")                             # the reference's schedule (train_flow.py:20)
    assert every5.matches(dt.datetime(2026, 1, 1, 10, 35)) and not every5.matches(dt.datetime(2026, 1, 1, 10, 36))
    assert every5.next_after(dt.datetime(2026, 1, 1, 10, 36, 12)) == dt.datetime(2026, 1, 1, 10, 40)
    wk = CronSchedule("30 2 * * 1-5")
    assert wk.matches(dt.datetime(2026, 11, 6, 2, 30)) and not wk.matches(dt.datetime(2026, 11, 7, 2, 30))    # Fri yes, Sat no
    assert CronSchedule("0 0 1,15 * 0").matches(dt.datetime(2026, 11, 8, 0, 0))                                   # Sunday OR day-of-month
    with pytest.raises(ValueError):
        CronSchedule("61 * * * *")
    # the serve loop fires the scheduled flow at a matching minute and chains the deployed eval flow after it
    env = dict(os.environ, RTDC_FLOW_DATASTORE=datastore, RTDC_TRAIN_SIZE="256", RTDC_TEST_SIZE="64", RTDC_FLOW_GPU="0")
    _run([os.path.join(EX, "eval_flow.py"), "argo-workflows", "create"], env)
    out = _run([os.path.join(EX, "train_flow.py"), "argo-workflows", "serve", "--max-runs", "1", "--poll-seconds", "0.05",
                "--now", "2026-01-01T10:35:00", "--epochs", "1", "--batch_size", "64"], env)
    assert "cron slot 2026-01-01 10:35 -> running RayTorchTrain" in out and "triggering RayTorchEval" in out
    assert flow.Flow("RayTorchTrain").latest_successful_run is not None
