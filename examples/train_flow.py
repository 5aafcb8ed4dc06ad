"""Training flow — counterpart of the reference's ``train_flow.py`` on the framework's flow runner.

    python examples/train_flow.py run [--epochs 3 --batch_size 32 --learning_rate 1e-3]
    python examples/train_flow.py run --from-run RayTorchTrain/<run_id>       # warm start
    python examples/train_flow.py argo-workflows create | trigger             # local deployment emulation

Set ``RTDC_FLOW_GPU=0`` to run the workers on CPU/gloo (the reference's flow hard-codes ``use_gpu=True``,
train_flow.py:64; here it follows the hardware unless overridden).
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from ray_torch_distributed_checkpoint_b200.flow import (FlowSpec, Parameter, Run, Task, current, gpu_profile, kubernetes,
                                                         metaflow_ray, pypi, retry, schedule, step)

N_PARALLEL = int(os.environ.get("RTDC_FLOW_N_PARALLEL", 2))
N_GPU_PER_WORKER = 1


def _use_gpu() -> bool:
    if "RTDC_FLOW_GPU" in os.environ:
        return os.environ["RTDC_FLOW_GPU"] not in ("0", "false", "")
    import torch
    return torch.cuda.is_available() and torch.cuda.device_count() >= N_PARALLEL * N_GPU_PER_WORKER


@schedule(cron="*/5 * * * *")
class RayTorchTrain(FlowSpec):
    epochs = Parameter("epochs", default=3)
    global_batch_size = Parameter("batch_size", default=32)
    learning_rate = Parameter("learning_rate", default=1e-3)
    upstream_task_pathspec = Parameter(
        "from-task", default=None,
        help="Task pathspec flow/run_id/step/task_id whose .result artifact holds a checkpoint to start from.")
    upstream_run_pathspec = Parameter(
        "from-run", default=None,
        help="Run pathspec flow/run_id whose .result artifact holds a checkpoint to start from.")

    @step
    def start(self):
        self.next(self.train, num_parallel=N_PARALLEL)

    @retry(times=3)
    @metaflow_ray(all_nodes_started_timeout=60 * 5)
    @pypi(packages={"torch": "2.11.0"})
    @gpu_profile(interval=1)
    @kubernetes(gpu=N_GPU_PER_WORKER, compute_pool="obp-gpu")
    @step
    def train(self):
        from my_module import train_fashion_mnist

        args = dict(num_workers=N_PARALLEL * N_GPU_PER_WORKER, use_gpu=_use_gpu(),
                    checkpoint_storage_path=current.ray_storage_path, epochs=self.epochs,
                    global_batch_size=self.global_batch_size, learning_rate=self.learning_rate)
        if self.upstream_task_pathspec not in (None, "null"):
            args["checkpoint"] = Task(self.upstream_task_pathspec).data.result.checkpoint
        elif self.upstream_run_pathspec not in (None, "null"):
            args["checkpoint"] = Run(self.upstream_run_pathspec).data.result.checkpoint
        else:
            print("Training from newly initialized")
        self.result = train_fashion_mnist(**args)
        self.next(self.join)

    @kubernetes
    @step
    def join(self, inputs):
        # only the head task of the gang ran the body, so only it carries `result`
        for i in inputs:
            if hasattr(i, "result"):
                self.result = i.result
        self.next(self.end)

    @kubernetes
    @step
    def end(self):
        print(self.result)


if __name__ == "__main__":
    RayTorchTrain()
