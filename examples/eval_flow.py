"""Evaluation flow — counterpart of the reference's ``eval_flow.py``: restore the checkpoint of a
training run, batch-predict the 10 000-image validation split, store predictions and
misclassifications as artifacts and render an error-analysis card.

    python examples/eval_flow.py run --from-run RayTorchTrain/<run_id>
    python examples/eval_flow.py argo-workflows create      # then `train_flow.py argo-workflows trigger` chains it
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from ray_torch_distributed_checkpoint_b200.flow import (FlowSpec, Parameter, Run, Task, card, current, gpu_profile,
                                                         kubernetes, pypi, step, trigger_on_finish)
from ray_torch_distributed_checkpoint_b200.flow.cards import Image, Markdown, Table

N_GPU = 1


@trigger_on_finish(flow="RayTorchTrain")
class RayTorchEval(FlowSpec):
    upstream_task_pathspec = Parameter("from-task", default=None, help="Task pathspec holding a .result with a checkpoint.")
    upstream_run_pathspec = Parameter("from-run", default=None, help="Run pathspec holding a .result with a checkpoint.")
    upstream_namespace = Parameter("from-namespace", default=None, help="Accepted for parity; the local datastore has one namespace.")
    batch_size = Parameter("batch_size", default=512)
    n_error_samples = 50

    def _get_checkpoint(self):
        try:
            return current.trigger.run.data.result.checkpoint       # started by RayTorchTrain finishing
        except AttributeError:
            pass
        if self.upstream_task_pathspec not in (None, "null"):
            return Task(self.upstream_task_pathspec).data.result.checkpoint
        if self.upstream_run_pathspec not in (None, "null"):
            return Run(self.upstream_run_pathspec).data.result.checkpoint
        raise ValueError("If this run is not being triggered by RayTorchTrain, you must specify an upstream run or task id.")

    @card(type="blank", id="error_analysis")
    @gpu_profile(interval=1)
    @kubernetes(gpu=N_GPU, compute_pool="obp-gpu")
    @pypi(packages={"torch": "2.11.0", "pandas": "2.2.3"})
    @step
    def start(self):
        import numpy as np
        import pandas as pd
        import torch
        from my_module import TorchPredictor, get_dataloaders, get_labels_map
        from ray_torch_distributed_checkpoint_b200.data import DataContext

        DataContext.get_current().enable_tensor_extension_casting = False
        self.upstream_checkpoint = self._get_checkpoint()
        ds = get_dataloaders(batch_size=self.batch_size, val_only=True, as_ray_ds=True)
        predictor = TorchPredictor(checkpoint=self.upstream_checkpoint, cpu_only=not torch.cuda.is_available())
        result = ds.map_batches(predictor, concurrency=N_GPU, batch_size=self.batch_size, num_gpus=N_GPU).take_all()
        self.predictions = pd.concat([ds.to_pandas(), pd.DataFrame(result)], axis=1)
        wrong = self.predictions.labels != self.predictions.predicted_values
        self.misclassifications = self.predictions[wrong]
        self.accuracy = 1.0 - float(wrong.mean())

        labels_map = get_labels_map()
        n = min(self.n_error_samples, len(self.misclassifications))
        sample = self.misclassifications.sample(n, random_state=0) if n else self.misclassifications
        current.card["error_analysis"].append(
            Markdown(f"### Misclassifications {self.misclassifications.shape[0]} out of {self.predictions.shape[0]}"))
        rows = []
        for _, row in sample.iterrows():
            pixels = ((np.asarray(row.features).reshape(28, 28) * 0.5 + 0.5) * 255).clip(0, 255).astype(np.uint8)
            rows.append([Image.from_array(pixels), labels_map[int(row.labels)], labels_map[int(row.predicted_values)],
                         Image.bar_chart(list(labels_map.values()), row.logits, title="Logits")])
        current.card["error_analysis"].append(Table(headers=["Image", "True label", "Predicted label", "Logits"], data=rows))
        self.next(self.end)

    @step
    def end(self):
        print(f"accuracy {self.accuracy:.4f}; {len(self.misclassifications)} misclassified of {len(self.predictions)}")


if __name__ == "__main__":
    RayTorchEval()
