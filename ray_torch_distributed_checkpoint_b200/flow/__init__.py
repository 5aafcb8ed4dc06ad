"""A Metaflow-shaped flow runner for one box.

Covers what the reference's two flows use from Metaflow (``/root/reference/train_flow.py:1-15``,
``/root/reference/eval_flow.py:1-15``): ``FlowSpec``, ``@step``, ``self.next(..., num_parallel=N)``,
``Parameter`` (-> CLI flags), ``current`` (incl. ``current.ray_storage_path``, ``current.trigger``,
``current.card``), ``Run`` / ``Task`` / ``Flow`` over a local datastore with pickled artifacts, and the
decorators ``retry``, ``metaflow_ray``, ``kubernetes``, ``pypi``, ``gpu_profile``, ``card``, ``schedule``,
``trigger_on_finish``. CLI: ``python flow.py run|evaluate [--flag v]``, ``argo-workflows create|trigger|serve`` (serve
honours the ``@schedule`` cron expression on this box),
``show``; ``--environment=...`` is accepted and ignored.
"""
from .core import (FlowSpec, Parameter, Run, Task, Flow, Step, current, step, retry, metaflow_ray, kubernetes, pypi,
                   gpu_profile, card, schedule, trigger_on_finish, namespace, datastore_root, FlowError, CronSchedule)
from . import cards

__all__ = ["FlowSpec", "Parameter", "Run", "Task", "Flow", "Step", "current", "step", "retry", "metaflow_ray",
           "kubernetes", "pypi", "gpu_profile", "card", "schedule", "trigger_on_finish", "namespace",
           "datastore_root", "FlowError", "CronSchedule", "cards"]
