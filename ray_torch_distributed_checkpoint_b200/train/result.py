"""``Result`` of ``TorchTrainer.fit()``: same fields the reference reads from Ray's Result
(``.checkpoint`` at ``/root/reference/train_flow.py:70,73`` and ``eval_flow.py:42-49``; printed at
``train_flow.py:95``)."""
from __future__ import annotations

import dataclasses
import json
import os
from typing import Any, Dict, List, Optional, Tuple

from .checkpoint import Checkpoint


@dataclasses.dataclass
class Result:
    metrics: Optional[Dict[str, Any]]
    checkpoint: Optional[Checkpoint]
    error: Optional[Exception] = None
    path: Optional[str] = None
    metrics_history: Optional[List[Dict[str, Any]]] = None
    best_checkpoints: Optional[List[Tuple[Checkpoint, Dict[str, Any]]]] = None
    filesystem: Optional[Any] = None

    @property
    def config(self) -> Optional[Dict[str, Any]]:
        return (self.metrics or {}).get("config")

    @property
    def metrics_dataframe(self):
        """All reported metric rows (rank 0's, one per ``report`` call) as a pandas DataFrame."""
        import pandas as pd
        return pd.DataFrame(self.metrics_history or [])

    def get_best_checkpoint(self, metric: str, mode: str = "max") -> Optional[Checkpoint]:
        if mode not in ("max", "min"):
            raise ValueError("mode must be 'max' or 'min'")
        cands = [(c, m) for c, m in (self.best_checkpoints or []) if metric in m]
        if not cands:
            return None
        pick = max if mode == "max" else min
        return pick(cands, key=lambda cm: cm[1][metric])[0]

    @classmethod
    def from_path(cls, path: str) -> "Result":
        """Rebuild a Result from a trial directory written by ``fit()`` (result.json + checkpoint_*/)."""
        path = os.path.abspath(path)
        rows = []
        rj = os.path.join(path, "result.json")
        if os.path.exists(rj):
            with open(rj) as f:
                rows = [json.loads(line) for line in f if line.strip()]
        ckpts = sorted(d for d in os.listdir(path) if d.startswith("checkpoint_") and os.path.isdir(os.path.join(path, d)))
        by_name = {r.get("checkpoint_dir_name"): r for r in rows if r.get("checkpoint_dir_name")}
        best = [(Checkpoint.from_directory(os.path.join(path, d)), by_name.get(d, {})) for d in ckpts]
        return cls(metrics=rows[-1] if rows else None, checkpoint=best[-1][0] if best else None, path=path,
                   metrics_history=rows, best_checkpoints=best)

    def __repr__(self) -> str:
        shown = {k: v for k, v in (self.metrics or {}).items() if not isinstance(v, (dict, list))}
        parts = [f"  metrics={shown}", f"  path={self.path!r}", f"  checkpoint={self.checkpoint}"]
        if self.error is not None:
            parts.append(f"  error={type(self.error).__name__}({self.error})")
        return "Result(\n" + ",\n".join(parts) + "\n)"
