"""A small columnar ``Dataset`` with the Ray Data calls the reference makes:
``ray.data.from_items(list_of_dicts)`` (``/root/reference/my_ray_module.py:50,72``),
``ds.map_batches(callable, concurrency=, batch_size=, num_gpus=).take_all()`` and ``ds.to_pandas()``
(``/root/reference/eval_flow.py:85-91``), ``DataContext.get_current().enable_tensor_extension_casting``
(``eval_flow.py:78-80``).

Batches handed to the callable are dicts of numpy arrays (``{"features": float32[B,1,28,28],
"labels": int64[B]}``); outputs are concatenated row-wise. With ``concurrency > 1`` and a callable
class, one instance per worker thread is built (each pinned to its own CUDA device when
``num_gpus`` is set) and batches are dealt round-robin — inference data parallelism.
"""
from __future__ import annotations

import concurrent.futures as _fut
import inspect
from typing import Any, Callable, Dict, Iterable, List, Optional

import numpy as np


class DataContext:
    _current: Optional["DataContext"] = None

    def __init__(self):
        self.enable_tensor_extension_casting = True
        self.target_max_block_size = None

    @classmethod
    def get_current(cls) -> "DataContext":
        if cls._current is None:
            cls._current = DataContext()
        return cls._current


def _stack(values: List[Any]) -> np.ndarray:
    first = values[0]
    if isinstance(first, np.ndarray):
        return np.stack(values)
    try:
        import torch
        if isinstance(first, torch.Tensor):
            return np.stack([v.numpy() for v in values])
    except ImportError:
        pass
    return np.asarray(values)


class Dataset:
    def __init__(self, columns: Dict[str, np.ndarray]):
        lens = {len(v) for v in columns.values()}
        if len(lens) > 1:
            raise ValueError(f"columns have different lengths: {lens}")
        self._cols = dict(columns)
        self._n = lens.pop() if lens else 0

    # ------------------------------------------------------------------ basics
    def count(self) -> int:
        return self._n

    def __len__(self) -> int:
        return self._n

    def columns(self) -> List[str]:
        return list(self._cols)

    def schema(self) -> Dict[str, Any]:
        return {k: (v.dtype, v.shape[1:]) for k, v in self._cols.items()}

    def iter_batches(self, batch_size: int = 256) -> Iterable[Dict[str, np.ndarray]]:
        for i in range(0, self._n, batch_size):
            yield {k: v[i:i + batch_size] for k, v in self._cols.items()}

    def take(self, limit: int = 20) -> List[Dict[str, Any]]:
        return [{k: v[i] for k, v in self._cols.items()} for i in range(min(limit, self._n))]

    def take_all(self) -> List[Dict[str, Any]]:
        return self.take(self._n)

    def to_pandas(self):
        import pandas as pd
        data = {}
        for k, v in self._cols.items():
            data[k] = list(v) if v.ndim > 1 else v        # ndarray cells stay plain numpy objects
        return pd.DataFrame(data)

    def select_columns(self, cols: List[str]) -> "Dataset":
        return Dataset({k: self._cols[k] for k in cols})

    def limit(self, n: int) -> "Dataset":
        return Dataset({k: v[:n] for k, v in self._cols.items()})

    # ------------------------------------------------------------------ map_batches
    def map_batches(self, fn: Callable, *, batch_size: Optional[int] = 256, concurrency: Optional[int] = None,
                    num_gpus: Optional[float] = None, fn_constructor_args: tuple = (),
                    fn_constructor_kwargs: Optional[dict] = None, **_ignored) -> "Dataset":
        batch_size = batch_size or max(self._n, 1)
        workers = int(concurrency or 1)
        batches = list(self.iter_batches(batch_size))
        is_class = inspect.isclass(fn)

        def make(worker_idx: int):
            if not is_class:
                return fn
            if num_gpus:
                try:
                    import torch
                    if torch.cuda.is_available():
                        torch.cuda.set_device(worker_idx % torch.cuda.device_count())
                except ImportError:
                    pass
            return fn(*fn_constructor_args, **(fn_constructor_kwargs or {}))

        if workers <= 1 or len(batches) <= 1:
            f = make(0)
            outs = [f(b) for b in batches]
        else:
            def run(worker_idx: int):
                f = make(worker_idx)
                return [(i, f(batches[i])) for i in range(worker_idx, len(batches), workers)]
            with _fut.ThreadPoolExecutor(max_workers=workers) as ex:
                parts = list(ex.map(run, range(workers)))
            outs = [o for _, o in sorted((p for part in parts for p in part), key=lambda t: t[0])]
        if not outs:
            return Dataset({})
        keys = list(outs[0].keys())
        return Dataset({k: np.concatenate([np.asarray(o[k]) for o in outs], axis=0) for k in keys})


def from_items(items: List[Any]) -> Dataset:
    """Dataset from a list of dicts (one row each); non-dict items land in an ``item`` column."""
    if not items:
        return Dataset({})
    if not isinstance(items[0], dict):
        return Dataset({"item": _stack(list(items))})
    keys = list(items[0].keys())
    return Dataset({k: _stack([it[k] for it in items]) for k in keys})


def from_numpy(columns: Dict[str, np.ndarray]) -> Dataset:
    return Dataset({k: np.asarray(v) for k, v in columns.items()})
