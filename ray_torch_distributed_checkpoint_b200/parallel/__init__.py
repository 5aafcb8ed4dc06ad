"""Data parallelism (the reference's only strategy): bucket planning, the DDP
wrapper, symmetric memory + NVLink collectives, and the fused MLP train-step engine."""
from . import buckets

__all__ = ["buckets", "DistributedDataParallel", "SymmetricArena", "MLPTrainEngine", "EngineCore", "EngineModel", "EvalAccumulator"]


def __getattr__(name):
    if name == "DistributedDataParallel":
        from .ddp import DistributedDataParallel
        return DistributedDataParallel
    if name == "SymmetricArena":
        from .symm import SymmetricArena
        return SymmetricArena
    if name == "MLPTrainEngine":
        from .engine import MLPTrainEngine
        return MLPTrainEngine
    if name == "EngineCore":
        from .engine import EngineCore
        return EngineCore
    if name in ("EngineModel", "EvalAccumulator", "EngineSGD", "LazyLogits", "LazyLoss"):
        from . import engine_api
        return getattr(engine_api, name)
    raise AttributeError(name)
