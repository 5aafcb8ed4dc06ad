"""ray_torch_distributed_checkpoint_b200 — a B200-native data-parallel training + checkpointing
framework with the public surface of the Metaflow + Ray Train example
``outerbounds/ray-torch-distributed-checkpoint`` (see SURVEY.md / DESIGN.md).
"""
__version__ = "0.1.0"
