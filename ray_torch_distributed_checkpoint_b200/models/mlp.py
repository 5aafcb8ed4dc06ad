"""The reference model (``NeuralNetwork``, ``/root/reference/my_ray_module.py:94-112``):
``Flatten -> Linear(784,512) -> ReLU -> Dropout(.25) -> Linear(512,512) -> ReLU -> Dropout(.25) ->
Linear(512,10) -> ReLU`` — note the trailing ReLU on the logits. The ``nn.Sequential`` indices 0, 3, 6
hold the Linears, so ``state_dict`` keys are ``linear_relu_stack.{0,3,6}.{weight,bias}``; checkpoints
are interchangeable with the reference's.
"""
from __future__ import annotations

from torch import nn


class NeuralNetwork(nn.Module):
    def __init__(self, in_features: int = 28 * 28, hidden: int = 512, classes: int = 10, p_drop: float = 0.25):
        super().__init__()
        self.flatten = nn.Flatten()
        self.linear_relu_stack = nn.Sequential(
            nn.Linear(in_features, hidden),
            nn.ReLU(),
            nn.Dropout(p_drop),
            nn.Linear(hidden, hidden),
            nn.ReLU(),
            nn.Dropout(p_drop),
            nn.Linear(hidden, classes),
            nn.ReLU(),
        )

    def forward(self, x):
        return self.linear_relu_stack(self.flatten(x))


def mlp_flops_per_sample(in_features: int = 784, hidden: int = 512, classes: int = 10, training: bool = True) -> int:
    """Multiply-add FLOPs per sample: forward, plus (training) wgrad for all layers and dgrad for all
    but the first (its input needs no gradient)."""
    fwd = 2 * (in_features * hidden + hidden * hidden + hidden * classes)
    if not training:
        return fwd
    wgrad = fwd
    dgrad = 2 * (hidden * hidden + hidden * classes)
    return fwd + wgrad + dgrad
