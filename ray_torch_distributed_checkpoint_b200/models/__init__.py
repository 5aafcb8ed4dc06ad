"""Model families: the reference's FashionMNIST MLP plus the north-star configurations
(ResNet-50, GPT-2 small). ``fused`` holds the modules that route Linear(+ReLU+Dropout) chains
through the tcgen05 kernels."""
from .mlp import NeuralNetwork, mlp_flops_per_sample
from .fused import fuse_linear_chains

__all__ = ["NeuralNetwork", "mlp_flops_per_sample", "fuse_linear_chains", "resnet50", "GPT2", "GPT2Config"]


def __getattr__(name):
    if name == "resnet50":
        from .resnet import resnet50
        return resnet50
    if name in ("GPT2", "GPT2Config"):
        from . import gpt2
        return getattr(gpt2, name)
    raise AttributeError(name)
