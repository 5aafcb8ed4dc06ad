"""Data layer: synthetic datasets of the reference's shapes, sharded samplers, device prefetch, and
a mini ``Dataset`` with ``from_items / map_batches / take_all / to_pandas`` (the Ray Data surface the
eval flow uses, ``/root/reference/eval_flow.py:83-91``)."""
from .datasets import (FashionMNISTSynthetic, ImageNetSynthetic, TokenSynthetic, get_labels_map, write_fashion_mnist_idx)
from .loader import DeviceLoader, DistributedSampler, U8BatchLoader, U8Images, shard_data_loader
from .dataset import DataContext, Dataset, from_items, from_numpy

__all__ = ["FashionMNISTSynthetic", "ImageNetSynthetic", "TokenSynthetic", "get_labels_map", "write_fashion_mnist_idx",
           "DeviceLoader", "DistributedSampler", "U8BatchLoader", "U8Images", "shard_data_loader",
           "DataContext", "Dataset", "from_items", "from_numpy"]
