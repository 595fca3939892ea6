from .datasets import CIFAR10TorchDataset, DeviceResidentLoader, WikiText2TorchDataset
from .prefetch import DevicePrefetcher
from . import prepare, synth

__all__ = ["CIFAR10TorchDataset", "DeviceResidentLoader", "DevicePrefetcher", "WikiText2TorchDataset", "prepare", "synth"]
