from .datasets import CIFAR10TorchDataset, DeviceResidentLoader, WikiText2TorchDataset
from . import synth

__all__ = ["CIFAR10TorchDataset", "DeviceResidentLoader", "WikiText2TorchDataset", "synth"]
