"""Drop-in module for code written against the reference's ``02_development/distributed_utils.py``.

``import distributed_utils as du`` keeps working after the switch: the twelve public names of the reference module
(distributed_utils.py:43, 58, 75, 96, 101, 122, 132, 208, 290, 415, 563, 780) resolve to this framework's implementations, with the
reference's positional arguments.  Nothing is implemented here; see ``hyperion_b200/{data,models,parallel,train,utils}``.

    du.setup(rank, world)                                   -> hyperion_b200.parallel.runtime.setup (returns the device as well)
    du.train_language_model_ddp(rank, world, epochs, dir)   -> hyperion_b200.train.train_language_model_ddp
    du.train_llama_fsdp(rank, world, epochs=1, lora=True)   -> hyperion_b200.train.train_llama_fsdp
    du.create_scaling_report(dist_dir)                      -> hyperion_b200.utils.report.create_scaling_report
    du.run_scaling_experiment("cifar", [1, 2, 4], 5, dir)   -> hyperion_b200.utils.scaling.run_scaling_experiment

Differences a caller can see: the default ``base_dir`` is the current directory instead of ``/home/aac/project-hyperion``; the trainers
read ``data/processed`` when it exists and ``synthetic=False`` is passed, and otherwise train on seeded data of the same shape;
``hf_token`` is accepted and unused (nothing is downloaded); ``create_scaling_report`` never invents sample curves.
"""
from __future__ import annotations

from hyperion_b200.data import CIFAR10TorchDataset, WikiText2TorchDataset
from hyperion_b200.models import SimpleTransformerLM
from hyperion_b200.parallel.runtime import _local_gpu, cleanup, setup
from hyperion_b200.train import (train_cifar_model_ddp, train_language_model_ddp, train_language_model_fsdp,
                                 train_llama_fsdp)
from hyperion_b200.utils.report import create_scaling_report
from hyperion_b200.utils.scaling import run_scaling_experiment as _run_scaling_experiment

__all__ = ["WikiText2TorchDataset", "CIFAR10TorchDataset", "SimpleTransformerLM", "_local_gpu", "setup", "cleanup",
           "train_language_model_ddp", "train_cifar_model_ddp", "train_language_model_fsdp", "train_llama_fsdp",
           "create_scaling_report", "run_scaling_experiment"]


def run_scaling_experiment(model_type: str, gpu_counts: list, epochs: int = 5, base_dir: str = ".", hf_token: str | None = None, **kw):
    """The reference's argument order (distributed_utils.py:780-786); ``hf_token`` is passed on to the launcher's flag of that name."""
    extra = list(kw.pop("extra_args", ()))
    if hf_token:
        extra.append(f"--hf_token={hf_token}")
    return _run_scaling_experiment(model_type, list(gpu_counts), epochs, base_dir, extra, **kw)
