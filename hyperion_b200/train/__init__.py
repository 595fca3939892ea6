"""Trainers with the reference's names and signatures (02_development/distributed_utils.py, core_framework.ipynb)."""
from .ddp import train_cifar_model_ddp, train_language_model_ddp
from .engine import Engine, build_engine
from .fsdp import train_gpt2_fsdp, train_language_model_fsdp
from .llama import train_llama_fsdp
from .single import profile_amp_training, train_cifar_model, train_language_model, train_one_batch

__all__ = ["train_cifar_model_ddp", "train_language_model_ddp", "Engine", "build_engine", "train_gpt2_fsdp",
           "train_language_model_fsdp", "train_llama_fsdp", "profile_amp_training", "train_cifar_model",
           "train_language_model", "train_one_batch"]
