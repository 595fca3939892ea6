"""Typed configuration with exactly the keys of the reference's ``Phase 1/default_config.json``
(which no reference code ever loads -- SURVEY.md C20, A-14).  Here it is JSON-loadable and
actually wired: the CLIs take ``--config`` and use these values as defaults."""
from __future__ import annotations

import json
from dataclasses import asdict, dataclass, field
from pathlib import Path


@dataclass
class HardwareConfig:
    gpu_type: str = "B200"
    num_gpus: int = 8
    memory_per_gpu_gb: int = 180


@dataclass
class OptimizationConfig:
    enable_amp: bool = True
    amp_dtype: str = "bf16"          # addition: the reference only had the boolean
    enable_compile: bool = False     # kept for schema parity; the hot path is hand-written kernels, never torch.compile
    compile_mode: str = "reduce-overhead"
    optimize_dataloader: bool = True
    set_omp_threads: bool = True
    enable_channels_last: bool = True
    gradient_accumulation_steps: int = 1
    enable_gradient_checkpointing: bool = False
    distributed_strategy: str = "ddp"


@dataclass
class BenchmarkingConfig:
    batch_sizes: list = field(default_factory=lambda: [1, 2, 4, 8, 16, 32, 64, 128])
    models: list = field(default_factory=lambda: ["resnet50", "vit_small", "transformer"])
    precision_formats: list = field(default_factory=lambda: ["fp32", "fp16", "bf16"])
    num_iterations: int = 50
    warmup_iterations: int = 10


@dataclass
class DistributedConfig:
    backend: str = "nccl"
    init_method: str = "env://"


@dataclass
class HyperionConfig:
    hardware: HardwareConfig = field(default_factory=HardwareConfig)
    optimization: OptimizationConfig = field(default_factory=OptimizationConfig)
    benchmarking: BenchmarkingConfig = field(default_factory=BenchmarkingConfig)
    distributed: DistributedConfig = field(default_factory=DistributedConfig)

    def to_dict(self) -> dict:
        return asdict(self)

    def save(self, path) -> None:
        Path(path).write_text(json.dumps(self.to_dict(), indent=2))

    @staticmethod
    def from_dict(d: dict) -> "HyperionConfig":
        def sub(cls, key):
            known = {f for f in cls.__dataclass_fields__}
            return cls(**{k: v for k, v in d.get(key, {}).items() if k in known})
        return HyperionConfig(sub(HardwareConfig, "hardware"), sub(OptimizationConfig, "optimization"),
                              sub(BenchmarkingConfig, "benchmarking"), sub(DistributedConfig, "distributed"))

    @staticmethod
    def load(path) -> "HyperionConfig":
        return HyperionConfig.from_dict(json.loads(Path(path).read_text()))


DEFAULT_CONFIG_PATH = Path(__file__).resolve().parent / "default_config.json"


def load_default() -> HyperionConfig:
    return HyperionConfig.load(DEFAULT_CONFIG_PATH) if DEFAULT_CONFIG_PATH.exists() else HyperionConfig()
