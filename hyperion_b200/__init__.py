"""hyperion_b200 -- a B200-native (sm_100a) training and benchmarking framework with the capabilities of the MI250X-era
Hyperion project: hardware probe, baseline model benchmarks, AMP + activation checkpointing + memory stats, DDP / FSDP /
LoRA trainers driven by ``run_distributed.py``, scaling reports.  See DESIGN.md and INTEGRATION.md."""
__version__ = "0.1.0"
