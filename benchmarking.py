"""Drop-in module for code written against the reference's ``Phase 1/benchmarking.py``: its four public functions
(benchmarking.py:12, 57, 153, 223) resolve to ``hyperion_b200.bench.utils``, which accepts the reference's call forms
(``benchmark_forward_pass(model, input_tensor, num_iterations=..., warmup=...)`` etc.) and returns the same keys
(``total_time, avg_time, throughput, memory_usage``), timed with CUDA events after a warm-up."""
from hyperion_b200.bench.utils import (benchmark_forward_pass, benchmark_training_step, compare_precision_formats,
                                       save_benchmark_results)

__all__ = ["benchmark_forward_pass", "benchmark_training_step", "compare_precision_formats", "save_benchmark_results"]
