"""Device timing helpers: CUDA events after warm-up, max over ranks, L2 flush between iterations,
and an nvidia-smi clock sampler (B200_PROFILING.md "Timing hygiene")."""
from __future__ import annotations

import statistics
import subprocess
import threading
import time

import torch

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


class L2Flusher:
    """Writes a buffer larger than the 126 MB L2 so the next timed iteration starts cold."""

    def __init__(self, device, nbytes: int = 256 << 20):
        self.buf = torch.empty(nbytes, device=device, dtype=torch.uint8) if torch.device(device).type == "cuda" else None

    def __call__(self):
        if self.buf is not None:
            self.buf.zero_()


class CudaTimer:
    def __init__(self):
        self.e0 = torch.cuda.Event(enable_timing=True)
        self.e1 = torch.cuda.Event(enable_timing=True)

    def __enter__(self):
        self.e0.record()
        return self

    def __exit__(self, *a):
        self.e1.record()

    def ms(self) -> float:
        self.e1.synchronize()
        return self.e0.elapsed_time(self.e1)


def time_fn(fn, iters: int = 20, warmup: int = 5, flush_l2: bool = True):
    """(median ms, best ms) of ``fn`` on the current stream."""
    if not torch.cuda.is_available():
        for _ in range(warmup):
            fn()
        ts = []
        for _ in range(iters):
            t0 = time.perf_counter(); fn(); ts.append((time.perf_counter() - t0) * 1e3)
        return statistics.median(ts), min(ts)
    flusher = L2Flusher("cuda") if flush_l2 else None
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flusher:
            flusher()
        with CudaTimer() as t:
            fn()
        ts.append(t.ms())
    return statistics.median(ts), min(ts)


class ClockSampler:
    """Samples SM clock, power and throttle reasons in a background thread during a timed region.

    NVML (``pynvml``) is polled every few milliseconds, so even a 30 ms region (20 steps of the 1.4 ms reference LM) gets
    samples *during* the region; when NVML is not importable it falls back to ``nvidia-smi -lms``."""

    _REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
                0x80: "hw_power_brake_slowdown"}

    def __init__(self, gpu_index: int = 0, period_ms: int = 200):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self.proc = None
        self.lines: list[str] = []
        self._thr = None
        self._stop = threading.Event()
        self._nvml = None
        self._samples: list[tuple[float, float, int]] = []   # (sm MHz, W, reasons bitmask)
        self._max_mhz = None

    def _nvml_handle(self):
        import pynvml

        pynvml.nvmlInit()
        try:
            uuid = torch.cuda.get_device_properties(self.gpu_index).uuid
            return pynvml, pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)).encode())
        except Exception:  # noqa: BLE001 -- older torch / masked devices: index order
            return pynvml, pynvml.nvmlDeviceGetHandleByIndex(self.gpu_index)

    def start(self):
        try:
            nv, h = self._nvml_handle()
            self._max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            self._nvml = (nv, h)

            def poll():
                while not self._stop.is_set():
                    try:
                        mhz = float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                        watts = nv.nvmlDeviceGetPowerUsage(h) / 1e3
                        try:
                            reasons = int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
                        except AttributeError:
                            reasons = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                        self._samples.append((mhz, watts, reasons))
                    except Exception:  # noqa: BLE001
                        pass
                    self._stop.wait(0.003)
            self._thr = threading.Thread(target=poll, daemon=True)
            self._thr.start()
            return self
        except Exception:  # noqa: BLE001 -- no NVML: nvidia-smi below
            self._nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={_QUERY}",
                                          "--format=csv,noheader,nounits", "-lms", str(self.period_ms)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except (FileNotFoundError, OSError):
            self.proc = None
            return self

        def pump():
            for line in self.proc.stdout:
                self.lines.append(line.strip())
        self._thr = threading.Thread(target=pump, daemon=True)
        self._thr.start()
        return self

    def stop(self) -> dict:
        if self._nvml is not None:
            self._stop.set()
            self._thr.join(timeout=1.0)
            sm = [s[0] for s in self._samples]
            reasons = set()
            for _, _, bits in self._samples:
                for bit, name in self._REASONS.items():
                    if bits & bit:
                        reasons.add(name)
            return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": self._max_mhz,
                    "power_w_max": max((s[1] for s in self._samples), default=None), "samples": len(sm),
                    "reasons": sorted(reasons), "source": "nvml"}
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1])); smax.append(float(parts[2])); power.append(float(parts[3]))
            except ValueError:
                continue
            for nm, val in zip(names, parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons),
                "source": "nvidia-smi"}
