"""SM-clock / throttle-reason sampling DURING a timed region (B200_PROFILING.md "clocks line").

Uses NVML through nvidia_ml_py (poll every few ms from a background thread — benchmark regions here can be only
milliseconds long, far shorter than `nvidia-smi -lms 200` can resolve) and falls back to an `nvidia-smi` subprocess.
"""
from __future__ import annotations

import statistics
import subprocess
import threading
import time

_REASONS = {
    0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x10: "sync_boost",
    0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting",
}
BAD = {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}


class ClockSampler:
    def __init__(self, device_index: int = 0, period_s: float = 0.0005):
        self.device_index = device_index
        self.period_s = period_s
        self.samples: list[tuple[float, int, int]] = []  # (t, sm_mhz, reasons bitmask)
        self.sm_max_mhz = 0
        self._stop = threading.Event()
        self._thread: threading.Thread | None = None
        self._nvml = None
        self._handle = None
        self.backend = "none"
        self.paused = False  # while True the thread sleeps without touching NVML (see pause())
        self.first_error: str | None = None  # the first exception a poll raised (round 1 swallowed them all and reported nothing)
        self.polls = 0

    def start(self) -> "ClockSampler":
        try:
            import pynvml

            pynvml.nvmlInit()
            self._nvml = pynvml
            self._handle = pynvml.nvmlDeviceGetHandleByIndex(self.device_index)
            self.sm_max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self._handle, pynvml.NVML_CLOCK_SM))
            self.backend = "nvml"
            self._thread = threading.Thread(target=self._run_nvml, daemon=True)
        except Exception as e:  # noqa: BLE001
            self.first_error = f"nvml init: {type(e).__name__}: {e}"
            self.backend = "nvidia-smi"
            self._thread = threading.Thread(target=self._run_smi, daemon=True)
        self._thread.start()
        return self

    def _run_nvml(self) -> None:
        n = self._nvml
        while not self._stop.is_set():
            if self.paused:
                time.sleep(self.period_s)
                continue
            try:
                mhz = int(n.nvmlDeviceGetClockInfo(self._handle, n.NVML_CLOCK_SM))
                try:
                    reasons = int(n.nvmlDeviceGetCurrentClocksEventReasons(self._handle))
                except Exception:
                    reasons = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self._handle))
                self.samples.append((time.perf_counter(), mhz, reasons))
            except Exception as e:  # noqa: BLE001
                if self.first_error is None:
                    self.first_error = f"nvml poll: {type(e).__name__}: {e}"
            self.polls += 1
            time.sleep(self.period_s)

    def _run_smi(self) -> None:
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.active"
        try:
            p = subprocess.Popen(["nvidia-smi", f"--id={self.device_index}", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                  "-lms", "100"], stdout=subprocess.PIPE, text=True)
        except Exception:
            return
        try:
            while not self._stop.is_set():
                line = p.stdout.readline()
                if not line:
                    break
                parts = [s.strip() for s in line.split(",")]
                try:
                    self.sm_max_mhz = int(parts[1])
                    self.samples.append((time.perf_counter(), int(parts[0]), int(parts[2], 16)))
                except Exception:
                    continue
        finally:
            p.kill()

    def pause(self, on: bool = True) -> None:
        """NVML queries take driver-wide locks; a query in flight can hold up a kernel or graph launch of the same process for
        tens of microseconds. A timed region that is only a few hundred microseconds long is therefore measured with the sampler
        paused, and its clock record comes from the samples right before it and from an untimed repeat of the same work."""
        self.paused = on

    def wait_for_samples(self, n: int = 1, timeout_s: float = 2.0) -> bool:
        """Blocks until at least `n` samples exist (the first NVML query of a process can take milliseconds — longer than a
        whole 20-step timed region)."""
        t_end = time.perf_counter() + timeout_s
        while len(self.samples) < n and time.perf_counter() < t_end:
            time.sleep(0.001)
        return len(self.samples) >= n

    def stop(self) -> None:
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=2)

    def summary(self, t0: float | None = None, t1: float | None = None) -> dict:
        """Median SM clock and the union of event reasons over samples inside [t0, t1] (perf_counter times); when the
        region was too short to catch a sample, the nearest samples around it are used and `window` says so."""
        inside = [s for s in self.samples if (t0 is None or s[0] >= t0) and (t1 is None or s[0] <= t1)]
        window = "timed_region"
        if not inside and self.samples:
            mid = ((t0 or 0) + (t1 or 0)) / 2
            inside = sorted(self.samples, key=lambda s: abs(s[0] - mid))[:3]
            window = "nearest_samples"
        if not inside:
            return {"sm_mhz": None, "sm_max_mhz": self.sm_max_mhz or None, "reasons": [], "samples": 0, "window": "none",
                    "backend": self.backend, "polls": self.polls, "error": self.first_error}
        mask = 0
        for s in inside:
            mask |= s[2]
        reasons = sorted(name for bit, name in _REASONS.items() if mask & bit and name != "gpu_idle")
        return {"sm_mhz": int(statistics.median(s[1] for s in inside)), "sm_max_mhz": self.sm_max_mhz or None, "reasons": reasons,
                "samples": len(inside), "window": window, "backend": self.backend, "sm_min_mhz": int(min(s[1] for s in inside)),
                "rejected": bool(BAD & set(reasons)), "error": self.first_error}
