"""NVML clock sampling for benchmark lines.

The reference times nothing (its only timing helper, ``cpuSecond`` at CUDACG.cu:35-39, is never called).
Every number this repository reports carries the SM clock seen *during* the timed region and the throttle
reasons that were active, so a reader can reject a run that was throttled or clock-locked.

NVML is initialised and the device handle resolved in ``start()`` (or ``__enter__``), i.e. BEFORE the timed
window opens; the sampling thread then only reads two counters per period. ``sample_now()`` takes one sample
synchronously so that even a window shorter than the period has a record.
"""
from __future__ import annotations

import statistics
import threading
import time


class ClockSampler:
    _NAMES = None

    def __init__(self, index: int, period_s: float = 0.002):
        self.index, self.period = index, period_s
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._stop = threading.Event()
        self._t = None
        self._nv = None
        self._h = None

    # ---- life cycle -------------------------------------------------------------------------
    def start(self):
        """Initialise NVML and resolve the handle. Call well before the timed region."""
        if self._h is not None:
            return self
        try:
            import pynvml as nv
            nv.nvmlInit()
            self._nv = nv
            self._h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(self._h, nv.NVML_CLOCK_SM)
            self._names = {
                nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
                nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake",
            }
        except Exception as e:  # NVML missing: report that instead of inventing clocks
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")
        return self

    def sample_now(self):
        if self._h is None:
            return
        nv = self._nv
        try:
            self.samples.append(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
            for bit, name in self._names.items():
                if r & bit:
                    self.reasons.add(name)
        except Exception as e:
            self.reasons.add(f"nvml_error:{type(e).__name__}")

    def _run(self):
        while not self._stop.is_set():
            self.sample_now()
            time.sleep(self.period)

    def __enter__(self):
        self.start()
        self.samples.clear()
        self._stop.clear()
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self.sample_now()          # at least one sample while the GPU is still busy / just finished
        self._stop.set()
        self._t.join(timeout=2)

    def summary(self):
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}
