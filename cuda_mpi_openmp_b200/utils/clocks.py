"""Samples nvidia-smi clocks / throttle reasons while a timed region runs."""
from __future__ import annotations

import os
import shutil
import statistics
import subprocess
import threading
from typing import Dict, List

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
          "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
          "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
_REASONS = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]


class ClockSampler:
    """Polls NVML in a thread (every ~5 ms) while the with-block runs; falls back to
    an `nvidia-smi -lms` subprocess when the NVML binding is unavailable."""

    def __init__(self, gpu_index: int = 0, period_ms: int = 5):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self.rows: List[List[str]] = []
        self._proc = None
        self._thread = None
        self._stop = threading.Event()

    def _nvml_loop(self, pynvml, handle):
        R = {
            "hw_slowdown": getattr(pynvml, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(pynvml, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(pynvml, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(pynvml, "nvmlClocksEventReasonSwPowerCap", 0x4),
        }
        get_reasons = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
        mx = pynvml.nvmlDeviceGetMaxClockInfo(handle, pynvml.NVML_CLOCK_SM)
        while not self._stop.is_set():
            try:
                sm = pynvml.nvmlDeviceGetClockInfo(handle, pynvml.NVML_CLOCK_SM)
                mask = get_reasons(handle)
                cells = ["Active" if mask & R[k] else "Not Active" for k in _REASONS]
                self.rows.append([str(self.gpu_index), str(sm), str(mx), "", ""] + cells)
            except Exception:  # noqa: BLE001
                pass
            self._stop.wait(self.period_ms / 1000.0)

    def __enter__(self):
        try:
            import pynvml

            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = self.gpu_index
            if vis:
                parts = [p for p in vis.split(",") if p.strip()]
                if self.gpu_index < len(parts) and parts[self.gpu_index].strip().isdigit():
                    idx = int(parts[self.gpu_index])
            handle = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self._thread = threading.Thread(target=self._nvml_loop, args=(pynvml, handle), daemon=True)
            self._thread.start()
            return self
        except Exception:  # noqa: BLE001 - fall back to the command-line tool
            self._thread = None
        exe = shutil.which("nvidia-smi")
        if exe is None:
            return self
        self._proc = subprocess.Popen(
            [exe, f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-lms",
             str(max(50, self.period_ms)), "-i", str(self.gpu_index)],
            stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        self._thread = threading.Thread(target=self._pump, daemon=True)
        self._thread.start()
        return self

    def _pump(self):
        for line in self._proc.stdout:
            cells = [c.strip() for c in line.split(",")]
            if len(cells) >= 9:
                self.rows.append(cells)

    def __exit__(self, *exc):
        self._stop.set()
        if self._proc is None and self._thread is not None:
            self._thread.join(timeout=2)
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self._proc.kill()
            self._thread.join(timeout=2)
        return False

    def summary(self) -> Dict:
        sm, mx, reasons = [], 0.0, set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx = max(mx, float(r[2]))
            except ValueError:
                continue
            for name, cell in zip(_REASONS, r[5:9]):
                if cell.lower().startswith("active"):
                    reasons.add(name)
        return {
            "sm_mhz": statistics.median(sm) if sm else None,
            "sm_max_mhz": mx or None,
            "reasons": sorted(reasons),
            "samples": len(sm),
        }
