"""Samples nvidia-smi clocks / throttle reasons while a timed region runs."""
from __future__ import annotations

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
    def __init__(self, gpu_index: int = 0, period_ms: int = 50):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self.rows: List[List[str]] = []
        self._proc = None
        self._thread = None

    def __enter__(self):
        exe = shutil.which("nvidia-smi")
        if exe is None:
            return self
        self._proc = subprocess.Popen(
            [exe, f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-lms",
             str(self.period_ms), "-i", str(self.gpu_index)],
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
