"""GPU clock / throttle-reason sampler shared by every ``bench.py`` arm (B200_PROFILING.md "clocks line").

Round 1 polled ``nvidia-smi -lms 50`` around a 23 ms timed region and got 0-1 samples.  This sampler is a
child process that polls NVML directly every ~2 ms, stamps each sample with ``time.time()`` and is started
BEFORE warm-up; the parent marks the timed window with :meth:`ClockSampler.window` and the summary is computed
over the samples that fall inside it (plus the whole-run summary, so a short window still has context).  It
falls back to ``nvidia-smi -lms 10`` when NVML cannot be loaded.

Run as a script it is the child: ``python bench/clocks.py <gpu index> <out file>``.
"""
from __future__ import annotations

import os
import statistics
import subprocess
import sys
import time

REASONS = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40,
           "hw_power_brake": 0x80}
SMI_Q = ("timestamp,index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")


def _child(index: int, path: str) -> None:
    import pynvml
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(index)
    mx = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
    get_reasons = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
    with open(path, "w", buffering=1) as f:
        while True:
            try:
                sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                rs = get_reasons(h)
                pw = pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0
            except Exception:
                time.sleep(0.01)
                continue
            f.write(f"{time.time():.6f},{sm},{mx},{pw:.1f},{rs}\n")
            time.sleep(0.002)


class ClockSampler:
    def __init__(self, gpu_index: int):
        self.gpu_index, self.proc, self.mode = gpu_index, None, None
        self.path = f"/tmp/ewdml_clocks_{os.getpid()}_{gpu_index}.csv"
        self.windows = {}

    def start(self) -> "ClockSampler":
        # CUDA_VISIBLE_DEVICES renumbers devices for CUDA, not for NVML: translate when it is a plain index list
        idx = self.gpu_index
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            try:
                idx = int(vis.split(",")[self.gpu_index])
            except (ValueError, IndexError):
                pass
        try:
            self.proc = subprocess.Popen([sys.executable, os.path.abspath(__file__), str(idx), self.path],
                                         stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            time.sleep(0.3)
            if self.proc.poll() is None:
                self.mode = "nvml"
                return self
        except Exception:
            pass
        try:
            self._f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={SMI_Q}", "--format=csv,noheader,nounits", "-lms", "10",
                                          "-i", str(idx)], stdout=self._f, stderr=subprocess.DEVNULL)
            self.mode = "nvidia-smi"
        except Exception:
            self.proc = None
        return self

    def window(self, name: str, t0: float, t1: float) -> None:
        """Mark [t0, t1] (``time.time()`` stamps) as a timed region."""
        self.windows[name] = (t0, t1)

    def _rows(self):
        rows = []
        try:
            for line in open(self.path):
                p = [x.strip() for x in line.split(",")]
                try:
                    if self.mode == "nvml":
                        rows.append((float(p[0]), float(p[1]), float(p[2]), float(p[3]), int(p[4])))
                    else:
                        ts = time.mktime(time.strptime(p[0].split(".")[0], "%Y/%m/%d %H:%M:%S")) + float("0." + p[0].split(".")[1])
                        bits = 0
                        for nm, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[6:10]):
                            if v.lower().startswith("active"):
                                bits |= REASONS[nm]
                        rows.append((ts, float(p[2]), float(p[3]), float(p[4]), bits))
                except (ValueError, IndexError):
                    continue
        except OSError:
            pass
        return rows

    @staticmethod
    def _summary(rows):
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        bits = 0
        for r in rows:
            bits |= r[4]
        return {"sm_mhz": statistics.median(r[1] for r in rows), "sm_min_mhz": min(r[1] for r in rows),
                "sm_max_mhz": max(r[2] for r in rows), "power_w_max": max(r[3] for r in rows),
                "reasons": sorted(n for n, b in REASONS.items() if bits & b), "samples": len(rows)}

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "sampler": None}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        if self.mode == "nvidia-smi":
            self._f.close()
        rows = self._rows()
        try:
            os.remove(self.path)
        except OSError:
            pass
        # headline = samples inside the first marked window (the device-timed region); fall back to every sample under
        # load (SM clock above idle) when the window was too short to catch three of them
        out = None
        per = {}
        for name, (t0, t1) in self.windows.items():
            per[name] = self._summary([r for r in rows if t0 <= r[0] <= t1])
        first = next(iter(per.values()), None)
        if first and first["samples"] >= 3:
            out = dict(first)
        else:
            lo = min((w[0] for w in self.windows.values()), default=0.0)
            hi = max((w[1] for w in self.windows.values()), default=float("inf"))
            out = self._summary([r for r in rows if lo <= r[0] <= hi] or rows)
            out["note"] = "window shorter than three samples: summary over all timed regions of this run"
        out["sampler"] = f"{self.mode}, {len(rows)} samples over the whole run"
        out["windows"] = per
        return out


if __name__ == "__main__":
    _child(int(sys.argv[1]), sys.argv[2])
