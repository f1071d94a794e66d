"""Helpers shared by bench.py and bench_workloads.py: measured peaks, clock sampling."""
from __future__ import annotations

import json
import os
import statistics
import subprocess
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d.get("hbm_gbs", 6650.0), "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}


def _int_peak_file():
    p = os.path.join(ROOT, "profiles", "int_peak.json")
    return json.load(open(p)) if os.path.exists(p) else None


def load_int_peak():
    """(wide-MAD/s peak of the integer pipe, where it comes from, Fp multiplications/s of the fe_mul microbenchmark),
    measured by tools/microbench (profiles/)."""
    d = _int_peak_file()
    if d:
        return (d["wide_mad_per_s"], "measured (profiles/int_peak.json: IMAD.WIDE.U32.X carry chains)",
                d.get("fp_mul_per_s", d["wide_mad_per_s"] / 288))
    nominal = 148 * 32 * 1.965e9
    return nominal, "nominal 148 SM x 32 IMAD.WIDE/clk x 1.965 GHz", nominal / 288


def load_fr_mul_rate():
    d = _int_peak_file()
    return d.get("fr_mul_per_s", d["wide_mad_per_s"] / 112) if d else 148 * 32 * 1.965e9 / 112


def host_cores() -> int:
    """Cores this process may run on (never the inherited OMP_NUM_THREADS: torchrun exports 1)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.QUERY,
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1]))
                mx.append(float(c[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def load_traffic():
    """{phase: dram bytes per MSM} from the committed ncu --set full capture (profiles/traffic.json)."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        return {k: v for k, v in json.load(open(p)).items() if not k.startswith("_")}
    return {}
