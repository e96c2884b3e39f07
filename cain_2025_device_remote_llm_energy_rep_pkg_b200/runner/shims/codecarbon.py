"""Stand-in for `codecarbon` (missing in this image) with the surface the reference's plugin touches
(Plugins/Profilers/CodecarbonWrapper.py:53-57, 89-97): `OfflineEmissionsTracker(project_name=, output_dir=,
country_iso_code=)` / `EmissionsTracker`, `.start()`, `.stop()`, `._output_dir`, `._output_file`, and a ONE-row
`emissions.csv` whose `energy_consumed` column is kWh.  Energy here = GPU energy from
nvmlDeviceGetTotalEnergyConsumption (devices in $B2_NVML_DEVICES, default 0) + CPU package energy from RAPL
when /sys/class/powercap is readable; both 0 when neither source exists (e.g. a CPU-only test box)."""
import csv
import glob
import os
import time


def _rapl_uj():
    tot, ok = 0, False
    for p in glob.glob("/sys/class/powercap/intel-rapl:[0-9]*/energy_uj"):
        try:
            with open(p) as f:
                tot += int(f.read().strip())
                ok = True
        except OSError:
            pass
    return tot if ok else None


def _server_mj(devs):
    """The server's side channel first: it holds NVML open, this process was forked for one run and would pay an NVML start-up."""
    try:
        import json
        import urllib.request
        host = os.environ.get("B2_SERVER_HOST") or "127.0.0.1"
        with urllib.request.urlopen(f"http://{host}:{os.environ.get('B2_SERVER_PORT', '11434')}/b2/gpus", timeout=1.0) as r:
            gpus = json.loads(r.read().decode()).get("gpus", {})
        vals = [gpus[str(d)]["energy_mj"] for d in devs]
        return None if any(v is None for v in vals) else sum(vals)
    except Exception:
        return None


def _nvml_mj():
    devs = [int(d) for d in os.environ.get("B2_NVML_DEVICES", "0").split(",") if d != ""]
    got = _server_mj(devs)
    if got is not None:
        return got
    try:
        import pynvml
        pynvml.nvmlInit()
        return sum(pynvml.nvmlDeviceGetTotalEnergyConsumption(pynvml.nvmlDeviceGetHandleByIndex(d)) for d in devs)
    except Exception:
        return None


class OfflineEmissionsTracker:
    def __init__(self, *args, project_name="codecarbon", output_dir=".", output_file="emissions.csv", country_iso_code=None, **kwargs):
        self._project_name, self._output_dir, self._output_file = project_name, output_dir, output_file
        self._country_iso_code = country_iso_code
        self._t0 = self._g0 = self._c0 = None

    def start(self):
        self._t0, self._g0, self._c0 = time.time(), _nvml_mj(), _rapl_uj()

    def stop(self):
        t1, g1, c1 = time.time(), _nvml_mj(), _rapl_uj()
        gpu_kwh = ((g1 - self._g0) / 1000.0 / 3.6e6) if (g1 is not None and self._g0 is not None) else 0.0
        cpu_kwh = ((c1 - self._c0) / 1e6 / 3.6e6) if (c1 is not None and self._c0 is not None and c1 >= self._c0) else 0.0
        dur = t1 - (self._t0 or t1)
        row = {"timestamp": time.strftime("%Y-%m-%dT%H:%M:%S"), "project_name": self._project_name, "duration": dur,
               "emissions": 0.0, "emissions_rate": 0.0, "cpu_power": 0.0, "gpu_power": (gpu_kwh * 3.6e6 / dur) if dur > 0 else 0.0,
               "ram_power": 0.0, "cpu_energy": cpu_kwh, "gpu_energy": gpu_kwh, "ram_energy": 0.0,
               "energy_consumed": cpu_kwh + gpu_kwh, "country_iso_code": self._country_iso_code or ""}
        os.makedirs(self._output_dir, exist_ok=True)
        with open(os.path.join(self._output_dir, self._output_file), "w", newline="") as f:
            w = csv.DictWriter(f, fieldnames=list(row))
            w.writeheader()
            w.writerow(row)
        return row["emissions"]


EmissionsTracker = OfflineEmissionsTracker
