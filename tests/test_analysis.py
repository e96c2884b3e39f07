"""runner/analysis.py against the numbers the reference's own notebook prints for its own run table
(`data-analysis/analysis-visualization.ipynb`, outputs of code cells 33, 37 and 42 over `data-analysis/run_table.csv`).
The CSV stays in the reference checkout (nothing is copied); without it only the closed-form checks run."""
import os

import numpy as np
import pytest

from cain_2025_device_remote_llm_energy_rep_pkg_b200.runner import analysis as A

REF_TABLE = "/root/reference/data-analysis/run_table.csv"
needs_ref = pytest.mark.skipif(not os.path.exists(REF_TABLE), reason="reference checkout not present")


@needs_ref
def test_notebook_tables_are_reproduced():
    res = A.analyse(REF_TABLE)
    # cell 33, printed data frame
    shapiro = {"on_device_short": (0.9547488, 3.202963e-05), "on_device_medium": (0.8745747, 3.521535e-11),
               "on_device_long": (0.9290627, 5.084372e-08), "remote_short": (0.8594111, 1.116692e-11),
               "remote_medium": (0.9094642, 2.101168e-08), "remote_long": (0.9322779, 6.199450e-07)}
    for k, (w, p) in shapiro.items():
        assert res["shapiro"][k][0] == pytest.approx(w, abs=5e-8) and res["shapiro"][k][1] == pytest.approx(p, rel=1e-6)
    # cell 37, LaTeX table (W printed as an integer, delta and interval to three decimals; p "< 2.2e-16")
    h1 = {"short": (28370, 0.941, 0.905, 0.964), "medium": (28486, 0.956, 0.920, 0.976), "long": (29587, 0.912, 0.868, 0.942)}
    for k, (w, d, lo, hi) in h1.items():
        r = res["h1"][k]
        assert round(r["W"] + 1e-9) == w or abs(r["W"] - w) <= 0.5
        assert round(r["delta"], 3) == d and round(r["lower"], 3) == lo and round(r["upper"], 3) == hi
        assert r["p"] < 2.2e-16 and r["magnitude"] == "Large"
    # cell 42, first rows of the LaTeX table
    sp = res["spearman"]
    assert round(sp["on_device_short"]["execution_time"][0], 3) == 0.991
    assert round(sp["on_device_short"]["cpu_usage"][0], 3) == -0.708
    assert round(sp["on_device_short"]["gpu_usage"][0], 3) == -0.176 and round(sp["on_device_short"]["gpu_usage"][1], 3) == 0.023
    assert round(sp["on_device_medium"]["memory_usage"][0], 3) == 0.520
    assert round(sp["remote_short"]["memory_usage"][0], 3) == 0.010 and round(sp["remote_short"]["memory_usage"][1], 3) == 0.893
    assert round(sp["remote_medium"]["gpu_usage"][0], 3) == -0.723


def test_iqr_filter_is_sequential_and_type7():
    rows = [{"a": float(v), "b": 0.0} for v in [1, 2, 3, 4, 100]] + [{"a": 2.5, "b": 50.0}]
    kept = A.remove_outliers(rows, ["a", "b"])
    assert [r["a"] for r in kept] == [1.0, 2.0, 3.0, 4.0]            # 100 goes on column a, the b outlier on column b afterwards


def test_cliffs_delta_closed_forms():
    d, lo, hi, mag = A.cliffs_delta([5, 6, 7, 8], [1, 2, 3, 4])
    assert d == 1.0 and mag == "Large"
    d, *_ = A.cliffs_delta([1, 2, 3, 4], [1, 2, 3, 4])
    assert d == 0.0
    d, lo, hi, _ = A.cliffs_delta([1, 3, 5, 7, 9], [2, 4, 6, 8, 10])
    assert d == pytest.approx(-0.2) and lo < d < hi


def test_b200_table_with_nvml_energy(tmp_path):
    """A table as tools/sweep_replay.py / the NVML plugin write it: CPU / GPU / RAM columns empty, energy in nvml__energy_J."""
    rng = np.random.default_rng(0)
    p = tmp_path / "run_table.csv"
    with open(p, "w") as f:
        f.write("__run_id,__done,model,method,length,topic,execution_time,cpu_usage,gpu_usage,memory_usage,codecarbon__energy_consumed,nvml__energy_J\n")
        for i in range(120):
            ln = (100, 500, 1000)[i % 3]
            t = ln / 600.0 * (1 + 0.05 * rng.standard_normal())
            f.write(f"run_{i},DONE,llama3.1:8b,{('remote', 'on_device')[i % 2]},{ln},X,{t},,,,,{t * 700.0}\n")
    res = A.analyse(str(p), energy="nvml__energy_J")
    assert set(res["n"]) == {f"{m}_{l}" for m in ("on_device", "remote") for l in ("short", "medium", "long")}
    assert all(abs(v["delta"]) < 0.5 for v in res["h1"].values())                 # the two methods are the same server here
    assert all(v["execution_time"][0] > 0.99 for v in res["spearman"].values())
