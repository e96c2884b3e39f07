"""tools/sweep_replay.py restates the reference's run table (experiment/RunnerConfig.py:80-88) and its request grammar
(RunnerConfig.py:120-131) for boxes without the reference checkout; these checks pin that restatement to the reference's
own numbers (no GPU, no server)."""
import os
import sys

from conftest import ROOT

sys.path.insert(0, os.path.join(ROOT, "tools"))
import sweep_replay as sw  # noqa: E402


def test_run_table_has_the_reference_shape():
    rows = sw.run_table(30, 0, sw.MODELS, sw.LENGTHS)
    assert len(rows) == 7 * 2 * 3 * 30 == 1260
    assert {r["model"] for r in rows} == set(sw.MODELS) and {r["method"] for r in rows} == {"remote", "on_device"}
    assert {r["length"] for r in rows} == {"100", "500", "1000"}
    assert len({r["__run_id"] for r in rows}) == 1260
    assert rows == sw.run_table(30, 0, sw.MODELS, sw.LENGTHS)   # seeded, unlike the reference
    assert rows != sw.run_table(30, 1, sw.MODELS, sw.LENGTHS)


def test_columns_start_with_the_reference_csv_header():
    ref = ["__run_id", "__done", "model", "method", "length", "topic", "execution_time", "cpu_usage", "gpu_usage", "memory_usage",
           "codecarbon__energy_consumed"]                  # data-analysis/run_table.csv:1 (without the post-hoc energy_usage_J)
    assert sw.COLUMNS[:len(ref)] == ref
    assert all(c.startswith("nvml__") for c in sw.COLUMNS[len(ref):])


def test_model_list_matches_the_package():
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.runner import launch
    assert sw.MODELS == launch.MODELS
