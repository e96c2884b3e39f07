"""Drives the UNCHANGED reference runner + experiment/RunnerConfig.py (from /root/reference, never copied)
through runner/launch.py against the HTTP server with a stub backend: checks the shims (jsonpickle, codecarbon,
sudo/powermetrics, psutil zombie fix), the wrapper config, the NVML plugin columns and the CSV schema
(data-analysis/run_table.csv:1).  Skipped where the reference checkout is absent (the GPU box)."""
import csv
import json
import os
import socket
import sys
from pathlib import Path

import pytest

from cain_2025_device_remote_llm_energy_rep_pkg_b200 import server
from cain_2025_device_remote_llm_energy_rep_pkg_b200.runner import launch, nvml_energy
from stub_backend import StubBackend

REF = Path("/root/reference")
pytestmark = pytest.mark.skipif(not (REF / "experiment-runner" / "__main__.py").exists(), reason="reference checkout not present")


def port_free(p):
    with socket.socket() as s:
        s.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)   # a previous run may have left TIME_WAIT sockets
        try:
            s.bind(("127.0.0.1", p))
            return True
        except OSError:
            return False


@pytest.mark.timeout(240)
def test_unchanged_runner_end_to_end(tmp_path):
    if not port_free(11434):
        pytest.skip("port 11434 is busy")
    srv = server.serve(StubBackend(delay_s=0.3), host="127.0.0.1", port=11434, verbose=False)
    try:
        rc = launch.main(["--reference-dir", str(REF), "--work-dir", str(tmp_path), "--name", "e2e", "--cooldown-ms", "0",
                          "--repetitions", "1", "--models", "qwen2:1.5b", "--lengths", "100", "--seed", "1"])
        log = (tmp_path / "runner.log").read_text()
        assert rc == 0, log[-3000:]
    finally:
        srv.shutdown()
        srv.server_close()
    table = tmp_path / "experiments_output" / "e2e" / "run_table.csv"
    rows = list(csv.DictReader(open(table)))
    ref_header = open(REF / "data-analysis" / "run_table.csv").readline().strip().split(",")
    # the reference's columns in its order, then the plugin's; `energy_usage_J` is appended last by the
    # reference's own after_experiment hook (RunnerConfig.py:253), after every data column
    cols = list(rows[0].keys())
    assert ref_header[-1] == "energy_usage_J" and cols[-1] == "energy_usage_J"
    assert cols[: len(ref_header) - 1] == ref_header[:-1]
    assert [c for c in rows[0] if c.startswith(("nvml__", "rapl__"))] == [c.column for c in nvml_energy.ALL_COLUMNS]
    assert len(rows) == 2 and {r["method"] for r in rows} == {"remote", "on_device"}
    for r in rows:
        assert r["__done"] == "DONE" and r["model"] == "qwen2:1.5b" and r["length"] == "100"
        assert float(r["execution_time"]) > 0.3 and r["topic"] and float(r["cpu_usage"]) >= 0
        assert float(r["nvml__eval_count"]) == 133 and float(r["nvml__tokens_per_s"]) > 0
        assert float(r["codecarbon__energy_consumed"]) >= 0 and r["energy_usage_J"] != ""
    reqs = srv.backend.requests
    assert len(reqs) == 2 and all(q[1].startswith("In 100 words, please give me information about ") for q in reqs)
    meta = json.load(open(tmp_path / "experiments_output" / "e2e" / "metadata.json"))
    assert meta["py/object"] == "ConfigValidator.Config.Models.Metadata.Metadata" and "py/b64" in meta["_md5sum"]
    # no reference file was modified or copied: topics.csv is a symlink into the checkout
    assert (tmp_path / "experiment" / "topics.csv").is_symlink()


@pytest.mark.timeout(300)
def test_two_runner_shards_share_port_11434_and_split_the_rows(tmp_path):
    """K unchanged runner processes all call :11434 (RunnerConfig.py:130); the rows of ONE seeded run table
    are dealt over them, the server spreads concurrent requests over its backends, the merged CSV holds every row once and
    each row's token count comes from its own request (matched by model and prompt)."""
    if not port_free(11434):
        pytest.skip("port 11434 is busy")
    b0, b1 = StubBackend(delay_s=0.6), StubBackend(delay_s=0.6)
    b0.device, b1.device = 0, 1
    srv = server.serve(server.MultiBackend([b0, b1]), host="127.0.0.1", port=11434, verbose=False)
    try:
        rc = launch.main(["--reference-dir", str(REF), "--work-dir", str(tmp_path), "--name", "e2e", "--cooldown-ms", "0", "--repetitions", "1",
                          "--models", "qwen2:1.5b,gemma:2b", "--lengths", "100,500", "--shards", "2", "--nvml-devices", "0,1"])
        assert rc == 0, (tmp_path / "shard0" / "runner.log").read_text()[-3000:]
    finally:
        srv.shutdown()
        srv.server_close()
    rows = list(csv.DictReader(open(tmp_path / "run_table.csv")))
    assert len(rows) == 8 and len({r["__run_id"] for r in rows}) == 8 and all(r["__done"] == "DONE" for r in rows)
    assert {(r["model"], r["method"], r["length"]) for r in rows} == {(m, me, l) for m in ("qwen2:1.5b", "gemma:2b") for me in ("remote", "on_device") for l in ("100", "500")}
    per_shard = [list(csv.DictReader(open(tmp_path / f"shard{i}" / "experiments_output" / f"e2e_s{i}" / "run_table.csv"))) for i in range(2)]
    assert [len(t) for t in per_shard] == [4, 4] and not ({r["__run_id"] for r in per_shard[0]} & {r["__run_id"] for r in per_shard[1]})
    for r in rows:   # 100 words -> 133 tokens, 500 -> 665: the record of this row's own request, not a neighbour's
        assert float(r["nvml__eval_count"]) == {"100": 133, "500": 665}[r["length"]]
        # energy and utilisation came through the server's side channel (no NVML in this container): the stub's counters; the
        # sampler's first line is 0 % while it starts, so the mean sits a little below the stub's 42 %
        assert float(r["nvml__energy_J"]) > 0 and 30.0 < float(r["gpu_usage"]) <= 42.0 and float(r["codecarbon__energy_consumed"]) > 0
    assert len(b0.requests) + len(b1.requests) == 8 and len(b0.requests) >= 2 and len(b1.requests) >= 2   # both backends worked


def test_shard_assignment_is_a_partition():
    groups = launch.shard_models(launch.MODELS, 8)
    assert sorted(sum(groups, [])) == sorted(launch.MODELS) and len(groups) == 7
    g2 = launch.shard_models(launch.MODELS, 2)
    assert sorted(sum(g2, [])) == sorted(launch.MODELS) and len(g2) == 2


def test_merge_run_tables(tmp_path):
    for i in range(2):
        with open(tmp_path / f"t{i}.csv", "w") as f:
            f.write("__run_id,__done,model\n" + f"run_{i},DONE,m{i}\n")
    n = launch.merge_run_tables([tmp_path / "t0.csv", tmp_path / "t1.csv"], tmp_path / "out.csv")
    assert n == 2 and len(open(tmp_path / "out.csv").read().splitlines()) == 3


def test_jsonpickle_shim_round_trip():
    sys.path.insert(0, str(launch.SHIMS))
    try:
        import importlib
        jp = importlib.import_module("jsonpickle")
        sys.path.insert(0, str(REF / "experiment-runner"))
        from ConfigValidator.Config.Models.Metadata import Metadata
        s = jp.encode(Metadata(b"\x01\x02\xff" * 5 + b"\x00"), indent=2)
        ref = json.load(open(REF / "experiment" / "experiments_output" / "new_runner_experiment" / "metadata.json"))
        assert json.loads(s).keys() == ref.keys()
        back = jp.decode(s)
        assert isinstance(back, Metadata) and back.md5sum == b"\x01\x02\xff" * 5 + b"\x00"
        assert jp.decode(json.dumps(ref)).md5sum and len(jp.decode(json.dumps(ref)).md5sum) == 16
    finally:
        sys.path.remove(str(launch.SHIMS))
        if str(REF / "experiment-runner") in sys.path:
            sys.path.remove(str(REF / "experiment-runner"))
        sys.modules.pop("jsonpickle", None)


def test_rapl_reader_sums_packages_and_handles_wraparound(tmp_path):
    """Host CPU energy for the plugin's rapl__cpu_energy_J column, on a fake powercap tree (two packages, one sub-domain that
    must be ignored, one counter wrapping inside the window)."""
    for name, e, mx in (("intel-rapl:0", 1_000_000, 10_000_000), ("intel-rapl:1", 9_500_000, 10_000_000), ("intel-rapl:0:0", 5, 100)):
        d = tmp_path / name
        d.mkdir()
        (d / "energy_uj").write_text(str(e))
        (d / "max_energy_range_uj").write_text(str(mx))
    r = nvml_energy.RaplReader(str(tmp_path))
    assert len(r.domains) == 2
    a = r.read_uj()
    (tmp_path / "intel-rapl:0" / "energy_uj").write_text("3000000")     # + 2 J
    (tmp_path / "intel-rapl:1" / "energy_uj").write_text("500000")      # wrapped: 9.5 -> 10 -> 0.5 = + 1 J
    b = r.read_uj()
    assert r.delta_j(a, b) == pytest.approx(3.0)
    assert nvml_energy.RaplReader(str(tmp_path / "missing")).read_uj() is None
