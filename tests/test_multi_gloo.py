"""N>1 host logic on CPU (gloo, world_size 2): the rank aggregation of bench.py (max of times, sum of tokens), the
run-table sharding by model, and the reference arm's "rank 0 only" rule.  The data path itself has no collective."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    import bench
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # rank r: wall 1+r s, device time 0.5+r s, 100*(r+1) tokens, 10 launches, 2.5 J
    (wall, ev), (tok, launches, joules) = bench.aggregate(dist, 1.0 + rank, 0.5 + rank, 100 * (rank + 1), 10, 2.5, device="cpu")
    dist.barrier()
    if rank == 0:
        json.dump({"wall": wall, "eval": ev, "tokens": tok, "launches": launches, "joules": joules}, open(out, "w"))
    dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_aggregate_max_time_sum_tokens_world2(tmp_path):
    out = str(tmp_path / "agg.json")
    mp.spawn(_worker, args=(2, free_port(), out), nprocs=2, join=True)
    r = json.load(open(out))
    assert r == {"wall": 2.0, "eval": 1.5, "tokens": 300.0, "launches": 20.0, "joules": 5.0}
    # value = all tokens / max time: whole-job throughput, not a per-GPU mean
    assert r["tokens"] / r["eval"] == 200.0


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, env=env, timeout=120)
    assert p.returncode == 0 and p.stdout.strip() == ""


def test_replica_sharding_covers_every_model_once():
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.runner import launch
    for n in (2, 4, 8):
        groups = launch.shard_models(launch.MODELS, n)
        flat = sum(groups, [])
        assert sorted(flat) == sorted(launch.MODELS) and len(groups) == min(n, 7)
    # heaviest models are spread first: gemma:7b and llama3.1:8b never share a shard when there are >= 2
    g2 = launch.shard_models(launch.MODELS, 2)
    assert not any({"gemma:7b", "llama3.1:8b"} <= set(g) for g in g2)


def _handle_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.engine import TP_HANDLE_BYTES, gather_tp_handles
    dist.init_process_group("gloo", rank=rank, world_size=world)

    def all_gather(obj):
        got = [None] * world
        dist.all_gather_object(got, obj)
        return got[::-1]          # a transport may deliver in any order: the result must still be rank-ordered

    blob = gather_tp_handles(bytes([rank + 1]) * TP_HANDLE_BYTES, rank, world, all_gather)
    open(out + str(rank), "wb").write(blob)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_tp_handle_exchange_is_rank_ordered_world2(tmp_path):
    """Host side of the tensor-parallel set-up: every rank ends up with the same rank-ordered block of handles."""
    out = str(tmp_path / "handles")
    mp.spawn(_handle_worker, args=(2, free_port(), out), nprocs=2, join=True)
    want = bytes([1]) * 64 + bytes([2]) * 64
    assert open(out + "0", "rb").read() == want and open(out + "1", "rb").read() == want


def test_tp_handle_exchange_rejects_bad_input():
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.engine import gather_tp_handles
    ok = bytes(64)
    with pytest.raises(ValueError):
        gather_tp_handles(bytes(10), 0, 2, lambda o: [o, (1, ok)])
    with pytest.raises(ValueError):
        gather_tp_handles(ok, 0, 2, lambda o: [o])                       # a rank is missing
    with pytest.raises(ValueError):
        gather_tp_handles(ok, 0, 2, lambda o: [o, (0, ok)])              # the same rank twice
    with pytest.raises(ValueError):
        gather_tp_handles(ok, 0, 2, lambda o: [o, (1, bytes(8))])        # short handle from a peer


@pytest.mark.timeout(600)
def test_reference_arm_prints_one_json_line():
    """`bench.py --impl reference` (the oracle on the host cores, bounded sample) end to end: one JSON line with the contract's keys."""
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    env["B2_LIB"] = "/nonexistent/libb2llm.so"   # the reference arm must not load product code: with this it would fail loudly if it tried
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", "--cpu-tokens", "2"],
                       capture_output=True, text=True, env=env, timeout=580)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1, p.stderr[-2000:]
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "decode_tokens_per_s" and d["value"] > 0
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    # both arms name the workload with the same dict (the driver compares them)
    import argparse
    import bench
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.archs import ARCHS
    ours = bench.workload_config(argparse.Namespace(model="llama3.1:8b", qtype="q4_0", head_qtype=None, new_tokens=256, prompt_len=32), ARCHS["llama3.1:8b"])
    assert d["config"] == ours
