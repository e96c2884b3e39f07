"""Replay of the reference's run table against N B200 servers, one per GPU.

The GPU box has no checkout of the replication package, so the unchanged runner cannot run there (it does in
tests/test_runner_e2e.py, against a stub backend).  This tool restates only what the runner does on the wire and on disk:
  * the run table of experiment/RunnerConfig.py:80-88 (7 models x 2 methods x 3 lengths x repetitions, shuffled, seeded here),
  * the request of RunnerConfig.py:120-131 (`curl http://<host>:11434/api/generate -d '{"model", "prompt", "stream": false}'`:
    a JSON body under curl's form content type), one request at a time per server,
  * run_table.csv with the reference's columns (data-analysis/run_table.csv:1) plus the nvml__* columns of the energy plugin.
Rows are independent requests (ExperimentController.py:119-132): with --gpus N the models are dealt to N servers (heaviest
first, runner/launch.py:shard_models) and N clients work through their rows concurrently; no collective anywhere.
The topics are a fixed sample of Wikipedia-style titles (the reference draws from its 101-line experiment/topics.csv with an
unseeded random.choice; a topic only changes a few prompt bytes).

    python tools/sweep_replay.py --gpus 8 --repetitions 1 --out gpurun_out/sweep8
prints one JSON line: wall time, requests, tokens, whole-sweep tokens/s, joules, per-model tokens/s and J/token.
"""
from __future__ import annotations

import argparse
import csv
import http.client
import json
import os
import random
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
MODELS = ['llama3.1:8b', 'gemma:2b', 'gemma:7b', 'phi3:3.8b', 'qwen2:1.5b', 'qwen2:7b', 'mistral:7b']   # RunnerConfig.py:80
METHODS = ['remote', 'on_device']                                                                        # RunnerConfig.py:81
LENGTHS = ['100', '500', '1000']                                                                         # RunnerConfig.py:82
TOPICS = ["United States", "Prince Philip, Duke of Edinburgh", "India", "World War II", "Elizabeth II", "Cristiano Ronaldo",
          "The Beatles", "Albert Einstein", "Game of Thrones", "Periodic table"]
COLUMNS = ["__run_id", "__done", "model", "method", "length", "topic", "execution_time", "cpu_usage", "gpu_usage", "memory_usage",
           "codecarbon__energy_consumed", "nvml__energy_J", "nvml__eval_count", "nvml__tokens_per_s", "nvml__joules_per_token"]


def run_table(repetitions: int, seed: int, models, lengths):
    rows = [{"model": m, "method": me, "length": ln} for m in models for me in METHODS for ln in lengths for _ in range(repetitions)]
    rng = random.Random(seed)
    rng.shuffle(rows)                                       # RunTableModel.py:95-96
    for i, r in enumerate(rows):
        r["__run_id"] = f"run_{i}_repetition_{i % max(1, repetitions)}"
        r["topic"] = rng.choice(TOPICS)                     # RunnerConfig.py:118
    return rows


def post_generate(port: int, model: str, prompt: str) -> dict:
    body = json.dumps({"model": model, "prompt": prompt, "stream": False})
    c = http.client.HTTPConnection("127.0.0.1", port, timeout=600)
    c.request("POST", "/api/generate", body=body, headers={"Content-Type": "application/x-www-form-urlencoded"})   # what curl -d sends
    resp = c.getresponse()
    data = json.loads(resp.read())
    c.close()
    if resp.status != 200:
        raise RuntimeError(f"{model}: HTTP {resp.status}: {data}")
    return data


def wait_ready(port: int, deadline: float):
    while time.time() < deadline:
        try:
            c = http.client.HTTPConnection("127.0.0.1", port, timeout=2)
            c.request("GET", "/api/version")
            if c.getresponse().status == 200:
                return
        except OSError:
            time.sleep(0.5)
    raise TimeoutError(f"server on port {port} did not come up")


def main():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--repetitions", type=int, default=1, help="reference: 30")
    ap.add_argument("--models", default=",".join(MODELS))
    ap.add_argument("--lengths", default=",".join(LENGTHS))
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--port", type=int, default=11434)
    ap.add_argument("--out", default="gpurun_out/sweep")
    args = ap.parse_args()
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import _lib
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.runner.launch import shard_models
    import ctypes as C
    lib = _lib.load()
    models = args.models.split(",")
    shards = shard_models(models, args.gpus)
    # GPUs left over (8 GPUs, 7 models): a second replica of the models that take longest per request
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.archs import ARCHS
    spare = args.gpus - len(shards)
    for m in sorted(models, key=lambda m: -ARCHS[m].weight_bytes_q4())[:max(0, spare)]:
        shards.append([m])
    rows = run_table(args.repetitions, args.seed, models, args.lengths.split(","))
    os.makedirs(args.out, exist_ok=True)
    procs = []
    try:
        for i, ms in enumerate(shards):     # one server per GPU, its models loaded before the clock starts (Ollama keeps models resident too)
            log = open(os.path.join(args.out, f"server{i}.log"), "w")
            procs.append(subprocess.Popen([sys.executable, "-m", "cain_2025_device_remote_llm_energy_rep_pkg_b200.server", "--device", str(i),
                                           "--port", str(args.port + i), "--preload", ",".join(ms)], cwd=ROOT, stdout=log, stderr=log))
        for i in range(len(shards)):
            wait_ready(args.port + i, time.time() + 300)
        # rows of a model go to its server; a model with two replicas has its rows dealt alternately
        where = {}
        for i, ms in enumerate(shards):
            for m in ms:
                where.setdefault(m, []).append(i)
        seen = {}
        for r in rows:
            k = seen.get(r["model"], 0)
            seen[r["model"]] = k + 1
            r["_server"] = where[r["model"]][k % len(where[r["model"]])]
        done, lock = [], threading.Lock()

        def client(i):
            for r in [r for r in rows if r["_server"] == i]:
                prompt = f"In {r['length']} words, please give me information about " + r["topic"]      # RunnerConfig.py:120
                e0, e1 = C.c_uint64(), C.c_uint64()
                lib.b2_nvml_energy_mj(i, C.byref(e0))
                t0 = time.time()
                resp = post_generate(args.port + i, r["model"], prompt)
                dt = time.time() - t0
                have = lib.b2_nvml_energy_mj(i, C.byref(e1)) == 0
                joules = (e1.value - e0.value) / 1000.0 if have else None
                n = int(resp.get("eval_count", 0))
                out = dict(r, __done="DONE", execution_time=round(dt, 4), cpu_usage="", gpu_usage="", memory_usage="",
                           codecarbon__energy_consumed="", nvml__energy_J=joules, nvml__eval_count=n,
                           nvml__tokens_per_s=round(n / (resp.get("eval_duration", 0) * 1e-9), 2) if resp.get("eval_duration") else "",
                           nvml__joules_per_token=round(joules / n, 4) if (joules is not None and n) else "")
                with lock:
                    done.append(out)

        threads = [threading.Thread(target=client, args=(i,)) for i in range(len(shards))]
        t0 = time.time()
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        wall = time.time() - t0
    finally:
        for p in procs:     # the exact processes started above
            p.terminate()
        for p in procs:
            try:
                p.wait(timeout=20)
            except subprocess.TimeoutExpired:
                p.kill()
    done.sort(key=lambda r: int(r["__run_id"].split("_")[1]))
    with open(os.path.join(args.out, "run_table.csv"), "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=COLUMNS)
        w.writeheader()
        for r in done:
            w.writerow({k: r.get(k, "") for k in COLUMNS})
    per = {}
    for r in done:
        d = per.setdefault(r["model"], {"tokens": 0, "seconds": 0.0, "joules": 0.0, "requests": 0})
        d["tokens"] += r["nvml__eval_count"]; d["seconds"] += r["execution_time"]; d["joules"] += r["nvml__energy_J"] or 0.0; d["requests"] += 1
    tokens = sum(d["tokens"] for d in per.values())
    print(json.dumps({"sweep": "7 models x 2 methods x 3 lengths" if len(models) == 7 else ",".join(models), "repetitions": args.repetitions,
                      "gpus": args.gpus, "servers": len(shards), "requests": len(done), "expected_requests": len(rows), "wall_s": round(wall, 2),
                      "tokens": tokens, "sweep_tokens_per_s": round(tokens / wall, 1), "joules": round(sum(d["joules"] for d in per.values()), 1),
                      "per_model": {m: {"requests": d["requests"], "tokens_per_s_incl_http": round(d["tokens"] / d["seconds"], 1),
                                        "joules_per_token": round(d["joules"] / d["tokens"], 4) if d["tokens"] else None} for m, d in per.items()},
                      "shards": shards}), flush=True)
    return 0 if len(done) == len(rows) else 1


if __name__ == "__main__":
    sys.exit(main())
