#!/usr/bin/env python
"""bench.py — decode tokens/s (and NVML joules/token) of the generation path, BASELINE.json's metric.

A "step" is one /api/generate-shaped request: reset the session, evaluate a synthetic prompt, decode
`--new-tokens` tokens greedily, batch 1, on random-init weights of the named architecture (default: the
north-star config, llama3.1:8b Q4_0).  `value` is device-timed decode throughput (CUDA events on the
engine's stream, weights and KV resident in HBM); `e2e` is the same request timed on the host through the
C ABI with host token buffers (prompt H2D, tokens D2H, prefill included).  With --gpus N every rank runs an
independent replica of the model (the run table shards by row): weak scaling, no collective
on the data path; torch.distributed is used only for the barrier and the max-over-ranks time.

`--impl reference` times the CPU restatement of the same path (oracle/, PyTorch fp32, all host threads) on
a bounded sample (2 real-size layers + the LM head, extrapolated to the full depth).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="llama3.1:8b")
    ap.add_argument("--qtype", default="q4_0", help="format of the layer matrices: q4_0 (default, the north-star config) or q4_k")
    ap.add_argument("--head-qtype", default=None, help="format of the LM head / embedding (default: as --qtype; q6_k is what llama.cpp writes)")
    ap.add_argument("--new-tokens", type=int, default=256)
    ap.add_argument("--prompt-len", type=int, default=32)
    ap.add_argument("--all-models", action="store_true", help="also report per-model tokens/s and J/token for all seven tags")
    ap.add_argument("--tp", action="store_true", help="the N ranks form ONE tensor-parallel group that serves one request stream "
                    "(default: N independent model replicas, the run table's natural grain)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-tokens", type=int, default=6)
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, copy)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def prompt_tokens(n, vocab):
    # synthetic prompt: bytes of the reference's prompt grammar (experiment/RunnerConfig.py:120)
    text = ("In 100 words, please give me information about United States " * 8).encode()
    return [(b % vocab) for b in text[:n]]


# ------------------------------------------------------------------------------------------ CPU arm
class CpuSample:
    """The oracle on a bounded sample of the workload: 2 real-size layers + the LM head of the named architecture, built
    once (model load, not timed); every `run` decodes a few tokens and extrapolates the layer time to the full depth."""

    def __init__(self, model_key: str, threads: int | None = None):
        import dataclasses
        import torch
        from cain_2025_device_remote_llm_energy_rep_pkg_b200.archs import ARCHS
        from cain_2025_device_remote_llm_energy_rep_pkg_b200.engine import host_random_tensors
        from oracle.archs import ARCHS as OARCHS
        from oracle.model import OracleModel
        self.torch = torch
        self.cores = threads or os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        self.key, self.a = model_key, ARCHS[model_key]
        self.sub_l = min(2, self.a.L)
        sub = dataclasses.replace(self.a, L=self.sub_l)
        self.osub = dataclasses.replace(OARCHS[model_key], L=self.sub_l)
        self.om = OracleModel(self.osub, host_random_tensors(sub, seed=1))
        self.sess = self.om.session(2048)
        self.tok = 1
        self.om.head(self.sess.hidden(self.tok))   # first call dequantises and caches the weights

    def run(self, n_tokens: int) -> dict:
        torch = self.torch
        if self.sess.pos + n_tokens >= self.sess.max_ctx:
            self.sess.reset()
        t_layers = t_head = 0.0
        for _ in range(n_tokens):
            t0 = time.perf_counter()
            h = self.sess.hidden(self.tok)
            t1 = time.perf_counter()
            self.tok = int(torch.argmax(self.om.head(h)))
            t2 = time.perf_counter()
            t_layers += t1 - t0
            t_head += t2 - t1
        per_token = (t_layers / n_tokens) / self.sub_l * self.a.L + t_head / n_tokens
        return {"value": 1.0 / per_token, "unit": "tokens/s", "cores": self.cores, "kind": "port",
                "sample": f"{self.sub_l} of {self.a.L} real-size layers + LM head of {self.key} ({self.osub.matmul_weights() / 1e6:.0f}M weights), "
                          f"{n_tokens} decode tokens, PyTorch fp32 oracle, extrapolated to {self.a.L} layers"}


def cpu_sample(model_key: str, n_tokens: int, threads: int | None = None):
    return CpuSample(model_key, threads).run(n_tokens)


def run_reference(args):
    """The CPU restatement of the path on all host cores; under torchrun only rank 0 works, the others exit 0."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cs = CpuSample(args.model)
    vals, info = [], None
    t0 = time.perf_counter()
    for i in range(args.warmup + args.steps):
        info = cs.run(args.cpu_tokens)
        if i >= args.warmup:
            vals.append(info["value"])
    v = sum(vals) / len(vals)
    info["value"] = v
    line = {"metric": "decode_tokens_per_s", "value": v, "unit": "tokens/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1000.0 * (time.perf_counter() - t0) / (args.warmup + args.steps),
            "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": {"workload": f"{args.model} {args.qtype}{'' if not args.head_qtype else ' + ' + args.head_qtype + ' head'} batch-1 greedy decode, random-init weights, one replica per GPU",
                       "new_tokens": args.new_tokens, "prompt_len": args.prompt_len, "requests_per_step_per_gpu": 1,
                       "step": f"bounded sample: {args.cpu_tokens} decode tokens of the oracle per step"},
            "cpu_baseline": info, "e2e": {"value": v, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------ GPU arm
class ClockSampler(threading.Thread):
    def __init__(self, lib, dev):
        super().__init__(daemon=True)
        self.lib, self.dev, self.stop_flag, self.samples = lib, dev, False, []

    def run(self):
        import ctypes as C
        sm, mx, rs = C.c_uint32(), C.c_uint32(), C.c_uint64()
        pw = C.c_uint32()
        while not self.stop_flag:
            if self.lib.b2_nvml_clocks(self.dev, C.byref(sm), C.byref(mx), C.byref(rs)) == 0:
                self.lib.b2_nvml_power_mw(self.dev, C.byref(pw))
                self.samples.append((sm.value, mx.value, rs.value, pw.value))
            time.sleep(0.1)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml_unavailable"]}
        sms = sorted(s[0] for s in self.samples)
        names = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x10: "sync_boost",
                 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}
        bits = 0
        for s in self.samples:
            bits |= s[2]
        reasons = [n for b, n in names.items() if bits & b and n != "gpu_idle"]
        return {"sm_mhz": sms[len(sms) // 2], "sm_max_mhz": self.samples[0][1], "reasons": reasons,
                "power_w_max": max(s[3] for s in self.samples) / 1000.0, "samples": len(self.samples)}


def bench_model(eng, key, args, lib, dev, dist, steps, warmup):
    """Returns a dict of measurements for one architecture on this rank."""
    import ctypes as C
    import torch
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.archs import ARCHS
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.engine import Model
    a = ARCHS[key]
    model = Model.random(eng, a, seed=1, qtype=args.qtype, head_qtype=args.head_qtype)
    sess = model.session(a.max_ctx)
    prompt = prompt_tokens(args.prompt_len, a.V)
    T = args.new_tokens

    def request():
        sess.reset()
        sess.prefill(prompt)
        return sess.decode(T)

    for _ in range(warmup):
        request()
    sampler = ClockSampler(lib, dev)
    sampler.start()
    e0, e1 = C.c_uint64(), C.c_uint64()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    have_energy = lib.b2_nvml_energy_mj(dev, C.byref(e0)) == 0
    st0 = sess.stats()
    t0 = time.perf_counter()
    eval_ns = 0
    prompt_ns = 0
    for _ in range(steps):
        request()
        st = sess.stats()
        eval_ns += st["eval_ns"]
        prompt_ns += st["prompt_eval_ns"]
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    lib.b2_nvml_energy_mj(dev, C.byref(e1))
    if dist is not None:
        dist.barrier()
    sampler.stop_flag = True
    sampler.join()
    st1 = sess.stats()
    launches = st1["kernel_launches"] - st0["kernel_launches"]
    wbytes = model.weight_bytes()
    # KV bytes read per decoded token, averaged over the decode positions of one request
    avg_ctx = args.prompt_len + (T - 1) / 2.0
    kv_bytes = a.kv_bytes_per_ctx_token() * avg_ctx
    res = {"key": key, "wall_s": wall, "eval_s": eval_ns / 1e9, "prompt_s": prompt_ns / 1e9, "tokens": steps * T,
           "launches": launches, "weight_bytes": wbytes, "kv_bytes_per_token": kv_bytes,
           "energy_j": (e1.value - e0.value) / 1000.0 if have_energy else None, "clocks": sampler.summary()}
    sess.close()
    model.close()
    return res


def aggregate(dist, wall_s, eval_s, tokens, launches, energy_j, device="cuda"):
    """Whole-job numbers from per-rank ones: MAX of the timed quantities, SUM of tokens / launches / joules.
    No collective is on the data path (replicas are independent); this is the only place ranks talk."""
    import torch
    vec = torch.tensor([wall_s, eval_s], dtype=torch.float64, device=device)
    tot = torch.tensor([tokens, launches, energy_j or 0.0], dtype=torch.float64, device=device)
    if dist is not None:
        dist.all_reduce(vec, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    return vec.tolist(), tot.tolist()


def ncu_traffic_per_token(model: str):
    """DRAM bytes (read + write) of one decoded token from the committed ncu launch list (profiles/r01d_launches_summary.csv,
    captured on llama3.1:8b): per-launch bytes of every kernel x its launches per token.  None for other workloads."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r01d_launches_summary.csv")
    if model != "llama3.1:8b" or not os.path.exists(path):
        return None
    import csv
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.archs import ARCHS
    per_layer = ("k_gemv<0, 2,", "k_gemv<2, 1,", "k_gemv<0, 0,", "k_gemv<1, 1,", "k_attn_fused")   # gate/up, down, QKV, O, attention
    once = ("k_gemv<0, 3,", "k_embed_row", "k_argmax_partials", "k_commit_sample", "k_advance")
    total = 0.0
    for r in csv.DictReader(open(path)):
        b = (float(r["dram_read_MB_per_launch"]) + float(r["dram_write_MB_per_launch"])) * 1e6
        if any(k in r["kernel"] for k in per_layer):
            total += b * ARCHS[model].L
        elif any(k in r["kernel"] for k in once):
            total += b
    return total or None


def run_ours(args):
    import torch
    import __graft_entry__ as g
    g.build()
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import _lib
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.archs import ARCHS
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.engine import Engine
    lib = _lib.load()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist_mod
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = dist_mod
    tp = args.tp and world > 1
    if tp:   # one group, one request stream: every rank holds 1/N of each layer and returns the same tokens
        eng = Engine(local, tp_rank=rank, tp_size=world)
        eng.tp_connect_torch()
    else:
        eng = Engine(local)
    r = bench_model(eng, args.model, args, lib, local, dist, args.steps, args.warmup)
    if tp:
        r["kv_bytes_per_token"] /= world
        a_ = ARCHS[args.model]
        if a_.V % (16 * world) == 0 and os.environ.get("B2_TP_REPLICATED_HEAD") != "1":
            # greedy steps multiply only this rank's rows of the LM head: count 1/N of it, not the whole head the model holds
            hq = args.head_qtype or args.qtype
            head = a_.V * a_.H // 256 * 210 if hq == "q6_k" else a_.V * a_.H // 32 * 18
            r["weight_bytes"] -= head - head // world
    (wall, eval_s), (tokens, launches, energy) = aggregate(dist, r["wall_s"], r["eval_s"], r["tokens"], r["launches"], r["energy_j"])
    if tp:
        tokens = r["tokens"]   # the ranks produced the same tokens once, not N times
    per_model = None
    if args.all_models and rank == 0 and world == 1:
        per_model = {}
        for key in ARCHS:
            m = r if key == args.model else bench_model(eng, key, args, lib, local, None, max(2, args.steps // 2), 3)
            a = ARCHS[key]
            tps = m["tokens"] / m["eval_s"]
            per_model[key] = {"decode_tokens_per_s": round(tps, 1),
                              "joules_per_token": round(m["energy_j"] / m["tokens"], 4) if m["energy_j"] else None,
                              "roofline_frac": round((m["weight_bytes"] + m["kv_bytes_per_token"]) * tps / 1e9 / peaks()[0], 4),
                              "e2e_tokens_per_s": round(m["tokens"] / m["wall_s"], 1)}
    if rank == 0:
        a = ARCHS[args.model]
        peak, peak_src = peaks()
        value = tokens / eval_s
        traffic = ncu_traffic_per_token(args.model)
        step_bytes = r["weight_bytes"] + r["kv_bytes_per_token"]
        ach = step_bytes * (r["tokens"] / r["eval_s"]) / 1e9   # per GPU: every replica reads its own copy
        line = {
            "metric": "decode_tokens_per_s", "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1000.0 * wall / args.steps, "higher_is_better": True, "scaling": "strong" if tp else "weak",
            "vs_baseline": None, "dtype": "int4 weights x 21-bit fixed-point activations, int32/fp32 accumulate", "data": "synthetic",
            "config": {"workload": f"{args.model} {args.qtype}{'' if not args.head_qtype else ' + ' + args.head_qtype + ' head'} batch-1 greedy decode, random-init weights, " + (f"one request stream over a tensor-parallel group of {world} GPUs (partial sums exchanged through peer memory inside the GEMV kernels)" if tp else "one replica per GPU"),
                       "new_tokens": args.new_tokens, "prompt_len": args.prompt_len, "requests_per_step_per_gpu": 1,
                       "l2": "inputs larger than L2 (weights %.0f MB per token)" % (r["weight_bytes"] / 1e6)},
            "e2e": {"value": tokens / wall, "unit": "tokens/s", "h2d_bytes_per_step": args.prompt_len * 4 + 16 * (1 + 1),
                    "d2h_bytes_per_step": args.new_tokens * 4,
                    "note": "host-timed b2_session_reset+b2_prefill+b2_decode per request; includes prompt evaluation"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "decode token step (CUDA graph; GEMV k_gemv_q4_0 dominates)", "achieved": ach,
                         "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic, "peak_source": peak_src,
                         "bytes_per_launch": step_bytes, "launch": "one decoded token", "avg_launch_us": 1e6 * r["eval_s"] / r["tokens"]},
            "clocks": r["clocks"],
            "joules_per_token": (energy / tokens) if energy else None,
        }
        if per_model is not None:
            line["per_model"] = per_model
        if world == 1 and not args.no_cpu_baseline:
            try:
                line["cpu_baseline"] = cpu_sample(args.model, args.cpu_tokens)
            except Exception as ex:  # the baseline is reported, never required for the GPU number
                line["cpu_baseline"] = {"value": None, "unit": "tokens/s", "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {ex}"}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
