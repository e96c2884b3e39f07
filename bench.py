#!/usr/bin/env python
"""bench.py — decode tokens/s (and NVML joules/token) of the generation path, BASELINE.json's metric.

A "step" is one /api/generate-shaped request: reset the session, evaluate a synthetic prompt, decode
`--new-tokens` tokens greedily, batch 1, on random-init weights of the named architecture (default: the
north-star config, llama3.1:8b Q4_0).  `value` is device-timed decode throughput (CUDA events on the
engine's stream, weights and KV resident in HBM); `e2e` is the same request timed on the host through the
C ABI with host token buffers (prompt H2D, tokens D2H, prefill included).  With --gpus N every rank runs an
independent replica of the model (the run table shards by row): weak scaling, no collective
on the data path; torch.distributed is used only for the barrier and the max-over-ranks time.  After the
replica measurement the N ranks also form ONE tensor-parallel group and serve the same request once more
(`tp` in the line; the run fails when its tokens differ from the replica's).

The default line also carries `per_length` (the reference's three length budgets, experiment/RunnerConfig.py:82,120:
133 / 667 / 1 333 new tokens), `per_model` (the seven tags of RunnerConfig.py:80) and `roofline.per_kernel`
(an in-kernel %globaltimer timeline of one decoded token).

`--impl reference` times the CPU restatement of the same path (oracle/, PyTorch fp32, all host threads) on
a bounded sample (2 real-size layers + the LM head, extrapolated to the full depth).  That arm imports
nothing of the product: weights come from oracle/randinit.py.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LENGTHS = (133, 667, 1333)   # ceil(1.33 * words) for the reference's 100 / 500 / 1000-word requests


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
    ap.add_argument("--all-models", action="store_true", help="per-model numbers with the full step count (the default line carries a short per_model pass)")
    ap.add_argument("--tp", action="store_true", help="the N ranks form ONE tensor-parallel group that serves one request stream "
                    "(default: N independent model replicas, the run table's natural grain, followed by a short tensor-parallel leg)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip per_length / per_model / per_kernel / the tensor-parallel leg (A/B runs)")
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


def weight_bytes_q(a, qtype: str, head_qtype: str | None) -> int:
    """Algorithmic bytes per decoded token: every matmul weight once (Q4_0 / Q4_K 4.5 bit, Q6_K 6.5625 bit)."""
    per_layer = (a.q_dim + 2 * a.kv_dim) * a.H + a.H * a.q_dim + 3 * a.I * a.H
    head = a.V * a.H
    hq = head_qtype or qtype
    return a.L * per_layer * 18 // 32 + (head * 210 // 256 if hq == "q6_k" else head * 18 // 32)


def workload_config(args, a, tp_world: int = 0) -> dict:
    """The workload both arms name: identical keys and values in `--impl ours` and `--impl reference`."""
    wl = f"{args.model} {args.qtype}{'' if not args.head_qtype else ' + ' + args.head_qtype + ' head'} batch-1 greedy decode, random-init weights, "
    wl += (f"one request stream over a tensor-parallel group of {tp_world} GPUs (partial sums exchanged through peer memory inside the GEMV kernels)"
           if tp_world else "one replica per GPU")
    return {"workload": wl, "new_tokens": args.new_tokens, "prompt_len": args.prompt_len, "requests_per_step_per_gpu": 1,
            "l2": "inputs larger than L2 (weights %.0f MB per token)" % (weight_bytes_q(a, args.qtype, args.head_qtype) / 1e6)}


# ------------------------------------------------------------------------------------------ CPU arm
class CpuSample:
    """The oracle on a bounded sample of the workload: 2 real-size layers + the LM head of the named architecture, built
    once (model load, not timed) from oracle/randinit.py (no product code); every `run` decodes a few tokens and
    extrapolates the layer time to the full depth."""

    def __init__(self, model_key: str, threads: int | None = None):
        import dataclasses
        import numpy as np
        import torch
        from oracle import randinit
        from oracle.archs import ARCHS as OARCHS, tensor_table
        from oracle.model import OracleModel
        self.torch = torch
        self.cores = threads or os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        self.key, self.a = model_key, OARCHS[model_key]
        self.sub_l = min(2, self.a.L)
        self.osub = dataclasses.replace(self.a, L=self.sub_l)
        from concurrent.futures import ThreadPoolExecutor
        tensors = {}
        with ThreadPoolExecutor(max_workers=min(8, self.cores)) as pool:   # numpy releases the GIL; row chunks bound the temporaries
            for name, kind, layer, rows, cols in tensor_table(self.osub):
                if cols == 0:
                    tensors[name] = ("f32", randinit.make_tensor(self.osub, 1, kind, layer, rows, 0, "f32"))
                else:
                    step = max(1, (1 << 23) // cols)
                    parts = pool.map(lambda r0: randinit.make_tensor(self.osub, 1, kind, layer, rows, cols, "q4_0", r0, min(step, rows - r0)), range(0, rows, step))
                    tensors[name] = ("q4_0", np.concatenate(list(parts), axis=0))
        self.om = OracleModel(self.osub, tensors)
        self.sess = self.om.session(2048)
        self.tok = 1
        self.om.head(self.sess.hidden(self.tok))   # first call dequantises and caches the weights

    def run(self, n_tokens: int) -> float:
        """tokens/s of one pass of n_tokens decode steps, layer time extrapolated to the full depth."""
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
        return 1.0 / per_token

    def info(self, value: float, n_tokens: int, passes: str) -> dict:
        return {"value": value, "unit": "tokens/s", "cores": self.cores, "kind": "port",
                "sample": f"{self.sub_l} of {self.a.L} real-size layers + LM head of {self.key} ({self.osub.matmul_weights() / 1e6:.0f}M weights), "
                          f"{n_tokens} decode tokens per pass, {passes}, PyTorch fp32 oracle, extrapolated to {self.a.L} layers"}


def cpu_sample(model_key: str, n_tokens: int, threads: int | None = None, passes: int = 5) -> dict:
    cs = CpuSample(model_key, threads)
    cs.run(n_tokens)   # untimed
    vals = [cs.run(n_tokens) for _ in range(passes)]
    return cs.info(statistics.median(vals), n_tokens, f"median of {passes} passes after one untimed pass")


def run_reference(args):
    """The CPU restatement of the path on all host cores; under torchrun only rank 0 works, the others exit 0."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle.archs import ARCHS as OARCHS
    cs = CpuSample(args.model)
    vals = []
    t0 = time.perf_counter()
    for i in range(args.warmup + args.steps):
        v = cs.run(args.cpu_tokens)
        if i >= args.warmup:
            vals.append(v)
    v = statistics.median(vals)
    info = cs.info(v, args.cpu_tokens, f"median of {args.steps} steps after {args.warmup} warm-up steps (one step = one pass)")
    line = {"metric": "decode_tokens_per_s", "value": v, "unit": "tokens/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1000.0 * (time.perf_counter() - t0) / (args.warmup + args.steps),
            "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": workload_config(args, OARCHS[args.model]),
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


def measure(sess, model, a, args, lib, dev, dist, steps, warmup, new_tokens=None, keep_tokens=False):
    """`steps` requests on one session after `warmup` untimed ones; returns the per-rank measurements."""
    import ctypes as C
    import torch
    T = new_tokens or args.new_tokens
    prompt = prompt_tokens(args.prompt_len, a.V)
    last = [None]

    def request():
        sess.reset()
        sess.prefill(prompt)
        last[0] = sess.decode(T)

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
    # KV bytes read per decoded token, averaged over the decode positions of one request
    avg_ctx = args.prompt_len + (T - 1) / 2.0
    res = {"wall_s": wall, "eval_s": eval_ns / 1e9, "prompt_s": prompt_ns / 1e9, "tokens": steps * T,
           "launches": st1["kernel_launches"] - st0["kernel_launches"], "weight_bytes": model.weight_bytes(),
           "kv_bytes_per_token": a.kv_bytes_per_ctx_token() * avg_ctx,
           "energy_j": (e1.value - e0.value) / 1000.0 if have_energy else None, "clocks": sampler.summary()}
    if keep_tokens:
        res["last_tokens"] = [int(t) for t in last[0]]
    return res


def bench_model(eng, key, args, lib, dev, dist, steps, warmup, keep_tokens=False, lengths=None):
    """One architecture on this rank: the main measurement and, with `lengths`, one short pass per length budget."""
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.archs import ARCHS
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.engine import Model
    a = ARCHS[key]
    model = Model.random(eng, a, seed=1, qtype=args.qtype, head_qtype=args.head_qtype)
    sess = model.session(a.max_ctx)
    res = measure(sess, model, a, args, lib, dev, dist, steps, warmup, keep_tokens=keep_tokens)
    res["key"] = key
    if lengths:
        res["per_length"] = {}
        for n in lengths:
            m = measure(sess, model, a, args, lib, dev, None, 2, 1, new_tokens=n)
            tps = m["tokens"] / m["eval_s"]
            res["per_length"][str(n)] = {"decode_tokens_per_s": round(tps, 1), "e2e_tokens_per_s": round(m["tokens"] / m["wall_s"], 1),
                                         "roofline_frac": round((m["weight_bytes"] + m["kv_bytes_per_token"]) * tps / 1e9 / peaks()[0], 4),
                                         "joules_per_token": round(m["energy_j"] / m["tokens"], 4) if m["energy_j"] else None,
                                         "mean_context": round(args.prompt_len + (n - 1) / 2.0, 1)}
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


TRAFFIC_FILE = "profiles/r02_launches_summary.csv"


def ncu_traffic_per_token(model: str):
    """DRAM bytes (read + write) of one decoded token from the committed ncu launch list (captured on llama3.1:8b): per-launch
    bytes of every kernel x its launches per token.  None for other workloads.  Replayed from the file, not measured in this run."""
    path = os.path.join(ROOT, TRAFFIC_FILE)
    if model != "llama3.1:8b" or not os.path.exists(path):
        return None, None
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
    commit = None
    try:
        commit = subprocess.run(["git", "log", "-1", "--format=%h", "--", TRAFFIC_FILE], cwd=ROOT, capture_output=True, text=True, timeout=5).stdout.strip() or None
    except Exception:
        pass
    src = f"{TRAFFIC_FILE}" + (f" @ {commit}" if commit else "") + " (ncu launch list of an earlier run of this workload, replayed; not measured in this run)"
    return (total or None), src


def per_kernel_profile(eng, args, lib):
    """In-kernel timeline of one decoded token (graphs off, %globaltimer stamps of every GEMV CTA through the development hook
    b2dbg_gemv_profile_ring): for each stage of a layer the share of the token it occupies on the critical path, i.e. the
    time from the end of the previous stage to its own end, with its algorithmic bytes and the fraction of the HBM peak that
    share corresponds to.  Attention has no stamps of its own: its share runs from the end of the QKV GEMV to the moment the O
    projection's dependency wait returns."""
    import ctypes as C
    import numpy as np
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.archs import ARCHS
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.devmem import DevBuf
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.engine import Model
    if not hasattr(lib, "b2dbg_gemv_profile_ring"):
        return None
    a = ARCHS[args.model]
    old = os.environ.get("B2_NO_GRAPH")
    os.environ["B2_NO_GRAPH"] = "1"
    try:
        model = Model.random(eng, a, seed=1, qtype=args.qtype, head_qtype=args.head_qtype)
        sess = model.session(512)
    finally:
        if old is None:
            os.environ.pop("B2_NO_GRAPH", None)
        else:
            os.environ["B2_NO_GRAPH"] = old
    ctx = args.prompt_len + args.new_tokens // 2
    sess.prefill(prompt_tokens(args.prompt_len, a.V))
    sess.decode(max(4, args.new_tokens // 2))        # to the mean context of the benchmark
    per_tok = 4 * a.L + 1
    n_tok = 4
    slots = n_tok * per_tok
    dbg = DevBuf.from_numpy(np.zeros(slots * 148 * 8, dtype=np.int64))
    lib.b2dbg_gemv_profile_ring.argtypes = [C.c_void_p, C.c_int]
    lib.b2dbg_gemv_profile_ring(dbg.ptr, slots)
    sess.decode(n_tok)
    lib.b2dbg_gemv_profile_ring(None, 1)
    d = dbg.to_numpy(np.int64, (slots, 148, 8)).astype(np.float64)
    sess.close()
    model.close()
    peak = peaks()[0]
    blk = lambda rows, cols: rows * cols // 32 * 18
    hq = args.head_qtype or args.qtype
    kv = 2 * ctx * a.kv_dim * 2
    nbytes = {"qkv": blk(a.q_dim + 2 * a.kv_dim, a.H), "attention": kv, "o": blk(a.H, a.q_dim), "gate_up": blk(2 * a.I, a.H),
              "down": blk(a.H, a.I), "lm_head": a.V * a.H * 210 // 256 if hq == "q6_k" else blk(a.V, a.H)}
    spans = {k: [] for k in nbytes}

    def grid_of(slot):   # CTAs that left stamps
        return d[slot][:, 0] > 0

    def end(slot):
        return d[slot][grid_of(slot), 5].max()

    for t in range(1, n_tok):   # the first profiled token starts behind an idle stream
        base = t * per_tok
        prev_end = end(base - 1)   # the LM head of the previous token
        tok_start = d[base][grid_of(base), 0].min()
        for l in range(a.L):
            s = base + 4 * l
            if l > 0:
                prev_end = end(s - 1)
            else:
                prev_end = tok_start   # sampling, embedding and bookkeeping between the tokens are not attributed to a stage
            qkv_end = end(s)
            o_wait = np.median(d[s + 1][grid_of(s + 1), 2])
            spans["qkv"].append(qkv_end - prev_end)
            spans["attention"].append(o_wait - qkv_end)
            spans["o"].append(end(s + 1) - o_wait)
            spans["gate_up"].append(end(s + 2) - end(s + 1))
            spans["down"].append(end(s + 3) - end(s + 2))
        h = base + 4 * a.L
        spans["lm_head"].append(end(h) - end(h - 1))
    out = []
    for k in ("qkv", "attention", "o", "gate_up", "down", "lm_head"):
        us = float(np.median(spans[k])) / 1e3
        out.append({"name": k, "bytes": int(nbytes[k]), "us": round(us, 2), "launches_per_token": 1 if k == "lm_head" else a.L,
                    "frac": round(nbytes[k] / (us * 1e-6) / 1e9 / peak, 4) if us > 0 else None})
    tok_us = sum(o["us"] * o["launches_per_token"] for o in out)
    return {"stages": out, "token_us_attributed": round(tok_us, 1), "context": ctx,
            "note": "graphs off for the stamps; a stage's time runs from the end of the previous stage to its own end (median over layers and 3 tokens)"}


def tp_leg(args, lib, local, rank, world, dist, replica_tokens, replica_tps):
    """The N ranks as ONE tensor-parallel group on the same request: tokens must equal the replica's."""
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.archs import ARCHS
    from cain_2025_device_remote_llm_energy_rep_pkg_b200.engine import Engine
    a = ARCHS[args.model]
    if a.nkv % world or a.I % (32 * world):
        return {"skipped": f"{args.model}: {a.nkv} KV heads / ffn {a.I} do not split over {world} ranks"}
    eng = Engine(local, tp_rank=rank, tp_size=world)
    eng.tp_connect_torch()
    r = bench_model(eng, args.model, args, lib, local, dist, max(2, min(args.steps, 5)), 2, keep_tokens=True)
    eng.close()
    (wall, eval_s), (_, _, energy) = aggregate(dist, r["wall_s"], r["eval_s"], r["tokens"], r["launches"], r["energy_j"])
    import torch
    same = torch.tensor([1 if r["last_tokens"] == replica_tokens else 0], dtype=torch.int32, device="cuda")
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    head = weight_bytes_q(a, args.qtype, args.head_qtype) - a.L * ((a.q_dim + 2 * a.kv_dim) * a.H + a.H * a.q_dim + 3 * a.I * a.H) * 18 // 32
    vp = a.V % (16 * world) == 0 and os.environ.get("B2_TP_REPLICATED_HEAD") != "1"
    rank_bytes = (weight_bytes_q(a, args.qtype, args.head_qtype) - head) / world + (head / world if vp else head) + r["kv_bytes_per_token"] / world
    tps = r["tokens"] / eval_s
    return {"gpus": world, "decode_tokens_per_s": round(tps, 1), "e2e_tokens_per_s": round(r["tokens"] / wall, 1),
            "joules_per_token": round(energy / r["tokens"], 4) if energy else None,
            "roofline_frac_per_gpu": round(rank_bytes * tps / 1e9 / peaks()[0], 4),
            "speedup_vs_one_replica": round(tps / replica_tps, 3), "tokens_match_replica": bool(int(same.item())),
            "exchange": "partial sums as tagged 8-byte pairs stored into every rank's memory from inside the GEMV epilogues; loads read local memory only"}


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
    extras = not args.no_extras and not tp
    a = ARCHS[args.model]
    r = bench_model(eng, args.model, args, lib, local, dist, args.steps, args.warmup, keep_tokens=True,
                    lengths=LENGTHS if (extras and rank == 0 and world == 1) else None)
    if tp:
        r["kv_bytes_per_token"] /= world
        if a.V % (16 * world) == 0 and os.environ.get("B2_TP_REPLICATED_HEAD") != "1":
            # greedy steps multiply only this rank's rows of the LM head: count 1/N of it, not the whole head the model holds
            hq = args.head_qtype or args.qtype
            head = a.V * a.H // 256 * 210 if hq == "q6_k" else a.V * a.H // 32 * 18
            r["weight_bytes"] -= head - head // world
    (wall, eval_s), (tokens, launches, energy) = aggregate(dist, r["wall_s"], r["eval_s"], r["tokens"], r["launches"], r["energy_j"])
    if tp:
        tokens = r["tokens"]   # the ranks produced the same tokens once, not N times
    per_model = None
    if rank == 0 and world == 1 and (args.all_models or extras):
        per_model = {}
        steps_m, warm_m = (max(2, args.steps // 2), 3) if args.all_models else (2, 2)
        for key in ARCHS:
            m = r if key == args.model else bench_model(eng, key, args, lib, local, None, steps_m, warm_m)
            tps = m["tokens"] / m["eval_s"]
            per_model[key] = {"decode_tokens_per_s": round(tps, 1),
                              "joules_per_token": round(m["energy_j"] / m["tokens"], 4) if m["energy_j"] else None,
                              "roofline_frac": round((m["weight_bytes"] + m["kv_bytes_per_token"]) * tps / 1e9 / peaks()[0], 4),
                              "e2e_tokens_per_s": round(m["tokens"] / m["wall_s"], 1)}
    per_kernel = None
    if rank == 0 and world == 1 and extras:
        try:
            per_kernel = per_kernel_profile(eng, args, lib)
        except Exception as ex:   # a development hook: never costs the line
            per_kernel = {"error": str(ex)}
    tp_rec = None
    if world > 1 and not tp and not args.no_extras:
        tp_rec = tp_leg(args, lib, local, rank, world, dist, r["last_tokens"], r["tokens"] / r["eval_s"])
    rc = 0
    if rank == 0:
        peak, peak_src = peaks()
        value = tokens / eval_s
        traffic, traffic_src = ncu_traffic_per_token(args.model)
        step_bytes = r["weight_bytes"] + r["kv_bytes_per_token"]
        ach = step_bytes * (r["tokens"] / r["eval_s"]) / 1e9   # per GPU: every replica reads its own copy
        line = {
            "metric": "decode_tokens_per_s", "value": value, "unit": "tokens/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1000.0 * wall / args.steps, "higher_is_better": True, "scaling": "strong" if tp else "weak",
            "vs_baseline": None, "dtype": "int4 weights x 21-bit fixed-point activations, int32/fp32 accumulate", "data": "synthetic",
            "config": workload_config(args, a, world if tp else 0),
            "e2e": {"value": tokens / wall, "unit": "tokens/s", "h2d_bytes_per_step": args.prompt_len * 4 + 16 * (1 + 1),
                    "d2h_bytes_per_step": args.new_tokens * 4,
                    "note": "host-timed b2_session_reset+b2_prefill+b2_decode per request; includes prompt evaluation"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "decode token step (CUDA graph; GEMV k_gemv_q4_0 dominates)", "achieved": ach,
                         "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": traffic, "traffic_source": traffic_src,
                         "peak_source": peak_src, "bytes_per_launch": step_bytes, "launch": "one decoded token",
                         "avg_launch_us": 1e6 * r["eval_s"] / r["tokens"]},
            "clocks": r["clocks"],
            "joules_per_token": (energy / tokens) if energy else None,
        }
        if per_kernel is not None:
            line["roofline"]["per_kernel"] = per_kernel
        if r.get("per_length"):
            line["per_length"] = r["per_length"]
        if per_model is not None:
            line["per_model"] = per_model
        if tp_rec is not None:
            line["tp"] = tp_rec
            if tp_rec.get("tokens_match_replica") is False:
                rc = 3
        if world == 1 and not args.no_cpu_baseline:
            try:
                line["cpu_baseline"] = cpu_sample(args.model, args.cpu_tokens)
            except Exception as ex:  # the baseline is reported, never required for the GPU number
                line["cpu_baseline"] = {"value": None, "unit": "tokens/s", "cores": os.cpu_count(), "kind": "port", "sample": f"failed: {ex}"}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()
    if rc:
        print("tensor-parallel tokens differ from the replica's", file=sys.stderr)
        sys.exit(rc)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
