"""Tensor-parallel decode: 2 ranks = 2 processes = 2 GPUs, partial sums exchanged through
peer memory inside the GEMV kernels.  Needs two GPUs (`gpurun --gpus 2 -- python -m pytest tests/test_tp_gpu.py -m gpu`);
on a one-GPU box the tests skip.  Every check goes through the C ABI; the oracle is the PyTorch CPU forward (tiny
configurations) and the committed full-size fixture (llama3.1:8b, 256 greedy tokens)."""
import os
import socket
import sys
from dataclasses import replace

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu
PROMPT = [1, 5, 9, 200, 31, 77, 999, 3]


def _n_gpus() -> int:
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


needs2 = pytest.mark.skipif(_n_gpus() < 2, reason="tensor-parallel tests need two GPUs")


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def tp_arch(key):
    """Tiny variants whose KV heads split over two ranks."""
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import archs
    from oracle.archs import TINY as OTINY
    over = {"llama3.1:8b": dict(nh=4, nkv=2), "qwen2:7b": dict(H=512, nh=4, nkv=2, I=1216), "gemma:7b": dict(nh=2, nkv=2),
            "phi3:3.8b": dict(H=384, nh=4, nkv=4, I=448)}[key]
    return replace(archs.TINY[key], **over), replace(OTINY[key], **over)


def _worker(rank, world, port, case, out_dir):
    """One rank: connect, build the model, run the case, save what the parent checks."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import archs, engine, tokenizer
    dist.init_process_group("gloo", rank=rank, world_size=world)
    eng = engine.Engine(rank, tp_rank=rank, tp_size=world)
    eng.tp_connect_torch()
    res = {}
    if case["kind"] == "tiny":
        a, _ = tp_arch(case["key"])
        tensors = engine.host_random_tensors(a, seed=case["seed"])
        model = engine.Model.from_tensors(eng, a, tensors, max_ctx=320)
        sess = model.session(320)
        res["logits"] = sess.forward_tokens(PROMPT + [3, 4, 5, 6, 7, 8])
        sess.reset()
        res["greedy"] = sess.generate_greedy(PROMPT, case["n_new"])
        # the device-side initialiser fills each rank's window of every tensor with the same bytes
        model2 = engine.Model.random(eng, a, seed=case["seed"], max_ctx=320)
        res["logits_devinit"] = model2.session(320).forward_tokens(PROMPT)
        sess.reset()
        res["prefill_logits"] = sess.prefill(PROMPT, want_logits=True)
        samp = engine.SamplerCfg(8, 0.8, 0.95, 1234)
        res["sampled"] = sess.decode(24, sampler=samp)
        # a greedy call (LM head split over the ranks in all steps but its last) must leave complete logits for whatever comes next:
        # greedy, then sampled, then logits -- against the same calls on a single-GPU engine with the same weights
        sess.reset()
        sess.prefill(PROMPT)
        g1 = sess.decode(12)
        s1 = sess.decode(6, sampler=samp)
        g2, lg2 = sess.decode(3, want_logits=True)
        res["mixed"] = np.concatenate([g1, s1, g2])
        res["mixed_logits"] = lg2
        solo = engine.Model.from_tensors(engine.Engine(rank), a, tensors, max_ctx=320).session(320)
        solo.prefill(PROMPT)
        h1 = solo.decode(12)
        t1 = solo.decode(6, sampler=samp)
        h2, lh2 = solo.decode(3, want_logits=True)
        res["solo_mixed"] = np.concatenate([h1, t1, h2])
        res["solo_mixed_logits"] = lh2
    elif case["kind"] == "stress":
        # passes without the LM head (prompt body), vocabulary-parallel greedy steps, sampled steps (no arg-max exchange)
        # and replicated steps alternate for more than a thousand forward passes; the slot buffers and tags must never mix rounds.
        if case.get("no_pdl"):
            os.environ["B2_NO_PDL"] = "1"       # read at b2_session_create: plain stream order instead of programmatic launches
        a, _ = tp_arch(case["key"])
        tensors = engine.host_random_tensors(a, seed=case["seed"])
        sess = engine.Model.from_tensors(eng, a, tensors, max_ctx=160).session(160)
        solo = engine.Model.from_tensors(engine.Engine(rank), a, tensors, max_ctx=160).session(160)
        samp = engine.SamplerCfg(8, 0.8, 0.95, 99)
        got, want, passes = [], [], 0
        for rnd in range(case["rounds"]):
            prompt = [1 + (7 * i + 13 * rnd) % (a.V - 1) for i in range(40 + rnd % 5)]
            for s_, out in ((sess, got), (solo, want)):
                s_.reset()
                s_.prefill(prompt)                       # tensor parallel: token by token, every pass but the last without the LM head
                out.append(s_.decode(24))                # greedy: LM head split over the ranks, last step replicated
                out.append(s_.decode(16, sampler=samp))  # sampled: whole head on every rank, no arg-max exchange
                out.append(s_.decode(1))                 # a single greedy step: replicated
                out.append(s_.decode(9))
            passes += len(prompt) + 50
        res["got"], res["want"], res["passes"] = np.concatenate(got), np.concatenate(want), np.array(passes)
    elif case["kind"] == "missing_peer":   # rank 1 never makes the call: rank 0 must get an error in seconds, not hang
        import time
        from cain_2025_device_remote_llm_energy_rep_pkg_b200._lib import B2Error
        a, _ = tp_arch(case["key"])
        model = engine.Model.random(eng, a, seed=3, max_ctx=64)
        sess = model.session(64)
        dist.barrier()
        code, dt = 0, 0.0
        if rank == 0:
            t0 = time.time()
            try:
                sess.prefill(PROMPT)
            except B2Error as ex:
                code = ex.code
            dt = time.time() - t0
        res["code"], res["seconds"] = np.array(code), np.array(dt)
    elif case["kind"] == "tiny_kq":   # K-quant layer matrices (column windows on super-block borders) and a Q6_K head, from host tensors
        from oracle import randinit as orand
        a, oa = tp_arch(case["key"])
        tensors = orand.make_model(oa, seed=case["seed"], qtype="q4_k", head_qtype="q6_k")
        model = engine.Model.from_tensors(eng, a, tensors, max_ctx=96)
        sess = model.session(96)
        res["logits"] = sess.forward_tokens(PROMPT)
        sess.reset()
        res["greedy"] = sess.generate_greedy(PROMPT, 40)
    else:
        a = archs.ARCHS[case["key"]]
        prompt = tokenizer.encode("In 100 words, please give me information about United States", a.V, a.family)
        model = engine.Model.random(eng, a, seed=1, max_ctx=512)
        sess = model.session(512)
        sess.prefill(prompt)
        toks, lg = sess.decode(256, want_logits=True)
        res["tokens"] = toks
        res["logits_sub"] = lg[[0, 1, 127, 255]][:, ::97]
        sess.reset()
        sess.prefill(prompt)
        res["tokens_vp"] = sess.decode(256)       # no logits asked for: the LM head runs vocabulary-parallel in all steps but the last
        st = sess.stats()
        res["tok_per_s"] = np.array(st["eval_count"] / (st["eval_ns"] * 1e-9))
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), **res)
    dist.barrier()
    dist.destroy_process_group()


def run_case(tmp_path, case, world=2):
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(world, free_port(), case, str(tmp_path)), nprocs=world, join=True)
    return [np.load(os.path.join(str(tmp_path), f"rank{r}.npz")) for r in range(world)]


@needs2
@pytest.mark.timeout(600)
@pytest.mark.parametrize("key", ["llama3.1:8b", "qwen2:7b", "gemma:7b", "phi3:3.8b"])
def test_tp2_tiny_matches_oracle_and_ranks_agree(lib, tmp_path, key):
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import engine
    from oracle.model import OracleModel
    a, oa = tp_arch(key)
    r0, r1 = run_case(tmp_path, {"kind": "tiny", "key": key, "seed": 42, "n_new": 128})
    # every rank forms the same x in the same order and runs the whole LM head: bit-identical outputs
    for k in ("logits", "greedy", "logits_devinit", "prefill_logits", "sampled", "mixed", "mixed_logits"):
        assert np.array_equal(r0[k], r1[k]), k
    assert np.array_equal(r0["mixed"], r0["solo_mixed"])          # the tensor-parallel group and one GPU pick the same tokens
    assert np.abs(r0["mixed_logits"] - r0["solo_mixed_logits"]).max() <= 1e-4 * np.abs(r0["solo_mixed_logits"]).max()
    tensors = engine.host_random_tensors(a, seed=42)
    om = OracleModel(oa, tensors)
    os_ = om.session(320)
    ref = np.stack([os_.step(t).numpy() for t in PROMPT + [3, 4, 5, 6, 7, 8]])
    # same bar as the single-GPU engine (test_model_gpu.py): 1e-4 of the logit scale; the K split only reorders fp32 sums
    assert np.abs(r0["logits"] - ref).max() <= 1e-4 * np.abs(ref).max()
    assert np.array_equal(r0["logits"].argmax(1), ref.argmax(1))
    want = om.session(320).generate_greedy(PROMPT, 128)
    assert list(map(int, r0["greedy"])) == want
    assert np.array_equal(r0["logits_devinit"], r0["logits"][:len(PROMPT)])
    # the prompt as one batch (tcgen05 GEMM, partial sums of the row-parallel products exchanged per batch) against token by token: the
    # same terms in another order (bar of the single-GPU engine, test_model_gpu.py)
    want_last = r0["logits"][len(PROMPT) - 1]
    assert np.abs(r0["prefill_logits"] - want_last).max() <= 2e-4 * np.abs(want_last).max()


@needs2
@pytest.mark.timeout(900)
@pytest.mark.parametrize("key", ["llama3.1:8b", "mistral:7b", "qwen2:7b", "gemma:7b", "phi3:3.8b", "qwen2:1.5b"])
def test_tp2_full_size_greedy_256(lib, tmp_path, key):
    """The north-star bar on two GPUs, for every architecture whose KV heads split in two (gemma:2b has one): the 256 greedy
    tokens of the committed oracle fixture, logits within 1e-3 of the scale."""
    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "full_size.npz"))
    r0, r1 = run_case(tmp_path, {"kind": "full", "key": key})
    assert np.array_equal(r0["tokens"], r1["tokens"]) and np.array_equal(r0["logits_sub"], r1["logits_sub"])
    assert np.array_equal(r0["tokens"], G[key + "/tokens"])
    assert np.array_equal(r0["tokens_vp"], G[key + "/tokens"]) and np.array_equal(r1["tokens_vp"], G[key + "/tokens"])
    scale = G[key + "/absmax"][[0, 1, 127, 255]]
    assert np.abs(r0["logits_sub"] - G[key + "/logits_sub"]).max() <= 1e-3 * scale.max()
    print(f"tp2 {key} decode {float(r0['tok_per_s']):.1f} tokens/s")


@needs2
@pytest.mark.timeout(600)
def test_tp2_q4k_layers_q6k_head_match_oracle(lib, tmp_path):
    """Q4_K layer matrices and a Q6_K LM head split over two ranks (what config 4 of SURVEY.md 8d names): each rank packs its
    column window of attn_output / ffn_down from the host GGML super-blocks; logits and greedy tokens against the oracle."""
    from oracle import randinit as orand
    from oracle.model import OracleModel
    key = "llama3.1:8b"
    a, oa = tp_arch(key)
    assert a.H % 256 == 0 and (a.I // 2) % 256 == 0 and (a.nh // 2 * a.hd) % 256 == 0
    r0, r1 = run_case(tmp_path, {"kind": "tiny_kq", "key": key, "seed": 21})
    assert np.array_equal(r0["logits"], r1["logits"]) and np.array_equal(r0["greedy"], r1["greedy"])
    tensors = orand.make_model(oa, seed=21, qtype="q4_k", head_qtype="q6_k")
    om = OracleModel(oa, tensors)
    os_ = om.session(96)
    want = np.stack([os_.step(t).numpy() for t in PROMPT])
    assert np.abs(r0["logits"] - want).max() <= 2e-4 * np.abs(want).max()      # bar of the single-GPU K-quant test (test_q4k_gpu.py)
    assert list(map(int, r0["greedy"])) == om.session(96).generate_greedy(PROMPT, 40)


@needs2
@pytest.mark.timeout(900)
@pytest.mark.parametrize("no_pdl", [False, True])
def test_tp2_alternating_calls_for_a_thousand_passes(lib, tmp_path, no_pdl):
    """Long prompts (passes without the LM head), vocabulary-parallel greedy, sampled and replicated steps in turn, > 1 000 forward
    passes on one pair of engines: every token equals the single-GPU engine's on the same weights, on both ranks, and no wait on a
    peer ever timed out (a call would have failed with B2_ENCCL).  Also with programmatic launches off (B2_NO_PDL=1), which would
    tell an ordering problem of the launch chain from one of the exchange itself."""
    r0, r1 = run_case(tmp_path, {"kind": "stress", "key": "llama3.1:8b", "seed": 17, "rounds": 12, "no_pdl": no_pdl})
    assert int(r0["passes"]) >= 1000
    assert np.array_equal(r0["got"], r1["got"])
    assert np.array_equal(r0["got"], r0["want"]) and np.array_equal(r1["got"], r1["want"])


@needs2
@pytest.mark.timeout(300)
def test_tp2_missing_peer_is_an_error_not_a_hang(lib, tmp_path):
    """Every wait on a peer inside the kernels is bounded (2 s) and leaves a mark: the call of the rank that waited in vain fails
    with B2_ENCCL, its GPU stays usable (the process goes on to write its result and leave the barrier)."""
    r0, r1 = run_case(tmp_path, {"kind": "missing_peer", "key": "llama3.1:8b"})
    assert int(r0["code"]) == -6, int(r0["code"])          # B2_ENCCL
    assert 1.5 <= float(r0["seconds"]) <= 30.0, float(r0["seconds"])
    assert int(r1["code"]) == 0
