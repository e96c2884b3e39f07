"""HTTP -> B2Backend -> CUDA (SURVEY.md §8b.1 on the real engine): the request the unchanged reference sends with curl
(experiment/RunnerConfig.py:128-131) answered by the B200 engine behind the Ollama-compatible server, tokens checked against
the committed full-size oracle fixture (tests/golden/make_full_golden.py).  Also: the stats side channel names the device,
streaming returns the same tokens, num_ctx / keep_alive behave."""
import json
import os
import urllib.request

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu
GOLD = os.path.join(ROOT, "tests", "golden", "full_size.npz")
TAG = "qwen2:1.5b"
PROMPT = "In 100 words, please give me information about United States"   # RunnerConfig.py:120 + topics.csv:2, as the fixture


@pytest.fixture(scope="module")
def srv(lib):
    from cain_2025_device_remote_llm_energy_rep_pkg_b200 import server
    s = server.serve(server.B2Backend(device=0, seed=1), host="127.0.0.1", port=0, verbose=False)
    yield s
    s.shutdown()
    s.server_close()


def post(s, body: dict):
    req = urllib.request.Request(f"http://127.0.0.1:{s.server_address[1]}/api/generate", data=json.dumps(body).encode(),
                                 headers={"Content-Type": "application/x-www-form-urlencoded"})   # what curl -d sends
    return urllib.request.urlopen(req, timeout=300).read()


def test_reference_request_returns_the_oracle_tokens(srv):
    gold = np.load(GOLD)
    rec = json.loads(post(srv, {"model": TAG, "prompt": PROMPT, "stream": False}))
    n_prompt = rec["prompt_eval_count"]
    assert rec["done"] is True and rec["done_reason"] == "length" and rec["eval_count"] == 133      # ceil(1.33 * 100 words)
    assert rec["context"][:n_prompt] == gold[TAG + "/prompt"].tolist()
    assert rec["context"][n_prompt:] == gold[TAG + "/tokens"][:133].tolist()                          # greedy, identical to the CPU oracle
    assert rec["eval_duration"] > 0 and rec["prompt_eval_duration"] > 0 and rec["total_duration"] >= rec["eval_duration"]
    stats = json.loads(urllib.request.urlopen(f"http://127.0.0.1:{srv.server_address[1]}/b2/stats", timeout=5).read())["requests"][-1]
    assert stats["model"] == TAG and stats["device"] == 0 and stats["eval_count"] == 133 and stats["prompt"] == PROMPT
    assert stats["energy_mj"] is None or stats["energy_mj"] >= 0


def test_streaming_returns_the_same_tokens_in_chunks(srv):
    gold = np.load(GOLD)
    lines = [json.loads(l) for l in post(srv, {"model": TAG, "prompt": PROMPT, "options": {"num_predict": 40}}).splitlines()]
    assert len(lines) >= 3 and lines[-1]["done"] is True and all(not l["done"] for l in lines[:-1])
    n_prompt = lines[-1]["prompt_eval_count"]
    assert lines[-1]["context"][n_prompt:] == gold[TAG + "/tokens"][:40].tolist()
    assert "".join(l["response"] for l in lines[:-1]) != ""


def test_num_ctx_and_keep_alive(srv):
    n_prompt = json.loads(post(srv, {"model": TAG, "prompt": PROMPT, "stream": False, "options": {"num_predict": 1}}))["prompt_eval_count"]
    rec = json.loads(post(srv, {"model": TAG, "prompt": PROMPT, "stream": False, "options": {"num_ctx": n_prompt + 7}}))
    assert rec["eval_count"] == 7 and rec["done_reason"] == "length"                                  # the window ends the generation
    assert TAG in srv.backend.loaded()
    rec = json.loads(post(srv, {"model": TAG, "prompt": PROMPT, "stream": False, "keep_alive": 0, "options": {"num_predict": 3}}))
    assert rec["eval_count"] == 3 and TAG not in srv.backend.loaded()                                 # keep_alive 0: unloaded at once
    rec = json.loads(post(srv, {"model": TAG, "prompt": PROMPT, "stream": False, "options": {"num_predict": 3}}))
    assert rec["load_duration"] > 0 and TAG in srv.backend.loaded()                                   # ... and loaded again on demand
