"""HTTP boundary: the exact request the unchanged reference sends with curl, the Ollama
response shape, error conventions, streaming, and the stats side channel.  Uses a stub backend (no compute)."""
import json
import shlex
import socket
import subprocess
import urllib.request

import pytest

from cain_2025_device_remote_llm_energy_rep_pkg_b200 import server, tokenizer
from stub_backend import StubBackend

GOLDEN_BODY = ('{"model": "llama3.1:8b", "prompt": "In 100 words, please give me information about Prince Philip, '
               'Duke of Edinburgh", "stream": false}')


@pytest.fixture()
def srv():
    s = server.serve(StubBackend(), host="127.0.0.1", port=0, verbose=False)
    yield s
    s.shutdown()
    s.server_close()


def url(s, path):
    return f"http://127.0.0.1:{s.server_address[1]}{path}"


def test_reference_curl_command_line(srv):
    """The reference builds `curl http://host:11434/api/generate -d '{...}'` with shlex (RunnerConfig.py:128-131):
    curl -d posts the JSON with Content-Type application/x-www-form-urlencoded; the server must parse it anyway."""
    model, prompt = "llama3.1:8b", "In 100 words, please give me information about Prince Philip, Duke of Edinburgh"
    data = "'{" + f'"model": "{model}", "prompt": "{prompt}", "stream": false' + "}'"
    cmd = f'curl -s http://127.0.0.1:{srv.server_address[1]}/api/generate -d {data}'
    out = subprocess.run(shlex.split(cmd), capture_output=True, timeout=30)
    assert out.returncode == 0
    rec = json.loads(out.stdout)
    for k in ("model", "created_at", "response", "done", "done_reason", "context", "total_duration", "load_duration",
              "prompt_eval_count", "prompt_eval_duration", "eval_count", "eval_duration"):
        assert k in rec, k
    assert rec["model"] == model and rec["done"] is True
    assert rec["eval_count"] == 133 == len(rec["response"])          # ceil(1.33 * 100)
    assert srv.backend.requests[-1][1] == prompt


def test_golden_wire_request(srv):
    """Byte-level replay of the captured request: headers as curl 8.5.0 sends them."""
    body = GOLDEN_BODY.encode()
    assert len(body) == 134
    req = (b"POST /api/generate HTTP/1.1\r\nHost: 127.0.0.1\r\nUser-Agent: curl/8.5.0\r\nAccept: */*\r\n"
           b"Content-Length: 134\r\nContent-Type: application/x-www-form-urlencoded\r\n\r\n" + body)
    with socket.create_connection(("127.0.0.1", srv.server_address[1]), timeout=10) as c:
        c.sendall(req)
        buf = b""
        while b"\r\n\r\n" not in buf:
            buf += c.recv(65536)
        head, _, rest = buf.partition(b"\r\n\r\n")
        assert head.startswith(b"HTTP/1.1 200")
        n = int([l for l in head.split(b"\r\n") if l.lower().startswith(b"content-length")][0].split(b":")[1])
        while len(rest) < n:
            rest += c.recv(65536)
    rec = json.loads(rest[:n])
    assert b"application/json" in head.lower() and rec["done"] and rec["model"] == "llama3.1:8b"


def test_errors_follow_ollama_convention(srv):
    def post(body):
        r = urllib.request.Request(url(srv, "/api/generate"), data=body, method="POST")
        try:
            with urllib.request.urlopen(r, timeout=10) as resp:
                return resp.status, json.loads(resp.read())
        except urllib.error.HTTPError as e:
            return e.code, json.loads(e.read())
    assert post(b"not json")[0] == 400
    code, rec = post(json.dumps({"model": "nope:1b", "prompt": "x", "stream": False}).encode())
    assert code == 404 and "not found" in rec["error"]
    assert post(json.dumps({"prompt": "x"}).encode())[0] == 400
    assert post(b"[1,2]")[0] == 400


def test_num_predict_option_and_default_budget():
    assert server.default_num_predict("In 500 words, please give me information about India") == 665
    assert server.default_num_predict("In 1000 words, please give me information about X") == 1330
    assert server.default_num_predict("hello") == 128


def test_streaming_ndjson_and_aux_endpoints(srv):
    body = json.dumps({"model": "qwen2:1.5b", "prompt": "hi", "options": {"num_predict": 40}}).encode()   # stream defaults to true
    with urllib.request.urlopen(urllib.request.Request(url(srv, "/api/generate"), data=body, method="POST"), timeout=10) as r:
        lines = [json.loads(l) for l in r.read().decode().splitlines() if l.strip()]
    assert [l["done"] for l in lines] == [False, False, False, True]
    assert "".join(l["response"] for l in lines[:-1]) and lines[-1]["eval_count"] == 40
    with urllib.request.urlopen(url(srv, "/api/tags"), timeout=10) as r:
        tags = {m["name"] for m in json.loads(r.read())["models"]}
    assert tags == {"llama3.1:8b", "gemma:2b", "gemma:7b", "phi3:3.8b", "qwen2:1.5b", "qwen2:7b", "mistral:7b"}
    with urllib.request.urlopen(url(srv, "/"), timeout=10) as r:
        assert r.read() == b"Ollama is running"
    with urllib.request.urlopen(url(srv, "/b2/stats"), timeout=10) as r:
        recs = json.loads(r.read())["requests"]
    assert recs[-1]["model"] == "qwen2:1.5b" and recs[-1]["eval_count"] == 40 and recs[-1]["energy_mj"] == 1234


def test_tokenizer_stand_in():
    ids = tokenizer.encode("In 100 words", 32064, "phi3")
    assert ids[0] == tokenizer.BOS and all(0 <= i < 32064 for i in ids)
    assert tokenizer.encode("é", 1000, "", template=False) == [1, 3 + 0xC3, 3 + 0xA9]
    txt = tokenizer.decode(range(300))
    assert len(txt) == 300 and json.loads(json.dumps(txt)) == txt and all(32 <= ord(c) < 127 for c in txt)


def test_sampling_options_map_to_the_device_sampler():
    f = server.sampler_from_options
    assert f({}) is None                                            # no options: greedy (what parity pins)
    assert f({}, default_sampling=True) == (40, 0.8, 0.9, 0)        # --ollama-sampling: Ollama's stochastic defaults
    assert f({"temperature": 0}) is None                            # Ollama: temperature 0 = greedy
    assert f({"top_k": 1, "temperature": 0.7}) is None
    assert f({"seed": 42}) == (40, 0.8, 0.9, 42)
    assert f({"temperature": 1.5, "top_k": 500, "top_p": 2.0, "seed": -1}) == (64, 1.5, 1.0, 0xFFFFFFFFFFFFFFFF)
    with pytest.raises(ValueError):
        f({"temperature": "hot"})


def test_chat_endpoint(srv):
    body = json.dumps({"model": "qwen2:1.5b", "stream": False, "options": {"num_predict": 5},
                       "messages": [{"role": "system", "content": "be brief"}, {"role": "user", "content": "hello"}]}).encode()
    rec = json.loads(urllib.request.urlopen(urllib.request.Request(url(srv, "/api/chat"), data=body), timeout=30).read())
    assert rec["done"] is True and rec["message"]["role"] == "assistant" and isinstance(rec["message"]["content"], str)
    assert "response" not in rec and rec["eval_count"] == 5
    tag, prompt, _ = srv.backend.requests[-1]
    # the model family's chat form (qwen2: ChatML), not a made-up "role: content" flattening; sent to the backend as a raw prompt
    assert prompt == "<|im_start|>system\nbe brief<|im_end|>\n<|im_start|>user\nhello<|im_end|>\n<|im_start|>assistant\n"
    bad = urllib.request.Request(url(srv, "/api/chat"), data=json.dumps({"model": "qwen2:1.5b", "messages": []}).encode())
    with pytest.raises(urllib.error.HTTPError) as e:
        urllib.request.urlopen(bad, timeout=30)
    assert e.value.code == 400 and "error" in json.loads(e.value.read())
    # streaming chat: NDJSON lines carry message objects
    body = json.dumps({"model": "qwen2:1.5b", "options": {"num_predict": 20}, "messages": [{"role": "user", "content": "hi"}]}).encode()
    lines = [json.loads(l) for l in urllib.request.urlopen(urllib.request.Request(url(srv, "/api/chat"), data=body), timeout=30).read().splitlines()]
    assert lines[-1]["done"] is True and all("message" in l for l in lines) and len(lines) >= 3


def test_chat_forms_of_the_five_families():
    msgs = [{"role": "system", "content": "S"}, {"role": "user", "content": "U"}, {"role": "assistant", "content": "A"}, {"role": "user", "content": "Q"}]
    f = server.chat_prompt
    assert f(msgs, "llama") == ("<|start_header_id|>system<|end_header_id|>\n\nS<|eot_id|><|start_header_id|>user<|end_header_id|>\n\nU<|eot_id|>"
                                "<|start_header_id|>assistant<|end_header_id|>\n\nA<|eot_id|><|start_header_id|>user<|end_header_id|>\n\nQ<|eot_id|>"
                                "<|start_header_id|>assistant<|end_header_id|>\n\n")
    assert f(msgs[1:2], "gemma") == "<start_of_turn>user\nU<end_of_turn>\n<start_of_turn>model\n"
    assert f(msgs[1:3], "gemma").endswith("<start_of_turn>model\nA<end_of_turn>\n<start_of_turn>model\n")
    assert f(msgs[1:2], "phi3") == "<|user|>\nU<|end|>\n<|assistant|>\n"
    assert f(msgs[1:2], "mistral") == "[INST] U [/INST]"
    # a single user turn renders exactly as /api/generate's template for the family
    for fam in ("llama", "qwen2", "gemma", "phi3", "mistral"):
        assert f([{"role": "user", "content": "hello"}], fam) == tokenizer.TEMPLATES[fam].format(prompt="hello")
    # the GGUF file's own tokenizer.chat_template decides when present (its marker tokens), whatever the architecture says
    assert server.chat_family("llama", "{% for m in messages %}<|im_start|>{{ m.role }}...") == "qwen2"
    assert server.chat_family("llama", "") == "llama"
    assert f(msgs[1:2], "somethingelse") == "user: U\nassistant:"


def test_repeat_options():
    f = server.repeat_from_options
    assert f({}) == (1.0, 64) and f({}, default_sampling=True) == (1.1, 64)          # Ollama's defaults only with --ollama-sampling
    assert f({"repeat_penalty": 1.3, "repeat_last_n": 16}) == (1.3, 16)
    assert f({"repeat_penalty": 1.3, "repeat_last_n": 0}) == (1.0, 0) and f({"repeat_penalty": 1.2, "repeat_last_n": -1})[1] > 4096
    with pytest.raises(ValueError):
        f({"repeat_penalty": "strong"})


def test_keep_alive_and_devices_parsing():
    assert server.parse_keep_alive("5m", 1.0) == 300.0 and server.parse_keep_alive(0, 1.0) == 0.0 and server.parse_keep_alive("-1", 1.0) == -1.0
    assert server.parse_keep_alive(None, 7.0) == 7.0 and server.parse_keep_alive("1h", 0) == 3600.0 and server.parse_keep_alive("250ms", 0) == 0.25
    with pytest.raises(ValueError):
        server.parse_keep_alive("soon", 1.0)
    assert server.parse_devices("0-3,6") == [0, 1, 2, 3, 6] and server.parse_devices("2") == [2]
    with pytest.raises(ValueError):
        server.parse_devices("1,1")


def test_two_concurrent_requests_land_on_different_backends():
    """SURVEY.md §8e / VERDICT r1 #3: K runner processes all call :11434 (RunnerConfig.py:130); the one server owns a backend per
    GPU and spreads concurrent requests, preferring an idle backend that already holds the model."""
    import threading
    b0, b1 = StubBackend(delay_s=0.5), StubBackend(delay_s=0.5)
    b0.device, b1.device = 0, 1
    s = server.serve(server.MultiBackend([b0, b1]), host="127.0.0.1", port=0, verbose=False)
    try:
        def post(model, n):
            body = json.dumps({"model": model, "prompt": f"In {n} words, please give me information about X", "stream": False}).encode()
            return json.loads(urllib.request.urlopen(urllib.request.Request(url(s, "/api/generate"), data=body), timeout=30).read())
        out = [None, None]
        ts = [threading.Thread(target=lambda i=i: out.__setitem__(i, post("gemma:2b", 100 + i))) for i in range(2)]
        for t in ts:
            t.start()
        import time as _t
        _t.sleep(0.25)
        busy = json.loads(urllib.request.urlopen(url(s, "/b2/active"), timeout=5).read())["devices"]
        for t in ts:
            t.join()
        assert sorted(busy) == [0, 1]
        assert len(b0.requests) == 1 and len(b1.requests) == 1 and all(o["done"] for o in out)
        # a lone request for a model goes to the backend that already holds it; a new model to the one with fewer residents
        post("gemma:2b", 100)
        post("gemma:2b", 100)
        assert len(b0.requests) == 3 and len(b1.requests) == 1
        post("phi3:3.8b", 100)
        assert b0.requests[-1][0] == "phi3:3.8b" and len(b1.requests) == 1
        post("phi3:3.8b", 100)                                         # ... and stays where it is resident
        assert len(b0.requests) == 5 and len(b1.requests) == 1
        recs = json.loads(urllib.request.urlopen(url(s, "/b2/stats"), timeout=5).read())["requests"]
        assert {r["device"] for r in recs} == {0, 1} and all("prompt" in r for r in recs)
    finally:
        s.shutdown()
        s.server_close()


def test_error_in_mid_stream_ends_the_chunked_answer_cleanly():
    class Failing(StubBackend):
        def generate(self, tag, prompt, options, raw=False):
            yield {"tokens": [1, 2, 3]}
            raise RuntimeError("engine fell over")
    s = server.serve(Failing(), host="127.0.0.1", port=0, verbose=False)
    try:
        body = json.dumps({"model": "gemma:2b", "prompt": "hi"}).encode()
        data = urllib.request.urlopen(urllib.request.Request(url(s, "/api/generate"), data=body), timeout=30).read()
        lines = [json.loads(l) for l in data.splitlines()]            # every line is valid JSON: no second status line in the stream
        assert lines[0]["done"] is False and "engine fell over" in lines[-1]["error"]
    finally:
        s.shutdown()
        s.server_close()


def test_tags_lists_a_model_outside_the_seven_architectures():
    class WithGguf(StubBackend):
        def list_models(self):
            return super().list_models() + ["custom:latest"]
    s = server.serve(WithGguf(), host="127.0.0.1", port=0, verbose=False)
    try:
        tags = json.loads(urllib.request.urlopen(url(s, "/api/tags"), timeout=5).read())["models"]
        assert {"custom:latest", "llama3.1:8b"} <= {m["name"] for m in tags}
        assert [m for m in tags if m["name"] == "custom:latest"][0]["details"]["family"] == "unknown"
    finally:
        s.shutdown()
        s.server_close()
