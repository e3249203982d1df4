"""HTTP contract of llmd (the drop-in for Ollama at OLLAMA_URL): replays exactly what the reference sends
(web/streamlit_app.py:91-95) and checks what it reads back (:96-98), plus the Ollama surface north_star adds."""
import json
import os
import socket
import subprocess
import threading
import time

import numpy as np
import pytest
import requests

import __graft_entry__ as ge
from tests.util import tiny_cfg

pytestmark = pytest.mark.gpu
LLMD = os.path.join(ge.PKG_DIR, "llmd")


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.fixture(scope="module")
def server():
    port = _free_port()
    env = dict(os.environ, LLMD_ADDR=f"127.0.0.1:{port}", LLMD_TINY="2", LLMD_MAX_CTX="512", LLMD_SEED="7", LLMD_MODEL="llama3.1")
    proc = subprocess.Popen([LLMD], env=env, stderr=subprocess.PIPE, text=True)
    url = f"http://127.0.0.1:{port}"
    for _ in range(600):
        if proc.poll() is not None:
            raise RuntimeError("llmd exited: " + proc.stderr.read())
        try:
            if requests.get(url + "/", timeout=1).status_code == 200:
                break
        except requests.RequestException:
            time.sleep(0.1)
    else:
        proc.kill()
        raise RuntimeError("llmd did not come up")
    yield url
    proc.terminate()
    proc.wait(timeout=10)


def test_reference_request_replay(server):
    """The exact call of web/streamlit_app.py:91-95 and the reads of :96-98."""
    pkg = ge.load_package()
    from importlib import import_module
    client = import_module("p2p_llm_chat_go_b200.client")
    body = client.suggestion_request("Hey! How's it going?")
    assert set(body) == {"model", "prompt", "stream"} and body["stream"] is False and body["model"] == "llama3.1"
    resp = requests.post(f"{server}/api/generate", json=body, timeout=60)
    assert resp.status_code == 200
    data = resp.json()
    assert isinstance(data.get("response"), str) and data["done"] is True and data["model"] == "llama3.1"
    for k in ("total_duration", "load_duration", "prompt_eval_count", "prompt_eval_duration", "eval_count", "eval_duration"):
        assert isinstance(data[k], int) and data[k] >= 0
    assert data["eval_count"] >= 1 and data["done_reason"] in ("stop", "length")
    os.environ["OLLAMA_URL"] = server
    out = client.ai_suggest("Hey! How's it going?")
    assert isinstance(out, str) and not out.startswith("(LLM")
    os.environ["LLM_MODEL"] = "no-such-model"
    try:
        assert client.ai_suggest("hi") == "(LLM error)"          # non-200 -> the reference's error string
    finally:
        os.environ["LLM_MODEL"] = "llama3.1"
    os.environ["OLLAMA_URL"] = "http://127.0.0.1:1"
    assert client.ai_suggest("hi", timeout=2).startswith("(LLM unavailable:")
    os.environ["OLLAMA_URL"] = server


def test_greedy_tokens_match_the_engine_in_process(server):
    pkg = ge.load_package()
    body = {"model": "llama3.1", "prompt": "abc é中", "raw": True, "stream": False,
            "options": {"temperature": 0, "num_predict": 12, "seed": 1}}
    data = requests.post(f"{server}/api/generate", json=body, timeout=60).json()
    prompt_ids = list("abc é中".encode())
    assert data["context"][:len(prompt_ids)] == prompt_ids and data["eval_count"] == 12 and data["done_reason"] == "length"
    cfg = tiny_cfg(pkg, pkg.WEIGHTS_BF16, n_layers=2, max_ctx=512, seed=7)
    with pkg.Engine(cfg) as eng:
        ids, _, _ = eng.generate_greedy(np.array(prompt_ids, np.int32), 12, pkg.MODE_GEMM)
        assert data["context"][len(prompt_ids):] == ids.tolist()
        assert data["response"] == eng.detokenize(ids)


def test_streaming_ndjson_equals_non_stream(server):
    body = {"model": "llama3.1", "prompt": "stream me", "options": {"temperature": 0, "num_predict": 10}}
    one = requests.post(f"{server}/api/generate", json=dict(body, stream=False), timeout=60).json()
    with requests.post(f"{server}/api/generate", json=body, stream=True, timeout=60) as r:   # stream defaults to true
        assert r.status_code == 200 and r.headers["Content-Type"].startswith("application/x-ndjson")
        objs = [json.loads(line) for line in r.iter_lines() if line]
    assert all(o["done"] is False for o in objs[:-1]) and objs[-1]["done"] is True
    assert "".join(o["response"] for o in objs) == one["response"]
    assert objs[-1]["eval_count"] == one["eval_count"] == 10


def test_chat_endpoint_and_misc_routes(server):
    r = requests.post(f"{server}/api/chat", json={"model": "llama3.1:latest", "stream": False, "options": {"num_predict": 5},
                                                  "messages": [{"role": "system", "content": "be brief"}, {"role": "user", "content": "hi"}]}, timeout=60)
    assert r.status_code == 200
    m = r.json()
    assert m["message"]["role"] == "assistant" and isinstance(m["message"]["content"], str) and m["done"] is True
    assert requests.get(server + "/").text == "Ollama is running"
    assert "version" in requests.get(server + "/api/version").json()
    assert requests.get(server + "/api/tags").json()["models"][0]["name"] == "llama3.1:latest"
    bad = requests.post(f"{server}/api/generate", data="{not json", timeout=10)
    assert bad.status_code == 400 and "error" in bad.json()
    nf = requests.post(f"{server}/api/generate", json={"model": "mistral", "prompt": "x"}, timeout=10)
    assert nf.status_code == 404 and "not found" in nf.json()["error"]


def test_concurrent_peers(server):
    """Several UI sessions clicking 'Suggest a reply' at once (one blocking request each)."""
    out = [None] * 8

    def peer(i):
        out[i] = requests.post(f"{server}/api/generate", json={"model": "llama3.1", "prompt": f"peer {i}", "stream": False,
                                                                "options": {"temperature": 0, "num_predict": 6}}, timeout=60)
    ts = [threading.Thread(target=peer, args=(i,)) for i in range(8)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert all(o is not None and o.status_code == 200 and o.json()["eval_count"] == 6 for o in out)
    same = requests.post(f"{server}/api/generate", json={"model": "llama3.1", "prompt": "peer 3", "stream": False,
                                                          "options": {"temperature": 0, "num_predict": 6}}, timeout=60).json()
    assert same["response"] == out[3].json()["response"]


def test_metrics_endpoint_counts_requests(server):
    """GET /metrics (Prometheus text format): the operational counters of SURVEY 8(f) rank 4."""
    def scrape():
        txt = requests.get(f"{server}/metrics", timeout=10).text
        return {ln.split()[0]: float(ln.split()[1]) for ln in txt.splitlines() if ln and not ln.startswith("#")}
    before = scrape()
    body = {"model": "llama3.1", "prompt": "hello", "stream": False, "options": {"temperature": 0, "num_predict": 5}}
    assert requests.post(f"{server}/api/generate", json=body, timeout=60).status_code == 200
    after = scrape()
    assert after["llmd_requests_total"] == before["llmd_requests_total"] + 1
    assert after["llmd_generated_tokens_total"] >= before["llmd_generated_tokens_total"] + 1
    assert after["llmd_prompt_tokens_total"] > before["llmd_prompt_tokens_total"]
    assert after["llmd_requests_in_flight"] == 0 and after["llmd_generation_seconds_total"] >= before["llmd_generation_seconds_total"]


def test_malformed_requests_do_not_take_the_server_down(server):
    """Untrusted network input: broken request lines, wrong lengths, invalid JSON, absurd options.  Each gets an error
    reply or a closed connection; afterwards the server still serves the reference's request."""
    host, port = server.replace("http://", "").split(":")
    raw_cases = [
        b"", b"\r\n\r\n", b"GARBAGE\r\n\r\n", b"POST /api/generate HTTP/1.1\r\nContent-Length: 999999999999\r\n\r\n{}",
        b"POST /api/generate HTTP/1.1\r\nContent-Length: -5\r\n\r\n{}", b"POST /api/generate HTTP/1.1\r\nContent-Length: 10\r\n\r\n{",
        b"POST /api/generate HTTP/1.1\r\n" + b"X: " + b"a" * 100000 + b"\r\n\r\n", b"GET /" + b"a" * 70000 + b" HTTP/1.1\r\n\r\n",
        b"POST /api/generate HTTP/1.1\r\nContent-Length: 7\r\n\r\n[[[[[[[", b"\x00\xff\xfe" * 100,
    ]
    for raw in raw_cases:
        s = socket.create_connection((host, int(port)), timeout=5)
        try:
            s.sendall(raw)
            s.settimeout(2)
            try:
                s.recv(4096)
            except (socket.timeout, ConnectionError):
                pass
        finally:
            s.close()
    bad_bodies = ["{", "[]", "null", '{"model":5}', '{"model":"llama3.1"}', '{"model":"llama3.1","prompt":5,"options":"x"}',
                  '{"model":"llama3.1","prompt":"x","options":{"num_predict":-7,"top_k":-1,"temperature":-3,"top_p":7,"repeat_last_n":99999999}}',
                  '{"model":"llama3.1","messages":"nope"}', '{"model":"llama3.1","prompt":"' + "z" * 5000 + '"}']
    for b in bad_bodies:
        for path in ("/api/generate", "/api/chat"):
            r = requests.post(server + path, data=b, headers={"Content-Type": "application/json"}, timeout=60)
            assert r.status_code in (200, 400, 404, 500), (b[:40], r.status_code)
    ok = requests.post(f"{server}/api/generate", json={"model": "llama3.1", "prompt": "still there?", "stream": False,
                                                        "options": {"num_predict": 4}}, timeout=60)
    assert ok.status_code == 200 and ok.json()["done"] is True
