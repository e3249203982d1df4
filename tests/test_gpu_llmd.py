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
    from tests import ref_client as client
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


# ---- Ollama request fields beyond what the reference sends ------------------------------
def _gen(server, **kw):
    body = {"model": "llama3.1", "prompt": "tell me something", "stream": False, "options": {"temperature": 0, "num_predict": 24}}
    opts = kw.pop("options", {})
    body.update(kw)
    body["options"] = dict(body["options"], **opts)
    return requests.post(f"{server}/api/generate", json=body, timeout=60)


def test_stop_sequences_cut_the_reply(server):
    base = _gen(server).json()
    text = base["response"]
    assert base["done_reason"] == "length" and len(text) > 12
    stop = text[7:11]
    cut = text.index(stop)                                 # the first place that sequence occurs
    for stop_opt in (stop, ["\x00never", stop]):
        one = _gen(server, options={"stop": stop_opt}).json()
        assert one["response"] == text[:cut] and one["done_reason"] == "stop" and one["eval_count"] <= base["eval_count"]
    with requests.post(f"{server}/api/generate", json={"model": "llama3.1", "prompt": "tell me something",
                       "options": {"temperature": 0, "num_predict": 24, "stop": [stop]}}, stream=True, timeout=60) as r:
        objs = [json.loads(line) for line in r.iter_lines() if line]
    assert "".join(o["response"] for o in objs) == text[:cut] and objs[-1]["done_reason"] == "stop"
    assert _gen(server, options={"stop": 5}).status_code == 400


def test_num_ctx_limits_window_and_is_validated(server):
    assert _gen(server, options={"num_ctx": 100000}).status_code == 400        # beyond LLMD_MAX_CTX=512: refused, not ignored
    assert _gen(server, options={"num_ctx": "big"}).status_code == 400
    small = _gen(server, prompt="window test", options={"num_ctx": 64, "num_predict": -1}).json()   # until the 64-token window is full
    assert small["done_reason"] == "length" and len(small["context"]) == 64 + 1        # 64 in the cache + the last token
    # a prompt longer than the window is cut from the front (Ollama's behaviour), never an error
    long = _gen(server, prompt="x" * 400, options={"num_ctx": 64, "num_predict": 8}).json()
    assert long["done"] is True and long["eval_count"] == 8 and long["prompt_eval_count"] <= 64 - 8


def test_context_field_continues_a_conversation(server):
    first = _gen(server, prompt="first message").json()
    ctx = first["context"]
    second = _gen(server, prompt="and then?", context=ctx).json()
    assert second["context"][:len(ctx)] == ctx and len(second["context"]) > len(ctx) + second["eval_count"] - 1
    # the earlier turns were still in a slot's KV cache: only the new turn was evaluated
    assert second["prompt_eval_count"] < len(second["context"]) - second["eval_count"] - len(ctx) + 8
    assert _gen(server, context=[1, 2, 10 ** 9]).status_code == 400
    assert _gen(server, context="nope").status_code == 400


def test_template_field_and_its_limits(server):
    a = _gen(server, prompt="P", system="S", template="{{ if .System }}sys={{ .System }};{{ end }}q={{ .Prompt }};a={{ .Response }}ignored").json()
    b = _gen(server, prompt="sys=S;q=P;a=", raw=True).json()                   # the same text, passed raw
    assert a["response"] == b["response"] and a["prompt_eval_count"] + 0 >= 1
    assert a["context"][:12] == list(b"sys=S;q=P;a=")
    c = _gen(server, prompt="P", template="{{- if .System }}never{{ end -}}  q={{ .Prompt }}").json()
    assert c["context"][:3] == list(b"q=P")
    r = _gen(server, template="{{ range .Messages }}x{{ end }}")
    assert r.status_code == 400 and "unsupported" in r.json()["error"]
    assert _gen(server, template="{{ if .System }}open").status_code == 400


def test_keep_alive_format_and_ps(server):
    assert _gen(server, keep_alive="10m").status_code == 200 and _gen(server, keep_alive=-1).status_code == 200
    assert _gen(server, keep_alive="soon").status_code == 400 and _gen(server, keep_alive=[1]).status_code == 400
    load = requests.post(f"{server}/api/generate", json={"model": "llama3.1", "keep_alive": "1h"}, timeout=10).json()
    assert load["done"] is True and load["done_reason"] == "load" and load["response"] == ""
    unload = requests.post(f"{server}/api/generate", json={"model": "llama3.1", "keep_alive": 0}, timeout=10).json()
    assert unload["done_reason"] == "unload"
    assert _gen(server).status_code == 200                                     # ... and the model is of course still served
    ps = requests.get(f"{server}/api/ps", timeout=10).json()["models"][0]
    assert ps["name"] == "llama3.1:latest" and "expires_at" in ps and ps["size_vram"] > 0
    fmt = _gen(server, format="json")
    assert fmt.status_code == 400 and "format" in fmt.json()["error"]
    assert _gen(server, format="").status_code == 200


def test_suggest_endpoint_is_the_reference_call_behind_one_route(server):
    """POST /suggest {"content"} = what a node-side handler next to /send would forward.  It must give
    exactly what the reference's ai_suggest() computes from /api/generate (web/streamlit_app.py:89-98): the same
    prompt template, the stripped response — blocking and streamed."""
    from tests import ref_client as client
    msg = "Are we still on for lunch tomorrow?"
    opts = {"temperature": 0}
    want = requests.post(f"{server}/api/generate", json=dict(client.suggestion_request(msg), options=opts), timeout=60).json()["response"].strip()
    # greedy on both routes (llmd's defaults sample, as Ollama's do): the same text
    g = requests.post(f"{server}/suggest", json={"content": msg, "options": opts}, timeout=60)
    assert g.status_code == 200 and g.json()["suggestion"] == want
    # defaults, as the node-side handler would call it (sampled, as Ollama's defaults are: only the shape is checked)
    a = requests.post(f"{server}/suggest", json={"content": msg}, timeout=60)
    assert a.status_code == 200 and isinstance(a.json()["suggestion"], str) and a.json()["done"] is True
    assert "eval_count" in a.json() and "total_duration" in a.json()
    # streamed (greedy, so the text is the one above): one JSON object per line, the deltas concatenate to the final
    # suggestion up to the strip
    r = requests.post(f"{server}/suggest", json={"content": msg, "stream": True, "options": opts}, timeout=60, stream=True)
    assert r.status_code == 200 and r.headers["Content-Type"].startswith("application/x-ndjson")
    lines = [json.loads(l) for l in r.iter_lines() if l]
    assert lines[-1]["done"] is True and all(not l["done"] for l in lines[:-1])
    assert "".join(l["delta"] for l in lines[:-1]).strip() == lines[-1]["suggestion"] == want
    # malformed bodies
    assert requests.post(f"{server}/suggest", json={"text": "x"}, timeout=10).status_code == 400
    assert requests.post(f"{server}/suggest", data="{", timeout=10).status_code == 400


def test_one_url_spreads_sessions_over_several_engines():
    """the reference has ONE OLLAMA_URL (web/streamlit_app.py:27); llmd with LLMD_GPUS owns an engine per
    listed device behind it.  Two engines (both on GPU 0 here: the test box has one): concurrent requests land on
    both, replies equal a single-engine server's, and a peer's follow-up goes where its history is cached."""
    def start(gpus):
        port = _free_port()
        env = dict(os.environ, LLMD_ADDR=f"127.0.0.1:{port}", LLMD_TINY="2", LLMD_MAX_CTX="512", LLMD_SEED="7", LLMD_GPUS=gpus, LLMD_MAX_SEQS="8")
        proc = subprocess.Popen([LLMD], env=env, stderr=subprocess.PIPE, text=True)
        url = f"http://127.0.0.1:{port}"
        for _ in range(600):
            if proc.poll() is not None:
                raise RuntimeError("llmd exited: " + proc.stderr.read())
            try:
                if requests.get(url + "/", timeout=1).status_code == 200:
                    return proc, url
            except requests.RequestException:
                time.sleep(0.1)
        proc.kill()
        raise RuntimeError("llmd did not come up")

    def metrics(url):
        out = {}
        for ln in requests.get(url + "/metrics", timeout=10).text.splitlines():
            if ln and not ln.startswith("#"):
                k, v = ln.rsplit(" ", 1)
                out[k] = float(v)
        return out
    one_p, one = start("0")
    two_p, two = start("0,0")
    try:
        prompts = [f"peer {i} says hello to everyone in the room" for i in range(12)]
        replies = {}

        def peer(url, i):
            replies[(url, i)] = requests.post(f"{url}/api/generate", json={"model": "llama3.1", "prompt": prompts[i], "stream": False,
                                              "options": {"temperature": 0, "num_predict": 8}}, timeout=120).json()
        for url in (one, two):
            ts = [threading.Thread(target=peer, args=(url, i)) for i in range(12)]
            [t.start() for t in ts]
            [t.join() for t in ts]
        m = metrics(two)
        assert m["llmd_gpus"] == 2
        # both engines report device 0 here: the per-GPU line appears once per engine
        txt = requests.get(two + "/metrics", timeout=10).text
        counts = [float(ln.rsplit(" ", 1)[1]) for ln in txt.splitlines() if ln.startswith("llmd_gpu_requests_total{")]
        assert len(counts) == 2 and sum(counts) == 12 and min(counts) >= 3, counts       # spread, not piled on one engine
        same = sum(replies[(one, i)]["response"] == replies[(two, i)]["response"] for i in range(12))
        assert same >= 10, same     # batch composition differs between the servers; a near-tie may flip at most rarely
        # follow-up of peer 5 (history + reply + new text): evaluated where the history is cached -> few new tokens
        r5 = replies[(two, 5)]
        follow = requests.post(f"{two}/api/generate", json={"model": "llama3.1", "prompt": "and what next?", "context": r5["context"], "stream": False,
                               "options": {"temperature": 0, "num_predict": 4}}, timeout=60).json()
        new_turn = len(follow["context"]) - follow["eval_count"] - len(r5["context"])
        assert follow["prompt_eval_count"] <= new_turn + 1, (follow["prompt_eval_count"], new_turn, len(r5["context"]))
    finally:
        for p in (one_p, two_p):
            p.terminate()
            p.wait(timeout=10)
