"""oracle/serve.py: the /api/generate route over the CPU oracle that bench.py's CPU baseline goes through
("reported via the same /api/generate route").  Runs without a GPU."""
import os

import numpy as np
import requests

from oracle import serve
from tests.util import tiny_from_hf_fixture

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "hf_tiny_llama.npz")


def _post(url, ids, n, **opts):
    return requests.post(url + "/api/generate", timeout=120, json={
        "model": "llama3.1", "prompt": "", "context": ids, "stream": False,
        "options": dict({"temperature": 0, "num_predict": n, "seed": 0}, **opts)})


def test_route_serves_the_oracles_greedy_tokens_with_ollamas_reply_shape():
    z = np.load(GOLDEN, allow_pickle=True)
    model, cfg = tiny_from_hf_fixture(z, kv_bf16=1)
    twin, _ = tiny_from_hf_fixture(z, kv_bf16=1)
    prompt = [int(t) for t in np.random.default_rng(3).integers(0, cfg.vocab, 12)]
    want, _ = twin.generate_greedy(np.asarray(prompt, np.int32), 6, 1)
    srv = serve.OracleServer(model, cfg.vocab)
    try:
        assert requests.get(srv.url + "/", timeout=10).text == "Ollama is running"
        r = _post(srv.url, prompt, 6)
        assert r.status_code == 200
        d = r.json()
        for k in ("model", "created_at", "response", "done", "done_reason", "context", "total_duration", "load_duration",
                  "prompt_eval_count", "prompt_eval_duration", "eval_count", "eval_duration"):
            assert k in d, k
        assert d["done"] is True and d["eval_count"] == 6 and d["prompt_eval_count"] == len(prompt)
        assert d["context"][:len(prompt)] == prompt and d["context"][len(prompt):] == [int(t) for t in want]
        assert d["total_duration"] >= d["prompt_eval_duration"] + d["eval_duration"] > 0
        # the same conversation again: found cached, only the last prompt token is evaluated (Ollama's prompt cache)
        d2 = _post(srv.url, prompt, 3).json()
        assert d2["prompt_eval_count"] == 1 and d2["eval_count"] == 3
        # a continuation of the reply: the common prefix is kept
        d3 = _post(srv.url, d["context"][:len(prompt) + 2], 2).json()
        assert d3["prompt_eval_count"] <= 3
        # refusals: sampling, text prompts, unknown model, bad ids
        assert _post(srv.url, prompt, 2, temperature=0.8).status_code == 400
        assert requests.post(srv.url + "/api/generate", json={"model": "llama3.1", "prompt": "hi", "options": {"temperature": 0}},
                             timeout=10).status_code == 400
        assert requests.post(srv.url + "/api/generate", json={"model": "other", "context": prompt, "options": {"temperature": 0}},
                             timeout=10).status_code == 404
        assert _post(srv.url, [cfg.vocab + 5], 2).status_code == 400
        assert requests.post(srv.url + "/api/generate", data="{", timeout=10).status_code == 400
    finally:
        srv.close()
