"""`POST /api/generate` over the CPU oracle — TEST INFRASTRUCTURE ONLY (see oracle.cpp header).

SURVEY.md 8(d) config 1 wants the CPU baseline "reported via the same /api/generate route" the reference calls
(web/streamlit_app.py:91-95).  The product's server (`llmd`) has no CPU path by design, so this is the route served by
the oracle instead: bench.py's `cpu_baseline` leg and `--impl reference` arm post to it and read Ollama's own timing
fields.  Nothing under p2p-llm-chat-go_b200/ imports this file.

Scope: one model, greedy only (`options.temperature` must be 0), token ids in and out — the oracle has no tokenizer,
so the conversation travels in Ollama's `context` field (the token ids so far) and `response` stays empty; the reply
carries `context` (ids incl. the generated ones) and `prompt_eval_count / prompt_eval_duration / eval_count /
eval_duration / total_duration` in ns, as Ollama does.  Like Ollama's prompt cache, a request whose ids start with what
the model already holds evaluates only the rest (so a bounded decode sample does not repeat a 36 s CPU prefill).
"""
import json
import threading
import time
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

import numpy as np


class OracleServer:
    def __init__(self, model, vocab, name="llama3.1", port=0):
        self.model, self.vocab, self.name = model, vocab, name
        self.held = []            # token ids whose K/V the oracle model holds
        self.lock = threading.Lock()
        outer = self

        class Handler(BaseHTTPRequestHandler):
            def log_message(self, *a):  # quiet
                pass

            def _send(self, code, obj):
                body = json.dumps(obj).encode()
                self.send_response(code)
                self.send_header("Content-Type", "application/json")
                self.send_header("Content-Length", str(len(body)))
                self.end_headers()
                self.wfile.write(body)

            def do_GET(self):
                if self.path == "/":
                    self.send_response(200)
                    self.send_header("Content-Type", "text/plain")
                    self.end_headers()
                    self.wfile.write(b"Ollama is running")
                else:
                    self._send(404, {"error": "not found"})

            def do_POST(self):
                if self.path != "/api/generate":
                    return self._send(404, {"error": "not found"})
                try:
                    req = json.loads(self.rfile.read(int(self.headers.get("Content-Length", "0"))))
                except (ValueError, TypeError):
                    return self._send(400, {"error": "invalid JSON body"})
                code, obj = outer.generate(req)
                self._send(code, obj)

        self.httpd = ThreadingHTTPServer(("127.0.0.1", port), Handler)
        self.url = f"http://127.0.0.1:{self.httpd.server_address[1]}"
        self.thread = threading.Thread(target=self.httpd.serve_forever, daemon=True)
        self.thread.start()

    def close(self):
        self.httpd.shutdown()
        self.httpd.server_close()

    def generate(self, req):
        if req.get("model") not in (self.name, self.name + ":latest"):
            return 404, {"error": f"model '{req.get('model')}' not found"}
        opts = req.get("options") or {}
        if opts.get("temperature", 0.8) != 0:
            return 400, {"error": "the oracle route is greedy only: options.temperature must be 0"}
        ids = req.get("context") or []
        if req.get("prompt") or not ids or not all(isinstance(t, int) and 0 <= t < self.vocab for t in ids):
            return 400, {"error": "the oracle route takes token ids in `context` and an empty `prompt` (it has no tokenizer)"}
        n_new = int(opts.get("num_predict", 128))
        with self.lock:
            t0 = time.perf_counter()
            keep = 0
            while keep < min(len(self.held), len(ids) - 1) and self.held[keep] == ids[keep]:
                keep += 1
            self.model.truncate(keep)
            rest = ids[keep:]
            # prompt tokens: tensor-core numerics (mode 1) for a real prefill, the decode path's for a one-token tail
            lg = self.model.forward(rest, mode=1 if len(rest) > 1 else 0)
            t1 = time.perf_counter()
            out = [int(np.argmax(lg))]
            held = list(ids)
            for _ in range(n_new - 1):
                lg = self.model.forward([out[-1]], mode=0)
                held.append(out[-1])
                out.append(int(np.argmax(lg)))
            t2 = time.perf_counter()
            self.held = held
        ns = lambda s: int(s * 1e9)
        return 200, {"model": self.name, "created_at": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime()), "response": "",
                     "done": True, "done_reason": "length", "context": list(ids) + out,
                     "total_duration": ns(t2 - t0), "load_duration": 0, "prompt_eval_count": len(rest),
                     "prompt_eval_duration": ns(t1 - t0), "eval_count": n_new, "eval_duration": ns(t2 - t1)}
