"""256 UI sessions clicking 'Suggest a reply' at once, over HTTP against llmd (BASELINE configs[3] through the
boundary the reference uses).  python scratch/http_peers.py [n_peers]"""
import json, os, socket, subprocess, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import requests
import __graft_entry__ as ge

n_peers = int(sys.argv[1]) if len(sys.argv) > 1 else 256
s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
env = dict(os.environ, LLMD_ADDR=f"127.0.0.1:{port}", LLMD_MAX_CTX="640", LLMD_MAX_SEQS=str(n_peers), LLMD_EOS="-1")
proc = subprocess.Popen([os.path.join(ge.PKG_DIR, "llmd")], env=env, stderr=subprocess.DEVNULL)
url = f"http://127.0.0.1:{port}"
try:
    for _ in range(1200):
        try:
            if requests.get(url + "/", timeout=1).status_code == 200:
                break
        except requests.RequestException:
            time.sleep(0.1)
    text = ("userA: are we still on for friday? userB: yes, 7pm at the usual place. " * 8)[:511]
    def body(i):
        return {"model": "llama3.1", "prompt": text[:-3] + f"{i:03d}", "raw": True, "stream": False,
                "options": {"temperature": 0, "num_predict": 128}}
    requests.post(url + "/api/generate", json=body(999), timeout=120)  # warm-up
    out = [None] * n_peers
    def peer(i):
        t0 = time.perf_counter()
        r = requests.post(url + "/api/generate", json=body(i), timeout=120)
        out[i] = (r.status_code, r.json().get("eval_count", 0) if r.status_code == 200 else 0, time.perf_counter() - t0)
    ts = [threading.Thread(target=peer, args=(i,)) for i in range(n_peers)]
    t0 = time.perf_counter()
    [t.start() for t in ts]
    [t.join() for t in ts]
    wall = time.perf_counter() - t0
    ok = sum(1 for o in out if o and o[0] == 200)
    toks = sum(o[1] for o in out if o)
    lat = sorted(o[2] for o in out if o)
    print(json.dumps({"peers": n_peers, "ok": ok, "tokens": toks, "wall_s": wall, "aggregate_tok_per_s": toks / wall,
                      "latency_s_p50": lat[len(lat) // 2], "latency_s_max": lat[-1],
                      "metrics": requests.get(url + "/metrics", timeout=5).text.split("\n")[1::2][:7]}))
finally:
    proc.terminate()
