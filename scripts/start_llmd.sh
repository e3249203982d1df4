#!/usr/bin/env bash
# Starts llmd (the B200 drop-in for the Ollama container of the reference's README.md:62-70) the way the
# reference's start_all.sh:26-40 starts its own processes: background, log file, pid file.  ONE process and ONE
# address for every GPU listed — the reference knows a single OLLAMA_URL (web/streamlit_app.py:27); llmd keeps an
# engine per GPU and routes each request to the least loaded one (cache-aware, see csrc/llmd.cpp).
#   GPUS="0 1 2 3" MODEL_GGUF=/models/llama3.1.gguf scripts/start_llmd.sh   # real weights, four GPUs behind one URL
#   WEIGHTS=q4_k_m scripts/start_llmd.sh                                     # random-init 8B, Ollama's default mix
set -euo pipefail
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
BIN="$ROOT/p2p-llm-chat-go_b200/llmd"
[ -x "$BIN" ] || (cd "$ROOT" && python -c "import __graft_entry__ as g; g.build()")
GPUS="${GPUS:-0}"
PORT="${PORT:-11434}"          # web/streamlit_app.py:27 defaults OLLAMA_URL to this port
LOGDIR="${LOGDIR:-$ROOT/logs}"
mkdir -p "$LOGDIR"
LLMD_ADDR="127.0.0.1:$PORT" LLMD_GPUS="$(echo $GPUS | tr ' ' ',')" LLMD_MODEL="${LLM_MODEL:-llama3.1}" LLMD_GGUF="${MODEL_GGUF:-}" \
LLMD_WEIGHTS="${WEIGHTS:-bf16}" LLMD_MAX_CTX="${MAX_CTX:-4096}" LLMD_MAX_SEQS="${MAX_SEQS:-64}" \
  nohup "$BIN" > "$LOGDIR/llmd.log" 2>&1 &
echo $! > "$LOGDIR/llmd.pid"
echo "llmd on GPU(s) $GPUS: http://127.0.0.1:$PORT  (log $LOGDIR/llmd.log, Prometheus counters at /metrics, per GPU)"
echo "export OLLAMA_URL=http://127.0.0.1:$PORT   # then start the reference UI unchanged"
