#!/usr/bin/env bash
# Starts llmd (the B200 drop-in for the Ollama container of the reference's README.md:62-70) the way the
# reference's start_all.sh:26-40 starts its own processes: background, log file, pid file.  One engine per GPU;
# a front-end that spreads sessions over the ports is outside this script (replicas are independent).
#   GPUS="0 1" MODEL_GGUF=/models/llama3.1.gguf scripts/start_llmd.sh       # real weights
#   WEIGHTS=q4_k_m scripts/start_llmd.sh                                     # random-init 8B, Ollama's default mix
set -euo pipefail
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
BIN="$ROOT/p2p-llm-chat-go_b200/llmd"
[ -x "$BIN" ] || (cd "$ROOT" && python -c "import __graft_entry__ as g; g.build()")
GPUS="${GPUS:-0}"
BASE_PORT="${BASE_PORT:-11434}"          # web/streamlit_app.py:27 defaults OLLAMA_URL to this port
LOGDIR="${LOGDIR:-$ROOT/logs}"
mkdir -p "$LOGDIR"
i=0
for g in $GPUS; do
  port=$((BASE_PORT + i))
  LLMD_ADDR="127.0.0.1:$port" LLMD_GPU="$g" LLMD_MODEL="${LLM_MODEL:-llama3.1}" LLMD_GGUF="${MODEL_GGUF:-}" \
  LLMD_WEIGHTS="${WEIGHTS:-bf16}" LLMD_MAX_CTX="${MAX_CTX:-4096}" LLMD_MAX_SEQS="${MAX_SEQS:-64}" \
    nohup "$BIN" > "$LOGDIR/llmd_gpu$g.log" 2>&1 &
  echo $! > "$LOGDIR/llmd_gpu$g.pid"
  echo "llmd on GPU $g: http://127.0.0.1:$port  (log $LOGDIR/llmd_gpu$g.log, metrics /metrics)"
  i=$((i + 1))
done
echo "export OLLAMA_URL=http://127.0.0.1:$BASE_PORT   # then start the reference UI unchanged"
