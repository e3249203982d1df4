module github.com/NajyFannoun/P2P-LLM-Chat-Go/go/cmd/llmd

go 1.25.0
