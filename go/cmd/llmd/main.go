// Command llmd serves Ollama's HTTP API (/api/generate, /api/chat, /api/tags, /api/version, GET /) on top of
// libb200llm, the B200-native inference engine of this repository, through cgo.  Pointing OLLAMA_URL
// (web/streamlit_app.py:27) at it is the only change the chat application needs; the Streamlit UI and the
// libp2p node stay untouched.
//
// Layout follows the reference's convention of one module per command (go/cmd/node/go.mod,
// go/cmd/directory/go.mod) and its configuration style (environment variables with defaults,
// go/cmd/node/main.go:131-134, :286-291).  Standard library only.
//
// STATUS: written against include/b200llm.h but NOT compiled on the build image (no Go toolchain there).
// Its behaviour is mirrored line for line by p2p-llm-chat-go_b200/csrc/llmd.cpp, which is what the tests
// and benchmarks run.
package main

/*
#cgo CFLAGS: -I${SRCDIR}/../../../include
#cgo LDFLAGS: -L${SRCDIR}/../../../p2p-llm-chat-go_b200 -lb200llm -Wl,-rpath,${SRCDIR}/../../../p2p-llm-chat-go_b200
#include <stdlib.h>
#include "b200llm.h"
*/
import "C"

import (
	"encoding/json"
	"errors"
	"fmt"
	"log"
	"net/http"
	"os"
	"strconv"
	"strings"
	"sync/atomic"
	"time"
	"unicode/utf8"
	"unsafe"
)

// Prometheus counters served at /metrics
var mRequests, mErrors, mInFlight, mPromptTokens, mEvalTokens, mPromptNs, mEvalNs atomic.Int64

var (
	engine *C.b200_engine
	cfg    C.b200_config
	model  = getenv("LLMD_MODEL", "llama3.1")
)

func getenv(k, d string) string {
	if v := os.Getenv(k); v != "" {
		return v
	}
	return d
}

func lastError() error { return errors.New(C.GoString(C.b200_last_error())) }

// ---- request / reply shapes (Ollama API) -----------------------------------------------------------

type options struct {
	Temperature   *float32 `json:"temperature"`
	TopK          *int32   `json:"top_k"`
	TopP          *float32 `json:"top_p"`
	RepeatPenalty *float32 `json:"repeat_penalty"`
	RepeatLastN   *int32   `json:"repeat_last_n"`
	Seed          *int64   `json:"seed"`
	NumPredict    *int32   `json:"num_predict"`
}

type message struct {
	Role    string `json:"role"`
	Content string `json:"content"`
}

type request struct {
	Model    string    `json:"model"`
	Prompt   string    `json:"prompt"`
	System   string    `json:"system"`
	Raw      bool      `json:"raw"`
	Stream   *bool     `json:"stream"` // Ollama streams unless told otherwise; the reference sends false
	Messages []message `json:"messages"`
	Options  *options  `json:"options"`
}

type reply struct {
	Model              string   `json:"model"`
	CreatedAt          string   `json:"created_at"`
	Response           *string  `json:"response,omitempty"`
	Message            *message `json:"message,omitempty"`
	Done               bool     `json:"done"`
	DoneReason         string   `json:"done_reason,omitempty"`
	Context            []int32  `json:"context,omitempty"`
	TotalDuration      int64    `json:"total_duration,omitempty"`
	LoadDuration       int64    `json:"load_duration,omitempty"`
	PromptEvalCount    int32    `json:"prompt_eval_count,omitempty"`
	PromptEvalDuration int64    `json:"prompt_eval_duration,omitempty"`
	EvalCount          int32    `json:"eval_count,omitempty"`
	EvalDuration       int64    `json:"eval_duration,omitempty"`
}

// ---- prompt construction (Llama-3 chat framing, as Ollama's llama3.1 template) ------------------------

type promptBuilder struct {
	ids      []int32
	specials bool
}

func newPromptBuilder() *promptBuilder {
	p := &promptBuilder{specials: cfg.vocab > 128009}
	if p.specials {
		p.ids = append(p.ids, 128000) // <|begin_of_text|>
	}
	return p
}

func (p *promptBuilder) text(s string) {
	if len(s) == 0 {
		return
	}
	buf := make([]int32, len(s)+4)
	var n C.size_t
	cs := C.CString(s)
	defer C.free(unsafe.Pointer(cs))
	if C.b200_tokenize(engine, cs, C.size_t(len(s)), 0, (*C.int32_t)(unsafe.Pointer(&buf[0])), C.size_t(len(buf)), &n) == 0 {
		p.ids = append(p.ids, buf[:int(n)]...)
	}
}

func (p *promptBuilder) turn(role, content string) {
	if p.specials {
		p.ids = append(p.ids, 128006) // <|start_header_id|>
		p.text(role)
		p.ids = append(p.ids, 128007) // <|end_header_id|>
		p.text("\n\n")
		p.text(content)
		p.ids = append(p.ids, 128009) // <|eot_id|>
	} else {
		p.text(role + ": " + content + "\n")
	}
}

func (p *promptBuilder) openAssistant() {
	if p.specials {
		p.ids = append(p.ids, 128006)
		p.text("assistant")
		p.ids = append(p.ids, 128007)
		p.text("\n\n")
	} else {
		p.text("assistant: ")
	}
}

func sampling(o *options) C.b200_sampling {
	var sp C.b200_sampling
	C.b200_default_sampling(&sp) // Ollama's defaults — what the reference gets, it sends no options
	sp.num_predict = -1
	sp.seed = C.uint64_t(time.Now().UnixNano())
	if o == nil {
		return sp
	}
	if o.Temperature != nil {
		sp.temperature = C.float(*o.Temperature)
	}
	if o.TopP != nil {
		sp.top_p = C.float(*o.TopP)
	}
	if o.RepeatPenalty != nil {
		sp.repeat_penalty = C.float(*o.RepeatPenalty)
	}
	if o.TopK != nil {
		sp.top_k = C.int32_t(*o.TopK)
	}
	if o.RepeatLastN != nil {
		sp.repeat_last_n = C.int32_t(*o.RepeatLastN)
	}
	if o.NumPredict != nil {
		sp.num_predict = C.int32_t(*o.NumPredict)
	}
	if o.Seed != nil {
		sp.seed = C.uint64_t(*o.Seed)
	}
	return sp
}

func detok(id int32) []byte {
	var buf [64]C.char
	var n C.size_t
	cid := C.int32_t(id)
	if C.b200_detokenize(engine, &cid, 1, &buf[0], 64, &n) != 0 {
		return nil
	}
	return C.GoBytes(unsafe.Pointer(&buf[0]), C.int(n))
}

// utf8Stream releases only complete UTF-8 sequences (a byte-level vocabulary can split a character).
type utf8Stream struct{ pend []byte }

func (u *utf8Stream) push(b []byte) string {
	u.pend = append(u.pend, b...)
	var out strings.Builder
	for len(u.pend) > 0 {
		if !utf8.FullRune(u.pend) {
			break
		}
		r, size := utf8.DecodeRune(u.pend)
		out.WriteRune(r) // invalid bytes become U+FFFD
		u.pend = u.pend[size:]
	}
	return out.String()
}

func (u *utf8Stream) flush() string {
	s := strings.Repeat("�", len(u.pend))
	u.pend = nil
	return s
}

func writeError(w http.ResponseWriter, code int, msg string) {
	w.Header().Set("Content-Type", "application/json")
	w.WriteHeader(code)
	json.NewEncoder(w).Encode(map[string]string{"error": msg})
}

func generate(w http.ResponseWriter, r *http.Request, chat bool) {
	var req request
	if err := json.NewDecoder(r.Body).Decode(&req); err != nil {
		writeError(w, http.StatusBadRequest, "invalid JSON body")
		return
	}
	if req.Model != model && req.Model != model+":latest" {
		writeError(w, http.StatusNotFound, fmt.Sprintf("model '%s' not found", req.Model))
		return
	}
	stream := req.Stream == nil || *req.Stream
	pb := newPromptBuilder()
	switch {
	case chat:
		for _, m := range req.Messages {
			pb.turn(m.Role, m.Content)
		}
		pb.openAssistant()
	case req.Raw:
		pb.text(req.Prompt)
	default:
		if req.System != "" {
			pb.turn("system", req.System)
		}
		pb.turn("user", req.Prompt)
		pb.openAssistant()
	}
	if len(pb.ids) == 0 {
		pb.ids = []int32{0}
	}
	sp := sampling(req.Options)
	room := int(cfg.max_ctx) - len(pb.ids) + 1
	if room < 1 {
		writeError(w, http.StatusBadRequest, "prompt does not fit the context window")
		return
	}
	if sp.num_predict <= 0 || int(sp.num_predict) > room {
		sp.num_predict = C.int32_t(room)
	}
	var h *C.b200_req
	mRequests.Add(1)
	// the library copies the ids before b200_submit returns (cgo pointer rules)
	if C.b200_submit(engine, (*C.int32_t)(unsafe.Pointer(&pb.ids[0])), C.size_t(len(pb.ids)), &sp, &h) != 0 {
		mErrors.Add(1)
		writeError(w, http.StatusInternalServerError, lastError().Error())
		return
	}
	defer C.b200_free(h)
	mInFlight.Add(1)

	piece := func(text string, final bool, st *C.b200_stats, ctx []int32) reply {
		out := reply{Model: model, CreatedAt: time.Now().UTC().Format(time.RFC3339Nano), Done: final}
		if chat {
			out.Message = &message{Role: "assistant", Content: text}
		} else {
			out.Response = &text
		}
		if final {
			out.DoneReason = "length"
			if st.done_reason == 1 {
				out.DoneReason = "stop"
			}
			if !chat {
				out.Context = ctx
			}
			out.TotalDuration, out.LoadDuration = int64(st.total_ns), int64(st.load_ns)
			out.PromptEvalCount, out.PromptEvalDuration = int32(st.prompt_eval_count), int64(st.prompt_eval_ns)
			out.EvalCount, out.EvalDuration = int32(st.eval_count), int64(st.eval_ns)
		}
		return out
	}
	flusher, _ := w.(http.Flusher)
	enc := json.NewEncoder(w)
	if stream {
		w.Header().Set("Content-Type", "application/x-ndjson")
	}
	var u8 utf8Stream
	var full strings.Builder
	context := append([]int32(nil), pb.ids...)
	var failure error
	for {
		var id C.int32_t
		var done C.int
		if C.b200_next(h, &id, &done, -1) != 0 { // blocks this goroutine's OS thread until the next token
			failure = lastError()
			break
		}
		context = append(context, int32(id))
		text := ""
		if int32(id) != int32(cfg.eos_token) {
			text = u8.push(detok(int32(id)))
		}
		if done != 0 {
			text += u8.flush()
		}
		full.WriteString(text)
		if stream && text != "" {
			if err := enc.Encode(piece(text, false, nil, nil)); err != nil {
				C.b200_cancel(h) // client went away
			} else if flusher != nil {
				flusher.Flush()
			}
		}
		if done != 0 {
			break
		}
	}
	var st C.b200_stats
	C.b200_stats_get(h, &st)
	mInFlight.Add(-1)
	mPromptTokens.Add(int64(st.prompt_eval_count))
	mEvalTokens.Add(int64(st.eval_count))
	mPromptNs.Add(int64(st.load_ns + st.prompt_eval_ns))
	mEvalNs.Add(int64(st.eval_ns))
	if failure != nil {
		mErrors.Add(1)
		if !stream {
			writeError(w, http.StatusInternalServerError, failure.Error())
		} else {
			enc.Encode(map[string]string{"error": failure.Error()})
		}
		return
	}
	if stream {
		enc.Encode(piece("", true, &st, context))
	} else {
		w.Header().Set("Content-Type", "application/json")
		enc.Encode(piece(full.String(), true, &st, context))
	}
}

func main() {
	C.b200_llama31_8b(&cfg)
	if getenv("LLMD_WEIGHTS", "bf16") == "q4_k_m" {
		cfg.weight_format = C.B200_WEIGHTS_Q4_K_M
	}
	gpu, _ := strconv.Atoi(getenv("LLMD_GPU", "0"))
	ctx, _ := strconv.Atoi(getenv("LLMD_MAX_CTX", "2048"))
	seed, _ := strconv.ParseInt(getenv("LLMD_SEED", "0"), 10, 64)
	cfg.device, cfg.max_ctx, cfg.seed = C.int32_t(gpu), C.int32_t(ctx), C.int64_t(seed)
	seqs, _ := strconv.Atoi(getenv("LLMD_MAX_SEQS", "8"))
	if path := os.Getenv("LLMD_GGUF"); path != "" { // a real model file, e.g. the blob `ollama pull llama3.1` stored
		cpath := C.CString(path)
		defer C.free(unsafe.Pointer(cpath))
		if C.b200_load_gguf(cpath, cfg.device, cfg.max_ctx, C.int(seqs), &engine) != 0 {
			log.Fatalf("llmd: loading %s failed: %v", path, lastError())
		}
		C.b200_get_config(engine, &cfg)
	} else {
		cfg.max_seqs = C.int32_t(seqs)
		if C.b200_init(&cfg, &engine) != 0 {
			log.Fatalf("llmd: engine init failed: %v", lastError())
		}
	}
	defer C.b200_shutdown(engine)

	mux := http.NewServeMux()
	mux.HandleFunc("/", func(w http.ResponseWriter, r *http.Request) {
		if r.URL.Path != "/" {
			writeError(w, http.StatusNotFound, "not found")
			return
		}
		fmt.Fprint(w, "Ollama is running")
	})
	mux.HandleFunc("/api/version", func(w http.ResponseWriter, r *http.Request) {
		json.NewEncoder(w).Encode(map[string]string{"version": "0.5.7-b200llm", "engine": C.GoString(C.b200_version())})
	})
	mux.HandleFunc("/api/tags", func(w http.ResponseWriter, r *http.Request) {
		q := "BF16"
		if cfg.weight_format == C.B200_WEIGHTS_Q4_K_M {
			q = "Q4_K_M"
		}
		json.NewEncoder(w).Encode(map[string]any{"models": []map[string]any{{
			"name": model + ":latest", "model": model + ":latest", "size": int64(C.b200_weight_bytes(engine)),
			"details": map[string]string{"format": "gguf", "family": "llama", "quantization_level": q}}}})
	})
	post := func(chat bool) http.HandlerFunc {
		return func(w http.ResponseWriter, r *http.Request) {
			if r.Method != http.MethodPost {
				writeError(w, http.StatusMethodNotAllowed, "method not allowed")
				return
			}
			generate(w, r, chat)
		}
	}
	mux.HandleFunc("/metrics", func(w http.ResponseWriter, r *http.Request) {
		w.Header().Set("Content-Type", "text/plain; version=0.0.4")
		fmt.Fprintf(w, "llmd_requests_total %d\nllmd_request_errors_total %d\nllmd_requests_in_flight %d\n"+
			"llmd_prompt_tokens_total %d\nllmd_generated_tokens_total %d\nllmd_prompt_seconds_total %.6f\nllmd_generation_seconds_total %.6f\n",
			mRequests.Load(), mErrors.Load(), mInFlight.Load(), mPromptTokens.Load(), mEvalTokens.Load(),
			float64(mPromptNs.Load())/1e9, float64(mEvalNs.Load())/1e9)
	})
	mux.HandleFunc("/api/generate", post(false))
	mux.HandleFunc("/api/chat", post(true))

	addr := getenv("LLMD_ADDR", getenv("OLLAMA_HOST", "127.0.0.1:11434"))
	log.Printf("llmd: %s serving model '%s' on %s", C.GoString(C.b200_version()), model, addr)
	// a client that stops sending mid-request must not pin a goroutine for ever (no WriteTimeout: replies stream)
	srv := &http.Server{Addr: addr, Handler: mux, ReadHeaderTimeout: 30 * time.Second, ReadTimeout: 30 * time.Second}
	log.Fatal(srv.ListenAndServe())
}
