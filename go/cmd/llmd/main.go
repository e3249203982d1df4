// Command llmd serves Ollama's HTTP API (/api/generate, /api/chat, /api/tags, /api/ps, /api/version, GET /) on top
// of libb200llm, the B200-native inference engine of this repository, through cgo.  Pointing OLLAMA_URL
// (web/streamlit_app.py:27) at it is the only change the chat application needs; the Streamlit UI and the
// libp2p node stay untouched.
//
// One URL, N GPUs: the reference knows a single OLLAMA_URL (web/streamlit_app.py:27, start_all.sh:30-39), so this
// process owns every GPU named in LLMD_GPUS — one engine per device, no collective between them.  A request goes to
// the engine with the fewest outstanding tokens, where prompt tokens an engine still caches for this peer
// (b200_prefix_match) count as already done.
//
// Layout follows the reference's convention of one module per command (go/cmd/node/go.mod,
// go/cmd/directory/go.mod) and its configuration style (environment variables with defaults,
// go/cmd/node/main.go:131-134, :286-291).  Standard library only.
//
// STATUS: written against include/b200llm.h but NOT compiled on the build image (no Go toolchain there).
// Its behaviour is mirrored by p2p-llm-chat-go_b200/csrc/llmd.cpp, which is what the tests and benchmarks run.
package main

/*
#cgo CFLAGS: -I${SRCDIR}/../../../include
#cgo LDFLAGS: -L${SRCDIR}/../../../p2p-llm-chat-go_b200 -lb200llm -Wl,-rpath,${SRCDIR}/../../../p2p-llm-chat-go_b200
#include <stdlib.h>
#include "b200llm.h"
*/
import "C"

import (
	"bytes"
	"encoding/json"
	"errors"
	"fmt"
	"io"
	"log"
	"net/http"
	"os"
	"runtime"
	"strconv"
	"strings"
	"sync"
	"sync/atomic"
	"time"
	"unicode"
	"unicode/utf8"
	"unsafe"
)

// Prometheus counters served at /metrics
var mRequests, mErrors, mInFlight, mPromptTokens, mEvalTokens, mPromptNs, mEvalNs, mRejected, mConns atomic.Int64

// gpu is one engine on one device; the counters feed the router and /metrics.
type gpu struct {
	device                                                                    int
	engine                                                                    *C.b200_engine
	outstanding, inFlight, requests, genTokens, promptTokens, reusedTokens atomic.Int64
}

var (
	gpus          []*gpu
	engine        *C.b200_engine // gpus[0].engine: tokenizer / template / stop-id queries (the same on every GPU)
	cfg           C.b200_config
	model         = getenv("LLMD_MODEL", "llama3.1")
	numCtxDefault int
	maxConns      int64 = 1024
	rr            atomic.Uint32
	expiresUnix   atomic.Int64

	framing = framePlain
	stopIDs []int32
	bosID   int32 = -1
)

const (
	framePlain = iota
	frameLlama3
	frameChatML
)

func getenv(k, d string) string {
	if v := os.Getenv(k); v != "" {
		return v
	}
	return d
}

// withError runs a library call and, if it failed, reads the library's thread-local message on the SAME OS thread
// (a goroutine may otherwise be moved between the call and b200_last_error).
func withError(call func() C.int) error {
	runtime.LockOSThread()
	defer runtime.UnlockOSThread()
	if call() != 0 {
		return errors.New(C.GoString(C.b200_last_error()))
	}
	return nil
}

// ---- request / reply shapes (Ollama API) -----------------------------------------------------------

type options struct {
	Temperature   *float32        `json:"temperature"`
	TopK          *int32          `json:"top_k"`
	TopP          *float32        `json:"top_p"`
	RepeatPenalty *float32        `json:"repeat_penalty"`
	RepeatLastN   *int32          `json:"repeat_last_n"`
	Seed          *int64          `json:"seed"`
	NumPredict    *int32          `json:"num_predict"`
	NumCtx        json.RawMessage `json:"num_ctx"`
	Stop          json.RawMessage `json:"stop"` // a string or an array of strings
}

type message struct {
	Role    string `json:"role"`
	Content string `json:"content"`
}

type request struct {
	Model     string          `json:"model"`
	Prompt    string          `json:"prompt"`
	System    string          `json:"system"`
	Template  string          `json:"template"`
	Raw       bool            `json:"raw"`
	Stream    *bool           `json:"stream"` // Ollama streams unless told otherwise; the reference sends false
	Messages  json.RawMessage `json:"messages"`
	Context   json.RawMessage `json:"context"`
	KeepAlive json.RawMessage `json:"keep_alive"`
	Format    json.RawMessage `json:"format"`
	Options   *options        `json:"options"`
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

// ---- chat framing ------------------------------------------------------------------------------------
// Which framing the model expects is read from the GGUF file's tokenizer.chat_template (recognised by the control
// tokens it uses, not interpreted): Llama-3 headers, ChatML, or — for engines without a vocabulary — Llama-3 when the
// id space holds the published Llama-3 control ids, else plain text.

func special(text string, published int32) int32 {
	cs := C.CString(text)
	defer C.free(unsafe.Pointer(cs))
	if id := int32(C.b200_special_token(engine, cs)); id >= 0 {
		return id
	}
	if published >= 0 && published < int32(cfg.vocab) {
		return published
	}
	return -1
}

func detectFraming() {
	var n C.size_t
	C.b200_chat_template(engine, nil, 0, &n)
	tpl := ""
	if n > 0 {
		buf := make([]byte, int(n))
		if C.b200_chat_template(engine, (*C.char)(unsafe.Pointer(&buf[0])), n, &n) == 0 {
			tpl = string(buf[:int(n)])
		}
	}
	switch {
	case strings.Contains(tpl, "<|start_header_id|>"):
		framing = frameLlama3
	case strings.Contains(tpl, "<|im_start|>"):
		framing = frameChatML
	case tpl == "" && cfg.vocab > 128009:
		framing = frameLlama3
	}
	if framing == frameLlama3 {
		bosID = special("<|begin_of_text|>", 128000)
	}
	var ids [64]C.int32_t
	var ns C.size_t
	if C.b200_stop_ids(engine, &ids[0], 64, &ns) == 0 {
		for i := 0; i < int(ns); i++ {
			stopIDs = append(stopIDs, int32(ids[i]))
		}
	}
	if framing == frameLlama3 {
		for _, extra := range []int32{special("<|eot_id|>", 128009), special("<|eom_id|>", 128008), special("<|end_of_text|>", 128001)} {
			if extra >= 0 && !isStopID(extra) {
				stopIDs = append(stopIDs, extra)
			}
		}
	}
	if cfg.eos_token < 0 { // benchmarks / tests: never stop early
		stopIDs = nil
	}
}

func isStopID(id int32) bool {
	for _, s := range stopIDs {
		if s == id {
			return true
		}
	}
	return false
}

type promptBuilder struct {
	ids     []int32
	tooLong bool
}

func (p *promptBuilder) text(s string) {
	if len(s) == 0 {
		return
	}
	buf := make([]int32, int(cfg.max_ctx)+8) // a text that cannot fit the context is refused untokenised
	var n C.size_t
	cs := C.CString(s)
	defer C.free(unsafe.Pointer(cs))
	if C.b200_tokenize(engine, cs, C.size_t(len(s)), 0, (*C.int32_t)(unsafe.Pointer(&buf[0])), C.size_t(len(buf)), &n) == 0 {
		p.ids = append(p.ids, buf[:int(n)]...)
	} else {
		p.tooLong = true
	}
}

var llama3Known = map[string]int32{"<|begin_of_text|>": 128000, "<|end_of_text|>": 128001, "<|start_header_id|>": 128006,
	"<|end_header_id|>": 128007, "<|eom_id|>": 128008, "<|eot_id|>": 128009}

// textWithSpecials: text in which control tokens may be written out ("<|eot_id|>"): custom templates, raw prompts.
func (p *promptBuilder) textWithSpecials(s string) {
	for len(s) > 0 {
		open := strings.Index(s, "<|")
		if open < 0 {
			break
		}
		cl := strings.Index(s[open+2:], "|>")
		if cl < 0 {
			break
		}
		end := open + 2 + cl + 2
		name := s[open:end]
		id := special(name, -1)
		if id < 0 && framing == frameLlama3 {
			if k, ok := llama3Known[name]; ok && k < int32(cfg.vocab) {
				id = k
			}
		}
		if id < 0 { // not a control token: ordinary text
			p.text(s[:end])
		} else {
			p.text(s[:open])
			p.ids = append(p.ids, id)
		}
		s = s[end:]
	}
	p.text(s)
}

func (p *promptBuilder) begin() {
	if bosID >= 0 {
		p.ids = append(p.ids, bosID)
	}
}

func (p *promptBuilder) turn(role, content string) {
	switch framing {
	case frameLlama3:
		p.ids = append(p.ids, special("<|start_header_id|>", 128006))
		p.text(role)
		p.ids = append(p.ids, special("<|end_header_id|>", 128007))
		p.text("\n\n")
		p.text(content)
		p.ids = append(p.ids, special("<|eot_id|>", 128009))
	case frameChatML:
		p.textWithSpecials("<|im_start|>" + role + "\n")
		p.text(content)
		p.textWithSpecials("<|im_end|>\n")
	default:
		p.text(role + ": " + content + "\n")
	}
}

func (p *promptBuilder) openAssistant() {
	switch framing {
	case frameLlama3:
		p.ids = append(p.ids, special("<|start_header_id|>", 128006))
		p.text("assistant")
		p.ids = append(p.ids, special("<|end_header_id|>", 128007))
		p.text("\n\n")
	case frameChatML:
		p.textWithSpecials("<|im_start|>assistant\n")
	default:
		p.text("assistant: ")
	}
}

// renderTemplate renders the subset of Go templates Ollama's own model templates use for one turn: {{ .System }},
// {{ .Prompt }}, {{ .Response }} (rendering stops there), {{ if .System }} / {{ if .Prompt }} ... {{ end }}, with
// optional "-" trim markers.  Anything else is refused with a clear error rather than guessed at.  (text/template
// itself is not used so that both front ends accept exactly the same inputs.)
func renderTemplate(tpl, system, prompt string) (string, error) {
	var out strings.Builder
	var stack []bool
	emitting := func() bool {
		for _, b := range stack {
			if !b {
				return false
			}
		}
		return true
	}
	trimNext := false
	for len(tpl) > 0 {
		open := strings.Index(tpl, "{{")
		lit := tpl
		if open >= 0 {
			lit = tpl[:open]
		}
		if trimNext {
			lit = strings.TrimLeft(lit, " \t\r\n")
			trimNext = false
		}
		if open < 0 {
			if emitting() {
				out.WriteString(lit)
			}
			break
		}
		cl := strings.Index(tpl[open+2:], "}}")
		if cl < 0 {
			return "", errors.New("template: unterminated {{")
		}
		act := tpl[open+2 : open+2+cl]
		if strings.HasPrefix(act, "-") {
			lit = strings.TrimRight(lit, " \t\r\n")
			act = act[1:]
		}
		if strings.HasSuffix(act, "-") {
			trimNext = true
			act = act[:len(act)-1]
		}
		if emitting() {
			out.WriteString(lit)
		}
		switch act = strings.TrimSpace(act); act {
		case ".System":
			if emitting() {
				out.WriteString(system)
			}
		case ".Prompt":
			if emitting() {
				out.WriteString(prompt)
			}
		case ".Response":
			if emitting() {
				return out.String(), nil
			}
		case "if .System":
			stack = append(stack, system != "")
		case "if .Prompt":
			stack = append(stack, prompt != "")
		case "end":
			if len(stack) == 0 {
				return "", errors.New("template: {{ end }} without {{ if }}")
			}
			stack = stack[:len(stack)-1]
		default:
			return "", fmt.Errorf("template: unsupported action {{ %s }} (supported: .System .Prompt .Response, if .System/.Prompt, end)", act)
		}
		tpl = tpl[open+2+cl+2:]
	}
	if len(stack) != 0 {
		return "", errors.New("template: {{ if }} without {{ end }}")
	}
	return out.String(), nil
}

type genOptions struct {
	sp     C.b200_sampling
	numCtx int
	stop   []string
}

func fillOptions(o *options) (genOptions, error) {
	var g genOptions
	C.b200_default_sampling(&g.sp) // Ollama's defaults — what the reference gets, it sends no options
	g.sp.num_predict = -1
	g.sp.seed = C.uint64_t(time.Now().UnixNano())
	g.numCtx = int(cfg.max_ctx)
	if numCtxDefault > 0 {
		g.numCtx = numCtxDefault
	}
	if o == nil {
		return g, nil
	}
	if o.Temperature != nil {
		g.sp.temperature = C.float(*o.Temperature)
	}
	if o.TopP != nil {
		g.sp.top_p = C.float(*o.TopP)
	}
	if o.RepeatPenalty != nil {
		g.sp.repeat_penalty = C.float(*o.RepeatPenalty)
	}
	if o.TopK != nil {
		g.sp.top_k = C.int32_t(*o.TopK)
	}
	if o.RepeatLastN != nil {
		g.sp.repeat_last_n = C.int32_t(*o.RepeatLastN)
	}
	if o.NumPredict != nil {
		g.sp.num_predict = C.int32_t(*o.NumPredict)
	}
	if o.Seed != nil {
		g.sp.seed = C.uint64_t(*o.Seed)
	}
	if len(o.NumCtx) > 0 && string(o.NumCtx) != "null" {
		var nc float64
		if json.Unmarshal(o.NumCtx, &nc) != nil || nc < 4 {
			return g, errors.New("options.num_ctx must be a number >= 4")
		}
		// a window larger than the KV room this server was started with cannot be honoured: say so
		if nc > float64(cfg.max_ctx) {
			return g, fmt.Errorf("options.num_ctx %d exceeds this server's LLMD_MAX_CTX %d", int64(nc), int(cfg.max_ctx))
		}
		g.numCtx = int(nc)
	}
	if len(o.Stop) > 0 && string(o.Stop) != "null" {
		var one string
		var many []string
		if json.Unmarshal(o.Stop, &one) == nil {
			many = []string{one}
		} else if json.Unmarshal(o.Stop, &many) != nil {
			return g, errors.New("options.stop must be a string or an array of strings")
		}
		for _, s := range many {
			if s != "" {
				g.stop = append(g.stop, s)
			}
		}
	}
	return g, nil
}

// parseKeepAlive: a number of seconds or a Go duration string.  The weights of this server stay resident for its
// lifetime; the value is validated, answers load / unload probes as Ollama does, and sets expires_at in /api/ps.
func parseKeepAlive(raw json.RawMessage) (float64, error) {
	if len(raw) == 0 || string(raw) == "null" {
		return 300, nil // Ollama's default: five minutes
	}
	var num float64
	if json.Unmarshal(raw, &num) == nil {
		return num, nil
	}
	var s string
	if json.Unmarshal(raw, &s) != nil {
		return 0, errors.New("keep_alive must be a number of seconds or a duration string")
	}
	if v, err := strconv.ParseFloat(s, 64); err == nil {
		return v, nil // a bare number in a string: seconds
	}
	d, err := time.ParseDuration(s)
	if err != nil {
		return 0, fmt.Errorf("keep_alive: malformed duration '%s'", s)
	}
	return d.Seconds(), nil
}

func detok(id int32) ([]byte, error) {
	var small [64]C.char
	var n C.size_t
	cid := C.int32_t(id)
	var out []byte
	err := withError(func() C.int {
		rc := C.b200_detokenize(engine, &cid, 1, &small[0], 64, &n)
		if rc == 0 {
			out = C.GoBytes(unsafe.Pointer(&small[0]), C.int(n))
			return 0
		}
		if rc == C.B200_ERANGE && n > 64 { // Llama-3 has entries longer than 64 bytes: retry with the size the library reported
			big := make([]byte, int(n))
			rc = C.b200_detokenize(engine, &cid, 1, (*C.char)(unsafe.Pointer(&big[0])), n, &n)
			if rc == 0 {
				out = big[:int(n)]
			}
		}
		return rc
	})
	return out, err
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

// route: the engine whose queue is shortest once the tokens it already caches for this prompt are taken off the bill.
func route(ids []int32, newTokens int64) (*gpu, int) {
	var best *gpu
	var bestCost int64
	bestMatch := 0
	start := int(rr.Add(1) - 1)
	for k := range gpus {
		g := gpus[(start+k)%len(gpus)] // rotating start: ties spread round-robin
		match := 0
		if len(gpus) > 1 {
			match = int(C.b200_prefix_match(g.engine, (*C.int32_t)(unsafe.Pointer(&ids[0])), C.size_t(len(ids))))
		}
		cost := g.outstanding.Load() + int64(len(ids)) + newTokens - int64(match)
		if best == nil || cost < bestCost {
			best, bestCost, bestMatch = g, cost, match
		}
	}
	return best, bestMatch
}

// suggest serves POST /suggest — what a node-side handler next to /send (reference go/cmd/node/main.go:219) forwards
// to (SURVEY 8f-3).  Body {"content": string, "stream": bool (default false, like the reference's blocking call),
// "options"?}.  It is the reference's own call (web/streamlit_app.py:89-98: that prompt, that model, `.strip()` of the
// response) behind one route: the request is rewritten into the /api/generate body and handed to `gen`; the reply is
// rewritten on the way out into {"suggestion": text, ...stats} or, streaming, NDJSON {"delta": piece, "done": false}
// lines closed by an object that carries the whole suggestion.
func suggest(gen http.HandlerFunc) http.HandlerFunc {
	return func(w http.ResponseWriter, r *http.Request) {
		var body struct {
			Content *string         `json:"content"`
			Stream  bool            `json:"stream"`
			Options json.RawMessage `json:"options"`
		}
		if r.Method != http.MethodPost {
			writeError(w, http.StatusMethodNotAllowed, "method not allowed")
			return
		}
		if err := json.NewDecoder(r.Body).Decode(&body); err != nil || body.Content == nil {
			writeError(w, http.StatusBadRequest, `body must be {"content": string, "stream": bool}`)
			return
		}
		inner := map[string]any{"model": model, "stream": body.Stream,
			"prompt": "You are a helpful assistant. Draft a concise, friendly reply to the following message:\n\n" + *body.Content + "\n\nReply:"}
		if len(body.Options) > 0 {
			inner["options"] = body.Options
		}
		buf, _ := json.Marshal(inner)
		r2 := r.Clone(r.Context())
		r2.Body = io.NopCloser(bytes.NewReader(buf))
		gen(&suggestWriter{w: w}, r2)
	}
}

// suggestWriter turns the /api/generate reply lines into the /suggest ones as they are written.
type suggestWriter struct {
	w      http.ResponseWriter
	code   int
	line   []byte
	soFar  strings.Builder
}

func (s *suggestWriter) Header() http.Header { return s.w.Header() }
func (s *suggestWriter) WriteHeader(code int) {
	s.code = code
	s.w.WriteHeader(code)
}
func (s *suggestWriter) Flush() {
	if f, ok := s.w.(http.Flusher); ok {
		f.Flush()
	}
}
func (s *suggestWriter) Write(p []byte) (int, error) {
	if s.code != 0 && s.code != http.StatusOK { // an error body: passed through as it is
		return s.w.Write(p)
	}
	s.line = append(s.line, p...)
	for {
		i := bytes.IndexByte(s.line, '\n')
		if i < 0 {
			return len(p), nil
		}
		var in map[string]json.RawMessage
		if err := json.Unmarshal(s.line[:i], &in); err != nil {
			return 0, err
		}
		if _, isErr := in["error"]; isErr { // a failure after the stream began: the line goes out as it is
			if _, err := s.w.Write(s.line[:i+1]); err != nil {
				return 0, err
			}
			s.line = s.line[i+1:]
			continue
		}
		s.line = s.line[i+1:]
		var piece string
		json.Unmarshal(in["response"], &piece)
		s.soFar.WriteString(piece)
		out := map[string]any{"done": false, "delta": piece}
		if string(in["done"]) == "true" {
			// Python's str.strip() (web/streamlit_app.py:98): Unicode white space plus 0x1c-0x1f, nothing else
			out = map[string]any{"suggestion": strings.TrimFunc(s.soFar.String(), func(r rune) bool {
				return unicode.IsSpace(r) || (r >= 0x1c && r <= 0x1f)
			})}
			for k, v := range in { // done, done_reason and the *_duration / *_count fields
				if k != "response" && k != "model" && k != "created_at" && k != "context" {
					out[k] = v
				}
			}
		}
		line, _ := json.Marshal(out)
		if _, err := s.w.Write(append(line, '\n')); err != nil {
			return 0, err
		}
	}
}

func probeReply(w http.ResponseWriter, chat bool, keepAlive float64) {
	out := reply{Model: model, CreatedAt: time.Now().UTC().Format(time.RFC3339Nano), Done: true, DoneReason: "load"}
	if keepAlive == 0 {
		out.DoneReason = "unload"
	}
	empty := ""
	if chat {
		out.Message = &message{Role: "assistant", Content: ""}
	} else {
		out.Response = &empty
	}
	w.Header().Set("Content-Type", "application/json")
	json.NewEncoder(w).Encode(out)
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
	gopt, err := fillOptions(req.Options)
	if err != nil {
		writeError(w, http.StatusBadRequest, err.Error())
		return
	}
	keepAlive, err := parseKeepAlive(req.KeepAlive)
	if err != nil {
		writeError(w, http.StatusBadRequest, err.Error())
		return
	}
	if keepAlive < 0 {
		expiresUnix.Store(-1)
	} else {
		expiresUnix.Store(time.Now().Unix() + int64(keepAlive))
	}
	if f := strings.TrimSpace(string(req.Format)); f != "" && f != "null" && f != `""` {
		writeError(w, http.StatusBadRequest, "format: constrained (JSON / schema) output is not supported by this server")
		return
	}

	pb := &promptBuilder{}
	havePrompt := false
	if chat {
		var msgs []message
		if len(req.Messages) > 0 && string(req.Messages) != "null" {
			if json.Unmarshal(req.Messages, &msgs) != nil {
				writeError(w, http.StatusBadRequest, "messages must be an array")
				return
			}
		}
		pb.begin()
		for _, m := range msgs {
			role := m.Role
			if role == "" {
				role = "user"
			}
			pb.turn(role, m.Content)
			havePrompt = true
		}
		pb.openAssistant()
	} else {
		havePrompt = req.Prompt != ""
		if len(req.Context) > 0 && string(req.Context) != "null" { // the token ids a previous reply returned
			var ctx []float64
			if json.Unmarshal(req.Context, &ctx) != nil {
				writeError(w, http.StatusBadRequest, "context must be an array of token ids")
				return
			}
			for _, v := range ctx {
				if v < 0 || v >= float64(cfg.vocab) || v != float64(int64(v)) {
					writeError(w, http.StatusBadRequest, "context holds an invalid token id")
					return
				}
				pb.ids = append(pb.ids, int32(v))
			}
		}
		fresh := len(pb.ids) == 0
		switch {
		case req.Raw:
			pb.textWithSpecials(req.Prompt)
		case req.Template != "":
			rendered, err := renderTemplate(req.Template, req.System, req.Prompt)
			if err != nil {
				writeError(w, http.StatusBadRequest, err.Error())
				return
			}
			if fresh && !strings.HasPrefix(rendered, "<|begin_of_text|>") {
				pb.begin()
			}
			pb.textWithSpecials(rendered)
		default:
			if fresh {
				pb.begin()
			}
			if req.System != "" {
				pb.turn("system", req.System)
			}
			pb.turn("user", req.Prompt)
			pb.openAssistant()
		}
	}
	if pb.tooLong {
		writeError(w, http.StatusBadRequest, "prompt does not fit the context window")
		return
	}
	if !havePrompt { // Ollama's load / unload probe: an empty prompt only (re)sets the residency timer
		probeReply(w, chat, keepAlive)
		return
	}
	if len(pb.ids) == 0 {
		pb.ids = []int32{0}
	}
	// num_ctx: prompt + reply live in a window of num_ctx tokens.  A longer prompt is cut the way Ollama cuts it — the
	// first token (BOS) stays, the oldest tokens after it go — leaving room for the reply; the reply ends with
	// done_reason "length" when the window is full (this server does not shift the context while generating).
	limit := gopt.numCtx
	wantOut := limit / 4
	if wantOut < 1 {
		wantOut = 1
	}
	if gopt.sp.num_predict > 0 {
		wantOut = int(gopt.sp.num_predict)
		if wantOut > limit/2 {
			wantOut = limit / 2
		}
	}
	maxPrompt := limit - wantOut
	if maxPrompt < 1 {
		maxPrompt = 1
	}
	if len(pb.ids) > maxPrompt {
		keep := 0
		if bosID >= 0 && pb.ids[0] == bosID {
			keep = 1
		}
		drop := len(pb.ids) - maxPrompt
		pb.ids = append(pb.ids[:keep:keep], pb.ids[keep+drop:]...)
	}
	room := limit - len(pb.ids) + 1
	if room < 1 {
		writeError(w, http.StatusBadRequest, "prompt does not fit the context window")
		return
	}
	if gopt.sp.num_predict <= 0 || int(gopt.sp.num_predict) > room {
		gopt.sp.num_predict = C.int32_t(room)
	}

	g, reused := route(pb.ids, int64(gopt.sp.num_predict))
	bill := int64(len(pb.ids)) - int64(reused) + int64(gopt.sp.num_predict)
	var h *C.b200_req
	mRequests.Add(1)
	g.requests.Add(1)
	// the library copies the ids before b200_submit returns (cgo pointer rules)
	if err := withError(func() C.int {
		return C.b200_submit(g.engine, (*C.int32_t)(unsafe.Pointer(&pb.ids[0])), C.size_t(len(pb.ids)), &gopt.sp, &h)
	}); err != nil {
		mErrors.Add(1)
		writeError(w, http.StatusInternalServerError, err.Error())
		return
	}
	defer C.b200_free(h)
	g.outstanding.Add(bill)
	mInFlight.Add(1)
	g.inFlight.Add(1)

	piece := func(text string, final bool, st *C.b200_stats, ctx []int32, reason string) reply {
		out := reply{Model: model, CreatedAt: time.Now().UTC().Format(time.RFC3339Nano), Done: final}
		if chat {
			out.Message = &message{Role: "assistant", Content: text}
		} else {
			out.Response = &text
		}
		if final {
			out.DoneReason = reason
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
	full := ""  // the reply so far (complete characters only)
	sent := 0   // bytes of `full` already streamed
	maxStop := 0
	for _, s := range gopt.stop {
		if len(s) > maxStop {
			maxStop = len(s)
		}
	}
	context := append([]int32(nil), pb.ids...)
	var failure error
	alive, stopped := true, false
	for !stopped {
		var id C.int32_t
		var done C.int
		// blocks this goroutine's OS thread until the next token
		if err := withError(func() C.int { return C.b200_next(h, &id, &done, -1) }); err != nil {
			failure = err
			break
		}
		stopID := isStopID(int32(id))
		text := ""
		if !stopID {
			context = append(context, int32(id))
			raw, err := detok(int32(id))
			if err != nil {
				failure = err
				C.b200_cancel(h)
				break
			}
			text = u8.push(raw)
		}
		if done != 0 || stopID {
			text += u8.flush()
		}
		before := len(full)
		full += text
		safe := len(full) // bytes that can be shown: everything, unless a stop sequence may be forming at the end
		if len(gopt.stop) > 0 {
			from := 0
			if before > maxStop {
				from = before - maxStop
			}
			hit := -1
			for _, s := range gopt.stop {
				if i := strings.Index(full[from:], s); i >= 0 && (hit < 0 || from+i < hit) {
					hit = from + i
				}
			}
			if hit >= 0 {
				full = full[:hit]
				safe = hit
				stopped = true
			} else if done == 0 && !stopID {
				for _, s := range gopt.stop {
					for k := min(len(s)-1, len(full)); k > 0; k-- {
						if strings.HasSuffix(full, s[:k]) {
							if len(full)-k < safe {
								safe = len(full) - k
							}
							break
						}
					}
				}
			}
		}
		if stopID {
			stopped = true
		}
		if stream && alive && safe > sent {
			if err := enc.Encode(piece(full[sent:safe], false, nil, nil, "")); err != nil {
				alive = false
				C.b200_cancel(h) // client went away: stop generating
				break
			} else if flusher != nil {
				flusher.Flush()
			}
			sent = safe
		}
		if stopped && done == 0 {
			C.b200_cancel(h) // a stop sequence / secondary stop id ended the reply early
		}
		if done != 0 {
			break
		}
	}
	var st C.b200_stats
	C.b200_stats_get(h, &st)
	g.outstanding.Add(-bill)
	mInFlight.Add(-1)
	g.inFlight.Add(-1)
	mPromptTokens.Add(int64(st.prompt_eval_count))
	mEvalTokens.Add(int64(st.eval_count))
	g.promptTokens.Add(int64(st.prompt_eval_count))
	g.genTokens.Add(int64(st.eval_count))
	g.reusedTokens.Add(int64(len(pb.ids)) - int64(st.prompt_eval_count))
	mPromptNs.Add(int64(st.load_ns + st.prompt_eval_ns))
	mEvalNs.Add(int64(st.eval_ns))
	reason := "length"
	if stopped || st.done_reason == 1 {
		reason = "stop"
	}
	if failure != nil {
		mErrors.Add(1)
		if !stream {
			writeError(w, http.StatusInternalServerError, failure.Error())
		} else if alive {
			enc.Encode(map[string]string{"error": failure.Error()})
		}
		return
	}
	if stream {
		if alive {
			if len(full) > sent {
				enc.Encode(piece(full[sent:], false, nil, nil, ""))
			}
			enc.Encode(piece("", true, &st, context, reason))
		}
	} else {
		w.Header().Set("Content-Type", "application/json")
		enc.Encode(piece(full, true, &st, context, reason))
	}
}

func modelEntry(withExpiry bool) map[string]any {
	q := "BF16"
	if cfg.weight_format == C.B200_WEIGHTS_Q4_K_M {
		q = "Q4_K_M"
	}
	m := map[string]any{"name": model + ":latest", "model": model + ":latest", "size": int64(C.b200_weight_bytes(engine)),
		"details": map[string]string{"format": "gguf", "family": "llama", "quantization_level": q}}
	if withExpiry {
		ex := expiresUnix.Load()
		t := time.Unix(ex, 0)
		if ex < 0 {
			t = time.Now().AddDate(100, 0, 0)
		} else if ex == 0 {
			t = time.Now().Add(5 * time.Minute)
		}
		m["expires_at"] = t.UTC().Format(time.RFC3339Nano)
		m["size_vram"] = int64(C.b200_weight_bytes(engine)) * int64(len(gpus))
	}
	return m
}

func main() {
	C.b200_llama31_8b(&cfg)
	if tiny, _ := strconv.Atoi(getenv("LLMD_TINY", "0")); tiny > 0 { // small random-init model of the same architecture (tests)
		cfg.n_layers, cfg.d_model, cfg.n_heads, cfg.n_kv_heads, cfg.d_ffn, cfg.vocab = C.int32_t(tiny), 2048, 16, 4, 2048, 1024
		cfg.rope_orig_ctx, cfg.eos_token = 64, -1
	}
	if getenv("LLMD_WEIGHTS", "bf16") == "q4_k_m" {
		cfg.weight_format = C.B200_WEIGHTS_Q4_K_M
	}
	ctx, _ := strconv.Atoi(getenv("LLMD_MAX_CTX", "2048"))
	seed, _ := strconv.ParseInt(getenv("LLMD_SEED", "0"), 10, 64)
	cfg.max_ctx, cfg.seed = C.int32_t(ctx), C.int64_t(seed)
	numCtxDefault, _ = strconv.Atoi(getenv("LLMD_NUM_CTX", "0"))
	maxConns, _ = strconv.ParseInt(getenv("LLMD_MAX_CONNS", "1024"), 10, 64)
	seqs, _ := strconv.Atoi(getenv("LLMD_MAX_SEQS", "8"))
	eos, eosSet := os.LookupEnv("LLMD_EOS") // benchmarks: -1 = never stop early
	if eosSet {
		v, _ := strconv.Atoi(eos)
		cfg.eos_token = C.int32_t(v)
	}
	// LLMD_GPUS=0,1,...: one engine per device, all behind this one address; the engines load side by side
	for _, f := range strings.Split(getenv("LLMD_GPUS", getenv("LLMD_GPU", "0")), ",") {
		if f = strings.TrimSpace(f); f != "" {
			d, _ := strconv.Atoi(f)
			gpus = append(gpus, &gpu{device: d})
		}
	}
	if len(gpus) == 0 {
		gpus = append(gpus, &gpu{})
	}
	path := os.Getenv("LLMD_GGUF") // a real model file, e.g. the blob `ollama pull llama3.1` stored (README.md:70)
	cfgs := make([]C.b200_config, len(gpus))
	errs := make([]error, len(gpus))
	var wg sync.WaitGroup
	for i, g := range gpus {
		wg.Add(1)
		go func(i int, g *gpu) {
			defer wg.Done()
			cfgs[i] = cfg
			if path != "" {
				cpath := C.CString(path)
				defer C.free(unsafe.Pointer(cpath))
				errs[i] = withError(func() C.int { return C.b200_load_gguf(cpath, C.int(g.device), cfg.max_ctx, C.int(seqs), &g.engine) })
				if errs[i] == nil {
					C.b200_get_config(g.engine, &cfgs[i])
				}
			} else {
				cfgs[i].device, cfgs[i].max_seqs = C.int32_t(g.device), C.int32_t(seqs)
				errs[i] = withError(func() C.int { return C.b200_init(&cfgs[i], &g.engine) })
			}
		}(i, g)
	}
	wg.Wait()
	for i, g := range gpus {
		if errs[i] != nil {
			log.Fatalf("llmd: engine init failed on GPU %d: %v", g.device, errs[i])
		}
	}
	cfg = cfgs[0]
	if eosSet {
		v, _ := strconv.Atoi(eos)
		cfg.eos_token = C.int32_t(v)
	}
	engine = gpus[0].engine
	detectFraming()

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
		json.NewEncoder(w).Encode(map[string]any{"models": []map[string]any{modelEntry(false)}})
	})
	mux.HandleFunc("/api/ps", func(w http.ResponseWriter, r *http.Request) {
		json.NewEncoder(w).Encode(map[string]any{"models": []map[string]any{modelEntry(true)}})
	})
	post := func(chat bool) http.HandlerFunc {
		return func(w http.ResponseWriter, r *http.Request) {
			if r.Method != http.MethodPost {
				writeError(w, http.StatusMethodNotAllowed, "method not allowed")
				return
			}
			// bounded: beyond LLMD_MAX_CONNS a request is turned away at once (503) instead of piling up OS threads
			// blocked in b200_next — 256 peers with one blocking request each is the design load (BASELINE config 4)
			if mConns.Add(1) > maxConns {
				mConns.Add(-1)
				mRejected.Add(1)
				writeError(w, http.StatusServiceUnavailable, "server busy: too many open connections")
				return
			}
			defer mConns.Add(-1)
			generate(w, r, chat)
		}
	}
	mux.HandleFunc("/metrics", func(w http.ResponseWriter, r *http.Request) {
		w.Header().Set("Content-Type", "text/plain; version=0.0.4")
		fmt.Fprintf(w, "llmd_requests_total %d\nllmd_request_errors_total %d\nllmd_requests_in_flight %d\n"+
			"llmd_prompt_tokens_total %d\nllmd_generated_tokens_total %d\nllmd_prompt_seconds_total %.6f\nllmd_generation_seconds_total %.6f\n"+
			"llmd_connections %d\nllmd_connections_rejected_total %d\nllmd_gpus %d\n",
			mRequests.Load(), mErrors.Load(), mInFlight.Load(), mPromptTokens.Load(), mEvalTokens.Load(),
			float64(mPromptNs.Load())/1e9, float64(mEvalNs.Load())/1e9, mConns.Load(), mRejected.Load(), len(gpus))
		for _, g := range gpus {
			fmt.Fprintf(w, "llmd_gpu_requests_total{gpu=\"%d\"} %d\nllmd_gpu_requests_in_flight{gpu=\"%d\"} %d\nllmd_gpu_outstanding_tokens{gpu=\"%d\"} %d\n"+
				"llmd_gpu_generated_tokens_total{gpu=\"%d\"} %d\nllmd_gpu_prompt_tokens_total{gpu=\"%d\"} %d\nllmd_gpu_prefix_reused_tokens_total{gpu=\"%d\"} %d\n",
				g.device, g.requests.Load(), g.device, g.inFlight.Load(), g.device, g.outstanding.Load(),
				g.device, g.genTokens.Load(), g.device, g.promptTokens.Load(), g.device, g.reusedTokens.Load())
		}
	})
	mux.HandleFunc("/api/generate", post(false))
	mux.HandleFunc("/api/chat", post(true))
	mux.HandleFunc("/suggest", suggest(post(false)))

	addr := getenv("LLMD_ADDR", getenv("OLLAMA_HOST", "127.0.0.1:11434"))
	log.Printf("llmd: %s serving model '%s' on %s, %d GPU(s), %d sequences each", C.GoString(C.b200_version()), model, addr, len(gpus), seqs)
	// a client that stops sending mid-request must not pin a goroutine for ever (no WriteTimeout: replies stream)
	srv := &http.Server{Addr: addr, Handler: mux, ReadHeaderTimeout: 30 * time.Second, ReadTimeout: 30 * time.Second}
	log.Fatal(srv.ListenAndServe())
}
