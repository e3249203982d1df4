// llmd.cpp — Ollama-compatible HTTP front end over the C ABI of libb200llm (include/b200llm.h).
//
// This is the runnable twin of go/cmd/llmd/main.go (the image has no Go toolchain): same routes, same JSON,
// same use of the C entry points.  It replaces the server that web/streamlit_app.py:91 talks to
// (`requests.post(f"{OLLAMA_URL}/api/generate", json={"model","prompt","stream":False}, timeout=60)`):
//   POST /api/generate   {"model","prompt","stream"?,"raw"?,"system"?,"template"?,"context"?,"keep_alive"?,"format"?,
//                         "options"?{temperature,top_k,top_p,repeat_penalty,repeat_last_n,seed,num_predict,num_ctx,stop}}
//   POST /api/chat       {"model","messages":[{"role","content"}],"stream"?,"keep_alive"?,"options"?}
//   GET  /               "Ollama is running"      GET /api/version   /api/tags   /api/ps   GET /metrics (Prometheus)
// Non-stream replies are one JSON object; stream replies are NDJSON chunks (Transfer-Encoding: chunked),
// one per token, then a final {"done":true,...} object with Ollama's *_duration / *_count statistics (ns).
// Errors are a non-200 status with {"error": "..."} — the reference shows "(LLM error)" for any of them
// (web/streamlit_app.py:99).
//
// ONE URL, N GPUs.  The reference knows a single OLLAMA_URL (web/streamlit_app.py:27, start_all.sh:30-39), so this
// process owns every GPU named in LLMD_GPUS: one engine (weights + KV pool + scheduler thread) per device, no
// collective between them.  A request goes to the engine with the fewest outstanding tokens, counting the prompt
// tokens an engine would NOT have to evaluate because a slot of it still caches the peer's chat history
// (b200_prefix_match) as already done — so a peer's successive suggestions find their prefix, and load stays even.
//
// Configuration is by environment, like the reference's own services (go/cmd/node/main.go:131-134): LLMD_ADDR
// (default 127.0.0.1:11434, Ollama's port, web/streamlit_app.py:27; OLLAMA_HOST honoured), LLMD_MODEL (llama3.1),
// LLMD_GGUF (path of a real model file; else random-init weights: LLMD_WEIGHTS bf16|q4_k_m, LLMD_SEED),
// LLMD_GPUS (comma list, default LLMD_GPU or 0), LLMD_MAX_CTX (KV room per sequence), LLMD_NUM_CTX (default num_ctx
// of a request, <= LLMD_MAX_CTX), LLMD_MAX_SEQS (concurrent sequences per GPU, continuous batching), LLMD_MAX_CONNS;
// LLMD_TINY=layers for a small random-init model in tests.
#include <arpa/inet.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <signal.h>
#include <sys/socket.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "../../include/b200llm.h"
#include "json.h"

using mini_json::Value;

// One engine per GPU; the counters feed the router and GET /metrics.
struct Gpu {
  int device = 0;
  b200_engine* e = nullptr;
  std::atomic<long long> outstanding{0};  // tokens still to evaluate or generate for the requests it is serving
  std::atomic<long long> in_flight{0}, requests{0}, gen_tokens{0}, prompt_tokens{0}, reused_tokens{0};
};
static std::vector<std::unique_ptr<Gpu>> g_gpus;
static b200_engine* g_engine = nullptr;  // g_gpus[0]->e: tokenizer / template / stop-id queries (identical on every GPU)
static b200_config g_cfg;
static std::string g_model = "llama3.1";
static int g_num_ctx_default = 0;  // LLMD_NUM_CTX (0: the engines' max_ctx)
static std::atomic<long long> g_conns{0};
static long long g_max_conns = 1024;
static std::atomic<unsigned> g_rr{0};
static std::atomic<long long> g_expires_unix{0};  // keep_alive bookkeeping for /api/ps (the weights stay resident)

// Prometheus counters (GET /metrics): what an operator watches next to the reference's start_all.sh processes
static std::atomic<long long> m_requests{0}, m_errors{0}, m_in_flight{0}, m_prompt_tokens{0}, m_eval_tokens{0}, m_prompt_ns{0}, m_eval_ns{0},
    m_rejected{0};

static std::string env(const char* k, const char* d) {
  const char* v = getenv(k);
  return v && *v ? v : d;
}

static std::string rfc3339(std::chrono::system_clock::time_point tp) {
  using namespace std::chrono;
  const time_t t = system_clock::to_time_t(tp);
  const long ns = (long)(duration_cast<nanoseconds>(tp.time_since_epoch()).count() % 1000000000LL);
  struct tm g;
  gmtime_r(&t, &g);
  char b[64];
  snprintf(b, sizeof b, "%04d-%02d-%02dT%02d:%02d:%02d.%09ldZ", g.tm_year + 1900, g.tm_mon + 1, g.tm_mday, g.tm_hour, g.tm_min, g.tm_sec, ns);
  return b;
}
static std::string now_rfc3339() { return rfc3339(std::chrono::system_clock::now()); }

// ---- chat framing --------------------------------------------------------------------------------------
// Which framing the model expects is read from the GGUF file's tokenizer.chat_template (a Jinja program; it is
// recognised by the control tokens it uses, not interpreted): Llama-3 headers, ChatML, or — for engines without a
// vocabulary of their own — Llama-3 when the id space holds the published Llama-3 control ids, else plain text.
enum Framing { FRAME_PLAIN, FRAME_LLAMA3, FRAME_CHATML };
static Framing g_framing = FRAME_PLAIN;
static std::vector<int32_t> g_stop_ids;
static int32_t g_bos = -1;

static int32_t special(const char* text, int32_t published) {
  const int id = b200_special_token(g_engine, text);
  if (id >= 0) return id;
  return published >= 0 && published < g_cfg.vocab ? published : -1;
}

static void detect_framing() {
  std::string tpl;
  size_t n = 0;
  b200_chat_template(g_engine, nullptr, 0, &n);
  if (n) {
    tpl.resize(n);
    if (b200_chat_template(g_engine, &tpl[0], n, &n) != 0) tpl.clear();
  }
  if (tpl.find("<|start_header_id|>") != std::string::npos) g_framing = FRAME_LLAMA3;
  else if (tpl.find("<|im_start|>") != std::string::npos) g_framing = FRAME_CHATML;
  else if (tpl.empty() && g_cfg.vocab > 128009) g_framing = FRAME_LLAMA3;  // synthetic full-size engine: Llama-3 id space
  else g_framing = FRAME_PLAIN;
  g_bos = g_framing == FRAME_LLAMA3 ? special("<|begin_of_text|>", 128000) : -1;
  int32_t ids[64];
  size_t ns = 0;
  if (b200_stop_ids(g_engine, ids, 64, &ns) == 0) g_stop_ids.assign(ids, ids + ns);
  if (g_framing == FRAME_LLAMA3)
    for (int32_t extra : {special("<|eot_id|>", 128009), special("<|eom_id|>", 128008), special("<|end_of_text|>", 128001)})
      if (extra >= 0 && std::find(g_stop_ids.begin(), g_stop_ids.end(), extra) == g_stop_ids.end()) g_stop_ids.push_back(extra);
  if (g_cfg.eos_token < 0) g_stop_ids.clear();  // benchmarks / tests: never stop early
}

struct PromptBuilder {
  std::vector<int32_t> ids;
  bool too_long = false;
  void text(const std::string& s) {
    if (s.empty()) return;
    std::vector<int32_t> tmp((size_t)g_cfg.max_ctx + 8);  // a text that cannot fit the context is refused untokenised
    size_t n = 0;
    const int rc = b200_tokenize(g_engine, s.data(), s.size(), 0, tmp.data(), tmp.size(), &n);
    if (rc == 0) ids.insert(ids.end(), tmp.begin(), tmp.begin() + n);
    else too_long = true;
  }
  // text in which control tokens may be written out ("<|eot_id|>"): custom templates and raw prompts
  void text_with_specials(const std::string& s) {
    size_t i = 0;
    while (i < s.size()) {
      const size_t open = s.find("<|", i);
      if (open == std::string::npos) break;
      const size_t close = s.find("|>", open + 2);
      if (close == std::string::npos) break;
      const std::string name = s.substr(open, close + 2 - open);
      int32_t id = b200_special_token(g_engine, name.c_str());
      if (id < 0 && g_framing == FRAME_LLAMA3) {
        static const std::pair<const char*, int32_t> known[] = {{"<|begin_of_text|>", 128000}, {"<|end_of_text|>", 128001},
            {"<|start_header_id|>", 128006}, {"<|end_header_id|>", 128007}, {"<|eom_id|>", 128008}, {"<|eot_id|>", 128009}};
        for (auto& k : known) if (name == k.first && k.second < g_cfg.vocab) id = k.second;
      }
      if (id < 0) { text(s.substr(i, close + 2 - i)); i = close + 2; continue; }  // not a control token: ordinary text
      text(s.substr(i, open - i));
      ids.push_back(id);
      i = close + 2;
    }
    text(s.substr(i));
  }
  void begin() { if (g_bos >= 0) ids.push_back(g_bos); }
  void turn(const std::string& role, const std::string& content) {
    if (g_framing == FRAME_LLAMA3) {
      ids.push_back(special("<|start_header_id|>", 128006));
      text(role);
      ids.push_back(special("<|end_header_id|>", 128007));
      text("\n\n");
      text(content);
      ids.push_back(special("<|eot_id|>", 128009));
    } else if (g_framing == FRAME_CHATML) {
      text_with_specials("<|im_start|>" + role + "\n");
      text(content);
      text_with_specials("<|im_end|>\n");
    } else {
      text(role + ": " + content + "\n");
    }
  }
  void open_assistant() {
    if (g_framing == FRAME_LLAMA3) {
      ids.push_back(special("<|start_header_id|>", 128006));
      text("assistant");
      ids.push_back(special("<|end_header_id|>", 128007));
      text("\n\n");
    } else if (g_framing == FRAME_CHATML) {
      text_with_specials("<|im_start|>assistant\n");
    } else {
      text("assistant: ");
    }
  }
};

// Ollama's `template` request field is a Go template.  The subset its own model templates use for a single turn is
// rendered here: {{ .System }}, {{ .Prompt }}, {{ .Response }} (rendering stops there), {{ if .System }} / {{ if .Prompt }}
// ... {{ end }}, with optional "-" trim markers.  Anything else is refused with a clear error rather than guessed at.
static bool render_template(const std::string& tpl, const std::string& system, const std::string& prompt, std::string* out, std::string* err) {
  std::vector<bool> emit_stack;  // one entry per open {{ if }}
  auto emitting = [&] { for (bool b : emit_stack) if (!b) return false; return true; };
  size_t i = 0;
  bool trim_next = false;
  while (i < tpl.size()) {
    const size_t open = tpl.find("{{", i);
    std::string lit = tpl.substr(i, open == std::string::npos ? std::string::npos : open - i);
    if (trim_next) { size_t k = 0; while (k < lit.size() && isspace((unsigned char)lit[k])) ++k; lit.erase(0, k); trim_next = false; }
    if (open == std::string::npos) { if (emitting()) *out += lit; break; }
    const size_t close = tpl.find("}}", open + 2);
    if (close == std::string::npos) { *err = "template: unterminated {{"; return false; }
    std::string act = tpl.substr(open + 2, close - open - 2);
    if (!act.empty() && act[0] == '-') { while (!lit.empty() && isspace((unsigned char)lit.back())) lit.pop_back(); act.erase(0, 1); }
    if (!act.empty() && act.back() == '-') { trim_next = true; act.pop_back(); }
    if (emitting()) *out += lit;
    size_t b = 0, e = act.size();
    while (b < e && isspace((unsigned char)act[b])) ++b;
    while (e > b && isspace((unsigned char)act[e - 1])) --e;
    act = act.substr(b, e - b);
    if (act == ".System") { if (emitting()) *out += system; }
    else if (act == ".Prompt") { if (emitting()) *out += prompt; }
    else if (act == ".Response") { if (emitting()) return true; }
    else if (act == "if .System") emit_stack.push_back(!system.empty());
    else if (act == "if .Prompt") emit_stack.push_back(!prompt.empty());
    else if (act == "end") { if (emit_stack.empty()) { *err = "template: {{ end }} without {{ if }}"; return false; } emit_stack.pop_back(); }
    else { *err = "template: unsupported action {{ " + act + " }} (supported: .System .Prompt .Response, if .System/.Prompt, end)"; return false; }
    i = close + 2;
  }
  if (!emit_stack.empty()) { *err = "template: {{ if }} without {{ end }}"; return false; }
  return true;
}

// Emits only complete UTF-8 sequences (a byte-level vocabulary can split a character across tokens).
struct Utf8Stream {
  std::string pend;
  std::string push(const char* p, size_t n) {
    pend.append(p, n);
    std::string out;
    size_t i = 0;
    while (i < pend.size()) {
      const unsigned char c = (unsigned char)pend[i];
      const size_t need = c < 0x80 ? 1 : (c >> 5) == 6 ? 2 : (c >> 4) == 14 ? 3 : (c >> 3) == 30 ? 4 : 0;
      if (need == 0) { out += "\xEF\xBF\xBD"; ++i; continue; }
      if (i + need > pend.size()) {
        bool ok = true;
        for (size_t j = i + 1; j < pend.size(); ++j) ok = ok && (((unsigned char)pend[j] & 0xC0) == 0x80);
        if (ok) break;  // wait for the rest of this character
        out += "\xEF\xBF\xBD"; ++i; continue;
      }
      bool ok = true;
      for (size_t j = 1; j < need; ++j) ok = ok && (((unsigned char)pend[i + j] & 0xC0) == 0x80);
      if (ok) { out.append(pend, i, need); i += need; }
      else { out += "\xEF\xBF\xBD"; ++i; }
    }
    pend.erase(0, i);
    return out;
  }
  std::string flush() {
    std::string out;
    for (size_t i = 0; i < pend.size(); ++i) out += "\xEF\xBF\xBD";
    pend.clear();
    return out;
  }
};

// One token -> its bytes.  The buffer grows to whatever the vocabulary entry needs (Llama-3 has entries longer than 64
// bytes: runs of spaces, dashes, '='); any other failure is an error, not an empty piece.
static bool detok(int32_t id, std::string* out, std::string* err) {
  char small[64];
  size_t n = 0;
  int rc = b200_detokenize(g_engine, &id, 1, small, sizeof small, &n);
  if (rc == 0) { out->assign(small, n); return true; }
  if (rc == B200_ERANGE && n > sizeof small) {
    out->resize(n);
    rc = b200_detokenize(g_engine, &id, 1, &(*out)[0], n, &n);
    if (rc == 0) { out->resize(n); return true; }
  }
  *err = b200_last_error();
  return false;
}

// ---- HTTP plumbing -------------------------------------------------------------------------------------
static bool send_all(int fd, const char* p, size_t n) {
  while (n) {
    const ssize_t w = send(fd, p, n, MSG_NOSIGNAL);
    if (w <= 0) return false;
    p += w;
    n -= (size_t)w;
  }
  return true;
}
static bool send_str(int fd, const std::string& s) { return send_all(fd, s.data(), s.size()); }

static void reply(int fd, int code, const char* ctype, const std::string& body) {
  const char* reason = code == 200 ? "OK" : code == 400 ? "Bad Request" : code == 404 ? "Not Found" : code == 405 ? "Method Not Allowed"
                     : code == 503 ? "Service Unavailable" : "Internal Server Error";
  char h[256];
  snprintf(h, sizeof h, "HTTP/1.1 %d %s\r\nContent-Type: %s\r\nContent-Length: %zu\r\nConnection: close\r\n\r\n", code, reason, ctype, body.size());
  send_str(fd, h);
  send_str(fd, body);
}
static void reply_error(int fd, int code, const std::string& msg) { reply(fd, code, "application/json", "{\"error\":" + mini_json::quote(msg) + "}"); }
static bool send_chunk(int fd, const std::string& s) {
  char h[32];
  snprintf(h, sizeof h, "%zx\r\n", s.size());
  return send_str(fd, h) && send_str(fd, s) && send_str(fd, "\r\n");
}

struct HttpRequest {
  std::string method, path, body;
};
static bool read_request(int fd, HttpRequest* rq) {
  std::string buf;
  char tmp[8192];
  size_t hdr_end = std::string::npos;
  while (hdr_end == std::string::npos) {
    const ssize_t n = recv(fd, tmp, sizeof tmp, 0);
    if (n <= 0) return false;
    buf.append(tmp, (size_t)n);
    hdr_end = buf.find("\r\n\r\n");
    if (buf.size() > (1u << 20)) return false;
  }
  const size_t sp1 = buf.find(' '), sp2 = buf.find(' ', sp1 + 1);
  if (sp1 == std::string::npos || sp2 == std::string::npos) return false;
  rq->method = buf.substr(0, sp1);
  rq->path = buf.substr(sp1 + 1, sp2 - sp1 - 1);
  const size_t q = rq->path.find('?');
  if (q != std::string::npos) rq->path.resize(q);
  size_t clen = 0;
  {
    std::string low = buf.substr(0, hdr_end);
    for (auto& c : low) c = (char)tolower((unsigned char)c);
    const size_t p = low.find("content-length:");
    if (p != std::string::npos) clen = (size_t)strtoull(low.c_str() + p + 15, nullptr, 10);
    if (low.find("expect: 100-continue") != std::string::npos) send_str(fd, "HTTP/1.1 100 Continue\r\n\r\n");
  }
  if (clen > (64u << 20)) return false;
  rq->body = buf.substr(hdr_end + 4);
  while (rq->body.size() < clen) {
    const ssize_t n = recv(fd, tmp, sizeof tmp, 0);
    if (n <= 0) return false;
    rq->body.append(tmp, (size_t)n);
  }
  rq->body.resize(clen);
  return true;
}

// ---- request options -----------------------------------------------------------------------------------
struct GenOptions {
  b200_sampling sp;
  int num_ctx = 0;                 // context window of this request (<= the engines' max_ctx)
  std::vector<std::string> stop;   // stop sequences: generation ends before the first one that appears
};

static bool fill_options(const Value& req, GenOptions* go, std::string* err) {
  b200_sampling* sp = &go->sp;
  b200_default_sampling(sp);          // Ollama defaults: what the reference gets, it sends no options
  sp->num_predict = -1;               // Ollama: generate until EOS or the context is full
  sp->seed = (uint64_t)std::chrono::steady_clock::now().time_since_epoch().count();  // default seed is random
  go->num_ctx = g_num_ctx_default > 0 ? g_num_ctx_default : g_cfg.max_ctx;
  const Value* o = req.get("options");
  if (!o || o->kind != Value::Obj) return true;
  sp->temperature = (float)o->get_num("temperature", sp->temperature);
  sp->top_p = (float)o->get_num("top_p", sp->top_p);
  sp->repeat_penalty = (float)o->get_num("repeat_penalty", sp->repeat_penalty);
  sp->top_k = (int32_t)o->get_num("top_k", sp->top_k);
  sp->repeat_last_n = (int32_t)o->get_num("repeat_last_n", sp->repeat_last_n);
  sp->num_predict = (int32_t)o->get_num("num_predict", sp->num_predict);
  if (o->get("seed")) sp->seed = (uint64_t)(int64_t)o->get_num("seed", 0);
  if (const Value* nc = o->get("num_ctx")) {
    if (nc->kind != Value::Num || nc->num < 4) { *err = "options.num_ctx must be a number >= 4"; return false; }
    // a window larger than the KV room this server was started with cannot be honoured: say so instead of
    // silently answering from a shorter context
    if (nc->num > g_cfg.max_ctx) { *err = "options.num_ctx " + std::to_string((long long)nc->num) + " exceeds this server's LLMD_MAX_CTX " + std::to_string(g_cfg.max_ctx); return false; }
    go->num_ctx = (int)nc->num;
  }
  if (const Value* st = o->get("stop")) {
    if (st->kind == Value::Str) { if (!st->str.empty()) go->stop.push_back(st->str); }
    else if (st->kind == Value::Arr) {
      for (const Value& v : st->arr) {
        if (v.kind != Value::Str) { *err = "options.stop must be a string or an array of strings"; return false; }
        if (!v.str.empty()) go->stop.push_back(v.str);
      }
    } else if (st->kind != Value::Null) { *err = "options.stop must be a string or an array of strings"; return false; }
  }
  return true;
}

// keep_alive: a number of seconds or a Go duration string ("5m", "1h30m", "-1", "0").  Returns false on a malformed
// value.  The weights of this server stay resident for its lifetime (a B200 dedicated to llmd has nothing to make room
// for); the value is validated, answers load / unload probes as Ollama does, and sets expires_at in /api/ps.
static bool parse_keep_alive(const Value* v, double* seconds, std::string* err) {
  *seconds = 300;  // Ollama's default: five minutes
  if (!v || v->kind == Value::Null) return true;
  if (v->kind == Value::Num) { *seconds = v->num; return true; }
  if (v->kind != Value::Str) { *err = "keep_alive must be a number of seconds or a duration string"; return false; }
  const std::string& s = v->str;
  size_t i = 0;
  double total = 0, sign = 1;
  if (i < s.size() && (s[i] == '-' || s[i] == '+')) { sign = s[i] == '-' ? -1 : 1; ++i; }
  if (i >= s.size()) { *err = "keep_alive: empty duration"; return false; }
  bool any = false;
  while (i < s.size()) {
    char* endp = nullptr;
    const double val = strtod(s.c_str() + i, &endp);
    if (endp == s.c_str() + i) { *err = "keep_alive: malformed duration '" + s + "'"; return false; }
    i = (size_t)(endp - s.c_str());
    double unit = 1;
    if (s.compare(i, 2, "ms") == 0) { unit = 1e-3; i += 2; }
    else if (s.compare(i, 2, "us") == 0) { unit = 1e-6; i += 2; }
    else if (s.compare(i, 2, "ns") == 0) { unit = 1e-9; i += 2; }
    else if (i < s.size() && s[i] == 's') { unit = 1; ++i; }
    else if (i < s.size() && s[i] == 'm') { unit = 60; ++i; }
    else if (i < s.size() && s[i] == 'h') { unit = 3600; ++i; }
    else if (i == s.size() && !any) { unit = 1; }  // a bare number in a string: seconds
    else { *err = "keep_alive: unknown unit in '" + s + "'"; return false; }
    total += val * unit;
    any = true;
  }
  *seconds = sign * total;
  return true;
}

static std::string stats_json(const b200_stats& st, const char* reason) {
  char b[384];
  snprintf(b, sizeof b,
           "\"done\":true,\"done_reason\":\"%s\",\"total_duration\":%lld,\"load_duration\":%lld,\"prompt_eval_count\":%d,"
           "\"prompt_eval_duration\":%lld,\"eval_count\":%d,\"eval_duration\":%lld",
           reason, (long long)st.total_ns, (long long)st.load_ns, st.prompt_eval_count, (long long)st.prompt_eval_ns, st.eval_count,
           (long long)st.eval_ns);
  return b;
}

// The engine whose queue is shortest once the tokens it already caches for this prompt are taken off the bill.
static Gpu* route(const std::vector<int32_t>& ids, long long new_tokens, int* reused) {
  Gpu* best = nullptr;
  long long best_cost = 0;
  int best_match = 0;
  const unsigned start = g_rr.fetch_add(1);
  for (size_t k = 0; k < g_gpus.size(); ++k) {
    Gpu* g = g_gpus[(start + k) % g_gpus.size()].get();  // rotating start: ties spread round-robin
    const int match = g_gpus.size() > 1 ? b200_prefix_match(g->e, ids.data(), ids.size()) : 0;
    const long long cost = g->outstanding.load() + (long long)ids.size() + new_tokens - match;
    if (!best || cost < best_cost) { best = g; best_cost = cost; best_match = match; }
  }
  *reused = best_match;
  return best;
}

// ---- /api/generate and /api/chat ----------------------------------------------------------------------
// The reply text as the reference keeps it: Python's `str.strip()` of the response (web/streamlit_app.py:98), i.e.
// without leading / trailing characters for which str.isspace() holds: space, \t \n \v \f \r, 0x1c-0x1f, U+0085,
// U+00A0, U+1680, U+2000-200A, U+2028, U+2029, U+202F, U+205F, U+3000 — and nothing else (other control bytes stay).
static size_t py_space_at(const std::string& s, size_t i) {  // bytes of a whitespace character starting at i, or 0
  const unsigned char c = (unsigned char)s[i];
  if (c == ' ' || (c >= 0x09 && c <= 0x0d) || (c >= 0x1c && c <= 0x1f)) return 1;
  auto b = [&](size_t k) { return i + k < s.size() ? (unsigned char)s[i + k] : 0u; };
  if (c == 0xC2 && (b(1) == 0x85 || b(1) == 0xA0)) return 2;
  if (c == 0xE1 && b(1) == 0x9A && b(2) == 0x80) return 3;
  if (c == 0xE2 && b(1) == 0x80 && ((b(2) >= 0x80 && b(2) <= 0x8A) || b(2) == 0xA8 || b(2) == 0xA9 || b(2) == 0xAF)) return 3;
  if (c == 0xE2 && b(1) == 0x81 && b(2) == 0x9F) return 3;
  if (c == 0xE3 && b(1) == 0x80 && b(2) == 0x80) return 3;
  return 0;
}
static std::string strip_ws(const std::string& s) {
  size_t a = 0, b = s.size();
  for (size_t n; a < b && (n = py_space_at(s, a)) != 0;) a += n;
  for (;;) {  // from the end: a whitespace character is 1-3 bytes long
    bool cut = false;
    for (size_t n = 1; n <= 3 && n <= b - a; ++n)
      if (py_space_at(s, b - n) == n) { b -= n; cut = true; break; }
    if (!cut) break;
  }
  return s.substr(a, b - a);
}

// suggest = true: the request came through POST /suggest (below); same path, another reply shape
static void handle_generate(int fd, const Value& req, bool chat, bool suggest = false) {
  const std::string model = req.get_str("model");
  if (model != g_model && model != g_model + ":latest") {
    reply_error(fd, 404, "model '" + model + "' not found");
    return;
  }
  const bool stream = req.get_bool("stream", true);  // Ollama's default is to stream
  std::string err;
  GenOptions go;
  if (!fill_options(req, &go, &err)) { reply_error(fd, 400, err); return; }
  double keep_alive_s = 300;
  if (!parse_keep_alive(req.get("keep_alive"), &keep_alive_s, &err)) { reply_error(fd, 400, err); return; }
  g_expires_unix = keep_alive_s < 0 ? -1 : (long long)time(nullptr) + (long long)keep_alive_s;
  if (const Value* f = req.get("format")) {
    const bool empty = f->kind == Value::Null || (f->kind == Value::Str && f->str.empty());
    if (!empty) { reply_error(fd, 400, "format: constrained (JSON / schema) output is not supported by this server"); return; }
  }
  const std::string mq = mini_json::quote(g_model);

  PromptBuilder pb;
  bool have_prompt = false;
  if (chat) {
    const Value* msgs = req.get("messages");
    if (msgs && msgs->kind != Value::Arr) { reply_error(fd, 400, "messages must be an array"); return; }
    pb.begin();
    if (msgs)
      for (const Value& m : msgs->arr) { pb.turn(m.get_str("role", "user"), m.get_str("content")); have_prompt = true; }
    pb.open_assistant();
  } else {
    const std::string prompt = req.get_str("prompt"), sys = req.get_str("system");
    have_prompt = !prompt.empty();
    // `context`: the token ids a previous reply returned; the conversation so far, verbatim
    if (const Value* ctx = req.get("context")) {
      if (ctx->kind != Value::Arr && ctx->kind != Value::Null) { reply_error(fd, 400, "context must be an array of token ids"); return; }
      for (const Value& v : ctx->arr) {
        if (v.kind != Value::Num || v.num < 0 || v.num >= g_cfg.vocab || v.num != (double)(int64_t)v.num) { reply_error(fd, 400, "context holds an invalid token id"); return; }
        pb.ids.push_back((int32_t)v.num);
      }
    }
    const bool fresh = pb.ids.empty();
    const Value* tpl = req.get("template");
    if (req.get_bool("raw", false)) {
      pb.text_with_specials(prompt);
    } else if (tpl && tpl->kind == Value::Str && !tpl->str.empty()) {
      std::string rendered;
      if (!render_template(tpl->str, sys, prompt, &rendered, &err)) { reply_error(fd, 400, err); return; }
      if (fresh && rendered.compare(0, 17, "<|begin_of_text|>") != 0) pb.begin();
      pb.text_with_specials(rendered);
    } else {
      if (fresh) pb.begin();
      if (!sys.empty()) pb.turn("system", sys);
      pb.turn("user", prompt);
      pb.open_assistant();
    }
  }
  if (pb.too_long) { reply_error(fd, 400, "prompt does not fit the context window"); return; }
  if (!have_prompt && !chat) {
    // Ollama's load / unload probe: an empty prompt only (re)sets the residency timer
    reply(fd, 200, "application/json", "{\"model\":" + mq + ",\"created_at\":\"" + now_rfc3339() + "\",\"response\":\"\",\"done\":true,\"done_reason\":\"" +
                                           (keep_alive_s == 0 ? "unload" : "load") + "\"}");
    return;
  }
  if (chat && !have_prompt) {
    reply(fd, 200, "application/json", "{\"model\":" + mq + ",\"created_at\":\"" + now_rfc3339() +
                                           "\",\"message\":{\"role\":\"assistant\",\"content\":\"\"},\"done\":true,\"done_reason\":\"" + (keep_alive_s == 0 ? "unload" : "load") + "\"}");
    return;
  }
  if (pb.ids.empty()) pb.ids.push_back(0);
  // num_ctx: prompt + reply live in a window of num_ctx tokens.  A longer prompt is cut the way Ollama cuts it — the
  // first token (BOS) stays, the oldest tokens after it go — leaving room for the reply (this server does not shift
  // the context while generating: the reply ends with done_reason "length" when the window is full).
  const int limit = go.num_ctx;
  {
    const int want_out = go.sp.num_predict > 0 ? std::min(go.sp.num_predict, limit / 2) : std::max(1, limit / 4);
    const int max_prompt = std::max(1, limit - want_out);
    if ((int)pb.ids.size() > max_prompt) {
      const int keep = (g_bos >= 0 && pb.ids[0] == g_bos) ? 1 : 0;
      const size_t drop = pb.ids.size() - (size_t)max_prompt;
      pb.ids.erase(pb.ids.begin() + keep, pb.ids.begin() + keep + (long)drop);
    }
  }
  const int room = limit - (int)pb.ids.size() + 1;
  if (room < 1) { reply_error(fd, 400, "prompt does not fit the context window"); return; }
  if (go.sp.num_predict <= 0 || go.sp.num_predict > room) go.sp.num_predict = room;

  int reused = 0;
  Gpu* gpu = route(pb.ids, go.sp.num_predict, &reused);
  const long long bill = (long long)pb.ids.size() - reused + go.sp.num_predict;
  b200_req* r = nullptr;
  ++m_requests;
  ++gpu->requests;
  if (b200_submit(gpu->e, pb.ids.data(), pb.ids.size(), &go.sp, &r) != 0) { ++m_errors; reply_error(fd, 500, b200_last_error()); return; }
  gpu->outstanding += bill;
  ++m_in_flight;
  ++gpu->in_flight;
  std::string suggestion_so_far;  // /suggest: everything streamed so far (the final object carries the stripped whole)
  auto piece_obj = [&](const std::string& piece, bool final, const b200_stats* st, const std::vector<int32_t>* ctx, const char* reason) {
    if (suggest) {
      suggestion_so_far += piece;
      if (!final) return "{\"delta\":" + mini_json::quote(piece) + ",\"done\":false}";
      return "{\"suggestion\":" + mini_json::quote(strip_ws(suggestion_so_far)) + "," + stats_json(*st, reason) + "}";
    }
    std::string o = "{\"model\":" + mq + ",\"created_at\":\"" + now_rfc3339() + "\",";
    if (chat) o += "\"message\":{\"role\":\"assistant\",\"content\":" + mini_json::quote(piece) + "},";
    else o += "\"response\":" + mini_json::quote(piece) + ",";
    if (final) {
      if (!chat && ctx) {
        o += "\"context\":[";
        for (size_t i = 0; i < ctx->size(); ++i) o += (i ? "," : "") + std::to_string((*ctx)[i]);
        o += "],";
      }
      o += stats_json(*st, reason);
    } else {
      o += "\"done\":false";
    }
    return o + "}";
  };
  if (stream) send_str(fd, "HTTP/1.1 200 OK\r\nContent-Type: application/x-ndjson\r\nTransfer-Encoding: chunked\r\nConnection: close\r\n\r\n");
  Utf8Stream u8;
  std::string full;       // the reply so far (complete characters only)
  size_t sent = 0;        // bytes of `full` already streamed
  size_t max_stop = 0;
  for (auto& s : go.stop) max_stop = std::max(max_stop, s.size());
  std::vector<int32_t> context(pb.ids);
  int rc = 0;
  bool alive = true, stopped = false;
  while (!stopped) {
    int32_t id;
    int done = 0;
    rc = b200_next(r, &id, &done, -1);
    if (rc != 0) break;
    const bool is_stop_id = std::find(g_stop_ids.begin(), g_stop_ids.end(), id) != g_stop_ids.end();
    if (!is_stop_id) context.push_back(id);
    std::string piece;
    if (!is_stop_id) {
      std::string raw;
      if (!detok(id, &raw, &err)) { rc = B200_EINVAL; b200_cancel(r); break; }
      piece = u8.push(raw.data(), raw.size());
    }
    if (done || is_stop_id) piece += u8.flush();
    const size_t before = full.size();
    full += piece;
    size_t safe = full.size();  // bytes that can be shown: everything, unless a stop sequence may be forming at the end
    if (!go.stop.empty()) {
      const size_t from = before > max_stop ? before - max_stop : 0;
      size_t hit = std::string::npos;
      for (auto& s : go.stop) hit = std::min(hit, full.find(s, from));
      if (hit != std::string::npos) {
        full.resize(hit);
        safe = hit;
        stopped = true;
      } else if (!done && !is_stop_id) {
        for (auto& s : go.stop)
          for (size_t k = std::min(s.size() - 1, full.size()); k > 0; --k)
            if (full.compare(full.size() - k, k, s, 0, k) == 0) { safe = std::min(safe, full.size() - k); break; }
      }
    }
    if (is_stop_id) stopped = true;
    if (stream && alive && safe > sent) {
      alive = send_chunk(fd, piece_obj(full.substr(sent, safe - sent), false, nullptr, nullptr, "") + "\n");
      sent = safe;
      if (!alive) { b200_cancel(r); break; }  // client went away: stop generating
    }
    if (stopped && !done) b200_cancel(r);  // a stop sequence / secondary stop id ended the reply early
    if (done) break;
  }
  b200_stats st;
  b200_stats_get(r, &st);
  if (rc != 0 && err.empty()) err = b200_last_error();
  b200_free(r);
  gpu->outstanding -= bill;
  --m_in_flight;
  --gpu->in_flight;
  if (rc) ++m_errors;
  m_prompt_tokens += st.prompt_eval_count; m_eval_tokens += st.eval_count;
  gpu->prompt_tokens += st.prompt_eval_count; gpu->gen_tokens += st.eval_count;
  gpu->reused_tokens += (long long)pb.ids.size() - st.prompt_eval_count;
  m_prompt_ns += st.load_ns + st.prompt_eval_ns; m_eval_ns += st.eval_ns;
  const char* reason = (stopped || st.done_reason == 1) ? "stop" : "length";
  if (rc != 0 && !stream) { reply_error(fd, 500, err); return; }
  if (stream) {
    if (alive) {
      if (rc == 0 && full.size() > sent) send_chunk(fd, piece_obj(full.substr(sent), false, nullptr, nullptr, "") + "\n");
      send_chunk(fd, (rc ? "{\"error\":" + mini_json::quote(err) + "}" : piece_obj("", true, &st, &context, reason)) + "\n");
      send_str(fd, "0\r\n\r\n");
    }
  } else {
    reply(fd, 200, "application/json", piece_obj(full, true, &st, &context, reason));
  }
}

static std::string metrics_text() {
  char b[1536];
  snprintf(b, sizeof b,
           "# TYPE llmd_requests_total counter\nllmd_requests_total %lld\n# TYPE llmd_request_errors_total counter\nllmd_request_errors_total %lld\n"
           "# TYPE llmd_requests_in_flight gauge\nllmd_requests_in_flight %lld\n# TYPE llmd_prompt_tokens_total counter\nllmd_prompt_tokens_total %lld\n"
           "# TYPE llmd_generated_tokens_total counter\nllmd_generated_tokens_total %lld\n# TYPE llmd_prompt_seconds_total counter\n"
           "llmd_prompt_seconds_total %.6f\n# TYPE llmd_generation_seconds_total counter\nllmd_generation_seconds_total %.6f\n"
           "# TYPE llmd_connections gauge\nllmd_connections %lld\n# TYPE llmd_connections_rejected_total counter\nllmd_connections_rejected_total %lld\n"
           "# TYPE llmd_gpus gauge\nllmd_gpus %zu\n",
           m_requests.load(), m_errors.load(), m_in_flight.load(), m_prompt_tokens.load(), m_eval_tokens.load(), m_prompt_ns.load() / 1e9,
           m_eval_ns.load() / 1e9, g_conns.load(), m_rejected.load(), g_gpus.size());
  std::string out = b;
  out += "# TYPE llmd_gpu_requests_total counter\n# TYPE llmd_gpu_requests_in_flight gauge\n# TYPE llmd_gpu_outstanding_tokens gauge\n"
         "# TYPE llmd_gpu_generated_tokens_total counter\n# TYPE llmd_gpu_prompt_tokens_total counter\n# TYPE llmd_gpu_prefix_reused_tokens_total counter\n";
  for (auto& g : g_gpus) {
    snprintf(b, sizeof b,
             "llmd_gpu_requests_total{gpu=\"%d\"} %lld\nllmd_gpu_requests_in_flight{gpu=\"%d\"} %lld\nllmd_gpu_outstanding_tokens{gpu=\"%d\"} %lld\n"
             "llmd_gpu_generated_tokens_total{gpu=\"%d\"} %lld\nllmd_gpu_prompt_tokens_total{gpu=\"%d\"} %lld\nllmd_gpu_prefix_reused_tokens_total{gpu=\"%d\"} %lld\n",
             g->device, g->requests.load(), g->device, g->in_flight.load(), g->device, g->outstanding.load(), g->device, g->gen_tokens.load(),
             g->device, g->prompt_tokens.load(), g->device, g->reused_tokens.load());
    out += b;
  }
  return out;
}

static std::string model_entry(bool with_expiry) {
  const char* q = g_cfg.weight_format == B200_WEIGHTS_Q4_K_M ? "Q4_K_M" : "BF16";
  std::string o = "{\"name\":" + mini_json::quote(g_model + ":latest") + ",\"model\":" + mini_json::quote(g_model + ":latest") + ",\"size\":" +
                  std::to_string((long long)b200_weight_bytes(g_engine)) + ",\"details\":{\"format\":\"gguf\",\"family\":\"llama\",\"quantization_level\":\"" + q + "\"}";
  if (with_expiry) {
    const long long ex = g_expires_unix.load();
    const auto tp = ex < 0 ? std::chrono::system_clock::time_point::max() - std::chrono::hours(24 * 365)
                           : std::chrono::system_clock::from_time_t((time_t)(ex > 0 ? ex : time(nullptr) + 300));
    o += ",\"expires_at\":\"" + rfc3339(tp) + "\",\"size_vram\":" + std::to_string((long long)b200_weight_bytes(g_engine) * (long long)g_gpus.size());
  }
  return o + "}";
}

static void serve(int fd) {
  HttpRequest rq;
  if (read_request(fd, &rq)) {
    if (rq.method == "GET" || rq.method == "HEAD") {
      if (rq.path == "/") reply(fd, 200, "text/plain; charset=utf-8", "Ollama is running");
      else if (rq.path == "/api/version") reply(fd, 200, "application/json", std::string("{\"version\":\"0.5.7-b200llm\",\"engine\":") + mini_json::quote(b200_version()) + "}");
      else if (rq.path == "/metrics") reply(fd, 200, "text/plain; version=0.0.4", metrics_text());
      else if (rq.path == "/api/tags") reply(fd, 200, "application/json", "{\"models\":[" + model_entry(false) + "]}");
      else if (rq.path == "/api/ps") reply(fd, 200, "application/json", "{\"models\":[" + model_entry(true) + "]}");
      else reply_error(fd, 404, "not found");
    } else if (rq.method == "POST" && (rq.path == "/api/generate" || rq.path == "/api/chat")) {
      Value req;
      if (!mini_json::Parser(rq.body).parse(&req) || req.kind != Value::Obj) reply_error(fd, 400, "invalid JSON body");
      else handle_generate(fd, req, rq.path == "/api/chat");
    } else if (rq.method == "POST" && rq.path == "/suggest") {
      // what a node-side `/suggest` next to `/send` (reference go/cmd/node/main.go:219) forwards to.
      // {"content": <the peer's message>, "stream": bool (default false, like the reference's blocking call), "options"?} ->
      // the reference's own prompt (web/streamlit_app.py:93) and model; reply {"suggestion": <stripped text>, ...stats}
      // or, streaming, NDJSON {"delta": ...,"done":false} lines and a final object with the whole suggestion.
      Value body;
      if (!mini_json::Parser(rq.body).parse(&body) || body.kind != Value::Obj || !body.get("content") || body.get("content")->kind != Value::Str) {
        reply_error(fd, 400, "body must be {\"content\": string, \"stream\": bool}");
      } else {
        Value req, v;
        req.kind = Value::Obj;
        v.kind = Value::Str; v.str = g_model;
        req.obj.emplace_back("model", v);
        v.str = "You are a helpful assistant. Draft a concise, friendly reply to the following message:\n\n" + body.get_str("content") + "\n\nReply:";
        req.obj.emplace_back("prompt", v);
        v = Value(); v.kind = Value::Bool; v.b = body.get_bool("stream", false);
        req.obj.emplace_back("stream", v);
        if (const Value* o = body.get("options")) req.obj.emplace_back("options", *o);  // same meaning as in /api/generate
        handle_generate(fd, req, false, true);
      }
    } else {
      reply_error(fd, rq.method == "POST" ? 404 : 405, "not found");
    }
  }
  shutdown(fd, SHUT_RDWR);
  close(fd);
  --g_conns;
}

int main() {
  signal(SIGPIPE, SIG_IGN);
  g_model = env("LLMD_MODEL", "llama3.1");
  b200_llama31_8b(&g_cfg);
  const int tiny = atoi(env("LLMD_TINY", "0").c_str());
  if (tiny > 0) {  // small random-init model of the same architecture (tests)
    g_cfg.n_layers = tiny; g_cfg.d_model = 2048; g_cfg.n_heads = 16; g_cfg.n_kv_heads = 4; g_cfg.d_ffn = 2048; g_cfg.vocab = 1024;
    g_cfg.rope_orig_ctx = 64; g_cfg.eos_token = -1;
  }
  if (getenv("LLMD_EOS")) g_cfg.eos_token = atoi(env("LLMD_EOS", "-1").c_str());  // benchmarks: -1 = never stop early
  g_cfg.weight_format = env("LLMD_WEIGHTS", "bf16") == "q4_k_m" ? B200_WEIGHTS_Q4_K_M : B200_WEIGHTS_BF16;
  g_cfg.max_ctx = atoi(env("LLMD_MAX_CTX", "2048").c_str());
  g_cfg.seed = atoll(env("LLMD_SEED", "0").c_str());
  g_num_ctx_default = atoi(env("LLMD_NUM_CTX", "0").c_str());
  g_max_conns = atoll(env("LLMD_MAX_CONNS", "1024").c_str());
  const std::string gguf = env("LLMD_GGUF", "");
  const int max_seqs = atoi(env("LLMD_MAX_SEQS", "8").c_str());
  // LLMD_GPUS=0,1,...: one engine per device, all behind this one address
  std::vector<int> devices;
  {
    const std::string list = env("LLMD_GPUS", env("LLMD_GPU", "0").c_str());
    size_t i = 0;
    while (i < list.size()) {
      size_t j = list.find(',', i);
      if (j == std::string::npos) j = list.size();
      if (j > i) devices.push_back(atoi(list.substr(i, j - i).c_str()));
      i = j + 1;
    }
    if (devices.empty()) devices.push_back(0);
  }
  // the engines load side by side (each synthesises / uploads its own copy of the weights)
  std::vector<int> rcs(devices.size(), 0);
  std::vector<std::string> errs(devices.size());
  std::vector<b200_config> cfgs(devices.size(), g_cfg);
  for (size_t i = 0; i < devices.size(); ++i) {
    g_gpus.emplace_back(new Gpu());
    g_gpus[i]->device = devices[i];
  }
  std::vector<std::thread> loaders;
  for (size_t i = 0; i < devices.size(); ++i)
    loaders.emplace_back([&, i] {
      if (!gguf.empty()) {  // a real model file, e.g. the blob `ollama pull llama3.1` stored (README.md:70)
        rcs[i] = b200_load_gguf(gguf.c_str(), devices[i], g_cfg.max_ctx, max_seqs, &g_gpus[i]->e);
        if (rcs[i] == 0) b200_get_config(g_gpus[i]->e, &cfgs[i]);
      } else {
        cfgs[i].device = devices[i];
        cfgs[i].max_seqs = max_seqs;
        rcs[i] = b200_init(&cfgs[i], &g_gpus[i]->e);
      }
      if (rcs[i] != 0) errs[i] = b200_last_error();
    });
  for (auto& t : loaders) t.join();
  for (size_t i = 0; i < devices.size(); ++i)
    if (rcs[i] != 0) {
      fprintf(stderr, "llmd: engine init failed on GPU %d: %s\n", devices[i], errs[i].c_str());
      return 1;
    }
  g_cfg = cfgs[0];
  if (getenv("LLMD_EOS")) g_cfg.eos_token = atoi(env("LLMD_EOS", "-1").c_str());
  g_engine = g_gpus[0]->e;
  detect_framing();
  std::string addr = env("LLMD_ADDR", env("OLLAMA_HOST", "127.0.0.1:11434").c_str());
  const size_t colon = addr.rfind(':');
  const std::string host = colon == std::string::npos ? addr : addr.substr(0, colon);
  const int port = colon == std::string::npos ? 11434 : atoi(addr.c_str() + colon + 1);
  const int ls = socket(AF_INET, SOCK_STREAM, 0);
  int one = 1;
  setsockopt(ls, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
  sockaddr_in sa = {};
  sa.sin_family = AF_INET;
  sa.sin_port = htons((uint16_t)port);
  if (inet_pton(AF_INET, host.c_str(), &sa.sin_addr) != 1) sa.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
  if (bind(ls, (sockaddr*)&sa, sizeof sa) != 0 || listen(ls, 1024) != 0) {
    perror("llmd: bind/listen");
    return 1;
  }
  fprintf(stderr, "llmd: %s serving model '%s' on %s:%d, %zu GPU(s), %d sequences each (%lld weight bytes per token)\n", b200_version(),
          g_model.c_str(), host.c_str(), port, g_gpus.size(), max_seqs, (long long)b200_weight_bytes(g_engine));
  fflush(stderr);
  for (;;) {
    const int fd = accept(ls, nullptr, nullptr);
    if (fd < 0) continue;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
    timeval tv = {30, 0};  // a client that stops sending mid-request must not pin a thread for ever
    setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof tv);
    // one thread per connection, bounded: beyond LLMD_MAX_CONNS a connection is turned away at once (503) instead of
    // piling up threads — 256 peers with one blocking request each is the design load (BASELINE config 4)
    if (g_conns.load() >= g_max_conns) {
      ++m_rejected;
      reply_error(fd, 503, "server busy: too many open connections");
      close(fd);
      continue;
    }
    ++g_conns;
    std::thread(serve, fd).detach();
  }
}
