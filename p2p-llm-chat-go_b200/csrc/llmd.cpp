// llmd.cpp — Ollama-compatible HTTP front end over the C ABI of libb200llm (include/b200llm.h).
//
// This is the runnable twin of go/cmd/llmd/main.go (the image has no Go toolchain): same routes, same JSON,
// same use of the C entry points.  It replaces the server that web/streamlit_app.py:91 talks to
// (`requests.post(f"{OLLAMA_URL}/api/generate", json={"model","prompt","stream":False}, timeout=60)`):
//   POST /api/generate   {"model","prompt","stream"?,"raw"?,"system"?,"options"?{temperature,top_k,top_p,
//                         repeat_penalty,repeat_last_n,seed,num_predict}}
//   POST /api/chat       {"model","messages":[{"role","content"}],"stream"?,"options"?}
//   GET  /               "Ollama is running"      GET /api/version      GET /api/tags      GET /metrics (Prometheus)
// Non-stream replies are one JSON object; stream replies are NDJSON chunks (Transfer-Encoding: chunked),
// one per token, then a final {"done":true,...} object with Ollama's *_duration / *_count statistics (ns).
// Errors are a non-200 status with {"error": "..."} — the reference shows "(LLM error)" for any of them
// (web/streamlit_app.py:99).  Configuration is by environment, like the reference's own services
// (go/cmd/node/main.go:131-134): LLMD_ADDR (default 127.0.0.1:11434, Ollama's port, web/streamlit_app.py:27),
// LLMD_MODEL (llama3.1), LLMD_GGUF (path of a real model file; else random-init weights: LLMD_WEIGHTS
// bf16|q4_k_m, LLMD_SEED), LLMD_GPU, LLMD_MAX_CTX, LLMD_MAX_SEQS (concurrent sequences, continuous batching);
// LLMD_TINY=layers for a small random-init model in tests.
#include <arpa/inet.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <signal.h>
#include <sys/socket.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <string>
#include <thread>
#include <vector>

#include "../../include/b200llm.h"
#include "json.h"

using mini_json::Value;

static b200_engine* g_engine = nullptr;
static b200_config g_cfg;
static std::string g_model = "llama3.1";

// Prometheus counters (GET /metrics): what an operator watches next to the reference's start_all.sh processes
static std::atomic<long long> m_requests{0}, m_errors{0}, m_in_flight{0}, m_prompt_tokens{0}, m_eval_tokens{0}, m_prompt_ns{0}, m_eval_ns{0};

static std::string env(const char* k, const char* d) {
  const char* v = getenv(k);
  return v && *v ? v : d;
}

static std::string now_rfc3339() {
  using namespace std::chrono;
  const auto tp = system_clock::now();
  const time_t t = system_clock::to_time_t(tp);
  const long ns = (long)(duration_cast<nanoseconds>(tp.time_since_epoch()).count() % 1000000000LL);
  struct tm g;
  gmtime_r(&t, &g);
  char b[64];
  snprintf(b, sizeof b, "%04d-%02d-%02dT%02d:%02d:%02d.%09ldZ", g.tm_year + 1900, g.tm_mon + 1, g.tm_mday, g.tm_hour, g.tm_min, g.tm_sec, ns);
  return b;
}

// ---- prompt construction -----------------------------------------------------------------------------
// Llama-3 chat framing, as Ollama's llama3.1 template does it.  Special ids are the published Llama-3 ones;
// a model whose vocabulary is too small to contain them (tiny test models) gets the text only.
struct PromptBuilder {
  std::vector<int32_t> ids;
  bool specials;
  explicit PromptBuilder(const b200_config& c) : specials(c.vocab > 128009) {
    if (specials) ids.push_back(128000);  // <|begin_of_text|>
  }
  void text(const std::string& s) {
    std::vector<int32_t> tmp(s.size() + 4);
    size_t n = 0;
    if (b200_tokenize(g_engine, s.data(), s.size(), 0, tmp.data(), tmp.size(), &n) == 0) ids.insert(ids.end(), tmp.begin(), tmp.begin() + n);
  }
  void turn(const std::string& role, const std::string& content) {
    if (specials) {
      ids.push_back(128006);  // <|start_header_id|>
      text(role);
      ids.push_back(128007);  // <|end_header_id|>
      text("\n\n");
      text(content);
      ids.push_back(128009);  // <|eot_id|>
    } else {
      text(role + ": " + content + "\n");
    }
  }
  void open_assistant() {
    if (specials) {
      ids.push_back(128006);
      text("assistant");
      ids.push_back(128007);
      text("\n\n");
    } else {
      text("assistant: ");
    }
  }
};

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
  const char* reason = code == 200 ? "OK" : code == 400 ? "Bad Request" : code == 404 ? "Not Found" : code == 405 ? "Method Not Allowed" : "Internal Server Error";
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

// ---- /api/generate and /api/chat ----------------------------------------------------------------------
static void fill_sampling(const Value& req, b200_sampling* sp) {
  b200_default_sampling(sp);          // Ollama defaults: what the reference gets, it sends no options
  sp->num_predict = -1;               // Ollama: generate until EOS or the context is full
  sp->seed = (uint64_t)std::chrono::steady_clock::now().time_since_epoch().count();  // default seed is random
  const Value* o = req.get("options");
  if (!o || o->kind != Value::Obj) return;
  sp->temperature = (float)o->get_num("temperature", sp->temperature);
  sp->top_p = (float)o->get_num("top_p", sp->top_p);
  sp->repeat_penalty = (float)o->get_num("repeat_penalty", sp->repeat_penalty);
  sp->top_k = (int32_t)o->get_num("top_k", sp->top_k);
  sp->repeat_last_n = (int32_t)o->get_num("repeat_last_n", sp->repeat_last_n);
  sp->num_predict = (int32_t)o->get_num("num_predict", sp->num_predict);
  if (o->get("seed")) sp->seed = (uint64_t)(int64_t)o->get_num("seed", 0);
}

static std::string stats_json(const b200_stats& st) {
  char b[384];
  snprintf(b, sizeof b,
           "\"done\":true,\"done_reason\":\"%s\",\"total_duration\":%lld,\"load_duration\":%lld,\"prompt_eval_count\":%d,"
           "\"prompt_eval_duration\":%lld,\"eval_count\":%d,\"eval_duration\":%lld",
           st.done_reason == 1 ? "stop" : "length", (long long)st.total_ns, (long long)st.load_ns, st.prompt_eval_count,
           (long long)st.prompt_eval_ns, st.eval_count, (long long)st.eval_ns);
  return b;
}

static void handle_generate(int fd, const Value& req, bool chat) {
  const std::string model = req.get_str("model");
  if (model != g_model && model != g_model + ":latest") {
    reply_error(fd, 404, "model '" + model + "' not found");
    return;
  }
  const bool stream = req.get_bool("stream", true);  // Ollama's default is to stream
  PromptBuilder pb(g_cfg);
  if (chat) {
    const Value* msgs = req.get("messages");
    if (!msgs || msgs->kind != Value::Arr) { reply_error(fd, 400, "messages must be an array"); return; }
    for (const Value& m : msgs->arr) pb.turn(m.get_str("role", "user"), m.get_str("content"));
    pb.open_assistant();
  } else if (req.get_bool("raw", false)) {
    pb.text(req.get_str("prompt"));
  } else {
    const std::string sys = req.get_str("system");
    if (!sys.empty()) pb.turn("system", sys);
    pb.turn("user", req.get_str("prompt"));
    pb.open_assistant();
  }
  if (pb.ids.empty()) pb.ids.push_back(0);
  b200_sampling sp;
  fill_sampling(req, &sp);
  const int room = g_cfg.max_ctx - (int)pb.ids.size() + 1;
  if (room < 1) { reply_error(fd, 400, "prompt does not fit the context window"); return; }
  if (sp.num_predict <= 0 || sp.num_predict > room) sp.num_predict = room;

  b200_req* r = nullptr;
  ++m_requests;
  if (b200_submit(g_engine, pb.ids.data(), pb.ids.size(), &sp, &r) != 0) { ++m_errors; reply_error(fd, 500, b200_last_error()); return; }
  ++m_in_flight;
  const std::string mq = mini_json::quote(g_model);
  auto piece_obj = [&](const std::string& piece, bool final, const b200_stats* st, const std::vector<int32_t>* ctx) {
    std::string o = "{\"model\":" + mq + ",\"created_at\":\"" + now_rfc3339() + "\",";
    if (chat) o += "\"message\":{\"role\":\"assistant\",\"content\":" + mini_json::quote(piece) + "},";
    else o += "\"response\":" + mini_json::quote(piece) + ",";
    if (final) {
      if (!chat && ctx) {
        o += "\"context\":[";
        for (size_t i = 0; i < ctx->size(); ++i) o += (i ? "," : "") + std::to_string((*ctx)[i]);
        o += "],";
      }
      o += stats_json(*st);
    } else {
      o += "\"done\":false";
    }
    return o + "}";
  };
  if (stream) send_str(fd, "HTTP/1.1 200 OK\r\nContent-Type: application/x-ndjson\r\nTransfer-Encoding: chunked\r\nConnection: close\r\n\r\n");
  Utf8Stream u8;
  std::string full;
  std::vector<int32_t> context(pb.ids);
  int rc = 0;
  bool alive = true;
  for (;;) {
    int32_t id;
    int done = 0;
    rc = b200_next(r, &id, &done, -1);
    if (rc != 0) break;
    context.push_back(id);
    std::string piece;
    if (id != g_cfg.eos_token) {
      char buf[64];
      size_t n = 0;
      if (b200_detokenize(g_engine, &id, 1, buf, sizeof buf, &n) == 0) piece = u8.push(buf, n);
    }
    if (done) piece += u8.flush();
    full += piece;
    if (stream && alive && !piece.empty()) {
      alive = send_chunk(fd, piece_obj(piece, false, nullptr, nullptr) + "\n");
      if (!alive) b200_cancel(r);  // client went away: stop generating
    }
    if (done) break;
  }
  b200_stats st;
  b200_stats_get(r, &st);
  const std::string err = rc ? b200_last_error() : "";
  b200_free(r);
  --m_in_flight;
  if (rc) ++m_errors;
  m_prompt_tokens += st.prompt_eval_count; m_eval_tokens += st.eval_count;
  m_prompt_ns += st.load_ns + st.prompt_eval_ns; m_eval_ns += st.eval_ns;
  if (rc != 0 && !stream) { reply_error(fd, 500, err); return; }
  if (stream) {
    if (alive) {
      send_chunk(fd, (rc ? "{\"error\":" + mini_json::quote(err) + "}" : piece_obj("", true, &st, &context)) + "\n");
      send_str(fd, "0\r\n\r\n");
    }
  } else {
    reply(fd, 200, "application/json", piece_obj(full, true, &st, &context));
  }
}

static void serve(int fd) {
  HttpRequest rq;
  if (read_request(fd, &rq)) {
    if (rq.method == "GET" || rq.method == "HEAD") {
      if (rq.path == "/") reply(fd, 200, "text/plain; charset=utf-8", "Ollama is running");
      else if (rq.path == "/api/version") reply(fd, 200, "application/json", std::string("{\"version\":\"0.5.7-b200llm\",\"engine\":") + mini_json::quote(b200_version()) + "}");
      else if (rq.path == "/metrics") {
        char b[1024];
        snprintf(b, sizeof b,
                 "# TYPE llmd_requests_total counter\nllmd_requests_total %lld\n# TYPE llmd_request_errors_total counter\nllmd_request_errors_total %lld\n"
                 "# TYPE llmd_requests_in_flight gauge\nllmd_requests_in_flight %lld\n# TYPE llmd_prompt_tokens_total counter\nllmd_prompt_tokens_total %lld\n"
                 "# TYPE llmd_generated_tokens_total counter\nllmd_generated_tokens_total %lld\n# TYPE llmd_prompt_seconds_total counter\n"
                 "llmd_prompt_seconds_total %.6f\n# TYPE llmd_generation_seconds_total counter\nllmd_generation_seconds_total %.6f\n",
                 m_requests.load(), m_errors.load(), m_in_flight.load(), m_prompt_tokens.load(), m_eval_tokens.load(), m_prompt_ns.load() / 1e9,
                 m_eval_ns.load() / 1e9);
        reply(fd, 200, "text/plain; version=0.0.4", b);
      } else if (rq.path == "/api/tags") {
        const char* q = g_cfg.weight_format == B200_WEIGHTS_Q4_K_M ? "Q4_K_M" : "BF16";
        reply(fd, 200, "application/json",
              "{\"models\":[{\"name\":" + mini_json::quote(g_model + ":latest") + ",\"model\":" + mini_json::quote(g_model + ":latest") +
                  ",\"size\":" + std::to_string((long long)b200_weight_bytes(g_engine)) + ",\"details\":{\"format\":\"gguf\",\"family\":\"llama\",\"quantization_level\":\"" + q + "\"}}]}");
      } else reply_error(fd, 404, "not found");
    } else if (rq.method == "POST" && (rq.path == "/api/generate" || rq.path == "/api/chat")) {
      Value req;
      if (!mini_json::Parser(rq.body).parse(&req) || req.kind != Value::Obj) reply_error(fd, 400, "invalid JSON body");
      else handle_generate(fd, req, rq.path == "/api/chat");
    } else {
      reply_error(fd, rq.method == "POST" ? 404 : 405, "not found");
    }
  }
  shutdown(fd, SHUT_RDWR);
  close(fd);
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
  g_cfg.device = atoi(env("LLMD_GPU", "0").c_str());
  g_cfg.max_ctx = atoi(env("LLMD_MAX_CTX", "2048").c_str());
  g_cfg.seed = atoll(env("LLMD_SEED", "0").c_str());
  const std::string gguf = env("LLMD_GGUF", "");
  const int max_seqs = atoi(env("LLMD_MAX_SEQS", "8").c_str());
  int rc;
  if (!gguf.empty()) {  // a real model file, e.g. the blob `ollama pull llama3.1` stored (README.md:70)
    rc = b200_load_gguf(gguf.c_str(), g_cfg.device, g_cfg.max_ctx, max_seqs, &g_engine);
    if (rc == 0) b200_get_config(g_engine, &g_cfg);
  } else {
    g_cfg.max_seqs = max_seqs;
    rc = b200_init(&g_cfg, &g_engine);
  }
  if (rc != 0) {
    fprintf(stderr, "llmd: engine init failed: %s\n", b200_last_error());
    return 1;
  }
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
  if (bind(ls, (sockaddr*)&sa, sizeof sa) != 0 || listen(ls, 512) != 0) {
    perror("llmd: bind/listen");
    return 1;
  }
  fprintf(stderr, "llmd: %s serving model '%s' on %s:%d (%lld weight bytes per token)\n", b200_version(), g_model.c_str(), host.c_str(), port,
          (long long)b200_weight_bytes(g_engine));
  fflush(stderr);
  for (;;) {
    const int fd = accept(ls, nullptr, nullptr);
    if (fd < 0) continue;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
    timeval tv = {30, 0};  // a client that stops sending mid-request must not pin a thread for ever
    setsockopt(fd, SOL_SOCKET, SO_RCVTIMEO, &tv, sizeof tv);
    std::thread(serve, fd).detach();
  }
}
