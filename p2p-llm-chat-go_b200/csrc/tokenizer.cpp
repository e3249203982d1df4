// tokenizer.cpp — text <-> ids for the HTTP front end (see tokenizer.h).
//
// The reference sends text (web/streamlit_app.py:93) and Ollama tokenises it with the byte-level BPE
// vocabulary stored in the model's GGUF file.  When the engine was created from a GGUF file that vocabulary
// (tokens + merge ranks) is used; a synthetic random-init engine has no merge table and uses the base alphabet
// of the same scheme directly: ids 0..255 are the raw UTF-8 bytes (lossless round trip for any string).
#include "tokenizer.h"

#include <cstdio>
#include <algorithm>
#include <cstring>
#include <queue>

#include "../../include/b200llm.h"
#include "gguf.h"
#include "unicode_tables.h"

namespace b200 {
void set_error(const char* fmt, ...);

namespace {

// GPT-2 bytes_to_unicode: printable bytes map to themselves, the rest to U+0100.. in order.
struct ByteMap {
  int cp_of_byte[256];
  std::unordered_map<int, int> byte_of_cp;
  ByteMap() {
    int n = 0;
    for (int b = 0; b < 256; ++b) {
      const bool keep = (b >= 33 && b <= 126) || (b >= 161 && b <= 172) || (b >= 174 && b <= 255);
      cp_of_byte[b] = keep ? b : 256 + n++;
      byte_of_cp[cp_of_byte[b]] = b;
    }
  }
};
const ByteMap& bytemap() {
  static ByteMap m;
  return m;
}

// token string in the byte-level unicode alphabet -> raw bytes; false if it uses other code points
bool unicode_token_to_bytes(const std::string& s, std::string* out) {
  const ByteMap& m = bytemap();
  size_t i = 0;
  while (i < s.size()) {
    const unsigned char c = (unsigned char)s[i];
    int cp, len;
    if (c < 0x80) { cp = c; len = 1; }
    else if ((c >> 5) == 6 && i + 1 < s.size()) { cp = ((c & 0x1F) << 6) | (s[i + 1] & 0x3F); len = 2; }
    else if ((c >> 4) == 14 && i + 2 < s.size()) { cp = ((c & 0x0F) << 12) | ((s[i + 1] & 0x3F) << 6) | (s[i + 2] & 0x3F); len = 3; }
    else return false;
    auto it = m.byte_of_cp.find(cp);
    if (it == m.byte_of_cp.end()) return false;
    out->push_back((char)it->second);
    i += len;
  }
  return true;
}

// ---- code-point classes of the Llama-3 pre-tokeniser regex (what llama.cpp's unicode.cpp derives from its tables)
bool in_ranges(const uint32_t (*r)[2], int n, uint32_t cp) {
  int lo = 0, hi = n - 1;
  while (lo <= hi) {
    const int mid = (lo + hi) / 2;
    if (cp < r[mid][0]) hi = mid - 1;
    else if (cp > r[mid][1]) lo = mid + 1;
    else return true;
  }
  return false;
}
inline bool is_letter(uint32_t c) {  // \p{L}
  if (c < 0x80) return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z');
  return in_ranges(kUnicodeLetters, kUnicodeLetters_n, c);
}
inline bool is_digit(uint32_t c) {  // \p{N}
  if (c < 0x80) return c >= '0' && c <= '9';
  return in_ranges(kUnicodeNumbers, kUnicodeNumbers_n, c);
}
inline bool is_space(uint32_t c) {  // \s = White_Space
  return (c >= 0x9 && c <= 0xD) || c == 0x20 || c == 0x85 || c == 0xA0 || c == 0x1680 || (c >= 0x2000 && c <= 0x200A) || c == 0x2028 ||
         c == 0x2029 || c == 0x202F || c == 0x205F || c == 0x3000;
}
inline bool is_nl(uint32_t c) { return c == '\n' || c == '\r'; }
// lower case for the (?i:...) contractions: ASCII, plus U+017F (long s), which case-folds to 's'
inline uint32_t fold(uint32_t c) { return c == 0x17F ? 's' : ((c >= 'A' && c <= 'Z') ? c | 0x20 : c); }

struct Cp {
  uint32_t c;    // code point (0xFFFD class "other" for a byte that is not valid UTF-8)
  uint32_t off;  // byte offset in the text
};
// UTF-8 -> code points with byte offsets; an invalid byte becomes its own symbol so that no input byte is lost
void decode_utf8(const char* t, size_t n, std::vector<Cp>* out) {
  size_t i = 0;
  while (i < n) {
    const unsigned char b = (unsigned char)t[i];
    uint32_t c = 0xFFFD;
    size_t len = 1;
    auto cont = [&](size_t k) { return i + k < n && ((unsigned char)t[i + k] & 0xC0) == 0x80; };
    if (b < 0x80) c = b;
    else if ((b >> 5) == 6 && cont(1)) { c = ((b & 0x1Fu) << 6) | (t[i + 1] & 0x3F); len = 2; if (c < 0x80) { c = 0xFFFD; len = 1; } }
    else if ((b >> 4) == 14 && cont(1) && cont(2)) {
      c = ((b & 0x0Fu) << 12) | ((t[i + 1] & 0x3Fu) << 6) | (t[i + 2] & 0x3F); len = 3;
      if (c < 0x800 || (c >= 0xD800 && c <= 0xDFFF)) { c = 0xFFFD; len = 1; }
    } else if ((b >> 3) == 30 && cont(1) && cont(2) && cont(3)) {
      c = ((b & 0x07u) << 18) | ((t[i + 1] & 0x3Fu) << 12) | ((t[i + 2] & 0x3Fu) << 6) | (t[i + 3] & 0x3F); len = 4;
      if (c < 0x10000 || c > 0x10FFFF) { c = 0xFFFD; len = 1; }
    }
    out->push_back({c, (uint32_t)i});
    i += len;
  }
}

std::string merge_key(const std::string& a, const std::string& b) {
  std::string k;
  k.reserve(a.size() + b.size() + 5);
  const uint32_t n = (uint32_t)a.size();
  k.append((const char*)&n, 4);
  k += a;
  k += b;
  return k;
}

}  // namespace

// (?i:'s|'t|'re|'ve|'m|'ll|'d) | [^\r\n\p{L}\p{N}]?\p{L}+ | \p{N}{1,3} | ?[^\s\p{L}\p{N}]+[\r\n]* | \s*[\r\n]+ | \s+(?!\S) | \s+
// evaluated on code points, alternatives in order, each with the regex's greedy / back-tracking outcome
void pretokenize(const char* t, size_t nbytes, std::vector<std::string>* words) {
  std::vector<Cp> cp;
  decode_utf8(t, nbytes, &cp);
  const size_t n = cp.size();
  auto C = [&](size_t k) { return cp[k].c; };
  auto emit = [&](size_t i, size_t j) {  // code points [i, j)
    const size_t b0 = cp[i].off, b1 = j < n ? cp[j].off : nbytes;
    words->emplace_back(t + b0, b1 - b0);
  };
  size_t i = 0;
  while (i < n) {
    size_t j = i;
    // contractions
    if (C(i) == '\'' && i + 1 < n) {
      const uint32_t a = fold(C(i + 1)), b = i + 2 < n ? fold(C(i + 2)) : 0;
      size_t len = 0;
      if (a == 's' || a == 't' || a == 'm' || a == 'd') len = 2;
      if ((a == 'r' && b == 'e') || (a == 'v' && b == 'e') || (a == 'l' && b == 'l')) len = 3;
      if (len) { emit(i, i + len); i += len; continue; }
    }
    // optional single char that is no letter / number / newline, then letters
    {
      size_t k = i;
      if (!is_letter(C(k)) && !is_digit(C(k)) && !is_nl(C(k)) && k + 1 < n && is_letter(C(k + 1))) ++k;
      if (is_letter(C(k))) {
        j = k;
        while (j < n && is_letter(C(j))) ++j;
        emit(i, j); i = j; continue;
      }
    }
    if (is_digit(C(i))) {
      while (j < n && j - i < 3 && is_digit(C(j))) ++j;
      emit(i, j); i = j; continue;
    }
    {  // optional space, run of "other" characters, trailing newlines
      size_t k = i;
      if (C(k) == ' ' && k + 1 < n && !is_space(C(k + 1)) && !is_letter(C(k + 1)) && !is_digit(C(k + 1))) ++k;
      if (!is_space(C(k)) && !is_letter(C(k)) && !is_digit(C(k))) {
        j = k;
        while (j < n && !is_space(C(j)) && !is_letter(C(j)) && !is_digit(C(j))) ++j;
        while (j < n && is_nl(C(j))) ++j;
        emit(i, j); i = j; continue;
      }
    }
    // whitespace
    while (j < n && is_space(C(j))) ++j;
    size_t last_nl = 0;
    bool has_nl = false;
    for (size_t k = i; k < j; ++k)
      if (is_nl(C(k))) { has_nl = true; last_nl = k; }
    if (has_nl) { emit(i, last_nl + 1); i = last_nl + 1; continue; }  // \s*[\r\n]+
    if (j < n && j - i > 1) --j;  // \s+(?!\S): leave the last space to the following word
    emit(i, j);
    i = j;
  }
}

std::string Tokenizer::load(const GgufFile& f) {
  const GgufValue* toks = f.kv("tokenizer.ggml.tokens");
  if (!toks || toks->strs.empty()) return "no tokenizer.ggml.tokens in the GGUF file";
  const std::string model = f.get_s("tokenizer.ggml.model", "gpt2");
  if (model != "gpt2") return "tokenizer model '" + model + "' is not a byte-level BPE (gpt2) vocabulary";
  const GgufValue* types = f.kv("tokenizer.ggml.token_type");
  const size_t V = toks->strs.size();
  id_to_bytes_.assign(V, "");
  is_special_.assign(V, 0);
  for (size_t i = 0; i < V; ++i) {
    const int tt = types && i < types->ints.size() ? (int)types->ints[i] : 1;
    std::string raw;
    if (tt == 1 && unicode_token_to_bytes(toks->strs[i], &raw)) {  // normal token
      id_to_bytes_[i] = raw;
      bytes_to_id_.emplace(raw, (int32_t)i);
      max_token_bytes_ = std::max(max_token_bytes_, raw.size());
    } else {
      is_special_[i] = 1;  // control / user-defined / unused: never produced by encode, rendered as nothing
      special_to_id_.emplace(toks->strs[i], (int32_t)i);
    }
  }
  if (const GgufValue* merges = f.kv("tokenizer.ggml.merges")) {
    for (size_t r = 0; r < merges->strs.size(); ++r) {
      const std::string& m = merges->strs[r];
      const size_t sp = m.find(' ', 1);
      if (sp == std::string::npos) continue;
      std::string a, b;
      if (unicode_token_to_bytes(m.substr(0, sp), &a) && unicode_token_to_bytes(m.substr(sp + 1), &b))
        merge_rank_.emplace(merge_key(a, b), (int32_t)r);
    }
  }
  bos_id = (int32_t)f.get_u("tokenizer.ggml.bos_token_id", 128000);
  eos_id = (int32_t)f.get_u("tokenizer.ggml.eos_token_id", 128009);
  chat_template = f.get_s("tokenizer.chat_template", "");
  stop_ids.clear();
  auto add_stop = [&](int64_t id) {
    if (id < 0 || (size_t)id >= V) return;
    for (int32_t s : stop_ids) if (s == (int32_t)id) return;
    stop_ids.push_back((int32_t)id);
  };
  add_stop(eos_id);
  for (const char* key : {"tokenizer.ggml.eot_token_id", "tokenizer.ggml.eom_token_id"})
    if (f.kv(key)) add_stop((int64_t)f.get_u(key, 0));
  for (const char* text : {"<|eot_id|>", "<|eom_id|>", "<|end_of_text|>", "<|im_end|>", "<|endoftext|>", "<|end|>", "</s>"})
    add_stop(special_id(text));
  return "";
}

int32_t Tokenizer::special_id(const std::string& text) const {
  auto it = special_to_id_.find(text);
  return it == special_to_id_.end() ? -1 : it->second;
}

// One pre-token -> ids.  As llama.cpp does for "llama-bpe" (ignore_merges) and tiktoken does by construction, a
// pre-token that is itself a vocabulary entry is emitted as that id without consulting the merge table.  Otherwise
// the classic BPE loop — merge the adjacent pair of lowest rank, leftmost on ties, until no pair has a rank — run
// over a linked list of symbols with a priority queue of candidate pairs: O(n log n) per word instead of a full
// rescan per merge (a \p{L}+ run can be arbitrarily long).
void Tokenizer::bpe_word(const std::string& w, std::vector<int32_t>* out) const {
  if (w.empty()) return;
  {
    auto whole = bytes_to_id_.find(w);
    if (whole != bytes_to_id_.end()) { out->push_back(whole->second); return; }
  }
  struct Sym { int prev, next; uint32_t off, len; };
  struct Pair { int32_t rank; int left, right; uint32_t llen, rlen; };
  struct Later {  // priority_queue keeps the "largest": lowest rank first, then the leftmost pair
    bool operator()(const Pair& a, const Pair& b) const { return a.rank > b.rank || (a.rank == b.rank && a.left > b.left); }
  };
  const int n = (int)w.size();
  std::vector<Sym> sym((size_t)n);
  for (int i = 0; i < n; ++i) sym[i] = {i - 1, i + 1 < n ? i + 1 : -1, (uint32_t)i, 1u};
  std::priority_queue<Pair, std::vector<Pair>, Later> pq;
  auto push = [&](int l, int r) {
    if (l < 0 || r < 0) return;
    auto it = merge_rank_.find(merge_key(w.substr(sym[l].off, sym[l].len), w.substr(sym[r].off, sym[r].len)));
    if (it != merge_rank_.end()) pq.push({it->second, l, r, sym[l].len, sym[r].len});
  };
  for (int i = 0; i + 1 < n; ++i) push(i, i + 1);
  while (!pq.empty()) {
    const Pair p = pq.top();
    pq.pop();
    Sym& L = sym[p.left];
    Sym& R = sym[p.right];
    if (L.len != p.llen || R.len != p.rlen || L.next != p.right) continue;  // one side has changed since: stale
    L.len += R.len;  // symbols are contiguous in w, so the merged symbol is [L.off, L.off + L.len)
    R.len = 0;
    L.next = R.next;
    if (R.next >= 0) sym[R.next].prev = p.left;
    push(L.prev, p.left);
    push(p.left, L.next);
  }
  for (int i = 0; i >= 0; i = sym[i].next) {
    const std::string s = w.substr(sym[i].off, sym[i].len);
    auto it = bytes_to_id_.find(s);
    if (it != bytes_to_id_.end()) { out->push_back(it->second); continue; }
    for (char c : s) {  // a merge produced a string that is not a token: fall back to its bytes
      auto b2 = bytes_to_id_.find(std::string(1, c));
      if (b2 != bytes_to_id_.end()) out->push_back(b2->second);
    }
  }
}

void Tokenizer::encode(const char* text, size_t n, bool add_bos, int32_t vocab_limit, std::vector<int32_t>* out) const {
  if (add_bos && bos_id >= 0 && bos_id < vocab_limit) out->push_back(bos_id);
  if (!has_vocab()) {
    for (size_t i = 0; i < n; ++i) out->push_back((int32_t)((unsigned char)text[i] % vocab_limit));
    return;
  }
  std::vector<std::string> words;
  pretokenize(text, n, &words);
  for (const std::string& w : words) bpe_word(w, out);
}

void Tokenizer::decode(const int32_t* ids, size_t n, std::string* out) const {
  char tmp[24];
  for (size_t i = 0; i < n; ++i) {
    const int32_t id = ids[i];
    if (has_vocab()) {
      if (id >= 0 && (size_t)id < id_to_bytes_.size() && !is_special_[id]) *out += id_to_bytes_[id];
    } else if (id >= 0 && id < 256) {
      out->push_back((char)id);
    } else if (id >= 256 && id < 128000) {  // synthetic vocabulary: visible placeholder, valid UTF-8
      snprintf(tmp, sizeof tmp, " t%d", id);
      *out += tmp;
    }
  }
}

}  // namespace b200
