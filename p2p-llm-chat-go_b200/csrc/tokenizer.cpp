// tokenizer.cpp — text <-> ids for the HTTP front end.
//
// The reference sends text (web/streamlit_app.py:93) and Ollama tokenises it with the model's own
// byte-level BPE vocabulary, which ships inside the GGUF file.  No tokenizer asset exists offline, so the
// synthetic-weight engine uses the byte-level base alphabet of that scheme directly: ids 0..255 are the
// raw UTF-8 bytes (lossless round trip for any string), Llama-3 special ids keep their published values
// (128000 <|begin_of_text|>, 128009 <|eot_id|>).  Merges from a real GGUF plug in above this alphabet.
#include <cstring>

#include "../../include/b200llm.h"

namespace b200 {
void set_error(const char* fmt, ...);

int tokenizer_encode(const b200_config& cfg, const char* utf8, size_t n, int add_bos, int32_t* ids, size_t cap, size_t* n_out) {
  const bool bos = add_bos && cfg.vocab > 128000;
  const size_t need = n + (bos ? 1 : 0);
  *n_out = need;
  if (need > cap || (need && !ids)) {
    set_error("tokenize: need room for %zu ids", need);
    return B200_ERANGE;
  }
  size_t o = 0;
  if (bos) ids[o++] = 128000;
  for (size_t i = 0; i < n; ++i) ids[o++] = (int32_t)(unsigned char)utf8[i] % cfg.vocab;
  return 0;
}

int tokenizer_decode(const b200_config& cfg, const int32_t* ids, size_t n, char* buf, size_t cap, size_t* n_out) {
  (void)cfg;
  size_t o = 0;
  for (size_t i = 0; i < n; ++i)
    if (ids[i] >= 0 && ids[i] < 256) ++o;
  *n_out = o;
  if (o > cap || (o && !buf)) {
    set_error("detokenize: need room for %zu bytes", o);
    return B200_ERANGE;
  }
  o = 0;
  for (size_t i = 0; i < n; ++i)
    if (ids[i] >= 0 && ids[i] < 256) buf[o++] = (char)ids[i];  // ids >= 256 (specials / unmapped) render as nothing
  return 0;
}
}  // namespace b200
