// tokenizer.cpp — text <-> ids for the HTTP front end.
//
// The reference sends text (web/streamlit_app.py:93) and Ollama tokenises it with the model's own
// byte-level BPE vocabulary, which ships inside the GGUF file.  No tokenizer asset exists offline, so the
// synthetic-weight engine uses the byte-level base alphabet of that scheme directly: ids 0..255 are the
// raw UTF-8 bytes (lossless round trip for any string), Llama-3 special ids keep their published values
// (128000 <|begin_of_text|>, 128009 <|eot_id|>).  Merges from a real GGUF plug in above this alphabet.
#include <cstdio>
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

// Renders ids as text.  Byte tokens give their byte; special ids give nothing; any other id of the synthetic
// vocabulary (no merge table is loaded) is shown as " t<id>" so that a random-init model still answers with
// visible, valid UTF-8 text.
int tokenizer_decode(const b200_config& cfg, const int32_t* ids, size_t n, char* buf, size_t cap, size_t* n_out) {
  (void)cfg;
  size_t o = 0;
  char tmp[24];
  for (int pass = 0; pass < 2; ++pass) {
    o = 0;
    for (size_t i = 0; i < n; ++i) {
      const int32_t id = ids[i];
      if (id >= 0 && id < 256) {
        if (pass) buf[o] = (char)id;
        ++o;
      } else if (id >= 256 && id < 128000) {
        const int m = snprintf(tmp, sizeof tmp, " t%d", id);
        if (pass) memcpy(buf + o, tmp, (size_t)m);
        o += (size_t)m;
      }
    }
    if (!pass) {
      *n_out = o;
      if (o > cap || (o && !buf)) {
        set_error("detokenize: need room for %zu bytes", o);
        return B200_ERANGE;
      }
    }
  }
  return 0;
}
}  // namespace b200
