// tokenizer.h — text <-> token ids.
// With a GGUF vocabulary (tokenizer.ggml.tokens / merges / token_type): GPT-2-style byte-level BPE as used
// by Llama-3 ("llama-bpe").  Without one (synthetic random-init engines): the byte-level base alphabet only,
// ids 0..255 = raw bytes, Llama-3 special ids kept at their published values.
#pragma once
#include <cstdint>
#include <string>
#include <unordered_map>
#include <vector>

namespace b200 {

class GgufFile;

class Tokenizer {
 public:
  int32_t bos_id = 128000, eos_id = 128009;
  bool has_vocab() const { return !id_to_bytes_.empty(); }
  size_t vocab_size() const { return id_to_bytes_.size(); }
  std::string load(const GgufFile& f);  // "" on success
  void encode(const char* text, size_t n, bool add_bos, int32_t vocab_limit, std::vector<int32_t>* out) const;
  void decode(const int32_t* ids, size_t n, std::string* out) const;  // raw bytes (may split UTF-8 sequences)

 private:
  std::vector<std::string> id_to_bytes_;
  std::vector<uint8_t> is_special_;
  std::unordered_map<std::string, int32_t> bytes_to_id_;
  std::unordered_map<std::string, int32_t> merge_rank_;  // key: left + '\0' + right (lengths disambiguated below)
  void bpe_word(const std::string& w, std::vector<int32_t>* out) const;
};

// llama-bpe pre-tokenisation (approximation of the Llama-3 regex: ASCII classes exact, every non-ASCII
// byte counts as a letter).  Exposed for tests.
void pretokenize(const char* text, size_t n, std::vector<std::string>* words);

}  // namespace b200
