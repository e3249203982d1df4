#include <cstdarg>
#include <cstdio>
#include <random>
#include <string>
#include <vector>
#include "../p2p-llm-chat-go_b200/csrc/tokenizer.h"
namespace b200 { void set_error(const char*, ...) {} }
int main() {
  std::mt19937 rng(9);
  size_t total = 0;
  for (int it = 0; it < 200000; ++it) {
    std::string s(rng() % 40, 0);
    for (auto& c : s) { const int r = rng() % 10; c = r < 4 ? (char)(rng() % 128) : r < 7 ? (char)(0x80 + rng() % 128) : " '\n\r\t1a"[rng() % 7]; }
    std::vector<std::string> w;
    b200::pretokenize(s.data(), s.size(), &w);
    std::string j;
    for (auto& p : w) { if (p.empty()) { printf("empty piece\n"); return 1; } j += p; }
    if (j != s) { printf("bytes lost\n"); return 1; }
    total += w.size();
    // the byte-fallback tokenizer (no vocabulary) round trip
    b200::Tokenizer tk;
    std::vector<int32_t> ids;
    tk.encode(s.data(), s.size(), false, 128256, &ids);
    std::string back;
    tk.decode(ids.data(), ids.size(), &back);
    if (back != s) { printf("round trip differs\n"); return 1; }
  }
  printf("ok, %zu pieces\n", total);
  return 0;
}
