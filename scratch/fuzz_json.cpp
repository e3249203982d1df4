#include <cstdio>
#include <cstdlib>
#include <random>
#include <string>
#include "../p2p-llm-chat-go_b200/csrc/json.h"
int main() {
  std::mt19937 rng(5);
  const std::string seeds[] = {
    R"({"model":"llama3.1","prompt":"hi \"there\" é 😀","stream":false,"options":{"temperature":0.8,"top_k":40,"seed":1e3,"stop":["a","b"]}})",
    R"({"model":"m","messages":[{"role":"user","content":"x"},{"role":"assistant","content":"y"}],"stream":true})",
    R"([1,2,{"a":[[[[]]]]},null,true,false,-1.5e-3,"\\\/\b\f\n\r\t"])"};
  const char alphabet[] = "{}[]\",:\\u0123456789eE+-.tfn alsrue\xc3\xa9\x00\xff";
  long ok = 0, bad = 0;
  for (int it = 0; it < 300000; ++it) {
    std::string s = seeds[rng() % 3];
    const int n_mut = rng() % 6;
    for (int m = 0; m < n_mut && !s.empty(); ++m) {
      const size_t pos = rng() % s.size();
      switch (rng() % 4) {
        case 0: s[pos] = alphabet[rng() % (sizeof alphabet - 1)]; break;
        case 1: s.erase(pos, 1 + rng() % 4); break;
        case 2: s.insert(pos, 1, alphabet[rng() % (sizeof alphabet - 1)]); break;
        default: s.resize(pos); break;
      }
    }
    if (it % 1000 == 0) s = std::string(100000, '[');  // deep nesting
    mini_json::Value v;
    if (mini_json::Parser(s).parse(&v)) { ++ok; (void)v.get_str("model"); (void)v.get_bool("stream", true); (void)mini_json::quote(v.get_str("prompt")); }
    else ++bad;
  }
  printf("parsed %ld, rejected %ld\n", ok, bad);
  return 0;
}
