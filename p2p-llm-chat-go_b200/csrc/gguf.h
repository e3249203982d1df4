// gguf.h — reader for GGUF v2/v3 model files (the format `ollama pull llama3.1` leaves on disk): metadata
// key/values, tensor directory, memory-mapped tensor data.  Host-only; nothing here touches the GPU.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

namespace b200 {

struct GgufValue {
  uint32_t type = 0;  // GGUF value type; 9 = array (elem_type below)
  uint32_t elem_type = 0;
  uint64_t u = 0;     // any integer / bool
  double f = 0;       // any float
  std::string s;      // string
  std::vector<std::string> strs;  // array of strings
  std::vector<int64_t> ints;      // array of integers
  std::vector<float> floats;      // array of floats
};

struct GgufTensor {
  std::string name;
  uint32_t type = 0;          // ggml type id
  std::vector<int64_t> ne;    // ne[0] = innermost (columns)
  uint64_t offset = 0;        // from the start of the data section
  uint64_t nbytes = 0;
};

class GgufFile {
 public:
  ~GgufFile();
  // returns "" on success, else a message
  std::string open(const std::string& path);
  const GgufValue* kv(const std::string& key) const;
  uint64_t get_u(const std::string& key, uint64_t dflt) const;
  double get_f(const std::string& key, double dflt) const;
  std::string get_s(const std::string& key, const std::string& dflt) const;
  const GgufTensor* tensor(const std::string& name) const;
  const uint8_t* data(const GgufTensor& t) const { return base_ + data_off_ + t.offset; }
  const std::vector<GgufTensor>& tensors() const { return tensors_; }
  uint32_t version() const { return version_; }

 private:
  int fd_ = -1;
  const uint8_t* base_ = nullptr;
  uint64_t size_ = 0, data_off_ = 0;
  uint32_t version_ = 0;
  std::map<std::string, GgufValue> kvs_;
  std::vector<GgufTensor> tensors_;
  std::map<std::string, size_t> index_;
};

uint64_t ggml_tensor_bytes(uint32_t type, int64_t n_elements);  // 0 for types this library does not handle

}  // namespace b200
