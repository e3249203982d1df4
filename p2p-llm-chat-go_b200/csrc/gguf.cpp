// gguf.cpp — GGUF v2/v3 reader (see gguf.h).  Format per the public GGUF specification:
// "GGUF" | version u32 | tensor_count u64 | kv_count u64 | kv pairs | tensor infos | padding to
// general.alignment (default 32) | tensor data.  Strings are u64 length + bytes, no terminator.
#include "gguf.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstring>

namespace b200 {

namespace {
struct Cursor {
  const uint8_t* p;
  const uint8_t* end;
  bool ok = true;
  template <class T>
  T get() {
    T v{};
    if ((size_t)(end - p) < sizeof(T)) { ok = false; return v; }
    memcpy(&v, p, sizeof(T));
    p += sizeof(T);
    return v;
  }
  std::string str() {
    const uint64_t n = get<uint64_t>();
    if (!ok || (uint64_t)(end - p) < n) { ok = false; return ""; }
    std::string s((const char*)p, (size_t)n);
    p += n;
    return s;
  }
};

bool read_scalar(Cursor& c, uint32_t type, GgufValue* v) {
  switch (type) {
    case 0: v->u = c.get<uint8_t>(); break;
    case 1: v->u = (uint64_t)(int64_t)c.get<int8_t>(); break;
    case 2: v->u = c.get<uint16_t>(); break;
    case 3: v->u = (uint64_t)(int64_t)c.get<int16_t>(); break;
    case 4: v->u = c.get<uint32_t>(); break;
    case 5: v->u = (uint64_t)(int64_t)c.get<int32_t>(); break;
    case 6: v->f = c.get<float>(); v->u = (uint64_t)v->f; break;
    case 7: v->u = c.get<uint8_t>() != 0; break;
    case 8: v->s = c.str(); break;
    case 10: v->u = c.get<uint64_t>(); break;
    case 11: v->u = (uint64_t)c.get<int64_t>(); break;
    case 12: v->f = c.get<double>(); v->u = (uint64_t)v->f; break;
    default: return false;
  }
  if (type != 6 && type != 12 && type != 8) v->f = (double)(int64_t)v->u;
  return c.ok;
}
}  // namespace

uint64_t ggml_tensor_bytes(uint32_t type, int64_t n) {
  switch (type) {
    case 0: return (uint64_t)n * 4;                 // F32
    case 1: return (uint64_t)n * 2;                 // F16
    case 30: return (uint64_t)n * 2;                // BF16
    case 12: return (uint64_t)(n / 256) * 144;      // Q4_K
    case 14: return (uint64_t)(n / 256) * 210;      // Q6_K
    case 8: return (uint64_t)(n / 32) * 34;         // Q8_0 (size only)
  }
  return 0;
}

GgufFile::~GgufFile() {
  if (base_) munmap((void*)base_, size_);
  if (fd_ >= 0) close(fd_);
}

std::string GgufFile::open(const std::string& path) {
  fd_ = ::open(path.c_str(), O_RDONLY);
  if (fd_ < 0) return "cannot open " + path;
  struct stat st;
  if (fstat(fd_, &st) != 0 || st.st_size < 24) return "cannot stat / file too small: " + path;
  size_ = (uint64_t)st.st_size;
  void* m = mmap(nullptr, size_, PROT_READ, MAP_PRIVATE, fd_, 0);
  if (m == MAP_FAILED) return "mmap failed: " + path;
  base_ = (const uint8_t*)m;
  Cursor c{base_, base_ + size_};
  if (memcmp(c.p, "GGUF", 4) != 0) return "not a GGUF file (bad magic)";
  c.p += 4;
  version_ = c.get<uint32_t>();
  if (version_ < 2 || version_ > 3) return "unsupported GGUF version " + std::to_string(version_);
  const uint64_t n_tensors = c.get<uint64_t>(), n_kv = c.get<uint64_t>();
  if (!c.ok || n_tensors > (1u << 20) || n_kv > (1u << 20)) return "corrupt GGUF header";
  for (uint64_t i = 0; i < n_kv; ++i) {
    const std::string key = c.str();
    GgufValue v;
    v.type = c.get<uint32_t>();
    if (!c.ok) return "corrupt GGUF metadata";
    if (v.type == 9) {
      v.elem_type = c.get<uint32_t>();
      const uint64_t n = c.get<uint64_t>();
      if (!c.ok || n > (1ull << 28)) return "corrupt GGUF array '" + key + "'";
      for (uint64_t k = 0; k < n; ++k) {
        GgufValue e;
        if (!read_scalar(c, v.elem_type, &e)) return "unsupported GGUF array element type in '" + key + "'";
        if (v.elem_type == 8) v.strs.push_back(std::move(e.s));
        else if (v.elem_type == 6 || v.elem_type == 12) v.floats.push_back((float)e.f);
        else v.ints.push_back((int64_t)e.u);
      }
    } else if (!read_scalar(c, v.type, &v)) {
      return "unsupported GGUF value type for '" + key + "'";
    }
    kvs_[key] = std::move(v);
  }
  tensors_.resize((size_t)n_tensors);
  for (auto& t : tensors_) {
    t.name = c.str();
    const uint32_t nd = c.get<uint32_t>();
    if (!c.ok || nd > 4) return "corrupt GGUF tensor info";
    int64_t n = 1;
    for (uint32_t d = 0; d < nd; ++d) {
      const uint64_t dim = c.get<uint64_t>();
      // element counts beyond 2^40 cannot be real (and would overflow the product / the byte size)
      if (!c.ok || dim > (1ull << 40) || (dim != 0 && (uint64_t)n > (1ull << 40) / dim)) return "corrupt GGUF tensor info (dimensions)";
      t.ne.push_back((int64_t)dim);
      n *= (int64_t)dim;
    }
    t.type = c.get<uint32_t>();
    t.offset = c.get<uint64_t>();
    t.nbytes = ggml_tensor_bytes(t.type, n);
    if (!c.ok) return "corrupt GGUF tensor info";
  }
  const uint64_t align = get_u("general.alignment", 32);
  if (align == 0 || (align & (align - 1)) != 0 || align > (1u << 20)) return "corrupt GGUF header (general.alignment must be a power of two)";
  const uint64_t pos = (uint64_t)(c.p - base_);
  data_off_ = (pos + align - 1) / align * align;
  if (data_off_ > size_) data_off_ = size_;
  for (size_t i = 0; i < tensors_.size(); ++i) {
    const GgufTensor& t = tensors_[i];
    // written so that a huge offset cannot wrap the sum around
    const uint64_t room = size_ - data_off_;
    if (t.nbytes && (t.offset > room || t.nbytes > room - t.offset)) return "tensor '" + t.name + "' runs past the end of the file";
    index_[t.name] = i;
  }
  return "";
}

const GgufValue* GgufFile::kv(const std::string& key) const {
  auto it = kvs_.find(key);
  return it == kvs_.end() ? nullptr : &it->second;
}
uint64_t GgufFile::get_u(const std::string& key, uint64_t d) const {
  const GgufValue* v = kv(key);
  return v && v->type != 8 && v->type != 9 ? v->u : d;
}
double GgufFile::get_f(const std::string& key, double d) const {
  const GgufValue* v = kv(key);
  return v && v->type != 8 && v->type != 9 ? v->f : d;
}
std::string GgufFile::get_s(const std::string& key, const std::string& d) const {
  const GgufValue* v = kv(key);
  return v && v->type == 8 ? v->s : d;
}
const GgufTensor* GgufFile::tensor(const std::string& name) const {
  auto it = index_.find(name);
  return it == index_.end() ? nullptr : &tensors_[it->second];
}

}  // namespace b200
