// Native HF checkpoint loading behind the C ABI: config.json, safetensors files
// (single or sharded through model.safetensors.index.json), bf16 / f16 / f32 tensors, tied
// embeddings.  Replaces Qwen3Config::from_hf_dir (src/models/qwen3.rs:77-101) and
// Qwen3ForCausalLM::from_hf_dir (:515-536), which goes through candle's VarBuilder over ONE
// mmapped model.safetensors converted to f32 and has no tied-embedding fallback (:525-528).
// Tensors go mmap -> (dtype conversion on the host if needed) -> nvc_runner_load_weight, which
// slices this rank's shard; nothing is staged as f32.
//
// safetensors layout: u64 little-endian header length N, N bytes of JSON
// {name: {"dtype": "BF16", "shape": [..], "data_offsets": [begin, end]}, "__metadata__": {..}},
// then the byte buffer the offsets index.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <set>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/nvc_b200_engine.h"

namespace {

// ---------------------------------------------------------------------------------- tiny JSON
struct JValue {
  enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
  bool b = false;
  double num = 0;
  std::string str;
  std::vector<JValue> arr;
  std::vector<std::pair<std::string, JValue>> obj;   // insertion order kept

  const JValue* get(const std::string& key) const {
    if (kind != Obj) return nullptr;
    for (auto& kv : obj)
      if (kv.first == key) return &kv.second;
    return nullptr;
  }
};

struct JParser {
  const char* p;
  const char* end;
  std::string err;

  bool fail(const std::string& m) {
    if (err.empty()) err = m;
    return false;
  }
  void ws() {
    while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p;
  }
  bool lit(const char* s) {
    const size_t n = strlen(s);
    if ((size_t)(end - p) < n || memcmp(p, s, n) != 0) return fail(std::string("expected ") + s);
    p += n;
    return true;
  }
  static void utf8(std::string& out, uint32_t cp) {
    if (cp < 0x80) out += (char)cp;
    else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
    else if (cp < 0x10000) { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
    else { out += (char)(0xF0 | (cp >> 18)); out += (char)(0x80 | ((cp >> 12) & 0x3F)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
  }
  bool hex4(uint32_t& v) {
    if (end - p < 4) return fail("truncated \\u escape");
    v = 0;
    for (int i = 0; i < 4; ++i) {
      const char c = *p++;
      v <<= 4;
      if (c >= '0' && c <= '9') v |= c - '0';
      else if (c >= 'a' && c <= 'f') v |= c - 'a' + 10;
      else if (c >= 'A' && c <= 'F') v |= c - 'A' + 10;
      else return fail("bad \\u escape");
    }
    return true;
  }
  bool string(std::string& out) {
    if (p >= end || *p != '"') return fail("expected string");
    ++p;
    while (p < end && *p != '"') {
      if (*p == '\\') {
        if (++p >= end) return fail("truncated escape");
        const char c = *p++;
        switch (c) {
          case '"': out += '"'; break;
          case '\\': out += '\\'; break;
          case '/': out += '/'; break;
          case 'b': out += '\b'; break;
          case 'f': out += '\f'; break;
          case 'n': out += '\n'; break;
          case 'r': out += '\r'; break;
          case 't': out += '\t'; break;
          case 'u': {
            uint32_t cp;
            if (!hex4(cp)) return false;
            if (cp >= 0xDC00 && cp < 0xE000) return fail("lone trailing surrogate in \\u escape");
            if (cp >= 0xD800 && cp < 0xDC00) {
              if (end - p < 6 || p[0] != '\\' || p[1] != 'u') return fail("lone leading surrogate in \\u escape");
              p += 2;
              uint32_t lo;
              if (!hex4(lo)) return false;
              if (lo < 0xDC00 || lo >= 0xE000) return fail("lone leading surrogate in \\u escape");
              cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
            }
            utf8(out, cp);
            break;
          }
          default: return fail("bad escape");
        }
      } else {
        // raw bytes: no control characters, well-formed UTF-8 only (what serde_json / RFC 8259 accept)
        const unsigned char c = (unsigned char)*p;
        if (c < 0x20) return fail("control character in string");
        int extra = 0;
        uint32_t cp = c;
        if (c >= 0x80) {
          if (c >= 0xC2 && c <= 0xDF) { extra = 1; cp = c & 0x1F; }
          else if (c >= 0xE0 && c <= 0xEF) { extra = 2; cp = c & 0x0F; }
          else if (c >= 0xF0 && c <= 0xF4) { extra = 3; cp = c & 0x07; }
          else return fail("invalid UTF-8 in string");
          if (end - p <= extra) return fail("invalid UTF-8 in string");
          for (int i = 1; i <= extra; ++i) {
            const unsigned char cc = (unsigned char)p[i];
            if ((cc & 0xC0) != 0x80) return fail("invalid UTF-8 in string");
            cp = (cp << 6) | (cc & 0x3F);
          }
          if ((extra == 2 && (cp < 0x800 || (cp >= 0xD800 && cp < 0xE000))) || (extra == 3 && (cp < 0x10000 || cp > 0x10FFFF)))
            return fail("invalid UTF-8 in string");
        }
        out.append(p, (size_t)extra + 1);
        p += extra + 1;
      }
    }
    if (p >= end) return fail("unterminated string");
    ++p;
    return true;
  }
  bool value(JValue& v, int depth = 0) {
    if (depth > 64) return fail("nesting too deep");
    ws();
    if (p >= end) return fail("unexpected end of input");
    const char c = *p;
    if (c == '{') {
      v.kind = JValue::Obj;
      ++p;
      ws();
      if (p < end && *p == '}') { ++p; return true; }
      for (;;) {
        ws();
        std::string key;
        if (!string(key)) return false;
        ws();
        if (p >= end || *p != ':') return fail("expected ':'");
        ++p;
        JValue child;
        if (!value(child, depth + 1)) return false;
        v.obj.emplace_back(std::move(key), std::move(child));
        ws();
        if (p < end && *p == ',') { ++p; continue; }
        if (p < end && *p == '}') { ++p; return true; }
        return fail("expected ',' or '}'");
      }
    }
    if (c == '[') {
      v.kind = JValue::Arr;
      ++p;
      ws();
      if (p < end && *p == ']') { ++p; return true; }
      for (;;) {
        JValue child;
        if (!value(child, depth + 1)) return false;
        v.arr.push_back(std::move(child));
        ws();
        if (p < end && *p == ',') { ++p; continue; }
        if (p < end && *p == ']') { ++p; return true; }
        return fail("expected ',' or ']'");
      }
    }
    if (c == '"') { v.kind = JValue::Str; return string(v.str); }
    if (c == 't') { v.kind = JValue::Bool; v.b = true; return lit("true"); }
    if (c == 'f') { v.kind = JValue::Bool; v.b = false; return lit("false"); }
    if (c == 'n') { v.kind = JValue::Null; return lit("null"); }
    // number
    // number: -? (0 | [1-9][0-9]*) (. [0-9]+)? ([eE] [+-]? [0-9]+)?   (RFC 8259)
    const char* s = p;
    auto digits = [&]() { const char* d = p; while (p < end && *p >= '0' && *p <= '9') ++p; return p > d; };
    if (p < end && *p == '-') ++p;
    if (p < end && *p == '0') ++p;
    else if (!digits()) return fail(p == s ? "unexpected character" : "invalid number");
    if (p < end && *p == '.') { ++p; if (!digits()) return fail("invalid number"); }
    if (p < end && (*p == 'e' || *p == 'E')) {
      ++p;
      if (p < end && (*p == '+' || *p == '-')) ++p;
      if (!digits()) return fail("invalid number");
    }
    v.kind = JValue::Num;
    v.num = strtod(std::string(s, p).c_str(), nullptr);
    if (!std::isfinite(v.num)) return fail("number out of range");
    return true;
  }
};

bool parse_json(const char* data, size_t n, JValue& out, std::string& err) {
  JParser ps{data, data + n, ""};
  if (!ps.value(out)) { err = ps.err; return false; }
  ps.ws();
  if (ps.p != ps.end) { err = "trailing characters after JSON value"; return false; }
  return true;
}

bool read_file(const std::string& path, std::string& out) {
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  std::ostringstream ss;
  ss << f.rdbuf();
  out = ss.str();
  return true;
}

}  // namespace

// ---------------------------------------------------------------------------------- safetensors
struct nvc_safetensors {
  struct Entry {
    std::string name;
    int dtype;                 // NVC_ST_*
    std::string dtype_name;
    std::vector<int64_t> shape;
    uint64_t begin, end;       // byte offsets into the data section
  };
  std::string path;
  int fd = -1;
  const uint8_t* map = nullptr;
  size_t map_bytes = 0;
  const uint8_t* data = nullptr;   // start of the byte buffer
  size_t data_bytes = 0;
  std::vector<Entry> entries;
  std::map<std::string, int> by_name;

  ~nvc_safetensors() {
    if (map) munmap(const_cast<uint8_t*>(map), map_bytes);
    if (fd >= 0) close(fd);
  }
};

namespace {

thread_local std::string g_hf_error;
thread_local std::string g_hf_warning;   // non-fatal notes of the last nvc_runner_load_hf_dir (skipped tensors)

int hf_fail(int code, const std::string& msg) {
  g_hf_error = msg;
  return code;
}

int elem_bytes(int dtype) { return dtype == NVC_ST_F32 ? 4 : 2; }

// f32 -> bf16, round to nearest even (what torch's .to(bfloat16) and candle's bf16::from_f32 do); NaN stays NaN
inline uint16_t f32_to_bf16(uint32_t u) {
  if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x0040u);
  return (uint16_t)((u + 0x7fffu + ((u >> 16) & 1u)) >> 16);
}

inline uint32_t f16_to_f32_bits(uint16_t h) {
  const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
  uint32_t exp = (h >> 10) & 0x1fu, man = h & 0x3ffu;
  if (exp == 0) {
    if (man == 0) return sign;
    int e = -1;                                  // subnormal half: normalise
    do { ++e; man <<= 1; } while (!(man & 0x400u));
    return sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3ffu) << 13);
  }
  if (exp == 31) return sign | 0x7f800000u | (man << 13);
  return sign | ((exp + 112u) << 23) | (man << 13);
}

// converts `n` elements of `dtype` at src into bf16 at dst
void to_bf16(const uint8_t* src, int dtype, size_t n, uint16_t* dst) {
  if (dtype == NVC_ST_BF16) {
    memcpy(dst, src, n * 2);
  } else if (dtype == NVC_ST_F16) {
    for (size_t i = 0; i < n; ++i) {
      uint16_t h;
      memcpy(&h, src + 2 * i, 2);
      dst[i] = f32_to_bf16(f16_to_f32_bits(h));
    }
  } else {
    for (size_t i = 0; i < n; ++i) {
      uint32_t u;
      memcpy(&u, src + 4 * i, 4);
      dst[i] = f32_to_bf16(u);
    }
  }
}

int open_safetensors(const std::string& path, std::unique_ptr<nvc_safetensors>& out) {
  std::unique_ptr<nvc_safetensors> st(new nvc_safetensors());
  st->path = path;
  st->fd = open(path.c_str(), O_RDONLY);
  if (st->fd < 0) return hf_fail(NVC_EINVAL, "cannot open " + path);
  struct stat sb;
  if (fstat(st->fd, &sb) != 0 || sb.st_size < 8) return hf_fail(NVC_EINVAL, path + ": not a safetensors file (too small)");
  st->map_bytes = (size_t)sb.st_size;
  void* m = mmap(nullptr, st->map_bytes, PROT_READ, MAP_PRIVATE, st->fd, 0);
  if (m == MAP_FAILED) return hf_fail(NVC_ENOMEM, "mmap failed for " + path);
  st->map = static_cast<const uint8_t*>(m);
  uint64_t hlen = 0;
  for (int i = 0; i < 8; ++i) hlen |= (uint64_t)st->map[i] << (8 * i);
  if (hlen > st->map_bytes - 8 || hlen > (100ull << 20)) return hf_fail(NVC_EINVAL, path + ": bad safetensors header length");
  st->data = st->map + 8 + hlen;
  st->data_bytes = st->map_bytes - 8 - (size_t)hlen;
  JValue root;
  std::string err;
  if (!parse_json(reinterpret_cast<const char*>(st->map + 8), (size_t)hlen, root, err) || root.kind != JValue::Obj)
    return hf_fail(NVC_EINVAL, path + ": header is not a JSON object (" + err + ")");
  for (auto& kv : root.obj) {
    if (kv.first == "__metadata__") continue;
    const JValue *dt = kv.second.get("dtype"), *sh = kv.second.get("shape"), *off = kv.second.get("data_offsets");
    if (!dt || dt->kind != JValue::Str || !sh || sh->kind != JValue::Arr || !off || off->kind != JValue::Arr ||
        off->arr.size() != 2)
      return hf_fail(NVC_EINVAL, path + ": malformed entry for tensor " + kv.first);
    nvc_safetensors::Entry e;
    e.name = kv.first;
    e.dtype_name = dt->str;
    e.dtype = dt->str == "BF16" ? NVC_ST_BF16 : dt->str == "F16" ? NVC_ST_F16 : dt->str == "F32" ? NVC_ST_F32 : NVC_ST_OTHER;
    // sizes and offsets are unsigned integers; a double holds them exactly up to 2^53 (8 PiB), beyond any file
    auto as_u64 = [](const JValue& v, uint64_t& o) {
      if (v.kind != JValue::Num || !(v.num >= 0.0 && v.num <= 9007199254740992.0) || v.num != std::floor(v.num)) return false;
      o = (uint64_t)v.num;
      return true;
    };
    uint64_t count = 1;
    for (auto& d : sh->arr) {
      uint64_t dim;
      if (!as_u64(d, dim) || (dim != 0 && count > (1ull << 62) / dim)) return hf_fail(NVC_EINVAL, path + ": bad shape for " + kv.first);
      e.shape.push_back((int64_t)dim);
      count *= dim;
    }
    if (!as_u64(off->arr[0], e.begin) || !as_u64(off->arr[1], e.end))
      return hf_fail(NVC_EINVAL, path + ": data_offsets out of range for " + kv.first);
    if (e.begin > e.end || e.end > st->data_bytes) return hf_fail(NVC_EINVAL, path + ": data_offsets out of range for " + kv.first);
    if (e.dtype != NVC_ST_OTHER && e.end - e.begin != count * (uint64_t)elem_bytes(e.dtype))
      return hf_fail(NVC_EINVAL, path + ": byte size does not match shape for " + kv.first);
    if (!st->by_name.emplace(e.name, (int)st->entries.size()).second)
      return hf_fail(NVC_EINVAL, path + ": tensor " + kv.first + " is listed twice");
    st->entries.push_back(std::move(e));
  }
  {   // the format stores tensors back to back: two entries sharing bytes mean a damaged (or hostile) header
    std::vector<std::pair<uint64_t, uint64_t>> spans;
    for (auto& e : st->entries)
      if (e.end > e.begin) spans.emplace_back(e.begin, e.end);
    std::sort(spans.begin(), spans.end());
    for (size_t i = 1; i < spans.size(); ++i)
      if (spans[i].first < spans[i - 1].second) return hf_fail(NVC_EINVAL, path + ": overlapping data_offsets");
  }
  out = std::move(st);
  return NVC_OK;
}

bool file_exists(const std::string& p) {
  struct stat sb;
  return stat(p.c_str(), &sb) == 0 && S_ISREG(sb.st_mode);
}

}  // namespace

extern "C" {

const char* nvc_hf_last_error(void) { return g_hf_error.c_str(); }
const char* nvc_hf_last_warning(void) { return g_hf_warning.c_str(); }

int nvc_safetensors_open(const char* path, nvc_safetensors** out) {
  if (!path || !out) return hf_fail(NVC_EINVAL, "null pointer");
  *out = nullptr;
  std::unique_ptr<nvc_safetensors> st;
  int rc = open_safetensors(path, st);
  if (rc) return rc;
  *out = st.release();
  return NVC_OK;
}

int nvc_safetensors_close(nvc_safetensors* st) {
  if (!st) return NVC_EINVAL;
  delete st;
  return NVC_OK;
}

int nvc_safetensors_num_tensors(const nvc_safetensors* st) { return st ? (int)st->entries.size() : 0; }

int nvc_safetensors_tensor_info(const nvc_safetensors* st, int i, const char** name, int* dtype, int* ndim,
                                int64_t* shape4) {
  if (!st || i < 0 || i >= (int)st->entries.size()) return hf_fail(NVC_EINVAL, "tensor index out of range");
  const auto& e = st->entries[i];
  if (name) *name = e.name.c_str();
  if (dtype) *dtype = e.dtype;
  if (ndim) *ndim = (int)e.shape.size();
  if (shape4)
    for (int d = 0; d < 4; ++d) shape4[d] = d < (int)e.shape.size() ? e.shape[d] : 1;
  return NVC_OK;
}

int nvc_safetensors_read_bf16(const nvc_safetensors* st, const char* name, nvc_bf16* out, int64_t capacity_elems) {
  if (!st || !name || !out) return hf_fail(NVC_EINVAL, "null pointer");
  auto it = st->by_name.find(name);
  if (it == st->by_name.end()) return hf_fail(NVC_EINVAL, std::string("no tensor named ") + name + " in " + st->path);
  const auto& e = st->entries[it->second];
  if (e.dtype == NVC_ST_OTHER) return hf_fail(NVC_ENOTIMPL, e.name + ": unsupported dtype " + e.dtype_name);
  const size_t n = (size_t)(e.end - e.begin) / elem_bytes(e.dtype);
  if ((int64_t)n > capacity_elems) return hf_fail(NVC_ENOMEM, e.name + ": output buffer too small");
  to_bf16(st->data + e.begin, e.dtype, n, out);
  return NVC_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------- config.json + runner load
namespace {

bool jnum(const JValue& root, const char* key, double& out) {
  const JValue* v = root.get(key);
  if (!v || v->kind != JValue::Num) return false;
  out = v->num;
  return true;
}

// The tensors Qwen3ForCausalLM::from_hf_dir fetches (qwen3.rs:146-187,301-314,351-363,431-447,525-528).  The reference's
// VarBuilder only asks for these names, so anything else in a checkpoint (buffers of other exports) is skipped, with a
// note in nvc_hf_last_warning(); that every needed tensor arrived is checked by nvc_runner_finalize's count.
bool is_model_tensor(const std::string& name) {
  if (name == "model.embed_tokens.weight" || name == "model.norm.weight" || name == "lm_head.weight") return true;
  static const char* kPrefix = "model.layers.";
  if (name.compare(0, strlen(kPrefix), kPrefix) != 0) return false;
  size_t p = strlen(kPrefix), digits = 0;
  while (p < name.size() && name[p] >= '0' && name[p] <= '9') { ++p; ++digits; }
  if (digits == 0 || digits > 6 || p >= name.size() || name[p] != '.') return false;
  static const char* kSub[] = {"self_attn.q_proj.weight", "self_attn.k_proj.weight", "self_attn.v_proj.weight",
                               "self_attn.o_proj.weight", "self_attn.q_norm.weight", "self_attn.k_norm.weight",
                               "mlp.gate_proj.weight", "mlp.up_proj.weight", "mlp.down_proj.weight",
                               "input_layernorm.weight", "post_attention_layernorm.weight"};
  const std::string sub = name.substr(p + 1);
  for (const char* k : kSub)
    if (sub == k) return true;
  return false;
}

// names of a safetensors file -> runner, through one pinned-size-agnostic host staging buffer
int load_file_into_runner(nvc_runner* r, const nvc_safetensors& st, const std::set<std::string>* only,
                          std::set<std::string>& seen, std::vector<uint16_t>& stage) {
  for (const auto& e : st.entries) {
    if (only && !only->count(e.name)) continue;
    static const std::string kInvFreq = "rotary_emb.inv_freq";   // legacy buffer of older exports, not a weight
    if (e.name.size() >= kInvFreq.size() && e.name.compare(e.name.size() - kInvFreq.size(), kInvFreq.size(), kInvFreq) == 0) continue;
    if (!is_model_tensor(e.name)) {
      if (g_hf_warning.size() < 2000) g_hf_warning += "skipped unknown tensor " + e.name + " in " + st.path + "; ";
      continue;
    }
    if (e.dtype == NVC_ST_OTHER) return hf_fail(NVC_ENOTIMPL, e.name + ": unsupported dtype " + e.dtype_name + " in " + st.path);
    if (e.shape.empty() || e.shape.size() > 2) return hf_fail(NVC_ESHAPE, e.name + ": expected a vector or a matrix");
    const int64_t rows = e.shape[0], cols = e.shape.size() == 2 ? e.shape[1] : 1;
    const uint16_t* src;
    if (e.dtype == NVC_ST_BF16 && ((uintptr_t)(st.data + e.begin) % 2) == 0) {
      src = reinterpret_cast<const uint16_t*>(st.data + e.begin);   // straight out of the mapping
    } else {
      stage.resize((size_t)rows * cols);
      to_bf16(st.data + e.begin, e.dtype, (size_t)rows * cols, stage.data());
      src = stage.data();
    }
    int rc = nvc_runner_load_weight(r, e.name.c_str(), src, rows, cols);
    if (rc) return hf_fail(rc, "nvc_runner_load_weight failed for " + e.name + " from " + st.path);
    seen.insert(e.name);
  }
  return NVC_OK;
}

}  // namespace

extern "C" {

int nvc_qwen3_config_from_hf_dir(const char* model_dir, nvc_qwen3_config* out, int* tie_word_embeddings) {
  if (!model_dir || !out) return hf_fail(NVC_EINVAL, "null pointer");
  const std::string path = std::string(model_dir) + "/config.json";
  std::string text, err;
  if (!read_file(path, text)) return hf_fail(NVC_EINVAL, "Failed to read HF config " + path);
  JValue root;
  if (!parse_json(text.data(), text.size(), root, err) || root.kind != JValue::Obj)
    return hf_fail(NVC_EINVAL, "Failed to parse HF config " + path + ": " + err);
  // required fields of HfQwen3Config (qwen3.rs:36-53)
  struct { const char* key; int32_t* dst; } req[] = {
      {"vocab_size", &out->vocab_size}, {"hidden_size", &out->hidden_size},
      {"num_hidden_layers", &out->num_hidden_layers}, {"num_attention_heads", &out->num_attention_heads},
      {"num_key_value_heads", &out->num_key_value_heads}, {"intermediate_size", &out->intermediate_size},
      {"max_position_embeddings", &out->max_position_embeddings}, {"bos_token_id", &out->bos_token_id},
      {"eos_token_id", &out->eos_token_id}};
  double v;
  for (auto& f : req) {
    if (!jnum(root, f.key, v)) return hf_fail(NVC_EINVAL, "Failed to parse HF config " + path + ": missing field `" + f.key + "`");
    // serde reads these into usize / u32: negative, fractional or out-of-range numbers are parse errors there too
    if (!(v >= 0.0 && v <= 2147483647.0) || v != (double)(int64_t)v)
      return hf_fail(NVC_EINVAL, "Failed to parse HF config " + path + ": invalid value for `" + f.key + "`");
    *f.dst = (int32_t)v;
  }
  if (!jnum(root, "rms_norm_eps", v)) return hf_fail(NVC_EINVAL, "Failed to parse HF config " + path + ": missing field `rms_norm_eps`");
  out->rms_norm_eps = (float)v;
  const JValue* act = root.get("hidden_act");
  if (!act || act->kind != JValue::Str) return hf_fail(NVC_EINVAL, "Failed to parse HF config " + path + ": missing field `hidden_act`");
  if (act->str != "silu") return hf_fail(NVC_ENOTIMPL, "hidden_act " + act->str + " is not supported (SiluAndMul only, activation.rs:13-18)");
  if (out->num_attention_heads <= 0) return hf_fail(NVC_EINVAL, "num_attention_heads must be positive");
  if (jnum(root, "head_dim", v)) {                                                                        // :88-90
    if (!(v >= 1.0 && v <= 2147483647.0) || v != (double)(int64_t)v)
      return hf_fail(NVC_EINVAL, "Failed to parse HF config " + path + ": invalid value for `head_dim`");
    out->head_dim = (int32_t)v;
  } else {
    out->head_dim = out->hidden_size / out->num_attention_heads;
  }
  out->rope_theta = jnum(root, "rope_theta", v) ? (float)v : 1e6f;                                        // :96
  if (tie_word_embeddings) {
    const JValue* t = root.get("tie_word_embeddings");
    *tie_word_embeddings = (t && t->kind == JValue::Bool && t->b) ? 1 : 0;
  }
  return NVC_OK;
}

int nvc_runner_load_hf_dir(nvc_runner* r, const char* model_dir) {
  if (!r || !model_dir) return hf_fail(NVC_EINVAL, "null pointer");
  const std::string dir(model_dir);
  g_hf_warning.clear();
  std::set<std::string> seen;
  std::vector<uint16_t> stage;
  const std::string index = dir + "/model.safetensors.index.json";
  // file -> names to take from it (sharded checkpoints); a single file takes everything
  std::vector<std::pair<std::string, std::set<std::string>>> plan;
  if (file_exists(index)) {
    std::string text, err;
    JValue root;
    if (!read_file(index, text) || !parse_json(text.data(), text.size(), root, err))
      return hf_fail(NVC_EINVAL, "cannot parse " + index + ": " + err);
    const JValue* wm = root.get("weight_map");
    if (!wm || wm->kind != JValue::Obj) return hf_fail(NVC_EINVAL, index + ": no weight_map");
    for (auto& kv : wm->obj) {
      if (kv.second.kind != JValue::Str) return hf_fail(NVC_EINVAL, index + ": weight_map values must be file names");
      // shard names are joined onto the model directory: a plain file name only, never a path out of it
      const std::string& fn = kv.second.str;
      if (fn.empty() || fn.find('/') != std::string::npos || fn.find('\\') != std::string::npos || fn == "." || fn == ".." ||
          fn.find('\0') != std::string::npos)
        return hf_fail(NVC_EINVAL, index + ": shard name \"" + fn + "\" is not a plain file name");
      auto it = plan.begin();
      for (; it != plan.end(); ++it)
        if (it->first == kv.second.str) break;
      if (it == plan.end()) { plan.emplace_back(kv.second.str, std::set<std::string>()); it = plan.end() - 1; }
      it->second.insert(kv.first);
    }
  } else if (file_exists(dir + "/model.safetensors")) {
    plan.emplace_back("model.safetensors", std::set<std::string>());
  } else {
    return hf_fail(NVC_EINVAL, "no model.safetensors or model.safetensors.index.json in " + dir);
  }
  std::string embed_file;
  for (auto& f : plan) {
    std::unique_ptr<nvc_safetensors> st;
    int rc = open_safetensors(dir + "/" + f.first, st);
    if (rc) return rc;
    if (!f.second.empty())
      for (auto& name : f.second)
        if (!st->by_name.count(name)) return hf_fail(NVC_EINVAL, index + " maps " + name + " to " + f.first + ", which does not hold it");
    if (st->by_name.count("model.embed_tokens.weight")) embed_file = f.first;
    rc = load_file_into_runner(r, *st, f.second.empty() ? nullptr : &f.second, seen, stage);
    if (rc) return rc;
  }
  if (!seen.count("lm_head.weight")) {
    // tied embeddings (tie_word_embeddings = true, e.g. Qwen3-0.6B): the reference fails here
    // (qwen3.rs:525-528 requires lm_head.weight); lm_head = embed_tokens
    if (embed_file.empty()) return hf_fail(NVC_ESTATE, "checkpoint has neither lm_head.weight nor model.embed_tokens.weight");
    std::unique_ptr<nvc_safetensors> st;
    int rc = open_safetensors(dir + "/" + embed_file, st);
    if (rc) return rc;
    const auto& e = st->entries[st->by_name["model.embed_tokens.weight"]];
    if (e.shape.size() != 2 || e.dtype == NVC_ST_OTHER) return hf_fail(NVC_ESHAPE, "model.embed_tokens.weight: bad shape / dtype");
    stage.resize((size_t)e.shape[0] * e.shape[1]);
    to_bf16(st->data + e.begin, e.dtype, stage.size(), stage.data());
    rc = nvc_runner_load_weight(r, "lm_head.weight", stage.data(), e.shape[0], e.shape[1]);
    if (rc) return hf_fail(rc, "nvc_runner_load_weight failed for the tied lm_head.weight");
  }
  int rc = nvc_runner_finalize(r);
  if (rc) return hf_fail(rc, "nvc_runner_finalize failed after loading " + dir);
  return NVC_OK;
}

}  // extern "C"
