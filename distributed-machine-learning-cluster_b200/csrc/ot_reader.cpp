// See ot_reader.h.  Layout facts used here: general-purpose flag 0x808 means
// sizes live in the central directory, tensor payloads are STORED, `data.pkl` is STORED, names
// carry a `<stem>/` prefix, archives may or may not be ZIP64.
#include "ot_reader.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <memory>

namespace b2i {
namespace {

struct Blob {
  std::vector<uint8_t> bytes;
  bool load(const std::string& path, std::string* err) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) { *err = "cannot open " + path; return false; }
    std::fseek(f, 0, SEEK_END);
    long long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    if (n <= 0) { std::fclose(f); *err = "empty file " + path; return false; }
    bytes.resize(static_cast<size_t>(n));
    size_t got = std::fread(bytes.data(), 1, bytes.size(), f);
    std::fclose(f);
    if (got != bytes.size()) { *err = "short read on " + path; return false; }
    return true;
  }
};

inline uint16_t rd16(const uint8_t* p) { return static_cast<uint16_t>(p[0] | (p[1] << 8)); }
inline uint32_t rd32(const uint8_t* p) {
  return static_cast<uint32_t>(p[0]) | (static_cast<uint32_t>(p[1]) << 8) |
         (static_cast<uint32_t>(p[2]) << 16) | (static_cast<uint32_t>(p[3]) << 24);
}
inline uint64_t rd64(const uint8_t* p) {
  return static_cast<uint64_t>(rd32(p)) | (static_cast<uint64_t>(rd32(p + 4)) << 32);
}

struct ZipEntry {
  uint64_t offset = 0;  // payload offset in file
  uint64_t size = 0;
  int method = 0;
};

bool zip_index(const Blob& b, std::map<std::string, ZipEntry>* idx, std::string* err) {
  const std::vector<uint8_t>& v = b.bytes;
  const size_t n = v.size();
  if (n < 22) { *err = "not a zip archive (too small)"; return false; }
  // End-of-central-directory record: scan back over a possible trailing comment.
  size_t eocd = std::string::npos;
  const size_t lo = n > (22 + 65535) ? n - (22 + 65535) : 0;
  for (size_t i = n - 22 + 1; i-- > lo;) {
    if (rd32(&v[i]) == 0x06054b50u) { eocd = i; break; }
  }
  if (eocd == std::string::npos) { *err = "zip end-of-central-directory not found (is this a Git-LFS pointer?)"; return false; }
  uint64_t count = rd16(&v[eocd + 10]);
  uint64_t cd_size = rd32(&v[eocd + 12]);
  uint64_t cd_off = rd32(&v[eocd + 16]);
  if (eocd >= 20 && rd32(&v[eocd - 20]) == 0x07064b50u) {  // ZIP64 locator
    const uint64_t e64 = rd64(&v[eocd - 20 + 8]);
    if (e64 > n || n - e64 < 56 || rd32(&v[e64]) != 0x06064b50u) { *err = "bad zip64 end record"; return false; }
    count = rd64(&v[e64 + 32]);
    cd_size = rd64(&v[e64 + 40]);
    cd_off = rd64(&v[e64 + 48]);
  }
  // every range check below is written as "size > n || offset > n - size": the 64-bit fields come from the
  // file and a sum of two of them can wrap
  if (cd_size > n || cd_off > n - cd_size) { *err = "zip central directory out of range"; return false; }
  size_t p = static_cast<size_t>(cd_off);
  for (uint64_t i = 0; i < count; ++i) {
    if (p + 46 > n || rd32(&v[p]) != 0x02014b50u) { *err = "bad zip central directory entry"; return false; }
    const int method = rd16(&v[p + 10]);
    uint64_t csize = rd32(&v[p + 20]);
    uint64_t usize = rd32(&v[p + 24]);
    const size_t nlen = rd16(&v[p + 28]), elen = rd16(&v[p + 30]), clen = rd16(&v[p + 32]);
    uint64_t lho = rd32(&v[p + 42]);
    if (p + 46 + nlen + elen + clen > n) { *err = "zip entry overruns file"; return false; }
    std::string name(reinterpret_cast<const char*>(&v[p + 46]), nlen);
    // ZIP64 extended information (id 0x0001): present fields replace the 0xFFFFFFFF markers, in order
    size_t x = p + 46 + nlen;
    const size_t xend = x + elen;
    while (x + 4 <= xend) {
      const uint16_t id = rd16(&v[x]);
      const uint16_t sz = rd16(&v[x + 2]);
      const size_t rend = std::min(x + 4 + static_cast<size_t>(sz), xend);   // a record never reaches past the extra field
      if (id == 0x0001) {
        size_t q = x + 4;
        if (usize == 0xFFFFFFFFu && q + 8 <= rend) { usize = rd64(&v[q]); q += 8; }
        if (csize == 0xFFFFFFFFu && q + 8 <= rend) { csize = rd64(&v[q]); q += 8; }
        if (lho == 0xFFFFFFFFu && q + 8 <= rend) { lho = rd64(&v[q]); q += 8; }
      }
      x += 4 + static_cast<size_t>(sz);
    }
    if (lho > n || n - lho < 30 || rd32(&v[lho]) != 0x04034b50u) { *err = "bad zip local header for " + name; return false; }
    const size_t lnlen = rd16(&v[lho + 26]), lelen = rd16(&v[lho + 28]);
    if (n - lho < 30 + lnlen + lelen) { *err = "zip local header overruns file for " + name; return false; }
    ZipEntry e;
    e.offset = lho + 30 + lnlen + lelen;
    e.size = usize;
    e.method = method;
    // the stored size of a payload is its compressed size; for STORED entries the two agree
    const uint64_t stored = method == 0 ? usize : csize;
    if (stored > n || e.offset > n - stored) { *err = "zip payload out of range for " + name; return false; }
    (*idx)[name] = e;
    p += 46 + nlen + elen + clen;
  }
  return true;
}

// ------------------------------------------------------------------ minimal pickle VM
struct Val;
using VP = std::shared_ptr<Val>;
struct Val {
  enum Kind { NONE, INT, BOOL, FLOAT, STR, TUPLE, LIST, DICT, GLOBAL, STORAGE, TENSOR, OBJECT, MARK } kind = NONE;
  int64_t i = 0;
  double f = 0;
  std::string s, s2;              // STR: s; GLOBAL: module s, name s2; STORAGE: key s, dtype s2
  std::vector<VP> items;          // TUPLE / LIST; DICT: k0, v0, k1, v1, ...
  VP state;                       // OBJECT: BUILD state
  // TENSOR
  VP storage;
  int64_t offset = 0;
  std::vector<int64_t> sizes, strides;
};
VP mk(Val::Kind k) { auto v = std::make_shared<Val>(); v->kind = k; return v; }

struct Pickle {
  const uint8_t* p;
  const uint8_t* end;
  std::vector<VP> stack;
  std::map<uint32_t, VP> memo;
  std::string err;

  bool need(size_t n) {
    if (static_cast<size_t>(end - p) < n) { err = "pickle truncated"; return false; }
    return true;
  }
  bool pop(VP* v) {
    if (stack.empty()) { err = "pickle stack underflow"; return false; }
    *v = stack.back();
    stack.pop_back();
    return true;
  }
  bool pop_to_mark(std::vector<VP>* out) {
    size_t i = stack.size();
    while (i > 0 && stack[i - 1]->kind != Val::MARK) --i;
    if (i == 0) { err = "pickle MARK not found"; return false; }
    out->assign(stack.begin() + i, stack.end());
    stack.resize(i - 1);
    return true;
  }
  bool readline(std::string* s) {
    const uint8_t* q = p;
    while (q < end && *q != '\n') ++q;
    if (q == end) { err = "pickle truncated line"; return false; }
    s->assign(reinterpret_cast<const char*>(p), q - p);
    p = q + 1;
    return true;
  }
  static bool ints_of(const VP& t, std::vector<int64_t>* out) {
    if (t->kind != Val::TUPLE && t->kind != Val::LIST) return false;
    out->clear();
    for (const VP& e : t->items) {
      if (e->kind != Val::INT) return false;
      out->push_back(e->i);
    }
    return true;
  }

  bool reduce(const VP& fn, const VP& args, VP* res) {
    if (fn->kind != Val::GLOBAL || args->kind != Val::TUPLE) { err = "pickle REDUCE on non-global"; return false; }
    const std::string q = fn->s + "." + fn->s2;
    if (q == "torch._utils._rebuild_tensor_v2" || q == "torch._utils._rebuild_tensor") {
      if (args->items.size() < 4 || args->items[0]->kind != Val::STORAGE || args->items[1]->kind != Val::INT) {
        err = "unexpected _rebuild_tensor arguments"; return false;
      }
      VP t = mk(Val::TENSOR);
      t->storage = args->items[0];
      t->offset = args->items[1]->i;
      if (!ints_of(args->items[2], &t->sizes) || !ints_of(args->items[3], &t->strides) ||
          t->sizes.size() != t->strides.size()) {
        err = "unexpected tensor size/stride encoding"; return false;
      }
      *res = t;
      return true;
    }
    if (q == "torch._utils._rebuild_parameter") {
      if (args->items.empty()) { err = "bad _rebuild_parameter"; return false; }
      *res = args->items[0];
      return true;
    }
    if (q == "collections.OrderedDict") {
      *res = mk(Val::DICT);
      return true;
    }
    err = "unsupported pickle callable " + q;
    return false;
  }

  bool run(VP* result) {
    while (p < end) {
      const uint8_t op = *p++;
      switch (op) {
        case 0x80: if (!need(1)) return false; ++p; break;                       // PROTO
        case 0x95: if (!need(8)) return false; p += 8; break;                    // FRAME
        case 'c': {                                                              // GLOBAL
          VP g = mk(Val::GLOBAL);
          if (!readline(&g->s) || !readline(&g->s2)) return false;
          stack.push_back(g);
          break;
        }
        case 'q': if (!need(1)) return false; if (stack.empty()) { err = "BINPUT on empty stack"; return false; } memo[*p++] = stack.back(); break;
        case 'r': if (!need(4)) return false; if (stack.empty()) { err = "LONG_BINPUT on empty stack"; return false; } memo[rd32(p)] = stack.back(); p += 4; break;
        case 0x94: if (stack.empty()) { err = "MEMOIZE on empty stack"; return false; } { uint32_t k = static_cast<uint32_t>(memo.size()); memo[k] = stack.back(); } break;
        case 'h': { if (!need(1)) return false; auto it = memo.find(*p++); if (it == memo.end()) { err = "BINGET of unknown memo"; return false; } stack.push_back(it->second); break; }
        case 'j': { if (!need(4)) return false; auto it = memo.find(rd32(p)); p += 4; if (it == memo.end()) { err = "LONG_BINGET of unknown memo"; return false; } stack.push_back(it->second); break; }
        case '(': stack.push_back(mk(Val::MARK)); break;
        case ')': stack.push_back(mk(Val::TUPLE)); break;
        case ']': stack.push_back(mk(Val::LIST)); break;
        case '}': stack.push_back(mk(Val::DICT)); break;
        case 'N': stack.push_back(mk(Val::NONE)); break;
        case 0x88: { VP b = mk(Val::BOOL); b->i = 1; stack.push_back(b); break; }
        case 0x89: { VP b = mk(Val::BOOL); b->i = 0; stack.push_back(b); break; }
        case 'K': { if (!need(1)) return false; VP v = mk(Val::INT); v->i = *p++; stack.push_back(v); break; }
        case 'M': { if (!need(2)) return false; VP v = mk(Val::INT); v->i = rd16(p); p += 2; stack.push_back(v); break; }
        case 'J': { if (!need(4)) return false; VP v = mk(Val::INT); v->i = static_cast<int32_t>(rd32(p)); p += 4; stack.push_back(v); break; }
        case 0x8a: {                                                             // LONG1
          if (!need(1)) return false;
          const size_t n = *p++;
          if (n > 8 || !need(n)) { err = "LONG1 too wide"; return false; }
          uint64_t u = 0;
          for (size_t i = 0; i < n; ++i) u |= static_cast<uint64_t>(p[i]) << (8 * i);
          if (n > 0 && n < 8 && (p[n - 1] & 0x80)) u |= ~0ull << (8 * n);
          p += n;
          VP v = mk(Val::INT); v->i = static_cast<int64_t>(u); stack.push_back(v);
          break;
        }
        case 'G': {                                                              // BINFLOAT (big endian)
          if (!need(8)) return false;
          uint64_t u = 0;
          for (int i = 0; i < 8; ++i) u = (u << 8) | p[i];
          p += 8;
          VP v = mk(Val::FLOAT); std::memcpy(&v->f, &u, 8); stack.push_back(v);
          break;
        }
        case 'X': {                                                              // BINUNICODE
          if (!need(4)) return false;
          const uint32_t n = rd32(p); p += 4;
          if (!need(n)) return false;
          VP v = mk(Val::STR); v->s.assign(reinterpret_cast<const char*>(p), n); p += n; stack.push_back(v);
          break;
        }
        case 0x8c: {                                                             // SHORT_BINUNICODE
          if (!need(1)) return false;
          const uint32_t n = *p++;
          if (!need(n)) return false;
          VP v = mk(Val::STR); v->s.assign(reinterpret_cast<const char*>(p), n); p += n; stack.push_back(v);
          break;
        }
        case 't': { VP t = mk(Val::TUPLE); if (!pop_to_mark(&t->items)) return false; stack.push_back(t); break; }
        case 0x85: case 0x86: case 0x87: {
          const size_t n = op - 0x84;
          if (stack.size() < n) { err = "TUPLEn underflow"; return false; }
          VP t = mk(Val::TUPLE);
          t->items.assign(stack.end() - n, stack.end());
          stack.resize(stack.size() - n);
          stack.push_back(t);
          break;
        }
        case 'a': { VP v, l; if (!pop(&v)) return false; if (stack.empty() || stack.back()->kind != Val::LIST) { err = "APPEND to non-list"; return false; } stack.back()->items.push_back(v); (void)l; break; }
        case 'e': { std::vector<VP> it; if (!pop_to_mark(&it)) return false; if (stack.empty() || stack.back()->kind != Val::LIST) { err = "APPENDS to non-list"; return false; } for (auto& v : it) stack.back()->items.push_back(v); break; }
        case 0x81: {                                                             // NEWOBJ
          VP args, cls;
          if (!pop(&args) || !pop(&cls)) return false;
          VP o = mk(Val::OBJECT);
          if (cls->kind == Val::GLOBAL) { o->s = cls->s; o->s2 = cls->s2; }
          stack.push_back(o);
          break;
        }
        case 'Q': {                                                              // BINPERSID
          VP pid;
          if (!pop(&pid)) return false;
          if (pid->kind != Val::TUPLE || pid->items.size() < 5 || pid->items[0]->kind != Val::STR ||
              pid->items[0]->s != "storage" || pid->items[1]->kind != Val::GLOBAL ||
              pid->items[2]->kind != Val::STR || pid->items[4]->kind != Val::INT) {
            err = "unexpected persistent id (want ('storage', type, key, device, numel))"; return false;
          }
          VP st = mk(Val::STORAGE);
          st->s = pid->items[2]->s;
          st->s2 = pid->items[1]->s2;
          st->i = pid->items[4]->i;
          stack.push_back(st);
          break;
        }
        case 'R': { VP args, fn, res; if (!pop(&args) || !pop(&fn)) return false; if (!reduce(fn, args, &res)) return false; stack.push_back(res); break; }
        case 's': {
          VP v, k;
          if (!pop(&v) || !pop(&k)) return false;
          if (stack.empty() || stack.back()->kind != Val::DICT) { err = "SETITEM on non-dict"; return false; }
          stack.back()->items.push_back(k); stack.back()->items.push_back(v);
          break;
        }
        case 'u': {
          std::vector<VP> it;
          if (!pop_to_mark(&it)) return false;
          if (stack.empty() || stack.back()->kind != Val::DICT || (it.size() & 1)) { err = "SETITEMS on non-dict"; return false; }
          for (auto& v : it) stack.back()->items.push_back(v);
          break;
        }
        case 'b': {                                                              // BUILD
          VP st, obj;
          if (!pop(&st)) return false;
          if (stack.empty()) { err = "BUILD on empty stack"; return false; }
          obj = stack.back();
          if (obj->kind == Val::OBJECT) obj->state = st;
          else if (obj->kind == Val::DICT && st->kind == Val::DICT) for (auto& v : st->items) obj->items.push_back(v);
          break;
        }
        case '.': return pop(result);
        default: {
          char buf[64];
          std::snprintf(buf, sizeof buf, "unsupported pickle opcode 0x%02x", op);
          err = buf;
          return false;
        }
      }
    }
    err = "pickle ended without STOP";
    return false;
  }
};

float half_to_float(uint16_t h) {
  const uint32_t sign = (h & 0x8000u) << 16;
  uint32_t exp = (h >> 10) & 0x1F, man = h & 0x3FF, bits;
  if (exp == 0) {
    if (man == 0) bits = sign;
    else {
      exp = 127 - 15 + 1;
      while (!(man & 0x400)) { man <<= 1; --exp; }
      bits = sign | (exp << 23) | ((man & 0x3FF) << 13);
    }
  } else if (exp == 31) bits = sign | 0x7F800000u | (man << 13);
  else bits = sign | ((exp + 127 - 15) << 23) | (man << 13);
  float f;
  std::memcpy(&f, &bits, 4);
  return f;
}

int dtype_size(const std::string& st) {
  if (st == "FloatStorage") return 4;
  if (st == "DoubleStorage") return 8;
  if (st == "HalfStorage" || st == "BFloat16Storage") return 2;
  return 0;
}

float load_elem(const uint8_t* base, const std::string& st, int64_t i) {
  if (st == "FloatStorage") { float f; std::memcpy(&f, base + 4 * i, 4); return f; }
  if (st == "DoubleStorage") { double d; std::memcpy(&d, base + 8 * i, 8); return static_cast<float>(d); }
  if (st == "HalfStorage") return half_to_float(rd16(base + 2 * i));
  uint32_t b = static_cast<uint32_t>(rd16(base + 2 * i)) << 16;
  float f; std::memcpy(&f, &b, 4); return f;
}

void collect(const VP& v, const std::string& prefix, std::vector<std::pair<std::string, VP>>* out, int depth = 0) {
  if (!v || depth > 64) return;   // module trees are a few levels deep; a crafted archive must not recurse without bound
  if (v->kind == Val::OBJECT) { collect(v->state, prefix, out, depth + 1); return; }
  if (v->kind == Val::DICT) {
    for (size_t i = 0; i + 1 < v->items.size(); i += 2) {
      const VP& k = v->items[i];
      const VP& val = v->items[i + 1];
      if (k->kind != Val::STR) continue;
      const std::string name = prefix.empty() ? k->s : prefix + "." + k->s;
      if (val->kind == Val::TENSOR) out->emplace_back(name, val);
      else if (val->kind == Val::DICT || val->kind == Val::OBJECT) collect(val, name, out, depth + 1);
    }
  }
}

}  // namespace

bool read_ot_archive(const std::string& path, std::map<std::string, HostTensor>* out,
                     std::string* err) {
  Blob blob;
  if (!blob.load(path, err)) return false;
  if (blob.bytes.size() < 200 && std::memcmp(blob.bytes.data(), "version https://git-lfs", 23) == 0) {
    *err = path + " is a Git-LFS pointer, not the weight archive";
    return false;
  }
  std::map<std::string, ZipEntry> idx;
  if (!zip_index(blob, &idx, err)) return false;
  std::string stem;
  const ZipEntry* pkl = nullptr;
  for (auto& kv : idx) {
    const std::string& n = kv.first;
    if (n.size() >= 8 && n.compare(n.size() - 8, 8, "data.pkl") == 0 &&
        (n.size() == 8 || n[n.size() - 9] == '/')) {
      stem = n.substr(0, n.size() - 8);
      pkl = &kv.second;
      break;
    }
  }
  if (!pkl) { *err = "archive has no data.pkl"; return false; }
  if (pkl->method != 0) { *err = "data.pkl is compressed; expected STORED"; return false; }
  Pickle vm;
  vm.p = blob.bytes.data() + pkl->offset;
  vm.end = vm.p + pkl->size;
  VP root;
  if (!vm.run(&root)) { *err = "data.pkl: " + vm.err; return false; }
  std::vector<std::pair<std::string, VP>> tensors;
  collect(root, "", &tensors);
  if (tensors.empty()) { *err = "data.pkl holds no tensors"; return false; }
  for (auto& nt : tensors) {
    const VP& t = nt.second;
    const VP& st = t->storage;
    const int esz = dtype_size(st->s2);
    if (!esz) { *err = "unsupported storage type " + st->s2 + " for " + nt.first; return false; }
    auto it = idx.find(stem + "data/" + st->s);
    if (it == idx.end()) { *err = "missing payload data/" + st->s + " for " + nt.first; return false; }
    if (it->second.method != 0) { *err = "tensor payload is compressed: " + nt.first; return false; }
    const int64_t avail = static_cast<int64_t>(it->second.size / esz);
    HostTensor h;
    h.shape = t->sizes;
    // sizes, strides and the offset come from the file: every product / sum is checked against `avail`
    // (at most the file size) before it is formed, so nothing here can wrap
    int64_t numel = 1;
    bool empty = false;
    for (int64_t d : h.shape) {
      if (d < 0) { *err = "negative dimension in " + nt.first; return false; }
      if (d == 0) empty = true;
    }
    if (empty) numel = 0;
    else
      for (int64_t d : h.shape) {
        if (numel > avail / d) { *err = "tensor larger than its storage: " + nt.first; return false; }
        numel *= d;
      }
    if (t->offset < 0 || t->offset > avail) { *err = "tensor view exceeds its storage: " + nt.first; return false; }
    int64_t max_index = t->offset;
    for (size_t d = 0; d < h.shape.size() && !empty; ++d) {
      if (t->strides[d] < 0) { *err = "negative stride in " + nt.first; return false; }
      const int64_t steps = h.shape[d] - 1;
      if (steps > 0 && t->strides[d] > (avail - max_index) / steps) { *err = "tensor view exceeds its storage: " + nt.first; return false; }
      max_index += steps * t->strides[d];
    }
    if (numel > 0 && max_index >= avail) { *err = "tensor view exceeds its storage: " + nt.first; return false; }
    h.data.resize(static_cast<size_t>(numel));
    const uint8_t* base = blob.bytes.data() + it->second.offset;
    std::vector<int64_t> ix(h.shape.size(), 0);
    for (int64_t i = 0; i < numel; ++i) {
      int64_t src = t->offset;
      for (size_t d = 0; d < ix.size(); ++d) src += ix[d] * t->strides[d];
      h.data[static_cast<size_t>(i)] = load_elem(base, st->s2, src);
      for (size_t d = ix.size(); d-- > 0;) {
        if (++ix[d] < h.shape[d]) break;
        ix[d] = 0;
      }
    }
    std::string name = nt.first;
    for (char& c : name) if (c == '|') c = '.';
    (*out)[name] = std::move(h);
  }
  return true;
}

}  // namespace b2i
