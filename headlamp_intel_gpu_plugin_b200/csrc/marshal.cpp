// marshal.cpp — native host marshal: Kubernetes JSON text -> the columnar layout of igpu_fleet_cols
// (SURVEY.md §8(f) N1).  Host code only (no CUDA); part of libigpu_rollup.so.
//
// What it replaces on the reference side: the per-object field access that feeds every helper of
// src/api/k8s.ts — labels / capacity (k8s.ts:128-149), conditions (:330), containers + initContainers and
// their resources.requests / limits (:254-261, :291-297), spec.nodeName (NodesPage.tsx:155), status.phase
// (OverviewPage.tsx:106,123) — plus JSON.parse itself and `extractJsonData` (IntelGpuDataContext.tsx:85-90).
// The JS number semantics applied here (parseInt(v,10) || 0, truthiness, `=== 'true'`, ?? '0') are the
// same ones k8s.py applies; tests/test_marshal_native.py requires column-for-column equality with it.
//
// Method: no DOM.  One pass finds the top-level item boundaries (brace depth, string aware); the items are
// then parsed in parallel by a cursor that descends only into the handful of paths the rollup reads and
// skips everything else.  Strings are compared in place unless they contain a backslash.
// Documented deviations from JSON.parse (malformed input only): a key repeated inside one object is
// processed each time it occurs (JSON.parse keeps the last); integer-like object keys are not reordered.
#include <algorithm>
#include <atomic>
#include <memory>
#include <charconv>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <string_view>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../../include/igpu_rollup.h"

#if defined(__AVX2__)
#include <immintrin.h>
#endif

int32_t igpu_fail(struct igpu_ctx* ctx, int32_t status, const char* fmt, ...);  // ctx.cu (NULL ctx = thread-local text)

namespace {

using sv = std::string_view;

// This file is built with -mavx2 -mpclmul (csrc/Makefile): on a host without them it must refuse, not die of SIGILL.
bool host_cpu_ok() {
#if defined(__AVX2__) && defined(__x86_64__)
  static const bool ok = __builtin_cpu_supports("avx2") && __builtin_cpu_supports("pclmul");
  return ok;
#else
  return true;
#endif
}

constexpr sv K_PREFIX = "gpu.intel.com/";
constexpr sv K_I915 = "gpu.intel.com/i915", K_XE = "gpu.intel.com/xe";
constexpr sv L_NFD = "intel.feature.node.kubernetes.io/gpu", L_DISCRETE = "node-role.kubernetes.io/gpu",
             L_IGPU = "node-role.kubernetes.io/igpu";
const char* const DEFAULT_RES[5] = {"gpu.intel.com/i915", "gpu.intel.com/xe", "gpu.intel.com/millicores", "gpu.intel.com/memory.max",
                                    "gpu.intel.com/tiles"};

struct ParseError {
  std::string msg;
  size_t pos;
};

// ---- JSON cursor ------------------------------------------------------------------------------------
struct Cur {
  const char* p;
  const char* end;
  const char* base;
  [[noreturn]] void fail(const char* what) const { throw ParseError{what, (size_t)(p - base)}; }
  void ws() {
    while (p < end && (*p == ' ' || *p == '\n' || *p == '\r' || *p == '\t')) ++p;
  }
  char peek() {
    ws();
    if (p >= end) fail("unexpected end of JSON");
    return *p;
  }
  void expect(char c) {
    if (peek() != c) fail("unexpected character");
    ++p;
  }
  bool consume(char c) {
    if (peek() == c) {
      ++p;
      return true;
    }
    return false;
  }
};

void append_utf8(std::string& out, uint32_t cp) {
  if (cp < 0x80) out.push_back((char)cp);
  else if (cp < 0x800) out.push_back((char)(0xC0 | (cp >> 6))), out.push_back((char)(0x80 | (cp & 0x3F)));
  else if (cp < 0x10000)
    out.push_back((char)(0xE0 | (cp >> 12))), out.push_back((char)(0x80 | ((cp >> 6) & 0x3F))), out.push_back((char)(0x80 | (cp & 0x3F)));
  else
    out.push_back((char)(0xF0 | (cp >> 18))), out.push_back((char)(0x80 | ((cp >> 12) & 0x3F))),
        out.push_back((char)(0x80 | ((cp >> 6) & 0x3F))), out.push_back((char)(0x80 | (cp & 0x3F)));
}

uint32_t hex4(Cur& c) {
  if (c.end - c.p < 4) c.fail("truncated \\u escape");
  uint32_t v = 0;
  for (int i = 0; i < 4; ++i) {
    const char ch = *c.p++;
    v <<= 4;
    if (ch >= '0' && ch <= '9') v |= ch - '0';
    else if (ch >= 'a' && ch <= 'f') v |= ch - 'a' + 10;
    else if (ch >= 'A' && ch <= 'F') v |= ch - 'A' + 10;
    else c.fail("bad \\u escape");
  }
  return v;
}

inline const char* find_quote_or_backslash(const char* p, const char* end);
inline const char* find_structural(const char* p, const char* end);

// Parses a JSON string.  The view points into the input when there is no escape, else into `scratch`.
sv parse_string(Cur& c, std::string& scratch) {
  c.expect('"');
  const char* s = c.p;
  c.p = find_quote_or_backslash(c.p, c.end);
  if (c.p >= c.end) c.fail("unterminated string");
  if (*c.p == '"') {
    sv v(s, c.p - s);
    ++c.p;
    return v;
  }
  scratch.assign(s, c.p - s);
  while (c.p < c.end && *c.p != '"') {
    if (*c.p != '\\') {
      scratch.push_back(*c.p++);
      continue;
    }
    ++c.p;
    if (c.p >= c.end) c.fail("unterminated escape");
    const char e = *c.p++;
    switch (e) {
      case '"': scratch.push_back('"'); break;
      case '\\': scratch.push_back('\\'); break;
      case '/': scratch.push_back('/'); break;
      case 'b': scratch.push_back('\b'); break;
      case 'f': scratch.push_back('\f'); break;
      case 'n': scratch.push_back('\n'); break;
      case 'r': scratch.push_back('\r'); break;
      case 't': scratch.push_back('\t'); break;
      case 'u': {
        uint32_t cp = hex4(c);
        if (cp >= 0xD800 && cp < 0xDC00 && c.end - c.p >= 6 && c.p[0] == '\\' && c.p[1] == 'u') {
          Cur t = c;
          t.p += 2;
          const uint32_t lo = hex4(t);
          if (lo >= 0xDC00 && lo < 0xE000) {
            cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
            c = t;
          }
        }
        append_utf8(scratch, cp);
        break;
      }
      default: c.fail("bad escape");
    }
  }
  if (c.p >= c.end) c.fail("unterminated string");
  ++c.p;
  return sv(scratch);
}

// First byte in [p, end) that is a quote or a backslash (end if none): 32 bytes per step with AVX2.
inline const char* find_quote_or_backslash(const char* p, const char* end) {
#if defined(__AVX2__)
  const __m256i q = _mm256_set1_epi8('"'), b = _mm256_set1_epi8('\\');
  while (end - p >= 32) {
    const __m256i v = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(p));
    const uint32_t m = (uint32_t)_mm256_movemask_epi8(_mm256_or_si256(_mm256_cmpeq_epi8(v, q), _mm256_cmpeq_epi8(v, b)));
    if (m) return p + __builtin_ctz(m);
    p += 32;
  }
#endif
  while (p < end && *p != '"' && *p != '\\') ++p;
  return p;
}

// First byte in [p, end) that is one of  " { } [ ]  (end if none).
inline const char* find_structural(const char* p, const char* end) {
#if defined(__AVX2__)
  const __m256i q = _mm256_set1_epi8('"'), lb = _mm256_set1_epi8('['), rb = _mm256_set1_epi8(']');
  const __m256i lc = _mm256_set1_epi8('{'), rc = _mm256_set1_epi8('}');
  while (end - p >= 32) {
    const __m256i v = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(p));
    // '[' 0x5B / ']' 0x5D and '{' 0x7B / '}' 0x7D differ only in bit 1: fold each pair into one compare
    const __m256i v2 = _mm256_or_si256(v, _mm256_set1_epi8(0x06));
    const __m256i hit = _mm256_or_si256(_mm256_cmpeq_epi8(v, q),
                                        _mm256_or_si256(_mm256_cmpeq_epi8(v2, _mm256_or_si256(lb, _mm256_set1_epi8(0x06))),
                                                        _mm256_cmpeq_epi8(v2, _mm256_or_si256(lc, _mm256_set1_epi8(0x06)))));
    uint32_t m = (uint32_t)_mm256_movemask_epi8(hit);
    while (m) {  // v | 6 also matches 0x59..0x5F / 0x79..0x7F neighbours: confirm the candidate
      const char ch = p[__builtin_ctz(m)];
      if (ch == '"' || ch == '{' || ch == '}' || ch == '[' || ch == ']') return p + __builtin_ctz(m);
      m &= m - 1;
    }
    p += 32;
  }
  (void)rb;
  (void)rc;
#endif
  while (p < end && *p != '"' && *p != '{' && *p != '}' && *p != '[' && *p != ']') ++p;
  return p;
}

void skip_string(Cur& c) {
  c.expect('"');
  for (;;) {
    c.p = find_quote_or_backslash(c.p, c.end);
    if (c.p >= c.end) c.fail("unterminated string");
    if (*c.p == '"') break;
    c.p += 2;  // backslash + the escaped character (a \uXXXX tail contains neither quote nor backslash)
    if (c.p > c.end) c.fail("unterminated string");
  }
  ++c.p;
}

void skip_value(Cur& c) {
  const char ch = c.peek();
  if (ch == '"') return skip_string(c);
  if (ch == '{' || ch == '[') {
    int depth = 0;
    for (;;) {
      c.p = find_structural(c.p, c.end);
      if (c.p >= c.end) c.fail("unterminated container");
      const char x = *c.p;
      if (x == '"') {
        skip_string(c);
        continue;
      }
      if (x == '{' || x == '[') ++depth;
      else if (--depth == 0) {
        ++c.p;
        return;
      }
      ++c.p;
    }
  }
  while (c.p < c.end && *c.p != ',' && *c.p != '}' && *c.p != ']' && *c.p != ' ' && *c.p != '\n' && *c.p != '\r' && *c.p != '\t') ++c.p;
}

// A JSON value as far as the rollup cares: what parseInt / truthiness / === need.
struct Val {
  enum Kind : uint8_t { ABSENT, NUL, STR, NUM, TRUE_, FALSE_, OTHER } kind = ABSENT;
  std::string s;  // STR (unescaped copy: node / resource maps are tiny)
  double num = 0;
};

Val parse_val(Cur& c, std::string& scratch) {
  Val v;
  const char ch = c.peek();
  if (ch == '"') {
    v.kind = Val::STR;
    v.s = std::string(parse_string(c, scratch));
  } else if (ch == '{' || ch == '[') {
    v.kind = Val::OTHER;
    skip_value(c);
  } else {
    const char* s = c.p;
    skip_value(c);
    const sv tok(s, c.p - s);
    if (tok == "null") v.kind = Val::NUL;
    else if (tok == "true") v.kind = Val::TRUE_;
    else if (tok == "false") v.kind = Val::FALSE_;
    else {
      v.kind = Val::NUM;
      char* e = nullptr;
      std::string t(tok);
      v.num = std::strtod(t.c_str(), &e);
      if (e == t.c_str()) c.fail("bad JSON literal");
    }
  }
  return v;
}

bool truthy(const Val& v) {
  switch (v.kind) {
    case Val::ABSENT: case Val::NUL: case Val::FALSE_: return false;
    case Val::STR: return !v.s.empty();
    case Val::NUM: return !(v.num == 0 || v.num != v.num);
    default: return true;
  }
}

// ECMA-262 StrWhiteSpaceChar in UTF-8: returns bytes consumed (0 = not whitespace)
size_t js_ws_len(const char* p, const char* end) {
  const unsigned char c = (unsigned char)*p;
  if (c == ' ' || (c >= 9 && c <= 13)) return 1;
  if (c == 0xC2 && end - p >= 2 && (unsigned char)p[1] == 0xA0) return 2;  // U+00A0
  if (end - p >= 3) {
    const unsigned char b1 = (unsigned char)p[1], b2 = (unsigned char)p[2];
    if (c == 0xE1 && b1 == 0x9A && b2 == 0x80) return 3;                                   // U+1680
    if (c == 0xE2 && b1 == 0x80 && ((b2 >= 0x80 && b2 <= 0x8A) || b2 == 0xA8 || b2 == 0xA9 || b2 == 0xAF)) return 3;  // U+2000-200A, 2028, 2029, 202F
    if (c == 0xE2 && b1 == 0x81 && b2 == 0x9F) return 3;                                   // U+205F
    if (c == 0xE3 && b1 == 0x80 && b2 == 0x80) return 3;                                   // U+3000
    if (c == 0xEF && b1 == 0xBB && b2 == 0xBF) return 3;                                   // U+FEFF
  }
  return 0;
}

// parseInt(str, 10) as a double (NaN if no digits)
double parse_int_str(sv s) {
  const char* p = s.data();
  const char* end = p + s.size();
  while (p < end) {
    const size_t n = js_ws_len(p, end);
    if (!n) break;
    p += n;
  }
  double sign = 1;
  if (p < end && (*p == '+' || *p == '-')) sign = (*p++ == '-') ? -1 : 1;
  const char* d = p;
  double v = 0;
  while (p < end && *p >= '0' && *p <= '9') v = v * 10 + (*p++ - '0');  // exact below 2^53; larger values are rejected anyway
  if (p == d) return NAN;
  return sign * v;
}

// parseInt(String(number), 10)
double parse_int_num(double x) {
  if (x != x) return NAN;
  if (std::isinf(x)) return NAN;  // "Infinity"
  const double a = std::fabs(x);
  if (a != 0 && (a >= 1e21 || a < 1e-6)) {  // String() uses exponent notation: parseInt reads the leading digit(s)
    char buf[64];
    auto r = std::to_chars(buf, buf + sizeof buf, x, std::chars_format::scientific);
    return parse_int_str(sv(buf, r.ptr - buf).substr(0, sv(buf, r.ptr - buf).find_first_of(".e")));
  }
  return std::trunc(x);
}

// parseInt(v, 10) for any JSON value
double parse_int(const Val& v) {
  switch (v.kind) {
    case Val::STR: return parse_int_str(v.s);
    case Val::NUM: return parse_int_num(v.num);
    default: return NAN;  // "null", "true", "false", "undefined", "[object Object]"
  }
}

struct RangeError {
  std::string msg;
};

int64_t int_or_zero(const Val& v) {  // parseInt(v, 10) || 0
  const double x = parse_int(v);
  if (x != x || x == 0) return 0;
  if (std::fabs(x) >= 9007199254740992.0) throw RangeError{"a quantity is outside the exact-integer domain (|x| < 2^53) of the GPU path"};
  return (int64_t)x;
}

int32_t fits_i32(int64_t x) {
  if (x < INT32_MIN || x > INT32_MAX) throw RangeError{"a node quantity does not fit the int32 node columns"};
  return (int32_t)x;
}

// ---- item boundaries ---------------------------------------------------------------------------------
struct Span {
  const char* b;
  const char* e;
};

#if defined(__AVX2__) && defined(__PCLMUL__)
// 64 bytes -> bit masks of the bytes equal to c
inline uint64_t eq_mask(const __m256i lo, const __m256i hi, char c) {
  const __m256i k = _mm256_set1_epi8(c);
  return (uint64_t)(uint32_t)_mm256_movemask_epi8(_mm256_cmpeq_epi8(lo, k)) |
         ((uint64_t)(uint32_t)_mm256_movemask_epi8(_mm256_cmpeq_epi8(hi, k)) << 32);
}
// prefix XOR of the bits of m (bit i = XOR of bits 0..i): carry-less multiply by all ones
inline uint64_t prefix_xor(uint64_t m) {
  const __m128i r = _mm_clmulepi64_si128(_mm_set_epi64x(0, (long long)m), _mm_set1_epi8((char)0xFF), 0);
  return (uint64_t)_mm_cvtsi128_si64(r);
}

// Top-level elements of the JSON array whose '[' has just been consumed at `p`: a simdjson-style scan,
// 64 bytes at a time — unescaped quotes give the in-string mask by prefix XOR, braces and brackets outside
// strings drive a depth counter, and an element is [position where depth leaves 0, position where it
// returns to 0].  Elements must be objects or arrays (Kubernetes list items always are); returns false
// (caller falls back to the scalar splitter) if the array does not close.
template <class Sink>
bool split_array_fast_to(const char* p, const char* end, Sink&& sink, const char** close_pos = nullptr) {
  uint64_t in_string_carry = 0;  // all ones while inside a string at the block boundary
  bool esc_carry = false;        // previous block ended with an unpaired backslash
  int depth = 0;
  const char* begin = nullptr;
  // The block scan only sees containers.  Whatever lies between two of them at depth 0 — normally just a comma — is
  // walked by the scalar cursor, so scalar elements (null, numbers, strings: junk in a Kubernetes list, but JSON) become
  // items too, in order, exactly as JSON.parse would list them.
  const char* prev_end = p;
  auto emit_gap = [&](const char* gap_end) {
    Cur g{prev_end, gap_end, p};
    for (;;) {
      g.ws();
      if (g.p >= g.end) break;
      if (*g.p == ',') {
        ++g.p;
        continue;
      }
      const char* b = g.p;
      skip_value(g);
      if (g.p == b) break;  // nothing a scalar could start with (malformed): leave it to the item parsers
      sink(Span{b, g.p});
    }
  };
  alignas(32) char tail[64];
  for (const char* blk = p; blk < end; blk += 64) {
    const char* src = blk;
    if (end - blk < 64) {
      memset(tail, ' ', sizeof tail);
      memcpy(tail, blk, (size_t)(end - blk));
      src = tail;
    }
    const __m256i lo = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(src));
    const __m256i hi = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(src + 32));
    uint64_t quote = eq_mask(lo, hi, '"');
    const uint64_t bs = eq_mask(lo, hi, '\\');
    if (bs || esc_carry) {  // rare: work out which characters are escaped, byte by byte
      uint64_t escaped = 0;
      bool esc = esc_carry;
      for (int i = 0; i < 64; ++i) {
        if (esc) escaped |= 1ull << i, esc = false;
        else if ((bs >> i) & 1) esc = true;
      }
      esc_carry = esc;
      quote &= ~escaped;
    }
    const uint64_t in_string = prefix_xor(quote) ^ in_string_carry;
    in_string_carry = (uint64_t)((int64_t)in_string >> 63);
    const uint64_t open = (eq_mask(lo, hi, '{') | eq_mask(lo, hi, '[')) & ~in_string;
    const uint64_t close = (eq_mask(lo, hi, '}') | eq_mask(lo, hi, ']')) & ~in_string;
    uint64_t both = open | close;
    while (both) {
      const int i = __builtin_ctzll(both);
      both &= both - 1;
      if ((open >> i) & 1) {
        if (depth++ == 0) {
          begin = blk + i;
          bool plain = true;  // the usual gap: a comma and white space
          for (const char* q = prev_end; q < begin; ++q) plain &= *q == ',' || *q == ' ' || *q == '\n' || *q == '\r' || *q == '\t';
          if (!plain) emit_gap(begin);
        }
      } else {
        if (depth == 0) {  // the ']' that closes the array itself
          if (blk + i > prev_end) emit_gap(blk + i);
          if (close_pos) *close_pos = blk + i;
          return true;
        }
        if (--depth == 0) {
          sink(Span{begin, blk + i + 1});
          prev_end = blk + i + 1;
        }
      }
    }
  }
  return false;
}
inline bool split_array_fast(const char* p, const char* end, std::vector<Span>& out, const char** close_pos = nullptr) {
  return split_array_fast_to(p, end, [&](Span sp) { out.push_back(sp); }, close_pos);
}
#endif

// Accepts `[ {...}, ... ]` or `{ ..., "items": [ {...}, ... ], ... }` (a KubeList, k8s.ts:315-323).
template <class Sink>
void split_items_to(const char* text, size_t len, Sink&& sink) {
  Cur c{text, text + len, text};
  std::string scratch;
  if (len == 0) return;
  if (c.peek() == '{') {
    ++c.p;
    bool found = false;
    if (!c.consume('}')) {
      do {
        const sv key = parse_string(c, scratch);
        c.expect(':');
        if (key == "items" && c.peek() == '[') {
          found = true;
          break;
        }
        skip_value(c);
      } while (c.consume(','));
    }
    if (!found) return;
  }
  c.expect('[');
  if (c.consume(']')) return;
#if defined(__AVX2__) && defined(__PCLMUL__)
  {
    const char first = c.peek();
    if (first == '{' || first == '[') {
      // The fast scan emits as it goes; if it then fails to find the end of the array the slow path below would
      // emit everything again, so a failed fast scan is a hard error (the text is not a closed JSON array).
      if (split_array_fast_to(c.p, c.end, sink)) return;
      c.fail("unterminated JSON array");
    }
  }
#endif
  do {
    c.ws();
    const char* b = c.p;
    skip_value(c);
    sink(Span{b, c.p});
  } while (c.consume(','));
  c.expect(']');
}

std::vector<Span> split_items(const char* text, size_t len) {
  std::vector<Span> out;
  out.reserve(len / 400 + 16);
  split_items_to(text, len, [&](Span sp) { out.push_back(sp); });
  return out;
}

// ---- nodes ----------------------------------------------------------------------------------------------
struct NodeRow {
  uint8_t flags = 0;
  int32_t cap_i915 = 0, cap_xe = 0, alloc_i915 = 0, alloc_xe = 0, alloc_own = 0;
  std::string name;
  bool has_name = false;
};

template <class F>
void for_each_member(Cur& c, std::string& scratch, F f) {  // f(key) must consume the value
  if (c.peek() != '{') {
    skip_value(c);
    return;
  }
  ++c.p;
  if (c.consume('}')) return;
  do {
    std::string keybuf;
    const sv key = parse_string(c, keybuf);
    c.expect(':');
    f(key);
  } while (c.consume(','));
  c.expect('}');
  (void)scratch;
}

NodeRow parse_node(Span sp, const char* base) {
  NodeRow n;
  Cur c{sp.b, sp.e, base};
  std::string scratch;
  if (c.peek() != '{') return n;  // not an object: contributes an all-zero row, like k8s.py
  Val cap[2], alloc[2];
  bool gpu_cap = false;
  for_each_member(c, scratch, [&](sv key) {
    if (key == "metadata") {
      for_each_member(c, scratch, [&](sv k2) {
        if (k2 == "name") {
          Val v = parse_val(c, scratch);
          if (v.kind == Val::STR) n.name = v.s, n.has_name = true;
        } else if (k2 == "labels") {
          for_each_member(c, scratch, [&](sv lk) {
            const bool hit = lk == L_DISCRETE || lk == L_IGPU || lk == L_NFD;
            if (!hit) return skip_value(c);
            std::string lkey(lk);
            Val v = parse_val(c, scratch);
            if (v.kind == Val::STR && v.s == "true")
              n.flags |= lkey == L_DISCRETE ? IGPU_NODE_DISCRETE_ROLE : (lkey == L_IGPU ? IGPU_NODE_IGPU_ROLE : IGPU_NODE_NFD_LABEL);
          });
        } else skip_value(c);
      });
    } else if (key == "status") {
      for_each_member(c, scratch, [&](sv k2) {
        if (k2 == "capacity" || k2 == "allocatable") {
          const bool is_cap = k2 == "capacity";
          for_each_member(c, scratch, [&](sv rk) {
            if (rk.substr(0, K_PREFIX.size()) != K_PREFIX) return skip_value(c);
            if (is_cap) gpu_cap = true;
            const int which = rk == K_I915 ? 0 : (rk == K_XE ? 1 : -1);
            if (which < 0) return skip_value(c);
            Val v = parse_val(c, scratch);
            if (v.kind == Val::ABSENT) v.kind = Val::NUL;
            (is_cap ? cap : alloc)[which] = std::move(v);
          });
        } else if (k2 == "conditions" && c.peek() == '[') {
          ++c.p;
          if (!c.consume(']')) {
            do {
              std::string type, status;
              bool t_str = false, s_str = false;
              for_each_member(c, scratch, [&](sv ck) {
                if (ck == "type" || ck == "status") {
                  const bool is_type = ck == "type";
                  Val v = parse_val(c, scratch);
                  if (v.kind == Val::STR) (is_type ? type : status) = v.s, (is_type ? t_str : s_str) = true;
                } else skip_value(c);
              });
              if (t_str && s_str && type == "Ready" && status == "True") n.flags |= IGPU_NODE_READY;  // k8s.ts:330
            } while (c.consume(','));
            c.expect(']');
          }
        } else skip_value(c);
      });
    } else skip_value(c);
  });
  if (gpu_cap) n.flags |= IGPU_NODE_HAS_GPU_CAPACITY;  // k8s.ts:146-148
  int64_t own = 0;
  for (int w = 0; w < 2; ++w) {
    if (cap[w].kind != Val::ABSENT) {
      // getNodeGpuCount (k8s.ts:175-176): `value && (parseInt(value) || 0)`
      (w ? n.cap_xe : n.cap_i915) = fits_i32(truthy(cap[w]) ? int_or_zero(cap[w]) : 0);
      // OverviewPage.tsx:98-99: parseInt(x ?? '0') without `|| 0`
      Val zero;
      zero.kind = Val::STR;
      zero.s = "0";
      if (std::isnan(parse_int(cap[w].kind == Val::NUL ? zero : cap[w]))) n.flags |= IGPU_NODE_CAP_NAN;
      const Val& a = (alloc[w].kind == Val::ABSENT || alloc[w].kind == Val::NUL) ? zero : alloc[w];
      if (std::isnan(parse_int(a))) n.flags |= IGPU_NODE_ALLOC_NAN;
      else (w ? n.alloc_xe : n.alloc_i915) = fits_i32(int_or_zero(a));
    }
    if (alloc[w].kind != Val::ABSENT) own += int_or_zero(alloc[w]);  // NodeDetailSection.tsx:72-76
  }
  n.alloc_own = fits_i32(own);
  return n;
}

// ---- pods -------------------------------------------------------------------------------------------------
struct LocalEntry {
  uint32_t res_local;  // index into the thread's resource table
  uint8_t flags;
  uint8_t container;   // index of the container in its list
  int64_t value;
};

// raw value of a resource key, kept until the container's other map has been read (`req === lim`, PodsPage.tsx:69)
struct RawVal {
  uint8_t kind;
  double num;
  std::string s;
};

// a === b for two JSON-borne values: strings by content, numbers by value, true / false / null by kind; containers
// (objects, arrays) are distinct objects after JSON.parse and never equal
bool strict_equal(const RawVal& a, const RawVal& b) {
  if (a.kind != b.kind) return false;
  switch (a.kind) {
    case Val::STR: return a.s == b.s;
    case Val::NUM: return a.num == b.num;
    case Val::NUL: case Val::TRUE_: case Val::FALSE_: return true;
    default: return false;
  }
}

struct PodChunk {  // one thread's contiguous range of pods
  std::vector<uint32_t> node;
  std::vector<uint8_t> phase;
  std::vector<uint32_t> n_entries;
  std::vector<LocalEntry> entries;
  std::vector<std::string> res_names;  // thread-local dictionary, first-seen order; ids 0..4 are the defaults
  std::unordered_map<std::string, uint32_t> res_ids;
  std::string error;
  int32_t error_status = 0;
};

uint8_t phase_code(const Val& v) {
  if (v.kind != Val::STR) return IGPU_PHASE_OTHER;
  const std::string& s = v.s;
  if (s == "Running") return IGPU_PHASE_RUNNING;
  if (s == "Pending") return IGPU_PHASE_PENDING;
  if (s == "Succeeded") return IGPU_PHASE_SUCCEEDED;
  if (s == "Failed") return IGPU_PHASE_FAILED;
  if (s == "Other") return IGPU_PHASE_OTHER;
  static const char* const proto[] = {"constructor", "__defineGetter__", "__defineSetter__", "hasOwnProperty", "__lookupGetter__",
                                      "__lookupSetter__", "isPrototypeOf", "propertyIsEnumerable", "toString", "valueOf", "__proto__",
                                      "toLocaleString"};
  for (const char* p : proto)
    if (s == p) return IGPU_PHASE_PROTO;  // `phase in podPhaseCounts` is true for inherited names
  return IGPU_PHASE_OTHER;
}

// One resources map.  A key written twice keeps its first position and its last value, like JSON.parse.
void parse_resource_map(Cur& c, std::string& scratch, PodChunk& out, uint8_t flags, std::vector<LocalEntry>& dst, std::vector<RawVal>& raw) {
  dst.clear();  // a map written twice: the last one is the property's value
  raw.clear();
  if (c.peek() != '{') return skip_value(c);
  for_each_member(c, scratch, [&](sv rk) {
    if (rk.substr(0, K_PREFIX.size()) != K_PREFIX) return skip_value(c);
    std::string key(rk);
    Val v = parse_val(c, scratch);
    auto it = out.res_ids.find(key);
    uint32_t id;
    if (it == out.res_ids.end()) {
      id = (uint32_t)out.res_names.size();
      out.res_ids.emplace(key, id);
      out.res_names.push_back(key);
    } else id = it->second;
    const LocalEntry le{id, (uint8_t)(flags | (truthy(v) ? IGPU_ENTRY_TRUTHY : 0)), 0, int_or_zero(v)};
    RawVal rv{(uint8_t)v.kind, v.num, std::move(v.s)};
    for (size_t i = 0; i < dst.size(); ++i)
      if (dst[i].res_local == id) {
        dst[i] = le;
        raw[i] = std::move(rv);
        return;
      }
    dst.push_back(le);
    raw.push_back(std::move(rv));
  });
}

void parse_containers(Cur& c, std::string& scratch, PodChunk& out, uint8_t init_flag, std::vector<LocalEntry>& dst) {
  dst.clear();  // the key written twice: the last list is the property's value
  if (c.peek() != '[') return skip_value(c);
  ++c.p;
  if (c.consume(']')) return;
  std::vector<LocalEntry> req, lim;
  std::vector<RawVal> req_raw, lim_raw;
  size_t ordinal = 0;
  do {
    req.clear(), lim.clear(), req_raw.clear(), lim_raw.clear();
    if (c.peek() == '{') {
      for_each_member(c, scratch, [&](sv ck) {
        if (ck != "resources") return skip_value(c);
        req.clear(), lim.clear(), req_raw.clear(), lim_raw.clear();
        if (c.peek() != '{') return skip_value(c);
        for_each_member(c, scratch, [&](sv rk) {
          if (rk == "requests") parse_resource_map(c, scratch, out, init_flag, req, req_raw);
          else if (rk == "limits") parse_resource_map(c, scratch, out, (uint8_t)(init_flag | IGPU_ENTRY_LIMIT), lim, lim_raw);
          else skip_value(c);
        });
      });
    } else skip_value(c);
    if (!req.empty() || !lim.empty()) {
      if (ordinal >= IGPU_MAX_CONTAINERS) throw RangeError{"a pod lists more than " + std::to_string(IGPU_MAX_CONTAINERS) + " containers"};
      for (size_t i = 0; i < req.size(); ++i)
        for (size_t j = 0; j < lim.size(); ++j)
          if (req[i].res_local == lim[j].res_local && strict_equal(req_raw[i], lim_raw[j])) req[i].flags |= IGPU_ENTRY_SAME, lim[j].flags |= IGPU_ENTRY_SAME;
      for (auto& e : req) e.container = (uint8_t)ordinal;
      for (auto& e : lim) e.container = (uint8_t)ordinal;
      // the reference reads requests, then limits, whatever their order in the text (k8s.ts:258-260)
      dst.insert(dst.end(), req.begin(), req.end());
      dst.insert(dst.end(), lim.begin(), lim.end());
    }
    ++ordinal;
  } while (c.consume(','));
  c.expect(']');
}

void parse_pods(const Span* items, size_t lo, size_t hi, const char* base,
                const std::unordered_map<std::string, uint32_t>& node_index, PodChunk& out) {
  for (int i = 0; i < 5; ++i) out.res_ids.emplace(DEFAULT_RES[i], (uint32_t)i), out.res_names.push_back(DEFAULT_RES[i]);
  std::string scratch;
  std::vector<LocalEntry> main_e, init_e;
  try {
    for (size_t i = lo; i < hi; ++i) {
      Cur c{items[i].b, items[i].e, base};
      uint32_t node = IGPU_NODE_NONE;
      uint8_t phase = IGPU_PHASE_OTHER;
      main_e.clear();
      init_e.clear();
      if (c.peek() == '{') {
        for_each_member(c, scratch, [&](sv key) {
          if (key == "spec") {
            node = IGPU_NODE_NONE, main_e.clear(), init_e.clear();  // a key written twice: JSON.parse keeps the last value
            for_each_member(c, scratch, [&](sv k2) {
              if (k2 == "nodeName") {
                Val v = parse_val(c, scratch);
                node = IGPU_NODE_NONE;
                if (v.kind == Val::STR && !v.s.empty()) {
                  auto it = node_index.find(v.s);
                  node = it == node_index.end() ? IGPU_NODE_NONE : it->second;
                }
              } else if (k2 == "containers") parse_containers(c, scratch, out, 0, main_e);
              else if (k2 == "initContainers") parse_containers(c, scratch, out, IGPU_ENTRY_INIT, init_e);
              else skip_value(c);
            });
          } else if (key == "status") {
            phase = IGPU_PHASE_OTHER;
            for_each_member(c, scratch, [&](sv k2) {
              if (k2 == "phase") phase = phase_code(parse_val(c, scratch));
              else skip_value(c);
            });
          } else skip_value(c);
        });
      }
      out.node.push_back(node);
      out.phase.push_back(phase);
      out.n_entries.push_back((uint32_t)(main_e.size() + init_e.size()));
      // containers first, then initContainers (k8s.ts:257, :291), whatever their order in the text
      out.entries.insert(out.entries.end(), main_e.begin(), main_e.end());
      out.entries.insert(out.entries.end(), init_e.begin(), init_e.end());
    }
  } catch (const ParseError& e) {
    char buf[256];
    snprintf(buf, sizeof buf, "pods JSON: %s at byte %zu", e.msg.c_str(), e.pos);
    out.error = buf;
    out.error_status = IGPU_ERR_ARG;
  } catch (const RangeError& e) {
    out.error = e.msg;
    out.error_status = IGPU_ERR_RANGE;
  }
}

}  // namespace

struct igpu_marshal {
  std::vector<uint8_t> node_flags;
  std::vector<int32_t> cap_i915, cap_xe, alloc_i915, alloc_xe, alloc_own;
  std::vector<uint32_t> pod_node, pod_off;
  std::vector<uint8_t> pod_phase;
  std::vector<igpu_entry> entries;
  std::vector<std::string> res_names;
  std::vector<uint32_t> pod_order;  // IGPU_OPT_GROUP_PODS_BY_NODE: JSON position of the pod in each column position
  uint32_t options = 0;
};

extern "C" {

int32_t igpu_marshal_fleet_json(const char* nodes_json, uint64_t nodes_len, const char* pods_json, uint64_t pods_len, uint32_t options,
                                int32_t threads, igpu_marshal** out) {
  if (!out || (!nodes_json && nodes_len) || (!pods_json && pods_len)) return igpu_fail(nullptr, IGPU_ERR_ARG, "igpu_marshal_fleet_json: NULL argument");
  if (!host_cpu_ok()) return igpu_fail(nullptr, IGPU_ERR_NODEVICE, "igpu_marshal_fleet_json: this build needs a host CPU with AVX2 and PCLMULQDQ");
  *out = nullptr;
  if (threads < 1) threads = (int32_t)std::max(1u, std::thread::hardware_concurrency());
  igpu_marshal* m = new igpu_marshal();
  m->options = options & ~IGPU_OPT_GROUP_PODS_BY_NODE;  // a marshal-only flag: the fleet columns never carry it
  const bool timing = getenv("IGPU_MARSHAL_TIMING") != nullptr;
  auto t_last = std::chrono::steady_clock::now();
  auto lap = [&](const char* what) {
    if (!timing) return;
    const auto now = std::chrono::steady_clock::now();
    fprintf(stderr, "[marshal] %-18s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
    t_last = now;
  };
  try {
    // nodes (parallel over items; the name index is built afterwards in item order: first occurrence wins)
    const std::vector<Span> nitems = split_items(nodes_json, nodes_len);
    const size_t N = nitems.size();
    std::vector<NodeRow> rows(N);
    std::vector<std::string> errs((size_t)threads);
    std::vector<int32_t> err_status((size_t)threads, 0);
    {
      std::vector<std::thread> pool;
      for (int t = 0; t < threads; ++t)
        pool.emplace_back([&, t] {
          try {
            for (size_t i = N * t / threads; i < N * (t + 1) / threads; ++i) rows[i] = parse_node(nitems[i], nodes_json);
          } catch (const ParseError& e) {
            errs[t] = "nodes JSON: " + e.msg + " at byte " + std::to_string(e.pos);
            err_status[t] = IGPU_ERR_ARG;
          } catch (const RangeError& e) {
            errs[t] = e.msg;
            err_status[t] = IGPU_ERR_RANGE;
          }
        });
      for (auto& th : pool) th.join();
    }
    for (int t = 0; t < threads; ++t)
      if (err_status[t]) {
        const int32_t st = err_status[t];
        const std::string msg = errs[t];
        delete m;
        return igpu_fail(nullptr, st, "%s", msg.c_str());
      }
    std::unordered_map<std::string, uint32_t> node_index;
    node_index.reserve(N * 2);
    m->node_flags.resize(N);
    m->cap_i915.resize(N), m->cap_xe.resize(N), m->alloc_i915.resize(N), m->alloc_xe.resize(N), m->alloc_own.resize(N);
    for (size_t i = 0; i < N; ++i) {
      const NodeRow& r = rows[i];
      m->node_flags[i] = r.flags;
      m->cap_i915[i] = r.cap_i915, m->cap_xe[i] = r.cap_xe, m->alloc_i915[i] = r.alloc_i915, m->alloc_xe[i] = r.alloc_xe,
      m->alloc_own[i] = r.alloc_own;
      if (r.has_name && !r.name.empty()) node_index.emplace(r.name, (uint32_t)i);  // emplace keeps the first
    }
    // pods
    lap("nodes");
    // Splitting the pod list into items is one sequential scan of the text; parsing the items is the parallel part.
    // The two overlap: the calling thread scans and publishes item spans in blocks of POD_BLOCK, the workers claim
    // blocks as they complete.  A block is a contiguous, ordered range of pods, so visiting the blocks in order below
    // reproduces the sequential result whatever the interleaving was.
    constexpr size_t POD_BLOCK = 2048;
    // Scalar elements are items too ("0," is two bytes, and an item can never take less than one), so the tables are
    // sized for one item per byte of text: 16 bytes of table per 2 KB of JSON.  The sink checks the bound all the same.
    const size_t max_blocks = (size_t)pods_len / POD_BLOCK + 2;
    std::vector<std::unique_ptr<Span[]>> span_blocks(max_blocks);
    std::vector<std::unique_ptr<PodChunk>> block_chunks(max_blocks);
    std::atomic<size_t> published{0}, next_block{0};
    std::atomic<bool> split_done{false};
    auto worker = [&] {
      for (;;) {
        const size_t b = next_block.fetch_add(1, std::memory_order_relaxed);
        size_t have;
        for (;;) {  // wait until block b is complete, or the scan has ended
          have = published.load(std::memory_order_acquire);
          if (have >= (b + 1) * POD_BLOCK || split_done.load(std::memory_order_acquire)) break;
          std::this_thread::yield();
        }
        have = published.load(std::memory_order_acquire);
        if (have <= b * POD_BLOCK) return;
        const size_t cnt = std::min(POD_BLOCK, have - b * POD_BLOCK);
        block_chunks[b].reset(new PodChunk());
        parse_pods(span_blocks[b].get(), 0, cnt, pods_json, node_index, *block_chunks[b]);
      }
    };
    std::vector<std::thread> pool;
    for (int t = 0; t < threads - 1; ++t) pool.emplace_back(worker);  // threads == 1: the caller scans, then parses
    size_t P = 0;
    std::string split_error;
    size_t split_error_pos = 0;
    try {
      split_items_to(pods_json, pods_len, [&](Span sp) {
        const size_t b = P / POD_BLOCK, k = P % POD_BLOCK;
        if (b >= max_blocks) throw ParseError{"more list items than bytes of text", (size_t)pods_len};
        if (k == 0) span_blocks[b].reset(new Span[POD_BLOCK]);
        span_blocks[b][k] = sp;
        ++P;
        if (k + 1 == POD_BLOCK) published.store(P, std::memory_order_release);
      });
    } catch (const ParseError& e) {
      split_error = e.msg;
      split_error_pos = e.pos;
    }
    published.store(P, std::memory_order_release);
    split_done.store(true, std::memory_order_release);
    lap("split pods");
    worker();  // the scanning thread helps with what is left
    for (auto& th : pool) th.join();
    if (!split_error.empty()) throw ParseError{split_error, split_error_pos};
    const size_t n_chunks = (P + POD_BLOCK - 1) / POD_BLOCK;
    std::vector<PodChunk> chunks(n_chunks);
    for (size_t b = 0; b < n_chunks; ++b) chunks[b] = std::move(*block_chunks[b]);
    lap("parse pods");
    for (auto& ch : chunks)
      if (ch.error_status) {
        const int32_t st = ch.error_status;
        const std::string msg = ch.error;
        delete m;
        return igpu_fail(nullptr, st, "%s", msg.c_str());
      }
    // global resource dictionary: block ranges are contiguous and ordered, so visiting the blocks in
    // order and their new names in local first-seen order reproduces the sequential first-seen order
    for (int i = 0; i < 5; ++i) m->res_names.push_back(DEFAULT_RES[i]);
    std::unordered_map<std::string, uint32_t> gids;
    for (uint32_t i = 0; i < 5; ++i) gids.emplace(DEFAULT_RES[i], i);
    std::vector<std::vector<uint32_t>> remap(chunks.size());
    size_t E = 0;
    for (size_t t = 0; t < chunks.size(); ++t) {
      remap[t].resize(chunks[t].res_names.size());
      // order in which this thread FIRST USED each local name (not the order of its table, which starts with the defaults)
      std::vector<uint8_t> seen(chunks[t].res_names.size(), 0);
      for (const LocalEntry& e : chunks[t].entries)
        if (!seen[e.res_local]) {
          seen[e.res_local] = 1;
          const std::string& name = chunks[t].res_names[e.res_local];
          auto it = gids.find(name);
          if (it == gids.end()) {
            if (m->res_names.size() >= IGPU_MAX_RESOURCES) {
              delete m;
              return igpu_fail(nullptr, IGPU_ERR_RANGE, "more than %d distinct gpu.intel.com/ resource names", IGPU_MAX_RESOURCES);
            }
            it = gids.emplace(name, (uint32_t)m->res_names.size()).first;
            m->res_names.push_back(name);
          }
          remap[t][e.res_local] = it->second;
        }
      E += chunks[t].entries.size();
    }
    if (E >= 0xFFFFFFFFull) {
      delete m;
      return igpu_fail(nullptr, IGPU_ERR_RANGE, "too many resource entries for the uint32 offsets");
    }
    m->pod_node.reserve(P), m->pod_phase.reserve(P), m->pod_off.reserve(P + 1), m->entries.reserve(E);
    m->pod_off.push_back(0);
    uint32_t pod = 0;
    unsigned __int128 abs_sum = 0;
    for (size_t t = 0; t < chunks.size(); ++t) {
      const PodChunk& ch = chunks[t];
      size_t e = 0;
      for (size_t i = 0; i < ch.node.size(); ++i, ++pod) {
        m->pod_node.push_back(ch.node[i]);
        m->pod_phase.push_back(ch.phase[i]);
        for (uint32_t k = 0; k < ch.n_entries[i]; ++k, ++e) {
          const LocalEntry& le = ch.entries[e];
          igpu_entry ge;
          ge.pod = pod;
          ge.res = (uint16_t)remap[t][le.res_local];
          ge.flags = le.flags;
          ge.container = le.container;
          ge.value = le.value;
          abs_sum += (unsigned __int128)(le.value < 0 ? -le.value : le.value);
          m->entries.push_back(ge);
        }
        m->pod_off.push_back((uint32_t)m->entries.size());
      }
    }
    lap("merge");
    if (options & IGPU_OPT_GROUP_PODS_BY_NODE) {
      // stable counting sort of the pods on their node id (unscheduled pods, IGPU_NODE_NONE, form the last bucket),
      // then the CSR entry stream is rebuilt in the new order
      std::vector<uint32_t> start(N + 2, 0);
      for (size_t i = 0; i < P; ++i) ++start[std::min<size_t>(m->pod_node[i], N) + 1];
      for (size_t b = 0; b <= N; ++b) start[b + 1] += start[b];
      m->pod_order.resize(P);
      for (size_t i = 0; i < P; ++i) m->pod_order[start[std::min<size_t>(m->pod_node[i], N)]++] = (uint32_t)i;
      std::vector<uint32_t> node2(P), off2(P + 1);
      std::vector<uint8_t> phase2(P);
      std::vector<igpu_entry> ent2(m->entries.size());
      uint32_t w = 0;
      off2[0] = 0;
      for (size_t i = 0; i < P; ++i) {
        const uint32_t src = m->pod_order[i];
        node2[i] = m->pod_node[src];
        phase2[i] = m->pod_phase[src];
        for (uint32_t e = m->pod_off[src]; e < m->pod_off[src + 1]; ++e) {
          ent2[w] = m->entries[e];
          ent2[w].pod = (uint32_t)i;
          ++w;
        }
        off2[i + 1] = w;
      }
      m->pod_node.swap(node2), m->pod_phase.swap(phase2), m->pod_off.swap(off2), m->entries.swap(ent2);
      lap("group by node");
    }
    if (abs_sum >= ((unsigned __int128)1 << 53)) {
      delete m;
      return igpu_fail(nullptr, IGPU_ERR_RANGE, "sum of |entry.value| reaches 2^53: outside the exact-integer domain of the reference");
    }
  } catch (const ParseError& e) {
    delete m;
    return igpu_fail(nullptr, IGPU_ERR_ARG, "JSON: %s at byte %zu", e.msg.c_str(), e.pos);
  } catch (const std::bad_alloc&) {
    delete m;
    return igpu_fail(nullptr, IGPU_ERR_OOM, "out of host memory while marshalling");
  }
  *out = m;
  return IGPU_OK;
}

int32_t igpu_marshal_columns(const igpu_marshal* m, igpu_fleet_cols* c) {
  if (!m || !c) return IGPU_ERR_ARG;
  c->n_nodes = (uint32_t)m->node_flags.size();
  c->n_pods = (uint32_t)m->pod_node.size();
  c->n_entries = m->entries.size();
  c->n_resources = (uint32_t)std::max<size_t>(4, m->res_names.size());
  c->options = m->options;
  c->node_flags = m->node_flags.data();
  c->node_cap_i915 = m->cap_i915.data();
  c->node_cap_xe = m->cap_xe.data();
  c->node_alloc_i915 = m->alloc_i915.data();
  c->node_alloc_xe = m->alloc_xe.data();
  c->node_alloc_own = m->alloc_own.data();
  c->pod_node = m->pod_node.data();
  c->pod_phase = m->pod_phase.data();
  c->pod_entry_off = m->pod_off.data();
  c->entries = m->entries.data();
  return IGPU_OK;
}

const uint32_t* igpu_marshal_pod_order(const igpu_marshal* m) { return m && !m->pod_order.empty() ? m->pod_order.data() : nullptr; }

const char* igpu_marshal_resource_name(const igpu_marshal* m, uint32_t id) {
  return (m && id < m->res_names.size()) ? m->res_names[id].c_str() : nullptr;
}

int32_t igpu_marshal_free(igpu_marshal* m) {
  delete m;
  return IGPU_OK;
}

}  // extern "C"

// ======================================================================================================
// N3: Prometheus `query_range` matrix response -> float64 [series][time] slab.
// Replaces, for the range view, the JSON.parse + `parseFloat(r.value[1])` loops of metrics.ts:67-75,130-141
// (which handle instant vectors).  Wire format (Prometheus HTTP API v1):
//   {"status":"success","data":{"resultType":"matrix","result":[{"metric":{..labels..},"values":[[<unix s>,"<float>"],..]},..]}}
// `status !== 'success'` yields no series, like queryPrometheus (metrics.ts:73).  A bare `[ {...}, ... ]`
// result array is accepted too.  Samples land on the grid t = start_ms + i * step_ms; a sample whose timestamp
// is off the grid or outside [0, n_time) is counted in `dropped` and skipped; a grid step without a sample is NaN.
namespace {

// JSON number (unix seconds, at most millisecond digits as Prometheus writes them) -> integer milliseconds, exactly
bool parse_ts_ms(Cur& c, int64_t& out_ms) {
  c.ws();
  const char* s = c.p;
  bool neg = false;
  if (c.p < c.end && *c.p == '-') neg = true, ++c.p;
  int64_t sec = 0;
  const char* d = c.p;
  bool huge = false;  // more seconds than fit the grid arithmetic below: such a sample is simply off the grid
  while (c.p < c.end && *c.p >= '0' && *c.p <= '9') {
    if (sec > (int64_t)1e14) huge = true;
    else sec = sec * 10 + (*c.p - '0');
    ++c.p;
  }
  if (c.p == d) c.fail("bad timestamp");
  int64_t ms = 0;
  int digits = 0;
  bool inexact = false;
  if (c.p < c.end && *c.p == '.') {
    ++c.p;
    while (c.p < c.end && *c.p >= '0' && *c.p <= '9') {
      if (digits < 3) ms = ms * 10 + (*c.p - '0'), ++digits;
      else if (*c.p != '0') inexact = true;
      ++c.p;
    }
  }
  while (digits++ < 3) ms *= 10;
  if (c.p < c.end && (*c.p == 'e' || *c.p == 'E')) {  // exponent form: not something Prometheus writes; round via strtod
    while (c.p < c.end && *c.p != ',' && *c.p != ']' && *c.p != ' ') ++c.p;
    const std::string t(s, c.p - s);
    const double x = std::strtod(t.c_str(), nullptr) * 1000.0;
    if (!(std::fabs(x) < 1e17)) {
      out_ms = 0;
      return false;
    }
    out_ms = (int64_t)std::llround(x);
    return true;
  }
  out_ms = (neg ? -1 : 1) * (sec * 1000 + ms);  // |sec| <= 1e15: no overflow
  return !inexact && !huge;
}

// Sample value text -> double with the semantics of JS parseFloat, which is what the reference applies to a sample
// (`parseFloat(r.value[1])`, metrics.ts:130-137): leading white space, then the longest prefix that is a decimal literal
// ([+-] digits [. digits] [e[+-]digits], or [+-]Infinity); anything else — "NaN", "+Inf", "0x10" past its "0" — follows
// from that rule.  The prefix is converted exactly: Clinger's fast path when it applies, strtod otherwise.
double parse_sample_value(sv t) {
  static const double P10[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
  const char* p = t.data();
  const char* end = p + t.size();
  while (p < end) {
    const size_t n = js_ws_len(p, end);
    if (!n) break;
    p += n;
  }
  const char* start = p;
  bool neg = false;
  if (p < end && (*p == '-' || *p == '+')) neg = *p++ == '-';
  if ((size_t)(end - p) >= 8 && memcmp(p, "Infinity", 8) == 0) return neg ? -INFINITY : INFINITY;
  uint64_t mant = 0;
  int nd = 0, exp10 = 0, digits = 0;
  bool small = true;  // mantissa still exact in 64 bits / 15 significant digits
  while (p < end && *p >= '0' && *p <= '9') {
    if (mant || *p != '0') ++nd;
    if (nd <= 15) mant = mant * 10 + (uint64_t)(*p - '0'); else small = false;
    ++p, ++digits;
  }
  if (p < end && *p == '.') {
    const char* q = p + 1;
    while (q < end && *q >= '0' && *q <= '9') {
      if (mant || *q != '0') ++nd;
      if (nd <= 15) mant = mant * 10 + (uint64_t)(*q - '0'), --exp10; else small = false;
      ++q, ++digits;
    }
    if (digits) p = q;  // "5." and ".5" are literals, a lone "." is not
  }
  if (!digits) return NAN;
  if (p < end && (*p == 'e' || *p == 'E')) {  // the exponent belongs to the literal only if digits follow
    const char* q = p + 1;
    bool eneg = false;
    if (q < end && (*q == '-' || *q == '+')) eneg = *q++ == '-';
    const char* d = q;
    int e = 0;
    while (q < end && *q >= '0' && *q <= '9') {
      if (e < 100000) e = e * 10 + (*q - '0');
      ++q;
    }
    if (q > d) {
      exp10 += eneg ? -e : e;
      p = q;
    }
  }
  if (small && exp10 >= -22 && exp10 <= 22) {
    double v = (double)mant;
    v = exp10 < 0 ? v / P10[-exp10] : v * P10[exp10];
    return neg ? -v : v;
  }
  const std::string z(start, p - start);  // only [+-0-9.eE]: strtod reads exactly this literal
  return std::strtod(z.c_str(), nullptr);
}

}  // namespace

struct igpu_matrix {
  uint32_t n_series = 0, n_time = 0, n_labels = 0;
  std::vector<double> values;        // [n_series][n_time]
  std::vector<std::string> labels;   // [n_series][n_labels]
  std::vector<uint8_t> has_label;    // [n_series][n_labels]
  uint64_t dropped = 0;              // samples off the grid or outside the slab
};

extern "C" {

int32_t igpu_marshal_matrix_json(const char* json, uint64_t len, int64_t start_ms, uint32_t step_ms, uint32_t n_time,
                                 const char* const* label_names, uint32_t n_labels, int32_t threads, igpu_matrix** out) {
  if (!out || (!json && len) || (!label_names && n_labels)) return igpu_fail(nullptr, IGPU_ERR_ARG, "igpu_marshal_matrix_json: NULL argument");
  *out = nullptr;
  if (!step_ms || !n_time) return igpu_fail(nullptr, IGPU_ERR_ARG, "step_ms and n_time must be positive");
  if (!host_cpu_ok()) return igpu_fail(nullptr, IGPU_ERR_NODEVICE, "igpu_marshal_matrix_json: this build needs a host CPU with AVX2 and PCLMULQDQ");
  try {
    std::vector<Span> items;
    std::string scratch;
    Cur c{json, json + len, json};
    bool success = true;
    if (len && c.peek() == '{') {  // the API envelope: status, data.result
      ++c.p;
      bool found = false;
      if (!c.consume('}')) {
        do {
          const sv key = parse_string(c, scratch);
          const std::string k(key);
          c.expect(':');
          if (k == "status") {
            const Val v = parse_val(c, scratch);
            success = v.kind == Val::STR && v.s == "success";
          } else if (k == "data" && c.peek() == '{') {
            ++c.p;
            if (!c.consume('}')) {
              do {
                const std::string dk(parse_string(c, scratch));
                c.expect(':');
                if (dk == "result" && c.peek() == '[' && !found) {
                  found = true;
                  const char* b = c.p;
                  bool fast = false;
#if defined(__AVX2__) && defined(__PCLMUL__)
                  {  // one pass: the block splitter also tells where the array ends
                    Cur a = c;
                    ++a.p;
                    const char first = a.peek();
                    const char* close = nullptr;
                    if ((first == '{' || first == '[') && split_array_fast(a.p, a.end, items, &close)) {
                      c.p = close + 1;
                      fast = true;
                    } else {
                      items.clear();
                    }
                  }
#endif
                  if (!fast) {
                    skip_value(c);
                    items = split_items(b, (size_t)(c.p - b));
                  }
                } else {
                  skip_value(c);
                }
              } while (c.consume(','));
              c.expect('}');
            }
          } else {
            skip_value(c);
          }
        } while (c.consume(','));
      }
      if (!success) items.clear();
    } else if (len) {
      items = split_items(json, len);
    }
    igpu_matrix* m = new igpu_matrix();
    m->n_series = (uint32_t)items.size();
    m->n_time = n_time;
    m->n_labels = n_labels;
    m->values.assign((size_t)m->n_series * n_time, NAN);
    m->labels.resize((size_t)m->n_series * n_labels);
    m->has_label.assign((size_t)m->n_series * n_labels, 0);
    std::vector<std::string> names(n_labels);
    for (uint32_t i = 0; i < n_labels; ++i) names[i] = label_names[i] ? label_names[i] : "";
    int nt = threads > 0 ? threads : (int)std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 16u);
    nt = (int)std::min<size_t>((size_t)nt, std::max<size_t>(1, items.size() / 8));
    std::vector<uint64_t> dropped((size_t)nt, 0);
    std::vector<std::string> errors((size_t)nt);
    auto work = [&](int tid) {
      std::string scr, scr2;
      try {
        for (size_t si = (size_t)tid; si < items.size(); si += (size_t)nt) {
          Cur q{items[si].b, items[si].e, json};
          double* row = m->values.data() + si * n_time;
          q.expect('{');
          if (q.consume('}')) continue;
          do {
            const std::string key(parse_string(q, scr));
            q.expect(':');
            if (key == "metric" && q.peek() == '{') {
              ++q.p;
              if (!q.consume('}')) {
                do {
                  const std::string lk(parse_string(q, scr));
                  q.expect(':');
                  const Val v = parse_val(q, scr2);
                  for (uint32_t i = 0; i < n_labels; ++i)
                    if (lk == names[i] && v.kind == Val::STR) {
                      m->labels[si * n_labels + i] = v.s;
                      m->has_label[si * n_labels + i] = 1;
                    }
                } while (q.consume(','));
                q.expect('}');
              }
            } else if (key == "values" && q.peek() == '[') {
              ++q.p;
              if (!q.consume(']')) {
                do {
                  q.expect('[');
                  int64_t ts = 0;
                  const bool exact = parse_ts_ms(q, ts);
                  q.expect(',');
                  double v;
                  if (q.peek() == '"') {
                    v = parse_sample_value(parse_string(q, scr));
                  } else {  // not what Prometheus writes: a bare number is taken as it is, anything else is no sample
                    const Val bare = parse_val(q, scr);
                    v = bare.kind == Val::NUM ? bare.num : NAN;
                  }
                  q.expect(']');
                  const int64_t rel = exact ? (int64_t)((uint64_t)ts - (uint64_t)start_ms) : -1;  // wraps instead of overflowing
                  if (exact && rel >= 0 && rel % step_ms == 0 && rel / step_ms < (int64_t)n_time) row[rel / step_ms] = v;
                  else ++dropped[(size_t)tid];
                } while (q.consume(','));
                q.expect(']');
              }
            } else {
              skip_value(q);
            }
          } while (q.consume(','));
          q.expect('}');
        }
      } catch (const ParseError& e) {
        errors[(size_t)tid] = e.msg + " at byte " + std::to_string(e.pos);
      }
    };
    if (nt <= 1) {
      work(0);
    } else {
      std::vector<std::thread> pool;
      for (int t = 0; t < nt; ++t) pool.emplace_back(work, t);
      for (auto& t : pool) t.join();
    }
    for (const auto& e : errors)
      if (!e.empty()) {
        delete m;
        return igpu_fail(nullptr, IGPU_ERR_ARG, "matrix JSON: %s", e.c_str());
      }
    for (uint64_t d : dropped) m->dropped += d;
    *out = m;
    return IGPU_OK;
  } catch (const ParseError& e) {
    return igpu_fail(nullptr, IGPU_ERR_ARG, "matrix JSON: %s at byte %zu", e.msg.c_str(), e.pos);
  } catch (const std::bad_alloc&) {
    return igpu_fail(nullptr, IGPU_ERR_OOM, "matrix JSON: out of host memory");
  }
}

uint32_t igpu_matrix_n_series(const igpu_matrix* m) { return m ? m->n_series : 0; }
const double* igpu_matrix_values(const igpu_matrix* m) { return m ? m->values.data() : nullptr; }
const char* igpu_matrix_label(const igpu_matrix* m, uint32_t series, uint32_t label) {
  if (!m || series >= m->n_series || label >= m->n_labels) return nullptr;
  const size_t i = (size_t)series * m->n_labels + label;
  return m->has_label[i] ? m->labels[i].c_str() : nullptr;
}
uint64_t igpu_matrix_dropped(const igpu_matrix* m) { return m ? m->dropped : 0; }
int32_t igpu_matrix_free(igpu_matrix* m) {
  delete m;
  return IGPU_OK;
}

}  // extern "C"
