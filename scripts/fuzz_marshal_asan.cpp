// ASAN + UBSAN fuzz driver for the native marshals (csrc/marshal.cpp): mutated JSON in exact-size heap buffers.
// Build and run (seed files are written by scripts/fuzz_marshal.py --dump DIR):
//   g++ -O1 -g -mavx2 -mpclmul -std=c++17 -fsanitize=address,undefined -pthread -Iinclude -o /tmp/fuzz \
//       headlamp_intel_gpu_plugin_b200/csrc/marshal.cpp scripts/fuzz_marshal_asan_stub.cpp scripts/fuzz_marshal_asan.cpp
//   /tmp/fuzz SEED ITERATIONS [DIR]
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include "igpu_rollup.h"
static uint64_t rs = 88172645463325252ull;
static uint64_t rnd() { rs ^= rs << 13; rs ^= rs >> 7; rs ^= rs << 17; return rs; }
static std::string slurp(const char* p) { std::ifstream f(p); std::stringstream ss; ss << f.rdbuf(); return ss.str(); }
static const char SPECIAL[] = "\"\\{}[],:0-9eE.+ \n\ttruefalsn\xff\xc2\xa0";
static std::string mutate(std::string b) {
  int n = 1 + rnd() % 4;
  for (int k = 0; k < n && !b.empty(); ++k) {
    size_t i = rnd() % b.size();
    switch (rnd() % 6) {
      case 0: b[i] = SPECIAL[rnd() % (sizeof SPECIAL - 1)]; break;
      case 1: b.erase(i, 1 + rnd() % 20); break;
      case 2: { std::string ins; for (int j = 0, m = 1 + rnd() % 6; j < m; ++j) ins.push_back(SPECIAL[rnd() % (sizeof SPECIAL - 1)]); b.insert(i, ins); break; }
      case 3: b.resize(i); break;
      case 4: { size_t back = 1 + rnd() % 40; size_t from = i > back ? i - back : 0; b.insert(i, b.substr(from, i - from)); break; }
      default: b[i] = (char)(rnd() & 0xFF); break;
    }
  }
  return b;
}
static char* heap(const std::string& s) { char* p = (char*)malloc(s.size() ? s.size() : 1); memcpy(p, s.data(), s.size()); return p; }
int main(int argc, char** argv) {
  rs ^= (uint64_t)atoll(argv[1]) * 0x9E3779B97F4A7C15ull;
  const long iters = atol(argv[2]);
  const std::string dir = argc > 3 ? argv[3] : "/tmp/asan";
  const std::string nj = slurp((dir + "/nodes.json").c_str()), pj = slurp((dir + "/pods.json").c_str()), mx = slurp((dir + "/matrix.json").c_str());
  const char* names[2] = {"chip", "instance"};
  long ok = 0, rej = 0;
  // long lists of the shortest possible items (scalars, empty containers): the block tables are sized from the text length
  for (const char* item : {"0", "[]", "{}", "\"\"", "1 "}) {
    std::string t = "[";
    for (int i = 0; i < 200000; ++i) { if (i) t += ','; t += item; }
    t += ']';
    for (int th = 1; th <= 3; th += 2) {
      char *a = heap("[]"), *b = heap(t); igpu_marshal* h = nullptr;
      const int32_t st = igpu_marshal_fleet_json(a, 2, b, t.size(), 1u, th, &h);
      if (st == 0) igpu_marshal_free(h);
      ok += st == 0; rej += st != 0;
      free(a); free(b);
    }
  }
  for (long it = 0; it < iters; ++it) {
    int32_t st;
    if (it % 3 == 0) {
      const std::string m = mutate(pj); char *a = heap(nj), *b = heap(m); igpu_marshal* h = nullptr;
      st = igpu_marshal_fleet_json(a, nj.size(), b, m.size(), 1u | (it % 2 ? 0x100u : 0u), 1 + it % 3, &h);
      if (st == 0) { igpu_fleet_cols c; igpu_marshal_columns(h, &c); igpu_marshal_free(h); }
      free(a); free(b);
    } else if (it % 3 == 1) {
      const std::string m = mutate(nj); char *a = heap(m), *b = heap(pj); igpu_marshal* h = nullptr;
      st = igpu_marshal_fleet_json(a, m.size(), b, pj.size(), 1u, 2, &h);
      if (st == 0) igpu_marshal_free(h);
      free(a); free(b);
    } else {
      const std::string m = mutate(mx); char* a = heap(m); igpu_matrix* h = nullptr;
      st = igpu_marshal_matrix_json(a, m.size(), 1700000000000ll, 15000, 30, names, 2, 1 + it % 2, &h);
      if (st == 0) { (void)igpu_matrix_label(h, 0, 0); igpu_matrix_free(h); }
      free(a);
    }
    ok += st == 0; rej += st != 0;
  }
  printf("asan fuzz: %ld ok, %ld rejected\n", ok, rej);
  return 0;
}
