// Host build of canu_b200/csrc/oic_core.cuh for CPU unit tests of its scalar logic (not a product path).
//   g++ -O2 -std=c++17 -ffp-contract=off -DOIC_HOST_ONLY -shared -fPIC tests/core_host.cpp -o tests/libcore_host.so
#include <vector>
#include "../canu_b200/csrc/oic_core.cuh"
extern "C" {
// runs: n x (start, offset, len) in arrival order; out: n x (offset, len, start) in list order
int core_replay(const int32_t *runs, int n, int K, int32_t *out, int32_t *info /*consistent, ct, bgn, end*/) {
  std::vector<oic::Seed> in(n), o(n);
  std::vector<int32_t> scr(4 * (size_t)n + 4);
  for (int i = 0; i < n; i++) { in[i].start = runs[3 * i]; in[i].offset = runs[3 * i + 1]; in[i].len = runs[3 * i + 2]; in[i].str = 0; }
  oic::replay_add_match(in.data(), n, K, o.data(), scr.data(), info[0], info[1], info[2], info[3]);
  for (int i = 0; i < n; i++) { out[3 * i] = o[i].offset; out[3 * i + 1] = o[i].len; out[3 * i + 2] = o[i].start; }
  return 0;
}
void core_pack_record(uint32_t *w, const uint32_t *f) {
  oic::pack_record(w, f[0], f[1], f[2], f[3], f[4], f[5], f[6], f[7], f[8], f[9], f[10], f[11]);
}
}
