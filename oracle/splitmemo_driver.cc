// TEST INFRASTRUCTURE (oracle/): C entry points over the reference's own SplitFactorizationMemo, compiled from the
// reference source lines that oracle/Makefile extracts (class: search_policy/utils.h:668-685; GetFactorizationSchemes,
// DfsEnumerate, GetFactors: search_policy/utils.cc:414-463).  Only tests/ load the resulting oracle/_ref/libsplitmemo.so.
#include "splitmemo_shim.h"

namespace tvm {
namespace auto_scheduler {
#include "_ref/splitmemo_class.inc"
#include "_ref/splitmemo_impl.inc"
}  // namespace auto_scheduler
}  // namespace tvm

using tvm::auto_scheduler::SplitFactorizationMemo;
static SplitFactorizationMemo g_memo;

extern "C" {
// sorted divisors of n (GetFactors); returns the count, writes at most cap of them
int ref_get_factors(int n, int* out, int cap) {
  const std::vector<int>& f = g_memo.GetFactors(n);
  for (int i = 0; i < (int)f.size() && i < cap; ++i) out[i] = f[i];
  return (int)f.size();
}
// every scheme of GetFactorizationSchemes(extent, n_lengths, max_innermost) as n_lengths ints each, in the
// reference's enumeration order; returns the number of schemes, writes at most cap_schemes of them
int64_t ref_factorization_schemes(int extent, int n_lengths, int max_innermost, int* out, int64_t cap_schemes) {
  const auto& s = g_memo.GetFactorizationSchemes(extent, n_lengths, max_innermost);
  int64_t w = 0;
  for (const auto& row : s) {
    if (w >= cap_schemes) break;
    for (int j = 0; j < n_lengths; ++j) out[w * n_lengths + j] = (int)row[j].as<tvm::IntImmNode>()->value;
    ++w;
  }
  return (int64_t)s.size();
}
// member[i] = 1 iff rows[i, :] is one of the enumerated schemes (lengths < 2^16)
int ref_scheme_members(int extent, int n_lengths, int max_innermost, const int* rows, int64_t n_rows, unsigned char* member) {
  if (n_lengths < 1 || n_lengths > 4) return -1;
  const auto& s = g_memo.GetFactorizationSchemes(extent, n_lengths, max_innermost);
  std::vector<uint64_t> keys;
  keys.reserve(s.size());
  for (const auto& row : s) {
    uint64_t k = 0;
    for (int j = 0; j < n_lengths; ++j) k = (k << 16) | (uint64_t)row[j].as<tvm::IntImmNode>()->value;
    keys.push_back(k);
  }
  std::sort(keys.begin(), keys.end());
  for (int64_t i = 0; i < n_rows; ++i) {
    uint64_t k = 0;
    bool ok = true;
    for (int j = 0; j < n_lengths; ++j) {
      const int v = rows[i * n_lengths + j];
      if (v < 0 || v >= 65536) ok = false;
      k = (k << 16) | (uint64_t)(uint16_t)v;
    }
    member[i] = ok && std::binary_search(keys.begin(), keys.end(), k);
  }
  return 0;
}
}
