// TEST INFRASTRUCTURE (oracle/): stand-ins for the few TVM container types that the reference's
// SplitFactorizationMemo (src/auto_scheduler/search_policy/utils.{h,cc}) uses, so that the reference's OWN source
// lines -- extracted at build time by oracle/Makefile into oracle/_ref/, never committed -- compile without the TVM
// runtime.  Nothing here restates the algorithm: Array is a value-semantics vector (TVM's Array is copy-on-write, so
// `results_->push_back(tmp_stack_)` followed by `tmp_stack_.Set(...)` keeps the pushed copy intact in both),
// Integer holds an IntImmNode whose `value` the reference reads through `.as<IntImmNode>()`.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <tuple>
#include <unordered_map>
#include <vector>

namespace std {
template <>
struct hash<std::tuple<int, int, int>> {
  size_t operator()(const std::tuple<int, int, int>& k) const {
    return ((size_t)std::get<0>(k) * 1000003u) ^ ((size_t)std::get<1>(k) * 10007u) ^ (size_t)std::get<2>(k);
  }
};
}  // namespace std

namespace tvm {
struct IntImmNode { int64_t value; };
class Integer {
 public:
  Integer() : node_{0} {}
  Integer(int v) : node_{v} {}  // NOLINT: the reference writes Integer(f)
  template <class N> const N* as() const { return &node_; }
 private:
  IntImmNode node_;
};
template <class T>
class Array : public std::vector<T> {
 public:
  using std::vector<T>::vector;
  void Set(int i, T v) { (*this)[i] = v; }
};
}  // namespace tvm
