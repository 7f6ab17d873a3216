// TEST INFRASTRUCTURE ONLY (oracle build). The reference's ThreeTierAggrStore<T> buffers updates per target rank and
// ships them by RPC (upcxx-utils/include/upcxx_utils/three_tier_aggr_store.hpp:914-960). In the single-process
// oracle every update is delivered at once, either to the registered update function (all shards in one table)
// or to an optional tap that records (target, element) so that a harness can emulate several ranks.
#pragma once
#include <functional>
#include <cstdint>
namespace upcxx_utils {
template <typename T>
class ThreeTierAggrStore {
  std::function<void(T)> update_func;

 public:
  static std::function<bool(int, const T &)> &tap() {
    static std::function<bool(int, const T &)> t;
    return t;
  }
  void set_size(const char *, int64_t, int = 0, bool = false) {}
  template <typename F>
  void set_update_func(F f) { update_func = f; }
  void update(int target, const T &elem) {
    auto &t = tap();
    if (t && t(target, elem)) return;  // tap consumed it
    update_func(elem);
  }
  void flush_updates() {}
  void clear() {}
};
template <typename T>
using FlatAggrStore = ThreeTierAggrStore<T>;
}  // namespace upcxx_utils
