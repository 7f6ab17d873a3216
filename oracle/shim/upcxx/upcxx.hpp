// TEST INFRASTRUCTURE ONLY (oracle build). Single-process stand-in for the UPC++ API surface that the
// reference's kcount CPU path touches (kmer.hpp, kmer_dht.{hpp,cpp}, kcount.hpp, kcount_cpu.cpp).
// It lets those reference sources compile UNMODIFIED from /root/reference; it carries no arithmetic.
// rank_n() is settable so that the owner-rank function (kmer_dht.cpp:192-196) can be evaluated for N shards.
#pragma once
#include <cstdint>
#include <cstddef>
#include <functional>
#include <memory>
#include <utility>
#include <type_traits>
#include <tuple>

#define UPCXX_SERIALIZED_FIELDS(...)

namespace upcxx {
using intrank_t = int;

struct shim_state {
  static int &nranks() { static int n = 1; return n; }
  static int &me() { static int r = 0; return r; }
};

inline intrank_t rank_me() { return shim_state::me(); }
inline intrank_t rank_n() { return shim_state::nranks(); }

struct team {
  bool local;
  intrank_t rank_n() const { return local ? 1 : upcxx::rank_n(); }
  intrank_t rank_me() const { return local ? 0 : upcxx::rank_me(); }
};
inline team &world() { static team t{false}; return t; }
inline team &local_team() { static team t{true}; return t; }

inline void barrier() {}
inline void barrier(const team &) {}
inline void progress() {}
inline void init() {}
inline void finalize() {}

template <typename... T>
struct future;
template <>
struct future<> {
  void wait() const {}
  bool ready() const { return true; }
  template <typename F>
  auto then(F &&f) const { f(); return future<>(); }
};
template <typename T>
struct future<T> {
  T v;
  T wait() const { return v; }
  T result() const { return v; }
  bool ready() const { return true; }
  template <typename F>
  auto then(F &&f) const {
    using R = decltype(f(v));
    if constexpr (std::is_void<R>::value) { f(v); return future<>(); }
    else return future<R>{f(v)};
  }
};
inline future<> make_future() { return future<>(); }
template <typename T>
future<T> make_future(T v) { return future<T>{v}; }

// fastq.hpp (pulled in by packed_reads.cpp) declares members of these types; nothing on the kcount path calls them
template <typename... T>
struct promise {
  std::tuple<T...> v;
  template <typename... A>
  void fulfill_result(A &&...a) { v = std::tuple<T...>(std::forward<A>(a)...); }
  future<T...> get_future() const {
    if constexpr (sizeof...(T) == 0) return future<>();
    else return future<T...>{std::get<0>(v)};
  }
};
template <typename... F>
future<> when_all(F &&...) { return future<>(); }

struct op_add_t {}; struct op_max_t {}; struct op_min_t {};
static constexpr op_add_t op_fast_add{};
static constexpr op_max_t op_fast_max{};
static constexpr op_min_t op_fast_min{};

template <typename T, typename OP>
future<T> reduce_one(T v, OP, intrank_t, const team & = world()) { return future<T>{v}; }
template <typename T, typename OP>
future<T> reduce_all(T v, OP, const team & = world()) { return future<T>{v}; }
template <typename T>
future<T> broadcast(T v, intrank_t, const team & = world()) { return future<T>{v}; }

template <typename T>
struct global_ptr {
  T *p = nullptr;
  global_ptr() {}
  global_ptr(std::nullptr_t) {}
  T *local() const { return p; }
  bool is_null() const { return p == nullptr; }
  explicit operator bool() const { return p != nullptr; }
  bool operator==(const global_ptr &o) const { return p == o.p; }
  bool operator!=(const global_ptr &o) const { return p != o.p; }
};

template <typename T>
class dist_object {
  std::shared_ptr<T> p;

 public:
  dist_object(T v) : p(std::make_shared<T>(std::move(v))) {}
  template <typename A, typename B, typename... R>
  dist_object(A &&a, B &&b, R &&...r) : p(new T(std::forward<A>(a), std::forward<B>(b), std::forward<R>(r)...)) {}
  T *operator->() { return p.get(); }
  const T *operator->() const { return p.get(); }
  T &operator*() { return *p; }
  const T &operator*() const { return *p; }
};

// rpc to any rank == direct call in the single-process oracle
template <typename F, typename... A>
auto rpc(intrank_t, F &&f, A &&...a) {
  using R = decltype(f(std::forward<A>(a)...));
  if constexpr (std::is_void<R>::value) { f(std::forward<A>(a)...); return future<>(); }
  else return future<R>{f(std::forward<A>(a)...)};
}
}  // namespace upcxx
