// TEST INFRASTRUCTURE ONLY (oracle build). Stand-in for upcxx-utils' umbrella header: logging macros become
// no-ops (DIE aborts), the aggregating store delivers each update by a direct call. No arithmetic lives here.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>
#include <cctype>
#include <iostream>
#include <iomanip>
#include <sstream>
#include <fstream>
#include <string>
#include <vector>
#include <array>
#include <algorithm>
#include <limits>
#include <functional>
#include <upcxx/upcxx.hpp>

// the reference headers rely on these names having leaked into the global namespace (kmer_dht.hpp:63,120)
using upcxx::dist_object;
using upcxx::global_ptr;
using upcxx::rank_me;
using upcxx::rank_n;
using upcxx::barrier;
using upcxx::reduce_one;
using upcxx::reduce_all;
using upcxx::op_fast_add;
using upcxx::op_fast_max;
using upcxx::op_fast_min;
using upcxx::local_team;
using upcxx::world;
using upcxx::progress;
using upcxx::rpc;
using upcxx::future;

namespace upcxx_utils {
using std::string;

struct shim_mem {
  static double &free_mem() { static double m = 8e9; return m; }
};
inline double get_free_mem() { return shim_mem::free_mem(); }
inline string get_size_str(double) { return string(); }
template <typename A, typename B>
inline string perc_str(A, B) { return string(); }
inline bool get_rank_path(string &, int) { return true; }
inline string get_current_time() { return string(); }

template <typename... A>
[[noreturn]] inline void shim_die(const char *file, int line, A &&...a) {
  std::ostringstream os;
  (os << ... << a);
  fprintf(stderr, "DIE %s:%d: %s\n", file, line, os.str().c_str());
  abort();
}
}  // namespace upcxx_utils

#define KNORM ""
#define KLGREEN ""
#define KBLUE ""
#define SLOG(...) do { } while (0)
#define SLOG_VERBOSE(...) do { } while (0)
#define SOUT(...) do { } while (0)
#define WARN(...) do { } while (0)
#define SWARN(...) do { } while (0)
#define DBG(...) do { } while (0)
#define LOG(...) do { } while (0)
#define LOG_MEM(...) do { } while (0)
#define DIE(...) upcxx_utils::shim_die(__FILE__, __LINE__, __VA_ARGS__)
#define SDIE(...) upcxx_utils::shim_die(__FILE__, __LINE__, __VA_ARGS__)

#include "upcxx_utils/three_tier_aggr_store.hpp"
