// env.h — B2R_* tuning variables without a getenv on every launch.
//
// getenv walks the whole environment: ~0.13 us per lookup with 130 variables set, ~0.25 us with 280 (a torchrun worker).
// The single-GPU launch path looked up seven names and a collective up to fourteen, i.e. 1-3 us of host time per call — a
// third of a 9 us reduction, and more than the whole GPU side of a small LL collective. Every call site therefore keeps its
// own cached value, tagged with a process-wide generation; b2r_refresh_env() (also run by the introspection calls
// b2r_plan_local and b2r_select_algo, which must see the environment as it is now) bumps the generation, and the next
// use of each variable reads it again. A program that changes B2R_* variables at run time calls b2r_refresh_env() afterwards.
#pragma once
#include <atomic>
#include <cstdlib>

namespace b2r {

inline std::atomic<unsigned>& env_generation() {
  static std::atomic<unsigned> gen{1};
  return gen;
}
inline void env_refresh() { env_generation().fetch_add(1, std::memory_order_acq_rel); }

struct EnvSlot {
  std::atomic<unsigned> gen{0};  // generation the cached value belongs to (0 = never read)
  std::atomic<int> value{0};
  std::atomic<bool> set{false};
};

// Concurrent first uses of one slot read the same environment and store the same values, so the race is benign.
inline int env_int_cached(EnvSlot& s, const char* name, int dflt) {
  const unsigned g = env_generation().load(std::memory_order_acquire);
  if (s.gen.load(std::memory_order_acquire) != g) {
    const char* v = getenv(name);
    const bool is_set = v && *v;
    s.value.store(is_set ? atoi(v) : 0, std::memory_order_relaxed);
    s.set.store(is_set, std::memory_order_relaxed);
    s.gen.store(g, std::memory_order_release);
  }
  return s.set.load(std::memory_order_relaxed) ? s.value.load(std::memory_order_relaxed) : dflt;
}

}  // namespace b2r

// One static slot per call site.
#define B2R_ENV_INT(name, dflt)                        \
  ([&]() -> int {                                      \
    static ::b2r::EnvSlot b2r_env_slot_;               \
    return ::b2r::env_int_cached(b2r_env_slot_, name, dflt); \
  }())
