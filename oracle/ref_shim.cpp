// oracle/ref_shim.cpp — TEST INFRASTRUCTURE, not product code.
//
// A C-ABI window onto the UNMODIFIED reference sources under /root/reference
// (compiled in place by oracle/Makefile into oracle/_ref/libcog_ref.so).  It
// instantiates the reference's own containers
//   vec_cog_env<N>          include/vec_environment.h:12-81
//   vec_action_sampler<N>   include/vec_sampler.h:7-27
//   ThreadedRunner<N>       include/runner.h:21-134
// for a fixed set of N and exposes raw pointers to their buffers, so that
//   * tests/ can mint golden trajectories and pin oracle/cog_oracle.c to them,
//   * bench.py --impl reference can time the reference's C++ thread pool
//     with no Python inside the timed region.
// Nothing here restates game logic; every call forwards to reference code.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference
// arm may load this library.

#include <algorithm>
#include <array>
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <memory>
#include <new>
#include <optional>
#include <random>
#include <sstream>
#include <csignal>
#include <cstddef>
#include <numeric>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

// The reference keeps its per-env objects private; the shim needs to read
// hidden state (rng, player locations, turn counter) for debugging dumps.
// Access specifiers do not change layout or code generation.
#include "atomic_queues.hpp" // vendored third-party queue, included before the access hack
#define private public
#include "api.h"
#include "environment.h"
#include "map.h"
#include "runner.h"
#include "sampler.h"
#include "vec_environment.h"
#include "vec_sampler.h"
#undef private

#define SHIM_API extern "C" __attribute__((visibility("default")))

namespace {

thread_local std::string g_err;

struct IVec {
  virtual ~IVec() = default;
  virtual size_t n() const = 0;
  virtual void reset_default() = 0;
  virtual void reset(uint32_t seed, uint8_t n_players, uint8_t n_pieces,
                     int difficulty, unsigned max_steps) = 0;
  virtual void step(const void *actions) = 0;
  virtual void sample(const void *masks) = 0;
  virtual void *buffer(int which) = 0;
  virtual void make_runner(size_t n_threads) = 0;
  virtual void drop_runner() = 0;
  virtual double time_run(size_t steps, int mode) = 0;
  virtual double time_sample(size_t iters) = 0;
  virtual double time_reset(size_t iters) = 0;
  virtual void hidden(size_t i, uint32_t *out) = 0;
};

template <size_t N> struct VecImpl final : IVec {
  // heap: vec_cog_env<65536> is > 1 GiB
  struct EnvDeleter;
  std::unique_ptr<vec_cog_env<N>, EnvDeleter> envs;
  std::unique_ptr<vec_action_sampler<N>> samplers;
  std::unique_ptr<ThreadedRunner<N>> runner;

  // The reference never initialises the Player slots of seats >= n_players, yet get_reward sums
  // has_won over all four (src/environment.cpp:281-288, SURVEY A#10): with recycled heap memory a
  // 2-player game can report garbage rewards.  The storage is zeroed before the reference's own
  // constructor runs, i.e. every instance behaves like one on a fresh allocation.
  static vec_cog_env<N> *fresh_envs() {
    void *mem = ::operator new(sizeof(vec_cog_env<N>), std::align_val_t(alignof(vec_cog_env<N>)));
    std::memset(mem, 0, sizeof(vec_cog_env<N>));
    return new (mem) vec_cog_env<N>();
  }
  struct EnvDeleter {
    void operator()(vec_cog_env<N> *p) const {
      p->~vec_cog_env<N>();
      ::operator delete(p, std::align_val_t(alignof(vec_cog_env<N>)));
    }
  };
  explicit VecImpl(uint64_t sampler_seed)
      : envs(fresh_envs()),
        samplers(new vec_action_sampler<N>(static_cast<uint32_t>(sampler_seed))) {}

  size_t n() const override { return N; }
  void reset_default() override { envs->reset(); }
  void reset(uint32_t seed, uint8_t n_players, uint8_t n_pieces, int difficulty,
             unsigned max_steps) override {
    envs->reset(seed, n_players, n_pieces, static_cast<Difficulty>(difficulty),
                max_steps, false);
  }
  void step(const void *actions) override {
    // same reinterpretation as include/pybind/vectorized.h:60-68
    envs->step(*reinterpret_cast<const std::array<ActionData, N> *>(actions));
  }
  void sample(const void *masks) override {
    // include/pybind/vectorized.h:118-125
    samplers->sample(*reinterpret_cast<const std::array<ActionMask, N> *>(masks));
  }
  void *buffer(int which) override {
    switch (which) {
    case 0: return const_cast<ObsData *>(envs->get_observations().data());
    case 1: return const_cast<ActionMask *>(envs->get_selected_action_masks().data());
    case 2: return const_cast<float *>(envs->get_rewards()[0].data());
    case 3: return const_cast<bool *>(envs->get_dones().data());
    case 4: return const_cast<Info *>(envs->get_infos().data());
    case 5: return const_cast<u_char *>(envs->get_agent_selections().data());
    case 6: return const_cast<ActionData *>(samplers->get_actions().data());
    default: return nullptr;
    }
  }
  void make_runner(size_t n_threads) override {
    runner.reset();
    runner.reset(new ThreadedRunner<N>(*envs, *samplers, n_threads));
  }
  void drop_runner() override { runner.reset(); }

  // benchmarks/benchmarks.py:47-51 (time_run) with the Python loop replaced by
  // a C++ loop.  mode 0 = sequential, 1 = runner async, 2 = runner sync.
  double time_run(size_t steps, int mode) override {
    const auto &masks = envs->get_selected_action_masks();
    const auto &actions = samplers->get_actions();
    auto t0 = std::chrono::steady_clock::now();
    if (mode == 0) {
      for (size_t s = 0; s < steps; ++s) {
        samplers->sample(masks);
        envs->step(actions);
      }
    } else {
      if (!runner) throw std::runtime_error("runner not created");
      for (size_t s = 0; s < steps; ++s) {
        runner->sample();
        runner->step();
        if (mode == 2) runner->sync();
      }
      runner->sync();
    }
    auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
  }

  // benchmarks/benchmarks.py:53-57 (time_sample): the sampler alone on the current masks
  double time_sample(size_t iters) override {
    const auto &masks = envs->get_selected_action_masks();
    auto t0 = std::chrono::steady_clock::now();
    for (size_t s = 0; s < iters; ++s) samplers->sample(masks);
    auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
  }
  // benchmarks/benchmarks.py:59-63 (time_reset): vec.reset() of the whole batch, sequential
  double time_reset(size_t iters) override {
    auto t0 = std::chrono::steady_clock::now();
    for (size_t s = 0; s < iters; ++s) envs->reset();
    auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
  }

  // hidden per-env state for debugging dumps (never compared as "parity")
  void hidden(size_t i, uint32_t *out) override {
    cog_env &e = envs->environments[i];
    std::ostringstream ss;
    ss << e.rng;
    out[0] = static_cast<uint32_t>(std::stoul(ss.str()));
    out[1] = e.turn_counter;
    out[2] = e.agent_selection;
    out[3] = e.shop.n_in_market;
    for (int p = 0; p < 4; ++p) {
      point loc = p < (int)e.map.player_locations.size()
                      ? e.map.player_locations[p]
                      : point{0, 0};
      out[4 + 2 * p] = static_cast<uint32_t>(static_cast<int>(loc.x * 2));
      out[5 + 2 * p] = static_cast<uint32_t>(static_cast<int>(loc.y * 2));
    }
    out[12] = static_cast<uint32_t>(static_cast<int>(e.map.min_xy.x * 2));
    out[13] = static_cast<uint32_t>(static_cast<int>(e.map.min_xy.y * 2));
    out[14] = static_cast<uint32_t>(e.map.pieces.size());
    out[15] = static_cast<uint32_t>(e.map.xy.size());
  }
};

IVec *make(size_t n, uint64_t sampler_seed) {
  switch (n) {
  case 1: return new VecImpl<1>(sampler_seed);
  case 16: return new VecImpl<16>(sampler_seed);
  case 64: return new VecImpl<64>(sampler_seed);
  case 256: return new VecImpl<256>(sampler_seed);
  case 1024: return new VecImpl<1024>(sampler_seed);
  case 4096: return new VecImpl<4096>(sampler_seed);
  case 16384: return new VecImpl<16384>(sampler_seed);
  case 65536: return new VecImpl<65536>(sampler_seed);
  default: return nullptr;
  }
}

} // namespace

SHIM_API const char *ref_last_error() { return g_err.c_str(); }

SHIM_API void *ref_create(uint64_t n_envs, uint64_t sampler_seed) {
  try {
    IVec *v = make(n_envs, sampler_seed);
    if (!v) g_err = "unsupported n_envs (have 1,16,64,256,1024,4096,16384,65536)";
    return v;
  } catch (const std::exception &e) {
    g_err = e.what();
    return nullptr;
  }
}

SHIM_API void ref_destroy(void *h) { delete static_cast<IVec *>(h); }

SHIM_API int ref_reset(void *h, uint32_t seed, uint8_t n_players, uint8_t n_pieces,
                       int difficulty, uint32_t max_steps) {
  try {
    static_cast<IVec *>(h)->reset(seed, n_players, n_pieces, difficulty, max_steps);
    return 0;
  } catch (const generate_map_failure &e) {
    g_err = e.what();
    return -2;
  } catch (const std::exception &e) {
    g_err = e.what();
    return -1;
  }
}

SHIM_API int ref_reset_default(void *h) {
  try {
    static_cast<IVec *>(h)->reset_default();
    return 0;
  } catch (const generate_map_failure &e) {
    g_err = e.what();
    return -2;
  } catch (const std::exception &e) {
    g_err = e.what();
    return -1;
  }
}

SHIM_API int ref_step(void *h, const void *actions) {
  try {
    static_cast<IVec *>(h)->step(actions);
    return 0;
  } catch (const generate_map_failure &e) {
    g_err = e.what();
    return -2;
  } catch (const std::exception &e) {
    g_err = e.what();
    return -1;
  }
}

SHIM_API int ref_sample(void *h, const void *masks) {
  static_cast<IVec *>(h)->sample(masks);
  return 0;
}

// which: 0 obs, 1 selected masks, 2 rewards, 3 dones, 4 infos, 5 agent_selection,
//        6 sampler actions
SHIM_API void *ref_buffer(void *h, int which) {
  return static_cast<IVec *>(h)->buffer(which);
}

SHIM_API int ref_runner_create(void *h, uint64_t n_threads) {
  try {
    static_cast<IVec *>(h)->make_runner(n_threads);
    return 0;
  } catch (const std::exception &e) {
    g_err = e.what();
    return -1;
  }
}

SHIM_API void ref_runner_destroy(void *h) { static_cast<IVec *>(h)->drop_runner(); }

SHIM_API double ref_time_run(void *h, uint64_t steps, int mode) {
  try {
    return static_cast<IVec *>(h)->time_run(steps, mode);
  } catch (const std::exception &e) {
    g_err = e.what();
    return -1.0;
  }
}

SHIM_API double ref_time_sample(void *h, uint64_t iters) {
  try {
    return static_cast<IVec *>(h)->time_sample(iters);
  } catch (const std::exception &e) {
    g_err = e.what();
    return -1.0;
  }
}

SHIM_API double ref_time_reset(void *h, uint64_t iters) {
  try {
    return static_cast<IVec *>(h)->time_reset(iters);
  } catch (const std::exception &e) {
    g_err = e.what();
    return -1.0;
  }
}

SHIM_API void ref_hidden(void *h, uint64_t i, uint32_t *out16) {
  static_cast<IVec *>(h)->hidden(i, out16);
}

SHIM_API uint64_t ref_sizeof(int which) {
  switch (which) {
  case 0: return sizeof(ObsData);
  case 1: return sizeof(ActionMask);
  case 2: return sizeof(float) * MAX_N_PLAYERS;
  case 3: return sizeof(bool);
  case 4: return sizeof(Info);
  case 5: return sizeof(u_char);
  case 6: return sizeof(ActionData);
  case 7: return sizeof(PlayerData);
  case 8: return sizeof(cog_env);
  default: return 0;
  }
}

// ---- the reference's SINGLE env and sampler (src/pybind/single_env.cpp:13-32, common.cpp:25-27) ----
// cog_env needs caller-owned records (cog_env::init); the shim owns them, zeroed like a fresh allocation.
namespace {
struct Single {
  ObsData obs;
  Info info;
  std::array<float, MAX_N_PLAYERS> rewards;
  ActionMask selected;
  cog_env env;
  Single(uint32_t seed, u_char n_players, u_char n_pieces, int difficulty, unsigned max_steps)
      : obs{}, info{}, rewards{}, selected{},
        env(seed, n_players, n_pieces, static_cast<Difficulty>(difficulty), max_steps, false) {
    env.init(obs, info, rewards, selected);
  }
};
} // namespace

SHIM_API void *ref_single_create(uint32_t seed, uint8_t n_players, uint8_t n_pieces, int difficulty,
                                 uint32_t max_steps) {
  try {
    void *mem = ::operator new(sizeof(Single), std::align_val_t(alignof(Single)));
    std::memset(mem, 0, sizeof(Single));
    return new (mem) Single(seed, n_players, n_pieces, difficulty, max_steps);
  } catch (const std::exception &e) {
    g_err = e.what();
    return nullptr;
  }
}
SHIM_API void ref_single_destroy(void *h) {
  if (!h) return;
  static_cast<Single *>(h)->~Single();
  ::operator delete(h, std::align_val_t(alignof(Single)));
}
SHIM_API int ref_single_reset(void *h, int reseed, uint32_t seed, uint8_t n_players, uint8_t n_pieces,
                              int difficulty, uint32_t max_steps) {
  try {
    Single *s = static_cast<Single *>(h);
    if (reseed) s->env.reset(seed, n_players, n_pieces, static_cast<Difficulty>(difficulty), max_steps, false);
    else s->env.reset();
    return 0;
  } catch (const generate_map_failure &e) {
    g_err = e.what();
    return -2;
  } catch (const std::exception &e) {
    g_err = e.what();
    return -1;
  }
}
SHIM_API int ref_single_step(void *h, const void *action) {
  try {
    ActionData a{}; /* the caller's record need not be 64-byte aligned */
    std::memcpy(&a, action, 5);
    static_cast<Single *>(h)->env.step(a);
    return 0;
  } catch (const std::exception &e) {
    g_err = e.what();
    return -1;
  }
}
// which: 0 obs, 1 selected mask, 2 rewards, 4 info
SHIM_API void *ref_single_buffer(void *h, int which) {
  Single *s = static_cast<Single *>(h);
  switch (which) {
  case 0: return &s->obs;
  case 1: return &s->selected;
  case 2: return s->rewards.data();
  case 4: return &s->info;
  default: return nullptr;
  }
}
// what: 0 agent_selection, 1 done, 2 seed, 3 n_players
SHIM_API uint32_t ref_single_get(void *h, int what) {
  Single *s = static_cast<Single *>(h);
  switch (what) {
  case 0: return s->env.get_agent_selection();
  case 1: return s->env.get_done();
  case 2: return s->env.get_seed();
  case 3: return s->env.get_n_players();
  default: return 0;
  }
}
SHIM_API void *ref_single_sampler_create(uint32_t seed) { return new action_sampler(seed); }
SHIM_API void ref_single_sampler_destroy(void *h) { delete static_cast<action_sampler *>(h); }
SHIM_API void ref_single_sampler_sample(void *h, const void *mask, void *action_out) {
  ActionMask m{};
  std::memcpy(&m, mask, 92); /* payload bytes; the caller's record need not be 64-byte aligned */
  const ActionData a = static_cast<action_sampler *>(h)->sample(m);
  std::memcpy(action_out, &a, 5);
}

// ---- standalone Map::generate driver -------------------------------------
// Drives the reference's Map::generate (src/map.cpp:697-742) on a caller-owned
// observation buffer that may be LARGER than 48x48x7 so that boards with many
// pieces (which the reference would write out of bounds) can still be used to
// pin the retry/recursion logic.  rows = first dimension of the buffer actually
// allocated by the caller (>= 48); stride per row stays 48*7 as in the
// reference's MapObservation type, so out-of-range y spills into the next row
// exactly as it would in the reference.
SHIM_API int ref_map_generate(uint32_t seed, uint8_t n_pieces, int difficulty,
                              void *obs_buf, uint32_t *out_meta, int32_t *out_xy,
                              uint32_t max_hexes) {
  try {
    Map map;
    map.init(*reinterpret_cast<MapObservation *>(obs_buf));
    std::default_random_engine rng(seed);
    map.reset();
    map.generate(n_pieces, static_cast<Difficulty>(difficulty), 0, MAX_FAILURES, rng);
    out_meta[0] = static_cast<uint32_t>(map.pieces.size());
    out_meta[1] = static_cast<uint32_t>(map.xy.size());
    out_meta[2] = static_cast<uint32_t>(static_cast<int>(map.min_xy.x * 2));
    out_meta[3] = static_cast<uint32_t>(static_cast<int>(map.min_xy.y * 2));
    out_meta[4] = static_cast<uint32_t>(static_cast<int>(map.max_xy.x * 2));
    out_meta[5] = static_cast<uint32_t>(static_cast<int>(map.max_xy.y * 2));
    for (size_t i = 0; i < map.xy.size() && i < max_hexes; ++i) {
      out_xy[2 * i] = static_cast<int>(map.xy[i].x * 2);
      out_xy[2 * i + 1] = static_cast<int>(map.xy[i].y * 2);
    }
    return 0;
  } catch (const generate_map_failure &e) {
    g_err = e.what();
    return -2;
  } catch (const std::exception &e) {
    g_err = e.what();
    return -1;
  }
}
