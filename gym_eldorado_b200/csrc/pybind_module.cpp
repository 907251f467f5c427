/* pybind_module.cpp — the `city_of_gold` Python module, B200 edition.
 *
 * Host C++ behind the reference's own Python surface (reference src/pybind/common.cpp,
 * src/pybind/vectorized.cpp, include/pybind/vectorized.h): same module, class, method
 * and property names, same structured dtypes — but the containers live on the GPU and
 * every call goes through the C ABI of include/cog_b200.h (no CUDA or torch types here).
 *
 *   city_of_gold.Difficulty.{EASY,MEDIUM,HARD}
 *   city_of_gold.vec.get_vec_env(N) / get_vec_sampler(N) / get_runner(N) -> class objects
 *   Env().reset(), .reset(seed, n_players, n_pieces, difficulty, max_steps, render), .step(actions)
 *   Env.observations / selected_action_masks / rewards / dones / infos / agent_selection / num_envs
 *   Sampler(seed=None).get_actions() / .sample(action_mask)
 *   Runner(env, sampler, n_threads=None).sample() / step() / step_sync() / sync() / get_*()
 *
 * Where the reference hands out zero-copy numpy views of host arrays, this module hands
 * out `DeviceBuffer` objects: persistent, zero-copy views of DEVICE memory exposing
 * `__cuda_array_interface__` and `__dlpack__` (uint8[N, itemsize] for the structured
 * records, f32[N,4] / bool[N] / u8[N] for the plain ones), `.dtype` (the reference's
 * structured dtype) and `.numpy()` (an explicit device->host copy as a structured array).
 */
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <cstdint>
#include <memory>
#include <optional>
#include <random>
#include <stdexcept>
#include <string>
#include <vector>

#include "cog_b200.h"

#ifndef COG_MODULE_NAME
#define COG_MODULE_NAME city_of_gold
#endif

namespace py = pybind11;

namespace {

/* ---- minimal DLPack ABI (dlpack.h v0.8, stable C structs) ---- */
struct DLDevice { int32_t device_type, device_id; };
struct DLDataType { uint8_t code, bits; uint16_t lanes; };
struct DLTensor {
  void *data; DLDevice device; int32_t ndim; DLDataType dtype;
  int64_t *shape, *strides; uint64_t byte_offset;
};
struct DLManagedTensor {
  DLTensor dl_tensor; void *manager_ctx; void (*deleter)(DLManagedTensor *);
};
enum { kDLCPU = 1, kDLCUDA = 2, kDLUInt = 1, kDLFloat = 2, kDLBool = 6 };

[[noreturn]] void raise_last(int rc) {
  const std::string msg = cog_last_error();
  if (rc == COG_ERR_INVALID) throw py::value_error(msg);
  throw std::runtime_error(msg); /* generate_map_failure is a RuntimeError in the reference too */
}
inline void ck(int rc) {
  if (rc != COG_OK) raise_last(rc);
}

struct VecOwner { /* shared so that views keep the device memory alive */
  cog_vec *h = nullptr;
  ~VecOwner() { cog_vec_destroy(h); }
};
struct SamplerOwner {
  cog_sampler *s = nullptr;
  ~SamplerOwner() { cog_sampler_destroy(s); }
};

/* A persistent zero-copy view of one exported device buffer. */
class DeviceBuffer {
public:
  std::shared_ptr<VecOwner> vec;         /* exactly one of the two owners is set */
  std::shared_ptr<SamplerOwner> sampler;
  int which = 0;
  void *ptr = nullptr;
  size_t itemsize = 0, count = 0;
  int device = 0;
  py::object np_dtype; /* the reference's (structured) dtype of one record */
  std::string typestr; /* element type of the raw view */
  std::vector<int64_t> shape;
  DLDataType dl{};

  py::dict cuda_array_interface() const {
    py::dict d;
    py::tuple shp(shape.size());
    for (size_t i = 0; i < shape.size(); ++i) shp[i] = shape[i];
    d["shape"] = shp;
    d["typestr"] = typestr;
    d["data"] = py::make_tuple((uintptr_t)ptr, false);
    d["version"] = 3;
    d["strides"] = py::none();
    d["stream"] = 1; /* legacy default stream: where the library enqueues its kernels */
    return d;
  }

  /* the consumer's stream (DLPack / CUDA array interface convention: None or 1 = legacy default stream,
   * 2 = per-thread default stream, anything else = a cudaStream_t) waits for what the library has
   * enqueued so far: an event, no host synchronisation */
  void release_to(const py::object &stream) const {
    if (stream.is_none()) return;
    const intptr_t v = py::cast<intptr_t>(stream);
    if (v == 1 || v == 0) return; /* the library's own stream (the module creates every handle on it) */
    void *consumer = v == 2 ? (void *)(intptr_t)2 /* cudaStreamPerThread */ : (void *)v;
    if (vec) ck(cog_vec_release_to_stream(vec->h, consumer));
    else ck(cog_sampler_release_to_stream(sampler->s, consumer));
  }

  /* legacy (unversioned) capsule; extra protocol keywords (max_version, copy, dl_device) are accepted */
  py::capsule dlpack(py::object stream, py::kwargs /*unused*/) const {
    release_to(stream);
    struct Ctx { std::vector<int64_t> shape; std::shared_ptr<VecOwner> v; std::shared_ptr<SamplerOwner> s; };
    auto *ctx = new Ctx{shape, vec, sampler};
    auto *mt = new DLManagedTensor{};
    mt->dl_tensor.data = ptr;
    mt->dl_tensor.device = DLDevice{(int32_t)(device < 0 ? kDLCPU : kDLCUDA), device < 0 ? 0 : device};
    mt->dl_tensor.ndim = (int32_t)ctx->shape.size();
    mt->dl_tensor.dtype = dl;
    mt->dl_tensor.shape = ctx->shape.data();
    mt->dl_tensor.strides = nullptr;
    mt->dl_tensor.byte_offset = 0;
    mt->manager_ctx = ctx;
    mt->deleter = [](DLManagedTensor *self) {
      delete static_cast<Ctx *>(self->manager_ctx);
      delete self;
    };
    return py::capsule(mt, "dltensor", [](PyObject *cap) {
      if (PyCapsule_IsValid(cap, "dltensor")) { /* never consumed */
        auto *m = static_cast<DLManagedTensor *>(PyCapsule_GetPointer(cap, "dltensor"));
        m->deleter(m);
      }
    });
  }

  py::tuple dlpack_device() const { return py::make_tuple((int)(device < 0 ? kDLCPU : kDLCUDA), device < 0 ? 0 : device); }

  py::array numpy(py::object out) const {
    py::array arr;
    if (out.is_none()) {
      py::tuple shp(1);
      shp[0] = count;
      arr = py::module_::import("numpy").attr("empty")(shp, np_dtype);
    } else {
      arr = py::array::ensure(out);
      if (!arr || (size_t)arr.nbytes() != itemsize * count || !(arr.flags() & py::array::c_style))
        throw py::value_error("out must be a C-contiguous array of the buffer's size");
    }
    if (vec) ck(cog_vec_read(vec->h, which, arr.mutable_data(), itemsize * count));
    else ck(cog_sampler_read_actions(sampler->s, arr.mutable_data(), itemsize * count));
    if (out.is_none() && shape.size() == 3) return arr.attr("reshape")(shape[0], shape[1]);
    return arr;
  }
};

/* numpy's array protocol: np.asarray(buffer), np.nonzero(envs.dones), envs.infos[i] — what unmodified
 * reference user code does with the views (README.md / pytest/test_vec_env.py:97-105).  Each use is an
 * explicit device -> host copy of the whole buffer (the views are device memory here). */
py::object buffer_array(const DeviceBuffer &b, py::object dtype, py::object /*copy*/) {
  py::object arr = b.numpy(py::none());
  if (!dtype.is_none()) arr = arr.attr("astype")(dtype);
  return arr;
}

/* device pointer of an argument that may be a DeviceBuffer, anything exposing
 * __cuda_array_interface__ (torch / cupy), or a host numpy array (returns nullptr and
 * sets *host) — the reference's step()/sample() take the buffer unchecked as well */
void *resolve_pointer(const py::object &obj, size_t need_bytes, const void **host, void **producer_stream) {
  *host = nullptr;
  *producer_stream = nullptr;
  if (py::isinstance<DeviceBuffer>(obj)) {
    const DeviceBuffer &b = obj.cast<const DeviceBuffer &>();
    if (b.itemsize * b.count < need_bytes) throw py::value_error("device buffer holds fewer bytes than one record per env");
    return b.ptr;
  }
  if (py::hasattr(obj, "__cuda_array_interface__")) {
    py::dict d = obj.attr("__cuda_array_interface__");
    /* size check: product(shape) * itemsize from the typestr ("|u1", "<f4", "|V64", ...) */
    size_t count = 1;
    for (auto dim : py::tuple(d["shape"])) count *= py::cast<size_t>(dim);
    const std::string typestr = py::cast<std::string>(d["typestr"]);
    const size_t item = typestr.size() > 2 ? (size_t)std::stoul(typestr.substr(2)) : 1;
    if (count * item < need_bytes) throw py::value_error("CUDA array holds fewer bytes than one record per env");
    if (d.contains("strides") && !d["strides"].is_none()) {
      /* C-contiguous only: the last stride must be the item size and each earlier one the product below it */
      py::tuple st = d["strides"], sh = d["shape"];
      size_t expect = item;
      for (size_t k = st.size(); k-- > 0;) {
        if (py::cast<size_t>(sh[k]) > 1 && py::cast<size_t>(st[k]) != expect)
          throw py::value_error("CUDA array must be C-contiguous");
        expect *= py::cast<size_t>(sh[k]);
      }
    }
    if (d.contains("stream") && !d["stream"].is_none()) {
      const intptr_t v = py::cast<intptr_t>(d["stream"]);
      if (v != 0 && v != 1) *producer_stream = v == 2 ? (void *)(intptr_t)2 : (void *)v;
    }
    return (void *)py::cast<uintptr_t>(py::tuple(d["data"])[0]);
  }
  py::array arr = py::array::ensure(obj);
  if (!arr) throw py::type_error("expected a DeviceBuffer, a CUDA array or a numpy array");
  if (!(arr.flags() & py::array::c_style) || (size_t)arr.nbytes() < need_bytes)
    throw py::value_error("host array must be C-contiguous and hold one record per env");
  *host = arr.data();
  return nullptr;
}
/* optional host output array of exactly `bytes` bytes (None -> nullptr) */
void *host_out(const py::object &obj, size_t bytes, const char *what) {
  if (obj.is_none()) return nullptr;
  py::array arr = py::array::ensure(obj);
  if (!arr || !(arr.flags() & py::array::c_style) || !arr.writeable() || (size_t)arr.nbytes() != bytes)
    throw py::value_error(std::string(what) + ": expected a writable C-contiguous host array of one record per env");
  return arr.mutable_data();
}

/* tests/emu links this module against the emulated library, whose "device" pointers are
 * host pointers: views then advertise kDLCPU so numpy can consume the DLPack capsule */
inline int buffer_device(int device) {
#ifdef COG_HOST_POINTERS
  (void)device;
  return -1;
#else
  return device;
#endif
}

py::object layout_dtype(const char *name) {
  return py::module_::import("gym_eldorado_b200.layout").attr(name);
}

class VecEnv {
public:
  std::shared_ptr<VecOwner> own = std::make_shared<VecOwner>();
  uint64_t n;
  int device;

  VecEnv(uint64_t n_envs, int device_, uint64_t env_index_offset) : n(n_envs), device(device_) {
    ck(cog_vec_create(n_envs, env_index_offset, device_, nullptr, &own->h));
  }
  void reset_default() { ck(cog_vec_reset_default(own->h)); }
  void reset(uint32_t seed, uint8_t n_players, uint8_t n_pieces, int difficulty, uint32_t max_steps, bool render) {
    ck(cog_vec_reset(own->h, seed, n_players, n_pieces, difficulty, max_steps, render));
  }
  /* step(actions): a device array (DeviceBuffer / CUDA array interface: asynchronous, ordered after its
   * producer's stream) or a host numpy array like the reference's (include/pybind/vectorized.h:60-68).
   * With out_* host arrays the call goes through cog_vec_step_host_io: zero copy for pinned arrays, the
   * chunked copy/compute pipeline otherwise, and the results are in the arrays when it returns. */
  void step(const py::object &actions, const py::object &out_masks, const py::object &out_rewards,
            const py::object &out_dones, const py::object &out_agents) {
    const void *host;
    void *producer;
    void *dev = resolve_pointer(actions, n * COG_ACTION_SIZE, &host, &producer);
    const bool outs = !(out_masks.is_none() && out_rewards.is_none() && out_dones.is_none() && out_agents.is_none());
    if (host && outs) {
      ck(cog_vec_step_host_io(own->h, host, host_out(out_masks, n * COG_MASK_SIZE, "out_masks"),
                              host_out(out_rewards, n * 16, "out_rewards"), host_out(out_dones, n, "out_dones"),
                              host_out(out_agents, n, "out_agents")));
      return;
    }
    if (outs) throw py::value_error("out_* arrays need host actions (device callers read the device views)");
    if (producer) ck(cog_vec_wait_stream(own->h, producer));
    ck(host ? cog_vec_step_host(own->h, host) : cog_vec_step(own->h, dev));
  }
  DeviceBuffer view(int which) const {
    DeviceBuffer b;
    b.vec = own;
    b.which = which;
    b.device = buffer_device(device);
    ck(cog_vec_buffer(own->h, which, &b.ptr, &b.itemsize, &b.count));
    switch (which) {
    case COG_BUF_OBS: b.np_dtype = layout_dtype("obs_dtype"); break;
    case COG_BUF_SELECTED_MASKS:
    case COG_BUF_MERGED_MASKS: b.np_dtype = layout_dtype("action_mask_dtype"); break;
    case COG_BUF_INFOS: b.np_dtype = layout_dtype("info_dtype"); break;
    case COG_BUF_COMPACT:
    case COG_BUF_ROLLOUT: b.np_dtype = layout_dtype("compact_dtype"); break;
    case COG_BUF_GRID: b.np_dtype = py::dtype("2304u1"); break;
    default: break;
    }
    if (which == COG_BUF_REWARDS) {
      b.np_dtype = py::dtype("4<f4"); b.typestr = "<f4"; b.shape = {(int64_t)n, 4}; b.dl = {kDLFloat, 32, 1};
    } else if (which == COG_BUF_DONES) {
      b.np_dtype = py::dtype("bool"); b.typestr = "|b1"; b.shape = {(int64_t)n}; b.dl = {kDLBool, 8, 1};
    } else if (which == COG_BUF_AGENT_SELECTION) {
      b.np_dtype = py::dtype("uint8"); b.typestr = "|u1"; b.shape = {(int64_t)n}; b.dl = {kDLUInt, 8, 1};
    } else if (which == COG_BUF_ROLLOUT) { /* env-major [n][n_slots] records */
      b.typestr = "|u1"; b.shape = {(int64_t)n, (int64_t)(b.count / n), (int64_t)b.itemsize}; b.dl = {kDLUInt, 8, 1};
    } else {
      b.typestr = "|u1"; b.shape = {(int64_t)n, (int64_t)b.itemsize}; b.dl = {kDLUInt, 8, 1};
    }
    return b;
  }
  py::dict stats() const {
    uint64_t v[COG_STAT_COUNT];
    ck(cog_vec_stats(own->h, v, COG_STAT_COUNT));
    static const char *names[COG_STAT_COUNT] = {"env_steps", "episodes", "truncations", "total_length",
        "wins_seat0", "wins_seat1", "wins_seat2", "wins_seat3", "travelled_hexes", "cards_added", "resets", "reserved"};
    py::dict d;
    for (int i = 0; i < COG_STAT_COUNT; ++i) d[names[i]] = v[i];
    return d;
  }
};

class VecSampler {
public:
  std::shared_ptr<SamplerOwner> own = std::make_shared<SamplerOwner>();
  uint64_t n;
  int device;
  VecSampler(uint64_t n_envs, std::optional<uint64_t> seed, int device_, uint64_t env_index_offset)
      : n(n_envs), device(device_) {
    /* py_vec_action_sampler(uint32_t seed); None -> std::random_device
     * (include/pybind/vectorized.h:217-224) */
    const uint32_t s = seed ? (uint32_t)*seed : (uint32_t)std::random_device{}();
    ck(cog_sampler_create(n_envs, env_index_offset, s, device_, nullptr, &own->s));
  }
  DeviceBuffer get_actions() const {
    DeviceBuffer b;
    b.sampler = own;
    b.device = buffer_device(device);
    ck(cog_sampler_actions(own->s, &b.ptr));
    b.itemsize = COG_ACTION_SIZE;
    b.count = n;
    b.np_dtype = layout_dtype("action_dtype");
    b.typestr = "|u1";
    b.shape = {(int64_t)n, COG_ACTION_SIZE};
    b.dl = {kDLUInt, 8, 1};
    return b;
  }
  /* sample(action_mask[, out]): device or host masks (include/pybind/vectorized.h:118-125); with a host
   * `out` ActionData array the sampled actions are in it when the call returns (cog_sampler_sample_host_io) */
  void sample(const py::object &masks, const py::object &out) {
    const void *host;
    void *producer;
    void *dev = resolve_pointer(masks, n * COG_MASK_SIZE, &host, &producer);
    if (!out.is_none()) {
      if (!host) throw py::value_error("out= needs host masks (device callers read get_actions())");
      ck(cog_sampler_sample_host_io(own->s, host, host_out(out, n * COG_ACTION_SIZE, "out")));
      return;
    }
    if (producer) ck(cog_sampler_wait_stream(own->s, producer));
    ck(host ? cog_sampler_sample_host(own->s, host) : cog_sampler_sample(own->s, dev));
  }
};

/* ThreadedRunner<N> (include/runner.h:21-134, bound at include/pybind/vectorized.h:129-161):
 * the CPU thread pool becomes the GPU grid; sample()/step() enqueue kernels on the
 * stream, sync() waits; run(k) is the persistent fused kernel. */
class Runner {
public:
  py::object env_obj, sampler_obj; /* keep the Python objects (the reference keeps C++ refs) */
  VecEnv &env;
  VecSampler &smp;
  size_t n_threads;
  Runner(py::object e, py::object s, std::optional<size_t> threads)
      : env_obj(e), sampler_obj(s), env(e.cast<VecEnv &>()), smp(s.cast<VecSampler &>()),
        n_threads(threads.value_or(1)) {
    if (env.n != smp.n) throw py::value_error("env and sampler sizes differ");
  }
  /* runner.sample(); runner.step() — the reference's loop (include/runner.h:60-96) — is ONE kernel here:
   * sample() only notes that a draw from the env's own selected masks is due, step() then launches
   * cog_vec_step_sampled (sample + step, no mask / action round trip).  Anything that could observe the
   * difference — get_actions(), sync(), another sample() — first launches the sampling kernel on its own. */
  bool sample_due = false;
  void flush_sample() {
    if (!sample_due) return;
    sample_due = false;
    void *m; size_t a, b;
    ck(cog_vec_buffer(env.own->h, COG_BUF_SELECTED_MASKS, &m, &a, &b));
    ck(cog_sampler_sample(smp.own->s, m));
  }
  void sample() {
    flush_sample();
    sample_due = true;
  }
  void step() {
    if (sample_due) {
      sample_due = false;
      ck(cog_vec_step_sampled(env.own->h, smp.own->s));
      return;
    }
    void *a;
    ck(cog_sampler_actions(smp.own->s, &a));
    ck(cog_vec_step(env.own->h, a));
  }
  void sync() {
    flush_sample();
    ck(cog_sync(env.own->h));
  }
  void step_sync() { step(); sync(); }
  void run(uint64_t k, bool merged) {
    flush_sample();
    ck(cog_vec_run_fused_ex(env.own->h, smp.own->s, k, merged ? COG_MASKS_MERGED : COG_MASKS_SELECTED));
  }
};

/* ---- the single env and sampler (reference src/pybind/single_env.cpp:13-86, src/pybind/common.cpp:25-27) ----
 * cog_env / action_sampler are N = 1 handles of the same kernels with the single env's semantics (no
 * auto-reset: cog_vec_set_autoreset(h, 0)).  The reference's cog_env needs caller-owned ObsData / Info
 * records passed to init() — records Python cannot construct there, Info is not even bound — so here the env
 * owns its records and hands out host copies of them as the reference's record classes (ObsData ->
 * .shared / .player_data[i] -> .obs / .action_mask ...), each a view of one numpy structured record. */
struct RecView {
  py::object rec; /* numpy structured scalar (np.void) or 0-d structured array */
};
struct ObsDataView : RecView {};
struct SharedObsView : RecView {};
struct PlayerDataView : RecView {};
struct ActionMaskView : RecView {};
struct PlayerObsView : RecView {};
struct InfoView : RecView {};

py::object rec_get(const RecView &v, const char *field) { return v.rec.attr("__getitem__")(field); }
template <class T> T rec_wrap(py::object rec) {
  T t;
  t.rec = std::move(rec);
  return t;
}

struct ActionDataPy { /* ActionData (include/api.h:146-161): five indices, 0 = none */
  uint8_t play = 0, play_special = 0, remove = 0, move = 0, get_from_shop = 0;
};

class SingleEnv {
public:
  VecEnv vec;
  uint32_t seed;
  uint8_t n_players, n_pieces;
  int difficulty;
  uint32_t max_steps;
  bool render_flag, seeded = false;
  /* cog_env::cog_env() src/environment.cpp:10-15 / cog_env(seed, ...) :17-24 */
  SingleEnv() : SingleEnv((uint32_t)std::random_device{}(), 4, 3, COG_EASY, 100000u, false) {}
  SingleEnv(uint32_t seed_, uint8_t n_players_, uint8_t n_pieces_, int difficulty_, uint32_t max_steps_, bool render_)
      : vec(1, 0, 0), seed(seed_), n_players(n_players_), n_pieces(n_pieces_), difficulty(difficulty_),
        max_steps(max_steps_), render_flag(render_) {
    ck(cog_vec_set_autoreset(vec.own->h, 0));
  }
  void reset_default() { /* cog_env::reset() :42-64: the constructor seeded the engine, later resets continue it */
    if (!seeded) {
      seeded = true;
      vec.reset(seed, n_players, n_pieces, difficulty, max_steps, render_flag);
    } else {
      vec.reset_default();
    }
  }
  void reset(uint32_t seed_, uint8_t n_players_, uint8_t n_pieces_, int difficulty_, uint32_t max_steps_, bool render_) {
    seed = seed_; n_players = n_players_; n_pieces = n_pieces_; difficulty = difficulty_; max_steps = max_steps_;
    render_flag = render_;
    seeded = true;
    vec.reset(seed, n_players, n_pieces, difficulty, max_steps, render_flag);
  }
  void step(const ActionDataPy &a) { /* cog_env::step :91-224; ignored once the episode is over (dead_step) */
    alignas(8) uint8_t rec[COG_ACTION_SIZE] = {a.play, a.play_special, a.remove, a.move, a.get_from_shop};
    ck(cog_vec_step_host(vec.own->h, rec));
    ck(cog_sync(vec.own->h));
  }
  py::object record(int which) const { return vec.view(which).numpy(py::none()).attr("__getitem__")(0); }
  uint8_t agent_selection() const { return py::cast<uint8_t>(record(COG_BUF_AGENT_SELECTION)); }
  bool done() const { return py::cast<bool>(record(COG_BUF_DONES)); }
};

class SingleSampler { /* action_sampler(seed = 42) include/sampler.h:8-12 */
public:
  VecSampler smp;
  explicit SingleSampler(uint32_t seed) : smp(1, (uint64_t)seed, 0, 0) {}
  ActionDataPy sample(const ActionMaskView &mask) {
    py::array rec = py::array::ensure(py::module_::import("numpy").attr("ascontiguousarray")(mask.rec).attr("reshape")(1));
    if (!rec || (size_t)rec.nbytes() != COG_MASK_SIZE) throw py::value_error("expected one ActionMask record");
    alignas(8) uint8_t out[COG_ACTION_SIZE];
    ck(cog_sampler_sample_host_io(smp.own->s, rec.data(), out));
    ActionDataPy a;
    a.play = out[0]; a.play_special = out[1]; a.remove = out[2]; a.move = out[3]; a.get_from_shop = out[4];
    return a;
  }
};

/* get_vec_env(N) & co. return CLASS objects named like the reference's
 * (vec_cog_env_N, vec_sampler_N, vec_runner_N; src/pybind/vectorized.cpp:8-21) */
py::object sized_class(const py::object &base, const std::string &name, uint64_t n, bool is_runner) {
  /* the cache lives on the base class object, so it dies with the interpreter, not after it */
  if (!py::hasattr(base, "_sized")) py::setattr(base, "_sized", py::dict());
  py::dict cache = base.attr("_sized");
  py::str key(name);
  if (cache.contains(key)) return cache[key];
  py::dict ns;
  if (!is_runner) {
    ns["__init__"] = py::cpp_function(
        [base, n](py::object self, py::args args, py::kwargs kw) {
          py::tuple full(args.size() + 1);
          full[0] = n;
          for (size_t i = 0; i < args.size(); ++i) full[i + 1] = args[i];
          base.attr("__init__")(self, *full, **kw);
        },
        py::is_method(py::none()));
  }
  ns["N"] = n;
  py::object cls = py::module_::import("builtins").attr("type")(key, py::make_tuple(base), ns);
  cache[key] = cls;
  return cls;
}

}  // namespace

PYBIND11_MODULE(COG_MODULE_NAME, m) {
  m.doc() = "City of Gold vectorised environments on B200: the reference's city_of_gold module surface (same names, "
            "arguments, dtypes) over device memory; the views are DeviceBuffers that numpy reads through explicit copies";

  py::enum_<cog_difficulty>(m, "Difficulty") /* src/pybind/common.cpp:32-36 */
      .value("EASY", COG_EASY).value("MEDIUM", COG_MEDIUM).value("HARD", COG_HARD).export_values();

  py::class_<DeviceBuffer>(m, "DeviceBuffer")
      .def_property_readonly("__cuda_array_interface__", &DeviceBuffer::cuda_array_interface)
      .def("__dlpack__", &DeviceBuffer::dlpack, py::arg("stream") = py::none())
      .def("__dlpack_device__", &DeviceBuffer::dlpack_device)
      .def("numpy", &DeviceBuffer::numpy, py::arg("out") = py::none(),
           "device -> host copy as a numpy array with the reference's structured dtype")
      .def("__array__", &buffer_array, py::arg("dtype") = py::none(), py::arg("copy") = py::none(),
           "np.asarray(view): an explicit device -> host copy as the reference's structured array")
      .def("__getitem__", [](const DeviceBuffer &b, py::object idx) { return b.numpy(py::none()).attr("__getitem__")(idx); },
           "view[i] like the reference's numpy views (a device -> host copy of the buffer, then the item)")
      .def_property_readonly("dtype", [](const DeviceBuffer &b) { return b.np_dtype; })
      .def_property_readonly("shape", [](const DeviceBuffer &b) -> py::tuple {
        if (b.which == COG_BUF_REWARDS && b.vec) return py::make_tuple(b.count, 4);
        return py::make_tuple(b.count);
      })
      .def_property_readonly("itemsize", [](const DeviceBuffer &b) { return b.itemsize; })
      .def_property_readonly("device_pointer", [](const DeviceBuffer &b) { return (uintptr_t)b.ptr; })
      .def("__len__", [](const DeviceBuffer &b) { return b.count; });

  /* ---- record classes and the single env (src/pybind/single_env.cpp:34-86,13-32; common.cpp:25-27) ---- */
  py::class_<ActionDataPy>(m, "ActionData")
      .def(py::init<>())
      .def(py::init([](uint8_t play, uint8_t play_special, uint8_t remove, uint8_t move, uint8_t get_from_shop) {
             ActionDataPy a;
             a.play = play; a.play_special = play_special; a.remove = remove; a.move = move; a.get_from_shop = get_from_shop;
             return a;
           }),
           py::arg("play") = 0, py::arg("play_special") = 0, py::arg("remove") = 0, py::arg("move") = 0,
           py::arg("get_from_shop") = 0)
      .def_readwrite("play", &ActionDataPy::play)
      .def_readwrite("play_special", &ActionDataPy::play_special)
      .def_readwrite("move", &ActionDataPy::move)
      .def_readwrite("get_from_shop", &ActionDataPy::get_from_shop)
      .def_readwrite("remove", &ActionDataPy::remove);
  py::class_<ActionMaskView>(m, "ActionMask")
      .def(py::init([](py::object rec) { return rec_wrap<ActionMaskView>(rec); }), py::arg("record"),
           "wrap one numpy ActionMask record (e.g. envs.selected_action_masks.numpy()[i])")
      .def_property_readonly("play", [](const ActionMaskView &v) { return rec_get(v, "play"); })
      .def_property_readonly("play_special", [](const ActionMaskView &v) { return rec_get(v, "play_special"); })
      .def_property_readonly("remove", [](const ActionMaskView &v) { return rec_get(v, "remove"); })
      .def_property_readonly("move", [](const ActionMaskView &v) { return rec_get(v, "move"); })
      .def_property_readonly("get_from_shop", [](const ActionMaskView &v) { return rec_get(v, "get_from_shop"); })
      .def_property_readonly("record", [](const ActionMaskView &v) { return v.rec; });
  py::class_<PlayerObsView>(m, "PlayerObservation")
      .def_property_readonly("draw", [](const PlayerObsView &v) { return rec_get(v, "draw"); })
      .def_property_readonly("hand", [](const PlayerObsView &v) { return rec_get(v, "hand"); })
      .def_property_readonly("active", [](const PlayerObsView &v) { return rec_get(v, "active"); })
      .def_property_readonly("played", [](const PlayerObsView &v) { return rec_get(v, "played"); })
      .def_property_readonly("discard", [](const PlayerObsView &v) { return rec_get(v, "discard"); });
  py::class_<PlayerDataView>(m, "PlayerData")
      .def_property_readonly("action_mask", [](const PlayerDataView &v) { return rec_wrap<ActionMaskView>(rec_get(v, "action_mask")); })
      .def_property_readonly("obs", [](const PlayerDataView &v) { return rec_wrap<PlayerObsView>(rec_get(v, "obs")); });
  py::class_<SharedObsView>(m, "SharedObservation")
      .def_property_readonly("phase", [](const SharedObsView &v) { return rec_get(v, "phase"); })
      .def_property_readonly("map", [](const SharedObsView &v) { return rec_get(v, "map"); })
      .def_property_readonly("shop", [](const SharedObsView &v) { return rec_get(v, "shop"); })
      .def_property_readonly("resources", [](const SharedObsView &v) { return rec_get(v, "current_resources"); });
  py::class_<ObsDataView>(m, "ObsData")
      .def_property_readonly("shared", [](const ObsDataView &v) { return rec_wrap<SharedObsView>(rec_get(v, "shared")); })
      .def_property_readonly("player_data", [](const ObsDataView &v) {
        py::list out;
        py::object pd = rec_get(v, "player_data");
        for (int p = 0; p < COG_MAX_PLAYERS; ++p) out.append(rec_wrap<PlayerDataView>(pd.attr("__getitem__")(p)));
        return out;
      });
  py::class_<InfoView>(m, "Info")
      .def_property_readonly("total_length", [](const InfoView &v) { return rec_get(v, "total_length"); })
      .def_property_readonly("agent_infos", [](const InfoView &v) { return rec_get(v, "agent_infos"); });

  py::class_<SingleEnv>(m, "cog_env")
      .def(py::init<>())
      .def(py::init<uint32_t, uint8_t, uint8_t, cog_difficulty, uint32_t, bool>())
      .def_property_readonly("agent_selection", &SingleEnv::agent_selection)
      .def("init", [](SingleEnv &, py::args, py::kwargs) {},
           "accepted for source compatibility: the env owns its records (see .observations / .info / .rewards / "
           ".selected_action_mask)")
      .def("reset", &SingleEnv::reset_default)
      .def("reset", [](SingleEnv &e, uint32_t seed, uint8_t n_players, uint8_t n_pieces, cog_difficulty d, uint32_t max_steps,
                       bool render) { e.reset(seed, n_players, n_pieces, (int)d, max_steps, render); })
      .def("step", &SingleEnv::step)
      .def("render", [](SingleEnv &) {}, "the reference prints text stubs; nothing is rendered here")
      .def("get_seed", [](const SingleEnv &e) { return e.seed; })
      .def("get_n_players", [](const SingleEnv &e) { return e.n_players; })
      .def("get_done", &SingleEnv::done)
      .def_property_readonly("observations", [](const SingleEnv &e) { return rec_wrap<ObsDataView>(e.record(COG_BUF_OBS)); })
      .def_property_readonly("selected_action_mask",
                             [](const SingleEnv &e) { return rec_wrap<ActionMaskView>(e.record(COG_BUF_SELECTED_MASKS)); })
      .def_property_readonly("info", [](const SingleEnv &e) { return rec_wrap<InfoView>(e.record(COG_BUF_INFOS)); })
      .def_property_readonly("rewards", [](const SingleEnv &e) { return e.record(COG_BUF_REWARDS); });
  py::class_<SingleSampler>(m, "action_sampler")
      .def(py::init<uint32_t>(), py::arg("seed") = 42)
      .def("sample", &SingleSampler::sample);

  auto vec = m.def_submodule("vec", "vectorised containers (reference src/pybind/vectorized.cpp)");

  py::class_<VecEnv> env(vec, "vec_cog_env");
  env.def(py::init<uint64_t, int, uint64_t>(), py::arg("n_envs"), py::arg("device") = 0,
          py::arg("env_index_offset") = 0)
      .def("reset", &VecEnv::reset_default)
      .def("reset", &VecEnv::reset, py::arg("seed"), py::arg("n_players"), py::arg("n_pieces"),
           py::arg("difficulty"), py::arg("max_steps"), py::arg("render"))
      .def("step", &VecEnv::step, py::arg("actions"), py::arg("out_masks") = py::none(),
           py::arg("out_rewards") = py::none(), py::arg("out_dones") = py::none(), py::arg("out_agents") = py::none())
      .def_property_readonly("observations", [](const VecEnv &e) { return e.view(COG_BUF_OBS); })
      .def_property_readonly("selected_action_masks", [](const VecEnv &e) { return e.view(COG_BUF_SELECTED_MASKS); })
      .def_property_readonly("rewards", [](const VecEnv &e) { return e.view(COG_BUF_REWARDS); })
      .def_property_readonly("dones", [](const VecEnv &e) { return e.view(COG_BUF_DONES); })
      .def_property_readonly("infos", [](const VecEnv &e) { return e.view(COG_BUF_INFOS); })
      .def_property_readonly("agent_selection", [](const VecEnv &e) { return e.view(COG_BUF_AGENT_SELECTION); })
      .def_property_readonly("num_envs", [](const VecEnv &e) { return e.n; })
      .def("merged_action_masks",
           [](const VecEnv &e) {
             ck(cog_vec_merged_masks(e.own->h, nullptr));
             return e.view(COG_BUF_MERGED_MASKS);
           },
           "not in the reference: the merged legal mask of each env's current agent (play / play_special from "
           "selected_action_masks; move / get_from_shop / remove from observations.player_data[agent].action_mask), "
           "recomputed on every call into one persistent device buffer")
      .def_property_readonly("hex_grid", [](const VecEnv &e) { return e.view(COG_BUF_GRID); },
           "not in the reference: the board as one byte per hex, uint8[n][48*48] (bits 0-2 requirement + 1, "
           "bits 3-5 n_required, bit 6 is_end); layout.grid_to_map() expands it to observations.shared.map")
      .def("compact_observations",
           [](const VecEnv &e, bool merged) {
             ck(cog_vec_compact_obs(e.own->h, nullptr, merged ? COG_MASKS_MERGED : COG_MASKS_SELECTED));
             return e.view(COG_BUF_COMPACT);
           },
           py::arg("merged_masks") = true,
           "not in the reference: one 256-byte record per env with everything the current agent acts on "
           "(layout.compact_dtype), recomputed on every call into one persistent device buffer")
      .def("attach_rollout", [](VecEnv &e, uint32_t n_slots) { ck(cog_vec_rollout_attach(e.own->h, n_slots)); },
           py::arg("n_slots"),
           "not in the reference: every later runner.run() appends one compact record per env-step (state, "
           "mask sampled from, action, done, rewards) to a device ring [n][n_slots] (each env's trajectory is "
           "contiguous); 0 detaches")
      .def_property_readonly("rollout", [](const VecEnv &e) { return e.view(COG_BUF_ROLLOUT); })
      .def_property_readonly("rollout_steps",
                             [](const VecEnv &e) {
                               uint64_t steps = 0;
                               ck(cog_vec_rollout_info(e.own->h, nullptr, nullptr, &steps));
                               return steps;
                             })
      .def("sync", [](VecEnv &e) { ck(cog_sync(e.own->h)); })
      .def("episode_stats", &VecEnv::stats)
      .def("reset_episode_stats", [](VecEnv &e) { ck(cog_vec_stats_reset(e.own->h)); });

  py::class_<VecSampler> sampler(vec, "vec_sampler");
  sampler.def(py::init<uint64_t, std::optional<uint64_t>, int, uint64_t>(), py::arg("n_envs"),
              py::arg("seed") = py::none(), py::arg("device") = 0, py::arg("env_index_offset") = 0)
      .def("get_actions", &VecSampler::get_actions)
      .def("sample", &VecSampler::sample, py::arg("action_mask"), py::arg("out") = py::none());

  py::class_<Runner> runner(vec, "vec_runner");
  runner.def(py::init<py::object, py::object, std::optional<size_t>>(), py::arg("env"), py::arg("sampler"),
             py::arg("n_threads") = py::none())
      .def("sample", &Runner::sample)
      .def("step", &Runner::step)
      .def("step_sync", &Runner::step_sync)
      .def("sync", &Runner::sync)
      .def("run", &Runner::run, py::arg("k_steps"), py::arg("merged_masks") = false,
           "k_steps x {sample; step} on the device, no host round trip; merged_masks=True samples from the "
           "merged legal mask of the current agent instead of selected_action_masks")
      .def("get_envs", [](Runner &r) { return r.env_obj; })
      .def("get_samplers", [](Runner &r) { r.flush_sample(); return r.sampler_obj; })
      .def("get_actions", [](Runner &r) { r.flush_sample(); return r.smp.get_actions(); })
      .def("get_action_masks", [](Runner &r) { return r.env.view(COG_BUF_SELECTED_MASKS); })
      .def("get_n_threads", [](Runner &r) { return r.n_threads; });

  py::object env_base = env, sampler_base = sampler, runner_base = runner;
  vec.def("get_vec_env", [env_base](uint64_t n) {
    return sized_class(env_base, "vec_cog_env_" + std::to_string(n), n, false);
  }, py::arg("n_envs"));
  vec.def("get_vec_sampler", [sampler_base](uint64_t n) {
    return sized_class(sampler_base, "vec_sampler_" + std::to_string(n), n, false);
  }, py::arg("n_envs"));
  vec.def("get_runner", [runner_base](uint64_t n) {
    return sized_class(runner_base, "vec_runner_" + std::to_string(n), n, true);
  }, py::arg("n_envs"));
  /* the README calls the getters at top level (README.md:120-127) */
  m.attr("get_vec_env") = vec.attr("get_vec_env");
  m.attr("get_vec_sampler") = vec.attr("get_vec_sampler");
  m.attr("get_runner") = vec.attr("get_runner");
  m.def("kernel_launches", []() { return cog_kernel_launches(); });
}
