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

  /* legacy (unversioned) capsule; extra protocol keywords (max_version, copy, dl_device) are accepted */
  py::capsule dlpack(py::object /*stream*/, py::kwargs /*unused*/) const {
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

/* device pointer of an argument that may be a DeviceBuffer, anything exposing
 * __cuda_array_interface__ (torch / cupy), or a host numpy array (returns nullptr and
 * sets *host) — the reference's step()/sample() take the buffer unchecked as well */
void *resolve_pointer(const py::object &obj, size_t need_bytes, const void **host) {
  *host = nullptr;
  if (py::isinstance<DeviceBuffer>(obj)) return obj.cast<const DeviceBuffer &>().ptr;
  if (py::hasattr(obj, "__cuda_array_interface__")) {
    py::dict d = obj.attr("__cuda_array_interface__");
    return (void *)py::cast<uintptr_t>(py::tuple(d["data"])[0]);
  }
  py::array arr = py::array::ensure(obj);
  if (!arr) throw py::type_error("expected a DeviceBuffer, a CUDA array or a numpy array");
  if (!(arr.flags() & py::array::c_style) || (size_t)arr.nbytes() < need_bytes)
    throw py::value_error("host array must be C-contiguous and hold one record per env");
  *host = arr.data();
  return nullptr;
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
  void step(const py::object &actions) {
    const void *host;
    void *dev = resolve_pointer(actions, n * COG_ACTION_SIZE, &host);
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
  void sample(const py::object &masks) {
    const void *host;
    void *dev = resolve_pointer(masks, n * COG_MASK_SIZE, &host);
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
  void sample() {
    void *m; size_t a, b;
    ck(cog_vec_buffer(env.own->h, COG_BUF_SELECTED_MASKS, &m, &a, &b));
    ck(cog_sampler_sample(smp.own->s, m));
  }
  void step() {
    void *a;
    ck(cog_sampler_actions(smp.own->s, &a));
    ck(cog_vec_step(env.own->h, a));
  }
  void sync() { ck(cog_sync(env.own->h)); }
  void step_sync() { step(); sync(); }
  void run(uint64_t k, bool merged) {
    ck(cog_vec_run_fused_ex(env.own->h, smp.own->s, k, merged ? COG_MASKS_MERGED : COG_MASKS_SELECTED));
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
  m.doc() = "City of Gold vectorised environments on B200 (drop-in for the reference's city_of_gold module)";

  py::enum_<cog_difficulty>(m, "Difficulty") /* src/pybind/common.cpp:32-36 */
      .value("EASY", COG_EASY).value("MEDIUM", COG_MEDIUM).value("HARD", COG_HARD).export_values();

  py::class_<DeviceBuffer>(m, "DeviceBuffer")
      .def_property_readonly("__cuda_array_interface__", &DeviceBuffer::cuda_array_interface)
      .def("__dlpack__", &DeviceBuffer::dlpack, py::arg("stream") = py::none())
      .def("__dlpack_device__", &DeviceBuffer::dlpack_device)
      .def("numpy", &DeviceBuffer::numpy, py::arg("out") = py::none(),
           "device -> host copy as a numpy array with the reference's structured dtype")
      .def_property_readonly("dtype", [](const DeviceBuffer &b) { return b.np_dtype; })
      .def_property_readonly("shape", [](const DeviceBuffer &b) -> py::tuple {
        if (b.which == COG_BUF_REWARDS && b.vec) return py::make_tuple(b.count, 4);
        return py::make_tuple(b.count);
      })
      .def_property_readonly("itemsize", [](const DeviceBuffer &b) { return b.itemsize; })
      .def_property_readonly("device_pointer", [](const DeviceBuffer &b) { return (uintptr_t)b.ptr; })
      .def("__len__", [](const DeviceBuffer &b) { return b.count; });

  auto vec = m.def_submodule("vec", "vectorised containers (reference src/pybind/vectorized.cpp)");

  py::class_<VecEnv> env(vec, "vec_cog_env");
  env.def(py::init<uint64_t, int, uint64_t>(), py::arg("n_envs"), py::arg("device") = 0,
          py::arg("env_index_offset") = 0)
      .def("reset", &VecEnv::reset_default)
      .def("reset", &VecEnv::reset, py::arg("seed"), py::arg("n_players"), py::arg("n_pieces"),
           py::arg("difficulty"), py::arg("max_steps"), py::arg("render"))
      .def("step", &VecEnv::step, py::arg("actions"))
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
      .def("sample", &VecSampler::sample, py::arg("action_mask"));

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
      .def("get_samplers", [](Runner &r) { return r.sampler_obj; })
      .def("get_actions", [](Runner &r) { return r.smp.get_actions(); })
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
