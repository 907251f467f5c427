"""ctypes binding of the C ABI in include/cog_b200.h (libcog_b200.so).

This is the thin layer every Python-side caller goes through; the pybind11 module
(csrc/pybind_module.cpp) binds the same symbols from C++.  There is no fallback: if
the CUDA library is missing or cannot be loaded, importing the handles below raises.
"""
import ctypes
import os

import numpy as np

from . import layout

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcog_b200.so")

COG_OK = 0
NOALLOC_MAX_ENVS = 65536  # COG_NOALLOC_MAX_ENVS (include/cog_b200.h)
COG_ERR_INVALID, COG_ERR_MAP_FAILURE, COG_ERR_MAP_BOUNDS, COG_ERR_CUDA, COG_ERR_NOT_RESET = -1, -2, -3, -4, -5
STAT_NAMES = ("env_steps", "episodes", "truncations", "total_length", "wins_seat0", "wins_seat1",
              "wins_seat2", "wins_seat3", "travelled_hexes", "cards_added", "resets", "reserved")

# every symbol include/cog_b200.h declares: name -> (restype, argtypes)
_vp, _u64, _u32, _u8, _int, _sz = (ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_uint8,
                                    ctypes.c_int, ctypes.c_size_t)
_P = ctypes.POINTER
SYMBOLS = {
    "cog_last_error": (ctypes.c_char_p, []),
    "cog_device_count": (_int, [_P(_int)]),
    "cog_vec_create": (_int, [_u64, _u64, _int, _vp, _P(_vp)]),
    "cog_vec_destroy": (_int, [_vp]),
    "cog_vec_reset": (_int, [_vp, _u32, _u8, _u8, _int, _u32, _int]),
    "cog_vec_reset_default": (_int, [_vp]),
    "cog_vec_step": (_int, [_vp, _vp]),
    "cog_vec_step_host": (_int, [_vp, _vp]),
    "cog_vec_step_sampled": (_int, [_vp, _vp]),
    "cog_vec_set_autoreset": (_int, [_vp, _int]),
    "cog_vec_buffer": (_int, [_vp, _int, _P(_vp), _P(_sz), _P(_sz)]),
    "cog_vec_read": (_int, [_vp, _int, _vp, _sz]),
    "cog_vec_num_envs": (_int, [_vp, _P(_u64)]),
    "cog_sampler_create": (_int, [_u64, _u64, _u32, _int, _vp, _P(_vp)]),
    "cog_sampler_destroy": (_int, [_vp]),
    "cog_sampler_sample": (_int, [_vp, _vp]),
    "cog_sampler_sample_host": (_int, [_vp, _vp]),
    "cog_sampler_actions": (_int, [_vp, _P(_vp)]),
    "cog_sampler_read_actions": (_int, [_vp, _vp, _sz]),
    "cog_sampler_sample_host_io": (_int, [_vp, _vp, _vp]),
    "cog_vec_step_host_io": (_int, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "cog_io_path": (_int, [_vp, _vp, _P(_int), _P(_int)]),
    "cog_vec_wait_stream": (_int, [_vp, _vp]),
    "cog_vec_release_to_stream": (_int, [_vp, _vp]),
    "cog_sampler_wait_stream": (_int, [_vp, _vp]),
    "cog_sampler_release_to_stream": (_int, [_vp, _vp]),
    "cog_vec_stream": (_int, [_vp, _P(_vp)]),
    "cog_sampler_stream": (_int, [_vp, _P(_vp)]),
    "cog_sampler_sample_packed_host_io": (_int, [_vp, _vp, _vp]),
    "cog_sampler_sample_packed_noalloc": (_int, [_vp, _vp, _vp]),
    "cog_vec_step_packed_host_io": (_int, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "cog_vec_packed_masks": (_int, [_vp, _vp, _int, _int]),
    "cog_vec_run_fused": (_int, [_vp, _vp, _u64]),
    "cog_vec_run_fused_ex": (_int, [_vp, _vp, _u64, _int]),
    "cog_vec_merged_masks": (_int, [_vp, _vp]),
    "cog_vec_compact_obs": (_int, [_vp, _vp, _int]),
    "cog_vec_rollout_attach": (_int, [_vp, _u32]),
    "cog_vec_rollout_info": (_int, [_vp, _P(_vp), _P(_u32), _P(_u64)]),
    "cog_vec_stats": (_int, [_vp, _P(_u64), _sz]),
    "cog_vec_stats_reset": (_int, [_vp]),
    "cog_sync": (_int, [_vp]),
    "cog_kernel_launches": (_u64, []),
    "cog_vec_last_kernel_ms": (_int, [_vp, _P(ctypes.c_float)]),
}


class CogError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(message)
        self.code = code


class MapGenerationError(CogError):
    """generate_map_failure (reference include/map.h:13-15 -> Python RuntimeError)."""


def load(path=LIB_PATH):
    """dlopen the library and declare every prototype. Raises if it is absent."""
    if not os.path.exists(path):
        raise ImportError(
            "%s is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU fallback." % path)
    lib = ctypes.CDLL(path)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    return lib


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = load()
    return _lib


def check(L, rc):
    if rc == COG_OK:
        return
    msg = L.cog_last_error().decode()
    if rc in (COG_ERR_MAP_FAILURE, COG_ERR_MAP_BOUNDS):
        raise MapGenerationError(rc, msg)
    raise CogError(rc, msg)


class VecHandle:
    """cog_vec: one GPU's slice of vec_cog_env (reference include/vec_environment.h:12-81)."""

    def __init__(self, n_envs, env_index_offset=0, device=0, stream=None, library=None):
        self.L = library or lib()
        self.n = int(n_envs)
        self.device = device
        self.offset = env_index_offset
        h = _vp()
        check(self.L, self.L.cog_vec_create(self.n, env_index_offset, device, stream, ctypes.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            self.L.cog_vec_destroy(self._h)
            self._h = None

    __del__ = close

    def reset(self, seed=None, n_players=4, n_pieces=3, difficulty=0, max_steps=100000, render=False):
        if seed is None:
            check(self.L, self.L.cog_vec_reset_default(self._h))
        else:
            check(self.L, self.L.cog_vec_reset(self._h, int(seed) & 0xFFFFFFFF, n_players, n_pieces,
                                               int(difficulty), max_steps, int(bool(render))))

    def step_device(self, d_actions_ptr):
        check(self.L, self.L.cog_vec_step(self._h, d_actions_ptr))

    def set_autoreset(self, on):
        """False = single-env semantics: a finished episode keeps its terminal state, steps are ignored"""
        check(self.L, self.L.cog_vec_set_autoreset(self._h, 1 if on else 0))

    def step_sampled(self, sampler):
        """sample(selected_action_masks); step(sampled actions) as one kernel."""
        check(self.L, self.L.cog_vec_step_sampled(self._h, sampler._s))

    def step_host(self, actions):
        assert actions.dtype == layout.action_dtype and actions.flags.c_contiguous and len(actions) == self.n
        check(self.L, self.L.cog_vec_step_host(self._h, actions.ctypes.data))

    def step_host_io(self, actions, masks_out=None, rewards_out=None, dones_out=None, agents_out=None):
        """step(host actions) + read-back of the selected masks / rewards / dones / agent_selection into
        host arrays, copies and kernels pipelined chunk by chunk (pinned memory overlaps)."""
        assert actions.dtype == layout.action_dtype and actions.flags.c_contiguous and len(actions) == self.n
        ptr = lambda a: None if a is None else a.ctypes.data
        check(self.L, self.L.cog_vec_step_host_io(self._h, actions.ctypes.data, ptr(masks_out), ptr(rewards_out),
                                                  ptr(dones_out), ptr(agents_out)))

    def io_path(self, sampler=None):
        """(env path, sampler path) of the *_host_io entry points for pinned arrays:
        "zero_copy", "dma_pipeline" or "measuring"."""
        a, b = _int(-1), _int(-1)
        check(self.L, self.L.cog_io_path(self._h, sampler._s if sampler is not None else None,
                                         ctypes.byref(a), ctypes.byref(b)))
        name = {1: "zero_copy", 0: "dma_pipeline", -1: "measuring"}
        return name[a.value], name[b.value]

    def step_packed_host_io(self, pactions, pmasks_out=None, rewards_out=None, dones_out=None, agents_out=None):
        """step(packed host actions, layout.packed_action_dtype) + read-back of the packed selected masks
        (layout.packed_mask_dtype) / rewards / dones / agent_selection into host arrays."""
        assert pactions.dtype == layout.packed_action_dtype and pactions.flags.c_contiguous and len(pactions) == self.n
        ptr = lambda a: None if a is None else a.ctypes.data
        check(self.L, self.L.cog_vec_step_packed_host_io(self._h, pactions.ctypes.data, ptr(pmasks_out), ptr(rewards_out),
                                                         ptr(dones_out), ptr(agents_out)))

    def packed_masks(self, merged=False, out=None):
        """current selected (or merged) masks in packed form as a host array"""
        if out is None:
            out = np.empty(self.n, dtype=layout.packed_mask_dtype)
        check(self.L, self.L.cog_vec_packed_masks(self._h, out.ctypes.data, 1, 1 if merged else 0))
        return out

    def buffer(self, which):
        """(device pointer, itemsize, count) of an exported buffer."""
        p, item, cnt = _vp(), _sz(), _sz()
        check(self.L, self.L.cog_vec_buffer(self._h, which, ctypes.byref(p), ctypes.byref(item), ctypes.byref(cnt)))
        return p.value, item.value, cnt.value

    def read(self, which, out=None):
        """Device -> host copy of an exported buffer as a structured numpy array."""
        dt = layout.BUFFER_DTYPES[which]
        if out is None:
            out = np.empty(self.n, dtype=dt)
        check(self.L, self.L.cog_vec_read(self._h, which, out.ctypes.data, out.nbytes))
        return out

    def run_fused(self, sampler, k_steps, merged=False):
        """k_steps x {sample; step} on the device; merged=True samples from the merged legal mask."""
        check(self.L, self.L.cog_vec_run_fused_ex(self._h, sampler._s, int(k_steps), 1 if merged else 0))

    def merged_masks_device(self, d_out_ptr):
        check(self.L, self.L.cog_vec_merged_masks(self._h, d_out_ptr))

    def compact_obs_device(self, d_out_ptr=None, merged=True):
        """One 256-byte compact record per env (layout.compact_dtype) to device memory, or to
        the handle's BUF_COMPACT buffer when d_out_ptr is None."""
        check(self.L, self.L.cog_vec_compact_obs(self._h, d_out_ptr, 1 if merged else 0))

    def compact_obs(self, merged=True):
        """Compact records of the current state as a host array."""
        self.compact_obs_device(None, merged)
        return self.read(layout.BUF_COMPACT)

    def rollout_attach(self, n_slots):
        """Every later run_fused appends one compact record per env-step to a device ring of
        n_slots time slots; 0 detaches."""
        check(self.L, self.L.cog_vec_rollout_attach(self._h, int(n_slots)))

    def rollout_info(self):
        """(device pointer, n_slots, steps recorded per env since attach)"""
        p, t, c = _vp(), _u32(), _u64()
        check(self.L, self.L.cog_vec_rollout_info(self._h, ctypes.byref(p), ctypes.byref(t), ctypes.byref(c)))
        return p.value, t.value, c.value

    def read_rollout(self):
        """The whole ring as a host array [n][n_slots] of compact records (env-major)."""
        _, t, _ = self.rollout_info()
        out = np.empty((self.n, t), dtype=layout.compact_dtype)
        check(self.L, self.L.cog_vec_read(self._h, layout.BUF_ROLLOUT, out.ctypes.data, out.nbytes))
        return out

    def stats(self):
        out = (_u64 * len(STAT_NAMES))()
        check(self.L, self.L.cog_vec_stats(self._h, out, len(STAT_NAMES)))
        return dict(zip(STAT_NAMES, (int(x) for x in out)))

    def stats_reset(self):
        check(self.L, self.L.cog_vec_stats_reset(self._h))

    def sync(self):
        check(self.L, self.L.cog_sync(self._h))

    def last_kernel_ms(self):
        ms = ctypes.c_float()
        check(self.L, self.L.cog_vec_last_kernel_ms(self._h, ctypes.byref(ms)))
        return ms.value


class SamplerHandle:
    """cog_sampler: vec_action_sampler (reference include/vec_sampler.h:7-27)."""

    def __init__(self, n_envs, seed, env_index_offset=0, device=0, stream=None, library=None):
        self.L = library or lib()
        self.n = int(n_envs)
        s = _vp()
        check(self.L, self.L.cog_sampler_create(self.n, env_index_offset, int(seed) & 0xFFFFFFFF, device,
                                                stream, ctypes.byref(s)))
        self._s = s

    def close(self):
        if getattr(self, "_s", None):
            self.L.cog_sampler_destroy(self._s)
            self._s = None

    __del__ = close

    def sample_device(self, d_masks_ptr):
        check(self.L, self.L.cog_sampler_sample(self._s, d_masks_ptr))

    def sample_host(self, masks):
        assert masks.dtype == layout.action_mask_dtype and masks.flags.c_contiguous and len(masks) == self.n
        check(self.L, self.L.cog_sampler_sample_host(self._s, masks.ctypes.data))

    def sample_host_io(self, masks, actions_out):
        """sample(host masks) with the sampled actions copied back into `actions_out`, pipelined."""
        assert masks.dtype == layout.action_mask_dtype and masks.flags.c_contiguous and len(masks) == self.n
        assert actions_out.dtype == layout.action_dtype and actions_out.flags.c_contiguous and len(actions_out) == self.n
        check(self.L, self.L.cog_sampler_sample_host_io(self._s, masks.ctypes.data, actions_out.ctypes.data))

    def sample_packed_host_io(self, pmasks, pactions_out):
        """sample(packed host masks) with the sampled actions copied back in packed form"""
        assert pmasks.dtype == layout.packed_mask_dtype and pmasks.flags.c_contiguous and len(pmasks) == self.n
        assert pactions_out.dtype == layout.packed_action_dtype and len(pactions_out) == self.n
        check(self.L, self.L.cog_sampler_sample_packed_host_io(self._s, pmasks.ctypes.data, pactions_out.ctypes.data))

    def sample_packed_noalloc(self, pmasks, pactions_out):
        """sample_packed_host_io through the library's static device scratch: the call allocates nothing
        (at most NOALLOC_MAX_ENVS envs; all calls on a device share the scratch)"""
        assert pmasks.dtype == layout.packed_mask_dtype and pmasks.flags.c_contiguous and len(pmasks) == self.n
        assert pactions_out.dtype == layout.packed_action_dtype and len(pactions_out) == self.n
        check(self.L, self.L.cog_sampler_sample_packed_noalloc(self._s, pmasks.ctypes.data, pactions_out.ctypes.data))

    def actions_ptr(self):
        p = _vp()
        check(self.L, self.L.cog_sampler_actions(self._s, ctypes.byref(p)))
        return p.value

    def read_actions(self, out=None):
        if out is None:
            out = np.empty(self.n, dtype=layout.action_dtype)
        check(self.L, self.L.cog_sampler_read_actions(self._s, out.ctypes.data, out.nbytes))
        return out
