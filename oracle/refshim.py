"""ctypes loader for oracle/_ref/libcog_ref.so (the UNMODIFIED reference, compiled
by oracle/Makefile).  TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke()
and bench.py's reference / cpu_baseline leg may import this module.

The numpy views returned here alias the reference's own buffers, exactly like the
views its pybind module hands out (reference include/pybind/common.h:97-101).
"""
import ctypes
import os

import numpy as np

from gym_eldorado_b200 import layout

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libcog_ref.so")
SUPPORTED_N = (1, 16, 64, 256, 1024, 4096, 16384, 65536)

_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = ctypes.CDLL(LIB_PATH)
        L.ref_last_error.restype = ctypes.c_char_p
        L.ref_create.restype = ctypes.c_void_p
        L.ref_create.argtypes = [ctypes.c_uint64, ctypes.c_uint64]
        L.ref_destroy.argtypes = [ctypes.c_void_p]
        L.ref_reset.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint8,
                                ctypes.c_uint8, ctypes.c_int, ctypes.c_uint32]
        L.ref_reset_default.argtypes = [ctypes.c_void_p]
        L.ref_step.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.ref_sample.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.ref_buffer.restype = ctypes.c_void_p
        L.ref_buffer.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.ref_runner_create.argtypes = [ctypes.c_void_p, ctypes.c_uint64]
        L.ref_runner_destroy.argtypes = [ctypes.c_void_p]
        L.ref_time_run.restype = ctypes.c_double
        L.ref_time_run.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int]
        L.ref_time_sample.restype = ctypes.c_double
        L.ref_time_sample.argtypes = [ctypes.c_void_p, ctypes.c_uint64]
        L.ref_time_reset.restype = ctypes.c_double
        L.ref_time_reset.argtypes = [ctypes.c_void_p, ctypes.c_uint64]
        L.ref_hidden.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p]
        L.ref_sizeof.restype = ctypes.c_uint64
        L.ref_sizeof.argtypes = [ctypes.c_int]
        L.ref_map_generate.argtypes = [ctypes.c_uint32, ctypes.c_uint8, ctypes.c_int,
                                       ctypes.c_void_p, ctypes.c_void_p,
                                       ctypes.c_void_p, ctypes.c_uint32]
        _lib = L
    return _lib


def _view(ptr, dtype, shape):
    nbytes = int(np.prod(shape)) * dtype.itemsize
    buf = (ctypes.c_uint8 * nbytes).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype).reshape(shape)


class RefVec:
    """vec_cog_env<N> + vec_action_sampler<N> (+ optional ThreadedRunner<N>)."""

    def __init__(self, n_envs, sampler_seed):
        L = lib()
        self.n = n_envs
        self._h = L.ref_create(n_envs, sampler_seed)
        if not self._h:
            raise ValueError(L.ref_last_error().decode())
        b = lambda w: L.ref_buffer(self._h, w)
        self.observations = _view(b(0), layout.obs_dtype, (n_envs,))
        self.selected_action_masks = _view(b(1), layout.action_mask_dtype, (n_envs,))
        self.rewards = _view(b(2), np.dtype(np.float32), (n_envs, 4))
        self.dones = _view(b(3), np.dtype(np.bool_), (n_envs,))
        self.infos = _view(b(4), layout.info_dtype, (n_envs,))
        self.agent_selection = _view(b(5), np.dtype(np.uint8), (n_envs,))
        self.actions = _view(b(6), layout.action_dtype, (n_envs,))

    def close(self):
        if self._h:
            lib().ref_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise RuntimeError(lib().ref_last_error().decode())

    def reset(self, seed=None, n_players=4, n_pieces=3, difficulty=0,
              max_steps=100000):
        if seed is None:
            self._check(lib().ref_reset_default(self._h))
        else:
            self._check(lib().ref_reset(self._h, seed & 0xFFFFFFFF, n_players,
                                        n_pieces, int(difficulty), max_steps))

    def step(self, actions=None):
        a = self.actions if actions is None else actions
        assert a.dtype == layout.action_dtype and a.flags.c_contiguous
        self._check(lib().ref_step(self._h, a.ctypes.data))

    def sample(self, masks=None):
        m = self.selected_action_masks if masks is None else masks
        assert m.dtype == layout.action_mask_dtype and m.flags.c_contiguous
        lib().ref_sample(self._h, m.ctypes.data)

    def merged_masks(self):
        """Regime M: S.play/play_special, P[agent].move /
        get_from_shop, P[agent].remove AND hand>0."""
        n = self.n
        idx = np.arange(n)
        ag = self.agent_selection
        pd = self.observations["player_data"][idx, ag]
        m = np.zeros(n, dtype=layout.action_mask_dtype)
        S = self.selected_action_masks
        m["play"] = S["play"]
        m["play_special"] = S["play_special"]
        m["move"] = pd["action_mask"]["move"]
        m["get_from_shop"] = pd["action_mask"]["get_from_shop"]
        rem = pd["action_mask"]["remove"].copy()
        rem[:, 1:] &= pd["obs"]["hand"] > 0
        m["remove"] = rem
        return m

    def hidden(self, i):
        out = (ctypes.c_uint32 * 16)()
        lib().ref_hidden(self._h, i, out)
        return np.array(out, dtype=np.uint32).astype(np.int64)

    # ---- timing (bench.py reference arm) ----
    def runner(self, n_threads):
        self._check(lib().ref_runner_create(self._h, n_threads))

    def drop_runner(self):
        lib().ref_runner_destroy(self._h)

    def time_run(self, steps, mode):
        """mode: 0 sequential, 1 runner async, 2 runner sync. Returns seconds."""
        t = lib().ref_time_run(self._h, steps, mode)
        if t < 0:
            raise RuntimeError(lib().ref_last_error().decode())
        return t


    def time_sample(self, iters):
        """the sampler alone, `iters` calls on the current selected masks (benchmarks.py time_sample)"""
        t = lib().ref_time_sample(self._h, iters)
        if t < 0:
            raise RuntimeError(lib().ref_last_error().decode())
        return t

    def time_reset(self, iters):
        """`iters` sequential vec.reset() calls (benchmarks.py time_reset)"""
        t = lib().ref_time_reset(self._h, iters)
        if t < 0:
            raise RuntimeError(lib().ref_last_error().decode())
        return t


class RefSingle:
    """The reference's single cog_env (src/pybind/single_env.cpp:13-32) with shim-owned records."""

    def __init__(self, seed, n_players=4, n_pieces=3, difficulty=0, max_steps=100000):
        L = lib()
        L.ref_single_create.restype = ctypes.c_void_p
        L.ref_single_create.argtypes = [ctypes.c_uint32, ctypes.c_uint8, ctypes.c_uint8, ctypes.c_int, ctypes.c_uint32]
        L.ref_single_destroy.argtypes = [ctypes.c_void_p]
        L.ref_single_reset.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_uint32, ctypes.c_uint8, ctypes.c_uint8,
                                       ctypes.c_int, ctypes.c_uint32]
        L.ref_single_step.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.ref_single_buffer.restype = ctypes.c_void_p
        L.ref_single_buffer.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.ref_single_get.restype = ctypes.c_uint32
        L.ref_single_get.argtypes = [ctypes.c_void_p, ctypes.c_int]
        self._h = L.ref_single_create(seed & 0xFFFFFFFF, n_players, n_pieces, int(difficulty), max_steps)
        b = lambda w: L.ref_single_buffer(self._h, w)
        self.observations = _view(b(0), layout.obs_dtype, (1,))
        self.selected_action_masks = _view(b(1), layout.action_mask_dtype, (1,))
        self.rewards = _view(b(2), np.dtype(np.float32), (1, 4))
        self.infos = _view(b(4), layout.info_dtype, (1,))

    def close(self):
        if self._h:
            lib().ref_single_destroy(self._h)
            self._h = None

    def reset(self, seed=None, n_players=4, n_pieces=3, difficulty=0, max_steps=100000):
        rc = lib().ref_single_reset(self._h, 0 if seed is None else 1, (seed or 0) & 0xFFFFFFFF, n_players, n_pieces,
                                    int(difficulty), max_steps)
        if rc:
            raise RuntimeError(lib().ref_last_error().decode())

    def step(self, action):
        assert action.dtype == layout.action_dtype
        if lib().ref_single_step(self._h, action.ctypes.data):
            raise RuntimeError(lib().ref_last_error().decode())

    agent_selection = property(lambda self: lib().ref_single_get(self._h, 0))
    done = property(lambda self: bool(lib().ref_single_get(self._h, 1)))
    seed = property(lambda self: lib().ref_single_get(self._h, 2))
    n_players = property(lambda self: lib().ref_single_get(self._h, 3))


class RefSampler:
    """The reference's single action_sampler (include/sampler.h, bound at src/pybind/common.cpp:25-27)."""

    def __init__(self, seed):
        L = lib()
        L.ref_single_sampler_create.restype = ctypes.c_void_p
        L.ref_single_sampler_create.argtypes = [ctypes.c_uint32]
        L.ref_single_sampler_destroy.argtypes = [ctypes.c_void_p]
        L.ref_single_sampler_sample.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        self._h = L.ref_single_sampler_create(seed & 0xFFFFFFFF)

    def sample(self, mask):
        assert mask.dtype == layout.action_mask_dtype
        out = np.zeros(1, dtype=layout.action_dtype)
        lib().ref_single_sampler_sample(self._h, mask.ctypes.data, out.ctypes.data)
        return out

    def close(self):
        if self._h:
            lib().ref_single_sampler_destroy(self._h)
            self._h = None


def map_generate(seed, n_pieces, difficulty, rows=48, max_hexes=2048):
    """Reference Map::generate on a (rows,48,7) buffer; see ref_shim.cpp."""
    obs = np.zeros((rows, 48, 7), dtype=np.uint8)
    meta = np.zeros(8, dtype=np.uint32)
    xy = np.zeros((max_hexes, 2), dtype=np.int32)
    rc = lib().ref_map_generate(seed, n_pieces, int(difficulty), obs.ctypes.data,
                                meta.ctypes.data, xy.ctypes.data, max_hexes)
    if rc != 0:
        raise RuntimeError(lib().ref_last_error().decode())
    m = meta.astype(np.int32)
    return obs, m, xy[: int(meta[1])]
