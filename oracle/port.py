"""ctypes loader for oracle/libcog_oracle.so (the plain-C restatement,
oracle/cog_oracle.c).  TEST INFRASTRUCTURE: only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline leg may import this module.

`OracleVec` has the same surface as oracle.refshim.RefVec so the two can be
driven side by side.
"""
import ctypes
import os
import subprocess

import numpy as np

from gym_eldorado_b200 import layout

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcog_oracle.so")
_lib = None


def build(force=False):
    src = os.path.join(_HERE, "cog_oracle.c")
    if force or not os.path.exists(LIB_PATH) or \
            os.path.getmtime(LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "oracle"])


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(LIB_PATH)
        L.orc_create.restype = ctypes.c_void_p
        L.orc_create.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint64]
        L.orc_destroy.argtypes = [ctypes.c_void_p]
        L.orc_reset.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.c_uint8,
                                ctypes.c_uint8, ctypes.c_int, ctypes.c_uint32]
        L.orc_reset_again.argtypes = [ctypes.c_void_p]
        L.orc_step.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.orc_sample.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.orc_set_autoreset.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.orc_buffer.restype = ctypes.c_void_p
        L.orc_buffer.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.orc_time_run.restype = ctypes.c_double
        L.orc_time_run.argtypes = [ctypes.c_void_p, ctypes.c_uint64]
        L.orc_hidden.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_void_p]
        L.orc_map_generate.argtypes = [ctypes.c_uint32, ctypes.c_uint8, ctypes.c_int,
                                       ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p,
                                       ctypes.c_void_p, ctypes.c_uint32]
        L.orc_minstd_seed.restype = ctypes.c_uint32
        L.orc_minstd_seed.argtypes = [ctypes.c_uint64]
        L.orc_minstd_next.restype = ctypes.c_uint32
        L.orc_minstd_next.argtypes = [ctypes.c_void_p]
        L.orc_uniform.restype = ctypes.c_uint64
        L.orc_uniform.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64]
        _lib = L
    return _lib


def _view(ptr, dtype, shape):
    nbytes = int(np.prod(shape)) * dtype.itemsize
    buf = (ctypes.c_uint8 * nbytes).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype).reshape(shape)


class MapGenerationError(RuntimeError):
    pass


class OracleVec:
    def __init__(self, n_envs, sampler_seed, env_index_offset=0):
        L = lib()
        self.n = n_envs
        self._h = L.orc_create(n_envs, env_index_offset, sampler_seed)
        b = lambda w: L.orc_buffer(self._h, w)
        self.observations = _view(b(0), layout.obs_dtype, (n_envs,))
        self.selected_action_masks = _view(b(1), layout.action_mask_dtype, (n_envs,))
        self.rewards = _view(b(2), np.dtype(np.float32), (n_envs, 4))
        self.dones = _view(b(3), np.dtype(np.bool_), (n_envs,))
        self.infos = _view(b(4), layout.info_dtype, (n_envs,))
        self.agent_selection = _view(b(5), np.dtype(np.uint8), (n_envs,))
        self.actions = _view(b(6), layout.action_dtype, (n_envs,))

    def close(self):
        if self._h:
            lib().orc_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def _check(rc):
        if rc == -2:
            raise MapGenerationError(
                "Failed to generate map in specified maximum number of attempts")
        if rc == -3:
            raise MapGenerationError("board exceeds the 48x48 grid (UB in the reference)")
        if rc != 0:
            raise RuntimeError("oracle error %d" % rc)

    def reset(self, seed=None, n_players=4, n_pieces=3, difficulty=0, max_steps=100000):
        if seed is None:
            self._check(lib().orc_reset_again(self._h))
        else:
            self._check(lib().orc_reset(self._h, seed & 0xFFFFFFFF, n_players, n_pieces,
                                        int(difficulty), max_steps))

    def step(self, actions=None):
        a = self.actions if actions is None else actions
        assert a.dtype == layout.action_dtype and a.flags.c_contiguous
        self._check(lib().orc_step(self._h, a.ctypes.data))

    def sample(self, masks=None):
        m = self.selected_action_masks if masks is None else masks
        assert m.dtype == layout.action_mask_dtype and m.flags.c_contiguous
        lib().orc_sample(self._h, m.ctypes.data)

    def set_autoreset(self, on):
        lib().orc_set_autoreset(self._h, 1 if on else 0)

    def merged_masks(self):
        from oracle.refshim import RefVec
        return RefVec.merged_masks(self)

    def hidden(self, i):
        out = (ctypes.c_uint32 * 16)()
        lib().orc_hidden(self._h, i, out)
        return np.array(out, dtype=np.uint32).astype(np.int64)

    def time_run(self, steps):
        return lib().orc_time_run(self._h, steps)


def map_generate(seed, n_pieces, difficulty, rows=48, max_hexes=2048):
    obs = np.zeros((rows, 48, 7), dtype=np.uint8)
    meta = np.zeros(8, dtype=np.uint32)
    xy = np.zeros((max_hexes, 2), dtype=np.int32)
    rc = lib().orc_map_generate(seed, n_pieces, int(difficulty), obs.ctypes.data, rows,
                                meta.ctypes.data, xy.ctypes.data, max_hexes)
    OracleVec._check(rc)
    return obs, meta.astype(np.int32), xy[: int(meta[1])]


def minstd_stream(seed, n):
    st = ctypes.c_uint32(lib().orc_minstd_seed(seed))
    return [lib().orc_minstd_next(ctypes.byref(st)) for _ in range(n)]
