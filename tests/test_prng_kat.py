"""Known-answer tests of the device PRNG (tests/kat/prng_kat.cu): the product's division-free
minstd_rand0 / uniform_int_distribution / sampler fast paths against the plain arithmetic of
libstdc++ (/usr/include/c++/13/bits/uniform_int_dist.h:332-340).

GPU: exhaustive — every engine state, every range 1..32 against every engine output.
CPU: the same kernels on sub-ranges (around the four guard states of the skip-4 jump, the ends
of the engine's range and a rejection boundary) under the warp emulator, which also checks the
checker."""
import ctypes
import os
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
KAT = os.path.join(HERE, "kat")
NAMES = ["minstd_next", "table", "quotient", "rejection", "skip4_value", "skip4_false_states", "skip4_false_wrong",
         "uniform_draw", "sample_action", "nth_set_bit_warp"]
M31 = 2147483647
GUARDS = (739806647, 671875339, 939811632, 1785685951)


def test_prng_kat_subranges_emulated():
    subprocess.check_call(["make", "-s", "-C", KAT, "emu"])
    L = ctypes.CDLL(os.path.join(KAT, "libcog_kat_emu.so"))
    L.cog_kat_prng_emu.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint]
    out = (ctypes.c_uint64 * 16)()
    ranges = [(0, 2000), (M31 - 2000, M31), (2147483645 // 7 * 7 - 200, 2147483645 // 7 * 7 + 200)]
    ranges += [(g - 100, g + 100) for g in GUARDS]
    false_states = 0
    for v0, v1 in ranges:
        assert L.cog_kat_prng_emu(out, 16, v0, v1, 1500) == 0
        got = dict(zip(NAMES, list(out)))
        false_states += got.pop("skip4_false_states")
        assert not any(got.values()), ((v0, v1), got)
    assert false_states == 4  # exactly the four guard states fall back to the step-by-step path


@pytest.mark.gpu
def test_prng_kat_exhaustive_gpu():
    """oracle: libstdc++'s arithmetic restated with 64-bit modulo and divisions, on the device"""
    import torch
    assert torch.cuda.is_available()
    path = os.path.join(KAT, "libcog_kat.so")
    assert os.path.exists(path), "tests/kat/libcog_kat.so is missing: run __graft_entry__.build()"
    L = ctypes.CDLL(path)
    L.cog_kat_prng.argtypes = [ctypes.c_void_p, ctypes.c_int]
    out = (ctypes.c_uint64 * 16)()
    assert L.cog_kat_prng(out, 16) == 0
    got = dict(zip(NAMES, list(out)))
    assert got.pop("skip4_false_states") == 4, got
    assert not any(got.values()), got
