"""cog_sampler_sample_packed_noalloc (include/cog_b200.h): the no-allocation variant of
cog_sampler_sample_packed_host_io — staging in the library's static device array instead of buffers
allocated on first use.  No counterpart in the reference; checked against the existing entry point.

Collected last on purpose (the file name).  The test stands alone (its
own handles, nothing shared with other test modules) and also runs after the rest of the suite.

Integer / byte work: the tolerance is the one every parity test of this repository uses — bit-exact."""
import numpy as np
import pytest

from gym_eldorado_b200 import capi, layout


def _random_masks(rng, n):
    """packed masks with at least one legal index per head (bit 0 = the null action, as in the game)"""
    m = np.zeros(n, dtype=layout.packed_mask_dtype)
    m["play"] = (rng.integers(0, 1 << 22, n, dtype=np.uint32) & np.uint32(0x3FFFFF)) | np.uint32(1)
    m["remove"] = (rng.integers(0, 1 << 22, n, dtype=np.uint32) & np.uint32(0x3FFFFF)) | np.uint32(1)
    m["move_shop"] = (rng.integers(0, 1 << 26, n, dtype=np.uint32) & np.uint32(0x3FFFFFF)) | np.uint32(0x81)
    m["play_special"] = m["play"] & (rng.integers(0, 1 << 22, n, dtype=np.uint32) | np.uint32(1))
    return m


def _sample_pair(lib, n, seed, masks_of_call):
    """Two samplers with the same seed: one through the existing entry point, one through the new one,
    on the same masks, call after call; returns the number of calls compared."""
    a = capi.SamplerHandle(n, seed, 0, 0, None, lib)
    b = capi.SamplerHandle(n, seed, 0, 0, None, lib)
    try:
        for k, masks in enumerate(masks_of_call):
            out_a = np.full(n, 0xAB, dtype=np.uint8).repeat(8).view(layout.packed_action_dtype)
            out_b = np.full(n, 0xCD, dtype=np.uint8).repeat(8).view(layout.packed_action_dtype)
            a.sample_packed_host_io(masks, out_a)
            b.sample_packed_noalloc(masks, out_b)
            assert out_a.tobytes() == out_b.tobytes(), "packed actions differ on call %d" % k
            # the sampler's own ActionData buffer (what cog_vec_step consumes) and, through the next call,
            # the RNG streams
            assert a.read_actions().tobytes() == b.read_actions().tobytes(), "ActionData differs on call %d" % k
            assert np.all(out_b["play"] < 22) and np.all(out_b["get_from_shop"] < 19)
        return len(masks_of_call)
    finally:
        a.close()
        b.close()


def _game_masks(lib, n, seed, calls):
    """merged packed masks of a real game, one array per call: reset, then step with the sampled actions"""
    vec = capi.VecHandle(n, 0, 0, None, lib)
    smp = capi.SamplerHandle(n, seed + 1, 0, 0, None, lib)
    try:
        vec.reset(seed, 4, 3, layout.MEDIUM, 40)
        out = []
        act = np.zeros(n, dtype=layout.packed_action_dtype)
        for _ in range(calls):
            m = vec.packed_masks(merged=True)
            out.append(m.copy())
            smp.sample_packed_host_io(m, act)
            vec.step_packed_host_io(act)
        return out
    finally:
        vec.close()
        smp.close()


def _check_limit(lib):
    too_many = capi.SamplerHandle(capi.NOALLOC_MAX_ENVS + 1, 7, 0, 0, None, lib)
    try:
        n = capi.NOALLOC_MAX_ENVS + 1
        with pytest.raises(capi.CogError) as err:
            too_many.sample_packed_noalloc(np.zeros(n, dtype=layout.packed_mask_dtype),
                                           np.zeros(n, dtype=layout.packed_action_dtype))
        assert err.value.code == capi.COG_ERR_INVALID
        assert "COG_NOALLOC_MAX_ENVS" in str(err.value) and str(capi.NOALLOC_MAX_ENVS) in str(err.value)
    finally:
        too_many.close()


@pytest.mark.gpu
def test_zz_no_alloc_matches_existing_entry_point(cuda_lib):
    rng = np.random.default_rng(20261105)
    # masks of a real game (merged: every head has choices), five calls in a row on one pair of samplers
    assert _sample_pair(cuda_lib, 4096, 99, _game_masks(cuda_lib, 4096, 4242, 5)) == 5
    # the maximum size, random masks, two calls in a row; then a SMALLER sampler on the same device: the shared
    # array still holds the larger call's bytes and must not leak into the result
    big = [_random_masks(rng, capi.NOALLOC_MAX_ENVS) for _ in range(2)]
    assert _sample_pair(cuda_lib, capi.NOALLOC_MAX_ENVS, 12345, big) == 2
    assert _sample_pair(cuda_lib, 1000, 5, [_random_masks(rng, 1000) for _ in range(3)]) == 3
    _check_limit(cuda_lib)


def test_zz_no_alloc_emulated(emu):
    """the same host code under the warp emulator (tests/emu; CPU): argument checks, the split of the static
    array, two calls in a row"""
    rng = np.random.default_rng(7)
    assert _sample_pair(emu, 48, 3, [_random_masks(rng, 48) for _ in range(2)]) == 2
    assert _sample_pair(emu, 16, 3, _game_masks(emu, 16, 11, 2)) == 2
    _check_limit(emu)
