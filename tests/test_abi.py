"""CPU: the C-ABI library loads and exports every symbol include/cog_b200.h declares
(no compute calls without a GPU), and the layouts agree with the header."""
import os
import re

from gym_eldorado_b200 import capi, layout

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "cog_b200.h")).read()
    return set(re.findall(r"COG_API\s+[\w\s\*]+?\b(cog_\w+)\s*\(", text))


def test_header_and_binding_agree():
    assert _declared() == set(capi.SYMBOLS)


def test_library_exports_every_symbol():
    import __graft_entry__
    __graft_entry__.build()
    L = capi.load()  # raises if the .so is missing or lacks a symbol
    for name in _declared():
        assert hasattr(L, name)
    assert L.cog_kernel_launches() == 0


def test_record_sizes_match_header():
    text = open(os.path.join(ROOT, "include", "cog_b200.h")).read()
    defs = dict((k, int(v)) for k, v in re.findall(r"#define (COG_\w+) (\d+)", text))
    assert defs["COG_OBS_SIZE"] == layout.obs_dtype.itemsize == 17216
    assert defs["COG_MASK_SIZE"] == layout.action_mask_dtype.itemsize == 128
    assert defs["COG_ACTION_SIZE"] == layout.action_dtype.itemsize == 64
    assert defs["COG_INFO_SIZE"] == layout.info_dtype.itemsize == 192
    assert defs["COG_MASK_PAYLOAD"] == layout.ACTION_MASK_PAYLOAD == 92
    assert layout.player_data_dtype.itemsize == 256


def test_packed_masks_round_trip():
    """layout.pack_masks / unpack_masks (the host side of the packed host-I/O records, include/cog_b200.h):
    ActionMask -> 16-byte bit sets -> ActionMask is the identity on every payload byte, and the bit positions are
    the ones the header documents (bit k of `play` = play[k], move in bits 0-6, get_from_shop in bits 7-25)."""
    import numpy as np
    from hypothesis import given, settings, strategies as st
    from gym_eldorado_b200 import layout

    assert layout.packed_mask_dtype.itemsize == 16 and layout.packed_action_dtype.itemsize == 8

    @settings(max_examples=60, deadline=None)
    @given(st.integers(0, 2**32 - 1), st.integers(1, 40))
    def check(seed, n):
        rng = np.random.default_rng(seed)
        m = np.zeros(n, dtype=layout.action_mask_dtype)
        for name in ("play", "play_special", "remove", "move", "get_from_shop"):
            m[name] = rng.integers(0, 2, m[name].shape).astype(np.bool_)
        p = layout.pack_masks(m)
        back = layout.unpack_masks(p)
        assert layout.payload_bytes(back).tobytes() == layout.payload_bytes(m).tobytes()
        k = int(rng.integers(0, m["play"].shape[1]))
        assert np.array_equal((p["play"] >> np.uint32(k)) & 1, m["play"][:, k].astype(np.uint32))
        d = int(rng.integers(0, m["move"].shape[1]))
        assert np.array_equal((p["move_shop"] >> np.uint32(d)) & 1, m["move"][:, d].astype(np.uint32))
        c = int(rng.integers(0, m["get_from_shop"].shape[1]))
        assert np.array_equal((p["move_shop"] >> np.uint32(7 + c)) & 1, m["get_from_shop"][:, c].astype(np.uint32))
        assert not np.any(p["move_shop"] >> np.uint32(26)) and not np.any(p["play"] >> np.uint32(22))

    check()
