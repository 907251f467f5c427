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
