"""Exported record layouts (the wire format of the drop-in boundary).

These numpy structured dtypes restate the reference's POD records byte for byte
(reference include/api.h:67-161, dtype registration src/pybind/common.cpp:8-20;
offsets as measured on x86-64 with a 64-byte interference size).
The same numbers are `static_assert`ed on the device side in csrc/cog_layout.h.

Only payload bytes are meaningful: record padding is indeterminate in the
reference and is never compared.
"""
import numpy as np

GRIDSIZE = 48
N_MAP_FEATURES = 7
N_CARDTYPES = 21
N_BUYABLETYPES = 18
N_DIRECTIONS = 7
N_RESOURCETYPES = 3
MAX_N_PLAYERS = 4

# Difficulty (include/constants.h:11)
EASY, MEDIUM, HARD = 0, 1, 2
# TurnPhase (include/constants.h:76)
INACTIVE, MOVEMENT, BUYING = 0, 1, 2

deck_obs_dtype = np.dtype({
    "names": ["draw", "hand", "active", "played", "discard"],
    "formats": [(np.uint8, (N_CARDTYPES,))] * 5,
    "offsets": [0, 21, 42, 63, 84],
    "itemsize": 105,
})

action_mask_dtype = np.dtype({
    "names": ["play", "play_special", "remove", "move", "get_from_shop"],
    "formats": [(np.bool_, (N_CARDTYPES + 1,)), (np.bool_, (N_CARDTYPES + 1,)),
                (np.bool_, (N_CARDTYPES + 1,)), (np.bool_, (N_DIRECTIONS,)),
                (np.bool_, (N_BUYABLETYPES + 1,))],
    "offsets": [0, 22, 44, 66, 73],
    "itemsize": 128,
})
ACTION_MASK_PAYLOAD = 92

player_data_dtype = np.dtype({
    "names": ["obs", "action_mask"],
    "formats": [deck_obs_dtype, action_mask_dtype],
    "offsets": [0, 128],
    "itemsize": 256,
})

shared_obs_dtype = np.dtype({
    "names": ["map", "phase", "current_resources", "shop"],
    "formats": [(np.uint8, (GRIDSIZE, GRIDSIZE, N_MAP_FEATURES)), np.uint8,
                (np.float32, (N_RESOURCETYPES,)), (np.uint8, (N_BUYABLETYPES,))],
    "offsets": [0, 16128, 16132, 16144],
    "itemsize": 16164,
})

obs_dtype = np.dtype({
    "names": ["shared", "player_data"],
    "formats": [shared_obs_dtype, (player_data_dtype, (MAX_N_PLAYERS,))],
    "offsets": [0, 16192],
    "itemsize": 17216,
})

action_dtype = np.dtype({
    "names": ["play", "play_special", "remove", "move", "get_from_shop"],
    "formats": [np.uint8] * 5,
    "offsets": [0, 1, 2, 3, 4],
    "itemsize": 64,
})
ACTION_PAYLOAD = 5

agent_info_dtype = np.dtype({
    "names": ["steps_taken", "returns", "travelled_hexes", "cards_added",
              "cards_removed", "n_machete_uses", "n_paddle_uses", "n_coin_uses",
              "n_card_uses"],
    "formats": [np.uint8, np.float32, np.uint32, np.uint8, np.uint8, np.uint32,
                np.uint32, np.uint32, np.uint32],
    "offsets": [0, 4, 8, 12, 13, 16, 20, 24, 28],
    "itemsize": 32,
})

info_dtype = np.dtype({
    "names": ["total_length", "agent_infos"],
    "formats": [np.uint32, (agent_info_dtype, (MAX_N_PLAYERS,))],
    "offsets": [0, 4],
    "itemsize": 192,
})

# Not in the reference: the compact record (include/cog_b200.h, cog_vec_compact_obs) — the
# current agent's view of one env in 256 bytes, and the element of the rollout ring.
compact_card_dtype = np.dtype({
    "names": ["draw", "hand", "active", "played", "discard", "shop"],
    "formats": [np.uint8] * 6,
    "offsets": [0, 1, 2, 3, 4, 5],
    "itemsize": 8,
})
compact_loc_dtype = np.dtype({"names": ["row", "col"], "formats": [np.uint8, np.uint8], "offsets": [0, 1],
                              "itemsize": 2})
compact_dtype = np.dtype({
    "names": ["agent_selection", "phase", "resources_x2", "action", "mask_play", "mask_remove",
              "mask_move_shop", "mask_play_special", "loc", "turn_counter", "t", "done", "filled", "rewards",
              "cards"],
    "formats": [np.uint8, np.uint8, (np.uint16, (N_RESOURCETYPES,)), (np.uint8, (5,)), np.uint32, np.uint32,
                np.uint32, np.uint32, (compact_loc_dtype, (MAX_N_PLAYERS,)), np.uint32, np.uint32, np.uint8,
                np.uint8, (np.float32, (MAX_N_PLAYERS,)), (compact_card_dtype, (N_CARDTYPES,))],
    "offsets": [0, 1, 2, 8, 16, 20, 24, 28, 32, 40, 44, 48, 49, 64, 80],
    "itemsize": 256,
})
GRID_BYTES = GRIDSIZE * GRIDSIZE

# Not in the reference: the packed host-I/O records (include/cog_b200.h, cog_*_packed_host_io):
# an ActionMask as four bit sets (bit k of a word = bool k of the head) and the five ActionData
# bytes in eight.
packed_mask_dtype = np.dtype({
    "names": ["play", "remove", "move_shop", "play_special"],
    "formats": [np.uint32] * 4, "offsets": [0, 4, 8, 12], "itemsize": 16,
})
packed_action_dtype = np.dtype({
    "names": ["play", "play_special", "remove", "move", "get_from_shop"],
    "formats": [np.uint8] * 5, "offsets": [0, 1, 2, 3, 4], "itemsize": 8,
})


def _bits(a):
    a = np.asarray(a).astype(np.uint64)
    return (a << np.arange(a.shape[1], dtype=np.uint64)).sum(axis=1).astype(np.uint32)


def pack_masks(masks):
    """ActionMask[n] (bool per byte) -> packed_mask_dtype[n]"""
    out = np.zeros(len(masks), dtype=packed_mask_dtype)
    out["play"] = _bits(masks["play"])
    out["remove"] = _bits(masks["remove"])
    out["move_shop"] = _bits(masks["move"]) | (_bits(masks["get_from_shop"]) << np.uint32(7))
    out["play_special"] = _bits(masks["play_special"])
    return out


def unpack_masks(packed):
    """packed_mask_dtype[n] -> ActionMask[n]"""
    out = np.zeros(len(packed), dtype=action_mask_dtype)
    bit = lambda w, n, sh=0: ((w[:, None] >> (np.arange(n, dtype=np.uint32) + np.uint32(sh))) & 1).astype(np.bool_)
    out["play"] = bit(packed["play"], N_CARDTYPES + 1)
    out["remove"] = bit(packed["remove"], N_CARDTYPES + 1)
    out["play_special"] = bit(packed["play_special"], N_CARDTYPES + 1)
    out["move"] = bit(packed["move_shop"], N_DIRECTIONS)
    out["get_from_shop"] = bit(packed["move_shop"], N_BUYABLETYPES + 1, 7)
    return out


def grid_to_map(grid):
    """uint8[..., 48*48] hex codes (BUF_GRID) -> uint8[..., 48, 48, 7] map observation
    (Map::finalize, reference src/map.cpp:389-405; feature 0 is never written)."""
    g = np.asarray(grid, dtype=np.uint8).reshape(grid.shape[:-1] + (GRIDSIZE, GRIDSIZE))
    out = np.zeros(g.shape + (N_MAP_FEATURES,), dtype=np.uint8)
    f = g & 7
    nreq = (g >> 3) & 7
    for k in range(1, 7):
        out[..., k] = np.where(f == k, nreq, 0)
    out[..., 6] |= (g >> 6) & 1
    return out


# Which exported buffer is which (shared with include/cog_b200.h: enum cog_buffer)
(BUF_OBS, BUF_SELECTED_MASKS, BUF_REWARDS, BUF_DONES, BUF_INFOS, BUF_AGENT_SELECTION, BUF_MERGED_MASKS,
 BUF_GRID, BUF_COMPACT, BUF_ROLLOUT) = range(10)

BUFFER_DTYPES = {
    BUF_OBS: obs_dtype,
    BUF_SELECTED_MASKS: action_mask_dtype,
    BUF_REWARDS: np.dtype((np.float32, (MAX_N_PLAYERS,))),
    BUF_DONES: np.dtype(np.bool_),
    BUF_INFOS: info_dtype,
    BUF_AGENT_SELECTION: np.dtype(np.uint8),
    BUF_MERGED_MASKS: action_mask_dtype,
    BUF_GRID: np.dtype((np.uint8, (GRID_BYTES,))),
    BUF_COMPACT: compact_dtype,
    BUF_ROLLOUT: compact_dtype,
}


def payload_bytes(arr):
    """Concatenate the payload (non-padding) bytes of a structured array.

    Returns a uint8 array of shape (len(arr), n_payload_bytes); used by the
    parity tests so padding never takes part in a comparison.
    """
    arr = np.ascontiguousarray(arr)
    n = arr.shape[0] if arr.ndim else 1
    raw = arr.reshape(n, -1) if arr.dtype.names is None else arr.reshape(n)
    if arr.dtype.names is None:
        return raw.view(np.uint8).reshape(n, -1)
    cols = []

    def walk(dt, base, view):
        if dt.names is None:
            sub = dt.subdtype
            if sub is not None:
                elem, shape = sub
                count = int(np.prod(shape))
                if elem.names is None:
                    cols.append(view[:, base:base + elem.itemsize * count])
                else:
                    for k in range(count):
                        walk(elem, base + k * elem.itemsize, view)
            else:
                cols.append(view[:, base:base + dt.itemsize])
            return
        for name in dt.names:
            fdt, off = dt.fields[name][:2]
            walk(fdt, base + off, view)

    bytes_view = raw.view(np.uint8).reshape(n, arr.dtype.itemsize)
    walk(arr.dtype, 0, bytes_view)
    return np.concatenate(cols, axis=1)
