/* cog_layout.h — byte layouts shared by the kernels and the host C ABI.
 *
 * (1) Exported records: the reference's PODs, byte for byte (reference
 *     include/api.h:67-161, measured offsets in SURVEY App. B).
 * (2) The private per-env state record in HBM ("hidden" state):
 *     COG_HID_LINES lines of 32 words.  One warp owns one env and lane l loads
 *     word l of every line, so each line is ONE fully coalesced 128-byte
 *     transaction and lands directly in the register that lane keeps it in.
 */
#ifndef COG_LAYOUT_H
#define COG_LAYOUT_H
#include <stdint.h>

/* ---- exported records ---------------------------------------------------- */
enum {
  COGL_OBS_SIZE = 17216,
  COGL_MAP_BYTES = 48 * 48 * 7, /* 16128 */
  COGL_OBS_PHASE = 16128,
  COGL_OBS_RES = 16132,  /* float[3] machete, paddle, coin */
  COGL_OBS_SHOP = 16144, /* u8[18] */
  COGL_OBS_PLAYER = 16192,
  COGL_PLAYER_STRIDE = 256,
  COGL_DK_DRAW = 0, COGL_DK_HAND = 21, COGL_DK_ACTIVE = 42, COGL_DK_PLAYED = 63, COGL_DK_DISCARD = 84,
  COGL_PD_MASK = 128,
  COGL_MASK_SIZE = 128,
  COGL_AM_PLAY = 0, COGL_AM_SPECIAL = 22, COGL_AM_REMOVE = 44, COGL_AM_MOVE = 66, COGL_AM_SHOP = 73,
  COGL_AM_PAYLOAD = 92,
  COGL_ACTION_SIZE = 64,
  COGL_INFO_SIZE = 192,
  COGL_INFO_AGENT = 4, COGL_INFO_AGENT_STRIDE = 32,
  COGL_GRID = 48,
  COGL_GRID_BYTES = 48 * 48,
  /* packed host-I/O records (not in the reference; include/cog_b200.h): an ActionMask as four bit sets —
   * u32 play, u32 remove, u32 move | get_from_shop << 7, u32 play_special (the mask words of the compact
   * record) — and the five ActionData bytes in an 8-byte record */
  COGL_PACKED_MASK_SIZE = 16,
  COGL_PACKED_ACTION_SIZE = 8
};

/* ---- hidden per-env record ------------------------------------------------
 * line 0..4  card piles draw/hand/active/played/discard: lane l (1..21) holds card
 *            type l-1, one byte per player, ROTATED by the current agent: byte r of the
 *            word = the count of player (agent + r) & 3 (the agent is always byte 0)
 * line 5     shop: lane l (1..18) holds slot l-1: bits 0-7 n_available, bit 8 in_market
 * line 6     neighbour cache: lane d (0..6): byte r = hex code of the hex in direction d
 *            of player (agent + r) & 3 (d = 0: the hex the player stands on)
 * line 7     per-player fields, word 8*p + f = field f of player p (COG_PF_*): a player's
 *            fields are two 16-byte vectors
 * line 8     env scalars, word = COG_ES_* (four 16-byte vectors)
 */
enum {
  COG_HID_LINES = 9,
  COG_HID_WORDS = COG_HID_LINES * 32,
  COG_HL_DRAW = 0, COG_HL_HAND = 1, COG_HL_ACTIVE = 2, COG_HL_PLAYED = 3, COG_HL_DISCARD = 4,
  COG_HL_SHOP = 5, COG_HL_NBR = 6, COG_HL_PLAYER = 7, COG_HL_ENV = 8
};
/* per-player fields (line 7) */
enum {
  COG_PF_FLAGS = 0, /* bit0 has_won, bit1 movement_in_progress, bit2 next_card_free,
                       bit3 next_move_free; byte1 n_removes; byte2 n_added_cards; byte3 steps_taken */
  COG_PF_COUNTS = 1, /* byte0 n_in_hand, byte1 n_active, byte2 n_in_draw, byte3 idx_last_activated */
  COG_PF_MOVES = 2,  /* n_movements u32 */
  COG_PF_LOC = 3,    /* int16 x | int16 y << 16, in half-hex units (2x the reference's float) */
  COG_PF_MPLAY = 4, COG_PF_MSPEC = 5 /* unused: play_special is derived from play */,
  COG_PF_MREM = 6, /* P[p] mask heads as bit sets */
  COG_PF_MMS = 7,    /* P[p].move (bits 0-6) | P[p].get_from_shop (bits 7-25) */
  COG_PF_STRIDE = 8  /* words per player */
};
/* env scalars (line 8): four 16-byte vectors.  The first two hold what EVERY step changes (phase,
 * resources, the play head of S, the step counters), the last two what only a turn end, a draw or a
 * special action changes — the fast path of the single-step kernel writes back the first 32 bytes only */
enum {
  COG_ES_MISC = 0,       /* byte0 agent_selection, byte1 phase, byte2 n_in_market */
  COG_ES_RES0 = 1, COG_ES_RES1 = 2, COG_ES_RES2 = 3, /* current_resources * 2 (int32) */
  COG_ES_SPLAY = 4,      /* S.play (S.play_special is derived from it) */
  COG_ES_INFOSTEPS = 5,  /* Info.agent_infos[k].steps_taken, one byte per player (rotated) */
  COG_ES_CPF = 6, COG_ES_CPC = 7, /* live COG_PF_FLAGS / COG_PF_COUNTS of the current agent
                                     (that player's words in line 7 are stale) */
  COG_ES_RNG = 8,        /* minstd_rand0 state of the env */
  COG_ES_TURN = 9,       /* turn_counter */
  COG_ES_MIN = 10,       /* int16 min_x | int16 min_y << 16 (half-hex units) */
  COG_ES_SREM = 11, COG_ES_SMS = 12, /* S.remove, S.move | S.get_from_shop << 7 */
  COG_ES_PENDING = 14, /* only non-zero INSIDE a fused launch: 1 + the sample+step iterations this env
                          still owes after an episode end (k_step_tail resets the env, runs them and
                          clears the word) */
  COG_ES_DEAD = 15     /* single-env semantics only (auto-reset off): the episode is over, steps are
                          ignored until reset() */
};

/* ------------------------------------------------------------------------------------------
 * Compact record (not in the reference): one env as seen by its current
 * agent in 256 bytes.  It carries every dynamic field of ObsData that the acting agent needs
 * (the 48x48x7 map is static per episode and lives in COG_BUF_GRID, one byte per hex) plus the
 * players' grid positions, which the reference never exports (map feature 0 is never written,
 * SURVEY A#1).  The layout follows the warp that writes it: warp-uniform words sit in 16-byte
 * units that lane 0 stores whole, per-lane data where its lane stores it.
 *
 *    0  u8 agent_selection, u8 phase, u16 current_resources[3] * 2
 *    8  the action taken from this state: u8 play, play_special, remove, move, get_from_shop,
 *       3 zero bytes (all zero in cog_vec_compact_obs)
 *   16  u32 mask.play (bit k = play[k]), u32 mask.remove,
 *       u32 mask.move (bits 0-6) | mask.get_from_shop << 7 (bits 7-25), u32 mask.play_special
 *   32  4 x { u8 row, u8 col }: the cell of COG_BUF_GRID player p stands on (0, 0 for p >= n_players)
 *   40  u32 turn_counter, u32 index of the step since cog_vec_rollout_attach (low 32 bits)
 *   48  u8 done (the step taken from this state ended the episode), u8 filled (1 in rollout
 *       records: action, done, rewards, step index are valid), 14 zero bytes
 *   64  float rewards[4] of that step (the env's rewards when done, else 0)
 *   80  21 x { u8 draw, hand, active, played, discard (DeckObs of the CURRENT agent, card type c),
 *       u8 shop stock of slot c (c < 18, else 0), 2 zero bytes }
 *  248  8 zero bytes
 * The masks are the heads the action was (or will be) sampled from: selected_action_masks in
 * COG_MASKS_SELECTED mode, the merged legal mask in COG_MASKS_MERGED mode. */
enum {
  COGL_CR_SIZE = 256,
  COGL_CR_HEAD = 0, COGL_CR_ACTION = 8, COGL_CR_MASK_PLAY = 16, COGL_CR_MASK_REMOVE = 20,
  COGL_CR_MASK_MOVESHOP = 24, COGL_CR_MASK_SPECIAL = 28, COGL_CR_LOC = 32, COGL_CR_TURN = 40, COGL_CR_T = 44,
  COGL_CR_DONE = 48, COGL_CR_REWARDS = 64, COGL_CR_CARDS = 80
};

#endif
