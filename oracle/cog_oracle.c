/* oracle/cog_oracle.c — TEST INFRASTRUCTURE (the checker); see cog_oracle.h.
 *
 * Scalar C restatement of the reference's hot path.  Every function cites the
 * reference file:line it follows (paths relative to the reference tree).
 * Coordinates are the reference's float axial hex coordinates multiplied by 2
 * (exact: every coordinate the reference produces is a multiple of 0.5).
 * PARITY PINNED against the compiled reference and the golden vectors minted from it
 * (tests/test_oracle.py).
 */
#define _POSIX_C_SOURCE 200809L
#include "cog_oracle.h"
#include "cog_tables.h"

#include <stdlib.h>
#include <string.h>
#include <time.h>

/* ------------------------------------------------------------------ tables */
typedef struct { int8_t x, y; } xy8;
static const xy8 COORDS_LARGE[37] = COG_COORDS_LARGE_INIT;
static const xy8 COORDS_SMALL[16] = COG_COORDS_SMALL_INIT;
static const xy8 COORDS_END[3] = COG_COORDS_END_INIT;
typedef struct { uint8_t n_hexes, size, type, difficulty; } piece_meta;
static const piece_meta PIECE_META[COG_N_PIECES_TOTAL] = COG_PIECE_META_INIT;
static const uint8_t PIECE_HEX[COG_N_PIECES_TOTAL][COG_PIECE_MAX_HEXES] = COG_PIECE_HEX_INIT;
typedef struct { uint8_t cost, market, single, res0, res1, res2, special; } card_def;
static const card_def CARDS[21] = COG_CARDS_INIT;
static uint8_t card_res(int card, int k) {
  return k == 0 ? CARDS[card].res0 : k == 1 ? CARDS[card].res1 : CARDS[card].res2;
}
static const uint8_t SHOP_SLOT_CARD[18] = COG_SHOP_SLOT_CARD_INIT;

enum { SIZE_LARGE = 0, SIZE_SMALL = 1, SIZE_TRIPLE = 2 };
enum { TYPE_START = 0, TYPE_TRAVEL = 1, TYPE_END = 2 };
enum { PH_INACTIVE = 0, PH_MOVEMENT = 1, PH_BUYING = 2, PH_MAX = 3 };
enum { REQ_MACHETE = 0, REQ_PADDLE, REQ_COIN, REQ_DISCARD, REQ_REMOVE, REQ_NULL };
enum { SP_NONE = 0, SP_TRANSMIT, SP_DRAW2, SP_DRAW3, SP_DRAW1_REMOVE1, SP_DRAW2_REMOVE2,
       SP_NATIVE, SP_REMOVE_PENDING };

/* hex code helpers (encoding: tools/extract_tables.py) */
static int hex_req(uint8_t c) { return (c & 7) ? (c & 7) - 1 : REQ_NULL; }
static int hex_n(uint8_t c) { return (c >> 3) & 7; }
static int hex_end(uint8_t c) { return (c >> 6) & 1; }

static const xy8 *piece_coords(int piece) {
  switch (PIECE_META[piece].size) {
  case SIZE_LARGE: return COORDS_LARGE;
  case SIZE_SMALL: return COORDS_SMALL;
  default: return COORDS_END;
  }
}

/* -------------------------------------------------------------------- PRNG */
/* std::default_random_engine = minstd_rand0 (libstdc++ bits/random.h:1593);
 * seed(): bits/random.tcc:119-126 */
uint32_t orc_minstd_seed(uint64_t s) {
  uint64_t x = s % 2147483647ULL;
  return (uint32_t)(x ? x : 1);
}
uint32_t orc_minstd_next(uint32_t *state) {
  *state = (uint32_t)(((uint64_t)*state * 16807ULL) % 2147483647ULL);
  return *state;
}
/* std::uniform_int_distribution<size_t>(lo,hi)(minstd_rand0): the "2 divisions"
 * fallback of libstdc++ bits/uniform_int_dist.h:332-340 (engine range 2^31-3). */
uint64_t orc_uniform(uint32_t *state, uint64_t lo, uint64_t hi) {
  const uint64_t urngrange = 2147483645ULL;
  uint64_t urange = hi - lo;
  if (urange >= urngrange) abort(); /* upscaling path: unreachable for valid play */
  uint64_t uerange = urange + 1;
  uint64_t scaling = urngrange / uerange;
  uint64_t past = uerange * scaling;
  uint64_t ret;
  do {
    ret = (uint64_t)orc_minstd_next(state) - 1;
  } while (ret >= past);
  return lo + ret / scaling;
}

/* --------------------------------------------------------- board generation */
/* one +60 degree step in 2x axial coordinates: (x,y) -> (-y, x+y)
 * (src/map.cpp:17-37 via src/geometry.cpp:3-17, exact on the half lattice) */
static void rotate_pt(int *x, int *y, int times) {
  int t = times % 6;
  if (t < 0) t += 6;
  while (t--) {
    int nx = -*y, ny = *x + *y;
    *x = nx;
    *y = ny;
  }
}

#define BG_MAX_PIECES 512
#define BG_MAX_HEXES (BG_MAX_PIECES * 37)
#define BG_OCC_HALF 1024 /* occupancy bitmap covers 2x coords in [-1024,1024) */

typedef struct {
  int n_list;                       /* Map::pieces (pointers -> ids, may repeat) */
  uint8_t list[BG_MAX_PIECES];
  int cx[COG_N_PIECES_TOTAL], cy[COG_N_PIECES_TOTAL], rot[COG_N_PIECES_TOTAL]; /* MapPiece pose (global objects) */
  int n_hex;                        /* Map::xy / Map::hexes */
  int16_t hx[BG_MAX_HEXES], hy[BG_MAX_HEXES];
  uint8_t hcode[BG_MAX_HEXES];
  int minx, miny, maxx, maxy;       /* Map::min_xy / max_xy */
  uint8_t *occ;                     /* (2*BG_OCC_HALF)^2 bytes, 1 = some hex has this coordinate */
  int overflow;
} boardgen;

static void bg_map_reset(boardgen *b) { /* Map::reset src/map.cpp:744-752 */
  for (int i = 0; i < b->n_hex; ++i)
    b->occ[(b->hx[i] + BG_OCC_HALF) * 2 * BG_OCC_HALF + (b->hy[i] + BG_OCC_HALF)] = 0;
  b->n_list = 0;
  b->n_hex = 0;
  b->minx = b->miny = b->maxx = b->maxy = 0;
}

/* Map::add_piece src/map.cpp:309-341 (hex_array/hex_index are rebuilt from the
 * final bounds in bg_finalize; only the last rebuild is observable) */
static void bg_add_piece(boardgen *b, int piece, int px, int py, int rotation) {
  int r = rotation % 6; /* MapPiece::rotate :179-184 */
  b->rot[piece] = b->rot[piece] + r;
  b->cx[piece] += px; /* MapPiece::translate :186-191 */
  b->cy[piece] += py;
  if (b->n_list >= BG_MAX_PIECES || b->n_hex + 37 > BG_MAX_HEXES) {
    b->overflow = 1;
    return;
  }
  b->list[b->n_list++] = (uint8_t)piece;
  const xy8 *c = piece_coords(piece);
  for (int i = 0; i < PIECE_META[piece].n_hexes; ++i) {
    int x = c[i].x, y = c[i].y;
    rotate_pt(&x, &y, b->rot[piece]);
    x += b->cx[piece];
    y += b->cy[piece];
    if (x < -BG_OCC_HALF || x >= BG_OCC_HALF || y < -BG_OCC_HALF || y >= BG_OCC_HALF) {
      b->overflow = 1;
      return;
    }
    b->hx[b->n_hex] = (int16_t)x;
    b->hy[b->n_hex] = (int16_t)y;
    b->hcode[b->n_hex] = PIECE_HEX[piece][i];
    b->n_hex++;
    b->occ[(x + BG_OCC_HALF) * 2 * BG_OCC_HALF + (y + BG_OCC_HALF)] = 1;
    if (x > b->maxx) b->maxx = x;
    if (y > b->maxy) b->maxy = y;
    if (x < b->minx) b->minx = x;
    if (y < b->miny) b->miny = y;
  }
}

/* connection tables: MapPiece::get_ref_connection_points src/map.cpp:203-263 */
static const int8_t CONN_LL[2][2] = {{8, 6}, {6, 8}};
static const int8_t CONN_LS[3][2] = {{3, 7}, {5, 5}, {7, 3}};
static const int8_t CONN_LT[1][2] = {{0, 8}};
static const int8_t CONN_SL[6][2] = {{-7, 10}, {-5, 10}, {-3, 10}, {7, -10}, {5, -10}, {3, -10}};
static const int ROT6[6] = {-2, -1, 0, 1, 2, 3};
static const int ROT2[2] = {-1, 2};
static const int ROT1[1] = {-3};

typedef struct { int x, y, rot_add, n_opts; const int *opts; } candidate;

/* entries of host->get_centered_connections(new) (src/map.cpp:192-201), appended to out */
static int bg_connections(const boardgen *b, int host, int newp, candidate *out, int cap) {
  const int8_t(*base)[2] = NULL;
  int n_alt = 0, expand = 0, n_opts = 0;
  const int *opts = NULL;
  int hs = PIECE_META[host].size, ht = PIECE_META[host].type, ns = PIECE_META[newp].size;
  if (hs == SIZE_LARGE) {
    if (ns == SIZE_LARGE) { base = CONN_LL; n_alt = 2; expand = 1; opts = ROT6; n_opts = 6; }
    else if (ns == SIZE_SMALL) { base = CONN_LS; n_alt = 3; expand = 1; opts = ROT2; n_opts = 2; }
    else if (ns == SIZE_TRIPLE && ht != TYPE_START) { base = CONN_LT; n_alt = 1; expand = 1; opts = ROT1; n_opts = 1; }
  } else if (hs == SIZE_SMALL && ns == SIZE_LARGE) {
    base = CONN_SL; n_alt = 6; expand = 0; opts = ROT6; n_opts = 6;
  }
  int n = 0;
  int reps = expand ? 7 : 1; /* the base entries plus six derived copies :249-260 */
  for (int i = 0; i < reps && n_alt; ++i)
    for (int j = 0; j < n_alt; ++j) {
      if (n >= cap) return -1;
      int x = base[j][0], y = base[j][1];
      rotate_pt(&x, &y, i);                /* derived entry i = base rotated i times */
      rotate_pt(&x, &y, b->rot[host]);     /* ::rotate(coords, rotation) */
      out[n].x = x + b->cx[host];          /* ::translate(.., center) */
      out[n].y = y + b->cy[host];
      out[n].rot_add = i + b->rot[host];   /* rot += rotation */
      out[n].n_opts = n_opts;
      out[n].opts = opts;
      ++n;
    }
  return n;
}

/* Map::add_random_piece src/map.cpp:277-307 */
static int bg_add_random_piece(boardgen *b, int piece, uint32_t *rng) {
  static candidate cand[BG_MAX_PIECES * 21];
  /* MapPiece::reset :174-177 — acts on the global object, so entries of this
   * piece already in Map::pieces see the canonical pose while candidates are built */
  b->cx[piece] = b->cy[piece] = 0;
  b->rot[piece] = 0;
  int n_cand = 0;
  for (int k = 0; k < b->n_list; ++k) {
    int got = bg_connections(b, b->list[k], piece, cand + n_cand, BG_MAX_PIECES * 21 - n_cand);
    if (got < 0) { b->overflow = 1; return 0; }
    n_cand += got;
  }
  const xy8 *c = piece_coords(piece);
  int nh = PIECE_META[piece].n_hexes;
  int n_valid = 0;
  for (int k = 0; k < n_cand; ++k) {
    int hit = 0;
    for (int i = 0; i < nh && !hit; ++i) {
      int x = c[i].x, y = c[i].y;
      rotate_pt(&x, &y, cand[k].opts[0] + cand[k].rot_add); /* first rotation option only */
      x += cand[k].x;
      y += cand[k].y;
      if (x >= -BG_OCC_HALF && x < BG_OCC_HALF && y >= -BG_OCC_HALF && y < BG_OCC_HALF)
        hit = b->occ[(x + BG_OCC_HALF) * 2 * BG_OCC_HALF + (y + BG_OCC_HALF)];
    }
    if (!hit) cand[n_valid++] = cand[k]; /* ::overlap :53-74 = exact coordinate match */
  }
  if (!n_valid) return 0;
  int idx = (int)orc_uniform(rng, 0, (uint64_t)n_valid - 1);
  int rot_idx = (int)orc_uniform(rng, 0, (uint64_t)cand[idx].n_opts - 1);
  bg_add_piece(b, piece, cand[idx].x, cand[idx].y, cand[idx].opts[rot_idx] + cand[idx].rot_add);
  return 1;
}

/* Map::generate src/map.cpp:697-742.  rng is passed BY VALUE like the reference.
 * returns 0 ok, -2 generate_map_failure (the reference throws). */
static int bg_generate(boardgen *b, int n_pieces, int difficulty, int failures, uint32_t rng) {
  if (failures >= 5) return -2;
  int start = (int)orc_uniform(&rng, 0, 1);
  bg_add_piece(b, start, 0, 0, 0);
  int valid[16], n_valid = 0;
  for (int i = 0; i < 16; ++i)
    if (PIECE_META[2 + i].difficulty <= difficulty) valid[n_valid++] = i;
  for (int i = 0; i < n_pieces; ++i) {
    int success = 0, next_idx = 0;
    if (n_valid) {
      next_idx = valid[orc_uniform(&rng, 0, (uint64_t)n_valid - 1)];
      success = bg_add_random_piece(b, 2 + next_idx, &rng);
    }
    if (success) {
      /* valid_indices.erase(begin + next_idx): the VALUE is used as a position;
       * past-the-end positions degenerate to pop_back in libstdc++ 13
       * (bits/vector.tcc:182-190). */
      if (next_idx < n_valid - 1)
        memmove(valid + next_idx, valid + next_idx + 1, sizeof(int) * (size_t)(n_valid - 1 - next_idx));
      n_valid--;
    } else {
      int rc = bg_generate(b, n_pieces, difficulty, failures + 1, rng); /* no reset */
      if (rc) return rc;
    }
    if (b->overflow) return -3;
  }
  int endp = 18 + (int)orc_uniform(&rng, 0, 1);
  if (!bg_add_random_piece(b, endp, &rng)) {
    bg_map_reset(b);
    int rc = bg_generate(b, n_pieces, difficulty, failures + 1, rng);
    if (rc) return rc;
  }
  return b->overflow ? -3 : 0;
}

static boardgen *bg_new(void) {
  boardgen *b = (boardgen *)calloc(1, sizeof(boardgen));
  b->occ = (uint8_t *)calloc((size_t)4 * BG_OCC_HALF * BG_OCC_HALF, 1);
  return b;
}
static void bg_free(boardgen *b) {
  if (b) { free(b->occ); free(b); }
}

/* Map::finalize src/map.cpp:389-405 with hex_index from the last add_piece
 * (:335-340): cell = trunc(coord - min + 1).  obs has `rows` rows of 48*7 bytes.
 * Also fills the 1-byte-per-hex grid (same indexing as Map::hex_array). */
static int bg_finalize(const boardgen *b, uint8_t *obs, uint32_t rows, uint8_t *grid, int gdim) {
  memset(obs, 0, (size_t)rows * 48 * 7);
  if (grid) memset(grid, 0, (size_t)gdim * (size_t)gdim);
  int bad = 0;
  for (int i = 0; i < b->n_hex; ++i) {
    int ix = (b->hx[i] - b->minx + 2) / 2;
    int iy = (b->hy[i] - b->miny + 2) / 2;
    uint8_t code = b->hcode[i];
    size_t off = ((size_t)ix * 48 + (size_t)iy) * 7;
    if (off + 7 > (size_t)rows * 48 * 7) { bad = 1; continue; }
    uint8_t *cell = obs + off;
    memset(cell, 0, 7);
    if (hex_req(code) != REQ_NULL) cell[hex_req(code) + 1] = (uint8_t)hex_n(code);
    cell[6] = (uint8_t)hex_end(code);
    if (grid) {
      if (ix >= gdim || iy >= gdim) bad = 1;
      else grid[ix * gdim + iy] = code;
    }
  }
  return bad ? -3 : 0;
}

int orc_map_generate(uint32_t seed, uint8_t n_pieces, int difficulty, void *obs,
                     uint32_t rows, uint32_t *out_meta, int32_t *out_xy,
                     uint32_t max_hexes) {
  boardgen *b = bg_new();
  bg_map_reset(b);
  int rc = bg_generate(b, n_pieces, difficulty, 0, orc_minstd_seed(seed));
  if (rc == 0) {
    bg_finalize(b, (uint8_t *)obs, rows, NULL, 0);
    out_meta[0] = (uint32_t)b->n_list;
    out_meta[1] = (uint32_t)b->n_hex;
    out_meta[2] = (uint32_t)b->minx;
    out_meta[3] = (uint32_t)b->miny;
    out_meta[4] = (uint32_t)b->maxx;
    out_meta[5] = (uint32_t)b->maxy;
    for (int i = 0; i < b->n_hex && (uint32_t)i < max_hexes; ++i) {
      out_xy[2 * i] = b->hx[i];
      out_xy[2 * i + 1] = b->hy[i];
    }
  }
  bg_free(b);
  return rc;
}

/* ------------------------------------------------------- environment state */
/* exported-record offsets (include/api.h:67-161) */
enum {
  OBS_PHASE = 16128, OBS_RES = 16132, OBS_SHOP = 16144, OBS_PLAYER = 16192, PLAYER_STRIDE = 256,
  DK_DRAW = 0, DK_HAND = 21, DK_ACTIVE = 42, DK_PLAYED = 63, DK_DISCARD = 84, PD_MASK = 128,
  AM_PLAY = 0, AM_SPECIAL = 22, AM_REMOVE = 44, AM_MOVE = 66, AM_SHOP = 73, AM_PAYLOAD = 92,
  INFO_AGENT = 4, INFO_AGENT_STRIDE = 32
};
#define ORC_GRID 64

typedef struct { /* Player (include/player.h:12-16,66-74) + Deck (include/cards.h:70-77) */
  uint8_t has_won, movement_in_progress, n_removes, next_card_free, next_move_free;
  uint8_t n_added_cards, steps_taken;
  uint32_t n_movements;
  uint8_t n_in_hand, n_active, n_in_draw, idx_last_activated;
} orc_player;

typedef struct { /* cog_env (include/environment.h:13-33) + Shop + Map hidden state */
  uint32_t rng, seed, max_steps, turn_counter;
  uint8_t n_players, n_pieces, agent, done;
  int difficulty;
  uint8_t n_in_market, in_market[18];
  orc_player pl[4];
  int loc[4][2];
  int minx, miny;
  uint8_t grid[ORC_GRID * ORC_GRID];
  uint8_t *obs, *S, *info;
  float *rewards;
  uint64_t global_index; /* env_index_offset + i */
  uint32_t n_start_b;    /* boards so far whose piece list holds start piece B (tests: SURVEY A#13) */
} orc_env;

struct orc_vec {
  uint64_t n;
  orc_env *env;
  uint32_t *sampler_rng;
  uint8_t *obs, *masks, *dones, *infos, *agents, *actions;
  float *rewards;
  boardgen *bg;
  int no_autoreset; /* single-env semantics: cog_env::step does not reset (src/environment.cpp:187-207) */
};

static float *e_res(orc_env *e) { return (float *)(e->obs + OBS_RES); }
static uint8_t *e_shop(orc_env *e) { return e->obs + OBS_SHOP; }
static uint8_t *e_deck(orc_env *e, int p) { return e->obs + OBS_PLAYER + PLAYER_STRIDE * p; }
static uint8_t *e_pmask(orc_env *e, int p) { return e_deck(e, p) + PD_MASK; }
static int is_special(int card) { return CARDS[card].special != SP_NONE; }

/* ActionMask::reset include/api.h:104-118 (move[1..], get_from_shop[1..] untouched) */
static void mask_reset(uint8_t *m) {
  memset(m + AM_PLAY, 0, 22); m[AM_PLAY] = 1;
  memset(m + AM_REMOVE, 0, 22); m[AM_REMOVE] = 1;
  memset(m + AM_SPECIAL, 0, 22); m[AM_SPECIAL] = 1;
  m[AM_MOVE] = 1;
  m[AM_SHOP] = 1;
}

/* ---- Deck (src/cards.cpp:163-292); the deck's action_mask is always S */
static void deck_move_discard_to_draw(orc_env *e, int p) { /* :234-240 */
  uint8_t *d = e_deck(e, p);
  for (int i = 0; i < 21; ++i) {
    d[DK_DRAW + i] = (uint8_t)(d[DK_DRAW + i] + d[DK_DISCARD + i]);
    e->pl[p].n_in_draw = (uint8_t)(e->pl[p].n_in_draw + d[DK_DISCARD + i]);
    d[DK_DISCARD + i] = 0;
  }
}
static void deck_draw(orc_env *e, int p, uint8_t n) { /* :183-211 */
  orc_player *pl = &e->pl[p];
  uint8_t *d = e_deck(e, p);
  if (pl->n_in_draw < n) deck_move_discard_to_draw(e, p);
  if (pl->n_in_draw < n) n = pl->n_in_draw;
  for (int i = 0; i < n; ++i) {
    uint64_t target = orc_uniform(&e->rng, 0, (uint64_t)pl->n_in_draw - 1);
    int card = 0;
    while (target >= d[DK_DRAW + card]) { target -= d[DK_DRAW + card]; ++card; }
    --d[DK_DRAW + card];
    --pl->n_in_draw;
    ++d[DK_HAND + card];
    e->S[AM_PLAY + card + 1] = 1;
    e->S[AM_SPECIAL + card + 1] = (uint8_t)is_special(card);
  }
  pl->n_in_hand = (uint8_t)(pl->n_in_hand + n);
}
static void deck_reset(orc_env *e, int p) { /* :163-171 */
  uint8_t *d = e_deck(e, p);
  d[DK_DISCARD + 0] = 3;  /* EXPLORER */
  d[DK_DISCARD + 7] = 4;  /* TRAVELER */
  d[DK_DISCARD + 5] = 1;  /* SAILOR */
  e->pl[p].n_in_draw = 0;
  e->pl[p].n_in_hand = 0;
  e->pl[p].n_active = 0;
  deck_draw(e, p, 4);
}
static void deck_activate(orc_env *e, int p, int idx) { /* :242-253 */
  uint8_t *d = e_deck(e, p);
  --e->pl[p].n_in_hand;
  ++e->pl[p].n_active;
  e->pl[p].idx_last_activated = (uint8_t)idx;
  uint8_t prev = d[DK_HAND + idx]--;
  ++d[DK_ACTIVE + idx];
  e->S[AM_PLAY + idx + 1] = prev > 1;
  e->S[AM_SPECIAL + idx + 1] = e->S[AM_PLAY + idx + 1] && is_special(idx);
}
static void deck_play_last_activated(orc_env *e, int p) { /* :255-261 */
  uint8_t *d = e_deck(e, p);
  int idx = e->pl[p].idx_last_activated;
  --e->pl[p].n_active;
  --d[DK_ACTIVE + idx];
  if (!CARDS[idx].single) ++d[DK_PLAYED + idx];
}
static void deck_play_immediate(orc_env *e, int p, int idx) { /* :263-273 */
  uint8_t *d = e_deck(e, p);
  --e->pl[p].n_in_hand;
  uint8_t prev = d[DK_HAND + idx]--;
  ++d[DK_PLAYED + idx];
  e->S[AM_PLAY + idx + 1] = prev > 1;
  e->S[AM_SPECIAL + idx + 1] = e->S[AM_PLAY + idx + 1] && is_special(idx);
}
static void deck_remove_immediate(orc_env *e, int p, int idx) { /* :281-290 */
  uint8_t *d = e_deck(e, p);
  --e->pl[p].n_in_hand;
  uint8_t prev = d[DK_HAND + idx]--;
  e->S[AM_REMOVE + idx + 1] = e->S[AM_REMOVE + idx + 1] && prev > 1;
  e->S[AM_PLAY + idx + 1] = e->S[AM_PLAY + idx + 1] && prev > 1;
  e->S[AM_SPECIAL + idx + 1] = e->S[AM_PLAY + idx + 1] && is_special(idx);
}

/* ---- Player (src/player.cpp) */
static void player_disable_playing(orc_env *e) { /* :191-196 */
  memset(e->S + AM_PLAY, 0, 22); e->S[AM_PLAY] = 1;
  memset(e->S + AM_SPECIAL, 0, 22); e->S[AM_SPECIAL] = 1;
}
static void player_enable_playing(orc_env *e, int p) { /* :198-206 */
  uint8_t *d = e_deck(e, p);
  memset(e->S + AM_REMOVE, 0, 22); e->S[AM_REMOVE] = 1;
  for (int idx = 1; idx < 22; ++idx) {
    e->S[AM_PLAY + idx] = d[DK_HAND + idx - 1] > 0;
    e->S[AM_SPECIAL + idx] = e->S[AM_PLAY + idx] && is_special(idx - 1);
  }
}
static void player_reset(orc_env *e, int p) { /* :29-43 */
  orc_player *pl = &e->pl[p];
  pl->has_won = pl->movement_in_progress = pl->next_card_free = pl->next_move_free = 0;
  pl->n_removes = 0;
  pl->steps_taken = 0;
  pl->n_movements = 0;
  pl->n_added_cards = 0;
  uint8_t *d = e_deck(e, p); /* DeckObs::reset include/api.h:75-81: played is NOT cleared */
  memset(d + DK_DRAW, 0, 21); memset(d + DK_HAND, 0, 21);
  memset(d + DK_ACTIVE, 0, 21); memset(d + DK_DISCARD, 0, 21);
  mask_reset(e->S);
  deck_reset(e, p);
  memcpy(e_pmask(e, p), e->S, AM_PAYLOAD); /* save_actionmask :166 */
}
static void player_play_card(orc_env *e, int p, int card, int phase) { /* :47-69 */
  float *res = e_res(e);
  if (phase == PH_MOVEMENT) {
    for (int k = 0; k < 3; ++k) res[k] = (float)card_res(card, k);
  } else if (phase == PH_BUYING) {
    if (card_res(card, REQ_COIN) > 0) res[REQ_COIN] += (float)card_res(card, REQ_COIN);
    else res[REQ_COIN] += 0.5f;
  }
  deck_activate(e, p, card);
}
static int player_play_special(orc_env *e, int p, int card) { /* :71-79 */
  if (CARDS[card].single) deck_remove_immediate(e, p, card);
  else deck_play_immediate(e, p, card);
  return CARDS[card].special;
}
static void player_discard_cards(orc_env *e, int p, uint8_t n) { /* :85-104 (no clamp) */
  uint8_t *d = e_deck(e, p);
  uint8_t n_discardable = e->pl[p].n_active;
  for (uint8_t i = 0; i < n; ++i) {
    uint64_t target = orc_uniform(&e->rng, 0, (uint64_t)(int64_t)((int)n_discardable - 1 - (int)i));
    int card = 0;
    while (target >= d[DK_ACTIVE + card]) { target -= d[DK_ACTIVE + card]; ++card; }
    --e->pl[p].n_active; /* Deck::discard :213-217 */
    d[DK_ACTIVE + card]--;
    d[DK_DISCARD + card]++;
  }
}
static void player_remove_cards(orc_env *e, int p, uint8_t n) { /* :106-131 (clamped) */
  uint8_t *d = e_deck(e, p);
  uint8_t n_removable = e->pl[p].n_active;
  if (n > n_removable) n = n_removable;
  for (uint8_t i = 0; i < n; ++i) {
    uint64_t target = orc_uniform(&e->rng, 0, (uint64_t)(int64_t)((int)n_removable - 1 - (int)i));
    int card = 0;
    while (target >= d[DK_ACTIVE + card]) { target -= d[DK_ACTIVE + card]; ++card; }
    --e->pl[p].n_active; /* Deck::remove_active :275-279 */
    d[DK_ACTIVE + card]--;
  }
}
static void player_handle_requirement(orc_env *e, int p, int req, uint8_t n) { /* :141-162 */
  float *res = e_res(e);
  if (req < 3) {
    float left = res[req] - (float)n;
    res[0] = res[1] = res[2] = 0;
    res[req] = left;
    if (!e->pl[p].movement_in_progress) {
      deck_play_last_activated(e, p);
      e->pl[p].movement_in_progress = 1;
    }
  } else if (req == REQ_REMOVE) {
    player_remove_cards(e, p, n);
    res[0] = res[1] = res[2] = 0;
    e->pl[p].movement_in_progress = 0;
  } else if (req == REQ_DISCARD) {
    player_discard_cards(e, p, n);
    res[0] = res[1] = res[2] = 0;
    e->pl[p].movement_in_progress = 0;
  }
}
static void player_end_turn(orc_env *e, int p) { /* :170-180 */
  uint8_t *d = e_deck(e, p);
  e->pl[p].n_active = 0; /* Deck::discard_all_active / _played src/cards.cpp:219-232 */
  for (int i = 0; i < 21; ++i) {
    d[DK_DISCARD + i] = (uint8_t)(d[DK_DISCARD + i] + d[DK_ACTIVE + i]);
    d[DK_ACTIVE + i] = 0;
  }
  for (int i = 0; i < 21; ++i) {
    d[DK_DISCARD + i] = (uint8_t)(d[DK_DISCARD + i] + d[DK_PLAYED + i]);
    d[DK_PLAYED + i] = 0;
  }
  int n_draw = 4 - (int)e->pl[p].n_in_hand;
  if (n_draw > 0) deck_draw(e, p, (uint8_t)n_draw);
  float *res = e_res(e);
  res[0] = res[1] = res[2] = 0;
  memcpy(e_pmask(e, p), e->S, AM_PAYLOAD); /* save_actionmask */
}

/* ---- Shop (src/cards.cpp:94-142) */
static void shop_reset(orc_env *e) { /* :94-100: n_in_market is NOT reset */
  memset(e_shop(e), 3, 18);
  for (int i = 0; i < 18; ++i) e->in_market[i] = CARDS[SHOP_SLOT_CARD[i]].market;
}
static int shop_get(orc_env *e, int idx) { /* :136-142 */
  uint8_t *avail = e_shop(e);
  if (!--avail[idx] && e->in_market[idx]) {
    e->in_market[idx] = 0;
    e->n_in_market -= 1;
  }
  return SHOP_SLOT_CARD[idx];
}
static int shop_buy(orc_env *e, int idx) { /* :102-106 */
  e->n_in_market = (uint8_t)(e->n_in_market + (uint8_t)(1 - e->in_market[idx]));
  e->in_market[idx] = 1;
  return shop_get(e, idx);
}
static void shop_set_available_mask(orc_env *e, float coins, uint8_t *mask) { /* :109-121 */
  uint8_t *avail = e_shop(e);
  for (int i = 0; i < 18; ++i) {
    int cost = CARDS[SHOP_SLOT_CARD[i]].cost;
    if (e->n_in_market < 6) mask[i + 1] = (avail[i] > 0) && (coins > (float)cost);
    else mask[i + 1] = e->in_market[i] && (coins > (float)cost);
  }
}
static void shop_set_transmit_mask(orc_env *e, uint8_t *mask) { /* :123-127 */
  uint8_t *avail = e_shop(e);
  for (int i = 0; i < 18; ++i) mask[i + 1] = avail[i] > 0;
}

/* ---- Map runtime (src/map.cpp:273-275,356-387,407-410) */
static const int DIRS[7][2] = {{0, 0}, {2, 0}, {0, 2}, {-2, 2}, {-2, 0}, {0, -2}, {2, -2}};
static uint8_t map_lookup(const orc_env *e, int x, int y) { /* get_from_array :273-275 */
  int ix = (x - e->minx) / 2 + 1; /* C truncation toward zero, then +1 (SURVEY A#17) */
  int iy = (y - e->miny) / 2 + 1;
  if (ix < 0 || iy < 0 || ix >= ORC_GRID || iy >= ORC_GRID) return 0;
  return e->grid[ix * ORC_GRID + iy];
}
static void map_set_movement_mask(orc_env *e, uint8_t *mask, int player, const float *res, int n_active) {
  for (int i = 1; i < 7; ++i) {
    uint8_t c = map_lookup(e, e->loc[player][0] + DIRS[i][0], e->loc[player][1] + DIRS[i][1]);
    int req = hex_req(c), filled;
    if (req >= REQ_DISCARD) filled = n_active > hex_n(c);
    else filled = res[req] >= (float)hex_n(c);
    mask[AM_MOVE + i] = (req != REQ_NULL) && filled;
  }
}

/* ---- cog_env (src/environment.cpp) */
static void env_update_observation(orc_env *e, int agent) { /* :252-279 */
  uint8_t *am = e_pmask(e, agent);
  memset(am + AM_MOVE, 0, 7); am[AM_MOVE] = 1;
  memset(am + AM_SHOP, 0, 19); am[AM_SHOP] = 1;
  int phase = e->obs[OBS_PHASE];
  if (phase == PH_MOVEMENT)
    map_set_movement_mask(e, am, agent, e_res(e), e->pl[agent].n_active);
  else if (phase == PH_BUYING)
    shop_set_available_mask(e, e_res(e)[REQ_COIN], am + AM_SHOP);
}
static void env_cycle_phase(orc_env *e) { /* include/player.h:78-84 */
  uint8_t n = (uint8_t)(e->obs[OBS_PHASE] + 1);
  if (n >= PH_MAX) n = (uint8_t)(n - PH_MAX);
  e->obs[OBS_PHASE] = n;
}
static void env_next_agent(orc_env *e) { /* :79-89 */
  player_end_turn(e, e->agent);
  e->agent = (uint8_t)(e->agent + 1);
  if (e->agent >= e->n_players) e->agent = 0;
  memcpy(e->S, e_pmask(e, e->agent), AM_PAYLOAD); /* load_actionmask */
  float *res = e_res(e);
  res[0] = res[1] = res[2] = 0;
  e->turn_counter++;
}

static int env_reset(orc_env *e, boardgen *bg) { /* cog_env::reset() :42-64 */
  e->agent = 0;
  e->obs[OBS_PHASE] = PH_INACTIVE;
  bg_map_reset(bg);
  int rc = bg_generate(bg, e->n_pieces, e->difficulty, 0, e->rng); /* rng BY VALUE :48 */
  if (rc) return rc;
  for (int k = 0; k < bg->n_list; ++k)
    if (bg->list[k] == 1) { e->n_start_b++; break; }
  if (bg_finalize(bg, e->obs, 48, e->grid, ORC_GRID)) return -3;
  /* the reference would write outside its 48x48 observation (UB): report instead */
  if ((bg->maxx - bg->minx + 2) / 2 > 47 || (bg->maxy - bg->miny + 2) / 2 > 47) return -3;
  e->minx = bg->minx;
  e->miny = bg->miny;
  for (int i = 0; i < e->n_players; ++i) player_reset(e, i);
  /* Map::add_players :343-354.  For 4 players both start pieces give
   * player i = hex i of the start piece; fewer players is UB in the reference
   * for start piece B (SURVEY A#13) and is DEFINED here the same way. */
  for (int i = 0; i < e->n_players && i < 4; ++i) {
    e->loc[i][0] = bg->hx[i];
    e->loc[i][1] = bg->hy[i];
  }
  shop_reset(e);
  e->done = 0;
  e->turn_counter = 0;
  for (int i = 0; i < e->n_players; ++i) env_update_observation(e, i);
  memcpy(e->S, e_pmask(e, e->agent), AM_PAYLOAD);
  return 0;
}

static void env_step(orc_env *e, const uint8_t *a) { /* cog_env::step :91-224 */
  if (e->done) return;
  uint8_t a_play = a[0], a_special = a[1], a_remove = a[2], a_move = a[3], a_shop = a[4];
  e->info[INFO_AGENT + INFO_AGENT_STRIDE * e->agent] += 1; /* steps_taken, wrapping u8 */
  if (e->obs[OBS_PHASE] == PH_INACTIVE) env_cycle_phase(e);
  int p = e->agent;
  orc_player *pl = &e->pl[p];
  pl->steps_taken++;
  int special = SP_NONE;

  if (a_play) {
    player_play_card(e, p, a_play - 1, e->obs[OBS_PHASE]);
  } else if (a_special) {
    special = player_play_special(e, p, a_special - 1);
  } else if (a_move) {
    int tx = e->loc[p][0] + DIRS[a_move][0], ty = e->loc[p][1] + DIRS[a_move][1];
    uint8_t c = map_lookup(e, tx, ty); /* Map::_move :361-367 */
    e->loc[p][0] = tx;
    e->loc[p][1] = ty;
    if (!pl->next_move_free) {
      player_handle_requirement(e, p, hex_req(c), (uint8_t)hex_n(c));
    } else {
      pl->next_move_free = 0;
      player_enable_playing(e, p);
    }
    pl->n_movements++;
    pl->has_won = (uint8_t)hex_end(c);
  } else {
    pl->next_move_free = 0;
    if (a_shop) {
      int idx = a_shop - 1, card;
      if (pl->next_card_free) {
        card = shop_get(e, idx); /* Shop::transmit */
      } else {
        card = shop_buy(e, idx);
        e_res(e)[REQ_COIN] -= (float)CARDS[card].cost; /* Player::pay */
        env_cycle_phase(e);
      }
      e_deck(e, p)[DK_DISCARD + card]++; /* Deck::add */
      pl->n_added_cards++;
    } else if (a_remove) {
      deck_remove_immediate(e, p, a_remove - 1);
      if (!--pl->n_removes) player_enable_playing(e, p);
      else special = SP_REMOVE_PENDING;
    } else {
      env_cycle_phase(e);
      if (pl->n_removes > 0) {
        pl->n_removes = 0;
        player_enable_playing(e, p);
      }
    }
    if (pl->next_card_free) {
      pl->next_card_free = 0;
      player_enable_playing(e, p);
    }
  }

  if (pl->movement_in_progress && !a_move) {
    pl->movement_in_progress = 0;
    float *res = e_res(e);
    res[0] = res[1] = res[2] = 0;
  }

  /* maybe_end_turn :244-250 */
  if (pl->has_won || e->obs[OBS_PHASE] == PH_INACTIVE) env_next_agent(e);
  env_update_observation(e, e->agent);

  if (special != SP_NONE) { /* :183-186; cards.cpp:8-36; lambda :156-158 */
    uint8_t *mask = e_pmask(e, e->agent);
    static const float plenty[3] = {100, 100, 100};
    switch (special) {
    case SP_TRANSMIT:
      memset(mask + AM_MOVE, 0, 7); mask[AM_MOVE] = 1;
      player_disable_playing(e);
      shop_set_transmit_mask(e, mask + AM_SHOP);
      pl->next_card_free = 1;
      break;
    case SP_DRAW2: deck_draw(e, p, 2); break;
    case SP_DRAW3: deck_draw(e, p, 3); break;
    case SP_DRAW1_REMOVE1:
    case SP_DRAW2_REMOVE2: {
      uint8_t k = special == SP_DRAW1_REMOVE1 ? 1 : 2;
      deck_draw(e, p, k);
      pl->n_removes = k;
      memcpy(mask + AM_REMOVE, mask + AM_PLAY, 22);
      player_disable_playing(e);
      shop_set_available_mask(e, 0.0f, mask + AM_SHOP);
      break;
    }
    case SP_NATIVE:
      map_set_movement_mask(e, mask, p, plenty, 100);
      pl->next_move_free = 1;
      player_disable_playing(e);
      shop_set_available_mask(e, 0.0f, mask + AM_SHOP);
      break;
    case SP_REMOVE_PENDING:
      shop_set_available_mask(e, 0.0f, mask + AM_SHOP);
      break;
    }
  } else {
    uint8_t here = map_lookup(e, e->loc[e->agent][0], e->loc[e->agent][1]);
    /* Map::player_done :407-410 uses get_from_array(loc - min_xy) like any lookup */
    if (hex_end(here) || e->turn_counter >= e->max_steps) {
      e->done = 1;
      memcpy(e->info, &e->turn_counter, 4);
      float n_winners = 0; /* get_reward :281-288 sums all four Player slots */
      for (int k = 0; k < 4; ++k) n_winners += (float)e->pl[k].has_won;
      for (int k = 0; k < e->n_players; ++k) {
        uint8_t *ai = e->info + INFO_AGENT + INFO_AGENT_STRIDE * k;
        float ret = (float)((int)e->n_players * (int)e->pl[k].has_won) - n_winners;
        uint32_t zero = 0, n_added = e->pl[k].n_added_cards;
        ai[0] = e->pl[k].steps_taken;
        memcpy(ai + 4, &ret, 4);
        e->rewards[k] = ret;
        memcpy(ai + 8, &e->pl[k].n_movements, 4);
        ai[12] = e->pl[k].n_added_cards;
        ai[13] = e->pl[k].n_added_cards; /* get_n_removed returns n_added_cards (player.cpp:224) */
        memcpy(ai + 16, &zero, 4);
        memcpy(ai + 20, &zero, 4);
        memcpy(ai + 24, &zero, 4);
        memcpy(ai + 28, &n_added, 4);   /* get_n_discarded returns n_added_cards (:223) */
      }
    }
  }
}

/* ----------------------------------------------------------------- sampler */
/* action_sampler::sample include/sampler.h:14-79: five heads, one draw each
 * (index 0 is always valid so a draw is always consumed) */
static void sampler_sample(uint32_t *rng, const uint8_t *mask, uint8_t *action) {
  static const int head_off[5] = {AM_PLAY, AM_SPECIAL, AM_REMOVE, AM_MOVE, AM_SHOP};
  static const int head_len[5] = {22, 22, 22, 7, 19};
  for (int h = 0; h < 5; ++h) {
    uint8_t valid[22];
    int k = 0;
    for (int i = 0; i < head_len[h]; ++i)
      if (mask[head_off[h] + i]) valid[k++] = (uint8_t)i;
    action[h] = 0;
    if (k) action[h] = valid[orc_uniform(rng, 0, (uint64_t)k - 1)];
  }
}

/* -------------------------------------------------------- vector container */
/* vec_cog_env<N> include/vec_environment.h:12-81, vec_action_sampler<N>
 * include/vec_sampler.h:7-27, with a runtime N and a global env-index offset
 * (seed + offset + i) so that a shard reproduces its slice of a larger job. */
orc_vec *orc_create(uint64_t n_envs, uint64_t env_index_offset, uint64_t sampler_seed) {
  orc_vec *v = (orc_vec *)calloc(1, sizeof(orc_vec));
  v->n = n_envs;
  v->env = (orc_env *)calloc(n_envs, sizeof(orc_env));
  v->sampler_rng = (uint32_t *)calloc(n_envs, sizeof(uint32_t));
  v->obs = (uint8_t *)calloc(n_envs, ORC_OBS_SIZE);
  v->masks = (uint8_t *)calloc(n_envs, ORC_MASK_SIZE);
  v->dones = (uint8_t *)calloc(n_envs, 1);
  v->infos = (uint8_t *)calloc(n_envs, ORC_INFO_SIZE);
  v->agents = (uint8_t *)calloc(n_envs, 1);
  v->actions = (uint8_t *)calloc(n_envs, ORC_ACTION_SIZE);
  v->rewards = (float *)calloc(n_envs * 4, sizeof(float));
  v->bg = bg_new();
  for (uint64_t i = 0; i < n_envs; ++i) {
    orc_env *e = &v->env[i];
    e->obs = v->obs + i * ORC_OBS_SIZE;
    e->S = v->masks + i * ORC_MASK_SIZE;
    e->info = v->infos + i * ORC_INFO_SIZE;
    e->rewards = v->rewards + 4 * i;
    /* defaults of cog_env::cog_env() src/environment.cpp:10-15 (seed: random_device) */
    e->n_players = 4; e->n_pieces = 3; e->difficulty = 0; e->max_steps = 100000;
    e->seed = (uint32_t)rand();
    e->rng = orc_minstd_seed(e->seed);
    e->n_in_market = 6;                           /* Shop::Shop src/cards.cpp:85-92 */
    memset(e->obs + OBS_SHOP, 3, 18);             /* Shop::init :80-83 */
    mask_reset(e->S);                             /* ActionMask() include/api.h:101-103 */
    for (int p = 0; p < 4; ++p) mask_reset(e_pmask(e, p));
    /* set_seed(seed + i): 64-bit sum, include/vec_sampler.h:11, sampler.h:12 */
    v->sampler_rng[i] = orc_minstd_seed((uint64_t)(uint32_t)sampler_seed + env_index_offset + i);
    e->global_index = env_index_offset + i;
  }
  return v;
}

void orc_destroy(orc_vec *v) {
  if (!v) return;
  free(v->env); free(v->sampler_rng); free(v->obs); free(v->masks); free(v->dones);
  free(v->infos); free(v->agents); free(v->actions); free(v->rewards);
  bg_free(v->bg);
  free(v);
}

int orc_reset(orc_vec *v, uint32_t seed, uint8_t n_players, uint8_t n_pieces,
              int difficulty, uint32_t max_steps) {
  /* vec_cog_env::reset(seed,..) include/vec_environment.h:38-44 ->
   * cog_env::reset(seed,..) src/environment.cpp:66-77 */
  for (uint64_t i = 0; i < v->n; ++i) {
    orc_env *e = &v->env[i];
    e->n_players = n_players; e->n_pieces = n_pieces; e->difficulty = difficulty;
    e->max_steps = max_steps;
    e->seed = (uint32_t)((uint64_t)seed + e->global_index); /* seed + i, truncated to u32 */
    e->rng = orc_minstd_seed(e->seed);
    int rc = env_reset(e, v->bg);
    if (rc) return rc;
  }
  return 0;
}

int orc_reset_again(orc_vec *v) { /* vec_cog_env::reset() :33-37 */
  for (uint64_t i = 0; i < v->n; ++i) {
    int rc = env_reset(&v->env[i], v->bg);
    if (rc) return rc;
  }
  return 0;
}

void orc_set_autoreset(orc_vec *v, int on) { v->no_autoreset = !on; }

int orc_step(orc_vec *v, const void *actions) {
  /* vec_cog_env::step / step_single include/vec_environment.h:46-61 */
  const uint8_t *a = (const uint8_t *)actions;
  for (uint64_t i = 0; i < v->n; ++i) {
    orc_env *e = &v->env[i];
    env_step(e, a + i * ORC_ACTION_SIZE);
    v->dones[i] = e->done;
    if (e->done && !v->no_autoreset) {
      int rc = env_reset(e, v->bg); /* auto-reset, no reseed */
      if (rc) return rc;
    }
    v->agents[i] = e->agent;
  }
  return 0;
}

void orc_sample(orc_vec *v, const void *masks) {
  const uint8_t *m = (const uint8_t *)masks;
  for (uint64_t i = 0; i < v->n; ++i)
    sampler_sample(&v->sampler_rng[i], m + i * ORC_MASK_SIZE, v->actions + i * ORC_ACTION_SIZE);
}

void *orc_buffer(orc_vec *v, int which) {
  switch (which) {
  case 0: return v->obs;
  case 1: return v->masks;
  case 2: return v->rewards;
  case 3: return v->dones;
  case 4: return v->infos;
  case 5: return v->agents;
  case 6: return v->actions;
  default: return NULL;
  }
}

double orc_time_run(orc_vec *v, uint64_t steps) {
  struct timespec t0, t1;
  clock_gettime(CLOCK_MONOTONIC, &t0);
  for (uint64_t s = 0; s < steps; ++s) {
    orc_sample(v, v->masks);
    orc_step(v, v->actions);
  }
  clock_gettime(CLOCK_MONOTONIC, &t1);
  return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}

void orc_hidden(orc_vec *v, uint64_t i, uint32_t *out) {
  orc_env *e = &v->env[i];
  out[0] = e->rng; out[1] = e->turn_counter; out[2] = e->agent; out[3] = e->n_in_market;
  for (int p = 0; p < 4; ++p) {
    out[4 + 2 * p] = (uint32_t)e->loc[p][0];
    out[5 + 2 * p] = (uint32_t)e->loc[p][1];
  }
  out[12] = (uint32_t)e->minx; out[13] = (uint32_t)e->miny;
  out[14] = e->n_start_b; out[15] = 0;
}
