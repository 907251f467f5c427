/* cog_device.cuh — device-side game logic of the City of Gold hot path (sm_100a).
 *
 * Execution model: ONE WARP PER ENVIRONMENT.  Every branch of the game (phase,
 * action kind, turn end, done) depends only on per-env values, so control flow is
 * warp-uniform and the 32 lanes are used as a data-parallel axis:
 *   lane l (1..21)  <-> card type l-1   (the five card piles, play/special/remove bits)
 *   lane l (1..18)  <-> shop slot l-1   (stock, market flag, price)
 *   lane d (0..6)   <-> hex direction d (neighbour cache, movement mask)
 *   lane p (0..3)   <-> player p        (hidden per-player fields, P[p] masks)
 * Mask index k of the reference's ActionMask is therefore bit k / lane k with no
 * off-by-one shuffling, masks are bit sets built with __ballot_sync, and weighted
 * draws from a pile are a warp prefix sum + ballot.
 *
 * The code restates the reference's semantics (file:line cited per function,
 * paths relative to the reference tree); it shares no code with it.
 */
#ifndef COG_DEVICE_CUH
#define COG_DEVICE_CUH
#include <stdint.h>

#include "cog_layout.h"
#include "cog_tables.h"

namespace cog {

typedef uint8_t u8;
typedef uint32_t u32;
typedef uint64_t u64;
#define COG_FULL 0xffffffffu
/* branch-layout hints: keep rarely taken paths (episode end + board generation, PRNG
 * rejections) out of the hot instruction stream */
#define COG_UNLIKELY(x) __builtin_expect(!!(x), 0)
#define COG_LIKELY(x) __builtin_expect(!!(x), 1)

/* ------------------------------------------------------------------ tables */
struct card_def { u8 cost, market, single, res0, res1, res2, special; };
struct piece_meta { u8 n_hexes, size, type, difficulty; };
struct xy8 { int8_t x, y; };

namespace tab {
constexpr card_def CARDS[21] = COG_CARDS_INIT;
constexpr u8 SLOT_CARD[18] = COG_SHOP_SLOT_CARD_INIT;

/* per-lane constants, bit layout:
 *  0-2 machete  3-5 paddle  6-8 coin  9 single_use  10-12 special  13-15 card cost
 *  (card type lane-1);  16-18 price and 19 starts_in_market of shop slot lane-1 */
constexpr u32 lane_const(int l) {
  u32 v = 0;
  if (l >= 1 && l <= 21) {
    const card_def c = CARDS[l - 1];
    v |= (u32)c.res0 | ((u32)c.res1 << 3) | ((u32)c.res2 << 6) | ((u32)c.single << 9) |
         ((u32)c.special << 10) | ((u32)c.cost << 13);
  }
  if (l >= 1 && l <= 18) {
    const card_def c = CARDS[SLOT_CARD[l - 1]];
    v |= ((u32)c.cost << 16) | ((u32)c.market << 19);
  }
  return v;
}
constexpr u32 special_mask() { /* bit l: card l-1 has a special action (src/cards.cpp:311-323) */
  u32 m = 0;
  for (int c = 0; c < 21; ++c)
    if (CARDS[c].special) m |= 1u << (c + 1);
  return m;
}
}  // namespace tab

#define COG_LC(l) tab::lane_const(l)
__device__ const u32 LANE_CONST[32] = {
    COG_LC(0),  COG_LC(1),  COG_LC(2),  COG_LC(3),  COG_LC(4),  COG_LC(5),  COG_LC(6),  COG_LC(7),
    COG_LC(8),  COG_LC(9),  COG_LC(10), COG_LC(11), COG_LC(12), COG_LC(13), COG_LC(14), COG_LC(15),
    COG_LC(16), COG_LC(17), COG_LC(18), COG_LC(19), COG_LC(20), COG_LC(21), COG_LC(22), COG_LC(23),
    COG_LC(24), COG_LC(25), COG_LC(26), COG_LC(27), COG_LC(28), COG_LC(29), COG_LC(30), COG_LC(31)};
constexpr u32 COG_SPECIAL_MASK = tab::special_mask();
__constant__ const u8 SLOT_CARD_D[18] = COG_SHOP_SLOT_CARD_INIT;

/* piece geometry in half-hex units (src/map.cpp:446-462), hex codes and piece
 * attributes (:464-695); encodings in tools/extract_tables.py */
__device__ const xy8 COORDS_LARGE[37] = COG_COORDS_LARGE_INIT;
__device__ const xy8 COORDS_SMALL[16] = COG_COORDS_SMALL_INIT;
__device__ const xy8 COORDS_END[3] = COG_COORDS_END_INIT;
/* __constant__: read with warp-uniform indices and feeding control flow (uniform loads) */
__constant__ const piece_meta PIECE_META[COG_N_PIECES_TOTAL] = COG_PIECE_META_INIT;
__device__ const u8 PIECE_HEX[COG_N_PIECES_TOTAL][COG_PIECE_MAX_HEXES] = COG_PIECE_HEX_INIT;
/* connection points per (host size, new size): MapPiece::get_ref_connection_points
 * src/map.cpp:203-263.  kind 0 LARGE->LARGE, 1 LARGE->SMALL, 2 LARGE->TRIPLE, 3 SMALL->LARGE */
__device__ const int8_t CONN_BASE[4][6][2] = {
    {{8, 6}, {6, 8}, {0, 0}, {0, 0}, {0, 0}, {0, 0}},
    {{3, 7}, {5, 5}, {7, 3}, {0, 0}, {0, 0}, {0, 0}},
    {{0, 8}, {0, 0}, {0, 0}, {0, 0}, {0, 0}, {0, 0}},
    {{-7, 10}, {-5, 10}, {-3, 10}, {7, -10}, {5, -10}, {3, -10}}};

/* Footprints as bitmap rows.  The occupancy bitmap of board generation is indexed [x][y] in half-hex
 * units, so a piece shape at rotation r is at most seven rows (distinct x) of a few bits each:
 * row entry = dx (int8) | y0 (int8) << 8 | bits << 16, bit b <-> a hex at (dx, y0 + b).  A candidate
 * position is tested with one shifted AND per row instead of one lookup per hex (37 for a large
 * piece) — and every candidate takes the same seven steps, so the lanes never diverge.
 * bbox = min dx | max dx << 8 | min y << 16 | max y << 24 (int8 each).
 * conn[kind][j][r] = connection point j of table `kind` rotated r times (int8 x | int8 y << 8). */
struct foot_tab {
  u32 row[3][6][7];
  u32 bbox[3][6];
  unsigned short conn[4][6][6];
  u32 bad; /* a shape did not fit seven rows of sixteen bits (checked at compile time) */
};
namespace tab {
constexpr xy8 C_LARGE[37] = COG_COORDS_LARGE_INIT;
constexpr xy8 C_SMALL[16] = COG_COORDS_SMALL_INIT;
constexpr xy8 C_END[3] = COG_COORDS_END_INIT;
constexpr int C_CONN[4][6][2] = {
    {{8, 6}, {6, 8}, {0, 0}, {0, 0}, {0, 0}, {0, 0}},
    {{3, 7}, {5, 5}, {7, 3}, {0, 0}, {0, 0}, {0, 0}},
    {{0, 8}, {0, 0}, {0, 0}, {0, 0}, {0, 0}, {0, 0}},
    {{-7, 10}, {-5, 10}, {-3, 10}, {7, -10}, {5, -10}, {3, -10}}};
constexpr void rot60(int &x, int &y, int times) { /* (x, y) -> (-y, x + y), `times` times */
  for (int t = 0; t < times; ++t) {
    const int nx = -y, ny = x + y;
    x = nx;
    y = ny;
  }
}
constexpr foot_tab make_foot() {
  foot_tab f{};
  for (int s = 0; s < 3; ++s) {
    const xy8 *co = s == 0 ? C_LARGE : (s == 1 ? C_SMALL : C_END);
    const int nh = s == 0 ? 37 : (s == 1 ? 16 : 3);
    for (int r = 0; r < 6; ++r) {
      int xs[40] = {}, ys[40] = {};
      int minx = 127, maxx = -127, miny = 127, maxy = -127;
      for (int h = 0; h < nh; ++h) {
        int x = co[h].x, y = co[h].y;
        rot60(x, y, r);
        xs[h] = x;
        ys[h] = y;
        minx = x < minx ? x : minx; maxx = x > maxx ? x : maxx;
        miny = y < miny ? y : miny; maxy = y > maxy ? y : maxy;
      }
      f.bbox[s][r] = (u32)(minx & 0xff) | ((u32)(maxx & 0xff) << 8) | ((u32)(miny & 0xff) << 16) | ((u32)(maxy & 0xff) << 24);
      int n_rows = 0;
      for (int x = minx; x <= maxx; ++x) {
        int y0 = 127;
        for (int h = 0; h < nh; ++h)
          if (xs[h] == x && ys[h] < y0) y0 = ys[h];
        if (y0 == 127) continue; /* no hex in this row */
        u32 bits = 0;
        for (int h = 0; h < nh; ++h)
          if (xs[h] == x) bits |= 1u << (ys[h] - y0);
        if (n_rows >= 7 || bits > 0xffffu) {
          f.bad = 1;
          continue;
        }
        f.row[s][r][n_rows++] = (u32)(x & 0xff) | ((u32)(y0 & 0xff) << 8) | (bits << 16);
      }
    }
  }
  for (int k = 0; k < 4; ++k)
    for (int j = 0; j < 6; ++j)
      for (int r = 0; r < 6; ++r) {
        int x = C_CONN[k][j][0], y = C_CONN[k][j][1];
        rot60(x, y, r);
        f.conn[k][j][r] = (unsigned short)((x & 0xff) | ((y & 0xff) << 8));
      }
  return f;
}
}  // namespace tab
static_assert(tab::make_foot().bad == 0, "piece footprints must fit seven bitmap rows of sixteen bits");
__device__ const foot_tab FOOT = tab::make_foot();

enum { SIZE_LARGE = 0, SIZE_SMALL = 1, SIZE_TRIPLE = 2 };
enum { TYPE_START = 0, TYPE_TRAVEL = 1, TYPE_END = 2 };
enum { PH_INACTIVE = 0, PH_MOVEMENT = 1, PH_BUYING = 2, PH_MAX = 3 };
enum { REQ_MACHETE = 0, REQ_PADDLE, REQ_COIN, REQ_DISCARD, REQ_REMOVE, REQ_NULL };
enum { SP_NONE = 0, SP_TRANSMIT, SP_DRAW2, SP_DRAW3, SP_DRAW1_REMOVE1, SP_DRAW2_REMOVE2,
       SP_NATIVE, SP_REMOVE_PENDING };
enum { PFL_WON = 1, PFL_MIP = 2, PFL_CARD_FREE = 4, PFL_MOVE_FREE = 8 };

/* hex code: bits 0-2 requirement+1 (0 = none), bits 3-5 n_required, bit 6 is_end */
__device__ __forceinline__ int hex_req(u32 c) { return (c & 7) ? (int)(c & 7) - 1 : (int)REQ_NULL; }
__device__ __forceinline__ int hex_n(u32 c) { return (int)((c >> 3) & 7); }
__device__ __forceinline__ int hex_end(u32 c) { return (int)((c >> 6) & 1); }

/* ----------------------------------------------------------- warp helpers */
__device__ __forceinline__ u32 bcast(u32 v, int src) { return __shfl_sync(COG_FULL, v, src); }
/* inclusive prefix sum over the warp.  shfl.up reports in a predicate whether the source
 * lane exists, so every round is SHFL + one predicated IADD (no lane compares) */
__device__ __forceinline__ u32 warp_incl_scan(u32 v, int lane) {
#ifndef __CUDACC__ /* a host compiler (the warp emulator of tests/emu): the same scan without PTX */
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    u32 t = __shfl_up_sync(COG_FULL, v, d);
    if (lane >= d) v += t;
  }
#else
  (void)lane;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1)
    asm volatile("{ .reg .u32 t; .reg .pred q;\n\t"
                 "shfl.sync.up.b32 t|q, %0, %1, 0, 0xffffffff;\n\t"
                 "@q add.u32 %0, %0, t; }"
                 : "+r"(v) : "r"(d));
#endif
  return v;
}
/* (a & m) | (w & ~m) as ONE LOP3 (the compiler emits two for the C expression) */
__device__ __forceinline__ u32 bitsel(u32 a, u32 m, u32 w) {
#ifndef __CUDACC__
  return (a & m) | (w & ~m);
#else
  u32 d;
  asm("lop3.b32 %0, %1, %2, %3, 0xE2;" : "=r"(d) : "r"(a), "r"(m), "r"(w));
  return d;
#endif
}
/* byte p of a word.  Written with shifts and masks so that a literal p (almost always 0,
 * see "rotated player order" below) folds to one LOP3 / one IADD + one LOP3 */
__device__ __forceinline__ u32 getb(u32 w, int p) { return (w >> (8 * p)) & 0xffu; }
__device__ __forceinline__ u32 setb(u32 w, int p, u32 v) {
  return bitsel(v << (8 * p), 0xffu << (8 * p), w);
}
/* u8 arithmetic inside byte p (wraps like the reference's u_char counters) */
__device__ __forceinline__ u32 addb(u32 w, int p, int d) {
  return bitsel(w + ((u32)d << (8 * p)), 0xffu << (8 * p), w);
}
/* rotate the four bytes of w right by r positions: byte r moves to byte 0 */
__device__ __forceinline__ u32 rotb(u32 w, u32 r) { return __funnelshift_r(w, w, 8u * r); }
__device__ __forceinline__ u32 setbit(u32 w, int b, bool v) {
  return (w & ~(1u << b)) | ((u32)v << b);
}

/* -------------------------------------------------------------------- PRNG */
/* std::default_random_engine = minstd_rand0: x <- 16807 x mod (2^31 - 1)
 * (libstdc++ bits/random.h:1593); seed(): bits/random.tcc:119-126 */
__host__ __device__ __forceinline__ u32 minstd_seed(u64 s) {
  u64 x = s % 2147483647ull;
  return (u32)(x ? x : 1);
}
__device__ __forceinline__ u32 minstd_next(u32 &s) {
  u64 t = (u64)s * 16807ull; /* < 2^46: one Mersenne fold suffices */
  u32 r = (u32)(t & 0x7fffffffu) + (u32)(t >> 31);
  if (r >= 0x7fffffffu) r -= 0x7fffffffu;
  s = r;
  return r;
}
/* std::uniform_int_distribution<size_t>(0, n-1)(minstd_rand0): the "two divisions"
 * downscaling branch of libstdc++ bits/uniform_int_dist.h:332-340 (the engine's
 * range 2^31-3 is not 2^k-1, so Lemire's path is not taken):
 *     scaling = 2147483645 / n;  past = n * scaling;
 *     do r = g() - 1; while (r >= past);  return r / scaling;
 * A draw is consumed even for n == 1; rejection makes the count data dependent.
 * Both divisions are avoided for n <= 32 (every draw of the step/sample path): scaling
 * and past come from a constant table indexed by the warp-uniform n, and r / scaling is
 * a multiply-high by magic = floor(2^36 / scaling) + 1 — an over-estimate by less than
 * r / 2^36 < 1/32, so at most one too large — followed by one exact correction. */
struct alignas(16) ud_entry { u32 scaling, past, magic, pad; };
namespace tab {
constexpr ud_entry ud_make(u32 n) {
  if (n == 0) return ud_entry{0, 0, 0, 0};
  const u32 s = 2147483645u / n;
  return ud_entry{s, n * s, (u32)((1ull << 36) / s) + 1u, 0};
}
}  // namespace tab
#define COG_UD(n) tab::ud_make(n)
__constant__ const ud_entry UD_TAB[33] = {
    COG_UD(0),  COG_UD(1),  COG_UD(2),  COG_UD(3),  COG_UD(4),  COG_UD(5),  COG_UD(6),  COG_UD(7),  COG_UD(8),
    COG_UD(9),  COG_UD(10), COG_UD(11), COG_UD(12), COG_UD(13), COG_UD(14), COG_UD(15), COG_UD(16), COG_UD(17),
    COG_UD(18), COG_UD(19), COG_UD(20), COG_UD(21), COG_UD(22), COG_UD(23), COG_UD(24), COG_UD(25), COG_UD(26),
    COG_UD(27), COG_UD(28), COG_UD(29), COG_UD(30), COG_UD(31), COG_UD(32)};

/* r / e.scaling for r < e.past, without a division (exhaustively checked against the division for
 * every n <= 32 and every engine output by tests/kat/prng_kat.cu) */
__device__ __forceinline__ u32 ud_quotient(u32 r, const ud_entry &e) {
  u32 q = __umulhi(r, e.magic) >> 4;
  if (q * e.scaling > r) --q;
  return q;
}
__device__ __forceinline__ u32 uniform_draw(u32 &s, u32 n) { /* n >= 1 */
  if (COG_LIKELY(n <= 32)) { /* n == 1: scaling = past = 2147483645, the result is 0 */
    const ud_entry e = UD_TAB[n];
    u32 r;
    do {
      r = minstd_next(s) - 1u;
    } while (COG_UNLIKELY(r >= e.past));
    return ud_quotient(r, e);
  }
  const u32 scaling = 2147483645u / n;
  const u32 past = n * scaling;
  u32 r;
  do {
    r = minstd_next(s) - 1u;
  } while (r >= past);
  return r / scaling;
}

/* position of the (j+1)-th set bit of v (j < popc(v)): clear the j lowest set bits */
__device__ __forceinline__ u32 nth_set_bit(u32 v, u32 j) {
  for (u32 t = 0; t < j; ++t) v &= v - 1u;
  return (u32)__ffs((int)v) - 1u;
}
/* the same without a loop: lane l owns bit l and knows how many set bits lie below it */
__device__ __forceinline__ u32 nth_set_bit_warp(u32 v, u32 j, int lane) {
  const u32 below = v & ((1u << lane) - 1u);
  const u32 me = __ballot_sync(COG_FULL, ((v >> lane) & 1u) && (u32)__popc(below) == j);
  return (u32)__ffs((int)me) - 1u;
}

/* one +60 degree step in half-hex axial coordinates: (x, y) -> (-y, x + y)
 * (rotate_point src/map.cpp:17-37 through src/geometry.cpp:3-17; exact on the
 * half lattice).  `times` keeps C remainder semantics like the reference. */
/* The six rotations as integer matrices (a b; c d): k steps of (x, y) -> (-y, x + y).
 * Packed 2 bits per coefficient (value + 1) so that one constant serves all of them. */
struct rot6 { int a, b, c, d; };
__device__ __forceinline__ rot6 rot_matrix(int times) {
  int t = times % 6;
  if (t < 0) t += 6;
  /* t:      0         1          2          3          4          5
   * a b   1  0     0 -1      -1 -1      -1  0       0  1       1  1
   * c d   0  1     1  1       1  0       0 -1      -1 -1      -1  0   */
  const u32 A = 0x2 | (0x1 << 2) | (0x0 << 4) | (0x0 << 6) | (0x1 << 8) | (0x2 << 10);
  const u32 Bc = 0x1 | (0x0 << 2) | (0x0 << 4) | (0x1 << 6) | (0x2 << 8) | (0x2 << 10);
  const u32 C = 0x1 | (0x2 << 2) | (0x2 << 4) | (0x1 << 6) | (0x0 << 8) | (0x0 << 10);
  const u32 D = 0x2 | (0x2 << 2) | (0x1 << 4) | (0x0 << 6) | (0x0 << 8) | (0x1 << 10);
  rot6 m;
  m.a = (int)((A >> (2 * t)) & 3u) - 1;
  m.b = (int)((Bc >> (2 * t)) & 3u) - 1;
  m.c = (int)((C >> (2 * t)) & 3u) - 1;
  m.d = (int)((D >> (2 * t)) & 3u) - 1;
  return m;
}
__device__ __forceinline__ void rot_apply(const rot6 &m, int &x, int &y) {
  const int nx = m.a * x + m.b * y, ny = m.c * x + m.d * y;
  x = nx;
  y = ny;
}
__device__ __forceinline__ void rotate_pt(int &x, int &y, int times) { rot_apply(rot_matrix(times), x, y); }

/* ------------------------------------------------------------ env in regs */
/* parameters of vec_cog_env::reset(seed, ...) — identical for every env of a handle */
struct Params {
  u32 max_steps;
  u32 n_players, n_pieces, difficulty;
  u32 flags; /* COG_PRM_* */
};
/* single-env semantics (cog_env::step, src/environment.cpp:91-95,187-207): a finished episode is NOT
 * reset by the step that ends it; the env keeps its terminal state and ignores further steps
 * (dead_step) until reset() */
enum { COG_PRM_NO_AUTORESET = 1 };

/* device buffers of one handle */
struct Buffers {
  u32 *hidden;   /* [n][COG_HID_WORDS]  private state */
  u8 *grid;      /* [n][48*48]          one hex code per cell, indexed like the map obs */
  u8 *obs;       /* [n] ObsData */
  u8 *masks;     /* [n] ActionMask  (selected_action_masks, "S") */
  float *rewards;/* [n][4] */
  u8 *dones;     /* [n] */
  u8 *infos;     /* [n] Info */
  u8 *agents;    /* [n] */
  unsigned long long *stats; /* [COG_STAT_COUNT] */
  int *err;      /* [0] first error code, [1] env index */
  unsigned *park; /* [0] envs ever parked by k_step_fused, [1] envs k_step_tail picked up again,
                   * [2] env tickets handed out by the running k_step_fused, [3] by the running k_step_tail */
};

/* ROTATED PLAYER ORDER.  The seven words that hold one byte per player (the five piles,
 * the neighbour cache and info_steps) are kept — in registers and in the hidden record —
 * rotated by the current agent: byte r belongs to player (agent + r) & 3.  The player whose
 * turn it is therefore always sits in byte 0, every pile update of the hot path addresses a
 * literal byte, and the rotation is paid once per turn change (seven funnel shifts).
 * Functions below take `rp` = a player's RELATIVE byte index, rel(E, p). */
struct Env {
  /* lane l <-> card type l-1; byte r = player (agent + r) & 3 */
  u32 draw, hand, active, played, discard;
  u32 shop; /* lane l <-> slot l-1: bits 0-7 stock, bit 8 in_market */
  u32 nbr;  /* lane d <-> direction d: byte r = hex code next to player (agent + r) & 3 */
  u32 lc;   /* LANE_CONST[lane]: card / shop-slot constants of this lane (not stored) */
  /* lane p <-> player p.  pf / pc of the CURRENT agent are stale here: the live copies
   * are the uniform cpf / cpc below (written back when the turn passes on) */
  u32 pf, pc, pmv, ploc, Pplay, Prem, Pms;
  /* warp-uniform */
  u32 rng, turn, agent, phase, n_in_market;
  u32 cpf, cpc;   /* COG_PF_FLAGS / COG_PF_COUNTS words of the current agent */
  u32 info_steps; /* Info.agent_infos[k].steps_taken, rotated like the piles */
  /* S = selected_action_masks.  play_special is not stored: in S and in every P[p] it equals
   * play & (special cards | bit 0) — each writer of the reference sets play_special[k] =
   * play[k] && is_special(k - 1) and both [0] are always true (src/cards.cpp:183-290,
   * src/player.cpp:191-206, include/api.h:104-118) */
  u32 Splay, Srem, Sms;
  int res[3]; /* current_resources * 2 */
  int minx, miny;
  u32 dead; /* COG_ES_DEAD: ignore steps until reset() (finished single env / failed board generation) */
};

__device__ __forceinline__ u32 spec_of(u32 play) { return play & (COG_SPECIAL_MASK | 1u); }
__device__ __forceinline__ int rel(const Env &E, int p) { return (p - (int)E.agent) & 3; }
/* re-base the rotated words after the agent advanced by `r` seats (or, with
 * r = (4 - agent) & 3, back to absolute player order) */
__device__ __forceinline__ void env_rotate(Env &E, u32 r) {
  E.draw = rotb(E.draw, r);
  E.hand = rotb(E.hand, r);
  E.active = rotb(E.active, r);
  E.played = rotb(E.played, r);
  E.discard = rotb(E.discard, r);
  E.nbr = rotb(E.nbr, r);
  E.info_steps = rotb(E.info_steps, r);
}
__device__ __forceinline__ int loc_x(u32 w) { return (int)(int16_t)(w & 0xffffu); }
__device__ __forceinline__ int loc_y(u32 w) { return (int)(int16_t)(w >> 16); }
__device__ __forceinline__ u32 loc_pack(int x, int y) {
  return ((u32)x & 0xffffu) | ((u32)y << 16);
}

/* L2 = true: read through L2 (ld.global.cg).  For a record the SAME launch wrote earlier —
 * lines that other lanes stored may still sit, stale, in this SM's L1. */
template <bool L2 = false>
__device__ __forceinline__ u32 hid_word(const u32 *rec, int w) {
  if (L2) return __ldcg(rec + w);
  return rec[w];
}
template <bool L2 = false>
__device__ __forceinline__ uint4 hid_vec(const u32 *rec, int w) {
  if (L2) return __ldcg(reinterpret_cast<const uint4 *>(rec + w));
  return *reinterpret_cast<const uint4 *>(rec + w);
}
/* No shuffles: the seven per-lane lines are one coalesced load each; the player line and the
 * env scalars are read as 16-byte vectors straight from the addresses every lane needs (lane l
 * wants the fields of player l & 3 and all sixteen scalars) — same-address lanes share one
 * transaction, so this is 6 vector loads instead of 2 loads + 27 shuffles. */
/* the env scalars (line 8) and this lane's card / shop-slot constants */
template <bool L2 = false>
__device__ __forceinline__ void env_load_scalars(Env &E, const u32 *rec, int lane) {
  E.lc = LANE_CONST[lane];
  const uint4 e0 = hid_vec<L2>(rec, COG_HL_ENV * 32), e1 = hid_vec<L2>(rec, COG_HL_ENV * 32 + 4);
  const uint4 e2 = hid_vec<L2>(rec, COG_HL_ENV * 32 + 8), e3 = hid_vec<L2>(rec, COG_HL_ENV * 32 + 12);
  static_assert(COG_ES_MISC == 0 && COG_ES_RES2 == 3 && COG_ES_SPLAY == 4 && COG_ES_CPC == 7 && COG_ES_RNG == 8 &&
                    COG_ES_SREM == 11 && COG_ES_SMS == 12 && COG_ES_PENDING == 14 && COG_ES_DEAD == 15,
                "env_load_scalars / env_store_scalars follow cog_layout.h");
  E.agent = e0.x & 0xffu; /* COG_ES_MISC */
  E.phase = (e0.x >> 8) & 0xffu;
  E.n_in_market = (e0.x >> 16) & 0xffu;
  E.res[0] = (int)e0.y; /* COG_ES_RES0.. */
  E.res[1] = (int)e0.z;
  E.res[2] = (int)e0.w;
  E.Splay = e1.x;      /* COG_ES_SPLAY */
  E.info_steps = e1.y; /* COG_ES_INFOSTEPS */
  E.cpf = e1.z;        /* COG_ES_CPF */
  E.cpc = e1.w;        /* COG_ES_CPC */
  E.rng = e2.x;        /* COG_ES_RNG */
  E.turn = e2.y;       /* COG_ES_TURN */
  E.minx = loc_x(e2.z); /* COG_ES_MIN */
  E.miny = loc_y(e2.z);
  E.Srem = e2.w;       /* COG_ES_SREM */
  E.Sms = e3.x;        /* COG_ES_SMS */
  E.dead = e3.w;       /* COG_ES_DEAD */
}
/* the per-lane lines in `lines` (bit COG_HL_*) and, with bit COG_HL_PLAYER, the player line */
template <u32 LINES, bool L2 = false>
__device__ __forceinline__ void env_load_lines(Env &E, const u32 *rec, int lane) {
  if (LINES & (1u << COG_HL_DRAW)) E.draw = hid_word<L2>(rec, COG_HL_DRAW * 32 + lane);
  if (LINES & (1u << COG_HL_HAND)) E.hand = hid_word<L2>(rec, COG_HL_HAND * 32 + lane);
  if (LINES & (1u << COG_HL_ACTIVE)) E.active = hid_word<L2>(rec, COG_HL_ACTIVE * 32 + lane);
  if (LINES & (1u << COG_HL_PLAYED)) E.played = hid_word<L2>(rec, COG_HL_PLAYED * 32 + lane);
  if (LINES & (1u << COG_HL_DISCARD)) E.discard = hid_word<L2>(rec, COG_HL_DISCARD * 32 + lane);
  if (LINES & (1u << COG_HL_SHOP)) E.shop = hid_word<L2>(rec, COG_HL_SHOP * 32 + lane);
  if (LINES & (1u << COG_HL_NBR)) E.nbr = hid_word<L2>(rec, COG_HL_NBR * 32 + lane);
  if (LINES & (1u << COG_HL_PLAYER)) {
    const int pbase = COG_HL_PLAYER * 32 + COG_PF_STRIDE * (lane & 3);
    const uint4 p0 = hid_vec<L2>(rec, pbase), p1 = hid_vec<L2>(rec, pbase + 4);
    E.pf = p0.x;    /* COG_PF_FLAGS */
    E.pc = p0.y;    /* COG_PF_COUNTS */
    E.pmv = p0.z;   /* COG_PF_MOVES */
    E.ploc = p0.w;  /* COG_PF_LOC */
    E.Pplay = p1.x; /* COG_PF_MPLAY */
    E.Prem = p1.z;  /* COG_PF_MREM */
    E.Pms = p1.w;   /* COG_PF_MMS */
  }
}
enum : u32 { COG_LINES_ALL = 0xffu /* lines 0..7; the env scalars have their own loader */ };
/* No shuffles: the seven per-lane lines are one coalesced load each; the player line and the
 * env scalars are read as 16-byte vectors straight from the addresses every lane needs (lane l
 * wants the fields of player l & 3 and all sixteen scalars) — same-address lanes share one
 * transaction, so this is 6 vector loads instead of 2 loads + 27 shuffles. */
template <bool L2 = false>
__device__ __forceinline__ void env_load(Env &E, const u32 *rec, int lane) {
  env_load_lines<COG_LINES_ALL, L2>(E, rec, lane);
  env_load_scalars<L2>(E, rec, lane);
}
/* the same values, provably warp-uniform (a load is not, a shuffle from a fixed lane is): lets ptxas
 * drop the divergence guards around the step logic and move scalar work to the uniform datapath */
__device__ __forceinline__ void env_uniform_scalars_hot(Env &E) { /* what every kind of step reads */
#ifndef COG_NO_UNIFORM_PROOF
  E.turn = bcast(E.turn, 0); E.agent = bcast(E.agent, 0);
  E.phase = bcast(E.phase, 0); E.n_in_market = bcast(E.n_in_market, 0);
  E.cpf = bcast(E.cpf, 0); E.cpc = bcast(E.cpc, 0); E.info_steps = bcast(E.info_steps, 0);
  E.Splay = bcast(E.Splay, 0); E.Srem = bcast(E.Srem, 0); E.Sms = bcast(E.Sms, 0);
  E.res[0] = (int)bcast((u32)E.res[0], 0); E.res[1] = (int)bcast((u32)E.res[1], 0);
  E.res[2] = (int)bcast((u32)E.res[2], 0);
  E.dead = bcast(E.dead, 0);
#endif
}
__device__ __forceinline__ void env_uniform_scalars_cold(Env &E) { /* RNG and board origin: draws, moves, resets */
#ifndef COG_NO_UNIFORM_PROOF
  E.rng = bcast(E.rng, 0);
  E.minx = (int)bcast((u32)E.minx, 0); E.miny = (int)bcast((u32)E.miny, 0);
#endif
}
__device__ __forceinline__ void env_uniform_scalars(Env &E) {
  env_uniform_scalars_hot(E);
  env_uniform_scalars_cold(E);
}

/* HOT_ONLY: the first two vectors (what a card play / a phase change touches) */
template <bool HOT_ONLY = false>
__device__ __forceinline__ void env_store_scalars(const Env &E, u32 *rec, int lane, u32 pending = 0, u32 dead = 0) {
  if (lane == 0) {
    uint4 *ev = reinterpret_cast<uint4 *>(rec + COG_HL_ENV * 32);
    ev[0] = make_uint4(E.agent | (E.phase << 8) | (E.n_in_market << 16), (u32)E.res[0], (u32)E.res[1], (u32)E.res[2]);
    ev[1] = make_uint4(E.Splay, E.info_steps, E.cpf, E.cpc);
    if (!HOT_ONLY) {
      ev[2] = make_uint4(E.rng, E.turn, loc_pack(E.minx, E.miny), E.Srem);
      ev[3] = make_uint4(E.Sms, 0u, pending, dead);
    }
  }
}
__device__ __forceinline__ void env_store(const Env &E, u32 *rec, int lane, u32 pending = 0, u32 dead = 0) {
  rec[COG_HL_DRAW * 32 + lane] = E.draw;
  rec[COG_HL_HAND * 32 + lane] = E.hand;
  rec[COG_HL_ACTIVE * 32 + lane] = E.active;
  rec[COG_HL_PLAYED * 32 + lane] = E.played;
  rec[COG_HL_DISCARD * 32 + lane] = E.discard;
  rec[COG_HL_SHOP * 32 + lane] = E.shop;
  rec[COG_HL_NBR * 32 + lane] = E.nbr;
  if (lane < 4) {
    uint4 *pl = reinterpret_cast<uint4 *>(rec + COG_HL_PLAYER * 32 + COG_PF_STRIDE * lane);
    pl[0] = make_uint4(E.pf, E.pc, E.pmv, E.ploc);
    pl[1] = make_uint4(E.Pplay, 0u, E.Prem, E.Pms);
  }
  env_store_scalars(E, rec, lane, pending, dead);
}

/* ------------------------------------------------------------------ export */
/* one ActionMask record (92 payload bytes of bool) from its four bit sets:
 * lane w < 23 expands 4 mask bits into one 32-bit word -> one coalesced store */
__device__ __forceinline__ void export_mask(u8 *dst, u32 play, u32 spec, u32 rem, u32 ms, int lane) {
  const u64 lo = (u64)play | ((u64)spec << 22) | ((u64)rem << 44);
  const u32 hi = (rem >> 20) | (ms << 2);
  if (lane < 23) {
    const u32 n = lane < 16 ? (u32)(lo >> (4 * lane)) & 15u : (hi >> (4 * (lane - 16))) & 15u;
    reinterpret_cast<u32 *>(dst)[lane] = (n & 1u) | ((n & 2u) << 7) | ((n & 4u) << 14) | ((n & 8u) << 21);
  }
}

/* everything a step can change in ObsData/ActionMask/agent_selection/Info: phase, resources, shop
 * stock, S, and per player p DeckObs[p] (bit p of `pmask`) and P[p] (bit 4 + p) — a turn end changes
 * the old agent's piles and mask but only the mask of the new one */
/* with_agent = false: vec_cog_env::reset() / reset(seed, ...) leave agent_selections[] alone — the array is
 * only written by step_single (include/vec_environment.h:33-44,53-61), so after an explicit reset it
 * still shows the agent of the last step */
__device__ __forceinline__ void export_dynamic(const Env &E, const Buffers &B, u64 i, u32 pmask, int lane,
                                               bool with_agent = true) {
  u8 *obs = B.obs + i * COGL_OBS_SIZE;
  /* phase (+ 3 bytes of record padding) and the three resource floats are 16 contiguous,
   * 16-byte aligned bytes: one vector store of lane 0 */
  static_assert(COGL_OBS_PHASE % 16 == 0 && COGL_OBS_RES == COGL_OBS_PHASE + 4, "shared tail of ObsData");
  if (lane == 0)
    *reinterpret_cast<uint4 *>(obs + COGL_OBS_PHASE) =
        make_uint4(E.phase, __float_as_uint(0.5f * (float)E.res[0]), __float_as_uint(0.5f * (float)E.res[1]),
                   __float_as_uint(0.5f * (float)E.res[2]));
  if (lane >= 1 && lane <= 18) obs[COGL_OBS_SHOP + lane - 1] = (u8)(E.shop & 0xffu);
  for (int p = 0; p < 4; ++p) {
    if (!((pmask >> p) & 0x11u)) continue;
    u8 *pd = obs + COGL_OBS_PLAYER + COGL_PLAYER_STRIDE * p;
    const int rp = rel(E, p);
    if (((pmask >> p) & 1u) && lane >= 1 && lane <= 21) {
      pd[COGL_DK_DRAW + lane - 1] = (u8)getb(E.draw, rp);
      pd[COGL_DK_HAND + lane - 1] = (u8)getb(E.hand, rp);
      pd[COGL_DK_ACTIVE + lane - 1] = (u8)getb(E.active, rp);
      pd[COGL_DK_PLAYED + lane - 1] = (u8)getb(E.played, rp);
      pd[COGL_DK_DISCARD + lane - 1] = (u8)getb(E.discard, rp);
    }
    if ((pmask >> p) & 0x10u) {
      const u32 Pplay = bcast(E.Pplay, p);
      export_mask(pd + COGL_PD_MASK, Pplay, spec_of(Pplay), bcast(E.Prem, p), bcast(E.Pms, p), lane);
    }
  }
  export_mask(B.masks + i * COGL_MASK_SIZE, E.Splay, spec_of(E.Splay), E.Srem, E.Sms, lane);
  if (with_agent && lane == 0) B.agents[i] = (u8)E.agent;
  if (lane < 4)
    B.infos[i * COGL_INFO_SIZE + COGL_INFO_AGENT + COGL_INFO_AGENT_STRIDE * lane] =
        (u8)getb(E.info_steps, rel(E, lane));
}

/* ---------------------------------------------------------- compact records */
/* ring of compact records (layout: cog_layout.h) that a fused launch appends to, one record
 * per env-step, env-major: step t of env i lives at ring + (i * T + t % T) * COGL_CR_SIZE, so a
 * warp streams its env's trajectory into consecutive 256-byte records */
struct Rollout {
  u8 *ring;
  u32 T;     /* slots per env */
  u32 slot0; /* slot of the launch's first step */
  u32 t0;    /* index of that step since the ring was attached */
};

/* The part of a record that is known BEFORE the step: the state the current agent acts from,
 * the mask heads its action is sampled from and the action.  No lane ever selects between
 * values: the warp-uniform words leave as 16-byte stores of lane 0 alone, the per-lane data
 * (cards: lanes 1..21, positions: lanes 0..3) as one store each under its own lane predicate. */
__device__ __forceinline__ void compact_store_pre(u8 *rec, const Env &E, const Params &prm, u32 play, u32 spec,
                                                  u32 rem, u32 ms, u32 act, u32 a_shop, u32 t, int lane) {
  const u32 hx = __byte_perm(E.agent | (E.phase << 8), (u32)E.res[0], 0x5410);
  const u32 hy = __byte_perm((u32)E.res[1], (u32)E.res[2], 0x5410);
  if (lane == 0) *reinterpret_cast<uint4 *>(rec + COGL_CR_HEAD) = make_uint4(hx, hy, act, a_shop);
  if (lane == 0) *reinterpret_cast<uint4 *>(rec + COGL_CR_MASK_PLAY) = make_uint4(play, rem, ms, spec);
  if (lane == 0) *reinterpret_cast<uint2 *>(rec + COGL_CR_TURN) = make_uint2(E.turn, t);
  /* the cell of the hex grid a player stands on (Map::get_from_array indexing).  A player's
   * hex is never left of / below the board's minimum and both differences are even, so the
   * two 16-bit halves of the packed location subtract (offset-binary: no borrow between the
   * halves), halve and increment as one word. */
  const u32 mn = __byte_perm((u32)E.minx, (u32)E.miny, 0x5410);
  const u32 h = (((E.ploc ^ 0x80008000u) - (mn ^ 0x80008000u)) >> 1) + 0x00010001u;
  const u32 g = __byte_perm(h, 0u, 0x4420); /* row | col << 8 */
  if (lane < 4)
    reinterpret_cast<unsigned short *>(rec + COGL_CR_LOC)[lane] = (unsigned short)(lane < (int)prm.n_players ? g : 0u);
  /* byte 0 (the current agent) of the five pile words and of the shop word (zero in the lanes
   * that are no shop slot): three PRMT for the low word, one PRMT + mask for the high */
  const u32 x = __byte_perm(__byte_perm(E.draw, E.hand, 0x0040), __byte_perm(E.active, E.played, 0x0040), 0x5410);
  const u32 y = __byte_perm(E.discard, E.shop, 0x0040) & 0xffffu;
  if ((unsigned)(lane - 1) < 21u) reinterpret_cast<uint2 *>(rec + COGL_CR_CARDS - 8)[lane] = make_uint2(x, y);
}
/* the outcome of the step taken from that state: done (+ "filled"), rewards zero — the
 * kernel that finishes an episode overwrites them (compact_store_rewards) */
__device__ __forceinline__ void compact_store_post(u8 *rec, u32 done, u32 filled, int lane) {
  if (lane < 2)
    *reinterpret_cast<uint4 *>(rec + COGL_CR_DONE + 16 * lane) =
        make_uint4(lane == 0 ? done | (filled << 8) : 0u, 0u, 0u, 0u);
}
/* the rewards of an episode end (get_reward src/environment.cpp:281-288) into the record of
 * the step that ended it.  Call after env_finish_episode (absolute player order). */
__device__ __forceinline__ void compact_store_rewards(u8 *rec, const Env &E, const Params &prm, int lane) {
  const int np = (int)prm.n_players < 4 ? (int)prm.n_players : 4;
  const bool won = lane < 4 && (E.pf & PFL_WON);
  const int n_winners = __popc(__ballot_sync(COG_FULL, won));
  __syncwarp(); /* lane 1 zeroed these words in compact_store_post */
  if (lane < 4)
    reinterpret_cast<float *>(rec + COGL_CR_REWARDS)[lane] =
        lane < np ? (float)((int)prm.n_players * (won ? 1 : 0) - n_winners) : 0.0f;
}

/* ------------------------------------------------------------------- decks */
__device__ __forceinline__ int res_get(const Env &E, int k) {
  return k == 0 ? E.res[0] : (k == 1 ? E.res[1] : E.res[2]);
}
__device__ __forceinline__ void res_clear(Env &E) { E.res[0] = E.res[1] = E.res[2] = 0; }
__device__ __forceinline__ void res_set(Env &E, int k, int v) {
  if (k == 0) E.res[0] = v;
  else if (k == 1) E.res[1] = v;
  else E.res[2] = v;
}

/* weighted pick without replacement: the card whose cumulative count first exceeds
 * `target` (the `while (target >= pile[card])` scans of src/cards.cpp:192-197,
 * src/player.cpp:94-99,117-122) as a warp prefix sum + ballot.  Returns the lane
 * (= card type + 1) or 0 when target is beyond the pile (UB in the reference). */
__device__ __forceinline__ int pile_pick(u32 pile, int p, u32 target, int lane) {
  const u32 incl = warp_incl_scan(getb(pile, p), lane);
  const u32 hit = __ballot_sync(COG_FULL, incl > target);
  return hit ? __ffs((int)hit) - 1 : 0;
}

/* Deck::draw src/cards.cpp:183-211 (with move_discard_to_draw :234-240).
 * `pc` = the player's COG_PF_COUNTS word (uniform copy, updated in place); here and in
 * the deck/player functions below `p` is the player's RELATIVE byte index (0 = the agent). */
__device__ __forceinline__ void deck_draw(Env &E, u32 &pc, int p, u32 n, int lane) {
  u32 n_in_draw = getb(pc, 2);
  if (n_in_draw < n) {
    const u32 d = getb(E.discard, p);
    E.draw = addb(E.draw, p, (int)d);
    E.discard = setb(E.discard, p, 0);
    n_in_draw = (n_in_draw + __reduce_add_sync(COG_FULL, d)) & 0xffu;
  }
  if (n_in_draw < n) n = n_in_draw;
  /* one prefix sum for the whole batch: drawing a copy of card c lowers the cumulative
   * counts of the lanes from c upwards by one (the picked pile entry is >= 1) */
  u32 incl = n ? warp_incl_scan(getb(E.draw, p), lane) : 0u;
  for (u32 i = 0; i < n; ++i) {
    const u32 target = uniform_draw(E.rng, n_in_draw);
    const u32 hit = __ballot_sync(COG_FULL, incl > target);
    if (COG_LIKELY(hit)) { /* no hit: target beyond the pile (UB in the reference) */
      const int cl = __ffs((int)hit) - 1;
      if (lane == cl) {
        E.draw = addb(E.draw, p, -1);
        E.hand = addb(E.hand, p, 1);
      }
      if (lane >= cl) --incl;
      E.Splay |= 1u << cl;
    }
    n_in_draw = (n_in_draw - 1) & 0xffu;
  }
  pc = setb(pc, 2, n_in_draw);
  pc = setb(pc, 0, getb(pc, 0) + n);
}

/* hand -> active; Deck::activate src/cards.cpp:242-253 */
__device__ __forceinline__ void deck_activate(Env &E, u32 &pc, int p, int idx, int lane) {
  const int cl = idx + 1;
  pc = setb(pc, 0, getb(pc, 0) - 1);
  pc = setb(pc, 1, getb(pc, 1) + 1);
  pc = setb(pc, 3, (u32)idx);
  const u32 prev = bcast(getb(E.hand, p), cl);
  if (lane == cl) {
    E.hand = addb(E.hand, p, -1);
    E.active = addb(E.active, p, 1);
  }
  E.Splay = setbit(E.Splay, cl, prev > 1);
}
/* Deck::play_last_activated src/cards.cpp:255-261 */
__device__ __forceinline__ void deck_play_last_activated(Env &E, u32 &pc, int p, int lane) {
  const int cl = (int)getb(pc, 3) + 1;
  pc = setb(pc, 1, getb(pc, 1) - 1);
  if (lane == cl) {
    E.active = addb(E.active, p, -1);
    if (!((E.lc >> 9) & 1u)) E.played = addb(E.played, p, 1);
  }
}
/* Deck::play_immediate src/cards.cpp:263-273 */
__device__ __forceinline__ void deck_play_immediate(Env &E, u32 &pc, int p, int idx, int lane) {
  const int cl = idx + 1;
  pc = setb(pc, 0, getb(pc, 0) - 1);
  const u32 prev = bcast(getb(E.hand, p), cl);
  if (lane == cl) {
    E.hand = addb(E.hand, p, -1);
    E.played = addb(E.played, p, 1);
  }
  E.Splay = setbit(E.Splay, cl, prev > 1);
}
/* Deck::remove_immediate src/cards.cpp:281-290 */
__device__ __forceinline__ void deck_remove_immediate(Env &E, u32 &pc, int p, int idx, int lane) {
  const int cl = idx + 1;
  pc = setb(pc, 0, getb(pc, 0) - 1);
  const u32 prev = bcast(getb(E.hand, p), cl);
  if (lane == cl) E.hand = addb(E.hand, p, -1);
  const bool still = prev > 1;
  E.Srem = setbit(E.Srem, cl, ((E.Srem >> cl) & 1u) && still);
  if (!still) E.Splay &= ~(1u << cl);
}

/* ------------------------------------------------------------------ player */
/* Player::disable_playing src/player.cpp:191-196 */
__device__ __forceinline__ void player_disable_playing(Env &E) {
  E.Splay = 1u;
}
/* Player::enable_playing src/player.cpp:198-206: play[k] = hand[k-1] > 0 is one ballot */
__device__ __forceinline__ void player_enable_playing(Env &E, int p, int lane) {
  E.Srem = 1u;
  const u32 have = __ballot_sync(COG_FULL, lane >= 1 && lane <= 21 && getb(E.hand, p) > 0);
  E.Splay = (E.Splay & 1u) | have;
}
/* Player::discard_cards src/player.cpp:85-104 (no clamp) / remove_cards :106-131 (clamped) */
__device__ __forceinline__ void player_drop_active(Env &E, u32 &pc, int p, u32 n, bool remove, int lane) {
  const u32 n_act = getb(pc, 1);
  if (remove && n > n_act) n = n_act;
  for (u32 i = 0; i < n; ++i) {
    if (n_act <= i) break; /* the reference would draw from an empty range (UB) */
    const u32 target = uniform_draw(E.rng, n_act - i);
    const int cl = pile_pick(E.active, p, target, lane);
    pc = setb(pc, 1, getb(pc, 1) - 1);
    if (cl && lane == cl) {
      E.active = addb(E.active, p, -1);
      if (!remove) E.discard = addb(E.discard, p, 1);
    }
  }
}
/* Player::handle_requirement src/player.cpp:141-162 */
__device__ __forceinline__ void player_handle_requirement(Env &E, u32 &pf, u32 &pc, int p, int req,
                                                          u32 n, int lane) {
  if (req < 3) {
    const int left = res_get(E, req) - 2 * (int)n;
    res_clear(E);
    res_set(E, req, left);
    if (!(pf & PFL_MIP)) {
      deck_play_last_activated(E, pc, p, lane);
      pf |= PFL_MIP;
    }
  } else if (req == REQ_REMOVE) {
    player_drop_active(E, pc, p, n, true, lane);
    res_clear(E);
    pf &= ~(u32)PFL_MIP;
  } else if (req == REQ_DISCARD) {
    player_drop_active(E, pc, p, n, false, lane);
    res_clear(E);
    pf &= ~(u32)PFL_MIP;
  }
}
/* Player::end_turn src/player.cpp:170-180 (+ Deck::discard_all_active/_played
 * src/cards.cpp:219-232); P[p] = S is save_actionmask :166 */
__device__ __forceinline__ void player_end_turn(Env &E, u32 &pc, int p_abs, int p, int lane) {
  pc = setb(pc, 1, 0);
  E.discard = addb(E.discard, p, (int)(getb(E.active, p) + getb(E.played, p)));
  E.active = setb(E.active, p, 0);
  E.played = setb(E.played, p, 0);
  const int n_draw = 4 - (int)getb(pc, 0);
  if (n_draw > 0) deck_draw(E, pc, p, (u32)n_draw, lane);
  res_clear(E);
  if (lane == p_abs) {
    E.Pplay = E.Splay;
    E.Prem = E.Srem;
    E.Pms = E.Sms;
  }
}

/* -------------------------------------------------------------------- shop */
/* Shop::get src/cards.cpp:136-142; returns the card type */
__device__ __forceinline__ int shop_get(Env &E, int idx, int lane) {
  const int sl = idx + 1;
  const u32 w = bcast(E.shop, sl);
  const u32 stock = ((w & 0xffu) - 1u) & 0xffu;
  u32 inm = (w >> 8) & 1u;
  if (!stock && inm) {
    inm = 0;
    E.n_in_market = (E.n_in_market - 1u) & 0xffu;
  }
  if (lane == sl) E.shop = stock | (inm << 8);
  return idx < 18 ? (int)SLOT_CARD_D[idx] : 0;
}
/* Shop::buy src/cards.cpp:102-106 */
__device__ __forceinline__ int shop_buy(Env &E, int idx, int lane) {
  const int sl = idx + 1;
  const u32 inm = (bcast(E.shop, sl) >> 8) & 1u;
  E.n_in_market = (E.n_in_market + 1u - inm) & 0xffu;
  if (lane == sl) E.shop |= 0x100u;
  return shop_get(E, idx, lane);
}
/* Shop::set_available_mask src/cards.cpp:109-121: bits 1..18 of get_from_shop.
 * coin2 = coins * 2; the comparison is strict (coins > cost). */
__device__ __forceinline__ u32 shop_available_bits(const Env &E, int coin2, int lane) {
  const int cost = (int)((E.lc >> 16) & 7u);
  const bool open = E.n_in_market < 6 ? (E.shop & 0xffu) > 0 : ((E.shop >> 8) & 1u) != 0;
  return __ballot_sync(COG_FULL, lane >= 1 && lane <= 18 && open && coin2 > 2 * cost);
}
/* Shop::set_transmit_mask src/cards.cpp:123-127 */
__device__ __forceinline__ u32 shop_transmit_bits(const Env &E, int lane) {
  return __ballot_sync(COG_FULL, lane >= 1 && lane <= 18 && (E.shop & 0xffu) > 0);
}

/* --------------------------------------------------------------------- map */
/* Map::get_from_array src/map.cpp:273-275 on the one-byte-per-hex grid: the
 * reference truncates (coord - min) toward zero, then adds 1 (SURVEY A#17) */
__device__ __forceinline__ u32 grid_lookup(const u8 *__restrict__ grid, int x, int y, int minx, int miny) {
  const int ix = (x - minx) / 2 + 1;
  const int iy = (y - miny) / 2 + 1;
  if (ix < 0 || iy < 0 || ix >= COGL_GRID || iy >= COGL_GRID) return 0;
  return grid[ix * COGL_GRID + iy];
}
__device__ __forceinline__ void dir_delta(int d, int &dx, int &dy) {
  /* include/geometry.h:42-50: none, E, NE, NW, W, SW, SE (x2) */
  dx = d == 1 ? 2 : (d == 3 || d == 4) ? -2 : d == 6 ? 2 : 0;
  dy = d == 2 ? 2 : d == 3 ? 2 : (d == 5 || d == 6) ? -2 : 0;
}
/* refresh the neighbour cache of player p_abs (relative byte p) from the grid (lanes 0..6) */
__device__ __forceinline__ void nbr_refresh(Env &E, const u8 *__restrict__ grid, int p_abs, int p, int lane) {
  const u32 loc = bcast(E.ploc, p_abs);
  int dx, dy;
  dir_delta(lane, dx, dy);
  u32 code = 0;
  if (lane < 7) code = grid_lookup(grid, loc_x(loc) + dx, loc_y(loc) + dy, E.minx, E.miny);
  E.nbr = setb(E.nbr, p, code);
}
/* Map::set_movement_mask src/map.cpp:369-387: bits 1..6 of move for the player in
 * relative byte `player`, with explicit resources (x2) and active-card count */
__device__ __forceinline__ u32 movement_bits(const Env &E, int player, int r0, int r1, int r2,
                                             int n_active, int lane) {
  const u32 c = getb(E.nbr, player);
  const int req = hex_req(c), n = hex_n(c);
  bool filled;
  if (req >= REQ_DISCARD) filled = n_active > n;
  else filled = (req == 0 ? r0 : (req == 1 ? r1 : r2)) >= 2 * n;
  return __ballot_sync(COG_FULL, lane >= 1 && lane <= 6 && req != REQ_NULL && filled);
}

/* ----------------------------------------------------------------- cog_env */
/* cog_env::update_observation src/environment.cpp:252-279 */
/* `pc` = the COG_PF_COUNTS word of player `agent` (relative byte rp) */
/* returns P[agent].move | P[agent].get_from_shop << 7 (also stored in lane `agent`) */
__device__ __forceinline__ u32 env_update_observation(Env &E, int agent, int rp, u32 pc, int lane) {
  u32 move = 1u, shop = 1u;
  if (E.phase == PH_MOVEMENT) {
    const int n_active = (int)getb(pc, 1);
    move |= movement_bits(E, rp, E.res[0], E.res[1], E.res[2], n_active, lane);
  } else if (E.phase == PH_BUYING) {
    shop |= shop_available_bits(E, E.res[2], lane);
  }
  const u32 pms = move | (shop << 7);
  if (lane == agent) E.Pms = pms;
  return pms;
}
__device__ __forceinline__ void env_cycle_phase(Env &E) { /* include/player.h:78-84 */
  u32 n = E.phase + 1;
  if (n >= PH_MAX) n -= PH_MAX;
  E.phase = n;
}

/* --------------------------------------------------------- board generation */
/* Per-warp scratch in shared memory.  The occupancy bitmap is a 128 x 128 torus of
 * half-hex cells: a board that fits the 48 x 48 observation spans <= 94 half units,
 * and a candidate footprint reaches < 17 further, so two distinct live cells can
 * never alias (a generation whose intermediate board outgrows the observation, or piles up more than
 * BG_MAX_PLACED pieces, is repeated out of line by the piece list: bg_generate_wide). */
#ifndef COG_BG_MAX_PLACED
#define COG_BG_MAX_PLACED 32 /* one lane per piece of the map (tests build with less, to exercise bg_generate_wide) */
#endif
enum { BG_OCC_WORDS = 128 * 128 / 32, BG_MAX_PLACED = COG_BG_MAX_PLACED };
#ifdef COG_WIDE_BOARDS
enum { BG_WIDE = -100 }; /* internal: this generation needs the out-of-line path by the piece list (bg_generate_wide) */
#else
enum { BG_WIDE = -3 };   /* default build: reported as COG_ERR_MAP_BOUNDS (see env_reset) */
#endif
struct bg_pose { int16_t cx, cy, rot, pad; };
struct bg_placed { int16_t cx, cy, rot; u8 piece, pad; };
struct BoardScratch {
  u32 occ[BG_OCC_WORDS];
  bg_pose pose[COG_N_PIECES_TOTAL];  /* current pose of each piece (global objects in the reference) */
  bg_placed placed[BG_MAX_PLACED];   /* Map::pieces, with the pose each entry was placed at */
};
struct BoardState {
  int n_list;
  int minx, miny, maxx, maxy; /* Map::min_xy / max_xy */
};

__device__ __forceinline__ const xy8 *piece_coords(int size) {
  return size == SIZE_LARGE ? COORDS_LARGE : (size == SIZE_SMALL ? COORDS_SMALL : COORDS_END);
}
__device__ __forceinline__ u32 occ_index(int x, int y) { return ((u32)(x & 127) << 7) | (u32)(y & 127); }

/* Map::reset src/map.cpp:744-752 */
__device__ __forceinline__ void bg_map_reset(BoardScratch &sc, BoardState &st, int lane) {
  uint4 *o4 = reinterpret_cast<uint4 *>(sc.occ);
  for (int w = lane; w < BG_OCC_WORDS / 4; w += 32) o4[w] = make_uint4(0u, 0u, 0u, 0u);
  st.n_list = 0;
  st.minx = st.miny = st.maxx = st.maxy = 0;
  __syncwarp();
}
__device__ __forceinline__ int mod6(int v) { /* v in [0, 12) */
  return v >= 6 ? v - 6 : v;
}
__device__ __forceinline__ int sx8(u32 v) { return (int)(int8_t)(v & 0xffu); } /* sign-extended byte */

/* Map::add_piece src/map.cpp:309-341 with MapPiece::rotate/translate :179-191.  `rotation` is already
 * reduced to [0, 6) (only its value mod 6 is ever observable).  Lanes 0..6 set the piece's rows in the
 * occupancy bitmap; the bounds come from the shape's bounding box.  Returns 0 or COG error (-3). */
__device__ __forceinline__ int bg_add_piece(BoardScratch &sc, BoardState &st, int piece, int px, int py,
                                            int rotation, int lane) {
  if (st.n_list >= BG_MAX_PLACED) return BG_WIDE; /* one lane per piece of the map: see bg_generate_wide */
  /* (shared-memory reads that steer control flow go through a broadcast: provably uniform) */
  const int rot = mod6((int)bcast((u32)sc.pose[piece].rot, 0) + rotation),
            cx = (int)bcast((u32)(sc.pose[piece].cx + px), 0), cy = (int)bcast((u32)(sc.pose[piece].cy + py), 0);
  __syncwarp();
  if (lane == 0) {
    sc.pose[piece].rot = (int16_t)rot;
    sc.pose[piece].cx = (int16_t)cx;
    sc.pose[piece].cy = (int16_t)cy;
    bg_placed pl;
    pl.cx = (int16_t)cx; pl.cy = (int16_t)cy; pl.rot = (int16_t)rot; pl.piece = (u8)piece; pl.pad = 0;
    sc.placed[st.n_list] = pl;
  }
  st.n_list++;
  const int shape = (int)PIECE_META[piece].size;
  if (lane < 7) {
    const u32 e = FOOT.row[shape][rot][lane];
    const u32 bits = e >> 16;
    const u32 X = (u32)(cx + sx8(e)) & 127u, Y = (u32)(cy + sx8(e >> 8)) & 127u;
    const u32 w0 = Y >> 5, sh = Y & 31u;
    /* rows of one piece are distinct x: no two lanes touch the same word */
    if (bits) {
      sc.occ[X * 4 + w0] |= bits << sh;
      if (sh > 16) sc.occ[X * 4 + ((w0 + 1) & 3u)] |= bits >> (32u - sh);
    }
  }
  const u32 bb = FOOT.bbox[shape][rot];
  st.minx = min(st.minx, cx + sx8(bb));
  st.maxx = max(st.maxx, cx + sx8(bb >> 8));
  st.miny = min(st.miny, cy + sx8(bb >> 16));
  st.maxy = max(st.maxy, cy + sx8(bb >> 24));
  __syncwarp();
  /* A board whose largest cell index (coord - min + 2) / 2 leaves the 48 x 48 observation cannot be the final
   * one (bg_finalize reports that), but it can be an intermediate one: when the end piece then finds no place,
   * the reference resets the map and generates again (src/map.cpp:734-739), and that board may fit.  The
   * bitmap only holds boards that fit: the generation is repeated by the piece list (bg_generate_wide). */
  if ((st.maxx - st.minx + 2) / 2 > COGL_GRID - 1 || (st.maxy - st.miny + 2) / 2 > COGL_GRID - 1) return BG_WIDE;
  return 0;
}

/* rotation option `k` of a connection kind (:203-263): {-2..3}, {-1,2}, {-3}; as a value in [0, 6) */
__device__ __forceinline__ int bg_rot_option(int kind, int k) {
  return kind == 1 ? (k == 0 ? 5 : 2) : (kind == 2 ? 3 : mod6(k + 4));
}
/* entry e of host->get_centered_connections(new) (src/map.cpp:192-201): base point e % n_alt, rotated
 * e / n_alt times and then by the host's own rotation — one table lookup, since rotations add up.
 * rot_add = (e / n_alt + host rotation) mod 6 */
__device__ __forceinline__ void bg_conn_entry(int hcx, int hcy, int hrot, int kind, int n_alt, int e,
                                              int &x, int &y, int &rot_add) {
  /* e / n_alt for n_alt in {1, 2, 3, 6} and e < 21 as a multiply-shift */
  const u32 recip = n_alt == 1 ? 256u : (n_alt == 2 ? 128u : (n_alt == 3 ? 86u : 43u));
  const int i = (int)(((u32)e * recip) >> 8), j = e - i * n_alt;
  rot_add = mod6(i + hrot);
  const u32 c = FOOT.conn[kind][j][rot_add];
  x = sx8(c) + hcx;
  y = sx8(c >> 8) + hcy;
}

/* size / type of a piece from its index (the table order: 0-1 start, 2-13 large travel, 14-17 small travel,
 * 18-19 end) — arithmetic, so that each LANE can ask about a different piece */
__device__ __forceinline__ int piece_size(int p) { return p < 14 ? (int)SIZE_LARGE : (p < 18 ? (int)SIZE_SMALL : (int)SIZE_TRIPLE); }
namespace tab {
constexpr piece_meta C_META[COG_N_PIECES_TOTAL] = COG_PIECE_META_INIT;
constexpr bool piece_order_ok() {
  for (int p = 0; p < COG_N_PIECES_TOTAL; ++p) {
    const int size = p < 14 ? 0 : (p < 18 ? 1 : 2), type = p < 2 ? 0 : (p < 18 ? 1 : 2);
    if (C_META[p].size != size || C_META[p].type != type) return false;
  }
  return true;
}
}  // namespace tab
static_assert(tab::piece_order_ok(), "piece_size() / the start-piece test follow the order of the piece table");

/* Map::add_random_piece src/map.cpp:277-307 (+ overlap :53-74).  The candidates — for every piece of the
 * map in order, the entries of its connection table towards the new piece — are numbered consecutively and
 * tested 32 at a time, one per lane: lane c finds its host piece and entry, looks the entry's connection
 * point up (rotations add, so one table lookup) and tests the new piece's footprint at the entry's FIRST
 * rotation option against the occupancy bitmap, row by row.  A board of three travel pieces needs 5 rounds
 * (14, 28, 42 and 21 candidates) where one round per host piece took 10.
 * Returns 1 placed, 0 no candidate fits, < 0 error. */
__device__ __forceinline__ int bg_add_random_piece(BoardScratch &sc, BoardState &st, int piece, u32 &rng,
                                                   int lane) {
  __syncwarp();
  if (lane == 0) { /* MapPiece::reset :174-177 acts on the shared piece object */
    sc.pose[piece].cx = 0; sc.pose[piece].cy = 0; sc.pose[piece].rot = 0;
  }
  __syncwarp();
  const int ns = piece_size(piece);
  /* lane k <-> piece k of the map: which connection table it offers the new piece (src/map.cpp:203-263:
   * 0 LARGE->LARGE 14 entries, 1 LARGE->SMALL 21, 2 LARGE(not a start piece)->TRIPLE 7, 3 SMALL->LARGE 6) */
  const int my_host = lane < st.n_list ? (int)sc.placed[lane].piece : 0;
  const int hs = piece_size(my_host);
  int my_kind = -1;
  if (ns == SIZE_LARGE) my_kind = hs == SIZE_LARGE ? 0 : (hs == SIZE_SMALL ? 3 : -1);
  else if (ns == SIZE_SMALL) my_kind = hs == SIZE_LARGE ? 1 : -1;
  else my_kind = (hs == SIZE_LARGE && my_host >= 2) ? 2 : -1;
  if (lane >= st.n_list) my_kind = -1;
  const u32 my_n = my_kind < 0 ? 0u : (0x31eaeu >> (5 * my_kind)) & 31u; /* 14, 21, 7, 6 */
  const u32 my_incl = warp_incl_scan(my_n, lane);
  const u32 total = bcast(my_incl, 31);
  const int n_rows = ns == SIZE_TRIPLE ? 3 : 7; /* rows of the new piece's footprint */
  u32 my_valid = 0; /* lane r keeps the valid-candidate bit set of round r (<= 32 x 21 / 32 = 21 rounds) */
  int n_valid = 0, round = 0;
  for (u32 base = 0; base < total; base += 32, ++round) {
    const u32 c = base + (u32)lane;
    int k = 0; /* the piece of the map candidate c belongs to */
    for (int j = 0; j < st.n_list; ++j) k += c >= bcast(my_incl, j);
    k = k < 31 ? k : 31;
    const int kind = max(__shfl_sync(COG_FULL, my_kind, k), 0);
    const int e = (int)(c - (__shfl_sync(COG_FULL, my_incl, k) - __shfl_sync(COG_FULL, my_n, k)));
    const bg_pose hp = sc.pose[__shfl_sync(COG_FULL, my_host, k)];
    const int n_alt = (0x6132 >> (4 * kind)) & 15, first_opt = (0x4354 >> (4 * kind)) & 15; /* 2 3 1 6 / -2 -1 -3 -2 mod 6 */
    int x, y, rot_add;
    bg_conn_entry(hp.cx, hp.cy, hp.rot, kind, n_alt, c < total ? e : 0, x, y, rot_add);
    const u32 *rows = FOOT.row[ns][mod6(first_opt + rot_add)]; /* only the FIRST rotation option is tested */
    u32 hit = 0;
    for (int j = 0; j < n_rows; ++j) {
      const u32 r = rows[j];
      const u32 X = (u32)(x + sx8(r)) & 127u, Y = (u32)(y + sx8(r >> 8)) & 127u;
      const u32 w0 = Y >> 5;
      const u32 window = __funnelshift_r(sc.occ[X * 4 + w0], sc.occ[X * 4 + ((w0 + 1) & 3u)], Y & 31u);
      hit |= window & (r >> 16);
    }
    const u32 vm = __ballot_sync(COG_FULL, c < total && !hit);
    if (lane == round) my_valid = vm;
    n_valid += __popc(vm);
  }
  if (!n_valid) return 0;
  const u32 idx = uniform_draw(rng, (u32)n_valid);
  const u32 cnt = (u32)__popc(my_valid);
  const u32 incl = warp_incl_scan(cnt, lane);
  const u32 hitr = __ballot_sync(COG_FULL, incl > idx);
  const int r = __ffs((int)hitr) - 1;
  const u32 vm = bcast(my_valid, r);
  const u32 within = idx - (bcast(incl, r) - bcast(cnt, r));
  const u32 c = 32u * (u32)r + nth_set_bit_warp(vm, within, lane); /* the chosen candidate */
  int k = 0;
  for (int j = 0; j < st.n_list; ++j) k += c >= bcast(my_incl, j);
  const int kind = (int)bcast((u32)my_kind, k);
  const int e = (int)(c - (bcast(my_incl, k) - bcast(my_n, k)));
  const int host = (int)bcast((u32)my_host, k);
  const bg_pose hp = sc.pose[host];
  const int n_alt = (0x6132 >> (4 * kind)) & 15, n_opts = (0x6126 >> (4 * kind)) & 15; /* rotation options: 6 2 1 6 */
  int x, y, rot_add;
  bg_conn_entry((int)bcast((u32)hp.cx, 0), (int)bcast((u32)hp.cy, 0), (int)bcast((u32)hp.rot, 0), kind, n_alt, e, x, y,
                rot_add);
  const int rot_idx = (int)uniform_draw(rng, (u32)n_opts);
  const int rc = bg_add_piece(sc, st, piece, x, y, mod6(bg_rot_option(kind, rot_idx) + rot_add), lane);
  return rc ? rc : 1;
}

/* ------------------------------------------------ board generation by the piece list (rare, out of line)
 * The fast path above keeps the map as a bitmap that only holds boards that fit the observation and enumerates
 * candidates one lane per placed piece, so it stops (BG_WIDE) when an intermediate board grows wider than 48
 * cells or more than 32 pieces pile up (a travel piece that finds no place makes the reference recurse WITHOUT
 * Map::reset, src/map.cpp:726-731).  The reference goes on in both cases and may still end with a board that
 * fits — met once in about half a million boards of 6-8 pieces, never at <= 5.  Those generations are repeated from the same
 * engine state by the functions below: same candidate order, same draws, overlap tested piece against piece
 * (exact for any extent), up to 512 pieces (the C restatement's capacity), the list spread over the lanes'
 * local memory.  Slow and never on the path of the benchmarks. */
enum { BG_WIDE_SLOTS = 16 };
struct WideList { /* piece k of the map: lane k & 31, slot k >> 5 */
  u32 a[BG_WIDE_SLOTS]; /* cx | cy << 16 (int16 each), the pose it was placed at */
  u32 b[BG_WIDE_SLOTS]; /* rot | piece << 8 */
};
__device__ __forceinline__ void wide_get(const WideList &wl, int k, int &cx, int &cy, int &rot, int &piece) {
  const u32 a = __shfl_sync(COG_FULL, wl.a[k >> 5], k & 31), b = __shfl_sync(COG_FULL, wl.b[k >> 5], k & 31);
  cx = (int)(int16_t)(a & 0xffffu);
  cy = (int)(int16_t)(a >> 16);
  rot = (int)(b & 0xffu);
  piece = (int)(b >> 8);
}

/* Map::add_piece src/map.cpp:309-341, list only */
__device__ __forceinline__ int bg_add_piece_wide(BoardScratch &sc, BoardState &st, WideList &wl, int piece, int px,
                                                 int py, int rotation, int lane) {
  if (st.n_list >= 32 * BG_WIDE_SLOTS) return -3;
  const int rot = mod6((int)bcast((u32)sc.pose[piece].rot, 0) + rotation),
            cx = (int)bcast((u32)(sc.pose[piece].cx + px), 0), cy = (int)bcast((u32)(sc.pose[piece].cy + py), 0);
  __syncwarp();
  if (lane == 0) {
    sc.pose[piece].rot = (int16_t)rot;
    sc.pose[piece].cx = (int16_t)cx;
    sc.pose[piece].cy = (int16_t)cy;
  }
  if (lane == (st.n_list & 31)) {
    wl.a[st.n_list >> 5] = ((u32)cx & 0xffffu) | ((u32)cy << 16);
    wl.b[st.n_list >> 5] = (u32)rot | ((u32)piece << 8);
  }
  st.n_list++;
  const u32 bb = FOOT.bbox[PIECE_META[piece].size][rot];
  st.minx = min(st.minx, cx + sx8(bb));
  st.maxx = max(st.maxx, cx + sx8(bb >> 8));
  st.miny = min(st.miny, cy + sx8(bb >> 16));
  st.maxy = max(st.maxy, cy + sx8(bb >> 24));
  __syncwarp();
  return 0;
}

/* overlap (src/map.cpp:53-74) of this lane's footprint (`rows` at x, y) with every piece of the map.  Same row
 * encoding as the bitmap test: byte 0 dx, byte 1 dy of the row's first cell, bits 16.. one bit per cell along y. */
__device__ __forceinline__ u32 bg_overlap_wide(const WideList &wl, int n_list, const u32 *rows, int n_rows, int x,
                                               int y) {
  u32 hit = 0;
  for (int q = 0; q < n_list; ++q) {
    int qcx, qcy, qrot, qpiece;
    wide_get(wl, q, qcx, qcy, qrot, qpiece);
    const u32 *qrows = FOOT.row[piece_size(qpiece)][qrot];
    for (int j = 0; j < n_rows; ++j) {
      const u32 r = rows[j];
      const int cx = x + sx8(r), cy = y + sx8(r >> 8);
      const u64 cb = (u64)(r >> 16) << 16; /* room to shift by up to 15 either way */
      for (int k = 0; k < 7; ++k) {
        const u32 r2 = qrows[k];
        const int sh = cy - (qcy + sx8(r2 >> 8));
        if (qcx + sx8(r2) == cx && sh > -16 && sh < 16)
          hit |= (u32)(((sh >= 0 ? cb << sh : cb >> -sh) & ((u64)(r2 >> 16) << 16)) != 0);
      }
    }
  }
  return hit;
}

/* Map::add_random_piece src/map.cpp:277-307: the candidates in the order of bg_add_random_piece (pieces of the
 * map in order, the entries of each one's connection table), 32 hosts at a time; one pass to count the valid
 * ones, the draw, a second pass to find the one drawn.  Returns 1 placed, 0 no candidate fits, < 0 error. */
__device__ __forceinline__ int bg_add_random_piece_wide(BoardScratch &sc, BoardState &st, WideList &wl, int piece,
                                                        u32 &rng, int lane) {
  __syncwarp();
  if (lane == 0) { /* MapPiece::reset :174-177 acts on the shared piece object */
    sc.pose[piece].cx = 0; sc.pose[piece].cy = 0; sc.pose[piece].rot = 0;
  }
  __syncwarp();
  const int ns = piece_size(piece);
  const int n_rows = ns == SIZE_TRIPLE ? 3 : 7;
  u32 n_valid = 0, idx = 0;
  for (int pass = 0; pass < 2; ++pass) {
    u32 seen = 0;
    for (int hb = 0; hb < st.n_list; hb += 32) {
      /* lane l <-> piece hb + l of the map, which is this lane's own slot hb >> 5 */
      const bool have = hb + lane < st.n_list;
      const int my_host = have ? (int)(wl.b[hb >> 5] >> 8) : 0;
      const int hs = piece_size(my_host);
      int my_kind = -1;
      if (ns == SIZE_LARGE) my_kind = hs == SIZE_LARGE ? 0 : (hs == SIZE_SMALL ? 3 : -1);
      else if (ns == SIZE_SMALL) my_kind = hs == SIZE_LARGE ? 1 : -1;
      else my_kind = (hs == SIZE_LARGE && my_host >= 2) ? 2 : -1;
      if (!have) my_kind = -1;
      const u32 my_n = my_kind < 0 ? 0u : (0x31eaeu >> (5 * my_kind)) & 31u; /* 14, 21, 7, 6 */
      const u32 my_incl = warp_incl_scan(my_n, lane);
      const u32 total = bcast(my_incl, 31);
      const int n_here = st.n_list - hb < 32 ? st.n_list - hb : 32;
      for (u32 base = 0; base < total; base += 32) {
        const u32 c = base + (u32)lane;
        int k = 0; /* the piece of this chunk candidate c belongs to */
        for (int j = 0; j < n_here; ++j) k += c >= bcast(my_incl, j);
        k = k < 31 ? k : 31;
        const int kind = max(__shfl_sync(COG_FULL, my_kind, k), 0);
        const int e_raw = (int)(c - (__shfl_sync(COG_FULL, my_incl, k) - __shfl_sync(COG_FULL, my_n, k)));
        const int e = c < total ? e_raw : 0; /* lanes past the last candidate compute on entry 0 and are masked below */
        const int host = __shfl_sync(COG_FULL, my_host, k);
        const bg_pose hp = sc.pose[host];
        const int n_alt = (0x6132 >> (4 * kind)) & 15, first_opt = (0x4354 >> (4 * kind)) & 15;
        int x, y, rot_add;
        bg_conn_entry(hp.cx, hp.cy, hp.rot, kind, n_alt, e, x, y, rot_add);
        const u32 hit = bg_overlap_wide(wl, st.n_list, FOOT.row[ns][mod6(first_opt + rot_add)], n_rows, x, y);
        const u32 vm = __ballot_sync(COG_FULL, c < total && !hit);
        const u32 cnt = (u32)__popc(vm);
        if (pass == 0) {
          n_valid += cnt;
        } else if (idx - seen < cnt) { /* (unsigned: idx >= seen always holds here) the candidate drawn is in this round */
          const int src = (int)nth_set_bit_warp(vm, idx - seen, lane);
          const int ckind = __shfl_sync(COG_FULL, kind, src), ce = __shfl_sync(COG_FULL, e, src);
          const bg_pose chp = sc.pose[__shfl_sync(COG_FULL, host, src)];
          const int c_alt = (0x6132 >> (4 * ckind)) & 15, n_opts = (0x6126 >> (4 * ckind)) & 15; /* rotation options: 6 2 1 6 */
          int cx, cy, crot_add;
          bg_conn_entry((int)bcast((u32)chp.cx, 0), (int)bcast((u32)chp.cy, 0), (int)bcast((u32)chp.rot, 0), ckind, c_alt,
                        ce, cx, cy, crot_add);
          const int rot_idx = (int)uniform_draw(rng, (u32)n_opts);
          const int rc = bg_add_piece_wide(sc, st, wl, piece, cx, cy, mod6(bg_rot_option(ckind, rot_idx) + crot_add), lane);
          return rc ? rc : 1;
        }
        seen += cnt;
      }
    }
    if (pass == 0) {
      if (!n_valid) return 0;
      idx = uniform_draw(rng, n_valid);
    }
  }
  return -3; /* not reached: the second pass always finds candidate idx < n_valid */
}

/* Map::generate src/map.cpp:697-742.  The reference recurses (rng BY VALUE, no map
 * reset in the travel-piece branch); here the <= 5 frames live on an explicit stack.
 * Returns 0, -2 (generate_map_failure) or -3 (board outgrew the observation). */
template <bool WIDE = false>
__device__ __forceinline__ int bg_generate(BoardScratch &sc, BoardState &st, u32 rng0, int n_pieces,
                                           int difficulty, int lane, WideList *wl = nullptr) {
  enum { F_ENTER = 0, F_LOOP = 1, F_DONE = 2, MAX_DEPTH = 5 };
  /* the running frame lives in registers; parents wait on a small stack.  Popped values go
   * through a broadcast so that ptxas can still prove them warp-uniform (a shuffle from a fixed
   * lane is uniform by construction; a value that came back through local memory is not) */
  u32 s_rng[MAX_DEPTH];
  u64 s_valid[MAX_DEPTH];
  int s_nvalid[MAX_DEPTH], s_i[MAX_DEPTH], s_state[MAX_DEPTH];
  u32 rng = rng0;
  u64 valid = 0; /* valid_indices as a list of nibbles */
  int nvalid = 0, fi = 0, state = F_ENTER, depth = 0;
  if (lane < 2) { /* the start pieces are never moved */
    sc.pose[lane].cx = 0; sc.pose[lane].cy = 0; sc.pose[lane].rot = 0;
  }
  __syncwarp();
  for (;;) {
    if (state == F_ENTER) {
      const int start = (int)uniform_draw(rng, 2);
      int rc;
      if constexpr (WIDE) rc = bg_add_piece_wide(sc, st, *wl, start, 0, 0, 0, lane);
      else rc = bg_add_piece(sc, st, start, 0, 0, 0, lane);
      if (rc) return rc;
      valid = 0;
      nvalid = 0;
      for (int i = 0; i < 16; ++i)
        if ((int)PIECE_META[2 + i].difficulty <= difficulty) valid |= (u64)i << (4 * nvalid++);
      fi = 0;
      state = F_LOOP;
    } else if (state == F_LOOP) {
      bool descend = false;
      if (fi < n_pieces) {
        int success = 0, next_idx = 0;
        if (nvalid) {
          const u32 pos = uniform_draw(rng, (u32)nvalid);
          next_idx = (int)((valid >> (4 * pos)) & 15u);
          if constexpr (WIDE) success = bg_add_random_piece_wide(sc, st, *wl, 2 + next_idx, rng, lane);
          else success = bg_add_random_piece(sc, st, 2 + next_idx, rng, lane);
          if (success < 0) return success;
        }
        fi++;
        if (success) {
          /* valid_indices.erase(begin + next_idx): the piece index is used as a POSITION;
           * past-the-end degenerates to pop_back (libstdc++ bits/vector.tcc:182-190) */
          if (next_idx < nvalid - 1) {
            const u64 low = valid & ((1ull << (4 * next_idx)) - 1ull);
            valid = low | ((valid >> (4 * (next_idx + 1))) << (4 * next_idx));
          }
          nvalid--;
        } else {
          descend = true; /* generate(..., failures + 1, rng) WITHOUT reset, then carry on */
        }
      } else {
        const int endp = 18 + (int)uniform_draw(rng, 2);
        int ok;
        if constexpr (WIDE) ok = bg_add_random_piece_wide(sc, st, *wl, endp, rng, lane);
        else ok = bg_add_random_piece(sc, st, endp, rng, lane);
        if (ok < 0) return ok;
        state = F_DONE;
        if (!ok) {
          bg_map_reset(sc, st, lane);
          descend = true;
        }
      }
      if (descend) {
        if (depth + 1 >= MAX_DEPTH) return -2; /* failures >= MAX_FAILURES throws */
        s_rng[depth] = rng; /* the callee starts from a COPY of the caller's engine */
        s_valid[depth] = valid;
        s_nvalid[depth] = nvalid;
        s_i[depth] = fi;
        s_state[depth] = state;
        state = F_ENTER;
        depth++;
      }
    } else {
      if (depth == 0) return 0;
      depth--;
      rng = bcast(s_rng[depth], 0);
      valid = (u64)bcast((u32)s_valid[depth], 0) | ((u64)bcast((u32)(s_valid[depth] >> 32), 0) << 32);
      nvalid = (int)bcast((u32)s_nvalid[depth], 0);
      fi = (int)bcast((u32)s_i[depth], 0);
      state = (int)bcast((u32)s_state[depth], 0);
    }
  }
}

/* Map::finalize src/map.cpp:389-405 (cell index from the last add_piece :335-340): clear what the previous
 * board left in the private 48 x 48 grid and the 48 x 48 x 7 observation (16-byte stores), then scatter, piece by piece in
 * placement order (later pieces win index collisions like the reference's overwrite), one code per
 * hex into the grid and the hex's seven feature bytes into the observation: feature 1 + requirement =
 * n_required, feature 6 = is_end, feature 0 never written.  (The zero fill leaves the lines in L2, the
 * byte stores merge there; the encode used to be a pass over all 1008 16-byte chunks of the map that
 * re-derived every chunk from the grid — 3 100 warp instructions per reset against about 500.) */
__device__ __forceinline__ void bg_clear_previous(u8 *grid, u8 *mapobs, int lane) {
  uint4 *g4 = reinterpret_cast<uint4 *>(grid);
  uint4 *m4 = reinterpret_cast<uint4 *>(mapobs);
  /* The observation is the expansion of the private grid, cell by cell, at all times (both start zeroed —
   * cog_vec_create — and are written together, below and nowhere else): 16 grid cells <-> 7 x 16 bytes of
   * the observation.  So only the chunks the PREVIOUS board touched need clearing — about 35 of the 144 —
   * instead of all 18 KB: 1.30 -> 1.40e8 resets/s, reset-heavy configs[3] 2.47 -> 2.59e9 env-steps/s
   * (profiles/r02_w_clear_dirty.txt). */
  static_assert(COGL_MAP_BYTES == 7 * COGL_GRID_BYTES && COGL_GRID_BYTES % 16 == 0, "chunk correspondence");
  for (int w = lane; w < COGL_GRID_BYTES / 16; w += 32) {
    const uint4 g = g4[w];
    if (g.x | g.y | g.z | g.w) {
      g4[w] = make_uint4(0, 0, 0, 0);
#pragma unroll
      for (int b = 0; b < 7; ++b) m4[7 * w + b] = make_uint4(0, 0, 0, 0);
    }
  }
  __syncwarp();
}
/* one piece of the map into the grid and the observation; returns true when a hex falls outside */
__device__ __forceinline__ bool bg_scatter_piece(int piece, int rot, int ox, int oy, u8 *grid, u8 *mapobs, int lane) {
  bool bad = false;
  {
    const piece_meta pm = PIECE_META[piece];
    const xy8 *co = piece_coords(pm.size);
    const rot6 m = rot_matrix(rot);
    for (int h = lane; h < pm.n_hexes; h += 32) {
      int x = co[h].x, y = co[h].y;
      rot_apply(m, x, y);
      const int ix = (x + ox) / 2, iy = (y + oy) / 2;
      if (ix >= COGL_GRID || iy >= COGL_GRID) {
        bad = true;
      } else {
        const u32 code = PIECE_HEX[piece][h];
        const int cell = ix * COGL_GRID + iy;
        grid[cell] = (u8)code;
        const u32 f = code & 7u; /* requirement + 1, 0 = none */
        u64 v = (u64)((code >> 6) & 1u) << 48;
        if (f) v |= (u64)((code >> 3) & 7u) << (8 * f);
        u8 *rec = mapobs + 7 * cell;
#pragma unroll
        for (int b = 0; b < 7; ++b) rec[b] = (u8)(v >> (8 * b));
      }
    }
    __syncwarp();
  }
  return bad;
}
__device__ __forceinline__ int bg_finalize(const BoardScratch &sc, const BoardState &st, u8 *grid,
                                           u8 *mapobs, int lane) {
  bg_clear_previous(grid, mapobs, lane);
  bool bad = false;
  for (int k = 0; k < st.n_list; ++k) {
    const bg_placed pl = sc.placed[k];
    const int piece = (int)bcast((u32)pl.piece, 0), rot = (int)bcast((u32)pl.rot, 0);
    const int ox = (int)bcast((u32)(pl.cx - st.minx + 2), 0), oy = (int)bcast((u32)(pl.cy - st.miny + 2), 0);
    bad |= bg_scatter_piece(piece, rot, ox, oy, grid, mapobs, lane);
  }
  if (__any_sync(COG_FULL, bad)) return -3;
  return 0;
}

/* The whole generation once more from the same engine state, by the piece list, and the map encode from that
 * list.  Out of line: a path this cold must not shape the register allocation of the kernels that reset. */
/* (Everything goes in and out BY VALUE except the shared-memory scratch: a caller's BoardState handed over by
 * reference would have to live in local memory for the whole kernel.)  Returns the error code in the low word
 * and, on success, the board's min_xy (loc_pack) in the high word. */
__device__ __attribute__((noinline)) u64 bg_generate_wide(BoardScratch &sc, u32 rng0, int n_pieces, int difficulty,
                                                          u8 *grid, u8 *mapobs, int lane) {
  WideList wl;
  for (int k = 0; k < BG_WIDE_SLOTS; ++k) wl.a[k] = wl.b[k] = 0;
  BoardState st;
  bg_map_reset(sc, st, lane);
  const int rc = bg_generate<true>(sc, st, rng0, n_pieces, difficulty, lane, &wl);
  if (rc) return (u64)(u32)rc;
  /* a board that leaves the observation is reported before anything is written */
  if ((st.maxx - st.minx + 2) / 2 > COGL_GRID - 1 || (st.maxy - st.miny + 2) / 2 > COGL_GRID - 1) return (u64)(u32)-3;
  bg_clear_previous(grid, mapobs, lane);
  bool bad = false;
  for (int k = 0; k < st.n_list; ++k) {
    int cx, cy, rot, piece;
    wide_get(wl, k, cx, cy, rot, piece);
    bad |= bg_scatter_piece(piece, rot, cx - st.minx + 2, cy - st.miny + 2, grid, mapobs, lane);
  }
  if (__any_sync(COG_FULL, bad)) return (u64)(u32)-3;
  return (u64)loc_pack(st.minx, st.miny) << 32;
}

/* ------------------------------------------------------------------- reset */
/* back to absolute player order with agent 0: the current agent's flag / count words go
 * back to its lane, the rotated words are re-based, cpf / cpc follow player 0 */
__device__ __forceinline__ void env_to_absolute(Env &E, int lane) {
  if (lane == (int)E.agent) { E.pf = E.cpf; E.pc = E.cpc; }
  env_rotate(E, (4u - E.agent) & 3u);
  E.agent = 0;
  E.cpf = bcast(E.pf, 0);
  E.cpc = bcast(E.pc, 0);
}

/* the inverse: from absolute order (agent 0) back to the rotated order of agent `a` */
__device__ __forceinline__ void env_from_absolute(Env &E, u32 a, int lane) {
  env_rotate(E, a);
  E.agent = a;
  E.cpf = bcast(E.pf, (int)a);
  E.cpc = bcast(E.pc, (int)a);
}

__device__ __forceinline__ void report_error(const Buffers &B, int code, u64 i, int lane) {
  if (lane == 0 && atomicCAS(&B.err[0], 0, code) == 0) B.err[1] = (int)i;
}

/* Player::reset src/player.cpp:29-43 -> Deck::reset src/cards.cpp:163-171.
 * ActionMask::reset (include/api.h:104-118) leaves move[1..]/get_from_shop[1..] alone;
 * DeckObs::reset (:75-81) leaves `played` alone; idx_last_activated survives. */
__device__ __forceinline__ void player_reset(Env &E, int p, int lane) {
  /* Deck::reset puts the eight starting cards (3 EXPLORER = card 0, 1 SAILOR = card 5, 4 TRAVELER = card 7) on
   * the discard pile; the first Deck::draw moves them to the draw pile and draws four, each a uniform target
   * below the pile size resolved by the cumulative counts in card order (src/cards.cpp:183-211).  With three
   * card types that scan is scalar arithmetic — same draws, same result as deck_draw, a third of the
   * instructions (the reset kernel spent a fifth of its time here). */
  u32 left0 = 3, left5 = 1, left7 = 4, got0 = 0, got5 = 0, got7 = 0;
  for (u32 n_in_draw = 8; n_in_draw > 4; --n_in_draw) {
    const u32 target = uniform_draw(E.rng, n_in_draw);
    if (target < left0) { --left0; ++got0; }
    else if (target < left0 + left5) { --left5; ++got5; }
    else { --left7; ++got7; }
  }
  E.draw = setb(E.draw, p, lane == 1 ? left0 : lane == 6 ? left5 : lane == 8 ? left7 : 0u);
  E.hand = setb(E.hand, p, lane == 1 ? got0 : lane == 6 ? got5 : lane == 8 ? got7 : 0u);
  E.active = setb(E.active, p, 0);
  E.discard = setb(E.discard, p, 0);
  E.Splay = 1u | (got0 ? 1u << 1 : 0u) | (got5 ? 1u << 6 : 0u) | (got7 ? 1u << 8 : 0u);
  E.Srem = 1u;
  E.Sms |= 1u | (1u << 7);
  /* n_in_hand = 4, n_active = 0, n_in_draw = 4; idx_last_activated survives */
  const u32 pc = (bcast(E.pc, p) & 0xff000000u) | 4u | (4u << 16);
  if (lane == p) {
    E.pf = 0; E.pmv = 0; E.pc = pc;
    E.Pplay = E.Splay; E.Prem = E.Srem; E.Pms = E.Sms; /* save_actionmask */
  }
}

/* cog_env::reset() src/environment.cpp:42-64.  The board comes from a COPY of the
 * env RNG (:48), so the deck draws that follow replay the same stream prefix.
 * Returns 0 or the error that was reported. */
__device__ __forceinline__ int env_reset(Env &E, const Params &prm, const Buffers &B, u64 i,
                                         BoardScratch &sc, int lane) {
  env_to_absolute(E, lane); /* agent 0: relative byte index == player index below */
  E.phase = PH_INACTIVE;
  BoardState st;
  bg_map_reset(sc, st, lane);
  int rc = bg_generate(sc, st, E.rng, (int)prm.n_pieces, (int)prm.difficulty, lane);
  u8 *grid = B.grid + i * COGL_GRID_BYTES;
#ifdef COG_WIDE_BOARDS
  if (COG_UNLIKELY(rc == BG_WIDE)) {
    const u64 w = bg_generate_wide(sc, E.rng, (int)prm.n_pieces, (int)prm.difficulty, grid, B.obs + i * COGL_OBS_SIZE, lane);
    rc = (int)(u32)w;
    st.minx = loc_x((u32)(w >> 32));
    st.miny = loc_y((u32)(w >> 32));
  } else
#else
  /* Default build: such a generation is reported (COG_ERR_MAP_BOUNDS) instead of repeated by the piece list.
   * -DCOG_WIDE_BOARDS builds the exact path in — bit-exact with the restatement on the emulator, also when
   * EVERY board is forced through it (COG_BG_MAX_PLACED=1) — but the call changes the register allocation of
   * the three kernels that reset (more spill code in their bodies), and no GPU time was left to measure it. */
  static_assert(BG_WIDE == -3, "without the exact path BG_WIDE is the error itself");
#endif
  if (!rc) rc = bg_finalize(sc, st, grid, B.obs + i * COGL_OBS_SIZE, lane);
  if (rc) {
    report_error(B, rc, i, lane);
    return rc;
  }
  __syncwarp();
  E.minx = st.minx;
  E.miny = st.miny;
  const int np = (int)prm.n_players < 4 ? (int)prm.n_players : 4;
  for (int p = 0; p < np; ++p) player_reset(E, p, lane);
  /* Map::add_players src/map.cpp:343-354: player i stands on hex i of the start piece
   * (defined like that for fewer than 4 players too; SURVEY A#13) */
  if (lane < np) E.ploc = loc_pack(COORDS_LARGE[lane].x, COORDS_LARGE[lane].y);
  { /* the neighbour caches of all players at once: lane 8 p + d looks up direction d of player p, lane d
     * (0..6) then collects its four bytes */
    const int p = lane >> 3, d = lane & 7;
    const u32 loc = __shfl_sync(COG_FULL, E.ploc, p);
    int dx, dy;
    dir_delta(d, dx, dy);
    u32 code = 0;
    if (d < 7 && p < np) code = grid_lookup(grid, loc_x(loc) + dx, loc_y(loc) + dy, E.minx, E.miny);
    const u32 c1 = __shfl_sync(COG_FULL, code, d + 8), c2 = __shfl_sync(COG_FULL, code, d + 16),
              c3 = __shfl_sync(COG_FULL, code, d + 24);
    E.nbr = lane < 7 ? (code | (c1 << 8) | (c2 << 16) | (c3 << 24)) : 0u;
  }
  /* Shop::reset src/cards.cpp:94-100 (n_in_market is NOT reset) */
  E.shop = (lane >= 1 && lane <= 18) ? (3u | (((E.lc >> 19) & 1u) << 8)) : 0u;
  E.turn = 0;
  for (int p = 0; p < np; ++p) env_update_observation(E, p, p, bcast(E.pc, p), lane);
  E.Splay = bcast(E.Pplay, 0); /* load_actionmask of agent 0 :62-63 */
  E.Srem = bcast(E.Prem, 0);
  E.Sms = bcast(E.Pms, 0);
  E.cpf = bcast(E.pf, 0);
  E.cpc = bcast(E.pc, 0);
  if (lane == 0) atomicAdd(&B.stats[10 /* COG_STAT_RESETS */], 1ull);
  return 0;
}

/* -------------------------------------------------------------------- step */
/* cog_env::next_agent src/environment.cpp:79-89 */
__device__ __forceinline__ void env_next_agent(Env &E, u32 pf, u32 &pc, int p, u32 n_players, int lane) {
  player_end_turn(E, pc, p, 0, lane);
  if (lane == p) { E.pf = pf; E.pc = pc; } /* the old agent's words return to its lane */
  u32 delta = 1; /* seats the agent advances by, mod 4 */
  E.agent = E.agent + 1;
  if (E.agent >= n_players) {
    E.agent = 0;
    delta = (4u - (u32)p) & 3u;
  }
  env_rotate(E, delta);
  E.cpf = bcast(E.pf, (int)E.agent);
  E.cpc = bcast(E.pc, (int)E.agent);
  E.Splay = bcast(E.Pplay, (int)E.agent); /* load_actionmask */
  E.Srem = bcast(E.Prem, (int)E.agent);
  E.Sms = bcast(E.Pms, (int)E.agent);
  res_clear(E);
  E.turn++;
}

/* Episode end, src/environment.cpp:187-207 + get_reward :281-288: rewards, Info and the
 * episode aggregates.  Leaves the env in absolute player order, ready for env_reset().
 * Kept out of env_step so that the fused hot kernel carries no episode-end code at all. */
__device__ __forceinline__ void env_finish_episode(Env &E, const Params &prm, const Buffers &B, u64 i, int lane) {
  env_to_absolute(E, lane); /* absolute player order for the bookkeeping below */
  const u32 np = prm.n_players < 4 ? prm.n_players : 4;
  const bool won = lane < 4 && (E.pf & PFL_WON);
  const u32 winners = __ballot_sync(COG_FULL, won); /* get_reward :281-288 sums all 4 slots */
  const int n_winners = __popc(winners);
  u8 *info = B.infos + i * COGL_INFO_SIZE;
  if (lane == 0) *reinterpret_cast<u32 *>(info) = E.turn; /* total_length = turn_counter */
  if (lane < (int)np) {
    u8 *ai = info + COGL_INFO_AGENT + COGL_INFO_AGENT_STRIDE * lane;
    const float ret = (float)((int)prm.n_players * (won ? 1 : 0) - n_winners);
    const u32 added = getb(E.pf, 2);
    ai[0] = (u8)getb(E.pf, 3);
    *reinterpret_cast<float *>(ai + 4) = ret;
    *reinterpret_cast<u32 *>(ai + 8) = E.pmv;
    ai[12] = (u8)added;
    ai[13] = (u8)added; /* get_n_removed returns n_added_cards src/player.cpp:224 */
    *reinterpret_cast<u32 *>(ai + 16) = 0;
    *reinterpret_cast<u32 *>(ai + 20) = 0;
    *reinterpret_cast<u32 *>(ai + 24) = 0;
    *reinterpret_cast<u32 *>(ai + 28) = added; /* get_n_discarded also returns it :223 */
    B.rewards[i * 4 + lane] = ret;
  }
  /* info steps_taken bytes now mirror the players' counters */
  u32 steps = E.info_steps;
  for (u32 k = 0; k < np; ++k) steps = setb(steps, (int)k, getb(bcast(E.pf, (int)k), 3));
  E.info_steps = steps;
  /* episode aggregates (reduced across GPUs by the host) */
  const u32 moved = __reduce_add_sync(COG_FULL, lane < (int)np ? E.pmv : 0u);
  const u32 added = __reduce_add_sync(COG_FULL, lane < (int)np ? getb(E.pf, 2) : 0u);
  if (lane == 0) {
    atomicAdd(&B.stats[1], 1ull);
    if (!n_winners) atomicAdd(&B.stats[2], 1ull);
    atomicAdd(&B.stats[3], (unsigned long long)E.turn);
    atomicAdd(&B.stats[8], (unsigned long long)moved);
    atomicAdd(&B.stats[9], (unsigned long long)added);
  }
  if (lane < 4 && won) atomicAdd(&B.stats[4 + lane], 1ull);
}

/* cog_env::step src/environment.cpp:91-224 followed by vec_cog_env::step_single's
 * auto-reset (include/vec_environment.h:53-61).  `act` = the five ActionData bytes
 * packed play | special<<8 | remove<<16 | move<<24, shop separately.
 * Returns: bit0 done, bits 8-11 players whose DeckObs changed, bits 12-15 players whose mask changed.  When done, the
 * CALLER must run env_finish_episode() and env_reset() next (kept out of this function so
 * that each kernel holds ONE inlined copy of them — and the fused hot kernel none).
 *
 * LAZY (the fused kernel): update_observation is a pure function of the state the step
 * leaves behind and its result (P[agent].move / .get_from_shop) is only ever read by a
 * special action of the same step or by the export after the last step — every other value
 * is overwritten by the next step's update or by save_actionmask at the end of the turn.
 * So it is run only when a special action needs it; otherwise `obs_pending` is raised and
 * the caller runs it once after its last step. */
/* NO_END: the caller has established that this step neither ends the turn nor the episode (a card
 * play, or a null action in the movement phase, by a player who has not won, before truncation):
 * env_next_agent and the done check fold away.  `pms_out` (optional) receives the value
 * update_observation computed, when it ran. */
template <bool LAZY, bool PLAY_ONLY, bool NO_END = false>
__device__ __forceinline__ u32 env_step(Env &E, u32 act, u32 a_shop_in, const Params &prm, const Buffers &B,
                                        u64 i, bool &obs_pending, int lane, u32 *pms_out = nullptr) {
  /* PLAY_ONLY: the caller knows that only the `play` head can be non-zero (the sampler's
   * fast path) — every other branch of the step folds away at compile time */
  const u32 a_play = act & 0xffu;
  const u32 a_special = PLAY_ONLY ? 0u : (act >> 8) & 0xffu;
  const u32 a_remove = PLAY_ONLY ? 0u : (act >> 16) & 0xffu;
  const u32 a_move = PLAY_ONLY ? 0u : act >> 24;
  const u32 a_shop = PLAY_ONLY ? 0u : a_shop_in;
  const int p = (int)E.agent; /* relative byte 0 until the agent advances */
  E.info_steps = addb(E.info_steps, 0, 1); /* info.agent_infos[agent].steps_taken++ :97 */
  if (E.phase == PH_INACTIVE) env_cycle_phase(E);
  u32 pf = E.cpf, pc = E.cpc;
  pf += 1u << 24; /* Player::steps_taken (byte 3, wraps) */
  int special = SP_NONE;

  if (a_play) { /* Player::play_card src/player.cpp:47-69 */
    const u32 lc = bcast(E.lc, (int)a_play);
    if (E.phase == PH_MOVEMENT) {
      E.res[0] = 2 * (int)(lc & 7u);
      E.res[1] = 2 * (int)((lc >> 3) & 7u);
      E.res[2] = 2 * (int)((lc >> 6) & 7u);
    } else if (E.phase == PH_BUYING) {
      const int coin = (int)((lc >> 6) & 7u);
      E.res[2] += coin > 0 ? 2 * coin : 1;
    }
    deck_activate(E, pc, 0, (int)a_play - 1, lane);
  } else if (a_special) { /* Player::play_special src/player.cpp:71-79 */
    const u32 lc = bcast(E.lc, (int)a_special);
    if ((lc >> 9) & 1u) deck_remove_immediate(E, pc, 0, (int)a_special - 1, lane);
    else deck_play_immediate(E, pc, 0, (int)a_special - 1, lane);
    special = (int)((lc >> 10) & 7u);
  } else if (a_move) { /* Map::move_in_direction src/map.cpp:356-367 */
    const u32 c = bcast(getb(E.nbr, 0), (int)a_move);
    int dx, dy;
    dir_delta((int)a_move, dx, dy);
    const u32 loc = bcast(E.ploc, p);
    if (lane == p) E.ploc = loc_pack(loc_x(loc) + dx, loc_y(loc) + dy);
    nbr_refresh(E, B.grid + i * COGL_GRID_BYTES, p, 0, lane);
    if (!(pf & PFL_MOVE_FREE)) {
      player_handle_requirement(E, pf, pc, 0, hex_req(c), (u32)hex_n(c), lane);
    } else {
      pf &= ~(u32)PFL_MOVE_FREE;
      player_enable_playing(E, 0, lane);
    }
    if (lane == p) E.pmv++;
    pf = (pf & ~(u32)PFL_WON) | (u32)hex_end(c);
  } else {
    pf &= ~(u32)PFL_MOVE_FREE;
    if (a_shop) {
      const int idx = (int)a_shop - 1;
      int card;
      if (pf & PFL_CARD_FREE) {
        card = shop_get(E, idx, lane); /* Shop::transmit */
      } else {
        card = shop_buy(E, idx, lane);
        E.res[2] -= 2 * (int)((bcast(E.lc, card + 1) >> 13) & 7u); /* Player::pay */
        env_cycle_phase(E);
      }
      if (lane == card + 1) E.discard = addb(E.discard, 0, 1); /* Deck::add */
      pf = setb(pf, 2, getb(pf, 2) + 1);
    } else if (a_remove) {
      deck_remove_immediate(E, pc, 0, (int)a_remove - 1, lane);
      const u32 nr = (getb(pf, 1) - 1u) & 0xffu;
      pf = setb(pf, 1, nr);
      if (!nr) player_enable_playing(E, 0, lane);
      else special = SP_REMOVE_PENDING;
    } else {
      env_cycle_phase(E);
      if (getb(pf, 1) > 0) {
        pf = setb(pf, 1, 0);
        player_enable_playing(E, 0, lane);
      }
    }
    if (pf & PFL_CARD_FREE) {
      pf &= ~(u32)PFL_CARD_FREE;
      player_enable_playing(E, 0, lane);
    }
  }

  if ((pf & PFL_MIP) && !a_move) {
    pf &= ~(u32)PFL_MIP;
    res_clear(E);
  }

  /* maybe_end_turn :244-250 */
  const bool turn_ends = !NO_END && ((pf & PFL_WON) || E.phase == PH_INACTIVE);
  if (turn_ends) {
    env_next_agent(E, pf, pc, p, prm.n_players, lane);
  } else {
    E.cpf = pf;
    E.cpc = pc;
  }
  const int agent = (int)E.agent;
  u32 touched = LAZY ? 0xffu : (0x11u << p) | (0x10u << agent); /* piles: bit p, masks: bit 4 + p */
  u32 done = 0;

  if (COG_UNLIKELY(special != SP_NONE)) { /* :183-186; src/cards.cpp:8-36; remove lambda :156-158 */
    env_update_observation(E, agent, 0, E.cpc, lane);
    obs_pending = false;
    const int rp = rel(E, p); /* the player who acted; the agent may have advanced */
    u32 Pms = bcast(E.Pms, agent);
    switch (special) {
    case SP_TRANSMIT: /* transmit_action */
      Pms = 1u | ((1u | shop_transmit_bits(E, lane)) << 7);
      player_disable_playing(E);
      pf |= PFL_CARD_FREE;
      break;
    case SP_DRAW2: deck_draw(E, pc, rp, 2, lane); break;
    case SP_DRAW3: deck_draw(E, pc, rp, 3, lane); break;
    case SP_DRAW1_REMOVE1:
    case SP_DRAW2_REMOVE2: { /* draw_and_remove_action<n, m> */
      const u32 k = special == SP_DRAW1_REMOVE1 ? 1u : 2u;
      deck_draw(E, pc, rp, k, lane);
      pf = setb(pf, 1, k);
      if (lane == agent) E.Prem = E.Pplay; /* P.remove = (stale) P.play */
      player_disable_playing(E);
      Pms = (Pms & 0x7fu) | (((Pms >> 7) & 1u) << 7);
      break;
    }
    case SP_NATIVE: { /* native_action: any passable neighbour of the PLAYER who played it */
      const u32 mv = movement_bits(E, rp, 200, 200, 200, 100, lane);
      Pms = (Pms & 1u) | mv | (((Pms >> 7) & 1u) << 7);
      pf |= PFL_MOVE_FREE;
      player_disable_playing(E);
      break;
    }
    default: /* SP_REMOVE_PENDING: shop.set_available_mask(0, mask) */
      Pms = (Pms & 0x7fu) | (((Pms >> 7) & 1u) << 7);
      break;
    }
    if (lane == agent) E.Pms = Pms;
    if (p == agent) { /* still this player's turn: the live copies are the uniform ones */
      E.cpf = pf;
      E.cpc = pc;
    } else if (lane == p) {
      E.pf = pf;
      E.pc = pc;
    }
  } else {
    if (LAZY) {
      obs_pending = true;
    } else {
      const u32 pms = env_update_observation(E, agent, 0, E.cpc, lane);
      if (pms_out) *pms_out = pms;
    }
    /* Map::player_done src/map.cpp:407-410 for the NEW current agent, or truncation.
     * The current agent can only have ARRIVED on an end hex by moving or by the turn
     * passing to a player who already stands on one */
    bool finished = !NO_END && E.turn >= prm.max_steps;
    if (!NO_END && (turn_ends || a_move)) finished = finished || hex_end(bcast(getb(E.nbr, 0), 0));
    if (COG_UNLIKELY(finished)) {
      done = 1;
      obs_pending = false;
      touched = 0xffu; /* the caller runs env_reset next (auto-reset, the RNG stream continues) */
    }
  }
  return done | (touched << 8);
}

/* ----------------------------------------------------------------- sampler */
/* action_sampler::sample include/sampler.h:14-79: five heads in the order play,
 * play_special, remove, move, get_from_shop; each head draws uniformly among its set
 * bits (one draw even when only index 0 is valid).  `valid` are the head's bit sets. */
__device__ __forceinline__ u32 sample_head(u32 &rng, u32 valid) {
  if (valid == 1u) { /* only "none" is legal: the draw over [0, 0] is still consumed */
    while (COG_UNLIKELY(minstd_next(rng) == 2147483646u)) {}
    return 0;
  }
  if (!valid) return 0;
  const u32 j = uniform_draw(rng, (u32)__popc(valid));
  return nth_set_bit(valid, j);
}
/* The `play` head on the hot path: no branch on the number of legal indices (a head with
 * one legal index goes through the table like any other: scaling = past = 2147483645 gives
 * 0), the n-th set bit found by the lanes.  A rejected draw — or an empty head, whose table
 * entry has past = 0 — falls back to the step-by-step version with the state untouched. */
__device__ __forceinline__ u32 sample_head_warp(u32 &rng, u32 valid, int lane) {
  const ud_entry e = UD_TAB[__popc(valid)];
  u32 s1 = rng;
  const u32 r = minstd_next(s1) - 1u;
  if (COG_UNLIKELY(r >= e.past)) return sample_head(rng, valid);
  rng = s1;
  return nth_set_bit_warp(valid, ud_quotient(r, e), lane);
}
/* The MERGED legal mask of the current agent: the reference splits
 * the truth over two records — play / play_special live in S, move / get_from_shop / remove in
 * P[agent] — and a remove bit only makes sense for a card still in hand.  Needs an
 * up-to-date P[agent] (update_observation not deferred). */
__device__ __forceinline__ void merged_heads(const Env &E, int lane, u32 &rem, u32 &ms) {
  const int agent = (int)E.agent;
  const u32 have = __ballot_sync(COG_FULL, lane >= 1 && lane <= 21 && getb(E.hand, 0) > 0);
  rem = bcast(E.Prem, agent) & (have | 1u);
  ms = bcast(E.Pms, agent);
}

/* Four consecutive draws over [0, 0] (heads whose only legal index is "none") advance
 * the engine four times and reject only an output of 2147483646 = -1 (mod m).  Output i
 * equals x a^i, so no rejection happens unless x = -a^-i for some i in 1..4; otherwise
 * the four steps collapse into ONE modular multiply by a^4 mod m.  Returns false (state
 * untouched) in the 4-in-2^31 case that needs the step-by-step path. */
__device__ __forceinline__ bool minstd_skip4_unit_draws(u32 &s) {
  if (COG_UNLIKELY(s == 739806647u || s == 671875339u || s == 939811632u || s == 1785685951u)) return false;
  const u64 t = (u64)s * 984943658ull;                 /* 16807^4 mod (2^31 - 1) */
  const u64 f = (t & 0x7fffffffull) + (t >> 31);       /* < 2^32 */
  u32 r = (u32)(f & 0x7fffffffull) + (u32)(f >> 31);
  if (r >= 0x7fffffffu) r -= 0x7fffffffu;
  s = r;
  return true;
}

/* ---- all five heads at once (the general case: several heads have a choice) -----------------------
 * Draw k of a sample() call consumes engine output k + 1 as long as no earlier draw was rejected, and
 * output k + 1 is x a^(k+1) mod m: lane k (0..4) computes ITS output with one modular multiply by a
 * per-lane constant, takes its own head's table entry, and finds its own n-th set bit; one OR-reduction
 * and two shuffles collect the five indices.  A rejected draw in any lane (a few in 2^31) — or an empty
 * head — leaves the state untouched and returns false: the caller takes the step-by-step path. */
namespace tab {
constexpr u32 minstd_pow(int k) { /* 16807^k mod (2^31 - 1) */
  u64 v = 1;
  for (int i = 0; i < k; ++i) v = v * 16807ull % 2147483647ull;
  return (u32)v;
}
}  // namespace tab
__device__ const u32 MINSTD_POW_LANE[32] = {
    tab::minstd_pow(1), tab::minstd_pow(2), tab::minstd_pow(3), tab::minstd_pow(4), tab::minstd_pow(5),
    1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1};
static_assert(tab::minstd_pow(4) == 984943658u && tab::minstd_pow(5) == 1144108930u, "powers of the minstd multiplier");
/* the same table as UD_TAB for per-LANE indices (constant memory serialises divergent addresses) */
__device__ const ud_entry UD_TAB_LANE[33] = {
    COG_UD(0),  COG_UD(1),  COG_UD(2),  COG_UD(3),  COG_UD(4),  COG_UD(5),  COG_UD(6),  COG_UD(7),  COG_UD(8),
    COG_UD(9),  COG_UD(10), COG_UD(11), COG_UD(12), COG_UD(13), COG_UD(14), COG_UD(15), COG_UD(16), COG_UD(17),
    COG_UD(18), COG_UD(19), COG_UD(20), COG_UD(21), COG_UD(22), COG_UD(23), COG_UD(24), COG_UD(25), COG_UD(26),
    COG_UD(27), COG_UD(28), COG_UD(29), COG_UD(30), COG_UD(31), COG_UD(32)};

__device__ __forceinline__ bool sample_heads_parallel(u32 &rng, u32 play, u32 spec, u32 rem, u32 ms, u32 &act,
                                                      u32 &a_shop, int lane) {
  const u32 head = lane == 0 ? play : (lane == 1 ? spec : (lane == 2 ? rem : (lane == 3 ? (ms & 0x7fu) : (ms >> 7))));
  const u32 n = (u32)__popc(head); /* <= 22 */
  const u64 t = (u64)rng * MINSTD_POW_LANE[lane];   /* < 2^62 */
  const u64 f = (t & 0x7fffffffull) + (t >> 31);    /* < 2^32 */
  u32 o = (u32)(f & 0x7fffffffull) + (u32)(f >> 31);
  if (o >= 0x7fffffffu) o -= 0x7fffffffu;
  const u32 r = o - 1u;
  const ud_entry e = UD_TAB_LANE[n]; /* n == 0: past = 0, every r is "rejected" */
  if (COG_UNLIKELY(__any_sync(COG_FULL, lane < 5 && r >= e.past))) return false;
  u32 v = head;
  for (u32 j = ud_quotient(r, e); j; --j) v &= v - 1u; /* the j-th set bit of this lane's head */
  const u32 idx = (u32)__ffs((int)v) - 1u;
  act = __reduce_or_sync(COG_FULL, lane < 4 ? idx << (8 * lane) : 0u);
  a_shop = bcast(idx, 4);
  rng = bcast(o, 4);
  return true;
}

/* The heads are bit sets with nothing above bit 21 (play, special, remove) / bit 25 (ms =
 * move | get_from_shop << 7): every producer builds them from lane-guarded ballots. */
/* Returns true when only the `play` head had a choice (act < 256, a_shop = 0). */
__device__ __forceinline__ bool sample_action(u32 &rng, u32 play, u32 spec, u32 rem, u32 ms, u32 &act,
                                              u32 &a_shop, int lane) {
  u32 a0;
  if (spec == 1u && rem == 1u && ms == 0x81u) { /* the reference's own flow: only `play` ever has a choice (SURVEY A#2) */
    a0 = sample_head_warp(rng, play, lane);
    if (COG_LIKELY(minstd_skip4_unit_draws(rng))) {
      a_shop = 0;
      act = a0;
      return true;
    }
  } else {
    if (COG_LIKELY(sample_heads_parallel(rng, play, spec, rem, ms, act, a_shop, lane))) return false;
    a0 = sample_head(rng, play);
  }
  /* step by step: a draw was rejected somewhere (or a head is empty) */
  const u32 mv = ms & 0x7fu, shop = ms >> 7;
  const u32 a1 = sample_head(rng, spec);
  const u32 a2 = sample_head(rng, rem);
  const u32 a3 = sample_head(rng, mv);
  a_shop = sample_head(rng, shop);
  act = a0 | (a1 << 8) | (a2 << 16) | (a3 << 24);
  return false;
}

}  // namespace cog
#endif /* COG_DEVICE_CUH */
