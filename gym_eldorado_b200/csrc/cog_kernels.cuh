/* cog_kernels.cuh — the __global__ entry points (one warp per environment).
 *
 *   k_init        state of a default-constructed vec_cog_env<N>
 *   k_reset       vec_cog_env<N>::reset(...)  — board generation, decks, shop, masks, map encode
 *   k_step_single vec_cog_env<N>::step(actions)  — single step, matches the Python API
 *   k_step_fused  persistent fused K x {sample(selected masks); step} with no host round trip
 *   k_sample      vec_action_sampler<N>::sample(masks) on an arbitrary mask array
 */
#ifndef COG_KERNELS_CUH
#define COG_KERNELS_CUH
#include "cog_device.cuh"

namespace cog {

enum { WARPS_PER_BLOCK = 4, BLOCK_THREADS = WARPS_PER_BLOCK * 32 };

extern __shared__ uint4 cog_smem[];

/* Env indices are 32-bit (a handle holds fewer than 2^32 envs; checked by cog_vec_create); the
 * 64-bit value the address arithmetic uses is a plain zero extension, so every `i * record size`
 * is one IMAD.WIDE.U32 instead of a 64 x 64-bit multiply. */
__device__ __forceinline__ u64 env_index(u32 i) { return (u64)i; }

/* programmatic dependent launch (no-ops for a kernel launched without the attribute).  Used by k_step_single
 * only: 38.9 -> 36.5 us per sampled step of 65 536 envs.  (Tried on k_sample as well: the two-kernel loop
 * got slower, 49.2 -> 52.6 us — its 16 384 short blocks become resident early and wait.) */
__device__ __forceinline__ void grid_dependency_wait() {
#ifdef __CUDACC__
  asm volatile("griddepcontrol.wait;" ::: "memory");
#endif
}
__device__ __forceinline__ void grid_dependency_launch() {
#ifdef __CUDACC__
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif
}

__device__ __forceinline__ BoardScratch &warp_scratch() {
  return reinterpret_cast<BoardScratch *>(cog_smem)[threadIdx.x >> 5];
}

/* vec_cog_env<N>() include/vec_environment.h:23-31 -> cog_env::init src/environment.cpp:25-40:
 * Shop() puts 6 cards in the market and 3 copies in stock (src/cards.cpp:80-92), every
 * ActionMask is reset (include/api.h:101-103); everything else is zero. */
__global__ void k_init(Buffers B, u64 n) {
  const int lane = threadIdx.x & 31;
  const u64 i = env_index(blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5));
  if (i >= n) return;
  Env E;
  E.draw = E.hand = E.active = E.played = E.discard = 0;
  E.shop = (lane >= 1 && lane <= 18) ? 3u : 0u;
  E.nbr = 0;
  E.lc = LANE_CONST[lane];
  E.pf = E.pc = E.pmv = E.ploc = 0;
  E.cpf = E.cpc = 0;
  E.Pplay = E.Prem = 1u;
  E.Pms = 1u | (1u << 7);
  E.rng = 1;
  E.turn = 0; E.agent = 0; E.phase = 0; E.n_in_market = 6; E.info_steps = 0;
  E.Splay = E.Srem = 1u;
  E.Sms = 1u | (1u << 7);
  E.res[0] = E.res[1] = E.res[2] = 0;
  E.minx = E.miny = 0;
  E.dead = 0;
  env_store(E, B.hidden + i * COG_HID_WORDS, lane);
  export_dynamic(E, B, i, 0xffu, lane);
}

/* re-broadcast of the env scalars from lane 0: the same values, but provably warp-uniform for ptxas */
__device__ __forceinline__ void env_prove_uniform(Env &E) {
  E.rng = bcast(E.rng, 0); E.turn = bcast(E.turn, 0); E.agent = bcast(E.agent, 0);
  E.phase = bcast(E.phase, 0); E.n_in_market = bcast(E.n_in_market, 0);
  E.cpf = bcast(E.cpf, 0); E.cpc = bcast(E.cpc, 0); E.info_steps = bcast(E.info_steps, 0);
  E.Splay = bcast(E.Splay, 0); E.Srem = bcast(E.Srem, 0); E.Sms = bcast(E.Sms, 0);
  E.res[0] = (int)bcast((u32)E.res[0], 0); E.res[1] = (int)bcast((u32)E.res[1], 0);
  E.res[2] = (int)bcast((u32)E.res[2], 0);
  E.minx = (int)bcast((u32)E.minx, 0); E.miny = (int)bcast((u32)E.miny, 0);
}

/* vec_cog_env<N>::reset include/vec_environment.h:33-44.  reseed: env i restarts its RNG
 * from uint32(seed_base + i) (cog_env::reset(seed, ...) src/environment.cpp:66-77). */
/* warps (= envs) per block of the kernels whose blocks the hardware scheduler hands out one env per warp: a block
 * keeps its slot until its SLOWEST warp is done, and board generation times differ a lot between envs */
#ifndef COG_RESET_WARPS
#define COG_RESET_WARPS 4
#endif
#ifndef COG_SINGLE_WARPS
#define COG_SINGLE_WARPS 4
#endif
#ifndef COG_RESET_BLOCKS
#define COG_RESET_BLOCKS 12 /* with provably uniform control flow, occupancy beats spills: 7 blocks (72 registers) 1.39e8,
                             * 9 1.44e8, 12 (40 registers) 1.50e8, 14 / 16 (32) 1.49e8 resets/s (profiles/r02_y_resident_blocks.txt) */
#endif
__global__ void __launch_bounds__(COG_RESET_WARPS * 32, COG_RESET_BLOCKS) k_reset(Buffers B, Params prm, u64 n, u64 seed_base, int reseed) {
  const int lane = threadIdx.x & 31;
#ifndef COG_NO_UNIFORM_PROOF
  /* warp index and env scalars through shuffles from a fixed lane: board generation is all warp-uniform
   * control flow, and ptxas only compiles it that way when it can prove it (see k_step_fused): 4 016 -> 2 800
   * instructions, BSSY 112 -> 26, BRA.DIV 41 -> 0; 1.21 -> 1.29e8 resets/s */
  const u64 i = env_index(blockIdx.x * COG_RESET_WARPS + bcast(threadIdx.x >> 5, 0));
#else
  const u64 i = env_index(blockIdx.x * COG_RESET_WARPS + (threadIdx.x >> 5));
#endif
  if (i >= n) return;
  Env E;
  env_load(E, B.hidden + i * COG_HID_WORDS, lane);
#ifndef COG_NO_UNIFORM_PROOF
  env_prove_uniform(E);
#endif
  if (reseed) E.rng = minstd_seed((u64)(u32)(seed_base + i));
  const int rc = env_reset(E, prm, B, i, warp_scratch(), lane);
  env_store(E, B.hidden + i * COG_HID_WORDS, lane, 0, rc ? 1u : 0u); /* a failed env ignores steps until a reset succeeds */
  export_dynamic(E, B, i, 0xffu, lane, false);
  /* the single env's own flag: cog_env::reset() clears `done` (src/environment.cpp:57); the vector env's
   * dones[] array is only written by step_single (include/vec_environment.h:53-61) */
  if ((prm.flags & COG_PRM_NO_AUTORESET) && lane == 0) B.dones[i] = 0;
}

/* vec_cog_env<N>::step(actions): one step per env, auto-reset included; exports only what the step
 * changed.  SAMPLED = false: the actions come from `actions_in` (ActionData[n], 64-byte records; may be
 * mapped pinned HOST memory — each warp pulls its 8 action bytes over PCIe itself).  SAMPLED = true
 * (cog_vec_step_sampled): vec_action_sampler::sample(selected_action_masks) and the step in ONE
 * kernel — the action is drawn from the env's own S mask (already in its scalar line), written to
 * the sampler's ActionData buffer and applied; bit-identical to k_sample followed by the step.
 * Works on the envs [i0, n) (i0 a multiple of the warps per block): the host-buffer entry points run
 * it chunk by chunk so that copies and kernels of different chunks overlap.
 *
 * Two paths.  FAST: a card play, or a null action in the movement phase (78 % of the steps of the
 * reference's own flow): the turn cannot end, so the step reads the scalar line and four per-lane
 * lines (hand, active, shop, neighbour cache: 12 sectors instead of the record's 36), writes back the
 * scalar line, one word each of hand / active and the agent's move / shop mask word, and exports the
 * bytes that changed (two DeckObs bytes, two mask bytes of S, the 26 move / get_from_shop bytes of
 * P[agent], phase + resources, steps_taken).  GENERAL: everything else — the whole record, with the
 * episode end (rewards, Info, board generation) behind it. */
#ifndef COG_SINGLE_BLOCKS
#define COG_SINGLE_BLOCKS 10 /* measured: 8 blocks 43.3 us, 10 blocks 38.3 us, 12 blocks 41.3 us (spills) per sampled step of 65 536 envs */
#endif
/* act_stride: bytes between the action records of consecutive envs in `actions_in` — COGL_ACTION_SIZE
 * for ActionData[n], COGL_PACKED_ACTION_SIZE for the packed form (the same five bytes + 3 of padding).
 * pmasks (optional): selected_action_masks in packed form (COGL_PACKED_MASK_SIZE bytes per env: the four
 * heads as bit sets), for the packed host-I/O entry points. */
template <bool SAMPLED>
__device__ __forceinline__ void step_one_env(const Buffers &B, const Params &prm, u64 i, const u8 *__restrict__ actions_in,
                                             u32 act_stride, u8 *masks2, u8 *pmasks, u32 *srng, u8 *actions_out, int lane);

/* One warp per env; the hardware block scheduler hands out the envs four at a time.  (Tried and dropped: a
 * resident grid whose warps draw env indices from a device counter, the next index fetched while the
 * current env is stepped — 65 536 atomics on one address per launch serialise in L2: 66 us per step of
 * 65 536 envs against 40 us, gpurun_out/r02_d_api_loop.txt.) */
template <bool SAMPLED>
__global__ void __launch_bounds__(COG_SINGLE_WARPS * 32, COG_SINGLE_BLOCKS) k_step_single(Buffers B, Params prm, u64 i0, u64 n,
                                                                  const u8 *__restrict__ actions_in, u32 act_stride,
                                                                  u8 *masks2, u8 *pmasks, u32 *srng, u8 *actions_out) {
  const int lane = threadIdx.x & 31;
  grid_dependency_launch(); /* the next kernel of the stream may become resident as this grid drains ... */
  grid_dependency_wait();   /* ... and this one touches memory only after the previous one has finished */
#ifndef COG_NO_UNIFORM_PROOF
  /* as in k_step_fused: warp index and env scalars through shuffles from a fixed lane, so that
   * ptxas can prove them warp-uniform */
  const u64 i = env_index((u32)i0 + blockIdx.x * COG_SINGLE_WARPS + bcast(threadIdx.x >> 5, 0));
#else
  const u64 i = env_index((u32)i0 + blockIdx.x * COG_SINGLE_WARPS + (threadIdx.x >> 5));
#endif
  if (i >= n) return;
  step_one_env<SAMPLED>(B, prm, i, actions_in, act_stride, masks2, pmasks, srng, actions_out, lane);
}

template <bool SAMPLED>
__device__ __forceinline__ void step_one_env(const Buffers &B, const Params &prm, u64 i, const u8 *__restrict__ actions_in,
                                             u32 act_stride, u8 *masks2, u8 *pmasks, u32 *srng, u8 *actions_out, int lane) {
  u32 *rec = B.hidden + i * COG_HID_WORDS;
  Env E;
  E.draw = E.played = E.discard = 0;
  E.pf = E.pc = E.pmv = E.ploc = E.Pplay = E.Prem = E.Pms = 0;
  /* one round trip for everything the fast path reads (and every other step reads as well): the
   * scalar line, four per-lane lines and the action (or the sampler state) */
  env_load_scalars(E, rec, lane);
  enum : u32 { FAST_LINES = (1u << COG_HL_HAND) | (1u << COG_HL_ACTIVE) | (1u << COG_HL_SHOP) | (1u << COG_HL_NBR) };
  env_load_lines<FAST_LINES>(E, rec, lane);
  u32 act, a_shop;
  if (SAMPLED) {
    u32 sr = srng[i];
    env_uniform_scalars_hot(E);
#ifndef COG_NO_UNIFORM_PROOF
    sr = bcast(sr, 0);
#endif
    sample_action(sr, E.Splay, spec_of(E.Splay), E.Srem, E.Sms, act, a_shop, lane);
    __syncwarp(); /* every lane has read srng[i] before lane 0 overwrites it */
    if (lane == 0) {
      srng[i] = sr;
      *reinterpret_cast<uint2 *>(actions_out + i * COGL_ACTION_SIZE) = make_uint2(act, a_shop);
    }
  } else {
    /* the five ActionData bytes: every lane reads the same 8 bytes (one transaction) */
    const uint2 a = *reinterpret_cast<const uint2 *>(actions_in + i * act_stride);
    env_uniform_scalars_hot(E);
#ifndef COG_NO_UNIFORM_PROOF
    act = bcast(a.x, 0);
    a_shop = bcast(a.y, 0) & 0xffu;
#else
    act = a.x;
    a_shop = a.y & 0xffu;
#endif
  }
  /* dead_step src/environment.cpp:92-95: a finished single env (auto-reset off) — or an env whose board
   * generation failed — ignores steps until reset().  (In SAMPLED mode its draw is still consumed, like the
   * reference's sampler, which knows nothing about the env.) */
  if (COG_UNLIKELY(E.dead)) return;
  bool obs_pending = false;
  const u32 phase1 = E.phase == PH_INACTIVE ? (u32)PH_MOVEMENT : E.phase;
  const bool fast = act < 256u && !a_shop && !(E.cpf & PFL_WON) && E.turn < prm.max_steps &&
                    (act || phase1 == PH_MOVEMENT);
  u8 *obs = B.obs + i * COGL_OBS_SIZE;
  if (fast) {
    const u32 Splay0 = E.Splay, Srem0 = E.Srem;
    const int p = (int)E.agent;
    u32 pms = 0x81u;
    env_step<false, true, true>(E, act, 0u, prm, B, i, obs_pending, lane, &pms);
    /* write-back: only what a play / a phase change touches */
    u8 *pd = obs + COGL_OBS_PLAYER + COGL_PLAYER_STRIDE * p;
    if (act && lane == (int)act) { /* Deck::activate: hand -> active of one card type */
      rec[COG_HL_HAND * 32 + lane] = E.hand;
      rec[COG_HL_ACTIVE * 32 + lane] = E.active;
      pd[COGL_DK_HAND + lane - 1] = (u8)getb(E.hand, 0);
      pd[COGL_DK_ACTIVE + lane - 1] = (u8)getb(E.active, 0);
    }
    if (lane == p) rec[COG_HL_PLAYER * 32 + COG_PF_STRIDE * lane + COG_PF_MMS] = pms;
    if (COG_LIKELY(E.Srem == Srem0)) env_store_scalars<true>(E, rec, lane); /* rng, turn, S.move / shop: untouched */
    else env_store_scalars(E, rec, lane);
    /* export */
    if (lane == 0) {
      *reinterpret_cast<uint4 *>(obs + COGL_OBS_PHASE) =
          make_uint4(E.phase, __float_as_uint(0.5f * (float)E.res[0]), __float_as_uint(0.5f * (float)E.res[1]),
                     __float_as_uint(0.5f * (float)E.res[2]));
      (B.infos + i * COGL_INFO_SIZE)[COGL_INFO_AGENT + COGL_INFO_AGENT_STRIDE * p] = (u8)getb(E.info_steps, 0);
      B.dones[i] = 0;
    }
    /* P[agent].move (7 bytes at 66) and .get_from_shop (19 bytes at 73) are 26 contiguous bools:
     * byte 66 + k = bit k of the mask word */
    if (lane < 26) pd[COGL_PD_MASK + COGL_AM_MOVE + lane] = (u8)((pms >> lane) & 1u);
    /* S: a play clears at most one bit of `play` (and its play_special twin); anything else
     * (enable_playing after a pending removal / transmission) rewrites the record */
    const u32 d = E.Splay ^ Splay0;
    if (COG_LIKELY(E.Srem == Srem0 && !(d & (d - 1u)))) {
      if ((d >> lane) & 1u) {
        const u8 bit = (u8)((E.Splay >> lane) & 1u), sbit = (u8)((spec_of(E.Splay) >> lane) & 1u);
        u8 *m = B.masks + i * COGL_MASK_SIZE;
        m[COGL_AM_PLAY + lane] = bit;
        m[COGL_AM_SPECIAL + lane] = sbit;
        if (masks2) {
          u8 *m2 = masks2 + i * COGL_MASK_SIZE;
          m2[COGL_AM_PLAY + lane] = bit;
          m2[COGL_AM_SPECIAL + lane] = sbit;
        }
      }
    } else {
      export_mask(B.masks + i * COGL_MASK_SIZE, E.Splay, spec_of(E.Splay), E.Srem, E.Sms, lane);
      if (masks2) export_mask(masks2 + i * COGL_MASK_SIZE, E.Splay, spec_of(E.Splay), E.Srem, E.Sms, lane);
    }
  } else {
    env_load_lines<COG_LINES_ALL & ~FAST_LINES>(E, rec, lane);
    env_uniform_scalars_cold(E);
    const u32 status = env_step<false, false>(E, act, a_shop, prm, B, i, obs_pending, lane);
    u32 dead = 0;
    if (status & 1u) {
      const u32 agent = E.agent;
      env_finish_episode(E, prm, B, i, lane);
      if (COG_UNLIKELY(prm.flags & COG_PRM_NO_AUTORESET)) {
        env_from_absolute(E, agent, lane); /* the terminal state stays (cog_env::step does not reset) */
        dead = 1;
      } else if (env_reset(E, prm, B, i, warp_scratch(), lane)) {
        dead = 1; /* board generation failed (reported at the next synchronising call): a defined state */
      }
    }
    env_store(E, rec, lane, 0, dead);
    export_dynamic(E, B, i, status >> 8, lane);
    /* host-buffer entry point with mapped pinned memory: selected_action_masks goes straight to
     * the caller's host array as well (92 contiguous bytes per env, posted writes over PCIe) */
    if (masks2) export_mask(masks2 + i * COGL_MASK_SIZE, E.Splay, spec_of(E.Splay), E.Srem, E.Sms, lane);
    if (lane == 0) B.dones[i] = (u8)(status & 1u);
  }
  if (pmasks && lane == 0)
    *reinterpret_cast<uint4 *>(pmasks + i * COGL_PACKED_MASK_SIZE) = make_uint4(E.Splay, E.Srem, E.Sms, spec_of(E.Splay));
  /* (COG_STAT_ENV_STEPS of single steps is counted by the host: n per call) */
}

/* the sampler's five action bytes of env i */
__device__ __forceinline__ void store_action(u8 *actions_out, u64 i, u32 act, u32 a_shop, int lane) {
  if (lane < 4) actions_out[i * COGL_ACTION_SIZE + lane] = (u8)(act >> (8 * lane));
  if (lane == 4) actions_out[i * COGL_ACTION_SIZE + 4] = (u8)a_shop;
}

/* Persistent fused path: k_steps x {vec_sampler.sample(selected masks); vec_env.step} with
 * no host round trip.  Each warp takes envs from a ticket counter, one at a time, and keeps the env in
 * registers for all its iterations.
 *
 * Two kernels per launch.  k_step_fused is the hot loop and contains no episode-end code:
 * an env whose episode ends leaves the loop and parks the iterations it still owes in its
 * hidden record (COG_ES_PENDING).  k_step_tail, enqueued right behind it, picks those envs
 * up, does the episode-end bookkeeping (rewards, Info, aggregates), resets them (board
 * generation) and runs the rest with the general loop; it exits at once when nothing was
 * parked.  The hot kernel is small (about 3 000 instructions against 9 000 with board
 * generation inlined) and its register allocation is not shaped by the cold paths. */
#ifndef COG_TAIL_BLOCKS
#define COG_TAIL_BLOCKS 16 /* board generation is latency-bound: occupancy beats spills.  Reset-heavy configs[3] shape, with
                            * the uniform proof: 6 blocks 2.33e9, 8 2.41e9, 10 2.46e9 env-steps/s; with dirty-chunk clearing
                            * 10 / 12 / 16 blocks 2.60 / 2.54 / 2.60e9; regime M with episode ends (EASY, max_steps 40)
                            * 2.38 / 2.45 / 2.59e9 (profiles/r02_y_resident_blocks.txt) */
#endif
/* resident blocks per SM of the fused kernels that also write the rollout ring: the record
 * assembly wants registers more than a ninth block (measured: 7 blocks 3.31e9, 8 3.17e9, 9 2.97e9
 * env-steps/s sampling from selected_action_masks; 1.90 / 1.95 / 1.86e9 from merged masks) */
#ifndef COG_RING_BLOCKS
#define COG_RING_BLOCKS 7
#endif
#ifndef COG_RING_BLOCKS_MERGED
#define COG_RING_BLOCKS_MERGED 8
#endif
#ifndef COG_MIN_BLOCKS
#define COG_MIN_BLOCKS 16 /* full occupancy at 32 registers, no spills: 9 blocks (52 registers) 5.25e9, 12 (40) 5.58e9, 13-16 (32) 5.60e9 env-steps/s; regime M 3.19 / 3.41 / 3.53e9 (profiles/r02_y_resident_blocks.txt) */
#endif
/* MERGED = false: actions sampled from selected_action_masks (the reference's own flow,
 * "regime A": only `play` ever has a choice).  MERGED = true: from the merged legal mask of
 * the current agent ("regime M": moves, purchases, removals, wins — every branch of step). */
/* ROLLOUT = true: a compact record of the state, the mask heads sampled from, the action and
 * the outcome goes to `rec` — two coalesced stores per env-step. */
template <bool MERGED, bool ROLLOUT>
__device__ __forceinline__ u32 sample_and_step(Env &E, u32 &sr, u32 &act, u32 &a_shop, const Params &prm,
                                               const Buffers &B, u64 i, bool &obs_pending, u8 *rec, u32 t,
                                               int lane) {
  u32 status;
  if (MERGED) {
    u32 rem, ms;
    merged_heads(E, lane, rem, ms);
    sample_action(sr, E.Splay, spec_of(E.Splay), rem, ms, act, a_shop, lane);
    if (ROLLOUT) compact_store_pre(rec, E, prm, E.Splay, spec_of(E.Splay), rem, ms, act, a_shop, t, lane);
    /* most steps are still card plays and nulls: they take the instantiation whose other branches fold away */
#ifndef COG_NO_M_DISPATCH
    if (act < 256u && !a_shop) status = env_step<false, true>(E, act, 0u, prm, B, i, obs_pending, lane);
    else
#endif
      status = env_step<false, false>(E, act, a_shop, prm, B, i, obs_pending, lane);
  } else {
    const bool play_only = sample_action(sr, E.Splay, spec_of(E.Splay), E.Srem, E.Sms, act, a_shop, lane);
    if (ROLLOUT) compact_store_pre(rec, E, prm, E.Splay, spec_of(E.Splay), E.Srem, E.Sms, act, a_shop, t, lane);
    if (play_only) status = env_step<true, true>(E, act, 0, prm, B, i, obs_pending, lane);
    else status = env_step<true, false>(E, act, a_shop, prm, B, i, obs_pending, lane);
  }
  /* rewards stay zero here: the kernel that finishes the episode fills them in */
  if (ROLLOUT) compact_store_post(rec, status & 1u, 1u, lane);
  return status;
}
/* the record of slot `slot` of env i.  A 32-bit record index (the host checks n * T < 2^32)
 * keeps this at two instructions wherever the compiler chooses to recompute it. */
__device__ __forceinline__ u8 *rollout_rec(const Rollout &R, u32 i, u32 slot) {
  return R.ring + (u64)(i * R.T + slot) * COGL_CR_SIZE;
}

template <bool MERGED, bool ROLLOUT>
__global__ void __launch_bounds__(BLOCK_THREADS, ROLLOUT ? (MERGED ? COG_RING_BLOCKS_MERGED : COG_RING_BLOCKS) : COG_MIN_BLOCKS) k_step_fused(Buffers B, Params prm, u64 n, u32 *srng,
                                                                             u8 *actions_out, u32 k_steps, Rollout R) {
  const int lane = threadIdx.x & 31;
#ifndef COG_NO_UNIFORM_PROOF
  /* The warp index goes through a shuffle from a fixed lane.  The value is the same, but now
   * ptxas can PROVE the env loop — and, with the re-broadcast after env_load below, every env
   * scalar — warp-uniform: it drops the BSSY/BSYNC pairs and BRA.DIV guards from the step loop
   * (178 -> 24 and 51 -> 0 in the kernel, 3 976 -> 2 880 instructions) and moves scalar work
   * to the uniform datapath, off the ALU pipe that bounds this kernel (+5 % env-steps/s).
   * The proof is fragile (it also depends on the loop-carried counters below being there:
   * without them ptxas falls back to divergent code) and an earlier layout of the kernel
   * produced wrong counters in this mode, so the result is pinned by tests:
   * tests/test_gpu_parity.py::test_fused_equals_single_steps (12000 envs: several envs per warp,
   * episodes ending inside the launch) and tools/variant_check.py against a build with
   * -DCOG_NO_UNIFORM_PROOF. */
  const u32 warp0 = blockIdx.x * WARPS_PER_BLOCK + bcast(threadIdx.x >> 5, 0);
#else
  const u32 warp0 = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
#endif
  u64 steps_done = 0;
  u32 n_parked = 0;
#ifndef COG_STATIC_ENVS
  /* ENVS BY TICKET.  The warp schedulers do not advance the resident warps evenly: with a static grid-stride
   * split (7 envs per warp at 65 536 envs) the warps that are served first leave early and the launch ends
   * with a long stretch of few active warps per SM (ncu: 65 % of the resident warps active on average).  Here
   * one counter hands out the envs: a warp that finishes takes the next one.  Measured against the static
   * split (profiles/r02_ab_ticketed_units.txt): regime A 5.60 -> 5.99e9 env-steps/s, regime M 3.57 -> 3.91e9,
   * regime M with episode ends 2.75 -> 3.02e9, reset-heavy configs[3] 2.69 -> 2.86e9; results identical (which
   * warp steps an env changes nothing).  Splitting each env's steps into several smaller units to shorten the
   * end of the launch further was measured as well and lost (one record round trip per unit: 5.4-5.5e9).
   * The ticket comes back through a shuffle from lane 0, and the NEXT ticket is drawn in the loop body rather
   * than at the loop's head: this is the shape ptxas still proves warp-uniform (a draw at the head costs the
   * proof: BSSY 33 -> 204, BRA.DIV 0 -> 56, spills).  k_step_tail, which always follows, rewinds the counter. */
  (void)warp0;
  if (blockIdx.x == 0 && threadIdx.x == 0) B.park[3] = 0; /* k_step_tail's tickets */
  u32 next = 0;
  if (lane == 0) next = atomicAdd(&B.park[2], 1u);
  for (next = bcast(next, 0); next < (u32)n;) {
    const u32 i32 = next;
    if (lane == 0) next = atomicAdd(&B.park[2], 1u);
    next = bcast(next, 0);
#else
  const u32 stride = gridDim.x * WARPS_PER_BLOCK;
  for (u32 i32 = warp0; i32 < (u32)n; i32 += stride) {
#endif
    const u64 i = env_index(i32);
    Env E;
    env_load(E, B.hidden + i * COG_HID_WORDS, lane);
    u32 sr = srng[i];
#ifndef COG_NO_UNIFORM_PROOF
    /* the same values, provably warp-uniform again (loads are not, shuffles from lane 0 are) */
    env_prove_uniform(E);
    sr = bcast(sr, 0);
#endif
    if (COG_UNLIKELY(E.dead)) continue; /* failed board generation: ignored until reset() */
    u32 act = 0, a_shop = 0, status = 0;
    bool obs_pending = false;
    u32 k = k_steps;
    u32 slot = R.slot0, t = R.t0;
    do {
      status = sample_and_step<MERGED, ROLLOUT>(E, sr, act, a_shop, prm, B, i, obs_pending,
                                                ROLLOUT ? rollout_rec(R, (u32)i, slot) : nullptr, t, lane);
      if (ROLLOUT) {
        ++t;
        if (++slot == R.T) slot = 0;
      }
      --k;
    } while (k && !(status & 1u));
    steps_done += k_steps - k;
    const u32 parked = (status & 1u) ? k + 1u : 0u;
    if (obs_pending) env_update_observation(E, (int)E.agent, 0, E.cpc, lane);
    __syncwarp(); /* every lane has read srng[i] before lane 0 overwrites it */
    if (lane == 0) srng[i] = sr;
    store_action(actions_out, i, act, a_shop, lane);
    env_store(E, B.hidden + i * COG_HID_WORDS, lane, parked);
    if (parked) {
      ++n_parked;
    } else {
      export_dynamic(E, B, i, 0xffu, lane);
      if (lane == 0) B.dones[i] = 0;
    }
  }
  if (lane == 0) {
    if (steps_done) atomicAdd(&B.stats[0], (unsigned long long)steps_done);
    if (n_parked) atomicAdd(&B.park[0], n_parked);
  }
}

/* second kernel of a fused launch: the envs whose episode ended in k_step_fused.
 * park[0] counts envs ever parked, park[1] envs ever picked up again. */
template <bool MERGED, bool ROLLOUT>
__global__ void __launch_bounds__(BLOCK_THREADS, COG_TAIL_BLOCKS) k_step_tail(Buffers B, Params prm, u64 n, u32 *srng,
                                                                             u8 *actions_out, u32 k_steps, Rollout R) {
  const int lane = threadIdx.x & 31;
#ifndef COG_STATIC_ENVS
  if (blockIdx.x == 0 && threadIdx.x == 0) B.park[2] = 0; /* k_step_fused's tickets: rewound for the next launch */
#endif
  if (*(volatile unsigned *)&B.park[0] == *(volatile unsigned *)&B.park[1]) return; /* nothing parked */
  /* the uniform proof of k_step_fused, here as well — after the load AND after every reset (the scalars
   * board generation leaves behind are not provably uniform either): BSSY 322 -> 51, BRA.DIV 103 -> 0;
   * reset-heavy configs[3] 1.91 -> 2.42e9 env-steps/s, regime M with episode ends 1.66 -> 2.17e9 */
#ifndef COG_NO_UNIFORM_PROOF
  const u32 warp0 = blockIdx.x * WARPS_PER_BLOCK + bcast(threadIdx.x >> 5, 0);
#else
  const u32 warp0 = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
#endif
  const u32 stride = gridDim.x * WARPS_PER_BLOCK;
  BoardScratch &sc = warp_scratch();
  u64 steps_done = 0;
  u32 n_handled = 0;
#ifndef COG_STATIC_ENVS
  /* envs by ticket, as in k_step_fused: what a parked env still owes differs by orders of magnitude
   * (a reset and a few steps, or most of the launch), a static split leaves warps idle */
  (void)warp0; (void)stride;
  u32 next = 0;
  if (lane == 0) next = atomicAdd(&B.park[3], 1u);
  for (next = bcast(next, 0); next < (u32)n;) {
    const u32 i32 = next;
    if (lane == 0) next = atomicAdd(&B.park[3], 1u); /* drawn in the body: see k_step_fused */
    next = bcast(next, 0);
#else
  for (u32 i32 = warp0; i32 < (u32)n; i32 += stride) {
#endif
    const u64 i = env_index(i32);
#ifndef COG_NO_UNIFORM_PROOF
    const u32 parked = bcast(B.hidden[i * COG_HID_WORDS + COG_HL_ENV * 32 + COG_ES_PENDING], 0);
#else
    const u32 parked = B.hidden[i * COG_HID_WORDS + COG_HL_ENV * 32 + COG_ES_PENDING];
#endif
    if (!parked) continue;
    ++n_handled;
    Env E;
    env_load(E, B.hidden + i * COG_HID_WORDS, lane);
    u32 sr = srng[i];
#ifndef COG_NO_UNIFORM_PROOF
    env_prove_uniform(E);
    sr = bcast(sr, 0);
#endif
    u32 act = 0, a_shop = 0, status = 1;
    bool obs_pending = false;
    const u32 k = parked - 1u;
    /* the step that parked the env was step k_steps - parked of this launch */
    u32 slot = 0, t = 0;
    u8 *rec = nullptr;
    if (ROLLOUT) {
      const u32 j_done = k_steps - parked;
      slot = (u32)(((u64)R.slot0 + j_done) % R.T);
      t = R.t0 + j_done;
      rec = rollout_rec(R, (u32)i, slot);
    }
    u32 dead = 0;
    for (u32 j = 0;; ++j) {
      if (status & 1u) { /* rewards / Info / aggregates, then auto-reset: the RNG stream continues */
        env_finish_episode(E, prm, B, i, lane);
        if (ROLLOUT) compact_store_rewards(rec, E, prm, lane);
        if (COG_UNLIKELY(env_reset(E, prm, B, i, sc, lane) != 0)) {
          dead = 1; /* reported at the next synchronising call; the env ignores steps until reset() */
          break;
        }
#ifndef COG_NO_UNIFORM_PROOF
        env_prove_uniform(E);
#endif
      }
      if (j == k) break;
      if (ROLLOUT) {
        ++t;
        if (++slot == R.T) slot = 0;
        rec = rollout_rec(R, (u32)i, slot);
      }
      status = sample_and_step<MERGED, ROLLOUT>(E, sr, act, a_shop, prm, B, i, obs_pending, rec, t, lane);
    }
    steps_done += k;
    if (obs_pending) env_update_observation(E, (int)E.agent, 0, E.cpc, lane);
    __syncwarp();
    if (lane == 0) srng[i] = sr;
    if (k) store_action(actions_out, i, act, a_shop, lane);
    env_store(E, B.hidden + i * COG_HID_WORDS, lane, 0, dead);
    export_dynamic(E, B, i, 0xffu, lane);
    if (lane == 0) B.dones[i] = (u8)(status & 1u);
  }
  /* park[1] can only reach park[0] once EVERY parked env was picked up, so a warp that
   * starts late never sees the counters equal while it still has work */
  if (lane == 0) {
    if (steps_done) atomicAdd(&B.stats[0], (unsigned long long)steps_done);
    if (n_handled) atomicAdd(&B.park[1], n_handled);
  }
}

/* the merged legal mask of every env's current agent as ActionMask[n] records */
__global__ void k_merged_masks(Buffers B, u64 n, u8 *out) {
  const int lane = threadIdx.x & 31;
  const u64 i = env_index(blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5));
  if (i >= n) return;
  Env E;
  env_load(E, B.hidden + i * COG_HID_WORDS, lane);
  u32 rem, ms;
  merged_heads(E, lane, rem, ms);
  export_mask(out + i * COGL_MASK_SIZE, E.Splay, spec_of(E.Splay), rem, ms, lane);
}

/* selected_action_masks (merged = 0) or the merged legal mask (merged = 1) of every env in packed form */
__global__ void k_packed_masks(Buffers B, u64 n, u8 *out, int merged) {
  const int lane = threadIdx.x & 31;
  const u64 i = env_index(blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5));
  if (i >= n) return;
  Env E;
  env_load(E, B.hidden + i * COG_HID_WORDS, lane);
  u32 rem = E.Srem, ms = E.Sms;
  if (merged) merged_heads(E, lane, rem, ms);
  if (lane == 0)
    *reinterpret_cast<uint4 *>(out + i * COGL_PACKED_MASK_SIZE) = make_uint4(E.Splay, rem, ms, spec_of(E.Splay));
}

/* one compact record per env for its current state; merged = which mask heads */
__global__ void k_compact_obs(Buffers B, Params prm, u64 n, u8 *out, int merged) {
  const int lane = threadIdx.x & 31;
  const u64 i = env_index(blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5));
  if (i >= n) return;
  Env E;
  env_load(E, B.hidden + i * COG_HID_WORDS, lane);
  u32 rem = E.Srem, ms = E.Sms;
  if (merged) merged_heads(E, lane, rem, ms);
  u8 *rec = out + i * COGL_CR_SIZE;
  reinterpret_cast<uint2 *>(rec)[lane] = make_uint2(0u, 0u); /* caller memory: padding included */
  __syncwarp();
  compact_store_pre(rec, E, prm, E.Splay, spec_of(E.Splay), rem, ms, 0u, 0u, 0u, lane);
}

/* vec_action_sampler<N>::sample include/vec_sampler.h:14-21 on ActionMask[n] records.
 * The 92 payload bytes of a record are read ONCE, as 23 words by 23 lanes (three 32-byte
 * sectors — `masks` may be mapped pinned host memory, then these are the bytes that cross
 * PCIe); lane k then fetches bool k of each head from the lane that holds its word, and a
 * ballot turns the head into a bit set.  actions2 (optional): a second destination for the
 * ActionData bytes, e.g. the caller's mapped host array. */
/* PACKED: `masks` holds COGL_PACKED_MASK_SIZE bytes per env (the heads as bit sets, see cog_b200.h) and
 * actions2 gets COGL_PACKED_ACTION_SIZE bytes per env. */
template <bool PACKED>
__global__ void k_sample(const u8 *__restrict__ masks, u32 *srng, u8 *actions, u64 n, u8 *actions2) {
  const int lane = threadIdx.x & 31;
#ifndef COG_NO_UNIFORM_PROOF
  const u64 i = env_index(blockIdx.x * WARPS_PER_BLOCK + bcast(threadIdx.x >> 5, 0));
#else
  const u64 i = env_index(blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5));
#endif
  if (i >= n) return;
#ifndef COG_NO_UNIFORM_PROOF
  u32 sr = bcast(srng[i], 0);
#else
  u32 sr = srng[i];
#endif
  u32 act, a_shop;
  if (PACKED) {
    const uint4 pm = *reinterpret_cast<const uint4 *>(masks + i * COGL_PACKED_MASK_SIZE);
    /* (a caller's array: keep the heads inside their ranges, like the ballots of the bool form do) */
    const u32 play = bcast(pm.x, 0) & 0x3fffffu, rem = bcast(pm.y, 0) & 0x3fffffu, ms = bcast(pm.z, 0) & 0x3ffffffu,
              spec = bcast(pm.w, 0) & 0x3fffffu;
    sample_action(sr, play, spec, rem, ms, act, a_shop, lane);
    __syncwarp();
    if (lane == 0) {
      srng[i] = sr;
      *reinterpret_cast<uint2 *>(actions + i * COGL_ACTION_SIZE) = make_uint2(act, a_shop);
      if (actions2) *reinterpret_cast<uint2 *>(actions2 + i * COGL_PACKED_ACTION_SIZE) = make_uint2(act, a_shop);
    }
    return;
  }
  const u32 *m = reinterpret_cast<const u32 *>(masks + i * COGL_MASK_SIZE);
  const u32 w = lane < 23 ? m[lane] : 0u;
  /* bool (base + lane) of the record = byte ((base + lane) & 3) of word (base + lane) >> 2 */
#define COG_HEAD(base, count) \
  __ballot_sync(COG_FULL, ((__shfl_sync(COG_FULL, w, ((base) + lane) >> 2) >> (8 * (((base) + lane) & 3))) & 0xffu) && \
                              lane < (count))
  const u32 play = COG_HEAD(COGL_AM_PLAY, 22);
  const u32 spec = COG_HEAD(COGL_AM_SPECIAL, 22);
  const u32 rem = COG_HEAD(COGL_AM_REMOVE, 22);
  const u32 move = COG_HEAD(COGL_AM_MOVE, 7);
  const u32 shop = COG_HEAD(COGL_AM_SHOP, 19);
#undef COG_HEAD
  sample_action(sr, play, spec, rem, move | (shop << 7), act, a_shop, lane);
  __syncwarp(); /* every lane has read srng[i] before lane 0 overwrites it */
  if (lane == 0) {
    srng[i] = sr;
    *reinterpret_cast<uint2 *>(actions + i * COGL_ACTION_SIZE) = make_uint2(act, a_shop);
    if (actions2) *reinterpret_cast<uint2 *>(actions2 + i * COGL_ACTION_SIZE) = make_uint2(act, a_shop);
  }
}

/* vec_action_sampler<N>(seed) include/vec_sampler.h:9-13: set_seed(seed + i) on a
 * 64-bit sum (include/sampler.h:12) */
__global__ void k_sampler_seed(u32 *srng, u64 n, u64 seed_base) {
  const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) srng[i] = minstd_seed(seed_base + i);
}

}  // namespace cog
#endif
