/* cog_kernels.cuh — the __global__ entry points (one warp per environment).
 *
 *   k_init        state of a default-constructed vec_cog_env<N>
 *   k_reset       vec_cog_env<N>::reset(...)  — board generation, decks, shop, masks, map encode
 *   k_step<false> vec_cog_env<N>::step(actions)  — single step, matches the Python API
 *   k_step<true>  persistent fused K x {sample(selected masks); step} with no host round trip
 *   k_sample      vec_action_sampler<N>::sample(masks) on an arbitrary mask array
 */
#ifndef COG_KERNELS_CUH
#define COG_KERNELS_CUH
#include "cog_device.cuh"

namespace cog {

enum { WARPS_PER_BLOCK = 4, BLOCK_THREADS = WARPS_PER_BLOCK * 32 };

extern __shared__ uint4 cog_smem[];

__device__ __forceinline__ BoardScratch &warp_scratch() {
  return reinterpret_cast<BoardScratch *>(cog_smem)[threadIdx.x >> 5];
}

/* vec_cog_env<N>() include/vec_environment.h:23-31 -> cog_env::init src/environment.cpp:25-40:
 * Shop() puts 6 cards in the market and 3 copies in stock (src/cards.cpp:80-92), every
 * ActionMask is reset (include/api.h:101-103); everything else is zero. */
__global__ void k_init(Buffers B, u64 n) {
  const int lane = threadIdx.x & 31;
  const u64 i = (u64)blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
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
  env_store(E, B.hidden + i * COG_HID_WORDS, lane);
  export_dynamic(E, B, i, 0xfu, lane);
}

/* vec_cog_env<N>::reset include/vec_environment.h:33-44.  reseed: env i restarts its RNG
 * from uint32(seed_base + i) (cog_env::reset(seed, ...) src/environment.cpp:66-77). */
__global__ void k_reset(Buffers B, Params prm, u64 n, u64 seed_base, int reseed) {
  const int lane = threadIdx.x & 31;
  const u64 i = (u64)blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
  if (i >= n) return;
  Env E;
  env_load(E, B.hidden + i * COG_HID_WORDS, lane);
  if (reseed) E.rng = minstd_seed((u64)(u32)(seed_base + i));
  env_reset(E, prm, B, i, warp_scratch(), lane);
  env_store(E, B.hidden + i * COG_HID_WORDS, lane);
  export_dynamic(E, B, i, 0xfu, lane);
}

/* FUSED = false: one step per env from `actions` (ActionData[n], 64-byte records).
 * FUSED = true : persistent; each warp walks envs with a grid stride, keeps the env in
 *                registers for k_steps x {sample; step}, then writes it back once.
 *                `srng` is the sampler's RNG array, `actions` receives the last action. */
#ifndef COG_MIN_BLOCKS
#define COG_MIN_BLOCKS 8
#endif
template <bool FUSED>
__global__ void __launch_bounds__(BLOCK_THREADS, FUSED ? COG_MIN_BLOCKS : 6) k_step(Buffers B, Params prm, u64 n, const u8 *__restrict__ actions_in, u32 *srng,
                       u8 *actions_out, u64 k_steps) {
  const int lane = threadIdx.x & 31;
  const u64 warp0 = (u64)blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
  const u64 stride = (u64)gridDim.x * WARPS_PER_BLOCK;
  BoardScratch &sc = warp_scratch();
  u64 steps_done = 0;
  for (u64 i = warp0; i < n; i += stride) {
    Env E;
    env_load(E, B.hidden + i * COG_HID_WORDS, lane);
    u32 status = 0;
    if (FUSED) {
      u32 sr = srng[i];
      u32 act = 0, a_shop = 0, touched = 0;
      bool obs_pending = false;
      for (u32 k = 0; k < (u32)k_steps; ++k) {
        if (sample_action(sr, E.Splay, spec_of(E.Splay), E.Srem, E.Sms, act, a_shop, lane))
          status = env_step<true, true>(E, act, 0, prm, B, i, sc, obs_pending, lane);
        else
          status = env_step<true, false>(E, act, a_shop, prm, B, i, sc, obs_pending, lane);
        touched |= status >> 8;
      }
      if (obs_pending) env_update_observation(E, (int)E.agent, 0, E.cpc, lane);
      __syncwarp(); /* every lane has read srng[i] before lane 0 overwrites it */
      if (lane == 0) srng[i] = sr;
      if (k_steps) {
        if (lane < 4) actions_out[i * COGL_ACTION_SIZE + lane] = (u8)(act >> (8 * lane));
        if (lane == 4) actions_out[i * COGL_ACTION_SIZE + 4] = (u8)a_shop;
      }
      status = (status & 1u) | (touched << 8);
      steps_done += k_steps;
    } else {
      const u8 *a = actions_in + i * COGL_ACTION_SIZE;
      const u32 mine = lane < 5 ? (u32)a[lane] : 0u;
      const u32 act = bcast(mine, 0) | (bcast(mine, 1) << 8) | (bcast(mine, 2) << 16) | (bcast(mine, 3) << 24);
      bool obs_pending = false;
      status = env_step<false, false>(E, act, bcast(mine, 4), prm, B, i, sc, obs_pending, lane);
      steps_done += 1;
    }
    env_store(E, B.hidden + i * COG_HID_WORDS, lane);
    if (!FUSED || k_steps) {
      export_dynamic(E, B, i, status >> 8, lane);
      if (lane == 0) B.dones[i] = (u8)(status & 1u);
    }
  }
  if (lane == 0 && steps_done) atomicAdd(&B.stats[0], (unsigned long long)steps_done);
}

/* vec_action_sampler<N>::sample include/vec_sampler.h:14-21 on ActionMask[n] records:
 * lane k reads bool k of each head, a ballot turns the head into a bit set. */
__global__ void k_sample(const u8 *__restrict__ masks, u32 *srng, u8 *actions, u64 n) {
  const int lane = threadIdx.x & 31;
  const u64 i = (u64)blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
  if (i >= n) return;
  const u8 *m = masks + i * COGL_MASK_SIZE;
  const bool in22 = lane < 22;
  const u32 play = __ballot_sync(COG_FULL, in22 && m[COGL_AM_PLAY + lane]);
  const u32 spec = __ballot_sync(COG_FULL, in22 && m[COGL_AM_SPECIAL + lane]);
  const u32 rem = __ballot_sync(COG_FULL, in22 && m[COGL_AM_REMOVE + lane]);
  const u32 move = __ballot_sync(COG_FULL, lane < 7 && m[COGL_AM_MOVE + lane]);
  const u32 shop = __ballot_sync(COG_FULL, lane < 19 && m[COGL_AM_SHOP + lane]);
  u32 sr = srng[i], act, a_shop;
  sample_action(sr, play, spec, rem, move | (shop << 7), act, a_shop, lane);
  __syncwarp(); /* every lane has read srng[i] before lane 0 overwrites it */
  if (lane == 0) srng[i] = sr;
  if (lane < 4) actions[i * COGL_ACTION_SIZE + lane] = (u8)(act >> (8 * lane));
  if (lane == 4) actions[i * COGL_ACTION_SIZE + 4] = (u8)a_shop;
}

/* vec_action_sampler<N>(seed) include/vec_sampler.h:9-13: set_seed(seed + i) on a
 * 64-bit sum (include/sampler.h:12) */
__global__ void k_sampler_seed(u32 *srng, u64 n, u64 seed_base) {
  const u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) srng[i] = minstd_seed(seed_base + i);
}

}  // namespace cog
#endif
