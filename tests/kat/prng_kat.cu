/* tests/kat/prng_kat.cu — TEST INFRASTRUCTURE: exhaustive known-answer tests of the device PRNG.
 *
 * Includes the PRODUCT header (gym_eldorado_b200/csrc/cog_device.cuh) and checks its division-free
 * restatement of libstdc++'s minstd_rand0 + uniform_int_distribution ("two divisions" branch,
 * /usr/include/c++/13/bits/uniform_int_dist.h:332-340; draw sites of the reference:
 * include/sampler.h:25,37,49,61,74, src/cards.cpp:190-191, src/player.cpp:92-93,115-116,
 * src/map.cpp:298-301,705-734) against the plain 64-bit arithmetic it replaces:
 *
 *   [0] minstd_next(s) != s * 16807 mod (2^31 - 1)                for EVERY state s in [1, 2^31 - 2]
 *   [1] UD_TAB[n] != {2147483645 / n, n * scaling}                for n = 1..32
 *   [2] ud_quotient(r, UD_TAB[n]) != r / scaling                  for n = 1..32 and EVERY r in [0, past)
 *   [3] (r >= UD_TAB[n].past) != (r >= n * (2147483645 / n))      for n = 1..32 and EVERY r in [0, 2^31 - 3]
 *   [4] minstd_skip4_unit_draws(s) != four sample_head(s, 1) calls for EVERY state s (true branch)
 *   [5] number of states where it returns false (must be exactly the 4 guard states)
 *   [6] a false return whose four step-by-step draws saw NO rejection (must be 0)
 *   [7] uniform_draw(s, n) != the two-division loop               for 2^22 states x n = 1..64
 *   [8] sample_action fast path != five sample_head calls          for 2^22 (state, play mask) pairs
 *   [9] nth_set_bit_warp(v, j) != nth_set_bit(v, j)                for 2^16 values x every j
 * Every counter must come back 0 (except [5] == 4).  Built by __graft_entry__.build() into
 * tests/kat/libcog_kat.so; only tests/ load it.
 */
#ifdef COG_EMU /* tests/kat/kat_emu.cpp: the same kernels on sub-ranges under the warp emulator (CPU) */
#include "cuda_emu.h"
#else
#include <cuda_runtime.h>
#endif
#include <stdint.h>

#include "cog_device.cuh"

using namespace cog;

namespace {
enum { KAT_COUNTERS = 16 };
constexpr u64 M31 = 2147483647ull;

__device__ __forceinline__ void bump(unsigned long long *out, int k, u32 bad) {
  const u32 total = __reduce_add_sync(COG_FULL, bad);
  if (total && (threadIdx.x & 31) == 0) atomicAdd(&out[k], (unsigned long long)total);
}

/* one thread per engine state / engine output, grid-stride */
__global__ void k_kat_exhaustive(unsigned long long *out, u64 v0, u64 v1) {
  const u64 stride = (u64)gridDim.x * blockDim.x;
  u32 bad0 = 0, bad2 = 0, bad3 = 0, bad4 = 0, n5 = 0, bad6 = 0;
  for (u64 v = v0 + (u64)blockIdx.x * blockDim.x + threadIdx.x; v < v1; v += stride) {
    if (v >= 1 && v <= M31 - 1) { /* engine states 1 .. 2^31 - 2 */
      u32 s = (u32)v;
      const u32 got = minstd_next(s);
      const u32 want = (u32)((v * 16807ull) % M31);
      bad0 += (got != want) | (s != want);
      /* four unit draws: the collapsed jump against the step-by-step path */
      u32 a = (u32)v, b = (u32)v;
      const bool fast = minstd_skip4_unit_draws(a);
      bool rejected = false;
      {
        u64 x = v;
        for (int k = 0; k < 4; ++k) {
          x = (x * 16807ull) % M31;
          rejected |= x == 2147483646ull;
        }
      }
      for (int k = 0; k < 4; ++k) sample_head(b, 1u);
      if (fast) {
        bad4 += (a != b) | rejected;
      } else {
        n5 += 1;
        bad6 += (a != (u32)v) | !rejected;
      }
    }
    if (v <= M31 - 3) { /* r = g() - 1 in [0, 2^31 - 3] */
      const u32 r = (u32)v;
#pragma unroll 1
      for (u32 n = 1; n <= 32; ++n) {
        const ud_entry e = UD_TAB[n];
        const u32 scaling = 2147483645u / n, past = n * scaling;
        bad3 += (r >= e.past) != (r >= past);
        if (r < past) bad2 += ud_quotient(r, e) != r / scaling;
      }
    }
  }
  bump(out, 0, bad0);
  bump(out, 2, bad2);
  bump(out, 3, bad3);
  bump(out, 4, bad4);
  bump(out, 5, n5);
  bump(out, 6, bad6);
}

__global__ void k_kat_table(unsigned long long *out) {
  const u32 n = threadIdx.x + 1;
  u32 bad = 0;
  if (n <= 32) {
    const ud_entry e = UD_TAB[n];
    bad = (e.scaling != 2147483645u / n) | (e.past != n * (2147483645u / n));
  }
  bump(out, 1, bad);
}

__device__ __forceinline__ u32 ref_uniform(u32 &s, u32 n) { /* uniform_int_dist.h:332-340 verbatim */
  const u32 scaling = 2147483645u / n, past = n * scaling;
  u32 r;
  do {
    s = (u32)(((u64)s * 16807ull) % M31);
    r = s - 1u;
  } while (r >= past);
  return r / scaling;
}
__device__ __forceinline__ u32 ref_head(u32 &s, u32 valid) { /* include/sampler.h:14-79, one head */
  const u32 j = ref_uniform(s, (u32)__popc(valid));
  u32 v = valid;
  for (u32 t = 0; t < j; ++t) v &= v - 1u;
  return (u32)__ffs((int)v) - 1u;
}
__device__ __forceinline__ u32 mix(u64 x) { /* splitmix64 finaliser: test inputs only */
  x += 0x9e3779b97f4a7c15ull;
  x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
  x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
  return (u32)((x ^ (x >> 31)) >> 16);
}

/* one WARP per test case (the sampler's fast path is a warp collective) */
__global__ void k_kat_sampled(unsigned long long *out, u32 cases) {
  const int lane = threadIdx.x & 31;
  const u32 warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const u32 n_warps = (gridDim.x * blockDim.x) >> 5;
  u32 bad7 = 0, bad8 = 0, bad9 = 0;
  for (u32 c = warp; c < cases; c += n_warps) {
    u32 s0 = mix(c) % (u32)(M31 - 1) + 1u;
    if ((c & 1023u) < 4u) { /* the four guard states of the skip-4 jump, each many times */
      const u32 guard[4] = {739806647u, 671875339u, 939811632u, 1785685951u};
      s0 = guard[c & 3u];
    }
    /* [7] the whole draw, both branches of uniform_draw (n <= 32 table, n > 32 divisions) */
    for (u32 n = 1 + lane; n <= 64; n += 32) {
      u32 a = s0, b = s0;
      const u32 x = uniform_draw(a, n), y = ref_uniform(b, n);
      bad7 += (x != y) | (a != b);
    }
    /* [8] sample_action against five reference heads */
    const u32 m = mix((u64)c << 32 | 7u);
    u32 play = (m & 0x3ffffeu) | 1u;
    if ((c & 7u) == 0) play = 1u;             /* nothing to play */
    if ((c & 7u) == 1) play = 1u | (2u << (m % 21u)); /* one card */
    u32 spec = 1u, rem = 1u, ms = 0x81u;
    if ((c & 15u) == 3u) { /* general path: every head has a choice */
      const u32 m2 = mix((u64)c << 32 | 9u);
      spec = spec_of(play);
      rem = (m2 & 0x3ffffeu) | 1u;
      ms = ((m2 >> 3) & 0x7eu) | 1u | ((((m2 >> 9) & 0x7fffeu) | 1u) << 7);
    }
    u32 a = s0, b = s0, act = 0, a_shop = 0;
    sample_action(a, play, spec, rem, ms, act, a_shop, lane);
    const u32 w0 = ref_head(b, play), w1 = ref_head(b, spec), w2 = ref_head(b, rem), w3 = ref_head(b, ms & 0x7fu),
              w4 = ref_head(b, ms >> 7);
    if (lane == 0) bad8 += (act != (w0 | (w1 << 8) | (w2 << 16) | (w3 << 24))) | (a_shop != w4) | (a != b);
    /* [9] n-th set bit, warp version against the loop */
    if (c < (1u << 16)) {
      const u32 v = (mix((u64)c << 32 | 11u) & 0x3ffffffu) | 1u;
      const u32 pc = (u32)__popc(v);
      for (u32 j = 0; j < pc; ++j) {
        const u32 x = nth_set_bit_warp(v, j, lane), y = nth_set_bit(v, j);
        if (lane == 0) bad9 += x != y;
      }
    }
  }
  bump(out, 7, bad7);
  bump(out, 8, bad8);
  bump(out, 9, bad9);
}
}  // namespace

#ifdef COG_EMU
/* CPU check of the checker itself: the engine states / outputs in [v0, v1) and `cases` sampled cases */
extern "C" __attribute__((visibility("default"))) int cog_kat_prng_emu(unsigned long long *out, int n_out,
                                                                       unsigned long long v0, unsigned long long v1,
                                                                       unsigned cases) {
  if (!out || n_out < KAT_COUNTERS) return -1;
  for (int k = 0; k < KAT_COUNTERS; ++k) out[k] = 0;
  cuda_emu::launch(1, 32, 0, [&]() { k_kat_table(out); });
  cuda_emu::launch(2, 64, 0, [&]() { k_kat_exhaustive(out, v0, v1); });
  cuda_emu::launch(2, 64, 0, [&]() { k_kat_sampled(out, cases); });
  return 0;
}
#else
extern "C" __attribute__((visibility("default"))) int cog_kat_prng(unsigned long long *out, int n_out) {
  if (!out || n_out < KAT_COUNTERS) return -1;
  unsigned long long *d = nullptr;
  if (cudaMalloc(&d, KAT_COUNTERS * sizeof(*d)) != cudaSuccess) return -4;
  cudaMemset(d, 0, KAT_COUNTERS * sizeof(*d));
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  k_kat_table<<<1, 32>>>(d);
  k_kat_exhaustive<<<sms * 8, 256>>>(d, 0ull, M31);
  k_kat_sampled<<<sms * 8, 256>>>(d, 1u << 22);
  const cudaError_t e = cudaDeviceSynchronize();
  cudaMemcpy(out, d, KAT_COUNTERS * sizeof(*d), cudaMemcpyDeviceToHost);
  cudaFree(d);
  return e == cudaSuccess ? 0 : -4;
}
#endif
