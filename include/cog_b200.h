/* cog_b200.h — C ABI of the B200-native City of Gold vectorised hot path.
 *
 * This is the drop-in boundary: every entry point below replaces one member of
 * the reference's vectorised containers (aapo-kossi/gym-eldorado), which its
 * pybind module `city_of_gold` binds (reference include/pybind/vectorized.h).
 * Plain pointers and sizes only; no C++ or torch types cross this line.
 *
 * One `cog_vec` handle = one GPU's slice [env_index_offset, +n_envs) of a job.
 * Seeds derive from the GLOBAL env index, so any sharding of a job over GPUs
 * gives identical per-env trajectories.  The library owns all device memory;
 * exported buffers keep the reference's record layouts (include/api.h:67-161)
 * at device addresses that never change until cog_vec_destroy().
 *
 * Calls on one handle must come from one host thread at a time; work is
 * enqueued on the handle's CUDA stream and is asynchronous unless stated.
 * Return value: 0 on success, negative cog_status otherwise; the message for
 * the calling thread's last failure is cog_last_error().
 */
#ifndef COG_B200_H
#define COG_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define COG_API __declspec(dllexport)
#else
#define COG_API __attribute__((visibility("default")))
#endif

/* exported record sizes (reference include/api.h) */
#define COG_OBS_SIZE 17216
#define COG_MASK_SIZE 128
#define COG_MASK_PAYLOAD 92
#define COG_ACTION_SIZE 64
#define COG_ACTION_PAYLOAD 5
#define COG_INFO_SIZE 192
#define COG_MAX_PLAYERS 4

typedef enum cog_status {
  COG_OK = 0,
  COG_ERR_INVALID = -1,      /* bad argument / handle */
  COG_ERR_MAP_FAILURE = -2,  /* generate_map_failure (reference include/map.h:13-15) */
  COG_ERR_MAP_BOUNDS = -3,   /* board does not fit the 48x48 observation (UB in the reference); rarely (seen at 7 pieces) also:
                              * more than 32 pieces on the map during one generation, or an intermediate board
                              * wider than the observation ("known limit"; a library built with
                              * -DCOG_WIDE_BOARDS generates those boards exactly) */
  COG_ERR_CUDA = -4,         /* a CUDA runtime call failed */
  COG_ERR_NOT_RESET = -5     /* step before the first reset(seed, ...) */
} cog_status;

/* reference include/constants.h:11 */
typedef enum cog_difficulty { COG_EASY = 0, COG_MEDIUM = 1, COG_HARD = 2 } cog_difficulty;

/* which exported buffer (reference property in parentheses,
 * include/pybind/vectorized.h:70-102,200-214) */
typedef enum cog_buffer {
  COG_BUF_OBS = 0,             /* observations            ObsData[n]      17216 B */
  COG_BUF_SELECTED_MASKS = 1,  /* selected_action_masks   ActionMask[n]     128 B */
  COG_BUF_REWARDS = 2,         /* rewards                 float[n][4]        16 B */
  COG_BUF_DONES = 3,           /* dones                   bool[n]             1 B */
  COG_BUF_INFOS = 4,           /* infos                   Info[n]           192 B */
  COG_BUF_AGENT_SELECTION = 5, /* agent_selection         uint8[n]            1 B */
  COG_BUF_MERGED_MASKS = 6,    /* not in the reference: the merged legal mask of each env's current agent,
                                  ActionMask[n]; allocated on first use, refreshed by
                                  cog_vec_merged_masks(h, NULL) */
  COG_BUF_GRID = 7,            /* not in the reference: the board as ONE byte per hex, uint8[n][48*48], cell
                                  [row * 48 + col] describes the same hex as observations.shared.map[row][col]
                                  (Map::finalize src/map.cpp:389-405): bits 0-2 = requirement + 1 (0 = no hex;
                                  the map feature that is non-zero), bits 3-5 = n_required, bit 6 = is_end.
                                  Constant within an episode; rewritten by every (auto-)reset */
  COG_BUF_COMPACT = 8,         /* not in the reference: one 256-byte compact record per env (below);
                                  allocated on first use, refreshed by cog_vec_compact_obs(h, NULL, mode) */
  COG_BUF_ROLLOUT = 9          /* not in the reference: the rollout ring, [n][n_slots] compact records
                                  (count = n * n_slots); only after cog_vec_rollout_attach */
} cog_buffer;
#define COG_GRID_BYTES 2304
#define COG_COMPACT_SIZE 256

/* episode aggregates accumulated on the device since cog_vec_create / the last
 * cog_vec_stats_reset (the vector ranks all-reduce with NCCL) */
typedef enum cog_stat {
  COG_STAT_ENV_STEPS = 0,
  COG_STAT_EPISODES = 1,       /* finished episodes (= auto-resets) */
  COG_STAT_TRUNCATIONS = 2,    /* episodes ended by turn_counter >= max_steps without a winner */
  COG_STAT_TOTAL_LENGTH = 3,   /* sum of Info.total_length (turns) over finished episodes */
  COG_STAT_WINS_SEAT0 = 4,     /* +seat: episodes that seat won */
  COG_STAT_TRAVELLED_HEXES = 8,
  COG_STAT_CARDS_ADDED = 9,
  COG_STAT_RESETS = 10,        /* boards generated (explicit + auto) */
  COG_STAT_COUNT = 12
} cog_stat;

typedef struct cog_vec cog_vec;         /* vec_cog_env<N>        include/vec_environment.h:12-81 */
typedef struct cog_sampler cog_sampler; /* vec_action_sampler<N> include/vec_sampler.h:7-27 */

COG_API const char *cog_last_error(void);
COG_API int cog_device_count(int *count);

/* vec_cog_env<N>::vec_cog_env()  (include/vec_environment.h:23-31).
 * `stream` is a cudaStream_t; NULL = the device's default stream, which also orders the
 * handle's work against a sampler created with NULL and against torch's default stream. */
COG_API int cog_vec_create(uint64_t n_envs, uint64_t env_index_offset, int device,
                           void *stream, cog_vec **out);
COG_API int cog_vec_destroy(cog_vec *h);

/* vec_cog_env<N>::reset(seed, n_players, n_pieces, difficulty, max_steps, render)
 * (include/vec_environment.h:38-44 -> src/environment.cpp:42-77).  Env i is seeded
 * uint32(seed + env_index_offset + i).  Synchronous: returns COG_ERR_MAP_FAILURE
 * where the reference throws generate_map_failure.  `render` is accepted and
 * ignored (the reference's rendering is text stubs). */
COG_API int cog_vec_reset(cog_vec *h, uint32_t seed, uint8_t n_players, uint8_t n_pieces,
                          int difficulty, uint32_t max_steps, int render);
/* vec_cog_env<N>::reset()  (include/vec_environment.h:33-37): same parameters,
 * RNG streams continue. */
COG_API int cog_vec_reset_default(cog_vec *h);

/* vec_cog_env<N>::step(actions)  (include/vec_environment.h:46-61): one step per
 * env with auto-reset on done.  `d_actions` is a DEVICE pointer to n_envs
 * ActionData records (64 B each, 5 payload bytes). */
COG_API int cog_vec_step(cog_vec *h, const void *d_actions);
/* Auto-reset off = the semantics of the reference's SINGLE env (cog_env::step, src/environment.cpp:91-95,
 * 187-207, bound at src/pybind/single_env.cpp:13-32): the step that ends an episode writes rewards / Info
 * and sets done, but does not reset; the env keeps its terminal state and ignores further steps until
 * cog_vec_reset(_default).  On (the default) = vec_cog_env<N>::step_single (include/vec_environment.h:53-61).
 * The fused loop requires auto-reset. */
COG_API int cog_vec_set_autoreset(cog_vec *h, int on);
/* vec_action_sampler<N>::sample(envs.selected_action_masks) followed by vec_cog_env<N>::step(
 * sampler.get_actions()) — the two calls of the reference's README / pytest / asv loop
 * (include/vec_sampler.h:14-21, include/vec_environment.h:46-61) — as ONE kernel: the action of each env
 * is drawn from its own selected mask, written to the sampler's ActionData buffer and applied.
 * Bit-identical to cog_sampler_sample(s, <COG_BUF_SELECTED_MASKS>) + cog_vec_step(h, <sampler actions>),
 * without the second launch and the mask / action round trip through memory.  (Masks from anywhere
 * else — a policy's own array, the merged masks — go through cog_sampler_sample.) */
COG_API int cog_vec_step_sampled(cog_vec *h, cog_sampler *s);
/* same, but `h_actions` is a HOST pointer (what the reference's step(ndarray) gets).  Returns when the
 * array has been copied (the caller may rewrite it at once, like after the reference's synchronous
 * step()); the step itself stays asynchronous */
COG_API int cog_vec_step_host(cog_vec *h, const void *h_actions);

/* device address of an exported buffer; stable for the handle's lifetime */
COG_API int cog_vec_buffer(cog_vec *h, int which, void **d_ptr, size_t *itemsize,
                           size_t *count);
/* copy an exported buffer to host memory (synchronises the handle's stream) */
COG_API int cog_vec_read(cog_vec *h, int which, void *h_dst, size_t nbytes);
COG_API int cog_vec_num_envs(cog_vec *h, uint64_t *n);

/* vec_action_sampler<N>(seed)  (include/vec_sampler.h:9-13): sampler i is seeded
 * with the un-truncated 64-bit uint32(seed) + env_index_offset + i. */
COG_API int cog_sampler_create(uint64_t n_envs, uint64_t env_index_offset, uint32_t seed,
                               int device, void *stream, cog_sampler **out);
COG_API int cog_sampler_destroy(cog_sampler *s);
/* vec_action_sampler<N>::sample(masks)  (include/vec_sampler.h:14-21 ->
 * include/sampler.h:14-79).  `d_masks`: DEVICE pointer to n_envs ActionMask records. */
COG_API int cog_sampler_sample(cog_sampler *s, const void *d_masks);
COG_API int cog_sampler_sample_host(cog_sampler *s, const void *h_masks);
/* vec_action_sampler<N>::get_actions(): device address of ActionData[n_envs] */
COG_API int cog_sampler_actions(cog_sampler *s, void **d_ptr);
COG_API int cog_sampler_read_actions(cog_sampler *s, void *h_dst, size_t nbytes);

/* Host-buffer forms of sample() / step() for callers whose policy lives on the host (the
 * reference's Python loop hands numpy arrays to vec_sampler.sample and vec_cog_env.step,
 * include/pybind/vectorized.h:60-68,107-127, and reads the views afterwards).  The host buffers
 * are complete when the call returns.  Two implementations, same bytes:
 *  - staged: copy the inputs in, run, copy the results out, cut into two chunks (COG_IO_CHUNKS)
 *    on two internal streams so that H2D, kernel and D2H of different chunks overlap (pinned
 *    memory is needed for the overlap; pageable memory works, serialised);
 *  - zero copy (pinned, device-addressable arrays only — cudaHostAlloc / cudaHostRegister, e.g.
 *    torch's pin_memory()): the kernels read masks / actions from, and write actions / masks
 *    to, the caller's arrays over PCIe themselves; only the needed sectors cross the link.
 * For pinned arrays each handle times both on its first nine calls and keeps the faster one
 * (zero copy on one GPU; the staged pipeline when eight ranks share one host's memory).
 * COG_IO_ZEROCOPY=0 / 1 in the environment forces one.  The records must be aligned like the
 * reference's (ActionMask 4 bytes, ActionData 8 bytes — any array of whole records is).
 * Any *_out pointer of cog_vec_step_host_io may be NULL.
 *   h_masks / h_masks_out : ActionMask[n_envs]   h_actions / h_actions_out : ActionData[n_envs]
 *   h_rewards_out : float[n_envs][4]   h_dones_out : bool[n_envs]   h_agents_out : uint8[n_envs] */
COG_API int cog_sampler_sample_host_io(cog_sampler *s, const void *h_masks, void *h_actions_out);
COG_API int cog_vec_step_host_io(cog_vec *h, const void *h_actions, void *h_masks_out, void *h_rewards_out,
                                 void *h_dones_out, void *h_agents_out);

/* Packed host I/O (not in the reference; opt-in — the reference-layout arrays above stay the default).
 * Three quarters of the bytes of the host loop above are bool-per-byte masks (92 payload bytes in a
 * 128-byte record) and 59 of the 64 bytes of an ActionData record are padding.  The packed forms carry
 * the same information:
 *   packed mask   COG_PACKED_MASK_SIZE = 16 bytes: u32 play (bit k = play[k]), u32 remove,
 *                 u32 move (bits 0-6) | get_from_shop << 7 (bits 7-25), u32 play_special
 *                 — the mask words of the compact record, in the same order
 *   packed action COG_PACKED_ACTION_SIZE = 8 bytes: u8 play, play_special, remove, move, get_from_shop, 3 x 0
 *                 — the first eight bytes of an ActionData record
 * so one sample + step moves 16 + 8 + 8 + 16 + 18 = 66 bytes per env over PCIe instead of 246-402.
 * Both calls stage through device memory by DMA and return when the host arrays are complete; any
 * *_out pointer of the step call may be NULL.  cog_vec_packed_masks writes the current packed
 * selected (COG_MASKS_SELECTED) or merged (COG_MASKS_MERGED) masks to device (dst_is_host = 0) or host
 * memory (dst_is_host = 1, synchronous) — what a loop needs after reset(). */
#define COG_PACKED_MASK_SIZE 16
#define COG_PACKED_ACTION_SIZE 8
COG_API int cog_sampler_sample_packed_host_io(cog_sampler *s, const void *h_pmasks, void *h_pactions_out);
COG_API int cog_vec_step_packed_host_io(cog_vec *h, const void *h_pactions, void *h_pmasks_out,
                                        void *h_rewards_out, void *h_dones_out, void *h_agents_out);
COG_API int cog_vec_packed_masks(cog_vec *h, void *dst, int dst_is_host, int mask_mode);

/* cog_sampler_sample_packed_noalloc — NO COUNTERPART IN THE REFERENCE; added
 * for callers that must not allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).
 *
 * The same operation, argument meaning and results as cog_sampler_sample_packed_host_io above (bit-identical
 * packed actions, the same advance of the sampler's RNG streams, the sampler's ActionData buffer updated), but
 * the device staging of the packed masks and packed actions is NOT allocated on first use: it lives in a
 * file-scope `static __device__ __align__(256)` array inside the library
 *   COG_NOALLOC_MAX_ENVS * (COG_PACKED_MASK_SIZE + COG_PACKED_ACTION_SIZE) = 65 536 * 24 bytes = 1.5 MiB
 * that every GPU which loads the library holds until the process exits.  The call performs no cudaMalloc,
 * cudaFree, event / stream creation or any other allocating runtime or library routine (the array's device
 * address is resolved once per device with cudaGetSymbolAddress — that device current, return code checked —
 * on the first call made for that device, and cached).
 *   - a sampler of more than COG_NOALLOC_MAX_ENVS envs is rejected with COG_ERR_INVALID and a message that
 *     names the limit (cog_last_error); nothing is launched.
 *   - ALL calls on one device share the one array, whichever sampler or stream they use: two calls whose GPU
 *     work overlaps (different streams, different host threads) RACE on it.  The call returns after its stream
 *     has drained, so calls issued one after the other from one thread are safe.
 *   - the array keeps the previous call's contents; nothing in it has to start at zero (the first n_envs packed
 *     masks are overwritten by the host-to-device copy and all 8 bytes of each of the first n_envs packed
 *     actions by the kernel before they are copied out), bytes beyond n_envs are never read. */
#define COG_NOALLOC_MAX_ENVS 65536
COG_API int cog_sampler_sample_packed_noalloc(cog_sampler *s, const void *h_pmasks, void *h_pactions_out);

/* Streams.  Every handle enqueues on the stream it was created with; handles that exchange device buffers
 * (the sampler's actions into cog_vec_step, the env's masks into cog_sampler_sample) must share one stream
 * — the pybind module creates all of them on the legacy default stream — or be ordered explicitly:
 * cog_*_wait_stream makes the handle's later work wait for what `producer` has enqueued so far (a policy
 * network writing the action array on its own stream), cog_*_release_to_stream makes `consumer` wait for
 * what the handle has enqueued so far (a consumer of the exported buffers: the DLPack `stream` argument).
 * Events only, no host synchronisation.  cog_vec_run_fused(_ex) / cog_vec_step_sampled order the sampler's
 * stream before the env's themselves. */
COG_API int cog_vec_wait_stream(cog_vec *h, void *producer);
COG_API int cog_vec_release_to_stream(cog_vec *h, void *consumer);
COG_API int cog_sampler_wait_stream(cog_sampler *s, void *producer);
COG_API int cog_sampler_release_to_stream(cog_sampler *s, void *consumer);
COG_API int cog_vec_stream(cog_vec *h, void **stream);
COG_API int cog_sampler_stream(cog_sampler *s, void **stream);

/* Which implementation the *_host_io entry points use for PINNED arrays on these handles (either
 * may be NULL): 1 zero copy, 0 staged pipeline, -1 still measuring (fewer than nine calls so far).
 * Pageable arrays always take the staged pipeline. */
COG_API int cog_io_path(cog_vec *h, cog_sampler *s, int *vec_path, int *sampler_path);

/* ThreadedRunner<N> replacement (include/runner.h:21-134): k_steps x
 * { sample(selected_action_masks); step(actions) } in ONE persistent kernel with
 * no host round trip; bit-identical to k_steps single sample+step calls.
 * The sampler's action buffer holds the last sampled actions afterwards. */
/* What the exported buffers show afterwards is the state after the LAST of the k steps: `dones` holds the
 * done flags of that step only (an env whose episode ended earlier in the launch has been auto-reset and
 * stepped on, like in the reference's loop, where the caller would have seen dones[i] == true on that step
 * and here does not), `rewards` / `infos` keep the values of each env's most recent episode end, and
 * cog_vec_stats counts every episode.  A caller that needs per-step dones / rewards attaches a rollout ring
 * (cog_vec_rollout_attach: one record per env-step, done and rewards included) or steps with k_steps = 1. */
COG_API int cog_vec_run_fused(cog_vec *h, cog_sampler *s, uint64_t k_steps);

/* Which mask the fused loop samples from.  The reference keeps the legal actions of the
 * current agent in TWO records: play / play_special in
 * selected_action_masks, move / get_from_shop / remove in
 * observations.player_data[agent].action_mask (include/api.h:104-131,
 * src/environment.cpp:252-279).  COG_MASKS_SELECTED feeds the sampler
 * selected_action_masks, exactly like the reference's README / pytest / asv loop (only `play`
 * ever has a choice there).  COG_MASKS_MERGED feeds it the merged legal mask (play,
 * play_special from S; move, get_from_shop from P[agent]; remove from P[agent] AND card
 * still in hand): agents move, buy, remove and win. */
enum { COG_MASKS_SELECTED = 0, COG_MASKS_MERGED = 1 };
COG_API int cog_vec_run_fused_ex(cog_vec *h, cog_sampler *s, uint64_t k_steps, int mask_mode);
/* the merged legal mask of every env's current agent, written as n_envs ActionMask records
 * to DEVICE memory `d_out` (caller-owned, n_envs * COG_MASK_SIZE bytes), or, with d_out ==
 * NULL, to the handle's own COG_BUF_MERGED_MASKS buffer */
COG_API int cog_vec_merged_masks(cog_vec *h, void *d_out);

/* Compact observation.  The reference's ObsData is 17216 B per env, 94 % of it
 * the 48x48x7 map that never changes within an episode (and whose "occupying player" feature
 * is never written).  The compact record carries what the CURRENT agent acts
 * on in 256 bytes; together with COG_BUF_GRID it restates every field of ObsData the agent
 * can use, plus the players' positions (byte offsets):
 *     0  u8 agent_selection, u8 phase, u16 current_resources[3] * 2
 *     8  ActionData taken from this state: u8 play, play_special, remove, move, get_from_shop,
 *        3 zero bytes (all zero in cog_vec_compact_obs)
 *    16  u32 mask.play (bit k = play[k]), u32 mask.remove,
 *        u32 mask.move (bits 0-6) | mask.get_from_shop << 7 (bits 7-25), u32 mask.play_special
 *    32  4 x { u8 row, u8 col }: the cell of COG_BUF_GRID player p stands on (0, 0 for p >= n_players)
 *    40  u32 turn_counter, u32 step index since cog_vec_rollout_attach (low 32 bits)
 *    48  u8 done, u8 filled (1 in rollout records: action, done, rewards, step index are valid),
 *        14 zero bytes
 *    64  float rewards[4] (the env's rewards when that step ended the episode, else 0)
 *    80  21 x { u8 draw, hand, active, played, discard of card type c in the current agent's
 *        DeckObs (include/api.h:67-81), u8 shop stock of slot c (c < 18, else 0), 2 zero bytes }
 *   248  8 zero bytes
 * The mask words are selected_action_masks (COG_MASKS_SELECTED) or the merged legal mask
 * (COG_MASKS_MERGED).  Writes n_envs records to DEVICE memory d_out, or to COG_BUF_COMPACT
 * when d_out == NULL. */
COG_API int cog_vec_compact_obs(cog_vec *h, void *d_out, int mask_mode);

/* Rollout writer (the reference leaves this to the user: "copy the views into
 * your own buffers", docs/source/index.rst:20-26).  After cog_vec_rollout_attach(h, T) every
 * cog_vec_run_fused(_ex) launch appends one compact record per env-step to a device ring of
 * T slots per env, env-major: step t (counted per env since attach) of env i is at
 *   ring + (i * T + t % T) * COG_COMPACT_SIZE
 * holding the state BEFORE the step, the mask heads the action was sampled from, the action,
 * and done / rewards AFTER it — what a PPO-style trainer consumes; each env's trajectory is
 * contiguous (returns / advantages scan along it), and the warp that owns the env streams it
 * out as consecutive 256-byte records.  The ring costs n_envs * T * 256 bytes; n_slots == 0
 * detaches and frees it.  cog_vec_rollout_info returns the ring, its size and the number of
 * steps recorded so far (the next record goes to slot steps % T). */
COG_API int cog_vec_rollout_attach(cog_vec *h, uint32_t n_slots);
COG_API int cog_vec_rollout_info(cog_vec *h, void **d_ring, uint32_t *n_slots, uint64_t *steps_recorded);

/* episode aggregates (COG_STAT_COUNT uint64 values); synchronises the stream */
COG_API int cog_vec_stats(cog_vec *h, uint64_t *out, size_t n);
COG_API int cog_vec_stats_reset(cog_vec *h);

/* wait for everything enqueued on the handle; reports deferred device-side
 * failures (map generation during an auto-reset) */
COG_API int cog_sync(cog_vec *h);
/* number of kernels this library has launched on behalf of this process */
COG_API uint64_t cog_kernel_launches(void);
/* elapsed device time (ms) of the most recent cog_vec_run_fused / cog_vec_run_fused_ex launch (both kernels)
 * on this handle, measured with CUDA events on the handle's stream; synchronises.  Single steps
 * (cog_vec_step, cog_vec_step_sampled) are timed only when the process runs with COG_STEP_EVENTS=1: a timing
 * event between two kernels of a stream is a serialisation point that costs the per-step loop 12 %, and the
 * per-step kernels use programmatic dependent launch (COG_PDL=0 turns that off). */
COG_API int cog_vec_last_kernel_ms(cog_vec *h, float *ms);

#ifdef __cplusplus
}
#endif
#endif /* COG_B200_H */
