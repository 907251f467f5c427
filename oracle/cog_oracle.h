/* oracle/cog_oracle.h — TEST INFRASTRUCTURE (the checker), never shipped or timed
 * as the product.
 *
 * Plain-C, single-threaded restatement of the reference's vectorised hot path
 * (aapo-kossi/gym-eldorado "City of Gold"): procedural board generation, reset,
 * step (+ auto-reset), mask maintenance and the masked random sampler, operating
 * on buffers with the reference's exported record layouts.
 *
 * PARITY PINNED: tests/test_oracle.py checks this file against the
 * reference itself compiled here (oracle/_ref/libcog_ref.so, unmodified sources)
 * and against golden fixtures under tests/golden/ minted from that build.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
 * include, link or execute anything under oracle/.
 */
#ifndef COG_ORACLE_H
#define COG_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* exported record sizes (reference include/api.h:67-161) */
#define ORC_OBS_SIZE 17216
#define ORC_MASK_SIZE 128
#define ORC_ACTION_SIZE 64
#define ORC_INFO_SIZE 192

typedef struct orc_vec orc_vec;

/* which: 0 obs, 1 selected masks, 2 rewards, 3 dones, 4 infos, 5 agent_selection,
 *        6 sampler actions   (same numbering as oracle/ref_shim.cpp) */
orc_vec *orc_create(uint64_t n_envs, uint64_t env_index_offset, uint64_t sampler_seed);
void orc_destroy(orc_vec *v);
/* returns 0, or -2 when board generation failed (generate_map_failure) */
int orc_reset(orc_vec *v, uint32_t seed, uint8_t n_players, uint8_t n_pieces,
              int difficulty, uint32_t max_steps);
int orc_reset_again(orc_vec *v); /* vec_cog_env::reset() without parameters */
int orc_step(orc_vec *v, const void *actions);
/* on = 0: single-env semantics (no auto-reset, steps on a finished env are ignored) */
void orc_set_autoreset(orc_vec *v, int on);
void orc_sample(orc_vec *v, const void *masks);
void *orc_buffer(orc_vec *v, int which);
/* K x (sample(selected masks); step(actions)); returns seconds */
double orc_time_run(orc_vec *v, uint64_t steps);
/* debugging dump with the same 16 slots as ref_hidden() */
void orc_hidden(orc_vec *v, uint64_t i, uint32_t *out16);

/* standalone Map::generate twin of ref_map_generate(); obs has rows*48*7 bytes */
int orc_map_generate(uint32_t seed, uint8_t n_pieces, int difficulty, void *obs,
                     uint32_t rows, uint32_t *out_meta, int32_t *out_xy,
                     uint32_t max_hexes);

/* PRNG known-answer helpers */
uint32_t orc_minstd_seed(uint64_t s);
uint32_t orc_minstd_next(uint32_t *state);
uint64_t orc_uniform(uint32_t *state, uint64_t lo, uint64_t hi);

#ifdef __cplusplus
}
#endif
#endif
