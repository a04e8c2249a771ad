/* ORACLE / TEST INFRASTRUCTURE -- not part of the product path.  See mt_oracle.c. */
#ifndef MT_ORACLE_H
#define MT_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
#define MT_MAX_TEAMS 8
#define MT_ROW 17 /* ScenarioConfig.obs_vec_length, MISSION/dca_multiteam/collective_assault_parallel_run.py:62 */

/* POD snapshot of MISSION/dca_multiteam ScenarioConfig (collective_assault_parallel_run.py:27-77) */
typedef struct {
    int32_t n_team;
    int32_t team_size[MT_MAX_TEAMS];
    double size, random_jam_prob;
    int32_t introduce_terrain;
    double terrain_A, terrain_B;
    int32_t max_episode_step, steps_to_narrow;
    double init_distance;
    int32_t half_death_reward;
    uint64_t seed;
} MtConfig;

typedef struct {
    double *pos, *vel, *atk_rad, *terrain;         /* [E,N,2] [E,N,2] [E,N] [E,N] */
    uint8_t *alive, *just_died, *hit, *was_hit;    /* [E,N] */
    uint8_t *bsa;                                  /* [E,N] member of World.before_step_alive_agents */
    double *init_theta;                            /* [E] */
    int32_t *time_step;                            /* [E] */
    uint32_t *obs_counter, *episode;               /* [E] keyed-RNG counters */
} MtState;

typedef struct {
    float *obs, *tobs;   /* [E,N,12,17]; tobs (terminal obs) written where done, may be NULL */
    double *reward;      /* [E,N] */
    uint8_t *done;       /* [E] */
    int32_t *ranking;    /* [E,n_team] team_ranking, valid where done */
    int16_t *slot_id;    /* [E,N,12] or NULL */
} MtOut;

int mt_reset(const MtConfig *cfg, MtState *st, MtOut *out, int E, int64_t env_gid0);
int mt_step(const MtConfig *cfg, MtState *st, MtOut *out, int E, int64_t env_gid0, const double *actions,
            int autoreset);
#ifdef __cplusplus
}
#endif
#endif
