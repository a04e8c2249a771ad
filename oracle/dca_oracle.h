/* ORACLE / TEST INFRASTRUCTURE -- not part of the product path.
 *
 * CPU restatement (plain C, fp64/fp32 exactly as numpy evaluates it) of the reference's DCA
 * step + observation + reset + autoreset + GAE path.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library; the product
 * (hmp2g_b200/) never does.
 *
 * Parity status: PINNED against outputs of the reference itself, executed in the build
 * container under the keyed RNG stream (oracle/ref_harness/ref_env.py ->
 * oracle/make_golden.py -> tests/golden/ *.npz; checked by tests/test_oracle_golden.py).
 * The reference's own tests hold no golden vectors for this path.
 */
#ifndef DCA_ORACLE_H
#define DCA_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* POD snapshot of ScenarioConfig (MISSION/dca/collective_assault_parallel_run.py:30-101). */
typedef struct {
    int32_t num_guards, num_attackers;
    double agent_acc, agent_max_speed, agent_max_rot;
    double agent_shoot_base_radius, agent_shoot_win;
    double random_jam_prob;
    int32_t introduce_terrain;
    double terrain_A, terrain_B;
    int32_t max_episode_step;
    double init_distance;
    int32_t enable_attacker_fluctuation, attacker_fluctuation;
    double size;
    int32_t half_death_reward, reward_debug;
    double reward_debug_value;
    int32_t disable_red_function;
    uint64_t seed;
} OrcConfig;

/* Env state, SoA over [E, N] (N = G + A; guards first).  Caller-owned arrays. */
typedef struct {
    double *pos;      /* [E,N,2] */
    double *vel;      /* [E,N,2] */
    double *atk_rad;  /* [E,N]   */
    double *terrain;  /* [E,N]   */
    uint8_t *alive, *just_died, *hit, *was_hit; /* [E,N] */
    int32_t *num_hit, *num_was_hit;             /* [E,N] */
    double *init_theta;                         /* [E] */
    int32_t *time_step, *n_alive_guards, *n_alive_attackers; /* [E] */
    uint32_t *obs_counter, *episode;            /* [E] keyed-RNG counters */
} OrcState;

/* Per-step outputs.  Any pointer may be NULL (then that output is skipped). */
typedef struct {
    float *obs;        /* [E,G,12,16] */
    float *tobs;       /* [E,G,12,16] terminal obs, written only where done */
    double *reward;    /* [E,G] */
    uint8_t *done;     /* [E] */
    uint8_t *win;      /* [E] valid where done */
    int16_t *slot_id;  /* [E,G,12] entity id gathered into each obs row, -1 = zero row */
    int32_t *nearest;  /* [E,A] guard index each alive attacker aimed at, -1 for dead attackers */
} OrcOut;

/* reset envs [0,E): episode counter++, formation, terrain, observation.  env_gid0 = global id of env 0. */
int orc_reset(const OrcConfig *cfg, OrcState *st, OrcOut *out, int E, int64_t env_gid0, int nthreads);

/* One step.  actions: f64 [E,G] (any NaN in a row => env skipped, outputs untouched);
 * att_force: f64 [E,A,2] planner output (attacker.act[0:2]);  rng_words: u32 [E,G,2,12] or NULL (=> Philox);
 * autoreset != 0 applies EnvAutoReset semantics (UTIL/shm_env.py:33-74). */
int orc_step(const OrcConfig *cfg, OrcState *st, OrcOut *out, int E, int64_t env_gid0,
             const double *actions, const double *att_force, const uint32_t *rng_words,
             int autoreset, int nthreads);

/* Planner stand-in used by the benchmark: force = centroid(alive guards) - pos for alive attackers. */
int orc_plan_centroid(const OrcConfig *cfg, const OrcState *st, int E, double *att_force, int nthreads);

/* GAE / return over a rollout buffer laid out [T,E,G]; traj_len[E] (NULL => T).
 * Follows ALGORITHM/conc/trajectory.py:60-117 (reward_push_forward, threat, gae_finalize_return). */
int orc_gae(const double *reward, const float *value, const uint8_t *dead, const int32_t *traj_len,
            float *ret, float *threat, double *reward_pushed, int T, int E, int G,
            double gamma, double tau, double gamma_fwd, int threat_limit);

void orc_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]);

#ifdef __cplusplus
}
#endif
#endif
