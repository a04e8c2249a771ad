/* dca_b200.h -- C-ABI of libdca_b200.so: the B200-native replacement of HMAP's DCA batched
 * env step + observation + GAE hot path.
 *
 * Nothing comparable exists in the reference as a native interface: today's "ABI" is pickled Python
 * tuples (name, 'step', args) pushed through POSIX shm to worker processes
 * (/root/reference/UTIL/shm_pool.pyx:168-181, :307-376).  Each entry point below names the reference
 * Python interface it replaces.  Binding stubs (ctypes, as a maintainer of the reference would add
 * them in UTIL/shm_env.py) are in INTEGRATION.md.
 *
 * Conventions: return 0 = ok, negative = DCA_E_* ; never throws; kernels are enqueued on the given
 * stream and NOT synchronised; no allocation after dca_create; the caller owns every buffer passed
 * in DcaBuffers (normally torch CUDA tensors); the library owns only its private state store.
 * A handle is not thread-safe (the reference caller is single-threaded, task_runner.py:59-72);
 * distinct handles are independent.  All pointers are DEVICE pointers unless marked host.
 */
#ifndef DCA_B200_H
#define DCA_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DCA_OK 0
#define DCA_E_INVALID (-1)   /* bad argument / unsupported size */
#define DCA_E_CUDA (-2)      /* CUDA runtime error, see dca_last_error */
#define DCA_E_UNBOUND (-3)   /* dca_bind_buffers not called */
#define DCA_E_NODEVICE (-4)  /* no CUDA device: there is NO CPU fallback */

#define DCA_OBS_ROWS 12  /* 6 friend rows (row 0 = self) + 6 hostile rows, collective_assault_env.py:344 */
#define DCA_OBS_ROW 16   /* entity record length, ScenarioConfig.obs_vec_length parallel_run.py:70 */
#define DCA_RNG_WORDS 12 /* u32 words per (guard, list) in an injected observation stream */

/* POD snapshot of ScenarioConfig, taken at env construction like the reference does
 * (MISSION/dca/collective_assault_parallel_run.py:30-101; read at core.py:72-73,
 * collective_assault_env.py:10-13). */
typedef struct DcaConfig {
    int32_t num_guards, num_attackers;
    double agent_acc, agent_max_speed, agent_max_rot;
    double agent_shoot_base_radius, agent_shoot_win;
    double random_jam_prob;
    int32_t introduce_terrain;
    double terrain_A, terrain_B; /* terrain_parameters[0], [1] */
    int32_t max_episode_step;
    double init_distance;
    int32_t enable_attacker_fluctuation, attacker_fluctuation;
    double size;
    int32_t half_death_reward, reward_debug;
    double reward_debug_value;
    int32_t disable_red_function; /* DISALBE_RED_FUNCTION (sic) */
    uint64_t seed;                /* key of the counter-based RNG; replaces SmartPool's base_seed+index, shm_pool.pyx:186 */
    int64_t env_gid0;             /* global id of local env 0 (multi-GPU sharding: rank r owns [gid0, gid0+E)) */
} DcaConfig;

/* Output buffers, bound once; written in place every step (zero-copy hand-off to the torch policy).
 * Replaces the tuple (ob, reward, done, info) of EnvAutoReset.step UTIL/shm_env.py:33-74 and the
 * np.stack in SuperpoolEnv.step UTIL/shm_env.py:145-153. */
typedef struct DcaBuffers {
    float *obs;       /* [E,G,12,16] post-step obs (post-RESET obs where done, shm_env.py:58) */
    float *tobs;      /* [E,G,12,16] terminal obs = info['obs-echo'], written where done; may be NULL */
    float *reward;    /* [E,G] */
    uint8_t *done;    /* [E] */
    uint8_t *win;     /* [E] info['win'], valid where done */
    uint8_t *alive;   /* [E,N] alive mask after the step (after reset where done); may be NULL */
    int16_t *slot_id; /* [E,G,12] entity id gathered into each obs row, -1 = zero row; may be NULL */
    int32_t *nearest; /* [E,A] guard index each alive attacker aimed at (-1 dead); may be NULL */
} DcaBuffers;

/* Full env state in a plain SoA layout, for import/export (parity harness, checkpointing; the
 * reference never checkpoints env state).  Device pointers. */
typedef struct DcaState {
    double *pos;     /* [E,N,2] */
    double *vel;     /* [E,N,2] */
    double *atk_rad; /* [E,N] */
    double *terrain; /* [E,N] */
    uint8_t *alive, *just_died, *hit, *was_hit; /* [E,N] */
    int32_t *num_hit, *num_was_hit;             /* [E,N] */
    double *init_theta;                         /* [E] */
    int32_t *time_step, *n_alive_guards, *n_alive_attackers; /* [E] */
    uint32_t *obs_counter, *episode;            /* [E] */
} DcaState;

typedef struct DcaHandle DcaHandle;
typedef void *dca_stream_t; /* cudaStream_t */

/* replaces: SuperpoolEnv.__init__ -> SmartPool.add_target('env', EnvAutoReset, ...) UTIL/shm_env.py:126-141,
 * make_collective_assault_env parallel_run.py:103-117.  Allocates the state store for n_env envs on `device`. */
int dca_create(const DcaConfig *cfg, int n_env, int device, DcaHandle **out);
int dca_destroy(DcaHandle *h);
int dca_bind_buffers(DcaHandle *h, const DcaBuffers *bufs);

/* replaces: SuperpoolEnv.reset UTIL/shm_env.py:155-164 / EnvAutoReset._real_reset :95-97 ->
 * collective_assaultGlobalEnv.reset parallel_run.py:248-270.  env_mask: u8[E] (NULL = all envs). */
int dca_reset(DcaHandle *h, const uint8_t *env_mask, dca_stream_t stream);

/* replaces: SuperpoolEnv.step UTIL/shm_env.py:145-153 -> EnvAutoReset.step :33-74 ->
 * collective_assaultGlobalEnv.step parallel_run.py:192-246 (World.step core.py:189-207).
 *   actions   i32[E,G]   discrete 0..6 (parallel_run.py:337-343)
 *   att_force f64[E,A,2] scripted-attacker planner output attacker.act[0:2] (cheat_script_ai.py:183-184)
 *   skip      u8[E] or NULL: 1 = ENV-PAUSE (NaN action row, shm_env.py:37-41): env untouched, outputs keep last values
 *   rng_words u32[E,G,2,12] or NULL: injected observation stream; NULL = in-kernel Philox4x32-10 (same words)
 *   autoreset != 0: EnvAutoReset semantics (terminal obs -> tobs, reset, obs = reset obs) */
int dca_step(DcaHandle *h, const int32_t *actions, const double *att_force, const uint8_t *skip,
             const uint32_t *rng_words, int autoreset, dca_stream_t stream);

/* Observation pass only (the k-nearest-neighbour gather kernel on the resident state):
 * collective_assaultEnvV1.observation collective_assault_env.py:228-359 for every guard.
 * Advances the per-env observation counter like any other pass. */
int dca_observe(DcaHandle *h, const uint32_t *rng_words, dca_stream_t stream);

/* replaces: CheatScriptAI.attackers_policy_1 formation planner MISSION/dca/cheat_script_ai.py:19-194 + get_groups
 * :259-297 (DBSCAN targets, k-means teams, team->target and member->ring-slot assignments), which the reference runs
 * inside World.step core.py:189-196.  Fills att_force f64[E,A,2] for dca_step.  The team state (World.start_flag,
 * teams_result_step1) lives in the handle and survives resets like the reference's.  team_label i32[E,A] +
 * team_centroid f64[E,A/5,2], both or neither: inject a k-means result (parity harness) instead of the keyed one. */
int dca_plan(DcaHandle *h, double *att_force, const int32_t *team_label, const double *team_centroid,
             dca_stream_t stream);

/* Benchmark planner stand-in (NOT the reference planner): force = centroid(alive guards) - pos. */
int dca_plan_centroid(DcaHandle *h, double *att_force, dca_stream_t stream);

int dca_export_state(DcaHandle *h, const DcaState *dst, dca_stream_t stream);
int dca_import_state(DcaHandle *h, const DcaState *src, dca_stream_t stream);

/* replaces: trajectory.finalize / reward_push_forward / gae_finalize_return
 * ALGORITHM/conc/trajectory.py:60-117 (= ALGORITHM/common/traj_gae.py:54-133), batched over a
 * rollout buffer [T,E,G]; one trajectory per env column of length traj_len[e] (NULL = T).
 * reward f32 in; ret/threat f32 out; reward_pushed (f32, may be NULL) = reward after push-forward.
 * gamma_fwd: AlgorithmConfig.gamma_in_reward_forwarding_value when gamma_in_reward_forwarding is set -- a dead
 * frame's reward moves to the previous frame scaled by it (traj_gae.py:55-60, conc_4hist/trajectory.py:57-63,
 * ppo_ma/trajectory.py:59-60); 1.0 = the flag is off (plain push).  ABI version 2 added this argument. */
int dca_gae(const float *reward, const float *value, const uint8_t *dead, const int32_t *traj_len,
            float *ret, float *threat, float *reward_pushed, int T, int E, int G,
            double gamma, double tau, double gamma_fwd, int threat_limit, dca_stream_t stream);

/* ---- runner bookkeeping on the device -----------------------------------------------------------------------------
 * replaces: the per-env Python loop of task_runner.Runner.update_runner task_runner.py:95-138 (and its test-run twin
 * update_test_runner :191-217): ENV-PAUSE / Env-Suffered-Reset / Current-Obs-Step / Latest-Reward-Sum /
 * Thread-Episode-Cnt per env, current_n_frame / current_n_episode, and the Recent-Reward-Sum / Recent-Win lists,
 * kept as an episode log indexed by global episode number (episodes are numbered in env-index order within a step, as
 * the reference's loop appends them).  `pause` is the skip mask of the next dca_step (= the NaN action rows
 * MMPlatform.act writes for paused envs, multi_team.py:62-66).  All arrays are caller-owned device buffers. */
typedef struct DcaRunnerState {
    uint8_t *pause;       /* [E] ENV-PAUSE */
    uint8_t *suffered;    /* [E] Env-Suffered-Reset */
    int32_t *cur_step;    /* [E] Current-Obs-Step */
    int32_t *ep_cnt;      /* [E] Thread-Episode-Cnt */
    double *reward_sum;   /* [E,G] Latest-Reward-Sum */
    int64_t *counters;    /* [8]: 0 current_n_frame, 1 current_n_episode, 2 aligned restarts */
    int32_t *fin_slot;    /* [E] scratch: -1 paused, 0 running, slot+1 = finished an episode logged at `slot` */
    int64_t *log_meta;    /* [log_capacity,4]: env, win, steps, episode number (1-based) */
    double *log_reward;   /* [log_capacity,G] reward sums of the logged episodes */
    int32_t log_capacity; /* ring size: >= E + the longest window that is read back */
} DcaRunnerState;
/* init_runner (task_runner.py:77-93): zero everything, Env-Suffered-Reset = 1. */
int dca_runner_init(const DcaRunnerState *s, int n_env, int n_agent, dca_stream_t stream);
/* update_runner: done u8[E], win u8[E] (read where done), reward f32[E,G] of the step just taken. */
int dca_runner_update(const DcaRunnerState *s, int n_env, int n_agent, int align_episode, const uint8_t *done,
                      const uint8_t *win, const float *reward, dca_stream_t stream);
/* _checkout_interested_agents (task_runner.py:231-275) over episodes first_episode .. first_episode+n_episodes-1:
 * out f64[n_team+1] = mean episode reward of each team's agents (team_of_agent i32[G]), then the win rate. */
int dca_runner_report(const DcaRunnerState *s, int n_agent, int64_t first_episode, int n_episodes,
                      const int32_t *team_of_agent, int n_team, double *out, dca_stream_t stream);

/* ---- conc_4hist history pool -------------------------------------------------------------------------------------
 * replaces: ShellEnvWrapper.interact_with_env's '_Histpool_Obs_' handling + solve_duplicate
 * ALGORITHM/conc_4hist/shell_env.py:25-99 and roll_hisory ALGORITHM/conc_4hist/cython_func.pyx:32-53.
 * obs f32 [E,G,12,16] (the env's observation, empty rows 0 or NaN); his_pool f32 [E,G,12,16] in/out (the
 * '_Histpool_Obs_' state, zeros at start); pause / suffered_reset u8 [E] (ENV-PAUSE / Env-Suffered-Reset, NULL = all 0);
 * obs_feed f32 [E,G,24,16] out = what the policy net is fed (new rows ++ history rows, empty rows NaN, everything NaN
 * for a dead agent).  Envs are not compacted: a paused env keeps its slot, its obs_feed is NaN and its history is
 * cleared (shell_env.py:44). */
int dca_histpool_roll(const float *obs, float *his_pool, const uint8_t *pause, const uint8_t *suffered_reset,
                      float *obs_feed, int n_env, int n_agent, dca_stream_t stream);

/* ---- advantage normalisation -------------------------------------------------------------------------------------
 * replaces: adv = ret - value; adv = (adv - adv.mean()) / (adv.std() + 1e-5) over the pooled batch,
 * ALGORITHM/conc/ppo.py:53-55 (= ALGORITHM/common/ppo_sampler.py:67-77; numpy std, ddof 0).
 * dca_adv_moments: stats f64[3] = (sum adv, sum adv^2, count) over the n elements whose valid[i / valid_group] != 0
 * (valid NULL = all); scratch f64[1776] (device).  Across GPUs: all-reduce `stats` (3 doubles), then normalise.
 * dca_adv_normalize: adv f32[n] out (0 where not valid). */
int dca_adv_moments(const float *ret, const float *value, const uint8_t *valid, int valid_group, uint64_t n,
                    double *scratch, double *stats, dca_stream_t stream);
int dca_adv_normalize(const float *ret, const float *value, const uint8_t *valid, int valid_group, uint64_t n,
                      const double *stats, float *adv, dca_stream_t stream);
/* dca_adv_moments_static: NO counterpart in the reference -- added for callers
 * that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid allocation).  Same
 * kernels, arguments and result as dca_adv_moments, minus `scratch`: the per-CTA partials live in a library-owned
 * `static __device__` array (592 x 3 doubles = 14 208 B, held by every GPU that loads the library until exit; its
 * device address is resolved with cudaGetSymbolAddress once per device).  The call allocates nothing (no cudaMalloc /
 * cudaFree / allocating library routine).  n > DCA_ADV_STATIC_MAX_N is refused with DCA_E_INVALID.  Every call on a
 * device shares the one array (cleared at the start of each call): calls that overlap on the GPU -- two streams, or
 * two host threads -- race; serialise them on one stream or use dca_adv_moments with separate scratch buffers. */
#define DCA_ADV_STATIC_MAX_N (1ull << 28) /* 268 435 456 elements = 7 x a [180,4096,50] rollout buffer */
int dca_adv_moments_static(const float *ret, const float *value, const uint8_t *valid, int valid_group, uint64_t n,
                           double *stats, dca_stream_t stream);

/* ---- observation normalisation of the conc / conc_4hist policy -------------------------------------------------------
 * replaces: DynamicNorm.forward ALGORITHM/common/norm.py:31-69 (exclude_nan, exclude_one_hot) as called from Net._act
 * ALGORITHM/conc_4hist/net.py:194-198 (= ALGORITHM/conc/net.py:300-304), on the policy's input [rows, 16] float32.
 * dca_norm_moments: stats f64[49] = sum x[16], sum x^2[16], number of values outside {-1,0,1}[16], row count, over the
 *   rows that hold no NaN; scratch f64[592*49] (device).  The caller merges them into the running moments (and
 *   all-reduces the 49 doubles across GPUs first: every rank then holds the same moments).
 * dca_norm_apply: y = nan_to_num(clip((x - mean) / sqrt(var + 1e-8), -10, 10), 0) in float32 like torch; row_nan u8[rows] =
 *   isnan(x).any(-1) (the policy's entity mask); y may alias x. */
#define DCA_NORM_STATS 49
#define DCA_NORM_SCRATCH (592 * 49)
int dca_norm_moments(const float *x, uint64_t rows, int dim, double *scratch, double *stats, dca_stream_t stream);
int dca_norm_apply(const float *x, const float *mean, const float *var, uint64_t rows, int dim, float *y,
                   uint8_t *row_nan, dca_stream_t stream);

/* replaces: DynamicNorm apply + mask + nan_to_num + AT_obs_encoder (Linear 16->48, ReLU, Linear 48->48) of Net._act
 * ALGORITHM/conc_4hist/net.py:194-199 on every entity row, fused (float32 FFMA, nn.Linear layouts w1 [48,16], w2 [48,48]).
 *   x f32 [rows,16] raw observation rows (NaN rows = empty / dead); x_norm f32 [rows,16], row_nan u8 [rows] as
 *   dca_norm_apply; v f32 [rows,48] encodings. */
int dca_policy_encode(const float *x, const float *mean, const float *var, const float *w1, const float *b1,
                      const float *w2, const float *b2, uint64_t rows, float *x_norm, uint8_t *row_nan, float *v,
                      dca_stream_t stream);

/* replaces: one nn.Linear (+ the torch.cat in front of it, + ReLU) of the conc / conc_4hist policy behind the encoder --
 * Concentration's query / key / value / motivate / forward layers (ALGORITHM/common/conc.py:30-62) and the logit /
 * value / threat heads (ALGORITHM/conc_4hist/net.py:213-229): out = act(a Wa^T + b Wb^T + bias), float32 FFMA.
 *   a f32 [rows,ka] with row stride lda, b f32 [rows,kb] with row stride ldb (kb = 0: no second input); weights are read
 *   as W[n * sn + k * sk] (nn.Linear weight: sn = in_features, sk = 1; a transposed matrix: sn = 1, sk = row length);
 *   bias f32 [n] or NULL; n <= 48; relu: bit 0 = ReLU, bit 1 = float32 FFMA arithmetic instead of the default, three tf32
 *   tensor-core products per multiply with the operands split into head + tail (float32-class accuracy, error 2^-22 of
 *   a product; taken anyway when ka or kb is not a multiple of 8); out f32 [rows,n] with row stride ldo.  ka, kb, lda,
 *   ldb multiples of 4, a and b 16-byte aligned. */
int dca_policy_linear(const float *a, int lda, int ka, const float *wa, int wa_sn, int wa_sk, const float *b, int ldb, int kb,
                      const float *wb, int wb_sn, int wb_sk, const float *bias, int n, int relu, uint64_t rows, float *out,
                      int ldo, dca_stream_t stream);

/* replaces: the part of Concentration.forward between its matrix products, ALGORITHM/common/conc.py:30-58 (masked
 * softmax attention of the agent's own encoding over one entity list, attended encoding, top-n_focus gather with the
 * skip connection), for the conc / conc_4hist policy (ALGORITHM/conc_4hist/net.py:205-207).  One warp per agent.
 *   v f32 [B,24,48] entity encodings, x f32 [B,24,16] normalised observations, dead u8 [B,24] (dca_norm_apply's row mask),
 *   qk f32 [B,48] = (vs W_query) W_key^T; the list is entities first .. first+n-1 of the 24.
 *   va f32 [B,48] = sum_j score_j ve_j (multiply by W_val); stack f32 [B,(n_focus+1)*64] = [vs|zs] ++ the n_focus
 *   entities with the highest scores, [ve|ze] each, ties to the lowest entity index. */
int dca_conc_attend(const float *v, const float *x, const uint8_t *dead, const float *qk, int first, int n, int n_focus,
                    uint64_t n_agent, float *va, float *stack, dca_stream_t stream);

/* dca_conc_attend for two adjacent entity lists of every agent in one pass (conc_4hist: entities 1..11 and 12..23,
 * ALGORITHM/conc_4hist/net.py:170-181,205-207): list a = entities first_a .. first_a+n_a-1 with qk_a / focus_a, list b the
 * n_b entities behind it with qk_b / focus_b.  Outputs as two dca_conc_attend calls, bit for bit. */
int dca_conc_attend2(const float *v, const float *x, const uint8_t *dead, const float *qk_a, const float *qk_b, int first_a,
                     int n_a, int focus_a, int n_b, int focus_b, uint64_t n_agent, float *va_a, float *stack_a, float *va_b,
                     float *stack_b, dca_stream_t stream);

/* ---- dca_multiteam -------------------------------------------------------------------------------------------------
 * The sibling env MISSION/dca_multiteam (N teams, every agent policy-controlled, per-hit success draw, shrinking kill
 * circle, team ranking).  Same conventions as the dca_* calls above. */
#define DCA_MT_MAX_TEAMS 8
#define DCA_MT_OBS_ROW 17 /* ScenarioConfig.obs_vec_length, MISSION/dca_multiteam/collective_assault_parallel_run.py:62 */
/* POD snapshot of MISSION/dca_multiteam ScenarioConfig (collective_assault_parallel_run.py:27-77) */
typedef struct DcaMtConfig {
    int32_t n_team;
    int32_t team_size[DCA_MT_MAX_TEAMS]; /* N_AGENT_EACH_TEAM; uids are contiguous per team (arrange_id) */
    double size, random_jam_prob;
    int32_t introduce_terrain;
    double terrain_A, terrain_B;
    int32_t max_episode_step, steps_to_narrow; /* MaxEpisodeStep, StepsToNarrowCircle */
    double init_distance;
    int32_t half_death_reward;
    uint64_t seed;
    int64_t env_gid0;
} DcaMtConfig;
/* Output buffers (caller-owned, written in place).  replaces: the (ob, reward, done, info) tuple of EnvAutoReset.step
 * UTIL/shm_env.py:33-74 around collective_assaultGlobalEnv.step dca_multiteam/collective_assault_parallel_run.py:156-191 */
typedef struct DcaMtBuffers {
    float *obs;       /* [E,N,12,17] post-step obs (post-RESET obs where done) */
    float *tobs;      /* [E,N,12,17] terminal obs = info['obs-echo'], written where done; may be NULL */
    float *reward;    /* [E,N] */
    uint8_t *done;    /* [E] */
    int32_t *ranking; /* [E,n_team] info['team_ranking'] (0 = first, -1 = all teams draw), written where done */
    uint8_t *alive;   /* [E,N]; may be NULL */
    int16_t *slot_id; /* [E,N,12] uid gathered into each obs row, -1 = zero row; may be NULL */
} DcaMtBuffers;
/* SoA view of the env state for export / import (device pointers, all required) */
typedef struct DcaMtState {
    double *pos, *vel, *atk_rad, *terrain;              /* [E,N,2] [E,N,2] [E,N] [E,N] */
    uint8_t *alive, *just_died, *hit, *was_hit, *bsa;   /* [E,N]; bsa = member of World.before_step_alive_agents */
    double *init_theta;                                 /* [E] */
    int32_t *time_step;                                 /* [E] */
    uint32_t *obs_counter, *episode;                    /* [E] keyed-RNG counters */
} DcaMtState;
typedef struct DcaMtHandle DcaMtHandle;
/* replaces: make_collective_assault_env dca_multiteam/collective_assault_parallel_run.py:80-89 x n_env */
int dca_mt_create(const DcaMtConfig *cfg, int n_env, int device, DcaMtHandle **out);
int dca_mt_destroy(DcaMtHandle *h);
int dca_mt_bind_buffers(DcaMtHandle *h, const DcaMtBuffers *b);
/* replaces: collective_assaultGlobalEnv.reset :193-209 (env_mask u8[E], NULL = all) */
int dca_mt_reset(DcaMtHandle *h, const uint8_t *env_mask, dca_stream_t stream);
/* replaces: EnvAutoReset.step UTIL/shm_env.py:33-74 -> collective_assaultGlobalEnv.step :156-191.
 * actions i32[E,N] in 0..6 (one per agent, all teams); skip u8[E] (NULL = none) = NaN action rows. */
int dca_mt_step(DcaMtHandle *h, const int32_t *actions, const uint8_t *skip, int autoreset, dca_stream_t stream);
int dca_mt_export_state(DcaMtHandle *h, const DcaMtState *dst, dca_stream_t stream);
int dca_mt_import_state(DcaMtHandle *h, const DcaMtState *src, dca_stream_t stream);
/* Diagnostics, as dca_debug_phase_clocks: slots 0 start, 1 loaded, 2 terrain, 3 laser, 4 deaths+integrate, 5 done/ranking
 * (+ terminal obs and reset on done), 6 observation tasks, 7 gather, 8 state stored; inside the laser phase: 9 team
 * boxes, 10 scan done, 11 barrier passed. */
int dca_mt_debug_phase_clocks(DcaMtHandle *h, uint64_t *clocks);

/* Options.  DCA_OPT_EMPTY_ROWS_NAN: write invisible / jammed / empty observation rows (and the whole observation of a
 * dead guard) as NaN instead of 0 -- what ShellEnvWrapper does to every observation on the host before the policy sees
 * it (ALGORITHM/conc/shell_env.py:72-75); default 0 = the env's own zeros (collective_assault_env.py:313-316,340-343). */
#define DCA_OPT_EMPTY_ROWS_NAN 1
/* DCA_OPT_OVERLAP_PLANNER: a dca_step that directly follows dca_plan / dca_plan_centroid on the same stream (same
 * att_force, nothing enqueued between them) is launched as a programmatic dependent of the planner kernel: blob load,
 * terrain, aim and laser run while the planner is still computing the forces; the step orders itself behind the
 * planner right before it reads them.  Results are identical; `nearest` may be overwritten while the planner runs.
 * Default 0. */
#define DCA_OPT_OVERLAP_PLANNER 2
/* DCA_OPT_PIPELINE_STEPS: a loop that issues plan / step / plan / step ... back to back on one stream (nothing between
 * a dca_step and the next dca_plan* call, the same att_force) hands every env over individually instead of launch by
 * launch: the planner warp of env e waits only for the step CTA of env e (which publishes a per-env epoch word once its
 * blob is stored), the next step's CTA of env e only for that planner warp and for its own blob.  The kernels are
 * launched as programmatic dependents of each other, so the first CTAs of launch t+1 run in the SM slots the last,
 * partly filled wave of launch t leaves empty (a quarter of a 4096-env launch on 148 SMs).  Results are identical.
 * Anything else on the stream (a policy reading obs, dca_reset, a pause mask) is ordered behind the whole previous
 * launch as usual and restarts the pipeline.  The epoch counters live in the handle on the host, so launches on a
 * stream that is being captured into a CUDA graph are ordinary ones (a replay would publish the captured epochs again).
 * Default 0. */
#define DCA_OPT_PIPELINE_STEPS 4
int dca_set_option(DcaHandle *h, int option, int value);

const char *dca_last_error(const DcaHandle *h); /* host string; h may be NULL for create/gae errors */
int dca_kernel_launches(const DcaHandle *h, int64_t *count); /* kernels launched through this handle so far */
/* Diagnostics: u64[E,16] device buffer receiving clock64() of thread 0 at the phase boundaries of the next
 * dca_step launches (NULL switches it off).  Slots: 0 start, 1 blob loaded, 2 decode+terrain, 3 aim, 4 laser,
 * 5 deaths+integrate, 6 reward/done, 7 entity table, 8 kNN+shuffle, 9 gather, 10 reset path done, 11 blob stored. */
int dca_debug_phase_clocks(DcaHandle *h, uint64_t *clocks);
int dca_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif
