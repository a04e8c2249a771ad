/* ORACLE / TEST INFRASTRUCTURE -- not part of the product path.
 *
 * CPU restatement (plain C) of the reference's `dca_multiteam` env:
 *   MISSION/dca_multiteam/collective_assault_parallel_run.py:156-279 (step / reset / _get_done / _set_action)
 *   MISSION/dca_multiteam/core.py:100-240                            (World.step, laser, integrate, kill circle)
 *   MISSION/dca_multiteam/envs/collective_assault_env.py:42-239     (reset_world, reward, observation)
 *   UTIL/shm_env.py:33-97                                           (EnvAutoReset)
 * Parity status: PINNED against the reference itself, executed in the build container under the keyed RNG stream
 * (oracle/ref_harness/ref_mt_env.py -> oracle/make_golden_mt.py -> tests/golden/mt_*.npz).
 * Keyed draws (oracle/dca_rng.py layout): observation words as in DCA (item = agent*2 + list); reset world item
 * word0 = terrain theta, word1 = formation theta, agent item word0 = heading noise; laser hit draw of (shooter i,
 * target j) = word0 of Philox(gid, obs_counter, i*N + j, site 4).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "mt_oracle.h"

void orc_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]);

#define VIS_N 6
#define MT_ROWS 12
#define MT_LEN (MT_ROW * MT_ROWS)
#define MAX_N 512
static const double PI = 3.141592653589793;
static const int JAM_W[6] = {0, 1, 3, 5, 7, 9};
static const int SHUF_W[6] = {-1, 2, 4, 6, 8, 10};
enum { SITE_OBS = 1, SITE_RESET = 2, SITE_LASER = 4 };

static int mt_n(const MtConfig *c) {
    int n = 0;
    for (int t = 0; t < c->n_team; ++t) n += c->team_size[t];
    return n;
}
/* arrange_id (UTIL/tensor_ops.py): contiguous uid ranges per team */
static void mt_layout(const MtConfig *c, int *team, int *tid, int *first) {
    int k = 0;
    for (int t = 0; t < c->n_team; ++t) {
        first[t] = k;
        for (int i = 0; i < c->team_size[t]; ++i, ++k) { team[k] = t; tid[k] = i; }
    }
    first[c->n_team] = k;
}

static double py_fmod(double a, double b) {
    double m = fmod(a, b);
    if (m != 0.0) {
        if ((b < 0) != (m < 0)) m += b;
    } else {
        m = copysign(0.0, b);
    }
    return m;
}
static double reg_angle(double rad) { return py_fmod(rad + PI, 2 * PI) - PI; }

/* World.set_terrain_adv core.py:147-157 + get_terrain :226-234, for the agents of `before_step_alive_agents` */
static void mt_terrain(const MtConfig *cfg, MtState *st, int e, int N) {
    const double th = st->init_theta[e];
    const double c = cos(th), s = sin(th);
    for (int i = 0; i < N; ++i) {
        const int k = e * N + i;
        if (!st->bsa[k]) continue;
        if (!cfg->introduce_terrain) { st->terrain[k] = 1.0; continue; }
        const double X = st->pos[2 * k], Y = st->pos[2 * k + 1];
        const double X_ = X * c + Y * s;
        const double Y_ = -X * s + Y * c;
        const double a = 0.1 * X_, b = 0.1 * Y_;
        double Z = -1 + cfg->terrain_B * (a * a + b * b);
        Z = Z - cfg->terrain_A * cos(2 * PI * (0.3 * X_));
        Z = Z - cfg->terrain_A * cos(2 * PI * (0.5 * Y_));
        st->terrain[k] = -Z;
    }
}

/* reset_world envs/collective_assault_env.py:42-82 */
static void mt_reset_world(const MtConfig *cfg, MtState *st, int e, int64_t gid) {
    const int N = mt_n(cfg);
    int team[MAX_N], tid[MAX_N], first[MT_MAX_TEAMS + 1];
    mt_layout(cfg, team, tid, first);
    st->episode[e] += 1;
    const uint32_t ep = st->episode[e];
    const uint32_t k0 = (uint32_t)cfg->seed, k1 = (uint32_t)(cfg->seed >> 32);
    uint32_t w[4];
    orc_philox((uint32_t)gid, ep, 0, SITE_RESET, k0, k1, w);
    st->time_step[e] = 0;
    st->init_theta[e] = (2 * (w[0] * 0x1p-32) - 1) * PI;          /* :47 */
    const double random_theta = (2 * (w[1] * 0x1p-32) - 1) * PI;  /* :50 */
    const double xMin = -5, xMax = 5, yMin = -5, yMax = 5;         /* init_box :16 */
    const double xMid = xMin / 2 + xMax / 2, yMid = yMin / 2 + yMax / 2;
    const int row = 5;
    const double row_dis = 0.1;
    for (int i = 0; i < N; ++i) {
        const int k = e * N + i;
        st->alive[k] = 1;
        st->vel[2 * k] = st->vel[2 * k + 1] = 0.0;
        const int n_in_team = cfg->team_size[team[i]];
        const double team_pos_dir = (2 * PI / cfg->n_team) * team[i] + random_theta;
        const double cs = cos(team_pos_dir), sn = sin(team_pos_dir);
        const double team_face_dir = team_pos_dir + PI;
        const double x_ = xMid + cfg->init_distance / 2 + row_dis * (tid[i] % row);
        const double y_ = yMid + ((double)(tid[i] / row) / ((double)n_in_team / row) - 0.5) * (yMax - yMin) * 0.25;
        /* np.dot((x_,y_), team_rotate.T): [x c - y s, x s + y c] */
        st->pos[2 * k] = x_ * cs + y_ * (-sn);
        st->pos[2 * k + 1] = x_ * sn + y_ * cs;
        orc_philox((uint32_t)gid, ep, (uint32_t)(1 + i), SITE_RESET, k0, k1, w);
        const double noise = ((w[0] * 0x1p-32) - 0.5) / 12;       /* :72 */
        st->atk_rad[k] = team_face_dir + noise;
        st->hit[k] = 0;
        st->was_hit[k] = 0;
        st->bsa[k] = 1; /* reset(): before_step_alive_agents = all, parallel_run.py:201 */
    }
}

static float f32_dist(float xi, float yi, float xj, float yj) {
    volatile float dx = xj - xi, dy = yj - yi;
    volatile float sx = dx * dx, sy = dy * dy;
    volatile float s = sx + sy;
    return sqrtf(s);
}
static void top6(const float *d, int n, int idx[VIS_N]) {
    uint8_t used[MAX_N];
    memset(used, 0, (size_t)n);
    for (int k = 0; k < VIS_N; ++k) {
        int best = -1;
        for (int j = 0; j < n; ++j) {
            if (used[j]) continue;
            if (best < 0 || d[j] < d[best]) best = j;
        }
        idx[k] = best;
        if (best >= 0) used[best] = 1;
    }
}

/* observation of every agent, envs/collective_assault_env.py:131-226 */
static void mt_observation(const MtConfig *cfg, MtState *st, int e, int64_t gid, float *obs /*[N,204]*/,
                           int16_t *slot_id /*[N,12] or NULL*/) {
    const int N = mt_n(cfg);
    int team[MAX_N], tid[MAX_N], first[MT_MAX_TEAMS + 1];
    mt_layout(cfg, team, tid, first);
    st->obs_counter[e] += 1;
    const uint32_t counter = st->obs_counter[e];
    const uint32_t k0 = (uint32_t)cfg->seed, k1 = (uint32_t)(cfg->seed >> 32);
    /* prepare_common_data_for_obs :116-128: [alive, pos2, atk_rad, vel2, iden, terrain, bi_hot(tid) 8, team] */
    float tab[MAX_N][MT_ROW];
    for (int i = 0; i < N; ++i) {
        const int k = e * N + i;
        tab[i][0] = (float)st->alive[k];
        tab[i][1] = (float)st->pos[2 * k];
        tab[i][2] = (float)st->pos[2 * k + 1];
        tab[i][3] = (float)st->atk_rad[k];
        tab[i][4] = (float)st->vel[2 * k];
        tab[i][5] = (float)st->vel[2 * k + 1];
        tab[i][6] = (float)i;
        tab[i][7] = (float)st->terrain[k];
        int v = tid[i]; /* get_binary_n_rows: n_int/2 as float, then .astype(np.int8) */
        for (int b = 0; b < 8; ++b) {
            tab[i][8 + b] = (float)((((v % 2) + 2) % 2) == 1);
            v = (int)(int8_t)(int32_t)(v / 2.0);
        }
        tab[i][16] = (float)team[i];
    }
    const float obs_range = 2.0f;
    for (int g = 0; g < N; ++g) {
        float *o = obs + (size_t)g * MT_LEN;
        int16_t *sid = slot_id ? slot_id + g * MT_ROWS : NULL;
        memset(o, 0, sizeof(float) * MT_LEN);
        if (sid) for (int r = 0; r < MT_ROWS; ++r) sid[r] = -1;
        if (!st->alive[e * N + g]) continue; /* :176-179, no draws */
        /* foe list = other teams in team order, ally list = own team (:162-170) */
        int foe[MAX_N], nf = 0, ally[MAX_N], na = 0;
        for (int j = 0; j < N; ++j) {
            if (team[j] == team[g]) ally[na++] = j; else foe[nf++] = j;
        }
        float dh[MAX_N], df[MAX_N];
        for (int a = 0; a < nf; ++a)
            dh[a] = st->alive[e * N + foe[a]] ? f32_dist(tab[g][1], tab[g][2], tab[foe[a]][1], tab[foe[a]][2]) : INFINITY;
        for (int a = 0; a < na; ++a)
            df[a] = st->alive[e * N + ally[a]] ? f32_dist(tab[g][1], tab[g][2], tab[ally[a]][1], tab[ally[a]][2]) : INFINITY;
        int hs[VIS_N], fs[VIS_N];
        top6(dh, nf, hs);
        top6(df, na, fs);
        for (int lst = 0; lst < 2; ++lst) {
            int vis[VIS_N], n = 0;
            const int *srt = lst == 0 ? hs : fs;
            const float *d = lst == 0 ? dh : df;
            const int *ids = lst == 0 ? foe : ally;
            const float lim = lst == 0 ? obs_range : obs_range * 1.5f;
            for (int k = 0; k < VIS_N; ++k)
                if (srt[k] >= 0 && d[srt[k]] <= lim) vis[n++] = srt[k];
            int self_id = -1, firstv = 0;
            if (lst == 1) { self_id = vis[0]; firstv = 1; }
            int *src = vis + firstv;
            const int m = n - firstv;
            uint32_t w[12];
            for (int b = 0; b < 3; ++b)
                orc_philox((uint32_t)gid, counter, (uint32_t)(g * 2 + lst), SITE_OBS | (b << 8), k0, k1, w + 4 * b);
            for (int i = m - 1; i >= 1; --i) {
                int j = (int)(((uint64_t)w[SHUF_W[i]] * (uint64_t)(i + 1)) >> 32);
                int t = src[i]; src[i] = src[j]; src[j] = t;
            }
            int n_mv = 0;
            for (int k = 0; k < m; ++k) n_mv += (w[JAM_W[k]] * 0x1p-32) < cfg->random_jam_prob;
            const int keep = m - n_mv;
            int row = lst == 1 ? 0 : VIS_N; /* friend block first :222 */
            if (lst == 1) {
                memcpy(o + row * MT_ROW, tab[ids[self_id]], sizeof(float) * MT_ROW);
                if (sid) sid[row] = (int16_t)ids[self_id];
                row++;
            }
            for (int k = 0; k < keep; ++k, ++row) {
                memcpy(o + row * MT_ROW, tab[ids[src[k]]], sizeof(float) * MT_ROW);
                if (sid) sid[row] = (int16_t)ids[src[k]];
            }
        }
    }
}

static int laser_hit(double ox, double oy, double tx, double ty, double fanRadius, double fanOpenRad, double fanDirRad) {
    const double dx = tx - ox, dy = ty - oy;
    const double dis_square = dx * dx + dy * dy;
    if (dis_square > fanRadius * fanRadius) return 0;
    const double ori_rad_pos = fanDirRad + fanOpenRad / 2;
    const double ori_rad_neg = fanDirRad - fanOpenRad / 2;
    const double ori_2tgt = atan2(dy, dx);
    const double d1rad = fabs(reg_angle(ori_rad_pos - ori_2tgt));
    const double d2rad = fabs(reg_angle(ori_rad_neg - ori_2tgt));
    return d1rad <= fanOpenRad && d2rad <= fanOpenRad;
}

static void mt_env_step(const MtConfig *cfg, MtState *st, MtOut *out, int e, int64_t gid, const double *actions,
                        int autoreset) {
    const int N = mt_n(cfg);
    int team[MAX_N], tid[MAX_N], first[MT_MAX_TEAMS + 1];
    mt_layout(cfg, team, tid, first);
    const double *act_in = actions + (size_t)e * N;
    for (int i = 0; i < N; ++i)
        if (isnan(act_in[i])) return; /* ENV-PAUSE shm_env.py:37-41 */
    const uint32_t k0 = (uint32_t)cfg->seed, k1 = (uint32_t)(cfg->seed >> 32);
    double ax[MAX_N], ay[MAX_N], ar[MAX_N];
    for (int i = 0; i < N; ++i) { /* _set_action parallel_run.py:283-323; accel 3, max_rot 0.17 (env :27-29) */
        double a0 = 0, a1 = 0, a2 = 0;
        const double a = act_in[i];
        if (a == 1) a0 = +1.0;
        if (a == 2) a0 = -1.0;
        if (a == 3) a1 = +1.0;
        if (a == 4) a1 = -1.0;
        if (a == 5) a2 = +0.17;
        if (a == 6) a2 = -0.17;
        ax[i] = a0 * 3; ay[i] = a1 * 3; ar[i] = a2;
    }
    mt_terrain(cfg, st, e, N); /* parallel_run.py:164: the list of the PREVIOUS world.step (or of reset) */
    /* ---- World.step core.py:100-113 */
    uint8_t *alive = st->alive + (size_t)e * N, *bsa = st->bsa + (size_t)e * N;
    uint8_t *hit = st->hit + (size_t)e * N, *was = st->was_hit + (size_t)e * N, *jd = st->just_died + (size_t)e * N;
    double *pos = st->pos + (size_t)e * N * 2, *vel = st->vel + (size_t)e * N * 2;
    double *atk = st->atk_rad + (size_t)e * N, *ter = st->terrain + (size_t)e * N;
    for (int i = 0; i < N; ++i) bsa[i] = alive[i];
    for (int i = 0; i < N; ++i) if (bsa[i]) { hit[i] = 0; was[i] = 0; }  /* :161-165 */
    const uint32_t lcount = st->obs_counter[e];
    for (int i = 0; i < N; ++i) {                                         /* :167-190 */
        if (!bsa[i]) continue;
        for (int j = 0; j < N; ++j) {
            if (!bsa[j] || team[i] == team[j]) continue;
            if (!laser_hit(pos[2 * i], pos[2 * i + 1], pos[2 * j], pos[2 * j + 1], 0.4 * ter[i], PI / 4, atk[i])) continue;
            double adv = ter[i] - ter[j];
            adv = fmax(fmin(adv, 0.1), -0.1);
            const double prob = 0.8 + adv;
            uint32_t w[4];
            orc_philox((uint32_t)gid, lcount, (uint32_t)(i * N + j), SITE_LASER, k0, k1, w);
            if ((w[0] * 0x1p-32) < prob) { hit[i] = 1; was[j] = 1; }
        }
    }
    for (int i = 0; i < N; ++i) if (!alive[i]) jd[i] = 0;                  /* :193-195 */
    for (int i = 0; i < N; ++i) if (bsa[i] && was[i]) { alive[i] = 0; jd[i] = 1; } /* :198-201 */
    for (int i = 0; i < N; ++i) {                                         /* integrate_state :204-222 */
        if (!bsa[i]) continue;
        double vx = vel[2 * i] * (1 - 0.25), vy = vel[2 * i + 1] * (1 - 0.25);
        vx += (ax[i] / 1.0) * 0.1; vy += (ay[i] / 1.0) * 0.1;
        const double speed = sqrt(vx * vx + vy * vy);
        if (speed > 1.0) {
            const double sp2 = sqrt(vx * vx + vy * vy);
            vx = vx / sp2 * 1.0; vy = vy / sp2 * 1.0;
        }
        vel[2 * i] = vx; vel[2 * i + 1] = vy;
        if (alive[i]) atk[i] = reg_angle(atk[i] + ar[i]);
        pos[2 * i] += vx * 0.1; pos[2 * i + 1] += vy * 0.1;
    }
    {   /* update_kill_circle :115-132 */
        const double init_r = cfg->size, term_r = cfg->size / 3.5;
        double frac = (double)st->time_step[e] / (double)cfg->steps_to_narrow;
        if (frac > 1) frac = 1;
        const double cur = init_r + (term_r - init_r) * frac;
        const double cur2 = cur * cur;
        for (int i = 0; i < N; ++i) {
            if (!bsa[i] || !alive[i]) continue;
            const double d2 = pos[2 * i] * pos[2 * i] + pos[2 * i + 1] * pos[2 * i + 1];
            if (d2 > cur2) { alive[i] = 0; was[i] = 1; jd[i] = 1; }
        }
    }
    /* ---- obs, reward, done (parallel_run.py:170-191) */
    float *obs = out->obs + (size_t)e * N * MT_LEN;
    int16_t *sid = out->slot_id ? out->slot_id + (size_t)e * N * MT_ROWS : NULL;
    mt_observation(cfg, st, e, gid, obs, sid);
    for (int i = 0; i < N; ++i) { /* reward envs/collective_assault_env.py:87-95 */
        double r = 0;
        if (alive[i] || jd[i]) {
            if (hit[i]) r += 1;
            if (was[i]) r = cfg->half_death_reward ? -0.5 : -1;
        }
        out->reward[(size_t)e * N + i] = r;
    }
    int cnt[MT_MAX_TEAMS] = {0}, teams_alive = 0;
    for (int i = 0; i < N; ++i) cnt[team[i]] += alive[i];
    for (int t = 0; t < cfg->n_team; ++t) teams_alive += cnt[t] > 0;
    const int done = teams_alive <= 1 || st->time_step[e] == cfg->max_episode_step - 1; /* :229-230 */
    out->done[e] = (uint8_t)done;
    if (done) { /* get_rank :219-223 */
        int all0 = 1;
        for (int t = 0; t < cfg->n_team; ++t) {
            int r = 0;
            for (int u = 0; u < cfg->n_team; ++u) r += cnt[u] > cnt[t];
            out->ranking[(size_t)e * cfg->n_team + t] = r;
            all0 &= r == 0;
        }
        if (all0) for (int t = 0; t < cfg->n_team; ++t) out->ranking[(size_t)e * cfg->n_team + t] = -1;
    }
    st->time_step[e] += 1;
    if (done && autoreset) { /* EnvAutoReset shm_env.py:51-66 */
        if (out->tobs) memcpy(out->tobs + (size_t)e * N * MT_LEN, obs, sizeof(float) * N * MT_LEN);
        mt_reset_world(cfg, st, e, gid);
        mt_terrain(cfg, st, e, N);
        mt_observation(cfg, st, e, gid, obs, sid);
    }
}

int mt_reset(const MtConfig *cfg, MtState *st, MtOut *out, int E, int64_t env_gid0) {
    const int N = mt_n(cfg);
    if (N > MAX_N || cfg->n_team > MT_MAX_TEAMS) return -1;
    for (int e = 0; e < E; ++e) {
        mt_reset_world(cfg, st, e, env_gid0 + e);
        mt_terrain(cfg, st, e, N);
        mt_observation(cfg, st, e, env_gid0 + e, out->obs + (size_t)e * N * MT_LEN,
                       out->slot_id ? out->slot_id + (size_t)e * N * MT_ROWS : NULL);
    }
    return 0;
}

int mt_step(const MtConfig *cfg, MtState *st, MtOut *out, int E, int64_t env_gid0, const double *actions,
            int autoreset) {
    const int N = mt_n(cfg);
    if (N > MAX_N || cfg->n_team > MT_MAX_TEAMS) return -1;
#pragma omp parallel for schedule(static)
    for (int e = 0; e < E; ++e) mt_env_step(cfg, st, out, e, env_gid0 + e, actions, autoreset);
    return 0;
}
