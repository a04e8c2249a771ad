/* ORACLE / TEST INFRASTRUCTURE -- see dca_oracle.h.  Plain C restatement of the reference's
 * numpy/Cython DCA path; every function cites the reference file:line it follows
 * (paths relative to /root/reference).  Compile WITHOUT -ffast-math and with
 * -ffp-contract=off: numpy evaluates every product and sum as a separate rounding. */
#include "dca_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define VIS_N 6
#define OBS_ROW 16
#define OBS_ROWS 12
#define OBS_LEN (OBS_ROW * OBS_ROWS)
#define MAX_N 512

static const double PI = 3.141592653589793; /* np.pi */

/* ---------------------------------------------------------------- keyed RNG (oracle/dca_rng.py) */
void orc_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]) {
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
enum { SITE_OBS = 1, SITE_RESET = 2 };
static const int JAM_W[6] = {0, 1, 3, 5, 7, 9};
static const int SHUF_W[6] = {-1, 2, 4, 6, 8, 10};

static void obs_words(const OrcConfig *cfg, int64_t gid, uint32_t counter, int guard, int lst, uint32_t w[12]) {
    for (int b = 0; b < 3; ++b)
        orc_philox((uint32_t)gid, counter, (uint32_t)(guard * 2 + lst), SITE_OBS | (b << 8),
                   (uint32_t)cfg->seed, (uint32_t)(cfg->seed >> 32), w + 4 * b);
}

/* Python-style floored modulo of floats (numpy npy_divmod / CPython float_rem):
 * used by reg_angle core.py:284-285 and reg_rad cython_func.pyx:16-17. */
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

/* np.logaddexp(0, y)  (numpy/_core/src/npymath/npy_math_internal.h.src npy_logaddexp) */
static double logaddexp0(double y) {
    const double x = 0.0;
    if (x == y) return x + 0.693147180559945309417232121458176568;
    double tmp = x - y;
    if (tmp > 0) return x + log1p(exp(-tmp));
    else if (tmp <= 0) return y + log1p(exp(tmp));
    return tmp;
}

/* ---------------------------------------------------------------- terrain
 * World.set_terrain_adv core.py:224-234 + get_terrain core.py:306-314 (alive agents only). */
static void env_terrain(const OrcConfig *cfg, OrcState *st, int e) {
    const int N = cfg->num_guards + cfg->num_attackers;
    const double th = st->init_theta[e];
    const double c = cos(th), s = sin(th);
    for (int i = 0; i < N; ++i) {
        const int k = e * N + i;
        if (!st->alive[k]) continue;
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

/* ---------------------------------------------------------------- reset
 * collective_assaultEnvV1.reset_world collective_assault_env.py:55-136, draws keyed per dca_rng.py. */
static void env_reset_world(const OrcConfig *cfg, OrcState *st, int e, int64_t gid) {
    const int G = cfg->num_guards, A = cfg->num_attackers, N = G + A;
    const uint32_t k0 = (uint32_t)cfg->seed, k1 = (uint32_t)(cfg->seed >> 32);
    st->episode[e] += 1;
    const uint32_t ep = st->episode[e];
    uint32_t w[4];
    orc_philox((uint32_t)gid, ep, 0, SITE_RESET, k0, k1, w);
    st->time_step[e] = 0;
    st->n_alive_attackers[e] = A; /* :59-60 -- not corrected for excluded attackers (faithful) */
    st->n_alive_guards[e] = G;
    const double theta = (2 * (w[0] * 0x1p-32) - 1) * PI; /* :64 */
    st->init_theta[e] = theta;
    const double ct = cos(theta), sn = sin(theta);

    /* attacker fluctuation :67-77 : n_attacker = randint(max-fluct, max+1); exclude (max-n_attacker) of them */
    int n_excl = 0;
    uint32_t key[MAX_N];
    if (cfg->enable_attacker_fluctuation) {
        const int low = A - cfg->attacker_fluctuation, high = A + 1;
        const int n_att = low + (int)(((uint64_t)w[1] * (uint64_t)(high - low)) >> 32);
        n_excl = A - n_att;
    }
    for (int i = 0; i < N; ++i) {
        const int k = e * N + i;
        uint32_t a[4];
        orc_philox((uint32_t)gid, ep, (uint32_t)(1 + i), SITE_RESET, k0, k1, a);
        key[i] = a[3];
        const int attacker = i >= G;
        const double u = a[0] * 0x1p-32;
        const double u1 = ((double)a[1] + 1.0) * 0x1p-32, u2 = a[2] * 0x1p-32;
        const double r = sqrt(-2.0 * log(u1));
        const double n0 = r * cos(2.0 * PI * u2), n1 = r * sin(2.0 * PI * u2);
        st->alive[k] = 1;
        st->vel[2 * k] = 0; st->vel[2 * k + 1] = 0;
        st->atk_rad[k] = attacker ? (theta + PI) + (u - 0.5) / 12 : theta + (u - 0.5) / 12; /* :85 */
        const double xMid = -5.0 / 2 + 5.0 / 2, yMid = xMid, yMin = -5.0, yMax = 5.0;
        double x_, y_;
        if (attacker) { /* :92-96 */
            x_ = xMid + cfg->init_distance / 2;
            y_ = (yMax - yMin) / A * (i - G + 0.5) + yMin;
        } else { /* :115-119 */
            x_ = xMid - cfg->init_distance / 2;
            y_ = (yMax - yMin) / G * (i + 0.5) + yMin;
        }
        double px = x_ + (n0 - 0.5) / 10, py = y_ + (n1 - 0.5) / 10; /* :97 / :120 */
        if ((attacker ? A : G) > 48) {                               /* :99-106 / :122-129 */
            double ratio = 1;
            if (i % 3 == 0) ratio = 0.5;
            if (i % 3 == 1) ratio = 0.75;
            px = xMid + (px - xMid) * ratio;
            py = yMid + (py - yMid) * ratio;
        }
        /* np.dot(pos, rotate.T) :108 / :131 */
        st->pos[2 * k] = px * ct + py * (-sn);
        st->pos[2 * k + 1] = px * sn + py * ct;
        if (attacker) st->terrain[k] = NAN; /* :113 */
        st->num_hit[k] = 0; st->num_was_hit[k] = 0; st->hit[k] = 0; st->was_hit[k] = 0; /* :133-136 */
        /* justDied is NOT touched by reset_world (faithful) */
    }
    if (n_excl > 0) { /* keyed stand-in for np.random.choice(..., replace=False): smallest (key, index) */
        for (int a = 0; a < A; ++a) {
            int rank = 0;
            for (int b = 0; b < A; ++b)
                if (key[G + b] < key[G + a] || (key[G + b] == key[G + a] && b < a)) rank++;
            if (rank < n_excl) st->alive[e * N + G + a] = 0; /* :109-111 */
        }
    }
}

/* ---------------------------------------------------------------- observation
 * collective_assaultEnvV1.observation collective_assault_env.py:228-359 for all guards,
 * reward collective_assault_env.py:153-179 is evaluated by the caller. */
static float f32_dist(float xi, float yi, float xj, float yj) {
    /* distance_matrix :361-368: At-A in f32, np.linalg.norm = sqrt(add.reduce(x*x)) in f32 */
    volatile float dx = xj - xi, dy = yj - yi;
    volatile float sx = dx * dx, sy = dy * dy;
    volatile float s = sx + sy;
    return sqrtf(s);
}

/* stable top-k (k=6) of cand distances: ascending distance, ties -> lower index.  +inf = dead. */
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

static void env_observation(const OrcConfig *cfg, OrcState *st, int e, int64_t gid, const uint32_t *rng_words,
                            float *obs /*[G,192]*/, int16_t *slot_id /*[G,12] or NULL*/) {
    const int G = cfg->num_guards, A = cfg->num_attackers, N = G + A;
    st->obs_counter[e] += 1;
    const uint32_t counter = st->obs_counter[e];
    /* entity table :237-276 : [alive,x,y,atk_rad,vx,vy,iden,terrain, bit0..7] cast to float32 */
    float tab[MAX_N][OBS_ROW];
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
        /* get_binary_array :207-215: n_int is divided as float then cast to int8 each round */
        int v = i;
        for (int b = 0; b < 8; ++b) {
            tab[i][8 + b] = (float)((((v % 2) + 2) % 2) == 1); /* numpy % is floored */
            v = (int)(int8_t)(int32_t)(v / 2.0);                 /* .astype(np.int8) wraps */
        }
    }
    const float obs_range = 2.0f;
    for (int g = 0; g < G; ++g) {
        float *o = obs + (size_t)g * OBS_LEN;
        int16_t *sid = slot_id ? slot_id + g * OBS_ROWS : NULL;
        memset(o, 0, sizeof(float) * OBS_LEN);
        if (sid) for (int r = 0; r < OBS_ROWS; ++r) sid[r] = -1;
        if (!st->alive[e * N + g]) continue; /* :313-316 -- no RNG draws */
        float dh[MAX_N], df[MAX_N];
        for (int a = 0; a < A; ++a) /* :286-297 dead rows/cols = +inf */
            dh[a] = st->alive[e * N + G + a] ? f32_dist(tab[g][1], tab[g][2], tab[G + a][1], tab[G + a][2]) : INFINITY;
        for (int f = 0; f < G; ++f)
            df[f] = st->alive[e * N + f] ? f32_dist(tab[g][1], tab[g][2], tab[f][1], tab[f][2]) : INFINITY;
        int hs[VIS_N], fs[VIS_N];
        top6(dh, A, hs); /* :308-309 argsort()[:6] */
        top6(df, G, fs);
        for (int lst = 0; lst < 2; ++lst) {
            /* visible list in sorted order :320-322 (hostile, d<=2.0) / :331-333 (friend, d<=3.0) */
            int vis[VIS_N], n = 0;
            const int *srt = lst == 0 ? hs : fs;
            const float *d = lst == 0 ? dh : df;
            const float lim = lst == 0 ? obs_range : obs_range * 1.5f;
            for (int k = 0; k < VIS_N; ++k)
                if (srt[k] >= 0 && d[srt[k]] <= lim) vis[n++] = srt[k]; /* finite => alive */
            int self_id = -1, first = 0;
            if (lst == 1) { self_id = vis[0]; first = 1; } /* :334-335 self = first visible friend */
            int *src = vis + first;
            int m = n - first;
            /* item_random_mv :217-226 with keyed draws */
            uint32_t w[12];
            if (rng_words) memcpy(w, rng_words + (((size_t)e * G + g) * 2 + lst) * 12, sizeof(w));
            else obs_words(cfg, gid, counter, g, lst, w);
            for (int i = m - 1; i >= 1; --i) {
                int j = (int)(((uint64_t)w[SHUF_W[i]] * (uint64_t)(i + 1)) >> 32);
                int t = src[i]; src[i] = src[j]; src[j] = t;
            }
            int n_mv = 0;
            for (int k = 0; k < m; ++k) n_mv += (w[JAM_W[k]] * 0x1p-32) < cfg->random_jam_prob;
            const int keep = m - n_mv;
            /* rows: friend block first (rows 0..5, row 0 = self), hostile block rows 6..11 :344 */
            int row = lst == 1 ? 0 : VIS_N;
            const int base = lst == 1 ? 0 : G;
            if (lst == 1) {
                memcpy(o + row * OBS_ROW, tab[self_id], sizeof(float) * OBS_ROW);
                if (sid) sid[row] = (int16_t)self_id;
                row++;
            }
            for (int k = 0; k < keep; ++k, ++row) {
                memcpy(o + row * OBS_ROW, tab[base + src[k]], sizeof(float) * OBS_ROW);
                if (sid) sid[row] = (int16_t)(base + src[k]);
            }
        }
    }
}

/* ---------------------------------------------------------------- laser
 * laser_hit_improve3 cython_func.pyx:22-46 */
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

/* ---------------------------------------------------------------- one env step
 * collective_assaultGlobalEnv.step collective_assault_parallel_run.py:192-246 (without the planner's
 * formation part, whose output att_force is injected) + EnvAutoReset.step UTIL/shm_env.py:33-74. */
static void env_step(const OrcConfig *cfg, OrcState *st, OrcOut *out, int e, int64_t gid, const double *actions,
                     const double *att_force, const uint32_t *rng_words, int autoreset) {
    const int G = cfg->num_guards, A = cfg->num_attackers, N = G + A;
    const double *act_in = actions + (size_t)e * G;
    for (int g = 0; g < G; ++g)
        if (isnan(act_in[g])) return; /* skip command shm_env.py:37-41: cached outputs stay */

    double ax[MAX_N], ay[MAX_N], ar[MAX_N];
    uint8_t can_fire[MAX_N];
    /* _set_action :326-374 (every guard, dead ones included) */
    for (int g = 0; g < G; ++g) {
        double a0 = 0, a1 = 0, a2 = 0;
        const double a = act_in[g];
        if (a == 1) a0 = +1.0;
        if (a == 2) a0 = -1.0;
        if (a == 3) a1 = +1.0;
        if (a == 4) a1 = -1.0;
        if (a == 5) a2 = +cfg->agent_max_rot;
        if (a == 6) a2 = -cfg->agent_max_rot;
        ax[g] = a0 * cfg->agent_acc; ay[g] = a1 * cfg->agent_acc; ar[g] = a2;
        can_fire[g] = 1;
    }
    env_terrain(cfg, st, e); /* :202 */

    /* attackers_policy_1: init cheat_script_ai.py:31-34, force :183-184 (injected), aim :196-218 */
    for (int a = 0; a < A; ++a) {
        const int i = G + a, k = e * N + i;
        ax[i] = ay[i] = ar[i] = 0;
        can_fire[i] = 0;
        if (out->nearest) out->nearest[(size_t)e * A + a] = -1;
        if (!st->alive[k]) continue;
        ax[i] = att_force[((size_t)e * A + a) * 2];
        ay[i] = att_force[((size_t)e * A + a) * 2 + 1];
        /* argmin over the cached float32 distance matrix (dead guards = +inf), first minimum wins */
        int best = 0;
        float bd = INFINITY;
        const float xi = (float)st->pos[2 * k], yi = (float)st->pos[2 * k + 1];
        for (int g = 0; g < G; ++g) {
            const int kg = e * N + g;
            float d = st->alive[kg] ? f32_dist(xi, yi, (float)st->pos[2 * kg], (float)st->pos[2 * kg + 1]) : INFINITY;
            if (d < bd) { bd = d; best = g; }
        }
        if (out->nearest) out->nearest[(size_t)e * A + a] = best;
        const int kt = e * N + best;
        double th = atan2(st->pos[2 * kt + 1] - st->pos[2 * k + 1], st->pos[2 * kt] - st->pos[2 * k]);
        if (th < 0) th += 2 * PI;
        st->atk_rad[k] = th;
        can_fire[i] = cfg->disable_red_function ? 0 : 1;
    }

    /* apply_laser_effect core.py:237-279 */
    for (int i = 0; i < N; ++i)
        if (st->alive[e * N + i]) { st->hit[e * N + i] = 0; st->was_hit[e * N + i] = 0; }
    for (int i = 0; i < N; ++i) {
        const int ki = e * N + i;
        if (!st->alive[ki] || !can_fire[i]) continue;
        for (int j = 0; j < N; ++j) {
            const int kj = e * N + j;
            if (!st->alive[kj] || ((i >= G) == (j >= G))) continue;
            if (laser_hit(st->pos[2 * ki], st->pos[2 * ki + 1], st->pos[2 * kj], st->pos[2 * kj + 1],
                          cfg->agent_shoot_base_radius * st->terrain[ki], cfg->agent_shoot_win, st->atk_rad[ki])) {
                st->hit[ki] = 1; st->num_hit[ki] += 1;
                st->was_hit[kj] = 1; st->num_was_hit[kj] += 1;
            }
        }
    }
    for (int i = 0; i < N; ++i)
        if (!st->alive[e * N + i]) st->just_died[e * N + i] = 0;
    for (int i = 0; i < N; ++i) {
        const int k = e * N + i;
        if (st->alive[k] && st->was_hit[k]) {
            st->alive[k] = 0; st->just_died[k] = 1;
            if (i >= G) st->n_alive_attackers[e] -= 1; else st->n_alive_guards[e] -= 1;
        }
    }

    /* forces core.py:210-222,339-353 and integrate_state :282-302, surviving agents only */
    const double wall = cfg->size, esz = 0.05, kk = 1e-10, dt = 0.1;
    for (int i = 0; i < N; ++i) {
        const int k = e * N + i;
        if (!st->alive[k]) continue;
        const double x = st->pos[2 * k], y = st->pos[2 * k + 1];
        const double d0 = x - (esz + (-wall)), d1 = wall - x - esz, d2 = y - esz - (-wall), d3 = wall - y - esz;
        double fx = 0, fy = 0;
        if (!(d0 > 0 && d1 > 0 && d2 > 0 && d3 > 0)) {
            const double f1 = 100.0 * (logaddexp0(-d0 / kk) * kk), f2 = 100.0 * (logaddexp0(-d1 / kk) * kk);
            const double f3 = 100.0 * (logaddexp0(-d2 / kk) * kk), f4 = 100.0 * (logaddexp0(-d3 / kk) * kk);
            fx = f1 - f2; fy = f3 - f4;
        }
        const double Fx = fx + ax[i], Fy = fy + ay[i];
        double vx = st->vel[2 * k] * 0.75, vy = st->vel[2 * k + 1] * 0.75;
        vx += (Fx / 1.0) * dt; vy += (Fy / 1.0) * dt;
        const double speed = sqrt(vx * vx + vy * vy);
        if (speed > cfg->agent_max_speed) {
            vx = vx / speed * cfg->agent_max_speed;
            vy = vy / speed * cfg->agent_max_speed;
        }
        st->atk_rad[k] = reg_angle(st->atk_rad[k] + ar[i]);
        st->vel[2 * k] = vx; st->vel[2 * k + 1] = vy;
        st->pos[2 * k] = x + vx * dt; st->pos[2 * k + 1] = y + vy * dt;
    }

    /* observation + reward :213-218 */
    float *obs = out->obs + (size_t)e * G * OBS_LEN;
    int16_t *sid = out->slot_id ? out->slot_id + (size_t)e * G * OBS_ROWS : NULL;
    env_observation(cfg, st, e, gid, rng_words, obs, sid);
    double *rew = out->reward + (size_t)e * G;
    for (int g = 0; g < G; ++g) { /* reward/guard_reward collective_assault_env.py:153-179 */
        const int k = e * N + g;
        double r = 0;
        if (st->alive[k] || st->just_died[k]) {
            if (st->hit[k]) r += 1;
            if (st->was_hit[k]) r += cfg->half_death_reward ? -0.5 : -1;
        }
        rew[g] = r;
    }
    /* _get_done :286-316 */
    int done = 0, win = 0;
    const int nA = st->n_alive_attackers[e], nG = st->n_alive_guards[e];
    if (nA == 0) { done = 1; win = 1; }
    else if (nG == 0) { done = 1; win = 0; }
    else if (st->time_step[e] == cfg->max_episode_step - 1) { done = 1; win = (nA * G <= nG * A); }
    if (cfg->reward_debug && done && win) /* :228-231 */
        for (int g = 0; g < G; ++g) rew[g] += cfg->reward_debug_value;
    st->time_step[e] += 1; /* :238 */
    out->done[e] = (uint8_t)done;
    if (out->win) out->win[e] = (uint8_t)win;

    if (done && autoreset) { /* shm_env.py:51-66 */
        if (out->tobs) memcpy(out->tobs + (size_t)e * G * OBS_LEN, obs, sizeof(float) * G * OBS_LEN);
        env_reset_world(cfg, st, e, gid);
        env_terrain(cfg, st, e);
        env_observation(cfg, st, e, gid, NULL, obs, sid);
    }
}

int orc_reset(const OrcConfig *cfg, OrcState *st, OrcOut *out, int E, int64_t env_gid0, int nthreads) {
    const int G = cfg->num_guards;
    if (G + cfg->num_attackers > MAX_N || G < VIS_N || cfg->num_attackers < VIS_N) return -1;
#pragma omp parallel for num_threads(nthreads > 0 ? nthreads : 1) schedule(dynamic, 8)
    for (int e = 0; e < E; ++e) {
        env_reset_world(cfg, st, e, env_gid0 + e);
        env_terrain(cfg, st, e); /* collective_assault_parallel_run.py:258 */
        env_observation(cfg, st, e, env_gid0 + e, NULL, out->obs + (size_t)e * G * OBS_LEN,
                        out->slot_id ? out->slot_id + (size_t)e * G * OBS_ROWS : NULL);
    }
    return 0;
}

int orc_step(const OrcConfig *cfg, OrcState *st, OrcOut *out, int E, int64_t env_gid0, const double *actions,
             const double *att_force, const uint32_t *rng_words, int autoreset, int nthreads) {
    if (cfg->num_guards + cfg->num_attackers > MAX_N) return -1;
#pragma omp parallel for num_threads(nthreads > 0 ? nthreads : 1) schedule(dynamic, 8)
    for (int e = 0; e < E; ++e) env_step(cfg, st, out, e, env_gid0 + e, actions, att_force, rng_words, autoreset);
    return 0;
}

int orc_plan_centroid(const OrcConfig *cfg, const OrcState *st, int E, double *att_force, int nthreads) {
    const int G = cfg->num_guards, A = cfg->num_attackers, N = G + A;
#pragma omp parallel for num_threads(nthreads > 0 ? nthreads : 1) schedule(static)
    for (int e = 0; e < E; ++e) {
        double sx = 0, sy = 0;
        int n = 0;
        for (int g = 0; g < G; ++g)
            if (st->alive[e * N + g]) { sx += st->pos[2 * (e * N + g)]; sy += st->pos[2 * (e * N + g) + 1]; n++; }
        const double cx = n ? sx / n : 0.0, cy = n ? sy / n : 0.0;
        for (int a = 0; a < A; ++a) {
            const int k = e * N + G + a;
            att_force[((size_t)e * A + a) * 2] = st->alive[k] ? cx - st->pos[2 * k] : 0.0;
            att_force[((size_t)e * A + a) * 2 + 1] = st->alive[k] ? cy - st->pos[2 * k + 1] : 0.0;
        }
    }
    return 0;
}

/* ---------------------------------------------------------------- GAE / return
 * ALGORITHM/conc/trajectory.py:60-64 (reward_push_forward), :75-86 (threat), :92-117
 * (gae_finalize_return); ALGORITHM/common/traj_gae.py:54-133 is the same arithmetic with
 * SAFE_LIMIT 8.  One trajectory per env column, length traj_len[e] (<= T); rows past it untouched.
 * dtype walk (numpy): reward f64, value f32, gamma*value[t+1] evaluated in f32 (python scalar is
 * weak), delta/gae in f64, return stored f32. */
int orc_gae(const double *reward, const float *value, const uint8_t *dead, const int32_t *traj_len, float *ret,
            float *threat, double *reward_pushed, int T, int E, int G, double gamma, double tau, double gamma_fwd,
            int threat_limit) {
    const size_t S = (size_t)E * G;
    for (int e = 0; e < E; ++e) {
        const int L = traj_len ? traj_len[e] : T;
        if (L <= 0) continue;
        for (int g = 0; g < G; ++g) {
            const size_t c = (size_t)e * G + g;
            /* reward push: r[i-1] += r[i]*dead[i]; r[i] *= !dead[i], i = L-1 .. 1; with gamma_in_reward_forwarding
             * (traj_gae.py:55-60, conc_4hist/trajectory.py:58-63) the pushed amount is r[i]*gamma_fwd (1.0 = off) */
            double carry_r = 0; /* r[i] after receiving pushes from i+1 */
            double gae = 0;
            double th = -1;
            for (int i = L - 1; i >= 0; --i) {
                const size_t k = (size_t)i * S + c;
                double r = reward[k] + carry_r;
                const int d = dead[k] != 0;
                if (i > 0) {
                    carry_r = (r * gamma_fwd) * (double)d; /* np.where(dead_mask[i], reward[i]*gamma, 0) */
                    r = r * (double)(!d);
                }
                if (reward_pushed) reward_pushed[k] = r;
                /* threat[t] = -1 + sum_{i>=t} term(i), term(L-1) = !dead[L-1], term(i) = !(dead[i+1]&dead[i]) */
                if (i == L - 1) th += (double)(!d);
                else th += (double)(!(dead[k + S] && d));
                double tc = th < -1 ? -1 : (th > threat_limit ? threat_limit : th);
                if (threat) threat[k] = (float)tc;
                const double v = (double)value[k];
                double delta;
                if (i == L - 1) {
                    delta = r + gamma * 0 - v;
                    gae = delta;
                } else {
                    volatile float gv = (float)gamma * value[k + S];
                    delta = r + (double)gv - v;
                    gae = delta + gamma * tau * gae;
                }
                ret[k] = (float)(gae + v);
            }
        }
    }
    return 0;
}
