// dca_multiteam on the device: the sibling env of DCA with N teams, every agent policy-controlled,
// a per-hit success draw, a shrinking kill circle and a team ranking instead of win/lose.
//   MISSION/dca_multiteam/collective_assault_parallel_run.py:156-323  step / reset / _get_done / get_rank / _set_action
//   MISSION/dca_multiteam/core.py:100-240                             World.step, laser, integrate, kill circle, terrain
//   MISSION/dca_multiteam/envs/collective_assault_env.py:42-226      reset_world, reward, observation (12 x 17 floats)
//   UTIL/shm_env.py:33-97                                            EnvAutoReset (skip mask, terminal obs, reset obs)
// One CTA per env, state in SoA arrays [E,N] (loaded into shared memory per step), same keyed Philox stream as DCA plus
// one site for the laser draws: hit draw of (shooter i, target j) = word 0 of Philox(gid, obs_counter, i*N+j, site 4).
// Equal float32 distances are ordered by index (the reference's np.argsort leaves that order to its sort kernel).
#pragma once
#include "dca_device.cuh"

namespace dca {

constexpr int kMtRow = DCA_MT_OBS_ROW;          // 17 floats per entity record
constexpr int kMtLen = kObsRows * kMtRow;       // 204 floats per agent
constexpr int kMtThreads = 128;
constexpr int kMtPairCap = 512;                 // queued in-range laser pairs per env (overflow is tested in place)
enum : uint32_t { SITE_MT_LASER = 4 };
constexpr double kCosHalfWin2 = 0.85355339059327376220;  // cos^2(pi/8) = (2 + sqrt 2) / 4: half of shootWin = pi/4
#define MT_PROF(P, e, k)                                                                      \
    do {                                                                                      \
        if ((P).prof != nullptr && threadIdx.x == 0) (P).prof[(size_t)(e) * 16 + (k)] = clock64(); \
    } while (0)
enum : uint8_t { MT_ALIVE = 1, MT_JUST_DIED = 2, MT_HIT = 4, MT_WAS_HIT = 8, MT_BSA = 16 };

struct MtParams {
    DcaMtConfig cfg;
    int E, N;
    DcaMtState st;     // library-owned SoA state
    DcaMtBuffers out;  // caller-owned outputs
    unsigned long long jam_thresh;
    float smax[2];  // largest float32 s with sqrt_rn(s) <= 2.0 / 3.0: "s <= smax" is the reference's "d <= range"
    int idx_bits;   // key bits that hold the list position
    unsigned long long *prof;  // diagnostics: [E,16] clock64() of thread 0 at phase boundaries (NULL = off)
};

struct MtSmem {
    double *px, *py, *vx, *vy, *atk, *ter;
    double *ux, *uy;         // unit heading (cos, sin of atk_rad) of the agents that can shoot this step
    float *xs, *ys, *table;  // table: [(N+1)][17], record N = zeros
    int *hitf;               // per agent: bit 0 = hit someone, bit 1 = was hit (this step)
    int *pairq;              // [kMtPairCap] shooter * N + target
    float *bbox;             // [8][4] x0 x1 y0 y1 of each team's alive agents
    int16_t *sid;            // [N,12]
    uint8_t *flags, *team, *tid;
    int *misc;               // [0..7] alive count per team, [16] queued laser pairs, [20..28] first uid of each team
};

__host__ __device__ inline size_t mt_smem_bytes(int N) {
    size_t b = 8 * sizeof(double) * N;
    b += 2 * sizeof(float) * N + sizeof(float) * (size_t)(N + 1) * kMtRow;
    b += sizeof(int) * N + sizeof(int) * kMtPairCap + sizeof(int16_t) * (size_t)N * kObsRows + 3 * (size_t)N + sizeof(int) * 32 + sizeof(float) * 4 * DCA_MT_MAX_TEAMS;
    return (b + 63) & ~(size_t)15;
}

__device__ __forceinline__ MtSmem mt_carve(unsigned char *p, int N) {
    MtSmem s;
    s.px = (double *)p; p += sizeof(double) * N;
    s.py = (double *)p; p += sizeof(double) * N;
    s.vx = (double *)p; p += sizeof(double) * N;
    s.vy = (double *)p; p += sizeof(double) * N;
    s.atk = (double *)p; p += sizeof(double) * N;
    s.ter = (double *)p; p += sizeof(double) * N;
    s.ux = (double *)p; p += sizeof(double) * N;
    s.uy = (double *)p; p += sizeof(double) * N;
    s.xs = (float *)p; p += sizeof(float) * N;
    s.ys = (float *)p; p += sizeof(float) * N;
    s.table = (float *)p; p += sizeof(float) * (size_t)(N + 1) * kMtRow;
    s.hitf = (int *)p; p += sizeof(int) * N;
    s.misc = (int *)p; p += sizeof(int) * 32;
    s.pairq = (int *)p; p += sizeof(int) * kMtPairCap;
    s.bbox = (float *)p; p += sizeof(float) * 4 * DCA_MT_MAX_TEAMS;
    s.sid = (int16_t *)p; p += sizeof(int16_t) * (size_t)N * kObsRows;
    s.flags = p; p += N;
    s.team = p; p += N;
    s.tid = p;
    return s;
}

// entity record envs/collective_assault_env.py:116-128: [alive, x, y, atk_rad, vx, vy, iden, terrain, bi_hot(tid) x8, team]
__device__ __forceinline__ void mt_build_row(const MtSmem &s, int i) {
    float *r = s.table + (size_t)i * kMtRow;
    const bool alive = s.flags[i] & MT_ALIVE;
    const float x = (float)s.px[i], y = (float)s.py[i];
    r[0] = alive ? 1.f : 0.f; r[1] = x; r[2] = y; r[3] = (float)s.atk[i];
    r[4] = (float)s.vx[i]; r[5] = (float)s.vy[i]; r[6] = (float)i; r[7] = (float)s.ter[i];
    int v = s.tid[i];  // get_binary_n_rows: n_int / 2 as float, then .astype(np.int8)
#pragma unroll
    for (int b = 0; b < 8; ++b) {
        r[8 + b] = (float)((((v % 2) + 2) % 2) == 1);
        v = (int)(signed char)(int)((double)v / 2.0);
    }
    r[16] = (float)s.team[i];
    s.xs[i] = alive ? x : CUDART_INF_F;
    s.ys[i] = alive ? y : CUDART_INF_F;
}

// float32 bound of a shooter's squared laser reach: |coord| <= 64 and r < 1 keep the float32 squared distance within
// 1e-4 of the exact one, so a pair beyond r^2 * 1.01 + 1e-3 cannot pass the fp64 test (+inf = no bound)
__device__ __forceinline__ float mt_reach2f(double x, double y, double terrain) {
    const double radius = 0.4 * terrain, r2 = radius * radius;
    const bool ok = fabs(x) <= 64.0 && fabs(y) <= 64.0 && r2 < 1.0;
    return ok ? __fadd_ru(__fmul_ru((float)r2, 1.01f), 1.0e-3f) : CUDART_INF_F;
}

// exact laser test of the ordered pair p = shooter * N + target (already inside the shooter's radius):
// laser_hit_improve3 (dca_multiteam/cython_func.pyx:22-53), then the success draw core.py:178-190
// (plain pointers / scalars only: a struct passed by reference to an out-of-line function would live in local memory)
__device__ __noinline__ void mt_laser_pair(const double *px, const double *py, const double *atk, const double *ter,
                                           int *hitf, int N, uint32_t gid, uint32_t counter, uint32_t k0, uint32_t k1,
                                           int entry) {
    const int p = entry & 0x3fffffff;
    const int i = p / N, j = p - i * N;
    if (entry >> 30) {  // the cone test was too close to call: the reference's own arithmetic decides
        const double win = kPi / 4;
        const double dx = px[j] - px[i], dy = py[j] - py[i];
        const double dir = atk[i];
        const double tt = atan2(dy, dx);
        const double d1 = fabs(reg_angle((dir + win / 2) - tt)), d2 = fabs(reg_angle((dir - win / 2) - tt));
        if (!(d1 <= win && d2 <= win)) return;
    }
    const double adv = fmax(fmin(ter[i] - ter[j], 0.1), -0.1);
    const double prob = 0.8 + adv;
    const uint4 w = philox4x32_10(gid, counter, (uint32_t)p, SITE_MT_LASER, k0, k1);
    if (((double)w.x * 0x1p-32) < prob) { atomicOr(&hitf[i], 1); atomicOr(&hitf[j], 2); }
}

// Exact (rare) path of one observation task: the 6 nearest in-range candidates by (sqrt_rn distance, list position),
// taken when two keys of the fast path are too close to call.
__device__ __noinline__ unsigned long long mt_top6_exact(const float *xs, const float *ys, int g, int lst, int own0,
                                                          int own1, int N, float lim, int *n_out) {
    float bd[kVisN];
    int bi[kVisN];
#pragma unroll
    for (int k = 0; k < kVisN; ++k) { bd[k] = CUDART_INF_F; bi[k] = -1; }
    const float gx = xs[g], gy = ys[g];
    const int j0 = lst == 1 ? own0 : 0, j1 = lst == 1 ? own1 : N;
    for (int j = j0; j < j1; ++j) {
        if (lst == 0 && j >= own0 && j < own1) { j = own1 - 1; continue; }
        const float d = f32_dist(gx, gy, xs[j], ys[j]);  // dead = +inf
        if (d <= lim) {
            bool p[kVisN];
#pragma unroll
            for (int k = 0; k < kVisN; ++k) p[k] = d < bd[k];
#pragma unroll
            for (int k = kVisN - 1; k >= 1; --k) {
                bd[k] = p[k] ? (p[k - 1] ? bd[k - 1] : d) : bd[k];
                bi[k] = p[k] ? (p[k - 1] ? bi[k - 1] : j) : bi[k];
            }
            bd[0] = p[0] ? d : bd[0];
            bi[0] = p[0] ? j : bi[0];
        }
    }
    unsigned long long vis = 0;
    int n = 0;
#pragma unroll
    for (int k = 0; k < kVisN; ++k)
        if (bi[k] >= 0) { vis |= (unsigned long long)bi[k] << (8 * n); ++n; }
    *n_out = n;
    return vis;
}

// observation of every agent (envs/collective_assault_env.py:131-226) + coalesced gather into dest [N,12,17]
__device__ __forceinline__ void mt_obs_pass(const MtSmem &s, const MtParams &P, int e, float *dest, uint32_t counter) {
    const int tid = threadIdx.x, N = P.N;
    const uint32_t k0 = (uint32_t)P.cfg.seed, k1 = (uint32_t)(P.cfg.seed >> 32);
    const uint32_t gid = (uint32_t)(P.cfg.env_gid0 + e);
    if (dest != nullptr) {
        for (int task = tid; task < 2 * N; task += kMtThreads) {
            const int g = task >> 1, lst = task & 1;  // 0 = hostile (rows 6..11), 1 = friend (rows 0..5, row 0 = self)
            int16_t *sid = s.sid + g * kObsRows + (lst == 0 ? kVisN : 0);
#pragma unroll
            for (int k = 0; k < kVisN; ++k) sid[k] = -1;
            if (!(s.flags[g] & MT_ALIVE)) continue;  // :176-179 zeros, no draws
            const int tm = s.team[g];
            const int own0 = s.misc[20 + tm], own1 = s.misc[21 + tm];
            const float lim = lst == 0 ? 2.0f : 3.0f;  // obs_range, obs_range * 1.5 (P.smax = the same limits, squared)
            const float gx = s.xs[g], gy = s.ys[g];
            // candidates in uid order (ally list = own team, foe list = the other teams in team order, :162-170);
            // argsort()[:6] + range mask == the 6 nearest in range, ties by list position.
            // Fast path (as DCA's knn_task): 7-slot min/max network on 32-bit keys = float32 squared-distance bits with
            // the low idx_bits replaced by the list position; sqrt_rn is monotone, so key order == (distance, position)
            // order unless two of the kept keys are within one key step -- then the exact serial path decides.
            const int j0 = lst == 1 ? own0 : 0, j1 = lst == 1 ? own1 : N;
            const int nown = own1 - own0;
            const float q0 = P.smax[lst];
            const uint32_t im = (1u << P.idx_bits) - 1u;
            uint32_t b[kVisN + 1];
#pragma unroll
            for (int k = 0; k <= kVisN; ++k) b[k] = 0x80000000u + ((uint32_t)k << 8);  // empty: distinct, ascending
            for (int j = j0; j < j1; ++j) {  // (unrolling this loop costs registers -> resident CTAs: slower)
                if (lst == 0 && j >= own0 && j < own1) { j = own1 - 1; continue; }
                const float sq = f32_sqdist(gx, gy, s.xs[j], s.ys[j]);  // dead = +inf
                const uint32_t pos = (uint32_t)(lst == 1 ? j - own0 : (j < own0 ? j : j - nown));
                uint32_t key = sq <= q0 ? ((__float_as_uint(sq) & ~im) | pos) : 0xFFFFFFFFu;
#pragma unroll
                for (int k = 0; k <= kVisN; ++k) {
                    const uint32_t lo = min(b[k], key);
                    key = max(b[k], key);
                    b[k] = lo;
                }
            }
            bool close_call = false;
#pragma unroll
            for (int k = 0; k < kVisN; ++k)
                close_call |= ((int32_t)b[k + 1] >= 0) && ((b[k + 1] >> P.idx_bits) - (b[k] >> P.idx_bits) <= 1u);
            // visible uids packed one byte each, nearest first
            unsigned long long vis = 0;
            int n = 0;
            if (close_call) {
                vis = mt_top6_exact(s.xs, s.ys, g, lst, own0, own1, N, lim, &n);
            } else {
#pragma unroll
                for (int k = 0; k < kVisN; ++k) {
                    if ((int32_t)b[k] >= 0) {
                        const int pos = (int)(b[k] & im);
                        const int uid = lst == 1 ? own0 + pos : (pos < own0 ? pos : pos + nown);
                        vis |= (unsigned long long)uid << (8 * n);
                        ++n;
                    }
                }
            }
            int row = 0;
            if (lst == 1) {  // self = first visible friend (:207-208), excluded from shuffle / jamming
                sid[row++] = (int16_t)(vis & 0xff);
                vis >>= 8; --n;
            }
            if (n > 0) {  // item_random_mv :103-113 with the keyed words (shuffle i = n-1..1, then jam draws)
                uint32_t w[DCA_RNG_WORDS];
                const int nblk = n <= 2 ? 1 : (n <= 4 ? 2 : 3);
#pragma unroll
                for (int bb = 0; bb < 3; ++bb) {
                    if (bb < nblk) {
                        const uint4 q = philox4x32_10(gid, counter, (uint32_t)(g * 2 + lst), SITE_OBS | (bb << 8), k0, k1);
                        w[4 * bb] = q.x; w[4 * bb + 1] = q.y; w[4 * bb + 2] = q.z; w[4 * bb + 3] = q.w;
                    } else {
                        w[4 * bb] = w[4 * bb + 1] = w[4 * bb + 2] = w[4 * bb + 3] = 0;
                    }
                }
#pragma unroll
                for (int i = kVisN - 1; i >= 1; --i) {
                    if (i < n) {
                        const int j = (int)__umulhi(w[2 * i], (uint32_t)(i + 1));  // SHUF_W[i] = 2i
                        const unsigned long long x = ((vis >> (8 * i)) ^ (vis >> (8 * j))) & 0xffull;
                        vis ^= (x << (8 * i)) | (x << (8 * j));
                    }
                }
                int n_mv = 0;
#pragma unroll
                for (int k = 0; k < kVisN; ++k) {
                    const int wi = k == 0 ? 0 : 2 * k - 1;  // JAM_W = {0,1,3,5,7,9}
                    if (k < n) n_mv += ((unsigned long long)w[wi] < P.jam_thresh) ? 1 : 0;
                }
                n -= n_mv;
            }
#pragma unroll
            for (int k = 0; k < kVisN; ++k)
                if (k < n && row < kVisN) sid[row++] = (int16_t)((vis >> (8 * k)) & 0xff);
        }
    }
    __syncthreads();
    MT_PROF(P, e, 6);
    if (dest != nullptr) {
        // gather: float4 stores over the env's contiguous [N*204] floats; a float4 may straddle two records
        // (staging whole 68-byte records in shared memory first was slower: the extra 26 KB cost two resident CTAs)
        float4 *o4 = (float4 *)(dest + (size_t)e * N * kMtLen);
        const int total4 = N * kMtLen / 4;  // 204 = 4 * 51
        for (int q = tid; q < total4; q += kMtThreads) {
            float v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int f = 4 * q + u;
                const int slot = f / kMtRow, c = f - slot * kMtRow;  // slot = agent * 12 + row
                const int id = s.sid[slot];
                v[u] = s.table[(size_t)(id < 0 ? N : id) * kMtRow + c];
            }
#if DCA_L2_HINTS
            __stcs(&o4[q], make_float4(v[0], v[1], v[2], v[3]));  // written once, read by the policy: stream past L2
#else
            o4[q] = make_float4(v[0], v[1], v[2], v[3]);
#endif
        }
        if (P.out.slot_id) {
            int16_t *o = P.out.slot_id + (size_t)e * N * kObsRows;
            for (int t = tid; t < N * kObsRows; t += kMtThreads) o[t] = s.sid[t];
        }
    }
}

// reset_world envs/collective_assault_env.py:42-82 (keyed draws), then reset(): BSA = all, terrain (parallel_run.py:197-203)
__device__ __forceinline__ void mt_reset_world(const MtSmem &s, const MtParams &P, int e, uint32_t episode,
                                               double *theta_out) {
    const int tid = threadIdx.x, N = P.N;
    const DcaMtConfig &c = P.cfg;
    const uint32_t k0 = (uint32_t)c.seed, k1 = (uint32_t)(c.seed >> 32);
    const uint32_t gid = (uint32_t)(c.env_gid0 + e);
    const uint4 ww = philox4x32_10(gid, episode, 0u, SITE_RESET, k0, k1);
    const double theta = (2 * ((double)ww.x * 0x1p-32) - 1) * kPi;
    const double random_theta = (2 * ((double)ww.y * 0x1p-32) - 1) * kPi;
    *theta_out = theta;
    const double ct = cos(theta), st = sin(theta);
    for (int i = tid; i < N; i += kMtThreads) {
        const int tm = s.team[i], t = s.tid[i];
        const double team_pos_dir = (2 * kPi / c.n_team) * tm + random_theta;
        const double cs = cos(team_pos_dir), sn = sin(team_pos_dir);
        const double x_ = 0.0 + c.init_distance / 2 + 0.1 * (t % 5);
        const double y_ = 0.0 + ((double)(t / 5) / ((double)c.team_size[tm] / 5) - 0.5) * 10.0 * 0.25;
        const double x = x_ * cs + y_ * (-sn), y = x_ * sn + y_ * cs;
        const uint4 wa = philox4x32_10(gid, episode, (uint32_t)(1 + i), SITE_RESET, k0, k1);
        s.px[i] = x; s.py[i] = y; s.vx[i] = 0.0; s.vy[i] = 0.0;
        s.atk[i] = (team_pos_dir + kPi) + (((double)wa.x * 0x1p-32) - 0.5) / 12;
        s.flags[i] = (uint8_t)((s.flags[i] & MT_JUST_DIED) | MT_ALIVE | MT_BSA);  // hit / wasHit cleared, justDied kept
        s.ter[i] = c.introduce_terrain ? terrain_fn(c.terrain_A, c.terrain_B, x, y, ct, st) : 1.0;
        mt_build_row(s, i);
    }
}

__device__ __forceinline__ void mt_load(const MtSmem &s, const MtParams &P, int e) {
    const int N = P.N;
    if (threadIdx.x < 20) s.misc[threadIdx.x] = 0;
    if (threadIdx.x <= DCA_MT_MAX_TEAMS) {  // team uid ranges (arrange_id): misc[20 + t] = first uid of team t
        int k = 0;
#pragma unroll
        for (int t = 0; t < DCA_MT_MAX_TEAMS; ++t) k += (t < (int)threadIdx.x && t < P.cfg.n_team) ? P.cfg.team_size[t] : 0;
        s.misc[20 + threadIdx.x] = k;
    }
    for (int c = threadIdx.x; c < kMtRow; c += kMtThreads) s.table[(size_t)N * kMtRow + c] = 0.f;
    __syncthreads();
    for (int i = threadIdx.x; i < N; i += kMtThreads) {
        const size_t g = (size_t)e * N + i;
        const double2 p = ((const double2 *)P.st.pos)[g], v = ((const double2 *)P.st.vel)[g];
        s.px[i] = p.x; s.py[i] = p.y; s.vx[i] = v.x; s.vy[i] = v.y;
        s.atk[i] = P.st.atk_rad[g]; s.ter[i] = P.st.terrain[g];
        s.flags[i] = (uint8_t)((P.st.alive[g] ? MT_ALIVE : 0) | (P.st.just_died[g] ? MT_JUST_DIED : 0) |
                               (P.st.hit[g] ? MT_HIT : 0) | (P.st.was_hit[g] ? MT_WAS_HIT : 0) | (P.st.bsa[g] ? MT_BSA : 0));
        int tm = 0;
        while (i >= s.misc[21 + tm]) ++tm;
        s.team[i] = (uint8_t)tm; s.tid[i] = (uint8_t)(i - s.misc[20 + tm]);
    }
}

__device__ __forceinline__ void mt_store(const MtSmem &s, const MtParams &P, int e) {
    const int N = P.N;
    for (int i = threadIdx.x; i < N; i += kMtThreads) {
        const size_t g = (size_t)e * N + i;
        ((double2 *)P.st.pos)[g] = make_double2(s.px[i], s.py[i]);
        ((double2 *)P.st.vel)[g] = make_double2(s.vx[i], s.vy[i]);
        P.st.atk_rad[g] = s.atk[i]; P.st.terrain[g] = s.ter[i];
        const uint8_t f = s.flags[i];
        P.st.alive[g] = (f & MT_ALIVE) != 0; P.st.just_died[g] = (f & MT_JUST_DIED) != 0;
        P.st.hit[g] = (f & MT_HIT) != 0; P.st.was_hit[g] = (f & MT_WAS_HIT) != 0; P.st.bsa[g] = (f & MT_BSA) != 0;
        if (P.out.alive) P.out.alive[g] = (f & MT_ALIVE) != 0;
    }
}

__global__ void __launch_bounds__(kMtThreads, 8) k_mt_step(const MtParams P, const int32_t *__restrict__ actions,
                                                         const uint8_t *__restrict__ skip, int autoreset) {
    extern __shared__ __align__(16) unsigned char mt_raw[];
    const int e = blockIdx.x, tid = threadIdx.x, N = P.N;
    if (skip != nullptr && skip[e]) return;  // ENV-PAUSE shm_env.py:37-41: cached outputs stay
    const MtSmem s = mt_carve(mt_raw, N);
    const DcaMtConfig &c = P.cfg;
    MT_PROF(P, e, 0);
    mt_load(s, P, e);
    double theta = P.st.init_theta[e];
    const int time_step = P.st.time_step[e];
    uint32_t counter = P.st.obs_counter[e], episode = P.st.episode[e];
    const uint32_t k0 = (uint32_t)c.seed, k1 = (uint32_t)(c.seed >> 32);
    const uint32_t gid = (uint32_t)(c.env_gid0 + e);
    __syncthreads();
    MT_PROF(P, e, 1);
    // ---- set_terrain_adv (parallel_run.py:164) for the agents of the previous step's list; then World.step's new list,
    // hit flags cleared for its members (core.py:101, :161-165)
    {
        const double ct = cos(theta), st = sin(theta);
        for (int i = tid; i < N; i += kMtThreads) {
            uint8_t f = s.flags[i];
            if (f & MT_BSA) s.ter[i] = c.introduce_terrain ? terrain_fn(c.terrain_A, c.terrain_B, s.px[i], s.py[i], ct, st) : 1.0;
            f = (uint8_t)(f & ~MT_BSA);
            if (f & MT_ALIVE) {
                f = (uint8_t)((f | MT_BSA) & ~(MT_HIT | MT_WAS_HIT));
                double sn, cs;
                sincos(s.atk[i], &sn, &cs);
                s.ux[i] = cs; s.uy[i] = sn;
            }
            s.flags[i] = f;
            s.hitf[i] = 0;
            // float32 positions of the possible shooters / targets for the laser pre-filter (others = +inf)
            s.xs[i] = (f & MT_BSA) ? (float)s.px[i] : CUDART_INF_F;
            s.ys[i] = (f & MT_BSA) ? (float)s.py[i] : CUDART_INF_F;
        }
    }
    __syncthreads();
    MT_PROF(P, e, 2);
    // ---- laser core.py:167-190: every alive shooter against every alive agent of another team.  A warp takes a
    // shooter, its lanes the targets; pairs inside the shooter's radius are queued and the exact angle test + keyed
    // success draw then run one queued pair per thread (the keyed draw makes the outcome order-independent).
    {
        const int lane = tid & 31, warp = tid >> 5;
        // bounding box of each team's possible targets (float32 positions; dead = +inf are left out): a shooter only
        // scans the teams whose box is within its reach -- none at all while the teams are still apart
        for (int t = warp; t < c.n_team; t += kMtThreads / 32) {
            float x0 = CUDART_INF_F, x1 = -CUDART_INF_F, y0 = CUDART_INF_F, y1 = -CUDART_INF_F;
            for (int j = s.misc[20 + t] + lane; j < s.misc[21 + t]; j += 32) {
                const float x = s.xs[j], y = s.ys[j];
                if (x < CUDART_INF_F) { x0 = fminf(x0, x); x1 = fmaxf(x1, x); y0 = fminf(y0, y); y1 = fmaxf(y1, y); }
            }
#pragma unroll
            for (int o = 16; o; o >>= 1) {
                x0 = fminf(x0, __shfl_xor_sync(0xffffffffu, x0, o)); x1 = fmaxf(x1, __shfl_xor_sync(0xffffffffu, x1, o));
                y0 = fminf(y0, __shfl_xor_sync(0xffffffffu, y0, o)); y1 = fmaxf(y1, __shfl_xor_sync(0xffffffffu, y1, o));
            }
            if (lane == 0) { s.bbox[4 * t] = x0; s.bbox[4 * t + 1] = x1; s.bbox[4 * t + 2] = y0; s.bbox[4 * t + 3] = y1; }
        }
        __syncthreads();
        MT_PROF(P, e, 9);
        // (1) one thread per shooter: which other teams' boxes are within its reach?  -> work list of (shooter, team)
        // (lives in the sid rows, which are not written before the observation pass)
        int16_t *work = s.sid;
        for (int i = tid; i < N; i += kMtThreads) {
            if (!(s.flags[i] & MT_BSA)) continue;
            const float xif = s.xs[i], yif = s.ys[i];
            const float r2f = mt_reach2f(s.px[i], s.py[i], s.ter[i]);
            const int tm = s.team[i];
            for (int t = 0; t < c.n_team; ++t) {
                if (t == tm) continue;
                const float bx = fmaxf(fmaxf(s.bbox[4 * t] - xif, xif - s.bbox[4 * t + 1]), 0.f);
                const float by = fmaxf(fmaxf(s.bbox[4 * t + 2] - yif, yif - s.bbox[4 * t + 3]), 0.f);
                if (bx * bx + by * by <= r2f) work[atomicAdd(&s.misc[17], 1)] = (int16_t)(i | (t << 8));  // empty box: inf
            }
        }
        __syncthreads();
        // (2) one warp per work item, lanes over the members of that team
        const int n_work = s.misc[17];
        for (int k = warp; k < n_work; k += kMtThreads / 32) {
            const int i = work[k] & 0xff, t = work[k] >> 8;
            const double radius = 0.4 * s.ter[i], r2 = radius * radius;
            const double xi = s.px[i], yi = s.py[i], uxi = s.ux[i], uyi = s.uy[i];
            const float r2f = mt_reach2f(xi, yi, s.ter[i]);
            const float xif = s.xs[i], yif = s.ys[i];
            {
                for (int j = s.misc[20 + t] + lane; j < s.misc[21 + t]; j += 32) {
                const float dxf = s.xs[j] - xif, dyf = s.ys[j] - yif;
                if (!(dxf * dxf + dyf * dyf <= r2f)) continue;
                if (!(s.flags[j] & MT_BSA)) continue;
                const double dx = s.px[j] - xi, dy = s.py[j] - yi;
                const double len2 = dx * dx + dy * dy;
                if (len2 > r2) continue;
                // cone test of laser_hit_improve3: both |reg(dir +- w/2 - theta)| <= w  <=>  the target direction is
                // within w/2 of the heading  <=>  dot >= cos(w/2) |d|.  Decided here without atan2 unless it is within
                // 1e-9 (relative) of the boundary -- ~1e6 times the rounding error of either formulation.
                const double dot = dx * uxi + dy * uyi;
                const double margin = dot * dot - kCosHalfWin2 * len2;
                int amb = 0;
                if (!(len2 > 1e-100) || fabs(margin) <= 1e-9 * len2) amb = 1;
                else if (dot <= 0.0 || margin < 0.0) continue;  // outside the cone
                const int q = atomicAdd(&s.misc[16], 1);
                const int entry = (i * N + j) | (amb << 30);
                if (q < kMtPairCap) s.pairq[q] = entry;
                else mt_laser_pair(s.px, s.py, s.atk, s.ter, s.hitf, N, gid, counter, k0, k1, entry);  // queue full
                }
            }
        }
        MT_PROF(P, e, 10);
        __syncthreads();
        MT_PROF(P, e, 11);
        const int nq = min(s.misc[16], kMtPairCap);
        for (int q = tid; q < nq; q += kMtThreads)
            mt_laser_pair(s.px, s.py, s.atk, s.ter, s.hitf, N, gid, counter, k0, k1, s.pairq[q]);
    }
    __syncthreads();
    MT_PROF(P, e, 3);
    // ---- deaths (:193-201), integrate (:204-222, just-killed agents still move), kill circle (:115-132), records
    {
        const double frac = fmin((double)time_step / (double)c.steps_to_narrow, 1.0);
        const double cur = c.size + (c.size / 3.5 - c.size) * frac, cur2 = cur * cur;
        for (int i = tid; i < N; i += kMtThreads) {
            uint8_t f = s.flags[i];
            if (!(f & MT_ALIVE)) f = (uint8_t)(f & ~MT_JUST_DIED);
            if (f & MT_BSA) {
                const int h = s.hitf[i];
                if (h & 1) f |= MT_HIT;
                if (h & 2) { f |= MT_WAS_HIT; f = (uint8_t)((f & ~MT_ALIVE) | MT_JUST_DIED); }
                const int a = actions[(size_t)e * N + i];
                double a0 = 0, a1 = 0, ar = 0;
                if (a == 1) a0 = +1.0;
                if (a == 2) a0 = -1.0;
                if (a == 3) a1 = +1.0;
                if (a == 4) a1 = -1.0;
                if (a == 5) ar = +0.17;
                if (a == 6) ar = -0.17;
                double vx = s.vx[i] * (1 - 0.25), vy = s.vy[i] * (1 - 0.25);
                vx += ((a0 * 3) / 1.0) * 0.1; vy += ((a1 * 3) / 1.0) * 0.1;
                const double speed = sqrt(vx * vx + vy * vy);
                if (speed > 1.0) { vx = vx / speed * 1.0; vy = vy / speed * 1.0; }
                s.vx[i] = vx; s.vy[i] = vy;
                if (f & MT_ALIVE) s.atk[i] = reg_angle(s.atk[i] + ar);
                const double x = s.px[i] + vx * 0.1, y = s.py[i] + vy * 0.1;
                s.px[i] = x; s.py[i] = y;
                if ((f & MT_ALIVE) && x * x + y * y > cur2) f = (uint8_t)((f & ~MT_ALIVE) | MT_WAS_HIT | MT_JUST_DIED);
            }
            s.flags[i] = f;
            mt_build_row(s, i);
            if (f & MT_ALIVE) atomicAdd(&s.misc[s.team[i]], 1);
            // reward envs/collective_assault_env.py:87-95
            float r = 0.f;
            if (f & (MT_ALIVE | MT_JUST_DIED)) {
                if (f & MT_HIT) r += 1.f;
                if (f & MT_WAS_HIT) r = c.half_death_reward ? -0.5f : -1.f;
            }
            P.out.reward[(size_t)e * N + i] = r;
        }
    }
    __syncthreads();
    MT_PROF(P, e, 4);
    // ---- done / ranking parallel_run.py:219-245 (every thread computes the same flags)
    int teams_alive = 0;
    for (int t = 0; t < c.n_team; ++t) teams_alive += s.misc[t] > 0;
    const bool done = teams_alive <= 1 || time_step == c.max_episode_step - 1;
    if (tid == 0) {
        P.out.done[e] = done;
        if (done) {
            bool all0 = true;
            int rk[DCA_MT_MAX_TEAMS];
            for (int t = 0; t < c.n_team; ++t) {
                int r = 0;
                for (int u = 0; u < c.n_team; ++u) r += s.misc[u] > s.misc[t];
                rk[t] = r;
                all0 &= r == 0;
            }
            for (int t = 0; t < c.n_team; ++t) P.out.ranking[(size_t)e * c.n_team + t] = all0 ? -1 : rk[t];
        }
    }
    int new_time = time_step + 1;
    counter += 1;
    if (done && autoreset) {
        mt_obs_pass(s, P, e, P.out.tobs, counter);  // terminal obs = info['obs-echo'] (tobs may be NULL)
        __syncthreads();
        episode += 1;
        mt_reset_world(s, P, e, episode, &theta);
        new_time = 0;
        counter += 1;
        __syncthreads();
    }
    MT_PROF(P, e, 5);
    mt_obs_pass(s, P, e, P.out.obs, counter);
    MT_PROF(P, e, 7);
    mt_store(s, P, e);
    MT_PROF(P, e, 8);
    if (tid == 0) {
        P.st.init_theta[e] = theta; P.st.time_step[e] = new_time;
        P.st.obs_counter[e] = counter; P.st.episode[e] = episode;
    }
}

// reset(): masked envs get a fresh episode + observation (EnvAutoReset.reset / first manual reset)
__global__ void __launch_bounds__(kMtThreads) k_mt_reset(const MtParams P, const uint8_t *__restrict__ mask) {
    extern __shared__ __align__(16) unsigned char mt_raw[];
    const int e = blockIdx.x;
    if (mask != nullptr && !mask[e]) return;
    const MtSmem s = mt_carve(mt_raw, P.N);
    mt_load(s, P, e);
    const uint32_t episode = P.st.episode[e] + 1, counter = P.st.obs_counter[e] + 1;
    double theta;
    __syncthreads();
    mt_reset_world(s, P, e, episode, &theta);
    __syncthreads();
    mt_obs_pass(s, P, e, P.out.obs, counter);
    mt_store(s, P, e);
    if (threadIdx.x == 0) {
        P.st.init_theta[e] = theta; P.st.time_step[e] = 0;
        P.st.obs_counter[e] = counter; P.st.episode[e] = episode;
        if (P.out.done) P.out.done[e] = 0;
    }
}

__global__ void k_mt_init(MtParams P) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < P.E) { P.st.episode[e] = 0xFFFFFFFFu; P.st.obs_counter[e] = 0xFFFFFFFFu; }
}

}  // namespace dca
