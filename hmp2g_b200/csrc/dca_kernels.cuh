// dca_kernels.cuh -- the DCA step / observation / reset kernels (one CTA per env, state in shared memory).
// Reference semantics are cited per phase (paths relative to /root/reference).
#pragma once
#include "dca_device.cuh"

namespace dca {

// Shared-memory view of one env: the blob (TMA destination) followed by per-step work arrays.
struct Smem {
    unsigned char *blob;
    double *px, *py, *vx, *vy, *atk, *ter;
    int32_t *nhit, *nwas;
    uint8_t *flags;
    EnvScalars *sc;
    double *fx, *fy, *ar;  // force (act[0:2]) and rotation (act[2]) of this step
    float2 *xyg, *xya;     // float32-rounded positions of guards / attackers (distance matrix inputs), each padded
                           // to a multiple of 32 entries; dead agents and padding = (+inf, +inf)
    int32_t *hitcnt;       // times agent i was hit this step (also: exclusion keys during reset)
    float *table;          // [N,16] float32 entity records
    int16_t *sid;          // [G,12] entity id per obs row, -1 = zero row
    int32_t *misc;         // [0] guard deaths, [1] attacker deaths
    uint64_t *bar;
};

__host__ __device__ inline int smem_bytes(const BlobLayout &L) {
    int o = L.bytes;
    o += 3 * align_up(L.N * 8, 16);   // fx fy ar
    o += (align_up(L.G, 32) + align_up(L.A, 32)) * 8;  // xyg xya
    o += align_up(L.N * 4, 16);       // hitcnt
    o += L.N * kObsRow * 4;           // table
    o += align_up(L.G * kObsRows * 2, 16);  // sid
    o += 32;                          // misc
    o += 16;                          // mbarrier
    return o;
}

__device__ __forceinline__ Smem carve(unsigned char *base, const BlobLayout &L) {
    Smem s;
    s.blob = base;
    s.px = (double *)(base + L.px); s.py = (double *)(base + L.py);
    s.vx = (double *)(base + L.vx); s.vy = (double *)(base + L.vy);
    s.atk = (double *)(base + L.atk); s.ter = (double *)(base + L.terrain);
    s.nhit = (int32_t *)(base + L.num_hit); s.nwas = (int32_t *)(base + L.num_was_hit);
    s.flags = base + L.flags;
    s.sc = (EnvScalars *)(base + L.scalars);
    unsigned char *p = base + L.bytes;
    const int f64 = align_up(L.N * 8, 16), f32 = align_up(L.N * 4, 16);
    s.fx = (double *)p; p += f64; s.fy = (double *)p; p += f64; s.ar = (double *)p; p += f64;
    s.xyg = (float2 *)p; p += align_up(L.G, 32) * 8; s.xya = (float2 *)p; p += align_up(L.A, 32) * 8;
    s.hitcnt = (int32_t *)p; p += f32;
    s.table = (float *)p; p += L.N * kObsRow * 4;
    s.sid = (int16_t *)p; p += align_up(L.G * kObsRows * 2, 16);
    s.misc = (int32_t *)p; p += 32;
    s.bar = (uint64_t *)p;
    return s;
}

// Fetch the env blob with one TMA bulk copy; returns after the data has landed for every thread.
__device__ __forceinline__ void load_blob(const Smem &s, const Params &P, int e) {
    if (threadIdx.x == 0) {
        mbar_init(s.bar, 1);
        fence_barrier_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_expect_tx(s.bar, (uint32_t)P.L.bytes);
        tma_load_1d(s.blob, P.blobs + (size_t)e * P.L.bytes, (uint32_t)P.L.bytes, s.bar);
    }
    mbar_wait(s.bar, 0);
}

// Write the blob back with one TMA bulk store.  Generic-proxy writes are fenced into the async proxy first.
__device__ __forceinline__ void store_blob(const Smem &s, const Params &P, int e) {
    fence_proxy_async();
    __syncthreads();
    if (threadIdx.x == 0) {
        tma_store_1d(P.blobs + (size_t)e * P.L.bytes, s.blob, (uint32_t)P.L.bytes);
        tma_store_commit_wait();
    }
}

// Optional per-CTA phase clocks (diagnostics, dca_debug_phase_clocks): slot k of env e <- clock64() of thread 0.
#define DCA_PROF(P, e, k)                                                                     \
    do {                                                                                      \
        if ((P).prof != nullptr && threadIdx.x == 0) (P).prof[(size_t)(e) * 16 + (k)] = clock64(); \
    } while (0)

// ------------------------------------------------------------------ terrain
// World.set_terrain_adv core.py:224-234 + get_terrain core.py:306-314; alive agents only, dead keep old value.
__device__ __forceinline__ double terrain_at(const DcaConfig &c, double X, double Y, double ct, double st) {
    if (!c.introduce_terrain) return 1.0;
    const double X_ = X * ct + Y * st;
    const double Y_ = -X * st + Y * ct;
    const double a = 0.1 * X_, b = 0.1 * Y_;
    double Z = -1 + c.terrain_B * (a * a + b * b);
    Z = Z - c.terrain_A * cos(2 * kPi * (0.3 * X_));
    Z = Z - c.terrain_A * cos(2 * kPi * (0.5 * Y_));
    return -Z;
}

// Refresh the padded float32 position arrays from the fp64 state (dead agents and padding slots = +inf).
__device__ __forceinline__ void fill_xy(const Smem &s, const BlobLayout &L) {
    const int Gp = align_up(L.G, 32), Ap = align_up(L.A, 32);
    for (int k = threadIdx.x; k < Gp + Ap; k += blockDim.x) {
        const bool isg = k < Gp;
        const int a = isg ? k : k - Gp;
        const int i = isg ? a : L.G + a;
        float2 v = make_float2(CUDART_INF_F, CUDART_INF_F);
        if ((isg ? a < L.G : a < L.A) && (s.flags[i] & F_ALIVE)) v = make_float2((float)s.px[i], (float)s.py[i]);
        s.xyg[k] = v;  // xya follows xyg contiguously
    }
}

// Branch-free stable insert of (d, j) into the ascending top-6 list: position = first k with d < bd[k]
// (ties keep the earlier, lower-index entry in front), everything from there shifts down by one.
__device__ __forceinline__ void top6_insert(float (&bd)[kVisN], int (&bi)[kVisN], float d, int j) {
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

// ------------------------------------------------------------------ observation pass
// collective_assaultEnvV1.observation collective_assault_env.py:228-359 for every guard of the env.
// dest == nullptr: the pass is only counted (keeps the keyed RNG stream aligned), nothing is computed.
__device__ void obs_pass(const Smem &s, const Params &P, int e, float *dest, const uint32_t *rng_words,
                         int prof_slot = 12) {
    const BlobLayout &L = P.L;
    const int tid = threadIdx.x, nt = blockDim.x;
    const uint32_t counter = s.sc->obs_counter + 1;
    if (dest == nullptr) {
        __syncthreads();
        if (tid == 0) s.sc->obs_counter = counter;
        __syncthreads();
        return;
    }
    fill_xy(s, L);
    // entity table :237-276  [alive,x,y,atk_rad,vx,vy,iden,terrain,bit0..bit7] as float32
    for (int i = tid; i < L.N; i += nt) {
        float4 *row = (float4 *)(s.table + i * kObsRow);
        const float x = (float)s.px[i], y = (float)s.py[i];
        const bool alive_i = s.flags[i] & F_ALIVE;
        row[0] = make_float4(alive_i ? 1.f : 0.f, x, y, (float)s.atk[i]);
        row[1] = make_float4((float)s.vx[i], (float)s.vy[i], (float)i, (float)s.ter[i]);
        // get_binary_array :207-215 (float halving, int8 wrap, floored %); identical to plain bits for i < 256
        int v = i;
        float b[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            b[k] = (float)((((v % 2) + 2) % 2) == 1);
            v = (int)(signed char)(int)((double)v / 2.0);
        }
        row[2] = make_float4(b[0], b[1], b[2], b[3]);
        row[3] = make_float4(b[4], b[5], b[6], b[7]);
    }
    __syncthreads();
    DCA_PROF(P, e, prof_slot);
    if (tid == 0) s.sc->obs_counter = counter;

    const uint32_t k0 = (uint32_t)P.cfg.seed, k1 = (uint32_t)(P.cfg.seed >> 32);
    const uint32_t gid = (uint32_t)(P.cfg.env_gid0 + e);
    // One task per (guard, list): 2G tasks dealt out evenly over the CTA's warps (a warp lives on one SM
    // sub-partition; a plain thread-per-guard mapping would leave half of the sub-partitions idle here).
    const int ntask = 2 * L.G, nwarp = nt >> 5, quota = (ntask + nwarp - 1) / nwarp;
    for (int o = tid & 31; o < quota; o += 32) {
        const int task = (tid >> 5) * quota + o;
        if (task >= ntask) break;
        const int g = task >> 1, lst = task & 1;  // 0 = hostile (:318-328), 1 = friend (:330-342)
        int16_t *sid = s.sid + g * kObsRows + (lst == 0 ? kVisN : 0);  // friend rows 0..5, hostile rows 6..11 (:344)
        if (!(s.flags[g] & F_ALIVE)) {  // :313-316 dead guard: zeros, no RNG draws
#pragma unroll
            for (int r = 0; r < kVisN; ++r) sid[r] = -1;
            continue;
        }
        const float2 pg = s.xyg[g];
        {
            const int base = lst == 0 ? L.G : 0;
            // argsort()[:6] then the range mask (:308-309, :320-322, :331-333) == the (up to) 6 nearest IN-RANGE
            // candidates: anything farther than an out-of-range member of the top-6 is out of range too.
            // Two phases per chunk of 32 candidates so that the warp never diverges on a per-candidate basis:
            //  1. branch-free scan: squared distance (same roundings as distance_matrix :361-368) tested against
            //     three nested radii -> three bit masks.  q[l] = largest f32 whose sqrt_rn is <= range, range/2,
            //     range/4, so "in mask l" == "distance <= radius l" exactly; dead / padding = +inf never pass.
            //  2. the smallest radius holding >= 6 candidates contains this chunk's 6 nearest; only those get the
            //     sqrt_rn and the stable sorted insert (ties -> lower index, candidates visited in index order).
            const float q0 = P.smax[lst][0], q1 = P.smax[lst][1], q2 = P.smax[lst][2];
            const float2 *list = lst == 0 ? s.xya : s.xyg;
            const int cntp = align_up(lst == 0 ? L.A : L.G, 32);
            float bd[kVisN];
            int bi[kVisN];
#pragma unroll
            for (int k = 0; k < kVisN; ++k) { bd[k] = CUDART_INF_F; bi[k] = -1; }
            for (int c0 = 0; c0 < cntp; c0 += 32) {
                const float4 *c4 = (const float4 *)(list + c0);
                uint32_t m0 = 0, m1 = 0, m2 = 0;
#pragma unroll
                for (int jj = 0; jj < 16; ++jj) {
                    const float4 v = c4[jj];
                    const float dx0 = __fsub_rn(v.x, pg.x), dy0 = __fsub_rn(v.y, pg.y);
                    const float dx1 = __fsub_rn(v.z, pg.x), dy1 = __fsub_rn(v.w, pg.y);
                    const float s0 = __fadd_rn(__fmul_rn(dx0, dx0), __fmul_rn(dy0, dy0));
                    const float s1 = __fadd_rn(__fmul_rn(dx1, dx1), __fmul_rn(dy1, dy1));
                    m0 |= (s0 <= q0) ? (1u << (2 * jj)) : 0u;
                    m1 |= (s0 <= q1) ? (1u << (2 * jj)) : 0u;
                    m2 |= (s0 <= q2) ? (1u << (2 * jj)) : 0u;
                    m0 |= (s1 <= q0) ? (2u << (2 * jj)) : 0u;
                    m1 |= (s1 <= q1) ? (2u << (2 * jj)) : 0u;
                    m2 |= (s1 <= q2) ? (2u << (2 * jj)) : 0u;
                }
                uint32_t m = __popc(m2) >= kVisN ? m2 : (__popc(m1) >= kVisN ? m1 : m0);
                while (m) {
                    const int j = __ffs(m) - 1;
                    m &= m - 1;
                    const float2 cj = list[c0 + j];
                    const float dx = __fsub_rn(cj.x, pg.x), dy = __fsub_rn(cj.y, pg.y);
                    const float d = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
                    top6_insert(bd, bi, d, c0 + j);
                }
            }
            // visible list in sorted order, packed one byte per entry
            unsigned long long vis = 0;
            int n = 0;
#pragma unroll
            for (int k = 0; k < kVisN; ++k) {
                if (bi[k] >= 0) { vis |= (unsigned long long)bi[k] << (8 * n); ++n; }
            }
            int row = 0;
            if (lst == 1) {  // :334-335 self = first visible friend, excluded from shuffle / jamming
                sid[row++] = (int16_t)(vis & 0xff);
                vis >>= 8; --n;
            }
            // item_random_mv :217-226 with the keyed words (shuffle i = n-1..1, then jam draws)
            if (n > 0) {
                uint32_t w[DCA_RNG_WORDS];
                if (rng_words) {
                    const uint4 *src = (const uint4 *)(rng_words + (((size_t)e * L.G + g) * 2 + lst) * DCA_RNG_WORDS);
#pragma unroll
                    for (int b = 0; b < 3; ++b) {
                        const uint4 q = src[b];
                        w[4 * b] = q.x; w[4 * b + 1] = q.y; w[4 * b + 2] = q.z; w[4 * b + 3] = q.w;
                    }
                } else {
                    const int nblk = n <= 2 ? 1 : (n <= 4 ? 2 : 3);
#pragma unroll
                    for (int b = 0; b < 3; ++b) {
                        if (b < nblk) {
                            const uint4 q = philox4x32_10(gid, counter, (uint32_t)(g * 2 + lst), SITE_OBS | (b << 8), k0, k1);
                            w[4 * b] = q.x; w[4 * b + 1] = q.y; w[4 * b + 2] = q.z; w[4 * b + 3] = q.w;
                        } else {
                            w[4 * b] = w[4 * b + 1] = w[4 * b + 2] = w[4 * b + 3] = 0;
                        }
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
                    if (k < n) n_mv += ((double)w[wi] * 0x1p-32 < P.cfg.random_jam_prob) ? 1 : 0;
                }
                n -= n_mv;
            }
#pragma unroll
            for (int k = 0; k < kVisN; ++k) {
                if (row < kVisN) {
                    sid[row] = k < n ? (int16_t)(base + (int)((vis >> (8 * k)) & 0xff)) : (int16_t)-1;
                    ++row;
                }
            }
        }
    }
    __syncthreads();
    DCA_PROF(P, e, prof_slot + 1);
    // gather: 48 x 128-bit per guard, consecutive threads write consecutive 16 B (coalesced)
    float4 *out4 = (float4 *)(dest + (size_t)e * L.G * kObsLen);
    const float4 *tab4 = (const float4 *)s.table;
    const int total = L.G * kObsRows * 4;
    for (int t = tid; t < total; t += nt) {
        const int id = s.sid[t >> 2];
        out4[t] = id >= 0 ? tab4[id * 4 + (t & 3)] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (P.out.slot_id) {
        int16_t *o = P.out.slot_id + (size_t)e * L.G * kObsRows;
        for (int t = tid; t < L.G * kObsRows; t += nt) o[t] = s.sid[t];
    }
    __syncthreads();
    DCA_PROF(P, e, prof_slot + 2);
}

// ------------------------------------------------------------------ reset
// collective_assaultEnvV1.reset_world collective_assault_env.py:55-136 with keyed draws, then the terrain pass
// of collective_assaultGlobalEnv.reset parallel_run.py:258.  Caller runs obs_pass afterwards.
__device__ void reset_world(const Smem &s, const Params &P, int e) {
    const BlobLayout &L = P.L;
    const DcaConfig &c = P.cfg;
    const int tid = threadIdx.x, nt = blockDim.x;
    const uint32_t k0 = (uint32_t)c.seed, k1 = (uint32_t)(c.seed >> 32);
    const uint32_t gid = (uint32_t)(c.env_gid0 + e);
    const uint32_t ep = s.sc->episode + 1;
    const uint4 ww = philox4x32_10(gid, ep, 0u, SITE_RESET, k0, k1);
    const double theta = (2 * ((double)ww.x * 0x1p-32) - 1) * kPi;  // :64
    double sn, ct;
    sincos(theta, &sn, &ct);
    int n_excl = 0;
    if (c.enable_attacker_fluctuation) {  // :67-77
        const int low = L.A - c.attacker_fluctuation, high = L.A + 1;
        n_excl = L.A - (low + (int)__umulhi(ww.y, (uint32_t)(high - low)));
    }
    __syncthreads();  // everyone has read episode
    if (tid == 0) {
        s.sc->episode = ep;
        s.sc->time_step = 0;
        s.sc->n_alive_attackers = L.A;  // :59-60 (not corrected for excluded attackers, as in the reference)
        s.sc->n_alive_guards = L.G;
        s.sc->init_theta = theta;
    }
    for (int i = tid; i < L.N; i += nt) {
        const uint4 a = philox4x32_10(gid, ep, (uint32_t)(1 + i), SITE_RESET, k0, k1);
        const bool attacker = i >= L.G;
        const double u = (double)a.x * 0x1p-32;
        const double u1 = ((double)a.y + 1.0) * 0x1p-32, u2 = (double)a.z * 0x1p-32;
        const double r = sqrt(-2.0 * log(u1));
        double sb, cb;
        sincos(2.0 * kPi * u2, &sb, &cb);
        const double n0 = r * cb, n1 = r * sb;
        s.hitcnt[i] = (int32_t)a.w;  // exclusion sort key
        s.vx[i] = 0; s.vy[i] = 0;
        s.atk[i] = attacker ? (theta + kPi) + (u - 0.5) / 12 : theta + (u - 0.5) / 12;  // :85
        const double yMin = -5.0, yMax = 5.0;
        double x_, y_;
        if (attacker) {  // :92-96
            x_ = 0.0 + c.init_distance / 2;
            y_ = (yMax - yMin) / L.A * (i - L.G + 0.5) + yMin;
        } else {  // :115-119
            x_ = 0.0 - c.init_distance / 2;
            y_ = (yMax - yMin) / L.G * (i + 0.5) + yMin;
        }
        double px = x_ + (n0 - 0.5) / 10, py = y_ + (n1 - 0.5) / 10;  // :97 / :120
        if ((attacker ? L.A : L.G) > 48) {                             // :99-106 / :122-129
            double ratio = 1;
            if (i % 3 == 0) ratio = 0.5;
            if (i % 3 == 1) ratio = 0.75;
            px = 0.0 + (px - 0.0) * ratio;
            py = 0.0 + (py - 0.0) * ratio;
        }
        s.px[i] = px * ct + py * (-sn);  // np.dot(pos, rotate.T) :108 / :131
        s.py[i] = px * sn + py * ct;
        if (attacker) s.ter[i] = CUDART_NAN;  // :113
        s.nhit[i] = 0; s.nwas[i] = 0;
        // alive=True, hit=wasHit=False; justDied is not touched by reset_world
        s.flags[i] = (uint8_t)((s.flags[i] & F_JUST_DIED) | F_ALIVE);
    }
    __syncthreads();
    if (n_excl > 0) {  // keyed stand-in for np.random.choice(replace=False): the n_excl smallest (key, index)
        for (int a = tid; a < L.A; a += nt) {
            const uint32_t ka = (uint32_t)s.hitcnt[L.G + a];
            int rank = 0;
            for (int b = 0; b < L.A; ++b) {
                const uint32_t kb = (uint32_t)s.hitcnt[L.G + b];
                rank += (kb < ka || (kb == ka && b < a)) ? 1 : 0;
            }
            if (rank < n_excl) s.flags[L.G + a] &= (uint8_t)~F_ALIVE;  // :109-111
        }
        __syncthreads();
    }
    for (int i = tid; i < L.N; i += nt)  // set_terrain_adv after reset, parallel_run.py:258
        if (s.flags[i] & F_ALIVE) s.ter[i] = terrain_at(c, s.px[i], s.py[i], ct, sn);
    __syncthreads();
}

// ------------------------------------------------------------------ fused step kernel
// collective_assaultGlobalEnv.step parallel_run.py:192-246 + EnvAutoReset.step UTIL/shm_env.py:33-74.
__global__ void __launch_bounds__(256) k_step(const Params P, const int32_t *__restrict__ actions,
                                               const double *__restrict__ att_force, const uint8_t *__restrict__ skip,
                                               const uint32_t *__restrict__ rng_words, int autoreset) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int e = blockIdx.x;
    if (skip != nullptr && skip[e]) return;  // ENV-PAUSE shm_env.py:37-41: cached outputs stay in place
    const BlobLayout &L = P.L;
    const DcaConfig &c = P.cfg;
    const Smem s = carve(smem_raw, L);
    const int tid = threadIdx.x, nt = blockDim.x;
    DCA_PROF(P, e, 0);
    load_blob(s, P, e);
    DCA_PROF(P, e, 1);

    double sn, ct;
    sincos(s.sc->init_theta, &sn, &ct);
    if (tid < 2) s.misc[tid] = 0;
    fill_xy(s, L);
    // phase A: guard action decode (_set_action parallel_run.py:326-374), terrain (core.py:224-234),
    // attacker act init (cheat_script_ai.py:31-34) + injected planner force (:183-184), clear hit flags (core.py:239-242)
    for (int i = tid; i < L.N; i += nt) {
        uint8_t f = s.flags[i];
        const bool alive = f & F_ALIVE;
        double fx = 0, fy = 0, ar = 0;
        if (i < L.G) {
            const int a = actions[(size_t)e * L.G + i];
            double a0 = 0, a1 = 0;
            if (a == 1) a0 = +1.0;
            if (a == 2) a0 = -1.0;
            if (a == 3) a1 = +1.0;
            if (a == 4) a1 = -1.0;
            if (a == 5) ar = +c.agent_max_rot;
            if (a == 6) ar = -c.agent_max_rot;
            fx = a0 * c.agent_acc; fy = a1 * c.agent_acc;
        } else if (alive) {
            const double2 F = ((const double2 *)att_force)[(size_t)e * L.A + (i - L.G)];
            fx = F.x; fy = F.y;
        }
        s.fx[i] = fx; s.fy[i] = fy; s.ar[i] = ar;
        s.hitcnt[i] = 0;
        if (alive) {
            s.ter[i] = terrain_at(c, s.px[i], s.py[i], ct, sn);
            s.flags[i] = f & (uint8_t)~(F_HIT | F_WAS_HIT);
        }
    }
    __syncthreads();
    DCA_PROF(P, e, 2);
    DCA_PROF(P, e, 3);
    // phase B+C: scripted aim (cheat_script_ai.py:196-218) and apply_laser_effect (core.py:244-263 with
    // laser_hit_improve3 cython_func.pyx:22-46).  Warp roles: the first half of the CTA's warps own the attackers
    // (aim, then fire), the second half the guards (fire), so no warp mixes the two code paths.
    {
        const int nwarp = nt >> 5, wa = nwarp >> 1;
        const bool att_role = (tid >> 5) < wa;
        const int role_threads = (att_role ? wa : nwarp - wa) * 32;
        const int role_tid = att_role ? tid : tid - wa * 32;
        const int role_cnt = att_role ? L.A : L.G;
        const float2 *opp = att_role ? s.xyg : s.xya;       // float32 positions of the opposing team (+inf = dead)
        const int opp_cnt = att_role ? L.G : L.A, opp_base = att_role ? 0 : L.G;
        const int opp4 = (opp_cnt + 1) >> 1;                // float4 = 2 candidates; padding slots are +inf
        const float4 *opp_v4 = (const float4 *)opp;
        for (int t = role_tid; t < role_cnt; t += role_threads) {
            const int i = att_role ? L.G + t : t;
            const bool alive = s.flags[i] & F_ALIVE;
            const float2 po = att_role ? s.xya[t] : s.xyg[t];
            if (att_role) {
                int best = -1;
                if (alive) {
                    // np.argmin over the float32 distance row (first minimum; dead guards +inf).  One branch-free pass
                    // tracks the two smallest SQUARED distances and the first index of the smallest; sqrt_rn is
                    // monotone, so only when the runner-up is within 2^-20 of the minimum can the rounded distances
                    // tie -- then (rare) a second pass compares true sqrt_rn values, lowest index first.
                    float s1 = CUDART_INF_F, s2 = CUDART_INF_F;
                    best = 0;
#pragma unroll 4
                    for (int jj = 0; jj < opp4; ++jj) {
                        const float4 v = opp_v4[jj];
                        const float dx0 = __fsub_rn(v.x, po.x), dy0 = __fsub_rn(v.y, po.y);
                        const float dx1 = __fsub_rn(v.z, po.x), dy1 = __fsub_rn(v.w, po.y);
                        const float q0 = __fadd_rn(__fmul_rn(dx0, dx0), __fmul_rn(dy0, dy0));
                        const float q1 = __fadd_rn(__fmul_rn(dx1, dx1), __fmul_rn(dy1, dy1));
                        bool lt = q0 < s1;
                        s2 = fminf(s2, lt ? s1 : q0); best = lt ? 2 * jj : best; s1 = fminf(s1, q0);
                        lt = q1 < s1;
                        s2 = fminf(s2, lt ? s1 : q1); best = lt ? 2 * jj + 1 : best; s1 = fminf(s1, q1);
                    }
                    if (s2 <= __fmul_ru(s1, 1.000001f) && s1 < CUDART_INF_F) {
                        float bd = CUDART_INF_F;
                        for (int g = 0; g < opp_cnt; ++g) {
                            const float d = f32_dist(po.x, po.y, opp[g].x, opp[g].y);
                            if (d < bd) { bd = d; best = g; }
                        }
                    }
                    double th = atan2(s.py[best] - s.py[i], s.px[best] - s.px[i]);
                    if (th < 0) th += 2 * kPi;
                    s.atk[i] = th;
                }
                if (P.out.nearest) P.out.nearest[(size_t)e * L.A + t] = best;
                if (c.disable_red_function) continue;  // can_fire = False (cheat_script_ai.py:216-217)
            }
            if (!alive) continue;
            const double radius = c.agent_shoot_base_radius * s.ter[i];
            const double r2 = radius * radius;
            const double dir = s.atk[i];
            const double pos_r = dir + c.agent_shoot_win / 2, neg_r = dir - c.agent_shoot_win / 2;
            const double ox = s.px[i], oy = s.py[i];
            // conservative float32 pre-filter (dead = +inf): a pair whose f32 squared distance exceeds r^2 by
            // 1% + 1e-3 cannot pass the exact fp64 test (f32 rounding of |coord| <= 64 perturbs d^2 by < 1e-4).
            // Candidates are collected branch-free into a bit mask per 32 opponents; the exact fp64 test
            // (atan2, two reg_angle) then runs once per set bit, all lanes of the warp in parallel.
            const bool prefilter = fabs(ox) <= 64.0 && fabs(oy) <= 64.0 && r2 < 1.0e3;
            const float r2f = prefilter ? __fadd_ru(__fmul_ru((float)r2, 1.01f), 1.0e-3f) : CUDART_INF_F;
            int hits = 0;
            for (int c0 = 0; c0 < opp_cnt; c0 += 32) {
                uint32_t m = 0;
                const int n4 = min(16, opp4 - (c0 >> 1));
#pragma unroll 4
                for (int jj = 0; jj < n4; ++jj) {
                    const float4 v = opp_v4[(c0 >> 1) + jj];
                    const float dx0 = v.x - po.x, dy0 = v.y - po.y, dx1 = v.z - po.x, dy1 = v.w - po.y;
                    const float f0 = dx0 * dx0 + dy0 * dy0, f1 = dx1 * dx1 + dy1 * dy1;
                    m |= ((f0 <= r2f) ? 1u : 0u) << (2 * jj);
                    m |= ((f1 <= r2f) ? 2u : 0u) << (2 * jj);
                }
                while (m) {
                    const int j = opp_base + c0 + __ffs(m) - 1;
                    m &= m - 1;
                    if (j >= opp_base + opp_cnt) break;
                    if (!prefilter && !(s.flags[j] & F_ALIVE)) continue;  // r2f = +inf lets dead targets through
                    const double dx = s.px[j] - ox, dy = s.py[j] - oy;
                    const double d2 = dx * dx + dy * dy;
                    if (d2 > r2) continue;
                    const double tt = atan2(dy, dx);
                    const double d1 = fabs(reg_angle(pos_r - tt)), d2r = fabs(reg_angle(neg_r - tt));
                    if (d1 <= c.agent_shoot_win && d2r <= c.agent_shoot_win) {
                        ++hits;
                        atomicAdd(&s.hitcnt[j], 1);
                    }
                }
            }
            if (hits) { s.flags[i] |= F_HIT; s.nhit[i] += hits; }
        }
    }
    __syncthreads();
    DCA_PROF(P, e, 4);
    // phase D: deaths core.py:266-279, forces :210-222,339-353, integrate_state :282-302 (survivors only)
    for (int i = tid; i < L.N; i += nt) {
        uint8_t f = s.flags[i];
        if (!(f & F_ALIVE)) {
            s.flags[i] = f & (uint8_t)~F_JUST_DIED;
            continue;
        }
        const int hc = s.hitcnt[i];
        if (hc > 0) {
            s.nwas[i] += hc;
            s.flags[i] = (uint8_t)((f | F_WAS_HIT | F_JUST_DIED) & ~F_ALIVE);
            atomicAdd(&s.misc[i < L.G ? 0 : 1], 1);
            continue;
        }
        const double x = s.px[i], y = s.py[i];
        const double wall = c.size, esz = 0.05, kk = 1e-10, dt = 0.1;
        const double d0 = x - (esz + (-wall)), d1 = wall - x - esz, d2 = y - esz - (-wall), d3 = wall - y - esz;
        double wx = 0, wy = 0;
        if (!(d0 > 0 && d1 > 0 && d2 > 0 && d3 > 0)) {
            const double f1 = 100.0 * (logaddexp0(-d0 / kk) * kk), f2 = 100.0 * (logaddexp0(-d1 / kk) * kk);
            const double f3 = 100.0 * (logaddexp0(-d2 / kk) * kk), f4 = 100.0 * (logaddexp0(-d3 / kk) * kk);
            wx = f1 - f2; wy = f3 - f4;
        }
        const double Fx = wx + s.fx[i], Fy = wy + s.fy[i];
        double vx = s.vx[i] * 0.75, vy = s.vy[i] * 0.75;
        vx += (Fx / 1.0) * dt; vy += (Fy / 1.0) * dt;
        const double speed = sqrt(vx * vx + vy * vy);
        if (speed > c.agent_max_speed) {
            vx = vx / speed * c.agent_max_speed;
            vy = vy / speed * c.agent_max_speed;
        }
        s.atk[i] = reg_angle(s.atk[i] + s.ar[i]);
        s.vx[i] = vx; s.vy[i] = vy;
        s.px[i] = x + vx * dt; s.py[i] = y + vy * dt;
    }
    __syncthreads();
    DCA_PROF(P, e, 5);
    // phase E: done/win (_get_done parallel_run.py:286-316), reward (collective_assault_env.py:153-179, :228-231)
    const int nG = s.sc->n_alive_guards - s.misc[0], nA = s.sc->n_alive_attackers - s.misc[1];
    const int ts = s.sc->time_step;
    int done = 0, win = 0;
    if (nA == 0) { done = 1; win = 1; }
    else if (nG == 0) { done = 1; win = 0; }
    else if (ts == c.max_episode_step - 1) { done = 1; win = (nA * L.G <= nG * L.A) ? 1 : 0; }
    for (int g = tid; g < L.G; g += nt) {
        const uint8_t f = s.flags[g];
        float r = 0.f;
        if (f & (F_ALIVE | F_JUST_DIED)) {
            if (f & F_HIT) r += 1.f;
            if (f & F_WAS_HIT) r += c.half_death_reward ? -0.5f : -1.f;
        }
        if (c.reward_debug && done && win) r += (float)c.reward_debug_value;
        P.out.reward[(size_t)e * L.G + g] = r;
    }
    __syncthreads();  // all threads have read the scalars
    if (tid == 0) {
        s.sc->n_alive_guards = nG; s.sc->n_alive_attackers = nA;
        s.sc->time_step = ts + 1;  // :238
        P.out.done[e] = (uint8_t)done;
        P.out.win[e] = (uint8_t)win;
    }
    __syncthreads();
    DCA_PROF(P, e, 6);
    // phase F: observation; on done the terminal obs goes to tobs, the env is reset and observed again
    const bool do_reset = done && autoreset;
    obs_pass(s, P, e, do_reset ? P.out.tobs : P.out.obs, rng_words, 7);
    if (do_reset) {
        reset_world(s, P, e);
        obs_pass(s, P, e, P.out.obs, nullptr);
    }
    if (P.out.alive) {
        for (int i = tid; i < L.N; i += nt) P.out.alive[(size_t)e * L.N + i] = s.flags[i] & F_ALIVE;
    }
    DCA_PROF(P, e, 10);
    store_blob(s, P, e);
    DCA_PROF(P, e, 11);
}

// reset kernel: SuperpoolEnv.reset -> collective_assaultGlobalEnv.reset parallel_run.py:248-270
__global__ void __launch_bounds__(256) k_reset(const Params P, const uint8_t *__restrict__ env_mask) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int e = blockIdx.x;
    if (env_mask != nullptr && !env_mask[e]) return;
    const Smem s = carve(smem_raw, P.L);
    load_blob(s, P, e);
    reset_world(s, P, e);
    obs_pass(s, P, e, P.out.obs, nullptr);
    if (threadIdx.x == 0) { P.out.done[e] = 0; P.out.win[e] = 0; }
    for (int g = threadIdx.x; g < P.L.G; g += blockDim.x) P.out.reward[(size_t)e * P.L.G + g] = 0.f;
    if (P.out.alive)
        for (int i = threadIdx.x; i < P.L.N; i += blockDim.x) P.out.alive[(size_t)e * P.L.N + i] = s.flags[i] & F_ALIVE;
    store_blob(s, P, e);
}

// observation-only kernel (the kNN gather kernel on the resident state)
__global__ void __launch_bounds__(256) k_observe(const Params P, const uint32_t *__restrict__ rng_words) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int e = blockIdx.x;
    const Smem s = carve(smem_raw, P.L);
    load_blob(s, P, e);
    obs_pass(s, P, e, P.out.obs, rng_words);
    // only the observation counter changed: write the scalars back
    if (threadIdx.x == 0)
        ((EnvScalars *)(P.blobs + (size_t)e * P.L.bytes + P.L.scalars))->obs_counter = s.sc->obs_counter;
}

// ------------------------------------------------------------------ state import / export (plain SoA <-> blobs)
__global__ void k_state_io(const Params P, const DcaState st, int import) {
    const BlobLayout &L = P.L;
    const size_t total = (size_t)P.E * L.N;
    for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const int e = (int)(t / L.N), i = (int)(t % L.N);
        unsigned char *b = P.blobs + (size_t)e * L.bytes;
        double *px = (double *)(b + L.px), *py = (double *)(b + L.py), *vx = (double *)(b + L.vx),
               *vy = (double *)(b + L.vy), *atk = (double *)(b + L.atk), *ter = (double *)(b + L.terrain);
        int32_t *nh = (int32_t *)(b + L.num_hit), *nw = (int32_t *)(b + L.num_was_hit);
        uint8_t *fl = b + L.flags;
        EnvScalars *sc = (EnvScalars *)(b + L.scalars);
        if (import) {
            px[i] = st.pos[2 * t]; py[i] = st.pos[2 * t + 1];
            vx[i] = st.vel[2 * t]; vy[i] = st.vel[2 * t + 1];
            atk[i] = st.atk_rad[t]; ter[i] = st.terrain[t];
            nh[i] = st.num_hit[t]; nw[i] = st.num_was_hit[t];
            fl[i] = (uint8_t)((st.alive[t] ? F_ALIVE : 0) | (st.just_died[t] ? F_JUST_DIED : 0) |
                              (st.hit[t] ? F_HIT : 0) | (st.was_hit[t] ? F_WAS_HIT : 0));
            if (i == 0) {
                sc->init_theta = st.init_theta[e]; sc->time_step = st.time_step[e];
                sc->n_alive_guards = st.n_alive_guards[e]; sc->n_alive_attackers = st.n_alive_attackers[e];
                sc->obs_counter = st.obs_counter[e]; sc->episode = st.episode[e]; sc->pad = 0;
            }
        } else {
            st.pos[2 * t] = px[i]; st.pos[2 * t + 1] = py[i];
            st.vel[2 * t] = vx[i]; st.vel[2 * t + 1] = vy[i];
            st.atk_rad[t] = atk[i]; st.terrain[t] = ter[i];
            st.num_hit[t] = nh[i]; st.num_was_hit[t] = nw[i];
            const uint8_t f = fl[i];
            st.alive[t] = (f & F_ALIVE) ? 1 : 0; st.just_died[t] = (f & F_JUST_DIED) ? 1 : 0;
            st.hit[t] = (f & F_HIT) ? 1 : 0; st.was_hit[t] = (f & F_WAS_HIT) ? 1 : 0;
            if (i == 0) {
                st.init_theta[e] = sc->init_theta; st.time_step[e] = sc->time_step;
                st.n_alive_guards[e] = sc->n_alive_guards; st.n_alive_attackers[e] = sc->n_alive_attackers;
                st.obs_counter[e] = sc->obs_counter; st.episode[e] = sc->episode;
            }
        }
    }
}

// fresh handle: counters at -1 so the first reset is episode 0 / observation pass 0; everything else zero
__global__ void k_init_blobs(const Params P) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= P.E) return;
    EnvScalars *sc = (EnvScalars *)(P.blobs + (size_t)e * P.L.bytes + P.L.scalars);
    sc->obs_counter = 0xFFFFFFFFu;
    sc->episode = 0xFFFFFFFFu;
}

// ------------------------------------------------------------------ benchmark planner stand-in (one warp per env)
// force = centroid(alive guards) - pos for alive attackers.  NOT the reference planner (cheat_script_ai.py:28-194).
__global__ void k_plan_centroid(const Params P, double *__restrict__ att_force) {
    const BlobLayout &L = P.L;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= P.E) return;
    const unsigned char *b = P.blobs + (size_t)warp * L.bytes;
    const double *px = (const double *)(b + L.px), *py = (const double *)(b + L.py);
    const uint8_t *fl = b + L.flags;
    double sx = 0, sy = 0;
    int n = 0;
    for (int g = lane; g < L.G; g += 32)
        if (fl[g] & F_ALIVE) { sx += px[g]; sy += py[g]; ++n; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sx += __shfl_xor_sync(0xffffffffu, sx, o);
        sy += __shfl_xor_sync(0xffffffffu, sy, o);
        n += __shfl_xor_sync(0xffffffffu, n, o);
    }
    const double cx = n ? sx / n : 0.0, cy = n ? sy / n : 0.0;
    double2 *out = (double2 *)att_force + (size_t)warp * L.A;
    for (int a = lane; a < L.A; a += 32) {
        const bool alive = fl[L.G + a] & F_ALIVE;
        out[a] = alive ? make_double2(cx - px[L.G + a], cy - py[L.G + a]) : make_double2(0.0, 0.0);
    }
}

// ------------------------------------------------------------------ GAE / return
// trajectory.finalize ALGORITHM/conc/trajectory.py:67-117: reward_push_forward :60-64, threat :75-86,
// gae_finalize_return :92-117.  Rollout buffers are [T,E,G]; one thread owns one (env, guard) column and scans
// time backwards (loads coalesced across the E*G columns).  dtype walk as numpy: gamma*V[t+1] in f32,
// delta / gae in f64, return stored f32.
__global__ void k_gae(const float *__restrict__ reward, const float *__restrict__ value,
                      const uint8_t *__restrict__ dead, const int32_t *__restrict__ traj_len, float *__restrict__ ret,
                      float *__restrict__ threat, float *__restrict__ reward_pushed, int T, int E, int G, double gamma,
                      double tau, int threat_limit) {
    const size_t S = (size_t)E * G;
    const size_t c = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (c >= S) return;
    const int Lc = traj_len ? traj_len[c / G] : T;
    if (Lc <= 0) return;
    const float gamma_f = (float)gamma;
    const double gt = gamma * tau;
    double carry = 0, gae = 0, th = -1;
    float v_next = 0.f;
    int d_next = 0;
    for (int i = Lc - 1; i >= 0; --i) {
        const size_t k = (size_t)i * S + c;
        const int d = dead[k] != 0;
        const float vf = value[k];
        double r = (double)reward[k] + carry;
        if (i > 0) {
            carry = r * (double)d;
            r = r * (double)(!d);
        }
        if (reward_pushed) reward_pushed[k] = (float)r;
        th += (i == Lc - 1) ? (double)(!d) : (double)(!(d_next && d));
        const double tc = th < -1 ? -1 : (th > threat_limit ? (double)threat_limit : th);
        if (threat) threat[k] = (float)tc;
        const double v = (double)vf;
        if (i == Lc - 1) {
            gae = r + gamma * 0 - v;
        } else {
            const float gv = __fmul_rn(gamma_f, v_next);
            const double delta = r + (double)gv - v;
            gae = delta + gt * gae;
        }
        ret[k] = (float)(gae + v);
        v_next = vf; d_next = d;
    }
}

}  // namespace dca
