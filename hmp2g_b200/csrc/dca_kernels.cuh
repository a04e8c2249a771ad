// dca_kernels.cuh -- the DCA step / observation / reset kernels (one CTA per env, state in shared memory).
// Reference semantics are cited per phase (paths relative to /root/reference).
//
// Every kernel is a template on <threads, G, A>: the common team sizes are compiled with the blob layout as
// compile-time constants (all offsets, loop bounds and index arithmetic fold); <.., 0, 0> reads the layout from
// the kernel parameters.
#pragma once
#include "dca_device.cuh"

namespace dca {

#ifndef DCA_KNN_SORT8
#define DCA_KNN_SORT8 1  // kNN candidates: eight at a time through a sorting network (1) or one by one (0)
#endif
#ifndef DCA_AIM_SPLIT
#define DCA_AIM_SPLIT 0  // scripted aim: a lane pair per attacker (1) or one lane (0)
#endif
#ifndef DCA_CTAS_128
#define DCA_CTAS_128 8  // resident CTAs per SM the 128-thread kernels are register-budgeted for
#endif
#ifndef DCA_CTAS_256
#define DCA_CTAS_256 4  // ... and the 256-thread kernels (N > 128 entities per env): 64 registers like the 128-thread ones (3 CTAs / 80 registers: 100v100 7 % slower)
#endif
#ifdef DCA_DIAG_COUNT
__device__ unsigned long long g_diag[8];  // diagnostics build only: kNN task / close-call counters
#endif
constexpr int kPairCap = 256;  // queued (shooter, target) laser candidates per env; overflow is tested in place
constexpr int kHintN = kVisN + 1;  // neighbours remembered per (guard, list) for the next pass

// Shared-memory view of one env: the blob (TMA destination) followed by per-step work arrays.
struct Smem {
    unsigned char *blob;
    double *px, *py, *vx, *vy, *atk, *ter;
    int32_t *nhit, *nwas;
    uint8_t *flags;
    u64 *knn_ids;
    float *hmin;
    EnvScalars *sc;
    float *xg, *yg, *xa, *ya;  // float32-rounded positions of guards / attackers (distance matrix inputs), SoA,
                               // list_pad() entries each; dead agents and padding = +inf
    int32_t *hitcnt;       // times agent i was hit this step (also: exclusion keys during reset)
    int32_t *shots;        // hits agent i scored this step
    float4 *table;         // [4][N+1] float32 entity records, one plane per 16-byte quarter of the 64-byte record (so that
                           // consecutive agents are consecutive in a plane: conflict-free stores); record N is all zero
                           // (invisible / jammed / empty slots)
    int16_t *sid;          // [G,12] entity id per obs row, -1 = zero row          } the laser work lists are dead
    int16_t *shooters;     // [N] agents that may hit something this step           } before the observation pass
    uint32_t *pairs;       // [kPairCap] (shooter << 16 | target) laser candidates  } writes sid: same bytes
    int16_t *alist;        // [N] agents alive at the start of the step, index order (guards first)
    int16_t *glist;        // [G] guards alive for the observation pass (any order)
    int32_t *misc;         // [0] guard deaths, [1] attacker deaths, [2] #shooters, [3] #pairs, [4] #alist, [5] alive
                           // guards in alist, [6] #glist, [8..15] per-warp counts of the compaction
    uint64_t *bar;
};

__host__ __device__ constexpr int smem_bytes(const BlobLayout &L) {
    int o = L.bytes;
    o += (list_pad(L.G) + list_pad(L.A)) * 8;       // xg yg xa ya
    o += 2 * align_up(L.N * 4, 16);                 // hitcnt shots
    o += (L.N + 1) * kObsRow * 4;                   // table
    const int a = align_up(L.G * kObsRows * 2, 16), b = align_up(L.N * 2, 16) + kPairCap * 4;
    o += a > b ? a : b;                             // sid | shooters + pairs
    o += align_up(L.N * 2, 16) + align_up(L.G * 2, 16);  // alist glist
    o += 64;                                        // misc
    o += 16;                                        // mbarrier
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
    s.knn_ids = (u64 *)(base + L.knn_ids); s.hmin = (float *)(base + L.hmin);
    s.sc = (EnvScalars *)(base + L.scalars);
    unsigned char *p = base + L.bytes;
    const int f32 = align_up(L.N * 4, 16);
    const int Gp = list_pad(L.G), Ap = list_pad(L.A);
    s.xg = (float *)p; p += Gp * 4; s.yg = (float *)p; p += Gp * 4;
    s.xa = (float *)p; p += Ap * 4; s.ya = (float *)p; p += Ap * 4;
    s.hitcnt = (int32_t *)p; p += f32; s.shots = (int32_t *)p; p += f32;
    s.table = (float4 *)p; p += (L.N + 1) * kObsRow * 4;
    {
        const int a = align_up(L.G * kObsRows * 2, 16), b = align_up(L.N * 2, 16) + kPairCap * 4;
        s.sid = (int16_t *)p;
        s.shooters = (int16_t *)p;
        s.pairs = (uint32_t *)(p + align_up(L.N * 2, 16));
        p += a > b ? a : b;
    }
    s.alist = (int16_t *)p; p += align_up(L.N * 2, 16);
    s.glist = (int16_t *)p; p += align_up(L.G * 2, 16);
    s.misc = (int32_t *)p; p += 64;
    s.bar = (uint64_t *)p;
    return s;
}

// Fetch the env blob with one TMA bulk copy; returns after the data has landed for every thread.
// wait_on / want: DCA_OPT_PIPELINE_STEPS -- the blob is only fetched once the previous step launch has published env e.
__device__ __forceinline__ void load_blob(const Smem &s, const Params &P, const BlobLayout &L, int e,
                                          const uint32_t *wait_on = nullptr, uint32_t want = 0) {
    if (threadIdx.x == 0) {  // the copy is in flight before anybody else reaches the barrier below
        if (wait_on != nullptr) {
            pipe_wait(wait_on, want);
            fence_proxy_async_all();  // the blob was written through the async proxy (bulk store) and is read through it
        }
        mbar_init(s.bar, 1);
        fence_barrier_init();
        mbar_expect_tx(s.bar, (uint32_t)L.bytes);
        tma_load_1d(s.blob, P.blobs + (size_t)e * L.bytes, (uint32_t)L.bytes, s.bar);
    }
    __syncthreads();  // the mbarrier is initialised for every waiter
    mbar_wait(s.bar, 0);
}

// Write the blob back with one TMA bulk store.  Generic-proxy writes are fenced into the async proxy first.
// publish_to / value: the env's epoch word, stored with release semantics once the bulk store has completed (every
// thread's output stores precede the barrier, so the release covers them too).
__device__ __forceinline__ void store_blob(const Smem &s, const Params &P, const BlobLayout &L, int e,
                                           uint32_t *publish_to = nullptr, uint32_t value = 0) {
    fence_proxy_async();
    __syncthreads();
    if (threadIdx.x == 0) {
        tma_store_1d(P.blobs + (size_t)e * L.bytes, s.blob, (uint32_t)L.bytes);
        if (publish_to == nullptr) {
            tma_store_commit_wait();
        } else {
            tma_store_commit_wait_written();
            fence_proxy_async_all();
            __threadfence();
            st_release_u32(publish_to, value);
        }
    }
}

// End of a CTA whose blob store was issued early (obs_pass early_store): wait until shared memory has been read out --
// or, when the env is handed to the next launch (publish_to), until the blob is written, every thread's output stores
// are issued (barrier) and then release the env's epoch word.
__device__ __forceinline__ void finish_blob(uint32_t *publish_to, uint32_t value) {
    if (publish_to == nullptr) {
        if (threadIdx.x == 0) tma_store_wait_read();
        return;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        tma_store_wait_written();
        fence_proxy_async_all();
        __threadfence();
        st_release_u32(publish_to, value);
    }
}

// Optional per-CTA phase clocks (diagnostics, dca_debug_phase_clocks): slot k of env e <- clock64() of thread 0.
#define DCA_PROF(P, e, k)                                                                     \
    do {                                                                                      \
        if ((P).prof != nullptr && threadIdx.x == 0) (P).prof[(size_t)(e) * 16 + (k)] = clock64(); \
    } while (0)

// ------------------------------------------------------------------ terrain
// World.set_terrain_adv core.py:224-234 + get_terrain core.py:306-314; alive agents only, dead keep old value.
__device__ __forceinline__ double terrain_fn(double tA, double tB, double X, double Y, double ct, double st) {
    const double X_ = X * ct + Y * st;
    const double Y_ = -X * st + Y * ct;
    const double a = 0.1 * X_, b = 0.1 * Y_;
    double Z = -1 + tB * (a * a + b * b);
    // cos(2 pi t) as cospi(2 t): same value to ~1e-16 (the reference rounds 2 pi t first), cheaper range reduction
    Z = Z - tA * cospi(2 * (0.3 * X_));
    Z = Z - tA * cospi(2 * (0.5 * Y_));
    return -Z;
}
__device__ __forceinline__ double terrain_at(const DcaConfig &c, double X, double Y, double ct, double st) {
    if (!c.introduce_terrain) return 1.0;
    return terrain_fn(c.terrain_A, c.terrain_B, X, Y, ct, st);
}

// Padding slots of the float32 position lists (+inf) and the all-zero table row; once per kernel.
__device__ __forceinline__ void init_pads(const Smem &s, const BlobLayout &L, const Params &P, bool lists = true,
                                          bool zero_row = true) {
    const int t = threadIdx.x;
    const int ng = list_pad(L.G) - L.G, na = list_pad(L.A) - L.A;  // 1..4 each
    if (t < ng) { if (lists) { s.xg[L.G + t] = CUDART_INF_F; s.yg[L.G + t] = CUDART_INF_F; } }
    else if (t < ng + na) { if (lists) { s.xa[L.A + t - ng] = CUDART_INF_F; s.ya[L.A + t - ng] = CUDART_INF_F; } }
    else if (t < ng + na + 4 && zero_row) {
        // the conc / conc_4hist policies turn all-zero rows into NaN first thing (ALGORITHM/conc/shell_env.py:72-75):
        // with DCA_OPT_EMPTY_ROWS_NAN the kernel hands them over that way
        const float z = P.empty_nan ? __int_as_float(0x7fc00000) : 0.f;
        s.table[(t - ng - na) * (L.N + 1) + L.N] = make_float4(z, z, z, z);
    }
}

// float32 position of agent i in its team list (distance matrix input; dead = +inf)
__device__ __forceinline__ void put_xy(const Smem &s, const BlobLayout &L, int i, bool alive, float x, float y) {
    const float lx = alive ? x : CUDART_INF_F, ly = alive ? y : CUDART_INF_F;
    if (i < L.G) { s.xg[i] = lx; s.yg[i] = ly; }
    else { s.xa[i - L.G] = lx; s.ya[i - L.G] = ly; }
}


// Index-ordered list of the alive agents in [lo, hi): list[0..n), returns n (to every thread).
// Warps that hold no alive agent then issue nothing in the loops over the list -- dead agents cost no issue slots.
// CTA-wide; ends with a barrier (the list is visible to everyone on return).
// *n_below (optional): how many of them have an index below `split` (the alive guards, for split = G).
__device__ __forceinline__ int compact_alive(const Smem &s, int lo, int hi, int16_t *list, int split = 0,
                                             int *n_below = nullptr) {
    const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, warp = tid >> 5, nw = nt >> 5;
    int base = 0, below = 0;
    for (int r0 = lo; r0 < hi; r0 += nt) {  // uniform
        const int i = r0 + tid;
        const bool alive = i < hi && (s.flags[i] & F_ALIVE);
        const uint32_t b = __ballot_sync(0xffffffffu, alive);
        const uint32_t bb = __ballot_sync(0xffffffffu, alive && i < split);
        if (r0 != lo) __syncthreads();  // the previous round's counts have been read
        if (lane == 0) s.misc[8 + warp] = __popc(b) | (__popc(bb) << 16);
        __syncthreads();
        int off = base, tot = 0;
        for (int w = 0; w < nw; ++w) {
            const int cw = s.misc[8 + w];
            off += w < warp ? (cw & 0xffff) : 0;
            tot += cw;
        }
        if (alive) list[off + __popc(b & ((1u << lane) - 1u))] = (int16_t)i;
        base += tot & 0xffff;
        below += tot >> 16;
    }
    __syncthreads();
    if (n_below) *n_below = below;
    return base;
}

// Entity record of agent i, collective_assault_env.py:237-276: [alive,x,y,atk_rad,vx,vy,iden,terrain,bit0..bit7] as
// float32, plus its list position.
__device__ __forceinline__ void build_row(const Smem &s, const BlobLayout &L, int i) {
    float4 *row = s.table + i;
    const int ps = L.N + 1;  // plane stride
    const float x = (float)s.px[i], y = (float)s.py[i];
    const bool alive_i = s.flags[i] & F_ALIVE;
    put_xy(s, L, i, alive_i, x, y);
    row[0] = make_float4(alive_i ? 1.f : 0.f, x, y, (float)s.atk[i]);
    row[ps] = make_float4((float)s.vx[i], (float)s.vy[i], (float)i, (float)s.ter[i]);
    // get_binary_array :207-215 (float halving, int8 wrap, floored %) yields the plain bits for 0 <= i < 256
    float b[8];
    if (i < 256) {
#pragma unroll
        for (int k = 0; k < 8; ++k) b[k] = (float)((i >> k) & 1);
    } else {
        int v = i;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            b[k] = (float)((((v % 2) + 2) % 2) == 1);
            v = (int)(signed char)(int)((double)v / 2.0);
        }
    }
    row[2 * ps] = make_float4(b[0], b[1], b[2], b[3]);
    row[3 * ps] = make_float4(b[4], b[5], b[6], b[7]);
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

// Conservative candidate mask of list entries [c0, c0 + 4*nq), nq <= 8 quads: bit j set if the (approximate, packed)
// squared distance of candidate c0+j is <= thr.  Warp-uniform addresses -> shared-memory broadcasts.
__device__ __forceinline__ uint32_t scan_chunk(const float *lx, const float *ly, int c0, int nq, u64 px2, u64 py2,
                                               float thr) {
    const ulonglong2 *x4 = (const ulonglong2 *)(lx + c0), *y4 = (const ulonglong2 *)(ly + c0);
    uint32_t m = 0;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        if (q < nq) {
            const ulonglong2 xs = x4[q], ys = y4[q];
            const u64 s01 = sqdist2(xs.x, ys.x, px2, py2), s23 = sqdist2(xs.y, ys.y, px2, py2);
            m |= (lo2(s01) <= thr) ? (1u << (4 * q)) : 0u;
            m |= (hi2(s01) <= thr) ? (2u << (4 * q)) : 0u;
            m |= (lo2(s23) <= thr) ? (4u << (4 * q)) : 0u;
            m |= (hi2(s23) <= thr) ? (8u << (4 * q)) : 0u;
        }
    }
    return m;
}

// Exact (rare) path of one kNN task: the candidates inside `thr` ordered by (sqrt_rn distance, index) exactly as
// argsort()[:6] + range mask would (collective_assault_env.py:308-309,320-322,331-333); taken when two keys of the
// fast path are too close to call.  Returns the visible ids packed one byte each, nearest first.
__device__ __noinline__ unsigned long long knn_exact(const float *lx, const float *ly, int cnt, float pgx, float pgy,
                                                      float thr, float q0, int *n_out) {
    float bd[kVisN];
    int bi[kVisN];
#pragma unroll
    for (int k = 0; k < kVisN; ++k) { bd[k] = CUDART_INF_F; bi[k] = -1; }
    float lim = CUDART_INF_F;  // upper bound of the squares that can still enter the list
    const u64 px2 = pack2(pgx, pgx), py2 = pack2(pgy, pgy);
    const int nquad = (cnt + 3) >> 2;
    for (int c0 = 0; c0 < cnt; c0 += 32) {
        uint32_t m = scan_chunk(lx, ly, c0, min(8, nquad - (c0 >> 2)), px2, py2, thr);
        while (m) {
            const int j = c0 + __ffs(m) - 1;
            m &= m - 1;
            const float sq = f32_sqdist(pgx, pgy, lx[j], ly[j]);
            // a candidate enters only if it is strictly nearer than the current 6th (ties keep the earlier, lower
            // index): cheap bound on the square first, then the exact distance -- a stack of equidistant
            // candidates costs a compare each, not an insertion
            if (sq <= q0 && sq <= lim) {
                const float d = __fsqrt_rn(sq);
                if (d < bd[kVisN - 1]) {
                    top6_insert(bd, bi, d, j);
                    lim = __fmul_ru(__fmul_ru(bd[kVisN - 1], bd[kVisN - 1]), 1.000001f);  // inf while fewer than 6
                }
            }
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

// Close call among the 7 kept keys only (the usual one): their exact (sqrt_rn distance, index) order decides, no
// other candidate can reach the first six.  b0..b6 = the kept keys (index in the low bits, top bit = empty slot).
__device__ __noinline__ unsigned long long knn_local_exact(const float *lx, const float *ly, float pgx, float pgy,
                                                            uint32_t b0, uint32_t b1, uint32_t b2, uint32_t b3,
                                                            uint32_t b4, uint32_t b5, uint32_t b6, uint32_t im,
                                                            int *n_out) {
    const uint32_t b[kHintN] = {b0, b1, b2, b3, b4, b5, b6};
    u64 k[kHintN];
#pragma unroll
    for (int t = 0; t < kHintN; ++t) {
        const bool valid = (int32_t)b[t] >= 0;
        const int j = valid ? (int)(b[t] & im) : 0;
        const float d = __fsqrt_rn(f32_sqdist(pgx, pgy, lx[j], ly[j]));
        k[t] = valid ? (((u64)__float_as_uint(d) << 32) | (u64)j) : ~0ull;
    }
#define DCA_CEX(i, j) { const u64 lo_ = k[i] < k[j] ? k[i] : k[j]; k[j] = k[i] < k[j] ? k[j] : k[i]; k[i] = lo_; }
    DCA_CEX(0, 6) DCA_CEX(2, 3) DCA_CEX(4, 5)
    DCA_CEX(0, 2) DCA_CEX(1, 4) DCA_CEX(3, 6)
    DCA_CEX(0, 1) DCA_CEX(2, 5) DCA_CEX(3, 4)
    DCA_CEX(1, 2) DCA_CEX(4, 6)
    DCA_CEX(2, 3) DCA_CEX(4, 5)
    DCA_CEX(1, 2) DCA_CEX(3, 4) DCA_CEX(5, 6)
#undef DCA_CEX
    unsigned long long vis = 0;
    int n = 0;
#pragma unroll
    for (int t = 0; t < kVisN; ++t)
        if (k[t] != ~0ull) { vis |= (k[t] & 0xffull) << (8 * n); ++n; }
    *n_out = n;
    return vis;
}

// One kNN task: the (up to) 6 nearest in-range candidates of a list, nearest first, ties -> lower index.
//   argsort()[:6] then the range mask == the 6 nearest IN-RANGE candidates (anything farther than an out-of-range
//   member of the top-6 is out of range too).  q0 = largest f32 s with sqrt_rn(s) <= range, so "s <= q0" is the
//   reference's "d <= range" exactly.
// (1) Scan radius: the 7 neighbours remembered from the last pass (blob field knn_ids) are re-measured; if all are
//     still alive, the 6 nearest now lie within the farthest of them, so only that disc (or the obs range, if
//     smaller) has to be searched.  A dead / missing neighbour measures +inf -> the full range is searched.  The hint
//     only shrinks the search: any 7 live candidates bound the 6th nearest, whatever they are.
// (2) Packed conservative scan of the list against that radius -> bit mask (typically 7-9 bits).
// (3) The masked candidates get the exact squared distance and go through a 7-slot min/max insertion network on
//     32-bit keys (s bits with the low idx_bits replaced by the index).  sqrt_rn is monotone, so key order ==
//     (distance, index) order unless two of the 7 smallest keys are within one key step -- then knn_exact() decides.
__device__ __forceinline__ unsigned long long knn_task(const float *lx, const float *ly, int cnt, float pgx, float pgy,
                                                        u64 prev, float q0, int idx_bits, int *n_out, u64 *ids_out,
                                                        float *near_out) {
    const uint32_t im = (1u << idx_bits) - 1u;
    const float q0c = __fmul_ru(q0, 1.000001f);  // conservative range for the approximate scan
    float tub = 0.f;
#pragma unroll
    for (int k = 0; k < kHintN; ++k) {
        const int id = (int)((prev >> (8 * k)) & 0xffu);
        tub = fmaxf(tub, f32_sqdist(pgx, pgy, lx[id], ly[id]));
    }
    const float thr = fminf(__fmul_ru(tub, 1.000001f), q0c);
    const u64 px2 = pack2(pgx, pgx), py2 = pack2(pgy, pgy);
    uint32_t b[kHintN];
    uint32_t drop = 0xFFFFFFFFu;  // smallest key that did not make it into b[] (the 8th smallest overall)
    const int nquad = (cnt + 3) >> 2;
#if DCA_KNN_SORT8
    // Exact key of candidate j (list padding measures +inf -> "no key").
    auto key_of = [&](int j) -> uint32_t {
        const float sq = f32_sqdist(pgx, pgy, lx[j], ly[j]);
        return sq <= q0 ? ((__float_as_uint(sq) & ~im) | (uint32_t)j) : 0xFFFFFFFFu;
    };
    // The first 64 candidates: both chunk masks in one word.  When some task of the warp has more than a handful of
    // candidates (warp-uniform test: the callers keep all 32 lanes in the task loop), the first eight of each lane are
    // evaluated side by side -- independent loads and arithmetic instead of a loop-carried chain -- and ordered by a
    // 19-exchange sorting network; the 7 smallest seed the insertion slots.  Whatever is left (and any candidate
    // beyond the first 64) goes through the insertion chain below.
    u64 m = (u64)scan_chunk(lx, ly, 0, min(8, nquad), px2, py2, thr);
    if (cnt > 32) m |= (u64)scan_chunk(lx, ly, 32, min(8, nquad - 8), px2, py2, thr) << 32;
    if (__reduce_max_sync(0xffffffffu, (unsigned)__popcll(m)) > 3u) {
        uint32_t k[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) {
            const int j = m ? __ffsll((long long)m) - 1 : cnt;
            m &= m - 1;
            k[t] = key_of(j);
        }
#define DCA_CEX(i, j) { const uint32_t lo_ = min(k[i], k[j]); k[j] = max(k[i], k[j]); k[i] = lo_; }
        DCA_CEX(0, 2) DCA_CEX(1, 3) DCA_CEX(4, 6) DCA_CEX(5, 7)
        DCA_CEX(0, 4) DCA_CEX(1, 5) DCA_CEX(2, 6) DCA_CEX(3, 7)
        DCA_CEX(0, 1) DCA_CEX(2, 3) DCA_CEX(4, 5) DCA_CEX(6, 7)
        DCA_CEX(2, 4) DCA_CEX(3, 5)
        DCA_CEX(1, 4) DCA_CEX(3, 6)
        DCA_CEX(1, 2) DCA_CEX(3, 4) DCA_CEX(5, 6)
#undef DCA_CEX
#pragma unroll
        for (int t = 0; t < kHintN; ++t) b[t] = k[t];
        drop = k[7];
    } else {
#pragma unroll
        for (int t = 0; t < kHintN; ++t) b[t] = 0xFFFFFFFFu;
    }
    for (int c0 = 0; c0 < cnt; c0 += 64) {
        if (c0 > 0) {
            m = (u64)scan_chunk(lx, ly, c0, min(8, nquad - (c0 >> 2)), px2, py2, thr);
            if (cnt > c0 + 32) m |= (u64)scan_chunk(lx, ly, c0 + 32, min(8, nquad - 8 - (c0 >> 2)), px2, py2, thr) << 32;
        }
        while (m) {
            uint32_t key = key_of(c0 + __ffsll((long long)m) - 1);
            m &= m - 1;
            if (key < b[kHintN - 1]) {  // anything else would fall straight through the chain
#pragma unroll
                for (int t = 0; t < kHintN; ++t) {
                    const uint32_t lo = min(b[t], key);
                    key = max(b[t], key);
                    b[t] = lo;
                }
            }
            drop = min(drop, key);  // what fell off the end (or never entered)
        }
    }
#else
#pragma unroll
    for (int k = 0; k < kHintN; ++k) b[k] = 0x80000000u + ((uint32_t)k << 8);  // empty slots: distinct, ascending
    for (int c0 = 0; c0 < cnt; c0 += 32) {
        uint32_t m = scan_chunk(lx, ly, c0, min(8, nquad - (c0 >> 2)), px2, py2, thr);
        while (m) {
            const int j = c0 + __ffs(m) - 1;
            m &= m - 1;
            const float sq = f32_sqdist(pgx, pgy, lx[j], ly[j]);
            uint32_t key = sq <= q0 ? ((__float_as_uint(sq) & ~im) | (uint32_t)j) : 0xFFFFFFFFu;
#pragma unroll
            for (int k = 0; k < kHintN; ++k) {
                const uint32_t lo = min(b[k], key);
                key = max(b[k], key);
                b[k] = lo;
            }
            drop = min(drop, key);
        }
    }
#endif
    bool close_call = false;
#pragma unroll
    for (int k = 0; k < kVisN; ++k)
        close_call |= ((int32_t)b[k + 1] >= 0) && ((b[k + 1] >> idx_bits) - (b[k] >> idx_bits) <= 1u);
    unsigned long long vis = 0;
    int n = 0;
#ifdef DCA_DIAG_NOEXACT
    close_call = false;
#endif
#ifdef DCA_DIAG_NEVEREXACT
    close_call = close_call && q0 < 0.f;  // never true at run time, unknown at compile time
#endif
#ifdef DCA_DIAG_COUNT
    atomicAdd(&g_diag[0], 1ull);
    if (close_call) {
        atomicAdd(&g_diag[1], 1ull);
        int nc = 0;
        for (int j = 0; j < cnt; ++j) nc += f32_sqdist(pgx, pgy, lx[j], ly[j]) <= thr ? 1 : 0;
        atomicAdd(&g_diag[2], (unsigned long long)nc);
        int same = 0;
#pragma unroll
        for (int k = 0; k < kVisN; ++k) same += ((int32_t)b[k + 1] >= 0 && (b[k + 1] >> idx_bits) == (b[k] >> idx_bits)) ? 1 : 0;
        atomicAdd(&g_diag[3], same > 0 ? 1ull : 0ull);
        atomicAdd(&g_diag[4], cnt == 50 && lx != nullptr ? 1ull : 0ull);
    }
#endif
    if (close_call) {
        // Candidates outside b[] matter only if the tie chain runs through the 6th/7th kept key into the best dropped
        // one; otherwise every dropped candidate is strictly farther (>= one whole key step in the square) than the
        // first six, and the exact order of the kept keys settles it.
        const bool c56 = ((int32_t)b[6] >= 0) && ((b[6] >> idx_bits) - (b[5] >> idx_bits) <= 1u);
        const bool c6d = ((int32_t)drop >= 0) && ((drop >> idx_bits) - (b[6] >> idx_bits) <= 1u);
        if (c56 && c6d) vis = knn_exact(lx, ly, cnt, pgx, pgy, thr, q0, &n);
        else vis = knn_local_exact(lx, ly, pgx, pgy, b[0], b[1], b[2], b[3], b[4], b[5], b[6], im, &n);
    } else {
#pragma unroll
        for (int k = 0; k < kVisN; ++k)
            if ((int32_t)b[k] >= 0) { vis |= (unsigned long long)(b[k] & im) << (8 * n); ++n; }
    }
    *n_out = n;
    u64 ids = 0;
#pragma unroll
    for (int k = 0; k < kHintN; ++k)
        ids |= (u64)((int32_t)b[k] >= 0 ? (b[k] & im) : (uint32_t)cnt) << (8 * k);
    *ids_out = ids;
    *near_out = (int32_t)b[0] >= 0 ? __uint_as_float(b[0] & ~im) : q0;  // lower bound of the nearest squared distance
    return vis;
}

// ------------------------------------------------------------------ observation pass
// collective_assaultEnvV1.observation collective_assault_env.py:228-359 for every guard of the env.
// `counter` = number of this pass in the env's keyed RNG stream (stored as the new obs_counter).
// dest == nullptr: the pass is only counted (keeps the keyed RNG stream aligned), nothing is computed.
// prebuilt: the entity table and the position lists already describe the current state and the CTA has synchronised
// since they were written (k_step builds them while integrating).  No barrier at the end: the caller synchronises
// before shared memory is written again.
// early_store: the caller promises not to touch the env's blob image again; it is written back (one bulk store, issued
// by thread 0, not waited for) as soon as the observation tasks have left their hints in it, so the store runs under the
// gather; the caller ends with finish_blob().
__device__ __forceinline__ void obs_pass(const Smem &s, const Params &P, const BlobLayout &L, int e, float *dest,
                                         const uint32_t *rng_words, uint32_t counter, bool prebuilt, int prof_slot,
                                         bool early_store = false) {
    const int tid = threadIdx.x, nt = blockDim.x;
    if (dest == nullptr) {
        if (tid == 0) s.sc->obs_counter = counter;  // callers have synchronised since every thread read the old value
        if (early_store) {
            fence_proxy_async();
            __syncthreads();
            if (tid == 0) {
                fence_proxy_async();
                tma_store_1d(P.blobs + (size_t)e * L.bytes, s.blob, (uint32_t)L.bytes);
                tma_store_commit();
            }
        }
        return;
    }
    int n_obs_guards = s.misc[6];  // prebuilt: the caller's list of guards alive now
    if (!prebuilt) {
        init_pads(s, L, P);
        for (int i = tid; i < L.N; i += nt) build_row(s, L, i);
        {   // dead guards: zeros (:313-316), no RNG draws -- every row starts as "empty" (12 int16 = three 8-byte words)
            uint2 *w = (uint2 *)s.sid;
            for (int t = tid; t < L.G * (kObsRows / 4); t += nt) w[t] = make_uint2(0xFFFFFFFFu, 0xFFFFFFFFu);
        }
        __syncthreads();
        n_obs_guards = compact_alive(s, 0, L.G, s.glist);
    }
    DCA_PROF(P, e, prof_slot);

    const uint32_t k0 = (uint32_t)P.cfg.seed, k1 = (uint32_t)(P.cfg.seed >> 32);
    const uint32_t gid = (uint32_t)(P.cfg.env_gid0 + e);
    // One task per (alive guard, list), packed densely over the threads: warps beyond the last task issue nothing.
    // (Dead guards' rows were preset to "empty"; their hints are not needed before the next reset rewrites them.)
    // Hostile-list tasks first, friend-list tasks after them: a warp then holds tasks of one kind, and the hostile ones
    // (often no candidate in range, no RNG draws) do not sit idle in the lanes next to a friend list's candidate loop.
    // All 32 lanes of a warp stay in the loop (knn_task votes across the warp): lanes past the last task repeat it
    // and write nothing.
    const int n_tasks = 2 * n_obs_guards;
    for (int task0 = tid & ~31; task0 < n_tasks; task0 += nt) {
        const bool valid = task0 + (tid & 31) < n_tasks;
        const int task = valid ? task0 + (tid & 31) : n_tasks - 1;
        const int lst = task >= n_obs_guards ? 1 : 0;  // 0 = hostile (:318-328), 1 = friend (:330-342)
        const int g = s.glist[task - (lst ? n_obs_guards : 0)];
        int16_t *sid = s.sid + g * kObsRows + (lst == 0 ? kVisN : 0);  // friend rows 0..5, hostile rows 6..11 (:344)
        const int cnt = lst == 0 ? L.A : L.G;
        const int hint_slot = 2 * g + lst;
        const int base = lst == 0 ? L.G : 0;
        int n = 0;
        u64 ids;
        float nearest;
        unsigned long long vis = knn_task(lst == 0 ? s.xa : s.xg, lst == 0 ? s.ya : s.yg, cnt, s.xg[g], s.yg[g],
                                          s.knn_ids[hint_slot], P.smax[lst], P.idx_bits, &n, &ids, &nearest);
        if (!valid) continue;
        s.knn_ids[hint_slot] = ids;
        if (lst == 0) s.hmin[g] = nearest;
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
#ifdef DCA_DIAG_NOPHILOX
                const int nblk = 0;
#else
                const int nblk = n <= 2 ? 1 : (n <= 4 ? 2 : 3);
#endif
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
                // rand() < random_jam_prob with rand() = w * 2^-32, as an integer compare
                if (k < n) n_mv += ((unsigned long long)w[wi] < P.jam_thresh) ? 1 : 0;
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
    if (early_store) fence_proxy_async();  // this thread's writes to the blob image -> visible to the bulk store
    __syncthreads();
    DCA_PROF(P, e, prof_slot + 1);
    if (tid == 0) {
        s.sc->obs_counter = counter;
        if (early_store) {
            fence_proxy_async();
            tma_store_1d(P.blobs + (size_t)e * L.bytes, s.blob, (uint32_t)L.bytes);
            tma_store_commit();
        }
    }
    // gather: 12 records of 64 B per guard; a pair of threads copies one record, each thread its 32-byte half with
    // one 256-bit store (STG.256), so a warp instruction writes 1 KB of contiguous obs.  Empty slots read the zero
    // record.  Half-record index of the k-th round = tid + k * nt.
    {
        float *op = dest + (size_t)e * L.G * kObsLen + tid * 8;
        const float4 *tp = s.table + (tid & 1) * 2 * (L.N + 1);  // planes 0,1 (first half) or 2,3 (second half)
        const int16_t *sp = s.sid + (tid >> 1);
        const int total = L.G * kObsRows * 2, rstep = nt >> 1, ps = L.N + 1;
        int k = 0;
        for (; (k + 4) * nt <= total; k += 4) {  // 8 independent shared loads in flight before the stores
            float4 v[4][2];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int id = min((int)(uint16_t)sp[(k + u) * rstep], L.N);
                v[u][0] = tp[id]; v[u][1] = tp[id + ps];
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) st_global_256(op + (size_t)(k + u) * nt * 8, v[u][0], v[u][1]);
        }
        for (; k * nt + tid < total; ++k) {
            const int id = min((int)(uint16_t)sp[k * rstep], L.N);
            st_global_256(op + (size_t)k * nt * 8, tp[id], tp[id + ps]);
        }
    }
    if (P.out.slot_id) {
        int16_t *o = P.out.slot_id + (size_t)e * L.G * kObsRows;
        for (int t = tid; t < L.G * kObsRows; t += nt) o[t] = s.sid[t];
    }
    DCA_PROF(P, e, prof_slot + 2);
}

// ------------------------------------------------------------------ reset
// collective_assaultEnvV1.reset_world collective_assault_env.py:55-136 with keyed draws, then the terrain pass
// of collective_assaultGlobalEnv.reset parallel_run.py:258.  Caller runs obs_pass afterwards.
__device__ __forceinline__ void reset_world(const Smem &s, const Params &P, const BlobLayout &L, int e) {
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
        s.sc->cos_theta = ct;
        s.sc->sin_theta = sn;
    }
    // positions jump: the remembered neighbours are void.  Hostile lists: no hint (every entry = the marker slot, which
    // measures +inf).  Friend lists: the formation puts guards in index order along a line (:115-119), so the 7 index
    // neighbours of a guard are a good hint -- any 7 distinct live candidates bound the search radius, and all guards
    // are alive after a reset.
    for (int k = tid; k < 2 * L.G; k += nt) {
        u64 ids = 0x0101010101010101ull * (u64)((k & 1) ? L.G : L.A);
        if ((k & 1) && L.G >= kHintN) {
            const int g = k >> 1, lo = min(max(g - kHintN / 2, 0), L.G - kHintN);
            ids = 0;
#pragma unroll
            for (int t = 0; t < kHintN; ++t) ids |= (u64)(lo + t) << (8 * t);
        }
        s.knn_ids[k] = ids;
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

// alive mask output [E,N] u8 (4 agents per 32-bit store when the row is 4-byte aligned)
__device__ __forceinline__ void write_alive(const Smem &s, const Params &P, const BlobLayout &L, int e) {
    if (!P.out.alive) return;
    const int N = L.N;
    uint8_t *dst = P.out.alive + (size_t)e * N;
    if ((N & 3) == 0 && (((uintptr_t)dst) & 3) == 0) {
        const uint32_t *f4 = (const uint32_t *)s.flags;  // flags sit at a 16-byte aligned blob offset
        for (int k = threadIdx.x; k < N / 4; k += blockDim.x) ((uint32_t *)dst)[k] = f4[k] & 0x01010101u;
    } else {
        for (int i = threadIdx.x; i < N; i += blockDim.x) dst[i] = s.flags[i] & F_ALIVE;
    }
}

// soft wall, core.py:339-353: contact_force * softplus penetration (k = contact_margin); noinline: near-wall agents only
__device__ __noinline__ double wall_push(double d) {
    const double kk = 1e-10;
    if (d > 1e-7) return 0.0;  // -d/kk < -1000: softplus is exactly 0 (dca_device.cuh logaddexp0); skips the fp64 divide
    return 100.0 * (logaddexp0(-d / kk) * kk);
}

// ------------------------------------------------------------------ laser: exact pair test
// laser_hit_improve3 cython_func.pyx:22-46 for shooter i -> target j (both alive, opposite teams); records the hit.
__device__ __noinline__ void laser_test(const double *px, const double *py, const double *atk, const double *ter,
                                        int32_t *shots, int32_t *hitcnt, double base_radius, double win, int i, int j) {
    const double radius = base_radius * ter[i];
    const double r2 = radius * radius;
    const double dx = px[j] - px[i], dy = py[j] - py[i];
    const double d2 = dx * dx + dy * dy;
    if (d2 > r2) return;
    const double dir = atk[i];
    const double pos_r = dir + win / 2, neg_r = dir - win / 2;
    const double tt = atan2(dy, dx);
    const double d1 = fabs(reg_angle(pos_r - tt)), d2r = fabs(reg_angle(neg_r - tt));
    if (d1 <= win && d2r <= win) {
        atomicAdd(&shots[i], 1);
        atomicAdd(&hitcnt[j], 1);
    }
}
__device__ __forceinline__ void laser_pair(const Smem &s, const DcaConfig &c, int i, int j) {
    laser_test(s.px, s.py, s.atk, s.ter, s.shots, s.hitcnt, c.agent_shoot_base_radius, c.agent_shoot_win, i, j);
}

// Conservative float32 bound of shooter i's squared laser radius: a pair whose f32 squared distance exceeds r^2 by
// 1% + 1e-3 cannot pass the exact fp64 test (f32 rounding of |coord| <= 64 perturbs d^2 by < 1e-4).  +inf = no bound.
__device__ __forceinline__ float laser_r2f(const Smem &s, const DcaConfig &c, int i) {
    const double radius = c.agent_shoot_base_radius * s.ter[i];
    const double r2 = radius * radius;
    const bool ok = fabs(s.px[i]) <= 64.0 && fabs(s.py[i]) <= 64.0 && r2 < 1.0e3;
    return ok ? __fadd_ru(__fmul_ru((float)r2, 1.01f), 1.0e-3f) : CUDART_INF_F;
}

__device__ __forceinline__ void push_shooter(const Smem &s, int i) {
    s.shooters[atomicAdd(&s.misc[2], 1)] = (int16_t)i;
}

// ------------------------------------------------------------------ fused step kernel
// collective_assaultGlobalEnv.step parallel_run.py:192-246 + EnvAutoReset.step UTIL/shm_env.py:33-74.
template <int NT, int TG, int TA>
__global__ void __launch_bounds__(NT, NT == 128 ? DCA_CTAS_128 : DCA_CTAS_256) k_step(
    const Params P, const int32_t *__restrict__ actions, const double *__restrict__ att_force,
    const uint8_t *__restrict__ skip, const uint32_t *__restrict__ rng_words, int autoreset, const Pipe pipe) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int e = blockIdx.x;
#ifndef DCA_NO_STEP_TRIGGER
    pdl_trigger();  // DCA_OPT_PIPELINE_STEPS: the planner launched behind this step may become resident in its last wave
#endif
    uint32_t *const my_epoch = pipe.epoch != nullptr ? pipe.epoch + e : nullptr;
    if (skip != nullptr) {
        pdl_wait();  // the pause mask may come from the kernel right before this one: no overlap when it is given
        if (skip[e]) {  // ENV-PAUSE shm_env.py:37-41: cached outputs stay in place
            if (my_epoch != nullptr && threadIdx.x == 0) st_release_u32(my_epoch, pipe.publish);
            return;
        }
    }
    const BlobLayout L = TG > 0 ? make_layout(TG, TA) : P.L;
    const DcaConfig &c = P.cfg;
    const Smem s = carve(smem_raw, L);
    const int tid = threadIdx.x, nt = NT;
    DCA_PROF(P, e, 0);
    if (tid < 16) s.misc[tid] = 0;  // ordered before any use by the barrier inside load_blob
    load_blob(s, P, L, e, pipe.wait_step ? my_epoch : nullptr, pipe.wait_step);
    DCA_PROF(P, e, 1);

    // env scalars into registers: nobody reads them from shared memory again, so thread 0 may update them freely
    const double ct = s.sc->cos_theta, sn = s.sc->sin_theta;
    const int nG0 = s.sc->n_alive_guards, nA0 = s.sc->n_alive_attackers, ts = s.sc->time_step;
    const uint32_t obs_counter = s.sc->obs_counter;
    init_pads(s, L, P);
    // phase A0: dead agents are settled here once (no position in the distance lists; justDied cleared, core.py:266-268
    // -- nothing reads it before the reward, where a long-dead agent must show False); everything after this works on
    // the compact index-ordered list of the agents alive at the start of the step (guards first).
    for (int i = tid; i < L.N; i += nt) {
        const uint8_t f = s.flags[i];
        s.hitcnt[i] = 0; s.shots[i] = 0;
        if (!(f & F_ALIVE)) {
            put_xy(s, L, i, false, 0.f, 0.f);
            s.flags[i] = f & (uint8_t)~F_JUST_DIED;
            if (i >= L.G && P.out.nearest) P.out.nearest[(size_t)e * L.A + (i - L.G)] = -1;
        }
    }
    int n_alive_g = 0;  // alive guards = the entries of alist below G (index order: they come first)
    const int n_alive = compact_alive(s, 0, L.N, s.alist, L.G, &n_alive_g);
    // phase A: terrain (core.py:224-234), clear hit flags (core.py:239-242), float32 position lists of the pre-move
    // state.  (Action decode and the injected planner force are read where they are used, in phase D.)
    for (int q = tid; q < n_alive; q += nt) {
        const int i = s.alist[q];
        const double x = s.px[i], y = s.py[i];
        put_xy(s, L, i, true, (float)x, (float)y);
        s.ter[i] = terrain_at(c, x, y, ct, sn);
        s.flags[i] &= (uint8_t)~(F_HIT | F_WAS_HIT);
    }
    __syncthreads();
    DCA_PROF(P, e, 2);
    // phase B: scripted aim (cheat_script_ai.py:196-218), and the list of agents that may hit anything this step.
    // np.argmin over the float32 distance row (first minimum; dead guards +inf): one packed pass keeps the two
    // smallest keys (approximate squared distance, low idx_bits = guard index).  sqrt_rn is monotone, so the smallest
    // key names the argmin unless the runner-up is within one key step -- then (rare) an exact pass compares true
    // sqrt_rn distances, lowest index first.
    {
        const uint32_t im = (1u << P.idx_bits) - 1u;
        const int Gq = (L.G + 3) >> 2;
#if DCA_AIM_SPLIT
        // A lane pair per alive attacker: each lane scans half of the guard quads, the two (smallest, runner-up) key
        // pairs are merged with one shuffle each; then the even lane resolves the angle and the odd lane the shooter
        // test.  All 32 lanes of a warp run the same number of rounds (the shuffles need them).
        const int n_att = n_alive - n_alive_g, Gh = (Gq + 1) >> 1;
        for (int p0 = (tid & ~31) >> 1; p0 < n_att; p0 += nt >> 1) {
            const int p = p0 + ((tid & 31) >> 1), half = tid & 1;
            const bool active = p < n_att;
            const int i = active ? s.alist[n_alive_g + p] : L.G, t = i - L.G;
            const float pox = s.xa[t], poy = s.ya[t];
            const u64 px2 = pack2(pox, pox), py2 = pack2(poy, poy);
            const ulonglong2 *x4 = (const ulonglong2 *)s.xg, *y4 = (const ulonglong2 *)s.yg;
            uint32_t m1 = 0xFFFFFFFFu, m2 = 0xFFFFFFFFu;
            const int q_lo = half ? Gh : 0, q_hi = active ? (half ? Gq : Gh) : 0;
#pragma unroll 4
            for (int q = q_lo; q < q_hi; ++q) {
                const ulonglong2 xs = x4[q], ys = y4[q];
                const u64 s01 = sqdist2(xs.x, ys.x, px2, py2), s23 = sqdist2(xs.y, ys.y, px2, py2);
                const uint32_t ka = (__float_as_uint(lo2(s01)) & ~im) | (uint32_t)(4 * q);
                const uint32_t kb = (__float_as_uint(hi2(s01)) & ~im) | (uint32_t)(4 * q + 1);
                const uint32_t kc = (__float_as_uint(lo2(s23)) & ~im) | (uint32_t)(4 * q + 2);
                const uint32_t kd = (__float_as_uint(hi2(s23)) & ~im) | (uint32_t)(4 * q + 3);
                const uint32_t lab = min(ka, kb), hab = max(ka, kb), lcd = min(kc, kd), hcd = max(kc, kd);
                const uint32_t lo4 = min(lab, lcd), se4 = min(max(lab, lcd), min(hab, hcd));
                m2 = min(min(m2, se4), max(m1, lo4));
                m1 = min(m1, lo4);
            }
            {
                const uint32_t o1 = __shfl_xor_sync(0xffffffffu, m1, 1), o2 = __shfl_xor_sync(0xffffffffu, m2, 1);
                m2 = min(max(m1, o1), min(m2, o2));
                m1 = min(m1, o1);
            }
            if (!active) continue;
            if (half == 0) {
                int best = (int)(m1 & im);
                if ((m2 >> P.idx_bits) - (m1 >> P.idx_bits) <= 1u && m1 < 0x7F800000u) {
                    float bd = CUDART_INF_F;
                    best = 0;
                    for (int g = 0; g < L.G; ++g) {
                        const float d = f32_dist(pox, poy, s.xg[g], s.yg[g]);
                        if (d < bd) { bd = d; best = g; }
                    }
                }
                if (m1 >= 0x7F800000u) best = 0;  // no guard alive: argmin of an all-inf row is 0
                double th = atan2(s.py[best] - s.py[i], s.px[best] - s.px[i]);
                if (th < 0) th += 2 * kPi;
                s.atk[i] = th;
                if (P.out.nearest) P.out.nearest[(size_t)e * L.A + t] = best;
            } else {
                // can_fire (cheat_script_ai.py:214-217): candidate shooter if the nearest guard may be inside the radius
                if (!c.disable_red_function && __uint_as_float(m1 & ~im) <= laser_r2f(s, c, i)) push_shooter(s, i);
            }
        }
#else
        for (int q = n_alive_g + tid; q < n_alive; q += nt) {
            const int i = s.alist[q], t = i - L.G;
            int best = -1;
            {
                const float pox = s.xa[t], poy = s.ya[t];
                const u64 px2 = pack2(pox, pox), py2 = pack2(poy, poy);
                const ulonglong2 *x4 = (const ulonglong2 *)s.xg, *y4 = (const ulonglong2 *)s.yg;
                uint32_t m1 = 0xFFFFFFFFu, m2 = 0xFFFFFFFFu;
#pragma unroll 4
                for (int q = 0; q < Gq; ++q) {
                    const ulonglong2 xs = x4[q], ys = y4[q];
                    const u64 s01 = sqdist2(xs.x, ys.x, px2, py2), s23 = sqdist2(xs.y, ys.y, px2, py2);
                    const uint32_t ka = (__float_as_uint(lo2(s01)) & ~im) | (uint32_t)(4 * q);
                    const uint32_t kb = (__float_as_uint(hi2(s01)) & ~im) | (uint32_t)(4 * q + 1);
                    const uint32_t kc = (__float_as_uint(lo2(s23)) & ~im) | (uint32_t)(4 * q + 2);
                    const uint32_t kd = (__float_as_uint(hi2(s23)) & ~im) | (uint32_t)(4 * q + 3);
                    // two smallest of {m1, m2, ka..kd}: pairwise tournaments keep the chains short
                    const uint32_t lab = min(ka, kb), hab = max(ka, kb), lcd = min(kc, kd), hcd = max(kc, kd);
                    const uint32_t lo4 = min(lab, lcd), se4 = min(max(lab, lcd), min(hab, hcd));
                    m2 = min(min(m2, se4), max(m1, lo4));
                    m1 = min(m1, lo4);
                }
                best = (int)(m1 & im);
                if ((m2 >> P.idx_bits) - (m1 >> P.idx_bits) <= 1u && m1 < 0x7F800000u) {
                    float bd = CUDART_INF_F;
                    best = 0;
                    for (int g = 0; g < L.G; ++g) {
                        const float d = f32_dist(pox, poy, s.xg[g], s.yg[g]);
                        if (d < bd) { bd = d; best = g; }
                    }
                }
                if (m1 >= 0x7F800000u) best = 0;  // no guard alive: argmin of an all-inf row is 0
                double th = atan2(s.py[best] - s.py[i], s.px[best] - s.px[i]);
                if (th < 0) th += 2 * kPi;
                s.atk[i] = th;
                // can_fire (cheat_script_ai.py:214-217): candidate shooter if the nearest guard may be inside the radius
                if (!c.disable_red_function && __uint_as_float(m1 & ~im) <= laser_r2f(s, c, i)) push_shooter(s, i);
            }
            if (P.out.nearest) P.out.nearest[(size_t)e * L.A + t] = best;
        }
#endif
        // guards: hmin = lower bound of the squared distance to the nearest alive attacker, left by the last
        // observation pass (same positions, same alive set)
        for (int q = tid; q < n_alive_g; q += nt) {
            const int g = s.alist[q];
            if (s.hmin[g] <= laser_r2f(s, c, g)) push_shooter(s, g);
        }
    }
    __syncthreads();
    DCA_PROF(P, e, 3);
    // phase C: apply_laser_effect core.py:244-263.  C1: every listed shooter resolves its unit heading and scans the
    // opposing team with the packed float32 pre-filter, queueing (shooter, target) candidates; C2: the whole CTA works
    // through the queue, one pair per thread: the reference's range test (d2 > r2 -> miss, same arithmetic), then its
    // window test in cone form -- |reg(dir +- w/2 - theta)| <= w  <=>  dot(heading, d) >= cos(w/2) |d| -- unless that
    // is within 1e-9 (relative) of the boundary, ~1e6 times the rounding error of either formulation; those pairs
    // (and every pair when the window is too wide for the cone form) get the atan2 test.  A serial loop per shooter
    // (fp64 chains, divergent outcomes) was the longest thing in the kernel for envs where the teams are on top of each
    // other: 30 k cycles against a CTA lifetime of 29 k.  Headings and queue live in the entity-table region, which is
    // not written before phase D.
    const int nsh = s.misc[2];
    if (nsh > 0) {  // uniform over the CTA
        double *hx = (double *)s.table, *hy = hx + L.N;
        uint32_t *pairq = (uint32_t *)(hy + L.N);
        const int pair_cap = 12 * L.N;  // table region = (N + 1) * 64 B = 16 N (headings) + 48 N + 64
        const double cc2 = P.cone_cos2;
        // Every lane of a warp runs the same number of rounds (a guard shooter scans L.A opponents, an attacker L.G:
        // with ceil(G/32) != ceil(A/32) their chunk counts differ), lanes without a shooter or past their own list
        // contribute an empty mask: the warp-wide slot allocation below needs all 32 lanes.
        const int cmax = L.G > L.A ? L.G : L.A;
        for (int q0 = tid & ~31; q0 < nsh; q0 += nt) {
            const int lane = tid & 31, q = q0 + lane;
            const bool has = q < nsh;
            const int i = has ? (int)s.shooters[q] : 0;
            const bool att = i >= L.G;
            const float *ox = att ? s.xg : s.xa, *oy = att ? s.yg : s.ya;
            const int opp_cnt = has ? (att ? L.G : L.A) : 0, opp_base = att ? 0 : L.G;
            const float pox = att ? s.xa[i - L.G] : s.xg[i], poy = att ? s.ya[i - L.G] : s.yg[i];
            const u64 px2 = pack2(pox, pox), py2 = pack2(poy, poy);
            const float r2f = laser_r2f(s, c, i);
            const int nquad = (opp_cnt + 3) >> 2;
            if (has && cc2 > 0) {
                double sn_, cs_;
                sincos(s.atk[i], &sn_, &cs_);
                hx[i] = cs_; hy[i] = sn_;
            }
            for (int c0 = 0; c0 < cmax; c0 += 32) {  // warp-uniform
                uint32_t m = 0;
                if (c0 < opp_cnt) {
                    m = scan_chunk(ox, oy, c0, min(8, nquad - (c0 >> 2)), px2, py2, r2f);
                    if (r2f == CUDART_INF_F) {  // no bound: every alive opponent is a candidate (inf <= inf passes above)
                        m = 0;
                        for (int j = 0; j < 32 && c0 + j < opp_cnt; ++j)
                            if (s.flags[opp_base + c0 + j] & F_ALIVE) m |= 1u << j;
                    }
                }
                // Queue slots for the whole warp with ONE atomic: one atomic per candidate on the single counter word
                // serialised the CTA -- 500 pairs took 27 k cycles.
                const unsigned act = 0xffffffffu;
                const int cntm = __popc(m);
                int incl = cntm;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const int v = __shfl_up_sync(act, incl, o);
                    if (lane >= o) incl += v;
                }
                const int last = 31;
                int slot = 0;
                if (lane == last && incl > 0) slot = atomicAdd(&s.misc[3], incl);
                slot = __shfl_sync(act, slot, last) + incl - cntm;
                while (m) {
                    const int j = opp_base + c0 + __ffs(m) - 1;
                    m &= m - 1;
                    if (slot < pair_cap) pairq[slot] = ((uint32_t)i << 16) | (uint32_t)j;
                    else laser_pair(s, c, i, j);  // queue full: the reference's arithmetic, in place
                    ++slot;
                }
            }
        }
        __syncthreads();
        const int np = min(s.misc[3], pair_cap);
        for (int q = tid; q < np; q += nt) {
            const uint32_t pr = pairq[q];
            const int i = (int)(pr >> 16), j = (int)(pr & 0xffffu);
            if (cc2 > 0) {
                const double radius = c.agent_shoot_base_radius * s.ter[i], r2 = radius * radius;
                const double dx = s.px[j] - s.px[i], dy = s.py[j] - s.py[i];
                const double len2 = dx * dx + dy * dy;
                if (len2 > r2) continue;  // out of range (cython_func.pyx:30-31)
                const double dot = dx * hx[i] + dy * hy[i];
                const double margin = dot * dot - cc2 * len2;
                if (len2 > 1e-100 && fabs(margin) > 1e-9 * len2) {
                    if (dot > 0.0 && margin > 0.0) {
                        // piled-up agents hit the same target from neighbouring queue slots: lanes with the same
                        // target add once instead of fighting over one shared-memory word
                        const unsigned hit = __activemask();
                        const unsigned same_j = __match_any_sync(hit, j), same_i = __match_any_sync(hit, i);
                        if ((int)(__ffs(same_j) - 1) == (tid & 31)) atomicAdd(&s.hitcnt[j], __popc(same_j));
                        if ((int)(__ffs(same_i) - 1) == (tid & 31)) atomicAdd(&s.shots[i], __popc(same_i));
                    }
                    continue;
                }
            }
            laser_pair(s, c, i, j);
        }
        __syncthreads();
        // the table region is the observation pass's again: restore its all-zero record
        init_pads(s, L, P, false, true);
    }
    DCA_PROF(P, e, 4);

    // phase D: deaths core.py:266-279, _set_action parallel_run.py:326-374 / planner force cheat_script_ai.py:183-184,
    // wall force core.py:210-222,339-353, integrate_state :282-302 (survivors only); then the agent's float32 entity
    // record and list position for the observation pass.
    // Everything above read only the env's own blob (and `skip`); actions and attacker forces may come from the kernel
    // right before this one in the stream (DCA_OPT_OVERLAP_PLANNER), so order behind it here -- behind the whole
    // planner grid, or (DCA_OPT_PIPELINE_STEPS) behind the planner warp that filled this env's forces.
    if (pipe.wait_plan) pipe_wait(pipe.epoch + P.E + e, pipe.wait_plan);
    else pdl_wait();
    for (int q = tid; q < n_alive; q += nt) {
        const int i = s.alist[q];
        {
            uint8_t f = s.flags[i];
            const int sh = s.shots[i];
            if (sh > 0) { f |= F_HIT; s.nhit[i] += sh; s.flags[i] = f; }
            const int hc = s.hitcnt[i];
            if (hc > 0) {
                s.nwas[i] += hc;
                s.flags[i] = (uint8_t)((f | F_WAS_HIT | F_JUST_DIED) & ~F_ALIVE);
                atomicAdd(&s.misc[i < L.G ? 0 : 1], 1);
                put_xy(s, L, i, false, 0.f, 0.f);  // gone from the distance lists of the observation pass
            } else {
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
                } else {
                    const double2 F = ((const double2 *)att_force)[(size_t)e * L.A + (i - L.G)];
                    fx = F.x; fy = F.y;
                }
                const double x = s.px[i], y = s.py[i];
                const double wall = c.size, esz = 0.05, dt = 0.1;
                const double d0 = x - (esz + (-wall)), d1 = wall - x - esz, d2 = y - esz - (-wall), d3 = wall - y - esz;
                double wx = 0, wy = 0;
                if (!(d0 > 0 && d1 > 0 && d2 > 0 && d3 > 0)) {
                    const double f1 = wall_push(d0), f2 = wall_push(d1), f3 = wall_push(d2), f4 = wall_push(d3);
                    wx = f1 - f2; wy = f3 - f4;
                }
                const double Fx = wx + fx, Fy = wy + fy;
                double vx = s.vx[i] * 0.75, vy = s.vy[i] * 0.75;
                vx += (Fx / 1.0) * dt; vy += (Fy / 1.0) * dt;
                const double speed = sqrt(vx * vx + vy * vy);
                if (speed > c.agent_max_speed) {
                    vx = vx / speed * c.agent_max_speed;
                    vy = vy / speed * c.agent_max_speed;
                }
                s.atk[i] = reg_angle(s.atk[i] + ar);
                s.vx[i] = vx; s.vy[i] = vy;
                s.px[i] = x + vx * dt; s.py[i] = y + vy * dt;
                build_row(s, L, i);  // only alive entities are ever gathered into an observation
                if (i < L.G) s.glist[atomicAdd(&s.misc[6], 1)] = (int16_t)i;
            }
        }
    }
    __syncthreads();
    DCA_PROF(P, e, 5);
    // phase E: done/win (_get_done parallel_run.py:286-316), reward (collective_assault_env.py:153-179, :228-231)
    const int nG = nG0 - s.misc[0], nA = nA0 - s.misc[1];
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
        if (!(f & F_ALIVE)) {  // :313-316 dead guard: an observation of zeros (no RNG draws)
            uint32_t *row = (uint32_t *)(s.sid + g * kObsRows);
#pragma unroll
            for (int k = 0; k < kObsRows / 2; ++k) row[k] = 0xFFFFFFFFu;
        }
    }
    if (tid == 0) {
        s.sc->n_alive_guards = nG; s.sc->n_alive_attackers = nA;
        s.sc->time_step = ts + 1;  // :238
        P.out.done[e] = (uint8_t)done;
        P.out.win[e] = (uint8_t)win;
    }
    DCA_PROF(P, e, 6);
    // phase F: observation; on done the terminal obs goes to tobs, the env is reset and observed again
    const bool do_reset = done && autoreset;
    float *dest = do_reset ? P.out.tobs : P.out.obs;
    const uint32_t *words = rng_words;
    for (int pass = 0;; ++pass) {
        const bool last = pass == 1 || !do_reset;  // the blob image is final once this pass's tasks are through
#ifndef DCA_NO_EARLY_STORE
        const bool early = last;
#else
        const bool early = false;  // diagnostics build: blob stored at the very end, as before
#endif
        obs_pass(s, P, L, e, dest, words, obs_counter + 1 + pass, pass == 0, pass == 0 ? 7 : 12, early);
        if (last) break;
        reset_world(s, P, L, e);
        dest = P.out.obs;
        words = nullptr;
    }
    write_alive(s, P, L, e);
    DCA_PROF(P, e, 10);
#ifndef DCA_NO_EARLY_STORE
    finish_blob(my_epoch, pipe.publish);  // the bulk store has been in flight since the observation tasks ended
#else
    store_blob(s, P, L, e, my_epoch, pipe.publish);
#endif
    DCA_PROF(P, e, 11);
    if (P.prof != nullptr && threadIdx.x == 0) {  // diagnostics: wall-clock end of the CTA (ns) and the SM it ran on
        unsigned long long gt;
        unsigned sm;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
        asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
        P.prof[(size_t)e * 16 + 15] = (gt << 8) | (unsigned long long)(sm & 0xffu);
    }
}

// reset kernel: SuperpoolEnv.reset -> collective_assaultGlobalEnv.reset parallel_run.py:248-270
template <int NT, int TG, int TA>
__global__ void __launch_bounds__(NT, NT == 128 ? DCA_CTAS_128 : DCA_CTAS_256) k_reset(const Params P,
                                                                            const uint8_t *__restrict__ env_mask) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int e = blockIdx.x;
    if (env_mask != nullptr && !env_mask[e]) return;
    const BlobLayout L = TG > 0 ? make_layout(TG, TA) : P.L;
    const Smem s = carve(smem_raw, L);
    load_blob(s, P, L, e);
    const uint32_t obs_counter = s.sc->obs_counter;
    reset_world(s, P, L, e);
    obs_pass(s, P, L, e, P.out.obs, nullptr, obs_counter + 1, false, 12);
    if (threadIdx.x == 0) { P.out.done[e] = 0; P.out.win[e] = 0; }
    for (int g = threadIdx.x; g < L.G; g += blockDim.x) P.out.reward[(size_t)e * L.G + g] = 0.f;
    write_alive(s, P, L, e);
    store_blob(s, P, L, e);
}

// observation-only kernel (the kNN gather kernel on the resident state)
template <int NT, int TG, int TA>
__global__ void __launch_bounds__(NT, NT == 128 ? DCA_CTAS_128 : DCA_CTAS_256) k_observe(const Params P,
                                                                              const uint32_t *__restrict__ rng_words) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int e = blockIdx.x;
    const BlobLayout L = TG > 0 ? make_layout(TG, TA) : P.L;
    const Smem s = carve(smem_raw, L);
    load_blob(s, P, L, e);
    const uint32_t counter = s.sc->obs_counter + 1;
    obs_pass(s, P, L, e, P.out.obs, rng_words, counter, false, 12);
    // only the observation counter changed: write the scalars back (hints stay valid: the state did not move)
    if (threadIdx.x == 0) ((EnvScalars *)(P.blobs + (size_t)e * L.bytes + L.scalars))->obs_counter = counter;
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
                double sn_, ct_;
                sincos(st.init_theta[e], &sn_, &ct_);
                sc->cos_theta = ct_; sc->sin_theta = sn_;
            }
            // derived speed-up hints: none for a foreign state (full-range kNN scan, every guard a laser candidate)
            if (i < L.G) {
                u64 *ids = (u64 *)(b + L.knn_ids);
                ids[2 * i] = 0x0101010101010101ull * (u64)L.A; ids[2 * i + 1] = 0x0101010101010101ull * (u64)L.G;
                ((float *)(b + L.hmin))[i] = 0.f;
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
    u64 *ids = (u64 *)(P.blobs + (size_t)e * P.L.bytes + P.L.knn_ids);
    for (int g = 0; g < P.L.G; ++g) {  // no kNN hints yet (marker = list length)
        ids[2 * g] = 0x0101010101010101ull * (u64)P.L.A; ids[2 * g + 1] = 0x0101010101010101ull * (u64)P.L.G;
    }
}

// ------------------------------------------------------------------ benchmark planner stand-in (one warp per env)
// force = centroid(alive guards) - pos for alive attackers.  NOT the reference planner (cheat_script_ai.py:28-194).
__global__ void k_plan_centroid(const Params P, double *__restrict__ att_force, const Pipe pipe) {
    pdl_trigger();  // a dca_step launched behind this kernel may start its force-independent phases now
    const BlobLayout &L = P.L;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= P.E) return;
    if (pipe.wait_step) pipe_wait(pipe.epoch + warp, pipe.wait_step);  // this env's blob is final (every lane acquires)
    else pdl_wait();  // launched as a programmatic dependent without the per-env hand-over: behind the whole grid
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
    if (pipe.epoch != nullptr) {
        __syncwarp();
        if (lane == 0) { __threadfence(); st_release_u32(pipe.epoch + P.E + warp, pipe.publish); }
    }
}

// ------------------------------------------------------------------ GAE / return
// trajectory.finalize ALGORITHM/conc/trajectory.py:67-117: reward_push_forward :60-64, threat :75-86,
// gae_finalize_return :92-117.  gamma_fwd = gamma_in_reward_forwarding_value when gamma_in_reward_forwarding is set
// (ALGORITHM/common/traj_gae.py:54-60, conc_4hist/trajectory.py:57-63), else 1.  Rollout buffers are [T,E,G]; one thread owns one (env, guard) column and scans
// time backwards (loads coalesced across the E*G columns).  dtype walk as numpy: gamma*V[t+1] in f32,
// delta / gae in f64, return stored f32.
__global__ void k_gae(const float *__restrict__ reward, const float *__restrict__ value,
                      const uint8_t *__restrict__ dead, const int32_t *__restrict__ traj_len, float *__restrict__ ret,
                      float *__restrict__ threat, float *__restrict__ reward_pushed, int T, int E, int G, double gamma,
                      double tau, double gamma_fwd, int threat_limit) {
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
    // one backward step of the scan (numpy's dtype walk: gamma*V[t+1] in f32, delta / gae in f64, f32 stores)
    auto step = [&](int i, float rw, float vf, int d) {
        const size_t k = (size_t)i * S + c;
        double r = (double)rw + carry;
        if (i > 0) {
            carry = (r * gamma_fwd) * (double)d;  // np.where(dead_mask[i], reward[i] * gamma_fwd, 0); 1.0 = plain push
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
    };
    // the recurrence is serial in time but its inputs are not: fetch kU time steps ahead of the dependent chain
    constexpr int kU = 8;
    int i = Lc - 1;
    for (; i >= kU - 1; i -= kU) {
        float rw[kU], vf[kU];
        int dd[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const size_t k = (size_t)(i - u) * S + c;
            rw[u] = reward[k]; vf[u] = value[k]; dd[u] = dead[k] != 0;
        }
#pragma unroll
        for (int u = 0; u < kU; ++u) step(i - u, rw[u], vf[u], dd[u]);
    }
    for (; i >= 0; --i) {
        const size_t k = (size_t)i * S + c;
        step(i, reward[k], value[k], dead[k] != 0);
    }
}

}  // namespace dca
