// dca_planner.cuh -- the scripted attacker formation planner on the GPU, one warp (= one CTA) per env.
// CheatScriptAI.attackers_policy_1 MISSION/dca/cheat_script_ai.py:19-194 and get_groups :259-297 (paths relative to
// /root/reference): DBSCAN clusters of the alive guards -> targets; k-means teams of the attackers; team -> target and
// member -> ring-slot assignments; force = slot - position.  The reference calls sklearn / scipy for the three
// building blocks (unpinned versions, absent from the tree); their published algorithms are restated here, in fp64
// with every product and sum rounded separately (-fmad=false), the same arithmetic the test suite's CPU checker uses:
//   DBSCAN(eps .5, min_samples 2)  = connected components (>= 2 points) of the eps-graph, numbered by lowest index
//   kmeans2(k, iter 20, '++')      = k-means++ seeding + 20 Lloyd rounds, sequential-sum means, 'raise' -> redraw
//   linear_sum_assignment          = shortest augmenting paths with potentials, ties -> lowest column
// The k-means++ draws come from the keyed Philox stream (site 3), so team composition is statistically, not
// bit-wise, the reference's; with injected labels/centroids the kernel reproduces the reference's forces.
#pragma once
#include "dca_device.cuh"

namespace dca {

enum : uint32_t { SITE_PLAN = 3 };
constexpr int kPlanThreads = 32;  // one warp per env: no cross-warp waiting, every phase is warp-synchronous
constexpr int kPlanTeamPre = 8;   // teams up to this size are matched by subset DP (assign_dp8)
constexpr int kPlanTeamPre2 = 16; // teams up to this size get their cost matrix precomputed (independent sqrt chains)

// persistent team state of one env (World.start_flag / teams_result_step1 / team_centroid_step1 survive resets)
struct PlanState {
    uint8_t *flags;     // [E,2]  start_flag (0 = teams not formed yet), small-team flag
    uint32_t *counter;  // [E]    k-means calls so far (keyed RNG counter)
    uint8_t *label;     // [E,A]  team of every attacker
    double *cent;       // [E,K,2]
    double *big;        // [E, big_n^2] cost matrices of teams with more than kPlanTeamPre2 members (L2-resident scratch)
    int big_n;          // min(A, kPlanBigMax), 0 when A <= kPlanTeamPre2 (no such team can exist)
};
constexpr int kPlanBigMax = 64;

// Shared-memory plan of one env.  The env is worked on by ONE warp phase after phase, so workspaces of different
// phases share one region (`un`): k-means {d2, pre}, DBSCAN {gidx, adj, comp, csize, tsrc}, team->target {c1 (+ the
// h* workspaces when there are more than 32 targets)}, member->slot {cq, then cw | h*: small teams first, then the
// large ones}.
// Guard positions (dead once the targets exist) share their space with the ring slots.  50v50: 5.5 KB per env, so all
// 4096 envs of the benchmark are resident at once (28 per SM; 9.8 KB gave 21 per SM = 1.3 waves).
struct PlanSmem {
    double *gx, *gy, *ax, *ay;      // positions
    uint8_t *galive, *aalive;
    int16_t *gidx;                  // alive guards in index order
    unsigned long long *adj;        // [n][W] eps-graph rows
    int16_t *comp;                  // [n] lowest member of the component of alive guard a
    int32_t *csize;                 // [n] size of the component rooted at a
    int16_t *tsrc;                  // [K] target source: >= 0 cluster root, < 0 -> noise point ~value
    double *tx, *ty;                // [K] targets
    double *cx, *cy;                // [K] centroids (code book)
    int32_t *cnt;                   // [K]
    double *d2, *pre;               // [A] squared distance to the nearest centre so far / lane-local running sums
    int16_t *lab;                   // [A] team label in use
    uint32_t *lmask;                // [K][AW] members of team k as bit rows over the attackers
    int16_t *team_target;           // [K]
    int16_t *toff;                  // [K+1] member list offsets
    int16_t *mem, *tof;             // [A] members grouped by team / team of that list position
    double *c1;                     // [K*K] team-centroid <-> target distances
    double *cq, *dp;                // [64] member <-> slot distances of one team of up to 8 / [256] subset DP values
    uint8_t *dpc;                   // [256] subset DP choices
    double *cw;                     // [kPlanTeamPre2^2] member <-> slot distances of one team of 9..16
    double *hu, *hv, *hminv;        // [A+K+2] assignment workspaces of the shared-memory routine (> 32 columns)
    int16_t *hp, *hway;             // [A+K+2]
    uint8_t *hused;                 // [A+K+2]
    int16_t *hcol;                  // [A] slot of list position i
    double *sx, *sy;                // [A] ring slots by list position
    int32_t *misc;                  // [0] n alive guards, [2] n targets, [3] raised flag, [4] changed flag
};

__host__ __device__ constexpr int plan_max(int a, int b) { return a > b ? a : b; }
__host__ __device__ constexpr int plan_h_bytes(int H) {
    return 3 * align_up(H * 8, 16) + 2 * align_up(H * 2, 16) + align_up(H, 16);  // hu hv hminv hp hway hused
}
__host__ __device__ constexpr int plan_union_bytes(int G, int A) {
    const int K = A / 5 > 0 ? A / 5 : 1, H = A + K + 2, W = (G + 63) / 64;
    const int km = 2 * align_up(A * 8, 16);
    const int db = align_up(G * 2, 16) + G * W * 8 + align_up(G * 2, 16) + align_up(G * 4, 16) + align_up(K * 2, 16);
    const int tt = align_up(K * K * 8, 16) + (K > 32 ? plan_h_bytes(H) : 0);
    const int ms = plan_max(plan_max(64 * 8 + 256 * 8 + 256, kPlanTeamPre2 * kPlanTeamPre2 * 8), plan_h_bytes(H));
    return plan_max(plan_max(km, db), plan_max(tt, ms));
}
__host__ __device__ constexpr int plan_smem_bytes(int G, int A) {
    const int K = A / 5 > 0 ? A / 5 : 1, AW = (A + 31) / 32;
    int o = 0;
    o += 2 * plan_max(G, A) * 8 + 2 * A * 8;  // gx gy (later sx sy), ax ay
    o += align_up(G + A, 16);                 // alive
    o += 4 * align_up(K * 8, 16);             // tx ty cx cy
    o += align_up(K * 4, 16);                 // cnt
    o += align_up(A * 2, 16);                 // lab
    o += align_up(K * AW * 4, 16);            // lmask
    o += align_up(K * 2, 16) + align_up((K + 1) * 2, 16) + 2 * align_up(A * 2, 16);  // team_target toff mem tof
    o += align_up(A * 2, 16);                 // hcol
    o += 32;                                  // misc
    o += plan_union_bytes(G, A);
    return o;
}

__device__ __forceinline__ PlanSmem plan_carve(unsigned char *p, int G, int A) {
    const int K = A / 5 > 0 ? A / 5 : 1, H = A + K + 2, W = (G + 63) / 64, AW = (A + 31) / 32;
    const int GA = plan_max(G, A);
    PlanSmem s;
    s.gx = (double *)p; s.sx = s.gx; p += GA * 8; s.gy = (double *)p; s.sy = s.gy; p += GA * 8;
    s.ax = (double *)p; p += A * 8; s.ay = (double *)p; p += A * 8;
    s.galive = p; s.aalive = p + G; p += align_up(G + A, 16);
    const int k8 = align_up(K * 8, 16), a8 = align_up(A * 8, 16), a2 = align_up(A * 2, 16);
    s.tx = (double *)p; p += k8; s.ty = (double *)p; p += k8;
    s.cx = (double *)p; p += k8; s.cy = (double *)p; p += k8;
    s.cnt = (int32_t *)p; p += align_up(K * 4, 16);
    s.lab = (int16_t *)p; p += a2;
    s.lmask = (uint32_t *)p; p += align_up(K * AW * 4, 16);
    s.team_target = (int16_t *)p; p += align_up(K * 2, 16);
    s.toff = (int16_t *)p; p += align_up((K + 1) * 2, 16);
    s.mem = (int16_t *)p; p += a2; s.tof = (int16_t *)p; p += a2;
    s.hcol = (int16_t *)p; p += a2;
    s.misc = (int32_t *)p; p += 32;
    unsigned char *un = p;  // ---- the phase union
    {   // k-means
        unsigned char *q = un;
        s.d2 = (double *)q; q += a8; s.pre = (double *)q;
    }
    {   // DBSCAN / targets
        unsigned char *q = un;
        s.gidx = (int16_t *)q; q += align_up(G * 2, 16);
        s.adj = (unsigned long long *)q; q += G * W * 8;
        s.comp = (int16_t *)q; q += align_up(G * 2, 16);
        s.csize = (int32_t *)q; q += align_up(G * 4, 16);
        s.tsrc = (int16_t *)q;
    }
    unsigned char *hb = un;
    s.c1 = (double *)un;                      // team -> target
    if (K > 32) hb = un + align_up(K * K * 8, 16);
    s.cq = (double *)un; s.dp = s.cq + 64; s.dpc = (uint8_t *)(s.dp + 256);  // member -> slot: the small teams first ...
    s.cw = (double *)un;                      // ... then the larger teams one by one
    {
        unsigned char *q = hb;
        const int h8 = align_up(H * 8, 16), h2 = align_up(H * 2, 16);
        s.hu = (double *)q; q += h8; s.hv = (double *)q; q += h8; s.hminv = (double *)q; q += h8;
        s.hp = (int16_t *)q; q += h2; s.hway = (int16_t *)q; q += h2;
        s.hused = q;
    }
    return s;
}

// ------------------------------------------------------------------ assignment by one warp (rows n <= cols m)
// linear_sum_assignment restated: shortest augmenting paths with potentials (serial semantics preserved); the
// scan over the columns of one row is spread over the lanes, the argmin keeps the serial tie rule (lowest column).
// cost(i, j) is evaluated on the fly.  Workspaces hold m+1 entries.  col_of_row[n] receives the result.
template <typename CostFn>
__device__ __forceinline__ void assign_warp(CostFn cost, int n, int m, double *u, double *v, double *minv, int16_t *p,
                                            int16_t *way, uint8_t *used, int16_t *col_of_row) {
    const int lane = threadIdx.x & 31;
    for (int j = lane; j <= m; j += 32) { v[j] = 0; p[j] = 0; }
    for (int i = lane; i <= n; i += 32) u[i] = 0;
    __syncwarp();
    for (int i = 1; i <= n; ++i) {
        if (lane == 0) p[0] = (int16_t)i;
        for (int j = lane; j <= m; j += 32) { minv[j] = CUDART_INF; used[j] = 0; }
        int j0 = 0;
        __syncwarp();
        for (;;) {
            if (lane == 0) used[j0] = 1;
            __syncwarp();
            const int i0 = p[j0];
            const double ui0 = u[i0];
            double best = CUDART_INF;
            int bj = 0x7fffffff;
            for (int j = lane + 1; j <= m; j += 32) {
                if (used[j]) continue;
                const double cur = (cost(i0 - 1, j - 1) - ui0) - v[j];
                double mv = minv[j];
                if (cur < mv) { mv = cur; minv[j] = cur; way[j] = (int16_t)j0; }
                if (mv < best) { best = mv; bj = j; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const double ob = __shfl_xor_sync(0xffffffffu, best, o);
                const int oj = __shfl_xor_sync(0xffffffffu, bj, o);
                if (ob < best || (ob == best && oj < bj)) { best = ob; bj = oj; }
            }
            const double delta = best;
            const int j1 = bj == 0x7fffffff ? 0 : bj;
            __syncwarp();
            for (int j = lane; j <= m; j += 32) {
                if (used[j]) { u[p[j]] += delta; v[j] -= delta; }
                else minv[j] -= delta;
            }
            j0 = j1;
            __syncwarp();
            if (p[j0] == 0) break;
        }
        if (lane == 0) {
            do {
                const int j1 = way[j0];
                p[j0] = p[j1];
                j0 = j1;
            } while (j0);
        }
        __syncwarp();
    }
    for (int j = lane + 1; j <= m; j += 32)
        if (p[j]) col_of_row[p[j] - 1] = (int16_t)(j - 1);
    __syncwarp();
}

// order-preserving map double -> u64 (and back); -0.0 is folded into +0.0 first so that equal doubles get equal keys
__device__ __forceinline__ u64 plan_dkey(double x) {
    const long long b = __double_as_longlong(x + 0.0);
    return (u64)(b ^ ((b >> 63) | (long long)0x8000000000000000ull));
}
__device__ __forceinline__ double plan_dkey_inv(u64 k) {
    const long long b = (long long)k;
    return __longlong_as_double(b < 0 ? (long long)(k ^ 0x8000000000000000ull) : ~b);
}

// arg-min of 64-bit keys over the warp with two 32-bit REDUX steps (high words, then the low words of the lanes that
// hold the smallest high word); returns the key, `holders` = lanes holding it
__device__ __forceinline__ u64 plan_warp_min_key(u64 key, unsigned *holders) {
    const unsigned full = 0xffffffffu;
    const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
    const unsigned hmin = __reduce_min_sync(full, hi);
    const unsigned lmin = __reduce_min_sync(full, hi == hmin ? lo : 0xffffffffu);
    *holders = __ballot_sync(full, hi == hmin && lo == lmin);
    return ((u64)hmin << 32) | lmin;
}

// The same assignment with the whole state in registers (m <= 32 columns): lane j-1 owns column j -- v[j], minv[j],
// way[j], used[j], the row p[j] matched to it and that row's potential u[p[j]] (it travels with the row when the
// augmenting path is flipped); the row being inserted sits in the virtual column 0, held uniformly.  Same arithmetic in
// the same order as the serial routine (identical potentials, identical ties: lowest column), no shared-memory traffic.
// cost_col(i0) = cost(row i0, this lane's column).  The set of unmatched columns is a ballot kept per row.
template <typename CostFn>
__device__ __forceinline__ void assign_warp_reg(CostFn cost_col, int n, int m, int16_t *col_of_row) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const bool col = lane < m;
    double v = 0.0, uc = 0.0;
    int p = 0;
    unsigned freem = m >= 32 ? full : ((1u << m) - 1u);  // columns no row is matched to yet
    for (int i = 1; i <= n; ++i) {
        double u0 = 0.0, minv = CUDART_INF;
        int way = 0, j0 = 0;
        bool used = !col;  // lanes without a column never compete
        int i0 = i;        // row matched to column j0 (the inserted row sits in the virtual column 0) ...
        double ui0 = 0.0;  // ... and its potential
        for (;;) {
            u64 key = ~0ull;  // order-preserving integer image of minv[j]; used / missing columns never win
            if (!used) {
                const double cur = (cost_col(i0 - 1) - ui0) - v;
                if (cur < minv) { minv = cur; way = j0; }
                key = plan_dkey(minv);
            }
            unsigned holders;
            const u64 kmin = plan_warp_min_key(key, &holders);  // ties -> lowest column
            const double delta = plan_dkey_inv(kmin);
            if (col) {
                if (used) { uc += delta; v -= delta; }
                else minv -= delta;
            }
            u0 += delta;  // column 0 is used from the first round on: u[p[0]] = u[i]
            j0 = __ffs(holders);
            if ((freem >> (j0 - 1)) & 1u) break;
            if (lane == j0 - 1) used = true;
            i0 = __shfl_sync(full, p, j0 - 1);
            ui0 = __shfl_sync(full, uc, j0 - 1);
        }
        freem &= ~(1u << (j0 - 1));
        while (j0) {  // flip the augmenting path: p[j0] = p[way[j0]], the row's potential moves with it
            const int j1 = __shfl_sync(full, way, j0 - 1);
            const int s1 = j1 > 0 ? j1 - 1 : 0;
            const int pj1 = __shfl_sync(full, p, s1);
            const double uj1 = __shfl_sync(full, uc, s1);
            if (lane == j0 - 1) { p = j1 == 0 ? i : pj1; uc = j1 == 0 ? u0 : uj1; }
            j0 = j1;
        }
    }
    if (col && p) col_of_row[p - 1] = (int16_t)lane;
    __syncwarp();
}

// The register routine for 33..64 columns: lane l owns columns l + 1 (slot 0) and l + 33 (slot 1).  Same arithmetic, same
// order, ties -> lowest column (slot 0 before slot 1).  cost(i, j): 0-based row and column.
template <typename CostFn>
__device__ __forceinline__ void assign_warp_reg2(CostFn cost, int n, int m, int16_t *col_of_row) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const bool col0 = lane < m, col1 = lane + 32 < m;
    double v0 = 0.0, v1 = 0.0, uc0 = 0.0, uc1 = 0.0;
    int p0 = 0, p1 = 0;
    unsigned free0 = m >= 32 ? full : ((1u << m) - 1u);
    unsigned free1 = m >= 64 ? full : (m > 32 ? ((1u << (m - 32)) - 1u) : 0u);
    for (int i = 1; i <= n; ++i) {
        double u0 = 0.0, minv0 = CUDART_INF, minv1 = CUDART_INF, ui0 = 0.0;
        int way0 = 0, way1 = 0, j0 = 0, i0 = i;
        bool used0 = !col0, used1 = !col1;
        for (;;) {
            u64 key0 = ~0ull, key1 = ~0ull;
            if (!used0) {
                const double cur = (cost(i0 - 1, lane) - ui0) - v0;
                if (cur < minv0) { minv0 = cur; way0 = j0; }
                key0 = plan_dkey(minv0);
            }
            if (!used1) {
                const double cur = (cost(i0 - 1, lane + 32) - ui0) - v1;
                if (cur < minv1) { minv1 = cur; way1 = j0; }
                key1 = plan_dkey(minv1);
            }
            unsigned holders;
            const u64 kmin = plan_warp_min_key(key1 < key0 ? key1 : key0, &holders);
            const unsigned h0 = __ballot_sync(full, key0 == kmin);  // used / missing columns hold ~0, never the minimum
            int jn;
            if (h0) {
                jn = __ffs(h0);
            } else {
                const unsigned h1 = __ballot_sync(full, key1 == kmin);
                jn = 32 + __ffs(h1);
            }
            const double delta = plan_dkey_inv(kmin);
            if (col0) {
                if (used0) { uc0 += delta; v0 -= delta; }
                else minv0 -= delta;
            }
            if (col1) {
                if (used1) { uc1 += delta; v1 -= delta; }
                else minv1 -= delta;
            }
            u0 += delta;
            j0 = jn;
            const int sl = (j0 - 1) >> 5, src = (j0 - 1) & 31;
            if (((sl ? free1 : free0) >> src) & 1u) break;
            if (lane == src) {
                if (sl) used1 = true;
                else used0 = true;
            }
            i0 = __shfl_sync(full, sl ? p1 : p0, src);
            ui0 = __shfl_sync(full, sl ? uc1 : uc0, src);
        }
        if ((j0 - 1) >> 5) free1 &= ~(1u << ((j0 - 1) & 31));
        else free0 &= ~(1u << ((j0 - 1) & 31));
        while (j0) {  // flip the augmenting path
            const int sl = (j0 - 1) >> 5, src = (j0 - 1) & 31;
            const int j1 = __shfl_sync(full, sl ? way1 : way0, src);
            const int sl1 = j1 > 0 ? (j1 - 1) >> 5 : 0, s1 = j1 > 0 ? (j1 - 1) & 31 : 0;
            const int pj1 = __shfl_sync(full, sl1 ? p1 : p0, s1);
            const double uj1 = __shfl_sync(full, sl1 ? uc1 : uc0, s1);
            if (lane == src) {
                const int np = j1 == 0 ? i : pj1;
                const double nu = j1 == 0 ? u0 : uj1;
                if (sl) { p1 = np; uc1 = nu; }
                else { p0 = np; uc0 = nu; }
            }
            j0 = j1;
        }
    }
    if (col0 && p0) col_of_row[p0 - 1] = (int16_t)lane;
    if (col1 && p1) col_of_row[p1 - 1] = (int16_t)(lane + 32);
    __syncwarp();
}

// assign_warp_reg with the n x m costs evaluated up front into `scratch` (row-major): the square roots then run as
// independent chains instead of sitting inside the augmenting loop's dependency chain.
template <typename CostFn>
__device__ __forceinline__ void assign_warp_reg_pre(CostFn cost_col, int n, int m, double *scratch, int16_t *col_of_row) {
    const int lane = threadIdx.x & 31;
    if (lane < m)
        for (int r = 0; r < n; ++r) scratch[r * m + lane] = cost_col(r);
    __syncwarp();
    auto pre = [=](int r) { return scratch[r * m + lane]; };
    assign_warp_reg(pre, n, m, col_of_row);
}

// ------------------------------------------------------------------ tiny assignments by subset DP (n <= 8)
// linear_sum_assignment returns THE optimal matching; for up to 8 rows the warp finds it by dynamic programming over
// column subsets instead of augmenting paths: dp[S] = min over j in S of dp[S - j] + cost(|S| - 1, j) -- row r = |S| - 1
// takes column j -- layer by layer (|S| = 1 .. n, the C(n,|S|) subsets of a layer spread over the lanes), then the
// matching is read back from the recorded choices.  n 2^(n-1) additions (80 for n = 5, 1024 for n = 8) against the
// dozens of dependent search / flip rounds of the augmenting-path routine, whose rows all prefer the same ring slots.
// The optimum is unique unless two matchings tie to the last bit of a double sum (ring slots and members are in general
// position), so the result is the serial routine's; costs are added in row order.
// kPlanPopMask: the subsets of {0..7} sorted by (size, value) -- the first C(n,k) entries of size class k are exactly
// the subsets of {0..n-1}.
__device__ const uint8_t kPlanPopMask[256] = {
    0, 1, 2, 4, 8, 16, 32, 64, 128, 3, 5, 6, 9, 10, 12, 17, 18, 20, 24, 33, 34, 36, 40, 48, 65, 66, 68, 72, 80,
    96, 129, 130, 132, 136, 144, 160, 192, 7, 11, 13, 14, 19, 21, 22, 25, 26, 28, 35, 37, 38, 41, 42, 44, 49, 50,
    52, 56, 67, 69, 70, 73, 74, 76, 81, 82, 84, 88, 97, 98, 100, 104, 112, 131, 133, 134, 137, 138, 140, 145, 146,
    148, 152, 161, 162, 164, 168, 176, 193, 194, 196, 200, 208, 224, 15, 23, 27, 29, 30, 39, 43, 45, 46, 51, 53,
    54, 57, 58, 60, 71, 75, 77, 78, 83, 85, 86, 89, 90, 92, 99, 101, 102, 105, 106, 108, 113, 114, 116, 120, 135,
    139, 141, 142, 147, 149, 150, 153, 154, 156, 163, 165, 166, 169, 170, 172, 177, 178, 180, 184, 195, 197, 198,
    201, 202, 204, 209, 210, 212, 216, 225, 226, 228, 232, 240, 31, 47, 55, 59, 61, 62, 79, 87, 91, 93, 94, 103,
    107, 109, 110, 115, 117, 118, 121, 122, 124, 143, 151, 155, 157, 158, 167, 171, 173, 174, 179, 181, 182, 185,
    186, 188, 199, 203, 205, 206, 211, 213, 214, 217, 218, 220, 227, 229, 230, 233, 234, 236, 241, 242, 244, 248,
    63, 95, 111, 119, 123, 125, 126, 159, 175, 183, 187, 189, 190, 207, 215, 219, 221, 222, 231, 235, 237, 238,
    243, 245, 246, 249, 250, 252, 127, 191, 223, 239, 247, 251, 253, 254, 255};
__constant__ uint8_t kPlanPopOff[9] = {0, 1, 9, 37, 93, 163, 219, 247, 255};
__constant__ uint8_t kPlanBinom[9][9] = {
    {1, 0, 0, 0, 0, 0, 0, 0, 0},
    {1, 1, 0, 0, 0, 0, 0, 0, 0},
    {1, 2, 1, 0, 0, 0, 0, 0, 0},
    {1, 3, 3, 1, 0, 0, 0, 0, 0},
    {1, 4, 6, 4, 1, 0, 0, 0, 0},
    {1, 5, 10, 10, 5, 1, 0, 0, 0},
    {1, 6, 15, 20, 15, 6, 1, 0, 0},
    {1, 7, 21, 35, 35, 21, 7, 1, 0},
    {1, 8, 28, 56, 70, 56, 28, 8, 1},
};

// cost: row-major with stride 8; dp[256], choice[256]: workspaces; col_of_row[n] receives the matching.
__device__ __forceinline__ void assign_dp8(const double *cost, int n, double *dp, uint8_t *choice, int16_t *col_of_row) {
    const int lane = threadIdx.x & 31;
    if (lane == 0) dp[0] = 0.0;
    __syncwarp();
    for (int k = 1; k <= n; ++k) {
        const int cnt = kPlanBinom[n][k], first = kPlanPopOff[k];
        const double *crow = cost + (k - 1) * 8;
        for (int idx = lane; idx < cnt; idx += 32) {
            const unsigned mask = kPlanPopMask[first + idx];
            double best = CUDART_INF;
            int bj = 0;
            for (unsigned m = mask; m; m &= m - 1) {  // k columns: the same trip count on every lane
                const int j = __ffs(m) - 1;
                const double val = dp[mask ^ (1u << j)] + crow[j];
                if (val < best) { best = val; bj = j; }
            }
            dp[mask] = best;
            choice[mask] = (uint8_t)bj;
        }
        __syncwarp();
    }
    if (lane == 0) {
        unsigned mask = (1u << n) - 1u;
        for (int r = n - 1; r >= 0; --r) {
            const int j = choice[mask];
            col_of_row[r] = (int16_t)j;
            mask ^= 1u << j;
        }
    }
    __syncwarp();
}

// The same DP for several small teams at once: the warp is split into groups of LANES lanes, group g works on its own
// team (n members, list offset off; n = 0: nothing to do) with its own slab of shared memory -- n x 8 costs, 2^NMAX dp
// values, 2^NMAX choices -- so four teams of up to 5 (LANES = 8) or two teams of 6 (LANES = 16) cost one pass instead of
// four / two (a layer of a 5-member team has at most 10 subsets: a whole warp per team leaves two thirds of it idle).
template <int LANES, int NMAX>
__device__ __forceinline__ void assign_dp_groups(unsigned char *slabs, int n, int off, const double *ax, const double *ay,
                                                 const double *sx, const double *sy, const int16_t *mem,
                                                 int16_t *col_of_row) {
    constexpr int kSlab = (NMAX * 8 * 8 + (1 << NMAX) * 8 + (1 << NMAX) + 15) / 16 * 16;
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, gl = lane % LANES, grp = lane / LANES;
    unsigned char *slab = slabs + grp * kSlab;
    double *cost = (double *)slab, *dp = cost + NMAX * 8;
    uint8_t *choice = (uint8_t *)(dp + (1 << NMAX));
    for (int q = gl; q < n * 8; q += LANES) {
        const int r = q >> 3, j = q & 7;
        if (j < n) {
            const int a = mem[off + r];
            const double dx = ax[a] - sx[off + j], dy = ay[a] - sy[off + j];
            cost[q] = sqrt(dx * dx + dy * dy);
        }
    }
    if (gl == 0) dp[0] = 0.0;
    int nmax = n;
#pragma unroll
    for (int o = LANES; o < 32; o <<= 1) nmax = max(nmax, __shfl_xor_sync(full, nmax, o));
    __syncwarp();
    for (int k = 1; k <= nmax; ++k) {  // warp-uniform trip count; groups with n < k idle
        if (k <= n) {
            const int cnt = kPlanBinom[n][k], first = kPlanPopOff[k];
            const double *crow = cost + (k - 1) * 8;
            for (int idx = gl; idx < cnt; idx += LANES) {
                const unsigned mask = kPlanPopMask[first + idx];
                double best = CUDART_INF;
                int bj = 0;
                for (unsigned m = mask; m; m &= m - 1) {
                    const int j = __ffs(m) - 1;
                    const double val = dp[mask ^ (1u << j)] + crow[j];
                    if (val < best) { best = val; bj = j; }
                }
                dp[mask] = best;
                choice[mask] = (uint8_t)bj;
            }
        }
        __syncwarp();
    }
    if (gl == 0 && n > 0) {
        unsigned mask = (1u << n) - 1u;
        for (int r = n - 1; r >= 0; --r) {
            const int j = choice[mask];
            col_of_row[off + r] = (int16_t)j;
            mask ^= 1u << j;
        }
    }
    __syncwarp();
}

__device__ __forceinline__ double plan_sq2(double ax, double ay, double bx, double by) {
    const double dx = ax - bx, dy = ay - by;
    double d = 0;
    d += dx * dx;
    d += dy * dy;
    return d;
}

// Keyed draws of one k-means call: word d of (env, call counter, attempt); a Philox block holds four of them and is
// kept until the next block is needed.
struct PlanDraws {
    uint4 q;
    int block;
};
__device__ __forceinline__ uint32_t plan_word(PlanDraws &dr, const DcaConfig &c, uint32_t gid, uint32_t counter, int attempt,
                                              int d) {
    const int blk = attempt * 64 + d / 4;
    if (dr.block != blk) {
        dr.block = blk;
        dr.q = philox4x32_10(gid, counter, (uint32_t)blk, SITE_PLAN, (uint32_t)c.seed, (uint32_t)(c.seed >> 32));
    }
    const int k = d & 3;
    return k == 0 ? dr.q.x : (k == 1 ? dr.q.y : (k == 2 ? dr.q.z : dr.q.w));
}

// Bit rows of the team labels: lmask[k][w] bit b <=> attacker 32 w + b is in team k (warp ballots).  No sync inside.
__device__ __forceinline__ void plan_label_masks(const PlanSmem &s, int A, int K) {
    const int AW = (A + 31) / 32, lane = threadIdx.x & 31;
    for (int j0 = 0; j0 < A; j0 += 32) {  // warp-uniform
        const int j = j0 + lane;
        const int l = j < A ? (int)s.lab[j] : -1;
        for (int k = 0; k < K; ++k) {
            const uint32_t b = __ballot_sync(0xffffffffu, l == k);
            if (lane == 0) s.lmask[k * AW + (j0 >> 5)] = b;
        }
    }
}

// get_groups' kmeans2(attacker positions, K, iter=20, minit='++', missing='raise') with up to 5 redraws, then
// missing='warn' (cheat_script_ai.py:276-288), by one warp.  Result: s.lab[A], s.lmask, s.cx/s.cy[K].
// Seeding (_kpp): centre i is the first attacker whose running sum of squared distances to the nearest centre so far
// reaches r * total.  The running sums have a fixed, warp-shaped order (the draws are keyed, not the reference's MT19937
// stream, so the order is ours to choose; the CPU checker uses the same one): lane l owns ceil(A/32) consecutive
// attackers and sums them in index order, the lane totals go through a Kogge-Stone inclusive scan.
// Lloyd: the means are sequential sums in index order (update_cluster_means); a round that changes no label is a
// fixed point, so the remaining rounds of the 20 are skipped without changing the result.
__device__ __forceinline__ void plan_kmeans(const PlanSmem &s, const DcaConfig &c, int A, int K, uint32_t gid,
                                            uint32_t counter) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, AW = (A + 31) / 32;
    const int ch = (A + 31) / 32, jb = min(A, lane * ch), je = min(A, jb + ch);
    for (int attempt = 0;; ++attempt) {
        PlanDraws dr;
        dr.block = -1;
        int pick = (int)__umulhi(plan_word(dr, c, gid, counter, attempt, 0), (uint32_t)A);
        double ccx = s.ax[pick], ccy = s.ay[pick];  // the centre drawn last
        if (lane == 0) { s.cx[0] = ccx; s.cy[0] = ccy; }
        for (int j = lane; j < A; j += 32) { s.d2[j] = CUDART_INF; s.lab[j] = -1; }
        __syncwarp();
        for (int i = 1; i < K; ++i) {
            double run = 0.0;
            for (int j = jb; j < je; ++j) {
                const double d = plan_sq2(ccx, ccy, s.ax[j], s.ay[j]);
                double m = s.d2[j];
                if (d < m) { m = d; s.d2[j] = d; }
                run += m;
                s.pre[j] = run;
            }
            double sc = run;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const double t = __shfl_up_sync(full, sc, o);
                if (lane >= o) sc += t;
            }
            const double tot = __shfl_sync(full, sc, 31);
            double excl = __shfl_up_sync(full, sc, 1);
            if (lane == 0) excl = 0.0;
            const double thr = ((double)plan_word(dr, c, gid, counter, attempt, i) * 0x1p-32) * tot;
            int first = 0x7fffffff;
            for (int j = je - 1; j >= jb; --j)
                if (excl + s.pre[j] >= thr) first = j;
            first = __reduce_min_sync(full, first);
            pick = first == 0x7fffffff ? A - 1 : first;
            ccx = s.ax[pick]; ccy = s.ay[pick];
            if (lane == 0) { s.cx[i] = ccx; s.cy[i] = ccy; }
        }
        __syncwarp();
        // Lloyd rounds
        bool raised = false;
        for (int it = 0; it < 20; ++it) {
            bool changed = false;
            for (int j = lane; j < A; j += 32) {  // vq: nearest code, first minimum
                const double xj = s.ax[j], yj = s.ay[j];
                double best = plan_sq2(xj, yj, s.cx[0], s.cy[0]);
                int bl = 0;
                for (int k = 1; k < K; ++k) {
                    const double d = plan_sq2(xj, yj, s.cx[k], s.cy[k]);
                    if (d < best) { best = d; bl = k; }
                }
                if (bl != s.lab[j]) { s.lab[j] = (int16_t)bl; changed = true; }
            }
            if (!__any_sync(full, changed)) break;  // same labels as last round: the code book already holds their means
            plan_label_masks(s, A, K);
            __syncwarp();
            bool miss = false;
            for (int k = lane; k < K; k += 32) {  // update_cluster_means: sequential sums in index order
                double sx = 0, sy = 0;
                int n = 0;
                for (int w = 0; w < AW; ++w) {
                    uint32_t m = s.lmask[k * AW + w];
                    while (m) {
                        const int j = w * 32 + __ffs(m) - 1;
                        m &= m - 1;
                        sx += s.ax[j]; sy += s.ay[j]; ++n;
                    }
                }
                if (n) { s.cx[k] = sx / n; s.cy[k] = sy / n; }
                else if (attempt < 5) miss = true;  // missing='raise'; ('warn': the old centre stays)
            }
            raised = __any_sync(full, miss);
            if (raised) break;
        }
        __syncwarp();
        if (!raised) return;
    }
}

// ------------------------------------------------------------------ planner kernel
// inj_label (i32[E,A]) / inj_cent (f64[E,K,2]) != nullptr: use these teams (the parity harness injects the
// reference's own k-means result) and leave the persistent state alone.
// GC / AC: compile-time team sizes (0 = take them from the layout at run time).
template <int GC, int AC>
__global__ void __launch_bounds__(kPlanThreads, 32) k_plan(const Params P, const PlanState ps,
                                                          double *__restrict__ att_force,
                                                          const int32_t *__restrict__ inj_label,
                                                          const double *__restrict__ inj_cent, const Pipe pipe) {
    static_assert(kPlanThreads == 32, "the shared-memory plan assumes one warp per env");
    extern __shared__ __align__(16) unsigned char plan_raw[];
    pdl_trigger();  // a dca_step launched behind this kernel may start its force-independent phases now
    const unsigned full = 0xffffffffu;
    const BlobLayout &L = P.L;
    const DcaConfig &c = P.cfg;
    const int G = GC ? GC : L.G, A = AC ? AC : L.A, K = A / 5 > 0 ? A / 5 : 1, W = (G + 63) / 64, AW = (A + 31) / 32;
    const int e = blockIdx.x, lane = threadIdx.x;
    const PlanSmem s = plan_carve(plan_raw, G, A);
    const unsigned char *blob = P.blobs + (size_t)e * L.bytes;
    const double *bpx = (const double *)(blob + L.px), *bpy = (const double *)(blob + L.py);
    const uint8_t *bfl = blob + L.flags;
    double2 *out = (double2 *)att_force + (size_t)e * A;
    DCA_PROF(P, e, 0);
    if (pipe.wait_step) pipe_wait(pipe.epoch + e, pipe.wait_step);  // DCA_OPT_PIPELINE_STEPS: this env's blob is final
    else pdl_wait();
    for (int g = lane; g < G; g += 32) { s.gx[g] = bpx[g]; s.gy[g] = bpy[g]; s.galive[g] = bfl[g] & F_ALIVE; }
    for (int a = lane; a < A; a += 32) { s.ax[a] = bpx[G + a]; s.ay[a] = bpy[G + a]; s.aalive[a] = bfl[G + a] & F_ALIVE; }
    __syncwarp();

    DCA_PROF(P, e, 1);
    // ---- teams (cheat_script_ai.py:103-117)
    if (inj_label != nullptr) {
        for (int a = lane; a < A; a += 32) s.lab[a] = (int16_t)inj_label[(size_t)e * A + a];
        for (int k = lane; k < K; k += 32) { s.cx[k] = inj_cent[((size_t)e * K + k) * 2]; s.cy[k] = inj_cent[((size_t)e * K + k) * 2 + 1]; }
        __syncwarp();
        plan_label_masks(s, A, K);
    } else {
        const uint32_t gid = (uint32_t)(c.env_gid0 + e);
        uint32_t counter = ps.counter[e];
        const bool start = ps.flags[2 * e] == 0;
        bool small = ps.flags[2 * e + 1] != 0;
        if (start) {
            plan_kmeans(s, c, A, K, gid, counter++);
            for (int a = lane; a < A; a += 32) ps.label[(size_t)e * A + a] = (uint8_t)s.lab[a];
            bool tiny = false;
            for (int k = lane; k < K; k += 32) {
                ps.cent[((size_t)e * K + k) * 2] = s.cx[k]; ps.cent[((size_t)e * K + k) * 2 + 1] = s.cy[k];
                int m = 0;
                for (int w = 0; w < AW; ++w) m += __popc(s.lmask[k * AW + w]);
                if (m > 0 && m <= 3) tiny = true;  // a team of <= 3: regroup on every call from now on (:112-117)
            }
            small = __any_sync(full, tiny);
        } else if (!small) {
            for (int a = lane; a < A; a += 32) s.lab[a] = ps.label[(size_t)e * A + a];
            for (int k = lane; k < K; k += 32) { s.cx[k] = ps.cent[((size_t)e * K + k) * 2]; s.cy[k] = ps.cent[((size_t)e * K + k) * 2 + 1]; }
            __syncwarp();
            plan_label_masks(s, A, K);
        }
        if (small) plan_kmeans(s, c, A, K, gid, counter++);
        if (lane == 0) { ps.counter[e] = counter; ps.flags[2 * e] = 1; ps.flags[2 * e + 1] = small ? 1 : 0; }
    }
    __syncwarp();

    DCA_PROF(P, e, 2);
    // ---- targets: DBSCAN(eps .5, min_samples 2) of the alive guards (cheat_script_ai.py:49-101)
    // (the k-means workspaces are dead: the shared region now holds the guard list and the graph).  Alive guards are
    // compacted into index order with ballots.
    int n = 0;
    for (int g0 = 0; g0 < G; g0 += 32) {  // warp-uniform
        const int g = g0 + lane;
        const bool alive = g < G && s.galive[g];
        const unsigned b = __ballot_sync(full, alive);
        if (alive) s.gidx[n + __popc(b & ((1u << lane) - 1u))] = (int16_t)g;
        n += __popc(b);
    }
    __syncwarp();
    if (n == 0) {  // no guard alive: nothing to aim at (the reference would raise); forces are zero
        for (int a = lane; a < A; a += 32) out[a] = make_double2(0.0, 0.0);
        if (pipe.epoch != nullptr) {
            __syncwarp();
            if (lane == 0) { __threadfence(); st_release_u32(pipe.epoch + P.E + e, pipe.publish); }
        }
        return;
    }
    for (int a = lane; a < n; a += 32) {
        const double xa = s.gx[s.gidx[a]], ya = s.gy[s.gidx[a]];
        for (int w = 0; w < W; ++w) {
            unsigned long long m = 0;
            for (int b = w * 64; b < n && b < w * 64 + 64; ++b) {
                const double dx = s.gx[s.gidx[b]] - xa, dy = s.gy[s.gidx[b]] - ya;
                if (dx * dx + dy * dy <= 0.25) m |= 1ull << (b - w * 64);
            }
            s.adj[a * W + w] = m;
        }
        s.comp[a] = (int16_t)a;
        s.csize[a] = 0;
    }
    __syncwarp();
    DCA_PROF(P, e, 3);
    // connected components by min-label propagation with pointer jumping: comp[a] -> lowest member of a's component
    for (;;) {
        bool changed = false;
        for (int a = lane; a < n; a += 32) {
            int best = s.comp[a];
            for (int w = 0; w < W; ++w) {
                unsigned long long m = s.adj[a * W + w];
                while (m) {
                    const int b = w * 64 + __ffsll((long long)m) - 1;
                    m &= m - 1;
                    best = min(best, (int)s.comp[b]);
                }
            }
            best = min(best, (int)s.comp[best]);
            if (best < s.comp[a]) { s.comp[a] = (int16_t)best; changed = true; }
        }
        if (!__any_sync(full, changed)) break;
        __syncwarp();
    }
    __syncwarp();
    DCA_PROF(P, e, 4);
    for (int a = lane; a < n; a += 32) atomicAdd(&s.csize[s.comp[a]], 1);
    __syncwarp();
    if (lane == 0) {
        // target list in the reference's order: labels by first appearance (a cluster first appears at its lowest
        // member); the noise label expands to all its points where it first appears
        int ntg = 0;
        bool noise_done = false;
        for (int a = 0; a < n && ntg < K; ++a) {
            const int r = s.comp[a];
            if (s.csize[r] >= 2) {
                if (r == a) s.tsrc[ntg++] = (int16_t)r;
            } else if (!noise_done) {
                noise_done = true;
                for (int b = a; b < n && ntg < K; ++b)
                    if (s.csize[s.comp[b]] < 2) s.tsrc[ntg++] = (int16_t)~b;
            }
        }
        s.misc[2] = ntg;
    }
    __syncwarp();
    const int ntg = s.misc[2];
    for (int t = lane; t < K; t += 32) {  // :129-141 first len(teams) targets, or the last one repeated
        const int src = s.tsrc[t < ntg ? t : ntg - 1];
        if (src >= 0) {  // cluster mean = kmeans2(k=1): sequential sum in index order / count
            double sx = 0, sy = 0;
            for (int b = 0; b < n; ++b)
                if (s.comp[b] == src) { sx += s.gx[s.gidx[b]]; sy += s.gy[s.gidx[b]]; }
            s.tx[t] = sx / s.csize[src]; s.ty[t] = sy / s.csize[src];
        } else {
            s.tx[t] = s.gx[s.gidx[~src]]; s.ty[t] = s.gy[s.gidx[~src]];
        }
        int cntk = 0;  // team sizes for the member lists
        for (int w = 0; w < AW; ++w) cntk += __popc(s.lmask[t * AW + w]);
        s.cnt[t] = cntk;
    }
    __syncwarp();  // from here on the guard positions and the DBSCAN workspaces are dead (ring slots / c1 take their place)
    DCA_PROF(P, e, 5);
    // ---- team -> target (:153-157): distances centroid <-> target, then the assignment
    if (lane == 0) {
        int o = 0;
        for (int k = 0; k < K; ++k) { s.toff[k] = (int16_t)o; o += s.cnt[k]; }
        s.toff[K] = (int16_t)o;
    }
    __syncwarp();
    // member lists (attacker index order inside a team), one team per thread
    for (int k = lane; k < K; k += 32) {
        int o = s.toff[k];
        for (int w = 0; w < AW; ++w) {
            uint32_t m = s.lmask[k * AW + w];
            while (m) {
                const int a = w * 32 + __ffs(m) - 1;
                m &= m - 1;
                s.mem[o] = (int16_t)a; s.tof[o] = (int16_t)k; ++o;
            }
        }
    }
    if (K <= 32) {
        const double txj = lane < K ? s.tx[lane] : 0.0, tyj = lane < K ? s.ty[lane] : 0.0;
        const double *cx = s.cx, *cy = s.cy;
        auto cost1 = [=](int r) {
            const double dx = txj - cx[r], dy = tyj - cy[r];
            return sqrt(dx * dx + dy * dy);
        };
        assign_warp_reg_pre(cost1, K, K, s.c1, s.team_target);
    } else {  // cost matrix for the shared-memory routine (more than 32 targets)
        for (int q = lane; q < K * K; q += 32) {
            const int r = q / K, cc = q - r * K;
            const double dx = s.tx[cc] - s.cx[r], dy = s.ty[cc] - s.cy[r];
            s.c1[q] = sqrt(dx * dx + dy * dy);
        }
        __syncwarp();
        const double *c1 = s.c1;
        auto cost1 = [c1, K](int r, int cc) { return c1[r * K + cc]; };
        assign_warp(cost1, K, K, s.hu, s.hv, s.hminv, s.hp, s.hway, s.hused, s.team_target);
    }
    __syncwarp();
    DCA_PROF(P, e, 6);
    // ---- ring slots (:160-170), by list position
    for (int i = lane; i < A; i += 32) {
        const int k = s.tof[i], nm = s.cnt[k], j = i - s.toff[k];
        const double delta = kPi / nm - 1;  // (float)(np.pi / len(team_agents) - 1) :163
        const int tg = s.team_target[k];
        s.sx[i] = s.tx[tg] + 0.8 * cos(delta * j);
        s.sy[i] = s.ty[tg] + 0.8 * sin(delta * j);
    }
    __syncwarp();
    DCA_PROF(P, e, 7);
    // ---- member -> slot (:171-173): teams of up to 8 by subset DP (assign_dp8), larger ones by augmenting paths (up to 32
    // members entirely in registers, cost matrix precomputed up to 16)
    {
        const double *ax = s.ax, *ay = s.ay;
        // small teams by subset DP: up to 5 members four teams at a time (8 lanes each), 6 members two at a time
        // (16 lanes each), 7 and 8 members one by one with the whole warp
        for (int pass = 0; pass < 2; ++pass) {
            const int groups = pass == 0 ? 4 : 2, lo = pass == 0 ? 1 : 6, hi = pass == 0 ? 5 : 6;
            const int grp = lane / (32 / groups);
            for (int kk = 0;;) {  // warp-uniform scan for the next `groups` teams of this size class
                int myk = -1, c = 0;
                for (; kk < K && c < groups; ++kk) {
                    const int nm = s.cnt[kk];
                    if (nm >= lo && nm <= hi) {
                        if (c == grp) myk = kk;
                        ++c;
                    }
                }
                if (c == 0) break;
                const int nm = myk >= 0 ? s.cnt[myk] : 0, off = myk >= 0 ? s.toff[myk] : 0;
                if (pass == 0) assign_dp_groups<8, 5>((unsigned char *)s.cq, nm, off, ax, ay, s.sx, s.sy, s.mem, s.hcol);
                else assign_dp_groups<16, 6>((unsigned char *)s.cq, nm, off, ax, ay, s.sx, s.sy, s.mem, s.hcol);
            }
        }
        for (int kk = 0; kk < K; ++kk) {  // warp-uniform
            const int nm = s.cnt[kk];
            if (nm < 7 || nm > kPlanTeamPre) continue;
            const int off = s.toff[kk];
#pragma unroll
            for (int q = lane; q < 64; q += 32) {
                const int r = q >> 3, j = q & 7;
                if (r < nm && j < nm) {
                    const int a = s.mem[off + r];
                    const double dx = ax[a] - s.sx[off + j], dy = ay[a] - s.sy[off + j];
                    s.cq[q] = sqrt(dx * dx + dy * dy);
                }
            }
            __syncwarp();
            assign_dp8(s.cq, nm, s.dp, s.dpc, s.hcol + off);
        }
        for (int kk = 0; kk < K; ++kk) {  // warp-uniform
            const int nm = s.cnt[kk];
            if (nm <= kPlanTeamPre) continue;
            const int off = s.toff[kk], ho = off + kk;
            {   // a team whose members are all dead gets no forces (:175-188): its slot assignment is never read.  Late in
                // an episode the pile of dead attackers often forms one team of 30-45 -- the most expensive assignment.
                bool any_alive = false;
                for (int i = lane; i < nm; i += 32) any_alive |= s.aalive[s.mem[off + i]] != 0;
                if (!__any_sync(full, any_alive)) continue;
            }
            const double *sx = s.sx + off, *sy = s.sy + off;
            const int16_t *mem = s.mem + off;
            // Teams too large for the shared-memory cost matrix get theirs in the env's global scratch (it stays in L1 /
            // L2): the square roots would otherwise sit inside the augmenting loop, once per visited (row, column) --
            // a late-episode team of 33-45 attackers around the last guard took 0.9-1.2 M cycles that way.
            double *big = (nm > kPlanTeamPre2 && nm <= ps.big_n) ? ps.big + (size_t)e * ps.big_n * ps.big_n : nullptr;
            if (big != nullptr) {
                for (int q = lane; q < nm * nm; q += 32) {
                    const int r = q / nm, j = q - r * nm;
                    const double dx = ax[mem[r]] - sx[j], dy = ay[mem[r]] - sy[j];
                    big[q] = sqrt(dx * dx + dy * dy);
                }
                __syncwarp();
            }
            if (nm <= 32) {
                const double sxj = lane < nm ? sx[lane] : 0.0, syj = lane < nm ? sy[lane] : 0.0;
                auto cost2 = [=](int i) {
                    const double dx = ax[mem[i]] - sxj, dy = ay[mem[i]] - syj;
                    return sqrt(dx * dx + dy * dy);
                };
                if (nm <= kPlanTeamPre2) {
                    assign_warp_reg_pre(cost2, nm, nm, s.cw, s.hcol + off);
                } else if (big != nullptr) {
                    const double *col = big + (lane < nm ? lane : 0);
                    auto cost_big = [=](int i) { return col[i * nm]; };
                    assign_warp_reg(cost_big, nm, nm, s.hcol + off);
                } else {
                    assign_warp_reg(cost2, nm, nm, s.hcol + off);
                }
            } else if (big != nullptr && nm <= 64) {
                auto cost_big = [=](int i, int j) { return big[i * nm + j]; };
                assign_warp_reg2(cost_big, nm, nm, s.hcol + off);
            } else if (big != nullptr) {
                auto cost_big = [=](int i, int j) { return big[i * nm + j]; };
                assign_warp(cost_big, nm, nm, s.hu + ho, s.hv + ho, s.hminv + ho, s.hp + ho, s.hway + ho, s.hused + ho,
                            s.hcol + off);
            } else {
                auto cost2 = [=](int i, int j) {
                    const double dx = ax[mem[i]] - sx[j], dy = ay[mem[i]] - sy[j];
                    return sqrt(dx * dx + dy * dy);
                };
                assign_warp(cost2, nm, nm, s.hu + ho, s.hv + ho, s.hminv + ho, s.hp + ho, s.hway + ho, s.hused + ho,
                            s.hcol + off);
            }
        }
    }
    __syncwarp();
    DCA_PROF(P, e, 8);
    if (P.prof != nullptr && lane == 0) {  // diagnostics: team-size profile of this env (slots 10..13)
        int mx = 0, n_big = 0, sum_big = 0, n_small = 0;
        for (int kk = 0; kk < K; ++kk) {
            const int nm = s.cnt[kk];
            mx = max(mx, nm);
            if (nm > kPlanTeamPre) { ++n_big; sum_big += nm; }
            else if (nm > 0) ++n_small;
        }
        P.prof[(size_t)e * 16 + 10] = mx; P.prof[(size_t)e * 16 + 11] = n_big;
        P.prof[(size_t)e * 16 + 12] = sum_big; P.prof[(size_t)e * 16 + 13] = n_small;
    }
    // ---- forces (:175-188); dead attackers get zero
    for (int i = lane; i < A; i += 32) {
        const int a = s.mem[i];
        double fx = 0.0, fy = 0.0;
        if (s.aalive[a]) {
            const int slot = s.toff[s.tof[i]] + s.hcol[i];
            fx = -s.ax[a] + s.sx[slot]; fy = -s.ay[a] + s.sy[slot];
            if (c.disable_red_function) { fx = -fx; fy = -fy; }
        }
        out[a] = make_double2(fx, fy);
    }
    if (pipe.epoch != nullptr) {
        __syncwarp();
        if (lane == 0) { __threadfence(); st_release_u32(pipe.epoch + P.E + e, pipe.publish); }
    }
    DCA_PROF(P, e, 9);
}

}  // namespace dca
