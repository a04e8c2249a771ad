// dca_planner.cuh -- the scripted attacker formation planner on the GPU, one CTA per env.
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
constexpr int kPlanThreads = 128;

// persistent team state of one env (World.start_flag / teams_result_step1 / team_centroid_step1 survive resets)
struct PlanState {
    uint8_t *flags;     // [E,2]  start_flag (0 = teams not formed yet), small-team flag
    uint32_t *counter;  // [E]    k-means calls so far (keyed RNG counter)
    uint8_t *label;     // [E,A]  team of every attacker
    double *cent;       // [E,K,2]
};

struct PlanSmem {
    double *gx, *gy, *ax, *ay;      // positions
    uint8_t *galive, *aalive;
    int16_t *gidx;                  // alive guards in index order
    unsigned long long *adj;        // [n][W] eps-graph rows
    int16_t *comp;                  // [n] component of alive guard a
    int16_t *csize;                 // [n] size of component k
    int16_t *tsrc;                  // [K] target source: >= 0 cluster id, < 0 -> noise point ~value
    double *tx, *ty;                // [K] targets
    double *cx, *cy, *nx, *ny;      // [K] centroids (code book) / new code book
    int32_t *cnt;                   // [K]
    double *d2;                     // [A]
    int16_t *lab;                   // [A] team label in use
    int16_t *team_target;           // [K]
    int16_t *toff;                  // [K+1] member list offsets
    int16_t *mem;                   // [A] members grouped by team
    // per-warp assignment workspaces (4 warps)
    double *hu, *hv, *hminv;        // [4][M+1]
    int16_t *hp, *hway;             // [4][M+1]
    uint8_t *hused;                 // [4][M+1]
    int16_t *hcol;                  // [4][M]
    double *sx, *sy;                // [4][A] ring slots
    int32_t *misc;                  // [0] n alive guards, [1] n components, [2] n targets, [3] raised flag, [4] pick
};

__host__ __device__ inline int plan_smem_bytes(int G, int A) {
    const int K = A / 5 > 0 ? A / 5 : 1, M = (A > K ? A : K) + 1, W = (G + 63) / 64;
    int o = 0;
    o += 2 * G * 8 + 2 * A * 8;          // positions
    o += align_up(G + A, 16);            // alive
    o += align_up(G * 2, 16);            // gidx
    o += G * W * 8;                      // adj
    o += 2 * align_up(G * 2, 16);        // comp csize
    o += align_up(K * 2, 16);            // tsrc
    o += 6 * align_up(K * 8, 16);        // tx ty cx cy nx ny
    o += align_up(K * 4, 16);            // cnt
    o += align_up(A * 8, 16);            // d2
    o += align_up(A * 2, 16);            // lab
    o += align_up(K * 2, 16) + align_up((K + 1) * 2, 16) + align_up(A * 2, 16);  // team_target toff mem
    o += 4 * (3 * align_up(M * 8, 16) + 2 * align_up(M * 2, 16) + align_up(M, 16) + align_up(M * 2, 16));
    o += 4 * 2 * align_up(A * 8, 16);    // sx sy
    o += 32;
    return o;
}

__device__ __forceinline__ PlanSmem plan_carve(unsigned char *p, int G, int A) {
    const int K = A / 5 > 0 ? A / 5 : 1, M = (A > K ? A : K) + 1, W = (G + 63) / 64;
    PlanSmem s;
    s.gx = (double *)p; p += G * 8; s.gy = (double *)p; p += G * 8;
    s.ax = (double *)p; p += A * 8; s.ay = (double *)p; p += A * 8;
    s.galive = p; s.aalive = p + G; p += align_up(G + A, 16);
    s.gidx = (int16_t *)p; p += align_up(G * 2, 16);
    s.adj = (unsigned long long *)p; p += G * W * 8;
    s.comp = (int16_t *)p; p += align_up(G * 2, 16);
    s.csize = (int16_t *)p; p += align_up(G * 2, 16);
    s.tsrc = (int16_t *)p; p += align_up(K * 2, 16);
    const int k8 = align_up(K * 8, 16);
    s.tx = (double *)p; p += k8; s.ty = (double *)p; p += k8;
    s.cx = (double *)p; p += k8; s.cy = (double *)p; p += k8;
    s.nx = (double *)p; p += k8; s.ny = (double *)p; p += k8;
    s.cnt = (int32_t *)p; p += align_up(K * 4, 16);
    s.d2 = (double *)p; p += align_up(A * 8, 16);
    s.lab = (int16_t *)p; p += align_up(A * 2, 16);
    s.team_target = (int16_t *)p; p += align_up(K * 2, 16);
    s.toff = (int16_t *)p; p += align_up((K + 1) * 2, 16);
    s.mem = (int16_t *)p; p += align_up(A * 2, 16);
    const int m8 = align_up(M * 8, 16), m2 = align_up(M * 2, 16), m1 = align_up(M, 16);
    s.hu = (double *)p; p += 4 * m8; s.hv = (double *)p; p += 4 * m8; s.hminv = (double *)p; p += 4 * m8;
    s.hp = (int16_t *)p; p += 4 * m2; s.hway = (int16_t *)p; p += 4 * m2;
    s.hused = p; p += 4 * m1;
    s.hcol = (int16_t *)p; p += 4 * m2;
    s.sx = (double *)p; p += 4 * align_up(A * 8, 16); s.sy = (double *)p; p += 4 * align_up(A * 8, 16);
    s.misc = (int32_t *)p;
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

__device__ __forceinline__ double plan_sq2(double ax, double ay, double bx, double by) {
    const double dx = ax - bx, dy = ay - by;
    double d = 0;
    d += dx * dx;
    d += dy * dy;
    return d;
}

__device__ __forceinline__ uint32_t plan_word(const DcaConfig &c, uint32_t gid, uint32_t counter, int attempt, int d) {
    const uint4 q = philox4x32_10(gid, counter, (uint32_t)(attempt * 64 + d / 4), SITE_PLAN, (uint32_t)c.seed,
                                  (uint32_t)(c.seed >> 32));
    const int k = d & 3;
    return k == 0 ? q.x : (k == 1 ? q.y : (k == 2 ? q.z : q.w));
}

// get_groups' kmeans2(attacker positions, K, iter=20, minit='++', missing='raise') with up to 5 redraws, then
// missing='warn' (cheat_script_ai.py:276-288), CTA-wide.  Result: s.lab[A], s.cx/s.cy[K].
__device__ void plan_kmeans(const PlanSmem &s, const DcaConfig &c, int A, int K, uint32_t gid, uint32_t counter) {
    const int tid = threadIdx.x, nt = blockDim.x;
    for (int attempt = 0;; ++attempt) {
        // _kpp seeding
        if (tid == 0) {
            const int pick = (int)__umulhi(plan_word(c, gid, counter, attempt, 0), (uint32_t)A);
            s.cx[0] = s.ax[pick]; s.cy[0] = s.ay[pick];
        }
        for (int j = tid; j < A; j += nt) s.d2[j] = CUDART_INF;
        __syncthreads();
        for (int i = 1; i < K; ++i) {
            for (int j = tid; j < A; j += nt) {
                const double d = plan_sq2(s.cx[i - 1], s.cy[i - 1], s.ax[j], s.ay[j]);
                if (d < s.d2[j]) s.d2[j] = d;
            }
            __syncthreads();
            if (tid == 0) {
                double tot = 0;
                for (int j = 0; j < A; ++j) tot += s.d2[j];
                const double r = (double)plan_word(c, gid, counter, attempt, i) * 0x1p-32;
                double cum = 0;
                int pick = A - 1;
                for (int j = 0; j < A; ++j) {
                    cum += s.d2[j] / tot;
                    if (cum >= r) { pick = j; break; }
                }
                s.cx[i] = s.ax[pick]; s.cy[i] = s.ay[pick];
            }
            __syncthreads();
        }
        // 20 Lloyd rounds
        bool raised = false;
        for (int it = 0; it < 20; ++it) {
            if (tid == 0) s.misc[3] = 0;
            for (int j = tid; j < A; j += nt) {  // vq: nearest code, first minimum
                double best = plan_sq2(s.ax[j], s.ay[j], s.cx[0], s.cy[0]);
                int bl = 0;
                for (int k = 1; k < K; ++k) {
                    const double d = plan_sq2(s.ax[j], s.ay[j], s.cx[k], s.cy[k]);
                    if (d < best) { best = d; bl = k; }
                }
                s.lab[j] = (int16_t)bl;
            }
            __syncthreads();
            for (int k = tid; k < K; k += nt) {  // update_cluster_means: sequential sums in index order
                double sx = 0, sy = 0;
                int n = 0;
                for (int j = 0; j < A; ++j)
                    if (s.lab[j] == k) { sx += s.ax[j]; sy += s.ay[j]; ++n; }
                if (n) { s.nx[k] = sx / n; s.ny[k] = sy / n; }
                else if (attempt < 5) s.misc[3] = 1;               // missing='raise'
                else { s.nx[k] = s.cx[k]; s.ny[k] = s.cy[k]; }     // missing='warn'
            }
            __syncthreads();
            raised = s.misc[3] != 0;
            if (raised) break;
            for (int k = tid; k < K; k += nt) { s.cx[k] = s.nx[k]; s.cy[k] = s.ny[k]; }
            __syncthreads();
        }
        __syncthreads();
        if (!raised) return;
    }
}

// ------------------------------------------------------------------ planner kernel
// inj_label (i32[E,A]) / inj_cent (f64[E,K,2]) != nullptr: use these teams (the parity harness injects the
// reference's own k-means result) and leave the persistent state alone.
__global__ void __launch_bounds__(kPlanThreads) k_plan(const Params P, const PlanState ps, double *__restrict__ att_force,
                                                       const int32_t *__restrict__ inj_label,
                                                       const double *__restrict__ inj_cent) {
    extern __shared__ __align__(16) unsigned char plan_raw[];
    const BlobLayout &L = P.L;
    const DcaConfig &c = P.cfg;
    const int G = L.G, A = L.A, K = A / 5 > 0 ? A / 5 : 1, M = (A > K ? A : K) + 1, W = (G + 63) / 64;
    const int e = blockIdx.x, tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, warp = tid >> 5;
    const PlanSmem s = plan_carve(plan_raw, G, A);
    const unsigned char *blob = P.blobs + (size_t)e * L.bytes;
    const double *bpx = (const double *)(blob + L.px), *bpy = (const double *)(blob + L.py);
    const uint8_t *bfl = blob + L.flags;
    double2 *out = (double2 *)att_force + (size_t)e * A;
    for (int i = tid; i < L.N; i += nt) {
        const bool alive = bfl[i] & F_ALIVE;
        if (i < G) { s.gx[i] = bpx[i]; s.gy[i] = bpy[i]; s.galive[i] = alive; }
        else { s.ax[i - G] = bpx[i]; s.ay[i - G] = bpy[i]; s.aalive[i - G] = alive; out[i - G] = make_double2(0.0, 0.0); }
    }
    __syncthreads();

    // ---- teams (cheat_script_ai.py:103-117)
    if (inj_label != nullptr) {
        for (int a = tid; a < A; a += nt) s.lab[a] = (int16_t)inj_label[(size_t)e * A + a];
        for (int k = tid; k < K; k += nt) { s.cx[k] = inj_cent[((size_t)e * K + k) * 2]; s.cy[k] = inj_cent[((size_t)e * K + k) * 2 + 1]; }
        __syncthreads();
    } else {
        const uint32_t gid = (uint32_t)(c.env_gid0 + e);
        uint32_t counter = ps.counter[e];
        const bool start = ps.flags[2 * e] == 0;
        bool small = ps.flags[2 * e + 1] != 0;
        __syncthreads();
        if (start) {
            plan_kmeans(s, c, A, K, gid, counter++);
            for (int a = tid; a < A; a += nt) ps.label[(size_t)e * A + a] = (uint8_t)s.lab[a];
            for (int k = tid; k < K; k += nt) { ps.cent[((size_t)e * K + k) * 2] = s.cx[k]; ps.cent[((size_t)e * K + k) * 2 + 1] = s.cy[k]; }
            for (int k = tid; k < K; k += nt) s.cnt[k] = 0;
            if (tid == 0) s.misc[3] = 0;
            __syncthreads();
            for (int a = tid; a < A; a += nt) atomicAdd(&s.cnt[s.lab[a]], 1);
            __syncthreads();
            for (int k = tid; k < K; k += nt)
                if (s.cnt[k] > 0 && s.cnt[k] <= 3) s.misc[3] = 1;  // a team of <= 3: regroup on every call from now on
            __syncthreads();
            small = s.misc[3] != 0;
            __syncthreads();
        } else if (!small) {
            for (int a = tid; a < A; a += nt) s.lab[a] = ps.label[(size_t)e * A + a];
            for (int k = tid; k < K; k += nt) { s.cx[k] = ps.cent[((size_t)e * K + k) * 2]; s.cy[k] = ps.cent[((size_t)e * K + k) * 2 + 1]; }
            __syncthreads();
        }
        if (small) plan_kmeans(s, c, A, K, gid, counter++);
        if (tid == 0) { ps.counter[e] = counter; ps.flags[2 * e] = 1; ps.flags[2 * e + 1] = small ? 1 : 0; }
    }

    // ---- targets: DBSCAN components of the alive guards (cheat_script_ai.py:49-101)
    if (tid == 0) {
        int n = 0;
        for (int g = 0; g < G; ++g)
            if (s.galive[g]) s.gidx[n++] = (int16_t)g;
        s.misc[0] = n;
    }
    __syncthreads();
    const int n = s.misc[0];
    if (n == 0) return;  // no guard alive: nothing to aim at (the reference would raise); forces stay zero
    for (int a = tid; a < n; a += nt) {
        const double xa = s.gx[s.gidx[a]], ya = s.gy[s.gidx[a]];
        for (int w = 0; w < W; ++w) {
            unsigned long long m = 0;
            for (int b = w * 64; b < n && b < w * 64 + 64; ++b) {
                const double dx = s.gx[s.gidx[b]] - xa, dy = s.gy[s.gidx[b]] - ya;
                if (dx * dx + dy * dy <= 0.25) m |= 1ull << (b - w * 64);
            }
            s.adj[a * W + w] = m;
        }
    }
    __syncthreads();
    if (tid == 0) {  // components in order of their lowest member (serial flood fill over the bit rows)
        unsigned long long unv[4] = {0, 0, 0, 0};
        for (int a = 0; a < n; ++a) unv[a >> 6] |= 1ull << (a & 63);
        int nc = 0;
        for (int a = 0; a < n; ++a) {
            if (!((unv[a >> 6] >> (a & 63)) & 1ull)) continue;
            unsigned long long fr[4] = {0, 0, 0, 0};
            fr[a >> 6] = 1ull << (a & 63);
            unv[a >> 6] &= ~fr[a >> 6];
            int size = 0;
            for (;;) {
                int b = -1;
                for (int w = 0; w < W; ++w)
                    if (fr[w]) { b = w * 64 + __ffsll((long long)fr[w]) - 1; break; }
                if (b < 0) break;
                fr[b >> 6] &= ~(1ull << (b & 63));
                s.comp[b] = (int16_t)nc;
                ++size;
                for (int w = 0; w < W; ++w) {
                    const unsigned long long nb = s.adj[b * W + w] & unv[w];
                    unv[w] &= ~nb;
                    fr[w] |= nb;
                }
            }
            s.csize[nc++] = (int16_t)size;
        }
        // target list in the reference's order: labels by first appearance; the noise label expands to its points
        int ntg = 0;
        bool noise_done = false;
        unsigned long long seen[4] = {0, 0, 0, 0};
        for (int a = 0; a < n && ntg < K; ++a) {
            const int k = s.comp[a];
            if (s.csize[k] >= 2) {
                if ((seen[k >> 6] >> (k & 63)) & 1ull) continue;
                seen[k >> 6] |= 1ull << (k & 63);
                s.tsrc[ntg++] = (int16_t)k;
            } else if (!noise_done) {
                noise_done = true;
                for (int b = a; b < n && ntg < K; ++b)
                    if (s.csize[s.comp[b]] < 2) s.tsrc[ntg++] = (int16_t)~b;
            }
        }
        s.misc[2] = ntg;
    }
    __syncthreads();
    const int ntg = s.misc[2];
    for (int t = tid; t < K; t += nt) {  // :129-141 first len(teams) targets, or the last one repeated
        const int src = s.tsrc[t < ntg ? t : ntg - 1];
        if (src >= 0) {  // cluster mean = kmeans2(k=1): sequential sum in index order / count
            double sx = 0, sy = 0;
            for (int b = 0; b < n; ++b)
                if (s.comp[b] == src) { sx += s.gx[s.gidx[b]]; sy += s.gy[s.gidx[b]]; }
            s.tx[t] = sx / s.csize[src]; s.ty[t] = sy / s.csize[src];
        } else {
            s.tx[t] = s.gx[s.gidx[~src]]; s.ty[t] = s.gy[s.gidx[~src]];
        }
    }
    // member lists per team (attacker index order)
    if (tid == 0) {
        int o = 0;
        for (int k = 0; k < K; ++k) {
            s.toff[k] = (int16_t)o;
            for (int a = 0; a < A; ++a)
                if (s.lab[a] == k) s.mem[o++] = (int16_t)a;
        }
        s.toff[K] = (int16_t)o;
    }
    __syncthreads();
    const int m8 = align_up(M * 8, 16) / 8, m2 = align_up(M * 2, 16) / 2, m1 = align_up(M, 16), a8 = align_up(A * 8, 16) / 8;
    double *u = s.hu + warp * m8, *v = s.hv + warp * m8, *minv = s.hminv + warp * m8;
    int16_t *hp = s.hp + warp * m2, *hway = s.hway + warp * m2, *hcol = s.hcol + warp * m2;
    uint8_t *used = s.hused + warp * m1;
    // ---- team -> target (:153-157): assignment on the euclidean distances centroid <-> target
    if (warp == 0) {
        auto cost1 = [&](int r, int cc) {
            const double dx = s.tx[cc] - s.cx[r], dy = s.ty[cc] - s.cy[r];
            return sqrt(dx * dx + dy * dy);
        };
        assign_warp(cost1, K, K, u, v, minv, hp, hway, used, s.team_target);
    }
    __syncthreads();
    // ---- member -> ring slot (:160-184): one warp per team
    double *sx = s.sx + warp * a8, *sy = s.sy + warp * a8;
    for (int key = warp; key < K; key += (nt >> 5)) {
        const int off = s.toff[key], nm = s.toff[key + 1] - off;
        if (nm == 0) continue;
        const double delta = kPi / nm - 1;
        const double lx = s.tx[s.team_target[key]], ly = s.ty[s.team_target[key]];
        for (int j = lane; j < nm; j += 32) { sx[j] = lx + 0.8 * cos(delta * j); sy[j] = ly + 0.8 * sin(delta * j); }
        __syncwarp();
        auto cost2 = [&](int i, int j) {
            const double dx = s.ax[s.mem[off + i]] - sx[j], dy = s.ay[s.mem[off + i]] - sy[j];
            return sqrt(dx * dx + dy * dy);
        };
        assign_warp(cost2, nm, nm, u, v, minv, hp, hway, used, hcol);
        for (int i = lane; i < nm; i += 32) {
            const int a = s.mem[off + i];
            if (!s.aalive[a]) continue;
            double fx = -s.ax[a] + sx[hcol[i]], fy = -s.ay[a] + sy[hcol[i]];
            if (c.disable_red_function) { fx = -fx; fy = -fy; }
            out[a] = make_double2(fx, fy);
        }
        __syncwarp();
    }
}

}  // namespace dca
