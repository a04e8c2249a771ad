/* ORACLE / TEST INFRASTRUCTURE -- see dca_oracle.h.  Plain C restatement of the reference's scripted attacker
 * formation planner, CheatScriptAI.attackers_policy_1 (MISSION/dca/cheat_script_ai.py:19-194) and get_groups
 * (:259-297); every function cites the lines it follows (paths relative to /root/reference).
 *
 * Third-party arithmetic on this path (absent from /root/reference, versions unpinned by the reference; this container:
 * scikit-learn 1.9.0, scipy 1.18.1): sklearn.cluster.DBSCAN, scipy.cluster.vq.kmeans2, scipy.optimize.
 * linear_sum_assignment.  Their published algorithms are restated here:
 *   - DBSCAN(eps, min_samples=2): every point with another point within eps is a core point, so clusters are the
 *     connected components (>= 2 points) of the eps-graph, numbered in order of their lowest index; the rest is noise.
 *   - kmeans2(k, iter=20, minit='++'): k-means++ seeding (scipy/cluster/vq/_vq_impl.py _kpp), then 20 rounds of
 *     nearest-centroid labelling (strict <, first minimum) and sequential-sum means; the labels returned are those of
 *     the last round, the centroids the means of those labels.  missing='raise' = start over with fresh draws.
 *     The k-means++ DRAWS come from the keyed Philox stream, not from numpy's MT19937: team composition is
 *     statistically, not bit-wise, the reference's ("parity unpinned" for this sub-step; pinned for everything else
 *     by injecting the reference's own labels/centroids, tests/golden/planner_*.npz).
 *   - linear_sum_assignment: any exact solver returns the same assignment when the optimum is unique (costs are
 *     distances of generic points); here the O(n^3) shortest-augmenting-path method with potentials.
 * Compile WITHOUT -ffast-math and with -ffp-contract=off. */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "dca_oracle.h"

#define PL_MAX 256 /* max agents per team side */
static const double PI = 3.141592653589793;
enum { SITE_PLAN = 3 };

/* ---------------------------------------------------------------- assignment (rows n <= cols m), cost(i,j) callback-free */
/* cost: row-major [n][ld].  out: col_of_row[n].  Deterministic: ties -> lowest column index. */
void orc_assign(const double *cost, int n, int m, int ld, int *col_of_row) {
    double u[PL_MAX + 1], v[PL_MAX + 1], minv[PL_MAX + 1];
    int p[PL_MAX + 1], way[PL_MAX + 1];
    char used[PL_MAX + 1];
    for (int j = 0; j <= m; ++j) { v[j] = 0; p[j] = 0; }
    for (int i = 0; i <= n; ++i) u[i] = 0;
    for (int i = 1; i <= n; ++i) {
        p[0] = i;
        int j0 = 0;
        for (int j = 0; j <= m; ++j) { minv[j] = INFINITY; used[j] = 0; }
        do {
            used[j0] = 1;
            const int i0 = p[j0];
            double delta = INFINITY;
            int j1 = 0;
            for (int j = 1; j <= m; ++j) {
                if (used[j]) continue;
                const double cur = (cost[(size_t)(i0 - 1) * ld + (j - 1)] - u[i0]) - v[j];
                if (cur < minv[j]) { minv[j] = cur; way[j] = j0; }
                if (minv[j] < delta) { delta = minv[j]; j1 = j; }
            }
            for (int j = 0; j <= m; ++j) {
                if (used[j]) { u[p[j]] += delta; v[j] -= delta; }
                else minv[j] -= delta;
            }
            j0 = j1;
        } while (p[j0] != 0);
        do {
            const int j1 = way[j0];
            p[j0] = p[j1];
            j0 = j1;
        } while (j0);
    }
    for (int j = 1; j <= m; ++j)
        if (p[j]) col_of_row[p[j] - 1] = j - 1;
}

/* ---------------------------------------------------------------- targets: DBSCAN clusters of the alive guards */
/* cheat_script_ai.py:49-101.  Returns the number of targets written (at most max_t), in the reference's list order:
 * labels in order of first appearance over the alive guards; a cluster contributes its mean (kmeans2 with k=1 ==
 * sequential sum / count), the noise label contributes every noise point. */
int orc_plan_targets(const double *gpos, const uint8_t *galive, int G, double *target, int max_t) {
    int idx[PL_MAX], comp[PL_MAX], csize[PL_MAX], n = 0;
    for (int g = 0; g < G; ++g)
        if (galive[g]) idx[n++] = g;
    const double eps2 = 0.5 * 0.5;
    for (int a = 0; a < n; ++a) comp[a] = -1;
    int nc = 0;
    for (int a = 0; a < n; ++a) { /* components in order of lowest member == sklearn's label order */
        if (comp[a] >= 0) continue;
        int stack[PL_MAX], sp = 0, size = 0;
        stack[sp++] = a; comp[a] = nc;
        while (sp) {
            const int b = stack[--sp];
            ++size;
            for (int c = 0; c < n; ++c) {
                if (comp[c] >= 0) continue;
                const double dx = gpos[2 * idx[c]] - gpos[2 * idx[b]], dy = gpos[2 * idx[c] + 1] - gpos[2 * idx[b] + 1];
                if (dx * dx + dy * dy <= eps2) { comp[c] = nc; stack[sp++] = c; }
            }
        }
        csize[nc++] = size;
    }
    int nt = 0, noise_done = 0;
    char seen[PL_MAX];
    memset(seen, 0, sizeof(seen));
    for (int a = 0; a < n && nt < max_t; ++a) {
        const int k = comp[a];
        if (csize[k] >= 2) {
            if (seen[k]) continue;
            seen[k] = 1;
            double sx = 0, sy = 0;
            for (int b = 0; b < n; ++b)
                if (comp[b] == k) { sx += gpos[2 * idx[b]]; sy += gpos[2 * idx[b] + 1]; }
            target[2 * nt] = sx / csize[k]; target[2 * nt + 1] = sy / csize[k];
            ++nt;
        } else if (!noise_done) {
            noise_done = 1;
            for (int b = a; b < n && nt < max_t; ++b)
                if (csize[comp[b]] < 2) { target[2 * nt] = gpos[2 * idx[b]]; target[2 * nt + 1] = gpos[2 * idx[b] + 1]; ++nt; }
        }
    }
    return nt;
}

/* ---------------------------------------------------------------- deterministic core, cheat_script_ai.py:119-184 */
/* label[A] in [0,K): team of every attacker (dead ones included: teams never lose members);  tcent[K][2]: k-means
 * centroids indexed by label.  force[A][2] = attacker.act[0:2] (zeros for dead attackers, :31-33,:176). */
int orc_plan_core(int G, int A, const double *gpos, const uint8_t *galive, const double *apos, const uint8_t *aalive,
                  const int32_t *label, int K, const double *tcent, int disable_red, double *force) {
    if (G > PL_MAX || A > PL_MAX || K > PL_MAX || K < 1) return -1;
    memset(force, 0, sizeof(double) * 2 * A);
    double target[2 * PL_MAX];
    int nt = orc_plan_targets(gpos, galive, G, target, K); /* :129-141 keeps the first len(teams) targets ... */
    if (nt == 0) return 0;                                 /* (no guard alive: the reference would raise) */
    for (; nt < K; ++nt) { target[2 * nt] = target[2 * nt - 2]; target[2 * nt + 1] = target[2 * nt - 1]; } /* ... or repeats the last */
    /* :153-157 team centroid -> target, Hungarian on euclidean distances */
    static _Thread_local double cost[PL_MAX * PL_MAX];
    for (int r = 0; r < K; ++r)
        for (int c = 0; c < K; ++c) {
            const double dx = target[2 * c] - tcent[2 * r], dy = target[2 * c + 1] - tcent[2 * r + 1];
            cost[r * K + c] = sqrt(dx * dx + dy * dy);
        }
    int team_target[PL_MAX];
    orc_assign(cost, K, K, K, team_target);
    /* :160-184 per team: ring of radius 0.8 around its target, members -> slots by a second assignment */
    for (int key = 0; key < K; ++key) {
        int mem[PL_MAX], n = 0;
        for (int a = 0; a < A; ++a)
            if (label[a] == key) mem[n++] = a;
        if (n == 0) continue;
        const double delta = PI / n - 1; /* (float)(np.pi / len(team_agents) - 1) :163 */
        const double lx = target[2 * team_target[key]], ly = target[2 * team_target[key] + 1];
        double sx[PL_MAX], sy[PL_MAX];
        for (int j = 0; j < n; ++j) { sx[j] = lx + 0.8 * cos(delta * j); sy[j] = ly + 0.8 * sin(delta * j); }
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < n; ++j) {
                const double dx = apos[2 * mem[i]] - sx[j], dy = apos[2 * mem[i] + 1] - sy[j];
                cost[i * n + j] = sqrt(dx * dx + dy * dy);
            }
        int slot[PL_MAX];
        orc_assign(cost, n, n, n, slot);
        for (int i = 0; i < n; ++i) {
            const int a = mem[i];
            if (!aalive[a]) continue;
            double fx = -apos[2 * a] + sx[slot[i]], fy = -apos[2 * a + 1] + sy[slot[i]];
            if (disable_red) { fx = -fx; fy = -fy; } /* :186-188 */
            force[2 * a] = fx; force[2 * a + 1] = fy;
        }
    }
    return 0;
}

/* ---------------------------------------------------------------- k-means teams, get_groups cheat_script_ai.py:259-297 */
static double plan_draw(uint64_t seed, int64_t gid, uint32_t counter, int attempt, int d, uint32_t *word) {
    uint32_t w[4];
    orc_philox((uint32_t)gid, counter, (uint32_t)(attempt * 64 + d / 4), SITE_PLAN, (uint32_t)seed, (uint32_t)(seed >> 32), w);
    *word = w[d % 4];
    return (double)w[d % 4] * 0x1p-32;
}

static double sq2(const double *a, const double *b) {
    const double dx = a[0] - b[0], dy = a[1] - b[1];
    double d = 0;
    d += dx * dx;
    d += dy * dy;
    return d;
}

/* kmeans2(pts, k, iter=20, minit='++', missing='raise') retried up to 5 times, then missing='warn' (:276-288).
 * Returns the number of attempts that raised. */
int orc_kmeans(const double *pts, int n, int k, uint64_t seed, int64_t gid, uint32_t counter, int32_t *label, double *cent) {
    if (n > PL_MAX || k > n || k < 1) return -1;
    double cb[2 * PL_MAX], nb[2 * PL_MAX], d2[PL_MAX];
    int cnt[PL_MAX];
    for (int attempt = 0;; ++attempt) {
        /* _kpp: first centre uniform over the points, each next one drawn with probability ~ squared distance to
         * the nearest centre so far (searchsorted over the running sum of probabilities) */
        uint32_t w;
        plan_draw(seed, gid, counter, attempt, 0, &w);
        int pick = (int)(((uint64_t)w * (uint64_t)n) >> 32);
        cb[0] = pts[2 * pick]; cb[1] = pts[2 * pick + 1];
        for (int i = 1; i < k; ++i) {
            /* running sums of d2 in a fixed, warp-shaped order (the draws are keyed, so the order is this
             * restatement's to choose; the CUDA planner uses the same one): 32 "lanes" own ceil(n/32) consecutive
             * points each and sum them in index order, the lane totals go through a Kogge-Stone inclusive scan;
             * the pick is the first point whose running sum reaches r * total (= searchsorted(cumsum(p), r)) */
            const int ch = (n + 31) / 32;
            double pre[PL_MAX], sc[32], nx[32];
            for (int l = 0; l < 32; ++l) {
                double run = 0;
                for (int j = l * ch; j < n && j < (l + 1) * ch; ++j) {
                    double m = INFINITY;
                    for (int c = 0; c < i; ++c) { const double d = sq2(cb + 2 * c, pts + 2 * j); if (d < m) m = d; }
                    d2[j] = m; run += m; pre[j] = run;
                }
                sc[l] = run;
            }
            for (int o = 1; o < 32; o <<= 1) {
                for (int l = 0; l < 32; ++l) nx[l] = l >= o ? sc[l] + sc[l - o] : sc[l];
                memcpy(sc, nx, sizeof(sc));
            }
            const double r = plan_draw(seed, gid, counter, attempt, i, &w);
            const double thr = r * sc[31];
            pick = n - 1;
            for (int j = 0; j < n; ++j) {
                const int l = j / ch;
                const double excl = l > 0 ? sc[l - 1] : 0.0;
                if (excl + pre[j] >= thr) { pick = j; break; }
            }
            cb[2 * i] = pts[2 * pick]; cb[2 * i + 1] = pts[2 * pick + 1];
        }
        int raised = 0;
        for (int it = 0; it < 20 && !raised; ++it) {
            for (int j = 0; j < n; ++j) { /* vq: nearest code, first minimum */
                double best = sq2(pts + 2 * j, cb);
                int bl = 0;
                for (int c = 1; c < k; ++c) { const double d = sq2(pts + 2 * j, cb + 2 * c); if (d < best) { best = d; bl = c; } }
                label[j] = bl;
            }
            for (int c = 0; c < k; ++c) { nb[2 * c] = nb[2 * c + 1] = 0; cnt[c] = 0; } /* update_cluster_means */
            for (int j = 0; j < n; ++j) { nb[2 * label[j]] += pts[2 * j]; nb[2 * label[j] + 1] += pts[2 * j + 1]; ++cnt[label[j]]; }
            for (int c = 0; c < k; ++c) {
                if (cnt[c]) { nb[2 * c] /= cnt[c]; nb[2 * c + 1] /= cnt[c]; }
                else if (attempt < 5) raised = 1;                              /* missing='raise' */
                else { nb[2 * c] = cb[2 * c]; nb[2 * c + 1] = cb[2 * c + 1]; }  /* missing='warn' keeps the old centre */
            }
            if (!raised) memcpy(cb, nb, sizeof(double) * 2 * k);
        }
        if (!raised) { memcpy(cent, cb, sizeof(double) * 2 * k); return attempt; }
    }
}

/* ---------------------------------------------------------------- the planner with its persistent team state */
/* attackers_policy_1 :103-117: teams are formed on the World's first call ever (start_flag survives resets) and kept;
 * if one of those teams has <= 3 members, the grouping is recomputed on EVERY call (team lists never shrink, so the
 * condition never changes).  ps: start_flag u8[E], small u8[E], counter u32[E], label i32[E,A], cent f64[E,K,2]. */
int orc_plan(const OrcConfig *cfg, const OrcState *st, int E, int64_t env_gid0, uint8_t *start_flag, uint8_t *small,
             uint32_t *counter, int32_t *label1, double *cent1, double *att_force, int nthreads) {
    const int G = cfg->num_guards, A = cfg->num_attackers, N = G + A;
    const int K = A / 5 > 0 ? A / 5 : 1;
    (void)nthreads;
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthreads > 0 ? nthreads : 1)
    for (int e = 0; e < E; ++e) {
        const double *gpos = st->pos + (size_t)e * N * 2, *apos = gpos + 2 * G;
        const uint8_t *galive = st->alive + (size_t)e * N, *aalive = galive + G;
        int32_t *l1 = label1 + (size_t)e * A;
        double *c1 = cent1 + (size_t)e * K * 2;
        if (start_flag[e]) {
            orc_kmeans(apos, A, K, cfg->seed, env_gid0 + e, counter[e]++, l1, c1);
            start_flag[e] = 0;
            int cnt[PL_MAX];
            memset(cnt, 0, sizeof(cnt));
            for (int a = 0; a < A; ++a) ++cnt[l1[a]];
            small[e] = 0;
            for (int k = 0; k < K; ++k)
                if (cnt[k] > 0 && cnt[k] <= 3) small[e] = 1;
        }
        int32_t l2[PL_MAX];
        double c2[2 * PL_MAX];
        const int32_t *lab = l1;
        const double *cen = c1;
        if (small[e]) {
            orc_kmeans(apos, A, K, cfg->seed, env_gid0 + e, counter[e]++, l2, c2);
            lab = l2; cen = c2;
        }
        orc_plan_core(G, A, gpos, galive, apos, aalive, lab, K, cen, cfg->disable_red_function,
                      att_force + (size_t)e * A * 2);
    }
    return 0;
}
