// dca_policy.cuh -- observation normalisation for the conc / conc_4hist policy rollout on device tensors
// (SURVEY.md section 8 row f2).  DynamicNorm ALGORITHM/common/norm.py:14-69 (paths relative to /root/reference):
//   training: running mean / variance over the rows of the batch that hold no NaN (exclude_nan), merged with the stored
//             moments by the parallel-variance formula; features that only ever took the values -1 / 0 / 1 keep their
//             initial (0, 1) moments (exclude_one_hot);
//   always:   x <- clip((x - mean) / sqrt(var + 1e-8), -10, 10); Net._act then takes mask = isnan(x).any(-1) and
//             x <- nan_to_num(x, 0) (ALGORITHM/conc_4hist/net.py:196-198).
// Here: one pass that reduces (sum x, sum x^2, #values outside {-1,0,1}) per feature and the row count in fp64 with a
// fixed order (per-CTA partials, then one CTA) -- across GPUs the caller all-reduces those 49 doubles, which is the
// "(sum x, sum x^2, n)" merge SURVEY.md section 8(e) names -- and one pass that applies the stored moments, writes the
// row mask and zeroes the NaNs.  Rows are 16 floats (ScenarioConfig.obs_vec_length, parallel_run.py:70): a lane quad owns
// a row, one float4 per lane, so a warp reads 512 contiguous bytes per step.
#pragma once
#include "dca_device.cuh"

namespace dca {

constexpr int kNormBlocks = 592, kNormThreads = 256, kNormDim = 16, kNormStats = 3 * kNormDim + 1;

__device__ __forceinline__ bool norm_unit(float v) { return v == 0.f || v == 1.f || v == -1.f; }

__global__ void __launch_bounds__(kNormThreads) k_norm_partial(const float4 *__restrict__ x, size_t rows,
                                                               double *__restrict__ partial) {
    __shared__ double sh[kNormThreads / 32][4][13];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, quad = lane & 3;
    const unsigned full = 0xffffffffu;
    double s[4] = {0, 0, 0, 0}, q[4] = {0, 0, 0, 0}, nu[4] = {0, 0, 0, 0}, cnt = 0;
    const size_t rows_per_pass = (size_t)kNormBlocks * kNormThreads / 4;
    const size_t first = ((size_t)blockIdx.x * kNormThreads + (tid & ~31)) / 4;  // first row of this warp
    constexpr int U = 4;  // rows in flight per lane quad (one float4 per thread and pass kept HBM at 3 TB/s); the sums
                          // take the rows in the same order as a one-row loop
    for (size_t base = first; base < rows; base += U * rows_per_pass) {           // warp-uniform
        float4 v[U];
        bool in[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const size_t row = base + u * rows_per_pass + (lane >> 2);
            in[u] = row < rows;
            v[u] = in[u] ? x[row * 4 + quad] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const bool bad = !in[u] || (v[u].x != v[u].x) || (v[u].y != v[u].y) || (v[u].z != v[u].z) || (v[u].w != v[u].w);
            const unsigned b = __ballot_sync(full, bad);
            if ((b >> (lane & ~3)) & 0xfu) continue;  // exclude_nan: a row with any NaN contributes nothing
            const float e[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const double d = (double)e[k];
                s[k] += d; q[k] += d * d;
                if (!norm_unit(e[k])) nu[k] += 1.0;
            }
            if (quad == 0) cnt += 1.0;
        }
    }
    // lanes with the same quad hold the same four features: fold lane bits 2..4
#pragma unroll
    for (int o = 4; o < 32; o <<= 1) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            s[k] += __shfl_xor_sync(full, s[k], o);
            q[k] += __shfl_xor_sync(full, q[k], o);
            nu[k] += __shfl_xor_sync(full, nu[k], o);
        }
        cnt += __shfl_xor_sync(full, cnt, o);
    }
    if (lane < 4) {
#pragma unroll
        for (int k = 0; k < 4; ++k) { sh[warp][lane][k] = s[k]; sh[warp][lane][4 + k] = q[k]; sh[warp][lane][8 + k] = nu[k]; }
        sh[warp][lane][12] = cnt;
    }
    __syncthreads();
    if (tid < kNormStats) {  // stats layout: sum[16], sumsq[16], nonunit[16], count
        const int grp = tid / kNormDim, f = tid % kNormDim;
        double a = 0;
        for (int w = 0; w < kNormThreads / 32; ++w)
            a += tid == 3 * kNormDim ? sh[w][0][12] : sh[w][f >> 2][grp * 4 + (f & 3)];
        partial[(size_t)blockIdx.x * kNormStats + tid] = a;
    }
}

// 8 threads per statistic (49 x 8 = 392): each sums every eighth block's partial, then three shuffle steps in the group
// (one thread per statistic walked 592 dependent additions: 21 us)
constexpr int kNormFinalThreads = 8 * ((kNormStats * 8 + 63) / 64) * 8;  // 392 rounded up to whole warps
__global__ void __launch_bounds__(kNormFinalThreads) k_norm_final(const double *__restrict__ partial, double *__restrict__ stats) {
    const int t = threadIdx.x >> 3, part = threadIdx.x & 7;
    double a = 0;
    if (t < kNormStats) {
#pragma unroll 4
        for (int b = part; b < kNormBlocks; b += 8) a += partial[(size_t)b * kNormStats + t];
    }
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
    if (t < kNormStats && part == 0) stats[t] = a;
}

// y = nan_to_num(clip((x - mean) / sqrt(var + 1e-8), -10, 10), 0), row_nan[r] = any NaN in row r.  In place when y == x.
__global__ void __launch_bounds__(256) k_norm_apply(const float4 *x, const float *__restrict__ mean,
                                                    const float *__restrict__ var, size_t rows, float4 *y,
                                                    uint8_t *__restrict__ row_nan) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, quad = lane & 3;
    const float4 m = ((const float4 *)mean)[quad], vv = ((const float4 *)var)[quad];
    const float sd[4] = {__fsqrt_rn(__fadd_rn(vv.x, 1e-8f)), __fsqrt_rn(__fadd_rn(vv.y, 1e-8f)),
                         __fsqrt_rn(__fadd_rn(vv.z, 1e-8f)), __fsqrt_rn(__fadd_rn(vv.w, 1e-8f))};
    const float mu[4] = {m.x, m.y, m.z, m.w};
    const size_t rows_per_pass = (size_t)gridDim.x * blockDim.x / 4;
    const size_t first = ((size_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31)) / 4;
    for (size_t base = first; base < rows; base += rows_per_pass) {  // warp-uniform
        const size_t row = base + (lane >> 2);
        const bool in = row < rows;
        float e[4] = {0.f, 0.f, 0.f, 0.f};
        if (in) {
            const float4 v = x[row * 4 + quad];
            e[0] = v.x; e[1] = v.y; e[2] = v.z; e[3] = v.w;
        }
        bool bad = false;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (e[k] != e[k]) { bad = true; e[k] = 0.f; }
            else e[k] = fminf(fmaxf(__fdiv_rn(__fsub_rn(e[k], mu[k]), sd[k]), -10.f), 10.f);
        }
        const unsigned b = __ballot_sync(full, bad);
        if (in) {
            y[row * 4 + quad] = make_float4(e[0], e[1], e[2], e[3]);
            if (quad == 0) row_nan[row] = ((b >> lane) & 0xfu) != 0;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Concentration.forward ALGORITHM/common/conc.py:28-63 between its matrix products, for one list of entities of every
// agent (conc_4hist: entities 1..11 = "current", 12..23 = "history", ALGORITHM/conc_4hist/net.py:170-181): attention
// scores of the agent's own encoding against its entities, masked softmax (no live entity: all zero), the attended
// encoding sum_j score_j ve_j (the caller multiplies by W_val), and the forward branch's input -- the agent's own
// [encoding | observation] followed by those of the n_focus entities with the highest scores.  Ties (float32 softmax
// gives several live entities exactly 0 when the attention is peaked) go to the lowest entity index: torch.argsort's
// order among equal scores is unspecified and differs between its CPU and CUDA sorts; the stable order is what the
// reference computes on the CPU.
// One warp per agent.  v [B,24,48] encodings, x [B,24,16] normalised observations (NaN -> 0), dead u8 [B,24],
// qk [B,48] = (vs Wq) Wk^T; out: va [B,48], stack [B,(n_focus+1)*64].
constexpr int kPolH = 48, kPolD = 16, kPolEnt = 24;

__global__ void __launch_bounds__(256) k_conc_attend(const float *__restrict__ v, const float *__restrict__ x,
                                                      const uint8_t *__restrict__ dead, const float *__restrict__ qk,
                                                      int first, int n, int n_focus, size_t B, float *__restrict__ va,
                                                      float *__restrict__ stack) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const size_t b = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (b >= B) return;  // warp-uniform
    const float *vb = v + b * kPolEnt * kPolH;
    const float *xb = x + b * kPolEnt * kPolD;
    // scores: lane j < n owns entity first + j
    float compat = -CUDART_INF_F;
    if (lane < n && !dead[b * kPolEnt + first + lane]) {
        const float4 *row = (const float4 *)(vb + (first + lane) * kPolH);
        const float4 *q4 = (const float4 *)(qk + b * kPolH);
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < kPolH / 4; ++k) {
            const float4 a = row[k], q = q4[k];
            acc = fmaf(a.x, q.x, acc); acc = fmaf(a.y, q.y, acc); acc = fmaf(a.z, q.z, acc); acc = fmaf(a.w, q.w, acc);
        }
        compat = acc * 0.14433756729740643f;  // 1 / sqrt(48)
    }
    float mx = compat;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(full, mx, o));
    float ex = (lane < n && compat > -CUDART_INF_F) ? expf(compat - mx) : 0.f;
    float sum = ex;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(full, sum, o);
    const float score = sum > 0.f ? __fdiv_rn(ex, sum) : 0.f;  // nan_to_num(softmax(all -inf)) = 0
    // attended encoding: lanes own dims lane and lane + 32 (< 48)
    {
        float a0 = 0.f, a1 = 0.f;
        for (int j = 0; j < n; ++j) {
            const float sj = __shfl_sync(full, score, j);
            const float *row = vb + (first + j) * kPolH;
            a0 = fmaf(sj, row[lane], a0);
            if (lane < kPolH - 32) a1 = fmaf(sj, row[32 + lane], a1);
        }
        va[b * kPolH + lane] = a0;
        if (lane < kPolH - 32) va[b * kPolH + 32 + lane] = a1;
    }
    // forward branch input: [vs | zs], then the n_focus best entities in stable descending score order
    const int row_len = kPolH + kPolD;  // 64
    float *sb = stack + b * (size_t)(n_focus + 1) * row_len;
    {
        sb[lane] = vb[lane];
        if (lane < kPolH - 32) sb[32 + lane] = vb[32 + lane];
        else sb[32 + lane] = xb[lane - (kPolH - 32)];  // lanes 16..31 -> x[0..15]
    }
    unsigned taken = 0;
    for (int r = 1; r <= n_focus; ++r) {
        // the highest remaining score; scores are >= 0, so their bit patterns order like the values
        const unsigned key = (lane < n && !((taken >> lane) & 1u)) ? __float_as_uint(score) + 1u : 0u;
        const unsigned best = __reduce_max_sync(full, key);
        const int j = __ffs(__ballot_sync(full, key == best && key != 0u)) - 1;  // lowest index among equals
        taken |= 1u << j;
        const float *vr = vb + (first + j) * kPolH;
        const float *xr = xb + (first + j) * kPolD;
        float *o = sb + r * row_len;
        o[lane] = vr[lane];
        if (lane < kPolH - 32) o[32 + lane] = vr[32 + lane];
        else o[32 + lane] = xr[lane - (kPolH - 32)];
    }
}

// Both entity lists of an agent in one pass (conc_4hist: "current" then "history", adjacent in the 24 rows): the warp
// first pulls the agent's whole block -- 24 encodings (4.6 KB, contiguous) and 24 observation rows (1.5 KB) -- into its
// slice of shared memory with 16-byte cp.async copies, all in flight at once and fully coalesced (k_conc_attend's lanes
// each walk their own 192-byte row: 12 dependent, uncoalesced loads per lane, DRAM at 44 % of peak), then does exactly
// k_conc_attend's arithmetic in k_conc_attend's order for list a and for list b (same butterflies over the same lanes:
// the results are bit-identical), reading rows from shared memory (row stride 52 floats: conflict-free 16-byte reads).
// Lists: entities first_a .. first_a+n_a-1 and first_a+n_a .. first_a+n_a+n_b-1; n_a + n_b <= 32.
__device__ __forceinline__ void enc_cp_async16(void *smem_dst, const void *gsrc);
constexpr int kAttVs = 52;
struct AttWarpSmem {
    float v[kPolEnt][kAttVs];
    float x[kPolEnt][kPolD];
};
__global__ void __launch_bounds__(256) k_conc_attend2(const float *__restrict__ v, const float *__restrict__ x,
                                                       const uint8_t *__restrict__ dead, const float *__restrict__ qk_a,
                                                       const float *__restrict__ qk_b, int first_a, int n_a, int focus_a,
                                                       int n_b, int focus_b, size_t B, float *__restrict__ va_a,
                                                       float *__restrict__ stack_a, float *__restrict__ va_b,
                                                       float *__restrict__ stack_b) {
    extern __shared__ __align__(16) unsigned char att_raw[];
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const size_t b = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (b >= B) return;  // warp-uniform
    AttWarpSmem &s = reinterpret_cast<AttWarpSmem *>(att_raw)[warp];
    {
        const float *vb = v + b * kPolEnt * kPolH, *xb = x + b * kPolEnt * kPolD;
        for (int i = lane; i < kPolEnt * (kPolH / 4); i += 32) {
            const int r = i / (kPolH / 4), q = i - r * (kPolH / 4);
            enc_cp_async16(&s.v[r][q * 4], vb + i * 4);
        }
        for (int i = lane; i < kPolEnt * (kPolD / 4); i += 32) enc_cp_async16(&s.x[0][0] + i * 4, xb + i * 4);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    // dot products: lane L < n_a + n_b owns entity first_a + L (list a, then list b)
    const int n_ab = n_a + n_b;
    const bool mine = lane < n_ab && !dead[b * kPolEnt + first_a + lane];
    const float4 *q4 = (const float4 *)((lane < n_a ? qk_a : qk_b) + b * kPolH);
    float4 qv[kPolH / 4];
#pragma unroll
    for (int k = 0; k < kPolH / 4; ++k) qv[k] = q4[k];
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    float dot = -CUDART_INF_F;
    if (mine) {
        const float4 *row = (const float4 *)&s.v[first_a + lane][0];
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < kPolH / 4; ++k) {
            const float4 a = row[k], q = qv[k];
            acc = fmaf(a.x, q.x, acc); acc = fmaf(a.y, q.y, acc); acc = fmaf(a.z, q.z, acc); acc = fmaf(a.w, q.w, acc);
        }
        dot = acc * 0.14433756729740643f;  // 1 / sqrt(48)
    }
    const float dot_b = __shfl_sync(full, dot, (lane + n_a) & 31);  // list b's scores moved to lanes 0 .. n_b-1
#pragma unroll 1
    for (int list = 0; list < 2; ++list) {
        const int first = list == 0 ? first_a : first_a + n_a, n = list == 0 ? n_a : n_b, n_focus = list == 0 ? focus_a : focus_b;
        float *va = list == 0 ? va_a : va_b, *stack = list == 0 ? stack_a : stack_b;
        const float compat = lane < n ? (list == 0 ? dot : dot_b) : -CUDART_INF_F;
        float mx = compat;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(full, mx, o));
        const float ex = (lane < n && compat > -CUDART_INF_F) ? expf(compat - mx) : 0.f;
        float sum = ex;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(full, sum, o);
        const float score = sum > 0.f ? __fdiv_rn(ex, sum) : 0.f;
        {
            float a0 = 0.f, a1 = 0.f;
            for (int j = 0; j < n; ++j) {
                const float sj = __shfl_sync(full, score, j);
                const float *row = s.v[first + j];
                a0 = fmaf(sj, row[lane], a0);
                if (lane < kPolH - 32) a1 = fmaf(sj, row[32 + lane], a1);
            }
            va[b * kPolH + lane] = a0;
            if (lane < kPolH - 32) va[b * kPolH + 32 + lane] = a1;
        }
        const int row_len = kPolH + kPolD;  // 64
        float *sb = stack + b * (size_t)(n_focus + 1) * row_len;
        {
            sb[lane] = s.v[0][lane];
            if (lane < kPolH - 32) sb[32 + lane] = s.v[0][32 + lane];
            else sb[32 + lane] = s.x[0][lane - (kPolH - 32)];
        }
        unsigned taken = 0;
        for (int r = 1; r <= n_focus; ++r) {
            const unsigned key = (lane < n && !((taken >> lane) & 1u)) ? __float_as_uint(score) + 1u : 0u;
            const unsigned best = __reduce_max_sync(full, key);
            const int j = __ffs(__ballot_sync(full, key == best && key != 0u)) - 1;
            taken |= 1u << j;
            const float *vr = s.v[first + j], *xr = s.x[first + j];
            float *o = sb + r * row_len;
            o[lane] = vr[lane];
            if (lane < kPolH - 32) o[32 + lane] = vr[32 + lane];
            else o[32 + lane] = xr[lane - (kPolH - 32)];
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// The observation encoder of the conc / conc_4hist policy on every entity row (Net._act ALGORITHM/conc_4hist/net.py:
// 194-199): DynamicNorm apply + mask + nan_to_num (as k_norm_apply) + AT_obs_encoder = Linear(16,48), ReLU,
// Linear(48,48), in one pass.  This is 60 % of the policy's arithmetic (24 entity rows per agent, 3 072 multiply-adds
// each) and in torch four passes over a [rows, 48] activation plus two bias-epilogue kernels.  float32 FFMA (the policy's
// attention logits reach the hundreds: tf32 inputs would move its picks), register-blocked like an SGEMM: a CTA owns a
// tile of 128 rows, the normalised rows and the hidden activations sit in shared memory k-major, a thread accumulates
// 8 rows x 6 columns (5 shared loads per 48 FFMA), the grid is persistent so the 12 KB of weights are staged once.
// nn.Linear layout: w1 [48,16], w2 [48,48] (out, in).
constexpr int kEncTile = 128, kEncThreads = 128;
struct EncSmem {
    float xs[kPolD][kEncTile + 4];   // normalised observation rows, k-major (+4: the transposing stores of a warp hit
    float hs[kPolH][kEncTile + 4];   // hidden activations after the ReLU    2 banks' worth of rows instead of 8)
    float w1[kPolD][kPolH];      // w1^T
    float w2[kPolH][kPolH];      // w2^T
    float b1[kPolH], b2[kPolH];
    float4 stage[kEncTile * 4];      // the next tile's raw rows, fetched with cp.async while this tile is computed
};
__device__ __forceinline__ void enc_cp_async16(void *smem_dst, const void *gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc)
                 : "memory");
}

__global__ void __launch_bounds__(kEncThreads) k_policy_encode(const float4 *__restrict__ x, const float *__restrict__ mean,
                                                               const float *__restrict__ var,
                                                               const float *__restrict__ w1, const float *__restrict__ b1,
                                                               const float *__restrict__ w2, const float *__restrict__ b2,
                                                               size_t rows, float4 *__restrict__ x_norm,
                                                               uint8_t *__restrict__ row_nan, float *__restrict__ v) {
    extern __shared__ __align__(16) unsigned char enc_raw[];
    EncSmem &s = *reinterpret_cast<EncSmem *>(enc_raw);
    const unsigned full = 0xffffffffu;
    const int tid = threadIdx.x, lane = tid & 31, quad = tid & 3;
    for (int i = tid; i < kPolD * kPolH; i += kEncThreads) s.w1[i % kPolD][i / kPolD] = w1[i];  // w1[j][k] -> w1s[k][j]
    for (int i = tid; i < kPolH * kPolH; i += kEncThreads) s.w2[i % kPolH][i / kPolH] = w2[i];
    if (tid < kPolH) { s.b1[tid] = b1[tid]; s.b2[tid] = b2[tid]; }
    const float4 m = ((const float4 *)mean)[quad], vv = ((const float4 *)var)[quad];
    const float sd[4] = {__fsqrt_rn(__fadd_rn(vv.x, 1e-8f)), __fsqrt_rn(__fadd_rn(vv.y, 1e-8f)),
                         __fsqrt_rn(__fadd_rn(vv.z, 1e-8f)), __fsqrt_rn(__fadd_rn(vv.w, 1e-8f))};
    const float mu[4] = {m.x, m.y, m.z, m.w};
    const int rg = tid >> 3, cg = tid & 7;  // this thread's 8 rows (rg*8 ..) and 6 columns (cg*6 ..)
    const size_t tiles = (rows + kEncTile - 1) / kEncTile;
    auto prefetch = [&](size_t t) {  // every thread stages the four float4 it will read itself
        if (t < tiles) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int i = tid + q * kEncThreads;
                const size_t row = t * kEncTile + (i >> 2);
                if (row < rows) enc_cp_async16(&s.stage[i], &x[row * 4 + quad]);
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    prefetch(blockIdx.x);
    for (size_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const size_t r0 = tile * kEncTile;
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        float4 raw[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) raw[q] = s.stage[tid + q * kEncThreads];
        prefetch(tile + gridDim.x);  // own slots only: they are in registers now
        __syncthreads();  // the previous tile's shared activations have been consumed; weights are in place
        // ---- load + normalise 128 rows (4 float4 per thread; a lane quad holds one row)
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int i = tid + q * kEncThreads, lr = i >> 2;  // local row, quad = i & 3 = tid & 3
            const size_t row = r0 + lr;
            float e[4] = {0.f, 0.f, 0.f, 0.f};
            const bool in = row < rows;
            if (in) {
                const float4 t = raw[q];
                e[0] = t.x; e[1] = t.y; e[2] = t.z; e[3] = t.w;
            }
            bool bad = false;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (e[k] != e[k]) { bad = true; e[k] = 0.f; }
                else e[k] = fminf(fmaxf(__fdiv_rn(__fsub_rn(e[k], mu[k]), sd[k]), -10.f), 10.f);
            }
            const unsigned bb = __ballot_sync(full, bad);
            if (in) {
                x_norm[row * 4 + quad] = make_float4(e[0], e[1], e[2], e[3]);
                if (quad == 0) row_nan[row] = ((bb >> lane) & 0xfu) != 0;
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) s.xs[quad * 4 + k][lr] = e[k];
        }
        __syncthreads();
        // ---- layer 1: h = relu(x w1^T + b1)
        float acc[8][6];
#pragma unroll
        for (int c = 0; c < 6; ++c) {
            const float bias = s.b1[cg * 6 + c];
#pragma unroll
            for (int r = 0; r < 8; ++r) acc[r][c] = bias;
        }
#pragma unroll 4
        for (int k = 0; k < kPolD; ++k) {
            const float4 a0 = *(const float4 *)&s.xs[k][rg * 8], a1 = *(const float4 *)&s.xs[k][rg * 8 + 4];
            const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float2 w01 = *(const float2 *)&s.w1[k][cg * 6], w23 = *(const float2 *)&s.w1[k][cg * 6 + 2],
                         w45 = *(const float2 *)&s.w1[k][cg * 6 + 4];
            const float w[6] = {w01.x, w01.y, w23.x, w23.y, w45.x, w45.y};
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int c = 0; c < 6; ++c) acc[r][c] = fmaf(a[r], w[c], acc[r][c]);
        }
#pragma unroll
        for (int c = 0; c < 6; ++c) {
            *(float4 *)&s.hs[cg * 6 + c][rg * 8] = make_float4(fmaxf(acc[0][c], 0.f), fmaxf(acc[1][c], 0.f),
                                                              fmaxf(acc[2][c], 0.f), fmaxf(acc[3][c], 0.f));
            *(float4 *)&s.hs[cg * 6 + c][rg * 8 + 4] = make_float4(fmaxf(acc[4][c], 0.f), fmaxf(acc[5][c], 0.f),
                                                                  fmaxf(acc[6][c], 0.f), fmaxf(acc[7][c], 0.f));
        }
        __syncthreads();
        // ---- layer 2: v = h w2^T + b2
#pragma unroll
        for (int c = 0; c < 6; ++c) {
            const float bias = s.b2[cg * 6 + c];
#pragma unroll
            for (int r = 0; r < 8; ++r) acc[r][c] = bias;
        }
#pragma unroll 4
        for (int k = 0; k < kPolH; ++k) {
            const float4 a0 = *(const float4 *)&s.hs[k][rg * 8], a1 = *(const float4 *)&s.hs[k][rg * 8 + 4];
            const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float2 w01 = *(const float2 *)&s.w2[k][cg * 6], w23 = *(const float2 *)&s.w2[k][cg * 6 + 2],
                         w45 = *(const float2 *)&s.w2[k][cg * 6 + 4];
            const float w[6] = {w01.x, w01.y, w23.x, w23.y, w45.x, w45.y};
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int c = 0; c < 6; ++c) acc[r][c] = fmaf(a[r], w[c], acc[r][c]);
        }
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            const size_t row = r0 + rg * 8 + r;
            if (row < rows) {
                float2 *o = (float2 *)(v + row * kPolH + cg * 6);
                o[0] = make_float2(acc[r][0], acc[r][1]);
                o[1] = make_float2(acc[r][2], acc[r][3]);
                o[2] = make_float2(acc[r][4], acc[r][5]);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// One dense layer of the conc / conc_4hist policy on every agent row: out = act(a Wa^T + b Wb^T + bias), N <= 48 output
// features, for the layers behind the attention (Concentration's query / key / value / motivate / forward layers
// ALGORITHM/common/conc.py:30-62, the logit / value / threat heads ALGORITHM/conc_4hist/net.py:213-229).  The second
// input replaces the reference's torch.cat((a, b), -1) in front of a Linear.  Why not cuBLAS: its float32 SIMT tiles
// are 64 x 128 and N is 48 (24, 7, 1 in the heads), so 62 % of every tile is padding, and bias and ReLU are separate
// passes over the activation.  Here a CTA owns 128 rows x all N columns, register-blocked like k_policy_encode (a
// thread accumulates 8 rows x NC columns, N <= 8 NC); inputs and weights come through shared memory in chunks of 32
// features, two stages deep (cp.async: chunk c + 1 is in flight while chunk c is multiplied); float32 FFMA in ascending k.
// a: [rows, Ka] with row stride lda (a column slice of a wider tensor is fine), b likewise (Kb = 0: none); weights are
// addressed W[n * sn + k * sk], so nn.Linear weights, their column halves and transposed matrices need no copies.
constexpr int kLinTile = 128, kLinThreads = 128, kLinKC = 32, kLinStride = kLinKC + 4;
struct LinSmem {                              // 48 KB: two stages of {128 input rows x 32 features, 32 x 48 weights}
    float xs[2][kLinTile][kLinStride];        // row-major as in global memory (cp.async, no transposition); a thread's
    float ws[2][kLinKC][48];                  // rows are rg, rg + 16, ..: the 4 row groups of a warp sit 144 B apart
};
struct LinSrc {
    const float *x;
    const float *w;
    int ld, K, sn, sk;
};
__device__ __forceinline__ void lin_cp_async4(void *smem_dst, const void *gsrc) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc)
                 : "memory");
}

template <int NC>
__global__ void __launch_bounds__(kLinThreads) k_policy_linear(const LinSrc sa, const LinSrc sb, const float *__restrict__ bias,
                                                               int N, int relu, size_t rows, float *__restrict__ out, int ldo) {
    __shared__ __align__(16) LinSmem s;
    constexpr int NP = 8 * NC;  // padded width
    const int tid = threadIdx.x;
    const int rg = tid >> 3, cg = tid & 7;
    const size_t tiles = (rows + kLinTile - 1) / kLinTile;
    const int chunks_a = (sa.K + kLinKC - 1) / kLinKC, chunks = chunks_a + (sb.K + kLinKC - 1) / kLinKC;
    float bs[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) bs[c] = (bias != nullptr && cg * NC + c < N) ? bias[cg * NC + c] : 0.f;
    const bool vec2 = (ldo & 1) == 0 && ((uintptr_t)out & 7) == 0;
    // stage chunk c of tile t into buffer b (asynchronous; rows past the end repeat the last row, never stored)
    auto stage = [&](size_t t, int c, int b) {
        if (t < tiles) {
            const LinSrc &S = c < chunks_a ? sa : sb;
            const int k0 = (c < chunks_a ? c : c - chunks_a) * kLinKC, kc = min(kLinKC, S.K - k0), q4 = kc >> 2;
            const size_t r0 = t * kLinTile;
            if ((q4 & (q4 - 1)) == 0) {  // 8 / 4 / 2 pieces of 16 bytes per row (chunks of 32 / 16 / 8 features): shifts
                const int sh = __ffs(q4) - 1, q = tid & (q4 - 1), step = kLinThreads >> sh;
                const float *src = S.x + k0 + q * 4;
                for (int lr = tid >> sh; lr < kLinTile; lr += step) {
                    const size_t row = min(r0 + lr, rows - 1);
                    enc_cp_async16(&s.xs[b][lr][q * 4], src + row * (size_t)S.ld);
                }
            } else {
                for (int i = tid; i < kLinTile * q4; i += kLinThreads) {
                    const int lr = i / q4, q = i - lr * q4;
                    const size_t row = min(r0 + lr, rows - 1);
                    enc_cp_async16(&s.xs[b][lr][q * 4], S.x + row * (size_t)S.ld + k0 + q * 4);
                }
            }
            for (int i = tid; i < kc * NP; i += kLinThreads) {
                const int k = i / NP, n = i - k * NP;
                if (n < N) lin_cp_async4(&s.ws[b][k][n], S.w + (size_t)n * S.sn + (size_t)(k0 + k) * S.sk);
                else s.ws[b][k][n] = 0.f;
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    int buf = 0;
    stage(blockIdx.x, 0, 0);
    for (size_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const size_t r0 = tile * kLinTile;
        float acc[8][NC];
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int c = 0; c < NC; ++c) acc[r][c] = bs[c];
        for (int c = 0; c < chunks; ++c) {
            if (c + 1 < chunks) stage(tile, c + 1, buf ^ 1);
            else stage(tile + gridDim.x, 0, buf ^ 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
            __syncthreads();  // chunk c has landed for every thread
            const LinSrc &S = c < chunks_a ? sa : sb;
            const int k0 = (c < chunks_a ? c : c - chunks_a) * kLinKC, kc = min(kLinKC, S.K - k0);
            for (int k4 = 0; k4 < kc; k4 += 4) {
                float4 a[8];
#pragma unroll
                for (int r = 0; r < 8; ++r) a[r] = *(const float4 *)&s.xs[buf][rg + 16 * r][k4];
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                    float w[NC];
                    if constexpr (NC == 6) {
                        const float2 w01 = *(const float2 *)&s.ws[buf][k4 + kk][cg * 6],
                                     w23 = *(const float2 *)&s.ws[buf][k4 + kk][cg * 6 + 2],
                                     w45 = *(const float2 *)&s.ws[buf][k4 + kk][cg * 6 + 4];
                        w[0] = w01.x; w[1] = w01.y; w[2] = w23.x; w[3] = w23.y; w[4] = w45.x; w[5] = w45.y;
                    } else {
#pragma unroll
                        for (int cc = 0; cc < NC; ++cc) w[cc] = s.ws[buf][k4 + kk][cg * NC + cc];
                    }
#pragma unroll
                    for (int r = 0; r < 8; ++r) {
                        const float ar = kk == 0 ? a[r].x : (kk == 1 ? a[r].y : (kk == 2 ? a[r].z : a[r].w));
#pragma unroll
                        for (int cc = 0; cc < NC; ++cc) acc[r][cc] = fmaf(ar, w[cc], acc[r][cc]);
                    }
                }
            }
            __syncthreads();  // buffer `buf` is free for the chunk after next
            buf ^= 1;
        }
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            const size_t row = r0 + rg + 16 * r;
            if (row < rows) {
                float *o = out + row * (size_t)ldo + cg * NC;
                if (NC == 6 && N == 48 && vec2) {  // full-width layer, 8-byte aligned rows: three 8-byte stores
#pragma unroll
                    for (int c = 0; c < NC; c += 2) {
                        float2 v2 = make_float2(acc[r][c], acc[r][c + 1]);
                        if (relu) { v2.x = fmaxf(v2.x, 0.f); v2.y = fmaxf(v2.y, 0.f); }
                        *(float2 *)(o + c) = v2;
                    }
                } else {
#pragma unroll
                    for (int c = 0; c < NC; ++c)
                        if (cg * NC + c < N) o[c] = relu ? fmaxf(acc[r][c], 0.f) : acc[r][c];
                }
            }
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// The same layer on the tensor cores with float32-class accuracy ("3xTF32"): every operand is split into a tf32 head
// and a tf32 tail (x = hi + lo exactly to 2^-22 |x|), and a product is three mma.sync.m16n8k8 tf32 instructions --
// lo*hi + hi*lo + hi*hi, small terms first, float32 accumulators; the dropped lo*lo term is 2^-22 of the product, the
// size of float32 rounding itself (a single tf32 pass, 2^-11, would move the attention's picks: its logits reach the
// hundreds).  The FFMA kernel above is bound by shared-memory operand bandwidth (14 floats per 48 FFMA per thread) at a
// third of the FP32 peak; here a warp owns 32 rows x all N columns (2 x NT accumulator tiles) and loads 20 words per
// 8 k-steps.  Staging as above (row-major rows with a 36-float stride and k-major weights with a 56-float stride make
// every fragment load conflict-free).  Needs ka, kb multiples of 8.
constexpr int kLinWStride = 56;
struct LinSmemTc {
    float xs[2][kLinTile][kLinStride];
    float ws[2][kLinKC][kLinWStride];
};
// head: x rounded to tf32's 10 mantissa bits, nearest with ties away from zero -- what cvt.rna.tf32.f32 computes, done
// with two integer instructions (cvt runs on the quarter-rate conversion pipe, which bound the first version of this
// kernel at 54 % busy with the tensor pipe at 5 %); tail: x - head (exact), rounded the same way (left to the tensor
// core's truncation, 2^-21 |x| instead of 2^-22, two of the 1000 reference agents took another pick).
__device__ __forceinline__ uint32_t tf32_rna(float x) { return (__float_as_uint(x) + 0x1000u) & 0xffffe000u; }
__device__ __forceinline__ void tf32_split(float x, uint32_t &hi, uint32_t &lo) {
    hi = tf32_rna(x);
    lo = tf32_rna(x - __uint_as_float(hi));
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

template <int NT>
__global__ void __launch_bounds__(kLinThreads) k_policy_linear_tc(const LinSrc sa, const LinSrc sb,
                                                                  const float *__restrict__ bias, int N, int relu, size_t rows,
                                                                  float *__restrict__ out, int ldo) {
    extern __shared__ __align__(16) unsigned char lin_raw[];
    LinSmemTc &s = *reinterpret_cast<LinSmemTc *>(lin_raw);
    constexpr int NP = 8 * NT;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const size_t tiles = (rows + kLinTile - 1) / kLinTile;
    const int chunks_a = (sa.K + kLinKC - 1) / kLinKC, chunks = chunks_a + (sb.K + kLinKC - 1) / kLinKC;
    float bs[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int j = 0; j < 2; ++j) bs[nt][j] = (bias != nullptr && nt * 8 + 2 * t + j < N) ? bias[nt * 8 + 2 * t + j] : 0.f;
    const bool vec2 = (ldo & 1) == 0 && ((uintptr_t)out & 7) == 0;
    auto stage = [&](size_t tl, int c, int b) {
        if (tl < tiles) {
            const LinSrc &S = c < chunks_a ? sa : sb;
            const int k0 = (c < chunks_a ? c : c - chunks_a) * kLinKC, kc = min(kLinKC, S.K - k0), q4 = kc >> 2;
            const size_t r0 = tl * kLinTile;
            if ((q4 & (q4 - 1)) == 0) {  // 8 / 4 / 2 pieces of 16 bytes per row (chunks of 32 / 16 / 8 features): shifts
                const int sh = __ffs(q4) - 1, q = tid & (q4 - 1), step = kLinThreads >> sh;
                const float *src = S.x + k0 + q * 4;
                for (int lr = tid >> sh; lr < kLinTile; lr += step) {
                    const size_t row = min(r0 + lr, rows - 1);
                    enc_cp_async16(&s.xs[b][lr][q * 4], src + row * (size_t)S.ld);
                }
            } else {
                for (int i = tid; i < kLinTile * q4; i += kLinThreads) {
                    const int lr = i / q4, q = i - lr * q4;
                    const size_t row = min(r0 + lr, rows - 1);
                    enc_cp_async16(&s.xs[b][lr][q * 4], S.x + row * (size_t)S.ld + k0 + q * 4);
                }
            }
            for (int i = tid; i < kc * NP; i += kLinThreads) {
                const int k = i / NP, n = i - k * NP;
                if (n < N) lin_cp_async4(&s.ws[b][k][n], S.w + (size_t)n * S.sn + (size_t)(k0 + k) * S.sk);
                else s.ws[b][k][n] = 0.f;
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    int buf = 0;
    stage(blockIdx.x, 0, 0);
    for (size_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const size_t r0 = tile * kLinTile;
        float acc[2][NT][4];
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                acc[mt][nt][0] = acc[mt][nt][2] = bs[nt][0];
                acc[mt][nt][1] = acc[mt][nt][3] = bs[nt][1];
            }
        for (int c = 0; c < chunks; ++c) {
            if (c + 1 < chunks) stage(tile, c + 1, buf ^ 1);
            else stage(tile + gridDim.x, 0, buf ^ 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
            __syncthreads();
            const LinSrc &S = c < chunks_a ? sa : sb;
            const int k0 = (c < chunks_a ? c : c - chunks_a) * kLinKC, kc = min(kLinKC, S.K - k0);
            for (int ks = 0; ks < kc; ks += 8) {
                uint32_t bh[NT][2], bl[NT][2];
#pragma unroll
                for (int nt = 0; nt < NT; ++nt)
#pragma unroll
                    for (int j = 0; j < 2; ++j) tf32_split(s.ws[buf][ks + t + 4 * j][nt * 8 + g], bh[nt][j], bl[nt][j]);
#pragma unroll
                for (int mt = 0; mt < 2; ++mt) {
                    const int row = warp * 32 + mt * 16 + g;
                    uint32_t ah[4], al[4];
                    tf32_split(s.xs[buf][row][ks + t], ah[0], al[0]);
                    tf32_split(s.xs[buf][row + 8][ks + t], ah[1], al[1]);
                    tf32_split(s.xs[buf][row][ks + t + 4], ah[2], al[2]);
                    tf32_split(s.xs[buf][row + 8][ks + t + 4], ah[3], al[3]);
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt) {
                        mma_tf32(acc[mt][nt], al, bh[nt]);
                        mma_tf32(acc[mt][nt], ah, bl[nt]);
                        mma_tf32(acc[mt][nt], ah, bh[nt]);
                    }
                }
            }
            __syncthreads();
            buf ^= 1;
        }
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const size_t row = r0 + warp * 32 + mt * 16 + g + 8 * h;
                if (row >= rows) continue;
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    const int n = nt * 8 + 2 * t;
                    float v0 = acc[mt][nt][2 * h], v1 = acc[mt][nt][2 * h + 1];
                    if (relu) { v0 = fmaxf(v0, 0.f); v1 = fmaxf(v1, 0.f); }
                    float *o = out + row * (size_t)ldo + n;
                    if (vec2 && n + 1 < N) {
                        *(float2 *)o = make_float2(v0, v1);
                    } else {
                        if (n < N) o[0] = v0;
                        if (n + 1 < N) o[1] = v1;
                    }
                }
            }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
}

// ---------------------------------------------------------------------------------------------------------------
// k_policy_encode on the tensor cores (3xTF32 as in k_policy_linear_tc; same inputs, outputs and NaN / clip semantics).
// Everything per tile is warp-local: a warp owns 32 of the tile's 128 rows from the raw load to the store of their
// encodings -- it normalises them (a lane quad per row), keeps them row-major in its own slice of shared memory for
// the A fragments of layer 1, writes the ReLU'd hidden rows to another slice (the C fragment layout of an mma is not
// the A layout of the next) and runs layer 2 from there; only __syncwarp between the phases.  The weights are split
// into tf32 head / tail once per CTA and stay in shared memory k-major (stride 56: conflict-free B fragment loads).
constexpr int kEncXs = 20, kEncHs = 52;  // row strides (floats): g * stride + t hits 32 distinct banks for g < 8, t < 4
constexpr int kEncTcThreads = 256, kEncTcWarps = kEncTcThreads / 32, kEncTcTile = kEncTcWarps * 32;
struct EncTcSmem {  // 102 KB: two CTAs = 16 warps per SM
    float w1h[kPolD][kLinWStride], w1l[kPolD][kLinWStride];
    float w2h[kPolH][kLinWStride], w2l[kPolH][kLinWStride];
    float b1[kPolH], b2[kPolH];
    float xs[kEncTcWarps][32][kEncXs];  // per warp: its 32 rows of the tile in work
    float hs[kEncTcWarps][32][kEncHs];
};

__global__ void __launch_bounds__(kEncTcThreads, 2) k_policy_encode_tc(const float4 *__restrict__ x, const float *__restrict__ mean,
                                                                    const float *__restrict__ var,
                                                                    const float *__restrict__ w1, const float *__restrict__ b1,
                                                                    const float *__restrict__ w2, const float *__restrict__ b2,
                                                                    size_t rows, float4 *__restrict__ x_norm,
                                                                    uint8_t *__restrict__ row_nan, float *__restrict__ v) {
    extern __shared__ __align__(16) unsigned char enc_tc_raw[];
    EncTcSmem &s = *reinterpret_cast<EncTcSmem *>(enc_tc_raw);
    const unsigned full = 0xffffffffu;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, quad = lane & 3, g = lane >> 2, t = lane & 3;
    for (int i = tid; i < kPolD * kPolH; i += kEncTcThreads) {  // w1[j][k] -> w1h / w1l[k][j]
        uint32_t hi, lo;
        tf32_split(w1[i], hi, lo);
        s.w1h[i % kPolD][i / kPolD] = __uint_as_float(hi); s.w1l[i % kPolD][i / kPolD] = __uint_as_float(lo);
    }
    for (int i = tid; i < kPolH * kPolH; i += kEncTcThreads) {
        uint32_t hi, lo;
        tf32_split(w2[i], hi, lo);
        s.w2h[i % kPolH][i / kPolH] = __uint_as_float(hi); s.w2l[i % kPolH][i / kPolH] = __uint_as_float(lo);
    }
    if (tid < kPolH) { s.b1[tid] = b1[tid]; s.b2[tid] = b2[tid]; }
    const float4 m = ((const float4 *)mean)[quad], vv = ((const float4 *)var)[quad];
    const float sd[4] = {__fsqrt_rn(__fadd_rn(vv.x, 1e-8f)), __fsqrt_rn(__fadd_rn(vv.y, 1e-8f)),
                         __fsqrt_rn(__fadd_rn(vv.z, 1e-8f)), __fsqrt_rn(__fadd_rn(vv.w, 1e-8f))};
    const float mu[4] = {m.x, m.y, m.z, m.w};
    const size_t tiles = (rows + kEncTcTile - 1) / kEncTcTile;
    float (*xs)[kEncXs] = s.xs[warp];
    float (*hs)[kEncHs] = s.hs[warp];
    // slot q of this lane: row (warp's first row) + g + 8 q, float4 `quad` of it; the next tile's slots travel in registers
    float4 nxt[4];
    auto fetch = [&](size_t tl) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const size_t row = tl * kEncTcTile + warp * 32 + g + 8 * q;
            nxt[q] = (tl < tiles && row < rows) ? x[row * 4 + quad] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    };
    fetch(blockIdx.x);
    __syncthreads();  // weights in place
    for (size_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const size_t r0 = tile * kEncTcTile + warp * 32;  // this warp's first row
        float4 raw[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) raw[q] = nxt[q];
        fetch(tile + gridDim.x);
        // ---- normalise this warp's 32 rows (4 per lane quad); the values stay in raw[]
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const size_t row = r0 + g + 8 * q;
            float e[4] = {raw[q].x, raw[q].y, raw[q].z, raw[q].w};
            const bool in = row < rows;
            bool bad = false;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (e[k] != e[k]) { bad = true; e[k] = 0.f; }
                else e[k] = fminf(fmaxf(__fdiv_rn(__fsub_rn(e[k], mu[k]), sd[k]), -10.f), 10.f);
            }
            const unsigned bb = __ballot_sync(full, bad);
            raw[q] = make_float4(e[0], e[1], e[2], e[3]);
            if (in) {
                x_norm[row * 4 + quad] = raw[q];
                if (quad == 0) row_nan[row] = ((bb >> lane) & 0xfu) != 0;
            }
        }
        __syncwarp();  // the previous tile's xs / hs have been consumed
#pragma unroll
        for (int q = 0; q < 4; ++q) *(float4 *)&xs[g + 8 * q][quad * 4] = raw[q];
        __syncwarp();
        // Both 16-row halves go through a k-step together, so a B fragment (weights) is read from shared memory once
        // per 32 rows: with one half at a time the kernel sat at 73 % of the shared-memory throughput.
        float acc[2][6][4];
        // ---- layer 1: h = relu(x w1^T + b1), 32 rows x 48 columns, K = 16
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < 6; ++nt) {
                acc[mt][nt][0] = acc[mt][nt][2] = s.b1[nt * 8 + 2 * t];
                acc[mt][nt][1] = acc[mt][nt][3] = s.b1[nt * 8 + 2 * t + 1];
            }
#pragma unroll
        for (int ks = 0; ks < kPolD; ks += 8) {
            uint32_t bh[6][2], bl[6][2];
#pragma unroll
            for (int nt = 0; nt < 6; ++nt) {
                bh[nt][0] = __float_as_uint(s.w1h[ks + t][nt * 8 + g]); bh[nt][1] = __float_as_uint(s.w1h[ks + t + 4][nt * 8 + g]);
                bl[nt][0] = __float_as_uint(s.w1l[ks + t][nt * 8 + g]); bl[nt][1] = __float_as_uint(s.w1l[ks + t + 4][nt * 8 + g]);
            }
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) {
                const int row = mt * 16 + g;
                uint32_t ah[4], al[4];
                tf32_split(xs[row][ks + t], ah[0], al[0]);
                tf32_split(xs[row + 8][ks + t], ah[1], al[1]);
                tf32_split(xs[row][ks + t + 4], ah[2], al[2]);
                tf32_split(xs[row + 8][ks + t + 4], ah[3], al[3]);
#pragma unroll
                for (int nt = 0; nt < 6; ++nt) {
                    mma_tf32(acc[mt][nt], al, bh[nt]);
                    mma_tf32(acc[mt][nt], ah, bl[nt]);
                    mma_tf32(acc[mt][nt], ah, bh[nt]);
                }
            }
        }
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < 6; ++nt) {
                *(float2 *)&hs[mt * 16 + g][nt * 8 + 2 * t] = make_float2(fmaxf(acc[mt][nt][0], 0.f), fmaxf(acc[mt][nt][1], 0.f));
                *(float2 *)&hs[mt * 16 + g + 8][nt * 8 + 2 * t] = make_float2(fmaxf(acc[mt][nt][2], 0.f), fmaxf(acc[mt][nt][3], 0.f));
            }
        __syncwarp();
        // ---- layer 2: v = h w2^T + b2, K = 48
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int nt = 0; nt < 6; ++nt) {
                acc[mt][nt][0] = acc[mt][nt][2] = s.b2[nt * 8 + 2 * t];
                acc[mt][nt][1] = acc[mt][nt][3] = s.b2[nt * 8 + 2 * t + 1];
            }
#pragma unroll 1
        for (int ks = 0; ks < kPolH; ks += 8) {
            uint32_t bh[6][2], bl[6][2];
#pragma unroll
            for (int nt = 0; nt < 6; ++nt) {
                bh[nt][0] = __float_as_uint(s.w2h[ks + t][nt * 8 + g]); bh[nt][1] = __float_as_uint(s.w2h[ks + t + 4][nt * 8 + g]);
                bl[nt][0] = __float_as_uint(s.w2l[ks + t][nt * 8 + g]); bl[nt][1] = __float_as_uint(s.w2l[ks + t + 4][nt * 8 + g]);
            }
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) {
                const int row = mt * 16 + g;
                uint32_t ah[4], al[4];
                tf32_split(hs[row][ks + t], ah[0], al[0]);
                tf32_split(hs[row + 8][ks + t], ah[1], al[1]);
                tf32_split(hs[row][ks + t + 4], ah[2], al[2]);
                tf32_split(hs[row + 8][ks + t + 4], ah[3], al[3]);
#pragma unroll
                for (int nt = 0; nt < 6; ++nt) {
                    mma_tf32(acc[mt][nt], al, bh[nt]);
                    mma_tf32(acc[mt][nt], ah, bl[nt]);
                    mma_tf32(acc[mt][nt], ah, bh[nt]);
                }
            }
        }
#pragma unroll
        for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const size_t grow = r0 + mt * 16 + g + 8 * h;
                if (grow < rows) {
#pragma unroll
                    for (int nt = 0; nt < 6; ++nt)
                        *(float2 *)(v + grow * kPolH + nt * 8 + 2 * t) = make_float2(acc[mt][nt][2 * h], acc[mt][nt][2 * h + 1]);
                }
            }
    }
}

}  // namespace dca
