// conc_4hist history pool on the device: ShellEnvWrapper.interact_with_env's handling of
// '_Histpool_Obs_' + solve_duplicate (ALGORITHM/conc_4hist/shell_env.py:25-99) + roll_hisory
// (ALGORITHM/conc_4hist/cython_func.pyx:32-53), for every (env, agent) in one pass over HBM.
//
// Per (env e, agent a), rows are core_dim = 16 floats (64 B), n = 12 rows per observation:
//   prev      = 0 if Env-Suffered-Reset[e] else his_pool[e,a]                       (shell_env.py:37)
//   row valid = not (all 16 == 0) and no NaN; row 0 (self) is never buffered        (:78-83)
//   next_his  = [valid new rows in order] ++ prev[0 : n - n_valid]                   (cython_func.pyx:44-52)
//   obs_feed  = new ++ prev  (2n rows) with all-zero rows -> NaN; every row NaN if the self row is empty/NaN (:75-99)
//   paused env (ENV-PAUSE): his_pool[e] = 0 (:44), its obs_feed rows are written as NaN (the reference does not feed
//   paused envs to the policy at all)
// Pure data movement: 2 x 768 B read, 1536 + 768 B written per agent -> HBM-bound.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include <math_constants.h>

namespace dca {

constexpr int kHistRows = 12;               // DCA_OBS_ROWS
constexpr int kHistQuads = kHistRows * 4;   // float4s per observation (core_dim 16)

__device__ __forceinline__ bool quad_zero(const float4 v) { return v.x == 0.f && v.y == 0.f && v.z == 0.f && v.w == 0.f; }
__device__ __forceinline__ bool quad_nan(const float4 v) { return v.x != v.x || v.y != v.y || v.z != v.z || v.w != v.w; }

// 12-bit row masks from per-quad predicates: lanes hold quads q0 = lane (rows 0..7) and q1 = 32 + lane (rows 8..11)
__device__ __forceinline__ uint32_t rows_all(bool p0, bool p1) {
    const uint32_t b0 = __ballot_sync(0xffffffffu, p0), b1 = __ballot_sync(0xffffffffu, p1) & 0xffffu;
    uint32_t m = 0;
#pragma unroll
    for (int r = 0; r < 8; ++r) m |= (((b0 >> (4 * r)) & 0xfu) == 0xfu) ? (1u << r) : 0u;
#pragma unroll
    for (int r = 0; r < 4; ++r) m |= (((b1 >> (4 * r)) & 0xfu) == 0xfu) ? (1u << (8 + r)) : 0u;
    return m;
}
__device__ __forceinline__ uint32_t rows_any(bool p0, bool p1) {
    const uint32_t b0 = __ballot_sync(0xffffffffu, p0), b1 = __ballot_sync(0xffffffffu, p1) & 0xffffu;
    uint32_t m = 0;
#pragma unroll
    for (int r = 0; r < 8; ++r) m |= ((b0 >> (4 * r)) & 0xfu) ? (1u << r) : 0u;
#pragma unroll
    for (int r = 0; r < 4; ++r) m |= ((b1 >> (4 * r)) & 0xfu) ? (1u << (8 + r)) : 0u;
    return m;
}

// one warp per (env, agent)
__global__ void __launch_bounds__(256) k_histpool(const float4 *__restrict__ obs, float4 *__restrict__ his_pool,
                                                  const uint8_t *__restrict__ pause,
                                                  const uint8_t *__restrict__ suffered, float4 *__restrict__ feed,
                                                  int E, int G) {
    const int lane = threadIdx.x & 31;
    const size_t w = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (w >= (size_t)E * G) return;  // warp-uniform
    const int e = (int)(w / G);
    const float4 *o = obs + w * kHistQuads;
    float4 *h = his_pool + w * kHistQuads;
    float4 *f = feed + w * 2 * kHistQuads;
    const bool has1 = lane < kHistQuads - 32;  // lanes 0..15 hold a second quad
    const float4 qnan = make_float4(CUDART_NAN_F, CUDART_NAN_F, CUDART_NAN_F, CUDART_NAN_F);
    const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
    if (pause && pause[e]) {
        h[lane] = zero;
        if (has1) h[32 + lane] = zero;
        f[lane] = qnan; f[32 + lane] = qnan; f[64 + lane] = qnan;
        return;
    }
    const bool fresh = suffered && suffered[e] != 0;
    const float4 n0 = o[lane], n1 = has1 ? o[32 + lane] : zero;
    const float4 p0 = fresh ? zero : h[lane], p1 = (fresh || !has1) ? zero : h[32 + lane];
    const uint32_t new_zero = rows_all(quad_zero(n0), has1 && quad_zero(n1));
    const uint32_t new_nan = rows_any(quad_nan(n0), has1 && quad_nan(n1));
    const uint32_t prev_zero = rows_all(quad_zero(p0), has1 && quad_zero(p1));
    const uint32_t empty = new_zero | new_nan;              // rows that are (or become) NaN rows
    const uint32_t valid = ~empty & 0xffeu;                 // buffered rows: not the self row
    const int n_valid = __popc(valid);
    const bool dead = empty & 1u;                           // self row empty <=> the agent is dead (:94-97)
    // ---- obs_feed: new rows, then the history rows
    const int r0 = lane >> 2, r1 = 8 + (lane >> 2);
    f[lane] = (dead || ((new_zero >> r0) & 1u)) ? qnan : n0;
    if (has1) f[32 + lane] = (dead || ((new_zero >> r1) & 1u)) ? qnan : n1;
    f[kHistQuads + lane] = (dead || ((prev_zero >> r0) & 1u)) ? qnan : p0;
    if (has1) f[kHistQuads + 32 + lane] = (dead || ((prev_zero >> r1) & 1u)) ? qnan : p1;
    // ---- next history: valid new rows move to the front, the old rows shift back by n_valid
    const int part = lane & 3;
    if ((valid >> r0) & 1u) h[4 * __popc(valid & ((1u << r0) - 1u)) + part] = n0;
    if (has1 && ((valid >> r1) & 1u)) h[4 * __popc(valid & ((1u << r1) - 1u)) + part] = n1;
    if (r0 + n_valid < kHistRows) h[4 * (r0 + n_valid) + part] = p0;
    if (has1 && r1 + n_valid < kHistRows) h[4 * (r1 + n_valid) + part] = p1;
}

}  // namespace dca
