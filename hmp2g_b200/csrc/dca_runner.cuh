// Runner bookkeeping on the device: what task_runner.Runner.update_runner does per step with an O(E)
// Python loop (task_runner.py:95-138; the test-run twin :191-217), for E envs at once.  The pause masks it produces are
// the ENV-PAUSE rows that MMPlatform.act turns into NaN actions (multi_team.py:62-66) and that dca_step takes as `skip`.
//
// Per env (reference order): running envs (R = ~ENV-PAUSE) add the step's reward to Latest-Reward-Sum and count the
// step; Env-Suffered-Reset = done; a running env that finished logs one episode, in env-index order: the episode gets
// the next global episode number, (env, win, steps, number) and its reward sums go to the episode log, its sums / step
// counter are zeroed, Thread-Episode-Cnt += 1, and with align_episode the env pauses until every env has paused.
// The per-window reports (_checkout_interested_agents, task_runner.py:231-287) read the log by episode number, so the
// windows fall exactly where the reference's in-loop `current_n_episode % report_interval == 0` test puts them.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/dca_b200.h"

namespace dca {

constexpr int kRunnerThreads = 1024;

// One CTA walks the envs in index order (the episode numbering is a prefix sum over them).
__global__ void __launch_bounds__(kRunnerThreads) k_runner_scan(DcaRunnerState S, int E, int align,
                                                                 const uint8_t *__restrict__ done,
                                                                 const uint8_t *__restrict__ win) {
    __shared__ int wcnt[kRunnerThreads / 32];
    __shared__ long long wframes[kRunnerThreads / 32];
    __shared__ int any_running;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long n_ep0 = S.counters[1];
    if (tid == 0) any_running = 0;
    int base = 0;
    long long frames = 0;
    bool running_after = false;
    for (int e0 = 0; e0 < E; e0 += kRunnerThreads) {  // uniform
        const int e = e0 + tid;
        const bool in = e < E;
        const bool paused = in && S.pause[e] != 0, run = in && !paused;
        const bool d = in && done[e] != 0;
        const bool fin = run && d;
        const int step = run ? S.cur_step[e] + 1 : 0;
        const uint32_t b = __ballot_sync(0xffffffffu, fin);
        __syncthreads();  // wcnt of the previous round has been read
        if (lane == 0) wcnt[warp] = __popc(b);
        __syncthreads();
        int off = base, tot = 0;
        for (int w = 0; w < kRunnerThreads / 32; ++w) {
            off += w < warp ? wcnt[w] : 0;
            tot += wcnt[w];
        }
        if (in) S.suffered[e] = d ? 1 : 0;
        if (fin) {
            const long long n = n_ep0 + off + __popc(b & ((1u << lane) - 1u)) + 1;  // this episode's global number
            const int slot = (int)((n - 1) % S.log_capacity);
            int64_t *m = S.log_meta + 4 * (size_t)slot;
            m[0] = e; m[1] = win[e] != 0; m[2] = step; m[3] = n;
            S.fin_slot[e] = slot + 1;
            frames += step;
            S.cur_step[e] = 0;
            S.ep_cnt[e] += 1;
            if (align) S.pause[e] = 1;
        } else if (in) {
            S.fin_slot[e] = run ? 0 : -1;
            if (run) S.cur_step[e] = step;
        }
        running_after |= in && !(paused || (fin && align));
        base += tot;
    }
    for (int o = 16; o; o >>= 1) frames += __shfl_xor_sync(0xffffffffu, frames, o);
    if (lane == 0) wframes[warp] = frames;
    if (running_after) any_running = 1;
    __syncthreads();
    if (align && !any_running)  // every env has paused: release them all (task_runner.py:135)
        for (int e = tid; e < E; e += kRunnerThreads) S.pause[e] = 0;
    if (tid == 0) {
        long long f = 0;
        for (int w = 0; w < kRunnerThreads / 32; ++w) f += wframes[w];
        S.counters[0] += f;      // current_n_frame
        S.counters[1] += base;   // current_n_episode
        S.counters[2] += (align && !any_running) ? 1 : 0;  // aligned restarts
    }
}

// Latest-Reward-Sum[R] += reward[R]; finished envs move their sums into the episode log and start from 0.
__global__ void k_runner_rewards(DcaRunnerState S, int E, int G, const float *__restrict__ reward) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)E * G) return;
    const int e = (int)(idx / G), g = (int)(idx - (size_t)e * G);
    const int f = S.fin_slot[e];
    if (f < 0) return;  // paused
    double r = S.reward_sum[idx] + (double)reward[idx];
    if (f > 0) {
        S.log_reward[(size_t)(f - 1) * G + g] = r;
        r = 0.0;
    }
    S.reward_sum[idx] = r;
}

// One report window (_checkout_interested_agents, task_runner.py:231-275): episodes first .. first+count-1 of the log.
// out[t] = mean over those episodes and team t's agents of the episode reward sums, t < n_team; out[n_team] = win rate.
__global__ void __launch_bounds__(256) k_runner_report(DcaRunnerState S, int G, long long first, int count,
                                                       const int32_t *__restrict__ team_of_agent, int n_team,
                                                       double *__restrict__ out) {
    __shared__ double acc[256];
    __shared__ int cnt[256];
    const int tid = threadIdx.x;
    for (int t = 0; t <= n_team; ++t) {
        double a = 0.0;
        int c = 0;
        if (t < n_team) {
            for (int k = tid; k < count * G; k += 256) {
                const int ep = k / G, g = k - ep * G;
                if (team_of_agent[g] != t) continue;
                const int slot = (int)((first - 1 + ep) % S.log_capacity);
                a += S.log_reward[(size_t)slot * G + g];
                ++c;
            }
        } else {
            for (int ep = tid; ep < count; ep += 256) {
                const int slot = (int)((first - 1 + ep) % S.log_capacity);
                a += (double)S.log_meta[4 * (size_t)slot + 1];
                ++c;
            }
        }
        acc[tid] = a;
        cnt[tid] = c;
        __syncthreads();
        for (int o = 128; o; o >>= 1) {
            if (tid < o) { acc[tid] += acc[tid + o]; cnt[tid] += cnt[tid + o]; }
            __syncthreads();
        }
        if (tid == 0) out[t] = acc[0] / (double)cnt[0];
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Advantage normalisation over the pooled batch (ALGORITHM/conc/ppo.py:53-55, ALGORITHM/common/ppo_sampler.py:67-77):
// adv = ret - value; adv = (adv - adv.mean()) / (adv.std() + 1e-5), numpy std = population std (ddof 0).
// Moments are accumulated in fp64 in a fixed order (per-CTA partials, then one CTA): bit-reproducible run to run.
// stats = (sum adv, sum adv^2, count); across GPUs the caller all-reduces the three doubles before normalising.
constexpr int kAdvBlocks = 592, kAdvThreads = 256;  // 4 CTAs per SM on 148 SMs

__device__ __forceinline__ void adv_block_reduce(double (&v)[3], double *sh) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#pragma unroll
    for (int k = 0; k < 3; ++k)
        for (int o = 16; o; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
    if (lane == 0) { sh[warp * 3] = v[0]; sh[warp * 3 + 1] = v[1]; sh[warp * 3 + 2] = v[2]; }
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < kAdvThreads / 32; ++w) { v[0] += sh[w * 3]; v[1] += sh[w * 3 + 1]; v[2] += sh[w * 3 + 2]; }
    }
}

__global__ void __launch_bounds__(kAdvThreads) k_adv_partial(const float *__restrict__ ret, const float *__restrict__ value,
                                                             const uint8_t *__restrict__ valid, int valid_group, size_t n,
                                                             double *__restrict__ partial) {
    __shared__ double sh[3 * kAdvThreads / 32];
    double v[3] = {0.0, 0.0, 0.0};
    for (size_t i = (size_t)blockIdx.x * kAdvThreads + threadIdx.x; i < n; i += (size_t)kAdvBlocks * kAdvThreads) {
        if (valid && !valid[i / valid_group]) continue;
        const double a = (double)__fsub_rn(ret[i], value[i]);  // the f32 difference the reference forms first
        v[0] += a; v[1] += a * a; v[2] += 1.0;
    }
    adv_block_reduce(v, sh);
    if (threadIdx.x == 0) { partial[blockIdx.x * 3] = v[0]; partial[blockIdx.x * 3 + 1] = v[1]; partial[blockIdx.x * 3 + 2] = v[2]; }
}

__global__ void __launch_bounds__(kAdvThreads) k_adv_final(const double *__restrict__ partial, double *__restrict__ stats) {
    __shared__ double sh[3 * kAdvThreads / 32];
    double v[3] = {0.0, 0.0, 0.0};
    for (int b = threadIdx.x; b < kAdvBlocks; b += kAdvThreads) { v[0] += partial[b * 3]; v[1] += partial[b * 3 + 1]; v[2] += partial[b * 3 + 2]; }
    adv_block_reduce(v, sh);
    if (threadIdx.x == 0) { stats[0] = v[0]; stats[1] = v[1]; stats[2] = v[2]; }
}

__global__ void __launch_bounds__(kAdvThreads) k_adv_normalize(const float *__restrict__ ret, const float *__restrict__ value,
                                                               const uint8_t *__restrict__ valid, int valid_group, size_t n,
                                                               const double *__restrict__ stats, float *__restrict__ adv) {
    const double cnt = stats[2] > 0.0 ? stats[2] : 1.0;
    const double mean = stats[0] / cnt;
    const double var = fmax(stats[1] / cnt - mean * mean, 0.0);
    const float m = (float)mean, sd = __fadd_rn((float)sqrt(var), 1e-5f);
    for (size_t i = (size_t)blockIdx.x * kAdvThreads + threadIdx.x; i < n; i += (size_t)gridDim.x * kAdvThreads) {
        const bool ok = !valid || valid[i / valid_group];
        adv[i] = ok ? __fdiv_rn(__fsub_rn(__fsub_rn(ret[i], value[i]), m), sd) : 0.f;
    }
}

}  // namespace dca
