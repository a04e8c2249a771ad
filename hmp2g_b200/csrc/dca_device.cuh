// dca_device.cuh -- device-side building blocks of the DCA hot path (sm_100a).
//
// Data layout in HBM: one 128-byte aligned "env blob" per env holding the complete entity state as
// structure-of-arrays (field arrays of length N inside the blob) followed by the env scalars.  A CTA owns one
// env: the blob is fetched into shared memory with ONE 1-D TMA bulk copy (cp.async.bulk + mbarrier), every
// phase of the step works on shared memory, and the blob goes back with one bulk store.  The observation
// tile [G,12,16] f32 is written straight from the shared entity table with 128-bit coalesced stores.
//
// Arithmetic follows the reference rounding for rounding (compile with -fmad=false): fp64 state, fp32
// entity table, fp32 distance = sqrt_rn(fl(dx*dx)+fl(dy*dy)) (collective_assault_env.py:276,286,361-368).
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#include "../../include/dca_b200.h"

namespace dca {

constexpr int kVisN = 6;
constexpr int kObsRows = DCA_OBS_ROWS;
constexpr int kObsRow = DCA_OBS_ROW;
constexpr int kObsLen = kObsRows * kObsRow;  // 192 f32 per guard
constexpr int kMaxN = 320;                   // reference allows up to 150 + 150 (parallel_run.py:37-38)
constexpr double kPi = 3.141592653589793;    // np.pi

// flag bits of the per-agent state byte
enum : uint8_t { F_ALIVE = 1, F_JUST_DIED = 2, F_HIT = 4, F_WAS_HIT = 8 };

__host__ __device__ constexpr int align_up(int x, int a) { return (x + a - 1) / a * a; }

// Byte offsets of the fields inside one env blob.
struct BlobLayout {
    int N, G, A;
    int px, py, vx, vy, atk, terrain;  // f64[N]
    int num_hit, num_was_hit;          // i32[N]
    int flags;                         // u8[N]
    int knn_ids;                       // u64[2G] perf hint: the 7 nearest of the last pass per (guard, list), one byte each
                                       //         (marker = list length); bounds the next scan radius, never affects results
    int hmin;                          // f32[G]  lower bound of the f32 squared distance guard -> nearest alive attacker
    int scalars;                       // EnvScalars
    int bytes;                         // multiple of 128
};

struct EnvScalars {
    double init_theta;
    int32_t time_step, n_alive_guards, n_alive_attackers;
    uint32_t obs_counter, episode;
    uint32_t pad;
    double cos_theta, sin_theta;  // cos / sin of init_theta (derived; refreshed by reset and state import)
};
static_assert(sizeof(EnvScalars) == 48, "EnvScalars must stay 48 bytes");

// list arrays (float32 positions per team) are padded with +inf to a multiple of 4 entries and always hold at least
// one padding slot: index `cnt` is the "no entry" marker of the kNN hints
__host__ __device__ constexpr int list_pad(int cnt) { return (cnt + 1 + 3) / 4 * 4; }

__host__ __device__ constexpr BlobLayout make_layout(int G, int A) {
    BlobLayout L{};
    L.G = G; L.A = A; L.N = G + A;
    int o = 0;
    const int f64 = align_up(L.N * 8, 16), i32 = align_up(L.N * 4, 16), u8 = align_up(L.N, 16);
    L.px = o; o += f64; L.py = o; o += f64; L.vx = o; o += f64; L.vy = o; o += f64;
    L.atk = o; o += f64; L.terrain = o; o += f64;
    L.num_hit = o; o += i32; L.num_was_hit = o; o += i32;
    L.flags = o; o += u8;
    L.knn_ids = o; o += align_up(2 * L.G * 8, 16);
    L.hmin = o; o += align_up(L.G * 4, 16);
    L.scalars = o; o += (int)sizeof(EnvScalars);
    L.bytes = align_up(o, 128);
    return L;
}

// Kernel parameters shared by every kernel (passed by value: lands in constant bank).
struct Params {
    DcaConfig cfg;
    BlobLayout L;
    int E;
    unsigned char *blobs;  // [E, L.bytes]
    DcaBuffers out;
    unsigned long long *prof;  // optional [E,16] phase clocks (diagnostics), nullptr = off
    float smax[2];     // [hostile, friend]: largest f32 s with sqrt_rn(s) <= obs range (2.0 / 3.0)
    int idx_bits;      // low bits of a kNN / aim key that hold the candidate index
    unsigned long long jam_thresh;  // a jam draw w (u32) removes an item iff w < jam_thresh  (== w * 2^-32 < random_jam_prob)
    double cone_cos2;  // cos^2(agent_shoot_win / 2) when the laser window test can be done as a cone test, else 0
    int empty_nan;     // DCA_OPT_EMPTY_ROWS_NAN: invisible / jammed / empty obs rows (and dead guards' obs) are NaN, not 0
};

// ------------------------------------------------------------------ Philox4x32-10 (keyed RNG stream)
enum : uint32_t { SITE_OBS = 1, SITE_RESET = 2 };

__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                               uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
}

// ------------------------------------------------------------------ exact-semantics math helpers
// Python / numpy floored float modulo (CPython float_rem, numpy npy_divmod).
__device__ __forceinline__ double py_fmod(double a, double b) {
    // fmod is exact; for |a| < 2|b| it is a or a -/+ b, and that subtraction is exact (Sterbenz) -- skip the slow loop
    const double fa = fabs(a), fb = fabs(b);
    double m = fa < fb ? a : (fa < 2 * fb ? copysign(fa - fb, a) : fmod(a, b));
    if (m != 0.0) {
        if ((b < 0) != (m < 0)) m += b;
    } else {
        m = copysign(0.0, b);
    }
    return m;
}
// reg_angle core.py:284-285 == reg_rad cython_func.pyx:16-17
__device__ __forceinline__ double reg_angle(double rad) { return py_fmod(rad + kPi, 2 * kPi) - kPi; }

// np.logaddexp(0, y)
__device__ __forceinline__ double logaddexp0(double y) {
    // exp(-800) is 0 in double, so log1p(exp(-|y|)) is exactly 0 beyond it: the result is max(0, y) bit for bit
    if (y < -800.0) return 0.0;
    if (y > 800.0) return y;
    if (y == 0.0) return 0.693147180559945309417232121458176568;
    const double tmp = 0.0 - y;
    if (tmp > 0) return 0.0 + log1p(exp(-tmp));
    else if (tmp <= 0) return y + log1p(exp(tmp));
    return tmp;
}

// float32 distance exactly as distance_matrix() evaluates it (collective_assault_env.py:361-368)
__device__ __forceinline__ float f32_dist(float xi, float yi, float xj, float yj) {
    const float dx = __fsub_rn(xj, xi), dy = __fsub_rn(yj, yi);
    return __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
}

// ------------------------------------------------------------------ packed f32x2 arithmetic (sm_100 FADD2 / FMUL2 / FFMA2)
// Two candidates per instruction.  ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2, so the packed squared
// distance below may differ from the reference's fl(fl(dx*dx)+fl(dy*dy)) by one rounding: it is used ONLY for
// conservative pre-filters (thresholds carry a margin); every value that decides an output is recomputed with the
// scalar __fmul_rn/__fadd_rn sequence.
typedef unsigned long long u64;
__device__ __forceinline__ u64 pack2(float lo, float hi) {
    return (u64)__float_as_uint(lo) | ((u64)__float_as_uint(hi) << 32);
}
__device__ __forceinline__ float lo2(u64 v) { return __uint_as_float((uint32_t)v); }
__device__ __forceinline__ float hi2(u64 v) { return __uint_as_float((uint32_t)(v >> 32)); }
__device__ __forceinline__ u64 sub2(u64 a, u64 b) {
    u64 r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ u64 mul2(u64 a, u64 b) {
    u64 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
    u64 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
// approximate squared distances of two candidates (x0,x1),(y0,y1) to (px,px),(py,py): within 2 ulp of the exact one
__device__ __forceinline__ u64 sqdist2(u64 xs, u64 ys, u64 px2, u64 py2) {
    const u64 dx = sub2(xs, px2), dy = sub2(ys, py2);
    return fma2(dx, dx, mul2(dy, dy));
}
// exact float32 squared distance, same roundings as distance_matrix() before the sqrt (collective_assault_env.py:361-368)
__device__ __forceinline__ float f32_sqdist(float xi, float yi, float xj, float yj) {
    const float dx = __fsub_rn(xj, xi), dy = __fsub_rn(yj, yi);
    return __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
}

// L2 residency hints: the observation tile is written once per step and never read back by these kernels, while the
// env blobs (28 MB at 4096 x 50v50) are re-read by the next launch -- so obs stores carry L2::evict_first and the blob
// bulk copies an evict_last policy, and the blobs stay in the 126 MB L2 under the 157 MB/step obs stream.
#ifndef DCA_L2_HINTS
#define DCA_L2_HINTS 1
#endif

// 256-bit global store (sm_100: STG.E.{ENL2,EFL2}.256); p must be 32-byte aligned
__device__ __forceinline__ void st_global_256(float *p, float4 a, float4 b) {
#if DCA_L2_HINTS
    asm volatile("st.global.L2::evict_first.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(a.x), "f"(a.y), "f"(a.z),
                 "f"(a.w), "f"(b.x), "f"(b.y), "f"(b.z), "f"(b.w)
                 : "memory");
#else
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w),
                 "f"(b.x), "f"(b.y), "f"(b.z), "f"(b.w)
                 : "memory");
#endif
}

// ------------------------------------------------------------------ programmatic dependent launch (PDL)
// pdl_trigger: the next kernel of the stream, if launched with the programmatic-serialization attribute, may become
// resident from here on.  pdl_wait: returns once the preceding kernel has completed and its writes are visible
// (no-op for a normal launch).
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ------------------------------------------------------------------ per-env hand-over between launches
// DCA_OPT_PIPELINE_STEPS: consecutive launches on one handle (step -> planner -> step ...) only depend on each other
// env by env, so instead of the grid-wide griddepcontrol.wait the consumer CTA of env e waits for the epoch its
// producer CTA published for env e: the next launch's first CTAs run in the slots the previous launch's last wave leaves
// empty.  epoch words: [0,E) last dca_step launch that finished env e, [E,2E) last planner launch that filled e's forces.
struct Pipe {
    uint32_t *epoch;     // [2E] device words (nullptr: feature off)
    uint32_t wait_step;  // consumer: wait until epoch[e] has reached this value (0 = do not wait)
    uint32_t wait_plan;  // k_step only: wait until epoch[E + e] has reached this value before the forces are read
    uint32_t publish;    // producer: value stored with release semantics when env e is done
};
__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u32(uint32_t *p, uint32_t v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// The producer is resident or finished whenever a consumer can run (the dependent grid is only scheduled once every
// CTA of its predecessor has started), so the wait is short; a bounded spin turns a protocol bug into a trap, not a hang.
__device__ __forceinline__ void pipe_wait(const uint32_t *p, uint32_t want) {
    unsigned spins = 0;
    while ((int32_t)(ld_acquire_u32(p) - want) < 0) {
        __nanosleep(64);
        if (++spins > (1u << 24)) __trap();
    }
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

// ------------------------------------------------------------------ TMA bulk copy (1-D) + mbarrier
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t phase) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(phase)
            : "memory");
    } while (!ok);
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
// global -> shared, completion signalled on the mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void tma_load_1d(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar) {
#if DCA_L2_HINTS
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
            smem_u32(smem_dst)),
        "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(l2_policy_evict_last())
        : "memory");
#else
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
#endif
}
// shared -> global bulk store
__device__ __forceinline__ void tma_store_1d(void *gdst, const void *smem_src, uint32_t bytes) {
#if DCA_L2_HINTS
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(gdst),
                 "r"(smem_u32(smem_src)), "r"(bytes), "l"(l2_policy_evict_last())
                 : "memory");
#else
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)),
                 "r"(bytes)
                 : "memory");
#endif
}
__device__ __forceinline__ void tma_store_commit_wait() {
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_written() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// ... until the bulk store has been written (not just read out of shared memory): needed before the env is handed over
__device__ __forceinline__ void tma_store_commit_wait_written() {
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

}  // namespace dca
