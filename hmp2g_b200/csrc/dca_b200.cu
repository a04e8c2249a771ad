// dca_b200.cu -- C-ABI (include/dca_b200.h) over the kernels in dca_kernels.cuh.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -fmad=false -shared -Xcompiler -fPIC
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <string>

#include "dca_kernels.cuh"
#include "dca_planner.cuh"
#include "dca_runner.cuh"
#include "dca_histpool.cuh"
#include "dca_multiteam.cuh"
#include "dca_policy.cuh"

using namespace dca;

struct DcaHandle {
    Params P;
    int device;
    int threads;
    int smem;
    bool bound;
    int64_t launches;
    PlanState plan;
    int plan_smem;
    bool overlap_planner;           // DCA_OPT_OVERLAP_PLANNER
    const void *planned_force;      // att_force / stream of the planner launch, if that was the handle's last launch
    cudaStream_t planned_stream;
    bool pipeline_steps;            // DCA_OPT_PIPELINE_STEPS
    uint32_t *epoch;                // [2E] per-env hand-over words: step epochs, planner epochs
    uint32_t step_epoch, plan_epoch;  // launches published so far
    bool last_was_step;             // the handle's last launch was a dca_step without a pause mask ...
    cudaStream_t last_step_stream;  // ... on this stream
    std::string err;
};

struct DcaMtHandle {
    dca::MtParams P;
    int device;
    int smem;
    bool bound;
    void *slab;  // one allocation holding every state array
};

static thread_local std::string g_err;

static int fail(DcaHandle *h, int code, const std::string &msg) {
    if (h) h->err = msg;
    g_err = msg;
    return code;
}

#define CUDA_TRY(h, expr)                                                                      \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess)                                                                 \
            return fail(h, DCA_E_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));     \
    } while (0)

namespace {
struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
        else prev = -1;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};
// The handle-less entry points (dca_gae, dca_runner_*, dca_histpool_roll, dca_adv_*) launch on the device that owns
// their buffers, whatever device is current in the calling thread.
struct PtrDeviceGuard {
    int prev = -1;
    explicit PtrDeviceGuard(const void *p) {
        cudaPointerAttributes a;
        if (p && cudaPointerGetAttributes(&a, p) == cudaSuccess && a.type == cudaMemoryTypeDevice) {
            cudaGetDevice(&prev);
            if (prev != a.device) cudaSetDevice(a.device);
            else prev = -1;
        } else {
            cudaGetLastError();  // not a device pointer known to the runtime: leave the current device alone
        }
    }
    ~PtrDeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};
}  // namespace

// largest float s with sqrtf(s) <= lim (sqrtf is correctly rounded, like __fsqrt_rn and numpy's float32 sqrt)
static float sq_limit(float lim) {
    float s = lim * lim;
    while (sqrtf(s) > lim) s = nextafterf(s, 0.0f);
    while (sqrtf(nextafterf(s, INFINITY)) <= lim) s = nextafterf(s, INFINITY);
    return s;
}

// One launch path for the per-env kernels.  pdl: programmatic dependent launch (sm_90+) -- the grid may become resident
// while the preceding kernel of the stream is still running; the kernel orders itself behind that kernel's results
// with griddepcontrol.wait (k_step: just before it reads actions / attacker forces).
template <typename... KArgs, typename... Args>
static void launch_env_kernel(void (*kernel)(KArgs...), int grid, int threads, int smem, cudaStream_t st, bool pdl,
                              Args... args) {
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((unsigned)grid);
    lc.blockDim = dim3((unsigned)threads);
    lc.dynamicSmemBytes = (size_t)smem;
    lc.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at;
    lc.numAttrs = pdl ? 1 : 0;
    cudaLaunchKernelEx(&lc, kernel, KArgs(args)...);  // errors surface through cudaGetLastError() at the call site
}

// Kernel instantiations: the BASELINE team sizes get compile-time layouts, everything else the generic build.
#define DCA_DISPATCH(h, PDL, KERNEL, ...)                                                                      \
    do {                                                                                                       \
        const int G_ = (h)->P.L.G, A_ = (h)->P.L.A, E_ = (h)->P.E, sm_ = (h)->smem;                            \
        if ((h)->threads == 128) {                                                                             \
            if (G_ == 50 && A_ == 50) launch_env_kernel(KERNEL<128, 50, 50>, E_, 128, sm_, st_, PDL, __VA_ARGS__);      \
            else if (G_ == 50 && A_ == 55) launch_env_kernel(KERNEL<128, 50, 55>, E_, 128, sm_, st_, PDL, __VA_ARGS__); \
            else launch_env_kernel(KERNEL<128, 0, 0>, E_, 128, sm_, st_, PDL, __VA_ARGS__);                    \
        } else {                                                                                               \
            if (G_ == 100 && A_ == 100) launch_env_kernel(KERNEL<256, 100, 100>, E_, 256, sm_, st_, PDL, __VA_ARGS__);  \
            else launch_env_kernel(KERNEL<256, 0, 0>, E_, 256, sm_, st_, PDL, __VA_ARGS__);                    \
        }                                                                                                      \
    } while (0)

template <typename K>
static cudaError_t allow_smem(K kernel, int bytes) {
    return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}
#define DCA_ALLOW_SMEM(KERNEL, bytes)                                                                   \
    do {                                                                                                \
        if (e == cudaSuccess) e = allow_smem(KERNEL<128, 50, 50>, bytes);                               \
        if (e == cudaSuccess) e = allow_smem(KERNEL<128, 50, 55>, bytes);                               \
        if (e == cudaSuccess) e = allow_smem(KERNEL<128, 0, 0>, bytes);                                 \
        if (e == cudaSuccess) e = allow_smem(KERNEL<256, 100, 100>, bytes);                             \
        if (e == cudaSuccess) e = allow_smem(KERNEL<256, 0, 0>, bytes);                                 \
    } while (0)

// Hand-over words for a planner launch: it waits per env for the step launched right before it on the same stream
// (DCA_OPT_PIPELINE_STEPS) -- otherwise it is a normal launch behind everything earlier in the stream -- and always
// publishes its own epoch.
// A stream that is being captured into a CUDA graph gets ordinary launches: a replay would publish the captured epoch
// values again, so nothing may wait on them.
static bool stream_capturing(cudaStream_t st) {
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    return cudaStreamIsCapturing(st, &cs) == cudaSuccess && cs != cudaStreamCaptureStatusNone;
}

static Pipe planner_pipe(DcaHandle *h, cudaStream_t st) {
    Pipe pipe;
    if (h->pipeline_steps && stream_capturing(st)) {
        pipe.epoch = nullptr; pipe.wait_step = pipe.wait_plan = pipe.publish = 0u;
        h->last_was_step = false;
        return pipe;
    }
    pipe.epoch = h->pipeline_steps ? h->epoch : nullptr;
    pipe.wait_step = (h->pipeline_steps && h->last_was_step && h->last_step_stream == st) ? h->step_epoch : 0u;
    pipe.wait_plan = 0u;
    pipe.publish = h->pipeline_steps ? ++h->plan_epoch : 0u;
    return pipe;
}

extern "C" {

int dca_abi_version(void) { return 2; }

const char *dca_last_error(const DcaHandle *h) { return h ? h->err.c_str() : g_err.c_str(); }

int dca_create(const DcaConfig *cfg, int n_env, int device, DcaHandle **out) {
    if (!cfg || !out || n_env <= 0) return fail(nullptr, DCA_E_INVALID, "dca_create: null argument or n_env <= 0");
    *out = nullptr;
    const int G = cfg->num_guards, A = cfg->num_attackers;
    // the reference observation needs >= 6 candidates per team (argsort()[:6], collective_assault_env.py:308-309)
    if (G < kVisN || A < kVisN || G > 250 || A > 250 || G + A > kMaxN)
        return fail(nullptr, DCA_E_INVALID, "dca_create: need 6 <= num_guards, num_attackers <= 250 and N <= 320");
    if (cfg->enable_attacker_fluctuation && (cfg->attacker_fluctuation < 0 || cfg->attacker_fluctuation > A))
        return fail(nullptr, DCA_E_INVALID, "dca_create: attacker_fluctuation out of range");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(nullptr, DCA_E_NODEVICE, "dca_create: no CUDA device (this library has no CPU fallback)");
    if (device < 0 || device >= ndev) return fail(nullptr, DCA_E_INVALID, "dca_create: bad device index");
    DeviceGuard guard(device);
    DcaHandle *h = new (std::nothrow) DcaHandle();
    if (!h) return fail(nullptr, DCA_E_INVALID, "dca_create: out of host memory");
    h->device = device;
    h->bound = false;
    h->launches = 0;
    h->overlap_planner = false;
    h->planned_force = nullptr;
    h->planned_stream = nullptr;
    h->pipeline_steps = false;
    h->epoch = nullptr;
    h->step_epoch = h->plan_epoch = 0;
    h->last_was_step = false;
    h->last_step_stream = nullptr;
    std::memset(&h->P, 0, sizeof(Params));
    h->P.cfg = *cfg;
    h->P.L = make_layout(G, A);
    h->P.E = n_env;
    h->P.smax[0] = sq_limit(2.0f);  // obs_range collective_assault_env.py:284
    h->P.smax[1] = sq_limit(3.0f);  // obs_range * 1.5 :332
    {
        const int cap = (G > A ? G : A) + 1;  // candidate indices and the marker are <= cap - 1
        int bits = 5;
        while ((1 << bits) < cap) ++bits;
        h->P.idx_bits = bits;  // 6 for teams of <= 63, 7 for <= 127, 8 for <= 255
    }
    {
        // (double)w * 2^-32 < p  <=>  w < p * 2^32 (exact scaling)  <=>  w < ceil(p * 2^32) for integer w
        const double t = std::ceil(cfg->random_jam_prob * 4294967296.0);
        h->P.jam_thresh = t <= 0 ? 0ull : (t >= 4294967296.0 ? 4294967296ull : (unsigned long long)t);
    }
    {
        // laser_hit_improve3 (cython_func.pyx:22-46) accepts a target iff |reg(dir + w/2 - theta)| <= w and
        // |reg(dir - w/2 - theta)| <= w; for w < 2 pi / 3 the wrapped branches can never pass, and the test is "the
        // target direction lies within w/2 of the heading": dot(heading, d) >= cos(w/2) |d|
        const double w = cfg->agent_shoot_win;
        h->P.cone_cos2 = (w > 0.0 && w <= 2.0) ? std::cos(w / 2) * std::cos(w / 2) : 0.0;
    }
    h->threads = (G + A) <= 128 ? 128 : 256;
    h->smem = smem_bytes(h->P.L);
    const size_t bytes = (size_t)n_env * h->P.L.bytes;
    cudaError_t e = cudaMalloc((void **)&h->P.blobs, bytes);
    if (e == cudaSuccess) e = cudaMemset(h->P.blobs, 0, bytes);
    if (e == cudaSuccess) e = cudaMalloc((void **)&h->epoch, sizeof(uint32_t) * 2 * (size_t)n_env);
    if (e == cudaSuccess) e = cudaMemset(h->epoch, 0, sizeof(uint32_t) * 2 * (size_t)n_env);
    {   // persistent team state of the scripted attacker planner (zero = teams not formed yet)
        const int K = A / 5 > 0 ? A / 5 : 1;
        std::memset(&h->plan, 0, sizeof(PlanState));
        h->plan_smem = plan_smem_bytes(G, A);
        if (e == cudaSuccess) e = cudaMalloc((void **)&h->plan.flags, (size_t)n_env * 2);
        if (e == cudaSuccess) e = cudaMemset(h->plan.flags, 0, (size_t)n_env * 2);
        if (e == cudaSuccess) e = cudaMalloc((void **)&h->plan.counter, (size_t)n_env * 4);
        if (e == cudaSuccess) e = cudaMemset(h->plan.counter, 0, (size_t)n_env * 4);
        if (e == cudaSuccess) e = cudaMalloc((void **)&h->plan.label, (size_t)n_env * A);
        if (e == cudaSuccess) e = cudaMemset(h->plan.label, 0, (size_t)n_env * A);
        if (e == cudaSuccess) e = cudaMalloc((void **)&h->plan.cent, (size_t)n_env * K * 16);
        if (e == cudaSuccess) e = cudaMemset(h->plan.cent, 0, (size_t)n_env * K * 16);
        h->plan.big_n = A > kPlanTeamPre2 ? std::min(A, kPlanBigMax) : 0;  // cost-matrix scratch of the large teams
        if (e == cudaSuccess && h->plan.big_n > 0)
            e = cudaMalloc((void **)&h->plan.big, (size_t)n_env * h->plan.big_n * h->plan.big_n * sizeof(double));
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(k_plan<0, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     std::max(plan_smem_bytes(250, 250), plan_smem_bytes(kMaxN - 6, 6)));
    }
    // cudaFuncAttributeMaxDynamicSharedMemorySize is a per-function (per-device) setting shared by every handle: raise
    // it once to what the largest admissible layout needs instead of each handle overwriting it with its own size
    {
        const int max_smem = smem_bytes(make_layout(kMaxN / 2, kMaxN - kMaxN / 2));
        DCA_ALLOW_SMEM(k_step, max_smem);
        DCA_ALLOW_SMEM(k_reset, max_smem);
        DCA_ALLOW_SMEM(k_observe, max_smem);
    }
    if (e == cudaSuccess) {
        k_init_blobs<<<(n_env + 255) / 256, 256>>>(h->P);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        const int rc = fail(nullptr, DCA_E_CUDA, std::string("dca_create: ") + cudaGetErrorString(e));
        if (h->P.blobs) cudaFree(h->P.blobs);
        if (h->epoch) cudaFree(h->epoch);
        cudaFree(h->plan.flags); cudaFree(h->plan.counter); cudaFree(h->plan.label); cudaFree(h->plan.cent);
    cudaFree(h->plan.big);
        cudaFree(h->plan.big);
        delete h;
        return rc;
    }
    h->launches += 1;
    *out = h;
    return DCA_OK;
}

int dca_destroy(DcaHandle *h) {
    if (!h) return DCA_E_INVALID;
    DeviceGuard guard(h->device);
    cudaFree(h->P.blobs);
    cudaFree(h->epoch);
    cudaFree(h->plan.flags); cudaFree(h->plan.counter); cudaFree(h->plan.label); cudaFree(h->plan.cent);
    delete h;
    return DCA_OK;
}

int dca_bind_buffers(DcaHandle *h, const DcaBuffers *b) {
    if (!h || !b) return fail(h, DCA_E_INVALID, "dca_bind_buffers: null argument");
    if (!b->obs || !b->reward || !b->done || !b->win)
        return fail(h, DCA_E_INVALID, "dca_bind_buffers: obs, reward, done and win are required");
    if (((uintptr_t)b->obs & 31) || (b->tobs && ((uintptr_t)b->tobs & 31)))
        return fail(h, DCA_E_INVALID, "dca_bind_buffers: obs/tobs must be 32-byte aligned");
    h->P.out = *b;
    h->bound = true;
    return DCA_OK;
}

static int check_ready(DcaHandle *h, const char *who) {
    if (!h) return fail(nullptr, DCA_E_INVALID, std::string(who) + ": null handle");
    if (!h->bound) return fail(h, DCA_E_UNBOUND, std::string(who) + ": call dca_bind_buffers first");
    return DCA_OK;
}

int dca_reset(DcaHandle *h, const uint8_t *env_mask, dca_stream_t stream) {
    if (int rc = check_ready(h, "dca_reset")) return rc;
    DeviceGuard guard(h->device);
    cudaStream_t st_ = (cudaStream_t)stream;
    DCA_DISPATCH(h, false, k_reset, h->P, env_mask);
    CUDA_TRY(h, cudaGetLastError());
    h->planned_force = nullptr;
    h->last_was_step = false;
    h->launches += 1;
    return DCA_OK;
}

int dca_step(DcaHandle *h, const int32_t *actions, const double *att_force, const uint8_t *skip,
             const uint32_t *rng_words, int autoreset, dca_stream_t stream) {
    if (int rc = check_ready(h, "dca_step")) return rc;
    if (!actions || !att_force) return fail(h, DCA_E_INVALID, "dca_step: actions and att_force are required");
    if (((uintptr_t)att_force & 15) || (rng_words && ((uintptr_t)rng_words & 15)))
        return fail(h, DCA_E_INVALID, "dca_step: att_force / rng_words must be 16-byte aligned");
    DeviceGuard guard(h->device);
    cudaStream_t st_ = (cudaStream_t)stream;
    // Overlap with the planner (DCA_OPT_OVERLAP_PLANNER): only when this handle's previous launch was the planner that
    // filled this att_force on this stream -- then everything k_step touches before its pdl_wait() (its own blobs,
    // `skip`) is settled, and the forces are read after the wait.
    const bool behind_planner = h->planned_force == (const void *)att_force && h->planned_stream == st_;
    // DCA_OPT_PIPELINE_STEPS: the step takes its envs over one by one -- env e starts once the previous step launch has
    // published e (its blob; trivially true when the planner was a normal launch behind that step) and reads e's forces
    // once the planner warp of e has published them.  Every step / planner launch publishes, piped or not.
    const bool live = h->pipeline_steps && !stream_capturing(st_);  // captured launches neither wait nor publish
    const bool piped = live && behind_planner && skip == nullptr;
    const bool pdl = piped || (h->overlap_planner && behind_planner);
    Pipe pipe;
    pipe.epoch = live ? h->epoch : nullptr;  // epochs are published only while the option is on
    pipe.wait_step = piped ? h->step_epoch : 0u;  // 0 on the very first step: nothing to wait for
    pipe.wait_plan = piped ? h->plan_epoch : 0u;
    pipe.publish = live ? ++h->step_epoch : 0u;
    DCA_DISPATCH(h, pdl, k_step, h->P, actions, att_force, skip, rng_words, autoreset, pipe);
    CUDA_TRY(h, cudaGetLastError());
    h->planned_force = nullptr;
    h->last_was_step = live && skip == nullptr;  // a pause mask orders the launch behind the whole previous kernel anyway
    h->last_step_stream = st_;
    h->launches += 1;
    return DCA_OK;
}

int dca_observe(DcaHandle *h, const uint32_t *rng_words, dca_stream_t stream) {
    if (int rc = check_ready(h, "dca_observe")) return rc;
    DeviceGuard guard(h->device);
    cudaStream_t st_ = (cudaStream_t)stream;
    DCA_DISPATCH(h, false, k_observe, h->P, rng_words);
    CUDA_TRY(h, cudaGetLastError());
    h->planned_force = nullptr;
    h->last_was_step = false;
    h->launches += 1;
    return DCA_OK;
}

int dca_plan_centroid(DcaHandle *h, double *att_force, dca_stream_t stream) {
    if (!h || !att_force) return fail(h, DCA_E_INVALID, "dca_plan_centroid: null argument");
    DeviceGuard guard(h->device);
    const int threads = 128, warps_per_block = threads / 32;
    const Pipe pipe = planner_pipe(h, (cudaStream_t)stream);
    launch_env_kernel(k_plan_centroid, (h->P.E + warps_per_block - 1) / warps_per_block, threads, 0, (cudaStream_t)stream,
                      pipe.wait_step != 0, h->P, att_force, pipe);
    CUDA_TRY(h, cudaGetLastError());
    h->planned_force = att_force;
    h->planned_stream = (cudaStream_t)stream;
    h->last_was_step = false;
    h->launches += 1;
    return DCA_OK;
}

int dca_plan(DcaHandle *h, double *att_force, const int32_t *team_label, const double *team_centroid,
             dca_stream_t stream) {
    if (!h || !att_force) return fail(h, DCA_E_INVALID, "dca_plan: null argument");
    if ((team_label == nullptr) != (team_centroid == nullptr))
        return fail(h, DCA_E_INVALID, "dca_plan: team_label and team_centroid are injected together or not at all");
    if ((uintptr_t)att_force & 15) return fail(h, DCA_E_INVALID, "dca_plan: att_force must be 16-byte aligned");
    DeviceGuard guard(h->device);
    {   // compile-time team sizes for the BASELINE configurations, the generic build for everything else
        const int G = h->P.L.G, A = h->P.L.A;
        auto kernel = k_plan<0, 0>;
        if (G == 50 && A == 50) kernel = k_plan<50, 50>;
        else if (G == 50 && A == 55) kernel = k_plan<50, 55>;
        else if (G == 100 && A == 100) kernel = k_plan<100, 100>;
        const Pipe pipe = planner_pipe(h, (cudaStream_t)stream);
        launch_env_kernel(kernel, h->P.E, kPlanThreads, h->plan_smem, (cudaStream_t)stream, pipe.wait_step != 0, h->P,
                          h->plan, att_force, team_label, team_centroid, pipe);
        }
    CUDA_TRY(h, cudaGetLastError());
    h->planned_force = att_force;
    h->planned_stream = (cudaStream_t)stream;
    h->last_was_step = false;
    h->launches += 1;
    return DCA_OK;
}

static int state_io(DcaHandle *h, const DcaState *st, int import, dca_stream_t stream) {
    if (!h || !st) return fail(h, DCA_E_INVALID, "dca_state_io: null argument");
    const void *ptrs[] = {st->pos, st->vel, st->atk_rad, st->terrain, st->alive, st->just_died, st->hit, st->was_hit,
                          st->num_hit, st->num_was_hit, st->init_theta, st->time_step, st->n_alive_guards,
                          st->n_alive_attackers, st->obs_counter, st->episode};
    for (const void *p : ptrs)
        if (!p) return fail(h, DCA_E_INVALID, "dca_state_io: every DcaState pointer is required");
    DeviceGuard guard(h->device);
    const size_t total = (size_t)h->P.E * h->P.L.N;
    const int blocks = (int)((total + 255) / 256);
    k_state_io<<<blocks, 256, 0, (cudaStream_t)stream>>>(h->P, *st, import);
    CUDA_TRY(h, cudaGetLastError());
    h->planned_force = nullptr;
    h->last_was_step = false;
    h->launches += 1;
    return DCA_OK;
}

int dca_export_state(DcaHandle *h, const DcaState *dst, dca_stream_t stream) { return state_io(h, dst, 0, stream); }
int dca_import_state(DcaHandle *h, const DcaState *src, dca_stream_t stream) { return state_io(h, src, 1, stream); }

int dca_gae(const float *reward, const float *value, const uint8_t *dead, const int32_t *traj_len, float *ret,
            float *threat, float *reward_pushed, int T, int E, int G, double gamma, double tau, double gamma_fwd,
            int threat_limit, dca_stream_t stream) {
    if (!reward || !value || !dead || !ret || T <= 0 || E <= 0 || G <= 0)
        return fail(nullptr, DCA_E_INVALID, "dca_gae: null argument or empty shape");
    if (!(gamma_fwd == gamma_fwd))
        return fail(nullptr, DCA_E_INVALID, "dca_gae: gamma_fwd is NaN (pass 1.0 when gamma_in_reward_forwarding is off)");
    PtrDeviceGuard guard(reward);
    const size_t S = (size_t)E * G;
    const int threads = 128;
    k_gae<<<(unsigned)((S + threads - 1) / threads), threads, 0, (cudaStream_t)stream>>>(
        reward, value, dead, traj_len, ret, threat, reward_pushed, T, E, G, gamma, tau, gamma_fwd, threat_limit);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_gae: ") + cudaGetErrorString(e));
    return DCA_OK;
}

int dca_runner_init(const DcaRunnerState *s, int n_env, int n_agent, dca_stream_t stream) {
    if (!s || !s->pause || !s->suffered || !s->cur_step || !s->ep_cnt || !s->reward_sum || !s->counters || !s->fin_slot ||
        !s->log_meta || !s->log_reward || n_env <= 0 || n_agent <= 0 || s->log_capacity < n_env)
        return fail(nullptr, DCA_E_INVALID, "dca_runner_init: null buffer, empty shape or log_capacity < n_env");
    PtrDeviceGuard guard(s->pause);
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t e = cudaMemsetAsync(s->pause, 0, n_env, st);
    auto also = [&e](cudaError_t r) { if (e == cudaSuccess) e = r; };
    also(cudaMemsetAsync(s->suffered, 1, n_env, st));
    also(cudaMemsetAsync(s->cur_step, 0, sizeof(int32_t) * n_env, st));
    also(cudaMemsetAsync(s->ep_cnt, 0, sizeof(int32_t) * n_env, st));
    also(cudaMemsetAsync(s->reward_sum, 0, sizeof(double) * n_env * n_agent, st));
    also(cudaMemsetAsync(s->counters, 0, sizeof(int64_t) * 8, st));
    also(cudaMemsetAsync(s->fin_slot, 0, sizeof(int32_t) * n_env, st));
    also(cudaMemsetAsync(s->log_meta, 0, sizeof(int64_t) * 4 * s->log_capacity, st));
    also(cudaMemsetAsync(s->log_reward, 0, sizeof(double) * (size_t)s->log_capacity * n_agent, st));
    if (e != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_runner_init: ") + cudaGetErrorString(e));
    return DCA_OK;
}

int dca_runner_update(const DcaRunnerState *s, int n_env, int n_agent, int align_episode, const uint8_t *done,
                      const uint8_t *win, const float *reward, dca_stream_t stream) {
    if (!s || !done || !win || !reward || n_env <= 0 || n_agent <= 0 || s->log_capacity < n_env)
        return fail(nullptr, DCA_E_INVALID, "dca_runner_update: null argument, empty shape or log_capacity < n_env");
    PtrDeviceGuard guard(s->pause);
    cudaStream_t st = (cudaStream_t)stream;
    k_runner_scan<<<1, kRunnerThreads, 0, st>>>(*s, n_env, align_episode != 0, done, win);
    const size_t n = (size_t)n_env * n_agent;
    k_runner_rewards<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(*s, n_env, n_agent, reward);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_runner_update: ") + cudaGetErrorString(e));
    return DCA_OK;
}

int dca_runner_report(const DcaRunnerState *s, int n_agent, int64_t first_episode, int n_episodes,
                      const int32_t *team_of_agent, int n_team, double *out, dca_stream_t stream) {
    if (!s || !team_of_agent || !out || n_agent <= 0 || n_team <= 0 || first_episode < 1 || n_episodes <= 0 ||
        n_episodes > s->log_capacity)
        return fail(nullptr, DCA_E_INVALID, "dca_runner_report: bad window (1-based first episode, at most log_capacity)");
    PtrDeviceGuard guard(out);
    k_runner_report<<<1, 256, 0, (cudaStream_t)stream>>>(*s, n_agent, (long long)first_episode, n_episodes,
                                                         team_of_agent, n_team, out);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_runner_report: ") + cudaGetErrorString(e));
    return DCA_OK;
}

int dca_histpool_roll(const float *obs, float *his_pool, const uint8_t *pause, const uint8_t *suffered_reset,
                      float *obs_feed, int n_env, int n_agent, dca_stream_t stream) {
    if (!obs || !his_pool || !obs_feed || n_env <= 0 || n_agent <= 0)
        return fail(nullptr, DCA_E_INVALID, "dca_histpool_roll: null buffer or empty shape");
    if (((uintptr_t)obs | (uintptr_t)his_pool | (uintptr_t)obs_feed) & 15)
        return fail(nullptr, DCA_E_INVALID, "dca_histpool_roll: buffers must be 16-byte aligned");
    PtrDeviceGuard guard(obs);
    const size_t n = (size_t)n_env * n_agent;  // one warp each
    k_histpool<<<(unsigned)((n + 7) / 8), 256, 0, (cudaStream_t)stream>>>(
        (const float4 *)obs, (float4 *)his_pool, pause, suffered_reset, (float4 *)obs_feed, n_env, n_agent);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_histpool_roll: ") + cudaGetErrorString(e));
    return DCA_OK;
}

int dca_adv_moments(const float *ret, const float *value, const uint8_t *valid, int valid_group, uint64_t n,
                    double *scratch, double *stats, dca_stream_t stream) {
    if (!ret || !value || !scratch || !stats || n == 0 || (valid && valid_group <= 0))
        return fail(nullptr, DCA_E_INVALID, "dca_adv_moments: null buffer, empty batch or bad valid_group");
    PtrDeviceGuard guard(ret);
    cudaStream_t st = (cudaStream_t)stream;
    k_adv_partial<<<kAdvBlocks, kAdvThreads, 0, st>>>(ret, value, valid, valid ? valid_group : 1, (size_t)n, scratch);
    k_adv_final<<<1, kAdvThreads, 0, st>>>(scratch, stats);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_adv_moments: ") + cudaGetErrorString(e));
    return DCA_OK;
}

int dca_adv_normalize(const float *ret, const float *value, const uint8_t *valid, int valid_group, uint64_t n,
                      const double *stats, float *adv, dca_stream_t stream) {
    if (!ret || !value || !stats || !adv || n == 0 || (valid && valid_group <= 0))
        return fail(nullptr, DCA_E_INVALID, "dca_adv_normalize: null buffer, empty batch or bad valid_group");
    PtrDeviceGuard guard(ret);
    const size_t blocks = std::min<size_t>((n + kAdvThreads - 1) / kAdvThreads, 148 * 16);
    k_adv_normalize<<<(unsigned)blocks, kAdvThreads, 0, (cudaStream_t)stream>>>(ret, value, valid, valid ? valid_group : 1,
                                                                             (size_t)n, stats, adv);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_adv_normalize: ") + cudaGetErrorString(e));
    return DCA_OK;
}

int dca_norm_moments(const float *x, uint64_t rows, int dim, double *scratch, double *stats, dca_stream_t stream) {
    if (!x || !scratch || !stats || rows == 0)
        return fail(nullptr, DCA_E_INVALID, "dca_norm_moments: null buffer or empty batch");
    if (dim != kNormDim || ((uintptr_t)x & 15))
        return fail(nullptr, DCA_E_INVALID, "dca_norm_moments: rows must be 16 floats (obs_vec_length) and 16-byte aligned");
    PtrDeviceGuard guard(x);
    cudaStream_t st = (cudaStream_t)stream;
    k_norm_partial<<<kNormBlocks, kNormThreads, 0, st>>>((const float4 *)x, (size_t)rows, scratch);
    k_norm_final<<<1, kNormFinalThreads, 0, st>>>(scratch, stats);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_norm_moments: ") + cudaGetErrorString(e));
    return DCA_OK;
}

int dca_norm_apply(const float *x, const float *mean, const float *var, uint64_t rows, int dim, float *y,
                   uint8_t *row_nan, dca_stream_t stream) {
    if (!x || !mean || !var || !y || !row_nan || rows == 0)
        return fail(nullptr, DCA_E_INVALID, "dca_norm_apply: null buffer or empty batch");
    if (dim != kNormDim || (((uintptr_t)x | (uintptr_t)y | (uintptr_t)mean | (uintptr_t)var) & 15))
        return fail(nullptr, DCA_E_INVALID, "dca_norm_apply: rows must be 16 floats (obs_vec_length), buffers 16-byte aligned");
    PtrDeviceGuard guard(x);
    const size_t blocks = std::min<size_t>((rows * 4 + 255) / 256, 148 * 16);
    k_norm_apply<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const float4 *)x, mean, var, (size_t)rows, (float4 *)y,
                                                                    row_nan);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_norm_apply: ") + cudaGetErrorString(e));
    return DCA_OK;
}

int dca_conc_attend(const float *v, const float *x, const uint8_t *dead, const float *qk, int first, int n, int n_focus,
                    uint64_t n_agent, float *va, float *stack, dca_stream_t stream) {
    if (!v || !x || !dead || !qk || !va || !stack || n_agent == 0)
        return fail(nullptr, DCA_E_INVALID, "dca_conc_attend: null buffer or empty batch");
    if (first < 1 || n < 1 || n > 32 || first + n > kPolEnt || n_focus < 1 || n_focus > n)
        return fail(nullptr, DCA_E_INVALID, "dca_conc_attend: entity range must lie in [1, 24) and 1 <= n_focus <= n <= 32");
    if (((uintptr_t)v | (uintptr_t)qk) & 15)
        return fail(nullptr, DCA_E_INVALID, "dca_conc_attend: v and qk must be 16-byte aligned");
    PtrDeviceGuard guard(v);
    const size_t blocks = (n_agent * 32 + 255) / 256;
    k_conc_attend<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(v, x, dead, qk, first, n, n_focus, (size_t)n_agent, va,
                                                                       stack);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_conc_attend: ") + cudaGetErrorString(e));
    return DCA_OK;
}

int dca_conc_attend2(const float *v, const float *x, const uint8_t *dead, const float *qk_a, const float *qk_b, int first_a,
                     int n_a, int focus_a, int n_b, int focus_b, uint64_t n_agent, float *va_a, float *stack_a, float *va_b,
                     float *stack_b, dca_stream_t stream) {
    if (!v || !x || !dead || !qk_a || !qk_b || !va_a || !stack_a || !va_b || !stack_b || n_agent == 0)
        return fail(nullptr, DCA_E_INVALID, "dca_conc_attend2: null buffer or empty batch");
    if (first_a < 1 || n_a < 1 || n_b < 1 || first_a + n_a + n_b > 24 || focus_a < 1 || focus_a > n_a || focus_b < 1 || focus_b > n_b)
        return fail(nullptr, DCA_E_INVALID, "dca_conc_attend2: the two adjacent entity lists must lie in [1, 24), 1 <= n_focus <= n");
    if (((uintptr_t)v | (uintptr_t)x | (uintptr_t)qk_a | (uintptr_t)qk_b) & 15)
        return fail(nullptr, DCA_E_INVALID, "dca_conc_attend2: v, x and qk must be 16-byte aligned");
    PtrDeviceGuard guard(v);
    const int threads = 256, warps = threads / 32;
    const size_t smem = sizeof(AttWarpSmem) * warps;
    static bool attr_set[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
        cudaError_t ea = cudaFuncSetAttribute(k_conc_attend2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (ea != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_conc_attend2: ") + cudaGetErrorString(ea));
        attr_set[dev] = true;
    }
    const unsigned grid = (unsigned)((n_agent + warps - 1) / warps);
    k_conc_attend2<<<grid, threads, smem, (cudaStream_t)stream>>>(v, x, dead, qk_a, qk_b, first_a, n_a, focus_a, n_b, focus_b,
                                                                  (size_t)n_agent, va_a, stack_a, va_b, stack_b);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_conc_attend2: ") + cudaGetErrorString(e));
    return DCA_OK;
}

int dca_policy_encode(const float *x, const float *mean, const float *var, const float *w1, const float *b1,
                      const float *w2, const float *b2, uint64_t rows, float *x_norm, uint8_t *row_nan, float *v,
                      dca_stream_t stream) {
    if (!x || !mean || !var || !w1 || !b1 || !w2 || !b2 || !x_norm || !row_nan || !v || rows == 0)
        return fail(nullptr, DCA_E_INVALID, "dca_policy_encode: null buffer or empty batch");
    if (((uintptr_t)x | (uintptr_t)x_norm | (uintptr_t)mean | (uintptr_t)var | (uintptr_t)v) & 15)
        return fail(nullptr, DCA_E_INVALID, "dca_policy_encode: buffers must be 16-byte aligned");
    PtrDeviceGuard guard(x);
    static bool attr_set[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
        cudaError_t ea = cudaFuncSetAttribute(k_policy_encode, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(EncSmem));
        if (ea == cudaSuccess)
            ea = cudaFuncSetAttribute(k_policy_encode_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(EncTcSmem));
        if (ea != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_policy_encode: ") + cudaGetErrorString(ea));
        attr_set[dev] = true;
    }
    const size_t tiles = (rows + kEncTile - 1) / kEncTile;
#ifdef DCA_ENCODE_FFMA  // the float32 FFMA encoder (before the tensor-core one existed); kept for A/B builds
    const unsigned grid = (unsigned)std::min<size_t>(tiles, 148 * 4);
    k_policy_encode<<<grid, kEncThreads, sizeof(EncSmem), (cudaStream_t)stream>>>(
        (const float4 *)x, mean, var, w1, b1, w2, b2, (size_t)rows, (float4 *)x_norm, row_nan, v);
#else
    const unsigned grid = (unsigned)std::min<size_t>((rows + kEncTcTile - 1) / kEncTcTile, 148 * 2);
    k_policy_encode_tc<<<grid, kEncTcThreads, sizeof(EncTcSmem), (cudaStream_t)stream>>>(
        (const float4 *)x, mean, var, w1, b1, w2, b2, (size_t)rows, (float4 *)x_norm, row_nan, v);
#endif
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_policy_encode: ") + cudaGetErrorString(e));
    return DCA_OK;
}

int dca_policy_linear(const float *a, int lda, int ka, const float *wa, int wa_sn, int wa_sk, const float *b, int ldb, int kb,
                      const float *wb, int wb_sn, int wb_sk, const float *bias, int n, int relu, uint64_t rows, float *out,
                      int ldo, dca_stream_t stream) {
    if (!a || !wa || !out || rows == 0 || ka <= 0 || n <= 0 || (kb > 0 && (!b || !wb)))
        return fail(nullptr, DCA_E_INVALID, "dca_policy_linear: null buffer or empty batch");
    if (n > 48 || (ka & 3) || (kb & 3) || kb < 0 || (lda & 3) || (kb > 0 && (ldb & 3)) || ldo < n)
        return fail(nullptr, DCA_E_INVALID, "dca_policy_linear: n <= 48, ka / kb / lda / ldb multiples of 4, ldo >= n");
    if (((uintptr_t)a & 15) || (kb > 0 && ((uintptr_t)b & 15)))
        return fail(nullptr, DCA_E_INVALID, "dca_policy_linear: inputs must be 16-byte aligned");
    PtrDeviceGuard guard(a);
    const LinSrc sa{a, wa, lda, ka, wa_sn, wa_sk};
    const LinSrc sb{b, wb, ldb, kb > 0 ? kb : 0, wb_sn, wb_sk};
    const size_t tiles = (rows + kLinTile - 1) / kLinTile;
    const unsigned grid = (unsigned)std::min<size_t>(tiles, 148 * 4);
    cudaStream_t st = (cudaStream_t)stream;
    const int act = relu & 1;
    if ((relu & 2) || (ka & 7) || (kb & 7)) {  // float32 FFMA path (asked for, or a k that is not a multiple of 8)
        if (n > 24) k_policy_linear<6><<<grid, kLinThreads, 0, st>>>(sa, sb, bias, n, act, (size_t)rows, out, ldo);
        else if (n > 8) k_policy_linear<3><<<grid, kLinThreads, 0, st>>>(sa, sb, bias, n, act, (size_t)rows, out, ldo);
        else k_policy_linear<1><<<grid, kLinThreads, 0, st>>>(sa, sb, bias, n, act, (size_t)rows, out, ldo);
    } else {                                  // 3xTF32 on the tensor cores
        static bool attr_set[64] = {};
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev >= 0 && dev < 64 && !attr_set[dev]) {
            cudaError_t ea = cudaFuncSetAttribute(k_policy_linear_tc<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(LinSmemTc));
            if (ea == cudaSuccess) ea = cudaFuncSetAttribute(k_policy_linear_tc<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(LinSmemTc));
            if (ea == cudaSuccess) ea = cudaFuncSetAttribute(k_policy_linear_tc<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(LinSmemTc));
            if (ea != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_policy_linear: ") + cudaGetErrorString(ea));
            attr_set[dev] = true;
        }
        const size_t sm = sizeof(LinSmemTc);
        if (n > 24) k_policy_linear_tc<6><<<grid, kLinThreads, sm, st>>>(sa, sb, bias, n, act, (size_t)rows, out, ldo);
        else if (n > 8) k_policy_linear_tc<3><<<grid, kLinThreads, sm, st>>>(sa, sb, bias, n, act, (size_t)rows, out, ldo);
        else k_policy_linear_tc<1><<<grid, kLinThreads, sm, st>>>(sa, sb, bias, n, act, (size_t)rows, out, ldo);
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_policy_linear: ") + cudaGetErrorString(e));
    return DCA_OK;
}

// ---- dca_adv_moments_static: the same two kernels with the per-CTA partials in a library-owned array ---------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (graph capture, serving loops).  The array lives in every GPU context that loads the library until exit.
static __device__ __align__(256) double g_adv_static_scratch[dca::kAdvBlocks * 3];  // 592 x 3 doubles = 14 208 B

namespace {
constexpr int kStaticMaxDevices = 64;
double *g_adv_static_ptr[kStaticMaxDevices];  // device address of g_adv_static_scratch on each device, resolved once
bool g_adv_static_ready[kStaticMaxDevices];
std::mutex g_adv_static_mu;

// In host code `g_adv_static_scratch` is only a stand-in for the symbol, never a device address: resolve the address once per
// device, with that device current, and hand the kernels the resolved pointer.
cudaError_t adv_static_scratch(int device, double **out) {
    std::lock_guard<std::mutex> lock(g_adv_static_mu);
    if (!g_adv_static_ready[device]) {
        double *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void **)&p, g_adv_static_scratch);
        if (e != cudaSuccess) return e;
        g_adv_static_ptr[device] = p;
        g_adv_static_ready[device] = true;
    }
    *out = g_adv_static_ptr[device];
    return cudaSuccess;
}
}  // namespace

int dca_adv_moments_static(const float *ret, const float *value, const uint8_t *valid, int valid_group, uint64_t n,
                           double *stats, dca_stream_t stream) {
    if (!ret || !value || !stats || n == 0 || (valid && valid_group <= 0))
        return fail(nullptr, DCA_E_INVALID, "dca_adv_moments_static: null buffer, empty batch or bad valid_group");
    if (n > DCA_ADV_STATIC_MAX_N)
        return fail(nullptr, DCA_E_INVALID, "dca_adv_moments_static: n = " + std::to_string(n) + " exceeds DCA_ADV_STATIC_MAX_N = " +
                                                std::to_string((uint64_t)DCA_ADV_STATIC_MAX_N) +
                                                " elements; use dca_adv_moments with a caller-owned scratch buffer");
    PtrDeviceGuard guard(ret);
    int device = -1;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_adv_moments_static: cudaGetDevice: ") + cudaGetErrorString(e));
    if (device < 0 || device >= kStaticMaxDevices)
        return fail(nullptr, DCA_E_INVALID, "dca_adv_moments_static: device index beyond the per-device address table");
    double *scratch = nullptr;
    e = adv_static_scratch(device, &scratch);
    if (e != cudaSuccess || !scratch)
        return fail(nullptr, DCA_E_CUDA, std::string("dca_adv_moments_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
    cudaStream_t st = (cudaStream_t)stream;
    // the array keeps the previous call's partials; every slot is rewritten by k_adv_partial, cleared first all the same
    e = cudaMemsetAsync(scratch, 0, sizeof(double) * kAdvBlocks * 3, st);
    if (e != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_adv_moments_static: memset: ") + cudaGetErrorString(e));
    k_adv_partial<<<kAdvBlocks, kAdvThreads, 0, st>>>(ret, value, valid, valid ? valid_group : 1, (size_t)n, scratch);
    k_adv_final<<<1, kAdvThreads, 0, st>>>(scratch, stats);
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_adv_moments_static: ") + cudaGetErrorString(e));
    return DCA_OK;
}

// ------------------------------------------------------------------------------------------------ dca_multiteam
namespace {
struct MtField { size_t offset, bytes; };
// carve the SoA state out of one slab; returns total bytes
size_t mt_carve_state(DcaMtState *st, char *base, int E, int N) {
    size_t off = 0;
    auto take = [&](size_t bytes) { char *p = base ? base + off : nullptr; off += (bytes + 255) & ~(size_t)255; return (void *)p; };
    const size_t EN = (size_t)E * N;
    st->pos = (double *)take(EN * 16); st->vel = (double *)take(EN * 16);
    st->atk_rad = (double *)take(EN * 8); st->terrain = (double *)take(EN * 8);
    st->alive = (uint8_t *)take(EN); st->just_died = (uint8_t *)take(EN); st->hit = (uint8_t *)take(EN);
    st->was_hit = (uint8_t *)take(EN); st->bsa = (uint8_t *)take(EN);
    st->init_theta = (double *)take((size_t)E * 8); st->time_step = (int32_t *)take((size_t)E * 4);
    st->obs_counter = (uint32_t *)take((size_t)E * 4); st->episode = (uint32_t *)take((size_t)E * 4);
    return off;
}
}  // namespace

int dca_mt_create(const DcaMtConfig *cfg, int n_env, int device, DcaMtHandle **out) {
    if (!cfg || !out || n_env <= 0) return fail(nullptr, DCA_E_INVALID, "dca_mt_create: null argument or n_env <= 0");
    *out = nullptr;
    if (cfg->n_team < 2 || cfg->n_team > DCA_MT_MAX_TEAMS)
        return fail(nullptr, DCA_E_INVALID, "dca_mt_create: need 2 <= n_team <= 8");
    int N = 0;
    for (int t = 0; t < cfg->n_team; ++t) {
        // the reference observation takes argsort()[:6] of the ally list and of the foe list: teams of >= 6
        if (cfg->team_size[t] < kVisN) return fail(nullptr, DCA_E_INVALID, "dca_mt_create: every team needs >= 6 agents");
        N += cfg->team_size[t];
    }
    if (N > 252 || (N * kMtLen) % 4 != 0) return fail(nullptr, DCA_E_INVALID, "dca_mt_create: at most 252 agents");
    if (cfg->steps_to_narrow <= 0 || cfg->max_episode_step <= 0)
        return fail(nullptr, DCA_E_INVALID, "dca_mt_create: StepsToNarrowCircle and MaxEpisodeStep must be positive");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(nullptr, DCA_E_NODEVICE, "dca_mt_create: no CUDA device (this library has no CPU fallback)");
    if (device < 0 || device >= ndev) return fail(nullptr, DCA_E_INVALID, "dca_mt_create: bad device index");
    DeviceGuard guard(device);
    DcaMtHandle *h = new (std::nothrow) DcaMtHandle();
    if (!h) return fail(nullptr, DCA_E_INVALID, "dca_mt_create: out of host memory");
    std::memset(&h->P, 0, sizeof(h->P));
    h->device = device;
    h->bound = false;
    h->P.cfg = *cfg;
    h->P.E = n_env;
    h->P.N = N;
    {
        const double t = std::ceil(cfg->random_jam_prob * 4294967296.0);
        h->P.jam_thresh = t <= 0 ? 0ull : (t >= 4294967296.0 ? 4294967296ull : (unsigned long long)t);
    }
    h->P.smax[0] = sq_limit(2.0f);  // obs_range envs/collective_assault_env.py:148
    h->P.smax[1] = sq_limit(3.0f);  // obs_range * 1.5 :204
    {
        int longest = 0, smallest = N;
        for (int t = 0; t < cfg->n_team; ++t) {
            longest = cfg->team_size[t] > longest ? cfg->team_size[t] : longest;
            smallest = cfg->team_size[t] < smallest ? cfg->team_size[t] : smallest;
        }
        longest = (N - smallest) > longest ? (N - smallest) : longest;  // ally list / foe list
        int bits = 3;
        while ((1 << bits) < longest) ++bits;
        h->P.idx_bits = bits;
    }
    h->smem = (int)mt_smem_bytes(N);
    DcaMtState probe;
    const size_t bytes = mt_carve_state(&probe, nullptr, n_env, N);
    cudaError_t e = cudaMalloc(&h->slab, bytes);
    if (e != cudaSuccess) { delete h; return fail(nullptr, DCA_E_CUDA, std::string("dca_mt_create: cudaMalloc: ") + cudaGetErrorString(e)); }
    e = cudaMemset(h->slab, 0, bytes);
    mt_carve_state(&h->P.st, (char *)h->slab, n_env, N);
    {   // per-function setting shared by every handle: always the largest admissible env (252 agents)
        const int max_smem = (int)mt_smem_bytes(252);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_mt_step, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_mt_reset, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
    }
    if (e == cudaSuccess) {
        k_mt_init<<<(n_env + 127) / 128, 128>>>(h->P);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { cudaFree(h->slab); delete h; return fail(nullptr, DCA_E_CUDA, std::string("dca_mt_create: ") + cudaGetErrorString(e)); }
    *out = h;
    return DCA_OK;
}

int dca_mt_destroy(DcaMtHandle *h) {
    if (!h) return DCA_E_INVALID;
    DeviceGuard guard(h->device);
    cudaFree(h->slab);
    delete h;
    return DCA_OK;
}

int dca_mt_bind_buffers(DcaMtHandle *h, const DcaMtBuffers *b) {
    if (!h || !b) return DCA_E_INVALID;
    if (!b->obs || !b->reward || !b->done || !b->ranking) {
        g_err = "dca_mt_bind_buffers: obs, reward, done and ranking are required";
        return DCA_E_INVALID;
    }
    if (((uintptr_t)b->obs | (uintptr_t)b->tobs) & 15) { g_err = "dca_mt_bind_buffers: obs / tobs must be 16-byte aligned"; return DCA_E_INVALID; }
    h->P.out = *b;
    h->bound = true;
    return DCA_OK;
}

static int mt_launch_check(const char *what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
    return DCA_OK;
}

int dca_mt_reset(DcaMtHandle *h, const uint8_t *env_mask, dca_stream_t stream) {
    if (!h) return DCA_E_INVALID;
    if (!h->bound) return fail(nullptr, DCA_E_UNBOUND, "dca_mt_reset: call dca_mt_bind_buffers first");
    DeviceGuard guard(h->device);
    k_mt_reset<<<h->P.E, kMtThreads, h->smem, (cudaStream_t)stream>>>(h->P, env_mask);
    return mt_launch_check("dca_mt_reset");
}

int dca_mt_step(DcaMtHandle *h, const int32_t *actions, const uint8_t *skip, int autoreset, dca_stream_t stream) {
    if (!h || !actions) return fail(nullptr, DCA_E_INVALID, "dca_mt_step: null argument");
    if (!h->bound) return fail(nullptr, DCA_E_UNBOUND, "dca_mt_step: call dca_mt_bind_buffers first");
    DeviceGuard guard(h->device);
    k_mt_step<<<h->P.E, kMtThreads, h->smem, (cudaStream_t)stream>>>(h->P, actions, skip, autoreset);
    return mt_launch_check("dca_mt_step");
}

static int mt_state_io(DcaMtHandle *h, const DcaMtState *u, bool import, dca_stream_t stream) {
    if (!h || !u) return DCA_E_INVALID;
    DeviceGuard guard(h->device);
    const size_t EN = (size_t)h->P.E * h->P.N, E = (size_t)h->P.E;
    const DcaMtState &m = h->P.st;
    struct { void *mine; void *theirs; size_t bytes; } f[] = {
        {m.pos, u->pos, EN * 16}, {m.vel, u->vel, EN * 16}, {m.atk_rad, u->atk_rad, EN * 8}, {m.terrain, u->terrain, EN * 8},
        {m.alive, u->alive, EN}, {m.just_died, u->just_died, EN}, {m.hit, u->hit, EN}, {m.was_hit, u->was_hit, EN},
        {m.bsa, u->bsa, EN}, {m.init_theta, u->init_theta, E * 8}, {m.time_step, u->time_step, E * 4},
        {m.obs_counter, u->obs_counter, E * 4}, {m.episode, u->episode, E * 4}};
    for (auto &x : f) {
        if (!x.theirs) return fail(nullptr, DCA_E_INVALID, "dca_mt state io: every DcaMtState field is required");
        cudaError_t e = cudaMemcpyAsync(import ? x.mine : x.theirs, import ? x.theirs : x.mine, x.bytes,
                                        cudaMemcpyDeviceToDevice, (cudaStream_t)stream);
        if (e != cudaSuccess) return fail(nullptr, DCA_E_CUDA, std::string("dca_mt state io: ") + cudaGetErrorString(e));
    }
    return DCA_OK;
}
int dca_mt_debug_phase_clocks(DcaMtHandle *h, uint64_t *clocks) {
    if (!h) return DCA_E_INVALID;
    h->P.prof = (unsigned long long *)clocks;
    return DCA_OK;
}
int dca_mt_export_state(DcaMtHandle *h, const DcaMtState *dst, dca_stream_t stream) { return mt_state_io(h, dst, false, stream); }
int dca_mt_import_state(DcaMtHandle *h, const DcaMtState *src, dca_stream_t stream) { return mt_state_io(h, src, true, stream); }

int dca_set_option(DcaHandle *h, int option, int value) {
    if (!h) return DCA_E_INVALID;
    if (option == DCA_OPT_EMPTY_ROWS_NAN) { h->P.empty_nan = value != 0; return DCA_OK; }
    if (option == DCA_OPT_OVERLAP_PLANNER) { h->overlap_planner = value != 0; return DCA_OK; }
    if (option == DCA_OPT_PIPELINE_STEPS) {
        if ((value != 0) != h->pipeline_steps) {  // (re)start the epochs from zero: nothing is published while it is off
            DeviceGuard guard(h->device);
            CUDA_TRY(h, cudaDeviceSynchronize());
            CUDA_TRY(h, cudaMemset(h->epoch, 0, sizeof(uint32_t) * 2 * (size_t)h->P.E));
            h->step_epoch = h->plan_epoch = 0;
            h->last_was_step = false;
            h->planned_force = nullptr;
            h->pipeline_steps = value != 0;
        }
        return DCA_OK;
    }
    return fail(h, DCA_E_INVALID, "dca_set_option: unknown option");
}

int dca_debug_phase_clocks(DcaHandle *h, uint64_t *clocks) {
    if (!h) return DCA_E_INVALID;
    h->P.prof = (unsigned long long *)clocks;
    return DCA_OK;
}

#ifdef DCA_DIAG_COUNT
int dca_debug_counters(unsigned long long *out8, int reset) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out8, dca::g_diag, sizeof(unsigned long long) * 8);
    if (reset) { unsigned long long z[8] = {0}; cudaMemcpyToSymbol(dca::g_diag, z, sizeof(z)); }
    return 0;
}
#endif

int dca_kernel_launches(const DcaHandle *h, int64_t *count) {
    if (!h || !count) return DCA_E_INVALID;
    *count = h->launches;
    return DCA_OK;
}

}  // extern "C"
