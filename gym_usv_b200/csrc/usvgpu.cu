// usvgpu.cu -- host side of libusvgpu.so: the handle and the C ABI of include/usvgpu.h.
//
// The device code lives in the headers next to this file:
//   usv_model.cuh    one ASMC + 3-DOF sub-step (gym_usv/control/usv_asmc.py:53-244)
//   usv_state.cuh    SoA state layout          usv_philox.cuh  counter-based reset RNG
//   usv_reset.cuh    Env.reset()               usv_sensor.cuh  warp-cooperative range sensors
//   usv_kernels.cuh  dynamics_kernel (thread per env) + sensor_kernel (warp per env) = one VecEnv.step()
// No tensor cores: nothing on this path is a contraction.  Reference citations are relative to the
// reference root (romi2002/gym-usv).
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <new>
#include <string>

#include "../../include/usvgpu.h"
#include "usv_kernels.cuh"
#include "usv_legacy.cuh"
#include "usv_ca.cuh"
#include "usv_render.cuh"
#include "usv_info.cuh"

// =================================================================================================
// Host side: handle + C ABI
// =================================================================================================
using namespace usv;

constexpr int HOST_CHUNKS_MAX = 16;

struct UsvGpu {
    UsvConfig cfg;
    int device;
    int obs_dim, act_dim;  // 143 / 2 ('usv-asmc-simple') or 6 / 1 ('usv-asmc-v0')
    bool legacy;
    bool ca;               // env_kind == USV_ENV_ASMC_CA_V0: sp.dyn + cp.ca + cp.n_obs, the rest of sp as for legacy
    CaPtrs cp;
    StatePtrs sp;
    LegacyPtrs lp;         // env_kind == USV_ENV_ASMC_V0: sp holds only n, stats, ep_return, steps, episode, ep_done
    size_t obst_elem;
    std::string err;
    unsigned long long host_env_steps;
    unsigned long long launches;  // kernels launched by step / rollout so far
    int sm_count;
    // device staging for the host-buffer face (usvgpu_step_host), allocated on first use
    float* d_actions;
    float* d_obs;
    float* d_tobs;
    float* d_reward;
    uint8_t* d_flags;   // terminated[N] then truncated[N]
    int32_t* h_idx;     // page-locked, mapped: gather list of usvgpu_terminal_obs_host
    float* h_rows;      // page-locked, mapped: gathered rows (the gather kernel stores them over PCIe)
    double2* h_ep;      // page-locked, mapped: (return, length) of the finished episodes
    int64_t gather_cap;
    cudaStream_t side;      // the gather runs here, beside the observation copies
    cudaStream_t copy;      // observation chunks go home on this stream while the caller's stream computes the next chunk
    cudaEvent_t ev_chunk[HOST_CHUNKS_MAX];
    cudaEvent_t ev_copied;
    int host_chunks;
    cudaEvent_t ev_flags;   // reward / flags of the last usvgpu_step_host have reached the host
    bool own_staging;   // false after usvgpu_use_staging(): the caller's device buffers are used instead
};

static thread_local std::string g_create_err;
static double2* g_raytab[64] = {nullptr};  // per device ray table for the stand-alone sensors entry point

// Every ABI entry point makes the env's device current for its own duration and puts the caller's device back on
// the way out: a process that drives several GPUs (or whose torch default device differs from the env's) keeps its
// current device across calls into this library.
struct DeviceGuard {
    int prev = -1;
    bool switched = false;
    cudaError_t err = cudaSuccess;
    explicit DeviceGuard(int device) {
        err = cudaGetDevice(&prev);
        if (err == cudaSuccess && prev != device) {
            err = cudaSetDevice(device);
            switched = err == cudaSuccess;
        }
    }
    ~DeviceGuard() {
        if (switched) cudaSetDevice(prev);
    }
    DeviceGuard(const DeviceGuard&) = delete;
    DeviceGuard& operator=(const DeviceGuard&) = delete;
};

static int fail(UsvGpu* h, int code, const std::string& msg) {
    if (h) h->err = msg;
    else g_create_err = msg;
    return code;
}
#define CU_TRY(h, expr)                                                                        \
    do {                                                                                       \
        cudaError_t e_ = (expr);                                                               \
        if (e_ != cudaSuccess)                                                                 \
            return fail(h, USV_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); \
    } while (0)

static void host_raytab(double2* t) {
    // usv_asmc_ca_env.py:420-421: -np.pi * 2 / 3 + np.arange(S) * lidar_resolution
    for (int k = 0; k < N_SENS; ++k) {
        const double phi = -PI * 2 / 3 + (double)k * SENSOR_PITCH;
        t[k] = make_double2(cos(phi), sin(phi));
    }
}

static int check_device(UsvGpu* h, int device) {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(h, USV_ERR_NO_DEVICE, std::string("no CUDA device (there is no CPU fallback): ") +
                                              (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0"));
    if (device < 0 || device >= count) return fail(h, USV_ERR_INVALID, "device index out of range");
    return USV_OK;
}

// <<<grid, block, 0, st>>>, or the same launch with programmatic stream serialisation allowed (USV_PDL, usv_state.cuh)
template <typename... KArgs, typename... Args>
static void launch_k(void (*kern)(KArgs...), unsigned grid, unsigned block, cudaStream_t st, Args... args) {
#if USV_PDL
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
#else
    kern<<<grid, block, 0, st>>>(KArgs(args)...);
#endif
}

template <bool OBST, typename ObT>
static void launch_sensor(UsvGpu* h, const StepFlags& fl, float* obs, float* reward, uint8_t* term, uint8_t* trunc,
                          float* tobs, cudaStream_t st) {
    const long long need = (fl.hi - fl.lo + SENS_WARPS - 1) / SENS_WARPS;
    const long long resident = (long long)h->sm_count * SENS_MIN_CTAS;  // persistent: one wave, warps loop over envs
    const unsigned blocks = (unsigned)(need < resident ? need : resident);
    launch_k(sensor_kernel<OBST, ObT>, blocks, SENS_WARPS * 32, st, h->sp, fl, obs, reward, term, trunc, tobs);
}

// One VecEnv.step(): dynamics_kernel (thread per env) then sensor_kernel (warp per env), stream-ordered.
static int step_impl(UsvGpu* h, const float* actions, float* obs, float* reward, uint8_t* term, uint8_t* trunc,
                     float* tobs, cudaStream_t st, cudaEvent_t mid = nullptr, long long lo = 0, long long hi = -1) {
    if (hi < 0) hi = h->sp.n;
    const StepFlags fl = {h->cfg.do_perturb, h->cfg.skip_kinematic, h->cfg.max_episode_steps, h->cfg.auto_reset, lo, hi};
    if (h->ca) {  // 'usv-asmc-ca-v0': one kernel, thread per env
        launch_k(ca_kernel<false>, (unsigned)((hi - lo + CA_B - 1) / CA_B), CA_B, st, h->cp, fl, (const uint8_t*)nullptr, actions,
                 obs, reward, term, trunc, tobs, UsvCaResetOptions{});
        cudaError_t ce = cudaGetLastError();
        if (ce != cudaSuccess) return fail(h, USV_ERR_CUDA, std::string("ca step launch: ") + cudaGetErrorString(ce));
        if (mid) cudaEventRecord(mid, st);
        h->host_env_steps += (unsigned long long)(hi - lo);
        h->launches += 1;
        return USV_OK;
    }
    if (h->legacy) {  // 'usv-asmc-v0' and its float64 siblings: one kernel, thread per env
        const unsigned lblocks = (unsigned)((hi - lo + LEG_B - 1) / LEG_B);
        if (h->lp.variant == LEG_YE_INT)
            launch_k(legacy64_step_kernel<LEG_YE_INT>, lblocks, LEG_B, st, h->lp, fl, actions, obs, reward, term, trunc, tobs);
        else if (h->lp.variant == LEG_PID)
            launch_k(legacy64_step_kernel<LEG_PID>, lblocks, LEG_B, st, h->lp, fl, actions, obs, reward, term, trunc, tobs);
        else
            launch_k(legacy_step_kernel, lblocks, LEG_B, st, h->lp, fl, actions, obs, reward, term, trunc, tobs);
        cudaError_t le = cudaGetLastError();
        if (le != cudaSuccess) return fail(h, USV_ERR_CUDA, std::string("legacy step launch: ") + cudaGetErrorString(le));
        if (mid) cudaEventRecord(mid, st);
        h->host_env_steps += (unsigned long long)(hi - lo);
        h->launches += 1;
        return USV_OK;
    }
    const unsigned dblocks = (unsigned)((hi - lo + DYN_B - 1) / DYN_B);
    if (h->cfg.env_kind == USV_ENV_SIMPLE)
        launch_k(kinematic_kernel, (unsigned)((hi - lo + KIN_B - 1) / KIN_B), KIN_B, st, h->sp, fl, actions, reward, term, trunc);
    else if (h->cfg.do_perturb) launch_k(dynamics_kernel<true>, dblocks, DYN_B, st, h->sp, fl, actions, reward, term, trunc);
    else launch_k(dynamics_kernel<false>, dblocks, DYN_B, st, h->sp, fl, actions, reward, term, trunc);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(h, USV_ERR_CUDA, std::string("dynamics launch: ") + cudaGetErrorString(e));
    if (mid) cudaEventRecord(mid, st);
    const bool f32 = h->cfg.obstacles_f32 != 0;
    if (h->cfg.ignore_obstacles) {
        if (f32) launch_sensor<false, float>(h, fl, obs, reward, term, trunc, tobs, st);
        else launch_sensor<false, double>(h, fl, obs, reward, term, trunc, tobs, st);
    } else {
        if (f32) launch_sensor<true, float>(h, fl, obs, reward, term, trunc, tobs, st);
        else launch_sensor<true, double>(h, fl, obs, reward, term, trunc, tobs, st);
    }
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(h, USV_ERR_CUDA, std::string("sensor launch: ") + cudaGetErrorString(e));
    h->host_env_steps += (unsigned long long)(hi - lo);
    h->launches += 2;
    return USV_OK;
}

static bool bad_alignment(const UsvGpu* h, const void* obs, const void* actions) {
    return h->legacy ? (((uintptr_t)obs & 7) || ((uintptr_t)actions & 3)) : (((uintptr_t)obs & 3) || ((uintptr_t)actions & 7));
}

static int field_ptr(UsvGpu* h, int32_t field, void** p, int64_t* bytes) {
    const int64_t n = h->sp.n;
    switch (field) {
        case USV_FIELD_LEGACY: if (!h->lp.st) return -1; *p = h->lp.st; *bytes = n * LEG_ROWS * 4; return 0;
        case USV_FIELD_LEGACY_TARGET: if (!h->legacy) return -1; *p = h->lp.target; *bytes = n * 6 * 8; return 0;
        case USV_FIELD_LEGACY_POS0: if (!h->lp.pos0) return -1; *p = h->lp.pos0; *bytes = n * 3 * 8; return 0;
        case USV_FIELD_LEGACY64: if (!h->lp.st64) return -1; *p = h->lp.st64; *bytes = n * LEG64_ROWS * 8; return 0;
        case USV_FIELD_CA: if (!h->ca) return -1; *p = h->cp.ca; *bytes = n * CA_ROWS * 8; return 0;
        case USV_FIELD_CA_N_OBS: if (!h->ca) return -1; *p = h->cp.n_obs; *bytes = n * 4; return 0;
        case USV_FIELD_DYN: if (h->legacy) return -1; *p = h->sp.dyn; *bytes = n * N_DYN * 8; return 0;
        case USV_FIELD_EPI: if (h->legacy || h->ca) return -1; *p = h->sp.epi; *bytes = n * N_EPI * 8; return 0;
        // these two live inside the interleaved sensor-input rows: *p is the first element, copies are pitched
        case USV_FIELD_OBSTACLES: if (h->legacy || h->ca) return -1; *p = h->sp.sin; *bytes = n * MAX_OBST * 3 * (int64_t)h->obst_elem; return 0;
        case USV_FIELD_N_OBS: if (h->legacy || h->ca) return -1; *p = h->sp.sin + h->sp.sin_tail + sizeof(Hand) + HDR_PAD * sizeof(float); *bytes = n * 4; return 0;
        case USV_FIELD_TARGET0: if (h->legacy || h->ca) return -1; *p = h->sp.sin + h->sp.sin_tail + sizeof(Hand) + HDR_PAD * sizeof(float) + sizeof(int); *bytes = n * 8; return 0;
        case USV_FIELD_STEPS: *p = h->sp.steps; *bytes = n * 4; return 0;
        case USV_FIELD_EPISODE: *p = h->sp.episode; *bytes = n * 4; return 0;
        case USV_FIELD_EP_RETURN: *p = h->sp.ep_return; *bytes = n * 8; return 0;
        default: return -1;
    }
}

extern "C" {

int usvgpu_version(void) { return USVGPU_VERSION; }

#ifndef USVGPU_SOURCE_HASH
#define USVGPU_SOURCE_HASH "unknown"
#endif
const char* usvgpu_source_hash(void) { return USVGPU_SOURCE_HASH; }

const char* usvgpu_last_error(usvgpu_handle h) { return h ? h->err.c_str() : g_create_err.c_str(); }

int usvgpu_create(const UsvConfig* cfg, int device, int64_t n_envs, int64_t global_env_offset, uint64_t seed,
                  usvgpu_handle* out) {
    if (!cfg || !out) return fail(nullptr, USV_ERR_INVALID, "cfg/out is NULL");
    *out = nullptr;
    if (n_envs <= 0) return fail(nullptr, USV_ERR_INVALID, "n_envs must be positive");
    if (n_envs >= (1LL << 31)) return fail(nullptr, USV_ERR_INVALID, "n_envs must be below 2^31 (the kernels index envs with 32 bits)");
    if (cfg->env_kind < USV_ENV_ASMC_SIMPLE || cfg->env_kind > USV_ENV_ASMC_CA_V0)
        return fail(nullptr, USV_ERR_INVALID, "unknown env_kind");
    if (cfg->precision != USV_PREC_F64) return fail(nullptr, USV_ERR_INVALID, "precision not supported by this build");
    int rc = check_device(nullptr, device);
    if (rc) return rc;
    UsvGpu* h = new (std::nothrow) UsvGpu();
    if (!h) return fail(nullptr, USV_ERR_ALLOC, "host allocation failed");
    h->cfg = *cfg;
    h->device = device;
    const int variant = cfg->env_kind == USV_ENV_ASMC_YE_INT_V0 ? LEG_YE_INT : cfg->env_kind == USV_ENV_PID_V0 ? LEG_PID : LEG_ASMC_V0;
    h->legacy = cfg->env_kind == USV_ENV_ASMC_V0 || variant != LEG_ASMC_V0;
    h->ca = cfg->env_kind == USV_ENV_ASMC_CA_V0;
    h->obs_dim = h->legacy ? LEG_OBS : h->ca ? CA_OBS : OBS_DIM;
    h->act_dim = h->legacy ? 1 : 2;
    memset(&h->lp, 0, sizeof(h->lp));
    memset(&h->cp, 0, sizeof(h->cp));
    h->host_env_steps = 0;
    h->launches = 0;
    h->obst_elem = cfg->obstacles_f32 ? sizeof(float) : sizeof(double);
    h->sm_count = 148;
    h->d_actions = h->d_obs = h->d_tobs = h->d_reward = h->h_rows = nullptr;
    h->d_flags = nullptr;
    h->h_idx = nullptr;
    h->h_ep = nullptr;
    h->gather_cap = 0;
    h->side = nullptr;
    h->copy = nullptr;
    h->ev_flags = nullptr;
    h->ev_copied = nullptr;
    for (int c = 0; c < HOST_CHUNKS_MAX; ++c) h->ev_chunk[c] = nullptr;
    h->host_chunks = 2;
    h->own_staging = true;
    cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, device);
    memset(&h->sp, 0, sizeof(h->sp));
    h->sp.n = n_envs;
    h->sp.global_offset = global_env_offset;
    h->sp.seed = seed;
    h->sp.kinematic = cfg->env_kind == USV_ENV_SIMPLE;
    const size_t n = (size_t)n_envs;
    DeviceGuard dev_guard(device);
    cudaError_t e = dev_guard.err;
    auto alloc = [&](void** p, size_t bytes) {
        if (e == cudaSuccess) e = cudaMalloc(p, bytes);
        if (e == cudaSuccess) e = cudaMemset(*p, 0, bytes);
    };
    if (h->legacy) {
        if (variant == LEG_ASMC_V0) {
            alloc((void**)&h->lp.st, n * LEG_ROWS * sizeof(float));
            alloc((void**)&h->lp.pos0, n * 3 * sizeof(double));
        } else {
            alloc((void**)&h->lp.st64, n * LEG64_ROWS * sizeof(double));
        }
        alloc((void**)&h->lp.target, n * 6 * sizeof(double));
    } else if (h->ca) {
        alloc((void**)&h->sp.dyn, n * N_DYN * sizeof(double));
        alloc((void**)&h->cp.ca, n * CA_ROWS * sizeof(double));
        alloc((void**)&h->cp.n_obs, n * sizeof(int));
        double2 t[CA_SENS];
        for (int k = 0; k < CA_SENS; ++k) {  // usv_asmc_ca_env.py:47-49, 420-421
            const double phi = -PI * 2 / 3 + (double)k * (((2.0 / 3) * (2 * PI)) / CA_SENS);
            t[k] = make_double2(cos(phi), sin(phi));
        }
        if (e == cudaSuccess) e = cudaMemcpyToSymbol(CA_RAYTAB, t, sizeof(t));
    } else {
        alloc((void**)&h->sp.dyn, n * N_DYN * sizeof(double));
        alloc((void**)&h->sp.epi, n * N_EPI * sizeof(double));
        h->sp.sin_tail = (int)(MAX_OBST * 3 * h->obst_elem);
        h->sp.sin_row = h->sp.sin_tail + SIN_TAIL_BYTES;
        alloc((void**)&h->sp.sin, n * (size_t)h->sp.sin_row);
    }
    alloc((void**)&h->sp.steps, n * sizeof(int));
    alloc((void**)&h->sp.episode, n * sizeof(int));
    alloc((void**)&h->sp.ep_return, n * sizeof(double));
    alloc((void**)&h->sp.ep_done, n * sizeof(double2));
    alloc((void**)&h->sp.stats, sizeof(DeviceStats));
    double2* rt = nullptr;
    alloc((void**)&rt, N_SENS * sizeof(double2));
    h->sp.raytab = rt;
    if (e == cudaSuccess) e = cudaMemset(h->sp.episode, 0xFF, n * sizeof(int));  // episode = -1 before the first reset
    if (e == cudaSuccess) {
        double2 t[N_SENS];
        host_raytab(t);
        e = cudaMemcpy(rt, t, sizeof(t), cudaMemcpyHostToDevice);
    }
    if (e != cudaSuccess) {
        std::string msg = std::string("allocating env state: ") + cudaGetErrorString(e);
        usvgpu_destroy(h);
        return fail(nullptr, USV_ERR_CUDA, msg);
    }
    h->lp.steps = h->sp.steps;
    h->lp.episode = h->sp.episode;
    h->lp.ep_return = h->sp.ep_return;
    h->lp.ep_done = h->sp.ep_done;
    h->lp.stats = h->sp.stats;
    h->lp.n = n_envs;
    h->lp.global_offset = global_env_offset;
    h->lp.seed = seed;
    h->lp.variant = variant;
    h->cp.dyn = h->sp.dyn;
    h->cp.steps = h->sp.steps;
    h->cp.episode = h->sp.episode;
    h->cp.ep_return = h->sp.ep_return;
    h->cp.ep_done = h->sp.ep_done;
    h->cp.stats = h->sp.stats;
    h->cp.n = n_envs;
    h->cp.global_offset = global_env_offset;
    h->cp.seed = seed;
    *out = h;
    return USV_OK;
}

int usvgpu_obs_dim(usvgpu_handle h) { return h ? h->obs_dim : USV_ERR_INVALID; }
int usvgpu_act_dim(usvgpu_handle h) { return h ? h->act_dim : USV_ERR_INVALID; }

int usvgpu_destroy(usvgpu_handle h) {
    if (!h) return USV_OK;
    DeviceGuard dev_guard(h->device);
    cudaFree(h->lp.st);
    cudaFree(h->lp.pos0);
    cudaFree(h->lp.st64);
    cudaFree(h->lp.target);
    cudaFree(h->cp.ca);
    cudaFree(h->cp.n_obs);
    cudaFree(h->sp.dyn);
    cudaFree(h->sp.epi);
    cudaFree(h->sp.sin);
    cudaFree(h->sp.steps);
    cudaFree(h->sp.episode);
    cudaFree(h->sp.ep_return);
    cudaFree(h->sp.ep_done);
    cudaFree(h->sp.stats);
    cudaFree((void*)h->sp.raytab);
    if (h->own_staging) {
        cudaFree(h->d_actions);
        cudaFree(h->d_obs);
        cudaFree(h->d_tobs);
        cudaFree(h->d_reward);
        cudaFree(h->d_flags);
    }
    if (h->h_idx) cudaFreeHost(h->h_idx);
    if (h->h_rows) cudaFreeHost(h->h_rows);
    if (h->h_ep) cudaFreeHost(h->h_ep);
    if (h->side) cudaStreamDestroy(h->side);
    if (h->copy) cudaStreamDestroy(h->copy);
    if (h->ev_flags) cudaEventDestroy(h->ev_flags);
    if (h->ev_copied) cudaEventDestroy(h->ev_copied);
    for (int c = 0; c < HOST_CHUNKS_MAX; ++c)
        if (h->ev_chunk[c]) cudaEventDestroy(h->ev_chunk[c]);
    delete h;
    return USV_OK;
}

int usvgpu_reset(usvgpu_handle h, const uint8_t* mask, float* obs, void* stream) {
    if (!h) return USV_ERR_INVALID;
    DeviceGuard dev_guard(h->device);
    CU_TRY(h, dev_guard.err);
    const int threads = 128;
    const unsigned blocks = (unsigned)((h->sp.n + threads - 1) / threads);
    cudaStream_t st = (cudaStream_t)stream;
    if (h->ca) {
        const StepFlags fl = {0, 0, h->cfg.max_episode_steps, h->cfg.auto_reset, 0, h->sp.n};
        ca_kernel<true><<<(unsigned)((h->sp.n + CA_B - 1) / CA_B), CA_B, 0, st>>>(h->cp, fl, mask, nullptr, obs, nullptr, nullptr,
                                                                                 nullptr, nullptr, UsvCaResetOptions{});
    } else if (h->legacy) legacy_reset_kernel<<<blocks, threads, 0, st>>>(h->lp, mask, obs);
    else if (h->cfg.obstacles_f32) reset_kernel<float><<<blocks, threads, 0, st>>>(h->sp, mask, obs);
    else reset_kernel<double><<<blocks, threads, 0, st>>>(h->sp, mask, obs);
    CU_TRY(h, cudaGetLastError());
    return USV_OK;
}

int usvgpu_reset_ca_options(usvgpu_handle h, const uint8_t* mask, const UsvCaResetOptions* options, float* obs,
                            void* stream) {
    if (!h) return USV_ERR_INVALID;
    if (!h->ca) return fail(h, USV_ERR_INVALID, "usvgpu_reset_ca_options: the handle is not a usv-asmc-ca-v0 env");
    if (!options) return fail(h, USV_ERR_INVALID, "usvgpu_reset_ca_options: NULL options (usvgpu_reset is the plain reset)");
    if (options->has_obstacles && (options->n_obstacles < 0 || options->n_obstacles > USVGPU_CA_MAX_OBSTACLES))
        return fail(h, USV_ERR_INVALID, "usvgpu_reset_ca_options: n_obstacles outside 0..USVGPU_CA_MAX_OBSTACLES");
    DeviceGuard dev_guard(h->device);
    CU_TRY(h, dev_guard.err);
    const StepFlags fl = {0, 0, h->cfg.max_episode_steps, h->cfg.auto_reset, 0, h->sp.n};
    ca_kernel<true><<<(unsigned)((h->sp.n + CA_B - 1) / CA_B), CA_B, 0, (cudaStream_t)stream>>>(
        h->cp, fl, mask, nullptr, obs, nullptr, nullptr, nullptr, nullptr, *options);
    CU_TRY(h, cudaGetLastError());
    return USV_OK;
}

int usvgpu_step(usvgpu_handle h, const float* actions, float* obs, float* reward, uint8_t* terminated,
                uint8_t* truncated, float* terminal_obs, void* stream) {
    if (!h) return USV_ERR_INVALID;
    if (!actions || !obs || !reward || !terminated || !truncated) return fail(h, USV_ERR_INVALID, "NULL buffer");
    if (bad_alignment(h, obs, actions)) return fail(h, USV_ERR_INVALID, "obs / actions misaligned (simple: 4 / 8 bytes, legacy: 8 / 4 bytes)");
    DeviceGuard dev_guard(h->device);
    CU_TRY(h, dev_guard.err);
    return step_impl(h, actions, obs, reward, terminated, truncated, terminal_obs, (cudaStream_t)stream);
}

int usvgpu_step_timed(usvgpu_handle h, const float* actions, float* obs, float* reward, uint8_t* terminated,
                      uint8_t* truncated, float* terminal_obs, void* stream, float* kernel_ms) {
    if (!h) return USV_ERR_INVALID;
    if (!actions || !obs || !reward || !terminated || !truncated || !kernel_ms) return fail(h, USV_ERR_INVALID, "NULL buffer");
    if (bad_alignment(h, obs, actions)) return fail(h, USV_ERR_INVALID, "obs / actions misaligned (simple: 4 / 8 bytes, legacy: 8 / 4 bytes)");
    DeviceGuard dev_guard(h->device);
    CU_TRY(h, dev_guard.err);
    cudaStream_t st = (cudaStream_t)stream;
    cudaEvent_t ev[3];
    for (int k = 0; k < 3; ++k) CU_TRY(h, cudaEventCreate(&ev[k]));
    CU_TRY(h, cudaEventRecord(ev[0], st));
    int rc = step_impl(h, actions, obs, reward, terminated, truncated, terminal_obs, st, ev[1]);
    if (rc == USV_OK) {
        CU_TRY(h, cudaEventRecord(ev[2], st));
        CU_TRY(h, cudaEventSynchronize(ev[2]));
        CU_TRY(h, cudaEventElapsedTime(&kernel_ms[0], ev[0], ev[1]));
        CU_TRY(h, cudaEventElapsedTime(&kernel_ms[1], ev[1], ev[2]));
    }
    for (int k = 0; k < 3; ++k) cudaEventDestroy(ev[k]);
    return rc;
}

int usvgpu_rollout(usvgpu_handle h, const float* actions, int32_t n_steps, int32_t action_wrap, float* obs,
                   float* reward, uint8_t* terminated, uint8_t* truncated, void* stream) {
    if (!h) return USV_ERR_INVALID;
    if (!actions || !obs || !reward || !terminated || !truncated || n_steps < 0)
        return fail(h, USV_ERR_INVALID, "bad rollout arguments");
    if (bad_alignment(h, obs, actions)) return fail(h, USV_ERR_INVALID, "obs / actions misaligned (simple: 4 / 8 bytes, legacy: 8 / 4 bytes)");
    DeviceGuard dev_guard(h->device);
    CU_TRY(h, dev_guard.err);
    for (int k = 0; k < n_steps; ++k) {
        const int slab = action_wrap > 0 ? (k % action_wrap) : k;
        int rc = step_impl(h, actions + (size_t)slab * h->sp.n * h->act_dim, obs, reward, terminated, truncated, nullptr,
                           (cudaStream_t)stream);
        if (rc) return rc;
    }
    return USV_OK;
}

// ---- host-buffer face: what stock SB3 does with a VecEnv (numpy actions in, numpy observations out) ----------
static int ensure_staging(UsvGpu* h) {
    if (!h->side) CU_TRY(h, cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking));
    if (!h->ev_flags) CU_TRY(h, cudaEventCreateWithFlags(&h->ev_flags, cudaEventDisableTiming));
    if (!h->copy) CU_TRY(h, cudaStreamCreateWithFlags(&h->copy, cudaStreamNonBlocking));
    if (!h->ev_copied) CU_TRY(h, cudaEventCreateWithFlags(&h->ev_copied, cudaEventDisableTiming));
    for (int c = 0; c < HOST_CHUNKS_MAX; ++c)
        if (!h->ev_chunk[c]) CU_TRY(h, cudaEventCreateWithFlags(&h->ev_chunk[c], cudaEventDisableTiming));
    if (h->d_obs) return USV_OK;
    const size_t n = (size_t)h->sp.n;
    CU_TRY(h, cudaMalloc((void**)&h->d_actions, n * h->act_dim * sizeof(float)));
    CU_TRY(h, cudaMalloc((void**)&h->d_reward, n * sizeof(float)));
    CU_TRY(h, cudaMalloc((void**)&h->d_flags, 2 * n));
    CU_TRY(h, cudaMalloc((void**)&h->d_tobs, n * h->obs_dim * sizeof(float)));
    CU_TRY(h, cudaMemset(h->d_tobs, 0, n * h->obs_dim * sizeof(float)));
    CU_TRY(h, cudaMalloc((void**)&h->d_obs, n * h->obs_dim * sizeof(float)));
    CU_TRY(h, cudaMemset(h->d_obs, 0, n * h->obs_dim * sizeof(float)));
    return USV_OK;
}

int usvgpu_use_staging(usvgpu_handle h, float* actions, float* obs, float* reward, uint8_t* flags, float* terminal_obs) {
    if (!h) return USV_ERR_INVALID;
    if (!actions || !obs || !reward || !flags || !terminal_obs) return fail(h, USV_ERR_INVALID, "NULL staging buffer");
    if (h->d_obs && h->own_staging) {
        DeviceGuard dev_guard(h->device);
        cudaDeviceSynchronize();
        cudaFree(h->d_actions);
        cudaFree(h->d_obs);
        cudaFree(h->d_tobs);
        cudaFree(h->d_reward);
        cudaFree(h->d_flags);
    }
    h->own_staging = false;
    h->d_actions = actions;
    h->d_obs = obs;
    h->d_reward = reward;
    h->d_flags = flags;
    h->d_tobs = terminal_obs;
    return USV_OK;
}

int usvgpu_host_alloc(void** out, int64_t bytes) {
    if (!out || bytes <= 0) return fail(nullptr, USV_ERR_INVALID, "bad host_alloc arguments");
    CU_TRY(nullptr, cudaHostAlloc(out, (size_t)bytes, cudaHostAllocPortable | cudaHostAllocMapped));
    return USV_OK;
}

int usvgpu_host_free(void* p) {
    if (p) CU_TRY(nullptr, cudaFreeHost(p));
    return USV_OK;
}

int usvgpu_reset_host(usvgpu_handle h, const uint8_t* mask_host, float* obs_host, void* stream) {
    if (!h || !obs_host) return USV_ERR_INVALID;
    DeviceGuard dev_guard(h->device);
    CU_TRY(h, dev_guard.err);
    int rc = ensure_staging(h);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t n = (size_t)h->sp.n;
    const uint8_t* mask = nullptr;
    if (mask_host) {
        CU_TRY(h, cudaMemcpyAsync(h->d_flags, mask_host, n, cudaMemcpyHostToDevice, st));
        mask = h->d_flags;
    }
    rc = usvgpu_reset(h, mask, h->d_obs, stream);
    if (rc) return rc;
    CU_TRY(h, cudaMemcpyAsync(obs_host, h->d_obs, n * h->obs_dim * sizeof(float), cudaMemcpyDeviceToHost, st));
    CU_TRY(h, cudaStreamSynchronize(st));
    return USV_OK;
}

int usvgpu_step_host(usvgpu_handle h, const float* actions_host, float* obs_host, float* reward_host,
                     uint8_t* terminated_host, uint8_t* truncated_host, int32_t mapped_obs, void* stream) {
    if (!h) return USV_ERR_INVALID;
    if (!actions_host || !reward_host || !terminated_host || !truncated_host) return fail(h, USV_ERR_INVALID, "NULL buffer");
    if (!obs_host && mapped_obs) return fail(h, USV_ERR_INVALID, "mapped_obs needs a host observation buffer");
    DeviceGuard dev_guard(h->device);
    CU_TRY(h, dev_guard.err);
    int rc = ensure_staging(h);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t n = (size_t)h->sp.n;
    CU_TRY(h, cudaMemcpyAsync(h->d_actions, actions_host, n * h->act_dim * sizeof(float), cudaMemcpyHostToDevice, st));
    if (mapped_obs) {  // the sensor kernel stores the rows straight into the caller's page-locked, mapped buffer
        void* dp = nullptr;
        CU_TRY(h, cudaHostGetDevicePointer(&dp, obs_host, 0));
        rc = step_impl(h, h->d_actions, (float*)dp, h->d_reward, h->d_flags, h->d_flags + n, h->d_tobs, st);
        if (rc) return rc;
    } else {
        // Chunks of envs: while the copy engine takes chunk c's observation rows home (37.5 MB per 65 536 envs: the
        // PCIe link is what bounds this face), the SMs compute chunk c + 1.
        // The first chunk is small (1/8 of the shard) so that the copy engine starts early; the rest is split evenly.
        int chunks = h->host_chunks;
        if ((long long)n < 8192LL * chunks) chunks = 1;
        const long long first = chunks > 1 ? ((long long)n / 8 + 255) / 256 * 256 : (long long)n;
        const long long per = chunks > 1 ? (((long long)n - first + chunks - 2) / (chunks - 1) + 255) / 256 * 256 : 0;
        long long lo = 0;
        for (int c = 0; c < chunks && lo < (long long)n; ++c) {
            long long hi = lo + (c == 0 ? first : per);
            if (hi > (long long)n || c == chunks - 1) hi = (long long)n;
            rc = step_impl(h, h->d_actions, h->d_obs, h->d_reward, h->d_flags, h->d_flags + n, h->d_tobs, st, nullptr, lo, hi);
            if (rc) return rc;
            CU_TRY(h, cudaEventRecord(h->ev_chunk[c], st));
            CU_TRY(h, cudaStreamWaitEvent(h->copy, h->ev_chunk[c], 0));
            // the one device-to-host copy engine serves its queue in order: reward and flags of the whole shard go
            // just before the LAST chunk's rows, so the host has them (ev_flags) while those rows are in flight
            const size_t m = (size_t)(hi - lo);
            if (hi == (long long)n) {
                CU_TRY(h, cudaMemcpyAsync(reward_host, h->d_reward, n * sizeof(float), cudaMemcpyDeviceToHost, h->copy));
                CU_TRY(h, cudaMemcpyAsync(terminated_host, h->d_flags, n, cudaMemcpyDeviceToHost, h->copy));
                CU_TRY(h, cudaMemcpyAsync(truncated_host, h->d_flags + n, n, cudaMemcpyDeviceToHost, h->copy));
                CU_TRY(h, cudaEventRecord(h->ev_flags, h->copy));
            }
            if (obs_host)  // NULL: the rows stay on the device, only reward / flags go home
                CU_TRY(h, cudaMemcpyAsync(obs_host + lo * h->obs_dim, h->d_obs + lo * h->obs_dim, m * h->obs_dim * sizeof(float),
                                          cudaMemcpyDeviceToHost, h->copy));
            lo = hi;
        }
        CU_TRY(h, cudaEventRecord(h->ev_copied, h->copy));
    }
    if (mapped_obs) {
        CU_TRY(h, cudaMemcpyAsync(reward_host, h->d_reward, n * sizeof(float), cudaMemcpyDeviceToHost, st));
        CU_TRY(h, cudaMemcpyAsync(terminated_host, h->d_flags, n, cudaMemcpyDeviceToHost, st));
        CU_TRY(h, cudaMemcpyAsync(truncated_host, h->d_flags + n, n, cudaMemcpyDeviceToHost, st));
        CU_TRY(h, cudaEventRecord(h->ev_flags, st));
    }
    if (!mapped_obs) CU_TRY(h, cudaStreamWaitEvent(st, h->ev_copied, 0));  // `stream` is done when the rows are home
    return USV_OK;
}

int usvgpu_set_host_chunks(usvgpu_handle h, int32_t chunks) {
    if (!h) return USV_ERR_INVALID;
    if (chunks < 1 || chunks > HOST_CHUNKS_MAX) return fail(h, USV_ERR_INVALID, "host chunks must be in 1..16");
    h->host_chunks = chunks;
    return USV_OK;
}

int usvgpu_stream_sync(usvgpu_handle h, void* stream) {
    if (!h) return USV_ERR_INVALID;
    DeviceGuard dev_guard(h->device);
    CU_TRY(h, dev_guard.err);
    CU_TRY(h, cudaStreamSynchronize((cudaStream_t)stream));
    return USV_OK;
}

int usvgpu_flags_sync(usvgpu_handle h) {
    if (!h) return USV_ERR_INVALID;
    if (!h->ev_flags) return fail(h, USV_ERR_INVALID, "no usvgpu_step_host() has run on this handle");
    DeviceGuard dev_guard(h->device);
    CU_TRY(h, dev_guard.err);
    CU_TRY(h, cudaEventSynchronize(h->ev_flags));
    return USV_OK;
}

int usvgpu_terminal_obs_host(usvgpu_handle h, const int32_t* index_host, int64_t n_index, float* rows_host,
                             double* episode_host) {
    if (!h || !index_host || !rows_host || n_index < 0) return USV_ERR_INVALID;
    if (n_index == 0) return USV_OK;
    if (!h->d_tobs || !h->ev_flags) return fail(h, USV_ERR_INVALID, "no usvgpu_step_host() has run on this handle");
    if (n_index > h->sp.n) return fail(h, USV_ERR_INVALID, "more indices than envs");
    DeviceGuard dev_guard(h->device);
    CU_TRY(h, dev_guard.err);
    if (n_index > h->gather_cap) {
        int64_t cap = h->gather_cap ? h->gather_cap : 1024;
        while (cap < n_index) cap *= 2;
        if (cap > h->sp.n) cap = h->sp.n;
        CU_TRY(h, cudaStreamSynchronize(h->side));
        if (h->h_idx) cudaFreeHost(h->h_idx);
        if (h->h_rows) cudaFreeHost(h->h_rows);
        if (h->h_ep) cudaFreeHost(h->h_ep);
        h->h_idx = nullptr;
        h->h_rows = nullptr;
        h->h_ep = nullptr;
        h->gather_cap = 0;
        CU_TRY(h, cudaHostAlloc((void**)&h->h_idx, (size_t)cap * sizeof(int32_t), cudaHostAllocMapped));
        CU_TRY(h, cudaHostAlloc((void**)&h->h_rows, (size_t)cap * h->obs_dim * sizeof(float), cudaHostAllocMapped));
        CU_TRY(h, cudaHostAlloc((void**)&h->h_ep, (size_t)cap * sizeof(double2), cudaHostAllocMapped));
        h->gather_cap = cap;
    }
    // The kernels of the step are complete once the flags are on the host (the caller saw them); the gather runs on
    // the side stream, reading the list from and storing the rows to mapped host memory, while the copy engine is
    // still busy with the observations on the caller's stream.
    CU_TRY(h, cudaStreamWaitEvent(h->side, h->ev_flags, 0));
    memcpy(h->h_idx, index_host, (size_t)n_index * sizeof(int32_t));
    void *d_idx = nullptr, *d_rows = nullptr, *d_ep = nullptr;
    CU_TRY(h, cudaHostGetDevicePointer(&d_idx, h->h_idx, 0));
    CU_TRY(h, cudaHostGetDevicePointer(&d_rows, h->h_rows, 0));
    CU_TRY(h, cudaHostGetDevicePointer(&d_ep, h->h_ep, 0));
    gather_rows_kernel<<<(unsigned)((n_index + 7) / 8), 256, 0, h->side>>>(h->d_tobs, h->sp.ep_done, (const int*)d_idx, n_index,
                                                                         h->sp.n, h->obs_dim, (float*)d_rows, (double2*)d_ep);
    CU_TRY(h, cudaGetLastError());
    CU_TRY(h, cudaStreamSynchronize(h->side));
    memcpy(rows_host, h->h_rows, (size_t)n_index * h->obs_dim * sizeof(float));
    if (episode_host) memcpy(episode_host, h->h_ep, (size_t)n_index * sizeof(double2));
    return USV_OK;
}

int64_t usvgpu_field_bytes(usvgpu_handle h, int32_t field) {
    if (!h) return USV_ERR_INVALID;
    void* p;
    int64_t b;
    if (field_ptr(h, field, &p, &b)) return fail(h, USV_ERR_INVALID, "unknown field");
    return b;
}

int usvgpu_get_state(usvgpu_handle h, int32_t field, void* buf, int64_t bytes, void* stream) {
    if (!h || !buf) return USV_ERR_INVALID;
    void* p;
    int64_t b;
    if (field_ptr(h, field, &p, &b)) return fail(h, USV_ERR_INVALID, "unknown field");
    if (bytes != b) return fail(h, USV_ERR_INVALID, "buffer size does not match usvgpu_field_bytes()");
    DeviceGuard dev_guard(h->device);
    CU_TRY(h, dev_guard.err);
    if (field == USV_FIELD_OBSTACLES || field == USV_FIELD_N_OBS || field == USV_FIELD_TARGET0) {  // rows of the interleaved sensor-input block
        const size_t w = (size_t)(b / h->sp.n);
        CU_TRY(h, cudaMemcpy2DAsync(buf, w, p, (size_t)h->sp.sin_row, w, (size_t)h->sp.n, cudaMemcpyDefault, (cudaStream_t)stream));
    } else {
        CU_TRY(h, cudaMemcpyAsync(buf, p, (size_t)b, cudaMemcpyDefault, (cudaStream_t)stream));
    }
    CU_TRY(h, cudaStreamSynchronize((cudaStream_t)stream));
    return USV_OK;
}

int usvgpu_set_state(usvgpu_handle h, int32_t field, const void* buf, int64_t bytes, void* stream) {
    if (!h || !buf) return USV_ERR_INVALID;
    void* p;
    int64_t b;
    if (field_ptr(h, field, &p, &b)) return fail(h, USV_ERR_INVALID, "unknown field");
    if (bytes != b) return fail(h, USV_ERR_INVALID, "buffer size does not match usvgpu_field_bytes()");
    DeviceGuard dev_guard(h->device);
    CU_TRY(h, dev_guard.err);
    if (field == USV_FIELD_OBSTACLES || field == USV_FIELD_N_OBS || field == USV_FIELD_TARGET0) {
        const size_t w = (size_t)(b / h->sp.n);
        CU_TRY(h, cudaMemcpy2DAsync(p, (size_t)h->sp.sin_row, buf, w, w, (size_t)h->sp.n, cudaMemcpyDefault, (cudaStream_t)stream));
    } else {
        CU_TRY(h, cudaMemcpyAsync(p, buf, (size_t)b, cudaMemcpyDefault, (cudaStream_t)stream));
    }
    CU_TRY(h, cudaStreamSynchronize((cudaStream_t)stream));
    return USV_OK;
}

int usvgpu_episode_stats(usvgpu_handle h, UsvStats* out, int32_t reset, void* stream) {
    if (!h || !out) return USV_ERR_INVALID;
    DeviceGuard dev_guard(h->device);
    CU_TRY(h, dev_guard.err);
    cudaStream_t st = (cudaStream_t)stream;
    CU_TRY(h, cudaMemcpyAsync(out, h->sp.stats, sizeof(UsvStats), cudaMemcpyDeviceToHost, st));
    if (reset) CU_TRY(h, cudaMemsetAsync(h->sp.stats, 0, sizeof(DeviceStats), st));
    CU_TRY(h, cudaStreamSynchronize(st));
    out->env_steps = h->host_env_steps;
    if (reset) h->host_env_steps = 0;
    return USV_OK;
}

int usvgpu_stats_device_ptr(usvgpu_handle h, void** out) {
    if (!h || !out) return USV_ERR_INVALID;
    *out = h->sp.stats;
    return USV_OK;
}

int usvgpu_asmc_compute(int device, int64_t n, double* state, const double* actions, int32_t n_calls,
                        int32_t do_perturb, int64_t first_substep, int32_t precision, double* info, void* stream) {
    if (!state || !actions || n <= 0 || n_calls < 0) return fail(nullptr, USV_ERR_INVALID, "bad asmc_compute arguments");
    if (precision != USV_PREC_F64) return fail(nullptr, USV_ERR_INVALID, "precision not supported by this build");
    int rc = check_device(nullptr, device);
    if (rc) return rc;
    DeviceGuard dev_guard(device);
    CU_TRY(nullptr, dev_guard.err);
    const int threads = 128;
    const unsigned blocks = (unsigned)((n + threads - 1) / threads);
    cudaStream_t st = (cudaStream_t)stream;
    if (do_perturb) asmc_compute_kernel<true><<<blocks, threads, 0, st>>>(n, state, actions, n_calls, first_substep, info);
    else asmc_compute_kernel<false><<<blocks, threads, 0, st>>>(n, state, actions, n_calls, first_substep, info);
    CU_TRY(nullptr, cudaGetLastError());
    return USV_OK;
}

int usvgpu_sensors(int device, int64_t n, const double* pose, const double* obstacles, const int32_t* n_obs,
                   double* out, void* stream) {
    if (!pose || !obstacles || !n_obs || !out || n <= 0) return fail(nullptr, USV_ERR_INVALID, "bad sensors arguments");
    int rc = check_device(nullptr, device);
    if (rc) return rc;
    DeviceGuard dev_guard(device);
    CU_TRY(nullptr, dev_guard.err);
    if (!g_raytab[device & 63]) {
        double2 t[N_SENS];
        host_raytab(t);
        CU_TRY(nullptr, cudaMalloc((void**)&g_raytab[device & 63], sizeof(t)));
        CU_TRY(nullptr, cudaMemcpy(g_raytab[device & 63], t, sizeof(t), cudaMemcpyHostToDevice));
    }
    const unsigned blocks = (unsigned)((n + SENS_WARPS - 1) / SENS_WARPS);
    sensors_unit_kernel<<<blocks, SENS_WARPS * 32, 0, (cudaStream_t)stream>>>(n, g_raytab[device & 63], pose, obstacles,
                                                                            n_obs, out);
    CU_TRY(nullptr, cudaGetLastError());
    return USV_OK;
}

// ---------------------------------------------------------------------------------------------
// usvgpu_sensors_noalloc: the scratch of the no-allocation sensor entry point lives in the module image, not on the heap.
// Every GPU that loads this library holds the array (6 MiB + 2 KiB) until the process exits.  Layout: the staged
// obstacle rows of up to USVGPU_NOALLOC_MAX_ENVS envs (96 doubles each), then the 128-entry ray table.
// In HOST code the name `g_noalloc_scratch` is only a stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, once per device with that device current (noalloc_setup), and only that pointer is handed to
// kernels.
// ---------------------------------------------------------------------------------------------
constexpr size_t NOALLOC_ROW_DOUBLES = (size_t)USVGPU_NOALLOC_MAX_ENVS * OBST_ROW;
constexpr size_t NOALLOC_DOUBLES = NOALLOC_ROW_DOUBLES + 2 * (size_t)N_SENS;
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_DOUBLES];
static double* g_noalloc_ptr[64] = {nullptr};  // per device: device address of g_noalloc_scratch, ray table in place
static std::mutex g_noalloc_mutex;

// once per device, with that device current: symbol address, then the ray table (2 KiB, a plain copy: nothing is allocated)
static int noalloc_setup(int device, double** out) {
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[device & 63]) {
        double* ptr = nullptr;
        CU_TRY(nullptr, cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch));
        if (!ptr) return fail(nullptr, USV_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the no-alloc scratch");
        double2 t[N_SENS];
        host_raytab(t);
        CU_TRY(nullptr, cudaMemcpy(ptr + NOALLOC_ROW_DOUBLES, t, sizeof(t), cudaMemcpyHostToDevice));
        g_noalloc_ptr[device & 63] = ptr;
    }
    *out = g_noalloc_ptr[device & 63];
    return USV_OK;
}

int usvgpu_sensors_noalloc(int device, int64_t n, const double* pose, const double* obstacles, const int32_t* n_obs,
                           double* out, void* stream) {
    if (!pose || !obstacles || !n_obs || !out || n <= 0) return fail(nullptr, USV_ERR_INVALID, "bad sensors arguments");
    if (n > USVGPU_NOALLOC_MAX_ENVS)
        return fail(nullptr, USV_ERR_INVALID,
                    "usvgpu_sensors_noalloc: n = " + std::to_string((long long)n) + " envs exceeds the fixed scratch of " +
                        std::to_string(USVGPU_NOALLOC_MAX_ENVS) + " envs (USVGPU_NOALLOC_MAX_ENVS); split the batch or use usvgpu_sensors");
    if (device < 0 || device >= 64) return fail(nullptr, USV_ERR_INVALID, "usvgpu_sensors_noalloc: device index outside 0..63");
    int rc = check_device(nullptr, device);
    if (rc) return rc;
    DeviceGuard dev_guard(device);
    CU_TRY(nullptr, dev_guard.err);
    double* scratch = nullptr;
    rc = noalloc_setup(device, &scratch);
    if (rc) return rc;
    double* rows = scratch;
    const double2* raytab = reinterpret_cast<const double2*>(scratch + NOALLOC_ROW_DOUBLES);
    cudaStream_t st = (cudaStream_t)stream;
    sensors_stage_kernel<<<(unsigned)((n + STAGE_ENVS - 1) / STAGE_ENVS), 256, 0, st>>>(n, obstacles, n_obs, rows);
    CU_TRY(nullptr, cudaGetLastError());
    sensors_rows_kernel<<<(unsigned)((n + SENS_WARPS - 1) / SENS_WARPS), SENS_WARPS * 32, 0, st>>>(n, raytab, pose, rows, n_obs, out);
    CU_TRY(nullptr, cudaGetLastError());
    return USV_OK;
}

int usvgpu_frame_stack_push(int device, int64_t n, int32_t obs_dim, int32_t n_stack, int32_t pos, float* ring,
                            const float* obs, const uint8_t* terminated, const uint8_t* truncated,
                            const float* terminal_obs, float* terminal_stack, void* stream) {
    if (!ring || !obs || n <= 0 || obs_dim <= 0 || n_stack <= 0 || pos < 0 || pos >= n_stack)
        return fail(nullptr, USV_ERR_INVALID, "bad frame_stack_push arguments");
    int rc = check_device(nullptr, device);
    if (rc) return rc;
    DeviceGuard dev_guard(device);
    CU_TRY(nullptr, dev_guard.err);
    frame_push_kernel<<<(unsigned)((n + 7) / 8), 256, 0, (cudaStream_t)stream>>>(n, obs_dim, n_stack, pos, ring, obs, terminated,
                                                                              truncated, terminal_obs, terminal_stack);
    CU_TRY(nullptr, cudaGetLastError());
    return USV_OK;
}

int usvgpu_gae(int device, int32_t n_steps, int64_t n, const float* rewards, const float* values,
               const float* episode_starts, const float* last_values, const uint8_t* last_dones, double gamma,
               double gae_lambda, float* advantages, float* returns, void* stream) {
    if (!rewards || !values || !episode_starts || !last_values || !last_dones || !advantages || !returns || n <= 0 ||
        n_steps <= 0)
        return fail(nullptr, USV_ERR_INVALID, "bad gae arguments");
    int rc = check_device(nullptr, device);
    if (rc) return rc;
    DeviceGuard dev_guard(device);
    CU_TRY(nullptr, dev_guard.err);
    gae_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(n_steps, n, rewards, values, episode_starts,
                                                                           last_values, last_dones, (float)gamma,
                                                                           (float)(gamma * gae_lambda), advantages, returns);
    CU_TRY(nullptr, cudaGetLastError());
    return USV_OK;
}

int usvgpu_render(usvgpu_handle h, const int64_t* index, int64_t n_frames, const float* obs, int32_t size,
                  uint8_t* frames, void* stream) {
    if (!h) return fail(nullptr, USV_ERR_INVALID, "handle is NULL");
    if (h->legacy || h->ca) return fail(h, USV_ERR_INVALID, "render: only the simple envs draw with SimpleEnvVisualizer");
    if (!obs || !frames || n_frames <= 0 || n_frames > 65535 || size < 1 || size > 8192)
        return fail(h, USV_ERR_INVALID, "bad render arguments (obs / frames NULL, n_frames not in 1..65535, size not in 1..8192)");
    if (!index && n_frames > h->sp.n) return fail(h, USV_ERR_INVALID, "render: more frames than envs");
    if (((uintptr_t)frames & 3) || ((uintptr_t)obs & 3)) return fail(h, USV_ERR_INVALID, "render: frames / obs misaligned");
    DeviceGuard dev_guard(h->device);
    CU_TRY(h, dev_guard.err);
    const int tiles = (size + RND_TILE - 1) / RND_TILE;
    // tiles per CTA: the draw list is built once per CTA, so whole tile rows when the frames alone fill the machine,
    // shorter runs (down to single tiles) for a handful of frames
    const long long cta_target = 32LL * h->sm_count;  // tiles differ a lot in cost (all 128 rays meet near the boat): many CTAs per SM even out the tail
    int tpc = tiles;
    while (tpc > 1 && n_frames * tiles * ((tiles + tpc - 1) / tpc) < cta_target) tpc = (tpc + 1) / 2;
    const int groups = (tiles + tpc - 1) / tpc;
    render_kernel<<<dim3((unsigned)(tiles * groups), (unsigned)n_frames), RND_THREADS, 0, (cudaStream_t)stream>>>(
        h->sp, reinterpret_cast<const long long*>(index), obs, OBS_DIM, h->cfg.obstacles_f32, size, tpc, frames);
    CU_TRY(h, cudaGetLastError());
    h->launches += 1;
    return USV_OK;
}

int usvgpu_step_info(usvgpu_handle h, const int64_t* index, int64_t n_index, double* out, void* stream) {
    if (!h) return fail(nullptr, USV_ERR_INVALID, "handle is NULL");
    if (h->legacy || h->ca) return fail(h, USV_ERR_INVALID, "step_info: only the simple envs return the info dict of simple_env.py");
    if (!out || n_index <= 0 || (!index && n_index > h->sp.n)) return fail(h, USV_ERR_INVALID, "bad step_info arguments");
    DeviceGuard dev_guard(h->device);
    CU_TRY(h, dev_guard.err);
    info_kernel<<<(unsigned)((n_index + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        h->sp, reinterpret_cast<const long long*>(index), n_index, out);
    CU_TRY(h, cudaGetLastError());
    return USV_OK;
}

int usvgpu_reset_obs(usvgpu_handle h, const uint8_t* mask, float* obs, void* stream) {
    if (!h) return fail(nullptr, USV_ERR_INVALID, "handle is NULL");
    if (h->legacy || h->ca) return fail(h, USV_ERR_INVALID, "reset_obs: only the simple envs");
    if (!obs) return fail(h, USV_ERR_INVALID, "reset_obs: obs is NULL");
    DeviceGuard dev_guard(h->device);
    CU_TRY(h, dev_guard.err);
    reset_obs_kernel<<<(unsigned)((h->sp.n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(h->sp, mask, obs);
    CU_TRY(h, cudaGetLastError());
    return USV_OK;
}

int usvgpu_replay_add(int device, int64_t n, int32_t obs_dim, int32_t act_dim, int64_t capacity, int64_t pos,
                      const float* obs, int64_t obs_stride, const float* next_obs, int64_t next_stride,
                      const float* terminal_obs, int64_t terminal_stride, const float* actions, const float* rewards,
                      const uint8_t* terminated, const uint8_t* truncated, float* buf_obs, float* buf_next_obs,
                      float* buf_actions, float* buf_rewards, uint8_t* buf_dones, uint8_t* buf_timeouts, void* stream) {
    if ((!obs && !next_obs) || !buf_obs || !buf_next_obs || !buf_actions || !buf_rewards || !buf_dones || !buf_timeouts)
        return fail(nullptr, USV_ERR_INVALID, "replay_add: NULL argument");
    if (next_obs && (!actions || !rewards || !terminated || !truncated))
        return fail(nullptr, USV_ERR_INVALID, "replay_add: NULL argument");
    if (n <= 0 || obs_dim <= 0 || act_dim <= 0 || act_dim > 32 || capacity <= 0 || pos < 0 || pos >= capacity ||
        (obs && obs_stride < obs_dim) || (next_obs && next_stride < obs_dim) || (terminal_obs && terminal_stride < obs_dim))
        return fail(nullptr, USV_ERR_INVALID, "replay_add: bad sizes (act_dim <= 32, 0 <= pos < capacity, strides >= obs_dim)");
    int rc = check_device(nullptr, device);
    if (rc) return rc;
    DeviceGuard dev_guard(device);
    CU_TRY(nullptr, dev_guard.err);
    const ReplayPtrs b = {buf_obs, buf_next_obs, buf_actions, buf_rewards, buf_dones, buf_timeouts};
    replay_add_kernel<<<(unsigned)((n * 32 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        b, n, obs_dim, act_dim, pos, obs, obs_stride, next_obs, next_stride, terminal_obs, terminal_stride, actions,
        rewards, terminated, truncated);
    CU_TRY(nullptr, cudaGetLastError());
    return USV_OK;
}

int usvgpu_replay_sample(int device, int64_t n, int32_t obs_dim, int32_t act_dim, int64_t capacity, int64_t batch,
                         const int64_t* batch_inds, const int64_t* env_inds, const float* buf_obs,
                         const float* buf_next_obs, const float* buf_actions, const float* buf_rewards,
                         const uint8_t* buf_dones, const uint8_t* buf_timeouts, float* out_obs, float* out_next_obs,
                         float* out_actions, float* out_rewards, float* out_dones, void* stream) {
    if (!batch_inds || !env_inds || !buf_obs || !buf_next_obs || !buf_actions || !buf_rewards || !buf_dones ||
        !buf_timeouts || !out_obs || !out_next_obs || !out_actions || !out_rewards || !out_dones)
        return fail(nullptr, USV_ERR_INVALID, "replay_sample: NULL argument");
    if (n <= 0 || obs_dim <= 0 || act_dim <= 0 || act_dim > 32 || batch <= 0 || capacity <= 0)
        return fail(nullptr, USV_ERR_INVALID, "replay_sample: bad sizes");
    int rc = check_device(nullptr, device);
    if (rc) return rc;
    DeviceGuard dev_guard(device);
    CU_TRY(nullptr, dev_guard.err);
    const ReplayPtrs b = {const_cast<float*>(buf_obs), const_cast<float*>(buf_next_obs), const_cast<float*>(buf_actions),
                          const_cast<float*>(buf_rewards), const_cast<uint8_t*>(buf_dones),
                          const_cast<uint8_t*>(buf_timeouts)};
    replay_sample_kernel<<<(unsigned)((batch * 32 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        b, n, capacity, obs_dim, act_dim, batch, reinterpret_cast<const long long*>(batch_inds),
        reinterpret_cast<const long long*>(env_inds), out_obs, out_next_obs, out_actions, out_rewards, out_dones);
    CU_TRY(nullptr, cudaGetLastError());
    return USV_OK;
}

int usvgpu_math_probe(int device, int32_t which, int64_t n, const double* in, double* out, void* stream) {
    if (!in || !out || n <= 0 || which < 0 || which > 4) return fail(nullptr, USV_ERR_INVALID, "bad math_probe arguments");
    int rc = check_device(nullptr, device);
    if (rc) return rc;
    DeviceGuard dev_guard(device);
    CU_TRY(nullptr, dev_guard.err);
    const int threads = 128;
    math_probe_kernel<<<(unsigned)((n + threads - 1) / threads), threads, 0, (cudaStream_t)stream>>>(which, n, in, out);
    CU_TRY(nullptr, cudaGetLastError());
    return USV_OK;
}

int usvgpu_fp64_peak(int device, int32_t blocks, int32_t iters, double* dfma_per_sec, void* stream) {
    if (!dfma_per_sec || blocks == 0 || iters <= 0) return fail(nullptr, USV_ERR_INVALID, "bad fp64_peak arguments");
    int rc = check_device(nullptr, device);
    if (rc) return rc;
    DeviceGuard dev_guard(device);
    CU_TRY(nullptr, dev_guard.err);
    cudaStream_t st = (cudaStream_t)stream;
    double* sink = nullptr;
    CU_TRY(nullptr, cudaMalloc((void**)&sink, sizeof(double)));
    cudaEvent_t ev[2];
    CU_TRY(nullptr, cudaEventCreate(&ev[0]));
    CU_TRY(nullptr, cudaEventCreate(&ev[1]));
    // blocks > 0: 256 threads x 8 independent chains (throughput).  blocks < 0: |blocks| CTAs of 128 threads (one
    // warp per SM sub-partition when |blocks| <= SM count) x 1 dependent chain (latency: 8 FMAs per iteration).
    const bool lat = blocks < 0;
    if (lat) blocks = -blocks;
    if (lat) fp64_peak_kernel<1><<<blocks, 128, 0, st>>>(iters / 8 + 1, 1.0, sink);
    else fp64_peak_kernel<8><<<blocks, 256, 0, st>>>(iters / 8 + 1, 1.0, sink);  // warm-up
    CU_TRY(nullptr, cudaEventRecord(ev[0], st));
    if (lat) fp64_peak_kernel<1><<<blocks, 128, 0, st>>>(iters, 1.0, sink);
    else fp64_peak_kernel<8><<<blocks, 256, 0, st>>>(iters, 1.0, sink);
    CU_TRY(nullptr, cudaEventRecord(ev[1], st));
    CU_TRY(nullptr, cudaEventSynchronize(ev[1]));
    float ms = 0.f;
    CU_TRY(nullptr, cudaEventElapsedTime(&ms, ev[0], ev[1]));
    cudaEventDestroy(ev[0]);
    cudaEventDestroy(ev[1]);
    cudaFree(sink);
    *dfma_per_sec = (double)blocks * (lat ? 128.0 : 256.0) * 8.0 * (double)iters / ((double)ms * 1e-3);
    return USV_OK;
}

}  // extern "C"
