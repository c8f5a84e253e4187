// usv_state.cuh -- device-side state layout of one shard of 'usv-asmc-simple' envs.
//
// Structure of arrays, env index fastest, so that the thread-per-env dynamics kernel reads and
// writes every field fully coalesced.  The obstacle table is the exception: the sensor kernel
// gives one WARP to an env (lane = obstacle), so a row of the table is the 3 x 32 values of one
// env, contiguous (384 B in float32): three coalesced 128 B loads per env.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/usvgpu.h"
#include "usv_model.cuh"

namespace usv {

constexpr int OBS_DIM = USVGPU_OBS_DIM;
constexpr int N_SENS = USVGPU_SENSORS;
constexpr int MAX_OBST = USVGPU_MAX_OBSTACLES;
constexpr int HDR = 15;                // observation columns before the sensors (simple_env.py:91-96)
constexpr int RAYS_FULL_CIRCLE = 192;  // 2 pi / pitch, pitch = (4 pi / 3) / 128 (simple_env.py:12-14)
constexpr int OBST_ROW = 3 * MAX_OBST; // elements per env in the obstacle table: x[32] y[32] r[32]

// rows of USV_FIELD_DYN
enum DynRow {
    D_X = 0, D_Y, D_PSI, D_U, D_V, D_R,
    D_XD_L, D_YD_L, D_PSID_L, D_UD_L, D_VD_L, D_RD_L, D_EU_L, D_KADU_L, D_KADP_L,
    D_PSI_D_L, D_ODD_L, D_OD_L, D_O_L,
    D_EU_INT, D_KA_U, D_KA_PSI,
    D_PROGRESS,
    N_DYN
};
static_assert(N_DYN == USV_DYN_ROWS, "header/kernel row count mismatch");
// 'usv-simple' (UsvSimpleEnv itself, no ASMC controller): the controller rows are unused and two of them hold
// last_action[0] and last_action[2] (simple_env.py:41,343), which a reset does NOT clear
constexpr int D_LA0 = D_XD_L, D_LA2 = D_YD_L;
// ... and two more hold what only the info dict of 'usv-simple' needs (usv_info.cuh): sum |last_action - action| of the
// last step (simple_env.py:158,196 'delta_action') and last_action[0] before it (:198 'reward_velocity')
constexpr int D_DACT = D_PSID_L, D_LA0_PREV = D_UD_L;
// rows of USV_FIELD_EPI
enum EpiRow { E_PSX = 0, E_PSY, E_PEX, E_PEY, E_MA0, E_MA2, E_VREF, N_EPI };
static_assert(N_EPI == USV_EPI_ROWS, "header/kernel row count mismatch");

// Hand-over from the dynamics kernel to the sensor kernel (same launch pair, stays in L2): one 48-byte record per
// env (three 16-byte loads, all lanes of the env's warp read the same addresses) plus the observation header
// row, 15 float32 columns of simple_env.py:91-96 already divided and rounded, padded to 16 (one 64-byte line).
struct alignas(16) Hand {
    double x, y;        // pose the sensors see
    double cs, sn;      // cos / sin of its heading
    double rew_part;    // reward without the collision term
    unsigned flags;     // FLG_OOB | FLG_TLIMIT
    unsigned pad;
};
static_assert(sizeof(Hand) == 48, "hand-over record");
constexpr int HDR_PAD = 16;
constexpr unsigned FLG_OOB = 1u, FLG_TLIMIT = 2u;

// simple_env.py:34-35,56,13
constexpr double MAX_ACC_U = 1.75, MAX_ACC_R = 3.0, DT = 1.0 / 25, ENV_HI = 20.0, SENSOR_MAX_RANGE = 100.0;
constexpr double SENSOR_PITCH = ((2.0 / 3) * (2 * PI)) / N_SENS;
constexpr float INV_PITCH_F = (float)(1.0 / SENSOR_PITCH);
constexpr double DIST_NORM = 28.284271247461902;  // hypot(20, 20), simple_env.py:80

struct DeviceStats {
    double return_sum;
    double length_sum;
    unsigned long long episodes, terminated, out_of_bounds, time_limit, env_steps, nonfinite;
};
static_assert(sizeof(DeviceStats) == sizeof(UsvStats), "UsvStats layout");

struct StatePtrs {
    double* dyn;        // [N_DYN][n]
    double* epi;        // [N_EPI][n]
    // Sensor-input rows [n][sin_row bytes], 128-byte aligned: everything sensor_kernel needs for one env in one
    // contiguous block -- obstacle table x[32] y[32] radius[32] (float or double) | Hand 48 B | observation header row
    // 64 B | n_obs int32 | reset target 2 x float32 | 4 B pad.  512 B per env with a float32 table (one 16-byte cp.async per lane), 896 B with
    // float64.  The dynamics kernel writes Hand + header (one 128-byte line), the reset writes table + n_obs.
    char* sin;
    int sin_row;        // bytes per env row
    int sin_tail;       // byte offset of the Hand record inside a row (= table bytes)
    int* steps;         // [n]
    int* episode;       // [n]
    double* ep_return;  // [n]
    double2* ep_done;   // [n]  (return, length) of the env's last finished episode (SB3 infos[i]["episode"])
    DeviceStats* stats;
    const double2* raytab;  // [128] cos/sin of the ray angles relative to the bow
    long long n;
    long long global_offset;
    unsigned long long seed;
    int kinematic;          // env kind 'usv-simple': rows D_LA0 / D_LA2 survive resets
};

struct StepFlags {
    int do_perturb, skip_kinematic, max_episode_steps, auto_reset;
    long long lo, hi;  // env range [lo, hi) of this launch: the host-buffer face steps a shard in chunks so that the
                       // copy-out of one chunk overlaps the kernels of the next
};

constexpr int SIN_TAIL_BYTES = 128;  // Hand (48) + header row (64) + n_obs (4) + reset target x, y float32 (8) + pad (4)
template <typename ObT>
__device__ __forceinline__ ObT* obst_row(const StatePtrs& s, long long i) {
    return reinterpret_cast<ObT*>(s.sin + i * (long long)(sizeof(ObT) * OBST_ROW + SIN_TAIL_BYTES));
}
template <typename ObT>
__device__ __forceinline__ ObT* obst_row(const StatePtrs& s, unsigned i) {  // 32-bit env index: one IMAD.WIDE
    return reinterpret_cast<ObT*>(s.sin + (size_t)i * (unsigned)(sizeof(ObT) * OBST_ROW + SIN_TAIL_BYTES));
}
__device__ __forceinline__ char* sin_tail(const StatePtrs& s, long long i) { return s.sin + i * s.sin_row + s.sin_tail; }
__device__ __forceinline__ Hand* hand_of(const StatePtrs& s, long long i) { return reinterpret_cast<Hand*>(sin_tail(s, i)); }
__device__ __forceinline__ float* hdr_of(const StatePtrs& s, long long i) {
    return reinterpret_cast<float*>(sin_tail(s, i) + sizeof(Hand));
}
__device__ __forceinline__ int* nobs_of(const StatePtrs& s, long long i) {
    return reinterpret_cast<int*>(sin_tail(s, i) + sizeof(Hand) + HDR_PAD * sizeof(float));
}

// the random target_position a reset draws (simple_env.py:245); only the renderer reads it, for envs at step 0
__device__ __forceinline__ float* target0_of(const StatePtrs& s, long long i) {
    return reinterpret_cast<float*>(sin_tail(s, i) + sizeof(Hand) + HDR_PAD * sizeof(float) + sizeof(int));
}

// Programmatic dependent launch between the kernels of consecutive launches (build knob, 0 = off): 1 = the launches carry
// cudaLaunchAttributeProgrammaticStreamSerialization and every kernel runs griddepcontrol.wait before it touches anything
// another kernel wrote (its own prologue -- parameters, indices, the ray table -- overlaps the previous kernel's drain);
// 2 = the kernels also signal griddepcontrol.launch_dependents once their main loop is behind them.
// Measured at 65 536 envs (r02m, us per step): off 68.63, 1: 67.83, 2: 68.03.  (An early trigger -- at kernel start,
// r02d -- made it 92: dependents that become resident early only sit in the wait.)
#ifndef USV_PDL
#define USV_PDL 1
#endif
__device__ __forceinline__ void pdl_wait() {
#if USV_PDL
    asm volatile("griddepcontrol.wait;" ::: "memory");
#endif
}
__device__ __forceinline__ void pdl_trigger() {
#if USV_PDL >= 2
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif
}

}  // namespace usv
