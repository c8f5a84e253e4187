// usv_info.cuh -- the per-step `info` dict of the simple envs as a device matrix, and the reset observation recomputed
// from state (for resets whose state the caller edited: reset options, custom experiments).
//
// UsvSimpleEnv.step returns info = _get_info(reward, action) | reward_info (gym_usv/envs/simple_env.py:102-115,
// 189-199, 341-342).  Every entry is a function of the post-step state, so nothing is stored per step: info_kernel
// recomputes the rows of the envs the caller asks for (thread per env, float64, the reference's formulas), on demand.
// A row is only meaningful for an env that did NOT finish in the step: a finished env has already been reset (SB3
// semantics) and its row describes the fresh episode, like `reset()`'s info (simple_env.py:305).
#pragma once
#include "usv_state.cuh"

namespace usv {

// columns of one info row (include/usvgpu.h: USVGPU_INFO_DIM, UsvInfoCol)
enum InfoCol {
    I_X = 0, I_Y, I_PSI, I_U, I_V, I_R,      // 'position', 'velocity'
    I_PSX, I_PSY, I_PEX, I_PEY,              // 'path_start', 'path_end'
    I_ACTION0, I_ACTION1,                    // 'action0', 'action1': the scaled, filtered action of the step (:311-317)
    I_YE, I_ANGLE,                           // 'ye', 'angle_to_target' (= angle / pi, _get_target_state()[0])
    I_YE_REW, I_ANGLE_REW, I_ANGLE_ACT_REW,  // 'ye_reward', 'angle_to_target_reward', 'angle_action_reward' (= 0)
    I_DACT_REW, I_DACT,                      // 'delta_action_reward', 'delta_action'
    I_VEL_REW, I_VREF, I_REW_VEL, I_VREF_ERR,  // 'velocity_track_reward', 'reference_velocity', 'reward_velocity',
                                               // 'reference_velocity_error'
    I_PROGRESS,
    N_INFO
};
static_assert(N_INFO == USVGPU_INFO_DIM, "header/kernel info row mismatch");

struct TargetView {
    double tx, ty;
};
// target_position as the reference holds it: the point on the path after a step (:328), the random draw of the reset
// (:245) for an env that has not stepped yet
__device__ __forceinline__ TargetView target_of(const StatePtrs& s, long long i, int steps, double psx, double psy, double pex,
                                                double pey, double progress) {
    TargetView t;
    if (steps == 0) {
        const float* t0 = target0_of(s, i);
        t.tx = (double)t0[0];
        t.ty = (double)t0[1];
    } else {
        t.tx = psx + progress * (pex - psx);
        t.ty = psy + progress * (pey - psy);
    }
    return t;
}

__global__ void info_kernel(StatePtrs s, const long long* __restrict__ index, long long n_index, double* __restrict__ out) {
    const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_index) return;
    const long long i = index ? index[k] : k;
    double* row = out + k * N_INFO;
    if (i < 0 || i >= s.n) {
        for (int c = 0; c < N_INFO; ++c) row[c] = 0.0;
        return;
    }
    const long long n = s.n;
    const double* dyn = s.dyn + i;
    const double* epi = s.epi + i;
    const double x = dyn[(long long)D_X * n], y = dyn[(long long)D_Y * n], psi = dyn[(long long)D_PSI * n];
    const double u = dyn[(long long)D_U * n], v = dyn[(long long)D_V * n], r = dyn[(long long)D_R * n];
    const double progress = dyn[(long long)D_PROGRESS * n];
    const double psx = epi[(long long)E_PSX * n], psy = epi[(long long)E_PSY * n];
    const double pex = epi[(long long)E_PEX * n], pey = epi[(long long)E_PEY * n];
    const double vref = epi[(long long)E_VREF * n];
    const TargetView t = target_of(s, i, s.steps[i], psx, psy, pex, pey, progress);
    const double a_k = atan2(pey - psy, pex - psx);                       // :134
    const double ye = -(x - psx) * sin(a_k) + (y - psy) * cos(a_k);       // :135-137
    const double raw = atan2(t.ty - y, t.tx - x) - psi;                   // :67-69
    const double ang = atan2(sin(raw), cos(raw));                         // _wrap_angle, :64-65
    const double q = ye / 0.075;
    const double ye_rew = fmax(exp(-fabs(q)), exp(-(q * q)));             // :165-170
    double la0 = 0.0, la2 = 0.0, dact = 0.0, la0_prev = 0.0;
    if (s.kinematic) {
        la0 = dyn[(long long)D_LA0 * n];
        la2 = dyn[(long long)D_LA2 * n];
        dact = dyn[(long long)D_DACT * n];
        la0_prev = dyn[(long long)D_LA0_PREV * n];
    }
    row[I_X] = x;
    row[I_Y] = y;
    row[I_PSI] = psi;
    row[I_U] = u;
    row[I_V] = v;
    row[I_R] = r;
    row[I_PSX] = psx;
    row[I_PSY] = psy;
    row[I_PEX] = pex;
    row[I_PEY] = pey;
    row[I_ACTION0] = la0;
    row[I_ACTION1] = la2;
    row[I_YE] = ye;
    row[I_ANGLE] = ang / PI;
    row[I_YE_REW] = ye_rew;
    row[I_ANGLE_REW] = exp(-fabs(ang));                                   // :172
    row[I_ANGLE_ACT_REW] = 0.0;                                           // :178
    row[I_DACT_REW] = -(dact / 2) * 0.15;                                 // :176
    row[I_DACT] = dact;
    row[I_VEL_REW] = exp(-fabs(hypot(u, v) - vref)) * 0.05;               // :180
    row[I_VREF] = vref;
    row[I_REW_VEL] = la0_prev;                                            // :198 (last_action before :343)
    row[I_VREF_ERR] = la0_prev - vref;
    row[I_PROGRESS] = progress;
}

// obs[i][0..14] <- _get_obs(zeros(3)) of simple_env.py:302 from the env's CURRENT state (the sensor columns stay as they
// are: the reference shows the previous episode's stale sensor_data there).  For envs whose reset state was edited.
__global__ void reset_obs_kernel(StatePtrs s, const uint8_t* __restrict__ mask, float* __restrict__ obs) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s.n) return;
    if (mask && !mask[i]) return;
    const long long n = s.n;
    const double* dyn = s.dyn + i;
    const double* epi = s.epi + i;
    const double x = dyn[(long long)D_X * n], y = dyn[(long long)D_Y * n], psi = dyn[(long long)D_PSI * n];
    const double psx = epi[(long long)E_PSX * n], psy = epi[(long long)E_PSY * n];
    const double pex = epi[(long long)E_PEX * n], pey = epi[(long long)E_PEY * n];
    const TargetView t = target_of(s, i, s.steps[i], psx, psy, pex, pey, dyn[(long long)D_PROGRESS * n]);
    const double a_k = atan2(pey - psy, pex - psx);
    const double ye = -(x - psx) * sin(a_k) + (y - psy) * cos(a_k);
    const double raw = atan2(t.ty - y, t.tx - x) - psi;
    const double ang = atan2(sin(raw), cos(raw));
    const double ma0 = epi[(long long)E_MA0 * n], ma2 = epi[(long long)E_MA2 * n];
    float* o = obs + i * OBS_DIM;
    o[0] = (float)(dyn[(long long)D_U * n] / 10);
    o[1] = (float)(dyn[(long long)D_V * n] / 10);
    o[2] = (float)(dyn[(long long)D_R * n] / 10);
    o[3] = (float)(ang / PI);
    o[4] = (float)(hypot(x - t.tx, y - t.ty) / DIST_NORM);
    o[5] = (float)(ye / 10);
    o[6] = (float)(epi[(long long)E_VREF * n] / 10);
    o[7] = 0.0f;
    o[8] = 0.0f;
    o[9] = (float)(ma0 / 10);
    o[10] = 0.0f;
    o[11] = (float)(ma2 / 10);
    o[12] = (float)(MAX_ACC_U / 10);
    o[13] = 0.0f;
    o[14] = (float)(MAX_ACC_R / 10);
}

}  // namespace usv
