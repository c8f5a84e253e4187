// usv_reset.cuh -- Env.reset() of one env: simple_env_asmc.py:14-16 -> simple_env.py:228-308 with
// Philox draws (usv_philox.cuh).  Two forms that produce the same values:
//   reset_env       one thread does the whole reset (stand-alone reset kernel);
//   reset_env_warp  a warp does it (auto-reset inside the sensor kernel): lanes draw the obstacle
//                   candidates in parallel, so a finished episode does not leave one lane running
//                   ~100 serial Philox blocks while its warp waits.
// Draw order (draw k = half k&1 of block k>>1): 0,1 normal pair | 2 psi | 3 angle | 4 dist |
// 5,6 target | 7,8,9 nu | 10,11,12 max_action | 13 max_action[2] | 14 v_ref | 15 obstacle_n |
// 16+2j, 17+2j obstacle j | (2 more if none survived) | radii of the kept ones.
#pragma once
#include "usv_philox.cuh"
#include "usv_state.cuh"

namespace usv {

struct ResetScalars {
    double psx, psy, psi, pex, pey, tx, ty, u, v, r, ma0, ma2, vref;
    int nob;
};

__device__ __forceinline__ double lerp_u(double lo, double hi, double u01) { return lo + (hi - lo) * u01; }

__device__ __forceinline__ ResetScalars reset_scalars(const double (&un)[16]) {
    ResetScalars q;
    const double rad = sqrt(-2.0 * log(1.0 - un[0]));  // Box-Muller pair
    double sn, cs;
    sincos_bounded(6.283185307179586 * un[1], &sn, &cs);
    q.psx = 0.5 * (rad * cs) + ENV_HI / 2;  // :234-235
    q.psy = 0.5 * (rad * sn) + ENV_HI / 2;
    // :236-237 draws a position that :238 overwrites; not drawn
    q.psi = lerp_u(-PI, PI, un[2]);                    // :238
    const double angle = lerp_u(-PI, PI, un[3]);       // :241
    const double dist = lerp_u(100.0, 110.0, un[4]);   // :242
    double sa, ca;
    sincos_bounded(angle, &sa, &ca);
    q.pex = q.psx + ca * dist;                         // :243
    q.pey = q.psy + sa * dist;
    q.tx = lerp_u(0.0, ENV_HI, un[5]);                 // :245
    q.ty = lerp_u(0.0, ENV_HI, un[6]);
    q.u = lerp_u(0.0, 0.15, un[7]);                    // :246
    q.v = lerp_u(0.0, 0.15, un[8]);
    q.r = lerp_u(0.0, 0.15, un[9]);
    q.ma0 = lerp_u(1.50, 3.0, un[10]);                 // :249 (un[11], un[12]: max_action[1] zeroed at :254, [2] redrawn)
    q.ma2 = lerp_u(3.0, 6.0, un[13]);                  // :250
    q.vref = lerp_u(0.75, q.ma0, un[14]);              // :251
    q.nob = min(15 + (int)floor(un[15] * 15.0), 29);   // :257 integers(15, 30)
    return q;
}

// :302  _get_obs(zeros(3)) with the random target and position == path_start: column k of the header
__device__ __forceinline__ void reset_header(const ResetScalars& q, float (&hdr)[HDR]) {
    const double ang = wrap_pi(atan2(q.ty - q.psy, q.tx - q.psx) - q.psi);
    hdr[0] = (float)(q.u / 10);
    hdr[1] = (float)(q.v / 10);
    hdr[2] = (float)(q.r / 10);
    hdr[3] = (float)(ang / PI);
    hdr[4] = (float)(hypot(q.psx - q.tx, q.psy - q.ty) / DIST_NORM);
    hdr[5] = 0.0f;  // ye: the boat sits on path_start
    hdr[6] = (float)(q.vref / 10);
    hdr[7] = 0.0f;
    hdr[8] = 0.0f;
    hdr[9] = (float)(q.ma0 / 10);
    hdr[10] = 0.0f;
    hdr[11] = (float)(q.ma2 / 10);
    hdr[12] = (float)(MAX_ACC_U / 10);
    hdr[13] = 0.0f;
    hdr[14] = (float)(MAX_ACC_R / 10);
}

// value of USV_FIELD_DYN row k / USV_FIELD_EPI row k right after a reset
__device__ __forceinline__ double reset_dyn_row(const ResetScalars& q, int k) {
    return k == D_X ? q.psx : k == D_Y ? q.psy : k == D_PSI ? q.psi : k == D_U ? q.u : k == D_V ? q.v : k == D_R ? q.r : 0.0;
}
__device__ __forceinline__ double reset_epi_row(const ResetScalars& q, int k) {
    return k == E_PSX ? q.psx : k == E_PSY ? q.psy : k == E_PEX ? q.pex : k == E_PEY ? q.pey : k == E_MA0 ? q.ma0
           : k == E_MA2 ? q.ma2 : q.vref;
}

// obstacle candidate j survives unless its centre is within 0.5 of the boat or of the target (:261-268)
template <typename ObT>
__device__ __forceinline__ bool obstacle_kept(const ResetScalars& q, double a, double b, double& ox, double& oy) {
    ox = (double)(ObT)lerp_u(0.0, ENV_HI, a);
    oy = (double)(ObT)lerp_u(0.0, ENV_HI, b);
    const double dp = hypot(q.psx - ox, q.psy - oy), dt = hypot(q.tx - ox, q.ty - oy);
    return !(dp < 0.5 || dt < 0.5);
}

// ---------------------------------------------------------------------------------------------
// one thread resets env i; hdr (may be NULL) receives observation columns 0..14
// ---------------------------------------------------------------------------------------------
template <typename ObT>
__device__ void reset_env(const StatePtrs& s, long long i, float* hdr) {
    const long long n = s.n;
    const int ep = s.episode[i] + 1;
    s.episode[i] = ep;
    s.steps[i] = 0;
    s.ep_return[i] = 0.0;
    const DrawKey dk(s.seed, (unsigned long long)(s.global_offset + i), (uint32_t)ep);
    double un[16];
#pragma unroll
    for (int b = 0; b < 8; ++b) dk.pair(b, un[2 * b], un[2 * b + 1]);
    const ResetScalars q = reset_scalars(un);
    ObT* row = obst_row<ObT>(s, i);
    int kept = 0;
    for (int j = 0; j < q.nob; ++j) {  // :258-268
        double a, b, ox, oy;
        dk.pair(8 + j, a, b);
        if (!obstacle_kept<ObT>(q, a, b, ox, oy)) continue;
        row[kept] = (ObT)ox;
        row[MAX_OBST + kept] = (ObT)oy;
        ++kept;
    }
    int r0 = 16 + 2 * q.nob;
    if (kept == 0) {  // :270-274
        double a, b;
        dk.pair(8 + q.nob, a, b);
        row[0] = (ObT)lerp_u(0.0, ENV_HI, a);
        row[MAX_OBST] = (ObT)lerp_u(0.0, ENV_HI, b);
        kept = 1;
        r0 += 2;
    }
    for (int j = 0; j < kept; j += 2) {  // :290
        double a, b;
        dk.pair((r0 + j) >> 1, a, b);
        row[2 * MAX_OBST + j] = (ObT)lerp_u(0.15, 0.5, a);
        if (j + 1 < kept) row[2 * MAX_OBST + j + 1] = (ObT)lerp_u(0.15, 0.5, b);
    }
    *nobs_of(s, i) = kept;
    target0_of(s, i)[0] = (float)q.tx;
    target0_of(s, i)[1] = (float)q.ty;
#pragma unroll
    for (int k = 0; k < N_DYN; ++k) {  // fresh UsvAsmc(), progress = 0
        if (s.kinematic && (k == D_LA0 || k == D_LA2)) continue;
        s.dyn[(long long)k * n + i] = reset_dyn_row(q, k);
    }
#pragma unroll
    for (int k = 0; k < N_EPI; ++k) s.epi[(long long)k * n + i] = reset_epi_row(q, k);
    if (hdr) {
        float h[HDR];
        reset_header(q, h);
#pragma unroll
        for (int k = 0; k < HDR; ++k) hdr[k] = h[k];
    }
}

// ---------------------------------------------------------------------------------------------
// one warp resets env i; returns observation column `lane` of the reset observation (lane < 15)
// ---------------------------------------------------------------------------------------------
template <typename ObT>
__device__ float reset_env_warp(const StatePtrs& s, long long i, int lane) {
    constexpr unsigned FULL = 0xFFFFFFFFu;
    const long long n = s.n;
    const int ep = s.episode[i] + 1;
    __syncwarp();
    const DrawKey dk(s.seed, (unsigned long long)(s.global_offset + i), (uint32_t)ep);
    double a = 0.0, b = 0.0;
    if (lane < 8) dk.pair(lane, a, b);
    double un[16];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        un[2 * k] = __shfl_sync(FULL, a, k);
        un[2 * k + 1] = __shfl_sync(FULL, b, k);
    }
    const ResetScalars q = reset_scalars(un);
    ObT* row = obst_row<ObT>(s, i);
    bool ok = false;
    double ox = 0.0, oy = 0.0;
    if (lane < q.nob) {
        dk.pair(8 + lane, a, b);
        ok = obstacle_kept<ObT>(q, a, b, ox, oy);
    }
    const unsigned bal = __ballot_sync(FULL, ok);
    int kept = __popc(bal);
    const int rank = __popc(bal & ((1u << lane) - 1u));
    if (ok) {
        row[rank] = (ObT)ox;
        row[MAX_OBST + rank] = (ObT)oy;
    }
    int r0 = 16 + 2 * q.nob;
    if (kept == 0) {  // warp-uniform
        dk.pair(8 + q.nob, a, b);
        if (lane == 0) {
            row[0] = (ObT)lerp_u(0.0, ENV_HI, a);
            row[MAX_OBST] = (ObT)lerp_u(0.0, ENV_HI, b);
        }
        kept = 1;
        r0 += 2;
    }
    if (lane < kept) {
        dk.pair((r0 + lane) >> 1, a, b);
        row[2 * MAX_OBST + lane] = (ObT)lerp_u(0.15, 0.5, (lane & 1) ? b : a);  // r0 is even
    }
    if (lane == 0) {
        s.episode[i] = ep;
        s.steps[i] = 0;
        s.ep_return[i] = 0.0;
        *nobs_of(s, i) = kept;
        target0_of(s, i)[0] = (float)q.tx;
        target0_of(s, i)[1] = (float)q.ty;
    }
    if (lane < N_DYN && !(s.kinematic && (lane == D_LA0 || lane == D_LA2))) s.dyn[(long long)lane * n + i] = reset_dyn_row(q, lane);
    if (lane < N_EPI) s.epi[(long long)lane * n + i] = reset_epi_row(q, lane);
    float h[HDR];
    reset_header(q, h);
    float mine = 0.0f;
#pragma unroll
    for (int k = 0; k < HDR; ++k) mine = (lane == k) ? h[k] : mine;
    return mine;
}

}  // namespace usv
