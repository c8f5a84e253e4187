// usv_ca.cuh -- the collision-avoidance env 'usv-asmc-ca-v0' (gym_usv/envs/usv_asmc_ca_env.py: UsvAsmcCaEnv; point
// target, <= 9 obstacles, 16-ray lidar, observation float32[23], TimeLimit 5000) with the in-repo controller
// substituted for the absent usv_libs_py: one env step = one UsvAsmc.compute
// (gym_usv/control/usv_asmc.py:53-244, 10 sub-steps of 0.01 s = asmc_substep of usv_model.cuh) with
// action = [setpoint.velocity, setpoint.heading] -- the substitution the reference keeps as a commented-out line
// (usv_asmc_ca_env.py:210).  Thread per env; the obstacle geometry of the CTA's envs sits in shared memory while the
// 16 x n_obs ray tests run (the reference's test, usv_asmc_ca_env.py:439-461, in float64).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "usv_model.cuh"
#include "usv_philox.cuh"
#include "usv_state.cuh"

namespace usv {

constexpr int CA_SENS = 16;                 // usv_asmc_ca_env.py:44
constexpr int CA_OBS = USVGPU_CA_OBS_DIM;   // :99  7 + sensor_num
constexpr int CA_MAX_OBST = USVGPU_CA_MAX_OBSTACLES;  // :349 int(uniform(2, 10)) -> 2..9
constexpr int CA_ROWS = USV_CA_ROWS;        // target x y | previous action (2) | obstacle x[9] y[9] radius[9]
#ifndef USV_CA_B
#define USV_CA_B 64
#endif
#ifndef USV_CA_MIN_CTAS
#define USV_CA_MIN_CTAS 8  // 128 registers: 1.64e9 env-steps/s at 262 144 envs against 1.44e9 with 162 (profiles/r01z_ca_launch_bounds.txt)
#endif
constexpr int CA_B = USV_CA_B;              // threads (= envs) per CTA
constexpr int CA_MIN_CTAS = USV_CA_MIN_CTAS;
enum CaRow { C_TX = 0, C_TY, C_PA0, C_PA1, C_OX, C_OY = C_OX + CA_MAX_OBST, C_OR = C_OY + CA_MAX_OBST };
static_assert(C_OR + CA_MAX_OBST == CA_ROWS, "CA row count");

struct CaPtrs {
    double* dyn;        // [N_DYN][n]  the rows of USV_FIELD_DYN (pose, velocity, controller memory; row 22 unused)
    double* ca;         // [CA_ROWS][n]
    int* n_obs;         // [n]
    int* steps;         // [n]
    int* episode;       // [n]
    double* ep_return;  // [n]
    double2* ep_done;   // [n]
    DeviceStats* stats;
    long long n;
    long long global_offset;
    unsigned long long seed;
};

// cos / sin of the ray angles relative to the bow, -2 pi / 3 + k (4 pi / 3) / 16 (usv_asmc_ca_env.py:47-49, 420-421);
// written once per process by usvgpu_create
__constant__ double2 CA_RAYTAB[CA_SENS];

// usv_asmc_ca_env.py:134-144 with in_min = -1, in_max = 1 on a float32 action element: numpy keeps the whole expression
// in float32 (the Python scalars are weak, NEP 50)
__device__ __forceinline__ float ca_denorm_f32(float x, float span, float out_min) {
    float t = __fsub_rn(x, -1.0f);
    t = __fmul_rn(t, span);
    t = __fdiv_rn(t, 2.0f);
    return __fadd_rn(t, out_min);
}
__device__ __forceinline__ double ca_wrap(double a) {  // :131-132 arctan2(sin a, cos a)
    double s, c;
    sincos(a, &s, &c);
    return atan2(s, c);
}

__device__ __forceinline__ void ca_load(const double* dyn, long long n, Pose& p, Ctrl& c) {
    p.x = dyn[(long long)D_X * n]; p.y = dyn[(long long)D_Y * n]; p.psi = dyn[(long long)D_PSI * n];
    p.u = dyn[(long long)D_U * n]; p.v = dyn[(long long)D_V * n]; p.r = dyn[(long long)D_R * n];
    c.xd_l = dyn[(long long)D_XD_L * n]; c.yd_l = dyn[(long long)D_YD_L * n]; c.psid_l = dyn[(long long)D_PSID_L * n];
    c.ud_l = dyn[(long long)D_UD_L * n]; c.vd_l = dyn[(long long)D_VD_L * n]; c.rd_l = dyn[(long long)D_RD_L * n];
    c.eu_l = dyn[(long long)D_EU_L * n]; c.kadu_l = dyn[(long long)D_KADU_L * n]; c.kadp_l = dyn[(long long)D_KADP_L * n];
    c.psi_d_l = dyn[(long long)D_PSI_D_L * n]; c.odd_l = dyn[(long long)D_ODD_L * n]; c.od_l = dyn[(long long)D_OD_L * n];
    c.o_l = dyn[(long long)D_O_L * n]; c.eu_int = dyn[(long long)D_EU_INT * n]; c.ka_u = dyn[(long long)D_KA_U * n];
    c.ka_psi = dyn[(long long)D_KA_PSI * n];
}
__device__ __forceinline__ void ca_store(double* dyn, long long n, const Pose& p, const Ctrl& c) {
    dyn[(long long)D_X * n] = p.x; dyn[(long long)D_Y * n] = p.y; dyn[(long long)D_PSI * n] = p.psi;
    dyn[(long long)D_U * n] = p.u; dyn[(long long)D_V * n] = p.v; dyn[(long long)D_R * n] = p.r;
    dyn[(long long)D_XD_L * n] = c.xd_l; dyn[(long long)D_YD_L * n] = c.yd_l; dyn[(long long)D_PSID_L * n] = c.psid_l;
    dyn[(long long)D_UD_L * n] = c.ud_l; dyn[(long long)D_VD_L * n] = c.vd_l; dyn[(long long)D_RD_L * n] = c.rd_l;
    dyn[(long long)D_EU_L * n] = c.eu_l; dyn[(long long)D_KADU_L * n] = c.kadu_l; dyn[(long long)D_KADP_L * n] = c.kadp_l;
    dyn[(long long)D_PSI_D_L * n] = c.psi_d_l; dyn[(long long)D_ODD_L * n] = c.odd_l; dyn[(long long)D_OD_L * n] = c.od_l;
    dyn[(long long)D_O_L * n] = c.o_l; dyn[(long long)D_EU_INT * n] = c.eu_int; dyn[(long long)D_KA_U * n] = c.ka_u;
    dyn[(long long)D_KA_PSI * n] = c.ka_psi;
}

struct CaGeo {  // obstacle geometry of the CTA's envs during the ray tests: [slot][thread]
    double bx[CA_MAX_OBST][CA_B], by[CA_MAX_OBST][CA_B], rr2[CA_MAX_OBST][CA_B], key[CA_MAX_OBST][CA_B];
};

struct CaOut {
    float obs[CA_OBS];
    double reward;
    bool done, trunc;
};

// UsvAsmcCaEnv.step after the action has been denormalised (usv_asmc_ca_env.py:195-311)
__device__ __forceinline__ void ca_step_core(const CaPtrs& s, long long i, Pose& p, Ctrl& c, double& pa0, double& pa1,
                                             const double tx, const double ty, const int nob, const double a0,
                                             const double a1, CaGeo& geo, CaOut& out) {
    const long long n = s.n;
    const int t = threadIdx.x;
    // :195-205 -> usv_asmc.py:53-244
    double sn, cs;
    sincos(p.psi, &sn, &cs);
#pragma unroll 1
    for (int k = 0; k < 10; ++k) asmc_substep<false, false>(p, c, a0, a1, cs, sn, 0, nullptr);
    sincos(p.psi, &sn, &cs);  // the epilogue wants the exact pair of the final heading
    const double x = p.x, y = p.y, psi = p.psi;

    // :229-246 collision; keys of the sensor order (:236, 414-416: the shift by boat_radius does not change the order)
    double dmin = INFINITY;
    const double* ca = s.ca + i;
    for (int j = 0; j < nob; ++j) {
        const double ox = ca[(long long)(C_OX + j) * n], oy = ca[(long long)(C_OY + j) * n], rr = ca[(long long)(C_OR + j) * n];
        const double dx = ox - x, dy = oy - y;
        const double key = (sqrt_k(fma(dx, dx, dy * dy)) - rr) - 0.1;
        dmin = fmin(dmin, key);
        geo.bx[j][t] = fma(cs, dx, sn * dy);       // offset in the body frame
        geo.by[j][t] = fma(cs, dy, -(sn * dx));
        geo.rr2[j][t] = rr * rr;
        geo.key[j][t] = key;
    }
    const bool collision = nob > 0 && dmin < 0.0;

    // :249-259, 439-461: per ray the hitting obstacle with the smallest key (lower slot wins a tie: stable argsort)
    // :284-289 observation
    const double tdx = tx - x, tdy = ty - y;
    const double angle_to_target = ca_wrap(atan2(tdy, tdx) - psi);  // :405-409
    const double te0 = fma(cs, tdx, sn * tdy), te1 = fma(cs, tdy, -(sn * tdx));  // :264-266 (the psi column meets a 0)
    const double te2 = ca_wrap(angle_to_target);
    const double terr = sqrt_k(fma(te0, te0, te1 * te1));
    out.obs[0] = (float)(p.u / 1.25);  // max_u = 2.5 / 2
    out.obs[1] = (float)(p.r / 3.5);
    out.obs[2] = (float)(te0 / 1000.0);  // max_x^2 + max_y^2
    out.obs[3] = (float)(te1 / 1000.0);
    out.obs[4] = (float)(te2 / PI);
    out.obs[5] = (float)(pa0 / PI);  // the previous action: average of a 1-deep history / max(max_action0, max_action1)
    out.obs[6] = (float)(pa1 / PI);
#pragma unroll 1
    for (int k = 0; k < CA_SENS; ++k) {
        const double2 ray = CA_RAYTAB[k];
        double bkey = INFINITY, bh = 100.0;
        for (int j = 0; j < nob; ++j) {
            const double bx = geo.bx[j][t], by = geo.by[j][t];
            const double along = fma(ray.x, bx, ray.y * by);
            const double perp = fma(ray.y, bx, -(ray.x * by));
            const double delta = fma(-perp, perp, geo.rr2[j][t]);
            if (!(along >= 0.0 && delta >= 0.0)) continue;       // :448, :454
            const double h = along - sqrt_k(delta);              // :457
            const double key = geo.key[j][t];
            if (h < 100.0 && key < bkey) {                       // :458
                bkey = key;
                bh = h;
            }
        }
        out.obs[7 + k] = (float)(bh / 100.0);
    }
    pa0 = a0;  // :294
    pa1 = a1;

    // :276-282, 485-498 reward; :296-311 flags
    double reward = (-terr / 75 - fabs(angle_to_target / PI)) + sqrt_k(fma(p.u, p.u, p.v * p.v)) * 0.5;
    bool done = sqrt_k(fma(tdx, tdx, tdy * tdy)) < 1.5;  // arrived
    bool trunc = collision;
    if (terr > 40.0) {
        done = true;
        reward -= 100.0;
    }
    if (fmax(fmax(fabs(x), fabs(y)), fabs(psi)) > 100.0) {
        done = true;
        trunc = true;
    }
    out.reward = reward;
    out.done = done;
    out.trunc = trunc;
}

// reset() (:327-403) with Philox draws keyed by (seed, GLOBAL env id, episode): block 0 -> x, y | 1 -> theta, target x |
// 2 -> target y, num_obs | 3 + 2 j -> radius of candidate j | 4 + 2 j -> Box-Muller pair of its (x, y).  Leaves the
// fresh pose / controller in p, c and the target / obstacles in HBM; the caller then runs step([-1, 0]) (:402).
// opt: reset(options=...) (:358-372), has_* all 0 for a plain reset; applied after the draws, as in the reference.
__device__ __forceinline__ void ca_reset_draws(const CaPtrs& s, long long i, Pose& p, Ctrl& c, double& pa0, double& pa1,
                                               double& tx, double& ty, int& nob, const UsvCaResetOptions& opt) {
    const long long n = s.n;
    const int ep = s.episode[i] + 1;
    s.episode[i] = ep;
    const DrawKey dk(s.seed, (unsigned long long)(s.global_offset + i), (uint32_t)ep);
    double a, b;
    dk.pair(0, a, b);
    // np.random.uniform(lo, hi) = lo + (hi - lo) u, product and sum rounded separately (no fma contraction)
    auto lin = [](double lo, double span, double u01) { return __dadd_rn(lo, __dmul_rn(span, u01)); };
    const double x = lin(-10.0, 30.0 - -10.0, a);         // :331
    const double y = lin(-10.0, (-10.0 + 5) - -10.0, b);  // :332
    dk.pair(1, a, b);
    const double theta = lin(-PI / 4, PI / 4 - -PI / 4, a);  // :333
    tx = lin(-10.0, (30.0 - 10) - -10.0, b);                 // :343-346
    dk.pair(2, a, b);
    ty = lin(10.0 - 5, (10.0 - 1) - (10.0 - 5), a);
    int num_obs = (int)lin(2.0, 10 - 2, b);                  // :349
    const double cx = (x + tx) / 2, cy = (y + ty) / 2;      // :353 (the drawn target)
    if (opt.has_obstacles) num_obs = opt.n_obstacles;       // :361-365
    if (opt.has_target) {                                   // :367-368
        tx = opt.target[0];
        ty = opt.target[1];
    }
    const double px = opt.has_start ? opt.start[0] : x, py = opt.has_start ? opt.start[1] : y;  // :370-372 self.position
    double* ca = s.ca + i;
    int kept = 0;
    for (int j = 0; j < num_obs; ++j) {
        double rr, ox, oy;
        if (opt.has_obstacles) {
            rr = opt.obs_r[j];
            ox = opt.obs_x[j];
            oy = opt.obs_y[j];
        } else {
            double ur, unused, u1, u2;
            dk.pair(3 + 2 * j, ur, unused);
            dk.pair(4 + 2 * j, u1, u2);
            rr = lin(1.0, 2 - 1, ur);                       // :354
            const double rad = sqrt(-2.0 * log(1.0 - u1));
            double sa, ca_;
            sincos(2.0 * PI * u2, &sa, &ca_);
            ox = lin(cx, 10.0, __dmul_rn(rad, ca_));        // :355-356
            oy = lin(cy, 10.0, __dmul_rn(rad, sa));
        }
        const double margin = 0.3 + 0.35;  // :374-399 hypot - r - boat_radius - (safety_radius + 0.35), left to right
        if (((hypot(ox - px, oy - py) - rr) - 0.1) - margin < 0) continue;
        if (((hypot(ox - tx, oy - ty) - rr) - 0.1) - margin < 0) continue;
        ca[(long long)(C_OX + kept) * n] = ox;
        ca[(long long)(C_OY + kept) * n] = oy;
        ca[(long long)(C_OR + kept) * n] = rr;
        ++kept;
    }
    for (int j = kept; j < CA_MAX_OBST; ++j)
        ca[(long long)(C_OX + j) * n] = ca[(long long)(C_OY + j) * n] = ca[(long long)(C_OR + j) * n] = 0.0;
    ca[(long long)C_TX * n] = tx;
    ca[(long long)C_TY * n] = ty;
    s.n_obs[i] = kept;
    nob = kept;
    p = {x, y, theta, 0.0, 0.0, 0.0};  // :336 DynamicModel(x, y, theta), :385 a fresh controller
    c = {};
    pa0 = pa1 = 0.0;                    // :341
}

__device__ __forceinline__ void ca_write_obs(float* __restrict__ dst, const float* src) {
#pragma unroll
    for (int k = 0; k < CA_OBS; ++k) dst[k] = src[k];
}

// RESET: Env.reset() of the masked envs (draws + the internal step([-1, 0])), else one VecEnv.step() with SB3 auto-reset
template <bool RESET>
__global__ void __launch_bounds__(CA_B, CA_MIN_CTAS)
ca_kernel(CaPtrs s, StepFlags fl, const uint8_t* __restrict__ mask, const float* __restrict__ actions,
          float* __restrict__ obs, float* __restrict__ reward, uint8_t* __restrict__ terminated,
          uint8_t* __restrict__ truncated, float* __restrict__ terminal_obs, const UsvCaResetOptions opt) {
    __shared__ CaGeo geo;
    const long long n = s.n;
    const long long i = fl.lo + (long long)blockIdx.x * CA_B + threadIdx.x;
    if (i >= fl.hi) return;
    pdl_wait();  // usv_state.cuh
    if (RESET && mask && !mask[i]) return;
    Pose p;
    Ctrl c;
    double pa0, pa1, tx, ty;
    int nob, steps = 0;
    CaOut out;
    bool finished = RESET;
    if (!RESET) {
        ca_load(s.dyn + i, n, p, c);
        const double* ca = s.ca + i;
        tx = ca[(long long)C_TX * n];
        ty = ca[(long long)C_TY * n];
        pa0 = ca[(long long)C_PA0 * n];
        pa1 = ca[(long long)C_PA1 * n];
        nob = s.n_obs[i];
        steps = s.steps[i] + 1;
        const float2 af = reinterpret_cast<const float2*>(actions)[i];
        const double a0 = (double)ca_denorm_f32(af.x, 2.0f, -1.0f);
        const double a1 = (double)ca_denorm_f32(af.y, (float)(PI - -PI), (float)-PI);
        ca_step_core(s, i, p, c, pa0, pa1, tx, ty, nob, a0, a1, geo, out);
        const bool tlimit = fl.max_episode_steps > 0 && steps >= fl.max_episode_steps;  // TimeLimit, __init__.py:18-22
        const bool trunc = out.trunc || tlimit;
        finished = out.done || trunc;
        reward[i] = (float)out.reward;
        terminated[i] = out.done ? 1 : 0;
        truncated[i] = trunc ? 1 : 0;
        if (!(isfinite(p.x) && isfinite(p.y) && isfinite(p.psi))) atomicAdd(&s.stats->nonfinite, 1ull);
        if (finished) {
            const double ep_ret = s.ep_return[i] + out.reward;
            s.ep_done[i] = make_double2(ep_ret, (double)steps);
            DeviceStats* ds = s.stats;
            atomicAdd(&ds->return_sum, ep_ret);
            atomicAdd(&ds->length_sum, (double)steps);
            atomicAdd(&ds->episodes, 1ull);
            if (out.done) atomicAdd(&ds->terminated, 1ull);
            else if (out.trunc) atomicAdd(&ds->out_of_bounds, 1ull);  // collision: the env's own truncation
            else atomicAdd(&ds->time_limit, 1ull);
            if (fl.auto_reset) {
                if (terminal_obs) ca_write_obs(terminal_obs + i * CA_OBS, out.obs);
            } else {
                s.ep_return[i] = ep_ret;
                finished = false;
            }
        } else {
            s.ep_return[i] += out.reward;
        }
    }
    if (finished) {  // reset: draws, then the observation of step([-1, 0]) with Python ints -> action [-1.0, 0.0]
        ca_reset_draws(s, i, p, c, pa0, pa1, tx, ty, nob, opt);
        ca_step_core(s, i, p, c, pa0, pa1, tx, ty, nob, -1.0, 0.0, geo, out);
        steps = 0;
        s.ep_return[i] = 0.0;
    }
    ca_store(s.dyn + i, n, p, c);
    s.ca[(long long)C_PA0 * n + i] = pa0;
    s.ca[(long long)C_PA1 * n + i] = pa1;
    s.steps[i] = steps;
    if (obs) ca_write_obs(obs + i * CA_OBS, out.obs);
}

}  // namespace usv
