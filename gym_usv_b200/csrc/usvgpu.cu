// usvgpu.cu -- batched gym-usv environment step for sm_100a + the C ABI of include/usvgpu.h.
//
// One thread owns one environment: the 20 ASMC + 3-DOF sub-steps of a VecEnv.step() run
// register-resident (usv_model.cuh), followed by the kinematic post-step, closest-point /
// cross-track maths, the range-sensor pass, reward, flags and (SB3 semantics) the in-kernel
// reset of finished envs.  A CTA stages its B observation rows ([B][143] float32, contiguous in
// the caller's [N][143] tensor) in shared memory and streams them out as float4; the per-ray
// winner table of the sensor pass lives beside it.  No tensor cores: nothing here is a
// contraction.  Reference citations are relative to the reference root (romi2002/gym-usv).
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <new>
#include <string>

#include "../../include/usvgpu.h"
#include "usv_model.cuh"
#include "usv_philox.cuh"

namespace usv {

constexpr int OBS_DIM = USVGPU_OBS_DIM;
constexpr int N_SENS = USVGPU_SENSORS;
constexpr int MAX_OBST = USVGPU_MAX_OBSTACLES;
constexpr int HDR = 15;              // observation columns before the sensors (simple_env.py:91-96)
constexpr int WIDX_STRIDE = 132;     // bytes per env row of the winner table (33 words: conflict-free columns)
constexpr int RAYS_FULL_CIRCLE = 192;  // 2 pi / pitch, pitch = (4 pi / 3) / 128 (simple_env.py:12-14)

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
// rows of USV_FIELD_EPI
enum EpiRow { E_PSX = 0, E_PSY, E_PEX, E_PEY, E_MA0, E_MA2, E_VREF, N_EPI };
static_assert(N_EPI == USV_EPI_ROWS, "header/kernel row count mismatch");

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
    double* dyn;       // [N_DYN][n]
    double* epi;       // [N_EPI][n]
    void* obst;        // [MAX_OBST][3][n]  double or float
    int* n_obs;        // [n]
    int* steps;        // [n]
    int* episode;      // [n]
    double* ep_return; // [n]
    DeviceStats* stats;
    const double2* raytab;  // [128] cos/sin of the ray angles relative to the bow
    long long n;
    long long global_offset;
    unsigned long long seed;
};

struct StepFlags {
    int do_perturb, skip_kinematic, max_episode_steps, auto_reset;
};

template <typename ObT>
__device__ __forceinline__ double ld_obst(const void* base, long long n, int slot, int f, long long i) {
    return (double)static_cast<const ObT*>(base)[((long long)slot * 3 + f) * n + i];
}
template <typename ObT>
__device__ __forceinline__ void st_obst(void* base, long long n, int slot, int f, long long i, double v) {
    static_cast<ObT*>(base)[((long long)slot * 3 + f) * n + i] = (ObT)v;
}

// ---------------------------------------------------------------------------------------------
// Reset of one env: simple_env_asmc.py:14-16 -> simple_env.py:228-308 with Philox draws.
// Writes the whole state of env i and observation columns 0..14 into hdr[0..14].
// ---------------------------------------------------------------------------------------------
template <typename ObT>
__device__ void reset_env(const StatePtrs& s, long long i, float* hdr) {
    const long long n = s.n;
    const int ep = s.episode[i] + 1;
    s.episode[i] = ep;
    s.steps[i] = 0;
    s.ep_return[i] = 0.0;
    DrawStream d(s.seed, (unsigned long long)(s.global_offset + i), (uint32_t)ep);

    double z0, z1;
    d.normal2(z0, z1);
    const double psx = 0.5 * z0 + ENV_HI / 2, psy = 0.5 * z1 + ENV_HI / 2;  // :234-235
    // :236-237 draws a position that :238 overwrites; not drawn
    const double psi = d.uniform(-PI, PI);                                 // :238
    const double angle = d.uniform(-PI, PI);                               // :241
    const double dist = d.uniform(100.0, 110.0);                           // :242
    double sa, ca;
    sincos(angle, &sa, &ca);
    const double pex = psx + ca * dist, pey = psy + sa * dist;             // :243
    const double tx = d.uniform(0.0, ENV_HI), ty = d.uniform(0.0, ENV_HI); // :245
    const double u = d.uniform(0.0, 0.15), v = d.uniform(0.0, 0.15), r = d.uniform(0.0, 0.15);  // :246
    const double ma0 = d.uniform(1.50, 3.0);                               // :249
    (void)d.uniform(1.50, 3.0);                                            // max_action[1], zeroed at :254
    (void)d.uniform(1.50, 3.0);                                            // max_action[2], redrawn at :250
    const double ma2 = d.uniform(3.0, 6.0);                                // :250
    const double vref = d.uniform(0.75, ma0);                              // :251
    int nob = 15 + (int)floor(d.u01() * 15.0);                             // :257 integers(15, 30)
    nob = min(nob, 29);
    int kept = 0;
    for (int j = 0; j < nob; ++j) {                                        // :258-268
        double ox = d.uniform(0.0, ENV_HI), oy = d.uniform(0.0, ENV_HI);
        ox = (double)(ObT)ox;
        oy = (double)(ObT)oy;
        const double dp = hypot(psx - ox, psy - oy), dt = hypot(tx - ox, ty - oy);
        if (dp < 0.5 || dt < 0.5) continue;
        st_obst<ObT>(s.obst, n, kept, 0, i, ox);
        st_obst<ObT>(s.obst, n, kept, 1, i, oy);
        ++kept;
    }
    if (kept == 0) {                                                       // :270-274
        st_obst<ObT>(s.obst, n, 0, 0, i, d.uniform(0.0, ENV_HI));
        st_obst<ObT>(s.obst, n, 0, 1, i, d.uniform(0.0, ENV_HI));
        kept = 1;
    }
    for (int j = 0; j < kept; ++j) st_obst<ObT>(s.obst, n, j, 2, i, d.uniform(0.15, 0.5));  // :290
    s.n_obs[i] = kept;

    double* dyn = s.dyn + i;
#pragma unroll
    for (int k = 0; k < N_DYN; ++k) dyn[(long long)k * n] = 0.0;          // fresh UsvAsmc(), progress = 0
    dyn[(long long)D_X * n] = psx;
    dyn[(long long)D_Y * n] = psy;
    dyn[(long long)D_PSI * n] = psi;
    dyn[(long long)D_U * n] = u;
    dyn[(long long)D_V * n] = v;
    dyn[(long long)D_R * n] = r;
    double* epi = s.epi + i;
    epi[(long long)E_PSX * n] = psx;
    epi[(long long)E_PSY * n] = psy;
    epi[(long long)E_PEX * n] = pex;
    epi[(long long)E_PEY * n] = pey;
    epi[(long long)E_MA0 * n] = ma0;
    epi[(long long)E_MA2 * n] = ma2;
    epi[(long long)E_VREF * n] = vref;

    if (hdr) {  // :302  _get_obs(zeros(3)) with the random target and position == path_start
        const double ang = wrap_pi(atan2(ty - psy, tx - psx) - psi);
        hdr[0] = (float)(u / 10);
        hdr[1] = (float)(v / 10);
        hdr[2] = (float)(r / 10);
        hdr[3] = (float)(ang / PI);
        hdr[4] = (float)(hypot(psx - tx, psy - ty) / DIST_NORM);
        hdr[5] = 0.0f;  // ye: the boat sits on path_start
        hdr[6] = (float)(vref / 10);
        hdr[7] = 0.0f;
        hdr[8] = 0.0f;
        hdr[9] = (float)(ma0 / 10);
        hdr[10] = 0.0f;
        hdr[11] = (float)(ma2 / 10);
        hdr[12] = (float)(MAX_ACC_U / 10);
        hdr[13] = 0.0f;
        hdr[14] = (float)(MAX_ACC_R / 10);
    }
}

template <typename ObT>
__global__ void reset_kernel(StatePtrs s, const uint8_t* __restrict__ mask, float* __restrict__ obs) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s.n) return;
    if (mask && !mask[i]) return;
    float hdr[HDR];
    reset_env<ObT>(s, i, hdr);
    if (obs) {
#pragma unroll
        for (int k = 0; k < HDR; ++k) obs[i * OBS_DIM + k] = hdr[k];
    }
}

// ---------------------------------------------------------------------------------------------
// Range-sensor pass for one env (usv_asmc_ca_env.py:411-461,500-519 as called from
// simple_env.py:203-226).  The reference walks obstacles in ascending (centre distance - radius)
// order and takes the FIRST one each ray hits; the equivalent sort-free rule is "among the
// obstacles a ray hits, the one with the smallest key wins" (SURVEY.md H2).  Each obstacle only
// subtends +-asin(r/d) so only the few rays inside that wedge (found in float32, widened by a
// margin) get the reference's exact float64 hit test.
//   put(ray, hit)  stores a reading;  wrow = per-ray winner slot (0xFF = none).
// Returns min over obstacles of (centre distance - radius); min_reading <- smallest stored hit.
// ---------------------------------------------------------------------------------------------
template <typename ObT>
__device__ __forceinline__ bool ray_hit(const double2* __restrict__ raytab, int ri, double cps, double sps, double dx,
                                        double dy, double rr, double& hit) {
    const double2 t = raytab[ri];
    const double c = cps * t.x - sps * t.y;  // cos(psi + phi_i)
    const double s = sps * t.x + cps * t.y;  // sin(psi + phi_i)
    const double along = c * dx + s * dy;    // usv_asmc_ca_env.py:516
    if (along < 0.0) return false;           // :448
    const double perp = s * dx - c * dy;     // :518
    const double delta = rr * rr - perp * perp;  // :453
    if (delta < 0.0) return false;           // :454
    hit = along - sqrt(delta);               // :457
    return hit < SENSOR_MAX_RANGE;           // :458
}

template <typename ObT, typename Put>
__device__ __forceinline__ double sensor_pass(const StatePtrs& s, long long i, double x, double y, double psi,
                                              const double2* __restrict__ raytab, uint8_t* wrow, Put put,
                                              double& min_reading) {
    const long long n = s.n;
    const int nob = s.n_obs[i];
    double cps, sps;
    sincos(psi, &sps, &cps);
    const float fpsi = (float)wrap_pi(psi);
    double min_key = INFINITY;
    double min_written = SENSOR_MAX_RANGE;
    for (int j = 0; j < nob; ++j) {
        const double ox = ld_obst<ObT>(s.obst, n, j, 0, i), oy = ld_obst<ObT>(s.obst, n, j, 1, i);
        const double rr = ld_obst<ObT>(s.obst, n, j, 2, i);
        const double dx = ox - x, dy = oy - y;  // usv_asmc_ca_env.py:508
        const double d = sqrt(fma(dx, dx, dy * dy));
        const double key = d - rr;              // simple_env.py:205-206
        min_key = fmin(min_key, key);
        int klo, khi;
        if (d <= rr) {  // boat inside the circle (only on a terminal step): every ray may hit
            klo = 0;
            khi = N_SENS - 1;
        } else {
            float rel = atan2f((float)dy, (float)dx) - fpsi;  // bearing relative to the bow
            rel -= 6.2831853f * rintf(rel * 0.15915494f);
            const float cf = (rel + 2.0943951f) * INV_PITCH_F;  // ray index of the bearing, in [-32, 160]
            const float w = asinf(fminf((float)rr / (float)d, 1.0f)) * INV_PITCH_F + 0.05f;
            klo = (int)ceilf(cf - w);
            khi = (int)floorf(cf + w);
        }
        for (int k = klo; k <= khi; ++k) {
            int ri = k;
            if (ri < 0) ri += RAYS_FULL_CIRCLE;
            else if (ri >= RAYS_FULL_CIRCLE) ri -= RAYS_FULL_CIRCLE;
            if (ri >= N_SENS) continue;  // the fan only spans 4 pi / 3
            double hit;
            if (!ray_hit<ObT>(raytab, ri, cps, sps, dx, dy, rr, hit)) continue;
            const int cur = wrow[ri];
            bool take = true;
            if (cur != 0xFF) {  // another obstacle already claimed this ray: smaller key wins
                const double cx = ld_obst<ObT>(s.obst, n, cur, 0, i) - x, cy = ld_obst<ObT>(s.obst, n, cur, 1, i) - y;
                const double ckey = sqrt(fma(cx, cx, cy * cy)) - ld_obst<ObT>(s.obst, n, cur, 2, i);
                take = key < ckey;
            }
            if (take) {
                wrow[ri] = (uint8_t)j;
                put(ri, hit);
                min_written = fmin(min_written, hit);
            }
        }
    }
    // min over the FINAL readings.  min_written is a lower bound (it also saw overwritten
    // readings); only when it could matter for the collision reward (< 0.2, simple_env.py:155) is
    // the exact minimum recomputed from the winners.
    if (min_written < 0.2) {
        double m = SENSOR_MAX_RANGE;
        for (int ri = 0; ri < N_SENS; ++ri) {
            const int cur = wrow[ri];
            if (cur == 0xFF) continue;
            const double dx = ld_obst<ObT>(s.obst, n, cur, 0, i) - x, dy = ld_obst<ObT>(s.obst, n, cur, 1, i) - y;
            double hit;
            if (ray_hit<ObT>(raytab, ri, cps, sps, dx, dy, ld_obst<ObT>(s.obst, n, cur, 2, i), hit)) m = fmin(m, hit);
        }
        min_written = m;
    }
    min_reading = min_written;
    return min_key;
}

// ---------------------------------------------------------------------------------------------
// The step kernel.  B threads = B envs per CTA.
// ---------------------------------------------------------------------------------------------
template <int B, bool OBST>
constexpr size_t step_smem_bytes() {
    return (size_t)B * OBS_DIM * sizeof(float) + (OBST ? (N_SENS * sizeof(double2) + (size_t)B * WIDX_STRIDE) : 0);
}

template <int B, bool OBST, typename ObT>
__global__ void __launch_bounds__(B)
step_kernel(StatePtrs s, StepFlags fl, const float* __restrict__ actions, float* __restrict__ obs,
            float* __restrict__ reward, uint8_t* __restrict__ terminated, uint8_t* __restrict__ truncated,
            float* __restrict__ terminal_obs) {
    static_assert(B % 4 == 0, "B*143*4 bytes must be a multiple of 16 for the float4 copy-out");
    extern __shared__ __align__(16) unsigned char smem[];
    float* tile = reinterpret_cast<float*>(smem);
    double2* raytab = reinterpret_cast<double2*>(smem + (size_t)B * OBS_DIM * sizeof(float));
    uint8_t* widx = reinterpret_cast<uint8_t*>(raytab + N_SENS);

    const int tid = threadIdx.x;
    const long long n = s.n;
    const long long base = (long long)blockIdx.x * B;
    const long long i = base + tid;
    const int nvalid = (int)min((long long)B, n - base);

    // ---- stage: every sensor reads max range (1.0 after /100), no ray has a winner -----------
    {
        float4* t4 = reinterpret_cast<float4*>(tile);
        for (int k = tid; k < B * OBS_DIM / 4; k += B) t4[k] = make_float4(1.f, 1.f, 1.f, 1.f);
        if (OBST) {
            uint32_t* w4 = reinterpret_cast<uint32_t*>(widx);
            for (int k = tid; k < B * WIDX_STRIDE / 4; k += B) w4[k] = 0xFFFFFFFFu;
            for (int k = tid; k < N_SENS; k += B) raytab[k] = s.raytab[k];
        }
    }
    __syncthreads();

    if (i < n) {
        // ---- load state (coalesced SoA columns) ---------------------------------------------
        const double* dyn = s.dyn + i;
        Pose p;
        Ctrl c;
        p.x = dyn[(long long)D_X * n];
        p.y = dyn[(long long)D_Y * n];
        p.psi = dyn[(long long)D_PSI * n];
        p.u = dyn[(long long)D_U * n];
        p.v = dyn[(long long)D_V * n];
        p.r = dyn[(long long)D_R * n];
        c.xd_l = dyn[(long long)D_XD_L * n];
        c.yd_l = dyn[(long long)D_YD_L * n];
        c.psid_l = dyn[(long long)D_PSID_L * n];
        c.ud_l = dyn[(long long)D_UD_L * n];
        c.vd_l = dyn[(long long)D_VD_L * n];
        c.rd_l = dyn[(long long)D_RD_L * n];
        c.eu_l = dyn[(long long)D_EU_L * n];
        c.kadu_l = dyn[(long long)D_KADU_L * n];
        c.kadp_l = dyn[(long long)D_KADP_L * n];
        c.psi_d_l = dyn[(long long)D_PSI_D_L * n];
        c.odd_l = dyn[(long long)D_ODD_L * n];
        c.od_l = dyn[(long long)D_OD_L * n];
        c.o_l = dyn[(long long)D_O_L * n];
        c.eu_int = dyn[(long long)D_EU_INT * n];
        c.ka_u = dyn[(long long)D_KA_U * n];
        c.ka_psi = dyn[(long long)D_KA_PSI * n];
        double progress = dyn[(long long)D_PROGRESS * n];
        const double* epi = s.epi + i;
        const double psx = epi[(long long)E_PSX * n], psy = epi[(long long)E_PSY * n];
        const double pex = epi[(long long)E_PEX * n], pey = epi[(long long)E_PEY * n];
        const double ma0 = epi[(long long)E_MA0 * n], ma2 = epi[(long long)E_MA2 * n];
        const double vref = epi[(long long)E_VREF * n];
        int steps = s.steps[i];
        const float2 af = reinterpret_cast<const float2*>(actions)[i];
        const double a0 = (double)af.x, a1 = (double)af.y;

        // ---- 2 x UsvAsmc.compute = 20 sub-steps (simple_env_asmc.py:19-25) -------------------
        const long long pstep0 = (long long)steps * 20;
        if (fl.do_perturb) {
            for (int k = 0; k < 20; ++k) {
                double sn, cs;
                sincos(p.psi, &sn, &cs);
                asmc_substep<true, false>(p, c, a0, a1, cs, sn, pstep0 + k, nullptr);
            }
        } else {
            for (int k = 0; k < 20; ++k) {
                double sn, cs;
                sincos(p.psi, &sn, &cs);
                asmc_substep<false, false>(p, c, a0, a1, cs, sn, 0, nullptr);
            }
        }

        // ---- UsvSimpleEnv.step(zeros(2)): kinematic post-step, simple_env.py:311-324 --------
        // action = 0.8*last_action + 0.2*(max_action*0) and last_action is 0 for the whole life
        // of this env, so the commanded velocity is 0.
        if (!fl.skip_kinematic) {
            const double du = np_clip(0.0 - p.u, -MAX_ACC_U, MAX_ACC_U);
            p.u = np_clip(p.u + du, -ma0, ma0);
            p.v = 0.0;  // clip(v + clip(-v, -0, 0), -0, 0): max_acceleration[1] = max_action[1] = 0
            const double dr = np_clip(0.0 - p.r, -MAX_ACC_R, MAX_ACC_R);
            p.r = np_clip(p.r + dr, -ma2, ma2);
            if (p.u != 0.0) {
                double sn, cs;
                sincos(p.psi, &sn, &cs);
                p.x = p.x + (p.u * cs) * DT;
                p.y = p.y + (p.u * sn) * DT;
            }
            p.psi = p.psi + p.r * DT;
        }

        // ---- closest point on the path, simple_env.py:139-148 --------------------------------
        const double pdx = pex - psx, pdy = pey - psy;
        const double det = pdx * pdx + pdy * pdy;
        double a = (pdy * (p.y - psy) + pdx * (p.x - psx)) / det;
        a += (0.005 / 10) * ENV_HI;
        a = np_clip(a, progress, 1.0);
        const double tgx = psx + a * pdx, tgy = psy + a * pdy;
        progress = a;

        // ---- sensors, simple_env.py:203-226 --------------------------------------------------
        float* row = tile + tid * OBS_DIM;
        double min_key = 1.0, min_reading = SENSOR_MAX_RANGE;
        if (OBST) {
            float* srow = row + HDR;
            min_key = sensor_pass<ObT>(s, i, p.x, p.y, p.psi, raytab, widx + tid * WIDX_STRIDE,
                                       [srow](int ri, double hit) { srow[ri] = (float)(hit * 0.01); }, min_reading);
        }

        // ---- flags, simple_env.py:334-336 + TimeLimit ----------------------------------------
        const bool term = OBST && (min_key < 0.05);
        const bool oob = (p.x > ENV_HI) || (p.y > ENV_HI) || (p.x < 0.0) || (p.y < 0.0);
        steps += 1;
        const bool tlimit = (fl.max_episode_steps > 0) && (steps >= fl.max_episode_steps);
        const bool trunc = oob || tlimit;

        // ---- observation header, simple_env.py:72-96 ----------------------------------------
        const double inv_len = rsqrt(det);
        const double ye = -(p.x - psx) * (pdy * inv_len) + (p.y - psy) * (pdx * inv_len);  // :133-137
        const double ang = wrap_pi(atan2(tgy - p.y, tgx - p.x) - p.psi);                  // :67-69
        const double tdx = p.x - tgx, tdy = p.y - tgy;
        const double tdist = sqrt(fma(tdx, tdx, tdy * tdy));
        row[0] = (float)(p.u / 10);
        row[1] = (float)(p.v / 10);
        row[2] = (float)(p.r / 10);
        row[3] = (float)(ang / PI);
        row[4] = (float)(tdist / DIST_NORM);
        row[5] = (float)(ye / 10);
        row[6] = (float)(vref / 10);
        row[7] = 0.0f;  // last_action[[0,2]] / max_action[[0,2]]
        row[8] = 0.0f;
        row[9] = (float)(ma0 / 10);
        row[10] = 0.0f;
        row[11] = (float)(ma2 / 10);
        row[12] = (float)(MAX_ACC_U / 10);
        row[13] = 0.0f;
        row[14] = (float)(MAX_ACC_R / 10);

        // ---- reward, simple_env.py:150-201 ----------------------------------------------------
        const double t = ye / 0.075;
        const double at = fabs(t);
        const double ye_reward = exp(-fmin(at, t * t));  // max(exp(-|t|), exp(-t^2)): exp is monotone
        const double angle_reward = exp(-fabs(ang));
        const double vel_reward = exp(-fabs(sqrt(fma(p.u, p.u, p.v * p.v)) - vref)) * 0.05;
        const double collision = (OBST && min_reading < 0.2) ? -20.0 : 0.0;
        const double rew = 0.0 + collision + ye_reward + angle_reward + vel_reward;  // delta_action_reward = -0

        reward[i] = (float)rew;
        terminated[i] = term ? 1 : 0;
        truncated[i] = trunc ? 1 : 0;
        if (!(isfinite(p.x) && isfinite(p.y) && isfinite(p.psi))) atomicAdd(&s.stats->nonfinite, 1ull);

        const bool done = term || trunc;
        const double ep_ret = s.ep_return[i] + rew;
        if (done) {
            DeviceStats* st = s.stats;
            atomicAdd(&st->return_sum, ep_ret);
            atomicAdd(&st->length_sum, (double)steps);
            atomicAdd(&st->episodes, 1ull);
            if (term) atomicAdd(&st->terminated, 1ull);
            if (oob) atomicAdd(&st->out_of_bounds, 1ull);
            if (tlimit && !term && !oob) atomicAdd(&st->time_limit, 1ull);
        }
        if (done && fl.auto_reset) {
            if (terminal_obs) {
                float* dst = terminal_obs + i * OBS_DIM;
                for (int k = 0; k < OBS_DIM; ++k) dst[k] = row[k];
            }
            reset_env<ObT>(s, i, row);  // header <- reset obs; sensor columns stay (simple_env.py:302)
        } else {
            double* dw = s.dyn + i;
            dw[(long long)D_X * n] = p.x;
            dw[(long long)D_Y * n] = p.y;
            dw[(long long)D_PSI * n] = p.psi;
            dw[(long long)D_U * n] = p.u;
            dw[(long long)D_V * n] = p.v;
            dw[(long long)D_R * n] = p.r;
            dw[(long long)D_XD_L * n] = c.xd_l;
            dw[(long long)D_YD_L * n] = c.yd_l;
            dw[(long long)D_PSID_L * n] = c.psid_l;
            dw[(long long)D_UD_L * n] = c.ud_l;
            dw[(long long)D_VD_L * n] = c.vd_l;
            dw[(long long)D_RD_L * n] = c.rd_l;
            dw[(long long)D_EU_L * n] = c.eu_l;
            dw[(long long)D_KADU_L * n] = c.kadu_l;
            dw[(long long)D_KADP_L * n] = c.kadp_l;
            dw[(long long)D_PSI_D_L * n] = c.psi_d_l;
            dw[(long long)D_ODD_L * n] = c.odd_l;
            dw[(long long)D_OD_L * n] = c.od_l;
            dw[(long long)D_O_L * n] = c.o_l;
            dw[(long long)D_EU_INT * n] = c.eu_int;
            dw[(long long)D_KA_U * n] = c.ka_u;
            dw[(long long)D_KA_PSI * n] = c.ka_psi;
            dw[(long long)D_PROGRESS * n] = progress;
            s.steps[i] = steps;
            s.ep_return[i] = ep_ret;
        }
    }
    __syncthreads();

    // ---- stream the CTA's observation rows out: contiguous in obs[N][143] -------------------
    {
        float* gdst = obs + base * OBS_DIM;
        const int total = nvalid * OBS_DIM;
        const int n4 = total / 4;
        const float4* t4 = reinterpret_cast<const float4*>(tile);
        float4* g4 = reinterpret_cast<float4*>(gdst);
        for (int k = tid; k < n4; k += B) g4[k] = t4[k];
        for (int k = n4 * 4 + tid; k < total; k += B) gdst[k] = tile[k];
    }
}

// ---------------------------------------------------------------------------------------------
// Unit-parity kernels
// ---------------------------------------------------------------------------------------------
template <bool PERTURB>
__global__ void asmc_compute_kernel(long long n, double* __restrict__ state, const double* __restrict__ actions,
                                    int n_calls, long long first_substep, double* __restrict__ info) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double* dyn = state + i;
    Pose p;
    Ctrl c;
    double* pf[6] = {&p.x, &p.y, &p.psi, &p.u, &p.v, &p.r};
    double* cf[16] = {&c.xd_l, &c.yd_l, &c.psid_l, &c.ud_l, &c.vd_l, &c.rd_l, &c.eu_l, &c.kadu_l,
                      &c.kadp_l, &c.psi_d_l, &c.odd_l, &c.od_l, &c.o_l, &c.eu_int, &c.ka_u, &c.ka_psi};
#pragma unroll
    for (int k = 0; k < 6; ++k) *pf[k] = dyn[(long long)k * n];
#pragma unroll
    for (int k = 0; k < 16; ++k) *cf[k] = dyn[(long long)(6 + k) * n];
    const double a0 = actions[2 * i], a1 = actions[2 * i + 1];
    double inf[4] = {0, 0, 0, 0};
    long long ps = first_substep;
    for (int k = 0; k < n_calls * 10; ++k, ++ps) {
        double sn, cs;
        sincos(p.psi, &sn, &cs);
        asmc_substep<PERTURB, true>(p, c, a0, a1, cs, sn, ps, inf);
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) dyn[(long long)k * n] = *pf[k];
#pragma unroll
    for (int k = 0; k < 16; ++k) dyn[(long long)(6 + k) * n] = *cf[k];
    if (info) {
#pragma unroll
        for (int k = 0; k < 4; ++k) info[(long long)k * n + i] = inf[k];
    }
}

constexpr int SENS_B = 64;
__global__ void __launch_bounds__(SENS_B)
sensors_kernel(StatePtrs s, const double* __restrict__ pose, double* __restrict__ out) {
    __shared__ double2 raytab[N_SENS];
    __shared__ __align__(4) uint8_t widx[SENS_B * WIDX_STRIDE];
    const int tid = threadIdx.x;
    for (int k = tid; k < N_SENS; k += SENS_B) raytab[k] = s.raytab[k];
    for (int k = tid; k < SENS_B * WIDX_STRIDE / 4; k += SENS_B) reinterpret_cast<uint32_t*>(widx)[k] = 0xFFFFFFFFu;
    __syncthreads();
    const long long n = s.n;
    const long long i = (long long)blockIdx.x * SENS_B + tid;
    if (i >= n) return;
    for (int k = 0; k < N_SENS; ++k) out[(long long)k * n + i] = SENSOR_MAX_RANGE;
    double min_reading;
    double* col = out + i;
    sensor_pass<double>(s, i, pose[i], pose[n + i], pose[2 * n + i], raytab, widx + tid * WIDX_STRIDE,
                        [col, n](int ri, double hit) { col[(long long)ri * n] = hit; }, min_reading);
}

}  // namespace usv

// =================================================================================================
// Host side: handle + C ABI
// =================================================================================================
using namespace usv;

struct UsvGpu {
    UsvConfig cfg;
    int device;
    StatePtrs sp;
    size_t obst_elem;
    std::string err;
    unsigned long long host_env_steps;
};

static thread_local std::string g_create_err;
static double2* g_raytab[64] = {nullptr};  // per device ray table for the stand-alone sensors entry point

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

template <int B, bool OBST, typename ObT>
static cudaError_t launch_step(UsvGpu* h, const StepFlags& fl, const float* actions, float* obs, float* reward,
                               uint8_t* term, uint8_t* trunc, float* tobs, cudaStream_t st) {
    constexpr size_t smem = step_smem_bytes<B, OBST>();
    static bool configured[64] = {false};
    if (!configured[h->device & 63]) {
        cudaError_t e = cudaFuncSetAttribute(step_kernel<B, OBST, ObT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured[h->device & 63] = true;
    }
    const unsigned blocks = (unsigned)((h->sp.n + B - 1) / B);
    step_kernel<B, OBST, ObT><<<blocks, B, smem, st>>>(h->sp, fl, actions, obs, reward, term, trunc, tobs);
    return cudaGetLastError();
}

static int step_impl(UsvGpu* h, const float* actions, float* obs, float* reward, uint8_t* term, uint8_t* trunc,
                     float* tobs, cudaStream_t st) {
    const StepFlags fl = {h->cfg.do_perturb, h->cfg.skip_kinematic, h->cfg.max_episode_steps, h->cfg.auto_reset};
    cudaError_t e;
    if (h->cfg.ignore_obstacles) {
        e = h->cfg.obstacles_f32 ? launch_step<128, false, float>(h, fl, actions, obs, reward, term, trunc, tobs, st)
                                 : launch_step<128, false, double>(h, fl, actions, obs, reward, term, trunc, tobs, st);
    } else {
        e = h->cfg.obstacles_f32 ? launch_step<128, true, float>(h, fl, actions, obs, reward, term, trunc, tobs, st)
                                 : launch_step<128, true, double>(h, fl, actions, obs, reward, term, trunc, tobs, st);
    }
    if (e != cudaSuccess) return fail(h, USV_ERR_CUDA, std::string("step launch: ") + cudaGetErrorString(e));
    h->host_env_steps += (unsigned long long)h->sp.n;
    return USV_OK;
}

static int field_ptr(UsvGpu* h, int32_t field, void** p, int64_t* bytes) {
    const int64_t n = h->sp.n;
    switch (field) {
        case USV_FIELD_DYN: *p = h->sp.dyn; *bytes = n * N_DYN * 8; return 0;
        case USV_FIELD_EPI: *p = h->sp.epi; *bytes = n * N_EPI * 8; return 0;
        case USV_FIELD_OBSTACLES: *p = h->sp.obst; *bytes = n * MAX_OBST * 3 * (int64_t)h->obst_elem; return 0;
        case USV_FIELD_N_OBS: *p = h->sp.n_obs; *bytes = n * 4; return 0;
        case USV_FIELD_STEPS: *p = h->sp.steps; *bytes = n * 4; return 0;
        case USV_FIELD_EPISODE: *p = h->sp.episode; *bytes = n * 4; return 0;
        case USV_FIELD_EP_RETURN: *p = h->sp.ep_return; *bytes = n * 8; return 0;
        default: return -1;
    }
}

extern "C" {

int usvgpu_version(void) { return USVGPU_VERSION; }

const char* usvgpu_last_error(usvgpu_handle h) { return h ? h->err.c_str() : g_create_err.c_str(); }

int usvgpu_create(const UsvConfig* cfg, int device, int64_t n_envs, int64_t global_env_offset, uint64_t seed,
                  usvgpu_handle* out) {
    if (!cfg || !out) return fail(nullptr, USV_ERR_INVALID, "cfg/out is NULL");
    *out = nullptr;
    if (n_envs <= 0) return fail(nullptr, USV_ERR_INVALID, "n_envs must be positive");
    if (cfg->env_kind != USV_ENV_ASMC_SIMPLE) return fail(nullptr, USV_ERR_INVALID, "env_kind not supported by this build");
    if (cfg->precision != USV_PREC_F64) return fail(nullptr, USV_ERR_INVALID, "precision not supported by this build");
    int rc = check_device(nullptr, device);
    if (rc) return rc;
    UsvGpu* h = new (std::nothrow) UsvGpu();
    if (!h) return fail(nullptr, USV_ERR_ALLOC, "host allocation failed");
    h->cfg = *cfg;
    h->device = device;
    h->host_env_steps = 0;
    h->obst_elem = cfg->obstacles_f32 ? sizeof(float) : sizeof(double);
    memset(&h->sp, 0, sizeof(h->sp));
    h->sp.n = n_envs;
    h->sp.global_offset = global_env_offset;
    h->sp.seed = seed;
    const size_t n = (size_t)n_envs;
    cudaError_t e = cudaSetDevice(device);
    auto alloc = [&](void** p, size_t bytes) {
        if (e == cudaSuccess) e = cudaMalloc(p, bytes);
        if (e == cudaSuccess) e = cudaMemset(*p, 0, bytes);
    };
    alloc((void**)&h->sp.dyn, n * N_DYN * sizeof(double));
    alloc((void**)&h->sp.epi, n * N_EPI * sizeof(double));
    alloc((void**)&h->sp.obst, n * MAX_OBST * 3 * h->obst_elem);
    alloc((void**)&h->sp.n_obs, n * sizeof(int));
    alloc((void**)&h->sp.steps, n * sizeof(int));
    alloc((void**)&h->sp.episode, n * sizeof(int));
    alloc((void**)&h->sp.ep_return, n * sizeof(double));
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
    *out = h;
    return USV_OK;
}

int usvgpu_destroy(usvgpu_handle h) {
    if (!h) return USV_OK;
    cudaSetDevice(h->device);
    cudaFree(h->sp.dyn);
    cudaFree(h->sp.epi);
    cudaFree(h->sp.obst);
    cudaFree(h->sp.n_obs);
    cudaFree(h->sp.steps);
    cudaFree(h->sp.episode);
    cudaFree(h->sp.ep_return);
    cudaFree(h->sp.stats);
    cudaFree((void*)h->sp.raytab);
    delete h;
    return USV_OK;
}

int usvgpu_reset(usvgpu_handle h, const uint8_t* mask, float* obs, void* stream) {
    if (!h) return USV_ERR_INVALID;
    CU_TRY(h, cudaSetDevice(h->device));
    const int threads = 128;
    const unsigned blocks = (unsigned)((h->sp.n + threads - 1) / threads);
    cudaStream_t st = (cudaStream_t)stream;
    if (h->cfg.obstacles_f32) reset_kernel<float><<<blocks, threads, 0, st>>>(h->sp, mask, obs);
    else reset_kernel<double><<<blocks, threads, 0, st>>>(h->sp, mask, obs);
    CU_TRY(h, cudaGetLastError());
    return USV_OK;
}

int usvgpu_step(usvgpu_handle h, const float* actions, float* obs, float* reward, uint8_t* terminated,
                uint8_t* truncated, float* terminal_obs, void* stream) {
    if (!h) return USV_ERR_INVALID;
    if (!actions || !obs || !reward || !terminated || !truncated) return fail(h, USV_ERR_INVALID, "NULL buffer");
    if (((uintptr_t)obs & 15) || ((uintptr_t)actions & 7)) return fail(h, USV_ERR_INVALID, "obs must be 16-byte and actions 8-byte aligned");
    CU_TRY(h, cudaSetDevice(h->device));
    return step_impl(h, actions, obs, reward, terminated, truncated, terminal_obs, (cudaStream_t)stream);
}

int usvgpu_rollout(usvgpu_handle h, const float* actions, int32_t n_steps, int32_t action_wrap, float* obs,
                   float* reward, uint8_t* terminated, uint8_t* truncated, void* stream) {
    if (!h) return USV_ERR_INVALID;
    if (!actions || !obs || !reward || !terminated || !truncated || n_steps < 0)
        return fail(h, USV_ERR_INVALID, "bad rollout arguments");
    if (((uintptr_t)obs & 15) || ((uintptr_t)actions & 7)) return fail(h, USV_ERR_INVALID, "obs must be 16-byte and actions 8-byte aligned");
    CU_TRY(h, cudaSetDevice(h->device));
    for (int k = 0; k < n_steps; ++k) {
        const int slab = action_wrap > 0 ? (k % action_wrap) : k;
        int rc = step_impl(h, actions + (size_t)slab * h->sp.n * 2, obs, reward, terminated, truncated, nullptr,
                           (cudaStream_t)stream);
        if (rc) return rc;
    }
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
    CU_TRY(h, cudaSetDevice(h->device));
    CU_TRY(h, cudaMemcpyAsync(buf, p, (size_t)b, cudaMemcpyDefault, (cudaStream_t)stream));
    CU_TRY(h, cudaStreamSynchronize((cudaStream_t)stream));
    return USV_OK;
}

int usvgpu_set_state(usvgpu_handle h, int32_t field, const void* buf, int64_t bytes, void* stream) {
    if (!h || !buf) return USV_ERR_INVALID;
    void* p;
    int64_t b;
    if (field_ptr(h, field, &p, &b)) return fail(h, USV_ERR_INVALID, "unknown field");
    if (bytes != b) return fail(h, USV_ERR_INVALID, "buffer size does not match usvgpu_field_bytes()");
    CU_TRY(h, cudaSetDevice(h->device));
    CU_TRY(h, cudaMemcpyAsync(p, buf, (size_t)b, cudaMemcpyDefault, (cudaStream_t)stream));
    CU_TRY(h, cudaStreamSynchronize((cudaStream_t)stream));
    return USV_OK;
}

int usvgpu_episode_stats(usvgpu_handle h, UsvStats* out, int32_t reset, void* stream) {
    if (!h || !out) return USV_ERR_INVALID;
    CU_TRY(h, cudaSetDevice(h->device));
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
    CU_TRY(nullptr, cudaSetDevice(device));
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
    CU_TRY(nullptr, cudaSetDevice(device));
    if (!g_raytab[device & 63]) {
        double2 t[N_SENS];
        host_raytab(t);
        CU_TRY(nullptr, cudaMalloc((void**)&g_raytab[device & 63], sizeof(t)));
        CU_TRY(nullptr, cudaMemcpy(g_raytab[device & 63], t, sizeof(t), cudaMemcpyHostToDevice));
    }
    StatePtrs sp;
    memset(&sp, 0, sizeof(sp));
    sp.n = n;
    sp.obst = (void*)obstacles;
    sp.n_obs = (int*)n_obs;
    sp.raytab = g_raytab[device & 63];
    const unsigned blocks = (unsigned)((n + SENS_B - 1) / SENS_B);
    sensors_kernel<<<blocks, SENS_B, 0, (cudaStream_t)stream>>>(sp, pose, out);
    CU_TRY(nullptr, cudaGetLastError());
    return USV_OK;
}

}  // extern "C"
