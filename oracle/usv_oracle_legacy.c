/*
 * usv_oracle_legacy.c -- CPU restatement of the legacy path-following env 'usv-asmc-v0'
 * (gym_usv/envs/usv_asmc_env.py: UsvAsmcEnv, one 0.01 s ASMC + 3-DOF sub-step per step(), no obstacles).
 *
 * TEST INFRASTRUCTURE ONLY (see usv_oracle.h).
 *
 * The reference mixes float32 and float64: reset() stores float64 arrays, step() re-stores every
 * array as float32, and under numpy >= 2 a Python float times an np.float32 scalar stays float32.  The oracle is
 * "the reference as executed on this box's numpy (2.3.5)": usv_oracle_legacy_step.inc restates step() once and is
 * compiled for both flows (S = double for the first step of an episode, S = float afterwards).
 * Pinned by tests/golden/legacy_asmc_env.npz (oracle/gen_golden.py::gen_legacy, the reference itself run here);
 * numpy's own SIMD exp / cos / sin differ from glibc's in the last ulp on some inputs, so the pin is a tolerance,
 * stated in tests/test_oracle_golden.py, not bit equality.
 */
#include "usv_oracle.h"

#include <math.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* usv_asmc_env.py:19-51 */
#define L_H 0.01
#define L_MIN_SPEED 0.3
#define L_XUD (-2.25)
#define L_YVD (-23.13)
#define L_YRD (-1.31)
#define L_NVD (-16.41)
#define L_NRD (-2.79)
#define L_YVV (-99.99)
#define L_YVR (-5.49)
#define L_YRV (-5.49)
#define L_YRR (-8.8)
#define L_NVV (-5.49)
#define L_NVR (-8.8)
#define L_NRV (-8.8)
#define L_NRR (-3.49)
#define L_M 30.0
#define L_IZ 4.1
#define L_B 0.41
#define L_C 0.78
#define L_K_U 0.1
#define L_K_PSI 0.2
#define L_KMIN_U 0.05
#define L_KMIN_PSI 0.2
#define L_K2_U 0.02
#define L_K2_PSI 0.1
#define L_MU_U 0.05
#define L_MU_PSI 0.1
#define L_LAMBDA_U 0.001
#define L_LAMBDA_PSI 1.0
#define L_K_AK 5.72
#define L_K_YE 0.5
#define L_SIGMA_YE 1.0
/* usv_asmc_env.py:66-69, 77-78 */
#define L_MAX_YE 10.0
#define L_MAX_X 30.0
#define L_MIN_X (-10.0) /* usv_asmc_ye_int_env.py:63, usv_pid_env.py:60 */
#define L_W_ACTION 0.2

/* :132-133  1.1+0.0045*(1.01/0.09) - 0.1*(0.27/0.09)+0.016*(np.power((0.27/0.09), 2)) -- an np.float64 */
static const double L_YV_K = 1.1 + 0.0045 * (1.01 / 0.09) - 0.1 * (0.27 / 0.09) + 0.016 * ((0.27 / 0.09) * (0.27 / 0.09));
/* :77  1. / np.power((max_action/2 - min_action/2) / integral_step, 2) */
static const double L_C_ACTION =
    1. / ((((M_PI / 2) / 2 - (-M_PI / 2) / 2) / L_H) * (((M_PI / 2) / 2 - (-M_PI / 2) / 2) / L_H));

static inline double l_sign(double x) { return (double)((x > 0.0) - (x < 0.0)); }
/* :124,148,229,232  np.where(|a| > pi, sign(a) * (|a| - 2 pi), a) */
static inline double l_wrap(double a) { return (fabs(a) > M_PI) ? l_sign(a) * (fabs(a) - 2 * M_PI) : a; }
/* :187-189, 214  np.linalg.inv(M): block-diagonal, closed form */
static void l_mass_inverse(double Minv[3][3]) {
    const double m00 = L_M - L_XUD, m11 = L_M - L_YVD, m12 = 0 - L_YRD, m21 = 0 - L_NVD, m22 = L_IZ - L_NRD;
    const double det = m11 * m22 - m12 * m21;
    memset(Minv, 0, 9 * sizeof(double));
    Minv[0][0] = 1.0 / m00;
    Minv[1][1] = m22 / det;
    Minv[1][2] = -m12 / det;
    Minv[2][1] = -m21 / det;
    Minv[2][2] = m11 / det;
}

/* float64 flow: the first step after reset() */
#define LEGACY_VARIANT 0
#define ARR_T float
#define STORE(x) ((double)(float)(x))
#define S double
#define LEGACY_FN(n) n##_f64
#define S_ABS fabs
#define S_SQRT sqrt
#define S_COS cos
#define S_SIN sin
#include "usv_oracle_legacy_step.inc"
#undef S
#undef LEGACY_FN
#undef S_ABS
#undef S_SQRT
#undef S_COS
#undef S_SIN

/* float32 flow: all later steps */
#define S float
#define LEGACY_FN(n) n##_f32
#define S_ABS fabsf
#define S_SQRT sqrtf
#define S_COS cosf
#define S_SIN sinf
#include "usv_oracle_legacy_step.inc"
#undef S
#undef LEGACY_FN
#undef S_ABS
#undef S_SQRT
#undef S_COS
#undef S_SIN
#undef LEGACY_VARIANT
#undef ARR_T
#undef STORE

/* the two float64 siblings: no dtype=np.float32 anywhere, so S = ARR_T = double and nothing is rounded on store */
#define ARR_T double
#define STORE(x) ((double)(x))
#define S double
#define S_ABS fabs
#define S_SQRT sqrt
#define S_COS cos
#define S_SIN sin
#define LEGACY_VARIANT 1
#define LEGACY_FN(n) n##_ye_int
#include "usv_oracle_legacy_step.inc"
#undef LEGACY_VARIANT
#undef LEGACY_FN
#define LEGACY_VARIANT 2
#define LEGACY_FN(n) n##_pid
#include "usv_oracle_legacy_step.inc"
#undef LEGACY_VARIANT
#undef LEGACY_FN
#undef S
#undef S_ABS
#undef S_SQRT
#undef S_COS
#undef S_SIN
#undef ARR_T
#undef STORE

int usv_oracle_sizeof_legacy(void) { return (int)sizeof(UsvOracleLegacy); }

/* usv_asmc_env.py:99-255 */
void usv_oracle_legacy_step(UsvOracleLegacy* e, float action, float obs[6], double* reward, uint8_t* done) {
    if (e->variant == USV_ORACLE_LEGACY_YE_INT) legacy_step_ye_int(e, action, obs, reward, done);
    else if (e->variant == USV_ORACLE_LEGACY_PID) legacy_step_pid(e, action, obs, reward, done);
    else if (e->f32_state) legacy_step_f32(e, action, obs, reward, done);
    else legacy_step_f64(e, action, obs, reward, done);
    e->steps += 1;
    e->ep_return += *reward;
}

/* usv_asmc_env.py:258-300 given the seven uniform draws, in the reference's draw order
 * (x, y, psi, x_0, y_0, x_d, desired_speed) */
void usv_oracle_legacy_reset_from_draws(UsvOracleLegacy* e, const double d[7], float obs[6]) {
    const double x = d[0], y = d[1], psi = d[2], x_0 = d[3], y_0 = d[4], x_d = d[5], desired_speed = d[6];
    const double y_d = y_0;                                   /* :280 */
    const double ak = (double)(float)atan2(y_d - y_0, x_d - x_0); /* :283-284 np.float32(ak) */
    const double psi_ak = (double)(float)l_wrap(psi - ak);    /* :286-288 */
    const double ye = -(x - x_0) * sin(ak) + (y - y_0) * cos(ak); /* :289 */
    const double v_ak = sin(psi_ak) * 0. + cos(psi_ak) * 0.;  /* :291 body_to_path(0, 0, psi_ak) */
    memset(e->state, 0, sizeof(e->state));
    memset(e->velocity, 0, sizeof(e->velocity));
    memset(e->aux, 0, sizeof(e->aux));
    memset(e->last, 0, sizeof(e->last));
    e->state[1] = v_ak;
    e->state[3] = ye;
    e->state[4] = psi_ak;
    e->position[0] = x;
    e->position[1] = y;
    e->position[2] = psi;
    e->target[0] = x_0;
    e->target[1] = y_0;
    e->target[2] = desired_speed;
    e->target[3] = ak;
    e->target[4] = x_d;
    e->target[5] = y_d;
    e->ye_int = 0.0;   /* usv_asmc_ye_int_env.py:268, 273 */
    e->ye_last = 0.0;
    e->f32_state = 0;  /* float64 arrays until the first step() */
    e->steps = 0;
    e->ep_return = 0.0;
    if (obs)
        for (int k = 0; k < 6; ++k) obs[k] = (float)e->state[k];
}

static inline double l_u53(uint32_t a, uint32_t b) {
    return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) * (1.0 / 9007199254740992.0);
}

/* reset() with the global np.random (MT19937) replaced by Philox4x32-10 keyed like the other env:
 * draw k = half (k & 1) of block (k >> 1) of (seed, global env id, episode) */
void usv_oracle_legacy_reset(UsvOracleLegacy* e, uint64_t seed, uint64_t global_id, float obs[6]) {
    /* usv_asmc_env.py:260-281; usv_asmc_ye_int_env.py:258-279 (x, y in +-5, speed 0.4..1.4);
     * usv_pid_env.py:238-257 (speed 0.4..1.4) */
    const double xy = (e->variant == USV_ORACLE_LEGACY_YE_INT) ? 5.0 : 2.5;
    const double s0 = (e->variant == USV_ORACLE_LEGACY_ASMC_V0) ? 1.4 : 0.4;
    const double lo[7] = {-xy, -xy, -M_PI, -2.5, -2.5, 15, s0};
    const double hi[7] = {xy, xy, M_PI, 2.5, 2.5, 30, s0 + 1.0};
    double d[7];
    e->episode += 1;
    for (uint32_t b = 0; b < 4; ++b) {
        uint32_t o[4];
        usv_oracle_philox(seed, global_id, (uint32_t)e->episode, b, o);
        d[2 * b] = lo[2 * b] + (hi[2 * b] - lo[2 * b]) * l_u53(o[0], o[1]);
        if (2 * b + 1 < 7) d[2 * b + 1] = lo[2 * b + 1] + (hi[2 * b + 1] - lo[2 * b + 1]) * l_u53(o[2], o[3]);
    }
    usv_oracle_legacy_reset_from_draws(e, d, obs);
}

void usv_oracle_legacy_init(UsvOracleLegacy* e) {
    memset(e, 0, sizeof(*e));
    e->episode = -1;
}

void usv_oracle_legacy_vec_reset(UsvOracleLegacy* envs, int64_t n, uint64_t seed, int64_t global_offset, float* obs,
                                 int n_threads) {
    (void)n_threads;
#pragma omp parallel for schedule(static) num_threads(n_threads > 0 ? n_threads : 1)
    for (int64_t i = 0; i < n; ++i) usv_oracle_legacy_reset(&envs[i], seed, (uint64_t)(global_offset + i), obs + 6 * i);
}

/* N envs with SB3 VecEnv auto-reset semantics: done envs are reset in place, obs then holds the reset observation
 * and terminal_obs (may be NULL) the last observation of the finished episode */
void usv_oracle_legacy_vec_step(UsvOracleLegacy* envs, int64_t n, const float* actions, float* obs, double* reward,
                                uint8_t* done, float* terminal_obs, int auto_reset, uint64_t seed,
                                int64_t global_offset, int n_threads) {
    (void)n_threads;
#pragma omp parallel for schedule(static) num_threads(n_threads > 0 ? n_threads : 1)
    for (int64_t i = 0; i < n; ++i) {
        usv_oracle_legacy_step(&envs[i], actions[i], obs + 6 * i, &reward[i], &done[i]);
        if (done[i] && auto_reset) {
            if (terminal_obs) memcpy(terminal_obs + 6 * i, obs + 6 * i, 6 * sizeof(float));
            usv_oracle_legacy_reset(&envs[i], seed, (uint64_t)(global_offset + i), obs + 6 * i);
        }
    }
}
