// usv_legacy.cuh -- the legacy path-following env 'usv-asmc-v0' (gym_usv/envs/usv_asmc_env.py: UsvAsmcEnv):
// one 0.01 s ASMC + 3-DOF sub-step per step(), no obstacles, observation float32[6].
//
// The reference mixes float32 and float64: reset() stores float64 arrays (usv_asmc_env.py:289-294),
// step() re-stores every array as float32 (:247-251) and, under numpy >= 2, Python-float x np.float32 stays
// float32.  legacy_step<S> follows that dtype flow operation by operation: S = double for the first step of an
// episode, S = float afterwards; expressions that touch `target` (float64), an np.float64 constant or an
// np.where(..., python float) result are float64 in both; the arrays the reference builds with dtype=np.float32
// (T, CRB, CA, Dl, Dn, J) are rounded to float32.  Every add / mul / div goes through the round-to-nearest
// intrinsics so that nothing is contracted into an FMA (numpy does not contract).
//
// Storage follows the same observation: after a step every state value IS a float32, so the state lives in HBM as
// float32 [19][N] (half the traffic of a float64 layout); the only float64 state a fresh reset holds is the pose,
// kept in pos0 [3][N] and read by the first step of the episode only.
//
// The two registered siblings share this step() (gym_usv/__init__.py:8-16): 'usv-asmc-ye-int-v0'
// (usv_asmc_ye_int_env.py: integral of the cross-track error in the observation, other reward / termination) and
// 'usv-pid-v0' (usv_pid_env.py: PID law instead of the adaptive sliding mode).  Neither has a dtype=np.float32
// anywhere, so they run legacy_step<double, V> with no rounding and keep their state as float64 [21][N] (st64).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include <type_traits>

#include "usv_philox.cuh"
#include "usv_state.cuh"

namespace usv {

constexpr int LEG_ROWS = 19;  // action_last | u v r | x y psi | e_u_int Ka_uKa_psi | last[9]
constexpr int LEG64_ROWS = 21;  // the 19 rows above | ye_int | ye_last, float64 (ye-int / pid envs)
constexpr int LEG_OBS = USVGPU_LEGACY_OBS_DIM;
enum LegVariant { LEG_ASMC_V0 = 0, LEG_YE_INT = 1, LEG_PID = 2 };
enum LegRow {
    G_ALAST = 0, G_U, G_V, G_R, G_X, G_Y, G_PSI, G_EUI, G_KAU, G_KAP,
    G_XDL, G_YDL, G_PDL, G_UDL, G_VDL, G_RDL, G_EUL, G_KADUL, G_KADPL
};

struct LegacyPtrs {
    float* st;          // [19][n]  float32 state (valid when steps > 0; zeros + pos0 when steps == 0)
    double* pos0;       // [3][n]   float64 pose of a freshly reset env (usv_asmc_env.py:293)
    double* st64;       // [21][n]  float64 state of the ye-int / pid envs (st and pos0 are null for them)
    double* target;     // [6][n]   x_0 y_0 desired_speed ak x_d y_d (usv_asmc_env.py:296), float64
    int* steps;         // [n]
    int* episode;       // [n]
    double* ep_return;  // [n]
    double2* ep_done;   // [n]
    DeviceStats* stats;
    long long n;
    long long global_offset;
    unsigned long long seed;
    int variant;        // LegVariant
};

// ---- round-to-nearest arithmetic without contraction, float and double --------------------------------------
__device__ __forceinline__ float r_add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float r_sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float r_mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float r_div(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float r_sqrt(float a) { return __fsqrt_rn(a); }
__device__ __forceinline__ float r_abs(float a) { return fabsf(a); }
__device__ __forceinline__ double r_add(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double r_sub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double r_mul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double r_div(double a, double b) { return __ddiv_rn(a, b); }
__device__ __forceinline__ double r_sqrt(double a) { return __dsqrt_rn(a); }
__device__ __forceinline__ double r_abs(double a) { return fabs(a); }
// np.cos / np.sin of an S heading: for float32, the float64 routine rounded once (what a correctly rounded cosf gives)
__device__ __forceinline__ void r_sincos(float a, float& s, float& c) {
    double sd, cd;
    sincos((double)a, &sd, &cd);
    s = (float)sd;
    c = (float)cd;
}
__device__ __forceinline__ void r_sincos(double a, double& s, double& c) { sincos(a, &s, &c); }

__device__ __forceinline__ double l_sign(double x) { return (double)((x > 0.0) - (x < 0.0)); }
// usv_asmc_env.py:124,148,229,232  np.where(|a| > pi, sign(a) * (|a| - 2 pi), a)
__device__ __forceinline__ double l_wrap(double a) {
    return (fabs(a) > PI) ? r_mul(l_sign(a), r_sub(fabs(a), 2 * PI)) : a;
}

struct LegacyState {
    double alast, u, v, r, x, y, psi, eui, kau, kap, xdl, ydl, pdl, udl, vdl, rdl, eul, kadul, kadpl;
    double yeint, yelast;  // usv_asmc_ye_int_env.py: aux_vars[3], last[9]
};
struct LegacyOut {
    float obs[LEG_OBS];
    double reward;
    bool done;
};

// usv_asmc_env.py:19-51, 66-78
namespace leg {
constexpr double H = 0.01, MIN_SPEED = 0.3, XUD = -2.25, YVD = -23.13, YRD = -1.31, NVD = -16.41, NRD = -2.79;
constexpr double YVV = -99.99, YVR = -5.49, YRV = -5.49, YRR = -8.8, NVV = -5.49, NVR = -8.8, NRV = -8.8, NRR = -3.49;
constexpr double M = 30.0, IZ = 4.1, B = 0.41, C = 0.78;
constexpr double K_U = 0.1, K_PSI = 0.2, KMIN_U = 0.05, KMIN_PSI = 0.2, K2_U = 0.02, K2_PSI = 0.1, MU_U = 0.05,
                 MU_PSI = 0.1, LAMBDA_U = 0.001, LAMBDA_PSI = 1.0, K_AK = 5.72, K_YE = 0.5, SIGMA_YE = 1.0;
constexpr double MAX_YE = 10.0, MAX_X = 30.0, W_ACTION = 0.2;
constexpr double MIN_X = -10.0;                                    // usv_asmc_ye_int_env.py:63, usv_pid_env.py:60
constexpr double K_I = 0.001;                                      // usv_asmc_ye_int_env.py: k_i
constexpr double KP_U = 1.1, KI_U = 0.2, KD_U = 0.1, KP_PSI = 0.8, KD_PSI = 3.0;  // usv_pid_env.py:37-41
// :132-133  1.1+0.0045*(1.01/0.09) - 0.1*(0.27/0.09)+0.016*(np.power((0.27/0.09), 2))  (an np.float64)
constexpr double YV_K = 1.1 + 0.0045 * (1.01 / 0.09) - 0.1 * (0.27 / 0.09) + 0.016 * ((0.27 / 0.09) * (0.27 / 0.09));
// :77  1. / np.power((max_action/2 - min_action/2) / integral_step, 2)
constexpr double C_ACT_D = ((PI / 2) / 2 - (-PI / 2) / 2) / H;
constexpr double C_ACTION = 1. / (C_ACT_D * C_ACT_D);
// :187-189, 214  np.linalg.inv(M), block-diagonal closed form
constexpr double M00 = M - XUD, M11 = M - YVD, M12 = 0 - YRD, M21 = 0 - NVD, M22 = IZ - NRD;
constexpr double DET = M11 * M22 - M12 * M21;
constexpr double MI00 = 1.0 / M00, MI11 = M22 / DET, MI12 = -M12 / DET, MI21 = -M21 / DET, MI22 = M11 / DET;
// UsvAsmcEnv.step (usv_asmc_env.py:99-255).  S: the dtype numpy gives to expressions built from the unpacked state
// scalars and Python literals (double on the first step of an episode, float afterwards).  V: LegVariant; the
// siblings (V != 0) are S = double with arrays and stores left in float64.
template <typename S, int V = LEG_ASMC_V0>
__device__ __forceinline__ void legacy_step(LegacyState& e, const double x_0, const double y_0,
                                            const double desired_speed, const double ak, const float action,
                                            LegacyOut& out) {
    static_assert(V == LEG_ASMC_V0 || std::is_same<S, double>::value, "the siblings are float64 throughout");
    using A = typename std::conditional<V == LEG_ASMC_V0, float, double>::type;  // dtype of T, CRB, CA, Dl, Dn, J
    // :108-115 unpack
    const S action_last = (S)e.alast, u = (S)e.u, v = (S)e.v, r = (S)e.r, x = (S)e.x, y = (S)e.y, psi = (S)e.psi;
    const S e_u_int0 = (S)e.eui, Ka_u0 = (S)e.kau, Ka_psi0 = (S)e.kap;
    const S x_dot_last = (S)e.xdl, y_dot_last = (S)e.ydl, psi_dot_last = (S)e.pdl;
    const S u_dot_last = (S)e.udl, v_dot_last = (S)e.vdl, r_dot_last = (S)e.rdl;
    const S e_u_last = (S)e.eul, Ka_dot_u_last = (S)e.kadul, Ka_dot_psi_last = (S)e.kadpl;

    // :120-124
    const S action_dot = r_div(r_sub((S)action, action_last), (S)H);
    const double psi_d = l_wrap(r_add((double)action, ak));

    // :126-130
    double Xu = -25, Xuu = 0;
    if (r_abs(u) > (S)1.2) {
        Xu = 64.55;
        Xuu = -70.92;
    }
    // :132-139  Yv picks up the np.float64 constant; Yr, Nv, Nr stay in S
    const double Yv = r_mul((double)r_mul((S)0.5, r_mul((S)(-40 * 1000), r_abs(v))), YV_K);
    const S speed = r_sqrt(r_add(r_mul(u, u), r_mul(v, v)));
    const S Yr = r_mul(r_mul(r_mul(r_mul((S)(6 * (-3.141592 * 1000)), speed), (S)0.09), (S)0.09), (S)1.01);
    const S Nv = r_mul(r_mul(r_mul(r_mul((S)(0.06 * (-3.141592 * 1000)), speed), (S)0.09), (S)0.09), (S)1.01);
    const S Nr = r_mul(r_mul(r_mul(r_mul(r_mul((S)(0.02 * (-3.141592 * 1000)), speed), (S)0.09), (S)0.09), (S)1.01), (S)1.01);

    // :141-145
    constexpr double g_u = 1 / (M - XUD), g_psi = 1 / (IZ - NRD);
    const S f_u = r_div(r_add(r_mul(r_mul((S)(M - YVD), v), r), r_add(r_mul((S)Xuu, r_abs(u)), r_mul((S)Xu, u))),
                        (S)(M - XUD));
    const S f_psi = r_div(r_add(r_mul(r_mul((S)(-XUD + YVD), u), v), r_mul(Nr, r)), (S)(IZ - NRD));

    // :147-156
    const double e_psi = l_wrap(r_sub(psi_d, (double)psi));
    const S e_psi_dot = r_sub((S)0, r);
    const double abs_e_psi = fabs(e_psi);
    const double u_psi = r_div(1.0, r_add(1.0, exp(r_mul(10.0, r_sub(r_mul(abs_e_psi, 2 / PI), 0.5)))));
    const double u_d = r_add(r_mul(desired_speed - MIN_SPEED, u_psi), MIN_SPEED);

    // :158-162  e_u_last is read but never written by the reference: it stays 0
    const double e_u = r_sub(u_d, (double)u);
    const double e_u_int = r_add(r_div(r_mul(H, r_add(e_u, (double)e_u_last)), 2.0), (double)e_u_int0);
    double Ka_dot_u, Ka_dot_psi, Ka_u, Ka_psi, Tx, Tz;
    if constexpr (V == LEG_PID) {
        // usv_pid_env.py:149-155  PID law; the adaptive gains are carried along untouched (they stay 0)
        const double e_u_dot = r_div(r_sub(e_u, (double)e_u_last), H);
        const double ua_u = r_add(r_add(r_mul(KP_U, e_u), r_mul(KI_U, e_u_int)), r_mul(KD_U, e_u_dot));
        const double ua_psi = r_add(r_mul(KP_PSI, e_psi), r_mul(KD_PSI, (double)e_psi_dot));
        Ka_u = (double)Ka_u0, Ka_psi = (double)Ka_psi0;
        Ka_dot_u = (double)Ka_dot_u_last, Ka_dot_psi = (double)Ka_dot_psi_last;
        Tx = r_div(r_add(-(double)f_u, ua_u), g_u);
        Tz = r_div(r_add(-(double)f_psi, ua_psi), g_psi);
    } else {
        const double sigma_u = r_add(e_u, r_mul(LAMBDA_U, e_u_int));
        const double sigma_psi = r_add((double)e_psi_dot, r_mul(LAMBDA_PSI, e_psi));

        // :164-171
        Ka_dot_u = (Ka_u0 > (S)KMIN_U) ? r_mul(K_U, l_sign(r_sub(fabs(sigma_u), MU_U))) : KMIN_U;
        Ka_dot_psi = (Ka_psi0 > (S)KMIN_PSI) ? r_mul(K_PSI, l_sign(r_sub(fabs(sigma_psi), MU_PSI))) : KMIN_PSI;
        Ka_u = r_add(r_div(r_mul(H, r_add(Ka_dot_u, (double)Ka_dot_u_last)), 2.0), (double)Ka_u0);
        Ka_psi = r_add(r_div(r_mul(H, r_add(Ka_dot_psi, (double)Ka_dot_psi_last)), 2.0), (double)Ka_psi0);

        // :173-174  np.power(., 0.5) == sqrt
        const double ua_u = r_sub(r_mul(r_mul(-Ka_u, r_sqrt(fabs(sigma_u))), l_sign(sigma_u)), r_mul(K2_U, sigma_u));
        const double ua_psi = r_sub(r_mul(r_mul(-Ka_psi, r_sqrt(fabs(sigma_psi))), l_sign(sigma_psi)), r_mul(K2_PSI, sigma_psi));

        // :176-185
        Tx = r_div(r_sub(r_sub(r_mul(LAMBDA_U, e_u), (double)f_u), ua_u), g_u);
        Tz = r_div(r_sub(r_sub(r_mul(LAMBDA_PSI, e_psi), (double)f_psi), ua_psi), g_psi);
    }
    double Tport = r_add(r_div(Tx, 2.0), r_div(Tz, B));
    double Tstbd = r_sub(r_div(Tx, 2 * C), r_div(Tz, B * C));
    if (Tport > 36.5) Tport = 36.5;
    if (Tport < -30) Tport = -30;
    if (Tstbd > 36.5) Tstbd = 36.5;
    if (Tstbd < -30) Tstbd = -30;

    // :191  dtype=np.float32
    const A T0 = (A)r_add(Tport, r_mul(C, Tstbd));
    const A T2 = (A)r_mul(0.5 * B, r_sub(Tport, r_mul(C, Tstbd)));

    // :193-201  entries evaluated in S, stored float32; C = CRB + CA in float32
    const A CRB02 = (A)r_sub((S)0, r_mul((S)M, v)), CRB12 = (A)r_mul((S)M, u);
    const A CRB20 = (A)r_mul((S)M, v), CRB21 = (A)r_sub((S)0, r_mul((S)M, u));
    const S half_yn = (S)((YRD + NVD) / 2);
    const A CA02 = (A)r_mul((S)2, r_add(r_mul((S)YVD, v), r_mul(half_yn, r)));
    const A CA12 = (A)r_sub((S)0, r_mul((S)(XUD * M), u));
    const A CA20 = (A)r_mul((S)2, r_sub(r_mul((S)(0 - YVD), v), r_mul(half_yn, r)));
    const A CA21 = (A)r_mul((S)(XUD * M), u);
    const A C02 = r_add(CRB02, CA02), C12 = r_add(CRB12, CA12), C20 = r_add(CRB20, CA20), C21 = r_add(CRB21, CA21);

    // :203-212
    const A Dl00 = (A)(0 - Xu), Dl11 = (A)(0 - Yv), Dl12 = (A)r_sub((S)0, Yr);
    const A Dl21 = (A)r_sub((S)0, Nv), Dl22 = (A)r_sub((S)0, Nr);
    const S au = r_abs(u), av = r_abs(v), ar = r_abs(r);
    const A Dn00 = (A)r_mul((S)Xuu, au);
    const A Dn11 = (A)r_add(r_mul((S)YVV, av), r_mul((S)YVR, ar));
    const A Dn12 = (A)r_add(r_mul((S)YRV, av), r_mul((S)YRR, ar));
    const A Dn21 = (A)r_add(r_mul((S)NVV, av), r_mul((S)NVR, ar));
    const A Dn22 = (A)r_add(r_mul((S)NRV, av), r_mul((S)NRR, ar));
    const A D00 = r_sub(Dl00, Dn00), D11 = r_sub(Dl11, Dn11), D12 = r_sub(Dl12, Dn12);
    const A D21 = r_sub(Dl21, Dn21), D22 = r_sub(Dl22, Dn22);

    // :214-218  matmul(float32 3x3, S vector) and T - . - . are S; inv(M) @ . is float64
    const S Cv0 = r_mul((S)C02, r), Cv1 = r_mul((S)C12, r), Cv2 = r_add(r_mul((S)C20, u), r_mul((S)C21, v));
    const S Dv0 = r_mul((S)D00, u), Dv1 = r_add(r_mul((S)D11, v), r_mul((S)D12, r));
    const S Dv2 = r_add(r_mul((S)D21, v), r_mul((S)D22, r));
    const S rhs0 = r_sub(r_sub((S)T0, Cv0), Dv0), rhs1 = r_sub(r_sub((S)0, Cv1), Dv1), rhs2 = r_sub(r_sub((S)T2, Cv2), Dv2);
    const double ud0 = r_mul(MI00, (double)rhs0);
    const double ud1 = r_add(r_mul(MI11, (double)rhs1), r_mul(MI12, (double)rhs2));
    const double ud2 = r_add(r_mul(MI21, (double)rhs1), r_mul(MI22, (double)rhs2));
    const double un = r_add(r_div(r_mul(H, r_add(ud0, (double)u_dot_last)), 2.0), (double)u);
    const double vn = r_add(r_div(r_mul(H, r_add(ud1, (double)v_dot_last)), 2.0), (double)v);
    const double rn = r_add(r_div(r_mul(H, r_add(ud2, (double)r_dot_last)), 2.0), (double)r);

    // :220-229  J is float32: np.cos / np.sin of the S heading, rounded
    S sn_s, cs_s;
    r_sincos(psi, sn_s, cs_s);
    const A Jc = (A)cs_s, Js = (A)sn_s;
    const double xd = r_add(r_mul((double)Jc, un), r_mul((double)(-Js), vn));
    const double yd = r_add(r_mul((double)Js, un), r_mul((double)Jc, vn));
    const double pd = rn;
    const double xn = r_add(r_div(r_mul(H, r_add(xd, (double)x_dot_last)), 2.0), (double)x);
    const double yn = r_add(r_div(r_mul(H, r_add(yd, (double)y_dot_last)), 2.0), (double)y);
    const double psin = l_wrap(r_add(r_div(r_mul(H, r_add(pd, (double)psi_dot_last)), 2.0), (double)psi));

    // :231-235
    const double psi_ak = l_wrap(r_sub(psin, ak));
    double sak, cak;
    sincos(ak, &sak, &cak);
    const double ye = r_add(r_mul(-r_sub(xn, x_0), sak), r_mul(r_sub(yn, y_0), cak));
    const double ye_abs = fabs(ye);

    // :237, 364-374
    const double ad2 = (double)r_mul(action_dot, action_dot);
    const double a_psi_ak = fabs(psi_ak);
    const double reward_action = r_mul(W_ACTION, tanh(r_mul(-C_ACTION, ad2)));
    const double reward_ak = -exp(r_mul(K_AK, r_sub(a_psi_ak, PI)));
    double reward, ye_obs = ye;
    if constexpr (V == LEG_YE_INT) {
        // usv_asmc_ye_int_env.py:230-236, 350-360
        double ye_int = e.yeint;
        if (l_sign(ye) != l_sign(e.yelast)) ye_int = 0;
        ye_int = r_add(r_mul(H, r_add(ye, e.yelast)), ye_int);
        ye_obs = r_add(ye, r_mul(K_I, ye_int));
        e.yeint = ye_int;
        e.yelast = ye;
        const double reward_ye = exp(r_mul(-K_YE, ye_abs));
        reward = r_add(reward_action, (a_psi_ak < PI / 2) ? reward_ye : reward_ak);
    } else {
        const double reward_ye = (ye_abs > SIGMA_YE) ? exp(r_mul(-K_YE, ye_abs))
                                                     : exp(r_div(r_mul(-K_YE, r_mul(ye_abs, ye_abs)), SIGMA_YE));
        reward = (a_psi_ak < PI / 2) ? r_add(reward_action, reward_ye) : reward_ak;
    }

    // :239, 376-401
    double spa, cpa;
    sincos(psi_ak, &spa, &cpa);
    const double v_ak = r_add(r_mul(spa, un), r_mul(cpa, vn));

    // :241-245
    bool done = false;
    // siblings: usv_asmc_ye_int_env.py:241, usv_pid_env.py:219
    if (ye_abs > MAX_YE || (V == LEG_ASMC_V0 ? fabs(xn) > MAX_X : xn < MIN_X)) {
        done = true;
        reward = -1;
    }

    // :247-251  everything re-stored as float32 (the siblings keep float64)
    e.alast = (double)action;
    e.u = (double)(A)un;
    e.v = (double)(A)vn;
    e.r = (double)(A)rn;
    e.x = (double)(A)xn;
    e.y = (double)(A)yn;
    e.psi = (double)(A)psin;
    e.eui = (double)(A)e_u_int;
    e.kau = (double)(A)Ka_u;
    e.kap = (double)(A)Ka_psi;
    e.xdl = (double)(A)xd;
    e.ydl = (double)(A)yd;
    e.pdl = (double)(A)pd;
    e.udl = (double)(A)ud0;
    e.vdl = (double)(A)ud1;
    e.rdl = (double)(A)ud2;
    e.eul = (double)(A)(double)e_u_last;
    e.kadul = (double)(A)Ka_dot_u;
    e.kadpl = (double)(A)Ka_dot_psi;
    out.obs[0] = (float)un;
    out.obs[1] = (float)v_ak;
    out.obs[2] = (float)rn;
    out.obs[3] = (float)ye_obs;
    out.obs[4] = (float)psi_ak;
    out.obs[5] = action;
    out.reward = reward;
    out.done = done;
}
}  // namespace leg

// UsvAsmcEnv.reset (usv_asmc_env.py:258-300).  The reference draws x, y, psi, x_0, y_0, x_d, desired_speed from the
// global np.random (MT19937); here draw k is half (k & 1) of Philox block (k >> 1) of (seed, GLOBAL env id, episode).
// Writes the float64 pose to pos0, zeroes the float32 state, returns the reset observation.
__device__ __forceinline__ void legacy_reset_env(const LegacyPtrs& s, long long i, float obs[LEG_OBS]) {
    const long long n = s.n;
    const int ep = s.episode[i] + 1;
    s.episode[i] = ep;
    const DrawKey dk(s.seed, (unsigned long long)(s.global_offset + i), (uint32_t)ep);
    double d[8];
#pragma unroll
    for (int b = 0; b < 4; ++b) dk.pair(b, d[2 * b], d[2 * b + 1]);
    // usv_asmc_ye_int_env.py:258-259 draws x, y from +-5; both siblings draw the speed from 0.4..1.4 (:279, usv_pid_env.py:257)
    const double xy = (s.variant == LEG_YE_INT) ? 5.0 : 2.5, sp0 = (s.variant == LEG_ASMC_V0) ? 1.4 : 0.4;
    const double x = -xy + (xy - -xy) * d[0], y = -xy + (xy - -xy) * d[1], psi = -PI + (PI - -PI) * d[2];
    const double x_0 = -2.5 + (2.5 - -2.5) * d[3], y_0 = -2.5 + (2.5 - -2.5) * d[4];
    const double x_d = 15 + (30 - 15) * d[5], desired_speed = sp0 + ((sp0 + 1.0) - sp0) * d[6];
    const double y_d = y_0;                                                        // :280
    const double ak = (double)(float)atan2(r_sub(y_d, y_0), r_sub(x_d, x_0));      // :283-284 np.float32(ak)
    const double psi_ak = (double)(float)l_wrap(r_sub(psi, ak));                   // :286-288
    double sak, cak;
    sincos(ak, &sak, &cak);
    const double ye = r_add(r_mul(-r_sub(x, x_0), sak), r_mul(r_sub(y, y_0), cak));  // :289
    if (s.variant == LEG_ASMC_V0) {
#pragma unroll
        for (int k = 0; k < LEG_ROWS; ++k) s.st[(long long)k * n + i] = 0.0f;
        s.pos0[i] = x;
        s.pos0[n + i] = y;
        s.pos0[2 * n + i] = psi;
    } else {
#pragma unroll
        for (int k = 0; k < LEG64_ROWS; ++k)
            s.st64[(long long)k * n + i] = (k == G_X) ? x : (k == G_Y) ? y : (k == G_PSI) ? psi : 0.0;
    }
    double* tg = s.target + i;
    tg[0] = x_0;
    tg[n] = y_0;
    tg[2 * n] = desired_speed;
    tg[3 * n] = ak;
    tg[4 * n] = x_d;
    tg[5 * n] = y_d;
    s.steps[i] = 0;
    s.ep_return[i] = 0.0;
    obs[0] = 0.0f;
    obs[1] = 0.0f;  // body_to_path(0, 0, psi_ak)
    obs[2] = 0.0f;
    obs[3] = (float)ye;
    obs[4] = (float)psi_ak;
    obs[5] = 0.0f;
}

__global__ void legacy_reset_kernel(LegacyPtrs s, const uint8_t* __restrict__ mask, float* __restrict__ obs) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= s.n) return;
    if (mask && !mask[i]) return;
    float o[LEG_OBS];
    legacy_reset_env(s, i, o);
    if (obs) {
        float2* row = reinterpret_cast<float2*>(obs + i * LEG_OBS);
        row[0] = make_float2(o[0], o[1]);
        row[1] = make_float2(o[2], o[3]);
        row[2] = make_float2(o[4], o[5]);
    }
}

// One VecEnv.step() of N UsvAsmcEnv: thread per env, state columns coalesced (float32 SoA), the float64 pose of a
// freshly reset env is fetched only by the lanes whose episode is at step 0.  Bound: HBM.
constexpr int LEG_B = 128;  // occupancy sweep in profiles/r01w_legacy_launch_bounds.txt: 3 CTAs per SM (138 registers) 104 us per 1 M envs, 5 CTAs 80 us, 8 CTAs 67.5 us, 10 CTAs 72 us
#ifndef LEG_MIN_CTAS
#define LEG_MIN_CTAS 8  // 64 registers (96 B spill), 32 warps per SM: 67.5 us per 1 M envs; 5 CTAs / 96 registers ran 80 us
#endif
__global__ void __launch_bounds__(LEG_B, LEG_MIN_CTAS)
legacy_step_kernel(LegacyPtrs s, StepFlags fl, const float* __restrict__ actions, float* __restrict__ obs,
                   float* __restrict__ reward, uint8_t* __restrict__ terminated, uint8_t* __restrict__ truncated,
                   float* __restrict__ terminal_obs) {
    const long long n = s.n;
    const long long i = fl.lo + (long long)blockIdx.x * LEG_B + threadIdx.x;
    if (i >= fl.hi) return;
    pdl_wait();  // usv_state.cuh
    float* st = s.st + i;
    float f[LEG_ROWS];
#pragma unroll
    for (int k = 0; k < LEG_ROWS; ++k) f[k] = st[(long long)k * n];
    int steps = s.steps[i];
    const double* tg = s.target + i;
    const double x_0 = tg[0], y_0 = tg[n], desired_speed = tg[2 * n], ak = tg[3 * n];
    const float action = actions[i];
    LegacyState e = {f[0], f[1], f[2], f[3], f[4], f[5], f[6], f[7], f[8], f[9], f[10], f[11], f[12], f[13], f[14],
                     f[15], f[16], f[17], f[18]};
    LegacyOut out;
    if (steps == 0) {  // float64 arrays straight out of reset() (usv_asmc_env.py:289-294)
        e.x = s.pos0[i];
        e.y = s.pos0[n + i];
        e.psi = s.pos0[2 * n + i];
        leg::legacy_step<double>(e, x_0, y_0, desired_speed, ak, action, out);
    } else {
        leg::legacy_step<float>(e, x_0, y_0, desired_speed, ak, action, out);
    }
    steps += 1;
    const bool tlimit = (fl.max_episode_steps > 0) && (steps >= fl.max_episode_steps);
    const bool done = out.done || tlimit;
    reward[i] = (float)out.reward;
    terminated[i] = out.done ? 1 : 0;
    truncated[i] = (tlimit && !out.done) ? 1 : 0;
    if (!(isfinite(e.x) && isfinite(e.y) && isfinite(e.psi))) atomicAdd(&s.stats->nonfinite, 1ull);

    float* o = out.obs;
    if (done) {
        const double ep_ret = s.ep_return[i] + out.reward;
        s.ep_done[i] = make_double2(ep_ret, (double)steps);
        DeviceStats* ds = s.stats;
        atomicAdd(&ds->return_sum, ep_ret);
        atomicAdd(&ds->length_sum, (double)steps);
        atomicAdd(&ds->episodes, 1ull);
        if (out.done) atomicAdd(&ds->terminated, 1ull);
        else atomicAdd(&ds->time_limit, 1ull);
        if (fl.auto_reset) {
            if (terminal_obs) {
                float2* trow = reinterpret_cast<float2*>(terminal_obs + i * LEG_OBS);
                trow[0] = make_float2(o[0], o[1]);
                trow[1] = make_float2(o[2], o[3]);
                trow[2] = make_float2(o[4], o[5]);
            }
            legacy_reset_env(s, i, o);  // zeroes the state columns, steps and ep_return
        } else {
            s.ep_return[i] = ep_ret;
        }
    } else {
        s.ep_return[i] += out.reward;
    }
    if (!(done && fl.auto_reset)) {
        const float g[LEG_ROWS] = {(float)e.alast, (float)e.u, (float)e.v, (float)e.r, (float)e.x, (float)e.y,
                                   (float)e.psi, (float)e.eui, (float)e.kau, (float)e.kap, (float)e.xdl, (float)e.ydl,
                                   (float)e.pdl, (float)e.udl, (float)e.vdl, (float)e.rdl, (float)e.eul,
                                   (float)e.kadul, (float)e.kadpl};
#pragma unroll
        for (int k = 0; k < LEG_ROWS; ++k) st[(long long)k * n] = g[k];
        s.steps[i] = steps;
    }
    float2* row = reinterpret_cast<float2*>(obs + i * LEG_OBS);
    row[0] = make_float2(o[0], o[1]);
    row[1] = make_float2(o[2], o[3]);
    row[2] = make_float2(o[4], o[5]);
}

// One VecEnv.step() of N UsvAsmcYeIntEnv (V = LEG_YE_INT) or UsvPidEnv (V = LEG_PID): the same thread-per-env
// shape over the float64 state columns st64 [21][N] (the pid env never touches rows 19, 20).
#ifndef LEG64_MIN_CTAS
#define LEG64_MIN_CTAS 5  // 96 registers, no spills: 79 / 75 us per 1 M envs; 3 CTAs (146 / 156 registers) ran 97 / 95 us
#endif
template <int V>
__global__ void __launch_bounds__(LEG_B, LEG64_MIN_CTAS)
legacy64_step_kernel(LegacyPtrs s, StepFlags fl, const float* __restrict__ actions, float* __restrict__ obs,
                     float* __restrict__ reward, uint8_t* __restrict__ terminated, uint8_t* __restrict__ truncated,
                     float* __restrict__ terminal_obs) {
    constexpr int ROWS = (V == LEG_YE_INT) ? LEG64_ROWS : LEG_ROWS;
    const long long n = s.n;
    const long long i = fl.lo + (long long)blockIdx.x * LEG_B + threadIdx.x;
    if (i >= fl.hi) return;
    pdl_wait();  // usv_state.cuh
    double* st = s.st64 + i;
    double f[LEG64_ROWS];
#pragma unroll
    for (int k = 0; k < LEG64_ROWS; ++k) f[k] = (k < ROWS) ? st[(long long)k * n] : 0.0;
    int steps = s.steps[i];
    const double* tg = s.target + i;
    const double x_0 = tg[0], y_0 = tg[n], desired_speed = tg[2 * n], ak = tg[3 * n];
    const float action = actions[i];
    LegacyState e = {f[0], f[1], f[2], f[3], f[4], f[5], f[6], f[7], f[8], f[9], f[10], f[11], f[12], f[13], f[14],
                     f[15], f[16], f[17], f[18], f[19], f[20]};
    LegacyOut out;
    leg::legacy_step<double, V>(e, x_0, y_0, desired_speed, ak, action, out);
    steps += 1;
    const bool tlimit = (fl.max_episode_steps > 0) && (steps >= fl.max_episode_steps);
    const bool done = out.done || tlimit;
    reward[i] = (float)out.reward;
    terminated[i] = out.done ? 1 : 0;
    truncated[i] = (tlimit && !out.done) ? 1 : 0;
    if (!(isfinite(e.x) && isfinite(e.y) && isfinite(e.psi))) atomicAdd(&s.stats->nonfinite, 1ull);

    float* o = out.obs;
    if (done) {
        const double ep_ret = s.ep_return[i] + out.reward;
        s.ep_done[i] = make_double2(ep_ret, (double)steps);
        DeviceStats* ds = s.stats;
        atomicAdd(&ds->return_sum, ep_ret);
        atomicAdd(&ds->length_sum, (double)steps);
        atomicAdd(&ds->episodes, 1ull);
        if (out.done) atomicAdd(&ds->terminated, 1ull);
        else atomicAdd(&ds->time_limit, 1ull);
        if (fl.auto_reset) {
            if (terminal_obs) {
                float2* trow = reinterpret_cast<float2*>(terminal_obs + i * LEG_OBS);
                trow[0] = make_float2(o[0], o[1]);
                trow[1] = make_float2(o[2], o[3]);
                trow[2] = make_float2(o[4], o[5]);
            }
            legacy_reset_env(s, i, o);  // rewrites the state columns, steps and ep_return
        } else {
            s.ep_return[i] = ep_ret;
        }
    } else {
        s.ep_return[i] += out.reward;
    }
    if (!(done && fl.auto_reset)) {
        const double g[LEG64_ROWS] = {e.alast, e.u, e.v, e.r, e.x, e.y, e.psi, e.eui, e.kau, e.kap, e.xdl,
                                      e.ydl, e.pdl, e.udl, e.vdl, e.rdl, e.eul, e.kadul, e.kadpl, e.yeint, e.yelast};
#pragma unroll
        for (int k = 0; k < ROWS; ++k) st[(long long)k * n] = g[k];
        s.steps[i] = steps;
    }
    float2* row = reinterpret_cast<float2*>(obs + i * LEG_OBS);
    row[0] = make_float2(o[0], o[1]);
    row[1] = make_float2(o[2], o[3]);
    row[2] = make_float2(o[4], o[5]);
}

}  // namespace usv
