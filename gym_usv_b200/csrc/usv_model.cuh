// usv_model.cuh -- vessel/controller constants and one ASMC + 3-DOF sub-step (device code).
//
// Restates gym_usv/control/usv_asmc.py:53-244 (reference root-relative) for one environment
// held entirely in registers.  The 3x3 matrices of the reference are never materialised: the
// products C(nu)nu, D(nu)nu and M^-1 rhs are expanded in closed form (verified to
// reproduce the reference to all printed digits in float64).
#pragma once
#include <cuda_runtime.h>
#include <math.h>

#include "usv_math.cuh"

namespace usv {

// ---- usv_asmc.py:7-24 vessel coefficients -------------------------------------------------
constexpr double X_U_DOT = -2.25, Y_V_DOT = -23.13, Y_R_DOT = -1.31, N_V_DOT = -16.41, N_R_DOT = -2.79;
constexpr double YVV = -99.99, YVR = -5.49, YRV = -5.49, YRR = -8.8, NVV = -5.49, NVR = -8.8, NRV = -8.8, NRR = -3.49;
constexpr double MASS = 30.0, IZ = 4.1, BEAM = 0.41, C_THR = 0.78;
// ---- usv_asmc.py:27-36 ASMC gains ----------------------------------------------------------
constexpr double K_U = 0.1, K_PSI = 0.2, KMIN_U = 0.05, KMIN_PSI = 0.2, K2_U = 0.02, K2_PSI = 0.1;
constexpr double MU_U = 0.05, MU_PSI = 0.1, LAMBDA_U = 0.001, LAMBDA_PSI = 1.0;
// ---- usv_asmc.py:39-41,47 ------------------------------------------------------------------
constexpr double F1 = 2.0, F2 = 2.0, F3 = 2.0, H_STEP = 0.01;
constexpr double HALF_H = H_STEP / 2;  // h*(a+b)/2 == (h/2)*(a+b) bit for bit (halving is exact)

// mass matrix entries (usv_asmc.py:172-174) and its closed-form inverse
constexpr double M00 = MASS - X_U_DOT, M11 = MASS - Y_V_DOT, M12 = 0 - Y_R_DOT, M21 = 0 - N_V_DOT, M22 = IZ - N_R_DOT;
constexpr double M_DET = M11 * M22 - M12 * M21;
constexpr double MI00 = 1.0 / M00, MI11 = M22 / M_DET, MI12 = -M12 / M_DET, MI21 = -M21 / M_DET, MI22 = M11 / M_DET;

// hydrodynamic coefficient factors, usv_asmc.py:101-108 (same literals; pi is 3.141592 there)
constexpr double K_YV = 1.1 + 0.0045 * (1.01 / 0.09) - 0.1 * (0.27 / 0.09) + 0.016 * ((0.27 / 0.09) * (0.27 / 0.09));
constexpr double C_YR = 6 * (-3.141592 * 1000) * 0.09 * 0.09 * 1.01;
constexpr double C_NV = 0.06 * (-3.141592 * 1000) * 0.09 * 0.09 * 1.01;
constexpr double C_NR = 0.02 * (-3.141592 * 1000) * 0.09 * 0.09 * 1.01 * 1.01;

// Coriolis terms (usv_asmc.py:201-211) folded:  C02 = CV*v + CR*r, C12 = CU*u, C20 = -C02, C21 = -CU*u
constexpr double COR_V = -MASS + 2 * Y_V_DOT;          // -30 + 2*(-23.13)
constexpr double COR_R = Y_R_DOT + N_V_DOT;            // 2 * ((Y_r_dot + N_v_dot)/2)
constexpr double COR_U = MASS - X_U_DOT * MASS;        // 30 + 67.5

constexpr double PI = 3.14159265358979323846;
constexpr double TWO_PI = 2 * PI;
constexpr double TWO_PI_HI = 6.283185307179586232;     // double(2*pi)
constexpr double TWO_PI_LO = 2.4492935982947064e-16;   // 2*pi - double(2*pi)

// Constants of the sub-step for device code.  A float64 literal whose low 32 bits are not zero costs two
// UMOVs at every use on sm_100a, so those live in __constant__ memory (one LDCU.128 brings two of them into
// uniform registers) and equal values share one slot to keep the whole set inside the uniform register
// file; literals that fit the 32-bit immediate form (100, 4, -25, 32.25, 97.5, 0.03125 ...) stay literals.
#ifndef USV_FP64_DIET
#define USV_FP64_DIET 1
#endif
struct ModelK {
    double half_h;                  // h/2
    double c0001;                   // 0.001: epsilon at :72 and lambda_u
    double u_switch, xu_fast, xuu_fast;  // :95-99
    double yv_k, c_yr, c_nv, c_nr;  // :101-108
    double fu_vr, fpsi_uv;          // m - Y_v_dot, -X_u_dot + Y_v_dot (:113-116)
    double inv_m00, inv_m22, m22;
    double c005, c01, c02, c002;    // 0.05 = kmin_u = mu_u | 0.1 = k_u = mu_psi = k2_psi | 0.2 = k_psi = kmin_psi | 0.02 = k2_u
    double cor_v, cor_r;
    double yvv, d549, d88, nrr;     // Yvv | Yvr = Yrv = Nvv | Yrr = Nvr = Nrv | Nrr
    double mi11, mi12, mi21, mi22;
    double pi, two_pi;
    double inv_beam, inv_2c, inv_bc, c_thr, half_beam;  // INFO path only
    double yv_yvv;                  // -20000 K_YV + Yvv: the two |v| terms of -D11 (USV_FP64_DIET)
};
static_assert(K_U == 0.1 && MU_PSI == 0.1 && K2_PSI == 0.1 && KMIN_U == 0.05 && MU_U == 0.05 && K_PSI == 0.2 &&
              KMIN_PSI == 0.2 && K2_U == 0.02 && LAMBDA_U == 0.001 && LAMBDA_PSI == 1.0, "shared constant slots");
static_assert(YVR == -5.49 && YRV == -5.49 && NVV == -5.49 && YRR == -8.8 && NVR == -8.8 && NRV == -8.8, "shared slots");
static_assert(F1 == 2.0 && F2 == 2.0 && F3 == 2.0 && H_STEP == 0.01 && M00 == 32.25 && COR_U == 97.5, "immediates");
constexpr double ROT_MAX = 0.03125;  // largest heading change per sub-step the carried rotation handles

__constant__ ModelK MK = {
    HALF_H,
    0.001,
    1.2, 64.55, -70.92,
    -20000.0 * K_YV, C_YR, C_NV, C_NR,
    MASS - Y_V_DOT, -X_U_DOT + Y_V_DOT,
    1.0 / M00, 1.0 / M22, M22,
    0.05, 0.1, 0.2, 0.02,
    COR_V, COR_R,
    YVV, -5.49, -8.8, NRR,
    MI11, MI12, MI21, MI22,
    PI, TWO_PI,
    1.0 / BEAM, 1.0 / (2 * C_THR), 1.0 / (BEAM * C_THR), C_THR, 0.5 * BEAM,
    -20000.0 * K_YV + YVV};

// numpy semantics
__device__ __forceinline__ double np_sign(double x) { return (x > 0.0) ? 1.0 : ((x < 0.0) ? -1.0 : 0.0); }
__device__ __forceinline__ double np_clip(double x, double lo, double hi) { return fmin(fmax(x, lo), hi); }
// usv_asmc.py:120  np.where(|e| > pi, sign(e)*(|e| - 2 pi), e)
__device__ __forceinline__ double wrap_once(double e) {
    const double a = fabs(e);
    // sign(e) (|e| - 2 pi) == e - sign(e) 2 pi bit for bit: the sign flips for pi < |e| < 2 pi and stays beyond 2 pi
    return (a > MK.pi) ? e - copysign(MK.two_pi, e) : e;
}
// angle reduced to [-pi, pi]  (simple_env.py:64-65 computes atan2(sin a, cos a))
__device__ __forceinline__ double wrap_pi(double a) {
    const double k = rint(a * (1.0 / TWO_PI));
    return fma(-k, TWO_PI_LO, fma(-k, TWO_PI_HI, a));
}

// Controller + integrator memory of one env (usv_asmc.py:43-49); 16 independent scalars because
// so_filter's (o, o_dot, o_dot_dot) equal their *_last twins after every sub-step (:90-92).
struct Ctrl {
    double xd_l, yd_l, psid_l;      // eta_dot_last
    double ud_l, vd_l, rd_l;        // upsilon_dot_last
    double eu_l, kadu_l, kadp_l;    // e_u_last, ka_dot_u_last, ka_dot_psi_last
    double psi_d_l, odd_l, od_l, o_l;  // psi_d_last, o_dot_dot_last, o_dot_last, o_last
    double eu_int, ka_u, ka_psi;    // aux_vars
};

struct Pose {
    double x, y, psi, u, v, r;
};

// k * sign(|sigma| - mu)   (usv_asmc.py:137-140; np.sign(0) = 0)
__device__ __forceinline__ double k_sign(double abs_sigma, double mu, double k) {
    return (abs_sigma > mu) ? k : ((abs_sigma < mu) ? -k : 0.0);
}

// One iteration of the loop at usv_asmc.py:56-243.  cs / sn hold cos / sin of the CURRENT psi on entry
// (the reference builds J from the old psi, :179-181) and of the NEW psi on exit: they are carried by a
// small-angle rotation (usv_math.cuh) with a full sincos() only if the heading moved > 1/32 rad in one
// 0.01 s sub-step.  INFO: also produce psi_d, u_d, tport, tstbd and follow the thruster round trip.
template <bool PERTURB, bool INFO>
__device__ __forceinline__ void asmc_substep(Pose& p, Ctrl& c, const double a0, const double a1, double& cs,
                                             double& sn, const long long perturb_step, double* info) {
    const double u = p.u, v = p.v, r = p.r, psi = p.psi;
    const double au = fabs(u), av = fabs(v), ar = fabs(r);
    const double half_h = MK.half_h;

    // :72-77  sideslip, course, desired heading (no wrap: :74,:81 are commented out)
    const double sp = sqrt_m(fma(u, u, v * v));  // hypot(u,v) (:72) and sqrt(u^2+v^2) (:103)
    const double beta = asin_k(div_k(v, MK.c0001 + sp));
    const double psi_d = (psi + beta) + a1;

    // :84-92  second-order filter for the desired yaw rate; f1 = f2 = f3 = 2, so
    // ((r_d - o_last) f1 - f3 o_dot_last) f2 == 4 ((r_d - o_last) - o_dot_last) bit for bit
    double r_d = (psi_d - c.psi_d_l) * 100.0;  // 1/h
    c.psi_d_l = psi_d;
    const double odd = ((r_d - c.o_l) - c.od_l) * 4.0;
    const double od = fma(half_h, odd + c.odd_l, c.od_l);
    const double o = fma(half_h, od + c.od_l, c.o_l);
    r_d = o;
    c.o_l = o;
    c.od_l = od;
    c.odd_l = odd;

    // :95-108  speed-dependent hydrodynamic coefficients
    const bool fast = au > MK.u_switch;
    const double Xu = fast ? MK.xu_fast : -25.0;
    const double Xuu = fast ? MK.xuu_fast : 0.0;
#if !USV_FP64_DIET
    const double Yv = MK.yv_k * av;
#endif
    const double Yr = MK.c_yr * sp, Nv = MK.c_nv * sp, Nr = MK.c_nr * sp;

    // :111-116  f, g of the simplified model
#if USV_FP64_DIET
    // f_u g_u^-1 and f_psi g_psi^-1: the reference divides by (m - X_u_dot) / (Iz - N_r_dot) here (:113-116) and multiplies
    // by the same numbers at :154-155; the numerators go straight into Tx / Tz below
    const double fu_n = fma(MK.fu_vr * v, r, fma(Xuu, au, Xu * u));  // Xuu |u| + Xu u, as written at :113
    const double fp_n = fma(MK.fpsi_uv * u, v, Nr * r);
#else
    const double f_u = fma(MK.fu_vr * v, r, fma(Xuu, au, Xu * u)) * MK.inv_m00;  // Xuu |u| + Xu u, as written at :113
    const double f_psi = fma(MK.fpsi_uv * u, v, Nr * r) * MK.inv_m22;
#endif

    // :119-130  errors
    const double e_psi = wrap_once(psi_d - psi);
    const double e_psi_dot = r_d - r;
    const double e_u = a0 - u;
    c.eu_int = fma(half_h, e_u + c.eu_l, c.eu_int);
    c.eu_l = e_u;

    // :133-147  sliding surfaces and adaptive gains
    const double sig_u = fma(MK.c0001, c.eu_int, e_u);
    const double sig_p = e_psi_dot + e_psi;  // lambda_psi = 1
    const double asu = fabs(sig_u), asp = fabs(sig_p);
    const double kadu = (c.ka_u > MK.c005) ? k_sign(asu, MK.c005, MK.c01) : MK.c005;
    const double kadp = (c.ka_psi > MK.c02) ? k_sign(asp, MK.c01, MK.c02) : MK.c02;
    c.ka_u = fma(half_h, kadu + c.kadu_l, c.ka_u);
    c.kadu_l = kadu;
    c.ka_psi = fma(half_h, kadp + c.kadp_l, c.ka_psi);
    c.kadp_l = kadp;

    // :150-151  ua = -Ka sqrt|sigma| sign(sigma) - k2 sigma
    // :154-155  control forces; dividing by g = 1/M is multiplying by M
#if USV_FP64_DIET
    const double nua_u = fma(c.ka_u, copysign(sqrt_m(asu), sig_u), MK.c002 * sig_u);  // -ua_u
    const double nua_p = fma(c.ka_psi, copysign(sqrt_m(asp), sig_p), MK.c01 * sig_p);
    const double tx = fma(32.25, fma(MK.c0001, e_u, nua_u), -fu_n);  // (lambda_u e_u - ua_u) (m - X_u_dot) - f_u (m - X_u_dot)
    const double tz = fma(MK.m22, e_psi + nua_p, -fp_n);
#else
    const double ua_u = -fma(c.ka_u, copysign(sqrt_m(asu), sig_u), MK.c002 * sig_u);
    const double ua_p = -fma(c.ka_psi, copysign(sqrt_m(asp), sig_p), MK.c01 * sig_p);
    const double tx = (fma(MK.c0001, e_u, -f_u) - ua_u) * 32.25;  // m - X_u_dot
    const double tz = ((e_psi - f_psi) - ua_p) * MK.m22;
#endif

    // :158-176  thruster allocation and back: T[0] = tport + c tstbd = tx, T[2] = B/2 (tport - c tstbd) = tz
    double T0 = tx, T1 = 0.0, T2 = tz;
    if (INFO) {
        const double tport = fma(tz, MK.inv_beam, tx * 0.5);
        const double tstbd = fma(-tz, MK.inv_bc, tx * MK.inv_2c);
        info[0] = psi_d;
        info[1] = a0;
        info[2] = tport;
        info[3] = tstbd;
        T0 = fma(MK.c_thr, tstbd, tport);
        T2 = MK.half_beam * fma(-MK.c_thr, tstbd, tport);
    }
    // :184-199  sinusoidal disturbance, row vector [fx, fy, 0] @ J
    if (PERTURB) {
        const double t = (double)perturb_step * H_STEP;
        const double k = 10 * TWO_PI;
        const double fx = cos(t * k) * 5, fy = cos(t + k + 10) * 5;
        T0 += fx * cs + fy * sn;
        T1 += fy * cs - fx * sn;
    }

    // :201-227  nu_dot = M^-1 (T - C nu - D nu);  D = Dl - Dn with Dl = -diag/offdiag(X,Y,N), so
    // -(D nu) = (Xu + Xuu|u|) u etc.: the signs are folded instead of negating every entry
    const double c02 = fma(MK.cor_v, v, MK.cor_r * r);
    const double cu = 97.5 * u;  // m - X_u_dot m
#if USV_FP64_DIET
    // the same sums with the common factors of v and r taken out and the |v| terms of -D11 under one coefficient
    const double t_vr = fma(MK.d549, av, MK.d88 * ar);
    const double nd11 = fma(MK.yv_yvv, av, MK.d549 * ar);  // -D11
    const double nd12 = Yr + t_vr;
    const double nd21 = Nv + t_vr;
    const double nd22 = Nr + fma(MK.d88, av, MK.nrr * ar);
    const double rhs0 = fma(fma(Xuu, au, Xu), u, fma(-c02, r, T0));  // -D00 u = (Xu + Xuu |u|) u
    const double rhs1 = PERTURB ? T1 + fma(nd11, v, (nd12 - cu) * r) : fma(nd11, v, (nd12 - cu) * r);
    const double rhs2 = fma(nd22, r, fma(cu + nd21, v, fma(c02, u, T2)));
#else
    const double nd11 = Yv + fma(MK.yvv, av, MK.d549 * ar);  // -D11
    const double nd12 = Yr + fma(MK.d549, av, MK.d88 * ar);
    const double nd21 = Nv + fma(MK.d549, av, MK.d88 * ar);
    const double nd22 = Nr + fma(MK.d88, av, MK.nrr * ar);
    const double rhs0 = fma(fma(Xuu, au, Xu), u, fma(-c02, r, T0));  // -D00 u = (Xu + Xuu |u|) u
    const double rhs1 = fma(-cu, r, T1) + fma(nd11, v, nd12 * r);
    const double rhs2 = (T2 + fma(c02, u, cu * v)) + fma(nd21, v, nd22 * r);
#endif
    const double ud = MK.inv_m00 * rhs0;
    const double vd = fma(MK.mi11, rhs1, MK.mi12 * rhs2);
    const double rd = fma(MK.mi21, rhs1, MK.mi22 * rhs2);

    // :228-235  trapezoid integration of nu, then eta with J(old psi) * nu_new
    const double un = fma(half_h, ud + c.ud_l, u);
    const double vn = fma(half_h, vd + c.vd_l, v);
    const double rn = fma(half_h, rd + c.rd_l, r);
    c.ud_l = ud;
    c.vd_l = vd;
    c.rd_l = rd;
    const double xd = fma(cs, un, -sn * vn);
    const double yd = fma(sn, un, cs * vn);
    const double dpsi = half_h * (rn + c.psid_l);
    p.x = fma(half_h, xd + c.xd_l, p.x);
    p.y = fma(half_h, yd + c.yd_l, p.y);
    p.psi = psi + dpsi;
    c.xd_l = xd;
    c.yd_l = yd;
    c.psid_l = rn;
    p.u = un;
    p.v = vn;
    p.r = rn;
    if (fabs(dpsi) <= ROT_MAX) rot_small(dpsi, cs, sn);
    else sincos(p.psi, &sn, &cs);
}

}  // namespace usv
