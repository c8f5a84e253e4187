// usv_model.cuh -- vessel/controller constants and one ASMC + 3-DOF sub-step (device code).
//
// Restates gym_usv/control/usv_asmc.py:53-244 (reference root-relative) for one environment
// held entirely in registers.  The 3x3 matrices of the reference are never materialised: the
// products C(nu)nu, D(nu)nu and M^-1 rhs are expanded in closed form (verified to
// reproduce the reference to all printed digits in float64).
#pragma once
#include <cuda_runtime.h>
#include <math.h>

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

// numpy semantics
__device__ __forceinline__ double np_sign(double x) { return (double)((x > 0.0) - (x < 0.0)); }
__device__ __forceinline__ double np_clip(double x, double lo, double hi) { return fmin(fmax(x, lo), hi); }
// usv_asmc.py:120  np.where(|e| > pi, sign(e)*(|e| - 2 pi), e)
__device__ __forceinline__ double wrap_once(double e) {
    const double a = fabs(e);
    return (a > PI) ? copysign(a - TWO_PI, e) : e;
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

// One iteration of the loop at usv_asmc.py:56-243.  `cs`/`sn` are cos/sin of the CURRENT psi
// (the reference builds J from the old psi, :179-181).  INFO: also produce psi_d,u_d,tport,tstbd.
template <bool PERTURB, bool INFO>
__device__ __forceinline__ void asmc_substep(Pose& p, Ctrl& c, const double a0, const double a1, const double cs,
                                             const double sn, const long long perturb_step, double* info) {
    const double u = p.u, v = p.v, r = p.r, psi = p.psi;
    const double au = fabs(u), av = fabs(v), ar = fabs(r);

    // :72-77  sideslip, course, desired heading (no wrap: :74,:81 are commented out)
    const double sp = sqrt(fma(u, u, v * v));  // hypot(u,v) (:72) and sqrt(u^2+v^2) (:103)
    const double beta = asin(v / (0.001 + sp));
    const double psi_d = (psi + beta) + a1;

    // :84-92  second-order filter for the desired yaw rate
    double r_d = (psi_d - c.psi_d_l) * (1.0 / H_STEP);
    c.psi_d_l = psi_d;
    const double odd = (((r_d - c.o_l) * F1) - (F3 * c.od_l)) * F2;
    const double od = fma(HALF_H, odd + c.odd_l, c.od_l);
    const double o = fma(HALF_H, od + c.od_l, c.o_l);
    r_d = o;
    c.o_l = o;
    c.od_l = od;
    c.odd_l = odd;

    // :95-108  speed-dependent hydrodynamic coefficients
    const bool fast = au > 1.2;
    const double Xu = fast ? 64.55 : -25.0;
    const double Xuu = fast ? -70.92 : 0.0;
    const double Yv = (-20000.0 * av) * K_YV;
    const double Yr = C_YR * sp, Nv = C_NV * sp, Nr = C_NR * sp;

    // :111-116  f, g of the simplified model
    const double f_u = ((MASS - Y_V_DOT) * v * r + (Xuu * au + Xu * u)) * (1.0 / M00);
    const double f_psi = ((-X_U_DOT + Y_V_DOT) * u * v + Nr * r) * (1.0 / M22);

    // :119-130  errors
    const double e_psi = wrap_once(psi_d - psi);
    const double e_psi_dot = r_d - r;
    const double e_u = a0 - u;
    c.eu_int = fma(HALF_H, e_u + c.eu_l, c.eu_int);
    c.eu_l = e_u;

    // :133-147  sliding surfaces and adaptive gains
    const double sig_u = e_u + LAMBDA_U * c.eu_int;
    const double sig_p = e_psi_dot + LAMBDA_PSI * e_psi;
    const double kadu = (c.ka_u > KMIN_U) ? K_U * np_sign(fabs(sig_u) - MU_U) : KMIN_U;
    const double kadp = (c.ka_psi > KMIN_PSI) ? K_PSI * np_sign(fabs(sig_p) - MU_PSI) : KMIN_PSI;
    c.ka_u = fma(HALF_H, kadu + c.kadu_l, c.ka_u);
    c.kadu_l = kadu;
    c.ka_psi = fma(HALF_H, kadp + c.kadp_l, c.ka_psi);
    c.kadp_l = kadp;

    // :150-151  ua = -Ka sqrt|sigma| sign(sigma) - k2 sigma
    const double ua_u = -(c.ka_u * copysign(sqrt(fabs(sig_u)), sig_u)) - K2_U * sig_u;
    const double ua_p = -(c.ka_psi * copysign(sqrt(fabs(sig_p)), sig_p)) - K2_PSI * sig_p;

    // :154-155  control forces; dividing by g = 1/M is multiplying by M
    const double tx = ((LAMBDA_U * e_u) - f_u - ua_u) * M00;
    const double tz = ((LAMBDA_PSI * e_psi) - f_psi - ua_p) * M22;

    // :158-176  thruster allocation and back: T[0] = tport + c tstbd = tx, T[2] = B/2 (tport - c tstbd) = tz
    double T0 = tx, T1 = 0.0, T2 = tz;
    if (INFO) {
        const double tport = (tx / 2) + (tz / BEAM);
        const double tstbd = (tx / (2 * C_THR)) - (tz / (BEAM * C_THR));
        info[0] = psi_d;
        info[1] = a0;
        info[2] = tport;
        info[3] = tstbd;
        T0 = tport + C_THR * tstbd;
        T2 = 0.5 * BEAM * (tport - C_THR * tstbd);
    }
    // :184-199  sinusoidal disturbance, row vector [fx, fy, 0] @ J
    if (PERTURB) {
        const double t = (double)perturb_step * H_STEP;
        const double k = 10 * TWO_PI;
        const double fx = cos(t * k) * 5, fy = cos(t + k + 10) * 5;
        T0 += fx * cs + fy * sn;
        T1 += fy * cs - fx * sn;
    }

    // :201-227  nu_dot = M^-1 (T - C nu - D nu)
    const double c02 = COR_V * v + COR_R * r;
    const double cv0 = c02 * r;
    const double cv1 = (COR_U * u) * r;
    const double cv2 = -c02 * u - (COR_U * u) * v;
    const double d00 = -Xu - Xuu * au;
    const double d11 = -Yv - (YVV * av + YVR * ar);
    const double d12 = -Yr - (YRV * av + YRR * ar);
    const double d21 = -Nv - (NVV * av + NVR * ar);
    const double d22 = -Nr - (NRV * av + NRR * ar);
    const double rhs0 = T0 - cv0 - d00 * u;
    const double rhs1 = T1 - cv1 - (d11 * v + d12 * r);
    const double rhs2 = T2 - cv2 - (d21 * v + d22 * r);
    const double ud = MI00 * rhs0;
    const double vd = MI11 * rhs1 + MI12 * rhs2;
    const double rd = MI21 * rhs1 + MI22 * rhs2;

    // :228-235  trapezoid integration of nu, then eta with J(old psi) * nu_new
    const double un = fma(HALF_H, ud + c.ud_l, u);
    const double vn = fma(HALF_H, vd + c.vd_l, v);
    const double rn = fma(HALF_H, rd + c.rd_l, r);
    c.ud_l = ud;
    c.vd_l = vd;
    c.rd_l = rd;
    const double xd = cs * un - sn * vn;
    const double yd = sn * un + cs * vn;
    p.x = fma(HALF_H, xd + c.xd_l, p.x);
    p.y = fma(HALF_H, yd + c.yd_l, p.y);
    p.psi = fma(HALF_H, rn + c.psid_l, psi);
    c.xd_l = xd;
    c.yd_l = yd;
    c.psid_l = rn;
    p.u = un;
    p.v = vn;
    p.r = rn;
}

}  // namespace usv
