// usv_math.cuh -- float64 device math for the step kernels, written for the sm_100a FP64 pipe.
//
// Why not the CUDA math library: on sm_100a every float64 literal that is not representable in its
// high 32 bits costs two UMOV issue slots per use (the r01b profile of dynamics_kernel showed 26 % of
// all issued instructions were UMOVs feeding the polynomial coefficients of sincos()/asin(), and the
// general-purpose routines carry range / special-case branches this path never takes).  Here
//   * coefficients live in __constant__ memory (one LDCU.128 brings two of them into uniform registers),
//   * sqrt and reciprocal are a MUFU seed + Newton steps without the slow-path checks (arguments are
//     sums of squares and bounded denominators),
//   * the heading's cos/sin are carried through the 20 sub-steps by a small-angle rotation instead of
//     20 full-range sincos() calls.
// Accuracy (tests/test_gpu_parity.py::test_device_math_probe): sqrt_k, sqrt_m, div_k <= 1 ulp, asin_k <= 2 ulp,
// rot_small <= 1 ulp per application (|d| <= 1/32) -- the same order as the libm-vs-CUDA differences the parity
// tolerance already covers.
// The dynamics kernel is bound by the FP64 pipe, i.e. by the NUMBER of FP64 instructions (profiles/r02_experiments.txt 11-12):
// the build-time knobs below (USV_SQRT_CUBIC, USV_ROT_SHORT, USV_ASIN_HORNER, USV_HIWORD) each trade instructions for nothing
// but the place of a rounding; their defaults are the measured optimum, the other values are kept for A/B builds
// (tools/build_variants.sh).
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace usv {

struct MathK {
    double asin_c[13];  // tools/gen_asin_poly.py: (asin(sqrt z)/sqrt z - 1)/z on z in [0, 0.25], degree 12
    double sin_c[3];    // -1/3!, 1/5!, -1/7!
    double cos_c[3];    // 1/4!, -1/6!, 1/8!   (-1/2! is an immediate)
    double pio2_hi, pio2_lo;
    double ksin[6];     // fdlibm __kernel_sin S1..S6: sin r = r + r^3 (S1 + z (S2 + ...)), |r| <= pi/4
    double kcos[6];     // fdlibm __kernel_cos C1..C6: cos r = 1 - z/2 + z^2 (C1 + z (C2 + ...))
};

__constant__ MathK MATHK = {
    {0.16666666666666669, 0.07499999999998433, 0.04464285714635543, 0.030381944138531247, 0.02237217294214989,
     0.017352392720869973, 0.013971212973552933, 0.011479177415184906, 0.01032281435018578, 0.005457506718640358,
     0.01740087944269402, -0.014851887071247204, 0.028757851367421566},
    {-1.0 / 6, 1.0 / 120, -1.0 / 5040},
    {1.0 / 24, -1.0 / 720, 1.0 / 40320},
    1.5707963267948966, 6.123233995736766e-17,
    {-1.66666666666666324348e-01, 8.33333333332248946124e-03, -1.98412698298579493134e-04, 2.75573137070700676789e-06,
     -2.50507602534068634195e-08, 1.58969099521155010221e-10},
    {4.16666666666666019037e-02, -1.38888888888741095749e-03, 2.48015872894767294178e-05, -2.75573143513906633035e-07,
     2.08757232129817482790e-09, -1.13596475577881948265e-11}};

// MUFU.RSQ64H / MUFU.RCP64H: ~20-bit seeds computed from the high word
__device__ __forceinline__ double rsqrt_seed(double a) {
    double r;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(a));
    return r;
}
__device__ __forceinline__ double rcp_seed(double a) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(a));
    return r;
}

// USV_HIWORD: the zero / subnormal guard of the sub-step's square root compares the HIGH WORD of its (never negative)
// argument as an integer: one ISETP on the idle integer pipe instead of a DSETP on the FP64 pipe, which is what bounds the
// sub-step (r02u: -0.2 us per step).  The same trick on the range split of asin_k (a >= 0.5) and on the small-angle bound
// of rot_small made the kernel SLOWER (r02v: +0.35 us per step), and so did the model's exact thresholds as 64-bit integer
// compares (r02u3: +0.9 us): those stay DSETP.
#ifndef USV_HIWORD
#define USV_HIWORD 1
#endif

// sqrt(a) for 0 <= a << DBL_MAX (sums of squares, |sigma|, discriminants); exactly 0 for a == 0 and for subnormal a
// (rsqrt.approx.ftz flushes a subnormal argument to zero and returns +inf: without the guard a |sigma| or a
// discriminant that underflows would turn into NaN and poison the env's state; sqrt of a subnormal is < 1.5e-154).
// half_rinv receives ~1/(2 sqrt a) (garbage when a is below DBL_MIN), which callers reuse as a free reciprocal.
constexpr double SQRT_MIN_ARG = 2.2250738585072014e-308;  // DBL_MIN: its low word is zero, so it is a 32-bit immediate
__device__ __forceinline__ double sqrt_k(double a, double& half_rinv) {
    const double r = rsqrt_seed(a);
    double h = 0.5 * r;
    double g = a * r;
    g = fma(fma(-g, g, a), h, g);          // Newton on g: ~2^-40
    h = fma(fma(-r, g, 1.0), h, h);        // refine h with the new g
    g = fma(fma(-g, g, a), h, g);          // final correction: <= 1 ulp
    half_rinv = h;
    return (a >= SQRT_MIN_ARG) ? g : 0.0;
}
// The same without the refinement of h: with h fixed at the seed's 1/(2 sqrt a) (1 + d), |d| ~ 2^-20, one correction
// g + (a - g^2) h leaves the error -e d - e^2 / (2 sqrt a): 2^-20 -> 2^-40 -> 2^-60, then the rounding of the last
// fma.  Two FP64 instructions and 16 cycles of dependent latency less than the form above -- the sub-step's critical
// path runs through two of these.  h20 receives the unrefined h (good to ~2^-20, enough for a correction term).
__device__ __forceinline__ double sqrt_k20(double a, double& h20) {
    const double r = rsqrt_seed(a);
    const double h = 0.5 * r;
    double g = a * r;
    g = fma(fma(-g, g, a), h, g);
    g = fma(fma(-g, g, a), h, g);
    h20 = h;
    return (a >= SQRT_MIN_ARG) ? g : 0.0;
}
#ifndef USV_SQRT_CUBIC
#define USV_SQRT_CUBIC 2
#endif
#ifndef USV_ROT_SHORT
#define USV_ROT_SHORT 1
#endif
// sqrt(a) with one third-order step instead of two Newton steps.  With the seed r = (1 + d) / sqrt(a), |d| ~ 2^-20, and
// t = a r:  e = 1 - t r = 1 - a r^2 (one fma, exact up to its rounding), and sqrt(a) = t / sqrt(1 - e) =
// t (1 + e/2 + 3 e^2 / 8 + O(e^3)); the truncation 5 e^3 / 16 is ~2^-59.  Five FP64 instructions in a chain of four
// instead of six in a chain of five; one more live register (the sensor kernel at 64 registers spills with it).
__device__ __forceinline__ double sqrt_cubic(double a) {
    const double r = rsqrt_seed(a);
    const double t = a * r;
    const double e = fma(-t, r, 1.0);
    const double g = fma(t * e, fma(0.375, e, 0.5), t);
#if USV_HIWORD  // a >= DBL_MIN as an integer compare of the high word (a is never negative here)
    return (__double2hiint(a) >= 0x00100000) ? g : 0.0;
#else
    return (a >= SQRT_MIN_ARG) ? g : 0.0;
#endif
}
__device__ __forceinline__ double sqrt_k(double a) {
#if USV_SQRT_CUBIC >= 2
    return sqrt_cubic(a);
#else
    double h;
    return sqrt_k20(a, h);
#endif
}
// the three square roots of a sub-step (usv_model.cuh)
__device__ __forceinline__ double sqrt_m(double a) {
#if USV_SQRT_CUBIC >= 1
    return sqrt_cubic(a);
#else
    return sqrt_k(a);
#endif
}

// 1/d for a normal, finite d bounded away from 0 (no special cases)
__device__ __forceinline__ double rcp_k(double d) {
    double r = rcp_seed(d);
    double e = fma(-d, r, 1.0);
    e = fma(e, e, e);
    r = fma(r, e, r);  // cubic step
    e = fma(-d, r, 1.0);
    return fma(r, e, r);
}
// n/d (<= 1 ulp): one Newton step takes the seed to ~2^-40, q = n r is then corrected with its exact residual
// (n - d q) r, which leaves 2^-80 plus the rounding of the last fma.  (The reciprocal itself is never refined
// further: three FP64 instructions and 32 cycles of dependent latency less than refining r first.)
__device__ __forceinline__ double div_k(double n, double d) {
    double r = rcp_seed(d);
    r = fma(r, fma(-d, r, 1.0), r);
    const double q = n * r;
    return fma(fma(-d, q, n), r, q);
}

#ifndef USV_ASIN_HORNER
#define USV_ASIN_HORNER 1
#endif
// asin(x), |x| <= 1.  One polynomial serves both ranges: |x| <= 0.5 directly, otherwise through
// asin(x) = pi/2 - 2 asin(sqrt((1 - |x|)/2)).
__device__ __forceinline__ double asin_k(double x) {
    const double a = fabs(x);
    const bool big = a > 0.5;
    double z = a * a, s = a, corr = 0.0;
    if (big) {
        z = fma(-0.5, a, 0.5);  // exact for a in [0.5, 1]
        double h;
        s = sqrt_k20(z, h);
        corr = fma(-s, s, z) * h;  // sqrt(z) - s, so that s + corr carries the rounding error of s
        if (!(z > 0.0)) corr = 0.0;
    }
    // degree-12 polynomial by Estrin's scheme: 4 dependent fma levels instead of Horner's 12 (the sub-step's critical
    // path runs through it); z <= 0.25, so the regrouping costs nothing in accuracy (probe: <= 2 ulp as before)
    const double* c = MATHK.asin_c;
#if USV_ASIN_HORNER == 2  // plain Horner: 12 FP64 instructions in a chain of 12
    double p = fma(c[12], z, c[11]);
#pragma unroll
    for (int k = 10; k >= 0; --k) p = fma(p, z, c[k]);
#elif USV_ASIN_HORNER == 1  // two interleaved Horner chains in z^2 (even / odd coefficients): 13 FP64 instructions in a chain of 8
    const double z2 = z * z;
    double pe = fma(c[12], z2, c[10]), po = fma(c[11], z2, c[9]);
    pe = fma(pe, z2, c[8]);
    po = fma(po, z2, c[7]);
    pe = fma(pe, z2, c[6]);
    po = fma(po, z2, c[5]);
    pe = fma(pe, z2, c[4]);
    po = fma(po, z2, c[3]);
    pe = fma(pe, z2, c[2]);
    po = fma(po, z2, c[1]);
    pe = fma(pe, z2, c[0]);
    const double p = fma(po, z, pe);
#else
    const double z2 = z * z, z4 = z2 * z2, z8 = z4 * z4;
    const double a0 = fma(c[1], z, c[0]), a1 = fma(c[3], z, c[2]), a2 = fma(c[5], z, c[4]);
    const double a3 = fma(c[7], z, c[6]), a4 = fma(c[9], z, c[8]), a5 = fma(c[11], z, c[10]);
    const double b0 = fma(a1, z2, a0), b1 = fma(a3, z2, a2), b2 = fma(a5, z2, a4);
    const double d0 = fma(b1, z4, b0), d1 = fma(c[12], z4, b2);
    const double p = fma(d1, z8, d0);
#endif
    const double t = fma(s * z, p, corr);  // asin(s) - s
    double r = s + t;
    if (big) r = (MATHK.pio2_hi - 2.0 * s) - (fma(2.0, t, -MATHK.pio2_lo));
    return copysign(r, x);
}

// sin / cos of a BOUNDED argument, |x| <= 64 (the reset's angles lie in [-pi, 2 pi]): quadrant by rint(x 2/pi), two-word
// pi/2 with fma, fdlibm kernel polynomials; <= 1 ulp of 1.0 absolute.  No range-reduction slow path, hence no call into
// the CUDA math library: the fused step kernel runs sincos() in its register-rich dynamics role, and ptxas 12.9 crashes
// on a library routine that is reachable from both sides of a setmaxnreg split (usv_fused.cuh).
__device__ __forceinline__ void sincos_bounded(double x, double* sn, double* cs) {
    const double kf = rint(x * 0.63661977236758134308);  // 2/pi
    const int k = (int)kf;
    double r = fma(-kf, MATHK.pio2_hi, x);
    r = fma(-kf, MATHK.pio2_lo, r);
    const double z = r * r;
    const double* S = MATHK.ksin;
    const double* C = MATHK.kcos;
    double ps = fma(S[5], z, S[4]);
    ps = fma(ps, z, S[3]);
    ps = fma(ps, z, S[2]);
    ps = fma(ps, z, S[1]);
    ps = fma(ps, z, S[0]);
    const double sr = fma(r * z, ps, r);
    double pc = fma(C[5], z, C[4]);
    pc = fma(pc, z, C[3]);
    pc = fma(pc, z, C[2]);
    pc = fma(pc, z, C[1]);
    pc = fma(pc, z, C[0]);
    const double cr = fma(z * z, pc, fma(-0.5, z, 1.0));
    const double s1 = (k & 1) ? cr : sr, c1 = (k & 1) ? sr : cr;
    *sn = (k & 2) ? -s1 : s1;
    *cs = ((k + 1) & 2) ? -c1 : c1;
}

// (cs, sn) <- rotation of (cs, sn) by d, |d| <= 1/32 (Taylor to d^7 / d^8: truncation < 1e-19 relative)
__device__ __forceinline__ void rot_small(double d, double& cs, double& sn) {
    const double z = d * d;
    const double* S = MATHK.sin_c;
    const double* C = MATHK.cos_c;
    double ps = fma(S[2], z, S[1]);
    ps = fma(ps, z, S[0]);
    const double sd = fma(d * z, ps, d);  // sin d
#if USV_ROT_SHORT  // without the d^8 / 8! term: |d| <= 1/32 makes it < 2.3e-17 of cos d ~ 1, i.e. below half an ulp
    double pc = fma(C[1], z, C[0]);
#else
    double pc = fma(C[2], z, C[1]);
    pc = fma(pc, z, C[0]);
#endif
    pc = fma(pc, z, -0.5);
    const double cm1 = z * pc;  // cos d - 1
    const double c0 = cs, s0 = sn;
    cs = fma(-s0, sd, fma(c0, cm1, c0));
    sn = fma(c0, sd, fma(s0, cm1, s0));
}

}  // namespace usv
