// usv_sensor.cuh -- warp-cooperative range-sensor pass for ONE env.
//
// Restates UsvAsmcCaEnv.compute_sensor_measurments (gym_usv/envs/usv_asmc_ca_env.py:411-461,
// 500-519) as called from UsvSimpleEnv._compute_sensor_measurment (simple_env.py:203-226).
//
// The reference walks the obstacles in ascending (centre distance - radius) order and takes the
// FIRST one each ray hits; the equivalent sort-free rule is "among the obstacles a ray hits, the
// one with the smallest key wins" (SURVEY.md H2).  One warp owns one env:
//   phase A  lane = obstacle: offset rotated into the body frame, key, and the wedge of rays the circle
//            can touch (+-asin(r/d) around its bearing, found in float32 and widened by a margin),
//            clipped to the 4 pi / 3 fan; a warp scan turns the wedge sizes into offsets of a compacted
//            list of (obstacle, ray) candidate pairs (~100 per env instead of 29 x 128);
//   phase B  lane = candidate pair, 32 per pass, all lanes busy: the reference's float64 hit test; a
//            hit sets the obstacle's bit in the ray's hit mask and leaves its reading;
//   phase C  lane = ray (4 passes of 32): no hit -> max range; one hit -> that reading; several hits
//            (overlapping circles, rare) -> smallest key wins and its reading is recomputed.
// All hit / miss / ordering decisions are taken in float64.  In the body frame the ray directions are
// the constant table (cos phi_i, sin phi_i), so along = cos phi bx + sin phi by and
// perp = sin phi bx - cos phi by equal the reference's expressions (usv_asmc_ca_env.py:516-518) with
// cos / sin (psi + phi_i) expanded.
#pragma once
#include "usv_math.cuh"
#include "usv_state.cuh"

namespace usv {

struct SensorScratch {  // per warp
    double bx[MAX_OBST], by[MAX_OBST], rr2[MAX_OBST], key[MAX_OBST];
    double hval[N_SENS];        // reading left by a hitting pair (valid when the ray has exactly one hit)
    uint32_t hitmask[N_SENS];   // bit j: obstacle j hits this ray
    int off[MAX_OBST];          // exclusive prefix sum of the candidate counts
    uint32_t seg[MAX_OBST];     // first ray of piece 1 | rays in piece 1 << 8 | first ray of piece 2 << 16
};

// reading of ray ri on obstacle j, whatever it is (phase C recomputation)
__device__ __forceinline__ double ray_reading(const SensorScratch& sc, const double2* raytab, int ri, int j) {
    const double2 t = raytab[ri];
    const double bx = sc.bx[j], by = sc.by[j];
    const double along = fma(t.x, bx, t.y * by);
    const double perp = fma(t.y, bx, -(t.x * by));
    return along - sqrt_k(fma(-perp, perp, sc.rr2[j]));
}

// ox, oy, rr: this lane's obstacle (ignored when lane >= nob).  cps/sps = cos/sin(psi).
// out[p]: reading of ray (lane + 32 p), SENSOR_MAX_RANGE when nothing is hit; hit[p] says which.
// min_key_lt / min_hit_lt: warp-uniform "min over obstacles of key < key_thr" and "min over rays of
// reading < hit_thr" (simple_env.py:334 and :153-156).
__device__ __forceinline__ void sensor_warp(SensorScratch& sc, const double2* __restrict__ raytab, const int lane,
                                            const int nob, const double x, const double y, const double cps,
                                            const double sps, const double ox, const double oy, const double rr,
                                            double out[4], bool hit[4], const double key_thr, bool& min_key_lt,
                                            const double hit_thr, bool& min_hit_lt) {
    constexpr unsigned FULL = 0xFFFFFFFFu;
    reinterpret_cast<uint4*>(sc.hitmask)[lane] = make_uint4(0u, 0u, 0u, 0u);

    // ---- phase A: lane = obstacle ---------------------------------------------------------------
    bool key_lt = false;
    int lo1 = 0, n1 = 0, lo2 = 0, n2 = 0;
    if (lane < nob) {
        const double dx = ox - x, dy = oy - y;  // usv_asmc_ca_env.py:508
        const double d = sqrt_k(fma(dx, dx, dy * dy));
        const double key = d - rr;              // simple_env.py:205-206
        key_lt = key < key_thr;
        const double bx = fma(cps, dx, sps * dy), by = fma(cps, dy, -(sps * dx));  // offset in the body frame
        sc.bx[lane] = bx;
        sc.by[lane] = by;
        sc.rr2[lane] = rr * rr;
        sc.key[lane] = key;
        if (d <= rr) {  // boat inside the circle (only on a terminal step): every ray may hit
            n1 = N_SENS;
        } else {
            const float rel = atan2f((float)by, (float)bx);       // bearing relative to the bow
            const float cf = (rel + 2.0943951f) * INV_PITCH_F;    // its ray index, in [-32, 160]
            const float w = asinf(fminf((float)rr / (float)d, 1.0f)) * INV_PITCH_F + 0.05f;
            const int klo = (int)ceilf(cf - w), khi = (int)floorf(cf + w);
            // rays klo..khi (mod 192) that fall inside the fan 0..127: one piece, plus a second one when the
            // wedge wraps around the stern (khi >= 192 or klo <= -65); the wedge is < 98 rays wide, so the
            // pieces cannot overlap
            lo1 = max(klo, 0);
            n1 = max(min(khi, N_SENS - 1) - lo1 + 1, 0);
            if (khi >= RAYS_FULL_CIRCLE) {
                n2 = min(khi - RAYS_FULL_CIRCLE, N_SENS - 1) + 1;
            } else if (klo < N_SENS - RAYS_FULL_CIRCLE) {
                lo2 = klo + RAYS_FULL_CIRCLE;
                n2 = N_SENS - lo2;
            }
        }
    }
    min_key_lt = __any_sync(FULL, key_lt);
    const int cnt = n1 + n2;
    int incl = cnt;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
        const int v = __shfl_up_sync(FULL, incl, s);
        if (lane >= s) incl += v;
    }
    sc.off[lane] = incl - cnt;
    sc.seg[lane] = (uint32_t)lo1 | ((uint32_t)n1 << 8) | ((uint32_t)lo2 << 16);
    const int total = __shfl_sync(FULL, incl, 31);
    __syncwarp();

    // ---- phase B: lane = candidate pair ------------------------------------------------------------
    for (int t = lane; t < total; t += 32) {
        int j = 0;  // largest j with off[j] <= t: the obstacle whose piece(s) contain pair t
#pragma unroll
        for (int s = 16; s; s >>= 1)
            if (sc.off[j + s] <= t) j += s;
        const int p = t - sc.off[j];
        const uint32_t sg = sc.seg[j];
        const int m1 = (int)((sg >> 8) & 0xFFu);
        const int ri = (p < m1) ? (int)(sg & 0xFFu) + p : (int)(sg >> 16) + (p - m1);
        const double2 tb = raytab[ri];
        const double bx = sc.bx[j], by = sc.by[j];
        const double along = fma(tb.x, bx, tb.y * by);             // usv_asmc_ca_env.py:516
        const double perp = fma(tb.y, bx, -(tb.x * by));           // :518
        const double delta = fma(-perp, perp, sc.rr2[j]);          // :453
        if (along >= 0.0 && delta >= 0.0) {                        // :448, :454
            const double h = along - sqrt_k(delta);                // :457
            if (h < SENSOR_MAX_RANGE) {                            // :458
                atomicOr(&sc.hitmask[ri], 1u << j);
                sc.hval[ri] = h;  // several hits on one ray race here; phase C recomputes those rays
            }
        }
    }
    __syncwarp();

    // ---- phase C: lane = ray ---------------------------------------------------------------------
    bool hit_lt = false;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int ri = lane + 32 * p;
        uint32_t m = sc.hitmask[ri];
        out[p] = SENSOR_MAX_RANGE;
        hit[p] = m != 0u;
        if (m) {
            double h;
            if (m & (m - 1u)) {
                int bj = 0;
                double bkey = INFINITY;
                do {  // ascending j + strict '<': the lower slot wins a tie, as the reference's stable argsort does
                    const int j = __ffs(m) - 1;
                    m &= m - 1u;
                    const double kj = sc.key[j];
                    if (kj < bkey) {
                        bkey = kj;
                        bj = j;
                    }
                } while (m);
                h = ray_reading(sc, raytab, ri, bj);
            } else {
                h = sc.hval[ri];
            }
            out[p] = h;
            hit_lt = hit_lt || (h < hit_thr);
        }
    }
    min_hit_lt = __any_sync(FULL, hit_lt);
    __syncwarp();
}

}  // namespace usv
