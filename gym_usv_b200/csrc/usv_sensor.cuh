// usv_sensor.cuh -- warp-cooperative range-sensor pass for ONE env.
//
// Restates UsvAsmcCaEnv.compute_sensor_measurments (gym_usv/envs/usv_asmc_ca_env.py:411-461,
// 500-519) as called from UsvSimpleEnv._compute_sensor_measurment (simple_env.py:203-226).
//
// The reference walks the obstacles in ascending (centre distance - radius) order and takes the
// FIRST one each ray hits; the equivalent sort-free rule is "among the obstacles a ray hits, the
// one with the smallest key wins" (SURVEY.md H2).  One warp owns one env:
//   phase A  lane = obstacle: offset, key, and the wedge of rays the circle can touch
//            (+-asin(r/d) around its bearing, found in float32 and widened by a margin); the lane
//            sets its bit in a per-ray candidate mask in shared memory;
//   phase B  lane = ray (4 passes of 32): the reference's float64 hit test on the few candidate
//            obstacles of that ray, smallest key wins, one sqrt for the winner.
// All hit / miss / ordering decisions are taken in float64 with the reference's expressions.
#pragma once
#include "usv_state.cuh"

namespace usv {

struct SensorScratch {         // per warp
    uint32_t mask[N_SENS];     // bit j set: obstacle j may hit this ray
    double dx[MAX_OBST], dy[MAX_OBST], rr[MAX_OBST], key[MAX_OBST];
};

// ox, oy, rr: this lane's obstacle (ignored when lane >= nob).  cps/sps = cos/sin(psi).
// out[p]: reading of ray (lane + 32 p), SENSOR_MAX_RANGE when nothing is hit; hit[p] says which.
// min_key_lt / min_hit_lt: warp-uniform "min over obstacles of key < key_thr" and "min over rays of
// reading < hit_thr" (simple_env.py:334 and :153-156).
__device__ __forceinline__ void sensor_warp(SensorScratch& sc, const double2* __restrict__ raytab, const int lane,
                                            const int nob, const double x, const double y, const double cps,
                                            const double sps, const double ox, const double oy, const double rr,
                                            double out[4], bool hit[4], const double key_thr, bool& min_key_lt,
                                            const double hit_thr, bool& min_hit_lt) {
    reinterpret_cast<uint4*>(sc.mask)[lane] = make_uint4(0u, 0u, 0u, 0u);
    __syncwarp();

    // ---- phase A: lane = obstacle ---------------------------------------------------------------
    bool key_lt = false;
    if (lane < nob) {
        const double dx = ox - x, dy = oy - y;  // usv_asmc_ca_env.py:508
        const double d = sqrt_pos(fma(dx, dx, dy * dy));
        const double key = d - rr;              // simple_env.py:205-206
        key_lt = key < key_thr;
        sc.dx[lane] = dx;
        sc.dy[lane] = dy;
        sc.rr[lane] = rr;
        sc.key[lane] = key;
        int klo, khi;
        if (d <= rr) {  // boat inside the circle (only on a terminal step): every ray may hit
            klo = 0;
            khi = N_SENS - 1;
        } else {
            const float fdx = (float)dx, fdy = (float)dy, fc = (float)cps, fs = (float)sps;
            const float bx = fc * fdx + fs * fdy, by = fc * fdy - fs * fdx;  // offset in the body frame
            const float rel = atan2f(by, bx);                                // bearing relative to the bow
            const float cf = (rel + 2.0943951f) * INV_PITCH_F;               // its ray index, in [-32, 160]
            const float w = asinf(fminf((float)rr / (float)d, 1.0f)) * INV_PITCH_F + 0.05f;
            klo = (int)ceilf(cf - w);
            khi = (int)floorf(cf + w);
        }
        const uint32_t bit = 1u << lane;
        for (int k = klo; k <= khi; ++k) {
            int ri = k;
            if (ri < 0) ri += RAYS_FULL_CIRCLE;
            else if (ri >= RAYS_FULL_CIRCLE) ri -= RAYS_FULL_CIRCLE;
            if (ri < N_SENS) atomicOr(&sc.mask[ri], bit);  // the fan only spans 4 pi / 3
        }
    }
    __syncwarp();
    min_key_lt = __any_sync(0xFFFFFFFFu, key_lt);

    // ---- phase B: lane = ray ---------------------------------------------------------------------
    bool hit_lt = false;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int ri = lane + 32 * p;
        uint32_t m = sc.mask[ri];
        out[p] = SENSOR_MAX_RANGE;
        hit[p] = false;
        if (m) {
            const double2 t = raytab[ri];
            const double c = cps * t.x - sps * t.y;  // cos(psi + phi_i)
            const double s = sps * t.x + cps * t.y;  // sin(psi + phi_i)
            double bkey = INFINITY, balong = 0.0, bdelta = 0.0;
            do {
                const int j = __ffs(m) - 1;
                m &= m - 1;
                const double dx = sc.dx[j], dy = sc.dy[j], r = sc.rr[j];
                const double along = c * dx + s * dy;      // usv_asmc_ca_env.py:516
                const double perp = s * dx - c * dy;       // :518
                const double delta = r * r - perp * perp;  // :453
                bool ok = (along >= 0.0) && (delta >= 0.0);  // :448, :454
                // :457-458  d = along - sqrt(delta) < R_max; sqrt >= 0 settles it unless the obstacle is > 100 m away
                if (ok && along >= SENSOR_MAX_RANGE) ok = (along - sqrt(delta)) < SENSOR_MAX_RANGE;
                const double kj = sc.key[j];
                if (ok && kj < bkey) {  // ascending j + strict '<': the lower slot wins a tie, as a stable sort would
                    bkey = kj;
                    balong = along;
                    bdelta = delta;
                }
            } while (m);
            if (bkey < INFINITY) {
                const double h = balong - sqrt_pos(bdelta);  // :457
                out[p] = h;
                hit[p] = true;
                hit_lt = hit_lt || (h < hit_thr);
            }
        }
    }
    min_hit_lt = __any_sync(0xFFFFFFFFu, hit_lt);
    __syncwarp();
}

}  // namespace usv
