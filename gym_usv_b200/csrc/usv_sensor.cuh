// usv_sensor.cuh -- warp-cooperative range-sensor pass for ONE env.
//
// Restates UsvAsmcCaEnv.compute_sensor_measurments (gym_usv/envs/usv_asmc_ca_env.py:411-461,
// 500-519) as called from UsvSimpleEnv._compute_sensor_measurment (simple_env.py:203-226).
//
// The reference walks the obstacles in ascending (centre distance - radius) order and takes the
// FIRST one each ray hits; the equivalent sort-free rule is "among the obstacles a ray hits, the
// one with the smallest key wins" (SURVEY.md H2).  One warp owns one env:
//   phase A  lane = obstacle: offset rotated into the body frame, key, and the wedge of rays the circle
//            can touch (+-asin(r/d) around its bearing; cheap float32 approximations widened by a margin
//            -- a few extra candidates cost a test, a missing one would be an error), clipped to the
//            4 pi / 3 fan; a warp scan turns the wedge sizes into offsets of a compacted list of
//            (obstacle, ray) candidate pairs (~140 per env instead of 29 x 128);
//   phase B  lane = candidate pair, 32 per pass, all lanes busy: the reference's float64 hit test; a hit
//            stores its reading in the pair's slot and enters the ray's winner word with a shared-memory
//            atomicMin on (order-preserving float32 key bits, top 24 | pair slot, low 8);
//   phase C  lane = ray (4 passes of 32): winner word -> pair slot -> reading (slot 255 holds "max range").
// All hit / miss decisions are float64.  The winner order is decided on the top 24 bits of the float32-
// rounded keys, which is exact unless two obstacles that hit the SAME ray have keys equal to within 2^-16
// relative: phase B detects exactly that (the atomicMin returns a word with the same key bits) and the
// env then re-does its rays on the exact path (float64 keys, ascending slot order breaks ties like the
// reference's stable argsort); so does an env with more than 255 candidate pairs (boat inside circles).  In the body frame the ray directions are the constant table (cos phi_i, sin phi_i),
// so along = cos phi bx + sin phi by and perp = sin phi bx - cos phi by equal the reference's
// expressions (usv_asmc_ca_env.py:516-518) with cos / sin (psi + phi_i) expanded.
#pragma once
#include "usv_math.cuh"
#include "usv_state.cuh"

namespace usv {

constexpr int PAIR_SLOTS = 256;           // slot 255 is reserved: "no hit" -> SENSOR_MAX_RANGE
constexpr uint32_t NO_HIT = 0xFFFFFFFFu;

struct SensorScratch {  // per warp
    double4 geo[MAX_OBST];      // bx, by (offset in the body frame), radius^2, key
    int4 rec[MAX_OBST];         // split, rbase, rbase2, key bits: pair t of this obstacle is ray t + (t < split ? rbase : rbase2)
    double hpair[PAIR_SLOTS];   // reading of candidate pair t (valid where the pair hit)
    uint32_t best[N_SENS];      // min over hitting obstacles of (key bits & ~255 | pair slot)
    int ne[MAX_OBST];           // slots of the obstacles that have candidates, in order
};

// atan2(y, x) to 2.1e-4 rad (7th-order odd polynomial on the folded octant)
__device__ __forceinline__ float atan2_coarse(float y, float x) {
    const float ax = fabsf(x), ay = fabsf(y);
    const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
    const float a = __fdividef(mn, fmaxf(mx, 1e-30f));
    const float s = a * a;
    float r = fmaf(fmaf(fmaf(-0.0464964749f, s, 0.15931422f), s, -0.327622764f), s * a, a);
    r = (ay > ax) ? 1.57079637f - r : r;
    r = (x < 0.0f) ? 3.14159274f - r : r;
    return copysignf(r, y);
}
// asin(q) for 0 <= q <= 1, never more than 7.5e-5 rad below the true value
__device__ __forceinline__ float asin_coarse(float q) {
    if (q > 0.5f) return asinf(fminf(q, 1.0f));
    const float s = q * q;
    return q * fmaf(s, fmaf(s, fmaf(s, 15.0f / 336, 3.0f / 40), 1.0f / 6), 1.0f);
}

// the reference's test of a ray (direction t in the body frame) against an obstacle g: hit iff along >= 0,
// delta >= 0 and the reading is inside the range (usv_asmc_ca_env.py:448-458)
__device__ __forceinline__ bool ray_test(const double2 t, const double4 g, double& h) {
    const double along = fma(t.x, g.x, t.y * g.y);      // :516
    const double perp = fma(t.y, g.x, -(t.x * g.y));    // :518
    const double delta = fma(-perp, perp, g.z);         // :453
    if (!(along >= 0.0 && delta >= 0.0)) return false;  // :448, :454
    h = along - sqrt_k(delta);                          // :457
    return h < SENSOR_MAX_RANGE;                        // :458
}

// ox, oy, rr: this lane's obstacle (ignored when lane >= nob).  cps/sps = cos/sin(psi).
// The caller sets sc.hpair[PAIR_SLOTS - 1] = SENSOR_MAX_RANGE once (the "no hit" slot).
// out[p]: reading of ray (lane + 32 p), SENSOR_MAX_RANGE when nothing is hit.
// min_key_lt / min_hit_lt: warp-uniform "min over obstacles of key < key_thr" and "min over rays of
// reading < hit_thr" (simple_env.py:334 and :153-156).
__device__ __forceinline__ void sensor_warp(SensorScratch& sc, const double2* __restrict__ raytab, const int lane,
                                            const int nob, const double x, const double y, const double cps,
                                            const double sps, const double ox, const double oy, const double rr,
                                            double out[4], const double key_thr, bool& min_key_lt,
                                            const double hit_thr, bool& min_hit_lt) {
    constexpr unsigned FULL = 0xFFFFFFFFu;
    reinterpret_cast<uint4*>(sc.best)[lane] = make_uint4(NO_HIT, NO_HIT, NO_HIT, NO_HIT);

    // ---- phase A: lane = obstacle ---------------------------------------------------------------
    bool key_lt = false;
    int lo1 = 0, n1 = 0, lo2 = 0, n2 = 0;
    uint32_t kbits = 0u;
    if (lane < nob) {
        const double dx = ox - x, dy = oy - y;  // usv_asmc_ca_env.py:508
        const double d = sqrt_k(fma(dx, dx, dy * dy));
        const double key = d - rr;              // simple_env.py:205-206
        key_lt = key < key_thr;
        const double bx = fma(cps, dx, sps * dy), by = fma(cps, dy, -(sps * dx));  // offset in the body frame
        sc.geo[lane] = make_double4(bx, by, rr * rr, key);
        const uint32_t kb = __float_as_uint((float)key);  // float32 rounding is monotone: it can tie, never invert
        kbits = ((kb & 0x80000000u) ? ~kb : (kb | 0x80000000u)) & ~255u;
        if (d <= rr) {  // boat inside the circle (only on a terminal step): every ray may hit
            n1 = N_SENS;
        } else {
            const float cf = (atan2_coarse((float)by, (float)bx) + 2.0943951f) * INV_PITCH_F;  // ray index of the bearing, [-32, 160]
            const float w = asin_coarse(__fdividef((float)rr, (float)d)) * INV_PITCH_F + 0.05f;  // approximations: < 0.01 ray
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
    const int off = incl - cnt;
    const bool has = cnt > 0;
    const unsigned has_bal = __ballot_sync(FULL, has);
    if (has) sc.ne[__popc(has_bal & ((1u << lane) - 1u))] = lane;
    sc.rec[lane] = make_int4(off + n1, lo1 - off, lo2 - off - n1, (int)kbits);
    const int total = __shfl_sync(FULL, incl, 31);
    __syncwarp();

    // ---- phase B: lane = candidate pair ------------------------------------------------------------
    // Pair t belongs to the last obstacle (among those with candidates) whose offset is <= t.  Per pass of 32
    // pairs: bit k of `starts` marks an obstacle whose list starts at pair base + k, `before` counts those that
    // started earlier, so the owner of pair base + q is ne[before + popc(starts & bits 0..q) - 1].
    bool ambiguous = false;
    int before = 0;
    for (int base = 0; base < total; base += 32) {
        const unsigned rel = (unsigned)(off - base);
        const unsigned starts = __reduce_or_sync(FULL, (has && rel < 32u) ? (1u << rel) : 0u);
        const int t = base + lane;
        if (t < total) {
            const int j = sc.ne[before + __popc(starts & (0xFFFFFFFFu >> (31 - lane))) - 1];
            const int4 rc = sc.rec[j];
            const int ri = t + ((t < rc.x) ? rc.y : rc.z);
            double h;
            if (ray_test(raytab[ri], sc.geo[j], h)) {
                if (t < PAIR_SLOTS - 1) {
                    const uint32_t v = (uint32_t)rc.w | (uint32_t)t;
                    const uint32_t old = atomicMin(&sc.best[ri], v);
                    ambiguous = ambiguous || ((old >> 8) == (v >> 8));  // same ray, equal key bits: order undecided
                    sc.hpair[t] = h;
                } else {
                    ambiguous = true;  // out of pair slots
                }
            }
        }
        before += __popc(starts);
    }
    ambiguous = __any_sync(FULL, ambiguous);
    __syncwarp();

    // ---- phase C: lane = ray ---------------------------------------------------------------------
    double hmin = SENSOR_MAX_RANGE;
    if (!ambiguous) {
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            out[p] = sc.hpair[sc.best[lane + 32 * p] & 0xFFu];  // NO_HIT -> slot 255 -> SENSOR_MAX_RANGE
            hmin = (out[p] < hmin) ? out[p] : hmin;
        }
    } else {
        // exact path: every obstacle against this lane's rays, float64 keys, strict '<' in ascending slot order
        // (the lower slot wins a tie, as the reference's stable argsort does)
        for (int p = 0; p < 4; ++p) {
            const double2 tb = raytab[lane + 32 * p];
            double bkey = INFINITY, bh = SENSOR_MAX_RANGE;
            for (int j = 0; j < nob; ++j) {
                const double4 g = sc.geo[j];
                double h;
                if (ray_test(tb, g, h) && g.w < bkey) {
                    bkey = g.w;
                    bh = h;
                }
            }
            out[p] = bh;
            hmin = (bh < hmin) ? bh : hmin;
        }
    }
    const bool hit_lt = hmin < hit_thr;  // readings of rays without a hit are SENSOR_MAX_RANGE
    min_hit_lt = __any_sync(FULL, hit_lt);
    __syncwarp();
}

}  // namespace usv
