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
//            (obstacle, ray) candidate pairs (~100 per env instead of 29 x 128);
//   phase B  lane = candidate pair, 32 per pass, all lanes busy: the reference's float64 hit test; a hit
//            stores its reading in the pair's slot and enters the ray's winner word with a shared-memory
//            atomicMin on (order-preserving float32 key bits, top 24 | pair slot, low 8);
//   phase C  lane = ray (4 passes of 32): winner word -> pair slot -> reading (slot 255 holds "max range").
// All hit / miss decisions are float64.  The winner order is decided on the top 24 bits of the float32-
// rounded keys, which is exact unless two obstacles that hit the SAME ray have keys equal to within 2^-16
// relative: phase B detects exactly that (the atomicMin returns a word with the same key bits) and the
// env then re-does its rays on the exact path (float64 keys, ascending slot order breaks ties like the
// reference's stable argsort); so does an env with more than 255 candidate pairs (boat inside circles).
// In the body frame the ray directions are the constant table (cos phi_i, sin phi_i), so
// along = cos phi bx + sin phi by and perp = sin phi bx - cos phi by equal the reference's expressions
// (usv_asmc_ca_env.py:516-518) with cos / sin (psi + phi_i) expanded.
//
// Two output forms (template HT):
//   HT = double  raw readings in metres (the unit-parity entry point usvgpu_sensors);
//   HT = float   what the observation row stores, float32(reading * 0.01) (simple_env.py:83), rounded ONCE from the
//                float64 reading in phase B -- phase C is then four 4-byte look-ups per lane.  The collision test
//                "min reading < 0.2" (simple_env.py:153-156) is decided on those float32 values: float32 rounding
//                is monotone, so v < float32(0.002) implies reading < 0.2 and v > float32(0.002) implies
//                reading > 0.2; only v == float32(0.002) (the reading within 2.3e-8 m of 0.2) is undecided and
//                sends the env to the exact float64 path.
#pragma once
#include "usv_math.cuh"
#include "usv_state.cuh"

namespace usv {

#ifndef USV_SENS_ASINSEL
#define USV_SENS_ASINSEL 1
#endif
constexpr int PAIR_SLOTS = 256;           // slot 255 is reserved: "no hit" -> SENSOR_MAX_RANGE
constexpr uint32_t NO_HIT = 0xFFFFFFFFu;

template <typename HT>
struct SensorScratch {  // per warp: 3.0 KB (HT = float) / 4.0 KB (double)
    // entry k of geo / rec belongs to the k-th obstacle THAT HAS CANDIDATE RAYS (table order kept): phase B finds the
    // owner of a pair as a rank and reads both with it, without a slot look-up in between
    double4 geo[MAX_OBST];      // bx, by (offset in the body frame), radius^2, key
    int4 rec[MAX_OBST];         // split, rbase, rbase2, key bits: pair t of this obstacle is ray t + (t < split ? rbase : rbase2)
    HT hpair[PAIR_SLOTS];       // reading of candidate pair t (valid where the pair hit)
    uint32_t best[N_SENS];      // min over hitting obstacles of (key bits & ~255 | pair slot)
};

template <typename HT> struct Reading;
template <> struct Reading<double> {
    static __device__ __forceinline__ double of(double h) { return h; }
    static __device__ __forceinline__ double no_hit() { return SENSOR_MAX_RANGE; }
};
template <> struct Reading<float> {
    static __device__ __forceinline__ float of(double h) { return (float)(h * 0.01); }  // SENSOR_MAX_RANGE * 0.01 rounds to 1.0f
    static __device__ __forceinline__ float no_hit() { return 1.0f; }
};

// 1/x as one MUFU.RCP (__fdividef wraps the same instruction in range checks that cost six more)
__device__ __forceinline__ float rcp_fast(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
// atan2(y, x) to 2.1e-4 rad (7th-order odd polynomial on the folded octant); (x, y) != (0, 0)
__device__ __forceinline__ float atan2_coarse(float y, float x) {
    const float ax = fabsf(x), ay = fabsf(y);
    const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
    const float a = mn * rcp_fast(fmaxf(mx, 1e-30f));
    const float s = a * a;
    float r = fmaf(fmaf(fmaf(-0.0464964749f, s, 0.15931422f), s, -0.327622764f), s * a, a);
    r = (ay > ax) ? 1.57079637f - r : r;
    r = (x < 0.0f) ? 3.14159274f - r : r;
    return copysignf(r, y);
}
// asin(q) for 0 <= q <= 1, never more than 7.5e-5 rad below the true value
__device__ __forceinline__ float asin_coarse(float q) {
#if USV_SENS_ASINSEL
    // no branch: asin is convex on [0, 1], so above 0.5 the chord through (0.5, pi / 6) and (1, pi / 2) is an upper bound
    // (at most 0.09 rad = 3 rays too wide, and only for a circle whose centre is closer than its diameter)
    const float s = q * q;
    const float p = q * fmaf(s, fmaf(s, fmaf(s, 15.0f / 336, 3.0f / 40), 1.0f / 6), 1.0f);
    const float c = fmaf(fminf(q, 1.0f), 2.0943952f, -0.52359877f);  // pi/6 + (q - 1/2) (2 pi / 3)
    return (q > 0.5f) ? c : p;
#else
    if (q > 0.5f) return asinf(fminf(q, 1.0f));
    const float s = q * q;
    return q * fmaf(s, fmaf(s, fmaf(s, 15.0f / 336, 3.0f / 40), 1.0f / 6), 1.0f);
#endif
}

// the reference's test of a ray (direction t in the body frame) against an obstacle at body-frame offset (bx, by) with
// squared radius rr2: hit iff along >= 0, delta >= 0 and the reading is inside the range (usv_asmc_ca_env.py:448-458)
__device__ __forceinline__ bool ray_test(const double2 t, const double bx, const double by, const double rr2, double& h) {
    const double along = fma(t.x, bx, t.y * by);      // :516
    const double perp = fma(t.y, bx, -(t.x * by));    // :518
    const double delta = fma(-perp, perp, rr2);       // :453
    if (!(along >= 0.0 && delta >= 0.0)) return false;  // :448, :454
    h = along - sqrt_k(delta);                          // :457
    return h < SENSOR_MAX_RANGE;                        // :458
}

// Entry k of the ray table (cos phi_k, sin phi_k) in shared memory, addressed by the table's 32-bit shared-window address:
// the caller computes that address once per kernel and keeps it in a register (opaque_u32), where the compiler would
// otherwise rebuild it from the CTA's window base at every use (S2UR + UMOV + ULEA per pass of phase B, r02 final SASS).
__device__ __forceinline__ double2 ray_of(const uint32_t ray_addr, const int k) {
    double2 t;
    asm("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(t.x), "=d"(t.y) : "r"(ray_addr + 16u * (uint32_t)k));
    return t;
}
__device__ __forceinline__ uint32_t opaque_u32(uint32_t v) {  // a value the compiler keeps instead of rematerialising
    asm volatile("mov.u32 %0, %0;" : "+r"(v));
    return v;
}

// once per warp, before its first env: the "no hit" slot
template <typename HT>
__device__ __forceinline__ void sensor_scratch_init(SensorScratch<HT>& sc, const int lane) {
    if (lane == 0) sc.hpair[PAIR_SLOTS - 1] = Reading<HT>::no_hit();
}

// ox, oy, rr: this lane's obstacle (ignored when lane >= nob).  cps/sps = cos/sin(psi).  ray_addr: see ray_of.
// The caller runs sensor_scratch_init(sc, lane) once.
// out[p]: reading of ray (lane + 32 p) in the form of Reading<HT>.
// min_key_lt: warp-uniform "min over obstacles of key < key_thr" (simple_env.py:334); min_hit_lt: warp-uniform "min over
// rays of reading < 0.2" (simple_env.py:153-156).
template <typename HT>
__device__ __forceinline__ void sensor_warp(SensorScratch<HT>& sc, const uint32_t ray_addr, const int lane,
                                            const int nob, const double x, const double y, const double cps,
                                            const double sps, const double ox, const double oy, const double rr,
                                            HT out[4], const double key_thr, bool& min_key_lt, bool& min_hit_lt) {
    constexpr unsigned FULL = 0xFFFFFFFFu;
    constexpr double HIT_THR = 0.2;
    reinterpret_cast<uint4*>(sc.best)[lane] = make_uint4(NO_HIT, NO_HIT, NO_HIT, NO_HIT);

    // ---- phase A: lane = obstacle ---------------------------------------------------------------
    bool key_lt = false;
    int lo1 = 0, n1 = 0, lo2 = 0, n2 = 0;
    uint32_t kbits = 0u;
    double bx = 0.0, by = 0.0, key = 0.0;
    if (lane < nob) {
        const double dx = ox - x, dy = oy - y;  // usv_asmc_ca_env.py:508
        const double d = sqrt_k(fma(dx, dx, dy * dy));
        key = d - rr;                           // simple_env.py:205-206
        key_lt = key < key_thr;
        bx = fma(cps, dx, sps * dy);            // offset in the body frame
        by = fma(cps, dy, -(sps * dx));
        const uint32_t kb = __float_as_uint((float)key);  // float32 rounding is monotone: it can tie, never invert
        kbits = ((kb & 0x80000000u) ? ~kb : (kb | 0x80000000u)) & ~255u;
        if (d <= rr) {  // boat inside the circle (only on a terminal step): every ray may hit
            n1 = N_SENS;
        } else {
            const float cf = (atan2_coarse((float)by, (float)bx) + 2.0943951f) * INV_PITCH_F;  // ray index of the bearing, [-32, 160]
            const float w = asin_coarse((float)rr * rcp_fast((float)d)) * INV_PITCH_F + 0.05f;  // approximations: < 0.01 ray
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
    if (has) {  // an obstacle whose (conservative) wedge misses the fan cannot be hit by any ray: it is not stored
        const int k = __popc(has_bal & ((1u << lane) - 1u));
        sc.geo[k] = make_double4(bx, by, rr * rr, key);
        sc.rec[k] = make_int4(off + n1, lo1 - off, lo2 - off - n1, (int)kbits);
    }
    const int n_cand = __popc(has_bal);
    const int total = __shfl_sync(FULL, incl, 31);
    __syncwarp();

    // ---- phase B: lane = candidate pair ------------------------------------------------------------
    // Pair t belongs to the last obstacle (among those with candidates) whose offset is <= t.  Per pass of 32
    // pairs: bit k of `starts` marks an obstacle whose list starts at pair base + k, `before` counts those that
    // started earlier, so the owner of pair base + q is entry before + popc(starts & bits 0..q) - 1.
    bool ambiguous = total > PAIR_SLOTS - 1;  // out of pair slots (boat inside circles): exact path
    uint32_t tie = 0xFFFFFFFFu;               // min over this lane's hits of (previous winner word ^ own word)
    if (!ambiguous) {
        const unsigned le_mask = 0xFFFFFFFFu >> (31 - lane);
        int before = -1;
        for (int base = 0; base < total; base += 32) {
            const unsigned rel = (unsigned)(off - base);
            const unsigned starts = __reduce_or_sync(FULL, (has && rel < 32u) ? (1u << rel) : 0u);
            const int t = base + lane;
            if (t < total) {
                const int j = before + __popc(starts & le_mask);
                const int4 rc = sc.rec[j];
                const int ri = t + ((t < rc.x) ? rc.y : rc.z);
                const double2 bxy = *reinterpret_cast<const double2*>(&sc.geo[j]);
                const double rr2 = sc.geo[j].z;
                double h;
                if (ray_test(ray_of(ray_addr, ri), bxy.x, bxy.y, rr2, h)) {
                    const uint32_t v = (uint32_t)rc.w | (uint32_t)t;
                    const uint32_t old = atomicMin(&sc.best[ri], v);
                    tie = min(tie, old ^ v);  // < 256: same ray, equal key bits -- order undecided
                    sc.hpair[t] = Reading<HT>::of(h);
                }
            }
            before += __popc(starts);
        }
        ambiguous = tie < 256u;
    }
    __syncwarp();

    // ---- phase C: lane = ray ---------------------------------------------------------------------
    bool hit_lt = false;
    if (!__any_sync(FULL, ambiguous)) {
        HT vmin = Reading<HT>::no_hit();
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            out[p] = sc.hpair[sc.best[lane + 32 * p] & 0xFFu];  // NO_HIT -> slot 255 -> no_hit()
            vmin = (out[p] < vmin) ? out[p] : vmin;
        }
        const HT thr = Reading<HT>::of(HIT_THR);
        hit_lt = vmin < thr;
        // float32 form: a minimum that rounds onto the threshold itself is undecided (header of this file)
        ambiguous = (sizeof(HT) == sizeof(float)) && (vmin == thr);
    }
    if (__any_sync(FULL, ambiguous)) {
        // exact path: every obstacle against this lane's rays, float64 keys, strict '<' in ascending slot order
        // (the lower slot wins a tie, as the reference's stable argsort does)
        double hmin = SENSOR_MAX_RANGE;
        for (int p = 0; p < 4; ++p) {
            const double2 tb = ray_of(ray_addr, lane + 32 * p);
            double bkey = INFINITY, bh = SENSOR_MAX_RANGE;
            for (int j = 0; j < n_cand; ++j) {
                const double4 g = sc.geo[j];
                double h;
                if (ray_test(tb, g.x, g.y, g.z, h) && g.w < bkey) {
                    bkey = g.w;
                    bh = h;
                }
            }
            out[p] = Reading<HT>::of(bh);
            hmin = (bh < hmin) ? bh : hmin;
        }
        hit_lt = hmin < HIT_THR;  // readings of rays without a hit are SENSOR_MAX_RANGE
    }
    min_hit_lt = __any_sync(FULL, hit_lt);
    __syncwarp();
}

}  // namespace usv
