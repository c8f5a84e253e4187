// usv_render.cuh -- headless rasteriser for the `rgb_array` frames of UsvSimpleEnv / UsvSimpleASMCEnv
// (SURVEY.md 8f rank 4): the draw list of SimpleEnvVisualizer.render_frame
// (gym_usv/envs/simple_env_visualizer.py:34-113, called from simple_env.py:117-131) written into a device buffer
// uint8 [K][size][size][3] (row = canvas y, column = canvas x: the np.transpose of :112-114).
//
// The reference draws with pygame's software renderer, which is absent here (and there is no graphics stack on
// either box): windows / blits are left out, and the exact pixel coverage of pygame.draw cannot be pinned.  The
// coverage rules are therefore stated here and restated independently by the test-side CPU checker (under oracle/)
// (painter's loop + incremental Bresenham there, per-pixel closed forms here); "parity unpinned" against pygame.
//
// Draw list, in painter's order (a later primitive overwrites an earlier one), scale = size / env_bounds[1]:
//   0        target      filled circle (0,0,255)   centre target*scale           radius 10      :40-45
//   1..128   sensor rays line (0,255,0) width 1    (x,y)*scale -> (x + d cos a, y + d sin a)*scale, a = ray angle  :48-61
//   129      agent       filled circle (255,0,0)   centre (x,y)*scale            radius 10      :64-69
//   130      "front"     filled circle (100,100,0) centre (0.1 (cos psi, sin psi) + (x,y))*scale  radius 8  :72-79
//   131..    obstacles   filled circle (0,100,0)   centre pos*scale              radius r*scale :82-91
//   last     path        line (100,0,0) width 5    path_start*scale -> path_end*scale          :94-100
// Coverage rules:
//   * every coordinate / radius is converted to an integer by truncation toward zero (C cast), clamped to +-2^20;
//   * filled circle of integer centre (cx, cy) and radius r >= 1: pixel (px, py) is covered when its centre lies in
//     the disc, (2 px + 1 - 2 cx)^2 + (2 py + 1 - 2 cy)^2 <= 4 r^2 (2 r pixels wide, like pygame's);
//   * line of width 1: integer Bresenham from the first to the second end point, ties rounded up: along the major
//     axis step k = 0..b the minor offset is floor((2 a k + b) / (2 b)), a <= b the absolute deltas;
//   * line of width w: w Bresenham lines shifted by -(w-1)/2 .. w/2 along the minor axis.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "usv_state.cuh"

namespace usv {

constexpr int RND_TILE = 32;           // a CTA paints one 32 x 32 pixel tile of one frame
constexpr int RND_THREADS = 256;       // 4 pixels per thread
constexpr int RND_MAX_PRIMS = 1 + N_SENS + 2 + MAX_OBST + 1;  // 164
constexpr int RND_COORD_MAX = 1 << 20;

struct RenderPrim {  // circle: (x0, y0) centre, x1 = radius, y1 = -1;  line: end points, w = width (y1 >= 0 or any)
    int x0, y0, x1, y1;
    int kind;        // 0 circle, 1 line
    int width;
    unsigned rgb;    // r | g << 8 | b << 16
    int pad;
};

// What a pixel evaluates: every test is a pair of linear forms (lines) or one quadratic (circles) over the pixel
// indices, with integer coefficients below 2^45 held exactly in float64 -- a handful of DFMAs per pixel and primitive
// instead of 64-bit integer multiplies.
//   line  : K = d[0] px + d[1] py + d[2] (step along the major axis), covered when 0 <= K <= d[3] and
//           d[7] <= F < d[8] with F = d[4] px + d[5] py + d[6] = 2 a K + b -+ 2 b (minor offset): the division-free
//           form of "minor offset - floor((2 a K + b) / (2 b)) lies in the width band"
//   circle: (2 px - d[0])^2 + (2 py - d[1])^2 <= d[2]
struct PixelPrim {
    double d[9];
    int kind;
    unsigned rgb;
};

__device__ __forceinline__ PixelPrim rnd_pixel_form(const RenderPrim& p) {
    PixelPrim q;
    q.kind = p.kind;
    q.rgb = p.rgb;
#pragma unroll
    for (int c = 0; c < 9; ++c) q.d[c] = 0.0;
    if (p.kind == 0) {
        q.d[0] = 2.0 * p.x0 - 1.0;
        q.d[1] = 2.0 * p.y0 - 1.0;
        q.d[2] = 4.0 * (double)p.x1 * (double)p.x1;
        return q;
    }
    const int dx = p.x1 - p.x0, dy = p.y1 - p.y0;
    const int adx = dx < 0 ? -dx : dx, ady = dy < 0 ? -dy : dy;
    const double lo = -((p.width - 1) / 2), hi = p.width / 2;
    const bool xmaj = adx >= ady;
    const double a = xmaj ? ady : adx, b = xmaj ? adx : ady;
    const double su = ((xmaj ? dx : dy) < 0) ? -1.0 : 1.0;   // direction of the walk along the major axis
    const double sm = ((xmaj ? dy : dx) < 0) ? -1.0 : 1.0;   // ... and of the minor steps
    const double maj0 = xmaj ? p.x0 : p.y0, min0 = xmaj ? p.y0 : p.x0;
    // K = su (major - maj0)
    q.d[xmaj ? 0 : 1] = su;
    q.d[2] = -su * maj0;
    q.d[3] = b;
    if (b == 0.0) {  // a single point: F = minor - min0 in [lo, hi]
        q.d[xmaj ? 5 : 4] = 1.0;
        q.d[6] = -min0;
        q.d[7] = lo;
        q.d[8] = hi + 1.0;
        return q;
    }
    // F = 2 a K + b - sm 2 b (minor - min0)
    q.d[xmaj ? 4 : 5] = 2.0 * a * su;
    q.d[xmaj ? 5 : 4] = -sm * 2.0 * b;
    q.d[6] = -2.0 * a * su * maj0 + b + sm * 2.0 * b * min0;
    q.d[7] = sm > 0 ? -2.0 * b * hi : 2.0 * b * lo;
    q.d[8] = sm > 0 ? 2.0 * b * (1.0 - lo) : 2.0 * b * (hi + 1.0);
    return q;
}

__device__ __forceinline__ bool rnd_covers(const PixelPrim& q, double px, double py) {
    if (q.kind == 0) {
        const double ddx = 2.0 * px - q.d[0], ddy = 2.0 * py - q.d[1];
        return ddx * ddx + ddy * ddy <= q.d[2];
    }
    const double k = q.d[0] * px + q.d[1] * py + q.d[2];
    const double f = q.d[4] * px + q.d[5] * py + q.d[6];
    return k >= 0.0 && k <= q.d[3] && f >= q.d[7] && f < q.d[8];
}

__host__ __device__ __forceinline__ int rnd_trunc(double v) {
    if (!(v == v)) return 0;
    if (v > (double)RND_COORD_MAX) return RND_COORD_MAX;
    if (v < -(double)RND_COORD_MAX) return -RND_COORD_MAX;
    return (int)v;
}

// bounding box of a primitive against the tile [tx0, tx0 + 32) x [ty0, ty0 + 32)
__device__ __forceinline__ bool rnd_touches_tile(const RenderPrim& p, int tx0, int ty0) {
    int x_lo, x_hi, y_lo, y_hi;
    if (p.kind == 0) {
        x_lo = p.x0 - p.x1; x_hi = p.x0 + p.x1; y_lo = p.y0 - p.x1; y_hi = p.y0 + p.x1;
    } else {
        const int e = p.width;  // generous: the shift is along one axis only
        x_lo = min(p.x0, p.x1) - e; x_hi = max(p.x0, p.x1) + e;
        y_lo = min(p.y0, p.y1) - e; y_hi = max(p.y0, p.y1) + e;
    }
    if (x_hi < tx0 || x_lo >= tx0 + RND_TILE || y_hi < ty0 || y_lo >= ty0 + RND_TILE) return false;
    if (p.kind == 1) {  // a long thin line crosses few of the tiles of its bounding box: test the tile's corners
        const long long dx = p.x1 - p.x0, dy = p.y1 - p.y0;
        const long long e = (long long)(p.width + 2) * (llabs(dx) + llabs(dy));  // |cross| margin for the line's width
        long long cmin = 0, cmax = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const long long qx = tx0 + ((c & 1) ? RND_TILE : 0) - p.x0, qy = ty0 + ((c & 2) ? RND_TILE : 0) - p.y0;
            const long long cr = dx * qy - dy * qx;
            cmin = c ? min(cmin, cr) : cr;
            cmax = c ? max(cmax, cr) : cr;
        }
        if (cmin > e || cmax < -e) return false;  // all four corners strictly on one side, beyond the width
    }
    return true;
}

// grid (tiles_y * ceil(tiles_x / tpc), K): a CTA builds the frame's draw list once and paints tpc consecutive tiles of
// one tile row (the host picks tpc: whole rows when there are many frames, single tiles when there are few, so that
// the grid still fills the machine); frames [K][size][size][3] uint8
__global__ void __launch_bounds__(RND_THREADS)
render_kernel(StatePtrs s, const long long* __restrict__ index, const float* __restrict__ obs, int obs_stride,
              int obstacles_f32, int size, int tpc, uint8_t* __restrict__ frames) {
    __shared__ PixelPrim prims[RND_MAX_PRIMS];
    __shared__ unsigned binmask[(RND_MAX_PRIMS + 31) / 32];  // bit id set: primitive id touches this tile
    __shared__ unsigned tile[RND_TILE][RND_TILE * 3 / 4];  // 32 rows x 96 bytes
    const int tid = threadIdx.x;
    const int tiles_x = (size + RND_TILE - 1) / RND_TILE, groups = (tiles_x + tpc - 1) / tpc;
    const int ty0 = (blockIdx.x / groups) * RND_TILE;
    const int tx_begin = (blockIdx.x % groups) * tpc * RND_TILE, tx_end = min(size, tx_begin + tpc * RND_TILE);
    const long long f = blockIdx.y;
    const long long i = index ? index[f] : f;
    const long long n = s.n;
    if (i < 0 || i >= n) return;  // an index outside the shard: its frame is left untouched (CTA-uniform exit)
    const double scale = (double)size / ENV_HI;  // simple_env_visualizer.py:37
    const double x = s.dyn[(long long)D_X * n + i], y = s.dyn[(long long)D_Y * n + i], psi = s.dyn[(long long)D_PSI * n + i];
    const double psx = s.epi[(long long)E_PSX * n + i], psy = s.epi[(long long)E_PSY * n + i];
    const double pex = s.epi[(long long)E_PEX * n + i], pey = s.epi[(long long)E_PEY * n + i];
    const int nob = *nobs_of(s, i);
    // ---- the draw list, one primitive per thread -------------------------------------------------------------
    RenderPrim p = {0, 0, 0, -1, 0, 1, 0u, 0};
    bool valid = tid < RND_MAX_PRIMS;
    if (tid < RND_MAX_PRIMS) {
        if (tid == 0) {  // target: simple_env.py:139-148 keeps it at path_start + progress (path_end - path_start);
                         // a freshly reset env still has the random target of :245 (kept by the reset in the row's pad)
            double tx, ty;
            if (s.steps[i] == 0) {
                const float* t0 = target0_of(s, i);
                tx = (double)t0[0];
                ty = (double)t0[1];
            } else {
                const double a = s.dyn[(long long)D_PROGRESS * n + i];
                tx = __dadd_rn(psx, __dmul_rn(a, pex - psx));  // no FMA: the value is truncated to a pixel
                ty = __dadd_rn(psy, __dmul_rn(a, pey - psy));
            }
            p.x0 = rnd_trunc(tx * scale); p.y0 = rnd_trunc(ty * scale); p.x1 = 10; p.rgb = 255u << 16;
        } else if (tid <= N_SENS) {  // ray k: angle -2 pi / 3 + k pitch + psi (usv_asmc_ca_env.py:420-423)
            const int k = tid - 1;
            const double d = (double)obs[i * obs_stride + HDR + k] * SENSOR_MAX_RANGE;
            const double ang = __dadd_rn(__dadd_rn(-PI * 2 / 3, __dmul_rn((double)k, SENSOR_PITCH)), psi);
            double sn, cs;
            sincos(ang, &sn, &cs);
            p.kind = 1;
            p.x0 = rnd_trunc(x * scale); p.y0 = rnd_trunc(y * scale);
            p.x1 = rnd_trunc(__dadd_rn(__dmul_rn(d, cs), x) * scale); p.y1 = rnd_trunc(__dadd_rn(__dmul_rn(d, sn), y) * scale);
            p.rgb = 255u << 8;
        } else if (tid == N_SENS + 1) {
            p.x0 = rnd_trunc(x * scale); p.y0 = rnd_trunc(y * scale); p.x1 = 10; p.rgb = 255u;
        } else if (tid == N_SENS + 2) {
            double sn, cs;
            sincos(psi, &sn, &cs);
            p.x0 = rnd_trunc(__dadd_rn(__dmul_rn(cs, 0.1), x) * scale); p.y0 = rnd_trunc(__dadd_rn(__dmul_rn(sn, 0.1), y) * scale); p.x1 = 8;
            p.rgb = 100u | (100u << 8);
        } else if (tid < N_SENS + 3 + MAX_OBST) {
            const int j = tid - (N_SENS + 3);
            valid = j < nob;
            if (valid) {
                double ox, oy, orad;
                if (obstacles_f32) {
                    const float* row = obst_row<float>(s, i);
                    ox = row[j]; oy = row[MAX_OBST + j]; orad = row[2 * MAX_OBST + j];
                } else {
                    const double* row = obst_row<double>(s, i);
                    ox = row[j]; oy = row[MAX_OBST + j]; orad = row[2 * MAX_OBST + j];
                }
                p.x0 = rnd_trunc(ox * scale); p.y0 = rnd_trunc(oy * scale); p.x1 = rnd_trunc(orad * scale);
                p.rgb = 100u << 8;
                valid = p.x1 >= 1;
            }
        } else {
            p.kind = 1; p.width = 5;
            p.x0 = rnd_trunc(psx * scale); p.y0 = rnd_trunc(psy * scale);
            p.x1 = rnd_trunc(pex * scale); p.y1 = rnd_trunc(pey * scale);
            p.rgb = 100u;
        }
        prims[tid] = rnd_pixel_form(p);
    }
    uint8_t* frame = frames + f * (long long)size * size * 3;
  for (int tx0 = tx_begin; tx0 < tx_end; tx0 += RND_TILE) {
    // ---- which primitives touch this tile ---------------------------------------------------------------------
    __syncthreads();  // the previous tile's mask and staged pixels have been consumed (first round: prims are written)
    if (tid < (RND_MAX_PRIMS + 31) / 32) binmask[tid] = 0u;
    __syncthreads();
    if (valid && rnd_touches_tile(p, tx0, ty0)) atomicOr(&binmask[tid >> 5], 1u << (tid & 31));
    __syncthreads();
    // ---- per pixel: the covering primitive drawn last wins ---------------------------------------------------
    // walk the tile's primitives from the last drawn to the first: the first one that covers the pixel is its colour
    unsigned bm[(RND_MAX_PRIMS + 31) / 32];
#pragma unroll
    for (int w = 0; w < (RND_MAX_PRIMS + 31) / 32; ++w) bm[w] = binmask[w];
    uint8_t* tb = reinterpret_cast<uint8_t*>(&tile[0][0]);
    for (int q = tid; q < RND_TILE * RND_TILE; q += RND_THREADS) {
        const int lx = q % RND_TILE, ly = q / RND_TILE;
        const int px = tx0 + lx, py = ty0 + ly;
        int best = -1;
#pragma unroll
        for (int w = (RND_MAX_PRIMS + 31) / 32 - 1; w >= 0 && best < 0; --w) {
            unsigned m = bm[w];
            while (m) {
                const int bit = 31 - __clz(m);
                m &= ~(1u << bit);
                if (rnd_covers(prims[w * 32 + bit], (double)px, (double)py)) {
                    best = w * 32 + bit;
                    break;
                }
            }
        }
        const unsigned rgb = best < 0 ? 0xFFFFFFu : prims[best].rgb;  // canvas.fill((255, 255, 255)) :35
        uint8_t* o = tb + (ly * RND_TILE + lx) * 3;
        o[0] = (uint8_t)(rgb & 255u);
        o[1] = (uint8_t)((rgb >> 8) & 255u);
        o[2] = (uint8_t)((rgb >> 16) & 255u);
    }
    __syncthreads();
    // ---- tile rows out: 96 bytes = 24 words per row, 4-byte aligned whenever the frame row is -------------------
    const bool words = (size % 4) == 0 && tx0 + RND_TILE <= size;
    if (words) {
        for (int q = tid; q < RND_TILE * 24; q += RND_THREADS) {
            const int ly = q / 24, w = q % 24;
            if (ty0 + ly < size)
                reinterpret_cast<unsigned*>(frame + ((long long)(ty0 + ly) * size + tx0) * 3)[w] = tile[ly][w];
        }
    } else {
        for (int q = tid; q < RND_TILE * RND_TILE * 3; q += RND_THREADS) {
            const int ly = q / (RND_TILE * 3), bx = q % (RND_TILE * 3);
            if (ty0 + ly < size && tx0 + bx / 3 < size) frame[((long long)(ty0 + ly) * size + tx0) * 3 + bx] = tb[q];
        }
    }
  }
}

}  // namespace usv
