/*
 * usv_oracle_render.c -- CPU restatement of SimpleEnvVisualizer.render_frame
 * (gym_usv/envs/simple_env_visualizer.py:17-114, rgb_array mode) for the checker of usvgpu_render.
 *
 * TEST INFRASTRUCTURE ONLY (see usv_oracle.h).
 *
 * PARITY UNPINNED against the reference's pixels: the reference draws with pygame's software renderer, pygame is
 * absent from this image (and there is no graphics stack on either box), so no golden frame of the reference can be
 * produced.  What IS taken from the reference: the draw list, its order, colours, radii, widths, the scale and the
 * (H, W, 3) layout (:34-114).  The pixel-coverage rules are this repo's own statement (gym_usv_b200/csrc/usv_render.cuh
 * header): truncation of every coordinate toward zero, disc test on pixel centres, integer Bresenham with ties
 * rounded up, wide lines as parallel Bresenham lines shifted along the minor axis.
 *
 * Structure on purpose unlike the kernel's: a painter's loop over the primitives in draw order, circles by scanning
 * their bounding box, lines by the incremental error-accumulating Bresenham walk, clipped per pixel.
 */
#include "usv_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

#define RND_COORD_MAX (1 << 20)

static int r_trunc(double v) {
    if (!(v == v)) return 0;
    if (v > (double)RND_COORD_MAX) return RND_COORD_MAX;
    if (v < -(double)RND_COORD_MAX) return -RND_COORD_MAX;
    return (int)v;
}

static void put(uint8_t* f, int size, int px, int py, const uint8_t rgb[3]) {
    if (px < 0 || py < 0 || px >= size || py >= size) return;
    memcpy(f + ((size_t)py * size + px) * 3, rgb, 3);
}

/* pygame.draw.circle(canvas, color, center, radius) -- filled */
static void fill_circle(uint8_t* f, int size, int cx, int cy, int r, const uint8_t rgb[3]) {
    if (r < 1) return;
    int y_lo = cy - r < 0 ? 0 : cy - r, y_hi = cy + r >= size ? size - 1 : cy + r;
    int x_lo = cx - r < 0 ? 0 : cx - r, x_hi = cx + r >= size ? size - 1 : cx + r;
    for (int py = y_lo; py <= y_hi; ++py)
        for (int px = x_lo; px <= x_hi; ++px) {
            /* pixel centre (px + 1/2, py + 1/2) inside the disc of radius r around (cx, cy) */
            double ddx = px + 0.5 - cx, ddy = py + 0.5 - cy;
            if (ddx * ddx + ddy * ddy <= (double)r * r) put(f, size, px, py, rgb);
        }
}

/* pygame.draw.line(canvas, color, start, end, width) */
static void draw_line(uint8_t* f, int size, int x0, int y0, int x1, int y1, int width, const uint8_t rgb[3]) {
    const int dx = x1 - x0, dy = y1 - y0;
    const long long a_x = llabs((long long)dx), a_y = llabs((long long)dy);
    const int sx = dx < 0 ? -1 : 1, sy = dy < 0 ? -1 : 1;
    const int lo = -((width - 1) / 2), hi = width / 2;
    if (a_x >= a_y) {
        long long err = a_x; /* accumulates 2 a_y per step; a minor step is due when it reaches 2 a_x */
        int x = x0, y = y0;
        for (long long k = 0; k <= a_x; ++k) {
            /* only the part of the walk that can touch the canvas is painted (the walk itself is never clipped) */
            if (x >= 0 && x < size)
                for (int o = lo; o <= hi; ++o) put(f, size, x, y + o, rgb);
            x += sx;
            err += 2 * a_y;
            if (a_x && err >= 2 * a_x) {
                y += sy;
                err -= 2 * a_x;
            }
        }
    } else {
        long long err = a_y;
        int x = x0, y = y0;
        for (long long k = 0; k <= a_y; ++k) {
            if (y >= 0 && y < size)
                for (int o = lo; o <= hi; ++o) put(f, size, x + o, y, rgb);
            y += sy;
            err += 2 * a_x;
            if (err >= 2 * a_y) {
                x += sx;
                err -= 2 * a_y;
            }
        }
    }
}

/* simple_env_visualizer.py:17-114.  position x y psi; target_position; sensor_distance[sensor_count] in metres
 * (sensor_data[:, 1]); ray k points at -2 pi / 3 + k * pitch + psi (usv_asmc_ca_env.py:420-423, sensor_data[:, 0]);
 * frame uint8 [size][size][3], row = canvas y. */
void usv_oracle_render_frame(const double position[3], const double target[2], const double* sensor_distance,
                             int sensor_count, int n_obs, const double* obs_x, const double* obs_y,
                             const double* obs_r, const double path_start[2], const double path_end[2],
                             double env_hi, int size, uint8_t* frame) {
    static const uint8_t BLUE[3] = {0, 0, 255}, GREEN[3] = {0, 255, 0}, RED[3] = {255, 0, 0}, FRONT[3] = {100, 100, 0},
                         OBST[3] = {0, 100, 0}, PATH[3] = {100, 0, 0};
    memset(frame, 255, (size_t)size * size * 3);          /* :35 canvas.fill((255, 255, 255)) */
    const double scale = (double)size / env_hi;            /* :37 */
    const double x = position[0], y = position[1], psi = position[2];
    fill_circle(frame, size, r_trunc(target[0] * scale), r_trunc(target[1] * scale), 10, BLUE); /* :40-45 */
    const double pitch = ((2.0 / 3) * (2 * M_PI)) / sensor_count;
    for (int k = 0; k < sensor_count; ++k) {               /* :48-61 */
        const double angle = (-M_PI * 2 / 3 + k * pitch) + psi;
        const double x_f = sensor_distance[k] * cos(angle) + x, y_f = sensor_distance[k] * sin(angle) + y;
        draw_line(frame, size, r_trunc(x * scale), r_trunc(y * scale), r_trunc(x_f * scale), r_trunc(y_f * scale), 1, GREEN);
    }
    fill_circle(frame, size, r_trunc(x * scale), r_trunc(y * scale), 10, RED); /* :64-69 */
    fill_circle(frame, size, r_trunc((cos(psi) * 0.1 + x) * scale), r_trunc((sin(psi) * 0.1 + y) * scale), 8, FRONT); /* :72-79 */
    for (int j = 0; j < n_obs; ++j)                        /* :82-91 */
        fill_circle(frame, size, r_trunc(obs_x[j] * scale), r_trunc(obs_y[j] * scale), r_trunc(obs_r[j] * scale), OBST);
    draw_line(frame, size, r_trunc(path_start[0] * scale), r_trunc(path_start[1] * scale), r_trunc(path_end[0] * scale),
              r_trunc(path_end[1] * scale), 5, PATH);      /* :94-100 */
}
