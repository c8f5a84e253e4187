/*
 * usv_oracle_ca.c -- CPU restatement of the collision-avoidance env 'usv-asmc-ca-v0'
 * (gym_usv/envs/usv_asmc_ca_env.py: UsvAsmcCaEnv; 16-ray lidar, point target, <= 9 obstacles, TimeLimit 5000),
 * with the in-repo controller substituted for the absent usv_libs_py: one env step runs one
 * UsvAsmc.compute (gym_usv/control/usv_asmc.py:53-244, 10 sub-steps of 0.01 s) with
 * action = [setpoint.velocity, setpoint.heading] -- the substitution the reference keeps as a commented-out line
 * (usv_asmc_ca_env.py:210).
 *
 * TEST INFRASTRUCTURE ONLY (see usv_oracle.h).
 *
 * Pinned by tests/golden/ca_env.npz = the reference's UsvAsmcCaEnv.reset / step executed unmodified over a
 * usv_libs_py-shaped shim of gym_usv.control.UsvAsmc, with NUMBA_DISABLE_JIT=1 because the env's own @njit helper does
 * not type at HEAD (oracle/gen_golden_ca.py explains both).  "Parity pinned against the reference's env code, dynamics
 * substituted": the controller / model numbers are the in-repo UsvAsmc's, not usv_libs_py's.
 */
#include "usv_oracle.h"

#include <math.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* usv_asmc_ca_env.py:44-58, 60-63, 82-86 */
#define CA_SENSORS USV_ORACLE_CA_SENSORS
#define CA_MAX_RANGE 100.0
#define CA_BOAT_RADIUS 0.1
#define CA_SAFETY 0.3
#define CA_MAX_U (2.5 / 2)
#define CA_MAX_R 3.5
#define CA_MAX_X 30.0
#define CA_MIN_X (-10.0)
#define CA_MAX_Y 10.0
#define CA_MIN_Y (-10.0)

int usv_oracle_sizeof_ca(void) { return (int)sizeof(UsvOracleCa); }

static double ca_wrap(double a) { return atan2(sin(a), cos(a)); } /* :131-132 */

/* :134-144 with in_min = -1, in_max = 1, evaluated in the dtype numpy gives it: a float32 action element keeps the
 * whole expression in float32 (NEP 50: the Python scalars are weak) */
static float ca_denorm_f32(float x, double out_min, double out_max) {
    float t = x - (float)(-1);
    t = t * (float)(out_max - out_min);
    t = t / (float)(1 - (-1));
    return t + (float)out_min;
}

/* :146-325 after the action has been denormalised */
static void ca_step_core(UsvOracleCa* e, const double action[2], float obs[USV_ORACLE_CA_OBS_DIM], double* reward_out,
                         uint8_t* done_out, uint8_t* trunc_out) {
    /* :195-205 -> usv_asmc.py:53-244 */
    usv_oracle_asmc_compute(&e->asmc, action, 0, NULL);
    const double x = e->asmc.eta[0], y = e->asmc.eta[1], psi = e->asmc.eta[2];
    const double u = e->asmc.nu[0], v = e->asmc.nu[1], r = e->asmc.nu[2];

    /* :229-246 collision */
    double distance[USV_ORACLE_CA_MAX_OBS];
    int collision = 0;
    double dmin = INFINITY;
    for (int j = 0; j < e->n_obs; ++j) {
        distance[j] = hypot(e->obs_x[j] - x, e->obs_y[j] - y) - e->obs_r[j] - CA_BOAT_RADIUS;
        if (distance[j] < dmin) dmin = distance[j];
    }
    if (e->n_obs > 0) collision = dmin < 0;

    /* :249-259 */
    double sens[CA_SENSORS];
    const double pitch = ((2.0 / 3) * (2 * M_PI)) / CA_SENSORS; /* :47-49 */
    usv_oracle_sensors(e->asmc.eta, CA_SENSORS, CA_MAX_RANGE, e->obs_r, pitch, e->obs_x, e->obs_y, e->n_obs, distance, sens);

    /* :261-271 */
    const double tx = e->target[0], ty = e->target[1];
    const double distance_to_target = hypot(x - tx, y - ty);
    const double angle_to_target = ca_wrap(atan2(ty - y, tx - x) - psi); /* :405-409 */
    const int arrived = distance_to_target < 1.5;
    const double c = cos(psi), s = sin(psi);
    const double d0 = tx - x, d1 = ty - y, d2 = 0.0 - psi; /* target_point[2] = 0 (:347) */
    const double te0 = c * d0 + s * d1 + 0.0 * d2;
    const double te1 = -s * d0 + c * d1 + 0.0 * d2;
    const double te2 = ca_wrap(angle_to_target);
    const double div_fac = CA_MAX_X * CA_MAX_X + CA_MAX_Y * CA_MAX_Y;

    /* :276-282, 485-498 */
    const double r_tracking_error = -hypot(te0, te1) / 75 - fabs(angle_to_target / M_PI);
    const double r_velocity = hypot(u, v);
    double reward = r_tracking_error + r_velocity * 0.5;

    /* :284-289 */
    double st[USV_ORACLE_CA_OBS_DIM];
    st[0] = u / CA_MAX_U;
    st[1] = r / CA_MAX_R;
    st[2] = te0 / div_fac;
    st[3] = te1 / div_fac;
    st[4] = te2 / M_PI;
    st[5] = e->prev_action[0] / M_PI; /* average of a 1-deep history / max(max_action0, max_action1) */
    st[6] = e->prev_action[1] / M_PI;
    for (int k = 0; k < CA_SENSORS; ++k) st[7 + k] = sens[k] / CA_MAX_RANGE;

    /* :294 */
    e->prev_action[0] = action[0];
    e->prev_action[1] = action[1];

    /* :296-311 */
    uint8_t done = 0, trunc = 0;
    if (arrived) done = 1;
    if (collision) trunc = 1;
    if (hypot(te0, te1) > 40) {
        done = 1;
        reward -= 100;
    }
    for (int k = 0; k < USV_ORACLE_CA_OBS_DIM; ++k) obs[k] = (float)st[k]; /* :306 */
    if (fmax(fmax(fabs(x), fabs(y)), fabs(psi)) > 100) {
        done = 1;
        trunc = 1;
    }
    *reward_out = reward;
    *done_out = done;
    *trunc_out = trunc;
}

/* usv_asmc_ca_env.py:146-325; action_in: the float32 pair a VecEnv hands over */
void usv_oracle_ca_step(UsvOracleCa* e, const float action_in[2], float obs[USV_ORACLE_CA_OBS_DIM], double* reward,
                        uint8_t* done, uint8_t* trunc, int max_episode_steps) {
    const double action[2] = {(double)ca_denorm_f32(action_in[0], -1, 1), (double)ca_denorm_f32(action_in[1], -M_PI, M_PI)};
    ca_step_core(e, action, obs, reward, done, trunc);
    e->steps += 1;
    e->ep_return += *reward;
    if (max_episode_steps > 0 && e->steps >= max_episode_steps) *trunc = 1; /* TimeLimit, gym_usv/__init__.py:18-22 */
}

static inline double ca_u53(uint32_t a, uint32_t b) {
    return ((double)(a >> 5) * 67108864.0 + (double)(b >> 6)) * (1.0 / 9007199254740992.0);
}
static void ca_pair(uint64_t seed, uint64_t gid, uint32_t episode, uint32_t block, double* a, double* b) {
    uint32_t o[4];
    usv_oracle_philox(seed, gid, episode, block, o);
    *a = ca_u53(o[0], o[1]);
    *b = ca_u53(o[2], o[3]);
}

/* reset() (:327-403) with the global np.random replaced by Philox4x32-10 keyed by (seed, global env id, episode):
 * block 0 -> x, y | 1 -> theta, target x | 2 -> target y, num_obs | 3 + 2 j -> radius of candidate j |
 * 4 + 2 j -> Box-Muller pair for its (x, y).  Ends, like the reference, with step([-1, 0]) whose observation is the
 * reset observation. */
void usv_oracle_ca_reset(UsvOracleCa* e, uint64_t seed, uint64_t global_id, float obs[USV_ORACLE_CA_OBS_DIM]) {
    usv_oracle_ca_reset_opt(e, seed, global_id, NULL, obs);
}

/* The same with reset(options=...) (:358-372): the draws are made first, as in the reference, then the options replace
 * the obstacle candidates / the target / self.position, then the deletion rule (:374-399) runs on whatever is there. */
void usv_oracle_ca_reset_opt(UsvOracleCa* e, uint64_t seed, uint64_t global_id, const UsvOracleCaOptions* opt,
                             float obs[USV_ORACLE_CA_OBS_DIM]) {
    e->episode += 1;
    const uint32_t ep = (uint32_t)e->episode;
    double a, b;
    ca_pair(seed, global_id, ep, 0, &a, &b);
    const double x = CA_MIN_X + (CA_MAX_X - CA_MIN_X) * a;                 /* :331 */
    const double y = CA_MIN_Y + ((CA_MIN_Y + 5) - CA_MIN_Y) * b;           /* :332 */
    ca_pair(seed, global_id, ep, 1, &a, &b);
    const double theta = -M_PI / 4 + (M_PI / 4 - -M_PI / 4) * a;           /* :333 */
    double tx = CA_MIN_X + ((CA_MAX_X - 10) - CA_MIN_X) * b;               /* :343-346 */
    ca_pair(seed, global_id, ep, 2, &a, &b);
    double ty = (CA_MAX_Y - 5) + ((CA_MAX_Y - 1) - (CA_MAX_Y - 5)) * a;
    int num_obs = (int)(2 + (10 - 2) * b);                                 /* :349 int(uniform(2, 10)) */
    const double cx = (x + tx) / 2, cy = (y + ty) / 2;                     /* :353 np.average of two (drawn target) */
    const int given = opt && opt->has_obstacles;
    if (given) num_obs = opt->n_obstacles;                                 /* :361-365 */
    if (opt && opt->has_target) {                                          /* :367-368 */
        tx = opt->target[0];
        ty = opt->target[1];
    }
    const double px = (opt && opt->has_start) ? opt->start[0] : x;         /* :370-372 self.position (not the model) */
    const double py = (opt && opt->has_start) ? opt->start[1] : y;
    usv_oracle_asmc_init(&e->asmc);                                        /* :336, 385 fresh model and controller */
    e->asmc.eta[0] = x;
    e->asmc.eta[1] = y;
    e->asmc.eta[2] = theta;
    e->target[0] = tx;
    e->target[1] = ty;
    e->prev_action[0] = e->prev_action[1] = 0.0;                           /* :341 */
    int kept = 0;
    for (int j = 0; j < num_obs; ++j) {
        double rr, ox, oy;
        if (given) {
            rr = opt->obs_r[j];
            ox = opt->obs_x[j];
            oy = opt->obs_y[j];
        } else {
            double ur, unused, u1, u2;
            ca_pair(seed, global_id, ep, 3 + 2 * j, &ur, &unused);
            ca_pair(seed, global_id, ep, 4 + 2 * j, &u1, &u2);
            rr = 1 + (2 - 1) * ur;                                         /* :354 */
            const double rad = sqrt(-2.0 * log(1.0 - u1)), ang = 2.0 * M_PI * u2;
            ox = cx + 10 * (rad * cos(ang));                               /* :355-356 */
            oy = cy + 10 * (rad * sin(ang));
        }
        /* :379-399 obstacles too close to the boat or to the target are deleted */
        const double margin = CA_SAFETY + 0.35; /* hypot - r - boat_radius - (safety_radius + 0.35), left to right */
        if (((hypot(ox - px, oy - py) - rr) - CA_BOAT_RADIUS) - margin < 0) continue;
        if (((hypot(ox - tx, oy - ty) - rr) - CA_BOAT_RADIUS) - margin < 0) continue;
        e->obs_x[kept] = ox;
        e->obs_y[kept] = oy;
        e->obs_r[kept] = rr;
        ++kept;
    }
    for (int j = kept; j < USV_ORACLE_CA_MAX_OBS; ++j) e->obs_x[j] = e->obs_y[j] = e->obs_r[j] = 0.0;
    e->n_obs = kept;
    /* :402 state, _, _, _, _ = self.step([-1, 0]): Python ints, float64 arithmetic: action = [-1.0, 0.0] */
    const double action[2] = {-1.0, 0.0};
    double rew;
    uint8_t d, t;
    float tmp[USV_ORACLE_CA_OBS_DIM];
    ca_step_core(e, action, obs ? obs : tmp, &rew, &d, &t);
    e->steps = 0;
    e->ep_return = 0.0;
}

void usv_oracle_ca_init(UsvOracleCa* e) {
    memset(e, 0, sizeof(*e));
    e->episode = -1;
}

void usv_oracle_ca_vec_reset(UsvOracleCa* envs, int64_t n, uint64_t seed, int64_t global_offset, float* obs,
                             int n_threads) {
    (void)n_threads;
#pragma omp parallel for schedule(static) num_threads(n_threads > 0 ? n_threads : 1)
    for (int64_t i = 0; i < n; ++i)
        usv_oracle_ca_reset(&envs[i], seed, (uint64_t)(global_offset + i), obs + USV_ORACLE_CA_OBS_DIM * i);
}

void usv_oracle_ca_vec_reset_opt(UsvOracleCa* envs, int64_t n, uint64_t seed, int64_t global_offset,
                                 const UsvOracleCaOptions* opt, float* obs, int n_threads) {
    (void)n_threads;
#pragma omp parallel for schedule(static) num_threads(n_threads > 0 ? n_threads : 1)
    for (int64_t i = 0; i < n; ++i)
        usv_oracle_ca_reset_opt(&envs[i], seed, (uint64_t)(global_offset + i), opt, obs + USV_ORACLE_CA_OBS_DIM * i);
}

/* N envs, SB3 VecEnv auto-reset semantics; terminated = done, truncated = collision / TimeLimit (both may be set) */
void usv_oracle_ca_vec_step(UsvOracleCa* envs, int64_t n, const float* actions, float* obs, double* reward,
                            uint8_t* done, uint8_t* trunc, float* terminal_obs, int auto_reset, int max_episode_steps,
                            uint64_t seed, int64_t global_offset, int n_threads) {
    (void)n_threads;
#pragma omp parallel for schedule(static) num_threads(n_threads > 0 ? n_threads : 1)
    for (int64_t i = 0; i < n; ++i) {
        float* o = obs + USV_ORACLE_CA_OBS_DIM * i;
        usv_oracle_ca_step(&envs[i], actions + 2 * i, o, &reward[i], &done[i], &trunc[i], max_episode_steps);
        if ((done[i] || trunc[i]) && auto_reset) {
            if (terminal_obs) memcpy(terminal_obs + USV_ORACLE_CA_OBS_DIM * i, o, USV_ORACLE_CA_OBS_DIM * sizeof(float));
            usv_oracle_ca_reset(&envs[i], seed, (uint64_t)(global_offset + i), o);
        }
    }
}
