/*
 * usv_oracle.h -- CPU restatement (plain C, float64) of the gym-usv env step.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library.  The product path
 * (gym_usv_b200/) never links, imports or calls it.
 *
 * Pinned against the reference's own code executed in-process (oracle/gen_golden.py
 * -> tests/golden/ (npz files), tests/test_oracle_golden.py) and against the three
 * behavioural tests of reference tests/test_usv_asmc.py:8-37.
 *
 * Every function cites the reference file:line it restates (paths relative to the
 * reference root).
 */
#ifndef USV_ORACLE_H
#define USV_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define USV_ORACLE_MAX_OBS 32
#define USV_ORACLE_SENSORS 128
#define USV_ORACLE_OBS_DIM 143

/* gym_usv/control/usv_asmc.py:43-49 -- controller + integrator memory, plus the pose and
 * body velocity the env hands in and gets back (usv_asmc.py:53,244). */
typedef struct UsvOracleAsmc {
    double eta[3];          /* x, y, psi                                   */
    double nu[3];           /* u, v, r                                     */
    double last[9];         /* x_dot,y_dot,psi_dot,u_dot,v_dot,r_dot,e_u,ka_dot_u,ka_dot_psi (last) */
    double so[4];           /* psi_d_last, o_dot_dot_last, o_dot_last, o_last (usv_asmc.py:58,90-92) */
    double aux[3];          /* e_u_int, ka_u, ka_psi                       */
    int64_t perturb_step;   /* usv_asmc.py:49,199                          */
} UsvOracleAsmc;

/* gym_usv/envs/simple_env.py:10-60 + simple_env_asmc.py:10-16 + TimeLimit(1000). */
typedef struct UsvOracleEnv {
    UsvOracleAsmc asmc;
    double path_start[2];
    double path_end[2];
    double target[2];
    double max_action[3];
    double max_accel[3];
    double v_ref;
    double progress;
    double last_action[3];
    double obs_x[USV_ORACLE_MAX_OBS];
    double obs_y[USV_ORACLE_MAX_OBS];
    double obs_r[USV_ORACLE_MAX_OBS];
    double sensor[USV_ORACLE_SENSORS];   /* sensor_data[:,1], simple_env.py:47,329 */
    double ep_return;
    int32_t n_obs;
    int32_t steps;                        /* TimeLimit elapsed steps */
    int32_t episode;                      /* resets so far - 1       */
    int32_t pad_;
} UsvOracleEnv;

typedef struct UsvOracleConfig {
    int32_t ignore_obstacles;     /* simple_env.py:49                         */
    int32_t do_perturb;           /* usv_asmc.py:53 (env passes False)        */
    int32_t skip_kinematic;       /* NOT reference behaviour (opt-in)         */
    int32_t max_episode_steps;    /* gym_usv/__init__.py:33 ; 0 = none        */
    int32_t auto_reset;           /* SB3 VecEnv semantics                     */
    int32_t obstacles_f32;        /* round reset obstacle tables to float32   */
    int32_t n_threads;            /* OpenMP threads for the vec entry points  */
    int32_t kinematic;            /* 1: 'usv-simple' = UsvSimpleEnv.step itself (simple_env.py:310-346, the action
                                   * drives the kinematic update); 0: 'usv-asmc-simple' (simple_env_asmc.py:18-27) */
} UsvOracleConfig;

int  usv_oracle_sizeof_env(void);
int  usv_oracle_sizeof_asmc(void);

/* usv_asmc.py:5-49 */
void usv_oracle_asmc_init(UsvOracleAsmc* s);
/* usv_asmc.py:53-244; info = psi_d,u_d,tport,tstbd of the LAST of the 10 sub-steps */
void usv_oracle_asmc_compute(UsvOracleAsmc* s, const double action[2], int do_perturb, double info[4]);
void usv_oracle_asmc_compute_batch(UsvOracleAsmc* s, int64_t n, const double* actions /*n x 2*/,
                                   int do_perturb, int n_calls, int n_threads);

/* usv_asmc_ca_env.py:411-461,500-519 (static compute_sensor_measurments and helpers) */
void usv_oracle_sensors(const double position[3], int sensor_count, double sensor_max_range,
                        const double* radius, double pitch, const double* posx, const double* posy,
                        int num_obs, const double* distance, double* out_dist);

/* simple_env.py:228-308 via simple_env_asmc.py:14-16, PCG64 replaced by Philox4x32-10
 * (key = seed, counter = (global env id, episode, block)); writes obs[0..14], leaves the
 * sensor part of the caller's obs row stale like simple_env.py:302 does. */
void usv_oracle_simple_reset(UsvOracleEnv* e, const UsvOracleConfig* cfg, uint64_t seed,
                             uint64_t global_id, float* obs);
/* simple_env.py:302: reset observation of the current (e.g. injected) state, all 143 columns */
void usv_oracle_simple_reset_obs(const UsvOracleEnv* e, float* obs);
/* first construction: simple_env.py:32-54 defaults */
void usv_oracle_simple_init(UsvOracleEnv* e);

/* simple_env_asmc.py:18-27 -> simple_env.py:310-346 (+ TimeLimit); cfg->kinematic: simple_env.py:310-346 alone */
void usv_oracle_simple_step(UsvOracleEnv* e, const UsvOracleConfig* cfg, const float action[2],
                            float* obs, double* reward, uint8_t* terminated, uint8_t* truncated);

/* N envs, SB3 VecEnv auto-reset semantics; terminal_obs may be NULL */
void usv_oracle_vec_reset(UsvOracleEnv* envs, int64_t n, const UsvOracleConfig* cfg, uint64_t seed,
                          int64_t global_offset, float* obs);
void usv_oracle_vec_step(UsvOracleEnv* envs, int64_t n, const UsvOracleConfig* cfg, const float* actions,
                         float* obs, double* reward, uint8_t* terminated, uint8_t* truncated,
                         float* terminal_obs, uint64_t seed, int64_t global_offset);

/* ---- legacy path-following env 'usv-asmc-v0' (gym_usv/envs/usv_asmc_env.py), usv_oracle_legacy.c ---- */
/* usv_asmc_env.py:53-58: the six arrays of UsvAsmcEnv.  Values are held in double; after the first step() they are
 * float32 values (usv_asmc_env.py:247-251) and f32_state selects the float32 arithmetic flow. */
typedef struct UsvOracleLegacy {
    double state[6];     /* u, v_ak, r, ye, psi_ak, action_last                      */
    double velocity[3];  /* u, v, r                                                  */
    double position[3];  /* x, y, psi                                                */
    double aux[3];       /* e_u_int, Ka_u, Ka_psi                                    */
    double last[9];      /* eta_dot_last[3], upsilon_dot_last[3], e_u_last, Ka_dot_u_last, Ka_dot_psi_last */
    double target[6];    /* x_0, y_0, desired_speed, ak, x_d, y_d                    */
    double ep_return;
    double ye_int;       /* usv_asmc_ye_int_env.py: aux_vars[3]                      */
    double ye_last;      /* usv_asmc_ye_int_env.py: last[9]                          */
    int32_t f32_state;   /* 0 after reset() (float64 arrays), 1 after a step() of variant 0 */
    int32_t steps;
    int32_t episode;
    int32_t variant;     /* USV_ORACLE_LEGACY_* ; set by the caller after _init      */
} UsvOracleLegacy;

/* the three envs that share UsvAsmcEnv's step(): gym_usv/__init__.py:3-16 */
#define USV_ORACLE_LEGACY_ASMC_V0 0    /* usv_asmc_env.py        (float32 / float64 mixed flow)      */
#define USV_ORACLE_LEGACY_YE_INT 1     /* usv_asmc_ye_int_env.py (float64 throughout, ye integral)   */
#define USV_ORACLE_LEGACY_PID 2        /* usv_pid_env.py         (float64 throughout, PID law)       */

int  usv_oracle_sizeof_legacy(void);
void usv_oracle_legacy_init(UsvOracleLegacy* e);
/* usv_asmc_env.py:258-300 from the seven uniform draws x, y, psi, x_0, y_0, x_d, desired_speed */
void usv_oracle_legacy_reset_from_draws(UsvOracleLegacy* e, const double draws[7], float obs[6]);
/* the same with Philox4x32-10 draws keyed by (seed, global env id, episode) instead of the global np.random */
void usv_oracle_legacy_reset(UsvOracleLegacy* e, uint64_t seed, uint64_t global_id, float obs[6]);
/* usv_asmc_env.py:99-255 (action: the float32 scalar the caller passes; done => reward = -1) */
void usv_oracle_legacy_step(UsvOracleLegacy* e, float action, float obs[6], double* reward, uint8_t* done);
void usv_oracle_legacy_vec_reset(UsvOracleLegacy* envs, int64_t n, uint64_t seed, int64_t global_offset, float* obs,
                                 int n_threads);
void usv_oracle_legacy_vec_step(UsvOracleLegacy* envs, int64_t n, const float* actions, float* obs, double* reward,
                                uint8_t* done, float* terminal_obs, int auto_reset, uint64_t seed,
                                int64_t global_offset, int n_threads);

/* ---- collision-avoidance env 'usv-asmc-ca-v0' (gym_usv/envs/usv_asmc_ca_env.py) with the in-repo UsvAsmc substituted
 * for the absent usv_libs_py, usv_oracle_ca.c ---- */
#define USV_ORACLE_CA_SENSORS 16   /* usv_asmc_ca_env.py:44 */
#define USV_ORACLE_CA_OBS_DIM 23   /* :99  7 + sensor_num */
#define USV_ORACLE_CA_MAX_OBS 9    /* :349 int(uniform(2, 10)) -> 2..9 */
typedef struct UsvOracleCa {
    UsvOracleAsmc asmc;                        /* model pose / velocity + controller memory */
    double target[2];                          /* target_point[:2] (:343-347)               */
    double prev_action[2];                     /* action_history (maxlen 1), denormalised   */
    double obs_x[USV_ORACLE_CA_MAX_OBS];
    double obs_y[USV_ORACLE_CA_MAX_OBS];
    double obs_r[USV_ORACLE_CA_MAX_OBS];
    double ep_return;
    int32_t n_obs;
    int32_t steps;
    int32_t episode;
    int32_t pad_;
} UsvOracleCa;
int  usv_oracle_sizeof_ca(void);
/* reset(options=...) (usv_asmc_ca_env.py:358-372; tools/test_env.py:8-44 builds such dicts): every part is optional */
typedef struct UsvOracleCaOptions {
    int32_t has_obstacles;  /* 'obs_x' in options: obs_x / obs_y / obs_r replace the drawn candidates */
    int32_t n_obstacles;    /* <= USV_ORACLE_CA_MAX_OBS */
    double obs_x[USV_ORACLE_CA_MAX_OBS];
    double obs_y[USV_ORACLE_CA_MAX_OBS];
    double obs_r[USV_ORACLE_CA_MAX_OBS];
    int32_t has_target;     /* 'target_point' */
    int32_t has_start;      /* 'start_position': moves self.position for the deletion rule only -- the DynamicModel was
                               built from the drawn pose before (:336) and step() reads the model (:199-204) */
    double target[2];
    double start[2];
} UsvOracleCaOptions;

void usv_oracle_ca_init(UsvOracleCa* e);
void usv_oracle_ca_reset(UsvOracleCa* e, uint64_t seed, uint64_t global_id, float obs[USV_ORACLE_CA_OBS_DIM]);
void usv_oracle_ca_reset_opt(UsvOracleCa* e, uint64_t seed, uint64_t global_id, const UsvOracleCaOptions* opt,
                             float obs[USV_ORACLE_CA_OBS_DIM]);
void usv_oracle_ca_vec_reset_opt(UsvOracleCa* envs, int64_t n, uint64_t seed, int64_t global_offset,
                                 const UsvOracleCaOptions* opt, float* obs, int n_threads);
void usv_oracle_ca_step(UsvOracleCa* e, const float action_in[2], float obs[USV_ORACLE_CA_OBS_DIM], double* reward,
                        uint8_t* done, uint8_t* trunc, int max_episode_steps);
void usv_oracle_ca_vec_reset(UsvOracleCa* envs, int64_t n, uint64_t seed, int64_t global_offset, float* obs,
                             int n_threads);
void usv_oracle_ca_vec_step(UsvOracleCa* envs, int64_t n, const float* actions, float* obs, double* reward,
                            uint8_t* done, uint8_t* trunc, float* terminal_obs, int auto_reset, int max_episode_steps,
                            uint64_t seed, int64_t global_offset, int n_threads);

/* ---- rgb_array frame of SimpleEnvVisualizer.render_frame (simple_env_visualizer.py:17-114), usv_oracle_render.c;
 * PARITY UNPINNED against pygame's pixels (pygame absent): draw list from the reference, coverage rules ours ---- */
void usv_oracle_render_frame(const double position[3], const double target[2], const double* sensor_distance,
                             int sensor_count, int n_obs, const double* obs_x, const double* obs_y,
                             const double* obs_r, const double path_start[2], const double path_end[2],
                             double env_hi, int size, uint8_t* frame);

/* RNG primitives exposed for tests (Philox4x32-10) */
void usv_oracle_philox(uint64_t seed, uint64_t global_id, uint32_t episode, uint32_t block, uint32_t out[4]);

#ifdef __cplusplus
}
#endif
#endif
