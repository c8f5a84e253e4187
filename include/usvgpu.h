/*
 * usvgpu.h -- C ABI of libusvgpu.so: the batched gym-usv environment step on one B200.
 *
 * One handle owns the structure-of-arrays state of N lock-step environments on one
 * device.  The CALLER owns the action / observation / reward / flag buffers (device
 * memory, e.g. torch CUDA tensors passed as raw pointers).  All work is stream-ordered on
 * the cudaStream_t the caller passes; nothing here synchronises unless it says so.
 * Every function returns 0 on success or a negative UsvStatus; usvgpu_last_error() gives
 * the text.  There is NO CPU fallback: without a CUDA device every call fails.
 *
 * Each entry point cites the reference interface it replaces (paths relative to the
 * reference root, romi2002/gym-usv).
 */
#ifndef USVGPU_H
#define USVGPU_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define USVGPU_VERSION 100         /* 0.1.0 */
#define USVGPU_OBS_DIM 143         /* gym_usv/envs/simple_env.py:27  (15 + 128 sensors)      */
#define USVGPU_ACT_DIM 2           /* gym_usv/envs/simple_env.py:30                           */
#define USVGPU_SENSORS 128         /* gym_usv/envs/simple_env.py:11                           */
#define USVGPU_MAX_OBSTACLES 32    /* reset draws 15..29 (simple_env.py:257); padded to 32    */
#define USVGPU_CA_OBS_DIM 23       /* gym_usv/envs/usv_asmc_ca_env.py:99 (7 + 16 sensors)     */
#define USVGPU_LEGACY_OBS_DIM 6    /* gym_usv/envs/usv_asmc_env.py:89-96                      */

typedef struct UsvGpu* usvgpu_handle;

typedef enum UsvStatus {
    USV_OK = 0,
    USV_ERR_INVALID = -1,   /* bad argument                          */
    USV_ERR_CUDA = -2,      /* CUDA runtime error (text in last_error) */
    USV_ERR_NO_DEVICE = -3, /* no usable CUDA device: there is no CPU path */
    USV_ERR_ALLOC = -4
} UsvStatus;

typedef enum UsvEnvKind {
    USV_ENV_ASMC_SIMPLE = 0, /* 'usv-asmc-simple' gym_usv/__init__.py:30-34 -> envs/simple_env_asmc.py */
    USV_ENV_ASMC_V0 = 1,     /* 'usv-asmc-v0'     gym_usv/__init__.py:3-6   -> envs/usv_asmc_env.py    */
    USV_ENV_SIMPLE = 2,      /* 'usv-simple'      gym_usv/__init__.py:24-28 -> envs/simple_env.py: the kinematic
                              * env train_test/sb3_train_vec.py:55 trains on (TimeLimit 500).  Same spaces, state
                              * fields and entry points as 'usv-asmc-simple'; USV_FIELD_DYN rows 6 and 7 hold
                              * last_action[0] and last_action[2] (simple_env.py:41,343), rows 8..21 are unused */
    /* the two float64 siblings of 'usv-asmc-v0': same spaces (obs [N][6], actions [N][1]) and entry points; their
     * state is USV_FIELD_LEGACY64 (+ USV_FIELD_LEGACY_TARGET), USV_FIELD_LEGACY / _POS0 do not exist for them */
    USV_ENV_ASMC_YE_INT_V0 = 3, /* 'usv-asmc-ye-int-v0' gym_usv/__init__.py:13-16 -> envs/usv_asmc_ye_int_env.py */
    USV_ENV_PID_V0 = 4,         /* 'usv-pid-v0'         gym_usv/__init__.py:8-11  -> envs/usv_pid_env.py         */
    /* 'usv-asmc-ca-v0' gym_usv/__init__.py:18-22 -> envs/usv_asmc_ca_env.py: point target, <= 9 obstacles, 16-ray lidar,
     * observation float32[23], action float32[2] in [-1, 1]^2, TimeLimit 5000.  usv_libs_py (its model / controller)
     * is not part of the reference tree: the env runs the in-repo UsvAsmc (one compute() of 10 sub-steps per step),
     * the substitution usv_asmc_ca_env.py:210 keeps as a comment.  State: USV_FIELD_DYN (row 22 unused) + USV_FIELD_CA
     * + USV_FIELD_CA_N_OBS.  terminated = arrived | tracking error > 40 | out of range; truncated = collision |
     * out of range | TimeLimit.  No render, no perturbation, no info rows. */
    USV_ENV_ASMC_CA_V0 = 5
} UsvEnvKind;

typedef enum UsvPrecision {
    USV_PREC_F64 = 0,  /* float64 state + arithmetic, as the reference (numpy float64)                 */
    USV_PREC_MIXED = 1 /* float64 integrator state, float32 force/transcendental terms */
} UsvPrecision;

/* Mirrors the knobs of the reference envs; literals that the reference hard-codes
 * (vessel coefficients, gains, sensor fan, bounds) are compile-time constants. */
typedef struct UsvConfig {
    int32_t env_kind;          /* UsvEnvKind                                                      */
    int32_t ignore_obstacles;  /* simple_env.py:49,155,222-224,334                                */
    int32_t do_perturb;        /* usv_asmc.py:53,184-199 (UsvSimpleASMCEnv passes False)          */
    int32_t skip_kinematic;    /* simple_env.py:310 update_position=False; NOT the default        */
    int32_t max_episode_steps; /* TimeLimit: 1000 for usv-asmc-simple; 0 = none                   */
    int32_t auto_reset;        /* SB3 VecEnv semantics: done envs are reset inside step()         */
    int32_t precision;         /* UsvPrecision                                                    */
    int32_t obstacles_f32;     /* in-kernel reset draws obstacle tables on the float32 grid and   *
                                * stores them as float32 (halves the table traffic)               */
} UsvConfig;

/* Episode statistics accumulated on the device (what SB3 Monitor / RecordEpisodeStatistics
 * would log); sums are over episodes finished since the last usvgpu_episode_stats(reset=1). */
typedef struct UsvStats {
    double   return_sum;
    double   length_sum;
    uint64_t episodes;
    uint64_t terminated;      /* collisions   simple_env.py:334 */
    uint64_t out_of_bounds;   /* simple_env.py:336              */
    uint64_t time_limit;      /* TimeLimit truncations          */
    uint64_t env_steps;       /* total env-steps executed       */
    uint64_t nonfinite;       /* env-steps that produced a non-finite pose (guard counter) */
} UsvStats;

/* State field groups for usvgpu_get_state / usvgpu_set_state.  Layout of every group is
 * [rows][n_envs] (structure of arrays, env index fastest) except the obstacle table, whose
 * row is one env (the sensor kernel gives a warp to an env, lane = obstacle slot). */
typedef enum UsvField {
    /* float64 [23][N]: x y psi | u v r | x_dot y_dot psi_dot u_dot v_dot r_dot e_u ka_dot_u ka_dot_psi (last,
     * usv_asmc.py:57) | psi_d_last o_dot_dot_last o_dot_last o_last (usv_asmc.py:58) |
     * e_u_int ka_u ka_psi (usv_asmc.py:63) | progress (simple_env.py:53) */
    USV_FIELD_DYN = 0,
    /* float64 [7][N]: path_start x y | path_end x y | max_action[0] max_action[2] | reference_velocity */
    USV_FIELD_EPI = 1,
    /* float64 (float32 when obstacles_f32) [N][3][32]: per env x[32], y[32], radius[32]; slots >= n_obs
     * are ignored (simple_env.py:258,290) */
    USV_FIELD_OBSTACLES = 2,
    USV_FIELD_N_OBS = 3,     /* int32 [N]   simple_env.py:268                    */
    USV_FIELD_STEPS = 4,     /* int32 [N]   TimeLimit elapsed steps              */
    USV_FIELD_EPISODE = 5,   /* int32 [N]   resets so far (Philox counter word)  */
    USV_FIELD_EP_RETURN = 6, /* float64 [N] running episode return               */
    /* legacy env (USV_ENV_ASMC_V0), float32 [19][N]: action_last (state[5]) | velocity u v r | position x y psi |
     * aux_vars e_u_int Ka_u Ka_psi | last[9] (usv_asmc_env.py:108-113).  After a step every one of them IS a float32
     * (usv_asmc_env.py:247-251). */
    USV_FIELD_LEGACY = 7,
    USV_FIELD_LEGACY_TARGET = 8, /* float64 [6][N]: x_0 y_0 desired_speed ak x_d y_d (usv_asmc_env.py:296) */
    /* float64 [3][N]: pose x y psi of an env that was reset and has not stepped yet (STEPS == 0): reset() stores
     * float64 arrays (usv_asmc_env.py:293) and the first step computes in float64 from them */
    USV_FIELD_LEGACY_POS0 = 9,
    /* USV_ENV_ASMC_YE_INT_V0 / USV_ENV_PID_V0, float64 [21][N]: the 19 rows of USV_FIELD_LEGACY in float64 (these
     * envs never round to float32) | ye_int (aux_vars[3]) | ye_last (last[9]) of usv_asmc_ye_int_env.py:247-251;
     * the last two rows stay 0 for the pid env */
    USV_FIELD_LEGACY64 = 10,
    /* USV_ENV_ASMC_CA_V0, float64 [31][N]: target x y | previous action (2, denormalised) | obstacle x[9] | y[9] |
     * radius[9] (usv_asmc_ca_env.py:341-356) */
    USV_FIELD_CA = 11,
    USV_FIELD_CA_N_OBS = 12, /* int32 [N] obstacles left after the reset's deletions (:379-399) */
    /* float32 [N][2]: the random target_position a reset draws (simple_env.py:245).  It is what the reset observation,
     * reset()'s info and the renderer show for an env that has not stepped yet (USV_FIELD_STEPS == 0); the first
     * step replaces it by the closest point of the path (:328).  Simple envs only. */
    USV_FIELD_TARGET0 = 13
} UsvField;

#define USV_LEGACY_ROWS 19
#define USV_LEGACY64_ROWS 21
#define USV_CA_ROWS 31
#define USV_DYN_ROWS 23
#define USV_EPI_ROWS 7

int usvgpu_version(void);
/* sha256 (16 hex digits) over the CUDA sources this binary was compiled from, as gym_usv_b200/_native.py::source_hash()
 * computes it for the tree; "unknown" for a build that did not pass -DUSVGPU_SOURCE_HASH.  No reference counterpart:
 * it makes "the tested binary is this tree" checkable (the .so is git-ignored and travels to the GPU box prebuilt). */
const char* usvgpu_source_hash(void);
/* text of the last error on this handle (or of the last failed usvgpu_create when h == NULL) */
const char* usvgpu_last_error(usvgpu_handle h);

/* Replaces N x `gym.make('usv-asmc-simple')` / SB3 make_vec_env (train_test/sb3_train_vec.py:67):
 * allocates the SoA state for n_envs on `device`.  global_env_offset is the index of this
 * shard's first env in the whole job: the reset RNG is Philox4x32-10 keyed by (seed, global
 * env id, episode), so trajectories do not depend on how envs are sharded over GPUs. */
int usvgpu_create(const UsvConfig* cfg, int device, int64_t n_envs, int64_t global_env_offset, uint64_t seed,
                  usvgpu_handle* out);
int usvgpu_destroy(usvgpu_handle h);
/* observation / action width of the handle's env kind: 143 / 2 ('usv-asmc-simple', simple_env.py:27,30) or
 * 6 / 1 ('usv-asmc-v0', usv_asmc_env.py:89-96).  Every obs / action buffer below is [N][obs_dim] / [N][act_dim].
 * For 'usv-asmc-v0' usvgpu_reset / usvgpu_step replace UsvAsmcEnv.reset / .step (usv_asmc_env.py:258-300, 99-255):
 * terminated = the reference's `done`, truncated only with max_episode_steps > 0 (none is registered:
 * gym_usv/__init__.py:3-6). */
int usvgpu_obs_dim(usvgpu_handle h);
int usvgpu_act_dim(usvgpu_handle h);

/* Replaces Env.reset() (simple_env_asmc.py:14-16 -> simple_env.py:228-308) for the envs whose
 * mask byte is non-zero (all envs when mask == NULL).  Writes observation columns 0..14 of the
 * selected rows of obs[N][143] and leaves columns 15..142 untouched -- the reference builds
 * the reset observation from the previous episode's stale sensor data (simple_env.py:302). */
int usvgpu_reset(usvgpu_handle h, const uint8_t* mask, float* obs, void* stream);

/* Replaces VecEnv.step(): for every env 2 x UsvAsmc.compute (usv_asmc.py:53-244, 20 sub-steps)
 * then UsvSimpleEnv.step(zeros) (simple_env.py:310-346) then TimeLimit.  actions float32 [N][2],
 * obs float32 [N][143], reward float32 [N], terminated/truncated uint8 [N] (all device pointers).
 * With auto_reset, done envs are reset in the same launch: obs then holds the reset observation
 * and, when terminal_obs != NULL, rows of done envs in terminal_obs[N][143] hold the last
 * observation of the finished episode (SB3 infos[i]["terminal_observation"]). */
int usvgpu_step(usvgpu_handle h, const float* actions, float* obs, float* reward, uint8_t* terminated,
                uint8_t* truncated, float* terminal_obs, void* stream);

/* usvgpu_step() that also reports the device time of its two kernels (CUDA events on `stream`):
 * kernel_ms[0] = dynamics_kernel, kernel_ms[1] = sensor_kernel.  Synchronises `stream`.  For
 * bench.py's roofline line; not for the training loop. */
int usvgpu_step_timed(usvgpu_handle h, const float* actions, float* obs, float* reward, uint8_t* terminated,
                      uint8_t* truncated, float* terminal_obs, void* stream, float* kernel_ms);

/* K back-to-back steps (two launches each; state round-trips HBM every step) with pre-staged
 * actions float32 [K][N][2]; outputs are overwritten every step.  action_wrap > 0 re-uses the
 * first action_wrap slabs cyclically.  This is the loop of tools/profile_env.py:5-7. */
int usvgpu_rollout(usvgpu_handle h, const float* actions, int32_t n_steps, int32_t action_wrap, float* obs,
                   float* reward, uint8_t* terminated, uint8_t* truncated, void* stream);

/* ---- host-buffer face: what stock SB3 does with a VecEnv (numpy actions in, numpy observations out;
 * DummyVecEnv.step_wait, train_test/sb3_train_vec.py:67).  All pointers below are HOST memory.  The handle owns the
 * device staging buffers.  Buffers should be page-locked (usvgpu_host_alloc, cudaHostRegister or torch
 * pin_memory) so that the copies run asynchronously on `stream`; pageable memory works but is staged by the
 * driver.  usvgpu_step_host only ENQUEUES (H2D actions, the two kernels, D2H reward / flags / observations):
 * read the outputs after usvgpu_stream_sync().  mapped_obs != 0: obs_host must come from usvgpu_host_alloc; the
 * sensor kernel then stores the observation rows straight into it over PCIe (no staging copy of the 572 B rows).
 * obs_host == NULL: the observation rows stay on the device (in the staging tensor of usvgpu_use_staging, i.e. the
 * caller's own CUDA tensor) and only reward / flags cross PCIe -- for a caller whose policy reads the observations on
 * the GPU but whose bookkeeping (episode statistics, terminal observations of the finished envs through
 * usvgpu_terminal_obs_host) lives on the host: 6 bytes per env-step over the bus instead of 578. */
/* Optional: make the host-buffer face stage through the CALLER's device buffers (actions [N][2], obs [N][143],
 * reward [N] float32, flags uint8 [2][N] = terminated then truncated, terminal_obs [N][143]) instead of
 * handle-owned ones, so that a caller mixing usvgpu_step and usvgpu_step_host sees one set of buffers. */
int usvgpu_use_staging(usvgpu_handle h, float* actions, float* obs, float* reward, uint8_t* flags, float* terminal_obs);
int usvgpu_host_alloc(void** out, int64_t bytes);
int usvgpu_host_free(void* p);
int usvgpu_reset_host(usvgpu_handle h, const uint8_t* mask_host, float* obs_host, void* stream); /* synchronises */
int usvgpu_step_host(usvgpu_handle h, const float* actions_host, float* obs_host, float* reward_host,
                     uint8_t* terminated_host, uint8_t* truncated_host, int32_t mapped_obs, void* stream);
int usvgpu_stream_sync(usvgpu_handle h, void* stream);
/* usvgpu_step_host steps the shard in `chunks` (1..16, default 2) env ranges -- a first small one (1/8 of the
 * shard) and chunks - 1 even ones -- so that the copy-out of one range overlaps the kernels of the next; results do
 * not depend on it. */
int usvgpu_set_host_chunks(usvgpu_handle h, int32_t chunks);
/* Blocks until reward / terminated / truncated of the last usvgpu_step_host are on the host (the observation copy
 * may still be in flight: the caller can do its episode bookkeeping meanwhile, then usvgpu_stream_sync). */
int usvgpu_flags_sync(usvgpu_handle h);
/* SB3 infos[i]["terminal_observation"] for the n_index envs listed in index_host (those whose flags came back set
 * from the last usvgpu_step_host): rows_host float32 [n_index][143] (any host memory); episode_host (may be NULL)
 * float64 [n_index][2] = return and length of the episode that just ended (SB3 Monitor's infos[i]["episode"]
 * r, l).  Runs on a handle-owned side stream beside the observation copy; returns when the rows are in rows_host. */
int usvgpu_terminal_obs_host(usvgpu_handle h, const int32_t* index_host, int64_t n_index, float* rows_host,
                             double* episode_host);

/* Copy a whole field group out of / into the handle.  buf may be host or device memory
 * (cudaMemcpyDefault); size is checked against usvgpu_field_bytes(). */
int64_t usvgpu_field_bytes(usvgpu_handle h, int32_t field);
int usvgpu_get_state(usvgpu_handle h, int32_t field, void* buf, int64_t bytes, void* stream);
int usvgpu_set_state(usvgpu_handle h, int32_t field, const void* buf, int64_t bytes, void* stream);

/* Device-side statistics -> host struct (synchronises `stream`).  reset != 0 zeroes them after. */
int usvgpu_episode_stats(usvgpu_handle h, UsvStats* out, int32_t reset, void* stream);
/* device address of the UsvStats block (for an NCCL all-reduce issued by the caller) */
int usvgpu_stats_device_ptr(usvgpu_handle h, void** out);

/* Unit-parity entry point: UsvAsmc.compute (usv_asmc.py:53-244) for n independent
 * controller+vessel states.  state float64 [23][n] uses the USV_FIELD_DYN row order (row 22
 * is ignored); actions float64 [n][2]; n_calls consecutive calls of 10 sub-steps each;
 * first_substep is UsvAsmc.perturb_step at entry; info (may be NULL) float64 [4][n] receives
 * psi_d, u_d, tport, tstbd of the last sub-step (usv_asmc.py:78,125,168-169). */
int usvgpu_asmc_compute(int device, int64_t n, double* state, const double* actions, int32_t n_calls,
                        int32_t do_perturb, int64_t first_substep, int32_t precision, double* info, void* stream);

/* Unit-parity entry point: UsvAsmcCaEnv.compute_sensor_measurments (usv_asmc_ca_env.py:411-437)
 * with the 128-ray fan of simple_env.py:11-14.  pose float64 [3][n]; obstacles float64
 * [32][3][n]; n_obs int32 [n]; out float64 [128][n] distances. */
int usvgpu_sensors(int device, int64_t n, const double* pose, const double* obstacles, const int32_t* n_obs,
                   double* out, void* stream);

/* usvgpu_sensors for callers that must not allocate (code under CUDA graph capture, serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same arguments,
 * same results as usvgpu_sensors (which allocates its ray table on a device's first call).  This one calls no
 * cudaMalloc / cudaFree and no allocating library routine: its scratch -- the obstacle table re-laid as one 768-byte
 * row per env, plus the 2 KiB ray table -- is a file-scope `static __device__` array of
 * USVGPU_NOALLOC_MAX_ENVS * 768 + 2048 bytes (6 MiB + 2 KiB) that every GPU loading the library holds until the
 * process exits.  n > USVGPU_NOALLOC_MAX_ENVS is refused with USV_ERR_INVALID and a message (usvgpu_last_error(NULL)).
 * The first call on a device looks the array's address up (cudaGetSymbolAddress) and copies the ray table in; make
 * that call outside stream capture.  ALL calls share the one array, which keeps the last call's contents: every row a
 * call reads it has written itself first (unused obstacle slots as 0), so nothing needs clearing between calls, but
 * two calls that overlap on the GPU (different streams, or graphs replayed concurrently) race on it -- order them. */
#define USVGPU_NOALLOC_MAX_ENVS 8192
int usvgpu_sensors_noalloc(int device, int64_t n, const double* pose, const double* obstacles, const int32_t* n_obs,
                           double* out, void* stream);

/* Replaces stable_baselines3 VecFrameStack / StackedObservations.update (train_test/sb3_train_vec.py:70,
 * train_test/sb3_train.py:50-51: n_stack = 5) for device-resident observations, without moving old frames.
 * ring: float32 [N][2 * n_stack][obs_dim], zero-initialised by the caller.  Call k (k = 0 for the reset observation)
 * passes pos = k % n_stack; afterwards the stacked observation of env i (oldest frame first, SB3's layout for 1-D
 * observations) is the contiguous run ring[i][pos + 1 .. pos + n_stack] -- a [N][n_stack * obs_dim] view with row
 * stride 2 * n_stack * obs_dim.  Envs with terminated | truncated set (either pointer may be NULL; pass both NULL for
 * the reset observation) get their older frames zeroed, and, when terminal_obs and terminal_stack are given,
 * terminal_stack[i] = the previous n_stack - 1 frames followed by terminal_obs[i] (SB3's stacked
 * infos[i]["terminal_observation"]); terminal_stack float32 [N][n_stack * obs_dim], rows of other envs untouched. */
int usvgpu_frame_stack_push(int device, int64_t n, int32_t obs_dim, int32_t n_stack, int32_t pos, float* ring,
                            const float* obs, const uint8_t* terminated, const uint8_t* truncated,
                            const float* terminal_obs, float* terminal_stack, void* stream);

/* Replaces UsvSimpleEnv.render() / _render_frame() with render_mode == "rgb_array" (gym_usv/envs/simple_env.py:117-131)
 * -> SimpleEnvVisualizer.render_frame (gym_usv/envs/simple_env_visualizer.py:17-114), what SB3's VecVideoRecorder
 * records (train_test/sb3_train_vec.py:69): the draw list of :34-100 (target, 128 sensor rays, agent, "front" marker,
 * obstacles, path line) rasterised by a CUDA kernel into frames uint8 [n_frames][size][size][3] (row = canvas y,
 * column = canvas x, RGB: the np.transpose of :112-114), DEVICE memory, 4-byte aligned.  No window, no pygame: the
 * reference's human mode (:24-31, :102-110) is unavailable on this box.
 * index int64 [n_frames] DEVICE (env indices of this shard, each in [0, N)) or NULL for envs 0 .. n_frames - 1;
 * obs float32 [N][143] DEVICE: the observation the last step / reset returned (the ray lengths are read from its
 * columns 15..142, like self.sensor_data).  size: the reference's window_size (512).
 * USV_ENV_ASMC_SIMPLE and USV_ENV_SIMPLE only.  Coverage rules: gym_usv_b200/csrc/usv_render.cuh.  The frame of an
 * index outside [0, N) is left untouched. */
int usvgpu_render(usvgpu_handle h, const int64_t* index, int64_t n_frames, const float* obs, int32_t size,
                  uint8_t* frames, void* stream);

/* Replaces the `info` dict UsvSimpleEnv.step / .reset return (gym_usv/envs/simple_env.py:102-115 _get_info and the
 * reward_info of :189-199 merged at :341-342): out float64 [n_index][USVGPU_INFO_DIM] DEVICE, row k for env index[k]
 * (int64 DEVICE, or NULL for envs 0 .. n_index - 1), recomputed from the env's current state with the reference's
 * formulas -- nothing is stored per step.  Columns: UsvInfoCol.  'reward' is what usvgpu_step wrote; 'left_thruster' /
 * 'right_thruster' are the constant 0 of :112-113.  An env that finished in the last step has been reset already: its
 * row describes the new episode, like reset()'s info (:305).  USV_ENV_ASMC_SIMPLE and USV_ENV_SIMPLE only. */
#define USVGPU_INFO_DIM 24
typedef enum UsvInfoCol {
    USV_INFO_POSITION = 0,     /* x y psi                                  */
    USV_INFO_VELOCITY = 3,     /* u v r                                    */
    USV_INFO_PATH_START = 6, USV_INFO_PATH_END = 8,
    USV_INFO_ACTION0 = 10, USV_INFO_ACTION1 = 11,
    USV_INFO_YE = 12, USV_INFO_ANGLE_TO_TARGET = 13,
    USV_INFO_YE_REWARD = 14, USV_INFO_ANGLE_TO_TARGET_REWARD = 15, USV_INFO_ANGLE_ACTION_REWARD = 16,
    USV_INFO_DELTA_ACTION_REWARD = 17, USV_INFO_DELTA_ACTION = 18, USV_INFO_VELOCITY_TRACK_REWARD = 19,
    USV_INFO_REFERENCE_VELOCITY = 20, USV_INFO_REWARD_VELOCITY = 21, USV_INFO_REFERENCE_VELOCITY_ERROR = 22,
    USV_INFO_PROGRESS = 23     /* self.progress (simple_env.py:53), not part of the reference's dict */
} UsvInfoCol;
int usvgpu_step_info(usvgpu_handle h, const int64_t* index, int64_t n_index, double* out, void* stream);

/* obs[i][0..14] <- the reset observation _get_obs(zeros(3)) of simple_env.py:302 recomputed from the CURRENT state of
 * every env with mask[i] != 0 (mask uint8 [N] DEVICE or NULL for all): for a reset whose state the caller then edited
 * through usvgpu_set_state -- the reference's reset options (simple_env.py:276-288 'place_obstacles_on_path') and the
 * constructor's custom experiment (:292-300) are host-side edits of a fresh reset.  The 128 sensor columns are left as
 * they are (the reference shows stale sensor_data there).  USV_ENV_ASMC_SIMPLE and USV_ENV_SIMPLE only. */
int usvgpu_reset_obs(usvgpu_handle h, const uint8_t* mask, float* obs, void* stream);

/* UsvAsmcCaEnv.reset(options=...) (gym_usv/envs/usv_asmc_ca_env.py:358-372; tools/test_env.py:8-21 builds such a
 * dict): usvgpu_reset for USV_ENV_ASMC_CA_V0 with the option dict of the reference applied to every masked env, in
 * the reference's order -- the draws (:331-356), then the options replace the obstacle candidates / the target /
 * self.position, then the deletion rule (:374-399), then the internal step([-1, 0]) (:402).  As in the reference,
 * 'start_position' moves self.position for the deletion rule only: the model was built from the drawn pose (:336) and
 * step() reads the model (:199-204).  At most USVGPU_CA_MAX_OBSTACLES obstacles (what the observation model of this
 * build holds per env; tools/test_env.py's second layout, 78 circles, is refused by the host side). */
#define USVGPU_CA_MAX_OBSTACLES 9
typedef struct UsvCaResetOptions {
    int32_t has_obstacles;   /* 'obs_x' in options */
    int32_t n_obstacles;     /* <= USVGPU_CA_MAX_OBSTACLES */
    double obs_x[USVGPU_CA_MAX_OBSTACLES];
    double obs_y[USVGPU_CA_MAX_OBSTACLES];
    double obs_r[USVGPU_CA_MAX_OBSTACLES];
    int32_t has_target;      /* 'target_point' */
    int32_t has_start;       /* 'start_position' */
    double target[2];
    double start[2];
} UsvCaResetOptions;
int usvgpu_reset_ca_options(usvgpu_handle h, const uint8_t* mask, const UsvCaResetOptions* options, float* obs,
                            void* stream);

/* Replaces stable_baselines3 RolloutBuffer.compute_returns_and_advantage (what PPO.learn of
 * train_test/config.py:3-15 runs after every n_steps = 2048 rollout) on device-resident buffers: rewards, values,
 * episode_starts, advantages, returns float32 [n_steps][N]; last_values float32 [N]; last_dones uint8 [N].
 * gamma and gae_lambda are DOUBLES as SB3 holds them (Python floats): numpy rounds gamma to float32 against the
 * arrays, and the product gamma * gae_lambda is taken in float64 first and rounded once. */
int usvgpu_gae(int device, int32_t n_steps, int64_t n, const float* rewards, const float* values,
               const float* episode_starts, const float* last_values, const uint8_t* last_dones, double gamma,
               double gae_lambda, float* advantages, float* returns, void* stream);

/* Replaces stable_baselines3 OffPolicyAlgorithm._store_transition + ReplayBuffer.add for SAC
 * (train_test/sb3_train_vec.py:72 with config_sac, train_test/config.py:17-37: buffer_size 400 000) on device-resident
 * buffers laid out like SB3's: buf_obs / buf_next_obs float32 [capacity][N][obs_dim], buf_actions
 * [capacity][N][act_dim], buf_rewards [capacity][N], buf_dones / buf_timeouts uint8 [capacity][N].  Writes slot `pos`
 * of every env: obs (the observation the action was computed from), the next observation -- terminal_obs[i] for envs
 * with terminated | truncated set, as the VecEnv has already reset them (terminal_obs may be NULL: next_obs is stored
 * as is) --, the action, the reward, dones = terminated | truncated, timeouts = truncated & ~terminated
 * (infos[i]["TimeLimit.truncated"]).  obs / next_obs / terminal_obs are [N] rows with the given row strides in
 * elements (a UsvGpuVecFrameStack view has row stride 2 * n_stack * obs_dim).  All pointers DEVICE.
 * Split form for envs that overwrite their observation tensor in place: a first call with next_obs == NULL stores
 * only obs (before env.step), a second call with obs == NULL stores everything else (after it). */
int usvgpu_replay_add(int device, int64_t n, int32_t obs_dim, int32_t act_dim, int64_t capacity, int64_t pos,
                      const float* obs, int64_t obs_stride, const float* next_obs, int64_t next_stride,
                      const float* terminal_obs, int64_t terminal_stride, const float* actions, const float* rewards,
                      const uint8_t* terminated, const uint8_t* truncated, float* buf_obs, float* buf_next_obs,
                      float* buf_actions, float* buf_rewards, uint8_t* buf_dones, uint8_t* buf_timeouts, void* stream);

/* Replaces ReplayBuffer._get_samples (what SAC.train draws gradient_steps = 8 times per train_freq = 8 steps,
 * batch_size 256: train_test/config.py:21-25): sample k = (batch_inds[k], env_inds[k]), int64 DEVICE arrays drawn by
 * the caller (SB3: np.random.randint); out_obs / out_next_obs float32 [batch][obs_dim], out_actions [batch][act_dim],
 * out_rewards [batch], out_dones [batch] = dones * (1 - timeouts).  capacity: slots per env of the buffers; a sample whose
 * indices fall outside [0, capacity) x [0, N) comes out as zeros. */
int usvgpu_replay_sample(int device, int64_t n, int32_t obs_dim, int32_t act_dim, int64_t capacity, int64_t batch,
                         const int64_t* batch_inds, const int64_t* env_inds, const float* buf_obs,
                         const float* buf_next_obs, const float* buf_actions, const float* buf_rewards,
                         const uint8_t* buf_dones, const uint8_t* buf_timeouts, float* out_obs, float* out_next_obs,
                         float* out_actions, float* out_rewards, float* out_dones, void* stream);

/* Device-math probe (test / measurement surface, no reference counterpart): evaluates one of the float64
 * routines of csrc/usv_math.cuh that replace libm calls of the reference (np.sqrt, '/', np.math.asin at
 * usv_asmc.py:72,103,150; np.cos / np.sin at :179-181) on n inputs.  which: 0 sqrt_k(in[i]) | 1 div_k(in[i],
 * in[n+i]) | 2 asin_k(in[i]) | 3 rot_small(in[i]) applied to (cos, sin) = (in[n+i], in[2n+i]) -> out[i], out[n+i] |
 * 4 sqrt_m(in[i]), the square root the ASMC sub-step uses. */
int usvgpu_math_probe(int device, int32_t which, int64_t n, const double* in, double* out, void* stream);

/* Measured float64 FMA rate of this GPU (the pipe that bounds both step kernels): a grid of
 * `blocks` x 256 threads runs `iters` x 8 independent DFMA chains per thread; *dfma_per_sec receives
 * thread-level FMAs per second (CUDA events on `stream`; synchronises).  blocks < 0 measures the dependent-issue
 * latency instead: |blocks| CTAs of 128 threads, ONE dependent chain per thread (seconds per FMA of a chain =
 * |blocks| * 128 / *dfma_per_sec). */
int usvgpu_fp64_peak(int device, int32_t blocks, int32_t iters, double* dfma_per_sec, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* USVGPU_H */
