/* lt_b200.h -- C ABI of liblt_b200.so: the B200-native (sm_100a) per-environment step path of
 * Daffan/legged_tracking.
 *
 * The reference has no FFI layer: its boundary is the Python object protocol of
 * go1_gym/envs/base/legged_robot_trajectory_tracking.py (LR below), class LeggedRobot.  Every entry point
 * here cites the reference method(s) it replaces.  A maintainer binds these with ctypes from that class
 * (see INTEGRATION.md); legged_tracking_b200/_abi.py is exactly such a binding.
 *
 * Conventions
 *   - every call returns LT_OK (0) or a negative lt_status; nothing throws across the boundary;
 *   - all buffer pointers are DEVICE pointers owned by the caller (torch tensors in the Python host) unless a
 *     parameter is documented as host memory; the library allocates only private scratch, freed by lt_destroy;
 *   - every compute call is asynchronous on the cudaStream_t passed as `void* stream` (NULL = legacy default
 *     stream) and never synchronises the host, so a step (4 x lt_compute_torques + lt_post_physics_step) can be
 *     captured in a CUDA graph;
 *   - one handle per GPU / process rank; a handle is not thread-safe;
 *   - structs hold only 4-byte ints/floats (lt_config) or only pointers (lt_sim_ptrs, lt_state_ptrs, ...):
 *     no padding, trivially mirrored by ctypes.
 */
#ifndef LT_B200_H
#define LT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LT_ABI_VERSION 7

#define LT_MAX_DOF 12
#define LT_MAX_FEET 4
#define LT_MAX_BODIES 32
#define LT_MAX_POINTS 32
#define LT_MAX_TERMS 32
#define LT_ACT_HIDDEN 32
#define LT_ACT_IN 6

typedef enum {
    LT_OK = 0,
    LT_ERR_INVALID = -1,     /* bad argument / unsupported configuration (message in lt_last_error) */
    LT_ERR_UNBOUND = -2,     /* a required buffer was never bound */
    LT_ERR_CUDA = -3,        /* CUDA runtime error (message in lt_last_error) */
    LT_ERR_NO_DEVICE = -4    /* no sm_100 device: there is NO CPU fallback */
} lt_status;

/* reward term ids: the `_reward_<name>` plugin methods of
 * go1_gym/envs/rewards/reward_crawling.py:18-123 and trajectory_tracking_reward.py:17-171 */
enum {
    LT_TERM_DOF_ACC = 0, LT_TERM_TORQUES, LT_TERM_DOF_POS_LIMITS, LT_TERM_COLLISION, LT_TERM_ACTION_RATE,
    LT_TERM_BASE_HEIGHT, LT_TERM_ANG_VEL_XY, LT_TERM_ORIENTATION, LT_TERM_LARGE_VEL, LT_TERM_E2E,
    LT_TERM_EXPLORATION_LIN, LT_TERM_EXPLORATION_YAW, LT_TERM_REACHING_Z, LT_TERM_REACHING_ROLL,
    LT_TERM_REACHING_PITCH, LT_TERM_DOF_VEL, LT_TERM_DOF_POS, LT_TERM_TASK_OLD, LT_TERM_REACH_GOAL,
    LT_TERM_REACH_GOAL_T, LT_TERM_REACH_GOAL_TT, LT_TERM_TASK, LT_TERM_EXPLORATION, LT_TERM_STALLING,
    LT_TERM_LINEAR_VEL, LT_TERM_LIN_VEL_Z, LT_TERM_FEET_AIR_TIME, LT_TERM_SURVIVE, LT_TERM_REACHING_LINEAR_VEL,
    LT_TERM_REACHING_YAW_ABS, LT_TERM_REACHING_YAW,
    LT_TERM_EXTERNAL = 31,   /* user plugin evaluated by the host in torch on the device: value read from extra_reward */
    LT_TERM_REACHING_LOCAL_GOAL = 32     /* trajectory_tracking_reward.py:95-97: plan_buf & replan (sampling-based planning only) */
};
enum { LT_FORM_EXP = 0, LT_FORM_L1 = 1, LT_FORM_L2 = 2, LT_FORM_PROD = 3 };
enum { LT_TRAJ_FIXED_TARGET = 0, LT_TRAJ_RANDOM_GOAL = 1, LT_TRAJ_RANDOM_TARGET = 2 };
enum { LT_CONTROL_ACTUATOR_NET = 0, LT_CONTROL_P = 1 };
/* per-term bucket for rew_buf_pos / rew_buf_neg (LR:332-335): decided by the sign of the ALL-ENV sum */
enum { LT_BUCKET_POS = 1, LT_BUCKET_NEG = -1, LT_BUCKET_BY_SUM = 0 };

/* Everything _parse_cfg (LR:1860-1878), _init_buffers (LR:1169-1327), _process_dof_props (LR:680-708) and
 * _prepare_reward_function (LR:1368-1418) derive from `Cfg`.  Only int32_t and float members. */
typedef struct lt_config {
    int32_t struct_bytes;            /* sizeof(lt_config): ABI check */
    int32_t abi_version;             /* LT_ABI_VERSION */
    /* sizes */
    int32_t num_envs, num_actors, num_bodies, num_dof, num_actions, num_obs, num_privileged_obs, obs_history_length;
    /* control (LR:957-996) */
    int32_t decimation, control_type, randomize_lag_timesteps, lag_timesteps;
    float dt, clip_actions, clip_observations, action_scale, hip_scale_reduction;
    float default_dof_pos[LT_MAX_DOF], torque_limits[LT_MAX_DOF], p_gains[LT_MAX_DOF], d_gains[LT_MAX_DOF];
    float dof_pos_lower[LT_MAX_DOF], dof_pos_upper[LT_MAX_DOF];
    /* body index sets (LR:1647-1664) */
    int32_t n_feet, feet_indices[LT_MAX_FEET];
    uint32_t penalised_mask, termination_mask;      /* bit b = rigid body b of the env */
    /* heights (LR:1902-1965) */
    int32_t plane, observe_heights, n_points_x, n_points_y, measure_front_half, camera_zero, tile_rows, tile_cols;
    float points_x[LT_MAX_POINTS], points_y[LT_MAX_POINTS], camera_offset[3], horizontal_scale, ceiling_height;
    /* goals (trajectory_function.py:14-93, LR:836-848) */
    int32_t traj_function, traj_length, num_interpolation, switch_upon_reach;
    float base_x, base_y, base_z, base_roll, base_pitch, base_yaw;
    float x_mean, x_range, y_mean, y_range, z_range, roll_range, pitch_range, yaw_range;
    float switch_dist, switch_yaw, switch_interval;
    /* termination (LR:198-216) */
    int32_t use_terminal_body_height, terminate_end_of_trajectory, use_terminal_body_rotation;
    float max_episode_length, terminal_body_height, T_reach;
    /* rewards (LR:320-355): evaluation order = term_ids[0..n_terms) */
    int32_t n_terms, term_ids[LT_MAX_TERMS], term_bucket[LT_MAX_TERMS];
    int32_t only_positive_rewards, only_positive_rewards_ji22_style, lin_vel_form;
    float term_scales[LT_MAX_TERMS];                /* already multiplied by dt (LR:1385) */
    float sigma_rew_neg, base_height_target, tracking_sigma_lin, tracking_sigma_ang, target_lin_vel, target_ang_vel;
    float lin_reaching_criterion, ang_reaching_criterion, small_vel_threshold, large_dist_threshold;
    /* observations (LR:357-590) */
    int32_t observe_command, timestep_in_obs, observe_two_prev_actions, observe_vel, observe_only_ang_vel;
    int32_t observe_only_lin_vel, observe_yaw, observe_contact_states, global_reference, add_noise;
    float obs_scale_lin_vel, obs_scale_ang_vel, obs_scale_dof_pos, obs_scale_dof_vel, obs_scale_height;
    int32_t priv_observe_friction, priv_observe_restitution, priv_observe_base_mass, priv_observe_com_displacement;
    int32_t priv_observe_motor_strength, priv_observe_motor_offset, priv_observe_body_height;
    int32_t priv_observe_body_velocity, priv_observe_gravity;
    float priv_scale[9], priv_shift[9];             /* get_scale_shift (math_utils.py:35-38), in the order above */
    /* reset (LR:218-296, 998-1052) and domain randomisation (LR:744-764) */
    int32_t custom_origins, randomize_motor_strength, randomize_motor_offset, randomize_Kp_factor;
    int32_t randomize_Kd_factor, rand_interval, send_timeouts;
    float base_init_state[13], x_init_range, y_init_range, x_init_offset, y_init_offset, yaw_init_range;
    float motor_strength_range[2], motor_offset_range[2], Kp_factor_range[2], Kd_factor_range[2];
    /* tuning: environments per warp of the fused step kernel (4, 8, 16 or 32; 0 = choose from num_envs) */
    int32_t envs_per_warp;
    /* seed of the in-kernel counter-based generator (used when no uniforms are injected) */
    uint32_t seed_lo, seed_hi;
} lt_config;

/* HOST arrays copied at lt_create. */
typedef struct lt_tables {
    const float* noise_scale_vec;    /* [num_obs]  _get_noise_scale_vec (LR:1086-1166) */
    const float* act_w0;             /* [32,6]  resources/actuator_nets/unitree_go1.pt, Linear 0 (LR:1307-1322) */
    const float* act_b0;             /* [32] */
    const float* act_w1;             /* [32,32] */
    const float* act_b1;             /* [32] */
    const float* act_w2;             /* [1,32] */
    const float* act_b2;             /* [1] */
} lt_tables;

/* Isaac Gym acquire_*_tensor layouts (LR:1173-1201).  Borrowed, never freed. */
typedef struct lt_sim_ptrs {
    float* root_states;              /* [num_envs*num_actors, 13]  pos3 quat(xyzw) linvel3 angvel3; robot = actor 0 */
    float* dof_state;                /* [num_envs*num_dof, 2]      (pos, vel) */
    const float* contact_forces;     /* [num_envs, num_bodies, 3] */
    const float* rigid_body_state;   /* [num_envs*num_bodies, 13]  (only read by host-side extras) */
} lt_sim_ptrs;

/* Terrain: the per-env copy env_height_samples[N,2,R,C] of LR:1827-1840 is replaced by the tile table it is
 * copied from plus a per-env tile index (same values, 26 MB instead of 1.7 GB at 65,536 envs). */
typedef struct lt_terrain_ptrs {
    const float* tile_heights;       /* [n_tiles, 2, tile_rows, tile_cols] metres; layer 0 ceiling, 1 floor; NULL on plane */
    const int32_t* env_tile;         /* [num_envs] */
    const float* env_terrain_origin; /* [num_envs, 3]  LR:1842-1847 */
    const float* env_origins;        /* [num_envs, 3]  LR:1808-1825 */
} lt_terrain_ptrs;

/* Struct-of-arrays env state: one row-major [num_envs, d] array per LeggedRobot attribute of the same name. */
typedef struct lt_state_ptrs {
    /* fp32 */
    float *actions, *last_actions, *last_last_actions;                                  /* [N,12] LR:71-72,148-149 */
    float *joint_pos_target, *last_joint_pos_target, *last_last_joint_pos_target;       /* [N,12] LR:150-151 */
    float *last_dof_vel, *last_root_vel, *torques;                                      /* [N,12] [N,6] [N,12] */
    float *joint_pos_err_last, *joint_pos_err_last_last, *joint_vel_last, *joint_vel_last_last;   /* [N,12] LR:985-988 */
    float *motor_strengths, *motor_offsets, *Kp_factors, *Kd_factors;                   /* [N,12] LR:744-764 */
    float *friction_coeffs, *restitutions, *payloads, *com_displacements, *gravities;   /* [N] [N] [N] [N,3] [N,3] */
    float *gravity_vec;                                                                 /* [3] (LR:1221 keeps N equal rows) */
    float *base_lin_vel, *base_ang_vel, *projected_gravity, *base_rotation;             /* [N,3] */
    float *relative_linear, *relative_rotation, *local_relative_linear, *local_relative_rotation;   /* [N,3] */
    float *target_poses, *feet_air_time;                                                /* [N,6] [N,n_feet] */
    float *rew_buf, *rew_buf_pos, *rew_buf_neg, *camera_pitch_angle;                    /* [N] */
    float *trajectories;                                                                /* [N, traj_length, 6] */
    float *lag_buffer;                                                                  /* [lag_timesteps+1, N, 12] ring, see lt_get_counters */
    float *episode_sums;                                                                /* [n_terms+3, N]  rows: terms, total, total_pos, total_neg */
    float *command_sums;                                                                /* [n_terms+1, N] or NULL: terms, ep_timesteps */
    float *obs_buf, *privileged_obs_buf, *obs_history;                                  /* [N,num_obs] [N,num_priv] [N,H*num_obs] */
    float *finished_sums, *finished_episode_length, *finished_goal_distance;            /* [n_terms+3,N] [N] [N]: episode record of the env's last reset (LR:256-273) */
    /* int64 (torch.long attributes) */
    int64_t *episode_length_buf, *curr_pose_index, *collision_count, *collision_env_buf;
    int64_t *reset_env_ids;                                                             /* [N] out: reset_buf.nonzero() ascending (LR:143) */
    /* bool (1 byte, torch.bool attributes) */
    uint8_t *reset_buf, *time_out_buf, *body_height_buf, *switched_buf, *reached_buf, *plan_buf;
    uint8_t *reached_env_buf, *time_outs_latched, *last_contacts;                       /* last_contacts [N,n_feet] */
    /* int32 */
    int32_t *reset_count;                                                               /* [1] out: len(env_ids) */
    /* optional debug/parity exports (NULL = skip) */
    int32_t *height_px, *height_py;                                                     /* [N,n_points_x] [N,n_points_y] cell indices (LR:1955-1956) */
} lt_state_ptrs;

/* Per-step inputs of lt_post_physics_step. */
typedef struct lt_step_args {
    const float* extra_reward;       /* [n_external, N] or NULL: unscaled values of the LT_TERM_EXTERNAL plugin terms, row k = k-th such term slot */
    const float* uniforms;           /* [N, lt_num_uniform_slots] in [0,1) or NULL -> in-kernel Philox4x32-10 */
    const float* noise_uniforms;     /* [N, num_obs] or NULL -> in-kernel Philox (only read if add_noise) */
    const float* new_gravity;        /* HOST float[3] or NULL: _randomize_gravity fired this step (LR:645-660): written
                                        to `gravities` before the observations are assembled */
    /* sampling-based planning (LR:856-916): this step's local target and its relative pose, as lt_plan_target_pose left them
     * ([N,6], [N,3], [N,3]); all three NULL = the local target is the global one (LR:917-920) */
    const float *local_target_poses, *local_relative_linear, *local_relative_rotation;
    const uint8_t* replan;           /* [N] or NULL: this step's `replan` flags (LR:861) as lt_plan_target_pose wrote them; read by
                                        LT_TERM_REACHING_LOCAL_GOAL only */
} lt_step_args;

typedef struct lt_handle lt_handle;

/* -------------------------------------------------------------------------------------------------------------
 * life cycle */
int lt_abi_version(void);
int lt_config_bytes(void);                                  /* sizeof(lt_config) as compiled */
int lt_num_uniform_slots(const lt_config* cfg);             /* row width of lt_step_args.uniforms */
int lt_create(const lt_config* cfg, const lt_tables* tables, lt_handle** out);      /* LR:25-62 (__init__) */
void lt_destroy(lt_handle* h);
const char* lt_last_error(const lt_handle* h);              /* h may be NULL: error of the last failed lt_create */

/* LR:1173-1201 (gym.acquire_*_tensor + gymtorch.wrap_tensor) */
int lt_bind_sim(lt_handle* h, const lt_sim_ptrs* sim);
/* LR:1808-1847 (_get_env_origins) */
int lt_bind_terrain(lt_handle* h, const lt_terrain_ptrs* terrain, int32_t n_tiles);
/* LR:1203-1305 (_init_buffers: the torch.zeros(...) attributes) */
int lt_bind_state(lt_handle* h, const lt_state_ptrs* state);

/* -------------------------------------------------------------------------------------------------------------
 * the step path */

/* LeggedRobot._compute_torques (LR:957-996) incl. eval_actuator_network (LR:1311-1320), called once per physics
 * sub-step from LeggedRobot.step (LR:82-88).  `actions` is the RAW policy output [N,12]; clipping to
 * +-clip_actions (LR:71-72) is fused here and, when substep == 0, stored to state.actions.  Reads sim.dof_state,
 * advances the lag FIFO and the actuator-net history, writes state.joint_pos_target and state.torques.  The call with
 * substep == decimation - 1 also performs the joint-target history shift of post_physics_step (LR:151-152:
 * last_last_joint_pos_target <- last_joint_pos_target <- joint_pos_target), so a step must issue sub-steps 0..decimation-1. */
int lt_compute_torques(lt_handle* h, const float* actions, int32_t substep, void* stream);

/* LeggedRobot.post_physics_step (LR:114-169) with everything it calls: _post_physics_step_callback (LR:774-848),
 * _get_heights (LR:1918-1970), _plan_target_pose / _compute_relative_target_pose (LR:850-932),
 * check_termination (LR:198-216), compute_reward (LR:320-355), reset_idx (LR:218-296, 998-1052),
 * compute_observations (LR:357-590), the history shifts (LR:148-153), the +-clip_observations clamp of
 * LeggedRobot.step (LR:108-111) and HistoryWrapper.step's obs_history shift (history_wrapper.py:18-24).
 * Three launches (+ the sign pre-pass when a term needs it): lt_step_kernel (per-env scalar logic), lt_obs_kernel
 * (heightfield gathers, noise, observation rows) and the ordered reset-id compaction (reset_buf.nonzero(), LR:143). */
int lt_post_physics_step(lt_handle* h, const lt_step_args* args, void* stream);

/* LeggedRobot.reset_idx(env_ids) (LR:218-296) called from OUTSIDE a step, e.g. by reset() (trajectory_tracking/__init__.py:46-56):
 * re-spawns the listed envs exactly as the step kernel does for envs whose reset_buf fired -- DR resample, dof / root state
 * written into the bound simulator tensors, new trajectory, histories / lag ring / episode sums zeroed, finished-episode
 * record filled.  env_ids: DEVICE int64[count]; count_dev: optional DEVICE int32 holding the live count (<= count) so that
 * a compacted id list can be used without a host read; uniforms: [N, lt_num_uniform_slots] or NULL (Philox). */
int lt_reset_idx(lt_handle* h, const int64_t* env_ids, int32_t count, const int32_t* count_dev, const float* uniforms,
                 void* stream);

/* Only when a term has LT_BUCKET_BY_SUM: evaluates those terms without side effects and leaves their all-env
 * sums in the device array returned by lt_term_sum_buffer (float[LT_MAX_TERMS]); a multi-GPU host all-reduces
 * that array before lt_post_physics_step.  lt_post_physics_step runs it itself when the host did not. */
int lt_reward_sign_prepass(lt_handle* h, const lt_step_args* args, void* stream);
float* lt_term_sum_buffer(lt_handle* h);

/* LeggedRobot.update_curriculum (LR:171-183) decays the exploration_lin / exploration_yaw scales on the host once
 * common_step_counter exceeds rewards.exploration_steps: `scales` (HOST float[n_terms], same order as
 * lt_config.term_scales, already multiplied by dt) replaces the term scales from the next launch on.  A CUDA graph
 * captured earlier keeps the scales it was captured with. */
int lt_set_term_scales(lt_handle* h, const float* scales, int32_t n);

/* Device int32[4]: {policy steps done, physics sub-steps done (= pushes into the lag FIFO), lag ring head, value of [0] the
 * last lt_post_physics_step started with}.
 * Logical lag slot j of LR:972-974 lives in physical slot (head + j) % (lag_timesteps + 1). */
int32_t* lt_counters(lt_handle* h);

/* -------------------------------------------------------------------------------------------------------------
 * rollout path: go1_gym_learn/ppo_cse/rollout_storage.py */

/* RolloutStorage.add_transitions (rollout_storage.py:57-71): one fused copy of a transition into row `t` of the
 * [T,N,*] buffers.  Any source may be NULL (skipped). */
typedef struct lt_rollout_ptrs {
    float *observations, *privileged_observations, *observation_histories, *actions, *rewards, *values;
    float *actions_log_prob, *mu, *sigma, *returns, *advantages;
    uint8_t* dones;
} lt_rollout_ptrs;
typedef struct lt_transition_ptrs {
    const float *observations, *privileged_observations, *observation_histories, *actions, *rewards, *values;
    const float *actions_log_prob, *action_mean, *action_sigma;
    const uint8_t *dones, *time_outs;    /* time_outs: extras["time_outs"]; bootstrap rewards += gamma*values*time_outs (ppo.py:85-87) */
} lt_transition_ptrs;
int lt_rollout_store(const lt_rollout_ptrs* storage, const lt_transition_ptrs* tr, int32_t t, int32_t num_envs,
                     int32_t num_obs, int32_t num_privileged_obs, int32_t num_obs_history, int32_t num_actions,
                     float gamma, void* stream);

/* RolloutStorage.compute_returns (rollout_storage.py:76-90): GAE over T steps for N envs.  Writes returns and
 * raw advantages, and accumulates {sum A, sum A^2, count} into stats (device double[3], zeroed by the call) so that
 * a multi-GPU host can all-reduce them; lt_normalize_advantages then applies (A - mean) / (std + 1e-8) with the
 * unbiased std of torch.Tensor.std(). */
int lt_gae(const float* rewards, const float* values, const uint8_t* dones, const float* last_values,
           float* returns, float* advantages, double* stats, int32_t T, int32_t num_envs, float gamma, float lam,
           void* stream);
int lt_normalize_advantages(float* advantages, const double* stats, int64_t count, void* stream);

/* -------------------------------------------------------------------------------------------------------------
 * terrain: device-side tile-table builder for terrain_type = single_path.
 * Replaces Terrain.__init__ / make_terrain / add_terrain_to_map (go1_gym/utils/tunnel.py:51-217),
 * vec_plane_from_points (go1_gym/utils/tunnel_fn.py:3-21), TerrainFunctions.single_path (tunnel_fn.py:99-163) and the
 * float32 conversion of LeggedRobot._get_env_origins (legged_robot_trajectory_tracking.py:1827-1840).  The random draws are
 * made by the caller (numpy generator, reference order); all arguments are DEVICE pointers:
 *   n_pyramids [tiles,2] int32 (1 or 2 wedges per layer; layer 0 ceiling, 1 floor)
 *   wedges     [tiles,2,2,5] float64 (mean_x, mean_y, mean_z, pw, pl) per wedge
 *   window     [tiles,2] int32 tile-local (row, col) of the sub-terrain patch
 *   xs [sub_rows], ys [sub_cols] float64 sample coordinates (np.linspace tables of tunnel_fn.py:152)
 *   consts     float64[5] = {vertical_scale, ceiling_height / vertical_scale, 0.05 / vertical_scale, border value of the
 *              ceiling layer, border value of the floor layer} (the borders in units of vertical_scale, tunnel.py:81-82)
 *   tile_heights [tiles,2,tile_rows,tile_cols] float32 out, metres.  Bit-identical to the reference's table. */
int lt_build_tunnel_tiles(const int32_t* n_pyramids, const double* wedges, const int32_t* window, const double* xs,
                          const double* ys, const double* consts, float* tile_heights, int32_t n_tiles, int32_t tile_rows,
                          int32_t tile_cols, int32_t sub_rows, int32_t sub_cols, void* stream);

/* _get_heights over the FULL measured_points grid (LR:1918-1965; the observation path gathers the front half only): out
 * [N, 2, n_points_x * n_points_y] ceiling / floor heights in metres from the bound terrain, the robot root rows of the bound
 * simulator tensors and the PREVIOUS step's pitch (state base_rotation, LR:1938: the camera offset uses the not yet refreshed
 * rotation).  Input of lt_plan_target_pose; call it before lt_post_physics_step of the same step. */
int lt_measure_heights(lt_handle* h, float* out, void* stream);

/* sampling-based local planner, LeggedRobot._plan_target_pose with cfg.commands.sampling_based_planning
 * (legged_robot_trajectory_tracking.py:856-916) and the relative pose of the resulting local target (:914-916, 922-932).
 * The reference runs a Python loop over the planning envs; this is one CTA per env, threads over the candidate poses.
 *   target_poses      [N,6]  world-frame goal = trajectories[env, curr_pose_index]
 *   root_states       robot root row of env e at root_states + e * root_stride (Isaac Gym layout: pos3, quat xyzw4, ...)
 *   points_xy         [n_points,2] base-frame sample offsets (_init_height_points, :1902-1916; 21 x 11 = 231)
 *   measured_heights  [N,2,n_points] ceiling / floor heights of _get_heights (:1918-1965) from the previous step
 *   candidates        [n_cands,6] cfg.commands.candidate_target_poses as float32 (config :167-175; 1,125 poses)
 *   episode_length_buf, plan_buf [N]: inputs of the replan rule (:858-868); plan_length_buf [N] is advanced / cleared in place
 *   local_target_poses [N,6] in / out (envs that do not plan keep theirs); local_relative_linear / rotation [N,3] out (may be
 *   NULL); chosen [N] out (may be NULL): winning candidate index, -1 no valid candidate (global goal used), -2 close to the goal,
 *   -3 no plan this step.  Ties between equal sort keys go to the lower candidate index (a stable argsort). */
typedef struct lt_plan_args {
    const float *target_poses, *root_states, *points_xy, *measured_heights, *candidates;
    const int64_t* episode_length_buf;
    const uint8_t* plan_buf;
    int64_t* plan_length_buf;
    float *local_target_poses, *local_relative_linear, *local_relative_rotation;
    int32_t* chosen;
    uint8_t* replan;                 /* [N] out or NULL: (plan_length_buf % plan_interval) == 0 after the increment (LR:858-861) */
} lt_plan_args;
int lt_plan_target_pose(const lt_plan_args* args, int32_t num_envs, int32_t n_points, int32_t n_cands, int32_t plan_interval,
                        int32_t root_stride, const float* robot_size, void* stream);

/* -------------------------------------------------------------------------------------------------------------
 * velocity-tracking specifics (LRV = go1_gym/envs/base/legged_robot_velocity_tracking.py,
 * CR = go1_gym/envs/rewards/corl_rewards.py) */

/* LeggedRobot._step_contact_targets (LRV:844-923): gait phase advance, per-foot phase, duration-warped clock inputs
 * (sin of 2 pi / 4 pi / pi x warped phase) and the Normal-CDF smoothed desired contact states.  commands [N, num_commands]
 * (columns 4..8 = frequency, phase, offset, bound, duration); gait_indices [N] in / out; the other five [N,4] out. */
typedef struct lt_gait_ptrs {
    const float* commands;
    float *gait_indices, *foot_indices, *clock_inputs, *doubletime_clock_inputs, *halftime_clock_inputs, *desired_contact_states;
} lt_gait_ptrs;
int lt_step_contact_targets(const lt_gait_ptrs* ptrs, int32_t num_envs, int32_t num_commands, float dt, float kappa,
                            int32_t pacing_offset, void* stream);

/* CoRLRewards._reward_<name> (CR:15-201): ids in the order the container defines them */
#define LT_CORL_TRACKING_LIN_VEL 0
#define LT_CORL_TRACKING_ANG_VEL 1
#define LT_CORL_LIN_VEL_Z 2
#define LT_CORL_ANG_VEL_XY 3
#define LT_CORL_ORIENTATION 4
#define LT_CORL_TORQUES 5
#define LT_CORL_DOF_ACC 6
#define LT_CORL_ACTION_RATE 7
#define LT_CORL_COLLISION 8
#define LT_CORL_DOF_POS_LIMITS 9
#define LT_CORL_JUMP 10
#define LT_CORL_TRACKING_CONTACTS_SHAPED_FORCE 11
#define LT_CORL_TRACKING_CONTACTS_SHAPED_VEL 12
#define LT_CORL_DOF_POS 13
#define LT_CORL_DOF_VEL 14
#define LT_CORL_ACTION_SMOOTHNESS_1 15
#define LT_CORL_ACTION_SMOOTHNESS_2 16
#define LT_CORL_FEET_SLIP 17
#define LT_CORL_FEET_CONTACT_VEL 18
#define LT_CORL_FEET_CONTACT_FORCES 19
#define LT_CORL_FEET_CLEARANCE_CMD_LINEAR 20
#define LT_CORL_FEET_IMPACT_VEL 21
#define LT_CORL_ORIENTATION_CONTROL 22
#define LT_CORL_RAIBERT_HEURISTIC 23
#define LT_CORL_MAX_TERMS 32
typedef struct lt_corl_cfg {
    int32_t num_envs, num_commands, num_bodies, n_terms, n_penalised;
    int32_t term_ids[LT_CORL_MAX_TERMS];              /* LT_CORL_*: row t of `out` is term term_ids[t] */
    int32_t feet_indices[4], penalised[LT_CORL_MAX_TERMS];
    float dt, tracking_sigma, tracking_sigma_yaw, base_height_target, gait_force_sigma, gait_vel_sigma, max_contact_force;   /* cfg.rewards */
    int32_t dof_stride;                               /* element stride of dof_pos / dof_vel: 0 or 1 = [N,12] rows; 2 = the columns of Isaac
                                                         Gym's interleaved dof_state [N,12,2] (dof_pos = dof_state, dof_vel = dof_state + 1) */
} lt_corl_cfg;
/* env attributes the container reads (CR; shapes per env): commands [num_commands], base_lin_vel / base_ang_vel /
 * projected_gravity / gravity_vec / base_pos [3], base_quat [4], the [12] dof / action / joint-target rows, contact_forces
 * [num_bodies,3], foot_positions / foot_velocities / prev_foot_velocities [4,3], foot_indices / desired_contact_states [4];
 * default_dof_pos [12] and dof_pos_limits [12,2] are shared by all envs.  last_contacts [N,4] (bool) is read AND written by
 * feet_slip (CR:107-113).  out [n_terms, N]: unscaled term values, the layout of lt_step_args.extra_reward. */
typedef struct lt_corl_ptrs {
    const float *commands, *base_lin_vel, *base_ang_vel, *projected_gravity, *gravity_vec, *base_pos, *base_quat;
    const float *torques, *dof_pos, *dof_vel, *last_dof_vel, *actions, *last_actions, *last_last_actions;
    const float *joint_pos_target, *last_joint_pos_target, *last_last_joint_pos_target, *contact_forces;
    const float *foot_positions, *foot_velocities, *prev_foot_velocities, *foot_indices, *desired_contact_states;
    const float *default_dof_pos, *dof_pos_limits;
    uint8_t* last_contacts;
    float* out;
} lt_corl_ptrs;
int lt_corl_rewards(const lt_corl_cfg* cfg, const lt_corl_ptrs* ptrs, void* stream);

/* LeggedRobot._resample_commands (LRV:728-842), the device half.  The curricula (RewardThresholdCurriculum, curriculum.py:113-159:
 * numpy weight grids, updated from the finished episodes' command_sums and sampled with numpy's RandomState) stay on the host, as
 * in the reference; what they return -- one sampled command row and one category per resampled env -- comes in here and the call
 * does everything the reference then does to the env tensors, in ONE launch instead of ~40 indexed torch ops:
 *   commands[env] = new row (LRV:776-777); the gait-category rewrite of the phase / offset / bound columns 5..7 -- by the env's
 *   curriculum category when gaitwise_curricula (LRV:780-797), else by a uniform draw per env for exclusive_phase_offset
 *   (LRV:799-809) or balance_gait_distribution (LRV:811-828); binary_phases rounding round(2c)/2 % 1 (LRV:830-833); zeroing of
 *   the planar velocity command when its norm is <= 0.2 (LRV:836); command_sums[:, env] = 0 (LRV:839-840). */
#define LT_GAIT_PRONK 0              /* category_gait values: the reference's category_names (LRV:1319-1321) */
#define LT_GAIT_TROT 1
#define LT_GAIT_PACE 2
#define LT_GAIT_BOUND 3
#define LT_GAIT_NONE -1              /* 'nominal': no rewrite */
#define LT_CMD_MAX_CATEGORIES 8
typedef struct lt_cmd_cfg {
    int32_t num_envs, num_commands, n_sums;
    int32_t gaitwise_curricula, exclusive_phase_offset, balance_gait_distribution, binary_phases;     /* cfg.commands.* */
    int32_t n_categories, category_gait[LT_CMD_MAX_CATEGORIES];
} lt_cmd_cfg;
typedef struct lt_cmd_ptrs {
    const int64_t* env_ids;          /* [count] envs being resampled */
    const float* new_commands;       /* [count, num_commands]: curriculum.sample()[0][:, :num_commands], row k for env_ids[k] */
    const int32_t* category;         /* [count] curriculum category of each env (LRV:760-766) */
    const float* gait_uniform;       /* [count] the second torch.rand draw (LRV:800, 812) or NULL when gaitwise_curricula */
    float* commands;                 /* [N, num_commands] in / out */
    float* command_sums;             /* [n_sums, N] or NULL: rows zeroed at env_ids */
} lt_cmd_ptrs;
int lt_resample_commands(const lt_cmd_cfg* cfg, const lt_cmd_ptrs* ptrs, int32_t count, void* stream);

/* -------------------------------------------------------------------------------------------------------------
 * cfg 4: LeggedRobot.post_physics_step of the velocity-tracking env (LRV:108-154) as a chain of launches around the three
 * calls above.  Order of one step (legged_tracking_b200/vel_env.py is the host mirror):
 *   lt_vel_pre                episode_length_buf += 1, body-frame velocities, projected gravity, foot positions / velocities,
 *                             check_termination (LRV:156-167), the callback's two interval masks (LRV:702-703, 715-716)
 *   [host]                    _resample_commands of the due envs: curricula on the host, lt_resample_commands
 *   lt_step_contact_targets   LRV:705
 *   lt_corl_rewards           every active CoRLRewards term, unscaled, [n_terms, N]
 *   lt_vel_reward             compute_reward (LRV:281-318): scales, all-env sign buckets, episode / command sums, ji22 or clip
 *   [host]                    reset_idx of reset_buf.nonzero() (LRV:169-265): rare path, indexed writes
 *   lt_vel_observe            compute_observations (LRV:320-470) + noise, the last_* shifts (LRV:143-148), the obs clip of
 *                             step() (LRV:102-105) and HistoryWrapper's shift-and-append (history_wrapper.py:23) */
typedef struct lt_vel_cfg {
    int32_t num_envs, num_bodies, num_commands, num_obs, num_privileged_obs, obs_history_length;
    int32_t feet_indices[4];
    int32_t n_termination, termination_indices[8];            /* termination_contact_indices (LRV:158) */
    int32_t use_terminal_body_height;
    float terminal_body_height, max_episode_length;
    int32_t sample_interval, rand_interval;                   /* int(resampling_time / dt), int(domain_rand.rand_interval) */
    int32_t decimation, lag_slots;                            /* for torque_counters (below) */
    int32_t n_terms;                                          /* rows of the lt_corl_rewards block, in reward_names order */
    int32_t command_sum_adds_scale[LT_CORL_MAX_TERMS];        /* 1 for tracking_contacts_shaped_force / _vel (LRV:296-297) */
    float term_scales[LT_CORL_MAX_TERMS];                     /* reward_scales[name] * dt (LRV:1243) */
    int32_t only_positive_rewards, only_positive_rewards_ji22_style;
    float sigma_rew_neg;
    int32_t observe_command, observe_two_prev_actions, observe_timing_parameter, observe_clock_inputs, add_noise;
    float clip_observations, obs_scale_dof_pos, obs_scale_dof_vel;
    int32_t priv_observe_friction, priv_observe_restitution;
    float friction_scale, friction_shift, restitution_scale, restitution_shift;     /* get_scale_shift of cfg.normalization.* */
    uint32_t seed_lo, seed_hi;                                /* Philox key of the in-kernel observation noise */
} lt_vel_cfg;
typedef struct lt_vel_ptrs {
    /* simulator tensors (LRV:1124-1150): root_states [N,13], rigid_body_state [N,B,13], contact_forces [N,B,3], dof_state [N,12,2] */
    const float *root_states, *rigid_body_state, *contact_forces, *dof_state;
    const float *gravity_vec;                                 /* [N,3] */
    const float *measured_height_mean;                        /* [N]: mean over the height points of root_z - measured height
                                                                 (lt_vel_measure_heights' height_mean), or NULL: plane, root_z itself */
    int64_t* episode_length_buf;                              /* [N] in / out */
    float *base_pos, *base_quat, *base_lin_vel, *base_ang_vel, *projected_gravity;        /* [N,3] [N,4] [N,3] x3 */
    float *foot_positions, *foot_velocities, *prev_foot_velocities;                       /* [N,4,3] */
    uint8_t *reset_buf, *time_out_buf, *body_height_buf, *resample_due, *dr_due;          /* [N] */
    /* rewards */
    const float* commands;                                    /* [N, num_commands] */
    float *rew_buf, *rew_buf_pos, *rew_buf_neg;               /* [N] */
    float* episode_sums;                                      /* [n_terms + 1, N]: terms, total */
    float* command_sums;                                      /* [n_terms + 5, N]: terms, lin_vel_raw, ang_vel_raw, lin_vel_residual, ang_vel_residual, ep_timesteps */
    double* term_sums;                                        /* [n_terms] scratch: all-env sums of the scaled terms */
    /* observations */
    const float *commands_scale, *default_dof_pos, *noise_scale_vec;      /* [num_commands] [12] [num_obs] */
    const float *actions, *joint_pos_target, *gait_indices, *clock_inputs, *friction_coeffs, *restitutions;   /* [N,12] [N,12] [N] [N,4] [N] [N] */
    float *last_actions, *last_last_actions, *last_joint_pos_target, *last_last_joint_pos_target, *last_dof_vel, *last_root_vel;
    float *obs_buf, *privileged_obs_buf;                      /* [N,num_obs] [N,num_privileged_obs] */
    const float* obs_history_in;                              /* [N, H * num_obs] */
    float* obs_history_out;                                   /* [N, H * num_obs], a different buffer: the host ping-pongs the two */
    int32_t* torque_counters;                                 /* lt_counters() of the handle whose lt_compute_torques ran this step, or NULL:
                                                                 lt_vel_pre advances them as lt_post_physics_step would ([0] += 1,
                                                                 [1] += decimation, [2] = ([2] + decimation) % lag_slots: the lag ring's head) */
    uint32_t* step_counter;                                   /* device uint32[1] or NULL.  Non-NULL: lt_vel_indexed_paths and lt_vel_observe key
                                                                 their Philox draws with *step_counter instead of their `step` argument, and
                                                                 lt_vel_observe adds 1 to it when it is done -- so that a CUDA graph of the
                                                                 step can be replayed (a by-value step would be frozen at capture) */
} lt_vel_ptrs;
/* LeggedRobot._get_heights of the velocity env (LRV:1790-1824) for heightfield / trimesh terrains: the P grid points
 * (height_points [P,3], base frame, the same for every env: _init_height_points LRV:1773-1788) rotated by the base yaw and
 * offset by the base position; cell = trunc((p + border_size) / horizontal_scale) clipped to the field; height = the MINIMUM of
 * the cells (px, py), (px + 1, py), (px, py + 1) of height_samples (int16 [rows, cols]) x vertical_scale.
 * measured_heights [N,P] out; height_mean [N] out or NULL: mean over the points of root_z - height, what check_termination
 * compares with terminal_body_height (LRV:163-165) -- pass it as lt_vel_ptrs.measured_height_mean. */
int lt_vel_measure_heights(const float* root_states, const float* height_points, const int16_t* height_samples, int32_t num_envs,
                           int32_t n_points, int32_t rows, int32_t cols, float border_size, float horizontal_scale, float vertical_scale,
                           float* measured_heights, float* height_mean, void* stream);
int lt_vel_pre(const lt_vel_cfg* cfg, const lt_vel_ptrs* ptrs, void* stream);

/* The three indexed paths of the cfg 4 step -- _resample_commands of the due envs and the interval _randomize_dof_props (the
 * callback, LRV:701-717: which = LT_VEL_PATH_CALLBACK) and reset_idx of reset_buf (LRV:169-265: which = LT_VEL_PATH_RESET) -- without
 * the host: mask-driven over all envs (no nonzero(), no id list), with the command curriculum itself on the device.
 *   curriculum (go1_gym/envs/base/curriculum.py:113-152 RewardThresholdCurriculum, one weight row per command category): update =
 *   every due env whose four task rewards (command_sums / ep_len) beat their thresholds bumps its own bin once and every bin within
 *   local_range of it once per env, +0.2 clipped to 1 (same result as the reference's sequential numpy updates: the increments
 *   are equal and the clip is monotone); sample = a bin drawn in proportion to the weights (Curriculum.sample_bins), a point uniform
 *   inside the bin (sample_uniform_from_cell).  The draws are Philox4x32-10 keyed (seed, step, env, slot), or injected through
 *   uniform_rows [N, LT_VEL_SLOTS] (tests): they replace numpy's RandomState, so samples are equal in distribution, not bit-wise.
 *   reset: _randomize_dof_props (LRV:663-673), _reset_dofs (:966-976), _reset_root_states (:983-1015, flat origins), the buffers
 *   (:192-198, 258-261); finished episodes' sums are accumulated into episode_acc [n_terms + 2] (sum of every episode_sums row over
 *   the reset envs, then their count) for the extras the reference averages on the host (:200-205). */
#define LT_VEL_PATH_CALLBACK 0
#define LT_VEL_PATH_RESET 1
#define LT_VEL_SLOTS 52              /* uniform slots per env and path: 0 category, 1 bin, 2..16 cell, 17 gait, 18 strength, 19..30 offsets,
                                        31..42 dof positions, 43 yaw, 44..49 root velocity, 50 / 51 x / y init offset */
#define LT_VEL_MAX_DIMS 16
typedef struct lt_vel_dev_cfg {
    int32_t n_bins, n_dims, n_keys;                  /* curriculum bins per category, command dims of the grid (15), task-reward keys (<= 4) */
    int32_t key_rows[4];                             /* command_sums rows of tracking_lin_vel, tracking_ang_vel, tracking_contacts_shaped_force / _vel */
    float key_thresholds[4];                         /* curriculum_thresholds[key] * reward_scales[key] (LRV:750) */
    float ep_len;                                    /* min(max_episode_length, int(resampling_time / dt)) (LRV:732-733) */
    double local_range[LT_VEL_MAX_DIMS];             /* LRV:755-757 */
    int32_t randomize_motor_strength, randomize_motor_offset, lag_slots;
    int32_t custom_origins;                          /* heightfield / trimesh terrains: x / y init offsets drawn at reset (LRV:991-1001, slots 50, 51) */
    float x_init_range, y_init_range, x_init_offset, y_init_offset;
    float motor_strength_lo, motor_strength_span, motor_offset_lo, motor_offset_span, yaw_init_range;   /* span = (float)(high - low), the
                                                        difference taken in double as python does before it meets the float32 tensor */
    float base_init_state[13];
    lt_cmd_cfg cmd;                                  /* the gait switches and category table of lt_resample_commands */
} lt_vel_dev_cfg;
typedef struct lt_vel_dev_ptrs {
    const double *grid, *bin_sizes;                  /* [n_dims, n_bins] bin centres, [n_dims] */
    double* weights;                                 /* [n_categories, n_bins] */
    int32_t *hit_self, *hit_adj;                     /* [n_categories, n_bins] scratch, zero between calls */
    int32_t *env_bins, *env_categories;              /* [N] */
    float *commands;                                 /* [N, num_commands] (lt_vel_ptrs.commands, writable) */
    float *dof_state_rw, *root_states_rw;            /* [N,12,2], [N,13]: the simulator rows a reset rewrites */
    const float* env_origins;                        /* [N,3] */
    float *motor_strengths, *motor_offsets;          /* [N,12] each: the torque kernel's tables */
    float *lag_buffer;                               /* [lag_slots, N, 12] */
    float *gait_indices, *feet_air_time;             /* [N], [N,4] */
    float* episode_acc;                              /* [n_terms + 2] */
    int32_t *work_env, *work_code, *work_count;      /* [N], [N], [3] scratch: the envs with indexed work this call, compacted;
                                                        work_count = {collecting, published length, block ticket}, ZERO before the
                                                        first call (the kernels leave [0] and [2] zero again) */
} lt_vel_dev_ptrs;
/* sizeof of the cfg 4 structs as compiled (0 lt_vel_cfg, 1 lt_vel_ptrs, 2 lt_vel_dev_cfg, 3 lt_vel_dev_ptrs, 4 lt_cmd_cfg): binding check */
int lt_vel_struct_bytes(int32_t which);
int lt_vel_indexed_paths(const lt_vel_cfg* cfg, const lt_vel_ptrs* ptrs, const lt_vel_dev_cfg* dcfg, const lt_vel_dev_ptrs* dptrs, int32_t which,
                         const float* uniform_rows, uint32_t step, void* stream);
/* term_values: the [n_terms, N] block lt_corl_rewards wrote */
int lt_vel_reward(const lt_vel_cfg* cfg, const lt_vel_ptrs* ptrs, const float* term_values, void* stream);
/* noise_uniforms [N, num_obs] in [0,1) or NULL -> Philox4x32-10 keyed (seed, step, env, column / 4); read only if add_noise */
int lt_vel_observe(const lt_vel_cfg* cfg, const lt_vel_ptrs* ptrs, const float* noise_uniforms, uint32_t step, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* LT_B200_H */
