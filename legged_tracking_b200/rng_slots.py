"""Per-env uniform-random slot layout of one policy step.

The reference draws `torch.rand(len(env_ids), ...)` on *compacted* env lists (reference
legged_robot_trajectory_tracking.py:744-764, 1006, 1026-1047; trajectory_function.py:28-93), so its
random stream depends on how many envs reset.  The fused step instead consumes, for env i, slot s of a
`[num_envs, n_slots]` array of uniforms in [0,1): either injected by the caller (parity tests: the
oracle harness records the reference's draws and scatters them into this layout) or produced in-kernel
by a counter-based Philox keyed (seed, step, env, slot).
"""

INTERVAL_STRENGTH = 0      # 1   _randomize_dof_props called from _post_physics_step_callback (LR:822-824)
INTERVAL_OFFSET = 1        # 12
INTERVAL_KP = 13           # 1
INTERVAL_KD = 14           # 1
RESET_STRENGTH = 15        # 1   _randomize_dof_props called from reset_idx (LR:232)
RESET_OFFSET = 16          # 12
RESET_KP = 28              # 1
RESET_KD = 29              # 1
RESET_DOF = 30             # 12  _reset_dofs (LR:1006)
RESET_X = 42               # 1   _reset_root_states (LR:1026-1031)
RESET_Y = 43               # 1
RESET_YAW = 44             # 1   (LR:1040-1042)
RESET_VEL = 45             # 6   (LR:1047)
TRAJ = 51                  # trajectory function draws (trajectory_function.py)
NOISE_GRAVITY = None       # observation noise uses a separate [num_envs, num_obs] array


def traj_slots(flat):
    fn = flat["traj_function"]
    if fn == "fixed_target":
        return 0
    if fn == "random_goal":
        return 3               # x, y, yaw  (trajectory_function.py:32-38)
    if fn == "random_target":
        nt = flat["traj_length"] // flat["num_interpolation"] + 1
        return 6 * nt          # x, y, z, yaw, pitch, roll blocks of nt (trajectory_function.py:77-82)
    raise NotImplementedError(fn)


def n_slots(flat):
    n = TRAJ + traj_slots(flat)
    return (n + 3) // 4 * 4    # keep rows 16-byte aligned
