// lt_cmd.cuh -- what LeggedRobot._resample_commands (go1_gym/envs/base/legged_robot_velocity_tracking.py:780-836, LRV) does to ONE
// env's freshly written command row: the gait-category rewrite of columns 5..7, binary-phase rounding, small-velocity zeroing.
// Shared by lt_resample_commands (csrc/lt_gait.cu) and the device-side indexed paths of the cfg 4 step (csrc/lt_vel.cu).
#pragma once
#include "lt_core.cuh"

// torch's `x % 1` for float tensors: fmod, then one addition when the signs differ (python modulo)
__device__ __forceinline__ float lt_mod1(float x) {
    float r = fmodf(x, 1.0f);
    if (r != 0.0f && r < 0.0f) r += 1.0f;
    return r;
}

__device__ __forceinline__ void lt_cmd_rewrite(const lt_cmd_cfg& c, float* cmd, int cat, float u) {
    const int nc = c.num_commands;
    if (nc > 5) {
        float c5 = cmd[5], c6 = cmd[6], c7 = cmd[7];
        int gait = LT_GAIT_NONE;
        bool rewrite = false;
        if (c.gaitwise_curricula) {                                                  // LRV:780-797: by curriculum category
            gait = (cat >= 0 && cat < c.n_categories) ? c.category_gait[cat] : LT_GAIT_NONE;
            rewrite = true;
        } else if (c.exclusive_phase_offset) {                                       // LRV:799-809: one non-zero column, untouched
            if (u < 0.34f) { c6 = 0.0f; c7 = 0.0f; }                                 // trotting
            else if (u < 0.67f) { c5 = 0.0f; c7 = 0.0f; }                            // pacing
            else { c5 = 0.0f; c6 = 0.0f; }                                           // bounding
        } else if (c.balance_gait_distribution) {                                    // LRV:811-828
            // pronking: u <= 0.25; trotting: 0.25 <= u < 0.5 (u == 0.25 belongs to both index sets: the pronk rewrite first, then
            // the trot rewrite on top of it, in the reference's statement order)
            if (u <= 0.25f) { c5 = lt_mod1(c5 / 2.0f - 0.25f); c6 = lt_mod1(c6 / 2.0f - 0.25f); c7 = lt_mod1(c7 / 2.0f - 0.25f); }
            const bool trot = 0.25f <= u && u < 0.5f, pace = 0.5f <= u && u < 0.75f, bound = 0.75f <= u;
            if (trot) { c6 = 0.0f; c7 = 0.0f; }
            if (pace) { c5 = 0.0f; c7 = 0.0f; }
            if (bound) { c5 = 0.0f; c6 = 0.0f; }
            if (trot) c5 = c5 / 2.0f + 0.25f;
            if (pace) c6 = c6 / 2.0f + 0.25f;
            if (bound) c7 = c7 / 2.0f + 0.25f;
        }
        if (rewrite) {
            if (gait == LT_GAIT_PRONK) { c5 = lt_mod1(c5 / 2.0f - 0.25f); c6 = lt_mod1(c6 / 2.0f - 0.25f); c7 = lt_mod1(c7 / 2.0f - 0.25f); }
            else if (gait == LT_GAIT_TROT) { c5 = c5 / 2.0f + 0.25f; c6 = 0.0f; c7 = 0.0f; }
            else if (gait == LT_GAIT_PACE) { c5 = 0.0f; c6 = c6 / 2.0f + 0.25f; c7 = 0.0f; }
            else if (gait == LT_GAIT_BOUND) { c5 = 0.0f; c6 = 0.0f; c7 = c7 / 2.0f + 0.25f; }
        }
        if (c.binary_phases) {                                                       // LRV:830-833: torch.round = half to even
            c5 = lt_mod1(rintf(2.0f * c5) / 2.0f);
            c6 = lt_mod1(rintf(2.0f * c6) / 2.0f);
            c7 = lt_mod1(rintf(2.0f * c7) / 2.0f);
        }
        cmd[5] = c5; cmd[6] = c6; cmd[7] = c7;
    }
    // LRV:836: commands[env_ids, :2] *= (norm(commands[env_ids, :2]) > 0.2)
    const float keep = sqrtf(__fadd_rn(__fmul_rn(cmd[0], cmd[0]), __fmul_rn(cmd[1], cmd[1]))) > 0.2f ? 1.0f : 0.0f;
    cmd[0] *= keep;
    cmd[1] *= keep;
}
