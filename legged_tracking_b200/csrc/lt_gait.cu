// lt_gait.cu -- velocity-tracking specifics of SURVEY.md section 8: row (f) 3, LeggedRobot._step_contact_targets
// (go1_gym/envs/base/legged_robot_velocity_tracking.py:844-923, LRV below) and row a9.3, the CoRLRewards container
// (go1_gym/envs/rewards/corl_rewards.py:15-201, CR below).  Thread per env; the reward kernel writes one UNSCALED row per
// requested term ([n_terms, N]), the layout lt_step_args.extra_reward takes, so the terms can enter the fused reward sum with
// the same scale / bucket / episode-sum handling as any other term.
#include <cuda_runtime.h>
#include <stdint.h>

#include "lt_core.cuh"

// ------------------------------------------------------------------------------------------------ _step_contact_targets
struct lt_gait_params {
    lt_gait_ptrs p;
    int32_t num_envs, num_commands, pacing_offset;
    float dt, kappa;
};

// torch.distributions.Normal(0, kappa).cdf(x) = 0.5 * (1 + erf(x * (1 / kappa) / sqrt(2)))
__device__ __forceinline__ float lt_normal_cdf(float x, float inv_kappa) {
    return LT_FMUL(0.5f, LT_FADD(1.0f, erff(LT_FDIV(LT_FMUL(x, inv_kappa), 1.41421356237309504880f))));
}

__global__ void __launch_bounds__(256) lt_gait_kernel(const __grid_constant__ lt_gait_params a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.num_envs) return;
    const float* c = a.p.commands + (size_t)i * a.num_commands;
    const float freq = c[4], phase = c[5], offset = c[6], bound = c[7], dur = c[8];
    const float gait = lt_remainder(LT_FADD(a.p.gait_indices[i], LT_FMUL(a.dt, freq)), 1.0f);          // LRV:851
    a.p.gait_indices[i] = gait;
    float idx[4];
    idx[0] = LT_FADD(LT_FADD(LT_FADD(gait, phase), offset), bound);                                       // LRV:853-862
    idx[1] = LT_FADD(gait, a.pacing_offset ? bound : offset);
    idx[2] = LT_FADD(gait, a.pacing_offset ? offset : bound);
    idx[3] = LT_FADD(gait, phase);
    const float inv_kappa = LT_FDIV(1.0f, a.kappa);
#pragma unroll
    for (int f = 0; f < 4; ++f) {
        const float rem = lt_remainder(idx[f], 1.0f);
        a.p.foot_indices[(size_t)i * 4 + f] = rem;                                                        // LRV:864
        float w = idx[f];                                                                                 // LRV:866-872
        if (rem < dur) w = LT_FMUL(rem, LT_FDIV(0.5f, dur));
        else if (rem > dur) w = LT_FADD(0.5f, LT_FMUL(LT_FSUB(rem, dur), LT_FDIV(0.5f, LT_FSUB(1.0f, dur))));
        a.p.clock_inputs[(size_t)i * 4 + f] = sinf(LT_FMUL(LT_TWO_PI_F, w));                              // LRV:876-889
        a.p.doubletime_clock_inputs[(size_t)i * 4 + f] = sinf(LT_FMUL(2.0f * LT_TWO_PI_F, w));
        a.p.halftime_clock_inputs[(size_t)i * 4 + f] = sinf(LT_FMUL(LT_PI_F, w));
        const float r = lt_remainder(w, 1.0f);                                                            // LRV:897-916
        const float s = LT_FADD(LT_FMUL(lt_normal_cdf(r, inv_kappa), LT_FSUB(1.0f, lt_normal_cdf(LT_FSUB(r, 0.5f), inv_kappa))),
                                LT_FMUL(lt_normal_cdf(LT_FSUB(r, 1.0f), inv_kappa),
                                        LT_FSUB(1.0f, lt_normal_cdf(LT_FSUB(LT_FSUB(r, 0.5f), 1.0f), inv_kappa))));
        a.p.desired_contact_states[(size_t)i * 4 + f] = s;
    }
}

// ------------------------------------------------------------------------------------------------ CoRLRewards
struct lt_corl_params {
    lt_corl_cfg c;
    lt_corl_ptrs p;
};

__device__ __forceinline__ float lt_sq(float x) { return LT_FMUL(x, x); }

// isaacgym.torch_utils.quat_mul
__device__ __forceinline__ void lt_quat_mul(const float a[4], const float b[4], float o[4]) {
    const float x1 = a[0], y1 = a[1], z1 = a[2], w1 = a[3], x2 = b[0], y2 = b[1], z2 = b[2], w2 = b[3];
    const float ww = LT_FMUL(LT_FADD(z1, x1), LT_FADD(x2, y2));
    const float yy = LT_FMUL(LT_FSUB(w1, y1), LT_FADD(w2, z2));
    const float zz = LT_FMUL(LT_FADD(w1, y1), LT_FSUB(w2, z2));
    const float xx = LT_FADD(LT_FADD(ww, yy), zz);
    const float qq = LT_FMUL(0.5f, LT_FADD(xx, LT_FMUL(LT_FSUB(z1, x1), LT_FSUB(x2, y2))));
    o[3] = LT_FADD(LT_FSUB(qq, ww), LT_FMUL(LT_FSUB(z1, y1), LT_FSUB(y2, z2)));
    o[0] = LT_FADD(LT_FSUB(qq, xx), LT_FMUL(LT_FADD(x1, w1), LT_FADD(x2, w2)));
    o[1] = LT_FADD(LT_FSUB(qq, yy), LT_FMUL(LT_FSUB(w1, x1), LT_FADD(y2, z2)));
    o[2] = LT_FADD(LT_FSUB(qq, zz), LT_FMUL(LT_FADD(z1, y1), LT_FSUB(w2, x2)));
}
// quat_from_angle_axis(angle, unit axis e_k) followed by quat_unit
__device__ __forceinline__ void lt_quat_axis(float angle, int k, float q[4]) {
    float s, c;
    sincosf(LT_FMUL(angle, 0.5f), &s, &c);
    q[0] = q[1] = q[2] = 0.0f;
    q[k] = s;
    q[3] = c;
    const float n = fmaxf(LT_FSQRT(LT_FADD(LT_FMUL(s, s), LT_FMUL(c, c))), 1e-9f);
    q[k] = LT_FDIV(q[k], n);
    q[3] = LT_FDIV(q[3], n);
}
// quat_apply_yaw(q, v) (math_utils.py:12-16)
__device__ __forceinline__ void lt_apply_yaw(const float q[4], const float v[3], float o[3]) {
    float n = LT_FSQRT(LT_FADD(LT_FMUL(q[2], q[2]), LT_FMUL(q[3], q[3])));
    n = fmaxf(n, 1e-9f);
    const float qy[4] = {0.0f, 0.0f, LT_FDIV(q[2], n), LT_FDIV(q[3], n)};
    float t[3], u[3];
    lt_cross(qy, v, t);
#pragma unroll
    for (int k = 0; k < 3; ++k) t[k] = LT_FMUL(t[k], 2.0f);
    lt_cross(qy, t, u);
#pragma unroll
    for (int k = 0; k < 3; ++k) o[k] = LT_FADD(LT_FADD(v[k], LT_FMUL(qy[3], t[k])), u[k]);
}

__global__ void __launch_bounds__(128) lt_corl_kernel(const __grid_constant__ lt_corl_params a) {
    const lt_corl_cfg& c = a.c;
    const size_t ds = c.dof_stride > 1 ? (size_t)c.dof_stride : 1;
    const lt_corl_ptrs& p = a.p;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int N = c.num_envs;
    if (i >= N) return;
    constexpr int D = LT_ND;
    const float* cmd = p.commands + (size_t)i * c.num_commands;
    const float* cf = p.contact_forces + (size_t)i * c.num_bodies * 3;
    // evaluated by the terms that use them: with one term per thread (blockIdx.y) most threads need neither
    auto foot_force = [&](int f) {
        const float* F = cf + 3 * c.feet_indices[f];
        return lt_norm3(F[0], F[1], F[2]);
    };
    auto foot_contact_z = [&](int f) { return cf[3 * c.feet_indices[f] + 2] > 1.0f; };
    const float* dcs = p.desired_contact_states + (size_t)i * 4;
    const float* fi = p.foot_indices + (size_t)i * 4;
    const float* fp = p.foot_positions + (size_t)i * 12;
    const float* fv = p.foot_velocities + (size_t)i * 12;
    // blockIdx.y strides the terms: launched with one row of blocks per term, every term of every env is its own thread, so the
    // loads of different terms are in flight together instead of one memory round trip per term of a per-env loop
    for (int t = blockIdx.y; t < c.n_terms; t += gridDim.y) {
        float v = 0.0f;
        switch (c.term_ids[t]) {
        case LT_CORL_TRACKING_LIN_VEL: {                                                                  // CR:15-18
            const float* b = p.base_lin_vel + (size_t)i * 3;
            const float e = LT_FADD(lt_sq(LT_FSUB(cmd[0], b[0])), lt_sq(LT_FSUB(cmd[1], b[1])));
            v = expf(LT_FDIV(-e, c.tracking_sigma));
        } break;
        case LT_CORL_TRACKING_ANG_VEL:                                                                    // CR:20-23
            v = expf(LT_FDIV(-lt_sq(LT_FSUB(cmd[2], p.base_ang_vel[(size_t)i * 3 + 2])), c.tracking_sigma_yaw));
            break;
        case LT_CORL_LIN_VEL_Z: v = lt_sq(p.base_lin_vel[(size_t)i * 3 + 2]); break;
        case LT_CORL_ANG_VEL_XY: v = LT_FADD(lt_sq(p.base_ang_vel[(size_t)i * 3]), lt_sq(p.base_ang_vel[(size_t)i * 3 + 1])); break;
        case LT_CORL_ORIENTATION: v = LT_FADD(lt_sq(p.projected_gravity[(size_t)i * 3]), lt_sq(p.projected_gravity[(size_t)i * 3 + 1])); break;
        case LT_CORL_TORQUES:
            for (int d = 0; d < D; ++d) v = LT_FADD(v, lt_sq(p.torques[(size_t)i * D + d]));
            break;
        case LT_CORL_DOF_ACC:
            for (int d = 0; d < D; ++d) v = LT_FADD(v, lt_sq(LT_FDIV(LT_FSUB(p.last_dof_vel[(size_t)i * D + d], p.dof_vel[((size_t)i * D + d) * ds]), c.dt)));
            break;
        case LT_CORL_ACTION_RATE:
            for (int d = 0; d < D; ++d) v = LT_FADD(v, lt_sq(LT_FSUB(p.last_actions[(size_t)i * D + d], p.actions[(size_t)i * D + d])));
            break;
        case LT_CORL_COLLISION:                                                                           // CR:49-52
            for (int k = 0; k < c.n_penalised; ++k) {
                const float* F = cf + 3 * c.penalised[k];
                if (lt_norm3(F[0], F[1], F[2]) > 0.1f) v = LT_FADD(v, 1.0f);
            }
            break;
        case LT_CORL_DOF_POS_LIMITS:                                                                      // CR:54-58
            for (int d = 0; d < D; ++d) {
                const float q = p.dof_pos[((size_t)i * D + d) * ds];
                const float lo = -fminf(LT_FSUB(q, p.dof_pos_limits[2 * d]), 0.0f);
                v = LT_FADD(v, LT_FADD(lo, fmaxf(LT_FSUB(q, p.dof_pos_limits[2 * d + 1]), 0.0f)));
            }
            break;
        case LT_CORL_JUMP:                                                                                // CR:60-65
            v = -lt_sq(LT_FSUB(p.base_pos[(size_t)i * 3 + 2], LT_FADD(cmd[3], c.base_height_target)));
            break;
        case LT_CORL_TRACKING_CONTACTS_SHAPED_FORCE:                                                      // CR:67-75
            for (int f = 0; f < 4; ++f)
                v = LT_FADD(v, LT_FMUL(-LT_FSUB(1.0f, dcs[f]), LT_FSUB(1.0f, expf(LT_FDIV(LT_FMUL(-1.0f, lt_sq(foot_force(f))), c.gait_force_sigma)))));
            v = LT_FDIV(v, 4.0f);
            break;
        case LT_CORL_TRACKING_CONTACTS_SHAPED_VEL:                                                        // CR:77-84
            for (int f = 0; f < 4; ++f) {
                const float s = lt_norm3(fv[3 * f], fv[3 * f + 1], fv[3 * f + 2]);
                v = LT_FADD(v, -LT_FMUL(dcs[f], LT_FSUB(1.0f, expf(LT_FDIV(LT_FMUL(-1.0f, lt_sq(s)), c.gait_vel_sigma)))));
            }
            v = LT_FDIV(v, 4.0f);
            break;
        case LT_CORL_DOF_POS:
            for (int d = 0; d < D; ++d) v = LT_FADD(v, lt_sq(LT_FSUB(p.dof_pos[((size_t)i * D + d) * ds], p.default_dof_pos[d])));
            break;
        case LT_CORL_DOF_VEL:
            for (int d = 0; d < D; ++d) v = LT_FADD(v, lt_sq(p.dof_vel[((size_t)i * D + d) * ds]));
            break;
        case LT_CORL_ACTION_SMOOTHNESS_1:                                                                 // CR:94-98
            for (int d = 0; d < D; ++d) {
                const float e = lt_sq(LT_FSUB(p.joint_pos_target[(size_t)i * D + d], p.last_joint_pos_target[(size_t)i * D + d]));
                v = LT_FADD(v, p.last_actions[(size_t)i * D + d] != 0.0f ? e : 0.0f);
            }
            break;
        case LT_CORL_ACTION_SMOOTHNESS_2:                                                                 // CR:100-105
            for (int d = 0; d < D; ++d) {
                const size_t k = (size_t)i * D + d;
                const float e = lt_sq(LT_FADD(LT_FSUB(p.joint_pos_target[k], LT_FMUL(2.0f, p.last_joint_pos_target[k])), p.last_last_joint_pos_target[k]));
                v = LT_FADD(v, (p.last_actions[k] != 0.0f && p.last_last_actions[k] != 0.0f) ? e : 0.0f);
            }
            break;
        case LT_CORL_FEET_SLIP:                                                                           // CR:107-113 (stateful)
            for (int f = 0; f < 4; ++f) {
                const bool cz = foot_contact_z(f);
                const bool filt = cz || p.last_contacts[(size_t)i * 4 + f] != 0;
                p.last_contacts[(size_t)i * 4 + f] = cz ? 1 : 0;
                const float s = lt_sq(lt_norm2(fv[3 * f], fv[3 * f + 1]));
                v = LT_FADD(v, filt ? s : 0.0f);
            }
            break;
        case LT_CORL_FEET_CONTACT_VEL:                                                                    // CR:115-120
            for (int f = 0; f < 4; ++f) {
                const float s = lt_sq(lt_norm3(fv[3 * f], fv[3 * f + 1], fv[3 * f + 2]));
                v = LT_FADD(v, fp[3 * f + 2] < 0.03f ? s : 0.0f);
            }
            break;
        case LT_CORL_FEET_CONTACT_FORCES:                                                                 // CR:122-125
            for (int f = 0; f < 4; ++f) v = LT_FADD(v, fmaxf(LT_FSUB(foot_force(f), c.max_contact_force), 0.0f));
            break;
        case LT_CORL_FEET_CLEARANCE_CMD_LINEAR:                                                           // CR:127-132
            for (int f = 0; f < 4; ++f) {
                const float cl = fminf(fmaxf(LT_FSUB(LT_FMUL(fi[f], 2.0f), 1.0f), 0.0f), 1.0f);
                const float ph = LT_FSUB(1.0f, fabsf(LT_FSUB(1.0f, LT_FMUL(cl, 2.0f))));
                const float target = LT_FADD(LT_FMUL(cmd[9], ph), 0.02f);
                v = LT_FADD(v, LT_FMUL(lt_sq(LT_FSUB(target, fp[3 * f + 2])), LT_FSUB(1.0f, dcs[f])));
            }
            break;
        case LT_CORL_FEET_IMPACT_VEL:                                                                     // CR:134-141
            for (int f = 0; f < 4; ++f) {
                const float pv = fminf(fmaxf(p.prev_foot_velocities[(size_t)i * 12 + 3 * f + 2], -100.0f), 0.0f);
                v = LT_FADD(v, foot_force(f) > 1.0f ? lt_sq(pv) : 0.0f);
            }
            break;
        case LT_CORL_ORIENTATION_CONTROL: {                                                               // CR:149-161
            float qr[4], qp[4], q[4], want[3];
            lt_quat_axis(-cmd[11], 0, qr);
            lt_quat_axis(-cmd[10], 1, qp);
            lt_quat_mul(qr, qp, q);
            lt_quat_rotate_inverse(q, p.gravity_vec + (size_t)i * 3, want);
            const float* pg = p.projected_gravity + (size_t)i * 3;
            v = LT_FADD(lt_sq(LT_FSUB(pg[0], want[0])), lt_sq(LT_FSUB(pg[1], want[1])));
        } break;
        case LT_CORL_RAIBERT_HEURISTIC: {                                                                 // CR:163-201
            const float* bq = p.base_quat + (size_t)i * 4;
            const float conj[4] = {-bq[0], -bq[1], -bq[2], bq[3]};
            const float width = cmd[12], length = cmd[13], freq = cmd[4];
            const float half_over_f = LT_FDIV(0.5f, freq);
            const float y_vel = LT_FDIV(LT_FMUL(cmd[2], length), 2.0f);
            for (int f = 0; f < 4; ++f) {
                const float rel[3] = {LT_FSUB(fp[3 * f], p.base_pos[(size_t)i * 3]), LT_FSUB(fp[3 * f + 1], p.base_pos[(size_t)i * 3 + 1]),
                                      LT_FSUB(fp[3 * f + 2], p.base_pos[(size_t)i * 3 + 2])};
                float body[3];
                lt_apply_yaw(conj, rel, body);
                const float ph = LT_FSUB(LT_FMUL(fabsf(LT_FSUB(1.0f, LT_FMUL(fi[f], 2.0f))), 1.0f), 0.5f);
                float yo = LT_FMUL(LT_FMUL(ph, y_vel), half_over_f);
                if (f >= 2) yo = LT_FMUL(yo, -1.0f);
                const float xo = LT_FMUL(LT_FMUL(ph, cmd[0]), half_over_f);
                const float ys = LT_FADD((f & 1) ? -LT_FDIV(width, 2.0f) : LT_FDIV(width, 2.0f), yo);
                const float xs = LT_FADD((f >= 2) ? -LT_FDIV(length, 2.0f) : LT_FDIV(length, 2.0f), xo);
                v = LT_FADD(v, LT_FADD(lt_sq(fabsf(LT_FSUB(xs, body[0]))), lt_sq(fabsf(LT_FSUB(ys, body[1])))));
            }
        } break;
        default: break;
        }
        p.out[(size_t)t * N + i] = v;
    }
}

// ------------------------------------------------------------------------------------------------ command resampling (LRV:776-840)
struct lt_cmd_params {
    lt_cmd_cfg c;
    lt_cmd_ptrs p;
    int count;
};

#include "lt_cmd.cuh"

__global__ void __launch_bounds__(128) lt_cmd_kernel(const lt_cmd_params a) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= a.count) return;
    const lt_cmd_cfg& c = a.c;
    const int64_t env = a.p.env_ids[k];
    const int nc = c.num_commands;
    float* cmd = a.p.commands + (size_t)env * nc;
    const float* src = a.p.new_commands + (size_t)k * nc;
    for (int j = 0; j < nc; ++j) cmd[j] = src[j];                                   // LRV:776-777
    lt_cmd_rewrite(c, cmd, a.p.category ? a.p.category[k] : 0, a.p.gait_uniform ? a.p.gait_uniform[k] : 0.0f);
    if (a.p.command_sums)
        for (int s = 0; s < c.n_sums; ++s) a.p.command_sums[(size_t)s * c.num_envs + env] = 0.0f;     // LRV:839-840
}

extern "C" {

int lt_resample_commands(const lt_cmd_cfg* cfg, const lt_cmd_ptrs* ptrs, int32_t count, void* stream) {
    if (!cfg || !ptrs || count < 0 || cfg->num_envs <= 0 || cfg->num_commands < 3 || cfg->n_categories < 0 ||
        cfg->n_categories > LT_CMD_MAX_CATEGORIES || cfg->n_sums < 0)
        return LT_ERR_INVALID;
    if (count == 0) return LT_OK;                                                   // LRV:730
    if (!ptrs->env_ids || !ptrs->new_commands || !ptrs->commands) return LT_ERR_INVALID;
    if (cfg->gaitwise_curricula && !ptrs->category) return LT_ERR_INVALID;
    if (!cfg->gaitwise_curricula && (cfg->exclusive_phase_offset || cfg->balance_gait_distribution) && cfg->num_commands > 5 && !ptrs->gait_uniform)
        return LT_ERR_INVALID;
    lt_cmd_params a;
    a.c = *cfg;
    a.p = *ptrs;
    a.count = count;
    lt_cmd_kernel<<<(count + 127) / 128, 128, 0, (cudaStream_t)stream>>>(a);
    return cudaGetLastError() == cudaSuccess ? LT_OK : LT_ERR_CUDA;
}

int lt_step_contact_targets(const lt_gait_ptrs* ptrs, int32_t num_envs, int32_t num_commands, float dt, float kappa, int32_t pacing_offset,
                            void* stream) {
    if (!ptrs || num_envs <= 0 || num_commands < 9 || !(kappa > 0.0f)) return LT_ERR_INVALID;
    if (!ptrs->commands || !ptrs->gait_indices || !ptrs->foot_indices || !ptrs->clock_inputs || !ptrs->doubletime_clock_inputs ||
        !ptrs->halftime_clock_inputs || !ptrs->desired_contact_states)
        return LT_ERR_INVALID;
    lt_gait_params a;
    a.p = *ptrs;
    a.num_envs = num_envs;
    a.num_commands = num_commands;
    a.pacing_offset = pacing_offset;
    a.dt = dt;
    a.kappa = kappa;
    lt_gait_kernel<<<(num_envs + 255) / 256, 256, 0, (cudaStream_t)stream>>>(a);
    return cudaGetLastError() == cudaSuccess ? LT_OK : LT_ERR_CUDA;
}

int lt_corl_rewards(const lt_corl_cfg* cfg, const lt_corl_ptrs* ptrs, void* stream) {
    if (!cfg || !ptrs || cfg->num_envs <= 0 || cfg->n_terms <= 0 || cfg->n_terms > LT_CORL_MAX_TERMS || cfg->num_commands < 14 ||
        cfg->n_penalised < 0 || cfg->n_penalised > LT_CORL_MAX_TERMS || !ptrs->out)
        return LT_ERR_INVALID;
    const void* const* pp = reinterpret_cast<const void* const*>(ptrs);
    for (size_t k = 0; k < sizeof(lt_corl_ptrs) / sizeof(void*); ++k)
        if (!pp[k]) return LT_ERR_INVALID;
    lt_corl_params a;
    a.c = *cfg;
    a.p = *ptrs;
    // a row of blocks per term while that is what hides the per-term memory round trips (BASELINE config 4, 4,096 envs: 25 -> 10 us);
    // from 16,384 envs the kernel is bound by its 48-byte-stride row reads (42 us at 65,536 envs for 1..12 rows, 46 for 24)
    int rows = cfg->n_terms > 0 ? cfg->n_terms : 1;
    if (cfg->num_envs > 16384 && rows > 8) rows = 8;
    lt_corl_kernel<<<dim3((cfg->num_envs + 127) / 128, rows), 128, 0, (cudaStream_t)stream>>>(a);
    return cudaGetLastError() == cudaSuccess ? LT_OK : LT_ERR_CUDA;
}

}   // extern "C"
