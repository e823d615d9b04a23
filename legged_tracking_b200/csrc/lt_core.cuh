// lt_core.cuh -- the per-environment arithmetic of the step path, shared by the CUDA kernels (lt_kernels.cu)
// and by the host-compiled test mirror (tests/hostsim).  Everything here is one env's worth of scalar work; the
// kernels add the warp choreography (tile staging, coalesced row copies, heightfield gathers).
//
// LR = reference go1_gym/envs/base/legged_robot_trajectory_tracking.py.  Operation ORDER follows the reference's
// torch ops one by one (each torch elementwise op rounds separately), because termination masks, waypoint
// switches and heightfield cell indices must be bit-exact: fmul/fadd/fsub below never contract into FMAs, and
// ffma is used only where torch's CPU kernel itself fuses (cross products and 2/3-element norms, verified
// against torch 2.11 CPU).
#pragma once
#include <math.h>
#include <stdint.h>

#include "../../include/lt_b200.h"

#if defined(__CUDACC__)
#define LT_HD __host__ __device__ __forceinline__
#define LT_HD_COLD inline __host__ __device__ __noinline__      /* rare paths (reset, domain randomisation): keep them out of the hot code */
#else
#define LT_HD inline
#define LT_HD_COLD inline
#endif

#if defined(__CUDA_ARCH__)
#define LT_FMUL(a, b) __fmul_rn((a), (b))
#define LT_FADD(a, b) __fadd_rn((a), (b))
#define LT_FSUB(a, b) __fsub_rn((a), (b))
#define LT_FFMA(a, b, c) __fmaf_rn((a), (b), (c))
#define LT_FDIV(a, b) __fdiv_rn((a), (b))
#define LT_FSQRT(a) __fsqrt_rn((a))
#else
// host mirror: compiled with -ffp-contract=off, so these are individually rounded as well
#define LT_FMUL(a, b) ((a) * (b))
#define LT_FADD(a, b) ((a) + (b))
#define LT_FSUB(a, b) ((a) - (b))
#define LT_FFMA(a, b, c) fmaf((a), (b), (c))
#define LT_FDIV(a, b) ((a) / (b))
#define LT_FSQRT(a) sqrtf((a))
#endif

// tuning switches (profiles/variants.py builds and times the alternatives)
#ifndef LT_OPT_UNROLL1
#define LT_OPT_UNROLL1 1        /* keep the small runtime-bounded loops rolled (instruction-cache footprint) */
#endif
#if LT_OPT_UNROLL1
#define LT_ROLLED _Pragma("unroll 1")
#else
#define LT_ROLLED
#endif

#define LT_ND 12   /* Go1: 12 actuated dofs = 12 actions; lt_create validates num_dof == num_actions == LT_ND */

#define LT_TWO_PI_F 6.28318530717958647692f   /* float(2*np.pi) */
#define LT_PI_F 3.14159265358979323846f       /* float(np.pi) */
#define LT_HALF_PI_F 1.57079632679489661923f
#define LT_EPSILON 1e-6f                      /* reward_crawling.py:7 */

// ---------------------------------------------------------------------------------------------------------------
// uniform-random slot layout of one policy step (mirrors legged_tracking_b200/rng_slots.py)
enum {
    LT_SLOT_INTERVAL_STRENGTH = 0, LT_SLOT_INTERVAL_OFFSET = 1, LT_SLOT_INTERVAL_KP = 13, LT_SLOT_INTERVAL_KD = 14,
    LT_SLOT_RESET_STRENGTH = 15, LT_SLOT_RESET_OFFSET = 16, LT_SLOT_RESET_KP = 28, LT_SLOT_RESET_KD = 29,
    LT_SLOT_RESET_DOF = 30, LT_SLOT_RESET_X = 42, LT_SLOT_RESET_Y = 43, LT_SLOT_RESET_YAW = 44, LT_SLOT_RESET_VEL = 45,
    LT_SLOT_TRAJ = 51
};
enum { LT_STREAM_SLOTS = 0x10000000u, LT_STREAM_NOISE = 0x20000000u };

LT_HD int lt_traj_slots(const lt_config& c) {
    if (c.traj_function == LT_TRAJ_RANDOM_GOAL) return 3;
    if (c.traj_function == LT_TRAJ_RANDOM_TARGET) return 6 * (c.traj_length / c.num_interpolation + 1);
    return 0;
}
LT_HD int lt_n_slots(const lt_config& c) { return (LT_SLOT_TRAJ + lt_traj_slots(c) + 3) / 4 * 4; }

// Derived layout (computed once on the host from lt_config).
struct lt_layout {
    int32_t x_start, nxf, n_heights, h0;        // first measured x row, rows used, height columns, first height column
    int32_t head_w;                             // num_obs - n_heights (staged, non-height columns)
    int32_t col_lin_only, col_ang_only, col_vel, col_gravity, col_commands, col_dof_pos, col_dof_vel, col_actions;
    int32_t col_timestep, col_last_actions, col_yaw, col_contacts;   // obs column of each block, -1 = absent
    int32_t n_sums, n_slots, lag_slots, tile_cells;                  // n_terms+3; uniform slots; lag_timesteps+1; rows*cols
    int32_t any_bucket_by_sum;
    float cam_off_norm;                         // ||camera_offset||  (LR:402)
};

LT_HD void lt_make_layout(const lt_config& c, lt_layout& L) {
    const int nd = c.num_actions;
    L.x_start = (c.observe_heights && c.measure_front_half) ? c.n_points_x / 2 + 1 : 0;
    L.nxf = c.observe_heights ? c.n_points_x - L.x_start : 0;
    L.n_heights = 2 * L.nxf * c.n_points_y;
    int col = 0;
    L.col_lin_only = L.col_ang_only = L.col_vel = L.col_timestep = L.col_last_actions = L.col_yaw = L.col_contacts = -1;
    L.col_commands = -1;
    if (c.observe_only_lin_vel) { L.col_lin_only = col; col += 3; }
    if (c.observe_only_ang_vel) { L.col_ang_only = col; col += 3; }
    if (c.observe_vel) { L.col_vel = col; col += 6; }
    L.col_gravity = col; col += 3;
    if (c.observe_command) { L.col_commands = col; col += 2; }
    L.col_dof_pos = col; col += nd;
    L.col_dof_vel = col; col += nd;
    L.col_actions = col; col += nd;
    if (c.observe_command && c.timestep_in_obs) { L.col_timestep = col; col += 1; }
    L.h0 = col; col += L.n_heights;
    if (c.observe_two_prev_actions) { L.col_last_actions = col; col += nd; }
    if (c.observe_yaw) { L.col_yaw = col; col += 1; }
    if (c.observe_contact_states) { L.col_contacts = col; col += c.n_feet; }
    L.head_w = col - L.n_heights;
    L.n_sums = c.n_terms + 3;
    L.n_slots = lt_n_slots(c);
    L.lag_slots = c.lag_timesteps + 1;
    L.tile_cells = c.tile_rows * c.tile_cols;
    L.any_bucket_by_sum = 0;
    for (int t = 0; t < c.n_terms; ++t) if (c.term_bucket[t] == LT_BUCKET_BY_SUM) L.any_bucket_by_sum = 1;
    // torch.norm over 3 elements: sqrt(fma(z,z,fma(y,y,x*x)))
    float s = c.camera_offset[0] * c.camera_offset[0];
    s = fmaf(c.camera_offset[1], c.camera_offset[1], s);
    s = fmaf(c.camera_offset[2], c.camera_offset[2], s);
    L.cam_off_norm = sqrtf(s);
}
LT_HD int lt_obs_total_width(const lt_layout& L) { return L.head_w + L.n_heights; }
// staged index of a non-height obs column
LT_HD int lt_head_index(const lt_layout& L, int col) { return col < L.h0 ? col : col - L.n_heights; }

// ---------------------------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al. 2011), counter = (env, block | stream, step, 0), key = seed
LT_HD void lt_philox(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        const uint32_t n1 = (uint32_t)p1;
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        const uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
LT_HD float lt_u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }   // [0,1), 24 bits like torch.rand

// Observation noise: ONE Philox4x32-10 call per (env, lane, call) yields ten 12-bit uniforms in (0,1) (120 of its 128
// bits).  Value v of a lane (heights: v = 2 k + layer for point lane + 32 k; other columns: v = 2 KP + m for staged
// column lane + 32 m) uses call v / 10, slot v % 10.
#define LT_NOISE_PER_CALL 10
LT_HD float lt_u12(const uint32_t rnd[4], int slot) {
    const int bit = 12 * slot, w = bit >> 5, sh = bit & 31;
    const uint64_t pair = (uint64_t)rnd[w] | ((uint64_t)rnd[w < 3 ? w + 1 : 3] << 32);
    const uint32_t x = (uint32_t)(pair >> sh) & 0xfffu;
    return LT_FMUL((float)x + 0.5f, 1.0f / 4096.0f);
}

// Per-env source of the step's uniform slots: injected row (parity) or Philox.
struct lt_uniform_src {
    const float* row;            // [n_slots] or nullptr
    uint32_t seed_lo, seed_hi, step, env;
    mutable uint32_t blk = 0xffffffffu, o[4] = {0u, 0u, 0u, 0u};      // last Philox block: consecutive slots share one call
    LT_HD_COLD float get(int slot) const {          // only reached on reset / interval randomisation
        if (row) return row[slot];
        const uint32_t b = (uint32_t)(slot >> 2);
        if (b != blk) {
            lt_philox(env, b | LT_STREAM_SLOTS, step, 0u, seed_lo, seed_hi, o);
            blk = b;
        }
        return lt_u01(o[slot & 3]);
    }
};

// ---------------------------------------------------------------------------------------------------------------
// quaternion helpers (isaacgym.torch_utils restated in oracle/shims; go1_gym/utils/math_utils.py:12-61)

// torch.cross kernel: a1*b2 - a2*b1 is compiled to fma(a1, b2, -(a2*b1))
LT_HD void lt_cross(const float a[3], const float b[3], float o[3]) {
    o[0] = LT_FFMA(a[1], b[2], -LT_FMUL(a[2], b[1]));
    o[1] = LT_FFMA(a[2], b[0], -LT_FMUL(a[0], b[2]));
    o[2] = LT_FFMA(a[0], b[1], -LT_FMUL(a[1], b[0]));
}
// quat_rotate_inverse(q, v) (call sites LR:132-134): a - b + c with
//   a = v * (2 w^2 - 1),  b = cross(q_vec, v) * w * 2,  c = q_vec * (q_vec . v) * 2
LT_HD void lt_quat_rotate_inverse(const float q[4], const float v[3], float o[3]) {
    const float w = q[3];
    const float s = LT_FSUB(LT_FMUL(2.0f, LT_FMUL(w, w)), 1.0f);
    float cr[3];
    lt_cross(q, v, cr);
    // torch.bmm of [1,3]x[3,1]: ((x0*v0 + x1*v1) + x2*v2), unfused
    const float dot = LT_FADD(LT_FADD(LT_FMUL(q[0], v[0]), LT_FMUL(q[1], v[1])), LT_FMUL(q[2], v[2]));
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float a = LT_FMUL(v[k], s);
        const float b = LT_FMUL(LT_FMUL(cr[k], w), 2.0f);
        const float c = LT_FMUL(LT_FMUL(q[k], dot), 2.0f);
        o[k] = LT_FADD(LT_FSUB(a, b), c);
    }
}
// python-style float remainder (torch.remainder): sign of the divisor
LT_HD_COLD float lt_fmod_cold(float a, float b) { return fmodf(a, b); }
LT_HD float lt_remainder(float a, float b) {
    float m = fabsf(a) < b ? a : lt_fmod_cold(a, b);      // fmod(a, b) == a exactly when |a| < b (b > 0 at every call site)
    if (m != 0.0f && ((b < 0.0f) != (m < 0.0f))) m = LT_FADD(m, b);
    return m;
}
// wrap_to_pi (math_utils.py:20-23)
LT_HD float lt_wrap_to_pi(float a) {
    a = lt_remainder(a, LT_TWO_PI_F);
    return LT_FSUB(a, (a > LT_PI_F) ? LT_TWO_PI_F : 0.0f);
}
// quaternion_to_roll_pitch_yaw (math_utils.py:40-46) = wrap_to_pi(get_euler_xyz(q) each % 2pi)
LT_HD void lt_quat_to_rpy(const float q[4], float rpy[3]) {
    const float qx = q[0], qy = q[1], qz = q[2], qw = q[3];
    const float sinr_cosp = LT_FMUL(2.0f, LT_FADD(LT_FMUL(qw, qx), LT_FMUL(qy, qz)));
    const float cosr_cosp = LT_FADD(LT_FSUB(LT_FSUB(LT_FMUL(qw, qw), LT_FMUL(qx, qx)), LT_FMUL(qy, qy)), LT_FMUL(qz, qz));
    const float roll = atan2f(sinr_cosp, cosr_cosp);
    const float sinp = LT_FMUL(2.0f, LT_FSUB(LT_FMUL(qw, qy), LT_FMUL(qz, qx)));
    float pitch;
    if (fabsf(sinp) >= 1.0f) pitch = LT_FMUL(LT_HALF_PI_F, (sinp > 0.0f) ? 1.0f : ((sinp < 0.0f) ? -1.0f : 0.0f));
    else pitch = asinf(sinp);
    const float siny_cosp = LT_FMUL(2.0f, LT_FADD(LT_FMUL(qw, qz), LT_FMUL(qx, qy)));
    const float cosy_cosp = LT_FSUB(LT_FSUB(LT_FADD(LT_FMUL(qw, qw), LT_FMUL(qx, qx)), LT_FMUL(qy, qy)), LT_FMUL(qz, qz));
    const float yaw = atan2f(siny_cosp, cosy_cosp);
    rpy[0] = lt_wrap_to_pi(lt_remainder(roll, LT_TWO_PI_F));
    rpy[1] = lt_wrap_to_pi(lt_remainder(pitch, LT_TWO_PI_F));
    rpy[2] = lt_wrap_to_pi(lt_remainder(yaw, LT_TWO_PI_F));
}
// quat_apply_yaw_inverse (math_utils.py:57-61): zero x,y, normalise (4-element norm: unfused, sequential), rotate inverse
LT_HD void lt_quat_apply_yaw_inverse(const float q[4], const float v[3], float o[3]) {
    float n = LT_FSQRT(LT_FADD(LT_FMUL(q[2], q[2]), LT_FMUL(q[3], q[3])));
    n = fmaxf(n, 1e-9f);
    const float qy4[4] = {LT_FDIV(0.0f, n), LT_FDIV(0.0f, n), LT_FDIV(q[2], n), LT_FDIV(q[3], n)};
    lt_quat_rotate_inverse(qy4, v, o);
}
// quat_from_euler_xyz (LR:942; go1_gym/utils/planner.py:25-38)
LT_HD void lt_quat_from_euler(float roll, float pitch, float yaw, float q[4]) {
    const float hy = LT_FMUL(yaw, 0.5f), hr = LT_FMUL(roll, 0.5f), hp = LT_FMUL(pitch, 0.5f);
    float cy, sy, cr, sr, cp, sp;
    sincosf(hy, &sy, &cy); sincosf(hr, &sr, &cr); sincosf(hp, &sp, &cp);
    q[3] = LT_FADD(LT_FMUL(LT_FMUL(cy, cr), cp), LT_FMUL(LT_FMUL(sy, sr), sp));
    q[0] = LT_FSUB(LT_FMUL(LT_FMUL(cy, sr), cp), LT_FMUL(LT_FMUL(sy, cr), sp));
    q[1] = LT_FADD(LT_FMUL(LT_FMUL(cy, cr), sp), LT_FMUL(LT_FMUL(sy, sr), cp));
    q[2] = LT_FSUB(LT_FMUL(LT_FMUL(sy, cr), cp), LT_FMUL(LT_FMUL(cy, sr), sp));
}
// torch.norm / torch.linalg.norm over 2 or 3 elements: fma chain
LT_HD float lt_norm2(float x, float y) { return LT_FSQRT(LT_FFMA(y, y, LT_FMUL(x, x))); }
LT_HD float lt_norm3(float x, float y, float z) { return LT_FSQRT(LT_FFMA(z, z, LT_FFMA(y, y, LT_FMUL(x, x)))); }
LT_HD float lt_square(float x) { return LT_FMUL(x, x); }

// ---------------------------------------------------------------------------------------------------------------
// actuator network: Sequential(Linear(6,32), softsign, Linear(32,32), softsign, Linear(32,1))  (LR:1307-1322)
struct alignas(16) lt_act_weights {
    float w0[LT_ACT_HIDDEN * LT_ACT_IN], b0[LT_ACT_HIDDEN];
    float w1[LT_ACT_HIDDEN * LT_ACT_HIDDEN], b1[LT_ACT_HIDDEN];
    float w2[LT_ACT_HIDDEN], b2[1], pad_[3];      // 16-byte multiple: the next kernel parameter stays aligned
};
#if defined(__CUDACC__)
// Blackwell packed fp32: one FFMA2 issues two independent IEEE fp32 FMAs (d.lo = a.lo*b.lo + c.lo, d.hi likewise), so the
// result of every output unit is bit-identical to the scalar chain while the FMA issue count halves.
__device__ __forceinline__ unsigned long long lt_pack2(float lo, float hi) {
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void lt_unpack2(unsigned long long v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ unsigned long long lt_fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
#endif
// softsign x / (1 + |x|).  Device: MUFU.RCP-based divide (<= 2 ulp; the denominator is >= 1, never denormal) -- an IEEE
// divide would be a third of the torque kernel's instructions; the parity budget of the torques is 1e-5.
LT_HD float lt_softsign(float x) {
#if defined(__CUDA_ARCH__)
    return __fdividef(x, fabsf(x) + 1.0f);
#else
    return x / (fabsf(x) + 1.0f);
#endif
}
LT_HD float lt_actuator_mlp(const lt_act_weights& W, const float x[LT_ACT_IN]) {
    float h1[LT_ACT_HIDDEN];
#pragma unroll
    for (int j = 0; j < LT_ACT_HIDDEN; ++j) {
        float a = W.b0[j];
#pragma unroll
        for (int k = 0; k < LT_ACT_IN; ++k) a = fmaf(W.w0[j * LT_ACT_IN + k], x[k], a);
        h1[j] = lt_softsign(a);
    }
    float out = W.b2[0];
#pragma unroll
    for (int j = 0; j < LT_ACT_HIDDEN; ++j) {
        float a = W.b1[j];
#pragma unroll
        for (int k = 0; k < LT_ACT_HIDDEN; ++k) a = fmaf(W.w1[j * LT_ACT_HIDDEN + k], h1[k], a);
        out = fmaf(W.w2[j], lt_softsign(a), out);
    }
    return out;
}

// One (env, dof) sample of _compute_torques (LR:957-996).  Returns the clipped torque; updates the histories.
struct lt_torque_io {
    float action_raw;                 // policy output (unclipped)
    float dof_pos, dof_vel;
    float lag_out;                    // lag_buffer[0] after the push (only if randomize_lag_timesteps)
    float motor_offset, motor_strength, kp_factor, kd_factor;
    float err_last, err_last_last, vel_last, vel_last_last;    // in: history; out: shifted history
    float action_clipped, action_scaled, joint_pos_target, torque;   // out
};
LT_HD float lt_scaled_action(const lt_config& c, float action_raw, int dof, float& clipped) {
    clipped = fminf(fmaxf(action_raw, -c.clip_actions), c.clip_actions);
    float a = LT_FMUL(clipped, c.action_scale);
    if (dof % 3 == 0) a = LT_FMUL(a, c.hip_scale_reduction);       // columns 0,3,6,9 (LR:962)
    return a;
}
// _compute_torques for one (env, dof) sample, split around the network so that the device kernel can run the network for
// several samples of one thread at once.  pre: joint target, network input x (returns false for the PD controller, whose
// torque is then already in tq); post: history shift, motor strength, clip.
LT_HD bool lt_torque_pre(const lt_config& c, int dof, lt_torque_io& s, float x[LT_ACT_IN], float& tq) {
    const float tgt_in = c.randomize_lag_timesteps ? s.lag_out : s.action_scaled;
    s.joint_pos_target = LT_FADD(tgt_in, c.default_dof_pos[dof]);
    if (c.control_type == LT_CONTROL_ACTUATOR_NET) {
        const float err = LT_FADD(LT_FSUB(s.dof_pos, s.joint_pos_target), s.motor_offset);
        x[0] = err; x[1] = s.err_last; x[2] = s.err_last_last; x[3] = s.dof_vel; x[4] = s.vel_last; x[5] = s.vel_last_last;
        return true;
    }
    const float p = LT_FMUL(LT_FMUL(c.p_gains[dof], s.kp_factor), LT_FADD(LT_FSUB(s.joint_pos_target, s.dof_pos), s.motor_offset));
    const float d = LT_FMUL(LT_FMUL(c.d_gains[dof], s.kd_factor), s.dof_vel);
    tq = LT_FSUB(p, d);
    return false;
}
LT_HD void lt_torque_post(const lt_config& c, int dof, lt_torque_io& s, bool net, const float x[LT_ACT_IN], float tq) {
    if (net) {
        s.err_last_last = s.err_last;
        s.err_last = x[0];
        s.vel_last_last = s.vel_last;
        s.vel_last = s.dof_vel;
    }
    tq = LT_FMUL(tq, s.motor_strength);
    s.torque = fminf(fmaxf(tq, -c.torque_limits[dof]), c.torque_limits[dof]);
}
LT_HD void lt_torque_sample(const lt_config& c, const lt_act_weights& W, int dof, lt_torque_io& s) {
    float x[LT_ACT_IN] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
    float tq = 0.0f;
    const bool net = lt_torque_pre(c, dof, s, x, tq);
    if (net) tq = lt_actuator_mlp(W, x);
    lt_torque_post(c, dof, s, net, x, tq);
}

#if defined(__CUDACC__)
// The network for 2*P samples of one thread: sample pairs ride in the two halves of FFMA2 operands and every weight is a
// broadcast uniform-register operand, so one 128-bit constant load feeds 8*P FMAs (the one-sample chain needs one constant
// load per 4 FMAs and is bound by the uniform datapath, not by the FMA pipe).  Per-sample op order = the scalar chain.
template <int P>
__device__ __forceinline__ void lt_actuator_mlp_pairs(const lt_act_weights& W, const unsigned long long (&x)[P][LT_ACT_IN],
                                                      unsigned long long (&out)[P]) {
    unsigned long long h1[P][LT_ACT_HIDDEN];
#pragma unroll
    for (int j = 0; j < LT_ACT_HIDDEN; ++j) {
        unsigned long long a[P];
#pragma unroll
        for (int q = 0; q < P; ++q) a[q] = lt_pack2(W.b0[j], W.b0[j]);
#pragma unroll
        for (int k = 0; k < LT_ACT_IN; ++k) {
            const unsigned long long w = lt_pack2(W.w0[j * LT_ACT_IN + k], W.w0[j * LT_ACT_IN + k]);
#pragma unroll
            for (int q = 0; q < P; ++q) a[q] = lt_fma2(w, x[q][k], a[q]);
        }
#pragma unroll
        for (int q = 0; q < P; ++q) {
            float lo, hi;
            lt_unpack2(a[q], lo, hi);
            h1[q][j] = lt_pack2(lt_softsign(lo), lt_softsign(hi));
        }
    }
#pragma unroll
    for (int q = 0; q < P; ++q) out[q] = lt_pack2(W.b2[0], W.b2[0]);
#pragma unroll
    for (int j = 0; j < LT_ACT_HIDDEN; ++j) {
        unsigned long long a[P];
#pragma unroll
        for (int q = 0; q < P; ++q) a[q] = lt_pack2(W.b1[j], W.b1[j]);
#pragma unroll
        for (int k = 0; k < LT_ACT_HIDDEN; ++k) {
            const unsigned long long w = lt_pack2(W.w1[j * LT_ACT_HIDDEN + k], W.w1[j * LT_ACT_HIDDEN + k]);
#pragma unroll
            for (int q = 0; q < P; ++q) a[q] = lt_fma2(w, h1[q][k], a[q]);
        }
        const unsigned long long w2 = lt_pack2(W.w2[j], W.w2[j]);
#pragma unroll
        for (int q = 0; q < P; ++q) {
            float lo, hi;
            lt_unpack2(a[q], lo, hi);
            out[q] = lt_fma2(w2, lt_pack2(lt_softsign(lo), lt_softsign(hi)), out[q]);
        }
    }
}
#endif

// ---------------------------------------------------------------------------------------------------------------
// One env of post_physics_step up to and including reset_idx and the non-height observation columns.

// Rows of this env in the tile-staged (kernel: shared memory; host mirror: the global arrays) sim / state tensors.
struct lt_env_rows {
    float* root;                 // [13] robot actor root state (LR:1189); overwritten on reset (LR:1015-1052)
    float* arrow;                // [13] goal-arrow actor or nullptr; pose written every step (LR:934-947)
    float* dof;                  // [num_dof*2] (pos, vel) interleaved; overwritten on reset (LR:998-1013)
    uint32_t cinfo;              // lt_contact_summary of this env's contact_forces rows
    const float* actions;        // [12] clipped actions of this step (state.actions)
    const float* last_actions;   // [12] before this step's shift
    const float* last_dof_vel;   // [12]
    const float* torques;        // [12]
    float* obs_row;              // this env's obs_buf row: the non-height columns are written here, before noise / clipping
    float* es;                   // this env's episode_sums column: es[t * sum_stride], t < n_terms + 3
    float* cs;                   // this env's command_sums column or nullptr
    size_t sum_stride;
    uint8_t* ix;                 // [nxf] out: heightfield row index of each measured x (LR:1955)
    uint8_t* iy;                 // [n_points_y] out: column index (LR:1956)
};
struct lt_env_result {
    int reset;                   // reset_buf
    float z_post;                // root z seen by compute_observations (post reset, LR:401)
    float cam_z;                 // sin(previous pitch) * ||camera_offset|| (LR:402)
};
struct lt_step_scalars {         // per-launch scalars
    uint32_t step;               // policy-step counter (Philox)
    int lag_head;                // not used by the scalar step (reset zeroes every slot)
    int set_gravity;             // new_gravity valid
    float new_gravity[3];
    const float* term_sums;      // device float[LT_MAX_TERMS] all-env sums for LT_BUCKET_BY_SUM terms (or nullptr)
};

LT_HD_COLD void lt_randomize_dof_props(const lt_config& c, const lt_state_ptrs& st, int i, const lt_uniform_src& U, int base) {
    // LR:744-764: torch.rand(K,1)*(hi-lo)+lo broadcast over the 12 dofs; offsets draw 12
    constexpr int nd = LT_ND;
    if (c.randomize_motor_strength) {
        const float v = LT_FADD(LT_FMUL(U.get(base + 0), LT_FSUB(c.motor_strength_range[1], c.motor_strength_range[0])),
                                c.motor_strength_range[0]);
        _Pragma("unroll") for (int d = 0; d < nd; ++d) st.motor_strengths[(size_t)i * nd + d] = v;
    }
    if (c.randomize_motor_offset) {
        _Pragma("unroll") for (int d = 0; d < nd; ++d)
            st.motor_offsets[(size_t)i * nd + d] =
                LT_FADD(LT_FMUL(U.get(base + 1 + d), LT_FSUB(c.motor_offset_range[1], c.motor_offset_range[0])),
                        c.motor_offset_range[0]);
    }
    if (c.randomize_Kp_factor) {
        const float v = LT_FADD(LT_FMUL(U.get(base + 13), LT_FSUB(c.Kp_factor_range[1], c.Kp_factor_range[0])), c.Kp_factor_range[0]);
        _Pragma("unroll") for (int d = 0; d < nd; ++d) st.Kp_factors[(size_t)i * nd + d] = v;
    }
    if (c.randomize_Kd_factor) {
        const float v = LT_FADD(LT_FMUL(U.get(base + 14), LT_FSUB(c.Kd_factor_range[1], c.Kd_factor_range[0])), c.Kd_factor_range[0]);
        _Pragma("unroll") for (int d = 0; d < nd; ++d) st.Kd_factors[(size_t)i * nd + d] = v;
    }
}

// target linear velocity of the exploration / reaching terms (reward_crawling.py:80-84)
LT_HD void lt_target_lin_vel(const lt_config& c, const float rel[3], float tv[2], float& mag) {
    mag = lt_norm2(rel[0], rel[1]);
    const float den = LT_FADD(mag, LT_EPSILON);
    const float gate = (mag > c.lin_reaching_criterion) ? 1.0f : 0.0f;
    tv[0] = LT_FMUL(LT_FMUL(LT_FDIV(rel[0], den), c.target_lin_vel), gate);
    tv[1] = LT_FMUL(LT_FMUL(LT_FDIV(rel[1], den), c.target_lin_vel), gate);
}
LT_HD float lt_lin_vel_form(const lt_config& c, const float rel[3], const float blv[3], bool allow_prod) {
    float tv[2], mag;
    lt_target_lin_vel(c, rel, tv, mag);
    const float dx = LT_FSUB(tv[0], blv[0]), dy = LT_FSUB(tv[1], blv[1]);
    switch (c.lin_vel_form) {
    case LT_FORM_EXP: return expf(LT_FDIV(-LT_FADD(lt_square(dx), lt_square(dy)), c.tracking_sigma_lin));
    case LT_FORM_L1: return LT_FADD(fabsf(dx), fabsf(dy));
    case LT_FORM_L2: return LT_FADD(lt_square(dx), lt_square(dy));
    default: break;
    }
    if (!allow_prod) return NAN;     // reward_crawling.py:102 / trajectory_tracking_reward.py: not implemented there either
    // "prod" (reward_crawling.py:95-101)
    const float bm = lt_norm2(blv[0], blv[1]);
    const float bden = LT_FADD(bm, LT_EPSILON);
    const float r0 = LT_FDIV(LT_FMUL(LT_FDIV(tv[0], c.target_lin_vel), blv[0]), bden);
    const float r1 = LT_FDIV(LT_FMUL(LT_FDIV(tv[1], c.target_lin_vel), blv[1]), bden);
    float rew = LT_FADD(r0, r1);
    rew = LT_FMUL(rew, (bm > c.small_vel_threshold) ? 1.0f : 0.0f);
    rew = LT_FADD(rew, LT_FMUL(expf(LT_FDIV(-LT_FMUL(bm, bm), c.tracking_sigma_lin)),
                               (mag < c.lin_reaching_criterion) ? 1.0f : 0.0f));
    return rew;
}

// Values the reward terms read (LR:320-355 evaluates them on PRE-reset state).
struct lt_reward_ctx {
    const float* dof;            // interleaved (pos, vel)
    const float* last_dof_vel;
    const float* torques;
    const float* actions;
    const float* last_actions;
    uint32_t cinfo;
    float root_z;
    float blv[3], bav[3], pg[3], rel_lin[3], rel_rot[3], loc_lin[3];
    float ep_f;                  // episode_length_buf as float
    int reached, n_collisions;
    int local_goal;              // plan_buf & replan of this step (LR:847, 861)
};

// Row of lt_step_args.extra_reward ([n_external, N]) that belongs to term slot t: the k-th LT_TERM_EXTERNAL slot reads row k.
LT_HD const float* lt_ext_row(const lt_config& c, const float* ext, int t) {
    if (!ext || c.term_ids[t] != LT_TERM_EXTERNAL) return ext;
    int k = 0;
    for (int q = 0; q < t; ++q) k += (c.term_ids[q] == LT_TERM_EXTERNAL) ? 1 : 0;
    return ext + (size_t)k * (size_t)c.num_envs;
}

// value of reward term `id` (unscaled).  `dry`: evaluate without touching state (sign pre-pass).
LT_HD float lt_reward_term(const lt_config& c, const lt_state_ptrs& st, int i, int id, const lt_reward_ctx& r,
                           const float* extra_reward, bool dry) {
    constexpr int nd = LT_ND;
    switch (id) {
    case LT_TERM_DOF_ACC: {           // reward_crawling.py:18-20
        float s = 0.0f;
        _Pragma("unroll") for (int d = 0; d < nd; ++d) s = LT_FADD(s, lt_square(LT_FDIV(LT_FSUB(r.last_dof_vel[d], r.dof[2 * d + 1]), c.dt)));
        return s;
    }
    case LT_TERM_TORQUES: {           // :22-24
        float s = 0.0f;
        _Pragma("unroll") for (int d = 0; d < nd; ++d) s = LT_FADD(s, lt_square(r.torques[d]));
        return s;
    }
    case LT_TERM_DOF_POS_LIMITS: {    // :26-30
        float s = 0.0f;
        _Pragma("unroll") for (int d = 0; d < nd; ++d) {
            const float q = r.dof[2 * d];
            float o = -fminf(LT_FSUB(q, c.dof_pos_lower[d]), 0.0f);
            o = LT_FADD(o, fmaxf(LT_FSUB(q, c.dof_pos_upper[d]), 0.0f));
            s = LT_FADD(s, o);
        }
        return s;
    }
    case LT_TERM_COLLISION: return (float)r.n_collisions;                       // :32-35
    case LT_TERM_ACTION_RATE: {       // :37-39
        float s = 0.0f;
        _Pragma("unroll") for (int d = 0; d < LT_ND; ++d) s = LT_FADD(s, lt_square(LT_FSUB(r.last_actions[d], r.actions[d])));
        return s;
    }
    case LT_TERM_BASE_HEIGHT: return lt_square(LT_FSUB(r.root_z, c.base_height_target));          // :41-42
    case LT_TERM_ANG_VEL_XY: return LT_FADD(lt_square(r.bav[0]), lt_square(r.bav[1]));             // :44-46
    case LT_TERM_ORIENTATION: return LT_FADD(lt_square(r.pg[0]), lt_square(r.pg[1]));              // :48-50
    case LT_TERM_LARGE_VEL: {         // :54-57
        const float g = (lt_norm2(r.blv[0], r.blv[1]) > 0.5f) ? 1.0f : 0.0f;
        return LT_FMUL(LT_FADD(lt_square(r.blv[0]), lt_square(r.blv[1])), g);
    }
    case LT_TERM_E2E: {               // :61-75
        const float mag = lt_norm2(r.rel_lin[0], r.rel_lin[1]);
        if (c.terminate_end_of_trajectory) return (mag < c.switch_dist) ? c.max_episode_length : 0.0f;
        const float reached = (mag < c.switch_dist) ? 1.0f : 0.0f;
        const float after = (r.ep_f > c.T_reach) ? 1.0f : 0.0f;
        const float err = LT_FADD(lt_square(r.blv[0]), lt_square(r.blv[1]));
        return LT_FMUL(LT_FMUL(expf(LT_FDIV(-err, c.tracking_sigma_lin)), reached), after);
    }
    case LT_TERM_EXPLORATION_LIN: return lt_lin_vel_form(c, r.rel_lin, r.blv, true);               // :77-103
    case LT_TERM_REACHING_LINEAR_VEL: return lt_lin_vel_form(c, r.rel_lin, r.blv, false);          // trajectory_tracking_reward.py
    case LT_TERM_EXPLORATION_YAW:
    case LT_TERM_REACHING_YAW: {      // reward_crawling.py:105-114
        const float t = r.rel_rot[2];
        const float mag = fabsf(t);
        float tv = LT_FMUL(LT_FDIV(t, LT_FADD(mag, LT_EPSILON)), c.target_ang_vel);
        tv = LT_FMUL(tv, (mag > c.ang_reaching_criterion) ? 1.0f : 0.0f);
        return expf(LT_FDIV(-lt_square(LT_FSUB(tv, r.bav[2])), c.tracking_sigma_ang));
    }
    case LT_TERM_REACHING_Z: return lt_square(r.rel_lin[2]);                                        // :116-117
    case LT_TERM_REACHING_ROLL: return lt_square(r.rel_rot[0]);
    case LT_TERM_REACHING_PITCH: return lt_square(r.rel_rot[1]);
    case LT_TERM_REACHING_YAW_ABS: return lt_square(r.rel_rot[2]);
    case LT_TERM_DOF_VEL: {
        float s = 0.0f;
        _Pragma("unroll") for (int d = 0; d < nd; ++d) s = LT_FADD(s, lt_square(r.dof[2 * d + 1]));
        return s;
    }
    case LT_TERM_DOF_POS: {
        float s = 0.0f;
        _Pragma("unroll") for (int d = 0; d < nd; ++d) s = LT_FADD(s, lt_square(LT_FSUB(r.dof[2 * d], c.default_dof_pos[d])));
        return s;
    }
    case LT_TERM_TASK_OLD: {          // trajectory_tracking_reward.py:55-62
        float v = LT_FDIV(LT_FDIV(0.5f, LT_FADD(0.5f, lt_norm2(r.rel_lin[0], r.rel_lin[1]))), c.T_reach);
        return LT_FMUL(v, (r.ep_f > c.T_reach) ? 1.0f : 0.0f);
    }
    case LT_TERM_REACH_GOAL: return r.reached ? 1.0f : 0.0f;
    case LT_TERM_REACH_GOAL_T: return r.reached ? r.ep_f : 0.0f;
    case LT_TERM_REACH_GOAL_TT: return (r.reached && r.ep_f > c.T_reach) ? 1.0f : 0.0f;
    case LT_TERM_TASK: {
        float tv[2], mag;
        lt_target_lin_vel(c, r.rel_lin, tv, mag);
        const float err = LT_FADD(lt_square(LT_FSUB(tv[0], r.blv[0])), lt_square(LT_FSUB(tv[1], r.blv[1])));
        return LT_FMUL(expf(LT_FDIV(-err, c.tracking_sigma_lin)), (mag < c.large_dist_threshold) ? 1.0f : 0.0f);
    }
    case LT_TERM_EXPLORATION: {       // trajectory_tracking_reward.py:86-94
        float v = LT_FADD(LT_FMUL(r.blv[0], r.loc_lin[0]), LT_FMUL(r.blv[1], r.loc_lin[1]));
        v = LT_FDIV(v, LT_FADD(lt_norm2(r.loc_lin[0], r.loc_lin[1]), LT_EPSILON));
        const float bm = lt_norm2(r.blv[0], r.blv[1]);
        v = LT_FDIV(v, LT_FADD(bm, LT_EPSILON));
        return LT_FMUL(v, (bm > c.small_vel_threshold) ? 1.0f : 0.0f);
    }
    case LT_TERM_STALLING: {
        const bool sv = lt_norm2(r.blv[0], r.blv[1]) < c.small_vel_threshold;
        const bool ld = lt_norm2(r.rel_lin[0], r.rel_lin[1]) > c.large_dist_threshold;
        return (sv && ld) ? -1.0f : -0.0f;
    }
    case LT_TERM_LINEAR_VEL: return (lt_norm3(r.blv[0], r.blv[1], r.blv[2]) > 0.7f) ? 1.0f : 0.0f;
    case LT_TERM_LIN_VEL_Z: return lt_square(r.blv[2]);
    case LT_TERM_FEET_AIR_TIME: {     // trajectory_tracking_reward.py:115-126 (stateful)
        float rew = 0.0f;
        for (int f = 0; f < c.n_feet; ++f) {
            const size_t k = (size_t)i * c.n_feet + f;
            const bool contact = (r.cinfo >> (16 + f)) & 1u;
            const bool filt = contact || (st.last_contacts[k] != 0);
            float air = st.feet_air_time[k];
            const bool first = (air > 0.0f) && filt;
            air = LT_FADD(air, c.dt);
            rew = LT_FADD(rew, LT_FMUL(LT_FSUB(air, 0.5f), first ? 1.0f : 0.0f));
            if (!dry) {
                st.last_contacts[k] = contact ? 1 : 0;
                st.feet_air_time[k] = LT_FMUL(air, filt ? 0.0f : 1.0f);
            }
        }
        return rew;
    }
    case LT_TERM_SURVIVE: return 1.0f;
    case LT_TERM_EXTERNAL: return extra_reward ? extra_reward[i] : 0.0f;
    case LT_TERM_REACHING_LOCAL_GOAL: return r.local_goal ? 1.0f : 0.0f;      // trajectory_tracking_reward.py:95-97
    default: return 0.0f;
    }
}

// Base-frame quantities + goal-relative pose (LR:130-134, 922-932); no side effects.
struct lt_pose {
    float blv[3], bav[3], pg[3], base_rot[3], rel_lin[3], rel_rot[3], target[6];
};
LT_HD void lt_derive_pose(const lt_state_ptrs& st, const lt_config& c, int i, const float* root, int64_t cpi, lt_pose& p) {
    const float* q = root + 3;
    lt_quat_rotate_inverse(q, root + 7, p.blv);
    lt_quat_rotate_inverse(q, root + 10, p.bav);
    const float g[3] = {st.gravity_vec[0], st.gravity_vec[1], st.gravity_vec[2]};
    lt_quat_rotate_inverse(q, g, p.pg);
    const float* tp = st.trajectories + ((size_t)i * c.traj_length + (size_t)cpi) * 6;
    for (int k = 0; k < 6; ++k) p.target[k] = tp[k];
    const float d[3] = {LT_FSUB(p.target[0], root[0]), LT_FSUB(p.target[1], root[1]), LT_FSUB(p.target[2], root[2])};
    lt_quat_apply_yaw_inverse(q, d, p.rel_lin);
    lt_quat_to_rpy(q, p.base_rot);
    for (int k = 0; k < 3; ++k) p.rel_rot[k] = lt_wrap_to_pi(LT_FSUB(p.target[3 + k], p.base_rot[k]));
}
// Everything the step needs from one env's net contact forces, packed: bits 0-7 number of penalised bodies with
// |F| > 0.1 (LR:846-848, reward_crawling.py:32-35), bit 8 any termination body with |F| > 1 (LR:201-203),
// bits 16+f foot f has F_z > 1 (trajectory_tracking_reward.py:116, LR:464-467).
LT_HD uint32_t lt_contact_summary(const lt_config& c, const float* contact) {
    uint32_t w = 0;
    const uint32_t any = c.penalised_mask | c.termination_mask;
    for (int b = 0; b < c.num_bodies; ++b) {
        if (!((any >> b) & 1u)) continue;                       // uniform across envs: only listed bodies cost a norm
        const float f = lt_norm3(contact[3 * b], contact[3 * b + 1], contact[3 * b + 2]);
        if (((c.penalised_mask >> b) & 1u) && f > 0.1f) w += 1u;
        if (((c.termination_mask >> b) & 1u) && f > 1.0f) w |= 0x100u;
    }
    for (int k = 0; k < c.n_feet; ++k)
        if (contact[3 * c.feet_indices[k] + 2] > 1.0f) w |= 0x10000u << k;
    return w;
}

// trajectory functions (go1_gym/envs/trajectories/trajectory_function.py:14-93) for one reset env
LT_HD_COLD void lt_resample_trajectory(const lt_config& c, const lt_state_ptrs& st, int i, const float* root, const lt_uniform_src& U) {
    const int L = c.traj_length;
    float* T = st.trajectories + (size_t)i * L * 6;
    if (c.traj_function == LT_TRAJ_FIXED_TARGET) {                 // :14-26
        for (int k = 0; k < L; ++k) {
            const float kk = (float)(k + 1);
            T[k * 6 + 0] = LT_FADD(LT_FMUL(kk, c.base_x), root[0]);
            T[k * 6 + 1] = LT_FADD(LT_FMUL(kk, c.base_y), root[1]);
            T[k * 6 + 2] = LT_FADD(0.0f, c.base_z);
            T[k * 6 + 3] = LT_FADD(0.0f, c.base_roll);
            T[k * 6 + 4] = LT_FADD(0.0f, c.base_pitch);
            T[k * 6 + 5] = LT_FMUL(0.0f, c.base_yaw);              // sic: zeros * base_yaw (:23)
        }
    } else if (c.traj_function == LT_TRAJ_RANDOM_GOAL) {           // :28-41 (traj_length 1)
        for (int k = 0; k < L; ++k) {
            T[k * 6 + 0] = LT_FADD(LT_FADD(LT_FMUL(LT_FSUB(U.get(LT_SLOT_TRAJ + 0), 0.5f), c.x_range), c.x_mean), root[0]);
            T[k * 6 + 1] = LT_FADD(LT_FADD(LT_FMUL(LT_FSUB(U.get(LT_SLOT_TRAJ + 1), 0.5f), c.y_range), c.y_mean), root[1]);
            T[k * 6 + 2] = LT_FADD(0.0f, c.base_z);
            T[k * 6 + 3] = 0.0f;
            T[k * 6 + 4] = 0.0f;
            T[k * 6 + 5] = LT_FSUB(LT_FMUL(LT_FMUL(U.get(LT_SLOT_TRAJ + 2), 2.0f), c.yaw_range), c.yaw_range);
        }
    } else {                                                       // random_target :70-93
        const int ni = c.num_interpolation;
        const int nt = L / ni + 1;
        const float rng[6] = {c.x_range, c.y_range, c.z_range, c.roll_range, c.pitch_range, c.yaw_range};
        const int blk[6] = {0, 1, 2, 5, 4, 3};      // column -> draw block: x, y, z, roll(5), pitch(4), yaw(3)
        for (int col = 0; col < 6; ++col) {
            float prev = 0.0f;                      // target_pose[:, 0, :] = 0
            for (int t = 1; t < nt; ++t) {
                const float u = U.get(LT_SLOT_TRAJ + blk[col] * nt + t);
                const float cur = LT_FSUB(LT_FMUL(LT_FMUL(u, 2.0f), rng[col]), rng[col]);
                const float delta = LT_FDIV(LT_FSUB(cur, prev), (float)ni);
                for (int s = 0; s < ni; ++s) {
                    const int k = (t - 1) * ni + s;
                    if (k < L) {
                        float v = LT_FADD(prev, LT_FMUL((float)(s + 1), delta));
                        if (col < 3) v = LT_FADD(v, root[col]);
                        T[k * 6 + col] = v;
                    }
                }
                prev = cur;
            }
        }
    }
}

// The simulator-state part of reset_idx for one env (rare path, kept out of line): _randomize_dof_props (LR:744-764),
// _reset_dofs (LR:998-1013), _reset_root_states (LR:1015-1052), _resample_trajectory (LR:949-953).
LT_HD_COLD void lt_reset_sim_state(const lt_config& c, const lt_state_ptrs& st, const lt_terrain_ptrs& ter,
                                   const lt_uniform_src& U, int i, float* root, float* dof) {
    constexpr int nd = LT_ND;
    lt_randomize_dof_props(c, st, i, U, LT_SLOT_RESET_STRENGTH);
    // _reset_dofs LR:998-1013
    _Pragma("unroll") for (int d = 0; d < nd; ++d) {
        dof[2 * d] = LT_FMUL(c.default_dof_pos[d], LT_FADD(LT_FMUL(1.0f, U.get(LT_SLOT_RESET_DOF + d)), 0.5f));
        dof[2 * d + 1] = 0.0f;
    }
    // _reset_root_states LR:1015-1052
    for (int k = 0; k < 13; ++k) root[k] = c.base_init_state[k];
    for (int k = 0; k < 3; ++k) root[k] = LT_FADD(root[k], ter.env_origins[(size_t)i * 3 + k]);
    if (c.custom_origins) {
        const float xr = c.x_init_range, yr = c.y_init_range;
        root[0] = LT_FADD(root[0], LT_FADD(LT_FMUL(LT_FSUB(xr, -xr), U.get(LT_SLOT_RESET_X)), -xr));
        root[1] = LT_FADD(root[1], LT_FADD(LT_FMUL(LT_FSUB(yr, -yr), U.get(LT_SLOT_RESET_Y)), -yr));
        root[0] = LT_FADD(root[0], c.x_init_offset);
        root[1] = LT_FADD(root[1], c.y_init_offset);
    }
    {
        const float yr = c.yaw_init_range;
        const float yaw = LT_FADD(LT_FMUL(LT_FSUB(yr, -yr), U.get(LT_SLOT_RESET_YAW)), -yr);
        const float th = LT_FDIV(yaw, 2.0f);
        const float s = sinf(th), w = cosf(th);
        // quat_from_angle_axis: normalise([axis*sin, cos]) with axis (0,0,1); 4-norm unfused sequential
        float n = LT_FSQRT(LT_FADD(LT_FADD(LT_FADD(LT_FMUL(0.0f, 0.0f), LT_FMUL(0.0f, 0.0f)), LT_FMUL(s, s)), LT_FMUL(w, w)));
        n = fmaxf(n, 1e-9f);
        root[3] = LT_FDIV(LT_FMUL(0.0f, s), n); root[4] = LT_FDIV(LT_FMUL(0.0f, s), n);
        root[5] = LT_FDIV(LT_FMUL(1.0f, s), n); root[6] = LT_FDIV(w, n);
    }
    for (int k = 0; k < 6; ++k) root[7 + k] = LT_FADD(LT_FMUL(1.0f, U.get(LT_SLOT_RESET_VEL + k)), -0.5f);
    // _resample_trajectory LR:949-953
    lt_resample_trajectory(c, st, i, root, U);
}

// The whole scalar part of one env's post_physics_step.  Returns what the vector phases need.
LT_HD lt_env_result lt_env_step(const lt_config& c, const lt_layout& L, const lt_state_ptrs& st,
                                const lt_terrain_ptrs& ter, const lt_step_args& args, const lt_step_scalars& sc,
                                int i, const lt_env_rows& R) {
    lt_env_result res;
    constexpr int nd = LT_ND;
    lt_uniform_src U;
    U.row = args.uniforms ? args.uniforms + (size_t)i * L.n_slots : nullptr;
    U.seed_lo = c.seed_lo; U.seed_hi = c.seed_hi; U.step = sc.step; U.env = (uint32_t)i;

    // ---- every per-env word the step reads, loaded before the first store so that the loads overlap (the compiler
    // cannot move a load above a store that may alias it)
    const size_t N = (size_t)c.num_envs;
    int64_t ep = st.episode_length_buf[i] + 1;      // LR:126
    int64_t cpi = st.curr_pose_index[i];
    const int64_t coll_count0 = st.collision_count[i];
    const bool body_h0 = st.body_height_buf[i] != 0;
    const float cam_pitch_prev = st.base_rotation[(size_t)i * 3 + 1];
    const float cam_pitch_state = st.camera_pitch_angle[i];
    const float friction0 = c.priv_observe_friction ? st.friction_coeffs[i] : 0.0f;          // read early: used last
    const float restitution0 = c.priv_observe_restitution ? st.restitutions[i] : 0.0f;
    // this env's episode_sums / command_sums rows: staged in shared memory by the kernel (stride = envs per warp), the
    // global arrays themselves on the host mirror (stride = num_envs)
#define LT_ES(t) R.es[(size_t)(t) * R.sum_stride]
#define LT_CS(t) R.cs[(size_t)(t) * R.sum_stride]
    float* root = R.root;

    // ---- _get_heights index part (LR:1918-1956): previous step's pitch (base_rotation not refreshed yet)
    const float cam_pitch = cam_pitch_prev;
    if (c.observe_heights && !c.plane) {
        st.camera_pitch_angle[i] = cam_pitch;
        float offx = 0.0f, offy = 0.0f;
        if (c.camera_zero) {
            const float cp = cosf(cam_pitch);
            offx = LT_FMUL(c.camera_offset[0], cp);
            offy = LT_FMUL(c.camera_offset[1], cp);
        }
        const float ox = ter.env_terrain_origin[(size_t)i * 3 + 0], oy = ter.env_terrain_origin[(size_t)i * 3 + 1];
        for (int a = 0; a < L.nxf; ++a) {
            float p = LT_FADD(c.points_x[L.x_start + a], root[0]);
            if (c.camera_zero) p = LT_FADD(p, offx);
            p = LT_FDIV(LT_FSUB(p, ox), c.horizontal_scale);
            int v = (int)p;                                       // .long(): truncation toward zero
            v = v < 0 ? 0 : (v > c.tile_rows - 2 ? c.tile_rows - 2 : v);
            R.ix[a] = (uint8_t)v;
            if (st.height_px) st.height_px[(size_t)i * c.n_points_x + L.x_start + a] = v;
        }
        for (int b = 0; b < c.n_points_y; ++b) {
            float p = LT_FADD(c.points_y[b], root[1]);
            if (c.camera_zero) p = LT_FADD(p, offy);
            p = LT_FDIV(LT_FSUB(p, oy), c.horizontal_scale);
            int v = (int)p;
            v = v < 0 ? 0 : (v > c.tile_cols - 2 ? c.tile_cols - 2 : v);
            R.iy[b] = (uint8_t)v;
            if (st.height_py) st.height_py[(size_t)i * c.n_points_y + b] = v;
        }
    }
    const float cam_pitch_obs = (c.observe_heights && !c.plane) ? cam_pitch : cam_pitch_state;
    res.cam_z = LT_FMUL(sinf(cam_pitch_obs), L.cam_off_norm);

    // ---- base-frame quantities and goal-relative pose (LR:130-134, 850-932)
    lt_pose P;
    lt_derive_pose(st, c, i, root, cpi, P);
    for (int k = 0; k < 3; ++k) {
        st.base_lin_vel[(size_t)i * 3 + k] = P.blv[k];
        st.base_ang_vel[(size_t)i * 3 + k] = P.bav[k];
        st.projected_gravity[(size_t)i * 3 + k] = P.pg[k];
        st.base_rotation[(size_t)i * 3 + k] = P.base_rot[k];
        st.relative_linear[(size_t)i * 3 + k] = P.rel_lin[k];
        st.relative_rotation[(size_t)i * 3 + k] = P.rel_rot[k];
    }
    for (int k = 0; k < 6; ++k) st.target_poses[(size_t)i * 6 + k] = P.target[k];
    float loc_lin[3] = {P.rel_lin[0], P.rel_lin[1], P.rel_lin[2]};     // local_* = same pose (LR:917-920)
    float loc_rot[3] = {P.rel_rot[0], P.rel_rot[1], P.rel_rot[2]};
    const float* local_target = P.target;
    if (args.local_relative_linear) {                                  // sampling-based planning: lt_plan_target_pose ran first (LR:856-916)
        for (int k = 0; k < 3; ++k) {
            loc_lin[k] = args.local_relative_linear[(size_t)i * 3 + k];
            loc_rot[k] = args.local_relative_rotation[(size_t)i * 3 + k];
        }
        local_target = args.local_target_poses + (size_t)i * 6;
    }
    // goal arrow actor (LR:800, 934-947): shows the LOCAL target
    if (R.arrow) {
        R.arrow[0] = local_target[0]; R.arrow[1] = local_target[1]; R.arrow[2] = local_target[2];
        lt_quat_from_euler(local_target[3], local_target[4], local_target[5], R.arrow + 3);
    }
    // ---- interval domain randomisation (LR:822-824)
    if (ep % (int64_t)c.rand_interval == 0) lt_randomize_dof_props(c, st, i, U, LT_SLOT_INTERVAL_STRENGTH);
    // ---- gravity (LR:826-830): the host decides when; rows written here so that the privileged obs see them
    if (sc.set_gravity) for (int k = 0; k < 3; ++k) st.gravities[(size_t)i * 3 + k] = sc.new_gravity[k];
    // ---- waypoint switching (LR:836-848)
    const float ep_f = (float)ep;
    bool switched;
    if (c.switch_upon_reach) switched = lt_norm2(P.rel_lin[0], P.rel_lin[1]) < c.switch_dist;
    else switched = lt_remainder(ep_f, c.switch_interval) == 0.0f;
    if (switched) { cpi += 1; if (cpi > c.traj_length - 1) cpi = c.traj_length - 1; }
    const bool reached = switched && (cpi == c.traj_length - 1);
    bool plan = (lt_norm2(loc_lin[0], loc_lin[1]) < c.switch_dist) && (fabsf(loc_rot[2]) < c.switch_yaw);
    const int n_coll = (int)(R.cinfo & 0xffu);
    int64_t coll_count = coll_count0 + n_coll;
    // ---- check_termination (LR:198-216)
    bool reset = (R.cinfo & 0x100u) != 0;
    const bool time_out = ep_f > c.max_episode_length;
    reset = reset || time_out;
    bool body_h = body_h0;
    if (c.use_terminal_body_height) { body_h = root[2] < c.terminal_body_height; reset = reset || body_h; }
    if (c.terminate_end_of_trajectory) reset = reset || (reached && ep_f > c.T_reach);
    if (c.use_terminal_body_rotation) reset = reset || (P.pg[2] > 0.0f);
    st.time_out_buf[i] = time_out ? 1 : 0;
    st.body_height_buf[i] = body_h ? 1 : 0;
    st.switched_buf[i] = switched ? 1 : 0;
    st.reached_buf[i] = reached ? 1 : 0;

    // ---- compute_reward (LR:320-355), on pre-reset state
    lt_reward_ctx rc;
    rc.dof = R.dof; rc.last_dof_vel = R.last_dof_vel; rc.torques = R.torques; rc.actions = R.actions;
    rc.last_actions = R.last_actions; rc.cinfo = R.cinfo; rc.root_z = root[2];
    for (int k = 0; k < 3; ++k) { rc.blv[k] = P.blv[k]; rc.bav[k] = P.bav[k]; rc.pg[k] = P.pg[k];
                                  rc.rel_lin[k] = P.rel_lin[k]; rc.rel_rot[k] = P.rel_rot[k]; rc.loc_lin[k] = loc_lin[k]; }
    rc.ep_f = ep_f; rc.reached = reached; rc.n_collisions = n_coll;
    rc.local_goal = (plan && args.replan && args.replan[i]) ? 1 : 0;
    float rew = 0.0f, rew_pos = 0.0f, rew_neg = 0.0f;
    for (int t = 0; t < c.n_terms; ++t) {
        const float v = lt_reward_term(c, st, i, c.term_ids[t], rc, lt_ext_row(c, args.extra_reward, t), false);
        const float r = LT_FMUL(v, c.term_scales[t]);
        rew = LT_FADD(rew, r);
        int bucket = c.term_bucket[t];
        if (bucket == LT_BUCKET_BY_SUM) {
            const float s = sc.term_sums ? sc.term_sums[t] : 0.0f;
            bucket = (s >= 0.0f) ? LT_BUCKET_POS : ((s <= 0.0f) ? LT_BUCKET_NEG : 2);
        }
        if (bucket == LT_BUCKET_POS) rew_pos = LT_FADD(rew_pos, r);
        else if (bucket == LT_BUCKET_NEG) rew_neg = LT_FADD(rew_neg, r);
        LT_ES(t) = LT_FADD(LT_ES(t), r);
        if (st.command_sums) LT_CS(t) = LT_FADD(LT_CS(t), r);
    }
    if (c.only_positive_rewards) rew = fmaxf(rew, 0.0f);
    else if (c.only_positive_rewards_ji22_style) rew = LT_FMUL(rew_pos, expf(LT_FDIV(rew_neg, c.sigma_rew_neg)));
    st.rew_buf[i] = rew; st.rew_buf_pos[i] = rew_pos; st.rew_buf_neg[i] = rew_neg;
    LT_ES(c.n_terms) = LT_FADD(LT_ES(c.n_terms), rew);
    LT_ES(c.n_terms + 1) = LT_FADD(LT_ES(c.n_terms + 1), rew_pos);
    LT_ES(c.n_terms + 2) = LT_FADD(LT_ES(c.n_terms + 2), rew_neg);
    if (st.command_sums) {
        LT_CS(c.n_terms) = LT_FADD(LT_CS(c.n_terms), 1.0f);
    }

    // ---- reset_idx (LR:218-296) for this env
    float last_actions_post[LT_ND];
    _Pragma("unroll") for (int d = 0; d < LT_ND; ++d) last_actions_post[d] = R.last_actions[d];
    if (reset) {
        lt_reset_sim_state(c, st, ter, U, i, root, R.dof);       // DR, dofs, root pose / velocity, new trajectory
        cpi = 0;
        // buffers LR:246-254
        _Pragma("unroll") for (int d = 0; d < LT_ND; ++d) last_actions_post[d] = 0.0f;
        for (int f = 0; f < c.n_feet; ++f) st.feet_air_time[(size_t)i * c.n_feet + f] = 0.0f;
        for (int k = 0; k < 3; ++k) { loc_lin[k] = 0.0f; loc_rot[k] = 0.0f; }
        plan = true;
        // finished-episode record (LR:256-273) then zero the sums
        st.finished_episode_length[i] = ep_f;
        st.finished_goal_distance[i] = lt_norm3(P.rel_lin[0], P.rel_lin[1], P.rel_lin[2]);
        LT_ROLLED for (int s = 0; s < L.n_sums; ++s) {
            st.finished_sums[(size_t)s * N + i] = LT_ES(s);
            LT_ES(s) = 0.0f;
        }
        st.reached_env_buf[i] = reached ? 1 : 0;
        st.collision_env_buf[i] = reached ? coll_count : 0;
        coll_count = 0;
        ep = 0;
        LT_ROLLED for (int s = 0; s < L.lag_slots; ++s)
            _Pragma("unroll") for (int d = 0; d < nd; ++d) st.lag_buffer[((size_t)s * N + i) * nd + d] = 0.0f;
    }
    st.episode_length_buf[i] = ep;
    st.curr_pose_index[i] = cpi;
    st.collision_count[i] = coll_count;
    st.reset_buf[i] = reset ? 1 : 0;
    st.plan_buf[i] = plan ? 1 : 0;
    for (int k = 0; k < 3; ++k) {
        st.local_relative_linear[(size_t)i * 3 + k] = loc_lin[k];
        st.local_relative_rotation[(size_t)i * 3 + k] = loc_rot[k];
    }
    res.reset = reset ? 1 : 0;
    res.z_post = root[2];

    // ---- compute_observations, non-height columns (LR:357-475); state is post-reset except projected_gravity
    float* H = R.obs_row;
    if (L.col_lin_only >= 0) for (int k = 0; k < 3; ++k) H[L.col_lin_only + k] = LT_FMUL(P.blv[k], c.obs_scale_lin_vel);
    if (L.col_ang_only >= 0) for (int k = 0; k < 3; ++k) H[L.col_ang_only + k] = LT_FMUL(P.bav[k], c.obs_scale_ang_vel);
    if (L.col_vel >= 0) {
        for (int k = 0; k < 3; ++k) {
            const float lv = c.global_reference ? root[7 + k] : P.blv[k];
            H[L.col_vel + k] = LT_FMUL(lv, c.obs_scale_lin_vel);
            H[L.col_vel + 3 + k] = LT_FMUL(P.bav[k], c.obs_scale_ang_vel);
        }
    }
    for (int k = 0; k < 3; ++k) H[L.col_gravity + k] = P.pg[k];
    if (L.col_commands >= 0) for (int k = 0; k < 2; ++k) H[L.col_commands + k] = LT_FMUL(loc_lin[k], 1.0f);
    _Pragma("unroll") for (int d = 0; d < nd; ++d) {
        H[L.col_dof_pos + d] = LT_FMUL(LT_FSUB(R.dof[2 * d], c.default_dof_pos[d]), c.obs_scale_dof_pos);
        H[L.col_dof_vel + d] = LT_FMUL(R.dof[2 * d + 1], c.obs_scale_dof_vel);
    }
    _Pragma("unroll") for (int d = 0; d < LT_ND; ++d) H[L.col_actions + d] = R.actions[d];
    if (L.col_timestep >= 0) H[L.col_timestep] = LT_FDIV((float)ep, c.max_episode_length);
    if (L.col_last_actions >= 0) _Pragma("unroll") for (int d = 0; d < LT_ND; ++d) H[L.col_last_actions + d] = last_actions_post[d];
    if (L.col_yaw >= 0) {
        // quat_apply(base_quat, [1,0,0]) then atan2 (LR:459-462; planner.py:17-23)
        const float* q = root + 3;
        const float fwd[3] = {1.0f, 0.0f, 0.0f};
        float t[3], u[3];
        lt_cross(q, fwd, t);
        for (int k = 0; k < 3; ++k) t[k] = LT_FMUL(t[k], 2.0f);
        lt_cross(q, t, u);
        const float fx = LT_FADD(LT_FADD(fwd[0], LT_FMUL(q[3], t[0])), u[0]);
        const float fy = LT_FADD(LT_FADD(fwd[1], LT_FMUL(q[3], t[1])), u[1]);
        H[L.col_yaw] = atan2f(fy, fx);
    }
    if (L.col_contacts >= 0)
        for (int f = 0; f < c.n_feet; ++f)
            H[L.col_contacts + f] = ((R.cinfo >> (16 + f)) & 1u) ? 1.0f : 0.0f;

    // ---- privileged observations (LR:477-590), clipped like LR:110-111
    if (c.num_privileged_obs > 0) {
        float* po = st.privileged_obs_buf + (size_t)i * c.num_privileged_obs;
        int k = 0;
        const float co = c.clip_observations;
#define LT_PRIV(idx, x) po[k++] = fminf(fmaxf(LT_FMUL(LT_FSUB((x), c.priv_shift[idx]), c.priv_scale[idx]), -co), co)
        if (c.priv_observe_friction) LT_PRIV(0, friction0);
        if (c.priv_observe_restitution) LT_PRIV(1, restitution0);
        if (c.priv_observe_base_mass) LT_PRIV(2, st.payloads[i]);
        if (c.priv_observe_com_displacement) for (int a = 0; a < 3; ++a) LT_PRIV(3, st.com_displacements[(size_t)i * 3 + a]);
        if (c.priv_observe_motor_strength) _Pragma("unroll") for (int d = 0; d < nd; ++d) LT_PRIV(4, st.motor_strengths[(size_t)i * nd + d]);
        if (c.priv_observe_motor_offset) _Pragma("unroll") for (int d = 0; d < nd; ++d) LT_PRIV(5, st.motor_offsets[(size_t)i * nd + d]);
        if (c.priv_observe_body_height) LT_PRIV(6, root[2]);
        if (c.priv_observe_body_velocity) for (int a = 0; a < 3; ++a) LT_PRIV(7, P.blv[a]);
        if (c.priv_observe_gravity)       // LR:574-579 DIVIDES by the scale
            for (int a = 0; a < 3; ++a)
                po[k++] = fminf(fmaxf(LT_FDIV(LT_FSUB(st.gravities[(size_t)i * 3 + a], c.priv_shift[8]), c.priv_scale[8]), -co), co);
#undef LT_PRIV
    }
    return res;
}

// height observation of one sampled cell (LR:395-409): raw metres -> observation value (before noise / clip)
LT_HD float lt_height_obs(const lt_config& c, float h, float z_post, float cam_z) {
    if (c.camera_zero) {
        h = LT_FSUB(LT_FSUB(h, z_post), cam_z);
        h = fminf(fmaxf(h, -0.3f), 0.3f);
    } else {
        h = fminf(fmaxf(h, 0.0f), c.ceiling_height);
        h = LT_FSUB(LT_FDIV(h, c.ceiling_height), 0.5f);
    }
    return LT_FMUL(h, c.obs_scale_height);
}
// noise (LR:473-474) + clamp of LeggedRobot.step (LR:108-109)
LT_HD float lt_obs_finish(const lt_config& c, float v, float u, float noise_scale) {
    if (c.add_noise) v = LT_FADD(v, LT_FMUL(LT_FSUB(LT_FMUL(2.0f, u), 1.0f), noise_scale));
    return fminf(fmaxf(v, -c.clip_observations), c.clip_observations);
}

// Observation-noise uniform of (env, obs column) when none is injected: lt_obs_kernel's lane / call / slot assignment
// restated per column (used by the host mirror).
LT_HD float lt_noise_uniform(const lt_config& c, const lt_layout& L, uint32_t env, int col, uint32_t step) {
    const int per_layer = L.nxf * c.n_points_y;
    const int kp = per_layer <= 128 ? 4 : 8;
    const int hc = col - L.h0;
    uint32_t lane;
    int v;
    if (hc >= 0 && hc < L.n_heights) {
        const int layer = hc / per_layer, pt = hc - layer * per_layer;
        lane = (uint32_t)(pt & 31);
        v = 2 * (pt >> 5) + layer;
    } else {
        const int hi = lt_head_index(L, col);
        lane = (uint32_t)(hi & 31);
        v = 2 * kp + (hi >> 5);
    }
    uint32_t o[4];
    lt_philox(env, (lane + 32u * (uint32_t)(v / LT_NOISE_PER_CALL)) | LT_STREAM_NOISE, step, 0u, c.seed_lo, c.seed_hi, o);
    return lt_u12(o, v % LT_NOISE_PER_CALL);
}
