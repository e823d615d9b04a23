// lt_vel.cu -- cfg 4: the dense per-step parts of LeggedRobot.post_physics_step of the velocity-tracking env
// (go1_gym/envs/base/legged_robot_velocity_tracking.py:108-154, LRV below) that are not already lt_step_contact_targets /
// lt_corl_rewards / lt_resample_commands (csrc/lt_gait.cu).  All of it is HBM-bound row work: thread per env for the scalar
// logic, thread per element for the observation rows and the 30-row history shift (the one large transfer: 8.4 KB read + 8.4 KB
// written per env at cfg 4's 70 x 30 history).
#include <cuda_runtime.h>
#include <stdint.h>

#include "lt_core.cuh"

struct lt_vel_params {
    lt_vel_cfg c;
    lt_vel_ptrs p;
    const float* aux;          // term values (reward) / noise uniforms (observe)
    uint32_t step;
};

// ------------------------------------------------------------------------------------------------ LRV:117-132, 156-167, 702-716
__global__ void __launch_bounds__(128) lt_vel_pre_kernel(const __grid_constant__ lt_vel_params a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const lt_vel_cfg& c = a.c;
    if (i >= c.num_envs) return;
    const lt_vel_ptrs& p = a.p;
    if (i == 0 && p.torque_counters) {                                                // what lt_compact_kernel does in the trajectory step
        p.torque_counters[0] += 1;
        p.torque_counters[1] += c.decimation;
        p.torque_counters[2] = (p.torque_counters[2] + c.decimation) % c.lag_slots;
    }
    // every load first, then the arithmetic and the stores: the output arrays are not known to be distinct from the inputs, so a load
    // written after a store waits for it -- as first written this kernel was a chain of ~10 memory round trips per env
    const int64_t ep = p.episode_length_buf[i] + 1;                                   // LRV:117
    float root[13], g[3], fnew[4][6], fold[12];
    {
        const float* rp = p.root_states + (size_t)i * 13;
#pragma unroll
        for (int k = 0; k < 13; ++k) root[k] = rp[k];
#pragma unroll
        for (int k = 0; k < 3; ++k) g[k] = p.gravity_vec[(size_t)i * 3 + k];
#pragma unroll
        for (int f = 0; f < 4; ++f) {
            const float* rb = p.rigid_body_state + ((size_t)i * c.num_bodies + c.feet_indices[f]) * 13;
#pragma unroll
            for (int k = 0; k < 3; ++k) { fnew[f][k] = rb[k]; fnew[f][3 + k] = rb[7 + k]; }
        }
#pragma unroll
        for (int k = 0; k < 12; ++k) fold[k] = p.foot_velocities[(size_t)i * 12 + k];
    }
    bool reset = false;                                                               // check_termination (LRV:156-167)
    for (int t = 0; t < c.n_termination; ++t) {
        const float* f = p.contact_forces + ((size_t)i * c.num_bodies + c.termination_indices[t]) * 3;
        reset = reset || (lt_norm3(f[0], f[1], f[2]) > 1.0f);
    }
    // mean(root_z - measured_heights) (LRV:163-165): lt_vel_measure_heights' height_mean, or root_z itself on a plane
    const float clearance = (c.use_terminal_body_height && p.measured_height_mean) ? p.measured_height_mean[i] : root[2];

    p.episode_length_buf[i] = ep;
    float q[4], o[3];
    for (int k = 0; k < 3; ++k) p.base_pos[(size_t)i * 3 + k] = root[k];
    for (int k = 0; k < 4; ++k) { q[k] = root[3 + k]; p.base_quat[(size_t)i * 4 + k] = q[k]; }
    lt_quat_rotate_inverse(q, root + 7, o);
    for (int k = 0; k < 3; ++k) p.base_lin_vel[(size_t)i * 3 + k] = o[k];
    lt_quat_rotate_inverse(q, root + 10, o);
    for (int k = 0; k < 3; ++k) p.base_ang_vel[(size_t)i * 3 + k] = o[k];
    lt_quat_rotate_inverse(q, g, o);
    for (int k = 0; k < 3; ++k) p.projected_gravity[(size_t)i * 3 + k] = o[k];
#pragma unroll
    for (int f = 0; f < 4; ++f) {                                                     // LRV:127-130; prev_foot_velocities: LRV:74
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const size_t d = ((size_t)i * 4 + f) * 3 + k;
            p.prev_foot_velocities[d] = fold[3 * f + k];
            p.foot_positions[d] = fnew[f][k];
            p.foot_velocities[d] = fnew[f][3 + k];
        }
    }
    const bool time_out = (float)ep > c.max_episode_length;
    reset = reset || time_out;
    bool low = false;
    if (c.use_terminal_body_height) {
        low = clearance < c.terminal_body_height;
        reset = reset || low;
    }
    p.time_out_buf[i] = time_out ? 1 : 0;
    p.body_height_buf[i] = low ? 1 : 0;
    p.reset_buf[i] = reset ? 1 : 0;
    p.resample_due[i] = (ep % c.sample_interval) == 0 ? 1 : 0;                        // LRV:702-703
    p.dr_due[i] = (ep % c.rand_interval) == 0 ? 1 : 0;                                // LRV:715-716
}

// ------------------------------------------------------------------------------------------------ compute_reward (LRV:281-318)
// all-env sum of every scaled term: only its sign is used (LRV:291-294)
__global__ void __launch_bounds__(256) lt_vel_term_sums_kernel(const __grid_constant__ lt_vel_params a) {
    __shared__ double s_part[8];
    const lt_vel_cfg& c = a.c;
    const int t = blockIdx.y;
    double v = 0.0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < c.num_envs; i += gridDim.x * blockDim.x)
        v += (double)LT_FMUL(a.aux[(size_t)t * c.num_envs + i], c.term_scales[t]);
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < 8; ++w) s += s_part[w];
        atomicAdd(a.p.term_sums + t, s);
    }
}

__global__ void __launch_bounds__(128) lt_vel_reward_kernel(const __grid_constant__ lt_vel_params a) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const lt_vel_cfg& c = a.c;
    if (i >= c.num_envs) return;
    const lt_vel_ptrs& p = a.p;
    const size_t N = (size_t)c.num_envs;
    float rew = 0.0f, pos = 0.0f, neg = 0.0f;
    // terms in blocks of CH: every load of a block is issued before its first store (the three arrays are not known to be distinct, so
    // a rolled loop would wait out one memory round trip per term); the sums themselves stay in term order
    constexpr int CH = 8;
    for (int t0 = 0; t0 < c.n_terms; t0 += CH) {
        float v[CH], es[CH], cs[CH];
        double ts[CH];
#pragma unroll
        for (int k = 0; k < CH; ++k) {
            const int t = t0 + k;
            const bool on = t < c.n_terms;
            const size_t at = (size_t)(on ? t : 0) * N + i;
            v[k] = on ? a.aux[at] : 0.0f;
            es[k] = on ? p.episode_sums[at] : 0.0f;
            cs[k] = on ? p.command_sums[at] : 0.0f;
            ts[k] = on ? p.term_sums[t] : 0.0;
        }
#pragma unroll
        for (int k = 0; k < CH; ++k) {
            const int t = t0 + k;
            if (t < c.n_terms) {
                const float r = LT_FMUL(v[k], c.term_scales[t]);
                rew = LT_FADD(rew, r);
                if (ts[k] >= 0.0) pos = LT_FADD(pos, r);
                else if (ts[k] <= 0.0) neg = LT_FADD(neg, r);
                p.episode_sums[(size_t)t * N + i] = LT_FADD(es[k], r);
                const float add = c.command_sum_adds_scale[t] ? LT_FADD(c.term_scales[t], r) : r;     // LRV:296-299
                p.command_sums[(size_t)t * N + i] = LT_FADD(cs[k], add);
            }
        }
    }
    if (c.only_positive_rewards) rew = fmaxf(rew, 0.0f);                                          // LRV:300-301
    else if (c.only_positive_rewards_ji22_style) rew = LT_FMUL(pos, expf(LT_FDIV(neg, c.sigma_rew_neg)));      // LRV:302-303
    p.rew_buf[i] = rew;
    p.rew_buf_pos[i] = pos;
    p.rew_buf_neg[i] = neg;
    const size_t T = (size_t)c.n_terms;
    p.episode_sums[T * N + i] = LT_FADD(p.episode_sums[T * N + i], rew);                          // "total" (LRV:304)
    const float vx = p.base_lin_vel[(size_t)i * 3], wz = p.base_ang_vel[(size_t)i * 3 + 2];
    const float* cmd = p.commands + (size_t)i * c.num_commands;
    float* cs = p.command_sums + T * N + i;                                                       // LRV:312-316
    cs[0] = LT_FADD(cs[0], vx);
    cs[N] = LT_FADD(cs[N], wz);
    const float dx = LT_FSUB(vx, cmd[0]), dz = LT_FSUB(wz, cmd[2]);
    cs[2 * N] = LT_FADD(cs[2 * N], LT_FMUL(dx, dx));
    cs[3 * N] = LT_FADD(cs[3 * N], LT_FMUL(dz, dz));
    cs[4 * N] = LT_FADD(cs[4 * N], 1.0f);
}

// ------------------------------------------------------------------------------------------------ compute_observations (LRV:320-470)
// One observation column of env i (before noise and clip): LRV:320-370 for the switches cfg 4 has.
__device__ __forceinline__ float lt_vel_obs_value(const lt_vel_cfg& c, const lt_vel_ptrs& p, int i, int col) {
    const int nc = c.observe_command ? c.num_commands : 0;
    int k = col;
    if (k < 3) return p.projected_gravity[(size_t)i * 3 + k];
    if ((k -= 3) < nc) return LT_FMUL(p.commands[(size_t)i * c.num_commands + k], p.commands_scale[k]);
    if ((k -= nc) < 12) return LT_FMUL(LT_FSUB(p.dof_state[((size_t)i * 12 + k) * 2], p.default_dof_pos[k]), c.obs_scale_dof_pos);
    if ((k -= 12) < 12) return LT_FMUL(p.dof_state[((size_t)i * 12 + k) * 2 + 1], c.obs_scale_dof_vel);
    if ((k -= 12) < 12) return p.actions[(size_t)i * 12 + k];
    k -= 12;
    const int n_prev = c.observe_two_prev_actions ? 12 : 0, n_tim = c.observe_timing_parameter ? 1 : 0;
    if (k < n_prev) return p.last_actions[(size_t)i * 12 + k];                        // the pre-shift last_actions (LRV:347-349, 143)
    if ((k -= n_prev) < n_tim) return p.gait_indices[i];
    return p.clock_inputs[(size_t)i * 4 + (k - n_tim)];
}
__device__ __forceinline__ float lt_vel_obs_finish(const lt_vel_params& a, int i, int col, float v) {
    const lt_vel_cfg& c = a.c;
    if (c.add_noise) {                                                                // LRV:391-392
        float u;
        if (a.aux) u = a.aux[(size_t)i * c.num_obs + col];
        else {
            uint32_t r[4];
            const uint32_t step = a.p.step_counter ? *a.p.step_counter : a.step;
            lt_philox((uint32_t)i, (uint32_t)(col >> 2), step, 0x56454c4fu, c.seed_lo, c.seed_hi, r);
            u = lt_u01(r[col & 3]);
        }
        v = LT_FADD(v, LT_FMUL(LT_FSUB(LT_FMUL(2.0f, u), 1.0f), a.p.noise_scale_vec[col]));
    }
    return fminf(fmaxf(v, -c.clip_observations), c.clip_observations);                // LRV:102-103
}

// One CTA per env: its threads sweep the env's history row -- the H - 1 older observations move one slot towards the past
// (history_wrapper.py:23), the newest slot and obs_buf get this step's observation.  VEC = 2: float2 accesses (num_obs even, so the
// shift by one observation keeps 8-byte alignment); VEC = 1: any num_obs.
template <int VEC>
__global__ void __launch_bounds__(256) lt_vel_observe_kernel(const __grid_constant__ lt_vel_params a) {
    const lt_vel_cfg& c = a.c;
    const lt_vel_ptrs& p = a.p;
    const int no = c.num_obs, H = c.obs_history_length, i = blockIdx.x;
    const int row = H * no, old = (H - 1) * no;
    const float* in = p.obs_history_in + (size_t)i * row;
    float* out = p.obs_history_out + (size_t)i * row;
    if (VEC == 2) {                                                                   // four independent 8-byte loads in flight per thread
        const float2* in2 = reinterpret_cast<const float2*>(in + no);
        float2* out2 = reinterpret_cast<float2*>(out);
        const int n2 = old >> 1;
        for (int base = 0; base < n2; base += 4 * (int)blockDim.x) {
            float2 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int j = base + u * (int)blockDim.x + (int)threadIdx.x;
                if (j < n2) v[u] = __ldcs(in2 + j);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int j = base + u * (int)blockDim.x + (int)threadIdx.x;
                if (j < n2) __stcs(out2 + j, v[u]);
            }
        }
    } else {
        for (int w = threadIdx.x; w < old; w += blockDim.x) out[w] = in[w + no];
    }
    for (int col = threadIdx.x * VEC; col < no; col += blockDim.x * VEC) {
        float v[VEC];
#pragma unroll
        for (int e = 0; e < VEC; ++e) v[e] = lt_vel_obs_finish(a, i, col + e, lt_vel_obs_value(c, p, i, col + e));
        if (VEC == 2) {
            const float2 t = make_float2(v[0], v[VEC - 1]);
            *reinterpret_cast<float2*>(out + old + col) = t;
            *reinterpret_cast<float2*>(p.obs_buf + (size_t)i * no + col) = t;
        } else {
            out[old + col] = v[0];
            p.obs_buf[(size_t)i * no + col] = v[0];
        }
    }
}

// privileged observations (LRV:396-426 for the two flags cfg 4 sets) and the last_* shifts (LRV:143-148).  Thread per (env, dof)
// element: coalesced, and all of an element's loads are issued before its stores; the first 6 N threads also copy one root-velocity
// component, the first N one env's privileged row.
__global__ void __launch_bounds__(256) lt_vel_shift_kernel(const __grid_constant__ lt_vel_params a) {
    const size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const lt_vel_cfg& c = a.c;
    const size_t N = (size_t)c.num_envs;
    if (j >= N * 12) return;
    const lt_vel_ptrs& p = a.p;
    // last launch of a step: the device-side step counter moves on (its readers of this step are all behind this launch on the stream,
    // and nothing in this kernel reads it)
    if (j == 0 && p.step_counter) *p.step_counter += 1u;
    const float la = p.last_actions[j], ac = p.actions[j], lt = p.last_joint_pos_target[j], jt = p.joint_pos_target[j], dv = p.dof_state[j * 2 + 1];
    float rv = 0.0f, fr = 0.0f, re = 0.0f;
    if (j < N * 6) rv = p.root_states[(j / 6) * 13 + 7 + (j % 6)];
    if (j < N) {
        if (c.priv_observe_friction) fr = p.friction_coeffs[j];
        if (c.priv_observe_restitution) re = p.restitutions[j];
    }
    p.last_last_actions[j] = la;
    p.last_actions[j] = ac;
    p.last_last_joint_pos_target[j] = lt;
    p.last_joint_pos_target[j] = jt;
    p.last_dof_vel[j] = dv;
    if (j < N * 6) p.last_root_vel[j] = rv;
    if (j < N) {
        int k = 0;
        float* pr = p.privileged_obs_buf + j * c.num_privileged_obs;
        const float clip = c.clip_observations;
        if (c.priv_observe_friction) pr[k++] = fminf(fmaxf(LT_FMUL(LT_FSUB(fr, c.friction_shift), c.friction_scale), -clip), clip);
        if (c.priv_observe_restitution) pr[k++] = fminf(fmaxf(LT_FMUL(LT_FSUB(re, c.restitution_shift), c.restitution_scale), -clip), clip);
    }
}

// ------------------------------------------------------------------------------------------------ _get_heights (LRV:1790-1824)
struct lt_vel_height_params {
    const float *root_states, *height_points;
    const int16_t* height_samples;
    int32_t num_envs, n_points, rows, cols;
    float border_size, horizontal_scale, vertical_scale;
    float *measured_heights, *height_mean;
};

// quat_apply_yaw (math_utils.py:12-16): zero x, y; normalise; quat_apply
__device__ __forceinline__ void lt_vel_apply_yaw(const float q[4], const float v[3], float o[3]) {
    float n = LT_FSQRT(LT_FADD(LT_FMUL(q[2], q[2]), LT_FMUL(q[3], q[3])));
    n = fmaxf(n, 1e-9f);
    const float qy[4] = {0.0f, 0.0f, LT_FDIV(q[2], n), LT_FDIV(q[3], n)};
    float t[3], u[3];
    lt_cross(qy, v, t);
    for (int k = 0; k < 3; ++k) t[k] = LT_FMUL(t[k], 2.0f);
    lt_cross(qy, t, u);
    for (int k = 0; k < 3; ++k) o[k] = LT_FADD(LT_FADD(v[k], LT_FMUL(qy[3], t[k])), u[k]);
}

// warp per env: lanes sweep the grid points (gathers from the int16 field are L2 hits), then one shuffle reduction for the mean
__global__ void __launch_bounds__(256) lt_vel_heights_kernel(const __grid_constant__ lt_vel_height_params a) {
    const int lane = threadIdx.x & 31;
    const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i >= a.num_envs) return;
    const float* root = a.root_states + (size_t)i * 13;
    const float q[4] = {root[3], root[4], root[5], root[6]};
    float acc = 0.0f;
    for (int j = lane; j < a.n_points; j += 32) {
        float o[3];
        lt_vel_apply_yaw(q, a.height_points + (size_t)j * 3, o);
        const float x = LT_FDIV(LT_FADD(LT_FADD(o[0], root[0]), a.border_size), a.horizontal_scale);
        const float y = LT_FDIV(LT_FADD(LT_FADD(o[1], root[1]), a.border_size), a.horizontal_scale);
        long long px = (long long)x, py = (long long)y;                              // .long(): truncation towards zero
        px = px < 0 ? 0 : (px > a.rows - 2 ? a.rows - 2 : px);
        py = py < 0 ? 0 : (py > a.cols - 2 ? a.cols - 2 : py);
        const int16_t h1 = a.height_samples[px * a.cols + py], h2 = a.height_samples[(px + 1) * a.cols + py],
                      h3 = a.height_samples[px * a.cols + py + 1];
        const int16_t hm = min(min(h1, h2), h3);
        const float h = LT_FMUL((float)hm, a.vertical_scale);
        a.measured_heights[(size_t)i * a.n_points + j] = h;
        acc += LT_FSUB(root[2], h);
    }
    if (a.height_mean) {
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) a.height_mean[i] = acc / (float)a.n_points;
    }
}

// ------------------------------------------------------------------------------------------------ device-side indexed paths
#include "lt_cmd.cuh"

struct lt_vel_dev_params {
    lt_vel_cfg c;
    lt_vel_ptrs p;
    lt_vel_dev_cfg d;
    lt_vel_dev_ptrs q;
    const float* uniform_rows;
    uint32_t step;
    int32_t which;
};

__device__ __forceinline__ bool lt_vel_mask_resample(const lt_vel_dev_params& a, int i) {
    return (a.which == LT_VEL_PATH_RESET ? a.p.reset_buf[i] : a.p.resample_due[i]) != 0;
}

// curriculum.update (curriculum.py:133-152) for env i: a success marks its own bin and counts into every adjacent bin (phase A);
// phase B -- weights[b] = clip(weights[b] + 0.2, 0, 1), once for the bin itself and once per adjacent success -- runs in the last
// block of lt_vel_scan_kernel, when every env's marks are in.
__device__ __forceinline__ void lt_cur_hits(const lt_vel_dev_params& a, int i) {
    if (a.d.n_keys == 0) return;
    const size_t N = (size_t)a.c.num_envs;
    bool ok = true;
    for (int k = 0; k < a.d.n_keys; ++k)
        ok = ok && (LT_FDIV(a.p.command_sums[(size_t)a.d.key_rows[k] * N + i], a.d.ep_len) > a.d.key_thresholds[k]);
    if (!ok) return;
    const int cat = a.q.env_categories[i], b = a.q.env_bins[i], nb = a.d.n_bins;
    a.q.hit_self[cat * nb + b] = 1;
    for (int g = 0; g < nb; ++g) {                                                  // get_local_bins (curriculum.py:122-131)
        bool adj = true;
        for (int d = 0; d < a.d.n_dims && adj; ++d) {
            const double cg = a.q.grid[(size_t)d * nb + g], cb = a.q.grid[(size_t)d * nb + b];
            adj = (cg >= cb - a.d.local_range[d]) && (cg <= cb + a.d.local_range[d]);
        }
        if (adj) atomicAdd(a.q.hit_adj + cat * nb + g, 1);
    }
}

// Philox uniform of (env, slot) for the indexed paths, or the injected row
__device__ __forceinline__ float lt_vel_u(const lt_vel_dev_params& a, int i, int slot) {
    if (a.uniform_rows) return a.uniform_rows[(size_t)i * LT_VEL_SLOTS + slot];
    uint32_t o[4];
    const uint32_t step = ((a.p.step_counter ? *a.p.step_counter : a.step) << 1) | (uint32_t)a.which;
    lt_philox((uint32_t)i, (uint32_t)(slot >> 2), step, 0x52455345u, a.c.seed_lo, a.c.seed_hi, o);
    return lt_u01(o[slot & 3]);
}

// Pass 1, thread per env, ONE launch: the curriculum's success marks of the envs about to be resampled, and which envs have indexed
// work this call, compacted (unordered) into work_env / work_code (code bit 0 resample, bit 1 interval / reset DR, bit 2 reset).  The
// block that finishes last (ticket in work_count[2]) then applies the marks to the weights and publishes the list length:
// work_count[0] collects, work_count[1] is what lt_vel_paths_kernel reads; [0] and [2] are left zero for the next call, so the step
// needs neither a memset nor separate hit / apply / mark launches.
__global__ void __launch_bounds__(256) lt_vel_scan_kernel(const __grid_constant__ lt_vel_dev_params a) {
    __shared__ int s_last;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool in = i < a.c.num_envs;
    const bool reset = in && a.which == LT_VEL_PATH_RESET && a.p.reset_buf[i];
    const bool resample = in && lt_vel_mask_resample(a, i);
    const bool dr = in && (a.which == LT_VEL_PATH_RESET ? reset : (a.p.dr_due[i] != 0));
    if (resample) lt_cur_hits(a, i);
    const int code = (resample ? 1 : 0) | (dr ? 2 : 0) | (reset ? 4 : 0);
    const unsigned m = __ballot_sync(0xffffffffu, code != 0);
    if (m != 0u) {
        const int lane = threadIdx.x & 31;
        int base = 0;
        if (lane == 0) base = atomicAdd(a.q.work_count, __popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (code) {
            const int slot = base + __popc(m & ((1u << lane) - 1u));
            a.q.work_env[slot] = i;
            a.q.work_code[slot] = code;
        }
    }
    __threadfence();                                                                 // this thread's marks and list entries before its block's ticket
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(a.q.work_count + 2, 1) == (int)gridDim.x - 1);
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const int nw = a.d.cmd.n_categories * a.d.n_bins;
    for (int j = threadIdx.x; j < nw; j += blockDim.x) {
        const int k = __ldcg(a.q.hit_self + j) + __ldcg(a.q.hit_adj + j);
        if (k) {
            double v = a.q.weights[j];
            for (int r = 0; r < k; ++r) v = fmin(fmax(v + 0.2, 0.0), 1.0);
            a.q.weights[j] = v;
            a.q.hit_self[j] = 0;
            a.q.hit_adj[j] = 0;
        }
    }
    if (threadIdx.x == 0) {
        a.q.work_count[1] = __ldcg(a.q.work_count);
        a.q.work_count[0] = 0;
        a.q.work_count[2] = 0;
    }
}

// Pass 2, warp per listed env: the lanes share the env's work -- the weighted bin draw scans the category's weights 32 at a time,
// the command columns, motor tables, dof rows, lag ring and sum rows are written one element per lane.
__global__ void __launch_bounds__(256) lt_vel_paths_kernel(const __grid_constant__ lt_vel_dev_params a) {
    __shared__ float s_acc[LT_CORL_MAX_TERMS + 2];
    const lt_vel_cfg& c = a.c;
    const lt_vel_dev_cfg& d = a.d;
    const lt_vel_dev_ptrs& q = a.q;
    const int count = q.work_count[1];
    const int warps = blockDim.x >> 5;
    if ((int)blockIdx.x * warps >= count) return;                                    // block-uniform
    const int lane = threadIdx.x & 31, wi = blockIdx.x * warps + (threadIdx.x >> 5);
    const size_t N = (size_t)c.num_envs;
    if (threadIdx.x < LT_CORL_MAX_TERMS + 2) s_acc[threadIdx.x] = 0.0f;
    __syncthreads();
    if (wi < count) {
        const int i = q.work_env[wi], code = q.work_code[wi];
        if (code & 1) {                                                              // _resample_commands (LRV:759-840)
            const int n_cat = d.cmd.n_categories, nb = d.n_bins, nc = c.num_commands;
            const float u = lt_vel_u(a, i, 0);
            const double pc = 1.0 / (double)n_cat;
            int cat = n_cat - 1;
            for (int k = 0; k < n_cat; ++k)
                if ((float)(pc * k) <= u && u < (float)(pc * (k + 1))) { cat = k; break; }
            const double* w = q.weights + (size_t)cat * nb;
            // Curriculum.sample_bins: choice(p = w / sum).  Lane l owns the bins [l * per, (l + 1) * per)
            const int per = (nb + 31) / 32, g0 = lane * per, g1 = min(nb, g0 + per);
            double local = 0.0;
            for (int g = g0; g < g1; ++g) local += fmax(w[g], 0.0);
            double incl = local;
            for (int o = 1; o < 32; o <<= 1) {
                const double t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            const double total = __shfl_sync(0xffffffffu, incl, 31);
            const double target = (double)lt_vel_u(a, i, 1) * total;
            int bin = -1;
            if (target < incl && local > 0.0) {                                       // the draw may fall into this lane's bins
                double acc = incl - local;
                for (int g = g0; g < g1; ++g) {
                    if (w[g] <= 0.0) continue;
                    acc += w[g];
                    if (target < acc) { bin = g; break; }
                }
            }
            const unsigned hit = __ballot_sync(0xffffffffu, bin >= 0);
            if (hit) bin = __shfl_sync(0xffffffffu, bin, __ffs(hit) - 1);             // first lane in bin order
            else {                                                                    // rounding at the very end: the last positive bin
                int last = -1;
                for (int g = g0; g < g1; ++g) if (w[g] > 0.0) last = g;
                const unsigned has = __ballot_sync(0xffffffffu, last >= 0);
                bin = has ? __shfl_sync(0xffffffffu, last, 31 - __clz(has)) : 0;
            }
            float* cmd = q.commands + (size_t)i * nc;
            if (lane < nc && lane < d.n_dims) {                                       // sample_uniform_from_cell: uniform(c + s / 2, c - s / 2)
                const double ctr = q.grid[(size_t)lane * nb + bin], lo = ctr + q.bin_sizes[lane] / 2, hi = ctr - q.bin_sizes[lane] / 2;
                cmd[lane] = (float)(lo + (hi - lo) * (double)lt_vel_u(a, i, 2 + lane));
            }
            __syncwarp();
            if (lane == 0) {
                lt_cmd_rewrite(d.cmd, cmd, cat, lt_vel_u(a, i, 17));
                q.env_bins[i] = bin;
                q.env_categories[i] = cat;
            }
            for (int r = lane; r < c.n_terms + 5; r += 32) a.p.command_sums[(size_t)r * N + i] = 0.0f;
        }
        if (code & 2) {                                                              // _randomize_dof_props (LRV:663-673)
            if (d.randomize_motor_strength && lane < 12)
                q.motor_strengths[(size_t)i * 12 + lane] = LT_FADD(LT_FMUL(lt_vel_u(a, i, 18), d.motor_strength_span), d.motor_strength_lo);
            if (d.randomize_motor_offset && lane < 12)
                q.motor_offsets[(size_t)i * 12 + lane] = LT_FADD(LT_FMUL(lt_vel_u(a, i, 19 + lane), d.motor_offset_span), d.motor_offset_lo);
        }
        if (code & 4) {
            if (lane < 12) {                                                         // _reset_dofs (LRV:966-976) and the [N,12] buffers (:192-198)
                q.dof_state_rw[((size_t)i * 12 + lane) * 2] = LT_FMUL(a.p.default_dof_pos[lane], LT_FADD(LT_FMUL(1.0f, lt_vel_u(a, i, 31 + lane)), 0.5f));
                q.dof_state_rw[((size_t)i * 12 + lane) * 2 + 1] = 0.0f;
                a.p.last_actions[(size_t)i * 12 + lane] = 0.0f;
                a.p.last_last_actions[(size_t)i * 12 + lane] = 0.0f;
                a.p.last_dof_vel[(size_t)i * 12 + lane] = 0.0f;
            }
            if (lane < 4) q.feet_air_time[(size_t)i * 4 + lane] = 0.0f;
            if (lane == 0) {                                                         // _reset_root_states (LRV:983-1015)
                float* root = q.root_states_rw + (size_t)i * 13;
                for (int k = 0; k < 13; ++k) root[k] = d.base_init_state[k];
                for (int k = 0; k < 3; ++k) root[k] = LT_FADD(root[k], q.env_origins[(size_t)i * 3 + k]);
                if (d.custom_origins) {                                              // LRV:991-1001
                    root[0] = LT_FADD(root[0], LT_FADD(LT_FMUL(LT_FSUB(d.x_init_range, -d.x_init_range), lt_vel_u(a, i, 50)), -d.x_init_range));
                    root[1] = LT_FADD(root[1], LT_FADD(LT_FMUL(LT_FSUB(d.y_init_range, -d.y_init_range), lt_vel_u(a, i, 51)), -d.y_init_range));
                    root[0] = LT_FADD(root[0], d.x_init_offset);
                    root[1] = LT_FADD(root[1], d.y_init_offset);
                }
                const float yaw = LT_FADD(LT_FMUL(LT_FSUB(d.yaw_init_range, -d.yaw_init_range), lt_vel_u(a, i, 43)), -d.yaw_init_range);
                const float half = LT_FDIV(yaw, 2.0f);
                root[3] = 0.0f; root[4] = 0.0f; root[5] = sinf(half); root[6] = cosf(half);      // quat_from_angle_axis(yaw, z)
                for (int k = 0; k < 6; ++k) root[7 + k] = LT_FADD(LT_FMUL(1.0f, lt_vel_u(a, i, 44 + k)), -0.5f);
                a.p.episode_length_buf[i] = 0;
                q.gait_indices[i] = 0.0f;                                            // LRV:258
                atomicAdd(&s_acc[c.n_terms + 1], 1.0f);
            }
            for (int k = lane; k < d.lag_slots * 12; k += 32)                        // LRV:260-261
                q.lag_buffer[((size_t)(k / 12) * N + i) * 12 + (k % 12)] = 0.0f;
            for (int r = lane; r <= c.n_terms; r += 32) {                            // LRV:200-205
                atomicAdd(&s_acc[r], a.p.episode_sums[(size_t)r * N + i]);
                a.p.episode_sums[(size_t)r * N + i] = 0.0f;
            }
        }
    }
    __syncthreads();
    if (threadIdx.x <= c.n_terms + 1 && s_acc[threadIdx.x] != 0.0f) atomicAdd(q.episode_acc + threadIdx.x, s_acc[threadIdx.x]);
}

static bool lt_vel_cfg_ok(const lt_vel_cfg* c) {
    return c && c->num_envs > 0 && c->num_bodies > 0 && c->num_commands >= 3 && c->num_obs > 0 && c->obs_history_length >= 1 &&
           c->n_terms >= 0 && c->n_terms <= LT_CORL_MAX_TERMS && c->n_termination >= 0 && c->n_termination <= 8 && c->sample_interval > 0 &&
           c->rand_interval > 0;
}

extern "C" {

int lt_vel_measure_heights(const float* root_states, const float* height_points, const int16_t* height_samples, int32_t num_envs,
                           int32_t n_points, int32_t rows, int32_t cols, float border_size, float horizontal_scale, float vertical_scale,
                           float* measured_heights, float* height_mean, void* stream) {
    if (!root_states || !height_points || !height_samples || !measured_heights || num_envs <= 0 || n_points <= 0 || rows < 2 || cols < 2 ||
        !(horizontal_scale > 0.0f))
        return LT_ERR_INVALID;
    lt_vel_height_params a{root_states, height_points, height_samples, num_envs, n_points, rows, cols, border_size, horizontal_scale,
                           vertical_scale, measured_heights, height_mean};
    lt_vel_heights_kernel<<<(num_envs + 7) / 8, 256, 0, (cudaStream_t)stream>>>(a);
    return cudaGetLastError() == cudaSuccess ? LT_OK : LT_ERR_CUDA;
}

int lt_vel_pre(const lt_vel_cfg* cfg, const lt_vel_ptrs* ptrs, void* stream) {
    if (!lt_vel_cfg_ok(cfg) || !ptrs) return LT_ERR_INVALID;
    const lt_vel_ptrs& p = *ptrs;
    if (!p.root_states || !p.rigid_body_state || !p.contact_forces || !p.gravity_vec || !p.episode_length_buf || !p.base_pos || !p.base_quat ||
        !p.base_lin_vel || !p.base_ang_vel || !p.projected_gravity || !p.foot_positions || !p.foot_velocities || !p.prev_foot_velocities ||
        !p.reset_buf || !p.time_out_buf || !p.body_height_buf || !p.resample_due || !p.dr_due)
        return LT_ERR_INVALID;
    lt_vel_params a;
    a.c = *cfg; a.p = p; a.aux = nullptr; a.step = 0;
    lt_vel_pre_kernel<<<(cfg->num_envs + 127) / 128, 128, 0, (cudaStream_t)stream>>>(a);
    return cudaGetLastError() == cudaSuccess ? LT_OK : LT_ERR_CUDA;
}

int lt_vel_struct_bytes(int32_t which) {
    const size_t sz[5] = {sizeof(lt_vel_cfg), sizeof(lt_vel_ptrs), sizeof(lt_vel_dev_cfg), sizeof(lt_vel_dev_ptrs), sizeof(lt_cmd_cfg)};
    return which >= 0 && which < 5 ? (int)sz[which] : -1;
}

int lt_vel_indexed_paths(const lt_vel_cfg* cfg, const lt_vel_ptrs* ptrs, const lt_vel_dev_cfg* dcfg, const lt_vel_dev_ptrs* dptrs, int32_t which,
                         const float* uniform_rows, uint32_t step, void* stream) {
    if (!lt_vel_cfg_ok(cfg) || !ptrs || !dcfg || !dptrs || (which != LT_VEL_PATH_CALLBACK && which != LT_VEL_PATH_RESET)) return LT_ERR_INVALID;
    const lt_vel_dev_cfg& d = *dcfg;
    if (d.n_bins <= 0 || d.n_dims <= 0 || d.n_dims > LT_VEL_MAX_DIMS || d.n_dims < cfg->num_commands || d.n_keys < 0 || d.n_keys > 4 ||
        d.cmd.n_categories <= 0 || d.cmd.n_categories > LT_CMD_MAX_CATEGORIES || d.cmd.num_commands != cfg->num_commands || d.lag_slots < 0 || !(d.ep_len > 0.0f))
        return LT_ERR_INVALID;
    const void* const* pp = reinterpret_cast<const void* const*>(dptrs);
    for (size_t k = 0; k < sizeof(lt_vel_dev_ptrs) / sizeof(void*); ++k)
        if (!pp[k]) return LT_ERR_INVALID;
    const lt_vel_ptrs& p = *ptrs;
    if (!p.reset_buf || !p.resample_due || !p.dr_due || !p.command_sums || !p.episode_sums || !p.default_dof_pos || !p.last_actions ||
        !p.last_last_actions || !p.last_dof_vel || !p.episode_length_buf)
        return LT_ERR_INVALID;
    cudaStream_t s = (cudaStream_t)stream;
    lt_vel_dev_params a;
    a.c = *cfg; a.p = p; a.d = d; a.q = *dptrs; a.uniform_rows = uniform_rows; a.step = step; a.which = which;
    const int n = cfg->num_envs, nw = d.cmd.n_categories * d.n_bins;
    (void)nw;
    lt_vel_scan_kernel<<<(n + 255) / 256, 256, 0, s>>>(a);
    lt_vel_paths_kernel<<<(n + 7) / 8, 256, 0, s>>>(a);              // a warp per listed env; blocks past the list's end return at once
    return cudaGetLastError() == cudaSuccess ? LT_OK : LT_ERR_CUDA;
}

int lt_vel_reward(const lt_vel_cfg* cfg, const lt_vel_ptrs* ptrs, const float* term_values, void* stream) {
    if (!lt_vel_cfg_ok(cfg) || !ptrs) return LT_ERR_INVALID;
    const lt_vel_ptrs& p = *ptrs;
    if ((cfg->n_terms > 0 && !term_values) || !p.term_sums || !p.rew_buf || !p.rew_buf_pos || !p.rew_buf_neg || !p.episode_sums || !p.command_sums ||
        !p.base_lin_vel || !p.base_ang_vel || !p.commands)
        return LT_ERR_INVALID;
    cudaStream_t s = (cudaStream_t)stream;
    lt_vel_params a;
    a.c = *cfg; a.p = p; a.aux = term_values; a.step = 0;
    if (cfg->n_terms > 0) {
        if (cudaMemsetAsync(p.term_sums, 0, sizeof(double) * cfg->n_terms, s) != cudaSuccess) return LT_ERR_CUDA;
        int bx = (cfg->num_envs + 255) / 256;
        if (bx > 64) bx = 64;
        lt_vel_term_sums_kernel<<<dim3(bx, cfg->n_terms), 256, 0, s>>>(a);
    }
    lt_vel_reward_kernel<<<(cfg->num_envs + 127) / 128, 128, 0, s>>>(a);
    return cudaGetLastError() == cudaSuccess ? LT_OK : LT_ERR_CUDA;
}

int lt_vel_observe(const lt_vel_cfg* cfg, const lt_vel_ptrs* ptrs, const float* noise_uniforms, uint32_t step, void* stream) {
    if (!lt_vel_cfg_ok(cfg) || !ptrs) return LT_ERR_INVALID;
    const lt_vel_ptrs& p = *ptrs;
    const int width = 3 + (cfg->observe_command ? cfg->num_commands : 0) + 36 + (cfg->observe_two_prev_actions ? 12 : 0) +
                      (cfg->observe_timing_parameter ? 1 : 0) + (cfg->observe_clock_inputs ? 4 : 0);
    if (width != cfg->num_obs) return LT_ERR_INVALID;                                  // LRV:1119 (assert on the layout)
    if (cfg->num_privileged_obs != (cfg->priv_observe_friction ? 1 : 0) + (cfg->priv_observe_restitution ? 1 : 0)) return LT_ERR_INVALID;
    if (!p.projected_gravity || !p.commands || !p.commands_scale || !p.dof_state || !p.default_dof_pos || !p.actions || !p.last_actions ||
        !p.last_last_actions || !p.joint_pos_target || !p.last_joint_pos_target || !p.last_last_joint_pos_target || !p.last_dof_vel ||
        !p.last_root_vel || !p.root_states || !p.gait_indices || !p.clock_inputs || !p.obs_buf || !p.privileged_obs_buf || !p.obs_history_in ||
        !p.obs_history_out || p.obs_history_in == p.obs_history_out || (cfg->add_noise && !p.noise_scale_vec) ||
        (cfg->priv_observe_friction && !p.friction_coeffs) || (cfg->priv_observe_restitution && !p.restitutions))
        return LT_ERR_INVALID;
    cudaStream_t s = (cudaStream_t)stream;
    lt_vel_params a;
    a.c = *cfg; a.p = p; a.aux = noise_uniforms; a.step = step;
    const int row = cfg->obs_history_length * cfg->num_obs;
    const bool vec2 = (cfg->num_obs % 2) == 0 && (reinterpret_cast<uintptr_t>(p.obs_history_in) % 8) == 0 &&
                      (reinterpret_cast<uintptr_t>(p.obs_history_out) % 8) == 0 && (reinterpret_cast<uintptr_t>(p.obs_buf) % 8) == 0;
    (void)row;
    if (vec2) lt_vel_observe_kernel<2><<<cfg->num_envs, 256, 0, s>>>(a);
    else lt_vel_observe_kernel<1><<<cfg->num_envs, 256, 0, s>>>(a);
    lt_vel_shift_kernel<<<(unsigned)(((size_t)cfg->num_envs * 12 + 255) / 256), 256, 0, s>>>(a);      // after the observation read last_actions
    return cudaGetLastError() == cudaSuccess ? LT_OK : LT_ERR_CUDA;
}

}   // extern "C"
