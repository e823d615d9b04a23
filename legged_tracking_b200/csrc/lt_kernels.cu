// lt_kernels.cu -- sm_100a kernels of the per-environment step path + the C ABI of include/lt_b200.h.
//
// Kernels (LR = reference go1_gym/envs/base/legged_robot_trajectory_tracking.py):
//   lt_torque_kernel     _compute_torques + actuator net (LR:957-996, 1311-1320); one thread per (env, dof)
//   lt_step_kernel<EPW>  post_physics_step (LR:114-169) fused: each warp owns a tile of EPW consecutive envs;
//                        tile rows are staged in shared memory with 16-byte cp.async copies, one lane per env runs
//                        the scalar logic (lt_core.cuh), then the whole warp streams the wide rows (observation
//                        row incl. the heightfield gathers, history shifts) with coalesced accesses
//   lt_compact_kernel    reset_buf.nonzero() (LR:143), extras["time_outs"] latch (LR:289-291), step counters
//   lt_sign_kernel       all-env sums of the sign-indefinite reward terms (LR:332-335), only when configured
// There is no CPU fallback: lt_create fails without an sm_100 device.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>

#include "lt_core.cuh"

#define LT_ND 12                       // Go1: 12 actuated dofs = 12 actions (validated in lt_create)
#define LT_STEP_THREADS 128
#define LT_TORQUE_THREADS 128
#define LT_COMPACT_THREADS 1024

// ---------------------------------------------------------------------------------------------------------------
struct lt_step_params {
    lt_config c;
    lt_layout L;
    lt_state_ptrs st;
    lt_sim_ptrs sim;
    lt_terrain_ptrs ter;
    lt_step_args args;               // device pointers only (new_gravity is copied into sc)
    lt_step_scalars sc;
    const float* noise_scale_vec;    // device [num_obs]
    const int32_t* counters;         // device int32[4]
    int32_t smem_floats_per_warp;
};

struct lt_torque_params {
    lt_config c;
    lt_act_weights W;
    const float* actions_raw;
    const float* dof_state;
    lt_state_ptrs st;
    const int32_t* counters;
    int32_t substep;
};

struct lt_handle {
    lt_config c;
    lt_layout L;
    lt_act_weights W;
    lt_state_ptrs st;
    lt_sim_ptrs sim;
    lt_terrain_ptrs ter;
    int32_t n_tiles;
    bool have_state, have_sim, have_terrain;
    float* d_noise_scale;
    float* d_term_sums;
    int32_t* d_counters;
    int32_t substeps_pending;        // lt_compute_torques calls since the last lt_post_physics_step
    bool sign_prepass_done;
    int epw;
    int device;
    std::string err;
};

static std::string g_create_error;

static int lt_fail(lt_handle* h, int code, const std::string& msg) {
    if (h) h->err = msg; else g_create_error = msg;
    return code;
}
#define LT_CUDA(h, call)                                                                        \
    do {                                                                                        \
        cudaError_t e_ = (call);                                                                \
        if (e_ != cudaSuccess)                                                                  \
            return lt_fail((h), LT_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

// ---------------------------------------------------------------------------------------------------------------
// device helpers
__device__ __forceinline__ void lt_cp_async16(void* smem, const void* gmem) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void lt_cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
}
// stage `n` floats (16-byte aligned source, n % 4 == 0 when `vec`) global -> shared, one warp
__device__ __forceinline__ void lt_tile_load(float* s, const float* g, int n, int lane, bool vec) {
    if (vec) {
        for (int k = lane * 4; k < n; k += 128) lt_cp_async16(s + k, g + k);
    } else {
        for (int k = lane; k < n; k += 32) s[k] = g[k];
    }
}
__device__ __forceinline__ void lt_tile_store(float* g, const float* s, int n, int lane, bool vec) {
    if (vec) {
        for (int k = lane * 4; k < n; k += 128) *reinterpret_cast<float4*>(g + k) = *reinterpret_cast<const float4*>(s + k);
    } else {
        for (int k = lane; k < n; k += 32) g[k] = s[k];
    }
}
__device__ __forceinline__ bool lt_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ---------------------------------------------------------------------------------------------------------------
// _compute_torques: one thread per (env, dof)
__global__ void __launch_bounds__(LT_TORQUE_THREADS) lt_torque_kernel(const __grid_constant__ lt_torque_params p) {
    const lt_config& c = p.c;
    const size_t n = (size_t)c.num_envs * LT_ND;
    const size_t s = (size_t)blockIdx.x * LT_TORQUE_THREADS + threadIdx.x;
    if (s >= n) return;
    const int dof = (int)(s % LT_ND);
    lt_torque_io io;
    io.action_raw = p.actions_raw[s];
    const float2 ds = reinterpret_cast<const float2*>(p.dof_state)[s];
    io.dof_pos = ds.x;
    io.dof_vel = ds.y;
    io.action_scaled = lt_scaled_action(c, io.action_raw, dof, io.action_clipped);
    io.lag_out = 0.0f;
    if (c.randomize_lag_timesteps) {
        // FIFO of lag_timesteps+1 slots (LR:972-974): push at the old head, read the new head
        const int Lg = c.lag_timesteps + 1;
        const int head = (p.counters[2] + p.substep) % Lg;
        const int rd = (head + 1) % Lg;
        p.st.lag_buffer[(size_t)head * n + s] = io.action_scaled;
        io.lag_out = (rd == head) ? io.action_scaled : p.st.lag_buffer[(size_t)rd * n + s];
    }
    io.motor_offset = p.st.motor_offsets[s];
    io.motor_strength = p.st.motor_strengths[s];
    const bool net = c.control_type == LT_CONTROL_ACTUATOR_NET;
    if (net) {
        io.err_last = p.st.joint_pos_err_last[s];
        io.err_last_last = p.st.joint_pos_err_last_last[s];
        io.vel_last = p.st.joint_vel_last[s];
        io.vel_last_last = p.st.joint_vel_last_last[s];
        io.kp_factor = io.kd_factor = 1.0f;
    } else {
        io.kp_factor = p.st.Kp_factors[s];
        io.kd_factor = p.st.Kd_factors[s];
        io.err_last = io.err_last_last = io.vel_last = io.vel_last_last = 0.0f;
    }
    lt_torque_sample(c, p.W, dof, io);
    if (net) {
        p.st.joint_pos_err_last[s] = io.err_last;
        p.st.joint_pos_err_last_last[s] = io.err_last_last;
        p.st.joint_vel_last[s] = io.vel_last;
        p.st.joint_vel_last_last[s] = io.vel_last_last;
    }
    if (p.substep == 0) p.st.actions[s] = io.action_clipped;
    p.st.joint_pos_target[s] = io.joint_pos_target;
    p.st.torques[s] = io.torque;
}

// ---------------------------------------------------------------------------------------------------------------
// post_physics_step, fused.  Shared memory per warp (floats), all regions 16-byte aligned for EPW % 4 == 0:
//   root[EPW*A*13] dof[EPW*24] contact[EPW*B*3] actions[EPW*12] last_actions[EPW*12] last_dof_vel[EPW*12]
//   torques[EPW*12] head[EPW*head_w4] ix[EPW*nxf] iy[EPW*ny] meta[EPW*4] (reset, z_post, cam_z, tile)
struct lt_smem_plan {
    int root, dof, contact, actions, last_actions, last_dof_vel, torques, head, ix, iy, meta, total, head_w4;
};
__host__ __device__ inline int lt_align4(int x) { return (x + 3) & ~3; }
__host__ __device__ inline lt_smem_plan lt_plan_smem(const lt_config& c, const lt_layout& L, int epw) {
    lt_smem_plan P;
    int o = 0;
    P.root = o; o += lt_align4(epw * c.num_actors * 13);
    P.dof = o; o += lt_align4(epw * LT_ND * 2);
    P.contact = o; o += lt_align4(epw * c.num_bodies * 3);
    P.actions = o; o += lt_align4(epw * LT_ND);
    P.last_actions = o; o += lt_align4(epw * LT_ND);
    P.last_dof_vel = o; o += lt_align4(epw * LT_ND);
    P.torques = o; o += lt_align4(epw * LT_ND);
    P.head_w4 = L.head_w | 1;                       // odd stride: conflict-free lane-per-env writes
    P.head = o; o += lt_align4(epw * P.head_w4);
    P.ix = o; o += lt_align4(epw * (L.nxf > 0 ? L.nxf : 1));
    P.iy = o; o += lt_align4(epw * (c.n_points_y > 0 ? c.n_points_y : 1));
    P.meta = o; o += lt_align4(epw * 4);
    P.total = o;
    return P;
}

template <int EPW>
__global__ void __launch_bounds__(LT_STEP_THREADS) lt_step_kernel(const __grid_constant__ lt_step_params p) {
    extern __shared__ __align__(16) float smem[];
    const lt_config& c = p.c;
    const lt_layout& L = p.L;
    const lt_state_ptrs& st = p.st;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    constexpr int WARPS = LT_STEP_THREADS / 32;
    const int N = c.num_envs;
    const int A13 = c.num_actors * 13, B3 = c.num_bodies * 3;

    // per-CTA lookup: height column -> (layer, x row, y column) packed
    uint32_t* lut = reinterpret_cast<uint32_t*>(smem);
    const int lut_words = lt_align4(L.n_heights);
    for (int k = threadIdx.x; k < L.n_heights; k += LT_STEP_THREADS) {
        const int per_layer = L.nxf * c.n_points_y;
        const int layer = k / per_layer, rem = k - layer * per_layer;
        const int a = rem / c.n_points_y, b = rem - a * c.n_points_y;
        lut[k] = (uint32_t)a | ((uint32_t)b << 8) | ((uint32_t)layer << 16);
    }
    __syncthreads();

    const lt_smem_plan P = lt_plan_smem(c, L, EPW);
    float* ws = smem + lut_words + (size_t)warp * P.total;
    float* s_root = ws + P.root;
    float* s_dof = ws + P.dof;
    float* s_contact = ws + P.contact;
    float* s_actions = ws + P.actions;
    float* s_last_actions = ws + P.last_actions;
    float* s_last_dof_vel = ws + P.last_dof_vel;
    float* s_torques = ws + P.torques;
    float* s_head = ws + P.head;
    int32_t* s_ix = reinterpret_cast<int32_t*>(ws + P.ix);
    int32_t* s_iy = reinterpret_cast<int32_t*>(ws + P.iy);
    float* s_meta = ws + P.meta;

    const int tile = blockIdx.x * WARPS + warp;
    const int env0 = tile * EPW;
    if (env0 >= N) return;
    const int nv = min(EPW, N - env0);            // valid envs in this tile
    const bool full = (nv == EPW) && (EPW % 4 == 0);
    const uint32_t step = (uint32_t)p.counters[0];

    // ---- phase 0: stage the tile rows (flat, contiguous in every source tensor)
    lt_tile_load(s_root, p.sim.root_states + (size_t)env0 * A13, nv * A13, lane, full);
    lt_tile_load(s_dof, p.sim.dof_state + (size_t)env0 * LT_ND * 2, nv * LT_ND * 2, lane, full);
    lt_tile_load(s_contact, p.sim.contact_forces + (size_t)env0 * B3, nv * B3, lane, full);
    lt_tile_load(s_actions, st.actions + (size_t)env0 * LT_ND, nv * LT_ND, lane, full);
    lt_tile_load(s_last_actions, st.last_actions + (size_t)env0 * LT_ND, nv * LT_ND, lane, full);
    lt_tile_load(s_last_dof_vel, st.last_dof_vel + (size_t)env0 * LT_ND, nv * LT_ND, lane, full);
    lt_tile_load(s_torques, st.torques + (size_t)env0 * LT_ND, nv * LT_ND, lane, full);
    lt_cp_async_wait_all();
    __syncwarp();

    // ---- phase 1: one lane per env, the scalar logic of the step
    if (lane < nv) {
        const int i = env0 + lane;
        lt_env_rows R;
        R.root = s_root + lane * A13;
        R.arrow = (c.num_actors > 1) ? s_root + lane * A13 + (c.num_actors - 1) * 13 : nullptr;
        R.dof = s_dof + lane * LT_ND * 2;
        R.contact = s_contact + lane * B3;
        R.actions = s_actions + lane * LT_ND;
        R.last_actions = s_last_actions + lane * LT_ND;
        R.last_dof_vel = s_last_dof_vel + lane * LT_ND;
        R.torques = s_torques + lane * LT_ND;
        R.head = s_head + lane * P.head_w4;
        R.ix = s_ix + lane * (L.nxf > 0 ? L.nxf : 1);
        R.iy = s_iy + lane * (c.n_points_y > 0 ? c.n_points_y : 1);
        lt_step_scalars sc = p.sc;
        sc.step = step;
        const lt_env_result r = lt_env_step(c, L, st, p.ter, p.args, sc, i, R);
        s_meta[lane * 4 + 0] = r.reset ? 1.0f : 0.0f;
        s_meta[lane * 4 + 1] = r.z_post;
        s_meta[lane * 4 + 2] = r.cam_z;
        s_meta[lane * 4 + 3] = __int_as_float((c.observe_heights && !c.plane) ? p.ter.env_tile[i] : 0);
    }
    __syncwarp();
    const unsigned reset_mask = __ballot_sync(0xffffffffu, lane < nv && s_meta[lane * 4] != 0.0f);

    // ---- phase 2: write back the sim rows the step owns (arrow pose every step; robot root / dofs on reset)
    if (c.num_actors > 1 || reset_mask) lt_tile_store(p.sim.root_states + (size_t)env0 * A13, s_root, nv * A13, lane, full);
    if (reset_mask) lt_tile_store(p.sim.dof_state + (size_t)env0 * LT_ND * 2, s_dof, nv * LT_ND * 2, lane, full);

    // ---- phase 3: history shifts (LR:148-153), float4 over the [EPW, 12] tiles
    {
        const size_t g0 = (size_t)env0 * LT_ND;
        for (int k = lane * 4; k < nv * LT_ND; k += 128) {
            const int e = k / LT_ND, d = k - e * LT_ND;
            const bool rs = (reset_mask >> e) & 1u;
            float4 la = *reinterpret_cast<const float4*>(s_last_actions + k);
            if (rs) la = make_float4(0.f, 0.f, 0.f, 0.f);          // last_actions[env_ids] = 0 before the shift (LR:246)
            *reinterpret_cast<float4*>(st.last_last_actions + g0 + k) = la;
            *reinterpret_cast<float4*>(st.last_actions + g0 + k) = *reinterpret_cast<const float4*>(s_actions + k);
            const float4 jt = *reinterpret_cast<const float4*>(st.joint_pos_target + g0 + k);
            const float4 ljt = *reinterpret_cast<const float4*>(st.last_joint_pos_target + g0 + k);
            *reinterpret_cast<float4*>(st.last_last_joint_pos_target + g0 + k) = ljt;
            *reinterpret_cast<float4*>(st.last_joint_pos_target + g0 + k) = jt;
            const float* dv = s_dof + e * LT_ND * 2 + 2 * d + 1;   // post-reset dof_vel
            *reinterpret_cast<float4*>(st.last_dof_vel + g0 + k) = make_float4(dv[0], dv[2], dv[4], dv[6]);
        }
        for (int k = lane; k < nv * 6; k += 32) {
            const int e = k / 6, j = k - e * 6;
            st.last_root_vel[(size_t)env0 * 6 + k] = s_root[e * A13 + 7 + j];
        }
    }

    // ---- phase 4: observation rows (LR:357-475) + obs_history (history_wrapper.py:18-24); warp per env
    const int no = c.num_obs;
    const int H = c.obs_history_length;
    const bool plane = c.plane != 0;
    for (int e = 0; e < nv; ++e) {
        const int i = env0 + e;
        const float z_post = s_meta[e * 4 + 1], cam_z = s_meta[e * 4 + 2];
        const int tile_id = __float_as_int(s_meta[e * 4 + 3]);
        const float* tile_base = plane ? nullptr : p.ter.tile_heights + (size_t)tile_id * 2 * L.tile_cells;
        const float* head = s_head + e * P.head_w4;
        const int32_t* ix = s_ix + e * (L.nxf > 0 ? L.nxf : 1);
        const int32_t* iy = s_iy + e * (c.n_points_y > 0 ? c.n_points_y : 1);
        float* obs_row = st.obs_buf + (size_t)i * no;
        float* hist_row = st.obs_history + (size_t)i * no * H;
        if (H > 1) {
            const int keep = (H - 1) * no;
            for (int k = lane; k < keep; k += 32) { const float v = hist_row[k + no]; hist_row[k] = v; }
        }
        float* hist_new = hist_row + (size_t)(H - 1) * no;
        const float* nz = (c.add_noise && p.args.noise_uniforms) ? p.args.noise_uniforms + (size_t)i * no : nullptr;
        uint32_t rnd[4];
        int it = 0;
        for (int col = lane; col < no; col += 32, ++it) {
            float v;
            const int hc = col - L.h0;
            if (hc >= 0 && hc < L.n_heights) {
                const uint32_t w = lut[hc];
                const int a = w & 0xff, b = (w >> 8) & 0xff, layer = w >> 16;
                float h;
                if (plane) h = (layer == 0) ? 1.0f : 0.0f;          // LR:1928-1932
                else h = __ldg(tile_base + (size_t)layer * L.tile_cells + ix[a] * c.tile_cols + iy[b]);
                v = lt_height_obs(c, h, z_post, cam_z);
            } else {
                v = head[hc < 0 ? col : col - L.n_heights];
            }
            float u = 0.5f;
            if (c.add_noise) {
                if (nz) u = nz[col];
                else {
                    if ((it & 3) == 0) lt_philox((uint32_t)i, (uint32_t)(lane + 32 * (it >> 2)) | LT_STREAM_NOISE, step, 0u,
                                                 c.seed_lo, c.seed_hi, rnd);
                    u = lt_u01(rnd[it & 3]);
                }
            }
            v = lt_obs_finish(c, v, u, c.add_noise ? p.noise_scale_vec[col] : 0.0f);
            obs_row[col] = v;
            hist_new[col] = v;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// ordered compaction of reset_buf + time-out latch + counters
__global__ void __launch_bounds__(LT_COMPACT_THREADS) lt_compact_kernel(const uint8_t* __restrict__ reset_buf,
                                                                        const uint8_t* __restrict__ time_out_buf,
                                                                        uint8_t* __restrict__ latched,
                                                                        int64_t* __restrict__ ids, int32_t* __restrict__ count,
                                                                        int32_t* __restrict__ counters, int n, int send_timeouts,
                                                                        int n_substeps, int lag_slots) {
    __shared__ int s_base[32], s_total[32], s_warp[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int start = blockIdx.x * LT_COMPACT_THREADS;
    int base = 0, total = 0;
    const int words = n >> 2;
    const uint32_t* rw = reinterpret_cast<const uint32_t*>(reset_buf);
    for (int k = tid; k < words; k += LT_COMPACT_THREADS) {
        const int cnt = __popc(rw[k] & 0x01010101u);
        total += cnt;
        if (4 * k < start) base += cnt;
    }
    if (tid == 0)
        for (int k = words * 4; k < n; ++k) { total += reset_buf[k] ? 1 : 0; if (k < start) base += reset_buf[k] ? 1 : 0; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { base += __shfl_xor_sync(0xffffffffu, base, o); total += __shfl_xor_sync(0xffffffffu, total, o); }
    if (lane == 0) { s_base[warp] = base; s_total[warp] = total; }
    const int env = start + tid;
    const bool flag = env < n && reset_buf[env] != 0;
    const unsigned bal = __ballot_sync(0xffffffffu, flag);
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    if (warp == 0) {
        int b = s_base[lane], t = s_total[lane], w = s_warp[lane];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { b += __shfl_xor_sync(0xffffffffu, b, o); t += __shfl_xor_sync(0xffffffffu, t, o); }
        int incl = w;                                   // inclusive scan of per-warp counts
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
        s_warp[lane] = incl - w;
        if (lane == 0) { s_base[0] = b; s_total[0] = t; }
    }
    __syncthreads();
    base = s_base[0];
    total = s_total[0];
    if (flag) ids[base + s_warp[warp] + __popc(bal & ((1u << lane) - 1u))] = env;
    if (send_timeouts && total > 0 && env < n) latched[env] = time_out_buf[env];
    if (blockIdx.x == 0 && tid == 0) {
        count[0] = total;
        counters[0] += 1;
        counters[1] += n_substeps;
        counters[2] = (counters[2] + n_substeps) % lag_slots;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// all-env sums of sign-indefinite reward terms (exploration, feet_air_time, external), no side effects
__global__ void __launch_bounds__(256) lt_sign_kernel(const __grid_constant__ lt_step_params p, float* __restrict__ sums) {
    const lt_config& c = p.c;
    const int i = blockIdx.x * 256 + threadIdx.x;
    __shared__ float s_part[8];
    for (int t = 0; t < c.n_terms; ++t) {
        if (c.term_bucket[t] != LT_BUCKET_BY_SUM) continue;
        float v = 0.0f;
        if (i < c.num_envs) {
            const float* root = p.sim.root_states + (size_t)i * c.num_actors * 13;
            lt_pose P;
            lt_derive_pose(p.st, c, i, root, p.st.curr_pose_index[i], P);
            lt_reward_ctx rc;
            rc.dof = p.sim.dof_state + (size_t)i * LT_ND * 2;
            rc.last_dof_vel = p.st.last_dof_vel + (size_t)i * LT_ND;
            rc.torques = p.st.torques + (size_t)i * LT_ND;
            rc.actions = p.st.actions + (size_t)i * LT_ND;
            rc.last_actions = p.st.last_actions + (size_t)i * LT_ND;
            rc.contact = p.sim.contact_forces + (size_t)i * c.num_bodies * 3;
            rc.root_z = root[2];
            for (int k = 0; k < 3; ++k) { rc.blv[k] = P.blv[k]; rc.bav[k] = P.bav[k]; rc.pg[k] = P.pg[k];
                                          rc.rel_lin[k] = P.rel_lin[k]; rc.rel_rot[k] = P.rel_rot[k]; rc.loc_lin[k] = P.rel_lin[k]; }
            rc.ep_f = (float)(p.st.episode_length_buf[i] + 1);
            rc.reached = 0;
            rc.n_collisions = 0;
            v = lt_reward_term(c, p.st, i, c.term_ids[t], rc, p.args.extra_reward, true) * c.term_scales[t];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = v;
        __syncthreads();
        if (threadIdx.x == 0) {
            float s = 0.0f;
            for (int w = 0; w < 8; ++w) s += s_part[w];
            atomicAdd(sums + t, s);
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------------------------
// launch helper (C++ linkage)
template <int EPW>
static cudaError_t lt_launch_step(const lt_step_params& p, cudaStream_t stream) {
    constexpr int WARPS = LT_STEP_THREADS / 32;
    const lt_smem_plan P = lt_plan_smem(p.c, p.L, EPW);
    const size_t smem = sizeof(float) * ((size_t)lt_align4(p.L.n_heights) + (size_t)WARPS * P.total);
    static size_t configured = 0;
    if (smem > 48 * 1024 && smem > configured) {
        cudaError_t e = cudaFuncSetAttribute(lt_step_kernel<EPW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured = smem;
    }
    const int tiles = (p.c.num_envs + EPW - 1) / EPW;
    lt_step_kernel<EPW><<<(tiles + WARPS - 1) / WARPS, LT_STEP_THREADS, smem, stream>>>(p);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------
// C ABI
extern "C" {

int lt_abi_version(void) { return LT_ABI_VERSION; }
int lt_config_bytes(void) { return (int)sizeof(lt_config); }
int lt_num_uniform_slots(const lt_config* cfg) { return cfg ? lt_n_slots(*cfg) : 0; }
const char* lt_last_error(const lt_handle* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

static int lt_validate(const lt_config& c, std::string& why) {
    char buf[256];
#define LT_REQ(cond)                                                     \
    if (!(cond)) { snprintf(buf, sizeof buf, "lt_config: requirement failed: %s", #cond); why = buf; return 0; }
    LT_REQ(c.struct_bytes == (int)sizeof(lt_config));
    LT_REQ(c.abi_version == LT_ABI_VERSION);
    LT_REQ(c.num_envs > 0);
    LT_REQ(c.num_dof == LT_ND && c.num_actions == LT_ND);
    LT_REQ(c.num_actors >= 1 && c.num_actors <= 4);
    LT_REQ(c.num_bodies >= 1 && c.num_bodies <= LT_MAX_BODIES);
    LT_REQ(c.n_feet >= 0 && c.n_feet <= LT_MAX_FEET);
    LT_REQ(c.n_terms >= 0 && c.n_terms <= LT_MAX_TERMS);
    LT_REQ(c.n_points_x >= 0 && c.n_points_x <= LT_MAX_POINTS && c.n_points_y >= 0 && c.n_points_y <= LT_MAX_POINTS);
    LT_REQ(c.obs_history_length >= 1);
    LT_REQ(c.lag_timesteps >= 0);
    LT_REQ(c.traj_length >= 1);
    LT_REQ(c.traj_function >= LT_TRAJ_FIXED_TARGET && c.traj_function <= LT_TRAJ_RANDOM_TARGET);
    LT_REQ(c.traj_function != LT_TRAJ_RANDOM_TARGET || (c.num_interpolation >= 1 && c.traj_length % c.num_interpolation == 0));
    LT_REQ(c.control_type == LT_CONTROL_ACTUATOR_NET || c.control_type == LT_CONTROL_P);
    LT_REQ(c.rand_interval >= 1);
    LT_REQ(c.plane || !c.observe_heights || (c.tile_rows >= 2 && c.tile_cols >= 2));
    LT_REQ(c.envs_per_warp == 0 || c.envs_per_warp == 4 || c.envs_per_warp == 8 || c.envs_per_warp == 16 || c.envs_per_warp == 32);
#undef LT_REQ
    lt_layout L;
    lt_make_layout(c, L);
    if (lt_obs_total_width(L) != c.num_obs) {
        snprintf(buf, sizeof buf, "observation layout is %d wide but num_obs = %d (LR:475)", lt_obs_total_width(L), c.num_obs);
        why = buf;
        return 0;
    }
    int pw = c.priv_observe_friction + c.priv_observe_restitution + c.priv_observe_base_mass + 3 * c.priv_observe_com_displacement +
             LT_ND * c.priv_observe_motor_strength + LT_ND * c.priv_observe_motor_offset + c.priv_observe_body_height +
             3 * c.priv_observe_body_velocity + 3 * c.priv_observe_gravity;
    if (pw != c.num_privileged_obs) {
        snprintf(buf, sizeof buf, "privileged layout is %d wide but num_privileged_obs = %d (LR:589)", pw, c.num_privileged_obs);
        why = buf;
        return 0;
    }
    return 1;
}

int lt_create(const lt_config* cfg, const lt_tables* tables, lt_handle** out) {
    if (!cfg || !tables || !out) return lt_fail(nullptr, LT_ERR_INVALID, "lt_create: null argument");
    *out = nullptr;
    std::string why;
    if (cfg->struct_bytes != (int)sizeof(lt_config) || !lt_validate(*cfg, why))
        return lt_fail(nullptr, LT_ERR_INVALID, why.empty() ? "lt_config.struct_bytes mismatch (ABI drift)" : why);
    int dev = 0, ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return lt_fail(nullptr, LT_ERR_NO_DEVICE, "no CUDA device: liblt_b200 has no CPU fallback");
    LT_CUDA(nullptr, cudaGetDevice(&dev));
    cudaDeviceProp prop;
    LT_CUDA(nullptr, cudaGetDeviceProperties(&prop, dev));
    if (prop.major < 10)
        return lt_fail(nullptr, LT_ERR_NO_DEVICE, std::string("device ") + prop.name + " is not sm_100: liblt_b200 is built for sm_100a only");
    lt_handle* h = new lt_handle();
    h->c = *cfg;
    lt_make_layout(h->c, h->L);
    h->device = dev;
    memset(&h->st, 0, sizeof h->st);
    memset(&h->sim, 0, sizeof h->sim);
    memset(&h->ter, 0, sizeof h->ter);
    h->have_state = h->have_sim = h->have_terrain = false;
    h->substeps_pending = 0;
    h->sign_prepass_done = false;
    h->n_tiles = 0;
    memcpy(h->W.w0, tables->act_w0, sizeof h->W.w0);
    memcpy(h->W.b0, tables->act_b0, sizeof h->W.b0);
    memcpy(h->W.w1, tables->act_w1, sizeof h->W.w1);
    memcpy(h->W.b1, tables->act_b1, sizeof h->W.b1);
    memcpy(h->W.w2, tables->act_w2, sizeof h->W.w2);
    memcpy(h->W.b2, tables->act_b2, sizeof h->W.b2);
    // envs per warp: small shards need more warps in flight than lanes-per-env efficiency
    int epw = cfg->envs_per_warp;
    if (epw == 0) {
        const int per_sm = (cfg->num_envs + prop.multiProcessorCount - 1) / prop.multiProcessorCount;
        epw = per_sm >= 384 ? 16 : (per_sm >= 96 ? 8 : 4);
    }
    h->epw = epw;
    cudaError_t e = cudaMalloc(&h->d_noise_scale, sizeof(float) * (size_t)cfg->num_obs);
    if (e == cudaSuccess) e = cudaMemcpy(h->d_noise_scale, tables->noise_scale_vec, sizeof(float) * (size_t)cfg->num_obs, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMalloc(&h->d_term_sums, sizeof(float) * LT_MAX_TERMS);
    if (e == cudaSuccess) e = cudaMemset(h->d_term_sums, 0, sizeof(float) * LT_MAX_TERMS);
    if (e == cudaSuccess) e = cudaMalloc(&h->d_counters, sizeof(int32_t) * 4);
    if (e == cudaSuccess) e = cudaMemset(h->d_counters, 0, sizeof(int32_t) * 4);
    if (e != cudaSuccess) {
        const std::string msg = std::string("lt_create: ") + cudaGetErrorString(e);
        delete h;
        return lt_fail(nullptr, LT_ERR_CUDA, msg);
    }
    *out = h;
    return LT_OK;
}

void lt_destroy(lt_handle* h) {
    if (!h) return;
    cudaFree(h->d_noise_scale);
    cudaFree(h->d_term_sums);
    cudaFree(h->d_counters);
    delete h;
}

int lt_bind_sim(lt_handle* h, const lt_sim_ptrs* sim) {
    if (!h || !sim) return LT_ERR_INVALID;
    if (!sim->root_states || !sim->dof_state || !sim->contact_forces) return lt_fail(h, LT_ERR_INVALID, "lt_bind_sim: null tensor");
    if (((uintptr_t)sim->root_states | (uintptr_t)sim->dof_state | (uintptr_t)sim->contact_forces) & 15u)
        return lt_fail(h, LT_ERR_INVALID, "lt_bind_sim: tensors must be 16-byte aligned");
    h->sim = *sim;
    h->have_sim = true;
    return LT_OK;
}

int lt_bind_terrain(lt_handle* h, const lt_terrain_ptrs* t, int32_t n_tiles) {
    if (!h || !t) return LT_ERR_INVALID;
    if (!t->env_origins) return lt_fail(h, LT_ERR_INVALID, "lt_bind_terrain: env_origins is required");
    if (h->c.observe_heights && !h->c.plane && (!t->tile_heights || !t->env_tile || !t->env_terrain_origin || n_tiles <= 0))
        return lt_fail(h, LT_ERR_INVALID, "lt_bind_terrain: tile table required unless mesh_type is plane");
    h->ter = *t;
    h->n_tiles = n_tiles;
    h->have_terrain = true;
    return LT_OK;
}

int lt_bind_state(lt_handle* h, const lt_state_ptrs* s) {
    if (!h || !s) return LT_ERR_INVALID;
    const void* const* p = reinterpret_cast<const void* const*>(s);
    const size_t n = sizeof(lt_state_ptrs) / sizeof(void*);
    // optional members: command_sums, height_px, height_py
    for (size_t k = 0; k < n; ++k) {
        const void* const* opt1 = reinterpret_cast<const void* const*>(&s->command_sums);
        const void* const* opt2 = reinterpret_cast<const void* const*>(&s->height_px);
        const void* const* opt3 = reinterpret_cast<const void* const*>(&s->height_py);
        if (p + k == opt1 || p + k == opt2 || p + k == opt3) continue;
        if (!p[k]) {
            char buf[96];
            snprintf(buf, sizeof buf, "lt_bind_state: pointer member #%zu of lt_state_ptrs is null", k);
            return lt_fail(h, LT_ERR_INVALID, buf);
        }
    }
    const uintptr_t al = (uintptr_t)s->actions | (uintptr_t)s->last_actions | (uintptr_t)s->last_last_actions |
                         (uintptr_t)s->joint_pos_target | (uintptr_t)s->last_joint_pos_target |
                         (uintptr_t)s->last_last_joint_pos_target | (uintptr_t)s->last_dof_vel | (uintptr_t)s->torques |
                         (uintptr_t)s->reset_buf;
    if (al & 15u) return lt_fail(h, LT_ERR_INVALID, "lt_bind_state: [N,12] state arrays and reset_buf must be 16-byte aligned");
    h->st = *s;
    h->have_state = true;
    return LT_OK;
}

int32_t* lt_counters(lt_handle* h) { return h ? h->d_counters : nullptr; }
float* lt_term_sum_buffer(lt_handle* h) { return h ? h->d_term_sums : nullptr; }

static int lt_ready(lt_handle* h) {
    if (!h) return LT_ERR_INVALID;
    if (!h->have_state || !h->have_sim || !h->have_terrain)
        return lt_fail(h, LT_ERR_UNBOUND, "bind sim, terrain and state before stepping (lt_bind_sim / lt_bind_terrain / lt_bind_state)");
    return LT_OK;
}

int lt_compute_torques(lt_handle* h, const float* actions, int32_t substep, void* stream) {
    int rc = lt_ready(h);
    if (rc) return rc;
    if (!actions) return lt_fail(h, LT_ERR_INVALID, "lt_compute_torques: actions is null");
    lt_torque_params p;
    p.c = h->c;
    p.W = h->W;
    p.actions_raw = actions;
    p.dof_state = h->sim.dof_state;
    p.st = h->st;
    p.counters = h->d_counters;
    p.substep = substep;
    const size_t n = (size_t)h->c.num_envs * LT_ND;
    const unsigned grid = (unsigned)((n + LT_TORQUE_THREADS - 1) / LT_TORQUE_THREADS);
    lt_torque_kernel<<<grid, LT_TORQUE_THREADS, 0, (cudaStream_t)stream>>>(p);
    LT_CUDA(h, cudaGetLastError());
    h->substeps_pending += 1;
    return LT_OK;
}

static void lt_fill_step_params(lt_handle* h, const lt_step_args* args, lt_step_params& p) {
    p.c = h->c;
    p.L = h->L;
    p.st = h->st;
    p.sim = h->sim;
    p.ter = h->ter;
    memset(&p.args, 0, sizeof p.args);
    memset(&p.sc, 0, sizeof p.sc);
    if (args) {
        p.args.extra_reward = args->extra_reward;
        p.args.uniforms = args->uniforms;
        p.args.noise_uniforms = args->noise_uniforms;
        if (args->new_gravity) {
            p.sc.set_gravity = 1;
            for (int k = 0; k < 3; ++k) p.sc.new_gravity[k] = args->new_gravity[k];
        }
    }
    p.sc.term_sums = h->d_term_sums;
    p.noise_scale_vec = h->d_noise_scale;
    p.counters = h->d_counters;
}

int lt_reward_sign_prepass(lt_handle* h, const lt_step_args* args, void* stream) {
    int rc = lt_ready(h);
    if (rc) return rc;
    if (!h->L.any_bucket_by_sum) return LT_OK;
    lt_step_params p;
    lt_fill_step_params(h, args, p);
    LT_CUDA(h, cudaMemsetAsync(h->d_term_sums, 0, sizeof(float) * LT_MAX_TERMS, (cudaStream_t)stream));
    lt_sign_kernel<<<(h->c.num_envs + 255) / 256, 256, 0, (cudaStream_t)stream>>>(p, h->d_term_sums);
    LT_CUDA(h, cudaGetLastError());
    h->sign_prepass_done = true;
    return LT_OK;
}

int lt_post_physics_step(lt_handle* h, const lt_step_args* args, void* stream) {
    int rc = lt_ready(h);
    if (rc) return rc;
    if (h->L.any_bucket_by_sum && !h->sign_prepass_done) {
        rc = lt_reward_sign_prepass(h, args, stream);
        if (rc) return rc;
    }
    h->sign_prepass_done = false;
    lt_step_params p;
    lt_fill_step_params(h, args, p);
    cudaError_t e;
    switch (h->epw) {
    case 4: e = lt_launch_step<4>(p, (cudaStream_t)stream); break;
    case 8: e = lt_launch_step<8>(p, (cudaStream_t)stream); break;
    case 16: e = lt_launch_step<16>(p, (cudaStream_t)stream); break;
    default: e = lt_launch_step<32>(p, (cudaStream_t)stream); break;
    }
    if (e != cudaSuccess) return lt_fail(h, LT_ERR_CUDA, std::string("lt_step_kernel: ") + cudaGetErrorString(e));
    const int n = h->c.num_envs;
    lt_compact_kernel<<<(n + LT_COMPACT_THREADS - 1) / LT_COMPACT_THREADS, LT_COMPACT_THREADS, 0, (cudaStream_t)stream>>>(
        h->st.reset_buf, h->st.time_out_buf, h->st.time_outs_latched, h->st.reset_env_ids, h->st.reset_count, h->d_counters, n,
        h->c.send_timeouts, h->substeps_pending, h->L.lag_slots);
    LT_CUDA(h, cudaGetLastError());
    h->substeps_pending = 0;
    return LT_OK;
}

}  // extern "C"
