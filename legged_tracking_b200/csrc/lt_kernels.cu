// lt_kernels.cu -- sm_100a kernels of the per-environment step path + the C ABI of include/lt_b200.h.
//
// Kernels (LR = reference go1_gym/envs/base/legged_robot_trajectory_tracking.py):
//   lt_torque_kernel     _compute_torques + actuator net (LR:957-996, 1311-1320); one thread per (env, dof)
//   lt_step_kernel<EPW>  post_physics_step (LR:114-169) up to the observation row: each warp owns a tile of EPW
//                        consecutive envs; root/dof tile rows are staged in shared memory with 16-byte cp.async
//                        copies, one lane per env runs the scalar logic (lt_core.cuh) and shifts the history rows
//   lt_obs_kernel<KP,N>  compute_observations (LR:357-475): one warp per env streams the observation row incl. the
//                        heightfield gathers, noise, clamp, obs_history
//   lt_compact_kernel    reset_buf.nonzero() (LR:143), extras["time_outs"] latch (LR:289-291), step counters
//   lt_sign_kernel       all-env sums of the sign-indefinite reward terms (LR:332-335), only when configured
// There is no CPU fallback: lt_create fails without an sm_100 device.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>

#include "lt_core.cuh"
#include "lt_torque_mma.h"

#ifndef LT_STEP_THREADS
#define LT_STEP_THREADS 64
#endif
#ifndef LT_TORQUE_THREADS
#define LT_TORQUE_THREADS 128
#endif
#ifndef LT_TORQUE_MINB
#define LT_TORQUE_MINB 1           // min resident CTAs per SM asked of ptxas for the multi-sample torque kernel
#endif
#ifndef LT_TORQUE_SPT
#define LT_TORQUE_SPT 2            // (env, dof) samples per thread of the torque kernel: 1, 2 or 4 (tuning: profiles/variants.py)
#endif
#ifndef LT_TORQUE_MMA
#define LT_TORQUE_MMA 1            // actuator net on the tensor pipe (lt_torque_mma.cu); 0 = FP32-pipe kernels below
#endif
#ifndef LT_TC5_MIN_ENVS
#define LT_TC5_MIN_ENVS 40960      // shard size from which lt_compute_torques picks the tcgen05 kernel
#endif
#define LT_COMPACT_THREADS 1024
#define LT_OBS_THREADS 256

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
    uint8_t* scratch;                // [num_envs, rec_bytes] private: what lt_obs_kernel needs from lt_step_kernel
    int32_t rec_bytes;
};

struct lt_obs_params {
    int32_t num_envs, num_obs, obs_history_length, per_layer, nxf, h0, n_heights, head_w, plane, camera_zero;
    int32_t tile_cols, tile_cells, rec_bytes, height_noise;
    uint32_t seed_lo, seed_hi;
    float clip_observations, ceiling_height, obs_scale_height;
    const uint8_t* scratch;
    const uint32_t* ptlut;           // [per_layer] x row | y column << 8 of each measured point
    const float* tile_heights;
    const int32_t* env_tile;
    const float* noise_scale_vec;
    const float* noise_uniforms;
    const int32_t* counters;
    float *obs_buf, *obs_history;
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
    int height_noise;                // 1 if any height column has a nonzero noise scale
    float* d_term_sums;
    int32_t* d_counters;
    uint8_t* d_scratch;
    uint32_t* d_ptlut;
    float* d_act_frag;               // [LT_MMA_FRAG_REGS, 32] pre-split fp16 weight fragments of the actuator net
    uint8_t* d_tc5_blob;             // B operands of the tcgen05 torque kernel (canonical K-major core-matrix layout)
    int torque_impl;                 // LT_TORQUE_TC5: unset = by shard size, 0 = mma.sync kernel, 1 = tcgen05 kernel
    cudaStream_t side;               // lt_compact_kernel runs here, beside lt_obs_kernel (both depend only on lt_step_kernel)
    cudaEvent_t ev_fork, ev_join;
    bool fork_compact;
    int sm_count;
    int32_t rec_bytes;
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
__device__ __forceinline__ void lt_cp_async4(void* smem, const void* gmem) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void lt_cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
}
// stage `n` floats (16-byte aligned source, n % 4 == 0 when `vec`) global -> shared, one warp
__device__ __forceinline__ void lt_tile_load(float* s, const float* g, int n, int lane, bool vec) {
    if (vec) {
        LT_ROLLED for (int k = lane * 4; k < n; k += 128) lt_cp_async16(s + k, g + k);
    } else {
        LT_ROLLED for (int k = lane; k < n; k += 32) s[k] = g[k];
    }
}
__device__ __forceinline__ void lt_tile_store(float* g, const float* s, int n, int lane, bool vec) {
    if (vec) {
        LT_ROLLED for (int k = lane * 4; k < n; k += 128) *reinterpret_cast<float4*>(g + k) = *reinterpret_cast<const float4*>(s + k);
    } else {
        LT_ROLLED for (int k = lane; k < n; k += 32) g[k] = s[k];
    }
}
#ifndef LT_OPT_PREFETCH
#define LT_OPT_PREFETCH 1
#endif
__device__ __forceinline__ void lt_prefetch(const void* g) {
#if LT_OPT_PREFETCH
    asm volatile("prefetch.global.L1 [%0];" ::"l"(g));
#endif
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
    if (p.substep == c.decimation - 1) {         // LR:151-152, done here where the rows are walked coalesced anyway
        p.st.last_last_joint_pos_target[s] = p.st.last_joint_pos_target[s];
        p.st.last_joint_pos_target[s] = io.joint_pos_target;
    }
    p.st.torques[s] = io.torque;
}

// Multi-sample variant: S (2 or 4) consecutive (env, dof) samples per thread.  The [N,12] rows are contiguous, so the S
// samples of a thread are S adjacent floats of every buffer: vector loads/stores, and the network runs on sample PAIRS in
// the two halves of FFMA2 (lt_actuator_mlp_pairs).  12 % S == 0, so a thread never straddles the end of the arrays.
template <int S> struct lt_vec;
template <> struct lt_vec<2> { typedef float2 T; };
template <> struct lt_vec<4> { typedef float4 T; };
template <int S> __device__ __forceinline__ void lt_ldv(const float* g, float (&r)[S]) {
    const typename lt_vec<S>::T v = *reinterpret_cast<const typename lt_vec<S>::T*>(g);
    const float* f = reinterpret_cast<const float*>(&v);
#pragma unroll
    for (int i = 0; i < S; ++i) r[i] = f[i];
}
template <int S> __device__ __forceinline__ void lt_stv(float* g, const float (&r)[S]) {
    typename lt_vec<S>::T v;
    float* f = reinterpret_cast<float*>(&v);
#pragma unroll
    for (int i = 0; i < S; ++i) f[i] = r[i];
    *reinterpret_cast<typename lt_vec<S>::T*>(g) = v;
}

template <int S>
__global__ void __launch_bounds__(LT_TORQUE_THREADS, LT_TORQUE_MINB) lt_torque_kernel_v(const __grid_constant__ lt_torque_params p) {
    const lt_config& c = p.c;
    const size_t n = (size_t)c.num_envs * LT_ND;
    const size_t s0 = ((size_t)blockIdx.x * LT_TORQUE_THREADS + threadIdx.x) * S;
    if (s0 >= n) return;
    const int dof0 = (int)(s0 % LT_ND);
    const bool net = c.control_type == LT_CONTROL_ACTUATOR_NET;
    float act[S], pos[S], vel[S], lag[S], moff[S], mstr[S], el[S], ell[S], vl[S], vll[S], kp[S], kd[S];
    lt_ldv<S>(p.actions_raw + s0, act);
    {
        float ds[2][S];
        lt_ldv<S>(p.dof_state + 2 * s0, ds[0]);
        lt_ldv<S>(p.dof_state + 2 * s0 + S, ds[1]);
#pragma unroll
        for (int i = 0; i < S; ++i) { pos[i] = ds[(2 * i) / S][(2 * i) % S]; vel[i] = ds[(2 * i + 1) / S][(2 * i + 1) % S]; }
    }
    lt_ldv<S>(p.st.motor_offsets + s0, moff);
    lt_ldv<S>(p.st.motor_strengths + s0, mstr);
    if (net) {
        lt_ldv<S>(p.st.joint_pos_err_last + s0, el);
        lt_ldv<S>(p.st.joint_pos_err_last_last + s0, ell);
        lt_ldv<S>(p.st.joint_vel_last + s0, vl);
        lt_ldv<S>(p.st.joint_vel_last_last + s0, vll);
    } else {
        lt_ldv<S>(p.st.Kp_factors + s0, kp);
        lt_ldv<S>(p.st.Kd_factors + s0, kd);
    }
    lt_torque_io io[S];
    float clipped[S], scaled[S];
#pragma unroll
    for (int i = 0; i < S; ++i) {
        io[i].action_raw = act[i];
        scaled[i] = lt_scaled_action(c, act[i], dof0 + i, clipped[i]);
        lag[i] = 0.0f;
    }
    if (c.randomize_lag_timesteps) {
        const int Lg = c.lag_timesteps + 1;
        const int head = (p.counters[2] + p.substep) % Lg;
        const int rd = (head + 1) % Lg;
        lt_stv<S>(p.st.lag_buffer + (size_t)head * n + s0, scaled);
        if (rd == head) {
#pragma unroll
            for (int i = 0; i < S; ++i) lag[i] = scaled[i];
        } else {
            lt_ldv<S>(p.st.lag_buffer + (size_t)rd * n + s0, lag);
        }
    }
    float x[S][LT_ACT_IN], tq[S];
#pragma unroll
    for (int i = 0; i < S; ++i) {
        lt_torque_io& o = io[i];
        o.dof_pos = pos[i]; o.dof_vel = vel[i]; o.lag_out = lag[i];
        o.action_clipped = clipped[i]; o.action_scaled = scaled[i];
        o.motor_offset = moff[i]; o.motor_strength = mstr[i];
        o.kp_factor = net ? 1.0f : kp[i]; o.kd_factor = net ? 1.0f : kd[i];
        o.err_last = net ? el[i] : 0.0f; o.err_last_last = net ? ell[i] : 0.0f;
        o.vel_last = net ? vl[i] : 0.0f; o.vel_last_last = net ? vll[i] : 0.0f;
        tq[i] = 0.0f;
#pragma unroll
        for (int k = 0; k < LT_ACT_IN; ++k) x[i][k] = 0.0f;
        lt_torque_pre(c, dof0 + i, o, x[i], tq[i]);
    }
    if (net) {
        constexpr int P = S / 2;
        unsigned long long xp[P][LT_ACT_IN], op[P];
#pragma unroll
        for (int q = 0; q < P; ++q)
#pragma unroll
            for (int k = 0; k < LT_ACT_IN; ++k) xp[q][k] = lt_pack2(x[2 * q][k], x[2 * q + 1][k]);
        lt_actuator_mlp_pairs<P>(p.W, xp, op);
#pragma unroll
        for (int q = 0; q < P; ++q) lt_unpack2(op[q], tq[2 * q], tq[2 * q + 1]);
    }
    float tgt[S], out[S];
#pragma unroll
    for (int i = 0; i < S; ++i) {
        lt_torque_post(c, dof0 + i, io[i], net, x[i], tq[i]);
        el[i] = io[i].err_last; ell[i] = io[i].err_last_last; vl[i] = io[i].vel_last; vll[i] = io[i].vel_last_last;
        tgt[i] = io[i].joint_pos_target; out[i] = io[i].torque;
    }
    if (net) {
        lt_stv<S>(p.st.joint_pos_err_last + s0, el);
        lt_stv<S>(p.st.joint_pos_err_last_last + s0, ell);
        lt_stv<S>(p.st.joint_vel_last + s0, vl);
        lt_stv<S>(p.st.joint_vel_last_last + s0, vll);
    }
    if (p.substep == 0) lt_stv<S>(p.st.actions + s0, clipped);
    lt_stv<S>(p.st.joint_pos_target + s0, tgt);
    if (p.substep == c.decimation - 1) {         // LR:151-152, done here where the rows are walked coalesced anyway
        float prev[S];
        lt_ldv<S>(p.st.last_joint_pos_target + s0, prev);
        lt_stv<S>(p.st.last_last_joint_pos_target + s0, prev);
        lt_stv<S>(p.st.last_joint_pos_target + s0, tgt);
    }
    lt_stv<S>(p.st.torques + s0, out);
}

// ---------------------------------------------------------------------------------------------------------------
// post_physics_step up to the observation row.  One warp owns EPW consecutive envs, one lane per env.
//   phase 0  cp.async (16 B) stages the tile's root_states and dof_state rows in shared memory (both are written
//            back: arrow pose every step, robot root / dofs on reset); each lane reads its env's [12]-rows straight
//            into registers and prefetches the per-env words the scalar logic will touch; the contact-force tile is
//            staged too and reduced to one packed word per env
//   phase 1  the scalar step logic (lt_core.cuh); non-height observation columns go straight to obs_buf, the
//            heightfield cell indices + post-reset z + camera z into a packed per-env record for lt_obs_kernel
//   phase 2  flat coalesced write-back of the root / dof / record tiles
// Shared memory per warp: root[EPW*A*13] dof[EPW*24] cinfo[EPW] es[EPW*(terms+3)] cs[EPW*(terms+1)] floats + rec[EPW*rec_bytes] bytes.
struct lt_smem_plan {
    int root, dof, contact, rec, es, cs, total;       // float offsets
};
__host__ __device__ inline int lt_align4(int x) { return (x + 3) & ~3; }
__host__ __device__ inline lt_smem_plan lt_plan_smem(const lt_config& c, int rec_bytes, int epw) {
    lt_smem_plan P;
    int o = 0;
    P.root = o; o += lt_align4(epw * c.num_actors * 13);
    P.dof = o; o += lt_align4(epw * LT_ND * 2);
    P.contact = o; o += lt_align4(epw);              // one packed contact word per env (lt_contact_summary)
    P.rec = o; o += lt_align4(epw * rec_bytes / 4);
    P.es = o; o += epw * (c.n_terms + 3);            // episode_sums rows [t][lane]
    P.cs = o; o += epw * (c.n_terms + 1);            // command_sums rows [t][lane]
    P.total = lt_align4(o);
    return P;
}

__device__ __forceinline__ void lt_load_row12(const float* g, float* r) {
    const float4* g4 = reinterpret_cast<const float4*>(g);
#pragma unroll
    for (int k = 0; k < 3; ++k) { const float4 v = g4[k]; r[4 * k] = v.x; r[4 * k + 1] = v.y; r[4 * k + 2] = v.z; r[4 * k + 3] = v.w; }
}
__device__ __forceinline__ void lt_store_row12(float* g, const float* r) {
    float4* g4 = reinterpret_cast<float4*>(g);
#pragma unroll
    for (int k = 0; k < 3; ++k) g4[k] = make_float4(r[4 * k], r[4 * k + 1], r[4 * k + 2], r[4 * k + 3]);
}

template <int EPW>
__global__ void __launch_bounds__(LT_STEP_THREADS) lt_step_kernel(const __grid_constant__ lt_step_params p) {
    extern __shared__ __align__(16) float smem[];
    lt_pdl_launch_dependents();
    lt_pdl_wait();                                 // everything below reads what the torque kernels / PhysX wrote
    const lt_config& c = p.c;
    const lt_layout& L = p.L;
    const lt_state_ptrs& st = p.st;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    constexpr int WARPS = LT_STEP_THREADS / 32;
    const int N = c.num_envs;
    const int A13 = c.num_actors * 13, B3 = c.num_bodies * 3;
    const int rec_bytes = p.rec_bytes;

    const lt_smem_plan P = lt_plan_smem(c, rec_bytes, EPW);
    float* ws = smem + (size_t)warp * P.total;
    float* s_root = ws + P.root;
    float* s_dof = ws + P.dof;
    uint32_t* s_cinfo = reinterpret_cast<uint32_t*>(ws + P.contact);
    uint8_t* s_rec = reinterpret_cast<uint8_t*>(ws + P.rec);
    float* s_es = ws + P.es;
    float* s_cs = ws + P.cs;

    const int tile = blockIdx.x * WARPS + warp;
    const int env0 = tile * EPW;
    if (env0 >= N) return;
    const int nv = min(EPW, N - env0);            // valid envs in this tile
    const bool full = (nv == EPW) && (EPW % 4 == 0);
    const bool mine = lane < nv;
    const int i = env0 + lane;

    // ---- phase 0
    {   // the contact-force tile (read by the sweep below in three dependent batches): its lines requested now, under the tile loads
        const char* cgb = reinterpret_cast<const char*>(p.sim.contact_forces + (size_t)env0 * B3);
        const int bytes = nv * B3 * 4;
        for (int o = lane * 128; o < bytes; o += 32 * 128) lt_prefetch(cgb + o);
    }
    lt_tile_load(s_root, p.sim.root_states + (size_t)env0 * A13, nv * A13, lane, full);
    lt_tile_load(s_dof, p.sim.dof_state + (size_t)env0 * LT_ND * 2, nv * LT_ND * 2, lane, full);
    // net contact forces: never staged.  Lane <-> (env, body) pair, so that the tile is read as one coalesced stream, and
    // each pair's flags are merged into its env's packed word (same bits as lt_contact_summary; integer, order-free)
    if (lane < EPW) s_cinfo[lane] = 0u;
    __syncwarp();
    {
        const float* cg = p.sim.contact_forces + (size_t)env0 * B3;
        const int nb = c.num_bodies, pairs = nv * nb;
        const int step_e = 32 / nb, step_b = 32 % nb;
        int e = lane / nb, b = lane % nb;
        const uint32_t listed = c.penalised_mask | c.termination_mask;
        uint32_t foot_bit[2] = {0u, 0u};          // body -> foot slot + 1, 4 bits per body (LT_MAX_BODIES = 32 -> 2 x 64 bits as 4 x 32)
        unsigned long long foot_lut0 = 0ull, foot_lut1 = 0ull;
        for (int k = 0; k < c.n_feet; ++k) {
            const int fb = c.feet_indices[k];
            if (fb < 16) foot_lut0 |= (unsigned long long)(k + 1) << (4 * fb);
            else foot_lut1 |= (unsigned long long)(k + 1) << (4 * (fb - 16));
        }
        (void)foot_bit;
#pragma unroll 6
        for (int q = lane; q < pairs; q += 32) {
            const float fx = cg[3 * q], fy = cg[3 * q + 1], fz = cg[3 * q + 2];
            uint32_t cnt = 0u, flags = 0u;
            if ((listed >> b) & 1u) {
                // |F| > 0.1 and |F| > 1 are decided on |F|^2 away from the thresholds (sqrt is monotonic; the 2 % bands are far wider
                // than its rounding); inside a band the exact lt_norm3 comparison runs
                const float ss = LT_FFMA(fz, fz, LT_FFMA(fy, fy, LT_FMUL(fx, fx)));
                bool g01 = ss > 0.0102f, g1 = ss > 1.02f;
                if ((ss > 0.0098f && ss <= 0.0102f) || (ss > 0.98f && ss <= 1.02f)) {
                    const float f = LT_FSQRT(ss);
                    g01 = f > 0.1f;
                    g1 = f > 1.0f;
                }
                if (((c.penalised_mask >> b) & 1u) && g01) cnt = 1u;
                if (((c.termination_mask >> b) & 1u) && g1) flags |= 0x100u;
            }
            const uint32_t slot = (uint32_t)((b < 16 ? foot_lut0 >> (4 * b) : foot_lut1 >> (4 * (b - 16))) & 0xfull);
            if (slot && fz > 1.0f) flags |= 0x8000u << slot;          // 0x10000 << (slot - 1)
            if (cnt) atomicAdd(s_cinfo + e, cnt);
            if (flags) atomicOr(s_cinfo + e, flags);
            e += step_e; b += step_b;
            if (b >= nb) { b -= nb; e += 1; }
        }
    }
    float r_act[LT_ND], r_last_act[LT_ND], r_last_dof_vel[LT_ND], r_torques[LT_ND];
    uint32_t cinfo = 0u;
    if (mine) {
        lt_load_row12(st.actions + (size_t)i * LT_ND, r_act);
        lt_load_row12(st.last_actions + (size_t)i * LT_ND, r_last_act);
        lt_load_row12(st.last_dof_vel + (size_t)i * LT_ND, r_last_dof_vel);
        lt_load_row12(st.torques + (size_t)i * LT_ND, r_torques);
        // the scalar phase walks these per-env words one dependent round trip at a time: start them all now
        lt_prefetch(st.episode_length_buf + i);
        lt_prefetch(st.curr_pose_index + i);
        lt_prefetch(st.collision_count + i);
        lt_prefetch(st.base_rotation + (size_t)i * 3);
        lt_prefetch(st.trajectories + (size_t)i * c.traj_length * 6);
        lt_prefetch(st.friction_coeffs + i);
        lt_prefetch(st.restitutions + i);
        lt_prefetch(st.body_height_buf + i);
        if (c.observe_heights && !c.plane) lt_prefetch(p.ter.env_terrain_origin + (size_t)i * 3);
        // episode / command sums: asynchronous 4-byte copies into [t][lane] shared rows (no register, no stall until the wait)
        LT_ROLLED for (int t = 0; t < L.n_sums; ++t) lt_cp_async4(s_es + t * EPW + lane, st.episode_sums + (size_t)t * N + i);
        if (st.command_sums)
            LT_ROLLED for (int t = 0; t <= c.n_terms; ++t) lt_cp_async4(s_cs + t * EPW + lane, st.command_sums + (size_t)t * N + i);
    }
    lt_cp_async_wait_all();
    __syncwarp();
    if (mine) cinfo = s_cinfo[lane];

    // ---- phase 1: the scalar logic of the step
    int my_reset = 0;
    if (mine) {
        lt_env_rows R;
        R.root = s_root + lane * A13;
        R.arrow = (c.num_actors > 1) ? s_root + lane * A13 + (c.num_actors - 1) * 13 : nullptr;
        R.dof = s_dof + lane * LT_ND * 2;
        R.cinfo = cinfo;
        R.actions = r_act;
        R.last_actions = r_last_act;
        R.last_dof_vel = r_last_dof_vel;
        R.torques = r_torques;
        R.obs_row = st.obs_buf + (size_t)i * c.num_obs;
        R.es = s_es + lane;
        R.cs = st.command_sums ? s_cs + lane : nullptr;
        R.sum_stride = EPW;
        uint8_t* rec = s_rec + lane * rec_bytes;
        R.ix = rec;
        R.iy = rec + L.nxf;
        lt_step_scalars sc = p.sc;
        sc.step = (uint32_t)p.counters[0];
        // lt_obs_kernel runs beside lt_compact_kernel, which advances counters[0]: leave it this step's value in counters[3]
        if (tile == 0 && lane == 0) const_cast<int32_t*>(p.counters)[3] = (int32_t)sc.step;
        const lt_env_result r = lt_env_step(c, L, st, p.ter, p.args, sc, i, R);
        my_reset = r.reset;
        *reinterpret_cast<float2*>(rec + rec_bytes - 8) = make_float2(r.z_post, r.cam_z);
        // history shifts (LR:148-153) of this env's rows, straight from registers
        float row[LT_ND];
#pragma unroll
        for (int d = 0; d < LT_ND; ++d) row[d] = my_reset ? 0.0f : r_last_act[d];      // last_actions[env_ids] = 0 precedes the shift (LR:246)
        lt_store_row12(st.last_last_actions + (size_t)i * LT_ND, row);
        lt_store_row12(st.last_actions + (size_t)i * LT_ND, r_act);
        // (the joint_pos_target history shift of LR:151-152 is done by the torque kernel's last sub-step)
#pragma unroll
        for (int d = 0; d < LT_ND; ++d) row[d] = R.dof[2 * d + 1];                      // post-reset dof_vel
        lt_store_row12(st.last_dof_vel + (size_t)i * LT_ND, row);
        float2* lrv = reinterpret_cast<float2*>(st.last_root_vel + (size_t)i * 6);
#pragma unroll
        for (int k = 0; k < 3; ++k) lrv[k] = make_float2(R.root[7 + 2 * k], R.root[8 + 2 * k]);
        // sums back to their term-major rows (coalesced across the warp)
        _Pragma("unroll 4") for (int t = 0; t < L.n_sums; ++t) st.episode_sums[(size_t)t * N + i] = s_es[t * EPW + lane];
        if (st.command_sums)
            _Pragma("unroll 4") for (int t = 0; t <= c.n_terms; ++t) st.command_sums[(size_t)t * N + i] = s_cs[t * EPW + lane];
    }
    __syncwarp();
    const unsigned reset_mask = __ballot_sync(0xffffffffu, my_reset != 0);

    // ---- phase 2: write back the sim rows the step owns (arrow pose every step; robot root / dofs on reset) and the
    // records of the observation pass
    if (c.num_actors > 1 || reset_mask) lt_tile_store(p.sim.root_states + (size_t)env0 * A13, s_root, nv * A13, lane, full);
    if (reset_mask) lt_tile_store(p.sim.dof_state + (size_t)env0 * LT_ND * 2, s_dof, nv * LT_ND * 2, lane, full);
    lt_tile_store(reinterpret_cast<float*>(p.scratch + (size_t)env0 * rec_bytes), reinterpret_cast<const float*>(s_rec),
                  nv * rec_bytes / 4, lane, true);
}

// ---------------------------------------------------------------------------------------------------------------
// Observation pass: one warp per env.  Reads the env's packed record + raw non-height columns, gathers the 2 x nxf x ny
// heightfield samples from the (L2-resident) tile table, applies camera / clip / scale (LR:395-409), observation
// noise (LR:473-474; NOISE 0 = off, 1 = injected uniforms, 2 = in-kernel Philox4x32-10, ten 12-bit values per call), the
// +-clip_observations clamp (LR:108-109) and writes obs_buf and the newest obs_history block (history_wrapper.py:22).
// Lane <-> measured point pt = lane + 32 k (k < KP); one cell offset serves both layers.  CAM = env.camera_zero.
template <int NOISE>
__device__ __forceinline__ float lt_obs_finish_dev(float v, float u, float scale, float clip) {
    if (NOISE) v = LT_FADD(v, LT_FMUL(LT_FSUB(LT_FMUL(2.0f, u), 1.0f), scale));
    return fminf(fmaxf(v, -clip), clip);
}
template <int SLOT>
__device__ __forceinline__ float lt_u12_static(const uint32_t (&rnd)[4]) {      // == lt_u12(rnd, SLOT)
    constexpr int bit = 12 * SLOT, w = bit >> 5, sh = bit & 31;
    uint32_t x = rnd[w] >> sh;
    if (sh > 20) x |= rnd[w < 3 ? w + 1 : 3] << (32 - sh);
    return LT_FMUL((float)(x & 0xfffu) + 0.5f, 1.0f / 4096.0f);
}
// HN = 1 only when a height column has a nonzero noise scale; the reference's _get_noise_scale_vec hard-codes zeros there
// (LR:1125-1131), and v + (2u-1)*0 == v, so the usual instantiation skips noise for the 220 height columns altogether.
template <int KP, int NOISE, int CAM, int HN>
__global__ void __launch_bounds__(LT_OBS_THREADS) lt_obs_kernel(const __grid_constant__ lt_obs_params p) {
    lt_pdl_launch_dependents();
    lt_pdl_wait();                                 // the record and the head columns come from lt_step_kernel
    const int lane = threadIdx.x & 31;
    const int ie = blockIdx.x * (LT_OBS_THREADS / 32) + (threadIdx.x >> 5);
    if (ie >= p.num_envs) return;
    const int no = p.num_obs, H = p.obs_history_length, per_layer = p.per_layer;
    const float clip = p.clip_observations, hscale = p.obs_scale_height, ceil_h = p.ceiling_height;
    const uint32_t step = NOISE == 2 ? (uint32_t)p.counters[3] : 0u;        // = counters[0] as lt_step_kernel saw it
    const uint8_t* rec = p.scratch + (size_t)ie * p.rec_bytes;
    // the env's packed record (cell-index bytes, then (z_post, cam_z)) in ONE coalesced load: lane l holds word l; a
    // point's row / column byte is fetched from the owning lane by shuffle instead of a dependent byte load
    const int rec_words = p.rec_bytes >> 2;
    const uint32_t rec_w = lane < rec_words ? __ldg(reinterpret_cast<const uint32_t*>(rec) + lane) : 0u;
    float* obs_row = p.obs_buf + (size_t)ie * no;
    // raw non-height columns (written by lt_step_kernel) and their noise scales: issued now, used after the height loop
    const int hcol0 = lane < p.h0 ? lane : lane + p.n_heights;
    const int hcol1 = (lane + 32) < p.h0 ? lane + 32 : lane + 32 + p.n_heights;
    const float raw0 = lane < p.head_w ? obs_row[hcol0] : 0.0f;
    const float raw1 = lane + 32 < p.head_w ? obs_row[hcol1] : 0.0f;
    const float ns0 = (NOISE && lane < p.head_w) ? __ldg(p.noise_scale_vec + hcol0) : 0.0f;
    const float ns1 = (NOISE && lane + 32 < p.head_w) ? __ldg(p.noise_scale_vec + hcol1) : 0.0f;
    const float2 zc = make_float2(__uint_as_float(__shfl_sync(0xffffffffu, rec_w, rec_words - 2)),
                                  __uint_as_float(__shfl_sync(0xffffffffu, rec_w, rec_words - 1)));        // (z_post, cam_z)
    float hc[KP], hf[KP];
    {
        const float* tile_base = p.plane ? nullptr : p.tile_heights + (size_t)__ldg(p.env_tile + ie) * 2 * p.tile_cells;
#pragma unroll
        for (int k = 0; k < KP; ++k) {
            const int pt = lane + 32 * k;
            hc[k] = 1.0f;                         // plane: ceiling 1, floor 0 (LR:1928-1932)
            hf[k] = 0.0f;
            const bool on = pt < per_layer;
            const uint32_t w = on ? __ldg(p.ptlut + pt) : 0u;
            const int bx = (int)(w & 0xffu), by = p.nxf + (int)(w >> 8);           // byte positions in the record
            const uint32_t wx = __shfl_sync(0xffffffffu, rec_w, bx >> 2), wy = __shfl_sync(0xffffffffu, rec_w, by >> 2);
            if (tile_base && on) {
                // one PRMT per byte: selector 0x444s puts byte s of the word in the low byte and zeros (byte 4 of {word, 0}) above it
                const int off = (int)__byte_perm(wx, 0u, 0x4440u | (uint32_t)(bx & 3)) * p.tile_cols +
                                (int)__byte_perm(wy, 0u, 0x4440u | (uint32_t)(by & 3));
                hc[k] = __ldg(tile_base + off);
                hf[k] = __ldg(tile_base + p.tile_cells + off);
            }
        }
    }
    uint32_t rnd[4] = {0u, 0u, 0u, 0u};
    if (HN && NOISE == 2) lt_philox((uint32_t)ie, (uint32_t)lane | LT_STREAM_NOISE, step, 0u, p.seed_lo, p.seed_hi, rnd);
    const bool hist_alias = p.obs_history == p.obs_buf;          // H == 1: the history block IS the observation row (engine.py)
    float* hist_row = p.obs_history + (size_t)ie * no * H;
    if (H > 1) {
        const int keep = (H - 1) * no;
        // in-place left shift by one block: every lane's read of a round completes before any lane's write of that round, and
        // a round completes before the next one starts (no reliance on the warp staying converged)
        for (int k0 = 0; k0 < keep; k0 += 32) {
            const int k = k0 + lane;
            const float v = k < keep ? hist_row[k + no] : 0.0f;
            __syncwarp();
            if (k < keep) hist_row[k] = v;
            __syncwarp();
        }
    }
    // per-lane base pointers: column h0 + lane (ceiling layer) and h0 + per_layer + lane (floor layer)
    float* hist_new = hist_row + (size_t)(H - 1) * no;
    const int c0 = p.h0 + lane, c1 = c0 + per_layer;
    // four per-lane row pointers, so that every store below is base + compile-time offset (no 64-bit address math per store)
    float* const po[2] = {obs_row + c0, obs_row + c1};
    float* const ph[2] = {hist_new + c0, hist_new + c1};
    const float* nz = NOISE == 1 ? p.noise_uniforms + (size_t)ie * no : nullptr;
#pragma unroll
    for (int k = 0; k < KP; ++k) {
        if (lane + 32 * k < per_layer) {
#pragma unroll
            for (int layer = 0; layer < 2; ++layer) {
                constexpr int dummy = 0; (void)dummy;
                const int col = (layer ? c1 : c0) + 32 * k;
                float h = layer ? hf[k] : hc[k];
                if (CAM) h = fminf(fmaxf(LT_FSUB(LT_FSUB(h, zc.x), zc.y), -0.3f), 0.3f);
                else h = LT_FSUB(LT_FDIV(fminf(fmaxf(h, 0.0f), ceil_h), ceil_h), 0.5f);
                h = LT_FMUL(h, hscale);
                float u = 0.5f;
                if (HN && NOISE == 1) u = nz[col];
                if (HN && NOISE == 2) {
                    const int v = 2 * k + layer;                 // compile-time after unrolling
                    if (v == LT_NOISE_PER_CALL)                  // KP = 8 only: values 10.. come from the second call
                        lt_philox((uint32_t)ie, (uint32_t)(lane + 32) | LT_STREAM_NOISE, step, 0u, p.seed_lo, p.seed_hi, rnd);
                    switch (v % LT_NOISE_PER_CALL) {
                    case 0: u = lt_u12_static<0>(rnd); break; case 1: u = lt_u12_static<1>(rnd); break;
                    case 2: u = lt_u12_static<2>(rnd); break; case 3: u = lt_u12_static<3>(rnd); break;
                    case 4: u = lt_u12_static<4>(rnd); break; case 5: u = lt_u12_static<5>(rnd); break;
                    case 6: u = lt_u12_static<6>(rnd); break; case 7: u = lt_u12_static<7>(rnd); break;
                    case 8: u = lt_u12_static<8>(rnd); break; default: u = lt_u12_static<9>(rnd); break;
                    }
                }
                const float o = lt_obs_finish_dev<(HN ? NOISE : 0)>(h, u, (HN && NOISE) ? __ldg(p.noise_scale_vec + col) : 0.0f, clip);
                po[layer][32 * k] = o;
                if (!hist_alias) ph[layer][32 * k] = o;
            }
        }
    }
    const int head_iters = (p.head_w + 31) >> 5;
    auto head_call = [&](int m) {                  // Philox call boundary: value index v = 2 KP + m uses call v / 10, slot v % 10
        const int v = 2 * KP + m;
        if (NOISE == 2 && (v % LT_NOISE_PER_CALL == 0 || (!HN && m == 0)))      // !HN: the height loop made no call yet
            lt_philox((uint32_t)ie, (uint32_t)(lane + 32 * (v / LT_NOISE_PER_CALL)) | LT_STREAM_NOISE, step, 0u, p.seed_lo, p.seed_hi, rnd);
    };
    auto head_col = [&](int m, float u) {
        const int hi = lane + 32 * m;
        if (hi < p.head_w) {
            const int col = hi < p.h0 ? hi : hi + p.n_heights;
            if (NOISE == 1) u = nz[col];
            const float raw = m == 0 ? raw0 : (m == 1 ? raw1 : obs_row[col]);
            const float ns = !NOISE ? 0.0f : (m == 0 ? ns0 : (m == 1 ? ns1 : __ldg(p.noise_scale_vec + col)));
            const float o = lt_obs_finish_dev<NOISE>(raw, u, ns, clip);
            obs_row[col] = o;
            if (!hist_alias) hist_new[col] = o;
        }
    };
    // the first two rounds (64 columns: every shipped configuration) with compile-time noise slots; the rest rolled
#pragma unroll
    for (int m = 0; m < 2; ++m) {
        if (m < head_iters) {
            head_call(m);
            float u = 0.5f;
            if (NOISE == 2) {
                switch ((2 * KP + m) % LT_NOISE_PER_CALL) {          // compile-time after unrolling
                case 0: u = lt_u12_static<0>(rnd); break; case 1: u = lt_u12_static<1>(rnd); break;
                case 2: u = lt_u12_static<2>(rnd); break; case 3: u = lt_u12_static<3>(rnd); break;
                case 4: u = lt_u12_static<4>(rnd); break; case 5: u = lt_u12_static<5>(rnd); break;
                case 6: u = lt_u12_static<6>(rnd); break; case 7: u = lt_u12_static<7>(rnd); break;
                case 8: u = lt_u12_static<8>(rnd); break; default: u = lt_u12_static<9>(rnd); break;
                }
            }
            head_col(m, u);
        }
    }
#pragma unroll 1
    for (int m = 2; m < head_iters; ++m) {
        head_call(m);
        head_col(m, NOISE == 2 ? lt_u12(rnd, (2 * KP + m) % LT_NOISE_PER_CALL) : 0.5f);
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
    lt_pdl_launch_dependents();
    lt_pdl_wait();
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
            rc.cinfo = lt_contact_summary(c, p.sim.contact_forces + (size_t)i * c.num_bodies * 3);
            rc.root_z = root[2];
            for (int k = 0; k < 3; ++k) { rc.blv[k] = P.blv[k]; rc.bav[k] = P.bav[k]; rc.pg[k] = P.pg[k];
                                          rc.rel_lin[k] = P.rel_lin[k]; rc.rel_rot[k] = P.rel_rot[k]; rc.loc_lin[k] = P.rel_lin[k]; }
            rc.ep_f = (float)(p.st.episode_length_buf[i] + 1);
            rc.reached = 0;
            rc.n_collisions = 0;
            rc.local_goal = 0;
            v = lt_reward_term(c, p.st, i, c.term_ids[t], rc, lt_ext_row(c, p.args.extra_reward, t), true) * c.term_scales[t];
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

// ---------------------------------------------------------------------------------------------------------------
// LeggedRobot.reset_idx(env_ids) called from OUTSIDE a step (env.reset(), LR:218-296): the same per-env reset the step
// kernel performs in line, for an explicit id list.  Rare path: thread per listed env, straight to global memory.
__global__ void __launch_bounds__(128) lt_reset_idx_kernel(const __grid_constant__ lt_step_params p, const int64_t* __restrict__ ids,
                                                           int count, const int32_t* __restrict__ count_dev) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    const int n_ids = count_dev ? *count_dev : count;
    if (k >= n_ids) return;
    const lt_config& c = p.c;
    const lt_state_ptrs& st = p.st;
    const int i = (int)ids[k];
    if (i < 0 || i >= c.num_envs) return;
    const size_t N = (size_t)c.num_envs;
    constexpr int nd = LT_ND;
    lt_uniform_src U;
    U.row = p.args.uniforms ? p.args.uniforms + (size_t)i * p.L.n_slots : nullptr;
    U.seed_lo = c.seed_lo; U.seed_hi = c.seed_hi; U.step = (uint32_t)p.counters[0] | 0x80000000u; U.env = (uint32_t)i;
    float* root = p.sim.root_states + (size_t)i * c.num_actors * 13;
    float* dof = p.sim.dof_state + (size_t)i * nd * 2;
    // finished-episode record (LR:256-273) before anything is zeroed
    st.finished_episode_length[i] = (float)st.episode_length_buf[i];
    const float* rl = st.relative_linear + (size_t)i * 3;
    st.finished_goal_distance[i] = lt_norm3(rl[0], rl[1], rl[2]);
    for (int s = 0; s < p.L.n_sums; ++s) {
        st.finished_sums[(size_t)s * N + i] = st.episode_sums[(size_t)s * N + i];
        st.episode_sums[(size_t)s * N + i] = 0.0f;
    }
    const bool reached = st.reached_buf[i] != 0;
    st.reached_env_buf[i] = reached ? 1 : 0;
    st.collision_env_buf[i] = reached ? st.collision_count[i] : 0;
    st.collision_count[i] = 0;
    lt_reset_sim_state(c, st, p.ter, U, i, root, dof);
    st.curr_pose_index[i] = 0;
    for (int d = 0; d < nd; ++d) {
        st.last_actions[(size_t)i * nd + d] = 0.0f;
        st.last_last_actions[(size_t)i * nd + d] = 0.0f;
        st.last_dof_vel[(size_t)i * nd + d] = 0.0f;
    }
    for (int f = 0; f < c.n_feet; ++f) st.feet_air_time[(size_t)i * c.n_feet + f] = 0.0f;
    st.episode_length_buf[i] = 0;
    st.reset_buf[i] = 1;
    st.plan_buf[i] = 1;
    for (int q = 0; q < 3; ++q) {
        st.local_relative_linear[(size_t)i * 3 + q] = 0.0f;
        st.local_relative_rotation[(size_t)i * 3 + q] = 0.0f;
    }
    for (int s = 0; s < p.L.lag_slots; ++s)
        for (int d = 0; d < nd; ++d) st.lag_buffer[((size_t)s * N + i) * nd + d] = 0.0f;
}

template <int EPW>
static cudaError_t lt_launch_step(const lt_step_params& p, cudaStream_t stream) {
    constexpr int WARPS = LT_STEP_THREADS / 32;
    const lt_smem_plan P = lt_plan_smem(p.c, p.rec_bytes, EPW);
    const size_t smem = sizeof(float) * (size_t)WARPS * P.total;
    static size_t configured = 0;
    if (smem > 48 * 1024 && smem > configured) {
        cudaError_t e = cudaFuncSetAttribute(lt_step_kernel<EPW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        configured = smem;
    }
    const int tiles = (p.c.num_envs + EPW - 1) / EPW;
    return lt_launch_chain(lt_step_kernel<EPW>, (unsigned)((tiles + WARPS - 1) / WARPS), LT_STEP_THREADS, smem, stream, p);
}
template <int KP, int CAM>
static cudaError_t lt_launch_obs_cam(const lt_obs_params& p, int noise_mode, cudaStream_t stream) {
    const unsigned grid = (unsigned)((p.num_envs + LT_OBS_THREADS / 32 - 1) / (LT_OBS_THREADS / 32));
    const bool hn = p.height_noise != 0;
    if (noise_mode == 0) return lt_launch_chain(lt_obs_kernel<KP, 0, CAM, 0>, grid, LT_OBS_THREADS, 0, stream, p);
    else if (noise_mode == 1 && hn) return lt_launch_chain(lt_obs_kernel<KP, 1, CAM, 1>, grid, LT_OBS_THREADS, 0, stream, p);
    else if (noise_mode == 1) return lt_launch_chain(lt_obs_kernel<KP, 1, CAM, 0>, grid, LT_OBS_THREADS, 0, stream, p);
    else if (hn) return lt_launch_chain(lt_obs_kernel<KP, 2, CAM, 1>, grid, LT_OBS_THREADS, 0, stream, p);
    else return lt_launch_chain(lt_obs_kernel<KP, 2, CAM, 0>, grid, LT_OBS_THREADS, 0, stream, p);
}
template <int KP>
static cudaError_t lt_launch_obs(const lt_obs_params& p, int noise_mode, cudaStream_t stream) {
    return p.camera_zero ? lt_launch_obs_cam<KP, 1>(p, noise_mode, stream) : lt_launch_obs_cam<KP, 0>(p, noise_mode, stream);
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
    LT_REQ(c.n_points_x * c.n_points_y <= 256);
    LT_REQ(c.plane || !c.observe_heights || (c.tile_rows <= 256 && c.tile_cols <= 256));
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
        epw = per_sm >= 256 ? 32 : 8;      // measured (B200): 8 wins at 8,192 and 16,384 envs, 32 at 65,536
    }
    h->epw = epw;
    h->height_noise = 0;
    for (int q = 0; q < h->L.n_heights; ++q)
        if (tables->noise_scale_vec[h->L.h0 + q] != 0.0f) h->height_noise = 1;
    cudaError_t e = cudaMalloc(&h->d_noise_scale, sizeof(float) * (size_t)cfg->num_obs);
    if (e == cudaSuccess) e = cudaMemcpy(h->d_noise_scale, tables->noise_scale_vec, sizeof(float) * (size_t)cfg->num_obs, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMalloc(&h->d_term_sums, sizeof(float) * LT_MAX_TERMS);
    if (e == cudaSuccess) e = cudaMemset(h->d_term_sums, 0, sizeof(float) * LT_MAX_TERMS);
    if (e == cudaSuccess) e = cudaMalloc(&h->d_counters, sizeof(int32_t) * 4);
    if (e == cudaSuccess) e = cudaMemset(h->d_counters, 0, sizeof(int32_t) * 4);
    // private scratch between lt_step_kernel and lt_obs_kernel: per env nxf + ny cell indices (bytes) + 2 floats
    h->rec_bytes = ((h->L.nxf + cfg->n_points_y + 3) / 4 * 4 + 8 + 15) / 16 * 16;
    if (e == cudaSuccess) e = cudaMalloc(&h->d_scratch, (size_t)h->rec_bytes * (size_t)cfg->num_envs);
    if (e == cudaSuccess) e = cudaMemset(h->d_scratch, 0, (size_t)h->rec_bytes * (size_t)cfg->num_envs);
    {
        const int per_layer = h->L.nxf * cfg->n_points_y;
        std::string lut((size_t)(per_layer > 0 ? per_layer : 1) * 4, '\0');
        uint32_t* w = reinterpret_cast<uint32_t*>(&lut[0]);
        for (int pt = 0; pt < per_layer; ++pt) w[pt] = (uint32_t)(pt / cfg->n_points_y) | ((uint32_t)(pt % cfg->n_points_y) << 8);
        if (e == cudaSuccess) e = cudaMalloc(&h->d_ptlut, lut.size());
        if (e == cudaSuccess) e = cudaMemcpy(h->d_ptlut, lut.data(), lut.size(), cudaMemcpyHostToDevice);
    }
    {
        std::string fr((size_t)LT_MMA_FRAG_REGS * 32 * sizeof(float), '\0');
        lt_torque_mma_build_frags(h->W, reinterpret_cast<float*>(&fr[0]));
        if (e == cudaSuccess) e = cudaMalloc(&h->d_act_frag, fr.size());
        if (e == cudaSuccess) e = cudaMemcpy(h->d_act_frag, fr.data(), fr.size(), cudaMemcpyHostToDevice);
    }
    {
        std::string blob((size_t)LT_TC5_BLOB_BYTES, '\0');
        lt_torque_tc5_build_blob(h->W, reinterpret_cast<uint8_t*>(&blob[0]));
        if (e == cudaSuccess) e = cudaMalloc(&h->d_tc5_blob, blob.size());
        if (e == cudaSuccess) e = cudaMemcpy(h->d_tc5_blob, blob.data(), blob.size(), cudaMemcpyHostToDevice);
        const char* t5 = getenv("LT_TORQUE_TC5");
        h->torque_impl = t5 ? atoi(t5) : -1;
    }
    h->sm_count = prop.multiProcessorCount;
    {
        const char* f = getenv("LT_FORK_COMPACT");
        h->fork_compact = !(f && f[0] == '0');
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming);
    }
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
    cudaFree(h->d_scratch);
    cudaFree(h->d_ptlut);
    cudaFree(h->d_act_frag);
    cudaFree(h->d_tc5_blob);
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    if (h->ev_join) cudaEventDestroy(h->ev_join);
    if (h->side) cudaStreamDestroy(h->side);
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
int lt_reset_idx(lt_handle* h, const int64_t* env_ids, int32_t count, const int32_t* count_dev, const float* uniforms, void* stream) {
    if (!h) return LT_ERR_INVALID;
    if (!h->have_sim || !h->have_terrain || !h->have_state) return lt_fail(h, LT_ERR_UNBOUND, "lt_reset_idx: bind sim, terrain and state first");
    if (!env_ids || count < 0) return lt_fail(h, LT_ERR_INVALID, "lt_reset_idx: bad id list");
    if (count == 0) return LT_OK;
    lt_step_params p;
    memset(&p, 0, sizeof p);
    p.c = h->c; p.L = h->L; p.st = h->st; p.sim = h->sim; p.ter = h->ter;
    p.args.uniforms = uniforms;
    p.counters = h->d_counters;
    lt_reset_idx_kernel<<<(count + 127) / 128, 128, 0, (cudaStream_t)stream>>>(p, env_ids, count, count_dev);
    LT_CUDA(h, cudaGetLastError());
    return LT_OK;
}

float* lt_term_sum_buffer(lt_handle* h) { return h ? h->d_term_sums : nullptr; }

int lt_set_term_scales(lt_handle* h, const float* scales, int32_t n) {
    if (!h) return LT_ERR_INVALID;
    if (!scales || n != h->c.n_terms) return lt_fail(h, LT_ERR_INVALID, "lt_set_term_scales: expected " + std::to_string(h->c.n_terms) + " scales");
    for (int t = 0; t < n; ++t) h->c.term_scales[t] = scales[t];      // the config travels as a kernel parameter: effective from the next launch
    return LT_OK;
}

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
#if LT_TORQUE_MMA
    // tcgen05 kernel for large shards, mma.sync kernel for small ones: the former pays ~1.3 us per launch for the TMEM allocation,
    // the weight copy to shared memory and the barrier set-up and only wins once a CTA has many tiles (profiles/r01e_tc5.md)
    const bool tc5 = h->torque_impl > 0 || (h->torque_impl < 0 && h->c.num_envs >= LT_TC5_MIN_ENVS);
    if (h->c.control_type == LT_CONTROL_ACTUATOR_NET && tc5) {
        LT_CUDA(h, lt_torque_tc5_launch(h->c, h->W, h->d_tc5_blob, actions, h->sim.dof_state, h->st, h->d_counters, substep, h->sm_count,
                                        0, (cudaStream_t)stream));
        h->substeps_pending += 1;
        return LT_OK;
    }
    if (h->c.control_type == LT_CONTROL_ACTUATOR_NET) {
        LT_CUDA(h, lt_torque_mma_launch(h->c, h->d_act_frag, actions, h->sim.dof_state, h->st, h->d_counters, substep, h->sm_count,
                                        (cudaStream_t)stream));
        h->substeps_pending += 1;
        return LT_OK;
    }
#endif
    // samples per thread: the vector kernel needs 16-byte aligned rows (torch allocations are); else one sample per thread
    uintptr_t al = (uintptr_t)actions | (uintptr_t)p.dof_state | (uintptr_t)p.st.torques | (uintptr_t)p.st.joint_pos_target |
                   (uintptr_t)p.st.actions | (uintptr_t)p.st.motor_offsets | (uintptr_t)p.st.motor_strengths |
                   (uintptr_t)p.st.joint_pos_err_last | (uintptr_t)p.st.joint_pos_err_last_last | (uintptr_t)p.st.joint_vel_last |
                   (uintptr_t)p.st.joint_vel_last_last | (uintptr_t)p.st.Kp_factors | (uintptr_t)p.st.Kd_factors |
                   (uintptr_t)p.st.lag_buffer | (uintptr_t)p.st.last_joint_pos_target | (uintptr_t)p.st.last_last_joint_pos_target;
    const int spt = (al & 15u) ? 1 : LT_TORQUE_SPT;
    const unsigned grid = (unsigned)((n / spt + LT_TORQUE_THREADS - 1) / LT_TORQUE_THREADS);
#if LT_TORQUE_SPT > 1
    if (spt > 1) lt_torque_kernel_v<LT_TORQUE_SPT><<<grid, LT_TORQUE_THREADS, 0, (cudaStream_t)stream>>>(p);
    else
#endif
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
        if (args->local_relative_linear && args->local_relative_rotation && args->local_target_poses) {
            p.args.local_target_poses = args->local_target_poses;
            p.args.local_relative_linear = args->local_relative_linear;
            p.args.local_relative_rotation = args->local_relative_rotation;
            p.args.replan = args->replan;
        }
        if (args->new_gravity) {
            p.sc.set_gravity = 1;
            for (int k = 0; k < 3; ++k) p.sc.new_gravity[k] = args->new_gravity[k];
        }
    }
    p.sc.term_sums = h->d_term_sums;
    p.noise_scale_vec = h->d_noise_scale;
    p.counters = h->d_counters;
    p.scratch = h->d_scratch;
    p.rec_bytes = h->rec_bytes;
}

// _get_heights over the full measured_points grid (LR:1918-1965): thread per (env, point); same index arithmetic as the
// observation path (lt_env_step: float32 add, subtract, DIVIDE by the cell size, truncate, clamp to [0, cells - 2])
struct lt_heights_params {
    lt_config c;
    lt_terrain_ptrs ter;
    const float* root_states;
    const float* base_rotation;
    float* out;
};
__global__ void __launch_bounds__(256) lt_heights_kernel(const __grid_constant__ lt_heights_params p) {
    const lt_config& c = p.c;
    const int np = c.n_points_x * c.n_points_y;
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= (int64_t)c.num_envs * np) return;
    const int i = (int)(q / np), pt = (int)(q % np), a = pt / c.n_points_y, b = pt % c.n_points_y;
    float* o = p.out + (size_t)i * 2 * np;
    if (c.plane) { o[pt] = 1.0f; o[np + pt] = 0.0f; return; }          // LR:1928-1932
    const float* root = p.root_states + (size_t)i * c.num_actors * 13;
    float offx = 0.0f, offy = 0.0f;
    if (c.camera_zero) {
        const float cp = cosf(p.base_rotation[(size_t)i * 3 + 1]);
        offx = LT_FMUL(c.camera_offset[0], cp);
        offy = LT_FMUL(c.camera_offset[1], cp);
    }
    float px = LT_FADD(c.points_x[a], root[0]), py = LT_FADD(c.points_y[b], root[1]);
    if (c.camera_zero) { px = LT_FADD(px, offx); py = LT_FADD(py, offy); }
    px = LT_FDIV(LT_FSUB(px, p.ter.env_terrain_origin[(size_t)i * 3 + 0]), c.horizontal_scale);
    py = LT_FDIV(LT_FSUB(py, p.ter.env_terrain_origin[(size_t)i * 3 + 1]), c.horizontal_scale);
    int ix = (int)px, iy = (int)py;
    ix = ix < 0 ? 0 : (ix > c.tile_rows - 2 ? c.tile_rows - 2 : ix);
    iy = iy < 0 ? 0 : (iy > c.tile_cols - 2 ? c.tile_cols - 2 : iy);
    const size_t cells = (size_t)c.tile_rows * c.tile_cols;
    const float* tile = p.ter.tile_heights + (size_t)p.ter.env_tile[i] * 2 * cells;
    o[pt] = tile[(size_t)ix * c.tile_cols + iy];
    o[np + pt] = tile[cells + (size_t)ix * c.tile_cols + iy];
}

int lt_measure_heights(lt_handle* h, float* out, void* stream) {
    const int rc = lt_ready(h);
    if (rc) return rc;
    if (!out) return lt_fail(h, LT_ERR_INVALID, "lt_measure_heights: out is NULL");
    lt_heights_params p;
    p.c = h->c;
    p.ter = h->ter;
    p.root_states = h->sim.root_states;
    p.base_rotation = h->st.base_rotation;
    p.out = out;
    const int64_t n = (int64_t)h->c.num_envs * h->c.n_points_x * h->c.n_points_y;
    lt_heights_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(p);
    return cudaGetLastError() == cudaSuccess ? LT_OK : lt_fail(h, LT_ERR_CUDA, "lt_heights_kernel launch failed");
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
    // ordered reset compaction + counters: needs only lt_step_kernel's masks, so it is forked onto the handle's side stream and
    // joins after lt_obs_kernel (under stream capture the side stream becomes a parallel branch of the graph)
    const int n = h->c.num_envs;
    cudaStream_t cs = (cudaStream_t)stream;
    if (h->fork_compact) {
        LT_CUDA(h, cudaEventRecord(h->ev_fork, (cudaStream_t)stream));
        LT_CUDA(h, cudaStreamWaitEvent(h->side, h->ev_fork, 0));
        cs = h->side;
    }
    LT_CUDA(h, lt_launch_chain(lt_compact_kernel, (unsigned)((n + LT_COMPACT_THREADS - 1) / LT_COMPACT_THREADS), LT_COMPACT_THREADS, 0,
                               cs, h->st.reset_buf, h->st.time_out_buf, h->st.time_outs_latched, h->st.reset_env_ids,
                               h->st.reset_count, h->d_counters, n, h->c.send_timeouts, h->substeps_pending, h->L.lag_slots));
    if (h->fork_compact) LT_CUDA(h, cudaEventRecord(h->ev_join, h->side));
    {
        const lt_config& c = h->c;
        lt_obs_params o;
        o.num_envs = c.num_envs; o.num_obs = c.num_obs; o.obs_history_length = c.obs_history_length;
        o.per_layer = h->L.nxf * c.n_points_y; o.nxf = h->L.nxf; o.h0 = h->L.h0; o.n_heights = h->L.n_heights; o.head_w = h->L.head_w;
        o.plane = c.plane; o.camera_zero = c.camera_zero; o.tile_cols = c.tile_cols; o.tile_cells = h->L.tile_cells;
        o.rec_bytes = h->rec_bytes; o.seed_lo = c.seed_lo; o.seed_hi = c.seed_hi;
        o.clip_observations = c.clip_observations; o.ceiling_height = c.ceiling_height; o.obs_scale_height = c.obs_scale_height;
        o.scratch = h->d_scratch; o.ptlut = h->d_ptlut; o.tile_heights = h->ter.tile_heights; o.env_tile = h->ter.env_tile;
        o.height_noise = h->height_noise;
        o.noise_scale_vec = h->d_noise_scale; o.noise_uniforms = p.args.noise_uniforms; o.counters = h->d_counters;
        o.obs_buf = h->st.obs_buf; o.obs_history = h->st.obs_history;
        const int mode = !c.add_noise ? 0 : (p.args.noise_uniforms ? 1 : 2);
        e = o.per_layer <= 128 ? lt_launch_obs<4>(o, mode, (cudaStream_t)stream) : lt_launch_obs<8>(o, mode, (cudaStream_t)stream);
        if (e != cudaSuccess) return lt_fail(h, LT_ERR_CUDA, std::string("lt_obs_kernel: ") + cudaGetErrorString(e));
    }
    if (h->fork_compact) LT_CUDA(h, cudaStreamWaitEvent((cudaStream_t)stream, h->ev_join, 0));
    h->substeps_pending = 0;
    return LT_OK;
}

}  // extern "C"
