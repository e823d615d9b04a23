// lt_planner.cu -- sampling-based local planner of LeggedRobot._plan_target_pose (go1_gym/envs/base/
// legged_robot_trajectory_tracking.py:856-916, LR below) + the relative pose of the local target (LR:914-916, 922-932).
//
// The reference loops over the planning envs in Python: per env it sorts the 1,125 candidate poses of
// legged_robot_trajectory_tracking_config.py:167-175 by distance to the goal, transforms all 2 x 21 x 11 measured height points
// into every candidate's yaw frame, and takes the first candidate whose robot-sized ellipsoid contains no point.  Here: one CTA
// per env, threads over candidates (early exit on the first point inside), an ordered 64-bit atomicMin over (sort key, index)
// picks the winner -- "first valid in sorted order" = "valid candidate with the smallest key" (ties: lowest candidate index, what a
// stable argsort gives).  Arithmetic goes through lt_core.cuh's helpers (same operation order as the torch expressions).
#include <cuda_runtime.h>
#include <stdint.h>

#include "lt_core.cuh"

struct lt_plan_params {
    lt_plan_args a;
    int32_t num_envs, n_points, n_cands, plan_interval, root_stride;
    float robot_size[3];
};

// quat_apply_yaw (math_utils.py:12-16): zero x, y; normalise; quat_apply (planner.py:17-23: b + w t + cross(xyz, t), t = 2 cross(xyz, b))
__device__ __forceinline__ void lt_quat_apply_yaw(const float q[4], const float v[3], float o[3]) {
    float n = LT_FSQRT(LT_FADD(LT_FMUL(q[2], q[2]), LT_FMUL(q[3], q[3])));
    n = fmaxf(n, 1e-9f);
    const float qy[4] = {LT_FDIV(0.0f, n), LT_FDIV(0.0f, n), LT_FDIV(q[2], n), LT_FDIV(q[3], n)};
    float t[3], u[3];
    lt_cross(qy, v, t);
#pragma unroll
    for (int k = 0; k < 3; ++k) t[k] = LT_FMUL(t[k], 2.0f);
    lt_cross(qy, t, u);
#pragma unroll
    for (int k = 0; k < 3; ++k) o[k] = LT_FADD(LT_FADD(v[k], LT_FMUL(qy[3], t[k])), u[k]);
}

#define LT_PLAN_THREADS 256

__global__ void __launch_bounds__(LT_PLAN_THREADS) lt_plan_kernel(const __grid_constant__ lt_plan_params p) {
    extern __shared__ float s_pts[];                  // [2 * n_points][3]: (x, y, measured height) of this env
    __shared__ unsigned long long s_best;
    __shared__ int s_mode;                            // 0 keep the local target, 1 take the global target, 2 search
    __shared__ float s_base[3 + 4 + 3], s_goal[2];
    const int e = blockIdx.x;
    if (e >= p.num_envs) return;
    const lt_plan_args& a = p.a;
    const float* root = a.root_states + (size_t)e * p.root_stride;
    const float* target = a.target_poses + (size_t)e * 6;
    float* local = a.local_target_poses + (size_t)e * 6;
    if (threadIdx.x == 0) {
        s_best = ~0ull;
        float rpy[3];
        lt_quat_to_rpy(root + 3, rpy);
        for (int k = 0; k < 3; ++k) { s_base[k] = root[k]; s_base[7 + k] = rpy[k]; }
        for (int k = 0; k < 4; ++k) s_base[3 + k] = root[3 + k];
        // relative pose of the GLOBAL target (LR:854-855): close_to_goal = |rel.xy| < 1 (LR:859)
        const float d[3] = {LT_FSUB(target[0], root[0]), LT_FSUB(target[1], root[1]), LT_FSUB(target[2], root[2])};
        float rel[3];
        lt_quat_apply_yaw_inverse(root + 3, d, rel);
        const bool close = lt_norm2(rel[0], rel[1]) < 1.0f;
        // LR:858-868: plan_length_buf += 1; replan on the interval; always at the first step of an episode
        long long pl = a.plan_length_buf[e] + 1;
        const bool ep_start = a.episode_length_buf[e] == 1;
        const bool replan = (pl % p.plan_interval) == 0;
        if (a.replan) a.replan[e] = replan ? 1 : 0;
        bool plan = ep_start || (replan && a.plan_buf[e] != 0);
        if (plan) pl = 0;                             // LR:874
        a.plan_length_buf[e] = pl;
        s_mode = !plan ? 0 : (close ? 1 : 2);
        s_goal[0] = LT_FSUB(target[0], root[0]);      // goal_pose[:2] -= base_pos[:2] (LR:881): world axes, not rotated
        s_goal[1] = LT_FSUB(target[1], root[1]);
        if (a.chosen) a.chosen[e] = !plan ? -3 : (close ? -2 : -1);
    }
    __syncthreads();
    const int mode = s_mode;
    if (mode == 2) {
        const int np = p.n_points;
        const float* mh = a.measured_heights + (size_t)e * 2 * np;
        for (int k = threadIdx.x; k < 2 * np; k += blockDim.x) {
            const int pt = k % np;
            s_pts[3 * k] = a.points_xy[2 * pt];
            s_pts[3 * k + 1] = a.points_xy[2 * pt + 1];
            s_pts[3 * k + 2] = mh[k];                                         // height_points[..., 2] = measured_heights[env] (LR:894)
        }
        __syncthreads();
        for (int c = threadIdx.x; c < p.n_cands; c += blockDim.x) {
            const float* cd = a.candidates + (size_t)c * 6;
            float q[4];
            lt_quat_from_euler(cd[3], cd[4], cd[5], q);                       // LR:892
            bool valid = true;
            for (int k = 0; k < 2 * np && valid; ++k) {
                const float d[3] = {LT_FSUB(s_pts[3 * k], cd[0]), LT_FSUB(s_pts[3 * k + 1], cd[1]), LT_FSUB(s_pts[3 * k + 2], cd[2])};
                float r[3];
                lt_quat_apply_yaw_inverse(q, d, r);                           // LR:897
                valid = lt_norm3(LT_FDIV(r[0], p.robot_size[0]), LT_FDIV(r[1], p.robot_size[1]), LT_FDIV(r[2], p.robot_size[2])) > 1.0f;   // LR:898
            }
            if (valid) {
                // sort key of LR:887-890 (non-negative: its bit pattern orders like the float)
                const float key = LT_FADD(lt_norm2(LT_FSUB(cd[0], s_goal[0]), LT_FSUB(cd[1], s_goal[1])),
                                          LT_FMUL(lt_norm3(cd[3], cd[4], cd[5]), 0.1f));
                atomicMin(&s_best, ((unsigned long long)__float_as_uint(key) << 32) | (unsigned)c);
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        float out[6];
        if (mode == 0) { for (int k = 0; k < 6; ++k) out[k] = local[k]; }
        else {
            for (int k = 0; k < 6; ++k) out[k] = target[k];                  // close to the goal, or no valid candidate (LR:878, 910)
            if (mode == 2 && s_best != ~0ull) {
                const int c = (int)(s_best & 0xffffffffu);
                const float* cd = a.candidates + (size_t)c * 6;
                const float lin[3] = {cd[0], cd[1], cd[2]};
                float w[3];
                lt_quat_apply_yaw(s_base + 3, lin, w);                        // LR:905
                out[0] = LT_FADD(w[0], s_base[0]);
                out[1] = LT_FADD(w[1], s_base[1]);
                out[2] = cd[2];
                for (int k = 0; k < 3; ++k) out[3 + k] = lt_wrap_to_pi(LT_FADD(cd[3 + k], s_base[7 + k]));       // LR:906
                if (a.chosen) a.chosen[e] = c;
            }
            for (int k = 0; k < 6; ++k) local[k] = out[k];
        }
        // _compute_relative_target_pose(local_target_poses) (LR:914-916, 922-932)
        const float d[3] = {LT_FSUB(out[0], s_base[0]), LT_FSUB(out[1], s_base[1]), LT_FSUB(out[2], s_base[2])};
        float rel[3];
        lt_quat_apply_yaw_inverse(s_base + 3, d, rel);
        for (int k = 0; k < 3; ++k) {
            if (a.local_relative_linear) a.local_relative_linear[(size_t)e * 3 + k] = rel[k];
            if (a.local_relative_rotation) a.local_relative_rotation[(size_t)e * 3 + k] = lt_wrap_to_pi(LT_FSUB(out[3 + k], s_base[7 + k]));
        }
    }
}

extern "C" int lt_plan_target_pose(const lt_plan_args* args, int32_t num_envs, int32_t n_points, int32_t n_cands, int32_t plan_interval,
                                   int32_t root_stride, const float* robot_size, void* stream) {
    if (!args || !robot_size || num_envs <= 0 || n_points <= 0 || n_cands <= 0 || plan_interval <= 0 || root_stride < 13) return LT_ERR_INVALID;
    if (!args->root_states || !args->target_poses || !args->points_xy || !args->measured_heights || !args->candidates ||
        !args->episode_length_buf || !args->plan_buf || !args->plan_length_buf || !args->local_target_poses)
        return LT_ERR_INVALID;
    lt_plan_params p;
    p.a = *args;
    p.num_envs = num_envs;
    p.n_points = n_points;
    p.n_cands = n_cands;
    p.plan_interval = plan_interval;
    p.root_stride = root_stride;
    for (int k = 0; k < 3; ++k) p.robot_size[k] = robot_size[k];
    const size_t smem = (size_t)2 * n_points * 3 * sizeof(float);
    if (smem > 48 * 1024) return LT_ERR_INVALID;
    lt_plan_kernel<<<num_envs, LT_PLAN_THREADS, smem, (cudaStream_t)stream>>>(p);
    return cudaGetLastError() == cudaSuccess ? LT_OK : LT_ERR_CUDA;
}
