// lt_rollout.cu -- rollout storage + GAE kernels (reference go1_gym_learn/ppo_cse/rollout_storage.py).
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/lt_b200.h"

// ---------------------------------------------------------------------------------------------------------------
// RolloutStorage.add_transitions (rollout_storage.py:57-71): every field of one transition copied into row t of
// its [T, N, w] buffer by ONE launch (the reference issues 11 copy_ kernels).  Work is laid out as a flat list of
// chunks over the concatenated fields: 16-byte float4 copies for aligned segments, scalar otherwise.
struct lt_copy_seg { const float* src; float* dst; long long n4; int vec; };   // n4 chunks: float4 when vec (16 B aligned), else float
struct lt_store_params {
    lt_copy_seg seg[9];
    int n_seg;
    long long total4;
    // rewards with the time-out bootstrap (ppo.py:85-87) and dones (bool -> uint8 buffer)
    const float *rewards, *values;
    const uint8_t *time_outs, *dones;
    float* rewards_dst;
    uint8_t* dones_dst;
    float gamma;
    int n;
};

__global__ void __launch_bounds__(256) lt_rollout_store_kernel(const __grid_constant__ lt_store_params p) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (long long k = tid; k < p.total4; k += stride) {
        long long r = k;
        int s = 0;
        while (r >= p.seg[s].n4) { r -= p.seg[s].n4; ++s; }
        if (p.seg[s].vec) reinterpret_cast<float4*>(p.seg[s].dst)[r] = reinterpret_cast<const float4*>(p.seg[s].src)[r];
        else p.seg[s].dst[r] = p.seg[s].src[r];
    }
    if (tid < p.n) {
        if (p.rewards_dst) {
            float r = p.rewards[tid];
            if (p.time_outs && p.values) r = __fadd_rn(r, __fmul_rn(__fmul_rn(p.gamma, p.values[tid]), p.time_outs[tid] ? 1.0f : 0.0f));
            p.rewards_dst[tid] = r;
        }
        if (p.dones_dst) p.dones_dst[tid] = p.dones[tid];
    }
}

// ---------------------------------------------------------------------------------------------------------------
// RolloutStorage.compute_returns (rollout_storage.py:76-90): one thread per env walks T backwards (the [T,N]
// layout makes every step's access coalesced across the warp); block-reduced sum / sum-of-squares of the raw
// advantages feed the normalisation.
__global__ void __launch_bounds__(256) lt_gae_kernel(const float* __restrict__ rewards, const float* __restrict__ values,
                                                     const uint8_t* __restrict__ dones, const float* __restrict__ last_values,
                                                     float* __restrict__ returns, float* __restrict__ advantages,
                                                     double* __restrict__ stats, int T, int n, float gamma, float lam) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    double s1 = 0.0, s2 = 0.0;
    if (i < n) {
        float adv = 0.0f;
        float next_v = last_values[i];
        for (int t = T - 1; t >= 0; --t) {
            const size_t k = (size_t)t * n + i;
            const float nt = __fsub_rn(1.0f, dones[k] ? 1.0f : 0.0f);                       // next_is_not_terminal
            const float v = values[k];
            // delta = r + next_is_not_terminal * gamma * next_values - v     (rollout_storage.py:84)
            const float delta = __fsub_rn(__fadd_rn(rewards[k], __fmul_rn(__fmul_rn(nt, gamma), next_v)), v);
            // advantage = delta + next_is_not_terminal * gamma * lam * advantage   (:85)
            adv = __fadd_rn(delta, __fmul_rn(__fmul_rn(__fmul_rn(nt, gamma), lam), adv));
            const float ret = __fadd_rn(adv, v);                                             // :86
            returns[k] = ret;
            const float a = __fsub_rn(ret, v);                                               // :89 advantages = returns - values
            advantages[k] = a;
            s1 += (double)a;
            s2 += (double)a * (double)a;
            next_v = v;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
    __shared__ double sh1[8], sh2[8];
    if ((threadIdx.x & 31) == 0) { sh1[threadIdx.x >> 5] = s1; sh2[threadIdx.x >> 5] = s2; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { a += sh1[w]; b += sh2[w]; }
        atomicAdd(stats + 0, a);
        atomicAdd(stats + 1, b);
        if (blockIdx.x == 0) stats[2] = (double)T * (double)n;
    }
}

// (A - mean) / (std + 1e-8) with the unbiased std of torch.Tensor.std() (rollout_storage.py:90)
__global__ void __launch_bounds__(256) lt_normalize_kernel(float* __restrict__ adv, const double* __restrict__ stats, long long count) {
    const double n = stats[2];
    const double mean = stats[0] / n;
    double var = (stats[1] - n * mean * mean) / (n - 1.0);
    if (var < 0.0) var = 0.0;
    const float m = (float)mean;
    const float den = __fadd_rn((float)sqrt(var), 1e-8f);
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < count; k += stride)
        adv[k] = __fdiv_rn(__fsub_rn(adv[k], m), den);
}

extern "C" {

int lt_rollout_store(const lt_rollout_ptrs* st, const lt_transition_ptrs* tr, int32_t t, int32_t n, int32_t num_obs,
                     int32_t num_priv, int32_t num_hist, int32_t num_actions, float gamma, void* stream) {
    if (!st || !tr || n <= 0 || t < 0) return LT_ERR_INVALID;
    lt_store_params p;
    p.n_seg = 0;
    p.total4 = 0;
    auto add = [&](const float* src, float* dst, long long w) {
        if (!src || !dst || w <= 0) return;
        float* d = dst + (size_t)t * n * w;
        const long long cnt = (long long)n * w;
        lt_copy_seg& s = p.seg[p.n_seg++];
        s.src = src;
        s.dst = d;
        s.vec = ((((uintptr_t)src | (uintptr_t)d) & 15u) == 0 && cnt % 4 == 0) ? 1 : 0;
        s.n4 = s.vec ? cnt / 4 : cnt;
        p.total4 += s.n4;
    };
    add(tr->observations, st->observations, num_obs);
    add(tr->privileged_observations, st->privileged_observations, num_priv);
    add(tr->observation_histories, st->observation_histories, num_hist);
    add(tr->actions, st->actions, num_actions);
    add(tr->values, st->values, 1);
    add(tr->actions_log_prob, st->actions_log_prob, 1);
    add(tr->action_mean, st->mu, num_actions);
    add(tr->action_sigma, st->sigma, num_actions);
    p.rewards = tr->rewards;
    p.values = tr->values;
    p.time_outs = tr->time_outs;
    p.dones = tr->dones;
    p.rewards_dst = (tr->rewards && st->rewards) ? st->rewards + (size_t)t * n : nullptr;
    p.dones_dst = (tr->dones && st->dones) ? st->dones + (size_t)t * n : nullptr;
    p.gamma = gamma;
    p.n = n;
    long long work = p.total4 > n ? p.total4 : n;
    long long blocks = (work + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (blocks * 256 < n) blocks = (n + 255) / 256;
    lt_rollout_store_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(p);
    return cudaGetLastError() == cudaSuccess ? LT_OK : LT_ERR_CUDA;
}

int lt_gae(const float* rewards, const float* values, const uint8_t* dones, const float* last_values, float* returns,
           float* advantages, double* stats, int32_t T, int32_t n, float gamma, float lam, void* stream) {
    if (!rewards || !values || !dones || !last_values || !returns || !advantages || !stats || T <= 0 || n <= 0) return LT_ERR_INVALID;
    if (cudaMemsetAsync(stats, 0, 3 * sizeof(double), (cudaStream_t)stream) != cudaSuccess) return LT_ERR_CUDA;
    lt_gae_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(rewards, values, dones, last_values, returns, advantages, stats, T, n,
                                                                       gamma, lam);
    return cudaGetLastError() == cudaSuccess ? LT_OK : LT_ERR_CUDA;
}

int lt_normalize_advantages(float* advantages, const double* stats, int64_t count, void* stream) {
    if (!advantages || !stats || count <= 0) return LT_ERR_INVALID;
    long long blocks = (count + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    lt_normalize_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(advantages, stats, (long long)count);
    return cudaGetLastError() == cudaSuccess ? LT_OK : LT_ERR_CUDA;
}

}  // extern "C"
