// lt_policy.cu -- actor-critic forward and the PPO minibatch step (include/lt_policy.h) on top of the tcgen05 contraction
// kernel of lt_gemm.cuh.  AC = go1_gym_learn/ppo_cse/actor_critic.py, PPO = go1_gym_learn/ppo_cse/ppo.py of the reference.
//
// Data layout.  Every GEMM operand is a pair of fp16 planes [2][rows][ld] (value * scale = hi + lo).  One shared input
// buffer X0 = [obs_history | latent | privileged_obs] feeds the first layer of all three MLPs; their first-layer weights
// are stored expanded to that width with zeros in the columns a network does not read (adaptation: obs_history only;
// actor: obs_history | latent, AC:130; critic: obs_history | privileged_obs, AC:158).  Hidden activations Y_l are written by
// the GEMM epilogue already split, row-major, with one extra column right of the last feature that holds the constant one
// (written once, never overwritten): the weight-gradient GEMM dW_l = dZ_l^T * [Y_{l-1} | 1] reads both row-major buffers as
// MN-major tensor-core operands (the contraction runs over rows), so its last output column is the bias gradient and nothing
// is ever stored transposed.
// Scales are powers of two, undone exactly in the epilogues: activations 2^4, weights 2^6.
//
// Gradient scales.  A gradient tensor dZ_l spans many orders of magnitude across layers (the adaptation module sees the actor's
// gradient through four weight matrices), and an fp16 "lo" plane keeps its 11 bits only while value * scale stays above 2^-14
// * 2^11: so every dZ tensor gets its OWN power-of-two scale, kept in a small device table and chosen so that the tensor's
// largest entry lands in [32, 64) -- three decades of headroom below the 60000 clamp, five above the subnormal floor.  The
// largest entry is measured by the kernels that write the tensor (atomicMax in the epilogue / head); the scales used by a call
// are the ones measured by the previous call of the same kind, and the very first call runs its backward pass twice (once at
// 2^ceil(log2(rows)) to measure).  No host synchronisation anywhere.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/lt_b200.h"
#include "../../include/lt_policy.h"
#include "lt_gemm.cuh"

namespace {

constexpr float S_X = 16.0f;          // activations / inputs
constexpr float S_W = 64.0f;          // weights
constexpr int HEAD_THREADS = 256;
constexpr int STAT_SLOTS = 48;        // per-block partial row of the head kernels: 3 loss sums + up to 45 action columns

std::string g_policy_error;

struct Planes {
    __half* p = nullptr;
    int64_t rows = 0, ld = 0, plane = 0;          // [2][rows][ld]; plane = rows * ld
    __half* lo() const { return p + plane; }
};

// Column map of a first layer: parameter columns [p, p + n) of the nn.Linear <-> GEMM columns [g, g + n) of the shared input
// planes, up to three runs; GEMM columns outside every run carry zero weights (and their weight gradient is dropped).
struct SegMap {
    int n = 0;
    int p[3] = {0, 0, 0}, g[3] = {0, 0, 0}, len[3] = {0, 0, 0};
    void add(int pp, int gg, int ll) {
        if (ll <= 0) return;
        p[n] = pp; g[n] = gg; len[n] = ll; ++n;
    }
};
__host__ __device__ inline int seg_param_col(const SegMap& m, int k) {           // GEMM column -> parameter column or -1
    for (int i = 0; i < m.n; ++i)
        if (k >= m.g[i] && k < m.g[i] + m.len[i]) return m.p[i] + (k - m.g[i]);
    return -1;
}

struct Layer {
    int K = 0, N = 0;                 // GEMM reduction width (expanded for first layers) and outputs
    int k_real = 0;                   // in_features of the nn.Linear
    SegMap map;
    int64_t w_off = 0, b_off = 0;     // offsets in the flat parameter vector
    float* bias16 = nullptr;          // 16-byte aligned copy of the bias (vector loads in the GEMM epilogue), refreshed with W
    Planes W, WT;                     // [2][N][ld(K)] and (training) [2][K][ld(N)] for the input-gradient GEMM
};

struct Mlp {
    int n = 0;                                    // number of Linear layers
    Layer L[LT_POLICY_MAX_HIDDEN + 1];
    Planes Y[LT_POLICY_MAX_HIDDEN];               // hidden activations [R][N_l + 1]: column N_l is the constant one
    Planes dZ[LT_POLICY_MAX_HIDDEN + 1];
    float* out = nullptr;                         // [R][N_last] fp32
    int idx = 0;                                  // 0 adaptation, 1 actor, 2 critic, 3 encoder: row of the gradient-scale table
    const Planes* X = nullptr;                    // input planes of layer 0
    const Planes* final_act = nullptr;            // != NULL: the LAST layer is followed by the activation too and lands, split, here
};

constexpr int GS_PER_MLP = LT_POLICY_MAX_HIDDEN + 1;
constexpr int GS_N = 4 * GS_PER_MLP;              // one scale per dZ tensor
enum { CTX_PPO = 0, CTX_ADAPT = 1 };              // lt_ppo_minibatch_grad and lt_adaptation_grad see different magnitudes

}   // namespace

struct lt_policy {
    lt_policy_config c;
    int sm_count = 0;
    int64_t R = 0;                                // row capacity (multiple of 128)
    int Din = 0, P = 0, A = 0, K0 = 0;            // Din: width of what the bodies see in place of obs_history inside X0
    int Dhist = 0;                                // width of an obs_history row as the caller stores it
    Mlp adapt, actor, critic;
    Planes X0;                                    // [2][R][ld(K0 + 1)]: column K0 is the constant one
    // ppo_cse_cnn height-map encoder (include/lt_policy.h): X0 = [emb (E) | s (E0) | s (E0) | latent | privileged | 1]
    bool has_enc = false;
    int HM = 0, E = 0, E0 = 0, last_off = 0;      // last_off: first column of the last observation row inside obs_history
    Mlp enc;
    Planes XH;                                    // [2][R][ld(HM + 1)]: the height columns of the last observation row
    float* enc_part = nullptr;                    // [3][R][E] fp32: emb-column input gradients of the three bodies
    const void* packed_src = nullptr;             // what X0 currently holds: (obs_history, indices, count) of the last gather
    const void* packed_idx = nullptr;
    int packed_count = 0;
    float* partial = nullptr;                     // split-K workspace
    int64_t partial_floats = 0;
    float* head_partial = nullptr;                // [blocks][STAT_SLOTS]
    float* norm_partial = nullptr;                // [blocks]
    float* scalars = nullptr;                     // device float[8]: [0] clip coefficient, [1] total norm
    float* gscale = nullptr;                      // device float[2][GS_N]: gradient scales per call kind
    uint32_t* gamax = nullptr;                    // device uint32[2][GS_N]: measured max |dz| (fp32 bits) of the running call
    bool calibrated[2] = {false, false};
    float* gs(int ctx, const Mlp& m, int l) const { return gscale + ctx * GS_N + m.idx * GS_PER_MLP + l; }
    uint32_t* ga(int ctx, const Mlp& m, int l) const { return gamax + ctx * GS_N + m.idx * GS_PER_MLP + l; }
    int64_t n_params = 0;
    float *params = nullptr, *grads = nullptr;
    std::vector<void*> allocs;
    std::string err;
};

namespace {

void set_error(lt_policy* P, const std::string& m) {
    if (P) P->err = m; else g_policy_error = m;
}

#define LTP_CUDA(P, call)                                                                                       \
    do {                                                                                                        \
        cudaError_t e_ = (call);                                                                                \
        if (e_ != cudaSuccess) {                                                                                \
            std::string m_ = std::string(#call) + ": " + cudaGetErrorString(e_);                                \
            set_error((lt_policy*)(P), m_);                                                                     \
            return LT_ERR_CUDA;                                                                                 \
        }                                                                                                       \
    } while (0)

int fail(lt_policy* P, int code, const std::string& msg) {
    set_error(P, msg);
    return code;
}

inline int64_t up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

int alloc_planes(lt_policy* P, Planes& pl, int64_t rows, int64_t width) {
    pl.rows = rows;
    pl.ld = up(width, 8);
    pl.plane = pl.rows * pl.ld;
    void* d = nullptr;
    LTP_CUDA(P, cudaMalloc(&d, (size_t)(2 * pl.plane) * sizeof(__half)));
    LTP_CUDA(P, cudaMemset(d, 0, (size_t)(2 * pl.plane) * sizeof(__half)));
    P->allocs.push_back(d);
    pl.p = (__half*)d;
    return LT_OK;
}
template <typename T>
int alloc_dev(lt_policy* P, T*& out, int64_t n) {
    void* d = nullptr;
    LTP_CUDA(P, cudaMalloc(&d, (size_t)n * sizeof(T)));
    LTP_CUDA(P, cudaMemset(d, 0, (size_t)n * sizeof(T)));
    P->allocs.push_back(d);
    out = (T*)d;
    return LT_OK;
}

inline int pick_bn(int n) {                      // smallest multiple of 16 that covers n in ceil(n / 256) equal tiles
    const int tiles = (n + 255) / 256;
    return (int)up((n + tiles - 1) / tiles, 16);
}

// ------------------------------------------------------------------------------------------------------------ kernels
__device__ __forceinline__ void split1(float x, __half& hi, __half& lo) {
    x = fminf(fmaxf(x, -ltg::OPERAND_MAX), ltg::OPERAND_MAX);
    hi = __float2half_rn(x);
    lo = __float2half_rn(x - __half2float(hi));
}

// master weights -> operand planes W [N][K] and WT [K][N] of one layer (expanded column map for first layers)
__global__ void split_weights_kernel(const float* __restrict__ w, int N, int K, int k_real, const SegMap map, __half* W, int64_t ldW,
                                     int64_t planeW, __half* WT, int64_t ldT, int64_t planeT, const float* __restrict__ bias,
                                     float* __restrict__ bias16) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N) bias16[i] = bias[i];
    if (i >= (int64_t)N * K) return;
    const int n = (int)(i / K), k = (int)(i % K);
    const int src = seg_param_col(map, k);
    const float v = src >= 0 ? w[(int64_t)n * k_real + src] * S_W : 0.0f;
    __half hi, lo;
    split1(v, hi, lo);
    W[(int64_t)n * ldW + k] = hi;
    W[planeW + (int64_t)n * ldW + k] = lo;
    if (WT) {
        WT[(int64_t)k * ldT + n] = hi;
        WT[planeT + (int64_t)k * ldT + n] = lo;
    }
}

// Gather + split the shared first-layer input: X0[r] = [obs_history[src] | 0 (or privileged_obs, teacher mode) | privileged_obs[src]].
// One warp per row: lanes sweep the columns, so both the fp32 reads and the fp16 writes are coalesced.
__global__ void __launch_bounds__(256) pack_x0_kernel(const float* __restrict__ obs_history, const float* __restrict__ priv,
                                                      const int64_t* __restrict__ indices, int rows, int Din, int P, int teacher,
                                                      __half* __restrict__ x, int64_t ldx, int64_t planex) {
    const int lane = threadIdx.x & 31;
    const int r = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (r >= rows) return;
    const int64_t src = indices ? indices[r] : (int64_t)r;
    const float* oh = obs_history + src * Din;
    __half* xh = x + (int64_t)r * ldx;
    __half* xl = xh + planex;
    for (int c = lane; c < Din; c += 32) {
        __half hi, lo;
        split1(__ldg(oh + c) * S_X, hi, lo);
        xh[c] = hi;
        xl[c] = lo;
    }
    if (lane < 2 * P) {
        const int j = lane < P ? lane : lane - P;
        float v = 0.0f;
        if (priv && (lane >= P || teacher)) v = priv[src * P + j];
        __half hi, lo;
        split1(v * S_X, hi, lo);
        xh[Din + lane] = hi;
        xl[Din + lane] = lo;
    }
}

// constant-one column (value 1 * S_X, hi plane) of an activation buffer: carries the bias gradient through the dW GEMM
__global__ void ones_col_kernel(__half* col_hi, int64_t rows, int64_t ld) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < rows) col_hi[i * ld] = __float2half_rn(S_X);
}

// latent (adaptation module output, fp32) -> its columns of X0 (AC:129-130: actor input = cat(obs_history, latent))
__global__ void scatter_latent_kernel(const float* __restrict__ latent, int rows, int P, int Din, __half* x, int64_t ldx, int64_t planex) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    for (int j = 0; j < P; ++j) {
        __half hi, lo;
        split1(latent[(int64_t)r * P + j] * S_X, hi, lo);
        const int64_t o = (int64_t)r * ldx + Din + j;
        x[o] = hi;
        x[planex + o] = lo;
    }
}

// ppo_cse_cnn gather (CNN:181-198): the last observation row of obs_history[src] goes, split, to X0 = [. (emb, written by the
// encoder) | s | s | . (latent) | privileged] and its height columns to XH.  One warp per row.
__global__ void __launch_bounds__(256) pack_enc_kernel(const float* __restrict__ obs_history, const float* __restrict__ priv,
                                                       const int64_t* __restrict__ indices, int rows, int Dhist, int last_off, int E0, int HM, int E,
                                                       int P, int teacher, __half* __restrict__ x, int64_t ldx, int64_t planex,
                                                       __half* __restrict__ xh, int64_t ldh, int64_t planeh) {
    const int lane = threadIdx.x & 31;
    const int r = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (r >= rows) return;
    const int64_t src = indices ? indices[r] : (int64_t)r;
    const float* o = obs_history + src * Dhist + last_off;
    __half* x_hi = x + (int64_t)r * ldx;
    __half* x_lo = x_hi + planex;
    for (int c = lane; c < E0; c += 32) {
        __half hi, lo;
        split1(__ldg(o + c) * S_X, hi, lo);
        x_hi[E + c] = hi;
        x_lo[E + c] = lo;
        x_hi[E + E0 + c] = hi;
        x_lo[E + E0 + c] = lo;
    }
    __half* h_hi = xh + (int64_t)r * ldh;
    __half* h_lo = h_hi + planeh;
    for (int c = lane; c < HM; c += 32) {
        __half hi, lo;
        split1(__ldg(o + E0 + c) * S_X, hi, lo);
        h_hi[c] = hi;
        h_lo[c] = lo;
    }
    if (lane < 2 * P) {
        const int Din = E + 2 * E0;
        const int j = lane < P ? lane : lane - P;
        float v = 0.0f;
        if (priv && (lane >= P || teacher)) v = priv[src * P + j];
        __half hi, lo;
        split1(v * S_X, hi, lo);
        x_hi[Din + lane] = hi;
        x_lo[Din + lane] = lo;
    }
}

// split-K partials of dW^T ... summed in a fixed order into the flat gradient: partial[s][n][k], k < K -> weight (through the
// first-layer column map), k == K -> bias
__global__ void reduce_dw_kernel(const float* __restrict__ partial, int splits, int64_t split_stride, int64_t ldp, int N, int K, int k_real,
                                 const SegMap map, float* __restrict__ gw, float* __restrict__ gb) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)N * (K + 1)) return;
    const int n = (int)(i / (K + 1)), k = (int)(i % (K + 1));
    const int dst = k == K ? -2 : seg_param_col(map, k);
    if (dst == -1) return;
    float acc = 0.0f;
    const float* src = partial + (int64_t)n * ldp + k;
    for (int s = 0; s < splits; ++s) acc += src[(int64_t)s * split_stride];
    if (dst == -2) gb[n] = acc;
    else gw[(int64_t)n * k_real + dst] = acc;
}

__global__ void sum_partials_kernel(const float* __restrict__ partial, int splits, int64_t n, int N, const float* __restrict__ bias, float* __restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float acc = 0.0f;
    for (int s = 0; s < splits; ++s) acc += partial[(int64_t)s * n + i];
    out[i] = acc + (bias ? bias[i % N] : 0.0f);
}

__device__ __forceinline__ float block_sum(float v, float* sh) {            // deterministic: fixed tree
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) sh[w] = v;
    __syncthreads();
    float t = 0.0f;
    if (threadIdx.x < 32) {
        t = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.0f;
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    }
    return t;                                                               // valid in thread 0
}

// warp max of |v| -> one atomicMax per warp on the fp32 bit pattern (non-negative floats order like unsigned integers)
__device__ __forceinline__ void record_amax(uint32_t* dst, float v) {
    const uint32_t m = __reduce_max_sync(0xffffffffu, __float_as_uint(fminf(v, 3.0e38f)));
    if ((threadIdx.x & 31) == 0 && m) atomicMax(dst, m);
}

// scale <- the power of two that puts the measured largest entry into [32, 64); the measurement is cleared for the next call
__global__ void update_scales_kernel(uint32_t* __restrict__ amax, float* __restrict__ scale, int n) {
    const int i = threadIdx.x;
    if (i >= n) return;
    const float a = __uint_as_float(amax[i]);
    if (a > 0.0f && a < 3.0e38f) {
        int e;
        frexpf(a, &e);                                            // a = m * 2^e, m in [0.5, 1)
        scale[i] = ldexpf(1.0f, min(max(6 - e, -100), 100));
    }
    amax[i] = 0u;
}
__global__ void fill_kernel(float* __restrict__ dst, float v, int n) {
    if ((int)threadIdx.x < n) dst[threadIdx.x] = v;
}

// gradient of the encoder's (activated) output: dz = (sum over the bodies of their emb-column input gradients) .* act'(emb),
// measured (amax) and split with the tensor's scale into the encoder's last dZ planes.  Thread per (row, 4 columns).
__global__ void __launch_bounds__(256) enc_grad_kernel(const float* __restrict__ part, int n_parts, int64_t part_stride, int count, int E, int act,
                                                       const __half* __restrict__ y, int64_t ldy, int64_t planey,
                                                       const float* __restrict__ scale_ptr, uint32_t* __restrict__ amax,
                                                       __half* __restrict__ dz, int64_t ldz, int64_t planez) {
    using namespace ltg;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int q = E >> 2;
    float m = 0.0f;
    if (i < (int64_t)count * q) {
        const int r = (int)(i / q), c = (int)(i % q) * 4;
        float4 g = *reinterpret_cast<const float4*>(part + (int64_t)r * E + c);
        for (int s = 1; s < n_parts; ++s) {
            const float4 t = *reinterpret_cast<const float4*>(part + s * part_stride + (int64_t)r * E + c);
            g.x += t.x; g.y += t.y; g.z += t.z; g.w += t.w;
        }
        const float scale = __ldg(scale_ptr);
        const float gv[4] = {g.x, g.y, g.z, g.w};
        const __half* yh = y + (int64_t)r * ldy + c;
        __half* zh = dz + (int64_t)r * ldz + c;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float yv = (__half2float(yh[j]) + __half2float(yh[planey + j])) * (1.0f / S_X);
            float d = gv[j];
            LTG_ACT_SWITCH(act, d *= act_bwd<A_>(yv);)
            m = fmaxf(m, fabsf(d));
            __half hi, lo;
            split1(d * scale, hi, lo);
            zh[j] = hi;
            zh[planez + j] = lo;
        }
    }
    record_amax(amax, m);
}


struct HeadParams {
    lt_ppo_batch_ptrs b;
    const int64_t* indices;
    int count, A;
    const float* mu;            // [count][A] actor output
    const float* value;         // [count]
    const float* std;           // [A]
    lt_ppo_hparams hp;
    float inv_count;            // 1 / (count * grad_world)
    const float *s_mu, *s_v;    // device: scales of the two head gradient tensors
    uint32_t *amax_mu, *amax_v;
    __half *dmu, *dv;                                 // planes
    int64_t ld_dmu, plane_dmu, ld_dv, plane_dv;
    float* partial;             // [blocks][STAT_SLOTS]
};

// Loss head of one minibatch (PPO:104-141) and its gradient with respect to the actor mean, the critic value and std.
__global__ void __launch_bounds__(HEAD_THREADS) ppo_head_kernel(const __grid_constant__ HeadParams p) {
    __shared__ float sh[HEAD_THREADS / 32];
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    const bool ok = r < p.count;
    float surr = 0.0f, vl = 0.0f, kl = 0.0f, g_logp = 0.0f;
    float dstd[32];
    float amax_mu = 0.0f, amax_v = 0.0f;
    const float s_mu = __ldg(p.s_mu), s_v = __ldg(p.s_v);
    const int A = p.A;
    int64_t s = 0;
    if (ok) {
        s = p.indices ? p.indices[r] : (int64_t)r;
        float logp = 0.0f;
        for (int j = 0; j < A; ++j) {
            const float mu = p.mu[(int64_t)r * A + j], sg = p.std[j], a = p.b.actions[s * A + j];
            const float d = a - mu;
            logp += -(d * d) / (2.0f * sg * sg) - logf(sg) - 0.91893853320467274178f;            // Normal.log_prob
            const float omu = p.b.old_mu[s * A + j], osg = p.b.old_sigma[s * A + j];
            kl += logf(sg / osg + 1.e-5f) + (osg * osg + (omu - mu) * (omu - mu)) / (2.0f * sg * sg) - 0.5f;     // PPO:112-116
        }
        const float adv = p.b.advantages[s], ratio = expf(logp - p.b.old_actions_log_prob[s]);
        const float lo = 1.0f - p.hp.clip_param, hi = 1.0f + p.hp.clip_param;
        const float s1 = -adv * ratio, s2 = -adv * fminf(fmaxf(ratio, lo), hi);
        surr = fmaxf(s1, s2);
        const bool inside = ratio >= lo && ratio <= hi;
        g_logp = (inside || s1 > s2) ? s1 : 0.0f;                    // d surrogate / d logp = -adv * ratio where the max picks the unclipped branch
        const float v = p.value[r], R = p.b.returns[s];
        float gv;
        if (p.hp.use_clipped_value_loss) {
            const float V = p.b.values[s];
            const float dv = v - V;
            const float vc = V + fminf(fmaxf(dv, -p.hp.clip_param), p.hp.clip_param);
            const float l1 = (v - R) * (v - R), l2 = (vc - R) * (vc - R);
            vl = fmaxf(l1, l2);
            const bool in_v = dv >= -p.hp.clip_param && dv <= p.hp.clip_param;
            if (in_v || l1 > l2) gv = 2.0f * (v - R);
            else if (l1 < l2) gv = 0.0f;
            else gv = (v - R);                                       // tie with a clipped value: torch.max splits the gradient
        } else {
            vl = (R - v) * (R - v);
            gv = 2.0f * (v - R);
        }
        // gradients of loss = mean(surr) + c_v * mean(vl) - c_e * mean(entropy)
        const float gs = g_logp * p.inv_count;
        for (int j = 0; j < A; ++j) {
            const float mu = p.mu[(int64_t)r * A + j], sg = p.std[j], a = p.b.actions[s * A + j];
            const float d = a - mu;
            const float gmu = gs * d / (sg * sg);
            amax_mu = fmaxf(amax_mu, fabsf(gmu));
            __half h, l;
            split1(gmu * s_mu, h, l);
            p.dmu[(int64_t)r * p.ld_dmu + j] = h;
            p.dmu[p.plane_dmu + (int64_t)r * p.ld_dmu + j] = l;
            dstd[j] = gs * (d * d - sg * sg) / (sg * sg * sg);
        }
        const float gval = p.hp.value_loss_coef * gv * p.inv_count;
        amax_v = fabsf(gval);
        __half h, l;
        split1(gval * s_v, h, l);
        p.dv[(int64_t)r * p.ld_dv] = h;
        p.dv[p.plane_dv + (int64_t)r * p.ld_dv] = l;
    } else {
        for (int j = 0; j < A; ++j) dstd[j] = 0.0f;
    }
    record_amax(p.amax_mu, amax_mu);
    record_amax(p.amax_v, amax_v);
    float* out = p.partial + (int64_t)blockIdx.x * STAT_SLOTS;
    float t = block_sum(surr, sh);
    if (threadIdx.x == 0) out[0] = t;
    t = block_sum(vl, sh);
    if (threadIdx.x == 0) out[1] = t;
    t = block_sum(kl, sh);
    if (threadIdx.x == 0) out[2] = t;
    for (int j = 0; j < A; ++j) {
        t = block_sum(dstd[j], sh);
        if (threadIdx.x == 0) out[3 + j] = t;
    }
}

// sums the per-block rows in block order; stats[0..3] and the std gradient (entropy term: d/dsigma of -c_e * mean(entropy) = -c_e / sigma)
__global__ void ppo_head_reduce_kernel(const float* __restrict__ partial, int blocks, int A, const float* __restrict__ std, float entropy_coef,
                                       float world_inv, int count, float* __restrict__ stats, float* __restrict__ gstd) {
    const int j = threadIdx.x;
    if (j >= 3 + A) return;
    float acc = 0.0f;
    for (int b = 0; b < blocks; ++b) acc += partial[(int64_t)b * STAT_SLOTS + j];
    if (j < 3) stats[j] = acc;
    else gstd[j - 3] = acc - entropy_coef * world_inv / std[j - 3];
    if (j == 0) stats[3] = (float)count;
}

struct AdaptHeadParams {
    const float* pred;          // [count][P]
    const float* priv;          // storage [*, P]
    const int64_t* indices;
    int count, P, num_train;
    float inv_train;            // 1 / (num_train * P * grad_world)
    const float* s_g;           // device: scale of the head gradient tensor
    uint32_t* amax;
    __half* dz;
    int64_t ld_dz, plane_dz;
    float* partial;
};

// F.mse_loss(pred[:num_train], target[:num_train]) and its gradient; test loss on the remaining rows (PPO:166-167)
__global__ void __launch_bounds__(HEAD_THREADS) adapt_head_kernel(const __grid_constant__ AdaptHeadParams p) {
    __shared__ float sh[HEAD_THREADS / 32];
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    float tr = 0.0f, te = 0.0f, amax = 0.0f;
    const float s_g = __ldg(p.s_g);
    if (r < p.count) {
        const int64_t s = p.indices ? p.indices[r] : (int64_t)r;
        for (int j = 0; j < p.P; ++j) {
            const float d = p.pred[(int64_t)r * p.P + j] - p.priv[s * p.P + j];
            const float g = r < p.num_train ? 2.0f * d * p.inv_train : 0.0f;
            if (r < p.num_train) tr += d * d; else te += d * d;
            amax = fmaxf(amax, fabsf(g));
            __half h, l;
            split1(g * s_g, h, l);
            p.dz[(int64_t)r * p.ld_dz + j] = h;
            p.dz[p.plane_dz + (int64_t)r * p.ld_dz + j] = l;
        }
    }
    record_amax(p.amax, amax);
    float* out = p.partial + (int64_t)blockIdx.x * STAT_SLOTS;
    float t = block_sum(tr, sh);
    if (threadIdx.x == 0) out[0] = t;
    t = block_sum(te, sh);
    if (threadIdx.x == 0) out[1] = t;
}
__global__ void adapt_head_reduce_kernel(const float* __restrict__ partial, int blocks, int count, int num_train, int P, float* __restrict__ stats) {
    if (threadIdx.x >= 2) return;
    float acc = 0.0f;
    for (int b = 0; b < blocks; ++b) acc += partial[(int64_t)b * STAT_SLOTS + threadIdx.x];
    stats[4 + 2 * threadIdx.x] = acc;
    stats[5 + 2 * threadIdx.x] = threadIdx.x == 0 ? (float)num_train * P : (float)(count - num_train) * P;
}

// Philox4x32-10 (Salmon et al. 2011) for the action sample when no noise is injected
__device__ __forceinline__ void philox4(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t (&o)[4]) {
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        const uint32_t h0 = __umulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
        const uint32_t h1 = __umulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = h1 ^ c1 ^ k0, n2 = h0 ^ c3 ^ k1;
        c0 = n0; c1 = l1; c2 = n2; c3 = l0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    o[0] = c0; o[1] = c1; o[2] = c2; o[3] = c3;
}

// PPO.act head (PPO:65-77, AC:131-138): sample, log-prob, sigma
__global__ void act_head_kernel(const float* __restrict__ mean, const float* __restrict__ std, const float* __restrict__ noise,
                                const float* __restrict__ actions_in, int rows, int A, uint64_t seed, uint64_t step, float* __restrict__ actions,
                                float* __restrict__ logp_out, float* __restrict__ mean_out, float* __restrict__ sigma_out) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    float logp = 0.0f;
    for (int j = 0; j < A; ++j) {
        const float mu = mean[(int64_t)r * A + j], sg = std[j];
        float a;
        if (actions_in) a = actions_in[(int64_t)r * A + j];
        else {
            float z;
            if (noise) z = noise[(int64_t)r * A + j];
            else {
                uint32_t o[4];
                philox4((uint32_t)r, (uint32_t)j, (uint32_t)step, (uint32_t)(step >> 32), (uint32_t)seed, (uint32_t)(seed >> 32), o);
                const float u1 = ((float)(o[0] >> 8) + 0.5f) * (1.0f / 16777216.0f), u2 = ((float)(o[1] >> 8) + 0.5f) * (1.0f / 16777216.0f);
                z = sqrtf(-2.0f * logf(u1)) * cospif(2.0f * u2);
            }
            a = mu + sg * z;
        }
        const float d = a - mu;
        logp += -(d * d) / (2.0f * sg * sg) - logf(sg) - 0.91893853320467274178f;
        if (actions) actions[(int64_t)r * A + j] = a;
        if (mean_out && mean_out != mean) mean_out[(int64_t)r * A + j] = mu;
        if (sigma_out) sigma_out[(int64_t)r * A + j] = sg;
    }
    if (logp_out) logp_out[r] = logp;
}

// sum of squares of a range of the gradient, per block (fixed order), then the finalisation below
__global__ void __launch_bounds__(256) sumsq_kernel(const float* __restrict__ g, int64_t n, float* __restrict__ partial) {
    __shared__ float sh[8];
    float acc = 0.0f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) acc += g[i] * g[i];
    const float t = block_sum(acc, sh);
    if (threadIdx.x == 0) partial[blockIdx.x] = t;
}
// total norm, clip coefficient (nn.utils.clip_grad_norm_) and the adaptive-KL learning-rate decision (PPO:111-125)
__global__ void prepare_step_kernel(const float* __restrict__ partial, int blocks, lt_ppo_hparams hp, const float* __restrict__ stats,
                                    float* __restrict__ lr, float* __restrict__ scalars, float* __restrict__ grad_norm_out) {
    if (threadIdx.x != 0) return;
    float ss = 0.0f;
    for (int b = 0; b < blocks; ++b) ss += partial[b];
    const float norm = sqrtf(ss);
    scalars[0] = fminf(hp.max_grad_norm / (norm + 1e-6f), 1.0f);
    scalars[1] = norm;
    if (grad_norm_out) *grad_norm_out = norm;
    if (hp.adaptive_schedule && stats) {
        const float kl_mean = stats[2] / stats[3];
        float v = *lr;
        if (kl_mean > hp.desired_kl * 2.0f) v = fmaxf(hp.lr_min, v / hp.lr_factor);
        else if (kl_mean < hp.desired_kl / 2.0f && kl_mean > 0.0f) v = fminf(hp.lr_max, v * hp.lr_factor);
        *lr = v;
    }
}
// torch.optim.Adam (no weight decay, no amsgrad) over params[lo, hi)
__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ params, const float* __restrict__ grads, float* __restrict__ m,
                                                   float* __restrict__ v, int64_t lo, int64_t hi, const float* __restrict__ lr_dev, float lr_host,
                                                   const float* __restrict__ clip, float beta1, float beta2, float eps, float bc1, float bc2_sqrt) {
    const int64_t i = lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= hi) return;
    const float lr = lr_dev ? *lr_dev : lr_host;
    const float g = grads[i] * (clip ? *clip : 1.0f);
    const float mi = beta1 * m[i] + (1.0f - beta1) * g;
    const float vi = beta2 * v[i] + (1.0f - beta2) * g * g;
    m[i] = mi;
    v[i] = vi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    params[i] -= (lr / bc1) * (mi / denom);
}

__global__ void split_f32_kernel(const float* __restrict__ src, int64_t rows, int cols, float scale, __half* dst, int64_t ld, int64_t plane) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows * cols) return;
    const int64_t r = i / cols;
    const int c = (int)(i % cols);
    __half hi, lo;
    split1(src[i] * scale, hi, lo);
    dst[r * ld + c] = hi;
    dst[plane + r * ld + c] = lo;
}

// ------------------------------------------------------------------------------------------------------------ host side
struct Operand {
    const __half* base;
    int64_t rows, ld, plane;
};
inline Operand op(const Planes& pl, int64_t rows) { return Operand{pl.p, rows, pl.ld, pl.plane}; }

// D[M, N] = A[M, K] * B[N, K]^T with the epilogue fields of `p` already set
int run_gemm(lt_policy* P, const Operand& A, int M, const Operand& B, int N, int K, int splits, ltg::Params& p, cudaStream_t stream) {
    if (M <= 0 || N <= 0) return LT_OK;
    p.M = M;
    p.N = N;
    p.K = K;
    p.bn = pick_bn(N);
    const int kb_total = (K + ltg::BK - 1) / ltg::BK;
    splits = splits < 1 ? 1 : (splits > kb_total ? kb_total : splits);
    p.kb_per_split = (kb_total + splits - 1) / splits;
    p.splits = (kb_total + p.kb_per_split - 1) / p.kb_per_split;
    if (!ltg::make_operand_map(&p.tma_a, A.base, M, K, A.ld, A.plane, ltg::BM) ||
        !ltg::make_operand_map(&p.tma_b, B.base, N, K, B.ld, B.plane, p.bn))
        return fail(P, LT_ERR_CUDA, "cuTensorMapEncodeTiled failed (driver entry point missing or bad operand geometry)");
    p.use_tma_out = 0;
    if ((p.mode == ltg::EPI_ELU || p.mode == ltg::EPI_BWD) && p.out_h && (p.ld_h & 7) == 0 && (p.plane_h & 7) == 0 &&
        (reinterpret_cast<uintptr_t>(p.out_h) & 15) == 0 && !getenv("LT_GEMM_NO_TMA_STORE"))
        p.use_tma_out = ltg::make_out_map(&p.tma_out, p.out_h, M, N, p.ld_h, p.plane_h) ? 1 : 0;
    LTP_CUDA(P, ltg::launch(p, P ? P->sm_count : 148, stream));
    return LT_OK;
}

// D[M, N] = A[K, M]^T * B[K, N]: both operands row-major [K rows][cols] planes read MN-major (contraction over rows)
int run_gemm_mn(lt_policy* P, const Operand& A, int M, const Operand& B, int N, int K, int splits, ltg::Params& p, cudaStream_t stream) {
    if (M <= 0 || N <= 0) return LT_OK;
    p.M = M;
    p.N = N;
    p.K = K;
    p.mn_major = 1;
    p.bn = pick_bn(N);
    const int kb_total = (K + ltg::BK - 1) / ltg::BK;
    splits = splits < 1 ? 1 : (splits > kb_total ? kb_total : splits);
    p.kb_per_split = (kb_total + splits - 1) / splits;
    p.splits = (kb_total + p.kb_per_split - 1) / p.kb_per_split;
    if (!ltg::make_operand_map_mn(&p.tma_a, A.base, K, M, A.ld, A.plane) || !ltg::make_operand_map_mn(&p.tma_b, B.base, K, N, B.ld, B.plane))
        return fail(P, LT_ERR_CUDA, "cuTensorMapEncodeTiled failed (driver entry point missing or bad operand geometry)");
    LTP_CUDA(P, ltg::launch(p, P ? P->sm_count : 148, stream));
    return LT_OK;
}

// `X`: input planes of layer 0 (GEMM width K_in, columns mapped onto the k_real0 parameter columns by map0);
// final_act != NULL: the output layer is activated too and written, split, into those planes instead of m.out
int build_mlp(lt_policy* P, Mlp& m, int n_hidden, const int32_t* hidden, int n_out, int k_real0, const Planes* X, int K_in, const SegMap& map0,
              const Planes* final_act, int64_t& off) {
    m.n = n_hidden + 1;
    m.X = X;
    m.final_act = final_act;
    int prev = k_real0;
    for (int l = 0; l < m.n; ++l) {
        Layer& L = m.L[l];
        L.N = l < n_hidden ? hidden[l] : n_out;
        L.k_real = prev;
        if (l == 0) {
            L.K = K_in;
            L.map = map0;
        } else {
            L.K = prev;
            L.map = SegMap();
            L.map.add(0, 0, prev);
        }
        L.w_off = off;
        off += (int64_t)L.N * L.k_real;
        L.b_off = off;
        off += L.N;
        int rc = alloc_planes(P, L.W, L.N, L.K);
        if (!rc) rc = alloc_dev(P, L.bias16, up(L.N, 4));
        if (rc) return rc;
        if (P->c.training) {
            rc = alloc_planes(P, L.WT, L.K, L.N);
            if (rc) return rc;
        }
        if (l < n_hidden) {
            rc = alloc_planes(P, m.Y[l], P->R, L.N + 1);
            if (rc) return rc;
            ones_col_kernel<<<(unsigned)((P->R + 255) / 256), 256>>>(m.Y[l].p + L.N, P->R, m.Y[l].ld);
        }
        if (P->c.training) {
            rc = alloc_planes(P, m.dZ[l], P->R, up(L.N, 16));
            if (rc) return rc;
        }
        prev = L.N;
    }
    return final_act ? LT_OK : alloc_dev(P, m.out, P->R * m.L[m.n - 1].N);
}

int refresh_mlp(lt_policy* P, Mlp& m, cudaStream_t s) {
    for (int l = 0; l < m.n; ++l) {
        Layer& L = m.L[l];
        const int64_t n = (int64_t)L.N * L.K;
        split_weights_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(P->params + L.w_off, L.N, L.K, L.k_real, L.map, L.W.p, L.W.ld,
                                                                        L.W.plane, L.WT.p, L.WT.ld, L.WT.plane, P->params + L.b_off, L.bias16);
    }
    LTP_CUDA(P, cudaGetLastError());
    return LT_OK;
}

// forward of one MLP on `rows` rows of X0; the last layer's fp32 output goes to m.out
int forward_mlp(lt_policy* P, Mlp& m, int rows, bool train, cudaStream_t s) {
    for (int l = 0; l < m.n; ++l) {
        Layer& L = m.L[l];
        ltg::Params p;
        memset(&p, 0, sizeof p);
        p.out_scale = 1.0f / (S_X * S_W);
        p.bias = L.bias16;
        const Operand A = l == 0 ? op(*m.X, rows) : op(m.Y[l - 1], rows);
        if (l < m.n - 1 || m.final_act) {
            const Planes& Y = l < m.n - 1 ? m.Y[l] : *m.final_act;
            p.mode = ltg::EPI_ELU;
            p.act = P->c.activation;
            p.out_h = Y.p;
            p.ld_h = Y.ld;
            p.plane_h = Y.plane;
            p.split_scale = S_X;
        } else {
            p.mode = ltg::EPI_F32;
            p.out_f32 = m.out;
            p.ld_f32 = L.N;
        }
        const int rc = run_gemm(P, A, rows, op(L.W, L.N), L.N, L.K, 1, p, s);
        if (rc) return rc;
    }
    return LT_OK;
}

int dw_splits(lt_policy* P, int N, int Kp1, int count) {
    const int tiles = ((N + ltg::BM - 1) / ltg::BM) * ((Kp1 + pick_bn(Kp1) - 1) / pick_bn(Kp1));
    const int kb_total = (count + ltg::BK - 1) / ltg::BK;
    int s = (2 * P->sm_count + tiles - 1) / tiles;
    if (s > kb_total) s = kb_total;
    return s < 1 ? 1 : s;
}

// backward of one MLP: m.dZ[n-1] holds the (scaled) gradient of the last layer's output
int backward_mlp(lt_policy* P, Mlp& m, int count, int ctx, cudaStream_t s) {
    for (int l = m.n - 1; l >= 0; --l) {
        Layer& L = m.L[l];
        {   // dW_l and db_l:  [N_l, K_l + 1] = dZ_l[count, N_l]^T * [X_{l-1} | 1][count, K_l + 1]
            const Planes& X = l == 0 ? *m.X : m.Y[l - 1];
            const int Kp1 = L.K + 1;
            ltg::Params p;
            memset(&p, 0, sizeof p);
            p.mode = ltg::EPI_PARTIAL;
            p.out_scale = 1.0f / S_X;
            p.a_scale_ptr = P->gs(ctx, m, l);
            p.out_f32 = P->partial;
            p.ld_f32 = up(Kp1, 4);
            p.split_stride = (int64_t)L.N * p.ld_f32;
            const int splits = dw_splits(P, L.N, Kp1, count);
            if ((int64_t)splits * p.split_stride > P->partial_floats) return fail(P, LT_ERR_INVALID, "split-K workspace too small");
            int rc = run_gemm_mn(P, op(m.dZ[l], count), L.N, op(X, count), Kp1, count, splits, p, s);
            if (rc) return rc;
            const int64_t n = (int64_t)L.N * Kp1;
            reduce_dw_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(P->partial, p.splits, p.split_stride, p.ld_f32, L.N, L.K, L.k_real, L.map,
                                                                        P->grads + L.w_off, P->grads + L.b_off);
        }
        if (l > 0) {   // dZ_{l-1} = (dZ_l * W_l) .* ELU'(Y_{l-1})
            ltg::Params p;
            memset(&p, 0, sizeof p);
            p.mode = ltg::EPI_BWD;
            p.act = P->c.activation;
            p.out_scale = 1.0f / S_W;
            p.a_scale_ptr = P->gs(ctx, m, l);
            p.split_scale_ptr = P->gs(ctx, m, l - 1);
            p.amax_out = P->ga(ctx, m, l - 1);
            p.out_h = m.dZ[l - 1].p;
            p.ld_h = m.dZ[l - 1].ld;
            p.plane_h = m.dZ[l - 1].plane;
            p.act_h = m.Y[l - 1].p;
            p.ld_act = m.Y[l - 1].ld;
            p.plane_act = m.Y[l - 1].plane;
            p.act_inv_scale = 1.0f / S_X;
            const int rc = run_gemm(P, op(m.dZ[l], count), count, op(L.WT, L.K), L.K, L.N, 1, p, s);
            if (rc) return rc;
        }
    }
    LTP_CUDA(P, cudaGetLastError());
    return LT_OK;
}

int pack(lt_policy* P, const float* obs_history, const float* priv, const int64_t* indices, int rows, int teacher, bool train, cudaStream_t s) {
    (void)train;
    if (2 * P->P > 32) return fail(P, LT_ERR_INVALID, "num_privileged_obs > 16 is not supported by the gather kernel");
    if (P->has_enc)
        pack_enc_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, s>>>(obs_history, priv, indices, rows, P->Dhist, P->last_off, P->E0, P->HM, P->E, P->P,
                                                                   teacher, P->X0.p, P->X0.ld, P->X0.plane, P->XH.p, P->XH.ld, P->XH.plane);
    else
        pack_x0_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, s>>>(obs_history, priv, indices, rows, P->Din, P->P, teacher, P->X0.p, P->X0.ld, P->X0.plane);
    LTP_CUDA(P, cudaGetLastError());
    P->packed_src = nullptr;
    return LT_OK;
}
int scatter_latent(lt_policy* P, int rows, bool train, cudaStream_t s) {
    (void)train;
    scatter_latent_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, s>>>(P->adapt.out, rows, P->P, P->Din, P->X0.p, P->X0.ld, P->X0.plane);
    LTP_CUDA(P, cudaGetLastError());
    return LT_OK;
}

// encoder forward (CNN:181-187): XH -> emb columns of X0.  No-op without an encoder.
int forward_enc(lt_policy* P, int rows, bool train, cudaStream_t s) {
    return P->has_enc ? forward_mlp(P, P->enc, rows, train, s) : LT_OK;
}

// encoder backward: the emb columns of the first-layer input gradients of `nets` (dZ_0 * W_0[:, emb]) summed in fp32, times the
// activation derivative -> the encoder's last dZ; then the encoder's own weight gradients.  No-op without an encoder.
int backward_enc(lt_policy* P, Mlp* const* nets, int n_nets, int count, int ctx, cudaStream_t s) {
    if (!P->has_enc) return LT_OK;
    Mlp& e = P->enc;
    const int64_t stride = P->R * (int64_t)P->E;
    for (int i = 0; i < n_nets; ++i) {
        Mlp& m = *nets[i];
        Layer& L = m.L[0];
        ltg::Params g;
        memset(&g, 0, sizeof g);
        g.mode = ltg::EPI_PARTIAL;
        g.out_scale = 1.0f / S_W;
        g.a_scale_ptr = P->gs(ctx, m, 0);
        g.out_f32 = P->enc_part + i * stride;
        g.ld_f32 = P->E;
        g.split_stride = 0;
        const Operand B{L.WT.p, P->E, L.WT.ld, L.WT.plane};             // rows [0, E) of W_0^T: the emb columns
        const int rc = run_gemm(P, op(m.dZ[0], count), count, B, P->E, L.N, 1, g, s);
        if (rc) return rc;
    }
    Planes& dz = e.dZ[e.n - 1];
    const int64_t n = (int64_t)count * (P->E / 4);
    enc_grad_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(P->enc_part, n_nets, stride, count, P->E, P->c.activation, P->X0.p, P->X0.ld, P->X0.plane,
                                                               P->gs(ctx, e, e.n - 1), P->ga(ctx, e, e.n - 1), dz.p, dz.ld, dz.plane);
    LTP_CUDA(P, cudaGetLastError());
    return backward_mlp(P, e, count, ctx, s);
}

int check_ready(lt_policy* P, int rows) {
    if (!P) return LT_ERR_INVALID;
    if (!P->params) return fail(P, LT_ERR_UNBOUND, "lt_policy_bind_params was never called");
    if (rows <= 0 || rows > P->R) return fail(P, LT_ERR_INVALID, "rows outside (0, max_rows]");
    return LT_OK;
}

float grad_scale_for(int count) {
    int e = 0;
    while ((1ll << e) < count) ++e;
    return ldexpf(1.0f, e);
}

// scales of call kind `ctx` for this call: measured by the previous call, or (first call) the row-count default
int begin_grad_scales(lt_policy* P, int ctx, int count_world, cudaStream_t s) {
    if (P->calibrated[ctx]) return 0;
    fill_kernel<<<1, 32, 0, s>>>(P->gscale + ctx * GS_N, grad_scale_for(count_world), GS_N);
    LTP_CUDA(P, cudaMemsetAsync(P->gamax + ctx * GS_N, 0, GS_N * sizeof(uint32_t), s));
    return 1;                                                   // caller runs its backward pass once more after end_grad_scales
}
int end_grad_scales(lt_policy* P, int ctx, cudaStream_t s) {
    update_scales_kernel<<<1, 32, 0, s>>>(P->gamax + ctx * GS_N, P->gscale + ctx * GS_N, GS_N);
    LTP_CUDA(P, cudaGetLastError());
    P->calibrated[ctx] = true;
    return LT_OK;
}

}   // namespace

// ---------------------------------------------------------------------------------------------------------------- static scratch
// Scratch of lt_linear_f32_static (include/lt_policy.h): the fp16 operand planes of ONE call, in a fixed array instead of
// cudaMalloc'd buffers.  Every GPU that loads the library holds these LT_LINEAR_STATIC_SCRATCH_BYTES until the process exits.
// In HOST code the name below is only a stand-in for the symbol, not a device address: the address handed to kernels and TMA
// maps comes from cudaGetSymbolAddress, looked up once per device with that device current (static_scratch_ptr).
static __device__ __align__(256) unsigned char g_linear_static_scratch[LT_LINEAR_STATIC_SCRATCH_BYTES];

namespace {
constexpr int MAX_DEVICES = 64;
unsigned char* g_static_scratch_ptr[MAX_DEVICES] = {nullptr};

// device address of g_linear_static_scratch on the CURRENT device (cached per device ordinal); NULL + message on failure
unsigned char* static_scratch_ptr(int dev) {
    if (dev < 0 || dev >= MAX_DEVICES) {
        set_error(nullptr, "lt_linear_f32_static: device ordinal outside the per-device address table");
        return nullptr;
    }
    if (!g_static_scratch_ptr[dev]) {
        void* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&ptr, g_linear_static_scratch);
        if (e != cudaSuccess || !ptr) {
            set_error(nullptr, std::string("cudaGetSymbolAddress(g_linear_static_scratch): ") + cudaGetErrorString(e));
            return nullptr;
        }
        g_static_scratch_ptr[dev] = static_cast<unsigned char*>(ptr);
    }
    return g_static_scratch_ptr[dev];
}
}   // namespace

// ================================================================================================================ C ABI
extern "C" {

int lt_policy_config_bytes(void) { return (int)sizeof(lt_policy_config); }

const char* lt_policy_last_error(const lt_policy* p) { return p ? p->err.c_str() : g_policy_error.c_str(); }

int lt_policy_create(const lt_policy_config* cfg, lt_policy** out) {
    if (!cfg || !out) return LT_ERR_INVALID;
    *out = nullptr;
    if (cfg->struct_bytes != (int)sizeof(lt_policy_config)) return fail(nullptr, LT_ERR_INVALID, "lt_policy_config size mismatch: rebuild");
    int dev = 0;
    cudaDeviceProp prop;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess || prop.major != 10)
        return fail(nullptr, LT_ERR_NO_DEVICE, "no sm_100 device: the actor-critic kernels have no CPU fallback");
    if (cfg->num_obs_history <= 0 || cfg->num_privileged_obs <= 0 || cfg->num_actions <= 0 || cfg->num_actions > STAT_SLOTS - 3 ||
        cfg->num_actions > 32 || GS_N > 32 || cfg->max_rows <= 0 || cfg->n_adaptation_hidden < 1 || cfg->n_actor_hidden < 1 || cfg->n_critic_hidden < 1 ||
        cfg->n_adaptation_hidden > LT_POLICY_MAX_HIDDEN || cfg->n_actor_hidden > LT_POLICY_MAX_HIDDEN || cfg->n_critic_hidden > LT_POLICY_MAX_HIDDEN ||
        cfg->activation < LT_ACT_ELU || cfg->activation > LT_ACT_SIGMOID)
        return fail(nullptr, LT_ERR_INVALID, "unsupported actor-critic geometry");
    if (cfg->height_map_size < 0 ||
        (cfg->height_map_size > 0 &&
         (cfg->num_obs <= 0 || cfg->num_obs_history % cfg->num_obs != 0 || cfg->height_map_size > cfg->num_obs || cfg->encoder_hidden <= 0 ||
          cfg->encoder_embedding <= 0 || cfg->encoder_embedding % 8 != 0)))
        return fail(nullptr, LT_ERR_INVALID,
                    "unsupported height-map encoder geometry (height_map_size <= num_obs, num_obs divides num_obs_history, "
                    "encoder_embedding a multiple of 8)");
    if (!ltg::encode_fn()) return fail(nullptr, LT_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
    lt_policy* P = new lt_policy();
    P->c = *cfg;
    P->sm_count = prop.multiProcessorCount;
    P->R = up(cfg->max_rows, 128);
    P->Dhist = cfg->num_obs_history;
    P->P = cfg->num_privileged_obs;
    P->A = cfg->num_actions;
    P->has_enc = cfg->height_map_size > 0;
    if (P->has_enc) {
        P->HM = cfg->height_map_size;
        P->E = cfg->encoder_embedding;
        P->E0 = cfg->num_obs - cfg->height_map_size;
        P->last_off = cfg->num_obs_history - cfg->num_obs;
        P->Din = P->E + 2 * P->E0;
    } else {
        P->Din = P->Dhist;
    }
    P->K0 = P->Din + 2 * P->P;
    P->adapt.idx = 0;
    P->actor.idx = 1;
    P->critic.idx = 2;
    P->enc.idx = 3;
    int rc = alloc_planes(P, P->X0, P->R, P->K0 + 1);
    if (!rc) ones_col_kernel<<<(unsigned)((P->R + 255) / 256), 256>>>(P->X0.p + P->K0, P->R, P->X0.ld);
    int64_t off = P->A;                                         // std first
    // first-layer column maps: parameter columns of the reference's input (obs_history | latent or privileged; with the encoder
    // [s | s | emb] in place of obs_history) -> columns of X0
    SegMap base, m_adapt, m_actor, m_critic;
    if (P->has_enc) {
        base.add(0, P->E, 2 * P->E0);
        base.add(2 * P->E0, 0, P->E);
        SegMap hmap;
        hmap.add(0, 0, P->HM);
        const int32_t enc_hidden[1] = {cfg->encoder_hidden};
        rc = rc ? rc : alloc_planes(P, P->XH, P->R, P->HM + 1);
        if (!rc) ones_col_kernel<<<(unsigned)((P->R + 255) / 256), 256>>>(P->XH.p + P->HM, P->R, P->XH.ld);
        if (!rc) rc = build_mlp(P, P->enc, 1, enc_hidden, P->E, P->HM, &P->XH, P->HM, hmap, &P->X0, off);
        if (!rc && cfg->training) rc = alloc_dev(P, P->enc_part, 3 * P->R * (int64_t)P->E);
    } else {
        base.add(0, 0, P->Din);
    }
    m_adapt = base;
    m_actor = base;
    m_actor.add(P->Din, P->Din, P->P);                          // latent columns
    m_critic = base;
    m_critic.add(P->Din, P->Din + P->P, P->P);                  // privileged columns
    if (!rc) rc = build_mlp(P, P->adapt, cfg->n_adaptation_hidden, cfg->adaptation_hidden, P->P, P->Din, &P->X0, P->K0, m_adapt, nullptr, off);
    if (!rc) rc = build_mlp(P, P->actor, cfg->n_actor_hidden, cfg->actor_hidden, P->A, P->Din + P->P, &P->X0, P->K0, m_actor, nullptr, off);
    if (!rc) rc = build_mlp(P, P->critic, cfg->n_critic_hidden, cfg->critic_hidden, 1, P->Din + P->P, &P->X0, P->K0, m_critic, nullptr, off);
    P->n_params = off;
    if (!rc && cfg->training) {
        int64_t need = 0;
        Mlp* ms[4] = {&P->adapt, &P->actor, &P->critic, &P->enc};
        for (Mlp* m : ms)
            for (int l = 0; l < m->n; ++l) {
                const int Kp1 = m->L[l].K + 1;
                const int64_t f = (int64_t)dw_splits(P, m->L[l].N, Kp1, (int)P->R) * m->L[l].N * up(Kp1, 4);
                if (f > need) need = f;
            }
        P->partial_floats = need;
        rc = alloc_dev(P, P->partial, need);
    }
    const int64_t head_blocks = (P->R + HEAD_THREADS - 1) / HEAD_THREADS;
    if (!rc) rc = alloc_dev(P, P->head_partial, head_blocks * STAT_SLOTS);
    if (!rc) rc = alloc_dev(P, P->norm_partial, 1024);
    if (!rc) rc = alloc_dev(P, P->scalars, 8);
    if (!rc) rc = alloc_dev(P, P->gscale, 2 * GS_N);
    if (!rc) rc = alloc_dev(P, P->gamax, 2 * GS_N);
    if (!rc && cudaDeviceSynchronize() != cudaSuccess) rc = fail(P, LT_ERR_CUDA, "initialisation kernels failed");
    if (rc) {
        g_policy_error = P->err;
        lt_policy_destroy(P);
        return rc;
    }
    *out = P;
    return LT_OK;
}

void lt_policy_destroy(lt_policy* p) {
    if (!p) return;
    for (void* d : p->allocs) cudaFree(d);
    delete p;
}

int64_t lt_policy_num_params(const lt_policy* p) { return p ? p->n_params : 0; }

int lt_policy_adaptation_slice(const lt_policy* p, int64_t* lo, int64_t* hi) {
    if (!p || !lo || !hi) return LT_ERR_INVALID;
    *lo = p->has_enc ? p->enc.L[0].w_off : p->adapt.L[0].w_off;
    *hi = p->actor.L[0].w_off;
    return LT_OK;
}

const float* lt_policy_grad_scales(const lt_policy* p) { return p ? p->gscale : nullptr; }

int lt_policy_bind_params(lt_policy* p, float* params, float* grads) {
    if (!p || !params) return LT_ERR_INVALID;
    if (p->c.training && !grads) return fail(p, LT_ERR_INVALID, "a training policy needs a gradient buffer");
    p->params = params;
    p->grads = grads;
    return LT_OK;
}

int lt_policy_refresh(lt_policy* p, void* stream) {
    if (!p) return LT_ERR_INVALID;
    if (!p->params) return fail(p, LT_ERR_UNBOUND, "lt_policy_bind_params was never called");
    int rc = refresh_mlp(p, p->adapt, (cudaStream_t)stream);
    if (!rc && p->has_enc) rc = refresh_mlp(p, p->enc, (cudaStream_t)stream);
    if (!rc) rc = refresh_mlp(p, p->actor, (cudaStream_t)stream);
    if (!rc) rc = refresh_mlp(p, p->critic, (cudaStream_t)stream);
    return rc;
}

int lt_policy_forward(lt_policy* p, const float* obs_history, const float* privileged_obs, int32_t rows, int32_t teacher_latent, float* latent,
                      float* action_mean, float* value, void* stream) {
    int rc = check_ready(p, rows);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (!obs_history) return fail(p, LT_ERR_INVALID, "obs_history is NULL");
    if ((value || teacher_latent) && !privileged_obs) return fail(p, LT_ERR_INVALID, "privileged_obs is NULL but the critic / teacher needs it");
    rc = pack(p, obs_history, privileged_obs, nullptr, rows, teacher_latent, false, s);
    if (!rc) rc = forward_enc(p, rows, false, s);
    if (rc) return rc;
    if (latent || (action_mean && !teacher_latent)) {
        rc = forward_mlp(p, p->adapt, rows, false, s);
        if (rc) return rc;
        if (latent) LTP_CUDA(p, cudaMemcpyAsync(latent, p->adapt.out, (size_t)rows * p->P * sizeof(float), cudaMemcpyDeviceToDevice, s));
        if (action_mean && !teacher_latent) {
            rc = scatter_latent(p, rows, false, s);
            if (rc) return rc;
        }
    }
    if (action_mean) {
        rc = forward_mlp(p, p->actor, rows, false, s);
        if (rc) return rc;
        LTP_CUDA(p, cudaMemcpyAsync(action_mean, p->actor.out, (size_t)rows * p->A * sizeof(float), cudaMemcpyDeviceToDevice, s));
    }
    if (value) {
        rc = forward_mlp(p, p->critic, rows, false, s);
        if (rc) return rc;
        LTP_CUDA(p, cudaMemcpyAsync(value, p->critic.out, (size_t)rows * sizeof(float), cudaMemcpyDeviceToDevice, s));
    }
    return LT_OK;
}

int lt_policy_act(lt_policy* p, const float* obs_history, const float* privileged_obs, int32_t rows, const float* normal_noise,
                  const float* actions_in, uint64_t seed, uint64_t step, float* actions, float* actions_log_prob, float* action_mean,
                  float* action_sigma, float* value, void* stream) {
    int rc = check_ready(p, rows);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (!obs_history) return fail(p, LT_ERR_INVALID, "obs_history is NULL");
    if (value && !privileged_obs) return fail(p, LT_ERR_INVALID, "privileged_obs is NULL but the critic needs it");
    rc = pack(p, obs_history, privileged_obs, nullptr, rows, 0, false, s);
    if (!rc) rc = forward_enc(p, rows, false, s);
    if (!rc) rc = forward_mlp(p, p->adapt, rows, false, s);
    if (!rc) rc = scatter_latent(p, rows, false, s);
    if (!rc) rc = forward_mlp(p, p->actor, rows, false, s);
    if (!rc && value) rc = forward_mlp(p, p->critic, rows, false, s);
    if (rc) return rc;
    act_head_kernel<<<(unsigned)((rows + 255) / 256), 256, 0, s>>>(p->actor.out, p->params, normal_noise, actions_in, rows, p->A, seed, step, actions,
                                                                  actions_log_prob, action_mean, action_sigma);
    LTP_CUDA(p, cudaGetLastError());
    if (value) LTP_CUDA(p, cudaMemcpyAsync(value, p->critic.out, (size_t)rows * sizeof(float), cudaMemcpyDeviceToDevice, s));
    return LT_OK;
}

int lt_ppo_minibatch_grad(lt_policy* p, const lt_ppo_batch_ptrs* batch, const int64_t* indices, int32_t count, const lt_ppo_hparams* hp,
                          int32_t grad_world, float* stats, void* stream) {
    int rc = check_ready(p, count);
    if (rc) return rc;
    if (!p->c.training) return fail(p, LT_ERR_INVALID, "policy was created with training = 0");
    if (!batch || !hp || !stats) return LT_ERR_INVALID;
    cudaStream_t s = (cudaStream_t)stream;
    const int world = grad_world > 1 ? grad_world : 1;
    rc = pack(p, batch->obs_history, batch->privileged_obs, indices, count, 0, true, s);
    if (!rc) {
        p->packed_src = batch->obs_history;
        p->packed_idx = indices;
        p->packed_count = count;
    }
    if (!rc) rc = forward_enc(p, count, true, s);
    if (!rc) rc = forward_mlp(p, p->adapt, count, true, s);
    if (!rc) rc = scatter_latent(p, count, true, s);
    if (!rc) rc = forward_mlp(p, p->actor, count, true, s);
    if (!rc) rc = forward_mlp(p, p->critic, count, true, s);
    if (rc) return rc;
    const int blocks = (count + HEAD_THREADS - 1) / HEAD_THREADS;
    const int passes = 1 + begin_grad_scales(p, CTX_PPO, count * world, s);
    for (int pass = 0; pass < passes; ++pass) {
        Mlp &a = p->actor, &c = p->critic, &ad = p->adapt;
        HeadParams h;
        memset(&h, 0, sizeof h);
        h.b = *batch;
        h.indices = indices;
        h.count = count;
        h.A = p->A;
        h.mu = a.out;
        h.value = c.out;
        h.std = p->params;
        h.hp = *hp;
        h.inv_count = 1.0f / ((float)count * (float)world);
        h.s_mu = p->gs(CTX_PPO, a, a.n - 1);
        h.s_v = p->gs(CTX_PPO, c, c.n - 1);
        h.amax_mu = p->ga(CTX_PPO, a, a.n - 1);
        h.amax_v = p->ga(CTX_PPO, c, c.n - 1);
        h.dmu = a.dZ[a.n - 1].p; h.ld_dmu = a.dZ[a.n - 1].ld; h.plane_dmu = a.dZ[a.n - 1].plane;
        h.dv = c.dZ[c.n - 1].p; h.ld_dv = c.dZ[c.n - 1].ld; h.plane_dv = c.dZ[c.n - 1].plane;
        h.partial = p->head_partial;
        ppo_head_kernel<<<blocks, HEAD_THREADS, 0, s>>>(h);
        ppo_head_reduce_kernel<<<1, 64, 0, s>>>(p->head_partial, blocks, p->A, p->params, hp->entropy_coef, 1.0f / (float)world, count, stats,
                                                p->grads);
        LTP_CUDA(p, cudaGetLastError());
        rc = backward_mlp(p, c, count, CTX_PPO, s);
        if (!rc) rc = backward_mlp(p, a, count, CTX_PPO, s);
        if (rc) return rc;
        {   // the latent columns of the actor's input gradient are the adaptation module's output gradient (AC:129-130)
            Layer& L = a.L[0];
            ltg::Params g;
            memset(&g, 0, sizeof g);
            g.mode = ltg::EPI_BWD;                                   // act_h == NULL: no activation derivative (last layer is linear)
            g.out_scale = 1.0f / S_W;
            g.a_scale_ptr = p->gs(CTX_PPO, a, 0);
            g.split_scale_ptr = p->gs(CTX_PPO, ad, ad.n - 1);
            g.amax_out = p->ga(CTX_PPO, ad, ad.n - 1);
            Planes& dz = ad.dZ[ad.n - 1];
            g.out_h = dz.p; g.ld_h = dz.ld; g.plane_h = dz.plane;
            Operand B{L.WT.p + (int64_t)p->Din * L.WT.ld, p->P, L.WT.ld, L.WT.plane};
            rc = run_gemm(p, op(a.dZ[0], count), count, B, p->P, L.N, 1, g, s);
            if (!rc) rc = backward_mlp(p, ad, count, CTX_PPO, s);
        }
        if (!rc) {
            Mlp* nets[3] = {&a, &c, &ad};
            rc = backward_enc(p, nets, 3, count, CTX_PPO, s);
        }
        if (!rc) rc = end_grad_scales(p, CTX_PPO, s);
        if (rc) return rc;
    }
    return LT_OK;
}

int lt_ppo_minibatch_apply(lt_policy* p, const lt_ppo_hparams* hp, const float* stats, float* lr, float* adam_m, float* adam_v, int32_t adam_step,
                           float* grad_norm_out, void* stream) {
    if (!p || !hp || !lr || !adam_m || !adam_v || adam_step < 1) return LT_ERR_INVALID;
    if (!p->params || !p->grads) return fail(p, LT_ERR_UNBOUND, "lt_policy_bind_params was never called");
    cudaStream_t s = (cudaStream_t)stream;
    const int nb = 512;
    sumsq_kernel<<<nb, 256, 0, s>>>(p->grads, p->n_params, p->norm_partial);
    prepare_step_kernel<<<1, 32, 0, s>>>(p->norm_partial, nb, *hp, stats, lr, p->scalars, grad_norm_out);
    const float bc1 = 1.0f - powf(hp->adam_beta1, (float)adam_step);
    const float bc2 = sqrtf(1.0f - powf(hp->adam_beta2, (float)adam_step));
    adam_kernel<<<(unsigned)((p->n_params + 255) / 256), 256, 0, s>>>(p->params, p->grads, adam_m, adam_v, 0, p->n_params, lr, 0.0f, p->scalars,
                                                                     hp->adam_beta1, hp->adam_beta2, hp->adam_eps, bc1, bc2);
    LTP_CUDA(p, cudaGetLastError());
    return lt_policy_refresh(p, stream);
}

int lt_adaptation_grad(lt_policy* p, const lt_ppo_batch_ptrs* batch, const int64_t* indices, int32_t count, int32_t num_train, int32_t grad_world,
                       float* stats, void* stream) {
    int rc = check_ready(p, count);
    if (rc) return rc;
    if (!p->c.training) return fail(p, LT_ERR_INVALID, "policy was created with training = 0");
    if (!batch || !stats || num_train < 0 || num_train > count) return LT_ERR_INVALID;
    cudaStream_t s = (cudaStream_t)stream;
    const int world = grad_world > 1 ? grad_world : 1;
    // PPO:150-176 runs on the minibatch lt_ppo_minibatch_grad just gathered: its split copy is still in X0 (the adaptation
    // module reads only the obs_history columns; its first-layer weights are zero on the latent / privileged columns)
    const bool reuse = p->packed_src == batch->obs_history && p->packed_idx == indices && p->packed_count == count && indices != nullptr;
    if (!reuse) rc = pack(p, batch->obs_history, batch->privileged_obs, indices, count, 0, true, s);
    if (!rc) rc = forward_enc(p, count, true, s);                // the encoder's weights moved since the PPO half (CNN PPO:170)
    if (!rc) rc = forward_mlp(p, p->adapt, count, true, s);
    if (rc) return rc;
    const int blocks = (count + HEAD_THREADS - 1) / HEAD_THREADS;
    Mlp& ad = p->adapt;
    const int passes = 1 + begin_grad_scales(p, CTX_ADAPT, num_train > 0 ? num_train * p->P * world : 1, s);
    for (int pass = 0; pass < passes; ++pass) {
        AdaptHeadParams h;
        memset(&h, 0, sizeof h);
        h.pred = ad.out;
        h.priv = batch->privileged_obs;
        h.indices = indices;
        h.count = count;
        h.P = p->P;
        h.num_train = num_train;
        h.inv_train = num_train > 0 ? 1.0f / ((float)num_train * (float)p->P * (float)world) : 0.0f;
        h.s_g = p->gs(CTX_ADAPT, ad, ad.n - 1);
        h.amax = p->ga(CTX_ADAPT, ad, ad.n - 1);
        h.dz = ad.dZ[ad.n - 1].p; h.ld_dz = ad.dZ[ad.n - 1].ld; h.plane_dz = ad.dZ[ad.n - 1].plane;
        h.partial = p->head_partial;
        adapt_head_kernel<<<blocks, HEAD_THREADS, 0, s>>>(h);
        adapt_head_reduce_kernel<<<1, 32, 0, s>>>(p->head_partial, blocks, count, num_train, p->P, stats);
        LTP_CUDA(p, cudaGetLastError());
        rc = backward_mlp(p, ad, count, CTX_ADAPT, s);
        if (!rc) {
            Mlp* nets[1] = {&ad};
            rc = backward_enc(p, nets, 1, count, CTX_ADAPT, s);
        }
        if (!rc) rc = end_grad_scales(p, CTX_ADAPT, s);
        if (rc) return rc;
    }
    return LT_OK;
}

int lt_adaptation_apply(lt_policy* p, const lt_ppo_hparams* hp, float lr, float* adam_m, float* adam_v, int32_t adam_step, void* stream) {
    if (!p || !hp || !adam_m || !adam_v || adam_step < 1) return LT_ERR_INVALID;
    if (!p->params || !p->grads) return fail(p, LT_ERR_UNBOUND, "lt_policy_bind_params was never called");
    cudaStream_t s = (cudaStream_t)stream;
    int64_t lo, hi;
    lt_policy_adaptation_slice(p, &lo, &hi);
    const float bc1 = 1.0f - powf(hp->adam_beta1, (float)adam_step);
    const float bc2 = sqrtf(1.0f - powf(hp->adam_beta2, (float)adam_step));
    adam_kernel<<<(unsigned)((hi - lo + 255) / 256), 256, 0, s>>>(p->params, p->grads, adam_m, adam_v, lo, hi, nullptr, lr, nullptr, hp->adam_beta1,
                                                                 hp->adam_beta2, hp->adam_eps, bc1, bc2);
    LTP_CUDA(p, cudaGetLastError());
    if (p->has_enc) {
        const int rc = refresh_mlp(p, p->enc, s);
        if (rc) return rc;
    }
    return refresh_mlp(p, p->adapt, s);
}

int lt_linear_f32(const float* a, const float* b, const float* bias, float* out, int32_t M, int32_t N, int32_t K, int32_t splits, void* stream) {
    if (!a || !b || !out || M <= 0 || N <= 0 || K <= 0) return LT_ERR_INVALID;
    int dev = 0;
    cudaDeviceProp prop;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess || prop.major != 10)
        return fail(nullptr, LT_ERR_NO_DEVICE, "no sm_100 device");
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t ld = up(K, 8);
    __half *pa = nullptr, *pb = nullptr;
    float* part = nullptr;
    LTP_CUDA(nullptr, cudaMalloc(&pa, (size_t)2 * M * ld * 2));
    LTP_CUDA(nullptr, cudaMalloc(&pb, (size_t)2 * N * ld * 2));
    LTP_CUDA(nullptr, cudaMemsetAsync(pa, 0, (size_t)2 * M * ld * 2, s));
    LTP_CUDA(nullptr, cudaMemsetAsync(pb, 0, (size_t)2 * N * ld * 2, s));
    split_f32_kernel<<<(unsigned)(((int64_t)M * K + 255) / 256), 256, 0, s>>>(a, M, K, S_X, pa, ld, (int64_t)M * ld);
    split_f32_kernel<<<(unsigned)(((int64_t)N * K + 255) / 256), 256, 0, s>>>(b, N, K, S_W, pb, ld, (int64_t)N * ld);
    ltg::Params p;
    memset(&p, 0, sizeof p);
    p.out_scale = 1.0f / (S_X * S_W);
    int rc;
    if (splits <= 1) {
        p.mode = ltg::EPI_F32;
        p.bias = bias;
        p.out_f32 = out;
        p.ld_f32 = N;
        rc = run_gemm(nullptr, Operand{pa, M, ld, (int64_t)M * ld}, M, Operand{pb, N, ld, (int64_t)N * ld}, N, K, 1, p, s);
    } else {
        LTP_CUDA(nullptr, cudaMalloc(&part, (size_t)splits * M * N * 4));
        p.mode = ltg::EPI_PARTIAL;
        p.out_f32 = part;
        p.ld_f32 = N;
        p.split_stride = (int64_t)M * N;
        rc = run_gemm(nullptr, Operand{pa, M, ld, (int64_t)M * ld}, M, Operand{pb, N, ld, (int64_t)N * ld}, N, K, splits, p, s);
        if (!rc) sum_partials_kernel<<<(unsigned)(((int64_t)M * N + 255) / 256), 256, 0, s>>>(part, p.splits, (int64_t)M * N, N, bias, out);
    }
    cudaError_t e = cudaStreamSynchronize(s);
    cudaFree(pa);
    cudaFree(pb);
    if (part) cudaFree(part);
    if (rc) return rc;
    LTP_CUDA(nullptr, e);
    return LT_OK;
}

int lt_linear_tn_f32(const float* a, const float* b, float* out, int32_t M, int32_t N, int32_t K, int32_t splits, void* stream) {
    if (!a || !b || !out || M <= 0 || N <= 0 || K <= 0) return LT_ERR_INVALID;
    int dev = 0;
    cudaDeviceProp prop;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess || prop.major != 10)
        return fail(nullptr, LT_ERR_NO_DEVICE, "no sm_100 device");
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t lda = up(M, 8), ldb = up(N, 8);
    __half *pa = nullptr, *pb = nullptr;
    float* part = nullptr;
    LTP_CUDA(nullptr, cudaMalloc(&pa, (size_t)2 * K * lda * 2));
    LTP_CUDA(nullptr, cudaMalloc(&pb, (size_t)2 * K * ldb * 2));
    LTP_CUDA(nullptr, cudaMemsetAsync(pa, 0, (size_t)2 * K * lda * 2, s));
    LTP_CUDA(nullptr, cudaMemsetAsync(pb, 0, (size_t)2 * K * ldb * 2, s));
    split_f32_kernel<<<(unsigned)(((int64_t)K * M + 255) / 256), 256, 0, s>>>(a, K, M, S_X, pa, lda, (int64_t)K * lda);
    split_f32_kernel<<<(unsigned)(((int64_t)K * N + 255) / 256), 256, 0, s>>>(b, K, N, S_X, pb, ldb, (int64_t)K * ldb);
    ltg::Params p;
    memset(&p, 0, sizeof p);
    p.out_scale = 1.0f / (S_X * S_X);
    if (splits < 1) splits = 1;
    LTP_CUDA(nullptr, cudaMalloc(&part, (size_t)splits * M * N * 4));
    p.mode = ltg::EPI_PARTIAL;
    p.out_f32 = part;
    p.ld_f32 = N;
    p.split_stride = (int64_t)M * N;
    int rc = run_gemm_mn(nullptr, Operand{pa, K, lda, (int64_t)K * lda}, M, Operand{pb, K, ldb, (int64_t)K * ldb}, N, K, splits, p, s);
    if (!rc) sum_partials_kernel<<<(unsigned)(((int64_t)M * N + 255) / 256), 256, 0, s>>>(part, p.splits, (int64_t)M * N, N, nullptr, out);
    cudaError_t e = cudaStreamSynchronize(s);
    cudaFree(pa);
    cudaFree(pb);
    cudaFree(part);
    if (rc) return rc;
    LTP_CUDA(nullptr, e);
    return LT_OK;
}

int lt_linear_f32_static(const float* a, const float* b, const float* bias, float* out, int32_t M, int32_t N, int32_t K, void* stream) {
    if (!a || !b || !out || M <= 0 || N <= 0 || K <= 0) return fail(nullptr, LT_ERR_INVALID, "lt_linear_f32_static: NULL operand or empty shape");
    int dev = 0;
    cudaDeviceProp prop;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&prop, dev) != cudaSuccess || prop.major != 10)
        return fail(nullptr, LT_ERR_NO_DEVICE, "no sm_100 device");
    const int64_t ld = up(K, 8);
    const int64_t a_bytes = up((int64_t)2 * M * ld * 2, 256), b_bytes = up((int64_t)2 * N * ld * 2, 256);
    if (a_bytes + b_bytes > (int64_t)LT_LINEAR_STATIC_SCRATCH_BYTES)
        return fail(nullptr, LT_ERR_INVALID,
                    "lt_linear_f32_static: the fp16 operand planes of this shape need " + std::to_string(a_bytes + b_bytes) +
                        " bytes of scratch, the static array holds " + std::to_string((long long)LT_LINEAR_STATIC_SCRATCH_BYTES) +
                        " (4 * (M + N) * roundup(K, 8) must fit); use lt_linear_f32");
    unsigned char* scratch = static_scratch_ptr(dev);
    if (!scratch) return LT_ERR_CUDA;
    cudaStream_t s = (cudaStream_t)stream;
    __half* pa = reinterpret_cast<__half*>(scratch);
    __half* pb = reinterpret_cast<__half*>(scratch + a_bytes);
    // the array keeps the previous call's planes: the pad columns K..ld of this call's rows must read as zero
    LTP_CUDA(nullptr, cudaMemsetAsync(scratch, 0, (size_t)(a_bytes + b_bytes), s));
    split_f32_kernel<<<(unsigned)(((int64_t)M * K + 255) / 256), 256, 0, s>>>(a, M, K, S_X, pa, ld, (int64_t)M * ld);
    split_f32_kernel<<<(unsigned)(((int64_t)N * K + 255) / 256), 256, 0, s>>>(b, N, K, S_W, pb, ld, (int64_t)N * ld);
    LTP_CUDA(nullptr, cudaGetLastError());
    ltg::Params p;
    memset(&p, 0, sizeof p);
    p.out_scale = 1.0f / (S_X * S_W);
    p.mode = ltg::EPI_F32;
    p.bias = bias;
    p.out_f32 = out;
    p.ld_f32 = N;
    return run_gemm(nullptr, Operand{pa, M, ld, (int64_t)M * ld}, M, Operand{pb, N, ld, (int64_t)N * ld}, N, K, 1, p, s);
}

}   // extern "C"
