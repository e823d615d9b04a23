// lt_gemm.cuh -- the dense contraction of the actor-critic MLPs (go1_gym_learn/ppo_cse/actor_critic.py:43-82 nn.Linear
// stacks; forward, input-gradient and weight-gradient products of ppo.py:104-148) on the 5th-generation tensor cores.
//
//   D[M, N] = A[M, K] * B[N, K]^T        A, B: K-major, each stored as TWO fp16 planes (hi, lo) with a * s = hi + lo
//   D[M, N] = A[K, M]^T * B[K, N]        Params::mn_major: both operands MN-major (the contraction index is the ROW of the stored
//                                        matrices) -- the weight-gradient product dW^T = dZ^T * [X | 1] reads the row-major
//                                        activation / gradient planes as they are, so nothing is ever stored transposed
//
// fp32-equivalent arithmetic on fp16 operands: every product is three tcgen05.mma (hi*hi + hi*lo + lo*hi, the lo*lo term is
// below 2^-22 of the product) accumulated in fp32 in tensor memory; the power-of-two operand scales s are undone exactly in
// the epilogue.  The reference runs these GEMMs in fp32 (TF32 off), so single-pass fp16/bf16/tf32 would be a precision drop.
//
// Structure (one persistent CTA per SM, 320 threads):
//   warp 0      TMA producer: cp.async.bulk.tensor.3d loads a [2 planes, 128 rows, 64 k] box of A and a [2, bn, 64] box of B
//               per k-block into a 128B-swizzled STAGES-deep shared-memory ring (mbarrier complete_tx)
//   warp 1      MMA issuer: one elected lane, 4 k16 steps x 3 products per k-block, tcgen05.commit frees the ring slot;
//               accumulators double-buffered in TMEM (2 x BN_MAX columns) so tile i+1's main loop overlaps tile i's epilogue
//   warps 2..9  epilogue: thread <-> accumulator row (TMEM lane), two warps per lane quarter on alternating 32-column chunks;
//               tcgen05.ld 32 columns at a time, then one of
//               EPI_F32      out = acc * scale + bias                                    -> fp32 [M, ld]
//               EPI_ELU      y = act(acc * scale + bias), act = ELU unless Params::act says else -> fp16 planes, row-major
//               EPI_BWD      dz = acc * scale * act'(y_prev), e.g. ELU'(y) = y > 0 ? 1 : y+1   -> fp16 planes, row-major
//               EPI_PARTIAL  acc * scale -> fp32 split-K partial [split][M, ld]
// Tiles are (split, m-block, n-block) triples, n fastest, so that the CTAs running at the same time share A rows in L2.
// Rows / columns / k beyond the tensor bounds are zero-filled by TMA and masked in the epilogue: no padding requirements
// beyond 16-byte row pitches.
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace ltg {

constexpr int BM = 128;          // rows per tile = TMEM lanes
constexpr int BK = 64;           // k per ring slot = one 128-byte swizzle row of fp16
constexpr int EPI_WARPS = 8;         // two per TMEM lane quarter, each taking every other 32-column chunk
constexpr int THREADS = 64 + 32 * EPI_WARPS;
constexpr float OPERAND_MAX = 60000.0f;      // fp16 planes saturate here instead of overflowing to inf

enum { EPI_F32 = 0, EPI_ELU = 1, EPI_BWD = 2, EPI_PARTIAL = 3 };        // EPI_ELU: "activation" epilogue (ELU is the reference default)
// get_activation (go1_gym_learn/ppo_cse/actor_critic.py:162-178); values = LT_ACT_* of include/lt_policy.h
enum { ACT_ELU = 0, ACT_SELU = 1, ACT_RELU = 2, ACT_LRELU = 3, ACT_TANH = 4, ACT_SIGMOID = 5 };

struct Params {
    CUtensorMap tma_a, tma_b;            // 3-D maps {k, row, plane}; box {64, 128, 2} / {64, bn, 2}; SWIZZLE_128B; zero OOB fill
    CUtensorMap tma_out;                 // EPI_ELU / EPI_BWD output planes {col, row, plane}; box {32, 32, 2}; SWIZZLE_64B (store)
    int32_t use_tma_out;
    int32_t M, N, K;
    int32_t bn;                          // N tile = UMMA N (multiple of 16, <= BN_MAX)
    int32_t splits, kb_per_split;        // split-K: k-blocks [s * kb_per_split, (s+1) * kb_per_split) go to partial s
    int32_t mode;
    int32_t mn_major;                    // 1: tma_a / tma_b are {col, row, plane} maps with box {64, 64, 2}; contraction over rows
    int32_t act;                         // ACT_*: activation of EPI_ELU, its derivative (from the stored OUTPUT y) in EPI_BWD
    float out_scale;                     // 1 / (s_a * s_b)
    const float* bias;                   // [N] or NULL
    float* out_f32;                      // EPI_F32 / EPI_PARTIAL
    int64_t ld_f32, split_stride;
    __half* out_h;                       // row-major planes [2][rows][ld_h]
    int64_t ld_h, plane_h;
    float split_scale;                   // scale of the fp16 planes written by EPI_ELU / EPI_BWD
    // gradient operands carry a per-tensor power-of-two scale that lives on the DEVICE (chosen from the tensor's measured
    // largest entry, see lt_policy.cu "gradient scales"), so that no host synchronisation is needed to pick it:
    const float* a_scale_ptr;            // != NULL: the A planes were written with this scale -> out_scale is divided by it
    const float* split_scale_ptr;        // != NULL: overrides split_scale for the planes EPI_BWD writes
    uint32_t* amax_out;                  // != NULL: EPI_BWD records max |dz| (fp32 bit pattern, atomicMax) of what it writes
    const __half* act_h;                 // EPI_BWD: y_prev planes (row-major, scale 1 / act_inv_scale)
    int64_t ld_act, plane_act;
    float act_inv_scale;
};

// ---------------------------------------------------------------------------------------------------------------- PTX
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int32_t c0, int32_t c1, int32_t c2) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
                 "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, uint32_t src, int32_t c0, int32_t c1, int32_t c2) {
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(reinterpret_cast<uint64_t>(map)),
                 "r"(src), "r"(c0), "r"(c1), "r"(c2)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]
__device__ __forceinline__ void umma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// shared-memory matrix descriptor: K-major, SWIZZLE_128B (8-row x 128-byte atoms, 1024 bytes between 8-row groups), version 1
__device__ __forceinline__ uint64_t sw128_desc(uint32_t smem_addr) {
    return (uint64_t)((smem_addr & 0x3ffffu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// MN-major, SWIZZLE_128B: atoms of 64 mn-elements (128 B) x 8 k-rows; 8-row k groups 1024 B apart (stride byte offset), 64-wide mn
// chunks `chunk_bytes` apart (leading byte offset)
__device__ __forceinline__ uint64_t sw128_mn_desc(uint32_t smem_addr, uint32_t chunk_bytes) {
    return (uint64_t)((smem_addr & 0x3ffffu) >> 4) | ((uint64_t)(chunk_bytes >> 4) << 16) | ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) |
           (2ull << 61);
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];\n\t"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
          "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
          "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]),
          "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];\n\t"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
          "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}

// a * s = hi + lo in fp16 (hi rounded to nearest, lo the rounded residual: 22 significand bits in total)
__device__ __forceinline__ void split_pair(float x0, float x1, uint32_t& hi, uint32_t& lo) {
    x0 = fminf(fmaxf(x0, -OPERAND_MAX), OPERAND_MAX);
    x1 = fminf(fmaxf(x1, -OPERAND_MAX), OPERAND_MAX);
    const __half2 h = __floats2half2_rn(x0, x1);
    const float2 hf = __half22float2(h);
    const __half2 l = __floats2half2_rn(x0 - hf.x, x1 - hf.y);
    hi = *reinterpret_cast<const uint32_t*>(&h);
    lo = *reinterpret_cast<const uint32_t*>(&l);
}
// fast-path variant: one saturating pack (F2FP.SATFINITE) per plane instead of four clamps; x - hi is exact in fp32
__device__ __forceinline__ void split_pair_sat(float x0, float x1, uint32_t& hi, uint32_t& lo) {
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(hi) : "f"(x1), "f"(x0));
    const float2 hf = __half22float2(*reinterpret_cast<const __half2*>(&hi));
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(lo) : "f"(x1 - hf.y), "f"(x0 - hf.x));
}
// ELU for the fast path: ex2.approx - 1 below zero.  Absolute error <= 1.2e-7 (one fp32 ulp of the O(1) activations it is added
// to in the next layer); the series branch of act_fwd<ACT_ELU> only improves the RELATIVE error of outputs below 0.03.
__device__ __forceinline__ float elu_fast(float z) { return z > 0.0f ? z : __expf(z) - 1.0f; }
constexpr float SELU_ALPHA = 1.6732632423543772848170429916717f, SELU_SCALE = 1.0507009873554804934193349852946f;
template <int ACT>
__device__ __forceinline__ float act_fwd(float z) {
    if (ACT == ACT_ELU) {                                                                  // nn.ELU(alpha = 1)
        // expm1 for z <= 0 without the slow path: 4th-order series below 1/32 (truncation 3e-10), ex2.approx - 1 beyond (absolute
        // error 1e-7 on a value of magnitude >= 0.03)
        const float series = z * (1.0f + z * (0.5f + z * (0.16666667f + z * 0.041666668f)));
        return z > 0.0f ? z : (z > -0.03125f ? series : __expf(z) - 1.0f);
    }
    if (ACT == ACT_SELU) return SELU_SCALE * (z > 0.0f ? z : SELU_ALPHA * expm1f(z));
    if (ACT == ACT_RELU) return fmaxf(z, 0.0f);
    if (ACT == ACT_LRELU) return z > 0.0f ? z : 0.01f * z;                                  // nn.LeakyReLU default slope
    if (ACT == ACT_TANH) return tanhf(z);
    return 1.0f / (1.0f + expf(-z));
}
// derivative with respect to the pre-activation, written in terms of the OUTPUT y (the only thing the forward pass keeps)
template <int ACT>
__device__ __forceinline__ float act_bwd(float y) {
    if (ACT == ACT_ELU) return y > 0.0f ? 1.0f : y + 1.0f;
    if (ACT == ACT_SELU) return y > 0.0f ? SELU_SCALE : y + SELU_SCALE * SELU_ALPHA;
    if (ACT == ACT_RELU) return y > 0.0f ? 1.0f : 0.0f;
    if (ACT == ACT_LRELU) return y > 0.0f ? 1.0f : 0.01f;
    if (ACT == ACT_TANH) return 1.0f - y * y;
    return y * (1.0f - y);
}
#define LTG_ACT_SWITCH(act, BODY)                                        \
    switch (act) {                                                       \
        case ACT_SELU: { constexpr int A_ = ACT_SELU; BODY } break;      \
        case ACT_RELU: { constexpr int A_ = ACT_RELU; BODY } break;      \
        case ACT_LRELU: { constexpr int A_ = ACT_LRELU; BODY } break;    \
        case ACT_TANH: { constexpr int A_ = ACT_TANH; BODY } break;      \
        case ACT_SIGMOID: { constexpr int A_ = ACT_SIGMOID; BODY } break;\
        default: { constexpr int A_ = ACT_ELU; BODY } break;             \
    }

template <int BN_MAX, int STAGES>
struct Smem {
    static constexpr uint32_t A_BYTES = 2u * BM * BK * 2u;           // both planes of the A box
    static constexpr uint32_t B_BYTES = 2u * BN_MAX * BK * 2u;
    static constexpr uint32_t STAGE = A_BYTES + B_BYTES;
    static constexpr uint32_t OUT = STAGES * STAGE;                  // epilogue staging: per warp [2 planes][32 rows][64 B], 64B-swizzled
    static constexpr uint32_t OUT_WARP = 2u * 32u * 64u;
    static constexpr uint32_t BARS = OUT + EPI_WARPS * OUT_WARP;     // full[STAGES] empty[STAGES] tfull[2] tempty[2] tmem slot
    static constexpr uint32_t TOTAL = BARS + (2 * STAGES + 4) * 8 + 16 + 1024;     // + alignment slack
};

template <int BN_MAX, int STAGES>
__global__ void __launch_bounds__(THREADS, 1) lt_gemm_kernel(const __grid_constant__ Params p) {
    using S = Smem<BN_MAX, STAGES>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bars = base + S::BARS;
    auto full = [&](int s) { return bars + 8u * s; };
    auto empty = [&](int s) { return bars + 8u * (STAGES + s); };
    auto tfull = [&](int s) { return bars + 8u * (2 * STAGES + s); };
    auto tempty = [&](int s) { return bars + 8u * (2 * STAGES + 2 + s); };
    constexpr uint32_t CHUNK = 2u * 64u * BK * 2u;                   // MN-major: one {64 mn, 64 k, 2 planes} box = 16 KB
    const uint32_t tmem_slot = bars + 8u * (2 * STAGES + 4);
    const uint32_t out_stage = base + S::OUT;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(full(s), 1);
            mbar_init(empty(s), 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(tfull(s), 1);
            mbar_init(tempty(s), EPI_WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(2 * BN_MAX) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_before();
    __syncthreads();
    fence_after();
    uint32_t tmem;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem) : "r"(tmem_slot));

    const int m_tiles = (p.M + BM - 1) / BM, n_tiles = (p.N + p.bn - 1) / p.bn;
    const int kb_total = (p.K + BK - 1) / BK;
    const int n_work = m_tiles * n_tiles * p.splits;

    if (warp == 0) {
        // ------------------------------------------------------------------------------------------ TMA producer
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&p.tma_a)) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&p.tma_b)) : "memory");
            const int b_chunks = (p.bn + 63) / 64;
            const uint32_t tx = p.mn_major ? (uint32_t)(BM / 64 + b_chunks) * CHUNK : S::A_BYTES + 2u * (uint32_t)p.bn * BK * 2u;
            int stage = 0;
            uint32_t phase = 0;
            for (int w = blockIdx.x; w < n_work; w += gridDim.x) {
                const int n_blk = w % n_tiles, m_blk = (w / n_tiles) % m_tiles, split = w / (n_tiles * m_tiles);
                const int kb0 = split * p.kb_per_split, kb1 = min(kb_total, kb0 + p.kb_per_split);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(empty(stage), phase ^ 1u);
                    mbar_expect_tx(full(stage), tx);
                    const uint32_t sa = base + (uint32_t)stage * S::STAGE;
                    if (p.mn_major) {
#pragma unroll
                        for (int c = 0; c < BM / 64; ++c) tma_load_3d(sa + c * CHUNK, &p.tma_a, full(stage), m_blk * BM + c * 64, kb * BK, 0);
                        for (int c = 0; c < b_chunks; ++c)
                            tma_load_3d(sa + S::A_BYTES + c * CHUNK, &p.tma_b, full(stage), n_blk * p.bn + c * 64, kb * BK, 0);
                    } else {
                        tma_load_3d(sa, &p.tma_a, full(stage), kb * BK, m_blk * BM, 0);
                        tma_load_3d(sa + S::A_BYTES, &p.tma_b, full(stage), kb * BK, n_blk * p.bn, 0);
                    }
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1u;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------------------------------ MMA issuer
        // instruction descriptor: D = F32, A = B = F16, both K-major, N = bn, M = 128
        // (bits 15 / 16: A / B are MN-major)
        const uint32_t idesc = (1u << 4) | ((uint32_t)(p.bn >> 3) << 17) | ((uint32_t)(BM >> 4) << 24) | (p.mn_major ? (3u << 15) : 0u);
        const uint32_t b_plane = (uint32_t)p.bn * BK * 2u;
        int stage = 0, as = 0;
        uint32_t phase = 0, aphase = 0;
        for (int w = blockIdx.x; w < n_work; w += gridDim.x) {
            const int split = w / (n_tiles * m_tiles);
            const int kb0 = split * p.kb_per_split, kb1 = min(kb_total, kb0 + p.kb_per_split);
            mbar_wait(tempty(as), aphase ^ 1u);
            fence_after();
            const uint32_t d_tmem = tmem + (uint32_t)as * BN_MAX;
            for (int kb = kb0; kb < kb1; ++kb) {
                mbar_wait(full(stage), phase);
                fence_after();
                if (lane == 0 && p.mn_major) {
                    // smem per operand: [64-wide mn chunk][plane][64 k rows][128 B]; one MMA reads 16 k rows = 2048 B further on
                    const uint32_t sa = base + (uint32_t)stage * S::STAGE, sb = sa + S::A_BYTES;
                    const uint64_t a_hi = sw128_mn_desc(sa, CHUNK), a_lo = sw128_mn_desc(sa + CHUNK / 2, CHUNK);
                    const uint64_t b_hi = sw128_mn_desc(sb, CHUNK), b_lo = sw128_mn_desc(sb + CHUNK / 2, CHUNK);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {
                        const uint64_t adv = (uint64_t)((k * 16 * 128) >> 4);
                        umma_ss(d_tmem, a_lo + adv, b_hi + adv, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
                        umma_ss(d_tmem, a_hi + adv, b_lo + adv, idesc, 1u);
                        umma_ss(d_tmem, a_hi + adv, b_hi + adv, idesc, 1u);
                    }
                    umma_commit(empty(stage));
                    if (kb == kb1 - 1) umma_commit(tfull(as));
                } else if (lane == 0) {
                    const uint32_t sa = base + (uint32_t)stage * S::STAGE, sb = sa + S::A_BYTES;
                    const uint64_t a_hi = sw128_desc(sa), a_lo = sw128_desc(sa + BM * BK * 2u);
                    const uint64_t b_hi = sw128_desc(sb), b_lo = sw128_desc(sb + b_plane);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {
                        const uint64_t adv = (uint64_t)((k * 32) >> 4);          // 16 fp16 = 32 bytes inside the swizzle row
                        umma_ss(d_tmem, a_lo + adv, b_hi + adv, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
                        umma_ss(d_tmem, a_hi + adv, b_lo + adv, idesc, 1u);
                        umma_ss(d_tmem, a_hi + adv, b_hi + adv, idesc, 1u);
                    }
                    umma_commit(empty(stage));                                    // slot free once these MMAs have read it
                    if (kb == kb1 - 1) umma_commit(tfull(as));
                }
                __syncwarp();
                if (++stage == STAGES) {
                    stage = 0;
                    phase ^= 1u;
                }
            }
            if (kb1 <= kb0 && lane == 0) umma_commit(tfull(as));                  // empty k range (never with a valid split plan)
            if (++as == 2) {
                as = 0;
                aphase ^= 1u;
            }
        }
    } else {
        // ------------------------------------------------------------------------------------------ epilogue
        const int q = warp & 3;                                                   // TMEM lane quarter this warp may read
        const int half = (warp - 2) >> 2;                                         // which of the two warps of that quarter
        int as = 0;
        uint32_t aphase = 0;
        for (int w = blockIdx.x; w < n_work; w += gridDim.x) {
            const int n_blk = w % n_tiles, m_blk = (w / n_tiles) % m_tiles, split = w / (n_tiles * m_tiles);
            const int row = m_blk * BM + q * 32 + lane;
            const bool row_ok = row < p.M;
            const int n0 = n_blk * p.bn;
            mbar_wait(tfull(as), aphase);
            fence_after();
            const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)as * BN_MAX;
            const float out_scale = p.a_scale_ptr ? p.out_scale / __ldg(p.a_scale_ptr) : p.out_scale;
            const float split_scale = p.split_scale_ptr ? __ldg(p.split_scale_ptr) : p.split_scale;
            float tile_amax = 0.0f;
            // fast path: ELU bodies (the reference default), whole 32-column chunks inside the matrix, 16-byte aligned rows / bias
            const bool fast_tile = p.act == ACT_ELU && (p.ld_h & 7) == 0 && (n0 & 7) == 0 &&
                                   ((p.mode == EPI_ELU && p.bias && (reinterpret_cast<uintptr_t>(p.bias) & 15) == 0) ||
                                    (p.mode == EPI_BWD && (!p.act_h || (p.ld_act & 7) == 0)));
            for (int c = 32 * half; c < p.bn; c += 64) {
                uint32_t r[32];
                const int width = min(32, p.bn - c);                              // 32 or 16 (bn is a multiple of 16)
                const int col0 = n0 + c;
                if (fast_tile && width == 32 && col0 + 32 <= p.N) {
                    uint32_t hi[16], lo[16];
                    if (p.mode == EPI_ELU) {
                        float4 b4[8];
#pragma unroll
                        for (int g = 0; g < 8; ++g) b4[g] = __ldg(reinterpret_cast<const float4*>(p.bias + col0) + g);
                        tmem_ld32(taddr + c, r);
#pragma unroll
                        for (int g = 0; g < 8; ++g) {
                            const float y0 = elu_fast(fmaf(__uint_as_float(r[4 * g]), out_scale, b4[g].x));
                            const float y1 = elu_fast(fmaf(__uint_as_float(r[4 * g + 1]), out_scale, b4[g].y));
                            const float y2 = elu_fast(fmaf(__uint_as_float(r[4 * g + 2]), out_scale, b4[g].z));
                            const float y3 = elu_fast(fmaf(__uint_as_float(r[4 * g + 3]), out_scale, b4[g].w));
                            split_pair_sat(y0 * split_scale, y1 * split_scale, hi[2 * g], lo[2 * g]);
                            split_pair_sat(y2 * split_scale, y3 * split_scale, hi[2 * g + 1], lo[2 * g + 1]);
                        }
                    } else {
                        uint4 qh[4], ql[4];
                        const bool have_act = p.act_h != nullptr;
                        if (have_act && row_ok) {
                            const uint4* ah = reinterpret_cast<const uint4*>(p.act_h + (int64_t)row * p.ld_act + col0);
                            const uint4* al = reinterpret_cast<const uint4*>(p.act_h + p.plane_act + (int64_t)row * p.ld_act + col0);
#pragma unroll
                            for (int g = 0; g < 4; ++g) {
                                qh[g] = __ldg(ah + g);
                                ql[g] = __ldg(al + g);
                            }
                        } else {
#pragma unroll
                            for (int g = 0; g < 4; ++g) qh[g] = ql[g] = make_uint4(0u, 0u, 0u, 0u);
                        }
                        tmem_ld32(taddr + c, r);
                        const float sc = split_scale * out_scale;
                        const float amax_scale = 1.0f / split_scale;
                        float m = 0.0f;
#pragma unroll
                        for (int g = 0; g < 4; ++g) {
                            const uint32_t wh[4] = {qh[g].x, qh[g].y, qh[g].z, qh[g].w}, wl[4] = {ql[g].x, ql[g].y, ql[g].z, ql[g].w};
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                const float2 fh = __half22float2(*reinterpret_cast<const __half2*>(&wh[e]));
                                const float2 fl = __half22float2(*reinterpret_cast<const __half2*>(&wl[e]));
                                const float y0 = (fh.x + fl.x) * p.act_inv_scale, y1 = (fh.y + fl.y) * p.act_inv_scale;
                                const float d0 = (have_act && y0 <= 0.0f) ? y0 + 1.0f : 1.0f, d1 = (have_act && y1 <= 0.0f) ? y1 + 1.0f : 1.0f;
                                const float x0 = __uint_as_float(r[g * 8 + 2 * e]) * sc * d0, x1 = __uint_as_float(r[g * 8 + 2 * e + 1]) * sc * d1;
                                m = fmaxf(m, fmaxf(fabsf(x0), fabsf(x1)));
                                split_pair_sat(x0, x1, hi[g * 4 + e], lo[g * 4 + e]);
                            }
                        }
                        tile_amax = fmaxf(tile_amax, m * amax_scale);
                    }
                    if (p.use_tma_out) {
                        // stage this warp's [32 rows x 32 columns] x 2 planes in shared memory (64B swizzle: 16-byte chunk j of row r
                        // sits at chunk j ^ ((r >> 1) & 3): conflict-free 16-byte stores) and hand it to the TMA: one bulk tensor
                        // store per chunk instead of 8 row-per-lane store instructions of 32 cache lines each; rows / columns
                        // beyond the matrix are clipped by the tensor map
                        const uint32_t buf = out_stage + (uint32_t)(warp - 2) * S::OUT_WARP;
                        if (lane == 0) tma_store_wait_read();                      // the previous chunk's store has read the buffer
                        __syncwarp();
                        const uint32_t rowb = buf + (uint32_t)lane * 64u, sw = (uint32_t)((lane >> 1) & 3);
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const uint32_t o = (((uint32_t)j ^ sw) << 4);
                            st_shared_v4(rowb + o, hi[4 * j], hi[4 * j + 1], hi[4 * j + 2], hi[4 * j + 3]);
                            st_shared_v4(rowb + 2048u + o, lo[4 * j], lo[4 * j + 1], lo[4 * j + 2], lo[4 * j + 3]);
                        }
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        __syncwarp();
                        if (lane == 0) tma_store_3d(&p.tma_out, buf, col0, m_blk * BM + q * 32, 0);
                    } else if (row_ok) {
                        uint4* oh = reinterpret_cast<uint4*>(p.out_h + (int64_t)row * p.ld_h + col0);
                        uint4* ol = reinterpret_cast<uint4*>(p.out_h + p.plane_h + (int64_t)row * p.ld_h + col0);
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            oh[j] = make_uint4(hi[4 * j], hi[4 * j + 1], hi[4 * j + 2], hi[4 * j + 3]);
                            ol[j] = make_uint4(lo[4 * j], lo[4 * j + 1], lo[4 * j + 2], lo[4 * j + 3]);
                        }
                    }
                    continue;
                }
                if (width == 32) tmem_ld32(taddr + c, r);
                else tmem_ld16(taddr + c, r);
                float v[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]) * out_scale;
                if (p.mode == EPI_F32 || p.mode == EPI_PARTIAL) {
                    float* out = p.out_f32 + (p.mode == EPI_PARTIAL ? (int64_t)split * p.split_stride : 0) + (int64_t)row * p.ld_f32;
                    if (row_ok) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const int n = col0 + j;
                            if (j < width && n < p.N) out[n] = v[j] + ((p.bias && p.mode == EPI_F32) ? __ldg(p.bias + n) : 0.0f);
                        }
                    }
                } else {
                    if (p.mode == EPI_ELU) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const int n = col0 + j;
                            v[j] += (p.bias && j < width && n < p.N) ? __ldg(p.bias + n) : 0.0f;
                        }
                        LTG_ACT_SWITCH(p.act, _Pragma("unroll") for (int j = 0; j < 32; ++j) v[j] = act_fwd<A_>(v[j]);)
                    } else if (p.act_h) {                                         // EPI_BWD: times ELU'(y_prev); act_h == NULL: linear layer
                        if (row_ok) {
                            const __half* ah = p.act_h + (int64_t)row * p.ld_act + col0;
                            const __half* al = ah + p.plane_act;
                            if (col0 + width <= p.N && (p.ld_act & 7) == 0 && (col0 & 7) == 0) {
#pragma unroll
                                for (int g = 0; g < 4; ++g) {
                                    if (g * 8 < width) {
                                        const uint4 qh = __ldg(reinterpret_cast<const uint4*>(ah) + g), ql = __ldg(reinterpret_cast<const uint4*>(al) + g);
                                        const uint32_t wh[4] = {qh.x, qh.y, qh.z, qh.w}, wl[4] = {ql.x, ql.y, ql.z, ql.w};
#pragma unroll
                                        for (int e = 0; e < 4; ++e) {
                                            const float2 fh = __half22float2(*reinterpret_cast<const __half2*>(&wh[e]));
                                            const float2 fl = __half22float2(*reinterpret_cast<const __half2*>(&wl[e]));
                                            const float y0 = (fh.x + fl.x) * p.act_inv_scale, y1 = (fh.y + fl.y) * p.act_inv_scale;
                                            LTG_ACT_SWITCH(p.act, v[g * 8 + 2 * e] *= act_bwd<A_>(y0); v[g * 8 + 2 * e + 1] *= act_bwd<A_>(y1);)
                                        }
                                    }
                                }
                            } else {
#pragma unroll
                                for (int j = 0; j < 32; ++j) {
                                    const int n = col0 + j;
                                    if (j < width && n < p.N) {
                                        const float y = (__half2float(ah[j]) + __half2float(al[j])) * p.act_inv_scale;
                                        LTG_ACT_SWITCH(p.act, v[j] *= act_bwd<A_>(y);)
                                    }
                                }
                            }
                        }
                    }
                    if (p.amax_out) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) tile_amax = fmaxf(tile_amax, j < width ? fabsf(v[j]) : 0.0f);      // fmaxf drops NaNs; x16 loads leave r[16..] unset
                    }
                    uint32_t hi[16], lo[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) split_pair(v[2 * j] * split_scale, v[2 * j + 1] * split_scale, hi[j], lo[j]);
                    if (row_ok) {
                        // row-major planes: this thread's 32 (or 16) consecutive columns, 16-byte stores where the row pitch allows
                        __half* oh = p.out_h + (int64_t)row * p.ld_h + col0;
                        __half* ol = oh + p.plane_h;
                        if (col0 + width <= p.N && (p.ld_h & 7) == 0 && (col0 & 7) == 0) {
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                if (j * 8 < width) {
                                    reinterpret_cast<uint4*>(oh)[j] = make_uint4(hi[4 * j], hi[4 * j + 1], hi[4 * j + 2], hi[4 * j + 3]);
                                    reinterpret_cast<uint4*>(ol)[j] = make_uint4(lo[4 * j], lo[4 * j + 1], lo[4 * j + 2], lo[4 * j + 3]);
                                }
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j) {
                                if (j < width && col0 + j < p.N) {
                                    reinterpret_cast<uint16_t*>(oh)[j] = (uint16_t)(hi[j >> 1] >> ((j & 1) * 16));
                                    reinterpret_cast<uint16_t*>(ol)[j] = (uint16_t)(lo[j >> 1] >> ((j & 1) * 16));
                                }
                            }
                        }
                    }
                }
            }
            if (p.amax_out) {
                const uint32_t m = __reduce_max_sync(0xffffffffu, __float_as_uint(fminf(tile_amax, 3.0e38f)));
                if (lane == 0 && m) atomicMax(p.amax_out, m);
            }
            fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty(as));
            if (++as == 2) {
                as = 0;
                aphase ^= 1u;
            }
        }
    }
    if (warp >= 2 && lane == 0 && p.use_tma_out) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");      // stores complete before exit
    fence_before();
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(2 * BN_MAX) : "memory");
}

// ---------------------------------------------------------------------------------------------------------------- host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(ptr);
    }
    return fn;
}

// planes [2][rows][ld] fp16 at `base`; logical extent rows x k; box {64, box_rows, 2}
inline bool make_operand_map(CUtensorMap* map, const __half* base, int64_t rows, int64_t k, int64_t ld, int64_t plane, int box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)k, (cuuint64_t)rows, 2};
    const cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)plane * 2};
    const cuuint32_t box[3] = {(cuuint32_t)BK, (cuuint32_t)box_rows, 2};
    const cuuint32_t estr[3] = {1, 1, 1};
    return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<__half*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// MN-major operand: planes [2][rows][ld] fp16 at `base`, logical extent rows (contraction) x cols (mn); box {64 cols, 64 rows, 2}
inline bool make_operand_map_mn(CUtensorMap* map, const __half* base, int64_t rows, int64_t cols, int64_t ld, int64_t plane) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, 2};
    const cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)plane * 2};
    const cuuint32_t box[3] = {64, (cuuint32_t)BK, 2};
    const cuuint32_t estr[3] = {1, 1, 1};
    return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<__half*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// output planes [2][rows][ld] fp16: store map {col, row, plane}, box {32, 32, 2}, SWIZZLE_64B
inline bool make_out_map(CUtensorMap* map, __half* base, int64_t rows, int64_t cols, int64_t ld, int64_t plane) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return false;
    const cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, 2};
    const cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)plane * 2};
    const cuuint32_t box[3] = {32, 32, 2};
    const cuuint32_t estr[3] = {1, 1, 1};
    return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <int BN_MAX, int STAGES>
inline cudaError_t launch_t(const Params& p, int sm_count, cudaStream_t stream) {
    using S = Smem<BN_MAX, STAGES>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(lt_gemm_kernel<BN_MAX, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::TOTAL);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    const int m_tiles = (p.M + BM - 1) / BM, n_tiles = (p.N + p.bn - 1) / p.bn;
    const int n_work = m_tiles * n_tiles * p.splits;
    if (n_work <= 0) return cudaSuccess;
    lt_gemm_kernel<BN_MAX, STAGES><<<n_work < sm_count ? n_work : sm_count, THREADS, S::TOTAL, stream>>>(p);
    return cudaGetLastError();
}

inline cudaError_t launch(const Params& p, int sm_count, cudaStream_t stream) {
    if (p.bn > 128) return launch_t<256, 2>(p, sm_count, stream);
    return launch_t<128, 3>(p, sm_count, stream);
}

}   // namespace ltg
