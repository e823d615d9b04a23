// lt_torque_mma.cu -- _compute_torques (LR:957-996) with the actuator network (LR:1307-1322) on the tensor pipe.
//
// The network is Linear(6,32) -> softsign -> Linear(32,32) -> softsign -> Linear(32,1) per (env, dof) sample: 1,248 MACs,
// 82 % of them in the 32x32 layer.  On the FP32 pipe the kernel was FMA-issue bound (71 % fma-pipe active, 45 us per
// sub-step at 65,536 envs; profiles/r01_ncu_full.md).  Here a warp owns a tile of 32 consecutive samples = two m16 row
// blocks of `mma.sync.aligned.m16n8k16 .f16` with fp32 accumulators (samples on M, units on N):
//   * the 1e-5 torque tolerance rules out single-pass half/TF32 operands, so every operand is split in two fp16 parts,
//     a = a_hi + a_lo with a_hi = the leading 11 significand bits of a, a_lo = rn_f16(a - a_hi) (22 significand bits; fp16 subnormals keep the absolute
//     error of the low part at 2^-25, and every operand here is O(1): weights |w| <= 1.06, activations in (-1, 1)), and
//     every product is  a*w = a_lo*w_hi + a_hi*w_lo + a_hi*w_hi  accumulated in fp32, small terms first.  The dropped
//     a_lo*w_lo term is 2^-22 relative.  (A first version used the same split on m16n8k8 .tf32: HMMA.1688.F32.TF32 issues
//     at ~8 cycles per SMSP = 1,024 FLOP/clk/SM, 120 of them per tile kept the tensor pipe 56 % busy and the kernel at
//     43 us, no faster than the FP32 pipe; profiles/r01_ncu_torque_tf32x3.md.  The fp16 shape does twice the K per issue.)
//   * the accumulator fragments of two adjacent n-tiles of layer 1 ARE, after softsign and packing, the A fragment of one
//     k16 step of layer 2 (row g | g+8, columns 2t, 2t+1 | 2t+8, 2t+9), so no shuffle or shared-memory round trip sits
//     between the layers;
//   * the 32->1 output layer is 8 FMAs per row per thread + two xor-shuffles;
//   * all weight fragments (57 registers per thread, hi and lo parts pre-split on the host) are loaded once per warp; the
//     kernel is persistent (grid = resident CTAs), the next tile's ten input rows are fetched before the current tile's
//     network runs.
// The per-sample scalar work around the network (action scaling, lag ring, joint target, history shift, motor strength,
// clip) is lt_core.cuh's lt_torque_pre / lt_torque_post, one lane per sample, exactly as in the FP32-pipe kernel.
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "lt_core.cuh"
#include "lt_torque_mma.h"

#define LT_MMA_THREADS 128
#define LT_MMA_WARPS (LT_MMA_THREADS / 32)
#ifndef LT_MMA_MB
#define LT_MMA_MB 2                 // 16-row blocks of a tile in flight together (independent HMMA chains: 4 per block)
#endif
#ifndef LT_MMA_MINB
#define LT_MMA_MINB 3
#endif

// fragment register map (per lane; the table in global memory is [LT_MMA_FRAG_REGS][32] 32-bit words, lane-minor).
// W0/W1 words hold two fp16 (low half = lower k); the rest are fp32.
enum { FR_W0_HI = 0, FR_W0_LO = 4, FR_W1_HI = 8, FR_W1_LO = 24, FR_B1 = 40, FR_W2 = 48, FR_B2 = 56 };
static_assert(FR_B2 + 1 == LT_MMA_FRAG_REGS, "fragment map");

static uint32_t lt_host_pack_h2(float lo_k, float hi_k) {
    const __half a = __float2half_rn(lo_k), b = __float2half_rn(hi_k);
    uint16_t ua, ub;
    memcpy(&ua, &a, 2);
    memcpy(&ub, &b, 2);
    return (uint32_t)ua | ((uint32_t)ub << 16);
}
static float lt_host_h_residual(float w) { return w - __half2float(__float2half_rn(w)); }

void lt_torque_mma_build_frags(const lt_act_weights& W, float* out_f) {
    uint32_t* out = reinterpret_cast<uint32_t*>(out_f);
    for (int lane = 0; lane < 32; ++lane) {
        const int g = lane >> 2, t = lane & 3;
        auto put = [&](int reg, uint32_t v) { out[reg * 32 + lane] = v; };
        auto putf = [&](int reg, float v) { uint32_t u; memcpy(&u, &v, 4); out[reg * 32 + lane] = u; };
        auto put_split = [&](int hi_reg, int lo_reg, float wa, float wb) {       // two consecutive k of one unit
            put(hi_reg, lt_host_pack_h2(wa, wb));
            put(lo_reg, lt_host_pack_h2(lt_host_h_residual(wa), lt_host_h_residual(wb)));
        };
        for (int nt = 0; nt < 4; ++nt) {
            const int unit = 8 * nt + g;                       // B[k][n = g] of n-tile nt
            // layer 1 (m16n8k8): b0 = (k = 2t, 2t+1); input 6 is the constant 1 that carries the bias, input 7 is zero padding
            const float* w0 = W.w0 + unit * LT_ACT_IN;
            put_split(FR_W0_HI + nt, FR_W0_LO + nt, 2 * t < LT_ACT_IN ? w0[2 * t] : W.b0[unit], 2 * t + 1 < LT_ACT_IN ? w0[2 * t + 1] : 0.0f);
            // layer 2 (m16n8k16), k-step q: b0 = (k = 16q + 2t, +1), b1 = (k = 16q + 8 + 2t, +1)
            const float* w1 = W.w1 + unit * LT_ACT_HIDDEN;
            for (int q = 0; q < 2; ++q) {
                const int r = 2 * (4 * q + nt);
                put_split(FR_W1_HI + r, FR_W1_LO + r, w1[16 * q + 2 * t], w1[16 * q + 2 * t + 1]);
                put_split(FR_W1_HI + r + 1, FR_W1_LO + r + 1, w1[16 * q + 8 + 2 * t], w1[16 * q + 8 + 2 * t + 1]);
            }
            // accumulator columns of thread (g, t) in n-tile nt: 8nt + 2t, 8nt + 2t + 1
            putf(FR_B1 + 2 * nt, W.b1[8 * nt + 2 * t]);
            putf(FR_B1 + 2 * nt + 1, W.b1[8 * nt + 2 * t + 1]);
            putf(FR_W2 + 2 * nt, W.w2[8 * nt + 2 * t]);
            putf(FR_W2 + 2 * nt + 1, W.w2[8 * nt + 2 * t + 1]);
        }
        putf(FR_B2, W.b2[0]);
    }
}

struct lt_torque_mma_params {
    lt_config c;
    const uint32_t* frag;
    const float* actions_raw;
    const float* dof_state;
    lt_state_ptrs st;
    const int32_t* counters;
    int32_t substep;
    int32_t n_tiles;
};

__device__ __forceinline__ void lt_mma_k16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1, bool main_term = true) {
#if LT_MMA_EXP & 2
    if (!main_term) return;
#endif
    asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// first link of an accumulation chain: D = A*B + C with C a loop-invariant register quad (bias, or zeros)
__device__ __forceinline__ void lt_mma_k16_c(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1, const float (&c)[4]) {
    asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%12,%13};"
                 : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1), "f"(c[0]), "f"(c[1]), "f"(c[2]), "f"(c[3]));
}
__device__ __forceinline__ void lt_mma_k8_z(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
    asm("mma.sync.aligned.m16n8k8.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%7,%7,%7,%7};"
                 : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
                 : "r"(a0), "r"(a1), "r"(b0), "f"(0.0f));
}
__device__ __forceinline__ void lt_mma_k8(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
    asm("mma.sync.aligned.m16n8k8.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a0), "r"(a1), "r"(b0));
}
// (x0, x1) -> packed fp16 pairs hi, lo with x = hi + lo to 22+ bits; x0 (the lower k) goes to the low half.
// hi = the leading 11 significand bits of x (one mask), so it converts to fp16 exactly (below the fp16 normal range the
// conversion rounds at 2^-25 absolute, far under the budget); lo = rn_f16(x - hi) carries the next 11 bits (|lo| < 2^-10 |x|,
// rounded at 2^-22 |x|).  One LOP3 + one FADD per element and one F2FP per pair and part -- no fp16 -> fp32 unpack on the FMA
// pipe.  (Rounding hi to nearest instead costs one more ALU op per element and 1.8 us per launch for 2 e-6 of torque error.)
__device__ __forceinline__ void lt_split_h2(float x0, float x1, uint32_t& hi, uint32_t& lo) {
    const float h0 = __uint_as_float(__float_as_uint(x0) & 0xffffe000u), h1 = __uint_as_float(__float_as_uint(x1) & 0xffffe000u);
    const __half2 h = __floats2half2_rn(h0, h1);
    const __half2 l = __floats2half2_rn(__fsub_rn(x0, h0), __fsub_rn(x1, h1));
    hi = *reinterpret_cast<const uint32_t*>(&h);
    lo = *reinterpret_cast<const uint32_t*>(&l);
}
// softsign x / (1 + |x|): the denominator is in [1, inf), so a bare MUFU.RCP (1 ulp) needs no range scaling
#ifndef LT_MMA_EXP
#define LT_MMA_EXP 0                // timing experiments only (profiles/variants.py): 1 = no MUFU, 2 = hi*hi products only, 3 = both
#endif
__device__ __forceinline__ float lt_softsign_fast(float x) {
#if LT_MMA_EXP & 1
    return x * (1.5f - 0.5f * fabsf(x));
#endif
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fabsf(x) + 1.0f));
    return x * r;
}

struct lt_torque_raw {                 // the ten words a sample reads (the prefetch unit)
    float act, pos, vel, moff, mstr, el, ell, vl, vll, lag;
};

__device__ __forceinline__ void lt_torque_fetch(const lt_torque_mma_params& p, size_t s, size_t n, int rd_slot, lt_torque_raw& r) {
    r.act = __ldg(p.actions_raw + s);
    const float2 ds = __ldg(reinterpret_cast<const float2*>(p.dof_state) + s);
    r.pos = ds.x;
    r.vel = ds.y;
    r.moff = p.st.motor_offsets[s];
    r.mstr = p.st.motor_strengths[s];
    r.el = p.st.joint_pos_err_last[s];
    r.ell = p.st.joint_pos_err_last_last[s];
    r.vl = p.st.joint_vel_last[s];
    r.vll = p.st.joint_vel_last_last[s];
    r.lag = rd_slot >= 0 ? p.st.lag_buffer[(size_t)rd_slot * n + s] : 0.0f;
}

__global__ void __launch_bounds__(LT_MMA_THREADS, LT_MMA_MINB) lt_torque_mma_kernel(const __grid_constant__ lt_torque_mma_params p) {
    __shared__ __align__(16) uint32_t xs_hi[LT_MMA_WARPS][32 * 4], xs_lo[LT_MMA_WARPS][32 * 4];
    const lt_config& c = p.c;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    uint32_t* xh = xs_hi[warp];
    uint32_t* xl = xs_lo[warp];
    const size_t n = (size_t)c.num_envs * LT_ND;
    const int n_warps = gridDim.x * LT_MMA_WARPS;
    int tile = blockIdx.x * LT_MMA_WARPS + warp;
    if (tile >= p.n_tiles) return;

    // lag FIFO of lag_timesteps+1 slots as a ring (LR:972-974): push at the old head, read the new head
    int head = -1, rd = -1;
    if (c.randomize_lag_timesteps) {
        const int Lg = c.lag_timesteps + 1;
        head = (p.counters[2] + p.substep) % Lg;
        rd = (head + 1) % Lg;
    }
    const int rd_fetch = (rd >= 0 && rd != head) ? rd : -1;
    const bool last_sub = p.substep == c.decimation - 1;

    // the first tile's rows go out before the weight fragments so that both round trips overlap
    lt_torque_raw cur;
    {
        const size_t s = (size_t)tile * 32 + lane;
        lt_torque_fetch(p, s < n ? s : n - 1, n, rd_fetch, cur);
    }
    // weight fragments: one coalesced pass over the [regs][32] table (L2-resident, 7 KB).  (Programmatic dependent launch was
    // tried here -- fragments before griddepcontrol.wait -- and cost 0.6 us per launch; the post-physics kernels keep it.)
    uint32_t w0h[4], w0l[4], w1h[16], w1l[16];
    float bq1[4][4], w2[8];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        w0h[r] = __ldg(p.frag + (FR_W0_HI + r) * 32 + lane);
        w0l[r] = __ldg(p.frag + (FR_W0_LO + r) * 32 + lane);
        bq1[r][0] = bq1[r][2] = __uint_as_float(__ldg(p.frag + (FR_B1 + 2 * r) * 32 + lane));
        bq1[r][1] = bq1[r][3] = __uint_as_float(__ldg(p.frag + (FR_B1 + 2 * r + 1) * 32 + lane));
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) w2[r] = __uint_as_float(__ldg(p.frag + (FR_W2 + r) * 32 + lane));
#pragma unroll
    for (int r = 0; r < 16; ++r) {
        w1h[r] = __ldg(p.frag + (FR_W1_HI + r) * 32 + lane);
        w1l[r] = __ldg(p.frag + (FR_W1_LO + r) * 32 + lane);
    }
    const float b2 = __uint_as_float(__ldg(p.frag + FR_B2 * 32 + lane));


    for (; tile < p.n_tiles; tile += n_warps) {
        const size_t s_raw = (size_t)tile * 32 + lane;
        const bool valid = s_raw < n;
        const size_t s = valid ? s_raw : n - 1;
        const int dof = (int)(s % LT_ND);
        lt_torque_io io;
        io.action_raw = cur.act;
        io.dof_pos = cur.pos;
        io.dof_vel = cur.vel;
        io.motor_offset = cur.moff;
        io.motor_strength = cur.mstr;
        io.err_last = cur.el;
        io.err_last_last = cur.ell;
        io.vel_last = cur.vl;
        io.vel_last_last = cur.vll;
        io.kp_factor = io.kd_factor = 1.0f;
        io.action_scaled = lt_scaled_action(c, io.action_raw, dof, io.action_clipped);
        io.lag_out = rd_fetch >= 0 ? cur.lag : io.action_scaled;
        {   // next tile's rows: in flight while this tile's network runs
            const int nt_ = tile + n_warps;
            if (nt_ < p.n_tiles) {
                const size_t sn = (size_t)nt_ * 32 + lane;
                lt_torque_fetch(p, sn < n ? sn : n - 1, n, rd_fetch, cur);
            }
        }
        float x[LT_ACT_IN], tq = 0.0f;
        lt_torque_pre(c, dof, io, x, tq);
        if (head >= 0 && valid) p.st.lag_buffer[(size_t)head * n + s] = io.action_scaled;

        // stage the sample's 6 network inputs (zero padded to 8) as split fp16 pairs for the fragment loads
        {
            uint4 h, l;
            lt_split_h2(x[0], x[1], h.x, l.x);
            lt_split_h2(x[2], x[3], h.y, l.y);
            lt_split_h2(x[4], x[5], h.z, l.z);
            h.w = 0x00003c00u;                             // (1.0, 0): the bias input of layer 1
            l.w = 0u;
            *reinterpret_cast<uint4*>(xh + lane * 4) = h;
            *reinterpret_cast<uint4*>(xl + lane * 4) = l;
        }
        __syncwarp();
        float res[2][2];                                   // network output of rows g, g+8 of each 16-row block (partial over t)
#pragma unroll
        for (int m0 = 0; m0 < 2; m0 += LT_MMA_MB) {
            constexpr int MB = LT_MMA_MB;
            uint32_t a0h[MB], a1h[MB], a0l[MB], a1l[MB];
#pragma unroll
            for (int m = 0; m < MB; ++m) {
                a0h[m] = xh[(16 * (m0 + m) + g) * 4 + t];
                a1h[m] = xh[(16 * (m0 + m) + g + 8) * 4 + t];
                a0l[m] = xl[(16 * (m0 + m) + g) * 4 + t];
                a1l[m] = xl[(16 * (m0 + m) + g + 8) * 4 + t];
            }
            float h[MB][4][4];
#pragma unroll
            for (int nt = 0; nt < 4; ++nt)
#pragma unroll
                for (int m = 0; m < MB; ++m) lt_mma_k8_z(h[m][nt], a0l[m], a1l[m], w0h[nt]);
#pragma unroll
            for (int nt = 0; nt < 4; ++nt)
#pragma unroll
                for (int m = 0; m < MB; ++m) lt_mma_k8(h[m][nt], a0h[m], a1h[m], w0l[nt]);
#pragma unroll
            for (int nt = 0; nt < 4; ++nt)
#pragma unroll
                for (int m = 0; m < MB; ++m) lt_mma_k8(h[m][nt], a0h[m], a1h[m], w0h[nt]);
            float a2[MB][4][4];
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                // C fragments of n-tiles 2q, 2q+1 -> A fragment of k16-step q:
                //   a0 = (row g, k 2t|2t+1) a1 = (row g+8, same)   from tile 2q;   a2, a3 = the same rows, k + 8, from tile 2q+1
                uint32_t ah[MB][4], al[MB][4];
#pragma unroll
                for (int m = 0; m < MB; ++m) {
                    lt_split_h2(lt_softsign_fast(h[m][2 * q][0]), lt_softsign_fast(h[m][2 * q][1]), ah[m][0], al[m][0]);
                    lt_split_h2(lt_softsign_fast(h[m][2 * q][2]), lt_softsign_fast(h[m][2 * q][3]), ah[m][1], al[m][1]);
                    lt_split_h2(lt_softsign_fast(h[m][2 * q + 1][0]), lt_softsign_fast(h[m][2 * q + 1][1]), ah[m][2], al[m][2]);
                    lt_split_h2(lt_softsign_fast(h[m][2 * q + 1][2]), lt_softsign_fast(h[m][2 * q + 1][3]), ah[m][3], al[m][3]);
                }
#pragma unroll
                for (int nt = 0; nt < 4; ++nt) {
                    const int r = 2 * (4 * q + nt);
#pragma unroll
                    for (int m = 0; m < MB; ++m) {
                        if (q == 0) lt_mma_k16_c(a2[m][nt], al[m], w1h[r], w1h[r + 1], bq1[nt]);
                        else lt_mma_k16(a2[m][nt], al[m], w1h[r], w1h[r + 1], false);
                    }
                }
#pragma unroll
                for (int nt = 0; nt < 4; ++nt) {
                    const int r = 2 * (4 * q + nt);
#pragma unroll
                    for (int m = 0; m < MB; ++m) lt_mma_k16(a2[m][nt], ah[m], w1l[r], w1l[r + 1], false);
                }
#pragma unroll
                for (int nt = 0; nt < 4; ++nt) {
                    const int r = 2 * (4 * q + nt);
#pragma unroll
                    for (int m = 0; m < MB; ++m) lt_mma_k16(a2[m][nt], ah[m], w1h[r], w1h[r + 1]);
                }
            }
#pragma unroll
            for (int m = 0; m < MB; ++m) {
                float o0 = 0.0f, o1 = 0.0f;
#pragma unroll
                for (int nt = 0; nt < 4; ++nt) {
                    o0 = fmaf(w2[2 * nt], lt_softsign_fast(a2[m][nt][0]), o0);
                    o0 = fmaf(w2[2 * nt + 1], lt_softsign_fast(a2[m][nt][1]), o0);
                    o1 = fmaf(w2[2 * nt], lt_softsign_fast(a2[m][nt][2]), o1);
                    o1 = fmaf(w2[2 * nt + 1], lt_softsign_fast(a2[m][nt][3]), o1);
                }
                res[m0 + m][0] = o0;
                res[m0 + m][1] = o1;
            }
        }
        __syncwarp();                                      // every lane is done reading xs before the next tile overwrites it
        // sum the four column groups (t) of a row, then hand sample `lane` its value: row lane%16 of block lane/16
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                float v = res[m][q];
                v += __shfl_xor_sync(0xffffffffu, v, 1);
                v += __shfl_xor_sync(0xffffffffu, v, 2);
                res[m][q] = v;
            }
        const int src = (lane & 7) * 4;
        const float v00 = __shfl_sync(0xffffffffu, res[0][0], src), v01 = __shfl_sync(0xffffffffu, res[0][1], src);
        const float v10 = __shfl_sync(0xffffffffu, res[1][0], src), v11 = __shfl_sync(0xffffffffu, res[1][1], src);
        const float lo_blk = (lane & 8) ? v01 : v00, hi_blk = (lane & 8) ? v11 : v10;
        tq = ((lane & 16) ? hi_blk : lo_blk) + b2;

        lt_torque_post(c, dof, io, true, x, tq);
        if (valid) {
            p.st.joint_pos_err_last[s] = io.err_last;
            p.st.joint_pos_err_last_last[s] = io.err_last_last;
            p.st.joint_vel_last[s] = io.vel_last;
            p.st.joint_vel_last_last[s] = io.vel_last_last;
            if (p.substep == 0) p.st.actions[s] = io.action_clipped;
            p.st.joint_pos_target[s] = io.joint_pos_target;
            if (last_sub) {                                // LR:151-152, done here where the rows are walked coalesced anyway
                p.st.last_last_joint_pos_target[s] = p.st.last_joint_pos_target[s];
                p.st.last_joint_pos_target[s] = io.joint_pos_target;
            }
            p.st.torques[s] = io.torque;
        }
    }
}

cudaError_t lt_torque_mma_launch(const lt_config& c, const float* frag, const float* actions, const float* dof_state,
                                 const lt_state_ptrs& st, const int32_t* counters, int substep, int sm_count, cudaStream_t stream) {
    lt_torque_mma_params p;
    p.c = c;
    p.frag = reinterpret_cast<const uint32_t*>(frag);
    p.actions_raw = actions;
    p.dof_state = dof_state;
    p.st = st;
    p.counters = counters;
    p.substep = substep;
    const size_t n = (size_t)c.num_envs * LT_ND;
    p.n_tiles = (int)((n + 31) / 32);
    static int ctas_per_sm = 0;
    if (!ctas_per_sm) {
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, lt_torque_mma_kernel, LT_MMA_THREADS, 0);
        if (e != cudaSuccess) return e;
        if (ctas_per_sm < 1) ctas_per_sm = 1;
    }
    const int need = (p.n_tiles + LT_MMA_WARPS - 1) / LT_MMA_WARPS;
    const int resident = sm_count * ctas_per_sm;
    const unsigned grid = (unsigned)(need < resident ? need : resident);
    lt_torque_mma_kernel<<<grid, LT_MMA_THREADS, 0, stream>>>(p);
    return cudaGetLastError();
}
