// lt_torque_tc5.cu -- _compute_torques (LR:957-996) with the actuator network (LR:1307-1322) on the 5th-generation tensor
// cores: tcgen05.mma with the accumulators AND the A operand in tensor memory (TMEM).
//
// One CTA (4 warps) owns tiles of 128 consecutive (env, dof) samples; thread t <-> sample row t <-> TMEM lane t, so nothing
// is ever exchanged between threads: a thread writes its own operand row with tcgen05.st, reads its own accumulator row
// with tcgen05.ld, and ends up holding its own sample's torque (the mma.sync kernel needs shared-memory staging, fragment
// bookkeeping and three shuffle rounds for that).  Same numerics as lt_torque_mma.cu: every operand split in two fp16 parts,
// a*w = a_lo*w_hi + a_hi*w_lo + a_hi*w_hi accumulated in fp32 (kind::f16, D in F32).
//   layer 1   D1[128x32] = A1[128x16] * W0^T   K = 16: 6 inputs, a constant 1 that carries the bias, zero padding
//   layer 2   D2[128x32] = A2[128x32] * W1^T   two K16 steps + one K16 step of a constant-one operand block against the bias
//   layer 3   32 FMAs per thread on its own row
// The weights (B operands, K-major, no swizzle: 8x8 core matrices of 128 B) are built on the host in the canonical layout
// and copied to shared memory once per CTA.  TMEM: 128 columns per CTA (A1 16, D1 32, A2 32, D2 32, constant-one block 8).
// One elected thread issues the MMAs; completion comes back through an mbarrier (tcgen05.commit).
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "lt_core.cuh"
#include "lt_torque_mma.h"

#define TC5_THREADS 128
#define TC5_COLS 128                 // TMEM columns allocated per CTA
#define TC5_A1 0                     // column offsets inside the allocation
#define TC5_A1_LO 8
#define TC5_D1 16
#define TC5_A2 48
#define TC5_A2_LO 64
#define TC5_D2 80
#define TC5_ONE 112                  // K16 operand block whose k = 0 is the constant 1: carries the layer-2 bias through the MMA
// shared-memory weight blob: eight [32 x 16] fp16 B blocks of 1 KB: W0hi W0lo W1hi(k0) W1hi(k1) W1lo(k0) W1lo(k1) b1hi b1lo
#define TC5_BLOB_BYTES 8192

static uint16_t tc5_h(float v) {
    const __half h = __float2half_rn(v);
    uint16_t u;
    memcpy(&u, &h, 2);
    return u;
}
static float tc5_res(float v) { return v - __half2float(__float2half_rn(v)); }

// canonical K-major no-swizzle layout of a [32 (n) x 16 (k)] block: core matrix (n / 8, k / 8) of 8 rows x 16 bytes
static inline int tc5_off(int n, int k) { return ((n >> 3) * 2 + (k >> 3)) * 128 + (n & 7) * 16 + (k & 7) * 2; }

void lt_torque_tc5_build_blob(const lt_act_weights& W, uint8_t* blob) {
    memset(blob, 0, TC5_BLOB_BYTES);
    auto put = [&](int block, int n, int k, float v, bool lo) {
        const uint16_t u = tc5_h(lo ? tc5_res(v) : v);
        memcpy(blob + block * 1024 + tc5_off(n, k), &u, 2);
    };
    for (int n = 0; n < 32; ++n) {
        for (int k = 0; k < LT_ACT_IN; ++k) { put(0, n, k, W.w0[n * LT_ACT_IN + k], false); put(1, n, k, W.w0[n * LT_ACT_IN + k], true); }
        put(0, n, LT_ACT_IN, W.b0[n], false);
        put(1, n, LT_ACT_IN, W.b0[n], true);
        for (int k = 0; k < 32; ++k) {
            put(2 + (k >> 4), n, k & 15, W.w1[n * 32 + k], false);
            put(4 + (k >> 4), n, k & 15, W.w1[n * 32 + k], true);
        }
        put(6, n, 0, W.b1[n], false);
        put(7, n, 0, W.b1[n], true);
    }
}

struct lt_tc5_tail { float w2[32], b2, pad_[3]; };

struct lt_torque_tc5_params {
    lt_config c;
    lt_tc5_tail T;
    const uint8_t* blob;
    const float* actions_raw;
    const float* dof_state;
    lt_state_ptrs st;
    const int32_t* counters;
    int32_t substep;
    int32_t n_tiles;
    int32_t swap_lbo_sbo;
};

// ---- PTX wrappers
__device__ __forceinline__ uint32_t tc5_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void tc5_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc5_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc5_mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tc5_smem(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void tc5_mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(tc5_smem(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tc5_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(tc5_smem(bar)) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem descriptor]
__device__ __forceinline__ void tc5_mma(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
#if TC5_EXP & 2
    return;
#endif
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tc5_st8(uint32_t taddr, const uint32_t (&v)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]),
                 "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
                 : "memory");
}
__device__ __forceinline__ void tc5_st16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
        "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]),
        "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
        : "memory");
}
__device__ __forceinline__ void tc5_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc5_ld32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
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
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
// shared-memory matrix descriptor, K-major, no swizzle: start address, leading / stride byte offsets (all >> 4), version 1
__device__ __forceinline__ uint64_t tc5_desc(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((smem_addr & 0x3ffffu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}
__device__ __forceinline__ void tc5_split(float x0, float x1, uint32_t& hi, uint32_t& lo) {      // as lt_torque_mma.cu
#if TC5_EXP & 1
    hi = __float_as_uint(x0) ^ __float_as_uint(x1);
    lo = 0u;
    return;
#endif
    const float h0 = __uint_as_float(__float_as_uint(x0) & 0xffffe000u), h1 = __uint_as_float(__float_as_uint(x1) & 0xffffe000u);
    const __half2 h = __floats2half2_rn(h0, h1);
    const __half2 l = __floats2half2_rn(__fsub_rn(x0, h0), __fsub_rn(x1, h1));
    hi = *reinterpret_cast<const uint32_t*>(&h);
    lo = *reinterpret_cast<const uint32_t*>(&l);
}
#ifndef TC5_EXP
#define TC5_EXP 0                   // timing ablations only (profiles/variants.py): 1 = no softsign / split arithmetic, 2 = no MMAs
#endif
__device__ __forceinline__ float tc5_softsign(float x) {
#if TC5_EXP & 1
    return x;
#endif
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fabsf(x) + 1.0f));
    return x * r;
}

#ifdef TC5_TRACE
// timing study only (profiles/tc5_trace.py, profiles/r02_tc5.md): %globaltimer stamps of each CTA's phases
__device__ unsigned long long g_tc5_trace[2048 * 8];
__device__ __forceinline__ unsigned long long tc5_now() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define TC5_STAMP(k) do { if (threadIdx.x == 0 && blockIdx.x < 2048) g_tc5_trace[blockIdx.x * 8 + (k)] = tc5_now(); } while (0)
extern "C" int lt_tc5_trace_read(unsigned long long* out) { return (int)cudaMemcpyFromSymbol(out, g_tc5_trace, sizeof(unsigned long long) * 2048 * 8); }
#else
#define TC5_STAMP(k) do {} while (0)
#endif

struct tc5_raw {                       // the ten words a sample reads (the prefetch unit)
    float act, pos, vel, moff, mstr, el, ell, vl, vll, lag;
};
__device__ __forceinline__ void tc5_fetch(const lt_torque_tc5_params& p, uint32_t s, uint32_t n, int rd_slot, tc5_raw& r) {
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

__global__ void __launch_bounds__(TC5_THREADS, 4) lt_torque_tc5_kernel(const __grid_constant__ lt_torque_tc5_params p) {
    __shared__ __align__(128) uint8_t s_blob[TC5_BLOB_BYTES];
    __shared__ __align__(8) uint64_t s_bar[2];
    __shared__ uint32_t s_tmem;
    const lt_config& c = p.c;
    const int tid = threadIdx.x, warp = tid >> 5;
    const uint32_t n = (uint32_t)c.num_envs * LT_ND;
    TC5_STAMP(0);

    // one-time setup: weights to shared memory, mbarriers, TMEM allocation (warp 0)
    for (int k = tid; k < TC5_BLOB_BYTES / 16; k += TC5_THREADS)
        reinterpret_cast<uint4*>(s_blob)[k] = __ldg(reinterpret_cast<const uint4*>(p.blob) + k);
    if (tid == 0) {
        tc5_mbar_init(&s_bar[0], 1);
        tc5_mbar_init(&s_bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tc5_smem(&s_tmem)), "n"(TC5_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");          // the blob was written by generic stores, the MMA reads it through the async proxy
    tc5_fence_before();
    __syncthreads();
    tc5_fence_after();
    TC5_STAMP(1);
    const uint32_t tmem = s_tmem;
    const uint32_t lane_addr = tmem + ((uint32_t)(warp * 32) << 16);      // this warp's 32 TMEM lanes

    // instruction descriptor: D = F32, A = B = F16, both K-major, N = 32, M = 128
    const uint32_t idesc = (1u << 4) | ((32u >> 3) << 17) | ((128u >> 4) << 24);
    const uint32_t blob_addr = tc5_smem(s_blob);
    const uint32_t lbo = p.swap_lbo_sbo ? 256u : 128u, sbo = p.swap_lbo_sbo ? 128u : 256u;
    uint64_t bdesc[8];
#pragma unroll
    for (int b = 0; b < 8; ++b) bdesc[b] = tc5_desc(blob_addr + b * 1024, lbo, sbo);
    {   // the constant-one operand block is tile-invariant: written once
        const uint32_t one[8] = {0x00003c00u, 0, 0, 0, 0, 0, 0, 0};
        tc5_st8(lane_addr + TC5_ONE, one);
        tc5_wait_st();
    }

    int head = -1, rd = -1;
    if (c.randomize_lag_timesteps) {
        const int Lg = c.lag_timesteps + 1;
        head = (p.counters[2] + p.substep) % Lg;
        rd = (head + 1) % Lg;
    }
    const int rd_fetch = (rd >= 0 && rd != head) ? rd : -1;
    const bool last_sub = p.substep == c.decimation - 1;
    uint32_t phase = 0;
    tc5_raw cur;
    {
        const uint32_t s0 = (uint32_t)blockIdx.x * TC5_THREADS + tid;
        tc5_fetch(p, s0 < n ? s0 : n - 1, n, rd_fetch, cur);
    }

    TC5_STAMP(2);
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        if (tile == (int)blockIdx.x + (int)gridDim.x) TC5_STAMP(4);
        const uint32_t s_raw = (uint32_t)tile * TC5_THREADS + tid;
        const bool valid = s_raw < n;
        const uint32_t s = valid ? s_raw : n - 1;
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
        {   // next tile's rows: in flight while this tile's two MMA round trips run
            const int nt_ = tile + (int)gridDim.x;
            if (nt_ < p.n_tiles) {
                const uint32_t sn = (uint32_t)nt_ * TC5_THREADS + tid;
                tc5_fetch(p, sn < n ? sn : n - 1, n, rd_fetch, cur);
            }
        }
        float x[LT_ACT_IN], tq = 0.0f;
        lt_torque_pre(c, dof, io, x, tq);
        if (tile == (int)blockIdx.x) TC5_STAMP(3);
        if (head >= 0 && valid) p.st.lag_buffer[(size_t)head * n + s] = io.action_scaled;

        // ---- layer 1 operand row: k = 0..5 inputs, k = 6 the constant 1 (bias), rest zero; hi and lo parts
        {
            uint32_t ah[8] = {0, 0, 0, 0, 0, 0, 0, 0}, al[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            tc5_split(x[0], x[1], ah[0], al[0]);
            tc5_split(x[2], x[3], ah[1], al[1]);
            tc5_split(x[4], x[5], ah[2], al[2]);
            ah[3] = 0x00003c00u;
            tc5_st8(lane_addr + TC5_A1, ah);
            tc5_st8(lane_addr + TC5_A1_LO, al);
            tc5_wait_st();
        }
        tc5_fence_before();
        __syncthreads();
        if (tid == 0) {
            tc5_fence_after();
            tc5_mma(tmem + TC5_D1, tmem + TC5_A1_LO, bdesc[0], idesc, 0u);
            tc5_mma(tmem + TC5_D1, tmem + TC5_A1, bdesc[1], idesc, 1u);
            tc5_mma(tmem + TC5_D1, tmem + TC5_A1, bdesc[0], idesc, 1u);
            tc5_commit(&s_bar[0]);
        }
        tc5_mbar_wait(&s_bar[0], phase);
        tc5_fence_after();

        // ---- layer 1 -> layer 2: read the row, softsign, split, write the operand row
        {
            float h[32];
            tc5_ld32(lane_addr + TC5_D1, h);
            uint32_t ah[16], al[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) tc5_split(tc5_softsign(h[2 * j]), tc5_softsign(h[2 * j + 1]), ah[j], al[j]);
            tc5_st16(lane_addr + TC5_A2, ah);
            tc5_st16(lane_addr + TC5_A2_LO, al);
            tc5_wait_st();
        }
        tc5_fence_before();
        __syncthreads();
        if (tid == 0) {
            tc5_fence_after();
            tc5_mma(tmem + TC5_D2, tmem + TC5_ONE, bdesc[7], idesc, 0u);
            tc5_mma(tmem + TC5_D2, tmem + TC5_ONE, bdesc[6], idesc, 1u);
            tc5_mma(tmem + TC5_D2, tmem + TC5_A2_LO, bdesc[2], idesc, 1u);
            tc5_mma(tmem + TC5_D2, tmem + TC5_A2_LO + 8, bdesc[3], idesc, 1u);
            tc5_mma(tmem + TC5_D2, tmem + TC5_A2, bdesc[4], idesc, 1u);
            tc5_mma(tmem + TC5_D2, tmem + TC5_A2 + 8, bdesc[5], idesc, 1u);
            tc5_mma(tmem + TC5_D2, tmem + TC5_A2, bdesc[2], idesc, 1u);
            tc5_mma(tmem + TC5_D2, tmem + TC5_A2 + 8, bdesc[3], idesc, 1u);
            tc5_commit(&s_bar[1]);
        }
        tc5_mbar_wait(&s_bar[1], phase);
        tc5_fence_after();
        phase ^= 1u;

        // ---- softsign, layer 3 on the thread's own row
        {
            float h[32];
            tc5_ld32(lane_addr + TC5_D2, h);
            float o0 = p.T.b2, o1 = 0.0f;
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
                o0 = fmaf(p.T.w2[j], tc5_softsign(h[j]), o0);
                o1 = fmaf(p.T.w2[j + 1], tc5_softsign(h[j + 1]), o1);
            }
            tq = o0 + o1;
        }
        lt_torque_post(c, dof, io, true, x, tq);
        if (valid) {
            p.st.joint_pos_err_last[s] = io.err_last;
            p.st.joint_pos_err_last_last[s] = io.err_last_last;
            p.st.joint_vel_last[s] = io.vel_last;
            p.st.joint_vel_last_last[s] = io.vel_last_last;
            if (p.substep == 0) p.st.actions[s] = io.action_clipped;
            p.st.joint_pos_target[s] = io.joint_pos_target;
            if (last_sub) {
                p.st.last_last_joint_pos_target[s] = p.st.last_joint_pos_target[s];
                p.st.last_joint_pos_target[s] = io.joint_pos_target;
            }
            p.st.torques[s] = io.torque;
        }
        // the next tile's tcgen05.st must not overtake this tile's tcgen05.ld of other warps' lanes: lanes are private, the MMA
        // of the next tile is issued only after the __syncthreads that follows every warp's loads -- nothing more is needed
    }
    TC5_STAMP(5);
    tc5_fence_before();
    __syncthreads();
    TC5_STAMP(6);
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TC5_COLS) : "memory");
}

cudaError_t lt_torque_tc5_launch(const lt_config& c, const lt_act_weights& W, const uint8_t* blob, const float* actions,
                                 const float* dof_state, const lt_state_ptrs& st, const int32_t* counters, int substep, int sm_count,
                                 int swap_lbo_sbo, cudaStream_t stream) {
    lt_torque_tc5_params p;
    p.c = c;
    memcpy(p.T.w2, W.w2, sizeof p.T.w2);
    p.T.b2 = W.b2[0];
    p.blob = blob;
    p.actions_raw = actions;
    p.dof_state = dof_state;
    p.st = st;
    p.counters = counters;
    p.substep = substep;
    p.swap_lbo_sbo = swap_lbo_sbo;
    const size_t n = (size_t)c.num_envs * LT_ND;
    if (n >= (1ull << 31)) return cudaErrorInvalidValue;
    p.n_tiles = (int)((n + TC5_THREADS - 1) / TC5_THREADS);
    const int resident = sm_count * 4;
    const unsigned grid = (unsigned)(p.n_tiles < resident ? p.n_tiles : resident);
    lt_torque_tc5_kernel<<<grid, TC5_THREADS, 0, stream>>>(p);
    return cudaGetLastError();
}
