// lt_torque_mma.h -- internal interface between lt_kernels.cu (C ABI, handle) and lt_torque_mma.cu (tensor-pipe torque kernel).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/lt_b200.h"

#define LT_MMA_FRAG_REGS 57                 /* per-lane registers of the pre-split weight fragment table */
#define LT_HIDDEN_SYM __attribute__((visibility("hidden")))

struct lt_act_weights;
// host: fills out[LT_MMA_FRAG_REGS * 32] (register-major, lane-minor) from the fp32 weights
LT_HIDDEN_SYM void lt_torque_mma_build_frags(const lt_act_weights& W, float* out);
// one sub-step of _compute_torques with control_type == actuator_net; asynchronous on `stream`
LT_HIDDEN_SYM cudaError_t lt_torque_mma_launch(const lt_config& c, const float* frag, const float* actions, const float* dof_state,
                                               const lt_state_ptrs& st, const int32_t* counters, int substep, int sm_count,
                                               cudaStream_t stream);

// ---- tcgen05 variant (lt_torque_tc5.cu): accumulators and the A operand in tensor memory, weights (B) in shared memory
#define LT_TC5_BLOB_BYTES 8192
LT_HIDDEN_SYM void lt_torque_tc5_build_blob(const lt_act_weights& W, uint8_t* blob);
LT_HIDDEN_SYM cudaError_t lt_torque_tc5_launch(const lt_config& c, const lt_act_weights& W, const uint8_t* blob, const float* actions,
                                               const float* dof_state, const lt_state_ptrs& st, const int32_t* counters, int substep,
                                               int sm_count, int swap_lbo_sbo, cudaStream_t stream);

// ---- programmatic dependent launch (PDL): the next kernel of the post-physics chain may be launched while the previous one
// drains; every kernel of the chain calls lt_pdl_wait() before its first access to memory a predecessor writes, so stream
// order is kept.  OFF by default, LT_PDL=1 in the environment turns it on.  Measured on a B200 (profiles/r01c_pdl.md): it
// shortens the eagerly launched post-physics chain at small env counts (46.3 -> 42.2 us at 8,192 envs) but inside the
// captured CUDA graph of a full step it gains nothing at 8,192 envs and loses 4 us at 65,536 (early-resident dependent
// CTAs); without the attribute the griddepcontrol instructions are no-ops.
#if defined(__CUDACC__)
__device__ __forceinline__ void lt_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void lt_pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif
#include <stdlib.h>
static inline bool lt_pdl_enabled() {
    static int on = -1;
    if (on < 0) { const char* e = getenv("LT_PDL"); on = (e && e[0] == '1') ? 1 : 0; }
    return on != 0;
}
template <typename... KArgs, typename... Args>
static inline cudaError_t lt_launch_chain(void (*kernel)(KArgs...), unsigned grid, unsigned block, size_t smem, cudaStream_t stream,
                                          Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = lt_pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
