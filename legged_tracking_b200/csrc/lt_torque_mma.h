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
