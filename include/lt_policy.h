/* lt_policy.h -- C ABI of the actor-critic / PPO half of liblt_b200.so.
 *
 * Replaces, for the hot path only, go1_gym_learn/ppo_cse/actor_critic.py:23-159 (ActorCritic: adaptation module, actor body,
 * critic body, Normal(mean, std) head) and go1_gym_learn/ppo_cse/ppo.py:65-77 (PPO.act) and :98-206 (PPO.update: one call
 * per minibatch).  AC below = actor_critic.py, PPO = ppo.py of the reference.
 *
 * All dense contractions run on the tcgen05 tensor cores as fp16-split x3 products with fp32 accumulation in tensor memory
 * (csrc/lt_gemm.cuh): fp32-equivalent arithmetic, the reference computes these GEMMs in fp32 with TF32 off.
 *
 * Conventions as in lt_b200.h: 0 or a negative lt_status; device pointers owned by the caller; asynchronous on `stream`; no
 * host synchronisation; no CPU fallback (LT_ERR_NO_DEVICE without an sm_100 GPU).  The library owns only its private
 * workspaces (operand planes of weights and activations), freed by lt_policy_destroy.
 *
 * Flat parameter layout (float32; the Python host maps state_dict keys onto it, see legged_tracking_b200/ppo.py):
 *     std[num_actions]
 *     adaptation_module: for each Linear:  weight[out, in] row-major, bias[out]           (AC:43-52)
 *     actor_body:        same                                                              (AC:57-68)
 *     critic_body:       same                                                              (AC:71-82)
 * With height_map_size > 0 (go1_gym_learn/ppo_cse_cnn, CNN: = ppo_cse_cnn/actor_critic.py) the two Linear layers of
 * HeightMapEncoder.model (CNN:45-52) sit between std and the adaptation module:
 *     std | height_map_encoder.model.0 | height_map_encoder.model.2 | adaptation_module | actor_body | critic_body
 */
#ifndef LT_POLICY_H
#define LT_POLICY_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LT_POLICY_MAX_HIDDEN 4

/* AC_Args.activation through get_activation (AC:162-178); 'crelu' maps to ReLU there */
#define LT_ACT_ELU 0
#define LT_ACT_SELU 1
#define LT_ACT_RELU 2
#define LT_ACT_LRELU 3
#define LT_ACT_TANH 4
#define LT_ACT_SIGMOID 5

typedef struct lt_policy_config {
    int32_t struct_bytes;                 /* sizeof(lt_policy_config): ABI check */
    int32_t num_obs_history;              /* H * num_obs: width of the observation history (AC:43, 57) */
    int32_t num_privileged_obs;           /* width of the privileged observation = adaptation target (AC:52) */
    int32_t num_actions;
    int32_t n_adaptation_hidden, adaptation_hidden[LT_POLICY_MAX_HIDDEN];     /* AC_Args.adaptation_module_branch_hidden_dims */
    int32_t n_actor_hidden, actor_hidden[LT_POLICY_MAX_HIDDEN];               /* AC_Args.actor_hidden_dims */
    int32_t n_critic_hidden, critic_hidden[LT_POLICY_MAX_HIDDEN];             /* AC_Args.critic_hidden_dims */
    int32_t max_rows;                     /* row capacity of the workspaces: >= envs per act() and >= minibatch size */
    int32_t training;                     /* 0: inference only (no gradient workspaces) */
    int32_t activation;                   /* LT_ACT_*: between the Linear layers of all three bodies (AC:40) */
    /* ppo_cse_cnn, MLP form of HeightMapEncoder with recurrent_latent_embedding = Identity
     * (AC_Args.use_cnn = use_gru = False, the defaults, CNN:19-21).  height_map_size = 0 selects ppo_cse and the three fields
     * after it are ignored.  With the encoder (CNN:181-198, process_obs_history):
     *     o   = last num_obs row of obs_history;   h = its last height_map_size columns,  s = the columns before them
     *     emb = act(W2 act(W1 h + b1) + b2)                                    (CNN:45-52: Linear, act, Linear, act)
     *     policy_input = [s, s, emb]      (s twice: once from lstm_input's last row, once from the explicit cat, CNN:194-197)
     * and policy_input takes the place of obs_history in the three bodies.  Only the last history row is read: the rows
     * before it feed the Identity 'recurrent' module and are dropped by recurrent_latent[:, -1, :]. */
    int32_t height_map_size;              /* prod(AC_Args.height_map_shape); 0 = no encoder */
    int32_t num_obs;                      /* width of one observation row: num_obs_history = H * num_obs */
    int32_t encoder_hidden;               /* 256 (CNN:47) */
    int32_t encoder_embedding;            /* AC_Args.cnn_num_embedding */
} lt_policy_config;

/* PPO_Args (PPO:13-30) + torch.optim.Adam defaults */
typedef struct lt_ppo_hparams {
    float clip_param, value_loss_coef, entropy_coef, max_grad_norm, desired_kl;
    float adam_beta1, adam_beta2, adam_eps;
    float lr_min, lr_max, lr_factor;      /* adaptive schedule bounds 1e-5, 1e-2 and factor 1.5 (PPO:118-121) */
    int32_t use_clipped_value_loss;
    int32_t adaptive_schedule;            /* desired_kl != None and schedule == 'adaptive' */
} lt_ppo_hparams;

/* Flat [T*N, w] views of RolloutStorage (rollout_storage.py:100-139): a minibatch is the rows `indices[0..count)` of these. */
typedef struct lt_ppo_batch_ptrs {
    const float *obs_history, *privileged_obs, *actions, *values, *returns, *advantages, *old_actions_log_prob, *old_mu, *old_sigma;
} lt_ppo_batch_ptrs;

/* Device float[16] written by lt_ppo_minibatch_grad / lt_adaptation_grad (sums over the rows of the minibatch; the host
 * all-reduces this array together with the gradient when the job is env-sharded over several GPUs):
 *   [0] sum surrogate loss  [1] sum value loss  [2] sum KL  [3] row count  [4] sum adaptation train loss terms
 *   [5] train element count  [6] sum adaptation test loss terms  [7] test element count */
#define LT_PPO_STATS 16

typedef struct lt_policy lt_policy;

int lt_policy_config_bytes(void);
int lt_policy_create(const lt_policy_config* cfg, lt_policy** out);                 /* AC:23-88 (__init__) */
void lt_policy_destroy(lt_policy* p);
const char* lt_policy_last_error(const lt_policy* p);                              /* p may be NULL: error of the last failed create */
int64_t lt_policy_num_params(const lt_policy* p);

/* params / grads: device float[lt_policy_num_params], caller-owned (the host keeps state_dict views into params and
 * all-reduces grads).  lt_policy_refresh re-derives the tensor-core operand planes from params: call it after every
 * external write to params (load_state_dict, broadcast); the apply calls below do it themselves. */
int lt_policy_bind_params(lt_policy* p, float* params, float* grads);
int lt_policy_refresh(lt_policy* p, void* stream);

/* ActorCritic.act_student / evaluate / get_student_latent (AC:146-161) for `rows` samples:
 *   latent = adaptation_module(obs_history); action_mean = actor_body(cat(obs_history, latent));
 *   value = critic_body(cat(obs_history, privileged_obs)).
 * Any output may be NULL (its sub-network is then skipped where nothing else needs it).  privileged_obs may be NULL when
 * value is NULL.  teacher_latent != 0: act_teacher (AC:152-155) -- the actor reads privileged_obs instead of the latent. */
int lt_policy_forward(lt_policy* p, const float* obs_history, const float* privileged_obs, int32_t rows, int32_t teacher_latent,
                      float* latent, float* action_mean, float* value, void* stream);

/* PPO.act (PPO:65-77) fused: forward, Normal(mean, std) sample, log-prob, value.  normal_noise: [rows, num_actions] standard
 * normal draws or NULL -> in-kernel Philox4x32-10 + Box-Muller keyed (seed, step, row, action).  actions_in != NULL replays
 * given actions instead of sampling (parity tests).  Outputs: actions, action_mean, action_sigma [rows, A]; log_prob, value [rows];
 * any of them may be NULL.  value == NULL skips the critic (ActorCritic.act alone, AC:133-135) and privileged_obs may then be NULL. */
int lt_policy_act(lt_policy* p, const float* obs_history, const float* privileged_obs, int32_t rows, const float* normal_noise,
                  const float* actions_in, uint64_t seed, uint64_t step, float* actions, float* actions_log_prob, float* action_mean,
                  float* action_sigma, float* value, void* stream);

/* One minibatch of PPO.update, first half (PPO:108-146): gather rows `indices` (DEVICE int64[count]; NULL = rows 0..count),
 * forward, surrogate / clipped value / entropy loss, KL, full backward into `grads` (mean over `count` rows, i.e. the
 * gradient loss.backward() leaves), sums into stats.  grad_world > 1 divides the loss by count * grad_world instead (every
 * rank then SUM-all-reduces grads and stats). */
int lt_ppo_minibatch_grad(lt_policy* p, const lt_ppo_batch_ptrs* batch, const int64_t* indices, int32_t count,
                          const lt_ppo_hparams* hp, int32_t grad_world, float* stats, void* stream);

/* Second half (PPO:111-125, 147-148): adaptive-KL learning-rate decision from stats (kl mean = stats[2] / stats[3]),
 * clip_grad_norm_ over all parameters, Adam step number `adam_step` (1-based) with state adam_m / adam_v (device
 * float[num_params], caller-owned), operand planes refreshed.  lr: DEVICE float[1], read and updated in place.
 * grad_norm_out: DEVICE float[1] or NULL, receives the pre-clip total norm. */
int lt_ppo_minibatch_apply(lt_policy* p, const lt_ppo_hparams* hp, const float* stats, float* lr, float* adam_m, float* adam_v,
                           int32_t adam_step, float* grad_norm_out, void* stream);

/* Adaptation-module regression step (PPO:150-176), first half: pred = adaptation_module(obs_history[indices]);
 * loss = mse(pred[:num_train], privileged_obs[:num_train]) back-propagated into the adaptation slice of grads (the rest of
 * grads is left untouched); test loss on rows num_train..count.  Sums into stats[4..8).
 * When batch->obs_history, indices (non-NULL) and count are the ones of the immediately preceding lt_ppo_minibatch_grad on this
 * handle -- the reference's case, PPO:150-176 runs on the minibatch just used -- the rows gathered by that call are reused:
 * do not rewrite the index array or the rollout buffers between the two calls. */
int lt_adaptation_grad(lt_policy* p, const lt_ppo_batch_ptrs* batch, const int64_t* indices, int32_t count, int32_t num_train,
                       int32_t grad_world, float* stats, void* stream);
/* Second half: Adam step (no clipping) on the adaptation slice with its own state (same flat indexing as params), lr host value. */
int lt_adaptation_apply(lt_policy* p, const lt_ppo_hparams* hp, float lr, float* adam_m, float* adam_v, int32_t adam_step, void* stream);
/* Diagnostic: DEVICE float[2][4 * (LT_POLICY_MAX_HIDDEN + 1)], the per-tensor power-of-two scales the backward passes of
 * lt_ppo_minibatch_grad ([0]) and lt_adaptation_grad ([1]) currently use for the gradient operand planes (rows: adaptation
 * module, actor, critic, height-map encoder; columns: layer).  See "gradient scales" in csrc/lt_policy.cu. */
const float* lt_policy_grad_scales(const lt_policy* p);
/* [lo, hi) of what the adaptation optimiser steps inside the flat parameter vector: the adaptation module, preceded by the
 * height-map encoder when there is one (with ppo_cse_cnn the regression loss reaches the encoder too, CNN PPO:170-186) */
int lt_policy_adaptation_slice(const lt_policy* p, int64_t* lo, int64_t* hi);

/* The contraction kernel on its own (tests, roofline measurement): out[M, N] = a[M, K] * b[N, K]^T (+ bias), fp32 in / out on
 * the device; operands are split into fp16 planes internally with the scales the policy uses. */
int lt_linear_f32(const float* a, const float* b, const float* bias, float* out, int32_t M, int32_t N, int32_t K, int32_t splits,
                  void* stream);

/* Same kernel with both operands MN-major, as the weight-gradient products use it: out[M, N] = a[K, M]^T * b[K, N] (the
 * contraction runs over the ROWS of the two stored matrices), summed over `splits` split-K partials in a fixed order. */
int lt_linear_tn_f32(const float* a, const float* b, float* out, int32_t M, int32_t N, int32_t K, int32_t splits, void* stream);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same contraction as lt_linear_f32 with splits = 1 --
 * out[M, N] = a[M, K] * b[N, K]^T (+ bias), fp32 in / out on the device, same kernel, same operand scales, bit-identical
 * result -- but the call allocates NOTHING (no cudaMalloc / cudaFree / allocating library routine): the fp16 operand planes
 * live in a static device array of LT_LINEAR_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit.
 *   - limit: 4 * (M + N) * roundup(K, 8) bytes (each operand rounded up to 256) must fit in the array; larger shapes are
 *     rejected with LT_ERR_INVALID and a message in lt_policy_last_error(NULL) -- use lt_linear_f32 for those;
 *   - asynchronous on `stream`, no host synchronisation (lt_linear_f32 synchronises because it frees its buffers);
 *   - ALL calls on a device share the one array and it keeps the last call's contents (the call clears what it uses first):
 *     two calls that overlap on the GPU (different streams, or graph branches) race -- serialise them. */
#define LT_LINEAR_STATIC_SCRATCH_BYTES 8388608
int lt_linear_f32_static(const float* a, const float* b, const float* bias, float* out, int32_t M, int32_t N, int32_t K, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* LT_POLICY_H */
