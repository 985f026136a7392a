/* ocppo.h -- C ABI of libocppo.so: the B200 (sm_100a) PPO_OBJ learner hot path.
 *
 * The reference (BluemlJ/oc_cleanrl) is 100% Python and has no FFI for this
 * path; every entry point below replaces a span of eager-torch code and cites
 * it (paths relative to the reference root).  INTEGRATION.md shows the ctypes
 * binding a reference maintainer would add.
 *
 * Conventions
 *  - plain C types only; all data pointers are DEVICE pointers owned by the
 *    caller (torch tensors), fp32 unless stated; `stream` is a cudaStream_t
 *    passed as void*; calls are asynchronous on that stream and never
 *    synchronise, allocate or start threads (so they can be captured into a
 *    CUDA graph).  Every call makes the device of its handle (or of its first
 *    data pointer) current for its own duration and restores the caller's.
 *  - every function returns 0 on success or a negative OCPPO_E* code;
 *    ocppo_last_error() returns a thread-local message.  Nothing throws.
 *  - there is no CPU fallback: a device that is not sm_100 is an error.
 */
#ifndef OCPPO_H
#define OCPPO_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OCPPO_OK 0
#define OCPPO_EINVAL (-1)   /* bad argument / unsupported shape          */
#define OCPPO_ECUDA (-2)    /* CUDA runtime error (see last_error)       */
#define OCPPO_EARCH (-3)    /* device is not compute capability 10.x     */
#define OCPPO_ESTATE (-4)   /* arena / workspace not bound               */

#define OCPPO_MAX_LAYERS 8
#define OCPPO_MAX_ACTIONS 31
#define OCPPO_NUM_STATS 8

typedef struct ocppo_handle ocppo_handle;

/* Shape of PPObj: cleanrl/architectures/ppo.py:61-84.  observation_space.shape
 * = (frames, obs_dim); encoder acts on the last dim of (B, frames, obs_dim),
 * Flatten, decoder, then actor (n_actions) and critic (1) heads. */
typedef struct ocppo_config {
    int32_t frames;
    int32_t obs_dim;
    int32_t n_actions;
    int32_t n_encoder;
    int32_t encoder_dims[OCPPO_MAX_LAYERS];
    int32_t n_decoder;
    int32_t decoder_dims[OCPPO_MAX_LAYERS];
    int32_t device;          /* CUDA ordinal */
    int32_t reserved;
} ocppo_config;

/* Hyper-parameters of one minibatch: cleanrl/ppo_atari_oc.py:127-151. */
typedef struct ocppo_hparams {
    double clip_coef;        /* python floats: kept in double, rounded per op */
    double ent_coef;
    double vf_coef;
    int32_t norm_adv;        /* :577-579 */
    int32_t clip_vloss;      /* :586-597 */
} ocppo_hparams;

/* stats[] slots written by ocppo_minibatch_fwd_bwd (:573-575,:599-602) */
enum {
    OCPPO_STAT_PG_LOSS = 0,
    OCPPO_STAT_V_LOSS = 1,
    OCPPO_STAT_ENTROPY = 2,
    OCPPO_STAT_LOSS = 3,
    OCPPO_STAT_OLD_APPROX_KL = 4,
    OCPPO_STAT_APPROX_KL = 5,
    OCPPO_STAT_CLIPFRAC = 6,
    OCPPO_STAT_RESERVED = 7
};

const char *ocppo_last_error(void);
int ocppo_version(void);
/* 1 if the library was compiled with -DOCPPO_BRINGUP (the bisecting knobs OCPPO_TC_DBG / OCPPO_TC_OPS / OCPPO_TC_TRUNC /
 * OCPPO_NO_MAPCACHE are then read from the environment), 0 for the release build, which has none of them. */
int ocppo_bringup_build(void);

/* ---- launch accounting / live kernel timing (measurement only) ---------- */
/* Kernel classes: 0 GAE, 1 trunk GEMM fwd (tcgen05), 2 trunk GEMM dgrad, 3 trunk
 * GEMM wgrad, 4 heads, 5 aux (gather/reduce/stats/pair passes), 6 clip+Adam,
 * 7 CUDA-core contraction kernels (first-layer FMA GEMM, skinny weight
 * gradients, every shape the tensor-core tile rejects).  ocppo_launch_count()
 * = kernels launched by this library since load.  With profiling enabled every
 * launch is bracketed by CUDA events on its own stream; the event ring is
 * drained into per-class sums whenever it fills, so NO launch is dropped:
 * ocppo_profile_read() returns the summed milliseconds and the number of
 * kernels timed per class (8 slots) since the last read / enable, then clears
 * them.  Launches made while their stream is being captured into a CUDA graph
 * cannot carry events; ocppo_launch_counts() reports them as `untimed`. */
#define OCPPO_NUM_KERNEL_CLASSES 8
int64_t ocppo_launch_count(void);
int ocppo_launch_counts(int64_t *launches_by_class, int64_t *untimed_by_class, int32_t n_classes);
int ocppo_profile_enable(int32_t on);
int ocppo_profile_read(double *ms_by_class, int64_t *timed_by_class, int32_t n_classes);
/* Synchronous micro-benchmark: achievable fp32 FMA TFLOP/s of the current device
 * (the roofline denominator of the CUDA-core GEMMs; measurement only). */
int ocppo_measure_fma_peak(double *tflops, void *stream);

/* ---- GEMM engine selection + raw GEMM (tests / measurement) -------------- */
/* The trunk GEMMs (nn.Linear forward, dgrad, wgrad of architectures/ppo.py:70-84) run on one of eight engines
 * (process-wide switch; shapes outside the 256x256x32 CTA-pair tile grid always use engine 0):
 *   0  fp32 CUDA-core FMA kernels (exact fp32 products);
 *   1  tcgen05 tensor cores, 3xTF32: x = hi + lo, hi.hi + hi.lo + lo.hi, fp32 accumulate in TMEM, main and correction
 *      products in separate accumulators (the TMEM accumulator rounds toward zero; the epilogue undoes the expected loss);
 *   2  as 1 with a single, double-buffered accumulator per tile: ~10 % faster, ~3x the rounding error;
 *   3  as 1, but the two correction products run as BF16 MMAs at twice the TF32 rate:
 *      trunc_tf32(a).trunc_tf32(b) + bf16(a - trunc_tf32(a)).bf16(b) + bf16(a).bf16(b - trunc_tf32(b));
 *   4  as 3 with a single, double-buffered accumulator (fastest of 1-4; misses the 1e-4 bar at the full minibatch);
 *   5  three FP16 passes: each operand tensor is scaled by the power of two that puts its largest magnitude
 *      in [2^14, 2^15) (exact), split x s = h + l with h = fp16(x s), l = fp16(x s - h), and the product is
 *      h.h + (h.l + l.h) in two TMEM accumulators, unscaled by the epilogue.  Same 11 significant bits per part as
 *      TF32 at the BF16 rate (3 instead of 4 BF16-equivalent passes), per-product error 2^-22.  The max-abs of every
 *      operand is a device scalar: tracked by the epilogue of the GEMM that produced the tensor, measured once per call
 *      for the weights (which are also split once per call and TMA-loaded as FP16 tiles), one reduction pass otherwise.
 *      Elements below 2^-29 of their tensor's maximum keep an absolute, not a relative, accuracy (2^-40 of the maximum).
 *   6  engine 5's arithmetic on fp16-PAIR STORAGE: every activation and every gradient that flows between two trunk GEMMs
 *      lives in global memory as (hi, lo) fp16 tensors -- written directly by the epilogue of the GEMM (or FMA layer) that
 *      produces it, scaled by the power of two of an a-priori bound (max|input| x largest row / column L1 norm of the
 *      weights + max|bias|, all device scalars) -- and is TMA-loaded straight into the MMA tiles by a conversion-free,
 *      six-stage kernel.  Same bytes as fp32 storage, no fp32 copy is written; bias gradients come out of the DGRAD
 *      epilogue (column sums); the loss head's fp32 gradient is split (and column-summed) in one extra pass;
 *   8  (default) as 6 with ONE 256-column accumulator per tile for FWD / DGRAD, double-buffered in TMEM so the epilogue of
 *      a tile overlaps the mainloop of the next (WGRAD keeps the split accumulators).  (7 is reserved: the test hook's
 *      "pair output" variant of ocppo_gemm.)
 * Measured error of one GEMM on the trunk shapes, relative to the result's max-norm (tools/acc.sh, zero-mean operands):
 * engines 5 / 6 0.7e-6 .. 0.9e-6 (FWD, DGRAD) and 2.5e-6 (65,536-row weight gradient), engine 8 1.3e-6 .. 2.3e-6 and 2.5e-6,
 * engine 3 1.3e-6 .. 2.4e-6 (4e-6 / 5e-6 for the weight gradient); every parity test holds each engine to the same 1e-4 bar. */
int ocppo_set_gemm_mode(int32_t mode);
int ocppo_get_gemm_mode(void);
/* One GEMM of the family.  op 0: C[M,N] = relu(A[M,K] B[N,K]^T + bias[N]);
 * op 1: C[M,N] = (mask[M,N] > 0) * (A[M,K] B[K,N]);
 * op 2: C[s][M,N] = sum over rows k of split s of A[k,M]^T B[k,N] (+ bias_part[s][M]
 * = column sums of A).  engine as above (1-8: the shape must be supported by the tensor-core kernel; engines 5-8 measure
 * the operand maxima themselves, 6-8 also split contiguous operands into transient fp16 pairs, 7 writes C as a pair scaled
 * by its bound and merges it back -- all synchronously: this is a test hook). */
int ocppo_gemm(int32_t op, int32_t engine, int32_t M, int32_t N, int64_t K, const float *A, int64_t lda,
               const float *B, int64_t ldb, float *C, int64_t ldc, const float *bias, const float *mask,
               float *bias_part, int32_t splits, int64_t k_per_split, void *stream);

/* ---- handle ------------------------------------------------------------ */
/* Replaces PPObj.__init__ bookkeeping (architectures/ppo.py:61-84). */
int ocppo_create(const ocppo_config *cfg, ocppo_handle **out);
int ocppo_destroy(ocppo_handle *h);

/* Number of parameter tensors (2 per Linear, nn.Module.parameters() order:
 * network.0.weight, network.0.bias, ..., actor.weight, actor.bias,
 * critic.weight, critic.bias) and the padded arena length in floats. */
int ocppo_num_tensors(const ocppo_handle *h, int32_t *n_tensors, int64_t *arena_floats,
                      int64_t *param_floats);
/* Tensor i lives at arena[offset .. offset+rows*cols), row-major (out,in) as
 * nn.Linear.weight; biases have cols == 1.  Gaps between tensors are padding
 * (kept zero) so every tensor starts 128-byte aligned. */
int ocppo_tensor_layout(const ocppo_handle *h, int32_t i, int64_t *offset, int32_t *rows,
                        int32_t *cols);
/* Bind the four flat fp32 arenas (params, grads, Adam exp_avg, exp_avg_sq),
 * each arena_floats long.  grads is the buffer ppo_atari_multigpu.py:362-374
 * builds with torch.cat and all-reduces. */
int ocppo_bind_arena(ocppo_handle *h, float *params, float *grads, float *exp_avg,
                     float *exp_avg_sq);
/* Workspace for at most `max_rows` samples per forward/backward chunk. */
int ocppo_workspace_bytes(const ocppo_handle *h, int64_t max_rows, int32_t training,
                          int64_t *bytes);
int ocppo_bind_workspace(ocppo_handle *h, void *workspace, int64_t bytes, int64_t max_rows,
                         int32_t training);

/* ---- rollout inference -------------------------------------------------- */
/* PPObj.get_action_and_value(x) under no_grad, cleanrl/ppo_atari_oc.py:505-509
 * + architectures/ppo.py:89-95.  obs (B,frames,obs_dim).  Sampling follows
 * torch-CUDA Categorical.sample(): argmax(softmax(l)/q), q = Exp(1) from
 * Philox4x32-10 at (seed, philox_offset) in torch's element layout.  If
 * `given_action` != NULL it is used instead of sampling (ppo.py:93-94).
 * Outputs: action int64 (B,), logp/entropy/value fp32 (B,); any may be NULL. */
int ocppo_policy_step(ocppo_handle *h, const float *obs, int64_t B, uint64_t seed,
                      uint64_t philox_offset, const int64_t *given_action, int64_t *action,
                      float *logp, float *entropy, float *value, void *stream);
/* The same call with the Philox (seed, offset) pair read from DEVICE memory
 * (philox_state[0] = seed, [1] = offset) by the head kernel instead of being
 * kernel arguments: a CUDA graph captured around this call can be replayed
 * for every rollout step (ppo_atari_oc.py:500-509 calls the policy num_steps
 * times per iteration), the caller refreshing the two words before each
 * replay. */
int ocppo_policy_step_dev(ocppo_handle *h, const float *obs, int64_t B, const uint64_t *philox_state,
                          const int64_t *given_action, int64_t *action, float *logp, float *entropy,
                          float *value, void *stream);
/* philox_state[0] = seed, [1] = offset (device memory), stream-ordered. */
int ocppo_philox_set(uint64_t *philox_state, uint64_t seed, uint64_t philox_offset, void *stream);
/* PPObj.get_value(x), architectures/ppo.py:86-87 (ppo_atari_oc.py:534). */
int ocppo_value(ocppo_handle *h, const float *obs, int64_t B, float *value, void *stream);
/* Optional promise that lets forward calls skip the per-call weight preparation of the FP16 engines (max-abs, fp16 split,
 * L1 norms: three small launches): `version` is any counter the caller bumps whenever it writes the parameter arena
 * (the host side passes the sum of torch's version counters of the parameters, which every in-place op increments).
 * Equal to the last value = unchanged; a negative value withdraws the promise.  Optimizer steps taken through
 * ocppo_clip_adam and ocppo_bind_arena invalidate the preparation by themselves.  (The rollout loop of
 * ppo_atari_oc.py:500-514 calls the policy 128 times per iteration on unchanged weights.) */
int ocppo_weights_version(ocppo_handle *h, int64_t version);
/* 1 when the next forward call will reuse the prepared weights (see above), else 0: the launch sequence of a forward
 * call depends on it, so a caller that replays captured CUDA graphs keys them on this. */
int ocppo_weights_prepared(const ocppo_handle *h);
/* CUDA-graph support.  Every compute call is capturable (asynchronous on its stream, no allocation, no sync), but the
 * handle also keeps HOST-side bookkeeping per call (which representation of each activation the last forward left,
 * whether the weights are prepared, the row count of the last forward).  A replayed graph redoes the device work only:
 * save the bookkeeping right after capturing a call sequence and restore it after every replay of that graph.
 * `buf` is host memory of ocppo_host_state_bytes() bytes; the content is opaque and only valid for the handle,
 * workspace binding and GEMM engine it was saved under. */
int64_t ocppo_host_state_bytes(void);
int ocppo_host_state_save(const ocppo_handle *h, void *buf, int64_t bytes);
int ocppo_host_state_restore(ocppo_handle *h, const void *buf, int64_t bytes);
/* torch.distributions.Categorical(logits=...) on given logits (B,A):
 * sample / log_prob / entropy (architectures/ppo.py:92-95). */
int ocppo_categorical(const float *logits, int64_t B, int32_t A, uint64_t seed,
                      uint64_t philox_offset, const int64_t *given_action, int64_t *action,
                      float *logp, float *entropy, void *stream);
/* Offset increment torch's generator applies for one exponential_ of `numel`. */
uint64_t ocppo_philox_increment(int64_t numel, int32_t sm_count, int32_t max_threads_per_sm);
/* The Exp(1) variates themselves (parity test vs torch.empty().exponential_()). */
int ocppo_exponential(float *out, int64_t numel, uint64_t seed, uint64_t philox_offset,
                      void *stream);

/* ---- GAE ---------------------------------------------------------------- */
/* cleanrl/ppo_atari_oc.py:535-547.  rewards/values/dones/adv/ret are (T,N)
 * row-major; next_value, next_done (N,).  Stateless. */
int ocppo_gae(const float *rewards, const float *values, const float *dones,
              const float *next_value, const float *next_done, int64_t T, int64_t N,
              double gamma, double gae_lambda, float *advantages, float *returns,
              void *stream);

/* The same scan launched as a PROGRAMMATIC DEPENDENT of the kernel before it on the stream: it may start while that
 * kernel is still draining, puts its tile loads of rewards / values / dones in flight, and waits for that kernel's
 * completion only before it reads next_value.  The caller guarantees that the preceding kernel writes none of
 * rewards / values / dones / advantages / returns -- true for the bootstrap ocppo_value call of ppo_atari_oc.py:534,
 * whose only output is next_value.  Hides the launch latency and most of the first DRAM round trip of a kernel whose
 * whole run time is a few microseconds.  Falls back to a plain launch where the TMA kernel is not eligible. */
int ocppo_gae_overlapped(const float *rewards, const float *values, const float *dones,
                         const float *next_value, const float *next_done, int64_t T, int64_t N,
                         double gamma, double gae_lambda, float *advantages, float *returns,
                         void *stream);
/* Kernel selection for tests / measurement: 0 (default) = the TMA, warp-specialised kernel whenever N % 4 == 0 and the
 * buffers are 16-byte aligned; 1 = always the first-generation cp.async kernel (the fallback for ragged buffers).
 * Both are bit-exact. */
int ocppo_set_gae_variant(int32_t variant);

/* ---- PPO update ---------------------------------------------------------- */
/* One minibatch of cleanrl/ppo_atari_oc.py:563-608: gather rows `idx` (int64,
 * mb of them; flat index t*N+n) of the flattened rollout buffers, forward,
 * losses, backward; overwrites the grad arena with d(loss)/d(params) and
 * writes OCPPO_NUM_STATS floats to `stats` (device). */
int ocppo_minibatch_fwd_bwd(ocppo_handle *h, const float *b_obs, const int64_t *b_actions,
                            const float *b_logprobs, const float *b_advantages,
                            const float *b_returns, const float *b_values, const int64_t *idx,
                            int64_t mb, const ocppo_hparams *hp, float *stats, void *stream);
/* nn.utils.clip_grad_norm_ + optim.Adam.step (cleanrl/ppo_atari_oc.py:609-610,
 * :444), with the 1/world_size of ppo_atari_multigpu.py:368-374 folded in.
 * `step` is the 1-based Adam step count.  grad_norm (device, 1 float) gets
 * the pre-clip total norm. */
int ocppo_clip_adam(ocppo_handle *h, float inv_world, float max_grad_norm, double lr,
                    double beta1, double beta2, double eps, int64_t step, float *grad_norm,
                    void *stream);

/* Writes the eight fp32 scalars the clip+Adam kernels consume to DEVICE memory
 * (scalars8), stream-ordered; they are formed in double on the host exactly as
 * torch/optim/adam.py does and travel as kernel arguments of a one-block
 * kernel (no host buffer to keep alive): inv_world, max_norm, 1-beta1, beta2,
 * 1-beta2, lr/(1-beta1^t), sqrt(1-beta2^t), eps. */
int ocppo_adam_scalars_set(float *scalars8, float inv_world, float max_grad_norm, double lr, double beta1,
                           double beta2, double eps, int64_t step, void *stream);
/* ocppo_clip_adam with those scalars read from DEVICE memory (scalars8) by the
 * kernels: graph-replayable -- LR annealing (ppo_atari_oc.py:471-474) and the
 * step count change every call, kernel arguments of a captured graph do not. */
int ocppo_clip_adam_dev(ocppo_handle *h, const float *scalars8, float *grad_norm, void *stream);

/* ---- logged statistics -------------------------------------------------- */
/* The two population variances of the script's explained-variance line (ppo_atari_oc.py:620-623) over the flat
 * rollout: out2[0] = var(returns), out2[1] = var(returns - values), both device floats (the script moves both
 * (T*N) buffers to the host and calls np.var twice).  Needs a training workspace (uses its partial-sum scratch). */
int ocppo_explained_variance(ocppo_handle *h, const float *values, const float *returns, int64_t n, float *out2,
                             void *stream);

/* Allocation-free variant of ocppo_explained_variance.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same kernels and the same result (out2[0] = var(returns), out2[1] = var(returns - values)), but the
 * per-block partial sums live in a file-scope `static __device__` array inside the library (32 KB per GPU, held by
 * every GPU that loads the library until the process exits) instead of a caller-bound workspace: the call takes no
 * handle and needs no workspace; it calls no cudaMalloc / cudaFree / allocating routine.  values / returns / out2 are
 * device pointers on one device (the call makes that device current for its duration).  Limits: 1 <= n <=
 * OCPPO_EV_STATIC_MAX_N, beyond that OCPPO_EINVAL with a message naming the limit.  All calls on a device SHARE the
 * array: it keeps the previous call's partial sums (every slot a call reads it has written first, so nothing needs
 * clearing), and two calls that overlap on the same GPU -- different streams without an ordering between them --
 * race.  Calls on one stream, or stream-ordered calls, are safe. */
#define OCPPO_EV_STATIC_MAX_N (1LL << 28)
int ocppo_explained_variance_static(const float *values, const float *returns, int64_t n, float *out2, void *stream);

/* ---- host env feed (pure host code, host pointers) ---------- */
/* Batched replacement of the per-env Python observation assembly and of VecNormalize's reward scaling:
 *  - submodules/OC_Atari/ocatari/core.py:444-448 (ns_state: chain of every slot's _nsrepr),
 *    ocatari/ram/game_objects.py:184-188 (_nsrepr = [x, y], [0, 0] when not visible),
 *    core.py:152-153,274-283 (deque(maxlen=buffer_window_size), append per step), core.py:241-244 (a reset fills every
 *    entry with the initial state), core.py:213-214 (np.array(deque): oldest frame first);
 *  - cleanrl/ppo_atari_oc.py:411-414 VecNormalize(envs, norm_obs=False, norm_reward=True) -- stable-baselines3==2.0.0
 *    (requirements pin; not vendored): returns = returns*gamma + reward; RunningMeanStd.update(returns);
 *    reward / sqrt(var + 1e-8) clipped to +-10; returns[done] = 0;
 *  - cleanrl/ppo_atari_oc.py:512-514 (float32 casts of reward / next_obs / next_done).
 * ALE / OCAtari are not available on this box: the emulator side (RAM -> object slots) stays with the env workers;
 * these calls take the slots of all N envs of one step and write straight into the caller's (pinned) buffers. */
typedef struct ocppo_feed ocppo_feed;
/* slots = object slots per env (Pong 3, SpaceInvaders 44, Breakout 110); obs_dim = 2 * slots.  gamma / epsilon /
 * clip_reward are VecNormalize's (0.99, 1e-8, 10 by default). */
int ocppo_feed_create(int32_t num_envs, int32_t frames, int32_t slots, double gamma, double epsilon, double clip_reward,
                      ocppo_feed **out);
int ocppo_feed_destroy(ocppo_feed *f);
/* One env step of all envs.  xy int32 [N, slots, 2]; visible uint8 [N, slots]; refill uint8 [N] or NULL (1: the env was
 * reset, all `frames` entries become this state; an env's first push always fills).  obs_out fp32 [N, frames, 2*slots]. */
int ocppo_feed_push(ocppo_feed *f, const int32_t *xy, const uint8_t *visible, const uint8_t *refill, float *obs_out);
/* reward float64 [N] as the envs return it, done uint8 [N] -> reward_out fp32 [N] (normalised, clipped), done_out fp32 [N]
 * (may be NULL).  Updates the running statistics (training mode). */
int ocppo_feed_normalize_reward(ocppo_feed *f, const double *reward, const uint8_t *done, float *reward_out,
                                float *done_out);
/* ret_rms (mean, var, count) and the per-env discounted returns (either may be NULL). */
int ocppo_feed_state(const ocppo_feed *f, double *mean_var_count3, double *returns);

/* ---- autograd bridge (compat tier: script's own loss/optimizer) --------- */
/* Trunk forward keeping activations: hidden (B, decoder_dims[-1]). */
int ocppo_trunk_forward(ocppo_handle *h, const float *obs, int64_t B, float *hidden,
                        void *stream);
/* Trunk backward from d(hidden) of the LAST ocppo_trunk_forward call; writes
 * trunk weight/bias grads into the grad arena (heads untouched). */
int ocppo_trunk_backward(ocppo_handle *h, const float *obs, int64_t B, const float *d_hidden,
                         void *stream);

/* Test hook: post-ReLU activation of trunk layer `layer` (0-based, encoder then decoder layers) as the LAST
 * forward left it in the workspace -- (B*frames, width) for encoder layers, (B, width) for decoder layers.
 * The backward pass takes its ReLU masks from exactly these buffers, so a parity test can hand the same
 * masks to the reference (tests/test_gpu_parity.py: decisions within fp32 rounding of zero). */
int ocppo_read_activation(const ocppo_handle *h, int32_t layer, int64_t B, float *out, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* OCPPO_H */
