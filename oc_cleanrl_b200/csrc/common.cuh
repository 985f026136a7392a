// Shared declarations for libocppo.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/ocppo.h"

namespace ocppo {

void set_error(const char *fmt, ...);

#define OCPPO_CHECK_ARG(cond, ...)                \
    do {                                          \
        if (!(cond)) {                            \
            ::ocppo::set_error(__VA_ARGS__);      \
            return OCPPO_EINVAL;                  \
        }                                         \
    } while (0)

#define OCPPO_CUDA(expr)                                                              \
    do {                                                                              \
        cudaError_t _e = (expr);                                                      \
        if (_e != cudaSuccess) {                                                      \
            ::ocppo::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                               __FILE__, __LINE__);                                   \
            return OCPPO_ECUDA;                                                       \
        }                                                                             \
    } while (0)

#define OCPPO_LAUNCH_CHECK(name)                                                        \
    do {                                                                                \
        cudaError_t _e = cudaGetLastError();                                            \
        if (_e != cudaSuccess) {                                                        \
            ::ocppo::set_error("launch of %s failed: %s", name, cudaGetErrorString(_e)); \
            return OCPPO_ECUDA;                                                         \
        }                                                                               \
    } while (0)

#define OCPPO_TRY(expr)          \
    do {                         \
        int _r = (expr);         \
        if (_r != OCPPO_OK) return _r; \
    } while (0)

// ---- launch accounting / live per-class kernel timing (bench.py roofline) ----
enum KernelClass { KC_GAE = 0, KC_SGEMM_FWD = 1, KC_SGEMM_DGRAD = 2, KC_SGEMM_WGRAD = 3, KC_HEAD = 4,
                   KC_AUX = 5, KC_ADAM = 6, KC_GEMM_CUDA_CORE = 7, KC_COUNT = 8 };
void prof_begin(int cls, int n_kernels, cudaStream_t s);
void prof_end(cudaStream_t s);
struct ProfScope {
    cudaStream_t s;
    ProfScope(int cls, int n_kernels, cudaStream_t st) : s(st) { prof_begin(cls, n_kernels, st); }
    ~ProfScope() { prof_end(s); }
};

static inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }
static inline int64_t ceil_div(int64_t x, int64_t m) { return (x + m - 1) / m; }

constexpr int kMaxDevices = 64;           // per-device caches of function attributes / SM counts
constexpr int kArenaAlignFloats = 32;      // every tensor starts 128 B aligned
constexpr int kSkinnyMaxIn = 96;           // widest layer input the skinny weight-gradient route takes (32 columns per pass; beyond 3 passes the tiled GEMM wins)
constexpr int kSplitMax = 32;              // max split-K factor of the wgrad kernels
constexpr int kLossBlocksMax = 4096;       // partial-sum slots of the loss head
constexpr int kNormBlocks = 592;           // 148 SMs x 4 CTAs for the grad-norm pass

struct Layer {
    int in_dim;       // K of the forward GEMM
    int out_dim;      // N
    int row_mult;     // rows per sample: frames for encoder layers, 1 after Flatten
    int64_t w_off;    // arena offsets (floats)
    int64_t b_off;
};

}  // namespace ocppo

struct ocppo_handle {
    ocppo_config cfg;
    int n_layers;                       // trunk Linear layers (encoder + decoder)
    ocppo::Layer layers[2 * OCPPO_MAX_LAYERS];
    ocppo::Layer actor, critic;
    int hidden;                         // width fed to the heads
    int n_tensors;
    int64_t tensor_off[4 * OCPPO_MAX_LAYERS + 4];
    int32_t tensor_rows[4 * OCPPO_MAX_LAYERS + 4];
    int32_t tensor_cols[4 * OCPPO_MAX_LAYERS + 4];
    int64_t arena_floats, param_floats;
    int sm_count, max_threads_per_sm;

    float *params, *grads, *exp_avg, *exp_avg_sq;

    // workspace carve-up (all device pointers into the caller's buffer)
    void *ws;
    int64_t ws_bytes, max_rows;
    int training;
    float *x0;                          // gathered observations [max_rows*frames, obs_dim]
    float *act[2 * OCPPO_MAX_LAYERS];   // post-ReLU activations per trunk layer
    uint32_t *relu_bits[2 * OCPPO_MAX_LAYERS];   // their ReLU decisions, one bit each (training workspaces)
    int relu_bits_valid[2 * OCPPO_MAX_LAYERS];   // written by the last forward (tensor-core engine only)
    float *gbuf[2];                     // ping-pong gradient buffers
    float *gz;                          // head gradients [max_rows, AP]
    int32_t *mb_action;
    float *mb_logp, *mb_adv, *mb_ret, *mb_val;
    float *splitk;                      // [kSplitMax, max(out*in)]
    float *splitk_bias;                 // [kSplitMax, max(out)]
    float *skinny;                      // slab partials of the skinny wgrad kernels
    double *loss_part;                  // [kLossBlocksMax, 8]
    double *adv_part;                   // [kLossBlocksMax, 2]
    float *adv_stat;                    // mean, std
    double *norm_part;                  // [kNormBlocks]
    int last_forward_rows;
    // FP16x3 tensor-core engine: max |.| per tensor as device scalars.  [0,1] scratch, then per trunk layer l:
    // kAmaxAct + l = activations act[l], kAmaxW + l = weights, kAmaxGrad + l = gradient entering layer l's backward
    float *amax;
    int act_amax_valid[2 * OCPPO_MAX_LAYERS];
    int w_amax_valid;
    // weight preparation of the FP16 engines (max-abs, fp16 split, norms) may be skipped by a forward call when the caller
    // vouches that the arena is unchanged (ocppo_weights_version) and no optimizer step ran in between
    int64_t w_version; int w_prepared_mode;
    void *w16_hi, *w16_lo;              // the trunk weights split into scaled fp16 hi / lo (indexed like params)
    // fp16-pair storage (GEMM engine 6): activations of every trunk layer but the last and the gradients flowing down the
    // trunk exist as (hi, lo) fp16 tensors; the fp32 buffers above are filled on demand only (merge16)
    void *act16_hi[2 * OCPPO_MAX_LAYERS], *act16_lo[2 * OCPPO_MAX_LAYERS];
    int act16_valid[2 * OCPPO_MAX_LAYERS], act32_valid[2 * OCPPO_MAX_LAYERS];
    const float *act_scale[2 * OCPPO_MAX_LAYERS];   // device scalar the pair's power-of-two scale derives from
    void *g16_hi[2], *g16_lo[2];        // pair twins of gbuf[]
    float *colsum_part;                 // [kColsumBlocksMax, max(out)] bias-gradient partials
    // first layer with a wide object vector (32 < D <= 256, e.g. SpaceInvaders 88, Breakout 220): input and weight rows
    // zero-padded to kPadCols in pair storage, so the layer's FWD and WGRAD run on the conversion-free tensor-core kernel
    void *x0p_hi, *x0p_lo;              // [max_rows * frames, kPadCols]
    void *w0p_hi, *w0p_lo;              // [out_0, kPadCols]
    int x0p_valid;
};
// device scalars in ocppo_handle::amax: true maxima (Act / W / Grad), the bounds the epilogues scale their fp16-pair outputs
// by (BndAct + l: act[l]; BndGrad + l: gradient entering layer l's backward), and per layer the weights' largest row / column
// L1 norm and largest |bias|
constexpr int kAmaxAct = 2, kAmaxW = 2 + 2 * OCPPO_MAX_LAYERS, kAmaxGrad = 2 + 4 * OCPPO_MAX_LAYERS;
constexpr int kBndAct = 2 + 6 * OCPPO_MAX_LAYERS, kBndGrad = 2 + 8 * OCPPO_MAX_LAYERS;
constexpr int kWRowL1 = 2 + 10 * OCPPO_MAX_LAYERS, kWColL1 = 2 + 12 * OCPPO_MAX_LAYERS, kBMax = 2 + 14 * OCPPO_MAX_LAYERS;
constexpr int kAmaxObs = 2 + 16 * OCPPO_MAX_LAYERS;           // max |observation| of the current chunk
constexpr int kAmaxSlots = 4 + 16 * OCPPO_MAX_LAYERS;

namespace ocppo {

// ---- kernels.cu / sgemm.cu entry points used across files -----------------
enum SgemmMode { SGEMM_FWD = 0, SGEMM_DGRAD = 1, SGEMM_WGRAD = 2 };

struct SgemmArgs {
    const float *A; int64_t lda;
    const float *B; int64_t ldb;
    float *C; int64_t ldc;
    const float *bias;        // FWD: [N]
    const float *mask;        // DGRAD: same layout as C; C = mask > 0 ? acc : 0
    uint32_t *relu_bits;      // optional, [M, N/32] words, bit j of word w = (element 32 w + j > 0).  FWD on the tensor-core
                              // engine writes them next to C; DGRAD reads them INSTEAD of `mask` (1/32 of the bytes)
    float *bias_part;         // WGRAD: [splits, M] column sums of A
    int M, N;
    int64_t K;                // contraction length (rows for WGRAD)
    int64_t k_per_split;
    int splits;
    int dbg;                  // tc_gemm bring-up variant (0 in production)
    float trunc_scale;        // tc_gemm: multiplies the accumulator-truncation compensation (1 in production; set by the launcher)
    int single_acc;           // tc_gemm: 1 = one TMEM accumulator per tile (faster, ~3x the rounding error)
    int corr_bf16;            // tc_gemm: 1 = the two correction products run as BF16 MMAs (split accumulators only)
    int f16x3;                // tc_gemm: 1 = three FP16 products on per-tensor power-of-two scaled operands (engine 5)
    const float *amax_a;      // f16x3: device scalars, max |A| and max |B| (>= the true maxima); null = sgemm_launch
    const float *amax_b;      //        measures them with a reduction pass into amax_scratch[0..1]
    float *amax_out;          // optional (tensor-core FWD / DGRAD): atomicMax of |C| as float bits; caller zeroes it
    float *amax_scratch;      // two floats of device scratch for the fallback reduction
    const void *b16_hi;       // f16x3, FWD / DGRAD: B already split into fp16 hi / lo tensors (same shape and ldb, scaled by
    const void *b16_lo;       //        the power of two that amax_b implies) -- the weights, split once per call
    // fp16-pair storage (engine 6): activations / gradients live in global memory as (hi, lo) fp16 tensors of the fp32
    // tensor's shape and stride.  With A and B both pre-split the conversion-free kernel runs (B also for WGRAD).
    const void *a16_hi, *a16_lo;   // A pre-split, scaled by the power of two that *amax_a implies
    void *c16_hi, *c16_lo;         // FWD / DGRAD: write C as an fp16 pair INSTEAD of fp32 C, scaled by the power of two of
    const float *bnd_amax;         //   bound = *bnd_amax (true max |A|) * *bnd_wl1 (largest row / column L1 norm of the
    const float *bnd_wl1;          //   weights) + *bnd_bmax (largest |bias|, may be null);
    const float *bnd_bmax;
    float *c_scale_src;            //   the kernel stores the bound here: the consumers' amax_a / amax_b
    int epi2;                      // conversion-free kernel: two epilogue staging tiles per warp + five-stage ring (set by the launcher)
    int want_pre;                  // planning hint for tc_gemm_supported(): the launch will carry both operands as fp16 pairs
                                   // (that alone admits the 128-column tile for N % 256 != 0)
    int fma_bits;                  // FMA kernel, FWD: write relu_bits too (full 128-column tiles only)
    float *colsum_part;            // DGRAD, conversion-free kernel: [tc_gemm_ctas() * 4, N] column sums of C, pre-zeroed
};

int sgemm_launch(SgemmMode mode, const SgemmArgs &a, cudaStream_t stream);       // dispatches FMA / tensor
int fma_gemm_launch(SgemmMode mode, const SgemmArgs &a, cudaStream_t stream);    // sgemm.cu, fp32 CUDA cores
bool tc_gemm_supported(SgemmMode mode, const SgemmArgs &a);                      // tc_gemm.cu, tcgen05 3xTF32
int tc_gemm_launch(SgemmMode mode, const SgemmArgs &a, cudaStream_t stream);
int tc_gemm_ctas(SgemmMode mode, const SgemmArgs &a, int sm_count);                 // grid of that launch
void sgemm_engine_flags(int engine, SgemmArgs *b);                               // engine number -> tc_gemm variant flags
int sgemm_fill_amax(SgemmMode mode, SgemmArgs *b, cudaStream_t stream);          // FP16x3: measure missing operand maxima
bool sgemm_uses_tc(SgemmMode mode, const SgemmArgs &a);                          // what sgemm_launch will pick
int tc_tile_n(const SgemmArgs &a);                                                // N tile of the CTA pair: 256, 128 (pair operands, N % 256 != 0) or 0 (unsupported)
// 0: fp32 FMA kernels only; 1: tcgen05 3xTF32 wherever the shape fits (fp32-grade accuracy; main and
// correction products in separate TMEM accumulators); 2: same with a single, double-buffered accumulator
extern int g_gemm_mode;

// ---- FP16x3 operand split (tc_gemm.cu producers, absmax.cu weight pre-split) ----
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t pack_f16(float lo_elem, float hi_elem) {
    uint32_t r;
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi_elem), "f"(lo_elem));
    return r;
}
// scale 2^k with amax * 2^k in [2^14, 2^15), k clamped to [-100, 126]; inv = 2^-k
__device__ __forceinline__ void f16_scale_of(float amax, float &scale, float &inv) {
    const int e = (int)((__float_as_uint(amax) >> 23) & 0xFFu);        // biased exponent (0 for zero / subnormal)
    int k = 14 - (e - 127);
    k = k > 126 ? 126 : (k < -100 ? -100 : k);
    if (e == 0 || e == 255) k = 0;                                     // all-zero tensor, or Inf / NaN: results follow the data
    scale = __uint_as_float((uint32_t)(127 + k) << 23);
    inv = __uint_as_float((uint32_t)(127 - k) << 23);
}
__device__ __forceinline__ void split_f16x2(float x, float y, float scale, uint32_t &hi, uint32_t &lo) {
    x *= scale; y *= scale;
    hi = pack_f16(x, y);
    const float2 hf = __half22float2(*reinterpret_cast<const __half2 *>(&hi));
    lo = pack_f16(x - hf.x, y - hf.y);
}
#endif

struct AbsmaxSegs { const float *x[2 * OCPPO_MAX_LAYERS]; int64_t n[2 * OCPPO_MAX_LAYERS]; float *out[2 * OCPPO_MAX_LAYERS]; int count; };
// hi[i] = fp16(x[i] s), lo[i] = fp16(x[i] s - hi[i]) per segment, s from the segment's max-abs (device scalar out[seg]);
// hi / lo are indexed like x relative to segs.x[0]'s arena base `base`
int launch_split_f16_multi(const AbsmaxSegs &segs, const float *base, void *hi, void *lo, cudaStream_t stream);
// ---- fp16-pair storage (pair16.cu; GEMM engine 6) ----
struct WnormSegs {
    const float *w[2 * OCPPO_MAX_LAYERS]; const float *b[2 * OCPPO_MAX_LAYERS]; int out_dim[2 * OCPPO_MAX_LAYERS], in_dim[2 * OCPPO_MAX_LAYERS];
    float *row_l1[2 * OCPPO_MAX_LAYERS], *col_l1[2 * OCPPO_MAX_LAYERS], *b_max[2 * OCPPO_MAX_LAYERS]; int count;
    float *w_amax[2 * OCPPO_MAX_LAYERS];   // optional (null: skipped): max |W| too, saving launch_absmax_multi's pass over the weights
};
constexpr int kColsumBlocksMax = 1024;
constexpr int kPadCols = 256;              // padded width of the first layer's input / weight rows in pair storage (one N tile)
int launch_pad_split16(const float *x, int64_t rows, int cols, int64_t ld, const float *scale_src, void *hi, void *lo, int sm_count,
                       cudaStream_t stream);
int launch_splitk_reduce_pad(const float *part, int splits, int rows, int cols, float *out, int accumulate, int sm_count, cudaStream_t stream);
int launch_split16(const float *x, int64_t n, const float *scale_src, void *hi, void *lo, int sm_count, cudaStream_t stream);
int launch_merge16(const void *hi, const void *lo, int64_t n, const float *scale_src, float *out, int sm_count, cudaStream_t stream);
int launch_colsum16(const void *hi, const void *lo, int64_t rows, int cols, const float *scale_src, float *part, int *blocks,
                    int sm_count, cudaStream_t stream);     // part: [kColsumBlocksMax][cols] at most; *blocks partial rows written
int launch_split16_colsum(const float *x, int64_t rows, int cols, const float *scale_src, void *hi, void *lo, float *part, int *blocks,
                          int sm_count, cudaStream_t stream);   // pair + column-sum partials of an fp32 tensor in one read
int launch_colsum_reduce(const float *part, int parts, int cols, float *out, int accumulate, cudaStream_t stream);
// split-K fold of a weight gradient + column-sum fold of its bias gradient in one launch (same sums as the two kernels)
int launch_wgrad_fold(const float *part, int splits, int64_t stride, int64_t count, float *out, int acc_w, const float *cpart,
                      int parts, int cols, float *cout, int acc_b, int sm_count, cudaStream_t stream);
int launch_wnorm(const WnormSegs &segs, cudaStream_t stream);                     // outputs pre-zeroed
int launch_absmax_multi(const AbsmaxSegs &segs, cudaStream_t stream);             // several small contiguous tensors, one launch; outs pre-zeroed
// out[0] = max |x[r, c]| over an [rows, cols] view with row stride ld (zeroed first); absmax.cu
int launch_absmax(const float *x, int64_t rows, int64_t cols, int64_t ld, float *out, int sm_count, cudaStream_t stream);

// tc_gemm.cu: plain 2-D fp32 TMA tile map (128-byte CUtensorMap written to tm_out), memoised
int tma_map_2d_f32_plain(void *tm_out, const float *base, int64_t rows, int64_t cols, int box_rows, int box_cols);
extern int g_gae_variant;
int launch_gae(const float *rewards, const float *values, const float *dones,
               const float *next_value, const float *next_done, int64_t T, int64_t N, float g,
               float c, float *adv, float *ret, int overlap_previous, cudaStream_t stream);

struct LossArgs {
    const float *hidden;        // [rows, H] post-ReLU output of the last trunk layer
    const int32_t *action;      // gathered minibatch columns, [rows]
    const float *logp_old, *adv, *ret, *val;
    const float *adv_stat;      // {mean, unbiased std} of the whole minibatch
    const float *Wa, *ba, *Wc, *bc;
    int64_t rows, mb;
    int H, A, AP;
    float clip_lo, clip_hi, clip_f, ent_coef, vf_coef;
    int norm_adv, clip_vloss;
    float *dh;                  // [rows, H]  d loss / d (pre-activation of last trunk layer)
    float *gz;                  // [rows, AP] d loss / d (logits, value), zero padded
    float *dh_amax;             // optional: atomicMax of |dh| as float bits (FP16x3 engine's operand scale); caller zeroes it
    double *part;               // [gridDim.x, 8] partial sums of the logged statistics
    // two-kernel variant (fp16-pair storage): the first kernel stops at gz and records max_i (|dv_i| + sum_k |dz_ik|) here
    // (float bits, caller zeroes it); the second kernel (launch_loss_head_bwd_pair) turns gz into d(hidden) as an fp16 pair
    float *gz_l1max;
};

struct LossBwdPairArgs {
    const float *gz;            // [rows, AP] from the forward kernel
    const uint32_t *relu_bits;  // ReLU decisions of the last trunk layer, [rows, H/32] words; null -> `hidden` is read instead
    const float *hidden;        // [rows, H] (only read when relu_bits is null)
    const float *Wa, *Wc;       // head weights [A, H], [1, H]
    const float *gz_l1max;      // device scalar written by the forward kernel
    int64_t rows;
    int H, A, AP;
    void *dh_hi, *dh_lo;        // [rows, H] fp16 each, scaled by the power of two of  *bound = gz_l1max * max|W_head|
    float *bound;               // written by block 0: what the consumers derive the pair's scale from
    float *dh_amax;             // atomicMax of the true |dh| (float bits; caller zeroes it)
    float *colsum_part;         // [blocks, H] column sums of dh (bias gradient of the last trunk layer)
};

int launch_policy_head(const ocppo_handle *h, const float *hidden, int64_t B, int64_t row_base, int64_t total_B,
                       uint64_t seed, uint64_t offset, const uint64_t *philox_dev, const int64_t *given, int64_t *action,
                       float *logp, float *entropy, float *value, int want_policy, cudaStream_t stream);
int launch_categorical(const float *logits, int64_t B, int A, uint64_t seed, uint64_t offset, const int64_t *given,
                       int64_t *action, float *logp, float *entropy, cudaStream_t stream);
int launch_exponential(float *out, int64_t numel, uint64_t seed, uint64_t offset, cudaStream_t stream);
int launch_gather(const ocppo_handle *h, const float *b_obs, const int64_t *b_actions, const float *b_logp,
                  const float *b_adv, const float *b_ret, const float *b_val, const int64_t *idx, int64_t rows,
                  cudaStream_t stream);
int launch_adv_stats(const ocppo_handle *h, const float *b_adv, const int64_t *idx, int64_t mb, cudaStream_t stream);
int launch_loss_head(const ocppo_handle *h, LossArgs a, int *blocks_used, int blocks_avail, cudaStream_t stream,
                     bool forward_only = false);
bool loss_head_pair_supported(int H, int A);
int launch_loss_head_bwd_pair(const ocppo_handle *h, LossBwdPairArgs a, int *colsum_blocks, cudaStream_t stream);
int launch_explained_var(const ocppo_handle *h, const float *y_pred, const float *y_true, int64_t n, float *out2,
                         cudaStream_t stream);
int launch_loss_finalize(const double *part, int n_part, int64_t mb, float ent_coef, float vf_coef, float *stats,
                         cudaStream_t stream);
int launch_splitk_reduce(const float *part, int splits, int64_t stride, int64_t count, float *out, int accumulate,
                         int sm_count, cudaStream_t stream);
int skinny_slabs(int64_t rows, int sm_count, int64_t *rows_per_cta, int per_sm = 4);
static inline int64_t skinny_slabs_max(int64_t rows, int sm_count) { int64_t per; return skinny_slabs(rows, sm_count, &per); }
int launch_skinny_wgrad(const float *wide, int64_t ldw, int nw, const float *skinny, int64_t lds, int ns, int64_t rows,
                        int sm_count, float *part, float *wsum_part, float *ssum_part, int *G_out, int *NS_out,
                        cudaStream_t stream);
int launch_skinny_reduce(const float *part, int G, int nw, int NS, int i0, int i1, float *out, int64_t si, int64_t sj,
                         int accumulate, cudaStream_t stream);
struct SkinnyReduceJob { const float *part; int G, nw, NS, i0, i1; float *out; int64_t si, sj; int accumulate; };
int launch_skinny_reduce_multi(const SkinnyReduceJob *jobs, int n, cudaStream_t stream);     // same sums, one launch
int launch_relu_mask(const float *g, const float *act, int64_t n, float *out, int sm_count, cudaStream_t stream);
int measure_fma_peak(double *tflops, cudaStream_t stream);
int launch_store_words(void *dst, const void *words, int n, cudaStream_t stream);     // n <= 8 32-bit words
void adam_scalars(float inv_world, float max_norm, double lr, double beta1, double beta2, double eps, int64_t step, float *out8);
// exactly one of host8 / dev8 (device memory, read by the kernels: graph-replayable) carries the eight scalars
int launch_clip_adam(ocppo_handle *h, const float *host8, const float *dev8, float *grad_norm, cudaStream_t stream);

}  // namespace ocppo
