// C-ABI entry points of libocppo.so (declared in include/ocppo.h) and the
// per-layer orchestration of the PPObj trunk forward / backward.
#include "common.cuh"
#include <math.h>
#include <new>
#include <string.h>

namespace ocppo {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// Live timing: when enabled every launch wrapper is bracketed by a pair of events on its own stream.  The pairs sit
// in a small ring; whenever it fills (and on ocppo_profile_read) the ring is DRAINED into per-class sums -- a host-side
// wait, taken in measurement runs only -- so no launch is ever dropped: timed[] always equals the launches made while
// profiling was on.  Launches issued while their stream is being captured into a CUDA graph cannot carry events; they
// are counted in uncaptured[] so a reader can tell.
static const int kProfRing = 2048;
static struct {
    bool on = false;
    int n = 0;
    bool open = false;                 // prof_begin recorded a start event that prof_end has to close
    cudaEvent_t start[kProfRing], stop[kProfRing];
    int cls[kProfRing], nk[kProfRing];
    bool created = false;
    int64_t launches[KC_COUNT] = {0};
    double ms[KC_COUNT] = {0};
    int64_t timed[KC_COUNT] = {0};
    int64_t skipped[KC_COUNT] = {0};   // profiling on, but the stream was capturing
} g_prof;

static int prof_drain() {
    for (int i = 0; i < g_prof.n; ++i) {
        OCPPO_CUDA(cudaEventSynchronize(g_prof.stop[i]));
        float ms = 0;
        OCPPO_CUDA(cudaEventElapsedTime(&ms, g_prof.start[i], g_prof.stop[i]));
        g_prof.ms[g_prof.cls[i]] += ms;
        g_prof.timed[g_prof.cls[i]] += g_prof.nk[i];
    }
    g_prof.n = 0;
    return OCPPO_OK;
}

void prof_begin(int cls, int n_kernels, cudaStream_t s) {
    g_prof.launches[cls] += n_kernels;
    g_prof.open = false;
    if (!g_prof.on) return;
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &cap) != cudaSuccess || cap != cudaStreamCaptureStatusNone) {
        g_prof.skipped[cls] += n_kernels;
        return;
    }
    if (g_prof.n >= kProfRing) prof_drain();
    g_prof.cls[g_prof.n] = cls;
    g_prof.nk[g_prof.n] = n_kernels;
    cudaEventRecord(g_prof.start[g_prof.n], s);
    g_prof.open = true;
}
void prof_end(cudaStream_t s) {
    if (!g_prof.open) return;
    cudaEventRecord(g_prof.stop[g_prof.n], s);
    ++g_prof.n;
    g_prof.open = false;
}

// Every entry point that takes a handle makes the handle's device current for the duration of the call (and restores
// the caller's): a drop-in agent on cuda:k must work without the caller ever calling cudaSetDevice (the reference's
// multi-GPU script never does, ppo_atari_multigpu.py:216-222).
struct DeviceGuard {
    int prev = -1;
    bool switched = false;
    explicit DeviceGuard(int dev) {
        if (dev < 0) return;
        if (cudaGetDevice(&prev) == cudaSuccess && prev != dev) switched = cudaSetDevice(dev) == cudaSuccess;
    }
    ~DeviceGuard() { if (switched) cudaSetDevice(prev); }
};
// stateless entry points (no handle): the device that owns the first data pointer
static int device_of(const void *p) {
    cudaPointerAttributes at;
    if (p != nullptr && cudaPointerGetAttributes(&at, p) == cudaSuccess && at.type == cudaMemoryTypeDevice) return at.device;
    cudaGetLastError();
    return -1;
}

static inline cudaStream_t as_stream(void *s) { return reinterpret_cast<cudaStream_t>(s); }

// ---- layout ---------------------------------------------------------------
static int build_layout(ocppo_handle *h) {
    const ocppo_config &c = h->cfg;
    OCPPO_CHECK_ARG(c.frames >= 1 && c.obs_dim >= 1, "frames/obs_dim must be >= 1");
    OCPPO_CHECK_ARG(c.n_actions >= 1 && c.n_actions <= OCPPO_MAX_ACTIONS, "n_actions must be in [1,%d]",
                    OCPPO_MAX_ACTIONS);
    OCPPO_CHECK_ARG(c.n_encoder >= 0 && c.n_encoder <= OCPPO_MAX_LAYERS, "bad n_encoder");
    OCPPO_CHECK_ARG(c.n_decoder >= 0 && c.n_decoder <= OCPPO_MAX_LAYERS, "bad n_decoder");
    OCPPO_CHECK_ARG(c.n_encoder + c.n_decoder >= 1, "need at least one trunk layer");
    int64_t off = 0;
    int nt = 0, nl = 0;
    auto add_tensor = [&](int rows, int cols) {
        h->tensor_off[nt] = off;
        h->tensor_rows[nt] = rows;
        h->tensor_cols[nt] = cols;
        ++nt;
        const int64_t o = off;
        off = round_up(off + (int64_t)rows * cols, kArenaAlignFloats);
        return o;
    };
    auto add_linear = [&](Layer &L, int in, int out, int mult) {
        L.in_dim = in; L.out_dim = out; L.row_mult = mult;
        L.w_off = add_tensor(out, in);
        L.b_off = add_tensor(out, 1);
    };
    int width = c.obs_dim;
    for (int i = 0; i < c.n_encoder; ++i) {
        OCPPO_CHECK_ARG(c.encoder_dims[i] >= 1, "encoder dim must be >= 1");
        add_linear(h->layers[nl++], width, c.encoder_dims[i], c.frames);
        width = c.encoder_dims[i];
    }
    width *= c.frames;                               // nn.Flatten, architectures/ppo.py:74-75
    for (int i = 0; i < c.n_decoder; ++i) {
        OCPPO_CHECK_ARG(c.decoder_dims[i] >= 1, "decoder dim must be >= 1");
        add_linear(h->layers[nl++], width, c.decoder_dims[i], 1);
        width = c.decoder_dims[i];
    }
    // With no decoder the heads would read the flattened encoder output; the
    // reference supports it (ppo.py:76,83) and so do we: hidden = frames * enc[-1].
    h->n_layers = nl;
    h->hidden = width;
    add_linear(h->actor, width, c.n_actions, 1);
    add_linear(h->critic, width, 1, 1);
    h->n_tensors = nt;
    h->arena_floats = off;
    int64_t pf = 0;
    for (int i = 0; i < nt; ++i) pf += (int64_t)h->tensor_rows[i] * h->tensor_cols[i];
    h->param_floats = pf;
    return OCPPO_OK;
}

struct Carver {
    char *base;
    int64_t off;
    template <typename T>
    T *take(int64_t count) {
        T *p = base ? reinterpret_cast<T *>(base + off) : nullptr;
        off = round_up(off + count * (int64_t)sizeof(T), 256);
        return p;
    }
};

static int64_t carve(ocppo_handle *h, void *base, int64_t R, int training, bool assign) {
    Carver cv{reinterpret_cast<char *>(base), 0};
    ocppo_handle tmp = *h;
    ocppo_handle *t = assign ? h : &tmp;
    const ocppo_config &c = h->cfg;
    int64_t max_act = 0, max_w = 0, max_out = 0, max_dim = 0;
    for (int l = 0; l < h->n_layers; ++l) {
        const Layer &L = h->layers[l];
        const int64_t n = R * L.row_mult * L.out_dim;
        t->act[l] = cv.take<float>(n);
        const bool pair = l + 1 < h->n_layers && n % 8 == 0;      // every activation a trunk GEMM may consume
        t->act16_hi[l] = pair ? cv.take<uint16_t>(n) : nullptr;
        t->act16_lo[l] = pair ? cv.take<uint16_t>(n) : nullptr;
        t->act16_valid[l] = 0; t->act32_valid[l] = 0; t->act_scale[l] = nullptr;
        t->relu_bits[l] = (training && L.out_dim % 32 == 0) ? cv.take<uint32_t>(n / 32) : nullptr;
        t->relu_bits_valid[l] = 0;
        if (n > max_act) max_act = n;
        if ((int64_t)L.out_dim * L.in_dim > max_w) max_w = (int64_t)L.out_dim * L.in_dim;
        if (L.out_dim > max_out) max_out = L.out_dim;
        if (L.in_dim > max_dim) max_dim = L.in_dim;
        if (L.out_dim > max_dim) max_dim = L.out_dim;
    }
    {   // wide first layer (32 < D <= kPadCols): padded pair copies of its input and weight
        const Layer &L0 = h->layers[0];
        const bool pad0 = h->n_layers > 1 && L0.in_dim > 32 && L0.in_dim <= kPadCols && L0.out_dim % 128 == 0;
        t->x0p_hi = pad0 ? cv.take<uint16_t>(R * L0.row_mult * kPadCols) : nullptr;
        t->x0p_lo = pad0 ? cv.take<uint16_t>(R * L0.row_mult * kPadCols) : nullptr;
        t->w0p_hi = pad0 ? cv.take<uint16_t>((int64_t)L0.out_dim * kPadCols) : nullptr;
        t->w0p_lo = pad0 ? cv.take<uint16_t>((int64_t)L0.out_dim * kPadCols) : nullptr;
        t->x0p_valid = 0;
        if (pad0 && (int64_t)L0.out_dim * kPadCols > max_w) max_w = (int64_t)L0.out_dim * kPadCols;
    }
    t->amax = cv.take<float>(kAmaxSlots);
    t->w_amax_valid = 0;
    t->w16_hi = cv.take<uint16_t>(h->arena_floats);     // indexed by ARENA offsets (alignment padding included)
    t->w16_lo = cv.take<uint16_t>(h->arena_floats);
    const int AP = (int)round_up(c.n_actions + 1, 4);
    if ((int64_t)AP * h->hidden > max_w) max_w = (int64_t)AP * h->hidden;
    if (AP > max_out) max_out = AP;
    if (training) {
        t->x0 = cv.take<float>(R * c.frames * c.obs_dim);
        t->gbuf[0] = cv.take<float>(max_act);
        t->gbuf[1] = cv.take<float>(max_act);
        for (int i = 0; i < 2; ++i) {
            t->g16_hi[i] = cv.take<uint16_t>(max_act);
            t->g16_lo[i] = cv.take<uint16_t>(max_act);
        }
        t->colsum_part = cv.take<float>((int64_t)kColsumBlocksMax * max_dim);
        t->gz = cv.take<float>(R * AP);
        t->mb_action = cv.take<int32_t>(R);
        t->mb_logp = cv.take<float>(R);
        t->mb_adv = cv.take<float>(R);
        t->mb_ret = cv.take<float>(R);
        t->mb_val = cv.take<float>(R);
        t->splitk = cv.take<float>((int64_t)kSplitMax * max_w);
        t->splitk_bias = cv.take<float>((int64_t)kSplitMax * max_out);
        {   // slab partials of the skinny weight-gradient kernels (first trunk layer, heads)
            const Layer &L0 = h->layers[0];
            const int64_t g0 = (int64_t)h->sm_count * 4, gh = g0;     // upper bound of the slab count
            int64_t need = gh * ((int64_t)h->hidden * 33 + 32);
            (void)L0;
            for (int l = 0; l < h->n_layers; ++l) {            // every layer whose wgrad may take the skinny route
                const Layer &L = h->layers[l];
                if (L.in_dim <= kSkinnyMaxIn) {
                    const int64_t n0 = g0 * (int64_t)L.out_dim * 33;
                    if (n0 > need) need = n0;
                }
            }
            t->skinny = cv.take<float>(need);
        }
        t->loss_part = cv.take<double>((int64_t)kLossBlocksMax * 8);
        t->adv_part = cv.take<double>((int64_t)kLossBlocksMax * 2);
        t->adv_stat = cv.take<float>(8);
        t->norm_part = cv.take<double>(kNormBlocks);
    }
    return cv.off;
}

static int require_ready(const ocppo_handle *h, bool need_training) {
    if (h == nullptr) { set_error("null handle"); return OCPPO_EINVAL; }
    if (h->cfg.device < 0) { set_error("layout-only handle (device < 0) cannot compute"); return OCPPO_ESTATE; }
    if (h->params == nullptr) { set_error("arena not bound (ocppo_bind_arena)"); return OCPPO_ESTATE; }
    if (h->ws == nullptr) { set_error("workspace not bound (ocppo_bind_workspace)"); return OCPPO_ESTATE; }
    if (need_training && !h->training) { set_error("workspace was bound without training buffers"); return OCPPO_ESTATE; }
    return OCPPO_OK;
}

// ---- trunk forward / backward ---------------------------------------------
// x: [B*frames, obs_dim] contiguous.  Writes h->act[l] for every trunk layer.
// FP16x3 engine: max |act[l]| measured on demand when the kernel that wrote act[l] did not track it (FMA layers)
static int ensure_act_amax(ocppo_handle *h, int l, int64_t B, cudaStream_t s) {
    if (h->act_amax_valid[l]) return OCPPO_OK;
    const Layer &L = h->layers[l];
    OCPPO_TRY(launch_absmax(h->act[l], B * L.row_mult, L.out_dim, L.out_dim, h->amax + kAmaxAct + l, h->sm_count, s));
    h->act_amax_valid[l] = 1;
    return OCPPO_OK;
}

// fp16-pair storage (engine 6): whichever representation of act[l] the last forward did not leave is made on demand
static int ensure_act_fp32(ocppo_handle *h, int l, int64_t B, cudaStream_t s) {
    if (h->act32_valid[l]) return OCPPO_OK;
    const Layer &L = h->layers[l];
    OCPPO_CHECK_ARG(h->act16_valid[l], "activation %d exists in neither representation", l);
    OCPPO_TRY(launch_merge16(h->act16_hi[l], h->act16_lo[l], B * L.row_mult * L.out_dim, h->act_scale[l], h->act[l], h->sm_count, s));
    h->act32_valid[l] = 1;
    return OCPPO_OK;
}
static int ensure_act16(ocppo_handle *h, int l, int64_t B, cudaStream_t s) {
    if (h->act16_valid[l]) return OCPPO_OK;
    const Layer &L = h->layers[l];
    OCPPO_CHECK_ARG(h->act32_valid[l] && h->act16_hi[l] != nullptr, "activation %d cannot be split", l);
    OCPPO_TRY(ensure_act_amax(h, l, B, s));
    h->act_scale[l] = h->amax + kAmaxAct + l;                      // the true maximum
    OCPPO_TRY(launch_split16(h->act[l], B * L.row_mult * L.out_dim, h->act_scale[l], h->act16_hi[l], h->act16_lo[l], h->sm_count, s));
    h->act16_valid[l] = 1;
    return OCPPO_OK;
}
static bool pair_ok(const ocppo_handle *h, int l, int64_t B) {      // act[l] can live as a pair
    const Layer &L = h->layers[l];
    return h->act16_hi[l] != nullptr && (B * L.row_mult * L.out_dim) % 4 == 0 && L.out_dim % 8 == 0;
}

static void fwd_args(const ocppo_handle *h, int l, int64_t B, const float *in, SgemmArgs *a) {
    const Layer &L = h->layers[l];
    a->A = in; a->lda = L.in_dim;
    a->B = h->params + L.w_off; a->ldb = L.in_dim;
    a->C = h->act[l]; a->ldc = L.out_dim;
    a->bias = h->params + L.b_off;
    a->M = (int)(B * L.row_mult); a->N = L.out_dim; a->K = L.in_dim;
    a->k_per_split = round_up(L.in_dim, 8); a->splits = 1;
}

static int trunk_forward(ocppo_handle *h, const float *x, int64_t B, cudaStream_t s) {
    const float *in = x;
    const bool f16 = g_gemm_mode >= 5, pre16 = g_gemm_mode >= 6;
    // operand scales of the FP16x3 engine: activations are tracked by the epilogue that writes them (slots zeroed here);
    // the weights are measured, split and normed now -- the arena is the caller's to modify between calls -- unless the
    // caller vouched for an unchanged arena (ocppo_weights_version) and this engine prepared it last
    const bool prepared = f16 && h->w_version >= 0 && h->w_prepared_mode == g_gemm_mode && h->w_amax_valid;
    if (f16 && prepared) {
        OCPPO_CUDA(cudaMemsetAsync(h->amax, 0, kAmaxW * sizeof(float), s));
    } else if (f16) {
        OCPPO_CUDA(cudaMemsetAsync(h->amax, 0, kAmaxGrad * sizeof(float), s));
        AbsmaxSegs segs{};
        for (int l = 0; l < h->n_layers; ++l) {
            const Layer &L = h->layers[l];
            segs.x[l] = h->params + L.w_off; segs.n[l] = (int64_t)L.out_dim * L.in_dim; segs.out[l] = h->amax + kAmaxW + l;
        }
        segs.count = h->n_layers;
        if (pre16) {        // the weight norms the epilogues bound their fp16-pair outputs with, and max |W| in the same pass
            OCPPO_CUDA(cudaMemsetAsync(h->amax + kWRowL1, 0, (kAmaxSlots - kWRowL1) * sizeof(float), s));
            WnormSegs w{};
            for (int l = 0; l < h->n_layers; ++l) {
                const Layer &L = h->layers[l];
                w.w[l] = h->params + L.w_off; w.b[l] = h->params + L.b_off; w.out_dim[l] = L.out_dim; w.in_dim[l] = L.in_dim;
                w.row_l1[l] = h->amax + kWRowL1 + l; w.col_l1[l] = h->amax + kWColL1 + l; w.b_max[l] = h->amax + kBMax + l;
                w.w_amax[l] = h->amax + kAmaxW + l;
            }
            w.count = h->n_layers;
            OCPPO_TRY(launch_wnorm(w, s));
        } else {
            OCPPO_TRY(launch_absmax_multi(segs, s));
        }
        OCPPO_TRY(launch_split_f16_multi(segs, h->params, h->w16_hi, h->w16_lo, s));
        h->w_amax_valid = 1;
        h->w_prepared_mode = g_gemm_mode;
    } else {
        h->w_amax_valid = 0;
    }
    for (int l = 0; l < h->n_layers; ++l) {
        const Layer &L = h->layers[l];
        h->act_amax_valid[l] = 0;
        h->act16_valid[l] = 0; h->act32_valid[l] = 0;
        SgemmArgs a{};
        fwd_args(h, l, B, in, &a);
        a.relu_bits = h->training ? h->relu_bits[l] : nullptr;
        const bool w16 = L.w_off % 8 == 0 && L.in_dim % 8 == 0;
        a.want_pre = pre16 && l > 0 && w16 && pair_ok(h, l - 1, B);     // both operands will arrive as pairs (below)
        const bool tc = sgemm_uses_tc(SGEMM_FWD, a);
        h->relu_bits_valid[l] = a.relu_bits != nullptr && tc;
        bool pair_out = false;
        if (l == 0) h->x0p_valid = 0;
        if (pre16 && l == 0 && h->x0p_hi != nullptr && h->w_amax_valid) {
            // wide object vectors (SpaceInvaders 88, Breakout 220 columns): the FMA route costs 0.5 - 1.2 ms forward and
            // 2.8 ms backward per minibatch here.  Input and weight rows are zero-padded to one 256-wide tile in pair
            // storage instead, and the layer runs on the conversion-free tensor-core kernel like the others.
            const int64_t rows0 = B * L.row_mult;
            OCPPO_TRY(launch_absmax(x, rows0, L.in_dim, L.in_dim, h->amax + kAmaxObs, h->sm_count, s));
            OCPPO_TRY(launch_pad_split16(x, rows0, L.in_dim, L.in_dim, h->amax + kAmaxObs, h->x0p_hi, h->x0p_lo, h->sm_count, s));
            if (!prepared)
                OCPPO_TRY(launch_pad_split16(h->params + L.w_off, L.out_dim, L.in_dim, L.in_dim, h->amax + kAmaxW, h->w0p_hi,
                                             h->w0p_lo, h->sm_count, s));
            a.A = nullptr; a.lda = kPadCols; a.ldb = kPadCols; a.K = kPadCols; a.k_per_split = kPadCols;
            a.a16_hi = h->x0p_hi; a.a16_lo = h->x0p_lo; a.amax_a = h->amax + kAmaxObs;
            a.b16_hi = h->w0p_hi; a.b16_lo = h->w0p_lo; a.amax_b = h->amax + kAmaxW;
            a.amax_scratch = h->amax;
            a.amax_out = h->amax + kAmaxAct;
            OCPPO_CHECK_ARG(sgemm_uses_tc(SGEMM_FWD, a), "padded first layer rejected by the tensor-core kernel");
            h->relu_bits_valid[0] = a.relu_bits != nullptr;
            SgemmArgs n{};
            fwd_args(h, 1, B, h->act[0], &n);
            n.want_pre = pair_ok(h, 0, B) && h->layers[1].w_off % 8 == 0;
            if (pair_ok(h, 0, B) && sgemm_uses_tc(SGEMM_FWD, n)) {
                a.c16_hi = h->act16_hi[0]; a.c16_lo = h->act16_lo[0];
                a.bnd_amax = h->amax + kAmaxObs; a.bnd_wl1 = h->amax + kWRowL1; a.bnd_bmax = h->amax + kBMax;
                a.c_scale_src = h->amax + kBndAct;
                pair_out = true;
            }
            h->x0p_valid = 1;
        } else
        if (f16) {
            a.amax_scratch = h->amax;
            a.amax_b = h->amax + kAmaxW + l;
            if (w16) {
                a.b16_hi = reinterpret_cast<const uint16_t *>(h->w16_hi) + L.w_off;
                a.b16_lo = reinterpret_cast<const uint16_t *>(h->w16_lo) + L.w_off;
            }
            a.amax_out = h->amax + kAmaxAct + l;          // both kernels track it
            if (pre16 && tc && l > 0 && w16 && pair_ok(h, l - 1, B)) {
                // the input arrives as an fp16 pair (written by the layer below, or split here once when an FMA layer made it)
                OCPPO_TRY(ensure_act16(h, l - 1, B, s));
                a.a16_hi = h->act16_hi[l - 1]; a.a16_lo = h->act16_lo[l - 1];
                a.amax_a = h->act_scale[l - 1];
                bool next_tc = false;
                if (l + 1 < h->n_layers && pair_ok(h, l, B)) {
                    SgemmArgs n{};
                    fwd_args(h, l + 1, B, h->act[l], &n);
                    n.want_pre = h->layers[l + 1].w_off % 8 == 0 && L.out_dim % 8 == 0;
                    next_tc = sgemm_uses_tc(SGEMM_FWD, n);
                }
                if (next_tc) {                            // ... and so does the output: no fp32 copy of it is ever written
                    OCPPO_TRY(ensure_act_amax(h, l - 1, B, s));
                    a.c16_hi = h->act16_hi[l]; a.c16_lo = h->act16_lo[l];
                    a.bnd_amax = h->amax + kAmaxAct + l - 1;
                    a.bnd_wl1 = h->amax + kWRowL1 + l; a.bnd_bmax = h->amax + kBMax + l;
                    a.c_scale_src = h->amax + kBndAct + l;
                    pair_out = true;
                }
            } else {
                if (l > 0) OCPPO_TRY(ensure_act_fp32(h, l - 1, B, s));
                if (tc && l > 0) {
                    OCPPO_TRY(ensure_act_amax(h, l - 1, B, s));
                    a.amax_a = h->amax + kAmaxAct + l - 1;
                }
                if (pre16 && !tc && l + 1 < h->n_layers && pair_ok(h, l, B) && L.out_dim % 128 == 0) {
                    // an FMA layer (the K = frames x features first layer) feeding a tensor-core one: its epilogue writes the
                    // fp16 pair and the ReLU decision bits itself, and no fp32 copy of the activation exists
                    SgemmArgs n{};
                    fwd_args(h, l + 1, B, h->act[l], &n);
                    n.want_pre = h->layers[l + 1].w_off % 8 == 0 && L.out_dim % 8 == 0;
                    if (sgemm_uses_tc(SGEMM_FWD, n)) {
                        if (l == 0) OCPPO_TRY(launch_absmax(x, B * L.row_mult, L.in_dim, L.in_dim, h->amax + kAmaxObs, h->sm_count, s));
                        else OCPPO_TRY(ensure_act_amax(h, l - 1, B, s));
                        a.c16_hi = h->act16_hi[l]; a.c16_lo = h->act16_lo[l];
                        a.bnd_amax = l == 0 ? h->amax + kAmaxObs : h->amax + kAmaxAct + l - 1;
                        a.bnd_wl1 = h->amax + kWRowL1 + l; a.bnd_bmax = h->amax + kBMax + l;
                        a.c_scale_src = h->amax + kBndAct + l;
                        a.fma_bits = a.relu_bits != nullptr;
                        h->relu_bits_valid[l] = a.relu_bits != nullptr;
                        pair_out = true;
                    }
                }
            }
        }
        OCPPO_TRY(sgemm_launch(SGEMM_FWD, a, s));
        if (f16) h->act_amax_valid[l] = 1;
        if (pair_out) { h->act16_valid[l] = 1; h->act_scale[l] = h->amax + kBndAct + l; }
        else h->act32_valid[l] = 1;
        in = h->act[l];
    }
    h->last_forward_rows = (int)B;
    return OCPPO_OK;
}

static void pick_split(const ocppo_handle *h, int M, int N, int64_t K, int *splits, int64_t *k_per, bool pre = false) {
    // N tile: 256, or 128 for the conversion-free kernel (`pre`: both operands will be fp16 pairs) when N % 256 != 0
    const int tn = N % 256 == 0 ? 256 : ((pre && g_gemm_mode >= 6 && N % 128 == 0) ? 128 : 0);
    const bool tc = g_gemm_mode >= 1 && M % 128 == 0 && tn != 0;
    int64_t S;
    const int64_t s_cap0 = K / 256 > 1 ? K / 256 : 1;
    const int64_t s_cap = s_cap0 < kSplitMax ? s_cap0 : kSplitMax;
    if (tc) {
        // persistent kernel, one CTA pair per 2 SMs, 256 x tn tiles: the split count that fills the last wave best
        // (fewest splits on ties -- less partial traffic); accuracy does not depend on it, the kernel
        // drains its accumulator every 1024 rows.
        const int64_t tiles = ceil_div(M, 256) * (N / tn), sms = h->sm_count / 2;   // one tile per CTA pair at a time
        S = 1;
        double best = 0.0;
        for (int64_t c = 1; c <= s_cap; ++c) {
            const int64_t w = tiles * c;
            const double util = (double)w / (double)(ceil_div(w, sms) * sms);
            if (util > best + 1e-9) { best = util; S = c; }
        }
    } else {
        // FMA path: 128x128 tiles, 2 CTAs per SM, aim at two full waves
        const int64_t tiles = ceil_div(M, 128) * ceil_div(N, 128);
        S = ceil_div((int64_t)h->sm_count * 4, tiles);
        if (S > s_cap) S = s_cap;
    }
    if (S < 1) S = 1;
    int64_t kp = round_up(ceil_div(K, S), 32);
    if (kp < 32) kp = 32;
    *k_per = kp;
    *splits = (int)ceil_div(K, kp);
    if (*splits < 1) *splits = 1;
}

// dW[out,in] (+)= dY[rows,out]^T X[rows,in];  db[out] (+)= colsum(dY)
static int wgrad(ocppo_handle *h, const float *dY, int out, const float *X, int in, int64_t rows, float *dW,
                 int64_t w_count, float *db, int64_t b_count, int accumulate, cudaStream_t s,
                 const float *amax_dy = nullptr, const float *amax_x = nullptr) {
    SgemmArgs a{};
    a.A = dY; a.lda = out;
    a.B = X; a.ldb = in;
    a.C = h->splitk; a.ldc = in;
    a.bias_part = h->splitk_bias;
    a.M = out; a.N = in; a.K = rows;
    a.amax_a = amax_dy; a.amax_b = amax_x; a.amax_scratch = h->amax;
    pick_split(h, out, in, rows, &a.splits, &a.k_per_split);
    if (in <= 32 || (in <= kSkinnyMaxIn && !sgemm_uses_tc(SGEMM_WGRAD, a))) {
        // skinny contraction (the first trunk layer: in = object-vector width): one thread per dY column, 32 input columns
        // per pass -- HBM-bound for Pong's 6 columns, at the FMA peak for SpaceInvaders' 88 / Breakout's 220 (3 / 7 passes;
        // the 128 x 128-tiled FMA GEMM took 3x as long on these)
        float *part = h->skinny, *wsum = h->skinny + (int64_t)h->sm_count * 4 * out * 32;
        for (int c0 = 0; c0 < in; c0 += 32) {
            const int n = in - c0 < 32 ? in - c0 : 32;
            int G = 0, NS = 0;
            OCPPO_TRY(launch_skinny_wgrad(dY, out, out, X + c0, in, n, rows, h->sm_count, part, c0 == 0 ? wsum : nullptr, nullptr, &G, &NS, s));
            const SkinnyReduceJob jobs[2] = {{part, G, out, NS, 0, n, dW + c0, 1, in, accumulate},
                                             {wsum, G, out, 1, 0, 1, db, 1, 1, accumulate}};
            OCPPO_TRY(launch_skinny_reduce_multi(jobs, c0 == 0 ? 2 : 1, s));
        }
        return OCPPO_OK;
    }
    OCPPO_TRY(sgemm_launch(SGEMM_WGRAD, a, s));
    OCPPO_TRY(launch_splitk_reduce(h->splitk, a.splits, (int64_t)out * in, w_count, dW, accumulate, h->sm_count, s));
    OCPPO_TRY(launch_splitk_reduce(h->splitk_bias, a.splits, out, b_count, db, accumulate, h->sm_count, s));
    return OCPPO_OK;
}

// fp16-pair WGRAD (engine 6): dY and X both as pairs -> the conversion-free kernel; the bias gradient is a column-sum pass
static void wgrad_args(const ocppo_handle *h, const float *dY, int out, const float *X, int in, int64_t rows, SgemmArgs *a) {
    a->A = dY; a->lda = out;
    a->B = X; a->ldb = in;
    a->C = h->splitk; a->ldc = in;
    a->M = out; a->N = in; a->K = rows;
    a->want_pre = 1;                  // planned for the conversion-free kernel; the caller launches it only with both pairs
    pick_split(h, out, in, rows, &a->splits, &a->k_per_split, true);
}

// g0 holds d loss / d pre-activation of the last trunk layer, [B, hidden].
// top_pair_scale != nullptr: the loss head left the top gradient as an fp16 pair in g16[cur] (scale from this device scalar)
// and `top_colsum_blocks` partial rows of its column sums in colsum_part; no fp32 copy of it exists.
static int trunk_backward(ocppo_handle *h, const float *x, int64_t B, int cur, int accumulate, cudaStream_t s,
                          bool top_amax_tracked = false, const float *top_pair_scale = nullptr, int top_colsum_blocks = 0) {
    const bool f16 = g_gemm_mode >= 5, pre16 = g_gemm_mode >= 6;
    bool dy_amax_valid = f16 && top_amax_tracked;   // the caller zeroed the gradient slots and the loss head filled the top one
    if (f16 && !top_amax_tracked) {
        // max |dY| of the top layer comes from one reduction pass; the dgrad epilogues track the layers below
        OCPPO_CUDA(cudaMemsetAsync(h->amax + kAmaxGrad, 0, 2 * OCPPO_MAX_LAYERS * sizeof(float), s));
        const Layer &LT = h->layers[h->n_layers - 1];
        OCPPO_TRY(launch_absmax(h->gbuf[cur], B * LT.row_mult, LT.out_dim, LT.out_dim, h->amax + kAmaxGrad + h->n_layers - 1,
                                h->sm_count, s));
        dy_amax_valid = true;
    }
    // which representations of the gradient entering layer l exist: fp32 in gbuf[cur], fp16 pair in g16[cur]
    bool dy32 = top_pair_scale == nullptr, dy16 = top_pair_scale != nullptr;
    bool db_done = false;                 // the dgrad above already formed this layer's bias gradient (fused column sums) ...
    int pending_parts = 0;                // ... as this many partial rows in colsum_part that still have to be folded
    const float *dy_scale = top_pair_scale;
    for (int l = h->n_layers - 1; l >= 0; --l) {
        const Layer &L = h->layers[l];
        const int64_t rows = B * L.row_mult;
        const float *X = (l == 0) ? x : h->act[l - 1];
        const float *dY = h->gbuf[cur];
        const float *amax_dy = (f16 && dy_amax_valid) ? h->amax + kAmaxGrad + l : nullptr;
        const bool w16 = L.w_off % 8 == 0 && L.in_dim % 8 == 0;
        const bool dy_pair_ok = pre16 && dy_amax_valid && (rows * L.out_dim) % 4 == 0 && L.out_dim % 8 == 0;

        // ---- plan: which of this layer's two GEMMs take fp16 pairs
        bool pre_w = false, pre_d = false;
        SgemmArgs wa{}, da{};
        const bool pad0 = l == 0 && pre16 && h->x0p_valid;         // first layer on padded pair storage (see trunk_forward)
        if (l > 0 && L.in_dim > 32) {
            wgrad_args(h, dY, L.out_dim, X, L.in_dim, rows, &wa);
            pre_w = dy_pair_ok && sgemm_uses_tc(SGEMM_WGRAD, wa) && L.out_dim % 64 == 0 && pair_ok(h, l - 1, B);
        } else if (pad0) {
            wgrad_args(h, dY, L.out_dim, nullptr, kPadCols, rows, &wa);
            pre_w = dy_pair_ok && sgemm_uses_tc(SGEMM_WGRAD, wa) && L.out_dim % 64 == 0;
        }
        if (l > 0) {
            da.A = dY; da.lda = L.out_dim;
            da.B = h->params + L.w_off; da.ldb = L.in_dim;
            da.C = h->gbuf[cur ^ 1]; da.ldc = L.in_dim;
            da.mask = h->act[l - 1];
            da.M = (int)rows; da.N = L.in_dim; da.K = L.out_dim;
            da.k_per_split = round_up(L.out_dim, 8); da.splits = 1;
            // the layer below left its ReLU decisions as one bit per element (same row-major element order, so the
            // encoder -> decoder flatten does not matter): the tensor-core dgrad reads those instead of the activations
            if (h->relu_bits_valid[l - 1] && L.in_dim % 128 == 0) da.relu_bits = h->relu_bits[l - 1];
            da.want_pre = dy_pair_ok && w16 && h->w_amax_valid;
            pre_d = da.want_pre && sgemm_uses_tc(SGEMM_DGRAD, da);
        }
        OCPPO_CHECK_ARG(!db_done || pre_w, "layer %d: bias gradient formed twice", l);
        if ((pre_w || pre_d) && !dy16) {        // the loss head's (or an FMA layer's) fp32 gradient, split once for both GEMMs
            dy_scale = amax_dy;
            if (pre_w && !db_done && L.out_dim % 4 == 0) {      // ... and the bias gradient falls out of the same read
                int blocks = 0;
                OCPPO_TRY(launch_split16_colsum(dY, rows, L.out_dim, dy_scale, h->g16_hi[cur], h->g16_lo[cur], h->colsum_part, &blocks,
                                                h->sm_count, s));
                OCPPO_TRY(launch_colsum_reduce(h->colsum_part, blocks, L.out_dim, h->grads + L.b_off, accumulate, s));
                db_done = true;
            } else {
                OCPPO_TRY(launch_split16(dY, rows * L.out_dim, dy_scale, h->g16_hi[cur], h->g16_lo[cur], h->sm_count, s));
            }
            dy16 = true;
        }
        if ((!pre_w || (l > 0 && !pre_d)) && !dy32) {
            OCPPO_TRY(launch_merge16(h->g16_hi[cur], h->g16_lo[cur], rows * L.out_dim, dy_scale, h->gbuf[cur], h->sm_count, s));
            dy32 = true;
        }

        // ---- weight and bias gradients
        if (pre_w) {
            wa.a16_hi = h->g16_hi[cur]; wa.a16_lo = h->g16_lo[cur]; wa.amax_a = dy_scale;
            wa.amax_scratch = h->amax;
            if (l == 0) {          // padded input pair; the partial tiles are kPadCols wide, the gradient rows in_dim
                wa.b16_hi = h->x0p_hi; wa.b16_lo = h->x0p_lo; wa.amax_b = h->amax + kAmaxObs;
                OCPPO_TRY(sgemm_launch(SGEMM_WGRAD, wa, s));
                OCPPO_TRY(launch_splitk_reduce_pad(h->splitk, wa.splits, L.out_dim, L.in_dim, h->grads + L.w_off, accumulate,
                                                   h->sm_count, s));
            } else {
                OCPPO_TRY(ensure_act16(h, l - 1, B, s));
                wa.b16_hi = h->act16_hi[l - 1]; wa.b16_lo = h->act16_lo[l - 1]; wa.amax_b = h->act_scale[l - 1];
                OCPPO_TRY(sgemm_launch(SGEMM_WGRAD, wa, s));
                // the bias gradient's column-sum partials are at hand -- left by the loss head, by the dgrad above
                // (pending_parts) or formed now -- and fold in the same launch as the split-K partials
                int blocks = pending_parts;
                if (!db_done) {
                    if (l == h->n_layers - 1 && top_colsum_blocks > 0) blocks = top_colsum_blocks;      // left by the loss head
                    else OCPPO_TRY(launch_colsum16(h->g16_hi[cur], h->g16_lo[cur], rows, L.out_dim, dy_scale, h->colsum_part, &blocks, h->sm_count, s));
                }
                if (blocks > 0) {
                    OCPPO_TRY(launch_wgrad_fold(h->splitk, wa.splits, (int64_t)L.out_dim * L.in_dim, (int64_t)L.out_dim * L.in_dim,
                                                h->grads + L.w_off, accumulate, h->colsum_part, blocks, L.out_dim,
                                                h->grads + L.b_off, accumulate, h->sm_count, s));
                    db_done = true;
                } else {
                    OCPPO_TRY(launch_splitk_reduce(h->splitk, wa.splits, (int64_t)L.out_dim * L.in_dim, (int64_t)L.out_dim * L.in_dim,
                                                   h->grads + L.w_off, accumulate, h->sm_count, s));
                }
                pending_parts = 0;
            }
            if (!db_done) {
                int blocks = 0;
                if (l == h->n_layers - 1 && top_colsum_blocks > 0) blocks = top_colsum_blocks;      // left by the loss head
                else OCPPO_TRY(launch_colsum16(h->g16_hi[cur], h->g16_lo[cur], rows, L.out_dim, dy_scale, h->colsum_part, &blocks, h->sm_count, s));
                OCPPO_TRY(launch_colsum_reduce(h->colsum_part, blocks, L.out_dim, h->grads + L.b_off, accumulate, s));
            }
            db_done = false;
        } else {
            const float *amax_x = nullptr;
            if (l > 0) OCPPO_TRY(ensure_act_fp32(h, l - 1, B, s));
            if (f16 && l > 0 && L.in_dim > 32) {
                OCPPO_TRY(ensure_act_amax(h, l - 1, B, s));
                amax_x = h->amax + kAmaxAct + l - 1;
            }
            OCPPO_TRY(wgrad(h, dY, L.out_dim, X, L.in_dim, rows, h->grads + L.w_off, (int64_t)L.out_dim * L.in_dim,
                            h->grads + L.b_off, L.out_dim, accumulate, s, amax_dy, amax_x));
        }

        // ---- gradient of the layer's input
        if (l > 0) {
            SgemmArgs &a = da;
            bool pair_out = false;
            if (da.relu_bits == nullptr) OCPPO_TRY(ensure_act_fp32(h, l - 1, B, s));      // fp32 mask
            if (f16) {
                a.amax_scratch = h->amax;
                a.amax_a = amax_dy;
                if (h->w_amax_valid) {
                    a.amax_b = h->amax + kAmaxW + l;
                    if (w16) {
                        a.b16_hi = reinterpret_cast<const uint16_t *>(h->w16_hi) + L.w_off;
                        a.b16_lo = reinterpret_cast<const uint16_t *>(h->w16_lo) + L.w_off;
                    }
                }
                dy_amax_valid = sgemm_uses_tc(SGEMM_DGRAD, a);
                if (dy_amax_valid) a.amax_out = h->amax + kAmaxGrad + l - 1;
                if (pre_d) {
                    a.a16_hi = h->g16_hi[cur]; a.a16_lo = h->g16_lo[cur]; a.amax_a = dy_scale;
                    // the gradient leaves as a pair when both GEMMs of the layer below will take it that way
                    const Layer &Ln = h->layers[l - 1];
                    bool next_pair = Ln.out_dim % 64 == 0 && (rows * L.in_dim) % 4 == 0 &&
                                     ((l - 1 >= 1 && Ln.in_dim > 32 && pair_ok(h, l - 2, B)) || (l - 1 == 0 && h->x0p_valid));
                    if (next_pair) {
                        SgemmArgs nw{};
                        wgrad_args(h, h->gbuf[cur ^ 1], Ln.out_dim, nullptr, l - 1 == 0 ? kPadCols : Ln.in_dim, B * Ln.row_mult, &nw);
                        next_pair = sgemm_uses_tc(SGEMM_WGRAD, nw);
                    }
                    if (next_pair) {
                        a.c16_hi = h->g16_hi[cur ^ 1]; a.c16_lo = h->g16_lo[cur ^ 1];
                        a.bnd_amax = h->amax + kAmaxGrad + l; a.bnd_wl1 = h->amax + kWColL1 + l; a.bnd_bmax = nullptr;
                        a.c_scale_src = h->amax + kBndGrad + l - 1;
                        pair_out = true;
                        // ... and its column sums (the bias gradient of the layer below) come out of the same epilogue
                        const int parts = tc_gemm_ctas(SGEMM_DGRAD, a, h->sm_count) * 4;
                        if (parts <= kColsumBlocksMax && L.in_dim % Ln.out_dim == 0) {
                            a.colsum_part = h->colsum_part;
                            OCPPO_CUDA(cudaMemsetAsync(h->colsum_part, 0, (size_t)parts * L.in_dim * sizeof(float), s));
                        }
                    }
                }
            }
            OCPPO_TRY(sgemm_launch(SGEMM_DGRAD, a, s));
            if (a.colsum_part != nullptr) {
                // [parts, in] partial rows read as [parts * in / out', out']: the flatten's frame groups fold with the rest
                const Layer &Ln = h->layers[l - 1];
                const int parts = tc_gemm_ctas(SGEMM_DGRAD, a, h->sm_count) * 4 * (L.in_dim / Ln.out_dim);
                // layers above the first fold these partials together with their weight gradient's (launch_wgrad_fold, after
                // the layer's WGRAD and before its DGRAD reuses colsum_part); the padded first layer folds them now
                if (l - 1 > 0) pending_parts = parts;
                else OCPPO_TRY(launch_colsum_reduce(h->colsum_part, parts, Ln.out_dim, h->grads + Ln.b_off, accumulate, s));
                db_done = true;
            }
            if (pair_out) { dy16 = true; dy32 = false; dy_scale = h->amax + kBndGrad + l - 1; }
            else { dy16 = false; dy32 = true; }
            cur ^= 1;
        }
    }
    return OCPPO_OK;
}

}  // namespace ocppo

using namespace ocppo;

extern "C" {

const char *ocppo_last_error(void) { return g_err; }
int ocppo_version(void) { return 200; }
int ocppo_bringup_build(void) {
#ifdef OCPPO_BRINGUP
    return 1;
#else
    return 0;
#endif
}

int64_t ocppo_launch_count(void) {
    int64_t t = 0;
    for (int i = 0; i < KC_COUNT; ++i) t += g_prof.launches[i];
    return t;
}

int ocppo_set_gemm_mode(int32_t mode) {
    OCPPO_CHECK_ARG((mode >= 0 && mode <= 6) || mode == 8, "gemm mode must be 0 (fp32 FMA), 1 / 2 (tcgen05 3xTF32, split / single accumulator), 3 / 4 (TF32 + BF16 corrections, split / single), 5 (FP16 x 3, fp32 storage) or 6 (FP16 x 3, fp16-pair storage)");
    g_gemm_mode = mode;
    return OCPPO_OK;
}
int ocppo_get_gemm_mode(void) { return g_gemm_mode; }

int ocppo_gemm(int32_t op, int32_t engine, int32_t M, int32_t N, int64_t K, const float *A, int64_t lda,
               const float *B, int64_t ldb, float *C, int64_t ldc, const float *bias, const float *mask,
               float *bias_part, int32_t splits, int64_t k_per_split, void *stream) {
    OCPPO_CHECK_ARG(op >= 0 && op <= 2 && engine >= 0 && engine <= 8, "bad op/engine");
    OCPPO_CHECK_ARG(A && B && C, "null pointer");
    DeviceGuard dg(device_of(C));
    SgemmArgs a{};
    a.A = A; a.lda = lda; a.B = B; a.ldb = ldb; a.C = C; a.ldc = ldc; a.bias = bias; a.mask = mask;
    a.bias_part = bias_part; a.M = M; a.N = N; a.K = K; a.splits = splits < 1 ? 1 : splits;
    a.k_per_split = k_per_split > 0 ? k_per_split : round_up(K, 32);
    if (engine >= 1) {
        cudaStream_t s = as_stream(stream);
        sgemm_engine_flags(engine, &a);
        float *scratch = nullptr;
        uint16_t *pairs = nullptr;
        int r = OCPPO_OK;
        if (engine >= 5) {              // test hook: operand maxima (and, engines 6 / 7, the fp16 pairs) live in transient allocations
            OCPPO_CUDA(cudaMalloc(&scratch, 8 * sizeof(float)));
            OCPPO_CUDA(cudaMemsetAsync(scratch, 0, 8 * sizeof(float), s));
            a.amax_scratch = scratch;
            r = sgemm_fill_amax((SgemmMode)op, &a, s);
        }
        if (r == OCPPO_OK && engine >= 6) {
            // engine 6: A and B split into pairs, fp32 C.  engine 7 (FWD / DGRAD): C leaves as a pair too and is merged back.
            const int64_t a_rows = op == 2 ? K : M, a_cols = op == 2 ? M : K;
            const int64_t b_rows = op == 0 ? N : K, b_cols = op == 0 ? K : N;
            const int64_t na = a_rows * a_cols, nb = b_rows * b_cols, nc = (int64_t)M * N;
            const bool ok = lda == a_cols && ldb == b_cols && ldc == N && na % 8 == 0 && nb % 8 == 0 && nc % 8 == 0;
            if (!ok) { set_error("ocppo_gemm: engines 6 / 7 take contiguous operands"); r = OCPPO_EINVAL; }
            a.bias_part = nullptr;      // WGRAD bias sums: a column-sum pass over the pair (below), not the GEMM kernel
            const int64_t n_part = (op == 2 && bias_part != nullptr) ? (int64_t)kColsumBlocksMax * M * 2 : 0;   // floats, as uint16 pairs
            if (r == OCPPO_OK && cudaMalloc(&pairs, (size_t)(2 * (na + nb + nc) + 2 * n_part) * sizeof(uint16_t)) != cudaSuccess) {
                set_error("ocppo_gemm: transient allocation failed"); r = OCPPO_ECUDA;
            }
            if (r == OCPPO_OK) {
                uint16_t *a_hi = pairs, *a_lo = pairs + na, *b_hi = a_lo + na, *b_lo = b_hi + nb, *c_hi = b_lo + nb, *c_lo = c_hi + nc;
                r = launch_split16(A, na, a.amax_a, a_hi, a_lo, 0, s);
                if (r == OCPPO_OK) r = launch_split16(B, nb, a.amax_b, b_hi, b_lo, 0, s);
                a.a16_hi = a_hi; a.a16_lo = a_lo; a.b16_hi = b_hi; a.b16_lo = b_lo;
                if (r == OCPPO_OK && engine == 7 && op != 2) {
                    WnormSegs w{};
                    w.w[0] = B; w.b[0] = op == 0 ? bias : nullptr; w.out_dim[0] = (int)b_rows; w.in_dim[0] = (int)b_cols;
                    w.row_l1[0] = scratch + 2; w.col_l1[0] = scratch + 3; w.b_max[0] = scratch + 4; w.count = 1;
                    r = launch_wnorm(w, s);
                    a.c16_hi = c_hi; a.c16_lo = c_lo;
                    a.bnd_amax = a.amax_a; a.bnd_wl1 = op == 0 ? scratch + 2 : scratch + 3; a.bnd_bmax = op == 0 && bias ? scratch + 4 : nullptr;
                    a.c_scale_src = scratch + 5;
                }
                if (r == OCPPO_OK) r = tc_gemm_launch((SgemmMode)op, a, s);
                if (r == OCPPO_OK && a.c16_hi != nullptr) r = launch_merge16(c_hi, c_lo, nc, scratch + 5, C, 0, s);
                if (r == OCPPO_OK && n_part > 0) {          // bias_part[0] = column sums of A, the other splits' rows zero
                    float *part = reinterpret_cast<float *>(c_lo + nc);
                    int blocks = 0;
                    cudaMemsetAsync(bias_part, 0, (size_t)a.splits * M * sizeof(float), s);
                    r = launch_colsum16(a_hi, a_lo, K, M, a.amax_a, part, &blocks, 0, s);
                    if (r == OCPPO_OK) r = launch_colsum_reduce(part, blocks, M, bias_part, 0, s);
                }
            }
        } else if (r == OCPPO_OK) {
            r = tc_gemm_launch((SgemmMode)op, a, s);
        }
        if (scratch != nullptr || pairs != nullptr) { cudaStreamSynchronize(s); cudaFree(scratch); cudaFree(pairs); }
        return r;
    }
    return fma_gemm_launch((SgemmMode)op, a, as_stream(stream));
}

int ocppo_measure_fma_peak(double *tflops, void *stream) {
    OCPPO_CHECK_ARG(tflops != nullptr, "null argument");
    return measure_fma_peak(tflops, as_stream(stream));
}

int ocppo_profile_enable(int32_t on) {
    if (on && !g_prof.created) {
        for (int i = 0; i < kProfRing; ++i) {
            OCPPO_CUDA(cudaEventCreate(&g_prof.start[i]));
            OCPPO_CUDA(cudaEventCreate(&g_prof.stop[i]));
        }
        g_prof.created = true;
    }
    if (g_prof.n > 0) OCPPO_TRY(prof_drain());
    g_prof.on = on != 0;
    g_prof.open = false;
    for (int i = 0; i < KC_COUNT; ++i) { g_prof.ms[i] = 0; g_prof.timed[i] = 0; g_prof.skipped[i] = 0; }
    return OCPPO_OK;
}

int ocppo_profile_read(double *ms_by_class, int64_t *timed_by_class, int32_t n_classes) {
    OCPPO_CHECK_ARG(ms_by_class != nullptr && n_classes >= KC_COUNT, "need %d class slots", (int)KC_COUNT);
    OCPPO_TRY(prof_drain());
    for (int i = 0; i < n_classes; ++i) {
        ms_by_class[i] = i < KC_COUNT ? g_prof.ms[i] : 0;
        if (timed_by_class) timed_by_class[i] = i < KC_COUNT ? g_prof.timed[i] : 0;
    }
    for (int i = 0; i < KC_COUNT; ++i) { g_prof.ms[i] = 0; g_prof.timed[i] = 0; }
    return OCPPO_OK;
}

int ocppo_launch_counts(int64_t *launches_by_class, int64_t *untimed_by_class, int32_t n_classes) {
    OCPPO_CHECK_ARG(launches_by_class != nullptr && n_classes >= KC_COUNT, "need %d class slots", (int)KC_COUNT);
    for (int i = 0; i < n_classes; ++i) {
        launches_by_class[i] = i < KC_COUNT ? g_prof.launches[i] : 0;
        if (untimed_by_class) untimed_by_class[i] = i < KC_COUNT ? g_prof.skipped[i] : 0;
    }
    return OCPPO_OK;
}

int ocppo_create(const ocppo_config *cfg, ocppo_handle **out) {
    OCPPO_CHECK_ARG(cfg != nullptr && out != nullptr, "null argument");
    ocppo_handle *h = new (std::nothrow) ocppo_handle();
    OCPPO_CHECK_ARG(h != nullptr, "out of host memory");
    memset(h, 0, sizeof(*h));
    h->cfg = *cfg;
    int r = build_layout(h);
    if (r != OCPPO_OK) { delete h; return r; }
    h->sm_count = 148;
    h->max_threads_per_sm = 2048;
    if (cfg->device >= 0) {
        cudaDeviceProp prop;
        cudaError_t e = cudaGetDeviceProperties(&prop, cfg->device);
        if (e != cudaSuccess) {
            set_error("cudaGetDeviceProperties(%d): %s", cfg->device, cudaGetErrorString(e));
            delete h;
            return OCPPO_ECUDA;
        }
        if (prop.major != 10) {
            set_error("device %d is sm_%d%d; libocppo is built for sm_100a only (no fallback)", cfg->device,
                      prop.major, prop.minor);
            delete h;
            return OCPPO_EARCH;
        }
        h->sm_count = prop.multiProcessorCount;
        h->max_threads_per_sm = prop.maxThreadsPerMultiProcessor;
    }
    *out = h;
    return OCPPO_OK;
}

int ocppo_destroy(ocppo_handle *h) {
    delete h;
    return OCPPO_OK;
}

int ocppo_num_tensors(const ocppo_handle *h, int32_t *n_tensors, int64_t *arena_floats, int64_t *param_floats) {
    OCPPO_CHECK_ARG(h != nullptr, "null handle");
    if (n_tensors) *n_tensors = h->n_tensors;
    if (arena_floats) *arena_floats = h->arena_floats;
    if (param_floats) *param_floats = h->param_floats;
    return OCPPO_OK;
}

int ocppo_tensor_layout(const ocppo_handle *h, int32_t i, int64_t *offset, int32_t *rows, int32_t *cols) {
    OCPPO_CHECK_ARG(h != nullptr, "null handle");
    OCPPO_CHECK_ARG(i >= 0 && i < h->n_tensors, "tensor index %d out of range [0,%d)", i, h->n_tensors);
    if (offset) *offset = h->tensor_off[i];
    if (rows) *rows = h->tensor_rows[i];
    if (cols) *cols = h->tensor_cols[i];
    return OCPPO_OK;
}

int ocppo_bind_arena(ocppo_handle *h, float *params, float *grads, float *exp_avg, float *exp_avg_sq) {
    OCPPO_CHECK_ARG(h != nullptr && params != nullptr, "null argument");
    OCPPO_CHECK_ARG(((uintptr_t)params & 127) == 0 && ((uintptr_t)grads & 127) == 0 &&
                        ((uintptr_t)exp_avg & 127) == 0 && ((uintptr_t)exp_avg_sq & 127) == 0,
                    "arenas must be 128-byte aligned");
    h->params = params; h->grads = grads; h->exp_avg = exp_avg; h->exp_avg_sq = exp_avg_sq;
    h->w_amax_valid = 0; h->w_version = -1; h->w_prepared_mode = -1;
    return OCPPO_OK;
}

int ocppo_weights_version(ocppo_handle *h, int64_t version) {
    OCPPO_CHECK_ARG(h != nullptr, "null handle");
    if (version < 0 || version != h->w_version) h->w_amax_valid = 0;     // changed (or no promise): prepare again
    h->w_version = version;
    return OCPPO_OK;
}

int ocppo_weights_prepared(const ocppo_handle *h) {
    if (h == nullptr) return 0;
    return g_gemm_mode >= 5 && h->w_version >= 0 && h->w_prepared_mode == g_gemm_mode && h->w_amax_valid;
}

// Host-side bookkeeping a forward / backward / optimizer call leaves in the handle (which representation of each
// activation is valid, whether the weights are prepared, ...).  A CUDA graph replays the device work of a captured
// call but not this; the caller saves it right after capture and restores it after every replay.
namespace {
struct HostState {
    int w_amax_valid, w_prepared_mode, last_forward_rows, x0p_valid;
    int act_amax_valid[2 * OCPPO_MAX_LAYERS], act16_valid[2 * OCPPO_MAX_LAYERS], act32_valid[2 * OCPPO_MAX_LAYERS];
    int relu_bits_valid[2 * OCPPO_MAX_LAYERS];
    const float *act_scale[2 * OCPPO_MAX_LAYERS];
};
}  // namespace

int64_t ocppo_host_state_bytes(void) { return (int64_t)sizeof(HostState); }

int ocppo_host_state_save(const ocppo_handle *h, void *buf, int64_t bytes) {
    OCPPO_CHECK_ARG(h != nullptr && buf != nullptr && bytes >= (int64_t)sizeof(HostState), "host_state_save: bad argument");
    HostState st{};
    st.w_amax_valid = h->w_amax_valid; st.w_prepared_mode = h->w_prepared_mode;
    st.last_forward_rows = h->last_forward_rows; st.x0p_valid = h->x0p_valid;
    for (int l = 0; l < 2 * OCPPO_MAX_LAYERS; ++l) {
        st.act_amax_valid[l] = h->act_amax_valid[l]; st.act16_valid[l] = h->act16_valid[l];
        st.act32_valid[l] = h->act32_valid[l]; st.relu_bits_valid[l] = h->relu_bits_valid[l];
        st.act_scale[l] = h->act_scale[l];
    }
    memcpy(buf, &st, sizeof(st));
    return OCPPO_OK;
}

int ocppo_host_state_restore(ocppo_handle *h, const void *buf, int64_t bytes) {
    OCPPO_CHECK_ARG(h != nullptr && buf != nullptr && bytes >= (int64_t)sizeof(HostState), "host_state_restore: bad argument");
    HostState st;
    memcpy(&st, buf, sizeof(st));
    h->w_amax_valid = st.w_amax_valid; h->w_prepared_mode = st.w_prepared_mode;
    h->last_forward_rows = st.last_forward_rows; h->x0p_valid = st.x0p_valid;
    for (int l = 0; l < 2 * OCPPO_MAX_LAYERS; ++l) {
        h->act_amax_valid[l] = st.act_amax_valid[l]; h->act16_valid[l] = st.act16_valid[l];
        h->act32_valid[l] = st.act32_valid[l]; h->relu_bits_valid[l] = st.relu_bits_valid[l];
        h->act_scale[l] = st.act_scale[l];
    }
    return OCPPO_OK;
}

int ocppo_workspace_bytes(const ocppo_handle *h, int64_t max_rows, int32_t training, int64_t *bytes) {
    OCPPO_CHECK_ARG(h != nullptr && bytes != nullptr, "null argument");
    OCPPO_CHECK_ARG(max_rows >= 1, "max_rows must be >= 1");
    OCPPO_CHECK_ARG(max_rows * h->cfg.frames <= 65535LL * 128, "max_rows too large (%lld)", (long long)max_rows);
    *bytes = carve(const_cast<ocppo_handle *>(h), nullptr, max_rows, training, false);
    return OCPPO_OK;
}

int ocppo_bind_workspace(ocppo_handle *h, void *workspace, int64_t bytes, int64_t max_rows, int32_t training) {
    OCPPO_CHECK_ARG(h != nullptr && workspace != nullptr, "null argument");
    OCPPO_CHECK_ARG(((uintptr_t)workspace & 255) == 0, "workspace must be 256-byte aligned");
    int64_t need = 0;
    OCPPO_TRY(ocppo_workspace_bytes(h, max_rows, training, &need));
    OCPPO_CHECK_ARG(bytes >= need, "workspace too small: %lld < %lld", (long long)bytes, (long long)need);
    carve(h, workspace, max_rows, training, true);
    h->ws = workspace; h->ws_bytes = bytes; h->max_rows = max_rows; h->training = training;
    h->last_forward_rows = 0;
    return OCPPO_OK;
}

static int policy_step(ocppo_handle *h, const float *obs, int64_t B, uint64_t seed, uint64_t philox_offset,
                       const uint64_t *philox_dev, const int64_t *given_action, int64_t *action, float *logp,
                       float *entropy, float *value, void *stream) {
    OCPPO_TRY(require_ready(h, false));
    DeviceGuard dg(h->cfg.device);
    OCPPO_CHECK_ARG(B >= 0 && (obs != nullptr || B == 0), "bad obs/B");
    cudaStream_t s = as_stream(stream);
    const int64_t row_floats = (int64_t)h->cfg.frames * h->cfg.obs_dim;
    for (int64_t c0 = 0; c0 < B; c0 += h->max_rows) {
        const int64_t rows = B - c0 < h->max_rows ? B - c0 : h->max_rows;
        OCPPO_TRY(trunk_forward(h, obs + c0 * row_floats, rows, s));
        OCPPO_TRY(launch_policy_head(h, h->act[h->n_layers - 1], rows, c0, B, seed, philox_offset, philox_dev,
                                     given_action ? given_action + c0 : nullptr, action ? action + c0 : nullptr,
                                     logp ? logp + c0 : nullptr, entropy ? entropy + c0 : nullptr,
                                     value ? value + c0 : nullptr, 1, s));
    }
    return OCPPO_OK;
}

int ocppo_policy_step(ocppo_handle *h, const float *obs, int64_t B, uint64_t seed, uint64_t philox_offset,
                      const int64_t *given_action, int64_t *action, float *logp, float *entropy, float *value,
                      void *stream) {
    return policy_step(h, obs, B, seed, philox_offset, nullptr, given_action, action, logp, entropy, value, stream);
}

int ocppo_policy_step_dev(ocppo_handle *h, const float *obs, int64_t B, const uint64_t *philox_state,
                          const int64_t *given_action, int64_t *action, float *logp, float *entropy, float *value,
                          void *stream) {
    OCPPO_CHECK_ARG(philox_state != nullptr || given_action != nullptr, "policy_step_dev: null philox_state");
    return policy_step(h, obs, B, 0, 0, philox_state, given_action, action, logp, entropy, value, stream);
}

int ocppo_value(ocppo_handle *h, const float *obs, int64_t B, float *value, void *stream) {
    OCPPO_TRY(require_ready(h, false));
    DeviceGuard dg(h->cfg.device);
    OCPPO_CHECK_ARG(B >= 0 && value != nullptr && (obs != nullptr || B == 0), "bad argument");
    cudaStream_t s = as_stream(stream);
    const int64_t row_floats = (int64_t)h->cfg.frames * h->cfg.obs_dim;
    for (int64_t c0 = 0; c0 < B; c0 += h->max_rows) {
        const int64_t rows = B - c0 < h->max_rows ? B - c0 : h->max_rows;
        OCPPO_TRY(trunk_forward(h, obs + c0 * row_floats, rows, s));
        OCPPO_TRY(launch_policy_head(h, h->act[h->n_layers - 1], rows, c0, B, 0, 0, nullptr, nullptr, nullptr, nullptr,
                                     nullptr, value + c0, 0, s));
    }
    return OCPPO_OK;
}

int ocppo_categorical(const float *logits, int64_t B, int32_t A, uint64_t seed, uint64_t philox_offset,
                      const int64_t *given_action, int64_t *action, float *logp, float *entropy, void *stream) {
    OCPPO_CHECK_ARG(B >= 0 && A >= 1 && A <= 32, "bad B/A");
    OCPPO_CHECK_ARG(logits != nullptr || B == 0, "null logits");
    DeviceGuard dg(device_of(logits));
    return launch_categorical(logits, B, A, seed, philox_offset, given_action, action, logp, entropy,
                              as_stream(stream));
}

uint64_t ocppo_philox_increment(int64_t numel, int32_t sm_count, int32_t max_threads_per_sm) {
    if (numel <= 0) return 0;
    uint64_t grid = ((uint64_t)numel + 255) / 256;
    const uint64_t cap = (uint64_t)sm_count * (uint64_t)(max_threads_per_sm / 256);
    if (grid > cap) grid = cap;
    return (((uint64_t)numel - 1) / (256 * grid * 4) + 1) * 4;
}

int ocppo_exponential(float *out, int64_t numel, uint64_t seed, uint64_t philox_offset, void *stream) {
    OCPPO_CHECK_ARG(numel >= 0 && (out != nullptr || numel == 0), "bad argument");
    DeviceGuard dg(device_of(out));
    return launch_exponential(out, numel, seed, philox_offset, as_stream(stream));
}

static int gae_entry(const float *rewards, const float *values, const float *dones, const float *next_value,
                     const float *next_done, int64_t T, int64_t N, double gamma, double gae_lambda, float *advantages,
                     float *returns, int overlap_previous, void *stream) {
    OCPPO_CHECK_ARG(T >= 0 && N >= 0, "negative T/N");
    if (T == 0 || N == 0) return OCPPO_OK;
    OCPPO_CHECK_ARG(rewards && values && dones && next_value && next_done && advantages && returns,
                    "null pointer");
    DeviceGuard dg(device_of(rewards));
    return launch_gae(rewards, values, dones, next_value, next_done, T, N, (float)gamma,
                      (float)(gamma * gae_lambda), advantages, returns, overlap_previous, as_stream(stream));
}

int ocppo_gae(const float *rewards, const float *values, const float *dones, const float *next_value,
              const float *next_done, int64_t T, int64_t N, double gamma, double gae_lambda, float *advantages,
              float *returns, void *stream) {
    return gae_entry(rewards, values, dones, next_value, next_done, T, N, gamma, gae_lambda, advantages, returns, 0, stream);
}

int ocppo_gae_overlapped(const float *rewards, const float *values, const float *dones, const float *next_value,
                         const float *next_done, int64_t T, int64_t N, double gamma, double gae_lambda,
                         float *advantages, float *returns, void *stream) {
    return gae_entry(rewards, values, dones, next_value, next_done, T, N, gamma, gae_lambda, advantages, returns, 1, stream);
}

int ocppo_set_gae_variant(int32_t variant) {
    OCPPO_CHECK_ARG(variant == 0 || variant == 1, "gae variant must be 0 (TMA kernel where eligible) or 1 (cp.async kernel)");
    g_gae_variant = variant;
    return OCPPO_OK;
}

int ocppo_minibatch_fwd_bwd(ocppo_handle *h, const float *b_obs, const int64_t *b_actions, const float *b_logprobs,
                            const float *b_advantages, const float *b_returns, const float *b_values,
                            const int64_t *idx, int64_t mb, const ocppo_hparams *hp, float *stats, void *stream) {
    OCPPO_TRY(require_ready(h, true));
    DeviceGuard dg(h->cfg.device);
    OCPPO_CHECK_ARG(mb >= 1, "minibatch size must be >= 1");
    OCPPO_CHECK_ARG(b_obs && b_actions && b_logprobs && b_advantages && b_returns && b_values && idx && hp && stats,
                    "null pointer");
    cudaStream_t s = as_stream(stream);
    const int A = h->cfg.n_actions, H = h->hidden, AP = (int)round_up(A + 1, 4);
    if (hp->norm_adv) OCPPO_TRY(launch_adv_stats(h, b_advantages, idx, mb, s));
    const int64_t n_chunks = ceil_div(mb, h->max_rows);
    OCPPO_CHECK_ARG(n_chunks <= kLossBlocksMax, "too many chunks");
    int used = 0;
    for (int64_t ci = 0; ci < n_chunks; ++ci) {
        const int64_t c0 = ci * h->max_rows;
        const int64_t rows = mb - c0 < h->max_rows ? mb - c0 : h->max_rows;
        OCPPO_TRY(launch_gather(h, b_obs, b_actions, b_logprobs, b_advantages, b_returns, b_values, idx + c0, rows, s));
        OCPPO_TRY(trunk_forward(h, h->x0, rows, s));
        LossArgs a{};
        a.hidden = h->act[h->n_layers - 1];
        a.action = h->mb_action; a.logp_old = h->mb_logp; a.adv = h->mb_adv; a.ret = h->mb_ret; a.val = h->mb_val;
        a.adv_stat = h->adv_stat;
        a.Wa = h->params + h->actor.w_off; a.ba = h->params + h->actor.b_off;
        a.Wc = h->params + h->critic.w_off; a.bc = h->params + h->critic.b_off;
        a.rows = rows; a.mb = mb; a.H = H; a.A = A; a.AP = AP;
        a.clip_lo = (float)(1.0 - hp->clip_coef); a.clip_hi = (float)(1.0 + hp->clip_coef);
        a.clip_f = (float)hp->clip_coef; a.ent_coef = (float)hp->ent_coef; a.vf_coef = (float)hp->vf_coef;
        a.norm_adv = hp->norm_adv; a.clip_vloss = hp->clip_vloss;
        a.dh = h->gbuf[0]; a.gz = h->gz;
        const bool f16 = g_gemm_mode >= 5;
        // fp16-pair storage: the loss head runs as two kernels -- forward (gz, statistics, the bound ingredient), then
        // d(hidden) written straight as the pair the top layer's dgrad / wgrad load, with its column sums -- whenever
        // the top layer's backward will run on pairs; otherwise one kernel leaves the fp32 gradient as before
        const Layer &LT = h->layers[h->n_layers - 1];
        const bool pair_top = g_gemm_mode >= 6 && h->n_layers >= 2 && loss_head_pair_supported(H, A) && h->g16_hi[0] != nullptr &&
                              LT.out_dim % 64 == 0 && h->w_amax_valid && pair_ok(h, h->n_layers - 2, rows) &&
                              (((uintptr_t)a.hidden) & 15) == 0;
        if (f16) {
            OCPPO_CUDA(cudaMemsetAsync(h->amax + kAmaxGrad, 0, 2 * OCPPO_MAX_LAYERS * sizeof(float), s));
            a.dh_amax = h->amax + kAmaxGrad + h->n_layers - 1;
        }
        a.part = h->loss_part + (int64_t)used * 8;
        int blocks = 0, top_colsum_blocks = 0;
        const int avail = (int)((kLossBlocksMax - used) / (n_chunks - ci));
        if (pair_top) {
            OCPPO_CUDA(cudaMemsetAsync(h->amax, 0, sizeof(float), s));           // scratch slot 0: max (|dv| + sum |dz|)
            a.gz_l1max = h->amax;
            OCPPO_TRY(launch_loss_head(h, a, &blocks, avail, s, true));
            LossBwdPairArgs b{};
            b.gz = h->gz;
            b.relu_bits = h->relu_bits_valid[h->n_layers - 1] ? h->relu_bits[h->n_layers - 1] : nullptr;
            b.hidden = a.hidden;
            b.Wa = a.Wa; b.Wc = a.Wc; b.gz_l1max = h->amax;
            b.rows = rows; b.H = H; b.A = A; b.AP = AP;
            b.dh_hi = h->g16_hi[0]; b.dh_lo = h->g16_lo[0];
            b.bound = h->amax + kBndGrad + h->n_layers - 1;
            b.dh_amax = a.dh_amax;
            b.colsum_part = h->colsum_part;
            OCPPO_TRY(launch_loss_head_bwd_pair(h, b, &top_colsum_blocks, s));
        } else {
            OCPPO_TRY(launch_loss_head(h, a, &blocks, avail, s));
        }
        used += blocks;
        const int acc = ci > 0;
        // heads: d[Wa;Wc] = gz^T hidden, d[ba;bc] = colsum(gz)
        if (A + 1 <= 32) {
            int G = 0, NS = 0;
            float *part = h->skinny, *ssum = h->skinny + (int64_t)h->sm_count * 4 * H * 32;
            OCPPO_TRY(launch_skinny_wgrad(a.hidden, H, H, h->gz, AP, A + 1, rows, h->sm_count, part, nullptr, ssum, &G, &NS, s));
            const SkinnyReduceJob jobs[4] = {{part, G, H, NS, 0, A, h->grads + h->actor.w_off, H, 1, acc},
                                             {part, G, H, NS, A, A + 1, h->grads + h->critic.w_off, H, 1, acc},
                                             {ssum, G, 1, NS, 0, A, h->grads + h->actor.b_off, 1, 0, acc},
                                             {ssum, G, 1, NS, A, A + 1, h->grads + h->critic.b_off, 1, 0, acc}};
            OCPPO_TRY(launch_skinny_reduce_multi(jobs, 4, s));
        } else
        {
            SgemmArgs g{};
            g.A = h->gz; g.lda = AP;
            g.B = a.hidden; g.ldb = H;
            g.C = h->splitk; g.ldc = H;
            g.bias_part = h->splitk_bias;
            g.M = AP; g.N = H; g.K = rows;
            pick_split(h, AP, H, rows, &g.splits, &g.k_per_split);
            OCPPO_TRY(sgemm_launch(SGEMM_WGRAD, g, s));
            const int64_t stride = (int64_t)AP * H;
            OCPPO_TRY(launch_splitk_reduce(h->splitk, g.splits, stride, (int64_t)A * H, h->grads + h->actor.w_off, acc, h->sm_count, s));
            OCPPO_TRY(launch_splitk_reduce(h->splitk + (int64_t)A * H, g.splits, stride, H, h->grads + h->critic.w_off, acc, h->sm_count, s));
            OCPPO_TRY(launch_splitk_reduce(h->splitk_bias, g.splits, AP, A, h->grads + h->actor.b_off, acc, h->sm_count, s));
            OCPPO_TRY(launch_splitk_reduce(h->splitk_bias + A, g.splits, AP, 1, h->grads + h->critic.b_off, acc, h->sm_count, s));
        }
        OCPPO_TRY(trunk_backward(h, h->x0, rows, 0, acc, s, g_gemm_mode >= 5,
                                 pair_top ? h->amax + kBndGrad + h->n_layers - 1 : nullptr, top_colsum_blocks));
    }
    OCPPO_TRY(launch_loss_finalize(h->loss_part, used, mb, (float)hp->ent_coef, (float)hp->vf_coef, stats, s));
    return OCPPO_OK;
}

int ocppo_clip_adam(ocppo_handle *h, float inv_world, float max_grad_norm, double lr, double beta1, double beta2,
                    double eps, int64_t step, float *grad_norm, void *stream) {
    OCPPO_TRY(require_ready(h, true));
    DeviceGuard dg(h->cfg.device);
    OCPPO_CHECK_ARG(step >= 1, "Adam step is 1-based");
    OCPPO_CHECK_ARG(h->grads && h->exp_avg && h->exp_avg_sq, "grad/Adam arenas not bound");
    h->w_amax_valid = 0;
    float sc[8];
    adam_scalars(inv_world, max_grad_norm, lr, beta1, beta2, eps, step, sc);
    return launch_clip_adam(h, sc, nullptr, grad_norm, as_stream(stream));
}

int ocppo_adam_scalars_set(float *scalars8, float inv_world, float max_grad_norm, double lr, double beta1, double beta2,
                           double eps, int64_t step, void *stream) {
    OCPPO_CHECK_ARG(scalars8 != nullptr && step >= 1, "adam_scalars_set: bad argument");
    DeviceGuard dg(device_of(scalars8));
    float sc[8];
    adam_scalars(inv_world, max_grad_norm, lr, beta1, beta2, eps, step, sc);
    return launch_store_words(scalars8, sc, 8, as_stream(stream));
}

int ocppo_philox_set(uint64_t *philox_state, uint64_t seed, uint64_t philox_offset, void *stream) {
    OCPPO_CHECK_ARG(philox_state != nullptr, "philox_set: null pointer");
    DeviceGuard dg(device_of(philox_state));
    const uint64_t w[2] = {seed, philox_offset};
    return launch_store_words(philox_state, w, 4, as_stream(stream));
}

int ocppo_clip_adam_dev(ocppo_handle *h, const float *scalars8, float *grad_norm, void *stream) {
    OCPPO_TRY(require_ready(h, true));
    DeviceGuard dg(h->cfg.device);
    OCPPO_CHECK_ARG(scalars8 != nullptr, "clip_adam_dev: null scalars");
    OCPPO_CHECK_ARG(h->grads && h->exp_avg && h->exp_avg_sq, "grad/Adam arenas not bound");
    h->w_amax_valid = 0;
    return launch_clip_adam(h, nullptr, scalars8, grad_norm, as_stream(stream));
}

int ocppo_explained_variance(ocppo_handle *h, const float *values, const float *returns, int64_t n, float *out2,
                             void *stream) {
    OCPPO_TRY(require_ready(h, true));
    DeviceGuard dg(h->cfg.device);
    OCPPO_CHECK_ARG(values && returns && out2 && n >= 1, "explained_variance: bad argument");
    return launch_explained_var(h, values, returns, n, out2, as_stream(stream));
}

int ocppo_trunk_forward(ocppo_handle *h, const float *obs, int64_t B, float *hidden, void *stream) {
    OCPPO_TRY(require_ready(h, false));
    DeviceGuard dg(h->cfg.device);
    OCPPO_CHECK_ARG(B >= 1 && B <= h->max_rows, "trunk_forward: B=%lld must be in [1,%lld]", (long long)B,
                    (long long)h->max_rows);
    OCPPO_CHECK_ARG(obs && hidden, "null pointer");
    cudaStream_t s = as_stream(stream);
    OCPPO_TRY(trunk_forward(h, obs, B, s));
    OCPPO_CUDA(cudaMemcpyAsync(hidden, h->act[h->n_layers - 1], (size_t)B * h->hidden * sizeof(float),
                               cudaMemcpyDeviceToDevice, s));
    return OCPPO_OK;
}

int ocppo_trunk_backward(ocppo_handle *h, const float *obs, int64_t B, const float *d_hidden, void *stream) {
    OCPPO_TRY(require_ready(h, true));
    DeviceGuard dg(h->cfg.device);
    OCPPO_CHECK_ARG(B >= 1 && B == h->last_forward_rows, "trunk_backward: B=%lld does not match last forward (%d)",
                    (long long)B, h->last_forward_rows);
    OCPPO_CHECK_ARG(obs && d_hidden, "null pointer");
    cudaStream_t s = as_stream(stream);
    OCPPO_TRY(launch_relu_mask(d_hidden, h->act[h->n_layers - 1], B * h->hidden, h->gbuf[0], h->sm_count, s));
    return trunk_backward(h, obs, B, 0, 0, s);
}

int ocppo_read_activation(const ocppo_handle *h, int32_t layer, int64_t B, float *out, void *stream) {
    OCPPO_TRY(require_ready(h, false));
    DeviceGuard dg(h->cfg.device);
    OCPPO_CHECK_ARG(out, "null pointer");
    OCPPO_CHECK_ARG(layer >= 0 && layer < h->n_layers, "read_activation: layer %d not in [0,%d)", layer, h->n_layers);
    OCPPO_CHECK_ARG(B >= 1 && B == h->last_forward_rows, "read_activation: B=%lld does not match last forward (%d)",
                    (long long)B, h->last_forward_rows);
    const Layer &L = h->layers[layer];
    if (!h->act32_valid[layer] && h->act16_valid[layer])     // fp16-pair storage: the tensor only exists as (hi, lo)
        return launch_merge16(h->act16_hi[layer], h->act16_lo[layer], B * L.row_mult * L.out_dim, h->act_scale[layer], out,
                              h->sm_count, as_stream(stream));
    OCPPO_CUDA(cudaMemcpyAsync(out, h->act[layer], (size_t)B * L.row_mult * L.out_dim * sizeof(float),
                               cudaMemcpyDeviceToDevice, as_stream(stream)));
    return OCPPO_OK;
}

}  // extern "C"
