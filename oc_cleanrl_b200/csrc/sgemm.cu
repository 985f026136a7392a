// FP32 CUDA-core GEMM family for the PPObj trunk (the reference runs full-fp32
// cuBLAS SGEMM through nn.Linear, architectures/ppo.py:70-84; loss.backward(),
// ppo_atari_oc.py:608).  One templated 128x128x8 register-tiled kernel covers
//   FWD   : Y[M,N]  = relu(X[M,K] . W[N,K]^T + b[N])
//   DGRAD : dX[M,N] = (mask[M,N] > 0) * (dY[M,K] . W[K,N])
//   WGRAD : dW[M,N] = dY[K,M]^T . X[K,N], split over K into deterministic
//           partials (+ column sums of dY for the bias gradient)
// Fixed accumulation order => run-to-run deterministic (the reference sets
// torch.use_deterministic_algorithms(True), ppo_atari_oc.py:209).
#include "common.cuh"
#include <stdlib.h>

namespace ocppo {

namespace {

constexpr int BM = 128, BN = 128, BK = 8, PAD = 4;
constexpr int kThreads = 256;

// Load one BMxBK (or BNxBK) operand slab into registers.
//  KC  : element (r, k) at base[r*ld + k]   (k contiguous)
//  !KC : element (r, k) at base[k*ld + r]   (r contiguous)
template <bool KC, bool VEC>
__device__ __forceinline__ void load_slab(float (&reg)[4], const float *__restrict__ base, int64_t ld,
                                          int r0, int R, int64_t k0, int64_t Kend, int tid) {
    if (KC) {
        const int r = r0 + (tid >> 1);
        const int64_t k = k0 + (tid & 1) * 4;
        if (VEC) {
            if (r < R && k < Kend) {
                const float4 v = *reinterpret_cast<const float4 *>(base + (int64_t)r * ld + k);
                reg[0] = v.x; reg[1] = v.y; reg[2] = v.z; reg[3] = v.w;
            } else {
                reg[0] = reg[1] = reg[2] = reg[3] = 0.f;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
                reg[i] = (r < R && k + i < Kend) ? base[(int64_t)r * ld + k + i] : 0.f;
        }
    } else {
        const int64_t k = k0 + (tid >> 5);
        const int r = r0 + (tid & 31) * 4;
        if (VEC) {
            if (r < R && k < Kend) {
                const float4 v = *reinterpret_cast<const float4 *>(base + k * ld + r);
                reg[0] = v.x; reg[1] = v.y; reg[2] = v.z; reg[3] = v.w;
            } else {
                reg[0] = reg[1] = reg[2] = reg[3] = 0.f;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i)
                reg[i] = (r + i < R && k < Kend) ? base[k * ld + r + i] : 0.f;
        }
    }
}

template <bool KC>
__device__ __forceinline__ void store_slab(float (*s)[BM + PAD], const float (&reg)[4], int tid) {
    if (KC) {
        const int r = tid >> 1, k = (tid & 1) * 4;
#pragma unroll
        for (int i = 0; i < 4; ++i) s[k + i][r] = reg[i];
    } else {
        const int k = tid >> 5, r = (tid & 31) * 4;
        *reinterpret_cast<float4 *>(&s[k][r]) = make_float4(reg[0], reg[1], reg[2], reg[3]);
    }
}

template <int MODE, bool A_VEC, bool B_VEC>
__global__ void __launch_bounds__(kThreads, 2) sgemm_kernel(SgemmArgs p) {
    constexpr bool A_KC = (MODE != SGEMM_WGRAD);
    constexpr bool B_KC = (MODE == SGEMM_FWD);
    __shared__ __align__(16) float As[2][BK][BM + PAD];
    __shared__ __align__(16) float Bs[2][BK][BN + PAD];

    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int split = blockIdx.z;
    const int64_t kbeg = (int64_t)split * p.k_per_split;
    const int64_t kend = (kbeg + p.k_per_split < p.K) ? kbeg + p.k_per_split : p.K;

    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
    float bsum[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) bsum[i] = 0.f;
    const bool do_bias = (MODE == SGEMM_WGRAD) && blockIdx.x == 0 && tx == 0 && p.bias_part != nullptr;

    float ra[4], rb[4];
    const int64_t n_tiles = (kend - kbeg + BK - 1) / BK;
    if (n_tiles > 0) {
        load_slab<A_KC, A_VEC>(ra, p.A, p.lda, m0, p.M, kbeg, kend, tid);
        load_slab<B_KC, B_VEC>(rb, p.B, p.ldb, n0, p.N, kbeg, kend, tid);
        store_slab<A_KC>(As[0], ra, tid);
        store_slab<B_KC>(Bs[0], rb, tid);
    }
    __syncthreads();

    for (int64_t t = 0; t < n_tiles; ++t) {
        const int cur = (int)(t & 1);
        if (t + 1 < n_tiles) {
            const int64_t k0 = kbeg + (t + 1) * BK;
            load_slab<A_KC, A_VEC>(ra, p.A, p.lda, m0, p.M, k0, kend, tid);
            load_slab<B_KC, B_VEC>(rb, p.B, p.ldb, n0, p.N, k0, kend, tid);
        }
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            const float4 a0 = *reinterpret_cast<const float4 *>(&As[cur][kk][ty * 4]);
            const float4 a1 = *reinterpret_cast<const float4 *>(&As[cur][kk][64 + ty * 4]);
            const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[cur][kk][tx * 4]);
            const float4 b1 = *reinterpret_cast<const float4 *>(&Bs[cur][kk][64 + tx * 4]);
            const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
            if (MODE == SGEMM_WGRAD) {
                if (do_bias) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) bsum[i] += a[i];
                }
            }
        }
        if (t + 1 < n_tiles) {
            store_slab<A_KC>(As[cur ^ 1], ra, tid);
            store_slab<B_KC>(Bs[cur ^ 1], rb, tid);
        }
        __syncthreads();
    }

    // ---- epilogue ----
    float *C = p.C;
    if (MODE == SGEMM_WGRAD) C += (int64_t)split * p.M * p.N;
    const bool c_vec = ((p.N & 3) == 0) && ((p.ldc & 3) == 0);
    // fp16-pair storage (engine 6 / 8; the launcher only asks for it on full 128-column tiles): the FWD output leaves as
    // (hi, lo) fp16 scaled by its a-priori bound plus one ReLU decision bit per element, and no fp32 copy is written
    const bool pair = (MODE == SGEMM_FWD) && p.c16_hi != nullptr;
    const bool bits_out = (MODE == SGEMM_FWD) && p.relu_bits != nullptr && p.fma_bits;
    float scale_c = 1.f;
    if (pair) {
        const float bound = __ldg(p.bnd_amax) * __ldg(p.bnd_wl1) + (p.bnd_bmax != nullptr ? __ldg(p.bnd_bmax) : 0.f);
        float inv_c;
        f16_scale_of(bound, scale_c, inv_c);
        if (blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) *p.c_scale_src = bound;
    }
    float run_max = 0.f;                  // max |C| of this thread's stores (p.amax_out: the consumer's FP16 scale)
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (bits_out) {
            // the 8 threads tx = 8 w .. 8 w + 7 of a row hold the 32 columns of one word: OR their nibbles together
            // (all lanes take part; rows / columns past the edge contribute zeros and store nothing)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int n = n0 + h * 64 + tx * 4;
                uint32_t word = 0;
                if (m < p.M && n < p.N) {
                    const float4 bb = *reinterpret_cast<const float4 *>(p.bias + n);
                    word = ((acc[i][h * 4 + 0] + bb.x > 0.f ? 1u : 0u) | (acc[i][h * 4 + 1] + bb.y > 0.f ? 2u : 0u) |
                            (acc[i][h * 4 + 2] + bb.z > 0.f ? 4u : 0u) | (acc[i][h * 4 + 3] + bb.w > 0.f ? 8u : 0u)) << (4 * (tx & 7));
                }
                word |= __shfl_xor_sync(0xffffffffu, word, 1);
                word |= __shfl_xor_sync(0xffffffffu, word, 2);
                word |= __shfl_xor_sync(0xffffffffu, word, 4);
                if ((tx & 7) == 0 && m < p.M && n < p.N) p.relu_bits[(int64_t)m * (p.N >> 5) + (n >> 5)] = word;
            }
        }
        if (m >= p.M) continue;
        if (MODE == SGEMM_WGRAD && do_bias) p.bias_part[(int64_t)split * p.M + m] = bsum[i];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int n = n0 + h * 64 + tx * 4;
            if (n >= p.N) continue;
            float v[4] = {acc[i][h * 4 + 0], acc[i][h * 4 + 1], acc[i][h * 4 + 2], acc[i][h * 4 + 3]};
            float *dst = C + (int64_t)m * p.ldc + n;
            if (c_vec) {
                if (MODE == SGEMM_FWD) {
                    const float4 bb = *reinterpret_cast<const float4 *>(p.bias + n);
                    v[0] = fmaxf(v[0] + bb.x, 0.f); v[1] = fmaxf(v[1] + bb.y, 0.f);
                    v[2] = fmaxf(v[2] + bb.z, 0.f); v[3] = fmaxf(v[3] + bb.w, 0.f);
                    if (pair) {
                        uint32_t h0, l0, h1, l1;
                        split_f16x2(v[0], v[1], scale_c, h0, l0);
                        split_f16x2(v[2], v[3], scale_c, h1, l1);
                        const int64_t e = (int64_t)m * p.ldc + n;          // element offset; 8-byte aligned as fp16
                        *reinterpret_cast<uint2 *>(reinterpret_cast<uint16_t *>(p.c16_hi) + e) = make_uint2(h0, h1);
                        *reinterpret_cast<uint2 *>(reinterpret_cast<uint16_t *>(p.c16_lo) + e) = make_uint2(l0, l1);
                        run_max = fmaxf(fmaxf(run_max, fmaxf(fabsf(v[0]), fabsf(v[1]))), fmaxf(fabsf(v[2]), fabsf(v[3])));
                        continue;
                    }
                } else if (MODE == SGEMM_DGRAD) {
                    const float4 mk = *reinterpret_cast<const float4 *>(p.mask + (int64_t)m * p.ldc + n);
                    v[0] = mk.x > 0.f ? v[0] : 0.f; v[1] = mk.y > 0.f ? v[1] : 0.f;
                    v[2] = mk.z > 0.f ? v[2] : 0.f; v[3] = mk.w > 0.f ? v[3] : 0.f;
                }
                *reinterpret_cast<float4 *>(dst) = make_float4(v[0], v[1], v[2], v[3]);
                if (MODE != SGEMM_WGRAD) run_max = fmaxf(fmaxf(run_max, fmaxf(fabsf(v[0]), fabsf(v[1]))), fmaxf(fabsf(v[2]), fabsf(v[3])));
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if (n + j >= p.N) continue;
                    float o = v[j];
                    if (MODE == SGEMM_FWD) o = fmaxf(o + p.bias[n + j], 0.f);
                    else if (MODE == SGEMM_DGRAD) o = p.mask[(int64_t)m * p.ldc + n + j] > 0.f ? o : 0.f;
                    dst[j] = o;
                    if (MODE != SGEMM_WGRAD) run_max = fmaxf(run_max, fabsf(o));
                }
            }
        }
    }
    if (MODE != SGEMM_WGRAD && p.amax_out != nullptr) {          // uniform across the grid
        const unsigned m = __reduce_max_sync(0xffffffffu, __float_as_uint(run_max));   // non-negative floats order as uints
        if ((tid & 31) == 0 && m != 0u) atomicMax(reinterpret_cast<unsigned int *>(p.amax_out), m);
    }
}

// First trunk layer with a handful of input features (Pong: K = frames' object vector of 6 floats), fp16-pair output.
// The 128x128x8 tile kernel above pads K to 8, stages both operands through shared memory and takes 210 us for the
// 524,288 x 256 layer of a full-size minibatch; the work is really one 537 MB write (HBM floor 85 us).  Here a thread keeps
// its four output columns' weights and biases in registers, reads the K inputs of a row as warp-uniform loads, and
// streams out (hi, lo) fp16 + the ReLU decision words.  Same products in the same order (k ascending from zero), so the
// activations are bit-identical to the tile kernel's.
// one row of four outputs per thread: products k ascending from zero, + bias, ReLU; the decision nibble ORed over the 8
// lanes of a 32-column word; (hi, lo) fp16 pair out
template <int KD>
__device__ __forceinline__ void thin_fwd_row(const float (&x)[KD], const float (&w)[4][KD], const float4 &bb, float scale_c,
                                             int lane8, bool live, uint32_t *bits, uint2 *hi, uint2 *lo, float &run_max) {
    // packed fp32 FMAs (FFMA2: two IEEE round-to-nearest FMAs per issue slot -- this kernel is issue-bound): columns (0, 1)
    // and (2, 3) share an instruction, same products in the same order as the scalar form
    float2 a01 = make_float2(0.f, 0.f), a23 = make_float2(0.f, 0.f);
#pragma unroll
    for (int k = 0; k < KD; ++k) {
        const float2 xx = make_float2(x[k], x[k]);
        a01 = __ffma2_rn(xx, make_float2(w[0][k], w[1][k]), a01);
        a23 = __ffma2_rn(xx, make_float2(w[2][k], w[3][k]), a23);
    }
    a01 = __fadd2_rn(a01, make_float2(bb.x, bb.y));
    a23 = __fadd2_rn(a23, make_float2(bb.z, bb.w));
    float v[4] = {fmaxf(a01.x, 0.f), fmaxf(a01.y, 0.f), fmaxf(a23.x, 0.f), fmaxf(a23.y, 0.f)};
    if (bits != nullptr) {                                        // (uniform) 8 lanes x 4 columns = one 32-column word
        uint32_t word = ((v[0] > 0.f ? 1u : 0u) | (v[1] > 0.f ? 2u : 0u) | (v[2] > 0.f ? 4u : 0u) | (v[3] > 0.f ? 8u : 0u))
                        << (4 * lane8);
        word |= __shfl_xor_sync(0xffffffffu, word, 1);
        word |= __shfl_xor_sync(0xffffffffu, word, 2);
        word |= __shfl_xor_sync(0xffffffffu, word, 4);
        if (lane8 == 0 && live) *bits = word;
    }
    if (!live) return;
    uint32_t h0, l0, h1, l1;
    split_f16x2(v[0], v[1], scale_c, h0, l0);
    split_f16x2(v[2], v[3], scale_c, h1, l1);
    __stcs(hi, make_uint2(h0, h1));
    __stcs(lo, make_uint2(l0, l1));
    run_max = fmaxf(fmaxf(run_max, fmaxf(v[0], v[1])), fmaxf(v[2], v[3]));
}

// FAST: K == KD (even), lda even, A 8-byte aligned and the block's row slab complete -- no per-element predicates, a row
// is read as KD / 2 warp-uniform 8-byte loads and every pointer advances by a constant.
template <int KD, bool FAST>
__device__ __forceinline__ void thin_fwd_rows(const SgemmArgs &p, const float (&w)[4][KD], const float4 &bb, float scale_c,
                                              int64_t r0, int64_t r1, int rl, int rpp, int c, int lane8, float &run_max) {
    const int wpr = p.N >> 5;
    uint16_t *const hi16 = reinterpret_cast<uint16_t *>(p.c16_hi), *const lo16 = reinterpret_cast<uint16_t *>(p.c16_lo);
    if (FAST) {
        const float2 *xa = reinterpret_cast<const float2 *>(p.A + (r0 + rl) * p.lda);
        const int64_t xrow = ((int64_t)rpp * p.lda) >> 1, crow = ((int64_t)rpp * p.ldc) >> 2, brow = (int64_t)rpp * wpr;
        uint2 *ha = reinterpret_cast<uint2 *>(hi16 + (r0 + rl) * p.ldc + c), *la = reinterpret_cast<uint2 *>(lo16 + (r0 + rl) * p.ldc + c);
        uint32_t *ba = p.relu_bits != nullptr ? p.relu_bits + (r0 + rl) * wpr + (c >> 5) : nullptr;
        // the two rows of the next pass are requested before this pass's arithmetic (ncu: long-scoreboard 5.7 per issue
        // without -- the loads are warp-uniform 24-byte rows, nothing but latency)
        float xn[2][KD];
        auto load_rows = [&](const float2 *src) {
#pragma unroll
            for (int u = 0; u < 2; ++u)
#pragma unroll
                for (int k = 0; k < KD; k += 2) {
                    const float2 t = __ldg(src + u * xrow + (k >> 1));
                    xn[u][k] = t.x; xn[u][k + 1] = t.y;
                }
        };
        if (r0 < r1) load_rows(xa);
        for (int64_t mb = r0; mb < r1; mb += 2 * rpp) {
            float x[2][KD];
#pragma unroll
            for (int u = 0; u < 2; ++u)
#pragma unroll
                for (int k = 0; k < KD; ++k) x[u][k] = xn[u][k];
            xa += 2 * xrow;
            if (mb + 2 * rpp < r1) load_rows(xa);
            thin_fwd_row<KD>(x[0], w, bb, scale_c, lane8, true, ba, ha, la, run_max);
            thin_fwd_row<KD>(x[1], w, bb, scale_c, lane8, true, ba != nullptr ? ba + brow : nullptr, ha + crow, la + crow, run_max);
            ha += 2 * crow; la += 2 * crow;
            if (ba != nullptr) ba += 2 * brow;
        }
        return;
    }
    for (int64_t mb = r0; mb < r1; mb += 2 * rpp) {               // warp-uniform trip count (a warp sits inside one row)
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int64_t m = mb + u * rpp + rl;
            const bool live = m < r1;
            float x[KD];
#pragma unroll
            for (int k = 0; k < KD; ++k) x[k] = (live && k < p.K) ? __ldg(p.A + m * p.lda + k) : 0.f;
            thin_fwd_row<KD>(x, w, bb, scale_c, lane8, live, p.relu_bits != nullptr ? p.relu_bits + m * wpr + (c >> 5) : nullptr,
                             reinterpret_cast<uint2 *>(hi16 + m * p.ldc + c), reinterpret_cast<uint2 *>(lo16 + m * p.ldc + c), run_max);
        }
    }
}

template <int KD>
__global__ void __launch_bounds__(256) thin_fwd_pair_kernel(SgemmArgs p, int64_t rows_per_cta, int fast_ok) {
    const int tid = threadIdx.x, tpr = p.N >> 2;                  // threads per row: N / 4 in {32, 64, 128, 256}
    const int rl = tid / tpr, c = (tid - rl * tpr) * 4, rpp = 256 / tpr;
    float w[4][KD];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int k = 0; k < KD; ++k) w[j][k] = k < p.K ? __ldg(p.B + (int64_t)(c + j) * p.ldb + k) : 0.f;
    const float4 bb = *reinterpret_cast<const float4 *>(p.bias + c);
    const float bound = __ldg(p.bnd_amax) * __ldg(p.bnd_wl1) + (p.bnd_bmax != nullptr ? __ldg(p.bnd_bmax) : 0.f);
    float scale_c, inv_c;
    f16_scale_of(bound, scale_c, inv_c);
    if (blockIdx.x == 0 && tid == 0) *p.c_scale_src = bound;
    float run_max = 0.f;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_cta;
    const int64_t r1 = r0 + rows_per_cta < p.M ? r0 + rows_per_cta : p.M;
    // (rows_per_cta is a multiple of the 2 * rpp rows of one pass, so a complete slab needs no row guards)
    if (fast_ok && r0 + rows_per_cta <= p.M) thin_fwd_rows<KD, true>(p, w, bb, scale_c, r0, r1, rl, rpp, c, tid & 7, run_max);
    else thin_fwd_rows<KD, false>(p, w, bb, scale_c, r0, r1, rl, rpp, c, tid & 7, run_max);
    if (p.amax_out != nullptr) {
        const unsigned mx = __reduce_max_sync(0xffffffffu, __float_as_uint(run_max));
        if ((tid & 31) == 0 && mx != 0u) atomicMax(reinterpret_cast<unsigned int *>(p.amax_out), mx);
    }
}

inline bool aligned16(const void *p) { return (((uintptr_t)p) & 15) == 0; }

template <int MODE>
int dispatch(const SgemmArgs &a, bool av, bool bv, dim3 grid, cudaStream_t s) {
    ProfScope ps(KC_GEMM_CUDA_CORE, 1, s);
    if (av && bv) sgemm_kernel<MODE, true, true><<<grid, kThreads, 0, s>>>(a);
    else if (av) sgemm_kernel<MODE, true, false><<<grid, kThreads, 0, s>>>(a);
    else if (bv) sgemm_kernel<MODE, false, true><<<grid, kThreads, 0, s>>>(a);
    else sgemm_kernel<MODE, false, false><<<grid, kThreads, 0, s>>>(a);
    OCPPO_LAUNCH_CHECK("sgemm_kernel");
    return OCPPO_OK;
}

}  // namespace

int g_gemm_mode = 8;   // tcgen05: three FP16 passes on fp16-pair stored operands, one accumulator; see ocppo_set_gemm_mode

static int tc_ops_mask() {
#ifdef OCPPO_BRINGUP
    static const int ops_mask = getenv("OCPPO_TC_OPS") ? atoi(getenv("OCPPO_TC_OPS")) : 7;   // bring-up: bit per op
    return ops_mask;
#else
    return 7;        // release build: no environment knobs
#endif
}

void sgemm_engine_flags(int engine, SgemmArgs *b) {
    b->single_acc = engine == 2 || engine == 4 || engine == 8;    // 8: fp16-pair storage, one double-buffered accumulator
    b->corr_bf16 = engine >= 3;
    b->f16x3 = engine >= 5;       // 6, 7: same kernels, operands / outputs stored as fp16 pairs
}

bool sgemm_uses_tc(SgemmMode mode, const SgemmArgs &a) {
    if (g_gemm_mode < 1 || !((tc_ops_mask() >> (int)mode) & 1)) return false;
    SgemmArgs b = a;
    sgemm_engine_flags(g_gemm_mode, &b);
    return tc_gemm_supported(mode, b);
}

// FP16x3 engine: operands whose max-abs the caller did not supply are measured here (one extra read of the operand)
int sgemm_fill_amax(SgemmMode mode, SgemmArgs *b, cudaStream_t stream) {
    if (!b->f16x3 || (b->amax_a != nullptr && b->amax_b != nullptr)) return OCPPO_OK;
    OCPPO_CHECK_ARG(b->amax_scratch != nullptr, "sgemm: FP16x3 engine without operand maxima needs amax_scratch");
    if (b->amax_a == nullptr) {
        const bool kfirst = mode == SGEMM_WGRAD;                        // A is [K, M] for WGRAD, [M, K] otherwise
        OCPPO_TRY(launch_absmax(b->A, kfirst ? b->K : b->M, kfirst ? b->M : b->K, b->lda, b->amax_scratch, 0, stream));
        b->amax_a = b->amax_scratch;
    }
    if (b->amax_b == nullptr) {
        const bool kfirst = mode != SGEMM_FWD;                          // B is [N, K] for FWD, [K, N] otherwise
        OCPPO_TRY(launch_absmax(b->B, kfirst ? b->K : b->N, kfirst ? b->N : b->K, b->ldb, b->amax_scratch + 1, 0, stream));
        b->amax_b = b->amax_scratch + 1;
    }
    return OCPPO_OK;
}

int sgemm_launch(SgemmMode mode, const SgemmArgs &a, cudaStream_t stream) {
    if (sgemm_uses_tc(mode, a)) {
        SgemmArgs b = a;
        sgemm_engine_flags(g_gemm_mode, &b);
        OCPPO_TRY(sgemm_fill_amax(mode, &b, stream));
        return tc_gemm_launch(mode, b, stream);
    }
    OCPPO_CHECK_ARG(mode != SGEMM_DGRAD || a.mask != nullptr, "sgemm: the FMA dgrad kernel needs the fp32 mask");
    return fma_gemm_launch(mode, a, stream);
}

static bool thin_fwd_enabled() {
#ifdef OCPPO_BRINGUP
    static const bool on = getenv("OCPPO_THIN_FWD") ? atoi(getenv("OCPPO_THIN_FWD")) != 0 : true;   // bring-up A/B
    return on;
#else
    return true;
#endif
}

int fma_gemm_launch(SgemmMode mode, const SgemmArgs &a, cudaStream_t stream) {
    if (a.M <= 0 || a.N <= 0) return OCPPO_OK;
    if (mode == SGEMM_FWD && a.c16_hi != nullptr && a.K <= 8 && (a.N == 128 || a.N == 256 || a.N == 512 || a.N == 1024) &&
        (a.ldc & 3) == 0 && aligned16(a.bias) && (a.relu_bits == nullptr || a.fma_bits) && thin_fwd_enabled()) {
        int dev = 0, sms = 148;
        OCPPO_CUDA(cudaGetDevice(&dev));
        OCPPO_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        const int rpp2 = 2 * (256 / (a.N >> 2));                  // rows one block pass covers
        int64_t per = round_up(ceil_div((int64_t)a.M, (int64_t)sms * 8), rpp2);
        if (per < rpp2) per = rpp2;
        const unsigned grid = (unsigned)ceil_div((int64_t)a.M, per);
        ProfScope ps(KC_GEMM_CUDA_CORE, 1, stream);
        const int kd = a.K <= 2 ? 2 : a.K <= 4 ? 4 : a.K <= 6 ? 6 : 8;
        const int fast = a.K == kd && a.lda % 2 == 0 && (((uintptr_t)a.A) & 7) == 0;
        if (kd == 2) thin_fwd_pair_kernel<2><<<grid, 256, 0, stream>>>(a, per, fast);
        else if (kd == 4) thin_fwd_pair_kernel<4><<<grid, 256, 0, stream>>>(a, per, fast);
        else if (kd == 6) thin_fwd_pair_kernel<6><<<grid, 256, 0, stream>>>(a, per, fast);
        else thin_fwd_pair_kernel<8><<<grid, 256, 0, stream>>>(a, per, fast);
        OCPPO_LAUNCH_CHECK("thin_fwd_pair_kernel");
        return OCPPO_OK;
    }
    const int splits = mode == SGEMM_WGRAD ? a.splits : 1;
    OCPPO_CHECK_ARG(splits >= 1 && splits <= kSplitMax, "sgemm: bad split count %d", splits);
    dim3 grid((unsigned)ceil_div(a.N, BN), (unsigned)ceil_div(a.M, BM), (unsigned)splits);
    OCPPO_CHECK_ARG(grid.y <= 65535, "sgemm: M too large for one launch (%d)", a.M);
    // vector-load eligibility per operand
    bool av, bv;
    const bool kvec = (a.K % 4 == 0) && (a.k_per_split % 4 == 0);
    if (mode == SGEMM_WGRAD) {
        av = (a.M % 4 == 0) && (a.lda % 4 == 0) && aligned16(a.A);
        bv = (a.N % 4 == 0) && (a.ldb % 4 == 0) && aligned16(a.B);
    } else {
        av = kvec && (a.lda % 4 == 0) && aligned16(a.A);
        if (mode == SGEMM_FWD) bv = kvec && (a.ldb % 4 == 0) && aligned16(a.B);
        else bv = (a.N % 4 == 0) && (a.ldb % 4 == 0) && aligned16(a.B);
    }
    switch (mode) {
        case SGEMM_FWD: return dispatch<SGEMM_FWD>(a, av, bv, grid, stream);
        case SGEMM_DGRAD: return dispatch<SGEMM_DGRAD>(a, av, bv, grid, stream);
        default: return dispatch<SGEMM_WGRAD>(a, av, bv, grid, stream);
    }
}

}  // namespace ocppo
