// fp16-pair storage (GEMM engine 6): a tensor x lives in global memory as hi = fp16(x s), lo = fp16(x s - hi) with s the
// power of two that a device scalar (the tensor's max-abs, or a rigorous bound on it) implies -- the layout the tensor-core
// kernel (tc_gemm.cu, PRE variant) loads with TMA straight into its MMA tiles.  The GEMM epilogues write pairs themselves;
// these HBM-bound passes cover the tensors that other kernels produce in fp32 (first-layer activations, the loss head's
// gradient), the bias gradients (column sums of a pair) and the weight norms the epilogues' scale bounds come from.
#include "common.cuh"
#include <cuda_fp16.h>

namespace ocppo {

namespace {

// x[n] fp32 -> pair; n % 4 == 0, 16-byte aligned
__global__ void __launch_bounds__(256) split16_kernel(const float *__restrict__ x, int64_t n4, const float *__restrict__ scale_src,
                                                      uint2 *__restrict__ hi, uint2 *__restrict__ lo) {
    float scale, inv;
    f16_scale_of(__ldg(scale_src), scale, inv);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 v = __ldcs(reinterpret_cast<const float4 *>(x) + i);
        uint32_t h0, l0, h1, l1;
        split_f16x2(v.x, v.y, scale, h0, l0);
        split_f16x2(v.z, v.w, scale, h1, l1);
        hi[i] = make_uint2(h0, h1);
        lo[i] = make_uint2(l0, l1);
    }
}

__device__ __forceinline__ float4 merge4(const uint2 h, const uint2 l, float inv) {
    const float2 h0 = __half22float2(*reinterpret_cast<const __half2 *>(&h.x)), h1 = __half22float2(*reinterpret_cast<const __half2 *>(&h.y));
    const float2 l0 = __half22float2(*reinterpret_cast<const __half2 *>(&l.x)), l1 = __half22float2(*reinterpret_cast<const __half2 *>(&l.y));
    return make_float4((h0.x + l0.x) * inv, (h0.y + l0.y) * inv, (h1.x + l1.x) * inv, (h1.y + l1.y) * inv);
}

// pair -> fp32 (test hooks, fp32 consumers of a tensor that only exists as a pair)
__global__ void __launch_bounds__(256) merge16_kernel(const uint2 *__restrict__ hi, const uint2 *__restrict__ lo, int64_t n4,
                                                      const float *__restrict__ scale_src, float4 *__restrict__ out) {
    float scale, inv;
    f16_scale_of(__ldg(scale_src), scale, inv);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = merge4(__ldg(hi + i), __ldg(lo + i), inv);
}

// column sums of a pair [rows, cols] (cols % 4 == 0): block g sums its fixed slab of rows into part[g][cols] -- a fixed
// order, so the bias gradients stay bitwise repeatable.  Thread t owns the 4 columns 4 (t % (cols / 4)) and every
// (256 / (cols / 4))-th row of the slab when cols < 1024.
__global__ void __launch_bounds__(256) colsum16_kernel(const uint2 *__restrict__ hi, const uint2 *__restrict__ lo, int64_t rows, int cols,
                                                       int64_t rows_per_block, const float *__restrict__ scale_src,
                                                       float *__restrict__ part) {
    float scale, inv;
    f16_scale_of(__ldg(scale_src), scale, inv);
    const int cq = cols >> 2;                                   // float4 groups per row
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
    const int64_t r1 = r0 + rows_per_block < rows ? r0 + rows_per_block : rows;
    __shared__ float4 red[256];
    for (int c0 = 0; c0 < cq; c0 += 256) {
        const int lanes = cq - c0 < 256 ? cq - c0 : 256;        // column groups handled in this pass
        const int rstep = 256 / lanes > 0 ? 256 / lanes : 1;    // rows processed side by side
        const int cg = threadIdx.x % lanes, rl = threadIdx.x / lanes;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        if (rl < rstep)
            for (int64_t r = r0 + rl; r < r1; r += rstep) {
                const int64_t i = r * cq + c0 + cg;
                const float4 v = merge4(__ldcs(hi + i), __ldcs(lo + i), 1.f);
                acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
            }
        red[threadIdx.x] = acc;
        __syncthreads();
        if (threadIdx.x < lanes) {
            float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int k = 0; k < rstep; ++k) {
                const float4 v = red[k * lanes + threadIdx.x];
                s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
            }
            reinterpret_cast<float4 *>(part + (int64_t)blockIdx.x * cols)[c0 + threadIdx.x] =
                make_float4(s.x * inv, s.y * inv, s.z * inv, s.w * inv);
        }
        __syncthreads();
    }
}

// fp32 [rows, cols] (row stride ld) -> pair [rows, kPadCols] with zero padding: the first trunk layer's input (and its weight)
// when the object vector is wider than the skinny route likes -- padded to the tensor-core tile so that layer, too, runs
// as a conversion-free GEMM.  One thread per 8 output columns.
__global__ void __launch_bounds__(256) pad_split16_kernel(const float *__restrict__ x, int64_t rows, int cols, int64_t ld,
                                                          const float *__restrict__ scale_src, uint4 *__restrict__ hi,
                                                          uint4 *__restrict__ lo) {
    float scale, inv;
    f16_scale_of(__ldg(scale_src), scale, inv);
    constexpr int G = kPadCols / 8;
    const int64_t total = rows * G;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / G;
        const int c0 = (int)(i - r * G) * 8;
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = (c0 + j < cols) ? __ldg(x + r * ld + c0 + j) : 0.f;
        uint32_t h[4], l[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) split_f16x2(v[2 * j], v[2 * j + 1], scale, h[j], l[j]);
        hi[i] = make_uint4(h[0], h[1], h[2], h[3]);
        lo[i] = make_uint4(l[0], l[1], l[2], l[3]);
    }
}

// out[r, c] (+)= sum_s part[s][r][c] for c < cols: split-K partials computed on kPadCols-wide rows, result rows `cols` wide
__global__ void __launch_bounds__(256) splitk_reduce_pad_kernel(const float *__restrict__ part, int splits, int rows, int cols,
                                                                float *__restrict__ out, int accumulate) {
    const int64_t total = (int64_t)rows * cols;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / cols;
        const int c = (int)(i - r * cols);
        float s = 0.f;
        for (int k = 0; k < splits; ++k) s += part[((int64_t)k * rows + r) * kPadCols + c];
        out[i] = accumulate ? out[i] + s : s;
    }
}

// the loss head's fp32 gradient [rows, cols]: pair AND column-sum partials in one read (same slab / order rules as above)
__global__ void __launch_bounds__(256) split16_colsum_kernel(const float4 *__restrict__ x, int64_t rows, int cols, int64_t rows_per_block,
                                                             const float *__restrict__ scale_src, uint2 *__restrict__ hi,
                                                             uint2 *__restrict__ lo, float *__restrict__ part) {
    float scale, inv;
    f16_scale_of(__ldg(scale_src), scale, inv);
    const int cq = cols >> 2;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
    const int64_t r1 = r0 + rows_per_block < rows ? r0 + rows_per_block : rows;
    __shared__ float4 red[256];
    for (int c0 = 0; c0 < cq; c0 += 256) {
        const int lanes = cq - c0 < 256 ? cq - c0 : 256;
        const int rstep = 256 / lanes > 0 ? 256 / lanes : 1;
        const int cg = threadIdx.x % lanes, rl = threadIdx.x / lanes;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        if (rl < rstep)
            for (int64_t r = r0 + rl; r < r1; r += rstep) {
                const int64_t i = r * cq + c0 + cg;
                const float4 v = __ldcs(x + i);
                uint32_t h0, l0, h1, l1;
                split_f16x2(v.x, v.y, scale, h0, l0);
                split_f16x2(v.z, v.w, scale, h1, l1);
                hi[i] = make_uint2(h0, h1);
                lo[i] = make_uint2(l0, l1);
                acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
            }
        red[threadIdx.x] = acc;
        __syncthreads();
        if (threadIdx.x < lanes) {
            float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int k = 0; k < rstep; ++k) {
                const float4 v = red[k * lanes + threadIdx.x];
                s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
            }
            reinterpret_cast<float4 *>(part + (int64_t)blockIdx.x * cols)[c0 + threadIdx.x] = s;
        }
        __syncthreads();
    }
}

// per weight matrix W[out, in] (+ bias[out]): largest row L1 norm, largest column L1 norm, largest |bias| as float bits
// (atomicMax on pre-zeroed slots).  grid (slices, matrices)
__global__ void __launch_bounds__(256) wnorm_kernel(WnormSegs segs) {
    const int m = blockIdx.y;
    const float *W = segs.w[m];
    const int out = segs.out_dim[m], in = segs.in_dim[m];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float rmax = 0.f, amax = 0.f;
    for (int j = blockIdx.x * 8 + warp; j < out; j += gridDim.x * 8) {          // warp per row
        float s = 0.f;
        for (int k = lane; k < in; k += 32) {
            const float a = fabsf(__ldg(W + (int64_t)j * in + k));
            s += a;
            amax = fmaxf(amax, a);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        rmax = fmaxf(rmax, s);
    }
    // columns: block = 32 columns at a time, warp w adds rows w, w + 8, ... (eight loads in flight: two at a time left this
    // pass latency-bound), the 8 partial sums fold in order
    __shared__ float red[8][33];
    float cmax = 0.f;
    for (int k0 = blockIdx.x * 32; k0 < in; k0 += gridDim.x * 32) {
        const int k = k0 + lane;
        float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (k < in) {
            int j = warp;
            for (; j + 56 < out; j += 64) {
                float v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) v[u] = fabsf(__ldg(W + (int64_t)(j + 8 * u) * in + k));
#pragma unroll
                for (int u = 0; u < 8; ++u) s[u] += v[u];
            }
            for (; j < out; j += 8) s[0] += fabsf(__ldg(W + (int64_t)j * in + k));
        }
        red[warp][lane] = ((s[0] + s[1]) + (s[2] + s[3])) + ((s[4] + s[5]) + (s[6] + s[7]));
        __syncthreads();
        if (warp == 0) {
            float t = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) t += red[w][lane];
            cmax = fmaxf(cmax, t);
        }
        __syncthreads();
    }
    float bmax = 0.f;
    if (blockIdx.x == 0 && segs.b[m] != nullptr)
        for (int j = threadIdx.x; j < out; j += 256) bmax = fmaxf(bmax, fabsf(__ldg(segs.b[m] + j)));
    // sums of non-negative terms in fp32 can come out a few ulp low: pad the norms so the bound stays a bound
    rmax *= 1.0001f; cmax *= 1.0001f;
    const unsigned r = __reduce_max_sync(0xffffffffu, __float_as_uint(rmax));
    const unsigned c = __reduce_max_sync(0xffffffffu, __float_as_uint(cmax));
    const unsigned b = __reduce_max_sync(0xffffffffu, __float_as_uint(bmax));
    if (segs.w_amax[m] != nullptr) {
        const unsigned a = __reduce_max_sync(0xffffffffu, __float_as_uint(amax));
        if (lane == 0 && a) atomicMax(reinterpret_cast<unsigned *>(segs.w_amax[m]), a);
    }
    if (lane == 0) {
        if (r) atomicMax(reinterpret_cast<unsigned *>(segs.row_l1[m]), r);
        if (c) atomicMax(reinterpret_cast<unsigned *>(segs.col_l1[m]), c);
        if (b) atomicMax(reinterpret_cast<unsigned *>(segs.b_max[m]), b);
    }
}

// out[c] (+)= sum over p of part[p][c]: block = 32 columns, warp w adds rows w, w + 32, ... in order, then the 32 warps fold
// in order -- a fixed summation tree whatever the launch looks like
__global__ void __launch_bounds__(1024) colsum_reduce_kernel(const float *__restrict__ part, int parts, int cols, float *__restrict__ out,
                                                            int accumulate) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, c = blockIdx.x * 32 + lane;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    if (c < cols) {
        int r = warp;
        for (; r + 96 < parts; r += 128) {
            s0 += __ldg(part + (int64_t)r * cols + c);
            s1 += __ldg(part + (int64_t)(r + 32) * cols + c);
            s2 += __ldg(part + (int64_t)(r + 64) * cols + c);
            s3 += __ldg(part + (int64_t)(r + 96) * cols + c);
        }
        for (; r < parts; r += 32) s0 += __ldg(part + (int64_t)r * cols + c);
    }
    __shared__ float red[32][33];
    red[warp][lane] = (s0 + s1) + (s2 + s3);
    __syncthreads();
    if (warp == 0 && c < cols) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < 32; ++w) s += red[w][lane];
        out[c] = accumulate ? out[c] + s : s;
    }
}

// The two folds that follow a weight-gradient GEMM on pair storage, in ONE launch: blockIdx.y == 0 sums the split-K partial
// tiles into dW (splitk_reduce_kernel's loop), blockIdx.y == 1 folds the column-sum partials of the bias gradient
// (colsum_reduce_kernel's fixed tree).  Same sums in the same order as the two kernels, so the gradients are bit-identical;
// four launches per minibatch fewer.
__global__ void __launch_bounds__(1024) wgrad_fold_kernel(const float *__restrict__ part, int splits, int64_t stride, int64_t count,
                                                         float *__restrict__ out, int acc_w, const float *__restrict__ cpart,
                                                         int parts, int cols, float *__restrict__ cout, int acc_b) {
    if (blockIdx.y == 0) {
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
            float s = 0.f;
            for (int k = 0; k < splits; ++k) s += part[(int64_t)k * stride + i];
            out[i] = acc_w ? out[i] + s : s;
        }
        return;
    }
    if ((int)blockIdx.x * 32 >= cols) return;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, c = blockIdx.x * 32 + lane;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    if (c < cols) {
        int r = warp;
        for (; r + 96 < parts; r += 128) {
            s0 += __ldg(cpart + (int64_t)r * cols + c);
            s1 += __ldg(cpart + (int64_t)(r + 32) * cols + c);
            s2 += __ldg(cpart + (int64_t)(r + 64) * cols + c);
            s3 += __ldg(cpart + (int64_t)(r + 96) * cols + c);
        }
        for (; r < parts; r += 32) s0 += __ldg(cpart + (int64_t)r * cols + c);
    }
    __shared__ float red[32][33];
    red[warp][lane] = (s0 + s1) + (s2 + s3);
    __syncthreads();
    if (warp == 0 && c < cols) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < 32; ++w) s += red[w][lane];
        cout[c] = acc_b ? cout[c] + s : s;
    }
}

inline int grid_of(int64_t items, int sm_count) {
    int64_t g = ceil_div(items, 256 * 4);
    const int64_t cap = (int64_t)(sm_count > 0 ? sm_count : 148) * 8;
    if (g > cap) g = cap;
    return g < 1 ? 1 : (int)g;
}

}  // namespace

int launch_split16(const float *x, int64_t n, const float *scale_src, void *hi, void *lo, int sm_count, cudaStream_t stream) {
    if (n == 0) return OCPPO_OK;
    OCPPO_CHECK_ARG(x && scale_src && hi && lo && n % 4 == 0 && ((((uintptr_t)x) | ((uintptr_t)hi) | ((uintptr_t)lo)) & 15) == 0,
                    "split16: bad arguments");
    ProfScope ps(KC_AUX, 1, stream);
    split16_kernel<<<grid_of(n / 4, sm_count), 256, 0, stream>>>(x, n / 4, scale_src, reinterpret_cast<uint2 *>(hi), reinterpret_cast<uint2 *>(lo));
    OCPPO_LAUNCH_CHECK("split16_kernel");
    return OCPPO_OK;
}

int launch_merge16(const void *hi, const void *lo, int64_t n, const float *scale_src, float *out, int sm_count, cudaStream_t stream) {
    if (n == 0) return OCPPO_OK;
    OCPPO_CHECK_ARG(hi && lo && scale_src && out && n % 4 == 0, "merge16: bad arguments");
    ProfScope ps(KC_AUX, 1, stream);
    merge16_kernel<<<grid_of(n / 4, sm_count), 256, 0, stream>>>(reinterpret_cast<const uint2 *>(hi), reinterpret_cast<const uint2 *>(lo), n / 4,
                                                                 scale_src, reinterpret_cast<float4 *>(out));
    OCPPO_LAUNCH_CHECK("merge16_kernel");
    return OCPPO_OK;
}

int colsum16_blocks(int64_t rows, int sm_count) {
    int64_t g = (int64_t)(sm_count > 0 ? sm_count : 148) * 4;
    if (g > kColsumBlocksMax) g = kColsumBlocksMax;
    if (g > rows) g = rows;
    return g < 1 ? 1 : (int)g;
}

int launch_colsum16(const void *hi, const void *lo, int64_t rows, int cols, const float *scale_src, float *part, int *blocks,
                    int sm_count, cudaStream_t stream) {
    OCPPO_CHECK_ARG(hi && lo && scale_src && part && blocks && rows >= 1 && cols >= 4 && cols % 4 == 0, "colsum16: bad arguments");
    const int g = colsum16_blocks(rows, sm_count);
    const int64_t rpb = ceil_div(rows, g);
    *blocks = (int)ceil_div(rows, rpb);
    ProfScope ps(KC_AUX, 1, stream);
    colsum16_kernel<<<*blocks, 256, 0, stream>>>(reinterpret_cast<const uint2 *>(hi), reinterpret_cast<const uint2 *>(lo), rows, cols, rpb,
                                                 scale_src, part);
    OCPPO_LAUNCH_CHECK("colsum16_kernel");
    return OCPPO_OK;
}

int launch_split16_colsum(const float *x, int64_t rows, int cols, const float *scale_src, void *hi, void *lo, float *part, int *blocks,
                          int sm_count, cudaStream_t stream) {
    OCPPO_CHECK_ARG(x && hi && lo && scale_src && part && blocks && rows >= 1 && cols >= 4 && cols % 4 == 0 && (((uintptr_t)x) & 15) == 0,
                    "split16_colsum: bad arguments");
    const int g = colsum16_blocks(rows, sm_count);
    const int64_t rpb = ceil_div(rows, g);
    *blocks = (int)ceil_div(rows, rpb);
    ProfScope ps(KC_AUX, 1, stream);
    split16_colsum_kernel<<<*blocks, 256, 0, stream>>>(reinterpret_cast<const float4 *>(x), rows, cols, rpb, scale_src,
                                                       reinterpret_cast<uint2 *>(hi), reinterpret_cast<uint2 *>(lo), part);
    OCPPO_LAUNCH_CHECK("split16_colsum_kernel");
    return OCPPO_OK;
}

int launch_pad_split16(const float *x, int64_t rows, int cols, int64_t ld, const float *scale_src, void *hi, void *lo, int sm_count,
                       cudaStream_t stream) {
    if (rows == 0) return OCPPO_OK;
    OCPPO_CHECK_ARG(x && scale_src && hi && lo && cols >= 1 && cols <= kPadCols && ld >= cols, "pad_split16: bad arguments");
    ProfScope ps(KC_AUX, 1, stream);
    pad_split16_kernel<<<grid_of(rows * (kPadCols / 8) / 2, sm_count), 256, 0, stream>>>(x, rows, cols, ld, scale_src,
                                                                                         reinterpret_cast<uint4 *>(hi), reinterpret_cast<uint4 *>(lo));
    OCPPO_LAUNCH_CHECK("pad_split16_kernel");
    return OCPPO_OK;
}

int launch_splitk_reduce_pad(const float *part, int splits, int rows, int cols, float *out, int accumulate, int sm_count, cudaStream_t stream) {
    OCPPO_CHECK_ARG(part && out && splits >= 1 && rows >= 1 && cols >= 1 && cols <= kPadCols, "splitk_reduce_pad: bad arguments");
    ProfScope ps(KC_AUX, 1, stream);
    splitk_reduce_pad_kernel<<<grid_of((int64_t)rows * cols, sm_count), 256, 0, stream>>>(part, splits, rows, cols, out, accumulate);
    OCPPO_LAUNCH_CHECK("splitk_reduce_pad_kernel");
    return OCPPO_OK;
}

int launch_colsum_reduce(const float *part, int parts, int cols, float *out, int accumulate, cudaStream_t stream) {
    OCPPO_CHECK_ARG(part && out && parts >= 1 && cols >= 1, "colsum_reduce: bad arguments");
    ProfScope ps(KC_AUX, 1, stream);
    colsum_reduce_kernel<<<(unsigned)ceil_div(cols, 32), 1024, 0, stream>>>(part, parts, cols, out, accumulate);
    OCPPO_LAUNCH_CHECK("colsum_reduce_kernel");
    return OCPPO_OK;
}

int launch_wgrad_fold(const float *part, int splits, int64_t stride, int64_t count, float *out, int acc_w, const float *cpart,
                      int parts, int cols, float *cout, int acc_b, int sm_count, cudaStream_t stream) {
    OCPPO_CHECK_ARG(part && out && cpart && cout && splits >= 1 && count >= 1 && parts >= 1 && cols >= 1, "wgrad_fold: bad arguments");
    int64_t gx = ceil_div(count, 1024);
    const int64_t cap = (int64_t)(sm_count > 0 ? sm_count : 148) * 2;
    if (gx > cap) gx = cap;
    const int64_t cb = ceil_div(cols, 32);
    if (gx < cb) gx = cb;
    ProfScope ps(KC_AUX, 1, stream);
    wgrad_fold_kernel<<<dim3((unsigned)gx, 2), 1024, 0, stream>>>(part, splits, stride, count, out, acc_w, cpart, parts, cols, cout, acc_b);
    OCPPO_LAUNCH_CHECK("wgrad_fold_kernel");
    return OCPPO_OK;
}

int launch_wnorm(const WnormSegs &segs, cudaStream_t stream) {
    if (segs.count <= 0) return OCPPO_OK;
    ProfScope ps(KC_AUX, 1, stream);
    wnorm_kernel<<<dim3(128, (unsigned)segs.count), 256, 0, stream>>>(segs);     // (32 slices per matrix were latency-bound: 36 us for 9 MB)
    OCPPO_LAUNCH_CHECK("wnorm_kernel");
    return OCPPO_OK;
}

}  // namespace ocppo
