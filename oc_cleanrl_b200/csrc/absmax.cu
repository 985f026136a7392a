// max |x| of a tensor as a device scalar: the power-of-two operand scale of the FP16x3 tensor-core engine
// (tc_gemm.cu) when the kernel that produced the tensor did not track it.  HBM-bound: one read of the tensor.
#include "common.cuh"

namespace ocppo {

namespace {

template <bool VEC>
__global__ void __launch_bounds__(256) absmax_kernel(const float *__restrict__ x, int64_t rows, int64_t cols, int64_t ld,
                                                     float *__restrict__ out) {
    float m = 0.f;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x, t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (VEC) {
        const int64_t cq = cols >> 2, n = rows * cq;
        const bool flat = (ld == cols);
        for (int64_t i = t0; i < n; i += stride) {
            const int64_t r = flat ? 0 : i / cq, c = flat ? i : i - r * cq;
            const float4 v = __ldcs(reinterpret_cast<const float4 *>(x + r * ld) + c);
            m = fmaxf(fmaxf(m, fmaxf(fabsf(v.x), fabsf(v.y))), fmaxf(fabsf(v.z), fabsf(v.w)));
        }
    } else {
        const int64_t n = rows * cols;
        for (int64_t i = t0; i < n; i += stride) {
            const int64_t r = i / cols, c = i - r * cols;
            m = fmaxf(m, fabsf(x[r * ld + c]));
        }
    }
    const unsigned wm = __reduce_max_sync(0xffffffffu, __float_as_uint(m));      // non-negative floats order as uints
    __shared__ unsigned part[8];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = wm;
    __syncthreads();
    if (threadIdx.x < 32) {
        const unsigned v = threadIdx.x < 8 ? part[threadIdx.x] : 0u;
        const unsigned bm = __reduce_max_sync(0xffffffffu, v);
        if (threadIdx.x == 0 && bm != 0u) atomicMax(reinterpret_cast<unsigned *>(out), bm);
    }
}

__global__ void __launch_bounds__(256) absmax_multi_kernel(AbsmaxSegs segs) {
    const float *x = segs.x[blockIdx.y];
    const int64_t n = segs.n[blockIdx.y];
    float m = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        m = fmaxf(m, fabsf(__ldg(x + i)));
    const unsigned wm = __reduce_max_sync(0xffffffffu, __float_as_uint(m));
    __shared__ unsigned part[8];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = wm;
    __syncthreads();
    if (threadIdx.x < 32) {
        const unsigned v = threadIdx.x < 8 ? part[threadIdx.x] : 0u;
        const unsigned bm = __reduce_max_sync(0xffffffffu, v);
        if (threadIdx.x == 0 && bm != 0u) atomicMax(reinterpret_cast<unsigned *>(segs.out[blockIdx.y]), bm);
    }
}

// two elements per thread; segments start at even offsets or fall back to the scalar tail
__global__ void __launch_bounds__(256) split_f16_multi_kernel(AbsmaxSegs segs, const float *base, uint16_t *hi, uint16_t *lo) {
    const float *x = segs.x[blockIdx.y];
    const int64_t n = segs.n[blockIdx.y], off = x - base;
    float scale, inv;
    f16_scale_of(__ldg(segs.out[blockIdx.y]), scale, inv);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        uint32_t h, l;
        split_f16x2(__ldg(x + i), 0.f, scale, h, l);
        hi[off + i] = (uint16_t)(h & 0xFFFFu);
        lo[off + i] = (uint16_t)(l & 0xFFFFu);
    }
}

}  // namespace

int launch_split_f16_multi(const AbsmaxSegs &segs, const float *base, void *hi, void *lo, cudaStream_t stream) {
    if (segs.count <= 0) return OCPPO_OK;
    ProfScope ps(KC_AUX, 1, stream);
    split_f16_multi_kernel<<<dim3(296, (unsigned)segs.count), 256, 0, stream>>>(segs, base, reinterpret_cast<uint16_t *>(hi),
                                                                              reinterpret_cast<uint16_t *>(lo));
    OCPPO_LAUNCH_CHECK("split_f16_multi_kernel");
    return OCPPO_OK;
}

int launch_absmax_multi(const AbsmaxSegs &segs, cudaStream_t stream) {
    if (segs.count <= 0) return OCPPO_OK;
    ProfScope ps(KC_AUX, 1, stream);
    absmax_multi_kernel<<<dim3(148, (unsigned)segs.count), 256, 0, stream>>>(segs);     // (32 blocks per tensor were latency-bound: 21 us for 9 MB)
    OCPPO_LAUNCH_CHECK("absmax_multi_kernel");
    return OCPPO_OK;
}

int launch_absmax(const float *x, int64_t rows, int64_t cols, int64_t ld, float *out, int sm_count, cudaStream_t stream) {
    OCPPO_CHECK_ARG(x != nullptr && out != nullptr && rows >= 0 && cols >= 0 && ld >= cols, "absmax: bad arguments");
    OCPPO_CUDA(cudaMemsetAsync(out, 0, sizeof(float), stream));
    if (rows == 0 || cols == 0) return OCPPO_OK;
    const bool vec = (cols % 4 == 0) && (ld % 4 == 0) && ((((uintptr_t)x) & 15) == 0);
    const int64_t items = vec ? rows * (cols / 4) : rows * cols;
    int64_t grid = ceil_div(items, 256 * 4);
    const int64_t cap = (int64_t)(sm_count > 0 ? sm_count : 148) * 8;
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    ProfScope ps(KC_AUX, 1, stream);
    if (vec) absmax_kernel<true><<<(unsigned)grid, 256, 0, stream>>>(x, rows, cols, ld, out);
    else absmax_kernel<false><<<(unsigned)grid, 256, 0, stream>>>(x, rows, cols, ld, out);
    OCPPO_LAUNCH_CHECK("absmax_kernel");
    return OCPPO_OK;
}

}  // namespace ocppo
