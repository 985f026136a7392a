// Weight gradients with one skinny side: dW = Wide[rows, nw]^T . Skinny[rows, ns], ns <= 32.
//   first trunk layer : Wide = dY0 [4B, 256],  Skinny = obs [4B, D]      (D = 6 for Pong)   -> dW0 [256, D]
//   actor/critic heads: Wide = hidden [B, H],  Skinny = d(logits,value) [B, A+1]            -> [A+1, H] (transposed)
// A 128x128-tiled GEMM wastes > 90 % of its FMAs here (ncu: 3 ms per launch, 14 % of the step); the
// contraction is HBM-bound instead: it reads Wide once.  One thread per Wide column keeps ns accumulators
// in registers, Skinny rows are staged in shared memory and broadcast, rows are split over the grid into
// fixed slabs and a second kernel sums the slab partials in fixed order (deterministic).
#include "common.cuh"

namespace ocppo {

namespace {

constexpr int kSkThreads = 256;
constexpr int kSkRows = 64;            // skinny rows staged per iteration

template <int NS>
__global__ void __launch_bounds__(kSkThreads)
skinny_wgrad_kernel(const float *__restrict__ wide, int64_t ldw, int nw, const float *__restrict__ skinny,
                    int64_t lds, int ns, int64_t rows, int64_t rows_per_cta, float *__restrict__ part,
                    float *__restrict__ wsum_part, float *__restrict__ ssum_part) {
    __shared__ __align__(16) float sk[kSkRows][NS];
    const int tid = threadIdx.x;
    const int j = blockIdx.y * kSkThreads + tid;
    const bool active = j < nw;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_cta;
    const int64_t r1 = r0 + rows_per_cta < rows ? r0 + rows_per_cta : rows;
    float acc[NS];
#pragma unroll
    for (int i = 0; i < NS; ++i) acc[i] = 0.f;
    float wsum = 0.f, ssum = 0.f;
    for (int64_t rc = r0; rc < r1; rc += kSkRows) {
        const int n = (int)(r1 - rc < kSkRows ? r1 - rc : kSkRows);
        __syncthreads();
        for (int e = tid; e < kSkRows * NS; e += kSkThreads) {
            const int r = e / NS, i = e - r * NS;
            sk[r][i] = (r < n && i < ns) ? skinny[(rc + r) * lds + i] : 0.f;
        }
        __syncthreads();
        if (blockIdx.y == 0 && tid < NS) {
            for (int r = 0; r < n; ++r) ssum += sk[r][tid];
        }
        if (active) {
            const float *wp = wide + rc * ldw + j;
            for (int rb = 0; rb < kSkRows; rb += 8) {
                float w[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) w[u] = (rb + u < n) ? __ldg(wp + (int64_t)(rb + u) * ldw) : 0.f;
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    wsum += w[u];
#pragma unroll
                    for (int i = 0; i < NS; i += 4) {
                        const float4 s = *reinterpret_cast<const float4 *>(&sk[rb + u][i]);
                        acc[i] = fmaf(w[u], s.x, acc[i]); acc[i + 1] = fmaf(w[u], s.y, acc[i + 1]);
                        acc[i + 2] = fmaf(w[u], s.z, acc[i + 2]); acc[i + 3] = fmaf(w[u], s.w, acc[i + 3]);
                    }
                }
            }
        }
    }
    if (active) {
        float *dst = part + ((int64_t)blockIdx.x * nw + j) * NS;
#pragma unroll
        for (int i = 0; i < NS; i += 4)
            *reinterpret_cast<float4 *>(dst + i) = make_float4(acc[i], acc[i + 1], acc[i + 2], acc[i + 3]);
        if (wsum_part != nullptr) wsum_part[(int64_t)blockIdx.x * nw + j] = wsum;
    }
    if (blockIdx.y == 0 && tid < NS && ssum_part != nullptr) ssum_part[(int64_t)blockIdx.x * NS + tid] = ssum;
}

// out[i * si + j * sj] (+)= sum_g part[(g * nw + j) * NS + i]   for i in [i0, i1), j in [0, nw).
// One warp per output: lane l sums slabs l, l+32, ... in order, then a fixed shuffle tree (deterministic).
__global__ void __launch_bounds__(256)
skinny_reduce_kernel(const float *__restrict__ part, int G, int nw, int NS, int i0, int i1,
                     float *__restrict__ out, int64_t si, int64_t sj, int accumulate) {
    const int ni = i1 - i0, lane = threadIdx.x & 31;
    const int64_t total = (int64_t)nw * ni;
    for (int64_t e = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); e < total; e += (int64_t)gridDim.x * 8) {
        const int j = (int)(e / ni), i = i0 + (int)(e - (int64_t)j * ni);
        float s = 0.f;
        for (int g = lane; g < G; g += 32) s += part[((int64_t)g * nw + j) * NS + i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) {
            float *o = out + (int64_t)(i - i0) * si + (int64_t)j * sj;
            *o = accumulate ? *o + s : s;
        }
    }
}

// up to four reductions of the kind above in ONE launch (blockIdx.y selects the job): the heads' four outputs (actor /
// critic weights and biases) used to cost four launches per minibatch chunk
struct SkinnyJobs { SkinnyReduceJob j[4]; int n; };
__global__ void __launch_bounds__(256) skinny_reduce_multi_kernel(SkinnyJobs jobs) {
    const SkinnyReduceJob &J = jobs.j[blockIdx.y];
    const int ni = J.i1 - J.i0, lane = threadIdx.x & 31;
    const int64_t total = (int64_t)J.nw * ni;
    for (int64_t e = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); e < total; e += (int64_t)gridDim.x * 8) {
        const int j = (int)(e / ni), i = J.i0 + (int)(e - (int64_t)j * ni);
        float s = 0.f;
        for (int g = lane; g < J.G; g += 32) s += J.part[((int64_t)g * J.nw + j) * J.NS + i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) {
            float *o = J.out + (int64_t)(i - J.i0) * J.si + (int64_t)j * J.sj;
            *o = J.accumulate ? *o + s : s;
        }
    }
}

}  // namespace

int launch_skinny_reduce_multi(const SkinnyReduceJob *jobs, int n, cudaStream_t stream) {
    OCPPO_CHECK_ARG(jobs != nullptr && n >= 1 && n <= 4, "skinny_reduce_multi: 1..4 jobs");
    SkinnyJobs J{};
    J.n = n;
    int64_t most = 0;
    for (int k = 0; k < n; ++k) {
        J.j[k] = jobs[k];
        const int64_t total = (int64_t)jobs[k].nw * (jobs[k].i1 - jobs[k].i0);
        if (total > most) most = total;
    }
    if (most <= 0) return OCPPO_OK;
    ProfScope ps(KC_AUX, 1, stream);
    skinny_reduce_multi_kernel<<<dim3((unsigned)ceil_div(most, 8), (unsigned)n), 256, 0, stream>>>(J);
    OCPPO_LAUNCH_CHECK("skinny_reduce_multi_kernel");
    return OCPPO_OK;
}

int skinny_slabs(int64_t rows, int sm_count, int64_t *rows_per_cta) {
    int64_t G = (int64_t)sm_count * 4;
    int64_t per = round_up(ceil_div(rows, G), kSkRows);
    if (per < kSkRows) per = kSkRows;
    *rows_per_cta = per;
    return (int)ceil_div(rows, per);
}

// part: [G][nw][NS] floats, wsum_part: [G][nw] or null, ssum_part: [G][NS] or null.  NS = 8 if ns <= 8 else 32.
int launch_skinny_wgrad(const float *wide, int64_t ldw, int nw, const float *skinny, int64_t lds, int ns, int64_t rows,
                        int sm_count, float *part, float *wsum_part, float *ssum_part, int *G_out, int *NS_out,
                        cudaStream_t stream) {
    OCPPO_CHECK_ARG(ns >= 1 && ns <= 32 && nw >= 1 && rows >= 1, "skinny wgrad: bad shape nw=%d ns=%d", nw, ns);
    int64_t per = 0;
    const int G = skinny_slabs(rows, sm_count, &per);
    const int NS = ns <= 8 ? 8 : 32;
    dim3 grid((unsigned)G, (unsigned)ceil_div(nw, kSkThreads));
    ProfScope ps(KC_GEMM_CUDA_CORE, 1, stream);
    if (NS == 8)
        skinny_wgrad_kernel<8><<<grid, kSkThreads, 0, stream>>>(wide, ldw, nw, skinny, lds, ns, rows, per, part, wsum_part, ssum_part);
    else
        skinny_wgrad_kernel<32><<<grid, kSkThreads, 0, stream>>>(wide, ldw, nw, skinny, lds, ns, rows, per, part, wsum_part, ssum_part);
    OCPPO_LAUNCH_CHECK("skinny_wgrad_kernel");
    *G_out = G;
    *NS_out = NS;
    return OCPPO_OK;
}

int launch_skinny_reduce(const float *part, int G, int nw, int NS, int i0, int i1, float *out, int64_t si, int64_t sj,
                         int accumulate, cudaStream_t stream) {
    const int64_t total = (int64_t)nw * (i1 - i0);
    if (total <= 0) return OCPPO_OK;
    ProfScope ps(KC_AUX, 1, stream);
    skinny_reduce_kernel<<<(unsigned)ceil_div(total, 8), 256, 0, stream>>>(part, G, nw, NS, i0, i1, out, si, sj, accumulate);
    OCPPO_LAUNCH_CHECK("skinny_reduce_kernel");
    return OCPPO_OK;
}

}  // namespace ocppo
