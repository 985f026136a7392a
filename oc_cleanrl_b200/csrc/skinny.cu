// Weight gradients with one skinny side: dW = Wide[rows, nw]^T . Skinny[rows, ns], ns <= 32.
//   first trunk layer : Wide = dY0 [4B, 256],  Skinny = obs [4B, D]      (D = 6 for Pong)   -> dW0 [256, D]
//   actor/critic heads: Wide = hidden [B, H],  Skinny = d(logits,value) [B, A+1]            -> [A+1, H] (transposed)
// A 128x128-tiled GEMM wastes > 90 % of its FMAs here (ncu: 3 ms per launch, 14 % of the step); the
// contraction is HBM-bound instead: it reads Wide once.  One thread per Wide column keeps ns accumulators
// in registers, Skinny rows are staged in shared memory and broadcast, rows are split over the grid into
// fixed slabs and a second kernel sums the slab partials in fixed order (deterministic).
#include "common.cuh"
#include <stdlib.h>

namespace ocppo {

namespace {

constexpr int kSkThreads = 256;
constexpr int kSkRows = 128;           // skinny rows staged per iteration
constexpr int kSk4DepthDefault = 4;
constexpr int kSkBatch = 16;           // Wide rows a thread has in flight at once (8 left the kernel latency-bound at ~3 TB/s)

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
            for (int rb = 0; rb < kSkRows; rb += kSkBatch) {
                if (rb >= n) break;                                  // (zero rows would add nothing)
                float w[kSkBatch];
#pragma unroll
                for (int u = 0; u < kSkBatch; ++u) w[u] = (rb + u < n) ? __ldcs(wp + (int64_t)(rb + u) * ldw) : 0.f;
#pragma unroll
                for (int u = 0; u < kSkBatch; ++u) {
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

// Same contraction, FOUR Wide columns per thread (16-byte loads) and the block's threads split into row groups that take
// every RG-th row of the slab; the groups' sums fold through shared memory in fixed order at the end.  The one-column kernel
// above issues 29 instructions per 4-byte element (ncu: issue-bound at 3 TB/s, 64 % issue-active); here a row costs one
// LDG.128, two broadcast LDS.128 and 36 FP instructions for four elements.
// kSk4Depth: rows (16-byte loads) a thread has in flight; 4 with three resident blocks per SM, 8 with two
template <int NS, int kSk4Depth>
__global__ void __launch_bounds__(kSkThreads, kSk4Depth == 4 ? 3 : 2)
skinny_wgrad4_kernel(const float *__restrict__ wide, int64_t ldw, int nw, const float *__restrict__ skinny,
                     int64_t lds, int ns, int64_t rows, int64_t rows_per_cta, int tpr_log2, float *__restrict__ part,
                     float *__restrict__ wsum_part, float *__restrict__ ssum_part) {
    static_assert(NS == 8, "four columns x NS accumulators per thread");
    __shared__ __align__(16) float sk[kSkRows][NS];
    __shared__ __align__(16) float fold[kSkThreads][4 * NS + 4];
    const int tid = threadIdx.x;
    const int tpr = 1 << tpr_log2, rg = tid >> tpr_log2, ct = tid & (tpr - 1), RG = kSkThreads >> tpr_log2;
    const int j4 = (blockIdx.y * tpr + ct) * 4;
    const bool active = j4 < nw;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_cta;
    const int64_t r1 = r0 + rows_per_cta < rows ? r0 + rows_per_cta : rows;
    // accumulators as fp32 pairs: FFMA2 does two IEEE round-to-nearest FMAs per issue slot (same sums as scalar FMAs)
    float2 acc[4][NS / 2];
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int i = 0; i < NS / 2; ++i) acc[c][i] = make_float2(0.f, 0.f);
    float wsum[4] = {0.f, 0.f, 0.f, 0.f};
    float ssum = 0.f;
    auto row = [&](const float4 &w, int r) {
        const float4 s0 = *reinterpret_cast<const float4 *>(&sk[r][0]);
        const float4 s1 = *reinterpret_cast<const float4 *>(&sk[r][4]);
        const float2 p0 = make_float2(s0.x, s0.y), p1 = make_float2(s0.z, s0.w), p2 = make_float2(s1.x, s1.y), p3 = make_float2(s1.z, s1.w);
        const float wv[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            wsum[c] += wv[c];
            const float2 ww = make_float2(wv[c], wv[c]);
            acc[c][0] = __ffma2_rn(ww, p0, acc[c][0]); acc[c][1] = __ffma2_rn(ww, p1, acc[c][1]);
            acc[c][2] = __ffma2_rn(ww, p2, acc[c][2]); acc[c][3] = __ffma2_rn(ww, p3, acc[c][3]);
        }
    };
    for (int64_t rc = r0; rc < r1; rc += kSkRows) {
        const int n = (int)(r1 - rc < kSkRows ? r1 - rc : kSkRows);
        __syncthreads();
        for (int e = tid; e < kSkRows * NS; e += kSkThreads) {
            const int r = e / NS, i = e - r * NS;
            sk[r][i] = (r < n && i < ns) ? skinny[(rc + r) * lds + i] : 0.f;
        }
        __syncthreads();
        if (blockIdx.y == 0 && tid < NS && ssum_part != nullptr) {
            for (int r = 0; r < n; ++r) ssum += sk[r][tid];
        }
        if (active) {
            const float *wp = wide + (rc + rg) * ldw + j4;
            const int64_t step = (int64_t)RG * ldw;
            int r = rg;
            for (; r + (kSk4Depth - 1) * RG < n; r += kSk4Depth * RG) {
                float4 w[kSk4Depth];
#pragma unroll
                for (int u = 0; u < kSk4Depth; ++u) w[u] = __ldcs(reinterpret_cast<const float4 *>(wp + u * step));
                wp += kSk4Depth * step;
#pragma unroll
                for (int u = 0; u < kSk4Depth; ++u) row(w[u], r + u * RG);
            }
            for (; r < n; r += RG) {
                const float4 w = __ldcs(reinterpret_cast<const float4 *>(wp));
                wp += step;
                row(w, r);
            }
        }
    }
    // fold the row groups in order 0, 1, ..., RG - 1 (group 0's threads own the outputs)
    __syncthreads();
    if (rg > 0) {
        float *f = fold[tid];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            *reinterpret_cast<float4 *>(f + c * NS) = make_float4(acc[c][0].x, acc[c][0].y, acc[c][1].x, acc[c][1].y);
            *reinterpret_cast<float4 *>(f + c * NS + 4) = make_float4(acc[c][2].x, acc[c][2].y, acc[c][3].x, acc[c][3].y);
        }
        *reinterpret_cast<float4 *>(f + 4 * NS) = make_float4(wsum[0], wsum[1], wsum[2], wsum[3]);
    }
    __syncthreads();
    if (rg == 0 && active) {
        for (int g = 1; g < RG; ++g) {
            const float *f = fold[g * tpr + ct];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const float4 a0 = *reinterpret_cast<const float4 *>(f + c * NS), a1 = *reinterpret_cast<const float4 *>(f + c * NS + 4);
                acc[c][0].x += a0.x; acc[c][0].y += a0.y; acc[c][1].x += a0.z; acc[c][1].y += a0.w;
                acc[c][2].x += a1.x; acc[c][2].y += a1.y; acc[c][3].x += a1.z; acc[c][3].y += a1.w;
            }
            const float4 ws = *reinterpret_cast<const float4 *>(f + 4 * NS);
            wsum[0] += ws.x; wsum[1] += ws.y; wsum[2] += ws.z; wsum[3] += ws.w;
        }
        float *dst = part + ((int64_t)blockIdx.x * nw + j4) * NS;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            *reinterpret_cast<float4 *>(dst + c * NS) = make_float4(acc[c][0].x, acc[c][0].y, acc[c][1].x, acc[c][1].y);
            *reinterpret_cast<float4 *>(dst + c * NS + 4) = make_float4(acc[c][2].x, acc[c][2].y, acc[c][3].x, acc[c][3].y);
        }
        if (wsum_part != nullptr)
            *reinterpret_cast<float4 *>(wsum_part + (int64_t)blockIdx.x * nw + j4) = make_float4(wsum[0], wsum[1], wsum[2], wsum[3]);
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

// slabs: 4 per SM for the one-column kernel, 3 (its resident blocks) for the four-column one -- never more than the
// 4 per SM the partial-sum workspace is sized for
int skinny_slabs(int64_t rows, int sm_count, int64_t *rows_per_cta, int per_sm) {
    int64_t G = (int64_t)sm_count * per_sm;
    // (the four-column kernel handles a partial last chunk at no cost: a finer slab size keeps all its blocks equally loaded)
    int64_t per = round_up(ceil_div(rows, G), per_sm == 4 ? kSkRows : 32);
    if (per < kSkRows) per = kSkRows;
    *rows_per_cta = per;
    return (int)ceil_div(rows, per);
}

static bool skinny_vec4_enabled() {
#ifdef OCPPO_BRINGUP
    static const bool on = getenv("OCPPO_SKINNY_VEC4") ? atoi(getenv("OCPPO_SKINNY_VEC4")) != 0 : true;   // bring-up A/B
    return on;
#else
    return true;
#endif
}

static int skinny_vec4_depth() {
#ifdef OCPPO_BRINGUP
    static const int d = getenv("OCPPO_SK4_DEPTH") ? atoi(getenv("OCPPO_SK4_DEPTH")) : kSk4DepthDefault;   // bring-up A/B
    return d;
#else
    return kSk4DepthDefault;
#endif
}

// part: [G][nw][NS] floats, wsum_part: [G][nw] or null, ssum_part: [G][NS] or null.  NS = 8 if ns <= 8 else 32.
int launch_skinny_wgrad(const float *wide, int64_t ldw, int nw, const float *skinny, int64_t lds, int ns, int64_t rows,
                        int sm_count, float *part, float *wsum_part, float *ssum_part, int *G_out, int *NS_out,
                        cudaStream_t stream) {
    OCPPO_CHECK_ARG(ns >= 1 && ns <= 32 && nw >= 1 && rows >= 1, "skinny wgrad: bad shape nw=%d ns=%d", nw, ns);
    int64_t per = 0;
    const int NS = ns <= 8 ? 8 : 32;
    const bool vec4 = NS == 8 && nw % 4 == 0 && ldw % 4 == 0 && (((uintptr_t)wide) & 15) == 0 && skinny_vec4_enabled();
    const int G = skinny_slabs(rows, sm_count, &per, vec4 ? (skinny_vec4_depth() == 8 ? 2 : 3) : 4);
    ProfScope ps(KC_GEMM_CUDA_CORE, 1, stream);
    if (vec4) {
        int tpr_log2 = 5;                                  // threads per row: the power of two covering nw / 4, 32 .. 256
        while ((1 << tpr_log2) < nw / 4 && tpr_log2 < 8) ++tpr_log2;
        dim3 grid4((unsigned)G, (unsigned)ceil_div(nw / 4, 1 << tpr_log2));
        if (skinny_vec4_depth() == 8)
            skinny_wgrad4_kernel<8, 8><<<grid4, kSkThreads, 0, stream>>>(wide, ldw, nw, skinny, lds, ns, rows, per, tpr_log2, part,
                                                                        wsum_part, ssum_part);
        else
            skinny_wgrad4_kernel<8, 4><<<grid4, kSkThreads, 0, stream>>>(wide, ldw, nw, skinny, lds, ns, rows, per, tpr_log2, part,
                                                                        wsum_part, ssum_part);
        OCPPO_LAUNCH_CHECK("skinny_wgrad4_kernel");
        *G_out = G;
        *NS_out = NS;
        return OCPPO_OK;
    }
    dim3 grid((unsigned)G, (unsigned)ceil_div(nw, kSkThreads));
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
