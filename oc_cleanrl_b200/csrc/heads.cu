// Actor/critic heads fused with the Categorical distribution:
//   rollout:  hidden -> logits,value -> normalise -> sample (Philox) -> logp, entropy
//             (architectures/ppo.py:89-95 under no_grad, ppo_atari_oc.py:505-509)
//   update :  hidden -> logits,value -> PPO losses -> d(loss)/d(logits,value,hidden)
//             (ppo_atari_oc.py:566-602 forward, closed-form backward SURVEY.md A.4)
// One warp per sample; lane k holds logit k (n_actions <= 32).
#include "common.cuh"
#include <stdlib.h>
#include "philox.cuh"
#include <math_constants.h>
#include <string.h>

namespace ocppo {

namespace {

constexpr unsigned kFull = 0xffffffffu;
constexpr int kHeadThreads = 256;
constexpr int kWarpsPerBlock = kHeadThreads / 32;

// butterfly reductions over aligned groups of G lanes (G = 32: the whole warp)
template <int G = 32>
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    return v;
}
template <int G = 32>
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(kFull, v, o));
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    return v;
}

// torch/distributions/categorical.py:78 (logits - logsumexp), logits_to_probs
// (softmax of the normalised logits), :159-163 (entropy).
// G < 32: one distribution per aligned group of G lanes (`lane` is then the lane's index inside its group); the terms of
// lanes >= A are exact zeros, so the sums are bit-identical to the whole-warp form.
template <int G = 32>
__device__ __forceinline__ void warp_categorical(float z, int lane, int A, float &ell, float &p, float &H) {
    const bool on = lane < A;
    const float m = warp_max<G>(on ? z : -CUDART_INF_F);
    const float s = warp_sum<G>(on ? expf(z - m) : 0.f);
    ell = z - (m + logf(s));
    const float m2 = warp_max<G>(on ? ell : -CUDART_INF_F);
    const float e2 = on ? expf(ell - m2) : 0.f;
    const float s2 = warp_sum<G>(e2);
    p = __fdiv_rn(e2, s2);
    H = -warp_sum<G>(on ? fmaxf(ell, -3.402823466e+38f) * p : 0.f);
}

// torch multinomial single-draw path: argmax_k p_k / q_k, first index on ties.
__device__ __forceinline__ int warp_sample(float p, int lane, int A, const PhiloxLayout &L, uint64_t row) {
    float r = -CUDART_INF_F;
    if (lane < A) {
        const uint32_t w = torch_philox_word(L, row * (uint64_t)A + (uint64_t)lane);
        const float q = torch_exponential_from_uniform(curand_uniform_from_word(w));
        r = __fdiv_rn(p, q);
    }
    int k = lane;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float r2 = __shfl_xor_sync(kFull, r, o);
        const int k2 = __shfl_xor_sync(kFull, k, o);
        if (r2 > r || (r2 == r && k2 < k)) { r = r2; k = k2; }
    }
    return k;
}

// Stage actor/critic weights: sW[(A+1)][H] (row A = critic), sB[A+1].
__device__ __forceinline__ void stage_head_weights(float *sW, float *sB, const float *Wa, const float *ba,
                                                   const float *Wc, const float *bc, int A, int H) {
    for (int i = threadIdx.x; i < A * H; i += blockDim.x) sW[i] = Wa[i];
    for (int i = threadIdx.x; i < H; i += blockDim.x) sW[A * H + i] = Wc[i];
    for (int i = threadIdx.x; i < A; i += blockDim.x) sB[i] = ba[i];
    if (threadIdx.x == 0) sB[A] = bc[0];
    __syncthreads();
}

// lane k (k < A) returns logit k; value broadcast to all lanes.
__device__ __forceinline__ void warp_head_forward(const float *__restrict__ h, const float *sW, const float *sB,
                                                  int A, int H, int lane, float &z, float &v) {
    z = 0.f;
    for (int k = 0; k <= A; ++k) {
        float part = 0.f;
        for (int j = lane; j < H; j += 32) part = fmaf(sW[k * H + j], h[j], part);
        part = warp_sum(part) + sB[k];
        if (k == A) v = part;
        else if (lane == k) z = part;
    }
}

// Head forward from a hidden row held in registers (NV float4 per lane): rows k_first..A of sW, eight at a time with
// independent accumulators so that the shared-memory loads, FMA chains and shuffle reductions of different rows overlap
// (one row after the other was a 250-cycle dependent chain per row).  Lane k keeps logit k; v is broadcast.
template <int NV>
__device__ __forceinline__ void head_forward_vec(const float4 (&hv)[NV], const float *sW, const float *sB, int A, int H,
                                                 int lane, int k_first, float &z, float &v) {
    for (int k0 = k_first; k0 <= A; k0 += 8) {
        float part[8];
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) part[kk] = 0.f;
#pragma unroll
        for (int g = 0; g < NV; ++g) {
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) {
                const int k = k0 + kk <= A ? k0 + kk : A;            // rows past A repeat row A and are ignored
                const float4 w = *(reinterpret_cast<const float4 *>(sW + k * H + g * 128) + lane);
                part[kk] = fmaf(w.x, hv[g].x, part[kk]);
                part[kk] = fmaf(w.y, hv[g].y, part[kk]);
                part[kk] = fmaf(w.z, hv[g].z, part[kk]);
                part[kk] = fmaf(w.w, hv[g].w, part[kk]);
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) part[kk] += __shfl_xor_sync(kFull, part[kk], o);
        }
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
            const int k = k0 + kk;
            if (k == A) v = part[kk] + sB[A];
            else if (k < A && lane == k) z = part[kk] + sB[k];
        }
    }
}

// HV > 0: H == 128 * HV, the hidden row is read once with 16-byte loads (next row prefetched); HV == 0: any H.
template <int HV>
__global__ void __launch_bounds__(kHeadThreads)
policy_head_kernel(const float *__restrict__ hidden, int64_t B, int H, int A, const float *Wa, const float *ba,
                   const float *Wc, const float *bc, PhiloxLayout L, const uint64_t *__restrict__ philox_dev, int64_t row_base,
                   const int64_t *__restrict__ given, int64_t *__restrict__ action, float *__restrict__ logp,
                   float *__restrict__ entropy, float *__restrict__ value, int want_policy) {
    extern __shared__ __align__(16) float sm[];
    // a kernel launched as a programmatic dependent of this one (the GAE scan after the bootstrap ocppo_value call) may
    // start its prologue now; it waits for this grid's completion before it reads the values written below
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    // graph-replayable calls keep (seed, offset) in device memory: kernel arguments are frozen at capture time
    if (philox_dev != nullptr) { L.seed = __ldg(philox_dev); L.offset = __ldg(philox_dev + 1); }
    float *sW = sm, *sB = sm + (size_t)(A + 1) * H;
    stage_head_weights(sW, sB, Wa, ba, Wc, bc, A, H);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t i0 = (int64_t)blockIdx.x * kWarpsPerBlock + warp, istep = (int64_t)gridDim.x * kWarpsPerBlock;
    constexpr int NV = HV > 0 ? HV : 1;
    float4 hn[NV];
    if (HV > 0 && i0 < B) {
#pragma unroll
        for (int g = 0; g < NV; ++g) hn[g] = __ldcs(reinterpret_cast<const float4 *>(hidden + i0 * H + g * 128) + lane);
    }
    for (int64_t i = i0; i < B; i += istep) {
        float z = 0.f, v = 0.f;
        if (HV > 0) {
            float4 hv[NV];
#pragma unroll
            for (int g = 0; g < NV; ++g) hv[g] = hn[g];
            if (i + istep < B) {
#pragma unroll
                for (int g = 0; g < NV; ++g)
                    hn[g] = __ldcs(reinterpret_cast<const float4 *>(hidden + (i + istep) * H + g * 128) + lane);
            }
            head_forward_vec<NV>(hv, sW, sB, A, H, lane, want_policy ? 0 : A, z, v);
        } else {
            warp_head_forward(hidden + i * H, sW, sB, A, H, lane, z, v);
        }
        if (value != nullptr && lane == 0) value[i] = v;
        if (!want_policy) continue;
        float ell, p, Hh;
        warp_categorical(z, lane, A, ell, p, Hh);
        int a;
        if (given != nullptr) a = (int)given[i];
        else a = warp_sample(p, lane, A, L, (uint64_t)(row_base + i));
        const float lp = __shfl_sync(kFull, ell, a & 31);
        if (lane == 0) {
            if (action != nullptr) action[i] = (int64_t)a;
            if (logp != nullptr) logp[i] = lp;
            if (entropy != nullptr) entropy[i] = Hh;
        }
    }
}

__global__ void __launch_bounds__(kHeadThreads)
categorical_kernel(const float *__restrict__ logits, int64_t B, int A, PhiloxLayout L,
                   const int64_t *__restrict__ given, int64_t *__restrict__ action,
                   float *__restrict__ logp, float *__restrict__ entropy) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int64_t i = (int64_t)blockIdx.x * kWarpsPerBlock + warp; i < B; i += (int64_t)gridDim.x * kWarpsPerBlock) {
        const float z = lane < A ? logits[i * A + lane] : 0.f;
        float ell, p, Hh;
        warp_categorical(z, lane, A, ell, p, Hh);
        int a;
        if (given != nullptr) a = (int)given[i];
        else a = warp_sample(p, lane, A, L, (uint64_t)i);
        const float lp = __shfl_sync(kFull, ell, a & 31);
        if (lane == 0) {
            if (action != nullptr) action[i] = (int64_t)a;
            if (logp != nullptr) logp[i] = lp;
            if (entropy != nullptr) entropy[i] = Hh;
        }
    }
}

__global__ void exponential_kernel(float *__restrict__ out, int64_t numel, PhiloxLayout L) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < numel; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = torch_exponential_from_uniform(curand_uniform_from_word(torch_philox_word(L, (uint64_t)i)));
}

}  // namespace

// ---------------------------------------------------------------------------
// PPO loss head, forward + backward in one pass (ppo_atari_oc.py:566-602).
// ---------------------------------------------------------------------------

namespace {

// Per-sample PPO loss terms from (logit of this lane, value): returns d loss / d logit of this lane and d loss / d value,
// adds the logged sums into acc[] on lane 0.
// The sample's five gathered scalars, one per lane (0 action, 1 old logp, 2 advantage, 3 return, 4 old value): a single
// coalesced-ish load per sample that the vector kernel issues one sample ahead (as five dependent scalar loads they were
// 20 % of the kernel's stall samples).
__device__ __forceinline__ float load_sample_scalar(const LossArgs &p, int64_t i, int lane) {
    if (lane == 0) return __int_as_float(p.action[i]);
    if (lane == 1) return p.logp_old[i];
    if (lane == 2) return p.adv[i];
    if (lane == 3) return p.ret[i];
    if (lane == 4) return p.val[i];
    return 0.f;
}

// G < 32: one sample per aligned group of G lanes (G >= 8 > the five scalars); `valid` = this group's sample exists
// (a group past the end still takes part in the shuffles but stores and accumulates nothing).
template <int G = 32>
__device__ __forceinline__ void ppo_sample_loss(const LossArgs &p, int64_t i, float sc, float z, float v, int wlane, float inv_mb,
                                                float mean, float stdv, float &dz, float &dv, double (&acc)[6],
                                                bool valid = true) {
    const int A = p.A;
    const int lane = wlane & (G - 1), base = wlane & ~(G - 1);
    float ell, pk, Hh;
    warp_categorical<G>(z, lane, A, ell, pk, Hh);
    const int a = __float_as_int(__shfl_sync(kFull, sc, base));
    const float newlogp = __shfl_sync(kFull, ell, base + (a & (G - 1)));
    const float logratio = newlogp - __shfl_sync(kFull, sc, base + 1);
    const float ratio = expf(logratio);
    float advn = __shfl_sync(kFull, sc, base + 2);
    if (p.norm_adv) advn = __fdiv_rn(advn - mean, stdv + 1e-8f);
    // clipped surrogate (:581-583)
    const float l1 = -advn * ratio;
    const float rc = fminf(fmaxf(ratio, p.clip_lo), p.clip_hi);
    const float l2 = -advn * rc;
    const float pg = fmaxf(l1, l2);
    const bool in_range = (ratio >= p.clip_lo) && (ratio <= p.clip_hi);
    float g_ratio;                       // d pg / d ratio
    if (l1 > l2) g_ratio = -advn;
    else if (l2 > l1) g_ratio = in_range ? -advn : 0.f;
    else g_ratio = 0.5f * (-advn) + (in_range ? 0.5f * (-advn) : 0.f);
    // value loss (:585-597)
    const float R = __shfl_sync(kFull, sc, base + 3), Vold = __shfl_sync(kFull, sc, base + 4);
    const float du = v - R;
    const float u = du * du;
    float vterm = u, g_v = du;           // d (0.5*vterm) / d v
    if (p.clip_vloss) {
        const float d = v - Vold;
        const float dc = fminf(fmaxf(d, -p.clip_f), p.clip_f);
        const float vc = Vold + dc;
        const float dw = vc - R;
        const float w = dw * dw;
        const bool v_in = (d >= -p.clip_f) && (d <= p.clip_f);
        const float gw = v_in ? dw : 0.f;
        if (u > w) { vterm = u; g_v = du; }
        else if (w > u) { vterm = w; g_v = gw; }
        else { vterm = u; g_v = 0.5f * du + 0.5f * gw; }
    }
    // gradients wrt logits / value
    const float coef_a = g_ratio * ratio * inv_mb;
    const float coef_e = p.ent_coef * inv_mb;
    dz = 0.f;
    if (lane < A) dz = coef_a * ((lane == a ? 1.f : 0.f) - pk) + coef_e * pk * (ell + Hh);
    dv = p.vf_coef * inv_mb * g_v;
    if (valid && lane < p.AP) p.gz[i * p.AP + lane] = lane < A ? dz : (lane == A ? dv : 0.f);
    if (valid && lane == 0) {
        acc[0] += (double)pg;
        acc[1] += (double)vterm;
        acc[2] += (double)Hh;
        acc[3] += (double)(-logratio);
        acc[4] += (double)((ratio - 1.0f) - logratio);
        acc[5] += (fabsf(ratio - 1.0f) > p.clip_f) ? 1.0 : 0.0;
    }
}

// HV > 0: H == 128 * HV, the sample's hidden row lives in registers (HV float4 per lane, read once with 16-byte
// loads, the next sample's row prefetched while this one is processed) and d(hidden) leaves as float4 stores: the
// kernel then runs at HBM rate (8*H bytes per sample).  HV == 0: any H, scalar loops.
// BWD = false: the forward half only (logits, value, losses, gz = d loss / d (logits, value)) plus the bound ingredient
// max_i (|dv_i| + sum_k |dz_ik|); loss_bwd_pair_kernel then writes d(hidden) straight as an fp16 pair.
template <int HV, bool BWD>
__global__ void __launch_bounds__(kHeadThreads) loss_head_kernel(LossArgs p) {
    extern __shared__ __align__(16) float sm[];
    const int A = p.A, H = p.H;
    float *sW = sm, *sB = sm + (size_t)(A + 1) * H;
    stage_head_weights(sW, sB, p.Wa, p.ba, p.Wc, p.bc, A, H);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float inv_mb = 1.0f / (float)p.mb;
    float mean = 0.f, stdv = 1.f;
    if (p.norm_adv) { mean = p.adv_stat[0]; stdv = p.adv_stat[1]; }

    double acc[6] = {0, 0, 0, 0, 0, 0};   // pg, vmax, entropy, -logratio, (ratio-1)-logratio, clipped
    float dmax = 0.f;                     // max |dh| written by this lane
    float l1max = 0.f;                    // forward-only: max over this warp's samples of |dv| + sum_k |dz_k|
    const int64_t i0 = (int64_t)blockIdx.x * kWarpsPerBlock + warp, istep = (int64_t)gridDim.x * kWarpsPerBlock;
    if (HV > 0) {
        constexpr int NV = HV > 0 ? HV : 1;
        float4 hn[NV];
        float sc_next = 0.f;
        if (i0 < p.rows) {
#pragma unroll
            for (int g = 0; g < NV; ++g) hn[g] = __ldcs(reinterpret_cast<const float4 *>(p.hidden + i0 * H + g * 128) + lane);
            sc_next = load_sample_scalar(p, i0, lane);
        }
        for (int64_t i = i0; i < p.rows; i += istep) {
            float4 hv[NV];
#pragma unroll
            for (int g = 0; g < NV; ++g) hv[g] = hn[g];
            const float sc = sc_next;
            if (i + istep < p.rows) {
#pragma unroll
                for (int g = 0; g < NV; ++g)
                    hn[g] = __ldcs(reinterpret_cast<const float4 *>(p.hidden + (i + istep) * H + g * 128) + lane);
                sc_next = load_sample_scalar(p, i + istep, lane);
            }
            float z = 0.f, v = 0.f;
            head_forward_vec<NV>(hv, sW, sB, A, H, lane, 0, z, v);
            float dz, dv;
            ppo_sample_loss(p, i, sc, z, v, lane, inv_mb, mean, stdv, dz, dv, acc);
            if (!BWD) {
                l1max = fmaxf(l1max, fabsf(dv) + warp_sum(fabsf(dz)));
                continue;
            }
            // d hidden = dv * Wc + sum_k dz_k * Wa[k], masked by the ReLU of the last trunk layer
            float4 s[NV];
#pragma unroll
            for (int g = 0; g < NV; ++g) {
                const float4 w = *(reinterpret_cast<const float4 *>(sW + A * H + g * 128) + lane);
                s[g] = make_float4(dv * w.x, dv * w.y, dv * w.z, dv * w.w);
            }
            for (int k0 = 0; k0 < A; k0 += 4) {      // four head rows at a time: their loads are issued together
#pragma unroll
                for (int kk = 0; kk < 4; ++kk) {
                    const int k = k0 + kk < A ? k0 + kk : A - 1;
                    const float dzk = k0 + kk < A ? __shfl_sync(kFull, dz, k) : 0.f;     // (uniform condition)
#pragma unroll
                    for (int g = 0; g < NV; ++g) {
                        const float4 w = *(reinterpret_cast<const float4 *>(sW + k * H + g * 128) + lane);
                        s[g].x = fmaf(dzk, w.x, s[g].x);
                        s[g].y = fmaf(dzk, w.y, s[g].y);
                        s[g].z = fmaf(dzk, w.z, s[g].z);
                        s[g].w = fmaf(dzk, w.w, s[g].w);
                    }
                }
            }
#pragma unroll
            for (int g = 0; g < NV; ++g) {
                const float4 o = make_float4(hv[g].x > 0.f ? s[g].x : 0.f, hv[g].y > 0.f ? s[g].y : 0.f,
                                             hv[g].z > 0.f ? s[g].z : 0.f, hv[g].w > 0.f ? s[g].w : 0.f);
                *(reinterpret_cast<float4 *>(p.dh + i * H + g * 128) + lane) = o;
                dmax = fmaxf(fmaxf(dmax, fmaxf(fabsf(o.x), fabsf(o.y))), fmaxf(fabsf(o.z), fabsf(o.w)));
            }
        }
    } else {
        for (int64_t i = i0; i < p.rows; i += istep) {
            const float *h = p.hidden + i * H;
            float z, v;
            warp_head_forward(h, sW, sB, A, H, lane, z, v);
            float dz, dv;
            ppo_sample_loss(p, i, load_sample_scalar(p, i, lane), z, v, lane, inv_mb, mean, stdv, dz, dv, acc);
            if (!BWD) {
                l1max = fmaxf(l1max, fabsf(dv) + warp_sum(fabsf(dz)));
                continue;
            }
            // d hidden, masked by ReLU of the last trunk layer
            // (warp-uniform trip count: the shuffles below need all 32 lanes, also when H % 32 != 0)
            for (int j0 = 0; j0 < H; j0 += 32) {
                const int j = j0 + lane;
                const bool on = j < H;
                float s = on ? dv * sW[A * H + j] : 0.f;
                for (int k = 0; k < A; ++k) {
                    const float dzk = __shfl_sync(kFull, dz, k);
                    if (on) s = fmaf(dzk, sW[k * H + j], s);
                }
                if (on) {
                    const float o = h[j] > 0.f ? s : 0.f;
                    p.dh[i * H + j] = o;
                    dmax = fmaxf(dmax, fabsf(o));
                }
            }
        }
    }
    if (BWD && p.dh_amax != nullptr) {
        const unsigned m = __reduce_max_sync(kFull, __float_as_uint(dmax));
        if (lane == 0 && m != 0u) atomicMax(reinterpret_cast<unsigned int *>(p.dh_amax), m);
    }
    if (!BWD && p.gz_l1max != nullptr) {
        const unsigned m = __reduce_max_sync(kFull, __float_as_uint(l1max));
        if (lane == 0 && m != 0u) atomicMax(reinterpret_cast<unsigned int *>(p.gz_l1max), m);
    }
    __shared__ double red[kWarpsPerBlock][6];
    if (lane == 0)
        for (int q = 0; q < 6; ++q) red[warp][q] = acc[q];
    __syncthreads();
    if (threadIdx.x < 6) {
        double s = 0;
        for (int w = 0; w < kWarpsPerBlock; ++w) s += red[w][threadIdx.x];
        p.part[(int64_t)blockIdx.x * 8 + threadIdx.x] = s;
    }
}

// Forward half of the loss head, FOUR samples per warp pass (A + 1 <= 8 outputs, H = 128 * HV <= 512).
// The one-sample-per-warp kernel above spends 780 warp instructions and 16 KB of shared-memory weight reads on every sample
// (ncu: 43 % issue-active, short-scoreboard stalls, 5x the HBM time of the 2 KB row it reads).  Here each head-weight
// float4 is read from shared memory once per four samples, the 32 (sample, output) partial sums are reduced with a
// transposed butterfly (31 shuffles for four samples instead of 40 per sample; lane 8 s + k ends up owning output k of
// sample s -- the same pairwise sums as the plain butterfly, so the logits are bit-identical), and the categorical /
// PPO-loss arithmetic then runs for four samples at once in the four 8-lane groups.  The next four rows are requested as
// soon as the dot products are done, so the reduction and loss arithmetic hide the HBM latency without a second register
// set.  Forward only: d(hidden) is written by loss_bwd_pair_kernel.
template <int HV>
__global__ void __launch_bounds__(kHeadThreads, 2) loss_head_quad_kernel(LossArgs p) {
    extern __shared__ __align__(16) float sm[];
    const int A = p.A, H = p.H;
    float *sW = sm, *sB = sm + (size_t)(A + 1) * H;
    stage_head_weights(sW, sB, p.Wa, p.ba, p.Wc, p.bc, A, H);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int grp = lane >> 3, sub = lane & 7;
    const float inv_mb = 1.0f / (float)p.mb;
    float mean = 0.f, stdv = 1.f;
    if (p.norm_adv) { mean = p.adv_stat[0]; stdv = p.adv_stat[1]; }
    const float bias = sB[sub < A ? sub : A];
    const int64_t last = p.rows - 1;

    double acc[6] = {0, 0, 0, 0, 0, 0};   // lanes 0, 8, 16, 24: the sums of their group's samples
    float l1max = 0.f;
    const int64_t q0 = (int64_t)blockIdx.x * kWarpsPerBlock + warp, qstep = (int64_t)gridDim.x * kWarpsPerBlock;
    const int64_t quads = (p.rows + 3) >> 2;
    float4 hv[4][HV];
    if (q0 < quads) {
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            const int64_t r = q0 * 4 + s < last ? q0 * 4 + s : last;
#pragma unroll
            for (int g = 0; g < HV; ++g) hv[s][g] = __ldcs(reinterpret_cast<const float4 *>(p.hidden + r * H + g * 128) + lane);
        }
    }
    for (int64_t q = q0; q < quads; q += qstep) {
        const int64_t i = q * 4 + grp;                       // this group's sample
        const bool valid = i <= last;
        const float sc = (valid && sub < 5) ? load_sample_scalar(p, i, sub) : 0.f;
        float part[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) part[j] = 0.f;
#pragma unroll
        for (int g = 0; g < HV; ++g) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int kr = k <= A ? k : A;               // rows past A repeat row A and are ignored
                const float4 w = *(reinterpret_cast<const float4 *>(sW + kr * H + g * 128) + lane);
#pragma unroll
                for (int s = 0; s < 4; ++s) {
                    float a = part[s * 8 + k];
                    a = fmaf(w.x, hv[s][g].x, a);
                    a = fmaf(w.y, hv[s][g].y, a);
                    a = fmaf(w.z, hv[s][g].z, a);
                    a = fmaf(w.w, hv[s][g].w, a);
                    part[s * 8 + k] = a;
                }
            }
        }
        if (q + qstep < quads) {                             // the rows of the next pass, into the registers just freed
#pragma unroll
            for (int s = 0; s < 4; ++s) {
                const int64_t r = (q + qstep) * 4 + s < last ? (q + qstep) * 4 + s : last;
#pragma unroll
                for (int g = 0; g < HV; ++g)
                    hv[s][g] = __ldcs(reinterpret_cast<const float4 *>(p.hidden + r * H + g * 128) + lane);
            }
        }
        // transposed butterfly: after the step with offset o a lane keeps the values whose index agrees with its own
        // lane number in bit o, having added its partner's partial sums of them
        float v16[16], v8[8], v4[4], v2[2];
        {
            const bool up = (lane & 16) != 0;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const float keep = up ? part[j + 16] : part[j], send = up ? part[j] : part[j + 16];
                v16[j] = keep + __shfl_xor_sync(kFull, send, 16);
            }
        }
        {
            const bool up = (lane & 8) != 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float keep = up ? v16[j + 8] : v16[j], send = up ? v16[j] : v16[j + 8];
                v8[j] = keep + __shfl_xor_sync(kFull, send, 8);
            }
        }
        {
            const bool up = (lane & 4) != 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float keep = up ? v8[j + 4] : v8[j], send = up ? v8[j] : v8[j + 4];
                v4[j] = keep + __shfl_xor_sync(kFull, send, 4);
            }
        }
        {
            const bool up = (lane & 2) != 0;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const float keep = up ? v4[j + 2] : v4[j], send = up ? v4[j] : v4[j + 2];
                v2[j] = keep + __shfl_xor_sync(kFull, send, 2);
            }
        }
        float tot;
        {
            const bool up = (lane & 1) != 0;
            const float keep = up ? v2[1] : v2[0], send = up ? v2[0] : v2[1];
            tot = keep + __shfl_xor_sync(kFull, send, 1);
        }
        tot += bias;                                         // lane 8 s + k: logit k (k < A) / value (k == A) of sample s
        const float z = tot;
        const float v = __shfl_sync(kFull, tot, (lane & ~7) + A);
        float dz, dv;
        ppo_sample_loss<8>(p, i, sc, z, v, lane, inv_mb, mean, stdv, dz, dv, acc, valid);
        const float l1 = fabsf(dv) + warp_sum<8>(fabsf(dz));
        if (valid) l1max = fmaxf(l1max, l1);
    }
    if (p.gz_l1max != nullptr) {
        const unsigned m = __reduce_max_sync(kFull, __float_as_uint(l1max));
        if (lane == 0 && m != 0u) atomicMax(reinterpret_cast<unsigned int *>(p.gz_l1max), m);
    }
    __shared__ double red[kWarpsPerBlock][6];
#pragma unroll
    for (int qq = 0; qq < 6; ++qq) {                         // the four group leaders, in order
        const double a0 = __shfl_sync(kFull, acc[qq], 0), a1 = __shfl_sync(kFull, acc[qq], 8);
        const double a2 = __shfl_sync(kFull, acc[qq], 16), a3 = __shfl_sync(kFull, acc[qq], 24);
        if (lane == 0) red[warp][qq] = ((a0 + a1) + a2) + a3;
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        double s = 0;
        for (int w = 0; w < kWarpsPerBlock; ++w) s += red[w][threadIdx.x];
        p.part[(int64_t)blockIdx.x * 8 + threadIdx.x] = s;
    }
}

// d(hidden) = mask * (dv * Wc + sum_k dz_k * Wa[k]) written as the fp16 pair the trunk's backward GEMMs load, plus its column
// sums (the last trunk layer's bias gradient) -- no fp32 copy of the gradient exists and no separate split pass runs.
// The pair's scale must be fixed before the first element exists: |dh_j| <= (|dv| + sum_k |dz_k|) * max|W_head|, with the
// first factor maximised over the minibatch chunk by the forward kernel.  Warp per sample, lane <-> 4 columns per
// 128-column group; ReLU decisions come as the packed words the last trunk GEMM left (16 bytes per sample instead of
// re-reading 2 KB of activations).  HBM-bound: 8 * H / 2 ... it writes 4 * H bytes and reads 4 * AP + H / 8 per sample.
template <int HV>
__global__ void __launch_bounds__(kHeadThreads) loss_bwd_pair_kernel(LossBwdPairArgs p) {
    extern __shared__ __align__(16) float sm[];
    const int A = p.A, H = p.H;
    float *sW = sm;                                            // [(A + 1)][H], row A = critic
    __shared__ float s_wmax[kWarpsPerBlock];
    float wmax = 0.f;
    for (int i = threadIdx.x; i < (A + 1) * H; i += blockDim.x) {
        const float w = i < A * H ? p.Wa[i] : p.Wc[i - A * H];
        sW[i] = w;
        wmax = fmaxf(wmax, fabsf(w));
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    wmax = __uint_as_float(__reduce_max_sync(kFull, __float_as_uint(wmax)));
    if (lane == 0) s_wmax[warp] = wmax;
    __syncthreads();
    wmax = 0.f;
#pragma unroll
    for (int w = 0; w < kWarpsPerBlock; ++w) wmax = fmaxf(wmax, s_wmax[w]);
    const float bound = __ldg(p.gz_l1max) * wmax;
    float scale, inv;
    f16_scale_of(bound, scale, inv);
    if (blockIdx.x == 0 && threadIdx.x == 0) *p.bound = bound;
    constexpr int NV = HV;
    const int wpr = H >> 5;                                    // ReLU decision words per sample
    float4 csum[NV];
#pragma unroll
    for (int g = 0; g < NV; ++g) csum[g] = make_float4(0.f, 0.f, 0.f, 0.f);
    float dmax = 0.f;
    const int64_t i0 = (int64_t)blockIdx.x * kWarpsPerBlock + warp, istep = (int64_t)gridDim.x * kWarpsPerBlock;
    float gz_next = 0.f;
    uint32_t bits_next = 0u;
    if (i0 < p.rows) {
        if (lane < p.AP) gz_next = __ldg(p.gz + i0 * p.AP + lane);
        if (p.relu_bits != nullptr && lane < wpr) bits_next = __ldg(p.relu_bits + i0 * wpr + lane);
    }
    for (int64_t i = i0; i < p.rows; i += istep) {
        const float gzv = gz_next;
        const uint32_t bitsv = bits_next;
        if (i + istep < p.rows) {                              // next sample's 48 bytes while this one is processed
            if (lane < p.AP) gz_next = __ldg(p.gz + (i + istep) * p.AP + lane);
            if (p.relu_bits != nullptr && lane < wpr) bits_next = __ldg(p.relu_bits + (i + istep) * wpr + lane);
        }
        const float dv = __shfl_sync(kFull, gzv, A);
        float4 s[NV];
#pragma unroll
        for (int g = 0; g < NV; ++g) {
            const float4 w = *(reinterpret_cast<const float4 *>(sW + A * H + g * 128) + lane);
            s[g] = make_float4(dv * w.x, dv * w.y, dv * w.z, dv * w.w);
        }
        for (int k = 0; k < A; ++k) {
            const float dzk = __shfl_sync(kFull, gzv, k);
#pragma unroll
            for (int g = 0; g < NV; ++g) {
                const float4 w = *(reinterpret_cast<const float4 *>(sW + k * H + g * 128) + lane);
                s[g].x = fmaf(dzk, w.x, s[g].x);
                s[g].y = fmaf(dzk, w.y, s[g].y);
                s[g].z = fmaf(dzk, w.z, s[g].z);
                s[g].w = fmaf(dzk, w.w, s[g].w);
            }
        }
#pragma unroll
        for (int g = 0; g < NV; ++g) {
            uint32_t nib;
            if (p.relu_bits != nullptr) {
                // columns g * 128 + 4 * lane .. + 3 sit in word g * 4 + lane / 8, bits 4 * (lane % 8) ..
                nib = (__shfl_sync(kFull, bitsv, g * 4 + (lane >> 3)) >> ((lane & 7) * 4)) & 0xFu;
            } else {
                const float4 hv = __ldcs(reinterpret_cast<const float4 *>(p.hidden + i * H + g * 128) + lane);
                nib = (hv.x > 0.f ? 1u : 0u) | (hv.y > 0.f ? 2u : 0u) | (hv.z > 0.f ? 4u : 0u) | (hv.w > 0.f ? 8u : 0u);
            }
            const float4 o = make_float4((nib & 1u) ? s[g].x : 0.f, (nib & 2u) ? s[g].y : 0.f, (nib & 4u) ? s[g].z : 0.f,
                                         (nib & 8u) ? s[g].w : 0.f);
            uint32_t h0, l0, h1, l1;
            split_f16x2(o.x, o.y, scale, h0, l0);
            split_f16x2(o.z, o.w, scale, h1, l1);
            const int64_t e = (i * H + g * 128) / 4 + lane;    // uint2 index: 4 fp16 per lane
            __stcs(reinterpret_cast<uint2 *>(p.dh_hi) + e, make_uint2(h0, h1));
            __stcs(reinterpret_cast<uint2 *>(p.dh_lo) + e, make_uint2(l0, l1));
            csum[g].x += o.x; csum[g].y += o.y; csum[g].z += o.z; csum[g].w += o.w;
            dmax = fmaxf(fmaxf(dmax, fmaxf(fabsf(o.x), fabsf(o.y))), fmaxf(fabsf(o.z), fabsf(o.w)));
        }
    }
    if (p.dh_amax != nullptr) {
        const unsigned m = __reduce_max_sync(kFull, __float_as_uint(dmax));
        if (lane == 0 && m != 0u) atomicMax(reinterpret_cast<unsigned int *>(p.dh_amax), m);
    }
    // column sums: the block's eight warps fold in fixed order through shared memory (the head weights are dead)
    __syncthreads();
    float4 *red = reinterpret_cast<float4 *>(sm);              // [kWarpsPerBlock][H / 4]
#pragma unroll
    for (int g = 0; g < NV; ++g) red[warp * (H >> 2) + g * 32 + lane] = csum[g];
    __syncthreads();
    for (int c = threadIdx.x; c < (H >> 2); c += blockDim.x) {
        float4 t = red[c];
        for (int w = 1; w < kWarpsPerBlock; ++w) {
            const float4 v = red[w * (H >> 2) + c];
            t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w;
        }
        reinterpret_cast<float4 *>(p.colsum_part + (int64_t)blockIdx.x * H)[c] = t;
    }
}

// The same pass for H = 512 with the head weights in REGISTERS: a lane always writes the same 16 columns, so its NR x 16
// weights (NR = A + 1 rows, critic last) are loaded once per block instead of 28 shared-memory float4 reads per sample
// (ncu on the kernel above: 61 % issue-active, short-scoreboard stalls on those reads), and the rank-1 updates run as packed
// FFMA2.  One block per SM (the weights alone take 16 NR registers); same products, same order, so d(hidden), its
// bound and its column-sum partial layout [blocks][H] are what the kernel above writes.
template <int NR>
__global__ void __launch_bounds__(kHeadThreads, 1) loss_bwd_pair_reg_kernel(LossBwdPairArgs p) {
    extern __shared__ __align__(16) float sm[];                // [kWarpsPerBlock][H / 4] float4 column-sum fold
    constexpr int H = 512, NV = 4;
    const int A = NR - 1;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float2 w2[NR][2 * NV];                                     // row k, columns g * 128 + 4 lane .. + 3 as two pairs per g
    float wmax = 0.f;
#pragma unroll
    for (int k = 0; k < NR; ++k) {
        const float *row = k < A ? p.Wa + (size_t)k * H : p.Wc;
#pragma unroll
        for (int g = 0; g < NV; ++g) {
            const float4 w = __ldg(reinterpret_cast<const float4 *>(row + g * 128) + lane);
            w2[k][2 * g] = make_float2(w.x, w.y);
            w2[k][2 * g + 1] = make_float2(w.z, w.w);
            wmax = fmaxf(fmaxf(wmax, fmaxf(fabsf(w.x), fabsf(w.y))), fmaxf(fabsf(w.z), fabsf(w.w)));
        }
    }
    wmax = __uint_as_float(__reduce_max_sync(kFull, __float_as_uint(wmax)));      // a warp covers all NR x 512 weights
    const float bound = __ldg(p.gz_l1max) * wmax;
    float scale, inv;
    f16_scale_of(bound, scale, inv);
    if (blockIdx.x == 0 && threadIdx.x == 0) *p.bound = bound;
    constexpr int wpr = H >> 5;                                // ReLU decision words per sample
    float4 csum[NV];
#pragma unroll
    for (int g = 0; g < NV; ++g) csum[g] = make_float4(0.f, 0.f, 0.f, 0.f);
    float dmax = 0.f;
    const int64_t i0 = (int64_t)blockIdx.x * kWarpsPerBlock + warp, istep = (int64_t)gridDim.x * kWarpsPerBlock;
    // the sample's 48 input bytes are requested two samples ahead (one ahead left the eight warps of the block waiting:
    // ncu long-scoreboard 2.6 per issue at 40 % issue-active)
    float gz_n1 = 0.f, gz_n2 = 0.f;
    uint32_t bits_n1 = 0u, bits_n2 = 0u;
    if (i0 < p.rows) {
        if (lane < p.AP) gz_n1 = __ldg(p.gz + i0 * p.AP + lane);
        if (lane < wpr) bits_n1 = __ldg(p.relu_bits + i0 * wpr + lane);
    }
    if (i0 + istep < p.rows) {
        if (lane < p.AP) gz_n2 = __ldg(p.gz + (i0 + istep) * p.AP + lane);
        if (lane < wpr) bits_n2 = __ldg(p.relu_bits + (i0 + istep) * wpr + lane);
    }
    for (int64_t i = i0; i < p.rows; i += istep) {
        const float gzv = gz_n1;
        const uint32_t bitsv = bits_n1;
        gz_n1 = gz_n2; bits_n1 = bits_n2;
        if (i + 2 * istep < p.rows) {
            if (lane < p.AP) gz_n2 = __ldg(p.gz + (i + 2 * istep) * p.AP + lane);
            if (lane < wpr) bits_n2 = __ldg(p.relu_bits + (i + 2 * istep) * wpr + lane);
        }
        const float dv = __shfl_sync(kFull, gzv, A);
        const float2 dv2 = make_float2(dv, dv);
        float2 s2[2 * NV];
#pragma unroll
        for (int j = 0; j < 2 * NV; ++j) s2[j] = __fmul2_rn(dv2, w2[NR - 1][j]);
#pragma unroll
        for (int k = 0; k < NR - 1; ++k) {
            const float dzk = __shfl_sync(kFull, gzv, k);
            const float2 dz2 = make_float2(dzk, dzk);
#pragma unroll
            for (int j = 0; j < 2 * NV; ++j) s2[j] = __ffma2_rn(dz2, w2[k][j], s2[j]);
        }
#pragma unroll
        for (int g = 0; g < NV; ++g) {
            // columns g * 128 + 4 * lane .. + 3 sit in word g * 4 + lane / 8, bits 4 * (lane % 8) ..
            const uint32_t nib = (__shfl_sync(kFull, bitsv, g * 4 + (lane >> 3)) >> ((lane & 7) * 4)) & 0xFu;
            const float4 o = make_float4((nib & 1u) ? s2[2 * g].x : 0.f, (nib & 2u) ? s2[2 * g].y : 0.f,
                                         (nib & 4u) ? s2[2 * g + 1].x : 0.f, (nib & 8u) ? s2[2 * g + 1].y : 0.f);
            uint32_t h0, l0, h1, l1;
            split_f16x2(o.x, o.y, scale, h0, l0);
            split_f16x2(o.z, o.w, scale, h1, l1);
            const int64_t e = (i * H + g * 128) / 4 + lane;    // uint2 index: 4 fp16 per lane
            __stcs(reinterpret_cast<uint2 *>(p.dh_hi) + e, make_uint2(h0, h1));
            __stcs(reinterpret_cast<uint2 *>(p.dh_lo) + e, make_uint2(l0, l1));
            csum[g].x += o.x; csum[g].y += o.y; csum[g].z += o.z; csum[g].w += o.w;
            dmax = fmaxf(fmaxf(dmax, fmaxf(fabsf(o.x), fabsf(o.y))), fmaxf(fabsf(o.z), fabsf(o.w)));
        }
    }
    if (p.dh_amax != nullptr) {
        const unsigned m = __reduce_max_sync(kFull, __float_as_uint(dmax));
        if (lane == 0 && m != 0u) atomicMax(reinterpret_cast<unsigned int *>(p.dh_amax), m);
    }
    float4 *red = reinterpret_cast<float4 *>(sm);              // [kWarpsPerBlock][H / 4]
#pragma unroll
    for (int g = 0; g < NV; ++g) red[warp * (H >> 2) + g * 32 + lane] = csum[g];
    __syncthreads();
    for (int c = threadIdx.x; c < (H >> 2); c += blockDim.x) {
        float4 t = red[c];
        for (int w = 1; w < kWarpsPerBlock; ++w) {
            const float4 v = red[w * (H >> 2) + c];
            t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w;
        }
        reinterpret_cast<float4 *>(p.colsum_part + (int64_t)blockIdx.x * H)[c] = t;
    }
}

// Single block: ordered sum of the block partials -> the seven logged scalars.
__global__ void loss_finalize_kernel(const double *__restrict__ part, int n_part, int64_t mb, float ent_coef,
                                     float vf_coef, float *__restrict__ stats) {
    // one warp per statistic (blockDim = 6 x 32): lane l sums partials l, l + 32, ... in order, then a fixed shuffle
    // tree -- deterministic, and ~30 dependent global loads instead of ~600 in one thread
    __shared__ double s[6];
    const int q = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (q < 6) {
        double a = 0;
        for (int b = lane; b < n_part; b += 32) a += part[(int64_t)b * 8 + q];
        a = warp_sum_d(a);
        if (lane == 0) s[q] = a / (double)mb;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const float pg = (float)s[0], vl = 0.5f * (float)s[1], en = (float)s[2];
        stats[OCPPO_STAT_PG_LOSS] = pg;
        stats[OCPPO_STAT_V_LOSS] = vl;
        stats[OCPPO_STAT_ENTROPY] = en;
        stats[OCPPO_STAT_LOSS] = pg - ent_coef * en + vl * vf_coef;
        stats[OCPPO_STAT_OLD_APPROX_KL] = (float)s[3];
        stats[OCPPO_STAT_APPROX_KL] = (float)s[4];
        stats[OCPPO_STAT_CLIPFRAC] = (float)s[5];
        stats[OCPPO_STAT_RESERVED] = 0.f;
    }
}

// ---------------------------------------------------------------------------
// minibatch gather (ppo_atari_oc.py:566-567,569,577,587-593) + advantage stats (:579)
// ---------------------------------------------------------------------------
__global__ void gather_kernel(const float *__restrict__ b_obs, const int64_t *__restrict__ b_actions,
                              const float *__restrict__ b_logp, const float *__restrict__ b_adv,
                              const float *__restrict__ b_ret, const float *__restrict__ b_val,
                              const int64_t *__restrict__ idx, int64_t rows, int obs_floats,
                              float *__restrict__ x0, int32_t *__restrict__ action, float *__restrict__ logp,
                              float *__restrict__ adv, float *__restrict__ ret, float *__restrict__ val) {
    const int64_t total = rows * obs_floats;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = e / obs_floats;
        const int c = (int)(e - r * obs_floats);
        const int64_t src = idx[r];
        x0[e] = b_obs[src * obs_floats + c];
        if (c == 0) {
            action[r] = (int32_t)b_actions[src];
            logp[r] = b_logp[src];
            adv[r] = b_adv[src];
            ret[r] = b_ret[src];
            val[r] = b_val[src];
        }
    }
}

__global__ void __launch_bounds__(256)
adv_stats_partial_kernel(const float *__restrict__ b_adv, const int64_t *__restrict__ idx, int64_t mb,
                         double *__restrict__ part) {
    double s = 0, s2 = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < mb; i += (int64_t)gridDim.x * blockDim.x) {
        const double x = (double)b_adv[idx[i]];
        s += x;
        s2 += x * x;
    }
    s = warp_sum_d(s);
    s2 = warp_sum_d(s2);
    __shared__ double red[8][2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { red[warp][0] = s; red[warp][1] = s2; }
    __syncthreads();
    if (threadIdx.x < 2) {
        double a = 0;
        for (int w = 0; w < 8; ++w) a += red[w][threadIdx.x];
        part[(int64_t)blockIdx.x * 2 + threadIdx.x] = a;
    }
}

// Explained variance of the value function (ppo_atari_oc.py:620-623: np.var(y_true), np.var(y_true - y_pred), population
// variances; y_true - y_pred is an fp32 subtraction there too).  Sums in fp64, ordered second stage.
__global__ void __launch_bounds__(256)
explained_var_partial_kernel(const float *__restrict__ y_pred, const float *__restrict__ y_true, int64_t n,
                             double *__restrict__ part) {
    double s[4] = {0, 0, 0, 0};          // sum y, sum y^2, sum d, sum d^2
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float yt = y_true[i];
        const double y = (double)yt, d = (double)__fsub_rn(yt, y_pred[i]);
        s[0] += y; s[1] += y * y; s[2] += d; s[3] += d * d;
    }
    __shared__ double red[8][4];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const double r = warp_sum_d(s[q]);
        if (lane == 0) red[warp][q] = r;
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        double a = 0;
        for (int w = 0; w < 8; ++w) a += red[w][threadIdx.x];
        part[(int64_t)blockIdx.x * 4 + threadIdx.x] = a;
    }
}

__global__ void explained_var_finalize_kernel(const double *__restrict__ part, int n_part, int64_t n,
                                              float *__restrict__ out) {
    if (threadIdx.x == 0) {
        double s[4] = {0, 0, 0, 0};
        for (int b = 0; b < n_part; ++b)
            for (int q = 0; q < 4; ++q) s[q] += part[4 * b + q];
        const double my = s[0] / (double)n, md = s[2] / (double)n;
        double vy = s[1] / (double)n - my * my, vd = s[3] / (double)n - md * md;
        if (vy < 0) vy = 0;
        if (vd < 0) vd = 0;
        out[0] = (float)vy;               // np.var(y_true)
        out[1] = (float)vd;               // np.var(y_true - y_pred)
    }
}

__global__ void adv_stats_finalize_kernel(const double *__restrict__ part, int n_part, int64_t mb,
                                          float *__restrict__ stat) {
    if (threadIdx.x == 0) {
        double s = 0, s2 = 0;
        for (int b = 0; b < n_part; ++b) { s += part[2 * b]; s2 += part[2 * b + 1]; }
        const double mean = s / (double)mb;
        double var = (s2 - s * mean) / (double)(mb - 1);     // unbiased, torch .std() default
        if (var < 0 && mb > 1) var = 0;
        stat[0] = (float)mean;
        stat[1] = (float)sqrt(var);
    }
}

// out[i] = (accumulate ? out[i] : 0) + sum_s part[s*stride + i], fixed order.
__global__ void splitk_reduce_kernel(const float *__restrict__ part, int splits, int64_t stride, int64_t count,
                                     float *__restrict__ out, int accumulate) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int k = 0; k < splits; ++k) s += part[(int64_t)k * stride + i];
        out[i] = accumulate ? out[i] + s : s;
    }
}

__global__ void relu_mask_kernel(const float *__restrict__ g, const float *__restrict__ act, int64_t n,
                                 float *__restrict__ out) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = act[i] > 0.f ? g[i] : 0.f;
}

}  // namespace

// ---------------------------------------------------------------------------
// clip_grad_norm_ + Adam over the flat arenas (ppo_atari_oc.py:609-610; torch
// nn/utils/clip_grad.py, optim/adam.py single-tensor path), with
// the 1/world_size of ppo_atari_multigpu.py:368-374 folded in.
// HBM-bound: reads g (twice), p, m, v; writes g, p, m, v  = 8 x 4 B per param.
// ---------------------------------------------------------------------------
namespace {

__global__ void __launch_bounds__(256)
sumsq_kernel(const float *__restrict__ g, int64_t n, float inv_world, const float *__restrict__ dev_scalars,
             double *__restrict__ part) {
    if (dev_scalars != nullptr) inv_world = __ldg(dev_scalars);
    double s = 0;
    const int64_t n4 = n >> 2;
    const float4 *g4 = reinterpret_cast<const float4 *>(g);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 v = g4[i];
        const float a = v.x * inv_world, b = v.y * inv_world, c = v.z * inv_world, d = v.w * inv_world;
        s += (double)a * a + (double)b * b + (double)c * c + (double)d * d;
    }
    s = warp_sum_d(s);
    __shared__ double red[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) red[warp] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0;
        for (int w = 0; w < 8; ++w) a += red[w];
        part[blockIdx.x] = a;
    }
}

struct AdamArgs {
    float *p, *g, *m, *v;
    int64_t n;
    const double *norm_part;
    int n_part;
    float inv_world, max_norm, one_minus_b1, b2, one_minus_b2, step_size, bc2_sqrt, eps;
    float *grad_norm;
    const float *dev_scalars;     // graph-replayable calls: the eight scalars above in device memory (ocppo_adam_scalars order)
};

__global__ void __launch_bounds__(256) adam_kernel(AdamArgs a) {
    __shared__ float s_coef;
    if (a.dev_scalars != nullptr) {
        a.inv_world = __ldg(a.dev_scalars + 0); a.max_norm = __ldg(a.dev_scalars + 1); a.one_minus_b1 = __ldg(a.dev_scalars + 2);
        a.b2 = __ldg(a.dev_scalars + 3); a.one_minus_b2 = __ldg(a.dev_scalars + 4); a.step_size = __ldg(a.dev_scalars + 5);
        a.bc2_sqrt = __ldg(a.dev_scalars + 6); a.eps = __ldg(a.dev_scalars + 7);
    }
    if (threadIdx.x < 32) {
        double s = 0;
        for (int i = threadIdx.x; i < a.n_part; i += 32) s += a.norm_part[i];
        s = warp_sum_d(s);
        if (threadIdx.x == 0) {
            const float total = (float)sqrt(s);
            float coef = a.max_norm / (total + 1e-6f);
            coef = fminf(coef, 1.0f);
            s_coef = coef;
            if (blockIdx.x == 0 && a.grad_norm != nullptr) *a.grad_norm = total;
        }
    }
    __syncthreads();
    const float coef = s_coef;
    const int64_t n4 = a.n >> 2;
    float4 *p4 = reinterpret_cast<float4 *>(a.p), *g4 = reinterpret_cast<float4 *>(a.g);
    float4 *m4 = reinterpret_cast<float4 *>(a.m), *v4 = reinterpret_cast<float4 *>(a.v);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        float4 G = g4[i], P = p4[i], M = m4[i], V = v4[i];
        float *gg = &G.x, *pp = &P.x, *mm = &M.x, *vv = &V.x;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float gr = (gg[q] * a.inv_world) * coef;
            gg[q] = gr;
            mm[q] = mm[q] + (gr - mm[q]) * a.one_minus_b1;
            vv[q] = vv[q] * a.b2 + a.one_minus_b2 * gr * gr;
            const float denom = sqrtf(vv[q]) / a.bc2_sqrt + a.eps;
            pp[q] = pp[q] - a.step_size * (mm[q] / denom);
        }
        g4[i] = G; p4[i] = P; m4[i] = M; v4[i] = V;
    }
}

}  // namespace

// FP32 FMA peak probe (roofline denominator for the CUDA-core GEMMs; MEASURED_PEAKS.json
// has no fp32 figure).  16 independent FMA chains per thread.
namespace {
__global__ void __launch_bounds__(256) fma_peak_kernel(float *out, int iters, float a, float b) {
    float acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = (float)(threadIdx.x + i);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = fmaf(acc[i], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i];
    if (s == 123.456f) out[0] = s;          // never true; keeps the chains alive
}
}  // namespace

int measure_fma_peak(double *tflops, cudaStream_t stream) {
    int sms, mt;
    {
        int dev = 0;
        OCPPO_CUDA(cudaGetDevice(&dev));
        OCPPO_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        mt = 0;
    }
    (void)mt;
    float *dummy = nullptr;
    OCPPO_CUDA(cudaMalloc(&dummy, 256));
    cudaEvent_t e0, e1;
    OCPPO_CUDA(cudaEventCreate(&e0));
    OCPPO_CUDA(cudaEventCreate(&e1));
    const int iters = 8192, blocks = sms * 8;
    double best = 0;
    for (int rep = 0; rep < 5; ++rep) {
        OCPPO_CUDA(cudaEventRecord(e0, stream));
        fma_peak_kernel<<<blocks, 256, 0, stream>>>(dummy, iters, 0.999f, 0.001f);
        OCPPO_CUDA(cudaEventRecord(e1, stream));
        OCPPO_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        OCPPO_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double tf = 2.0 * blocks * 256.0 * 16.0 * iters / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(dummy);
    *tflops = best;
    return OCPPO_OK;
}

// ---------------------------------------------------------------------------
// host-side launchers
// ---------------------------------------------------------------------------
static inline unsigned grid_for(int64_t work_items, int per_block, int cap) {
    int64_t g = ceil_div(work_items, per_block);
    if (g < 1) g = 1;
    if (g > cap) g = cap;
    return (unsigned)g;
}

static int head_smem_bytes(int A, int H) { return (int)(((size_t)(A + 1) * H + (A + 1)) * sizeof(float)); }

static int ensure_head_smem(const void *fn, int bytes) {
    if (bytes > 48 * 1024) OCPPO_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    return OCPPO_OK;
}

int launch_policy_head(const ocppo_handle *h, const float *hidden, int64_t B, int64_t row_base, int64_t total_B,
                       uint64_t seed, uint64_t offset, const uint64_t *philox_dev, const int64_t *given, int64_t *action,
                       float *logp, float *entropy, float *value, int want_policy, cudaStream_t stream) {
    if (B == 0) return OCPPO_OK;
    const int A = h->cfg.n_actions, H = h->hidden;
    PhiloxLayout L{seed, offset, torch_philox_grid(total_B * A, h->sm_count, h->max_threads_per_sm) * 256};
    const int smem = head_smem_bytes(A, H);
    const bool aligned = (((uintptr_t)hidden & 15) == 0);
    auto kern = policy_head_kernel<0>;
    if (aligned && H == 128) kern = policy_head_kernel<1>;
    else if (aligned && H == 256) kern = policy_head_kernel<2>;
    else if (aligned && H == 512) kern = policy_head_kernel<4>;
    else if (aligned && H == 1024) kern = policy_head_kernel<8>;
    OCPPO_TRY(ensure_head_smem((const void *)kern, smem));
    const unsigned grid = grid_for(B, kWarpsPerBlock, h->sm_count * 4);
    ProfScope ps(KC_HEAD, 1, stream);
    kern<<<grid, kHeadThreads, smem, stream>>>(
        hidden, B, H, A, h->params + h->actor.w_off, h->params + h->actor.b_off, h->params + h->critic.w_off,
        h->params + h->critic.b_off, L, philox_dev, row_base, given, action, logp, entropy, value, want_policy);
    OCPPO_LAUNCH_CHECK("policy_head_kernel");
    return OCPPO_OK;
}

static int device_geometry(int *sm_count, int *max_threads) {
    int dev = 0;
    OCPPO_CUDA(cudaGetDevice(&dev));
    OCPPO_CUDA(cudaDeviceGetAttribute(sm_count, cudaDevAttrMultiProcessorCount, dev));
    OCPPO_CUDA(cudaDeviceGetAttribute(max_threads, cudaDevAttrMaxThreadsPerMultiProcessor, dev));
    return OCPPO_OK;
}

int launch_categorical(const float *logits, int64_t B, int A, uint64_t seed, uint64_t offset, const int64_t *given,
                       int64_t *action, float *logp, float *entropy, cudaStream_t stream) {
    if (B == 0) return OCPPO_OK;
    int sms, mt;
    OCPPO_TRY(device_geometry(&sms, &mt));
    PhiloxLayout L{seed, offset, torch_philox_grid(B * A, sms, mt) * 256};
    ProfScope ps(KC_HEAD, 1, stream);
    categorical_kernel<<<grid_for(B, kWarpsPerBlock, sms * 8), kHeadThreads, 0, stream>>>(logits, B, A, L, given,
                                                                                         action, logp, entropy);
    OCPPO_LAUNCH_CHECK("categorical_kernel");
    return OCPPO_OK;
}

int launch_exponential(float *out, int64_t numel, uint64_t seed, uint64_t offset, cudaStream_t stream) {
    if (numel == 0) return OCPPO_OK;
    int sms, mt;
    OCPPO_TRY(device_geometry(&sms, &mt));
    PhiloxLayout L{seed, offset, torch_philox_grid(numel, sms, mt) * 256};
    ProfScope ps(KC_AUX, 1, stream);
    exponential_kernel<<<grid_for(numel, 256, sms * 8), 256, 0, stream>>>(out, numel, L);
    OCPPO_LAUNCH_CHECK("exponential_kernel");
    return OCPPO_OK;
}

int launch_gather(const ocppo_handle *h, const float *b_obs, const int64_t *b_actions, const float *b_logp,
                  const float *b_adv, const float *b_ret, const float *b_val, const int64_t *idx, int64_t rows,
                  cudaStream_t stream) {
    const int obs_floats = h->cfg.frames * h->cfg.obs_dim;
    ProfScope ps(KC_AUX, 1, stream);
    gather_kernel<<<grid_for(rows * obs_floats, 256, h->sm_count * 8), 256, 0, stream>>>(
        b_obs, b_actions, b_logp, b_adv, b_ret, b_val, idx, rows, obs_floats, h->x0, h->mb_action, h->mb_logp,
        h->mb_adv, h->mb_ret, h->mb_val);
    OCPPO_LAUNCH_CHECK("gather_kernel");
    return OCPPO_OK;
}

int launch_adv_stats(const ocppo_handle *h, const float *b_adv, const int64_t *idx, int64_t mb, cudaStream_t stream) {
    const unsigned grid = grid_for(mb, 256 * 8, h->sm_count * 2);
    ProfScope ps(KC_AUX, 2, stream);
    adv_stats_partial_kernel<<<grid, 256, 0, stream>>>(b_adv, idx, mb, h->adv_part);
    OCPPO_LAUNCH_CHECK("adv_stats_partial_kernel");
    adv_stats_finalize_kernel<<<1, 32, 0, stream>>>(h->adv_part, (int)grid, mb, h->adv_stat);
    OCPPO_LAUNCH_CHECK("adv_stats_finalize_kernel");
    return OCPPO_OK;
}

int launch_explained_var(const ocppo_handle *h, const float *y_pred, const float *y_true, int64_t n, float *out2,
                         cudaStream_t stream) {
    const unsigned grid = grid_for(n, 256 * 8, h->sm_count * 2);       // <= kLossBlocksMax partial slots of 8 doubles
    ProfScope ps(KC_AUX, 2, stream);
    explained_var_partial_kernel<<<grid, 256, 0, stream>>>(y_pred, y_true, n, h->loss_part);
    OCPPO_LAUNCH_CHECK("explained_var_partial_kernel");
    explained_var_finalize_kernel<<<1, 32, 0, stream>>>(h->loss_part, (int)grid, n, out2);
    OCPPO_LAUNCH_CHECK("explained_var_finalize_kernel");
    return OCPPO_OK;
}

static bool loss_head_quad_enabled() {
#ifdef OCPPO_BRINGUP
    static const bool on = getenv("OCPPO_LOSS_HEAD_QUAD") ? atoi(getenv("OCPPO_LOSS_HEAD_QUAD")) != 0 : true;   // bring-up A/B
    return on;
#else
    return true;
#endif
}

int launch_loss_head(const ocppo_handle *h, LossArgs a, int *blocks_used, int blocks_avail, cudaStream_t stream,
                     bool forward_only) {
    const int smem = head_smem_bytes(a.A, a.H);
    const bool aligned = ((((uintptr_t)a.hidden | (uintptr_t)a.dh) & 15) == 0);
    void (*kern)(LossArgs) = forward_only ? loss_head_kernel<0, false> : loss_head_kernel<0, true>;
    if (aligned && a.H == 128) kern = forward_only ? loss_head_kernel<1, false> : loss_head_kernel<1, true>;
    else if (aligned && a.H == 256) kern = forward_only ? loss_head_kernel<2, false> : loss_head_kernel<2, true>;
    else if (aligned && a.H == 512) kern = forward_only ? loss_head_kernel<4, false> : loss_head_kernel<4, true>;
    else if (aligned && a.H == 1024) kern = forward_only ? loss_head_kernel<8, false> : loss_head_kernel<8, true>;
    // four samples per warp pass when the outputs fit an 8-lane group (Pong / SpaceInvaders 6 + 1, Breakout 4 + 1)
    const bool quad = forward_only && aligned && a.A + 1 <= 8 && (a.H == 128 || a.H == 256 || a.H == 512) && loss_head_quad_enabled();
    if (quad) kern = a.H == 128 ? loss_head_quad_kernel<1> : a.H == 256 ? loss_head_quad_kernel<2> : loss_head_quad_kernel<4>;
    OCPPO_TRY(ensure_head_smem((const void *)kern, smem));
    int cap = h->sm_count * (quad ? 2 : forward_only ? 8 : 4);       // quad: the two resident blocks per SM, ~14 passes each
    if (cap > blocks_avail) cap = blocks_avail;
    OCPPO_CHECK_ARG(cap >= 1, "loss head: out of partial-sum slots");
    const unsigned grid = grid_for(quad ? (a.rows + 3) / 4 : a.rows, kWarpsPerBlock, cap);
    ProfScope ps(KC_HEAD, 1, stream);
    kern<<<grid, kHeadThreads, smem, stream>>>(a);
    OCPPO_LAUNCH_CHECK("loss_head_kernel");
    *blocks_used = (int)grid;
    return OCPPO_OK;
}

static bool loss_bwd_reg_enabled() {
#ifdef OCPPO_BRINGUP
    static const bool on = getenv("OCPPO_LOSS_BWD_REG") ? atoi(getenv("OCPPO_LOSS_BWD_REG")) != 0 : true;   // bring-up A/B
    return on;
#else
    return true;
#endif
}

bool loss_head_pair_supported(int H, int A) {
    return (H == 128 || H == 256 || H == 512 || H == 1024) && A >= 1 && A + 1 <= 32;
}

int launch_loss_head_bwd_pair(const ocppo_handle *h, LossBwdPairArgs a, int *colsum_blocks, cudaStream_t stream) {
    OCPPO_CHECK_ARG(loss_head_pair_supported(a.H, a.A), "loss head (pair output): unsupported H=%d A=%d", a.H, a.A);
    // shared memory: the head weights, later the 8 x H column-sum fold
    const int smem = (int)(sizeof(float) * (size_t)a.H * (a.A + 1 > kWarpsPerBlock ? a.A + 1 : kWarpsPerBlock));
    void (*kern)(LossBwdPairArgs) = loss_bwd_pair_kernel<1>;
    if (a.H == 256) kern = loss_bwd_pair_kernel<2>;
    else if (a.H == 512) kern = loss_bwd_pair_kernel<4>;
    else if (a.H == 1024) kern = loss_bwd_pair_kernel<8>;
    // H = 512 with the usual action counts and the ReLU decision words at hand: head weights in registers, one block per SM
    const bool reg = a.H == 512 && a.relu_bits != nullptr && (a.A + 1 == 5 || a.A + 1 == 7) && loss_bwd_reg_enabled();
    if (reg) kern = a.A + 1 == 5 ? loss_bwd_pair_reg_kernel<5> : loss_bwd_pair_reg_kernel<7>;
    OCPPO_TRY(ensure_head_smem((const void *)kern, smem));
    int cap = h->sm_count * (reg ? 1 : 4);
    if (cap > kColsumBlocksMax) cap = kColsumBlocksMax;
    const unsigned grid = grid_for(a.rows, kWarpsPerBlock, cap);
    ProfScope ps(KC_HEAD, 1, stream);
    kern<<<grid, kHeadThreads, smem, stream>>>(a);
    OCPPO_LAUNCH_CHECK("loss_bwd_pair_kernel");
    *colsum_blocks = (int)grid;
    return OCPPO_OK;
}

int launch_loss_finalize(const double *part, int n_part, int64_t mb, float ent_coef, float vf_coef, float *stats,
                         cudaStream_t stream) {
    ProfScope ps(KC_AUX, 1, stream);
    loss_finalize_kernel<<<1, 192, 0, stream>>>(part, n_part, mb, ent_coef, vf_coef, stats);
    OCPPO_LAUNCH_CHECK("loss_finalize_kernel");
    return OCPPO_OK;
}

int launch_splitk_reduce(const float *part, int splits, int64_t stride, int64_t count, float *out, int accumulate,
                         int sm_count, cudaStream_t stream) {
    if (count == 0) return OCPPO_OK;
    ProfScope ps(KC_AUX, 1, stream);
    splitk_reduce_kernel<<<grid_for(count, 256, sm_count * 8), 256, 0, stream>>>(part, splits, stride, count, out,
                                                                               accumulate);
    OCPPO_LAUNCH_CHECK("splitk_reduce_kernel");
    return OCPPO_OK;
}

int launch_relu_mask(const float *g, const float *act, int64_t n, float *out, int sm_count, cudaStream_t stream) {
    if (n == 0) return OCPPO_OK;
    ProfScope ps(KC_AUX, 1, stream);
    relu_mask_kernel<<<grid_for(n, 256, sm_count * 8), 256, 0, stream>>>(g, act, n, out);
    OCPPO_LAUNCH_CHECK("relu_mask_kernel");
    return OCPPO_OK;
}

// torch/optim/adam.py _single_tensor_adam forms its scalars in python double; so does this (host) helper, in the
// order adam_kernel reads them: inv_world, max_norm, 1 - beta1, beta2, 1 - beta2, lr / (1 - beta1^t), sqrt(1 - beta2^t), eps
void adam_scalars(float inv_world, float max_norm, double lr, double beta1, double beta2, double eps, int64_t step, float *out8) {
    const double bc1 = 1.0 - pow(beta1, (double)step);
    const double bc2 = 1.0 - pow(beta2, (double)step);
    out8[0] = inv_world; out8[1] = max_norm;
    out8[2] = (float)(1.0 - beta1); out8[3] = (float)beta2; out8[4] = (float)(1.0 - beta2);
    out8[5] = (float)(lr / bc1); out8[6] = (float)sqrt(bc2); out8[7] = (float)eps;
}

namespace {
struct Words8 { uint32_t w[8]; };
__global__ void store_words_kernel(uint32_t *dst, Words8 v, int n) {
    if (threadIdx.x < n) dst[threadIdx.x] = v.w[threadIdx.x];
}
}  // namespace

// dst[0..n) = words (n <= 8), stream-ordered: the values travel as kernel arguments, so the caller's host memory is
// free again when this returns (an H2D copy from pageable memory would synchronise, one from pinned memory would need
// a ring of slots).  Refreshes the device-resident scalars of graph-replayed calls.
int launch_store_words(void *dst, const void *words, int n, cudaStream_t stream) {
    Words8 v{};
    memcpy(v.w, words, (size_t)n * 4);
    ProfScope ps(KC_AUX, 1, stream);
    store_words_kernel<<<1, 32, 0, stream>>>(reinterpret_cast<uint32_t *>(dst), v, n);
    OCPPO_LAUNCH_CHECK("store_words_kernel");
    return OCPPO_OK;
}

int launch_clip_adam(ocppo_handle *h, const float *host8, const float *dev8, float *grad_norm, cudaStream_t stream) {
    const int64_t n = h->arena_floats;   // multiple of kArenaAlignFloats, padding is zero
    const int blocks = kNormBlocks < h->sm_count * 4 ? kNormBlocks : h->sm_count * 4;
    ProfScope ps(KC_ADAM, 2, stream);
    sumsq_kernel<<<blocks, 256, 0, stream>>>(h->grads, n, host8 ? host8[0] : 1.f, dev8, h->norm_part);
    OCPPO_LAUNCH_CHECK("sumsq_kernel");
    AdamArgs a{};
    a.p = h->params; a.g = h->grads; a.m = h->exp_avg; a.v = h->exp_avg_sq;
    a.n = n; a.norm_part = h->norm_part; a.n_part = blocks;
    if (host8 != nullptr) {
        a.inv_world = host8[0]; a.max_norm = host8[1]; a.one_minus_b1 = host8[2]; a.b2 = host8[3]; a.one_minus_b2 = host8[4];
        a.step_size = host8[5]; a.bc2_sqrt = host8[6]; a.eps = host8[7];
    }
    a.dev_scalars = dev8;
    a.grad_norm = grad_norm;
    adam_kernel<<<h->sm_count * 4, 256, 0, stream>>>(a);
    OCPPO_LAUNCH_CHECK("adam_kernel");
    return OCPPO_OK;
}

}  // namespace ocppo

// ---------------------------------------------------------------------------------------------------------------
// Allocation-free explained variance (ocppo_explained_variance_static, include/ocppo.h).  No counterpart in the
// reference: added for callers that cannot call cudaMalloc (code under CUDA graph
// capture, tight serving loops, harnesses that forbid it).  Same two kernels as ocppo_explained_variance; the per-block
// partial sums live in the file-scope array below instead of a caller-bound workspace, so the call needs no handle, no
// workspace and no allocation of any kind.
//   * 32 KB per GPU (1024 blocks x 4 doubles), held by every GPU that loads the library until the process exits;
//   * in HOST code the array's name is a host-side stand-in, not a device address: the address the kernels get comes
//     from cudaGetSymbolAddress, once per device, with that device current;
//   * every call on a device shares the array: a call leaves its partial sums behind, and the next call overwrites
//     every slot it later reads (block b writes slots 4b..4b+3, the finalize kernel reads exactly the blocks launched),
//     so nothing has to start at zero -- but two calls that OVERLAP on the same GPU (different streams) race.
// ---------------------------------------------------------------------------------------------------------------
#define OCPPO_EV_STATIC_BLOCKS 1024
static __device__ __align__(256) double g_ev_static_scratch[OCPPO_EV_STATIC_BLOCKS * 4];

namespace ocppo {

static int explained_variance_static(const float *values, const float *returns, int64_t n, float *out2, cudaStream_t stream) {
    OCPPO_CHECK_ARG(values != nullptr && returns != nullptr && out2 != nullptr, "explained_variance_static: null pointer");
    OCPPO_CHECK_ARG(n >= 1, "explained_variance_static: n must be >= 1 (got %lld)", (long long)n);
    OCPPO_CHECK_ARG(n <= (int64_t)OCPPO_EV_STATIC_MAX_N,
                    "explained_variance_static: n = %lld exceeds the static-scratch variant's maximum of %lld elements "
                    "(use ocppo_explained_variance with a bound workspace)", (long long)n, (long long)OCPPO_EV_STATIC_MAX_N);
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, values) != cudaSuccess || at.type != cudaMemoryTypeDevice) {
        cudaGetLastError();
        set_error("explained_variance_static: `values` is not device memory");
        return OCPPO_EINVAL;
    }
    const int dev = at.device;
    OCPPO_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "explained_variance_static: device ordinal %d out of range", dev);
    int prev = -1;
    OCPPO_CUDA(cudaGetDevice(&prev));
    if (prev != dev) OCPPO_CUDA(cudaSetDevice(dev));
    struct Restore { int prev, dev; ~Restore() { if (prev != dev) cudaSetDevice(prev); } } restore{prev, dev};
    // set-up, once per device, with that device current: the array's DEVICE address and the SM count
    static double *scratch_of[kMaxDevices] = {nullptr};
    static int sms_of[kMaxDevices] = {0};
    if (scratch_of[dev] == nullptr) {
        double *ptr = nullptr;
        OCPPO_CUDA(cudaGetSymbolAddress((void **)&ptr, g_ev_static_scratch));
        int sms = 0;
        OCPPO_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        sms_of[dev] = sms;
        scratch_of[dev] = ptr;
    }
    double *scratch = scratch_of[dev];
    // the grid of ocppo_explained_variance (so both entry points sum in the same order), capped by the array
    const unsigned grid = grid_for(n, 256 * 8, sms_of[dev] * 2);
    OCPPO_CHECK_ARG(grid <= OCPPO_EV_STATIC_BLOCKS, "explained_variance_static: %u blocks exceed the %d-block static scratch",
                    grid, OCPPO_EV_STATIC_BLOCKS);
    ProfScope ps(KC_AUX, 2, stream);
    explained_var_partial_kernel<<<grid, 256, 0, stream>>>(values, returns, n, scratch);
    OCPPO_LAUNCH_CHECK("explained_var_partial_kernel (static scratch)");
    explained_var_finalize_kernel<<<1, 32, 0, stream>>>(scratch, (int)grid, n, out2);
    OCPPO_LAUNCH_CHECK("explained_var_finalize_kernel (static scratch)");
    return OCPPO_OK;
}

}  // namespace ocppo

extern "C" int ocppo_explained_variance_static(const float *values, const float *returns, int64_t n, float *out2, void *stream) {
    return ocppo::explained_variance_static(values, returns, n, out2, reinterpret_cast<cudaStream_t>(stream));
}
