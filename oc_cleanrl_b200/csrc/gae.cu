// GAE reverse-time scan for (T, N) rollout buffers -- replaces the python loop of
// cleanrl/ppo_atari_oc.py:535-547 (same math ppo_atari_multigpu.py:288-301).
//
// HBM-bound: 20*T*N + 8*N algorithmic bytes (read r, V, done; write A, R).
// Layout: one CTA owns a slab of kEnvs consecutive env columns (one 128 B line
// per row) and walks time backwards.  All loads of a [kRows x kEnvs] tile of the
// three inputs are issued with cp.async (LDGSTS) into a kStages-deep shared-memory
// ring, so the DRAM requests of the whole slab are in flight before any
// arithmetic starts.  Per tile: (A) all warps turn r/V/done into delta_t and the
// coefficient c*nnt_{t+1} in place, (B) one warp runs the 2-flop recurrence, one
// lane per env, (C) all warps add V and write A_t / R_t as full 128 B lines.  Only
// (B) is serial, and it is kept to 5 instructions per step: with several CTAs per
// SM the scan warps share issue slots, and a fat scan loop was the bottleneck.
//
// Arithmetic follows torch's fp32 evaluation order with no FMA contraction
// (SURVEY.md appendix A.1), so results are bit-identical to the reference loop:
//   nnt = 1 - done_{t+1};  delta = (r_t + (g*V_{t+1})*nnt) - V_t;
//   A_t = delta + ((c*nnt)*A_{t+1});  R_t = A_t + V_t,   g=f32(gamma), c=f32(gamma*lambda)
//
// Two kernels.  gae_tma_kernel (below, the default: N % 4 == 0, 16-byte aligned buffers) moves every tile with TMA -- one
// bulk tensor load per (array, 32-step tile) and one bulk tensor store per output tile, issued by a single thread -- and
// splits the CTA into four elementwise warps and one scan warp that work on DIFFERENT tiles at the same time:
//   elementwise warps:  A(c): delta_t in place          C(c-1): R_t = A_t + V_t, stores       (A runs one tile ahead)
//   scan warp        :  B(c): A_t = delta_t + (c*nnt_{t+1}) * A_{t+1}, one lane per env, as soon as A(c) is done
// so the only serial part, the 2-flop recurrence (8 dependent cycles per time step), is never waited for with idle
// warps, and the per-thread address arithmetic of cp.async loads / STG stores (most of the first kernel's instructions:
// ncu showed it issue-bound at 25 % occupancy, not DRAM-bound, at 16,384 envs) is gone.  gae_scan_kernel is the first
// version (cp.async ring, block-wide phases); it stays for ragged / unaligned buffers.
#include "common.cuh"
#include <cuda.h>

namespace ocppo {

namespace {

constexpr int kEnvs = 32;      // env columns per CTA (one warp scans them)
constexpr int kRows = 32;      // time steps per pipeline stage
constexpr int kStages = 4;     // ring depth: 4 x 3 x 32 x 32 x 4 B = 48 KB
constexpr int kThreads = 128;

template <int BYTES>
__device__ __forceinline__ void cp_async(void *smem, const void *gmem) {
    unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    if (BYTES == 16)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
    else
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}

__device__ __forceinline__ float4 ld4(const float *p) { return *reinterpret_cast<const float4 *>(p); }
__device__ __forceinline__ void st4(float *p, float4 v) { *reinterpret_cast<float4 *>(p) = v; }
// nnt = 1 - done
__device__ __forceinline__ float4 nnt4(float4 d) {
    return make_float4(__fsub_rn(1.0f, d.x), __fsub_rn(1.0f, d.y), __fsub_rn(1.0f, d.z), __fsub_rn(1.0f, d.w));
}
// delta = (r + (g*V')*nnt') - V, in torch's evaluation order
__device__ __forceinline__ float delta1(float rw, float v, float pv, float pn, float g) {
    return __fsub_rn(__fadd_rn(rw, __fmul_rn(__fmul_rn(g, pv), pn)), v);
}

template <int VEC>
__global__ void __launch_bounds__(kThreads)
gae_scan_kernel(const float *__restrict__ rewards, const float *__restrict__ values,
                const float *__restrict__ dones, const float *__restrict__ next_value,
                const float *__restrict__ next_done, int64_t T, int64_t N, float g, float c,
                float *__restrict__ adv, float *__restrict__ ret) {
    extern __shared__ __align__(16) float smem[];   // [kStages][3][kRows][kEnvs], then bV[kEnvs], bN[kEnvs]
    float *bV = smem + kStages * 3 * kRows * kEnvs;  // V and nnt of the time step after the current chunk
    float *bN = bV + kEnvs;
    const int tid = threadIdx.x;
    const int64_t e0 = (int64_t)blockIdx.x * kEnvs;
    const int64_t n_chunks = (T + kRows - 1) / kRows;
    const float *src[3] = {rewards, values, dones};

    auto issue = [&](int64_t chunk) {
        if (chunk < n_chunks) {
            float *stage = smem + (size_t)(chunk % kStages) * 3 * kRows * kEnvs;
            const int64_t t_hi = T - 1 - chunk * kRows;
            constexpr int kVecPerRow = kEnvs / VEC;
            constexpr int kPerArr = kRows * kVecPerRow;
#pragma unroll
            for (int q = tid; q < 3 * kPerArr; q += kThreads) {
                const int arr = q / kPerArr;
                const int rem = q - arr * kPerArr;
                const int r = rem / kVecPerRow;
                const int col = (rem - r * kVecPerRow) * VEC;
                const int64_t t = t_hi - r;
                const int64_t env = e0 + col;
                if (t >= 0 && env < N)
                    cp_async<VEC * 4>(stage + (arr * kRows + r) * kEnvs + col,
                                      src[arr] + t * N + env);
            }
        }
        cp_async_commit();
    };

    // A rollout of T <= kStages * kRows steps (the script's default 128) fits the ring entirely: put the whole
    // slab in flight at once, so the kernel pays one DRAM round trip and never refills a stage.
    const bool all_upfront = n_chunks <= kStages;
#pragma unroll
    for (int s = 0; s < kStages - 1; ++s) issue(s);
    if (all_upfront) issue(kStages - 1);

    if (tid < kEnvs) {
        const int64_t env = e0 + tid;
        bV[tid] = env < N ? next_value[env] : 0.0f;
        bN[tid] = env < N ? __fsub_rn(1.0f, next_done[env]) : 0.0f;
    }
    // The recurrence is the only serial part: one warp per CTA runs it, 2 flops per step.  Which warp is rotated
    // with the slab index so that co-resident CTAs do not queue their scans on one SM sub-partition.
    const int lane = tid & 31;
    const bool scanner = (tid >> 5) == (int)(blockIdx.x & 3);
    float last = 0.0f;
    // elementwise phases: thread -> (row, 4 envs); a chunk is 2 such items per thread
    const int q4 = (tid & 7) * 4;
    const int r_lo = tid >> 3;                       // rows r_lo and r_lo + 16

    for (int64_t chunk = 0; chunk < n_chunks; ++chunk) {
        if (!all_upfront) issue(chunk + kStages - 1);
        else cp_async_commit();                      // keep the group count per iteration uniform
        cp_async_wait<kStages - 1>();
        __syncthreads();
        float *stage = smem + (size_t)(chunk % kStages) * 3 * kRows * kEnvs;
        float *sR = stage, *sV = stage + kRows * kEnvs, *sD = stage + 2 * kRows * kEnvs;
        const int64_t t_hi = T - 1 - chunk * kRows;

        // phase A (all warps): delta_t and the recurrence coefficient c*nnt_{t+1}, read into registers first because
        // row r needs V and done of row r-1 (time t+1) and the results overwrite r_t / done_t in place.
        float4 dl[2], cf[2], vkeep = make_float4(0.f, 0.f, 0.f, 0.f), nkeep = vkeep;
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int r = r_lo + 16 * i;
            const float4 rw = ld4(sR + r * kEnvs + q4), v = ld4(sV + r * kEnvs + q4);
            float4 pv, pn;
            if (r == 0) {
                pv = ld4(bV + q4);
                pn = ld4(bN + q4);
            } else {
                pv = ld4(sV + (r - 1) * kEnvs + q4);
                pn = nnt4(ld4(sD + (r - 1) * kEnvs + q4));
            }
            dl[i] = make_float4(delta1(rw.x, v.x, pv.x, pn.x, g), delta1(rw.y, v.y, pv.y, pn.y, g),
                                delta1(rw.z, v.z, pv.z, pn.z, g), delta1(rw.w, v.w, pv.w, pn.w, g));
            cf[i] = make_float4(__fmul_rn(c, pn.x), __fmul_rn(c, pn.y), __fmul_rn(c, pn.z), __fmul_rn(c, pn.w));
            if (r == kRows - 1) {                    // becomes the boundary of the next chunk
                vkeep = v;
                nkeep = nnt4(ld4(sD + r * kEnvs + q4));
            }
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int r = r_lo + 16 * i;
            st4(sR + r * kEnvs + q4, dl[i]);
            st4(sD + r * kEnvs + q4, cf[i]);
            if (r == kRows - 1) {
                st4(bV + q4, vkeep);
                st4(bN + q4, nkeep);
            }
        }
        __syncthreads();

        // phase B (one warp, one lane per env): A_t = delta_t + (c*nnt_{t+1}) * A_{t+1}.  Rows past t = 0 of a
        // ragged last chunk hold garbage, come last in scan order and are never stored.
        if (scanner) {
#pragma unroll
            for (int r = 0; r < kRows; ++r) {
                last = __fadd_rn(sR[r * kEnvs + lane], __fmul_rn(sD[r * kEnvs + lane], last));
                sR[r * kEnvs + lane] = last;
            }
        }
        __syncthreads();

        // phase C (all warps): R_t = A_t + V_t, both written as full 128 B lines
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int r = r_lo + 16 * i;
            const int64_t t = t_hi - r;
            const int64_t env = e0 + q4;
            if (t >= 0 && env < N) {
                const float4 a = ld4(sR + r * kEnvs + q4), v = ld4(sV + r * kEnvs + q4);
                const float4 rt = make_float4(__fadd_rn(a.x, v.x), __fadd_rn(a.y, v.y), __fadd_rn(a.z, v.z),
                                              __fadd_rn(a.w, v.w));
                const int64_t o = t * N + env;
                if (VEC == 4) {
                    __stcs(reinterpret_cast<float4 *>(adv + o), a);
                    __stcs(reinterpret_cast<float4 *>(ret + o), rt);
                } else {
                    const float av[4] = {a.x, a.y, a.z, a.w}, rv[4] = {rt.x, rt.y, rt.z, rt.w};
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (env + j < N) {
                            adv[o + j] = av[j];
                            ret[o + j] = rv[j];
                        }
                }
            }
        }
        if (!all_upfront) __syncthreads();           // the next iteration's issue refills this stage
    }
    cp_async_wait<0>();
}


// ---------------------------------------------------------------------------------------------------------------
// TMA version
// ---------------------------------------------------------------------------------------------------------------
constexpr int kEwThreads = 128;                  // elementwise warps 0-3
constexpr int kTmaThreads = kEwThreads + 32;     // + the scan warp
constexpr int kTileFloats = kRows * kEnvs;       // 4 KB per array and tile
constexpr int kStageFloats = 3 * kTileFloats;    // r (-> delta -> A), V (-> R), done
constexpr int kTmaSmemBytes = kStages * kStageFloats * 4 + 4 * kEnvs * 4 + 3 * kStages * 8 + 128;

__device__ __forceinline__ uint32_t s_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mb_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mb_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mb_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mb_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    const long long t0 = clock64();
    while (true) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (done) break;
        if (clock64() - t0 > 2000000000LL) __trap();      // never hang the device on a protocol bug
    }
}
__device__ __forceinline__ void tma_ld(uint32_t dst, const CUtensorMap *tm, int col, int row, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tm)), "r"(col), "r"(row), "r"(bar)
                 : "memory");
}
__device__ __forceinline__ void tma_st(const CUtensorMap *tm, uint32_t src, int col, int row) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(tm)), "r"(src), "r"(col), "r"(row)
                 : "memory");
}
__device__ __forceinline__ void ew_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kEwThreads) : "memory"); }

__global__ void __launch_bounds__(kTmaThreads, 4)
gae_tma_kernel(const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmV,
               const __grid_constant__ CUtensorMap tmD, const __grid_constant__ CUtensorMap tmA,
               const __grid_constant__ CUtensorMap tmRet, const float *__restrict__ next_value,
               const float *__restrict__ next_done, int T, int N, float g, float c) {
    extern __shared__ __align__(128) float sm[];
    float *bV = sm + kStages * kStageFloats;            // [2][kEnvs]  V of the time step after the tile (per tile parity)
    float *bN = bV + 2 * kEnvs;                         // [2][kEnvs]  1 - done of that step
    uint64_t *bars = reinterpret_cast<uint64_t *>(bN + 2 * kEnvs);
    const uint32_t bar_load = s_u32(bars), bar_adone = s_u32(bars + kStages), bar_sdone = s_u32(bars + 2 * kStages);
    const int tid = threadIdx.x;
    const int e0 = (int)blockIdx.x * kEnvs;
    const int n_chunks = (T + kRows - 1) / kRows;
    // Tiles are aligned to multiples of 32 time steps from t = 0 and walked from the last one down: chunk ch holds
    // t0 .. t0 + 31 with t0 = 32 (n_chunks - 1 - ch), smem row r <-> t0 + r.  Only the FIRST chunk can be ragged (rows
    // past T - 1: zero-filled by the load, forced to delta = 0 below, clipped by the store) -- TMA coordinates stay
    // non-negative (tiles hanging over the t < 0 side faulted on the B200).
    const int r_top = (T - 1) - kRows * (n_chunks - 1);  // row of time step T - 1 in chunk 0
    auto issue_loads = [&](int chunk) {                  // thread 0 only
        const int s = chunk % kStages;
        const uint32_t st = s_u32(sm + s * kStageFloats), bar = bar_load + 8 * s;
        const int t0 = kRows * (n_chunks - 1 - chunk);
        mb_expect_tx(bar, kStageFloats * 4);
        tma_ld(st, &tmR, e0, t0, bar);
        tma_ld(st + kTileFloats * 4, &tmV, e0, t0, bar);
        tma_ld(st + 2 * kTileFloats * 4, &tmD, e0, t0, bar);
    };
    if (tid == 0) {
        for (int s = 0; s < kStages; ++s) {
            mb_init(bar_load + 8 * s, 1);
            mb_init(bar_adone + 8 * s, 1);
            mb_init(bar_sdone + 8 * s, 1);
        }
        // (no cluster: the generic-proxy init only has to be visible to this CTA's async proxy -- the cluster-scope
        // fence.mbarrier_init compiles to MEMBAR.ALL.GPU, a microsecond-class cost in a 4 us kernel)
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int ch = 0; ch < kStages && ch < n_chunks; ++ch) issue_loads(ch);
    }
    // Programmatic dependent launch (no-ops in a plain launch): the NEXT kernel on the stream may start its prologue now,
    // and when THIS kernel was launched with the overlap flag it has been running since the previous kernel allowed it --
    // its tile loads above are already in flight.  Only next_value (the previous kernel's output, the bootstrap of
    // ppo_atari_oc.py:534) has to wait for that kernel to have completed and flushed.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (tid < kEnvs) {                                   // boundary of tile 0: the bootstrap value and next_done
        const int env = e0 + tid;
        bV[tid] = env < N ? next_value[env] : 0.0f;
        bN[tid] = env < N ? __fsub_rn(1.0f, next_done[env]) : 0.0f;
    }
    __syncthreads();                                     // barriers initialised, boundary visible

    if (tid >= kEwThreads) {
        // ===================== scan warp =====================
        const int lane = tid - kEwThreads;
        float last = 0.0f;
        float nnt = bN[lane];                            // 1 - done of the step after the current one
        for (int ch = 0; ch < n_chunks; ++ch) {
            const int s = ch % kStages;
            float *sR = sm + s * kStageFloats, *sD = sR + 2 * kTileFloats;
            mb_wait(bar_adone + 8 * s, (uint32_t)((ch / kStages) & 1));
            // the whole column into registers first (64 independent shared-memory loads in flight), the coefficients
            // c * nnt_{t+1} off the dependent chain, then the chain itself: one multiply and one add per time step
            float d[kRows], q[kRows];
#pragma unroll
            for (int r = 0; r < kRows; ++r) {
                d[r] = sR[r * kEnvs + lane];
                q[r] = sD[r * kEnvs + lane];
            }
            const float nnt_next = __fsub_rn(1.0f, q[0]);
#pragma unroll
            for (int r = 0; r < kRows - 1; ++r) q[r] = __fmul_rn(c, __fsub_rn(1.0f, q[r + 1]));     // coefficient of row r
            q[kRows - 1] = __fmul_rn(c, nnt);
            nnt = nnt_next;
#pragma unroll
            for (int r = kRows - 1; r >= 0; --r) {
                last = __fadd_rn(d[r], __fmul_rn(q[r], last));
                sR[r * kEnvs + lane] = last;
            }
            __syncwarp();
            if (lane == 0) mb_arrive(bar_sdone + 8 * s);
        }
    } else {
        // ===================== elementwise warps =====================
        const int q4 = (tid & 7) * 4;                    // thread -> (row, 4 envs); two rows per tile and thread
        const int r_lo = tid >> 3;                       // rows r_lo and r_lo + 16
        auto phase_a = [&](int ch) {
            const int s = ch % kStages;
            float *sR = sm + s * kStageFloats, *sV = sR + kTileFloats, *sD = sR + 2 * kTileFloats;
            mb_wait(bar_load + 8 * s, (uint32_t)((ch / kStages) & 1));
            const float *cbV = bV + (ch & 1) * kEnvs, *cbN = bN + (ch & 1) * kEnvs;
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const int r = r_lo + 16 * i;
                const float4 rw = ld4(sR + r * kEnvs + q4), v = ld4(sV + r * kEnvs + q4);
                float4 pv, pn;
                const int r_bnd = ch == 0 ? r_top : kRows - 1;      // the row whose successor in time lives outside the tile
                if (r == r_bnd) {
                    pv = ld4(cbV + q4);
                    pn = ld4(cbN + q4);
                } else {
                    pv = ld4(sV + (r + 1) * kEnvs + q4);
                    pn = nnt4(ld4(sD + (r + 1) * kEnvs + q4));
                }
                float4 dl = make_float4(delta1(rw.x, v.x, pv.x, pn.x, g), delta1(rw.y, v.y, pv.y, pn.y, g),
                                        delta1(rw.z, v.z, pv.z, pn.z, g), delta1(rw.w, v.w, pv.w, pn.w, g));
                if (r > r_bnd) dl = make_float4(0.f, 0.f, 0.f, 0.f);   // rows past T - 1: the scan starts from A_T = 0
                st4(sR + r * kEnvs + q4, dl);
                if (r == 0) {                            // becomes the boundary of the next tile
                    st4(bV + ((ch + 1) & 1) * kEnvs + q4, v);
                    st4(bN + ((ch + 1) & 1) * kEnvs + q4, nnt4(ld4(sD + q4)));
                }
            }
            ew_sync();                                   // every delta of the tile written, boundary visible
            if (tid == 0) mb_arrive(bar_adone + 8 * s);
        };
        auto phase_c = [&](int ch) {
            const int s = ch % kStages;
            float *sR = sm + s * kStageFloats, *sV = sR + kTileFloats;
            mb_wait(bar_sdone + 8 * s, (uint32_t)((ch / kStages) & 1));
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const int r = r_lo + 16 * i;
                const float4 a = ld4(sR + r * kEnvs + q4), v = ld4(sV + r * kEnvs + q4);
                st4(sV + r * kEnvs + q4, make_float4(__fadd_rn(a.x, v.x), __fadd_rn(a.y, v.y), __fadd_rn(a.z, v.z),
                                                     __fadd_rn(a.w, v.w)));
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            ew_sync();
            if (tid == 0) {
                const int t0 = kRows * (n_chunks - 1 - ch);
                tma_st(&tmA, s_u32(sR), e0, t0);
                tma_st(&tmRet, s_u32(sV), e0, t0);
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                if (ch + kStages < n_chunks) {           // T > 128: refill the stage once the stores have read it
                    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                    issue_loads(ch + kStages);
                }
            }
        };
        for (int ch = 0; ch < n_chunks; ++ch) {
            phase_a(ch);
            if (ch >= 1) phase_c(ch - 1);
        }
        phase_c(n_chunks - 1);
        // the stores only have to have READ shared memory before the CTA exits; the kernel boundary completes them
        if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
}
}  // namespace

int g_gae_variant = 0;      // 0: TMA kernel where eligible (default); 1: always the cp.async kernel (ocppo_set_gae_variant)

int launch_gae(const float *rewards, const float *values, const float *dones,
               const float *next_value, const float *next_done, int64_t T, int64_t N, float g,
               float c, float *adv, float *ret, int overlap_previous, cudaStream_t stream) {
    if (T == 0 || N == 0) return OCPPO_OK;
    const size_t smem = ((size_t)kStages * 3 * kRows * kEnvs + 2 * kEnvs) * sizeof(float);
    const int64_t grid = ceil_div(N, kEnvs);
    OCPPO_CHECK_ARG(grid <= 0x7fffffff, "gae: num_envs too large");
    const bool vec = (N % 4 == 0) && ((((uintptr_t)rewards | (uintptr_t)values | (uintptr_t)dones | (uintptr_t)adv |
                                       (uintptr_t)ret) & 15) == 0);
    static bool attr_done_of[kMaxDevices] = {false};     // a function attribute lives in the device's context
    int dev = 0;
    OCPPO_CUDA(cudaGetDevice(&dev));
    OCPPO_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "gae: device ordinal %d out of range", dev);
    bool &attr_done = attr_done_of[dev];
    if (!attr_done) {
        OCPPO_CUDA(cudaFuncSetAttribute(gae_scan_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        OCPPO_CUDA(cudaFuncSetAttribute(gae_scan_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        OCPPO_CUDA(cudaFuncSetAttribute(gae_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kTmaSmemBytes));
        attr_done = true;
    }
    if (vec && g_gae_variant != 1 && T < (1 << 30) && N < (1 << 30)) {
        alignas(64) CUtensorMap tm[5];
        const float *bases[5] = {rewards, values, dones, adv, ret};
        for (int i = 0; i < 5; ++i) OCPPO_TRY(tma_map_2d_f32_plain(&tm[i], bases[i], T, N, kRows, kEnvs));
        ProfScope ps(KC_GAE, 1, stream);
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3(kTmaThreads); cfg.dynamicSmemBytes = kTmaSmemBytes; cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        // (measured: with fewer CTAs than SMs the scan is pure latency and gains nothing from starting early -- 6.0 vs
        // 5.4 us per launch at 4096 envs -- while 16,384 envs go from 10.9 to 8.9 us and 65,536 from 32.6 to 29.5 us)
        int sms = 0;
        OCPPO_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        cfg.attrs = attr; cfg.numAttrs = (overlap_previous && grid > sms) ? 1 : 0;
        const int Ti = (int)T, Ni = (int)N;
        OCPPO_CUDA(cudaLaunchKernelEx(&cfg, gae_tma_kernel, tm[0], tm[1], tm[2], tm[3], tm[4], next_value, next_done, Ti, Ni, g, c));
        return OCPPO_OK;
    }
    ProfScope ps(KC_GAE, 1, stream);
    if (vec)
        gae_scan_kernel<4><<<(unsigned)grid, kThreads, smem, stream>>>(rewards, values, dones, next_value,
                                                                      next_done, T, N, g, c, adv, ret);
    else
        gae_scan_kernel<1><<<(unsigned)grid, kThreads, smem, stream>>>(rewards, values, dones, next_value,
                                                                      next_done, T, N, g, c, adv, ret);
    OCPPO_LAUNCH_CHECK("gae_scan_kernel");
    return OCPPO_OK;
}

}  // namespace ocppo
