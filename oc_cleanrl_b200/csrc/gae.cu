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
#include "common.cuh"

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

}  // namespace

int launch_gae(const float *rewards, const float *values, const float *dones,
               const float *next_value, const float *next_done, int64_t T, int64_t N, float g,
               float c, float *adv, float *ret, cudaStream_t stream) {
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
        attr_done = true;
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
