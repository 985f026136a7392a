// GAE reverse-time scan for (T, N) rollout buffers -- replaces the python loop of
// cleanrl/ppo_atari_oc.py:535-547 (same math ppo_atari_multigpu.py:288-301).
//
// HBM-bound: 20*T*N + 8*N algorithmic bytes (read r, V, done; write A, R).
// Layout: one CTA owns a slab of kEnvs consecutive env columns (one 128 B line
// per row) and walks time backwards.  All loads of a [kRows x kEnvs] tile of the
// three inputs are issued with cp.async (LDGSTS) into a kStages-deep shared-memory
// ring, so the DRAM requests of the whole slab are in flight before the
// sequential recurrence starts; one warp then runs the recurrence out of shared
// memory, one lane per env, and writes A_t / R_t as full 128 B lines.
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

template <int VEC>
__global__ void __launch_bounds__(kThreads)
gae_scan_kernel(const float *__restrict__ rewards, const float *__restrict__ values,
                const float *__restrict__ dones, const float *__restrict__ next_value,
                const float *__restrict__ next_done, int64_t T, int64_t N, float g, float c,
                float *__restrict__ adv, float *__restrict__ ret) {
    extern __shared__ __align__(16) float smem[];   // [kStages][3][kRows][kEnvs]
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
    // slab in flight at once, so the kernel pays one DRAM round trip instead of two.
    const bool all_upfront = n_chunks <= kStages;
#pragma unroll
    for (int s = 0; s < kStages - 1; ++s) issue(s);
    if (all_upfront) issue(kStages - 1);

    const int lane = tid & 31;
    const bool scanner = tid < 32;
    const int64_t env = e0 + lane;
    const bool live = scanner && env < N;
    float last = 0.0f, nv = 0.0f, nnt = 0.0f;
    if (live) {
        nv = next_value[env];
        nnt = __fsub_rn(1.0f, next_done[env]);
    }

    for (int64_t chunk = 0; chunk < n_chunks; ++chunk) {
        if (!all_upfront) issue(chunk + kStages - 1);
        else cp_async_commit();                   // keep the group count per iteration uniform
        cp_async_wait<kStages - 1>();
        __syncthreads();
        if (scanner) {
            const float *stage = smem + (size_t)(chunk % kStages) * 3 * kRows * kEnvs;
            const float *sR = stage, *sV = stage + kRows * kEnvs, *sD = stage + 2 * kRows * kEnvs;
            const int64_t t_hi = T - 1 - chunk * kRows;
            const int rows = (int)(t_hi + 1 < kRows ? t_hi + 1 : kRows);
            if (rows == kRows) {
#pragma unroll 8
                for (int r = 0; r < kRows; ++r) {
                    const float rw = sR[r * kEnvs + lane];
                    const float v = sV[r * kEnvs + lane];
                    const float d = sD[r * kEnvs + lane];
                    const float delta = __fsub_rn(__fadd_rn(rw, __fmul_rn(__fmul_rn(g, nv), nnt)), v);
                    last = __fadd_rn(delta, __fmul_rn(__fmul_rn(c, nnt), last));
                    if (live) {
                        const int64_t o = (t_hi - r) * N + env;
                        adv[o] = last;
                        ret[o] = __fadd_rn(last, v);
                    }
                    nv = v;
                    nnt = __fsub_rn(1.0f, d);
                }
            } else {
                for (int r = 0; r < rows; ++r) {
                    const float rw = sR[r * kEnvs + lane];
                    const float v = sV[r * kEnvs + lane];
                    const float d = sD[r * kEnvs + lane];
                    const float delta = __fsub_rn(__fadd_rn(rw, __fmul_rn(__fmul_rn(g, nv), nnt)), v);
                    last = __fadd_rn(delta, __fmul_rn(__fmul_rn(c, nnt), last));
                    if (live) {
                        const int64_t o = (t_hi - r) * N + env;
                        adv[o] = last;
                        ret[o] = __fadd_rn(last, v);
                    }
                    nv = v;
                    nnt = __fsub_rn(1.0f, d);
                }
            }
        }
        __syncthreads();
    }
    cp_async_wait<0>();
}

}  // namespace

int launch_gae(const float *rewards, const float *values, const float *dones,
               const float *next_value, const float *next_done, int64_t T, int64_t N, float g,
               float c, float *adv, float *ret, cudaStream_t stream) {
    if (T == 0 || N == 0) return OCPPO_OK;
    const size_t smem = (size_t)kStages * 3 * kRows * kEnvs * sizeof(float);
    const int64_t grid = ceil_div(N, kEnvs);
    OCPPO_CHECK_ARG(grid <= 0x7fffffff, "gae: num_envs too large");
    const bool vec = (N % 4 == 0) && ((((uintptr_t)rewards | (uintptr_t)values | (uintptr_t)dones) & 15) == 0);
    static bool attr_done = false;
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
