// Hardware probe for the tcgen05 kind::tf32 design decisions of tc_gemm.cu (bring-up tool, not
// part of libocppo.so).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/tc_probe tools/tc_probe.cu
//   Q1  does kind::tf32 truncate or round the low 13 mantissa bits of its fp32 operands?
//   Q2  cycles per 128x256x8 tf32 MMA when nothing else touches shared memory
//   Q3  the same while the other warps stream STS.128 / LDS.128 at the producer's rate
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    const long long t0 = clock64();
    uint32_t done = 0;
    while (true) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (done) break;
        if (clock64() - t0 > 2000000000LL) __trap();
    }
}
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t type) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)type << 61;
    return d;
}
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, bool a_mn, bool b_mn) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_tf32(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

constexpr int TM = 128, TN = 256, TK = 32;
constexpr int kABytes = TM * TK * 4, kBBytes = TN * TK * 4;

// ---------------------------------------------------------------- Q1 ----------
__global__ void __launch_bounds__(128, 1) q1_kernel(const float *A, const float *B, float *D) {
    extern __shared__ uint8_t raw[];
    uint8_t *smem = (uint8_t *)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    uint8_t *sa = smem, *sb = smem + kABytes;
    uint64_t *bar = (uint64_t *)(sb + kBBytes);
    uint32_t *slot = (uint32_t *)(bar + 1);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < TM * 8; i += 128) {
        const int r = i >> 3, c = i & 7;
        *(float4 *)(sa + (r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4)) = *(const float4 *)(A + r * TK + c * 4);
    }
    for (int i = tid; i < TN * 8; i += 128) {
        const int r = i >> 3, c = i & 7;
        *(float4 *)(sb + (r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4)) = *(const float4 *)(B + r * TK + c * 4);
    }
    if (tid == 0) { mbar_init(smem_u32(bar), 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(256u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = *slot;
    if (tid == 0) {
        const uint64_t da = make_desc(smem_u32(sa), 16, 1024, 2), db = make_desc(smem_u32(sb), 16, 1024, 2);
        for (int ks = 0; ks < 4; ++ks) umma_tf32(tm, da + ks * 2, db + ks * 2, make_idesc(TM, TN, false, false), ks > 0);
        umma_commit(smem_u32(bar));
    }
    mbar_wait(smem_u32(bar), 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    for (int cc = 0; cc < 8; ++cc) {
        uint32_t v[32];
        const uint32_t taddr = tm + ((uint32_t)(warp * 32) << 16) + cc * 32;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                     "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                     "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                       "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                       "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                       "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                     : "r"(taddr) : "memory");
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int j = 0; j < 32; ++j) D[(warp * 32 + lane) * TN + cc * 32 + j] = __uint_as_float(v[j]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(256u) : "memory");
}

// ------------------------------------------------------------ Q2 / Q3 ----------
// mode bit0: issue MMAs; bit1: other warps STS.128 at the producer's rate (28 per thread per stage);
// bit2: other warps LDS.128 (14 per thread per stage).  Two 96 KB operand stages + 16 KB scratch.
constexpr int kStage = 2 * (kABytes + kBBytes);
__global__ void __launch_bounds__(256, 1) q23_kernel(int mode, int niter, int passes, long long *out) {
    extern __shared__ uint8_t raw[];
    uint8_t *smem = (uint8_t *)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    uint8_t *scratch = smem + 2 * kStage;
    uint64_t *bars = (uint64_t *)(scratch + 16384);
    uint32_t *slot = (uint32_t *)(bars + 2);
    volatile int *issued = (volatile int *)(slot + 1);
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (2 * kStage + 16384) / 16; i += 256) ((float4 *)smem)[i] = make_float4(1.f, 0.5f, 0.25f, 2.f);
    if (tid == 0) {
        mbar_init(smem_u32(&bars[0]), 1); mbar_init(smem_u32(&bars[1]), 1);
        *issued = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(256u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tm = *slot;
    long long t_mma = 0, t_traffic = 0;
    if (tid == 0) {
        const long long t0 = clock64();
        const uint32_t idesc = make_idesc(TM, TN, false, false);
        for (int it = 0; it < niter; ++it) {
            const int s = it & 1;
            if ((mode & 1) && it >= 2) mbar_wait(smem_u32(&bars[s]), (uint32_t)(((it >> 1) - 1) & 1));
            if (mode & 1) {
                uint8_t *st = smem + s * kStage;
                const uint64_t da = make_desc(smem_u32(st), 16, 1024, 2), da2 = make_desc(smem_u32(st + kABytes), 16, 1024, 2);
                const uint64_t db = make_desc(smem_u32(st + 2 * kABytes), 16, 1024, 2);
                const uint64_t db2 = make_desc(smem_u32(st + 2 * kABytes + kBBytes), 16, 1024, 2);
                for (int ks = 0; ks < 4; ++ks) {
                    umma_tf32(tm, da2 + ks * 2, db + ks * 2, idesc, (it | ks) ? 1u : 0u);
                    if (passes >= 2) umma_tf32(tm, da + ks * 2, db2 + ks * 2, idesc, 1u);
                    if (passes >= 3) umma_tf32(tm, da + ks * 2, db + ks * 2, idesc, 1u);
                }
                umma_commit(smem_u32(&bars[s]));
            }
            *issued = it + 1;
        }
        if (mode & 1) {
            for (int it = niter - 2 > 0 ? niter - 2 : 0; it < niter; ++it) mbar_wait(smem_u32(&bars[it & 1]), (uint32_t)((it >> 1) & 1));
        }
        t_mma = clock64() - t0;
    } else if (warp >= 1 && (mode & 6)) {
        const int t = tid - 32;
        const long long t0 = clock64();
        float4 acc = make_float4(0, 0, 0, 0);
        for (int it = 0; it < niter; ++it) {
            if (mode & 1) { while (*issued <= it) { } }
            if (mode & 2) {
#pragma unroll
                for (int j = 0; j < 28; ++j) *(float4 *)(scratch + (((j * 224 + t) * 16) & 16383)) = make_float4((float)it, (float)j, 1.f, 2.f);
            }
            if (mode & 4) {
#pragma unroll
                for (int j = 0; j < 14; ++j) {
                    float4 v;
                    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                                 : "r"(smem_u32(scratch + (((j * 224 + t) * 16) & 16383))));
                    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
                }
            }
        }
        t_traffic = clock64() - t0;
        if (acc.x == 123.456f) out[7] = 1;
    }
    if (blockIdx.x == 0) {
        if (tid == 0) out[0] = t_mma;
        if (tid == 32) out[1] = t_traffic;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tm), "r"(256u) : "memory");
}

static float trunc_tf32(float x) { uint32_t u; memcpy(&u, &x, 4); u &= 0xFFFFE000u; memcpy(&x, &u, 4); return x; }
static float rna_tf32(float x) { uint32_t u; memcpy(&u, &x, 4); u += 0x1000u; u &= 0xFFFFE000u; memcpy(&x, &u, 4); return x; }

int main() {
    // ---- Q1
    std::vector<float> A(TM * TK), B(TN * TK), D(TM * TN);
    srand(1);
    for (auto &x : A) x = (float)rand() / RAND_MAX * 2.f - 1.f;
    for (auto &x : B) x = (float)rand() / RAND_MAX * 2.f - 1.f;
    float *dA, *dB, *dD;
    CK(cudaMalloc(&dA, A.size() * 4)); CK(cudaMalloc(&dB, B.size() * 4)); CK(cudaMalloc(&dD, D.size() * 4));
    CK(cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice));
    const int smem1 = kABytes + kBBytes + 1024 + 64;
    CK(cudaFuncSetAttribute(q1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem1));
    q1_kernel<<<1, 128, smem1>>>(dA, dB, dD);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
    double e_tr = 0, e_rn = 0, e_full = 0;
    for (int m = 0; m < TM; ++m)
        for (int n = 0; n < TN; ++n) {
            double st = 0, sr = 0, sf = 0;
            for (int k = 0; k < TK; ++k) {
                st += (double)trunc_tf32(A[m * TK + k]) * trunc_tf32(B[n * TK + k]);
                sr += (double)rna_tf32(A[m * TK + k]) * rna_tf32(B[n * TK + k]);
                sf += (double)A[m * TK + k] * B[n * TK + k];
            }
            const double d = D[m * TN + n];
            e_tr = fmax(e_tr, fabs(d - st)); e_rn = fmax(e_rn, fabs(d - sr)); e_full = fmax(e_full, fabs(d - sf));
        }
    printf("Q1 raw fp32 operands into kind::tf32: max|D - trunc model| = %.3e, max|D - rna model| = %.3e, max|D - fp32 exact| = %.3e\n",
           e_tr, e_rn, e_full);
    // ---- Q4: how does the TMEM accumulator round when a small product is added to a large one?
    {
        std::vector<float> A2(TM * TK, 0.f), B2(TN * TK, 0.f), D2(TM * TN);
        for (int m = 0; m < TM; ++m) { A2[m * TK + 0] = 4096.f; A2[m * TK + 8] = 1.5f; }
        const float big[4] = {4096.f, -4096.f, -4096.f, 4096.f}, small[4] = {1.f, -1.f, 1.f, -1.f};
        for (int n = 0; n < 4; ++n) { B2[n * TK + 0] = big[n]; B2[n * TK + 8] = small[n]; }
        CK(cudaMemcpy(dA, A2.data(), A2.size() * 4, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dB, B2.data(), B2.size() * 4, cudaMemcpyHostToDevice));
        q1_kernel<<<1, 128, smem1>>>(dA, dB, dD);
        CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(D2.data(), dD, D2.size() * 4, cudaMemcpyDeviceToHost));
        printf("Q4 accumulator rounding (exact / got): %.1f / %.1f   %.1f / %.1f   %.1f / %.1f   %.1f / %.1f\n",
               16777217.5, D2[0], -16777217.5, D2[1], -16777214.5, D2[2], 16777214.5, D2[3]);
        printf("   (RN: 16777218 -16777218 -16777214 16777214; toward zero: 16777216 -16777216 -16777214 16777214; floor: 16777216 -16777218 -16777215 16777214)\n");
    }
    // ---- Q2/Q3
    long long *dout, hout[8];
    CK(cudaMalloc(&dout, 64));
    const int smem2 = 2 * kStage + 16384 + 1024 + 64;
    CK(cudaFuncSetAttribute(q23_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem2));
    int dev_sms = 0;
    CK(cudaDeviceGetAttribute(&dev_sms, cudaDevAttrMultiProcessorCount, 0));
    const int niter = 2000;
    const int cfgs[][2] = {{1, 3}, {7, 3}};
    for (auto &c : cfgs) {
        for (int grid : {1, dev_sms}) {
            CK(cudaMemset(dout, 0, 64));
            cudaEvent_t e0, e1;
            CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
            CK(cudaEventRecord(e0));
            q23_kernel<<<grid, 256, smem2>>>(c[0], niter, c[1], dout);
            CK(cudaEventRecord(e1));
            CK(cudaDeviceSynchronize());
            float ms = 0;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            CK(cudaMemcpy(hout, dout, 64, cudaMemcpyDeviceToHost));
            const double mmas = (double)niter * 4 * c[1];
            printf("Q23 mode=%d (mma=%d sts=%d lds=%d) passes=%d grid=%3d: mma-thread %.1f cyc/stage (%.1f cyc/MMA), traffic %.1f cyc/stage, kernel %.3f ms",
                   c[0], c[0] & 1, (c[0] >> 1) & 1, (c[0] >> 2) & 1, c[1], grid, (double)hout[0] / niter,
                   (c[0] & 1) ? (double)hout[0] / mmas : 0.0, (double)hout[1] / niter, ms);
            if (c[0] & 1) printf(", %.1f TFLOP/s tf32 chip-wide", 2.0 * TM * TN * 8 * mmas * grid / (ms * 1e-3) / 1e12);
            printf("\n");
        }
    }
    return 0;
}
