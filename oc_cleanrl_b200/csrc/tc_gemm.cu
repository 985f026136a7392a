// fp32-grade GEMM on the 5th-gen tensor cores: tcgen05.mma kind::tf32, three passes
// (a_lo*b_hi + a_hi*b_lo + a_hi*b_hi, with a = a_hi + a_lo split to TF32 precision),
// fp32 accumulation in TMEM.  Same three contractions as sgemm.cu:
//   FWD   : Y[M,N]  = relu(X[M,K] . W[N,K]^T + b[N])      A K-major,  B K-major
//   DGRAD : dX[M,N] = (mask > 0) * (dY[M,K] . W[K,N])      A K-major,  B MN-major
//   WGRAD : dW[M,N] = dY[K,M]^T . X[K,N]  (split-K)        A MN-major, B MN-major
// Tile 128 x 256 x 32 per stage, 2 stages, one CTA per SM (192 KB smem, 256 TMEM columns).
// Operands travel global -> registers (LDG.128, split into hi/lo) -> shared memory in the
// canonical UMMA SWIZZLE_128B layouts -> tcgen05.mma issued by one thread; completion is
// tracked with tcgen05.commit on an mbarrier per stage.  Epilogue: tcgen05.ld 32x32b.
#include "common.cuh"
#include <stdlib.h>

namespace ocppo {

namespace {

constexpr int TM = 128, TN = 256, TK = 32;
constexpr int kTcThreads = 256;
constexpr int kStageBytesA = TM * TK * 4;                 // 16 KB
constexpr int kStageBytesB = TN * TK * 4;                 // 32 KB
constexpr int kStageBytes = 2 * (kStageBytesA + kStageBytesB);   // hi + lo: 96 KB
constexpr int kNumStages = 2;
constexpr int kTmemCols = 256;
constexpr int kTcSmemBytes = kNumStages * kStageBytes + 1024 /*align slack*/ + 64;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    const long long t0 = clock64();
    uint32_t done = 0;
    while (true) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (done) break;
        if (clock64() - t0 > 4000000000LL) __trap();      // never hang the device on a protocol bug
    }
}

// 64-bit shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout).
// layout_type: 2 = SWIZZLE_128B (K-major tiles), 1 = SWIZZLE_128B_BASE32B -- the only
// layout the hardware takes for MN-major 32-bit (tf32) operands (CUTLASS
// sm100_common.inl: "for mn-major tf32 operands, SW128_32B is the only available smem layout").
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                              uint32_t layout_type) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);               // start address, bits [0,14)
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;     // leading byte offset, bits [16,30)
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;     // stride byte offset, bits [32,46)
    d |= (uint64_t)1 << 46;                               // version = 1 (Blackwell)
    d |= (uint64_t)layout_type << 61;
    return d;
}

// kind::tf32 instruction descriptor (cute::UMMA::InstrDescriptor): F32 accumulate, TF32 A/B.
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, bool a_mn_major, bool b_mn_major) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((a_mn_major ? 1u : 0u) << 15) | ((b_mn_major ? 1u : 0u) << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

__device__ __forceinline__ void split_tf32(float x, float &hi, float &lo) {
    uint32_t h;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));
    hi = __uint_as_float(h);
    const float r = x - hi;                                // exact
    uint32_t l;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(l) : "f"(r));
    lo = __uint_as_float(l);
}

// ---- operand staging -------------------------------------------------------
// K-major tile of ROWS rows x 32 floats: 8 threads per row (16 B each), 32 rows per pass.
template <int ROWS>
__device__ __forceinline__ void ldg_kmajor(float4 (&r)[ROWS / 32], const float *__restrict__ base, int64_t ld,
                                           int row0, int row_limit, int64_t k0, int64_t k_limit, int tid) {
    const int rr = tid >> 3, c = tid & 7;
#pragma unroll
    for (int p = 0; p < ROWS / 32; ++p) {
        const int row = row0 + p * 32 + rr;
        const int64_t k = k0 + c * 4;
        if (row < row_limit && k < k_limit)
            r[p] = *reinterpret_cast<const float4 *>(base + (int64_t)row * ld + k);
        else
            r[p] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
}
template <int ROWS>
__device__ __forceinline__ void sts_kmajor(uint8_t *hi, uint8_t *lo, const float4 (&r)[ROWS / 32], int tid) {
    const int rr = tid >> 3, c = tid & 7;
#pragma unroll
    for (int p = 0; p < ROWS / 32; ++p) {
        const int row = p * 32 + rr;
        const int off = (row >> 3) * 1024 + (row & 7) * 128 + ((c ^ (row & 7)) << 4);
        float4 h, l;
        split_tf32(r[p].x, h.x, l.x); split_tf32(r[p].y, h.y, l.y);
        split_tf32(r[p].z, h.z, l.z); split_tf32(r[p].w, h.w, l.w);
        *reinterpret_cast<float4 *>(hi + off) = h;
        *reinterpret_cast<float4 *>(lo + off) = l;
    }
}
// MN-major tile of 32 k-rows x W contiguous floats: W/4 chunks per k-row.
template <int W>
__device__ __forceinline__ void ldg_mnmajor(float4 (&r)[W / 32], const float *__restrict__ base, int64_t ld,
                                            int col0, int col_limit, int64_t k0, int64_t k_limit, int tid) {
    constexpr int kChunks = W / 4;                 // per k-row
    constexpr int kRowsPerPass = kTcThreads / kChunks;
    const int q = tid % kChunks, kr = tid / kChunks;
#pragma unroll
    for (int p = 0; p < W / 32; ++p) {
        const int64_t k = k0 + p * kRowsPerPass + kr;
        const int col = col0 + q * 4;
        if (k < k_limit && col < col_limit)
            r[p] = *reinterpret_cast<const float4 *>(base + k * ld + col);
        else
            r[p] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
}
template <int W>
__device__ __forceinline__ void sts_mnmajor(uint8_t *hi, uint8_t *lo, const float4 (&r)[W / 32], int tid) {
    constexpr int kChunks = W / 4;
    constexpr int kRowsPerPass = kTcThreads / kChunks;
    const int q = tid % kChunks, kr = tid / kChunks;
    const int mb = q >> 3, c = q & 7;
#pragma unroll
    for (int p = 0; p < W / 32; ++p) {
        const int k = p * kRowsPerPass + kr;
        // SW128_BASE32B atom: 32 MN floats (128 B) x 4 k-rows; 32 B chunks XOR (k & 3)
        const int off = mb * 4096 + (k >> 2) * 512 + (k & 3) * 128 + ((((c >> 1) ^ (k & 3)) << 5) | ((c & 1) << 4));
        float4 h, l;
        split_tf32(r[p].x, h.x, l.x); split_tf32(r[p].y, h.y, l.y);
        split_tf32(r[p].z, h.z, l.z); split_tf32(r[p].w, h.w, l.w);
        *reinterpret_cast<float4 *>(hi + off) = h;
        *reinterpret_cast<float4 *>(lo + off) = l;
    }
}

template <int MODE>
__global__ void __launch_bounds__(kTcThreads, 1) tc_gemm_kernel(SgemmArgs p) {
    constexpr bool A_MN = (MODE == SGEMM_WGRAD);
    constexpr bool B_MN = (MODE != SGEMM_FWD);
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + kNumStages * kStageBytes);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + kNumStages);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
    const int split = blockIdx.z;
    const int64_t kbeg = (int64_t)split * p.k_per_split;
    const int64_t kend = (kbeg + p.k_per_split < p.K) ? kbeg + p.k_per_split : p.K;
    const int nk = (int)((kend - kbeg + TK - 1) / TK);

    if (tid == 0) {
        for (int s = 0; s < kNumStages; ++s) mbar_init(smem_u32(&bars[s]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"((uint32_t)kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;
    constexpr uint32_t idesc = make_idesc(TM, TN, A_MN, B_MN);

    float4 ra[4], rb[8];
    auto load_regs = [&](int kb) {
        const int64_t k0 = kbeg + (int64_t)kb * TK;
        if (A_MN) ldg_mnmajor<TM>(ra, p.A, p.lda, m0, p.M, k0, kend, tid);
        else ldg_kmajor<TM>(ra, p.A, p.lda, m0, p.M, k0, kend, tid);
        if (B_MN) ldg_mnmajor<TN>(rb, p.B, p.ldb, n0, p.N, k0, kend, tid);
        else ldg_kmajor<TN>(rb, p.B, p.ldb, n0, p.N, k0, kend, tid);
    };

    // WGRAD: column sums of dY (the bias gradient) ride along in the A-operand registers
    const bool do_bias = (MODE == SGEMM_WGRAD) && blockIdx.x == 0 && p.bias_part != nullptr;
    float4 bsum = make_float4(0.f, 0.f, 0.f, 0.f);

    if (nk > 0) load_regs(0);
    for (int kb = 0; kb < nk; ++kb) {
        const int s = kb & 1;
        uint8_t *st = smem + s * kStageBytes;
        uint8_t *a_hi = st, *a_lo = st + kStageBytesA, *b_hi = st + 2 * kStageBytesA, *b_lo = b_hi + kStageBytesB;
        if (kb >= kNumStages) mbar_wait(smem_u32(&bars[s]), (uint32_t)(((kb >> 1) - 1) & 1));
        if (MODE == SGEMM_WGRAD) {
            if (do_bias) {
#pragma unroll
                for (int q = 0; q < 4; ++q) { bsum.x += ra[q].x; bsum.y += ra[q].y; bsum.z += ra[q].z; bsum.w += ra[q].w; }
            }
        }
        if (A_MN) sts_mnmajor<TM>(a_hi, a_lo, ra, tid);
        else sts_kmajor<TM>(a_hi, a_lo, ra, tid);
        if (B_MN) sts_mnmajor<TN>(b_hi, b_lo, rb, tid);
        else sts_kmajor<TN>(b_hi, b_lo, rb, tid);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            // K-major: SBO = 8-row group stride (1024 B), LBO unused.  MN-major: LBO = stride between
            // 32-float MN chunks (4096 B), SBO = stride between 4-row k groups (512 B).
            const uint32_t mn_lbo = p.dbg == 1 ? 512 : 4096, mn_sbo = p.dbg == 1 ? 4096 : 512;
            const uint64_t da_hi = make_desc(smem_u32(a_hi), A_MN ? mn_lbo : 16, A_MN ? mn_sbo : 1024, A_MN ? 1 : 2);
            const uint64_t da_lo = make_desc(smem_u32(a_lo), A_MN ? mn_lbo : 16, A_MN ? mn_sbo : 1024, A_MN ? 1 : 2);
            const uint64_t db_hi = make_desc(smem_u32(b_hi), B_MN ? mn_lbo : 16, B_MN ? mn_sbo : 1024, B_MN ? 1 : 2);
            const uint64_t db_lo = make_desc(smem_u32(b_lo), B_MN ? mn_lbo : 16, B_MN ? mn_sbo : 1024, B_MN ? 1 : 2);
            // per k-step of 8: K-major advances 32 B inside the 128 B row, MN-major one 1024 B atom
            constexpr uint64_t a_step = (A_MN ? 1024 : 32) >> 4, b_step = (B_MN ? 1024 : 32) >> 4;
#pragma unroll
            for (int ks = 0; ks < TK / 8; ++ks) {
                const uint32_t acc0 = (kb > 0 || ks > 0) ? 1u : 0u;
                umma_tf32(tmem_base, da_lo + ks * a_step, db_hi + ks * b_step, idesc, acc0);
                umma_tf32(tmem_base, da_hi + ks * a_step, db_lo + ks * b_step, idesc, 1u);
                umma_tf32(tmem_base, da_hi + ks * a_step, db_hi + ks * b_step, idesc, 1u);
            }
            umma_commit(smem_u32(&bars[s]));
        }
        if (kb + 1 < nk) load_regs(kb + 1);
    }
    // ---- epilogue: TMEM -> registers -> global ----
    if (nk > 0) {
        const int last = nk - 1;
        mbar_wait(smem_u32(&bars[last & 1]), (uint32_t)((last >> 1) & 1));
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    {
        const int g = warp & 3, hh = warp >> 2;
        const int m = m0 + g * 32 + lane;
        float *Cb = p.C;
        if (MODE == SGEMM_WGRAD) Cb += (int64_t)split * p.M * p.N;
#pragma unroll 1
        for (int cc = 0; cc < 4; ++cc) {
            const int ncol = hh * 128 + cc * 32;
            uint32_t v[32];
            const uint32_t taddr = tmem_base + ((uint32_t)(g * 32) << 16) + (uint32_t)ncol;
            if (nk > 0) {
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                    : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                      "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]),
                      "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]),
                      "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]),
                      "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                    : "r"(taddr)
                    : "memory");
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] = 0u;
            }
            const int n = n0 + ncol;
            if (m < p.M && n < p.N) {
                float *dst = Cb + (int64_t)m * p.ldc + n;
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                    float4 o = make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                                           __uint_as_float(v[j + 3]));
                    if (MODE == SGEMM_FWD) {
                        const float4 bb = *reinterpret_cast<const float4 *>(p.bias + n + j);
                        o.x = fmaxf(o.x + bb.x, 0.f); o.y = fmaxf(o.y + bb.y, 0.f);
                        o.z = fmaxf(o.z + bb.z, 0.f); o.w = fmaxf(o.w + bb.w, 0.f);
                    } else if (MODE == SGEMM_DGRAD) {
                        const float4 mk = *reinterpret_cast<const float4 *>(p.mask + (int64_t)m * p.ldc + n + j);
                        o.x = mk.x > 0.f ? o.x : 0.f; o.y = mk.y > 0.f ? o.y : 0.f;
                        o.z = mk.z > 0.f ? o.z : 0.f; o.w = mk.w > 0.f ? o.w : 0.f;
                    }
                    *reinterpret_cast<float4 *>(dst + j) = o;
                }
            }
        }
    }
    if (MODE == SGEMM_WGRAD) {
        if (do_bias) {          // CTA-uniform; all MMAs have completed, stage memory is free
            float4 *red = reinterpret_cast<float4 *>(smem);      // [8 k-lanes][32 column quads]
            red[(tid >> 5) * 32 + (tid & 31)] = bsum;
            __syncthreads();
            if (tid < TM) {
                const float *rf = reinterpret_cast<const float *>(red);
                float s = 0.f;
#pragma unroll
                for (int kr = 0; kr < 8; ++kr) s += rf[kr * 128 + tid];
                if (m0 + tid < p.M) p.bias_part[(int64_t)split * p.M + m0 + tid] = s;
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)kTmemCols)
                     : "memory");
    }
}

inline bool al16(const void *p) { return (((uintptr_t)p) & 15) == 0; }

}  // namespace

// Shapes the tensor-core kernel takes; everything else stays on the FMA kernel.
bool tc_gemm_supported(SgemmMode mode, const SgemmArgs &a) {
    if (a.M <= 0 || a.N <= 0 || a.K <= 0) return false;
    if (a.N % TN != 0) return false;
    if ((a.lda & 3) || (a.ldb & 3) || (a.ldc & 3) || !al16(a.A) || !al16(a.B) || !al16(a.C)) return false;
    if (mode == SGEMM_WGRAD) {
        if (a.M % TM != 0) return false;
        if (a.k_per_split % TK != 0 || a.K % 4 != 0) return false;
    } else {
        if (a.K % TK != 0) return false;
        if (mode == SGEMM_FWD && !al16(a.bias)) return false;
        if (mode == SGEMM_DGRAD && !al16(a.mask)) return false;
    }
    if (ceil_div(a.M, TM) > 65535) return false;
    return true;
}

int tc_gemm_launch(SgemmMode mode, const SgemmArgs &a_in, cudaStream_t stream) {
    static const int dbg = getenv("OCPPO_TC_DBG") ? atoi(getenv("OCPPO_TC_DBG")) : 0;   // bring-up knob only
    SgemmArgs a = a_in;
    a.dbg = dbg;
    OCPPO_CHECK_ARG(tc_gemm_supported(mode, a), "tc_gemm: unsupported shape M=%d N=%d K=%lld", a.M, a.N, (long long)a.K);
    const int splits = mode == SGEMM_WGRAD ? a.splits : 1;
    dim3 grid((unsigned)(a.N / TN), (unsigned)ceil_div(a.M, TM), (unsigned)splits);
    static bool attr_done = false;
    if (!attr_done) {
        OCPPO_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<SGEMM_FWD>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes));
        OCPPO_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<SGEMM_DGRAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes));
        OCPPO_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<SGEMM_WGRAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes));
        attr_done = true;
    }
    ProfScope ps(mode == SGEMM_FWD ? KC_SGEMM_FWD : (mode == SGEMM_DGRAD ? KC_SGEMM_DGRAD : KC_SGEMM_WGRAD), 1, stream);
    if (mode == SGEMM_FWD) tc_gemm_kernel<SGEMM_FWD><<<grid, kTcThreads, kTcSmemBytes, stream>>>(a);
    else if (mode == SGEMM_DGRAD) tc_gemm_kernel<SGEMM_DGRAD><<<grid, kTcThreads, kTcSmemBytes, stream>>>(a);
    else tc_gemm_kernel<SGEMM_WGRAD><<<grid, kTcThreads, kTcSmemBytes, stream>>>(a);
    OCPPO_LAUNCH_CHECK("tc_gemm_kernel");
    return OCPPO_OK;
}

}  // namespace ocppo
