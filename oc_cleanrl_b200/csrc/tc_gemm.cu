// fp32-grade GEMM on the 5th-gen tensor cores: tcgen05.mma kind::tf32, three passes
// (a_lo*b_hi + a_hi*b_lo + a_hi*b_hi, with a = a_hi + a_lo split to TF32 precision),
// fp32 accumulation in TMEM.  Same three contractions as sgemm.cu:
//   FWD   : Y[M,N]  = relu(X[M,K] . W[N,K]^T + b[N])      A K-major,  B K-major
//   DGRAD : dX[M,N] = (mask > 0) * (dY[M,K] . W[K,N])      A K-major,  B MN-major
//   WGRAD : dW[M,N] = dY[K,M]^T . X[K,N]  (split-K)        A MN-major, B MN-major
//
// Persistent, warp-specialised: one CTA per SM walks 128 x 256 output tiles.
//   warps 0-3   epilogue : tcgen05.ld the finished accumulator, bias/ReLU or ReLU-mask, transpose
//                          through a per-warp smem pad, 128 B-coalesced STG
//   warp  4     MMA      : one thread issues tcgen05.mma; accumulators double-buffered in TMEM
//                          (2 x 256 columns) so the epilogue of tile i overlaps the mainloop of i+1
//   warps 5-7   idle (pad the MMA warp to a warpgroup so setmaxnreg can move registers to the producers)
//   warps 8-15  producers: LDG.128 -> registers (two k-blocks in flight) -> hi/lo split -> STS into
//                          the canonical UMMA layouts (K-major: SWIZZLE_128B, MN-major tf32:
//                          SWIZZLE_128B_BASE32B), 2 stages x 96 KB
// All hand-offs are mbarriers: full/empty per smem stage, tmem_full/tmem_empty per accumulator.
#include "common.cuh"
#include <stdlib.h>

namespace ocppo {

namespace {

constexpr int TM = 128, TN = 256, TK = 32;
constexpr int kEpiWarps = 4, kProdWarps = 8;
constexpr int kProdThreads = kProdWarps * 32;               // 256
constexpr int kProdTid0 = 256;                              // warpgroups 2-3
constexpr int kTcThreads = kProdTid0 + kProdThreads;        // 512 = 4 warpgroups (setmaxnreg is per warpgroup)
constexpr int kStageBytesA = TM * TK * 4;                   // 16 KB
constexpr int kStageBytesB = TN * TK * 4;                   // 32 KB
constexpr int kStageBytes = 2 * (kStageBytesA + kStageBytesB);   // hi + lo: 96 KB
constexpr int kNumStages = 2;
constexpr int kTmemCols = 512;                              // two 256-column accumulators
constexpr int kEpiPadBytes = kEpiWarps * 32 * 33 * 4;       // per-warp 32 x 33 transpose pads
constexpr int kBiasRedBytes = 8 * TM * 4;                   // WGRAD bias partials [8 k-lanes][128]
constexpr int kTcSmemBytes = kNumStages * kStageBytes + kEpiPadBytes + kBiasRedBytes + 128 /*barriers*/ + 1024 /*align*/;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
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
// layout the hardware takes for MN-major 32-bit (tf32) operands.
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
__device__ __forceinline__ void split4(const float4 &v, float4 &h, float4 &l) {
    split_tf32(v.x, h.x, l.x); split_tf32(v.y, h.y, l.y);
    split_tf32(v.z, h.z, l.z); split_tf32(v.w, h.w, l.w);
}

// ---- operand staging (p = producer-local thread id, 0..255) ------------------
// K-major tile of ROWS rows x 32 floats: 8 threads per row (16 B each), 32 rows per pass.
template <int ROWS>
__device__ __forceinline__ void ldg_kmajor(float4 (&r)[ROWS / 32], const float *__restrict__ base, int64_t ld,
                                           int row0, int row_limit, int64_t k0, int64_t k_limit, int p) {
    const int rr = p >> 3, c = p & 7;
#pragma unroll
    for (int q = 0; q < ROWS / 32; ++q) {
        const int row = row0 + q * 32 + rr;
        const int64_t k = k0 + c * 4;
        if (row < row_limit && k < k_limit)
            r[q] = __ldg(reinterpret_cast<const float4 *>(base + (int64_t)row * ld + k));
        else
            r[q] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
}
template <int ROWS>
__device__ __forceinline__ void sts_kmajor(uint8_t *hi, uint8_t *lo, const float4 (&r)[ROWS / 32], int p) {
    const int rr = p >> 3, c = p & 7;
#pragma unroll
    for (int q = 0; q < ROWS / 32; ++q) {
        const int row = q * 32 + rr;
        const int off = (row >> 3) * 1024 + (row & 7) * 128 + ((c ^ (row & 7)) << 4);
        float4 h, l;
        split4(r[q], h, l);
        *reinterpret_cast<float4 *>(hi + off) = h;
        *reinterpret_cast<float4 *>(lo + off) = l;
    }
}
// MN-major tile of 32 k-rows x W contiguous floats: W/4 chunks per k-row.
template <int W>
__device__ __forceinline__ void ldg_mnmajor(float4 (&r)[W / 32], const float *__restrict__ base, int64_t ld,
                                            int col0, int col_limit, int64_t k0, int64_t k_limit, int p) {
    constexpr int kChunks = W / 4;                 // per k-row
    constexpr int kRowsPerPass = kProdThreads / kChunks;
    const int q = p % kChunks, kr = p / kChunks;
#pragma unroll
    for (int i = 0; i < W / 32; ++i) {
        const int64_t k = k0 + i * kRowsPerPass + kr;
        const int col = col0 + q * 4;
        if (k < k_limit && col < col_limit)
            r[i] = __ldg(reinterpret_cast<const float4 *>(base + k * ld + col));
        else
            r[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
}
template <int W>
__device__ __forceinline__ void sts_mnmajor(uint8_t *hi, uint8_t *lo, const float4 (&r)[W / 32], int p) {
    constexpr int kChunks = W / 4;
    constexpr int kRowsPerPass = kProdThreads / kChunks;
    const int q = p % kChunks, kr = p / kChunks;
    const int mb = q >> 3, c = q & 7;
#pragma unroll
    for (int i = 0; i < W / 32; ++i) {
        const int k = i * kRowsPerPass + kr;
        // SW128_BASE32B atom: 32 MN floats (128 B) x 4 k-rows; 32 B chunks XOR (k & 3)
        const int off = mb * 4096 + (k >> 2) * 512 + (k & 3) * 128 + ((((c >> 1) ^ (k & 3)) << 5) | ((c & 1) << 4));
        float4 h, l;
        split4(r[i], h, l);
        *reinterpret_cast<float4 *>(hi + off) = h;
        *reinterpret_cast<float4 *>(lo + off) = l;
    }
}

struct TileCoord { int m0, n0, split; int64_t kbeg, kend; int nk; };

__device__ __forceinline__ TileCoord tile_coord(const SgemmArgs &p, int tile, int nt_n, int nt_m) {
    TileCoord t;
    const int nb = tile % nt_n;
    const int rest = tile / nt_n;
    const int mb = rest % nt_m;
    t.split = rest / nt_m;
    t.m0 = mb * TM;
    t.n0 = nb * TN;
    t.kbeg = (int64_t)t.split * p.k_per_split;
    t.kend = (t.kbeg + p.k_per_split < p.K) ? t.kbeg + p.k_per_split : p.K;
    t.nk = (int)((t.kend - t.kbeg + TK - 1) / TK);
    return t;
}

#define OCPPO_TMEM_LD32(v, taddr)                                                                                   \
    asm volatile(                                                                                                   \
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                                   \
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "                                   \
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"                   \
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),           \
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),     \
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),   \
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])    \
        : "r"(taddr)                                                                                                \
        : "memory")

template <int MODE>
__global__ void __launch_bounds__(kTcThreads, 1) tc_gemm_kernel(SgemmArgs p, int nt_n, int nt_m, int ntiles) {
    constexpr bool A_MN = (MODE == SGEMM_WGRAD);
    constexpr bool B_MN = (MODE != SGEMM_FWD);
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    float *epi_pad = reinterpret_cast<float *>(smem + kNumStages * kStageBytes);
    float *bias_red = reinterpret_cast<float *>(smem + kNumStages * kStageBytes + kEpiPadBytes);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + kNumStages * kStageBytes + kEpiPadBytes + kBiasRedBytes);
    // bars[0..1] full, [2..3] empty, [4..5] tmem_full, [6..7] tmem_empty
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 8);
    const uint32_t bar_full = smem_u32(&bars[0]), bar_empty = smem_u32(&bars[2]);
    const uint32_t bar_tfull = smem_u32(&bars[4]), bar_tempty = smem_u32(&bars[6]);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (tid == 0) {
        for (int s = 0; s < kNumStages; ++s) {
            mbar_init(bar_full + 8 * s, kProdWarps);
            mbar_init(bar_empty + 8 * s, 1);
            mbar_init(bar_tfull + 8 * s, 1);
            mbar_init(bar_tempty + 8 * s, kEpiWarps);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kEpiWarps) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"((uint32_t)kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;

    if (warp < kEpiWarps) {
        // ===================== epilogue =====================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 96;");
        float *pad = epi_pad + warp * 32 * 33;
        int ti = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++ti) {
            const TileCoord t = tile_coord(p, tile, nt_n, nt_m);
            const int acc = ti & 1;
            mbar_wait(bar_tfull + 8 * acc, (uint32_t)((ti >> 1) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            float *Cb = p.C;
            if (MODE == SGEMM_WGRAD) Cb += (int64_t)t.split * p.M * p.N;
            const int r_sub = lane >> 3, col = (lane & 7) * 4;
#pragma unroll 1
            for (int cc = 0; cc < TN / 32; ++cc) {
                uint32_t v[32];
                const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(acc * TN + cc * 32);
                OCPPO_TMEM_LD32(v, taddr);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (cc == TN / 32 - 1) {       // accumulator drained: hand it back to the MMA warp
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_tempty + 8 * acc);
                }
#pragma unroll
                for (int j = 0; j < 32; ++j) pad[lane * 33 + j] = __uint_as_float(v[j]);
                __syncwarp();
                const int n = t.n0 + cc * 32 + col;
                float4 bb = make_float4(0.f, 0.f, 0.f, 0.f);
                if (MODE == SGEMM_FWD) bb = __ldg(reinterpret_cast<const float4 *>(p.bias + n));
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int row = i * 4 + r_sub;
                    const int m = t.m0 + warp * 32 + row;
                    const float *src = pad + row * 33 + col;
                    float4 o = make_float4(src[0], src[1], src[2], src[3]);
                    if (m < p.M) {
                        if (MODE == SGEMM_FWD) {
                            o.x = fmaxf(o.x + bb.x, 0.f); o.y = fmaxf(o.y + bb.y, 0.f);
                            o.z = fmaxf(o.z + bb.z, 0.f); o.w = fmaxf(o.w + bb.w, 0.f);
                        } else if (MODE == SGEMM_DGRAD) {
                            const float4 mk = __ldg(reinterpret_cast<const float4 *>(p.mask + (int64_t)m * p.ldc + n));
                            o.x = mk.x > 0.f ? o.x : 0.f; o.y = mk.y > 0.f ? o.y : 0.f;
                            o.z = mk.z > 0.f ? o.z : 0.f; o.w = mk.w > 0.f ? o.w : 0.f;
                        }
                        *reinterpret_cast<float4 *>(Cb + (int64_t)m * p.ldc + n) = o;
                    }
                }
                __syncwarp();
            }
        }
    } else if (warp < 2 * kEpiWarps) {
        // ===================== MMA issuer (warp 4; warps 5-7 idle) =====================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        if (warp == kEpiWarps && lane == 0) {
            constexpr uint32_t idesc = make_idesc(TM, TN, A_MN, B_MN);
            // K-major: SBO = 8-row group stride (1024 B), LBO unused.  MN-major: LBO = stride between
            // 32-float MN chunks (4096 B), SBO = stride between 4-row k groups (512 B).
            constexpr uint32_t a_lbo = A_MN ? 4096 : 16, a_sbo = A_MN ? 512 : 1024, a_ty = A_MN ? 1 : 2;
            constexpr uint32_t b_lbo = B_MN ? 4096 : 16, b_sbo = B_MN ? 512 : 1024, b_ty = B_MN ? 1 : 2;
            // per k-step of 8: K-major advances 32 B inside the 128 B row, MN-major two 512 B atoms
            constexpr uint64_t a_step = (A_MN ? 1024 : 32) >> 4, b_step = (B_MN ? 1024 : 32) >> 4;
            int it = 0, ti = 0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++ti) {
                const TileCoord t = tile_coord(p, tile, nt_n, nt_m);
                const int acc = ti & 1;
                mbar_wait(bar_tempty + 8 * acc, (uint32_t)(((ti >> 1) & 1) ^ 1));
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * TN);
                for (int kb = 0; kb < t.nk; ++kb, ++it) {
                    const int s = it & 1;
                    mbar_wait(bar_full + 8 * s, (uint32_t)((it >> 1) & 1));
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    uint8_t *st = smem + s * kStageBytes;
                    const uint32_t a_hi = smem_u32(st), a_lo = a_hi + kStageBytesA;
                    const uint32_t b_hi = a_hi + 2 * kStageBytesA, b_lo = b_hi + kStageBytesB;
                    const uint64_t da_hi = make_desc(a_hi, a_lbo, a_sbo, a_ty), da_lo = make_desc(a_lo, a_lbo, a_sbo, a_ty);
                    const uint64_t db_hi = make_desc(b_hi, b_lbo, b_sbo, b_ty), db_lo = make_desc(b_lo, b_lbo, b_sbo, b_ty);
#pragma unroll
                    for (int ks = 0; ks < TK / 8; ++ks) {
                        const uint32_t acc0 = (kb > 0 || ks > 0) ? 1u : 0u;
                        umma_tf32(d_tmem, da_lo + ks * a_step, db_hi + ks * b_step, idesc, acc0);
                        umma_tf32(d_tmem, da_hi + ks * a_step, db_lo + ks * b_step, idesc, 1u);
                        umma_tf32(d_tmem, da_hi + ks * a_step, db_hi + ks * b_step, idesc, 1u);
                    }
                    umma_commit(bar_empty + 8 * s);          // stage free once these MMAs have read it
                }
                umma_commit(bar_tfull + 8 * acc);            // accumulator complete
            }
        }
        __syncwarp();
    } else {
        // ===================== producers =====================
        asm volatile("setmaxnreg.inc.sync.aligned.u32 168;");
        const int pt = tid - kProdTid0;
        const int pwarp_lane0 = (lane == 0);
        int it = 0;
        float4 ra0[4], rb0[8], ra1[4], rb1[8];
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            const TileCoord t = tile_coord(p, tile, nt_n, nt_m);
            const bool do_bias = (MODE == SGEMM_WGRAD) && t.n0 == 0 && p.bias_part != nullptr;
            float4 bsum = make_float4(0.f, 0.f, 0.f, 0.f);
            auto load = [&](int kb, float4 (&ra)[4], float4 (&rb)[8]) {
                const int64_t k0 = t.kbeg + (int64_t)kb * TK;
                if (A_MN) ldg_mnmajor<TM>(ra, p.A, p.lda, t.m0, p.M, k0, t.kend, pt);
                else ldg_kmajor<TM>(ra, p.A, p.lda, t.m0, p.M, k0, t.kend, pt);
                if (B_MN) ldg_mnmajor<TN>(rb, p.B, p.ldb, t.n0, p.N, k0, t.kend, pt);
                else ldg_kmajor<TN>(rb, p.B, p.ldb, t.n0, p.N, k0, t.kend, pt);
            };
            auto store = [&](const float4 (&ra)[4], const float4 (&rb)[8]) {
                const int s = it & 1;
                mbar_wait(bar_empty + 8 * s, (uint32_t)(((it >> 1) & 1) ^ 1));
                uint8_t *st = smem + s * kStageBytes;
                uint8_t *a_hi = st, *a_lo = st + kStageBytesA, *b_hi = st + 2 * kStageBytesA, *b_lo = b_hi + kStageBytesB;
                if (MODE == SGEMM_WGRAD) {
                    if (do_bias) {
#pragma unroll
                        for (int q = 0; q < 4; ++q) { bsum.x += ra[q].x; bsum.y += ra[q].y; bsum.z += ra[q].z; bsum.w += ra[q].w; }
                    }
                }
                if (A_MN) sts_mnmajor<TM>(a_hi, a_lo, ra, pt);
                else sts_kmajor<TM>(a_hi, a_lo, ra, pt);
                if (B_MN) sts_mnmajor<TN>(b_hi, b_lo, rb, pt);
                else sts_kmajor<TN>(b_hi, b_lo, rb, pt);
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (pwarp_lane0) mbar_arrive(bar_full + 8 * s);
                ++it;
            };
            load(0, ra0, rb0);
            if (t.nk > 1) load(1, ra1, rb1);
            for (int kb = 0; kb < t.nk; kb += 2) {
                store(ra0, rb0);
                if (kb + 2 < t.nk) load(kb + 2, ra0, rb0);
                if (kb + 1 < t.nk) {
                    store(ra1, rb1);
                    if (kb + 3 < t.nk) load(kb + 3, ra1, rb1);
                }
            }
            if (MODE == SGEMM_WGRAD) {
                if (do_bias) {           // tile-uniform across the producer warps
                    // [8 k-lanes][128 columns]; thread pt holds columns (pt % 32) * 4.. of k-lane pt / 32
                    *reinterpret_cast<float4 *>(bias_red + (pt >> 5) * TM + (pt & 31) * 4) = bsum;
                    asm volatile("bar.sync 1, %0;" ::"n"(kProdThreads) : "memory");
                    if (pt < TM) {
                        float sum = 0.f;
#pragma unroll
                        for (int kr = 0; kr < 8; ++kr) sum += bias_red[kr * TM + pt];
                        if (t.m0 + pt < p.M) p.bias_part[(int64_t)t.split * p.M + t.m0 + pt] = sum;
                    }
                    asm volatile("bar.sync 1, %0;" ::"n"(kProdThreads) : "memory");
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == kEpiWarps) {
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
    return true;
}

int tc_gemm_launch(SgemmMode mode, const SgemmArgs &a, cudaStream_t stream) {
    OCPPO_CHECK_ARG(tc_gemm_supported(mode, a), "tc_gemm: unsupported shape M=%d N=%d K=%lld", a.M, a.N, (long long)a.K);
    const int splits = mode == SGEMM_WGRAD ? a.splits : 1;
    const int nt_n = a.N / TN, nt_m = (int)ceil_div(a.M, TM);
    const int64_t ntiles64 = (int64_t)nt_n * nt_m * splits;
    OCPPO_CHECK_ARG(ntiles64 < (1LL << 31), "tc_gemm: too many tiles");
    const int ntiles = (int)ntiles64;
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        OCPPO_CUDA(cudaGetDevice(&dev));
        OCPPO_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        OCPPO_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<SGEMM_FWD>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes));
        OCPPO_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<SGEMM_DGRAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes));
        OCPPO_CUDA(cudaFuncSetAttribute(tc_gemm_kernel<SGEMM_WGRAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes));
    }
    const int grid = ntiles < sms ? ntiles : sms;
    ProfScope ps(mode == SGEMM_FWD ? KC_SGEMM_FWD : (mode == SGEMM_DGRAD ? KC_SGEMM_DGRAD : KC_SGEMM_WGRAD), 1, stream);
    if (mode == SGEMM_FWD) tc_gemm_kernel<SGEMM_FWD><<<grid, kTcThreads, kTcSmemBytes, stream>>>(a, nt_n, nt_m, ntiles);
    else if (mode == SGEMM_DGRAD) tc_gemm_kernel<SGEMM_DGRAD><<<grid, kTcThreads, kTcSmemBytes, stream>>>(a, nt_n, nt_m, ntiles);
    else tc_gemm_kernel<SGEMM_WGRAD><<<grid, kTcThreads, kTcSmemBytes, stream>>>(a, nt_n, nt_m, ntiles);
    OCPPO_LAUNCH_CHECK("tc_gemm_kernel");
    return OCPPO_OK;
}

}  // namespace ocppo
