// fp32-grade GEMM on the 5th-gen tensor cores (tcgen05.mma, fp32 accumulation in TMEM).  Every operand is split
// x = hi + lo and the product is hi.hi plus two correction products (lo.lo, 2^-22, is dropped).  Same three contractions
// as sgemm.cu:
//   FWD   : Y[M,N]  = relu(X[M,K] . W[N,K]^T + b[N])      A K-major,  B K-major
//   DGRAD : dX[M,N] = (mask > 0) * (dY[M,K] . W[K,N])      A K-major,  B MN-major
//   WGRAD : dW[M,N] = dY[K,M]^T . X[K,N]  (split-K)        A MN-major, B MN-major
// Engines (template parameter CORR; ocppo_set_gemm_mode):
//   CORR 0  3xTF32: hi = trunc_tf32(x) (kind::tf32 on the raw fp32 tile: the hardware truncates), lo.hi + hi.lo as TF32 MMAs
//   CORR 1  TF32 main product + two BF16 correction products
//   CORR 2  FP16 x 3 on fp32-stored operands: x.s = h + l in fp16 (s a per-tensor power of two), three kind::f16 products;
//           the weights of FWD / DGRAD arrive pre-split by TMA (four 48 KB stages), the rest is converted by producer warps
//   CORR 3  FP16 x 3 on fp16-PAIR storage (the default, engines 6 / 8): both operands live in global memory as (hi, lo)
//           fp16 tensors and are TMA-loaded straight into the MMA tiles -- six 32 KB stages, no producer warps (320 threads);
//           FWD / DGRAD can write their output as a pair too (scaled by an a-priori bound), DGRAD adds the bias column sums
// SPLIT: main product in TMEM columns [0,256), corrections in [256,512) (3x less rounding, one tile in flight) | one
// 256-column accumulator per tile, two tiles in flight (the epilogue overlaps the next mainloop).
//
// Persistent, warp-specialised, CTA-pair kernel: a cluster of two CTAs (one SM each, same TPC) walks
// 256 x 256 output tiles with tcgen05.mma.cta_group::2.  Each CTA stages its own 128 rows of A and its own
// 128 columns of B (so the pair fetches every operand byte once), and holds the 128 x 256 fp32 accumulators of
// its rows in its own TMEM.  Per CTA:
//   warps 0-7   epilogue : two per TMEM lane quarter; tcgen05.ld, unscale, bias/ReLU/decision bits | ReLU mask + column
//                          sums | k-segment accumulate, STS into a swizzled tile, bulk tensor store(s) per 32 x 32 chunk
//   warp  8     MMA      : one thread of the LEADER CTA issues the pair's MMAs; in the peer CTA the same thread forwards
//                          "my tiles have landed" to the leader (engines without conversion on the critical path)
//   warp  9     TMA      : one thread per CTA loads this CTA's tiles (2-D maps for K-major operands, 3-D maps for MN-major
//                          ones: nothing is ever transposed in memory)
//   warps 10-17 producers: (CORR 0-2 only) read the raw fp32 tiles back and write the MMA operand tiles in the canonical
//                          UMMA layouts
// Hand-offs are mbarriers: tma (per CTA), raw_full / full (leader), empty (tcgen05.commit multicast to both CTAs),
// tmem_full / tmem_empty per accumulator.
#include "common.cuh"
#include <cuda.h>
#include <cuda_fp16.h>
#include <stdlib.h>

namespace ocppo {

namespace {

constexpr int TM = 128;                                     // rows of A (and of the accumulator) per CTA
// N tile of the CTA pair (= columns of one accumulator) and TMEM use, by accumulator scheme:
//   SPLIT (mode 1): main accumulator in TMEM columns [0,256), corrections in [256,512): one tile in flight;
//   single (mode 2): one 256-column accumulator per tile, two tiles in flight (epilogue overlaps the next tile).
// A 128-column tile would let the split scheme double-buffer too, and was measured: 10-50 % SLOWER on every trunk
// shape with K >= 512 -- at N = 128 the MMA re-reads its A operand from shared memory twice as often per flop
// (SS mode: 128 B/cycle of operand traffic, the whole shared-memory bandwidth), so the kernel becomes smem-bound.
// (A 128-column tile exists all the same, for the conversion-free kernel only: layer widths that are a multiple of 128 but
// not of 256 -- the reference class's own default encoder_dims=(128, 64), 384-wide layers -- would otherwise leave the
// tensor cores altogether and run on the 9x slower FMA kernel.)
template <bool SPLIT, int TN_> struct Tile {
    static constexpr int TN = TN_;
    static constexpr int TBW = TN / 2;                      // columns of B staged per CTA
    static constexpr int kStageBytesB = TBW * 32 * 4;       // 16 KB (8 KB for the 128-column tile)
    static constexpr int kStageBytes = 2 * (128 * 32 * 4 + kStageBytesB);   // hi + lo: 64 KB
    static constexpr int kNumStages = 3;                    // 192 KB
    static constexpr int kAccStride = SPLIT ? 2 * TN : TN;  // TMEM columns per tile
    static constexpr int kAccBufs = 512 / kAccStride >= 2 ? 2 : 1;   // tiles whose accumulators fit in 512 TMEM columns
};
constexpr int TK = 32;
constexpr int kPairM = 2 * TM;                              // 256-row tile per CTA pair
// 16 warps: 0-7 epilogue (two per TMEM lane quarter, each half of the tile's columns), 8 MMA issuer, 9 TMA issuer,
// 10-11 idle, 12-15 lo-tile producers
constexpr int kEpiWarps = 8, kProdWarps = 8;
constexpr int kProdThreads = kProdWarps * 32;               // 256
constexpr int kProdTid0 = (kEpiWarps + 2) * 32;             // 320
constexpr int kTcThreads = kProdTid0 + kProdThreads;        // 576: at most 112 registers per thread
constexpr int kProdQ = 1024 / kProdThreads;                 // 16-byte chunks of a 16 KB raw tile per producer thread (4)
constexpr int kStageBytesA = TM * TK * 4;                   // 16 KB
constexpr int kRingBytes = 192 * 1024;                      // 3 x 64 KB, 4 x 48 KB or 6 x 32 KB stages
constexpr int kMaxStages = 6;
constexpr int kEpi2Policy = 1, kEpi2MaxK = 512;            // see tc_gemm_launch (measured: profiles/r02_epi2_ab.txt)
constexpr int kTmemCols = 512;                              // two tiles' accumulators
// The TMEM accumulator rounds TOWARD ZERO on every accumulate (tools/tc_probe.cu Q4): each chained MMA shrinks the running
// sum by half an ulp on average, a systematic loss that does not average out -- through five layers forward and
// backward it put the full-size gradient at 8e-5 of the 1e-4 bar.  For a sum that grows roughly linearly over n chained
// MMAs the expected loss is  sum_k ulp(D_k)/2 = n * 2^-24 * (1/m - 2/(3 m^2)) * D  for a final mantissa m in [1,2), i.e.
// (2.0 .. 2.2)e-8 * n relative whatever m is; the measured slope on all-positive WGRAD inputs is 2.4e-8 per MMA in
// max-norm.  The epilogue multiplies the drained main accumulator by 1 + kTruncPerMma * n: a first-order correction of a
// known bias, not an exact one (a sum dominated by its first or its last terms loses up to twice / nothing).
constexpr float kTruncPerMma = 2.1e-8f;
constexpr int kEpiChunkBytes = 32 * 32 * 4;                 // one warp's 32 rows x 32 columns, SWIZZLE_128B
constexpr int kEpiPadBytes = kEpiWarps * kEpiChunkBytes;    // per warp: one chunk, incoming aux then outgoing result (32 KB)
constexpr int kBiasRedBytes = 4 * TM * 4;                   // WGRAD bias partials [4][128] (2 KB): the 8 k-lanes fold in two steps
constexpr int kBarBytes = 512;                               // 36 mbarriers + the TMEM slot
constexpr int kTcSmemBytes = kRingBytes + kEpiPadBytes + kBiasRedBytes + kBarBytes;   // base is 1024-aligned
static_assert(kTcSmemBytes <= 232448, "shared memory budget (227 KB per CTA)");

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at the same smem offset in CTA `cta` of the pair.  Plain (default-semantics) arrive
// and try_wait, as CUTLASS's ClusterBarrier uses for its transform -> 2-SM MMA pipelines: the explicit
// .release.cluster / .acquire.cluster forms compile to MEMBAR.ALL.GPU + ERRBAR per arrive and a CCTL.IVALL
// per poll, which ncu showed as 35% of all stall samples.  Writers fence.proxy.async before arriving.
__device__ __forceinline__ void mbar_arrive_on(uint32_t bar, uint32_t cta) {
    asm volatile(
        "{\n\t.reg .b32 ra;\n\t"
        "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
        "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t}"
        ::"r"(bar), "r"(cta)
        : "memory");
}
// (A suspend-time hint on try_wait was measured: no gain on FWD/WGRAD, 10-15 % slower DGRAD -- later wake-ups.)
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

// Long waits of warps that share an SM sub-partition with the producers (F16 engine: the conversion is on the
// critical path and every poll takes an issue slot from it)
__device__ __forceinline__ void mbar_wait_sleep(uint32_t bar, uint32_t parity, unsigned ns) {
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
        __nanosleep(ns);
        if (clock64() - t0 > 4000000000LL) __trap();
    }
}

// (Releasing the raw tiles of a stage to the TMA right after the main product -- half a stage earlier than the
// converted tiles -- was measured as well: no gain.)
// (nanosleep back-off in the long waits was measured too: the eight epilogue warps polling tmem_full issue 95 % of
// the kernel's instructions, yet sleeping between polls changes nothing, and 32 ns in the TMA issuer's wait costs
// DGRAD 10 % -- the mainloop is bound by the latency of the empty -> TMA -> convert -> full loop, not by issue slots.)
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

// the address-independent part of such a descriptor (everything but the 14-bit start address)
__host__ __device__ constexpr uint64_t make_desc_const(uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout_type) {
    return ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) | ((uint64_t)1 << 46) |
           ((uint64_t)layout_type << 61);
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
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// kind::f16 instruction descriptor with BF16 A/B and F32 accumulate (K = 16 per instruction)
__host__ __device__ constexpr uint32_t make_idesc_bf16(int M, int N, bool a_mn_major, bool b_mn_major) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((a_mn_major ? 1u : 0u) << 15) | ((b_mn_major ? 1u : 0u) << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// The same MMA with the two 64-bit shared-memory descriptors handed over as (low word, high word): only the 14-bit
// address field in the low word changes from stage to stage, so the issuer keeps 32-bit running values instead of
// rebuilding 64-bit descriptors (the MMA issuer is ONE thread whose instruction stream paces the tensor pipe).
__device__ __forceinline__ void umma_f16_lohi(uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                              uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
        "setp.ne.b32 p, %6, 0;\n\t"
        "mov.b64 da, {%1, %2};\n\t"
        "mov.b64 db, {%3, %4};\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %5, p;\n\t}"
        ::"r"(d_tmem), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
        : "memory");
}
// Warp-convergent variants: the WHOLE warp executes the issuer loop (every value in it is warp-uniform) and one elected
// lane issues.  Inside an `if (lane == 0)` region the compiler must assume arbitrary divergence and wraps every
// tcgen05 instruction in an ELECT / R2UR.BROADCAST / BRA.U.ANY loop (9 extra instructions per MMA).
__device__ __forceinline__ void umma_f16_lohi_elect(uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                                    uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p, q;\n\t.reg .b64 da, db;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "setp.ne.b32 p, %6, 0;\n\t"
        "mov.b64 da, {%1, %2};\n\t"
        "mov.b64 db, {%3, %4};\n\t"
        "@q tcgen05.mma.cta_group::2.kind::f16 [%0], da, db, %5, p;\n\t}"
        ::"r"(d_tmem), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit_pair_elect(uint32_t bar) {
    asm volatile(
        "{\n\t.reg .pred q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n\t}"
        ::"r"(bar), "h"((uint16_t)3)
        : "memory");
}
// (Measured on top of this and rejected, profiles/r02_issue_experiments.txt: the TMA issuer made warp-convergent the same
// way, and .cta_group::2 TMA loads counting both CTAs' bytes on the leader's barrier -- one wait per stage in the MMA
// issuer and no forwarder thread in the peer: correct, but 0 .. +5 % cycles per launch.)
// mainloop wait of the two single-thread roles: no clock reads in the polling loop (a spin counter bounds it instead)
__device__ __forceinline__ void mbar_wait_lean(uint32_t bar, uint32_t parity) {
    uint32_t done = 0, spins = 0;
    while (true) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (done) break;
        if (++spins > (1u << 28)) __trap();               // never hang the device on a protocol bug
    }
}
// arrives (count 1) on the barrier at this smem offset in BOTH CTAs once all prior MMAs have completed
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3)
                 : "memory");
}

// ---- lo-tile generation (p = producer-local thread id, 0..255) ---------------
// TMA has landed the raw fp32 tile in the canonical UMMA layout; kind::tf32 reads only its top 19 bits, so
// the raw tile IS the hi operand.  The producers read it back, form lo = x - trunc_tf32(x) rounded to the TF32
// grid and write it at the same offset of the lo tile (same layout, so no index arithmetic beyond the
// thread -> 16-byte chunk map below, chosen so every warp access is bank-conflict free).
__device__ __forceinline__ float lo_of(float x) {
    const float r = x - __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);      // exact, |r| < 2^-10 |x|
    return __uint_as_float((__float_as_uint(r) + 0x1000u) & 0xFFFFE000u);       // to the TF32 grid, half-up
}
__device__ __forceinline__ float4 lo4(const float4 &v) {
    return make_float4(lo_of(v.x), lo_of(v.y), lo_of(v.z), lo_of(v.w));
}
// offsets of the 4 chunks thread p owns in a 16 KB operand tile
// K-major tile of 128 rows x 32 floats (SWIZZLE_128B): 8 threads per row, 16 rows per pass.
__device__ __forceinline__ int off_kmajor(int p, int q) {
    const int row = q * (kProdThreads / 8) + (p >> 3), c = p & 7;
    return row * 128 + ((c ^ (row & 7)) << 4);
}
// MN-major tile of 32 k-rows x 128 floats (4 chunks of [32 k][32 floats], SWIZZLE_128B_BASE32B).  Thread p
// owns the 4 columns (p % 32) * 4.. in k-rows p / 32 + 4 q (so its values sum into one float4 of column sums).
__device__ __forceinline__ int off_mnmajor(int p, int q) {
    const int g = p & 31, k = q * kProdWarps + (p >> 5);
    const int mb = g >> 3, c = g & 7;
    return mb * 4096 + k * 128 + ((((c >> 1) ^ (k & 3)) << 5) | ((c & 1) << 4));
}
// One stage: this thread's 8 raw chunks of A, then of B: read, derive lo, store (two batches of 8 loads in flight).
template <bool A_MN, bool B_MN, bool SUM, int BW>
__device__ __forceinline__ void convert_stage(const uint8_t *a_hi, uint8_t *a_lo, const uint8_t *b_hi, uint8_t *b_lo,
                                              int p, float4 &colsum) {
    static_assert(BW == 128, "B tile width");
    int o[kProdQ];
    float4 v[kProdQ];
#pragma unroll
    for (int q = 0; q < kProdQ; ++q) {
        o[q] = A_MN ? off_mnmajor(p, q) : off_kmajor(p, q);
        v[q] = *reinterpret_cast<const float4 *>(a_hi + o[q]);
    }
#pragma unroll
    for (int q = 0; q < kProdQ; ++q) {
        if (SUM) { colsum.x += v[q].x; colsum.y += v[q].y; colsum.z += v[q].z; colsum.w += v[q].w; }
        *reinterpret_cast<float4 *>(a_lo + o[q]) = lo4(v[q]);
    }
#pragma unroll
    for (int q = 0; q < kProdQ; ++q) {
        o[q] = B_MN ? off_mnmajor(p, q) : off_kmajor(p, q);
        v[q] = *reinterpret_cast<const float4 *>(b_hi + o[q]);
    }
#pragma unroll
    for (int q = 0; q < kProdQ; ++q) *reinterpret_cast<float4 *>(b_lo + o[q]) = lo4(v[q]);
}

// ---- BF16 correction operands -----------------------------------------------------------------------------
// The two correction products only have to be good to 2^-9 of THEIR magnitude (2^-11 of the main product's):
//   a.b ~= trunc_tf32(a).trunc_tf32(b)  [kind::tf32 on the raw tile]  +  bf16(a - trunc_tf32(a)).bf16(b)
//        + bf16(a).bf16(b - trunc_tf32(b))                               [kind::f16, twice the TF32 rate],
// worst-case 2^-19, typically 3e-7 relative per product with zero mean -- about fp32's own dot-product error
// (tests/test_gpu_gemm.py holds it to 2e-5 of the result's max-norm at every trunk shape).
// Tiles: K-major [128 rows][32 k] bf16 = 64-byte rows, SWIZZLE_64B (16-byte unit u of row r at u ^ ((r >> 1) & 3));
// MN-major [2 chunks][32 k][64 bf16] = 128-byte rows, SWIZZLE_128B (unit u of k-row k at u ^ (k & 7)).
__device__ __forceinline__ uint32_t pack_bf16(float lo_elem, float hi_elem) {
    uint32_t r;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi_elem), "f"(lo_elem));
    return r;
}
__device__ __forceinline__ float resid_tf32(float x) { return x - __uint_as_float(__float_as_uint(x) & 0xFFFFE000u); }
// byte offset of this thread's 4 consecutive elements (8 bytes) in the bf16 tile, for the raw chunk (p, q)
__device__ __forceinline__ int off16_kmajor(int p, int q) {
    const int row = q * (kProdThreads / 8) + (p >> 3), c = p & 7;
    return row * 64 + ((((c >> 1) ^ (row >> 1)) & 3) << 4) + ((c & 1) << 3);
}
__device__ __forceinline__ int off16_mnmajor(int p, int q) {
    const int g = p & 31, k = q * kProdWarps + (p >> 5);
    const int e = g * 4;                                   // first MN element of the chunk, 0..124
    return (e >> 6) * 4096 + k * 128 + (((((e & 63) >> 3) ^ (k & 7))) << 4) + (((e >> 2) & 1) << 3);
}
template <bool MN, bool SUM>
__device__ __forceinline__ void convert_operand_bf16(const uint8_t *raw, uint8_t *full16, uint8_t *lo16, int p, float4 &colsum) {
    float4 v[kProdQ];
#pragma unroll
    for (int q = 0; q < kProdQ; ++q) v[q] = *reinterpret_cast<const float4 *>(raw + (MN ? off_mnmajor(p, q) : off_kmajor(p, q)));
#pragma unroll
    for (int q = 0; q < kProdQ; ++q) {
        if (SUM) { colsum.x += v[q].x; colsum.y += v[q].y; colsum.z += v[q].z; colsum.w += v[q].w; }
        const int o = MN ? off16_mnmajor(p, q) : off16_kmajor(p, q);
        *reinterpret_cast<uint2 *>(full16 + o) = make_uint2(pack_bf16(v[q].x, v[q].y), pack_bf16(v[q].z, v[q].w));
        *reinterpret_cast<uint2 *>(lo16 + o) = make_uint2(pack_bf16(resid_tf32(v[q].x), resid_tf32(v[q].y)),
                                                          pack_bf16(resid_tf32(v[q].z), resid_tf32(v[q].w)));
    }
}

// ---- FP16 x 3 (engine 5) ----------------------------------------------------------------------------------
// fp16 carries the same 11 significant bits as TF32 at twice the MMA rate, so with x.s = h + l (h = fp16(x.s),
// l = fp16(x.s - h), s a power of two that puts the tensor's largest magnitude in [2^14, 2^15)) the product is
//   a.b = [ h_a.h_b + (h_a.l_b + l_a.h_b) ] / (s_a s_b)               -- three kind::f16 passes instead of TF32 + 2 BF16
// (3 instead of 4 BF16-rate passes).  Dropped: l_a.l_b, 2^-22 relative.  The scale is exact (powers of two) and comes
// from the tensor's max-abs, which the producing kernel's epilogue tracks (SgemmArgs::amax_*): elements below
// 2^-29 of the tensor's maximum go subnormal in h (below 2^-18 in l), i.e. keep an ABSOLUTE accuracy of 2^-40 of the
// maximum -- sixteen binary orders below the fp32 rounding of the large elements of the same tensor.
__host__ __device__ constexpr uint32_t make_idesc_f16(int M, int N, bool a_mn_major, bool b_mn_major) {
    return (1u << 4) | ((a_mn_major ? 1u : 0u) << 15) | ((b_mn_major ? 1u : 0u) << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
template <bool MN, bool SUM>
__device__ __forceinline__ void convert_operand_f16(const uint8_t *raw, uint8_t *hi16, uint8_t *lo16, int p, float scale,
                                                    float4 &colsum) {
    float4 v[kProdQ];
#pragma unroll
    for (int q = 0; q < kProdQ; ++q) v[q] = *reinterpret_cast<const float4 *>(raw + (MN ? off_mnmajor(p, q) : off_kmajor(p, q)));
#pragma unroll
    for (int q = 0; q < kProdQ; ++q) {
        if (SUM) { colsum.x += v[q].x; colsum.y += v[q].y; colsum.z += v[q].z; colsum.w += v[q].w; }
        const int o = MN ? off16_mnmajor(p, q) : off16_kmajor(p, q);
        uint32_t h0, l0, h1, l1;
        split_f16x2(v[q].x, v[q].y, scale, h0, l0);
        split_f16x2(v[q].z, v[q].w, scale, h1, l1);
        *reinterpret_cast<uint2 *>(hi16 + o) = make_uint2(h0, h1);
        *reinterpret_cast<uint2 *>(lo16 + o) = make_uint2(l0, l1);
    }
}

// ---- TMA ----------------------------------------------------------------------
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *tm, int c0, int c1, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tm)), "r"(c0), "r"(c1), "r"(bar)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *tm, int c0, int c1, int c2, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tm)), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
        : "memory");
}
// warp-convergent forms for the TMA issuer of the conversion-free kernel (the whole warp runs the loop, one elected lane
// issues; see umma_f16_lohi_elect): inside `if (lane == 0)` every UTMALDG costs an ELECT / 4 x R2UR.BROADCAST /
// BRA.U.ANY loop, and ncu's source view showed the issuing lane busy 72 % of a 978-cycle stage (768 of them tensor work).
// Same-box A/B (profiles/r02_tma_issue_ab.txt): forward GEMMs -2 % cycles, the others unchanged, results bitwise equal.
__device__ __forceinline__ void mbar_expect_tx_elect(uint32_t bar, uint32_t bytes) {
    asm volatile(
        "{\n\t.reg .pred q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n\t}"
        ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_2d_elect(uint32_t dst, const CUtensorMap *tm, int c0, int c1, uint32_t bar) {
    asm volatile(
        "{\n\t.reg .pred q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];\n\t}"
        ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tm)), "r"(c0), "r"(c1), "r"(bar)
        : "memory");
}
__device__ __forceinline__ void tma_load_3d_elect(uint32_t dst, const CUtensorMap *tm, int c0, int c1, int c2, uint32_t bar) {
    asm volatile(
        "{\n\t.reg .pred q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];\n\t}"
        ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tm)), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
        : "memory");
}
#ifdef OCPPO_TMA_LANE0
constexpr bool kTmaConv = false;      // the single-lane issuer this replaced (A/B: profiles/r02_tma_issue_ab.txt)
#else
constexpr bool kTmaConv = true;
#endif

// epilogue side: a finished 32 x 32 chunk leaves through a swizzled 4 KB smem tile and one bulk tensor store
__device__ __forceinline__ void tma_store_2d(const CUtensorMap *tm, uint32_t src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(reinterpret_cast<uint64_t>(tm)), "r"(src), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void tma_prefetch_2d(const CUtensorMap *tm, int c0, int c1) {
    asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global [%0, {%1, %2}];"
                 ::"l"(reinterpret_cast<uint64_t>(tm)), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

struct TileCoord { int m0, n0, split; int64_t kbeg, kend; int nk; };

__device__ __forceinline__ TileCoord tile_coord(const SgemmArgs &p, int tile, int nt_n, int nt_m, int tn) {
    TileCoord t;
    const int nb = tile % nt_n;
    const int rest = tile / nt_n;
    const int mb = rest % nt_m;
    t.split = rest / nt_m;
    t.m0 = mb * kPairM;
    t.n0 = nb * tn;
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

#define OCPPO_TMEM_LD16(v, taddr)                                                                                   \
    asm volatile(                                                                                                   \
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "                                                                   \
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"                            \
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),           \
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])      \
        : "r"(taddr)                                                                                                \
        : "memory")

template <int MODE, bool SPLIT, int CORR, int TN_ = 256>      // CORR: 0 = TF32 corrections, 1 = BF16 corrections, 2 = FP16 x 3 (no TF32 pass), 3 = FP16 x 3 on pair storage
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(CORR == 3 ? kProdTid0 : kTcThreads, 1) tc_gemm_kernel(SgemmArgs p, int nt_n, int nt_m, int ntiles, const __grid_constant__ CUtensorMap tmA,
                                                                                         const __grid_constant__ CUtensorMap tmB,
                                                                                         const __grid_constant__ CUtensorMap tmC,
                                                                                         const __grid_constant__ CUtensorMap tmX,
                                                                                         const __grid_constant__ CUtensorMap tmB2,
                                                                                         const __grid_constant__ CUtensorMap tmA2,
                                                                                         const __grid_constant__ CUtensorMap tmC2) {
    constexpr bool A_MN = (MODE == SGEMM_WGRAD);
    constexpr bool B_MN = (MODE != SGEMM_FWD);
    constexpr bool CORR16 = CORR >= 1, F16 = CORR >= 2, PRE = CORR == 3;   // PRE: both operands arrive as fp16 hi / lo pairs
    typedef Tile<SPLIT, TN_> TileT;
    static_assert(TN_ == 256 || (TN_ == 128 && CORR == 3), "the 128-column tile exists for the conversion-free kernel only");
    constexpr int TN = TileT::TN, TBW = TileT::TBW, kStageBytesB = TileT::kStageBytesB;
    constexpr int kStageBytes = TileT::kStageBytes, kNumStages = TileT::kNumStages;
    constexpr int kAccStride = TileT::kAccStride, kAccBufs = TileT::kAccBufs;
    // MMA instructions chained into the (main) accumulator per 32-deep stage: 4 main products, plus the corrections when
    // they share the accumulator (8 TF32 or 4 BF16; a K = 16 BF16 instruction loses about 1.5x what a K = 8 TF32 one does)
    // FP16 x 3: calibrated on zero-mean products (tools/acc_sweep.sh) -- every GEMM of the trunk multiplies a signed operand,
    // so the running sum is a random walk and loses ~0.6 of what a linearly growing sum would: 2 K = 16 instructions per
    // stage in the main accumulator (split), 6 when the corrections share it
    constexpr float kMmaPerStage = F16 ? (SPLIT ? 1.9f : 5.6f) : (SPLIT ? 4.f : (CORR16 ? 10.f : 12.f));
    // stage layout.  3xTF32: [A raw = hi 16K][A lo 16K][B raw = hi 16K][B lo 16K];
    // CORR16:  [A raw 16K][B raw 16K][A bf16 8K][A lo bf16 8K][B bf16 8K][B lo bf16 8K]
    constexpr int kOffA = 0, kOffB = CORR16 ? kStageBytesA : 2 * kStageBytesA;
    constexpr int kOffALo_ = CORR16 ? 2 * kStageBytesA + kStageBytesA / 2 : kStageBytesA;
    constexpr int kOffBLo_ = CORR16 ? 3 * kStageBytesA + kStageBytesB / 2 : 2 * kStageBytesA + kStageBytesB;
    constexpr int kOffA16_ = 2 * kStageBytesA, kOffB16_ = 3 * kStageBytesA;     // CORR16 only
    // FP16 x 3 with the weights pre-split in global memory (FWD / DGRAD of the learner): nothing ever lands in the raw-B
    // slot, so the stage shrinks to [A raw 16K][A hi 8K][A lo 8K][B hi 8K][B lo 8K] = 48 KB and the ring holds FOUR stages
    // PRE: both operands live in global memory as fp16 pairs: [A hi 8K][A lo 8K][B hi 8K][B lo 8K] = 32 KB, SIX stages,
    // no conversion warps at all (the kernel is launched with the first ten warps only)
    const bool compact = !PRE && F16 && MODE != SGEMM_WGRAD && p.b16_hi != nullptr && !(p.dbg & 64);
    const int stage_bytes = PRE ? 32 * 1024 : (compact ? 48 * 1024 : kStageBytes);
    // epi2 (conversion-free kernel): each epilogue warp owns TWO staging tiles, so a chunk's bulk store overlaps the next
    // chunk's arithmetic instead of being waited for (ncu: the K <= 512 layers ran at 52-75 % tensor-pipe activity, their
    // 17 k-cycle epilogue longer than the 12 k-cycle mainloop it should hide behind); the 32 KB come out of the stage ring:
    // five 32 KB stages instead of six
    const bool epi2 = PRE && p.epi2 != 0;
    const int ring_bytes = epi2 ? kRingBytes - kEpiPadBytes : kRingBytes;
    const int nstages = PRE ? ((p.dbg & 128) ? 3 : (epi2 ? 5 : 6)) : (compact ? 4 : kNumStages);
    const int kOffA16 = PRE ? 0 : (compact ? kStageBytesA : kOffA16_);
    const int kOffALo = PRE ? kStageBytesA / 2 : (compact ? kStageBytesA + kStageBytesA / 2 : kOffALo_);
    const int kOffB16 = PRE ? kStageBytesA : (compact ? 2 * kStageBytesA : kOffB16_);
    const int kOffBLo = PRE ? kStageBytesA + kStageBytesB / 2 : (compact ? 2 * kStageBytesA + kStageBytesB / 2 : kOffBLo_);

    static_assert(kNumStages * kStageBytes <= kRingBytes && kNumStages <= kMaxStages, "stage ring");
    // The TMEM accumulator adds with truncation (measured: error grows linearly with the number of MMAs
    // chained into one accumulator), so WGRAD, whose contraction runs over >= 10^4 rows, is cut into
    // 1024-row segments; each is drained by the epilogue and summed in fp32 round-to-nearest in fixed order.
    constexpr int kSegStages = (MODE == SGEMM_WGRAD) ? 128 : (1 << 30);
    extern __shared__ __align__(1024) uint8_t smem[];       // SWIZZLE_128B tiles need 1024-byte aligned bases
    if ((smem_u32(smem) & 1023u) != 0) __trap();
    float *epi_pad = reinterpret_cast<float *>(smem + ring_bytes);
    float *bias_red = reinterpret_cast<float *>(smem + kRingBytes + kEpiPadBytes);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + kRingBytes + kEpiPadBytes + kBiasRedBytes);
    // bars[0..5] full (used in the leader CTA), [6..11] empty, [12..13] tmem_full, [14..15] tmem_empty (leader),
    // [16..21] tma (tiles landed, per CTA), [22..29] aux chunk landed (one per epilogue warp),
    // [30..35] raw_full (leader: the PEER's tiles landed; the leader polls its own tma barrier)
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 36);
    const uint32_t bar_rawfull = smem_u32(&bars[30]);
    const uint32_t bar_aux = smem_u32(&bars[22]);
    const uint32_t bar_tma = smem_u32(&bars[16]);
    const uint32_t bar_full = smem_u32(&bars[0]), bar_empty = smem_u32(&bars[6]);
    const uint32_t bar_tfull = smem_u32(&bars[12]), bar_tempty = smem_u32(&bars[14]);
    const uint32_t cta_rank = cluster_ctarank();             // 0 = leader of the pair
    const int ncl = gridDim.x >> 1, cl = blockIdx.x >> 1;    // clusters in the grid, this cluster

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    float scale_a = 1.f, scale_b = 1.f, inv_a = 1.f, inv_b = 1.f;      // FP16 x 3: per-tensor power-of-two scales
    if (F16) {
        f16_scale_of(__ldg(p.amax_a), scale_a, inv_a);
        f16_scale_of(__ldg(p.amax_b), scale_b, inv_b);
    }
    // fp16-pair output (FWD / DGRAD): the scale must be fixed before the first element exists, so it comes from a rigorous
    // bound, max|A| * (largest row / column L1 norm of the weights) + max|bias| -- typically 2^3 .. 2^6 above the true
    // maximum, which costs nothing that matters: elements keep an absolute accuracy of 2^-40 of the BOUND
    const bool out16 = F16 && MODE != SGEMM_WGRAD && p.c16_hi != nullptr;
    float scale_c = 1.f;
    if (out16) {
        const float bound = __ldg(p.bnd_amax) * __ldg(p.bnd_wl1) + (p.bnd_bmax != nullptr ? __ldg(p.bnd_bmax) : 0.f);
        float inv_c;
        f16_scale_of(bound, scale_c, inv_c);
        if (blockIdx.x == 0 && tid == 0) *p.c_scale_src = bound;       // what the consumers derive the same scale from
    }
    // The TMEM accumulator adds with round-toward-zero (tools/tc_probe.cu Q4): every MMA chained into it costs
    // about half an ulp of bias, and two of the three passes only add tiny correction products.  With SPLIT
    // those go to a second accumulator (the tile's columns [TN, 2 TN)) that the epilogue adds in fp32
    // round-to-nearest: 3x less rounding error.
    constexpr bool split_acc = SPLIT;

    if (tid == 0) {
        for (int s = 0; s < kMaxStages; ++s) {
            mbar_init(bar_full + 8 * s, 2 * kProdWarps);     // both CTAs' producer warps
            mbar_init(bar_empty + 8 * s, 1);
            mbar_init(bar_tma + 8 * s, 1);
            mbar_init(bar_rawfull + 8 * s, (F16 && !PRE) ? kProdWarps : 1);   // the peer's forwarder | F16: raw tile consumed (own producers)
        }
        for (int w = 0; w < kEpiWarps; ++w) mbar_init(bar_aux + 8 * w, 1);
        for (int a = 0; a < 2; ++a) {
            mbar_init(bar_tfull + 8 * a, 1);
            mbar_init(bar_tempty + 8 * a, 2 * kEpiWarps);    // both CTAs' epilogue warps
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == kEpiWarps) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"((uint32_t)kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    cluster_sync_all();          // barriers of BOTH CTAs initialised before anyone arrives remotely
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;

    if (warp < kEpiWarps) {
        // ===================== epilogue =====================
        // Warp w drains TMEM lanes 32 (w % 4).. (its rows) of the column half w / 4, one 32-column chunk at a time:
        // tcgen05.ld (lane = row, 32 consecutive columns in registers), bias/ReLU | ReLU decision bits | k-segment
        // accumulate, 8 x STS.128 into the warp's SWIZZLE_128B tile (conflict-free: the 8 lanes of a quarter-warp hit
        // 8 different 16-byte columns), one bulk tensor store.  Nothing waits for a store; the tile is re-used after
        // cp.async.bulk.wait_group.read.  Two warps share each SM sub-partition and hide each other's latencies.
        // "aux" = fp32 ReLU mask (DGRAD without decision bits) or the running partial (WGRAD k-segments after the
        // first): it arrives by TMA in the same tile, is read to registers, and the result overwrites it.
        const int quarter = warp & 3, chalf = warp >> 2;
        constexpr int kChunks = TN / 64;                     // chunks per warp
        uint8_t *tile0 = reinterpret_cast<uint8_t *>(epi_pad) + warp * (epi2 ? 2 : 1) * kEpiChunkBytes;
        const uint32_t tile0_s = smem_u32(tile0), my_aux_bar = bar_aux + 8 * warp;
        uint32_t ec = 0;                      // chunks staged so far (epi2: staging tile = ec & 1)
        bool prev_aux = false;
        const int swz = lane & 7;
        uint32_t aux_phase = 0;
        int ai = 0;                           // accumulator passes handled so far (tile x k-segment)
        float run_max = 0.f;                  // max |output| of this lane's rows (p.amax_out: the consumer's FP16 scale)
        for (int tile = cl; tile < ntiles; tile += ncl) {
          const TileCoord t = tile_coord(p, tile, nt_n, nt_m, TN);
          for (int kb0 = 0; kb0 < t.nk; kb0 += kSegStages, ++ai) {
            const int nbuf = (p.dbg & 4) ? 1 : kAccBufs, acc = ai % nbuf;
            const bool rmw = (MODE == SGEMM_WGRAD) && kb0 > 0;     // later k-segments add onto the first
            const int seg_stages = t.nk - kb0 < kSegStages ? t.nk - kb0 : kSegStages;
            const float trunc_comp = (p.dbg & 8) ? 0.f : kTruncPerMma * p.trunc_scale * ((float)seg_stages * kMmaPerStage);
            const bool bits = (MODE == SGEMM_DGRAD) && p.relu_bits != nullptr;   // ReLU decisions as packed words
            const bool aux = ((MODE == SGEMM_DGRAD) && !bits) || rmw;
            const int m_loc = t.m0 + (int)cta_rank * TM + quarter * 32;     // first of this warp's 32 rows
            const int n_loc = t.n0 + chalf * (TN / 2);                      // first of this warp's columns
            // WGRAD partial tiles are stacked [split][M][N] in one map: rows past M belong to the next split
            const bool live = (MODE != SGEMM_WGRAD) || m_loc < p.M;
            const int row0 = (MODE == SGEMM_WGRAD ? t.split * p.M : 0) + m_loc;
            const CUtensorMap *tmAux = (MODE == SGEMM_DGRAD) ? &tmX : &tmC;
            if (aux && live && lane == 0) {
                if (rmw) {
                    bulk_wait_all0();          // this warp's stores of the previous k-segment have landed
                    asm volatile("fence.proxy.async;" ::: "memory");
                } else {
                    bulk_wait_read0();
#pragma unroll
                    for (int cc = 1; cc < kChunks; ++cc) tma_prefetch_2d(tmAux, n_loc + cc * 32, row0);
                }
                mbar_expect_tx(my_aux_bar, kEpiChunkBytes);
                tma_load_2d(tile0_s, tmAux, n_loc, row0, my_aux_bar);
            }
            // aux passes (the landing zone of the TMA load is tile 0) stay single-buffered; coming back from one, the
            // last aux store may still be reading tile 0
            if (epi2 && !aux && prev_aux && lane == 0) bulk_wait_read0();
            prev_aux = aux;
            // packed ReLU decisions of this lane's row: the warp's 128 columns are 4 words = 16 bytes, fetched before
            // the accumulator is complete (DGRAD), or written chunk by chunk (FWD)
            uint4 rb = make_uint4(0u, 0u, 0u, 0u);
            const int64_t rb_off = (int64_t)(m_loc + lane) * (p.N >> 5) + (n_loc >> 5);
            if (bits && m_loc + lane < p.M) {
                if (TN == 256) rb = __ldg(reinterpret_cast<const uint4 *>(p.relu_bits + rb_off));
                else { const uint2 r2 = __ldg(reinterpret_cast<const uint2 *>(p.relu_bits + rb_off)); rb.x = r2.x; rb.y = r2.y; }
            }
            if (F16) mbar_wait_sleep(bar_tfull + 8 * acc, (uint32_t)((ai / nbuf) & 1), 200);
            else mbar_wait(bar_tfull + 8 * acc, (uint32_t)((ai / nbuf) & 1));
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (!live) {
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive_on(bar_tempty + 8 * acc, 0);
                continue;
            }
#pragma unroll 1
            for (int cc = 0; cc < kChunks; ++cc) {
                uint32_t v[32];
                const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) +
                                       (uint32_t)(acc * kAccStride + chalf * (TN / 2) + cc * 32);
                OCPPO_TMEM_LD32(v, taddr);
                uint32_t w[32];
                if (split_acc) OCPPO_TMEM_LD32(w, taddr + TN);      // main in columns [0,TN), corrections in [TN,2TN)
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (cc == kChunks - 1) {       // accumulator drained: hand it back to the MMA warp
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) mbar_arrive_on(bar_tempty + 8 * acc, 0);
                }
                // Undo the accumulator's expected round-toward-zero loss (see kTruncPerMma), then add the corrections.
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    float x = __uint_as_float(v[j]);
                    x = fmaf(x, trunc_comp, x);
                    if (F16) x = ((split_acc ? x + __uint_as_float(w[j]) : x) * inv_a) * inv_b;
                    else if (split_acc) x += __uint_as_float(w[j]);
                    v[j] = __float_as_uint(x);
                }
                // (the aux tile is read only now, when the correction registers are dead: 96 registers per thread)
                const uint32_t bsel = (epi2 && !aux) ? (ec & 1u) : 0u;
                ++ec;
                uint8_t *out_tile = tile0 + bsel * kEpiChunkBytes;
                const uint32_t out_s = tile0_s + bsel * kEpiChunkBytes;
                float4 ax[8];
                if (aux) {
                    mbar_wait(my_aux_bar, aux_phase);
                    aux_phase ^= 1;
#pragma unroll
                    for (int c = 0; c < 8; ++c)
                        ax[c] = *reinterpret_cast<const float4 *>(out_tile + lane * 128 + ((c ^ swz) << 4));
                } else if (lane == 0) {
                    if (epi2) bulk_wait_read1();   // the store issued two chunks ago has read THIS tile; the last one may still be reading the other
                    else bulk_wait_read0();        // the previous chunk's store has read the tile
                }
                __syncwarp();
                const int n = n_loc + cc * 32;
                uint32_t word = 0;
                if (bits) word = (cc & 2) ? ((cc & 1) ? rb.w : rb.z) : ((cc & 1) ? rb.y : rb.x);
                // each lane reads (aux) and writes only its own row of the tile, and every value written depends on
                // the aux value read from the same place, so no lane can overwrite data that is still to be read
                // (fp16-pair output: rows move, but every lane's aux row is in registers before the __syncwarp above)
                uint32_t ph0 = 0, ph1 = 0, pl0 = 0, pl1 = 0;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    float4 o = make_float4(__uint_as_float(v[4 * c]), __uint_as_float(v[4 * c + 1]),
                                           __uint_as_float(v[4 * c + 2]), __uint_as_float(v[4 * c + 3]));
                    if (MODE == SGEMM_FWD) {
                        const float4 bb = __ldg(reinterpret_cast<const float4 *>(p.bias + n) + c);
                        o.x = fmaxf(o.x + bb.x, 0.f); o.y = fmaxf(o.y + bb.y, 0.f);
                        o.z = fmaxf(o.z + bb.z, 0.f); o.w = fmaxf(o.w + bb.w, 0.f);
                        if (p.relu_bits != nullptr && !(p.dbg & 4096))      // (4096: timing experiment without the decision bits)
                            word |= ((o.x > 0.f ? 1u : 0u) | (o.y > 0.f ? 2u : 0u) | (o.z > 0.f ? 4u : 0u) | (o.w > 0.f ? 8u : 0u)) << (4 * c);
                    } else if (MODE == SGEMM_DGRAD) {
                        if (bits) {
                            const uint32_t nib = word >> (4 * c);
                            o.x = (nib & 1u) ? o.x : 0.f; o.y = (nib & 2u) ? o.y : 0.f;
                            o.z = (nib & 4u) ? o.z : 0.f; o.w = (nib & 8u) ? o.w : 0.f;
                        } else {
                            o.x = ax[c].x > 0.f ? o.x : 0.f; o.y = ax[c].y > 0.f ? o.y : 0.f;
                            o.z = ax[c].z > 0.f ? o.z : 0.f; o.w = ax[c].w > 0.f ? o.w : 0.f;
                        }
                    } else if (rmw) {
                        o.x += ax[c].x; o.y += ax[c].y; o.z += ax[c].z; o.w += ax[c].w;
                    }
                    if (PRE && MODE == SGEMM_DGRAD) {          // kept for the column sums below
                        v[4 * c] = __float_as_uint(o.x); v[4 * c + 1] = __float_as_uint(o.y);
                        v[4 * c + 2] = __float_as_uint(o.z); v[4 * c + 3] = __float_as_uint(o.w);
                    }
                    if (out16) {
                        // fp16 pair: two [32 rows][32 fp16] tiles (hi at 0, lo at 2 KB), 64-byte rows, SWIZZLE_64B; one
                        // 16-byte unit = 8 columns = two of these float4 groups (conflict-free per quarter-warp)
                        uint32_t h0, l0, h1, l1;
                        split_f16x2(o.x, o.y, scale_c, h0, l0);
                        split_f16x2(o.z, o.w, scale_c, h1, l1);
                        if ((c & 1) == 0) { ph0 = h0; ph1 = h1; pl0 = l0; pl1 = l1; }
                        else {
                            const int off16 = lane * 64 + ((((c >> 1) ^ (lane >> 1)) & 3) << 4);
                            *reinterpret_cast<uint4 *>(out_tile + off16) = make_uint4(ph0, ph1, h0, h1);
                            *reinterpret_cast<uint4 *>(out_tile + 2048 + off16) = make_uint4(pl0, pl1, l0, l1);
                        }
                    } else
                    *reinterpret_cast<float4 *>(out_tile + lane * 128 + ((c ^ swz) << 4)) = o;
                    if (MODE != SGEMM_WGRAD && m_loc + lane < p.M) run_max = fmaxf(fmaxf(run_max, fmaxf(fabsf(o.x), fabsf(o.y))), fmaxf(fabsf(o.z), fabsf(o.w)));
                }
                if (MODE == SGEMM_FWD && p.relu_bits != nullptr && !(p.dbg & 4096)) {
                    // the row's decision words of this warp's columns leave as ONE 16-byte (TN = 128: 8-byte) store per tile
                    // instead of one 4-byte store per chunk: a quarter of the partial-sector writes
                    if (cc == 0) rb.x = word; else if (cc == 1) rb.y = word; else if (cc == 2) rb.z = word; else rb.w = word;
                    if (cc == kChunks - 1 && m_loc + lane < p.M) {
                        if (TN == 256) *reinterpret_cast<uint4 *>(p.relu_bits + rb_off) = rb;
                        else *reinterpret_cast<uint2 *>(p.relu_bits + rb_off) = make_uint2(rb.x, rb.y);
                    }
                }
                if (PRE && MODE == SGEMM_DGRAD && p.colsum_part != nullptr) {
                    // bias gradient of the layer below = column sums of this output: the warp's 32 rows fold in a butterfly
                    // (31 shuffles; lane j ends with column j), then onto the warp's PRIVATE partial row in global memory
                    // (rows past M are exact zeros).  Fixed order everywhere: bitwise repeatable.
#pragma unroll
                    for (int off = 16; off >= 1; off >>= 1) {
                        const bool up = (lane & off) != 0;
#pragma unroll
                        for (int j = 0; j < off; ++j) {
                            const float lo_v = __uint_as_float(v[j]), hi_v = __uint_as_float(v[j + off]);
                            const float got = __shfl_xor_sync(0xffffffffu, up ? lo_v : hi_v, off);
                            v[j] = __float_as_uint((up ? hi_v : lo_v) + got);
                        }
                    }
                    float *cp = p.colsum_part + ((int64_t)blockIdx.x * 4 + quarter) * p.N + n + lane;
                    *cp += __uint_as_float(v[0]);
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) {
                    if (!(p.dbg & 2)) {
                        // (8192: timing experiment, every store lands in the tensor's first 4096 rows (L2-resident) -- the
                        // store engine's shared-memory reads without the HBM write traffic; profiles/r02_store_experiments.txt)
                        const int srow = (p.dbg & 8192) ? (row0 & 4095) : row0;
                        tma_store_2d(&tmC, out_s, n, srow);
                        if (out16) tma_store_2d(&tmC2, out_s + 2048, n, srow);
                        bulk_commit();
                    }
                    if (aux && cc + 1 < kChunks) {           // next aux chunk, once the store has read the tile
                        bulk_wait_read0();
                        mbar_expect_tx(my_aux_bar, kEpiChunkBytes);
                        tma_load_2d(out_s, tmAux, n + 32, row0, my_aux_bar);
                    }
                }
            }
          }
        }
        if (lane == 0) bulk_wait_all0();       // all stores of this warp complete before the CTA may exit
        if (MODE != SGEMM_WGRAD && p.amax_out != nullptr) {
            const uint32_t m = __reduce_max_sync(0xffffffffu, __float_as_uint(run_max));   // non-negative floats order as uints
            if (lane == 0) atomicMax(reinterpret_cast<unsigned int *>(p.amax_out), m);
        }
    } else if (warp < kEpiWarps + 2) {
        // ===================== MMA issuer (warp 8 of the leader; the peer's warp 8 forwards) =====================
        if (warp == kEpiWarps && cta_rank == 0 && (PRE || lane == 0)) {      // PRE: the whole warp, convergent (see umma_f16_lohi_elect)
            constexpr uint32_t idesc = make_idesc(kPairM, TN, A_MN, B_MN);   // M = 256 across the pair
            // K-major: SBO = 8-row group stride (1024 B), LBO unused.  MN-major: LBO = stride between
            // 32-float MN chunks (4096 B), SBO = stride between 4-row k groups (512 B).
            constexpr uint32_t a_lbo = A_MN ? 4096 : 16, a_sbo = A_MN ? 512 : 1024, a_ty = A_MN ? 1 : 2;
            constexpr uint32_t b_lbo = B_MN ? 4096 : 16, b_sbo = B_MN ? 512 : 1024, b_ty = B_MN ? 1 : 2;
            // per k-step of 8: K-major advances 32 B inside the 128 B row, MN-major two 512 B atoms
            constexpr uint64_t a_step = (A_MN ? 1024 : 32) >> 4, b_step = (B_MN ? 1024 : 32) >> 4;
            int stage = 0, sphase = 0, ai = 0;
            for (int tile = cl; tile < ntiles; tile += ncl) {
              const TileCoord t = tile_coord(p, tile, nt_n, nt_m, TN);
              for (int kb0 = 0; kb0 < t.nk; kb0 += kSegStages, ++ai) {
                const int nbuf = (p.dbg & 4) ? 1 : kAccBufs, acc = ai % nbuf;
                mbar_wait(bar_tempty + 8 * acc, (uint32_t)(((ai / nbuf) & 1) ^ 1));
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * kAccStride);
                const int kb1 = (t.nk - kb0 < kSegStages) ? t.nk : kb0 + kSegStages;
                for (int kb = kb0; kb < kb1; ++kb) {
                    // The main product reads only the raw tiles: it starts as soon as both CTAs' TMA loads have landed
                    // and runs while the producers are still deriving the correction operands of the same stage.
                    const int s = stage, fphase = sphase;
                    if (!F16) {
                        mbar_wait(bar_tma + 8 * s, (uint32_t)fphase);        // own raw tiles
                        mbar_wait(bar_rawfull + 8 * s, (uint32_t)fphase);    // the peer's
                    }
                    if (++stage == nstages) { stage = 0; sphase ^= 1; }
                    if constexpr (PRE) {
                        // conversion-free kernel: lean issue path.  Descriptor HIGH words are constants of the operand
                        // layout; LOW words = (address >> 4) | lbo << 16 with address = ring base + s * 32 KB + tile offset.
                        constexpr uint32_t a16_lbo = A_MN ? 4096 : 16, a16_sbo = A_MN ? 1024 : 512, a16_ty = A_MN ? 2 : 4;
                        constexpr uint32_t b16_lbo = B_MN ? 4096 : 16, b16_sbo = B_MN ? 1024 : 512, b16_ty = B_MN ? 2 : 4;
                        constexpr uint32_t a_step = (A_MN ? 2048 : 32) >> 4, b_step = (B_MN ? 2048 : 32) >> 4;
                        constexpr uint32_t a_hiw = (uint32_t)(make_desc_const(a16_lbo, a16_sbo, a16_ty) >> 32);
                        constexpr uint32_t b_hiw = (uint32_t)(make_desc_const(b16_lbo, b16_sbo, b16_ty) >> 32);
                        constexpr uint32_t a_low = (uint32_t)make_desc_const(a16_lbo, a16_sbo, a16_ty);
                        constexpr uint32_t b_low = (uint32_t)make_desc_const(b16_lbo, b16_sbo, b16_ty);
                        constexpr uint32_t idescf = make_idesc_f16(kPairM, TN, A_MN, B_MN);
                        const uint32_t sa = (smem_u32(smem) >> 4) + (uint32_t)s * (32768u >> 4);      // stage base, 16-byte units
                        const uint32_t ah = a_low + sa, al = a_low + sa + ((kStageBytesA / 2) >> 4);
                        const uint32_t bh = b_low + sa + (kStageBytesA >> 4), bl = b_low + sa + ((kStageBytesA + kStageBytesB / 2) >> 4);
                        mbar_wait_lean(bar_tma + 8 * s, (uint32_t)fphase);        // own fp16 tiles (TMA)
                        mbar_wait_lean(bar_rawfull + 8 * s, (uint32_t)fphase);    // the peer's
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        const uint32_t d_corrf = split_acc ? d_tmem + TN : d_tmem;
                        const uint32_t first = kb > kb0 ? 1u : 0u;
                        umma_f16_lohi_elect(d_tmem, ah, a_hiw, bh, b_hiw, idescf, first);
                        umma_f16_lohi_elect(d_tmem, ah + a_step, a_hiw, bh + b_step, b_hiw, idescf, 1u);
                        if (!(p.dbg & 32)) {       // (bring-up timing experiment: main product only)
                            umma_f16_lohi_elect(d_corrf, al, a_hiw, bh, b_hiw, idescf, split_acc ? first : 1u);
                            umma_f16_lohi_elect(d_corrf, ah, a_hiw, bl, b_hiw, idescf, 1u);
                            umma_f16_lohi_elect(d_corrf, al + a_step, a_hiw, bh + b_step, b_hiw, idescf, 1u);
                            umma_f16_lohi_elect(d_corrf, ah + a_step, a_hiw, bl + b_step, b_hiw, idescf, 1u);
                        }
                        umma_commit_pair_elect(bar_empty + 8 * s);
                        continue;
                    }
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    uint8_t *st = smem + s * stage_bytes;
                    const uint32_t a_hi = smem_u32(st) + kOffA, a_lo = smem_u32(st) + kOffALo;
                    const uint32_t b_hi = smem_u32(st) + kOffB, b_lo = smem_u32(st) + kOffBLo;
                    const uint64_t da_hi = make_desc(a_hi, a_lbo, a_sbo, a_ty), db_hi = make_desc(b_hi, b_lbo, b_sbo, b_ty);
                    if (CORR16) {
                        // bf16 tiles: K-major SWIZZLE_64B (8-row groups 512 B apart, k-step of 16 = 32 B inside the row);
                        // MN-major SWIZZLE_128B (64-element chunks 4096 B apart, 8-k-row groups 1024 B, k-step = 2048 B)
                        constexpr uint32_t a16_lbo = A_MN ? 4096 : 16, a16_sbo = A_MN ? 1024 : 512, a16_ty = A_MN ? 2 : 4;
                        constexpr uint32_t b16_lbo = B_MN ? 4096 : 16, b16_sbo = B_MN ? 1024 : 512, b16_ty = B_MN ? 2 : 4;
                        constexpr uint64_t a16_step = (A_MN ? 2048 : 32) >> 4, b16_step = (B_MN ? 2048 : 32) >> 4;
                        constexpr uint32_t idesc16 = make_idesc_bf16(kPairM, TN, A_MN, B_MN);
                        const uint64_t da16 = make_desc(smem_u32(st) + kOffA16, a16_lbo, a16_sbo, a16_ty);
                        const uint64_t db16 = make_desc(smem_u32(st) + kOffB16, b16_lbo, b16_sbo, b16_ty);
                        const uint64_t da16lo = make_desc(a_lo, a16_lbo, a16_sbo, a16_ty);
                        const uint64_t db16lo = make_desc(b_lo, b16_lbo, b16_sbo, b16_ty);
                        if (F16) {
                            // all three products need the producers' tiles: h.h into the main accumulator, h.l + l.h
                            // (2^-11 of it, applied by the epilogue) into the correction accumulator
                            constexpr uint32_t idescf = make_idesc_f16(kPairM, TN, A_MN, B_MN);
                            if (PRE) {
                                mbar_wait(bar_tma + 8 * s, (uint32_t)fphase);        // own fp16 tiles (TMA)
                                mbar_wait(bar_rawfull + 8 * s, (uint32_t)fphase);    // the peer's
                            } else {
                                mbar_wait(bar_full + 8 * s, (uint32_t)fphase);
                            }
                            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                            for (int k2 = 0; k2 < TK / 16; ++k2)
                                umma_bf16(d_tmem, da16 + k2 * a16_step, db16 + k2 * b16_step, idescf, (kb > kb0 || k2 > 0) ? 1u : 0u);
                            if (!(p.dbg & 32))       // timing experiment: main product only
#pragma unroll
                            for (int k2 = 0; k2 < TK / 16; ++k2) {
                                const uint32_t d_corrf = split_acc ? d_tmem + TN : d_tmem;
                                umma_bf16(d_corrf, da16lo + k2 * a16_step, db16 + k2 * b16_step, idescf, (!split_acc || kb > kb0 || k2 > 0) ? 1u : 0u);
                                umma_bf16(d_corrf, da16 + k2 * a16_step, db16lo + k2 * b16_step, idescf, 1u);
                            }
                            umma_commit_pair(bar_empty + 8 * s);
                            continue;
                        }
#pragma unroll
                        for (int ks = 0; ks < TK / 8; ++ks)
                            umma_tf32(d_tmem, da_hi + ks * a_step, db_hi + ks * b_step, idesc, (kb > kb0 || ks > 0) ? 1u : 0u);
                        mbar_wait(bar_full + 8 * s, (uint32_t)fphase);      // the producers' bf16 tiles
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                        for (int k2 = 0; k2 < TK / 16; ++k2) {
                            const uint32_t d_corr16 = split_acc ? d_tmem + TN : d_tmem;
                            umma_bf16(d_corr16, da16lo + k2 * a16_step, db16 + k2 * b16_step, idesc16, (!split_acc || kb > kb0 || k2 > 0) ? 1u : 0u);
                            umma_bf16(d_corr16, da16 + k2 * a16_step, db16lo + k2 * b16_step, idesc16, 1u);
                        }
                        umma_commit_pair(bar_empty + 8 * s);
                        continue;
                    }
                    const uint64_t da_lo = make_desc(a_lo, a_lbo, a_sbo, a_ty), db_lo = make_desc(b_lo, b_lbo, b_sbo, b_ty);
#pragma unroll
                    for (int ks = 0; ks < TK / 8; ++ks)
                        umma_tf32(d_tmem, da_hi + ks * a_step, db_hi + ks * b_step, idesc, (kb > kb0 || ks > 0) ? 1u : 0u);
                    mbar_wait(bar_full + 8 * s, (uint32_t)fphase);          // the producers' lo tiles
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t d_corr = split_acc ? d_tmem + TN : d_tmem;
#pragma unroll
                    for (int ks = 0; ks < TK / 8; ++ks) {
                        umma_tf32(d_corr, da_lo + ks * a_step, db_hi + ks * b_step, idesc, (!split_acc || kb > kb0 || ks > 0) ? 1u : 0u);
                        umma_tf32(d_corr, da_hi + ks * a_step, db_lo + ks * b_step, idesc, 1u);
                    }
                    umma_commit_pair(bar_empty + 8 * s);     // stage free (in both CTAs) once these MMAs have read it
                }
                if constexpr (PRE) umma_commit_pair_elect(bar_tfull + 8 * acc);
                else umma_commit_pair(bar_tfull + 8 * acc);       // accumulator (k-segment) complete in both CTAs
              }
            }
        }
        else if (warp == kEpiWarps + 1 && ((PRE && kTmaConv) || lane == 0)) {
            // ===================== TMA issuer (warp 9, both CTAs; PRE + kTmaConv: the whole warp, convergent) =====================
            // raw fp32 tiles of this CTA's A rows / B columns, straight into the hi slots of the stage ring
            const int rank_m = (int)cta_rank * TM, rank_n = (int)cta_rank * TBW;
            int stage = 0, sphase = 0;
            for (int tile = cl; tile < ntiles; tile += ncl) {
                const TileCoord t = tile_coord(p, tile, nt_n, nt_m, TN);
                for (int kb = 0; kb < t.nk; ++kb) {
                    const int s = stage;
                    if (PRE) {
                        // both operands as fp16 hi / lo pairs, straight into the MMA's tiles: four loads per stage
                        const uint32_t st = smem_u32(smem + s * stage_bytes), bar = bar_tma + 8 * s;
                        const int k0 = (int)(t.kbeg + (int64_t)kb * TK);
                        mbar_wait_lean(bar_empty + 8 * s, (uint32_t)(sphase ^ 1));
                        // (bring-up timing experiments, results are garbage: 512 = no B loads, 1024 = no A loads)
                        const bool skip_b = (p.dbg & 512) != 0, skip_a = (p.dbg & 1024) != 0;
                        if constexpr (kTmaConv) {
                            mbar_expect_tx_elect(bar, kStageBytesA + kStageBytesB);
                            if (A_MN) {
                                tma_load_3d_elect(st + kOffA16, &tmA, 0, k0, (t.m0 + rank_m) >> 6, bar);
                                tma_load_3d_elect(st + kOffALo, &tmA2, 0, k0, (t.m0 + rank_m) >> 6, bar);
                            } else {
                                tma_load_2d_elect(st + kOffA16, &tmA, k0, t.m0 + rank_m, bar);
                                tma_load_2d_elect(st + kOffALo, &tmA2, k0, t.m0 + rank_m, bar);
                            }
                            if (B_MN) {
                                tma_load_3d_elect(st + kOffB16, &tmB, 0, k0, (t.n0 + rank_n) >> 6, bar);
                                tma_load_3d_elect(st + kOffBLo, &tmB2, 0, k0, (t.n0 + rank_n) >> 6, bar);
                            } else {
                                tma_load_2d_elect(st + kOffB16, &tmB, k0, t.n0 + rank_n, bar);
                                tma_load_2d_elect(st + kOffBLo, &tmB2, k0, t.n0 + rank_n, bar);
                            }
                            if (++stage == nstages) { stage = 0; sphase ^= 1; }
                            continue;
                        }
                        mbar_expect_tx(bar, (skip_a ? 0 : kStageBytesA) + (skip_b ? 0 : kStageBytesB));
                        if (skip_a) {
                        } else if (A_MN) {
                            tma_load_3d(st + kOffA16, &tmA, 0, k0, (t.m0 + rank_m) >> 6, bar);
                            tma_load_3d(st + kOffALo, &tmA2, 0, k0, (t.m0 + rank_m) >> 6, bar);
                        } else {
                            tma_load_2d(st + kOffA16, &tmA, k0, t.m0 + rank_m, bar);
                            tma_load_2d(st + kOffALo, &tmA2, k0, t.m0 + rank_m, bar);
                        }
                        if (skip_b) {
                        } else if (B_MN) {
                            tma_load_3d(st + kOffB16, &tmB, 0, k0, (t.n0 + rank_n) >> 6, bar);
                            tma_load_3d(st + kOffBLo, &tmB2, 0, k0, (t.n0 + rank_n) >> 6, bar);
                        } else {
                            tma_load_2d(st + kOffB16, &tmB, k0, t.n0 + rank_n, bar);
                            tma_load_2d(st + kOffBLo, &tmB2, k0, t.n0 + rank_n, bar);
                        }
                        if (++stage == nstages) { stage = 0; sphase ^= 1; }
                        continue;
                    }
                    // F16: nothing but the producers reads the raw tiles, so they are refilled as soon as the conversion
                    // has consumed them -- one MMA stage earlier than the converted tiles come free
                    mbar_wait((F16 ? bar_rawfull : bar_empty) + 8 * s, (uint32_t)(sphase ^ 1));
                    const uint32_t a_hi = smem_u32(smem + s * stage_bytes) + kOffA, b_hi = smem_u32(smem + s * stage_bytes) + kOffB;
                    const uint32_t bar = bar_tma + 8 * s;
                    const int k0 = (int)(t.kbeg + (int64_t)kb * TK);
                    mbar_expect_tx(bar, kStageBytesA + kStageBytesB);
                    if (A_MN) tma_load_3d(a_hi, &tmA, 0, k0, (t.m0 + rank_m) >> 5, bar);
                    else tma_load_2d(a_hi, &tmA, k0, t.m0 + rank_m, bar);
                    if (F16 && MODE != SGEMM_WGRAD && p.b16_hi != nullptr) {
                        // B (the layer's weights) was split into fp16 hi / lo tensors once per call: its tiles arrive in
                        // the MMA's layout (tmB = hi, tmB2 = lo) and wait for the MMAs of the stage's previous use
                        const uint32_t b16h = smem_u32(smem + s * stage_bytes) + kOffB16, b16l = smem_u32(smem + s * stage_bytes) + kOffBLo;
                        mbar_wait(bar_empty + 8 * s, (uint32_t)(sphase ^ 1));
                        if (B_MN) {
                            tma_load_3d(b16h, &tmB, 0, k0, (t.n0 + rank_n) >> 6, bar);
                            tma_load_3d(b16l, &tmB2, 0, k0, (t.n0 + rank_n) >> 6, bar);
                        } else {
                            tma_load_2d(b16h, &tmB, k0, t.n0 + rank_n, bar);
                            tma_load_2d(b16l, &tmB2, k0, t.n0 + rank_n, bar);
                        }
                    } else if (B_MN) tma_load_3d(b_hi, &tmB, 0, k0, (t.n0 + rank_n) >> 5, bar);
                    else tma_load_2d(b_hi, &tmB, k0, t.n0 + rank_n, bar);
                    if (++stage == nstages) { stage = 0; sphase ^= 1; }
                }
            }
        }
        else if ((!F16 || PRE) && warp == kEpiWarps && lane == 0 && cta_rank == 1) {
            // ===================== raw-tile forwarder (warp 8 of the peer CTA) =====================
            // tells the leader's MMA issuer that this CTA's raw tiles of a stage have landed (the leader polls its own)
            int stage = 0, sphase = 0;
            for (int tile = cl; tile < ntiles; tile += ncl) {
                const TileCoord t = tile_coord(p, tile, nt_n, nt_m, TN);
                for (int kb = 0; kb < t.nk; ++kb) {
                    mbar_wait(bar_tma + 8 * stage, (uint32_t)sphase);
                    mbar_arrive_on(bar_rawfull + 8 * stage, 0);
                    if (++stage == nstages) { stage = 0; sphase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if constexpr (!PRE) {
        // ===================== producers =====================
        const int pt = tid - kProdTid0;
        const int pwarp_lane0 = (lane == 0);
        int stage = 0, sphase = 0;
        const int rank_m = (int)cta_rank * TM;
        for (int tile = cl; tile < ntiles; tile += ncl) {
            const TileCoord t = tile_coord(p, tile, nt_n, nt_m, TN);
            const bool do_bias = (MODE == SGEMM_WGRAD) && t.n0 == 0 && p.bias_part != nullptr;
            float4 bsum = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int kb = 0; kb < t.nk; ++kb) {
                const int s = stage;
                mbar_wait(bar_tma + 8 * s, (uint32_t)sphase);           // raw tile landed (TMA complete_tx)
                uint8_t *st = smem + s * stage_bytes;
                uint8_t *a_hi = st + kOffA, *a_lo = st + kOffALo, *b_hi = st + kOffB, *b_lo = st + kOffBLo;
                if (F16) {
                    mbar_wait(bar_empty + 8 * s, (uint32_t)(sphase ^ 1));    // the MMAs of the previous use have read the fp16 tiles
                    if (p.dbg & 16) { /* timing experiment: no conversion */ } else
                    if (A_MN && do_bias) convert_operand_f16<A_MN, true>(a_hi, st + kOffA16, a_lo, pt, scale_a, bsum);
                    else convert_operand_f16<A_MN, false>(a_hi, st + kOffA16, a_lo, pt, scale_a, bsum);
                    if (MODE == SGEMM_WGRAD || p.b16_hi == nullptr) convert_operand_f16<B_MN, false>(b_hi, st + kOffB16, b_lo, pt, scale_b, bsum);
                } else if (CORR16) {
                    if (A_MN && do_bias) convert_operand_bf16<A_MN, true>(a_hi, st + kOffA16, a_lo, pt, bsum);
                    else convert_operand_bf16<A_MN, false>(a_hi, st + kOffA16, a_lo, pt, bsum);
                    convert_operand_bf16<B_MN, false>(b_hi, st + kOffB16, b_lo, pt, bsum);
                } else if (A_MN && do_bias) convert_stage<A_MN, B_MN, true, TBW>(a_hi, a_lo, b_hi, b_lo, pt, bsum);
                else convert_stage<A_MN, B_MN, false, TBW>(a_hi, a_lo, b_hi, b_lo, pt, bsum);
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (pwarp_lane0) {
                    mbar_arrive_on(bar_full + 8 * s, 0);      // the leader CTA's barrier
                    if (F16) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_rawfull + 8 * s) : "memory");
                }
                if (++stage == nstages) { stage = 0; sphase ^= 1; }
            }
            if (MODE == SGEMM_WGRAD) {
                if (do_bias) {           // tile-uniform across the producer warps
                    // thread pt holds columns (pt % 32) * 4.. of k-lane pt / 32: lanes 4-7 park theirs in [4][128 columns],
                    // lanes 0-3 add their own on top, then 128 threads sum the four rows
                    float4 *slot = reinterpret_cast<float4 *>(bias_red + ((pt >> 5) & 3) * TM + (pt & 31) * 4);
                    if (pt >= 128) *slot = bsum;
                    asm volatile("bar.sync 1, %0;" ::"n"(kProdThreads) : "memory");
                    if (pt < 128) {
                        const float4 o = *slot;
                        *slot = make_float4(bsum.x + o.x, bsum.y + o.y, bsum.z + o.z, bsum.w + o.w);
                    }
                    asm volatile("bar.sync 1, %0;" ::"n"(kProdThreads) : "memory");
                    if (pt < TM) {
                        float sum = 0.f;
#pragma unroll
                        for (int kr = 0; kr < 4; ++kr) sum += bias_red[kr * TM + pt];
                        if (t.m0 + rank_m + pt < p.M) p.bias_part[(int64_t)t.split * p.M + t.m0 + rank_m + pt] = sum;
                    }
                    asm volatile("bar.sync 1, %0;" ::"n"(kProdThreads) : "memory");
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    cluster_sync_all();          // neither CTA leaves (or frees TMEM) while its peer can still touch it
    if (warp == kEpiWarps) {
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)kTmemCols)
                     : "memory");
    }
}

inline bool al16(const void *p) { return (((uintptr_t)p) & 15) == 0; }

}  // namespace

// Shapes the tensor-core kernel takes; everything else stays on the FMA kernel.
// N tile of the CTA pair for a GEMM with N output columns: 256, or 128 when N is a multiple of 128 only AND the launch will
// be the conversion-free kernel (both operands as fp16 pairs: present already, or promised by the planner with want_pre)
int tc_tile_n(const SgemmArgs &a) {
    if (a.N % 256 == 0) return 256;
    const bool pre = a.f16x3 && ((a.a16_hi != nullptr && a.b16_hi != nullptr) || a.want_pre);
    return (pre && a.N % 128 == 0) ? 128 : 0;
}

bool tc_gemm_supported(SgemmMode mode, const SgemmArgs &a) {
    if (a.M <= 0 || a.N <= 0 || a.K <= 0) return false;
    if (tc_tile_n(a) == 0) return false;
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

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (fn == nullptr) {
        void *sym = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(sym);
    }
    return fn;
}

// Encoding a tensor map costs a driver call (~1-2 us); a rollout-inference step makes ~30 of them for the same few
// (pointer, shape) combinations, so the encoded maps are memoised.  A map is a pure function of its key.
struct MapKey { int kind; const void *base; int64_t a, b, ld; int box; };
struct MapSlot { MapKey key; CUtensorMap map; bool used; };
constexpr int kMapSlots = 256;
MapSlot g_maps[kMapSlots];
int g_map_next = 0;
bool map_lookup(const MapKey &k, CUtensorMap *tm) {
#ifdef OCPPO_BRINGUP
    static const bool off = getenv("OCPPO_NO_MAPCACHE") != nullptr;     // bisecting knob
    if (off) return false;
#endif
    for (int i = 0; i < kMapSlots; ++i) {
        const MapSlot &m = g_maps[i];
        if (m.used && m.key.kind == k.kind && m.key.base == k.base && m.key.a == k.a && m.key.b == k.b && m.key.ld == k.ld &&
            m.key.box == k.box) { *tm = m.map; return true; }
    }
    return false;
}
void map_store(const MapKey &k, const CUtensorMap &tm) {
    MapSlot &m = g_maps[g_map_next];
    g_map_next = (g_map_next + 1) % kMapSlots;
    m.key = k; m.map = tm; m.used = true;
}

// K-major operand [rows, K] (row stride ld floats): 2-D map {K, rows}, box 32 x box_mn, SWIZZLE_128B.
// MN-major operand [K, cols]: 3-D map {32, K, cols/32}, box 32 x 32 x box_mn/32 -> smem [chunk][k][32 floats],
// SWIZZLE_128B_ATOM_32B (the TMA name of UMMA's SWIZZLE_128B_BASE32B).
int make_map(CUtensorMap *tm, const float *base, bool mn_major, int64_t rows_or_cols, int64_t K, int64_t ld, int box_mn) {
    const MapKey key{0, base, rows_or_cols * 2 + (mn_major ? 1 : 0), K, ld, box_mn};
    if (map_lookup(key, tm)) return OCPPO_OK;
    EncodeTiledFn fn = encode_fn();
    if (fn == nullptr) { set_error("cuTensorMapEncodeTiled entry point not available"); return OCPPO_ECUDA; }
    CUresult r;
    if (!mn_major) {
        const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows_or_cols};
        const cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
        const cuuint32_t box[2] = {32, (cuuint32_t)box_mn}, estr[2] = {1, 1};
        r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else {
        const cuuint64_t dims[3] = {32, (cuuint64_t)K, (cuuint64_t)(rows_or_cols / 32)};
        const cuuint64_t strides[2] = {(cuuint64_t)ld * 4, 128};
        const cuuint32_t box[3] = {32, 32, (cuuint32_t)(box_mn / 32)}, estr[3] = {1, 1, 1};
        r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float *>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    }
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (%d)", (int)r); return OCPPO_ECUDA; }
    map_store(key, *tm);
    return OCPPO_OK;
}

// fp16 operand that is already split and scaled in global memory (same shape / stride as the fp32 one): tiles arrive in
// the layouts the producers write -- K-major [rows][32 k] = 64-byte rows, SWIZZLE_64B; MN-major [chunk][32 k][64], SWIZZLE_128B
int make_map16(CUtensorMap *tm, const void *base, bool mn_major, int64_t rows_or_cols, int64_t K, int64_t ld, int box_mn) {
    const MapKey key{1, base, rows_or_cols * 2 + (mn_major ? 1 : 0), K, ld, box_mn};
    if (map_lookup(key, tm)) return OCPPO_OK;
    EncodeTiledFn fn = encode_fn();
    if (fn == nullptr) { set_error("cuTensorMapEncodeTiled entry point not available"); return OCPPO_ECUDA; }
    CUresult r;
    if (!mn_major) {
        const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows_or_cols};
        const cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
        const cuuint32_t box[2] = {32, (cuuint32_t)box_mn}, estr[2] = {1, 1};
        r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else {
        const cuuint64_t dims[3] = {64, (cuuint64_t)K, (cuuint64_t)(rows_or_cols / 64)};
        const cuuint64_t strides[2] = {(cuuint64_t)ld * 2, 128};
        const cuuint32_t box[3] = {64, 32, (cuuint32_t)(box_mn / 64)}, estr[3] = {1, 1, 1};
        r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void *>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    }
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (fp16 operand) failed (%d)", (int)r); return OCPPO_ECUDA; }
    map_store(key, *tm);
    return OCPPO_OK;
}

// Output (and aux: DGRAD mask, WGRAD running partial) [rows, cols] row-major: 2-D map {cols, rows}, box 32 x 32,
// SWIZZLE_128B -- one epilogue warp's chunk.  Rows past `rows` are clipped on store and zero-filled on load.
int make_map_c(CUtensorMap *tm, const float *base, int64_t rows, int64_t cols, int64_t ld) {
    const MapKey key{2, base, rows, cols, ld, 0};
    if (map_lookup(key, tm)) return OCPPO_OK;
    EncodeTiledFn fn = encode_fn();
    if (fn == nullptr) { set_error("cuTensorMapEncodeTiled entry point not available"); return OCPPO_ECUDA; }
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    const cuuint32_t box[2] = {32, 32}, estr[2] = {1, 1};
    const CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(base), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (output map) failed (%d)", (int)r); return OCPPO_ECUDA; }
    map_store(key, *tm);
    return OCPPO_OK;
}

// fp16 half of an fp16-pair output [rows, cols]: box 32 x 32 = 64-byte rows, SWIZZLE_64B (what the epilogue writes)
int make_map_c16(CUtensorMap *tm, const void *base, int64_t rows, int64_t cols, int64_t ld) {
    const MapKey key{3, base, rows, cols, ld, 0};
    if (map_lookup(key, tm)) return OCPPO_OK;
    EncodeTiledFn fn = encode_fn();
    if (fn == nullptr) { set_error("cuTensorMapEncodeTiled entry point not available"); return OCPPO_ECUDA; }
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
    const cuuint32_t box[2] = {32, 32}, estr[2] = {1, 1};
    const CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (fp16 output map) failed (%d)", (int)r); return OCPPO_ECUDA; }
    map_store(key, *tm);
    return OCPPO_OK;
}

}  // namespace

// Plain (unswizzled) 2-D fp32 tile map over a row-major [rows, cols] tensor, box = box_rows x box_cols, zero fill outside:
// the GAE kernel's loads and stores (gae.cu).  `tm_out` is a CUtensorMap (128 bytes, 64-byte aligned); memoised like the
// GEMM maps.
int tma_map_2d_f32_plain(void *tm_out, const float *base, int64_t rows, int64_t cols, int box_rows, int box_cols) {
    CUtensorMap *tm = reinterpret_cast<CUtensorMap *>(tm_out);
    const MapKey key{4, base, rows, cols, cols, box_rows * 65536 + box_cols};
    if (map_lookup(key, tm)) return OCPPO_OK;
    EncodeTiledFn fn = encode_fn();
    if (fn == nullptr) { set_error("cuTensorMapEncodeTiled entry point not available"); return OCPPO_ECUDA; }
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)cols * 4};
    const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows}, estr[2] = {1, 1};
    const CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float *>(base), dims, strides, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (plain 2-D map) failed (%d)", (int)r); return OCPPO_ECUDA; }
    map_store(key, *tm);
    return OCPPO_OK;
}

int tc_gemm_ctas(SgemmMode mode, const SgemmArgs &a_in, int sm_count) {
    SgemmArgs a = a_in;
    if (!a.f16x3) sgemm_engine_flags(g_gemm_mode, &a);       // planning structs carry no engine flags yet
    const int tn = tc_tile_n(a);
    if (tn == 0) return 0;
    const int64_t ntiles = (int64_t)(a.N / tn) * ceil_div(a.M, kPairM) * (mode == SGEMM_WGRAD ? a.splits : 1);
    return (int)(2 * (ntiles < sm_count / 2 ? ntiles : sm_count / 2));
}

int tc_gemm_launch(SgemmMode mode, const SgemmArgs &a_in, cudaStream_t stream) {
#ifdef OCPPO_BRINGUP
    static const int dbg = getenv("OCPPO_TC_DBG") ? atoi(getenv("OCPPO_TC_DBG")) : 0;   // bring-up knob only
    static const float trunc_scale = getenv("OCPPO_TC_TRUNC") ? (float)atof(getenv("OCPPO_TC_TRUNC")) : 1.f;   // calibration knob
#else
    constexpr int dbg = 0;               // release build: the knobs do not exist
    constexpr float trunc_scale = 1.f;
#endif
    SgemmArgs a = a_in;
    a.dbg = dbg;
    a.trunc_scale = trunc_scale;
    // epilogue staging policy of the conversion-free kernel: 0 = one tile per warp + six stages everywhere; 1 = two tiles +
    // five stages for FWD / DGRAD with K <= kEpi2MaxK (the epilogue-bound layers); 2 = for every FWD / DGRAD; 3 = WGRAD too
#ifdef OCPPO_BRINGUP
    static const int epi2_policy = getenv("OCPPO_TC_EPI2") ? atoi(getenv("OCPPO_TC_EPI2")) : kEpi2Policy;
#else
    constexpr int epi2_policy = kEpi2Policy;
#endif
    a.epi2 = epi2_policy >= 3 || (mode != SGEMM_WGRAD && (epi2_policy == 2 || (epi2_policy == 1 && a.K <= kEpi2MaxK)));
    OCPPO_CHECK_ARG(tc_gemm_supported(mode, a), "tc_gemm: unsupported shape M=%d N=%d K=%lld", a.M, a.N, (long long)a.K);
    const int splits = mode == SGEMM_WGRAD ? a.splits : 1;
    const bool split = a.single_acc == 0;
    typedef void (*Kern)(SgemmArgs, int, int, int, const CUtensorMap, const CUtensorMap, const CUtensorMap, const CUtensorMap, const CUtensorMap,
                         const CUtensorMap, const CUtensorMap);
    constexpr int kVariants = 9;
    static const Kern kerns[kVariants][3] = {
        {tc_gemm_kernel<SGEMM_FWD, false, 0>, tc_gemm_kernel<SGEMM_DGRAD, false, 0>, tc_gemm_kernel<SGEMM_WGRAD, false, 0>},
        {tc_gemm_kernel<SGEMM_FWD, true, 0>, tc_gemm_kernel<SGEMM_DGRAD, true, 0>, tc_gemm_kernel<SGEMM_WGRAD, true, 0>},
        {tc_gemm_kernel<SGEMM_FWD, true, 1>, tc_gemm_kernel<SGEMM_DGRAD, true, 1>, tc_gemm_kernel<SGEMM_WGRAD, true, 1>},
        {tc_gemm_kernel<SGEMM_FWD, false, 1>, tc_gemm_kernel<SGEMM_DGRAD, false, 1>, tc_gemm_kernel<SGEMM_WGRAD, false, 1>},
        {tc_gemm_kernel<SGEMM_FWD, true, 2>, tc_gemm_kernel<SGEMM_DGRAD, true, 2>, tc_gemm_kernel<SGEMM_WGRAD, true, 2>},
        {tc_gemm_kernel<SGEMM_FWD, true, 3>, tc_gemm_kernel<SGEMM_DGRAD, true, 3>, tc_gemm_kernel<SGEMM_WGRAD, true, 3>},
        {tc_gemm_kernel<SGEMM_FWD, false, 3>, tc_gemm_kernel<SGEMM_DGRAD, false, 3>, tc_gemm_kernel<SGEMM_WGRAD, false, 3>},
        {tc_gemm_kernel<SGEMM_FWD, true, 3, 128>, tc_gemm_kernel<SGEMM_DGRAD, true, 3, 128>, tc_gemm_kernel<SGEMM_WGRAD, true, 3, 128>},
        {tc_gemm_kernel<SGEMM_FWD, false, 3, 128>, tc_gemm_kernel<SGEMM_DGRAD, false, 3, 128>, tc_gemm_kernel<SGEMM_WGRAD, false, 3, 128>}};
    // per device: SM count, and the opt-in shared-memory size (a function attribute lives in the device's context)
    static int sms_of[kMaxDevices] = {0};
    int dev = 0;
    OCPPO_CUDA(cudaGetDevice(&dev));
    OCPPO_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "tc_gemm: device ordinal %d out of range", dev);
    if (sms_of[dev] == 0) {
        int n = 0;
        OCPPO_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
        for (int i = 0; i < kVariants; ++i)
            for (int j = 0; j < 3; ++j)
                OCPPO_CUDA(cudaFuncSetAttribute(kerns[i][j], cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes));
        sms_of[dev] = n;
    }
    const int sms = sms_of[dev];
    int tn = tc_tile_n(a);
    // a forward launch whose 256-column tiles would leave half of the CTA pairs idle (rollout inference: the decoder layer
    // of a 4096-env batch has 32 tiles for 74 pairs) takes the conversion-free kernel's 128-column tile instead: twice
    // the tiles, half the MMA work per stage, the same sums per output element (results are bit-identical)
    if (mode == SGEMM_FWD && tn == 256 && a.f16x3 && a.a16_hi != nullptr && a.b16_hi != nullptr &&
        2 * (int64_t)(a.N / 256) * ceil_div(a.M, kPairM) <= sms / 2)
        tn = 128;
    const int nt_n = a.N / tn, nt_m = (int)ceil_div(a.M, kPairM);
    const int64_t ntiles64 = (int64_t)nt_n * nt_m * splits;
    OCPPO_CHECK_ARG(ntiles64 < (1LL << 31), "tc_gemm: too many tiles");
    const int ntiles = (int)ntiles64;
    const int grid = 2 * (ntiles < sms / 2 ? ntiles : sms / 2);     // CTA pairs
    CUtensorMap tmA, tmB, tmC, tmX, tmB2, tmA2, tmC2;
    // operands that already live in global memory as fp16 hi / lo pairs: both (PRE kernel: no conversion warps), or only
    // the weights of FWD / DGRAD (the conversion kernel loads them straight into the MMA tiles)
    if (!a.f16x3) { a.a16_hi = a.a16_lo = a.b16_hi = a.b16_lo = nullptr; a.c16_hi = a.c16_lo = nullptr; }
    const bool pre = a.a16_hi != nullptr && a.b16_hi != nullptr;
    if (!pre) {
        a.a16_hi = a.a16_lo = nullptr;
        if (mode == SGEMM_WGRAD) { a.b16_hi = nullptr; a.b16_lo = nullptr; }
        OCPPO_CHECK_ARG(a.A != nullptr && (a.B != nullptr || a.b16_hi != nullptr), "tc_gemm: operand neither fp32 nor pre-split");
    } else {
        OCPPO_CHECK_ARG(a.a16_lo != nullptr && (a.lda & 7) == 0 && al16(a.a16_hi) && al16(a.a16_lo), "tc_gemm: bad pre-split A operand");
        OCPPO_CHECK_ARG(mode != SGEMM_WGRAD || a.bias_part == nullptr, "tc_gemm: the pre-split WGRAD kernel does not form bias sums");
        OCPPO_CHECK_ARG(mode != SGEMM_WGRAD || a.M % 64 == 0, "tc_gemm: pre-split WGRAD needs M %% 64 == 0");
    }
    if (pre) {
        OCPPO_TRY(make_map16(&tmA, a.a16_hi, mode == SGEMM_WGRAD, a.M, a.K, a.lda, TM));
        OCPPO_TRY(make_map16(&tmA2, a.a16_lo, mode == SGEMM_WGRAD, a.M, a.K, a.lda, TM));
    } else {
        OCPPO_TRY(make_map(&tmA, a.A, mode == SGEMM_WGRAD, a.M, a.K, a.lda, TM));
        tmA2 = tmA;
    }
    if (a.b16_hi != nullptr) {
        OCPPO_CHECK_ARG(a.b16_lo != nullptr && (a.ldb & 7) == 0 && al16(a.b16_hi) && al16(a.b16_lo), "tc_gemm: bad pre-split B operand");
        OCPPO_TRY(make_map16(&tmB, a.b16_hi, mode != SGEMM_FWD, a.N, a.K, a.ldb, tn / 2));
        OCPPO_TRY(make_map16(&tmB2, a.b16_lo, mode != SGEMM_FWD, a.N, a.K, a.ldb, tn / 2));
    } else {
        OCPPO_TRY(make_map(&tmB, a.B, mode != SGEMM_FWD, a.N, a.K, a.ldb, tn / 2));
        tmB2 = tmB;
    }
    if (mode == SGEMM_WGRAD) { a.c16_hi = nullptr; a.c16_lo = nullptr; }
    if (!(pre && mode == SGEMM_DGRAD)) a.colsum_part = nullptr;
    if (a.c16_hi != nullptr) {
        OCPPO_CHECK_ARG(a.c16_lo != nullptr && (a.ldc & 7) == 0 && al16(a.c16_hi) && al16(a.c16_lo) && a.bnd_amax != nullptr &&
                        a.bnd_wl1 != nullptr && a.c_scale_src != nullptr, "tc_gemm: bad fp16-pair output arguments");
        OCPPO_TRY(make_map_c16(&tmC, a.c16_hi, a.M, a.N, a.ldc));
        OCPPO_TRY(make_map_c16(&tmC2, a.c16_lo, a.M, a.N, a.ldc));
    } else {
        OCPPO_CHECK_ARG(a.C != nullptr, "tc_gemm: no output");
        OCPPO_TRY(make_map_c(&tmC, a.C, (int64_t)a.M * splits, a.N, a.ldc));
        tmC2 = tmC;
    }
    if (mode == SGEMM_DGRAD && a.mask != nullptr) OCPPO_TRY(make_map_c(&tmX, a.mask, a.M, a.N, a.ldc));
    else tmX = tmC;
    OCPPO_CHECK_ARG(mode != SGEMM_DGRAD || a.mask != nullptr || a.relu_bits != nullptr, "tc_gemm: DGRAD needs a mask or ReLU decision bits");
    const int mi = mode == SGEMM_FWD ? 0 : (mode == SGEMM_DGRAD ? 1 : 2);
    ProfScope ps(mode == SGEMM_FWD ? KC_SGEMM_FWD : (mode == SGEMM_DGRAD ? KC_SGEMM_DGRAD : KC_SGEMM_WGRAD), 1, stream);
    OCPPO_CHECK_ARG(!a.f16x3 || ((split || pre) && a.amax_a != nullptr && a.amax_b != nullptr), "tc_gemm: the FP16x3 engine needs both operands' max-abs");
    // fp16-pair storage with one accumulator (engine 8): FWD / DGRAD overlap their epilogue with the next tile's mainloop;
    // WGRAD drains only every 128 stages and keeps the split accumulators' 3x smaller rounding error
    OCPPO_CHECK_ARG(tn == 256 || pre, "tc_gemm: the 128-column tile needs both operands as fp16 pairs (N=%d)", a.N);
    int variant = a.f16x3 ? (pre ? ((split || (mode == SGEMM_WGRAD && !(dbg & 256))) ? 5 : 6) : 4) : (split ? (a.corr_bf16 ? 2 : 1) : (a.corr_bf16 ? 3 : 0));
    if (tn == 128) variant += 2;            // 5 / 6 -> 7 / 8
    kerns[variant][mi]<<<grid, pre ? kProdTid0 : kTcThreads, kTcSmemBytes, stream>>>(a, nt_n, nt_m, ntiles, tmA, tmB, tmC, tmX, tmB2, tmA2, tmC2);
    OCPPO_LAUNCH_CHECK("tc_gemm_kernel");
    return OCPPO_OK;
}

}  // namespace ocppo
