// orbx_match_tc.cuh -- brute-force Hamming best / second on the 5th-generation tensor cores (sm_100a: tcgen05 + TMEM + TMA).
// Included by orbx_match.cu (same translation unit as orbm_context).
//
// The reference's inner loop (ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:1983-2003, called from every search loop)
// is a 256-bit Hamming distance.  With both descriptors expanded to +-1 bytes, dot = sum a_k b_k = 256 - 2 * hamming, an
// exact integer, so a query tile x database tile of distances is one int8 GEMM:
//   D[128 queries x 256 rows] (int32, TMEM) = A[128 x 256 (K)] . B[256 x 256 (K)]^T     (tcgen05.mma kind::i8, M128 N256 K32)
// and the best / second selection of k_bruteforce runs as the epilogue out of TMEM on the integer pipe while the next tile
// multiplies.  Results are integers: bit-identical to the scalar loop.
//
// CTA = (query tile of 128, database chunk).  576 threads, warp-specialised:
//   warp 0       TMA producer: the CTA's query tile once (32 KB), then the chunk's database rows as stages of 256 rows x 128
//                bytes of K (32 KB, SWIZZLE_128B), 6 stages in flight, mbarrier full / empty per stage
//   warp 1       allocates TMEM (512 columns = two 128 x 256 int32 accumulators); one lane issues the MMAs: per stage four K=32
//                steps, tcgen05.commit releases the stage, after the second K half another commit hands the accumulator over
//   warps 2..17  epilogue: warp w reads TMEM lanes 32 (w % 4) .. +31 (its rows = queries) and columns 64 p .. 64 p + 63 of the
//                accumulator, p = (w - 2) / 4: two tcgen05.ld 32x32b.x32 back to back, one wait, and the accumulator is handed
//                back BEFORE the key arithmetic (the registers are the third buffer).  Two columns per instruction: 16-bit keys
//                hamming << 7 | column packed by two IMADs; (best, second) of two words at a time in five VIMNMX.U16x2 /
//                VIMNMX3.U16x2 (lo, hi, best = min(best, lo), second = min3(max(best, lo), second, hi)), two independent
//                accumulator pairs per warp; keys are rebased to chunk rows once per tile
// Measured on B200 (profiles/r2_bruteforce_tc.md): 0.373 ms per 2000 x 1M search = 2.81e15 int8 op/s; with the key arithmetic
// removed (TMA + MMA + TMEM loads only) the same pipeline takes 0.282 ms, so the epilogue costs 24 %.  Variants measured equal
// within 3 %: 8 epilogue warps x 128 columns with the hand-over after the arithmetic; one-word (three-operation) updates.
// Partials are written in k_bruteforce's format, so k_merge_partials / k_merge_exchange are shared.
#pragma once
#include <cuda.h>

namespace tc {

constexpr int BM = 128;                 // queries per CTA (UMMA M)
constexpr int BN = 256;                 // database rows per accumulator (UMMA N)
constexpr int KHALF = 128;              // bytes of K per stage row = swizzle span
constexpr int STAGES = 6;
constexpr int STAGE_BYTES = BN * KHALF; // 32768
constexpr int A_BYTES = BM * 256;       // 32768: two K halves of 128 rows x 128 B
constexpr int EPI_PARTS = 4;            // column parts of the accumulator, one epilogue warp per (lane group, part)
constexpr int EPI_COLS = BN / EPI_PARTS;
constexpr int THREADS = 32 * (2 + 4 * EPI_PARTS);
constexpr int SMEM_BYTES = 1024 + A_BYTES + STAGES * STAGE_BYTES + 256;
constexpr int EPI_WARPS = 4 * EPI_PARTS;
constexpr int THREADS2 = 320;           // pair kernel: 8 epilogue warps
constexpr int EPI_WARPS2 = 8;
constexpr int KEY_SENTINEL = 0x3fffffff;

// instruction descriptor (cute::UMMA::InstrDescriptor): D = S32 (2 << 4), A = B = signed 8 bit (1 << 7, 1 << 10), both K-major,
// N >> 3 at bit 17, M >> 4 at bit 24
constexpr uint32_t IDESC = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    uint32_t done;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
    } while (!done);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// shared-memory matrix descriptor (cute::UMMA::SmemDescriptor), K-major, SWIZZLE_128B: start >> 4, LBO unused, SBO = 8 rows x
// 128 B = 1024 (>> 4) at bit 32, version 1 at bit 46, layout type 2 at bit 61
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr)
{
    return (uint64_t)((addr & 0x3ffffu) >> 4) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void mma_i8(uint32_t tmemD, uint64_t descA, uint64_t descB, uint32_t accumulate)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
        ::"r"(tmemD), "l"(descA), "l"(descB), "r"(IDESC), "r"(accumulate), "r"(0u)
        : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
          "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
          "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
          "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}
// tcgen05.wait::ld with the loaded registers as in/out operands: the data dependency keeps the compiler from scheduling uses of
// v[] above the wait (the asynchronous load itself only names them as outputs)
__device__ __forceinline__ void tmem_ld_wait(uint32_t (&v)[32])
{
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]), "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]), "+r"(v[16]), "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]), "+r"(v[23]), "+r"(v[24]), "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]), "+r"(v[30]), "+r"(v[31])
                 :
                 : "memory");
}

// the same for two register sets loaded back to back
__device__ __forceinline__ void tmem_ld_wait2(uint32_t (&v)[32], uint32_t (&w)[32])
{
    tmem_ld_wait(v);
    // the wait above covers every outstanding load; this one only ties w[] to it for the compiler
    asm volatile(""
                 : "+r"(w[0]), "+r"(w[1]), "+r"(w[2]), "+r"(w[3]), "+r"(w[4]), "+r"(w[5]), "+r"(w[6]), "+r"(w[7]), "+r"(w[8]), "+r"(w[9]), "+r"(w[10]), "+r"(w[11]), "+r"(w[12]), "+r"(w[13]), "+r"(w[14]), "+r"(w[15]), "+r"(w[16]), "+r"(w[17]), "+r"(w[18]), "+r"(w[19]), "+r"(w[20]), "+r"(w[21]), "+r"(w[22]), "+r"(w[23]), "+r"(w[24]), "+r"(w[25]), "+r"(w[26]), "+r"(w[27]), "+r"(w[28]), "+r"(w[29]), "+r"(w[30]), "+r"(w[31])
                 :
                 : "memory");
}

// packed 256-bit descriptors -> +-1 bytes (bit k of byte i = element 8 i + k; 1 -> +1, 0 -> -1); rows beyond n are zero
__global__ void __launch_bounds__(256) k_expand_pm1(const uint8_t* __restrict__ src, long long n, long long nPad, int8_t* __restrict__ dst)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // one input byte -> 8 output bytes
    if (i >= nPad * 32) return;
    uint2 o = make_uint2(0u, 0u);
    if (i < n * 32) {
        const uint32_t b = src[i];
        // byte k = bit ? 0x01 : 0xFF
        uint32_t lo = 0, hi = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            lo |= (((b >> k) & 1u) ? 0x01u : 0xFFu) << (8 * k);
            hi |= (((b >> (k + 4)) & 1u) ? 0x01u : 0xFFu) << (8 * k);
        }
        o = make_uint2(lo, hi);
    }
    reinterpret_cast<uint2*>(dst)[i] = o;
}

struct EpiAcc {
    int b0, s0, b1, s1;  // two independent (best, second) pairs of tile-local signed keys
};

template <bool MASKED>
__device__ __forceinline__ void epi_chunk(const uint32_t (&v)[32], int colBase, int nvalidCols, EpiAcc& a)
{
    const int negC = -(1 << 21);
#pragma unroll
    for (int j = 0; j < 32; j += 2) {
        int k0 = (int)v[j] * negC + (colBase + j);
        int k1 = (int)v[j + 1] * negC + (colBase + j + 1);
        if (MASKED) {
            if (colBase + j >= nvalidCols) k0 = KEY_SENTINEL;
            if (colBase + j + 1 >= nvalidCols) k1 = KEY_SENTINEL;
        }
        a.s0 = min(a.s0, max(k0, a.b0));
        a.b0 = min(a.b0, k0);
        a.s1 = min(a.s1, max(k1, a.b1));
        a.b1 = min(a.b1, k1);
    }
}

// Packed epilogue: two columns per instruction.  key16 = hamming << 7 | column (7 bits, column inside the warp's part) =
// 16384 - 64 dot + column fits 16 bits (hamming <= 256), so a pair of columns is two IMADs (the second one shifts its key into
// the upper half) and the best / second update works on both halves at once (epi_push2).  The two halves of the word are independent
// (best, second) pairs, folded into chunk-relative 32-bit keys once per tile.
struct EpiAcc16 {
    uint32_t b, s;     // accumulator A: packed (best, second) of two independent column streams
    uint32_t b2, s2;   // accumulator B: the same for the other half of the words (two dependency chains in flight per warp)
};

// one packed key word: columns (col, col + 1) of the accumulator row
template <bool MASKED>
__device__ __forceinline__ uint32_t epi_key16(uint32_t d0, uint32_t d1, int col, int nvalidCols)
{
    const uint32_t c = (uint32_t)(16384 + col) | ((uint32_t)(16384 + col + 1) << 16);
    const uint32_t t = (uint32_t)((int)d0 * -64) + c;
    uint32_t p = (uint32_t)((int)d1 * -(64 << 16)) + t;
    if (MASKED) {
        if (col >= nvalidCols) p |= 0x0000ffffu;
        if (col + 1 >= nvalidCols) p |= 0xffff0000u;
    }
    return p;
}
// (best, second) <- two new words: lo / hi of the pair, best = min(best, lo), second = min3(max(best, lo), second, hi):
// five ALU operations per two words (the one-word update costs three each), dependency depth three
__device__ __forceinline__ void epi_push2(uint32_t& b, uint32_t& s, uint32_t p, uint32_t q)
{
    const uint32_t lo = __vminu2(p, q), hi = __vmaxu2(p, q);
    const uint32_t t = __vmaxu2(b, lo);
    b = __vminu2(b, lo);
    s = __vimin3_u16x2(t, s, hi);
}

template <bool MASKED>
__device__ __forceinline__ void epi_chunk16(const uint32_t (&v)[32], int colBase, int nvalidCols, EpiAcc16& a)
{
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
        const uint32_t p0 = epi_key16<MASKED>(v[j], v[j + 1], colBase + j, nvalidCols);
        const uint32_t p1 = epi_key16<MASKED>(v[j + 2], v[j + 3], colBase + j + 2, nvalidCols);
        const uint32_t p2 = epi_key16<MASKED>(v[j + 4], v[j + 5], colBase + j + 4, nvalidCols);
        const uint32_t p3 = epi_key16<MASKED>(v[j + 6], v[j + 7], colBase + j + 6, nvalidCols);
        epi_push2(a.b, a.s, p0, p1);
        epi_push2(a.b2, a.s2, p2, p3);
    }
}
// the two accumulators of a tile -> one packed (best, second)
__device__ __forceinline__ void epi_merge(EpiAcc16& a)
{
    const uint32_t t = __vmaxu2(a.b, a.b2);
    a.s = __vimin3_u16x2(t, a.s, a.s2);
    a.b = __vminu2(a.b, a.b2);
}

// 16-bit tile key -> chunk-relative 32-bit key (0xffff = nothing seen)
__device__ __forceinline__ int key16_to_32(uint32_t k, int rowBase)
{
    return k == 0xffffu ? 0x7fffffff : (int)(((k >> 7) << 22) | (uint32_t)(rowBase + (int)(k & 127u)));
}

__global__ void __launch_bounds__(THREADS, 1)
k_bruteforce_tc(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, int nq, int ndb, int chunkRows,
                int idxOffset, int32_t* __restrict__ pBest, int32_t* __restrict__ pIdx, int32_t* __restrict__ pSecond)
{
    extern __shared__ uint8_t smemRaw[];
    const uint32_t base = (smem_u32(smemRaw) + 1023u) & ~1023u;   // SWIZZLE_128B tiles need 1024-byte alignment
    const uint32_t sA = base;
    const uint32_t sB = base + A_BYTES;
    const uint32_t bars = sB + STAGES * STAGE_BYTES;
    const uint32_t barFull = bars;                       // STAGES x 8 B
    const uint32_t barEmpty = bars + 8 * STAGES;         // STAGES x 8 B
    const uint32_t barA = bars + 16 * STAGES;            // 8 B
    const uint32_t barTmemFull = barA + 8;               // 2 x 8 B
    const uint32_t barTmemEmpty = barTmemFull + 16;      // 2 x 8 B
    const uint32_t tmemSlot = barTmemEmpty + 16;         // 4 B
    uint8_t* const smemGeneric = smemRaw + (base - smem_u32(smemRaw));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int chunk = blockIdx.x, qt = blockIdx.y;
    const int c0 = chunk * chunkRows, c1 = min(c0 + chunkRows, ndb);
    const int nTiles = (c1 - c0 + BN - 1) / BN;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; s++) { mbar_init(barFull + 8 * s, 1); mbar_init(barEmpty + 8 * s, 1); }
        mbar_init(barA, 1);
        for (int b = 0; b < 2; b++) { mbar_init(barTmemFull + 8 * b, 1); mbar_init(barTmemEmpty + 8 * b, EPI_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmemSlot), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmemBase;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmemBase) : "r"(tmemSlot));

    if (warp == 0) {
        if (lane == 0) {
            // ===== TMA producer =====
            mbar_expect_tx(barA, A_BYTES);
            tma_load_2d(sA, &mapA, barA, 0, qt * BM);
            tma_load_2d(sA + A_BYTES / 2, &mapA, barA, KHALF, qt * BM);
            const int nIt = nTiles * 2;
            for (int it = 0; it < nIt; it++) {
                const int s = it % STAGES;
                mbar_wait(barEmpty + 8 * s, ((it / STAGES) & 1) ^ 1);
                mbar_expect_tx(barFull + 8 * s, STAGE_BYTES);
                tma_load_2d(sB + s * STAGE_BYTES, &mapB, barFull + 8 * s, (it & 1) * KHALF, c0 + (it >> 1) * BN);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // ===== MMA issuer =====
            mbar_wait(barA, 0);
            tc_fence_after();
            for (int t = 0; t < nTiles; t++) {
                const int b = t & 1;
                mbar_wait(barTmemEmpty + 8 * b, ((t >> 1) & 1) ^ 1);
                tc_fence_after();
                const uint32_t tmemD = tmemBase + (uint32_t)(b * BN);
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int it = 2 * t + h, s = it % STAGES;
                    mbar_wait(barFull + 8 * s, (it / STAGES) & 1);
                    tc_fence_after();
#pragma unroll
                    for (int kk = 0; kk < 4; kk++) {
                        const uint64_t dA = smem_desc(sA + h * (A_BYTES / 2) + kk * 32);
                        const uint64_t dB = smem_desc(sB + s * STAGE_BYTES + kk * 32);
                        mma_i8(tmemD, dA, dB, (h | kk) != 0 ? 1u : 0u);
                    }
                    tc_commit(barEmpty + 8 * s);   // stage reusable once these MMAs have read it
                }
                tc_commit(barTmemFull + 8 * b);    // accumulator complete
            }
        }
    } else {
        // ===== epilogue =====
        const int lg = warp & 3;                 // TMEM lane group this warp may read
        const int part = (warp - 2) >> 2;        // column part of the accumulator
        int B = 0x7fffffff, S = 0x7fffffff;      // chunk-relative keys: hamming << 22 | row in chunk
        for (int t = 0; t < nTiles; t++) {
            const int b = t & 1;
            mbar_wait(barTmemFull + 8 * b, (t >> 1) & 1);
            tc_fence_after();
            const uint32_t taddr = tmemBase + ((uint32_t)(lg * 32) << 16) + (uint32_t)(b * BN + part * EPI_COLS);
            const int rowBase = t * BN + part * EPI_COLS;
            const int nvalid = (c1 - c0) - rowBase;  // valid columns of this warp's part
            EpiAcc16 a = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
            // the warp's 64 columns go to registers in one go and the accumulator is handed back BEFORE the key arithmetic: the
            // MMA of tile t + 2 waits for the TMEM reads of tile t only, the registers act as a third buffer
            static_assert(EPI_COLS == 64, "two 32-column loads per warp and tile");
            uint32_t v0[32], v1[32];
            tmem_ld32(taddr, v0);
            tmem_ld32(taddr + 32, v1);
            tmem_ld_wait2(v0, v1);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(barTmemEmpty + 8 * b);
#ifdef ORBX_TC_EXPERIMENT_NO_EPILOGUE
            a.b += v0[0] + v1[0];   // experiment: MMA + TMEM-load ceiling (no key arithmetic); results are wrong
#else
            if (nvalid >= EPI_COLS) {
                epi_chunk16<false>(v0, 0, EPI_COLS, a);
                epi_chunk16<false>(v1, 32, EPI_COLS, a);
            } else {
                epi_chunk16<true>(v0, 0, nvalid, a);
                epi_chunk16<true>(v1, 32, nvalid, a);
            }
#endif
            // fold the tile (two accumulators of two independent 16-bit pairs) into the running chunk-relative pair
            epi_merge(a);
            const int b0 = key16_to_32(a.b & 0xffffu, rowBase), b1 = key16_to_32(a.b >> 16, rowBase);
            const int s0 = key16_to_32(a.s & 0xffffu, rowBase), s1 = key16_to_32(a.s >> 16, rowBase);
            const int tb = min(b0, b1);
            const int ts = min(max(b0, b1), min(s0, s1));
            S = min(min(S, ts), max(B, tb));
            B = min(B, tb);
        }
        // merge the column parts through shared memory (stage 0 is idle: every MMA has completed) and write partials
        int2* xch = reinterpret_cast<int2*>(smemGeneric + A_BYTES);
        const int row = lg * 32 + lane;
        if (part > 0) xch[(part - 1) * BM + row] = make_int2(B, S);
        asm volatile("bar.sync 1, %0;" ::"n"(32 * EPI_WARPS) : "memory");
        if (part == 0) {
#pragma unroll
            for (int pp = 0; pp < EPI_PARTS - 1; pp++) {
                const int2 o = xch[pp * BM + row];
                S = min(min(S, o.y), max(B, o.x));
                B = min(B, o.x);
            }
            const int qi = qt * BM + row;
            if (qi < nq) {
                const size_t oidx = (size_t)chunk * nq + qi;
                pBest[oidx] = B >> 22;
                pIdx[oidx] = (B & ((1 << 22) - 1)) + c0 + idxOffset;
                pSecond[oidx] = min(S >> 22, 256);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmemBase), "r"(512u) : "memory");
    }
}

// ------------------------------------------------------------------------------------------------------------------
// CTA-pair variant (cta_group::2): the two CTAs of a cluster (the two SMs of a TPC) multiply 256 queries x 256 database rows
// per tile as ONE UMMA of M = 256.  CTA r holds queries [128 r, 128 r + 128) of the pair's query tile (its A operand) and loads
// HALF of every database stage (rows [128 r, +128) of the 256-row tile, its half of the B operand); the MMA, issued by the
// leader CTA only, reads A and B from both CTAs' shared memory and writes each CTA's own TMEM (its 128 queries x all 256 rows).
// Per SM that is half the L2 -> shared-memory stream (measured 2.05 GB instead of 4.10 GB per 2000 x 1M search) and half the
// shared-memory operand reads of k_bruteforce_tc.
// MEASURED (profiles/r2_bruteforce_tc.md): 0.42 ms per 2000 x 1M search against 0.37 ms for the single-CTA kernel, tensor pipe
// 60 % against 71 % active.  The single-CTA kernel is not L2-bound (lts throughput 42 %); its limit is the epilogue keeping
// pace with the MMAs, and the pair adds a cross-CTA round trip (remote arrive -> leader -> multicast commit) to every
// accumulator hand-over.  So this is NOT the default; ORBX_TC_PAIR=1 selects it (bit-identical results, same tests).
// Barriers: full[s] lives in the leader (both CTAs' TMA loads complete_tx on it, .cta_group::2); empty[s] and tmemFull[b] are
// arrived in BOTH CTAs by multicast tcgen05.commit; tmemEmpty[b] lives in the leader and collects the epilogue warps of both CTAs.
// The epilogue is the single-CTA one.
// ------------------------------------------------------------------------------------------------------------------
constexpr int STAGE2_BYTES = 128 * KHALF;   // per CTA: 128 rows x 128 B of K
constexpr int STAGES2 = 8;
constexpr int SMEM2_BYTES = 1024 + A_BYTES + STAGES2 * STAGE2_BYTES + 256;
constexpr uint32_t IDESC2 = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(256 >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
constexpr uint32_t PEER_BIT_MASK = 0xFEFFFFFFu;   // shared::cluster address of the same offset in CTA 0 of the pair

__device__ __forceinline__ uint32_t cluster_ctarank()
{
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all()
{
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// bounded wait: a protocol error traps instead of hanging the GPU
__device__ __forceinline__ void mbar_wait_bounded(uint32_t bar, uint32_t parity)
{
    uint32_t done;
    uint32_t tries = 0;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (!done && ++tries > (1u << 26)) __trap();
    } while (!done);
}
__device__ __forceinline__ void tma_load_2d_2sm(uint32_t dst, const CUtensorMap* map, uint32_t barInLeader, int c0, int c1)
{
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(map), "r"(barInLeader), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void tc_commit_pair(uint32_t bar)   // arrives on the barrier at this offset in BOTH CTAs
{
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive_in_leader(uint32_t bar)
{
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(bar), "r"(0u));
    // default (.release.cta) semantics: the TMEM reads are ordered by tcgen05.fence::before_thread_sync; a cluster-scope release
    // would add a MEMBAR per warp and tile (measured: the top stall, 0.63 instead of 0.37 ms per search)
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}
__device__ __forceinline__ void mma_i8_pair(uint32_t tmemD, uint64_t descA, uint64_t descB, uint32_t accumulate)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5, %5, %5, %5, %5}, p;\n\t}"
        ::"r"(tmemD), "l"(descA), "l"(descB), "r"(IDESC2), "r"(accumulate), "r"(0u)
        : "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(THREADS2, 1)
k_bruteforce_tc2(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB, int nq, int ndb, int chunkRows,
                 int idxOffset, int32_t* __restrict__ pBest, int32_t* __restrict__ pIdx, int32_t* __restrict__ pSecond)
{
    extern __shared__ uint8_t smemRaw[];
    const uint32_t base = (smem_u32(smemRaw) + 1023u) & ~1023u;   // identical in both CTAs of the pair (same kernel, same layout)
    const uint32_t sA = base;
    const uint32_t sB = base + A_BYTES;
    const uint32_t bars = sB + STAGES2 * STAGE2_BYTES;
    const uint32_t barFull = bars;                       // STAGES2 x 8 B   (used in the leader)
    const uint32_t barEmpty = bars + 8 * STAGES2;        // STAGES2 x 8 B   (one per CTA)
    const uint32_t barA = bars + 16 * STAGES2;           // 8 B             (leader)
    const uint32_t barTmemFull = barA + 8;               // 2 x 8 B         (one per CTA)
    const uint32_t barTmemEmpty = barTmemFull + 16;      // 2 x 8 B         (leader)
    const uint32_t tmemSlot = barTmemEmpty + 16;         // 4 B
    uint8_t* const smemGeneric = smemRaw + (base - smem_u32(smemRaw));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;
    const int chunk = blockIdx.x >> 1, qt = blockIdx.y;
    const int c0 = chunk * chunkRows, c1 = min(c0 + chunkRows, ndb);
    const int nTiles = (c1 - c0 + 255) / 256;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES2; s++) { mbar_init(barFull + 8 * s, 1); mbar_init(barEmpty + 8 * s, 1); }
        mbar_init(barA, 1);
        for (int b = 0; b < 2; b++) { mbar_init(barTmemFull + 8 * b, 1); mbar_init(barTmemEmpty + 8 * b, 2 * EPI_WARPS2); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmemSlot), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    cluster_sync_all();      // barriers of both CTAs initialised, TMEM of both allocated
    tc_fence_after();
    uint32_t tmemBase;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmemBase) : "r"(tmemSlot));

    if (warp == 0) {
        if (lane == 0) {
            // ===== TMA producer (both CTAs): own query rows, own half of every database stage; bytes are counted in the leader =====
            if (leader) mbar_expect_tx(barA, 2 * A_BYTES);
            const int qrow = (qt * 2 + (int)rank) * BM;
            tma_load_2d_2sm(sA, &mapA, barA & PEER_BIT_MASK, 0, qrow);
            tma_load_2d_2sm(sA + A_BYTES / 2, &mapA, barA & PEER_BIT_MASK, KHALF, qrow);
            const int nIt = nTiles * 2;
            for (int it = 0; it < nIt; it++) {
                const int s = it % STAGES2;
                mbar_wait_bounded(barEmpty + 8 * s, ((it / STAGES2) & 1) ^ 1);
                if (leader) mbar_expect_tx(barFull + 8 * s, 2 * STAGE2_BYTES);
                tma_load_2d_2sm(sB + s * STAGE2_BYTES, &mapB, (barFull + 8 * s) & PEER_BIT_MASK, (it & 1) * KHALF,
                                c0 + (it >> 1) * 256 + (int)rank * 128);
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && leader) {
            // ===== MMA issuer (leader CTA only) =====
            mbar_wait_bounded(barA, 0);
            tc_fence_after();
            for (int t = 0; t < nTiles; t++) {
                const int b = t & 1;
                mbar_wait_bounded(barTmemEmpty + 8 * b, ((t >> 1) & 1) ^ 1);   // epilogues of BOTH CTAs have drained buffer b
                tc_fence_after();
                const uint32_t tmemD = tmemBase + (uint32_t)(b * 256);
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int it = 2 * t + h, s = it % STAGES2;
                    mbar_wait_bounded(barFull + 8 * s, (it / STAGES2) & 1);
                    tc_fence_after();
#pragma unroll
                    for (int kk = 0; kk < 4; kk++) {
                        const uint64_t dA = smem_desc(sA + h * (A_BYTES / 2) + kk * 32);
                        const uint64_t dB = smem_desc(sB + s * STAGE2_BYTES + kk * 32);
                        mma_i8_pair(tmemD, dA, dB, (h | kk) != 0 ? 1u : 0u);
                    }
                    tc_commit_pair(barEmpty + 8 * s);   // stage reusable in both CTAs once these MMAs have read it
                }
                tc_commit_pair(barTmemFull + 8 * b);    // accumulators of both CTAs complete
            }
        }
    } else {
        // ===== epilogue (both CTAs): as k_bruteforce_tc =====
        const int lg = warp & 3;
        const int half = (warp - 2) >> 2;
        int B = 0x7fffffff, S = 0x7fffffff;
        for (int t = 0; t < nTiles; t++) {
            const int b = t & 1;
            mbar_wait_bounded(barTmemFull + 8 * b, (t >> 1) & 1);
            tc_fence_after();
            const uint32_t taddr = tmemBase + ((uint32_t)(lg * 32) << 16) + (uint32_t)(b * 256 + half * 128);
            const int rowBase = t * 256 + half * 128;
            const int nvalid = (c1 - c0) - rowBase;
            EpiAcc16 a = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
            uint32_t v0[32], v1[32];
            tmem_ld32(taddr, v0);
            tmem_ld_wait(v0);
            tmem_ld32(taddr + 32, v1);
            if (nvalid >= 128) {
                epi_chunk16<false>(v0, 0, 128, a);
                tmem_ld_wait(v1);
                tmem_ld32(taddr + 64, v0);
                epi_chunk16<false>(v1, 32, 128, a);
                tmem_ld_wait(v0);
                tmem_ld32(taddr + 96, v1);
                epi_chunk16<false>(v0, 64, 128, a);
                tmem_ld_wait(v1);
                epi_chunk16<false>(v1, 96, 128, a);
            } else {
                epi_chunk16<true>(v0, 0, nvalid, a);
                tmem_ld_wait(v1);
                tmem_ld32(taddr + 64, v0);
                epi_chunk16<true>(v1, 32, nvalid, a);
                tmem_ld_wait(v0);
                tmem_ld32(taddr + 96, v1);
                epi_chunk16<true>(v0, 64, nvalid, a);
                tmem_ld_wait(v1);
                epi_chunk16<true>(v1, 96, nvalid, a);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_in_leader(barTmemEmpty + 8 * b);
            epi_merge(a);
            const int b0 = key16_to_32(a.b & 0xffffu, rowBase), b1 = key16_to_32(a.b >> 16, rowBase);
            const int s0 = key16_to_32(a.s & 0xffffu, rowBase), s1 = key16_to_32(a.s >> 16, rowBase);
            const int tb = min(b0, b1);
            const int ts = min(max(b0, b1), min(s0, s1));
            S = min(min(S, ts), max(B, tb));
            B = min(B, tb);
        }
        // merge the two column halves through shared memory (the A tile is idle: every MMA has completed) and write partials
        int2* xch = reinterpret_cast<int2*>(smemGeneric);
        const int row = lg * 32 + lane;
        if (half == 1) xch[row] = make_int2(B, S);
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (half == 0) {
            const int2 o = xch[row];
            const int s2 = min(min(S, o.y), max(B, o.x));
            const int b2 = min(B, o.x);
            const int qi = (qt * 2 + (int)rank) * BM + row;
            if (qi < nq) {
                const size_t oidx = (size_t)chunk * nq + qi;
                pBest[oidx] = b2 >> 22;
                pIdx[oidx] = (b2 & ((1 << 22) - 1)) + c0 + idxOffset;
                pSecond[oidx] = min(s2 >> 22, 256);
            }
        }
    }
    tc_fence_before();
    cluster_sync_all();      // both CTAs are done with both TMEMs and with each other's shared memory
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmemBase), "r"(512u) : "memory");
    }
}

typedef CUresult (*PFN_tensorMapEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_tensorMapEncodeTiled encode_fn()
{
    static PFN_tensorMapEncodeTiled fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (PFN_tensorMapEncodeTiled)p;
    }();
    return fn;
}

// rows x 256 bytes, row-major; box = boxRows x 128 bytes, SWIZZLE_128B, out-of-range rows read as zero
static int make_map(CUtensorMap* map, const void* ptr, long long rows, int boxRows)
{
    PFN_tensorMapEncodeTiled fn = encode_fn();
    if (!fn) { set_error("cuTensorMapEncodeTiled is not available from this driver"); return ORBX_ERR_CUDA; }
    const cuuint64_t gdim[2] = {256, (cuuint64_t)rows};
    const cuuint64_t gstride[1] = {256};
    const cuuint32_t box[2] = {(cuuint32_t)KHALF, (cuuint32_t)boxRows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with %d", (int)r); return ORBX_ERR_CUDA; }
    return ORBX_OK;
}

}  // namespace tc
