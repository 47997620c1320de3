// orbx_match.cu -- B200 (sm_100a) Hamming matchers behind the orbm_* C-ABI (include/orbx.h).
//
//   k_bruteforce / k_merge_partials   brute-force best / second-best over a descriptor database (north_star (6),
//                                     config 5), distance = ORBmatcher::DescriptorDistance src/ORBmatcher.cc:1983-2003
//   k_proj_search                     ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th)
//                                     src/ORBmatcher.cc:61-183 + Frame::GetFeaturesInArea src/Frame.cc:488-548
//   k_stereo_rows                     coarse search of Frame::ComputeStereoMatches src/Frame.cc:684-765
//   k_bow_transform                   DBoW2 TemplatedVocabulary::transform (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1218-1259)
//   k_bow_nodes / k_bow_rotation      ORBmatcher::SearchByBoW(KeyFrame*, Frame&, ...) src/ORBmatcher.cc:228-392,
//                                     ComputeThreeMaxima :1935-1977
// 256-bit distances are 8 x (LOP3 xor, POPC) on the integer pipe; best / second-best selection is a min over packed
// (distance, order) keys, reduced across lanes with warp shuffles.  sm_100a has no native b1 AND-POPC MMA
// (SURVEY.md F10), so there is no tensor-core variant.
#include "orbx_common.cuh"
#include "orbx_internal.h"
#include <cooperative_groups.h>
#include <algorithm>
#include <cstdlib>
#include <mutex>
#include <vector>

namespace cg = cooperative_groups;

namespace orbx {

constexpr int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;
constexpr int GRID_COLS = 64, GRID_ROWS = 48;

struct Desc { uint32_t w[8]; };

__device__ __forceinline__ Desc load_desc(const uint8_t* p)
{
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(p));
    const uint4 b = __ldg(reinterpret_cast<const uint4*>(p) + 1);
    Desc d;
    d.w[0] = a.x; d.w[1] = a.y; d.w[2] = a.z; d.w[3] = a.w;
    d.w[4] = b.x; d.w[5] = b.y; d.w[6] = b.z; d.w[7] = b.w;
    return d;
}
__device__ __forceinline__ int hamming(const Desc& a, const Desc& b)
{
    int d = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) d += __popc(a.w[i] ^ b.w[i]);
    return d;
}

// ------------------------------------------------------------------------------------------------------
// K7a: brute force.  CTA = 256 threads x BF_RQ queries in registers; the database chunk of the CTA streams through
// shared memory in tiles and is read with broadcast 128-bit loads.  Per (query, chunk): packed best key
// (dist << 22 | local index) and second-best distance.  Chunks are merged in index order by k_merge_partials, which
// reproduces the sequential scan (first minimum wins, second = second smallest of the multiset).
// ------------------------------------------------------------------------------------------------------
constexpr int BF_T = 256;
constexpr int BF_RQ = 4;
constexpr int BF_TILE = 256;  // descriptors per shared-memory tile (8 KB)
constexpr int BF_IDX_BITS = 22;
constexpr int BF_DEFAULT_VARIANT = 1;  // 6-POPC variant: fastest on B200 (profiles/r1_bruteforce_variants.md)

// 256-bit Hamming distance of (q ^ db) with fewer POPC than words: a carry-save adder on bit planes
// (sum = a^b^c, carry = maj(a,b,c), both single LOP3) turns three population counts into two.
//   VARIANT 0: 8 POPC (plain)          VARIANT 1: two 3:2 compressors -> 6 POPC
//   VARIANT 2: full 8 -> 4 tree (ones, twos, fours, eights planes)
template <int VARIANT>
__device__ __forceinline__ int hamming_regs(const Desc& Q, const uint4& a, const uint4& b)
{
    const uint32_t x0 = Q.w[0] ^ a.x, x1 = Q.w[1] ^ a.y, x2 = Q.w[2] ^ a.z, x3 = Q.w[3] ^ a.w;
    const uint32_t x4 = Q.w[4] ^ b.x, x5 = Q.w[5] ^ b.y, x6 = Q.w[6] ^ b.z, x7 = Q.w[7] ^ b.w;
    if (VARIANT == 0) {
        return __popc(x0) + __popc(x1) + __popc(x2) + __popc(x3) + __popc(x4) + __popc(x5) + __popc(x6) + __popc(x7);
    } else if (VARIANT == 1) {
        const uint32_t s1 = x0 ^ x1 ^ x2, c1 = (x0 & x1) | (x2 & (x0 | x1));
        const uint32_t s2 = x3 ^ x4 ^ x5, c2 = (x3 & x4) | (x5 & (x3 | x4));
        return __popc(s1) + __popc(s2) + __popc(x6) + __popc(x7) + 2 * (__popc(c1) + __popc(c2));
    } else if (VARIANT == 3) {  // three compressors -> 5 POPC
        const uint32_t s1 = x0 ^ x1 ^ x2, c1 = (x0 & x1) | (x2 & (x0 | x1));
        const uint32_t s2 = x3 ^ x4 ^ x5, c2 = (x3 & x4) | (x5 & (x3 | x4));
        const uint32_t s3 = s1 ^ s2 ^ x6, c3 = (s1 & s2) | (x6 & (s1 | s2));
        return __popc(s3) + __popc(x7) + 2 * (__popc(c1) + __popc(c2) + __popc(c3));
    } else if (VARIANT == 4) {  // four compressors -> 4 POPC
        const uint32_t s1 = x0 ^ x1 ^ x2, c1 = (x0 & x1) | (x2 & (x0 | x1));
        const uint32_t s2 = x3 ^ x4 ^ x5, c2 = (x3 & x4) | (x5 & (x3 | x4));
        const uint32_t s3 = s1 ^ s2 ^ x6, c3 = (s1 & s2) | (x6 & (s1 | s2));
        const uint32_t t = c1 ^ c2 ^ c3, f = (c1 & c2) | (c3 & (c1 | c2));
        return __popc(s3) + __popc(x7) + 2 * __popc(t) + 4 * __popc(f);
    } else {
        const uint32_t s1 = x0 ^ x1 ^ x2, c1 = (x0 & x1) | (x2 & (x0 | x1));
        const uint32_t s2 = x3 ^ x4 ^ x5, c2 = (x3 & x4) | (x5 & (x3 | x4));
        const uint32_t s3 = s1 ^ s2 ^ x6, c3 = (s1 & s2) | (x6 & (s1 | s2));
        const uint32_t ones = s3 ^ x7, c4 = s3 & x7;
        const uint32_t t1 = c1 ^ c2 ^ c3, f1 = (c1 & c2) | (c3 & (c1 | c2));
        const uint32_t twos = t1 ^ c4, f2 = t1 & c4;
        const uint32_t fours = f1 ^ f2, eights = f1 & f2;
        return __popc(ones) + 2 * __popc(twos) + 4 * __popc(fours) + 8 * __popc(eights);
    }
}

template <int VARIANT>
__global__ void __launch_bounds__(BF_T) k_bruteforce(const uint8_t* __restrict__ q, int nq, const uint8_t* __restrict__ db,
                                                     int ndb, int chunk, int idxOffset, int32_t* __restrict__ pBest,
                                                     int32_t* __restrict__ pIdx, int32_t* __restrict__ pSecond)
{
    __shared__ __align__(16) uint4 tile[BF_TILE * 2];
    const int c0 = blockIdx.x * chunk, c1 = min(c0 + chunk, ndb);
    const int qBase = blockIdx.y * (BF_T * BF_RQ);
    Desc Q[BF_RQ];
    uint32_t best[BF_RQ], second[BF_RQ];
#pragma unroll
    for (int r = 0; r < BF_RQ; r++) {
        const int qi = qBase + r * BF_T + threadIdx.x;
        if (qi < nq) Q[r] = load_desc(q + 32 * (size_t)qi);
        else {
#pragma unroll
            for (int i = 0; i < 8; i++) Q[r].w[i] = 0;
        }
        best[r] = (256u << BF_IDX_BITS) | ((1u << BF_IDX_BITS) - 1);
        second[r] = 256u << BF_IDX_BITS;
    }
    for (int t0 = c0; t0 < c1; t0 += BF_TILE) {
        const int nt = min(BF_TILE, c1 - t0);
        __syncthreads();
        for (int i = threadIdx.x; i < nt * 2; i += BF_T)
            tile[i] = __ldg(reinterpret_cast<const uint4*>(db + 32 * (size_t)t0) + i);
        __syncthreads();
        const uint32_t jBase = (uint32_t)(t0 - c0);
#pragma unroll 2
        for (int j = 0; j < nt; j++) {
            const uint4 a = tile[2 * j], b = tile[2 * j + 1];
#pragma unroll
            for (int r = 0; r < BF_RQ; r++) {
                const int d = hamming_regs<VARIANT>(Q[r], a, b);
                const uint32_t key = ((uint32_t)d << BF_IDX_BITS) | (jBase + j);
                second[r] = min(second[r], max(key, best[r]));
                best[r] = min(best[r], key);
            }
        }
    }
#pragma unroll
    for (int r = 0; r < BF_RQ; r++) {
        const int qi = qBase + r * BF_T + threadIdx.x;
        if (qi >= nq) continue;
        const size_t o = (size_t)blockIdx.x * nq + qi;
        const int bd = best[r] >> BF_IDX_BITS;
        pBest[o] = bd;
        pIdx[o] = bd < 256 || c1 > c0 ? (int)(best[r] & ((1u << BF_IDX_BITS) - 1)) + c0 + idxOffset : -1;
        pSecond[o] = min((int)(second[r] >> BF_IDX_BITS), 256);
    }
}

// sequential-scan-equivalent merge of partials ordered by ascending index range.  (best, idx) is the lexicographic
// minimum of (distance, global index) -- parts cover ascending index ranges, so "first minimum in scan order" is the
// smallest index -- and second is the second smallest distance of the multiset; that merge is associative and
// commutative, so one warp per query reduces its lanes' partial results with shuffles.
struct BfTriple { int b, id, s; };
__device__ __forceinline__ BfTriple bf_merge(BfTriple x, BfTriple y)
{
    if (y.id < 0) return x;
    if (x.id < 0) return y;
    const bool yWins = y.b < x.b || (y.b == x.b && y.id < x.id);
    BfTriple o;
    o.b = yWins ? y.b : x.b;
    o.id = yWins ? y.id : x.id;
    o.s = min(min(x.s, y.s), yWins ? x.b : y.b);
    return o;
}

__global__ void __launch_bounds__(256) k_merge_partials(const int32_t* __restrict__ pBest, const int32_t* __restrict__ pIdx,
                                                        const int32_t* __restrict__ pSecond, int nparts, int nq,
                                                        int32_t* __restrict__ oBest, int32_t* __restrict__ oIdx,
                                                        int32_t* __restrict__ oSecond)
{
    const int lane = threadIdx.x & 31;
    const int qi = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (qi >= nq) return;
    BfTriple acc = {256, -1, 256};
    for (int p = lane; p < nparts; p += 32) {
        const size_t o = (size_t)p * nq + qi;
        BfTriple t = {pBest[o], pIdx[o], pSecond[o]};
        acc = bf_merge(acc, t);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        BfTriple t;
        t.b = __shfl_xor_sync(0xffffffffu, acc.b, o);
        t.id = __shfl_xor_sync(0xffffffffu, acc.id, o);
        t.s = __shfl_xor_sync(0xffffffffu, acc.s, o);
        acc = bf_merge(acc, t);
    }
    if (lane == 0) { oBest[qi] = acc.b; oIdx[qi] = acc.id; oSecond[qi] = acc.s; }
}

// ------------------------------------------------------------------------------------------------------
// Frame::UndistortKeyPoints (src/Frame.cc:586-628) = cv::undistortPoints(pts, K, dist, R = I, P = K): 5 fixed-point
// iterations of the inverse Brown model in double, one thread per keypoint; every operation is an explicitly rounded
// intrinsic (no FMA contraction) so that the result is the CPU's, bit for bit (oracle: orbo_undistort_points).
// ------------------------------------------------------------------------------------------------------
struct UndistortCam { double fx, fy, cx, cy, k[12]; };

__global__ void __launch_bounds__(128) k_undistort(const float* __restrict__ px, const float* __restrict__ py, int n,
                                                   UndistortCam c, float* __restrict__ ox, float* __restrict__ oy)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double* k = c.k;
    const double ifx = __ddiv_rn(1.0, c.fx), ify = __ddiv_rn(1.0, c.fy);
    const double u = (double)px[i], v = (double)py[i];
    double x = __dmul_rn(__dsub_rn(u, c.cx), ifx), y = __dmul_rn(__dsub_rn(v, c.cy), ify);
    const double x0 = x, y0 = y;
#define M_(a, b) __dmul_rn(a, b)
#define A_(a, b) __dadd_rn(a, b)
    for (int j = 0; j < 5; j++) {
        const double r2 = A_(M_(x, x), M_(y, y));
        const double num = A_(1.0, M_(A_(M_(A_(M_(k[7], r2), k[6]), r2), k[5]), r2));
        const double den = A_(1.0, M_(A_(M_(A_(M_(k[4], r2), k[1]), r2), k[0]), r2));
        const double icdist = __ddiv_rn(num, den);
        if (icdist < 0) { x = x0; y = y0; break; }
        // 2*k2*x*y + k3*(r2 + 2*x*x) + k8*r2 + k9*r2*r2, left to right as the C++ expression evaluates
        const double deltaX = A_(A_(A_(M_(M_(M_(2.0, k[2]), x), y), M_(k[3], A_(r2, M_(M_(2.0, x), x)))), M_(k[8], r2)), M_(M_(k[9], r2), r2));
        const double deltaY = A_(A_(A_(M_(k[2], A_(r2, M_(M_(2.0, y), y))), M_(M_(M_(2.0, k[3]), x), y)), M_(k[10], r2)), M_(M_(k[11], r2), r2));
        x = M_(__dsub_rn(x0, deltaX), icdist);
        y = M_(__dsub_rn(y0, deltaY), icdist);
    }
    ox[i] = (float)A_(M_(c.fx, x), c.cx);
    oy[i] = (float)A_(M_(c.fy, y), c.cy);
#undef M_
#undef A_
}

// ------------------------------------------------------------------------------------------------------
// Frame::AssignFeaturesToGrid / PosInGrid (src/Frame.cc:351-370, :558-567): 64 x 48 grid as CSR, cell (ix, iy) at
// ix * 48 + iy, keypoint indices ascending inside a cell (the reference pushes them in index order).  One CTA: cell
// counts by shared-memory atomics, exclusive scan, scatter in arbitrary order, then every cell's few entries are put
// in ascending order by one thread.
// ------------------------------------------------------------------------------------------------------
constexpr int GRID_CELLS = 64 * 48;

__global__ void __launch_bounds__(1024) k_assign_grid(const float* __restrict__ kx, const float* __restrict__ ky, int n, float minX,
                                                      float minY, float wInv, float hInv, int32_t* __restrict__ cellStart,
                                                      int32_t* __restrict__ cellItems)
{
    __shared__ int cnt[GRID_CELLS];
    __shared__ int fill[GRID_CELLS];
    __shared__ int wsum[33];
    const int tid = threadIdx.x;
    for (int c = tid; c < GRID_CELLS; c += 1024) cnt[c] = 0;
    __syncthreads();
    auto cell_of = [&](int i) {
        // posX = round((x - mnMinX) * mfGridElementWidthInv): float arithmetic, round half away from zero
        const int gx = (int)roundf(__fmul_rn(__fsub_rn(kx[i], minX), wInv));
        const int gy = (int)roundf(__fmul_rn(__fsub_rn(ky[i], minY), hInv));
        return (gx < 0 || gx >= 64 || gy < 0 || gy >= 48) ? -1 : gx * 48 + gy;
    };
    for (int i = tid; i < n; i += 1024) {
        const int c = cell_of(i);
        if (c >= 0) atomicAdd(&cnt[c], 1);
    }
    __syncthreads();
    // exclusive scan of 3072 counts: 3 per thread
    int v[3], sum = 0;
#pragma unroll
    for (int j = 0; j < 3; j++) { v[j] = cnt[3 * tid + j]; sum += v[j]; }
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if ((tid & 31) >= o) incl += t;
    }
    if ((tid & 31) == 31) wsum[tid >> 5] = incl;
    __syncthreads();
    if (tid < 32) {
        const int w = wsum[tid];
        int wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, wi, o);
            if (tid >= o) wi += t;
        }
        wsum[tid] = wi - w;
        if (tid == 31) wsum[32] = wi;
    }
    __syncthreads();
    int run = wsum[tid >> 5] + incl - sum;
#pragma unroll
    for (int j = 0; j < 3; j++) { fill[3 * tid + j] = run; cellStart[3 * tid + j] = run; run += v[j]; }
    if (tid == 0) cellStart[GRID_CELLS] = wsum[32];
    __syncthreads();
    for (int i = tid; i < n; i += 1024) {
        const int c = cell_of(i);
        if (c >= 0) cellItems[atomicAdd(&fill[c], 1)] = i;
    }
    __syncthreads();
    for (int c = tid; c < GRID_CELLS; c += 1024) {
        const int m = cnt[c];
        if (m < 2) continue;
        int32_t* a = cellItems + (fill[c] - m);
        for (int p = 1; p < m; p++) {  // insertion sort: a cell holds a handful of keypoints
            const int key = a[p];
            int q = p - 1;
            while (q >= 0 && a[q] > key) { a[q + 1] = a[q]; q--; }
            a[q + 1] = key;
        }
    }
}

// ------------------------------------------------------------------------------------------------------
// Sharded search, fused merge + exchange over peer memory (NVLink / NVSwitch P2P stores, no NCCL call on the data path).
// Every rank owns an exchange buffer that all peers have mapped: slots[parity][source rank][query] of one 64-bit word
//   epoch (14 bits) << 50 | best (9) << 41 | second (9) << 32 | index (32).
// One warp per query merges the rank's chunk partials, lanes 0..world-1 store the packed triple into the slot of this
// rank in every peer's buffer (one 64-bit store each: payload and "ready" flag travel in the same word, so no fence or
// separate signal is needed), then spin on the `world` slots of the query in the local buffer until they carry the
// current epoch and merge them in the same deterministic way as k_merge_partials.  Slots alternate between two sets by
// epoch parity: a peer may already be writing call k+1 while this rank still reads call k.
// The 14-bit tag wraps every 16384 calls and nq may differ from call to call, so a slot that recent calls did not use
// (query index >= their nq) could still carry a tag that becomes current again.  Every publish therefore also rewrites a
// rotating window of BFX scrub slots of its set with (current tag, empty triple): each slot of a set is rewritten at
// least once per 2048 calls of that parity, so the tag found in a slot is never more than 4096 epochs old and cannot
// alias the current one (slots below nq carry the real triple of this call instead).  The grid is one CTA per SM
// (persistent, always co-resident), the spin is bounded by a cycle budget and reports a timeout instead of hanging.
// ------------------------------------------------------------------------------------------------------
constexpr int BFX_MAX_WORLD = 16;
struct BfxPeers { unsigned long long* p[BFX_MAX_WORLD]; };

__global__ void __launch_bounds__(256) k_merge_exchange(const int32_t* __restrict__ pBest, const int32_t* __restrict__ pIdx,
                                                        const int32_t* __restrict__ pSecond, int nparts, int nq, int nqMax,
                                                        const __grid_constant__ BfxPeers peers, int rank, int world, uint32_t epoch,
                                                        long long spinBudget, int* __restrict__ timeoutFlag,
                                                        int32_t* __restrict__ oBest, int32_t* __restrict__ oIdx,
                                                        int32_t* __restrict__ oSecond, int phases, int scrubFirst, int scrubCount)
{
    // phases: 1 = publish (merge the chunk partials, store to the peers), 2 = collect (wait, merge, write the result),
    // 3 = both (the normal, fused call)
    const int lane = threadIdx.x & 31;
    const int warpsTotal = gridDim.x * (blockDim.x >> 5);
    const unsigned long long ep = (unsigned long long)(epoch & 0x3fffu);
    const size_t setOff = (size_t)(epoch & 1u) * world * nqMax;
    if ((phases & 1) && ((blockIdx.x * blockDim.x + threadIdx.x) >> 5) == warpsTotal - 1 && lane < world) {
        // tag scrub (see above): unused slots of this call's set get the current tag with an empty triple
        const unsigned long long empty = (ep << 50) | (256ull << 41) | (256ull << 32) | 0xffffffffull;
        for (int k = 0; k < scrubCount; k++) {
            int sl = scrubFirst + k;
            if (sl >= nqMax) sl -= nqMax;
            if (sl >= nq)
                asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(peers.p[lane] + setOff + (size_t)rank * nqMax + sl), "l"(empty) : "memory");
        }
    }
    for (int qi = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; qi < nq; qi += warpsTotal) {
        BfTriple acc = {256, -1, 256};
        if (phases & 1) {
            for (int p = lane; p < nparts; p += 32) {
                const size_t o = (size_t)p * nq + qi;
                BfTriple t = {pBest[o], pIdx[o], pSecond[o]};
                acc = bf_merge(acc, t);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                BfTriple t;
                t.b = __shfl_xor_sync(0xffffffffu, acc.b, o);
                t.id = __shfl_xor_sync(0xffffffffu, acc.id, o);
                t.s = __shfl_xor_sync(0xffffffffu, acc.s, o);
                acc = bf_merge(acc, t);
            }
            if (lane < world) {
                const unsigned long long word = (ep << 50) | ((unsigned long long)(uint32_t)acc.b << 41) |
                                                ((unsigned long long)(uint32_t)acc.s << 32) | (unsigned long long)(uint32_t)acc.id;
                // lane p -> peer p: slot of `rank` for this query
                asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(peers.p[lane] + setOff + (size_t)rank * nqMax + qi), "l"(word) : "memory");
            }
        }
        if (!(phases & 2)) continue;
        BfTriple mine = {256, -1, 256};
        if (lane < world) {
            // lane p <- slot of rank p in the local buffer
            const unsigned long long* slot = peers.p[rank] + setOff + (size_t)lane * nqMax + qi;
            unsigned long long v;
            const long long t0 = clock64();
            for (;;) {
                asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(slot) : "memory");
                if ((v >> 50) == ep) break;
                if (clock64() - t0 > spinBudget) { *timeoutFlag = 1; v = (ep << 50) | (256ull << 41) | (256ull << 32) | 0xffffffffull; break; }
                __nanosleep(64);
            }
            mine.b = (int)((v >> 41) & 0x1ffu);
            mine.s = (int)((v >> 32) & 0x1ffu);
            mine.id = (int)(uint32_t)(v & 0xffffffffull);
        }
        __syncwarp();
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            BfTriple t;
            t.b = __shfl_xor_sync(0xffffffffu, mine.b, o);
            t.id = __shfl_xor_sync(0xffffffffu, mine.id, o);
            t.s = __shfl_xor_sync(0xffffffffu, mine.s, o);
            mine = bf_merge(mine, t);
        }
        if (lane == 0) { oBest[qi] = mine.b; oIdx[qi] = mine.id; oSecond[qi] = mine.s; }
    }
}

// ------------------------------------------------------------------------------------------------------
// two-smallest tracking on packed keys (dist << 20 | visit order) with a payload for the best entry
// ------------------------------------------------------------------------------------------------------
struct Top2 {
    uint32_t k1, k2;  // smallest and second smallest key
    int p1, p2;       // payloads
};
__device__ __forceinline__ void top2_init(Top2& t) { t.k1 = t.k2 = 0xFFFFFFFFu; t.p1 = t.p2 = -1; }
__device__ __forceinline__ void top2_push(Top2& t, uint32_t k, int p)
{
    if (k < t.k1) { t.k2 = t.k1; t.p2 = t.p1; t.k1 = k; t.p1 = p; }
    else if (k < t.k2) { t.k2 = k; t.p2 = p; }
}
__device__ __forceinline__ void top2_warp_reduce(Top2& t)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const uint32_t ok1 = __shfl_xor_sync(0xffffffffu, t.k1, o), ok2 = __shfl_xor_sync(0xffffffffu, t.k2, o);
        const int op1 = __shfl_xor_sync(0xffffffffu, t.p1, o), op2 = __shfl_xor_sync(0xffffffffu, t.p2, o);
        top2_push(t, ok1, op1);
        top2_push(t, ok2, op2);
    }
}

// ------------------------------------------------------------------------------------------------------
// K7b: SearchByProjection(Frame, MapPoints, th).  The reference is a greedy sequential loop: a keypoint claimed by
// an earlier map point (with Observations() > 0) is skipped by later ones.  result(m) depends only on the claims of
// map points j < m, so the sequential result is the unique fixed point of "recompute every map point against the
// claims of the previous sweep"; sweeps are fully parallel (one warp per map point) and repeat inside one cooperative
// kernel until nothing changes (at most the longest claim chain + 1 sweeps, typically 2-3).
// ------------------------------------------------------------------------------------------------------
struct ProjArgs {
    // frame
    int nkp;
    const float* kx; const float* ky; const int32_t* koct; const float* uright; const uint8_t* kdesc;
    const uint8_t* occupied; const int32_t* cellStart; const int32_t* cellItems;
    float minX, minY, wInv, hInv;
    const float* kangle;  // mvKeysUn[i].angle (orientation histogram only)
    // queries (map points / source features after projection)
    int nmp;
    const float* px; const float* py; const float* pxr; const float* prr; const int32_t* pmin; const int32_t* pmax;
    const float* pangle;
    const uint8_t* pvalid; const uint8_t* pobs; const uint8_t* pdesc;
    float nnratio;
    int maxDist, ratioMode, checkOri;
    int* hist;          // [HISTO_LENGTH] rotation histogram
    // work / output
    int32_t* res;       // per map point: claimed keypoint or -1
    int32_t* claimA;    // per keypoint: smallest blocking map point index claiming it (INT_MAX = none)
    int32_t* claimB;
    int32_t* claimC;    // third buffer: reset while the other two are read / written (one grid barrier per sweep)
    // candidate lists of sweep 0 (every keypoint that passes a query's static gates, with its distance / order key): the
    // later sweeps only re-apply the claims to them
    uint32_t* listKey; int32_t* listIdx; int32_t* listCnt;
    int32_t* outMatch;  // per keypoint: last writer or -1
    int* changed;       // [2]
    int* nmatches;
    int* sweeps;
};

// rotation-histogram bin of a match (src/ORBmatcher.cc:1731-1738): round((a1 - a2 [+360]) * HISTO_LENGTH/360), 30 -> 0
__device__ __forceinline__ int rot_bin(float a1, float a2)
{
    float rot = __fsub_rn(a1, a2);
    if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
    int bin = (int)roundf(__fmul_rn(rot, 0.0833333358f));
    return bin == HISTO_LENGTH ? 0 : bin;
}

constexpr int PROJ_LIST = 32;  // cached candidates per query (one per lane); a query with more is recomputed in full every sweep

__device__ __forceinline__ int proj_finish(const ProjArgs& A, Top2& t)
{
    top2_warp_reduce(t);
    if (t.p1 < 0) return -1;
    const int bestDist = t.k1 >> 23;
    if (bestDist > A.maxDist) return -1;
    if (A.ratioMode == 1) {
        const int bestLevel = A.koct[t.p1];
        const int bestLevel2 = t.p2 >= 0 ? A.koct[t.p2] : -1;
        const int bestDist2 = t.p2 >= 0 ? (int)(t.k2 >> 23) : 256;
        if (bestLevel == bestLevel2 && (float)bestDist > __fmul_rn(A.nnratio, (float)bestDist2)) return -1;
    }
    return t.p1;
}

// sweeps >= 1: the query's cached candidates (lane = list entry) against the previous sweep's claims
__device__ __forceinline__ int proj_cached(const ProjArgs& A, int m, const int32_t* claimPrev, int lane, int cnt)
{
    Top2 t;
    top2_init(t);
    if (lane < cnt) {
        const int idx = A.listIdx[(size_t)m * PROJ_LIST + lane];
        if (!(claimPrev[idx] < m)) top2_push(t, A.listKey[(size_t)m * PROJ_LIST + lane], idx);
    }
    return proj_finish(A, t);
}

template <bool BUILD>
__device__ int proj_one(const ProjArgs& A, int m, const int32_t* claimPrev, int lane)
{
    if (!A.pvalid[m]) return -1;
    const float rr = A.prr[m];
    const float x = A.px[m], y = A.py[m];
    const int c0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, A.minX), rr), A.wInv)));
    if (c0 >= GRID_COLS) return -1;
    const int c1 = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, A.minX), rr), A.wInv)));
    if (c1 < 0) return -1;
    const int r0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, A.minY), rr), A.hInv)));
    if (r0 >= GRID_ROWS) return -1;
    const int r1 = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, A.minY), rr), A.hInv)));
    if (r1 < 0) return -1;
    const int minLevel = A.pmin[m], maxLevel = A.pmax[m];
    const bool checkLevels = (minLevel > 0) || (maxLevel >= 0);
    const Desc D = load_desc(A.pdesc + 32 * (size_t)m);
    const float xr = A.pxr[m];
    Top2 t;
    top2_init(t);
    // one lane per grid cell of the window (ix outer, iy inner = the reference's visiting order); the order key is
    // (cell sequence number, index inside the cell): 12 + 11 bits below the 9-bit distance
    const int nr = r1 - r0 + 1, ncell = (c1 - c0 + 1) * nr;
    for (int qc = lane; qc < ncell; qc += 32) {
        const int ix = c0 + qc / nr, iy = r0 + qc % nr;
        const int c = ix * GRID_ROWS + iy;
        const int s = A.cellStart[c], e = A.cellStart[c + 1];
        for (int k = s; k < e; k++) {
            const int idx = A.cellItems[k];
            const int oct = A.koct[idx];
            if (checkLevels) {
                if (oct < minLevel) continue;
                if (maxLevel >= 0 && oct > maxLevel) continue;
            }
            const float dx = __fsub_rn(A.kx[idx], x), dy = __fsub_rn(A.ky[idx], y);
            if (!(fabsf(dx) < rr && fabsf(dy) < rr)) continue;
            if (A.occupied[idx] || claimPrev[idx] < m) continue;
            if (A.uright) {
                const float ur = A.uright[idx];
                if (ur > 0 && fabsf(__fsub_rn(xr, ur)) > rr) continue;
            }
            const int dist = hamming(D, load_desc(A.kdesc + 32 * (size_t)idx));
            const uint32_t key = ((uint32_t)dist << 23) | ((uint32_t)qc << 11) | (uint32_t)(k - s);
            if (BUILD) {  // sweep 0: nothing is claimed yet, every gated candidate is recorded
                const int pos = atomicAdd(&A.listCnt[m], 1);
                if (pos < PROJ_LIST) { A.listKey[(size_t)m * PROJ_LIST + pos] = key; A.listIdx[(size_t)m * PROJ_LIST + pos] = idx; }
            }
            top2_push(t, key, idx);
        }
    }
    return proj_finish(A, t);
}

// threads per CTA: few large CTAs keep the grid-wide barriers between the sweeps cheap (one arrival per CTA)
static const int PROJ_T = [] { const char* e = std::getenv("ORBX_PROJ_T"); const int v = e ? std::atoi(e) : 0; return v >= 32 && v <= 1024 && v % 32 == 0 ? v : 1024; }();

__global__ void __launch_bounds__(1024) k_proj_search(ProjArgs A)
{
    cg::grid_group grid = cg::this_grid();
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x, gthreads = gridDim.x * blockDim.x;
    int32_t* claimPrev = A.claimA;   // claims of the previous sweep's results (read)
    int32_t* claimNext = A.claimB;   // claims of this sweep's results (atomicMin as the results are produced)
    int32_t* claimSpare = A.claimC;  // reset during this sweep, the next sweep's claimNext
    for (int i = gtid; i < A.nkp; i += gthreads) { claimPrev[i] = 0x7fffffff; claimNext[i] = 0x7fffffff; A.outMatch[i] = -1; }
    for (int m = gtid; m < A.nmp; m += gthreads) { A.res[m] = -1; A.listCnt[m] = 0; }
    if (gtid == 0) { A.changed[0] = 0; A.changed[1] = 0; A.changed[2] = 0; *A.nmatches = 0; }
    if (gtid < HISTO_LENGTH) A.hist[gtid] = 0;
    grid.sync();
    int sweep = 0;
    for (;; sweep++) {
        // three flags in rotation: with ONE barrier per sweep a thread may still be reading sweep k's flag (after that sweep's
        // barrier) while a faster one is already in sweep k + 1, so the flag reset there is the one of sweep k + 2
        int* chg = A.changed + sweep % 3;
        for (int m = warp; m < A.nmp; m += nwarps) {
            int r;
            if (sweep == 0) r = proj_one<true>(A, m, claimPrev, lane);
            else {
                const int cnt = A.listCnt[m];  // warp-uniform
                r = cnt <= PROJ_LIST ? proj_cached(A, m, claimPrev, lane, cnt) : proj_one<false>(A, m, claimPrev, lane);
            }
            if (lane == 0) {
                if (r != A.res[m]) { A.res[m] = r; *chg = 1; }
                if (r >= 0 && A.pobs[m]) atomicMin(&claimNext[r], m);
            }
        }
        for (int i = gtid; i < A.nkp; i += gthreads) claimSpare[i] = 0x7fffffff;
        if (gtid == 0) A.changed[(sweep + 1) % 3] = 0;
        grid.sync();
        if (*chg == 0) break;  // fixed point: res == sequential result
        int32_t* t = claimPrev; claimPrev = claimNext; claimNext = claimSpare; claimSpare = t;
    }
    // F.mvpMapPoints[bestIdx] = pMP: the last writer wins
    int local = 0;
    for (int m = gtid; m < A.nmp; m += gthreads) {
        const int r = A.res[m];
        if (r >= 0) {
            atomicMax(&A.outMatch[r], m);
            local++;
            if (A.checkOri) atomicAdd(&A.hist[rot_bin(A.pangle[m], A.kangle[r])], 1);
        }
    }
    if (local) atomicAdd(A.nmatches, local);
    if (gtid == 0) *A.sweeps = sweep + 1;
    if (!A.checkOri) return;
    grid.sync();
    // ComputeThreeMaxima + removal of the matches outside the three main rotation bins (set to NULL: -2)
    __shared__ int keepBin[HISTO_LENGTH];
    if (threadIdx.x == 0) {
        int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            const int sN = A.hist[i];
            if (sN > max1) { max3 = max2; max2 = max1; max1 = sN; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (sN > max2) { max3 = max2; max2 = sN; ind3 = ind2; ind2 = i; }
            else if (sN > max3) { max3 = sN; ind3 = i; }
        }
        if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
        else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) ind3 = -1;
        for (int i = 0; i < HISTO_LENGTH; i++) keepBin[i] = (i == ind1 || i == ind2 || i == ind3);
    }
    __syncthreads();
    int removed = 0;
    for (int m = gtid; m < A.nmp; m += gthreads) {
        const int r = A.res[m];
        if (r >= 0 && !keepBin[rot_bin(A.pangle[m], A.kangle[r])]) { A.outMatch[r] = -2; removed++; }
    }
    if (removed) atomicSub(A.nmatches, removed);
}

// ------------------------------------------------------------------------------------------------------
// K7c: stereo row search.  One warp per left keypoint; the reference's per-row candidate lists hold right keypoint
// iR in every row of [floor(y - r), ceil(y + r)], r = 2*scale[octave], in ascending iR, so "first minimum in list
// order" is the minimum over (distance, iR) of the right keypoints whose band contains row (int)vL.
// ------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_stereo_rows(int nl, const float* __restrict__ lx, const float* __restrict__ ly,
                                                     const int32_t* __restrict__ loct, const uint8_t* __restrict__ ldesc,
                                                     int nr, const float* __restrict__ rx, const float* __restrict__ ry,
                                                     const int32_t* __restrict__ roct, const uint8_t* __restrict__ rdesc,
                                                     const float* __restrict__ scaleFactors, int nRows, float minD,
                                                     float maxD, int32_t* __restrict__ bestR, int32_t* __restrict__ bestDist)
{
    const int lane = threadIdx.x & 31;
    const int iL = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (iL >= nl) return;
    const float uL = lx[iL];
    const int row = (int)ly[iL];
    const int levelL = loct[iL];
    const float minU = __fsub_rn(uL, maxD), maxU = __fsub_rn(uL, minD);
    uint32_t best = 0xFFFFFFFFu;
    if (row >= 0 && row < nRows && !(maxU < 0)) {
        const Desc D = load_desc(ldesc + 32 * (size_t)iL);
        for (int iR = lane; iR < nr; iR += 32) {
            const int oR = roct[iR];
            const float r = __fmul_rn(2.0f, scaleFactors[oR]);
            const float y = ry[iR];
            const int maxr = (int)ceilf(__fadd_rn(y, r)), minr = (int)floorf(__fsub_rn(y, r));
            if (row < minr || row > maxr) continue;
            if (oR < levelL - 1 || oR > levelL + 1) continue;
            const float uR = rx[iR];
            if (!(uR >= minU && uR <= maxU)) continue;
            const int d = hamming(D, load_desc(rdesc + 32 * (size_t)iR));
            best = min(best, ((uint32_t)d << 22) | (uint32_t)iR);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    if (lane == 0) {
        int d = TH_HIGH, idx = -1;
        if (best != 0xFFFFFFFFu && (int)(best >> 22) < TH_HIGH) { d = best >> 22; idx = best & ((1u << 22) - 1); }
        bestDist[iL] = d;
        bestR[iL] = (d < (TH_HIGH + TH_LOW) / 2) ? idx : -1;
    }
}

// ------------------------------------------------------------------------------------------------------
// K7c': SAD refinement of Frame::ComputeStereoMatches (src/Frame.cc:767-866) on the device-resident pyramids of the two
// extractor handles: for every coarse match, an 11x11 window (centre-subtracted) slides over 11 horizontal shifts in the
// right level image; L1 distances are integers (exact in float, as in the reference's cv::norm of CV_32F patches);
// parabola fit of the minimum in unfused float arithmetic; depth = mbf / disparity.  One warp per left keypoint.
// ------------------------------------------------------------------------------------------------------
struct StereoPyr {
    const uint8_t* interior[ORBX_MAX_LEVELS];
    int pitch[ORBX_MAX_LEVELS];
    int w[ORBX_MAX_LEVELS];
    float scale[ORBX_MAX_LEVELS], invScale[ORBX_MAX_LEVELS];
};

__global__ void __launch_bounds__(256) k_stereo_refine(int nl, const float* __restrict__ lx, const float* __restrict__ ly,
                                                       const int32_t* __restrict__ loct, const float* __restrict__ rx,
                                                       const int32_t* __restrict__ bestR, StereoPyr PLv, StereoPyr PRv,
                                                       float mbf, float minD, float maxD, float* __restrict__ uRight,
                                                       float* __restrict__ depth, int32_t* __restrict__ sad)
{
    const int lane = threadIdx.x & 31;
    const int iL = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (iL >= nl) return;
    float outU = -1.0f, outD = -1.0f;
    int outS = -1;
    const int br = bestR[iL];
    if (br >= 0) {
        const int oct = loct[iL];
        const float uL = lx[iL];
        const float sF = PLv.invScale[oct];
        const float scaleduL = roundf(__fmul_rn(uL, sF));
        const float scaledvL = roundf(__fmul_rn(ly[iL], sF));
        const float scaleduR0 = roundf(__fmul_rn(rx[br], sF));
        const float iniu = scaleduR0, endu = __fadd_rn(scaleduR0, 11.0f);  // scaleduR0 + L - w, scaleduR0 + L + w + 1
        if (!(iniu < 0 || endu >= (float)PRv.w[oct])) {
            const int pl = PLv.pitch[oct], pr = PRv.pitch[oct];
            const uint8_t* cL = PLv.interior[oct] + (long long)(int)scaledvL * pl + (int)scaleduL;
            const uint8_t* cR = PRv.interior[oct] + (long long)(int)scaledvL * pr + (int)scaleduR0;
            const int centerL = cL[0];
            int acc[11];
#pragma unroll
            for (int s = 0; s < 11; s++) acc[s] = 0;
            for (int p = lane; p < 121; p += 32) {
                const int dy = p / 11 - 5, dx = p - (p / 11) * 11 - 5;
                const int a = (int)cL[dy * pl + dx] - centerL;
                const uint8_t* rrow = cR + dy * pr + dx;
#pragma unroll
                for (int s = 0; s < 11; s++) {
                    const int b = (int)rrow[s - 5] - (int)cR[s - 5];
                    acc[s] += abs(a - b);
                }
            }
            int bestDist = 0x7fffffff, bestinc = 0;
            float vd[11];
#pragma unroll
            for (int s = 0; s < 11; s++) {
                const int d = __reduce_add_sync(0xffffffffu, acc[s]);
                vd[s] = (float)d;
                if (d < bestDist) { bestDist = d; bestinc = s - 5; }
            }
            if (bestinc != -5 && bestinc != 5) {
                float dist1 = 0, dist2 = 0, dist3 = 0;
#pragma unroll
                for (int s = 1; s < 10; s++)
                    if (s - 5 == bestinc) { dist1 = vd[s - 1]; dist2 = vd[s]; dist3 = vd[s + 1]; }
                const float den = __fmul_rn(2.0f, __fsub_rn(__fadd_rn(dist1, dist3), __fmul_rn(2.0f, dist2)));
                const float deltaR = __fdiv_rn(__fsub_rn(dist1, dist3), den);
                if (!(deltaR < -1 || deltaR > 1)) {
                    float bestuR = __fmul_rn(PLv.scale[oct], __fadd_rn(__fadd_rn(scaleduR0, (float)bestinc), deltaR));
                    float disparity = __fsub_rn(uL, bestuR);
                    if (disparity >= minD && disparity < maxD) {
                        if (disparity <= 0) {
                            disparity = 0.01f;                                           // (float)0.01
                            bestuR = __double2float_rn(__dsub_rn((double)uL, 0.01));    // float - double literal
                        }
                        outD = __fdiv_rn(mbf, disparity);
                        outU = bestuR;
                        outS = bestDist;
                    }
                }
            }
        }
    }
    if (lane == 0) { uRight[iL] = outU; depth[iL] = outD; sad[iL] = outS; }
}

// Outlier removal (src/Frame.cc:868-884).  In THIS reference the sort / median / reject block sits INSIDE the
// per-keypoint loop (upstream ORB-SLAM2 has it after the loop): after every accepted match the running list of (SAD, iL)
// pairs is sorted, median = element size/2, and every match so far with SAD >= 1.5f * 1.4f * median is rejected for good.
// A match inserted as the (i+1)-th is therefore rejected iff SAD_i >= min over t > i of thDist_t, where thDist_t comes from
// the median of the first t accepted matches (the block is idempotent on iterations that add nothing, and a no-op while
// the list is empty: its loop does not run, the out-of-range read of element 0 is never used).
//   k_stereo_medians: every CTA compacts the accepted SADs (iL order) into shared memory; warp per prefix length t finds
//                     the median VALUE by bisection on the value (counting entries <= v across the lanes, 17 steps).
//   k_stereo_reject:  suffix minimum of thDist (one CTA), then the per-match test.  mode 1 = upstream placement: every
//                     match is tested against thDist of the full list only.
__global__ void __launch_bounds__(256) k_stereo_medians(int nl, const int32_t* __restrict__ sad, float* __restrict__ thDist,
                                                        int32_t* __restrict__ entryIdx, int32_t* __restrict__ nEntries)
{
    extern __shared__ int32_t sE[];  // compacted SADs, nl ints
    __shared__ int sWarpCnt[8], sBase;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) sBase = 0;
    __syncthreads();
    for (int i0 = 0; i0 < nl; i0 += 256) {
        const int i = i0 + threadIdx.x;
        const int d = i < nl ? sad[i] : -1;
        const unsigned bal = __ballot_sync(0xffffffffu, d >= 0);
        if (lane == 0) sWarpCnt[warp] = __popc(bal);
        __syncthreads();
        int off = sBase;
        for (int w = 0; w < warp; w++) off += sWarpCnt[w];
        if (d >= 0) {
            const int pos = off + __popc(bal & ((1u << lane) - 1u));
            sE[pos] = d;
            if (blockIdx.x == 0) entryIdx[pos] = i;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            int tot = 0;
            for (int w = 0; w < 8; w++) tot += sWarpCnt[w];
            sBase += tot;
        }
        __syncthreads();
    }
    const int n = sBase;
    if (blockIdx.x == 0 && threadIdx.x == 0) *nEntries = n;
    const int t = blockIdx.x * 8 + warp + 1;  // prefix length
    if (t > n) return;
    const int target = t / 2 + 1;             // smallest v with #(e <= v) >= target is element t/2 of the sorted prefix
    int lo = 0, hi = 1 << 17;                 // SAD <= 121 * 510 < 2^16
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        int c = 0;
        for (int j = lane; j < t; j += 32) c += sE[j] <= mid;
        c = __reduce_add_sync(0xffffffffu, c);
        if (c >= target) hi = mid; else lo = mid + 1;
    }
    if (lane == 0) thDist[t - 1] = __fmul_rn(__fmul_rn(1.5f, 1.4f), (float)lo);
}

__global__ void __launch_bounds__(1024) k_stereo_reject(const int32_t* __restrict__ sad, const float* __restrict__ thDist,
                                                        const int32_t* __restrict__ entryIdx, const int32_t* __restrict__ nEntries,
                                                        int mode, float* uRight, float* depth)
{
    __shared__ float sMin[1024];
    const int n = *nEntries;
    if (n == 0) return;
    const int per = (n + 1023) / 1024;
    const int b = threadIdx.x * per, e = min(b + per, n);
    float m = 3.0e38f;
    for (int i = e - 1; i >= b; i--) m = fminf(m, thDist[i]);
    sMin[threadIdx.x] = m;
    __syncthreads();
    for (int s = 1; s < 1024; s <<= 1) {   // inclusive suffix minimum over the chunk minima
        const float o = threadIdx.x + s < 1024 ? sMin[threadIdx.x + s] : 3.0e38f;
        __syncthreads();
        sMin[threadIdx.x] = fminf(sMin[threadIdx.x], o);
        __syncthreads();
    }
    float run = threadIdx.x + 1 < 1024 ? sMin[threadIdx.x + 1] : 3.0e38f;  // minimum over the later chunks
    const float last = thDist[n - 1];
    for (int i = e - 1; i >= b; i--) {
        run = fminf(run, thDist[i]);       // min over t > i of thDist_t (thDist[i] belongs to prefix length i + 1)
        const int iL = entryIdx[i];
        const float th = mode == 1 ? last : run;
        if (!((float)sad[iL] < th)) { uRight[iL] = -1.0f; depth[iL] = -1.0f; }
    }
}

// ------------------------------------------------------------------------------------------------------
// K8: DBoW2 vocabulary descent (TemplatedVocabulary::transform, Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1218-1259).
// One warp per descriptor walks the k-ary tree: the lanes take the children of the current node (k = 10 in ORBvoc),
// compute their Hamming distances and the warp keeps the first minimum in child order (strict <).
// ------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_bow_transform(int L, const int32_t* __restrict__ childStart,
                                                       const int32_t* __restrict__ childIds, const uint8_t* __restrict__ nodeDesc,
                                                       const int32_t* __restrict__ wordId, const double* __restrict__ weight,
                                                       const uint8_t* __restrict__ desc, int n, int levelsup,
                                                       int32_t* __restrict__ outWord, int32_t* __restrict__ outNode,
                                                       double* __restrict__ outWeight)
{
    const int lane = threadIdx.x & 31;
    const int f = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (f >= n) return;
    const Desc D = load_desc(desc + 32 * (size_t)f);
    const int nidLevel = L - levelsup;
    int nid = 0, cur = 0, level = 0;
    for (;;) {
        ++level;
        const int cs = childStart[cur], ce = childStart[cur + 1];
        uint32_t best = 0xFFFFFFFFu;  // dist << 22 | child position
        for (int c = cs + lane; c < ce; c += 32) {
            const int id = childIds[c];
            const int d = hamming(D, load_desc(nodeDesc + 32 * (size_t)id));
            best = min(best, ((uint32_t)d << 22) | (uint32_t)(c - cs));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
        cur = childIds[cs + (int)(best & ((1u << 22) - 1))];
        if (level == nidLevel) nid = cur;
        if (childStart[cur + 1] <= childStart[cur]) break;  // leaf
    }
    if (lane == 0) { outWord[f] = wordId[cur]; outNode[f] = nid; outWeight[f] = weight[cur]; }
}

// ------------------------------------------------------------------------------------------------------
// K9: MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:295-360), batched: one warp per map point.  For each
// observation descriptor i the median of its distances to the set is the value at sorted position (int)(0.5*(N-1));
// distances are in [0, 256], so the order statistic comes from a 257-bin histogram in shared memory instead of a sort.
// ------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_distinctive(int nSets, const int32_t* __restrict__ setStart,
                                                     const uint8_t* __restrict__ desc, int32_t* __restrict__ bestIdx)
{
    __shared__ int hist[4][288];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int s = blockIdx.x * 4 + wid;
    if (s >= nSets) return;
    const int b = setStart[s], N = setStart[s + 1] - b;
    if (N <= 0) { if (lane == 0) bestIdx[s] = -1; return; }
    const int k = (int)(0.5 * (N - 1));
    int* H = hist[wid];
    int bestMedian = 0x7fffffff, best = 0;
    for (int i = 0; i < N; i++) {
#pragma unroll
        for (int t = 0; t < 9; t++) H[lane * 9 + t] = 0;
        __syncwarp();
        const Desc D = load_desc(desc + 32 * (size_t)(b + i));
        for (int j = lane; j < N; j += 32) atomicAdd(&H[hamming(D, load_desc(desc + 32 * (size_t)(b + j)))], 1);
        __syncwarp();
        int mine = 0;
#pragma unroll
        for (int t = 0; t < 9; t++) mine += H[lane * 9 + t];
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        // the lane whose bins contain sorted position k
        const bool has = (incl - mine) <= k && k < incl;
        int median = 0;
        if (has) {
            int cum = incl - mine;
            for (int t = 0; t < 9; t++) {
                cum += H[lane * 9 + t];
                if (k < cum) { median = lane * 9 + t; break; }
            }
        }
        median = __reduce_max_sync(0xffffffffu, has ? median : 0);
        if (median < bestMedian) { bestMedian = median; best = i; }
        __syncwarp();
    }
    if (lane == 0) bestIdx[s] = best;
}

// ------------------------------------------------------------------------------------------------------
// K7d: SearchByBoW(KF, F).  A frame feature belongs to exactly one vocabulary node (DBoW2 FeatureVector invariant),
// so the order-dependent "skip frame features already matched" only couples KF features of the same node: one warp
// per shared node walks its KF features in order, lanes scan the node's frame features.
// ------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_bow_nodes(const uint8_t* __restrict__ kfDesc, const float* __restrict__ kfAngle,
                                                   const uint8_t* __restrict__ kfHasMp, int nNodesKF,
                                                   const int32_t* __restrict__ kfNodeId, const int32_t* __restrict__ kfNodeStart,
                                                   const int32_t* __restrict__ kfNodeFeat, const uint8_t* __restrict__ fDesc,
                                                   const float* __restrict__ fAngle, int nNodesF,
                                                   const int32_t* __restrict__ fNodeId, const int32_t* __restrict__ fNodeStart,
                                                   const int32_t* __restrict__ fNodeFeat, float nnratio,
                                                   const uint8_t* __restrict__ fHasMp, int strictLow,
                                                   int32_t* outMatch, int32_t* binOf)
{
    const int lane = threadIdx.x & 31;
    const int a = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (a >= nNodesKF) return;
    // locate the same node id in the frame's feature vector
    const int id = kfNodeId[a];
    int lo = 0, hi = nNodesF;
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (fNodeId[mid] < id) lo = mid + 1; else hi = mid; }
    if (lo >= nNodesF || fNodeId[lo] != id) return;
    const int fs = fNodeStart[lo], fe = fNodeStart[lo + 1];
    const float factor = 0.0833333358f;  // HISTO_LENGTH / 360.0f
    for (int ik = kfNodeStart[a]; ik < kfNodeStart[a + 1]; ik++) {
        const int realIdxKF = kfNodeFeat[ik];
        if (!kfHasMp[realIdxKF]) continue;
        const Desc D = load_desc(kfDesc + 32 * (size_t)realIdxKF);
        Top2 t;
        top2_init(t);
        for (int jf = fs + lane; jf < fe; jf += 32) {
            const int realIdxF = fNodeFeat[jf];
            if (outMatch[realIdxF] >= 0) continue;
            if (fHasMp && !fHasMp[realIdxF]) continue;
            const int dist = hamming(D, load_desc(fDesc + 32 * (size_t)realIdxF));
            top2_push(t, ((uint32_t)dist << 20) | (uint32_t)(jf - fs), realIdxF);
        }
        top2_warp_reduce(t);
        if (t.p1 >= 0) {
            const int bestDist1 = t.k1 >> 20;
            const int bestDist2 = t.p2 >= 0 ? (int)(t.k2 >> 20) : 256;
            if ((strictLow ? bestDist1 < TH_LOW : bestDist1 <= TH_LOW) && (float)bestDist1 < __fmul_rn(nnratio, (float)bestDist2)) {
                if (lane == 0) {
                    outMatch[t.p1] = realIdxKF;
                    float rot = __fsub_rn(kfAngle[realIdxKF], fAngle[t.p1]);
                    if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
                    int bin = (int)roundf(__fmul_rn(rot, factor));
                    if (bin == HISTO_LENGTH) bin = 0;
                    binOf[t.p1] = bin;
                }
            }
        }
        __syncwarp();
        __threadfence_block();
    }
}

// histogram of rotation bins, ComputeThreeMaxima, removal of the matches outside the three main bins, final count
__global__ void __launch_bounds__(256) k_bow_rotation(int nf, int32_t* outMatch, const int32_t* __restrict__ binOf,
                                                      int checkOrientation, int* nmatches)
{
    __shared__ int hist[HISTO_LENGTH];
    __shared__ int keepBin[HISTO_LENGTH];
    __shared__ int total;
    if (threadIdx.x < HISTO_LENGTH) { hist[threadIdx.x] = 0; keepBin[threadIdx.x] = 0; }
    if (threadIdx.x == 0) total = 0;
    __syncthreads();
    int mine = 0;
    for (int f = threadIdx.x; f < nf; f += blockDim.x)
        if (outMatch[f] >= 0) { mine++; atomicAdd(&hist[binOf[f]], 1); }
    if (mine) atomicAdd(&total, mine);
    __syncthreads();
    if (!checkOrientation) {
        if (threadIdx.x == 0) *nmatches = total;
        return;
    }
    if (threadIdx.x == 0) {
        int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            const int s = hist[i];
            if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
            else if (s > max3) { max3 = s; ind3 = i; }
        }
        if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
        else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) ind3 = -1;
        int kept = 0;
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i == ind1 || i == ind2 || i == ind3) { keepBin[i] = 1; kept += hist[i]; }
        *nmatches = kept;
    }
    __syncthreads();
    for (int f = threadIdx.x; f < nf; f += blockDim.x)
        if (outMatch[f] >= 0 && !keepBin[binOf[f]]) outMatch[f] = -1;
}

// (KeyFrame, KeyFrame) overload: turn "feature of KF2 -> feature of KF1" into vpMatches12 (feature of KF1 -> feature of KF2)
__global__ void k_bow_scatter12(int n2, const int32_t* __restrict__ match21, int32_t* __restrict__ out12)
{
    const int i2 = blockIdx.x * blockDim.x + threadIdx.x;
    if (i2 < n2 && match21[i2] >= 0) out12[match21[i2]] = i2;
}


// ------------------------------------------------------------------------------------------------------
// K7e: claim-free windowed search -- Fuse(KF, MapPoints, th) src/ORBmatcher.cc:1067-1222 (gate 1: reprojection chi2),
// Fuse(KF, Scw, ...) :1225-1357 and the two passes of SearchBySim3 :1361-1602 (gate 0).  One warp per query, one lane
// per grid cell of KeyFrame::GetFeaturesInArea's window (src/KeyFrame.cc:623-658); first minimum in visiting order.
// ------------------------------------------------------------------------------------------------------
struct BestArgs {
    const float* kx; const float* ky; const int32_t* koct; const float* uright; const uint8_t* kdesc;
    const int32_t* cellStart; const int32_t* cellItems;
    float minX, minY, wInv, hInv;
    int nq;
    const float* qx; const float* qy; const float* qr; const int32_t* qmin; const int32_t* qmax; const float* qxr;
    const uint8_t* qvalid; const uint8_t* qdesc;
    int gate, maxDist;
    const float* invSigma2;
    int32_t* bestIdx; int32_t* bestDist;
};

__global__ void __launch_bounds__(256) k_search_best(BestArgs A)
{
    const int lane = threadIdx.x & 31;
    const int m = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (m >= A.nq) return;
    int outIdx = -1, outDist = 256;
    if (A.qvalid[m]) {
        const float rr = A.qr[m], u = A.qx[m], v = A.qy[m];
        const int c0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(u, A.minX), rr), A.wInv)));
        const int c1 = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(u, A.minX), rr), A.wInv)));
        const int r0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(v, A.minY), rr), A.hInv)));
        const int r1 = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(v, A.minY), rr), A.hInv)));
        if (c0 < GRID_COLS && c1 >= 0 && r0 < GRID_ROWS && r1 >= 0) {
            const int minLevel = A.qmin[m], maxLevel = A.qmax[m];
            const Desc D = load_desc(A.qdesc + 32 * (size_t)m);
            const float ur = A.qxr ? A.qxr[m] : 0.f;
            uint32_t bestKey = 0xFFFFFFFFu;
            int best = -1;
            const int nr = r1 - r0 + 1, ncell = (c1 - c0 + 1) * nr;
            for (int qc = lane; qc < ncell; qc += 32) {
                const int ix = c0 + qc / nr, iy = r0 + qc % nr;
                const int c = ix * GRID_ROWS + iy;
                const int s = A.cellStart[c], e = A.cellStart[c + 1];
                for (int k = s; k < e; k++) {
                    const int idx = A.cellItems[k];
                    const float kpx = A.kx[idx], kpy = A.ky[idx];
                    if (!(fabsf(__fsub_rn(kpx, u)) < rr && fabsf(__fsub_rn(kpy, v)) < rr)) continue;
                    const int oct = A.koct[idx];
                    if (oct < minLevel || oct > maxLevel) continue;
                    if (A.gate == 1) {
                        const float ex = __fsub_rn(u, kpx), ey = __fsub_rn(v, kpy);
                        float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
                        const float kpr = A.uright ? A.uright[idx] : -1.f;
                        double lim = 5.99;
                        if (kpr >= 0) {
                            const float er = __fsub_rn(ur, kpr);
                            e2 = __fadd_rn(e2, __fmul_rn(er, er));
                            lim = 7.8;
                        }
                        if ((double)__fmul_rn(e2, A.invSigma2[oct]) > lim) continue;
                    }
                    const int dist = hamming(D, load_desc(A.kdesc + 32 * (size_t)idx));
                    const uint32_t key = ((uint32_t)dist << 23) | ((uint32_t)qc << 11) | (uint32_t)(k - s);
                    if (key < bestKey) { bestKey = key; best = idx; }
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const uint32_t ok = __shfl_xor_sync(0xffffffffu, bestKey, o);
                const int ob = __shfl_xor_sync(0xffffffffu, best, o);
                if (ok < bestKey) { bestKey = ok; best = ob; }
            }
            if (best >= 0) {
                outDist = (int)(bestKey >> 23);
                if (outDist <= A.maxDist) outIdx = best;
            }
        }
    }
    if (lane == 0) { A.bestIdx[m] = outIdx; A.bestDist[m] = outDist; }
}


// ------------------------------------------------------------------------------------------------------
// K7f: SearchForInitialization src/ORBmatcher.cc:539-683.  The reference loop is sequential in i1: a candidate i2 is
// skipped when an EARLIER F1 feature already holds it with a distance <= this one (vMatchedDistance), and a later,
// better match steals it.  result(i1) depends only on the results of j < i1, so -- as in k_proj_search -- the
// sequential answer is the unique fixed point of "recompute every i1 against the previous sweep's results".  One CTA
// (the problem is a few hundred level-0 features x ~50 candidates), one warp per i1, one lane per grid cell; the
// previous sweep's accepted matches hang off their F2 feature as linked lists (head / next).
// ------------------------------------------------------------------------------------------------------
struct InitArgs {
    int n1; const int32_t* oct1; const float* angle1; const uint8_t* desc1; const float* prevx; const float* prevy;
    int n2; const float* kx; const float* ky; const int32_t* koct; const float* kangle; const uint8_t* kdesc;
    const int32_t* cellStart; const int32_t* cellItems;
    float minX, minY, wInv, hInv, window, nnratio;
    int checkOri;
    int32_t* resA; int32_t* resB;   // [n1] accepted F2 feature or -1 (double buffered across sweeps)
    int32_t* distA; int32_t* distB; // [n1] Hamming distance of the accepted match
    int32_t* head; int32_t* headB; int32_t* next;   // [n2] x 2 (the list heads the sweep reads / the next sweep's), [n1]
    int32_t* out12; int* nmatches; int* sweeps; int* misc;
};

__device__ int init_one(const InitArgs& A, int i1, const int32_t* distPrev, int lane, int* distOut)
{
    if (A.oct1[i1] > 0) return -1;
    const float rr = A.window, x = A.prevx[i1], y = A.prevy[i1];
    const int c0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, A.minX), rr), A.wInv)));
    if (c0 >= GRID_COLS) return -1;
    const int c1 = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, A.minX), rr), A.wInv)));
    if (c1 < 0) return -1;
    const int r0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, A.minY), rr), A.hInv)));
    if (r0 >= GRID_ROWS) return -1;
    const int r1 = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, A.minY), rr), A.hInv)));
    if (r1 < 0) return -1;
    const Desc D = load_desc(A.desc1 + 32 * (size_t)i1);
    Top2 t;
    top2_init(t);
    const int nr = r1 - r0 + 1, ncell = (c1 - c0 + 1) * nr;
    for (int qc = lane; qc < ncell; qc += 32) {
        const int ix = c0 + qc / nr, iy = r0 + qc % nr;
        const int c = ix * GRID_ROWS + iy;
        const int s = A.cellStart[c], e = A.cellStart[c + 1];
        for (int k = s; k < e; k++) {
            const int i2 = A.cellItems[k];
            if (A.koct[i2] != 0) continue;
            if (!(fabsf(__fsub_rn(A.kx[i2], x)) < rr && fabsf(__fsub_rn(A.ky[i2], y)) < rr)) continue;
            const int d = hamming(D, load_desc(A.kdesc + 32 * (size_t)i2));
            bool blocked = false;  // vMatchedDistance[i2] <= dist, as left by the features j < i1
            for (int j = A.head[i2]; j >= 0; j = A.next[j])
                if (j < i1 && distPrev[j] <= d) { blocked = true; break; }
            if (blocked) continue;
            top2_push(t, ((uint32_t)d << 23) | ((uint32_t)qc << 11) | (uint32_t)(k - s), i2);
        }
    }
    top2_warp_reduce(t);
    if (t.p1 < 0) return -1;
    const int bestDist = t.k1 >> 23;
    const int bestDist2 = t.p2 >= 0 ? (int)(t.k2 >> 23) : 0x7fffffff;
    if (bestDist > TH_LOW) return -1;
    if (!((float)bestDist < __fmul_rn((float)bestDist2, A.nnratio))) return -1;
    *distOut = bestDist;
    return t.p1;
}

static const int INIT_T = [] { const char* e = std::getenv("ORBX_INIT_T"); const int v = e ? std::atoi(e) : 0; return v >= 32 && v <= 1024 && v % 32 == 0 ? v : 1024; }();  // threads per CTA (grid barriers cost one arrival per CTA; measured 256 / 512 / 1024: 0.288 / 0.263 / 0.242 ms per call)

__global__ void __launch_bounds__(1024) k_init_search(InitArgs A)
{
    // cooperative grid (one warp per F1 feature and sweep, as many CTAs as level-0 features need): the sweeps are separated by
    // grid-wide barriers; flags, histogram and the kept bins live in A.misc
    cg::grid_group grid = cg::this_grid();
    int* changed = A.misc;             // [2], alternating by sweep
    int* hist = A.misc + 8;            // [HISTO_LENGTH]
    int* keepBin = A.misc + 8 + 32;    // [HISTO_LENGTH]
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x, gthreads = gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31, warp = gtid >> 5, nwarps = gthreads >> 5;
    int32_t* resPrev = A.resA;
    int32_t* resNext = A.resB;
    int32_t* distPrev = A.distA;
    int32_t* distNext = A.distB;
    for (int i = gtid; i < A.n2; i += gthreads) A.head[i] = -1;   // no accepted matches yet: empty lists
    for (int i = gtid; i < A.n1; i += gthreads) { resPrev[i] = -1; A.next[i] = -1; distPrev[i] = 0; }
    if (gtid < HISTO_LENGTH) hist[gtid] = 0;
    if (gtid == 0) { changed[0] = 0; changed[1] = 0; *A.nmatches = 0; }
    grid.sync();
    int sweep = 0;
    for (;; sweep++) {
        int* chg = changed + (sweep & 1);
        for (int i1 = warp; i1 < A.n1; i1 += nwarps) {
            int d = 0;
            const int r = init_one(A, i1, distPrev, lane, &d);
            if (lane == 0) {
                resNext[i1] = r;
                distNext[i1] = d;
                if (r != resPrev[i1]) *chg = 1;
            }
        }
        // the next sweep's list heads are prepared in the OTHER array (this sweep's lists are still being read elsewhere)
        for (int i = gtid; i < A.n2; i += gthreads) A.headB[i] = -1;
        grid.sync();
        if (*chg == 0) break;
        if (gtid == 0) changed[(sweep + 1) & 1] = 0;
        for (int j = gtid; j < A.n1; j += gthreads) {
            const int r = resNext[j];
            A.next[j] = r >= 0 ? atomicExch(&A.headB[r], j) : -1;
        }
        grid.sync();
        int32_t* t = resPrev; resPrev = resNext; resNext = t;
        t = distPrev; distPrev = distNext; distNext = t;
        t = A.head; A.head = A.headB; A.headB = t;
    }
    // resNext == resPrev == the sequential result.  vnMatches21[i2] = the LAST accepted i1 (later matches steal).
    // (A.headB was reset to -1 everywhere before the last barrier or at start-up and has not been written since)
    for (int j = gtid; j < A.n1; j += gthreads) {
        const int r = resNext[j];
        if (r >= 0) {
            atomicMax(&A.headB[r], j);
            if (A.checkOri) atomicAdd(&hist[rot_bin(A.angle1[j], A.kangle[r])], 1);   // stolen matches stay in rotHist
        }
    }
    grid.sync();
    if (gtid == 0) {
        int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < HISTO_LENGTH; i++) {
            const int sN = hist[i];
            if (sN > max1) { max3 = max2; max2 = max1; max1 = sN; ind3 = ind2; ind2 = ind1; ind1 = i; }
            else if (sN > max2) { max3 = max2; max2 = sN; ind3 = ind2; ind2 = i; }
            else if (sN > max3) { max3 = sN; ind3 = i; }
        }
        if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
        else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) ind3 = -1;
        for (int i = 0; i < HISTO_LENGTH; i++) keepBin[i] = !A.checkOri || (i == ind1 || i == ind2 || i == ind3);
        *A.sweeps = sweep + 1;
    }
    grid.sync();
    int mine = 0;
    for (int j = gtid; j < A.n1; j += gthreads) {
        const int r = resNext[j];
        int o = -1;
        if (r >= 0 && A.headB[r] == j && keepBin[A.checkOri ? rot_bin(A.angle1[j], A.kangle[r]) : 0]) { o = r; mine++; }
        A.out12[j] = o;
    }
    if (mine) atomicAdd(A.nmatches, mine);
}


// ------------------------------------------------------------------------------------------------------
// K7g: SearchForTriangulation src/ORBmatcher.cc:850-1056 (+ CheckDistEpipolarLine :186-213).  As in SearchByBoW the
// vbMatched2 claims only couple features of the same vocabulary node: one warp per shared node walks KF1's features in
// order, lanes scan KF2's.  `dist > bestDist` skips, so among the candidates passing the gates the minimum distance
// wins and ties go to the LAST one: key = dist << 20 | (0xFFFFF - position).
// ------------------------------------------------------------------------------------------------------
struct TriArgs {
    const float* x1; const float* y1; const float* angle1; const float* ur1; const uint8_t* desc1; const uint8_t* hasMp1;
    int nNodes1; const int32_t* nodeId1; const int32_t* nodeStart1; const int32_t* nodeFeat1;
    const float* x2; const float* y2; const int32_t* oct2; const float* angle2; const float* ur2; const uint8_t* desc2;
    const uint8_t* hasMp2;
    int nNodes2; const int32_t* nodeId2; const int32_t* nodeStart2; const int32_t* nodeFeat2;
    float F[9];
    float ex, ey;
    const float* scale2; const float* sigma2_2;
    int onlyStereo;
    int32_t* matched2;  // [n2] 0/1
    int32_t* out12;     // [n1]
    int32_t* binOf;     // [n1]
};

__global__ void __launch_bounds__(128) k_triang_nodes(TriArgs A)
{
    const int lane = threadIdx.x & 31;
    const int a = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (a >= A.nNodes1) return;
    const int id = A.nodeId1[a];
    int lo = 0, hi = A.nNodes2;
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (A.nodeId2[mid] < id) lo = mid + 1; else hi = mid; }
    if (lo >= A.nNodes2 || A.nodeId2[lo] != id) return;
    const int fs = A.nodeStart2[lo], fe = A.nodeStart2[lo + 1];
    for (int ik = A.nodeStart1[a]; ik < A.nodeStart1[a + 1]; ik++) {
        const int idx1 = A.nodeFeat1[ik];
        if (A.hasMp1[idx1]) continue;
        const bool bStereo1 = A.ur1[idx1] >= 0;
        if (A.onlyStereo && !bStereo1) continue;
        const Desc D = load_desc(A.desc1 + 32 * (size_t)idx1);
        // epipolar line of kp1 in image 2: l = x1^T F12
        const float px = A.x1[idx1], py = A.y1[idx1];
        const float la = __fadd_rn(__fadd_rn(__fmul_rn(px, A.F[0]), __fmul_rn(py, A.F[3])), A.F[6]);
        const float lb = __fadd_rn(__fadd_rn(__fmul_rn(px, A.F[1]), __fmul_rn(py, A.F[4])), A.F[7]);
        const float lc = __fadd_rn(__fadd_rn(__fmul_rn(px, A.F[2]), __fmul_rn(py, A.F[5])), A.F[8]);
        const float den = __fadd_rn(__fmul_rn(la, la), __fmul_rn(lb, lb));
        uint32_t bestKey = 0xFFFFFFFFu;
        int best = -1;
        for (int jf = fs + lane; jf < fe; jf += 32) {
            const int idx2 = A.nodeFeat2[jf];
            if (A.matched2[idx2] || A.hasMp2[idx2]) continue;
            const bool bStereo2 = A.ur2[idx2] >= 0;
            if (A.onlyStereo && !bStereo2) continue;
            const int dist = hamming(D, load_desc(A.desc2 + 32 * (size_t)idx2));
            if (dist > TH_LOW) continue;
            const float qx = A.x2[idx2], qy = A.y2[idx2];
            const int oct = A.oct2[idx2];
            if (!bStereo1 && !bStereo2) {
                const float dex = __fsub_rn(A.ex, qx), dey = __fsub_rn(A.ey, qy);
                if (__fadd_rn(__fmul_rn(dex, dex), __fmul_rn(dey, dey)) < __fmul_rn(100.0f, A.scale2[oct])) continue;
            }
            const float num = __fadd_rn(__fadd_rn(__fmul_rn(la, qx), __fmul_rn(lb, qy)), lc);
            if (den == 0) continue;
            const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
            if (!((double)dsqr < 3.84 * (double)A.sigma2_2[oct])) continue;
            const uint32_t key = ((uint32_t)dist << 20) | (0xFFFFFu - (uint32_t)(jf - fs));
            if (key < bestKey) { bestKey = key; best = idx2; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const uint32_t ok = __shfl_xor_sync(0xffffffffu, bestKey, o);
            const int ob = __shfl_xor_sync(0xffffffffu, best, o);
            if (ok < bestKey) { bestKey = ok; best = ob; }
        }
        if (best >= 0 && lane == 0) {
            A.out12[idx1] = best;
            A.matched2[best] = 1;
            A.binOf[idx1] = rot_bin(A.angle1[idx1], A.angle2[best]);
        }
        __syncwarp();
        __threadfence_block();
    }
}

}  // namespace orbx

using namespace orbx;

struct orbm_context {
    int device = 0;
    cudaStream_t stream = nullptr;
    int numSMs = 0;
    // growable device scratch
    char* arena = nullptr;
    size_t arenaBytes = 0;
    // page-locked mirror of the arena: the host arrays of one call are gathered here (same offsets) and go to the device in
    // ONE copy (a search uploads ~20 small pageable arrays; one cudaMemcpyAsync each costs more than the search kernel)
    char* hStage = nullptr;
    size_t hStageBytes = 0;
    size_t upLo = ~size_t(0), upHi = 0;
    cudaEvent_t evUp = nullptr;   // the last flush has read hStage
    // partial results of the brute-force kernel
    int32_t* partials = nullptr;
    size_t partialsBytes = 0;
    void* pinned = nullptr;
    size_t pinnedBytes = 0;
    // sharded search: exchange buffer mapped by every peer (orbm_exchange_*)
    unsigned long long* xchg = nullptr;
    size_t xchgBytes = 0;
    unsigned long long* xchgPeers[16] = {};
    bool xchgOpenedIpc[16] = {};
    bool xchgConnected = false;
    int world = 0, rank = 0, nqMax = 0;
    uint32_t epoch = 0;
    int* xchgTimeout = nullptr;  // mapped pinned host int, written by the kernel when a peer never answered
    // tensor-core search: +-1 expansion of the queries of the current call
    int8_t* qExp = nullptr;
    size_t qExpBytes = 0;
    bool tcAttrSet = false;      // >48 KB dynamic shared memory opt-in of k_bruteforce_tc / tc2 done on this context's device
    int stereoOutlierMode = 0;   // 0 = this reference (block inside the keypoint loop), 1 = upstream ORB-SLAM2 (after the loop)
};

namespace {

struct Carver {
    size_t off = 0;
    size_t take(size_t bytes) { size_t o = off; off += (bytes + 255) & ~size_t(255); return o; }
};

int ensure_arena(orbm_context* m, size_t bytes)
{
    // every staging function calls this before its first up(): a call that returned early between up() and flush_up()
    // must not leave its range behind for the next call to flush
    m->upLo = ~size_t(0); m->upHi = 0;
    if (bytes <= m->arenaBytes) return ORBX_OK;
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    cudaFree(m->arena); m->arena = nullptr; m->arenaBytes = 0;
    ORBX_CUDA_TRY(cudaMalloc(&m->arena, bytes));
    if (!m->evUp) ORBX_CUDA_TRY(cudaEventCreateWithFlags(&m->evUp, cudaEventDisableTiming));
    m->arenaBytes = bytes;
    return ORBX_OK;
}
// The page-locked mirror covers only the span that up() stages (offsets [0, need) of the arena): entry points that copy
// straight from caller memory (orbm_bruteforce with a large host database, undistort, assign_grid) never pin anything.
int ensure_stage(orbm_context* m, size_t need)
{
    if (need <= m->hStageBytes) return ORBX_OK;
    const size_t nb = std::min(m->arenaBytes, std::max(std::max(need, 2 * m->hStageBytes), size_t(256) << 10));
    char* p = nullptr;
    ORBX_CUDA_TRY(cudaHostAlloc((void**)&p, nb, cudaHostAllocDefault));
    if (m->hStage) {
        if (m->evUp) cudaEventSynchronize(m->evUp);  // the last flush no longer reads the old mirror
        if (m->upHi > m->upLo) std::memcpy(p + m->upLo, m->hStage + m->upLo, m->upHi - m->upLo);  // arrays staged earlier in this call
        cudaFreeHost(m->hStage);
    }
    m->hStage = p;
    m->hStageBytes = nb;
    return ORBX_OK;
}
int ensure_partials(orbm_context* m, size_t bytes)
{
    if (bytes <= m->partialsBytes) return ORBX_OK;
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    cudaFree(m->partials); m->partials = nullptr; m->partialsBytes = 0;
    ORBX_CUDA_TRY(cudaMalloc(&m->partials, bytes));
    m->partialsBytes = bytes;
    return ORBX_OK;
}

// Host -> device upload of one array.  Destinations inside the arena are gathered in the page-locked mirror and sent by
// flush_up() as one copy (everything between the lowest and the highest staged byte: the carve-up places a call's uploads
// next to each other; work areas that fall in between are overwritten with the mirror's stale bytes, which no kernel reads
// before writing).  Every function that stages calls flush_up() before its first launch.
template <typename T> int up(orbm_context* m, T* dst, const T* src, size_t n)
{
    if (!n) return ORBX_OK;
    const char* d = reinterpret_cast<const char*>(dst);
    const size_t bytes = n * sizeof(T);
    if (m->arena && d >= m->arena && d + bytes <= m->arena + m->arenaBytes) {
        const size_t off = (size_t)(d - m->arena);
        const int rc = ensure_stage(m, off + bytes);
        if (rc != ORBX_OK) return rc;
        if (m->upHi == 0 && m->evUp) ORBX_CUDA_TRY(cudaEventSynchronize(m->evUp));  // first array of a call: the previous flush is done
        std::memcpy(m->hStage + off, src, bytes);
        m->upLo = std::min(m->upLo, off);
        m->upHi = std::max(m->upHi, off + bytes);
        return ORBX_OK;
    }
    ORBX_CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, m->stream));
    return ORBX_OK;
}
int flush_up(orbm_context* m)
{
    if (m->upHi > m->upLo) {
        ORBX_CUDA_TRY(cudaMemcpyAsync(m->arena + m->upLo, m->hStage + m->upLo, m->upHi - m->upLo, cudaMemcpyHostToDevice, m->stream));
        ORBX_CUDA_TRY(cudaEventRecord(m->evUp, m->stream));
    }
    m->upLo = ~size_t(0); m->upHi = 0;
    return ORBX_OK;
}

}  // namespace

#include "orbx_match_tc.cuh"

// A database prepared for the tensor-core search: the +-1 byte expansion (256 B per row, rows padded to a multiple of 256
// with zeros) and its TMA descriptor.
struct orbm_db {
    int device = 0;
    int ndb = 0;
    long long rowsPad = 0;
    int8_t* exp = nullptr;
    CUtensorMap map;      // stages of 256 rows (k_bruteforce_tc)
    CUtensorMap map128;   // stages of 128 rows per CTA (k_bruteforce_tc2, CTA pairs)
};

namespace {

// chunk partials of one prepared database into m->partials, same layout as bruteforce_partials
int bruteforce_partials_tc(orbm_context* m, const orbm_db* d, const uint8_t* dq, int nq, int idxOffset, int* nChunksOut)
{
    // default: one CTA per (128 queries, chunk).  ORBX_TC_PAIR=1 selects the CTA-pair kernel (cta_group::2, 256 queries per
    // pair; measured slower on B200, see orbx_match_tc.cuh) when there is more than one query tile.
    const char* pairEnv = std::getenv("ORBX_TC_PAIR");
    const bool pair = pairEnv && std::atoi(pairEnv) != 0 && nq > tc::BM;
    const int rowsPerTile = pair ? 2 * tc::BM : tc::BM;
    const int qTiles = (nq + rowsPerTile - 1) / rowsPerTile;
    const long long nqPad = (long long)qTiles * rowsPerTile;
    if ((size_t)nqPad * 256 > m->qExpBytes) {
        ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
        cudaFree(m->qExp); m->qExp = nullptr; m->qExpBytes = 0;
        ORBX_CUDA_TRY(cudaMalloc(&m->qExp, (size_t)nqPad * 256));
        m->qExpBytes = (size_t)nqPad * 256;
    }
    tc::k_expand_pm1<<<(unsigned)((nqPad * 32 + 255) / 256), 256, 0, m->stream>>>(dq, nq, nqPad, m->qExp);
    CUtensorMap mapA;
    int rc = tc::make_map(&mapA, m->qExp, nqPad, tc::BM);
    if (rc != ORBX_OK) return rc;
    // one CTA (or CTA pair) per SM (TPC): (query tiles) x (database chunks); chunks are multiples of the 256-row tile
    const int units = pair ? m->numSMs / 2 : m->numSMs;
    const int target = std::max(1, units / qTiles);
    long long chunk = (d->ndb + target - 1) / target;
    chunk = std::max<long long>(tc::BN, (chunk + tc::BN - 1) / tc::BN * tc::BN);
    chunk = std::min<long long>(chunk, 1 << 21);
    const int nChunks = (int)std::max<long long>(1, (d->ndb + chunk - 1) / chunk);
    const size_t pBytes = sizeof(int32_t) * 3 * (size_t)nChunks * nq;
    rc = ensure_partials(m, pBytes);
    if (rc != ORBX_OK) return rc;
    int32_t* pBest = m->partials;
    int32_t* pIdx = pBest + (size_t)nChunks * nq;
    int32_t* pSecond = pIdx + (size_t)nChunks * nq;
    if (!m->tcAttrSet) {  // the opt-in is per device / context, and contexts may live on several threads: one flag per context
        ORBX_CUDA_TRY(cudaFuncSetAttribute(tc::k_bruteforce_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::SMEM_BYTES));
        ORBX_CUDA_TRY(cudaFuncSetAttribute(tc::k_bruteforce_tc2, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::SMEM2_BYTES));
        m->tcAttrSet = true;
    }
    if (pair)
        tc::k_bruteforce_tc2<<<dim3(2 * nChunks, qTiles), tc::THREADS2, tc::SMEM2_BYTES, m->stream>>>(mapA, d->map128, nq, d->ndb, (int)chunk,
                                                                                                 idxOffset, pBest, pIdx, pSecond);
    else
        tc::k_bruteforce_tc<<<dim3(nChunks, qTiles), tc::THREADS, tc::SMEM_BYTES, m->stream>>>(mapA, d->map, nq, d->ndb, (int)chunk, idxOffset,
                                                                                           pBest, pIdx, pSecond);
    ORBX_CUDA_TRY(cudaGetLastError());
    *nChunksOut = nChunks;
    return ORBX_OK;
}

}  // namespace

// ------------------------------------------------------------------------------------------------------
// Scratch of orbm_bruteforce_noalloc (include/orbx.h): the per-chunk partial results of k_bruteforce live in this
// array, which is part of the loaded module, instead of in a cudaMalloc'ed buffer of a context.  It costs
// ORBM_NOALLOC_SCRATCH_BYTES (12 MiB) of device memory on every GPU that loads liborbx.so, from the load until the
// process exits, whether the entry point is ever called or not.
//
// In HOST code the name g_bfNoAllocScratch is only a handle for the symbol, not a device address: handing it to a
// kernel would fault (or, on a system where the GPU can reach host memory, scribble over host memory without any error).
// noalloc_setup() asks the runtime for the real address, once per device, with that device current.
// ------------------------------------------------------------------------------------------------------
static __device__ __align__(256) int32_t g_bfNoAllocScratch[3 * ORBM_NOALLOC_SCRATCH_SLOTS];
static_assert(sizeof(g_bfNoAllocScratch) == ORBM_NOALLOC_SCRATCH_BYTES, "header and scratch array disagree");

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noallocMutex;
int32_t* g_noallocPtr[NOALLOC_MAX_DEVICES] = {};  // device address of g_bfNoAllocScratch on each GPU (nullptr = not set up)
int g_noallocSMs[NOALLOC_MAX_DEVICES] = {};

// First call on `device` (which must be current): resolves the scratch symbol and caches the SM count.  No allocation.
int noalloc_setup(int device, int32_t** scratch, int* numSMs)
{
    std::lock_guard<std::mutex> lock(g_noallocMutex);
    if (!g_noallocPtr[device]) {
        void* p = nullptr;
        ORBX_CUDA_TRY(cudaGetSymbolAddress(&p, g_bfNoAllocScratch));
        int sms = 0;
        ORBX_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
        if (!p || sms <= 0) { set_error("orbm_bruteforce_noalloc: could not resolve the scratch symbol on device %d", device); return ORBX_ERR_CUDA; }
        g_noallocSMs[device] = sms;
        g_noallocPtr[device] = static_cast<int32_t*>(p);
    }
    *scratch = g_noallocPtr[device];
    *numSMs = g_noallocSMs[device];
    return ORBX_OK;
}
}  // namespace

extern "C" {

int orbm_descriptor_distance(const uint8_t* a, const uint8_t* b)
{
    int d = 0;
    for (int i = 0; i < 8; i++) {
        uint32_t x, y;
        memcpy(&x, a + 4 * i, 4);
        memcpy(&y, b + 4 * i, 4);
        d += __builtin_popcount(x ^ y);
    }
    return d;
}

int orbm_create(int device, orbm_context** out)
{
    if (!out) { set_error("orbm_create: null argument"); return ORBX_ERR_BAD_ARG; }
    int ndev = orbx_device_count();
    if (ndev <= 0) { if (ndev == 0) set_error("no CUDA device"); return ORBX_ERR_NO_DEVICE; }
    if (device < 0 || device >= ndev) { set_error("device %d out of range", device); return ORBX_ERR_BAD_ARG; }
    orbm_context* m = new orbm_context;
    m->device = device;
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&m->numSMs, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess) { set_error("orbm_create: %s", cudaGetErrorString(e)); delete m; return ORBX_ERR_CUDA; }
    *out = m;
    return ORBX_OK;
}

static void exchange_release(orbm_context* m);

void orbm_destroy(orbm_context* m)
{
    if (!m) return;
    cudaSetDevice(m->device);
    if (m->stream) cudaStreamSynchronize(m->stream);
    exchange_release(m);
    cudaFree(m->arena); cudaFree(m->partials); cudaFree(m->qExp);
    if (m->hStage) cudaFreeHost(m->hStage);
    if (m->evUp) cudaEventDestroy(m->evUp);
    if (m->pinned) cudaFreeHost(m->pinned);
    if (m->stream) cudaStreamDestroy(m->stream);
    delete m;
}

void* orbm_stream(orbm_context* m) { return m ? (void*)m->stream : nullptr; }

int orbm_sync(orbm_context* m)
{
    if (!m) { set_error("null context"); return ORBX_ERR_BAD_ARG; }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

// chunk partials of one database (shard) into m->partials: [best | idx | second][chunk][nq]
static int bruteforce_partials(orbm_context* m, const uint8_t* dq, int nq, const uint8_t* ddb, int ndb, int idxOffset, int* nChunksOut)
{
    const int qBlocks = (nq + BF_T * BF_RQ - 1) / (BF_T * BF_RQ);
    // enough chunks to fill the GPU ~4 CTAs per SM, each chunk a multiple of the tile and below 2^22 rows
    int target = std::max(1, (m->numSMs * 4) / qBlocks);
    int chunk = (ndb + target - 1) / target;
    chunk = std::max(BF_TILE, (chunk + BF_TILE - 1) / BF_TILE * BF_TILE);
    chunk = std::min(chunk, 1 << 21);
    const int nChunks = std::max(1, (ndb + chunk - 1) / chunk);
    const size_t pBytes = sizeof(int32_t) * 3 * (size_t)nChunks * nq;
    int rc = ensure_partials(m, pBytes);
    if (rc != ORBX_OK) return rc;
    int32_t* pBest = m->partials;
    int32_t* pIdx = pBest + (size_t)nChunks * nq;
    int32_t* pSecond = pIdx + (size_t)nChunks * nq;
    static const int variant = [] { const char* e = std::getenv("ORBX_BF_VARIANT"); return e ? std::atoi(e) : BF_DEFAULT_VARIANT; }();
    const dim3 grid(nChunks, qBlocks);
    if (variant == 0) k_bruteforce<0><<<grid, BF_T, 0, m->stream>>>(dq, nq, ddb, ndb, chunk, idxOffset, pBest, pIdx, pSecond);
    else if (variant == 1) k_bruteforce<1><<<grid, BF_T, 0, m->stream>>>(dq, nq, ddb, ndb, chunk, idxOffset, pBest, pIdx, pSecond);
    else if (variant == 3) k_bruteforce<3><<<grid, BF_T, 0, m->stream>>>(dq, nq, ddb, ndb, chunk, idxOffset, pBest, pIdx, pSecond);
    else if (variant == 4) k_bruteforce<4><<<grid, BF_T, 0, m->stream>>>(dq, nq, ddb, ndb, chunk, idxOffset, pBest, pIdx, pSecond);
    else k_bruteforce<2><<<grid, BF_T, 0, m->stream>>>(dq, nq, ddb, ndb, chunk, idxOffset, pBest, pIdx, pSecond);
    ORBX_CUDA_TRY(cudaGetLastError());
    *nChunksOut = nChunks;
    return ORBX_OK;
}

static int bruteforce_device(orbm_context* m, const uint8_t* dq, int nq, const uint8_t* ddb, int ndb, int idxOffset,
                             int32_t* dBest, int32_t* dIdx, int32_t* dSecond)
{
    int nChunks = 0;
    int rc = bruteforce_partials(m, dq, nq, ddb, ndb, idxOffset, &nChunks);
    if (rc != ORBX_OK) return rc;
    const int32_t* pBest = m->partials;
    const int32_t* pIdx = pBest + (size_t)nChunks * nq;
    const int32_t* pSecond = pIdx + (size_t)nChunks * nq;
    k_merge_partials<<<(nq + 7) / 8, 256, 0, m->stream>>>(pBest, pIdx, pSecond, nChunks, nq, dBest, dIdx, dSecond);
    ORBX_CUDA_TRY(cudaGetLastError());
    return ORBX_OK;
}

int orbm_bruteforce(orbm_context* m, const uint8_t* q, int nq, const uint8_t* db, int ndb, int idx_offset, int32_t* best,
                    int32_t* best_idx, int32_t* second, int mem_kind)
{
    if (!m || !q || nq <= 0 || ndb < 0 || (ndb > 0 && !db) || !best || !best_idx || !second) {
        set_error("orbm_bruteforce: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    if (mem_kind == 1) return bruteforce_device(m, q, nq, db, ndb, idx_offset, best, best_idx, second);
    Carver c;
    const size_t oq = c.take(32 * (size_t)nq), odb = c.take(32 * (size_t)std::max(ndb, 1)), ob = c.take(4 * (size_t)nq),
                 oi = c.take(4 * (size_t)nq), os = c.take(4 * (size_t)nq);
    int rc = ensure_arena(m, c.off);
    if (rc != ORBX_OK) return rc;
    char* A = m->arena;
    ORBX_CUDA_TRY(cudaMemcpyAsync(A + oq, q, 32 * (size_t)nq, cudaMemcpyHostToDevice, m->stream));
    if (ndb) ORBX_CUDA_TRY(cudaMemcpyAsync(A + odb, db, 32 * (size_t)ndb, cudaMemcpyHostToDevice, m->stream));
    rc = bruteforce_device(m, (uint8_t*)(A + oq), nq, (uint8_t*)(A + odb), ndb, idx_offset, (int32_t*)(A + ob),
                           (int32_t*)(A + oi), (int32_t*)(A + os));
    if (rc != ORBX_OK) return rc;
    ORBX_CUDA_TRY(cudaMemcpyAsync(best, A + ob, 4 * (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaMemcpyAsync(best_idx, A + oi, 4 * (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaMemcpyAsync(second, A + os, 4 * (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

int orbm_merge_partials(orbm_context* m, const int32_t* best, const int32_t* best_idx, const int32_t* second, int nparts,
                        int nq, int32_t* out_best, int32_t* out_idx, int32_t* out_second)
{
    if (!m || !best || !best_idx || !second || nparts <= 0 || nq <= 0 || !out_best || !out_idx || !out_second) {
        set_error("orbm_merge_partials: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    k_merge_partials<<<(nq + 7) / 8, 256, 0, m->stream>>>(best, best_idx, second, nparts, nq, out_best, out_idx, out_second);
    ORBX_CUDA_TRY(cudaGetLastError());
    return ORBX_OK;
}

// orbm_bruteforce for callers that must not allocate: same two kernels (k_bruteforce<1>, k_merge_partials), the chunk
// partials in the module's static scratch array instead of orbm_context::partials.  No context, no cudaMalloc / cudaFree.
int orbm_bruteforce_noalloc(int device, void* stream, const uint8_t* q, int nq, const uint8_t* db, int ndb, int idx_offset,
                            int32_t* best, int32_t* best_idx, int32_t* second)
{
    if (device < 0 || device >= NOALLOC_MAX_DEVICES || !q || nq <= 0 || ndb < 0 || (ndb > 0 && !db) || !best || !best_idx || !second) {
        set_error("orbm_bruteforce_noalloc: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    if (nq > ORBM_NOALLOC_MAX_NQ || ndb > ORBM_NOALLOC_MAX_NDB) {
        set_error("orbm_bruteforce_noalloc: nq = %d / ndb = %d exceed the fixed scratch (at most %d queries x %d database rows); "
                  "use orbm_bruteforce, which grows its scratch with cudaMalloc", nq, ndb, ORBM_NOALLOC_MAX_NQ, ORBM_NOALLOC_MAX_NDB);
        return ORBX_ERR_CAPACITY;
    }
    ORBX_CUDA_TRY(cudaSetDevice(device));
    int32_t* scratch = nullptr;
    int numSMs = 0;
    int rc = noalloc_setup(device, &scratch, &numSMs);
    if (rc != ORBX_OK) return rc;
    // chunking of bruteforce_partials, with the chunk count also bounded by the scratch slots
    const int qBlocks = (nq + BF_T * BF_RQ - 1) / (BF_T * BF_RQ);
    const int target = std::max(1, std::min((numSMs * 4) / qBlocks, ORBM_NOALLOC_SCRATCH_SLOTS / nq));
    int chunk = (ndb + target - 1) / target;
    chunk = std::max(BF_TILE, (chunk + BF_TILE - 1) / BF_TILE * BF_TILE);
    const int nChunks = std::max(1, (ndb + chunk - 1) / chunk);
    if (chunk > (1 << 21) || (size_t)nChunks * nq > (size_t)ORBM_NOALLOC_SCRATCH_SLOTS) {
        set_error("orbm_bruteforce_noalloc: %d chunks of %d rows x %d queries do not fit the fixed scratch", nChunks, chunk, nq);
        return ORBX_ERR_CAPACITY;
    }
    // every slot k_merge_partials reads ([chunk < nChunks][query < nq] of the three planes) is written by k_bruteforce of
    // this same call, so what the previous call left in the array never needs clearing
    int32_t* pBest = scratch;
    int32_t* pIdx = pBest + (size_t)nChunks * nq;
    int32_t* pSecond = pIdx + (size_t)nChunks * nq;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    k_bruteforce<BF_DEFAULT_VARIANT><<<dim3(nChunks, qBlocks), BF_T, 0, st>>>(q, nq, db, ndb, chunk, idx_offset, pBest, pIdx, pSecond);
    ORBX_CUDA_TRY(cudaGetLastError());
    k_merge_partials<<<(nq + 7) / 8, 256, 0, st>>>(pBest, pIdx, pSecond, nChunks, nq, best, best_idx, second);
    ORBX_CUDA_TRY(cudaGetLastError());
    return ORBX_OK;
}

// ---- tensor-core brute force on a prepared database (see orbx_match_tc.cuh) ----
int orbm_db_create(orbm_context* m, const uint8_t* db, int ndb, int mem_kind, orbm_db** out)
{
    if (!m || !db || ndb <= 0 || !out || (mem_kind != 0 && mem_kind != 1)) {
        set_error("orbm_db_create: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    cudaDeviceProp prop;
    ORBX_CUDA_TRY(cudaGetDeviceProperties(&prop, m->device));
    if (prop.major != 10) {
        set_error("orbm_db_create: the tensor-core search needs an sm_100 device (tcgen05), found sm_%d%d", prop.major, prop.minor);
        return ORBX_ERR_NO_DEVICE;
    }
    orbm_db* d = new orbm_db;
    d->device = m->device;
    d->ndb = ndb;
    d->rowsPad = ((long long)ndb + tc::BN - 1) / tc::BN * tc::BN;
    cudaError_t e = cudaMalloc(&d->exp, (size_t)d->rowsPad * 256);
    if (e != cudaSuccess) { delete d; set_error("orbm_db_create: %s", cudaGetErrorString(e)); return ORBX_ERR_CUDA; }
    const uint8_t* src = db;
    uint8_t* tmp = nullptr;
    if (mem_kind == 0) {
        e = cudaMalloc(&tmp, 32 * (size_t)ndb);
        if (e == cudaSuccess) e = cudaMemcpyAsync(tmp, db, 32 * (size_t)ndb, cudaMemcpyHostToDevice, m->stream);
        if (e != cudaSuccess) { cudaFree(tmp); cudaFree(d->exp); delete d; set_error("orbm_db_create: %s", cudaGetErrorString(e)); return ORBX_ERR_CUDA; }
        src = tmp;
    }
    tc::k_expand_pm1<<<(unsigned)((d->rowsPad * 32 + 255) / 256), 256, 0, m->stream>>>(src, ndb, d->rowsPad, d->exp);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(m->stream);
    cudaFree(tmp);
    int rc = e == cudaSuccess ? tc::make_map(&d->map, d->exp, d->rowsPad, tc::BN) : ORBX_ERR_CUDA;
    if (rc == ORBX_OK) rc = tc::make_map(&d->map128, d->exp, d->rowsPad, 128);
    if (e != cudaSuccess) set_error("orbm_db_create: %s", cudaGetErrorString(e));
    if (rc != ORBX_OK) { cudaFree(d->exp); delete d; return rc; }
    *out = d;
    return ORBX_OK;
}

void orbm_db_destroy(orbm_db* d)
{
    if (!d) return;
    cudaSetDevice(d->device);
    cudaFree(d->exp);
    delete d;
}

int orbm_bruteforce_db(orbm_context* m, const orbm_db* d, const uint8_t* q, int nq, int idx_offset, int32_t* best, int32_t* best_idx,
                       int32_t* second, int mem_kind)
{
    if (!m || !d || !q || nq <= 0 || !best || !best_idx || !second || d->device != m->device) {
        set_error("orbm_bruteforce_db: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    const uint8_t* dq = q;
    int32_t *dB = best, *dI = best_idx, *dS = second;
    char* A = nullptr;
    size_t ob = 0, oi = 0, os = 0;
    if (mem_kind != 1) {
        Carver c;
        const size_t oq = c.take(32 * (size_t)nq);
        ob = c.take(4 * (size_t)nq); oi = c.take(4 * (size_t)nq); os = c.take(4 * (size_t)nq);
        int rc = ensure_arena(m, c.off);
        if (rc != ORBX_OK) return rc;
        A = m->arena;
        ORBX_CUDA_TRY(cudaMemcpyAsync(A + oq, q, 32 * (size_t)nq, cudaMemcpyHostToDevice, m->stream));
        dq = (uint8_t*)(A + oq); dB = (int32_t*)(A + ob); dI = (int32_t*)(A + oi); dS = (int32_t*)(A + os);
    }
    int nChunks = 0;
    int rc = bruteforce_partials_tc(m, d, dq, nq, idx_offset, &nChunks);
    if (rc != ORBX_OK) return rc;
    const int32_t* pBest = m->partials;
    const int32_t* pIdx = pBest + (size_t)nChunks * nq;
    const int32_t* pSecond = pIdx + (size_t)nChunks * nq;
    k_merge_partials<<<(nq + 7) / 8, 256, 0, m->stream>>>(pBest, pIdx, pSecond, nChunks, nq, dB, dI, dS);
    ORBX_CUDA_TRY(cudaGetLastError());
    if (mem_kind != 1) {
        ORBX_CUDA_TRY(cudaMemcpyAsync(best, A + ob, 4 * (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
        ORBX_CUDA_TRY(cudaMemcpyAsync(best_idx, A + oi, 4 * (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
        ORBX_CUDA_TRY(cudaMemcpyAsync(second, A + os, 4 * (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
        ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    }
    return ORBX_OK;
}

// ---- frame preparation between extraction and matching ("next" row) -------------------------------
int orbm_undistort_keypoints(orbm_context* m, const float* x, const float* y, int n, float fx, float fy, float cx, float cy,
                             const float* dist, int ndist, float* xu, float* yu, int mem_kind)
{
    if (!m || n < 0 || (n > 0 && (!x || !y || !xu || !yu)) || ndist < 0 || ndist > 12 || (ndist > 0 && !dist) || fx == 0.f || fy == 0.f) {
        set_error("orbm_undistort_keypoints: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    if (n == 0) return ORBX_OK;
    UndistortCam c;
    c.fx = fx; c.fy = fy; c.cx = cx; c.cy = cy;
    for (int i = 0; i < 12; i++) c.k[i] = i < ndist ? (double)dist[i] : 0.0;   // dist is a HOST array in both modes
    const bool identity = ndist == 0 || dist[0] == 0.0f;                        // src/Frame.cc:589-592: mvKeysUn = mvKeys
    const size_t bytes = sizeof(float) * (size_t)n;
    if (mem_kind == 1) {
        if (identity) {
            if (xu != x) ORBX_CUDA_TRY(cudaMemcpyAsync(xu, x, bytes, cudaMemcpyDeviceToDevice, m->stream));
            if (yu != y) ORBX_CUDA_TRY(cudaMemcpyAsync(yu, y, bytes, cudaMemcpyDeviceToDevice, m->stream));
        } else {
            k_undistort<<<(n + 127) / 128, 128, 0, m->stream>>>(x, y, n, c, xu, yu);
            ORBX_CUDA_TRY(cudaGetLastError());
        }
        return ORBX_OK;
    }
    if (identity) {
        if (xu != x) memcpy(xu, x, bytes);
        if (yu != y) memcpy(yu, y, bytes);
        return ORBX_OK;
    }
    Carver cv;
    const size_t oxi = cv.take(bytes), oyi = cv.take(bytes), oxo = cv.take(bytes), oyo = cv.take(bytes);
    int rc = ensure_arena(m, cv.off);
    if (rc != ORBX_OK) return rc;
    char* A = m->arena;
    ORBX_CUDA_TRY(cudaMemcpyAsync(A + oxi, x, bytes, cudaMemcpyHostToDevice, m->stream));
    ORBX_CUDA_TRY(cudaMemcpyAsync(A + oyi, y, bytes, cudaMemcpyHostToDevice, m->stream));
    k_undistort<<<(n + 127) / 128, 128, 0, m->stream>>>((float*)(A + oxi), (float*)(A + oyi), n, c, (float*)(A + oxo), (float*)(A + oyo));
    ORBX_CUDA_TRY(cudaGetLastError());
    ORBX_CUDA_TRY(cudaMemcpyAsync(xu, A + oxo, bytes, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaMemcpyAsync(yu, A + oyo, bytes, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

int orbm_assign_grid(orbm_context* m, const float* x, const float* y, int n, float min_x, float min_y, float grid_w_inv,
                     float grid_h_inv, int32_t* cell_start, int32_t* cell_items, int mem_kind)
{
    if (!m || n < 0 || (n > 0 && (!x || !y || !cell_items)) || !cell_start) {
        set_error("orbm_assign_grid: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    if (mem_kind == 1) {
        k_assign_grid<<<1, 1024, 0, m->stream>>>(x, y, n, min_x, min_y, grid_w_inv, grid_h_inv, cell_start, cell_items);
        ORBX_CUDA_TRY(cudaGetLastError());
        return ORBX_OK;
    }
    Carver cv;
    const size_t bytes = sizeof(float) * (size_t)std::max(n, 1);
    const size_t ox = cv.take(bytes), oy = cv.take(bytes), os = cv.take(4 * (GRID_CELLS + 1)), oi = cv.take(bytes);
    int rc = ensure_arena(m, cv.off);
    if (rc != ORBX_OK) return rc;
    char* A = m->arena;
    if (n) {
        ORBX_CUDA_TRY(cudaMemcpyAsync(A + ox, x, sizeof(float) * (size_t)n, cudaMemcpyHostToDevice, m->stream));
        ORBX_CUDA_TRY(cudaMemcpyAsync(A + oy, y, sizeof(float) * (size_t)n, cudaMemcpyHostToDevice, m->stream));
    }
    k_assign_grid<<<1, 1024, 0, m->stream>>>((float*)(A + ox), (float*)(A + oy), n, min_x, min_y, grid_w_inv, grid_h_inv,
                                             (int32_t*)(A + os), (int32_t*)(A + oi));
    ORBX_CUDA_TRY(cudaGetLastError());
    ORBX_CUDA_TRY(cudaMemcpyAsync(cell_start, A + os, 4 * (GRID_CELLS + 1), cudaMemcpyDeviceToHost, m->stream));
    if (n) ORBX_CUDA_TRY(cudaMemcpyAsync(cell_items, A + oi, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

// ---- sharded search over peer memory -------------------------------------------------------------------------------
static void exchange_release(orbm_context* m)
{
    for (int r = 0; r < 16; r++) {
        if (m->xchgOpenedIpc[r] && m->xchgPeers[r]) cudaIpcCloseMemHandle(m->xchgPeers[r]);
        m->xchgPeers[r] = nullptr;
        m->xchgOpenedIpc[r] = false;
    }
    if (m->xchg) cudaFree(m->xchg);
    m->xchg = nullptr;
    m->xchgBytes = 0;
    if (m->xchgTimeout) cudaFreeHost(m->xchgTimeout);
    m->xchgTimeout = nullptr;
    m->xchgConnected = false;
    m->world = 0;
}

int orbm_exchange_create(orbm_context* m, int world, int rank, int nq_max, uint8_t* handle_out, void** local_ptr_out)
{
    if (!m || world < 1 || world > BFX_MAX_WORLD || rank < 0 || rank >= world || nq_max <= 0) {
        set_error("orbm_exchange_create: bad argument (world 1..%d)", BFX_MAX_WORLD);
        return ORBX_ERR_BAD_ARG;
    }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    exchange_release(m);
    m->xchgBytes = sizeof(unsigned long long) * 2 * (size_t)world * nq_max;
    ORBX_CUDA_TRY(cudaMalloc(&m->xchg, m->xchgBytes));
    ORBX_CUDA_TRY(cudaMemset(m->xchg, 0, m->xchgBytes));  // epoch 0 = "never written"; calls start at epoch 1
    ORBX_CUDA_TRY(cudaHostAlloc(&m->xchgTimeout, sizeof(int), cudaHostAllocMapped));
    *m->xchgTimeout = 0;
    m->world = world; m->rank = rank; m->nqMax = nq_max; m->epoch = 0;
    m->xchgPeers[rank] = m->xchg;
    if (handle_out) {
        cudaIpcMemHandle_t hnd;
        static_assert(sizeof(hnd) == ORBM_IPC_HANDLE_BYTES, "handle size");
        ORBX_CUDA_TRY(cudaIpcGetMemHandle(&hnd, m->xchg));
        memcpy(handle_out, &hnd, sizeof hnd);
    }
    if (local_ptr_out) *local_ptr_out = m->xchg;
    m->xchgConnected = world == 1;
    return ORBX_OK;
}

int orbm_exchange_connect(orbm_context* m, const uint8_t* handles)
{
    if (!m || !m->xchg || !handles) { set_error("orbm_exchange_connect: create the exchange first"); return ORBX_ERR_BAD_ARG; }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    for (int r = 0; r < m->world; r++) {
        if (r == m->rank) continue;
        cudaIpcMemHandle_t hnd;
        memcpy(&hnd, handles + (size_t)r * ORBM_IPC_HANDLE_BYTES, sizeof hnd);
        void* p = nullptr;
        ORBX_CUDA_TRY(cudaIpcOpenMemHandle(&p, hnd, cudaIpcMemLazyEnablePeerAccess));
        m->xchgPeers[r] = (unsigned long long*)p;
        m->xchgOpenedIpc[r] = true;
    }
    m->xchgConnected = true;
    return ORBX_OK;
}

int orbm_exchange_connect_local(orbm_context* m, void* const* peer_ptrs)
{
    if (!m || !m->xchg || !peer_ptrs) { set_error("orbm_exchange_connect_local: create the exchange first"); return ORBX_ERR_BAD_ARG; }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    for (int r = 0; r < m->world; r++) {
        if (r == m->rank) continue;
        if (!peer_ptrs[r]) { set_error("orbm_exchange_connect_local: null pointer for rank %d", r); return ORBX_ERR_BAD_ARG; }
        cudaPointerAttributes a;
        ORBX_CUDA_TRY(cudaPointerGetAttributes(&a, peer_ptrs[r]));
        if (a.type == cudaMemoryTypeDevice && a.device != m->device) {
            cudaError_t e = cudaDeviceEnablePeerAccess(a.device, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else if (e != cudaSuccess) { set_error("peer access %d -> %d: %s", m->device, a.device, cudaGetErrorString(e)); return ORBX_ERR_CUDA; }
        }
        m->xchgPeers[r] = (unsigned long long*)peer_ptrs[r];
    }
    m->xchgConnected = true;
    return ORBX_OK;
}

void orbm_exchange_destroy(orbm_context* m)
{
    if (!m) return;
    cudaSetDevice(m->device);
    if (m->stream) cudaStreamSynchronize(m->stream);
    exchange_release(m);
}

int orbm_bruteforce_sharded(orbm_context* m, const uint8_t* q, int nq, const uint8_t* db, int ndb, int idx_offset,
                            int32_t* best, int32_t* best_idx, int32_t* second)
{
    return orbm_bruteforce_sharded_phases(m, q, nq, db, ndb, idx_offset, best, best_idx, second, 3);
}

static int sharded_impl(orbm_context* m, const uint8_t* q, int nq, const uint8_t* db, int ndb, const orbm_db* prepared, int idx_offset,
                        int32_t* best, int32_t* best_idx, int32_t* second, int phases);

int orbm_bruteforce_sharded_phases(orbm_context* m, const uint8_t* q, int nq, const uint8_t* db, int ndb, int idx_offset,
                                   int32_t* best, int32_t* best_idx, int32_t* second, int phases)
{
    return sharded_impl(m, q, nq, db, ndb, nullptr, idx_offset, best, best_idx, second, phases);
}

int orbm_bruteforce_sharded_db(orbm_context* m, const orbm_db* db_shard, const uint8_t* q, int nq, int idx_offset, int32_t* best,
                               int32_t* best_idx, int32_t* second)
{
    if (!db_shard || !m || db_shard->device != m->device) { set_error("orbm_bruteforce_sharded_db: bad argument"); return ORBX_ERR_BAD_ARG; }
    return sharded_impl(m, q, nq, nullptr, db_shard->ndb, db_shard, idx_offset, best, best_idx, second, 3);
}

static int sharded_impl(orbm_context* m, const uint8_t* q, int nq, const uint8_t* db, int ndb, const orbm_db* prepared, int idx_offset,
                        int32_t* best, int32_t* best_idx, int32_t* second, int phases)
{
    const bool pub = phases & 1, col = phases & 2;
    if (!m || nq <= 0 || !(pub || col) || (pub && (!q || ndb < 0 || (ndb > 0 && !db && !prepared))) || (col && (!best || !best_idx || !second))) {
        set_error("orbm_bruteforce_sharded: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    if (!m->xchg || !m->xchgConnected) { set_error("orbm_bruteforce_sharded: exchange not created / connected"); return ORBX_ERR_BAD_ARG; }
    if (nq > m->nqMax) { set_error("orbm_bruteforce_sharded: nq %d > nq_max %d of the exchange", nq, m->nqMax); return ORBX_ERR_CAPACITY; }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    if (*m->xchgTimeout) { set_error("orbm_bruteforce_sharded: an earlier call timed out waiting for a peer"); return ORBX_ERR_CUDA; }
    int nChunks = 0;
    if (pub) {
        int rc = prepared ? bruteforce_partials_tc(m, prepared, q, nq, idx_offset, &nChunks)
                          : bruteforce_partials(m, q, nq, db, ndb, idx_offset, &nChunks);
        if (rc != ORBX_OK) return rc;
        m->epoch++;
    } else if (m->epoch == 0) {
        set_error("orbm_bruteforce_sharded: collect without a preceding publish");
        return ORBX_ERR_BAD_ARG;
    }
    const int32_t* pBest = m->partials;
    const int32_t* pIdx = pBest + (size_t)nChunks * nq;
    const int32_t* pSecond = pIdx + (size_t)nChunks * nq;
    BfxPeers peers;
    for (int r = 0; r < BFX_MAX_WORLD; r++) peers.p[r] = r < m->world ? m->xchgPeers[r] : nullptr;
    int* dTimeout = nullptr;
    ORBX_CUDA_TRY(cudaHostGetDevicePointer(&dTimeout, m->xchgTimeout, 0));
    static const long long budget = [] { const char* e = std::getenv("ORBX_EXCHANGE_SPIN_CYCLES"); return e ? std::atoll(e) : 20000000000LL; }();
    const int grid = std::max(1, std::min(m->numSMs, (nq + 7) / 8));
    // tag scrub window of this call: W slots starting at (calls of this parity so far) * W, so that a set's nq_max slots
    // are covered once per <= 2048 calls of its parity (the tag wraps after 8192 of them)
    const int scrubW = std::min(m->nqMax, (m->nqMax + 2047) / 2048);
    const int scrubFirst = (int)(((unsigned long long)(m->epoch >> 1) * (unsigned)scrubW) % (unsigned)m->nqMax);
    k_merge_exchange<<<grid, 256, 0, m->stream>>>(pBest, pIdx, pSecond, nChunks, nq, m->nqMax, peers, m->rank, m->world, m->epoch, budget,
                                                  dTimeout, best, best_idx, second, phases, scrubFirst, pub ? scrubW : 0);
    ORBX_CUDA_TRY(cudaGetLastError());
    return ORBX_OK;
}

int orbm_exchange_timed_out(orbm_context* m) { return m && m->xchgTimeout ? *m->xchgTimeout : 0; }

int orbm_debug_exchange_snapshot(orbm_context* m, uint64_t* words, size_t nwords, uint32_t* epoch)
{
    if (!m || !m->xchg || !words || nwords * 8 > m->xchgBytes) { set_error("orbm_debug_exchange_snapshot: bad argument"); return ORBX_ERR_BAD_ARG; }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    ORBX_CUDA_TRY(cudaMemcpy(words, m->xchg, nwords * 8, cudaMemcpyDeviceToHost));
    if (epoch) *epoch = m->epoch;
    return ORBX_OK;
}

int orbm_search_window(orbm_context* m, const orbm_frame_view* f, const float* f_angle, const orbm_window_query_view* q,
                       int max_dist, int ratio_mode, float nnratio, int check_orientation, int32_t* out_match, int* nmatches)
{
    if (!m || !f || !q || !out_match || !nmatches || f->n < 0 || q->n < 0 || (check_orientation && (!f_angle || !q->angle)) ||
        f->n >= (1 << 30)) {
        set_error("orbm_search_window: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    const int nkp = f->n, nmp = q->n;
    if (nkp == 0 || nmp == 0) {
        for (int i = 0; i < nkp; i++) out_match[i] = -1;
        *nmatches = 0;
        return ORBX_OK;
    }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    const int ncell = GRID_COLS * GRID_ROWS;
    const int nitems = f->cell_start[ncell];
    for (int ci = 0; ci < ncell; ci++)
        if (f->cell_start[ci + 1] - f->cell_start[ci] >= 2048) {
            set_error("orbm_search_window: a grid cell holds %d keypoints (limit 2047)", f->cell_start[ci + 1] - f->cell_start[ci]);
            return ORBX_ERR_BAD_ARG;
        }
    Carver c;
    const size_t oKx = c.take(4 * (size_t)nkp), oKy = c.take(4 * (size_t)nkp), oKo = c.take(4 * (size_t)nkp),
                 oUr = c.take(4 * (size_t)nkp), oKa = c.take(4 * (size_t)nkp), oKd = c.take(32 * (size_t)nkp), oOcc = c.take(nkp),
                 oCs = c.take(4 * (size_t)(ncell + 1)), oCi = c.take(4 * (size_t)std::max(nitems, 1));
    const size_t oPx = c.take(4 * (size_t)nmp), oPy = c.take(4 * (size_t)nmp), oPxr = c.take(4 * (size_t)nmp),
                 oPr = c.take(4 * (size_t)nmp), oPmin = c.take(4 * (size_t)nmp), oPmax = c.take(4 * (size_t)nmp),
                 oPa = c.take(4 * (size_t)nmp), oPva = c.take(nmp), oPo = c.take(nmp), oPd = c.take(32 * (size_t)nmp);
    const size_t oRes = c.take(4 * (size_t)nmp), oCa = c.take(4 * (size_t)nkp), oCb = c.take(4 * (size_t)nkp),
                 oCc = c.take(4 * (size_t)nkp), oLk = c.take(4 * (size_t)nmp * PROJ_LIST), oLi = c.take(4 * (size_t)nmp * PROJ_LIST),
                 oLc = c.take(4 * (size_t)nmp), oOut = c.take(4 * (size_t)nkp), oMisc = c.take(256);
    int rc = ensure_arena(m, c.off);
    if (rc != ORBX_OK) return rc;
    char* A = m->arena;
#define UP(off, ptr, count, T) if ((rc = up<T>(m, (T*)(A + off), ptr, count)) != ORBX_OK) return rc
    UP(oKx, f->x, nkp, float); UP(oKy, f->y, nkp, float); UP(oKo, f->octave, nkp, int32_t);
    if (f->u_right) UP(oUr, f->u_right, nkp, float);
    if (f_angle) UP(oKa, f_angle, nkp, float);
    UP(oKd, f->desc, 32 * (size_t)nkp, uint8_t); UP(oOcc, f->occupied, nkp, uint8_t);
    UP(oCs, f->cell_start, ncell + 1, int32_t); UP(oCi, f->cell_items, nitems, int32_t);
    UP(oPx, q->x, nmp, float); UP(oPy, q->y, nmp, float);
    if (q->xr) UP(oPxr, q->xr, nmp, float);
    UP(oPr, q->radius, nmp, float); UP(oPmin, q->min_level, nmp, int32_t); UP(oPmax, q->max_level, nmp, int32_t);
    if (q->angle) UP(oPa, q->angle, nmp, float);
    UP(oPva, q->valid, nmp, uint8_t); UP(oPo, q->claims_block, nmp, uint8_t); UP(oPd, q->desc, 32 * (size_t)nmp, uint8_t);
#undef UP
    if ((rc = flush_up(m)) != ORBX_OK) return rc;
    ProjArgs a;
    a.nkp = nkp;
    a.kx = (float*)(A + oKx); a.ky = (float*)(A + oKy); a.koct = (int32_t*)(A + oKo);
    a.uright = (f->u_right && q->xr) ? (float*)(A + oUr) : nullptr;
    a.kdesc = (uint8_t*)(A + oKd); a.occupied = (uint8_t*)(A + oOcc);
    a.cellStart = (int32_t*)(A + oCs); a.cellItems = (int32_t*)(A + oCi);
    a.minX = f->min_x; a.minY = f->min_y; a.wInv = f->grid_w_inv; a.hInv = f->grid_h_inv;
    a.kangle = (float*)(A + oKa);
    a.nmp = nmp;
    a.px = (float*)(A + oPx); a.py = (float*)(A + oPy); a.pxr = (float*)(A + oPxr); a.prr = (float*)(A + oPr);
    a.pmin = (int32_t*)(A + oPmin); a.pmax = (int32_t*)(A + oPmax); a.pangle = (float*)(A + oPa);
    a.pvalid = (uint8_t*)(A + oPva); a.pobs = (uint8_t*)(A + oPo); a.pdesc = (uint8_t*)(A + oPd);
    a.nnratio = nnratio; a.maxDist = max_dist; a.ratioMode = ratio_mode; a.checkOri = check_orientation ? 1 : 0;
    a.res = (int32_t*)(A + oRes); a.claimA = (int32_t*)(A + oCa); a.claimB = (int32_t*)(A + oCb); a.claimC = (int32_t*)(A + oCc);
    a.listKey = (uint32_t*)(A + oLk); a.listIdx = (int32_t*)(A + oLi); a.listCnt = (int32_t*)(A + oLc);
    a.outMatch = (int32_t*)(A + oOut);
    a.changed = (int*)(A + oMisc); a.nmatches = a.changed + 3; a.sweeps = a.changed + 4; a.hist = a.changed + 8;  // changed[3]
    int perSM = 0;
    ORBX_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, k_proj_search, PROJ_T, 0));
    const int want = (nmp + PROJ_T / 32 - 1) / (PROJ_T / 32);  // one warp per query
    const int grid = std::max(1, std::min(want, perSM * m->numSMs));
    void* args[] = {&a};
    ORBX_CUDA_TRY(cudaLaunchCooperativeKernel((void*)k_proj_search, dim3(grid), dim3(PROJ_T), args, 0, m->stream));
    ORBX_CUDA_TRY(cudaMemcpyAsync(out_match, a.outMatch, 4 * (size_t)nkp, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaMemcpyAsync(nmatches, a.nmatches, sizeof(int), cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    static const bool projDebug = std::getenv("ORBX_PROJ_DEBUG") != nullptr;
    if (projDebug) {
        int sw = 0;
        cudaMemcpy(&sw, a.sweeps, sizeof(int), cudaMemcpyDeviceToHost);
        std::fprintf(stderr, "k_proj_search: %d queries, %d keypoints, grid %d x %d, %d sweeps\n", nmp, nkp, grid, PROJ_T, sw);
    }
    return ORBX_OK;
}

int orbm_search_by_projection(orbm_context* m, const orbm_frame_view* f, const orbm_mappoint_view* p, float th,
                              float nnratio, int32_t* out_match, int* nmatches)
{
    if (!m || !f || !p || !out_match || !nmatches || f->n < 0 || p->n < 0 || !f->scale_factors || f->nlevels <= 0) {
        set_error("orbm_search_by_projection: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    // window radius and level range per map point (src/ORBmatcher.cc:84-96, RadiusByViewingCos :176-183)
    const int n = p->n;
    std::vector<float> radius(n > 0 ? n : 1);
    std::vector<int32_t> lmin(n > 0 ? n : 1), lmax(n > 0 ? n : 1);
    std::vector<uint8_t> valid(n > 0 ? n : 1);
    const bool bFactor = th != 1.0f;
    for (int i = 0; i < n; i++) {
        const int level = p->level[i];
        valid[i] = p->valid[i] && level >= 0 && level < f->nlevels;
        float r = p->view_cos[i] > 0.998 ? 2.5f : 4.0f;
        if (bFactor) r *= th;
        radius[i] = valid[i] ? r * f->scale_factors[level] : 0.f;
        lmin[i] = level - 1;
        lmax[i] = level;
    }
    orbm_window_query_view q;
    q.n = n; q.x = p->proj_x; q.y = p->proj_y; q.radius = radius.data(); q.min_level = lmin.data(); q.max_level = lmax.data();
    q.xr = p->proj_xr; q.angle = nullptr; q.valid = valid.data(); q.claims_block = p->obs_pos; q.desc = p->desc;
    return orbm_search_window(m, f, nullptr, &q, TH_HIGH, 1, nnratio, 0, out_match, nmatches);
}

int orbm_stereo_rows(orbm_context* m, int nl, const float* lx, const float* ly, const int32_t* loct, const uint8_t* ldesc,
                     int nr, const float* rx, const float* ry, const int32_t* roct, const uint8_t* rdesc,
                     const float* scale_factors, int nlevels, int n_rows, float min_d, float max_d, int32_t* best_r,
                     int32_t* best_dist)
{
    if (!m || nl < 0 || nr < 0 || !best_r || !best_dist || !scale_factors || nlevels <= 0 || nr >= (1 << 22)) {
        set_error("orbm_stereo_rows: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    if (nl == 0) return ORBX_OK;
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    Carver c;
    const size_t oLx = c.take(4 * (size_t)nl), oLy = c.take(4 * (size_t)nl), oLo = c.take(4 * (size_t)nl),
                 oLd = c.take(32 * (size_t)nl);
    const int nrA = std::max(nr, 1);
    const size_t oRx = c.take(4 * (size_t)nrA), oRy = c.take(4 * (size_t)nrA), oRo = c.take(4 * (size_t)nrA),
                 oRd = c.take(32 * (size_t)nrA), oSf = c.take(4 * (size_t)nlevels);
    const size_t oBr = c.take(4 * (size_t)nl), oBd = c.take(4 * (size_t)nl);
    int rc = ensure_arena(m, c.off);
    if (rc != ORBX_OK) return rc;
    char* A = m->arena;
#define UP(off, ptr, count, T) if ((rc = up<T>(m, (T*)(A + off), ptr, count)) != ORBX_OK) return rc
    UP(oLx, lx, nl, float); UP(oLy, ly, nl, float); UP(oLo, loct, nl, int32_t); UP(oLd, ldesc, 32 * (size_t)nl, uint8_t);
    UP(oRx, rx, nr, float); UP(oRy, ry, nr, float); UP(oRo, roct, nr, int32_t); UP(oRd, rdesc, 32 * (size_t)nr, uint8_t);
    UP(oSf, scale_factors, nlevels, float);
#undef UP
    if ((rc = flush_up(m)) != ORBX_OK) return rc;
    k_stereo_rows<<<(nl + 7) / 8, 256, 0, m->stream>>>(nl, (float*)(A + oLx), (float*)(A + oLy), (int32_t*)(A + oLo),
                                                      (uint8_t*)(A + oLd), nr, (float*)(A + oRx), (float*)(A + oRy),
                                                      (int32_t*)(A + oRo), (uint8_t*)(A + oRd), (float*)(A + oSf), n_rows,
                                                      min_d, max_d, (int32_t*)(A + oBr), (int32_t*)(A + oBd));
    ORBX_CUDA_TRY(cudaGetLastError());
    ORBX_CUDA_TRY(cudaMemcpyAsync(best_r, A + oBr, 4 * (size_t)nl, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaMemcpyAsync(best_dist, A + oBd, 4 * (size_t)nl, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

int orbm_search_by_bow(orbm_context* m, int nkf, const uint8_t* kf_desc, const float* kf_angle, const uint8_t* kf_has_mp,
                       const orbm_featvec_view* kf_fv, int nf, const uint8_t* f_desc, const float* f_angle,
                       const orbm_featvec_view* f_fv, float nnratio, int check_orientation, int32_t* out_match,
                       int* nmatches)
{
    if (!m || !kf_fv || !f_fv || !out_match || !nmatches || nkf < 0 || nf < 0) {
        set_error("orbm_search_by_bow: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    for (int i = 0; i < nf; i++) out_match[i] = -1;
    *nmatches = 0;
    if (nkf == 0 || nf == 0 || kf_fv->n_nodes == 0 || f_fv->n_nodes == 0) return ORBX_OK;
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    const int nnk = kf_fv->n_nodes, nnf = f_fv->n_nodes;
    const int nfk = kf_fv->node_start[nnk], nff = f_fv->node_start[nnf];
    Carver c;
    const size_t oKd = c.take(32 * (size_t)nkf), oKa = c.take(4 * (size_t)nkf), oKm = c.take(nkf),
                 oKi = c.take(4 * (size_t)nnk), oKs = c.take(4 * (size_t)(nnk + 1)), oKf = c.take(4 * (size_t)std::max(nfk, 1));
    const size_t oFd = c.take(32 * (size_t)nf), oFa = c.take(4 * (size_t)nf), oFi = c.take(4 * (size_t)nnf),
                 oFs = c.take(4 * (size_t)(nnf + 1)), oFf = c.take(4 * (size_t)std::max(nff, 1));
    const size_t oOut = c.take(4 * (size_t)nf), oBin = c.take(4 * (size_t)nf), oN = c.take(64);
    int rc = ensure_arena(m, c.off);
    if (rc != ORBX_OK) return rc;
    char* A = m->arena;
#define UP(off, ptr, count, T) if ((rc = up<T>(m, (T*)(A + off), ptr, count)) != ORBX_OK) return rc
    UP(oKd, kf_desc, 32 * (size_t)nkf, uint8_t); UP(oKa, kf_angle, nkf, float); UP(oKm, kf_has_mp, nkf, uint8_t);
    UP(oKi, kf_fv->node_id, nnk, int32_t); UP(oKs, kf_fv->node_start, nnk + 1, int32_t); UP(oKf, kf_fv->node_feat, nfk, int32_t);
    UP(oFd, f_desc, 32 * (size_t)nf, uint8_t); UP(oFa, f_angle, nf, float);
    UP(oFi, f_fv->node_id, nnf, int32_t); UP(oFs, f_fv->node_start, nnf + 1, int32_t); UP(oFf, f_fv->node_feat, nff, int32_t);
#undef UP
    if ((rc = flush_up(m)) != ORBX_OK) return rc;
    ORBX_CUDA_TRY(cudaMemsetAsync(A + oOut, 0xff, 4 * (size_t)nf, m->stream));
    ORBX_CUDA_TRY(cudaMemsetAsync(A + oBin, 0, 4 * (size_t)nf, m->stream));
    k_bow_nodes<<<(nnk + 3) / 4, 128, 0, m->stream>>>((uint8_t*)(A + oKd), (float*)(A + oKa), (uint8_t*)(A + oKm), nnk,
                                                     (int32_t*)(A + oKi), (int32_t*)(A + oKs), (int32_t*)(A + oKf),
                                                     (uint8_t*)(A + oFd), (float*)(A + oFa), nnf, (int32_t*)(A + oFi),
                                                     (int32_t*)(A + oFs), (int32_t*)(A + oFf), nnratio, nullptr, 0,
                                                     (int32_t*)(A + oOut), (int32_t*)(A + oBin));
    k_bow_rotation<<<1, 256, 0, m->stream>>>(nf, (int32_t*)(A + oOut), (int32_t*)(A + oBin), check_orientation, (int*)(A + oN));
    ORBX_CUDA_TRY(cudaGetLastError());
    ORBX_CUDA_TRY(cudaMemcpyAsync(out_match, A + oOut, 4 * (size_t)nf, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaMemcpyAsync(nmatches, A + oN, sizeof(int), cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

int orbm_set_stereo_outlier_mode(orbm_context* m, int mode)
{
    if (!m || (mode != 0 && mode != 1)) {
        set_error("orbm_set_stereo_outlier_mode: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    m->stereoOutlierMode = mode;
    return ORBX_OK;
}

int orbm_stereo_match(orbm_context* m, orbx_handle* ex_left, int frame_left, orbx_handle* ex_right, int frame_right, int nl,
                      const float* lx, const float* ly, const int32_t* loct, const uint8_t* ldesc, int nr, const float* rx,
                      const float* ry, const int32_t* roct, const uint8_t* rdesc, float mbf, float mb, float* u_right,
                      float* depth, int* n_matched)
{
    if (!m || !ex_left || !ex_right || nl < 0 || nr < 0 || !u_right || !depth || nr >= (1 << 22) || !(mb > 0)) {
        set_error("orbm_stereo_match: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    if (n_matched) *n_matched = 0;
    if (nl == 0) return ORBX_OK;
    PyramidView vl, vr;
    int rc = pyramid_view(ex_left, frame_left, &vl);
    if (rc == ORBX_OK) rc = pyramid_view(ex_right, frame_right, &vr);
    if (rc != ORBX_OK) return rc;
    if (vl.device != m->device || vr.device != m->device || vl.nlevels != vr.nlevels) {
        set_error("orbm_stereo_match: extractor handles and matcher context must share a device and a level count");
        return ORBX_ERR_BAD_ARG;
    }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    StereoPyr pl, pr;
    for (int l = 0; l < ORBX_MAX_LEVELS; l++) {
        const bool ok = l < vl.nlevels;
        pl.interior[l] = ok ? vl.interior[l] : nullptr; pr.interior[l] = ok ? vr.interior[l] : nullptr;
        pl.pitch[l] = ok ? vl.pitch[l] : 0; pr.pitch[l] = ok ? vr.pitch[l] : 0;
        pl.w[l] = ok ? vl.w[l] : 0; pr.w[l] = ok ? vr.w[l] : 0;
        pl.scale[l] = ok ? vl.scale[l] : 0; pr.scale[l] = ok ? vr.scale[l] : 0;
        pl.invScale[l] = ok ? vl.invScale[l] : 0; pr.invScale[l] = ok ? vr.invScale[l] : 0;
    }
    const float minD = 0.0f, maxD = mbf / mb;  // src/Frame.cc:711-713 with decision D5 (minZ = mb)
    const int nRows = vl.h[0];
    Carver c;
    const size_t oLx = c.take(4 * (size_t)nl), oLy = c.take(4 * (size_t)nl), oLo = c.take(4 * (size_t)nl), oLd = c.take(32 * (size_t)nl);
    const int nrA = std::max(nr, 1);
    const size_t oRx = c.take(4 * (size_t)nrA), oRy = c.take(4 * (size_t)nrA), oRo = c.take(4 * (size_t)nrA), oRd = c.take(32 * (size_t)nrA),
                 oSf = c.take(4 * (size_t)ORBX_MAX_LEVELS);
    const size_t oBr = c.take(4 * (size_t)nl), oBd = c.take(4 * (size_t)nl), oU = c.take(4 * (size_t)nl), oD = c.take(4 * (size_t)nl),
                 oS = c.take(4 * (size_t)nl), oTh = c.take(4 * (size_t)nl), oEi = c.take(4 * (size_t)nl), oNe = c.take(16);
    if (4 * (size_t)nl > 200 * 1024) {
        set_error("orbm_stereo_match: more than 51200 left keypoints");
        return ORBX_ERR_BAD_ARG;
    }
    rc = ensure_arena(m, c.off);
    if (rc != ORBX_OK) return rc;
    char* A = m->arena;
#define UP(off, ptr, count, T) if ((rc = up<T>(m, (T*)(A + off), ptr, count)) != ORBX_OK) return rc
    UP(oLx, lx, nl, float); UP(oLy, ly, nl, float); UP(oLo, loct, nl, int32_t); UP(oLd, ldesc, 32 * (size_t)nl, uint8_t);
    UP(oRx, rx, nr, float); UP(oRy, ry, nr, float); UP(oRo, roct, nr, int32_t); UP(oRd, rdesc, 32 * (size_t)nr, uint8_t);
    UP(oSf, vl.scale, vl.nlevels, float);
#undef UP
    if ((rc = flush_up(m)) != ORBX_OK) return rc;
    k_stereo_rows<<<(nl + 7) / 8, 256, 0, m->stream>>>(nl, (float*)(A + oLx), (float*)(A + oLy), (int32_t*)(A + oLo),
                                                      (uint8_t*)(A + oLd), nr, (float*)(A + oRx), (float*)(A + oRy),
                                                      (int32_t*)(A + oRo), (uint8_t*)(A + oRd), (float*)(A + oSf), nRows, minD,
                                                      maxD, (int32_t*)(A + oBr), (int32_t*)(A + oBd));
    k_stereo_refine<<<(nl + 7) / 8, 256, 0, m->stream>>>(nl, (float*)(A + oLx), (float*)(A + oLy), (int32_t*)(A + oLo),
                                                        (float*)(A + oRx), (int32_t*)(A + oBr), pl, pr, mbf, minD, maxD,
                                                        (float*)(A + oU), (float*)(A + oD), (int32_t*)(A + oS));
    if (4 * (size_t)nl > 48 * 1024)
        ORBX_CUDA_TRY(cudaFuncSetAttribute(k_stereo_medians, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    k_stereo_medians<<<(nl + 7) / 8, 256, 4 * (size_t)nl, m->stream>>>(nl, (int32_t*)(A + oS), (float*)(A + oTh), (int32_t*)(A + oEi),
                                                                      (int32_t*)(A + oNe));
    k_stereo_reject<<<1, 1024, 0, m->stream>>>((int32_t*)(A + oS), (float*)(A + oTh), (int32_t*)(A + oEi), (int32_t*)(A + oNe),
                                              m->stereoOutlierMode, (float*)(A + oU), (float*)(A + oD));
    ORBX_CUDA_TRY(cudaGetLastError());
    ORBX_CUDA_TRY(cudaMemcpyAsync(u_right, A + oU, 4 * (size_t)nl, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaMemcpyAsync(depth, A + oD, 4 * (size_t)nl, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    if (n_matched) {
        int k = 0;
        for (int i = 0; i < nl; i++) k += u_right[i] >= 0;
        *n_matched = k;
    }
    return ORBX_OK;
}

struct orbm_vocab {
    int device, nNodes, L;
    char* mem;
    int32_t* childStart; int32_t* childIds; uint8_t* nodeDesc; int32_t* wordId; double* weight;
};

int orbm_vocab_create(orbm_context* m, int n_nodes, int L, const int32_t* child_start, const int32_t* child_ids,
                      const uint8_t* node_desc, const int32_t* word_id, const double* weight, orbm_vocab** out)
{
    if (!m || !out || n_nodes <= 0 || L <= 0 || !child_start || !child_ids || !node_desc || !word_id || !weight ||
        child_start[1] <= child_start[0]) {
        set_error("orbm_vocab_create: bad argument (the root, node 0, must have children)");
        return ORBX_ERR_BAD_ARG;
    }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    const int nEdges = child_start[n_nodes];
    Carver c;
    const size_t oW = c.take(8 * (size_t)n_nodes), oD = c.take(32 * (size_t)n_nodes), oS = c.take(4 * (size_t)(n_nodes + 1)),
                 oC = c.take(4 * (size_t)std::max(nEdges, 1)), oI = c.take(4 * (size_t)n_nodes);
    orbm_vocab* v = new orbm_vocab;
    v->device = m->device; v->nNodes = n_nodes; v->L = L; v->mem = nullptr;
    cudaError_t e = cudaMalloc(&v->mem, c.off);
    if (e != cudaSuccess) { set_error("orbm_vocab_create: %s", cudaGetErrorString(e)); delete v; return ORBX_ERR_CUDA; }
    v->weight = (double*)(v->mem + oW); v->nodeDesc = (uint8_t*)(v->mem + oD); v->childStart = (int32_t*)(v->mem + oS);
    v->childIds = (int32_t*)(v->mem + oC); v->wordId = (int32_t*)(v->mem + oI);
    int rc;
    if ((rc = up<double>(m, v->weight, weight, n_nodes)) != ORBX_OK || (rc = up<uint8_t>(m, v->nodeDesc, node_desc, 32 * (size_t)n_nodes)) != ORBX_OK ||
        (rc = up<int32_t>(m, v->childStart, child_start, n_nodes + 1)) != ORBX_OK || (rc = up<int32_t>(m, v->childIds, child_ids, nEdges)) != ORBX_OK ||
        (rc = up<int32_t>(m, v->wordId, word_id, n_nodes)) != ORBX_OK) {
        cudaFree(v->mem); delete v; return rc;
    }
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    *out = v;
    return ORBX_OK;
}

void orbm_vocab_destroy(orbm_vocab* v)
{
    if (!v) return;
    cudaSetDevice(v->device);
    cudaFree(v->mem);
    delete v;
}

int orbm_bow_transform(orbm_context* m, const orbm_vocab* v, const uint8_t* desc, int n, int levelsup, int32_t* word_id,
                       int32_t* node_id, double* weight)
{
    if (!m || !v || n < 0 || (n > 0 && (!desc || !word_id || !node_id || !weight)) || v->device != m->device) {
        set_error("orbm_bow_transform: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    if (n == 0) return ORBX_OK;
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    Carver c;
    const size_t oWt = c.take(8 * (size_t)n), oD = c.take(32 * (size_t)n), oW = c.take(4 * (size_t)n), oN = c.take(4 * (size_t)n);
    int rc = ensure_arena(m, c.off);
    if (rc != ORBX_OK) return rc;
    char* A = m->arena;
    if ((rc = up<uint8_t>(m, (uint8_t*)(A + oD), desc, 32 * (size_t)n)) != ORBX_OK) return rc;
    if ((rc = flush_up(m)) != ORBX_OK) return rc;
    k_bow_transform<<<(n + 7) / 8, 256, 0, m->stream>>>(v->L, v->childStart, v->childIds, v->nodeDesc, v->wordId, v->weight,
                                                       (uint8_t*)(A + oD), n, levelsup, (int32_t*)(A + oW), (int32_t*)(A + oN),
                                                       (double*)(A + oWt));
    ORBX_CUDA_TRY(cudaGetLastError());
    ORBX_CUDA_TRY(cudaMemcpyAsync(word_id, A + oW, 4 * (size_t)n, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaMemcpyAsync(node_id, A + oN, 4 * (size_t)n, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaMemcpyAsync(weight, A + oWt, 8 * (size_t)n, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

int orbm_distinctive_descriptors(orbm_context* m, int n_sets, const int32_t* set_start, const uint8_t* desc, int32_t* best_idx)
{
    if (!m || n_sets < 0 || (n_sets > 0 && (!set_start || !best_idx))) {
        set_error("orbm_distinctive_descriptors: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    if (n_sets == 0) return ORBX_OK;
    const int total = set_start[n_sets];
    if (total > 0 && !desc) { set_error("orbm_distinctive_descriptors: null descriptors"); return ORBX_ERR_BAD_ARG; }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    Carver c;
    const size_t oS = c.take(4 * (size_t)(n_sets + 1)), oD = c.take(32 * (size_t)std::max(total, 1)), oB = c.take(4 * (size_t)n_sets);
    int rc = ensure_arena(m, c.off);
    if (rc != ORBX_OK) return rc;
    char* A = m->arena;
    if ((rc = up<int32_t>(m, (int32_t*)(A + oS), set_start, n_sets + 1)) != ORBX_OK) return rc;
    if ((rc = up<uint8_t>(m, (uint8_t*)(A + oD), desc, 32 * (size_t)total)) != ORBX_OK) return rc;
    if ((rc = flush_up(m)) != ORBX_OK) return rc;
    k_distinctive<<<(n_sets + 3) / 4, 128, 0, m->stream>>>(n_sets, (int32_t*)(A + oS), (uint8_t*)(A + oD), (int32_t*)(A + oB));
    ORBX_CUDA_TRY(cudaGetLastError());
    ORBX_CUDA_TRY(cudaMemcpyAsync(best_idx, A + oB, 4 * (size_t)n_sets, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

int orbm_search_by_bow_kf(orbm_context* m, int n1, const uint8_t* desc1, const float* angle1, const uint8_t* has_mp1,
                          const orbm_featvec_view* fv1, int n2, const uint8_t* desc2, const float* angle2,
                          const uint8_t* has_mp2, const orbm_featvec_view* fv2, float nnratio, int check_orientation,
                          int32_t* out_match12, int* nmatches)
{
    if (!m || !fv1 || !fv2 || !out_match12 || !nmatches || n1 < 0 || n2 < 0) {
        set_error("orbm_search_by_bow_kf: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    for (int i = 0; i < n1; i++) out_match12[i] = -1;
    *nmatches = 0;
    if (n1 == 0 || n2 == 0 || fv1->n_nodes == 0 || fv2->n_nodes == 0) return ORBX_OK;
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    const int nn1 = fv1->n_nodes, nn2 = fv2->n_nodes;
    const int nf1 = fv1->node_start[nn1], nf2 = fv2->node_start[nn2];
    Carver c;
    const size_t o1d = c.take(32 * (size_t)n1), o1a = c.take(4 * (size_t)n1), o1m = c.take(n1), o1i = c.take(4 * (size_t)nn1),
                 o1s = c.take(4 * (size_t)(nn1 + 1)), o1f = c.take(4 * (size_t)std::max(nf1, 1));
    const size_t o2d = c.take(32 * (size_t)n2), o2a = c.take(4 * (size_t)n2), o2m = c.take(n2), o2i = c.take(4 * (size_t)nn2),
                 o2s = c.take(4 * (size_t)(nn2 + 1)), o2f = c.take(4 * (size_t)std::max(nf2, 1));
    const size_t o21 = c.take(4 * (size_t)n2), oBin = c.take(4 * (size_t)n2), o12 = c.take(4 * (size_t)n1), oN = c.take(64);
    int rc = ensure_arena(m, c.off);
    if (rc != ORBX_OK) return rc;
    char* A = m->arena;
#define UP(off, ptr, count, T) if ((rc = up<T>(m, (T*)(A + off), ptr, count)) != ORBX_OK) return rc
    UP(o1d, desc1, 32 * (size_t)n1, uint8_t); UP(o1a, angle1, n1, float); UP(o1m, has_mp1, n1, uint8_t);
    UP(o1i, fv1->node_id, nn1, int32_t); UP(o1s, fv1->node_start, nn1 + 1, int32_t); UP(o1f, fv1->node_feat, nf1, int32_t);
    UP(o2d, desc2, 32 * (size_t)n2, uint8_t); UP(o2a, angle2, n2, float); UP(o2m, has_mp2, n2, uint8_t);
    UP(o2i, fv2->node_id, nn2, int32_t); UP(o2s, fv2->node_start, nn2 + 1, int32_t); UP(o2f, fv2->node_feat, nf2, int32_t);
#undef UP
    if ((rc = flush_up(m)) != ORBX_OK) return rc;
    ORBX_CUDA_TRY(cudaMemsetAsync(A + o21, 0xff, 4 * (size_t)n2, m->stream));
    ORBX_CUDA_TRY(cudaMemsetAsync(A + oBin, 0, 4 * (size_t)n2, m->stream));
    ORBX_CUDA_TRY(cudaMemsetAsync(A + o12, 0xff, 4 * (size_t)n1, m->stream));
    k_bow_nodes<<<(nn1 + 3) / 4, 128, 0, m->stream>>>((uint8_t*)(A + o1d), (float*)(A + o1a), (uint8_t*)(A + o1m), nn1,
                                                     (int32_t*)(A + o1i), (int32_t*)(A + o1s), (int32_t*)(A + o1f),
                                                     (uint8_t*)(A + o2d), (float*)(A + o2a), nn2, (int32_t*)(A + o2i),
                                                     (int32_t*)(A + o2s), (int32_t*)(A + o2f), nnratio, (uint8_t*)(A + o2m), 1,
                                                     (int32_t*)(A + o21), (int32_t*)(A + oBin));
    k_bow_rotation<<<1, 256, 0, m->stream>>>(n2, (int32_t*)(A + o21), (int32_t*)(A + oBin), check_orientation, (int*)(A + oN));
    k_bow_scatter12<<<(n2 + 255) / 256, 256, 0, m->stream>>>(n2, (int32_t*)(A + o21), (int32_t*)(A + o12));
    ORBX_CUDA_TRY(cudaGetLastError());
    ORBX_CUDA_TRY(cudaMemcpyAsync(out_match12, A + o12, 4 * (size_t)n1, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaMemcpyAsync(nmatches, A + oN, sizeof(int), cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}


// uploads one frame view + query view, launches k_search_best, queues the D2H of the results (no synchronisation)
static int search_best_enqueue(orbm_context* m, const char* who, const orbm_frame_view* f, const orbm_window_query_view* q,
                               int gate, const float* inv_level_sigma2, int max_dist, int32_t* best_idx, int32_t* best_dist)
{
    const int nkp = f->n, nq = q->n;
    const int ncell = GRID_COLS * GRID_ROWS;
    const int nitems = f->cell_start[ncell];
    for (int ci = 0; ci < ncell; ci++)
        if (f->cell_start[ci + 1] - f->cell_start[ci] >= 2048) {
            set_error("%s: a grid cell holds %d keypoints (limit 2047)", who, f->cell_start[ci + 1] - f->cell_start[ci]);
            return ORBX_ERR_BAD_ARG;
        }
    const int nlev = std::max(f->nlevels, 1);
    Carver c;
    const size_t oKx = c.take(4 * (size_t)nkp), oKy = c.take(4 * (size_t)nkp), oKo = c.take(4 * (size_t)nkp),
                 oUr = c.take(4 * (size_t)nkp), oKd = c.take(32 * (size_t)nkp), oCs = c.take(4 * (size_t)(ncell + 1)),
                 oCi = c.take(4 * (size_t)std::max(nitems, 1)), oSig = c.take(4 * (size_t)nlev);
    const size_t oPx = c.take(4 * (size_t)nq), oPy = c.take(4 * (size_t)nq), oPxr = c.take(4 * (size_t)nq),
                 oPr = c.take(4 * (size_t)nq), oPmin = c.take(4 * (size_t)nq), oPmax = c.take(4 * (size_t)nq),
                 oPva = c.take(nq), oPd = c.take(32 * (size_t)nq), oBi = c.take(4 * (size_t)nq), oBd = c.take(4 * (size_t)nq);
    int rc = ensure_arena(m, c.off);
    if (rc != ORBX_OK) return rc;
    char* A = m->arena;
#define UP(off, ptr, count, T) if ((rc = up<T>(m, (T*)(A + off), ptr, count)) != ORBX_OK) return rc
    UP(oKx, f->x, nkp, float); UP(oKy, f->y, nkp, float); UP(oKo, f->octave, nkp, int32_t);
    if (f->u_right) UP(oUr, f->u_right, nkp, float);
    UP(oKd, f->desc, 32 * (size_t)nkp, uint8_t);
    UP(oCs, f->cell_start, ncell + 1, int32_t); UP(oCi, f->cell_items, nitems, int32_t);
    if (gate == 1) UP(oSig, inv_level_sigma2, f->nlevels, float);
    UP(oPx, q->x, nq, float); UP(oPy, q->y, nq, float);
    if (q->xr) UP(oPxr, q->xr, nq, float);
    UP(oPr, q->radius, nq, float); UP(oPmin, q->min_level, nq, int32_t); UP(oPmax, q->max_level, nq, int32_t);
    UP(oPva, q->valid, nq, uint8_t); UP(oPd, q->desc, 32 * (size_t)nq, uint8_t);
#undef UP
    if ((rc = flush_up(m)) != ORBX_OK) return rc;
    BestArgs a;
    a.kx = (float*)(A + oKx); a.ky = (float*)(A + oKy); a.koct = (int32_t*)(A + oKo);
    a.uright = f->u_right ? (float*)(A + oUr) : nullptr;
    a.kdesc = (uint8_t*)(A + oKd); a.cellStart = (int32_t*)(A + oCs); a.cellItems = (int32_t*)(A + oCi);
    a.minX = f->min_x; a.minY = f->min_y; a.wInv = f->grid_w_inv; a.hInv = f->grid_h_inv;
    a.nq = nq;
    a.qx = (float*)(A + oPx); a.qy = (float*)(A + oPy); a.qr = (float*)(A + oPr); a.qmin = (int32_t*)(A + oPmin);
    a.qmax = (int32_t*)(A + oPmax); a.qxr = q->xr ? (float*)(A + oPxr) : nullptr;
    a.qvalid = (uint8_t*)(A + oPva); a.qdesc = (uint8_t*)(A + oPd);
    a.gate = gate; a.maxDist = max_dist; a.invSigma2 = (float*)(A + oSig);
    a.bestIdx = (int32_t*)(A + oBi); a.bestDist = (int32_t*)(A + oBd);
    k_search_best<<<(nq + 7) / 8, 256, 0, m->stream>>>(a);
    ORBX_CUDA_TRY(cudaGetLastError());
    ORBX_CUDA_TRY(cudaMemcpyAsync(best_idx, a.bestIdx, 4 * (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
    if (best_dist) ORBX_CUDA_TRY(cudaMemcpyAsync(best_dist, a.bestDist, 4 * (size_t)nq, cudaMemcpyDeviceToHost, m->stream));
    return ORBX_OK;
}

static bool best_args_ok(const orbm_frame_view* f, const orbm_window_query_view* q, int gate, const float* sig)
{
    if (!f || !q || f->n < 0 || q->n < 0) return false;
    if (f->n > 0 && (!f->x || !f->y || !f->octave || !f->desc || !f->cell_start || !f->cell_items)) return false;
    if (q->n > 0 && (!q->x || !q->y || !q->radius || !q->min_level || !q->max_level || !q->valid || !q->desc)) return false;
    if (gate != 0 && gate != 1) return false;
    if (gate == 1 && (!sig || f->nlevels <= 0 || (f->u_right && !q->xr))) return false;
    return true;
}

int orbm_search_best(orbm_context* m, const orbm_frame_view* f, const orbm_window_query_view* q, int gate,
                     const float* inv_level_sigma2, int max_dist, int32_t* best_idx, int32_t* best_dist)
{
    if (!m || !best_idx || !best_args_ok(f, q, gate, inv_level_sigma2)) {
        set_error("orbm_search_best: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    if (q->n == 0) return ORBX_OK;
    if (f->n == 0) {
        for (int i = 0; i < q->n; i++) { best_idx[i] = -1; if (best_dist) best_dist[i] = 256; }
        return ORBX_OK;
    }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    int rc = search_best_enqueue(m, "orbm_search_best", f, q, gate, inv_level_sigma2, max_dist, best_idx, best_dist);
    if (rc != ORBX_OK) return rc;
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

int orbm_search_by_sim3(orbm_context* m, const orbm_frame_view* kf1, const orbm_window_query_view* q2_in_1,
                        const orbm_frame_view* kf2, const orbm_window_query_view* q1_in_2, int32_t* out_match12, int* nfound)
{
    if (!m || !out_match12 || !nfound || !best_args_ok(kf1, q2_in_1, 0, nullptr) || !best_args_ok(kf2, q1_in_2, 0, nullptr) ||
        q1_in_2->n != kf1->n || q2_in_1->n != kf2->n) {
        set_error("orbm_search_by_sim3: bad argument (q1_in_2 needs one entry per KF1 feature, q2_in_1 one per KF2 feature)");
        return ORBX_ERR_BAD_ARG;
    }
    const int n1 = kf1->n, n2 = kf2->n;
    *nfound = 0;
    for (int i = 0; i < n1; i++) out_match12[i] = -1;
    if (n1 == 0 || n2 == 0) return ORBX_OK;
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    std::vector<int32_t> vnMatch1(n1), vnMatch2(n2);
    int rc = search_best_enqueue(m, "orbm_search_by_sim3", kf2, q1_in_2, 0, nullptr, TH_HIGH, vnMatch1.data(), nullptr);
    if (rc != ORBX_OK) return rc;
    rc = search_best_enqueue(m, "orbm_search_by_sim3", kf1, q2_in_1, 0, nullptr, TH_HIGH, vnMatch2.data(), nullptr);
    if (rc != ORBX_OK) return rc;
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    // agreement check, src/ORBmatcher.cc:1585-1602
    int found = 0;
    for (int i1 = 0; i1 < n1; i1++) {
        const int idx2 = vnMatch1[i1];
        if (idx2 >= 0 && vnMatch2[idx2] == i1) { out_match12[i1] = idx2; found++; }
    }
    *nfound = found;
    return ORBX_OK;
}


int orbm_search_for_initialization(orbm_context* m, int n1, const int32_t* octave1, const float* angle1, const uint8_t* desc1,
                                   const float* prev_x, const float* prev_y, const orbm_frame_view* f2, const float* f2_angle,
                                   int window_size, float nnratio, int check_orientation, int32_t* out_match12, int* nmatches)
{
    if (!m || n1 < 0 || !f2 || f2->n < 0 || !out_match12 || !nmatches ||
        (n1 > 0 && (!octave1 || !desc1 || !prev_x || !prev_y)) || (check_orientation && n1 > 0 && f2->n > 0 && (!angle1 || !f2_angle))) {
        set_error("orbm_search_for_initialization: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    const int n2 = f2->n;
    *nmatches = 0;
    for (int i = 0; i < n1; i++) out_match12[i] = -1;
    if (n1 == 0 || n2 == 0) return ORBX_OK;
    const int ncell = GRID_COLS * GRID_ROWS;
    const int nitems = f2->cell_start[ncell];
    for (int ci = 0; ci < ncell; ci++)
        if (f2->cell_start[ci + 1] - f2->cell_start[ci] >= 2048) {
            set_error("orbm_search_for_initialization: a grid cell holds %d keypoints (limit 2047)", f2->cell_start[ci + 1] - f2->cell_start[ci]);
            return ORBX_ERR_BAD_ARG;
        }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    Carver c;
    const size_t o1o = c.take(4 * (size_t)n1), o1a = c.take(4 * (size_t)n1), o1d = c.take(32 * (size_t)n1), o1x = c.take(4 * (size_t)n1),
                 o1y = c.take(4 * (size_t)n1);
    const size_t oKx = c.take(4 * (size_t)n2), oKy = c.take(4 * (size_t)n2), oKo = c.take(4 * (size_t)n2), oKa = c.take(4 * (size_t)n2),
                 oKd = c.take(32 * (size_t)n2), oCs = c.take(4 * (size_t)(ncell + 1)), oCi = c.take(4 * (size_t)std::max(nitems, 1));
    const size_t oRa = c.take(4 * (size_t)n1), oRb = c.take(4 * (size_t)n1), oDa = c.take(4 * (size_t)n1), oDb = c.take(4 * (size_t)n1),
                 oHead = c.take(4 * (size_t)n2), oHeadB = c.take(4 * (size_t)n2), oNext = c.take(4 * (size_t)n1), oOut = c.take(4 * (size_t)n1), oMisc = c.take(512);
    int rc = ensure_arena(m, c.off);
    if (rc != ORBX_OK) return rc;
    char* A = m->arena;
#define UP(off, ptr, count, T) if ((rc = up<T>(m, (T*)(A + off), ptr, count)) != ORBX_OK) return rc
    UP(o1o, octave1, n1, int32_t); UP(o1d, desc1, 32 * (size_t)n1, uint8_t); UP(o1x, prev_x, n1, float); UP(o1y, prev_y, n1, float);
    if (check_orientation) { UP(o1a, angle1, n1, float); UP(oKa, f2_angle, n2, float); }
    UP(oKx, f2->x, n2, float); UP(oKy, f2->y, n2, float); UP(oKo, f2->octave, n2, int32_t); UP(oKd, f2->desc, 32 * (size_t)n2, uint8_t);
    UP(oCs, f2->cell_start, ncell + 1, int32_t); UP(oCi, f2->cell_items, nitems, int32_t);
#undef UP
    if ((rc = flush_up(m)) != ORBX_OK) return rc;
    InitArgs a;
    a.n1 = n1; a.oct1 = (int32_t*)(A + o1o); a.angle1 = (float*)(A + o1a); a.desc1 = (uint8_t*)(A + o1d);
    a.prevx = (float*)(A + o1x); a.prevy = (float*)(A + o1y);
    a.n2 = n2; a.kx = (float*)(A + oKx); a.ky = (float*)(A + oKy); a.koct = (int32_t*)(A + oKo); a.kangle = (float*)(A + oKa);
    a.kdesc = (uint8_t*)(A + oKd); a.cellStart = (int32_t*)(A + oCs); a.cellItems = (int32_t*)(A + oCi);
    a.minX = f2->min_x; a.minY = f2->min_y; a.wInv = f2->grid_w_inv; a.hInv = f2->grid_h_inv;
    a.window = (float)window_size; a.nnratio = nnratio; a.checkOri = check_orientation ? 1 : 0;
    a.resA = (int32_t*)(A + oRa); a.resB = (int32_t*)(A + oRb); a.distA = (int32_t*)(A + oDa); a.distB = (int32_t*)(A + oDb);
    a.head = (int32_t*)(A + oHead); a.headB = (int32_t*)(A + oHeadB); a.next = (int32_t*)(A + oNext); a.out12 = (int32_t*)(A + oOut);
    a.nmatches = (int*)(A + oMisc); a.sweeps = a.nmatches + 1; a.misc = a.nmatches + 8;
    {
        int perSM = 0;
        ORBX_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, k_init_search, INIT_T, 0));
        const int want = (n1 + INIT_T / 32 - 1) / (INIT_T / 32);  // one warp per F1 feature
        const int grid = std::max(1, std::min(want, perSM * m->numSMs));
        void* args[] = {&a};
        ORBX_CUDA_TRY(cudaLaunchCooperativeKernel((void*)k_init_search, dim3(grid), dim3(INIT_T), args, 0, m->stream));
    }
    ORBX_CUDA_TRY(cudaMemcpyAsync(out_match12, a.out12, 4 * (size_t)n1, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaMemcpyAsync(nmatches, a.nmatches, sizeof(int), cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

int orbm_search_for_triangulation(orbm_context* m, const orbm_kf_view* kf1, const orbm_kf_view* kf2, const float* F12, float ex,
                                  float ey, const float* scale_factors2, const float* level_sigma2_2, int nlevels2,
                                  int only_stereo, int check_orientation, int32_t* out_match12, int* nmatches)
{
    if (!m || !kf1 || !kf2 || !F12 || !scale_factors2 || !level_sigma2_2 || nlevels2 <= 0 || !out_match12 || !nmatches ||
        kf1->n < 0 || kf2->n < 0 || !kf1->fv || !kf2->fv) {
        set_error("orbm_search_for_triangulation: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    const int n1 = kf1->n, n2 = kf2->n;
    *nmatches = 0;
    for (int i = 0; i < n1; i++) out_match12[i] = -1;
    const int nn1 = kf1->fv->n_nodes, nn2 = kf2->fv->n_nodes;
    if (n1 == 0 || n2 == 0 || nn1 == 0 || nn2 == 0) return ORBX_OK;
    for (int i = 0; i < n2; i++)
        if (kf2->octave[i] < 0 || kf2->octave[i] >= nlevels2) {
            set_error("orbm_search_for_triangulation: KF2 octave %d outside [0, %d)", kf2->octave[i], nlevels2);
            return ORBX_ERR_BAD_ARG;
        }
    ORBX_CUDA_TRY(cudaSetDevice(m->device));
    const int nf1 = kf1->fv->node_start[nn1], nf2 = kf2->fv->node_start[nn2];
    Carver c;
    const size_t o1x = c.take(4 * (size_t)n1), o1y = c.take(4 * (size_t)n1), o1a = c.take(4 * (size_t)n1), o1u = c.take(4 * (size_t)n1),
                 o1d = c.take(32 * (size_t)n1), o1m = c.take(n1), o1i = c.take(4 * (size_t)nn1), o1s = c.take(4 * (size_t)(nn1 + 1)),
                 o1f = c.take(4 * (size_t)std::max(nf1, 1));
    const size_t o2x = c.take(4 * (size_t)n2), o2y = c.take(4 * (size_t)n2), o2o = c.take(4 * (size_t)n2), o2a = c.take(4 * (size_t)n2),
                 o2u = c.take(4 * (size_t)n2), o2d = c.take(32 * (size_t)n2), o2m = c.take(n2), o2i = c.take(4 * (size_t)nn2),
                 o2s = c.take(4 * (size_t)(nn2 + 1)), o2f = c.take(4 * (size_t)std::max(nf2, 1));
    const size_t oSc = c.take(4 * (size_t)nlevels2), oSg = c.take(4 * (size_t)nlevels2), oM2 = c.take(4 * (size_t)n2),
                 o12 = c.take(4 * (size_t)n1), oBin = c.take(4 * (size_t)n1), oN = c.take(64);
    int rc = ensure_arena(m, c.off);
    if (rc != ORBX_OK) return rc;
    char* A = m->arena;
#define UP(off, ptr, count, T) if ((rc = up<T>(m, (T*)(A + off), ptr, count)) != ORBX_OK) return rc
    UP(o1x, kf1->x, n1, float); UP(o1y, kf1->y, n1, float); UP(o1a, kf1->angle, n1, float); UP(o1u, kf1->u_right, n1, float);
    UP(o1d, kf1->desc, 32 * (size_t)n1, uint8_t); UP(o1m, kf1->has_mp, n1, uint8_t);
    UP(o1i, kf1->fv->node_id, nn1, int32_t); UP(o1s, kf1->fv->node_start, nn1 + 1, int32_t); UP(o1f, kf1->fv->node_feat, nf1, int32_t);
    UP(o2x, kf2->x, n2, float); UP(o2y, kf2->y, n2, float); UP(o2o, kf2->octave, n2, int32_t); UP(o2a, kf2->angle, n2, float);
    UP(o2u, kf2->u_right, n2, float); UP(o2d, kf2->desc, 32 * (size_t)n2, uint8_t); UP(o2m, kf2->has_mp, n2, uint8_t);
    UP(o2i, kf2->fv->node_id, nn2, int32_t); UP(o2s, kf2->fv->node_start, nn2 + 1, int32_t); UP(o2f, kf2->fv->node_feat, nf2, int32_t);
    UP(oSc, scale_factors2, nlevels2, float); UP(oSg, level_sigma2_2, nlevels2, float);
#undef UP
    if ((rc = flush_up(m)) != ORBX_OK) return rc;
    ORBX_CUDA_TRY(cudaMemsetAsync(A + oM2, 0, 4 * (size_t)n2, m->stream));
    ORBX_CUDA_TRY(cudaMemsetAsync(A + o12, 0xff, 4 * (size_t)n1, m->stream));
    ORBX_CUDA_TRY(cudaMemsetAsync(A + oBin, 0, 4 * (size_t)n1, m->stream));
    TriArgs a;
    a.x1 = (float*)(A + o1x); a.y1 = (float*)(A + o1y); a.angle1 = (float*)(A + o1a); a.ur1 = (float*)(A + o1u);
    a.desc1 = (uint8_t*)(A + o1d); a.hasMp1 = (uint8_t*)(A + o1m);
    a.nNodes1 = nn1; a.nodeId1 = (int32_t*)(A + o1i); a.nodeStart1 = (int32_t*)(A + o1s); a.nodeFeat1 = (int32_t*)(A + o1f);
    a.x2 = (float*)(A + o2x); a.y2 = (float*)(A + o2y); a.oct2 = (int32_t*)(A + o2o); a.angle2 = (float*)(A + o2a);
    a.ur2 = (float*)(A + o2u); a.desc2 = (uint8_t*)(A + o2d); a.hasMp2 = (uint8_t*)(A + o2m);
    a.nNodes2 = nn2; a.nodeId2 = (int32_t*)(A + o2i); a.nodeStart2 = (int32_t*)(A + o2s); a.nodeFeat2 = (int32_t*)(A + o2f);
    for (int i = 0; i < 9; i++) a.F[i] = F12[i];
    a.ex = ex; a.ey = ey; a.scale2 = (float*)(A + oSc); a.sigma2_2 = (float*)(A + oSg); a.onlyStereo = only_stereo ? 1 : 0;
    a.matched2 = (int32_t*)(A + oM2); a.out12 = (int32_t*)(A + o12); a.binOf = (int32_t*)(A + oBin);
    k_triang_nodes<<<(nn1 + 3) / 4, 128, 0, m->stream>>>(a);
    k_bow_rotation<<<1, 256, 0, m->stream>>>(n1, a.out12, a.binOf, check_orientation, (int*)(A + oN));
    ORBX_CUDA_TRY(cudaGetLastError());
    ORBX_CUDA_TRY(cudaMemcpyAsync(out_match12, a.out12, 4 * (size_t)n1, cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaMemcpyAsync(nmatches, A + oN, sizeof(int), cudaMemcpyDeviceToHost, m->stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(m->stream));
    return ORBX_OK;
}

}  // extern "C"
