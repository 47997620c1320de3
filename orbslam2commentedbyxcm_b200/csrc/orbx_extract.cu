// orbx_extract.cu -- B200 (sm_100a) ORB extractor: pyramid, per-cell FAST, quad-tree distribution + orientation,
// 7x7 Gaussian blur, steered BRIEF.  One launch per stage covers every frame of a batch (grid.y / grid.z = frame)
// and every pyramid level that stage can process concurrently.
//
// Reference behaviour reproduced (paths in the reference tree):
//   k_level0 / k_resize_rows(_tma)  ORBextractor::ComputePyramid src/ORBextractor.cc:1635-1694
//                        (fallbacks k_resize_words, k_resize)
//   k_fast_band          cell loop of ComputeKeyPointsOctTree    src/ORBextractor.cc:1017-1122 (+ cv::FAST 9/16, NMS)
//   k_octree             DistributeOctTree / DivideNode          src/ORBextractor.cc:581-1013, :1124-1164
//                        + IC_Angle / cv::fastAtan2              src/ORBextractor.cc:59-106, :559-570
//   k_blur_tma / k_blur  cv::GaussianBlur 7x7 sigma 2            src/ORBextractor.cc:1587-1595
//   k_describe           computeOrbDescriptor                    src/ORBextractor.cc:118-172, :1613-1622
// Results are bit-identical to the CPU oracle (oracle/orb_oracle.cc); tests/test_gpu_extractor.py checks that.
// Blackwell paths: the FAST bands, the blur tiles and the resize source windows are staged by TMA bulk tensor copies
// (cp.async.bulk.tensor.3d + mbarrier, per-level tensor maps as __grid_constant__ arguments); every such kernel has a
// non-TMA variant of the same arithmetic (environment switches in INTEGRATION.md section 5, tested equal).
#include "orbx_common.cuh"
#include "sincosf_dev.cuh"
#include "orbx_internal.h"
#include <cuda.h>   // CUtensorMap (the encoder is fetched through cudaGetDriverEntryPoint: no link against libcuda)
#include <cmath>
#include <cfloat>
#include <cstdarg>
#include <cstdlib>
#include <vector>
#include <utility>
#include <algorithm>
#include <mutex>

namespace orbx {

std::string& last_error()
{
    static thread_local std::string e;
    return e;
}
void set_error(const char* fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    last_error() = buf;
}

// ------------------------------------------------------------------------------------------------------
// Plan: everything that depends only on (config, image shape).  Passed to kernels by value.
// ------------------------------------------------------------------------------------------------------
struct LevelGeom {
    int w, h;            // interior size of the level
    int pitch;           // bytes per buffer row (multiple of 128)
    int planeRows;       // h + 2*EDGE
    long long planeOff;  // byte offset of frame 0's plane of this level inside the pyramid (and blurred) arena
    long long planeBytes;
    int nCols, nRows, wCell, hCell;  // 30-px cell grid (src/ORBextractor.cc:1043-1051)
    int wCellMagic;      // 65536 / wCell + 1: (x * wCellMagic) >> 16 == x / wCell for 0 <= x < 1024 (wCell >= 30)
    int maxBX, maxBY;    // w - 16, h - 16
    int blurOff;         // first CTA index of this level in k_blur
    int fastG, fastGroups, fastOff;  // k_fast_band: cells per CTA, CTAs per cell row, first CTA index of the level
    int quota;           // mnFeaturesPerLevel
    int candCap;         // capacity of the candidate list (worst case after 3x3 NMS)
    long long candOff;   // offset (in candidates) of frame 0's list of this level
    int nodeCap;         // quad-tree node capacity = per-level keypoint capacity
    int kpOff;           // offset of this level in a frame's keypoint staging
    int nIni;            // quad-tree roots
    float hX;            // root width
    float scale;         // mvScaleFactor[level]
    int patchSize;       // (int)(31 * scale)
    int xTab, yTab;      // offsets into the resize coefficient tables
    int resizeWords;     // 1: k_resize_words applies (4-column groups read within 8 source columns)
    int resizeRows;      // 1: k_resize_rows applies too (every destination row reads source rows r0, r0 + 1, r0 strictly increasing)
    int sTab;            // offset of this level in the source-row table (k_resize_rows)
    int resizeTma;       // 1: every CTA's source window of k_resize_rows_tma fits its RZ_TP x RZ_TROWS tile
};

struct Plan {
    int nlevels, rows, cols;
    int iniTh, minTh;
    int totalFastUnits, totalBlurUnits;
    int fastTileBytes, fastQCap;   // k_fast_band shared-memory carve-up (fastQCap = capacity of the post-NMS queue)
    int kpCapFrame;          // sum of nodeCap = keypoint capacity per frame
    long long pyrFrameStride;  // not used for addressing (planes are [level][frame]); kept for sizing
    long long candFrame;       // candidates per frame (sum of candCap)
    LevelGeom lv[ORBX_MAX_LEVELS];
};

constexpr int FAST_TMA_LEVELS = 8;  // levels served by the TMA-staged kernels (per-level tensor maps passed as kernel arguments)

struct Cand {  // one FAST candidate
    uint32_t xy;  // x | y << 16, relative to (minBorderX, minBorderY) = (16, 16) like the reference's vToDistributeKeys
    uint32_t ro;  // response << 24 | (0xFFFFFF - ord); ord = rank key of the reference's candidate order
};

struct FastUnit {  // one k_fast_band CTA (16 bytes, geometry precomputed on the host): level, cell row, first cell column,
    short level, ci, j0, iniX, iniY, rh, bw, pad;  // band origin in level pixels, band rows, band width (rh, bw >= 7)
};

struct StagedKp {  // per-level keypoint between k_octree and k_describe
    float x, y;    // level pixel coordinates (integers)
    float angle;
    float response;
};

struct Buffers {
    uint8_t* pyr;        // [level][frame] padded planes
    uint8_t* blur;       // same layout; only interiors are written
    Cand* cand;          // [frame][level] lists
    int* candCount;      // [frame][nlevels] atomic counters (reset by k_octree)
    int* candCountDbg;   // [frame][nlevels] copy kept for introspection
    StagedKp* staged;    // [frame][kpCapFrame]
    int* levelCount;     // [frame][nlevels]
    const short* xofs;   // resize tables
    const short2* xw;
    const short* yofs0;
    const short* yofs1;
    const short2* yw;
    const int4* yrow;    // k_resize_words row table: {r0 * srcPitch, r1 * srcPitch, b0 | b1 << 16, mirrored frame rows as deltas}
    const int2* srow;    // k_resize_rows table, per SOURCE row r: {b0 | b1 << 16 of the destination row whose lower source row is r
                         // (0: none), mirrored frame rows of that destination row as deltas}
    const FastUnit* fastUnits;  // [totalFastUnits]
    const uint8_t* zeros;  // fastTileBytes zero bytes (k_fast_band<TMA> clears its score tile with a bulk copy of them)
    int* errFlag;        // device-side overflow flag
    int frames;
    int frameBase;       // arena slot offset: arena frame = launch-local frame index + frameBase
};

// rBRIEF pattern: 256 pairs x (x0,y0,x1,y1), 1024 signed bytes
__device__ const signed char d_pattern[1024] __attribute__((aligned(16))) = {
#include "orb_pattern.inc"
};

// Programmatic dependent launch (the kernels of a batch are a chain; bench-sized launches last 20-100 us and the hand-over
// between two of them costs a few us): every kernel lets its successor be scheduled as soon as all of its own CTAs have
// started (pdl_trigger at entry), does the part of its prologue that reads only kernel arguments and plan-time tables, and
// waits for its predecessor to COMPLETE (pdl_wait: all memory of the predecessor visible; also orders this kernel's writes
// after the predecessor's reads) before it touches anything a batch kernel reads or writes.  Both are no-ops when the launch
// does not carry the attribute (ORBX_PDL=0, profiling mode).
__device__ __forceinline__ void pdl_trigger()
{
#ifdef ORBX_PDL_EARLY_TRIGGER
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

__device__ __forceinline__ uint8_t* plane_ptr(uint8_t* arena, const LevelGeom& g, int frame)
{
    return arena + g.planeOff + (long long)frame * g.planeBytes;
}
__device__ __forceinline__ const uint8_t* interior_ptr(const uint8_t* arena, const LevelGeom& g, int frame)
{
    return arena + g.planeOff + (long long)frame * g.planeBytes + (long long)PAD_TOP * g.pitch + PAD_LEFT;
}

// ------------------------------------------------------------------------------------------------------
// K1a: level 0 = input image + REFLECT_101 frame.  A CTA writes L0_ROWS buffer rows of one frame, 16 lanes per row:
// first the interior as 128-bit copies (when the source rows are 16-byte aligned), then the few 32-bit words that hold
// the 19-px frame (and an unaligned interior tail) byte by byte through the reflected column.  Buffer columns outside
// [12, 32 + w + 19] are never read by any kernel and stay unwritten.
// ------------------------------------------------------------------------------------------------------
constexpr int L0_ROWS = 8;

__global__ void __launch_bounds__(16 * L0_ROWS) k_level0(Plan plan, Buffers buf, const uint8_t* __restrict__ in, long long inStep,
                                                         long long inFrameStride)
{
    pdl_trigger();
    const LevelGeom& g = plan.lv[0];
    const int lframe = blockIdx.y, frame = lframe + buf.frameBase;
    const int by = blockIdx.x * L0_ROWS + (threadIdx.x >> 4);
    pdl_wait();
    if (by >= g.planeRows) return;
    const int l16 = threadIdx.x & 15;
    const int y = reflect101(by - PAD_TOP, g.h);
    const uint8_t* src = in + (long long)lframe * inFrameStride + (long long)y * inStep;
    uint8_t* dst = plane_ptr(buf.pyr, g, frame) + (long long)by * g.pitch + PAD_LEFT;
    const bool aligned = ((reinterpret_cast<uintptr_t>(in) | (uintptr_t)inStep | (uintptr_t)inFrameStride) & 15) == 0;
    const int nfast = aligned ? g.w >> 4 : 0;
    for (int c = l16; c < nfast; c += 16) reinterpret_cast<uint4*>(dst)[c] = __ldg(reinterpret_cast<const uint4*>(src) + c);
    const int xr = 16 * nfast;                              // first interior column not covered by the 128-bit copies
    const int nEdge = 5 + ((g.w - xr + EDGE + 3) >> 2);     // 5 words left of the interior, the rest from column xr on
    for (int e = l16; e < nEdge; e += 16) {
        const int x0 = e < 5 ? 4 * e - 20 : xr + 4 * (e - 5);
        uint32_t wv = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            int x = min(max(x0 + j, -EDGE), g.w - 1 + EDGE);
            x = reflect101(x, g.w);
            wv |= (uint32_t)__ldg(src + x) << (8 * j);
        }
        *reinterpret_cast<uint32_t*>(dst + x0) = wv;
    }
}

// ------------------------------------------------------------------------------------------------------
// K1b: level l = cv::resize(level l-1, INTER_LINEAR) + REFLECT_101 frame, fixed point (11-bit weights):
//   T = S[sx]*a0 + S[sx+1]*a1;  dst = (((b0*(T0>>4))>>16) + ((b1*(T1>>4))>>16) + 2) >> 2
// A thread owns one aligned 32-bit word (4 columns) of the destination buffer row -- frame columns included, their
// source column being the reflected one -- and walks rowsPerThread interior rows: column taps and weights stay in
// registers (rowsPerThread rows per thread, fewer for small grids), the horizontally interpolated source row T1 of one output row is reused as T0 of the next whenever the
// source row advances by one (80 % of rows at scale 1.2), and rows that the frame mirrors are stored twice.
// ------------------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(128) k_resize(Plan plan, Buffers buf, int level, int rowsPerThread)
{
    const LevelGeom& g = plan.lv[level];
    const LevelGeom& gp = plan.lv[level - 1];
    const int bx = 4 * (blockIdx.x * 32 + (threadIdx.x & 31));  // pitch is a multiple of 128: always inside the row
    const int frame = blockIdx.z + buf.frameBase;
    const int dy0 = (blockIdx.y * 4 + (threadIdx.x >> 5)) * rowsPerThread;
    if (dy0 >= g.h) return;
    const int dy1 = min(dy0 + rowsPerThread, g.h);
    int sx[4], sx1[4];
    short2 a[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        int dx = min(max(bx + j - PAD_LEFT, -EDGE), g.w - 1 + EDGE);
        dx = reflect101(dx, g.w);
        sx[j] = buf.xofs[g.xTab + dx];
        sx1[j] = min(sx[j] + 1, gp.w - 1);
        a[j] = buf.xw[g.xTab + dx];
    }
    const uint8_t* S = interior_ptr(buf.pyr, gp, frame);
    uint8_t* D = plane_ptr(buf.pyr, g, frame) + bx;
    // T rows are kept as T >> 4 (what the vertical pass consumes); ((b * (T >> 4)) >> 16) is the high half of
    // (b << 16) * (T >> 4), one IMAD.HI per tap.
    uint32_t t0[4], t1[4];
    int rowInT1 = -1;
    for (int dy = dy0; dy < dy1; dy++) {
        const int r0 = buf.yofs0[g.yTab + dy], r1 = buf.yofs1[g.yTab + dy];
        const short2 b = buf.yw[g.yTab + dy];
        if (r0 == rowInT1) {
#pragma unroll
            for (int j = 0; j < 4; j++) t0[j] = t1[j];
        } else {
            const uint8_t* R = S + (long long)r0 * gp.pitch;
#pragma unroll
            for (int j = 0; j < 4; j++) t0[j] = (uint32_t)(R[sx[j]] * a[j].x + R[sx1[j]] * a[j].y) >> 4;
        }
        if (r1 == r0) {
#pragma unroll
            for (int j = 0; j < 4; j++) t1[j] = t0[j];
        } else {
            const uint8_t* R = S + (long long)r1 * gp.pitch;
#pragma unroll
            for (int j = 0; j < 4; j++) t1[j] = (uint32_t)(R[sx[j]] * a[j].x + R[sx1[j]] * a[j].y) >> 4;
        }
        rowInT1 = r1;
        const uint32_t b0 = (uint32_t)b.x << 16, b1 = (uint32_t)b.y << 16;
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; j++) o[j] = (__umulhi(b0, t0[j]) + __umulhi(b1, t1[j]) + 2u) >> 2;  // in [0, 255]
        const uint32_t v = __byte_perm(__byte_perm(o[0], o[1], 0x0040), __byte_perm(o[2], o[3], 0x0040), 0x5410);
        *reinterpret_cast<uint32_t*>(D + (long long)(PAD_TOP + dy) * g.pitch) = v;
        if (dy >= 1 && dy <= EDGE) *reinterpret_cast<uint32_t*>(D + (long long)(PAD_TOP - dy) * g.pitch) = v;
        if (dy >= g.h - 1 - EDGE && dy <= g.h - 2)
            *reinterpret_cast<uint32_t*>(D + (long long)(PAD_TOP + 2 * (g.h - 1) - dy) * g.pitch) = v;
    }
}

// K1b, word variant (used when every aligned group of 4 destination columns reads source bytes that lie within 8
// consecutive source columns, i.e. scale factors up to 2; build_plan checks the tables and sets LevelGeom::resizeWords):
// instead of 8 byte loads per source row a thread loads the 3 aligned words covering its window, funnel-shifts them to
// the window start and picks each column's (S[sx], S[sx+1]) pair with one PRMT (selector precomputed per thread); the
// horizontal tap pair is one IDP.2A.  Addresses are 32-bit offsets from a CTA-uniform base.
__global__ void __launch_bounds__(128) k_resize_words(Plan plan, Buffers buf, int level, int rowsPerThread)
{
    const LevelGeom& g = plan.lv[level];
    const LevelGeom& gp = plan.lv[level - 1];
    const int bx = 4 * (blockIdx.x * 32 + (threadIdx.x & 31));  // pitch is a multiple of 128: always inside the row
    const int frame = blockIdx.z + buf.frameBase;
    const int dy0 = (blockIdx.y * 4 + (threadIdx.x >> 5)) * rowsPerThread;
    if (dy0 >= g.h) return;
    const int dy1 = min(dy0 + rowsPerThread, g.h);
    int sx[4], sx1[4];
    uint32_t a[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        int dx = min(max(bx + j - PAD_LEFT, -EDGE), g.w - 1 + EDGE);
        dx = reflect101(dx, g.w);
        sx[j] = buf.xofs[g.xTab + dx];
        sx1[j] = min(sx[j] + 1, gp.w - 1);
        const short2 w = buf.xw[g.xTab + dx];
        a[j] = (uint32_t)(unsigned short)w.x | ((uint32_t)(unsigned short)w.y << 16);
    }
    const int base = min(min(sx[0], sx[1]), min(sx[2], sx[3]));
    uint32_t sel[4];
#pragma unroll
    for (int j = 0; j < 4; j++) sel[j] = (uint32_t)(sx[j] - base) | ((uint32_t)(sx1[j] - base) << 4);
    const uint32_t colOff = (uint32_t)(PAD_LEFT + base) & ~3u, sh = 8u * ((uint32_t)(PAD_LEFT + base) & 3u);
    const uint8_t* S = plane_ptr(buf.pyr, gp, frame) + (long long)PAD_TOP * gp.pitch;  // row 0 of the interior, buffer column 0
    uint8_t* D = plane_ptr(buf.pyr, g, frame);
    const uint32_t spitch = (uint32_t)gp.pitch, dpitch = (uint32_t)g.pitch;
    const uint8_t* Sb = S + colOff;
    uint8_t* Dp = D + ((uint32_t)(PAD_TOP + dy0) * dpitch + bx);
    const int4* yrow = buf.yrow + g.yTab;
    // horizontal pass of one source row (byte offset off): T >> 4 for the thread's 4 columns
    auto hrow = [&](uint32_t off, uint32_t* t) {
        const uint32_t* R = reinterpret_cast<const uint32_t*>(Sb + off);
        const uint32_t w0 = __ldg(R), w1 = __ldg(R + 1), w2 = __ldg(R + 2);
        const uint32_t u0 = __funnelshift_r(w0, w1, sh), u1 = __funnelshift_r(w1, w2, sh);
#pragma unroll
        for (int j = 0; j < 4; j++) t[j] = __dp2a_lo(a[j], __byte_perm(u0, u1, sel[j]), 0u) >> 4;
    };
    // Both source rows are interpolated for every output row (no row cache: the copies and branches it needs cost as
    // much as the second horizontal pass).  Vertical pass: p = b * (T >> 4) < 2^27; the two high halves are summed by
    // one IDP.2A with weights 64 and accumulator 2 * 64, which leaves ((p0 >> 16) + (p1 >> 16) + 2) >> 2 in byte 1.
    for (int dy = dy0; dy < dy1; dy++) {
        const int4 e = __ldg(yrow + dy);
        uint32_t t0[4], t1[4];
        hrow((uint32_t)e.x, t0);
        hrow((uint32_t)e.y, t1);
        const uint32_t b0 = (uint32_t)e.z & 0xffffu, b1 = (uint32_t)e.z >> 16;
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; j++) o[j] = __dp2a_lo(__byte_perm(b0 * t0[j], b1 * t1[j], 0x7632), 0x4040u, 128u);
        const uint32_t v = __byte_perm(__byte_perm(o[0], o[1], 0x0051), __byte_perm(o[2], o[3], 0x0051), 0x5410);
        *reinterpret_cast<uint32_t*>(Dp) = v;
        if (e.w) {  // this row is also a row of the REFLECT_101 frame (above and / or below the interior)
            const int up = (short)(e.w & 0xffff), down = e.w >> 16;
            if (up) *reinterpret_cast<uint32_t*>(Dp + (long long)up * (int)dpitch) = v;
            if (down) *reinterpret_cast<uint32_t*>(Dp + (long long)down * (int)dpitch) = v;
        }
        Dp += dpitch;
    }
}

// K1b, source-row walk (default for the pyramid's scale factors, 1 < f <= 2: LevelGeom::resizeRows): the same arithmetic as
// k_resize_words, but a thread walks the SOURCE rows its destination rows read and interpolates each of them horizontally
// ONCE -- at scale 1.2 five destination rows read six source rows, which k_resize_words interpolates ten times.  Source row r
// is the lower row (r1) of at most one destination row; the per-source-row table says which weights that row has (0 = none:
// warp-uniform branch) and where the REFLECT_101 frame repeats it.  The three source words of the next row are loaded before
// the current row is processed (two register sets alternating by name; addresses are sequential, one pitch apart).
template <bool DUP, bool SMEM>
__device__ __forceinline__ void resize_rows_walk(const uint8_t* __restrict__ S, uint8_t* __restrict__ D, const int2* __restrict__ srow,
                                                 uint32_t soff, uint32_t doff, int n, const uint32_t spitch, const uint32_t dpitch,
                                                 const uint32_t (&a)[4], const uint32_t (&sel)[4], const uint32_t sh)
{
    // S / D are CTA-uniform plane bases, soff / doff the thread's 32-bit byte offsets (uniform base + 32-bit offset addressing);
    // SMEM: S is the CTA's TMA-staged source tile in shared memory
    auto load3 = [&](uint32_t off, uint32_t* w) {
        const uint32_t* R = reinterpret_cast<const uint32_t*>(S + off);
        if (SMEM) { w[0] = R[0]; w[1] = R[1]; w[2] = R[2]; }
        else { w[0] = __ldg(R); w[1] = __ldg(R + 1); w[2] = __ldg(R + 2); }
    };
    auto hpass = [&](const uint32_t* w, uint32_t* t) {  // T >> 4 of the thread's 4 columns
        const uint32_t u0 = __funnelshift_r(w[0], w[1], sh), u1 = __funnelshift_r(w[1], w[2], sh);
#pragma unroll
        for (int j = 0; j < 4; j++) t[j] = __dp2a_lo(a[j], __byte_perm(u0, u1, sel[j]), 0u) >> 4;
    };
    auto emit = [&](const uint32_t* t0, const uint32_t* t1, int2 e) {  // vertical pass (see k_resize_words) + stores
        const uint32_t b0 = (uint32_t)e.x & 0xffffu, b1 = (uint32_t)e.x >> 16;
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; j++) o[j] = __dp2a_lo(__byte_perm(b0 * t0[j], b1 * t1[j], 0x7632), 0x4040u, 128u);
        const uint32_t v = __byte_perm(__byte_perm(o[0], o[1], 0x0051), __byte_perm(o[2], o[3], 0x0051), 0x5410);
        *reinterpret_cast<uint32_t*>(D + doff) = v;
        if (DUP && e.y) {  // this row is also a row of the REFLECT_101 frame (above and / or below the interior)
            const int up = (short)(e.y & 0xffff), down = e.y >> 16;
            if (up) *reinterpret_cast<uint32_t*>(D + (doff + (uint32_t)(up * (int)dpitch))) = v;
            if (down) *reinterpret_cast<uint32_t*>(D + (doff + (uint32_t)(down * (int)dpitch))) = v;
        }
        doff += dpitch;
    };
    // (one row below the last one is read ahead and never used: the first row of the lower frame at worst, inside the plane)
    uint32_t wa[3], wb[3], ta[4], tb[4];
    load3(soff, wa);
    load3(soff + spitch, wb);
    soff += 2 * spitch;
    hpass(wa, ta);
    while (n > 0) {  // n = source rows still to process (the first one is only the upper row of the first destination row)
        load3(soff, wa);
        hpass(wb, tb);
        {
            const int2 e = __ldg(srow + 1);
            if (e.x) emit(ta, tb, e);
        }
        if (n == 1) break;
        load3(soff + spitch, wb);
        hpass(wa, ta);
        {
            const int2 e = __ldg(srow + 2);
            if (e.x) emit(tb, ta, e);
        }
        soff += 2 * spitch;
        srow += 2;
        n -= 2;
    }
}

__global__ void __launch_bounds__(128) k_resize_rows(Plan plan, Buffers buf, int level, int rowsPerThread, int spitchArg, int dpitchArg)
{
    pdl_trigger();
    const LevelGeom& g = plan.lv[level];
    const LevelGeom& gp = plan.lv[level - 1];
    const int bx = 4 * (blockIdx.x * 32 + (threadIdx.x & 31));  // pitch is a multiple of 128: always inside the row
    const int frame = blockIdx.z + buf.frameBase;
    const int dy0 = (blockIdx.y * 4 + (threadIdx.x >> 5)) * rowsPerThread;
    if (dy0 >= g.h) return;
    const int dy1 = min(dy0 + rowsPerThread, g.h);
    int sx[4], sx1[4];
    uint32_t a[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        int dx = min(max(bx + j - PAD_LEFT, -EDGE), g.w - 1 + EDGE);
        dx = reflect101(dx, g.w);
        sx[j] = buf.xofs[g.xTab + dx];
        sx1[j] = min(sx[j] + 1, gp.w - 1);
        const short2 w = buf.xw[g.xTab + dx];
        a[j] = (uint32_t)(unsigned short)w.x | ((uint32_t)(unsigned short)w.y << 16);
    }
    const int base = min(min(sx[0], sx[1]), min(sx[2], sx[3]));
    uint32_t sel[4];
#pragma unroll
    for (int j = 0; j < 4; j++) sel[j] = (uint32_t)(sx[j] - base) | ((uint32_t)(sx1[j] - base) << 4);
    const uint32_t colOff = (uint32_t)(PAD_LEFT + base) & ~3u, sh = 8u * ((uint32_t)(PAD_LEFT + base) & 3u);
    const uint32_t spitch = (uint32_t)spitchArg, dpitch = (uint32_t)dpitchArg;  // = gp.pitch, g.pitch (kernel arguments: constant-bank operands)
    const int rFirst = buf.yofs0[g.yTab + dy0], rLast = buf.yofs1[g.yTab + dy1 - 1];  // source rows rFirst .. rLast
    const uint8_t* S = plane_ptr(buf.pyr, gp, frame);
    uint8_t* D = plane_ptr(buf.pyr, g, frame);
    const uint32_t soff = (uint32_t)(PAD_TOP + rFirst) * spitch + colOff, doff = (uint32_t)(PAD_TOP + dy0) * dpitch + bx;
    pdl_wait();
    const int2* srow = buf.srow + g.sTab + rFirst;
    // destination rows 1..19 and h-20..h-2 are repeated in the REFLECT_101 frame: only the threads that own such rows
    // (warp-uniform) run the walk with the extra stores
    const bool dup = dy0 <= EDGE || dy1 - 1 >= g.h - 1 - EDGE;
    if (dup) resize_rows_walk<true, false>(S, D, srow, soff, doff, rLast - rFirst, spitch, dpitch, a, sel, sh);
    else resize_rows_walk<false, false>(S, D, srow, soff, doff, rLast - rFirst, spitch, dpitch, a, sel, sh);
}

// K1b, source-row walk over a TMA-staged source tile (batches: rowsPerThread = RZ_RPT; LevelGeom::resizeTma says the tile of
// every CTA of the level fits): the CTA's source window -- RZ_TP bytes from the 16-byte grid column left of its leftmost
// source word x RZ_TROWS rows from the first source row of its first destination row -- arrives through ONE
// cp.async.bulk.tensor.3d; the walk then reads shared memory (k_resize_rows' top stall was global-load latency).
constexpr int RZ_RPT = 16;          // destination rows per thread
constexpr int RZ_TP = 192;          // tile pitch = box width: 128 destination columns x scale <= 1.25 + 16-byte grid + word slack
constexpr int RZ_TROWS = 84;        // 64 destination rows x scale <= 1.25 + 2, + the row the walk reads ahead
struct ResizeMaps { CUtensorMap m[FAST_TMA_LEVELS]; };  // indexed by SOURCE level

__global__ void __launch_bounds__(128) k_resize_rows_tma(Plan plan, Buffers buf, int level, int dpitchArg, const __grid_constant__ ResizeMaps maps)
{
    pdl_trigger();
    extern __shared__ __align__(128) unsigned char rsm[];
    __shared__ __align__(8) unsigned long long sBar;
    const LevelGeom& g = plan.lv[level];
    const LevelGeom& gp = plan.lv[level - 1];
    const int bx = 4 * (blockIdx.x * 32 + (threadIdx.x & 31));  // pitch is a multiple of 128: always inside the row
    const int frame = blockIdx.z + buf.frameBase;
    const int dyCta = blockIdx.y * 4 * RZ_RPT;                   // < g.h by the grid
    const int dy0 = dyCta + (threadIdx.x >> 5) * RZ_RPT;
    int sx[4], sx1[4];
    uint32_t a[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        int dx = min(max(bx + j - PAD_LEFT, -EDGE), g.w - 1 + EDGE);
        dx = reflect101(dx, g.w);
        sx[j] = buf.xofs[g.xTab + dx];
        sx1[j] = min(sx[j] + 1, gp.w - 1);
        const short2 w = buf.xw[g.xTab + dx];
        a[j] = (uint32_t)(unsigned short)w.x | ((uint32_t)(unsigned short)w.y << 16);
    }
    const int base = min(min(sx[0], sx[1]), min(sx[2], sx[3]));
    uint32_t sel[4];
#pragma unroll
    for (int j = 0; j < 4; j++) sel[j] = (uint32_t)(sx[j] - base) | ((uint32_t)(sx1[j] - base) << 4);
    const uint32_t colOff = (uint32_t)(PAD_LEFT + base) & ~3u, sh = 8u * ((uint32_t)(PAD_LEFT + base) & 3u);
    // every warp of the CTA covers the same columns: the warp-wide minimum is the CTA's leftmost source word
    const uint32_t tileX0 = __reduce_min_sync(0xffffffffu, colOff) & ~15u;
    const int rTile0 = buf.yofs0[g.yTab + dyCta];
    pdl_wait();
    if (threadIdx.x == 0) {
        const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&sBar);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)(RZ_TROWS * RZ_TP)) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                     ::"r"((uint32_t)__cvta_generic_to_shared(rsm)), "l"(&maps.m[level - 1]), "r"(bar), "r"((int)tileX0),
                       "r"(PAD_TOP + rTile0), "r"(frame)
                     : "memory");
    }
    __syncthreads();  // orders the mbarrier's initialisation before the waits
    {
        const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&sBar);
        uint32_t done;
        do {
            asm volatile(
                "{\n\t.reg .pred p;\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                "selp.u32 %0, 1, 0, p;\n\t}"
                : "=r"(done)
                : "r"(bar), "r"(0u)
                : "memory");
        } while (!done);
    }
    if (dy0 >= g.h) return;
    const int dy1 = min(dy0 + RZ_RPT, g.h);
    const uint32_t dpitch = (uint32_t)dpitchArg;
    const int rFirst = buf.yofs0[g.yTab + dy0], rLast = buf.yofs1[g.yTab + dy1 - 1];  // source rows rFirst .. rLast
    uint8_t* D = plane_ptr(buf.pyr, g, frame);
    const uint32_t soff = (uint32_t)(rFirst - rTile0) * RZ_TP + (colOff - tileX0), doff = (uint32_t)(PAD_TOP + dy0) * dpitch + bx;
    const int2* srow = buf.srow + g.sTab + rFirst;
    const bool dup = dy0 <= EDGE || dy1 - 1 >= g.h - 1 - EDGE;
    if (dup) resize_rows_walk<true, true>(rsm, D, srow, soff, doff, rLast - rFirst, RZ_TP, dpitch, a, sel, sh);
    else resize_rows_walk<false, true>(rsm, D, srow, soff, doff, rLast - rFirst, RZ_TP, dpitch, a, sel, sh);
}

// ------------------------------------------------------------------------------------------------------
// K2: FAST-9/16 per 30-px cell with the iniThFAST -> minThFAST retry.  One CTA per band = up to fastG horizontally
// adjacent cells of one cell row (<= 124 px wide), 128 threads, lane = one aligned 32-bit word (4 pixels) of a row.
//   score S(p) = max over the 16 contiguous 9-arcs of max(min d, min -d) - 1, corner at t <=> S >= t;
//   3x3 NMS inside each cell's evaluated region only (cv::FAST is run on the cell ROI);
//   cell output = NMS survivors with S >= ini if there is one, else survivors with S >= min.
// Passes (all through shared memory):
//   1. SWAR rejection on 4 pixels at a time, at both thresholds in one sweep: a 9-arc contains one pixel of every opposite
//      ring pair, so a corner at t needs |I(p)-I(q)| > t for q in {N or S} and q in {E or W}  (VABSDIFF4 + a carry trick);
//      the survivors of each threshold are kept as nibble masks in registers.
//   Then two phases, as the reference calls cv::FAST twice per cell: phase 0 at iniTh, phase 1 at minTh for the columns of
//   the cells that kept nothing at iniTh (NMS at iniTh is exact on its own: a kept pixel beats every neighbour below
//   iniTh anyway).  Per phase:
//   2. the phase's survivors are compacted into a queue and scored exactly: the 16 ring differences packed as 8 x biased
//      u16x2, sliding min/max over windows of 9 as min3(min3) with VIMNMX3.U16x2.  The queue holds FAST_Q1CAP pixels
//      (shared memory bounds residency); a band with more survivors than that is scored in rounds of FAST_GIT row sweeps
//      from the masks.
//   3. NMS for corners only (sparse; a dense scan of the score tile after a multi-round pass 2), per-cell record of
//      "kept a corner >= iniTh".
//   4. per-cell ini/min decision, unordered append to the level's candidate list.
// The reference's candidate order (cell row-major, then pixel row-major) is carried in `ord` and only matters for
// response ties inside a quad-tree node.
// ------------------------------------------------------------------------------------------------------
constexpr int FAST_T = 128;
constexpr int FAST_P = 160;   // tile pitch in bytes: <= 124 + 6 px band + 15 alignment bytes, multiple of 16
constexpr int FAST_P4 = FAST_P / 4;
constexpr int FAST_MAXG = 4;

constexpr int FAST_Q1CAP = 1280;  // pass-1 queue entries; bands with more survivors are scored in rounds of 2 row sweeps (<= 1024)
constexpr int FAST_GIT = 2;       // row sweeps (4 rows each, 4 px x 32 lanes per row) per round
static_assert(FAST_T == 128, "k_fast_band hard-codes 4 warps (queue quarters, sweeps of 4 rows)");

__device__ __forceinline__ uint32_t swap16(uint32_t v) { return __byte_perm(v, 0, 0x1032); }
// a | b | c and a & b & c as ONE LOP3 each (opaque to the optimiser, which otherwise re-associates the shared terms of the two
// thresholds of pass 1 into more instructions)
__device__ __forceinline__ uint32_t or3(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xFE;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t and3(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0x80;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// exact arc measure M(p) = max over 9-arcs of max(min d, min -d), d_k = I(p) - I(ring_k).  The differences are kept as
// biased u16x2 pairs (d_k + 0x8000, d_{k+8} + 0x8000): one 32-bit subtraction makes both (no borrow can cross the
// halves), and unsigned 16-bit min / max order the biased values like the signed differences.
__device__ __forceinline__ int fast_measure_packed(const uint8_t* p)
{
    constexpr int O[16] = {
        3 * FAST_P + 0, 3 * FAST_P + 1, 2 * FAST_P + 2, 1 * FAST_P + 3, 0 * FAST_P + 3, -1 * FAST_P + 3,
        -2 * FAST_P + 2, -3 * FAST_P + 1, -3 * FAST_P + 0, -3 * FAST_P - 1, -2 * FAST_P - 2, -1 * FAST_P - 3,
        0 * FAST_P - 3, 1 * FAST_P - 3, 2 * FAST_P - 2, 3 * FAST_P - 1};
    const uint32_t cc = (uint32_t)p[0] * 0x10001u + 0x80008000u;
    uint32_t dp[10];  // dp[k] = (d[k], d[k+8]) biased
#pragma unroll
    for (int k = 0; k < 8; k++) dp[k] = cc - ((uint32_t)p[O[k]] | ((uint32_t)p[O[k + 8]] << 16));
    dp[8] = swap16(dp[0]);
    dp[9] = swap16(dp[1]);
    uint32_t tn[14], tx[14];  // min / max over 3 consecutive differences, (k, k+8) packed
#pragma unroll
    for (int k = 0; k < 8; k++) {
        tn[k] = __vimin3_u16x2(dp[k], dp[k + 1], dp[k + 2]);
        tx[k] = __vimax3_u16x2(dp[k], dp[k + 1], dp[k + 2]);
    }
#pragma unroll
    for (int k = 0; k < 6; k++) { tn[8 + k] = swap16(tn[k]); tx[8 + k] = swap16(tx[k]); }
    uint32_t wn[8], wx[8];  // min / max over the window of 9 starting at k (and k + 8)
#pragma unroll
    for (int k = 0; k < 8; k++) {
        wn[k] = __vimin3_u16x2(tn[k], tn[k + 3], tn[k + 6]);
        wx[k] = __vimax3_u16x2(tx[k], tx[k + 3], tx[k + 6]);
    }
    const uint32_t a = __vimax3_u16x2(__vimax3_u16x2(wn[0], wn[1], wn[2]), __vimax3_u16x2(wn[3], wn[4], wn[5]), __vmaxu2(wn[6], wn[7]));
    const uint32_t b = __vimin3_u16x2(__vimin3_u16x2(wx[0], wx[1], wx[2]), __vimin3_u16x2(wx[3], wx[4], wx[5]), __vminu2(wx[6], wx[7]));
    const int A = (int)max(a & 0xffffu, a >> 16) - 0x8000;
    const int B = (int)min(b & 0xffffu, b >> 16) - 0x8000;
    return max(A, -B);
}

// Band staging, two variants of the same kernel (identical tile contents wherever the kernel reads them):
//   TMA = true   one elected thread issues ONE cp.async.bulk.tensor.3d (box FAST_P bytes x (hCell + 6) rows x 1 frame of the
//                level's [frame][row][byte] tensor map, completion on an mbarrier).  Default when nlevels <= FAST_TMA_LEVELS.
//   TMA = false  16-byte cp.async copies issued by every thread (16 lanes per row).  ORBX_FAST_TMA=0 selects it (and it
//                serves more than FAST_TMA_LEVELS levels).
// Both start at the 16-byte aligned column at or left of the band (the tile is shifted by the band's misalignment): cp.async
// needs it, and a TMA box whose innermost coordinate is off the 16-byte grid faults (tools/microbench/tma_probe.cu).
struct FastMaps { CUtensorMap m[FAST_TMA_LEVELS]; };

template <bool TMA>
__global__ void __launch_bounds__(FAST_T) k_fast_band(Plan plan, Buffers buf, const __grid_constant__ FastMaps maps)
{
    extern __shared__ __align__(128) unsigned char fsm[];
    __shared__ __align__(8) unsigned long long sBar;
    __shared__ int sQ1[2], sQ3, sBase;  // sQ1: one survivor counter per phase (no reset barrier between them)
    __shared__ __align__(4) unsigned char sHasIni[FAST_MAXG];  // byte c = cell c kept a corner >= iniTh (read as one word)
    __shared__ int sSeg[FAST_T / 32];
    static_assert(FAST_MAXG == 4, "sHasIni is read as one 32-bit word");

    pdl_trigger();
    const int frame = blockIdx.y + buf.frameBase;
    const int4 fuw = __ldg(reinterpret_cast<const int4*>(buf.fastUnits) + blockIdx.x);
    const int level = (short)(fuw.x & 0xffff), ci = fuw.x >> 16, j0 = (short)(fuw.y & 0xffff), iniX = fuw.y >> 16;
    const int iniY = (short)(fuw.z & 0xffff), rh = fuw.z >> 16, bw = (short)(fuw.w & 0xffff);
    const LevelGeom& g = plan.lv[level];
    const int evalW = bw - 6, eh = rh - 6;

    uint8_t* tile = fsm;
    uint8_t* sc = fsm + plan.fastTileBytes;
    unsigned short* q1 = reinterpret_cast<unsigned short*>(fsm + 2 * plan.fastTileBytes);  // FAST_Q1CAP entries
    unsigned short* q3 = q1 + FAST_Q1CAP;                                                   // plan.fastQCap entries

    // stage the band (TMA: one bulk tensor copy; otherwise 16-byte asynchronous copies, 16 lanes per row, 8 rows per pass),
    // zero the score tile meanwhile
    pdl_wait();  // the pyramid is complete (and the previous batch's readers of the candidate lists are done)
    const int shift = (PAD_LEFT + iniX) & 15;  // both variants copy from a 16-byte aligned column (TMA: a coordinate off that grid is an illegal instruction)
    if (TMA) {
        if (threadIdx.x == 0) {
            const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&sBar);
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)((g.hCell + 6) * FAST_P + plan.fastTileBytes)) : "memory");
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                         ::"r"((uint32_t)__cvta_generic_to_shared(tile)), "l"(&maps.m[level]), "r"(bar), "r"(PAD_LEFT + iniX - shift),
                           "r"(PAD_TOP + iniY), "r"(frame)
                         : "memory");
            // the score tile is zeroed by a second bulk copy (from a zero buffer that lives in L2) instead of by stores
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"((uint32_t)__cvta_generic_to_shared(sc)), "l"(buf.zeros), "r"((uint32_t)plan.fastTileBytes), "r"(bar)
                         : "memory");
        }
    } else {
        const int nw16 = (shift + bw + 15) >> 4;  // <= 10
        const int c16 = threadIdx.x & 15, r0 = threadIdx.x >> 4;
        if (c16 < nw16) {
            const uint8_t* src = interior_ptr(buf.pyr, g, frame) + (long long)(iniY + r0) * g.pitch + (iniX - shift) + 16 * c16;
            uint32_t dst = (uint32_t)__cvta_generic_to_shared(tile) + r0 * FAST_P + 16 * c16;
            const long long step = (long long)(FAST_T / 16) * g.pitch;
            for (int r = r0; r < rh; r += FAST_T / 16) {
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
                src += step;
                dst += (FAST_T / 16) * FAST_P;
            }
        }
    }
    if (!TMA) {
        uint4* z = reinterpret_cast<uint4*>(sc) + threadIdx.x;
        for (int n = rh * (FAST_P / 16) - threadIdx.x; n > 0; n -= FAST_T, z += FAST_T) *z = make_uint4(0, 0, 0, 0);
    }
    if (threadIdx.x == 0) { *reinterpret_cast<uint32_t*>(sHasIni) = 0u; sQ1[0] = 0; sQ1[1] = 0; sQ3 = 0; }
    if (!TMA) asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();
    if (TMA) {  // (the barrier above orders the mbarrier's initialisation before these waits)
        const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&sBar);
        uint32_t done;
        do {
            asm volatile(
                "{\n\t.reg .pred p;\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                "selp.u32 %0, 1, 0, p;\n\t}"
                : "=r"(done)
                : "r"(bar), "r"(0u)
                : "memory");
        } while (!done);
    }

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int minTh = plan.minTh, iniTh = plan.iniTh;
    const int wCell = g.wCell;
    const int txFirst = shift + 3, txEnd = shift + 3 + evalW;  // evaluated tile columns [txFirst, txEnd)
    const int gw0 = txFirst >> 2, gw = gw0 + lane;                      // ((txEnd-1)>>2) - (txFirst>>2) + 1 <= 32 words
    const int ncells = min(4, (int)(((uint32_t)(evalW - 1) * (uint32_t)g.wCellMagic) >> 16) + 1);  // = min(4, (evalW-1)/wCell + 1), no division
    constexpr int FAST_RPW = (59 + FAST_T / 32 - 1) / (FAST_T / 32);  // sweeps: eh <= 59
    static_assert(FAST_RPW <= 16, "survivor nibbles must fit 64 bits");
    // Survivor masks: sweep `it` (rows warp + 3 + 4 * it) owns nibble 7 - it of the low word (it < 8) or nibble 15 - it of the
    // high word; bit b of the nibble is pixel column 4 * gw + b.  Pass-1 queue entry = (expanding) thread << 6 | code, code =
    // bit index (+ 32 for the low word), so it = 15 - (code >> 2): one constant addition per entry on the divergent producer
    // side, decoded by the dense consumer (the column's lane = expanding lane - 8 * (code >> 4), see the quarter hand-over).
    const uint32_t e0 = (uint32_t)threadIdx.x << 6;
    auto push = [&](uint32_t m, uint32_t ebase, int base) {
        unsigned short* qp = q1 + base;
        while (m) {
            uint32_t bit;
            asm("bfind.u32 %0, %1;" : "=r"(bit) : "r"(m));  // position of the highest set bit (one FLO)
            m ^= 1u << bit;
            *qp++ = (unsigned short)(ebase + bit);
        }
    };
    // ---- pass 1: SWAR rejection, 4 pixels per lane, at both thresholds; the survivors of a lane's rows are kept as 4-bit
    //      groups of a 60-bit mask per threshold, shifted in from below (one funnel shift per sweep and threshold) ----
    uint32_t iniLo = 0, iniHi = 0, minLo = 0, minHi = 0;
    {
        const uint32_t Ki = (uint32_t)(0x7f - min(iniTh, 0x7f)) * 0x01010101u, Km = (uint32_t)(0x7f - min(minTh, 0x7f)) * 0x01010101u;
        // (a threshold >= 127 gives K = 0: the test degrades to |d| >= 128, still necessary for |d| > threshold)
        // bytes of this word that are evaluated pixels: columns [txFirst, txEnd) -> bytes [lo, hi) of the word
        const int vlo = min(max(txFirst - 4 * gw, 0), 4), vhi = min(max(txEnd - 4 * gw, 0), 4);
        const uint32_t vmask = 0x80808080u & __funnelshift_lc(0xffffffffu, 0u, 8 * vhi) & ~__funnelshift_lc(0xffffffffu, 0u, 8 * vlo);
        const uint32_t* rowp = reinterpret_cast<const uint32_t*>(tile) + (warp + 3) * FAST_P4 + gw;
        const int nIt = (eh - warp + FAST_T / 32 - 1) / (FAST_T / 32);  // sweeps of this warp (warp-uniform, <= FAST_RPW; <= 0: none)
#pragma unroll
        for (int it = 0; it < FAST_RPW; it++) {
            if (it >= nIt) break;  // warp-uniform
            const uint32_t* cp = rowp + it * (FAST_T / 32) * FAST_P4;
            const uint32_t C = cp[0], Lw = cp[-1], Rw = cp[1], N = cp[-3 * FAST_P4], S = cp[3 * FAST_P4];
            const uint32_t W4 = __byte_perm(Lw, C, 0x4321);  // pixels x-3 .. x
            const uint32_t E4 = __byte_perm(C, Rw, 0x6543);  // pixels x+3 .. x+6
            const uint32_t an = __vabsdiffu4(C, N), as = __vabsdiffu4(C, S), aw = __vabsdiffu4(C, W4), ae = __vabsdiffu4(C, E4);
            const uint32_t ln = an & 0x7f7f7f7fu, ls = as & 0x7f7f7f7fu, lw = aw & 0x7f7f7f7fu, le = ae & 0x7f7f7f7fu;
            uint32_t ns7, we7;  // bit 7 of a byte: one of the two differences is >= 128 (above any threshold < 128)
            asm("lop3.b32 %0, %1, %2, 0, 0xFC;" : "=r"(ns7) : "r"(an), "r"(as));
            asm("lop3.b32 %0, %1, %2, 0, 0xFC;" : "=r"(we7) : "r"(aw), "r"(ae));
            // bit 7 of each byte: |d| > threshold
            const uint32_t mi = and3(or3(ln + Ki, ls + Ki, ns7), or3(lw + Ki, le + Ki, we7), vmask);
            const uint32_t mm = and3(or3(ln + Km, ls + Km, ns7), or3(lw + Km, le + Km, we7), vmask);
            const uint32_t pi = mi * 0x00204081u, pm = mm * 0x00204081u;  // bits 7, 15, 23, 31 -> the top nibble
            if (it < 8) { iniLo = __funnelshift_l(pi, iniLo, 4); minLo = __funnelshift_l(pm, minLo, 4); }
            else { iniHi = __funnelshift_l(pi, iniHi, 4); minHi = __funnelshift_l(pm, minHi, 4); }
        }
        // canonical positions: sweep 0 at the top nibble of the low word, sweep 8 at the top nibble of the high word
        const int padLo = 4 * max(8 - max(nIt, 0), 0), padHi = 4 * min(16 - max(nIt, 8), 8);
        // (clamped funnel shift: x << pad for pad < 32, 0 for pad = 32)
        iniLo = __funnelshift_lc(0u, iniLo, padLo); minLo = __funnelshift_lc(0u, minLo, padLo);
        iniHi = __funnelshift_lc(0u, iniHi, padHi); minHi = __funnelshift_lc(0u, minHi, padHi);
    }
    // Two phases, as the reference calls cv::FAST twice: phase 0 scores the corners at iniTh (NMS among them is exact: a
    // kept pixel beats every neighbour below iniTh anyway); phase 1 repeats passes 2 and 3 at minTh for the cells that kept
    // nothing, restricted to their columns.  The pixels with minTh <= S < iniTh of textured cells (the majority of what
    // a minTh filter lets through) are never scored.
    uint32_t retry = 0;  // phase 1: bit c = cell c of the band kept nothing at iniTh
#pragma unroll 1
    for (int phase = 0; phase < 2; phase++) {
    const int th = phase ? minTh : iniTh;
    uint32_t mlo = iniLo, mhi = iniHi;
    if (phase) {
        // pixels of this lane's word that lie in a cell to retry: the word's first column is in cell c0, a cell border can
        // fall on at most one of its four pixels (wCell >= 30): the pixels from kb on belong to cell c0 + 1
        const int e0 = 4 * gw - txFirst;
        const int c0 = (int)(((uint32_t)max(e0, 0) * (uint32_t)g.wCellMagic) >> 16);
        const int kb = (c0 + 1) * wCell - e0;                       // >= 1
        const uint32_t low = (1u << min(kb, 4)) - 1u, r0 = retry >> c0;
        const uint32_t nibOK = ((r0 & 1u) ? low : 0u) | ((r0 & 2u) ? (low ^ 15u) : 0u);
        mlo = minLo & (nibOK * 0x11111111u);
        mhi = minHi & (nibOK * 0x11111111u);
    }
    // Survivors cluster (a vertical edge puts a survivor in every sweep of ONE lane), and the expansion loop below runs as long
    // as the busiest lane of the warp: the four 16-bit quarters of a lane's mask are handed to the lanes +0, +8, +16, +24, so a
    // column's survivors are expanded by four lanes.  Quarter q = code >> 4 of processing lane P came from lane P - 8 q.
    {
        const uint32_t h8 = __shfl_sync(0xffffffffu, mhi, (lane - 8) & 31);
        const uint32_t l16 = __shfl_sync(0xffffffffu, mlo, (lane - 16) & 31);
        const uint32_t l24 = __shfl_sync(0xffffffffu, mlo, (lane - 24) & 31);
        mhi = (mhi & 0x0000ffffu) | (h8 & 0xffff0000u);
        mlo = (l16 & 0x0000ffffu) | (l24 & 0xffff0000u);
    }
    {
        const int cnt = __popc(mlo) + __popc(mhi);
        if (cnt) {
            const int base = atomicAdd(&sQ1[phase], cnt);
            if (base + cnt <= FAST_Q1CAP) {
                push(mlo, e0 + 32u, base);
                push(mhi, e0, base + __popc(mlo));
            }
        }
    }
    __syncthreads();
    // ---- pass 2: exact score of the queued pixels.  Each warp owns a contiguous quarter of the queue and compacts the
    //      entries that turn out to be corners in place at the head of its quarter (writes trail the reads) ----
    const int n1 = sQ1[phase];
    const bool sparse = n1 <= FAST_Q1CAP;  // CTA-uniform
    const int per = (((n1 + FAST_T / 32 - 1) / (FAST_T / 32)) + 31) & ~31;
    if (sparse) {
        const int wbase = warp * per, wend = min(wbase + per, n1);
        int cnt = 0;
        for (int q0 = wbase; q0 < wend; q0 += 32) {
            const int q = q0 + lane;
            const int e = q < wend ? q1[q] : 0;
            const int ty = (e >> 11) + 63 - 4 * ((e >> 2) & 15), tx = 4 * (gw0 + ((((e >> 6) & 31) - ((e >> 1) & 24)) & 31)) + (e & 3);
            const int idx = ty * FAST_P + tx;
            int mval = 0;
            if (q < wend) mval = fast_measure_packed(tile + idx);
            const bool corner = mval > th;
            if (corner) sc[idx] = (uint8_t)(mval - 1);
            const uint32_t bal = __ballot_sync(0xffffffffu, corner);
            if (corner) q1[wbase + cnt + __popc(bal & ((1u << lane) - 1u))] = (unsigned short)((ty << 8) | tx);
            cnt += __popc(bal);
        }
        if (lane == 0) sSeg[warp] = cnt;
    } else {
        // too many survivors for the queue (texture / noise everywhere): FAST_GIT sweeps per round
        for (int it0 = 0; it0 * (FAST_T / 32) < eh; it0 += FAST_GIT) {
            __syncthreads();
            if (threadIdx.x == 0) sQ1[phase] = 0;
            __syncthreads();
            static_assert(FAST_GIT == 2, "a round = the two nibbles of sweeps it0 (even), it0 + 1");
            const uint32_t m = it0 < 8 ? mlo & (0xffu << (4 * (6 - it0))) : mhi & (0xffu << (4 * (14 - it0)));
            if (m) push(m, e0 + (it0 < 8 ? 32u : 0u), atomicAdd(&sQ1[phase], __popc(m)));
            __syncthreads();
            const int n = sQ1[phase];
            for (int q = threadIdx.x; q < n; q += FAST_T) {
                const int e = q1[q];
                const int idx = ((e >> 11) + 63 - 4 * ((e >> 2) & 15)) * FAST_P + 4 * (gw0 + ((((e >> 6) & 31) - ((e >> 1) & 24)) & 31)) + (e & 3);
                const int mval = fast_measure_packed(tile + idx);
                if (mval > th) sc[idx] = (uint8_t)(mval - 1);
            }
        }
    }
    __syncthreads();
    // ---- pass 3: 3x3 NMS inside the cell, for corners only ----
    auto nms = [&](int ty, int tx) {
        const int ex = tx - (shift + 3);
        const int c = (ex >= wCell) + (ex >= 2 * wCell) + (ex >= 3 * wCell);
        if (phase && !((retry >> c) & 1u)) return;  // (dense scan only) this cell was finished in phase 0
        const int r = ex - c * wCell;
        const uint8_t* s0 = sc + ty * FAST_P + tx;
        const int s = s0[0];
        const bool hasL = r > 0, hasR = r < wCell - 1;  // neighbours in another cell's region count as 0
        bool keep = s > s0[-FAST_P] && s > s0[FAST_P];
        if (hasL) keep = keep && s > s0[-1] && s > s0[-FAST_P - 1] && s > s0[FAST_P - 1];
        if (hasR) keep = keep && s > s0[1] && s > s0[-FAST_P + 1] && s > s0[FAST_P + 1];
        if (keep) {
            const int p3 = atomicAdd(&sQ3, 1);
            if (p3 < plan.fastQCap) q3[p3] = (unsigned short)((ty << 8) | tx);
            else *reinterpret_cast<volatile int*>(buf.errFlag) = 1;  // cannot happen by the arithmetic of fastQCap; never write past the queue
            if (s >= iniTh) sHasIni[c] = 1;
        }
    };
    if (sparse) {
        const int c0 = sSeg[0], c1 = c0 + sSeg[1], c2 = c1 + sSeg[2], c3 = c2 + sSeg[3];
        for (int t = threadIdx.x; t < c3; t += FAST_T) {
            const int w = (t >= c0) + (t >= c1) + (t >= c2);
            const int e = q1[w * per + t - (w == 0 ? 0 : (w == 1 ? c0 : (w == 2 ? c1 : c2)))];
            nms(e >> 8, e & 0xff);
        }
    } else {
        const uint32_t* sc32 = reinterpret_cast<const uint32_t*>(sc);
        for (int idx = threadIdx.x; idx < eh * FAST_P4; idx += FAST_T) {
            const int ty = idx / FAST_P4 + 3, wx = idx - (ty - 3) * FAST_P4;
            uint32_t w = sc32[ty * FAST_P4 + wx];
            for (int bb = 0; w; bb++, w >>= 8)
                if (w & 0xffu) nms(ty, 4 * wx + bb);
        }
    }
    __syncthreads();
    if (phase || minTh >= iniTh) break;  // (a retry at minTh >= iniTh can only find a subset of nothing)
    // bytes 0/1 of the cells -> bits 0..3 (no two partial products share a bit below bit 28)
    retry = ~((*reinterpret_cast<const uint32_t*>(sHasIni) * 0x01020408u) >> 24) & ((1u << ncells) - 1u);
    if (!retry) break;  // CTA-uniform.  No barrier here: phase 1 writes q1 / sSeg / sHasIni only after its own barriers,
                        // which every thread reaches after it has read them above
    }
    // ---- pass 4: reserve, write.  Every NMS survivor is an output: phase 0 keeps only corners >= iniTh (their cell then
    //      "has an ini corner"), phase 1 runs only in the cells that kept nothing and keeps corners >= minTh -- the
    //      per-cell threshold of the reference's two cv::FAST calls is already applied by which phase scored the pixel ----
    const int n3 = min(sQ3, plan.fastQCap);
    if (threadIdx.x == 0) sBase = n3 ? atomicAdd(buf.candCount + frame * plan.nlevels + level, n3) : 0;
    __syncthreads();
    Cand* out = buf.cand + (long long)frame * plan.candFrame + g.candOff;
    for (int q = threadIdx.x; q < n3; q += FAST_T) {
        const int e = q3[q];
        const int ty = e >> 8, tx = e & 0xff;
        const int ex = tx - (shift + 3), ey = ty - 3;
        const int c = (ex >= wCell) + (ex >= 2 * wCell) + (ex >= 3 * wCell);
        const int s = sc[ty * FAST_P + tx];
        const int pos = sBase + q;
        if (pos >= g.candCap) { *reinterpret_cast<volatile int*>(buf.errFlag) = 1; break; }
        const int r = ex - c * wCell;
        const uint32_t xrel = (j0 + c) * wCell + r + 3, yrel = ci * g.hCell + ey + 3;
        const uint32_t cell = (uint32_t)ci * g.nCols + j0 + c;
        const uint32_t ord = (cell * g.hCell + ey) * wCell + r;
        Cand cd;
        cd.xy = xrel | (yrel << 16);
        cd.ro = ((uint32_t)s << 24) | (0xFFFFFFu - ord);
        out[pos] = cd;
    }
}

// ------------------------------------------------------------------------------------------------------
// K3: quad-tree distribution.  One CTA per (level, frame).
//
// The reference keeps a std::list of nodes and splits with push_front / erase; that list is reproduced as an
// array in list order, rebuilt once per pass with prefix sums:
//   phase 1 (expand every node with > 1 key):   L' = reverse(children in creation order) ++ [1-key nodes]
//   phase 2 (expand by size until >= N nodes):  candidates sorted by (size desc, creation index desc), all split
//            in parallel, the sequential "break when size >= N" becomes a prefix-sum cut; L' = reverse(children of
//            the processed candidates) ++ [old list without those candidates].
// Keys never move: each key carries the list position of its node, so node sizes are atomics and the final
// "max response, first in candidate order" pick is one atomicMax on (response << 24 | ~ord).
// ------------------------------------------------------------------------------------------------------
// Threads per (level, frame) CTA = template parameter OCT_T.  The kernel is barrier-latency bound: for batches, many small
// CTAs per SM (128 threads, 8 resident) give the best throughput (0.145 -> 0.092 ms per 256 frames against two CTAs of 512);
// a single frame is fastest with 512 threads per level.

struct OctSmem {
    short4* geoA; short4* geoB;       // node rectangles (ulx, uly, brx, bry), list order, double buffered
    int* cntA; int* cntB;             // keys per node
    int* childCnt;                    // 4 per split slot
    unsigned short* childPos;         // new list position of each child (0xFFFF = empty)
    unsigned short* nodePos;          // new list position of surviving nodes
    int* scanA; int* scanB;           // scratch for prefix sums
    unsigned short* V; unsigned short* Vn;  // phase-2 candidates (list positions), creation order
    short* slotOf;                    // list position -> split slot or -1
    unsigned short* procOfSlot;       // slot -> rank in processing order
    unsigned short* slotOfProc;       // processing rank -> slot
    unsigned int* best;               // per node max(ro)
    int* winner;                      // per node winning key index
    unsigned short* keyNode;          // per key: list position of its node (smem when it fits)
};

__device__ __forceinline__ int warp_incl_scan(int v)
{
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) >= o) v += t;
    }
    return v;
}

// in-place exclusive scan of data[0..n); returns the total.  All threads of the CTA must call it.
template <int OCT_T>
__device__ int block_exscan(int* data, int n, int* warpSums)
{
    const int ipt = (n + OCT_T - 1) / OCT_T;
    const int b = threadIdx.x * ipt, e = min(b + ipt, n);
    int sum = 0;
    for (int i = b; i < e; i++) sum += data[i];
    const int incl = warp_incl_scan(sum);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 31) warpSums[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int w = lane < OCT_T / 32 ? warpSums[lane] : 0;
        int wi = warp_incl_scan(w);
        warpSums[lane] = wi - w;
        if (lane == 31) warpSums[32] = wi;
    }
    __syncthreads();
    int run = warpSums[wid] + incl - sum;
    for (int i = b; i < e; i++) { int v = data[i]; data[i] = run; run += v; }
    const int total = warpSums[32];
    __syncthreads();
    return total;
}

__device__ __forceinline__ int child_of(short4 r, uint32_t xy, int& mx, int& my)
{
    mx = r.x + ((r.z - r.x + 1) >> 1);  // UL.x + ceil((UR.x-UL.x)/2.f)
    my = r.y + ((r.w - r.y + 1) >> 1);
    const int x = xy & 0xffff, y = xy >> 16;
    return (x < mx ? 0 : 1) + (y < my ? 0 : 2);  // n1 UL, n2 UR, n3 BL, n4 BR
}
__device__ __forceinline__ short4 child_rect(short4 r, int c, int mx, int my)
{
    short4 o;
    o.x = (c & 1) ? mx : r.x;
    o.z = (c & 1) ? r.z : mx;
    o.y = (c & 2) ? my : r.y;
    o.w = (c & 2) ? r.w : my;
    return o;
}

// cv::fastAtan2, scalar float path, no contraction
__device__ __forceinline__ float fast_atan2_dev(float y, float x)
{
    const float p1 = 57.2836266f, p3 = -18.6674461f, p5 = 8.91400051f, p7 = -2.53972459f;
    const float ax = fabsf(x), ay = fabsf(y);
    float a;
    if (ax >= ay) {
        const float c = __fdiv_rn(ay, __fadd_rn(ax, 2.22044605e-16f));
        const float c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        const float c = __fdiv_rn(ax, __fadd_rn(ay, 2.22044605e-16f));
        const float c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

// circular patch half-widths umax[|v|] (src/ORBextractor.cc:519-549)
__device__ __forceinline__ constexpr int umax_of(int av)
{
    // {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3} as nibbles (no local-memory table)
    return (int)((0x3689ABCDDEEEFFFFull >> (4 * av)) & 15ull);
}

template <int OCT_T>
__global__ void __launch_bounds__(OCT_T) k_octree(Plan plan, Buffers buf, int nodeCapMax, int keySmemCap,
                                                  unsigned short* keyNodeGlobal, long long keyNodeFrameStride)
{
    extern __shared__ __align__(16) unsigned char smemRaw[];
    __shared__ int warpSums[33];
    __shared__ int sN, sNV, sFinish, sCut, sCtot;

    pdl_trigger();
    const int level = blockIdx.x, frame = blockIdx.y + buf.frameBase;
    const LevelGeom& g = plan.lv[level];
    int* counter = buf.candCount + frame * plan.nlevels + level;
    pdl_wait();
    const int M = min(*counter, g.candCap);
    const Cand* __restrict__ cand = buf.cand + (long long)frame * plan.candFrame + g.candOff;
    StagedKp* staged = buf.staged + (long long)frame * plan.kpCapFrame + g.kpOff;
    __syncthreads();
    if (threadIdx.x == 0) {
        buf.candCountDbg[frame * plan.nlevels + level] = M;
        *counter = 0;  // ready for the next batch
    }
    if (M == 0) {
        if (threadIdx.x == 0) buf.levelCount[frame * plan.nlevels + level] = 0;
        return;
    }

    // carve shared memory
    OctSmem S;
    {
        unsigned char* p = smemRaw;
        const int cap = nodeCapMax;
        auto take = [&](size_t bytes) { unsigned char* r = p; p += (bytes + 15) & ~size_t(15); return r; };
        S.geoA = (short4*)take(sizeof(short4) * cap);
        S.geoB = (short4*)take(sizeof(short4) * cap);
        S.cntA = (int*)take(4 * cap);
        S.cntB = (int*)take(4 * cap);
        S.childCnt = (int*)take(16 * cap);
        S.scanA = (int*)take(4 * cap);
        S.scanB = (int*)take(4 * cap);
        S.best = (unsigned int*)take(4 * cap);
        S.winner = (int*)take(4 * cap);
        S.childPos = (unsigned short*)take(8 * cap);
        S.nodePos = (unsigned short*)take(2 * cap);
        S.V = (unsigned short*)take(2 * cap);
        S.Vn = (unsigned short*)take(2 * cap);
        S.slotOf = (short*)take(2 * cap);
        S.procOfSlot = (unsigned short*)take(2 * cap);
        S.slotOfProc = (unsigned short*)take(2 * cap);
        S.keyNode = (M <= keySmemCap) ? (unsigned short*)take(2 * (size_t)keySmemCap)
                                      : keyNodeGlobal + (long long)frame * keyNodeFrameStride + g.candOff;
    }
    const int N = g.quota;
    const int tid = threadIdx.x;

    // ---- roots (src/ORBextractor.cc:673-748) ----
    const int nIni = g.nIni;
    const float hX = g.hX;
    for (int i = tid; i < nIni; i += OCT_T) {
        short4 r;
        r.x = (short)(int)__fmul_rn(hX, (float)i);
        r.z = (short)(int)__fmul_rn(hX, (float)(i + 1));
        r.y = 0;
        r.w = (short)(g.maxBY - 16);
        S.geoB[i] = r;
        S.cntB[i] = 0;
    }
    __syncthreads();
    for (int k = tid; k < M; k += OCT_T) {
        const int x = cand[k].xy & 0xffff;
        const int root = (int)__fdiv_rn((float)x, hX);
        S.keyNode[k] = (unsigned short)root;
        atomicAdd(&S.cntB[root], 1);
    }
    __syncthreads();
    // drop empty roots, keep order
    for (int i = tid; i < nIni; i += OCT_T) S.scanA[i] = S.cntB[i] > 0;
    __syncthreads();
    int n = block_exscan<OCT_T>(S.scanA, nIni, warpSums);
    for (int i = tid; i < nIni; i += OCT_T)
        if (S.cntB[i] > 0) { S.geoA[S.scanA[i]] = S.geoB[i]; S.cntA[S.scanA[i]] = S.cntB[i]; S.nodePos[i] = (unsigned short)S.scanA[i]; }
    __syncthreads();
    for (int k = tid; k < M; k += OCT_T) S.keyNode[k] = S.nodePos[S.keyNode[k]];
    __syncthreads();

    short4* geo = S.geoA; short4* geoN = S.geoB;
    int* cnt = S.cntA; int* cntN = S.cntB;
    int nV = 0;
    bool finish = false;

    while (!finish) {
        // ================= phase 1 pass: split every node holding more than one key =================
        const int prevSize = n;
        for (int i = tid; i < 4 * n; i += OCT_T) S.childCnt[i] = 0;
        __syncthreads();
        for (int k = tid; k < M; k += OCT_T) {
            const int i = S.keyNode[k];
            if (cnt[i] > 1) {
                int mx, my;
                const int c = child_of(geo[i], cand[k].xy, mx, my);
                atomicAdd(&S.childCnt[4 * i + c], 1);
            }
        }
        __syncthreads();
        for (int i = tid; i < n; i += OCT_T) {
            int nch = 0, nex = 0;
            if (cnt[i] > 1) {
#pragma unroll
                for (int c = 0; c < 4; c++) { const int v = S.childCnt[4 * i + c]; nch += v > 0; nex += v > 1; }
            }
            S.scanA[i] = nch | (nex << 16);
            S.scanB[i] = cnt[i] > 1 ? 0 : 1;
        }
        __syncthreads();
        const int totA = block_exscan<OCT_T>(S.scanA, n, warpSums);
        const int totB = block_exscan<OCT_T>(S.scanB, n, warpSums);
        const int C = totA & 0xffff, nToExpand = totA >> 16;
        const int nNew = C + totB;
        // (capacity: nNew <= max(N, 4*nIni) <= nodeCap by construction)
        for (int i = tid; i < n; i += OCT_T) {
            if (cnt[i] > 1) {
                int ci = S.scanA[i] & 0xffff, vi = S.scanA[i] >> 16;
                const short4 r = geo[i];
                const int mx = r.x + ((r.z - r.x + 1) >> 1), my = r.y + ((r.w - r.y + 1) >> 1);
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const int v = S.childCnt[4 * i + c];
                    if (v > 0) {
                        const int pos = C - 1 - ci;
                        geoN[pos] = child_rect(r, c, mx, my);
                        cntN[pos] = v;
                        S.childPos[4 * i + c] = (unsigned short)pos;
                        if (v > 1) S.V[vi++] = (unsigned short)pos;
                        ci++;
                    }
                }
            } else {
                const int pos = C + S.scanB[i];
                geoN[pos] = geo[i];
                cntN[pos] = cnt[i];
                S.nodePos[i] = (unsigned short)pos;
            }
        }
        __syncthreads();
        for (int k = tid; k < M; k += OCT_T) {
            const int i = S.keyNode[k];
            if (cnt[i] > 1) {
                int mx, my;
                const int c = child_of(geo[i], cand[k].xy, mx, my);
                S.keyNode[k] = S.childPos[4 * i + c];
            } else {
                S.keyNode[k] = S.nodePos[i];
            }
        }
        __syncthreads();
        { short4* t = geo; geo = geoN; geoN = t; int* u = cnt; cnt = cntN; cntN = u; }
        n = nNew;
        nV = nToExpand;
        if (n >= N || n == prevSize) {
            finish = true;
        } else if (n + nToExpand * 3 > N) {
            // ================= phase 2: split by size, largest first, until the list reaches N =================
            unsigned short* V = S.V; unsigned short* Vn = S.Vn;
            while (!finish) {
                if (nV == 0) { finish = true; break; }  // nothing left to split: size unchanged (src/ORBextractor.cc:967)
                const int prev2 = n;
                // processing order: stable ascending sort on size, walked from the back
                //   = descending size, ties by descending creation index (decision D1)
                for (int s = tid; s < nV; s += OCT_T) {
                    const int cs = cnt[V[s]];
                    int rank = 0;
                    for (int t = 0; t < nV; t++) {
                        const int ct = cnt[V[t]];
                        rank += (ct > cs) || (ct == cs && t > s);
                    }
                    S.procOfSlot[s] = (unsigned short)rank;
                    S.slotOfProc[rank] = (unsigned short)s;
                }
                for (int i = tid; i < n; i += OCT_T) S.slotOf[i] = -1;
                for (int i = tid; i < 4 * nV; i += OCT_T) S.childCnt[i] = 0;
                __syncthreads();
                for (int s = tid; s < nV; s += OCT_T) S.slotOf[V[s]] = (short)s;
                __syncthreads();
                for (int k = tid; k < M; k += OCT_T) {
                    const int i = S.keyNode[k];
                    const int s = S.slotOf[i];
                    if (s >= 0) {
                        int mx, my;
                        const int c = child_of(geo[i], cand[k].xy, mx, my);
                        atomicAdd(&S.childCnt[4 * s + c], 1);
                    }
                }
                __syncthreads();
                // children per candidate in processing order -> prefix -> cut where the list reaches N
                for (int p = tid; p < nV; p += OCT_T) {
                    const int s = S.slotOfProc[p];
                    int nch = 0;
#pragma unroll
                    for (int c = 0; c < 4; c++) nch += S.childCnt[4 * s + c] > 0;
                    S.scanA[p] = nch;
                }
                if (tid == 0) sCut = nV - 1;
                __syncthreads();
                block_exscan<OCT_T>(S.scanA, nV, warpSums);
                // inclusive gain after processing p: (excl[p] + nch_p) - (p + 1)
                for (int p = tid; p < nV; p += OCT_T) {
                    const int s = S.slotOfProc[p];
                    int nch = 0;
#pragma unroll
                    for (int c = 0; c < 4; c++) nch += S.childCnt[4 * s + c] > 0;
                    if (n + S.scanA[p] + nch - (p + 1) >= N) atomicMin(&sCut, p);
                }
                __syncthreads();
                const int cut = sCut;  // last processed rank
                if (tid == 0) {
                    const int s = S.slotOfProc[cut];
                    int nch = 0;
                    for (int c = 0; c < 4; c++) nch += S.childCnt[4 * s + c] > 0;
                    sCtot = S.scanA[cut] + nch;
                }
                // surviving old nodes keep their order behind the new children
                for (int i = tid; i < n; i += OCT_T) {
                    const int s = S.slotOf[i];
                    S.scanB[i] = (s >= 0 && S.procOfSlot[s] <= cut) ? 0 : 1;
                }
                __syncthreads();
                const int Ctot = sCtot;
                const int keepOld = block_exscan<OCT_T>(S.scanB, n, warpSums);
                const int nNew2 = Ctot + keepOld;
                // new expandable list (creation order) needs its own prefix over processing ranks
                for (int p = tid; p < nV; p += OCT_T) {
                    const int s = S.slotOfProc[p];
                    int nex = 0;
                    if (p <= cut) {
#pragma unroll
                        for (int c = 0; c < 4; c++) nex += S.childCnt[4 * s + c] > 1;
                    }
                    S.winner[p] = nex;  // scratch
                }
                __syncthreads();
                const int nVnew = block_exscan<OCT_T>(S.winner, nV, warpSums);
                for (int p = tid; p <= cut; p += OCT_T) {
                    const int s = S.slotOfProc[p];
                    const int i = V[s];
                    const short4 r = geo[i];
                    const int mx = r.x + ((r.z - r.x + 1) >> 1), my = r.y + ((r.w - r.y + 1) >> 1);
                    int ci = S.scanA[p], vi = S.winner[p];
#pragma unroll
                    for (int c = 0; c < 4; c++) {
                        const int v = S.childCnt[4 * s + c];
                        if (v > 0) {
                            const int pos = Ctot - 1 - ci;
                            geoN[pos] = child_rect(r, c, mx, my);
                            cntN[pos] = v;
                            S.childPos[4 * s + c] = (unsigned short)pos;
                            if (v > 1) Vn[vi++] = (unsigned short)pos;
                            ci++;
                        }
                    }
                }
                for (int i = tid; i < n; i += OCT_T) {
                    const int s = S.slotOf[i];
                    if (!(s >= 0 && S.procOfSlot[s] <= cut)) {
                        const int pos = Ctot + S.scanB[i];
                        geoN[pos] = geo[i];
                        cntN[pos] = cnt[i];
                        S.nodePos[i] = (unsigned short)pos;
                    }
                }
                __syncthreads();
                for (int k = tid; k < M; k += OCT_T) {
                    const int i = S.keyNode[k];
                    const int s = S.slotOf[i];
                    if (s >= 0 && S.procOfSlot[s] <= cut) {
                        int mx, my;
                        const int c = child_of(geo[i], cand[k].xy, mx, my);
                        S.keyNode[k] = S.childPos[4 * s + c];
                    } else {
                        S.keyNode[k] = S.nodePos[i];
                    }
                }
                __syncthreads();
                { short4* t = geo; geo = geoN; geoN = t; int* u = cnt; cnt = cntN; cntN = u; }
                { unsigned short* t = V; V = Vn; Vn = t; }
                n = nNew2;
                nV = nVnew;
                if (n >= N || n == prev2) finish = true;
            }
        }
    }

    // ---- one keypoint per node: max response, first in candidate order on ties (src/ORBextractor.cc:984-1009) ----
    for (int i = tid; i < n; i += OCT_T) S.best[i] = 0;
    __syncthreads();
    for (int k = tid; k < M; k += OCT_T) atomicMax(&S.best[S.keyNode[k]], cand[k].ro);
    __syncthreads();
    for (int k = tid; k < M; k += OCT_T) {
        const int i = S.keyNode[k];
        if (cand[k].ro == S.best[i]) S.winner[i] = k;
    }
    __syncthreads();
    if (tid == 0) buf.levelCount[frame * plan.nlevels + level] = n;

    // ---- stage the picked keypoints in list order (orientation + descriptor follow in k_describe) ----
    for (int i = tid; i < n; i += OCT_T) {
        const Cand c = cand[S.winner[i]];
        StagedKp kp;
        kp.x = (float)((int)(c.xy & 0xffff) + 16);
        kp.y = (float)((int)(c.xy >> 16) + 16);
        kp.angle = -1.f;
        kp.response = (float)(c.ro >> 24);
        staged[i] = kp;
    }
}

// ------------------------------------------------------------------------------------------------------
// K5: 7x7 Gaussian blur, sigma 2, OpenCV 4.x fixed point: taps {18,34,48,56,48,34,18}/256, 16-bit horizontal sums,
// 32-bit vertical sums, dst = (v + 32768) >> 16.  Reads the padded plane (its 19-px REFLECT_101 frame makes the
// border handling implicit) and writes the interior of the blurred plane.  All levels in one launch.
// A thread owns 4 adjacent columns and slides down BLUR_ROWS rows, two rows per step:
//   horizontal pass on bytes with IDP.4A: h = dp4a(p[x-3..x], {18,34,48,56}) + dp4a(p[x+1..x+4], {48,34,18,0})
//   vertical pass on row pairs packed as u16x2 with IDP.2A (two taps per instruction, 32-bit accumulate); the
//   rounding constant is the accumulator's initial value.
// ------------------------------------------------------------------------------------------------------
constexpr int BLUR_ROWS = 32;  // output rows per thread (even; 64 measured no faster: the kernel is latency-, not issue-bound)

// the three words (columns x-4 .. x+7) a thread needs from one source row; row points at the thread's first column
__device__ __forceinline__ void blur_ldrow(const uint8_t* row, uint32_t (&n)[3])
{
    n[0] = __ldg(reinterpret_cast<const uint32_t*>(row) - 1);
    n[1] = __ldg(reinterpret_cast<const uint32_t*>(row));
    n[2] = __ldg(reinterpret_cast<const uint32_t*>(row) + 1);
}

__device__ __forceinline__ void blur_hrow(const uint32_t (&n)[3], uint32_t h[4])
{
    const uint32_t L = n[0], C = n[1], R = n[2];
    constexpr uint32_t TLO = 0x38302212u;  // taps 18,34,48,56 for bytes x-3..x
    constexpr uint32_t THI = 0x00122230u;  // taps 48,34,18,0  for bytes x+1..x+4
    h[0] = __dp4a(__funnelshift_r(C, R, 8), THI, __dp4a(__funnelshift_r(L, C, 8), TLO, 0u));
    h[1] = __dp4a(__funnelshift_r(C, R, 16), THI, __dp4a(__funnelshift_r(L, C, 16), TLO, 0u));
    h[2] = __dp4a(__funnelshift_r(C, R, 24), THI, __dp4a(__funnelshift_r(L, C, 24), TLO, 0u));
    h[3] = __dp4a(R, THI, __dp4a(C, TLO, 0u));
}

__global__ void __launch_bounds__(128) k_blur(Plan plan, Buffers buf)
{
    pdl_trigger();
    pdl_wait();
    int level = 0;
    while (level + 1 < plan.nlevels && (int)blockIdx.x >= plan.lv[level + 1].blurOff) level++;
    const LevelGeom& g = plan.lv[level];
    const int frame = blockIdx.y + buf.frameBase;
    const int unit = blockIdx.x - g.blurOff;
    const int cgroups = (g.w + 127) >> 7;  // CTAs across (32 threads x 4 columns)
    const int cy = unit / cgroups, cxg = unit - cy * cgroups;
    const int x = 4 * (cxg * 32 + (threadIdx.x & 31));
    const int y0 = (cy * 4 + (threadIdx.x >> 5)) * BLUR_ROWS;
    if (x >= g.w || y0 >= g.h) return;
    const int y1 = min(y0 + BLUR_ROWS, g.h);
    const uint8_t* src = interior_ptr(buf.pyr, g, frame) + x;
    uint8_t* dst = const_cast<uint8_t*>(interior_ptr(buf.blur, g, frame)) + x;
    const long long pitch = g.pitch;
    // window of 8 rows (y-3 .. y+4) as 4 vertical pairs per column: P[k][j] = h(y-3+2k, j) | h(y-2+2k, j) << 16
    uint32_t P[4][4];
    // the source words of a row pair are loaded two steps before they are used (buffers A and B alternate), so that their
    // latency hides behind the arithmetic of the steps in between
    uint32_t A0[3], A1[3], B0[3], B1[3];
    auto ldpair = [&](const uint8_t* row, uint32_t (&n0)[3], uint32_t (&n1)[3]) {  // rows `row` and `row + pitch`
        blur_ldrow(row, n0);
        blur_ldrow(row + pitch, n1);
    };
    auto hpair = [&](const uint32_t (&n0)[3], const uint32_t (&n1)[3], uint32_t (&Q)[4]) {
        uint32_t ha[4], hb[4];
        blur_hrow(n0, ha);
        blur_hrow(n1, hb);
#pragma unroll
        for (int j = 0; j < 4; j++) Q[j] = __byte_perm(ha[j], hb[j], 0x5410);
    };
    const uint8_t* sp = src + (long long)(y0 - 3) * pitch;
    ldpair(sp, A0, A1); sp += 2 * pitch;
    ldpair(sp, B0, B1); sp += 2 * pitch;
    hpair(A0, A1, P[0]); ldpair(sp, A0, A1); sp += 2 * pitch;
    hpair(B0, B1, P[1]); ldpair(sp, B0, B1); sp += 2 * pitch;
    hpair(A0, A1, P[2]); ldpair(sp, A0, A1); sp += 2 * pitch;
    hpair(B0, B1, P[3]); ldpair(sp, B0, B1); sp += 2 * pitch;   // A = rows y0+5/6, B = rows y0+7/8, sp -> row y0 + 9
    uint8_t* dp = dst + (long long)y0 * pitch;
    constexpr uint32_t TA = 0x12002212u;  // lo pair taps (18,34): row y   | hi pair taps (0,18):  row y+1
    constexpr uint32_t TB = 0x30223830u;  // (48,56) | (34,48)
    constexpr uint32_t TC = 0x30382230u;  // (48,34) | (56,48)
    constexpr uint32_t TD = 0x12220012u;  // (18,0)  | (34,18)
    // two output rows from the window (A,B,C,D); sums stay below 2^24, so the result is byte 2 of the accumulator
    auto vout = [&](const uint32_t (&A)[4], const uint32_t (&B)[4], const uint32_t (&C)[4], const uint32_t (&D)[4]) {
        uint32_t a0[4], a1[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            a0[j] = __dp2a_lo(D[j], TD, __dp2a_lo(C[j], TC, __dp2a_lo(B[j], TB, __dp2a_lo(A[j], TA, 32768u))));
            a1[j] = __dp2a_hi(D[j], TD, __dp2a_hi(C[j], TC, __dp2a_hi(B[j], TB, __dp2a_hi(A[j], TA, 32768u))));
        }
        // words may run up to 3 bytes / one row past the interior into the (unused) frame area of the blurred plane
        *reinterpret_cast<uint32_t*>(dp) = __byte_perm(__byte_perm(a0[0], a0[1], 0x0062), __byte_perm(a0[2], a0[3], 0x0062), 0x5410);
        *reinterpret_cast<uint32_t*>(dp + pitch) = __byte_perm(__byte_perm(a1[0], a1[1], 0x0062), __byte_perm(a1[2], a1[3], 0x0062), 0x5410);
        dp += 2 * pitch;
    };
    // four row pairs per trip, the window rotating through P[0..3] by name (no register moves)
    // (the last two pairs loaded are not used: at most rows y1 + 4 .. y1 + 7, inside the 19-row frame)
    for (int y = y0; y < y1; y += 8) {
        vout(P[0], P[1], P[2], P[3]);
        if (y + 2 >= y1) break;
        hpair(A0, A1, P[0]); ldpair(sp, A0, A1); sp += 2 * pitch;
        vout(P[1], P[2], P[3], P[0]);
        if (y + 4 >= y1) break;
        hpair(B0, B1, P[1]); ldpair(sp, B0, B1); sp += 2 * pitch;
        vout(P[2], P[3], P[0], P[1]);
        if (y + 6 >= y1) break;
        hpair(A0, A1, P[2]); ldpair(sp, A0, A1); sp += 2 * pitch;
        vout(P[3], P[0], P[1], P[2]);
        if (y + 8 >= y1) break;
        hpair(B0, B1, P[3]); ldpair(sp, B0, B1); sp += 2 * pitch;
    }
}

// K5, TMA-staged variant (default when nlevels <= FAST_TMA_LEVELS; ORBX_BLUR_TMA=0 keeps k_blur): the CTA's source tile --
// 128 + 2 x 16 columns (the box starts on the 16-byte grid left of the CTA's first column) x 4 x BLUR_ROWS + 8 rows -- arrives
// through ONE cp.async.bulk.tensor.3d, and the threads run the same arithmetic from shared memory: no per-thread global
// loads, no source words held two steps ahead in registers (k_blur's top stall was the latency of exactly those loads).
constexpr int BLUR_TP = 160;                        // tile pitch in bytes = box width
constexpr int BLUR_TROWS = 4 * BLUR_ROWS + 8;       // rows y - 3 .. y + 4 BLUR_ROWS + 4 (the last one belongs to an unused pair)
struct BlurMaps { CUtensorMap m[FAST_TMA_LEVELS]; };
constexpr int BT_WARPS = 4;                         // warps per CTA of k_blur_tma; a thread slides down BT_ROWS rows of the tile
constexpr int BT_ROWS = 4 * BLUR_ROWS / BT_WARPS;   // (measured: 2 warps x 64 rows -- a smaller window warm-up share, half the warps -- 0.139 against 0.131 ms / 256 frames)

__global__ void __launch_bounds__(32 * BT_WARPS) k_blur_tma(Plan plan, Buffers buf, const __grid_constant__ BlurMaps maps)
{
    extern __shared__ __align__(128) unsigned char bsm[];
    __shared__ __align__(8) unsigned long long sBar;
    pdl_trigger();
    int level = 0;
    while (level + 1 < plan.nlevels && (int)blockIdx.x >= plan.lv[level + 1].blurOff) level++;
    const LevelGeom& g = plan.lv[level];
    const int frame = blockIdx.y + buf.frameBase;
    const int unit = blockIdx.x - g.blurOff;
    const int cgroups = (g.w + 127) >> 7;  // CTAs across (32 threads x 4 columns)
    const int cy = unit / cgroups, cxg = unit - cy * cgroups;
    pdl_wait();
    if (threadIdx.x == 0) {
        const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&sBar);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)(BLUR_TROWS * BLUR_TP)) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                     ::"r"((uint32_t)__cvta_generic_to_shared(bsm)), "l"(&maps.m[level]), "r"(bar), "r"(PAD_LEFT + 128 * cxg - 16),
                       "r"(PAD_TOP + 4 * BLUR_ROWS * cy - 3), "r"(frame)
                     : "memory");
    }
    __syncthreads();  // orders the mbarrier's initialisation before the waits
    {
        const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&sBar);
        uint32_t done;
        do {
            asm volatile(
                "{\n\t.reg .pred p;\n\t"
                "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                "selp.u32 %0, 1, 0, p;\n\t}"
                : "=r"(done)
                : "r"(bar), "r"(0u)
                : "memory");
        } while (!done);
    }
    const int x = 4 * (cxg * 32 + (threadIdx.x & 31));
    const int y0 = cy * 4 * BLUR_ROWS + (threadIdx.x >> 5) * BT_ROWS;
    if (x >= g.w || y0 >= g.h) return;
    const int y1 = min(y0 + BT_ROWS, g.h);
    // tile row 0 = image row 4 BLUR_ROWS cy - 3, tile column 16 = the CTA's first column
    const uint32_t* sp = reinterpret_cast<const uint32_t*>(bsm + (threadIdx.x >> 5) * BT_ROWS * BLUR_TP + 16 + 4 * (threadIdx.x & 31));
    constexpr int RW = BLUR_TP / 4;  // words per tile row
    uint8_t* dp = const_cast<uint8_t*>(interior_ptr(buf.blur, g, frame)) + x + (long long)y0 * g.pitch;
    const long long pitch = g.pitch;
    uint32_t P[4][4];  // window of 8 rows as 4 vertical pairs per column (see k_blur)
    auto hpair = [&](const uint32_t* r, uint32_t (&Q)[4]) {  // rows r and r + 1 row
        uint32_t n0[3] = {r[-1], r[0], r[1]}, n1[3] = {r[RW - 1], r[RW], r[RW + 1]}, ha[4], hb[4];
        blur_hrow(n0, ha);
        blur_hrow(n1, hb);
#pragma unroll
        for (int j = 0; j < 4; j++) Q[j] = __byte_perm(ha[j], hb[j], 0x5410);
    };
    hpair(sp, P[0]); hpair(sp + 2 * RW, P[1]); hpair(sp + 4 * RW, P[2]); hpair(sp + 6 * RW, P[3]);
    sp += 8 * RW;
    constexpr uint32_t TA = 0x12002212u, TB = 0x30223830u, TC = 0x30382230u, TD = 0x12220012u;  // vertical tap pairs (see k_blur)
    auto vout = [&](const uint32_t (&A)[4], const uint32_t (&B)[4], const uint32_t (&C)[4], const uint32_t (&D)[4]) {
        uint32_t a0[4], a1[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
            a0[j] = __dp2a_lo(D[j], TD, __dp2a_lo(C[j], TC, __dp2a_lo(B[j], TB, __dp2a_lo(A[j], TA, 32768u))));
            a1[j] = __dp2a_hi(D[j], TD, __dp2a_hi(C[j], TC, __dp2a_hi(B[j], TB, __dp2a_hi(A[j], TA, 32768u))));
        }
        *reinterpret_cast<uint32_t*>(dp) = __byte_perm(__byte_perm(a0[0], a0[1], 0x0062), __byte_perm(a0[2], a0[3], 0x0062), 0x5410);
        *reinterpret_cast<uint32_t*>(dp + pitch) = __byte_perm(__byte_perm(a1[0], a1[1], 0x0062), __byte_perm(a1[2], a1[3], 0x0062), 0x5410);
        dp += 2 * pitch;
    };
    for (int y = y0; y < y1; y += 8) {
        vout(P[0], P[1], P[2], P[3]);
        if (y + 2 >= y1) break;
        hpair(sp, P[0]);
        vout(P[1], P[2], P[3], P[0]);
        if (y + 4 >= y1) break;
        hpair(sp + 2 * RW, P[1]);
        vout(P[2], P[3], P[0], P[1]);
        if (y + 6 >= y1) break;
        hpair(sp + 4 * RW, P[2]);
        vout(P[3], P[0], P[1], P[2]);
        if (y + 8 >= y1) break;
        hpair(sp + 6 * RW, P[3]);
        sp += 8 * RW;
    }
}

// ------------------------------------------------------------------------------------------------------
// K4+K6: orientation (IC_Angle) and steered BRIEF.  One warp per keypoint (a warp loops over several keypoints).
//   orientation: the 31 rows of the radius-15 disc are staged from the pyramid level into shared memory; lane r owns row
//     v = r - 15: 8 re-aligned words x IDP.4A against per-lane constant weight words (u masked by |u| <= umax[|v|], and
//     the 0/1 mask itself for the row sum), m10 = sum u*I, m01 = sum v*I, __reduce_add_sync, then cv::fastAtan2 in
//     unfused float arithmetic.
//   descriptor: a = cosf(angle*factorPI), b = sinf(...) (glibc-exact), sample (x,y) -> (round(x*a - y*b), round(x*b + y*a))
//     with unfused float arithmetic and round-half-even, read from the 37x37 blurred neighbourhood staged in shared
//     memory; lane i produces descriptor byte i (16 samples).
// Also finalises the keypoint record (level-0 coordinates, size, octave) and the per-frame count.
// ------------------------------------------------------------------------------------------------------
constexpr int DESC_WARPS = 4;
constexpr int DESC_PP = 64;   // blurred patch pitch in bytes: 37 columns at any 16-byte alignment span <= 15 + 37 = 52 bytes
constexpr int ORI_PB = 48;    // raw patch pitch in bytes: 31 columns at any 16-byte alignment span <= 15 + 31 = 46 bytes
constexpr int DESC_PBYTES = 37 * DESC_PP, ORI_BYTES = 31 * ORI_PB;

__device__ __forceinline__ int dp4a_u8s8(uint32_t a, uint32_t b, int c)
{
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// IC_Angle moments of one disc row held as 12 words starting at a 16-byte boundary; W4 = word offset of column cx-16
// (warp-uniform, so the switch in the caller does not diverge), sh8 = bit offset inside that word.
template <int W4>
__device__ __forceinline__ void ori_row(const uint4 q0, const uint4 q1, const uint4 q2, int sh8, const uint32_t (&wU)[8],
                                        const uint32_t (&wOne)[8], int& m10, int& rs)
{
    const uint32_t w[12] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w};
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const uint32_t px = __funnelshift_r(w[W4 + k], w[W4 + k + 1], sh8);  // columns 4k-16 .. 4k-13 relative to cx
        m10 = dp4a_u8s8(px, wU[k], m10);
        rs = (int)__dp4a(px, wOne[k], (uint32_t)rs);
    }
}

__global__ void __launch_bounds__(DESC_WARPS * 32, 7) k_describe(Plan plan, Buffers buf, orbx_keypoint* outKps,
                                                               uint8_t* outDesc, int* outCounts, int capacity)
{
    // per warp: two buffers for the blurred patch (keypoint i is sampled from one while cp.async fills the other for i + 1) and
    // ONE for the raw disc rows (IC_Angle reads them into registers first thing; the copy for i + 1 starts right after)
    __shared__ __align__(16) uint8_t stageB[DESC_WARPS][2][DESC_PBYTES];
    __shared__ __align__(16) uint8_t stageR[DESC_WARPS][ORI_BYTES];
    pdl_trigger();
    const int lframe = blockIdx.y, frame = lframe + buf.frameBase;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    __shared__ int cum[ORBX_MAX_LEVELS + 1];

    // per-lane constants of the orientation pass: row v = lane - 15, columns u = 4k + b - 16
    const int v = lane - 15;
    const int dmax = lane < 31 ? umax_of(v < 0 ? -v : v) : -1;
    uint32_t wU[8], wOne[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {
        uint32_t a = 0, o = 0;
#pragma unroll
        for (int bb = 0; bb < 4; bb++) {
            const int u = 4 * k + bb - 16;
            const bool in = (u < 0 ? -u : u) <= dmax;
            a |= (in ? (uint32_t)(u & 0xff) : 0u) << (8 * bb);
            o |= (in ? 1u : 0u) << (8 * bb);
        }
        wU[k] = a;
        wOne[k] = o;
    }
    // pattern coordinates as floats, [pair k][lane] as (x0, y0, x1, y1): conflict-free 128-bit reads, no I2F in the loop
    __shared__ __align__(16) float4 patF[8 * 32];
    for (int i = threadIdx.x; i < 256; i += DESC_WARPS * 32) {
        const int k = i >> 5, ln = i & 31;
        const signed char* pp = d_pattern + 32 * ln + 4 * k;
        patF[i] = make_float4((float)pp[0], (float)pp[1], (float)pp[2], (float)pp[3]);
    }
    // per-level constants of this CTA's frame in shared memory (the keypoint loop would otherwise index the kernel-argument
    // plan dynamically and redo 64-bit products per keypoint): record index offset, plane offsets of both arenas, pitch
    __shared__ int recOff[ORBX_MAX_LEVELS];           // staged record of keypoint gi of level l = gi + recOff[l]
    __shared__ long long lvPlane[ORBX_MAX_LEVELS];    // byte offset of interior pixel (0, 0) of the frame's level-l plane
    __shared__ int lvPitch[ORBX_MAX_LEVELS];
    pdl_wait();  // everything above reads only constants: it overlaps the tail of the previous kernel
    if (threadIdx.x == 0) {
        const int* lc = buf.levelCount + frame * plan.nlevels;
        int acc = 0;
        cum[0] = 0;
        for (int l = 0; l < plan.nlevels; l++) {
            recOff[l] = plan.lv[l].kpOff - acc;
            acc += lc[l];
            cum[l + 1] = acc;
            const LevelGeom& gl = plan.lv[l];
            lvPlane[l] = gl.planeOff + (long long)frame * gl.planeBytes + (long long)PAD_TOP * gl.pitch + PAD_LEFT;
            lvPitch[l] = gl.pitch;
        }
    }
    __syncthreads();  // cum, the per-level constants and patF
    const int total = min(cum[plan.nlevels], capacity);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        outCounts[lframe] = total;
        // a caller's capacity below the frame's keypoint count truncates the output: report it instead of hiding it
        if (cum[plan.nlevels] > capacity) *reinterpret_cast<volatile int*>(buf.errFlag) = 1;
    }
    const uint32_t stage0 = (uint32_t)__cvta_generic_to_shared(&stageB[wid][0][0]);
    const uint32_t stageR0 = (uint32_t)__cvta_generic_to_shared(&stageR[wid][0]);

    // staging maps: a lane copies 16-byte segments; blurred = 37 rows x 4 segments (lane -> segment lane & 3, rows
    // (lane >> 2) + 8 i), raw = 31 rows x 3 segments (lanes 0..29 -> segment lane % 3, rows lane / 3 + 10 i).  Segments
    // that the keypoint's alignment leaves unused are not fetched.
    const int bseg = lane & 3, brow = lane >> 2;
    const int rseg = lane % 3, rrow = lane / 3;
    // Three keypoints are in flight per warp: i is processed from shared memory, the patches of i+1 are being copied, and
    // the 16-byte record of i+2 is being loaded (the copy addresses of a keypoint depend on its record, so loading the
    // record one iteration earlier keeps that global-load latency off the warp's critical path).
    // The warp's keypoints are gidx0 + j * gstride, j < 32 (the grid gives a warp at most 12): lane j loads the 16-byte record
    // of keypoint j once, up front (level by a per-lane search), and the loop broadcasts record i / i + 1 with shuffles.
    const StagedKp* __restrict__ stagedFrame = buf.staged + (long long)frame * plan.kpCapFrame;
    const int gstride = gridDim.x * DESC_WARPS;
    const int gidx0 = blockIdx.x * DESC_WARPS + wid;
    float myX = 0.f, myY = 0.f, myResp = 0.f;
    int myLev = 0;
    {
        const int gi = gidx0 + lane * gstride;
        if (gi < total) {
            while (gi >= cum[myLev + 1]) myLev++;
            const StagedKp rec = stagedFrame[gi + recOff[myLev]];
            myX = rec.x; myY = rec.y; myResp = rec.response;
        }
    }
    auto load_record = [&](int it, StagedKp& rec, int& lev) {  // it = index of the keypoint within the warp's sequence (< 32)
        rec.x = __shfl_sync(0xffffffffu, myX, it);
        rec.y = __shfl_sync(0xffffffffu, myY, it);
        rec.response = __shfl_sync(0xffffffffu, myResp, it);
        lev = __shfl_sync(0xffffffffu, myLev, it);
    };
    auto fetch_blur = [&](const StagedKp& rec, int lev, int slot) {
        const int cx = (int)rec.x, cy = (int)rec.y;
        const int pitch = lvPitch[lev];
        // interior column 0 is 32-byte aligned in both arenas, so the alignment of a column is that of its index
        const int bsh = (cx - 18) & 15;
        if (bseg * 16 < bsh + 37) {
            const uint8_t* pb = buf.blur + lvPlane[lev] + ((cy - 18 + brow) * pitch + (cx - 18 - bsh) + 16 * bseg);
            uint32_t dst = stage0 + slot * DESC_PBYTES + brow * DESC_PP + 16 * bseg;
#pragma unroll
            for (int i = 0; i < 5; i++) {
                if (brow + 8 * i < 37)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 8 * i * DESC_PP), "l"(pb + (long long)(8 * i) * pitch) : "memory");
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    auto fetch_raw = [&](const StagedKp& rec, int lev) {
        const int cx = (int)rec.x, cy = (int)rec.y;
        const int pitch = lvPitch[lev];
        const int rsh = (cx - 16) & 15;
        if (lane < 30 && rseg * 16 < rsh + 32) {
            const uint8_t* pr = buf.pyr + lvPlane[lev] + ((cy - 15 + rrow) * pitch + (cx - 16 - rsh) + 16 * rseg);
            uint32_t dst = stageR0 + rrow * ORI_PB + 16 * rseg;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                if (rrow + 10 * i < 31)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 10 * i * ORI_PB), "l"(pr + (long long)(10 * i) * pitch) : "memory");
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    int gidx = gidx0;
    int slot = 0, it = 0;
    StagedKp kp, kpN;
    int level = 0, levelN = 0;
    kp.x = kp.y = kp.angle = kp.response = 0.f;
    kpN = kp;
    if (gidx < total) {
        load_record(0, kp, level);
        fetch_blur(kp, level, 0);
        fetch_raw(kp, level);
        load_record(1, kpN, levelN);   // (lanes without a keypoint hold zeros: used only when gidx + gstride < total)
    }
    for (; gidx < total; gidx += gstride, slot ^= 1, it++) {
        const LevelGeom& g = plan.lv[level];
        // cx and the warp index are the same in every lane; taking them through REDUX tells the compiler so, which keeps the
        // patch base + alignment offset in a uniform register (shared-memory loads then address [R + UR])
        const int cx = __reduce_min_sync(0xffffffffu, (int)kp.x);
        const int widu = __reduce_min_sync(0xffffffffu, wid);
        const uint8_t* P = &stageB[widu][slot][0];
        const uint8_t* R = &stageR[widu][0];
        asm volatile("cp.async.wait_all;" ::: "memory");
        __syncwarp();  // this keypoint's patches have landed (all lanes); the previous keypoint's reads of the other buffer are done
        if (gidx + gstride < total) fetch_blur(kpN, levelN, slot ^ 1);

        // ---- IC_Angle ----
        int m10 = 0, rs = 0;
        if (lane < 31) {
            const int rsh = (cx - 16) & 15;
            const int sh8 = 8 * (rsh & 3);
            const uint4* rq = reinterpret_cast<const uint4*>(R + lane * ORI_PB);
            const uint4 q0 = rq[0], q1 = rq[1], q2 = rq[2];
            switch (rsh >> 2) {  // warp-uniform
            case 0: ori_row<0>(q0, q1, q2, sh8, wU, wOne, m10, rs); break;
            case 1: ori_row<1>(q0, q1, q2, sh8, wU, wOne, m10, rs); break;
            case 2: ori_row<2>(q0, q1, q2, sh8, wU, wOne, m10, rs); break;
            default: ori_row<3>(q0, q1, q2, sh8, wU, wOne, m10, rs); break;
            }
        }
        m10 = __reduce_add_sync(0xffffffffu, m10);
        const int m01 = __reduce_add_sync(0xffffffffu, v * rs);
        // every lane has consumed its raw row (the reductions above are warp-wide): the raw buffer takes the next keypoint's rows
        if (gidx + gstride < total) fetch_raw(kpN, levelN);
        const float angle = fast_atan2_dev((float)m01, (float)m10);
        // ---- descriptor ----
        const float factorPI = 0.0174532924f;  // (float)(CV_PI/180.f)
        const SinCosOut sc = sincosf_glibc(__fmul_rn(angle, factorPI));
        const float a = sc.c, b = sc.s;
        int val = 0;
        // cvRound = round-half-even: adding 1.5 * 2^23 leaves the rounded integer in the low mantissa bits (|v| < 2^22);
        // the bias 0x4B400000 of both coordinates is folded into the patch base (32-bit wrap-around arithmetic)
        const float RND = 12582912.f;
        // (the shared-memory address of the patch is folded in as well: one 32-bit base, two instructions per sample address)
        const uint32_t cbase = (uint32_t)__cvta_generic_to_shared(P) + (uint32_t)(18 * DESC_PP + ((cx - 18) & 15) + 18) -
                               (uint32_t)(DESC_PP + 1) * 0x4B400000u;
        auto sample = [](uint32_t addr) {
            uint32_t v;
            asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
            return (int)v;
        };
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const float4 pt = patF[k * 32 + lane];
            const uint32_t iy0 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(pt.x, b), __fmul_rn(pt.y, a)), RND));
            const uint32_t ix0 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(pt.x, a), __fmul_rn(pt.y, b)), RND));
            const uint32_t iy1 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(pt.z, b), __fmul_rn(pt.w, a)), RND));
            const uint32_t ix1 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(pt.z, a), __fmul_rn(pt.w, b)), RND));
            const int t0 = sample(iy0 * DESC_PP + (ix0 + cbase)), t1 = sample(iy1 * DESC_PP + (ix1 + cbase));
            val |= (t0 < t1) << k;
        }
        const long long o = (long long)lframe * capacity + gidx;
        outDesc[o * 32 + lane] = (uint8_t)val;
        if (lane == 0) {
            orbx_keypoint r;
            r.x = level ? __fmul_rn(kp.x, g.scale) : kp.x;
            r.y = level ? __fmul_rn(kp.y, g.scale) : kp.y;
            r.size = (float)g.patchSize;
            r.angle = angle;
            r.response = kp.response;
            r.octave = level;
            outKps[o] = r;
        }
        kp = kpN; level = levelN;
        load_record(it + 2, kpN, levelN);   // keypoint it + 2 of this warp (zeros past the end; lane index wraps harmlessly)
    }
}

// ------------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------------
static inline int cv_round_f(float v) { return (int)lrintf(v); }
static inline int cv_floor_f(float v) { int i = (int)v; return i - (v < (float)i); }
static inline short sat_short(float v)
{
    int i = cv_round_f(v);
    return (short)(i < -32768 ? -32768 : (i > 32767 ? 32767 : i));
}

}  // namespace orbx

using namespace orbx;

struct orbx_handle {
    orbx_config cfg;
    int device;
    std::vector<float> scale, invScale, sigma2, invSigma2;
    std::vector<int> quota;
    cudaStream_t stream = nullptr, streamBlur = nullptr, streamH2D = nullptr, streamD2H = nullptr;
    static constexpr int PIPE_IN_SLOTS = 4;  // input staging slots of the pipelined host path (H2D runs ahead of the kernels)
    cudaEvent_t evInReady[PIPE_IN_SLOTS] = {}, evInFree[PIPE_IN_SLOTS] = {};
    cudaStream_t stream2 = nullptr, streamBlur2 = nullptr;   // second compute lane of the pipelined host path
    cudaEvent_t evFork2 = nullptr, evJoin2 = nullptr, evEntry = nullptr;
    std::vector<cudaEvent_t> evChunkDone;
    int* hErrFlag = nullptr;   // mapped pinned host word written by kernels on candidate overflow
    int* dErrFlag = nullptr;
    cudaEvent_t evFork = nullptr, evJoin = nullptr;
    cudaEvent_t evStage[ORBX_NUM_STAGES + 1] = {};
    // plan + arenas (valid for planRows x planCols, frameCap frames)
    bool havePlan = false;
    Plan plan;
    int frameCap = 0;
    Buffers buf = {};
    uint8_t* dIn = nullptr;      // staged input frames (host API)
    size_t dInBytes = 0;
    void* dOut = nullptr;        // [counts | kps | desc] for the host API
    size_t dOutBytes = 0;
    void* hOut = nullptr;        // pinned mirror
    size_t hOutBytes = 0;
    short* dTabs = nullptr;
    unsigned short* dKeyNode = nullptr;
    FastUnit* dFastUnits = nullptr;
    uint8_t* dZeros = nullptr;
    int octNodeCap = 0;
    int octKeySmemCap[3] = {0, 0, 0};  // per k_octree variant: 128, 256, 512 threads
    size_t fastSmem = 0;
    FastMaps fastMaps = {};      // per-level tensor maps of the pyramid arena (k_fast_band<true>)
    bool fastTma = false;
    BlurMaps blurMaps = {};      // same tensors, box BLUR_TP x BLUR_TROWS (k_blur_tma)
    bool blurTma = false;
    ResizeMaps resizeMaps = {};  // same tensors, box RZ_TP x RZ_TROWS (k_resize_rows_tma, indexed by source level)
    bool resizeTma = false;
    size_t octSmem[3] = {0, 0, 0};
    int lastFrames = 0;          // frames of the last batch
    // which frames of the last batch are still in the arena: the pipelined host path keeps only its last two chunks
    struct Resident { int first, count, arenaFirst; };
    Resident resident[2] = {{0, 0, 0}, {0, 0, 0}};
    int arena_frame(int frame) const   // arena index of batch frame `frame`, -1 when it has been overwritten
    {
        for (const Resident& r : resident)
            if (frame >= r.first && frame < r.first + r.count) return r.arenaFirst + (frame - r.first);
        return -1;
    }
    int plainRuns = 0;           // batches enqueued without a graph since the plan was (re)built
    bool profiling = false;
    bool useGraph = true;
    float stageMs[ORBX_NUM_STAGES] = {};
    int lastLaunches = 0;
    // graph cache for the device-resident path: key = (n, in ptr, steps, out ptrs)
    struct GraphKey { const void* in; long long step, fstride; void* kps; void* desc; void* counts; int n, cap, slot, frameBase; };
    struct GraphEntry { GraphKey key; cudaGraphExec_t exec; int launches; unsigned long long stamp; };
    std::vector<GraphEntry> graphs;   // small LRU cache (the pipelined host path uses one graph per chunk slot)
    unsigned long long graphStamp = 0;
};

static void free_arenas(orbx_handle* h)
{
    cudaFree(h->buf.pyr); cudaFree(h->buf.blur); cudaFree(h->buf.cand); cudaFree(h->buf.candCount);
    cudaFree(h->buf.candCountDbg); cudaFree(h->buf.staged); cudaFree(h->buf.levelCount);
    cudaFree(h->dTabs); cudaFree(h->dKeyNode); cudaFree(h->dFastUnits); cudaFree(h->dZeros);
    h->buf = Buffers{};
    h->dTabs = nullptr; h->dKeyNode = nullptr; h->dFastUnits = nullptr; h->dZeros = nullptr;
    for (auto& g : h->graphs) cudaGraphExecDestroy(g.exec);
    h->graphs.clear();
    h->havePlan = false;
    h->frameCap = 0;
}

// Level geometry, cell grids, capacities, resize tables for one image shape.
struct TabInfo { long long nx, ny, nxw, nyw, yrowOff, srowOff; std::vector<FastUnit> units; };  // padded lengths of xofs / yofs and true lengths of the weight arrays

static int build_plan(orbx_handle* h, int rows, int cols, Plan& P, std::vector<short>& tabs, long long& pyrBytesPerFrame,
                      TabInfo* ti = nullptr)
{
    const orbx_config& c = h->cfg;
    memset(&P, 0, sizeof P);
    P.nlevels = c.nlevels; P.rows = rows; P.cols = cols; P.iniTh = c.ini_th_fast; P.minTh = c.min_th_fast;
    long long planeOff = 0, candOff = 0;
    int fastOff = 0, kpOff = 0, fastRows = 0, fastQ = 0, blurOff = 0;
    // tables: per level >= 1: xofs[w], xw[2w], yofs0[h], yofs1[h], yw[2h] -- stored in 5 separate arrays
    std::vector<short> xofs, xw, y0v, y1v, yw;
    std::vector<int> yrow;  // 4 ints per destination row, indexed like y0v
    std::vector<int> srow;  // 2 ints per SOURCE row of every level >= 1's source level (k_resize_rows)
    for (int l = 0; l < c.nlevels; l++) {
        LevelGeom& g = P.lv[l];
        g.w = cv_round_f((float)cols * h->invScale[l]);
        g.h = cv_round_f((float)rows * h->invScale[l]);
        if (g.w >= (1 << 15) || g.h >= (1 << 15) || (long long)g.w * g.h >= (1 << 24)) {
            set_error("level %d is %dx%d: larger than the supported 2^24 pixels per level", l, g.w, g.h);
            return ORBX_ERR_BAD_ARG;
        }
        const int minB = EDGE - 3;
        g.maxBX = g.w - EDGE + 3; g.maxBY = g.h - EDGE + 3;
        const float width = (float)(g.maxBX - minB), height = (float)(g.maxBY - minB);
        g.nCols = (int)(width / 30.f); g.nRows = (int)(height / 30.f);
        if (g.nCols < 1 || g.nRows < 1) {
            set_error("level %d is %dx%d: smaller than one 30-px FAST cell inside the 16-px border", l, g.w, g.h);
            return ORBX_ERR_TOO_SMALL;
        }
        g.wCell = (int)std::ceil(width / g.nCols); g.hCell = (int)std::ceil(height / g.nRows);
        g.wCellMagic = 65536 / g.wCell + 1;
        g.nIni = (int)std::round((float)(g.maxBX - minB) / (g.maxBY - minB));
        if (g.nIni < 1) {
            set_error("level %d is %dx%d: aspect ratio gives zero quad-tree roots", l, g.w, g.h);
            return ORBX_ERR_TOO_SMALL;
        }
        g.hX = (float)(g.maxBX - minB) / g.nIni;
        g.pitch = align_up(PAD_LEFT + g.w + EDGE, 128);
        g.planeRows = g.h + 2 * EDGE;
        g.planeBytes = (long long)g.pitch * g.planeRows;
        g.planeOff = planeOff;  // finalised below once the frame capacity is known (layout [level][frame])
        planeOff += g.planeBytes;
        g.blurOff = blurOff; blurOff += ((g.w + 127) / 128) * ((g.h + 4 * BLUR_ROWS - 1) / (4 * BLUR_ROWS));
        g.fastG = std::max(1, std::min(FAST_MAXG, 124 / g.wCell));
        g.fastGroups = (g.nCols + g.fastG - 1) / g.fastG;
        g.fastOff = fastOff; fastOff += g.nRows * g.fastGroups;
        fastRows = std::max(fastRows, g.hCell + 6);
        // worst case after the 3x3 NMS, which runs per cell (neighbours across a cell border both survive): every other pixel
        // of every other row of each of the band's cells
        fastQ = std::max(fastQ, ((std::min(g.fastG * g.wCell, 124) + 1) / 2 + 1) * ((g.hCell + 1) / 2 + 1));
        fastQ = std::max(fastQ, g.fastG * ((g.wCell + 1) / 2) * ((g.hCell + 1) / 2));
        g.quota = h->quota[l];
        // capacity of the level's candidate list = the per-cell NMS worst case summed over the cell grid (the last row and
        // column of cells are clipped at maxBX - 3 / maxBY - 3)
        const int evalW = g.maxBX - minB - 6, evalH = g.maxBY - minB - 6;
        int capX = 0, capY = 0;
        for (int j = 0; j < g.nCols; j++) capX += (std::max(0, std::min(g.wCell, evalW - j * g.wCell)) + 1) / 2;
        for (int i = 0; i < g.nRows; i++) capY += (std::max(0, std::min(g.hCell, evalH - i * g.hCell)) + 1) / 2;
        g.candCap = std::max(capX * capY, ((evalW + 1) / 2 + 1) * ((evalH + 1) / 2 + 1));
        g.candOff = candOff; candOff += g.candCap;
        g.nodeCap = std::max(g.quota + 3, 4 * g.nIni) + 1;
        g.kpOff = kpOff; kpOff += g.nodeCap;
        g.scale = h->scale[l];
        g.patchSize = (int)(31 * h->scale[l]);
        g.xTab = (int)xofs.size(); g.yTab = (int)y0v.size();
        g.resizeWords = 0; g.resizeRows = 0; g.sTab = 0; g.resizeTma = 0;
        if (l > 0) {
            const LevelGeom& gp = P.lv[l - 1];
            // cv::resize INTER_LINEAR coefficient tables (see oracle/cvprim.h: linear_coeffs)
            {
                double scale = 1.0 / ((double)g.w / gp.w);
                for (int d = 0; d < g.w; d++) {
                    float f = (float)((d + 0.5) * scale - 0.5);
                    int s = cv_floor_f(f);
                    f -= s;
                    if (s < 0) { f = 0; s = 0; }
                    if (s >= gp.w - 1) { f = 0; s = gp.w - 1; }
                    xofs.push_back((short)s);
                    xw.push_back(sat_short((1.f - f) * 2048)); xw.push_back(sat_short(f * 2048));
                }
            }
            g.sTab = (int)(srow.size() / 2);
            g.resizeRows = 1;
            srow.resize(srow.size() + 2 * (size_t)gp.h, 0);
            int prevR0 = -1;
            {
                double scale = 1.0 / ((double)g.h / gp.h);
                for (int d = 0; d < g.h; d++) {
                    float f = (float)((d + 0.5) * scale - 0.5);
                    int s = cv_floor_f(f);
                    f -= s;
                    auto clip = [&](int x) { return x >= 0 ? (x < gp.h ? x : gp.h - 1) : 0; };
                    y0v.push_back((short)clip(s)); y1v.push_back((short)clip(s + 1));
                    const short wb0 = sat_short((1.f - f) * 2048), wb1 = sat_short(f * 2048);
                    yw.push_back(wb0); yw.push_back(wb1);
                    // destination rows 1..19 and h-20..h-2 are also the REFLECT_101 frame rows -d and 2(h-1)-d
                    // (as row deltas: low half = upper frame, high half = lower frame, 0 = none)
                    int dup = 0;
                    if (d >= 1 && d <= EDGE) dup |= (-2 * d) & 0xffff;
                    if (d >= g.h - 1 - EDGE && d <= g.h - 2) dup |= (2 * (g.h - 1 - d)) << 16;
                    yrow.push_back(clip(s) * gp.pitch); yrow.push_back(clip(s + 1) * gp.pitch);
                    yrow.push_back((int)((uint32_t)(unsigned short)wb0 | ((uint32_t)(unsigned short)wb1 << 16))); yrow.push_back(dup);
                    // source-row walk: r1 = r0 + 1 with r0 strictly increasing makes r1 the lower row of exactly this destination row
                    if (clip(s + 1) != clip(s) + 1 || clip(s) <= prevR0 || wb0 < 0 || wb1 < 0) g.resizeRows = 0;
                    prevR0 = clip(s);
                    srow[2 * ((size_t)g.sTab + clip(s + 1))] = (int)((uint32_t)(unsigned short)wb0 | ((uint32_t)(unsigned short)wb1 << 16));
                    srow[2 * ((size_t)g.sTab + clip(s + 1)) + 1] = dup;
                }
            }
            // k_resize_words eligibility: every aligned group of 4 buffer columns (frame columns map to the reflected
            // destination column) must read source columns inside [base, base + 7]
            g.resizeWords = 1;
            for (int bx = 0; bx < g.pitch && g.resizeWords; bx += 4) {
                int lo = 1 << 30, hi = -1;
                for (int j = 0; j < 4; j++) {
                    int dx = std::min(std::max(bx + j - PAD_LEFT, -EDGE), g.w - 1 + EDGE);
                    if (dx < 0) dx = -dx;
                    if (dx >= g.w) dx = 2 * (g.w - 1) - dx;
                    const int sxv = xofs[g.xTab + dx];
                    lo = std::min(lo, sxv); hi = std::max(hi, std::min(sxv + 1, gp.w - 1));
                }
                if (hi - lo > 7) g.resizeWords = 0;
            }
            // k_resize_rows_tma eligibility: the source window of every CTA (128 buffer columns x 4 RZ_RPT destination rows)
            // fits the RZ_TP x RZ_TROWS tile, and the box does not exceed the source plane
            g.resizeTma = g.resizeWords && g.resizeRows && gp.pitch >= RZ_TP && gp.planeRows >= RZ_TROWS;
            for (int bx0 = 0; bx0 < g.pitch && g.resizeTma; bx0 += 128) {
                int lo = 1 << 30, hi = -1;
                for (int bx = bx0; bx < bx0 + 128; bx += 4) {
                    int base = 1 << 30;
                    for (int j = 0; j < 4; j++) {
                        int dx = std::min(std::max(bx + j - PAD_LEFT, -EDGE), g.w - 1 + EDGE);
                        if (dx < 0) dx = -dx;
                        if (dx >= g.w) dx = 2 * (g.w - 1) - dx;
                        base = std::min(base, (int)xofs[g.xTab + dx]);
                    }
                    const int colOff = (PAD_LEFT + base) & ~3;
                    lo = std::min(lo, colOff & ~15); hi = std::max(hi, colOff + 12);
                }
                if (hi - lo > RZ_TP) g.resizeTma = 0;
            }
            for (int d0 = 0; d0 < g.h && g.resizeTma; d0 += 4 * RZ_RPT) {
                const int dl = std::min(d0 + 4 * RZ_RPT, g.h) - 1;
                if (y1v[g.yTab + dl] - y0v[g.yTab + d0] + 2 > RZ_TROWS) g.resizeTma = 0;
            }
        }
    }
    P.totalBlurUnits = blurOff;
    P.fastTileBytes = fastRows * FAST_P;
    P.fastQCap = align_up(fastQ + 8, 8);
    P.kpCapFrame = kpOff;
    P.candFrame = candOff;
    pyrBytesPerFrame = planeOff;
    // pack tables: [xofs | y0 | y1 | xw | yw] with 4-byte alignment for the short2 parts
    auto pad2 = [](std::vector<short>& v) { if (v.size() & 1) v.push_back(0); };
    pad2(xofs); pad2(y0v); pad2(y1v);
    tabs.clear();
    tabs.insert(tabs.end(), xofs.begin(), xofs.end());
    tabs.insert(tabs.end(), y0v.begin(), y0v.end());
    tabs.insert(tabs.end(), y1v.begin(), y1v.end());
    tabs.insert(tabs.end(), xw.begin(), xw.end());
    tabs.insert(tabs.end(), yw.begin(), yw.end());
    while (tabs.size() & 7) tabs.push_back(0);  // 16-byte alignment for the int4 row table
    const long long yrowOff = (long long)tabs.size();
    {
        const short* yr = reinterpret_cast<const short*>(yrow.data());
        tabs.insert(tabs.end(), yr, yr + 2 * yrow.size());
    }
    const long long srowOff = (long long)tabs.size();  // 16-byte aligned: the int4 table before it is
    {
        const short* sr = reinterpret_cast<const short*>(srow.data());
        tabs.insert(tabs.end(), sr, sr + 2 * srow.size());
    }
    std::vector<FastUnit> units;
    for (int l = 0; l < c.nlevels; l++) {
        const LevelGeom& g = P.lv[l];
        for (int i = 0; i < g.nRows; i++)
            for (int gi = 0; gi < g.fastGroups; gi++) {
                // band geometry: cells skipped by the reference (src/ORBextractor.cc:1064, :1081) never become units
                const int j0 = gi * g.fastG;
                const int iniY = 16 + i * g.hCell, iniX = 16 + j0 * g.wCell;
                if (iniY >= g.maxBY - 3 || iniX >= g.maxBX - 6) continue;
                const int rh = std::min(iniY + g.hCell + 6, g.maxBY) - iniY;
                if (rh < 7) continue;
                int j1 = std::min(j0 + g.fastG - 1, g.nCols - 1);
                while (16 + j1 * g.wCell >= g.maxBX - 6) j1--;
                const int bw = std::min(16 + j1 * g.wCell + g.wCell + 6, g.maxBX) - iniX;
                if (bw < 7) continue;
                units.push_back(FastUnit{(short)l, (short)i, (short)j0, (short)iniX, (short)iniY, (short)rh, (short)bw, 0});
            }
    }
    P.totalFastUnits = (int)units.size();
    if (ti) {
        ti->nx = (long long)xofs.size(); ti->ny = (long long)y0v.size(); ti->nxw = (long long)xw.size(); ti->nyw = (long long)yw.size(); ti->yrowOff = yrowOff; ti->srowOff = srowOff;
        ti->units.swap(units);
    }
    P.pyrFrameStride = pyrBytesPerFrame;
    return ORBX_OK;
}

// Plan (re)builds and graph captures are rare; they are serialised across handles: allocation, function-attribute
// changes and first-launch module loading in one host thread must not overlap a stream capture in another (two
// extractor objects are driven from two std::threads in stereo, src/Frame.cc:127-131).  Steady-state graph launches
// and copies of different handles run concurrently.
static std::mutex g_setupMutex;

static int ensure_plan(orbx_handle* h, int rows, int cols, int frames)
{
    if (h->havePlan && h->plan.rows == rows && h->plan.cols == cols && frames <= h->frameCap) return ORBX_OK;
    std::lock_guard<std::mutex> setupLock(g_setupMutex);
    h->plainRuns = 0;
    ORBX_CUDA_TRY(cudaSetDevice(h->device));
    ORBX_CUDA_TRY(cudaStreamSynchronize(h->stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(h->streamBlur));
    Plan P;
    std::vector<short> tabs;
    long long pyrPerFrame = 0;
    TabInfo ti;
    int rc = build_plan(h, rows, cols, P, tabs, pyrPerFrame, &ti);
    if (rc != ORBX_OK) return rc;
    // quad-tree kernel shared memory (validated before the old arenas are released)
    int nodeCapMax = 0, candCapMax = 0;
    for (int l = 0; l < P.nlevels; l++) { nodeCapMax = std::max(nodeCapMax, P.lv[l].nodeCap); candCapMax = std::max(candCapMax, P.lv[l].candCap); }
    nodeCapMax = align_up(nodeCapMax, 8);
    if (nodeCapMax > 16000) {
        set_error("per-level feature quota %d exceeds the quad-tree kernel's node capacity", nodeCapMax);
        return ORBX_ERR_BAD_ARG;
    }
    const size_t perNode = 8 + 8 + 4 + 4 + 16 + 4 + 4 + 4 + 4 + 8 + 2 + 2 + 2 + 2 + 2 + 2;
    const size_t nodeBytes = perNode * nodeCapMax + 16 * 16;
    int devSmem = 0;
    ORBX_CUDA_TRY(cudaDeviceGetAttribute(&devSmem, cudaDevAttrMaxSharedMemoryPerBlockOptin, h->device));
    if (nodeBytes + 1024 > (size_t)devSmem) {
        set_error("quad-tree node arrays (%zu B) exceed shared memory (%d B)", nodeBytes, devSmem);
        return ORBX_ERR_BAD_ARG;
    }
    const int cap = std::max(frames, (h->havePlan && h->plan.rows == rows && h->plan.cols == cols) ? h->frameCap : 0);
    free_arenas(h);
    // layout [level][frame]: level l starts at cap * (sum of planeBytes of lower levels)
    for (int l = 0; l < P.nlevels; l++) P.lv[l].planeOff *= cap;
    // allocate straight into the handle: a failure below leaves the partial set where free_arenas() (next plan / destroy)
    // releases it, and havePlan stays false
    Buffers& B = h->buf;
    B = Buffers{};
    B.frames = cap;
    const size_t pyrBytes = (size_t)pyrPerFrame * cap + 256;
    ORBX_CUDA_TRY(cudaMalloc(&B.pyr, pyrBytes));
    ORBX_CUDA_TRY(cudaMalloc(&B.blur, pyrBytes));
    ORBX_CUDA_TRY(cudaMemsetAsync(B.pyr, 0, pyrBytes, h->stream));
    ORBX_CUDA_TRY(cudaMemsetAsync(B.blur, 0, pyrBytes, h->stream));
    ORBX_CUDA_TRY(cudaMalloc(&B.cand, sizeof(Cand) * (size_t)P.candFrame * cap));
    ORBX_CUDA_TRY(cudaMalloc(&B.candCount, sizeof(int) * (size_t)P.nlevels * cap));
    ORBX_CUDA_TRY(cudaMalloc(&B.candCountDbg, sizeof(int) * (size_t)P.nlevels * cap));
    ORBX_CUDA_TRY(cudaMemsetAsync(B.candCount, 0, sizeof(int) * (size_t)P.nlevels * cap, h->stream));
    ORBX_CUDA_TRY(cudaMemsetAsync(B.candCountDbg, 0, sizeof(int) * (size_t)P.nlevels * cap, h->stream));
    ORBX_CUDA_TRY(cudaMalloc(&B.staged, sizeof(StagedKp) * (size_t)P.kpCapFrame * cap));
    ORBX_CUDA_TRY(cudaMalloc(&B.levelCount, sizeof(int) * (size_t)P.nlevels * cap));
    ORBX_CUDA_TRY(cudaMemsetAsync(B.levelCount, 0, sizeof(int) * (size_t)P.nlevels * cap, h->stream));
    B.errFlag = h->dErrFlag;
    ORBX_CUDA_TRY(cudaMalloc(&h->dTabs, sizeof(short) * std::max<size_t>(tabs.size(), 2)));
    ORBX_CUDA_TRY(cudaMemcpyAsync(h->dTabs, tabs.data(), sizeof(short) * tabs.size(), cudaMemcpyHostToDevice, h->stream));
    // xw / yw hold 2 shorts per destination column / row, indexed by xTab + dx / yTab + dy
    ORBX_CUDA_TRY(cudaMalloc(&h->dFastUnits, sizeof(FastUnit) * std::max<size_t>(ti.units.size(), 1)));
    ORBX_CUDA_TRY(cudaMemcpyAsync(h->dFastUnits, ti.units.data(), sizeof(FastUnit) * ti.units.size(), cudaMemcpyHostToDevice, h->stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(h->stream));  // ti.units is a local
    B.fastUnits = h->dFastUnits;
    ORBX_CUDA_TRY(cudaMalloc(&h->dZeros, (size_t)P.fastTileBytes));
    ORBX_CUDA_TRY(cudaMemsetAsync(h->dZeros, 0, (size_t)P.fastTileBytes, h->stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(h->stream));
    B.zeros = h->dZeros;
    B.xofs = h->dTabs;
    B.yofs0 = h->dTabs + ti.nx;
    B.yofs1 = h->dTabs + ti.nx + ti.ny;
    B.xw = reinterpret_cast<const short2*>(h->dTabs + ti.nx + 2 * ti.ny);
    B.yw = reinterpret_cast<const short2*>(h->dTabs + ti.nx + 2 * ti.ny + ti.nxw);
    B.yrow = reinterpret_cast<const int4*>(h->dTabs + ti.yrowOff);
    B.srow = reinterpret_cast<const int2*>(h->dTabs + ti.srowOff);
    // keys in shared memory when they fit in what is left, capped so that 1024 threads' worth of CTAs stay resident per SM
    // (the register file allows about that many at 58 registers per thread)
    h->octNodeCap = nodeCapMax;
    for (int v = 0; v < 3; v++) {
        const int T = 128 << v;
        const size_t ctasPerSM = 1024 / T;
        const size_t perCta = (size_t)(228 * 1024) / ctasPerSM - 1024;
        size_t keyBudget = perCta > nodeBytes + 4096 ? perCta - nodeBytes - 1024 : 4096;
        keyBudget = std::min<size_t>(std::min<size_t>(keyBudget, (size_t)devSmem - nodeBytes - 1024), 64 * 1024);
        int keyCap = (int)std::min<size_t>(keyBudget / 2, (size_t)candCapMax);
        keyCap &= ~7;
        h->octKeySmemCap[v] = keyCap;
        h->octSmem[v] = nodeBytes + 2 * (size_t)keyCap + 16;
    }
    ORBX_CUDA_TRY(cudaFuncSetAttribute(k_octree<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->octSmem[0]));
    ORBX_CUDA_TRY(cudaFuncSetAttribute(k_octree<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->octSmem[1]));
    ORBX_CUDA_TRY(cudaFuncSetAttribute(k_octree<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->octSmem[2]));
    h->fastSmem = 2 * (size_t)P.fastTileBytes + 2 * ((size_t)FAST_Q1CAP + (size_t)P.fastQCap);
    ORBX_CUDA_TRY(cudaFuncSetAttribute(k_fast_band<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->fastSmem));
    ORBX_CUDA_TRY(cudaFuncSetAttribute(k_fast_band<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->fastSmem));
    // tensor maps for the TMA-staged FAST bands: level l of the pyramid arena as a [frame][row][byte] tensor, box = FAST_P
    // bytes x (hCell + 6) rows x 1 frame, no swizzle, out-of-range bytes read as zero
    static const bool tmaOff = [] { const char* e = std::getenv("ORBX_FAST_TMA"); return e && std::atoi(e) == 0; }();
    h->fastTma = false;
    h->blurTma = false;
    h->resizeTma = false;
    if (!tmaOff && P.nlevels <= FAST_TMA_LEVELS) {
        typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                     const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
        void* fp = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess && fp) {
            bool ok = true;
            for (int l = 0; l < P.nlevels && ok; l++) {
                const LevelGeom& g = P.lv[l];
                const cuuint64_t gdim[3] = {(cuuint64_t)g.pitch, (cuuint64_t)g.planeRows, (cuuint64_t)cap};
                const cuuint64_t gstride[2] = {(cuuint64_t)g.pitch, (cuuint64_t)g.planeBytes};
                const cuuint32_t box[3] = {(cuuint32_t)FAST_P, (cuuint32_t)(g.hCell + 6), 1};
                const cuuint32_t estr[3] = {1, 1, 1};
                ok = g.hCell + 6 <= 256 &&
                     ((EncodeFn)fp)(&h->fastMaps.m[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, B.pyr + g.planeOff, gdim, gstride, box, estr,
                                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
            }
            h->fastTma = ok;
            static const bool blurOff = [] { const char* e = std::getenv("ORBX_BLUR_TMA"); return e && std::atoi(e) == 0; }();
            bool okb = !blurOff;
            for (int l = 0; l < P.nlevels && okb; l++) {
                const LevelGeom& g = P.lv[l];
                const cuuint64_t gdim[3] = {(cuuint64_t)g.pitch, (cuuint64_t)g.planeRows, (cuuint64_t)cap};
                const cuuint64_t gstride[2] = {(cuuint64_t)g.pitch, (cuuint64_t)g.planeBytes};
                const cuuint32_t box[3] = {(cuuint32_t)BLUR_TP, (cuuint32_t)BLUR_TROWS, 1};
                const cuuint32_t estr[3] = {1, 1, 1};
                okb = ((EncodeFn)fp)(&h->blurMaps.m[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, B.pyr + g.planeOff, gdim, gstride, box, estr,
                                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
            }
            static const bool rzOff = [] { const char* e = std::getenv("ORBX_RESIZE_TMA"); return e && std::atoi(e) == 0; }();
            bool okr = !rzOff;
            for (int l = 0; l + 1 < P.nlevels && okr; l++) {
                const LevelGeom& g = P.lv[l];
                const cuuint64_t gdim[3] = {(cuuint64_t)g.pitch, (cuuint64_t)g.planeRows, (cuuint64_t)cap};
                const cuuint64_t gstride[2] = {(cuuint64_t)g.pitch, (cuuint64_t)g.planeBytes};
                const cuuint32_t box[3] = {(cuuint32_t)RZ_TP, (cuuint32_t)RZ_TROWS, 1};
                const cuuint32_t estr[3] = {1, 1, 1};
                if (!P.lv[l + 1].resizeTma) { std::memset(&h->resizeMaps.m[l], 0, sizeof(CUtensorMap)); continue; }
                okr = ((EncodeFn)fp)(&h->resizeMaps.m[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, B.pyr + g.planeOff, gdim, gstride, box, estr,
                                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
            }
            h->resizeTma = okr;
            if (okr) ORBX_CUDA_TRY(cudaFuncSetAttribute(k_resize_rows_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, RZ_TROWS * RZ_TP));
            h->blurTma = okb;
            if (okb) ORBX_CUDA_TRY(cudaFuncSetAttribute(k_blur_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, BLUR_TROWS * BLUR_TP));
        }
    }
    ORBX_CUDA_TRY(cudaMalloc(&h->dKeyNode, sizeof(unsigned short) * (size_t)P.candFrame * cap));
    h->plan = P;
    h->frameCap = cap;
    h->havePlan = true;
    return ORBX_OK;
}

// Kernel launch with (pdl) or without the programmatic-stream-serialization attribute (see pdl_trigger / pdl_wait).
template <typename... KArgs, typename... Args>
static cudaError_t launch_k(bool pdl, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(std::forward<Args>(args))...);
}

// Enqueue every kernel of one batch on h->stream (+ the blur branch on h->streamBlur).
// slot 0 = the handle's main stream; slot 1 = a second compute lane (own streams / events, arena frames offset by
// frameBase) used by the pipelined host path so that consecutive chunks overlap.
static int enqueue(orbx_handle* h, const uint8_t* dIn, long long inStep, long long inFrameStride, int n,
                   orbx_keypoint* dKps, uint8_t* dDesc, int* dCounts, int capacity, bool profile, int* launches,
                   int slot = 0, int frameBase = 0)
{
    const Plan& P = h->plan;
    Buffers B = h->buf;
    B.frameBase = frameBase;
    cudaStream_t s = slot ? h->stream2 : h->stream, sb = profile ? s : (slot ? h->streamBlur2 : h->streamBlur);
    cudaEvent_t evFork = slot ? h->evFork2 : h->evFork, evJoin = slot ? h->evJoin2 : h->evJoin;
    int nl = 0;
    // ORBX_PDL=1 / 0 forces programmatic dependent launch on / off; default off: measured 287 k against 290 k frames/s with the
    // early trigger (the successor's CTAs take SM slots while they wait) -- see profiles/r2_state.md
    static const int pdlEnv = [] { const char* e = std::getenv("ORBX_PDL"); return e ? std::atoi(e) : -1; }();
    const bool pdl = (pdlEnv >= 0 ? pdlEnv != 0 : false) && !profile;
    // ORBX_DEBUG_SYNC=1 (never during capture: use with the graph disabled): synchronise after every launch and name the
    // kernel that failed
    static const bool dbgSync = [] { const char* e = std::getenv("ORBX_DEBUG_SYNC"); return e && std::atoi(e) != 0; }();
#define ORBX_DBG_SYNC(name)                                                                                   \
    do {                                                                                                      \
        if (dbgSync) {                                                                                        \
            cudaError_t _e = cudaStreamSynchronize(s);                                                        \
            if (_e == cudaSuccess) _e = cudaStreamSynchronize(sb);                                            \
            if (_e != cudaSuccess) { set_error("%s failed: %s", name, cudaGetErrorString(_e)); return ORBX_ERR_CUDA; } \
        }                                                                                                     \
    } while (0)
    if (profile) ORBX_CUDA_TRY(cudaEventRecord(h->evStage[ORBX_STAGE_PYRAMID], s));
    {
        const LevelGeom& g = P.lv[0];
        dim3 grid((g.planeRows + L0_ROWS - 1) / L0_ROWS, n);
        ORBX_CUDA_TRY(launch_k(pdl, k_level0, grid, dim3(16 * L0_ROWS), 0, s, P, B, dIn, inStep, inFrameStride));
        nl++;
    }
    for (int l = 1; l < P.nlevels; l++) {
        const LevelGeom& g = P.lv[l];
        // enough threads to fill the GPU: 16 rows per thread when the level is large, down to 4 for small grids
        int rpt = 16;
        while (rpt > 4 && (long long)(g.pitch / 4) * ((g.h + rpt - 1) / rpt) * n < 148LL * 2048) rpt >>= 1;
        dim3 grid(g.pitch / 128, (g.h + 4 * rpt - 1) / (4 * rpt), n);
        static const bool rowsOff = [] { const char* e = std::getenv("ORBX_RESIZE_ROWS"); return e && std::atoi(e) == 0; }();  // A/B switch
        if (h->resizeTma && g.resizeTma && rpt == RZ_RPT && !rowsOff)
            ORBX_CUDA_TRY(launch_k(pdl, k_resize_rows_tma, grid, dim3(128), RZ_TROWS * RZ_TP, s, P, B, l, g.pitch, h->resizeMaps));
        else if (g.resizeWords && g.resizeRows && !rowsOff) ORBX_CUDA_TRY(launch_k(pdl, k_resize_rows, grid, dim3(128), 0, s, P, B, l, rpt, P.lv[l - 1].pitch, g.pitch));
        else if (g.resizeWords) k_resize_words<<<grid, 128, 0, s>>>(P, B, l, rpt);
        else k_resize<<<grid, 128, 0, s>>>(P, B, l, rpt);
        nl++;
    }
    ORBX_DBG_SYNC("pyramid kernels");
    if (profile) ORBX_CUDA_TRY(cudaEventRecord(h->evStage[ORBX_STAGE_FAST], s));
    // The blur runs on its own branch of the graph.  A few frames: beside FAST (both grids fit the GPU at once, lowest
    // latency).  Batches: beside the quad-tree, whose barrier-bound CTAs leave most issue slots free (measured +2 % per
    // step against FAST || blur, where both kernels compete for the same issue slots).
    static const int blurFork = [] { const char* e = std::getenv("ORBX_BLUR_FORK"); return e ? std::atoi(e) : -1; }();  // experiments: 0 = beside FAST, 1 = after FAST
    const bool blurAfterFast = blurFork >= 0 ? blurFork != 0 : n >= 16;
    if (!profile && !blurAfterFast) {
        ORBX_CUDA_TRY(cudaEventRecord(evFork, s));
        ORBX_CUDA_TRY(cudaStreamWaitEvent(sb, evFork, 0));
    }
    if (h->fastTma) ORBX_CUDA_TRY(launch_k(pdl, k_fast_band<true>, dim3(P.totalFastUnits, n), dim3(FAST_T), h->fastSmem, s, P, B, h->fastMaps));
    else ORBX_CUDA_TRY(launch_k(pdl, k_fast_band<false>, dim3(P.totalFastUnits, n), dim3(FAST_T), h->fastSmem, s, P, B, h->fastMaps));
    nl++;
    ORBX_DBG_SYNC(h->fastTma ? "k_fast_band<TMA>" : "k_fast_band<cp.async>");
    if (!profile && blurAfterFast) {
        ORBX_CUDA_TRY(cudaEventRecord(evFork, s));
        ORBX_CUDA_TRY(cudaStreamWaitEvent(sb, evFork, 0));
    }
    if (profile) ORBX_CUDA_TRY(cudaEventRecord(h->evStage[ORBX_STAGE_OCTREE], s));
    // large batches: many small CTAs per SM (throughput); up to a few dozen frames: more threads per level (latency)
    static const int octForce = [] { const char* e = std::getenv("ORBX_OCT_T"); return e ? std::atoi(e) : 0; }();  // experiments only
    const int octT = octForce ? octForce : (n >= 112 ? 128 : (n >= 48 ? 256 : 512));  // measured: tools/sweep_octree.sh, profiles/r1h_sweep_octree.txt
    if (octT == 128) ORBX_CUDA_TRY(launch_k(pdl, k_octree<128>, dim3(P.nlevels, n), dim3(128), h->octSmem[0], s, P, B, h->octNodeCap, h->octKeySmemCap[0], h->dKeyNode, P.candFrame));
    else if (octT == 256) ORBX_CUDA_TRY(launch_k(pdl, k_octree<256>, dim3(P.nlevels, n), dim3(256), h->octSmem[1], s, P, B, h->octNodeCap, h->octKeySmemCap[1], h->dKeyNode, P.candFrame));
    else ORBX_CUDA_TRY(launch_k(pdl, k_octree<512>, dim3(P.nlevels, n), dim3(512), h->octSmem[2], s, P, B, h->octNodeCap, h->octKeySmemCap[2], h->dKeyNode, P.candFrame));
    nl++;
    if (profile) ORBX_CUDA_TRY(cudaEventRecord(h->evStage[ORBX_STAGE_BLUR], s));
    if (h->blurTma) ORBX_CUDA_TRY(launch_k(pdl, k_blur_tma, dim3(P.totalBlurUnits, n), dim3(32 * BT_WARPS), BLUR_TROWS * BLUR_TP, sb, P, B, h->blurMaps));
    else ORBX_CUDA_TRY(launch_k(pdl, k_blur, dim3(P.totalBlurUnits, n), dim3(128), 0, sb, P, B));
    nl++;
    if (!profile) {
        ORBX_CUDA_TRY(cudaEventRecord(evJoin, sb));
        ORBX_CUDA_TRY(cudaStreamWaitEvent(s, evJoin, 0));
    }
    if (profile) ORBX_CUDA_TRY(cudaEventRecord(h->evStage[ORBX_STAGE_DESC], s));
    {
        // keypoints per warp: the per-CTA constants are amortised over more keypoints when the batch alone fills the GPU
        const int perWarp = n >= 16 ? 12 : (n >= 4 ? 6 : 2);
        ORBX_CUDA_TRY(launch_k(pdl, k_describe, dim3((capacity + perWarp * DESC_WARPS - 1) / (perWarp * DESC_WARPS), n), dim3(DESC_WARPS * 32), 0, s, P, B, dKps, dDesc, dCounts, capacity));
    }
    nl++;
    if (profile) ORBX_CUDA_TRY(cudaEventRecord(h->evStage[ORBX_STAGE_D2H], s));
    ORBX_CUDA_TRY(cudaGetLastError());
    if (launches) *launches = nl;
    return ORBX_OK;
}

// Launch the kernel sequence of one batch through a cached CUDA graph (captured on first use of a
// (pointers, batch) combination; the first batch on a fresh plan runs with plain launches so that kernels are loaded
// outside any capture).
static const size_t GRAPH_CACHE = 24;

static int launch_cached(orbx_handle* h, const uint8_t* dIn, long long inStep, long long inFrameStride, int n, orbx_keypoint* dKps,
                         uint8_t* dDesc, int* dCounts, int capacity, int* launches, int slot = 0, int frameBase = 0)
{
    cudaStream_t cs = slot ? h->stream2 : h->stream;
    if (h->plainRuns == 0 || !h->useGraph) {
        std::lock_guard<std::mutex> setupLock(g_setupMutex);
        h->plainRuns = 1;
        return enqueue(h, dIn, inStep, inFrameStride, n, dKps, dDesc, dCounts, capacity, false, launches, slot, frameBase);
    }
    orbx_handle::GraphKey k;
    memset(&k, 0, sizeof k);
    k.in = dIn; k.step = inStep; k.fstride = inFrameStride; k.kps = dKps; k.desc = dDesc; k.counts = dCounts; k.n = n; k.cap = capacity; k.slot = slot; k.frameBase = frameBase;
    orbx_handle::GraphEntry* hit = nullptr;
    for (auto& g : h->graphs)
        if (memcmp(&g.key, &k, sizeof k) == 0) { hit = &g; break; }
    if (!hit) {
        std::lock_guard<std::mutex> setupLock(g_setupMutex);
        if (h->graphs.size() >= GRAPH_CACHE) {
            size_t victim = 0;
            for (size_t i = 1; i < h->graphs.size(); i++)
                if (h->graphs[i].stamp < h->graphs[victim].stamp) victim = i;
            cudaGraphExecDestroy(h->graphs[victim].exec);
            h->graphs.erase(h->graphs.begin() + victim);
        }
        cudaGraph_t graph = nullptr;
        int nl = 0;
        ORBX_CUDA_TRY(cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
        const int rc = enqueue(h, dIn, inStep, inFrameStride, n, dKps, dDesc, dCounts, capacity, false, &nl, slot, frameBase);
        cudaError_t e = cudaStreamEndCapture(cs, &graph);
        if (rc != ORBX_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
        if (e != cudaSuccess) { set_error("cudaStreamEndCapture: %s", cudaGetErrorString(e)); return ORBX_ERR_CUDA; }
        cudaGraphExec_t exec = nullptr;
        e = cudaGraphInstantiate(&exec, graph, 0);
        cudaGraphDestroy(graph);
        if (e != cudaSuccess) { set_error("cudaGraphInstantiate: %s", cudaGetErrorString(e)); return ORBX_ERR_CUDA; }
        h->graphs.push_back(orbx_handle::GraphEntry{k, exec, nl, 0});
        hit = &h->graphs.back();
    }
    hit->stamp = ++h->graphStamp;
    ORBX_CUDA_TRY(cudaGraphLaunch(hit->exec, cs));
    if (launches) *launches = hit->launches;
    return ORBX_OK;
}

namespace orbx {
int pyramid_view(orbx_handle* h, int frame, PyramidView* out)
{
    if (!h || !out || !h->havePlan || frame < 0 || frame >= h->lastFrames) { set_error("pyramid_view: nothing extracted / bad frame"); return ORBX_ERR_BAD_ARG; }
    const int batchFrame = frame;
    frame = h->arena_frame(batchFrame);
    if (frame < 0) { set_error("pyramid_view: frame %d of the last batch is no longer resident (pipelined host batch keeps its last two chunks)", batchFrame); return ORBX_ERR_BAD_ARG; }
    out->nlevels = h->cfg.nlevels;
    out->device = h->device;
    for (int l = 0; l < h->cfg.nlevels; l++) {
        const LevelGeom& g = h->plan.lv[l];
        out->interior[l] = h->buf.pyr + g.planeOff + (long long)frame * g.planeBytes + (long long)PAD_TOP * g.pitch + PAD_LEFT;
        out->pitch[l] = g.pitch;
        out->w[l] = g.w; out->h[l] = g.h;
        out->scale[l] = h->scale[l]; out->invScale[l] = h->invScale[l];
    }
    return ORBX_OK;
}
}  // namespace orbx

// ------------------------------------------------------------------------------------------------------
// Domain sweeps of the two transcendental device functions the descriptor path depends on (parity tests only):
// every launch folds a contiguous index range into one position-weighted 64-bit checksum per chunk, so that a test can
// compare ALL 1 135 869 953 float angles deg in [0, 360] (cosf/sinf of deg * factorPI, src/ORBextractor.cc:109,123-125)
// and >= 1e8 integer moment pairs (cv::fastAtan2((float)m01, (float)m10), :105) against the CPU oracle without moving
// the values.  checksum = sum_i bits(value_i) * (2 * i + 1) mod 2^64, i = index inside the chunk.
// ------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void dbg_atan2_pair(int mode, unsigned long long i, int& y, int& x)
{
    if (mode == 0) {  // dense grid [-H, H]^2, H = 5000, row-major
        const unsigned W = 10001u;
        y = (int)(i / W) - 5000;
        x = (int)(i % W) - 5000;
    } else {          // splitmix64(i) -> two signed 22-bit moments (|m| < 2^21 covers the 31x31 disc of u8 pixels)
        unsigned long long z = i + 0x9E3779B97F4A7C15ull * (unsigned long long)mode;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z ^= z >> 31;
        y = (int)(z & 0x3fffffu) - 0x200000;
        x = (int)((z >> 32) & 0x3fffffu) - 0x200000;
    }
}

__global__ void __launch_bounds__(256) k_dbg_sweep(int what, int mode, unsigned long long first, unsigned long long total,
                                                    unsigned long long chunk, unsigned long long* sumA,
                                                    unsigned long long* sumB)
{
    const unsigned long long c = blockIdx.x;
    const unsigned long long lo = c * chunk, hi = min(lo + chunk, total);
    unsigned long long a = 0, b = 0;
    const float factorPI = 0.0174532924f;
    for (unsigned long long i = lo + threadIdx.x; i < hi; i += blockDim.x) {
        const unsigned long long w = 2ull * (i - lo) + 1ull;
        if (what == 0) {
            const float deg = __uint_as_float((uint32_t)(first + i));
            const SinCosOut sc = sincosf_glibc(__fmul_rn(deg, factorPI));
            a += (unsigned long long)__float_as_uint(sc.s) * w;
            b += (unsigned long long)__float_as_uint(sc.c) * w;
        } else {
            int y, x;
            dbg_atan2_pair(mode, first + i, y, x);
            a += (unsigned long long)__float_as_uint(fast_atan2_dev((float)y, (float)x)) * w;
        }
    }
    __shared__ unsigned long long sa[256], sb[256];
    sa[threadIdx.x] = a; sb[threadIdx.x] = b;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) { sa[threadIdx.x] += sa[threadIdx.x + s]; sb[threadIdx.x] += sb[threadIdx.x + s]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { sumA[c] = sa[0]; if (sumB) sumB[c] = sb[0]; }
}

__global__ void k_dbg_values(int what, int mode, unsigned long long first, int n, float* outA, float* outB)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (what == 0) {
        const float deg = __uint_as_float((uint32_t)(first + i));
        const SinCosOut sc = sincosf_glibc(__fmul_rn(deg, 0.0174532924f));
        outA[i] = sc.s; outB[i] = sc.c;
    } else {
        int y, x;
        dbg_atan2_pair(mode, first + i, y, x);
        outA[i] = fast_atan2_dev((float)y, (float)x);
    }
}

static int dbg_sweep(int device, int what, int mode, unsigned long long first, unsigned long long total,
                     unsigned long long chunk, unsigned long long* sumA, unsigned long long* sumB)
{
    if (!sumA || chunk == 0 || (what == 0 && !sumB)) { set_error("orbx_debug sweep: bad argument"); return ORBX_ERR_BAD_ARG; }
    const unsigned long long nchunks = (total + chunk - 1) / chunk;
    if (nchunks == 0) return ORBX_OK;
    if (nchunks > 0x7fffffffull) { set_error("orbx_debug sweep: too many chunks"); return ORBX_ERR_BAD_ARG; }
    ORBX_CUDA_TRY(cudaSetDevice(device));
    unsigned long long *dA = nullptr, *dB = nullptr;
    ORBX_CUDA_TRY(cudaMalloc(&dA, nchunks * 8));
    ORBX_CUDA_TRY(cudaMalloc(&dB, nchunks * 8));
    k_dbg_sweep<<<(unsigned)nchunks, 256>>>(what, mode, first, total, chunk, dA, what == 0 ? dB : nullptr);
    cudaError_t e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaMemcpy(sumA, dA, nchunks * 8, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && what == 0) e = cudaMemcpy(sumB, dB, nchunks * 8, cudaMemcpyDeviceToHost);
    cudaFree(dA); cudaFree(dB);
    ORBX_CUDA_TRY(e);
    return ORBX_OK;
}

static int dbg_values(int device, int what, int mode, unsigned long long first, int n, float* outA, float* outB)
{
    if (n < 0 || !outA || (what == 0 && !outB)) { set_error("orbx_debug values: bad argument"); return ORBX_ERR_BAD_ARG; }
    if (n == 0) return ORBX_OK;
    ORBX_CUDA_TRY(cudaSetDevice(device));
    float *dA = nullptr, *dB = nullptr;
    ORBX_CUDA_TRY(cudaMalloc(&dA, (size_t)n * 4));
    ORBX_CUDA_TRY(cudaMalloc(&dB, (size_t)n * 4));
    k_dbg_values<<<(n + 255) / 256, 256>>>(what, mode, first, n, dA, dB);
    cudaError_t e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaMemcpy(outA, dA, (size_t)n * 4, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && what == 0) e = cudaMemcpy(outB, dB, (size_t)n * 4, cudaMemcpyDeviceToHost);
    cudaFree(dA); cudaFree(dB);
    ORBX_CUDA_TRY(e);
    return ORBX_OK;
}

extern "C" {

const char* orbx_last_error(void) { return last_error().c_str(); }

int orbx_device_count(void)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) { set_error("cudaGetDeviceCount: %s", cudaGetErrorString(e)); return ORBX_ERR_NO_DEVICE; }
    return n;
}

int orbx_create(const orbx_config* cfg, int device, orbx_handle** out)
{
    if (!cfg || !out) { set_error("orbx_create: null argument"); return ORBX_ERR_BAD_ARG; }
    if (cfg->nlevels < 1 || cfg->nlevels > ORBX_MAX_LEVELS || cfg->nfeatures < 0 || !(cfg->scale_factor > 1.0f) ||
        cfg->ini_th_fast < 0 || cfg->min_th_fast < 0 || cfg->ini_th_fast > 255 || cfg->min_th_fast > 255) {
        set_error("orbx_create: unsupported configuration (nlevels 1..%d, scale_factor > 1, thresholds 0..255)", ORBX_MAX_LEVELS);
        return ORBX_ERR_BAD_ARG;
    }
    int ndev = orbx_device_count();
    if (ndev <= 0) { if (ndev == 0) set_error("no CUDA device"); return ORBX_ERR_NO_DEVICE; }
    if (device < 0 || device >= ndev) { set_error("device %d out of range (%d devices)", device, ndev); return ORBX_ERR_BAD_ARG; }
    orbx_handle* h = new orbx_handle;
    h->cfg = *cfg;
    h->device = device;
    const int nl = cfg->nlevels;
    // src/ORBextractor.cc:445-500
    h->scale.resize(nl); h->invScale.resize(nl); h->sigma2.resize(nl); h->invSigma2.resize(nl); h->quota.resize(nl);
    h->scale[0] = 1.f; h->sigma2[0] = 1.f;
    // the reference keeps scaleFactor as a double member (include/ORBextractor.h): float * double -> float
    const double sfd = (double)cfg->scale_factor;
    for (int i = 1; i < nl; i++) { h->scale[i] = (float)(h->scale[i - 1] * sfd); h->sigma2[i] = h->scale[i] * h->scale[i]; }
    for (int i = 0; i < nl; i++) { h->invScale[i] = 1.f / h->scale[i]; h->invSigma2[i] = 1.f / h->sigma2[i]; }
    float factor = (float)(1.0f / sfd);
    float want = cfg->nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nl));
    int sum = 0;
    for (int l = 0; l < nl - 1; l++) { h->quota[l] = cv_round_f(want); sum += h->quota[l]; want *= factor; }
    h->quota[nl - 1] = std::max(cfg->nfeatures - sum, 0);
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->streamBlur, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->streamH2D, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->streamD2H, cudaStreamNonBlocking);
    for (int i = 0; i < orbx_handle::PIPE_IN_SLOTS && e == cudaSuccess; i++) {
        e = cudaEventCreateWithFlags(&h->evInReady[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->evInFree[i], cudaEventDisableTiming);
    }
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->stream2, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->streamBlur2, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->evFork2, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->evJoin2, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->evEntry, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaHostAlloc((void**)&h->hErrFlag, sizeof(int), cudaHostAllocMapped);
    if (e == cudaSuccess) { *h->hErrFlag = 0; e = cudaHostGetDevicePointer((void**)&h->dErrFlag, h->hErrFlag, 0); }
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->evFork, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->evJoin, cudaEventDisableTiming);
    for (int i = 0; i <= ORBX_NUM_STAGES && e == cudaSuccess; i++) e = cudaEventCreate(&h->evStage[i]);
    if (e != cudaSuccess) {
        set_error("orbx_create: %s", cudaGetErrorString(e));
        delete h;
        return ORBX_ERR_CUDA;
    }
    *out = h;
    return ORBX_OK;
}

void orbx_destroy(orbx_handle* h)
{
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->streamBlur) cudaStreamSynchronize(h->streamBlur);
    free_arenas(h);
    cudaFree(h->dIn); cudaFree(h->dOut);
    if (h->hOut) cudaFreeHost(h->hOut);
    for (auto& ev : h->evStage) if (ev) cudaEventDestroy(ev);
    if (h->evFork) cudaEventDestroy(h->evFork);
    if (h->evJoin) cudaEventDestroy(h->evJoin);
    if (h->stream) cudaStreamDestroy(h->stream);
    if (h->streamBlur) cudaStreamDestroy(h->streamBlur);
    if (h->streamH2D) cudaStreamDestroy(h->streamH2D);
    if (h->streamD2H) cudaStreamDestroy(h->streamD2H);
    for (int i = 0; i < orbx_handle::PIPE_IN_SLOTS; i++) { if (h->evInReady[i]) cudaEventDestroy(h->evInReady[i]); if (h->evInFree[i]) cudaEventDestroy(h->evInFree[i]); }
    for (auto& ev : h->evChunkDone) cudaEventDestroy(ev);
    if (h->evEntry) cudaEventDestroy(h->evEntry);
    if (h->stream2) { cudaStreamSynchronize(h->stream2); cudaStreamDestroy(h->stream2); }
    if (h->streamBlur2) cudaStreamDestroy(h->streamBlur2);
    if (h->evFork2) cudaEventDestroy(h->evFork2);
    if (h->evJoin2) cudaEventDestroy(h->evJoin2);
    if (h->hErrFlag) cudaFreeHost(h->hErrFlag);
    delete h;
}

int orbx_get_tables(const orbx_handle* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2, int* fpl)
{
    if (!h) { set_error("null handle"); return ORBX_ERR_BAD_ARG; }
    const int nl = h->cfg.nlevels;
    if (scale) std::copy(h->scale.begin(), h->scale.end(), scale);
    if (inv_scale) std::copy(h->invScale.begin(), h->invScale.end(), inv_scale);
    if (sigma2) std::copy(h->sigma2.begin(), h->sigma2.end(), sigma2);
    if (inv_sigma2) std::copy(h->invSigma2.begin(), h->invSigma2.end(), inv_sigma2);
    if (fpl) std::copy(h->quota.begin(), h->quota.end(), fpl);
    (void)nl;
    return ORBX_OK;
}

int orbx_max_keypoints(orbx_handle* h, int rows, int cols)
{
    if (!h || rows <= 0 || cols <= 0) { set_error("orbx_max_keypoints: bad argument"); return ORBX_ERR_BAD_ARG; }
    Plan P;
    std::vector<short> tabs;
    long long per = 0;
    int rc = build_plan(h, rows, cols, P, tabs, per);
    if (rc != ORBX_OK) return rc;
    return P.kpCapFrame;
}

int orbx_sync(orbx_handle* h)
{
    if (!h) { set_error("null handle"); return ORBX_ERR_BAD_ARG; }
    ORBX_CUDA_TRY(cudaSetDevice(h->device));
    ORBX_CUDA_TRY(cudaStreamSynchronize(h->stream));
    if (*h->hErrFlag) { *h->hErrFlag = 0; set_error("device reported an overflow: candidate list full, or more keypoints than the output capacity"); return ORBX_ERR_CAPACITY; }
    return ORBX_OK;
}

void* orbx_stream(orbx_handle* h) { return h ? (void*)h->stream : nullptr; }

int orbx_extract_batch_device(orbx_handle* h, const uint8_t* d_imgs, int n, int rows, int cols, size_t step,
                              size_t frame_stride, orbx_keypoint* d_kps, uint8_t* d_desc, int capacity, int* d_counts)
{
    if (!h || !d_imgs || !d_kps || !d_desc || !d_counts || n <= 0 || rows <= 0 || cols <= 0 || step < (size_t)cols || capacity <= 0) {
        set_error("orbx_extract_batch_device: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    ORBX_CUDA_TRY(cudaSetDevice(h->device));
    int rc = ensure_plan(h, rows, cols, n);
    if (rc != ORBX_OK) return rc;
    h->lastFrames = n;
    h->resident[0] = {0, n, 0};
    h->resident[1] = {0, 0, 0};
    if (h->profiling) {
        rc = enqueue(h, d_imgs, (long long)step, (long long)frame_stride, n, d_kps, d_desc, d_counts, capacity, h->profiling, &h->lastLaunches);
        if (rc != ORBX_OK) return rc;
        if (h->profiling) {
            ORBX_CUDA_TRY(cudaStreamSynchronize(h->stream));
            for (int sgi = ORBX_STAGE_PYRAMID; sgi <= ORBX_STAGE_DESC; sgi++)
                ORBX_CUDA_TRY(cudaEventElapsedTime(&h->stageMs[sgi], h->evStage[sgi], h->evStage[sgi + 1]));
        }
        return ORBX_OK;
    }
    return launch_cached(h, d_imgs, (long long)step, (long long)frame_stride, n, d_kps, d_desc, d_counts, capacity, &h->lastLaunches);
}

// Host entry point.  Small batches: one H2D, one graph launch, one D2H.  Large batches are cut into chunks of
// PIPE_CHUNK frames and software-pipelined over three streams (H2D of chunk k+1 | kernels of chunk k | D2H + copy-out of
// chunk k-1), so that PCIe transfers, which dominate the end-to-end time, overlap the kernels.
static const int PIPE_CHUNK = [] { const char* e = std::getenv("ORBX_PIPE_CHUNK"); const int v = e ? std::atoi(e) : 0; return v > 0 ? v : 32; }();

static int copy_in(orbx_handle* h, uint8_t* dDst, const uint8_t* imgs, int n, int rows, int cols, size_t step, size_t frame_stride,
                   cudaStream_t st)
{
    if (step == (size_t)cols && frame_stride == (size_t)rows * cols) {
        ORBX_CUDA_TRY(cudaMemcpyAsync(dDst, imgs, (size_t)n * rows * cols, cudaMemcpyHostToDevice, st));
    } else {
        for (int i = 0; i < n; i++)
            ORBX_CUDA_TRY(cudaMemcpy2DAsync(dDst + (size_t)i * rows * cols, cols, imgs + (size_t)i * frame_stride, step, cols, rows,
                                            cudaMemcpyHostToDevice, st));
    }
    return ORBX_OK;
}

int orbx_extract_batch(orbx_handle* h, const uint8_t* imgs, int n, int rows, int cols, size_t step, size_t frame_stride,
                       orbx_keypoint* kps, uint8_t* desc, int capacity, int* counts)
{
    if (!h || !imgs || !kps || !desc || !counts || n <= 0 || rows <= 0 || cols <= 0 || step < (size_t)cols || capacity <= 0) {
        set_error("orbx_extract_batch: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    ORBX_CUDA_TRY(cudaSetDevice(h->device));
    const bool piped = !h->profiling && n > PIPE_CHUNK + PIPE_CHUNK / 2;
    // chunk schedule: a short first chunk (its H2D is exposed), then growing to PIPE_CHUNK (larger launches are more efficient)
    std::vector<int> chunkStart;
    if (piped) {
        int f = 0, sz = std::max(8, PIPE_CHUNK / 4);
        while (f < n) {
            chunkStart.push_back(f);
            f += std::min(sz, n - f);
            sz = std::min(PIPE_CHUNK, sz * 3);
        }
    }
    chunkStart.push_back(n);
    const int chunk = piped ? PIPE_CHUNK : n;
    int rc = ensure_plan(h, rows, cols, piped ? 2 * chunk : chunk);  // pipelined: two arena slots of `chunk` frames
    if (rc != ORBX_OK) return rc;
    if (capacity < h->plan.kpCapFrame) {
        set_error("orbx_extract_batch: capacity %d < orbx_max_keypoints() = %d", capacity, h->plan.kpCapFrame);
        return ORBX_ERR_CAPACITY;
    }
    const int devCap = h->plan.kpCapFrame;
    // device staging for input (double buffered when pipelined) and packed output [counts | kps | desc]
    const size_t frameBytes = (size_t)rows * cols;
    const size_t inBytes = frameBytes * chunk * (piped ? orbx_handle::PIPE_IN_SLOTS : 1);
    if (inBytes > h->dInBytes) {
        std::lock_guard<std::mutex> setupLock(g_setupMutex);
        ORBX_CUDA_TRY(cudaDeviceSynchronize());
        cudaFree(h->dIn); h->dIn = nullptr; h->dInBytes = 0;
        ORBX_CUDA_TRY(cudaMalloc(&h->dIn, inBytes));
        h->dInBytes = inBytes;
    }
    const size_t cntBytes = ((size_t)n * sizeof(int) + 255) & ~size_t(255);
    const size_t kpBytes = ((size_t)n * devCap * sizeof(orbx_keypoint) + 255) & ~size_t(255);
    const size_t dsBytes = (size_t)n * devCap * 32;
    const size_t outBytes = cntBytes + kpBytes + dsBytes;
    if (outBytes > h->dOutBytes) {
        std::lock_guard<std::mutex> setupLock(g_setupMutex);
        ORBX_CUDA_TRY(cudaDeviceSynchronize());
        cudaFree(h->dOut); h->dOut = nullptr; h->dOutBytes = 0;
        if (h->hOut) { cudaFreeHost(h->hOut); h->hOut = nullptr; }
        ORBX_CUDA_TRY(cudaMalloc(&h->dOut, outBytes));
        ORBX_CUDA_TRY(cudaMallocHost(&h->hOut, outBytes));
        h->dOutBytes = outBytes;
        h->hOutBytes = outBytes;
    }
    int* dCounts = (int*)h->dOut;
    orbx_keypoint* dKps = (orbx_keypoint*)((char*)h->dOut + cntBytes);
    uint8_t* dDesc = (uint8_t*)h->dOut + cntBytes + kpBytes;
    const int* hc = (const int*)h->hOut;
    const orbx_keypoint* hk = (const orbx_keypoint*)((const char*)h->hOut + cntBytes);
    const uint8_t* hd = (const uint8_t*)h->hOut + cntBytes + kpBytes;
    auto copy_out = [&](int f0, int f1) {
        for (int i = f0; i < f1; i++) {
            const int c = hc[i];
            counts[i] = c;
            memcpy(kps + (size_t)i * capacity, hk + (size_t)i * devCap, sizeof(orbx_keypoint) * c);
            memcpy(desc + (size_t)i * capacity * 32, hd + (size_t)i * devCap * 32, (size_t)32 * c);
        }
    };
    *h->hErrFlag = 0;
    // caller buffers in page-locked memory with the device layout (capacity == orbx_max_keypoints): results are DMA'd
    // straight into them, no staging copy on the host
    auto is_pinned = [](const void* p) {
        cudaPointerAttributes a;
        if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
        return a.type == cudaMemoryTypeHost;
    };
    const bool direct = capacity == devCap && is_pinned(kps) && is_pinned(desc) && is_pinned(counts);
    char* hCnt = direct ? (char*)counts : (char*)h->hOut;
    char* hKps = direct ? (char*)kps : (char*)h->hOut + cntBytes;
    char* hDesc = direct ? (char*)desc : (char*)h->hOut + cntBytes + kpBytes;

    if (!piped) {
        if (h->profiling) ORBX_CUDA_TRY(cudaEventRecord(h->evStage[ORBX_STAGE_H2D], h->stream));
        if ((rc = copy_in(h, h->dIn, imgs, n, rows, cols, step, frame_stride, h->stream)) != ORBX_OK) return rc;
        rc = orbx_extract_batch_device(h, h->dIn, n, rows, cols, (size_t)cols, frameBytes, dKps, dDesc, devCap, dCounts);
        if (rc != ORBX_OK) return rc;
        if (direct) {
            ORBX_CUDA_TRY(cudaMemcpyAsync(hCnt, dCounts, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, h->stream));
            ORBX_CUDA_TRY(cudaMemcpyAsync(hKps, dKps, sizeof(orbx_keypoint) * (size_t)n * devCap, cudaMemcpyDeviceToHost, h->stream));
            ORBX_CUDA_TRY(cudaMemcpyAsync(hDesc, dDesc, (size_t)32 * n * devCap, cudaMemcpyDeviceToHost, h->stream));
        } else {
            ORBX_CUDA_TRY(cudaMemcpyAsync(h->hOut, h->dOut, outBytes, cudaMemcpyDeviceToHost, h->stream));
        }
        if (h->profiling) ORBX_CUDA_TRY(cudaEventRecord(h->evStage[ORBX_NUM_STAGES], h->stream));
        ORBX_CUDA_TRY(cudaStreamSynchronize(h->stream));
        if (h->profiling) {
            ORBX_CUDA_TRY(cudaEventElapsedTime(&h->stageMs[ORBX_STAGE_H2D], h->evStage[ORBX_STAGE_H2D], h->evStage[ORBX_STAGE_PYRAMID]));
            ORBX_CUDA_TRY(cudaEventElapsedTime(&h->stageMs[ORBX_STAGE_D2H], h->evStage[ORBX_STAGE_D2H], h->evStage[ORBX_NUM_STAGES]));
        }
        if (!direct) copy_out(0, n);
    } else {
        const int nchunks = (int)chunkStart.size() - 1;
        while ((int)h->evChunkDone.size() < nchunks) {
            cudaEvent_t e;
            ORBX_CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            h->evChunkDone.push_back(e);
        }
        h->lastFrames = n;
        h->resident[0] = h->resident[1] = {0, 0, 0};
        int launches = 0;
        // ORBX_PIPE_TRACE=1: per-chunk device timeline (ms from the first H2D) on stderr -- a debugging aid
        static const bool trace = std::getenv("ORBX_PIPE_TRACE") != nullptr;
        std::vector<cudaEvent_t> tev;
        auto mark = [&](cudaStream_t st) {
            if (!trace) return;
            cudaEvent_t e;
            cudaEventCreate(&e);
            cudaEventRecord(e, st);
            tev.push_back(e);
        };
        // an error in the middle of the pipeline must not leave copies into the caller's buffers in flight
        auto drain = [&](int code) {
            cudaStreamSynchronize(h->streamH2D); cudaStreamSynchronize(h->stream); cudaStreamSynchronize(h->stream2);
            cudaStreamSynchronize(h->streamD2H);
            return code;
        };
        // work a caller enqueued on h->stream (orbx_extract_batch_device is asynchronous) may still use the arena and dIn:
        // the other three lanes start behind it
        ORBX_CUDA_TRY(cudaEventRecord(h->evEntry, h->stream));
        ORBX_CUDA_TRY(cudaStreamWaitEvent(h->streamH2D, h->evEntry, 0));
        ORBX_CUDA_TRY(cudaStreamWaitEvent(h->stream2, h->evEntry, 0));
        for (int k = 0; k < nchunks; k++) {
            const int f0 = chunkStart[k], nk = chunkStart[k + 1] - f0, b = k & 1;
            // the input staging buffer has more slots than the arena: only k_level0 reads it, so the H2D engine may run
            // several chunks ahead of the kernels instead of waiting for chunk k-2 to finish
            const int ib = k % orbx_handle::PIPE_IN_SLOTS;
            uint8_t* dInB = h->dIn + (size_t)ib * chunk * frameBytes;
            if (k >= orbx_handle::PIPE_IN_SLOTS) ORBX_CUDA_TRY(cudaStreamWaitEvent(h->streamH2D, h->evInFree[ib], 0));
            mark(h->streamH2D);
            if ((rc = copy_in(h, dInB, imgs + (size_t)f0 * frame_stride, nk, rows, cols, step, frame_stride, h->streamH2D)) != ORBX_OK) return drain(rc);
            ORBX_CUDA_TRY(cudaEventRecord(h->evInReady[ib], h->streamH2D));
            mark(h->streamH2D);
            cudaStream_t cs = b ? h->stream2 : h->stream;  // chunk k computes in arena slot b on compute lane b
            ORBX_CUDA_TRY(cudaStreamWaitEvent(cs, h->evInReady[ib], 0));
            mark(cs);
            int nl = 0;
            rc = launch_cached(h, dInB, (long long)cols, (long long)frameBytes, nk, dKps + (size_t)f0 * devCap,
                               dDesc + (size_t)f0 * devCap * 32, dCounts + f0, devCap, &nl, b, b * chunk);
            if (rc != ORBX_OK) return drain(rc);
            launches += nl;
            h->resident[b] = {f0, nk, b * chunk};
            ORBX_CUDA_TRY(cudaEventRecord(h->evInFree[ib], cs));
            mark(cs);
            ORBX_CUDA_TRY(cudaStreamWaitEvent(h->streamD2H, h->evInFree[ib], 0));
            static const bool dbgNoD2H = std::getenv("ORBX_PIPE_NO_D2H") != nullptr;  // timeline experiments only
            if (!dbgNoD2H) {
            ORBX_CUDA_TRY(cudaMemcpyAsync(hCnt + (size_t)f0 * sizeof(int), dCounts + f0, sizeof(int) * nk, cudaMemcpyDeviceToHost, h->streamD2H));
            ORBX_CUDA_TRY(cudaMemcpyAsync(hKps + (size_t)f0 * devCap * sizeof(orbx_keypoint), dKps + (size_t)f0 * devCap,
                                          sizeof(orbx_keypoint) * (size_t)nk * devCap, cudaMemcpyDeviceToHost, h->streamD2H));
            ORBX_CUDA_TRY(cudaMemcpyAsync(hDesc + (size_t)f0 * devCap * 32, dDesc + (size_t)f0 * devCap * 32,
                                          (size_t)32 * nk * devCap, cudaMemcpyDeviceToHost, h->streamD2H));
            }
            ORBX_CUDA_TRY(cudaEventRecord(h->evChunkDone[k], h->streamD2H));
            mark(h->streamD2H);
        }
        h->lastLaunches = launches;
        if (trace) {
            cudaDeviceSynchronize();
            for (int k = 0; k < nchunks; k++) {
                float t[5];
                for (int i = 0; i < 5; i++) cudaEventElapsedTime(&t[i], tev[0], tev[5 * k + i]);
                fprintf(stderr, "chunk %2d (%3d frames): h2d %.3f-%.3f  kernels %.3f-%.3f  d2h done %.3f\n", k, chunkStart[k + 1] - chunkStart[k],
                        t[0], t[1], t[2], t[3], t[4]);
            }
            for (cudaEvent_t e : tev) cudaEventDestroy(e);
        }
        if (direct) {
            ORBX_CUDA_TRY(cudaStreamSynchronize(h->streamD2H));
        } else {
            for (int k = 0; k < nchunks; k++) {
                ORBX_CUDA_TRY(cudaEventSynchronize(h->evChunkDone[k]));
                copy_out(chunkStart[k], chunkStart[k + 1]);
            }
        }
        ORBX_CUDA_TRY(cudaStreamSynchronize(h->stream));
        ORBX_CUDA_TRY(cudaStreamSynchronize(h->stream2));
    }
    if (*h->hErrFlag) { set_error("device reported an overflow: candidate list full, or more keypoints than the output capacity"); return ORBX_ERR_CAPACITY; }
    return ORBX_OK;
}

int orbx_extract(orbx_handle* h, const uint8_t* img, int rows, int cols, size_t step, orbx_keypoint* kps, uint8_t* desc,
                 int capacity, int* count)
{
    return orbx_extract_batch(h, img, 1, rows, cols, step, step * (size_t)rows, kps, desc, capacity, count);
}

int orbx_get_level_size(orbx_handle* h, int level, int* rows_l, int* cols_l)
{
    if (!h || !h->havePlan || level < 0 || level >= h->cfg.nlevels) { set_error("orbx_get_level_size: no plan / bad level"); return ORBX_ERR_BAD_ARG; }
    if (rows_l) *rows_l = h->plan.lv[level].h;
    if (cols_l) *cols_l = h->plan.lv[level].w;
    return ORBX_OK;
}

int orbx_get_pyramid_level(orbx_handle* h, int frame, int level, uint8_t* dst, size_t dst_step)
{
    if (!h || !h->havePlan || !dst || level < 0 || level >= h->cfg.nlevels || frame < 0 || frame >= h->lastFrames) {
        set_error("orbx_get_pyramid_level: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    {
        const int batchFrame = frame;
        frame = h->arena_frame(batchFrame);
        if (frame < 0) {
            set_error("orbx_get_pyramid_level: frame %d of the last batch is no longer resident (a pipelined host batch keeps its last two chunks)", batchFrame);
            return ORBX_ERR_BAD_ARG;
        }
    }
    const LevelGeom& g = h->plan.lv[level];
    if (dst_step < (size_t)(g.w + 2 * EDGE)) { set_error("dst_step too small"); return ORBX_ERR_CAPACITY; }
    ORBX_CUDA_TRY(cudaSetDevice(h->device));
    const uint8_t* src = h->buf.pyr + g.planeOff + (long long)frame * g.planeBytes + (PAD_LEFT - EDGE);
    ORBX_CUDA_TRY(cudaMemcpy2DAsync(dst, dst_step, src, g.pitch, g.w + 2 * EDGE, g.planeRows, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(h->stream));
    return ORBX_OK;
}

int orbx_get_pyramid_all(orbx_handle* h, int frame, uint8_t* dst, const size_t* level_offsets, const size_t* level_steps)
{
    if (!h || !h->havePlan || !dst || !level_offsets || !level_steps || frame < 0 || frame >= h->lastFrames) {
        set_error("orbx_get_pyramid_all: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    const int arenaFrame = h->arena_frame(frame);
    if (arenaFrame < 0) {
        set_error("orbx_get_pyramid_all: frame %d of the last batch is no longer resident (a pipelined host batch keeps its last two chunks)", frame);
        return ORBX_ERR_BAD_ARG;
    }
    for (int l = 0; l < h->cfg.nlevels; l++)
        if (level_steps[l] < (size_t)(h->plan.lv[l].w + 2 * EDGE)) { set_error("orbx_get_pyramid_all: step of level %d too small", l); return ORBX_ERR_CAPACITY; }
    ORBX_CUDA_TRY(cudaSetDevice(h->device));
    // all levels are enqueued back to back and the stream is waited for ONCE (orbx_get_pyramid_level pays one wait per level)
    for (int l = 0; l < h->cfg.nlevels; l++) {
        const LevelGeom& g = h->plan.lv[l];
        const uint8_t* src = h->buf.pyr + g.planeOff + (long long)arenaFrame * g.planeBytes + (PAD_LEFT - EDGE);
        ORBX_CUDA_TRY(cudaMemcpy2DAsync(dst + level_offsets[l], level_steps[l], src, g.pitch, g.w + 2 * EDGE, g.planeRows,
                                        cudaMemcpyDeviceToHost, h->stream));
    }
    ORBX_CUDA_TRY(cudaStreamSynchronize(h->stream));
    return ORBX_OK;
}

void* orbx_alloc_pinned(size_t bytes)
{
    void* p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) { set_error("cudaMallocHost(%zu) failed", bytes); cudaGetLastError(); return nullptr; }
    return p;
}
void orbx_free_pinned(void* p) { if (p) cudaFreeHost(p); }

int orbx_set_profiling(orbx_handle* h, int on)
{
    if (!h) return ORBX_ERR_BAD_ARG;
    h->profiling = on != 0;
    return ORBX_OK;
}
int orbx_set_use_graph(orbx_handle* h, int on)
{
    if (!h) return ORBX_ERR_BAD_ARG;
    h->useGraph = on != 0;
    return ORBX_OK;
}
int orbx_stage_times(orbx_handle* h, float* ms)
{
    if (!h || !ms) return ORBX_ERR_BAD_ARG;
    for (int i = 0; i < ORBX_NUM_STAGES; i++) ms[i] = h->stageMs[i];
    return ORBX_OK;
}
int orbx_last_launch_count(orbx_handle* h) { return h ? h->lastLaunches : ORBX_ERR_BAD_ARG; }

int orbx_debug_get_candidates(orbx_handle* h, int frame, int level, int* xyr, int cap, int* n)
{
    if (!h || !h->havePlan || !n || level < 0 || level >= h->cfg.nlevels || frame < 0 || frame >= h->lastFrames) {
        set_error("orbx_debug_get_candidates: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    {
        const int batchFrame = frame;
        frame = h->arena_frame(batchFrame);
        if (frame < 0) {
            set_error("orbx_debug_get_candidates: frame %d of the last batch is no longer resident (a pipelined host batch keeps its last two chunks)", batchFrame);
            return ORBX_ERR_BAD_ARG;
        }
    }
    ORBX_CUDA_TRY(cudaSetDevice(h->device));
    ORBX_CUDA_TRY(cudaStreamSynchronize(h->stream));
    int m = 0;
    ORBX_CUDA_TRY(cudaMemcpy(&m, h->buf.candCountDbg + frame * h->plan.nlevels + level, sizeof(int), cudaMemcpyDeviceToHost));
    *n = m;
    if (!xyr || cap <= 0) return ORBX_OK;
    const int k = std::min(m, cap);
    std::vector<Cand> tmp(k);
    const LevelGeom& g = h->plan.lv[level];
    if (k) ORBX_CUDA_TRY(cudaMemcpy(tmp.data(), h->buf.cand + (long long)frame * h->plan.candFrame + g.candOff, sizeof(Cand) * k, cudaMemcpyDeviceToHost));
    for (int i = 0; i < k; i++) {
        xyr[3 * i] = (int)(tmp[i].xy & 0xffff) + 16;
        xyr[3 * i + 1] = (int)(tmp[i].xy >> 16) + 16;
        xyr[3 * i + 2] = (int)(tmp[i].ro >> 24);
    }
    return ORBX_OK;
}

int orbx_debug_get_blurred(orbx_handle* h, int frame, int level, uint8_t* dst)
{
    if (!h || !h->havePlan || !dst || level < 0 || level >= h->cfg.nlevels || frame < 0 || frame >= h->lastFrames) {
        set_error("orbx_debug_get_blurred: bad argument");
        return ORBX_ERR_BAD_ARG;
    }
    {
        const int batchFrame = frame;
        frame = h->arena_frame(batchFrame);
        if (frame < 0) {
            set_error("orbx_debug_get_blurred: frame %d of the last batch is no longer resident (a pipelined host batch keeps its last two chunks)", batchFrame);
            return ORBX_ERR_BAD_ARG;
        }
    }
    const LevelGeom& g = h->plan.lv[level];
    ORBX_CUDA_TRY(cudaSetDevice(h->device));
    const uint8_t* src = h->buf.blur + g.planeOff + (long long)frame * g.planeBytes + (long long)PAD_TOP * g.pitch + PAD_LEFT;
    ORBX_CUDA_TRY(cudaMemcpy2DAsync(dst, g.w, src, g.pitch, g.w, g.h, cudaMemcpyDeviceToHost, h->stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(h->stream));
    return ORBX_OK;
}

int orbx_debug_sincosf_checksum(int device, uint32_t first_bits, uint64_t count, uint64_t chunk, uint64_t* sum_sin,
                                uint64_t* sum_cos)
{
    return dbg_sweep(device, 0, 0, first_bits, count, chunk, (unsigned long long*)sum_sin, (unsigned long long*)sum_cos);
}
int orbx_debug_sincosf(int device, uint32_t first_bits, int n, float* out_sin, float* out_cos)
{
    return dbg_values(device, 0, 0, first_bits, n, out_sin, out_cos);
}
int orbx_debug_fast_atan2_checksum(int device, int mode, uint64_t first, uint64_t count, uint64_t chunk, uint64_t* sum)
{
    return dbg_sweep(device, 1, mode, first, count, chunk, (unsigned long long*)sum, nullptr);
}
int orbx_debug_fast_atan2(int device, int mode, uint64_t first, int n, float* out)
{
    return dbg_values(device, 1, mode, first, n, out, nullptr);
}

}  // extern "C"
