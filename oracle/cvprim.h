// oracle/cvprim.h -- TEST INFRASTRUCTURE (CPU oracle). Not part of the product path.
//
// Plain C++ restatements of the five OpenCV primitives the reference's ORB front-end calls.
// OpenCV itself is NOT in /root/reference (un-vendored, un-pinned: CMakeLists.txt:31-37,
// "Tested with OpenCV 2.4.11 and OpenCV 3.2" README.md:74).  The arithmetic below restates
// OpenCV's published algorithms and is pinned bit-for-bit against cv2 4.13.0 (the only OpenCV
// present in this image) by tests/test_oracle_primitives.py and the fixtures in tests/golden/.
//
//   resize_linear_u8      <- cv::resize(..., INTER_LINEAR) on CV_8UC1   (call site src/ORBextractor.cc:1656)
//   border_reflect101     <- cv::copyMakeBorder(..., BORDER_REFLECT_101) (call sites :1666, :1688)
//   fast9_16              <- cv::FAST(img, kps, threshold, true)         (call sites :1091, :1100)
//   gaussian_blur_7x7_s2  <- cv::GaussianBlur(m, m, Size(7,7), 2, 2, BORDER_REFLECT_101) (:1590)
//   fast_atan2            <- cv::fastAtan2(float y, float x)             (:105)
//   cv_round/cv_floor/cv_ceil <- cvRound / cvFloor / cvCeil
//
// Build with -ffp-contract=off: every float expression below is meant to be
// evaluated operation by operation, each rounded to nearest-even.
#pragma once
#include <cstdint>
#include <cstddef>
#include <cmath>
#include <cfloat>
#include <cstring>
#include <vector>
#include <algorithm>

namespace orbo {

// cvRound: round-half-to-even (SSE cvtss2si / lrint under the default rounding mode).
static inline int cv_round(float v) { return (int)lrintf(v); }
static inline int cv_round(double v) { return (int)lrint(v); }
static inline int cv_floor(float v) { int i = (int)v; return i - (v < (float)i); }
static inline int cv_floor(double v) { int i = (int)v; return i - (v < (double)i); }
static inline int cv_ceil(float v) { int i = (int)v; return i + (v > (float)i); }
static inline int cv_ceil(double v) { int i = (int)v; return i + (v > (double)i); }

static inline short sat_short_from_float(float v)
{
    int i = cv_round(v);
    return (short)(i < -32768 ? -32768 : (i > 32767 ? 32767 : i));
}

// BORDER_REFLECT_101 index map (gfedcb|abcdefgh|gfedcba), valid for any p once len > 1.
static inline int reflect101(int p, int len)
{
    if (len == 1) return 0;
    while (p < 0 || p >= len) {
        if (p < 0) p = -p;
        else p = 2 * (len - 1) - p;
    }
    return p;
}

// ---------------------------------------------------------------------------------------------
// cv::resize INTER_LINEAR, 8-bit single channel.  Fixed point with INTER_RESIZE_COEF_BITS = 11.
// Horizontal pass -> int32 (value * 2048 scale), vertical pass:
//   dst = ( ((b0 * (T0 >> 4)) >> 16) + ((b1 * (T1 >> 4)) >> 16) + 2 ) >> 2.
// Coefficients: fx = (float)((dx + 0.5) * scale_x - 0.5) with scale_x = 1 / (dw / sw) in double.
// ---------------------------------------------------------------------------------------------
struct LinearCoeffs {
    std::vector<int> ofs;      // source index of the first tap
    std::vector<short> w;      // 2 weights per destination sample (sum 2048)
};

static inline LinearCoeffs linear_coeffs(int ssize, int dsize)
{
    LinearCoeffs c;
    c.ofs.resize(dsize);
    c.w.resize(2 * (size_t)dsize);
    double inv_scale = (double)dsize / ssize;
    double scale = 1.0 / inv_scale;
    for (int d = 0; d < dsize; d++) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = cv_floor(f);
        f -= s;
        if (s < 0) { f = 0; s = 0; }
        if (s >= ssize - 1) { f = 0; s = ssize - 1; }
        c.ofs[d] = s;
        c.w[2 * d] = sat_short_from_float((1.f - f) * 2048);
        c.w[2 * d + 1] = sat_short_from_float(f * 2048);
    }
    return c;
}

static inline void resize_linear_u8(const uint8_t* src, int sw, int sh, size_t sstep,
                                    uint8_t* dst, int dw, int dh, size_t dstep)
{
    LinearCoeffs cx = linear_coeffs(sw, dw);
    // the vertical pass does not clamp the weight when the row index is clamped
    std::vector<int> yofs(dh);
    std::vector<short> beta(2 * (size_t)dh);
    {
        double inv_scale = (double)dh / sh;
        double scale = 1.0 / inv_scale;
        for (int d = 0; d < dh; d++) {
            float f = (float)((d + 0.5) * scale - 0.5);
            int s = cv_floor(f);
            f -= s;
            yofs[d] = s;
            beta[2 * d] = sat_short_from_float((1.f - f) * 2048);
            beta[2 * d + 1] = sat_short_from_float(f * 2048);
        }
    }
    std::vector<int> row0(dw), row1(dw);
    auto hpass = [&](int sy, std::vector<int>& out) {
        const uint8_t* S = src + (size_t)sy * sstep;
        for (int dx = 0; dx < dw; dx++) {
            int sx = cx.ofs[dx];
            if (sx + 1 < sw) out[dx] = S[sx] * cx.w[2 * dx] + S[sx + 1] * cx.w[2 * dx + 1];
            else out[dx] = S[sx] * 2048;
        }
    };
    auto clip = [](int x, int a, int b) { return x >= a ? (x < b ? x : b - 1) : a; };
    for (int dy = 0; dy < dh; dy++) {
        int sy0 = clip(yofs[dy], 0, sh), sy1 = clip(yofs[dy] + 1, 0, sh);
        hpass(sy0, row0);
        hpass(sy1, row1);
        int b0 = beta[2 * dy], b1 = beta[2 * dy + 1];
        uint8_t* D = dst + (size_t)dy * dstep;
        for (int dx = 0; dx < dw; dx++) {
            int v = (((b0 * (row0[dx] >> 4)) >> 16) + ((b1 * (row1[dx] >> 4)) >> 16) + 2) >> 2;
            D[dx] = (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
        }
    }
}

// Fill a `pad`-wide frame around the w x h interior that starts at `interior` (row stride `step`)
// with BORDER_REFLECT_101 copies of interior pixels (the memory around the interior must exist).
static inline void border_reflect101_inplace(uint8_t* interior, int w, int h, size_t step, int pad)
{
    for (int y = 0; y < h; y++) {
        uint8_t* r = interior + (ptrdiff_t)y * (ptrdiff_t)step;
        for (int i = 1; i <= pad; i++) {
            r[-i] = r[reflect101(-i, w)];
            r[w - 1 + i] = r[reflect101(w - 1 + i, w)];
        }
    }
    for (int i = 1; i <= pad; i++) {
        std::memcpy(interior + (ptrdiff_t)(-i) * (ptrdiff_t)step - pad,
                    interior + (ptrdiff_t)reflect101(-i, h) * (ptrdiff_t)step - pad, (size_t)w + 2 * pad);
        std::memcpy(interior + (ptrdiff_t)(h - 1 + i) * (ptrdiff_t)step - pad,
                    interior + (ptrdiff_t)reflect101(h - 1 + i, h) * (ptrdiff_t)step - pad, (size_t)w + 2 * pad);
    }
}

// ---------------------------------------------------------------------------------------------
// cv::FAST, TYPE_9_16.  Segment test on the 16-pixel Bresenham ring of radius 3.
//   d_k = I(p) - I(ring_k);  M(p) = max over the 16 contiguous 9-arcs of max(min d, min -d)
//   p is a corner at threshold t  <=>  M(p) > t;  score(p) = M(p) - 1  (independent of t)
// evaluated for x in [3, w-3), y in [3, h-3) only.  With NMS a corner survives iff its score is
// strictly greater than the stored score of all 8 neighbours, where non-corners and pixels outside
// the evaluated region store 0.  Output order: row-major.
// ---------------------------------------------------------------------------------------------
struct FastPoint { int x, y, score; };

static const int kFastRing[16][2] = {
    {0, 3}, {1, 3}, {2, 2}, {3, 1}, {3, 0}, {3, -1}, {2, -2}, {1, -3},
    {0, -3}, {-1, -3}, {-2, -2}, {-3, -1}, {-3, 0}, {-3, 1}, {-2, 2}, {-1, 3}};

// M(p) as defined above (can be <= 0 on flat patches).
static inline int fast_arc_measure(const uint8_t* p, ptrdiff_t step)
{
    int d[16];
    for (int k = 0; k < 16; k++) d[k] = (int)p[0] - (int)p[kFastRing[k][1] * step + kFastRing[k][0]];
    int best = -256;
    for (int s = 0; s < 16; s++) {
        int mn = 255, mx = -255;
        for (int k = 0; k < 9; k++) {
            int v = d[(s + k) & 15];
            mn = std::min(mn, v);
            mx = std::max(mx, v);
        }
        best = std::max(best, std::max(mn, -mx));
    }
    return best;
}

// plain transcription of the definition above; kept as the cross-check for the optimised fast9_16 below
static inline void fast9_16_simple(const uint8_t* img, int w, int h, size_t step, int threshold, bool nms,
                                   std::vector<FastPoint>& out)
{
    out.clear();
    threshold = std::min(std::max(threshold, 0), 255);
    if (w < 7 || h < 7) return;
    // stored score plane: 0 for non-corners / unevaluated pixels (what cv::FAST keeps in its row buffers)
    std::vector<int> sc((size_t)w * h, 0);
    std::vector<uint8_t> corner((size_t)w * h, 0);
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            int m = fast_arc_measure(img + (size_t)y * step + x, (ptrdiff_t)step);
            if (m > threshold) {
                corner[(size_t)y * w + x] = 1;
                sc[(size_t)y * w + x] = m - 1;
            }
        }
    for (int y = 3; y < h - 3; y++)
        for (int x = 3; x < w - 3; x++) {
            if (!corner[(size_t)y * w + x]) continue;
            int s = sc[(size_t)y * w + x];
            bool keep = true;
            if (nms) {
                for (int dy = -1; dy <= 1 && keep; dy++)
                    for (int dx = -1; dx <= 1; dx++) {
                        if (!dx && !dy) continue;
                        if (!(s > sc[(size_t)(y + dy) * w + (x + dx)])) { keep = false; break; }
                    }
            }
            if (keep) out.push_back({x, y, s});
        }
}

// Same result as fast9_16_simple, organised like cv::FAST itself so that the timed CPU baseline is not a strawman:
// high-speed rejection on the four compass pixels (every 9-arc contains at least two of them), sliding min/max by
// doubling for the arc measure, three rolling score rows for the 3x3 non-maximum suppression.
static inline int fast_arc_measure_fast(const uint8_t* p, const ptrdiff_t* off, int threshold)
{
    const int c = p[0];
    const int d0 = c - p[off[0]], d4 = c - p[off[4]], d8 = c - p[off[8]], d12 = c - p[off[12]];
    const int nb = (d0 > threshold) + (d4 > threshold) + (d8 > threshold) + (d12 > threshold);
    const int nd = (d0 < -threshold) + (d4 < -threshold) + (d8 < -threshold) + (d12 < -threshold);
    if (nb < 2 && nd < 2) return -256;
    int d[16], mn2[16], mx2[16], mn4[16], mx4[16];
    for (int k = 0; k < 16; k++) d[k] = c - p[off[k]];
    for (int k = 0; k < 16; k++) { mn2[k] = std::min(d[k], d[(k + 1) & 15]); mx2[k] = std::max(d[k], d[(k + 1) & 15]); }
    for (int k = 0; k < 16; k++) { mn4[k] = std::min(mn2[k], mn2[(k + 2) & 15]); mx4[k] = std::max(mx2[k], mx2[(k + 2) & 15]); }
    int a = -256, b = 256;
    for (int k = 0; k < 16; k++) {
        const int m8 = std::min(mn4[k], mn4[(k + 4) & 15]), x8 = std::max(mx4[k], mx4[(k + 4) & 15]);
        a = std::max(a, std::min(m8, d[(k + 8) & 15]));
        b = std::min(b, std::max(x8, d[(k + 8) & 15]));
    }
    return std::max(a, -b);
}

static inline void fast9_16(const uint8_t* img, int w, int h, size_t step, int threshold, bool nms,
                            std::vector<FastPoint>& out)
{
    out.clear();
    threshold = std::min(std::max(threshold, 0), 255);
    if (w < 7 || h < 7) return;
    ptrdiff_t off[16];
    for (int k = 0; k < 16; k++) off[k] = (ptrdiff_t)kFastRing[k][1] * (ptrdiff_t)step + kFastRing[k][0];
    thread_local std::vector<uint8_t> scoreRows;
    thread_local std::vector<int> cornerPos;
    scoreRows.assign((size_t)3 * w, 0);
    cornerPos.assign((size_t)3 * (w + 1), 0);
    uint8_t* buf[3] = {scoreRows.data(), scoreRows.data() + w, scoreRows.data() + 2 * (size_t)w};
    int* cpbuf[3] = {cornerPos.data() + 1, cornerPos.data() + (w + 1) + 1, cornerPos.data() + 2 * (size_t)(w + 1) + 1};
    for (int i = 3; i < h - 2; i++) {
        uint8_t* curr = buf[(i - 3) % 3];
        int* cpos = cpbuf[(i - 3) % 3];
        std::memset(curr, 0, (size_t)w);
        int ncorners = 0;
        if (i < h - 3) {
            const uint8_t* row = img + (size_t)i * step;
            for (int j = 3; j < w - 3; j++) {
                const int m = fast_arc_measure_fast(row + j, off, threshold);
                if (m > threshold) {
                    cpos[ncorners++] = j;
                    curr[j] = (uint8_t)(m - 1);
                }
            }
        }
        cpos[-1] = ncorners;
        if (i == 3) continue;
        const uint8_t* prev = buf[(i - 4 + 3) % 3];
        const uint8_t* pprev = buf[(i - 5 + 3) % 3];
        const int* ppos = cpbuf[(i - 4 + 3) % 3];
        const int np = ppos[-1];
        for (int k = 0; k < np; k++) {
            const int j = ppos[k];
            const int s = prev[j];
            if (!nms || (s > prev[j + 1] && s > prev[j - 1] && s > pprev[j - 1] && s > pprev[j] && s > pprev[j + 1] &&
                         s > curr[j - 1] && s > curr[j] && s > curr[j + 1]))
                out.push_back({j, i - 1, s});
        }
    }
}

// ---------------------------------------------------------------------------------------------
// cv::GaussianBlur 7x7, sigma 2, 8-bit, BORDER_REFLECT_101 -- OpenCV 4.x fixed-point path:
// taps (8.8 fixed point) {18,34,48,56,48,34,18}, sum 256.  h = sum T*src (16 bit), v = sum T*h
// (32 bit), dst = (v + 32768) >> 16.  `legacy` selects the OpenCV <= 3.2 separable-filter taps
// {18,34,49,55,49,34,18}/256 (not checkable in this image).
// src and dst may alias.
// ---------------------------------------------------------------------------------------------
static inline void gaussian_blur_7x7_s2(const uint8_t* src, int w, int h, size_t sstep,
                                        uint8_t* dst, size_t dstep)
{
    // horizontal pass into 16-bit rows (one padded source row at a time), vertical pass over 7 of them
    thread_local std::vector<uint16_t> hbuf;
    thread_local std::vector<uint8_t> prow;
    hbuf.resize((size_t)w * h);
    prow.resize((size_t)w + 6);
    for (int y = 0; y < h; y++) {
        const uint8_t* S = src + (size_t)y * sstep;
        for (int k = 0; k < 3; k++) { prow[k] = S[reflect101(k - 3, w)]; prow[w + 3 + k] = S[reflect101(w + k, w)]; }
        std::memcpy(prow.data() + 3, S, (size_t)w);
        uint16_t* H = &hbuf[(size_t)y * w];
        const uint8_t* P = prow.data();
        for (int x = 0; x < w; x++)
            H[x] = (uint16_t)(18 * (P[x] + P[x + 6]) + 34 * (P[x + 1] + P[x + 5]) + 48 * (P[x + 2] + P[x + 4]) + 56 * P[x + 3]);
    }
    thread_local std::vector<uint8_t> orow;
    orow.resize((size_t)w);
    for (int y = 0; y < h; y++) {
        const uint16_t* r[7];
        for (int k = 0; k < 7; k++) r[k] = &hbuf[(size_t)reflect101(y + k - 3, h) * w];
        for (int x = 0; x < w; x++) {
            const uint32_t a = 18u * (r[0][x] + r[6][x]) + 34u * (r[1][x] + r[5][x]) + 48u * (r[2][x] + r[4][x]) + 56u * r[3][x];
            orow[x] = (uint8_t)((a + 32768u) >> 16);
        }
        std::memcpy(dst + (size_t)y * dstep, orow.data(), (size_t)w);
    }
}

// ---------------------------------------------------------------------------------------------
// cv::fastAtan2(float y, float x), scalar path: degree-7 odd polynomial on min/max ratio,
// evaluated in float without FMA, result in degrees in [0, 360].
// ---------------------------------------------------------------------------------------------
static inline float fast_atan2(float y, float x)
{
    const float scale = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * scale;
    const float p3 = -0.3258083974640975f * scale;
    const float p5 = 0.1555786518463281f * scale;
    const float p7 = -0.04432655554792128f * scale;
    float ax = std::fabs(x), ay = std::fabs(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

// cv::undistortPoints(src, dst, K, distCoef, R = Mat(), P = K) as called by Frame::UndistortKeyPoints /
// ComputeImageBounds (src/Frame.cc:586-660): cvUndistortPointsInternal of OpenCV 4.x for that call (R = identity,
// P = K, default criteria = 5 iterations, no tilt terms); everything in double, K and the coefficients widened from
// float, result narrowed to float.  Pinned bit-for-bit against cv2 4.13 (tests/test_oracle_primitives.py,
// tests/golden/undistort_cv2.npz).
inline void undistort_points(const float* px, const float* py, int n, float fxf, float fyf, float cxf, float cyf,
                             const float* dist, int ndist, float* ox, float* oy)
{
    double k[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};  // k1 k2 p1 p2 k3 k4 k5 k6 s1 s2 s3 s4
    for (int i = 0; i < ndist && i < 12; i++) k[i] = (double)dist[i];
    const double fx = fxf, fy = fyf, cx = cxf, cy = cyf;
    const double ifx = 1. / fx, ify = 1. / fy;
    for (int i = 0; i < n; i++) {
        const double u = px[i], v = py[i];
        double x = (u - cx) * ifx, y = (v - cy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; j++) {
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
            if (icdist < 0) { x = (u - cx) * ifx; y = (v - cy) * ify; break; }
            const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
            const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
            x = (x0 - deltaX) * icdist;
            y = (y0 - deltaY) * icdist;
        }
        // P * R = K: xx = fx*x + 0*y + cx, yy = 0*x + fy*y + cy, ww = 1 / (0*x + 0*y + 1)
        ox[i] = (float)((fx * x + cx) * 1.0);
        oy[i] = (float)((fy * y + cy) * 1.0);
    }
}

}  // namespace orbo
