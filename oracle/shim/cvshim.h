// oracle/shim/cvshim.h -- TEST INFRASTRUCTURE.
//
// A minimal stand-in for the slice of the OpenCV C++ API that the reference's
// src/ORBextractor.cc + include/ORBextractor.h use, so that those two files can be compiled
// UNMODIFIED from /root/reference (oracle/Makefile target `_ref`).  OpenCV C++ is not installed
// in this image; the five numeric primitives delegate to oracle/cvprim.h, which is pinned against
// cv2 4.13.0.  Only what the reference touches is provided; nothing here is product code.
//
// One deliberate hook: the reference sorts vector<pair<int, ExtractorNode*>> with std::sort
// (src/ORBextractor.cc:906), i.e. ties on node size are broken by HEAP ADDRESS, which makes its
// output irreproducible.  When ORBX_CANONICAL_TIEBREAK is defined (default for the
// `_ref` build) cv::sort below is selected by overload resolution for exactly that iterator type
// and performs a stable sort on the size only (SURVEY.md decision D1).  Build with
// -DORBX_FAITHFUL_POINTER_SORT to get the raw behaviour instead.
#pragma once
#include <cstdint>
#include <cstring>
#include <cassert>
#include <cmath>
#include <memory>
#include <vector>
#include <list>
#include <algorithm>
#include <utility>
#include "../cvprim.h"

#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5
#define CV_PI 3.1415926535897932384626433832795

typedef unsigned char uchar;

static inline int cvRound(double v) { return orbo::cv_round(v); }
static inline int cvRound(float v) { return orbo::cv_round(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { return orbo::cv_floor(v); }
static inline int cvFloor(float v) { return orbo::cv_floor(v); }
static inline int cvCeil(double v) { return orbo::cv_ceil(v); }
static inline int cvCeil(float v) { return orbo::cv_ceil(v); }

namespace ORB_SLAM2 { class ExtractorNode; }

namespace cv {

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
    template <typename U> Point_(const Point_<U>& o) : x((T)o.x), y((T)o.y) {}
    Point_& operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }
};
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};

struct Rect {
    int x, y, width, height;
    Rect(int _x, int _y, int w, int h) : x(_x), y(_y), width(w), height(h) {}
};

struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0,
             int _class_id = -1)
        : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
};

struct MatZeros { int rows, cols, type; };

// 8-bit single-channel matrix with shared storage and ROI views.
class Mat {
public:
    int rows, cols;
    size_t step;
    uchar* data;
    std::shared_ptr<std::vector<uchar>> buf;

    Mat() : rows(0), cols(0), step(0), data(nullptr) {}
    Mat(int r, int c, int type) { alloc(r, c, type); }
    Mat(Size sz, int type) { alloc(sz.height, sz.width, type); }
    // wrap external memory (not owned)
    Mat(int r, int c, int type, void* ext, size_t _step) : rows(r), cols(c), step(_step), data((uchar*)ext)
    {
        assert(type == CV_8UC1);
    }

    void create(int r, int c, int type)
    {
        if (data && r == rows && c == cols) return;
        alloc(r, c, type);
    }
    void release() { rows = cols = 0; step = 0; data = nullptr; buf.reset(); }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return CV_8UC1; }
    size_t step1() const { return step; }
    bool isContinuous() const { return step == (size_t)cols; }

    Mat operator()(const Rect& r) const
    {
        Mat m;
        m.rows = r.height; m.cols = r.width; m.step = step;
        m.data = data + (size_t)r.y * step + r.x;
        m.buf = buf;
        return m;
    }
    Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
    Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }
    Mat clone() const
    {
        Mat m(rows, cols, CV_8UC1);
        for (int y = 0; y < rows; y++) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, cols);
        return m;
    }
    template <typename T> T& at(int y, int x) { return *(T*)(data + (size_t)y * step + x * sizeof(T)); }
    template <typename T> const T& at(int y, int x) const { return *(const T*)(data + (size_t)y * step + x * sizeof(T)); }
    uchar* ptr(int y = 0) { return data + (size_t)y * step; }
    const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }
    template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
    template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step); }

    static MatZeros zeros(int r, int c, int type) { return MatZeros{r, c, type}; }
    // MatExpr assignment semantics: when size/type already match, fill in place (keeps aliasing)
    Mat& operator=(const MatZeros& z)
    {
        create(z.rows, z.cols, z.type);
        for (int y = 0; y < rows; y++) std::memset(data + (size_t)y * step, 0, cols);
        return *this;
    }

private:
    void alloc(int r, int c, int type)
    {
        assert(type == CV_8UC1);
        (void)type;
        rows = r; cols = c; step = (size_t)c;
        buf = std::make_shared<std::vector<uchar>>((size_t)r * c + 1);
        data = buf->data();
    }
};

// InputArray / OutputArray: thin handles onto a Mat.
class _InputArray {
public:
    _InputArray() : m(nullptr) {}
    _InputArray(const Mat& _m) : m(const_cast<Mat*>(&_m)) {}
    Mat getMat() const { return m ? *m : Mat(); }
    bool empty() const { return !m || m->empty(); }
protected:
    Mat* m;
};
class _OutputArray : public _InputArray {
public:
    _OutputArray() {}
    _OutputArray(Mat& _m) : _InputArray(_m) {}
    void create(int r, int c, int type) const { m->create(r, c, type); }
    void release() const { m->release(); }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

enum { INTER_NEAREST = 0, INTER_LINEAR = 1 };
enum { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3, BORDER_REFLECT_101 = 4,
       BORDER_ISOLATED = 16 };

inline void resize(const Mat& src, Mat& dst, Size dsize, double = 0, double = 0, int interpolation = INTER_LINEAR)
{
    assert(interpolation == INTER_LINEAR);
    (void)interpolation;
    dst.create(dsize.height, dsize.width, CV_8UC1);
    orbo::resize_linear_u8(src.data, src.cols, src.rows, src.step, dst.data, dst.cols, dst.rows, dst.step);
}

// The reference only ever calls this with dst already sized (rows+top+bottom, cols+left+right) and, for
// levels >= 1, with src being the interior ROI of dst itself.  Input pixels outside src are never used
// (decision D6: the level-0 input is treated as isolated).
inline void copyMakeBorder(const Mat& src, Mat& dst, int top, int bottom, int left, int right, int borderType)
{
    assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101);
    assert(top == bottom && left == right && top == left);
    (void)borderType; (void)bottom; (void)right;
    dst.create(src.rows + 2 * top, src.cols + 2 * left, CV_8UC1);
    uchar* interior = dst.data + (size_t)top * dst.step + left;
    if (interior != src.data)
        for (int y = 0; y < src.rows; y++) std::memmove(interior + (size_t)y * dst.step, src.data + (size_t)y * src.step, src.cols);
    orbo::border_reflect101_inplace(interior, src.cols, src.rows, dst.step, top);
}

inline void FAST(const Mat& img, std::vector<KeyPoint>& kps, int threshold, bool nonmaxSuppression = true)
{
    std::vector<orbo::FastPoint> v;
    orbo::fast9_16(img.data, img.cols, img.rows, img.step, threshold, nonmaxSuppression, v);
    kps.clear();
    for (const auto& p : v) kps.push_back(KeyPoint((float)p.x, (float)p.y, 7.f, -1, (float)p.score));
}

inline void GaussianBlur(const Mat& src, Mat& dst, Size ksize, double sx, double sy = 0, int borderType = BORDER_REFLECT_101)
{
    assert(ksize.width == 7 && ksize.height == 7 && sx == 2 && sy == 2 && borderType == BORDER_REFLECT_101);
    (void)ksize; (void)sx; (void)sy; (void)borderType;
    dst.create(src.rows, src.cols, CV_8UC1);
    orbo::gaussian_blur_7x7_s2(src.data, src.cols, src.rows, src.step, dst.data, dst.step);
}

inline float fastAtan2(float y, float x) { return orbo::fast_atan2(y, x); }

// used only by the reference's dead ComputeKeyPointsOld (src/ORBextractor.cc:1168-1479); must compile.
struct KeyPointsFilter {
    static void retainBest(std::vector<KeyPoint>& kps, int n)
    {
        if (n >= 0 && (int)kps.size() > n) {
            std::stable_sort(kps.begin(), kps.end(), [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; });
            kps.resize(n);
        }
    }
};

#ifndef ORBX_FAITHFUL_POINTER_SORT
// Canonical tie-break hook (see header comment).  Non-template, so it wins overload resolution over
// std::sort for this exact iterator type; found through the reference's `using namespace cv`.
typedef std::vector<std::pair<int, ORB_SLAM2::ExtractorNode*>>::iterator OrbxNodeSizeIter;
inline void sort(OrbxNodeSizeIter a, OrbxNodeSizeIter b)
{
    std::stable_sort(a, b, [](const std::pair<int, ORB_SLAM2::ExtractorNode*>& l,
                              const std::pair<int, ORB_SLAM2::ExtractorNode*>& r) { return l.first < r.first; });
}
#endif

}  // namespace cv
