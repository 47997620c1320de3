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
#include <string>
#include <iostream>
#include <sstream>
#include <fstream>
#include <map>
#include <set>
#include <limits>
#include <cstdlib>
#include <climits>
#include "../cvprim.h"

#define CV_8U 0
#define CV_8UC1 0
#define CV_32S 4
#define CV_32F 5
#define CV_32FC1 5
#define CV_32FC2 13
#define CV_64F 6
#define CV_64FC1 6
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
    Rect() : x(0), y(0), width(0), height(0) {}
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

// Dense 2-D matrix with shared storage and ROI views.  Element type = CV_8U / CV_32F / CV_64F with 1..4 channels
// (type = depth + ((cn - 1) << 3), as in OpenCV).  Arithmetic is eager (no MatExpr): every operator returns a fresh
// matrix.  Float products accumulate in float in index order without contraction (-ffp-contract=off); see
// oracle/matchref_wrap.cc for how the matcher harness keeps the reference's pose arithmetic exact.
class Mat {
public:
    int rows, cols;
    size_t step;
    uchar* data;
    std::shared_ptr<std::vector<uchar>> buf;
    int flags_type;

    Mat() : rows(0), cols(0), step(0), data(nullptr), flags_type(CV_8UC1) {}
    Mat(int r, int c, int type) : data(nullptr) { alloc(r, c, type); }
    Mat(Size sz, int type) : data(nullptr) { alloc(sz.height, sz.width, type); }
    // wrap external memory (not owned)
    Mat(int r, int c, int type, void* ext, size_t _step = 0)
        : rows(r), cols(c), step(_step), data((uchar*)ext), flags_type(type)
    {
        if (step == 0) step = (size_t)c * elemSize();
    }

    static int depth_of(int type) { return type & 7; }
    static int channels_of(int type) { return (type >> 3) + 1; }
    static size_t depth_size(int depth) { return depth == CV_8U ? 1 : depth == CV_32F ? 4 : depth == CV_64F ? 8 : depth == CV_32S ? 4 : 2; }
    int type() const { return flags_type; }
    int depth() const { return depth_of(flags_type); }
    int channels() const { return channels_of(flags_type); }
    size_t elemSize() const { return depth_size(depth()) * channels(); }
    size_t elemSize1() const { return depth_size(depth()); }
    size_t total() const { return (size_t)rows * cols; }
    Size size() const { return Size(cols, rows); }

    void create(int r, int c, int type)
    {
        if (data && r == rows && c == cols && type == flags_type) return;
        alloc(r, c, type);
    }
    void create(Size sz, int type) { create(sz.height, sz.width, type); }
    void release() { rows = cols = 0; step = 0; data = nullptr; buf.reset(); }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    size_t step1() const { return step / elemSize1(); }
    bool isContinuous() const { return rows <= 1 || step == (size_t)cols * elemSize(); }

    Mat operator()(const Rect& r) const
    {
        Mat m;
        m.rows = r.height; m.cols = r.width; m.step = step; m.flags_type = flags_type;
        m.data = data + (size_t)r.y * step + (size_t)r.x * elemSize();
        m.buf = buf;
        return m;
    }
    Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
    Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }
    Mat row(int y) const { return rowRange(y, y + 1); }
    Mat col(int x) const { return colRange(x, x + 1); }
    Mat clone() const
    {
        Mat m;
        if (!data) return m;
        m.alloc(rows, cols, flags_type);
        for (int y = 0; y < rows; y++) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, (size_t)cols * elemSize());
        return m;
    }
    // copyTo: a destination of matching size/type is written in place (this is what makes
    // `R.copyTo(T.rowRange(0,3).colRange(0,3))` fill the view), anything else is reallocated.
    void copyTo(Mat& dst) const
    {
        if (empty()) { dst.release(); return; }
        dst.create(rows, cols, flags_type);
        for (int y = 0; y < rows; y++) std::memmove(dst.data + (size_t)y * dst.step, data + (size_t)y * step, (size_t)cols * elemSize());
    }
    void copyTo(const Mat& dst) const
    {
        assert(dst.rows == rows && dst.cols == cols && dst.flags_type == flags_type);
        for (int y = 0; y < rows; y++) std::memmove(dst.data + (size_t)y * dst.step, data + (size_t)y * step, (size_t)cols * elemSize());
    }
    // element access by value for the conversions below
    double get(int y, int x) const
    {
        const uchar* p = data + (size_t)y * step + (size_t)x * elemSize();
        switch (depth()) {
        case CV_8U: return *p;
        case CV_32F: { float v; std::memcpy(&v, p, 4); return v; }
        case CV_64F: { double v; std::memcpy(&v, p, 8); return v; }
        case CV_32S: { int v; std::memcpy(&v, p, 4); return v; }
        }
        return 0;
    }
    void convertTo(Mat& dst, int rtype) const
    {
        assert(channels() == 1);
        Mat out(rows, cols, rtype);
        for (int y = 0; y < rows; y++)
            for (int x = 0; x < cols; x++) {
                const double v = get(y, x);
                uchar* q = out.data + (size_t)y * out.step + (size_t)x * out.elemSize();
                switch (depth_of(rtype)) {
                case CV_8U: { int iv = (int)std::nearbyint(v); *q = (uchar)(iv < 0 ? 0 : iv > 255 ? 255 : iv); break; }
                case CV_32F: { float f = (float)v; std::memcpy(q, &f, 4); break; }
                case CV_64F: std::memcpy(q, &v, 8); break;
                default: assert(false);
                }
            }
        dst = out;
    }
    // reshape(cn): same storage, channel count changed, rows kept (continuous matrices only)
    Mat reshape(int cn, int = 0) const
    {
        assert(isContinuous());
        Mat m = *this;
        const int totalCh = cols * channels();
        assert(totalCh % cn == 0);
        m.cols = totalCh / cn;
        m.flags_type = depth() + ((cn - 1) << 3);
        m.step = (size_t)m.cols * m.elemSize();
        return m;
    }
    template <typename T> T& at(int y, int x) { return *(T*)(data + (size_t)y * step + x * sizeof(T)); }
    template <typename T> const T& at(int y, int x) const { return *(const T*)(data + (size_t)y * step + x * sizeof(T)); }
    // single index: row or column vectors (and continuous matrices)
    template <typename T> T& at(int i)
    {
        if (rows == 1) return *(T*)(data + i * sizeof(T));
        if (cols == 1) return *(T*)(data + (size_t)i * step);
        return at<T>(i / cols, i % cols);
    }
    template <typename T> const T& at(int i) const { return const_cast<Mat*>(this)->at<T>(i); }
    uchar* ptr(int y = 0) { return data + (size_t)y * step; }
    const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }
    template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
    template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step); }

    static MatZeros zeros(int r, int c, int type) { return MatZeros{r, c, type}; }
    static Mat ones(int r, int c, int type) { Mat m(r, c, type); m.fill(1.0); return m; }
    static Mat eye(int r, int c, int type)
    {
        Mat m(r, c, type);
        m.fill(0.0);
        for (int i = 0; i < std::min(r, c); i++) m.put(i, i, 1.0);
        return m;
    }
    Mat(const MatZeros& z) : data(nullptr) { alloc(z.rows, z.cols, z.type); fill(0.0); }
    // MatExpr assignment semantics: when size/type already match, fill in place (keeps aliasing)
    Mat& operator=(const MatZeros& z)
    {
        create(z.rows, z.cols, z.type);
        for (int y = 0; y < rows; y++) std::memset(data + (size_t)y * step, 0, (size_t)cols * elemSize());
        return *this;
    }
    void put(int y, int x, double v)
    {
        uchar* q = data + (size_t)y * step + (size_t)x * elemSize();
        switch (depth()) {
        case CV_8U: *q = (uchar)v; break;
        case CV_32F: { float f = (float)v; std::memcpy(q, &f, 4); break; }
        case CV_64F: std::memcpy(q, &v, 8); break;
        case CV_32S: { int iv = (int)v; std::memcpy(q, &iv, 4); break; }
        }
    }
    void fill(double v)
    {
        for (int y = 0; y < rows; y++)
            for (int x = 0; x < cols * channels(); x++) {
                uchar* q = data + (size_t)y * step + (size_t)x * elemSize1();
                switch (depth()) {
                case CV_8U: *q = (uchar)v; break;
                case CV_32F: { float f = (float)v; std::memcpy(q, &f, 4); break; }
                case CV_64F: std::memcpy(q, &v, 8); break;
                case CV_32S: { int iv = (int)v; std::memcpy(q, &iv, 4); break; }
                }
            }
    }

    Mat t() const
    {
        assert(depth() == CV_32F && channels() == 1);
        Mat m(cols, rows, flags_type);
        for (int y = 0; y < rows; y++)
            for (int x = 0; x < cols; x++) m.at<float>(x, y) = at<float>(y, x);
        return m;
    }
    // Mat::dot: double accumulation of float products (OpenCV dotProd_32f scalar tail)
    double dot(const Mat& o) const
    {
        assert(depth() == CV_32F && o.depth() == CV_32F && total() == o.total());
        double r = 0;
        const int n = (int)total();
        for (int i = 0; i < n; i++) r += (double)lin(i) * (double)o.lin(i);
        return r;
    }
    float lin(int i) const { return rows == 1 ? at<float>(0, i) : cols == 1 ? at<float>(i, 0) : at<float>(i / cols, i % cols); }

private:
    void alloc(int r, int c, int type)
    {
        flags_type = type;
        rows = r; cols = c; step = (size_t)c * elemSize();
        buf = std::make_shared<std::vector<uchar>>((size_t)r * step + 8);
        data = buf->data();
    }
};

// ---- eager float arithmetic (CV_32F, one channel) ----
inline Mat mat_binary(const Mat& a, const Mat& b, int op)
{
    assert(a.depth() == CV_32F && b.depth() == CV_32F && a.rows == b.rows && a.cols == b.cols);
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) {
            const float u = a.at<float>(y, x), v = b.at<float>(y, x);
            m.at<float>(y, x) = op == 0 ? u + v : u - v;
        }
    return m;
}
inline Mat operator+(const Mat& a, const Mat& b) { return mat_binary(a, b, 0); }
inline Mat operator-(const Mat& a, const Mat& b) { return mat_binary(a, b, 1); }
inline Mat mat_scale(const Mat& a, double s, bool divide)
{
    assert(a.depth() == CV_32F);
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++)
            m.at<float>(y, x) = divide ? (float)((double)a.at<float>(y, x) / s) : (float)((double)a.at<float>(y, x) * s);
    return m;
}
inline Mat operator-(const Mat& a) { return mat_scale(a, -1.0, false); }
inline Mat operator*(const Mat& a, double s) { return mat_scale(a, s, false); }
inline Mat operator*(double s, const Mat& a) { return mat_scale(a, s, false); }
inline Mat operator/(const Mat& a, double s) { return mat_scale(a, s, true); }
inline Mat operator*(const Mat& a, const Mat& b)
{
    assert(a.depth() == CV_32F && b.depth() == CV_32F && a.cols == b.rows);
    Mat m(a.rows, b.cols, CV_32F);
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < b.cols; x++) {
            float s = a.at<float>(y, 0) * b.at<float>(0, x);
            for (int k = 1; k < a.cols; k++) s = s + a.at<float>(y, k) * b.at<float>(k, x);
            m.at<float>(y, x) = s;
        }
    return m;
}

enum { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4 };
// cv::norm on CV_32F: accumulates in double (normL2_32f / normDiffL1_32f)
inline double norm(const Mat& a, int normType = NORM_L2)
{
    assert(a.depth() == CV_32F);
    double s = 0;
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) {
            const double v = a.at<float>(y, x);
            s += normType == NORM_L1 ? std::fabs(v) : v * v;
        }
    return normType == NORM_L1 ? s : std::sqrt(s);
}
inline double norm(const Mat& a, const Mat& b, int normType = NORM_L2)
{
    assert(a.depth() == CV_32F && b.depth() == CV_32F && a.rows == b.rows && a.cols == b.cols);
    double s = 0;
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) {
            const double v = (double)(a.at<float>(y, x) - b.at<float>(y, x));
            s += normType == NORM_L1 ? std::fabs(v) : v * v;
        }
    return normType == NORM_L1 ? s : std::sqrt(s);
}

// cv::Mat_<T>(r, c) << a, b, c ...
template <typename T> struct MatTypeOf;
template <> struct MatTypeOf<float> { enum { value = CV_32F }; };
template <> struct MatTypeOf<double> { enum { value = CV_64F }; };
template <> struct MatTypeOf<uchar> { enum { value = CV_8U }; };
template <typename T> class Mat_ : public Mat {
public:
    Mat_() {}
    Mat_(int r, int c) : Mat(r, c, MatTypeOf<T>::value) {}
    Mat_(const Mat& m) : Mat(m) {}
    T& operator()(int y, int x) { return this->template at<T>(y, x); }
    const T& operator()(int y, int x) const { return this->template at<T>(y, x); }
};
template <typename T> struct MatCommaInitializer_ {
    Mat_<T> m;
    int idx;
    MatCommaInitializer_(const Mat_<T>& _m) : m(_m), idx(0) {}
    template <typename U> MatCommaInitializer_& operator,(U v)
    {
        m.template at<T>(idx / m.cols, idx % m.cols) = (T)v;
        idx++;
        return *this;
    }
    operator Mat() const { return m; }
    operator Mat_<T>() const { return m; }
};
template <typename T, typename U> inline MatCommaInitializer_<T> operator<<(const Mat_<T>& m, U v)
{
    MatCommaInitializer_<T> ci(m);
    return (ci, v);
}

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
    _OutputArray(const Mat& _m) : _InputArray(_m) {}
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

// cv::undistortPoints for the one call shape the reference uses (src/Frame.cc:603-609, 641): N x 1 CV_32FC2 in place,
// R empty, P = K.  Arithmetic: orbo::undistort_points (oracle/cvprim.h, pinned against cv2 4.13).
inline void undistortPoints(const Mat& src, const Mat& dst, const Mat& K, const Mat& distCoef, const Mat& R, const Mat& P)
{
    assert(src.type() == CV_32FC2 && src.cols == 1 && src.data == dst.data && R.empty());
    assert(K.data == P.data && K.depth() == CV_32F);
    (void)R; (void)P; (void)dst;
    const int n = src.rows;
    std::vector<float> x(n), y(n), ox(n), oy(n), d;
    for (int i = 0; i < n; i++) { x[i] = src.ptr<float>(i)[0]; y[i] = src.ptr<float>(i)[1]; }
    const int nd = (int)distCoef.total();
    for (int i = 0; i < nd; i++) d.push_back(distCoef.at<float>(i));
    orbo::undistort_points(x.data(), y.data(), n, K.at<float>(0, 0), K.at<float>(1, 1), K.at<float>(0, 2), K.at<float>(1, 2),
                           d.data(), nd, ox.data(), oy.data());
    for (int i = 0; i < n; i++) { const_cast<Mat&>(src).ptr<float>(i)[0] = ox[i]; const_cast<Mat&>(src).ptr<float>(i)[1] = oy[i]; }
}

// cv::FileStorage / cv::FileNode: declarations only, so that TemplatedVocabulary.h's yml save/load members parse; the
// harness loads vocabularies through the reference's own loadFromTextFile.
class FileNode {
public:
    FileNode operator[](const char*) const { return FileNode(); }
    FileNode operator[](const std::string&) const { return FileNode(); }
    FileNode operator[](int) const { return FileNode(); }
    size_t size() const { return 0; }
    operator int() const { return 0; }
    operator double() const { return 0; }
    operator float() const { return 0; }
    operator std::string() const { return std::string(); }
};
class FileStorage {
public:
    enum { READ = 0, WRITE = 1 };
    FileStorage() {}
    FileStorage(const char*, int) {}
    FileStorage(const std::string&, int) {}
    bool isOpened() const { return false; }
    void release() {}
    FileNode operator[](const char*) const { return FileNode(); }
    FileNode operator[](const std::string&) const { return FileNode(); }
};
template <typename T> inline FileStorage& operator<<(FileStorage& fs, const T&) { return fs; }

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
