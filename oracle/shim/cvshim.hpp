// oracle/shim/cvshim.hpp — TEST INFRASTRUCTURE ONLY.
//
// A minimal stand-in for the OpenCV C++ API surface that /root/reference/src/ORBextractor.cc uses, so that the
// UNMODIFIED reference translation unit can be compiled where it lies (oracle/Makefile -> oracle/_ref/) without
// OpenCV's headers or libraries, which are not installed in this image. Types are written from scratch; the five
// image primitives forward to the integer restatements in oracle/cvprim.hpp, which tests pin against Python cv2.
// Only 8-bit single-channel matrices are supported (all the extractor touches).
#pragma once
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <memory>
#include <vector>

#include "../cvprim.hpp"

#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0
#define CV_32FC1 5

typedef unsigned char uchar;

static inline int cvRound(double v) { return cvprim::round_half_even(v); }
static inline int cvRound(float v) { return cvprim::round_half_even(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { return cvprim::floor_i(v); }
static inline int cvCeil(double v) { return cvprim::ceil_i(v); }

namespace cv {

template <typename T>
struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
    template <typename U>
    Point_(const Point_<U>& o) : x((T)o.x), y((T)o.y) {}
    Point_& operator*=(float s) {
        x = (T)(x * s);
        y = (T)(y * s);
        return *this;
    }
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
    float size;
    float angle;
    float response;
    int octave;
    int class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0, int _class_id = -1)
        : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
};

enum { BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16 };
enum { INTER_LINEAR = 1 };

class Mat;
struct ZerosExpr {
    int rows, cols, type;
};

class Mat {
public:
    int rows, cols;
    size_t step;
    uchar* data;
    std::shared_ptr<std::vector<uchar>> buf;  // owning storage (shared between headers)
    uchar* datastart;                         // start of the owning allocation (for locating ROIs)
    int wholeCols, wholeRows;

    Mat() : rows(0), cols(0), step(0), data(nullptr), datastart(nullptr), wholeCols(0), wholeRows(0) {}
    Mat(Size sz, int type) : Mat() { create(sz.height, sz.width, type); }
    Mat(int r, int c, int type) : Mat() { create(r, c, type); }
    // non-owning header over user memory
    Mat(int r, int c, int /*type*/, void* ptr, size_t _step)
        : rows(r), cols(c), step(_step), data((uchar*)ptr), datastart((uchar*)ptr), wholeCols(c), wholeRows(r) {}

    void create(int r, int c, int /*type*/) {
        if (data && r == rows && c == cols) return;
        buf = std::make_shared<std::vector<uchar>>((size_t)r * c);
        rows = r;
        cols = c;
        step = (size_t)c;
        data = buf->data();
        datastart = data;
        wholeCols = c;
        wholeRows = r;
    }
    void release() { *this = Mat(); }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return CV_8UC1; }
    size_t step1() const { return step; }
    size_t elemSize() const { return 1; }   // used by Thirdparty/fbow (descriptor rows are CV_8UC1)
    size_t elemSize1() const { return 1; }

    Mat operator()(const Rect& r) const {
        Mat m = *this;
        m.data = data + (size_t)r.y * step + r.x;
        m.rows = r.height;
        m.cols = r.width;
        return m;
    }
    Mat rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
    Mat colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }
    Mat clone() const {
        Mat m(rows, cols, CV_8UC1);
        for (int y = 0; y < rows; y++) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, cols);
        return m;
    }
    template <typename T>
    T& at(int y, int x) {
        return *(T*)(data + (size_t)y * step + x * sizeof(T));
    }
    template <typename T>
    const T& at(int y, int x) const {
        return *(const T*)(data + (size_t)y * step + x * sizeof(T));
    }
    uchar* ptr(int y = 0) { return data + (size_t)y * step; }
    const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }
    template <typename T>
    T* ptr(int y = 0) {
        return (T*)(data + (size_t)y * step);
    }
    template <typename T>
    const T* ptr(int y = 0) const {
        return (const T*)(data + (size_t)y * step);
    }

    static ZerosExpr zeros(int r, int c, int type) { return ZerosExpr{r, c, type}; }
    // Mat = MatExpr(zeros): create() is a no-op when the shape already matches, then the data is zero-filled IN PLACE
    // (this is what makes `descriptors = Mat::zeros(...)` at ORBextractor.cc:852 write through the rowRange header).
    Mat& operator=(const ZerosExpr& e) {
        create(e.rows, e.cols, e.type);
        for (int y = 0; y < rows; y++) std::memset(data + (size_t)y * step, 0, cols);
        return *this;
    }
};

struct _InputArray {
    const Mat* m;
    _InputArray() : m(nullptr) {}
    _InputArray(const Mat& _m) : m(&_m) {}
    bool empty() const { return !m || m->empty(); }
    Mat getMat() const { return m ? *m : Mat(); }
};
struct _OutputArray {
    Mat* m;
    _OutputArray(Mat& _m) : m(&_m) {}
    void release() const { m->release(); }
    void create(int r, int c, int type) const { m->create(r, c, type); }
    Mat getMat() const { return *m; }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;
static inline _InputArray noArray() { return _InputArray(); }

static inline float fastAtan2(float y, float x) { return cvprim::fast_atan2_deg(y, x); }

static inline void resize(const Mat& src, Mat& dst, Size dsize, double, double, int) {
    dst.create(dsize.height, dsize.width, CV_8UC1);
    cvprim::resize_linear_8u(src.data, src.cols, src.rows, src.step, dst.data, dst.cols, dst.rows, dst.step);
}

static inline void copyMakeBorder(const Mat& src, Mat& dst, int top, int bottom, int left, int right, int borderType) {
    assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101);
    (void)borderType;
    dst.create(src.rows + top + bottom, src.cols + left + right, CV_8UC1);
    cvprim::copy_make_border_101(src.data, src.cols, src.rows, src.step, dst.data, dst.step, top, bottom, left, right);
}

static inline void GaussianBlur(const Mat& src, Mat& dst, Size ksize, double sx, double sy, int borderType) {
    assert(ksize.width == 7 && ksize.height == 7 && sx == 2 && sy == 2 && borderType == BORDER_REFLECT_101);
    (void)ksize; (void)sx; (void)sy; (void)borderType;
    Mat tmp = src.clone();  // src and dst may be the same header
    dst.create(src.rows, src.cols, CV_8UC1);
    cvprim::gaussian_blur7_8u(tmp.data, tmp.cols, tmp.rows, tmp.step, dst.data, dst.step);
}

static inline void FAST(const Mat& img, std::vector<KeyPoint>& kps, int threshold, bool nonmax) {
    assert(nonmax);
    (void)nonmax;
    std::vector<cvprim::FastKp> v;
    cvprim::fast_9_16_nms(img.data, img.cols, img.rows, img.step, threshold, v);
    kps.clear();
    for (auto& k : v) kps.push_back(KeyPoint((float)k.x, (float)k.y, 7.f, -1, (float)k.score));
}

}  // namespace cv
