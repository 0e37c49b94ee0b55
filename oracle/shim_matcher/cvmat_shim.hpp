// oracle/shim_matcher/cvmat_shim.hpp — TEST INFRASTRUCTURE ONLY.
//
// Stand-in for the part of the OpenCV C++ API that /root/reference/src/ORBmatcher.cc and the extracted member
// functions of Frame / KeyFrame / MapPoint use, so that those UNMODIFIED reference sources compile where they lie
// (oracle/Makefile -> oracle/_ref/libref_orbmatcher.so). OpenCV itself (>= 4.5.1, CMakeLists.txt:46 of the
// reference) is a third-party dependency that is absent here; the arithmetic of each operator below restates what
// OpenCV 4.13 does and is pinned against the real library through Python cv2 (tests/golden/cv2_matops.npz,
// tests/test_cpu_refmatcher.py):
//   * A*B (+C) with no transpose flag, CV_32F, inner length 2..4 and an output dimension equal to it:
//       cv::gemm's small-matrix path — the products of a row are accumulated IN FLOAT, left to right, then
//       d = (float)((double)t*alpha + (double)c*beta).                        (core/src/matmul.simd.hpp, gemmImpl)
//   * every other product (here: A.t()*B, i.e. GEMM_1_T): GEMMSingleMul<float,double> — accumulation in DOUBLE,
//       d = (float)(s*alpha).
//   * Mat/s, s*Mat, -Mat, s*Mat.t(): MatExpr with a double alpha, realised by convertTo: d = src*(float)alpha.
//   * Mat - Mat, Mat + Mat: element-wise float.
//   * Mat::dot, cv::norm (L2) and cv::norm(a, b, NORM_L1) on CV_32F: accumulate in double, element order.
// Two element types only: CV_8U (descriptors, pyramid images) and CV_32F (poses, points, SAD patches).
#pragma once
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5
#define CV_32FC1 5

typedef unsigned char uchar;

namespace cv {

template <typename T>
struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
};
typedef Point_<float> Point2f;
typedef Point_<int> Point;

struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
};

enum { NORM_L1 = 2, NORM_L2 = 4 };

class Mat;
struct TExpr;       // a.t() * alpha
struct ScaledExpr;  // a * alpha
struct GemmExpr;    // alpha * op(a) * b + beta * c
struct OnesExpr;    // alpha * ones(rows, cols)

class Mat {
public:
    int rows = 0, cols = 0;
    int tp = CV_8U;
    size_t step = 0;
    uchar* data = nullptr;
    std::shared_ptr<std::vector<uchar>> buf;

    Mat() {}
    Mat(int r, int c, int type) { create(r, c, type); }
    Mat(int r, int c, int type, void* ptr, size_t _step = 0)
        : rows(r), cols(c), tp(type), step(_step ? _step : (size_t)c * esz(type)), data((uchar*)ptr) {}
    Mat(const TExpr& e);
    Mat(const ScaledExpr& e);
    Mat(const GemmExpr& e);
    Mat& operator=(const TExpr& e) { return *this = Mat(e); }
    Mat& operator=(const ScaledExpr& e) { return *this = Mat(e); }
    Mat& operator=(const GemmExpr& e) { return *this = Mat(e); }

    static size_t esz(int type) { return type == CV_32F ? 4 : 1; }
    size_t elemSize() const { return esz(tp); }
    int type() const { return tp; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    bool isContinuous() const { return step == (size_t)cols * elemSize() || rows == 1; }
    void create(int r, int c, int type) {
        buf = std::make_shared<std::vector<uchar>>((size_t)r * c * esz(type));
        rows = r; cols = c; tp = type; step = (size_t)c * esz(type); data = buf->data();
    }
    Mat clone() const {
        Mat m(rows, cols, tp);
        for (int y = 0; y < rows; y++) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, cols * elemSize());
        return m;
    }
    Mat roi(int y0, int y1, int x0, int x1) const {
        Mat m = *this;
        m.data = data + (size_t)y0 * step + (size_t)x0 * elemSize();
        m.rows = y1 - y0; m.cols = x1 - x0;
        return m;
    }
    Mat rowRange(int a, int b) const { return roi(a, b, 0, cols); }
    Mat colRange(int a, int b) const { return roi(0, rows, a, b); }
    Mat row(int y) const { return roi(y, y + 1, 0, cols); }
    Mat col(int x) const { return roi(0, rows, x, x + 1); }

    template <typename T> T& at(int y, int x) { return *(T*)(data + (size_t)y * step + (size_t)x * sizeof(T)); }
    template <typename T> const T& at(int y, int x) const { return *(const T*)(data + (size_t)y * step + (size_t)x * sizeof(T)); }
    // Mat::at(int i0): element i0 of a continuous or single-row matrix, else row i0 of a single-column matrix
    template <typename T> T& at(int i) {
        if (isContinuous() || rows == 1) return ((T*)data)[i];
        return *(T*)(data + step * (size_t)i);
    }
    template <typename T> const T& at(int i) const { return const_cast<Mat*>(this)->at<T>(i); }
    template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
    template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step); }
    uchar* ptr(int y = 0) { return data + (size_t)y * step; }
    const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }

    // convertTo without scale: 8U -> 32F (exact) or a copy; dst may be *this
    void convertTo(Mat& dst, int type) const {
        Mat out(rows, cols, type);
        for (int y = 0; y < rows; y++)
            for (int x = 0; x < cols; x++) {
                const float v = tp == CV_32F ? at<float>(y, x) : (float)at<uchar>(y, x);
                if (type == CV_32F) out.at<float>(y, x) = v;
                else out.at<uchar>(y, x) = (uchar)v;
            }
        dst = out;
    }
    void copyTo(Mat& dst) const { dst = clone(); }

    TExpr t() const;
    double dot(const Mat& m) const {  // dotProd_32f: products and sum in double, element order
        assert(tp == CV_32F && m.tp == CV_32F && rows * cols == m.rows * m.cols);
        double r = 0;
        const int n = rows * cols;
        for (int i = 0; i < n; i++) r += (double)lin(i) * (double)m.lin(i);
        return r;
    }
    float lin(int i) const { return at<float>(cols == 0 ? 0 : i / cols, cols == 0 ? 0 : i % cols); }

    static OnesExpr ones(int r, int c, int type);
};

struct TExpr { Mat a; double alpha; };
struct ScaledExpr { Mat a; double alpha; };
struct GemmExpr { Mat a, b, c; double alpha, beta; bool aT; };
struct OnesExpr { int rows, cols; double alpha; };

inline TExpr Mat::t() const { return TExpr{*this, 1.0}; }
inline OnesExpr Mat::ones(int r, int c, int) { return OnesExpr{r, c, 1.0}; }

// convertTo(m, -1, alpha): cvt_32f computes src*(float)alpha (+ (float)0) in float
inline Mat scaled_copy(const Mat& a, double alpha) {
    Mat m(a.rows, a.cols, CV_32F);
    const float fa = (float)alpha;
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) m.at<float>(y, x) = a.at<float>(y, x) * fa;
    return m;
}
inline Mat::Mat(const ScaledExpr& e) { *this = scaled_copy(e.a, e.alpha); }
inline Mat::Mat(const TExpr& e) {  // MatOp_T::assign: cv::transpose, then convertTo when alpha != 1
    Mat m(e.a.cols, e.a.rows, CV_32F);
    for (int y = 0; y < e.a.rows; y++)
        for (int x = 0; x < e.a.cols; x++) m.at<float>(x, y) = e.a.at<float>(y, x);
    *this = e.alpha != 1 ? scaled_copy(m, e.alpha) : m;
}
inline Mat::Mat(const GemmExpr& e) {
    const Mat &A = e.a, &B = e.b;
    const int M = e.aT ? A.cols : A.rows, len = e.aT ? A.rows : A.cols, N = B.cols;
    assert(A.tp == CV_32F && B.tp == CV_32F && B.rows == len);
    const bool hasC = !e.c.empty();
    Mat d(M, N, CV_32F);
    const bool smallPath = !e.aT && len >= 2 && len <= 4 && (len == N || len == M);
    for (int i = 0; i < M; i++)
        for (int j = 0; j < N; j++) {
            const double cv = hasC ? (double)e.c.at<float>(i, j) * e.beta : 0.0;
            if (smallPath) {
                float t = A.at<float>(i, 0) * B.at<float>(0, j);
                for (int k = 1; k < len; k++) t = t + A.at<float>(i, k) * B.at<float>(k, j);
                d.at<float>(i, j) = (float)((double)t * e.alpha + cv);
            } else {
                double s = 0;
                for (int k = 0; k < len; k++)
                    s += (double)(e.aT ? A.at<float>(k, i) : A.at<float>(i, k)) * (double)B.at<float>(k, j);
                d.at<float>(i, j) = hasC ? (float)(s * e.alpha + cv) : (float)(s * e.alpha);
            }
        }
    *this = d;
}

inline TExpr operator-(const TExpr& e) { return TExpr{e.a, -e.alpha}; }
inline TExpr operator*(double s, const TExpr& e) { return TExpr{e.a, e.alpha * s}; }
inline ScaledExpr operator-(const Mat& a) { return ScaledExpr{a, -1.0}; }
inline ScaledExpr operator*(double s, const Mat& a) { return ScaledExpr{a, s}; }
inline ScaledExpr operator*(const Mat& a, double s) { return ScaledExpr{a, s}; }
inline ScaledExpr operator/(const Mat& a, double s) { return ScaledExpr{a, 1. / s}; }
inline OnesExpr operator*(double s, const OnesExpr& e) { return OnesExpr{e.rows, e.cols, e.alpha * s}; }
inline GemmExpr operator*(const Mat& a, const Mat& b) { return GemmExpr{a, b, Mat(), 1.0, 0.0, false}; }
inline GemmExpr operator*(const TExpr& a, const Mat& b) { return GemmExpr{a.a, b, Mat(), a.alpha, 0.0, true}; }
inline GemmExpr operator*(const ScaledExpr& a, const Mat& b) { return GemmExpr{a.a, b, Mat(), a.alpha, 0.0, false}; }
inline GemmExpr operator+(const GemmExpr& g, const Mat& c) { return GemmExpr{g.a, g.b, c, g.alpha, 1.0, g.aT}; }
inline Mat operator-(const Mat& a, const Mat& b) {
    assert(a.tp == CV_32F && b.tp == CV_32F && a.rows == b.rows && a.cols == b.cols);
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) m.at<float>(y, x) = a.at<float>(y, x) - b.at<float>(y, x);
    return m;
}
inline Mat operator+(const Mat& a, const Mat& b) {
    Mat m(a.rows, a.cols, CV_32F);
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) m.at<float>(y, x) = a.at<float>(y, x) + b.at<float>(y, x);
    return m;
}
// Mat - alpha*ones: the initializer is realised as a matrix filled with (float)alpha, then cv::subtract
inline Mat operator-(const Mat& a, const OnesExpr& e) {
    Mat m(a.rows, a.cols, CV_32F);
    const float v = (float)e.alpha;
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) m.at<float>(y, x) = a.at<float>(y, x) - v;
    return m;
}
inline double norm(const Mat& a) {  // normL2_32f: squares and sum in double, then sqrt
    double s = 0;
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) {
            const double v = (double)a.at<float>(y, x);
            s += v * v;
        }
    return std::sqrt(s);
}
inline double norm(const Mat& a, const Mat& b, int normType) {  // normDiffL1_32f: |a-b| in float, sum in double
    assert(normType == NORM_L1);
    (void)normType;
    double s = 0;
    for (int y = 0; y < a.rows; y++)
        for (int x = 0; x < a.cols; x++) s += (double)std::abs(a.at<float>(y, x) - b.at<float>(y, x));
    return s;
}

}  // namespace cv
