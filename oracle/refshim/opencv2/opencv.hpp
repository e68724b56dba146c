// opencv.hpp — minimal stand-in for the OpenCV C++ API that /root/reference/src/ORBextractor.cc uses, so that the
// reference's OWN source file compiles here without OpenCV (TEST INFRASTRUCTURE: only oracle/_ref is built with it).
//
// What is real and what is restated when oracle/_ref runs:
//   * real: every line of ORBextractor.cc / ORBextractor.h — constructor tables, ComputePyramid's level sizes,
//     the FAST cell loop with its retry, DivideNode / DistributeOctTree (std::list, heap pointers and all), IC_Angle,
//     computeOrbDescriptor, the lapping-area output order of operator();
//   * restated: the OpenCV primitives behind cv::resize, cv::copyMakeBorder, cv::FAST, cv::GaussianBlur and
//     cv::fastAtan2, which forward to oracle/cpu_ref/cvprim.cc — each one held bit for bit to cv2 4.13.0 by
//     tests/test_oracle_vs_cv2.py.  cv::Mat here is a plain ref-counted 8-bit matrix with ROI views.
// Written from the documented OpenCV interface; no OpenCV source was consulted or copied.
#pragma once
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <type_traits>
#include <utility>
#include <vector>

#include "../../cpu_ref/cvprim.h"

typedef unsigned char uchar;

#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5        // only named by FORB::toMat32F, which nothing on the path calls
#define CV_PI 3.1415926535897932384626433832795

namespace cv {

typedef ::uchar uchar;

inline int cvRound(float v) { return (int)lrintf(v); }      // round half to even, like cvtss2si
inline int cvRound(double v) { return (int)lrint(v); }
inline int cvRound(int v) { return v; }
inline int cvFloor(double v) { return (int)std::floor(v); }
inline int cvCeil(double v) { return (int)std::ceil(v); }

template <class T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
    template <class S> Point_& operator*=(S s) { x = (T)(x * s); y = (T)(y * s); return *this; }
};
typedef Point_<int> Point2i;
typedef Point2i Point;
typedef Point_<float> Point2f;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};
struct Rect {
    int x, y, width, height;
    Rect(int x_, int y_, int w, int h) : x(x_), y(y_), width(w), height(h) {}
};

struct KeyPoint {          // 28 bytes, the layout of cv::KeyPoint
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float size_, float angle_ = -1, float response_ = 0, int octave_ = 0, int class_id_ = -1)
        : pt(x, y), size(size_), angle(angle_), response(response_), octave(octave_), class_id(class_id_) {}
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

struct Scalar {               // what KeyFrame.cc names: cv::Scalar(0.0f) as a fill value
    double val[4];
    Scalar(double v0 = 0, double v1 = 0, double v2 = 0, double v3 = 0) : val{v0, v1, v2, v3} {}
};
class Mat;
class MatExpr;
class _OutputArray;

class Mat {
public:
    int rows = 0, cols = 0;
    size_t step = 0;                 // bytes per row of the underlying buffer
    uchar* data = nullptr;

    Mat() {}
    Mat(int r, int c, int type) { create(r, c, type); }
    Mat(Size s, int type) { create(s.height, s.width, type); }
    Mat(int r, int c, int type, const Scalar& v) {
        create(r, c, type);
        for (int i = 0; i < r; ++i)
            for (int j = 0; j < c; ++j) {
                if (type == CV_32F) *reinterpret_cast<float*>(data + (size_t)i * step + 4 * (size_t)j) = (float)v.val[0];
                else data[(size_t)i * step + j] = (uchar)v.val[0];
            }
    }
    // user-allocated data (not owned)
    Mat(int r, int c, int type, void* d, size_t st) : rows(r), cols(c), step(st), data((uchar*)d), esz_(type == CV_32F ? 4 : 1) {
        assert(type == CV_8U || type == CV_32F);
    }

    void create(int r, int c, int type) {
        assert(type == CV_8U || type == CV_32F);
        const int e = type == CV_32F ? 4 : 1;
        if (data && r == rows && c == cols && e == esz_) return;   // like OpenCV: a matching header is kept (ROI views stay views)
        rows = r; cols = c; esz_ = e; step = (size_t)c * e;
        buf_.reset(new uchar[(size_t)r * c * e + 1], std::default_delete<uchar[]>());
        data = buf_.get();
    }
    void release() { buf_.reset(); data = nullptr; rows = cols = 0; step = 0; }
    static Mat zeros(int r, int c, int type) { Mat m(r, c, type); std::memset(m.data, 0, (size_t)r * m.step); return m; }
    static Mat eye(int r, int c, int type) { Mat m = zeros(r, c, type); for (int i = 0; i < r && i < c; ++i) m.at<float>(i, i) = 1.f; return m; }
    size_t total() const { return (size_t)rows * cols; }
    int channels() const { return esz_ >= 4 ? esz_ / 4 : 1; }
    // reshape(cn) of a continuous float matrix: same data, cn floats per element (Frame::UndistortKeyPoints, Frame.cc:748-750)
    Mat reshape(int cn) const {
        assert(esz_ >= 4 && step == (size_t)cols * esz_ && (cols * channels()) % cn == 0);
        Mat m = *this;
        m.cols = cols * channels() / cn; m.esz_ = 4 * cn; m.step = (size_t)m.cols * m.esz_;
        return m;
    }

    int type() const { return esz_ >= 4 ? CV_32F : CV_8U; }
    size_t elemSize() const { return (size_t)esz_; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    size_t step1() const { return step / esz_; }

    Mat operator()(const Rect& r) const { return view(r.y, r.y + r.height, r.x, r.x + r.width); }
    Mat rowRange(int a, int b) const { return view(a, b, 0, cols); }
    Mat colRange(int a, int b) const { return view(0, rows, a, b); }
    Mat row(int i) const { return view(i, i + 1, 0, cols); }
    Mat col(int j) const { return view(0, rows, j, j + 1); }
    // single index: element i of a row or column vector
    template <class T> T& at(int i) { return cols == 1 ? at<T>(i, 0) : at<T>(i / cols, i % cols); }
    template <class T> const T& at(int i) const { return cols == 1 ? at<T>(i, 0) : at<T>(i / cols, i % cols); }
    // float algebra (CV_32F), evaluated eagerly; see the operators below the class
    inline MatExpr t() const;
    inline MatExpr inv() const;
    inline double dot(const Mat& b) const;
    // M = <matrix expression>: OpenCV evaluates the expression INTO M when M already has the result's size and type (so
    // `A.row(0) = x * B.row(2) - B.row(0)` fills the row of A); M = <Mat> only rebinds the header.
    Mat(const Mat&) = default;
    Mat& operator=(const Mat&) = default;
    inline Mat& operator=(const MatExpr& e);

    template <class T> T& at(int y, int x) { return *reinterpret_cast<T*>(data + (size_t)y * step + x * sizeof(T)); }
    template <class T> const T& at(int y, int x) const { return *reinterpret_cast<const T*>(data + (size_t)y * step + x * sizeof(T)); }
    uchar* ptr(int y = 0) { return data + (size_t)y * step; }
    const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }
    template <class T> T* ptr(int y = 0) { return reinterpret_cast<T*>(data + (size_t)y * step); }
    template <class T> const T* ptr(int y = 0) const { return reinterpret_cast<const T*>(data + (size_t)y * step); }

    Mat clone() const {
        Mat m;
        if (empty()) return m;
        m.create(rows, cols, type());
        for (int y = 0; y < rows; ++y) std::memcpy(m.ptr(y), ptr(y), (size_t)cols * esz_);
        return m;
    }
    inline void copyTo(const _OutputArray& dst) const;

private:
    Mat view(int y0, int y1, int x0, int x1) const {
        assert(y0 >= 0 && y1 <= rows && x0 >= 0 && x1 <= cols && y0 <= y1 && x0 <= x1);
        Mat m;
        m.rows = y1 - y0; m.cols = x1 - x0; m.step = step; m.esz_ = esz_; m.data = data + (size_t)y0 * step + (size_t)x0 * esz_; m.buf_ = buf_;
        return m;
    }
    int esz_ = 1;
    std::shared_ptr<uchar> buf_;     // keeps the allocation alive for every view
};

class _InputArray {
public:
    _InputArray(const Mat& m) : m_(&m) {}
    bool empty() const { return m_->empty(); }
    Mat getMat() const { return *m_; }
private:
    const Mat* m_;
};
class _OutputArray {
public:
    _OutputArray(const Mat& m) : m_(const_cast<Mat*>(&m)) {}
    void release() const { m_->release(); }
    void create(int r, int c, int type) const { m_->create(r, c, type); }
    Mat getMat() const { return *m_; }
    Mat& ref() const { return *m_; }
private:
    Mat* m_;
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

inline void Mat::copyTo(const _OutputArray& dst) const {
    dst.create(rows, cols, type());
    Mat& d = dst.ref();
    for (int y = 0; y < rows; ++y) std::memmove(d.ptr(y), ptr(y), (size_t)cols * esz_);
}

// cv::undistortPoints(src, dst, cameraMatrix, distCoeffs, R = empty, P): forwards to the library's hook (defined by the
// wrapper that needs it from oracle/cpu_ref's undistortKeyPoints, which tests/test_oracle_vs_cv2.py holds to cv2 4.13).
void orbref_undistort_points_hook(const float* xy, int n, const float K4[4], const float* dist, int ndist, const float P4[4], float* out);
inline void undistortPoints(InputArray src, OutputArray dst, InputArray K, InputArray dist, InputArray R, InputArray P) {
    const Mat s = src.getMat(), k = K.getMat(), d = dist.getMat(), p = P.getMat();
    assert(R.empty() && s.channels() == 2 && s.cols == 1 && s.step == 8);
    const float K4[4] = {k.at<float>(0, 0), k.at<float>(1, 1), k.at<float>(0, 2), k.at<float>(1, 2)};
    const float P4[4] = {p.at<float>(0, 0), p.at<float>(1, 1), p.at<float>(0, 2), p.at<float>(1, 2)};
    std::vector<float> dc(d.total()), out((size_t)s.rows * 2);
    for (size_t i = 0; i < dc.size(); ++i) dc[i] = d.at<float>((int)i);
    orbref_undistort_points_hook(s.ptr<float>(0), s.rows, K4, dc.data(), (int)dc.size(), P4, out.data());
    Mat& o = dst.ref();
    if (o.data != s.data) o = Mat(s.rows, 2, CV_32F).reshape(2);
    std::memcpy(o.data, out.data(), out.size() * sizeof(float));
}

// ---- float matrices (CV_32F): the small dense algebra ORBmatcher.cc writes with cv::Mat.  OpenCV evaluates A*B (+C) of
// such sizes in plain float, products summed left to right, C added last (checked against cv2.gemm 4.13 on 20 000 random
// 3x3 * 3x1 + 3x1 and 3x3 * 3x3 cases: identical every time; a double accumulator matches only 14-19 % of them);
// cv::norm accumulates squares in double.  Expressions are evaluated operator by operator, which for A*B+C gives the
// same float operations in the same order as the fused GEMM.
class MatExpr : public Mat {       // the value of an arithmetic expression (evaluated eagerly); see Mat::operator=(const MatExpr&)
public:
    MatExpr(const Mat& m) : Mat(m) {}
};
inline Mat& Mat::operator=(const MatExpr& e) {
    if (data && e.data && rows == e.rows && cols == e.cols && elemSize() == e.elemSize() && data != e.data) {
        for (int y = 0; y < rows; ++y) std::memmove(ptr(y), e.ptr(y), (size_t)cols * elemSize());
        return *this;
    }
    return *this = static_cast<const Mat&>(e);
}
inline std::ostream& operator<<(std::ostream& os, const Mat& m) {
    os << "[";
    for (int i = 0; i < m.rows; ++i) { for (int j = 0; j < m.cols; ++j) os << (j ? ", " : "") << (m.type() == CV_32F ? m.at<float>(i, j) : (float)m.at<uchar>(i, j)); os << (i + 1 < m.rows ? ";\n " : "]"); }
    return os;
}

inline MatExpr operator*(const Mat& a, const Mat& b) {
    assert(a.type() == CV_32F && b.type() == CV_32F && a.cols == b.rows);
    Mat c(a.rows, b.cols, CV_32F);
    for (int i = 0; i < a.rows; ++i)
        for (int j = 0; j < b.cols; ++j) {
            float s = a.at<float>(i, 0) * b.at<float>(0, j);
            for (int k = 1; k < a.cols; ++k) s += a.at<float>(i, k) * b.at<float>(k, j);
            c.at<float>(i, j) = s;
        }
    return c;
}
#define ORBREF_MAT_EW(OP)                                                                          \
    inline MatExpr operator OP(const Mat& a, const Mat& b) {                                       \
        assert(a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows && a.cols == b.cols);  \
        Mat c(a.rows, a.cols, CV_32F);                                                             \
        for (int i = 0; i < a.rows; ++i)                                                           \
            for (int j = 0; j < a.cols; ++j) c.at<float>(i, j) = a.at<float>(i, j) OP b.at<float>(i, j); \
        return c;                                                                                  \
    }
ORBREF_MAT_EW(+)
ORBREF_MAT_EW(-)
#undef ORBREF_MAT_EW
// A * s, A / s: OpenCV scales with convertTo, whose scale factor is narrowed to float for CV_32F (A / s multiplies by (float)(1 / s))
inline MatExpr operator*(const Mat& a, double s) {
    Mat c(a.rows, a.cols, CV_32F);
    const float f = (float)s;
    for (int i = 0; i < a.rows; ++i)
        for (int j = 0; j < a.cols; ++j) c.at<float>(i, j) = a.at<float>(i, j) * f;
    return c;
}
inline MatExpr operator*(double s, const Mat& a) { return a * s; }
inline MatExpr operator/(const Mat& a, double s) { return a * (1.0 / s); }
inline MatExpr operator-(const Mat& a) { return a * -1.0; }
// Mat::inv() of a 3x3 float matrix (cofactors in double).  Only Pinhole::epipolarConstrain (the cv::Mat overload, which
// nothing on the hot path calls) reaches it; no parity claim rests on it.
inline MatExpr Mat::inv() const {
    assert(rows == 3 && cols == 3 && type() == CV_32F);
    double a[3][3];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) a[i][j] = at<float>(i, j);
    const double det = a[0][0] * (a[1][1] * a[2][2] - a[1][2] * a[2][1]) - a[0][1] * (a[1][0] * a[2][2] - a[1][2] * a[2][0]) +
                       a[0][2] * (a[1][0] * a[2][1] - a[1][1] * a[2][0]);
    Mat r = Mat::zeros(3, 3, CV_32F);
    if (det == 0) return r;
    const double d = 1.0 / det;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            const int i1 = (j + 1) % 3, i2 = (j + 2) % 3, j1 = (i + 1) % 3, j2 = (i + 2) % 3;
            r.at<float>(i, j) = (float)((a[i1][j1] * a[i2][j2] - a[i1][j2] * a[i2][j1]) * d);
        }
    return r;
}
inline MatExpr Mat::t() const {
    Mat c(cols, rows, CV_32F);
    for (int i = 0; i < rows; ++i)
        for (int j = 0; j < cols; ++j) c.at<float>(j, i) = at<float>(i, j);
    return c;
}
inline double Mat::dot(const Mat& b) const {
    double s = 0;
    for (int i = 0; i < rows; ++i)
        for (int j = 0; j < cols; ++j) s += (double)at<float>(i, j) * (double)b.at<float>(i, j);
    return s;
}
inline double norm(const Mat& a) { return std::sqrt(a.dot(a)); }

template <class T> struct Point3_ {
    T x, y, z;
    Point3_() : x(0), y(0), z(0) {}
    Point3_(T x_, T y_, T z_) : x(x_), y(y_), z(z_) {}
};
typedef Point3_<float> Point3f;

// cv::Matx: plain fixed-size arrays, products accumulate in the element type, k ascending
template <class T, int M, int N> struct Matx {
    T val[M * N];
    Matx() { for (int i = 0; i < M * N; ++i) val[i] = T(0); }
    // Matx(v0, v1, ...): the first values in row-major order, the rest zero (OpenCV's 2- to 16-argument constructors)
    template <class... A, class = typename std::enable_if<(sizeof...(A) >= 2 && sizeof...(A) <= M * N)>::type>
    Matx(A... a) : val{(T)a...} {}
    Matx<T, 1, N> row(int i) const { Matx<T, 1, N> r; for (int j = 0; j < N; ++j) r.val[j] = (*this)(i, j); return r; }
    template <int m, int n> Matx<T, m, n> get_minor(int i0, int j0) const {
        Matx<T, m, n> r; for (int i = 0; i < m; ++i) for (int j = 0; j < n; ++j) r(i, j) = (*this)(i0 + i, j0 + j); return r;
    }
    // 3x3 inverse by cofactors in T (restated from the documented closed form; the hot-path callers are given F12 directly)
    Matx inv() const {
        static_assert(M == 3 && N == 3, "3x3 only");
        const Matx& a = *this;
        Matx r;
        T det = a(0, 0) * (a(1, 1) * a(2, 2) - a(2, 1) * a(1, 2)) - a(0, 1) * (a(1, 0) * a(2, 2) - a(2, 0) * a(1, 2)) +
                a(0, 2) * (a(1, 0) * a(2, 1) - a(2, 0) * a(1, 1));
        if (det == 0) return r;
        const T d = 1 / det;
        r(0, 0) = (a(1, 1) * a(2, 2) - a(1, 2) * a(2, 1)) * d; r(0, 1) = (a(0, 2) * a(2, 1) - a(0, 1) * a(2, 2)) * d; r(0, 2) = (a(0, 1) * a(1, 2) - a(0, 2) * a(1, 1)) * d;
        r(1, 0) = (a(1, 2) * a(2, 0) - a(1, 0) * a(2, 2)) * d; r(1, 1) = (a(0, 0) * a(2, 2) - a(0, 2) * a(2, 0)) * d; r(1, 2) = (a(0, 2) * a(1, 0) - a(0, 0) * a(1, 2)) * d;
        r(2, 0) = (a(1, 0) * a(2, 1) - a(1, 1) * a(2, 0)) * d; r(2, 1) = (a(0, 1) * a(2, 0) - a(0, 0) * a(2, 1)) * d; r(2, 2) = (a(0, 0) * a(1, 1) - a(0, 1) * a(1, 0)) * d;
        return r;
    }
    T dot(const Matx& b) const { T s = 0; for (int i = 0; i < M * N; ++i) s += val[i] * b.val[i]; return s; }   // accumulates in T (ddot is the double one)
    operator Mat() const { Mat m(M, N, CV_32F); for (int i = 0; i < M * N; ++i) m.at<float>(i / N, i % N) = (float)val[i]; return m; }
    static Matx zeros() { return Matx(); }
    T& operator()(int i, int j) { return val[i * N + j]; }
    const T& operator()(int i, int j) const { return val[i * N + j]; }
    T& operator()(int i) { return val[i]; }
    const T& operator()(int i) const { return val[i]; }
    Matx<T, N, M> t() const { Matx<T, N, M> r; for (int i = 0; i < M; ++i) for (int j = 0; j < N; ++j) r(j, i) = (*this)(i, j); return r; }
};
template <class T, int M, int K, int N> inline Matx<T, M, N> operator*(const Matx<T, M, K>& a, const Matx<T, K, N>& b) {
    Matx<T, M, N> c;
    for (int i = 0; i < M; ++i)
        for (int j = 0; j < N; ++j) { T s = 0; for (int k = 0; k < K; ++k) s += a(i, k) * b(k, j); c(i, j) = s; }
    return c;
}
template <class T, int M, int N> inline Matx<T, M, N> operator+(const Matx<T, M, N>& a, const Matx<T, M, N>& b) {
    Matx<T, M, N> c; for (int i = 0; i < M * N; ++i) c.val[i] = a.val[i] + b.val[i]; return c;
}
template <class T, int M, int N> inline Matx<T, M, N> operator-(const Matx<T, M, N>& a, const Matx<T, M, N>& b) {
    Matx<T, M, N> c; for (int i = 0; i < M * N; ++i) c.val[i] = a.val[i] - b.val[i]; return c;
}
template <class T, int M, int N> inline Matx<T, M, N> operator-(const Matx<T, M, N>& a) {
    Matx<T, M, N> c; for (int i = 0; i < M * N; ++i) c.val[i] = -a.val[i]; return c;
}
template <class T, int M, int N> inline Matx<T, M, N> operator*(const Matx<T, M, N>& a, float s) {
    Matx<T, M, N> c; for (int i = 0; i < M * N; ++i) c.val[i] = (T)(a.val[i] * s); return c;
}
template <class T, int M, int N> inline Matx<T, M, N> operator*(float s, const Matx<T, M, N>& a) { return a * s; }
template <class T, int M, int N> inline double norm(const Matx<T, M, N>& a) {       // squares summed in double
    double s = 0; for (int i = 0; i < M * N; ++i) s += (double)a.val[i] * (double)a.val[i]; return std::sqrt(s);
}
typedef Matx<float, 3, 3> Matx33f;
typedef Matx<float, 3, 1> Matx31f;
typedef Matx<float, 4, 4> Matx44f;
typedef Matx<float, 3, 4> Matx34f;
typedef Matx<float, 1, 3> Matx13f;
typedef Matx<float, 1, 4> Matx14f;
typedef Matx<float, 4, 1> Matx41f;

// cv::Mat_<float>(r, c) << a, b, c ...
template <class T> class Mat_ : public Mat {
public:
    Mat_(int r, int c) : Mat(r, c, CV_32F) {}
};
template <class T> struct MatCommaInitializer_ {
    Mat m; int i;
    template <class V> MatCommaInitializer_(const Mat& mm, V v) : m(mm), i(0) { put(v); }
    template <class V> void put(V v) { assert(i < m.rows * m.cols); m.at<T>(i / m.cols, i % m.cols) = (T)v; ++i; }
    template <class V> MatCommaInitializer_& operator,(V v) { put(v); return *this; }
    operator Mat() const { return m; }
};
template <class T, class V> inline MatCommaInitializer_<T> operator<<(const Mat_<T>& m, V v) { return MatCommaInitializer_<T>(m, v); }

inline void hconcat(const Mat& a, const Mat& b, Mat& dst) {
    assert(a.rows == b.rows && a.type() == b.type());
    Mat c(a.rows, a.cols + b.cols, a.type());
    const size_t e = a.elemSize();
    for (int y = 0; y < a.rows; ++y) { std::memcpy(c.ptr(y), a.ptr(y), a.cols * e); std::memcpy(c.ptr(y) + a.cols * e, b.ptr(y), b.cols * e); }
    dst = c;
}
inline void vconcat(const Mat& a, const Mat& b, Mat& dst) {
    assert(a.cols == b.cols && a.type() == b.type());
    Mat c(a.rows + b.rows, a.cols, a.type());
    const size_t e = a.elemSize();
    for (int y = 0; y < a.rows; ++y) std::memcpy(c.ptr(y), a.ptr(y), a.cols * e);
    for (int y = 0; y < b.rows; ++y) std::memcpy(c.ptr(a.rows + y), b.ptr(y), b.cols * e);
    dst = c;
}

enum { NORM_L1 = 2, NORM_L2 = 4, NORM_HAMMING = 6 };
// cv::norm(A, B, NORM_L1) on 8-bit patches: the exact integer sum (Frame.cc:908)
inline double norm(const Mat& a, const Mat& b, int normType) {
    assert(normType == NORM_L1 && a.type() == CV_8U && b.type() == CV_8U && a.rows == b.rows && a.cols == b.cols);
    (void)normType;
    long s = 0;
    for (int y = 0; y < a.rows; ++y) for (int x = 0; x < a.cols; ++x) s += std::abs((int)a.at<uchar>(y, x) - (int)b.at<uchar>(y, x));
    return (double)s;
}

// cv::BFMatcher(NORM_HAMMING).knnMatch: the k smallest distances per query row, equal distances by ascending train index
// (held to cv2 4.13 by tests/test_oracle_vs_cv2.py)
struct DMatch {
    int queryIdx = -1, trainIdx = -1, imgIdx = -1;
    float distance = 0;
};
class BFMatcher {
public:
    BFMatcher(int normType = NORM_L2, bool crossCheck = false) { assert(normType == NORM_HAMMING && !crossCheck); (void)normType; (void)crossCheck; }
    void knnMatch(const Mat& q, const Mat& t, std::vector<std::vector<DMatch>>& matches, int k) const {
        matches.assign(q.rows, std::vector<DMatch>());
        std::vector<std::pair<int, int>> d(t.rows);
        for (int i = 0; i < q.rows; ++i) {
            for (int j = 0; j < t.rows; ++j) {
                int s = 0;
                for (int b = 0; b < q.cols; ++b) s += __builtin_popcount((unsigned)(q.at<uchar>(i, b) ^ t.at<uchar>(j, b)));
                d[j] = std::make_pair(s, j);
            }
            const int kk = std::min(k, t.rows);
            std::partial_sort(d.begin(), d.begin() + kk, d.end());
            for (int r = 0; r < kk; ++r) { DMatch m; m.queryIdx = i; m.trainIdx = d[r].second; m.imgIdx = 0; m.distance = (float)d[r].first; matches[i].push_back(m); }
        }
    }
};

// cv::SVD::compute: one-sided Jacobi in double on the float system, singular values descending, rows of vt = right
// singular vectors.  OpenCV's own float Jacobi differs in the last bits; the Python restatement calls the real
// cv2.SVDecomp and every test that goes through here compares decisions away from the thresholds.
struct SVD {
    enum { MODIFY_A = 1, NO_UV = 2, FULL_UV = 4 };
    static void jacobi(std::vector<double>& A, int m, int n, std::vector<double>& w, std::vector<double>& V) {
        V.assign((size_t)n * n, 0.0);
        for (int i = 0; i < n; ++i) V[(size_t)i * n + i] = 1.0;
        for (int sweep = 0; sweep < 60; ++sweep) {
            bool changed = false;
            for (int p = 0; p < n - 1; ++p)
                for (int q = p + 1; q < n; ++q) {
                    double al = 0, be = 0, ga = 0;
                    for (int i = 0; i < m; ++i) { const double x = A[(size_t)i * n + p], y = A[(size_t)i * n + q]; al += x * x; be += y * y; ga += x * y; }
                    if (std::fabs(ga) <= 1e-300 || std::fabs(ga) <= 1e-17 * std::sqrt(al * be)) continue;
                    changed = true;
                    const double zeta = (be - al) / (2.0 * ga);
                    const double t = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
                    const double c = 1.0 / std::sqrt(1.0 + t * t), sn = c * t;
                    for (int i = 0; i < m; ++i) { const double x = A[(size_t)i * n + p], y = A[(size_t)i * n + q]; A[(size_t)i * n + p] = c * x - sn * y; A[(size_t)i * n + q] = sn * x + c * y; }
                    for (int i = 0; i < n; ++i) { const double x = V[(size_t)i * n + p], y = V[(size_t)i * n + q]; V[(size_t)i * n + p] = c * x - sn * y; V[(size_t)i * n + q] = sn * x + c * y; }
                }
            if (!changed) break;
        }
        w.assign(n, 0.0);
        for (int j = 0; j < n; ++j) { double s = 0; for (int i = 0; i < m; ++i) s += A[(size_t)i * n + j] * A[(size_t)i * n + j]; w[j] = std::sqrt(s); }
    }
    // fills w (n), u (m x n, columns), vt (n x n, rows) in descending order of w
    static void solve(const float* a, int m, int n, float* w, float* u, float* vt) {
        std::vector<double> A((size_t)m * n), W, V;
        for (size_t i = 0; i < A.size(); ++i) A[i] = a[i];
        jacobi(A, m, n, W, V);
        std::vector<int> order(n);
        for (int j = 0; j < n; ++j) order[j] = j;
        std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return W[x] > W[y]; });
        for (int r = 0; r < n; ++r) {
            const int j = order[r];
            w[r] = (float)W[j];
            for (int i = 0; i < n; ++i) vt[(size_t)r * n + i] = (float)V[(size_t)i * n + j];
            for (int i = 0; i < m; ++i) u[(size_t)i * n + r] = (float)(W[j] > 0 ? A[(size_t)i * n + j] / W[j] : 0.0);
        }
    }
    static void compute(const Mat& A, Mat& w, Mat& u, Mat& vt, int flags = 0) {
        (void)flags;
        assert(A.type() == CV_32F);
        const Mat a = A.clone();
        w = Mat(A.cols, 1, CV_32F); u = Mat(A.rows, A.cols, CV_32F); vt = Mat(A.cols, A.cols, CV_32F);
        solve(a.ptr<float>(0), A.rows, A.cols, w.ptr<float>(0), u.ptr<float>(0), vt.ptr<float>(0));
    }
    template <int M, int N> static void compute(const Matx<float, M, N>& A, Matx<float, N, 1>& w, Matx<float, M, N>& u, Matx<float, N, N>& vt, int flags = 0) {
        (void)flags;
        solve(A.val, M, N, w.val, u.val, vt.val);
    }
};

namespace fisheye {   // named by KannalaBrandt8::ReconstructWithTwoViews (monocular initialisation, not on the path): compiles, never runs
inline void undistortPoints(const std::vector<Point2f>&, std::vector<Point2f>&, const Mat&, const Mat&, const Mat&, const Mat&) { assert(!"cv::fisheye::undistortPoints stand-in"); }
}

enum { BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16 };
enum { INTER_LINEAR = 1 };

inline float fastAtan2(float y, float x) { return orbref::fastAtan2(y, x); }

// cv::FAST(image, keypoints, threshold, nonmaxSuppression): type 9_16, keypoints in row-major order,
// KeyPoint(x, y, 7.f, -1, score).
inline void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression = true) {
    assert(nonmaxSuppression);
    (void)nonmaxSuppression;
    const Mat img = image.getMat();
    std::vector<orbref::FastKp> out;
    orbref::fast9_16(img.data, img.cols, img.rows, (ptrdiff_t)img.step, threshold, out);
    keypoints.clear();
    keypoints.reserve(out.size());
    for (const orbref::FastKp& k : out) keypoints.push_back(KeyPoint((float)k.x, (float)k.y, 7.f, -1, (float)k.score));
}

// cv::resize(src, dst, dsize, 0, 0, INTER_LINEAR), 8-bit single channel.
inline void resize(InputArray src_, OutputArray dst_, Size dsize, double fx = 0, double fy = 0, int interpolation = INTER_LINEAR) {
    assert(fx == 0 && fy == 0 && interpolation == INTER_LINEAR);
    (void)fx; (void)fy; (void)interpolation;
    const Mat src = src_.getMat();
    dst_.create(dsize.height, dsize.width, CV_8U);
    Mat dst = dst_.getMat();
    orbref::resizeLinearU8(src.data, src.cols, src.rows, (ptrdiff_t)src.step, dst.data, dst.cols, dst.rows, (ptrdiff_t)dst.step);
}

// cv::copyMakeBorder(src, dst, b, b, b, b, BORDER_REFLECT_101 [+ BORDER_ISOLATED]).
inline void copyMakeBorder(InputArray src_, OutputArray dst_, int top, int bottom, int left, int right, int borderType) {
    assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101 && top == bottom && top == left && top == right);
    (void)bottom; (void)left; (void)right; (void)borderType;
    const Mat src = src_.getMat();
    dst_.create(src.rows + 2 * top, src.cols + 2 * top, CV_8U);
    Mat dst = dst_.getMat();
    for (int y = 0; y < src.rows; ++y) {
        uchar* d = dst.ptr(y + top) + top;
        if (d != src.ptr(y)) std::memmove(d, src.ptr(y), (size_t)src.cols);
    }
    orbref::fillBorderReflect101(dst.data, src.cols, src.rows, top, (ptrdiff_t)dst.step);
}

// cv::GaussianBlur(src, dst, Size(7, 7), 2, 2, BORDER_REFLECT_101), 8-bit; src and dst may be the same matrix.
inline void GaussianBlur(InputArray src_, OutputArray dst_, Size ksize, double sigmaX, double sigmaY = 0, int borderType = BORDER_REFLECT_101) {
    assert(ksize.width == 7 && ksize.height == 7 && sigmaX == 2 && sigmaY == 2 && borderType == BORDER_REFLECT_101);
    (void)ksize; (void)sigmaX; (void)sigmaY; (void)borderType;
    const Mat src = src_.getMat().clone();
    dst_.create(src.rows, src.cols, CV_8U);
    Mat dst = dst_.getMat();
    orbref::gaussianBlur7x7(src.data, src.cols, src.rows, (ptrdiff_t)src.step, dst.data, (ptrdiff_t)dst.step);
}

// Only referenced by ORBextractor::ComputeKeyPointsOld, which the reference never calls (ORBextractor.cc:1083).
struct KeyPointsFilter {
    static void retainBest(std::vector<KeyPoint>& keypoints, int npoints) {
        if (npoints >= 0 && (int)keypoints.size() > npoints) keypoints.resize(npoints);
    }
};

// cv::FileStorage / cv::FileNode: TemplatedVocabulary's virtual save(fs) / load(fs) (YAML) must compile; ORB-SLAM3 only
// uses loadFromTextFile (System.cc:118), so these never open anything.
class FileNode {
public:
    FileNode operator[](const std::string&) const { return FileNode(); }
    FileNode operator[](const char*) const { return FileNode(); }
    FileNode operator[](int) const { return FileNode(); }
    size_t size() const { return 0; }
    operator int() const { return 0; }
    operator float() const { return 0.f; }
    operator double() const { return 0.0; }
    operator std::string() const { return std::string(); }
};
class FileStorage {
public:
    enum { READ = 0, WRITE = 1 };
    FileStorage() {}
    FileStorage(const std::string&, int) {}
    bool isOpened() const { return false; }
    FileNode operator[](const std::string&) const { return FileNode(); }
    FileNode operator[](const char*) const { return FileNode(); }
};
template <class T> inline FileStorage& operator<<(FileStorage& fs, const T&) { return fs; }

}  // namespace cv
