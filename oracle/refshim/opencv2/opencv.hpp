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

class Mat;
class _OutputArray;

class Mat {
public:
    int rows = 0, cols = 0;
    size_t step = 0;                 // bytes per row of the underlying buffer
    uchar* data = nullptr;

    Mat() {}
    Mat(int r, int c, int type) { create(r, c, type); }
    Mat(Size s, int type) { create(s.height, s.width, type); }
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

    int type() const { return esz_ == 4 ? CV_32F : CV_8U; }
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
    inline Mat t() const;
    inline double dot(const Mat& b) const;

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

// ---- float matrices (CV_32F): the small dense algebra ORBmatcher.cc writes with cv::Mat.  OpenCV evaluates A*B (+C) of
// such sizes in plain float, products summed left to right, C added last (checked against cv2.gemm 4.13 on 20 000 random
// 3x3 * 3x1 + 3x1 and 3x3 * 3x3 cases: identical every time; a double accumulator matches only 14-19 % of them);
// cv::norm accumulates squares in double.  Expressions are evaluated operator by operator, which for A*B+C gives the
// same float operations in the same order as the fused GEMM.
inline Mat operator*(const Mat& a, const Mat& b) {
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
    inline Mat operator OP(const Mat& a, const Mat& b) {                                           \
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
inline Mat operator*(const Mat& a, double s) {
    Mat c(a.rows, a.cols, CV_32F);
    const float f = (float)s;
    for (int i = 0; i < a.rows; ++i)
        for (int j = 0; j < a.cols; ++j) c.at<float>(i, j) = a.at<float>(i, j) * f;
    return c;
}
inline Mat operator*(double s, const Mat& a) { return a * s; }
inline Mat operator/(const Mat& a, double s) { return a * (1.0 / s); }
inline Mat operator-(const Mat& a) { return a * -1.0; }
inline Mat Mat::t() const {
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
    Matx(T a, T b, T c) { static_assert(M * N == 3, "three-element Matx"); val[0] = a; val[1] = b; val[2] = c; }
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
typedef Matx<float, 3, 3> Matx33f;
typedef Matx<float, 3, 1> Matx31f;
typedef Matx<float, 4, 4> Matx44f;

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
