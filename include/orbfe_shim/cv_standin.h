// cv_standin.h — the few OpenCV types the ORBextractor / ORBmatcher shims touch, for builds WITHOUT
// OpenCV (this repository's tests: the image has no OpenCV C++ headers).  When ORB-SLAM3 is built,
// the real <opencv2/core/core.hpp> is included instead and this file is not used.  Layouts match
// OpenCV's: cv::KeyPoint is 28 bytes (pt, size, angle, response, octave, class_id).
#pragma once
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#define CV_8U 0
#define CV_8UC1 0

namespace cv {

typedef unsigned char uchar;

struct Point2f {
    float x, y;
};

struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout");

struct Size {
    int width, height;
};

// A reference-counted 8-bit single-channel matrix or a view into one (what mvImagePyramid holds).
class Mat {
public:
    int rows = 0, cols = 0;
    size_t step = 0;
    uchar* data = nullptr;

    Mat() {}
    Mat(int r, int c, int /*type*/) { create(r, c, CV_8U); }
    // view of rows [y0, y0+h) x cols [x0, x0+w) of a parent
    Mat(const Mat& parent, int x0, int y0, int w, int h)
        : rows(h), cols(w), step(parent.step), data(parent.data + (size_t)y0 * parent.step + x0), owner_(parent.owner_) {}
    // wraps caller memory (no ownership), like cv::Mat(rows, cols, type, ptr, step)
    Mat(int r, int c, int /*type*/, void* p, size_t s) : rows(r), cols(c), step(s), data((uchar*)p) {}

    void create(int r, int c, int /*type*/) {
        if (r == rows && c == cols && owner_ && step == (size_t)c) return;
        owner_ = std::make_shared<std::vector<uchar>>((size_t)r * c);
        rows = r; cols = c; step = (size_t)c;
        data = owner_->data();
    }
    void release() { owner_.reset(); rows = cols = 0; step = 0; data = nullptr; }
    bool empty() const { return rows == 0 || cols == 0 || !data; }
    int type() const { return CV_8UC1; }
    bool isContinuous() const { return step == (size_t)cols; }
    uchar* ptr(int r = 0) { return data + (size_t)r * step; }
    const uchar* ptr(int r = 0) const { return data + (size_t)r * step; }
    template <class T> T* ptr(int r = 0) { return (T*)(data + (size_t)r * step); }
    template <class T> const T* ptr(int r = 0) const { return (const T*)(data + (size_t)r * step); }
    Mat row(int r) const { return Mat(*this, 0, r, cols, 1); }
    Size size() const { return Size{cols, rows}; }

private:
    std::shared_ptr<std::vector<uchar>> owner_;
};

typedef const Mat& InputArray;
typedef Mat& OutputArray;

}  // namespace cv
