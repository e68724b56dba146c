// ORBextractor.h — drop-in ORB_SLAM3::ORBextractor over the orbfe C ABI (include/orbfe.h).
//
// Same class name, constructor, operator(), getters and public mvImagePyramid member as the
// reference's include/ORBextractor.h:43-109, so Frame.cc / Tracking.cc compile unchanged when this
// directory precedes the reference's include/ on the include path.  All pixel work happens in
// liborbfe.so on the GPU; this header only moves bytes across the ABI.
//
// Builds against real OpenCV by default.  Define ORBFE_NO_OPENCV to use the small stand-in types
// of cv_standin.h (this repository's own tests: the image has no OpenCV C++ headers).
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <cstring>
#include <vector>

#ifdef ORBFE_NO_OPENCV
#include "cv_standin.h"
#else
#include <opencv2/core/core.hpp>
#endif

#include "../orbfe.h"

namespace ORB_SLAM3 {

namespace orbfe_shim {
#ifdef ORBFE_NO_OPENCV
inline cv::Mat get_mat(cv::InputArray a) { return a; }
inline cv::Mat create_out(cv::OutputArray a, int rows, int cols) { a.create(rows, cols, CV_8U); return a; }
inline void release_out(cv::OutputArray a) { a.release(); }
inline cv::Mat roi_view(const cv::Mat& parent, int x, int y, int w, int h) { return cv::Mat(parent, x, y, w, h); }
#else
inline cv::Mat get_mat(cv::InputArray a) { return a.getMat(); }
inline cv::Mat create_out(cv::OutputArray a, int rows, int cols) { a.create(rows, cols, CV_8U); return a.getMat(); }
inline void release_out(cv::OutputArray a) { a.release(); }
inline cv::Mat roi_view(const cv::Mat& parent, int x, int y, int w, int h) { return parent(cv::Rect(x, y, w, h)); }
#endif
}  // namespace orbfe_shim

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    ORBextractor(int nfeatures_, float scaleFactor_, int nlevels_, int iniThFAST_, int minThFAST_, int device = 0)
        : nfeatures(nfeatures_), scaleFactor(scaleFactor_), nlevels(nlevels_), iniThFAST(iniThFAST_), minThFAST(minThFAST_),
          device_(device) {
        mvImagePyramid.resize(nlevels);
    }

    ~ORBextractor() { orbfe_destroy(h_); }
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    // ORBextractor.cc:1068-1150.  The mask is ignored, as in the reference.
    int operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints,
                   cv::OutputArray _descriptors, std::vector<int>& vLappingArea) {
        cv::Mat image = orbfe_shim::get_mat(_image);
        if (image.empty()) return -1;
        if (!ensure(image.cols, image.rows)) return -1;
        const int cap = orbfe_max_keypoints(h_);
        kps_.resize(cap);
        desc_.resize((size_t)cap * 32);
        int n = 0;
        const int lap0 = vLappingArea.size() > 0 ? vLappingArea[0] : 0, lap1 = vLappingArea.size() > 1 ? vLappingArea[1] : 0;
        const int mono = orbfe_extract(h_, image.ptr(), image.cols, image.rows, (ptrdiff_t)image.step, lap0, lap1, kps_.data(),
                                       desc_.data(), cap, &n);
        pyramidValid_ = false;
        if (mono < 0) { lastStatus_ = mono; _keypoints.clear(); orbfe_shim::release_out(_descriptors); return -1; }
        lastStatus_ = ORBFE_OK;
        // the reference leaves mvImagePyramid filled after every call and Frame::ComputeStereoMatches (Frame.cc:801,891-906)
        // reads it on the host: by default so do we (one batched copy, ~1.3 MB for a 752x480 image)
        if (hostPyramid_ && !FetchImagePyramid()) { _keypoints.clear(); orbfe_shim::release_out(_descriptors); return -1; }
        static_assert(sizeof(cv::KeyPoint) == sizeof(orbfe_keypoint), "cv::KeyPoint must be 28 bytes");
        _keypoints.resize(n);
        if (n == 0) { orbfe_shim::release_out(_descriptors); return mono; }
        std::memcpy((void*)_keypoints.data(), kps_.data(), sizeof(orbfe_keypoint) * (size_t)n);
        cv::Mat d = orbfe_shim::create_out(_descriptors, n, 32);
        for (int i = 0; i < n; ++i) std::memcpy(d.ptr(i), desc_.data() + (size_t)i * 32, 32);
        return mono;
    }

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return (float)scaleFactor; }
    std::vector<float> inline GetScaleFactors() { tables(); return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { tables(); return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { tables(); return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { tables(); return mvInvLevelSigma2; }

    // Public member of the reference (ORBextractor.h:83), read by Frame::ComputeStereoMatches on the host.  The pyramid is
    // computed and kept on the GPU; after every operator() its bordered levels are copied to the host and mvImagePyramid[l]
    // points at the ROI inside each, exactly like the reference's cv::Mat views (ORBextractor.cc:1160-1163), so unmodified
    // host readers work.  An integration that routes Frame::ComputeStereoMatches to orbfe_stereo_match (which reads the
    // device-resident pyramid) switches the copy off with SetHostPyramid(false) and may still call FetchImagePyramid().
    std::vector<cv::Mat> mvImagePyramid;
    void SetHostPyramid(bool on) { hostPyramid_ = on; }

    bool FetchImagePyramid() {
        if (!h_) return false;
        if (pyramidValid_) return true;
        parents_.resize(nlevels);
        std::vector<uint8_t*> dst(nlevels);
        std::vector<ptrdiff_t> strides(nlevels);
        std::vector<int> w(nlevels), h(nlevels);
        for (int l = 0; l < nlevels; ++l) {
            if (orbfe_get_level_size(h_, l, &w[l], &h[l]) < 0) return false;
            parents_[l].create(h[l] + 38, w[l] + 38, CV_8U);
            dst[l] = parents_[l].ptr();
            strides[l] = (ptrdiff_t)parents_[l].step;
        }
        lastStatus_ = orbfe_get_pyramid(h_, 0, dst.data(), strides.data());
        if (lastStatus_ < 0) return false;
        for (int l = 0; l < nlevels; ++l) mvImagePyramid[l] = orbfe_shim::roi_view(parents_[l], 19, 19, w[l], h[l]);
        pyramidValid_ = true;
        return true;
    }

    orbfe_handle* handle() { return h_; }
    int lastStatus() const { return lastStatus_; }

protected:
    // The handle is created on the first image (the reference's ctor does not know the image size)
    // and re-created if a larger image arrives.
    bool ensure(int w, int h) {
        if (h_ && w <= maxW_ && h <= maxH_) return true;
        orbfe_destroy(h_);
        h_ = nullptr;
        orbfe_params p;
        p.nfeatures = nfeatures; p.scale_factor = (float)scaleFactor; p.nlevels = nlevels; p.ini_th_fast = iniThFAST;
        p.min_th_fast = minThFAST; p.device = device_; p.max_width = w > maxW_ ? w : maxW_; p.max_height = h > maxH_ ? h : maxH_;
        p.max_batch = 1;
        lastStatus_ = orbfe_create(&p, &h_);
        if (lastStatus_ != ORBFE_OK) { h_ = nullptr; return false; }
        maxW_ = p.max_width; maxH_ = p.max_height;
        return true;
    }

    // ORBextractor.cc:413-444: the tables depend only on the five constructor scalars; they are read
    // back from the library (which owns the float chain) the first time someone asks.
    void tables() {
        if (!mvScaleFactor.empty()) return;
        if (!h_ && !ensure(752, 480)) return;   // any shape whose top level holds a FAST cell
        mvScaleFactor.resize(nlevels); mvInvScaleFactor.resize(nlevels); mvLevelSigma2.resize(nlevels);
        mvInvLevelSigma2.resize(nlevels); mnFeaturesPerLevel.resize(nlevels);
        orbfe_get_tables(h_, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(), mvInvLevelSigma2.data(),
                         mnFeaturesPerLevel.data());
    }

    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;
    std::vector<int> mnFeaturesPerLevel;
    std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;

private:
    int device_;
    orbfe_handle* h_ = nullptr;
    int maxW_ = 0, maxH_ = 0, lastStatus_ = ORBFE_OK;
    bool pyramidValid_ = false, hostPyramid_ = true;
    std::vector<orbfe_keypoint> kps_;
    std::vector<uint8_t> desc_;
    std::vector<cv::Mat> parents_;
};

}  // namespace ORB_SLAM3

#endif  // ORBEXTRACTOR_H
