// TEST INFRASTRUCTURE (oracle/_ref): C entry points over the reference's own src/Frame.cc, src/MapPoint.cc,
// src/CameraModels/{Pinhole,KannalaBrandt8}.cpp and src/ORBextractor.cc, all compiled unmodified against
// refshim/frame_standin.h + refshim/opencv2.  ORBmatcher::DescriptorDistance / TH_HIGH / TH_LOW come from
// liborbslam3_orbmatcher.so (the reference's own ORBmatcher.cc).  Nothing here is linked into the product.
#define private public        // Frame::UndistortKeyPoints / AssignFeaturesToGrid / ComputeImageBounds are private members
#include "Frame.h"
#undef private
#include "KannalaBrandt8.h"
#include "MapPoint.h"
#include "ORBextractor.h"
#include "ORBVocabulary.h"
#include "Pinhole.h"

#include <cstdint>
#include <cstring>

#include "../cpu_ref/match_ref.h"

using namespace ORB_SLAM3;

// ---- a monotonic heap for this library only (linked -Bsymbolic), as in ref_capi.cc and for the same reason: DistributeOctTree
// orders equal-size nodes by their address (ORBextractor.cc:682) and the oracle / CUDA path define that tie as "created later
// compares greater".  Thread-safe (Frame's stereo constructor extracts the two eyes on two threads: each thread's own nodes
// still get increasing addresses); every C entry point below rewinds to where it started, all of its allocations being dead.
#include <sys/mman.h>

#include <atomic>
#include <new>
#ifdef ORBSLAM3_REF_SYSTEM_HEAP     // the build against the orbfe shim: no reference quadtree in this library, nothing to pin
namespace { struct HeapScope {}; }
#else
namespace {
const size_t kArena = (size_t)48 << 30;      // virtual; pages are touched on demand
char* arenaBase() {
    static char* base = [] {
        void* p = mmap(nullptr, kArena, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        return p == MAP_FAILED ? nullptr : (char*)p;
    }();
    return base;
}
std::atomic<size_t> g_top{0};
void* bump(size_t n) {
    char* base = arenaBase();
    const size_t need = ((n ? n : 1) + 15) & ~(size_t)15;
    const size_t at = g_top.fetch_add(need);
    if (!base || at + need > kArena) throw std::bad_alloc();
    return base + at;
}
struct HeapScope {
    size_t mark;
    HeapScope() : mark(g_top.load()) {}
    ~HeapScope() {
        const size_t top = g_top.load();
        if (top > mark) {
            const size_t a = (mark + 4095) & ~(size_t)4095, b = (top + 4095) & ~(size_t)4095;
            if (b > a) madvise(arenaBase() + a, b - a, MADV_DONTNEED);
            g_top.store(mark);
        }
    }
};
}  // namespace
void* operator new(size_t n) { return bump(n); }
void* operator new[](size_t n) { return bump(n); }
void operator delete(void*) noexcept {}
void operator delete[](void*) noexcept {}
void operator delete(void*, size_t) noexcept {}
void operator delete[](void*, size_t) noexcept {}
#endif

// cv::undistortPoints of the stand-in forwards here: oracle/cpu_ref's restatement, bit-identical to cv2.undistortPoints 4.13
namespace cv {
void orbref_undistort_points_hook(const float* xy, int n, const float K4[4], const float* dist, int ndist, const float P4[4], float* out) {
    assert(!std::memcmp(K4, P4, 4 * sizeof(float)));      // Frame.cc:749 passes the same intrinsics as camera matrix and as P
    std::vector<orbref::KeyPoint> in(n), o(n);
    for (int i = 0; i < n; ++i) { in[i].x = xy[2 * i]; in[i].y = xy[2 * i + 1]; }
    orbref::undistortKeyPoints(in.data(), n, K4, dist, ndist, o.data());
    for (int i = 0; i < n; ++i) { out[2 * i] = o[i].x; out[2 * i + 1] = o[i].y; }
}
}  // namespace cv

namespace {

std::vector<cv::KeyPoint> keypoints(const void* kps, int n) {
    std::vector<cv::KeyPoint> v(n);
    if (n) std::memcpy(v.data(), kps, (size_t)n * sizeof(cv::KeyPoint));
    return v;
}
cv::Mat descriptors(const uint8_t* d, int n) { return n ? cv::Mat(n, 32, CV_8U, const_cast<uint8_t*>(d), 32).clone() : cv::Mat(0, 32, CV_8U); }

// The tables a Frame copies from its extractor (Frame.cc:107-113)
void scaleTables(Frame& F, ORBextractor& ex) {
    F.mnScaleLevels = ex.GetLevels();
    F.mfScaleFactor = ex.GetScaleFactor();
    F.mfLogScaleFactor = log(F.mfScaleFactor);
    F.mvScaleFactors = ex.GetScaleFactors();
    F.mvInvScaleFactors = ex.GetInverseScaleFactors();
    F.mvLevelSigma2 = ex.GetScaleSigmaSquares();
    F.mvInvLevelSigma2 = ex.GetInverseScaleSigmaSquares();
}

// bordered level buffers (19 px each side, as ORBextractor::ComputePyramid leaves them) -> mvImagePyramid ROI views
void setPyramid(ORBextractor& ex, const uint8_t* const* levels, const int32_t* w, const int32_t* h) {
    for (int l = 0; l < ex.GetLevels(); ++l) {
        const size_t stride = (size_t)w[l] + 38;
        ex.mvImagePyramid[l] = cv::Mat(h[l], w[l], CV_8U, const_cast<uint8_t*>(levels[l]) + 19 * stride + 19, stride);
    }
}

}  // namespace

extern "C" {

// Frame::ComputeStereoMatches (Frame.cc:794-964) on a Frame holding the two key point sets and extractors whose mvImagePyramid
// are the given levels.  u_right / depth: mvuRight / mvDepth (N floats).
void orbslam3_frame_stereo_matches(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th,
                                   const uint8_t* const* levelsL, const uint8_t* const* levelsR, const int32_t* w, const int32_t* h,
                                   const void* kpL, const uint8_t* descL, int nL, const void* kpR, const uint8_t* descR, int nR,
                                   float bf, float b, float* u_right, float* depth) {
    HeapScope heap_scope;
    ORBextractor exL(nfeatures, scale_factor, nlevels, ini_th, min_th), exR(nfeatures, scale_factor, nlevels, ini_th, min_th);
    setPyramid(exL, levelsL, w, h);
    setPyramid(exR, levelsR, w, h);
    Frame F;
    F.mpORBextractorLeft = &exL;
    F.mpORBextractorRight = &exR;
    scaleTables(F, exL);
    F.mvKeys = keypoints(kpL, nL);
    F.mvKeysRight = keypoints(kpR, nR);
    F.mDescriptors = descriptors(descL, nL);
    F.mDescriptorsRight = descriptors(descR, nR);
    F.N = nL;
    F.mbf = bf;
    F.mb = b;
    F.ComputeStereoMatches();
    for (int i = 0; i < nL; ++i) { u_right[i] = F.mvuRight[i]; depth[i] = F.mvDepth[i]; }
}

// Frame::ComputeStereoFishEyeMatches (Frame.cc:1116-1156) with the reference's own KannalaBrandt8::TriangulateMatches
// (KannalaBrandt8.cpp:338-407).  cam1 / cam2: fx fy cx cy k1 k2 k3 k4.  R12 / t12 = mRlr / mtlr.  Returns the match count.
int orbslam3_frame_fisheye_matches(const float* cam1, const float* cam2, const void* kpL, const uint8_t* descL, int nL, int mono_left,
                                   const void* kpR, const uint8_t* descR, int nR, int mono_right, const float* R12, const float* t12,
                                   const float* sigma2, int nlevels, int32_t* l2r, int32_t* r2l, float* depth, float* p3d) {
    HeapScope heap_scope;
    KannalaBrandt8 c1(std::vector<float>(cam1, cam1 + 8)), c2(std::vector<float>(cam2, cam2 + 8));
    Frame F;
    F.mpCamera = &c1;
    F.mpCamera2 = &c2;
    F.mvKeys = keypoints(kpL, nL);
    F.mvKeysRight = keypoints(kpR, nR);
    F.mDescriptors = descriptors(descL, nL);
    F.mDescriptorsRight = descriptors(descR, nR);
    F.Nleft = nL;
    F.Nright = nR;
    F.monoLeft = mono_left;
    F.monoRight = mono_right;
    F.mvLevelSigma2.assign(sigma2, sigma2 + nlevels);
    F.mRlr = cv::Mat(3, 3, CV_32F, const_cast<float*>(R12), 12).clone();
    F.mtlr = cv::Mat(3, 1, CV_32F, const_cast<float*>(t12), 4).clone();
    F.ComputeStereoFishEyeMatches();
    int n = 0;
    for (int i = 0; i < nL; ++i) {
        l2r[i] = F.mvLeftToRightMatch[i];
        depth[i] = F.mvDepth[i];
        const cv::Mat& p = F.mvStereo3Dpoints[i];
        for (int k = 0; k < 3; ++k) p3d[3 * i + k] = p.empty() ? 0.f : p.at<float>(k);
        n += l2r[i] >= 0;
    }
    for (int i = 0; i < nR; ++i) r2l[i] = F.mvRightToLeftMatch[i];
    return n;
}

// Frame::UndistortKeyPoints (Frame.cc:730-763) with a Pinhole camera: mvKeys -> mvKeysUn.
void orbslam3_frame_undistort_keypoints(const void* kps, int n, const float* K4, const float* dist, int ndist, void* out) {
    HeapScope heap_scope;
    Pinhole cam(std::vector<float>(K4, K4 + 4));
    Frame F;
    F.mpCamera = &cam;
    F.mK = cam.toK();
    F.mDistCoef = cv::Mat(ndist, 1, CV_32F, const_cast<float*>(dist), 4).clone();
    F.mvKeys = keypoints(kps, n);
    F.N = n;
    F.UndistortKeyPoints();
    if (n) std::memcpy(out, F.mvKeysUn.data(), (size_t)n * sizeof(cv::KeyPoint));
}

// Frame::AssignFeaturesToGrid (Frame.cc:380-411) + PosInGrid (:708-719) + GetFeaturesInArea (:640-706).  kps: mvKeysUn (nleft ==
// -1) or mvKeys ++ mvKeysRight.  Window q = win[6q .. 6q+5] = (x, y, r, minLevel, maxLevel, bRight); the indices it returns are
// appended to `out` (right-camera indices as the reference returns them: relative to Nleft), counts[q] = how many.
int orbslam3_frame_features_in_area(const void* kps, int n, int nleft, float min_x, float min_y, float grid_w_inv, float grid_h_inv,
                                    const float* win, int nwin, int32_t* out, int out_cap, int32_t* counts) {
    HeapScope heap_scope;
    Frame F;
    Frame::mnMinX = min_x;
    Frame::mnMinY = min_y;
    Frame::mfGridElementWidthInv = grid_w_inv;
    Frame::mfGridElementHeightInv = grid_h_inv;
    const std::vector<cv::KeyPoint> all = keypoints(kps, n);
    F.N = n;
    F.Nleft = nleft;
    if (nleft == -1) F.mvKeysUn = all;
    else { F.mvKeys.assign(all.begin(), all.begin() + nleft); F.mvKeysRight.assign(all.begin() + nleft, all.end()); F.Nright = n - nleft; }
    F.AssignFeaturesToGrid();
    int total = 0;
    for (int q = 0; q < nwin; ++q) {
        const float* w = win + 6 * q;
        const std::vector<size_t> v = F.GetFeaturesInArea(w[0], w[1], w[2], (int)w[3], (int)w[4], w[5] != 0.f);
        counts[q] = (int)v.size();
        for (size_t i : v) { if (total < out_cap) out[total] = (int32_t)i; ++total; }
    }
    return total;
}

// KannalaBrandt8::project / unproject (KannalaBrandt8.cpp:24-41, 96-123) on n points.
void orbslam3_kb8_project(const float* cam, const float* p3d, int n, float* uv) {
    HeapScope heap_scope;
    KannalaBrandt8 c(std::vector<float>(cam, cam + 8));
    for (int i = 0; i < n; ++i) {
        const cv::Point2f q = c.project(cv::Point3f(p3d[3 * i], p3d[3 * i + 1], p3d[3 * i + 2]));
        uv[2 * i] = q.x; uv[2 * i + 1] = q.y;
    }
}
// The stereo constructor Frame(imLeft, imRight, ts, extractorLeft, extractorRight, voc, K, distCoef, bf, thDepth, pCamera)
// (Frame.cc:98-200) from two images: two extraction threads, UndistortKeyPoints, ComputeStereoMatches, image bounds, grid.
// `mb` is read by ComputeStereoMatches (:838) before the constructor assigns it (:176) and no initialiser mentions it: its value
// there is whatever the object's storage held.  The Frame is therefore constructed in storage where that member already
// holds bf / fx, the value the constructor itself assigns a few lines later.
// kps / kps_un / desc / u_right / depth: capacity cap (left eye); kps_r / desc_r: capacity cap (right eye).
// out_n[0] = N, out_n[1] = right key points.  bounds: mnMinX, mnMinY, mnMaxX, mnMaxY.  Returns 0, or -1 if cap is too small.
int orbslam3_frame_stereo_construct(const uint8_t* im_left, const uint8_t* im_right, int w, int h, int nfeatures, float scale_factor,
                                    int nlevels, int ini_th, int min_th, const float* K4, const float* dist, int ndist, float bf,
                                    float th_depth, int cap, void* kps, void* kps_un, uint8_t* desc, float* u_right, float* depth,
                                    void* kps_r, uint8_t* desc_r, int32_t* out_n, float* bounds) {
    HeapScope heap_scope;
    ORBextractor exL(nfeatures, scale_factor, nlevels, ini_th, min_th), exR(nfeatures, scale_factor, nlevels, ini_th, min_th);
    Pinhole cam(std::vector<float>(K4, K4 + 4));
    cv::Mat K = cam.toK();
    cv::Mat D = cv::Mat(ndist, 1, CV_32F, const_cast<float*>(dist), 4).clone();
    cv::Mat imL(h, w, CV_8UC1, const_cast<uint8_t*>(im_left), (size_t)w), imR(h, w, CV_8UC1, const_cast<uint8_t*>(im_right), (size_t)w);
    Frame::mbInitialComputations = true;
    alignas(64) static unsigned char storage[sizeof(Frame)];
    Frame* seed = new (storage) Frame();
    seed->mb = bf / K4[0];
    seed->~Frame();
    Frame* F = new (storage) Frame(imL, imR, 0.0, &exL, &exR, nullptr, K, D, bf, th_depth, &cam);
    int rc = 0;
    const int n = F->N, nr = (int)F->mvKeysRight.size();
    out_n[0] = n; out_n[1] = nr;
    if (n > cap || nr > cap) rc = -1;
    else {
        if (n) {
            std::memcpy(kps, F->mvKeys.data(), (size_t)n * sizeof(cv::KeyPoint));
            std::memcpy(kps_un, F->mvKeysUn.data(), (size_t)n * sizeof(cv::KeyPoint));
            for (int i = 0; i < n; ++i) { std::memcpy(desc + 32 * (size_t)i, F->mDescriptors.ptr(i), 32); u_right[i] = F->mvuRight[i]; depth[i] = F->mvDepth[i]; }
        }
        if (nr) std::memcpy(kps_r, F->mvKeysRight.data(), (size_t)nr * sizeof(cv::KeyPoint));
        for (int i = 0; i < nr; ++i) std::memcpy(desc_r + 32 * (size_t)i, F->mDescriptorsRight.ptr(i), 32);
        bounds[0] = Frame::mnMinX; bounds[1] = Frame::mnMinY; bounds[2] = Frame::mnMaxX; bounds[3] = Frame::mnMaxY;
    }
    F->~Frame();
    return rc;
}

// The stereo-fisheye constructor Frame(imLeft, imRight, ts, extractorLeft, extractorRight, voc, K, distCoef, bf, thDepth, pCamera,
// pCamera2, Tlr) (Frame.cc:1017-1114): extraction of both eyes with their lapping areas on two threads, ComputeStereoFishEyeMatches,
// descriptors stacked, grid, UndistortKeyPoints (zero distortion: a copy).  kps: mvKeys ++ mvKeysRight, desc: the stacked matrix
// (capacity cap rows each); out_n = Nleft, Nright, monoLeft, monoRight; l2r / depth / p3d: Nleft entries, r2l: Nright.
int orbslam3_frame_fisheye_construct(const uint8_t* im_left, const uint8_t* im_right, int w, int h, int nfeatures, float scale_factor,
                                     int nlevels, int ini_th, int min_th, const float* cam1, const float* cam2, const int32_t* lap_left,
                                     const int32_t* lap_right, const float* Tlr34, float bf, float th_depth, int cap, void* kps, uint8_t* desc,
                                     int32_t* out_n, int32_t* l2r, int32_t* r2l, float* depth, float* p3d) {
    HeapScope heap_scope;
    ORBextractor exL(nfeatures, scale_factor, nlevels, ini_th, min_th), exR(nfeatures, scale_factor, nlevels, ini_th, min_th);
    KannalaBrandt8 c1(std::vector<float>(cam1, cam1 + 8)), c2(std::vector<float>(cam2, cam2 + 8));
    c1.mvLappingArea[0] = lap_left[0]; c1.mvLappingArea[1] = lap_left[1];
    c2.mvLappingArea[0] = lap_right[0]; c2.mvLappingArea[1] = lap_right[1];
    cv::Mat K = c1.toK();
    cv::Mat D = cv::Mat::zeros(4, 1, CV_32F);                    // Tracking.cc: fisheye rigs carry zero mDistCoef
    cv::Mat Tlr = cv::Mat(3, 4, CV_32F, const_cast<float*>(Tlr34), 16).clone();
    cv::Mat imL(h, w, CV_8UC1, const_cast<uint8_t*>(im_left), (size_t)w), imR(h, w, CV_8UC1, const_cast<uint8_t*>(im_right), (size_t)w);
    Frame::mbInitialComputations = true;
    Frame F(imL, imR, 0.0, &exL, &exR, nullptr, K, D, bf, th_depth, &c1, &c2, Tlr);
    const int nl = F.Nleft, nr = F.Nright;
    out_n[0] = nl; out_n[1] = nr; out_n[2] = F.monoLeft; out_n[3] = F.monoRight;
    if (nl + nr > cap) return -1;
    if (nl) std::memcpy(kps, F.mvKeys.data(), (size_t)nl * sizeof(cv::KeyPoint));
    if (nr) std::memcpy((char*)kps + (size_t)nl * sizeof(cv::KeyPoint), F.mvKeysRight.data(), (size_t)nr * sizeof(cv::KeyPoint));
    for (int i = 0; i < F.mDescriptors.rows; ++i) std::memcpy(desc + 32 * (size_t)i, F.mDescriptors.ptr(i), 32);
    for (int i = 0; i < nl; ++i) {
        l2r[i] = F.mvLeftToRightMatch[i];
        depth[i] = F.mvDepth[i];
        const cv::Mat& p = F.mvStereo3Dpoints[i];
        for (int k = 0; k < 3; ++k) p3d[3 * i + k] = p.empty() ? 0.f : p.at<float>(k);
    }
    for (int i = 0; i < nr; ++i) r2l[i] = F.mvRightToLeftMatch[i];
    // the undistorted copy and the stacked descriptor matrix must be what the constructor promises
    if ((int)F.mvKeysUn.size() != nl || F.mDescriptors.rows != nl + nr) return -2;
    return 0;
}

// The monocular constructor Frame(imGray, ts, extractor, voc, pCamera, distCoef, bf, thDepth) (Frame.cc:284-378): ExtractORB with the
// lapping range (0, 1000), UndistortKeyPoints, bounds, grid; then Frame::ComputeBoW (:721-728) when voc_path names an ORBvoc text
// file (DBoW2's own loadFromTextFile + transform(features, mBowVec, mFeatVec, 4)).
// out_n: N, mono index, BowVector size, FeatureVector size.  words / values: capacity cap; nodes / offsets (+1) / indices: cap.
int orbslam3_frame_mono_construct(const uint8_t* im, int w, int h, int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th,
                                  const float* K4, const float* dist, int ndist, const char* voc_path, int cap, void* kps, void* kps_un,
                                  uint8_t* desc, int32_t* out_n, uint32_t* words, double* values, uint32_t* nodes, int32_t* offsets,
                                  int32_t* indices) {
    HeapScope heap_scope;
    ORBextractor ex(nfeatures, scale_factor, nlevels, ini_th, min_th);
    Pinhole cam(std::vector<float>(K4, K4 + 4));
    cv::Mat D = cv::Mat(ndist, 1, CV_32F, const_cast<float*>(dist), 4).clone();
    cv::Mat img(h, w, CV_8UC1, const_cast<uint8_t*>(im), (size_t)w);
    ORBVocabulary voc;
    if (voc_path && !voc.loadFromTextFile(voc_path)) return -3;
    Frame::mbInitialComputations = true;
    Frame F(img, 0.0, &ex, voc_path ? &voc : nullptr, &cam, D, 0.f, 0.f);
    const int n = F.N;
    out_n[0] = n; out_n[1] = F.monoLeft; out_n[2] = out_n[3] = 0;
    if (n > cap) return -1;
    if (n) {
        std::memcpy(kps, F.mvKeys.data(), (size_t)n * sizeof(cv::KeyPoint));
        std::memcpy(kps_un, F.mvKeysUn.data(), (size_t)n * sizeof(cv::KeyPoint));
        for (int i = 0; i < n; ++i) std::memcpy(desc + 32 * (size_t)i, F.mDescriptors.ptr(i), 32);
    }
    if (voc_path && n) {
        F.ComputeBoW();
        int k = 0;
        for (const auto& wv : F.mBowVec) { words[k] = wv.first; values[k] = wv.second; ++k; }
        out_n[2] = k;
        int nn = 0, at = 0;
        offsets[0] = 0;
        for (const auto& nf : F.mFeatVec) {
            nodes[nn] = nf.first;
            for (unsigned i : nf.second) indices[at++] = (int32_t)i;
            offsets[++nn] = at;
        }
        out_n[3] = nn;
    }
    return 0;
}

// KannalaBrandt8::epipolarConstrain_ (KannalaBrandt8.cpp:423-426) = TriangulateMatches_ (:409-480, the Matx code path that
// SearchForTriangulation_ reaches) > 0.0001, and the point it triangulates, per key point pair.
void orbslam3_kb8_epipolar_constrain_(const float* cam1, const float* cam2, const void* kps1, const void* kps2, int n, const float* R12,
                                      const float* t12, const float* sigma, const float* unc, uint8_t* ok, float* depth, float* p3d) {
    HeapScope heap_scope;
    KannalaBrandt8 c1(std::vector<float>(cam1, cam1 + 8)), c2(std::vector<float>(cam2, cam2 + 8));
    const std::vector<cv::KeyPoint> k1 = keypoints(kps1, n), k2 = keypoints(kps2, n);
    cv::Matx33f R(R12[0], R12[1], R12[2], R12[3], R12[4], R12[5], R12[6], R12[7], R12[8]);
    cv::Matx31f t(t12[0], t12[1], t12[2]);
    for (int i = 0; i < n; ++i) {
        ok[i] = c1.epipolarConstrain_(&c2, k1[i], k2[i], R, t, sigma[i], unc[i]) ? 1 : 0;
        cv::Matx31f p;
        depth[i] = c1.TriangulateMatches_(&c2, k1[i], k2[i], R, t, sigma[i], unc[i], p);
        for (int k = 0; k < 3; ++k) p3d[3 * i + k] = depth[i] > 0 ? p(k) : 0.f;
    }
}

// Pinhole::project (Pinhole.cpp:34-42) / unproject (:68-71) on n points.
void orbslam3_pinhole_project(const float* K4, const float* p3d, int n, float* uv) {
    HeapScope heap_scope;
    Pinhole c(std::vector<float>(K4, K4 + 4));
    for (int i = 0; i < n; ++i) {
        const cv::Point2f q = c.project(cv::Point3f(p3d[3 * i], p3d[3 * i + 1], p3d[3 * i + 2]));
        uv[2 * i] = q.x; uv[2 * i + 1] = q.y;
    }
}
void orbslam3_kb8_unproject(const float* cam, const float* uv, int n, float* ray) {
    HeapScope heap_scope;
    KannalaBrandt8 c(std::vector<float>(cam, cam + 8));
    for (int i = 0; i < n; ++i) {
        const cv::Point3f r = c.unproject(cv::Point2f(uv[2 * i], uv[2 * i + 1]));
        ray[3 * i] = r.x; ray[3 * i + 1] = r.y; ray[3 * i + 2] = r.z;
    }
}

}  // extern "C"
