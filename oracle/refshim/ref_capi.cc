// ref_capi.cc — C entry points around the reference's own ORB_SLAM3::ORBextractor, compiled from
// /root/reference/src/ORBextractor.cc against the OpenCV stand-in in oracle/refshim/opencv2 (TEST INFRASTRUCTURE).
// Used by tests/ to hold oracle/cpu_ref (and through it the CUDA path) to the reference's real control flow.
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include "ORBextractor.h"

// ---- a monotonic heap for this library only (linked -Bsymbolic, so the rest of the process keeps its allocator) ----
// DistributeOctTree sorts pair<int, ExtractorNode*> (ORBextractor.cc:682): equal sizes are ordered by the ADDRESS of
// the std::list node, i.e. by whatever the allocator returned.  That is not reproducible, so the oracle (and the CUDA
// path) define the tie rule as "the node created later compares greater" — what a heap that only grows gives.  To
// hold the reference's unmodified code to that rule, operator new inside this library hands out strictly increasing
// addresses from one reserved region and never reuses them; each extract() rewinds to the high-water mark left by
// the last constructor (everything an extraction allocates is dead by the next one).  Not thread-safe: tests only.
#ifndef ORBSLAM3_REF_SYSTEM_HEAP
#include <sys/mman.h>

#include <new>
namespace {
const size_t kArena = (size_t)48 << 30;      // virtual; pages are touched on demand
char* g_base = nullptr;
size_t g_top = 0, g_mark = 0;
void* bump(size_t n) {
    if (!g_base) {
        g_base = (char*)mmap(nullptr, kArena, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        if (g_base == (char*)MAP_FAILED) throw std::bad_alloc();
    }
    g_top = (g_top + 15) & ~(size_t)15;
    if (g_top + n > kArena) throw std::bad_alloc();
    void* p = g_base + g_top;
    g_top += n ? n : 1;
    return p;
}
}  // namespace
void* operator new(size_t n) { return bump(n); }
void* operator new[](size_t n) { return bump(n); }
void operator delete(void*) noexcept {}
void operator delete[](void*) noexcept {}
void operator delete(void*, size_t) noexcept {}
void operator delete[](void*, size_t) noexcept {}
static void heap_mark() { g_mark = g_top; }
static void heap_rewind() {
    if (g_top > g_mark) {                     // give the touched pages back, then reuse the addresses
        const size_t a = (g_mark + 4095) & ~(size_t)4095, b = (g_top + 4095) & ~(size_t)4095;
        if (b > a) madvise(g_base + a, b - a, MADV_DONTNEED);
        g_top = g_mark;
    }
}
#else       // the timing build (liborbslam3_orbextractor_mt.so): the process allocator, any number of threads
static void heap_mark() {}
static void heap_rewind() {}
#endif

extern "C" {
void* orbslam3_extractor_create(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST) {
    void* e = new ORB_SLAM3::ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST);
    heap_mark();
    return e;
}
void orbslam3_extractor_destroy(void* e) { delete static_cast<ORB_SLAM3::ORBextractor*>(e); }

// operator()(image, mask, keypoints, descriptors, vLappingArea).  kps: cap x 28 bytes, desc: cap x 32.
// Returns the number of keypoints (or -1 - n if cap is too small); *mono = the operator's return value.
int orbslam3_extractor_extract(void* e, const uint8_t* image, int w, int h, long stride, int lap0, int lap1, void* kps, uint8_t* desc,
                               int cap, int* mono) {
    ORB_SLAM3::ORBextractor& ex = *static_cast<ORB_SLAM3::ORBextractor*>(e);
    heap_rewind();
    cv::Mat img(h, w, CV_8UC1, const_cast<uint8_t*>(image), (size_t)stride), mask, descriptors;
    std::vector<cv::KeyPoint> keys;
    std::vector<int> lapping = {lap0, lap1};
    *mono = ex(img, mask, keys, descriptors, lapping);
    const int n = (int)keys.size();
    if (n > cap) return -1 - n;
    if (n > 0) {
        memcpy(kps, keys.data(), sizeof(cv::KeyPoint) * (size_t)n);
        for (int i = 0; i < n; ++i) memcpy(desc + 32 * (size_t)i, descriptors.ptr(i), 32);
    }
    return n;
}
// mvImagePyramid[level] (the ROI, without its border) after the last extract.
int orbslam3_extractor_pyramid(void* e, int level, uint8_t* dst, int cap, int* w, int* h) {
    ORB_SLAM3::ORBextractor& ex = *static_cast<ORB_SLAM3::ORBextractor*>(e);
    if (level < 0 || level >= (int)ex.mvImagePyramid.size()) return -1;
    const cv::Mat& m = ex.mvImagePyramid[level];
    *w = m.cols; *h = m.rows;
    if (m.cols * m.rows > cap) return -2;
    for (int y = 0; y < m.rows; ++y) memcpy(dst + (size_t)y * m.cols, m.ptr(y), (size_t)m.cols);
    return 0;
}
// The (rows + 38) x (cols + 38) buffer `temp` that mvImagePyramid[level] is a view of (ComputePyramid :1159-1160).
int orbslam3_extractor_pyramid_bordered(void* e, int level, uint8_t* dst, int cap, int* w, int* h) {
    ORB_SLAM3::ORBextractor& ex = *static_cast<ORB_SLAM3::ORBextractor*>(e);
    if (level < 0 || level >= (int)ex.mvImagePyramid.size()) return -1;
    const cv::Mat& m = ex.mvImagePyramid[level];
    const int B = 19, bw = m.cols + 2 * B, bh = m.rows + 2 * B;
    *w = bw; *h = bh;
    if ((int)m.step != bw) return -3;
    if (bw * bh > cap) return -2;
    memcpy(dst, m.data - (size_t)B * m.step - B, (size_t)bw * bh);
    return 0;
}
#ifdef ORBSLAM3_REF_SYSTEM_HEAP
// Frame-parallel timing loop: one ORBextractor per thread (instances are independent, as in Frame.cc:119-122).
long orbslam3_extract_many(int nf, float sf, int nl, int ini, int mn, const uint8_t* frames, int nframes, int w, int h, int lap0,
                           int lap1, int nthreads) {
    std::vector<long> counts(nthreads, 0);
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t)
        th.emplace_back([&, t]() {
            ORB_SLAM3::ORBextractor ex(nf, sf, nl, ini, mn);
            std::vector<int> lapping = {lap0, lap1};
            for (int f = t; f < nframes; f += nthreads) {
                cv::Mat img(h, w, CV_8UC1, const_cast<uint8_t*>(frames + (size_t)f * w * h), (size_t)w), mask, descriptors;
                std::vector<cv::KeyPoint> keys;
                ex(img, mask, keys, descriptors, lapping);
                counts[t] += (long)keys.size();
            }
        });
    for (auto& x : th) x.join();
    long total = 0;
    for (long c : counts) total += c;
    return total;
}
#endif
void orbslam3_extractor_tables(void* e, float* scale, float* invScale, float* sigma2, float* invSigma2) {
    ORB_SLAM3::ORBextractor& ex = *static_cast<ORB_SLAM3::ORBextractor*>(e);
    const int n = ex.GetLevels();
    memcpy(scale, ex.GetScaleFactors().data(), 4 * n);
    memcpy(invScale, ex.GetInverseScaleFactors().data(), 4 * n);
    memcpy(sigma2, ex.GetScaleSigmaSquares().data(), 4 * n);
    memcpy(invSigma2, ex.GetInverseScaleSigmaSquares().data(), 4 * n);
}
}
