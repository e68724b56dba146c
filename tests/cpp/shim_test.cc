// shim_test.cc — drives the C++ shim classes (include/orbfe_shim/) the way Frame.cc drives the
// reference's ORBextractor / ORBmatcher, with the OpenCV stand-in types (TEST INFRASTRUCTURE).
//
//   shim_test <left.raw> <right.raw> <w> <h> <nfeatures> <bf> <b> <out.bin> [<vocabulary.txt>]
//
// Writes: int32 nL, monoL, nR | nL keypoints (28 B) | nL x 32 descriptors | nL floats mvuRight | nL floats
// mvDepth | int32 nmatchesBoW | nR int32 (SearchByBoW KF=left, Frame=right with the node function of
// orbfe/synth.py and every left feature holding a MapPoint) | 8 floats scale factors | pyramid level 1 ROI bytes.
// Then: nL x (float x, y) undistorted left keypoints | int32 nproj | nR int32 window assignment | 3 int32 distinctive rows.
// With a vocabulary file: | int32 nwords, nnodes | nwords x (uint32 word, double value) | per node: uint32 id,
// int32 count, count x uint32 feature (ORBVocabulary::transform of the left descriptors, levelsup 4).
// Exit code 3 = no CUDA device (the CPU-side test expects exactly that), 0 = ok.
#define ORBFE_NO_OPENCV 1
#define ORBFE_NO_DBOW2 1
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <vector>

#include "ORBextractor.h"
#include "ORBVocabulary.h"
#include "ORBmatcher.h"

using namespace ORB_SLAM3;

static std::vector<unsigned char> read_file(const char* p, size_t n) {
    std::vector<unsigned char> v(n);
    FILE* f = fopen(p, "rb");
    if (!f || fread(v.data(), 1, n, f) != n) { fprintf(stderr, "cannot read %s\n", p); exit(2); }
    fclose(f);
    return v;
}

static std::map<unsigned, std::vector<unsigned>> feature_vector(const cv::Mat& d) {
    std::map<unsigned, std::vector<unsigned>> fv;
    for (int i = 0; i < d.rows; ++i) fv[1 + 10 * (d.ptr(i)[0] % 10) + (d.ptr(i)[1] % 10)].push_back(i);
    return fv;
}

int main(int argc, char** argv) {
    if (argc != 9 && argc != 10) return 2;
    const int w = atoi(argv[3]), h = atoi(argv[4]), nf = atoi(argv[5]);
    const float bf = (float)atof(argv[6]), b = (float)atof(argv[7]);
    std::vector<unsigned char> bl = read_file(argv[1], (size_t)w * h), br = read_file(argv[2], (size_t)w * h);
    cv::Mat imL(h, w, CV_8U, bl.data(), w), imR(h, w, CV_8U, br.data(), w);

    ORBextractor exL(nf, 1.2f, 8, 20, 7), exR(nf, 1.2f, 8, 20, 7);
    std::vector<cv::KeyPoint> kL, kR;
    cv::Mat dL, dR, none;
    std::vector<int> lap = {0, 0};
    const int monoL = exL(imL, none, kL, dL, lap);
    if (monoL < 0 && exL.lastStatus() == ORBFE_ERR_NO_DEVICE) {
        ORBVocabulary voc;           // must refuse as well, not fall back
        if (argc == 10 && (voc.loadFromTextFile(argv[9]) || voc.lastStatus() != ORBFE_ERR_NO_DEVICE)) return 1;
        printf("no CUDA device\n");
        return 3;
    }
    if (monoL < 0) { fprintf(stderr, "extract: %s\n", orbfe_strerror(exL.lastStatus())); return 1; }
    if (exR(imR, none, kR, dR, lap) < 0) return 1;
    cv::Mat empty;
    std::vector<cv::KeyPoint> kE;
    cv::Mat dE;
    if (exL.lastStatus() != ORBFE_OK || ORBextractor(nf, 1.2f, 8, 20, 7)(empty, none, kE, dE, lap) != -1) return 1;

    ORBmatcher matcher(0.7f, true);
    std::vector<float> uRight, depth;
    matcher.ComputeStereoMatches(exL.handle(), exR.handle(), bf, b, uRight, depth, (int)kL.size());
    if (matcher.lastStatus() != ORBFE_OK) { fprintf(stderr, "stereo: %s\n", orbfe_strerror(matcher.lastStatus())); return 1; }

    orbfe_shim::FeatureVectorCsr fvL(feature_vector(dL)), fvR(feature_vector(dR));
    std::vector<uint8_t> all(kL.size(), 1);
    orbfe_shim::Side a, f;
    a.desc = dL.data; a.n = dL.rows; a.fv = &fvL; a.mask = all.data(); a.kps = kL.data();
    f.desc = dR.data; f.n = dR.rows; f.fv = &fvR; f.kps = kR.data();
    std::vector<int32_t> bow;
    const int nb = matcher.SearchByBoWArrays(a, f, -1, bow);
    if (matcher.lastStatus() != ORBFE_OK) return 1;
    if (ORBmatcher::DescriptorDistance(dL.row(0), dL.row(0)) != 0) return 1;

    std::vector<float> sf = exL.GetScaleFactors();
    // like the reference's, the shim leaves mvImagePyramid filled after operator() (no explicit fetch): an unmodified
    // Frame::ComputeStereoMatches reads it right away (Frame.cc:801)
    if (exL.mvImagePyramid[0].rows != h || exL.mvImagePyramid[0].cols != w || exR.mvImagePyramid[0].rows != h) return 1;
    const cv::Mat& p1 = exL.mvImagePyramid[1];

    FILE* o = fopen(argv[8], "wb");
    if (!o) return 2;
    int32_t hdr[3] = {(int32_t)kL.size(), monoL, (int32_t)kR.size()};
    fwrite(hdr, 4, 3, o);
    fwrite(kL.data(), sizeof(cv::KeyPoint), kL.size(), o);
    for (int i = 0; i < dL.rows; ++i) fwrite(dL.ptr(i), 1, 32, o);
    fwrite(uRight.data(), 4, uRight.size(), o);
    fwrite(depth.data(), 4, depth.size(), o);
    int32_t nb32 = nb;
    fwrite(&nb32, 4, 1, o);
    fwrite(bow.data(), 4, bow.size(), o);
    fwrite(sf.data(), 4, 8, o);
    int32_t ps[2] = {p1.cols, p1.rows};
    fwrite(ps, 4, 2, o);
    for (int r = 0; r < p1.rows; ++r) fwrite(p1.ptr(r), 1, p1.cols, o);
    {   // Frame::UndistortKeyPoints with the EuRoC cam0 distortion; SearchByProjection windows around the left keypoints
        const float K4[4] = {458.654f, 457.296f, 367.215f, 248.375f}, D[4] = {-0.28340811f, 0.07395907f, 0.00019359f, 1.76187114e-05f};
        std::vector<cv::KeyPoint> kUn;
        if (!matcher.UndistortKeyPoints(kL, K4, D, 4, kUn)) return 1;
        for (const auto& k : kUn) { fwrite(&k.pt.x, 4, 1, o); fwrite(&k.pt.y, 4, 1, o); }
        orbfe_grid_frame F = {};
        F.kps = reinterpret_cast<const orbfe_keypoint*>(kR.data()); F.desc = dR.ptr(); F.n = dR.rows; F.nleft = -1;
        F.min_x = 0.f; F.min_y = 0.f; F.grid_w_inv = 64.f / (float)w; F.grid_h_inv = 48.f / (float)h;
        std::vector<orbfe_proj_query> win(kL.size());
        for (size_t i = 0; i < kL.size(); ++i) {       // the right image's content sits to the left: search around (x - 12, y)
            orbfe_proj_query& q = win[i];
            memset(&q, 0, sizeof(q));
            q.u = kL[i].pt.x - 12.f; q.v = kL[i].pt.y; q.r = 30.f; q.min_level = (int16_t)(kL[i].octave - 1); q.max_level = (int16_t)(kL[i].octave + 1);
            q.angle = kL[i].angle; q.obs = 1; q.valid = 1;
            memcpy(q.desc, dL.ptr((int)i), 32);
        }
        std::vector<int32_t> assign;
        const int32_t np = matcher.SearchByProjectionWindows(F, win, 0, ORBmatcher::TH_HIGH, false, true, assign);
        if (matcher.lastStatus() != ORBFE_OK) return 1;
        fwrite(&np, 4, 1, o);
        fwrite(assign.data(), 4, assign.size(), o);
        std::vector<int32_t> offs = {0, std::min(5, dL.rows), std::min(5, dL.rows), std::min(40, dL.rows)}, best;
        if (!matcher.ComputeDistinctiveDescriptors(dL, offs, best) || best.size() != 3 || best[1] != -1) return 1;
        fwrite(best.data(), 4, 3, o);
        // SearchForInitialization: left frame -> right frame, windows of 40 px around the left positions
        std::vector<cv::Point2f> prev(kL.size());
        for (size_t i = 0; i < kL.size(); ++i) prev[i] = kL[i].pt;
        std::vector<int> m12;
        const int32_t ni = matcher.SearchForInitializationArrays(kL, dL, F, prev, m12, 40);
        if (matcher.lastStatus() != ORBFE_OK || m12.size() != kL.size()) return 1;
        fwrite(&ni, 4, 1, o);
        fwrite(m12.data(), 4, m12.size(), o);
        fwrite(prev.data(), 8, prev.size(), o);
    }
    if (argc == 10) {
        ORBVocabulary voc;
        if (!voc.loadFromTextFile(argv[9]) || voc.empty()) { fprintf(stderr, "vocabulary: %s\n", orbfe_strerror(voc.lastStatus())); return 1; }
        std::vector<cv::Mat> rows;
        for (int i = 0; i < dL.rows; ++i) rows.push_back(dL.row(i));
        DBoW2::BowVector bv;
        DBoW2::FeatureVector fvv;
        voc.transform(rows, bv, fvv, 4);
        if (voc.lastStatus() != ORBFE_OK) return 1;
        int32_t cnt[2] = {(int32_t)bv.size(), (int32_t)fvv.size()};
        fwrite(cnt, 4, 2, o);
        for (const auto& kv : bv) { uint32_t w = kv.first; double x = kv.second; fwrite(&w, 4, 1, o); fwrite(&x, 8, 1, o); }
        for (const auto& kv : fvv) {
            uint32_t id = kv.first; int32_t c = (int32_t)kv.second.size();
            fwrite(&id, 4, 1, o); fwrite(&c, 4, 1, o); fwrite(kv.second.data(), 4, kv.second.size(), o);
        }
    }
    fclose(o);
    printf("shim ok: %zu left keypoints, %d stereo depths, %d BoW matches\n", kL.size(),
           (int)std::count_if(depth.begin(), depth.end(), [](float d) { return d > 0; }), nb);
    return 0;
}
