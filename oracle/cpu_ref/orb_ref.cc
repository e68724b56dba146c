// orb_ref.cc — see orb_ref.h.  TEST INFRASTRUCTURE: oracle + timed CPU baseline only.
#include "orb_ref.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <list>
#include <utility>

namespace orbref {

static const int8_t kPattern[1024] = {
#include "../../orb-slam3-noted_b200/csrc/orb_pattern.inc"
};

static const int PATCH_SIZE = 31, HALF_PATCH_SIZE = 15, EDGE_THRESHOLD = 19;

// ORBextractor.cc:408-468
OrbExtractorRef::OrbExtractorRef(int nf, float sf, int nl, int ini, int mn)
    : nfeatures(nf), nlevels(nl), iniThFAST(ini), minThFAST(mn), scaleFactor(sf) {
    std::memcpy(pattern, kPattern, sizeof(pattern));
    mvScaleFactor.resize(nlevels);
    mvLevelSigma2.resize(nlevels);
    mvScaleFactor[0] = 1.0f;
    mvLevelSigma2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {
        mvScaleFactor[i] = (float)(mvScaleFactor[i - 1] * scaleFactor);   // float * double
        mvLevelSigma2[i] = mvScaleFactor[i] * mvScaleFactor[i];
    }
    mvInvScaleFactor.resize(nlevels);
    mvInvLevelSigma2.resize(nlevels);
    for (int i = 0; i < nlevels; i++) {
        mvInvScaleFactor[i] = 1.0f / mvScaleFactor[i];
        mvInvLevelSigma2[i] = 1.0f / mvLevelSigma2[i];
    }
    mnFeaturesPerLevel.resize(nlevels);
    float factor = (float)(1.0f / scaleFactor);
    float nDesired = nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int level = 0; level < nlevels - 1; level++) {
        mnFeaturesPerLevel[level] = cvRoundF(nDesired);
        sum += mnFeaturesPerLevel[level];
        nDesired *= factor;
    }
    mnFeaturesPerLevel[nlevels - 1] = std::max(nfeatures - sum, 0);

    umax.assign(HALF_PATCH_SIZE + 1, 0);
    int v, v0;
    const int vmax = (int)std::floor(HALF_PATCH_SIZE * std::sqrt(2.f) / 2 + 1);
    const int vmin = (int)std::ceil(HALF_PATCH_SIZE * std::sqrt(2.f) / 2);
    const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
    for (v = 0; v <= vmax; ++v) umax[v] = cvRoundD(std::sqrt(hp2 - v * v));
    for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
        while (umax[v0] == umax[v0 + 1]) ++v0;
        umax[v] = v0;
        ++v0;
    }
    pyramid.resize(nlevels);
}

// ORBextractor.cc:1152-1177
void OrbExtractorRef::computePyramid(const uint8_t* image, int w, int h, ptrdiff_t stride) {
    for (int level = 0; level < nlevels; ++level) {
        const float scale = mvInvScaleFactor[level];
        Level& L = pyramid[level];
        L.w = cvRoundF((float)w * scale);
        L.h = cvRoundF((float)h * scale);
        L.step = L.w + 2 * EDGE_THRESHOLD;
        L.buf.assign((size_t)L.step * (L.h + 2 * EDGE_THRESHOLD), 0);
        if (level != 0) {
            const Level& P = pyramid[level - 1];
            resizeLinearU8(P.roi(), P.w, P.h, P.step, L.roi(), L.w, L.h, L.step);
        } else {
            for (int y = 0; y < h; ++y) std::memcpy(L.roi() + (ptrdiff_t)y * L.step, image + (ptrdiff_t)y * stride, (size_t)w);
        }
        fillBorderReflect101(L.buf.data(), L.w, L.h, EDGE_THRESHOLD, L.step);
    }
}

namespace {
struct Node {
    std::vector<Candidate> keys;
    int ULx, ULy, URx, URy, BLx, BLy, BRx, BRy;
    std::list<Node>::iterator lit;
    bool noMore = false;
    uint64_t uid = 0;
};

// ORBextractor.cc:479-535
void divideNode(const Node& p, Node& n1, Node& n2, Node& n3, Node& n4) {
    const int halfX = (int)std::ceil(static_cast<float>(p.URx - p.ULx) / 2);
    const int halfY = (int)std::ceil(static_cast<float>(p.BRy - p.ULy) / 2);
    n1.ULx = p.ULx; n1.ULy = p.ULy;
    n1.URx = p.ULx + halfX; n1.URy = p.ULy;
    n1.BLx = p.ULx; n1.BLy = p.ULy + halfY;
    n1.BRx = p.ULx + halfX; n1.BRy = p.ULy + halfY;
    n2.ULx = n1.URx; n2.ULy = n1.URy;
    n2.URx = p.URx; n2.URy = p.URy;
    n2.BLx = n1.BRx; n2.BLy = n1.BRy;
    n2.BRx = p.URx; n2.BRy = p.ULy + halfY;
    n3.ULx = n1.BLx; n3.ULy = n1.BLy;
    n3.URx = n1.BRx; n3.URy = n1.BRy;
    n3.BLx = p.BLx; n3.BLy = p.BLy;
    n3.BRx = n1.BRx; n3.BRy = p.BLy;
    n4.ULx = n3.URx; n4.ULy = n3.URy;
    n4.URx = n2.BRx; n4.URy = n2.BRy;
    n4.BLx = n3.BRx; n4.BLy = n3.BRy;
    n4.BRx = p.BRx; n4.BRy = p.BRy;
    for (const Candidate& kp : p.keys) {
        if (kp.x < n1.URx) {
            if (kp.y < n1.BRy) n1.keys.push_back(kp);
            else n3.keys.push_back(kp);
        } else if (kp.y < n1.BRy) n2.keys.push_back(kp);
        else n4.keys.push_back(kp);
    }
    if (n1.keys.size() == 1) n1.noMore = true;
    if (n2.keys.size() == 1) n2.noMore = true;
    if (n3.keys.size() == 1) n3.noMore = true;
    if (n4.keys.size() == 1) n4.noMore = true;
}
}  // namespace

// ORBextractor.cc:537-761.  The reference sorts pair<int size, ExtractorNode*>; pointer order is a
// malloc artefact, so the oracle's rule (pointerTieRule == 0) is "later-created node compares
// greater", which is what monotonically growing heap addresses give.  pointerTieRule == 1 keeps the
// real addresses of this process's std::list nodes (used only to measure how often the rule matters).
std::vector<Candidate> OrbExtractorRef::distributeOctTree(const std::vector<Candidate>& keys, int minX, int maxX,
                                                          int minY, int maxY, int N) const {
    const int nIni = (int)std::round(static_cast<float>(maxX - minX) / (maxY - minY));
    const float hX = static_cast<float>(maxX - minX) / nIni;
    std::list<Node> lNodes;
    std::vector<Node*> ini(nIni);
    uint64_t uid = 0;
    for (int i = 0; i < nIni; i++) {
        Node ni;
        ni.ULx = (int)(hX * static_cast<float>(i)); ni.ULy = 0;
        ni.URx = (int)(hX * static_cast<float>(i + 1)); ni.URy = 0;
        ni.BLx = ni.ULx; ni.BLy = maxY - minY;
        ni.BRx = ni.URx; ni.BRy = maxY - minY;
        ni.uid = ++uid;
        lNodes.push_back(ni);
        ini[i] = &lNodes.back();
    }
    for (const Candidate& kp : keys) ini[(size_t)(kp.x / hX)]->keys.push_back(kp);
    for (auto lit = lNodes.begin(); lit != lNodes.end();) {
        if (lit->keys.size() == 1) { lit->noMore = true; ++lit; }
        else if (lit->keys.empty()) lit = lNodes.erase(lit);
        else ++lit;
    }
    typedef std::pair<int, std::pair<uint64_t, Node*>> SizeNode;   // (size, (tie key, node))
    auto tieKey = [&](Node* n) { return pointerTieRule ? (uint64_t)(uintptr_t)n : n->uid; };
    std::vector<SizeNode> vSizeAndNode;
    auto pushChildren = [&](Node* ch[4], int& nToExpand) {
        for (int c = 0; c < 4; ++c) {
            if (ch[c]->keys.size() > 0) {
                ch[c]->uid = ++uid;
                lNodes.push_front(*ch[c]);
                if (ch[c]->keys.size() > 1) {
                    nToExpand++;
                    Node* f = &lNodes.front();
                    vSizeAndNode.push_back({(int)f->keys.size(), {tieKey(f), f}});
                    f->lit = lNodes.begin();
                }
            }
        }
    };
    bool bFinish = false;
    while (!bFinish) {
        int prevSize = (int)lNodes.size();
        auto lit = lNodes.begin();
        int nToExpand = 0;
        vSizeAndNode.clear();
        while (lit != lNodes.end()) {
            if (lit->noMore) { ++lit; continue; }
            Node n1, n2, n3, n4;
            divideNode(*lit, n1, n2, n3, n4);
            Node* ch[4] = {&n1, &n2, &n3, &n4};
            pushChildren(ch, nToExpand);
            lit = lNodes.erase(lit);
        }
        if ((int)lNodes.size() >= N || (int)lNodes.size() == prevSize) bFinish = true;
        else if (((int)lNodes.size() + nToExpand * 3) > N) {
            while (!bFinish) {
                prevSize = (int)lNodes.size();
                std::vector<SizeNode> prev = vSizeAndNode;
                vSizeAndNode.clear();
                std::sort(prev.begin(), prev.end());
                for (int j = (int)prev.size() - 1; j >= 0; j--) {
                    Node n1, n2, n3, n4;
                    Node* p = prev[j].second.second;
                    divideNode(*p, n1, n2, n3, n4);
                    Node* ch[4] = {&n1, &n2, &n3, &n4};
                    int dummy = 0;
                    pushChildren(ch, dummy);
                    lNodes.erase(p->lit);
                    if ((int)lNodes.size() >= N) break;
                }
                if ((int)lNodes.size() >= N || (int)lNodes.size() == prevSize) bFinish = true;
            }
        }
    }
    std::vector<Candidate> result;
    result.reserve(lNodes.size());
    for (auto& nd : lNodes) {
        const Candidate* best = &nd.keys[0];
        float maxResponse = best->response;
        for (size_t k = 1; k < nd.keys.size(); k++)
            if (nd.keys[k].response > maxResponse) { best = &nd.keys[k]; maxResponse = nd.keys[k].response; }
        result.push_back(*best);
    }
    return result;
}

// ORBextractor.cc:75-102
float OrbExtractorRef::icAngle(const Level& L, float x, float y) const {
    int m_01 = 0, m_10 = 0;
    const uint8_t* center = L.roi() + (ptrdiff_t)cvRoundF(y) * L.step + cvRoundF(x);
    for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m_10 += u * center[u];
    const int step = (int)L.step;
    for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
        int v_sum = 0;
        const int d = umax[v];
        for (int u = -d; u <= d; ++u) {
            const int val_plus = center[u + v * step], val_minus = center[u - v * step];
            v_sum += (val_plus - val_minus);
            m_10 += u * (val_plus + val_minus);
        }
        m_01 += v * v_sum;
    }
    return fastAtan2((float)m_01, (float)m_10);
}

// ORBextractor.cc:763-878
void OrbExtractorRef::computeKeyPointsOctTree(std::vector<std::vector<KeyPoint>>& all) {
    all.assign(nlevels, {});
    candidates.assign(nlevels, {});
    const float W = 35;
    std::vector<FastKp> cell;
    for (int level = 0; level < nlevels; ++level) {
        const Level& L = pyramid[level];
        const int minBorderX = EDGE_THRESHOLD - 3, minBorderY = minBorderX;
        const int maxBorderX = L.w - EDGE_THRESHOLD + 3, maxBorderY = L.h - EDGE_THRESHOLD + 3;
        std::vector<Candidate>& vToDistribute = candidates[level];
        vToDistribute.reserve((size_t)nfeatures * 10);
        const float width = (float)(maxBorderX - minBorderX), height = (float)(maxBorderY - minBorderY);
        const int nCols = (int)(width / W), nRows = (int)(height / W);
        const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
        for (int i = 0; i < nRows; i++) {
            const float iniY = (float)(minBorderY + i * hCell);
            float maxY = iniY + hCell + 6;
            if (iniY >= maxBorderY - 3) continue;
            if (maxY > maxBorderY) maxY = (float)maxBorderY;
            for (int j = 0; j < nCols; j++) {
                const float iniX = (float)(minBorderX + j * wCell);
                float maxX = iniX + wCell + 6;
                if (iniX >= maxBorderX - 6) continue;
                if (maxX > maxBorderX) maxX = (float)maxBorderX;
                const uint8_t* roi = L.roi() + (ptrdiff_t)iniY * L.step + (int)iniX;
                const int cw = (int)maxX - (int)iniX, chh = (int)maxY - (int)iniY;
                fast9_16(roi, cw, chh, L.step, iniThFAST, cell);
                if (cell.empty()) fast9_16(roi, cw, chh, L.step, minThFAST, cell);
                for (const FastKp& k : cell)
                    vToDistribute.push_back({(float)k.x + j * wCell, (float)k.y + i * hCell, (float)k.score});
            }
        }
        std::vector<Candidate> sel =
            distributeOctTree(vToDistribute, minBorderX, maxBorderX, minBorderY, maxBorderY, mnFeaturesPerLevel[level]);
        const int scaledPatchSize = (int)(PATCH_SIZE * mvScaleFactor[level]);
        std::vector<KeyPoint>& kps = all[level];
        kps.reserve(sel.size());
        for (const Candidate& c : sel)
            kps.push_back({c.x + minBorderX, c.y + minBorderY, (float)scaledPatchSize, -1.f, c.response, level, -1});
    }
    for (int level = 0; level < nlevels; ++level)
        for (KeyPoint& kp : all[level]) kp.angle = icAngle(pyramid[level], kp.x, kp.y);
}

// ORBextractor.cc:106-145
void OrbExtractorRef::computeDescriptor(const uint8_t* img, ptrdiff_t step, const KeyPoint& kpt, uint8_t* desc) const {
    const float factorPI = (float)(3.14159265358979323846 / 180.f);
    const float angle = (float)kpt.angle * factorPI;
    const float a = (float)cosf(angle), b = (float)sinf(angle);
    const uint8_t* center = img + (ptrdiff_t)cvRoundF(kpt.y) * step + cvRoundF(kpt.x);
    const int8_t* pat = pattern;
    auto val = [&](int idx) -> int {
        const float px = (float)pat[2 * idx], py = (float)pat[2 * idx + 1];
        return center[cvRoundF(px * b + py * a) * step + cvRoundF(px * a - py * b)];
    };
    for (int i = 0; i < 32; ++i, pat += 32) {
        int v = 0;
        for (int k = 0; k < 8; ++k) {
            const int t0 = val(2 * k), t1 = val(2 * k + 1);
            v |= (t0 < t1) << k;
        }
        desc[i] = (uint8_t)v;
    }
}

// ORBextractor.cc:1068-1150
int OrbExtractorRef::extract(const uint8_t* image, int w, int h, ptrdiff_t stride, int lap0, int lap1,
                             std::vector<KeyPoint>& out, std::vector<uint8_t>& descriptors) {
    if (!image || w <= 0 || h <= 0) return -1;
    computePyramid(image, w, h, stride);
    std::vector<std::vector<KeyPoint>> all;
    computeKeyPointsOctTree(all);
    int n = 0;
    for (int level = 0; level < nlevels; ++level) n += (int)all[level].size();
    descriptors.assign((size_t)n * 32, 0);
    out.assign((size_t)n, KeyPoint());
    int monoIndex = 0, stereoIndex = n - 1;
    std::vector<uint8_t> work;
    uint8_t d[32];
    for (int level = 0; level < nlevels; ++level) {
        std::vector<KeyPoint>& kps = all[level];
        if (kps.empty()) continue;
        const Level& L = pyramid[level];
        work.resize((size_t)L.w * L.h);
        gaussianBlur7x7(L.roi(), L.w, L.h, L.step, work.data(), L.w);
        const float scale = mvScaleFactor[level];
        for (KeyPoint& kp : kps) {
            computeDescriptor(work.data(), L.w, kp, d);
            if (level != 0) { kp.x *= scale; kp.y *= scale; }
            int dst;
            if (kp.x >= lap0 && kp.x <= lap1) dst = stereoIndex--;
            else dst = monoIndex++;
            out[dst] = kp;
            std::memcpy(&descriptors[(size_t)dst * 32], d, 32);
        }
    }
    return monoIndex;
}

}  // namespace orbref
