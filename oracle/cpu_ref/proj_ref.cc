// proj_ref.cc — see proj_ref.h (TEST INFRASTRUCTURE).
#include "proj_ref.h"

#include <algorithm>
#include <climits>
#include <cmath>

#include "match_ref.h"

namespace orbref {

FrameGrid::FrameGrid(const GridFrame& f) : F(f) {
    for (int i = 0; i < F.n; ++i) {                                              // Frame.cc:397-410
        const KeyPoint& kp = F.kps[i];
        const int posX = (int)std::round((kp.x - F.minX) * F.gridWInv);          // PosInGrid :710-711
        const int posY = (int)std::round((kp.y - F.minY) * F.gridHInv);
        if (posX < 0 || posX >= FRAME_GRID_COLS || posY < 0 || posY >= FRAME_GRID_ROWS) continue;
        if (F.nleft == -1 || i < F.nleft) cell[0][posX][posY].push_back(i);
        else cell[1][posX][posY].push_back(i);
    }
}

void FrameGrid::featuresInArea(float x, float y, float r, int minLevel, int maxLevel, bool bRight, std::vector<int>& out) const {
    out.clear();
    const float factorX = r, factorY = r;
    const int nMinCellX = std::max(0, (int)std::floor((x - F.minX - factorX) * F.gridWInv));
    if (nMinCellX >= FRAME_GRID_COLS) return;
    const int nMaxCellX = std::min(FRAME_GRID_COLS - 1, (int)std::ceil((x - F.minX + factorX) * F.gridWInv));
    if (nMaxCellX < 0) return;
    const int nMinCellY = std::max(0, (int)std::floor((y - F.minY - factorY) * F.gridHInv));
    if (nMinCellY >= FRAME_GRID_ROWS) return;
    const int nMaxCellY = std::min(FRAME_GRID_ROWS - 1, (int)std::ceil((y - F.minY + factorY) * F.gridHInv));
    if (nMaxCellY < 0) return;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
        for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
            for (int idx : cell[bRight ? 1 : 0][ix][iy]) {
                const KeyPoint& kpUn = F.kps[idx];
                if (bCheckLevels) {
                    if (kpUn.octave < minLevel) continue;
                    if (maxLevel >= 0 && kpUn.octave > maxLevel) continue;
                }
                const float distx = kpUn.x - x, disty = kpUn.y - y;
                if (std::fabs(distx) < factorX && std::fabs(disty) < factorY) out.push_back(idx);
            }
}

int searchByProjection(const GridFrame& F, const ProjQuery* q, int nq, const ProjParams& p, int32_t* assign, int32_t* choice) {
    FrameGrid grid(F);
    int nmatches = 0;
    // mvpMapPoints as "who sits there": -1 = the caller's original entry, else a query index
    std::vector<int32_t> holder(F.n, -1);
    auto isBlocked = [&](int idx) {               // mvpMapPoints[idx] && mvpMapPoints[idx]->Observations() > 0
        return holder[idx] >= 0 ? q[holder[idx]].obs != 0 : (F.blocked && F.blocked[idx]);
    };
    std::vector<int> rotHist[HISTO_LENGTH];
    std::vector<int> cand;
    bool prevRatioRejected = false, prevEmpty = true;
    for (int i = 0; i < nq; ++i) {
        choice[i] = -1;
        const ProjQuery& Q = q[i];
        const bool skip = (Q.skipIfPrevRatioRejected && prevRatioRejected) || (Q.skipIfPrevEmpty && prevEmpty);
        prevRatioRejected = false;
        prevEmpty = true;                 // of an invalid window: the reference `continue`d before searching
        if (!Q.valid) continue;
        grid.featuresInArea(Q.u, Q.v, Q.r, Q.minLevel, Q.maxLevel, Q.right != 0, cand);
        prevEmpty = cand.empty();         // SearchByProjection(Current, Last) :2236-2237 leaves the MapPoint, right camera included
        if (skip || cand.empty()) continue;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : cand) {
            if (isBlocked(idx)) continue;
            if (!Q.right && Q.checkUr && F.nleft == -1 && F.uRight && F.uRight[idx] > 0) {
                const float er = std::fabs(Q.ur - F.uRight[idx]);
                if (er > Q.r) continue;
            }
            if (Q.chi2 && F.invLevelSigma2) {                                    // Fuse :1744-1770
                const KeyPoint& kp = F.kps[idx];
                const float kpr = F.uRight ? F.uRight[Q.right ? idx - F.nleft : idx] : -1.f;     // mvuRight[idx], idx camera-local
                const float ex = Q.u - kp.x, ey = Q.v - kp.y;
                if (kpr >= 0) {
                    const float er = Q.ur - kpr;
                    const float e2 = ex * ex + ey * ey + er * er;
                    if (e2 * F.invLevelSigma2[kp.octave] > 7.8) continue;
                } else {
                    const float e2 = ex * ex + ey * ey;
                    if (e2 * F.invLevelSigma2[kp.octave] > 5.99) continue;
                }
            }
            const int dist = descriptorDistanceSwar(Q.desc, F.desc + 32 * (size_t)idx);
            if (dist < bestDist) {
                bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = F.kps[idx].octave; bestIdx = idx;
            } else if (dist < bestDist2) {
                bestLevel2 = F.kps[idx].octave; bestDist2 = dist;
            }
        }
        if (bestDist > p.th) continue;
        if (p.mode == 1 && bestLevel == bestLevel2 && (float)bestDist > p.nnRatio * (float)bestDist2) { prevRatioRejected = true; continue; }
        holder[bestIdx] = i;
        choice[i] = bestIdx;
        nmatches++;
        if (p.linkStereo && F.nleft != -1) {
            if (!Q.right) {
                if (F.leftToRight && F.leftToRight[bestIdx] != -1) { holder[F.leftToRight[bestIdx] + F.nleft] = i; nmatches++; }
            } else {
                const int r = bestIdx - F.nleft;
                if (F.rightToLeft && F.rightToLeft[r] != -1) { holder[F.rightToLeft[r]] = i; nmatches++; }
            }
        }
        if (p.rotHist) rotHist[rotBin(Q.angle, F.kps[bestIdx].angle)].push_back(bestIdx);
    }
    for (int i = 0; i < F.n; ++i) assign[i] = holder[i];
    if (p.rotHist) {
        int counts[HISTO_LENGTH], ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < HISTO_LENGTH; ++i) counts[i] = (int)rotHist[i].size();
        computeThreeMaxima(counts, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx : rotHist[i]) { assign[idx] = -2; nmatches--; }
        }
    }
    return nmatches;
}

int searchForInitialization(const KeyPoint* kps1, const uint8_t* desc1, int n1, const GridFrame& F2, float* prevMatched,
                            int windowSize, float nnRatio, bool checkOri, int32_t* matches12) {
    FrameGrid grid(F2);
    int nmatches = 0;
    for (int i = 0; i < n1; ++i) matches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    std::vector<int> vMatchedDistance(F2.n, INT_MAX), vnMatches21(F2.n, -1), cand;
    for (int i1 = 0; i1 < n1; ++i1) {
        const int level1 = kps1[i1].octave;
        if (level1 > 0) continue;                                                              // :724-725
        grid.featuresInArea(prevMatched[2 * i1], prevMatched[2 * i1 + 1], (float)windowSize, level1, level1, false, cand);
        if (cand.empty()) continue;
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (int i2 : cand) {
            const int dist = descriptorDistanceSwar(desc1 + 32 * (size_t)i1, F2.desc + 32 * (size_t)i2);
            if (vMatchedDistance[i2] <= dist) continue;                                        // :744
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist > TH_LOW) continue;
        if (!((float)bestDist < (float)bestDist2 * nnRatio)) continue;
        if (vnMatches21[bestIdx2] >= 0) { matches12[vnMatches21[bestIdx2]] = -1; nmatches--; }
        matches12[i1] = bestIdx2;
        vnMatches21[bestIdx2] = i1;
        vMatchedDistance[bestIdx2] = bestDist;
        nmatches++;
        if (checkOri) rotHist[rotBin(kps1[i1].angle, F2.kps[bestIdx2].angle)].push_back(i1);
    }
    if (checkOri) {
        int counts[HISTO_LENGTH], ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < HISTO_LENGTH; ++i) counts[i] = (int)rotHist[i].size();
        computeThreeMaxima(counts, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i])
                if (matches12[idx1] >= 0) { matches12[idx1] = -1; nmatches--; }
        }
    }
    for (int i1 = 0; i1 < n1; ++i1)
        if (matches12[i1] >= 0) { prevMatched[2 * i1] = F2.kps[matches12[i1]].x; prevMatched[2 * i1 + 1] = F2.kps[matches12[i1]].y; }
    return nmatches;
}

}  // namespace orbref
