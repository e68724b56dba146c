// match_ref.cc — see match_ref.h.  TEST INFRASTRUCTURE: oracle + timed CPU baseline only.
#include "match_ref.h"

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>

namespace orbref {

// ORBmatcher.cc:2561-2577
int descriptorDistanceSwar(const uint8_t* a, const uint8_t* b) {
    int32_t pa[8], pb[8];
    std::memcpy(pa, a, 32);
    std::memcpy(pb, b, 32);
    int dist = 0;
    for (int i = 0; i < 8; i++) {
        unsigned int v = (unsigned int)(pa[i] ^ pb[i]);
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}

// ORBmatcher.cc:2515-2556
void computeThreeMaxima(const int* counts, int L, int& ind1, int& ind2, int& ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; i++) {
        const int s = counts[i];
        if (s > max1) {
            max3 = max2; max2 = max1; max1 = s;
            ind3 = ind2; ind2 = ind1; ind1 = i;
        } else if (s > max2) {
            max3 = max2; max2 = s;
            ind3 = ind2; ind2 = i;
        } else if (s > max3) {
            max3 = s; ind3 = i;
        }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

// Exact L1 distance of two 11x11 uint8 windows (cv::norm(IL, IR, NORM_L1), Frame.cc:908).
static inline int sad11(const uint8_t* a, ptrdiff_t sa, const uint8_t* b, ptrdiff_t sb) {
    int s = 0;
    for (int r = 0; r < 11; ++r)
        for (int c = 0; c < 11; ++c) s += std::abs((int)a[r * sa + c] - (int)b[r * sb + c]);
    return s;
}

// Frame.cc:794-964
void computeStereoMatches(const OrbExtractorRef& left, const OrbExtractorRef& right, const KeyPoint* kpL,
                          const uint8_t* descL, int N, const KeyPoint* kpR, const uint8_t* descR, int Nr, float mbf,
                          float mb, float* mvuRight, float* mvDepth, int32_t* bestDistOut) {
    for (int i = 0; i < N; ++i) { mvuRight[i] = -1.0f; mvDepth[i] = -1.0f; if (bestDistOut) bestDistOut[i] = -1; }
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    const int nRows = left.pyramid[0].h;
    std::vector<std::vector<size_t>> vRowIndices(nRows);
    for (int iR = 0; iR < Nr; iR++) {
        const float kpY = kpR[iR].y;
        const float r = 2.0f * right.mvScaleFactor[kpR[iR].octave];
        const int maxr = (int)std::ceil(kpY + r);
        const int minr = (int)std::floor(kpY - r);
        // the reference indexes vRowIndices unchecked (Frame.cc:819-820); rows outside the image hold no left keypoint
        for (int yi = std::max(minr, 0); yi <= std::min(maxr, nRows - 1); yi++) vRowIndices[yi].push_back(iR);
    }
    const float minZ = mb, minD = 0, maxD = mbf / minZ;
    std::vector<std::pair<int, int>> vDistIdx;
    for (int iL = 0; iL < N; iL++) {
        const KeyPoint& kp = kpL[iL];
        const int levelL = kp.octave;
        const float vL = kp.y, uL = kp.x;
        const std::vector<size_t>& vCandidates = vRowIndices[(size_t)vL];
        if (vCandidates.empty()) continue;
        const float minU = uL - maxD, maxU = uL - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH;
        size_t bestIdxR = 0;
        for (size_t iC = 0; iC < vCandidates.size(); iC++) {
            const size_t iR = vCandidates[iC];
            const KeyPoint& kr = kpR[iR];
            if (kr.octave < levelL - 1 || kr.octave > levelL + 1) continue;
            const float uR = kr.x;
            if (uR >= minU && uR <= maxU) {
                const int dist = descriptorDistanceSwar(descL + 32 * (size_t)iL, descR + 32 * iR);
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
            }
        }
        if (bestDist < thOrbDist) {
            const float uR0 = kpR[bestIdxR].x;
            const float scaleFactor = left.mvInvScaleFactor[kp.octave];
            const float scaleduL = std::round(kp.x * scaleFactor);
            const float scaledvL = std::round(kp.y * scaleFactor);
            const float scaleduR0 = std::round(uR0 * scaleFactor);
            const int w = 5;
            const Level& PL = left.pyramid[kp.octave];
            const Level& PR = right.pyramid[kp.octave];
            const uint8_t* IL = PL.roi() + (ptrdiff_t)((int)(scaledvL - w)) * PL.step + (int)(scaleduL - w);
            int bestDistS = INT_MAX, bestincR = 0;
            const int L = 5;
            float vDists[2 * 5 + 1];
            const float iniu = scaleduR0 + L - w;
            const float endu = scaleduR0 + L + w + 1;
            if (iniu < 0 || endu >= PR.w) continue;
            for (int incR = -L; incR <= +L; incR++) {
                const uint8_t* IR = PR.roi() + (ptrdiff_t)((int)(scaledvL - w)) * PR.step + (int)(scaleduR0 + incR - w);
                const float dist = (float)sad11(IL, PL.step, IR, PR.step);
                if (dist < bestDistS) { bestDistS = (int)dist; bestincR = incR; }
                vDists[L + incR] = dist;
            }
            if (bestincR == -L || bestincR == L) continue;
            const float dist1 = vDists[L + bestincR - 1], dist2 = vDists[L + bestincR], dist3 = vDists[L + bestincR + 1];
            const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
            if (deltaR < -1 || deltaR > 1) continue;
            float bestuR = left.mvScaleFactor[kp.octave] * ((float)scaleduR0 + (float)bestincR + deltaR);
            float disparity = (uL - bestuR);
            if (disparity >= minD && disparity < maxD) {
                if (disparity <= 0) { disparity = 0.01; bestuR = uL - 0.01; }
                mvDepth[iL] = mbf / disparity;
                mvuRight[iL] = bestuR;
                vDistIdx.push_back(std::pair<int, int>(bestDistS, iL));
            }
        }
    }
    if (vDistIdx.empty()) return;   // the reference reads vDistIdx[0] of an empty vector here (Frame.cc:951)
    std::sort(vDistIdx.begin(), vDistIdx.end());
    const float median = vDistIdx[vDistIdx.size() / 2].first;
    const float thDist = 1.5f * 1.4f * median;
    if (bestDistOut) for (auto& p : vDistIdx) bestDistOut[p.second] = p.first;
    for (int i = (int)vDistIdx.size() - 1; i >= 0; i--) {
        if (vDistIdx[i].first < thDist) break;
        mvuRight[vDistIdx[i].second] = -1;
        mvDepth[vDistIdx[i].second] = -1;
    }
}

// Merge-join of two FeatureVectors (the while/lower_bound walk of ORBmatcher.cc:289-429): calls
// fn(kA, kB) for every node id present in both, ascending.
template <class Fn>
static void forSharedNodes(const FeatureVectorCsr& a, const FeatureVectorCsr& b, Fn fn) {
    int i = 0, j = 0;
    while (i < a.nnodes && j < b.nnodes) {
        if (a.nodeIds[i] == b.nodeIds[j]) { fn(i, j); ++i; ++j; }
        else if (a.nodeIds[i] < b.nodeIds[j]) ++i;   // lower_bound lands on the first id >= the other's
        else ++j;
    }
}

// ORBmatcher.cc:269-471 (KF,Frame) and :823-963 (KF,KF)
int searchByBoW(const uint8_t* descA, int nA, const FeatureVectorCsr& fvA, const uint8_t* maskA, const float* angleA,
                const uint8_t* descB, int nB, const FeatureVectorCsr& fvB, const uint8_t* maskB, const float* angleB,
                const BowParams& p, int32_t* matchAtoB, int32_t* matchBtoA) {
    for (int i = 0; i < nA; ++i) matchAtoB[i] = -1;
    for (int i = 0; i < nB; ++i) matchBtoA[i] = -1;
    int nmatches = 0;
    // rotHist[bin] holds (a, b) index pairs in push order
    std::vector<std::pair<int, int>> rotHist[HISTO_LENGTH];
    forSharedNodes(fvA, fvB, [&](int kA, int kB) {
        for (int ia = fvA.offsets[kA]; ia < fvA.offsets[kA + 1]; ++ia) {
            const int idxA = fvA.indices[ia];
            if (!maskA[idxA]) continue;
            const uint8_t* dA = descA + 32 * (size_t)idxA;
            int bestDist1 = 256, bestIdx = -1, bestDist2 = 256;
            int bestDist1R = 256, bestIdxR = -1, bestDist2R = 256;
            for (int ib = fvB.offsets[kB]; ib < fvB.offsets[kB + 1]; ++ib) {
                const int idxB = fvB.indices[ib];
                if (matchBtoA[idxB] >= 0) continue;            // vpMapPointMatches[realIdxF] / vbMatched2[idx2]
                if (maskB && !maskB[idxB]) continue;
                const int dist = descriptorDistanceSwar(dA, descB + 32 * (size_t)idxB);
                if (p.nleftB < 0 || idxB < p.nleftB) {
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx = idxB; }
                    else if (dist < bestDist2) bestDist2 = dist;
                } else {
                    if (dist < bestDist1R) { bestDist2R = bestDist1R; bestDist1R = dist; bestIdxR = idxB; }
                    else if (dist < bestDist2R) bestDist2R = dist;
                }
            }
            const bool under = p.strictTh ? bestDist1 < p.th : bestDist1 <= p.th;
            if (under) {
                if ((float)bestDist1 < p.nnRatio * (float)bestDist2) {
                    matchBtoA[bestIdx] = idxA;
                    matchAtoB[idxA] = bestIdx;
                    if (p.checkOri) rotHist[rotBin(angleA[idxA], angleB[bestIdx])].push_back({idxA, bestIdx});
                    nmatches++;
                }
                if (p.nleftB >= 0 && bestDist1R <= p.th) {     // :406-437, the ratio test there is "|| true"
                    matchBtoA[bestIdxR] = idxA;
                    if (matchAtoB[idxA] < 0) matchAtoB[idxA] = bestIdxR;
                    if (p.checkOri) rotHist[rotBin(angleA[idxA], angleB[bestIdxR])].push_back({idxA, bestIdxR});
                    nmatches++;
                }
            }
        }
    });
    if (p.checkOri) {
        int counts[HISTO_LENGTH], ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < HISTO_LENGTH; ++i) counts[i] = (int)rotHist[i].size();
        computeThreeMaxima(counts, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (auto& m : rotHist[i]) {
                matchBtoA[m.second] = -1;
                if (matchAtoB[m.first] == m.second) matchAtoB[m.first] = -1;
                nmatches--;
            }
        }
    }
    return nmatches;
}

// Pinhole::epipolarConstrain_ (Pinhole.cpp:159-181) with F12 already formed.
static inline bool epipolarOk(const float* F12, const KeyPoint& kp1, const KeyPoint& kp2, float unc) {
    const float a = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
    const float b = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
    const float c = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
    const float num = a * kp2.x + b * kp2.y + c;
    const float den = a * a + b * b;
    if (den == 0) return false;
    const float dsqr = num * num / den;
    return dsqr < 3.84 * unc;
}

// ORBmatcher.cc:1208-1432 (no second camera: mpCamera2 == nullptr, NLeft == -1)
int searchForTriangulation(const KeyPoint* kpA, const uint8_t* descA, int nA, const FeatureVectorCsr& fvA,
                           const uint8_t* hasMpA, const float* uRightA, const KeyPoint* kpB, const uint8_t* descB, int nB,
                           const FeatureVectorCsr& fvB, const uint8_t* hasMpB, const float* uRightB, const float* sigma2B,
                           const float* scaleB, const TriParams& p, int32_t* vMatches12) {
    (void)nB;
    for (int i = 0; i < nA; ++i) vMatches12[i] = -1;
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    forSharedNodes(fvA, fvB, [&](int kA, int kB) {
        for (int ia = fvA.offsets[kA]; ia < fvA.offsets[kA + 1]; ++ia) {
            const int idx1 = fvA.indices[ia];
            if (hasMpA[idx1]) continue;
            const bool bStereo1 = uRightA[idx1] >= 0;
            if (p.onlyStereo && !bStereo1) continue;
            const KeyPoint& kp1 = kpA[idx1];
            const uint8_t* d1 = descA + 32 * (size_t)idx1;
            int bestDist = TH_LOW, bestIdx2 = -1;
            for (int ib = fvB.offsets[kB]; ib < fvB.offsets[kB + 1]; ++ib) {
                const int idx2 = fvB.indices[ib];
                if (hasMpB[idx2]) continue;                     // vbMatched2 is never set in the reference
                const bool bStereo2 = uRightB[idx2] >= 0;
                if (p.onlyStereo && !bStereo2) continue;
                const int dist = descriptorDistanceSwar(d1, descB + 32 * (size_t)idx2);
                if (dist > TH_LOW || dist > bestDist) continue;
                const KeyPoint& kp2 = kpB[idx2];
                if (!bStereo1 && !bStereo2) {
                    const float distex = p.ex - kp2.x, distey = p.ey - kp2.y;
                    if (distex * distex + distey * distey < 100 * scaleB[kp2.octave]) continue;
                }
                if (epipolarOk(p.F12, kp1, kp2, sigma2B[kp2.octave]) || p.coarse) { bestIdx2 = idx2; bestDist = dist; }
            }
            if (bestIdx2 >= 0) {
                vMatches12[idx1] = bestIdx2;
                nmatches++;
                if (p.checkOri) rotHist[rotBin(kp1.angle, kpB[bestIdx2].angle)].push_back(idx1);
            }
        }
    });
    if (p.checkOri) {
        int counts[HISTO_LENGTH], ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < HISTO_LENGTH; ++i) counts[i] = (int)rotHist[i].size();
        computeThreeMaxima(counts, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx : rotHist[i]) { vMatches12[idx] = -1; nmatches--; }
        }
    }
    return nmatches;
}

// ORBmatcher.cc:1208-1432 with KannalaBrandt8::epipolarConstrain_ as the geometric test.
int searchForTriangulationKb8(const KeyPoint* kpA, const uint8_t* descA, int nA, const FeatureVectorCsr& fvA,
                              const uint8_t* hasMpA, const float* sigma2A, const KeyPoint* kpB, const uint8_t* descB, int nB,
                              const FeatureVectorCsr& fvB, const uint8_t* hasMpB, const float* sigma2B, const float* scaleB,
                              const TriKb8Params& p, int32_t* vMatches12, float* points12, float* margins) {
    (void)nB;
    for (int i = 0; i < nA; ++i) {
        vMatches12[i] = -1;
        if (margins) margins[i] = 1e30f;
        if (points12) points12[3 * i] = points12[3 * i + 1] = points12[3 * i + 2] = 0.f;
    }
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const bool dual = p.nleftA >= 0 && p.nleftB >= 0;          // pKF1->mpCamera2 && pKF2->mpCamera2 (:1332)
    forSharedNodes(fvA, fvB, [&](int kA, int kB) {
        for (int ia = fvA.offsets[kA]; ia < fvA.offsets[kA + 1]; ++ia) {
            const int idx1 = fvA.indices[ia];
            if (hasMpA[idx1]) continue;
            if (p.onlyStereo) continue;                            // bStereo1 is false for every fisheye feature
            const KeyPoint& kp1 = kpA[idx1];
            const bool bRight1 = !(p.nleftA == -1 || idx1 < p.nleftA);
            const uint8_t* d1 = descA + 32 * (size_t)idx1;
            int bestDist = TH_LOW, bestIdx2 = -1;
            float bestP[3] = {0, 0, 0};
            for (int ib = fvB.offsets[kB]; ib < fvB.offsets[kB + 1]; ++ib) {
                const int idx2 = fvB.indices[ib];
                if (hasMpB[idx2]) continue;
                const int dist = descriptorDistanceSwar(d1, descB + 32 * (size_t)idx2);
                if (dist > TH_LOW) continue;
                const KeyPoint& kp2 = kpB[idx2];
                const bool bRight2 = !(p.nleftB == -1 || idx2 < p.nleftB);
                if (p.nleftA < 0) {                                // !pKF1->mpCamera2
                    const float distex = p.ex - kp2.x, distey = p.ey - kp2.y;
                    if (distex * distex + distey * distey < 100 * scaleB[kp2.octave]) continue;
                }
                const int c = dual ? 2 * (int)bRight1 + (int)bRight2 : 0;
                const Kb8& cam1 = dual && bRight1 ? p.camA[1] : p.camA[0];
                const Kb8& cam2 = dual && bRight2 ? p.camB[1] : p.camB[0];
                float P[3] = {0, 0, 0}, mg = 1e30f;
                // evaluated for every candidate (the reference skips it when dist > bestDist; the test is pure,
                // so the outcome is the same) to report the margin of everything a parallel search evaluates
                const bool ok = kb8TriangulateMatches(cam1, cam2, kp1, kp2, p.R12[c], p.t12[c], sigma2A[kp1.octave],
                                                      sigma2B[kp2.octave], P, &mg) > 0.0001f;
                if (margins && !p.coarse && mg < margins[idx1]) margins[idx1] = mg;
                if (dist > bestDist) continue;
                if (ok || p.coarse) { bestIdx2 = idx2; bestDist = dist; bestP[0] = P[0]; bestP[1] = P[1]; bestP[2] = P[2]; }
            }
            if (bestIdx2 >= 0) {
                vMatches12[idx1] = bestIdx2;
                if (points12) { points12[3 * idx1] = bestP[0]; points12[3 * idx1 + 1] = bestP[1]; points12[3 * idx1 + 2] = bestP[2]; }
                nmatches++;
                if (p.checkOri) rotHist[rotBin(kp1.angle, kpB[bestIdx2].angle)].push_back(idx1);
            }
        }
    });
    if (p.checkOri) {
        int counts[HISTO_LENGTH], ind1 = -1, ind2 = -1, ind3 = -1;
        for (int i = 0; i < HISTO_LENGTH; ++i) counts[i] = (int)rotHist[i].size();
        computeThreeMaxima(counts, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx : rotHist[i]) { vMatches12[idx] = -1; nmatches--; }
        }
    }
    return nmatches;
}

// Frame.cc:730-763 + OpenCV's cvUndistortPointsInternal (imgproc/src/undistort.dispatch.cpp), iters = 5
void undistortKeyPoints(const KeyPoint* kps, int n, const float K4[4], const float* dist, int ndist, KeyPoint* out) {
    for (int i = 0; i < n; ++i) out[i] = kps[i];
    if (ndist < 1 || dist[0] == 0.0f) return;
    double k[5] = {0, 0, 0, 0, 0};                       // k1 k2 p1 p2 k3
    for (int j = 0; j < ndist && j < 5; ++j) k[j] = (double)dist[j];
    const double fx = K4[0], fy = K4[1], cx = K4[2], cy = K4[3], ifx = 1.0 / fx, ify = 1.0 / fy;
    for (int i = 0; i < n; ++i) {
        const double u = kps[i].x, v = kps[i].y;
        double x = (u - cx) * ifx, y = (v - cy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; ++j) {
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((0 * r2 + 0) * r2 + 0) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
            if (icdist < 0) { x = (u - cx) * ifx; y = (v - cy) * ify; break; }
            const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + 0 * r2 + 0 * r2 * r2;
            const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + 0 * r2 + 0 * r2 * r2;
            x = (x0 - deltaX) * icdist;
            y = (y0 - deltaY) * icdist;
        }
        out[i].x = (float)(fx * x + cx);
        out[i].y = (float)(fy * y + cy);
    }
}

// MapPoint.cc:388-417
int distinctiveDescriptor(const uint8_t* desc, int n) {
    if (n <= 0) return -1;
    std::vector<float> D((size_t)n * n);
    for (int i = 0; i < n; i++) {
        D[(size_t)i * n + i] = 0;
        for (int j = i + 1; j < n; j++) {
            const int distij = descriptorDistanceSwar(desc + 32 * (size_t)i, desc + 32 * (size_t)j);
            D[(size_t)i * n + j] = (float)distij;
            D[(size_t)j * n + i] = (float)distij;
        }
    }
    int BestMedian = 0x7fffffff, BestIdx = 0;
    for (int i = 0; i < n; i++) {
        std::vector<int> vDists(D.begin() + (size_t)i * n, D.begin() + (size_t)(i + 1) * n);
        std::sort(vDists.begin(), vDists.end());
        const int median = vDists[(size_t)(0.5 * (n - 1))];
        if (median < BestMedian) { BestMedian = median; BestIdx = i; }
    }
    return BestIdx;
}

// Frame.cc:1134-1141
void knn2RatioMatch(const uint8_t* q, int nq, const uint8_t* t, int nt, double ratio, int32_t* match) {
    std::vector<int32_t> idx(2 * (size_t)nq), dist(2 * (size_t)nq);
    knn2Hamming(q, nq, t, nt, idx.data(), dist.data());
    for (int i = 0; i < nq; ++i) {
        const bool two = idx[2 * i + 1] >= 0;
        // DMatch::distance is float; 0.7 is a double literal: (double)d0 < (double)d1 * 0.7
        match[i] = (two && (double)(float)dist[2 * i] < (double)(float)dist[2 * i + 1] * ratio) ? idx[2 * i] : -1;
    }
}

}  // namespace orbref
