// proj_ref.h — CPU restatement of the projection-guided window searches (TEST INFRASTRUCTURE).
//
// Follows /root/reference/src/Frame.cc (AssignFeaturesToGrid :380-411, GetFeaturesInArea :640-706, PosInGrid
// :708-719) and the search loops of /root/reference/src/ORBmatcher.cc that are built on them:
// SearchByProjection(Frame&, vector<MapPoint*>&, th, ...) :44-236 (best + second best with the scale-level
// rule, stereo links) and SearchByProjection(Frame& Current, const Frame& Last, th, bMono) :2173-2389 /
// SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) :2391-2510 (best only, rotation histogram).
// The KeyFrame-side searches reuse the loop (KeyFrame::GetFeaturesInArea, src/KeyFrame.cc:737-781, is the same cell
// walk without the level filter, which those callers apply themselves): SearchByProjection(KeyFrame*, Scw, ...)
// :473-585 / :588-700, Fuse :1629-1821 (with its chi-square gate) / :1823-1945, SearchBySim3 :1947-2171.
// The part of those functions that differs per overload — projecting a MapPoint with a pose and a camera,
// predicting its scale level — produces, per MapPoint, a window (u, v, r, level range); everything after that
// is the same loop, restated here over flattened arrays exactly as at the C ABI (include/orbfe.h).
#pragma once
#include <cstdint>
#include <vector>

#include "orb_ref.h"

namespace orbref {

static const int FRAME_GRID_ROWS = 48, FRAME_GRID_COLS = 64;       // Frame.h:45-46

struct GridFrame {                // the Frame being searched
    const KeyPoint* kps;          // mvKeysUn (nleft == -1) or mvKeys followed by mvKeysRight
    const uint8_t* desc;
    int n, nleft;                 // Frame::N, Frame::Nleft
    const uint8_t* blocked;       // mvpMapPoints[i] && mvpMapPoints[i]->Observations() > 0 (nullable)
    const float* uRight;          // mvuRight (nullable = all -1)
    const int32_t* leftToRight;   // mvLeftToRightMatch [nleft] (nullable)
    const int32_t* rightToLeft;   // mvRightToLeftMatch [n - nleft] (nullable)
    float minX, minY, gridWInv, gridHInv;     // mnMinX, mnMinY, mfGridElementWidthInv, mfGridElementHeightInv
    const float* invLevelSigma2;  // mvInvLevelSigma2 (windows with chi2; nullable)
};

#pragma pack(push, 1)
struct ProjQuery {                // one GetFeaturesInArea window + what is matched inside it; 64 bytes
    float u, v, r;                // GetFeaturesInArea(x, y, r, minLevel, maxLevel, bRight)
    float ur;                     // predicted right coordinate (stereo check), used iff checkUr
    float angle;                  // keypoint angle of the source observation (rotation histogram)
    int16_t minLevel, maxLevel;
    uint8_t right;                // search the right-camera grid
    uint8_t obs;                  // the MapPoint has Observations() > 0: once assigned it blocks the feature
    uint8_t checkUr;              // apply |ur - mvuRight[idx]| > r -> skip (only when nleft == -1 && mvuRight > 0)
    uint8_t valid;                // 0: skipped before the window search (not in view, bad, behind the camera ...)
    uint8_t skipIfPrevRatioRejected;   // :135: the `continue` after a failed ratio test also skips the same MapPoint's right-camera search
    uint8_t chi2;                 // Fuse's per-candidate reprojection gate (ORBmatcher.cc:1744-1770)
    uint8_t skipIfPrevEmpty;      // :2236-2237: an empty left window skips the same MapPoint's right-camera window too
    uint8_t pad;
    uint8_t desc[32];             // pMP->GetDescriptor()
};
#pragma pack(pop)
static_assert(sizeof(ProjQuery) == 64, "ProjQuery layout");

struct ProjParams {
    int mode;          // 0: best only, accept bestDist <= th (:2270); 1: best + second, level rule (:133-139)
    int th;            // TH_HIGH / ORBdist
    float nnRatio;     // mfNNratio (mode 1)
    int linkStereo;    // also assign the linked stereo observation (:141-147, :223-227)
    int rotHist;       // mbCheckOrientation of the best-only overloads (:2274-2292, :2368-2387)
};

// Frame::GetFeaturesInArea on a grid built by AssignFeaturesToGrid; indices are global (right features: + nleft).
class FrameGrid {
public:
    explicit FrameGrid(const GridFrame& f);
    void featuresInArea(float x, float y, float r, int minLevel, int maxLevel, bool bRight, std::vector<int>& out) const;
private:
    const GridFrame& F;
    std::vector<int> cell[2][FRAME_GRID_COLS][FRAME_GRID_ROWS];
};

// Processes the queries in order.  assign[n]: index of the query whose MapPoint ends up in mvpMapPoints[i],
// -1 = entry untouched, -2 = set to NULL by the rotation check.  choice[nq]: feature chosen by each query
// (global index) or -1.  Returns nmatches as the reference counts it.
int searchByProjection(const GridFrame& F, const ProjQuery* q, int nq, const ProjParams& p, int32_t* assign, int32_t* choice);

// ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:706-821): F1 = (kps1 = mvKeysUn, desc1), F2 = the Frame whose
// grid is searched; prevMatched [n1][2] = vbPrevMatched, updated in place (:815-817); matches12 [n1] = vnMatches12.
// Only level-0 features of both frames take part (:721-725).  A later F1 feature with a strictly smaller distance
// steals an F2 feature (:757-768); candidates whose recorded distance is <= the new one are invisible (:744).
int searchForInitialization(const KeyPoint* kps1, const uint8_t* desc1, int n1, const GridFrame& F2, float* prevMatched,
                            int windowSize, float nnRatio, bool checkOri, int32_t* matches12);

}  // namespace orbref
