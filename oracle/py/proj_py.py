"""Python restatement of ORB-SLAM3's projection-guided searches at OBJECT level (TEST INFRASTRUCTURE).

Follows src/Frame.cc (AssignFeaturesToGrid :380-411, GetFeaturesInArea :640-706, PosInGrid :708-719) and
src/ORBmatcher.cc: SearchByProjection(Frame&, const vector<MapPoint*>&, th, bFarPoints, thFarPoints) :44-236,
RadiusByViewingCos :238-244, SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, th, bMono)
:2173-2389, SearchForInitialization :706-821, and the KeyFrame-side searches SearchByProjection(KeyFrame*, Scw, ...) :473-585, Fuse :1629-1821 /
:1823-1945, SearchBySim3 :1947-2171.  Frames and MapPoints are small Python objects with the reference's member names; float32
arithmetic is spelled out with numpy scalars.  Written independently of oracle/cpu_ref/proj_ref.cc (which
works on the flattened windows); the golden fixtures come from this file.
"""
import math

import numpy as np

F32 = np.float32
FRAME_GRID_ROWS, FRAME_GRID_COLS = 48, 64
TH_HIGH, TH_LOW, HISTO_LENGTH = 100, 50, 30
_POP = np.array([bin(i).count("1") for i in range(256)], np.int32)


def descriptor_distance(a, b):
    return int(_POP[np.bitwise_xor(a, b)].sum())


def c_round(x):
    """C round(): half away from zero."""
    return int(math.floor(x + 0.5)) if x >= 0 else -int(math.floor(-x + 0.5))


class MapPoint:
    def __init__(self, descriptor, n_obs=1, bad=False, world_pos=None):
        self.descriptor, self.nObs, self.bad, self.mWorldPos = descriptor, n_obs, bad, world_pos
        self.mbTrackInView = self.mbTrackInViewR = False
        self.mTrackProjX = self.mTrackProjY = self.mTrackProjXR = self.mTrackProjYR = F32(0)
        self.mTrackDepth = F32(0)
        self.mnTrackScaleLevel = self.mnTrackScaleLevelR = 0
        self.mTrackViewCos = self.mTrackViewCosR = F32(0)

    def Observations(self):
        return self.nObs

    def isBad(self):
        return self.bad

    def GetDescriptor(self):
        return self.descriptor


class Frame:
    """keys: structured array (x, y, octave, angle ...) = mvKeysUn, or mvKeys ++ mvKeysRight when Nleft != -1."""

    def __init__(self, keys, descriptors, scale_factors, bounds, Nleft=-1, mvuRight=None, l2r=None, r2l=None, mbf=0.0, mb=0.0):
        self.keys, self.mDescriptors, self.N, self.Nleft = keys, descriptors, len(keys), Nleft
        self.mvScaleFactors = np.asarray(scale_factors, F32)
        self.mnMinX, self.mnMinY, self.mnMaxX, self.mnMaxY = (F32(b) for b in bounds)
        self.mfGridElementWidthInv = F32(F32(FRAME_GRID_COLS) / F32(self.mnMaxX - self.mnMinX))
        self.mfGridElementHeightInv = F32(F32(FRAME_GRID_ROWS) / F32(self.mnMaxY - self.mnMinY))
        self.mvuRight = np.full(self.N, -1, F32) if mvuRight is None else np.asarray(mvuRight, F32)
        self.mvLeftToRightMatch, self.mvRightToLeftMatch = l2r, r2l
        self.mbf, self.mb = F32(mbf), F32(mb)
        self.mvpMapPoints = [None] * self.N
        self.mvbOutlier = [False] * self.N
        self.mGrid = [[[] for _ in range(FRAME_GRID_ROWS)] for _ in range(FRAME_GRID_COLS)]
        self.mGridRight = [[[] for _ in range(FRAME_GRID_ROWS)] for _ in range(FRAME_GRID_COLS)]
        for i in range(self.N):                                            # AssignFeaturesToGrid
            ok, px, py = self.PosInGrid(self.keys[i])
            if ok:
                if Nleft == -1 or i < Nleft:
                    self.mGrid[px][py].append(i)
                else:
                    self.mGridRight[px][py].append(i - Nleft)

    def PosInGrid(self, kp):
        px = c_round(float(F32(F32(kp["x"] - self.mnMinX) * self.mfGridElementWidthInv)))
        py = c_round(float(F32(F32(kp["y"] - self.mnMinY) * self.mfGridElementHeightInv)))
        if px < 0 or px >= FRAME_GRID_COLS or py < 0 or py >= FRAME_GRID_ROWS:
            return False, px, py
        return True, px, py

    def kp_of(self, idx, bRight=False):
        return self.keys[idx + self.Nleft] if bRight else self.keys[idx]

    def GetFeaturesInArea(self, x, y, r, minLevel=-1, maxLevel=-1, bRight=False):
        x, y, r = F32(x), F32(y), F32(r)
        out = []
        nMinCellX = max(0, int(math.floor(float(F32(F32(F32(x - self.mnMinX) - r) * self.mfGridElementWidthInv)))))
        if nMinCellX >= FRAME_GRID_COLS:
            return out
        nMaxCellX = min(FRAME_GRID_COLS - 1, int(math.ceil(float(F32(F32(F32(x - self.mnMinX) + r) * self.mfGridElementWidthInv)))))
        if nMaxCellX < 0:
            return out
        nMinCellY = max(0, int(math.floor(float(F32(F32(F32(y - self.mnMinY) - r) * self.mfGridElementHeightInv)))))
        if nMinCellY >= FRAME_GRID_ROWS:
            return out
        nMaxCellY = min(FRAME_GRID_ROWS - 1, int(math.ceil(float(F32(F32(F32(y - self.mnMinY) + r) * self.mfGridElementHeightInv)))))
        if nMaxCellY < 0:
            return out
        bCheckLevels = (minLevel > 0) or (maxLevel >= 0)
        for ix in range(nMinCellX, nMaxCellX + 1):
            for iy in range(nMinCellY, nMaxCellY + 1):
                vCell = self.mGridRight[ix][iy] if bRight else self.mGrid[ix][iy]
                for j in vCell:
                    kpUn = self.kp_of(j, bRight and self.Nleft != -1)
                    if bCheckLevels:
                        if kpUn["octave"] < minLevel:
                            continue
                        if maxLevel >= 0 and kpUn["octave"] > maxLevel:
                            continue
                    distx, disty = F32(kpUn["x"] - x), F32(kpUn["y"] - y)
                    if abs(distx) < r and abs(disty) < r:
                        out.append(j)
        return out


def radius_by_viewing_cos(viewCos):
    return F32(2.5) if viewCos > F32(0.998) else F32(4.0)


def search_by_projection_mappoints(F, vpMapPoints, th, nnratio, bFarPoints=False, thFarPoints=50.0):
    """ORBmatcher.cc:44-236.  Mutates F.mvpMapPoints; returns nmatches."""
    nmatches = 0
    th, nnratio = F32(th), F32(nnratio)
    bFactor = th != F32(1.0)
    for pMP in vpMapPoints:
        if not pMP.mbTrackInView and not pMP.mbTrackInViewR:
            continue
        if bFarPoints and pMP.mTrackDepth > F32(thFarPoints):
            continue
        if pMP.isBad():
            continue
        if pMP.mbTrackInView:
            lev = pMP.mnTrackScaleLevel
            r = radius_by_viewing_cos(pMP.mTrackViewCos)
            if bFactor:
                r = F32(r * th)
            win = F32(r * F.mvScaleFactors[lev])
            vIndices = F.GetFeaturesInArea(pMP.mTrackProjX, pMP.mTrackProjY, win, lev - 1, lev)
            if vIndices:
                d_mp = pMP.GetDescriptor()
                bestDist, bestLevel, bestDist2, bestLevel2, bestIdx = 256, -1, 256, -1, -1
                for idx in vIndices:
                    if F.mvpMapPoints[idx] is not None and F.mvpMapPoints[idx].Observations() > 0:
                        continue
                    if F.Nleft == -1 and F.mvuRight[idx] > 0:
                        er = abs(F32(pMP.mTrackProjXR - F.mvuRight[idx]))
                        if er > win:
                            continue
                    dist = descriptor_distance(d_mp, F.mDescriptors[idx])
                    if dist < bestDist:
                        bestDist2, bestDist, bestLevel2 = bestDist, dist, bestLevel
                        bestLevel, bestIdx = int(F.keys[idx]["octave"]), idx
                    elif dist < bestDist2:
                        bestLevel2, bestDist2 = int(F.keys[idx]["octave"]), dist
                if bestDist <= TH_HIGH:
                    if bestLevel == bestLevel2 and F32(bestDist) > F32(nnratio * F32(bestDist2)):
                        pass                                   # `continue` of :135: skips the right-camera part too
                    else:
                        F.mvpMapPoints[bestIdx] = pMP
                        if F.Nleft != -1 and F.mvLeftToRightMatch[bestIdx] != -1:
                            F.mvpMapPoints[F.mvLeftToRightMatch[bestIdx] + F.Nleft] = pMP
                            nmatches += 1
                        nmatches += 1
                    if bestLevel == bestLevel2 and F32(bestDist) > F32(nnratio * F32(bestDist2)):
                        continue
        if F.Nleft != -1 and pMP.mbTrackInViewR:
            lev = pMP.mnTrackScaleLevelR
            if lev != -1:
                r = radius_by_viewing_cos(pMP.mTrackViewCosR)
                win = F32(r * F.mvScaleFactors[lev])
                vIndices = F.GetFeaturesInArea(pMP.mTrackProjXR, pMP.mTrackProjYR, win, lev - 1, lev, True)
                if not vIndices:
                    continue
                d_mp = pMP.GetDescriptor()
                bestDist, bestLevel, bestDist2, bestLevel2, bestIdx = 256, -1, 256, -1, -1
                for idx in vIndices:
                    mp = F.mvpMapPoints[idx + F.Nleft]
                    if mp is not None and mp.Observations() > 0:
                        continue
                    dist = descriptor_distance(d_mp, F.mDescriptors[idx + F.Nleft])
                    if dist < bestDist:
                        bestDist2, bestDist, bestLevel2 = bestDist, dist, bestLevel
                        bestLevel, bestIdx = int(F.keys[idx + F.Nleft]["octave"]), idx
                    elif dist < bestDist2:
                        bestLevel2, bestDist2 = int(F.keys[idx + F.Nleft]["octave"]), dist
                if bestDist <= TH_HIGH:
                    if bestLevel == bestLevel2 and F32(bestDist) > F32(nnratio * F32(bestDist2)):
                        continue
                    if F.Nleft != -1 and F.mvRightToLeftMatch[bestIdx] != -1:
                        F.mvpMapPoints[F.mvRightToLeftMatch[bestIdx]] = pMP
                        nmatches += 1
                    F.mvpMapPoints[bestIdx + F.Nleft] = pMP
                    nmatches += 1
    return nmatches


def compute_three_maxima(counts):
    max1 = max2 = max3 = 0
    ind1 = ind2 = ind3 = -1
    for i, s in enumerate(counts):
        if s > max1:
            max3, max2, max1, ind3, ind2, ind1 = max2, max1, s, ind2, ind1, i
        elif s > max2:
            max3, max2, ind3, ind2 = max2, s, ind2, i
        elif s > max3:
            max3, ind3 = s, i
    if max2 < F32(0.1) * F32(max1):
        ind2 = ind3 = -1
    elif max3 < F32(0.1) * F32(max1):
        ind3 = -1
    return ind1, ind2, ind3


def rot_bin(a1, a2):
    rot = F32(F32(a1) - F32(a2))
    if rot < 0.0:
        rot = F32(rot + F32(360.0))
    b = c_round(float(F32(rot * F32(F32(1.0) / F32(HISTO_LENGTH)))))
    return 0 if b == HISTO_LENGTH else b


def search_by_projection_lastframe(Cur, Last, th, bMono, check_ori, project, x3Dc_of, forward_backward, project_right=None):
    """ORBmatcher.cc:2173-2389.  `x3Dc_of(pMP)` = Rcw * GetWorldPos() + tcw (float32 triple), `project` = the
    camera model, `forward_backward` = (bForward, bBackward) (:2193-2194).  Mutates Cur.mvpMapPoints."""
    nmatches = 0
    rotHist = [[] for _ in range(HISTO_LENGTH)]
    th = F32(th)
    bForward, bBackward = forward_backward
    for i in range(Last.N):
        pMP = Last.mvpMapPoints[i]
        if pMP is None or Last.mvbOutlier[i]:
            continue
        xc, yc, zc = x3Dc_of(pMP)
        invzc = F32(1.0 / float(zc))                     # 1.0 / float -> double, stored in a float
        if invzc < 0:
            continue
        u, v = project((xc, yc, zc))
        if u < Cur.mnMinX or u > Cur.mnMaxX or v < Cur.mnMinY or v > Cur.mnMaxY:
            continue
        nLastOctave = int(Last.keys[i]["octave"])
        radius = F32(th * Cur.mvScaleFactors[nLastOctave])
        if bForward:
            vIndices2 = Cur.GetFeaturesInArea(u, v, radius, nLastOctave)
        elif bBackward:
            vIndices2 = Cur.GetFeaturesInArea(u, v, radius, 0, nLastOctave)
        else:
            vIndices2 = Cur.GetFeaturesInArea(u, v, radius, nLastOctave - 1, nLastOctave + 1)
        if not vIndices2:
            continue
        dMP = pMP.GetDescriptor()
        bestDist, bestIdx2 = 256, -1
        for i2 in vIndices2:
            mp = Cur.mvpMapPoints[i2]
            if mp is not None and mp.Observations() > 0:
                continue
            if Cur.Nleft == -1 and Cur.mvuRight[i2] > 0:
                ur = F32(u - F32(Cur.mbf * invzc))
                er = abs(F32(ur - Cur.mvuRight[i2]))
                if er > radius:
                    continue
            dist = descriptor_distance(dMP, Cur.mDescriptors[i2])
            if dist < bestDist:
                bestDist, bestIdx2 = dist, i2
        if bestDist <= TH_HIGH:
            Cur.mvpMapPoints[bestIdx2] = pMP
            nmatches += 1
            if check_ori:
                rotHist[rot_bin(Last.keys[i]["angle"], Cur.keys[bestIdx2]["angle"])].append(bestIdx2)
        if Cur.Nleft != -1:
            u, v = project_right((xc, yc, zc))
            radius = F32(th * Cur.mvScaleFactors[nLastOctave])
            if bForward:
                vIndices2 = Cur.GetFeaturesInArea(u, v, radius, nLastOctave, -1, True)
            elif bBackward:
                vIndices2 = Cur.GetFeaturesInArea(u, v, radius, 0, nLastOctave, True)
            else:
                vIndices2 = Cur.GetFeaturesInArea(u, v, radius, nLastOctave - 1, nLastOctave + 1, True)
            bestDist, bestIdx2 = 256, -1
            for i2 in vIndices2:
                mp = Cur.mvpMapPoints[i2 + Cur.Nleft]
                if mp is not None and mp.Observations() > 0:
                    continue
                dist = descriptor_distance(dMP, Cur.mDescriptors[i2 + Cur.Nleft])
                if dist < bestDist:
                    bestDist, bestIdx2 = dist, i2
            if bestDist <= TH_HIGH:
                Cur.mvpMapPoints[bestIdx2 + Cur.Nleft] = pMP
                nmatches += 1
                if check_ori:
                    rotHist[rot_bin(Last.keys[i]["angle"], Cur.keys[bestIdx2 + Cur.Nleft]["angle"])].append(bestIdx2 + Cur.Nleft)
    if check_ori:
        keep = compute_three_maxima([len(h) for h in rotHist])
        for i in range(HISTO_LENGTH):
            if i not in keep:
                for idx in rotHist[i]:
                    Cur.mvpMapPoints[idx] = None
                    nmatches -= 1
    return nmatches


def search_for_initialization(F1, F2, vbPrevMatched, windowSize, nnratio, check_ori):
    """ORBmatcher::SearchForInitialization, ORBmatcher.cc:706-821.  F1 / F2: Frames whose `keys` are mvKeysUn.
    vbPrevMatched: list of [x, y] float32 pairs, updated in place (:815-817).  Returns (nmatches, vnMatches12)."""
    INT_MAX = 2147483647
    nmatches = 0
    vnMatches12 = [-1] * F1.N
    rotHist = [[] for _ in range(HISTO_LENGTH)]
    vMatchedDistance = [INT_MAX] * F2.N
    vnMatches21 = [-1] * F2.N
    for i1 in range(F1.N):
        kp1 = F1.keys[i1]
        level1 = int(kp1["octave"])
        if level1 > 0:
            continue
        vIndices2 = F2.GetFeaturesInArea(vbPrevMatched[i1][0], vbPrevMatched[i1][1], windowSize, level1, level1)
        if not vIndices2:
            continue
        d1 = F1.mDescriptors[i1]
        bestDist, bestDist2, bestIdx2 = INT_MAX, INT_MAX, -1
        for i2 in vIndices2:
            dist = descriptor_distance(d1, F2.mDescriptors[i2])
            if vMatchedDistance[i2] <= dist:
                continue
            if dist < bestDist:
                bestDist2, bestDist, bestIdx2 = bestDist, dist, i2
            elif dist < bestDist2:
                bestDist2 = dist
        if bestDist <= TH_LOW:
            if F32(bestDist) < F32(F32(bestDist2) * F32(nnratio)):
                if vnMatches21[bestIdx2] >= 0:
                    vnMatches12[vnMatches21[bestIdx2]] = -1
                    nmatches -= 1
                vnMatches12[i1] = bestIdx2
                vnMatches21[bestIdx2] = i1
                vMatchedDistance[bestIdx2] = bestDist
                nmatches += 1
                if check_ori:
                    rotHist[rot_bin(F1.keys[i1]["angle"], F2.keys[bestIdx2]["angle"])].append(i1)
    if check_ori:
        keep = compute_three_maxima([len(h) for h in rotHist])
        for i in range(HISTO_LENGTH):
            if i in keep:
                continue
            for idx1 in rotHist[i]:
                if vnMatches12[idx1] >= 0:
                    vnMatches12[idx1] = -1
                    nmatches -= 1
    for i1 in range(F1.N):
        if vnMatches12[i1] >= 0:
            vbPrevMatched[i1] = [F2.keys[vnMatches12[i1]]["x"], F2.keys[vnMatches12[i1]]["y"]]
    return nmatches, vnMatches12


# ---- KeyFrame-side searches (loop closing, merging, local mapping) -------------------------------------------------
class KeyFrame(Frame):
    """What the searches below read of a KeyFrame: the Frame's grid (KeyFrame.cc:60 copies it), mvInvLevelSigma2,
    mvpMapPoints.  GetFeaturesInArea (KeyFrame.cc:737-781) has no level filter; IsInImage :783-786."""

    def __init__(self, *a, inv_level_sigma2=None, log_scale_factor=None, **k):
        super().__init__(*a, **k)
        self.mvInvLevelSigma2 = None if inv_level_sigma2 is None else np.asarray(inv_level_sigma2, F32)
        self.mfLogScaleFactor = F32(log_scale_factor)
        self.mnScaleLevels = len(self.mvScaleFactors)

    def GetFeaturesInAreaKF(self, x, y, r, bRight=False):
        return Frame.GetFeaturesInArea(self, x, y, r, -1, -1, bRight)

    def IsInImage(self, x, y):
        return bool(x >= self.mnMinX and x < self.mnMaxX and y >= self.mnMinY and y < self.mnMaxY)

    def GetMapPoint(self, idx):
        return self.mvpMapPoints[idx]


class MapPointKF(MapPoint):
    def __init__(self, descriptor, world_pos, normal, mf_min, mf_max, n_obs=1, bad=False):
        super().__init__(descriptor, n_obs, bad, np.asarray(world_pos, F32))
        self.mNormalVector = np.asarray(normal, F32)
        self.mfMinDistance, self.mfMaxDistance = F32(mf_min), F32(mf_max)
        self.in_kf = set()                       # ids of keyframes observing it

    def GetMinDistanceInvariance(self):
        return F32(F32(0.8) * self.mfMinDistance)

    def GetMaxDistanceInvariance(self):
        return F32(F32(1.2) * self.mfMaxDistance)

    def PredictScale(self, currentDist, pKF):                    # MapPoint.cc:534-549
        ratio = F32(self.mfMaxDistance / F32(currentDist))
        # float log / float division (std::log(float) under `using namespace std`); the correctly rounded float log agrees
        # with glibc's logf wherever the ceiling is concerned (tests/test_reference_mappoint.py)
        nScale = int(math.ceil(float(F32(F32(math.log(float(ratio))) / F32(pKF.mfLogScaleFactor)))))
        return 0 if nScale < 0 else min(nScale, pKF.mnScaleLevels - 1)

    def IsInKeyFrame(self, pKF):
        return id(pKF) in self.in_kf


def _norm(v):
    return F32(math.sqrt(sum(float(c) * float(c) for c in v)))


def _window_level(pMP, pKF, Ow):
    """Depth-invariance and viewing-angle gates + PredictScale (:523-541 and its copies)."""
    PO = (pMP.mWorldPos - Ow).astype(F32)
    dist = _norm(PO)
    if dist < pMP.GetMinDistanceInvariance() or dist > pMP.GetMaxDistanceInvariance():
        return None
    if sum(float(a) * float(b) for a, b in zip(PO, pMP.mNormalVector)) < 0.5 * float(dist):
        return None
    return pMP.PredictScale(dist, pKF)


def search_by_projection_sim3(pKF, vpPoints, vpMatched, th, ratioHamming, to_camera, project, Ow):
    """ORBmatcher.cc:473-585.  to_camera(p3Dw) = Rcw * p3Dw + tcw of the decomposed Scw.  Mutates vpMatched."""
    spAlreadyFound = {id(m) for m in vpMatched if m is not None}
    nmatches = 0
    for pMP in vpPoints:
        if pMP.isBad() or id(pMP) in spAlreadyFound:
            continue
        p3Dc = to_camera(pMP.mWorldPos)
        if p3Dc[2] < 0.0:
            continue
        u, v = project(p3Dc)
        if not pKF.IsInImage(u, v):
            continue
        nPredictedLevel = _window_level(pMP, pKF, Ow)
        if nPredictedLevel is None:
            continue
        radius = F32(F32(th) * pKF.mvScaleFactors[nPredictedLevel])
        vIndices = pKF.GetFeaturesInAreaKF(u, v, radius)
        if not vIndices:
            continue
        dMP = pMP.GetDescriptor()
        bestDist, bestIdx = 256, -1
        for idx in vIndices:
            if vpMatched[idx] is not None:
                continue
            kpLevel = int(pKF.keys[idx]["octave"])
            if kpLevel < nPredictedLevel - 1 or kpLevel > nPredictedLevel:
                continue
            dist = descriptor_distance(dMP, pKF.mDescriptors[idx])
            if dist < bestDist:
                bestDist, bestIdx = dist, idx
        if F32(bestDist) <= F32(F32(TH_LOW) * F32(ratioHamming)):
            vpMatched[bestIdx] = pMP
            nmatches += 1
    return nmatches


def fuse(pKF, vpMapPoints, th, to_camera, project, Ow, bRight=False):
    """ORBmatcher.cc:1629-1821.  Returns (nFused, [(index in vpMapPoints, bestIdx)]): the graph surgery of :1787-1812
    (Replace / AddObservation) is left to the caller, the searches do not depend on it."""
    nFused, picks = 0, []
    for i, pMP in enumerate(vpMapPoints):
        if pMP is None or pMP.isBad() or pMP.IsInKeyFrame(pKF):
            continue
        p3Dc = to_camera(pMP.mWorldPos)
        if p3Dc[2] < 0.0:
            continue
        invz = F32(F32(1) / p3Dc[2])
        u, v = project(p3Dc)
        if not pKF.IsInImage(u, v):
            continue
        ur = F32(u - F32(pKF.mbf * invz))
        nPredictedLevel = _window_level(pMP, pKF, Ow)
        if nPredictedLevel is None:
            continue
        radius = F32(F32(th) * pKF.mvScaleFactors[nPredictedLevel])
        vIndices = pKF.GetFeaturesInAreaKF(u, v, radius, bRight)
        if not vIndices:
            continue
        dMP = pMP.GetDescriptor()
        bestDist, bestIdx = 256, -1
        for idx in vIndices:
            kp = pKF.keys[idx] if (pKF.Nleft == -1 or not bRight) else pKF.keys[idx + pKF.Nleft]
            kpLevel = int(kp["octave"])
            if kpLevel < nPredictedLevel - 1 or kpLevel > nPredictedLevel:
                continue
            ex, ey = F32(u - kp["x"]), F32(v - kp["y"])
            if pKF.mvuRight[idx] >= 0:
                er = F32(ur - pKF.mvuRight[idx])
                e2 = F32(F32(F32(ex * ex) + F32(ey * ey)) + F32(er * er))
                if float(F32(e2 * pKF.mvInvLevelSigma2[kpLevel])) > 7.8:
                    continue
            else:
                e2 = F32(F32(ex * ex) + F32(ey * ey))
                if float(F32(e2 * pKF.mvInvLevelSigma2[kpLevel])) > 5.99:
                    continue
            if bRight:
                idx += pKF.Nleft
            dist = descriptor_distance(dMP, pKF.mDescriptors[idx])
            if dist < bestDist:
                bestDist, bestIdx = dist, idx
        if bestDist <= TH_LOW:
            picks.append((i, bestIdx))
            nFused += 1
    return nFused, picks


def fuse_sim3(pKF, vpPoints, th, to_camera, project, Ow, spAlreadyFound):
    """ORBmatcher.cc:1823-1945 (searches only, as in fuse())."""
    nFused, picks = 0, []
    for i, pMP in enumerate(vpPoints):
        if pMP.isBad() or id(pMP) in spAlreadyFound:
            continue
        p3Dc = to_camera(pMP.mWorldPos)
        if p3Dc[2] < 0.0:
            continue
        u, v = project(p3Dc)
        if not pKF.IsInImage(u, v):
            continue
        nPredictedLevel = _window_level(pMP, pKF, Ow)
        if nPredictedLevel is None:
            continue
        radius = F32(F32(th) * pKF.mvScaleFactors[nPredictedLevel])
        vIndices = pKF.GetFeaturesInAreaKF(u, v, radius)
        if not vIndices:
            continue
        dMP = pMP.GetDescriptor()
        bestDist, bestIdx = 2 ** 31 - 1, -1
        for idx in vIndices:
            kpLevel = int(pKF.keys[idx]["octave"])
            if kpLevel < nPredictedLevel - 1 or kpLevel > nPredictedLevel:
                continue
            dist = descriptor_distance(dMP, pKF.mDescriptors[idx])
            if dist < bestDist:
                bestDist, bestIdx = dist, idx
        if bestDist <= TH_LOW:
            picks.append((i, bestIdx))
            nFused += 1
    return nFused, picks


def search_by_sim3(pKF1, pKF2, vpMatches12, index_in_kf2, th, cam, to_c1, to_c2, c1_to_c2, c2_to_c1):
    """ORBmatcher.cc:1947-2171.  index_in_kf2(pMP) = get<0>(pMP->GetIndexInKeyFrame(pKF2)); to_c1 / to_c2 = camera
    from world; c1_to_c2(p) = sR21 * p + t21, c2_to_c1(p) = sR12 * p + t12.  Mutates vpMatches12; returns nFound and
    (vnMatch1, vnMatch2)."""
    fx, fy, cx, cy = (F32(c) for c in cam[:4])
    N1, N2 = pKF1.N, pKF2.N
    vbAlreadyMatched1, vbAlreadyMatched2 = [False] * N1, [False] * N2
    for i in range(N1):
        pMP = vpMatches12[i]
        if pMP is not None:
            vbAlreadyMatched1[i] = True
            idx2 = index_in_kf2(pMP)
            if 0 <= idx2 < N2:
                vbAlreadyMatched2[idx2] = True

    def one_way(src, dst, already, to_src, src_to_dst):
        vnMatch = [-1] * src.N
        for i in range(src.N):
            pMP = src.mvpMapPoints[i]
            if pMP is None or already[i] or pMP.isBad():
                continue
            pd = src_to_dst(to_src(pMP.mWorldPos))
            if pd[2] < 0.0:
                continue
            invz = F32(1.0 / float(pd[2]))
            x, y = F32(pd[0] * invz), F32(pd[1] * invz)
            u, v = F32(F32(fx * x) + cx), F32(F32(fy * y) + cy)
            if not dst.IsInImage(u, v):
                continue
            dist3D = _norm(pd)
            if dist3D < pMP.GetMinDistanceInvariance() or dist3D > pMP.GetMaxDistanceInvariance():
                continue
            nPredictedLevel = pMP.PredictScale(dist3D, dst)
            radius = F32(F32(th) * dst.mvScaleFactors[nPredictedLevel])
            vIndices = dst.GetFeaturesInAreaKF(u, v, radius)
            if not vIndices:
                continue
            dMP = pMP.GetDescriptor()
            bestDist, bestIdx = 2 ** 31 - 1, -1
            for idx in vIndices:
                kpo = int(dst.keys[idx]["octave"])
                if kpo < nPredictedLevel - 1 or kpo > nPredictedLevel:
                    continue
                dist = descriptor_distance(dMP, dst.mDescriptors[idx])
                if dist < bestDist:
                    bestDist, bestIdx = dist, idx
            if bestDist <= TH_HIGH:
                vnMatch[i] = bestIdx
        return vnMatch

    vnMatch1 = one_way(pKF1, pKF2, vbAlreadyMatched1, to_c1, c1_to_c2)
    vnMatch2 = one_way(pKF2, pKF1, vbAlreadyMatched2, to_c2, c2_to_c1)
    nFound = 0
    for i1 in range(N1):
        idx2 = vnMatch1[i1]
        if idx2 >= 0 and vnMatch2[idx2] == i1:
            vpMatches12[i1] = pKF2.mvpMapPoints[idx2]
            nFound += 1
    return nFound, vnMatch1, vnMatch2
