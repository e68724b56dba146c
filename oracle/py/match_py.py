"""Python restatement of the reference's Hamming matchers (TEST INFRASTRUCTURE, never shipped).

Follows /root/reference/src/ORBmatcher.cc and src/Frame.cc loop by loop, written independently of
oracle/cpu_ref/match_ref.cc, and calls OpenCV (cv2 4.13) where the reference does: cv2.norm(NORM_L1)
for the SAD window (Frame.cc:908) and cv2.BFMatcher(NORM_HAMMING).knnMatch for the fisheye lapping
match (Frame.cc:1134).  Its outputs are committed as tests/golden/match_*.npz by
oracle/py/make_golden_match.py and pin the C++ oracle.  float32 arithmetic is spelled out with
numpy scalars because the reference computes in `float`.
"""
import math

import numpy as np

try:
    import cv2
except Exception:  # pragma: no cover
    cv2 = None

F32 = np.float32
TH_HIGH, TH_LOW, HISTO_LENGTH = 100, 50, 30   # ORBmatcher.cc:36-38


def descriptor_distance(a, b):
    """ORBmatcher.cc:2561-2577 — the SWAR popcount, on 8 int32 words."""
    pa = np.frombuffer(np.ascontiguousarray(a).tobytes(), dtype="<u4")
    pb = np.frombuffer(np.ascontiguousarray(b).tobytes(), dtype="<u4")
    dist = 0
    for i in range(8):
        v = int(pa[i]) ^ int(pb[i])
        v = (v - ((v >> 1) & 0x55555555)) & 0xFFFFFFFF
        v = ((v & 0x33333333) + ((v >> 2) & 0x33333333)) & 0xFFFFFFFF
        dist += ((((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) & 0xFFFFFFFF) >> 24
    return dist


def hamming_matrix(A, B):
    """All-pairs distances (same numbers as descriptor_distance, vectorised for speed)."""
    x = np.bitwise_xor(A[:, None, :], B[None, :, :])
    return np.unpackbits(x, axis=2).sum(axis=2).astype(np.int32)


def c_round(x):
    """C round(): half away from zero (ORBmatcher.cc:394, Frame.cc:880-882)."""
    x = float(x)
    return math.floor(x + 0.5) if x >= 0 else -math.floor(-x + 0.5)


def compute_three_maxima(counts):
    """ORBmatcher.cc:2515-2556."""
    max1 = max2 = max3 = 0
    ind1 = ind2 = ind3 = -1
    for i, s in enumerate(counts):
        if s > max1:
            max3, max2, max1 = max2, max1, s
            ind3, ind2, ind1 = ind2, ind1, i
        elif s > max2:
            max3, max2 = max2, s
            ind3, ind2 = ind2, i
        elif s > max3:
            max3, ind3 = s, i
    if F32(max2) < F32(0.1) * F32(max1):
        ind2 = ind3 = -1
    elif F32(max3) < F32(0.1) * F32(max1):
        ind3 = -1
    return ind1, ind2, ind3


def rot_bin(a1, a2):
    factor = F32(1.0) / F32(HISTO_LENGTH)
    rot = F32(F32(a1) - F32(a2))
    if rot < 0.0:
        rot = F32(rot + F32(360.0))
    b = int(c_round(F32(rot * factor)))
    return 0 if b == HISTO_LENGTH else b


def compute_stereo_matches(pyrL, pyrR, level_size, scale, inv_scale, kpL, descL, kpR, descR, mbf, mb):
    """Frame.cc:794-964.  pyrL/pyrR: bordered levels (19 px) of the two extractors; level_size[l] = (w, h);
    kp*: structured arrays (x, y, octave).  Returns (mvuRight, mvDepth, bestDist per accepted iL or -1)."""
    E = 19
    N, Nr = len(kpL), len(kpR)
    mvuRight = np.full(N, -1.0, F32)
    mvDepth = np.full(N, -1.0, F32)
    sad = np.full(N, -1, np.int32)
    thOrbDist = (TH_HIGH + TH_LOW) // 2
    nRows = level_size[0][1]
    rows = [[] for _ in range(nRows)]
    for iR in range(Nr):
        kpY = F32(kpR["y"][iR])
        r = F32(F32(2.0) * scale[kpR["octave"][iR]])
        maxr = int(math.ceil(float(F32(kpY + r))))
        minr = int(math.floor(float(F32(kpY - r))))
        for yi in range(max(minr, 0), min(maxr, nRows - 1) + 1):
            rows[yi].append(iR)
    mbf, mb = F32(mbf), F32(mb)
    minD, maxD = F32(0), F32(mbf / mb)
    D = hamming_matrix(descL, descR) if N and Nr else np.zeros((N, Nr), np.int32)
    vDistIdx = []
    for iL in range(N):
        levelL = int(kpL["octave"][iL])
        vL, uL = F32(kpL["y"][iL]), F32(kpL["x"][iL])
        cands = rows[int(vL)]
        if not cands:
            continue
        minU, maxU = F32(uL - maxD), F32(uL - minD)
        if maxU < 0:
            continue
        bestDist, bestIdxR = TH_HIGH, 0
        for iR in cands:
            o = int(kpR["octave"][iR])
            if o < levelL - 1 or o > levelL + 1:
                continue
            uR = F32(kpR["x"][iR])
            if uR >= minU and uR <= maxU:
                d = int(D[iL, iR])
                if d < bestDist:
                    bestDist, bestIdxR = d, iR
        if bestDist >= thOrbDist:
            continue
        uR0 = F32(kpR["x"][bestIdxR])
        sf = inv_scale[levelL]
        scaleduL = c_round(F32(uL * sf))
        scaledvL = c_round(F32(vL * sf))
        scaleduR0 = c_round(F32(uR0 * sf))
        w = L = 5
        PL, PR = pyrL[levelL], pyrR[levelL]
        IL = PL[E + scaledvL - w:E + scaledvL + w + 1, E + scaleduL - w:E + scaleduL + w + 1]
        iniu, endu = scaleduR0 + L - w, scaleduR0 + L + w + 1
        if iniu < 0 or endu >= level_size[levelL][0]:
            continue
        best, bestinc = 2 ** 31 - 1, 0
        vD = [F32(0)] * (2 * L + 1)
        for inc in range(-L, L + 1):
            IR = PR[E + scaledvL - w:E + scaledvL + w + 1, E + scaleduR0 + inc - w:E + scaleduR0 + inc + w + 1]
            dist = F32(cv2.norm(np.ascontiguousarray(IL), np.ascontiguousarray(IR), cv2.NORM_L1))
            if dist < best:
                best, bestinc = int(dist), inc
            vD[L + inc] = dist
        if bestinc == -L or bestinc == L:
            continue
        d1, d2, d3 = vD[L + bestinc - 1], vD[L + bestinc], vD[L + bestinc + 1]
        with np.errstate(divide="ignore", invalid="ignore"):
            deltaR = F32(F32(d1 - d3) / F32(F32(2.0) * F32(F32(d1 + d3) - F32(F32(2.0) * d2))))
        if deltaR < -1 or deltaR > 1:
            continue
        bestuR = F32(scale[levelL] * F32(F32(F32(scaleduR0) + F32(bestinc)) + deltaR))
        disparity = F32(uL - bestuR)
        if disparity >= minD and disparity < maxD:
            if disparity <= 0:
                disparity = F32(0.01)
                bestuR = F32(float(uL) - 0.01)
            mvDepth[iL] = F32(mbf / disparity)
            mvuRight[iL] = bestuR
            vDistIdx.append((best, iL))
    if vDistIdx:
        vDistIdx.sort()
        median = F32(vDistIdx[len(vDistIdx) // 2][0])
        thDist = F32(F32(F32(1.5) * F32(1.4)) * median)
        for d, i in vDistIdx:
            sad[i] = d
        for d, i in reversed(vDistIdx):
            if F32(d) < thDist:
                break
            mvuRight[i] = -1
            mvDepth[i] = -1
    return mvuRight, mvDepth, sad


def shared_nodes(fvA, fvB):
    """The merge-join of ORBmatcher.cc:289-429.  fv: dict node id -> list of feature indices
    (a std::map iterates in ascending key order)."""
    ka, kb = sorted(fvA), sorted(fvB)
    i = j = 0
    while i < len(ka) and j < len(kb):
        if ka[i] == kb[j]:
            yield fvA[ka[i]], fvB[kb[j]]
            i += 1
            j += 1
        elif ka[i] < kb[j]:
            while i < len(ka) and ka[i] < kb[j]:   # lower_bound
                i += 1
        else:
            while j < len(kb) and kb[j] < ka[i]:
                j += 1


def search_by_bow_kf_frame(descKF, fvKF, mpKF, angKF, descF, fvF, angF, nnratio, check_ori, nleft=-1):
    """ORBmatcher.cc:269-471.  Returns (vpMapPointMatches as KF index per Frame feature or -1, nmatches)."""
    N = len(descF)
    matches = [-1] * N
    nmatches = 0
    rotHist = [[] for _ in range(HISTO_LENGTH)]
    for idsKF, idsF in shared_nodes(fvKF, fvF):
        for realIdxKF in idsKF:
            if not mpKF[realIdxKF]:
                continue
            bd1, bi, bd2 = 256, -1, 256
            bd1R, biR, bd2R = 256, -1, 256
            for realIdxF in idsF:
                if matches[realIdxF] >= 0:
                    continue
                dist = descriptor_distance(descKF[realIdxKF], descF[realIdxF])
                if nleft == -1 or realIdxF < nleft:
                    if dist < bd1:
                        bd2, bd1, bi = bd1, dist, realIdxF
                    elif dist < bd2:
                        bd2 = dist
                else:
                    if dist < bd1R:
                        bd2R, bd1R, biR = bd1R, dist, realIdxF
                    elif dist < bd2R:
                        bd2R = dist
            if bd1 <= TH_LOW:
                if F32(bd1) < F32(nnratio) * F32(bd2):
                    matches[bi] = realIdxKF
                    if check_ori:
                        rotHist[rot_bin(angKF[realIdxKF], angF[bi])].append(bi)
                    nmatches += 1
                if nleft != -1 and bd1R <= TH_LOW:
                    matches[biR] = realIdxKF
                    if check_ori:
                        rotHist[rot_bin(angKF[realIdxKF], angF[biR])].append(biR)
                    nmatches += 1
    if check_ori:
        keep = compute_three_maxima([len(h) for h in rotHist])
        for i in range(HISTO_LENGTH):
            if i in keep:
                continue
            for j in rotHist[i]:
                matches[j] = -1
                nmatches -= 1
    return np.array(matches, np.int32), nmatches


def search_by_bow_kf_kf(desc1, fv1, mp1, ang1, desc2, fv2, mp2, ang2, nnratio, check_ori):
    """ORBmatcher.cc:823-963.  Returns (vpMatches12 as index into KF2 or -1, nmatches)."""
    m12 = [-1] * len(desc1)
    matched2 = [False] * len(desc2)
    nmatches = 0
    rotHist = [[] for _ in range(HISTO_LENGTH)]
    for ids1, ids2 in shared_nodes(fv1, fv2):
        for idx1 in ids1:
            if not mp1[idx1]:
                continue
            bd1, bi2, bd2 = 256, -1, 256
            for idx2 in ids2:
                if matched2[idx2] or not mp2[idx2]:
                    continue
                dist = descriptor_distance(desc1[idx1], desc2[idx2])
                if dist < bd1:
                    bd2, bd1, bi2 = bd1, dist, idx2
                elif dist < bd2:
                    bd2 = dist
            if bd1 < TH_LOW:
                if F32(bd1) < F32(nnratio) * F32(bd2):
                    m12[idx1] = bi2
                    matched2[bi2] = True
                    if check_ori:
                        rotHist[rot_bin(ang1[idx1], ang2[bi2])].append(idx1)
                    nmatches += 1
    if check_ori:
        keep = compute_three_maxima([len(h) for h in rotHist])
        for i in range(HISTO_LENGTH):
            if i in keep:
                continue
            for j in rotHist[i]:
                m12[j] = -1
                nmatches -= 1
    return np.array(m12, np.int32), nmatches


def epipolar_constrain(F12, kp1, kp2, unc):
    """Pinhole.cpp:159-181 with F12 formed by the caller (row-major 3x3 float32)."""
    x1, y1, x2, y2 = F32(kp1[0]), F32(kp1[1]), F32(kp2[0]), F32(kp2[1])
    a = F32(F32(F32(x1 * F12[0]) + F32(y1 * F12[3])) + F12[6])
    b = F32(F32(F32(x1 * F12[1]) + F32(y1 * F12[4])) + F12[7])
    c = F32(F32(F32(x1 * F12[2]) + F32(y1 * F12[5])) + F12[8])
    num = F32(F32(F32(a * x2) + F32(b * y2)) + c)
    den = F32(F32(a * a) + F32(b * b))
    if den == 0:
        return False
    dsqr = F32(F32(num * num) / den)
    return float(dsqr) < 3.84 * float(F32(unc))


def search_for_triangulation(kp1, desc1, fv1, mp1, ur1, kp2, desc2, fv2, mp2, ur2, sigma2, scale, F12, ep, check_ori,
                             only_stereo, coarse):
    """ORBmatcher.cc:1208-1432, pinhole, no second camera.  kp: structured arrays.  Returns (vMatches12, n)."""
    F12 = np.asarray(F12, F32).reshape(9)
    m12 = [-1] * len(desc1)
    nmatches = 0
    rotHist = [[] for _ in range(HISTO_LENGTH)]
    for ids1, ids2 in shared_nodes(fv1, fv2):
        for idx1 in ids1:
            if mp1[idx1]:
                continue
            st1 = ur1[idx1] >= 0
            if only_stereo and not st1:
                continue
            bestDist, bestIdx2 = TH_LOW, -1
            for idx2 in ids2:
                if mp2[idx2]:
                    continue
                st2 = ur2[idx2] >= 0
                if only_stereo and not st2:
                    continue
                dist = descriptor_distance(desc1[idx1], desc2[idx2])
                if dist > TH_LOW or dist > bestDist:
                    continue
                if not st1 and not st2:
                    dx = F32(F32(ep[0]) - F32(kp2["x"][idx2]))
                    dy = F32(F32(ep[1]) - F32(kp2["y"][idx2]))
                    if F32(F32(dx * dx) + F32(dy * dy)) < F32(F32(100) * scale[kp2["octave"][idx2]]):
                        continue
                if coarse or epipolar_constrain(F12, (kp1["x"][idx1], kp1["y"][idx1]), (kp2["x"][idx2], kp2["y"][idx2]),
                                                sigma2[kp2["octave"][idx2]]):
                    bestIdx2, bestDist = idx2, dist
            if bestIdx2 >= 0:
                m12[idx1] = bestIdx2
                nmatches += 1
                if check_ori:
                    rotHist[rot_bin(kp1["angle"][idx1], kp2["angle"][bestIdx2])].append(idx1)
    if check_ori:
        keep = compute_three_maxima([len(h) for h in rotHist])
        for i in range(HISTO_LENGTH):
            if i in keep:
                continue
            for j in rotHist[i]:
                m12[j] = -1
                nmatches -= 1
    return np.array(m12, np.int32), nmatches


def search_for_triangulation_kb8(kp1, desc1, fv1, mp1, sigma2_1, kp2, desc2, fv2, mp2, sigma2_2, scale2, cam_a, cam_b, nleft_a,
                                 nleft_b, R12, t12, ep, check_ori, only_stereo, coarse):
    """ORBmatcher.cc:1208-1432 with Kannala-Brandt cameras: the geometric test is
    KannalaBrandt8::epipolarConstrain_ (KannalaBrandt8.cpp:239-242) = TriangulateMatches_ > 1e-4, the
    (R12, t12, cameras) are picked per pair from {ll, lr, rl, rr} when both keyframes have two cameras
    (:1332-1366).  mvuRight stays -1 with fisheye cameras, so bStereo1 = bStereo2 = false.
    Returns (vMatches12, n, point per match (n1, 3), margin per kp1)."""
    n1 = len(desc1)
    m12 = [-1] * n1
    pts = np.zeros((n1, 3), F32)
    margins = np.full(n1, 1e30, F32)
    nmatches = 0
    rotHist = [[] for _ in range(HISTO_LENGTH)]
    dual = nleft_a >= 0 and nleft_b >= 0
    for ids1, ids2 in shared_nodes(fv1, fv2):
        for idx1 in ids1:
            if mp1[idx1] or only_stereo:
                continue
            right1 = not (nleft_a == -1 or idx1 < nleft_a)
            bestDist, bestIdx2, bestP = TH_LOW, -1, None
            for idx2 in ids2:
                if mp2[idx2]:
                    continue
                dist = descriptor_distance(desc1[idx1], desc2[idx2])
                if dist > TH_LOW or dist > bestDist:
                    continue
                right2 = not (nleft_b == -1 or idx2 < nleft_b)
                if nleft_a < 0:
                    dx = F32(F32(ep[0]) - F32(kp2["x"][idx2]))
                    dy = F32(F32(ep[1]) - F32(kp2["y"][idx2]))
                    if F32(F32(dx * dx) + F32(dy * dy)) < F32(F32(100) * scale2[kp2["octave"][idx2]]):
                        continue
                c = 2 * int(right1) + int(right2) if dual else 0
                c1 = cam_a[1] if dual and right1 else cam_a[0]
                c2 = cam_b[1] if dual and right2 else cam_b[0]
                d, p, mg = kb8_triangulate_matches(c1, c2, (kp1["x"][idx1], kp1["y"][idx1]), (kp2["x"][idx2], kp2["y"][idx2]),
                                                   R12[c], t12[c], sigma2_1[kp1["octave"][idx1]], sigma2_2[kp2["octave"][idx2]])
                if not coarse:
                    margins[idx1] = min(margins[idx1], mg)
                if coarse or d > 0.0001:
                    bestIdx2, bestDist, bestP = idx2, dist, p
            if bestIdx2 >= 0:
                m12[idx1] = bestIdx2
                if bestP is not None:
                    pts[idx1] = np.asarray(bestP, F32).reshape(3)
                nmatches += 1
                if check_ori:
                    rotHist[rot_bin(kp1["angle"][idx1], kp2["angle"][bestIdx2])].append(idx1)
    if check_ori:
        keep = compute_three_maxima([len(h) for h in rotHist])
        for i in range(HISTO_LENGTH):
            if i in keep:
                continue
            for j in rotHist[i]:
                m12[j] = -1
                nmatches -= 1
    return np.array(m12, np.int32), nmatches, pts, margins


def fisheye_lapping_match(descL, descR):
    """Frame.cc:1134-1141: BFMatcher(NORM_HAMMING).knnMatch(k=2) + Lowe ratio 0.7.  Returns
    (match index per query or -1, knn idx (n,2), knn dist (n,2))."""
    m = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(np.ascontiguousarray(descL), np.ascontiguousarray(descR), k=2)
    out = np.full(len(descL), -1, np.int32)
    idx = np.full((len(descL), 2), -1, np.int32)
    dist = np.full((len(descL), 2), -1, np.int32)
    for i, it in enumerate(m):
        for k, dm in enumerate(it[:2]):
            idx[i, k], dist[i, k] = dm.trainIdx, int(dm.distance)
        if len(it) >= 2 and it[0].distance < it[1].distance * 0.7:
            out[i] = it[0].trainIdx
    return out, idx, dist


# ---- Kannala-Brandt fisheye stereo (Frame.cc:1116-1156, KannalaBrandt8.cpp) ------------------------------
def kb8_project(cam, p):
    """KannalaBrandt8.cpp:24-40.  cam = (fx, fy, cx, cy, k0..k3) float32."""
    x, y, z = (F32(v) for v in p)
    x2y2 = F32(F32(x * x) + F32(y * y))
    theta = F32(math.atan2(float(F32(np.sqrt(x2y2))), float(z)))          # atan2f
    psi = F32(math.atan2(float(y), float(x)))
    t2 = F32(theta * theta); t3 = F32(theta * t2); t5 = F32(t3 * t2); t7 = F32(t5 * t2); t9 = F32(t7 * t2)
    r = F32(F32(F32(F32(theta + F32(cam[4] * t3)) + F32(cam[5] * t5)) + F32(cam[6] * t7)) + F32(cam[7] * t9))
    u = F32(float(F32(cam[0] * r)) * math.cos(float(psi)) + float(cam[2]))     # cos(float) -> double arithmetic
    v = F32(float(F32(cam[1] * r)) * math.sin(float(psi)) + float(cam[3]))
    return u, v


def kb8_unproject(cam, u, v):
    """KannalaBrandt8.cpp:96-123, precision 1e-6."""
    pwx = F32(F32(F32(u) - cam[2]) / cam[0]); pwy = F32(F32(F32(v) - cam[3]) / cam[1])
    scale = F32(1)
    theta_d = F32(np.sqrt(F32(F32(pwx * pwx) + F32(pwy * pwy))))
    theta_d = min(max(F32(-math.pi / 2), theta_d), F32(math.pi / 2))
    if theta_d > 1e-8:
        theta = theta_d
        for _ in range(10):
            t2 = F32(theta * theta); t4 = F32(t2 * t2); t6 = F32(t4 * t2); t8 = F32(t4 * t4)
            a, b, c, d = F32(cam[4] * t2), F32(cam[5] * t4), F32(cam[6] * t6), F32(cam[7] * t8)
            num = F32(F32(theta * F32(F32(F32(F32(F32(1) + a) + b) + c) + d)) - theta_d)
            den = F32(F32(F32(F32(F32(1) + F32(F32(3) * a)) + F32(F32(5) * b)) + F32(F32(7) * c)) + F32(F32(9) * d))
            fix = F32(num / den)
            theta = F32(theta - fix)
            if abs(fix) < F32(1e-6):
                break
        scale = F32(F32(math.tan(float(theta))) / theta_d)
    return np.array([F32(pwx * scale), F32(pwy * scale), F32(1)], F32)


def kb8_triangulate_matches(cam1, cam2, pt1, pt2, R12, t12, sigma_level, unc):
    """KannalaBrandt8::TriangulateMatches (KannalaBrandt8.cpp:338-407) with the real cv2.SVDecomp on the float32
    4x4 system (Triangulate, :514-526).  Returns (z1 or -1, p3D, margin)."""
    R12 = np.asarray(R12, F32).reshape(3, 3)
    t12 = np.asarray(t12, F32).reshape(3, 1)
    r1 = kb8_unproject(cam1, pt1[0], pt1[1]).reshape(3, 1)
    r2 = kb8_unproject(cam2, pt2[0], pt2[1]).reshape(3, 1)
    r21 = (R12 @ r2).astype(F32)
    cosp = F32(float(r1.astype(np.float64).T @ r21.astype(np.float64)) / (cv2.norm(r1) * cv2.norm(r21)))
    margin = abs(float(cosp) - 0.9998) / 2e-4      # relative to 1 - threshold: cosines cluster next to 1
    if cosp > 0.9998:
        return -1.0, None, margin
    R21 = R12.T.copy()
    t21 = (-(R21 @ t12)).astype(F32)
    Tcw1 = np.hstack([np.eye(3, dtype=F32), np.zeros((3, 1), F32)])
    Tcw2 = np.hstack([R21, t21]).astype(F32)
    A = np.zeros((4, 4), F32)
    A[0] = r1[0, 0] * Tcw1[2] - Tcw1[0]
    A[1] = r1[1, 0] * Tcw1[2] - Tcw1[1]
    A[2] = r2[0, 0] * Tcw2[2] - Tcw2[0]
    A[3] = r2[1, 0] * Tcw2[2] - Tcw2[1]
    w, u, vt = cv2.SVDecomp(A, flags=cv2.SVD_MODIFY_A | cv2.SVD_FULL_UV)
    xh = vt[3].astype(F32)
    x3D = (xh[:3] / xh[3]).astype(F32)
    z1 = x3D[2]
    if z1 <= 0:
        return -1.0, None, margin
    z2 = F32(float(R21[2].astype(np.float64) @ x3D.astype(np.float64)) + float(t21[2, 0]))
    if z2 <= 0:
        return -1.0, None, margin
    u1, v1 = kb8_project(cam1, x3D)
    e1 = F32(F32(F32(u1 - F32(pt1[0])) ** 2) + F32(F32(v1 - F32(pt1[1])) ** 2))
    margin = min(margin, abs(float(e1) - 5.991 * float(sigma_level)) / (5.991 * float(sigma_level)))
    if float(e1) > 5.991 * float(F32(sigma_level)):
        return -1.0, None, margin
    x3D2 = ((R21 @ x3D.reshape(3, 1)) + t21).astype(F32).reshape(3)
    u2, v2 = kb8_project(cam2, x3D2)
    e2 = F32(F32(F32(u2 - F32(pt2[0])) ** 2) + F32(F32(v2 - F32(pt2[1])) ** 2))
    margin = min(margin, abs(float(e2) - 5.991 * float(unc)) / (5.991 * float(unc)))
    if float(e2) > 5.991 * float(F32(unc)):
        return -1.0, None, margin
    return float(z1), x3D, margin


def compute_stereo_fisheye_matches(cam1, cam2, kpL, descL, mono_left, kpR, descR, mono_right, R12, t12, sigma2):
    """Frame.cc:1116-1156.  Returns (mvLeftToRightMatch, mvRightToLeftMatch, mvDepth, p3D (n, 3), margins, nMatches)."""
    nL, nR = len(kpL), len(kpR)
    l2r = np.full(nL, -1, np.int32)
    r2l = np.full(nR, -1, np.int32)
    depth = np.full(nL, -1.0, F32)
    p3d = np.zeros((nL, 3), F32)
    margins = np.full(nL, 1e30, F32)
    match, _, _ = fisheye_lapping_match(descL[mono_left:], descR[mono_right:])
    n = 0
    for q, tr in enumerate(match):
        if tr < 0:
            continue
        iL, iR = q + mono_left, int(tr) + mono_right
        d, p, mg = kb8_triangulate_matches(cam1, cam2, (kpL["x"][iL], kpL["y"][iL]), (kpR["x"][iR], kpR["y"][iR]), R12, t12,
                                           sigma2[kpL["octave"][iL]], sigma2[kpR["octave"][iR]])
        margins[iL] = min(mg, abs(d - 1e-4) / 1e-4)
        if d > 0.0001:
            l2r[iL], r2l[iR], depth[iL], p3d[iL] = iR, iL, d, p
            n += 1
    return l2r, r2l, depth, p3d, margins, n


def compute_distinctive_descriptor(descs):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:388-417) on the list of observed descriptors.
    Returns the index of the chosen one (-1 for an empty list)."""
    N = len(descs)
    if N == 0:
        return -1
    D = np.zeros((N, N), np.float32)
    for i in range(N):
        for j in range(i + 1, N):
            D[i, j] = D[j, i] = cv2.norm(descs[i], descs[j], cv2.NORM_HAMMING)
    best_median, best_idx = 2 ** 31 - 1, 0
    for i in range(N):
        v = sorted(int(x) for x in D[i])
        median = v[int(0.5 * (N - 1))]
        if median < best_median:
            best_median, best_idx = median, i
    return best_idx
