"""cv2-backed restatement of the reference's ORBextractor (TEST INFRASTRUCTURE, never shipped).

Follows /root/reference/src/ORBextractor.cc function by function and calls OpenCV (cv2 4.13, the
pinned oracle version) for exactly the primitives the reference calls: resize,
copyMakeBorder, FAST, GaussianBlur, fastAtan2.  It is the closest thing to "the reference itself"
that can run in the build container (the C++ reference needs OpenCV C++ headers, absent here).
Its outputs pin oracle/cpu_ref (the OpenCV-free C++ restatement) and are committed as golden
vectors under tests/golden/ by oracle/py/make_golden.py.

Only tests/ and the golden-vector generator import this module.
"""
import ctypes
import math
import os
import re

import numpy as np

try:  # cv2 exists in the build image; guarded so that importing the module never fails
    import cv2
except Exception:  # pragma: no cover
    cv2 = None

_libm = ctypes.CDLL("libm.so.6")
_libm.cosf.restype = ctypes.c_float
_libm.cosf.argtypes = [ctypes.c_float]
_libm.sinf.restype = ctypes.c_float
_libm.sinf.argtypes = [ctypes.c_float]

F32 = np.float32
PATCH_SIZE = 31
HALF_PATCH_SIZE = 15
EDGE_THRESHOLD = 19

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATTERN_FILE = os.path.join(_HERE, "..", "..", "orb-slam3-noted_b200", "csrc", "orb_pattern.inc")


def load_pattern():
    txt = re.sub(r"/\*.*?\*/", "", open(_PATTERN_FILE).read(), flags=re.S)
    v = np.array([int(x) for x in re.findall(r"-?\d+", txt)], dtype=np.int32)
    assert v.size == 1024
    return v.reshape(512, 2)  # 512 points (x, y)


def cv_round(x):
    """cvRound: round-half-to-even (ORBextractor.cc:79,113,117-118,440,458,1157)."""
    return int(np.rint(x))


class Node:
    __slots__ = ("UL", "UR", "BL", "BR", "keys", "no_more", "uid")

    def __init__(self):
        self.keys = []
        self.no_more = False
        self.uid = -1


class ORBextractorCv2:
    """ORBextractor.cc:408-468 (ctor) and :1068-1150 (operator())."""

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST):
        self.nfeatures = nfeatures
        self.scaleFactor = float(F32(scaleFactor))  # member is double, assigned from float
        self.nlevels = nlevels
        self.iniThFAST = iniThFAST
        self.minThFAST = minThFAST
        sf = [F32(1.0)]
        s2 = [F32(1.0)]
        for i in range(1, nlevels):
            sf.append(F32(float(sf[i - 1]) * self.scaleFactor))  # float*double -> float
            s2.append(F32(sf[i] * sf[i]))
        self.mvScaleFactor = np.array(sf, dtype=F32)
        self.mvLevelSigma2 = np.array(s2, dtype=F32)
        self.mvInvScaleFactor = (F32(1.0) / self.mvScaleFactor).astype(F32)
        self.mvInvLevelSigma2 = (F32(1.0) / self.mvLevelSigma2).astype(F32)
        factor = F32(1.0 / self.scaleFactor)
        denom = F32(1) - F32(math.pow(float(factor), float(nlevels)))
        nd = F32(F32(F32(nfeatures) * F32(F32(1) - factor)) / denom)
        self.mnFeaturesPerLevel = []
        total = 0
        for _ in range(nlevels - 1):
            n = cv_round(nd)
            self.mnFeaturesPerLevel.append(n)
            total += n
            nd = F32(nd * factor)
        self.mnFeaturesPerLevel.append(max(nfeatures - total, 0))
        self.pattern = load_pattern()
        # umax (ORBextractor.cc:452-467)
        umax = [0] * (HALF_PATCH_SIZE + 1)
        vmax = int(math.floor(float(F32(HALF_PATCH_SIZE) * F32(np.sqrt(F32(2.0))) / F32(2) + F32(1))))
        vmin = int(math.ceil(float(F32(HALF_PATCH_SIZE) * F32(np.sqrt(F32(2.0))) / F32(2))))
        hp2 = float(HALF_PATCH_SIZE * HALF_PATCH_SIZE)
        for v in range(vmax + 1):
            umax[v] = cv_round(math.sqrt(hp2 - v * v))
        v0 = 0
        for v in range(HALF_PATCH_SIZE, vmin - 1, -1):
            while umax[v0] == umax[v0 + 1]:
                v0 += 1
            umax[v] = v0
            v0 += 1
        self.umax = umax
        self.pyramid = []  # bordered buffers
        self.tie_rule = "newest"

    # ---- ORBextractor.cc:1152-1177
    def compute_pyramid(self, image):
        self.pyramid = []
        self.level_size = []
        E = EDGE_THRESHOLD
        for level in range(self.nlevels):
            scale = self.mvInvScaleFactor[level]
            w = cv_round(F32(F32(image.shape[1]) * scale))
            h = cv_round(F32(F32(image.shape[0]) * scale))
            if level != 0:
                prev = self.roi(level - 1)
                lvl = cv2.resize(prev, (w, h), interpolation=cv2.INTER_LINEAR)
                temp = cv2.copyMakeBorder(lvl, E, E, E, E, cv2.BORDER_REFLECT_101)
            else:
                temp = cv2.copyMakeBorder(image, E, E, E, E, cv2.BORDER_REFLECT_101)
            self.pyramid.append(temp)
            self.level_size.append((w, h))

    def roi(self, level):
        E = EDGE_THRESHOLD
        w, h = self.level_size[level]
        return self.pyramid[level][E:E + h, E:E + w]

    # ---- ORBextractor.cc:479-535
    @staticmethod
    def divide_node(nd):
        halfX = int(math.ceil(float(F32(nd.UR[0] - nd.UL[0]) / F32(2))))
        halfY = int(math.ceil(float(F32(nd.BR[1] - nd.UL[1]) / F32(2))))
        n1, n2, n3, n4 = Node(), Node(), Node(), Node()
        n1.UL = nd.UL
        n1.UR = (nd.UL[0] + halfX, nd.UL[1])
        n1.BL = (nd.UL[0], nd.UL[1] + halfY)
        n1.BR = (nd.UL[0] + halfX, nd.UL[1] + halfY)
        n2.UL = n1.UR
        n2.UR = nd.UR
        n2.BL = n1.BR
        n2.BR = (nd.UR[0], nd.UL[1] + halfY)
        n3.UL = n1.BL
        n3.UR = n1.BR
        n3.BL = nd.BL
        n3.BR = (n1.BR[0], nd.BL[1])
        n4.UL = n3.UR
        n4.UR = n2.BR
        n4.BL = n3.BR
        n4.BR = nd.BR
        for kp in nd.keys:
            if kp[0] < n1.UR[0]:
                (n1 if kp[1] < n1.BR[1] else n3).keys.append(kp)
            elif kp[1] < n1.BR[1]:
                n2.keys.append(kp)
            else:
                n4.keys.append(kp)
        for n in (n1, n2, n3, n4):
            if len(n.keys) == 1:
                n.no_more = True
        return n1, n2, n3, n4

    # ---- ORBextractor.cc:537-761.  keys: list of (x, y, response)
    def distribute_octree(self, keys, minX, maxX, minY, maxY, N):
        nIni = int(math.floor(float(F32(maxX - minX) / F32(maxY - minY)) + 0.5))  # round()
        hX = F32(F32(maxX - minX) / F32(nIni))
        nodes = []  # python list used as the std::list; index 0 = front
        uid = [0]

        def new_uid():
            uid[0] += 1
            return uid[0]

        ini = []
        for i in range(nIni):
            ni = Node()
            ni.UL = (int(F32(hX * F32(i))), 0)
            ni.UR = (int(F32(hX * F32(i + 1))), 0)
            ni.BL = (ni.UL[0], maxY - minY)
            ni.BR = (ni.UR[0], maxY - minY)
            ni.uid = new_uid()
            nodes.append(ni)
            ini.append(ni)
        for kp in keys:
            ini[int(F32(kp[0]) / hX)].keys.append(kp)
        kept = []
        for n in nodes:
            if len(n.keys) == 1:
                n.no_more = True
                kept.append(n)
            elif len(n.keys) > 1:
                kept.append(n)
        nodes = kept

        def push_children(children, exp_list):
            nexp = 0
            for c in children:
                if len(c.keys) > 0:
                    c.uid = new_uid()
                    nodes.insert(0, c)
                    if len(c.keys) > 1:
                        nexp += 1
                        exp_list.append(c)
            return nexp

        finish = False
        expandable = []
        while not finish:
            prev_size = len(nodes)
            n_to_expand = 0
            expandable = []
            # one pass over the list; children go to the front and are not revisited
            snapshot = list(nodes)
            for nd in snapshot:
                if nd.no_more:
                    continue
                ch = self.divide_node(nd)
                n_to_expand += push_children(ch, expandable)
                nodes.remove(nd)
            if len(nodes) >= N or len(nodes) == prev_size:
                finish = True
            elif len(nodes) + n_to_expand * 3 > N:
                while not finish:
                    prev_size = len(nodes)
                    prev = expandable
                    expandable = []
                    # sort ascending by (size, pointer); pointer order := creation order (documented rule)
                    prev_sorted = sorted(prev, key=lambda n: (len(n.keys), n.uid))
                    for nd in reversed(prev_sorted):
                        ch = self.divide_node(nd)
                        push_children(ch, expandable)
                        nodes.remove(nd)
                        if len(nodes) >= N:
                            break
                    if len(nodes) >= N or len(nodes) == prev_size:
                        finish = True
        out = []
        for nd in nodes:
            best = nd.keys[0]
            for kp in nd.keys[1:]:
                if kp[2] > best[2]:
                    best = kp
            out.append(best)
        return out

    # ---- ORBextractor.cc:763-878
    def compute_keypoints_octree(self):
        W = F32(35)
        all_kps = []
        self.candidates = []
        for level in range(self.nlevels):
            w, h = self.level_size[level]
            img = self.pyramid[level]
            minBX = EDGE_THRESHOLD - 3
            minBY = minBX
            maxBX = w - EDGE_THRESHOLD + 3
            maxBY = h - EDGE_THRESHOLD + 3
            width = F32(maxBX - minBX)
            height = F32(maxBY - minBY)
            nCols = int(width / W)
            nRows = int(height / W)
            wCell = int(math.ceil(float(width / F32(nCols))))
            hCell = int(math.ceil(float(height / F32(nRows))))
            cand = []
            E = EDGE_THRESHOLD
            for i in range(nRows):
                iniY = minBY + i * hCell
                maxY = iniY + hCell + 6
                if iniY >= maxBY - 3:
                    continue
                if maxY > maxBY:
                    maxY = maxBY
                for j in range(nCols):
                    iniX = minBX + j * wCell
                    maxX = iniX + wCell + 6
                    if iniX >= maxBX - 6:
                        continue
                    if maxX > maxBX:
                        maxX = maxBX
                    # rowRange/colRange are relative to the level ROI inside the bordered buffer
                    cell = img[E + iniY:E + maxY, E + iniX:E + maxX]
                    kps = cv2.FastFeatureDetector_create(self.iniThFAST, True).detect(np.ascontiguousarray(cell))
                    if len(kps) == 0:
                        kps = cv2.FastFeatureDetector_create(self.minThFAST, True).detect(np.ascontiguousarray(cell))
                    for kp in kps:
                        cand.append((float(F32(kp.pt[0]) + F32(j * wCell)), float(F32(kp.pt[1]) + F32(i * hCell)),
                                     float(kp.response)))
            self.candidates.append(cand)
            sel = self.distribute_octree(cand, minBX, maxBX, minBY, maxBY, self.mnFeaturesPerLevel[level])
            size = float(int(F32(PATCH_SIZE) * self.mvScaleFactor[level]))
            kl = []
            for (x, y, r) in sel:
                kl.append([x + minBX, y + minBY, size, -1.0, r, level])
            all_kps.append(kl)
        for level in range(self.nlevels):
            roi_img = self.pyramid[level]
            for kp in all_kps[level]:
                kp[3] = self.ic_angle(roi_img, kp[0], kp[1])
        return all_kps

    # ---- ORBextractor.cc:75-102 (image = bordered buffer, pt relative to ROI)
    def ic_angle(self, bordered, x, y):
        cx = cv_round(x) + EDGE_THRESHOLD
        cy = cv_round(y) + EDGE_THRESHOLD
        m01 = 0
        m10 = 0
        row = bordered[cy].astype(np.int64)
        for u in range(-HALF_PATCH_SIZE, HALF_PATCH_SIZE + 1):
            m10 += u * int(row[cx + u])
        for v in range(1, HALF_PATCH_SIZE + 1):
            d = self.umax[v]
            p = bordered[cy + v, cx - d:cx + d + 1].astype(np.int64)
            m = bordered[cy - v, cx - d:cx + d + 1].astype(np.int64)
            u = np.arange(-d, d + 1, dtype=np.int64)
            m01 += v * int((p - m).sum())
            m10 += int((u * (p + m)).sum())
        return float(cv2.fastAtan2(float(m01), float(m10)))

    # ---- ORBextractor.cc:106-145
    def compute_descriptor(self, blurred, x, y, angle_deg):
        factorPI = F32(F32(math.pi) / F32(180.0))
        angle = F32(F32(angle_deg) * factorPI)
        a = F32(_libm.cosf(float(angle)))
        b = F32(_libm.sinf(float(angle)))
        cx = cv_round(x)
        cy = cv_round(y)
        px = self.pattern[:, 0].astype(F32)
        py = self.pattern[:, 1].astype(F32)
        ry = np.rint((px * b).astype(F32) + (py * a).astype(F32)).astype(np.int64)
        rx = np.rint((px * a).astype(F32) - (py * b).astype(F32)).astype(np.int64)
        vals = blurred[cy + ry, cx + rx].astype(np.int32)
        bits = (vals[0::2] < vals[1::2]).astype(np.uint8)  # 256 tests
        return np.packbits(bits.reshape(32, 8), axis=1, bitorder="little").reshape(32)

    # ---- ORBextractor.cc:1068-1150
    def __call__(self, image, lapping=(0, 0)):
        if image is None or image.size == 0:
            return None, None, -1
        assert image.dtype == np.uint8 and image.ndim == 2
        self.compute_pyramid(image)
        all_kps = self.compute_keypoints_octree()
        self.level_keypoints = [[list(k) for k in lv] for lv in all_kps]
        n = sum(len(k) for k in all_kps)
        kp_out = np.zeros((n, 7), dtype=np.float64)
        desc_out = np.zeros((n, 32), dtype=np.uint8)
        mono = 0
        stereo = n - 1
        for level in range(self.nlevels):
            kl = all_kps[level]
            if not kl:
                continue
            work = self.roi(level).copy()
            work = cv2.GaussianBlur(work, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
            scale = self.mvScaleFactor[level]
            for kp in kl:
                d = self.compute_descriptor(work, kp[0], kp[1], kp[3])
                x, y = F32(kp[0]), F32(kp[1])
                if level != 0:
                    x = F32(x * scale)
                    y = F32(y * scale)
                row = [float(x), float(y), kp[2], kp[3], kp[4], kp[5], -1]
                if x >= F32(lapping[0]) and x <= F32(lapping[1]):
                    kp_out[stereo] = row
                    desc_out[stereo] = d
                    stereo -= 1
                else:
                    kp_out[mono] = row
                    desc_out[mono] = d
                    mono += 1
        return kp_out, desc_out, mono
