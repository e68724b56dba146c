// cvprim.cc — see cvprim.h.  TEST INFRASTRUCTURE: oracle + timed CPU baseline only.
#include "cvprim.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>

namespace orbref {

// ---------------------------------------------------------------------------------------------
// resize, INTER_LINEAR, 8-bit single channel.  OpenCV: 11-bit fixed-point coefficients
// (INTER_RESIZE_COEF_BITS), horizontal pass into int32 rows, vertical pass
//   out = (((b0 * (r0 >> 4)) >> 16) + ((b1 * (r1 >> 4)) >> 16) + 2) >> 2.
// x: source index clamped AND fraction zeroed at the ends; y: rows clamped only.
// ---------------------------------------------------------------------------------------------
void resizeLinearU8(const uint8_t* src, int sw, int sh, ptrdiff_t sstep, uint8_t* dst, int dw, int dh,
                    ptrdiff_t dstep) {
    const double inv_sx = (double)dw / sw, inv_sy = (double)dh / sh;
    const double scale_x = 1. / inv_sx, scale_y = 1. / inv_sy;
    std::vector<int> xofs(dw), xofs1(dw);
    std::vector<short> xa0(dw), xa1(dw);
    for (int dx = 0; dx < dw; ++dx) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = (int)std::floor(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
        xofs[dx] = sx;
        xofs1[dx] = std::min(sx + 1, sw - 1);
        xa0[dx] = (short)cvRoundF((1.f - fx) * 2048.f);
        xa1[dx] = (short)cvRoundF(fx * 2048.f);
    }
    std::vector<int> row0(dw), row1(dw);
    int cached0 = -1, cached1 = -1;
    auto hresize = [&](int sy, std::vector<int>& row) {
        const uint8_t* S = src + (ptrdiff_t)sy * sstep;
        for (int dx = 0; dx < dw; ++dx) row[dx] = S[xofs[dx]] * xa0[dx] + S[xofs1[dx]] * xa1[dx];
    };
    for (int dy = 0; dy < dh; ++dy) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = (int)std::floor(fy);
        fy -= sy;
        const short b0 = (short)cvRoundF((1.f - fy) * 2048.f), b1 = (short)cvRoundF(fy * 2048.f);
        const int y0 = std::min(std::max(sy, 0), sh - 1), y1 = std::min(std::max(sy + 1, 0), sh - 1);
        if (y0 == cached1) { row0.swap(row1); std::swap(cached0, cached1); }
        if (y0 != cached0) { hresize(y0, row0); cached0 = y0; }
        if (y1 == y0) { row1 = row0; cached1 = y1; }
        else if (y1 != cached1) { hresize(y1, row1); cached1 = y1; }
        uint8_t* D = dst + (ptrdiff_t)dy * dstep;
        for (int dx = 0; dx < dw; ++dx) {
            int v = (((b0 * (row0[dx] >> 4)) >> 16) + ((b1 * (row1[dx] >> 4)) >> 16) + 2) >> 2;
            D[dx] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
        }
    }
}

void fillBorderReflect101(uint8_t* buf, int w, int h, int b, ptrdiff_t step) {
    // left/right of interior rows
    for (int y = 0; y < h; ++y) {
        uint8_t* row = buf + (ptrdiff_t)(y + b) * step + b;
        for (int x = -b; x < 0; ++x) row[x] = row[reflect101(x, w)];
        for (int x = w; x < w + b; ++x) row[x] = row[reflect101(x, w)];
    }
    // top/bottom rows copy whole (already horizontally extended) rows
    for (int y = -b; y < 0; ++y)
        std::memcpy(buf + (ptrdiff_t)(y + b) * step, buf + (ptrdiff_t)(reflect101(y, h) + b) * step, (size_t)(w + 2 * b));
    for (int y = h; y < h + b; ++y)
        std::memcpy(buf + (ptrdiff_t)(y + b) * step, buf + (ptrdiff_t)(reflect101(y, h) + b) * step, (size_t)(w + 2 * b));
}

// ---------------------------------------------------------------------------------------------
// FAST 9-16.  d_k = I(c) - I(ring_k).  a = max_k min(d_k..d_{k+8}), b = max_k min(-d_k..-d_{k+8}),
// corner iff max(a,b) > t, score = max(a,b) - 1 (largest threshold that keeps it a corner).
// NMS: keep iff score strictly greater than the 8 neighbours' scores (0 outside the tested area).
// ---------------------------------------------------------------------------------------------
static const int kRingDx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int kRingDy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

static inline int arcScore(const int* d) {  // d has 25 entries (wrapped)
    int best = -256;
    for (int k = 0; k < 16; ++k) {
        int m = d[k];
        for (int j = 1; j < 9; ++j) m = std::min(m, d[k + j]);
        best = std::max(best, m);
    }
    return best;
}

void fast9_16(const uint8_t* img, int w, int h, ptrdiff_t step, int t, std::vector<FastKp>& out) {
    out.clear();
    if (w < 7 || h < 7) return;
    int off[16];
    for (int k = 0; k < 16; ++k) off[k] = kRingDy[k] * (int)step + kRingDx[k];
    // three rolling score rows (prev-prev handled by emitting row y-1 once row y is known)
    std::vector<int> sc((size_t)3 * w, 0);
    auto rowp = [&](int y) { return sc.data() + (size_t)((y + 3) % 3) * w; };
    for (int y = 3; y <= h - 3; ++y) {
        int* cur = rowp(y);
        std::fill(cur, cur + w, 0);
        if (y < h - 3) {
            const uint8_t* p = img + (ptrdiff_t)y * step;
            for (int x = 3; x < w - 3; ++x) {
                const int c = p[x];
                const int lo = c - t, hi = c + t;
                // opposite-pair quick reject: a 9-arc contains ring k or ring k+8 for every k
                const int v0 = p[x + off[0]], v8 = p[x + off[8]];
                bool dark = (v0 < lo) || (v8 < lo), bright = (v0 > hi) || (v8 > hi);
                if (!dark && !bright) continue;
                const int v4 = p[x + off[4]], v12 = p[x + off[12]];
                dark = dark && ((v4 < lo) || (v12 < lo));
                bright = bright && ((v4 > hi) || (v12 > hi));
                if (!dark && !bright) continue;
                int d[25];
                unsigned md = 0, mb = 0;
                for (int k = 0; k < 16; ++k) {
                    const int v = p[x + off[k]];
                    d[k] = c - v;
                    md |= (unsigned)(v < lo) << k;
                    mb |= (unsigned)(v > hi) << k;
                }
                auto has9 = [](unsigned m) {
                    m |= m << 16;
                    unsigned r = m & (m >> 1);
                    r &= r >> 2;
                    r &= r >> 4;
                    r &= m >> 8;
                    return (r & 0xFFFFu) != 0;
                };
                if (!has9(md) && !has9(mb)) continue;
                for (int k = 16; k < 25; ++k) d[k] = d[k - 16];
                const int a = arcScore(d);
                int nd[25];
                for (int k = 0; k < 25; ++k) nd[k] = -d[k];
                const int b = arcScore(nd);
                cur[x] = std::max(a, b) - 1;
            }
        }
        if (y == 3) continue;
        // NMS for row y-1
        const int* pr = rowp(y - 1);
        const int* pp = rowp(y - 2);
        // (row 2 is never scored: its slot still holds the zeros it was initialised with)
        for (int x = 3; x < w - 3; ++x) {
            const int s = pr[x];
            if (s <= 0) continue;   // non-corners score 0 and can never be strictly greater
            if (s > pr[x - 1] && s > pr[x + 1] && s > pp[x - 1] && s > pp[x] && s > pp[x + 1] && s > cur[x - 1] &&
                s > cur[x] && s > cur[x + 1])
                out.push_back({x, y - 1, s});
        }
    }
}

// ---------------------------------------------------------------------------------------------
// GaussianBlur 7x7 sigma 2, 8-bit: OpenCV >= 3.4.2 fixed-point path, 8.8 kernel
// {18,34,48,56,48,34,18} (sum 256), dst = (sum_j q_j * (sum_i q_i * src) + 32768) >> 16.
// ---------------------------------------------------------------------------------------------
void gaussianBlur7x7(const uint8_t* src, int w, int h, ptrdiff_t sstep, uint8_t* dst, ptrdiff_t dstep) {
    static const int q[7] = {18, 34, 48, 56, 48, 34, 18};
    std::vector<uint16_t> hbuf((size_t)w * h);
    std::vector<int> xi((size_t)w + 6);
    for (int x = -3; x < w + 3; ++x) xi[x + 3] = reflect101(x, w);
    for (int y = 0; y < h; ++y) {
        const uint8_t* S = src + (ptrdiff_t)y * sstep;
        uint16_t* H = hbuf.data() + (size_t)y * w;
        if (w >= 7) {
            for (int x = 0; x < 3; ++x) { int a = 0; for (int i = 0; i < 7; ++i) a += q[i] * S[xi[x + i]]; H[x] = (uint16_t)a; }
            for (int x = 3; x < w - 3; ++x) {
                const uint8_t* s = S + x - 3;
                H[x] = (uint16_t)(18 * (s[0] + s[6]) + 34 * (s[1] + s[5]) + 48 * (s[2] + s[4]) + 56 * s[3]);
            }
            for (int x = std::max(3, w - 3); x < w; ++x) { int a = 0; for (int i = 0; i < 7; ++i) a += q[i] * S[xi[x + i]]; H[x] = (uint16_t)a; }
        } else {
            for (int x = 0; x < w; ++x) { int a = 0; for (int i = 0; i < 7; ++i) a += q[i] * S[xi[x + i]]; H[x] = (uint16_t)a; }
        }
    }
    for (int y = 0; y < h; ++y) {
        const uint16_t* r[7];
        for (int j = 0; j < 7; ++j) r[j] = hbuf.data() + (size_t)reflect101(y + j - 3, h) * w;
        uint8_t* D = dst + (ptrdiff_t)y * dstep;
        for (int x = 0; x < w; ++x) {
            uint32_t v = 18u * (r[0][x] + r[6][x]) + 34u * (r[1][x] + r[5][x]) + 48u * (r[2][x] + r[4][x]) + 56u * r[3][x];
            D[x] = (uint8_t)((v + 32768u) >> 16);
        }
    }
}

// fastAtan2 (degrees): OpenCV's 7th-order odd polynomial, unfused float32.
float fastAtan2(float y, float x) {
    const float s = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * s, p3 = -0.3258083974640975f * s, p5 = 0.1555786518463281f * s,
                p7 = -0.04432655554792128f * s;
    const float ax = std::fabs(x), ay = std::fabs(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

int hammingDistance256(const uint8_t* a, const uint8_t* b) {
    uint64_t x[4], y[4];
    std::memcpy(x, a, 32);
    std::memcpy(y, b, 32);
    return __builtin_popcountll(x[0] ^ y[0]) + __builtin_popcountll(x[1] ^ y[1]) + __builtin_popcountll(x[2] ^ y[2]) +
           __builtin_popcountll(x[3] ^ y[3]);
}

void knn2Hamming(const uint8_t* q, int nq, const uint8_t* t, int nt, int32_t* idx, int32_t* dist) {
    for (int i = 0; i < nq; ++i) {
        int b0 = 1 << 30, b1 = 1 << 30, i0 = -1, i1 = -1;
        for (int j = 0; j < nt; ++j) {
            const int d = hammingDistance256(q + (size_t)i * 32, t + (size_t)j * 32);
            if (d < b0) { b1 = b0; i1 = i0; b0 = d; i0 = j; }
            else if (d < b1) { b1 = d; i1 = j; }
        }
        idx[2 * i] = i0; idx[2 * i + 1] = i1;
        dist[2 * i] = i0 < 0 ? -1 : b0; dist[2 * i + 1] = i1 < 0 ? -1 : b1;
    }
}

}  // namespace orbref
