/*
 * synth.c — deterministic synthetic textured grayscale frames.
 *
 * Pure integer arithmetic (SplitMix64), no libm, so the same seed gives the same bytes on every
 * machine.  Used by tests, bench.py and the CPU baseline to feed *identical* inputs to the CUDA
 * path and to the oracle.  This is input generation, not part of the hot path.
 *
 * A frame is a window into a "canvas": mid-grey background, W*H/600 filled rectangles / discs /
 * triangles of random grey, +-6 LSB hash noise, one 3x3 box smooth.  A stereo right view is the
 * same canvas sampled at x + d(y) where d is one of three integer disparity planes (4, 17, 40 px)
 * selected by horizontal band, so left/right correspondences are known exactly.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline uint64_t sm64(uint64_t *s) {
    uint64_t z = (*s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static inline uint64_t hash3(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t s = a * 0xD6E8FEB86659FD93ull + b * 0xCA5A826395121157ull + c;
    return sm64(&s);
}

#define SYNTH_MARGIN 64 /* canvas is W + 64 wide so that disparities up to 64 never leave it */

/* Render the canvas (cw x h) for `seed`. */
static void render_canvas(uint64_t seed, int cw, int h, uint8_t *cv) {
    uint64_t st = seed * 0x2545F4914F6CDD1Dull + 0x1234567ull;
    memset(cv, 128, (size_t)cw * h);
    int nshapes = (int)(((int64_t)cw * h) / 600);
    for (int k = 0; k < nshapes; ++k) {
        uint64_t r = sm64(&st);
        int type = (int)(r % 3);          r >>= 2;
        int cx = (int)(r % (uint64_t)cw); r >>= 12;
        int cy = (int)(r % (uint64_t)h);  r >>= 12;
        int s = 5 + (int)(r % 36);        r >>= 8;
        int s2 = 5 + (int)(r % 36);       r >>= 8;
        uint8_t g = (uint8_t)(r & 0xFF);
        int x0 = cx - s, x1 = cx + s, y0 = cy - s2, y1 = cy + s2;
        if (type == 1) { y0 = cy - s; y1 = cy + s; }
        if (x0 < 0) x0 = 0;
        if (y0 < 0) y0 = 0;
        if (x1 >= cw) x1 = cw - 1;
        if (y1 >= h) y1 = h - 1;
        for (int y = y0; y <= y1; ++y) {
            uint8_t *row = cv + (size_t)y * cw;
            for (int x = x0; x <= x1; ++x) {
                int in;
                if (type == 0) in = 1;
                else if (type == 1) in = (x - cx) * (x - cx) + (y - cy) * (y - cy) <= s * s;
                else {
                    /* triangle (cx, cy-s2), (cx-s, cy+s2), (cx+s, cy+s2/2): half-plane tests */
                    int ax = cx, ay = cy - s2, bx = cx - s, by = cy + s2, ccx = cx + s, ccy = cy + s2 / 2;
                    int e0 = (bx - ax) * (y - ay) - (by - ay) * (x - ax);
                    int e1 = (ccx - bx) * (y - by) - (ccy - by) * (x - bx);
                    int e2 = (ax - ccx) * (y - ccy) - (ay - ccy) * (x - ccx);
                    in = (e0 <= 0 && e1 <= 0 && e2 <= 0) || (e0 >= 0 && e1 >= 0 && e2 >= 0);
                }
                if (in) row[x] = g;
            }
        }
    }
    /* hash noise, +-6 LSB */
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < cw; ++x) {
            int v = cv[(size_t)y * cw + x] + (int)(hash3(seed, (uint64_t)x, (uint64_t)y) % 13) - 6;
            cv[(size_t)y * cw + x] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
        }
    /* 3x3 box smooth, replicate border */
    uint8_t *tmp = (uint8_t *)malloc((size_t)cw * h);
    memcpy(tmp, cv, (size_t)cw * h);
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < cw; ++x) {
            int acc = 0;
            for (int dy = -1; dy <= 1; ++dy) {
                int yy = y + dy; yy = yy < 0 ? 0 : yy >= h ? h - 1 : yy;
                for (int dx = -1; dx <= 1; ++dx) {
                    int xx = x + dx; xx = xx < 0 ? 0 : xx >= cw ? cw - 1 : xx;
                    acc += tmp[(size_t)yy * cw + xx];
                }
            }
            cv[(size_t)y * cw + x] = (uint8_t)((acc + 4) / 9);
        }
    free(tmp);
}

/* Disparity plane of row y (three horizontal bands). */
static inline int band_disparity(int y, int h) {
    static const int d[3] = {4, 17, 40};
    int b = (3 * y) / h;
    return d[b > 2 ? 2 : b];
}

/*
 * orbfe_synth_frame: write one w x h frame (row stride = w) for `seed`.
 *   view = 0: left / mono view           out(x,y) = canvas(x + shift, y)
 *   view = 1: right view of a stereo pair out(x,y) = canvas(x + shift + d(y), y), plus +-1 LSB
 *             independent noise so the two views are not byte-identical.
 *   shift: horizontal camera shift in pixels, 0..(64 - 40); used to make "next keyframe" views.
 * Returns 0, or -1 on bad arguments.
 */
int orbfe_synth_frame(uint64_t seed, int w, int h, int view, int shift, uint8_t *out) {
    if (w <= 0 || h <= 0 || !out || shift < 0 || shift + 40 > SYNTH_MARGIN) return -1;
    int cw = w + SYNTH_MARGIN;
    uint8_t *cv = (uint8_t *)malloc((size_t)cw * h);
    if (!cv) return -1;
    render_canvas(seed, cw, h, cv);
    for (int y = 0; y < h; ++y) {
        int off = shift + (view ? band_disparity(y, h) : 0);
        const uint8_t *src = cv + (size_t)y * cw + off;
        uint8_t *dst = out + (size_t)y * w;
        if (!view) memcpy(dst, src, (size_t)w);
        else
            for (int x = 0; x < w; ++x) {
                int v = src[x] + (int)(hash3(seed ^ 0xABCDEFull, (uint64_t)x, (uint64_t)y) % 3) - 1;
                dst[x] = (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v);
            }
    }
    free(cv);
    return 0;
}

/* Ground-truth disparity of row y for the stereo pair produced above. */
int orbfe_synth_disparity(int y, int h) { return band_disparity(y, h); }

/* Deterministic random descriptors: n rows of 32 bytes. */
void orbfe_synth_descriptors(uint64_t seed, int n, uint8_t *out) {
    uint64_t st = seed ^ 0x5DEECE66Dull;
    for (int i = 0; i < n * 4; ++i) {
        uint64_t r = sm64(&st);
        memcpy(out + (size_t)i * 8, &r, 8);
    }
}
