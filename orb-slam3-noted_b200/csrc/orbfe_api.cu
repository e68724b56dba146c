// orbfe_api.cu — the C ABI (include/orbfe.h): handle life cycle, per-shape geometry, the extract
// entry points and the result / pyramid taps.  Host side of the hot path; all pixel work is in the
// k_*.cu kernels.  There is no CPU fallback: without a CUDA device every call returns an error.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <mutex>
#include <utility>
#include <vector>

#include "orbfe_internal.h"
// Measured (profiles/r02v_blur_sweep.txt): the blur on a side stream under FAST + quadtree gains nothing - 184.3 k frames/s
// serial, 184.5 k with the ordinary grid on the side stream (the block scheduler runs the two kernels one after the other),
// 149 / 173 / 179 / 180.5 k with 1 / 2 / 3 / 4 resident blur CTAs per SM (four blur warps per SM are too slow, more take the
// registers FAST needs).  Off by default; the knobs stay for other shapes.
#ifndef ORBFE_PYR_TMA_DEFAULT
#define ORBFE_PYR_TMA_DEFAULT 1    // resize levels through k_pyr_resize_tma (TMA-staged source): 0.659 -> 0.581 ms per 512 frames
#endif
#ifndef ORBFE_PIPE_STREAMS
#define ORBFE_PIPE_STREAMS 2       // compute streams of the host-buffer pipeline (orbfe_extract_batch)
#endif
#ifndef ORBFE_BLUR_DEFAULT
#define ORBFE_BLUR_DEFAULT ORBFE_BLUR_ALU
#endif
#ifndef ORBFE_BLUR_OVERLAP
#define ORBFE_BLUR_OVERLAP 0
#endif
#ifndef ORBFE_BLUR_PERSIST
#define ORBFE_BLUR_PERSIST 4
#endif
#ifndef ORBFE_FAST_BOX80
#define ORBFE_FAST_BOX80 0     // FAST tile pitch 80 instead of 64 bytes (bank spread of the tile rows); see k_fast.cu
#endif

static thread_local char g_cudaErr[256] = "";

static int cuda_fail(cudaError_t e, const char* what) {
    snprintf(g_cudaErr, sizeof(g_cudaErr), "%s: %s", what, cudaGetErrorString(e));
    return ORBFE_ERR_CUDA;
}
int orbfe_set_cuda_error(cudaError_t e, const char* what) { return cuda_fail(e, what); }   // for the other .cu files
#define CK(call)                                             \
    do {                                                     \
        cudaError_t e_ = (call);                             \
        if (e_ != cudaSuccess) return cuda_fail(e_, #call);  \
    } while (0)

static inline int cvRoundF(float v) { return (int)lrintf(v); }   // half-to-even

cudaError_t orbfe_ensure_dyn_smem(const void* func, size_t bytes) {
    if (bytes <= 48 * 1024) return cudaSuccess;
    static std::mutex mu;
    static std::map<std::pair<int, const void*>, size_t> granted;   // high-water mark per (device, kernel)
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lock(mu);
    size_t& g = granted[std::make_pair(dev, func)];
    if (bytes <= g) return cudaSuccess;
    e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e == cudaSuccess) g = bytes;
    return e;
}

// ---- reference tables, ORBextractor.cc:413-444 ----------------------------------------------------
static void build_tables(orbfe_handle* h) {
    const int nl = h->p.nlevels;
    const double sf = (double)h->p.scale_factor;   // the member is a double assigned from a float
    h->scale[0] = 1.0f;
    h->sigma2[0] = 1.0f;
    for (int i = 1; i < nl; i++) {
        h->scale[i] = (float)(h->scale[i - 1] * sf);
        h->sigma2[i] = h->scale[i] * h->scale[i];
    }
    for (int i = 0; i < nl; i++) {
        h->invScale[i] = 1.0f / h->scale[i];
        h->invSigma2[i] = 1.0f / h->sigma2[i];
    }
    const float factor = (float)(1.0f / sf);
    float nDesired = h->p.nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nl));
    int sum = 0;
    for (int l = 0; l < nl - 1; l++) {
        h->featPerLevel[l] = cvRoundF(nDesired);
        sum += h->featPerLevel[l];
        nDesired *= factor;
    }
    h->featPerLevel[nl - 1] = std::max(h->p.nfeatures - sum, 0);
}

// OpenCV resize(INTER_LINEAR) taps for one axis (src n -> dst m).  clampFrac: the x axis zeroes the
// fraction at the clamped ends, the y axis only clamps the row index.
static void build_taps(int n, int m, bool clampFrac, ResizeTap* out) {
    const double scale = 1. / ((double)m / n);
    for (int d = 0; d < m; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)floorf(f);
        f -= s;
        if (clampFrac) {
            if (s < 0) { f = 0; s = 0; }
            if (s >= n - 1) { f = 0; s = n - 1; }
        }
        ResizeTap t;
        t.s0 = (short)std::min(std::max(s, 0), n - 1);
        t.s1 = (short)std::min(std::max(s + 1, 0), n - 1);
        t.a0 = (short)cvRoundF((1.f - f) * 2048.f);
        t.a1 = (short)cvRoundF(f * 2048.f);
        out[d] = t;
    }
}

// Column plans of one level (see ColPlan / k_pyr_resize_plan).  Leaves g.planBase = -1 if some column
// pair's source bytes do not fit an 8-byte window (scale factors above ~2).
static void build_col_plans(LevelGeom& g, const ResizeTap* tx, std::vector<ColPlan>& plans) {
    const int nwords = g.pitch / 4;
    const size_t first = plans.size();
    plans.resize(first + nwords);
    for (int word = 0; word < nwords; ++word) {
        ColPlan cp;
        memset(&cp, 0, sizeof(cp));
        int s0[4], s1[4];
        for (int k = 0; k < 4; ++k) {
            const int x = word * 4 - ORBFE_XO + k;
            const bool in = x >= -ORBFE_EDGE && x < g.w + ORBFE_EDGE;
            int xr = x < 0 ? -x : x;
            if (xr >= g.w) xr = 2 * g.w - 2 - xr;
            // columns outside the border are masked off by `keep`; give them the tap of the nearest stored column
            // so that they do not widen the source window
            const int xc = in ? xr : (x < 0 ? ORBFE_EDGE : g.w - 2 - (ORBFE_EDGE - 1));
            const ResizeTap t = tx[std::min(std::max(xc, 0), g.w - 1)];
            s0[k] = t.s0; s1[k] = t.s1;
            cp.a[k] = (uint32_t)(uint16_t)t.a0 | (uint32_t)(uint16_t)t.a1 << 16;
            if (in) cp.keep |= 0xFFu << (8 * k);
        }
        uint32_t sel = 0;
        for (int pair = 0; pair < 2; ++pair) {
            const int b[4] = {s0[2 * pair], s1[2 * pair], s0[2 * pair + 1], s1[2 * pair + 1]};
            const int lo = std::min(std::min(b[0], b[1]), std::min(b[2], b[3]));
            const int hi = std::max(std::max(b[0], b[1]), std::max(b[2], b[3]));
            const int base = lo >> 2;
            if (hi - base * 4 > 7) { plans.resize(first); g.planBase = -1; return; }
            (pair ? cp.base23 : cp.base01) = base;
            uint32_t sl = 0;
            for (int j = 0; j < 4; ++j) sl |= (uint32_t)(b[j] - base * 4) << (4 * j);
            sel |= sl << (16 * pair);
        }
        cp.sel = sel;
        plans[first + word] = cp;
    }
    g.planBase = (int)first;
}

struct ShapePlan {
    LevelGeom geom[ORBFE_MAX_LEVELS];
    std::vector<ColPlan> plans;
    std::vector<CellDesc> cells;
    std::vector<ResizeTap> taps;
    size_t pyrBytesPerFrame = 0;
    long long candPerFrame = 0;
    int kpPerFrame = 0;
    int maxCellBytes = 0, maxCw = 0, maxCh = 0, maxKept = 0, boxW = 0, boxH = 0;
    int blurTiles = 0;
    int status = ORBFE_OK;
};

// Geometry for one image shape: level sizes (ORBextractor.cc:1156-1157), FAST grid (:771-805),
// quadtree roots (:541-543), arena offsets.  Pure host arithmetic.
static void plan_shape(const orbfe_handle* h, int W, int H, ShapePlan& sp) {
    const int nl = h->p.nlevels;
    long long pyrOff = 0, candOff = 0;
    int kpOff = 0, tapOff = 0, blurTiles = 0;
    for (int l = 0; l < nl; ++l) {
        LevelGeom& g = sp.geom[l];
        memset(&g, 0, sizeof(g));
        g.w = cvRoundF((float)W * h->invScale[l]);
        g.h = cvRoundF((float)H * h->invScale[l]);
        g.maxBX = g.w - ORBFE_EDGE + 3;
        g.maxBY = g.h - ORBFE_EDGE + 3;
        const float width = (float)(g.maxBX - ORBFE_MIN_BORDER), height = (float)(g.maxBY - ORBFE_MIN_BORDER);
        g.nCols = (int)(width / 35.f);
        g.nRows = (int)(height / 35.f);
        if (g.w < 64 || g.h < 64 || g.nCols < 1 || g.nRows < 1 || g.w > 2047 + ORBFE_MIN_BORDER || g.h > 2047 + ORBFE_MIN_BORDER) {
            sp.status = ORBFE_ERR_TOO_LARGE;   // level too small for a FAST cell / too large for the packed coordinates
            return;
        }
        g.wCell = (int)ceilf(width / g.nCols);
        g.hCell = (int)ceilf(height / g.nRows);
        if (g.wCell + 3 > 127 || g.hCell + 3 > 127) { sp.status = ORBFE_ERR_TOO_LARGE; return; }
        g.pitch = (ORBFE_XO + g.w + ORBFE_EDGE + 127) & ~127;
        g.rows = g.h + 2 * ORBFE_EDGE;
        g.plane = (long long)g.pitch * g.rows;
        g.base = pyrOff;   // frame stride applied later (needs batch): base = pyrOff * maxBatch
        pyrOff += g.plane;
        g.nfeat = h->featPerLevel[l];
        g.nIni = (int)roundf((float)(g.maxBX - ORBFE_MIN_BORDER) / (g.maxBY - ORBFE_MIN_BORDER));
        if (g.nIni < 1) { sp.status = ORBFE_ERR_TOO_LARGE; return; }
        g.hX = (float)(g.maxBX - ORBFE_MIN_BORDER) / g.nIni;
        g.scale = h->scale[l];
        g.size = (float)(int)(31 * h->scale[l]);
        g.kpCap = std::max(g.nfeat, 4 * g.nIni) + 3;
        if (g.kpCap > ORBFE_MAX_NODES) { sp.status = ORBFE_ERR_TOO_LARGE; return; }
        g.kpCap = (g.kpCap + 3) & ~3;
        g.kpBase = kpOff;
        kpOff += g.kpCap;
        // cells, in the reference's loop order
        g.cellBase = (int)sp.cells.size();
        int lvlMaxCh = 8;
        int cap = 0;
        for (int i = 0; i < g.nRows; i++) {
            const int iniY = ORBFE_MIN_BORDER + i * g.hCell;
            int maxY = iniY + g.hCell + 6;
            if (iniY >= g.maxBY - 3) continue;
            if (maxY > g.maxBY) maxY = g.maxBY;
            for (int j = 0; j < g.nCols; j++) {
                const int iniX = ORBFE_MIN_BORDER + j * g.wCell;
                int maxX = iniX + g.wCell + 6;
                if (iniX >= g.maxBX - 6) continue;
                if (maxX > g.maxBX) maxX = g.maxBX;
                CellDesc c;
                memset(&c, 0, sizeof(c));
                c.level = (short)l; c.x0 = (short)iniX; c.y0 = (short)iniY;
                c.cw = (short)(maxX - iniX); c.ch = (short)(maxY - iniY);
                c.ox = (short)(j * g.wCell); c.oy = (short)(i * g.hCell);
                const int tw = c.cw - 6, th = c.ch - 6;
                if (tw <= 0 || th <= 0) continue;   // FAST on a < 7 px ROI finds nothing
                {   // stage A of k_fast_cells walks the tile in 32-bit words (tile byte = sh + column), tested columns [3, cw - 3)
                    const int sh = (ORBFE_XO + c.x0) & 15, wLo = (sh + 3) >> 2, wHi = (sh + c.cw - 4) >> 2, nWords = wHi - wLo + 1;
                    c.sh = (short)sh; c.wLo = (short)wLo; c.nWords = (short)nWords; c.units = (short)(th * nWords);
                    c.inv = ((1u << 20) + nWords - 1) / nWords;
                    c.tx = (short)(ORBFE_XO + c.x0 - sh); c.ty = (short)(ORBFE_EDGE + c.y0);
                    const int kLo = std::max(3 - (4 * wLo - sh), 0), kHi = std::min(c.cw - 3 - (4 * wHi - sh), 4);
                    c.firstMask = kLo >= 4 ? 0u : 0x80808080u << (8 * kLo);
                    c.lastMask = kHi <= 0 ? 0u : 0x80808080u >> (8 * (4 - kHi));
                    if (nWords == 1) { c.firstMask &= c.lastMask; c.lastMask = c.firstMask; }
                }
                sp.cells.push_back(c);
                const int keptCap = ((tw + 1) / 2) * ((th + 1) / 2);
                cap += keptCap;
                sp.maxCw = std::max(sp.maxCw, (int)c.cw);
                sp.maxCh = std::max(sp.maxCh, (int)c.ch);
                lvlMaxCh = std::max(lvlMaxCh, (int)c.ch);
                sp.maxKept = std::max(sp.maxKept, keptCap);
            }
        }
        g.nCells = (int)sp.cells.size() - g.cellBase;
        g.fastBoxH = lvlMaxCh;
        g.candCap = (cap + 3) & ~3;
        g.candBase = candOff;
        candOff += g.candCap;
        // resize taps (level l from level l-1)
        g.blurTileBase = blurTiles;
        g.blurTilesX = (g.w + 127) / 128;
        blurTiles += g.blurTilesX * ((g.h + ORBFE_BLUR_TH - 1) / ORBFE_BLUR_TH);
        g.tabX = tapOff;
        g.tabY = tapOff + g.w;
        tapOff += g.w + g.h;
        sp.taps.resize(tapOff);
        g.planBase = -1;
        if (l > 0) {
            build_taps(sp.geom[l - 1].w, g.w, true, &sp.taps[g.tabX]);
            build_taps(sp.geom[l - 1].h, g.h, false, &sp.taps[g.tabY]);
            build_col_plans(g, &sp.taps[g.tabX], sp.plans);
        }
    }
    sp.pyrBytesPerFrame = (size_t)pyrOff;
    sp.candPerFrame = candOff;
    sp.kpPerFrame = kpOff;
    sp.blurTiles = blurTiles;
    // FAST tile = TMA box: inner extent a multiple of 16 bytes; layout of k_fast_cells' dynamic smem
    sp.boxW = (sp.maxCw + 15 + 15) & ~15;   // + up to 15 bytes of alignment shift (TMA start column is 16-byte aligned)
#if ORBFE_FAST_BOX80
    if (sp.boxW == 64) sp.boxW = 80;        // 20-word rows: rows r and r + 1 sit 20 banks apart, only rows 8 apart share banks
#endif
    sp.boxH = sp.maxCh;
    const int listCap = ((sp.boxW - 6) * (sp.boxH - 6) + 1) & ~1;
    sp.maxCellBytes = 2 * sp.boxW * sp.boxH + 2 * 2 * listCap + 4 * sp.maxKept + 16;
    if (sp.boxW > 256 || sp.boxH > 256) sp.status = ORBFE_ERR_TOO_LARGE;
}

// ---- handle life cycle ---------------------------------------------------------------------------
static void free_handle(orbfe_handle* h) {
    if (!h) return;
    cudaSetDevice(h->device);
    cudaFree(h->dIn); cudaFree(h->dPyr); cudaFree(h->dBlur); cudaFree(h->dCand); cudaFree(h->dNodeOf); cudaFree(h->dCandCount);
    cudaFree(h->dLvlKp); cudaFree(h->dLvlCount); cudaFree(h->dGeom); cudaFree(h->dCells); cudaFree(h->dTaps); cudaFree(h->dPlans); cudaFree(h->dPyrTma);
    cudaFree(h->dRes); cudaFree(h->dSlot);
    if (h->hRes) cudaFreeHost(h->hRes);
    for (int i = 0; i < ORBFE_GRAPH_SLOTS; ++i)
        if (h->graphs[i].exec) cudaGraphExecDestroy(h->graphs[i].exec);
    if (h->hN) cudaFreeHost(h->hN);
    if (h->hMono) cudaFreeHost(h->hMono);
    if (h->profEv) {
        for (int i = 0; i < ORBFE_PROF_RING; ++i)
            for (int s = 0; s <= ORBFE_NUM_STAGES; ++s) cudaEventDestroy(h->profEv[i][s]);
        delete[] h->profEv;
    }
    for (int c = 0; c < ORBFE_PIPE_MAX_CHUNKS; ++c)
        for (int k = 0; k < 3; ++k)
            if (h->evPipe[c][k]) cudaEventDestroy(h->evPipe[c][k]);
    if (h->evFork) cudaEventDestroy(h->evFork);
    if (h->evOrder) cudaEventDestroy(h->evOrder);
    for (int i = 0; i < ORBFE_MAX_COMPUTE_STREAMS; ++i) {
        if (h->evJoin[i]) cudaEventDestroy(h->evJoin[i]);
        if (h->evPyr[i]) cudaEventDestroy(h->evPyr[i]);
        if (h->evBlur[i]) cudaEventDestroy(h->evBlur[i]);
        if (h->sSide[i]) cudaStreamDestroy(h->sSide[i]);
    }
    if (h->sIn) cudaStreamDestroy(h->sIn);
    if (h->sOut) cudaStreamDestroy(h->sOut);
    for (int i = 1; i < ORBFE_MAX_COMPUTE_STREAMS; ++i)
        if (h->sComp[i]) cudaStreamDestroy(h->sComp[i]);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time dependency on libcuda).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int encode_fast_maps(orbfe_handle* h) {
    static EncodeTiledFn encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) return cuda_fail(e != cudaSuccess ? e : cudaErrorUnknown, "cuTensorMapEncodeTiled entry point");
        encode = (EncodeTiledFn)fn;
    }
    const size_t B = (size_t)h->p.max_batch;
    for (int l = 0; l < h->p.nlevels; ++l) {
        const LevelGeom& g = h->geom[l];
        const cuuint64_t dims[3] = {(cuuint64_t)g.pitch, (cuuint64_t)g.rows, (cuuint64_t)B};
        const cuuint64_t strides[2] = {(cuuint64_t)g.pitch, (cuuint64_t)g.plane};   // bytes, dims 1 and 2
        const cuuint32_t box[3] = {(cuuint32_t)h->fastBoxW, (cuuint32_t)g.fastBoxH, 1};   // per-level rows: most cells are ~44 rows, the coarsest levels up to ~60
        const cuuint32_t estr[3] = {1, 1, 1};
        CUresult r = encode(&h->fastMaps.m[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, h->dPyr + g.base, dims, strides, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            snprintf(g_cudaErr, sizeof(g_cudaErr), "cuTensorMapEncodeTiled(level %d) failed: CUresult %d", l, (int)r);
            return ORBFE_ERR_CUDA;
        }
    }
    return ORBFE_OK;
}

// Source boxes and tensor maps of k_pyr_resize_tma (see orbfe_internal.h); sp still holds the host copies of taps and plans.
static int encode_pyr_maps(orbfe_handle* h, const ShapePlan& sp) {
    memset(&h->pyrTma, 0, sizeof(h->pyrTma));
    if (!h->pyrUseTma) return ORBFE_OK;
    static EncodeTiledFn encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
        if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) return cuda_fail(e != cudaSuccess ? e : cudaErrorUnknown, "cuTensorMapEncodeTiled entry point");
        encode = (EncodeTiledFn)fn;
    }
    const size_t B = (size_t)h->p.max_batch;
    for (int l = 1; l < h->p.nlevels; ++l) {
        const LevelGeom& gd = h->geom[l];
        const LevelGeom& gs = h->geom[l - 1];
        PyrTmaLevel& L = h->pyrTma.lv[l];
        if (gd.planBase < 0) continue;
        L.tilesX = gd.pitch / 128;
        L.groupsY = (gd.rows + ORBFE_PYRTMA_ROWS - 1) / ORBFE_PYRTMA_ROWS;
        if (L.tilesX > 24 || L.groupsY > 96) continue;
        int spanW = 0, spanH = 0;
        for (int tx = 0; tx < L.tilesX; ++tx) {
            int lo = 1 << 30, hi = 0;
            for (int w = tx * 32; w < tx * 32 + 32; ++w) {
                const ColPlan& cp = sp.plans[gd.planBase + w];
                lo = std::min(lo, std::min(cp.base01, cp.base23));
                hi = std::max(hi, std::max(cp.base01, cp.base23) + 1);
            }
            L.colLo[tx] = lo & ~3;
            spanW = std::max(spanW, hi - L.colLo[tx] + 1);
        }
        for (int ty = 0; ty < L.groupsY; ++ty) {
            int lo = 1 << 30, hi = 0;
            for (int r = ty * ORBFE_PYRTMA_ROWS; r < std::min((ty + 1) * ORBFE_PYRTMA_ROWS, gd.rows); ++r) {
                int y = r - ORBFE_EDGE;
                y = y < 0 ? -y : y;
                if (y >= gd.h) y = 2 * gd.h - 2 - y;
                const ResizeTap& t = sp.taps[gd.tabY + y];
                lo = std::min(lo, (int)t.s0);
                hi = std::max(hi, (int)t.s1);
            }
            L.rowLo[ty] = lo;
            spanH = std::max(spanH, hi - lo + 1);
        }
        L.boxWords = (spanW + 3) & ~3;
        L.boxH = spanH;
        if (L.boxWords * 4 > 256 || L.boxH > 256) continue;
        const cuuint64_t dims[3] = {(cuuint64_t)gs.pitch, (cuuint64_t)gs.rows, (cuuint64_t)B};
        const cuuint64_t strides[2] = {(cuuint64_t)gs.pitch, (cuuint64_t)gs.plane};
        const cuuint32_t box[3] = {(cuuint32_t)(L.boxWords * 4), (cuuint32_t)L.boxH, 1};
        const cuuint32_t estr[3] = {1, 1, 1};
        CUresult r = encode(&h->pyrMaps.m[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, h->dPyr + gs.base, dims, strides, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            snprintf(g_cudaErr, sizeof(g_cudaErr), "cuTensorMapEncodeTiled(pyramid level %d) failed: CUresult %d", l, (int)r);
            return ORBFE_ERR_CUDA;
        }
        L.ok = 1;
    }
    return ORBFE_OK;
}

static void drop_graphs(orbfe_handle* h);

// Upload the geometry of a new image shape (no-op when the shape is unchanged).
static int configure_shape(orbfe_handle* h, int W, int H) {
    if (W == h->curW && H == h->curH) return ORBFE_OK;
    if (W > h->p.max_width || H > h->p.max_height) return ORBFE_ERR_TOO_LARGE;
    ShapePlan sp;
    plan_shape(h, W, H, sp);
    if (sp.status != ORBFE_OK) return sp.status;
    const size_t B = (size_t)h->p.max_batch;
    if (sp.pyrBytesPerFrame * B + (size_t)(ORBFE_BLUR_TH + 8) * sp.geom[0].pitch > h->pyrBytes || (size_t)sp.candPerFrame * B > h->candEntries ||
        (int)sp.cells.size() > h->cellCap || (int)sp.taps.size() > h->tabCap || sp.kpPerFrame > h->kpCapPerFrame ||
        (int)sp.plans.size() > h->planCap)
        return ORBFE_ERR_TOO_LARGE;
    for (int l = 0; l < h->p.nlevels; ++l) {
        h->geom[l] = sp.geom[l];
        h->geom[l].base = sp.geom[l].base * (long long)B;
    }
    h->nCellsTotal = (int)sp.cells.size();
    h->maxCellBytes = sp.maxCellBytes;
    h->fastBoxW = sp.boxW;
    h->fastBoxH = sp.boxH;
    h->fastMaxCw = sp.maxCw;
    {
        int st = encode_fast_maps(h);
        if (st == ORBFE_OK) st = encode_pyr_maps(h, sp);
        if (st != ORBFE_OK) return st;
    }
    h->candPerFrame = sp.candPerFrame;
    h->kpPerFrame = sp.kpPerFrame;
    h->blurTiles = sp.blurTiles;
    CK(cudaStreamSynchronize(h->stream));   // previous work may still read the old tables
    drop_graphs(h);                         // captured launches carry the old geometry and tensor maps
    CK(cudaMemcpyAsync(h->dGeom, h->geom, sizeof(LevelGeom) * h->p.nlevels, cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->dCells, sp.cells.data(), sizeof(CellDesc) * sp.cells.size(), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->dTaps, sp.taps.data(), sizeof(ResizeTap) * sp.taps.size(), cudaMemcpyHostToDevice, h->stream));
    if (!sp.plans.empty())
        CK(cudaMemcpyAsync(h->dPlans, sp.plans.data(), sizeof(ColPlan) * sp.plans.size(), cudaMemcpyHostToDevice, h->stream));
    CK(cudaMemcpyAsync(h->dPyrTma, &h->pyrTma, sizeof(PyrTmaPlan), cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));   // sp's vectors die at return
    h->curW = W;
    h->curH = H;
    h->lastBatch = 0;
    return ORBFE_OK;
}

extern "C" int orbfe_create(const orbfe_params* params, orbfe_handle** out) {
    if (!params || !out) return ORBFE_ERR_BAD_ARGUMENT;
    *out = nullptr;
    const orbfe_params& p = *params;
    if (p.nfeatures < 1 || p.nlevels < 1 || p.nlevels > ORBFE_MAX_LEVELS || !(p.scale_factor > 1.0f) || p.ini_th_fast < 1 ||
        p.min_th_fast < 1 || p.ini_th_fast > 254 || p.min_th_fast > 254 || p.max_width < 1 || p.max_height < 1 || p.max_batch < 1)
        return ORBFE_ERR_BAD_ARGUMENT;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) { cudaGetLastError(); return ORBFE_ERR_NO_DEVICE; }
    if (p.device < 0 || p.device >= ndev) return ORBFE_ERR_BAD_ARGUMENT;
    orbfe_handle* h = new orbfe_handle();
    memset(h, 0, sizeof(*h));
    h->p = p;
    h->device = p.device;
    build_tables(h);
    ShapePlan sp;
    plan_shape(h, p.max_width, p.max_height, sp);
    if (sp.status != ORBFE_OK) { delete h; return sp.status; }
    const size_t B = (size_t)p.max_batch;
    // a little slack so that smaller shapes whose rounding goes the other way still fit
    // slack: smaller shapes whose rounding goes the other way, and k_blur's unconditional reads of up to ORBFE_BLUR_TH rows past a
    // short last tile of the last level plane
    h->pyrBytes = sp.pyrBytesPerFrame * B + 4096 + (size_t)(ORBFE_BLUR_TH + 8) * sp.geom[0].pitch;
    h->candEntries = ((size_t)sp.candPerFrame + 256) * B;
    h->cellCap = (int)sp.cells.size() + 64;
    h->tabCap = (int)sp.taps.size() + 64;
    h->planCap = (int)sp.plans.size() + 32 * p.nlevels + 64;
    h->kpCapPerFrame = sp.kpPerFrame + 16 * p.nlevels;
    h->inBytes = (size_t)p.max_width * p.max_height * B;
    h->inFormat = ORBFE_FORMAT_GRAY8;
    h->inBpp = 1;
#define CKC(call)                                                              \
    do {                                                                       \
        cudaError_t e_ = (call);                                               \
        if (e_ != cudaSuccess) { cuda_fail(e_, #call); free_handle(h); return ORBFE_ERR_CUDA; } \
    } while (0)
    CKC(cudaSetDevice(h->device));
    CKC(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    CKC(cudaStreamCreateWithFlags(&h->sIn, cudaStreamNonBlocking));
    CKC(cudaStreamCreateWithFlags(&h->sOut, cudaStreamNonBlocking));
    h->sComp[0] = h->stream;
    h->cur = h->stream;
    for (int i = 1; i < ORBFE_MAX_COMPUTE_STREAMS; ++i) CKC(cudaStreamCreateWithFlags(&h->sComp[i], cudaStreamNonBlocking));
    CKC(cudaEventCreateWithFlags(&h->evFork, cudaEventDisableTiming));
    CKC(cudaEventCreateWithFlags(&h->evOrder, cudaEventDisableTiming));
    for (int i = 0; i < ORBFE_MAX_COMPUTE_STREAMS; ++i) CKC(cudaEventCreateWithFlags(&h->evJoin[i], cudaEventDisableTiming));
    for (int i = 0; i < ORBFE_MAX_COMPUTE_STREAMS; ++i) {
        CKC(cudaStreamCreateWithFlags(&h->sSide[i], cudaStreamNonBlocking));
        CKC(cudaEventCreateWithFlags(&h->evPyr[i], cudaEventDisableTiming));
        CKC(cudaEventCreateWithFlags(&h->evBlur[i], cudaEventDisableTiming));
    }
    CKC(cudaDeviceGetAttribute(&h->smCount, cudaDevAttrMultiProcessorCount, h->device));
    {   // device-resident pipeline tuning (see run_device_pipeline); the environment overrides are for sweeps
        const char* e1 = getenv("ORBFE_DEV_CHUNK");
        const char* e2 = getenv("ORBFE_DEV_STREAMS");
        const char* e3 = getenv("ORBFE_GRAPHS");
        h->devChunk = e1 && atoi(e1) > 0 ? atoi(e1) : ORBFE_DEV_CHUNK;
        h->devStreams = e2 && atoi(e2) > 0 ? std::min(atoi(e2), ORBFE_MAX_COMPUTE_STREAMS) : ORBFE_DEV_STREAMS;
        h->useGraphs = e3 ? atoi(e3) != 0 : 1;
        const char* e4 = getenv("ORBFE_BLUR_OVERLAP");
        const char* e5 = getenv("ORBFE_BLUR_PERSIST");
        const char* e7 = getenv("ORBFE_PYR_TMA");
        h->pyrUseTma = e7 ? atoi(e7) != 0 : ORBFE_PYR_TMA_DEFAULT;
        const char* e6 = getenv("ORBFE_BLUR");
        h->blurBackend = e6 ? (strcmp(e6, "alu") == 0 ? ORBFE_BLUR_ALU : ORBFE_BLUR_TENSOR) : ORBFE_BLUR_DEFAULT;
        h->blurOverlap = e4 ? atoi(e4) != 0 : ORBFE_BLUR_OVERLAP;
        h->blurPersist = e5 ? std::max(0, atoi(e5)) : ORBFE_BLUR_PERSIST;
    }
    for (int c = 0; c < ORBFE_PIPE_MAX_CHUNKS; ++c)
        for (int k = 0; k < 3; ++k) CKC(cudaEventCreateWithFlags(&h->evPipe[c][k], cudaEventDisableTiming));
    CKC(cudaMalloc(&h->dIn, h->inBytes));
    CKC(cudaMalloc(&h->dPyr, h->pyrBytes));
    CKC(cudaMalloc(&h->dBlur, h->pyrBytes));
    CKC(cudaMalloc(&h->dCand, h->candEntries * 4));
    CKC(cudaMalloc(&h->dNodeOf, h->candEntries * 4));
    CKC(cudaMalloc(&h->dCandCount, sizeof(int) * B * ORBFE_MAX_LEVELS));
    CKC(cudaMalloc(&h->dLvlKp, sizeof(uint32_t) * B * h->kpCapPerFrame));
    CKC(cudaMalloc(&h->dLvlCount, sizeof(int) * B * ORBFE_MAX_LEVELS));
    CKC(cudaMalloc(&h->dGeom, sizeof(LevelGeom) * ORBFE_MAX_LEVELS));
    CKC(cudaMalloc(&h->dCells, sizeof(CellDesc) * h->cellCap));
    CKC(cudaMalloc(&h->dTaps, sizeof(ResizeTap) * h->tabCap));
    CKC(cudaMalloc(&h->dPlans, sizeof(ColPlan) * h->planCap));
    CKC(cudaMalloc(&h->dPyrTma, sizeof(PyrTmaPlan)));
    {   // one result arena: counts | mono | keypoints | descriptors
        h->resKpsOff = (2 * sizeof(int) * B + 255) & ~(size_t)255;
        h->resDescOff = (h->resKpsOff + sizeof(orbfe_keypoint) * B * h->kpCapPerFrame + 255) & ~(size_t)255;
        h->resBytes = h->resDescOff + (size_t)32 * B * h->kpCapPerFrame;
        CKC(cudaMalloc(&h->dRes, h->resBytes));
        CKC(cudaMemset(h->dRes, 0, h->resBytes));
        h->dN = reinterpret_cast<int*>(h->dRes);
        h->dMono = h->dN + B;
        h->dKps = reinterpret_cast<orbfe_keypoint*>(h->dRes + h->resKpsOff);
        h->dDesc = h->dRes + h->resDescOff;
        if (p.max_batch <= ORBFE_SMALL_BATCH) CKC(cudaMallocHost(&h->hRes, h->resBytes));
    }
    CKC(cudaMalloc(&h->dSlot, sizeof(uint32_t) * B * h->kpCapPerFrame));
    CKC(cudaMallocHost(&h->hN, sizeof(int) * B));
    CKC(cudaMallocHost(&h->hMono, sizeof(int) * B));
    CKC(cudaMemset(h->dPyr, 0, h->pyrBytes));
    CKC(cudaMemset(h->dBlur, 0, h->pyrBytes));
#undef CKC
    *out = h;
    return ORBFE_OK;
}

extern "C" void orbfe_destroy(orbfe_handle* h) { free_handle(h); }

extern "C" int orbfe_get_tables(const orbfe_handle* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                                int* features_per_level) {
    if (!h) return ORBFE_ERR_BAD_ARGUMENT;
    for (int l = 0; l < h->p.nlevels; ++l) {
        if (scale) scale[l] = h->scale[l];
        if (inv_scale) inv_scale[l] = h->invScale[l];
        if (sigma2) sigma2[l] = h->sigma2[l];
        if (inv_sigma2) inv_sigma2[l] = h->invSigma2[l];
        if (features_per_level) features_per_level[l] = h->featPerLevel[l];
    }
    return ORBFE_OK;
}

extern "C" int orbfe_max_keypoints(const orbfe_handle* h) { return h ? h->kpCapPerFrame : ORBFE_ERR_BAD_ARGUMENT; }

// ---- extraction -------------------------------------------------------------------------------------
// ---- stage profiling ----------------------------------------------------------------------------------
static int prof_fold_oldest(orbfe_handle* h) {
    cudaEvent_t* ev = h->profEv[h->profHead];
    CK(cudaEventSynchronize(ev[ORBFE_NUM_STAGES]));
    for (int s = 0; s < ORBFE_NUM_STAGES; ++s) {
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, ev[s], ev[s + 1]));
        h->profMs[s] += ms;
    }
    h->profCalls++;
    h->profHead = (h->profHead + 1) % ORBFE_PROF_RING;
    h->profCount--;
    return ORBFE_OK;
}

// Enqueues the five stages for frames [frame0, frame0 + batch) of the arenas on h->cur; dImages = first of those frames.
static int enqueue_extract(orbfe_handle* h, const uint8_t* dImages, int frame0, int batch, int W, int H, long long stride,
                           long long frameStride, int lap0, int lap1) {
    cudaError_t e;
    cudaEvent_t* ev = nullptr;
    if (h->profiling) {
        if (h->profCount == ORBFE_PROF_RING) { int st = prof_fold_oldest(h); if (st != ORBFE_OK) return st; }
        ev = h->profEv[(h->profHead + h->profCount) % ORBFE_PROF_RING];
        CK(cudaEventRecord(ev[0], h->cur));
    }
    // FAST's few tall cells (the coarsest levels) are a second, small launch: it goes to the side stream and runs under the main
    // launch, and (mode 2) so do the resizes that produce the levels only those cells read.
    const int nl = h->p.nlevels, sideMode = orbfe_fast_side_mode();
    int fside = -1, tallFrom = nl;
    if (sideMode && !h->profiling && !h->blurOverlap && (tallFrom = orbfe_fast_tall_from(h, batch)) < nl)
        for (int i = 0; i < ORBFE_MAX_COMPUTE_STREAMS; ++i)
            if (h->sComp[i] == h->cur) fside = i;
    const int pyrOnCur = (fside >= 0 && sideMode >= 2) ? tallFrom : nl;
    if ((e = orbfe_launch_pyramid_levels(h, dImages, frame0, batch, stride, frameStride, 0, pyrOnCur, h->cur)) != cudaSuccess) return cuda_fail(e, "pyramid");
    if (ev) CK(cudaEventRecord(ev[1], h->cur));
    if ((e = orbfe_clear_candidates(h, frame0, batch)) != cudaSuccess) return cuda_fail(e, "fast");
    if (fside >= 0) {
        CK(cudaEventRecord(h->evPyr[fside], h->cur));
        CK(cudaStreamWaitEvent(h->sSide[fside], h->evPyr[fside], 0));
        if ((e = orbfe_launch_pyramid_levels(h, dImages, frame0, batch, stride, frameStride, pyrOnCur, nl, h->sSide[fside])) != cudaSuccess) return cuda_fail(e, "pyramid");
    }
    // The blur needs the pyramid and nothing else, FAST + quadtree + assemble do not need the blur: it goes to a side stream
    // (forked from and joined back into h->cur, so a capture sees one graph) as a few RESIDENT CTAs per SM that walk the
    // tiles - FAST is latency-bound at two thirds of the issue slots and leaves registers free, the blur is pure ALU work.
    int side = -1;
    if (!h->profiling && h->blurOverlap)
        for (int i = 0; i < ORBFE_MAX_COMPUTE_STREAMS; ++i)
            if (h->sComp[i] == h->cur) side = i;
    if (side >= 0) {
        CK(cudaEventRecord(h->evPyr[side], h->cur));
        CK(cudaStreamWaitEvent(h->sSide[side], h->evPyr[side], 0));
        if ((e = orbfe_launch_blur(h, frame0, batch, h->sSide[side], h->blurPersist)) != cudaSuccess) return cuda_fail(e, "blur");
        CK(cudaEventRecord(h->evBlur[side], h->sSide[side]));
    }
    if ((e = orbfe_launch_fast(h, frame0, batch, fside >= 0 ? h->sSide[fside] : nullptr)) != cudaSuccess) return cuda_fail(e, "fast");
    if (fside >= 0) {
        CK(cudaEventRecord(h->evBlur[fside], h->sSide[fside]));
        CK(cudaStreamWaitEvent(h->cur, h->evBlur[fside], 0));
    }
    if (ev) CK(cudaEventRecord(ev[2], h->cur));
    if ((e = orbfe_launch_octree(h, frame0, batch)) != cudaSuccess) return cuda_fail(e, "octree");
    if (ev) CK(cudaEventRecord(ev[3], h->cur));
    if ((e = orbfe_launch_assemble(h, frame0, batch, lap0, lap1)) != cudaSuccess) return cuda_fail(e, "assemble");
    if (ev) CK(cudaEventRecord(ev[4], h->cur));
    if (side >= 0) CK(cudaStreamWaitEvent(h->cur, h->evBlur[side], 0));
    else if ((e = orbfe_launch_blur(h, frame0, batch, h->cur, 0)) != cudaSuccess) return cuda_fail(e, "blur");
    if ((e = orbfe_launch_describe(h, frame0, batch)) != cudaSuccess) return cuda_fail(e, "describe");
    if (ev) { CK(cudaEventRecord(ev[5], h->cur)); h->profCount++; }
    h->lastBatch = std::max(h->lastBatch, frame0 + batch);
    return ORBFE_OK;
}

// The device-resident pipeline.  A batch is cut into chunks of at most h->devChunk frames; every chunk runs all five stages
// on one of h->devStreams compute streams (round-robin), so the narrow launches of one chunk (small pyramid levels, one
// quadtree CTA per level) run under the wide launches of its neighbours.  The side streams fork from and join back into
// the handle's public stream: callers see one ordered stream.  Measured on B200 (profiles/r02a_pipeline_sweep.txt, 512
// frames of 752x480): chunks small enough to keep a chunk's pyramid in the 126 MB L2 (16-64 frames) LOSE to whole-batch
// launches (150-157 k vs 158 k frames/s) - the stages are instruction-bound, not DRAM-bound, and short launches pay their
// tails - so the default chunk is 512 frames and chunking only matters for longer batches.  With stage profiling on,
// everything stays on the public stream (events between stages need a single timeline).
static int run_device_pipeline(orbfe_handle* h, const uint8_t* dImages, int batch, int W, int H, long long stride,
                               long long frameStride, int lap0, int lap1) {
    const int want = std::max(1, h->devChunk);
    const int nchunks = (batch + want - 1) / want;
    const int chunk = (batch + nchunks - 1) / nchunks;     // even chunks
    const int ns = h->profiling ? 1 : std::max(1, std::min(h->devStreams, nchunks));
    if (ns > 1) {
        CK(cudaEventRecord(h->evFork, h->stream));
        for (int i = 1; i < ns; ++i) CK(cudaStreamWaitEvent(h->sComp[i], h->evFork, 0));
    }
    int st = ORBFE_OK;
    for (int c = 0, f0 = 0; f0 < batch && st == ORBFE_OK; ++c, f0 += chunk) {
        h->cur = h->sComp[c % ns];
        st = enqueue_extract(h, dImages + (long long)f0 * frameStride, f0, std::min(chunk, batch - f0), W, H, stride, frameStride, lap0, lap1);
    }
    h->cur = h->stream;
    if (ns > 1)     // join even after an error: a capture must see every forked stream come back
        for (int i = 1; i < ns; ++i) {
            CK(cudaEventRecord(h->evJoin[i], h->sComp[i]));
            CK(cudaStreamWaitEvent(h->stream, h->evJoin[i], 0));
        }
    return st;
}

static void drop_graphs(orbfe_handle* h) {
    for (int i = 0; i < ORBFE_GRAPH_SLOTS; ++i) {
        if (h->graphs[i].exec) cudaGraphExecDestroy(h->graphs[i].exec);
        memset(&h->graphs[i], 0, sizeof(h->graphs[i]));
    }
}

// run_device_pipeline through a CUDA graph.  The first call with a given argument set runs eagerly (it also sets the
// per-device kernel attributes), the second one is captured (all chunks, all compute streams) and instantiated, later ones
// are a single cudaGraphLaunch on the public stream.  Keyed by the device image pointer and every argument that reaches a
// kernel; configure_shape drops the cache when the image shape changes.
static int device_pipeline(orbfe_handle* h, const uint8_t* dImages, int batch, int W, int H, long long stride,
                           long long frameStride, int lap0, int lap1) {
    h->lastBatch = 0;
    if (!h->useGraphs || h->profiling) return run_device_pipeline(h, dImages, batch, W, H, stride, frameStride, lap0, lap1);
    OrbfeGraphSlot* slot = nullptr;
    OrbfeGraphSlot* lru = &h->graphs[0];
    for (int i = 0; i < ORBFE_GRAPH_SLOTS; ++i) {
        OrbfeGraphSlot& g = h->graphs[i];
        if (g.seen && g.images == dImages && g.batch == batch && g.W == W && g.H == H && g.lap0 == lap0 && g.lap1 == lap1 &&
            g.format == h->inFormat && g.stride == stride && g.frameStride == frameStride) { slot = &g; break; }
        if (g.lastUse < lru->lastUse) lru = &g;
    }
    ++h->graphClock;
    if (slot && slot->exec) {
        slot->lastUse = h->graphClock;
        CK(cudaGraphLaunch(slot->exec, h->stream));
        h->launches += slot->kernels;
        h->graphLaunches++;
        h->lastBatch = batch;
        return ORBFE_OK;
    }
    if (!slot) {
        if (lru->exec) cudaGraphExecDestroy(lru->exec);
        memset(lru, 0, sizeof(*lru));
        lru->images = dImages; lru->batch = batch; lru->W = W; lru->H = H; lru->lap0 = lap0; lru->lap1 = lap1;
        lru->format = h->inFormat; lru->stride = stride; lru->frameStride = frameStride; lru->seen = 1;
        lru->lastUse = h->graphClock;
        return run_device_pipeline(h, dImages, batch, W, H, stride, frameStride, lap0, lap1);
    }
    slot->lastUse = h->graphClock;
    if (slot->seen < 0) return run_device_pipeline(h, dImages, batch, W, H, stride, frameStride, lap0, lap1);   // capture failed before
    const long long l0 = h->launches;
    cudaGraph_t graph = nullptr;
    cudaError_t e = cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeRelaxed);
    int st = ORBFE_OK;
    if (e == cudaSuccess) {
        st = run_device_pipeline(h, dImages, batch, W, H, stride, frameStride, lap0, lap1);
        e = cudaStreamEndCapture(h->stream, &graph);
    }
    if (e == cudaSuccess && st == ORBFE_OK) e = cudaGraphInstantiate(&slot->exec, graph, 0);
    if (graph) cudaGraphDestroy(graph);
    if (e != cudaSuccess || st != ORBFE_OK) {          // nothing ran: fall back to the eager pipeline for this argument set
        cudaGetLastError();
        slot->exec = nullptr;
        slot->seen = -1;
        h->launches = l0;
        if (st != ORBFE_OK) return st;
        return run_device_pipeline(h, dImages, batch, W, H, stride, frameStride, lap0, lap1);
    }
    slot->kernels = h->launches - l0;
    CK(cudaGraphLaunch(slot->exec, h->stream));
    h->graphLaunches++;
    h->lastBatch = batch;
    return ORBFE_OK;
}

extern "C" int orbfe_set_profiling(orbfe_handle* h, int enabled) {
    if (!h) return ORBFE_ERR_BAD_ARGUMENT;
    CK(cudaSetDevice(h->device));
    if (enabled && !h->profEv) {
        h->profEv = new cudaEvent_t[ORBFE_PROF_RING][ORBFE_NUM_STAGES + 1];
        for (int i = 0; i < ORBFE_PROF_RING; ++i)
            for (int s = 0; s <= ORBFE_NUM_STAGES; ++s) CK(cudaEventCreate(&h->profEv[i][s]));
    }
    h->profiling = enabled ? 1 : 0;
    return ORBFE_OK;
}

extern "C" int orbfe_get_stage_times(orbfe_handle* h, double* ms, long long* calls) {
    if (!h) return ORBFE_ERR_BAD_ARGUMENT;
    CK(cudaSetDevice(h->device));
    while (h->profCount > 0) { int st = prof_fold_oldest(h); if (st != ORBFE_OK) return st; }
    for (int s = 0; s < ORBFE_NUM_STAGES; ++s) { if (ms) ms[s] = h->profMs[s]; h->profMs[s] = 0; }
    if (calls) *calls = h->profCalls;
    h->profCalls = 0;
    return ORBFE_OK;
}

extern "C" const char* orbfe_stage_name(int stage) {
    static const char* names[ORBFE_NUM_STAGES] = {"pyramid", "fast", "octree", "assemble", "describe"};
    return stage >= 0 && stage < ORBFE_NUM_STAGES ? names[stage] : "?";
}

// Tracking::GrabImage* (src/Tracking.cc:1233-1258, 1325-1338) converts colour input with cv::cvtColor before the
// extractor sees it; here the conversion is fused into the level-0 kernel (k_pyramid.cu).  The staging arena of
// the host-buffer entry points grows to bytes-per-pixel times its grey size.
extern "C" int orbfe_set_input_format(orbfe_handle* h, int format) {
    if (!h || format < ORBFE_FORMAT_GRAY8 || format > ORBFE_FORMAT_BGRA8) return ORBFE_ERR_BAD_ARGUMENT;
    const int bpp = format == ORBFE_FORMAT_GRAY8 ? 1 : (format <= ORBFE_FORMAT_BGR8 ? 3 : 4);
    CK(cudaSetDevice(h->device));
    if (bpp > h->inBpp) {
        CK(cudaDeviceSynchronize());
        const size_t grey = (size_t)h->p.max_width * h->p.max_height * h->p.max_batch;
        cudaFree(h->dIn);
        h->dIn = nullptr;
        CK(cudaMalloc(&h->dIn, grey * bpp));
        h->inBytes = grey * bpp;
        h->inBpp = bpp;
    }
    h->inFormat = format;
    return ORBFE_OK;
}

static inline int format_bpp(int format) { return format == ORBFE_FORMAT_GRAY8 ? 1 : (format <= ORBFE_FORMAT_BGR8 ? 3 : 4); }

extern "C" int orbfe_extract_batch_device(orbfe_handle* h, const uint8_t* d_images, int batch, int width, int height,
                                          ptrdiff_t stride, ptrdiff_t frame_stride, int lap0, int lap1) {
    if (!h) return ORBFE_ERR_BAD_ARGUMENT;
    if (!d_images || width <= 0 || height <= 0 || batch <= 0) return ORBFE_ERR_EMPTY_IMAGE;
    if (stride < (ptrdiff_t)width * format_bpp(h->inFormat) || batch > h->p.max_batch) return batch > h->p.max_batch ? ORBFE_ERR_TOO_LARGE : ORBFE_ERR_BAD_ARGUMENT;
    if (batch > 1 && frame_stride < stride * (ptrdiff_t)height) return ORBFE_ERR_BAD_ARGUMENT;
    CK(cudaSetDevice(h->device));
    int st = configure_shape(h, width, height);
    if (st != ORBFE_OK) return st;
    return device_pipeline(h, d_images, batch, width, height, stride, frame_stride, lap0, lap1);
}

extern "C" int orbfe_fetch_results(orbfe_handle* h, int batch, orbfe_keypoint* kps, uint8_t* desc, int cap, int* n_out,
                                   int* mono_out) {
    if (!h || !kps || !desc || !n_out || batch <= 0) return ORBFE_ERR_BAD_ARGUMENT;
    if (h->lastBatch < batch) return ORBFE_ERR_NOT_READY;
    CK(cudaSetDevice(h->device));
    CK(cudaMemcpyAsync(h->hN, h->dN, sizeof(int) * batch, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(h->hMono, h->dMono, sizeof(int) * batch, cudaMemcpyDeviceToHost, h->stream));
    const int kpf = h->kpPerFrame;
    if (cap >= kpf) {
        // common case: one strided copy per array, rows of kpf entries into rows of cap entries
        CK(cudaMemcpy2DAsync(kps, sizeof(orbfe_keypoint) * (size_t)cap, h->dKps, sizeof(orbfe_keypoint) * (size_t)kpf,
                             sizeof(orbfe_keypoint) * (size_t)kpf, batch, cudaMemcpyDeviceToHost, h->stream));
        CK(cudaMemcpy2DAsync(desc, 32 * (size_t)cap, h->dDesc, 32 * (size_t)kpf, 32 * (size_t)kpf, batch, cudaMemcpyDeviceToHost,
                             h->stream));
        CK(cudaStreamSynchronize(h->stream));
        for (int b = 0; b < batch; ++b) { n_out[b] = h->hN[b]; if (mono_out) mono_out[b] = h->hMono[b]; }
        return ORBFE_OK;
    }
    CK(cudaStreamSynchronize(h->stream));
    for (int b = 0; b < batch; ++b) {
        n_out[b] = h->hN[b];
        if (mono_out) mono_out[b] = h->hMono[b];
        if (h->hN[b] > cap) return ORBFE_ERR_CAPACITY;
    }
    for (int b = 0; b < batch; ++b) {
        const int n = h->hN[b];
        if (n == 0) continue;
        CK(cudaMemcpyAsync(kps + (size_t)b * cap, h->dKps + (size_t)b * kpf, sizeof(orbfe_keypoint) * n, cudaMemcpyDeviceToHost, h->stream));
        CK(cudaMemcpyAsync(desc + (size_t)b * cap * 32, h->dDesc + (size_t)b * kpf * 32, 32 * (size_t)n, cudaMemcpyDeviceToHost, h->stream));
    }
    CK(cudaStreamSynchronize(h->stream));
    return ORBFE_OK;
}

// The drop-in call sizes (one image, or the two eyes of a stereo frame): one H2D copy, one graph launch, ONE D2H copy of
// the whole result arena into its pinned mirror, one synchronise; the rows go to the caller's buffers by memcpy.
static int extract_small(orbfe_handle* h, const uint8_t* images, int batch, int width, int height, ptrdiff_t stride,
                         ptrdiff_t frame_stride, int lap0, int lap1, orbfe_keypoint* kps, uint8_t* desc, int cap, int* n_out,
                         int* mono_out) {
    const ptrdiff_t rowBytes = (ptrdiff_t)width * format_bpp(h->inFormat);
    const size_t fbytes = (size_t)rowBytes * height;
    if (stride == rowBytes && (batch == 1 || frame_stride == (ptrdiff_t)fbytes)) {
        CK(cudaMemcpyAsync(h->dIn, images, fbytes * batch, cudaMemcpyHostToDevice, h->stream));
    } else {
        for (int b = 0; b < batch; ++b)
            CK(cudaMemcpy2DAsync(h->dIn + b * fbytes, rowBytes, images + (size_t)b * frame_stride, stride, rowBytes, height,
                                 cudaMemcpyHostToDevice, h->stream));
    }
    int st = device_pipeline(h, h->dIn, batch, width, height, rowBytes, (long long)fbytes, lap0, lap1);
    if (st != ORBFE_OK) { cudaStreamSynchronize(h->stream); return st; }
    const int kpf = h->kpPerFrame, B = h->p.max_batch;
    const size_t kpsBytes = sizeof(orbfe_keypoint) * (size_t)batch * kpf, descBytes = (size_t)32 * batch * kpf;
    if (batch == B) {
        CK(cudaMemcpyAsync(h->hRes, h->dRes, h->resDescOff + descBytes, cudaMemcpyDeviceToHost, h->stream));
    } else {
        CK(cudaMemcpyAsync(h->hRes, h->dRes, 2 * sizeof(int) * B, cudaMemcpyDeviceToHost, h->stream));
        CK(cudaMemcpyAsync(h->hRes + h->resKpsOff, h->dRes + h->resKpsOff, kpsBytes, cudaMemcpyDeviceToHost, h->stream));
        CK(cudaMemcpyAsync(h->hRes + h->resDescOff, h->dRes + h->resDescOff, descBytes, cudaMemcpyDeviceToHost, h->stream));
    }
    CK(cudaStreamSynchronize(h->stream));
    const int* hn = reinterpret_cast<const int*>(h->hRes);
    const int* hm = hn + B;
    for (int b = 0; b < batch; ++b) {
        n_out[b] = hn[b];
        if (mono_out) mono_out[b] = hm[b];
        if (hn[b] > cap) return ORBFE_ERR_CAPACITY;
    }
    for (int b = 0; b < batch; ++b) {
        memcpy(kps + (size_t)b * cap, h->hRes + h->resKpsOff + sizeof(orbfe_keypoint) * (size_t)b * kpf, sizeof(orbfe_keypoint) * (size_t)hn[b]);
        memcpy(desc + (size_t)b * cap * 32, h->hRes + h->resDescOff + (size_t)32 * b * kpf, (size_t)32 * hn[b]);
    }
    return ORBFE_OK;
}

extern "C" int orbfe_extract_batch(orbfe_handle* h, const uint8_t* images, int batch, int width, int height, ptrdiff_t stride,
                                   ptrdiff_t frame_stride, int lap0, int lap1, orbfe_keypoint* kps, uint8_t* desc, int cap,
                                   int* n_out, int* mono_out) {
    if (!h) return ORBFE_ERR_BAD_ARGUMENT;
    if (!images || width <= 0 || height <= 0 || batch <= 0) return ORBFE_ERR_EMPTY_IMAGE;
    const int bpp = format_bpp(h->inFormat);
    const ptrdiff_t rowBytes = (ptrdiff_t)width * bpp;
    if (stride < rowBytes) return ORBFE_ERR_BAD_ARGUMENT;
    if (batch > 1 && frame_stride < stride * (ptrdiff_t)height) return ORBFE_ERR_BAD_ARGUMENT;
    if (batch > h->p.max_batch || width > h->p.max_width || height > h->p.max_height) return ORBFE_ERR_TOO_LARGE;
    CK(cudaSetDevice(h->device));
    int st = configure_shape(h, width, height);
    if (st != ORBFE_OK) return st;
    if (!kps || !desc || !n_out) return ORBFE_ERR_BAD_ARGUMENT;
    if (h->hRes) return extract_small(h, images, batch, width, height, stride, frame_stride, lap0, lap1, kps, desc, cap, n_out, mono_out);
    // Chunked pipeline: the H2D copy of chunk k+1 (copy-in stream) overlaps the kernels of chunk k (main
    // stream) and the D2H copy of chunk k-1's results (copy-out stream).  Frames are staged densely
    // (row stride = width) in the input arena.
    static const int pipeChunkEnv = [] { const char* e = getenv("ORBFE_PIPE_CHUNK"); int v = e ? atoi(e) : 0; return v > 0 ? v : 0; }();
    // chunk = what overlaps: sized in BYTES (64 frames of 752x480 = 23 MB, measured best on B200, tools/sweep_chunk.sh), so larger
    // frames get proportionally fewer per chunk (11 at 1920x1080) and a 32-frame HD batch still pipelines its copies
    int pipeChunk = pipeChunkEnv ? pipeChunkEnv
                                 : (int)std::max<size_t>(2, std::min<size_t>(ORBFE_PIPE_CHUNK, (size_t)ORBFE_PIPE_CHUNK * 752 * 480 / ((size_t)rowBytes * height)));
    while ((batch + pipeChunk - 1) / pipeChunk + 2 > ORBFE_PIPE_MAX_CHUNKS) pipeChunk *= 2;    // very long batches: larger chunks, never a refusal
    // chunk boundaries: a short first chunk (its copy-in is the only one nothing overlaps), then full chunks
    int bounds[ORBFE_PIPE_MAX_CHUNKS + 1];
    int nchunks = 0;
    bounds[0] = 0;
    // ORBFE_PIPE_SCHEDULE="16,64,64,..." (debug / sweeps): explicit chunk sizes, used when they add up to the batch
    static const char* schedEnv = getenv("ORBFE_PIPE_SCHEDULE");
    if (schedEnv && *schedEnv) {
        int sum = 0, n = 0, tmp[ORBFE_PIPE_MAX_CHUNKS];
        for (const char* q = schedEnv; *q && n < ORBFE_PIPE_MAX_CHUNKS;) {
            const int v = atoi(q);
            if (v <= 0) break;
            tmp[n++] = v; sum += v;
            while (*q && *q != ',') ++q;
            if (*q == ',') ++q;
        }
        if (sum == batch)
            for (int i = 0; i < n; ++i) { bounds[nchunks + 1] = bounds[nchunks] + tmp[i]; ++nchunks; }
    }
    if (nchunks > 0) {
    } else if (batch <= pipeChunk + pipeChunk / 2) {
        bounds[++nchunks] = batch;
    } else {
        int f = std::min(pipeChunk / 4 > 0 ? pipeChunk / 4 : 1, batch);
        bounds[++nchunks] = f;
        // ... and optionally a tapering tail (half, then a quarter chunk; ORBFE_PIPE_TAPER=1).  Measured twice, no effect either
        // time (124.6 k vs 124.1 k frames/s, round 2): the timeline (ORBFE_PIPE_TRACE=1, tools/probe/pipe_trace.py) shows the
        // copy-in stream busy back to back for 3.48 of the 4.13 ms and every chunk's kernels finishing ~0.7 ms after its bytes
        // arrive - a chunk is a chain of fourteen dependent launches with a latency floor of 0.3-0.5 ms however small it is -
        // so what follows the last copy-in is that lag, not the size of the last chunk.
        static const int taper = [] { const char* e = getenv("ORBFE_PIPE_TAPER"); return e ? atoi(e) : 0; }();
        const int t2 = taper && pipeChunk >= 8 ? pipeChunk / 2 : 0, t4 = taper && pipeChunk >= 8 ? pipeChunk / 4 : 0;
        int body = batch - f - t2 - t4;                       // frames between the short first chunk and the tail
        int tail2 = t2, tail4 = t4;
        if (body < pipeChunk / 2) { body = batch - f; tail2 = tail4 = 0; }     // too short to taper
        int nbody = (body + pipeChunk - 1) / pipeChunk;
        if (nchunks + nbody + 2 > ORBFE_PIPE_MAX_CHUNKS) nbody = ORBFE_PIPE_MAX_CHUNKS - nchunks - 2;
        for (int k = 0; k < nbody; ++k) {                     // even body chunks
            f += body / nbody + (k < body % nbody ? 1 : 0);
            bounds[++nchunks] = f;
        }
        if (tail2) { f += tail2; bounds[++nchunks] = f; }
        if (tail4) { f += tail4; bounds[++nchunks] = f; }
    }
    const int kpf = h->kpPerFrame;
    const size_t fbytes = (size_t)rowBytes * height;
    h->lastBatch = 0;
    static const int pipeStreams = [] { const char* e = getenv("ORBFE_PIPE_STREAMS"); int v = e ? atoi(e) : 0; return v > 0 ? std::min(v, ORBFE_MAX_COMPUTE_STREAMS) : ORBFE_PIPE_STREAMS; }();
    // an error in the middle must not return while copies into the caller's buffers are still in flight
    auto fail = [&](int code) {
        h->cur = h->stream;
        cudaStreamSynchronize(h->sIn);
        for (int i = 0; i < pipeStreams; ++i) cudaStreamSynchronize(h->sComp[i]);
        cudaStreamSynchronize(h->sOut);
        return code;
    };
#define CKP(call)                                                   \
    do {                                                            \
        cudaError_t e_ = (call);                                    \
        if (e_ != cudaSuccess) return fail(cuda_fail(e_, #call));   \
    } while (0)
    // ORBFE_PIPE_TRACE=1 (debug): timed events after every copy-in, chunk of kernels and copy-out, printed to stderr
    static const int trace = [] { const char* e = getenv("ORBFE_PIPE_TRACE"); return e ? atoi(e) : 0; }();
    std::vector<cudaEvent_t> tev;
    auto mark = [&](cudaStream_t s_) { if (trace) { cudaEvent_t e_; cudaEventCreate(&e_); cudaEventRecord(e_, s_); tev.push_back(e_); } };
    mark(h->stream);
    CKP(cudaEventRecord(h->evPipe[0][0], h->stream));          // copy-in / the second compute stream must not overtake earlier work
    CKP(cudaStreamWaitEvent(h->sIn, h->evPipe[0][0], 0));
    for (int i = 1; i < pipeStreams; ++i) CKP(cudaStreamWaitEvent(h->sComp[i], h->evPipe[0][0], 0));
    // every copy-in is enqueued before the first kernel: the copy stream then runs back to back whatever the host needs to
    // enqueue a chunk's fourteen launches (the short first chunk is copied faster than its kernels are enqueued)
    for (int c = 0; c < nchunks; ++c) {
        const int f0 = bounds[c], nb = bounds[c + 1] - f0;
        if (stride == rowBytes && frame_stride == (ptrdiff_t)fbytes) {
            CKP(cudaMemcpyAsync(h->dIn + f0 * fbytes, images + f0 * fbytes, fbytes * nb, cudaMemcpyHostToDevice, h->sIn));
        } else {
            for (int b = f0; b < f0 + nb; ++b)
                CKP(cudaMemcpy2DAsync(h->dIn + b * fbytes, rowBytes, images + (size_t)b * frame_stride, stride, rowBytes, height,
                                      cudaMemcpyHostToDevice, h->sIn));
        }
        CKP(cudaEventRecord(h->evPipe[c][1], h->sIn));
        mark(h->sIn);
    }
    for (int c = 0; c < nchunks; ++c) {
        const int f0 = bounds[c], nb = bounds[c + 1] - f0;
        // chunks go round the compute streams, so the latency-bound stages of one chunk (small pyramid levels, quadtree)
        // run under the wide stages of its neighbours
        cudaStream_t compute = h->sComp[c % pipeStreams];
        CKP(cudaStreamWaitEvent(compute, h->evPipe[c][1], 0));
        h->cur = compute;
        st = enqueue_extract(h, h->dIn + f0 * fbytes, f0, nb, width, height, rowBytes, (long long)fbytes, lap0, lap1);
        h->cur = h->stream;
        if (st != ORBFE_OK) return fail(st);
        CKP(cudaEventRecord(h->evPipe[c][2], compute));
        mark(compute);
        CKP(cudaStreamWaitEvent(h->sOut, h->evPipe[c][2], 0));
        CKP(cudaMemcpyAsync(h->hN + f0, h->dN + f0, sizeof(int) * nb, cudaMemcpyDeviceToHost, h->sOut));
        CKP(cudaMemcpyAsync(h->hMono + f0, h->dMono + f0, sizeof(int) * nb, cudaMemcpyDeviceToHost, h->sOut));
        if (cap >= kpf) {
            CKP(cudaMemcpy2DAsync(kps + (size_t)f0 * cap, sizeof(orbfe_keypoint) * (size_t)cap, h->dKps + (size_t)f0 * kpf,
                                  sizeof(orbfe_keypoint) * (size_t)kpf, sizeof(orbfe_keypoint) * (size_t)kpf, nb,
                                  cudaMemcpyDeviceToHost, h->sOut));
            CKP(cudaMemcpy2DAsync(desc + (size_t)f0 * cap * 32, 32 * (size_t)cap, h->dDesc + (size_t)f0 * kpf * 32, 32 * (size_t)kpf,
                                  32 * (size_t)kpf, nb, cudaMemcpyDeviceToHost, h->sOut));
        }
        mark(h->sOut);
    }
    {
    // the public stream ends up ordered after everything this call enqueued (device results are valid on it)
    for (int i = 1; i < pipeStreams; ++i) {
        CKP(cudaEventRecord(h->evJoin[i], h->sComp[i]));
        CKP(cudaStreamWaitEvent(h->stream, h->evJoin[i], 0));
    }
    CKP(cudaStreamSynchronize(h->sOut));
    CKP(cudaStreamSynchronize(h->stream));
    }
    if (trace && !tev.empty()) {       // t0, copy-in x nchunks, then (kernels, copy-out) x nchunks
        fprintf(stderr, "[orbfe pipe trace] %d chunks:", nchunks);
        for (int c = 0; c < nchunks; ++c) fprintf(stderr, " %d", bounds[c + 1] - bounds[c]);
        fprintf(stderr, "\n");
        for (size_t i = 1; i < tev.size(); ++i) {
            float ms = 0.f;
            cudaEventElapsedTime(&ms, tev[0], tev[i]);
            const int c = (int)i - 1;
            if (c < nchunks) fprintf(stderr, "  copy-in %2d done %.3f ms\n", c, ms);
            else fprintf(stderr, "  %s %2d done %.3f ms\n", (c - nchunks) % 2 ? "copy-out" : "kernels ", (c - nchunks) / 2, ms);
        }
        for (cudaEvent_t e_ : tev) cudaEventDestroy(e_);
    }
#undef CKP
    if (cap >= kpf) {
        for (int b = 0; b < batch; ++b) { n_out[b] = h->hN[b]; if (mono_out) mono_out[b] = h->hMono[b]; }
        return ORBFE_OK;
    }
    return orbfe_fetch_results(h, batch, kps, desc, cap, n_out, mono_out);   // tight caller buffers: per-frame copies
}

extern "C" int orbfe_extract(orbfe_handle* h, const uint8_t* image, int width, int height, ptrdiff_t stride, int lap0, int lap1,
                             orbfe_keypoint* kps, uint8_t* desc, int cap, int* n_out) {
    int n = 0, mono = 0;
    int st = orbfe_extract_batch(h, image, 1, width, height, stride, (ptrdiff_t)stride * height, lap0, lap1, kps, desc, cap, &n, &mono);
    if (n_out) *n_out = n;
    return st < 0 ? st : mono;
}

// ---- ordering against other streams (the matchers and the vocabulary own their streams) ---------------------------
extern "C" int orbfe_stream_wait_extractor(orbfe_handle* h, void* stream) {
    if (!h) return ORBFE_ERR_BAD_ARGUMENT;
    CK(cudaSetDevice(h->device));
    CK(cudaEventRecord(h->evOrder, h->stream));
    CK(cudaStreamWaitEvent((cudaStream_t)stream, h->evOrder, 0));
    return ORBFE_OK;
}
extern "C" int orbfe_extractor_wait_stream(orbfe_handle* h, void* stream) {
    if (!h) return ORBFE_ERR_BAD_ARGUMENT;
    CK(cudaSetDevice(h->device));
    CK(cudaEventRecord(h->evOrder, (cudaStream_t)stream));
    CK(cudaStreamWaitEvent(h->stream, h->evOrder, 0));
    return ORBFE_OK;
}
extern "C" long long orbfe_graph_launch_count(const orbfe_handle* h) { return h ? h->graphLaunches : 0; }
extern "C" int orbfe_set_pipeline(orbfe_handle* h, int chunk_frames, int compute_streams, int use_graphs) {
    if (!h) return ORBFE_ERR_BAD_ARGUMENT;
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    if (chunk_frames > 0) h->devChunk = chunk_frames;
    if (compute_streams > 0) h->devStreams = std::min(compute_streams, ORBFE_MAX_COMPUTE_STREAMS);
    if (use_graphs >= 0) h->useGraphs = use_graphs != 0;
    drop_graphs(h);
    return ORBFE_OK;
}

extern "C" int orbfe_set_blur_backend(orbfe_handle* h, int backend) {
    if (!h || (backend != ORBFE_BLUR_ALU && backend != ORBFE_BLUR_TENSOR)) return ORBFE_ERR_BAD_ARGUMENT;
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    h->blurBackend = backend;
    drop_graphs(h);
    return ORBFE_OK;
}

extern "C" int orbfe_device_results(orbfe_handle* h, const orbfe_keypoint** d_kps, const uint8_t** d_desc, const int** d_n,
                                    const int** d_mono, int* cap) {
    if (!h) return ORBFE_ERR_BAD_ARGUMENT;
    if (h->lastBatch <= 0) return ORBFE_ERR_NOT_READY;
    if (d_kps) *d_kps = h->dKps;
    if (d_desc) *d_desc = h->dDesc;
    if (d_n) *d_n = h->dN;
    if (d_mono) *d_mono = h->dMono;
    if (cap) *cap = h->kpPerFrame;
    return ORBFE_OK;
}

// ---- taps ---------------------------------------------------------------------------------------------
extern "C" int orbfe_get_pyramid_level(orbfe_handle* h, int frame, int level, uint8_t* dst, ptrdiff_t dst_stride, int* w, int* hgt) {
    if (!h || level < 0 || level >= h->p.nlevels) return ORBFE_ERR_BAD_ARGUMENT;
    if (h->lastBatch <= 0) return ORBFE_ERR_NOT_READY;
    if (frame < 0 || frame >= h->lastBatch) return ORBFE_ERR_BAD_ARGUMENT;
    const LevelGeom& g = h->geom[level];
    if (w) *w = g.w;
    if (hgt) *hgt = g.h;
    if (!dst) return ORBFE_OK;
    const int bw = g.w + 2 * ORBFE_EDGE;
    if (dst_stride < bw) return ORBFE_ERR_BAD_ARGUMENT;
    CK(cudaSetDevice(h->device));
    const uint8_t* src = h->dPyr + g.base + (long long)frame * g.plane + (ORBFE_XO - ORBFE_EDGE);
    CK(cudaMemcpy2DAsync(dst, dst_stride, src, g.pitch, bw, g.rows, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return ORBFE_OK;
}

// The blurred level descriptors are sampled from (the reference's `workingMat` after GaussianBlur, ORBextractor.cc:1114-1115): ROI only.
extern "C" int orbfe_get_blurred_level(orbfe_handle* h, int frame, int level, uint8_t* dst, ptrdiff_t dst_stride) {
    if (!h || !dst || level < 0 || level >= h->p.nlevels) return ORBFE_ERR_BAD_ARGUMENT;
    if (h->lastBatch <= 0) return ORBFE_ERR_NOT_READY;
    if (frame < 0 || frame >= h->lastBatch) return ORBFE_ERR_BAD_ARGUMENT;
    const LevelGeom& g = h->geom[level];
    if (dst_stride < g.w) return ORBFE_ERR_BAD_ARGUMENT;
    CK(cudaSetDevice(h->device));
    const uint8_t* src = h->dBlur + g.base + (long long)frame * g.plane + (long long)ORBFE_EDGE * g.pitch + ORBFE_XO;
    CK(cudaMemcpy2DAsync(dst, dst_stride, src, g.pitch, g.w, g.h, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    return ORBFE_OK;
}

// All levels of one frame with one synchronise (what the shim's mvImagePyramid needs after every operator() when the host
// reads the pyramid, as the unmodified Frame::ComputeStereoMatches does).
extern "C" int orbfe_get_pyramid(orbfe_handle* h, int frame, uint8_t* const* dst, const ptrdiff_t* dst_stride) {
    if (!h || !dst || !dst_stride) return ORBFE_ERR_BAD_ARGUMENT;
    if (h->lastBatch <= 0) return ORBFE_ERR_NOT_READY;
    if (frame < 0 || frame >= h->lastBatch) return ORBFE_ERR_BAD_ARGUMENT;
    for (int l = 0; l < h->p.nlevels; ++l)
        if (!dst[l] || dst_stride[l] < h->geom[l].w + 2 * ORBFE_EDGE) return ORBFE_ERR_BAD_ARGUMENT;
    CK(cudaSetDevice(h->device));
    for (int l = 0; l < h->p.nlevels; ++l) {
        const LevelGeom& g = h->geom[l];
        const uint8_t* src = h->dPyr + g.base + (long long)frame * g.plane + (ORBFE_XO - ORBFE_EDGE);
        CK(cudaMemcpy2DAsync(dst[l], dst_stride[l], src, g.pitch, g.w + 2 * ORBFE_EDGE, g.rows, cudaMemcpyDeviceToHost, h->stream));
    }
    CK(cudaStreamSynchronize(h->stream));
    return ORBFE_OK;
}

extern "C" int orbfe_get_level_size(const orbfe_handle* h, int level, int* w, int* hgt) {
    if (!h || level < 0 || level >= h->p.nlevels || h->curW <= 0) return h && h->curW <= 0 ? ORBFE_ERR_NOT_READY : ORBFE_ERR_BAD_ARGUMENT;
    if (w) *w = h->geom[level].w;
    if (hgt) *hgt = h->geom[level].h;
    return ORBFE_OK;
}

extern "C" int orbfe_get_candidates(orbfe_handle* h, int frame, int level, int32_t* xyr, int cap) {
    if (!h || level < 0 || level >= h->p.nlevels) return ORBFE_ERR_BAD_ARGUMENT;
    if (h->lastBatch <= 0) return ORBFE_ERR_NOT_READY;
    if (frame < 0 || frame >= h->lastBatch) return ORBFE_ERR_BAD_ARGUMENT;
    CK(cudaSetDevice(h->device));
    const LevelGeom& g = h->geom[level];
    int n = 0;
    CK(cudaMemcpyAsync(&n, h->dCandCount + frame * h->p.nlevels + level, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (n <= 0 || !xyr) return n;
    std::vector<uint32_t> c((size_t)n);
    CK(cudaMemcpyAsync(c.data(), h->dCand + (long long)frame * h->candPerFrame + g.candBase, sizeof(uint32_t) * n,
                       cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    // restore the reference's push order: cell-row-major, then row-major inside the cell
    auto key = [&](uint32_t v) {
        const int x = ORBFE_CAND_X(v), y = ORBFE_CAND_Y(v);
        const long long cj = (x - 3) / g.wCell, ci = (y - 3) / g.hCell;
        return ((ci * g.nCols + cj) << 22) | ((long long)y << 11) | x;
    };
    std::sort(c.begin(), c.end(), [&](uint32_t a, uint32_t b) { return key(a) < key(b); });
    for (int i = 0; i < n && i < cap; ++i) {
        xyr[3 * i] = ORBFE_CAND_X(c[i]);
        xyr[3 * i + 1] = ORBFE_CAND_Y(c[i]);
        xyr[3 * i + 2] = ORBFE_CAND_R(c[i]);
    }
    return n;
}

extern "C" void* orbfe_stream(orbfe_handle* h) { return h ? (void*)h->stream : nullptr; }
extern "C" int orbfe_synchronize(orbfe_handle* h) {
    if (!h) return ORBFE_ERR_BAD_ARGUMENT;
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    return ORBFE_OK;
}
extern "C" long long orbfe_launch_count(const orbfe_handle* h) { return h ? h->launches : 0; }

extern "C" const char* orbfe_strerror(int status) {
    switch (status) {
        case ORBFE_OK: return "ok";
        case ORBFE_ERR_EMPTY_IMAGE: return "empty image";
        case ORBFE_ERR_BAD_ARGUMENT: return "bad argument";
        case ORBFE_ERR_TOO_LARGE: return "image, batch or feature count outside what the handle supports";
        case ORBFE_ERR_CAPACITY: return "output buffer too small";
        case ORBFE_ERR_CUDA: return "CUDA error";
        case ORBFE_ERR_NO_DEVICE: return "no CUDA device";
        case ORBFE_ERR_NOT_READY: return "no extraction result available";
        default: return status > 0 ? "ok" : "unknown error";
    }
}
extern "C" const char* orbfe_last_cuda_error(void) { return g_cudaErr; }
extern "C" const char* orbfe_version(void) { return "orbfe 0.1 sm_100a"; }
