// orbfe_internal.h — handle layout, level geometry and kernel launchers shared by the .cu files.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

#include "../../include/orbfe.h"

#define ORBFE_MAX_LEVELS 16
#define ORBFE_EDGE 19          // EDGE_THRESHOLD, ORBextractor.cc:71
#ifndef ORBFE_BLUR_TH
#define ORBFE_BLUR_TH 32       // rows per k_blur tile (tile = 128 x ORBFE_BLUR_TH, one warp)
#endif
#define ORBFE_XO 32            // byte offset of ROI column 0 inside a bordered row (keeps ROI 32B-aligned)
#define ORBFE_MIN_BORDER 16    // EDGE_THRESHOLD - 3, ORBextractor.cc:771
#define ORBFE_MAX_NODES 2048
#define ORBFE_PIPE_CHUNK 64        // frames per chunk of the pipelined host path (orbfe_extract_batch)
#define ORBFE_PIPE_MAX_CHUNKS 64
#define ORBFE_MAX_COMPUTE_STREAMS 4
#define ORBFE_GRAPH_SLOTS 4
#define ORBFE_DEV_CHUNK 512        // frames per chunk of the device-resident pipeline (profiles/r02a_pipeline_sweep.txt: L2-sized chunks lose)
#define ORBFE_DEV_STREAMS 2
#define ORBFE_SMALL_BATCH 8        // handles up to this max_batch mirror the whole result arena in pinned memory (one D2H per call)

// Packed FAST candidate: x | y << 11 | response << 22, (x, y) relative to (minBorderX, minBorderY).
#define ORBFE_CAND_X(c) ((int)((c) & 0x7FFu))
#define ORBFE_CAND_Y(c) ((int)(((c) >> 11) & 0x7FFu))
#define ORBFE_CAND_R(c) ((int)((c) >> 22))

// Geometry of one pyramid level (host-computed once per image shape, read by every kernel).
struct LevelGeom {
    int w, h;              // ROI size, ORBextractor.cc:1156-1157
    int pitch;             // bytes per row of the bordered device buffer
    int rows;              // h + 38
    long long base;        // byte offset of frame 0 of this level inside the pyramid arena
    long long plane;       // bytes per frame of this level (pitch * rows)
    // FAST grid, ORBextractor.cc:771-785
    int maxBX, maxBY, nCols, nRows, wCell, hCell;
    int cellBase, nCells;  // slice of the flattened cell table
    int fastBoxH;          // rows of this level's FAST TMA box = its tallest cell ROI (the box width is common, fastBoxW)
    // quadtree, ORBextractor.cc:541-543 and :433-444
    int nfeat, nIni;
    float hX;
    int candCap;           // worst-case candidates of this level
    long long candBase;    // offset (entries) of frame 0 / this level in the candidate arena
    int kpCap, kpBase;     // slots for selected keypoints of this level (per frame)
    float scale;           // mvScaleFactor[level]
    float size;            // (float)(int)(31 * scale)
    int tabX, tabY;        // offsets into the resize tables (entries)
    int blurTileBase, blurTilesX;   // 128 x 32 tiles of k_blur: first tile id of this level, tiles per tile row
    int planBase;          // first ColPlan of this level (pitch/4 entries), -1: no plan (generic resize kernel)
};

// One FAST cell = one CTA of k_fast (ORBextractor.cc:787-853), with everything the kernel would otherwise derive per CTA.
struct __align__(16) CellDesc {
    short level, x0, y0, cw, ch;   // ROI-relative origin (iniX, iniY) and size (maxX-iniX, maxY-iniY)
    short ox, oy;                  // j*wCell, i*hCell
    short sh;                      // (ORBFE_XO + x0) & 15: tile byte = sh + column
    short wLo, nWords;             // stage A: first tested word of a tile row, tested words per row
    short units;                   // (ch - 6) * nWords
    short tx, ty;                  // TMA box origin inside the bordered level (tx 16-byte aligned)
    short pad[3];
    unsigned inv;                  // ceil(2^20 / nWords): k_fast_cells splits a work unit index without dividing
    unsigned firstMask, lastMask;  // 0x80 in the bytes of the first / last tested word of a row that are tested columns
    unsigned pad2;
};
static_assert(sizeof(CellDesc) == 48, "CellDesc layout");

// Bilinear resize coefficients of one destination column/row (OpenCV 11-bit fixed point).
struct ResizeTap {
    short s0, s1, a0, a1;
};

// x side of the bilinear resize for one 32-bit word (4 columns) of a bordered destination row.
struct ColPlan {
    int base01, base23;    // source word offsets (from ROI column 0) of the 8-byte windows of columns 0-1 / 2-3
    uint32_t sel;          // PRMT selectors: low half gathers (S[s0_0], S[s1_0], S[s0_1], S[s1_1]), high half columns 2-3
    uint32_t a[4];         // a0 | a1 << 16 per column (11-bit fixed point, a0 + a1 = 2048)
    uint32_t keep;         // 0xFF per byte inside [-19, w + 19)
};

// TMA-staged form of the resize (k_pyr_resize_tma, the default; k_pyr_resize_plan is the L1 form): a CTA covers 128 destination
// bytes x 64 destination rows; the source rows / columns its taps touch are one TMA box of the parent level.
#ifndef ORBFE_PYRTMA_ROWS
#define ORBFE_PYRTMA_ROWS 64
#endif
struct PyrTmaLevel {
    int ok;                // 0: this level keeps the plain kernel (no column plan, or a box beyond 256)
    int boxWords, boxH;    // TMA box: source words per row, source rows
    int tilesX, groupsY;
    int colLo[24];         // first source word (from ROI column 0, a multiple of 4) per destination tile column
    int rowLo[96];         // first source row (ROI row) per group of ORBFE_PYRTMA_ROWS destination rows
};
struct PyrTmaPlan {
    PyrTmaLevel lv[ORBFE_MAX_LEVELS];
};

// One tiled tensor map per pyramid level: dims (pitch, rows, frames) of the bordered level, box = FAST tile.
struct FastMaps {
    CUtensorMap m[ORBFE_MAX_LEVELS];
};

// One captured extraction (all chunks, all compute streams) for one set of call arguments.
struct OrbfeGraphSlot {
    const uint8_t* images;
    int batch, W, H, lap0, lap1, format, seen;
    long long stride, frameStride, lastUse;
    cudaGraphExec_t exec;
    long long kernels;      // kernel nodes in the graph (launch accounting)
};

struct orbfe_handle {
    orbfe_params p;
    int device;
    cudaStream_t stream;                             // the handle's public stream: every call is ordered on it
    cudaStream_t cur;                                // stream the launchers enqueue on (stream or one of the side streams)
    cudaStream_t sIn, sOut;                          // copy-in / copy-out streams of the pipelined host path
    cudaStream_t sComp[ORBFE_MAX_COMPUTE_STREAMS];   // compute streams of the chunked pipelines; sComp[0] == stream
    cudaEvent_t evPipe[ORBFE_PIPE_MAX_CHUNKS][3];    // per chunk: [0][0] fence, [c][1] frames landed, [c][2] results ready
    cudaEvent_t evFork, evJoin[ORBFE_MAX_COMPUTE_STREAMS], evOrder;
    // the blur of a chunk only needs its pyramid: it runs on a side stream under FAST + quadtree + assemble (enqueue_extract)
    cudaStream_t sSide[ORBFE_MAX_COMPUTE_STREAMS];
    cudaEvent_t evPyr[ORBFE_MAX_COMPUTE_STREAMS], evBlur[ORBFE_MAX_COMPUTE_STREAMS];
    int blurBackend;                               // ORBFE_BLUR_ALU / ORBFE_BLUR_TENSOR (include/orbfe.h)
    int blurOverlap, blurPersist, smCount;         // overlap on/off; resident blur CTAs per SM (0: one CTA per four tiles)
    int devChunk, devStreams;                        // device-resident pipeline: frames per chunk, compute streams used
    int useGraphs;
    OrbfeGraphSlot graphs[ORBFE_GRAPH_SLOTS];
    long long graphLaunches, graphClock;
    long long launches;
    // reference tables (ORBextractor.cc:413-444)
    float scale[ORBFE_MAX_LEVELS], invScale[ORBFE_MAX_LEVELS], sigma2[ORBFE_MAX_LEVELS], invSigma2[ORBFE_MAX_LEVELS];
    int featPerLevel[ORBFE_MAX_LEVELS];
    // current shape
    int curW, curH;
    int lastBatch;
    LevelGeom geom[ORBFE_MAX_LEVELS];
    int nCellsTotal;
    int maxCellBytes;       // dynamic smem of k_fast_cells
    int fastBoxW, fastBoxH; // TMA box of the FAST tile (>= the largest cell, width a multiple of 16)
    int fastMaxCw;          // widest cell ROI (core + 6 px)
    FastMaps fastMaps;
    FastMaps pyrMaps;       // k_pyr_resize_tma: map [l] = the PARENT of level l with level l's source box
    PyrTmaPlan pyrTma;
    PyrTmaPlan* dPyrTma;
    int pyrUseTma;          // resize levels through k_pyr_resize_tma (default; ORBFE_PYR_TMA=0 selects the L1 form, see k_pyramid.cu)
    long long candPerFrame; // candidate entries per frame (all levels)
    int kpPerFrame;         // output slots per frame (= cap of device results)
    // capacities fixed at create
    size_t pyrBytes, candEntries, inBytes;
    int cellCap, tabCap, kpCapPerFrame;
    // device arenas
    uint8_t* dIn;           // staged input frames (host path)
    int inFormat, inBpp;    // ORBFE_FORMAT_* of the input images; bytes per pixel the staging arena is sized for
    uint8_t* dPyr;          // pyramid arena [level][frame][rows][pitch]
    uint8_t* dBlur;         // Gaussian-blurred levels, same layout (ROI pixels only are defined)
    int blurTiles;          // k_blur tiles per frame (all levels)
    uint32_t* dCand;        // packed candidates [frame][level slice]
    uint32_t* dNodeOf;      // quadtree scratch, same indexing as dCand
    int* dCandCount;        // [frame][level]
    uint32_t* dLvlKp;       // selected keypoints per (frame, level slice): packed like candidates
    int* dLvlCount;         // [frame][level]
    LevelGeom* dGeom;
    CellDesc* dCells;
    ResizeTap* dTaps;
    ColPlan* dPlans;
    int planCap;
    uint8_t* dRes;          // result arena: dN | dMono | dKps | dDesc, contiguous (small-batch handles fetch it with one copy)
    uint8_t* hRes;          // pinned mirror of the result arena (max_batch <= ORBFE_SMALL_BATCH), else NULL
    size_t resBytes, resKpsOff, resDescOff;
    orbfe_keypoint* dKps;   // [frame][kpPerFrame]
    uint8_t* dDesc;         // [frame][kpPerFrame][32]
    uint32_t* dSlot;        // [frame][kpPerFrame] : level << 28 | y << 14 | x (level coords) of the keypoint at dst slot
    int* dN;                // [frame]
    int* dMono;             // [frame]
    // pinned host staging for the small results
    int* hN;
    int* hMono;
    // stage profiling: a ring of event sets, one per extract call in flight
    int profiling;
    int profHead, profCount;                  // ring state
    cudaEvent_t (*profEv)[ORBFE_NUM_STAGES + 1];
    double profMs[ORBFE_NUM_STAGES];
    long long profCalls;
};
#define ORBFE_PROF_RING 64

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a per-device setting: raises it to `bytes` for the current device if this
// (device, kernel) pair has not been given that much yet.  Thread-safe; no-op up to 48 KB.
cudaError_t orbfe_ensure_dyn_smem(const void* func, size_t bytes);

// ---- launchers (each returns cudaGetLastError()) -------------------------------------------------
// Every launcher processes frames [frame0, frame0 + batch) of the handle's arenas on h->cur (chunked pipelines of
// the host and device paths); dImages points at the first of those frames.
cudaError_t orbfe_launch_pyramid(orbfe_handle* h, const uint8_t* dImages, int frame0, int batch, int width, int height,
                                 long long stride, long long frameStride);
cudaError_t orbfe_launch_pyramid_levels(orbfe_handle* h, const uint8_t* dImages, int frame0, int batch, long long stride,
                                        long long frameStride, int lFirst, int lEnd, cudaStream_t st);
int orbfe_fast_side_mode();
int orbfe_fast_tall_from(const orbfe_handle* h, int batch);
cudaError_t orbfe_clear_candidates(orbfe_handle* h, int frame0, int batch);
cudaError_t orbfe_launch_fast(orbfe_handle* h, int frame0, int batch, cudaStream_t sideStream);
cudaError_t orbfe_launch_octree(orbfe_handle* h, int frame0, int batch);
cudaError_t orbfe_launch_assemble(orbfe_handle* h, int frame0, int batch, int lap0, int lap1);
cudaError_t orbfe_launch_blur(orbfe_handle* h, int frame0, int batch, cudaStream_t st, int persistPerSm);
cudaError_t orbfe_launch_blur_tc(orbfe_handle* h, int frame0, int batch, cudaStream_t st);     // k_blur_tc.cu
cudaError_t orbfe_launch_describe(orbfe_handle* h, int frame0, int batch);
