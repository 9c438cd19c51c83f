// ygz_common.cuh — shared declarations of libygzorb_cuda.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <string>

#include "../../include/ygzorb.h"

namespace ygz {

constexpr int kMaxLevels = 16;
constexpr int kEdge = 19;        // EDGE_THRESHOLD, src/ORBextractor.cc:72
constexpr int kMinBorder = 16;   // EDGE_THRESHOLD - 3, :733
constexpr int kHalfPatch = 15;   // HALF_PATCH_SIZE, :71
constexpr int kFastCtaThreads = 64;  // k_fast CTA size (extract_kernels.cu: kFastThreads); sizes the per-warp lists
constexpr int kFastTilePitchNarrow = 48;   // k_fast tile pitch for cells up to 39 px wide
constexpr int kFastTilePitch = 76;   // k_fast image / score tile pitch: cell + 6 + up to 3 alignment bytes, multiple of 4
constexpr int kBlurTW = 256, kBlurTH = 32;  // blur tile (k_blur)

// Per-level geometry, computed once per (w, h) on the host with the reference's float arithmetic
// (src/ORBextractor.cc:728-745 cells, :537-556 quadtree roots, :1131-1133 sizes).
struct LevelGeom {
    int w, h, pitch;          // level size; device row pitch (level 0: the input's stride)
    long long off;            // byte offset of the level inside a frame's pyramid slab (levels >= 1)
    int nCols, nRows, wCell, hCell;  // FAST cell grid; nCols == 0: the level yields nothing
    int cellStart;            // index of the level's first cell in the frame-wide cell enumeration
    int nTarget;              // mnFeaturesPerLevel[level]
    int nIni;                 // quadtree roots
    float hX;                 // root width
    int candCap;              // candidate slots
    long long candOff;        // first candidate slot of the level inside the frame's candidate array
    int keptOff;              // first kept-list slot of the level inside the frame's kept array
    int listCap;              // quadtree list bound of the level
    float scale;              // mvScaleFactor[level]
    float kpSize;             // (float)(int)(31 * scale), :787
    int tabX, tabY;           // offsets of the resize tables (in 8-byte entries)
    int area2x;               // cv::resize switches exact 2x decimation to INTER_AREA
    int fastResize;           // the level's tables satisfy the assumptions of the word-based k_resize
    int tabXG;                // k_resize_band: pre-digested X table, 3 x int4 per column group (offset in 8-byte entries, 16-byte aligned)
    int rawPitch;             // k_resize_band<STAGED>: bytes of a source row one strip stages (multiple of 16, <= 256); 0 = not stageable
    int bandRows;             // k_resize_band: most source rows (y0(first) .. y1(last)) any 64-row destination tile touches; 0 = not usable
    int tileStart, tilesX, tilesY;  // blur tiles
};

struct FrameGeom {
    int nlevels;
    int w, h;
    int iniTh, minTh;
    int totalCells, totalTiles;   // totalCells: entries of the FAST cell table (cells with a non-empty detection area)
    int tileTab;                  // offset of the blur tile table in DevBuffers::tabs (x = x0 | y0<<12 | level<<24)
    int fastPitch;                // k_fast tile pitch of this geometry (kFastTilePitchNarrow or kFastTilePitch)
    int fastRegionA, fastScoreBytes, fastWarpCap;   // k_fast dynamic shared memory layout (bytes, bytes, list entries per warp)
    int cellTab;                  // offset of the cell table in DevBuffers::tabs (8-byte entries)
    int maxListCap;
    int keptCap;               // kept-list slots per frame (= max keypoints per frame)
    long long slabBytes;       // pyramid slab per frame (levels >= 1)
    long long candPerFrame;    // candidate slots per frame
    int umax[16];
    int vmaxOfU[16];   // rows |v| <= vmaxOfU[|u|] of the circular patch hold column u (umax is non-increasing)
    LevelGeom lv[kMaxLevels];
};

struct DevBuffers {
    const uint8_t *in;         // level 0 of every frame: in + f*inFrameStride, pitch lv[0].pitch
    long long inFrameStride;
    uint8_t *pyr;              // levels >= 1: pyr + f*slabBytes + lv[l].off
    uint8_t *blur;             // blurred levels, all levels: blur + f*blurSlab + blurOff[l]
    long long blurSlab;
    long long blurOff[kMaxLevels];
    int blurPitch[kMaxLevels];
    const int2 *tabs;          // resize tables + FAST cell table (x = iniX | iniY<<12 | level<<24, y = dw | dh<<7 | cell<<14)
    uint2 *cand;               // candidates: cand + f*candPerFrame + lv[l].candOff
    int *candCount;            // [frame][kMaxLevels]
    uint16_t *nodeOf;          // quadtree scratch, same indexing as cand
    uint32_t *kept;            // kept keypoints x | y<<12 | resp<<24: kept + f*keptCap + lv[l].keptOff
    int *keptCount;            // [frame][kMaxLevels]
    int *status;               // device-side error flags
    int pdl;                   // launched with programmatic dependent launch (small batches): kernels release their dependents early
};

#ifdef __CUDACC__
// Programmatic dependent launch (sm_90+).  Every kernel of the extraction chain waits for its predecessor before it touches
// anything a previous kernel wrote; with B.pdl it first lets the NEXT kernel of the stream start launching, so that kernel's
// launch latency and prologue overlap this kernel's execution.  Used for small batches (live tracking), where the chain of 11
// short kernels is launch-latency bound; a no-op for kernels launched the ordinary way.
__device__ __forceinline__ void pdlEnter(const DevBuffers &B) {
    if (B.pdl) asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}
#endif

__device__ __forceinline__ const uint8_t *levelPtr(const FrameGeom &G, const DevBuffers &B, int f, int l) {
    return l == 0 ? B.in + (long long)f * B.inFrameStride : B.pyr + (long long)f * G.slabBytes + G.lv[l].off;
}

void setError(const std::string &msg);
int cudaFail(cudaError_t e, const char *what, const char *file, int line);

#define YGZ_CUDA(call)                                                            \
    do {                                                                          \
        cudaError_t e__ = (call);                                                 \
        if (e__ != cudaSuccess) return ygz::cudaFail(e__, #call, __FILE__, __LINE__); \
    } while (0)

// Kernel launchers (extract_kernels.cu).  Each returns the number of kernel launches it enqueued.
int launchResizeLevel(const FrameGeom &G, const DevBuffers &B, int level, int nFrames, cudaStream_t s);
int launchFast(const FrameGeom &G, const DevBuffers &B, int nFrames, cudaStream_t s);
int launchQuadtree(const FrameGeom &G, const DevBuffers &B, int nFrames, cudaStream_t s);
int launchBlur(const FrameGeom &G, const DevBuffers &B, int nFrames, cudaStream_t s);
int launchDescribe(const FrameGeom &G, const DevBuffers &B, int nFrames, int cap, int32_t *counts,
                   ygzorb_keypoint *kps, uint8_t *desc, cudaStream_t s);
int launchRemap(const uint8_t *src, long long srcFrameStride, int srcPitch, int w, int h, int nFrames, const short2 *map1,
                const uint16_t *map2, uint8_t *dst, cudaStream_t s);
size_t quadtreeSmemBytes(int listCap);
int configureKernels();  // one-time cudaFuncSetAttribute calls

// Views into an extractor's device-resident pyramid for the matcher (extractor.cu).  Both drain the
// extractor's stream first, so the pointers are safe to read from another stream.
struct LevelView { const uint8_t *ptr; int pitch, w, h; };
int extractorLevelView(ygzorb_extractor *ex, int frame, int level, LevelView *out);
int extractorTables(ygzorb_extractor *ex, int *nlevels, const float **scale, const float **invScale);

}  // namespace ygz
