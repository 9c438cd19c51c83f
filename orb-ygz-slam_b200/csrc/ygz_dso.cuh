// ygz_dso.cuh — arguments of the DSO-style keypoint kernels (dso_kernels.cu) and of the list-driven orientation /
// descriptor kernel (extract_kernels.cu: k_describe_list).
#pragma once
#include "ygz_common.cuh"

namespace ygz {

constexpr int kDsoMaxGrid = 96;   // largest mnGridSize the cell kernel holds in shared memory

struct DsoArgs {
    int nFrames, w, h, pitch, nlevels, cap;
    long long frameStride;
    const uint8_t *img;            // level 0 of frame f at img + f*frameStride
    const int32_t *grid;           // mnGridSize per frame (device)
    const uint8_t *active;         // frames that run this iteration of the while (cnt < n) loop
    const int32_t *nExist;         // existing keypoints per frame (the first nExist[f] entries of kps)
    ygzorb_keypoint *kps;          // [nFrames * cap], in-out
    uint32_t *occ;                 // occupancy bitmap, occWords per frame, occPitch words per row
    int occPitch; long long occWords;
    int maxCells;                  // cell slots per frame
    uint32_t *cellKp;              // [nFrames * maxCells * 3]: x | y << 16
    uint8_t *cellCnt;              // [nFrames * maxCells]
    int32_t *total;                // [nFrames] new keypoints of the current iteration
    int32_t *counts;               // [nFrames] existing + new
    int4 *centre;                  // [nFrames * cap]: level, x, y (patch centre on that level), isNew
    float invScale[kMaxLevels];
};

int launchDsoOccupancy(const DsoArgs &A, int maxExist, cudaStream_t s);
int launchDsoCells(const DsoArgs &A, int maxGrid, int maxCellsNow, cudaStream_t s);
int launchDsoAssemble(const DsoArgs &A, cudaStream_t s);
int launchDescribeList(const FrameGeom &G, const DevBuffers &B, int nFrames, int cap, const int4 *centre, const int32_t *counts,
                       ygzorb_keypoint *kps, uint8_t *desc, cudaStream_t s, bool givenAngle = false);

}  // namespace ygz
