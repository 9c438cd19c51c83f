// dso_kernels.cu — the keypoint variant the system runs on frames that come out of direct tracking:
// ORBextractor::ComputeKeyPointsDSOSingleLevel (src/ORBextractor.cc:1275-1386) — grid of mnGridSize cells on level 0, FAST-10
// (Thirdparty/fast) at threshold 20 with a fallback to 5, Shi-Tomasi score (:1152-1187), the three best corners per cell —
// and the descriptor pass of operator()(Frame*, ..., DSO_KEYPOINT) (:1031-1127) over existing + new keypoints.
#include "ygz_common.cuh"
#include "ygz_dso.cuh"

namespace ygz {

constexpr int kDsoThreads = 128;

// 16-pixel Bresenham ring of Thirdparty/fast/src/fast_10.cpp:16-33, as (dx, dy)
__constant__ signed char kRing[16][2] = {{0, 3}, {1, 3}, {2, 2}, {3, 1}, {3, 0}, {3, -1}, {2, -2}, {1, -3},
                                         {0, -3}, {-1, -3}, {-2, -2}, {-3, -1}, {-3, 0}, {-3, 1}, {-2, 2}, {-1, 3}};

__device__ __forceinline__ bool run10(unsigned m) {   // 10 contiguous set bits in a circular 16-bit mask
    m |= m << 16;
    const unsigned x2 = m & (m >> 1);
    const unsigned x4 = x2 & (x2 >> 2);
    const unsigned x8 = x4 & (x4 >> 4);
    return (x8 & (x2 >> 8) & 0xffffu) != 0;
}

// K-dso-0: mOccupancy (:1281-1285): the pixels of the existing keypoints, one bit each
__global__ void k_dso_occupancy(DsoArgs A) {
    const int f = blockIdx.y;
    const int n = min(A.nExist[f], A.cap);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const ygzorb_keypoint &kp = A.kps[(size_t)f * A.cap + i];
        const int x = __float2int_rn(kp.x), y = __float2int_rn(kp.y);
        if (x < 0 || y < 0 || x >= A.w || y >= A.h) continue;
        atomicOr(&A.occ[(size_t)f * A.occWords + (size_t)y * A.occPitch + (x >> 5)], 1u << (x & 31));
    }
}

// K-dso-1: one CTA per (cell, frame) of the frames still looking for a grid size
__global__ void __launch_bounds__(kDsoThreads) k_dso_cells(DsoArgs A) {
    extern __shared__ unsigned char smem[];
    const int f = blockIdx.y, k = blockIdx.x, tid = threadIdx.x;
    if (!A.active[f]) return;
    const int g = A.grid[f];
    const int rows = A.h / g, cols = A.w / g;
    if (k >= rows * cols) return;
    uint8_t *cellCnt = A.cellCnt + (size_t)f * A.maxCells;
    const int nn = k / cols, cc = k - nn * cols;
    if (nn == 0 || nn == rows - 1 || cc == 0 || cc == cols - 1) { if (tid == 0) cellCnt[k] = 0; return; }   // :1320-1323
    const int x0 = cc * g, y0 = nn * g;
    // fast_corner_detect_10 (this fork's scalar version, used for widths < 22) tests every pixel of the cell and reads the ring
    // across the cell border; the SSE2 version (widths >= 22) keeps a 3-pixel margin (faster_corner_10_sse.cpp:24-29, 188-193)
    const int m = g < 22 ? 0 : 3;
    const int side = g - 2 * m, nPix = side * side;
    const int tp = g + 6;                           // tile = cell + 3-pixel halo
    unsigned char *tile = smem;                     // tp * tp
    unsigned char *flag = tile + ((tp * tp + 3) & ~3);          // nPix: bit0 corner at 20, bit1 corner at 5
    unsigned short *list = (unsigned short *)(flag + ((nPix + 3) & ~3));   // nPix
    float *score = (float *)(list + ((nPix + 1) & ~1));          // nPix
    __shared__ int nList;
    __shared__ unsigned long long bestKey;
    const uint8_t *img = A.img + (size_t)f * A.frameStride;
    for (int i = tid; i < tp * tp; i += kDsoThreads) {
        const int ty = i / tp, tx = i - ty * tp;
        tile[i] = img[(size_t)(y0 - 3 + ty) * A.pitch + (x0 - 3 + tx)];
    }
    __syncthreads();
    int any20 = 0;
    for (int i = tid; i < nPix; i += kDsoThreads) {
        const int py = i / side, px = i - py * side;
        const unsigned char *c = tile + (py + m + 3) * tp + (px + m + 3);
        const int p = *c;
        unsigned b20 = 0, d20 = 0, b5 = 0, d5 = 0;
#pragma unroll
        for (int r = 0; r < 16; r++) {
            const int v = c[kRing[r][1] * tp + kRing[r][0]];
            b20 |= (unsigned)(v > p + 20) << r; d20 |= (unsigned)(v < p - 20) << r;
            b5 |= (unsigned)(v > p + 5) << r; d5 |= (unsigned)(v < p - 5) << r;
        }
        const int c20 = run10(b20) || run10(d20), c5 = run10(b5) || run10(d5);
        flag[i] = (unsigned char)(c20 | (c5 << 1));
        any20 |= c20;
    }
    const int use = __syncthreads_or(any20) ? 1 : 2;     // :1337-1345 (empty at 20 -> threshold 5)
    // ---- corners in raster order that are far enough from the border and not occupied (:1352-1363)
    if (tid < 32) {
        int cnt = 0;
        const uint32_t *occ = A.occ + (size_t)f * A.occWords;
        for (int base = 0; base < nPix; base += 32) {
            const int i = base + tid;
            bool ok = false;
            if (i < nPix && (flag[i] & use)) {
                const int py = i / side, px = i - py * side;
                const int x = x0 + m + px, y = y0 + m + py;
                ok = !(x < 20 || y < 20 || x >= A.w - 20 || y >= A.h - 20) && !((occ[(size_t)y * A.occPitch + (x >> 5)] >> (x & 31)) & 1u);
            }
            const unsigned bal = __ballot_sync(0xffffffffu, ok);
            if (ok) list[cnt + __popc(bal & ((1u << tid) - 1))] = (unsigned short)i;
            cnt += __popc(bal);
        }
        if (tid == 0) nList = cnt;
    }
    __syncthreads();
    const int n = nList;
    // ---- ShiTomasiScore (:1152-1187): integer gradient sums over the 8x8 box (exact in float), float eigenvalue formula
    for (int j = tid; j < n; j += kDsoThreads) {
        const int i = list[j], py = i / side, px = i - py * side;
        const int u = x0 + m + px, v = y0 + m + py;
        float sc = 0.0f;
        if (!(u - 4 < 1 || u + 4 >= A.w - 1 || v - 4 < 1 || v + 4 >= A.h - 1)) {
            int dXX = 0, dYY = 0, dXY = 0;
            for (int y = v - 4; y < v + 4; y++) {
                const uint8_t *row = img + (size_t)y * A.pitch + (u - 4);
#pragma unroll
                for (int x = 0; x < 8; x++) {
                    const int dx = (int)row[x + 1] - (int)row[x - 1];
                    const int dy = (int)row[x + A.pitch] - (int)row[x - A.pitch];
                    dXX += dx * dx; dYY += dy * dy; dXY += dx * dy;
                }
            }
            const float a = (float)((double)dXX / 128.0), b = (float)((double)dYY / 128.0), c = (float)((double)dXY / 128.0);
            const float s = __fadd_rn(a, b);
            const float disc = __fsub_rn(__fmul_rn(s, s), __fmul_rn(4.0f, __fsub_rn(__fmul_rn(a, b), __fmul_rn(c, c))));
            sc = (float)(0.5 * (double)__fsub_rn(s, __fsqrt_rn(disc)));
        }
        score[j] = sc;
    }
    __syncthreads();
    // ---- the three best by score, earlier corner first on ties (std::sort's insertion sort for <= 16 elements is stable;
    // for longer lists the order of equal scores is unspecified in the reference: canonical = raster order)
    const int take = min(n, 3);
    uint32_t *out = A.cellKp + ((size_t)f * A.maxCells + k) * 3;
    for (int r = 0; r < take; r++) {
        if (tid == 0) bestKey = 0ull;
        __syncthreads();
        unsigned long long mine = 0ull;
        for (int j = tid; j < n; j += kDsoThreads) {
            const float sc = score[j];
            unsigned bits = __float_as_uint(sc);
            unsigned ord = (sc != sc) ? 1u : (bits & 0x80000000u ? ~bits : bits | 0x80000000u);
            if (list[j] == 0xffffu) ord = 0u;          // already taken
            const unsigned long long key = ((unsigned long long)ord << 32) | (unsigned)(0xffffffffu - (unsigned)j);
            if (ord && key > mine) mine = key;
        }
        if (mine) atomicMax(&bestKey, mine);
        __syncthreads();
        if (tid == 0) {
            const int j = (int)(0xffffffffu - (unsigned)(bestKey & 0xffffffffull));
            const int i = list[j], py = i / side, px = i - py * side;
            out[r] = (unsigned)(x0 + m + px) | ((unsigned)(y0 + m + py) << 16);
            list[j] = 0xffffu;
        }
        __syncthreads();
    }
    if (tid == 0) { cellCnt[k] = (uint8_t)take; if (take) atomicAdd(&A.total[f], take); }
}

// K-dso-2: cell-major assembly of the new keypoints behind the existing ones + the (level, x, y) centre list for the
// orientation / descriptor pass.  One CTA per frame.
__global__ void __launch_bounds__(512) k_dso_assemble(DsoArgs A) {
    __shared__ int wsum[16], carry;
    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int g = A.grid[f];
    const int cells = (A.h / g) * (A.w / g);
    const int nE = min(A.nExist[f], A.cap);
    const uint8_t *cellCnt = A.cellCnt + (size_t)f * A.maxCells;
    ygzorb_keypoint *kps = A.kps + (size_t)f * A.cap;
    int4 *centre = A.centre + (size_t)f * A.cap;
    // existing keypoints: centre on their own level, pt * mvInvScaleFactor[octave] rounded (:1094-1097, :1383-1385)
    for (int i = tid; i < nE; i += 512) {
        const ygzorb_keypoint kp = kps[i];
        const int oct = min(max(kp.octave, 0), A.nlevels - 1);
        const float s = A.invScale[oct];
        centre[i] = make_int4(oct, __float2int_rn(__fmul_rn(kp.x, s)), __float2int_rn(__fmul_rn(kp.y, s)), 0);
    }
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < cells; base += 512) {
        const int k = base + tid;
        const int c = k < cells ? cellCnt[k] : 0;
        int inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
        if (lane == 31) wsum[wid] = inc;
        __syncthreads();
        int before = carry;
        for (int w2 = 0; w2 < wid; w2++) before += wsum[w2];
        const int pos = nE + before + inc - c;
        const uint32_t *src = A.cellKp + ((size_t)f * A.maxCells + k) * 3;
        for (int r = 0; r < c; r++) {
            if (pos + r >= A.cap) break;
            const int x = src[r] & 0xffff, y = src[r] >> 16;
            ygzorb_keypoint kp;
            kp.x = (float)x; kp.y = (float)y; kp.size = 7.0f; kp.angle = -1.0f; kp.response = 0.0f; kp.octave = 0; kp.class_id = -1;
            kps[pos + r] = kp;
            centre[pos + r] = make_int4(0, x, y, 1);
        }
        __syncthreads();
        if (tid == 0) { int t = 0; for (int w2 = 0; w2 < 16; w2++) t += wsum[w2]; carry += t; }
        __syncthreads();
    }
    if (tid == 0) A.counts[f] = min(nE + carry, A.cap);
}

int launchDsoOccupancy(const DsoArgs &A, int maxExist, cudaStream_t s) {
    if (A.nFrames <= 0 || maxExist <= 0) return 0;
    k_dso_occupancy<<<dim3((maxExist + 255) / 256, A.nFrames), 256, 0, s>>>(A);
    return 1;
}

size_t dsoCellSmem(int g) {
    const int tp = g + 6, nPix = g * g;
    return (size_t)((tp * tp + 3) & ~3) + ((nPix + 3) & ~3) + (size_t)((nPix + 1) & ~1) * 2 + (size_t)nPix * 4;
}

int launchDsoCells(const DsoArgs &A, int maxGrid, int maxCellsNow, cudaStream_t s) {
    if (A.nFrames <= 0 || maxCellsNow <= 0) return 0;
    static bool configured[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        cudaFuncSetAttribute(k_dso_cells, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dsoCellSmem(kDsoMaxGrid));
        configured[dev] = true;
    }
    k_dso_cells<<<dim3(maxCellsNow, A.nFrames), kDsoThreads, dsoCellSmem(maxGrid), s>>>(A);
    return 1;
}

int launchDsoAssemble(const DsoArgs &A, cudaStream_t s) {
    if (A.nFrames <= 0) return 0;
    k_dso_assemble<<<A.nFrames, 512, 0, s>>>(A);
    return 1;
}

}  // namespace ygz
