// bow_kernels.cu — bag-of-words assignment of ORB descriptors against a DBoW2 vocabulary tree on the device
// (SURVEY.md §8(f) f1): DBoW2::TemplatedVocabulary<cv::Mat, FORB>::transform(features, BowVector&, FeatureVector&,
// levelsup), Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1149-1218 with the per-feature descent of :1242-1282, as
// Frame::ComputeBoW / KeyFrame::ComputeBoW call it (src/Frame.cc:495-500, src/KeyFrame.cc:295-302).
#include "ygz_match.cuh"

namespace ygz {

// ---- K11: tree descent, one thread per feature.  The children of a node are stored contiguously (child-list order:
// entry c holds the NodeId childList[c] and its 32-byte descriptor), so a level costs k x 2 16-byte loads from an
// L2-resident table.  First minimum wins (strict <, :1267); a node without children is a leaf (:118).
__global__ void __launch_bounds__(256) k_bow_descend(const __grid_constant__ BowArgs A) {
    const int f = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= min(A.nFeat[f], A.stride)) return;
    const size_t o = (size_t)f * A.stride + i;
    const uint4 *d = reinterpret_cast<const uint4 *>(A.desc + o * 32);
    const uint4 d0 = d[0], d1 = d[1];
    const uint4 *cdesc = reinterpret_cast<const uint4 *>(A.childDesc);
    const int nidLevel = A.L - A.levelsup;
    int nid = 0, node = 0, level = 0;
    int cs = A.childStart[0], ce = A.childStart[1];
    while (cs < ce) {
        ++level;
        int best = cs, bestD = hamming256(d0, d1, cdesc[2 * cs], cdesc[2 * cs + 1]);
        for (int c = cs + 1; c < ce; c++) {
            const int dist = hamming256(d0, d1, cdesc[2 * c], cdesc[2 * c + 1]);
            if (dist < bestD) { bestD = dist; best = c; }
        }
        node = A.childList[best];
        if (level == nidLevel) nid = node;
        cs = A.childStart[node]; ce = A.childStart[node + 1];
    }
    A.word[o] = A.wordId[node];
    A.node[o] = nid;
    A.weight[o] = A.nodeWeight[node];
}

// ---- K12: BowVector and FeatureVector of one frame, one CTA per frame.  Both are std::maps filled in feature
// order (BowVector::addWeight BowVector.cpp:33-46, FeatureVector::addFeature FeatureVector.cpp:33-46): sorting the
// live features (weight > 0, :1184) by (word, index) resp. (node, index) gives the maps' iteration order with the
// features of a key in insertion order.  A word's value is w + w + ... in that order (sequential double adds, not
// n * w); BowVector::normalize(L1) (BowVector.cpp:70-98) sums |value| over ascending words sequentially — one thread
// does it so that the doubles come out bit for bit — and divides.
constexpr int kBowThreads = 512;

__device__ void bowBitonicSort(unsigned long long *keys, int n2, int tid) {
    for (int k = 2; k <= n2; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < n2; i += kBowThreads) {
                const int p = i ^ j;
                if (p > i) {
                    const unsigned long long a = keys[i], b = keys[p];
                    const bool up = (i & k) == 0;
                    if ((a > b) == up) { keys[i] = b; keys[p] = a; }
                }
            }
            __syncthreads();
        }
}

// exclusive scan of flags[0..n) in place over the CTA (n <= n2, a power of two <= 8192); returns the total
__device__ int bowScan(int *a, int n, int *ws, int tid) {
    const int lane = tid & 31, wid = tid >> 5;
    const int chunk = (n + kBowThreads - 1) / kBowThreads;
    const int b = min(tid * chunk, n), e = min(b + chunk, n);
    int sum = 0;
    for (int i = b; i < e; i++) sum += a[i];
    int inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
    if (lane == 31) ws[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        const int w = lane < kBowThreads / 32 ? ws[lane] : 0;
        int winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, winc, o); if (lane >= o) winc += v; }
        if (lane < kBowThreads / 32) ws[lane] = winc - w;
        if (lane == 31) ws[32] = winc;
    }
    __syncthreads();
    int run = ws[wid] + inc - sum;
    for (int i = b; i < e; i++) { const int v = a[i]; a[i] = run; run += v; }
    const int total = ws[32];
    __syncthreads();
    return total;
}

__global__ void __launch_bounds__(kBowThreads) k_bow_vectors(const __grid_constant__ BowArgs A, int n2) {
    extern __shared__ __align__(16) unsigned char bowSmem[];
    unsigned long long *keys = reinterpret_cast<unsigned long long *>(bowSmem);   // n2 entries
    int *flag = reinterpret_cast<int *>(keys + n2);                                  // n2 entries
    __shared__ int ws[40];
    __shared__ double normS;
    const int f = blockIdx.x, tid = threadIdx.x;
    const int n = min(A.nFeat[f], A.stride);
    const size_t fo = (size_t)f * A.stride;
    const int32_t *word = A.word + fo, *node = A.node + fo;
    const double *weight = A.weight + fo;
    constexpr unsigned long long kDead = ~0ull;
    for (int pass = 0; pass < 2; pass++) {
        const int32_t *key = pass == 0 ? word : node;
        for (int i = tid; i < n2; i += kBowThreads)
            keys[i] = (i < n && weight[i] > 0) ? (((unsigned long long)(uint32_t)key[i] << 32) | (uint32_t)i) : kDead;
        __syncthreads();
        bowBitonicSort(keys, n2, tid);
        // segment heads
        for (int i = tid; i < n2; i += kBowThreads)
            flag[i] = keys[i] != kDead && (i == 0 || (keys[i] >> 32) != (keys[i - 1] >> 32));
        __syncthreads();
        const int total = bowScan(flag, n2, ws, tid);
        if (pass == 0) {
            double *val = A.bowValue + fo;
            int32_t *bw = A.bowWord + fo;
            for (int i = tid; i < n2; i += kBowThreads) {
                if (keys[i] == kDead) continue;
                const bool head = i == 0 || (keys[i] >> 32) != (keys[i - 1] >> 32);
                if (!head) continue;
                const int slot = flag[i];
                const unsigned long long wk = keys[i] >> 32;
                const double w = weight[(uint32_t)keys[i]];
                double v = w;
                for (int j = i + 1; j < n2 && (keys[j] >> 32) == wk; j++) v = __dadd_rn(v, w);   // value += w, feature order
                bw[slot] = (int32_t)wk;
                val[slot] = v;
            }
            __syncthreads();
            if (tid == 0) {
                double norm = 0.0;
                for (int s = 0; s < total; s++) norm = __dadd_rn(norm, fabs(val[s]));
                normS = norm;
                A.bowCount[f] = total;
            }
            __syncthreads();
            const double norm = normS;
            if (norm > 0.0)
                for (int s = tid; s < total; s += kBowThreads) val[s] = __ddiv_rn(val[s], norm);
        } else {
            uint32_t *fn = A.fvNode + fo, *ff = A.fvFeat + fo;
            int32_t *fs = A.fvStart + (size_t)f * (A.stride + 1);
            int live = 0;
            for (int i = tid; i < n2; i += kBowThreads) {
                if (keys[i] == kDead) continue;
                live++;
                ff[i] = (uint32_t)keys[i];                       // sorted position == CSR position
                const bool head = i == 0 || (keys[i] >> 32) != (keys[i - 1] >> 32);
                if (head) { fn[flag[i]] = (uint32_t)(keys[i] >> 32); fs[flag[i]] = i; }
            }
            if (tid == 0) ws[33] = 0;
            __syncthreads();
            if (live) atomicAdd(&ws[33], live);
            __syncthreads();
            if (tid == 0) { fs[total] = ws[33]; A.fvCount[f] = total; }
        }
        __syncthreads();
    }
}

int launchBowTransform(const BowArgs &A, cudaStream_t s) {
    if (A.nFrames <= 0) return 0;
    k_bow_descend<<<dim3((A.stride + 255) / 256, A.nFrames), 256, 0, s>>>(A);
    int n2 = 32;
    while (n2 < A.stride) n2 <<= 1;
    const size_t smem = (size_t)n2 * (sizeof(unsigned long long) + sizeof(int));
    static bool configured[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        cudaFuncSetAttribute(k_bow_vectors, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192 * 12);
        configured[dev] = true;
    }
    k_bow_vectors<<<A.nFrames, kBowThreads, smem, s>>>(A, n2);
    return 2;
}

}  // namespace ygz
