// matcher.cu — host side of the Hamming matching paths: handle (stream + scratch pool), argument
// marshalling between host/device buffers, and the C ABI of include/ygzorb.h.  No CPU compute path.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "ygz_match.cuh"

namespace ygz {

static int argFailM(const char *msg) { setError(msg); return YGZORB_E_ARG; }

struct Scratch { void *p = nullptr; size_t bytes = 0; };

struct Matcher {
    int device = 0, smCount = 148;
    cudaStream_t stream = nullptr, ownStream = nullptr;
    std::vector<Scratch> pool;
    long long launches = 0;
    int knnVariant = 2;
    int32_t *dStatus = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
};

// Device scratch slot `slot`, grown on demand (the stream is drained before a buffer is replaced).
static int scratch(Matcher *m, int slot, size_t bytes, void **out) {
    if ((int)m->pool.size() <= slot) m->pool.resize(slot + 1);
    Scratch &s = m->pool[slot];
    if (s.bytes < bytes || !s.p) {
        YGZ_CUDA(cudaStreamSynchronize(m->stream));
        if (s.p) YGZ_CUDA(cudaFree(s.p));
        s.p = nullptr; s.bytes = 0;
        const size_t want = std::max<size_t>(bytes + bytes / 4, 256);
        YGZ_CUDA(cudaMalloc(&s.p, want));
        s.bytes = want;
    }
    *out = s.p;
    return 0;
}

// Input array: device pointers are used in place, host arrays are staged into scratch slot `slot`.
template <class T> static int stageIn(Matcher *m, int slot, const T *src, size_t count, bool inDev, const T **out) {
    if (!src) { *out = nullptr; return 0; }
    if (inDev) { *out = src; return 0; }
    void *p = nullptr;
    int rc = scratch(m, slot, count * sizeof(T), &p);
    if (rc) return rc;
    if (count) YGZ_CUDA(cudaMemcpyAsync(p, src, count * sizeof(T), cudaMemcpyHostToDevice, m->stream));
    *out = (const T *)p;
    return 0;
}
// Output array: device pointers are written in place, host arrays get a scratch slot + a D2H copy later.
template <class T> static int stageOut(Matcher *m, int slot, T *dst, size_t count, bool outDev, T **out) {
    if (dst && outDev) { *out = dst; return 0; }
    void *p = nullptr;
    int rc = scratch(m, slot, std::max<size_t>(count, 1) * sizeof(T), &p);
    if (rc) return rc;
    *out = (T *)p;
    return 0;
}
template <class T> static int copyOut(Matcher *m, T *dst, const T *dev, size_t count, bool outDev) {
    if (!dst || outDev || count == 0) return 0;
    YGZ_CUDA(cudaMemcpyAsync(dst, dev, count * sizeof(T), cudaMemcpyDeviceToHost, m->stream));
    return 0;
}
static int finish(Matcher *m, unsigned flags) {
    YGZ_CUDA(cudaGetLastError());
    if (!(flags & YGZORB_NO_SYNC)) YGZ_CUDA(cudaStreamSynchronize(m->stream));
    return 0;
}

}  // namespace ygz

using namespace ygz;

extern "C" {

int ygzorb_matcher_create(int device, ygzorb_matcher **out) {
    if (!out) return argFailM("null argument");
    YGZ_CUDA(cudaSetDevice(device));
    Matcher *m = new Matcher();
    m->device = device;
    cudaDeviceProp prop;
    cudaError_t e = cudaGetDeviceProperties(&prop, device);
    if (e == cudaSuccess) m->smCount = prop.multiProcessorCount;
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&m->ownStream, cudaStreamNonBlocking);
    m->stream = m->ownStream;
    if (e == cudaSuccess) e = cudaMalloc((void **)&m->dStatus, 64);
    if (e == cudaSuccess) e = cudaMemset(m->dStatus, 0, 64);
    if (e == cudaSuccess) e = cudaEventCreate(&m->ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&m->ev1);
    if (e != cudaSuccess) { ygzorb_matcher_destroy((ygzorb_matcher *)m); return cudaFail(e, "matcher_create", __FILE__, __LINE__); }
    *out = (ygzorb_matcher *)m;
    return 0;
}

void ygzorb_matcher_destroy(ygzorb_matcher *h) {
    Matcher *m = (Matcher *)h;
    if (!m) return;
    cudaSetDevice(m->device);
    // a borrowed stream (ygzorb_matcher_set_stream) may already be gone: drain the device instead of the stream
    if (m->stream == m->ownStream) { if (m->ownStream) cudaStreamSynchronize(m->ownStream); }
    else cudaDeviceSynchronize();
    for (Scratch &s : m->pool) if (s.p) cudaFree(s.p);
    if (m->dStatus) cudaFree(m->dStatus);
    if (m->ev0) cudaEventDestroy(m->ev0);
    if (m->ev1) cudaEventDestroy(m->ev1);
    if (m->ownStream) cudaStreamDestroy(m->ownStream);
    delete m;
}

int ygzorb_matcher_sync(ygzorb_matcher *h) {
    Matcher *m = (Matcher *)h;
    if (!m) return argFailM("null matcher");
    YGZ_CUDA(cudaSetDevice(m->device));
    YGZ_CUDA(cudaStreamSynchronize(m->stream));
    return 0;
}
void *ygzorb_matcher_stream(ygzorb_matcher *h) { return h ? (void *)((Matcher *)h)->stream : nullptr; }
long long ygzorb_matcher_launch_count(const ygzorb_matcher *h) { return h ? ((const Matcher *)h)->launches : 0; }

int ygzorb_matcher_set_stream(ygzorb_matcher *h, void *stream) {
    Matcher *m = (Matcher *)h;
    if (!m) return argFailM("null matcher");
    YGZ_CUDA(cudaSetDevice(m->device));
    YGZ_CUDA(cudaStreamSynchronize(m->stream));
    m->stream = stream ? (cudaStream_t)stream : m->ownStream;
    return 0;
}

int ygzorb_matcher_set_knn_variant(ygzorb_matcher *h, int variant) {
    if (!h || variant < 0 || variant > 3) return argFailM("variant must be 0..3");
    ((Matcher *)h)->knnVariant = variant;
    return 0;
}

int ygzorb_measure_popc_peak(ygzorb_matcher *h, double *popcPerS) {
    Matcher *m = (Matcher *)h;
    if (!m || !popcPerS) return argFailM("null argument");
    YGZ_CUDA(cudaSetDevice(m->device));
    const int blocks = m->smCount * 8, threads = 256, iters = 4096;
    void *out = nullptr;
    int rc = scratch(m, 0, (size_t)blocks * threads * 4, &out);
    if (rc) return rc;
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
        YGZ_CUDA(cudaEventRecord(m->ev0, m->stream));
        m->launches += launchPopcPeak((unsigned *)out, blocks, threads, iters, m->stream);
        YGZ_CUDA(cudaEventRecord(m->ev1, m->stream));
        YGZ_CUDA(cudaStreamSynchronize(m->stream));
        float ms = 0;
        YGZ_CUDA(cudaEventElapsedTime(&ms, m->ev0, m->ev1));
        if (rep > 0) best = std::min(best, ms);
    }
    *popcPerS = (double)blocks * threads * iters * 32.0 / (best * 1e-3);
    return 0;
}

int ygzorb_descriptor_distance(ygzorb_matcher *h, const uint8_t *a, const uint8_t *b, int n, unsigned flags, int32_t *dist) {
    Matcher *m = (Matcher *)h;
    if (!m || n < 0 || (n > 0 && (!a || !b || !dist))) return argFailM("bad argument");
    if (n == 0) return 0;
    YGZ_CUDA(cudaSetDevice(m->device));
    const bool inDev = flags & YGZORB_IN_DEVICE, outDev = flags & YGZORB_OUT_DEVICE;
    const uint8_t *da, *db; int32_t *dd;
    int rc;
    if ((rc = stageIn(m, 0, a, (size_t)n * 32, inDev, &da))) return rc;
    if ((rc = stageIn(m, 1, b, (size_t)n * 32, inDev, &db))) return rc;
    if ((rc = stageOut(m, 2, dist, (size_t)n, outDev, &dd))) return rc;
    m->launches += launchDescriptorDistance(da, db, n, dd, m->stream);
    if ((rc = copyOut(m, dist, dd, (size_t)n, outDev))) return rc;
    return finish(m, flags);
}

int ygzorb_hamming_knn2(ygzorb_matcher *h, const uint8_t *query, int nq, const uint8_t *train, int nt, float ratio,
                        int th, unsigned flags, int32_t *bestIdx, int32_t *d1, int32_t *d2, int32_t *match) {
    Matcher *m = (Matcher *)h;
    if (!m || nq < 0 || nt < 0 || (nq > 0 && !query) || (nt > 0 && !train)) return argFailM("bad argument");
    if (nq == 0) return 0;
    YGZ_CUDA(cudaSetDevice(m->device));
    const bool inDev = flags & YGZORB_IN_DEVICE, outDev = flags & YGZORB_OUT_DEVICE;
    const uint8_t *dq, *dt;
    int rc;
    if ((rc = stageIn(m, 0, query, (size_t)nq * 32, inDev, &dq))) return rc;
    if (nt > 0) { if ((rc = stageIn(m, 1, train, (size_t)nt * 32, inDev, &dt))) return rc; }
    else { void *p; if ((rc = scratch(m, 1, 256, &p))) return rc; dt = (const uint8_t *)p; }
    const int splits = knnSplits(nq, nt, m->smCount, (size_t)256 << 20);
    void *partial = nullptr;
    if ((rc = scratch(m, 2, knnPartialBytes(nq, splits), &partial))) return rc;
    int32_t *oi = nullptr, *o1 = nullptr, *o2 = nullptr, *om = nullptr;
    if (bestIdx && (rc = stageOut(m, 3, bestIdx, (size_t)nq, outDev, &oi))) return rc;
    if (d1 && (rc = stageOut(m, 4, d1, (size_t)nq, outDev, &o1))) return rc;
    if (d2 && (rc = stageOut(m, 5, d2, (size_t)nq, outDev, &o2))) return rc;
    if (match && (rc = stageOut(m, 6, match, (size_t)nq, outDev, &om))) return rc;
    m->launches += launchKnn2(dq, nq, dt, nt, ratio, th, splits, partial, m->knnVariant, oi, o1, o2, om, m->stream);
    if ((rc = copyOut(m, bestIdx, oi, (size_t)nq, outDev))) return rc;
    if ((rc = copyOut(m, d1, o1, (size_t)nq, outDev))) return rc;
    if ((rc = copyOut(m, d2, o2, (size_t)nq, outDev))) return rc;
    if ((rc = copyOut(m, match, om, (size_t)nq, outDev))) return rc;
    return finish(m, flags);
}

int ygzorb_search_by_projection(ygzorb_matcher *h, const ygzorb_frame_features *F, const ygzorb_mappoint_proj *P, float th,
                                int checkLevel, float nnratio, unsigned flags, int32_t *kpToMp, int32_t *mpBestIdx,
                                int32_t *nMatches) {
    Matcher *m = (Matcher *)h;
    if (!m || !F || !P || !kpToMp || !nMatches) return argFailM("null argument");
    if (F->n_frames < 0 || F->kp_stride < 1 || P->mp_stride < 1) return argFailM("bad sizes");
    if (F->nlevels < 1 || F->nlevels > kMaxLevels || !F->scale_factors) return argFailM("nlevels must be 1..16 with scale_factors");
    if (!F->n_kp || !F->kps || !F->desc || !P->n_mp || !P->proj_x || !P->proj_y || !P->view_cos || !P->level || !P->in_view ||
        !P->has_obs || !P->desc) return argFailM("null array in frame/mappoint view");
    if (F->u_right && !P->proj_xr) return argFailM("proj_xr is required when u_right is given");
    if (F->n_frames == 0) return 0;
    YGZ_CUDA(cudaSetDevice(m->device));
    const bool inDev = flags & YGZORB_IN_DEVICE, outDev = flags & YGZORB_OUT_DEVICE;
    const size_t nf = F->n_frames, nk = nf * F->kp_stride, nm = nf * P->mp_stride;
    ProjArgs A;
    memset(&A, 0, sizeof A);
    A.nFrames = F->n_frames; A.kpStride = F->kp_stride; A.mpStride = P->mp_stride;
    int rc;
    if ((rc = stageIn(m, 0, F->n_kp, nf, inDev, &A.nKp))) return rc;
    if ((rc = stageIn(m, 1, P->n_mp, nf, inDev, &A.nMp))) return rc;
    if ((rc = stageIn(m, 2, F->kps, nk, inDev, &A.kps))) return rc;
    if ((rc = stageIn(m, 3, F->desc, nk * 32, inDev, &A.desc))) return rc;
    if ((rc = stageIn(m, 4, F->u_right, F->u_right ? nk : 0, inDev, &A.uRight))) return rc;
    if ((rc = stageIn(m, 5, F->claimed, F->claimed ? nk : 0, inDev, &A.claimed))) return rc;
    if ((rc = stageIn(m, 6, P->proj_x, nm, inDev, &A.projX))) return rc;
    if ((rc = stageIn(m, 7, P->proj_y, nm, inDev, &A.projY))) return rc;
    if ((rc = stageIn(m, 8, P->proj_xr, P->proj_xr ? nm : 0, inDev, &A.projXR))) return rc;
    if ((rc = stageIn(m, 9, P->view_cos, nm, inDev, &A.viewCos))) return rc;
    if ((rc = stageIn(m, 10, P->level, nm, inDev, &A.level))) return rc;
    if ((rc = stageIn(m, 11, P->in_view, nm, inDev, &A.inView))) return rc;
    if ((rc = stageIn(m, 12, P->has_obs, nm, inDev, &A.hasObs))) return rc;
    if ((rc = stageIn(m, 13, P->desc, nm * 32, inDev, &A.mpDesc))) return rc;
    A.minX = F->min_x; A.minY = F->min_y;
    A.invW = (float)kGridCols / (F->max_x - F->min_x);   // mfGridElementWidthInv, src/Frame.cc:296-297
    A.invH = (float)kGridRows / (F->max_y - F->min_y);
    for (int l = 0; l < F->nlevels; l++) A.scale[l] = F->scale_factors[l];
    A.th = th; A.nnratio = nnratio; A.checkLevel = checkLevel;
    if ((rc = stageOut(m, 14, kpToMp, nk, outDev, &A.kpToMp))) return rc;
    if ((rc = stageOut(m, 15, mpBestIdx, nm, outDev, &A.mpBestIdx))) return rc;
    if ((rc = stageOut(m, 16, nMatches, nf, outDev, &A.nMatches))) return rc;
    void *p;
    if ((rc = scratch(m, 17, nk * 4, &p))) return rc;
    A.claimTime = (int32_t *)p;
    if ((rc = scratch(m, 18, nk * 4, &p))) return rc;
    A.gridList = (int32_t *)p;
    if (F->kp_stride > 65535) return argFailM("kp_stride must be <= 65535");
    if ((rc = scratch(m, 19, nm * 12 * 4, &p))) return rc;
    A.candCache = (uint32_t *)p;
    if ((rc = scratch(m, 20, nm, &p))) return rc;
    A.candCnt = (uint8_t *)p;
    m->launches += launchSearchByProjection(A, F->kp_stride, m->stream);
    if ((rc = copyOut(m, kpToMp, A.kpToMp, nk, outDev))) return rc;
    if ((rc = copyOut(m, mpBestIdx, A.mpBestIdx, nm, outDev))) return rc;
    if ((rc = copyOut(m, nMatches, A.nMatches, nf, outDev))) return rc;
    return finish(m, flags);
}


int ygzorb_search_by_projection_last(ygzorb_matcher *h, const ygzorb_frame_features *F, const ygzorb_lastframe_points *L,
                                     float th, int mono, int checkLevel, int checkOrientation, unsigned flags,
                                     int32_t *kpToPt, int32_t *ptBestIdx, int32_t *nMatches) {
    Matcher *m = (Matcher *)h;
    if (!m || !F || !L || !kpToPt || !nMatches) return argFailM("null argument");
    if (F->n_frames < 0 || F->kp_stride < 1 || L->pt_stride < 1) return argFailM("bad sizes");
    if (F->nlevels < 1 || F->nlevels > kMaxLevels || !F->scale_factors) return argFailM("nlevels must be 1..16 with scale_factors");
    if (!F->n_kp || !F->kps || !F->desc || !L->n_pts || !L->valid || !L->has_obs || !L->world_pos || !L->desc || !L->octave ||
        !L->angle || !L->pose) return argFailM("null array in frame/lastframe view");
    if (F->kp_stride > 65535) return argFailM("kp_stride must be <= 65535");
    if (F->n_frames == 0) return 0;
    YGZ_CUDA(cudaSetDevice(m->device));
    const bool inDev = flags & YGZORB_IN_DEVICE, outDev = flags & YGZORB_OUT_DEVICE;
    const size_t nf = F->n_frames, nk = nf * F->kp_stride, nm = nf * L->pt_stride;
    ProjArgs A;
    memset(&A, 0, sizeof A);
    A.nFrames = F->n_frames; A.kpStride = F->kp_stride; A.mpStride = L->pt_stride;
    int rc;
    if ((rc = stageIn(m, 0, F->n_kp, nf, inDev, &A.nKp))) return rc;
    if ((rc = stageIn(m, 1, L->n_pts, nf, inDev, &A.nMp))) return rc;
    if ((rc = stageIn(m, 2, F->kps, nk, inDev, &A.kps))) return rc;
    if ((rc = stageIn(m, 3, F->desc, nk * 32, inDev, &A.desc))) return rc;
    if ((rc = stageIn(m, 4, F->u_right, F->u_right ? nk : 0, inDev, &A.uRight))) return rc;
    if ((rc = stageIn(m, 5, F->claimed, F->claimed ? nk : 0, inDev, &A.claimed))) return rc;
    if ((rc = stageIn(m, 6, L->world_pos, nm * 3, inDev, &A.worldPos))) return rc;
    if ((rc = stageIn(m, 7, L->angle, nm, inDev, &A.angleLast))) return rc;
    if ((rc = stageIn(m, 8, L->octave, nm, inDev, &A.octaveLast))) return rc;
    if ((rc = stageIn(m, 9, L->pose, nf * 30, inDev, &A.pose))) return rc;
    if ((rc = stageIn(m, 11, L->valid, nm, inDev, &A.inView))) return rc;
    if ((rc = stageIn(m, 12, L->has_obs, nm, inDev, &A.hasObs))) return rc;
    if ((rc = stageIn(m, 13, L->desc, nm * 32, inDev, &A.mpDesc))) return rc;
    A.minX = F->min_x; A.minY = F->min_y; A.maxX = F->max_x; A.maxY = F->max_y;
    A.invW = (float)kGridCols / (F->max_x - F->min_x);   // mfGridElementWidthInv, src/Frame.cc:296-297
    A.invH = (float)kGridRows / (F->max_y - F->min_y);
    for (int l = 0; l < F->nlevels; l++) A.scale[l] = F->scale_factors[l];
    A.th = th; A.checkLevel = checkLevel; A.bMono = mono; A.checkOrientation = checkOrientation;
    if ((rc = stageOut(m, 14, kpToPt, nk, outDev, &A.kpToMp))) return rc;
    if ((rc = stageOut(m, 15, ptBestIdx, nm, outDev, &A.mpBestIdx))) return rc;
    if ((rc = stageOut(m, 16, nMatches, nf, outDev, &A.nMatches))) return rc;
    void *p;
    if ((rc = scratch(m, 17, nk * 4, &p))) return rc;
    A.claimTime = (int32_t *)p;
    if ((rc = scratch(m, 18, nk * 4, &p))) return rc;
    A.gridList = (int32_t *)p;
    if ((rc = scratch(m, 19, nm * 12 * 4, &p))) return rc;
    A.candCache = (uint32_t *)p;
    if ((rc = scratch(m, 20, nm, &p))) return rc;
    A.candCnt = (uint8_t *)p;
    m->launches += launchSearchByProjectionLast(A, m->stream);
    if ((rc = copyOut(m, kpToPt, A.kpToMp, nk, outDev))) return rc;
    if ((rc = copyOut(m, ptBestIdx, A.mpBestIdx, nm, outDev))) return rc;
    if ((rc = copyOut(m, nMatches, A.nMatches, nf, outDev))) return rc;
    return finish(m, flags);
}

// Shared driver of the projection searches over MapPoints with a predicted level (kernel modes 2..6).  bestSlot = scratch
// slot of the per-point result, so that SearchBySim3 can keep the first direction's result while the second one runs.
static int searchByProjectionKeyFrame(ygzorb_matcher *h, const ygzorb_frame_features *F, const ygzorb_keyframe_points *L, float th,
                                      int bestDistMax, int checkOrientation, int mode, const float *invSigma2, unsigned flags,
                                      int32_t *kpToPt, int32_t *ptBestIdx, int32_t *nMatches, int bestSlot = 15,
                                      const int32_t **dBest = nullptr, const int32_t **dNPts = nullptr) {
    Matcher *m = (Matcher *)h;
    if (!m || !F || !L) return argFailM("null argument");
    if (F->n_frames < 0 || F->kp_stride < 1 || L->pt_stride < 1) return argFailM("bad sizes");
    if (F->nlevels < 1 || F->nlevels > kMaxLevels || !F->scale_factors) return argFailM("nlevels must be 1..16 with scale_factors");
    if (L->n_scale_levels < 1 || L->n_scale_levels > F->nlevels) return argFailM("n_scale_levels must be 1..nlevels");
    const bool needNormal = mode >= 3 && mode <= 5;
    if (!F->n_kp || !F->kps || !F->desc || !L->n_pts || !L->valid || !L->world_pos || !L->desc || !L->max_distance ||
        !L->min_distance || !L->pose || (needNormal && !L->normal) || (checkOrientation && !L->angle) || (mode == 4 && !invSigma2))
        return argFailM("null array in frame/keyframe-point view");
    if (F->kp_stride > 65535) return argFailM("kp_stride must be <= 65535");
    if (F->n_frames == 0) return 0;
    YGZ_CUDA(cudaSetDevice(m->device));
    const bool inDev = flags & YGZORB_IN_DEVICE, outDev = flags & YGZORB_OUT_DEVICE;
    const size_t nf = F->n_frames, nk = nf * F->kp_stride, nm = nf * L->pt_stride;
    ProjArgs A;
    memset(&A, 0, sizeof A);
    A.nFrames = F->n_frames; A.kpStride = F->kp_stride; A.mpStride = L->pt_stride;
    int rc;
    if ((rc = stageIn(m, 0, F->n_kp, nf, inDev, &A.nKp))) return rc;
    if ((rc = stageIn(m, 1, L->n_pts, nf, inDev, &A.nMp))) return rc;
    if ((rc = stageIn(m, 2, F->kps, nk, inDev, &A.kps))) return rc;
    if ((rc = stageIn(m, 3, F->desc, nk * 32, inDev, &A.desc))) return rc;
    if (mode == 4 && (rc = stageIn(m, 4, F->u_right, F->u_right ? nk : 0, inDev, &A.uRight))) return rc;
    if (mode < 4 && (rc = stageIn(m, 5, F->claimed, F->claimed ? nk : 0, inDev, &A.claimed))) return rc;
    if ((rc = stageIn(m, 6, L->world_pos, nm * 3, inDev, &A.worldPos))) return rc;
    if ((rc = stageIn(m, 7, L->angle, L->angle ? nm : 0, inDev, &A.angleLast))) return rc;
    if ((rc = stageIn(m, 8, L->normal, L->normal ? nm * 3 : 0, inDev, &A.normal))) return rc;
    if ((rc = stageIn(m, 9, L->pose, nf * 30, inDev, &A.pose))) return rc;
    if ((rc = stageIn(m, 10, L->max_distance, nm, inDev, &A.maxDist))) return rc;
    if ((rc = stageIn(m, 11, L->valid, nm, inDev, &A.inView))) return rc;
    if ((rc = stageIn(m, 12, L->min_distance, nm, inDev, &A.minDist))) return rc;
    if ((rc = stageIn(m, 13, L->desc, nm * 32, inDev, &A.mpDesc))) return rc;
    A.minX = F->min_x; A.minY = F->min_y; A.maxX = F->max_x; A.maxY = F->max_y;
    A.invW = (float)kGridCols / (F->max_x - F->min_x);
    A.invH = (float)kGridRows / (F->max_y - F->min_y);
    for (int l = 0; l < F->nlevels; l++) { A.scale[l] = F->scale_factors[l]; A.invSigma2[l] = invSigma2 ? invSigma2[l] : 1.0f; }
    A.th = th; A.checkLevel = 1; A.checkOrientation = checkOrientation;
    A.logScale = L->log_scale_factor; A.nScaleLevels = L->n_scale_levels; A.bestDistMax = bestDistMax;
    if ((rc = stageOut(m, 14, kpToPt, nk, outDev, &A.kpToMp))) return rc;
    if ((rc = stageOut(m, bestSlot, ptBestIdx, nm, outDev, &A.mpBestIdx))) return rc;
    if ((rc = stageOut(m, 16, nMatches, nf, outDev, &A.nMatches))) return rc;
    void *p;
    if ((rc = scratch(m, 17, nk * 4, &p))) return rc;
    A.claimTime = (int32_t *)p;
    if ((rc = scratch(m, 18, nk * 4, &p))) return rc;
    A.gridList = (int32_t *)p;
    if ((rc = scratch(m, 19, nm * 12 * 4, &p))) return rc;
    A.candCache = (uint32_t *)p;
    if ((rc = scratch(m, 20, nm, &p))) return rc;
    A.candCnt = (uint8_t *)p;
    m->launches += launchSearchByProjectionKeyFrame(A, mode, m->stream);
    if ((rc = copyOut(m, kpToPt, A.kpToMp, nk, outDev))) return rc;
    if ((rc = copyOut(m, ptBestIdx, A.mpBestIdx, nm, outDev))) return rc;
    if ((rc = copyOut(m, nMatches, A.nMatches, nf, outDev))) return rc;
    if (dBest) *dBest = A.mpBestIdx;
    if (dNPts) *dNPts = A.nMp;
    return 0;
}

int ygzorb_search_by_projection_reloc(ygzorb_matcher *h, const ygzorb_frame_features *F, const ygzorb_keyframe_points *L, float th,
                                      int orbDist, int checkOrientation, unsigned flags, int32_t *kpToPt, int32_t *ptBestIdx,
                                      int32_t *nMatches) {
    if (!kpToPt || !nMatches) return argFailM("null argument");
    int rc = searchByProjectionKeyFrame(h, F, L, th, orbDist, checkOrientation, 2, nullptr, flags, kpToPt, ptBestIdx, nMatches);
    return rc ? rc : (F->n_frames ? finish((Matcher *)h, flags) : 0);
}

int ygzorb_search_by_projection_sim3(ygzorb_matcher *h, const ygzorb_frame_features *F, const ygzorb_keyframe_points *L, int th,
                                     unsigned flags, int32_t *kpToPt, int32_t *ptBestIdx, int32_t *nMatches) {
    if (!kpToPt || !nMatches) return argFailM("null argument");
    int rc = searchByProjectionKeyFrame(h, F, L, (float)th, kThLow, 0, 3, nullptr, flags, kpToPt, ptBestIdx, nMatches);
    return rc ? rc : (F->n_frames ? finish((Matcher *)h, flags) : 0);
}

int ygzorb_fuse(ygzorb_matcher *h, const ygzorb_frame_features *KF, const ygzorb_keyframe_points *L, float th, int sim3,
                const float *invLevelSigma2, unsigned flags, int32_t *ptBestIdx, int32_t *nFused) {
    if (!ptBestIdx || !nFused) return argFailM("null argument");
    int rc = searchByProjectionKeyFrame(h, KF, L, th, kThLow, 0, sim3 ? 5 : 4, invLevelSigma2, flags, nullptr, ptBestIdx, nFused);
    return rc ? rc : (KF->n_frames ? finish((Matcher *)h, flags) : 0);
}

int ygzorb_search_by_sim3(ygzorb_matcher *h, const ygzorb_frame_features *KF1, const ygzorb_frame_features *KF2,
                          const ygzorb_keyframe_points *P1, const ygzorb_keyframe_points *P2, float th, unsigned flags,
                          int32_t *matches12, int32_t *nFound) {
    Matcher *m = (Matcher *)h;
    if (!m || !KF1 || !KF2 || !P1 || !P2 || !matches12 || !nFound) return argFailM("null argument");
    if (KF1->n_frames != KF2->n_frames) return argFailM("the two keyframe batches differ in length");
    if (P1->pt_stride != KF1->kp_stride || P2->pt_stride != KF2->kp_stride)
        return argFailM("MapPoint arrays are indexed by keypoint: pt_stride must equal kp_stride");
    if (KF1->n_frames == 0) return 0;
    const bool outDev = flags & YGZORB_OUT_DEVICE;
    const unsigned inner = (flags & YGZORB_IN_DEVICE) | YGZORB_OUT_DEVICE;
    const size_t nf = KF1->n_frames;
    void *b1 = nullptr, *b2 = nullptr, *c1 = nullptr;
    int rc;
    if ((rc = scratch(m, 30, nf * P1->pt_stride * 4, &b1))) return rc;
    if ((rc = scratch(m, 31, nf * P2->pt_stride * 4, &b2))) return rc;
    if ((rc = scratch(m, 32, nf * 4 * 2, &c1))) return rc;
    const int32_t *dN1 = nullptr, *dN2 = nullptr;
    // KF1's MapPoints searched in KF2 (:1043-1110), then KF2's in KF1 (:1113-1180)
    if ((rc = searchByProjectionKeyFrame(h, KF2, P1, th, kThHigh, 0, 6, nullptr, inner, nullptr, (int32_t *)b1, (int32_t *)c1, 15, nullptr, &dN1))) return rc;
    // n_pts of the first direction must outlive the second call's staging: keep a copy
    void *n1copy = nullptr;
    if ((rc = scratch(m, 33, nf * 4, &n1copy))) return rc;
    YGZ_CUDA(cudaMemcpyAsync(n1copy, dN1, nf * 4, cudaMemcpyDeviceToDevice, m->stream));
    if ((rc = searchByProjectionKeyFrame(h, KF1, P2, th, kThHigh, 0, 6, nullptr, inner, nullptr, (int32_t *)b2, (int32_t *)c1 + nf, 15, nullptr, &dN2))) return rc;
    int32_t *dM12 = nullptr, *dFound = nullptr;
    if ((rc = stageOut(m, 34, matches12, nf * P1->pt_stride, outDev, &dM12))) return rc;
    if ((rc = stageOut(m, 35, nFound, nf, outDev, &dFound))) return rc;
    m->launches += launchSim3Agreement((const int32_t *)b1, (const int32_t *)b2, (const int32_t *)n1copy, dN2, P1->pt_stride,
                                       P2->pt_stride, KF1->n_frames, dM12, dFound, m->stream);
    if ((rc = copyOut(m, matches12, dM12, nf * P1->pt_stride, outDev))) return rc;
    if ((rc = copyOut(m, nFound, dFound, nf, outDev))) return rc;
    return finish(m, flags);
}

int ygzorb_search_for_initialization(ygzorb_matcher *h, const ygzorb_frame_features *F1, const ygzorb_frame_features *F2,
                                     float *prevMatched, int windowSize, float nnratio, int checkOrientation, unsigned flags,
                                     int32_t *matches12, int32_t *nMatches) {
    Matcher *m = (Matcher *)h;
    if (!m || !F1 || !F2 || !prevMatched || !matches12 || !nMatches) return argFailM("null argument");
    if (F1->n_frames < 0 || F1->n_frames != F2->n_frames || F1->kp_stride < 1 || F2->kp_stride < 1) return argFailM("bad sizes");
    if (!F1->n_kp || !F1->kps || !F1->desc || !F2->n_kp || !F2->kps || !F2->desc) return argFailM("null array in frame view");
    if (F1->n_frames == 0) return 0;
    YGZ_CUDA(cudaSetDevice(m->device));
    const bool inDev = flags & YGZORB_IN_DEVICE, outDev = flags & YGZORB_OUT_DEVICE;
    const size_t nf = F1->n_frames, nk1 = nf * F1->kp_stride, nk2 = nf * F2->kp_stride;
    InitArgs A;
    memset(&A, 0, sizeof A);
    A.nFrames = F1->n_frames; A.stride1 = F1->kp_stride; A.stride2 = F2->kp_stride;
    int rc;
    if ((rc = stageIn(m, 0, F1->n_kp, nf, inDev, &A.n1))) return rc;
    if ((rc = stageIn(m, 1, F2->n_kp, nf, inDev, &A.n2))) return rc;
    if ((rc = stageIn(m, 2, F1->kps, nk1, inDev, &A.kps1))) return rc;
    if ((rc = stageIn(m, 3, F1->desc, nk1 * 32, inDev, &A.desc1))) return rc;
    if ((rc = stageIn(m, 4, F2->kps, nk2, inDev, &A.kps2))) return rc;
    if ((rc = stageIn(m, 5, F2->desc, nk2 * 32, inDev, &A.desc2))) return rc;
    const float *dPrev = nullptr;
    if ((rc = stageIn(m, 6, (const float *)prevMatched, nk1 * 2, inDev, &dPrev))) return rc;
    A.prevMatched = (float *)dPrev;
    A.minX = F2->min_x; A.minY = F2->min_y;
    A.invW = (float)kGridCols / (F2->max_x - F2->min_x);
    A.invH = (float)kGridRows / (F2->max_y - F2->min_y);
    A.window = (float)windowSize; A.nnratio = nnratio; A.checkOrientation = checkOrientation;
    if ((rc = stageOut(m, 14, matches12, nk1, outDev, &A.matches12))) return rc;
    if ((rc = stageOut(m, 16, nMatches, nf, outDev, &A.nMatches))) return rc;
    void *p;
    if ((rc = scratch(m, 18, nk2 * 4, &p))) return rc;
    A.gridList = (int32_t *)p;
    if ((rc = scratch(m, 19, (5 * nk1 + 2 * nk2) * 4, &p))) return rc;
    A.state = (int32_t *)p;
    m->launches += launchSearchForInitialization(A, m->stream);
    if ((rc = copyOut(m, matches12, A.matches12, nk1, outDev))) return rc;
    if ((rc = copyOut(m, nMatches, A.nMatches, nf, outDev))) return rc;
    if (!inDev) YGZ_CUDA(cudaMemcpyAsync(prevMatched, A.prevMatched, nk1 * 2 * sizeof(float), cudaMemcpyDeviceToHost, m->stream));
    return finish(m, flags);
}

static int stageKf(Matcher *m, int slot0, const ygzorb_keyframe_features *K, bool inDev, KfView *V) {
    int rc;
    V->n = K->n; V->nNodes = K->n_nodes;
    if ((rc = stageIn(m, slot0 + 0, K->kps, (size_t)K->n, inDev, &V->kps))) return rc;
    if ((rc = stageIn(m, slot0 + 1, K->desc, (size_t)K->n * 32, inDev, &V->desc))) return rc;
    if ((rc = stageIn(m, slot0 + 2, K->u_right, (size_t)K->n, inDev, &V->uRight))) return rc;
    if ((rc = stageIn(m, slot0 + 3, K->has_mappoint, (size_t)K->n, inDev, &V->hasMP))) return rc;
    if ((rc = stageIn(m, slot0 + 4, K->node_ids, (size_t)K->n_nodes, inDev, &V->nodeIds))) return rc;
    if ((rc = stageIn(m, slot0 + 5, K->node_start, (size_t)K->n_nodes + 1, inDev, &V->nodeStart))) return rc;
    // every keypoint sits in at most one node: node_feats has at most n entries (its exact length is node_start[n_nodes])
    size_t nFeats = (size_t)K->n;
    if (!inDev && K->n_nodes > 0) nFeats = (size_t)K->node_start[K->n_nodes];
    if ((rc = stageIn(m, slot0 + 6, K->node_feats, nFeats, inDev, &V->nodeFeats))) return rc;
    return 0;
}

int ygzorb_search_for_triangulation(ygzorb_matcher *h, const ygzorb_keyframe_features *K1, const ygzorb_keyframe_features *K2,
                                    const float *f12, float ex, float ey, int onlyStereo, int checkOrientation,
                                    unsigned flags, int32_t *matches12, int32_t *nMatches) {
    Matcher *m = (Matcher *)h;
    if (!m || !K1 || !K2 || !f12 || !matches12 || !nMatches) return argFailM("null argument");
    if (K1->n < 0 || K2->n < 0 || K1->n_nodes < 0 || K2->n_nodes < 0) return argFailM("bad sizes");
    if (K2->nlevels < 1 || K2->nlevels > kMaxLevels || !K2->scale_factors || !K2->level_sigma2) return argFailM("kf2 scale tables missing");
    for (const ygzorb_keyframe_features *K : {K1, K2})
        if (K->n > 0 && (!K->kps || !K->desc || !K->u_right || !K->has_mappoint)) return argFailM("null keyframe array");
    YGZ_CUDA(cudaSetDevice(m->device));
    const bool inDev = flags & YGZORB_IN_DEVICE, outDev = flags & YGZORB_OUT_DEVICE;
    if (K1->n == 0) { if (outDev) YGZ_CUDA(cudaMemsetAsync(nMatches, 0, 4, m->stream)); else *nMatches = 0; return finish(m, flags); }
    TriArgs A;
    memset(&A, 0, sizeof A);
    int rc;
    if ((rc = stageKf(m, 0, K1, inDev, &A.kf1))) return rc;
    if ((rc = stageKf(m, 7, K2, inDev, &A.kf2))) return rc;
    for (int l = 0; l < K2->nlevels; l++) { A.scale2[l] = K2->scale_factors[l]; A.sigma2_2[l] = K2->level_sigma2[l]; }
    for (int i = 0; i < 9; i++) A.F12[i] = f12[i];
    A.ex = ex; A.ey = ey; A.onlyStereo = onlyStereo; A.checkOrientation = checkOrientation;
    if ((rc = stageOut(m, 14, matches12, (size_t)K1->n, outDev, &A.matches12))) return rc;
    if ((rc = stageOut(m, 15, nMatches, 1, outDev, &A.nMatches))) return rc;
    void *p;
    if ((rc = scratch(m, 16, (size_t)K1->n * 4, &p))) return rc;
    A.nodeOf1 = (int32_t *)p;
    m->launches += launchTriangulation(A, m->stream);
    if ((rc = copyOut(m, matches12, A.matches12, (size_t)K1->n, outDev))) return rc;
    if ((rc = copyOut(m, nMatches, A.nMatches, 1, outDev))) return rc;
    return finish(m, flags);
}

int ygzorb_stereo_match(ygzorb_matcher *h, ygzorb_extractor *left, ygzorb_extractor *right, int frame,
                        const ygzorb_keypoint *kpsL, const uint8_t *descL, int nL, const ygzorb_keypoint *kpsR,
                        const uint8_t *descR, int nR, float mbf, float mb, unsigned flags, float *uRight, float *depth) {
    Matcher *m = (Matcher *)h;
    if (!m || !left || !right || nL < 0 || nR < 0 || (nL > 0 && (!kpsL || !descL || !uRight || !depth)) ||
        (nR > 0 && (!kpsR || !descR))) return argFailM("bad argument");
    if (nL == 0) return 0;
    YGZ_CUDA(cudaSetDevice(m->device));
    const bool inDev = flags & YGZORB_IN_DEVICE, outDev = flags & YGZORB_OUT_DEVICE;
    StereoArgs A;
    memset(&A, 0, sizeof A);
    int nlevels = 0, nlevelsR = 0;
    const float *sc = nullptr, *isc = nullptr, *scR = nullptr, *iscR = nullptr;
    int rc;
    if ((rc = extractorTables(left, &nlevels, &sc, &isc))) return rc;
    if ((rc = extractorTables(right, &nlevelsR, &scR, &iscR))) return rc;
    if (nlevels != nlevelsR) return argFailM("left and right extractors differ in nlevels");
    A.nlevels = nlevels;
    float maxScale = 1.f;
    for (int l = 0; l < nlevels; l++) {
        LevelView vl, vr;
        if ((rc = extractorLevelView(left, frame, l, &vl))) return rc;    // drains the extractor's stream
        if ((rc = extractorLevelView(right, frame, l, &vr))) return rc;
        if (vl.w != vr.w || vl.h != vr.h) return argFailM("left and right pyramids differ in size");
        A.lvL[l] = vl.ptr; A.lvR[l] = vr.ptr; A.pitchL[l] = vl.pitch; A.pitchR[l] = vr.pitch; A.lw[l] = vl.w; A.lh[l] = vl.h;
        A.scale[l] = sc[l]; A.invScale[l] = isc[l];
        maxScale = std::max(maxScale, sc[l]);
    }
    A.nL = nL; A.nR = nR; A.mbf = mbf; A.mb = mb;
    if ((rc = stageIn(m, 0, kpsL, (size_t)nL, inDev, &A.kpsL))) return rc;
    if ((rc = stageIn(m, 1, descL, (size_t)nL * 32, inDev, &A.descL))) return rc;
    if (nR > 0) {
        if ((rc = stageIn(m, 2, kpsR, (size_t)nR, inDev, &A.kpsR))) return rc;
        if ((rc = stageIn(m, 3, descR, (size_t)nR * 32, inDev, &A.descR))) return rc;
    }
    if ((rc = stageOut(m, 4, uRight, (size_t)nL, outDev, &A.uRight))) return rc;
    if ((rc = stageOut(m, 5, depth, (size_t)nL, outDev, &A.depth))) return rc;
    void *p;
    if ((rc = scratch(m, 6, (size_t)nL * 4, &p))) return rc;
    A.bestDist = (int32_t *)p;
    if ((rc = scratch(m, 7, ((size_t)A.lh[0] + 2) * 4, &p))) return rc;
    A.rowStart = (int32_t *)p;
    A.rowListCap = std::max(nR, 1) * ((int)std::ceil(4.f * maxScale) + 4);   // rows floor(y-2s)..ceil(y+2s) per right keypoint
    if ((rc = scratch(m, 8, (size_t)A.rowListCap * 4, &p))) return rc;
    A.rowList = (int32_t *)p;
    if ((rc = scratch(m, 9, (size_t)nL * 4, &p))) return rc;
    A.sortBuf = (int32_t *)p;
    A.status = m->dStatus;
    m->launches += launchStereo(A, m->stream);
    if ((rc = copyOut(m, uRight, A.uRight, (size_t)nL, outDev))) return rc;
    if ((rc = copyOut(m, depth, A.depth, (size_t)nL, outDev))) return rc;
    return finish(m, flags);
}

int ygzorb_search_by_bow(ygzorb_matcher *h, const ygzorb_keyframe_features *K1, const ygzorb_keyframe_features *K2,
                         int secondIsFrame, float nnratio, int checkOrientation, unsigned flags, int32_t *matches,
                         int32_t *nMatches) {
    Matcher *m = (Matcher *)h;
    if (!m || !K1 || !K2 || !matches || !nMatches) return argFailM("null argument");
    if (K1->n < 0 || K2->n < 0 || K1->n_nodes < 0 || K2->n_nodes < 0) return argFailM("bad sizes");
    if (K1->n > 0 && (!K1->kps || !K1->desc || !K1->has_mappoint)) return argFailM("null keyframe array");
    if (K2->n > 0 && (!K2->kps || !K2->desc || (!secondIsFrame && !K2->has_mappoint))) return argFailM("null keyframe array");
    for (const ygzorb_keyframe_features *K : {K1, K2})
        if (K->n_nodes > 0 && (!K->node_ids || !K->node_start || !K->node_feats)) return argFailM("null FeatureVector array");
    YGZ_CUDA(cudaSetDevice(m->device));
    const bool inDev = flags & YGZORB_IN_DEVICE, outDev = flags & YGZORB_OUT_DEVICE;
    const size_t nOut = (size_t)(secondIsFrame ? K2->n : K1->n);
    TriArgs A;
    memset(&A, 0, sizeof A);
    int rc;
    if ((rc = stageKf(m, 0, K1, inDev, &A.kf1))) return rc;
    if ((rc = stageKf(m, 7, K2, inDev, &A.kf2))) return rc;
    A.onlyStereo = secondIsFrame ? 1 : 0; A.checkOrientation = checkOrientation; A.nnratio = nnratio;
    if ((rc = stageOut(m, 14, matches, nOut, outDev, &A.matches12))) return rc;
    if ((rc = stageOut(m, 15, nMatches, 1, outDev, &A.nMatches))) return rc;
    void *p;
    if ((rc = scratch(m, 16, (size_t)std::max(K2->n, 1) * 4, &p))) return rc;
    A.nodeOf1 = (int32_t *)p;      // "already matched" flags of side 2
    m->launches += launchSearchByBoW(A, m->stream);
    if ((rc = copyOut(m, matches, A.matches12, nOut, outDev))) return rc;
    if ((rc = copyOut(m, nMatches, A.nMatches, 1, outDev))) return rc;
    return finish(m, flags);
}

// ================================================================================================
// SURVEY.md 8(f) row f4
int ygzorb_distinctive_descriptors(ygzorb_matcher *h, const uint8_t *obsDesc, const int32_t *obsStart, int nPoints, unsigned flags,
                                   int32_t *bestIdx, int32_t *bestMedian, uint8_t *bestDesc) {
    Matcher *m = (Matcher *)h;
    if (!m || !obsStart || !bestIdx || !bestMedian || !bestDesc) return argFailM("null argument");
    if (nPoints < 0) return argFailM("bad sizes");
    if (nPoints == 0) return 0;
    YGZ_CUDA(cudaSetDevice(m->device));
    const bool inDev = flags & YGZORB_IN_DEVICE, outDev = flags & YGZORB_OUT_DEVICE;
    long long nRows = -1;
    if (!inDev) {
        for (int p = 0; p < nPoints; p++) if (obsStart[p + 1] < obsStart[p]) return argFailM("obs_start must be non-decreasing");
        if (obsStart[0] < 0) return argFailM("obs_start must be non-negative");
        nRows = obsStart[nPoints];
        if (nRows > 0 && !obsDesc) return argFailM("null argument");
    }
    int rc;
    const uint8_t *dDesc = obsDesc; const int32_t *dStart = nullptr;
    if (!inDev && (rc = stageIn(m, 0, obsDesc, (size_t)std::max<long long>(nRows, 0) * 32, false, &dDesc))) return rc;
    if ((rc = stageIn(m, 1, obsStart, (size_t)nPoints + 1, inDev, &dStart))) return rc;
    int32_t *oi, *om; uint8_t *od;
    if ((rc = stageOut(m, 2, bestIdx, (size_t)nPoints, outDev, &oi))) return rc;
    if ((rc = stageOut(m, 3, bestMedian, (size_t)nPoints, outDev, &om))) return rc;
    if ((rc = stageOut(m, 4, bestDesc, (size_t)nPoints * 32, outDev, &od))) return rc;
    // rows of points without observations stay as the caller left them: seed the staged copy with the caller's bytes
    if (!outDev) YGZ_CUDA(cudaMemcpyAsync(od, bestDesc, (size_t)nPoints * 32, cudaMemcpyHostToDevice, m->stream));
    m->launches += launchDistinctive(dDesc, dStart, nPoints, oi, om, od, m->stream);
    if ((rc = copyOut(m, bestIdx, oi, (size_t)nPoints, outDev))) return rc;
    if ((rc = copyOut(m, bestMedian, om, (size_t)nPoints, outDev))) return rc;
    if ((rc = copyOut(m, bestDesc, od, (size_t)nPoints * 32, outDev))) return rc;
    return finish(m, flags);
}

namespace ygz {
struct Pyramid {
    int device = 0, nLevels = 0;
    uint8_t *base = nullptr;
    int w[kMaxLevels], h[kMaxLevels];
    size_t off[kMaxLevels];
};
static int pyramidAlloc(int device, const int32_t *w, const int32_t *h, int nLevels, Pyramid **out) {
    if (nLevels < 1 || nLevels > kMaxLevels) return argFailM("n_levels must be 1..16");
    Pyramid *P = new Pyramid();
    P->device = device; P->nLevels = nLevels;
    size_t total = 0;
    for (int l = 0; l < nLevels; l++) {
        if (w[l] < 1 || h[l] < 1) { delete P; return argFailM("bad level size"); }
        P->w[l] = w[l]; P->h[l] = h[l]; P->off[l] = total;
        total += ((size_t)w[l] * h[l] + 255) & ~(size_t)255;
    }
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaMalloc((void **)&P->base, total);
    if (e != cudaSuccess) { delete P; return cudaFail(e, "pyramid_create", __FILE__, __LINE__); }
    *out = P;
    return 0;
}
}  // namespace ygz

int ygzorb_pyramid_create(int device, const uint8_t *const *levels, const int32_t *w, const int32_t *h, const int32_t *pitch,
                          int nLevels, unsigned flags, ygzorb_pyramid **out) {
    if (!levels || !w || !h || !pitch || !out) return argFailM("null argument");
    for (int l = 0; l < nLevels && l < kMaxLevels; l++)
        if (!levels[l] || pitch[l] < w[l]) return argFailM("null level or pitch < width");
    Pyramid *P = nullptr;
    int rc = pyramidAlloc(device, w, h, nLevels, &P);
    if (rc) return rc;
    const cudaMemcpyKind kind = (flags & YGZORB_IN_DEVICE) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    for (int l = 0; l < nLevels; l++) {
        cudaError_t e = cudaMemcpy2D(P->base + P->off[l], (size_t)w[l], levels[l], (size_t)pitch[l], (size_t)w[l], (size_t)h[l], kind);
        if (e != cudaSuccess) { ygzorb_pyramid_destroy((ygzorb_pyramid *)P); return cudaFail(e, "pyramid_create copy", __FILE__, __LINE__); }
    }
    *out = (ygzorb_pyramid *)P;
    return 0;
}

int ygzorb_pyramid_from_extractor(ygzorb_extractor *ex, int frame, ygzorb_pyramid **out) {
    if (!ex || !out) return argFailM("null argument");
    int nl = 0; const float *sc, *isc;
    int rc = extractorTables(ex, &nl, &sc, &isc);
    if (rc) return rc;
    int32_t w[kMaxLevels], h[kMaxLevels];
    LevelView v[kMaxLevels];
    for (int l = 0; l < nl; l++) {
        if ((rc = extractorLevelView(ex, frame, l, &v[l]))) return rc;   // drains the extractor's stream
        w[l] = v[l].w; h[l] = v[l].h;
    }
    int device = 0;
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, v[0].ptr) == cudaSuccess) device = attr.device;
    Pyramid *P = nullptr;
    if ((rc = pyramidAlloc(device, w, h, nl, &P))) return rc;
    for (int l = 0; l < nl; l++) {
        cudaError_t e = cudaMemcpy2D(P->base + P->off[l], (size_t)w[l], v[l].ptr, (size_t)v[l].pitch, (size_t)w[l], (size_t)h[l], cudaMemcpyDeviceToDevice);
        if (e != cudaSuccess) { ygzorb_pyramid_destroy((ygzorb_pyramid *)P); return cudaFail(e, "pyramid_from_extractor copy", __FILE__, __LINE__); }
    }
    *out = (ygzorb_pyramid *)P;
    return 0;
}

void ygzorb_pyramid_destroy(ygzorb_pyramid *p) {
    Pyramid *P = (Pyramid *)p;
    if (!P) return;
    cudaSetDevice(P->device);
    if (P->base) cudaFree(P->base);
    delete P;
}

int ygzorb_pyramid_info(const ygzorb_pyramid *p, int32_t *nLevels, int32_t *w, int32_t *h) {
    const Pyramid *P = (const Pyramid *)p;
    if (!P || !nLevels) return argFailM("null argument");
    *nLevels = P->nLevels;
    for (int l = 0; l < P->nLevels; l++) { if (w) w[l] = P->w[l]; if (h) h[l] = P->h[l]; }
    return 0;
}

int ygzorb_pyramid_get_level(const ygzorb_pyramid *p, int level, uint8_t *out) {
    const Pyramid *P = (const Pyramid *)p;
    if (!P || !out) return argFailM("null argument");
    if (level < 0 || level >= P->nLevels) return argFailM("level out of range");
    YGZ_CUDA(cudaSetDevice(P->device));
    YGZ_CUDA(cudaMemcpy(out, P->base + P->off[level], (size_t)P->w[level] * P->h[level], cudaMemcpyDeviceToHost));
    return 0;
}

int ygzorb_find_direct_projection(ygzorb_matcher *h, const ygzorb_direct_keyframe *kfs, int nKfs, const ygzorb_pyramid *cur,
                                  const float *curIntr, const float *scale, const float *invScale, float invLevelSigma2_1, int n,
                                  const int32_t *kfIndex, const float *pxRef, const int32_t *octave, const float *worldPos,
                                  unsigned flags, float *pxCurr, int32_t *searchLevel, uint8_t *success) {
    Matcher *m = (Matcher *)h;
    const Pyramid *C = (const Pyramid *)cur;
    if (!m || !C || !curIntr || !scale || !invScale) return argFailM("null argument");
    if (n < 0 || nKfs < 0) return argFailM("bad sizes");
    if (n == 0) return 0;
    if (!kfs || nKfs == 0 || !kfIndex || !pxRef || !octave || !worldPos || !pxCurr || !searchLevel || !success) return argFailM("null argument");
    if (C->device != m->device) return argFailM("pyramid lives on another device");
    const bool inDev = flags & YGZORB_IN_DEVICE, outDev = flags & YGZORB_OUT_DEVICE;
    std::vector<DirectKf> hk((size_t)nKfs);
    for (int k = 0; k < nKfs; k++) {
        const Pyramid *P = (const Pyramid *)kfs[k].pyramid;
        if (!P) return argFailM("keyframe without a pyramid");
        if (P->device != m->device) return argFailM("pyramid lives on another device");
        DirectKf &D = hk[k];
        memset(&D, 0, sizeof D);
        for (int l = 0; l < P->nLevels; l++) { D.lv[l] = P->base + P->off[l]; D.w[l] = P->w[l]; D.h[l] = P->h[l]; D.pitch[l] = P->w[l]; }
        memcpy(D.poseRef, kfs[k].pose_ref, sizeof D.poseRef); memcpy(D.tCurRef, kfs[k].t_cur_ref, sizeof D.tCurRef);
        D.fx = kfs[k].fx; D.fy = kfs[k].fy; D.cx = kfs[k].cx; D.cy = kfs[k].cy;
    }
    if (!inDev)
        for (int i = 0; i < n; i++) {
            if (kfIndex[i] < 0 || kfIndex[i] >= nKfs) return argFailM("kf_index out of range");
            const Pyramid *P = (const Pyramid *)kfs[kfIndex[i]].pyramid;
            if (octave[i] < 0 || octave[i] >= P->nLevels || octave[i] >= C->nLevels) return argFailM("octave out of range");
        }
    YGZ_CUDA(cudaSetDevice(m->device));
    DirectArgs A;
    memset(&A, 0, sizeof A);
    int rc;
    const DirectKf *dk = nullptr;
    if ((rc = stageIn(m, 0, hk.data(), (size_t)nKfs, false, &dk))) return rc;
    YGZ_CUDA(cudaStreamSynchronize(m->stream));   // hk is a local: the staged copy must be complete before it goes away
    A.kfs = dk;
    for (int l = 0; l < C->nLevels; l++) {
        A.curLv[l] = C->base + C->off[l]; A.curW[l] = C->w[l]; A.curH[l] = C->h[l]; A.curPitch[l] = C->w[l];
        A.scale[l] = scale[l]; A.invScale[l] = invScale[l];
    }
    A.nLevels = C->nLevels; A.n = n;
    A.fx = curIntr[0]; A.fy = curIntr[1]; A.cx = curIntr[2]; A.cy = curIntr[3];
    A.invLevelSigma2_1 = invLevelSigma2_1;
    if ((rc = stageIn(m, 1, kfIndex, (size_t)n, inDev, &A.kfIndex))) return rc;
    if ((rc = stageIn(m, 2, pxRef, (size_t)2 * n, inDev, &A.pxRef))) return rc;
    if ((rc = stageIn(m, 3, octave, (size_t)n, inDev, &A.octave))) return rc;
    if ((rc = stageIn(m, 4, worldPos, (size_t)3 * n, inDev, &A.worldPos))) return rc;
    if ((rc = stageOut(m, 5, pxCurr, (size_t)2 * n, outDev, &A.pxCurr))) return rc;
    if (!outDev) YGZ_CUDA(cudaMemcpyAsync(A.pxCurr, pxCurr, (size_t)2 * n * sizeof(float), cudaMemcpyHostToDevice, m->stream));
    if ((rc = stageOut(m, 6, searchLevel, (size_t)n, outDev, &A.searchLevel))) return rc;
    if ((rc = stageOut(m, 7, success, (size_t)n, outDev, &A.success))) return rc;
    m->launches += launchFindDirectProjection(A, m->stream);
    if ((rc = copyOut(m, pxCurr, (const float *)A.pxCurr, (size_t)2 * n, outDev))) return rc;
    if ((rc = copyOut(m, searchLevel, (const int32_t *)A.searchLevel, (size_t)n, outDev))) return rc;
    if ((rc = copyOut(m, success, (const uint8_t *)A.success, (size_t)n, outDev))) return rc;
    return finish(m, flags);
}

namespace ygz {
struct SparseAlign {
    Matcher *m = nullptr;
    SparseAlignArgs A;
    int cap = 0, nLevels = 0;
    float *dKp = nullptr, *dWorld = nullptr, *dPose = nullptr, *dOut = nullptr, *dPatch = nullptr, *dJac = nullptr;
    uint8_t *dValid = nullptr, *dVisible = nullptr;
    float *hOut = nullptr;   // pinned
    bool begun = false;
};
static void sparseAlignFree(SparseAlign *S) {
    for (void *p : {(void *)S->dKp, (void *)S->dWorld, (void *)S->dPatch, (void *)S->dJac, (void *)S->dValid, (void *)S->dVisible})
        if (p) cudaFree(p);
    S->dKp = S->dWorld = S->dPatch = S->dJac = nullptr; S->dValid = S->dVisible = nullptr; S->cap = 0;
}
}  // namespace ygz

int ygzorb_sparse_align_create(ygzorb_matcher *h, ygzorb_sparse_align **out) {
    Matcher *m = (Matcher *)h;
    if (!m || !out) return argFailM("null argument");
    YGZ_CUDA(cudaSetDevice(m->device));
    SparseAlign *S = new SparseAlign();
    S->m = m;
    memset(&S->A, 0, sizeof S->A);
    cudaError_t e = cudaMalloc((void **)&S->dPose, 14 * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc((void **)&S->dOut, 64 * sizeof(float));
    if (e == cudaSuccess) e = cudaMallocHost((void **)&S->hOut, 64 * sizeof(float));
    if (e != cudaSuccess) { ygzorb_sparse_align_destroy((ygzorb_sparse_align *)S); return cudaFail(e, "sparse_align_create", __FILE__, __LINE__); }
    *out = (ygzorb_sparse_align *)S;
    return 0;
}

void ygzorb_sparse_align_destroy(ygzorb_sparse_align *sa) {
    SparseAlign *S = (SparseAlign *)sa;
    if (!S) return;
    cudaSetDevice(S->m->device);
    cudaStreamSynchronize(S->m->stream);
    sparseAlignFree(S);
    if (S->dPose) cudaFree(S->dPose);
    if (S->dOut) cudaFree(S->dOut);
    if (S->hOut) cudaFreeHost(S->hOut);
    delete S;
}

int ygzorb_sparse_align_begin(ygzorb_sparse_align *sa, const ygzorb_pyramid *ref, const ygzorb_pyramid *cur, int n, const float *kpXY,
                              const uint8_t *valid, const float *worldPos, const float *refPose, float focal, const float *curIntr,
                              const float *invScale) {
    SparseAlign *S = (SparseAlign *)sa;
    const Pyramid *R = (const Pyramid *)ref, *C = (const Pyramid *)cur;
    if (!S || !R || !C || !refPose || !curIntr || !invScale) return argFailM("null argument");
    if (n < 0 || (n > 0 && (!kpXY || !valid || !worldPos))) return argFailM("bad feature arrays");
    if (R->nLevels != C->nLevels) return argFailM("reference and current pyramids differ in depth");
    for (int l = 0; l < R->nLevels; l++)
        if (R->w[l] != C->w[l] || R->h[l] != C->h[l]) return argFailM("reference and current pyramids differ in size");
    Matcher *m = S->m;
    if (R->device != m->device || C->device != m->device) return argFailM("pyramid lives on another device");
    YGZ_CUDA(cudaSetDevice(m->device));
    if (n > S->cap) {
        YGZ_CUDA(cudaStreamSynchronize(m->stream));
        sparseAlignFree(S);
        const size_t cap = (((size_t)n + n / 4 + 64) + 31) & ~(size_t)31;
        YGZ_CUDA(cudaMalloc((void **)&S->dKp, cap * 2 * sizeof(float)));
        YGZ_CUDA(cudaMalloc((void **)&S->dWorld, cap * 3 * sizeof(float)));
        YGZ_CUDA(cudaMalloc((void **)&S->dPatch, cap * 16 * sizeof(float)));
        YGZ_CUDA(cudaMalloc((void **)&S->dJac, cap * 96 * sizeof(float)));
        YGZ_CUDA(cudaMalloc((void **)&S->dValid, cap));
        YGZ_CUDA(cudaMalloc((void **)&S->dVisible, cap));
        S->cap = (int)cap;
    }
    SparseAlignArgs &A = S->A;
    memset(&A, 0, sizeof A);
    for (int l = 0; l < R->nLevels; l++) {
        A.refLv[l] = R->base + R->off[l]; A.curLv[l] = C->base + C->off[l];
        A.w[l] = R->w[l]; A.h[l] = R->h[l]; A.refPitch[l] = R->w[l]; A.curPitch[l] = C->w[l];
        A.invScale[l] = invScale[l];
    }
    S->nLevels = R->nLevels;
    A.focal = focal; A.fx = curIntr[0]; A.fy = curIntr[1]; A.cx = curIntr[2]; A.cy = curIntr[3];
    A.n = n; A.stride = S->cap;
    A.kpXY = S->dKp; A.worldPos = S->dWorld; A.valid = S->dValid; A.refPose = S->dPose; A.tCurRef = S->dPose + 7;
    A.refPatch = S->dPatch; A.jac = S->dJac; A.visible = S->dVisible;
    if (n > 0) {
        YGZ_CUDA(cudaMemcpyAsync(S->dKp, kpXY, (size_t)n * 2 * sizeof(float), cudaMemcpyHostToDevice, m->stream));
        YGZ_CUDA(cudaMemcpyAsync(S->dWorld, worldPos, (size_t)n * 3 * sizeof(float), cudaMemcpyHostToDevice, m->stream));
        YGZ_CUDA(cudaMemcpyAsync(S->dValid, valid, (size_t)n, cudaMemcpyHostToDevice, m->stream));
        YGZ_CUDA(cudaMemsetAsync(S->dVisible, 0, (size_t)n, m->stream));
        YGZ_CUDA(cudaMemsetAsync(S->dPatch, 0, (size_t)S->cap * 16 * sizeof(float), m->stream));
    }
    YGZ_CUDA(cudaMemcpyAsync(S->dPose, refPose, 7 * sizeof(float), cudaMemcpyHostToDevice, m->stream));
    YGZ_CUDA(cudaStreamSynchronize(m->stream));   // the caller's arrays may go away after the call
    S->begun = true;
    return 0;
}

static int sparseAlignLevelOk(SparseAlign *S, int level) {
    if (!S) return argFailM("null argument");
    if (!S->begun) { setError("ygzorb_sparse_align_begin has not been called"); return YGZORB_E_STATE; }
    if (level < 0 || level >= S->nLevels) return argFailM("level out of range");
    return 0;
}

int ygzorb_sparse_align_run(ygzorb_sparse_align *sa, int maxLevel, int minLevel, float *tCurRef, int64_t *nTracked, float *stats) {
    SparseAlign *S = (SparseAlign *)sa;
    int rc;
    if ((rc = sparseAlignLevelOk(S, maxLevel)) || (rc = sparseAlignLevelOk(S, minLevel))) return rc;
    if (!tCurRef || !nTracked) return argFailM("null argument");
    Matcher *m = S->m;
    YGZ_CUDA(cudaSetDevice(m->device));
    if (S->A.n == 0) { *nTracked = 0; if (stats) stats[0] = stats[1] = 0; return 0; }   // "no features to track" (:24-27)
    YGZ_CUDA(cudaMemcpyAsync(S->dPose + 7, tCurRef, 7 * sizeof(float), cudaMemcpyHostToDevice, m->stream));
    m->launches += launchSparseAlignRun(S->A, maxLevel, minLevel, S->dOut, m->stream);
    YGZ_CUDA(cudaMemcpyAsync(S->hOut, S->dOut, 10 * sizeof(float), cudaMemcpyDeviceToHost, m->stream));
    YGZ_CUDA(cudaStreamSynchronize(m->stream));
    YGZ_CUDA(cudaGetLastError());
    memcpy(tCurRef, S->hOut, 7 * sizeof(float));
    *nTracked = (int64_t)S->hOut[7] / 16;
    if (stats) { stats[0] = S->hOut[8]; stats[1] = S->hOut[9]; }
    return 0;
}

int ygzorb_sparse_align_precompute(ygzorb_sparse_align *sa, int level) {
    SparseAlign *S = (SparseAlign *)sa;
    int rc;
    if ((rc = sparseAlignLevelOk(S, level))) return rc;
    YGZ_CUDA(cudaSetDevice(S->m->device));
    if (S->A.n > 0) S->m->launches += launchSparseAlignPrecompute(S->A, level, S->m->stream);
    YGZ_CUDA(cudaGetLastError());
    return 0;
}

int ygzorb_sparse_align_residuals(ygzorb_sparse_align *sa, int level, const float *tCurRef, float *h36, float *jres6, float *chi2Sum,
                                  int64_t *nMeas) {
    SparseAlign *S = (SparseAlign *)sa;
    int rc;
    if ((rc = sparseAlignLevelOk(S, level))) return rc;
    if (!tCurRef || !h36 || !jres6 || !chi2Sum || !nMeas) return argFailM("null argument");
    Matcher *m = S->m;
    YGZ_CUDA(cudaSetDevice(m->device));
    if (S->A.n == 0) { memset(h36, 0, 36 * sizeof(float)); memset(jres6, 0, 6 * sizeof(float)); *chi2Sum = 0; *nMeas = 0; return 0; }
    YGZ_CUDA(cudaMemcpyAsync(S->dPose + 7, tCurRef, 7 * sizeof(float), cudaMemcpyHostToDevice, m->stream));
    m->launches += launchSparseAlignResiduals(S->A, level, S->dOut, m->stream);
    YGZ_CUDA(cudaMemcpyAsync(S->hOut, S->dOut, 44 * sizeof(float), cudaMemcpyDeviceToHost, m->stream));
    YGZ_CUDA(cudaStreamSynchronize(m->stream));
    YGZ_CUDA(cudaGetLastError());
    memcpy(h36, S->hOut, 36 * sizeof(float)); memcpy(jres6, S->hOut + 36, 6 * sizeof(float));
    *chi2Sum = S->hOut[42]; *nMeas = (int64_t)S->hOut[43];
    return 0;
}

int ygzorb_sparse_align_caches(ygzorb_sparse_align *sa, uint8_t *visible, float *refPatch, float *jac) {
    SparseAlign *S = (SparseAlign *)sa;
    if (!S) return argFailM("null argument");
    if (!S->begun) { setError("ygzorb_sparse_align_begin has not been called"); return YGZORB_E_STATE; }
    YGZ_CUDA(cudaSetDevice(S->m->device));
    YGZ_CUDA(cudaStreamSynchronize(S->m->stream));
    const size_t n = (size_t)S->A.n;
    if (n == 0) return 0;
    if (visible) YGZ_CUDA(cudaMemcpy(visible, S->dVisible, n, cudaMemcpyDeviceToHost));
    // device layout is feature-minor (element e of feature i at e * cap + i); the caller gets the reference's [feature][element]
    const size_t cap = (size_t)S->cap;
    std::vector<float> tmp;
    if (refPatch) {
        tmp.resize(cap * 16);
        YGZ_CUDA(cudaMemcpy(tmp.data(), S->dPatch, cap * 16 * sizeof(float), cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < n; i++) for (size_t e = 0; e < 16; e++) refPatch[i * 16 + e] = tmp[e * cap + i];
    }
    if (jac) {
        tmp.resize(cap * 96);
        YGZ_CUDA(cudaMemcpy(tmp.data(), S->dJac, cap * 96 * sizeof(float), cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < n; i++) for (size_t e = 0; e < 96; e++) jac[i * 96 + e] = tmp[e * cap + i];
    }
    return 0;
}

// ================================================================================================
// Bag of words
}  // extern "C"

namespace ygz {

struct Vocabulary {
    int device = 0, k = 0, L = 0, scoring = 0, weighting = 0, nWords = 0;
    // host copy (tree as flat arrays)
    std::vector<int32_t> childStart, childList, wordId;
    std::vector<double> weight;
    std::vector<uint8_t> desc;
    // device copy
    int32_t *dChildStart = nullptr, *dChildList = nullptr, *dWordId = nullptr;
    double *dWeight = nullptr;
    uint8_t *dChildDesc = nullptr;
};

static int vocUpload(Vocabulary *v) {
    const size_t n = v->wordId.size(), nc = v->childList.size();
    YGZ_CUDA(cudaSetDevice(v->device));
    std::vector<uint8_t> cdesc(std::max<size_t>(nc, 1) * 32);
    for (size_t c = 0; c < nc; c++) memcpy(&cdesc[c * 32], &v->desc[(size_t)v->childList[c] * 32], 32);
    YGZ_CUDA(cudaMalloc((void **)&v->dChildStart, (n + 1) * 4));
    YGZ_CUDA(cudaMalloc((void **)&v->dChildList, std::max<size_t>(nc, 1) * 4));
    YGZ_CUDA(cudaMalloc((void **)&v->dWordId, n * 4));
    YGZ_CUDA(cudaMalloc((void **)&v->dWeight, n * 8));
    YGZ_CUDA(cudaMalloc((void **)&v->dChildDesc, cdesc.size()));
    YGZ_CUDA(cudaMemcpy(v->dChildStart, v->childStart.data(), (n + 1) * 4, cudaMemcpyHostToDevice));
    if (nc) YGZ_CUDA(cudaMemcpy(v->dChildList, v->childList.data(), nc * 4, cudaMemcpyHostToDevice));
    YGZ_CUDA(cudaMemcpy(v->dWordId, v->wordId.data(), n * 4, cudaMemcpyHostToDevice));
    YGZ_CUDA(cudaMemcpy(v->dWeight, v->weight.data(), n * 8, cudaMemcpyHostToDevice));
    YGZ_CUDA(cudaMemcpy(v->dChildDesc, cdesc.data(), cdesc.size(), cudaMemcpyHostToDevice));
    return 0;
}

// Validates the tree (every child id in range, reachable nodes form a tree of depth <= 32) and finishes the handle.
static int vocFinish(Vocabulary *v, ygzorb_vocabulary **out) {
    const int n = (int)v->wordId.size();
    if (n < 1 || (int)v->childStart.size() != n + 1) { delete v; return argFailM("vocabulary: bad array sizes"); }
    for (int i = 0; i < n; i++)
        if (v->childStart[i] > v->childStart[i + 1] || v->childStart[i] < 0) { delete v; return argFailM("vocabulary: child_start is not monotone"); }
    if (v->childStart[n] != (int)v->childList.size()) { delete v; return argFailM("vocabulary: child_start[n_nodes] != length of child_list"); }
    for (int32_t c : v->childList)
        if (c <= 0 || c >= n) { delete v; return argFailM("vocabulary: child id out of range"); }
    // depth check along every root-to-leaf path (a cycle would loop the descent kernel forever)
    std::vector<int> depth(n, -1);
    std::vector<int> stack{0};
    depth[0] = 0;
    while (!stack.empty()) {
        const int i = stack.back(); stack.pop_back();
        for (int c = v->childStart[i]; c < v->childStart[i + 1]; c++) {
            const int ch = v->childList[c];
            if (depth[ch] >= 0 || depth[i] >= 32) { delete v; return argFailM("vocabulary: not a tree (shared child, cycle or depth > 32)"); }
            depth[ch] = depth[i] + 1;
            stack.push_back(ch);
        }
    }
    v->nWords = 0;
    for (int i = 0; i < n; i++) if (v->wordId[i] >= 0) v->nWords = std::max(v->nWords, v->wordId[i] + 1);
    const int rc = vocUpload(v);
    if (rc) { ygzorb_vocabulary_destroy((ygzorb_vocabulary *)v); return rc; }
    *out = (ygzorb_vocabulary *)v;
    return 0;
}

// children as vectors -> CSR
static void vocFromChildren(Vocabulary *v, const std::vector<std::vector<int32_t>> &children) {
    const size_t n = children.size();
    v->childStart.assign(n + 1, 0);
    v->childList.clear();
    for (size_t i = 0; i < n; i++) {
        v->childStart[i] = (int32_t)v->childList.size();
        v->childList.insert(v->childList.end(), children[i].begin(), children[i].end());
    }
    v->childStart[n] = (int32_t)v->childList.size();
}

}  // namespace ygz

extern "C" {

int ygzorb_vocabulary_create(int nNodes, int k, int L, const int32_t *childStart, const int32_t *childList, const uint8_t *desc,
                             const int32_t *wordId, const double *weight, int device, ygzorb_vocabulary **out) {
    if (!out || !childStart || !desc || !wordId || !weight || nNodes < 1) return argFailM("null argument");
    if (L < 0 || L > 32) return argFailM("vocabulary: L must be in [0, 32]");
    Vocabulary *v = new Vocabulary();
    v->device = device; v->k = k; v->L = L;
    v->childStart.assign(childStart, childStart + nNodes + 1);
    const int nc = childStart[nNodes];
    if (nc < 0 || (nc > 0 && !childList)) { delete v; return argFailM("vocabulary: bad child list"); }
    v->childList.assign(childList, childList + nc);
    v->desc.assign(desc, desc + (size_t)nNodes * 32);
    v->wordId.assign(wordId, wordId + nNodes);
    v->weight.assign(weight, weight + nNodes);
    return vocFinish(v, out);
}

int ygzorb_vocabulary_load_binary(const char *path, int device, ygzorb_vocabulary **out) {
    if (!path || !out) return argFailM("null argument");
    FILE *f = fopen(path, "rb");
    if (!f) return argFailM("vocabulary: cannot open file");
    uint32_t nbNodes = 0, sizeNode = 0;
    int32_t hdr[4] = {0, 0, 0, 0};   // m_k, m_L, m_scoring, m_weighting (:1483-1486)
    if (fread(&nbNodes, 4, 1, f) != 1 || fread(&sizeNode, 4, 1, f) != 1 || fread(hdr, 4, 4, f) != 4 || sizeNode < 41 ||
        nbNodes > (1u << 28)) { fclose(f); return argFailM("vocabulary: not a binary vocabulary file"); }
    Vocabulary *v = new Vocabulary();
    v->device = device; v->k = hdr[0]; v->L = hdr[1]; v->scoring = hdr[2]; v->weighting = hdr[3];
    const size_t n = (size_t)nbNodes + 1;                  // m_nodes.resize(nb_nodes + 1) (:1492)
    std::vector<std::vector<int32_t>> children(n);
    v->wordId.assign(n, -1); v->weight.assign(n, 0.0); v->desc.assign(n * 32, 0);
    std::vector<char> buf(sizeNode, 0);
    size_t nid = 1;
    int nWords = 0;
    bool eof = false, bad = false;
    while (!eof && nid < n) {                              // while (!f.eof()) { f.read(buf, size_node); ... } (:1496-1515)
        if (fread(buf.data(), 1, sizeNode, f) < sizeNode) eof = true;   // the record of the previous pass stays in buf
        int32_t par; memcpy(&par, buf.data(), 4);
        if (par < 0 || (size_t)par >= n || (size_t)par == nid) { bad = true; break; }
        children[par].push_back((int32_t)nid);
        memcpy(&v->desc[nid * 32], buf.data() + 4, 32);
        float w; memcpy(&w, buf.data() + 36, 4);
        v->weight[nid] = w;
        if (buf[40]) v->wordId[nid] = nWords++;
        nid++;
    }
    fclose(f);
    if (bad) { delete v; return argFailM("vocabulary: parent id out of range"); }
    vocFromChildren(v, children);
    return vocFinish(v, out);
}

int ygzorb_vocabulary_load_text(const char *path, int device, ygzorb_vocabulary **out) {
    if (!path || !out) return argFailM("null argument");
    std::ifstream f(path);
    if (!f.is_open()) return argFailM("vocabulary: cannot open file");
    std::string line;
    std::getline(f, line);
    std::stringstream ss(line);
    int k = -1, L = -1, n1 = -1, n2 = -1;
    ss >> k >> L >> n1 >> n2;
    if (k < 0 || k > 20 || L < 1 || L > 10 || n1 < 0 || n1 > 5 || n2 < 0 || n2 > 3)   // :1383-1387
        return argFailM("vocabulary: this is not a correct text file");
    Vocabulary *v = new Vocabulary();
    v->device = device; v->k = k; v->L = L; v->scoring = n1; v->weighting = n2;
    std::vector<std::vector<int32_t>> children(1);
    v->wordId.assign(1, -1); v->weight.assign(1, 0.0); v->desc.assign(32, 0);
    int nWords = 0;
    bool bad = false;
    while (std::getline(f, line)) {
        if (line.find_first_not_of(" \t\r\n") == std::string::npos) continue;
        std::stringstream sn(line);
        int pid = -1, isLeaf = 0;
        sn >> pid >> isLeaf;
        const int nid = (int)children.size();
        if (sn.fail() || pid < 0 || pid >= nid) { bad = true; break; }
        children.emplace_back();
        children[pid].push_back(nid);
        v->desc.resize((size_t)(nid + 1) * 32, 0);
        for (int i = 0; i < 32; i++) {                    // F::fromString (FORB.cpp:121-138)
            int e = 0;
            sn >> e;
            if (!sn.fail()) v->desc[(size_t)nid * 32 + i] = (uint8_t)e;
        }
        double w = 0.0;
        sn >> w;
        v->weight.push_back(w);
        v->wordId.push_back(isLeaf > 0 ? nWords++ : -1);
    }
    if (bad) { delete v; return argFailM("vocabulary: malformed node line"); }
    vocFromChildren(v, children);
    return vocFinish(v, out);
}

void ygzorb_vocabulary_destroy(ygzorb_vocabulary *h) {
    Vocabulary *v = (Vocabulary *)h;
    if (!v) return;
    cudaSetDevice(v->device);
    void *bufs[] = {v->dChildStart, v->dChildList, v->dWordId, v->dWeight, v->dChildDesc};
    for (void *p : bufs) if (p) cudaFree(p);
    delete v;
}

int ygzorb_vocabulary_info(const ygzorb_vocabulary *h, int32_t *info6) {
    const Vocabulary *v = (const Vocabulary *)h;
    if (!v || !info6) return argFailM("null argument");
    info6[0] = (int32_t)v->wordId.size(); info6[1] = v->k; info6[2] = v->L; info6[3] = v->nWords;
    info6[4] = (int32_t)v->childList.size(); info6[5] = v->scoring * 16 + v->weighting;
    return 0;
}

int ygzorb_vocabulary_arrays(const ygzorb_vocabulary *h, int32_t *childStart, int32_t *childList, uint8_t *desc, int32_t *wordId,
                             double *weight) {
    const Vocabulary *v = (const Vocabulary *)h;
    if (!v) return argFailM("null argument");
    const size_t n = v->wordId.size();
    if (childStart) memcpy(childStart, v->childStart.data(), (n + 1) * 4);
    if (childList && !v->childList.empty()) memcpy(childList, v->childList.data(), v->childList.size() * 4);
    if (desc) memcpy(desc, v->desc.data(), n * 32);
    if (wordId) memcpy(wordId, v->wordId.data(), n * 4);
    if (weight) memcpy(weight, v->weight.data(), n * 8);
    return 0;
}

int ygzorb_bow_transform(ygzorb_matcher *h, const ygzorb_vocabulary *vh, const uint8_t *desc, const int32_t *nFeat, int nFrames,
                         int stride, int levelsup, unsigned flags, int32_t *word, int32_t *node, double *weight, int32_t *bowCount,
                         int32_t *bowWord, double *bowValue, int32_t *fvCount, uint32_t *fvNode, int32_t *fvStart, uint32_t *fvFeat) {
    Matcher *m = (Matcher *)h;
    const Vocabulary *v = (const Vocabulary *)vh;
    if (!m || !v || !nFeat || (nFrames > 0 && !desc)) return argFailM("null argument");
    if (!bowCount || !bowWord || !bowValue || !fvCount || !fvNode || !fvStart || !fvFeat) return argFailM("null output pointer");
    if (nFrames < 0 || stride < 1 || stride > 8192) return argFailM("stride must be in [1, 8192]");
    if (v->device != m->device) return argFailM("vocabulary and matcher live on different devices");
    if (nFrames == 0) return 0;
    YGZ_CUDA(cudaSetDevice(m->device));
    const bool inDev = flags & YGZORB_IN_DEVICE, outDev = flags & YGZORB_OUT_DEVICE;
    const size_t nf = nFrames, ns = nf * stride;
    BowArgs A;
    memset(&A, 0, sizeof A);
    A.nFrames = nFrames; A.stride = stride; A.L = v->L; A.levelsup = levelsup;
    A.childStart = v->dChildStart; A.childList = v->dChildList; A.childDesc = v->dChildDesc; A.wordId = v->dWordId;
    A.nodeWeight = v->dWeight;
    int rc;
    if ((rc = stageIn(m, 0, nFeat, nf, inDev, &A.nFeat))) return rc;
    if ((rc = stageIn(m, 1, desc, ns * 32, inDev, &A.desc))) return rc;
    if ((rc = stageOut(m, 2, word, ns, outDev, &A.word))) return rc;
    if ((rc = stageOut(m, 3, node, ns, outDev, &A.node))) return rc;
    if ((rc = stageOut(m, 4, weight, ns, outDev, &A.weight))) return rc;
    if ((rc = stageOut(m, 5, bowCount, nf, outDev, &A.bowCount))) return rc;
    if ((rc = stageOut(m, 6, bowWord, ns, outDev, &A.bowWord))) return rc;
    if ((rc = stageOut(m, 7, bowValue, ns, outDev, &A.bowValue))) return rc;
    if ((rc = stageOut(m, 8, fvCount, nf, outDev, &A.fvCount))) return rc;
    if ((rc = stageOut(m, 9, fvNode, ns, outDev, &A.fvNode))) return rc;
    if ((rc = stageOut(m, 10, fvStart, nf * (stride + 1), outDev, &A.fvStart))) return rc;
    if ((rc = stageOut(m, 11, fvFeat, ns, outDev, &A.fvFeat))) return rc;
    m->launches += launchBowTransform(A, m->stream);
    if ((rc = copyOut(m, word, A.word, ns, outDev))) return rc;
    if ((rc = copyOut(m, node, A.node, ns, outDev))) return rc;
    if ((rc = copyOut(m, weight, A.weight, ns, outDev))) return rc;
    if ((rc = copyOut(m, bowCount, A.bowCount, nf, outDev))) return rc;
    if ((rc = copyOut(m, bowWord, A.bowWord, ns, outDev))) return rc;
    if ((rc = copyOut(m, bowValue, A.bowValue, ns, outDev))) return rc;
    if ((rc = copyOut(m, fvCount, A.fvCount, nf, outDev))) return rc;
    if ((rc = copyOut(m, fvNode, A.fvNode, ns, outDev))) return rc;
    if ((rc = copyOut(m, fvStart, A.fvStart, nf * (stride + 1), outDev))) return rc;
    if ((rc = copyOut(m, fvFeat, A.fvFeat, ns, outDev))) return rc;
    return finish(m, flags);
}

}  // extern "C"
