// comm.cu — the multi-GPU layer of the two workloads that shard (BASELINE.json configs[3] and [4]):
//   * batched offline extraction, frame-sharded: every rank extracts a contiguous range of the frames with the single-GPU
//     pipeline; the ONLY communication is the gather of the results to the root rank — counts first, then the packed ragged
//     payload (keypoint records and descriptor rows of exactly the keypoints found), chunk by chunk on a side stream while the
//     next chunk extracts;
//   * brute-force Hamming kNN-2 + ratio test, query-sharded: the train rows are replicated with one ncclBroadcast, every rank
//     scans its query range and writes (idx, d1, d2, match) into its slice of the full-size result arrays; one grouped
//     ncclSend / ncclRecv gathers the slices on the root.
// There is no data-path collective: frames and queries are independent (north_star: "NCCL over NVLink used only to gather
// match results").  One process per GPU (torchrun) creates its communicator with ygzorb_comm_create_rank; a single process
// driving several GPUs uses ygzorb_comm_create (ncclCommInitAll) and one host thread per handle.
//
// NCCL is bound at run time (dlopen of libnccl.so.2, i.e. the copy the process already holds when PyTorch is loaded), so the
// library loads and every single-GPU entry point works where NCCL is absent.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/ygzorb.h"
#include "ygz_common.cuh"

namespace {

struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;
std::mutex g_ncclMutex;

int loadNccl() {
    std::lock_guard<std::mutex> lock(g_ncclMutex);
    if (g_nccl.lib) return 0;
    void *lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) { ygz::setError(std::string("NCCL is not available: ") + dlerror()); return YGZORB_E_STATE; }
#define YGZ_SYM(field, name)                                                                          \
    *(void **)(&g_nccl.field) = dlsym(lib, name);                                                     \
    if (!g_nccl.field) { ygz::setError(std::string("libnccl lacks ") + name); return YGZORB_E_STATE; }
    YGZ_SYM(GetUniqueId, "ncclGetUniqueId") YGZ_SYM(CommInitRank, "ncclCommInitRank") YGZ_SYM(CommInitAll, "ncclCommInitAll")
    YGZ_SYM(CommDestroy, "ncclCommDestroy") YGZ_SYM(Broadcast, "ncclBroadcast") YGZ_SYM(AllGather, "ncclAllGather")
    YGZ_SYM(Send, "ncclSend") YGZ_SYM(Recv, "ncclRecv") YGZ_SYM(GroupStart, "ncclGroupStart") YGZ_SYM(GroupEnd, "ncclGroupEnd")
    YGZ_SYM(GetErrorString, "ncclGetErrorString")
#undef YGZ_SYM
    g_nccl.lib = lib;
    return 0;
}

int ncclFail(ncclResult_t r, const char *what) {
    ygz::setError(std::string(what) + ": " + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "NCCL error"));
    return YGZORB_E_CUDA;
}
#define YGZ_NCCL(call)                                        \
    do {                                                      \
        ncclResult_t r__ = (call);                            \
        if (r__ != ncclSuccess) return ncclFail(r__, #call);  \
    } while (0)

struct Comm {
    int rank = 0, world = 1, device = 0;
    ncclComm_t nccl = nullptr;      // null when world == 1
    cudaStream_t stream = nullptr;  // the side stream the gathers run on
    cudaEvent_t evIn = nullptr, evDone = nullptr;
    // workspace of ygzorb_extract_batch_sharded (grown on demand)
    ygzorb_keypoint *dKps[2] = {}; uint8_t *dDesc[2] = {}; uint8_t *dPack[2] = {}; size_t chunkRecs = 0;
    int32_t *dCounts = nullptr; size_t countsCap = 0;
    int64_t *dTotals = nullptr;     // [world] chunk totals after the all-gather (+1: this rank's own)
    int64_t *hTotals = nullptr;     // pinned mirror
    cudaEvent_t evPacked[2] = {}, evSent[2] = {};
    long long launches = 0;
};

int argFailC(const char *msg) { ygz::setError(msg); return YGZORB_E_ARG; }

void shardRange(long long n, int rank, int world, long long *b, long long *e) {
    const long long base = n / world, rem = n % world;
    *b = rank * base + std::min<long long>(rank, rem);
    *e = *b + base + (rank < rem ? 1 : 0);
}

int finishCreate(Comm *c) {
    YGZ_CUDA(cudaSetDevice(c->device));
    YGZ_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    YGZ_CUDA(cudaEventCreateWithFlags(&c->evIn, cudaEventDisableTiming));
    YGZ_CUDA(cudaEventCreateWithFlags(&c->evDone, cudaEventDisableTiming));
    for (int i = 0; i < 2; i++) {
        YGZ_CUDA(cudaEventCreateWithFlags(&c->evPacked[i], cudaEventDisableTiming));
        YGZ_CUDA(cudaEventCreateWithFlags(&c->evSent[i], cudaEventDisableTiming));
    }
    YGZ_CUDA(cudaMalloc((void **)&c->dTotals, (size_t)(c->world + 1) * sizeof(int64_t)));
    YGZ_CUDA(cudaMallocHost((void **)&c->hTotals, (size_t)(c->world + 1) * sizeof(int64_t)));
    return 0;
}

// ---- packing: records of the frames of one chunk, fixed stride `cap` -> back to back ----------------------------------------
// offsets[f] = exclusive prefix of min(counts[f], cap) over the chunk; *total = the sum.  One CTA of 256 threads sweeping the frames.
__global__ void __launch_bounds__(256) k_pack_offsets(const int32_t *counts, int n, int cap, int64_t *offsets, int64_t *total) {
    __shared__ long long wsum[8];
    __shared__ long long carry;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int b = 0; b < n; b += 256) {
        const int f = b + tid;
        const long long v = f < n ? min(counts[f], cap) : 0;
        long long inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const long long t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
        if (lane == 31) wsum[wid] = inc;
        __syncthreads();
        long long woff = 0, sweep = 0;
        for (int i = 0; i < 8; i++) { if (i < wid) woff += wsum[i]; sweep += wsum[i]; }
        if (f < n) offsets[f] = carry + woff + inc - v;
        __syncthreads();
        if (tid == 0) carry += sweep;
        __syncthreads();
    }
    if (tid == 0) total[0] = carry;
}

// CTA = frame: copies counts[f] keypoint records (28 B) and descriptor rows (32 B) to their packed places.  Packed chunk
// layout: [total x 28 B keypoints][total x 32 B descriptors], both as 4-byte words.
__global__ void __launch_bounds__(128) k_pack_records(const int32_t *counts, const int64_t *offsets, const int64_t *total, int cap,
                                                      const uint32_t *kps, const uint32_t *desc, uint32_t *packed) {
    const int f = blockIdx.x;
    const long long n = min(counts[f], cap), o = offsets[f], tot = total[0];
    const uint32_t *sk = kps + (long long)f * cap * 7, *sd = desc + (long long)f * cap * 8;
    uint32_t *dk = packed + o * 7, *dd = packed + tot * 7 + o * 8;
    for (long long i = threadIdx.x; i < n * 7; i += blockDim.x) dk[i] = sk[i];
    for (long long i = threadIdx.x; i < n * 8; i += blockDim.x) dd[i] = sd[i];
}

// Root, after the last chunk: record offset of every frame in the gathered arrays.  Thread = rank shard (serial prefix).
__global__ void k_frame_offsets(const int32_t *counts, long long nTotal, int world, int cap, int64_t *frameOffset) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= world) return;
    const long long base = nTotal / world, rem = nTotal % world;
    const long long b = r * base + min((long long)r, rem), e = b + base + (r < rem ? 1 : 0);
    long long o = b * cap;                                   // region of rank r starts at its first frame's fixed-stride slot
    for (long long f = b; f < e; f++) { frameOffset[f] = o; o += min(counts[f], cap); }
}

template <class T> int grow(T *&p, size_t &have, size_t need) {
    if (need <= have && p) return 0;
    if (p) YGZ_CUDA(cudaFree(p));
    p = nullptr; have = 0;
    YGZ_CUDA(cudaMalloc((void **)&p, std::max<size_t>(need, 256)));
    have = need;
    return 0;
}

}  // namespace

extern "C" {

int ygzorb_shard_range(long long n, int rank, int world, long long *begin, long long *end) {
    if (n < 0 || world < 1 || rank < 0 || rank >= world || !begin || !end) return argFailC("bad argument");
    shardRange(n, rank, world, begin, end);
    return 0;
}

int ygzorb_comm_unique_id(uint8_t *id128) {
    if (!id128) return argFailC("null argument");
    if (int rc = loadNccl()) return rc;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    YGZ_NCCL(g_nccl.GetUniqueId(&id));
    memcpy(id128, &id, 128);
    return 0;
}

int ygzorb_comm_create_rank(const uint8_t *id128, int rank, int world, int device, ygzorb_comm **out) {
    if (!out || world < 1 || rank < 0 || rank >= world) return argFailC("bad argument");
    if (world > 1 && !id128) return argFailC("null unique id");
    Comm *c = new Comm();
    c->rank = rank; c->world = world; c->device = device;
    int rc = finishCreate(c);
    if (!rc && world > 1) {
        rc = loadNccl();
        if (!rc) {
            ncclUniqueId id;
            memcpy(&id, id128, 128);
            const ncclResult_t r = g_nccl.CommInitRank(&c->nccl, world, id, rank);
            if (r != ncclSuccess) rc = ncclFail(r, "ncclCommInitRank");
        }
    }
    if (rc) { ygzorb_comm_destroy((ygzorb_comm *)c); return rc; }
    *out = (ygzorb_comm *)c;
    return 0;
}

int ygzorb_comm_create(int n_gpus, ygzorb_comm **out) {
    if (!out || n_gpus < 1) return argFailC("bad argument");
    int have = 0;
    YGZ_CUDA(cudaGetDeviceCount(&have));
    if (n_gpus > have) return argFailC("n_gpus exceeds the visible devices");
    std::vector<ncclComm_t> comms((size_t)n_gpus, nullptr);
    if (n_gpus > 1) {
        if (int rc = loadNccl()) return rc;
        std::vector<int> devs((size_t)n_gpus);
        for (int i = 0; i < n_gpus; i++) devs[i] = i;
        YGZ_NCCL(g_nccl.CommInitAll(comms.data(), n_gpus, devs.data()));
    }
    for (int i = 0; i < n_gpus; i++) {
        Comm *c = new Comm();
        c->rank = i; c->world = n_gpus; c->device = i; c->nccl = comms[i];
        out[i] = (ygzorb_comm *)c;
        if (int rc = finishCreate(c)) { for (int k = 0; k <= i; k++) { ygzorb_comm_destroy(out[k]); out[k] = nullptr; } return rc; }
    }
    return 0;
}

void ygzorb_comm_destroy(ygzorb_comm *h) {
    Comm *c = (Comm *)h;
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->nccl && g_nccl.CommDestroy) g_nccl.CommDestroy(c->nccl);
    for (int i = 0; i < 2; i++) {
        if (c->dKps[i]) cudaFree(c->dKps[i]);
        if (c->dDesc[i]) cudaFree(c->dDesc[i]);
        if (c->dPack[i]) cudaFree(c->dPack[i]);
        if (c->evPacked[i]) cudaEventDestroy(c->evPacked[i]);
        if (c->evSent[i]) cudaEventDestroy(c->evSent[i]);
    }
    if (c->dCounts) cudaFree(c->dCounts);
    if (c->dTotals) cudaFree(c->dTotals);
    if (c->hTotals) cudaFreeHost(c->hTotals);
    if (c->evIn) cudaEventDestroy(c->evIn);
    if (c->evDone) cudaEventDestroy(c->evDone);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int ygzorb_comm_info(const ygzorb_comm *h, int *rank, int *world, int *device) {
    const Comm *c = (const Comm *)h;
    if (!c) return argFailC("null communicator");
    if (rank) *rank = c->rank;
    if (world) *world = c->world;
    if (device) *device = c->device;
    return 0;
}

long long ygzorb_comm_launch_count(const ygzorb_comm *h) { return h ? ((const Comm *)h)->launches : 0; }

int ygzorb_hamming_knn2_sharded(ygzorb_comm *hc, ygzorb_matcher *m, const uint8_t *query_shard, long long nq_total, uint8_t *train,
                                int nt, float ratio, int th, unsigned flags, int32_t *best_idx, int32_t *d1, int32_t *d2,
                                int32_t *match) {
    Comm *c = (Comm *)hc;
    if (!c || !m || nq_total < 0 || nt < 0 || (nt > 0 && !train)) return argFailC("bad argument");
    if (!best_idx || !d1 || !d2 || !match) return argFailC("all four result arrays are required (full size, device memory)");
    long long b, e;
    shardRange(nq_total, c->rank, c->world, &b, &e);
    const long long n = e - b;
    if (n > 0 && !query_shard) return argFailC("null query shard");
    if (n > 0x7fffffffLL) return argFailC("query shard too large");
    YGZ_CUDA(cudaSetDevice(c->device));
    cudaStream_t ms = (cudaStream_t)ygzorb_matcher_stream(m);
    // train rows: replicated once from the root (32 MB for 1M rows), on the matcher's stream so the scan follows it
    if (c->world > 1 && (flags & YGZORB_BROADCAST_TRAIN) && nt > 0)
        YGZ_NCCL(g_nccl.Broadcast(train, train, (size_t)nt * 32, ncclUint8, 0, c->nccl, ms));
    // every rank writes its slice of the full-size arrays in place
    if (n > 0) {
        const int rc = ygzorb_hamming_knn2(m, query_shard, (int)n, train, nt, ratio, th, YGZORB_IN_DEVICE | YGZORB_OUT_DEVICE | YGZORB_NO_SYNC,
                                           best_idx + b, d1 + b, d2 + b, match + b);
        if (rc) return rc;
    }
    if (c->world > 1) {   // gather of the slices on the root: one group of sends / receives on the same stream
        int32_t *arr[4] = {best_idx, d1, d2, match};
        YGZ_NCCL(g_nccl.GroupStart());
        if (c->rank == 0) {
            for (int r = 1; r < c->world; r++) {
                long long rb, re;
                shardRange(nq_total, r, c->world, &rb, &re);
                if (re > rb) for (int a = 0; a < 4; a++) YGZ_NCCL(g_nccl.Recv(arr[a] + rb, (size_t)(re - rb), ncclInt32, r, c->nccl, ms));
            }
        } else if (n > 0) {
            for (int a = 0; a < 4; a++) YGZ_NCCL(g_nccl.Send(arr[a] + b, (size_t)n, ncclInt32, 0, c->nccl, ms));
        }
        YGZ_NCCL(g_nccl.GroupEnd());
    }
    YGZ_CUDA(cudaGetLastError());
    if (!(flags & YGZORB_NO_SYNC)) return ygzorb_matcher_sync(m);
    return 0;
}

int ygzorb_extract_batch_sharded(ygzorb_comm *hc, ygzorb_extractor *ex, const uint8_t *frames_shard, long long n_total, int w, int h,
                                 int stride, size_t frame_stride, unsigned flags, int cap, int chunk_frames, int32_t *counts,
                                 int64_t *frame_offset, ygzorb_keypoint *keypoints, uint8_t *desc) {
    Comm *c = (Comm *)hc;
    if (!c || !ex || n_total < 0 || w <= 0 || h <= 0 || cap < 1 || chunk_frames < 1) return argFailC("bad argument");
    const bool root = c->rank == 0;
    if (root && (!counts || !frame_offset || !keypoints || !desc)) return argFailC("the root needs all four output arrays (device memory)");
    long long b, e;
    shardRange(n_total, c->rank, c->world, &b, &e);
    const long long n = e - b;
    if (n > 0 && !frames_shard) return argFailC("null frame shard");
    if (frame_stride == 0) frame_stride = (size_t)stride * h;
    YGZ_CUDA(cudaSetDevice(c->device));
    cudaStream_t xs = (cudaStream_t)ygzorb_extractor_stream(ex);
    // every rank runs the same number of chunk rounds (shard sizes differ by at most one frame)
    long long maxShard = 0;
    for (int r = 0; r < c->world; r++) { long long rb, re; shardRange(n_total, r, c->world, &rb, &re); maxShard = std::max(maxShard, re - rb); }
    const int rounds = (int)((maxShard + chunk_frames - 1) / chunk_frames);
    const size_t recs = (size_t)chunk_frames * cap;
    int rc;
    if (recs > c->chunkRecs) {
        YGZ_CUDA(cudaStreamSynchronize(c->stream));
        size_t dummy;
        for (int i = 0; i < 2; i++) {
            dummy = c->chunkRecs * sizeof(ygzorb_keypoint); if ((rc = grow(c->dKps[i], dummy, recs * sizeof(ygzorb_keypoint)))) return rc;
            dummy = c->chunkRecs * 32; if ((rc = grow(c->dDesc[i], dummy, recs * 32))) return rc;
            dummy = c->chunkRecs * 60; if ((rc = grow(c->dPack[i], dummy, recs * 60))) return rc;
        }
        c->chunkRecs = recs;
    }
    // per-chunk scratch: counts of the chunk, offsets, totals (one region per slot)
    const size_t cntBytes = ((size_t)chunk_frames * 4 + 7) & ~(size_t)7;
    const size_t perSlot = (cntBytes + 4 * sizeof(int64_t) + (size_t)chunk_frames * sizeof(int64_t) + 15) & ~(size_t)15;
    { size_t have = c->countsCap; int32_t *p = c->dCounts; if ((rc = grow(p, have, 2 * perSlot + 64))) return rc; c->dCounts = p; c->countsCap = have; }
    auto slotCounts = [&](int s) { return (int32_t *)((uint8_t *)c->dCounts + s * perSlot); };
    auto slotTotal = [&](int s) { return (int64_t *)((uint8_t *)c->dCounts + s * perSlot + cntBytes); };
    auto slotOffsets = [&](int s) { return slotTotal(s) + 4; };
    // the root's region bases: rank r's records start at its first frame's fixed-stride slot (r_begin * cap)
    std::vector<long long> rankBase((size_t)c->world), rankDone((size_t)c->world, 0), rankBegin((size_t)c->world), rankEnd((size_t)c->world);
    for (int r = 0; r < c->world; r++) { shardRange(n_total, r, c->world, &rankBegin[r], &rankEnd[r]); rankBase[r] = rankBegin[r] * cap; }

    auto gatherRound = [&](int k) -> int {   // exchange of round k's packed chunk; the host waits for the chunk's total only
        const int s = k & 1;
        const long long f0 = (long long)k * chunk_frames;
        YGZ_CUDA(cudaStreamWaitEvent(c->stream, c->evPacked[s], 0));
        int64_t *tot = slotTotal(s);
        if (c->world > 1) YGZ_NCCL(g_nccl.AllGather(tot, c->dTotals, 1, ncclInt64, c->nccl, c->stream));
        else YGZ_CUDA(cudaMemcpyAsync(c->dTotals, tot, sizeof(int64_t), cudaMemcpyDeviceToDevice, c->stream));
        YGZ_CUDA(cudaMemcpyAsync(c->hTotals, c->dTotals, (size_t)c->world * sizeof(int64_t), cudaMemcpyDeviceToHost, c->stream));
        YGZ_CUDA(cudaStreamSynchronize(c->stream));   // the main stream keeps running the next chunk meanwhile
        const long long myFrames = std::max<long long>(0, std::min<long long>(chunk_frames, n - f0));
        if (root) {
            // own chunk: device-to-device; others: receives straight into their places
            const long long t0 = c->hTotals[0];
            if (myFrames) YGZ_CUDA(cudaMemcpyAsync(counts + rankBegin[0] + f0, slotCounts(s), (size_t)myFrames * 4, cudaMemcpyDeviceToDevice, c->stream));
            if (t0) {
                YGZ_CUDA(cudaMemcpyAsync(keypoints + rankBase[0] + rankDone[0], c->dPack[s], (size_t)t0 * 28, cudaMemcpyDeviceToDevice, c->stream));
                YGZ_CUDA(cudaMemcpyAsync(desc + (rankBase[0] + rankDone[0]) * 32, c->dPack[s] + (size_t)t0 * 28, (size_t)t0 * 32, cudaMemcpyDeviceToDevice, c->stream));
            }
            rankDone[0] += t0;
            if (c->world > 1) {
                YGZ_NCCL(g_nccl.GroupStart());
                for (int r = 1; r < c->world; r++) {
                    const long long fr = std::max<long long>(0, std::min<long long>(chunk_frames, (rankEnd[r] - rankBegin[r]) - f0));
                    const long long t = c->hTotals[r];
                    if (fr) YGZ_NCCL(g_nccl.Recv(counts + rankBegin[r] + f0, (size_t)fr, ncclInt32, r, c->nccl, c->stream));
                    if (t) {
                        YGZ_NCCL(g_nccl.Recv(keypoints + rankBase[r] + rankDone[r], (size_t)t * 28, ncclUint8, r, c->nccl, c->stream));
                        YGZ_NCCL(g_nccl.Recv(desc + (rankBase[r] + rankDone[r]) * 32, (size_t)t * 32, ncclUint8, r, c->nccl, c->stream));
                    }
                    rankDone[r] += t;
                }
                YGZ_NCCL(g_nccl.GroupEnd());
            }
        } else {
            const long long t = c->hTotals[c->rank];
            YGZ_NCCL(g_nccl.GroupStart());
            if (myFrames) YGZ_NCCL(g_nccl.Send(slotCounts(s), (size_t)myFrames, ncclInt32, 0, c->nccl, c->stream));
            if (t) {
                YGZ_NCCL(g_nccl.Send(c->dPack[s], (size_t)t * 28, ncclUint8, 0, c->nccl, c->stream));
                YGZ_NCCL(g_nccl.Send(c->dPack[s] + (size_t)t * 28, (size_t)t * 32, ncclUint8, 0, c->nccl, c->stream));
            }
            YGZ_NCCL(g_nccl.GroupEnd());
        }
        YGZ_CUDA(cudaEventRecord(c->evSent[s], c->stream));
        return 0;
    };

    for (int k = 0; k < rounds; k++) {
        const int s = k & 1;
        const long long f0 = (long long)k * chunk_frames;
        const long long m = std::max<long long>(0, std::min<long long>(chunk_frames, n - f0));
        // the slot's buffers are free once round k - 2 has been sent
        if (k >= 2) YGZ_CUDA(cudaStreamWaitEvent(xs, c->evSent[s], 0));
        if (m > 0) {
            rc = ygzorb_extract_batch(ex, frames_shard + (size_t)f0 * frame_stride, (int)m, w, h, stride, frame_stride,
                                      (flags & YGZORB_IN_DEVICE) | YGZORB_OUT_DEVICE | YGZORB_NO_SYNC, cap, slotCounts(s), c->dKps[s], c->dDesc[s]);
            if (rc) return rc;
            k_pack_offsets<<<1, 256, 0, xs>>>(slotCounts(s), (int)m, cap, slotOffsets(s), slotTotal(s));
            k_pack_records<<<(unsigned)m, 128, 0, xs>>>(slotCounts(s), slotOffsets(s), slotTotal(s), cap, (const uint32_t *)c->dKps[s],
                                                        (const uint32_t *)c->dDesc[s], (uint32_t *)c->dPack[s]);
            c->launches += 2;
        } else {
            YGZ_CUDA(cudaMemsetAsync(slotTotal(s), 0, 2 * sizeof(int64_t), xs));
        }
        YGZ_CUDA(cudaEventRecord(c->evPacked[s], xs));
        // while this chunk runs: the exchange of the previous one
        if (k >= 1 && (rc = gatherRound(k - 1))) return rc;
    }
    if (rounds > 0 && (rc = gatherRound(rounds - 1))) return rc;
    if (root && n_total > 0) {
        k_frame_offsets<<<(c->world + 63) / 64, 64, 0, c->stream>>>(counts, n_total, c->world, cap, frame_offset);
        c->launches += 1;
    }
    // the caller's stream (the extractor's) sees the gather: events recorded after this call include it
    YGZ_CUDA(cudaEventRecord(c->evDone, c->stream));
    YGZ_CUDA(cudaStreamWaitEvent(xs, c->evDone, 0));
    YGZ_CUDA(cudaGetLastError());
    if (!(flags & YGZORB_NO_SYNC)) return ygzorb_extractor_sync(ex);
    return 0;
}

}  // extern "C"
