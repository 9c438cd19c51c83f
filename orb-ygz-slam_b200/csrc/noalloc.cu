// noalloc.cu — ygzorb_hamming_knn2_noalloc: the brute-force kNN-2 + ratio test of ygzorb_hamming_knn2 for callers that must
// not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// The kNN kernels (match_kernels.cu: k_knn2<V> + k_knn2_merge) need scratch for the per-split partial results
// (KnnPartial = 12 B per query and train split).  ygzorb_hamming_knn2 takes it from the matcher handle's growable pool
// (cudaMalloc on growth); this entry point takes it from the file-scope device array below, which every GPU that loads the
// library holds until the process exits.  Everything else the caller passes: device pointers in, device pointers out.
//
// In host code the array's NAME is a host-side stand-in, not a device address.  The address a kernel can use comes from
// cudaGetSymbolAddress, called here once per device with that device current; it differs per device.
#include <mutex>

#include "../../include/ygzorb.h"
#include "ygz_common.cuh"
#include "ygz_match.cuh"

namespace {

constexpr size_t kScratchBytes = (size_t)YGZORB_NOALLOC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_knnScratch[kScratchBytes];

constexpr int kMaxDevices = 64;
struct PerDevice {
    bool ready = false;
    void *scratch = nullptr;   // device address of g_knnScratch on this device
    int smCount = 0;
};
PerDevice g_dev[kMaxDevices];
std::mutex g_setupMutex;

// One-time set-up per device, with that device current: resolve the device address of the scratch array.
int setupDevice(int device, PerDevice **out) {
    if (device < 0 || device >= kMaxDevices) { ygz::setError("device index out of range"); return YGZORB_E_ARG; }
    PerDevice &D = g_dev[device];
    std::lock_guard<std::mutex> lock(g_setupMutex);
    if (!D.ready) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_knnScratch);
        if (e != cudaSuccess) return ygz::cudaFail(e, "cudaGetSymbolAddress(g_knnScratch)", __FILE__, __LINE__);
        if (!p) { ygz::setError("cudaGetSymbolAddress returned a null address"); return YGZORB_E_CUDA; }
        int sm = 0;
        e = cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, device);
        if (e != cudaSuccess) return ygz::cudaFail(e, "cudaDeviceGetAttribute", __FILE__, __LINE__);
        D.scratch = p; D.smCount = sm; D.ready = true;
    }
    *out = &D;
    return 0;
}

}  // namespace

extern "C" int ygzorb_hamming_knn2_noalloc(int device, void *stream, const uint8_t *query, int nq, const uint8_t *train, int nt,
                                           float ratio, int th, int32_t *best_idx, int32_t *d1, int32_t *d2, int32_t *match) {
    if (nq < 0 || nt < 0 || (nq > 0 && !query) || (nt > 0 && !train)) { ygz::setError("bad argument"); return YGZORB_E_ARG; }
    if (nq == 0) return 0;
    YGZ_CUDA(cudaSetDevice(device));
    PerDevice *D = nullptr;
    if (int rc = setupDevice(device, &D)) return rc;
    const int splits = ygz::knnSplits(nq, nt, D->smCount, kScratchBytes);   // never plans more scratch than the array holds
    const size_t need = ygz::knnPartialBytes(nq, splits);
    if (need > kScratchBytes) {
        ygz::setError("ygzorb_hamming_knn2_noalloc: " + std::to_string(nq) + " queries x " + std::to_string(splits) +
                      " train splits need " + std::to_string(need) + " B of scratch, the static array holds " +
                      std::to_string(kScratchBytes) + " B (YGZORB_NOALLOC_MAX_QUERIES queries always fit); use ygzorb_hamming_knn2");
        return YGZORB_E_CAPACITY;
    }
    // every (split, query) partial the merge reads is written by the scan kernel of the same call: nothing has to start at zero
    ygz::launchKnn2(query, nq, nt > 0 ? train : (const uint8_t *)D->scratch, nt, ratio, th, splits, D->scratch, 2, best_idx, d1, d2,
                    match, (cudaStream_t)stream);
    YGZ_CUDA(cudaGetLastError());
    return 0;
}
