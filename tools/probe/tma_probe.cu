// Stand-alone check of the TMA path k_fast_tma uses: 3-D byte tensor (x, y, frame), box 48 x R bytes, descriptor taken from a
// __grid_constant__ array with a run-time index, completion on an mbarrier.  nvcc -gencode arch=compute_100a,code=sm_100a
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
struct Maps { CUtensorMap lv[16]; int ok; };
struct Pad { int x[640]; };   // pushes the maps to the same parameter offset as in the library (2560)
__global__ void k(const __grid_constant__ Pad P, const __grid_constant__ Maps M, int level, int x, int y, int z, int rows, uint8_t *out) {
    extern __shared__ __align__(128) uint8_t sm[];
    const unsigned tile = (unsigned)__cvta_generic_to_shared(sm), bar = tile + 48 * 64;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(48u * rows) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(tile), "l"(&M.lv[level]), "r"(x), "r"(y), "r"(z), "r"(bar) : "memory");
    }
    __syncthreads();
    asm volatile("{\n\t.reg .pred P1;\n\tLAB_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t@P1 bra DONE;\n\tbra LAB_WAIT;\n\tDONE:\n\t}" ::"r"(bar), "r"(0u) : "memory");
    for (int i = threadIdx.x; i < 48 * rows; i += blockDim.x) out[i] = sm[i] + (uint8_t)P.x[0];
}
__global__ void k1(const __grid_constant__ CUtensorMap M, int x, int y, int z, int rows, uint8_t *out) {   // one map, offset 0
    extern __shared__ __align__(128) uint8_t sm[];
    const unsigned tile = (unsigned)__cvta_generic_to_shared(sm), bar = tile + 48 * 64;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(48u * rows) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(tile), "l"(&M), "r"(x), "r"(y), "r"(z), "r"(bar) : "memory");
    }
    __syncthreads();
    asm volatile("{\n\t.reg .pred P1;\n\tLAB_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t@P1 bra DONE;\n\tbra LAB_WAIT;\n\tDONE:\n\t}" ::"r"(bar), "r"(0u) : "memory");
    for (int i = threadIdx.x; i < 48 * rows; i += blockDim.x) out[i] = sm[i];
}
__global__ void k2(const CUtensorMap *M, int x, int y, int z, int rows, uint8_t *out) {   // map in global memory
    extern __shared__ __align__(128) uint8_t sm[];
    const unsigned tile = (unsigned)__cvta_generic_to_shared(sm), bar = tile + 48 * 64;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(48u * rows) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(tile), "l"(M), "r"(x), "r"(y), "r"(z), "r"(bar) : "memory");
    }
    __syncthreads();
    asm volatile("{\n\t.reg .pred P1;\n\tLAB_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t@P1 bra DONE;\n\tbra LAB_WAIT;\n\tDONE:\n\t}" ::"r"(bar), "r"(0u) : "memory");
    for (int i = threadIdx.x; i < 48 * rows; i += blockDim.x) out[i] = sm[i];
}
int main(int argc, char **argv) {
    const int mode = argc > 1 ? atoi(argv[1]) : 0;
    const int W = 752, H = 480, F = 4, rows = 42;
    std::vector<uint8_t> h((size_t)W * H * F);
    for (size_t i = 0; i < h.size(); i++) h[i] = (uint8_t)(i * 2654435761u >> 24);
    uint8_t *d, *o;
    cudaMalloc(&d, h.size()); cudaMalloc(&o, 48 * 64);
    cudaMemcpy(d, h.data(), h.size(), cudaMemcpyHostToDevice);
    void *fnp = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fnp, cudaEnableDefault, &q);
    typedef CUresult (*Fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                           const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    Maps M = {}; Pad P = {};
    const cuuint64_t dims[3] = {W, H, F}, strides[2] = {W, (cuuint64_t)W * H};
    const cuuint32_t box[3] = {48, rows, 1}, es[3] = {1, 1, 1};
    CUresult r = ((Fn)fnp)(&M.lv[3], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode %d (fn %p, q %d)\n", (int)r, fnp, (int)q);
    const int x = argc > 2 ? atoi(argv[2]) : 725, y = argc > 3 ? atoi(argv[3]) : 461, z = 2;   // box reaches past the right edge and the bottom: zero fill
    CUtensorMap *dM; cudaMalloc(&dM, sizeof(CUtensorMap)); cudaMemcpy(dM, &M.lv[3], sizeof(CUtensorMap), cudaMemcpyHostToDevice);
    if (mode == 0) k<<<1, 64, 48 * 64 + 16>>>(P, M, 3, x, y, z, rows, o);
    else if (mode == 1) k1<<<1, 64, 48 * 64 + 16>>>(M.lv[3], x, y, z, rows, o);
    else k2<<<1, 64, 48 * 64 + 16>>>(dM, x, y, z, rows, o);
    cudaError_t e = cudaDeviceSynchronize();
    printf("kernel: %s\n", cudaGetErrorString(e));
    std::vector<uint8_t> g(48 * rows);
    cudaMemcpy(g.data(), o, g.size(), cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int r2 = 0; r2 < rows; r2++) for (int c = 0; c < 48; c++) {
        const int xx = x + c, yy = y + r2;
        const uint8_t want = (xx < W && yy < H) ? h[((size_t)z * H + yy) * W + xx] : 0;
        bad += g[r2 * 48 + c] != want;
    }
    printf("mismatches %d of %d\n", bad, 48 * rows);
    return bad != 0 || e != cudaSuccess;
}
