// Issue-rate probe for the integer SIMD min/max instructions k_fast is built from (sm_100a).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_probe pipe_probe.cu && ./pipe_probe
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

constexpr int kIters = 4096, kChains = 8;

template <int OP> __global__ void __launch_bounds__(256) probe(unsigned *out, unsigned seed) {
    unsigned a[kChains], b = seed * 2654435761u + threadIdx.x, c = seed ^ 0x9e3779b9u;
#pragma unroll
    for (int k = 0; k < kChains; k++) a[k] = b * (k + 3) + c;
    __shared__ unsigned char sm[1024];
    sm[threadIdx.x] = (unsigned char)b; sm[threadIdx.x + 256] = (unsigned char)c;
    __syncthreads();
#pragma unroll 1
    for (int i = 0; i < kIters; i++) {
#pragma unroll
        for (int k = 0; k < kChains; k++) {
            if (OP == 0) a[k] = __vimin3_u16x2(a[k], b, c);
            if (OP == 1) a[k] = __vminu2(a[k], b);
            if (OP == 2) { __half2 h = __hmin2(*(__half2 *)&a[k], *(__half2 *)&b); a[k] = *(unsigned *)&h; }
            if (OP == 3) a[k] = a[k] * b + c;
            if (OP == 4) a[k] = (a[k] & b) ^ c;
            if (OP == 5) { a[k] = (k & 1) ? __vimin3_u16x2(a[k], b, c) : a[k] * b + c; }
            if (OP == 6) { if (k & 1) a[k] = __vimin3_u16x2(a[k], b, c); else { __half2 h = __hmin2(*(__half2 *)&a[k], *(__half2 *)&b); a[k] = *(unsigned *)&h; } }
            if (OP == 7) { a[k] = (k & 3) ? __vimin3_u16x2(a[k], b, c) : (unsigned)sm[(a[k] + k) & 511]; }
            if (OP == 8) a[k] = min(a[k], b);
            if (OP == 9) a[k] = a[k] + b;
            if (OP == 10) { a[k] = (k & 1) ? __vimin3_u16x2(a[k], b, c) : a[k] + b; }
            if (OP == 11) { a[k] = (k & 1) ? __vimin3_u16x2(a[k], b, c) : ((a[k] & b) ^ c); }
            if (OP == 12) { float f = fminf(__uint_as_float(a[k]), __uint_as_float(b)); a[k] = __float_as_uint(f); }
            if (OP == 13) { __half2 h = __hfma2(*(__half2 *)&a[k], *(__half2 *)&b, *(__half2 *)&c); a[k] = *(unsigned *)&h; }
        }
        b += 0x10001u;
    }
    unsigned r = 0;
#pragma unroll
    for (int k = 0; k < kChains; k++) r ^= a[k];
    if (r == 0x12345678u) out[0] = r;
}

template <int OP> void run(const char *name, unsigned *d) {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int ctas = sms * 8;
    probe<OP><<<ctas, 256>>>(d, 1);
    cudaEventRecord(e0);
    probe<OP><<<ctas, 256>>>(d, 2);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double warpInst = (double)ctas * 8 * kIters * kChains;
    const double clk = ms * 1e-3 * khz * 1e3;
    printf("%-28s %8.3f ms  %.3f warp-inst/clk/SMSP (at %d MHz nominal)\n", name, ms, warpInst / clk / sms / 4, khz / 1000);
}

int main() {
    unsigned *d; cudaMalloc(&d, 4);
    run<0>("VIMNMX3.U16x2", d); run<1>("VIMNMX.U16x2", d); run<2>("HMNMX2", d); run<3>("IMAD", d); run<4>("LOP3", d);
    run<8>("IMNMX.U32", d); run<9>("IADD", d); run<12>("FMNMX", d); run<13>("HFMA2", d);
    run<5>("VIMNMX3 + IMAD 1:1", d); run<6>("VIMNMX3 + HMNMX2 1:1", d); run<10>("VIMNMX3 + IADD 1:1", d);
    run<11>("VIMNMX3 + LOP3 1:1", d); run<7>("VIMNMX3 + LDS.U8 3:1", d);
    return 0;
}
